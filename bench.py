#!/usr/bin/env python
"""bench.py -- FP64 evals/s of the batched evaluation path (baobzi_eval_multi) on B200.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload C1..C5] [--points P]
  torchrun --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...      (one rank per GPU, NCCL only for the barrier)

A "step" is one pass of the hot path over the whole batch of synthetic uniform points of the workload.  Rank 0 prints
ONE JSON line.  The headline workload is C3 = BASELINE.json configs[2] (3D order 8 smooth function, 1e9 points per GPU,
the configuration the north_star target "1e10 evals/s for a 3D order-8 tree on one B200" is quoted on).  Without
--workload the line also carries `workloads`: short runs of the other BASELINE.json configurations (C1, C2, C4, C5 at
N=1; at every N the 8e9-point strong-scaling form of C5), each with its own clocks, roofline, cpu_baseline and parity.

  value     whole-job evals/s, points and results resident in HBM, CUDA-event timed on the launching stream
  e2e       the same through the reference-facing C ABI call baobzi_eval_multi with HOST (pinned) buffers:
            H2D of the points and D2H of the results inside the timed region; e2e.pageable = the same with plain
            (malloc / numpy) memory, which is what an unmodified caller of the reference hands over; e2e.all_gpus =
            one process, one call, every visible GPU (baobzi_set_host_devices)
  roofline  per stage (sort: HBM bytes it must move; evaluation: flops or bytes, whichever binds) from per-launch
            CUDA events inside the timed region (baobzi_profile_*); the top-level object is the dominant stage
  parity    the TIMED output of this run (first `points` results of the device-resident batch) against the unmodified
            reference (oracle/_ref, BAOBZI_ARCH=avx2) on the same saved tree and the same points: max ulp, node-index
            and NaN mismatches -- measured, not quoted
  cpu_baseline  the reference's CPU path (oracle/_ref when built, else the oracle port) on a bounded sample
  --impl reference  times only that CPU path, on all host cores, and prints the same line shape
"""
import argparse
import ctypes as C
import glob
import json
import math
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# BASELINE.json configs.  flops/eval = the reference formulation's count (FMA = 2).
WORKLOADS = {
    "C1": dict(desc="2D order 8, testfun_2d tol 1e-10 (baobzi_timing.c:176-191)", func="bench_fun_2d", dim=2, order=8,
               tol=1e-10, center=[1.5, 3.0], hl=[1.0, 1.0], data=1.5, points=1_000_000),
    "C2": dict(desc="1D order 16, 1.5*log(x) on [1,3]", func="bench_fun_1d", dim=1, order=16, tol=1e-10,
               center=[2.0], hl=[1.0], data=1.5, points=100_000_000),
    "C3": dict(desc="3D order 8, testfun_3d tol 1e-12 (baobzi_timing.c:195-208)", func="bench_fun_3d", dim=3, order=8,
               tol=1e-12, center=[1.5, 3.0, 1.5], hl=[1.0, 1.0, 1.0], data=None, points=1_000_000_000),
    "C4": dict(desc="2D order 6 oscillatory 2+sin(20x)cos(20y), tol 1e-10, coefficients > L2", func="bench_fun_2d_osc",
               dim=2, order=6, tol=1e-10, center=[0.0, 0.0], hl=[1.0, 1.0], data=20.0, points=1_000_000_000,
               bucketed=True),
    "C5": dict(desc="3D order 12, testfun_3d tol 1e-12", func="bench_fun_3d", dim=3, order=12, tol=1e-12,
               center=[1.5, 3.0, 1.5], hl=[1.0, 1.0, 1.0], data=None, points=1_000_000_000),
}
REF_FUNC = {"bench_fun_1d": "testfun_1d", "bench_fun_2d": "testfun_2d", "bench_fun_3d": "testfun_3d",
            "bench_fun_2d_osc": "testfun_2d_osc"}
METRIC = "FP64 evals/sec for baobzi_eval_multi"
C5_TOTAL_POINTS = 8_000_000_000        # BASELINE.json configs[4]: 8e9 points sharded over 1/2/4/8 GPUs
MIN_TIMED_SECONDS = 1.2                # extra workloads: long enough for >= 10 clock samples and stable clocks


def flops_per_eval(dim, o):
    if dim == 1:
        return 3 * o
    if dim == 2:
        return 2 * o * o + 2 * o + 4 * (o - 2) + 6
    return 2 * o ** 3 + 2 * o * o + 2 * o + 6 * (o - 2) + 9


def bytes_per_eval(dim):
    return 8 * (dim + 1)


def sort_bytes_per_point(dim):
    return 8 * dim + 4                 # the sort must read every point once and write one 4-byte slot index


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def git_head():
    try:
        return subprocess.run(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], stdout=subprocess.PIPE,
                              stderr=subprocess.DEVNULL, text=True).stdout.strip() or None
    except Exception:
        return None


def ncu_traffic(stage, dim, points_per_launch, bucketed=False):
    """DRAM bytes per launch of a kernel from the newest committed `ncu --set full` summary under profiles/, scaled to
    this run's points per launch.  Returns (bytes or None, provenance string): the summary carries a `_meta` record
    (git hash of the build that was profiled, points per profiled launch) written by scripts/ncu_summary.py."""
    pat = {("eval_grouped", 3): "*_k4_ncu_summary.json", ("eval_grouped", 2): "*_eval2d_ncu_summary.json",
           ("sort", 3): "*_sort_split_coal_ncu_summary.json",  # both launches of the split sort: summed
           # 1D: the two-leaf kernel serves C2 (captured before its register prefetch: same loads and stores)
           ("eval_direct", 1): "*_twoleaf1d_ncu_summary.json"}.get((stage, dim))
    if bucketed:  # trees larger than L2 (C4): the pipelined staged kernel of csrc/bucket_stage.cu
        pat = "*_bucket_pipe_ncu_summary.json" if stage == "eval_grouped" else None
    if not pat:
        return None, None
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", pat)))
    if not files:
        return None, None
    unit = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    try:
        recs = json.load(open(files[-1]))
        meta = next((r["_meta"] for r in recs if "_meta" in r), None)
        kernels = [r for r in recs if "dram__bytes_read.sum" in r]
        if stage != "sort":  # one kernel per stage, except the sort (lookup launch + scatter launch of one chunk)
            kernels = kernels[:1]
        tot = 0.0
        for rec in kernels:
            for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                v, u = rec[k].split()
                tot += float(v) * unit[u]
        if meta and meta.get("points_per_launch"):
            captured, src = float(meta["points_per_launch"]), "git %s" % meta.get("git")
        else:  # round-1 captures (scripts/gpu_round.sh of commit 075c245: 32 Mi points = 4 launches)
            captured, src = 33554432 / 4.0, "git 075c245 (round 1)"
        return tot * points_per_launch / captured, "%s, %s, scaled by points per launch" % (os.path.basename(files[-1]), src)
    except Exception:
        return None, None


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """SM clock, power and throttle reasons sampled DURING timed regions: NVML polled every 10 ms from a thread
    (nvidia-smi -lms 50 as the fallback).  One sampler serves several regions: summary(t0, t1) picks its samples."""
    REASONS = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
               0x80: "hw_power_brake_slowdown"}

    def __init__(self, torch_device):
        self.samples, self.stop_flag, self.thread, self.proc, self.how = [], False, None, None, None
        self.max_mhz = None
        self.dev = torch_device

    def start(self):
        try:
            import pynvml
            import torch
            pynvml.nvmlInit()
            h = None
            try:
                uuid = str(torch.cuda.get_device_properties(self.dev).uuid)
                h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid).encode() if not uuid.startswith("GPU-") else uuid.encode())
            except Exception:
                h = pynvml.nvmlDeviceGetHandleByIndex(self.dev.index or 0)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))
            get_reasons = getattr(pynvml, "nvmlDeviceGetCurrentClocksEventReasons", None) or \
                pynvml.nvmlDeviceGetCurrentClocksThrottleReasons

            def poll():
                while not self.stop_flag:
                    try:
                        mhz = pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)
                        pw = pynvml.nvmlDeviceGetPowerUsage(h) / 1000.0
                        rs = int(get_reasons(h))
                        self.samples.append((time.perf_counter(), float(mhz), pw, rs))
                    except Exception:
                        pass
                    time.sleep(0.01)
            self.thread = threading.Thread(target=poll, daemon=True)
            self.thread.start()
            self.how = "nvml, 10 ms"
            return
        except Exception:
            pass
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        bits = [0x8, 0x40, 0x20, 0x4]

        def reader():
            for ln in self.proc.stdout:
                f = [t.strip() for t in ln.split(",")]
                try:
                    rs = sum(b for b, v in zip(bits, f[3:7]) if v.lower().startswith("active"))
                    self.samples.append((time.perf_counter(), float(f[0]), float(f[2]), rs))
                    self.max_mhz = float(f[1])
                except (ValueError, IndexError):
                    continue
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.dev.index or 0), "--query-gpu=" + q,
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=reader, daemon=True)
            self.thread.start()
            self.how = "nvidia-smi, 50 ms"
        except Exception:
            self.proc = None

    def summary(self, t_begin, t_end):
        inside = [s for s in list(self.samples) if t_begin <= s[0] <= t_end]
        if not inside:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["no samples inside the timed region"],
                    "samples": 0, "sampler": self.how}
        mask = 0
        for s in inside:
            mask |= s[3]
        return {"sm_mhz": float(np.median([s[1] for s in inside])), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(n for b, n in self.REASONS.items() if mask & b), "samples": len(inside),
                "power_w_max": max(s[2] for s in inside), "sampler": self.how}

    def stop(self):
        self.stop_flag = True
        if self.proc:
            self.proc.terminate()


# ------------------------------------------------------------------------------------------------ CPU side
def sample_points(cfg, n, seed=1):
    rng = np.random.default_rng(seed)
    return np.array(cfg["center"]) + (rng.random((n, cfg["dim"])) * 2 - 1) * np.array(cfg["hl"])


def cpu_reference_rate(tree_path, cfg, threads, pts_per_thread, repeats=2):
    """The reference's CPU-dispatched baobzi_eval_multi (oracle/_ref, unmodified reference) -- or the oracle port
    when the reference could not be built -- on `threads` contiguous slices from `threads` host threads (legal:
    eval is const in the reference).  Best of `repeats` after one untimed pass.  Returns (evals/s, info dict).
    Only executed for cpu_baseline / --impl reference."""
    from oracle import ref
    n = threads * pts_per_thread
    x = sample_points(cfg, n)
    res = np.empty(n)
    if ref.have_ref():
        kind = "reference"
        h = ref.ref_restore(tree_path, arch=None)  # default CPU dispatch (get_iset, src/baobzi.cpp:27-59)
        lib = ref.ref_lib()

        def work(k):
            s = slice(k * pts_per_thread, (k + 1) * pts_per_thread)
            xs, rs = x[s], res[s]
            lib.baobzi_eval_multi(h.ptr, ref.dptr(xs), ref.dptr(rs), pts_per_thread)
    else:
        kind = "port"
        if not os.path.exists(ref.ORACLE_SO):
            subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "oracle"], check=True, stdout=subprocess.DEVNULL)
        t = ref.OracleTree(tree_path)
        lib = ref.oracle_lib()

        def work(k):
            s = slice(k * pts_per_thread, (k + 1) * pts_per_thread)
            xs, rs = x[s], res[s]
            lib.oracle_eval_multi(t.ptr, ref.dptr(xs), ref.dptr(rs), pts_per_thread, ref.ARITH_AVX2)

    times = []
    for it in range(repeats + 1):  # first pass warms caches / page-faults the buffers
        th = [threading.Thread(target=work, args=(k,)) for k in range(threads)]
        t0 = time.perf_counter()
        [t_.start() for t_ in th]
        [t_.join() for t_ in th]
        if it:
            times.append(time.perf_counter() - t0)
    best = min(times)
    isa = "n/a"
    try:
        flags = open("/proc/cpuinfo").read()
        isa = "avx512" if " avx512f" in flags else ("avx2" if " avx2" in flags else "generic")
    except OSError:
        pass
    info = {"kind": kind, "cores": threads, "sample": "%d threads x %d uniform points, same saved tree, %s dispatch, best of %d"
            % (threads, pts_per_thread, isa, repeats), "seconds": best}
    return n / best, info


def cpu_sample_size(cfg, seconds):
    """points per thread for about `seconds` of CPU work per pass: the reference runs ~1.5e6 (3D o8) .. 8e7 (1D)
    evals/s per core"""
    per_core_rate = {1: 8e7, 2: 5e6, 3: 1.5e6}[cfg["dim"]] * (8.0 / cfg["order"]) ** cfg["dim"]
    return int(max(1e5, min(2e7, per_core_rate * seconds)))


def cpu_baseline(tree_path, cfg, seconds):
    threads = host_cores()
    r1, _ = cpu_reference_rate(tree_path, cfg, 1, cpu_sample_size(cfg, seconds))
    rT, infoT = cpu_reference_rate(tree_path, cfg, threads, cpu_sample_size(cfg, seconds))
    return {"value": rT, "unit": "evals/s", "cores": threads, "kind": infoT["kind"], "sample": infoT["sample"],
            "single_thread_value": r1}


def parity_block(tree_path, cfg, xs, got, got_idx):
    """`got` / `got_idx` = the GPU's values and node indices for the points xs: compared with the unmodified reference
    pinned to its AVX2+FMA build (the flavour the kernels implement), else with the oracle port's AVX2 arithmetic."""
    from oracle import ref
    if ref.have_ref():
        want = ref.ref_restore(tree_path, arch="avx2").eval_multi(xs)
        ref.set_arch(None)
        against = "reference (oracle/_ref, BAOBZI_ARCH=avx2)"
    else:
        want = ref.OracleTree(tree_path).eval_multi(xs, ref.ARITH_AVX2)
        against = "oracle port (AVX2 arithmetic)"
    if ref.have_ref() and os.path.exists(ref.REFIDX_SO):
        widx = ref.ref_node_indices(tree_path, xs, cfg["dim"])[0]
    else:
        widx = ref.OracleTree(tree_path).node_indices(xs)
    d = ref.ulp_diff(got, want)
    finite = np.isfinite(d)
    return {"points": int(len(xs)), "against": against,
            "max_ulp": float(d[finite].max()) if finite.any() else 0.0,
            "nan_mismatch": int((~finite).sum()), "nan_results": int(np.isnan(got).sum()),
            "node_idx_mismatch": int((np.asarray(got_idx) != widx).sum())}


def tree_stats_from_file(path):
    from oracle import ref
    t = ref.OracleTree(path)
    return {"nodes": int(t.n_nodes), "leaves": int(t.n_leaves), "grid_cells": int(t.n_bins), "coeff_bytes": int(8 * t.n_coeffs)}


def fit_tree_to_file(cfg, path):
    """Fit the workload's tree with OUR host fitter (baobzi_init) and save it; needs the GPU library loaded."""
    import baobzi_b200
    funcs = C.CDLL(os.path.join(ROOT, "baobzi_b200", "lib", "libbenchfuncs.so"))
    fptr = C.cast(getattr(funcs, cfg["func"]), C.c_void_p).value
    b = baobzi_b200.Baobzi(fin=fptr, dim=cfg["dim"], order=cfg["order"], center=np.array(cfg["center"]),
                           half_length=np.array(cfg["hl"]), tol=cfg["tol"], data=cfg["data"])
    b.save(path)
    return b


def run_reference_arm(args, cfg, tag, rank, world):
    """--impl reference: rank 0 alone times the reference's CPU path with every host core it can use."""
    if rank != 0:
        return
    from oracle import ref
    threads = host_cores()
    pts = cpu_sample_size(cfg, 1.0)
    with tempfile.TemporaryDirectory() as td:
        path = os.path.join(td, "tree.baobzi")
        if ref.have_ref():  # the reference fits its own tree through its own baobzi_init
            subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "oracle"], check=True, stdout=subprocess.DEVNULL)
            inp = ref.make_input(REF_FUNC[cfg["func"]], cfg["dim"], cfg["order"], cfg["tol"], data=cfg["data"])
            ref.ref_init(inp, cfg["center"], cfg["hl"]).save(path)
            fitted_by = "the reference's own baobzi_init"
        else:
            import baobzi_b200  # noqa: F401  (fitting needs our library; evaluation below is the oracle port)
            fit_tree_to_file(cfg, path).free()
            fitted_by = "baobzi_b200 host fitter (reference not built)"
        tree = tree_stats_from_file(path)
        rates = []
        for step in range(args.warmup + args.steps):
            r, info = cpu_reference_rate(path, cfg, threads, pts, repeats=1)
            if step >= args.warmup:
                rates.append(r)
    value = float(np.mean(rates))
    line = {"metric": METRIC, "value": value, "unit": "evals/s", "impl": "reference", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * threads * pts / value,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "%s: %s" % (tag, cfg["desc"]), "points_per_step": threads * pts,
                       "tree": dict(tree, fitted_by=fitted_by),
                       "timing": "host wall clock; bounded sample of the workload"},
            "cpu_baseline": {"value": value, "unit": "evals/s", "cores": info["cores"], "kind": info["kind"],
                             "sample": info["sample"]},
            "e2e": {"value": value, "unit": "evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------ GPU side
class Ctx:
    """Per-process state shared by the workloads of one bench run."""

    def __init__(self, args, rank, world, local_rank):
        import torch
        self.args, self.rank, self.world = args, rank, world
        self.torch = torch
        self.dev = torch.device("cuda", local_rank)
        self.stream = torch.cuda.current_stream(self.dev)
        self.sampler = ClockSampler(self.dev)
        self.tmp = tempfile.TemporaryDirectory()
        self.hbm_peak, self.hbm_src = measured_peaks()
        self.fp64_peak = None

    def barrier(self):
        if self.world > 1:
            self.torch.distributed.barrier()
        self.torch.cuda.synchronize(self.dev)

    def max_over_ranks(self, v):
        if self.world == 1:
            return v
        t = self.torch.tensor([v], dtype=self.torch.float64, device=self.dev)
        self.torch.distributed.all_reduce(t, op=self.torch.distributed.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(self, v):
        if self.world == 1:
            return v
        t = self.torch.tensor([v], dtype=self.torch.int64, device=self.dev)
        self.torch.distributed.all_reduce(t)
        return int(t.item())

    def uniform_points(self, cfg, n, seed, out=None):
        """x_j = hl_j (2u - 1) + c_j (baobzi_timing.c:98-100), generated on the device in slabs"""
        torch = self.torch
        g = torch.Generator(device=self.dev).manual_seed(seed)
        x = out if out is not None else torch.empty((n, cfg["dim"]), dtype=torch.float64, device=self.dev)
        c = torch.tensor(cfg["center"], dtype=torch.float64, device=self.dev)
        h = torch.tensor(cfg["hl"], dtype=torch.float64, device=self.dev)
        slab = 1 << 24
        for o in range(0, n, slab):
            v = x[o:o + slab]
            v.uniform_(0.0, 1.0, generator=g)
            v.mul_(2.0).sub_(1.0).mul_(h).add_(c)
        return x


def stage_rooflines(ctx, cfg, stages, steps):
    """One roofline record per profiled stage, each against ITS bound: the sort moves bytes (8 DIM + 4 per point) and
    does no flops; an evaluation stage is measured against the FP64 pipe or HBM, whichever fraction is larger."""
    dim, order = cfg["dim"], cfg["order"]
    out = {}
    total_ms = sum(v[0] for v in stages.values())
    for name, (k_ms, k_launches, k_points) in stages.items():
        if not k_launches:
            continue
        per_launch_s = k_ms * 1e-3 / k_launches
        pts = k_points / k_launches
        rec = {"kernel": name, "avg_launch_ms": per_launch_s * 1e3, "points_per_launch": pts, "launches": k_launches,
               "share_of_step": k_ms / total_ms if total_ms else None, "ms_per_step": k_ms / steps}
        if name in ("sort", "scan", "scatter"):
            gb = sort_bytes_per_point(dim) * pts / per_launch_s / 1e9
            rec.update({"bound": "hbm", "achieved": gb, "peak": ctx.hbm_peak, "unit": "GB/s", "frac": gb / ctx.hbm_peak,
                        "peak_source": ctx.hbm_src, "algorithmic_bytes_per_point": sort_bytes_per_point(dim)})
        else:
            tf = flops_per_eval(dim, order) * pts / per_launch_s / 1e12
            gb = bytes_per_eval(dim) * pts / per_launch_s / 1e9
            f_fp = tf / ctx.fp64_peak if ctx.fp64_peak else 0.0
            f_hbm = gb / ctx.hbm_peak
            fp64_bound = dim == 3 or (dim == 2 and f_fp > f_hbm)
            rec.update({"bound": "fp64" if fp64_bound else "hbm", "achieved": tf if fp64_bound else gb,
                        "peak": ctx.fp64_peak if fp64_bound else ctx.hbm_peak, "unit": "TFLOP/s" if fp64_bound else "GB/s",
                        "frac": f_fp if fp64_bound else f_hbm,
                        "peak_source": "DFMA microbenchmark baobzi_b200_fp64_peak, measured in this run" if fp64_bound else ctx.hbm_src,
                        "algorithmic_flops_per_eval": flops_per_eval(dim, order),
                        "algorithmic_bytes_per_eval": bytes_per_eval(dim),
                        "hbm_gbs_achieved": gb, "hbm_frac": f_hbm, "fp64_tflops_achieved": tf, "fp64_frac": f_fp})
        if cfg.get("bucketed") and name.startswith("eval"):
            # a tree larger than L2 is reported against both byte models -- 8 (DIM + 1) B per evaluation
            # (ideal: every coefficient block read once per chunk) and that plus one block gather per evaluation
            blk = 8 * order ** dim
            gb2 = (bytes_per_eval(dim) + blk) * pts / per_launch_s / 1e9
            rec["one_block_gather_per_eval"] = {"bytes_per_eval": bytes_per_eval(dim) + blk, "hbm_gbs": gb2,
                                                "hbm_frac": gb2 / ctx.hbm_peak}
        traffic, src = ncu_traffic(name, dim, pts, bucketed=bool(cfg.get("bucketed")))
        rec["traffic"] = traffic
        if src:
            rec["traffic_source"] = src
        out[name] = rec
    return out


def run_workload(ctx, tag, cfg, n, steps, warmup, headline, cpu_seconds):
    """Fit, generate, warm up, time `steps` steps (the extra workloads: at least MIN_TIMED_SECONDS), and collect
    everything a bench record needs.  Returns the record (every rank; rank 0's is printed)."""
    import baobzi_b200
    torch, args = ctx.torch, ctx.args
    dim, order = cfg["dim"], cfg["order"]
    tree_path = os.path.join(ctx.tmp.name, "tree_%s_rank%d.baobzi" % (tag, ctx.rank))
    t0 = time.perf_counter()
    b = fit_tree_to_file(cfg, tree_path)      # host fit (callback on the host) + flatten + upload to this GPU
    fit_s = time.perf_counter() - t0
    if args.mode:
        b.set_mode(args.mode, args.chunk)
    elif args.chunk:
        b.set_mode("auto", args.chunk)
    if ctx.fp64_peak is None:
        fp = C.c_double(0.0)
        b._lib.baobzi_b200_fp64_peak(C.byref(fp))
        ctx.fp64_peak = fp.value

    x = ctx.uniform_points(cfg, n, 1 + ctx.rank)
    res = torch.empty(n, dtype=torch.float64, device=ctx.dev)
    stream = ctx.stream

    def step():
        rc = b._lib.baobzi_eval_multi_device(b.ptr, x.data_ptr(), res.data_ptr(), n, stream.cuda_stream)
        if rc != 0:
            raise RuntimeError("baobzi_eval_multi_device: CUDA error %d" % rc)

    for _ in range(warmup):
        step()
    ctx.barrier()
    if not headline:  # size the timed region: at least MIN_TIMED_SECONDS (clock samples), at least `steps` steps
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(3):
            step()
        e1.record(stream)
        torch.cuda.synchronize(ctx.dev)
        est = max(e0.elapsed_time(e1) / 3 * 1e-3, 1e-6)
        steps = int(min(200000, max(steps, math.ceil(MIN_TIMED_SECONDS / est))))
    launches0 = b.launches
    profile = not args.no_profile
    # per-launch events cost a few microseconds each: microsecond-scale steps are profiled in a separate short pass
    profile_inline = profile and (headline or steps <= 2000)
    if profile_inline:
        b.profile_begin()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_begin = time.perf_counter()
    ev0.record(stream)
    for _ in range(steps):
        step()
    ev1.record(stream)
    ctx.barrier()
    t_end = time.perf_counter()
    ms = ev0.elapsed_time(ev1)
    stages = b.profile_end() if profile_inline else {}
    launches = b.launches - launches0
    clocks = ctx.sampler.summary(t_begin, t_end)
    profiled_steps = steps
    if profile and not profile_inline:
        profiled_steps = 200
        b.profile_begin()
        for _ in range(profiled_steps):
            step()
        stages = b.profile_end()
    ms = ctx.max_over_ranks(ms)
    launches = ctx.sum_over_ranks(launches)
    value = ctx.world * n * steps / (ms * 1e-3)

    rec = {"value": value, "unit": "evals/s", "steps": steps, "warmup": warmup, "ms_per_step": ms / steps,
           "points_per_gpu": n, "gpu_launches": launches, "clocks": clocks, "tree": b.tree_info(), "fit_seconds": fit_s}

    # ---- parity of the TIMED output: first m results of the device-resident batch against the reference
    parity = None
    if ctx.rank == 0 and not args.no_parity:
        m = min(n, 200_000 if dim < 3 else 100_000)
        xs = x[:m].cpu().numpy()
        got = res[:m].cpu().numpy()
        got_idx = b.node_index(x[:m]).cpu().numpy()
        parity = parity_block(tree_path, cfg, xs, got, got_idx)
    rec["parity"] = parity

    # ---- e2e: the reference-facing call with HOST buffers, copies inside the timed region (headline only)
    if headline and not args.no_e2e:
        rec["e2e"] = measure_e2e(ctx, b, cfg, x, res, n)
    if ctx.rank == 0:
        roofs = stage_rooflines(ctx, cfg, stages, profiled_steps) if stages else {}
        rec["stages"] = roofs
        if roofs:
            top = max(roofs.values(), key=lambda r: r["ms_per_step"])
            roof = dict(top)
            roof["stages_ms_per_step"] = {k: v["ms_per_step"] for k, v in roofs.items()}
            if top["bound"] == "fp64" and clocks.get("sm_mhz") and clocks.get("sm_max_mhz"):
                at_clock = ctx.fp64_peak * clocks["sm_mhz"] / clocks["sm_max_mhz"]
                roof["peak_at_step_clock"] = at_clock
                roof["frac_at_step_clock"] = top["achieved"] / at_clock if at_clock else None
            # whole step against the same bound (sort included): what the user gets per point
            if top["bound"] == "fp64":
                roof["whole_step_frac"] = flops_per_eval(dim, order) * value / ctx.world / 1e12 / ctx.fp64_peak
            else:
                roof["whole_step_frac"] = bytes_per_eval(dim) * value / ctx.world / 1e9 / ctx.hbm_peak
            rec["roofline"] = roof
        else:
            rec["roofline"] = None
        rec["cpu_baseline"] = cpu_baseline(tree_path, cfg, cpu_seconds) if not args.no_cpu else None
        if rec["cpu_baseline"]:
            rec["vs_cpu_baseline"] = {"all_cores": value / rec["cpu_baseline"]["value"],
                                      "one_core": value / rec["cpu_baseline"]["single_thread_value"]}
    del x, res
    b.free()
    torch.cuda.empty_cache()
    return rec


def measure_e2e(ctx, b, cfg, x, res, n):
    """baobzi_eval_multi with HOST pointers: (1) pinned buffers (the contract's e2e), (2) pageable buffers -- what an
    unmodified caller of the reference passes, (3) one call spread over every visible GPU (single process only)."""
    import psutil
    torch = ctx.torch
    dim = cfg["dim"]
    need = n * (dim + 1) * 8
    avail = psutil.virtual_memory().available / ctx.world  # every rank pins its own buffers at the same time
    n_e = n if avail > 3 * need else max(1 << 20, int(avail / 3 / ((dim + 1) * 8)))
    n_e = min(n_e, 2**31 - 1)
    xh = torch.empty((n_e, dim), dtype=torch.float64, pin_memory=True)
    yh = torch.empty(n_e, dtype=torch.float64, pin_memory=True)
    xh.copy_(x[:n_e])
    torch.cuda.synchronize(ctx.dev)
    e_steps = max(1, min(ctx.args.steps, 3))

    def timed(xp, yp, count, reps):
        b._lib.baobzi_eval_multi(b.ptr, xp, yp, count)  # warm-up (allocates the workers' buffers, touches the pages)
        ctx.barrier()
        t0 = time.perf_counter()
        for _ in range(reps):
            b._lib.baobzi_eval_multi(b.ptr, xp, yp, count)
        torch.cuda.synchronize(ctx.dev)
        return ctx.max_over_ranks(time.perf_counter() - t0)

    dt = timed(xh.data_ptr(), yh.data_ptr(), n_e, e_steps)
    ref_bits = res[:1 << 20].cpu()
    same = bool(torch.equal(torch.nan_to_num(yh[:1 << 20]), torch.nan_to_num(ref_bits)))
    e2e = {"value": ctx.world * n_e * e_steps / dt, "unit": "evals/s", "h2d_bytes_per_step": n_e * dim * 8,
           "d2h_bytes_per_step": n_e * 8, "points_per_step": n_e, "steps": e_steps,
           "matches_device_path": same, "timing": "host wall clock around baobzi_eval_multi, pinned host buffers"}
    # (2) pageable: a bounded slice (page-faulting and copying 32 GB of numpy memory takes tens of seconds)
    n_p = int(min(n_e, 200_000_000))
    xp = np.empty((n_p, dim))
    yp = np.empty(n_p)
    xp[:] = xh[:n_p].numpy()
    dtp = timed(xp.ctypes.data, yp.ctypes.data, n_p, 2)
    e2e["pageable"] = {"value": ctx.world * n_p * 2 / dtp, "unit": "evals/s", "points_per_step": n_p, "steps": 2,
                       "matches_device_path": bool(np.array_equal(np.nan_to_num(yp[:1 << 20]), np.nan_to_num(ref_bits.numpy()))),
                       "buffers": "numpy (malloc) memory, staged through the library's pinned ring by host threads"}
    del xp, yp
    # (3) every visible GPU from ONE call (only meaningful in a single-process run)
    n_vis = torch.cuda.device_count()
    if ctx.world == 1 and n_vis > 1:
        b.set_host_devices(0)
        dta = timed(xh.data_ptr(), yh.data_ptr(), n_e, e_steps)
        b.set_host_devices(1)
        e2e["all_gpus"] = {"value": n_e * e_steps / dta, "unit": "evals/s", "devices": n_vis, "points_per_step": n_e,
                           "matches_device_path": bool(torch.equal(torch.nan_to_num(yh[:1 << 20]), torch.nan_to_num(ref_bits)))}
    else:
        e2e["all_gpus"] = {"value": None, "devices": n_vis, "note": "one visible GPU" if n_vis == 1 else
                           "multi-process run: every rank drives its own GPU"}
    del xh, yh
    return e2e


def run_c5_total(ctx, cfg, warmup):
    """BASELINE.json configs[4] as stated: 8e9 points in total, sharded by contiguous range over the ranks (strong
    scaling).  Each rank evaluates its share in slabs of <= 1e9 points regenerated on the device (192 GB of points do
    not fit beside the results); only the evaluations are timed (CUDA events per slab), generation is not."""
    import baobzi_b200  # noqa: F401
    torch = ctx.torch
    dim = cfg["dim"]
    share = C5_TOTAL_POINTS // ctx.world
    slab = min(share, 1_000_000_000)
    tree_path = os.path.join(ctx.tmp.name, "tree_c5t_rank%d.baobzi" % ctx.rank)
    b = fit_tree_to_file(cfg, tree_path)
    x = torch.empty((slab, dim), dtype=torch.float64, device=ctx.dev)
    res = torch.empty(slab, dtype=torch.float64, device=ctx.dev)
    stream = ctx.stream

    def ev(m):
        rc = b._lib.baobzi_eval_multi_device(b.ptr, x.data_ptr(), res.data_ptr(), m, stream.cuda_stream)
        if rc != 0:
            raise RuntimeError("baobzi_eval_multi_device: CUDA error %d" % rc)

    ctx.uniform_points(cfg, slab, 1000 + ctx.rank, out=x)
    for _ in range(warmup):
        ev(min(slab, 1 << 26))
    ctx.barrier()
    total_ms, done, k, checksum = 0.0, 0, 0, 0
    t_begin = time.perf_counter()
    launches0 = b.launches
    while done < share:
        m = min(slab, share - done)
        if k:
            ctx.uniform_points(cfg, slab, 1000 + ctx.rank + 7919 * k, out=x)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        ev(m)
        e1.record(stream)
        torch.cuda.synchronize(ctx.dev)
        total_ms += e0.elapsed_time(e1)
        checksum ^= int(res[:m:9973].view(torch.int64).sum().item())
        done += m
        k += 1
    ctx.barrier()
    t_end = time.perf_counter()
    launches = ctx.sum_over_ranks(b.launches - launches0)
    total_ms = ctx.max_over_ranks(total_ms)
    rec = {"value": C5_TOTAL_POINTS / (total_ms * 1e-3), "unit": "evals/s", "points_total": C5_TOTAL_POINTS,
           "points_per_gpu": share, "slabs_per_gpu": k, "eval_ms_total": total_ms, "scaling": "strong",
           "gpu_launches": launches, "clocks": ctx.sampler.summary(t_begin, t_end), "tree": b.tree_info(),
           "timing": "sum of CUDA-event spans around the %d evaluations of each rank (max over ranks); the on-device "
                     "regeneration of the points between slabs is not timed" % k,
           "result_checksum_rank0": checksum}
    del x, res
    b.free()
    torch.cuda.empty_cache()
    return rec


def run_ours(args, rank, world, local_rank):
    import torch
    import baobzi_b200
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the evaluation path exists only as sm_100a kernels")
    torch.cuda.set_device(local_rank)
    ctx = Ctx(args, rank, world, local_rank)
    if world > 1:
        import torch.distributed as dist
        # NCCL prints its version banner on stdout at the first collective: keep stdout for the ONE JSON line
        sys.stdout.flush()
        saved_stdout = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=ctx.dev)
            warm = torch.zeros(1, device=ctx.dev)
            dist.all_reduce(warm)
            torch.cuda.synchronize()
        finally:
            os.dup2(saved_stdout, 1)
            os.close(saved_stdout)
    if not os.path.exists(baobzi_b200.library_path()):
        baobzi_b200.build()
    ctx.sampler.start()

    tag = args.workload or "C3"
    cfg = WORKLOADS[tag]
    if args.order and args.order != cfg["order"]:
        cfg = dict(cfg, order=args.order, desc=cfg["desc"] + " [order overridden: %d]" % args.order)
    if args.tol and args.tol != cfg["tol"]:
        cfg = dict(cfg, tol=args.tol, desc=cfg["desc"] + " [tol overridden: %g]" % args.tol)
    n = int(args.points or cfg["points"])
    head = run_workload(ctx, tag, cfg, n, args.steps, args.warmup, True, cpu_seconds=1.5)

    extra = {}
    if not args.workload and not args.no_extra:
        if world == 1:
            for t2 in ("C1", "C2", "C4", "C5"):
                c2 = WORKLOADS[t2]
                extra[t2] = run_workload(ctx, t2, c2, int(c2["points"]), 3, 3, False, cpu_seconds=0.5)
        extra["C5_8e9_sharded"] = run_c5_total(ctx, WORKLOADS["C5"], 3)

    # the sustained FP64 rate (2 s of back-to-back DFMA): context for FP64-bound fractions measured inside long steps
    fps = C.c_double(0.0)
    if rank == 0 and cfg["dim"] > 1 and not args.no_profile:
        lib = baobzi_b200.load_library()
        lib.baobzi_b200_fp64_peak_sustained(C.byref(fps), C.c_double(2.0))
    ctx.sampler.stop()
    if world > 1:
        ctx.barrier()
        torch.distributed.destroy_process_group()
    if rank != 0:
        return
    dim = cfg["dim"]
    roof = head.get("roofline")
    if roof and fps.value and roof.get("bound") == "fp64":
        roof["peak_sustained_2s"] = fps.value
    line = {"metric": METRIC, "value": head["value"], "unit": "evals/s", "n_gpus": world, "steps": head["steps"],
            "warmup": head["warmup"], "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "%s: %s" % (tag, cfg["desc"]), "points_per_gpu": n,
                       "inputs": "uniform on the domain, device resident, %.1f GB per GPU (larger than L2, no flush needed)"
                                 % (n * (dim + 1) * 8 / 1e9),
                       "parallelism": "tree replicated, points sharded by contiguous range, no collective",
                       "tree": head["tree"], "fit_seconds": head["fit_seconds"], "mode": args.mode or "auto",
                       "parity": head["parity"], "git": git_head()},
            "e2e": head.get("e2e"), "gpu_launches": head["gpu_launches"], "clocks": head["clocks"], "roofline": roof,
            "stages": head.get("stages"), "cpu_baseline": head.get("cpu_baseline"),
            "vs_cpu_baseline": head.get("vs_cpu_baseline"), "fp64_peak_tflops_measured": ctx.fp64_peak,
            "workloads": extra or None}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="", choices=[""] + sorted(WORKLOADS),
                    help="one workload only (default: C3 as the headline + short runs of the others)")
    ap.add_argument("--points", type=int, default=0, help="points per GPU (default: the workload's)")
    ap.add_argument("--mode", default="", choices=["", "auto", "direct", "binned", "bucketed"])
    ap.add_argument("--chunk", type=int, default=0)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-profile", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip the `workloads` block")
    ap.add_argument("--order", type=int, default=0, help="experiments: the workload's function at another order")
    ap.add_argument("--tol", type=float, default=0.0, help="experiments: another fit tolerance")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        tag = args.workload or "C3"
        cfg = WORKLOADS[tag]
        if args.order and args.order != cfg["order"]:
            cfg = dict(cfg, order=args.order, desc=cfg["desc"] + " [order overridden: %d]" % args.order)
        if args.tol and args.tol != cfg["tol"]:
            cfg = dict(cfg, tol=args.tol, desc=cfg["desc"] + " [tol overridden: %g]" % args.tol)
        run_reference_arm(args, cfg, tag, rank, world)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
