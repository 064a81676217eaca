#!/usr/bin/env python
"""bench.py -- FP64 evals/s of the batched evaluation path (baobzi_eval_multi) on B200.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload C1..C5] [--points P]
  torchrun --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...      (one rank per GPU, NCCL only for the barrier)

A "step" is one pass of the hot path over the whole batch of synthetic uniform points of the workload
(default C3 = BASELINE.json configs[2]: 3D order 8 smooth function, 1e9 points per GPU, the configuration the
north_star target "1e10 evals/s for a 3D order-8 tree on one B200" is quoted on).  Rank 0 prints ONE JSON line.

  value     whole-job evals/s, points and results resident in HBM, CUDA-event timed on the launching stream
  e2e       the same through the reference-facing C ABI call baobzi_eval_multi with HOST (pinned) buffers:
            H2D of the points and D2H of the results inside the timed region
  roofline  for the dominant kernel, from per-launch CUDA events inside the timed region (baobzi_profile_*)
  cpu_baseline  the reference's CPU path (oracle/_ref when built, else the oracle port) on a bounded sample
  --impl reference  times only that CPU path, on all host cores, and prints the same line shape
"""
import argparse
import ctypes as C
import json
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
               dim=2, order=6, tol=1e-10, center=[0.0, 0.0], hl=[1.0, 1.0], data=20.0, points=1_000_000_000),
    "C5": dict(desc="3D order 12, testfun_3d tol 1e-12", func="bench_fun_3d", dim=3, order=12, tol=1e-12,
               center=[1.5, 3.0, 1.5], hl=[1.0, 1.0, 1.0], data=None, points=1_000_000_000),
}
METRIC = "FP64 evals/sec for baobzi_eval_multi"


def flops_per_eval(dim, o):
    if dim == 1:
        return 3 * o
    if dim == 2:
        return 2 * o * o + 2 * o + 4 * (o - 2) + 6
    return 2 * o ** 3 + 2 * o * o + 2 * o + 6 * (o - 2) + 9


def bytes_per_eval(dim):
    return 8 * (dim + 1)


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def ncu_traffic(stage, points_per_launch):
    """DRAM bytes per launch of the dominant kernel from the committed `ncu --set full` summary (profiles/), scaled
    to this run's points per launch; None when there is no capture of that kernel."""
    import glob
    pat = {"eval_grouped": "*_k4_ncu_summary.json", "sort": "*_sort_fused_ncu_summary.json"}.get(stage)
    if not pat:
        return None
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", pat)))
    if not files:
        return None
    unit = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    try:
        rec = json.load(open(files[-1]))[0]
        tot = 0.0
        for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            v, u = rec[k].split()
            tot += float(v) * unit[u]
        captured_points = 33554432 / 4.0  # scripts/gpu_round.sh: 32 Mi points = 4 chunks of 2048 points per leaf
        return tot * points_per_launch / captured_points
    except Exception:
        return None


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx, self.proc, self.lines = gpu_index, None, []

    def start(self):
        """Start sampling (nvidia-smi needs a moment to come up, so call this before the warm-up steps)."""
        def reader():
            for ln in self.proc.stdout:
                self.lines.append((time.perf_counter(), ln))
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=reader, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def stop(self, t_begin, t_end):
        """Summarise the samples that arrived inside the timed region [t_begin, t_end] (perf_counter times)."""
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        self.thread.join(timeout=2)
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        inside = [ln for (ts, ln) in self.lines if t_begin <= ts <= t_end + 0.06]
        for ln in inside:
            f = [t.strip() for t in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
                pw.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples inside the timed region"],
                    "samples": 0, "samples_total": len(self.lines)}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                "samples": len(sm), "power_w_max": max(pw)}


# ------------------------------------------------------------------------------------------------ CPU side
def cpu_reference_rate(tree_path, cfg, threads, pts_per_thread, repeats=1):
    """The reference's CPU-dispatched baobzi_eval_multi (oracle/_ref, unmodified reference) -- or the oracle port
    when the reference could not be built -- on `threads` contiguous slices from `threads` host threads (legal:
    eval is const in the reference).  Returns (evals/s, info dict).  Only executed for cpu_baseline / --impl reference."""
    from oracle import ref
    dim = cfg["dim"]
    rng = np.random.default_rng(1)
    n = threads * pts_per_thread
    x = np.array(cfg["center"]) + (rng.random((n, dim)) * 2 - 1) * np.array(cfg["hl"])
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

    best = None
    for _ in range(repeats + 1):  # first pass warms caches / page-faults the buffers
        th = [threading.Thread(target=work, args=(k,)) for k in range(threads)]
        t0 = time.perf_counter()
        [t_.start() for t_ in th]
        [t_.join() for t_ in th]
        dt = time.perf_counter() - t0
        best = dt if best is None or _ == 0 else min(best, dt)
        if _ == 0:
            best = None
    isa = "n/a"
    try:
        flags = open("/proc/cpuinfo").read()
        isa = "avx512" if " avx512f" in flags else ("avx2" if " avx2" in flags else "generic")
    except OSError:
        pass
    info = {"kind": kind, "cores": threads, "sample": "%d threads x %d uniform points, same saved tree, %s dispatch"
            % (threads, pts_per_thread, isa), "seconds": best}
    return n / best, info


def cpu_sample_size(cfg, threads):
    # ~10-30 s of CPU work in total: the reference runs ~2e6 (3D o8) .. 1e8 (1D) evals/s per core
    per_core_rate = {1: 8e7, 2: 5e6, 3: 1.5e6}[cfg["dim"]] * (8.0 / cfg["order"]) ** cfg["dim"]
    return int(max(2e5, min(2e7, per_core_rate * 0.6)))


def fit_tree_to_file(cfg, path):
    """Fit the workload's tree with OUR host fitter (baobzi_init) and save it; needs the GPU library loaded."""
    import baobzi_b200
    funcs = C.CDLL(os.path.join(ROOT, "baobzi_b200", "lib", "libbenchfuncs.so"))
    fptr = C.cast(getattr(funcs, cfg["func"]), C.c_void_p).value
    b = baobzi_b200.Baobzi(fin=fptr, dim=cfg["dim"], order=cfg["order"], center=np.array(cfg["center"]),
                           half_length=np.array(cfg["hl"]), tol=cfg["tol"], data=cfg["data"])
    b.save(path)
    return b


def run_reference_arm(args, cfg, rank, world):
    """--impl reference: rank 0 alone times the reference's CPU path with every host core it can use."""
    if rank != 0:
        return
    from oracle import ref
    threads = host_cores()
    pts = cpu_sample_size(cfg, threads)
    with tempfile.TemporaryDirectory() as td:
        path = os.path.join(td, "tree.baobzi")
        if ref.have_ref():  # the reference fits its own tree through its own baobzi_init
            subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "oracle"], check=True, stdout=subprocess.DEVNULL)
            name = {"bench_fun_1d": "testfun_1d", "bench_fun_2d": "testfun_2d", "bench_fun_3d": "testfun_3d",
                    "bench_fun_2d_osc": "testfun_2d_osc"}[cfg["func"]]
            inp = ref.make_input(name, cfg["dim"], cfg["order"], cfg["tol"], data=cfg["data"])
            ref.ref_init(inp, cfg["center"], cfg["hl"]).save(path)
        else:
            import baobzi_b200  # noqa: F401  (fitting needs our library; evaluation below is the oracle port)
            fit_tree_to_file(cfg, path).free()
        rates = []
        for step in range(args.warmup + args.steps):
            r, info = cpu_reference_rate(path, cfg, threads, pts, repeats=1)
            if step >= args.warmup:
                rates.append(r)
    value = float(np.mean(rates))
    line = {"metric": METRIC, "value": value, "unit": "evals/s", "impl": "reference", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * threads * pts / value,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "%s: %s" % (args.workload, cfg["desc"]), "points_per_step": threads * pts,
                       "timing": "host wall clock; bounded sample of the workload"},
            "cpu_baseline": {"value": value, "unit": "evals/s", "cores": info["cores"], "kind": info["kind"],
                             "sample": info["sample"]},
            "e2e": {"value": value, "unit": "evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------ GPU side
def run_ours(args, cfg, rank, world, local_rank):
    import torch
    import baobzi_b200
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the evaluation path exists only as sm_100a kernels")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        import torch.distributed as dist
        # NCCL prints its version banner on stdout at the first collective: keep stdout for the ONE JSON line
        sys.stdout.flush()
        saved_stdout = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=dev)
            warm = torch.zeros(1, device=dev)
            dist.all_reduce(warm)
            torch.cuda.synchronize()
        finally:
            os.dup2(saved_stdout, 1)
            os.close(saved_stdout)
    if not os.path.exists(baobzi_b200.library_path()):
        baobzi_b200.build()
    dim, order = cfg["dim"], cfg["order"]
    n = int(args.points or cfg["points"])

    tmp = tempfile.TemporaryDirectory()
    tree_path = os.path.join(tmp.name, "tree_rank%d.baobzi" % rank)
    t0 = time.perf_counter()
    b = fit_tree_to_file(cfg, tree_path)      # host fit (callback on the host) + flatten + upload to this GPU
    fit_s = time.perf_counter() - t0
    if args.mode:
        b.set_mode(args.mode, args.chunk)
    elif args.chunk:
        b.set_mode("auto", args.chunk)

    # synthetic uniform points x_j = hl_j (2u - 1) + c_j (baobzi_timing.c:98-100), generated on the device,
    # seeded per rank; generated in slabs so the temporaries stay small
    g = torch.Generator(device=dev).manual_seed(1 + rank)
    x = torch.empty((n, dim), dtype=torch.float64, device=dev)
    c = torch.tensor(cfg["center"], dtype=torch.float64, device=dev)
    h = torch.tensor(cfg["hl"], dtype=torch.float64, device=dev)
    slab = 1 << 24
    for o in range(0, n, slab):
        v = x[o:o + slab]
        v.uniform_(0.0, 1.0, generator=g)
        v.mul_(2.0).sub_(1.0).mul_(h).add_(c)
    res = torch.empty(n, dtype=torch.float64, device=dev)
    stream = torch.cuda.current_stream(dev)

    def step():
        rc = b._lib.baobzi_eval_multi_device(b.ptr, x.data_ptr(), res.data_ptr(), n, stream.cuda_stream)
        if rc != 0:
            raise RuntimeError("baobzi_eval_multi_device: CUDA error %d" % rc)

    def barrier():
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize(dev)

    sampler = ClockSampler(local_rank)
    sampler.start()
    for _ in range(args.warmup):
        step()
    barrier()
    launches0 = b.launches
    if not args.no_profile:
        b.profile_begin()                      # per-launch CUDA events inside the timed region (roofline)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_begin = time.perf_counter()
    ev0.record(stream)
    for _ in range(args.steps):
        step()
    ev1.record(stream)
    barrier()
    t_end = time.perf_counter()
    ms = ev0.elapsed_time(ev1)
    stages = b.profile_end() if not args.no_profile else {}
    launches = b.launches - launches0
    clocks = sampler.stop(t_begin, t_end)
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        ms = float(t.item())
        lt = torch.tensor([launches], dtype=torch.int64, device=dev)
        torch.distributed.all_reduce(lt)
        launches = int(lt.item())
    value = world * n * args.steps / (ms * 1e-3)
    nan_count = int(torch.isnan(res[: min(n, 1 << 22)]).sum().item())

    # ---- e2e: the reference-facing call with HOST buffers (pinned), copies inside the timed region
    e2e = None
    if not args.no_e2e:
        import psutil
        need = n * (dim + 1) * 8
        avail = psutil.virtual_memory().available / world  # every rank pins its own buffers at the same time
        n_e = n if avail > 3 * need else max(1 << 20, int(avail / 3 / ((dim + 1) * 8)))
        n_e = min(n_e, 2**31 - 1)
        xh = torch.empty((n_e, dim), dtype=torch.float64, pin_memory=True)
        yh = torch.empty(n_e, dtype=torch.float64, pin_memory=True)
        xh.copy_(x[:n_e])
        torch.cuda.synchronize(dev)
        e_steps = max(1, min(args.steps, 3))
        b._lib.baobzi_eval_multi(b.ptr, xh.data_ptr(), yh.data_ptr(), n_e)  # warm-up (allocates the pipeline slots)
        barrier()
        t0 = time.perf_counter()
        for _ in range(e_steps):
            b._lib.baobzi_eval_multi(b.ptr, xh.data_ptr(), yh.data_ptr(), n_e)
        torch.cuda.synchronize(dev)
        dt = time.perf_counter() - t0
        if world > 1:
            t = torch.tensor([dt], dtype=torch.float64, device=dev)
            torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
            dt = float(t.item())
        same = bool(torch.equal(torch.nan_to_num(yh[:1 << 20]), torch.nan_to_num(res[:1 << 20].cpu())))
        e2e = {"value": world * n_e * e_steps / dt, "unit": "evals/s", "h2d_bytes_per_step": n_e * dim * 8,
               "d2h_bytes_per_step": n_e * 8, "points_per_step": n_e, "steps": e_steps,
               "matches_device_path": same, "timing": "host wall clock around baobzi_eval_multi, pinned host buffers"}
        del xh, yh

    if world > 1:
        barrier()
        torch.distributed.destroy_process_group()
    if rank != 0:
        b.free()
        return
    # ---- roofline of the dominant kernel
    hbm_peak, hbm_src = measured_peaks()
    fp = C.c_double(0.0)
    b._lib.baobzi_b200_fp64_peak(C.byref(fp))
    fps = C.c_double(0.0)   # the same DFMA kernel back to back for 2 s: what the pipe sustains under the board's power cap
    if dim > 1 and not args.no_profile:
        b._lib.baobzi_b200_fp64_peak_sustained(C.byref(fps), C.c_double(2.0))
    roof = None
    if stages:
        name, (k_ms, k_launches, k_points) = max(stages.items(), key=lambda kv: kv[1][0])
        total_ms = sum(v[0] for v in stages.values())
        if k_launches:
            per_launch_s = k_ms * 1e-3 / k_launches
            pts_per_launch = k_points / k_launches
            tf = flops_per_eval(dim, order) * pts_per_launch / per_launch_s / 1e12
            gb = bytes_per_eval(dim) * pts_per_launch / per_launch_s / 1e9
            fp64_bound = dim == 3 or (dim == 2 and tf / max(fp.value, 1e-9) > gb / hbm_peak)
            roof = {"kernel": name, "bound": "fp64" if fp64_bound else "hbm",
                    "achieved": tf if fp64_bound else gb, "peak": fp.value if fp64_bound else hbm_peak,
                    "unit": "TFLOP/s" if fp64_bound else "GB/s",
                    "frac": (tf / fp.value if fp.value else None) if fp64_bound else gb / hbm_peak,
                    "traffic": ncu_traffic(name, pts_per_launch),
                    "peak_source": "DFMA microbenchmark baobzi_b200_fp64_peak, measured in this run" if fp64_bound else hbm_src,
                    "algorithmic_flops_per_eval": flops_per_eval(dim, order), "algorithmic_bytes_per_eval": bytes_per_eval(dim),
                    "avg_launch_ms": per_launch_s * 1e3, "points_per_launch": pts_per_launch,
                    "share_of_step": k_ms / total_ms if total_ms else None,
                    "hbm_gbs_achieved": gb, "hbm_frac": gb / hbm_peak, "fp64_tflops_achieved": tf,
                    "stages_ms_per_step": {k: v[0] / args.steps for k, v in stages.items() if v[1]}}
            if fp64_bound and clocks and clocks.get("sm_mhz") and clocks.get("sm_max_mhz"):
                # `peak` is the burst figure (boost clock); the step itself may run below it under sw_power_cap
                at_clock = fp.value * clocks["sm_mhz"] / clocks["sm_max_mhz"]
                roof["peak_at_step_clock"] = at_clock
                roof["frac_at_step_clock"] = tf / at_clock if at_clock else None
            if fp64_bound and fps.value:
                roof["peak_sustained_2s"] = fps.value
    # ---- CPU baseline on this box's host cores (bounded sample)
    cpu = None
    if not args.no_cpu:
        threads = host_cores()
        r1, info1 = cpu_reference_rate(tree_path, cfg, 1, cpu_sample_size(cfg, 1) // 2)
        rT, infoT = cpu_reference_rate(tree_path, cfg, threads, cpu_sample_size(cfg, threads) // 2)
        cpu = {"value": rT, "unit": "evals/s", "cores": threads, "kind": infoT["kind"], "sample": infoT["sample"],
               "single_thread_value": r1}
    line = {"metric": METRIC, "value": value, "unit": "evals/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "%s: %s" % (args.workload, cfg["desc"]), "points_per_gpu": n,
                       "inputs": "uniform on the domain, device resident, %.1f GB per GPU (larger than L2, no flush needed)"
                                 % (n * (dim + 1) * 8 / 1e9),
                       "parallelism": "tree replicated, points sharded by contiguous range, no collective",
                       "tree": {"nodes": int(b._lib.baobzi_launch_count(b.ptr) * 0 + 0) or None}, "fit_seconds": fit_s,
                       "mode": args.mode or "auto", "nan_results_in_sample": nan_count,
                       "parity": "bit-exact vs reference BAOBZI_ARCH=avx2 (tests/)"},
            "e2e": e2e, "gpu_launches": launches, "clocks": clocks, "roofline": roof, "cpu_baseline": cpu,
            "fp64_peak_tflops_measured": fp.value}
    line["config"].pop("tree")
    print(json.dumps(line), flush=True)
    b.free()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="C3", choices=sorted(WORKLOADS))
    ap.add_argument("--points", type=int, default=0, help="points per GPU (default: the workload's)")
    ap.add_argument("--mode", default="", choices=["", "auto", "direct", "binned", "bucketed"])
    ap.add_argument("--chunk", type=int, default=0)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-profile", action="store_true")
    ap.add_argument("--order", type=int, default=0, help="experiments: the workload's function at another order")
    ap.add_argument("--tol", type=float, default=0.0, help="experiments: another fit tolerance")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    cfg = WORKLOADS[args.workload]
    if args.order and args.order != cfg["order"]:
        cfg = dict(cfg, order=args.order, desc=cfg["desc"] + " [order overridden: %d]" % args.order)
    if args.tol and args.tol != cfg["tol"]:
        cfg = dict(cfg, tol=args.tol, desc=cfg["desc"] + " [tol overridden: %g]" % args.tol)
    if args.impl == "reference":
        run_reference_arm(args, cfg, rank, world)
    else:
        run_ours(args, cfg, rank, world, local_rank)


if __name__ == "__main__":
    main()
