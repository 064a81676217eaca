"""bench.py's output contract, checked on CPU: every committed bench line under profiles/ carries the keys the driver and
the judge read (roofline, cpu_baseline or its reference-arm twin, e2e, clocks, gpu_launches ...), and without a GPU
the product arm refuses to run instead of falling back to anything."""
import glob
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LINES = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_bench_*.json")))

BASE_KEYS = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
             "vs_baseline", "dtype", "data", "config"}


def _load(path):
    with open(path) as f:
        text = f.read().strip()
    return json.loads(text.splitlines()[-1])


@pytest.mark.parametrize("path", LINES, ids=[os.path.basename(p) for p in LINES])
def test_committed_bench_lines(path):
    d = _load(path)
    assert BASE_KEYS <= set(d), BASE_KEYS - set(d)
    assert d["metric"].startswith("FP64 evals/sec") and d["unit"] == "evals/s" and d["higher_is_better"] is True
    assert d["dtype"] == "f64" and d["data"] == "synthetic" and d["scaling"] == "weak" and d["vs_baseline"] is None
    assert "workload" in d["config"] and "model" not in d["config"]
    assert d["value"] > 0 and d["ms_per_step"] > 0
    if d.get("impl") == "reference":
        assert d["gpu_launches"] == 0
        assert d["cpu_baseline"]["kind"] in ("reference", "port") and d["cpu_baseline"]["cores"] >= 1
        assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["value"] == d["value"]
        return
    assert d["gpu_launches"] > 0
    assert d["warmup"] >= 3
    r = d["roofline"]
    assert r["bound"] in ("hbm", "fp64") and r["unit"] in ("GB/s", "TFLOP/s")
    assert 0 < r["frac"] <= 1.0 and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9
    assert "traffic" in r
    c = d["clocks"]
    assert not set(c["reasons"]) & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}
    if d.get("e2e"):
        e = d["e2e"]
        assert e["h2d_bytes_per_step"] > 0 and e["d2h_bytes_per_step"] > 0 and e["value"] < d["value"]
    # whole-job value = points of all ranks per second
    pts = d["config"]["points_per_gpu"] * d["n_gpus"]
    assert abs(d["value"] - pts / (d["ms_per_step"] * 1e-3)) / d["value"] < 1e-6


R02 = [p for p in LINES if os.path.basename(p).startswith("r02_bench_default")]


@pytest.mark.parametrize("path", R02, ids=[os.path.basename(p) for p in R02])
def test_round2_default_line_carries_every_configuration(path):
    """The default invocation: headline C3 plus short runs of the other BASELINE.json configurations, each with a clock
    record of >= 10 samples, its own roofline against its own bound, a cpu_baseline and a MEASURED parity block."""
    d = _load(path)
    par = d["config"]["parity"]
    assert par["max_ulp"] == 0 and par["node_idx_mismatch"] == 0 and par["nan_mismatch"] == 0 and par["points"] >= 100000
    assert d["config"]["tree"]["leaves"] == 4096 and d["config"]["tree"]["nodes"] == 4096
    assert d["clocks"]["samples"] >= 10
    assert d["e2e"]["pageable"]["value"] > 0 and d["e2e"]["pageable"]["matches_device_path"] is True
    w = d["workloads"]
    assert {"C1", "C2", "C4", "C5", "C5_8e9_sharded"} <= set(w)
    for tag in ("C1", "C2", "C4", "C5"):
        r = w[tag]
        assert r["clocks"]["samples"] >= 10, tag
        assert not set(r["clocks"]["reasons"]) & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}, tag
        assert r["parity"]["max_ulp"] == 0 and r["parity"]["node_idx_mismatch"] == 0, tag
        assert r["cpu_baseline"]["kind"] in ("reference", "port") and r["cpu_baseline"]["value"] > 0, tag
        assert r["tree"]["leaves"] > 0 and r["value"] > 0
        for name, st in r["stages"].items():          # a sort is measured against the bytes it moves, never against flops
            assert st["bound"] == ("hbm" if name in ("sort", "scan", "scatter") else st["bound"]), (tag, name)
            assert abs(st["frac"] - st["achieved"] / st["peak"]) < 1e-9
    assert w["C4"]["tree"]["leaves"] > 1_000_000
    assert w["C5_8e9_sharded"]["points_total"] == 8_000_000_000 and w["C5_8e9_sharded"]["scaling"] == "strong"


def test_there_are_committed_lines():
    names = [os.path.basename(p) for p in LINES]
    assert any("c3_1e9" in n for n in names) and any("ref" in n for n in names)


def test_product_arm_refuses_to_run_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("needs a machine without a GPU")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "1", "--warmup", "1"],
                       stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=300)
    assert r.returncode != 0
    assert r.stdout.strip() == ""                       # no JSON line: nothing was measured
    assert "CUDA" in r.stderr or "cuda" in r.stderr


def test_tree_larger_than_l2_is_reported_against_both_byte_models():
    """C4 reports the ideal 8 (DIM + 1) B per evaluation AND that plus one coefficient-block gather per
    evaluation; its `traffic` comes from the ncu capture of the kernel that serves it (csrc/bucket_stage.cu)."""
    import bench

    class Ctx:
        hbm_peak, hbm_src, fp64_peak = 6542.7, "test", 36.8
    stages = {"sort": (12.6, 156, 9e8), "eval_grouped": (33.3, 156, 9e8)}
    r = bench.stage_rooflines(Ctx(), dict(bench.WORKLOADS["C4"]), stages, 3)["eval_grouped"]
    assert r["bound"] == "hbm" and r["algorithmic_bytes_per_eval"] == 24
    g = r["one_block_gather_per_eval"]
    assert g["bytes_per_eval"] == 24 + 8 * 36 and abs(g["hbm_frac"] / r["hbm_frac"] - 312 / 24) < 1e-9
    assert "bucket_pipe" in r["traffic_source"] and r["traffic"] > 0
    c1 = bench.stage_rooflines(Ctx(), dict(bench.WORKLOADS["C1"]), stages, 3)["eval_grouped"]
    assert "one_block_gather_per_eval" not in c1 and "bucket_pipe" not in (c1.get("traffic_source") or "")
