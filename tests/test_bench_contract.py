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
