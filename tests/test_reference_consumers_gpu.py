"""The reference's own consumers of the C ABI, UNMODIFIED, run against libbaobzi.so (B200) -- the drop-in proof.

Built by `make -C oracle consumers` (called from __graft_entry__.build()) from the sources where they lie under
/root/reference, against include/baobzi.h and baobzi_b200/lib/libbaobzi.so of THIS repo; the binaries live under
oracle/_ref/consumers/ and travel to the GPU box (the reference tree itself does not exist there):

  baobzi_timing_c  examples/C/baobzi_timing.c:43-56,176-208 -- malloc'ed (pageable) buffers, srand(1) points; its
                   "Mevals/s" line of the 2D section IS BASELINE.json configs[0] through the reference's own harness
  test_c           tests/test_c.cpp:27-215 with the vendored Catch2 -- the reference's own C test-suite
  hpp_user         tests/consumers/hpp_user.cpp (ours) on the reference's include/baobzi.hpp:10-44
  pyref            the reference's Python binding src/python/__init__.py:77-138, imported as `baobzi`
"""
import os
import re
import shutil
import subprocess
import sys
import time

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CONS = os.path.join(ROOT, "oracle", "_ref", "consumers")
LIBDIR = os.path.join(ROOT, "baobzi_b200", "lib")


def _need(path):
    assert os.path.exists(path), "%s is missing: run __graft_entry__.build() where /root/reference exists" % path
    return path


def _env():
    env = dict(os.environ)
    env["LD_LIBRARY_PATH"] = LIBDIR + os.pathsep + env.get("LD_LIBRARY_PATH", "")
    return env


def _record(name, text):
    out = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(out):
        with open(os.path.join(out, name), "w") as f:
            f.write(text)


def test_baobzi_timing_c_unmodified(gpu_lib, tmp_path):
    """`baobzi_timing_c 10 8`: 1D, 1D2, 2D at order 8 (BASELINE config 1) and the 3D order-8 function, 1e6 points each,
    10 runs of baobzi_eval_multi + 1e6 scalar baobzi_eval calls per function for its error report."""
    exe = _need(os.path.join(CONS, "baobzi_timing_c"))
    t0 = time.time()
    r = subprocess.run([exe, "10", "8"], cwd=tmp_path, env=_env(), stdout=subprocess.PIPE, stderr=subprocess.STDOUT,
                       text=True, timeout=1500)
    dt = time.time() - t0
    _record("consumer_baobzi_timing.txt", r.stdout + "\nwall seconds: %.1f\n" % dt)
    assert r.returncode == 0, r.stdout[-2000:]
    rates = [float(v) for v in re.findall(r"Mevals/s: ([0-9.eE+-]+)", r.stdout)]
    assert len(rates) == 4 and all(v > 0 for v in rates), r.stdout
    errs = [float(v) for v in re.findall(r"rel error max, mean: ([0-9.eE+-]+)", r.stdout)]
    assert len(errs) == 4 and all(e < 1e-8 for e in errs), r.stdout       # fits meet their tolerance through baobzi_eval
    for d in (1, 2, 3):
        assert os.path.getsize(tmp_path / ("func_approx_%dd" % d)) > 1000   # baobzi_save wrote the trees
    assert "Baobzi function mapping 2 to 1" in r.stdout                      # baobzi_stats prints the reference's lines


def test_reference_test_c_suite(gpu_lib, tmp_path):
    exe = _need(os.path.join(CONS, "test_c"))
    r = subprocess.run([exe], cwd=tmp_path, env=_env(), stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True,
                       timeout=900)
    _record("consumer_test_c.txt", r.stdout)
    assert r.returncode == 0, r.stdout[-3000:]
    m = re.search(r"All tests passed \((\d+) assertions in (\d+) test cases\)", r.stdout)
    assert m and int(m.group(2)) == 4 and int(m.group(1)) >= 12, r.stdout[-500:]


def test_reference_cpp_wrapper(gpu_lib, tmp_path):
    exe = _need(os.path.join(CONS, "hpp_user"))
    r = subprocess.run([exe], cwd=tmp_path, env=_env(), stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True,
                       timeout=600)
    assert r.returncode == 0 and r.stdout.strip().endswith("OK 4097"), r.stdout[-1000:]


PY_USER = r"""
import sys, numpy as np
import baobzi                                   # the reference's src/python/__init__.py, unmodified
assert baobzi.baobzi_path.endswith('prefix/lib/libbaobzi.so'), baobzi.baobzi_path   # found by ITS OWN search
def f(x, y, data):                              # its INPUT_FUNC signature (ctypes pointers)
    y[0] = np.exp(np.cos(3.0 * x[0]) * np.sin(2.0 * x[1]))
c, hl = np.array([0.5, -0.25]), np.array([1.0, 0.5])
b = baobzi.Baobzi(f, 2, 8, c, hl, 1e-9)
rng = np.random.default_rng(11)
x = c + (rng.random((20000, 2)) * 2 - 1) * hl
y = b(x)
want = np.exp(np.cos(3.0 * x[:, 0]) * np.sin(2.0 * x[:, 1]))
assert y.shape == (20000,) and np.all(np.abs(y - want) <= 1e-8 * np.abs(want)), np.abs(y / want - 1).max()
b.save(sys.argv[1])
r = baobzi.Baobzi(filename=sys.argv[1])
assert (r.dim, r.order, r.output_dim) == (2, 8, 1)
assert np.array_equal(r(x).view(np.int64), y.view(np.int64))
b.stats()
print('PYREF OK')
"""


def test_reference_python_binding(gpu_lib, tmp_path):
    """`import baobzi` = the reference's ctypes binding; it finds libbaobzi.so through its own search (five directories
    above the module + lib/), i.e. a pip-style prefix whose lib/ holds OUR library."""
    src = _need(os.path.join(CONS, "pyref", "lib", "python3", "site-packages", "baobzi", "__init__.py"))
    site = tmp_path / "prefix" / "lib" / "python3" / "site-packages"
    os.makedirs(site / "baobzi")
    shutil.copy(src, site / "baobzi" / "__init__.py")
    os.symlink(os.path.join(LIBDIR, "libbaobzi.so"), tmp_path / "prefix" / "lib" / "libbaobzi.so")
    env = _env()
    env["PYTHONPATH"] = str(site)
    r = subprocess.run([sys.executable, "-c", PY_USER, str(tmp_path / "py.baobzi")], cwd=tmp_path, env=env,
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode == 0 and "PYREF OK" in r.stdout, r.stdout[-3000:]
    # and our own binding restores the file that went through the reference's binding, with the same values
    b = gpu_lib.Baobzi(filename=str(tmp_path / "py.baobzi"))
    assert (b.dim, b.order) == (2, 8)
    b.free()


def test_scalar_eval_latency_and_pageable_throughput(gpu_lib):
    """The two numbers a drop-in caller sees first (recorded for BASELINE.md / INTEGRATION.md): microseconds per
    baobzi_eval call, and baobzi_eval_multi throughput on plain numpy (pageable) buffers."""
    from conftest import load_golden
    path, _ = load_golden("t2d_o8")
    b = gpu_lib.Baobzi(filename=path)
    x1 = np.array([1.2, 3.3])
    for _ in range(200):
        b.eval(x1)
    t0 = time.perf_counter()
    n_calls = 5000
    import ctypes as C
    y = np.empty(1)
    xp, yp = x1.ctypes.data, y.ctypes.data
    for _ in range(n_calls):
        b._lib.baobzi_eval(b.ptr, xp, yp)
    us = (time.perf_counter() - t0) / n_calls * 1e6
    rng = np.random.default_rng(2)
    n = 20_000_000
    x = 0.5 + rng.random((n, 2)) * np.array([2.0, 2.0]) + np.array([0.0, 1.5])
    res = np.zeros(n)
    b(x, out=res)                       # first call of this size: the workers allocate their pinned / device buffers
    t0 = time.perf_counter()
    b(x, out=res)
    rate = n / (time.perf_counter() - t0)
    _record("consumer_latency.txt", "baobzi_eval: %.2f us/call (ctypes call overhead included)\n"
            "baobzi_eval_multi, pageable numpy buffers, 2D order 8, %d points: %.3e evals/s\n" % (us, n, rate))
    assert us < 200.0 and rate > 1e8
    b.free()
