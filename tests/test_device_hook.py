"""SURVEY.md 8f-4, fused producer / consumer hooks: (1) include/baobzi_device.cuh compiles in a user's translation unit
with stock nvcc flags and, on a GPU, a kernel that generates its own points gets the bits of the batched entry;
(2) baobzi_eval_multi_device_dot = sum of the bit-exact values, in every execution mode, reproducibly."""
import os
import shutil
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "device_hook", "fused_user_kernel.cu")
LIBDIR = os.path.join(ROOT, "baobzi_b200", "lib")


def _nvcc():
    return shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"


def _compile(out, extra=(), include=None):
    cmd = [_nvcc(), "-gencode", "arch=compute_100a,code=sm_100a", "-O2", "-std=c++17", "-I",
           include or os.path.join(ROOT, "include"), *extra, "-o", out, SRC]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert r.returncode == 0, r.stdout
    return out


def test_device_header_compiles_in_user_code(tmp_path):
    """No GPU needed: the public device header is self-contained (object file only, nothing runs)."""
    _compile(str(tmp_path / "fused_user_kernel.o"), extra=("-c",))


def test_public_headers_are_installable_on_their_own(tmp_path):
    """include/ copied somewhere else (as `make install` would) still compiles a user kernel: nothing in it reaches
    back into baobzi_b200/csrc."""
    inc = tmp_path / "prefix" / "include"
    shutil.copytree(os.path.join(ROOT, "include"), inc)
    _compile(str(tmp_path / "fused_user_kernel.o"), extra=("-c",), include=str(inc))
    for dirpath, _, files in os.walk(inc):
        for f in files:
            text = open(os.path.join(dirpath, f)).read()
            assert not [ln for ln in text.splitlines() if ln.lstrip().startswith("#include") and "csrc" in ln], f


@pytest.mark.gpu
@pytest.mark.parametrize("name,dim,order", [("t1d_o16", 1, 16), ("t2d_o8", 2, 8), ("t3d_o6", 3, 6)])
def test_user_kernel_matches_batched_entry(gpu_lib, tmp_path, name, dim, order):
    from conftest import load_golden
    path, _ = load_golden(name)
    exe = _compile(str(tmp_path / "fused_user_kernel"),
                   extra=("-L", LIBDIR, "-lbaobzi", "-Xlinker", "-rpath=" + LIBDIR))
    r = subprocess.run([exe, path, str(dim), str(order), "200003"], stdout=subprocess.PIPE, stderr=subprocess.STDOUT,
                       text=True, timeout=300)
    assert r.returncode == 0 and r.stdout.startswith("OK"), r.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("name,mode", [("t1d_o16", "direct"), ("t2d_o8", "direct"), ("t2d_o8", "binned"),
                                       ("osc_2d", "bucketed"), ("t3d_o8", "direct"), ("t3d_o8", "binned"),
                                       ("t3d_o12", "auto")])
def test_dot_is_the_sum_of_the_exact_values(gpu_lib, ref, name, mode):
    import torch
    from conftest import load_golden
    path, _ = load_golden(name)
    b = gpu_lib.Baobzi(filename=path, mode=mode)
    t = ref.OracleTree(path)
    a = t.arrays()
    rng = np.random.default_rng(17)
    n = 300_017
    x = a["lower_left"] + rng.random((n, b.dim)) * (a["upper_right"] - a["lower_left"])
    w = rng.standard_normal(n)
    y = t.eval_multi(x, ref.ARITH_AVX2)
    assert not np.isnan(y).any()
    xd, wd = torch.from_numpy(x).cuda(), torch.from_numpy(w).cuda()
    for chunk in (0, 65_537):
        b.set_mode(mode, chunk)
        s1, s2 = float(b.dot(xd)), float(b.dot(xd))
        assert s1 == s2                                            # fixed summation order
        assert abs(s1 - np.sum(y)) <= 1e-12 * np.sum(np.abs(y))
        d1 = float(b.dot(xd, wd))
        assert d1 == float(b.dot(xd, wd))
        assert abs(d1 - np.dot(w, y)) <= 1e-12 * np.sum(np.abs(w * y))
    # the empty sum, one point, and NaN from a point outside the domain
    assert float(b.dot(xd[:0])) == 0.0
    assert float(b.dot(xd[:1])) == y[0]
    x_out = x[:1000].copy()
    x_out[500] = a["upper_right"] + 1.0
    assert np.isnan(float(b.dot(torch.from_numpy(x_out).cuda())))
    with pytest.raises(TypeError):
        b.dot(xd, wd[:-1])
    b.free()


@pytest.mark.gpu
def test_device_view_bytes(gpu_lib):
    from conftest import load_golden
    path, _ = load_golden("t2d_o8")
    b = gpu_lib.Baobzi(filename=path)
    raw = b.device_view(0)
    assert len(raw) == b.VIEW_BYTES
    ll = np.frombuffer(raw[:24], dtype=np.float64)
    assert np.isfinite(ll).all()
    dim, order = np.frombuffer(raw[84:92], dtype=np.int32)
    assert (dim, order) == (2, 8)
    b.free()
