"""baobzi_eval_multi_device_noalloc (include/baobzi.h; no counterpart in the reference):
the grouped evaluation with its scratch in a fixed device array of the library, for callers that may not
allocate.  Checked through the C ABI against the existing entry point baobzi_eval_multi_device, bit for bit (the
tolerance of every parity test here: 0 ulp, NaN where the point is outside), on calls in a row that share the array.

Collected last on purpose: a failure here hides nothing behind it.  Runs alone in a
fresh process too -- baobzi_restore / baobzi_init put the replica on the current device, which is all the set-up the
entry point asks for."""
import ctypes as C

import numpy as np
import pytest

from conftest import golden_cases, load_golden

pytestmark = pytest.mark.gpu

CUDA_ERROR_INVALID_VALUE = 1
CUDA_ERROR_INVALID_DEVICE_POINTER = 17


def _bits(t):
    import torch
    return t.view(torch.int64)


def _both(b, xd, stream=0):
    """(existing entry point, new entry point) on the same device points, both through the C ABI."""
    import torch
    n = xd.shape[0]
    want = torch.full((n,), 7.0, dtype=torch.float64, device=xd.device)
    got = torch.full((n,), -7.0, dtype=torch.float64, device=xd.device)
    assert b._lib.baobzi_eval_multi_device(b.ptr, xd.data_ptr(), want.data_ptr(), n, stream) == 0
    assert b._lib.baobzi_eval_multi_device_noalloc(b.ptr, xd.data_ptr(), got.data_ptr(), n, stream) == 0
    torch.cuda.synchronize()
    return want, got


def _points(ref, path, n, seed):
    import torch
    a = ref.OracleTree(path).arrays()
    lo, hi = a["lower_left"], a["upper_right"]
    rng = np.random.default_rng(seed)
    x = lo + rng.random((n, len(lo))) * (hi - lo) * 1.04 - 0.02 * (hi - lo)  # a few percent outside the domain
    return torch.from_numpy(np.ascontiguousarray(x)).cuda()


@pytest.mark.parametrize("name", golden_cases())
def test_noalloc_matches_existing_entry_on_golden_trees(gpu_lib, ref, name):
    """Every golden tree (1D / 2D / 3D, orders 6-16, two-output 1D, deep and anisotropic ones): the reference's own
    points, then two more calls in a row of different sizes on the same scratch array."""
    import torch
    path, g = load_golden(name)
    b = gpu_lib.Baobzi(filename=path)
    xd = torch.from_numpy(g["x"]).cuda()
    want, got = _both(b, xd)
    assert torch.equal(_bits(got), _bits(want))
    assert ref.ulp_diff(got.cpu().numpy(), g["y_avx2"]).max() == 0       # and so the reference's values
    for n, seed in ((50000, 1), (777, 2), (50000, 3)):                     # big, small (stale tail behind it), big
        xr = _points(ref, path, n, seed)
        want, got = _both(b, xr)
        assert torch.equal(_bits(got), _bits(want)), (name, n)
        assert bool(torch.isnan(got).any()) and not bool(torch.isnan(got).all())
    b.free()


@pytest.mark.parametrize("name", ["t3d_o8", "t2d_o8", "t1d_o16"])
def test_noalloc_at_the_point_limit_twice_in_a_row(gpu_lib, ref, name):
    """BAOBZI_NOALLOC_MAX_POINTS points per call, two calls back to back on one stream with no host synchronisation
    between them (the second overwrites the first's keys / order / counters), through the Python wrapper; and the
    calls take no device memory."""
    import torch
    path, _ = load_golden(name)
    b = gpu_lib.Baobzi(filename=path)
    n = gpu_lib.NOALLOC_MAX_POINTS
    x1, x2 = _points(ref, path, n, 11), _points(ref, path, n, 12)
    want1, want2 = b(x1), b(x2)
    got1, got2 = torch.empty_like(want1), torch.empty_like(want2)
    b.eval_noalloc(x1, got1)                       # first use: loads the kernels, resolves the array's address
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    got1.fill_(0.0)
    b.eval_noalloc(x1, got1)
    b.eval_noalloc(x2, got2)
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] >= free0   # nothing was allocated by the two calls
    assert torch.equal(_bits(got1), _bits(want1))
    assert torch.equal(_bits(got2), _bits(want2))
    b.free()


def test_noalloc_two_handles_and_a_side_stream_share_the_array(gpu_lib, ref):
    """The array belongs to the device, not to a handle: alternate two trees of different dimension on a non-default
    stream (ordered by the stream, so no race)."""
    import torch
    pa, _ = load_golden("t3d_o10")
    pb, _ = load_golden("aniso_2d")
    a, b = gpu_lib.Baobzi(filename=pa), gpu_lib.Baobzi(filename=pb)
    xa, xb = _points(ref, pa, 200000, 5), _points(ref, pb, 300000, 6)
    wa, wb = a(xa), b(xb)
    torch.cuda.synchronize()
    st = torch.cuda.Stream()
    outs = []
    with torch.cuda.stream(st):
        for _ in range(3):
            ga, gb = torch.empty_like(wa), torch.empty_like(wb)
            a.eval_noalloc(xa, ga, stream=st.cuda_stream)
            b.eval_noalloc(xb, gb, stream=st.cuda_stream)
            outs.append((ga, gb))
    st.synchronize()
    for ga, gb in outs:
        assert torch.equal(_bits(ga), _bits(wa)) and torch.equal(_bits(gb), _bits(wb))
    a.free()
    b.free()


def test_noalloc_rejects_what_does_not_fit(gpu_lib, ref, capfd):
    import torch
    path, _ = load_golden("t2d_o8")
    b = gpu_lib.Baobzi(filename=path)
    fn = b._lib.baobzi_eval_multi_device_noalloc
    n = gpu_lib.NOALLOC_MAX_POINTS + 1
    xd = _points(ref, path, n, 3)
    out = torch.full((n,), 3.25, dtype=torch.float64, device="cuda")
    capfd.readouterr()
    assert fn(b.ptr, xd.data_ptr(), out.data_ptr(), n, 0) == CUDA_ERROR_INVALID_VALUE
    torch.cuda.synchronize()
    assert "BAOBZI_NOALLOC_MAX_POINTS" in capfd.readouterr().err
    assert bool((out == 3.25).all())                                       # rejected before anything was enqueued
    assert fn(b.ptr, xd.data_ptr(), out.data_ptr(), 0, 0) == 0             # empty batch: nothing to do
    assert fn(b.ptr, xd.data_ptr(), out.data_ptr(), n - 1, 0) == 0         # exactly the limit
    want = b(xd[: n - 1])
    torch.cuda.synchronize()
    assert torch.equal(_bits(out[: n - 1]), _bits(want)) and float(out[n - 1]) == 3.25
    host = np.zeros((16, 2))
    res = np.zeros(16)
    assert fn(b.ptr, host.ctypes.data, res.ctypes.data, 16, 0) == CUDA_ERROR_INVALID_DEVICE_POINTER
    assert "device pointers" in capfd.readouterr().err
    b.free()


def test_noalloc_leaf_limit(gpu_lib, ref):
    """A tree of exactly BAOBZI_NOALLOC_MAX_LEAVES leaves is served; the 1 041 664-leaf tree of BASELINE.json configs[3]
    is refused (and the ordinary entry point still serves it)."""
    import torch
    from test_scale_gpu import BIG, make, _uniform
    b3 = make(gpu_lib, ref, BIG["B3"])
    assert b3.tree_info()["leaves"] == gpu_lib.NOALLOC_MAX_LEAVES
    for seed in (1, 2):
        x = _uniform(BIG["B3"], 400000 + seed, seed)
        want, got = _both(b3, x)
        assert torch.equal(_bits(got), _bits(want))
    b3.free()
    c4 = make(gpu_lib, ref, BIG["C4"])
    assert c4.tree_info()["leaves"] > gpu_lib.NOALLOC_MAX_LEAVES
    x = _uniform(BIG["C4"], 1000, 4)
    out = torch.zeros(1000, dtype=torch.float64, device="cuda")
    assert c4._lib.baobzi_eval_multi_device_noalloc(c4.ptr, x.data_ptr(), out.data_ptr(), 1000, 0) == CUDA_ERROR_INVALID_VALUE
    assert not bool(torch.isnan(c4(x)).any())
    c4.free()
