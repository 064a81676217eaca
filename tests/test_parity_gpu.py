"""GPU suite: the CUDA path, called through the C ABI (ctypes -> libbaobzi.so), against the oracle and the
golden vectors.  Bar: node indices exact; values BIT-EXACT against the reference's AVX2+FMA build (tolerance
0 ulp, tighter than the 4 ulp north_star allows), NaN exactly where the reference gives NaN."""
import ctypes as C
import os

import numpy as np
import pytest

from conftest import golden_cases, load_golden

pytestmark = pytest.mark.gpu


def ulp(ref, a, b):
    return ref.ulp_diff(a, b).max() if len(a) else 0


@pytest.mark.parametrize("name", golden_cases())
def test_golden_host_pointers(gpu_lib, ref, name):
    """baobzi_restore + baobzi_eval_multi (host pointers) + node indices + baobzi_eval vs the reference's outputs."""
    path, g = load_golden(name)
    b = gpu_lib.Baobzi(filename=path)
    x = g["x"]
    assert np.array_equal(b.node_index(x), g["node_idx"])
    y = b(x)
    assert ulp(ref, y, g["y_avx2"]) == 0
    for row, ys in zip(g["scalar_rows"], g["ys_avx2"]):
        assert ulp(ref, b.eval(x[row]), ys) == 0
    assert b.launches > 0
    b.free()


@pytest.mark.parametrize("name", golden_cases())
@pytest.mark.parametrize("mode", ["direct", "binned", "bucketed"])
def test_golden_device_pointers(gpu_lib, ref, name, mode):
    import torch
    path, g = load_golden(name)
    b = gpu_lib.Baobzi(filename=path, mode=mode)
    xd = torch.from_numpy(g["x"]).cuda()
    yd = b(xd)
    torch.cuda.synchronize()
    assert ulp(ref, yd.cpu().numpy(), g["y_avx2"]) == 0
    assert np.array_equal(b.node_index(xd).cpu().numpy(), g["node_idx"])
    # the classic entry point also accepts device pointers (zero copy)
    y2 = torch.empty_like(yd)
    b._lib.baobzi_eval_multi(b.ptr, xd.data_ptr(), y2.data_ptr(), xd.shape[0])
    assert torch.equal(torch.nan_to_num(y2, nan=-1.0), torch.nan_to_num(yd, nan=-1.0))
    b.free()


@pytest.mark.parametrize("name", ["tc_1d1", "t2d_o8", "t3d_o8", "t3d_o12"])
@pytest.mark.parametrize("n", [0, 1, 31, 33, 1000, 100003])
def test_ragged_sizes_and_modes(gpu_lib, ref, name, n):
    """Empty and ragged batches; direct and binned kernels and any chunking give identical bits."""
    import torch
    path, g = load_golden(name)
    t = ref.OracleTree(path)
    a = t.arrays()
    rng = np.random.default_rng(n + 1)
    x = a["lower_left"] + rng.random((n, t.dim)) * (a["upper_right"] - a["lower_left"]) * 1.02 - 0.01  # some outside
    want = t.eval_multi(x, ref.ARITH_AVX2) if n else np.empty(0)
    b = gpu_lib.Baobzi(filename=path)
    assert ulp(ref, b(x), want) == 0
    if n:
        xd = torch.from_numpy(x).cuda()
        for mode, chunk in (("direct", 0), ("binned", 0), ("binned", 1000), ("binned", 4097), ("bucketed", 0), ("bucketed", 999)):
            b.set_mode(mode, chunk)
            yd = b(xd)
            torch.cuda.synchronize()
            assert ulp(ref, yd.cpu().numpy(), want) == 0, (mode, chunk)
    b.free()


def test_nan_and_inf_inputs(gpu_lib, ref):
    """NaN coordinates pass the reference's bounds test (all comparisons false) and poison the value; infinities
    are out of domain."""
    path, _ = load_golden("t2d_o8")
    b = gpu_lib.Baobzi(filename=path)
    x = np.array([[np.nan, 3.0], [1.5, np.nan], [np.inf, 3.0], [-np.inf, 3.0], [1.5, 3.0]])
    y = b(x)
    assert np.all(np.isnan(y[:4])) and not np.isnan(y[4])
    b.free()


def test_save_is_byte_identical_and_restorable(gpu_lib, ref, tmp_path):
    """baobzi_save of a restored handle reproduces the reference-written file byte for byte."""
    for name in ("tc_1d2", "t2d_o10", "t3d_o6"):
        path, g = load_golden(name)
        b = gpu_lib.Baobzi(filename=path)
        out = str(tmp_path / (name + ".out"))
        b.save(out)
        assert open(out, "rb").read() == open(path, "rb").read()
        b2 = gpu_lib.Baobzi(filename=out)
        assert ulp(ref, b2(g["x"]), g["y_avx2"]) == 0
        b.free()
        b2.free()


FIT_CASES = [  # reference tests/test_c.cpp: 27-71, 73-121, 123-168, 170-215 (3D tolerance relaxed to keep it quick)
    ("testfun_1d", 1, 8, 1e-10, [3.0], [1.0], 1.5, 1),
    ("testfun_1d2", 1, 8, 1e-10, [3.0], [1.0], 1.5, 2),
    ("testfun_2d", 2, 6, 1e-10, [-10.0, 3.0], [np.pi / 5, np.pi / 5], 1.5, 1),
    ("testfun_3d1", 3, 12, 1e-8, [3.0, -2.0, 1.0], [np.pi / 5, np.pi / 5, np.pi / 10], 1.5, 1),
]


@pytest.mark.parametrize("func,dim,order,tol,center,hl,data,od", FIT_CASES)
def test_reference_test_c_cases_through_init(gpu_lib, ref, tmp_path, func, dim, order, tol, center, hl, data, od):
    """The reference's own test cases, run against our library: baobzi_init, accuracy at the lower-left corner
    and the centre via baobzi_eval, then save -> restore -> bitwise equal (test_c.cpp:42-68 etc.)."""
    c, h = np.array(center), np.array(hl)
    b = gpu_lib.Baobzi(fin=ref.func_ptr(func), dim=dim, order=order, center=c, half_length=h, tol=tol, data=data,
                       output_dim=od)
    f = getattr(ref.funcs_lib(), func)
    f.restype, f.argtypes = None, [ref.c_double_p, ref.c_double_p, C.c_void_p]
    dat = C.c_double(data)
    for p in (c - h, c):
        p = np.ascontiguousarray(p)
        ye = np.zeros(od)
        f(p.ctypes.data_as(ref.c_double_p), ye.ctypes.data_as(ref.c_double_p), C.addressof(dat))
        assert np.all(np.abs((b.eval(p) - ye) / ye) < tol)
    out = str(tmp_path / "t.baobzi")
    b.save(out)
    b2 = gpu_lib.Baobzi(filename=out)
    assert np.array_equal(b.eval(c), b2.eval(c))
    # and the batched path equals the oracle on the same saved tree
    rng = np.random.default_rng(11)
    x = c + (rng.random((20000, dim)) * 2 - 1) * h
    assert ulp(ref, b2(x), ref.OracleTree(out).eval_multi(x, ref.ARITH_AVX2)) == 0
    b.stats()
    b.free()
    b2.free()


def test_python_callback_init(gpu_lib):
    b = gpu_lib.Baobzi(fin=lambda x: np.sin(x[0]) + 2.0, dim=1, order=10, center=np.array([1.0]),
                       half_length=np.array([2.0]), tol=1e-11)
    x = np.linspace(-0.99, 2.99, 1001)
    assert np.max(np.abs(b(x) - (np.sin(x) + 2))) < 1e-10
    b.free()


def test_concurrent_calls_on_one_handle(gpu_lib, ref):
    """eval is const in the reference, so concurrent calls on one handle are legal; they must stay so."""
    import threading
    path, g = load_golden("t2d_o12")
    b = gpu_lib.Baobzi(filename=path)
    rng = np.random.default_rng(0)
    xs = [np.ascontiguousarray(g["x"][rng.permutation(len(g["x"]))]) for _ in range(8)]
    t = ref.OracleTree(path)
    outs = [None] * 8

    def work(i):
        for _ in range(20):
            outs[i] = b(xs[i])

    th = [threading.Thread(target=work, args=(i,)) for i in range(8)]
    [t_.start() for t_ in th]
    [t_.join() for t_ in th]
    for i in range(8):
        assert ulp(ref, outs[i], t.eval_multi(xs[i], ref.ARITH_AVX2)) == 0
    b.free()


@pytest.mark.parametrize("name", ["t2d_o8", "osc_2d", "t2d_o12"])
def test_2d_points_that_are_only_8_byte_aligned(gpu_lib, ref, name):
    """The reference API promises nothing beyond `const double *`: a 2D point array may start on an odd double (a
    view one element into a buffer).  Every device entry must take it (no misaligned-address fault) and give the
    same bits -- whatever mode was asked for."""
    import torch
    path, _ = load_golden(name)
    t = ref.OracleTree(path)
    a = t.arrays()
    rng = np.random.default_rng(5)
    n = 1_200_001   # above the auto-mode thresholds of the grouped 2D paths
    x = a["lower_left"] + rng.random((n, 2)) * (a["upper_right"] - a["lower_left"])
    want = t.eval_multi(x[:50000], ref.ARITH_AVX2)
    flat = torch.empty(2 * n + 1, dtype=torch.float64, device="cuda")
    flat[1:] = torch.from_numpy(x).cuda().reshape(-1)
    xv = flat[1:].view(n, 2)
    assert xv.data_ptr() % 16 == 8 and xv.is_contiguous()
    b = gpu_lib.Baobzi(filename=path)
    aligned = b(torch.from_numpy(x).cuda()).clone()
    for mode in ("auto", "direct", "binned", "bucketed"):
        b.set_mode(mode)
        y = b(xv)
        torch.cuda.synchronize()
        assert torch.equal(y.view(torch.int64), aligned.view(torch.int64)), mode
    assert ulp(ref, aligned[:50000].cpu().numpy(), want) == 0
    assert np.array_equal(b.node_index(xv[:50000]).cpu().numpy(), t.node_indices(x[:50000]))
    w = torch.ones(n, dtype=torch.float64, device="cuda")
    b.set_mode("direct")
    assert torch.equal(b.dot(xv, w), b.dot(torch.from_numpy(x).cuda(), w))
    # the classic entry with device pointers
    y2 = torch.empty(n, dtype=torch.float64, device="cuda")
    b._lib.baobzi_eval_multi(b.ptr, xv.data_ptr(), y2.data_ptr(), n)
    assert torch.equal(y2.view(torch.int64), aligned.view(torch.int64))
    b.free()


class _CudaArray:
    """A CUDA buffer that is not a torch tensor: only __cuda_array_interface__ (what cupy / numba arrays expose)."""

    def __init__(self, tensor):
        self._keep = tensor
        self.__cuda_array_interface__ = {"shape": tuple(tensor.shape), "typestr": "<f%d" % tensor.element_size(),
                                         "data": (tensor.data_ptr(), False), "version": 3, "strides": None}


def test_cuda_array_interface_inputs(gpu_lib, ref):
    """Binding-level device entry for non-torch CUDA arrays."""
    import torch
    path, g = load_golden("t3d_o8")
    b = gpu_lib.Baobzi(filename=path)
    xd = torch.from_numpy(g["x"]).cuda()
    out = torch.full((xd.shape[0],), -7.0, dtype=torch.float64, device="cuda")
    ret = b(_CudaArray(xd), out=_CudaArray(out))
    torch.cuda.synchronize()
    assert isinstance(ret, _CudaArray)
    assert ulp(ref, out.cpu().numpy(), g["y_avx2"]) == 0
    with pytest.raises(TypeError):
        b(_CudaArray(xd))                                     # no torch tensor to size the output from
    with pytest.raises(TypeError):
        b(_CudaArray(xd), out=_CudaArray(out[:5]))            # too small
    with pytest.raises(TypeError):
        b(_CudaArray(xd.float()))                             # wrong dtype is refused, not reinterpreted
    b.free()


def test_host_out_argument_is_validated(gpu_lib):
    path, g = load_golden("t2d_o8")
    b = gpu_lib.Baobzi(filename=path)
    x = g["x"]
    for bad in (np.empty(len(x), dtype=np.float32), np.empty(len(x) - 1), np.empty(2 * len(x))[::2]):
        with pytest.raises(TypeError):
            b(x, out=bad)
    out = np.empty(len(x) + 3)
    assert b(x, out=out) is out
    b.free()


def test_python_callback_exception_propagates(gpu_lib):
    calls = []

    def f(x):
        calls.append(1)
        if len(calls) > 5:
            raise ZeroDivisionError("boom")
        return 1.0 + x[0]
    with pytest.raises(ZeroDivisionError):
        gpu_lib.Baobzi(fin=f, dim=1, order=8, center=np.array([0.0]), half_length=np.array([1.0]), tol=1e-10)
