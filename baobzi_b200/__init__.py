"""baobzi_b200 -- host-side mirror of the reference's Python binding (src/python/__init__.py) on top of the
B200-native libbaobzi.so built from baobzi_b200/csrc.

`Baobzi` keeps the reference class's constructor arguments, `__call__`, `save` and `stats`
(reference src/python/__init__.py:77-138) and adds the device entry points of include/baobzi.h:
torch CUDA tensors (or anything with ``__cuda_array_interface__``) are evaluated in place, zero copy.

There is no CPU evaluation path: importing works anywhere (so the ABI can be inspected), but creating a
`Baobzi` needs the CUDA library and a GPU and fails loudly otherwise.
"""
import ctypes as C
import os

import numpy as np

__all__ = ["Baobzi", "load_library", "library_path", "build", "BaobziInput", "BaobziStruct"]

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "lib", "libbaobzi.so")
_lib = None

c_double_p = C.POINTER(C.c_double)
INPUT_FUNC = C.CFUNCTYPE(None, c_double_p, c_double_p, C.c_void_p)


class BaobziInput(C.Structure):
    """baobzi_input_t, include/baobzi.h (reference baobzi.h:21-38; the reference's own ctypes mirror at
    src/python/__init__.py:40-44 is stale -- it lacks tol_type)."""
    _fields_ = [("func", C.c_void_p), ("data", C.c_void_p), ("dim", C.c_int), ("output_dim", C.c_int),
                ("order", C.c_int), ("tol", C.c_double), ("minimum_leaf_fraction", C.c_double),
                ("split_multi_eval", C.c_int), ("min_depth", C.c_int), ("max_depth", C.c_int),
                ("tol_type", C.c_int)]


class BaobziStruct(C.Structure):
    """baobzi_struct, include/baobzi.h (reference baobzi.h:47-57)."""
    _fields_ = [("obj", C.c_void_p), ("dim", C.c_int), ("output_dim", C.c_int), ("order", C.c_int),
                ("eval", C.c_void_p), ("eval_multi", C.c_void_p), ("save", C.c_void_p),
                ("stats", C.c_void_p), ("free", C.c_void_p)]


baobzi_t = C.POINTER(BaobziStruct)

# every symbol include/baobzi.h declares: name -> (restype, argtypes)
ABI = {
    "baobzi_init": (baobzi_t, [C.POINTER(BaobziInput), c_double_p, c_double_p]),
    "baobzi_restore": (baobzi_t, [C.c_char_p]),
    "baobzi_save": (None, [baobzi_t, C.c_char_p]),
    "baobzi_eval": (None, [baobzi_t, C.c_void_p, C.c_void_p]),
    "baobzi_eval_multi": (None, [baobzi_t, C.c_void_p, C.c_void_p, C.c_int]),
    "baobzi_stats": (None, [baobzi_t]),
    "baobzi_free": (baobzi_t, [baobzi_t]),
    "baobzi_eval_multi_device": (C.c_int, [baobzi_t, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "baobzi_eval_multi_device_noalloc": (C.c_int, [baobzi_t, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "baobzi_node_index_multi_device": (C.c_int, [baobzi_t, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "baobzi_node_index_multi": (C.c_int, [baobzi_t, C.c_void_p, C.c_void_p, C.c_int64]),
    "baobzi_eval_multi_sharded": (C.c_int, [baobzi_t, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_void_p),
                                            C.POINTER(C.c_void_p), C.POINTER(C.c_int64)]),
    "baobzi_sync": (C.c_int, [baobzi_t]),
    "baobzi_set_mode": (C.c_int, [baobzi_t, C.c_int, C.c_int64]),
    "baobzi_eval_multi_device_all": (C.c_int, [baobzi_t, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "baobzi_eval_multi_device_dot": (C.c_int, [baobzi_t, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]),
    "baobzi_device_view": (C.c_int, [baobzi_t, C.c_int, C.c_void_p, C.c_size_t]),
    "baobzi_tree_info": (C.c_int, [baobzi_t, C.POINTER(C.c_int64)]),
    "baobzi_set_host_devices": (C.c_int, [baobzi_t, C.c_int]),
    "baobzi_b200_guard_check": (C.c_int, [C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "baobzi_launch_count": (C.c_int64, [baobzi_t]),
    "baobzi_profile_begin": (C.c_int, [baobzi_t]),
    "baobzi_profile_end": (C.c_int, [baobzi_t, c_double_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "baobzi_b200_fp64_peak": (C.c_int, [c_double_p]),
    "baobzi_b200_fp64_peak_sustained": (C.c_int, [c_double_p, C.c_double]),
    "baobzi_b200_version": (C.c_char_p, []),
}
DATA_SYMBOLS = ["baobzi_input_default"]
NOALLOC_MAX_POINTS, NOALLOC_MAX_LEAVES = 1 << 20, 1 << 16  # include/baobzi.h
MODES = {"auto": 0, "direct": 1, "binned": 2, "bucketed": 3}


def library_path():
    return _LIB_PATH


def build(verbose=False):
    """Compile baobzi_b200/csrc into baobzi_b200/lib/libbaobzi.so with nvcc for sm_100a (no GPU needed)."""
    import subprocess
    cmd = ["make", "-C", os.path.join(_HERE, "csrc"), "-j", str(min(16, os.cpu_count() or 4))]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if verbose or r.returncode != 0:
        print(r.stdout)
    if r.returncode != 0:
        raise RuntimeError("building libbaobzi.so failed")
    return _LIB_PATH


def load_library():
    """dlopen the CUDA library and declare its ABI.  Raises if it has not been built -- nothing falls back."""
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            raise OSError("baobzi_b200: %s is missing; run baobzi_b200.build() (needs nvcc). "
                          "There is no CPU fallback." % _LIB_PATH)
        lib = C.CDLL(_LIB_PATH)
        for name, (res, args) in ABI.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        _lib = lib
    return _lib


def _device_pointer(a):
    """(pointer, element count, keepalive) for a CUDA array: torch tensor or __cuda_array_interface__."""
    if hasattr(a, "data_ptr") and hasattr(a, "is_cuda"):
        if not a.is_cuda:
            return None
        import torch
        if a.dtype != torch.float64 or not a.is_contiguous():
            raise TypeError("CUDA input must be a contiguous float64 tensor")
        return a.data_ptr(), a.numel(), a
    cai = getattr(a, "__cuda_array_interface__", None)
    if cai is not None:
        if cai["typestr"] not in ("<f8", "=f8") or cai.get("strides") is not None:
            raise TypeError("CUDA input must be contiguous float64")
        n = 1
        for s in cai["shape"]:
            n *= s
        return cai["data"][0], n, a
    return None


class Baobzi:
    """Same surface as the reference's ``Baobzi`` (src/python/__init__.py:77-138).

    ``fin`` is a Python callable f(x) -> float or sequence (x is a numpy view of `dim` doubles) or the address
    of a C function with the baobzi_input_func_t signature.
    """

    def __init__(self, fin=None, dim=None, order=None, center=None, half_length=None, tol=None, output_dim=1,
                 minimum_leaf_fraction=0.0, split_multi_eval=1, min_depth=0, max_depth=50, filename=None,
                 data=None, tol_type=0, mode=None):
        self.ptr = None
        lib = self._lib = load_library()
        if filename:
            self.ptr = lib.baobzi_restore(os.fsencode(filename))
        elif fin is not None:
            if not (dim and order and center is not None and half_length is not None and tol):
                raise ValueError("Baobzi: supply dim, order, center, half_length, and tol for init")
            inp = BaobziInput()
            self._cb_error = None
            if callable(fin):
                def trampoline(xp, yp, _data, _f=fin, _d=dim, _o=output_dim):
                    # ctypes swallows exceptions raised inside a callback: keep the first one, feed the fitter NaN
                    # (never uninitialised memory) and re-raise once baobzi_init has returned
                    try:
                        if self._cb_error is not None:
                            raise self._cb_error
                        y = np.atleast_1d(np.asarray(_f(np.ctypeslib.as_array(xp, (_d,))), dtype=np.float64))
                        for k in range(_o):
                            yp[k] = y[k]
                    except BaseException as exc:  # noqa: BLE001
                        if self._cb_error is None:
                            self._cb_error = exc
                        for k in range(_o):
                            yp[k] = float("nan")
                self._cb = INPUT_FUNC(trampoline)
                inp.func = C.cast(self._cb, C.c_void_p)
            else:
                inp.func = int(fin)
            self._data = None
            if data is not None:
                self._data = C.c_double(float(data))
                inp.data = C.cast(C.pointer(self._data), C.c_void_p)
            inp.dim, inp.output_dim, inp.order, inp.tol = dim, output_dim, order, tol
            inp.minimum_leaf_fraction, inp.split_multi_eval = minimum_leaf_fraction, split_multi_eval
            inp.min_depth, inp.max_depth, inp.tol_type = min_depth, max_depth, tol_type
            c = np.ascontiguousarray(center, dtype=np.float64)
            h = np.ascontiguousarray(half_length, dtype=np.float64)
            self.ptr = lib.baobzi_init(C.byref(inp), c.ctypes.data_as(c_double_p), h.ctypes.data_as(c_double_p))
            if self._cb_error is not None:
                err, self._cb_error = self._cb_error, None
                if self.ptr:
                    lib.baobzi_free(self.ptr)
                    self.ptr = None
                raise err
        else:
            raise ValueError("Baobzi requires either a 'filename' argument or a 'fin' argument")
        if not self.ptr or not self.ptr[0].obj:
            self.ptr = None
            raise RuntimeError("Unable to create baobzi object")
        self.dim, self.output_dim, self.order = self.ptr[0].dim, self.ptr[0].output_dim, self.ptr[0].order
        if mode is not None:
            self.set_mode(mode)

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass

    def free(self):
        if getattr(self, "ptr", None):
            self._lib.baobzi_free(self.ptr)
            self.ptr = None

    # ---- evaluation -------------------------------------------------------------------------------
    def __call__(self, x, out=None, stream=None):
        """Evaluate at points x ([n, dim] or flat).  numpy in -> numpy out through baobzi_eval_multi (host
        pointers); CUDA tensor in -> CUDA tensor out through baobzi_eval_multi_device (zero copy)."""
        dp = _device_pointer(x)
        if dp is not None:
            ptr, count, _keep = dp
            n = count // self.dim
            if out is None:
                if not hasattr(x, "is_cuda"):
                    raise TypeError("pass out= (a CUDA float64 array of n values) with a non-torch CUDA input")
                import torch
                out = torch.empty(n, dtype=torch.float64, device=x.device)
            op = _device_pointer(out)
            if op is None or op[1] < n:
                raise TypeError("out must be a contiguous CUDA float64 array with at least one value per point")
            if hasattr(x, "is_cuda") and hasattr(out, "is_cuda") and out.device != x.device:
                raise TypeError("out must live on the device that holds x")
            optr = op[0]
            if stream is None and hasattr(x, "is_cuda"):
                import torch
                stream = torch.cuda.current_stream(x.device).cuda_stream
            rc = self._lib.baobzi_eval_multi_device(self.ptr, ptr, optr, n, stream or 0)
            if rc != 0:
                raise RuntimeError("baobzi_eval_multi_device failed with CUDA error %d" % rc)
            return out
        xarr = np.ascontiguousarray(x, dtype=np.float64)
        n = xarr.size // self.dim
        if n > 2**31 - 1:
            raise ValueError("baobzi_eval_multi takes an int count; use the device entry for more points")
        res = out if out is not None else np.empty(n, dtype=np.float64)
        if not (isinstance(res, np.ndarray) and res.dtype == np.float64 and res.flags.c_contiguous
                and res.flags.writeable and res.size >= n):
            raise TypeError("out must be a writeable C-contiguous float64 numpy array with at least one value per point")
        self._lib.baobzi_eval_multi(self.ptr, xarr.ctypes.data, res.ctypes.data, n)
        return res

    def eval_noalloc(self, x, out, stream=None):
        """baobzi_eval_multi_device_noalloc: x and out are the caller's CUDA float64 arrays; neither this method nor the
        library call allocates (fixed scratch array in the library; at most NOALLOC_MAX_POINTS points per call, calls
        that overlap on the GPU race -- see include/baobzi.h).  No counterpart in the reference."""
        dp, op = _device_pointer(x), _device_pointer(out)
        if dp is None or op is None:
            raise TypeError("eval_noalloc takes contiguous CUDA float64 arrays for x and out")
        n = dp[1] // self.dim
        if op[1] < n:
            raise TypeError("out must hold at least one value per point")
        if stream is None and hasattr(x, "is_cuda"):
            import torch
            stream = torch.cuda.current_stream(x.device).cuda_stream
        rc = self._lib.baobzi_eval_multi_device_noalloc(self.ptr, dp[0], op[0], n, stream or 0)
        if rc != 0:
            raise RuntimeError("baobzi_eval_multi_device_noalloc failed with CUDA error %d" % rc)
        return out

    def eval(self, x):
        """One point through baobzi_eval: returns output_dim values."""
        xarr = np.ascontiguousarray(x, dtype=np.float64)
        y = np.empty(max(1, self.output_dim), dtype=np.float64)
        self._lib.baobzi_eval(self.ptr, xarr.ctypes.data, y.ctypes.data)
        return y

    def eval_all(self, x):
        """Every output component for a batch of device-resident points of a 1D tree: (n, output_dim) tensor whose
        row i is what eval(x[i]) returns (baobzi_eval_multi_device_all)."""
        import torch
        dp = _device_pointer(x)
        if dp is None:
            raise TypeError("eval_all takes a CUDA float64 tensor / array")
        ptr, count, _keep = dp
        n = count // self.dim
        out = torch.empty((n, max(1, self.output_dim)), dtype=torch.float64, device=x.device)
        rc = self._lib.baobzi_eval_multi_device_all(self.ptr, ptr, out.data_ptr(), n,
                                                    torch.cuda.current_stream(x.device).cuda_stream)
        if rc != 0:
            raise RuntimeError("baobzi_eval_multi_device_all failed with CUDA error %d" % rc)
        return out

    def dot(self, x, w=None):
        """sum_i w[i] * f(x[i]) over device-resident points (plain sum without w), evaluated and reduced on the GPU
        without an n-sized result array (baobzi_eval_multi_device_dot).  Returns a 0-dim CUDA tensor."""
        import torch
        dp = _device_pointer(x)
        if dp is None:
            raise TypeError("dot takes a CUDA float64 tensor / array")
        ptr, count, _keep = dp
        n = count // self.dim
        wptr = None
        if w is not None:
            wp = _device_pointer(w)
            if wp is None or wp[1] != n:
                raise TypeError("w must be a CUDA float64 tensor / array with one weight per point")
            wptr = wp[0]
        out = torch.empty((), dtype=torch.float64, device=x.device)
        rc = self._lib.baobzi_eval_multi_device_dot(self.ptr, ptr, wptr, n, out.data_ptr(),
                                                    torch.cuda.current_stream(x.device).cuda_stream)
        if rc != 0:
            raise RuntimeError("baobzi_eval_multi_device_dot failed with CUDA error %d" % rc)
        return out

    VIEW_BYTES = 160  # sizeof(baobzi_device::View), include/baobzi_device.cuh

    def device_view(self, device=0):
        """The bytes of the baobzi_device::View of the replica on `device`, to pass by value to a CUDA kernel that
        includes include/baobzi_device.cuh (baobzi_device_view)."""
        buf = C.create_string_buffer(self.VIEW_BYTES)
        rc = self._lib.baobzi_device_view(self.ptr, int(device), buf, self.VIEW_BYTES)
        if rc != 0:
            raise RuntimeError("baobzi_device_view failed with error %d" % rc)
        return buf.raw

    def node_index(self, x):
        """Global node index per point (the reference's Function::get_global_node_index), -1 outside."""
        dp = _device_pointer(x)
        if dp is not None:
            import torch
            ptr, count, _keep = dp
            n = count // self.dim
            out = torch.empty(n, dtype=torch.int64, device=x.device)
            rc = self._lib.baobzi_node_index_multi_device(self.ptr, ptr, out.data_ptr(), n,
                                                          torch.cuda.current_stream(x.device).cuda_stream)
            if rc != 0:
                raise RuntimeError("baobzi_node_index_multi_device failed with CUDA error %d" % rc)
            return out
        xarr = np.ascontiguousarray(x, dtype=np.float64)
        n = xarr.size // self.dim
        idx = np.empty(n, dtype=np.int64)
        rc = self._lib.baobzi_node_index_multi(self.ptr, xarr.ctypes.data, idx.ctypes.data, n)
        if rc != 0:
            raise RuntimeError("baobzi_node_index_multi failed with CUDA error %d" % rc)
        return idx

    # ---- the rest of the reference surface ----------------------------------------------------------
    def save(self, filename):
        self._lib.baobzi_save(self.ptr, os.fsencode(filename))

    def stats(self):
        self._lib.baobzi_stats(self.ptr)

    # ---- extension knobs ----------------------------------------------------------------------------
    def set_mode(self, mode, chunk=0):
        m = MODES[mode] if isinstance(mode, str) else int(mode)
        if self._lib.baobzi_set_mode(self.ptr, m, int(chunk)) != 0:
            raise ValueError("bad mode")

    def set_host_devices(self, n_devices):
        """GPUs a numpy (host-pointer) call is spread over: 1 = current device, 0 = all visible."""
        if self._lib.baobzi_set_host_devices(self.ptr, int(n_devices)) != 0:
            raise ValueError("bad device count")

    # "sort": fused lookup + scan + scatter (binned / bucketed paths); "scan", "scatter": only the unfused fallback
    STAGES = ("sort", "scan", "scatter", "eval_grouped", "eval_direct")

    def profile_begin(self):
        self._lib.baobzi_profile_begin(self.ptr)

    def profile_end(self):
        """{stage: (total ms, launches, points)} for the kernels launched since profile_begin (synchronises)."""
        ms = (C.c_double * 5)()
        la = (C.c_int64 * 5)()
        pts = (C.c_int64 * 5)()
        rc = self._lib.baobzi_profile_end(self.ptr, ms, la, pts)
        if rc != 0:
            raise RuntimeError("baobzi_profile_end failed with CUDA error %d" % rc)
        return {s: (ms[i], la[i], pts[i]) for i, s in enumerate(self.STAGES)}

    def tree_info(self):
        """{nodes, leaves, grid_cells, coeff_bytes, device_bytes} of the fitted tree (baobzi_tree_info)."""
        v = (C.c_int64 * 5)()
        self._lib.baobzi_tree_info(self.ptr, v)
        return {"nodes": v[0], "leaves": v[1], "grid_cells": v[2], "coeff_bytes": 8 * v[3], "device_bytes": v[4]}

    @property
    def launches(self):
        return int(self._lib.baobzi_launch_count(self.ptr))
