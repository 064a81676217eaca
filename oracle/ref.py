"""TEST INFRASTRUCTURE ONLY -- ctypes access to the checkers under oracle/.

May be imported by tests/, tests/golden/make_golden.py, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py.  Never by the product (baobzi_b200/).

Three libraries:
  * oracle/_ref/libbaobzi_ref.so     the UNMODIFIED reference (built by oracle/Makefile from
                                     /root/reference/src/*.cpp); C ABI of include/baobzi.h:62-99
  * oracle/_ref/libbaobzi_refidx.so  leaf-index probe (oracle/ref_probe.cpp)
  * oracle/_build/liboracle.so       the plain-C restatement (oracle/baobzi_oracle.c)
  * oracle/_build/libtestfuncs.so    C callbacks of examples/C/baobzi_timing.c:19-41 etc.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SO = os.path.join(HERE, "_ref", "libbaobzi_ref.so")
REFIDX_SO = os.path.join(HERE, "_ref", "libbaobzi_refidx.so")
ORACLE_SO = os.path.join(HERE, "_build", "liboracle.so")
FUNCS_SO = os.path.join(HERE, "_build", "libtestfuncs.so")

INPUT_FUNC = C.CFUNCTYPE(None, C.POINTER(C.c_double), C.POINTER(C.c_double), C.c_void_p)
c_double_p = C.POINTER(C.c_double)


class BaobziInput(C.Structure):
    """include/baobzi.h:21-38 (current header, incl. tol_type)."""
    _fields_ = [("func", C.c_void_p), ("data", C.c_void_p), ("dim", C.c_int), ("output_dim", C.c_int),
                ("order", C.c_int), ("tol", C.c_double), ("minimum_leaf_fraction", C.c_double),
                ("split_multi_eval", C.c_int), ("min_depth", C.c_int), ("max_depth", C.c_int),
                ("tol_type", C.c_int)]


class BaobziStruct(C.Structure):
    """include/baobzi.h:47-57."""
    _fields_ = [("obj", C.c_void_p), ("DIM", C.c_int), ("OUTPUT_DIM", C.c_int), ("ORDER", C.c_int),
                ("eval", C.c_void_p), ("eval_multi", C.c_void_p), ("save", C.c_void_p),
                ("stats", C.c_void_p), ("free", C.c_void_p)]


baobzi_t = C.POINTER(BaobziStruct)


def bind_abi(lib):
    """Declare the 7 ABI functions on a loaded library (reference or ours)."""
    lib.baobzi_init.restype = baobzi_t
    lib.baobzi_init.argtypes = [C.POINTER(BaobziInput), c_double_p, c_double_p]
    lib.baobzi_restore.restype = baobzi_t
    lib.baobzi_restore.argtypes = [C.c_char_p]
    lib.baobzi_eval.restype = None
    lib.baobzi_eval.argtypes = [baobzi_t, c_double_p, c_double_p]
    lib.baobzi_eval_multi.restype = None
    lib.baobzi_eval_multi.argtypes = [baobzi_t, c_double_p, c_double_p, C.c_int]
    lib.baobzi_save.restype = None
    lib.baobzi_save.argtypes = [baobzi_t, C.c_char_p]
    lib.baobzi_stats.restype = None
    lib.baobzi_stats.argtypes = [baobzi_t]
    lib.baobzi_free.restype = baobzi_t
    lib.baobzi_free.argtypes = [baobzi_t]
    return lib


_cache = {}


def have_ref():
    return os.path.exists(REF_SO)


def ref_lib():
    if "ref" not in _cache:
        _cache["ref"] = bind_abi(C.CDLL(REF_SO))
    return _cache["ref"]


def refidx_lib():
    if "refidx" not in _cache:
        lib = C.CDLL(REFIDX_SO)
        lib.refidx_node_indices.restype = C.c_int
        lib.refidx_node_indices.argtypes = [C.c_char_p, c_double_p, C.c_int64, C.POINTER(C.c_int64),
                                            C.POINTER(C.c_int64)]
        _cache["refidx"] = lib
    return _cache["refidx"]


def funcs_lib():
    if "funcs" not in _cache:
        _cache["funcs"] = C.CDLL(FUNCS_SO)
    return _cache["funcs"]


def func_ptr(name):
    return C.cast(getattr(funcs_lib(), name), C.c_void_p).value


def dptr(a):
    return a.ctypes.data_as(c_double_p)


def make_input(func, dim, order, tol, data=None, output_dim=1, minimum_leaf_fraction=0.0, split_multi_eval=1,
               min_depth=0, max_depth=50, tol_type=0):
    """baobzi_input_default (include/baobzi.h:34-36) with overrides; `func` = name in testfuncs.c or address."""
    inp = BaobziInput()
    inp.func = func_ptr(func) if isinstance(func, str) else func
    inp._keep = None
    if data is not None:
        inp._keep = C.c_double(float(data))
        inp.data = C.cast(C.pointer(inp._keep), C.c_void_p)
    inp.dim, inp.output_dim, inp.order = dim, output_dim, order
    inp.tol, inp.minimum_leaf_fraction = tol, minimum_leaf_fraction
    inp.split_multi_eval, inp.min_depth, inp.max_depth, inp.tol_type = split_multi_eval, min_depth, max_depth, tol_type
    return inp


class Handle:
    """A baobzi_t of `lib` (the reference by default) with numpy convenience."""

    def __init__(self, ptr, lib):
        if not ptr:
            raise RuntimeError("baobzi handle is NULL")
        self.ptr, self.lib = ptr, lib
        self.dim, self.order, self.output_dim = ptr[0].DIM, ptr[0].ORDER, ptr[0].OUTPUT_DIM

    def eval_multi(self, x):
        x = np.ascontiguousarray(x, dtype=np.float64)
        n = x.size // self.dim
        res = np.empty(n * max(1, self.output_dim), dtype=np.float64)
        self.lib.baobzi_eval_multi(self.ptr, dptr(x), dptr(res), n)
        return res[:n]

    def eval(self, x):
        x = np.ascontiguousarray(x, dtype=np.float64)
        y = np.empty(max(1, self.output_dim), dtype=np.float64)
        self.lib.baobzi_eval(self.ptr, dptr(x), dptr(y))
        return y

    def save(self, path):
        self.lib.baobzi_save(self.ptr, os.fsencode(path))

    def stats(self):
        self.lib.baobzi_stats(self.ptr)

    def free(self):
        if self.ptr:
            self.lib.baobzi_free(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def set_arch(arch):
    """Pin the reference's CPU dispatch (src/baobzi.cpp:39-55). Read at init/restore time."""
    if arch is None:
        os.environ.pop("BAOBZI_ARCH", None)
    else:
        os.environ["BAOBZI_ARCH"] = arch


def ref_init(inp, center, half_length, arch=None, lib=None):
    lib = lib or ref_lib()
    set_arch(arch)
    c = np.ascontiguousarray(center, dtype=np.float64)
    h = np.ascontiguousarray(half_length, dtype=np.float64)
    return Handle(lib.baobzi_init(C.byref(inp), dptr(c), dptr(h)), lib)


def ref_restore(path, arch=None, lib=None):
    lib = lib or ref_lib()
    set_arch(arch)
    return Handle(lib.baobzi_restore(os.fsencode(path)), lib)


def ref_node_indices(path, x, dim):
    """(global node index, coeff offset) per point from Function::get_global_node_index; -1 out of domain."""
    x = np.ascontiguousarray(x, dtype=np.float64)
    n = x.size // dim
    idx = np.empty(n, dtype=np.int64)
    off = np.empty(n, dtype=np.int64)
    rc = refidx_lib().refidx_node_indices(os.fsencode(path), dptr(x), n, idx.ctypes.data_as(C.POINTER(C.c_int64)),
                                          off.ctypes.data_as(C.POINTER(C.c_int64)))
    if rc != 0:
        raise RuntimeError("refidx_node_indices failed rc=%d" % rc)
    return idx, off


def ulp_diff(a, b):
    """Distance in units in the last place between two float64 arrays (NaN==NaN counts 0)."""
    a = np.ascontiguousarray(a, dtype=np.float64)
    b = np.ascontiguousarray(b, dtype=np.float64)
    ia = a.view(np.int64).copy()
    ib = b.view(np.int64).copy()
    ia = np.where(ia < 0, np.int64(-2**63) - ia, ia)
    ib = np.where(ib < 0, np.int64(-2**63) - ib, ib)
    d = np.abs(ia - ib).astype(np.float64)
    both_nan = np.isnan(a) & np.isnan(b)
    one_nan = np.isnan(a) ^ np.isnan(b)
    d[both_nan] = 0
    d[one_nan] = np.inf
    return d


# ---------------------------------------------------------------------------------------------
# plain-C restatement (oracle/baobzi_oracle.c)
# ---------------------------------------------------------------------------------------------
ARITH_GENERIC, ARITH_AVX, ARITH_AVX2, ARITH_AVX512 = 0, 1, 2, 3
ARITH_FMA = ARITH_AVX2
ARITH_OF_ARCH = {"generic": ARITH_GENERIC, "avx": ARITH_AVX, "avx2": ARITH_AVX2, "avx512": ARITH_AVX512}


class OracleTreeStruct(C.Structure):
    """oracle_tree_t (oracle/baobzi_oracle.h)."""
    _fields_ = [("dim", C.c_int), ("order", C.c_int), ("output_dim", C.c_int), ("split_multi_eval", C.c_int),
                ("tol", C.c_double),
                ("box_center", C.c_double * 3), ("box_inv_half_length", C.c_double * 3),
                ("lower_left", C.c_double * 3), ("upper_right", C.c_double * 3), ("inv_bin_size", C.c_double * 3),
                ("n_subtrees", C.c_int * 3),
                ("n_bins", C.c_int64), ("n_nodes", C.c_int64), ("n_leaves", C.c_int64), ("n_coeffs", C.c_int64),
                ("subtree_node_offsets", C.POINTER(C.c_int32)),
                ("center", c_double_p), ("inv_half_length", c_double_p),
                ("first_child_idx", C.POINTER(C.c_uint32)), ("coeff_offset", C.POINTER(C.c_uint64)),
                ("node_output_dim", C.POINTER(C.c_uint32)), ("coeffs", c_double_p)]


def oracle_lib():
    if "oracle" not in _cache:
        lib = C.CDLL(ORACLE_SO)
        P = C.POINTER(OracleTreeStruct)
        lib.oracle_restore.restype = P
        lib.oracle_restore.argtypes = [C.c_char_p]
        lib.oracle_free.restype = None
        lib.oracle_free.argtypes = [P]
        lib.oracle_node_index_multi.restype = None
        lib.oracle_node_index_multi.argtypes = [P, c_double_p, C.c_int64, C.POINTER(C.c_int64)]
        lib.oracle_eval_multi.restype = None
        lib.oracle_eval_multi.argtypes = [P, c_double_p, c_double_p, C.c_int64, C.c_int]
        lib.oracle_eval.restype = None
        lib.oracle_eval.argtypes = [P, c_double_p, c_double_p, C.c_int]
        lib.oracle_cheb_eval.restype = C.c_double
        lib.oracle_cheb_eval.argtypes = [C.c_int, C.c_int, c_double_p, c_double_p, C.c_int]
        _cache["oracle"] = lib
    return _cache["oracle"]


class OracleTree:
    def __init__(self, path):
        self.lib = oracle_lib()
        self.ptr = self.lib.oracle_restore(os.fsencode(path))
        if not self.ptr:
            raise RuntimeError("oracle_restore failed for %s" % path)
        s = self.ptr[0]
        self.dim, self.order, self.output_dim = s.dim, s.order, s.output_dim
        self.n_nodes, self.n_leaves, self.n_bins, self.n_coeffs = s.n_nodes, s.n_leaves, s.n_bins, s.n_coeffs

    def arrays(self):
        """numpy copies of every field (for checking the product's flatten against)."""
        s, d = self.ptr[0], self.dim
        A = np.ctypeslib.as_array
        return dict(
            lower_left=np.array(s.lower_left[:d]), upper_right=np.array(s.upper_right[:d]),
            inv_bin_size=np.array(s.inv_bin_size[:d]), n_subtrees=np.array(s.n_subtrees[:d]),
            box_center=np.array(s.box_center[:d]), box_inv_half_length=np.array(s.box_inv_half_length[:d]),
            tol=s.tol, split_multi_eval=s.split_multi_eval, output_dim=s.output_dim,
            subtree_node_offsets=A(s.subtree_node_offsets, (s.n_bins + 1,)).copy(),
            center=A(s.center, (s.n_nodes, d)).copy(), inv_half_length=A(s.inv_half_length, (s.n_nodes, d)).copy(),
            first_child_idx=A(s.first_child_idx, (s.n_nodes,)).copy(),
            coeff_offset=A(s.coeff_offset, (s.n_nodes,)).copy(),
            node_output_dim=A(s.node_output_dim, (s.n_nodes,)).copy(),
            coeffs=A(s.coeffs, (max(1, s.n_coeffs),)).copy()[:s.n_coeffs])

    def node_indices(self, x):
        x = np.ascontiguousarray(x, dtype=np.float64)
        n = x.size // self.dim
        idx = np.empty(n, dtype=np.int64)
        self.lib.oracle_node_index_multi(self.ptr, dptr(x), n, idx.ctypes.data_as(C.POINTER(C.c_int64)))
        return idx

    def eval_multi(self, x, arith=ARITH_FMA):
        x = np.ascontiguousarray(x, dtype=np.float64)
        n = x.size // self.dim
        res = np.empty(n, dtype=np.float64)
        self.lib.oracle_eval_multi(self.ptr, dptr(x), dptr(res), n, arith)
        return res

    def eval(self, x, arith=ARITH_FMA):
        x = np.ascontiguousarray(x, dtype=np.float64)
        y = np.empty(max(1, self.output_dim), dtype=np.float64)
        self.lib.oracle_eval(self.ptr, dptr(x), dptr(y), arith)
        return y

    def free(self):
        if self.ptr:
            self.lib.oracle_free(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass
