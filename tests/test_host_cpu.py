"""CPU suite, part 2: the product's host logic (reader/writer, fitter) and the C-ABI surface.
No compute calls: evaluation only exists on the GPU."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from conftest import ROOT, golden_cases, load_golden


@pytest.mark.parametrize("name", golden_cases())
def test_reader_writer_roundtrip_is_byte_identical(hostlib, tmp_path, name):
    """Files written by the reference, read by our reader and written by our writer: same bytes."""
    path, _ = load_golden(name)
    out = str(tmp_path / "rt.baobzi")
    assert hostlib.bzt_roundtrip(os.fsencode(path), os.fsencode(out)) == 0
    assert open(path, "rb").read() == open(out, "rb").read()


def test_reader_rejects_bad_input(hostlib, tmp_path):
    path, _ = load_golden("tc_1d1")
    raw = open(path, "rb").read()
    msg = C.create_string_buffer(256)
    cases = {
        "v4": raw[:3] + b"\x04" + raw[4:],                # version != 5 (reference baobzi.cpp:94-97)
        "order7": raw[:2] + b"\x07" + raw[3:],             # unsupported order (baobzi.cpp:102-108)
        "trunc": raw[: len(raw) // 3],
        "empty": b"",
        "junk": b"\xc1" * 64,
    }
    for tag, data in cases.items():
        p = tmp_path / (tag + ".baobzi")
        p.write_bytes(data)
        assert hostlib.bzt_read_status(os.fsencode(str(p)), msg, 256) != 0, tag
    assert hostlib.bzt_read_status(os.fsencode(str(tmp_path / "nope")), msg, 256) != 0
    assert hostlib.bzt_read_status(os.fsencode(path), msg, 256) == 0


def _repack(name, edit):
    """A reference-written golden file decoded with python-msgpack, edited, re-encoded (an unedited re-encode is
    byte-identical to the file, checked below)."""
    import copy
    import msgpack
    raw = open(load_golden(name)[0], "rb").read()
    u = msgpack.Unpacker(raw=False, strict_map_key=False)
    u.feed(raw)
    head, body = list(u)
    assert msgpack.packb(head) + msgpack.packb(body) == raw
    body = copy.deepcopy(body)
    edit(body)
    return msgpack.packb(head) + msgpack.packb(body)


def test_reader_rejects_structurally_bad_trees(hostlib, tmp_path):
    """Files that parse as msgpack but would make the kernels read out of bounds or loop for ever."""
    msg = C.create_string_buffer(256)

    def leaf_of(body, sub=0):
        return next(n for n in body["subtrees_"][sub][0] if n[1] == 0xFFFFFFFF)

    def deep_sub(body):
        return next(s[0] for s in body["subtrees_"] if len(s[0]) > 1)

    def wrap(body):                     # coeff_offset + need wraps around 2^64
        leaf_of(body)[2] = 2 ** 64 - 2

    def past_end(body):
        leaf_of(body)[2] = len(body["coeffs_"]) - 2

    def multi_output_short(body):       # a 1D leaf claiming more output blocks than coeffs_ holds
        n = leaf_of(body, len(body["subtrees_"]) - 1)
        n[3] = 64

    def cycle(body):                    # first_child pointing back at the node itself: the descent would never end
        nodes = deep_sub(body)
        parent = next(i for i, n in enumerate(nodes) if n[1] != 0xFFFFFFFF)
        nodes[parent][1] = parent if parent + 2 <= len(nodes) else 0

    def child_before_parent(body):
        nodes = deep_sub(body)
        inner = [i for i, n in enumerate(nodes) if n[1] != 0xFFFFFFFF]
        nodes[inner[-1]][1] = max(inner[-1] - 1, 0) if inner[-1] else 0

    def huge_output_dim(body):
        body["output_dim_"] = 2 ** 31

    for tag, name, edit in [("wrap", "t1d_o6", wrap), ("past_end", "t1d_o6", past_end),
                            ("multi_short", "t1d_o6", multi_output_short), ("cycle", "deep_1d", cycle),
                            ("backlink", "deep_1d", child_before_parent), ("outdim", "t1d_o6", huge_output_dim)]:
        p = tmp_path / (tag + ".baobzi")
        p.write_bytes(_repack(name, edit))
        assert hostlib.bzt_read_status(os.fsencode(str(p)), msg, 256) != 0, tag
    # declared lengths far beyond the bytes that follow must fail before any allocation of that size
    raw = open(load_golden("t1d_o6")[0], "rb").read()
    k = raw.index(b"coeffs_") + len(b"coeffs_")
    assert raw[k] == 0xdc                                   # array16 header of coeffs_
    lie = raw[:k] + b"\xdd\xff\xff\xff\xf0" + raw[k + 3:]     # array32 claiming 4 G entries
    p = tmp_path / "lie.baobzi"
    p.write_bytes(lie)
    assert hostlib.bzt_read_status(os.fsencode(str(p)), msg, 256) != 0
    k = raw.index(b"subtrees_") + len(b"subtrees_")
    assert raw[k] == 0xdc
    p.write_bytes(raw[:k] + b"\xdd\xff\xff\xff\xf0" + raw[k + 3:])
    assert hostlib.bzt_read_status(os.fsencode(str(p)), msg, 256) != 0


FIT_CASES = [  # the four fixtures of the reference's tests/test_c.cpp (27-71, 73-121, 123-168, 170-215)
    ("testfun_1d", 1, 8, 1e-10, [3.0], [1.0], 1.5, 1),
    ("testfun_1d2", 1, 8, 1e-10, [3.0], [1.0], 1.5, 2),
    ("testfun_2d", 2, 6, 1e-10, [-10.0, 3.0], [np.pi / 5, np.pi / 5], 1.5, 1),
    ("testfun_3d1", 3, 12, 1e-10, [3.0, -2.0, 1.0], [np.pi / 5, np.pi / 5, np.pi / 10], 1.5, 1),
]


@pytest.mark.parametrize("func,dim,order,tol,center,hl,data,od", FIT_CASES)
def test_fitter_meets_reference_test_tolerances(ref, hostlib, tmp_path, func, dim, order, tol, center, hl, data, od):
    """Our host fitter -> .baobzi -> evaluated by the oracle: the accuracy asserts of tests/test_c.cpp
    (relative error < tol at the lower-left corner and at the centre), plus a random sample."""
    inp = ref.make_input(func, dim, order, tol, data=data, output_dim=od)
    c = np.array(center)
    h = np.array(hl)
    out = str(tmp_path / "fit.baobzi")
    assert hostlib.bzt_fit_save(C.addressof(inp), c.ctypes.data, h.ctypes.data, os.fsencode(out)) == 0
    t = ref.OracleTree(out)
    f = getattr(ref.funcs_lib(), func)
    f.restype = None
    f.argtypes = [ref.c_double_p, ref.c_double_p, C.c_void_p]

    def exact(p):
        y = np.zeros(od)
        f(p.ctypes.data_as(ref.c_double_p), y.ctypes.data_as(ref.c_double_p), inp.data)
        return y

    rng = np.random.default_rng(3)
    for i, p in enumerate([c - h, c.copy()] + [c + (rng.random(dim) * 2 - 1) * h for _ in range(200)]):
        p = np.ascontiguousarray(p)
        rel = np.abs((t.eval(p) - exact(p)) / exact(p))
        assert np.all(rel < (tol if i < 2 else 10 * tol)), (i, p, rel)


STRUCT_CASES = FIT_CASES[:3] + [  # + the 3D benchmark function (BASELINE.json configs[2] at a looser tolerance)
    ("testfun_3d", 3, 8, 1e-6, [1.5, 3.0, 1.5], [1.0, 1.0, 1.0], None, 1),
]


@pytest.mark.parametrize("func,dim,order,tol,center,hl,data,od", STRUCT_CASES)
def test_fitter_tree_matches_reference_structure(ref, have_reference, hostlib, tmp_path, func, dim, order, tol,
                                                 center, hl, data, od):
    """Geometry is computed with the reference's expressions, so wherever accept/refine decisions agree the trees
    are identical in structure; coefficients agree to rounding of the dense solves."""
    if not have_reference:
        pytest.skip("needs oracle/_ref")
    inp = ref.make_input(func, dim, order, tol, data=data, output_dim=od)
    ours = str(tmp_path / "ours.baobzi")
    theirs = str(tmp_path / "ref.baobzi")
    c, h = np.array(center), np.array(hl)
    assert hostlib.bzt_fit_save(C.addressof(inp), c.ctypes.data, h.ctypes.data, os.fsencode(ours)) == 0
    # the reference's own geometry differs in the last bit between its builds (the FMA builds contract
    # (bin + 0.5) * bin_size + lower_left); ours follows the uncontracted (generic) build
    ref.ref_init(inp, center, hl, arch="generic").save(theirs)
    a, b = ref.OracleTree(ours).arrays(), ref.OracleTree(theirs).arrays()
    for k in ("lower_left", "upper_right", "inv_bin_size", "n_subtrees", "box_center", "box_inv_half_length"):
        assert np.array_equal(a[k], b[k]), k
    if a["first_child_idx"].shape == b["first_child_idx"].shape and np.array_equal(a["first_child_idx"], b["first_child_idx"]):
        assert np.array_equal(a["center"], b["center"]) and np.array_equal(a["inv_half_length"], b["inv_half_length"])
        assert np.array_equal(a["coeff_offset"], b["coeff_offset"])
        scale = np.abs(b["coeffs"]).max()
        assert np.abs(a["coeffs"] - b["coeffs"]).max() < 1e-11 * scale
    else:  # a borderline tolerance decision flipped: sizes must still be close
        assert abs(len(a["first_child_idx"]) - len(b["first_child_idx"])) < 0.05 * len(b["first_child_idx"]) + 2 ** dim + 1


def test_fitter_written_file_restores_in_reference(ref, have_reference, hostlib, tmp_path):
    """new-save -> reference-restore -> bitwise eval equality with the oracle (mirrors test_c.cpp:57-68)."""
    if not have_reference:
        pytest.skip("needs oracle/_ref")
    inp = ref.make_input("testfun_2d", 2, 8, 1e-8, data=1.5)
    c, h = np.array([1.5, 3.0]), np.array([1.0, 1.0])
    out = str(tmp_path / "ours.baobzi")
    assert hostlib.bzt_fit_save(C.addressof(inp), c.ctypes.data, h.ctypes.data, os.fsencode(out)) == 0
    x = np.random.default_rng(5).random((5000, 2)) * 2 - 1 + c
    yr = ref.ref_restore(out, arch="avx2").eval_multi(x)
    assert ref.ulp_diff(ref.OracleTree(out).eval_multi(x), yr).max() == 0


def test_fitter_max_depth_error(ref, hostlib, tmp_path, capfd):
    inp = ref.make_input("testfun_2d_osc", 2, 6, 1e-12, data=200.0, max_depth=3)
    c, h = np.array([0.0, 0.0]), np.array([1.0, 1.0])
    rc = hostlib.bzt_fit_save(C.addressof(inp), c.ctypes.data, h.ctypes.data, os.fsencode(str(tmp_path / "x")))
    assert rc == 1
    assert "depth exceeded" in capfd.readouterr().err  # same message as the reference's MaxDepthExceeded


# ---------------------------------------------------------------------------------------------- C ABI surface
def _declared_functions():
    text = open(os.path.join(ROOT, "include", "baobzi.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(baobzi_[a-z0-9_]+)\s*\(", text)) - {"baobzi_input_t"})


def test_library_exports_every_declared_symbol():
    """The shared library loads without a GPU and exports exactly what include/baobzi.h declares."""
    import baobzi_b200
    if not os.path.exists(baobzi_b200.library_path()):
        baobzi_b200.build()
    lib = C.CDLL(baobzi_b200.library_path())
    declared = _declared_functions()
    assert set(declared) == set(baobzi_b200.ABI), (declared, sorted(baobzi_b200.ABI))
    for name in declared + baobzi_b200.DATA_SYMBOLS:
        assert hasattr(lib, name), name
    # the drop-in set of the reference (include/baobzi.h:62-99)
    for name in ("baobzi_init", "baobzi_restore", "baobzi_save", "baobzi_eval", "baobzi_eval_multi", "baobzi_stats",
                 "baobzi_free", "baobzi_input_default"):
        assert hasattr(lib, name)
    lib.baobzi_b200_version.restype = C.c_char_p
    assert b"sm_100a" in lib.baobzi_b200_version()
    # baobzi_input_default carries the reference's defaults (baobzi.h:34-36)
    d = baobzi_b200.BaobziInput.in_dll(lib, "baobzi_input_default")
    assert (d.output_dim, d.order, d.split_multi_eval, d.max_depth, d.tol_type) == (1, 8, 1, 50, 0)
    assert d.tol == 0.0 and d.min_depth == 0 and not d.func
    # NULL-safe entry points (reference baobzi.cpp:118-131)
    lib.baobzi_free.restype = C.c_void_p
    lib.baobzi_free.argtypes = [C.c_void_p]
    assert lib.baobzi_free(None) is None
    lib.baobzi_stats.argtypes = [C.c_void_p]
    lib.baobzi_stats(None)


def test_struct_layouts_match_the_reference_header():
    import baobzi_b200
    assert C.sizeof(baobzi_b200.BaobziStruct) == 64 and baobzi_b200.BaobziStruct.eval.offset == 24
    assert C.sizeof(baobzi_b200.BaobziInput) == 64 and baobzi_b200.BaobziInput.tol.offset == 32


def test_library_contains_sm100a_code_only():
    import baobzi_b200
    if not os.path.exists(baobzi_b200.library_path()):
        baobzi_b200.build()
    r = subprocess.run(["cuobjdump", "-lelf", baobzi_b200.library_path()], stdout=subprocess.PIPE, text=True)
    archs = set(re.findall(r"sm_(\d+a?)", r.stdout))
    assert archs == {"100a"}, archs


def test_init_argument_errors_without_gpu(capfd):
    """tol <= 0, NULL func and unsupported (dim, order) are rejected before any CUDA call (baobzi.cpp:148-176)."""
    import baobzi_b200
    lib = baobzi_b200.load_library()
    inp = baobzi_b200.BaobziInput()
    inp.dim, inp.order, inp.output_dim, inp.tol = 2, 8, 1, 0.0
    c = (C.c_double * 2)(0, 0)
    assert not lib.baobzi_init(C.byref(inp), c, c)
    assert "invalid 'tol'" in capfd.readouterr().err
    inp.tol = 1e-8
    assert not lib.baobzi_init(C.byref(inp), c, c)
    assert "invalid 'func'" in capfd.readouterr().err
    cb = baobzi_b200.INPUT_FUNC(lambda x, y, d: None)
    inp.func = C.cast(cb, C.c_void_p)
    inp.order = 7
    assert not lib.baobzi_init(C.byref(inp), c, c)
    assert "(DIM, ORDER): (2, 7)" in capfd.readouterr().err
    assert not lib.baobzi_restore(b"/nonexistent/file.baobzi")
    assert "Unable to restore" in capfd.readouterr().err
