"""Every opt-in kernel variant of the 3D path gives the bits of the default one (and of the oracle): the FP64
tensor-core kernel (DMMA rounds like the FMA chain it replaces), the strided / run tile orders, the coalescing
scatter, the two-stream sort/evaluate overlap, the split sort (default) against the single cooperative launch, and the bucketed path on a tree with refined cells."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

VARIANTS = [
    {"BAOBZI_B200_K4": "4,16,1,1"},          # DMMA kernel, 16 warps
    {"BAOBZI_B200_K4": "5,16,1,1"},          # DMMA kernel, deferred quad reduction
    {"BAOBZI_B200_K4": "4,8,0,1,0"},         # DFMA kernel, strided tiles (coefficients staged per tile)
    {"BAOBZI_B200_K4": "3,12,0,0,1"},        # P = 3, 12 warps, run order
    {"BAOBZI_B200_SORT_COAL": "1"},          # coalescing scatter
    {"BAOBZI_B200_OVERLAP": "1"},            # sort of chunk c+1 beside the evaluation of chunk c
    {"BAOBZI_B200_FUSED_SORT": "0"},         # four-kernel sort (fallback for trees whose histogram exceeds smem)
    {"BAOBZI_B200_SORT_SKEY": "0"},          # keys through global memory
    {"BAOBZI_B200_SORT_SPLIT": "0"},         # one cooperative sort launch per chunk (no phase A / scatter overlap)
    {"BAOBZI_B200_SORT_SPLIT": "1"},         # split sort forced (the default only from 2048 leaves; fixtures are smaller)
    {"BAOBZI_B200_SORT_SPLIT": "2"},         # split sort, two CTAs per SM and phase
    {"BAOBZI_B200_SORT_SBIN": "0"},          # bin_leaf gathered through L1 instead of shared memory
    {"BAOBZI_B200_PIGGY": "1"},              # the evaluation launches carry the sort of the following chunks (opt-in)
    {"BAOBZI_B200_K4": "4,8,1,1,1"},         # T(z) advanced on the fly
    {"BAOBZI_B200_K4": "3,12,1,1,1"},
    {"BAOBZI_B200_K4": "4,12,1,0,1"},
    {"BAOBZI_B200_K4_EVICT": "1"},           # L2 evict-first hint on the gathered points and slot indices
]


@pytest.mark.parametrize("env", VARIANTS, ids=lambda e: ",".join("%s=%s" % kv for kv in e.items()))
def test_variant_bits(gpu_lib, ref, env):
    import torch
    from conftest import load_golden
    from test_parity_gpu import ulp
    path, g = load_golden("t3d_o8")
    t = ref.OracleTree(path)
    a = t.arrays()
    rng = np.random.default_rng(11)
    n = 300_000
    x = a["lower_left"] + rng.random((n, 3)) * (a["upper_right"] - a["lower_left"]) * 1.01 - 0.005
    want = t.eval_multi(x, ref.ARITH_AVX2)
    xd = torch.from_numpy(x).cuda()
    old = {k: os.environ.get(k) for k in env}
    os.environ.update(env)
    try:
        b = gpu_lib.Baobzi(filename=path, mode="binned")   # the knobs are read when the workspace is built
        for chunk in (0, 70_001):                          # one chunk, then several (exercises the buffer ping-pong)
            b.set_mode("binned", chunk)
            y = b(xd)
            torch.cuda.synchronize()
            assert ulp(ref, y.cpu().numpy(), want) == 0, (env, chunk)
        b.free()
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


SHARED = [("t3d_o12", "3,8,0,0,2"), ("t3d_o12", "2,12,0,1,2"), ("t3d_o12", "3,12,1,0,2"),
          ("t3d_o14", "2,8,0,1,2"), ("t3d_o14", "3,8,1,0,2"), ("t3d_o14", "2,12,1,0,2"),
          ("t3d_o16", "2,8,0,1,2"), ("t3d_o16", "2,8,0,0,2"), ("t3d_o16", "2,12,1,0,2")]


@pytest.mark.parametrize("name,k4", SHARED)
def test_cta_shared_tma_block_variants(gpu_lib, ref, name, k4, monkeypatch):
    """High orders with the leaf block staged once per CTA by the TMA engine (cp.async.bulk + mbarrier, run = 2): same
    bits as the oracle, over one chunk and several, with points outside the domain and ragged sizes."""
    import torch
    from conftest import load_golden
    from test_parity_gpu import ulp
    path, g = load_golden(name)
    t = ref.OracleTree(path)
    a = t.arrays()
    rng = np.random.default_rng(13)
    n = 150_001
    x = a["lower_left"] + rng.random((n, 3)) * (a["upper_right"] - a["lower_left"]) * 1.01 - 0.005
    want = t.eval_multi(x, ref.ARITH_AVX2)
    xd = torch.from_numpy(x).cuda()
    monkeypatch.setenv("BAOBZI_B200_K4", k4)
    b = gpu_lib.Baobzi(filename=path, mode="binned")
    for chunk, m in ((0, n), (40_003, n), (0, 97), (0, 1)):
        b.set_mode("binned", chunk)
        y = b(xd[:m].contiguous())
        torch.cuda.synchronize()
        assert ulp(ref, y.cpu().numpy(), want[:m]) == 0, (k4, chunk, m)
    assert ulp(ref, b(torch.from_numpy(g["x"]).cuda()).cpu().numpy(), g["y_avx2"]) == 0   # the golden edge points
    b.free()


def test_piggyback_schedule_ragged_chunks(gpu_lib, ref, monkeypatch):
    """The evaluation launches carry the sort of the following chunks (opt-in): chunk counts 2, 3, many; a last chunk of one
    point; most points outside the domain (few tiles: the side jobs finish in the tail loops); every point in one leaf."""
    import torch
    from conftest import load_golden
    from test_parity_gpu import ulp
    monkeypatch.setenv("BAOBZI_B200_PIGGY", "1")
    for name in ("t3d_o8", "t3d_o12", "tc_3d"):
        path, g = load_golden(name)
        t = ref.OracleTree(path)
        a = t.arrays()
        lo, hi = a["lower_left"], a["upper_right"]
        rng = np.random.default_rng(29)
        n = 200_001
        cases = {
            "uniform": lo + rng.random((n, 3)) * (hi - lo),
            "mostly_outside": lo + (rng.random((n, 3)) * 3.0 - 1.0) * (hi - lo),
            "one_leaf": lo + (0.4 + 1e-6 * rng.random((n, 3))) * (hi - lo),
        }
        b = gpu_lib.Baobzi(filename=path, mode="binned")
        for tag, x in cases.items():
            want = t.eval_multi(x, ref.ARITH_AVX2)
            xd = torch.from_numpy(np.ascontiguousarray(x)).cuda()
            for chunk in (100_001, 66_667, 50_000, 9_973):   # 2 chunks, 3 + one point, 4 + one point, 21
                b.set_mode("binned", chunk)
                y = b(xd)
                torch.cuda.synchronize()
                assert ulp(ref, y.cpu().numpy(), want) == 0, (name, tag, chunk)
                y2 = b(xd)                                    # again: the rotating buffers carry over between calls
                torch.cuda.synchronize()
                assert torch.equal(y2.view(torch.int64), y.view(torch.int64)), (name, tag, chunk)
        b.free()


def test_bucketed_on_refined_tree(gpu_lib, ref):
    """The bucketed path's per-bucket leaf sort on a tree whose grid cells are subdivided (2D oscillatory fixture)."""
    import torch
    from conftest import load_golden
    from test_parity_gpu import ulp
    path, g = load_golden("osc_2d")
    t = ref.OracleTree(path)
    a = t.arrays()
    rng = np.random.default_rng(5)
    n = 400_000
    x = a["lower_left"] + rng.random((n, 2)) * (a["upper_right"] - a["lower_left"])
    want = t.eval_multi(x, ref.ARITH_AVX2)
    b = gpu_lib.Baobzi(filename=path, mode="bucketed")
    xd = torch.from_numpy(x).cuda()
    for chunk in (0, 50_000):
        b.set_mode("bucketed", chunk)
        y = b(xd)
        torch.cuda.synchronize()
        assert ulp(ref, y.cpu().numpy(), want) == 0
    b.free()


STAGED = [("osc_2d", 1), ("osc_2d", 7), ("osc_2d", 64), ("osc_2d", 4096), ("aniso_2d", 16), ("t2d_o6", 2), ("t2d_o8", 16),
          ("t2d_o12", 5), ("t2d_o16", 4096), ("tc_2d", 8), ("t3d_o6", 4), ("t3d_o8", 64), ("t3d_o10", 4096)]


@pytest.mark.parametrize("name,buckets", STAGED)
def test_bucket_stage_tma(gpu_lib, ref, name, buckets, monkeypatch):
    """Bucketed path with each bucket's nodes + coefficient blocks staged in shared memory by cp.async.bulk
    (bucket_stage.cu; BAOBZI_B200_BUCKET_STAGE=2 makes a workspace that cannot use it an error): the oracle's bits, and
    the bits of the gathering kernel (STAGE=0), for buckets of one to many grid cells, one chunk and several, many
    batches per bucket, points outside the domain, NaN coordinates."""
    import torch
    from conftest import load_golden
    from test_parity_gpu import ulp
    path, g = load_golden(name)
    t = ref.OracleTree(path)
    a = t.arrays()
    rng = np.random.default_rng(buckets)
    n = 300_017
    x = a["lower_left"] + rng.random((n, t.dim)) * (a["upper_right"] - a["lower_left"]) * 1.01 - 0.005
    x[::1013, 0] = np.nan
    want = t.eval_multi(x, ref.ARITH_AVX2)
    xd = torch.from_numpy(x).cuda()
    monkeypatch.setenv("BAOBZI_B200_BUCKETS", str(buckets))
    got = {}
    # (stage, CTA-local order: by grid cell / by leaf / the workspace's choice, stages of the pipelined kernel or 0 = queue kernel)
    # ... "ov": both sort launches of chunk c+1 on a side stream beside the evaluation of chunk c (2D; needs the split sort)
    for stage, order, pipe in (("2", "2", "3"), ("2", "1", "3"), ("2", "2", "2"), ("2", "1", "2"), ("2", "2", "0"),
                               ("2", "1", "0"), ("2", None, None), ("2", None, "ov"), ("0", None, None)):
        monkeypatch.setenv("BAOBZI_B200_BUCKET_STAGE", stage)
        ov = pipe == "ov"
        pipe = None if ov else pipe
        for var, val in (("BAOBZI_B200_BUCKET_SORT", order), ("BAOBZI_B200_BUCKET_PIPE", pipe),
                         ("BAOBZI_B200_BUCKET_OVERLAP", "1" if ov else "0"), ("BAOBZI_B200_SORT_SPLIT", "1" if ov else None)):
            if val is None:
                monkeypatch.delenv(var, raising=False)
            else:
                monkeypatch.setenv(var, val)
        b = gpu_lib.Baobzi(filename=path, mode="bucketed")
        for chunk in (0, 65_537, 30_011):
            b.set_mode("bucketed", chunk)
            y = b(xd)
            torch.cuda.synchronize()
            assert ulp(ref, y.cpu().numpy(), want) == 0, (name, buckets, stage, order, pipe, ov, chunk)
            got[(stage, order, "ov" if ov else pipe, chunk)] = y
        b.free()
    assert torch.equal(got[("2", "2", "3", 0)].view(torch.int64), got[("0", None, None, 0)].view(torch.int64))


def test_bucket_stage_falls_back_when_a_bucket_does_not_fit(gpu_lib, ref, monkeypatch):
    """Two buckets of 256 KB of coefficients each exceed the stage: the workspace keeps the gathering kernel (same bits),
    and asking for the stage explicitly (=2) is an error instead."""
    import torch
    from conftest import load_golden
    from test_parity_gpu import ulp
    path, g = load_golden("aniso_2d")
    t = ref.OracleTree(path)
    a = t.arrays()
    rng = np.random.default_rng(2)
    x = a["lower_left"] + rng.random((100_000, 2)) * (a["upper_right"] - a["lower_left"])
    want = t.eval_multi(x, ref.ARITH_AVX2)
    xd = torch.from_numpy(x).cuda()
    monkeypatch.setenv("BAOBZI_B200_BUCKETS", "2")
    monkeypatch.setenv("BAOBZI_B200_BUCKET_STAGE", "1")
    b = gpu_lib.Baobzi(filename=path, mode="bucketed")
    y = b(xd)
    torch.cuda.synchronize()
    assert ulp(ref, y.cpu().numpy(), want) == 0
    b.free()
    monkeypatch.setenv("BAOBZI_B200_BUCKET_STAGE", "2")
    b = gpu_lib.Baobzi(filename=path, mode="bucketed")
    with pytest.raises(RuntimeError):
        b(xd)
    b.free()


def test_eval_all_multi_output_1d(gpu_lib, ref):
    """baobzi_eval_multi_device_all: row i == what baobzi_eval writes for point i (every output component), and
    column 0 == baobzi_eval_multi; NaN rows outside the domain; not supported for DIM > 1."""
    import torch
    from conftest import load_golden
    path, g = load_golden("tc_1d2")
    b = gpu_lib.Baobzi(filename=path)
    assert b.output_dim == 2
    t = ref.OracleTree(path)
    a = t.arrays()
    rng = np.random.default_rng(3)
    x = a["lower_left"] + rng.random((5000, 1)) * (a["upper_right"] - a["lower_left"]) * 1.02 - 0.01
    xd = torch.from_numpy(x).cuda()
    y = b.eval_all(xd).cpu().numpy()
    assert y.shape == (5000, 2)
    first = b(xd).cpu().numpy()
    assert np.array_equal(np.nan_to_num(y[:, 0], nan=-7.0), np.nan_to_num(first, nan=-7.0))
    for i in list(range(0, 5000, 97)):
        ys = b.eval(x[i])
        assert np.array_equal(np.nan_to_num(ys, nan=-7.0), np.nan_to_num(y[i], nan=-7.0)), i
    assert np.isnan(y).any() and not np.isnan(y).all()
    b.free()
    path3, _ = load_golden("t3d_o8")
    b3 = gpu_lib.Baobzi(filename=path3)
    with pytest.raises(RuntimeError):
        b3.eval_all(torch.zeros((4, 3), dtype=torch.float64, device="cuda"))
    b3.free()


ONE_D = ["t1d_o6", "t1d_o8", "t1d_o10", "t1d_o12", "t1d_o14", "t1d_o16", "deep_1d", "tc_1d2"]


@pytest.mark.parametrize("kernel", ["table", "sorted", "twoleaf"])
@pytest.mark.parametrize("name", ONE_D)
def test_table_1d_kernel_bits(gpu_lib, ref, name, kernel, monkeypatch):
    """The shared-memory 1D kernels (eval1d.cu) forced on from the first point -- "table": the whole tree in shared
    memory, coefficients per lane; "sorted": trees of at most 16 leaves (6 of these 8 fixtures; the others fall through
    to the table kernel), CTA-local sort by leaf and coefficients as constant-bank operands: golden points (edge cases included)
    and seeded points over and slightly beyond the domain, 0 ulp against the reference's AVX2+FMA values, ragged
    sizes around the tile size, an odd (8-byte aligned only) input offset, and the generic kernel's bits."""
    import torch
    from conftest import load_golden
    from test_parity_gpu import ulp
    path, g = load_golden(name)
    t = ref.OracleTree(path)
    a = t.arrays()
    rng = np.random.default_rng(21)
    n = 200_003
    x = a["lower_left"] + rng.random((n, 1)) * (a["upper_right"] - a["lower_left"]) * 1.02 - 0.01
    x[::1001] = np.nan
    want = t.eval_multi(x, ref.ARITH_AVX2)
    gx = np.ascontiguousarray(g["x"]).reshape(-1, 1)
    gwant = t.eval_multi(gx, ref.ARITH_AVX2)
    old = os.environ.get("BAOBZI_B200_TABLE1D_MIN")
    monkeypatch.setenv("BAOBZI_B200_SORTED1D", "1" if kernel == "sorted" else "0")
    monkeypatch.setenv("BAOBZI_B200_TWOLEAF1D", "1" if kernel == "twoleaf" else "0")   # trees of one or two leaves only
    try:
        b = gpu_lib.Baobzi(filename=path, mode="direct")
        os.environ["BAOBZI_B200_TABLE1D_MIN"] = "-1"       # generic direct kernel
        xd = torch.from_numpy(x).cuda()
        y_generic = b(xd).cpu().numpy()
        os.environ["BAOBZI_B200_TABLE1D_MIN"] = "0"        # table kernel from the first point
        launches = b.launches
        y = b(xd).cpu().numpy()
        assert b.launches == launches + 1
        assert ulp(ref, y, want) == 0
        assert np.array_equal(np.isnan(y), np.isnan(want))
        assert y.tobytes() == y_generic.tobytes()
        assert ulp(ref, b(torch.from_numpy(gx).cuda()).cpu().numpy(), gwant) == 0
        for m in (1, 31, 1023, 1024, 1025, 4097, 65_537):   # ragged tails, and an odd element offset
            ym = b(xd[1:1 + m]).cpu().numpy()
            assert ym.tobytes() == y[1:1 + m].tobytes(), m
        b.free()
    finally:
        if old is None:
            os.environ.pop("BAOBZI_B200_TABLE1D_MIN", None)
        else:
            os.environ["BAOBZI_B200_TABLE1D_MIN"] = old


@pytest.mark.parametrize("order", [6, 8, 10, 12, 14, 16])
@pytest.mark.parametrize("tol", [1e-3, 1e-9])
def test_two_leaf_1d_kernel_every_order(gpu_lib, ref, tmp_path, order, tol, monkeypatch):
    """Trees of one or two leaves (a smooth 1D function, BASELINE.json configs[1]) take the kernel whose coefficients
    are selected between two constant-bank candidates: oracle bits at every order, points outside and NaN included."""
    import torch
    from test_parity_gpu import ulp
    monkeypatch.setenv("BAOBZI_B200_TABLE1D_MIN", "0")
    b = gpu_lib.Baobzi(fin=ref.func_ptr("testfun_1d"), dim=1, order=order, center=np.array([2.0]),
                       half_length=np.array([1.0]), tol=tol, data=1.5)
    path = str(tmp_path / "t.baobzi")
    b.save(path)
    t = ref.OracleTree(path)
    if t.n_leaves > 2:
        b.free()
        pytest.skip("%d leaves at this order / tolerance" % t.n_leaves)
    rng = np.random.default_rng(order)
    n = 300_007
    x = 0.98 + rng.random((n, 1)) * 2.04
    x[::997] = np.nan
    want = t.eval_multi(x, ref.ARITH_AVX2)
    xd = torch.from_numpy(x).cuda()
    launches = b.launches
    y = b(xd).cpu().numpy()
    assert b.launches == launches + 1
    assert ulp(ref, y, want) == 0 and np.array_equal(np.isnan(y), np.isnan(want))
    monkeypatch.setenv("BAOBZI_B200_TWOLEAF1D", "0")
    monkeypatch.setenv("BAOBZI_B200_SORTED1D", "0")
    assert b(xd).cpu().numpy().tobytes() == y.tobytes()           # the table kernel's bytes
    for m in (1, 1023, 1025, 4097):
        assert b(xd[3:3 + m]).cpu().numpy().tobytes() == y[3:3 + m].tobytes()
    b.free()


@pytest.fixture(params=["0", "1"], ids=["single_launch_sort", "split_sort"])
def sort_split(request):
    """Both sort schedules: the golden trees are below the split sort's default leaf threshold, so it is forced."""
    old = os.environ.get("BAOBZI_B200_SORT_SPLIT")
    os.environ["BAOBZI_B200_SORT_SPLIT"] = request.param   # read when a handle builds its workspace
    yield request.param
    if old is None:
        os.environ.pop("BAOBZI_B200_SORT_SPLIT", None)
    else:
        os.environ["BAOBZI_B200_SORT_SPLIT"] = old


@pytest.mark.parametrize("name", ["t3d_o8", "t2d_o8", "tc_2d"])
def test_clustered_and_ordered_inputs(gpu_lib, ref, name, sort_split):
    """Inputs that are not uniform random: every point inside ONE leaf (its tile descriptors are then written by the
    whole grid and the histogram updates are aggregated per warp, binned.cu), two heavy leaves plus a sprinkle of
    others, and a lattice in C order -- same bits as the oracle in the binned path."""
    import torch
    from conftest import load_golden
    from test_parity_gpu import ulp
    path, g = load_golden(name)
    t = ref.OracleTree(path)
    a = t.arrays()
    dim = len(a["lower_left"])
    lo, hi = a["lower_left"], a["upper_right"]
    rng = np.random.default_rng(29)
    n = 400_000
    cases = {}
    cases["one_leaf"] = lo + (0.31 + 1e-4 * rng.random((n, dim))) * (hi - lo)
    mix = lo + rng.random((n, dim)) * (hi - lo)
    mix[: n // 2] = lo + (0.62 + 1e-4 * rng.random((n // 2, dim))) * (hi - lo)
    mix[n // 2: n - 1000] = lo + (0.17 + 1e-4 * rng.random((n - 1000 - n // 2, dim))) * (hi - lo)
    cases["two_heavy_leaves"] = mix[rng.permutation(n)]
    m = int(round(n ** (1.0 / dim)))
    axes = [lo[j] + (np.arange(m) + 0.5) / m * (hi[j] - lo[j]) for j in range(dim)]
    cases["lattice"] = np.stack(np.meshgrid(*axes, indexing="ij"), axis=-1).reshape(-1, dim)
    b = gpu_lib.Baobzi(filename=path, mode="binned")
    for label, x in cases.items():
        x = np.ascontiguousarray(x)
        want = t.eval_multi(x, ref.ARITH_AVX2)
        xd = torch.from_numpy(x).cuda()
        for chunk in (0, 150_001):
            b.set_mode("binned", chunk)
            y = b(xd).cpu().numpy()
            assert ulp(ref, y, want) == 0, (label, chunk)
    b.free()


def test_unsynchronised_call_sequences(gpu_lib, ref, sort_split):
    """Back-to-back calls without any synchronisation in between, alternating the device-pointer entries (caller's
    stream) and the host-pointer entry (the library's own non-blocking streams), across handles and modes: every
    workspace must be initialised before the first kernel that reads it, whichever stream that kernel runs on.
    (A cudaMemset on the default stream once raced with the first sort of a host-pointer call.)"""
    import torch
    from conftest import load_golden
    rng = np.random.default_rng(41)
    junk = [torch.full((1 << 22,), 0x7fffffff, dtype=torch.int32, device="cuda") for _ in range(8)]
    del junk                                   # recycled allocations are not zero
    for name in ("t2d_o8", "t3d_o8", "osc_2d"):
        path, g = load_golden(name)
        t = ref.OracleTree(path)
        a = t.arrays()
        n = 60_001
        x = a["lower_left"] + rng.random((n, len(a["lower_left"]))) * (a["upper_right"] - a["lower_left"])
        want = t.eval_multi(x, ref.ARITH_AVX2)
        xd = torch.from_numpy(x).cuda()
        xh = np.ascontiguousarray(x[:5000])
        for mode in ("binned", "bucketed"):
            b = gpu_lib.Baobzi(filename=path, mode=mode)
            outs = []
            for chunk in (0, 20_011):
                b.set_mode(mode, chunk)
                outs.append(b(xd))
                b.dot(xd)
            idx = b.node_index(xd)
            yh = b(xh)                          # host pointers: other streams, own workspaces
            outs.append(b(xd))
            torch.cuda.synchronize()
            for y in outs:
                assert y.cpu().numpy().tobytes() == want.tobytes(), (name, mode)
            assert np.asarray(yh).tobytes() == want[:5000].tobytes(), (name, mode)
            b.free()


def test_two_caller_streams_share_one_workspace(gpu_lib, ref, sort_split):
    """Calls on one handle from two CUDA streams, no synchronisation between them: the per-device workspace is chained
    through an event, so the second call's sort never overwrites what the first call's evaluation still reads."""
    import torch
    from conftest import load_golden
    path, g = load_golden("t3d_o8")
    t = ref.OracleTree(path)
    a = t.arrays()
    rng = np.random.default_rng(43)
    n = 400_000
    xs = [a["lower_left"] + rng.random((n, 3)) * (a["upper_right"] - a["lower_left"]) for _ in range(2)]
    wants = [t.eval_multi(x, ref.ARITH_AVX2) for x in xs]
    xd = [torch.from_numpy(x).cuda() for x in xs]
    b = gpu_lib.Baobzi(filename=path, mode="binned")
    if sort_split == "1":
        b.set_mode("binned", 100_003)          # several chunks per call: the side stream and both key buffers are in play
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    torch.cuda.synchronize()
    outs = []
    for rep in range(6):
        k = rep % 2
        with torch.cuda.stream(streams[k]):
            outs.append((k, b(xd[k])))
    torch.cuda.synchronize()
    for k, y in outs:
        assert y.cpu().numpy().tobytes() == wants[k].tobytes()
    b.free()
