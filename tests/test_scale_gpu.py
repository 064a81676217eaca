"""GPU suite at BASELINE.json sizes: trees of the benchmark configurations (fitted by our host fitter with the
reference's callbacks), checked against the oracle on a sample it finishes in seconds and, at full batch sizes,
through size-independent properties: kernel-strategy independence (direct == binned, any chunking), permutation
equivariance, shard-concatenation equivalence and a checksum of checksums."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

CONFIGS = {  # BASELINE.json configs
    "C1": dict(func="testfun_2d", dim=2, order=8, tol=1e-10, center=[1.5, 3.0], hl=[1.0, 1.0], data=1.5),
    "C2": dict(func="testfun_1d", dim=1, order=16, tol=1e-10, center=[2.0], hl=[1.0], data=1.5),
    "C3": dict(func="testfun_3d", dim=3, order=8, tol=1e-12, center=[1.5, 3.0, 1.5], hl=[1.0, 1.0, 1.0], data=None),
    "C5": dict(func="testfun_3d", dim=3, order=12, tol=1e-12, center=[1.5, 3.0, 1.5], hl=[1.0, 1.0, 1.0], data=None),
}
# The two trees that take the paths the configurations above never reach (fitted once per session, ~10 s each):
#   C4  BASELINE.json configs[3]: 2D order 6 oscillatory, 1 041 664 leaves, 300 MB of coefficients (> L2): the only
#       configuration that selects the bucketed path in auto mode
#   B3  a 3D tree above 40 960 leaves (C3's function on a 4 x 2 x 2 box, order 6: 65 536 leaves): the histogram no
#       longer fits in shared memory, so the binned path falls back to its four-kernel sort
BIG = {
    "C4": dict(func="testfun_2d_osc", dim=2, order=6, tol=1e-10, center=[0.0, 0.0], hl=[1.0, 1.0], data=20.0),
    "B3": dict(func="testfun_3d", dim=3, order=6, tol=3e-11, center=[2.0, 3.0, 1.5], hl=[2.0, 1.0, 1.0], data=None),
}


def make(gpu_lib, ref, cfg, **kw):
    return gpu_lib.Baobzi(fin=ref.func_ptr(cfg["func"]), dim=cfg["dim"], order=cfg["order"],
                          center=np.array(cfg["center"]), half_length=np.array(cfg["hl"]), tol=cfg["tol"],
                          data=cfg["data"], **kw)


@pytest.mark.parametrize("tag", list(CONFIGS))
def test_benchmark_trees_match_oracle(gpu_lib, ref, tmp_path, tag):
    cfg = CONFIGS[tag]
    b = make(gpu_lib, ref, cfg)
    path = str(tmp_path / (tag + ".baobzi"))
    b.save(path)
    t = ref.OracleTree(path)
    rng = np.random.default_rng(42)
    n = 200000 if cfg["dim"] < 3 else 100000
    x = np.array(cfg["center"]) + (rng.random((n, cfg["dim"])) * 2 - 1) * np.array(cfg["hl"])
    assert np.array_equal(b.node_index(x), t.node_indices(x))
    assert ref.ulp_diff(b(x), t.eval_multi(x, ref.ARITH_AVX2)).max() == 0
    b.free()


@pytest.fixture(scope="module")
def big_trees(gpu_lib, ref, tmp_path_factory):
    made = {}

    def get(tag):
        if tag not in made:
            b = make(gpu_lib, ref, BIG[tag])
            path = str(tmp_path_factory.mktemp("big") / (tag + ".baobzi"))
            b.save(path)
            made[tag] = (b, ref.OracleTree(path))
        return made[tag]
    yield get
    for b, _ in made.values():
        b.free()


def _uniform(cfg, n, seed):
    import torch
    g = torch.Generator(device="cuda").manual_seed(seed)
    c = torch.tensor(cfg["center"], dtype=torch.float64, device="cuda")
    h = torch.tensor(cfg["hl"], dtype=torch.float64, device="cuda")
    return c + (torch.rand((n, cfg["dim"]), generator=g, dtype=torch.float64, device="cuda") * 2 - 1) * h


def _stages(b, x):
    b.profile_begin()
    y = b(x)
    st = b.profile_end()
    return y, {k for k, v in st.items() if v[1] > 0}


def test_config4_tree_larger_than_l2(gpu_lib, ref, big_trees):
    """BASELINE.json configs[3] at its own tree size: auto mode must take the bucketed path and give the oracle's
    bits; node indices exact; direct and forced-bucketed agree bit for bit on the whole batch."""
    import torch
    b, t = big_trees("C4")
    assert t.n_leaves > 1_000_000 and t.n_nodes > 1_300_000
    n, m = 3_000_000, 200_000
    x = _uniform(BIG["C4"], n, 7)
    b.set_mode("auto")
    y_auto, used = _stages(b, x)
    assert used == {"sort", "eval_grouped"}, used            # bucketed: grouped evaluation, no direct kernel
    y_auto = y_auto.clone()
    xs = x[:m].cpu().numpy()
    want = t.eval_multi(xs, ref.ARITH_AVX2)
    assert ref.ulp_diff(y_auto[:m].cpu().numpy(), want).max() == 0
    assert np.array_equal(b.node_index(xs), t.node_indices(xs))
    b.set_mode("direct")
    y_dir, used = _stages(b, x)
    assert used == {"eval_direct"}, used
    assert torch.equal(y_dir.view(torch.int64), y_auto.view(torch.int64))
    b.set_mode("bucketed", 1 << 19)                           # several chunks: the split sort on bucket keys
    assert torch.equal(b(x).view(torch.int64), y_auto.view(torch.int64))
    b.set_mode("bucketed")
    assert ref.ulp_diff(b(x[:m].contiguous()).cpu().numpy(), want).max() == 0
    b.set_mode("auto")


def test_3d_tree_above_the_fused_sort_limit(gpu_lib, ref, big_trees):
    """65 536 leaves: auto mode selects the binned path, whose sort is the four-kernel fallback here."""
    import torch
    b, t = big_trees("B3")
    assert t.n_leaves > 40960
    n, m = 4_000_000, 100_000
    x = _uniform(BIG["B3"], n, 9)
    b.set_mode("auto")
    y_auto, used = _stages(b, x)
    assert {"scan", "scatter", "eval_grouped"} <= used and "eval_direct" not in used, used
    y_auto = y_auto.clone()
    xs = x[:m].cpu().numpy()
    want = t.eval_multi(xs, ref.ARITH_AVX2)
    assert ref.ulp_diff(y_auto[:m].cpu().numpy(), want).max() == 0
    assert np.array_equal(b.node_index(xs), t.node_indices(xs))
    b.set_mode("direct")
    assert torch.equal(b(x).view(torch.int64), y_auto.view(torch.int64))
    b.set_mode("binned", 1 << 20)                             # several chunks through the fallback sort
    assert torch.equal(b(x).view(torch.int64), y_auto.view(torch.int64))
    b.set_mode("bucketed")                                    # the same tree through the bucketed path
    assert torch.equal(b(x).view(torch.int64), y_auto.view(torch.int64))
    b.set_mode("auto")


@pytest.mark.parametrize("tag,n", [("C1", 10_000_000), ("C3", 20_000_000), ("C5", 4_000_000), ("C2", 50_000_000)])
def test_full_size_properties(gpu_lib, ref, tag, n):
    import torch
    cfg = CONFIGS[tag]
    d = cfg["dim"]
    b = make(gpu_lib, ref, cfg)
    g = torch.Generator(device="cuda").manual_seed(1)
    c = torch.tensor(cfg["center"], dtype=torch.float64, device="cuda")
    h = torch.tensor(cfg["hl"], dtype=torch.float64, device="cuda")
    x = c + (torch.rand((n, d), generator=g, dtype=torch.float64, device="cuda") * 2 - 1) * h
    b.set_mode("direct")
    y = b(x).clone()
    torch.cuda.synchronize()
    assert not torch.isnan(y).any()
    bits = y.view(torch.int64)
    # (1) strategy / chunking independence
    for mode, chunk in (("binned", 0), ("binned", 1 << 18), ("auto", 0)):
        b.set_mode(mode, chunk)
        y2 = b(x)
        torch.cuda.synchronize()
        assert torch.equal(y2.view(torch.int64), bits), (mode, chunk)
    # (2) permutation equivariance: eval(x[p]) == eval(x)[p]
    p = torch.randperm(n, generator=g, device="cuda")
    yp = b(x[p].contiguous())
    torch.cuda.synchronize()
    assert torch.equal(yp.view(torch.int64), bits[p])
    # (3) shards: evaluating two contiguous ranges separately equals the whole (the multi-GPU decomposition)
    k = n // 3
    y3 = torch.cat([b(x[:k].contiguous()), b(x[k:].contiguous())])
    torch.cuda.synchronize()
    assert torch.equal(y3.view(torch.int64), bits)
    # (4) checksum of checksums against the oracle on a strided sample of 2000 chunks x 16 points
    idx = (torch.arange(2000, device="cuda") * (n // 2000)).repeat_interleave(16) + torch.arange(16, device="cuda").repeat(2000)
    xs = x[idx].cpu().numpy()
    import tempfile, os
    with tempfile.TemporaryDirectory() as td:
        path = os.path.join(td, "t.baobzi")
        b.save(path)
        want = ref.OracleTree(path).eval_multi(xs, ref.ARITH_AVX2)
    got = y[idx].cpu().numpy()
    assert np.array_equal(got.view(np.int64).reshape(2000, 16).sum(1), want.view(np.int64).reshape(2000, 16).sum(1))
    b.free()


def test_sharded_entry_single_gpu(gpu_lib, ref):
    """baobzi_eval_multi_sharded with one shard per visible GPU (all of them when the box has several)."""
    import ctypes as C
    import torch
    cfg = CONFIGS["C3"]
    b = make(gpu_lib, ref, cfg)
    ndev = torch.cuda.device_count()
    n = 1_000_000
    xs, ys, whole = [], [], []
    for k in range(ndev):
        g = torch.Generator(device="cuda:%d" % k).manual_seed(100 + k)
        x = torch.rand((n, 3), generator=g, dtype=torch.float64, device="cuda:%d" % k) * 2 - 1
        x += torch.tensor(cfg["center"], dtype=torch.float64, device=x.device)
        xs.append(x)
        ys.append(torch.empty(n, dtype=torch.float64, device=x.device))
    devs = (C.c_int * ndev)(*range(ndev))
    px = (C.c_void_p * ndev)(*[x.data_ptr() for x in xs])
    py = (C.c_void_p * ndev)(*[y.data_ptr() for y in ys])
    cnt = (C.c_int64 * ndev)(*([n] * ndev))
    assert b._lib.baobzi_eval_multi_sharded(b.ptr, ndev, devs, px, py, cnt) == 0
    assert b._lib.baobzi_sync(b.ptr) == 0
    b.set_mode("direct")
    for k in range(ndev):
        with torch.cuda.device(k):
            want = b(xs[k])
            torch.cuda.synchronize()
            assert torch.equal(want.view(torch.int64), ys[k].view(torch.int64))
    b.free()


def test_host_call_spread_over_all_gpus(gpu_lib, ref):
    """ONE baobzi_eval_multi call with host pointers, chunks dealt to workers on every visible GPU
    (baobzi_set_host_devices(0)); pageable and pinned buffers; same bits as the single-GPU call."""
    import torch
    cfg = CONFIGS["C3"]
    b = make(gpu_lib, ref, cfg)
    n = 30_000_001
    rng = np.random.default_rng(77)
    x = np.array(cfg["center"]) + (rng.random((n, 3)) * 2 - 1) * np.array(cfg["hl"])
    one = b(x)
    b.set_host_devices(0)                                  # all visible devices
    many = b(x)
    assert one.tobytes() == many.tobytes()
    xp = torch.from_numpy(x).pin_memory()
    yp = torch.empty(n, dtype=torch.float64).pin_memory()
    b(xp.numpy(), out=yp.numpy())
    assert yp.numpy().tobytes() == one.tobytes()
    b.set_host_devices(1)
    xd = torch.from_numpy(x[:200000]).cuda()
    assert b(xd).cpu().numpy().tobytes() == one[:200000].tobytes()
    b.free()
