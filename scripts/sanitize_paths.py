"""Every kernel path once at small sizes, for compute-sanitizer (memcheck / racecheck / synccheck):
    compute-sanitizer --tool memcheck python scripts/sanitize_paths.py
Uses the golden fixtures only (no oracle): exercises direct, 1D table and few-leaf sorted kernels, binned (2D, 3D incl.
one heavy leaf; single-launch cooperative sort, the split sort over >= 3 chunks with its rotating histogram buffers,
and the four-kernel fallback), bucketed, node index, multi-output, dot, scalar, host-pointer entry points (direct and
staged through the pinned ring, several worker threads), an 8-byte-aligned 2D input and two caller streams on one
handle."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import baobzi_b200  # noqa: E402

G = os.path.join(ROOT, "tests", "golden")
GUARD = {"bad": 0, "buffers": 0, "checks": 0}


def release(b):
    """Guard-band check (BAOBZI_B200_GUARD=1: every workspace / chunk buffer sits between two 4 KB bands of a known
    pattern) of everything this handle allocated, then free it."""
    import ctypes as C
    bad, n = C.c_int64(0), C.c_int64(0)
    rc = b._lib.baobzi_b200_guard_check(C.byref(bad), C.byref(n))
    assert rc == 0, rc
    GUARD["bad"] += bad.value
    GUARD["buffers"] = max(GUARD["buffers"], n.value)
    GUARD["checks"] += 1
    b.free()


def domain(path):
    g = np.load(path.replace(".baobzi", ".npz"))
    x = g["x"].reshape(len(g["x"]), -1)
    return x.min(axis=0), x.max(axis=0)


def extra_cases(rng):
    """Paths the mode loop below does not reach."""
    # two caller streams on one handle (the per-device workspace is chained through an event)
    path = os.path.join(G, "t3d_o8.baobzi")
    lo, hi = domain(path)
    b = baobzi_b200.Baobzi(filename=path, mode="binned")
    xs = [torch.from_numpy(lo + rng.random((40_000, 3)) * (hi - lo)).cuda() for _ in range(2)]
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    torch.cuda.synchronize()
    outs = []
    for rep in range(3):
        for x, st in zip(xs, streams):
            with torch.cuda.stream(st):
                outs.append(b(x, stream=st.cuda_stream))
    torch.cuda.synchronize()
    print("two streams", float(outs[-1][-1]), flush=True)
    # host pointers staged through the pinned ring by several worker threads (forced on at this small size)
    os.environ["BAOBZI_B200_STAGE"] = "1"
    xh = lo + rng.random((300_000, 3)) * (hi - lo)
    yh = b(xh)
    os.environ.pop("BAOBZI_B200_STAGE")
    print("staged host path", float(yh[-1]), flush=True)
    release(b)
    # 2D points that start on an odd double
    path = os.path.join(G, "t2d_o8.baobzi")
    lo, hi = domain(path)
    b = baobzi_b200.Baobzi(filename=path)
    flat = torch.empty(2 * 50_000 + 1, dtype=torch.float64, device="cuda")
    flat[1:] = torch.from_numpy(lo + rng.random((50_000, 2)) * (hi - lo)).cuda().reshape(-1)
    for mode in ("auto", "binned", "bucketed"):
        b.set_mode(mode)
        y = b(flat[1:].view(50_000, 2))
    torch.cuda.synchronize()
    print("odd-aligned 2D", float(y[-1]), flush=True)
    release(b)
    # 3D binned with the four-kernel fallback sort
    os.environ["BAOBZI_B200_FUSED_SORT"] = "0"
    path = os.path.join(G, "t3d_o6.baobzi")
    lo, hi = domain(path)
    b = baobzi_b200.Baobzi(filename=path, mode="binned")
    b.set_mode("binned", 20_011)
    y = b(torch.from_numpy(lo + rng.random((60_001, 3)) * (hi - lo)).cuda())
    torch.cuda.synchronize()
    os.environ.pop("BAOBZI_B200_FUSED_SORT")
    print("four-kernel sort", float(y[-1]), flush=True)
    release(b)


def main():
    rng = np.random.default_rng(3)
    os.environ["BAOBZI_B200_TABLE1D_MIN"] = "0"
    extra_cases(rng)
    # the mode loop twice: single cooperative sort launch, then the split sort forced on (3 chunks: every one of the
    # three rotating histogram buffers is filled, scanned and cleared)
    for split in ("0", "1"):
        os.environ["BAOBZI_B200_SORT_SPLIT"] = split
        print("--- BAOBZI_B200_SORT_SPLIT=" + split, flush=True)
        mode_loop(rng, split == "1")
    print("guard bands: %d checks, up to %d buffers watched at once, %d bytes changed" % (GUARD["checks"], GUARD["buffers"], GUARD["bad"]))
    print("sanitize_paths: done")


def mode_loop(rng, binned_only):
    for name, modes in (("t1d_o16", ["direct"]), ("tc_1d2", ["direct"]), ("t2d_o8", ["direct", "binned", "bucketed"]),
                        ("osc_2d", ["bucketed", "binned"]), ("t3d_o8", ["direct", "binned", "bucketed"]),
                        ("t3d_o12", ["binned"]), ("t3d_o6", ["binned"])):
        path = os.path.join(G, name + ".baobzi")
        lo, hi = domain(path)
        for mode in modes:
            if binned_only and mode == "direct":
                continue
            b = baobzi_b200.Baobzi(filename=path, mode=mode)
            n = 60_001
            x = lo + rng.random((n, b.dim)) * (hi - lo)
            x[: n // 2] = lo + (0.4 + 1e-5 * rng.random((n // 2, b.dim))) * (hi - lo)   # a heavy leaf
            xd = torch.from_numpy(np.ascontiguousarray(x)).cuda()
            verbose = bool(os.environ.get("SANITIZE_VERBOSE"))

            def step(label, fn):
                out = fn()
                if verbose:
                    torch.cuda.synchronize()
                    print("   ", name, mode, label, "ok", flush=True)
                return out

            for chunk in (0, 20_011):
                b.set_mode(mode, chunk)
                y = step("eval chunk %d" % chunk, lambda: b(xd))
                s = step("dot chunk %d" % chunk, lambda: b.dot(xd))
            idx = step("node_index", lambda: b.node_index(xd))
            yh = step("host eval", lambda: b(np.ascontiguousarray(x[:5000])))
            ys = step("scalar", lambda: b.eval(x[0]))
            if b.dim == 1:
                ya = step("eval_all", lambda: b.eval_all(xd))
            torch.cuda.synchronize()
            print(name, mode, float(y[-1]), float(s), int(idx[-1]), float(np.asarray(yh)[-1]), np.asarray(ys)[0], flush=True)
            release(b)


if __name__ == "__main__":
    main()
