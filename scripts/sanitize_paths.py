"""Every kernel path once at small sizes, for compute-sanitizer (memcheck / racecheck / synccheck):
    compute-sanitizer --tool memcheck python scripts/sanitize_paths.py
Uses the golden fixtures only (no oracle): exercises direct, 1D table, binned (2D, 3D incl. one heavy leaf), bucketed,
node index, multi-output, dot, scalar and host-pointer entry points."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import baobzi_b200  # noqa: E402

G = os.path.join(ROOT, "tests", "golden")


def domain(path):
    g = np.load(path.replace(".baobzi", ".npz"))
    x = g["x"].reshape(len(g["x"]), -1)
    return x.min(axis=0), x.max(axis=0)


def main():
    rng = np.random.default_rng(3)
    os.environ["BAOBZI_B200_TABLE1D_MIN"] = "0"
    for name, modes in (("t1d_o16", ["direct"]), ("tc_1d2", ["direct"]), ("t2d_o8", ["direct", "binned", "bucketed"]),
                        ("osc_2d", ["bucketed", "binned"]), ("t3d_o8", ["direct", "binned", "bucketed"]),
                        ("t3d_o12", ["binned"]), ("t3d_o6", ["binned"])):
        path = os.path.join(G, name + ".baobzi")
        lo, hi = domain(path)
        for mode in modes:
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
            b.free()
    print("sanitize_paths: done")


if __name__ == "__main__":
    main()
