"""Generate the committed golden fixtures by RUNNING THE UNMODIFIED REFERENCE (oracle/_ref, built by
oracle/Makefile from /root/reference).  Run in the build container:  python tests/golden/make_golden.py

For every case:  <name>.baobzi  -- the tree as written by the reference's baobzi_save
                 <name>.npz     -- x (points), node_idx (Function::get_global_node_index, -1 outside),
                                   y_generic / y_avx / y_avx2 / y_avx512 (baobzi_eval_multi with BAOBZI_ARCH pinned),
                                   ys_avx2 (baobzi_eval at the first 16 in-domain points, all output components)
Points: seeded uniform samples plus the edge cases the domain has: corners, the semi-open upper boundary,
just-outside points, node centres (the strict `x > centre` comparison), grid-cell boundaries, +-1 ulp around them.
The reference's tests hold no golden vectors of their own; the four (function, domain,
order, tol) fixtures of tests/test_c.cpp are among the cases below (tc_*), tolerances loosened only where the
file would otherwise be megabytes (noted per case).
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import ref  # noqa: E402

PI = np.pi
CASES = []


def case(name, func, dim, order, tol, center, hl, data=None, **kw):
    CASES.append(dict(name=name, func=func, dim=dim, order=order, tol=tol, center=center, hl=hl, data=data, kw=kw))


# tests/test_c.cpp fixtures: 1D1 (27-71), 1D2 (73-121), 2D (123-168), 3D (170-215; tol 1e-10 there -> 3.5 MB, here 1e-3)
case("tc_1d1", "testfun_1d", 1, 8, 1e-10, [3.0], [1.0], data=1.5)
case("tc_1d2", "testfun_1d2", 1, 8, 1e-10, [3.0], [1.0], data=1.5, output_dim=2)
case("tc_2d", "testfun_2d", 2, 6, 1e-6, [-10.0, 3.0], [PI / 5, PI / 5], data=1.5)  # tol 1e-10 there
case("tc_3d", "testfun_3d1", 3, 12, 1e-3, [3.0, -2.0, 1.0], [PI / 5, PI / 5, PI / 10], data=1.5)
# examples/C/baobzi_timing.c cases at every supported order (dims 1-3 x orders 6..16), small tolerances relaxed
for o in (6, 8, 10, 12, 14, 16):
    case("t1d_o%d" % o, "testfun_1d", 1, o, 1e-10, [2.0], [1.0], data=1.5, minimum_leaf_fraction=1.0)
    case("t2d_o%d" % o, "testfun_2d", 2, o, {6: 1e-3, 8: 1e-4}.get(o, 1e-6), [1.5, 3.0], [1.0, 1.0], data=1.5)
    case("t3d_o%d" % o, "testfun_3d", 3, o, 1e-5, [1.5, 3.0, 1.5], [1.0, 1.0, 1.0], min_depth=1 if o >= 12 else 0)
# anisotropic domain (n_subtrees differ per axis) and a deeper 1D tree
case("aniso_2d", "testfun_2d_2", 2, 8, 1e-8, [1.5, 3.0], [2.0, 0.5])
case("deep_1d", "testfun_1d", 1, 6, 1e-10, [1.01], [1.0], data=1.5)  # log near 0: subtrees of uneven depth
case("osc_2d", "testfun_2d_osc", 2, 6, 1e-3, [0.0, 0.0], [1.0, 1.0], data=6.0)


def edge_points(tree, rng):
    a = tree.arrays()
    d = tree.dim
    ll, ur = a["lower_left"], a["upper_right"]
    pts = [ll.copy(), 0.5 * (ll + ur), np.nextafter(ur, ll), ur.copy(), np.nextafter(ll, ll - 1), ll - 0.1, ur + 0.1]
    for j in range(d):  # one coordinate out, the rest in
        p = 0.5 * (ll + ur)
        p[j] = ur[j]
        pts.append(p.copy())
        p[j] = np.nextafter(ll[j], ll[j] - 1)
        pts.append(p.copy())
    # node centres and +-1 ulp around them (strict `>` in the descent), for up to 200 nodes
    cen = a["center"]
    pick = rng.choice(len(cen), size=min(200, len(cen)), replace=False)
    for c in cen[pick]:
        pts += [c.copy(), np.nextafter(c, c + 1), np.nextafter(c, c - 1)]
    # grid-cell boundaries
    n_sub = a["n_subtrees"]
    for _ in range(50):
        b = np.array([rng.integers(0, n_sub[j] + 1) for j in range(d)])
        p = ll + b / a["inv_bin_size"]
        pts += [p, np.nextafter(p, p + 1), np.nextafter(p, p - 1)]
    return np.array(pts).reshape(-1, d)


def main():
    rng = np.random.default_rng(20261104)
    total = 0
    for c in CASES:
        d = c["dim"]
        inp = ref.make_input(c["func"], d, c["order"], c["tol"], data=c["data"], **c["kw"])
        h = ref.ref_init(inp, c["center"], c["hl"], arch="avx2")
        path = os.path.join(HERE, c["name"] + ".baobzi")
        h.save(path)
        tree = ref.OracleTree(path)
        x = (rng.random((1500, d)) * 2 - 1) * np.array(c["hl"]) + np.array(c["center"])
        x = np.vstack([x, edge_points(tree, rng)])
        # the reference reads out of bounds when rounding pushes a point just under upper_right into bin n
        # (SURVEY.md 7.2-3); keep only points whose bin is in range so the golden values are defined behaviour
        a = tree.arrays()
        bins = np.trunc((x - a["lower_left"]) * a["inv_bin_size"])
        inside = np.all((x >= a["lower_left"]) & (x < a["upper_right"]), axis=1)
        ok = ~inside | np.all(bins < a["n_subtrees"], axis=1)
        x = np.ascontiguousarray(x[ok])
        out = dict(x=x)
        out["node_idx"], _ = ref.ref_node_indices(path, x, d)
        for arch in ("generic", "avx", "avx2", "avx512"):
            hr = ref.ref_restore(path, arch=arch)
            out["y_" + arch] = hr.eval_multi(x)
            if arch == "avx2":
                ins = np.flatnonzero(out["node_idx"] >= 0)[:16]
                out["scalar_rows"] = ins
                out["ys_avx2"] = np.array([hr.eval(x[i]) for i in ins])
        np.savez_compressed(os.path.join(HERE, c["name"] + ".npz"), **out)
        sz = os.path.getsize(path) + os.path.getsize(os.path.join(HERE, c["name"] + ".npz"))
        total += sz
        print("%-10s dim %d order %2d nodes %6d leaves %6d points %5d  %7.1f KB" %
              (c["name"], d, c["order"], tree.n_nodes, tree.n_leaves, len(x), sz / 1024), flush=True)
    print("total %.2f MB" % (total / 2**20))


if __name__ == "__main__":
    main()
