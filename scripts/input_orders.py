"""Throughput of the device path for input ORDERS other than the benchmark's uniform random one (a user's lattice, a
cluster inside one leaf, points sorted along x): no distribution may fall off a cliff.
usage (GPU box): python scripts/input_orders.py [--workload C3] [--points 100000000]"""
import argparse
import json
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    import torch
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="C3")
    ap.add_argument("--points", type=int, default=100_000_000)
    ap.add_argument("--mode", default=None)
    a = ap.parse_args()
    cfg = bench.WORKLOADS[a.workload]
    dim = cfg["dim"]
    dev = torch.device("cuda", 0)
    tmp = tempfile.TemporaryDirectory()
    b = bench.fit_tree_to_file(cfg, os.path.join(tmp.name, "t.baobzi"))
    if a.mode:
        b.set_mode(a.mode, 0)
    c = torch.tensor(cfg["center"], dtype=torch.float64, device=dev)
    h = torch.tensor(cfg["hl"], dtype=torch.float64, device=dev)
    n = a.points
    g = torch.Generator(device=dev).manual_seed(7)

    def uniform():
        x = torch.empty((n, dim), dtype=torch.float64, device=dev)
        x.uniform_(0.0, 1.0, generator=g)
        return x.mul_(2.0).sub_(1.0).mul_(h).add_(c)

    def lattice():  # regular grid in C order (last coordinate fastest), cell-centred
        m = int(round(n ** (1.0 / dim)))
        axes = [((torch.arange(m, dtype=torch.float64, device=dev) + 0.5) / m * 2.0 - 1.0) * h[j] + c[j] for j in range(dim)]
        x = torch.stack(torch.meshgrid(*axes, indexing="ij"), dim=-1).reshape(-1, dim).contiguous()
        return x

    def cluster():  # every point inside a box 1e-3 of the domain wide: one leaf
        x = torch.empty((n, dim), dtype=torch.float64, device=dev)
        x.uniform_(0.0, 1.0, generator=g)
        return x.mul_(2.0).sub_(1.0).mul_(h * 1e-3).add_(c + 0.3 * h)

    def sorted_x():
        x = uniform()
        order = torch.argsort(x[:, 0])
        return x[order].contiguous()

    out = {}
    for name, make in (("uniform", uniform), ("lattice", lattice), ("cluster", cluster), ("sorted_x", sorted_x)):
        x = make()
        m = x.shape[0]
        res = torch.empty(m, dtype=torch.float64, device=dev)
        stream = torch.cuda.current_stream(dev)
        for _ in range(2):
            b._lib.baobzi_eval_multi_device(b.ptr, x.data_ptr(), res.data_ptr(), m, stream.cuda_stream)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        b.profile_begin()
        e0.record(stream)
        for _ in range(3):
            b._lib.baobzi_eval_multi_device(b.ptr, x.data_ptr(), res.data_ptr(), m, stream.cuda_stream)
        e1.record(stream)
        torch.cuda.synchronize()
        st = b.profile_end()
        ms = e0.elapsed_time(e1) / 3
        out[name] = {"points": m, "ms": ms, "evals_per_s": m / ms * 1e3,
                     "stages_ms": {k: round(v[0] / 3, 3) for k, v in st.items() if v[1]},
                     "nan": int(torch.isnan(res[:1 << 20]).sum().item())}
        print(name, json.dumps(out[name]), flush=True)
        del x, res
    b.free()


if __name__ == "__main__":
    main()
