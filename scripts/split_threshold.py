"""Where does the split sort start to pay?  Trees of different leaf counts (the C1 / C3 functions at several orders), 1e8
uniform points, single cooperative sort launch against the split sort on the same handle type.
usage (GPU box): python scripts/split_threshold.py"""
import os
import struct
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    import torch
    dev = torch.device("cuda", 0)
    n = 100_000_000
    tmp = tempfile.TemporaryDirectory()
    for wl, order in (("C1", 10), ("C1", 14), ("C1", 16), ("C3", 10), ("C3", 12)):
        cfg = dict(bench.WORKLOADS[wl])
        cfg["order"] = order
        dim = cfg["dim"]
        c = torch.tensor(cfg["center"], dtype=torch.float64, device=dev)
        h = torch.tensor(cfg["hl"], dtype=torch.float64, device=dev)
        x = torch.empty((n, dim), dtype=torch.float64, device=dev).uniform_(0.0, 1.0).mul_(2.0).sub_(1.0).mul_(h).add_(c)
        res = torch.empty(n, dtype=torch.float64, device=dev)
        stream = torch.cuda.current_stream(dev)
        line = {}
        for split in ("0", "1"):
            os.environ["BAOBZI_B200_SORT_SPLIT"] = split       # read when the handle builds its workspace
            b = bench.fit_tree_to_file(cfg, os.path.join(tmp.name, "t.baobzi"))
            leaves = struct.unpack_from("I", b.device_view(0), 100)[0]   # TreeView.n_leaves
            b.set_mode("binned", 0)
            for _ in range(2):
                b._lib.baobzi_eval_multi_device(b.ptr, x.data_ptr(), res.data_ptr(), n, stream.cuda_stream)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(3):
                b._lib.baobzi_eval_multi_device(b.ptr, x.data_ptr(), res.data_ptr(), n, stream.cuda_stream)
            e1.record(stream)
            torch.cuda.synchronize()
            line[split] = e0.elapsed_time(e1) / 3
            b.free()
        print("%s order %d: %d leaves, single launch %.3f ms, split %.3f ms (%+.1f %%)"
              % (wl, order, leaves, line["0"], line["1"], 100.0 * (line["0"] / line["1"] - 1.0)), flush=True)
        del x, res
    os.environ.pop("BAOBZI_B200_SORT_SPLIT", None)


if __name__ == "__main__":
    main()
