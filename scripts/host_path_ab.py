"""A/B of the host-pointer path on one box: pinned buffers, pageable buffers staged by 2 / 4 / 8 / 16 workers,
pageable buffers pinned for the call (cudaHostRegister), the driver's own staging (BAOBZI_B200_STAGE=0), and -- when
the box has several GPUs -- one call spread over all of them.  C3 tree, 2e8 points.  Prints one line per variant."""
import ctypes as C
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import baobzi_b200  # noqa: E402


def main():
    cfg = bench.WORKLOADS["C3"]
    n = int(os.environ.get("AB_POINTS", "200000000"))
    b = bench.fit_tree_to_file(cfg, "/tmp/ab_tree.baobzi")
    x = bench.sample_points(cfg, n, seed=3)
    res = np.empty(n)
    res[:] = 0.0
    xp = torch.empty((n, 3), dtype=torch.float64, pin_memory=True)
    xp.numpy()[:] = x
    rp = torch.empty(n, dtype=torch.float64, pin_memory=True)

    def timed(label, xptr, rptr, env, reps=2, devices=1):
        old = {k: os.environ.get(k) for k in env}
        os.environ.update(env)
        b.set_host_devices(devices)
        b._lib.baobzi_eval_multi(b.ptr, xptr, rptr, n)
        t0 = time.perf_counter()
        for _ in range(reps):
            b._lib.baobzi_eval_multi(b.ptr, xptr, rptr, n)
        dt = (time.perf_counter() - t0) / reps
        for k, v in old.items():
            os.environ.pop(k, None) if v is None else os.environ.__setitem__(k, v)
        print("%-44s %.3e evals/s  (%.1f GB/s host->device)" % (label, n / dt, n * 24 / dt / 1e9), flush=True)

    timed("pinned, 3 workers", xp.data_ptr(), rp.data_ptr(), {})
    timed("pinned, 6 workers", xp.data_ptr(), rp.data_ptr(), {"BAOBZI_B200_HOST_WORKERS": "6"})
    ref = rp.numpy().copy()
    for w in (2, 4, 8, 12, 16):
        timed("pageable, staged, %d workers" % w, x.ctypes.data, res.ctypes.data, {"BAOBZI_B200_HOST_WORKERS": str(w)})
    assert np.array_equal(np.nan_to_num(res), np.nan_to_num(ref))
    timed("pageable, cudaHostRegister per call", x.ctypes.data, res.ctypes.data, {"BAOBZI_B200_HOSTREG": "1"})
    timed("pageable, driver staging (round 1)", x.ctypes.data, res.ctypes.data, {"BAOBZI_B200_STAGE": "0"}, reps=1)
    nd = torch.cuda.device_count()
    if nd > 1:
        timed("pinned, all %d GPUs, 3 workers each" % nd, xp.data_ptr(), rp.data_ptr(), {}, devices=0)
        assert np.array_equal(np.nan_to_num(rp.numpy()), np.nan_to_num(ref))
        timed("pageable, all %d GPUs" % nd, x.ctypes.data, res.ctypes.data, {}, devices=0)
    print("cores", os.cpu_count())
    b.free()


if __name__ == "__main__":
    main()
