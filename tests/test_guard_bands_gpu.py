"""Memory check of every kernel path without compute-sanitizer (closed on the pool this was developed on):
scripts/sanitize_paths.py -- direct, table / two-leaf / sorted 1D, binned 2D / 3D with the single-launch, split and
four-kernel sorts over several chunks, the piggyback schedule, bucketed, dot, node index, multi-output, scalar, staged
and multi-worker host calls, odd-aligned 2D input, two caller streams -- run with BAOBZI_B200_GUARD=1: every device buffer
of the evaluation path sits between two 4 KB bands of a known pattern, and no byte of any band may change."""
import os
import re
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("piggy", ["0", "1"])
def test_no_out_of_bounds_writes_next_to_any_buffer(gpu_lib, piggy):
    env = dict(os.environ, BAOBZI_B200_GUARD="1", BAOBZI_B200_PIGGY=piggy)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "sanitize_paths.py")], env=env, cwd=ROOT,
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=1200)
    assert r.returncode == 0 and "sanitize_paths: done" in r.stdout, r.stdout[-3000:]
    m = re.search(r"guard bands: (\d+) checks, up to (\d+) buffers watched at once, (\d+) bytes changed", r.stdout)
    assert m, r.stdout[-1000:]
    assert int(m.group(1)) >= 20 and int(m.group(2)) >= 8 and int(m.group(3)) == 0, m.group(0)
    out = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(out):
        with open(os.path.join(out, "guard_bands_piggy%s.txt" % piggy), "w") as f:
            f.write(r.stdout[-2500:])
