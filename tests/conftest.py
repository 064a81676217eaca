import glob
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def _make(target_dir, *targets):
    r = subprocess.run(["make", "-C", target_dir, *targets], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError("make %s failed:\n%s" % (" ".join(targets), r.stdout))


@pytest.fixture(scope="session")
def ref():
    """oracle/ref.py with the C restatement + test callbacks built (and the reference, when its sources exist)."""
    _make(os.path.join(ROOT, "oracle"), "oracle")
    from oracle import ref as _ref
    return _ref


@pytest.fixture(scope="session")
def have_reference(ref):
    return ref.have_ref() and os.path.exists(ref.REFIDX_SO)


def golden_cases():
    return sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "*.npz")))


def load_golden(name):
    return os.path.join(GOLDEN, name + ".baobzi"), dict(np.load(os.path.join(GOLDEN, name + ".npz")))


@pytest.fixture(scope="session")
def hostlib():
    """Host-only helper library: the product's .baobzi reader/writer and fitter behind a tiny C surface
    (tests/host_helpers.cpp) so they can be exercised without a GPU."""
    import ctypes as C
    out = os.path.join(ROOT, "tests", "_build")
    os.makedirs(out, exist_ok=True)
    so = os.path.join(out, "libhosthelpers.so")
    csrc = os.path.join(ROOT, "baobzi_b200", "csrc")
    srcs = [os.path.join(ROOT, "tests", "host_helpers.cpp"), os.path.join(csrc, "msgpack_io.cpp"),
            os.path.join(csrc, "fit.cpp")]
    deps = srcs + glob.glob(os.path.join(csrc, "*.hpp")) + [os.path.join(ROOT, "include", "baobzi.h")]
    if not os.path.exists(so) or any(os.path.getmtime(d) > os.path.getmtime(so) for d in deps):
        cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
        cmd = [cxx, "-O2", "-std=c++17", "-fPIC", "-shared", "-ffp-contract=off", "-I", csrc, "-o", so] + srcs
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if r.returncode != 0:
            raise RuntimeError("building host helpers failed:\n" + r.stdout)
    lib = C.CDLL(so)
    lib.bzt_roundtrip.restype = C.c_int
    lib.bzt_roundtrip.argtypes = [C.c_char_p, C.c_char_p]
    lib.bzt_fit_save.restype = C.c_int
    lib.bzt_fit_save.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_char_p]
    lib.bzt_read_status.restype = C.c_int
    lib.bzt_read_status.argtypes = [C.c_char_p, C.c_char_p, C.c_int]
    return lib


@pytest.fixture(scope="session")
def gpu_lib():
    """The product library; GPU tests must run the CUDA path, so a missing library is an error, not a skip."""
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    import baobzi_b200
    if not os.path.exists(baobzi_b200.library_path()):
        baobzi_b200.build()
    return baobzi_b200
