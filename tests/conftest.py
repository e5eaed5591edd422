import ctypes
import os
import subprocess
import sys

import numpy as np
import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if REPO not in sys.path:
    sys.path.insert(0, REPO)
GOLDEN = os.path.join(REPO, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle_lib():
    """The C restatement of the BPS stage (oracle/bps_oracle.c) — TEST INFRASTRUCTURE, built on demand."""
    so = os.path.join(REPO, "oracle", "libdexgg_oracle.so")
    if not os.path.exists(so):
        subprocess.check_call(["make", "-C", os.path.join(REPO, "oracle")])
    return ctypes.CDLL(so)


@pytest.fixture(scope="session")
def bps_golden():
    return np.load(os.path.join(GOLDEN, "bps_real_objects.npz"))


@pytest.fixture(scope="session")
def eval_batch():
    return np.load(os.path.join(GOLDEN, "eval_batch.npz"))


@pytest.fixture(scope="session")
def nets_golden():
    return np.load(os.path.join(GOLDEN, "nets_golden.npz"))


def vp(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def oracle_bps(lib, basis, pts, nthreads=8):
    basis = np.ascontiguousarray(basis, np.float32)
    pts = np.ascontiguousarray(pts, np.float32)
    P = basis.shape[0]
    d = np.empty(P, np.float32)
    idx = np.empty(P, np.int32)
    lib.dexgg_oracle_bps_encode_mt(vp(basis), P, vp(pts), len(pts), vp(d), vp(idx), nthreads)
    return d, idx
