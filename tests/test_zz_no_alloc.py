"""dexgg_bps_encode_static: the BPS encoder with its scratch in a static device array inside the library (no workspace argument, no
allocation in the call; include/dexgg.h).  Called through the C ABI and compared with the existing entry point dexgg_bps_encode on
the same inputs - bit-exact, the tolerance of every BPS test in this repository - on two calls in a row (the second with other
clouds of another shape, so stale contents of the shared scratch would show), then the first inputs again; and the documented limit
is enforced with a clear error.  Collected last on purpose; passes alone in a fresh process too."""
import ctypes

import numpy as np
import pytest
import torch

from dex_gan_grasp_b200 import _native
from dex_gan_grasp_b200 import bps as bps_mod

pytestmark = pytest.mark.gpu


def _inputs(n_clouds, sizes_seed, scale):
    rng = np.random.default_rng(sizes_seed)
    sizes = [int(s) for s in rng.integers(1, 3000, n_clouds)]
    sizes[0] = 2048
    clouds = [(rng.normal(0, scale, (n, 3))).astype(np.float32) for n in sizes]
    off = np.zeros(n_clouds + 1, dtype=np.int64)
    off[1:] = np.cumsum(sizes)
    pts = torch.from_numpy(np.concatenate(clouds, 0)).cuda()
    return pts, torch.from_numpy(off).cuda(), max(sizes), int(off[-1])


def _existing(basis, pts, off, B, maxn, total, flags=0):
    L = _native.lib()
    P = basis.shape[0]
    d = torch.empty((B, P), dtype=torch.float32, device="cuda")
    i = torch.empty((B, P), dtype=torch.int32, device="cuda")
    nbytes = L.dexgg_bps_workspace_bytes(B, P)
    ws = torch.empty((nbytes,), dtype=torch.uint8, device="cuda")
    _native.check(L.dexgg_bps_encode(basis.data_ptr(), P, pts.data_ptr(), off.data_ptr(), B, maxn, total, d.data_ptr(), i.data_ptr(),
                                     flags, ws.data_ptr(), nbytes, None))
    torch.cuda.synchronize()
    return d, i


def _static(basis, pts, off, B, maxn, total, d, i, flags=0):
    """outputs are allocated by the caller BEFORE the call: the call itself gets only pointers"""
    L = _native.lib()
    rc = L.dexgg_bps_encode_static(basis.data_ptr(), basis.shape[0], pts.data_ptr(), off.data_ptr(), B, maxn, total, d.data_ptr(),
                                   i.data_ptr(), flags, None)
    _native.check(rc)
    torch.cuda.synchronize()


def test_static_scratch_entry_point_matches_existing_twice():
    basis = torch.from_numpy(np.load(bps_mod.BASIS_PATH)).cuda()
    P = basis.shape[0]
    assert _native.lib().dexgg_bps_encode_static_max_keys() >= 64 * P
    case_a = _inputs(5, 1, 0.08)
    case_b = _inputs(64, 2, 0.3)  # the maximum: 64 clouds x 4096 basis points
    for case in (case_a, case_b, case_a):
        pts, off, maxn, total = case
        B = off.numel() - 1
        d_ref, i_ref = _existing(basis, pts, off, B, maxn, total)
        d = torch.full((B, P), -1.0, dtype=torch.float32, device="cuda")
        i = torch.full((B, P), -7, dtype=torch.int32, device="cuda")
        before = torch.cuda.memory_allocated()
        _static(basis, pts, off, B, maxn, total, d, i)
        assert torch.cuda.memory_allocated() == before  # nothing taken from torch's allocator by the call
        assert torch.equal(d, d_ref) and torch.equal(i, i_ref)
    # every kernel variant works on the built-in scratch (FP32-pipe filter, direct scan, split ranges)
    pts, off, maxn, total = case_a
    d_ref, i_ref = _existing(basis, pts, off, 5, maxn, total)
    for flags in (4, 2, 3 << 8):
        d = torch.empty((5, P), dtype=torch.float32, device="cuda")
        i = torch.empty((5, P), dtype=torch.int32, device="cuda")
        _static(basis, pts, off, 5, maxn, total, d, i, flags)
        assert torch.equal(d, d_ref) and torch.equal(i, i_ref), flags


def test_static_scratch_limit_is_enforced():
    basis = torch.from_numpy(np.load(bps_mod.BASIS_PATH)).cuda()
    P = basis.shape[0]
    B = int(_native.lib().dexgg_bps_encode_static_max_keys()) // P + 1
    pts = torch.zeros((B * 4, 3), dtype=torch.float32, device="cuda")
    d = torch.empty((B, P), dtype=torch.float32, device="cuda")
    i = torch.empty((B, P), dtype=torch.int32, device="cuda")
    rc = _native.lib().dexgg_bps_encode_static(basis.data_ptr(), P, pts.data_ptr(), None, B, 4, B * 4, d.data_ptr(), i.data_ptr(), 0, None)
    assert rc == -1  # DEXGG_E_ARG
    msg = _native.lib().dexgg_last_error().decode()
    assert "exceeds the built-in scratch" in msg and "dexgg_bps_encode" in msg
    with pytest.raises(_native.DexggError):
        _native.check(rc)
