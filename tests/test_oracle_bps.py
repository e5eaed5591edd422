"""CPU: the BPS oracle (oracle/bps_oracle.c) against the reference's own fixtures
(data/real_objects/object/*.pcd -> data/real_objects/bps/*.npy, committed reduced under tests/golden/)."""
import ctypes

import numpy as np

from conftest import oracle_bps, vp


def test_fixtures_bit_exact(oracle_lib, bps_golden):
    basis = bps_golden["basis"]
    for name in bps_golden["names"]:
        d, idx = oracle_bps(oracle_lib, basis, bps_golden[f"{name}/cloud"])
        assert np.array_equal(d, bps_golden[f"{name}/bps"][0]), name  # bit-exact
        assert np.array_equal(idx, bps_golden[f"{name}/idx"]), name
        assert d.min() < 0.1  # bps_encoder.py:50


def test_only_one_summation_order_reproduces_fixtures(oracle_lib, bps_golden):
    basis = bps_golden["basis"]
    n_diff_xy_z = n_diff_fma = 0
    for name in bps_golden["names"]:
        cloud = np.ascontiguousarray(bps_golden[f"{name}/cloud"])
        idx = np.ascontiguousarray(bps_golden[f"{name}/idx"])
        a, b, c = (np.empty(4096, np.float32) for _ in range(3))
        oracle_lib.dexgg_oracle_bps_dist_variants(vp(basis), 4096, vp(cloud), vp(idx), vp(a), vp(b), vp(c))
        gold = bps_golden[f"{name}/bps"][0]
        assert np.array_equal(a, gold)
        n_diff_xy_z += int((b != gold).sum())
        n_diff_fma += int((c != gold).sum())
        assert np.max(np.abs(b - gold)) <= 1.5e-8 and np.max(np.abs(c - gold)) <= 1.5e-8  # 1 ulp
    assert n_diff_xy_z > 0 and n_diff_fma > 0


def test_indices_match_float64_brute_force(oracle_lib, bps_golden):
    basis = bps_golden["basis"].astype(np.float64)
    cloud = bps_golden["red_mug/cloud"].astype(np.float64)
    d2 = ((basis[:, None, :] - cloud[None, :, :]) ** 2).sum(-1)
    assert np.array_equal(d2.argmin(1).astype(np.int32), bps_golden["red_mug/idx"])


def test_ties_lowest_index_and_edges(oracle_lib):
    basis = np.array([[0, 0, 0], [1, 0, 0]], np.float32)
    pts = np.array([[0.5, 0, 0], [0.5, 0, 0], [-0.5, 0, 0]], np.float32)  # exact ties everywhere
    d, idx = oracle_bps(oracle_lib, basis, pts, 1)
    assert idx.tolist() == [0, 0] and d.tolist() == [0.5, 0.5]
    d, idx = oracle_bps(oracle_lib, basis, np.zeros((0, 3), np.float32), 1)  # empty cloud
    assert np.all(np.isinf(d)) and idx.tolist() == [-1, -1]
    d, idx = oracle_bps(oracle_lib, basis, np.array([[1, 0, 0]], np.float32), 3)  # single point, more threads than work
    assert idx.tolist() == [0, 0] and d.tolist() == [1.0, 0.0]


def test_ragged_batch_equals_per_cloud(oracle_lib, bps_golden):
    basis = np.ascontiguousarray(bps_golden["basis"][:257])
    rng = np.random.default_rng(1)
    sizes = [0, 1, 33, 700, 5]
    clouds = [rng.normal(0, 0.08, (n, 3)).astype(np.float32) for n in sizes]
    pts = np.ascontiguousarray(np.concatenate(clouds))
    off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
    d = np.empty((len(sizes), 257), np.float32)
    idx = np.empty((len(sizes), 257), np.int32)
    oracle_lib.dexgg_oracle_bps_encode_batch(vp(basis), 257, vp(pts), vp(off), len(sizes), vp(d), vp(idx), 4)
    for b, c in enumerate(clouds):
        d1, i1 = oracle_bps(oracle_lib, basis, c, 1)
        assert np.array_equal(d[b], d1) and np.array_equal(idx[b], i1)
