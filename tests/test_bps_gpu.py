"""GPU parity: the sm_100a BPS kernel (through the C ABI) against the CPU oracle and the reference's fixtures."""
import numpy as np
import pytest
import torch

from conftest import oracle_bps
from dex_gan_grasp_b200 import _native, bps as bps_mod, synth

pytestmark = pytest.mark.gpu

NSPLIT = lambda n: (n & 0xff) << 8
SCALAR = 1
DIRECT = 2  # direct-difference scan instead of the default expanded-form filter + exact candidates
FMA = 4     # the filter scan on the FP32 pipe instead of the default tensor-core scan


def dev_encode(basis, clouds, flags=0):
    """clouds: list of [n,3] f32 -> (dists [B,P], idx [B,P]) numpy via the device entry point (ragged)."""
    sizes = [len(c) for c in clouds]
    pts = np.ascontiguousarray(np.concatenate(clouds).astype(np.float32)) if sum(sizes) else np.zeros((0, 3), np.float32)
    off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
    t_pts = torch.from_numpy(pts).cuda() if len(pts) else torch.zeros((4, 3), device="cuda")[:0]
    d, i = bps_mod.bps_encode_device(torch.from_numpy(np.ascontiguousarray(basis, np.float32)).cuda(), t_pts,
                                     torch.from_numpy(off).cuda(), max_cloud_points=max(sizes + [0]), flags=flags)
    torch.cuda.synchronize()
    return d.cpu().numpy(), i.cpu().numpy()


@pytest.mark.parametrize("flags", [0, SCALAR, NSPLIT(3), DIRECT, DIRECT | NSPLIT(2), FMA, FMA | NSPLIT(3)])
def test_reference_fixtures_bit_exact(bps_golden, flags):
    basis = bps_golden["basis"]
    names = list(bps_golden["names"])
    clouds = [bps_golden[f"{n}/cloud"] for n in names]
    d, i = dev_encode(basis, clouds, flags)
    for b, n in enumerate(names):
        assert np.array_equal(i[b], bps_golden[f"{n}/idx"]), n
        assert np.array_equal(d[b], bps_golden[f"{n}/bps"][0]), n  # bit-exact with data/real_objects/bps/<n>.npy


def test_encoder_class_surface(bps_golden):
    enc = bps_mod.BPSEncoder()
    out = enc.encode_pcd_with_bps(bps_golden["red_mug/cloud"].astype(np.float64))
    assert out.shape == (4096,) and out.dtype == np.float32
    assert np.array_equal(out, bps_golden["red_mug/bps"][0])
    with pytest.raises(AssertionError):
        enc.encode_pcd_with_bps(bps_golden["red_mug/cloud"].astype(np.float64) + 5.0)  # not centred (bps_encoder.py:50)
    shim = bps_mod.bps_torch(custom_basis=bps_golden["basis"], device="cuda")
    r = shim.encode(torch.from_numpy(bps_golden["red_mug/cloud"]), feature_type=["dists", "ids", "deltas"])
    assert r["dists"].shape == (1, 4096) and r["ids"].dtype == torch.int64
    assert torch.allclose(r["deltas"].norm(dim=2), r["dists"], atol=1e-7)


@pytest.mark.parametrize("flags", [0, NSPLIT(2), NSPLIT(5), SCALAR, DIRECT, DIRECT | NSPLIT(5), FMA, FMA | NSPLIT(5)])
def test_ragged_random_vs_oracle(oracle_lib, bps_golden, flags):
    rng = np.random.default_rng(3)
    basis = bps_golden["basis"]
    sizes = [1, 2, 3, 5, 31, 32, 33, 0, 1023, 1024, 1025, 2048, 4099, 7, 64]
    clouds = [rng.normal(0, 0.07, (n, 3)).astype(np.float32) for n in sizes]
    d, i = dev_encode(basis, clouds, flags)
    for b, c in enumerate(clouds):
        do, io = oracle_bps(oracle_lib, basis, c)
        assert np.array_equal(i[b], io), (b, sizes[b])
        assert np.array_equal(d[b], do), (b, sizes[b])


def test_ties_duplicates_lowest_index(oracle_lib):
    rng = np.random.default_rng(5)
    basis = rng.uniform(-0.2, 0.2, (300, 3)).astype(np.float32)  # P not a multiple of anything
    base = (rng.integers(-8, 8, (40, 3)) / 64.0).astype(np.float32)  # coarse grid -> many exact ties
    cloud = base[rng.integers(0, 40, 3000)]
    for flags in (0, NSPLIT(4), SCALAR, DIRECT, FMA):
        d, i = dev_encode(basis, [cloud, cloud[::-1].copy()], flags)
        for b, c in enumerate([cloud, cloud[::-1].copy()]):
            do, io = oracle_bps(oracle_lib, basis, c)
            assert np.array_equal(i[b], io) and np.array_equal(d[b], do)


def test_filter_near_ties_and_off_centre_clouds(oracle_lib, bps_golden):
    """Stress of the default kernel's filter (expanded form |x|^2 - 2 b.x, then exact evaluation of the candidates):
    (a) thousands of points at almost the same distance from the basis points (a thin spherical shell around the basis
    centroid, radial jitter at the 1e-7 level) - more than two chunks fall inside the error margin, which takes the exact
    fallback scan; (b) clouds far from the origin (the margin grows with the coordinate magnitude); (c) a cloud holding the
    basis points themselves (zero distances).  Indices and distances must stay bit-exact with the oracle."""
    rng = np.random.default_rng(11)
    basis = bps_golden["basis"][:512]
    u = rng.normal(size=(6000, 3))
    u /= np.linalg.norm(u, axis=1, keepdims=True)
    shell = (basis.mean(0) + 0.05 * u * (1.0 + 1e-7 * rng.normal(size=(6000, 1)))).astype(np.float32)
    tiny = (basis[7] + 0.01 * u[:3000] * (1.0 + 3e-8 * rng.normal(size=(3000, 1)))).astype(np.float32)
    far = (rng.normal(0, 0.05, (5000, 3)) + np.array([4.0, -7.0, 2.5])).astype(np.float32)
    huge = (rng.normal(0, 30.0, (3000, 3))).astype(np.float32)
    onb = np.concatenate([basis[::3], rng.normal(0, 0.1, (1000, 3)).astype(np.float32)])
    clouds = [shell, tiny, far, huge, onb]
    for flags in (0, NSPLIT(3), FMA):
        d, i = dev_encode(basis, clouds, flags)
        for b, c in enumerate(clouds):
            do, io = oracle_bps(oracle_lib, basis, c)
            assert np.array_equal(i[b], io), (flags, b)
            assert np.array_equal(d[b], do), (flags, b)


def test_shared_memory_finish_small_ranges(oracle_lib, bps_golden):
    """Ranges of at most 2048 points (the C3 shape) finish their candidates from the shared-memory stages, one thread per
    basis point in a rotated visiting order: exact ties (coarse grid, duplicates), a thin shell (more than two candidate
    chunks -> whole-range fallback in that path), an off-centre cloud, sizes around the tile and chunk boundaries and
    unaligned cloud starts (ragged offsets), all bit-exact with the oracle."""
    rng = np.random.default_rng(23)
    basis = bps_golden["basis"][:640]
    grid = (rng.integers(-6, 6, (30, 3)) / 32.0).astype(np.float32)
    u = rng.normal(size=(2048, 3))
    u /= np.linalg.norm(u, axis=1, keepdims=True)
    shell = (basis.mean(0) + 0.05 * u * (1.0 + 1e-7 * rng.normal(size=(2048, 1)))).astype(np.float32)
    far = (rng.normal(0, 0.05, (1500, 3)) + np.array([4.0, -7.0, 2.5])).astype(np.float32)
    clouds = [grid[rng.integers(0, 30, 2047)], shell, far, rng.normal(0, 0.08, (1, 3)).astype(np.float32),
              rng.normal(0, 0.08, (33, 3)).astype(np.float32), rng.normal(0, 0.08, (1023, 3)).astype(np.float32),
              rng.normal(0, 0.08, (1025, 3)).astype(np.float32), rng.normal(0, 0.08, (2048, 3)).astype(np.float32)]
    for flags in (0, NSPLIT(2), FMA):
        d, i = dev_encode(basis, clouds, flags)
        for b, c in enumerate(clouds):
            do, io = oracle_bps(oracle_lib, basis, c)
            assert np.array_equal(i[b], io), (flags, b)
            assert np.array_equal(d[b], do), (flags, b)


def test_uniform_batch_and_large_cloud(oracle_lib, bps_golden):
    basis = bps_golden["basis"]
    clouds = synth.synth_clouds(6, 2048)
    d, i = bps_mod.bps_encode_device(torch.from_numpy(basis).cuda(), torch.from_numpy(clouds).cuda())
    big = synth.synth_cloud(100000, 5)
    db, ib = bps_mod.bps_encode_device(torch.from_numpy(basis).cuda(), torch.from_numpy(big[None]).cuda())
    torch.cuda.synchronize()
    for b in range(6):
        do, io = oracle_bps(oracle_lib, basis, clouds[b])
        assert np.array_equal(i[b].cpu().numpy(), io) and np.array_equal(d[b].cpu().numpy(), do)
    do, io = oracle_bps(oracle_lib, basis, big)
    assert np.array_equal(ib[0].cpu().numpy(), io) and np.array_equal(db[0].cpu().numpy(), do)


def test_full_size_properties():
    """C3 size (1024 clouds x 2048 points): size-independent checks — permutation invariance of the distances,
    idempotence, d == |x[idx] - basis|, and agreement of the packed and the scalar kernels."""
    basis = torch.from_numpy(np.load(bps_mod.BASIS_PATH)).cuda()
    g = torch.Generator(device="cuda").manual_seed(0)
    pts = 0.08 * torch.randn((1024, 2048, 3), device="cuda", generator=g)
    d, i = bps_mod.bps_encode_device(basis, pts)
    d2, i2 = bps_mod.bps_encode_device(basis, pts, flags=SCALAR)
    assert torch.equal(d, d2) and torch.equal(i, i2)
    d4, i4 = bps_mod.bps_encode_device(basis, pts, flags=DIRECT)
    assert torch.equal(d, d4) and torch.equal(i, i4)
    perm = torch.randperm(2048, device="cuda", generator=g)
    d3, i3 = bps_mod.bps_encode_device(basis, pts[:, perm].contiguous())
    assert torch.equal(d, d3)  # no exact ties in continuous random data -> same winner, same distance
    assert torch.equal(perm[i3.long()], i.long())
    win = torch.gather(pts, 1, i.long().unsqueeze(-1).expand(-1, -1, 3))
    assert torch.allclose((win - basis[None]).norm(dim=2), d, atol=1e-7)


def test_host_entry_point_and_errors(bps_golden):
    L = _native.lib()
    basis = np.ascontiguousarray(bps_golden["basis"])
    cloud = np.ascontiguousarray(bps_golden["soft_scrub/cloud"])
    off = np.array([0, len(cloud)], np.int64)
    d = np.empty((1, 4096), np.float32)
    i = np.empty((1, 4096), np.int32)
    _native.check(L.dexgg_bps_encode_host(_native.ptr(basis), 4096, _native.ptr(cloud), _native.ptr(off), 1, _native.ptr(d),
                                          _native.ptr(i), 0))
    assert np.array_equal(d, bps_golden["soft_scrub/bps"])
    rc = L.dexgg_bps_encode(None, 4096, None, None, 1, 10, 10, None, None, 0, None, 0, None)
    assert rc == -1 and b"null" in L.dexgg_last_error()
    t = torch.zeros((1, 8, 3), device="cuda")
    rc = L.dexgg_bps_encode(t.data_ptr(), 2, t.data_ptr(), None, 1, 8, 8, t.data_ptr(), None, 0, None, 0, None)
    assert rc == -3  # workspace too small
