"""dexgg_center_clouds (the `obj_pcd.translate(-obj_pcd.get_center())` in front of the encoder, dexgangrasp_real_robot.py:133-134)
against oracle/center_oracle.py: centres to 1e-15 relative (float64 sums in another order), centred float32 points identical except
where the float64 difference sits within that noise of a float32 rounding boundary (counted, < 1e-5 of the values, 1 ulp)."""
import numpy as np
import pytest
import torch

from oracle import center_oracle as co


def test_oracle_on_the_reference_fixtures():
    """The shipped real-object clouds were stored after the reference's centring: their float64 mean is ~1e-3 m (sanity of the restatement, not a pin)."""
    import os
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "real_clouds_full.npz"))
    worst = 0.0
    for n in z["names"]:
        pts = z[f"{n}/cloud"]
        out, c = co.center_cloud(pts)
        worst = max(worst, float(np.abs(c).max()))
        assert np.abs(out.astype(np.float64) - (pts.astype(np.float64) - c)).max() < 1e-8
    print("largest |centre| of the shipped clouds:", worst)
    assert worst < 5e-3, worst  # (centred before the last outlier removal: the mean of what is left is ~1e-3, not 0)


@pytest.mark.gpu
def test_center_clouds_vs_oracle_ragged_and_uniform():
    from dex_gan_grasp_b200 import bps as bps_mod
    rng = np.random.default_rng(3)
    sizes = [1, 7, 1000, 2048, 38663, 100000]
    clouds = [(rng.normal(0, 0.08, (n, 3)) + rng.normal(0, 0.6, 3)).astype(np.float32) for n in sizes]
    off = np.zeros(len(sizes) + 1, dtype=np.int64)
    off[1:] = np.cumsum(sizes)
    pts = torch.from_numpy(np.concatenate(clouds)).cuda()
    out, cen = bps_mod.center_clouds_device(pts, torch.from_numpy(off).cuda(), max_cloud_points=max(sizes))
    torch.cuda.synchronize()
    out, cen = out.cpu().numpy(), cen.cpu().numpy()
    diff_values = 0
    for b, c in enumerate(clouds):
        ro, rc = co.center_cloud(c)
        assert np.allclose(cen[b], rc, rtol=0, atol=1e-15 * max(1.0, np.abs(c).max()) * 8), (b, cen[b], rc)
        got = out[off[b]:off[b + 1]]
        neq = got != ro
        diff_values += int(neq.sum())
        if neq.any():  # a float64 value within summation noise of a float32 rounding boundary: one ulp apart at most
            assert np.abs(got[neq] - ro[neq]).max() <= np.spacing(np.abs(ro[neq]).max())
    assert diff_values <= 1e-5 * 3 * sum(sizes), diff_values
    # uniform layout, in place, and the encoder on the centred clouds equals the encoder on oracle-centred clouds
    u = (rng.normal(0, 0.05, (5, 2048, 3)) + np.array([0.4, -0.2, 0.7])).astype(np.float32)
    t = torch.from_numpy(u).cuda()
    o2, c2 = bps_mod.center_clouds_device(t, out=t)
    assert o2.data_ptr() == t.data_ptr()
    ref = np.stack([co.center_cloud(x)[0] for x in u])
    basis = torch.from_numpy(np.load(bps_mod.BASIS_PATH)).cuda()
    d_a, i_a = bps_mod.bps_encode_device(basis, t)
    d_b, i_b = bps_mod.bps_encode_device(basis, torch.from_numpy(ref).cuda())
    same = (t.cpu().numpy() == ref).all()
    assert (not same) or (torch.equal(d_a, d_b) and torch.equal(i_a, i_b))
    assert np.abs(t.cpu().numpy().astype(np.float64).mean(1)).max() < 1e-8
