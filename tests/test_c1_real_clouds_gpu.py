"""Config C1 on the reference's real clouds, UN-REDUCED (tests/golden/real_clouds_full.npz: the xyz of all seven
data/real_objects/object/*.pcd files, 15 009 - 72 641 points): BPS bit-exact against the stored data/real_objects/bps/*.npy for
every kernel variant, then the dexgangrasp_offline.py-style sequence (encode -> generate 100 -> evaluate + filter) through the
class surface against the oracle.  VERDICT r01 item 1(e) / row g2."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from dex_gan_grasp_b200 import bps as bps_mod
from dex_gan_grasp_b200 import synth, tolerances
from oracle import nets_oracle as no

pytestmark = pytest.mark.gpu
NSPLIT = lambda n: (n & 0xff) << 8
SCALAR, DIRECT = 1, 2
CFG = dict(latentD=5, n_hand_joints=12, is_train=False, gpu_ids=[0], model="dexgangrasp", is_wgan=False)


@pytest.fixture(scope="module")
def real_full():
    return np.load(os.path.join(GOLDEN, "real_clouds_full.npz"))


@pytest.mark.parametrize("flags", [0, SCALAR, NSPLIT(3), NSPLIT(37), DIRECT])
def test_full_real_clouds_bit_exact(real_full, bps_golden, flags):
    basis = torch.from_numpy(bps_golden["basis"]).cuda()
    names = list(real_full["names"])
    clouds = [real_full[f"{n}/cloud"] for n in names]
    assert sorted(len(c) for c in clouds)[0] == 15009 and max(len(c) for c in clouds) == 72641
    off = np.concatenate([[0], np.cumsum([len(c) for c in clouds])]).astype(np.int64)
    d, i = bps_mod.bps_encode_device(basis, torch.from_numpy(np.concatenate(clouds)).cuda(), torch.from_numpy(off).cuda(),
                                     max_cloud_points=72641, flags=flags)
    for b, n in enumerate(names):
        assert np.array_equal(i[b].cpu().numpy(), real_full[f"{n}/idx"]), n
        assert np.array_equal(d[b].cpu().numpy(), real_full[f"{n}/bps"][0]), n  # == data/real_objects/bps/<n>.npy
    # one cloud alone (B = 1: the online shape, dexgangrasp_real_robot.py:139)
    for n in ("red_mug", "cheez_it"):
        d1, i1 = bps_mod.bps_encode_device(basis, torch.from_numpy(real_full[f"{n}/cloud"]).cuda().unsqueeze(0), flags=flags)
        assert np.array_equal(d1[0].cpu().numpy(), real_full[f"{n}/bps"][0]) and np.array_equal(i1[0].cpu().numpy(), real_full[f"{n}/idx"])


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_c1_offline_sequence_on_full_clouds(real_full, precision):
    """dexgangrasp_offline.py:130-162 per object: BPSEncoder.encode_pcd_with_bps -> generate_grasps(100) -> evaluate_grasps ->
    filter_grasps(0.5), compared with the oracle on the same z."""
    from dex_gan_grasp_b200.dexgangrasp import DexGanGrasp
    sd_gan, sd_eva = synth.synth_gan_state_dict(7), synth.synth_eva_state_dict(7)
    m = DexGanGrasp(dict(CFG, precision=precision))
    m.DexGANGrasp.load_state_dict(sd_gan)
    m.DexEvaluator.load_state_dict(sd_eva)
    enc = bps_mod.BPSEncoder(dict(CFG))
    ptol = tolerances.FP32_TOL if precision == "fp32" else tolerances.BF16_POSE_TOL
    stol = tolerances.FP32_TOL if precision == "fp32" else tolerances.BF16_SCORE_TOL
    for n in ("red_mug", "cheez_it", "soft_scrub"):
        bps = enc.encode_pcd_with_bps(real_full[f"{n}/cloud"].astype(np.float64))
        assert np.array_equal(bps, real_full[f"{n}/bps"][0])
        torch.manual_seed(5)
        g = m.generate_grasps(bps, n_samples=100, return_arr=True)
        tiled = torch.from_numpy(np.tile(bps, (100, 1)))
        o = no.generate_poses(sd_gan, torch.from_numpy(g["z"]), tiled, 0.025)
        assert np.abs(g["transl"] - o["transl"]).max() < ptol and np.abs(g["joint_conf"] - o["joint_conf"]).max() < ptol
        # scores of OUR grasps by our evaluator vs the oracle evaluator on the same grasps, then the same kept set
        p = m.evaluate_grasps(bps, {k: g[k] for k in ("rot_matrix", "transl", "joint_conf", "z")})
        with torch.no_grad():
            po = no.evaluator_forward(sd_eva, tiled, torch.from_numpy(g["rot_matrix"]), torch.from_numpy(g["transl"])).squeeze(1).numpy()
        assert np.abs(p - po).max() < stol, (n, float(np.abs(p - po).max()))
        thr = float(np.median(po))
        f = m.filter_grasps(bps, {k: g[k] for k in ("rot_matrix", "transl", "joint_conf", "z")}, thresh=thr)
        kept = {int(np.where((g["z"] == row).all(1))[0][0]) for row in f["z"]}
        assert {i for i in range(100) if po[i] > thr + stol} <= kept
        assert not (kept & {i for i in range(100) if po[i] < thr - stol})
        assert f["bps_object"].shape == (len(kept), 4096)
