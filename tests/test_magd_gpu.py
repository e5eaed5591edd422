"""MAGD on the device (csrc/magd.cu through the C ABI) against the oracle and the reference's golden vectors, and the C2
check of SURVEY.md §8c: the MAGD of grasp sets generated on data/eval_batch.pth by the CUDA path and by the oracle must
agree within 1e-3 (north_star tolerance)."""
import os

import numpy as np
import pytest
import torch

from dex_gan_grasp_b200 import synth
from oracle import magd_oracle, nets_oracle as no
from test_oracle_magd import cases

pytestmark = pytest.mark.gpu
TOL = 1e-3  # north_star: "MAGD on data/eval_batch.pth unchanged within 1e-3"


@pytest.fixture(scope="module")
def magd_golden():
    from conftest import GOLDEN
    return np.load(os.path.join(GOLDEN, "magd_golden.npz"))


def test_device_magd_matches_reference_golden(magd_golden):
    from dex_gan_grasp_b200 import magd
    for i, pred, gt, want in cases(magd_golden):
        got = magd.magd_for_grasp_distribution(pred, gt)
        for a, b in zip(got[:3], want[:3]):
            assert abs(a - b) <= TOL * max(1.0, abs(b)), (i, got, want)
        assert got[3] == pytest.approx(want[3], abs=1e-12), (i, got, want)
        _, match = magd.magd_device(pred, gt, return_match=True)
        ref_match = magd_oracle.magd(pred, gt, return_match=True)[4]
        assert np.array_equal(match.cpu().numpy(), ref_match), i


def test_large_sets_vs_oracle():
    from dex_gan_grasp_b200 import magd
    rng = np.random.default_rng(3)
    n, m = 20000, 3000
    q, _ = np.linalg.qr(rng.standard_normal((n + m, 3, 3)))
    pred = {"transl": (0.1 * rng.standard_normal((n, 3))).astype(np.float32), "rot_matrix": q[:n].astype(np.float32),
            "joint_conf": rng.uniform(0, 0.5, (n, 12)).astype(np.float32)}
    gt = {"transl": (0.1 * rng.standard_normal((m, 3))).astype(np.float32), "rot_matrix": q[n:].astype(np.float32),
          "joint_conf": rng.uniform(0, 0.5, (m, 12)).astype(np.float32)}
    got = magd.magd_for_grasp_distribution(pred, gt)
    want = magd_oracle.magd({k: v.astype(np.float64) for k, v in pred.items()}, {k: v.astype(np.float64) for k, v in gt.items()})
    for a, b in zip(got, want):
        assert abs(a - b) <= 1e-5 * max(1.0, abs(b)), (got, want)


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_c2_magd_on_eval_batch_unchanged(eval_batch, precision):
    """C2: generator forward on the 64 BPS rows of eval_batch with shared z; MAGD of the generated set against the batch's
    own ground-truth grasps, per object (grouped by obj_name like eval.py:309-369) and over the whole batch."""
    from dex_gan_grasp_b200 import magd
    from dex_gan_grasp_b200.dexgangrasp import DexGanGrasp
    cfg = dict(latentD=5, n_hand_joints=12, is_train=False, gpu_ids=[0], model="dexgangrasp", is_wgan=False, precision=precision)
    model = DexGanGrasp(cfg)
    sd = synth.synth_gan_state_dict(7)
    model.DexGANGrasp.load_state_dict(sd)
    bps = torch.from_numpy(eval_batch["bps_object"])
    z = synth.synth_latents(64, 5, seed=21)
    ref = no.generate_poses(sd, z, bps, 0.025)
    rot, transl, joint, _ = model.DexGANGrasp.generator.native_forward(z.cuda(), bps.cuda(), 0.025, True)
    ours = {"rot_matrix": rot.reshape(-1, 3, 3), "transl": transl, "joint_conf": joint}
    gt_all = {k: eval_batch[k] for k in ("transl", "rot_matrix", "joint_conf")}
    names = eval_batch["obj_name"]
    groups = [np.arange(64)] + [np.nonzero(names == n)[0] for n in sorted(set(names.tolist()))]
    tol = TOL if precision == "fp32" else 2e-2  # bf16 tensor-core mode: hidden activations carry bf16 rounding
    for idx in groups:
        gt = {k: v[idx] for k, v in gt_all.items()}
        want = magd_oracle.magd({k: np.asarray(ref[k])[idx] for k in gt}, gt)
        got = magd.magd_for_grasp_distribution({k: v[torch.from_numpy(idx).cuda()] for k, v in ours.items()}, gt)
        for a, b in zip(got[:3], want[:3]):
            assert abs(a - b) <= tol * max(1.0, abs(b)), (precision, len(idx), got, want)
