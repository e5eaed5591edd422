"""The benchmarked pipeline pinned AS A PIPELINE: GraspPipeline.run_device on the C3 shape (2048-point
synthetic clouds x 1000 latent samples -> per-object top-10 above 0.5) against the fp32 oracle pipeline
(oracle/pipeline_check.py: C BPS -> tiled generator -> rot6d / palm offset -> evaluator -> sort / threshold, the reference's
per-object loop of dexgangrasp_real_robot.py:139-150).  BPS must be bit-exact; the top-k set per object must satisfy the
filtered-set contract of tolerances.BF16_PIPELINE_EPS (bf16 mode) / FP32_TOL (fp32 modes), and the measured end-to-end
max |dp| is printed and bounded by the same epsilon."""
import numpy as np
import pytest
import torch

from dex_gan_grasp_b200 import bps as bps_mod
from dex_gan_grasp_b200 import synth, tolerances
from oracle import pipeline_check as pc

SEED = 7
CFG = dict(latentD=5, n_hand_joints=12, is_train=False, gpu_ids=[0], model="dexgangrasp", is_wgan=False)
N_OBJ, N_POINTS, S, K, THRESH = 16, 2048, 1000, 10, 0.5


def test_compare_topk_contract_cpu():
    """The set comparison itself (no GPU): exact set passes, a sure row dropped / a sure-out row kept is counted, rows inside
    the eps band may fall either way."""
    rng = np.random.default_rng(0)
    p = rng.random(3 * 50).astype(np.float32)
    idx, cnt = pc.no.segmented_topk(p, np.arange(4) * 50, 0.5, 5)
    assert pc.compare_topk(p, idx, cnt, 50, 0.5, 5, 1e-3)["set_mismatches"] == 0
    bad = idx.copy()
    worst = int(np.argmin(p[:50]))
    bad[0, 0] = worst  # best row of object 0 replaced by its worst row
    r = pc.compare_topk(p, bad, cnt, 50, 0.5, 5, 1e-3)
    assert r["missing"] == 1 and r["extra"] == 1
    # a swap inside the band is tolerated
    order = np.argsort(-p[:50])
    near = idx.copy()
    near[0, 4] = order[5]
    gap = float(p[order[4]] - p[order[5]])
    assert pc.compare_topk(p, near, cnt, 50, 0.5, 5, gap + 1e-6)["set_mismatches"] == 0
    assert pc.compare_topk(p, near, cnt, 50, 0.5, 5, gap * 0.25)["set_mismatches"] >= 1


def _run(precision):
    from dex_gan_grasp_b200.dexgangrasp import DexGanGrasp
    from dex_gan_grasp_b200.pipeline import GraspPipeline
    sd_gan, sd_eva = synth.synth_gan_state_dict(SEED), synth.synth_eva_state_dict(SEED)
    m = DexGanGrasp(dict(CFG, precision=precision))
    m.DexGANGrasp.load_state_dict(sd_gan)
    m.DexEvaluator.load_state_dict(sd_eva)
    pipe = GraspPipeline(m.DexGANGrasp, m.DexEvaluator, k=K, thresh=THRESH, z_offset=0.025)
    clouds = synth.synth_clouds(N_OBJ, N_POINTS, seed0=1000)
    z = torch.randn((N_OBJ * S, 5), generator=torch.Generator().manual_seed(1234))
    out = pipe.run_device(torch.from_numpy(clouds).cuda(), z.cuda(), S)
    torch.cuda.synchronize()
    ref = pc.oracle_pipeline(sd_gan, sd_eva, np.load(bps_mod.BASIS_PATH), clouds, z, S, THRESH, K)
    return out, ref


@pytest.mark.gpu
def test_bf16_pipeline_topk_set_vs_fp32_oracle():
    out, ref = _run("bf16")
    rec = pc.check_pipeline(out, ref, S, THRESH, K, tolerances.BF16_PIPELINE_EPS)
    print("bf16 pipeline vs fp32 oracle pipeline:", rec)
    assert rec["bps_bit_exact"]
    assert rec["max_dp"] < tolerances.BF16_PIPELINE_EPS, rec
    assert rec["set_mismatches"] == 0, rec
    assert rec["oracle_kept"] > 0 and rec["kept"] > 0
    # the gathered top-k rows are the generator's rows at those indices
    idx = out["idx"].cpu().numpy()
    sel = idx >= 0
    assert np.allclose(out["score"].cpu().numpy()[sel], out["p"].cpu().numpy().reshape(-1)[idx[sel]])
    # poses of the kept rows against the oracle's poses of the SAME rows (bf16 generator drift)
    d_t = np.abs(out["transl"].cpu().numpy()[sel] - ref["transl"][idx[sel]]).max()
    d_j = np.abs(out["joint_conf"].cpu().numpy()[sel] - ref["joint"][idx[sel]]).max()
    print("kept rows: max |d transl| %.2e, max |d joint| %.2e" % (d_t, d_j))
    assert d_t < tolerances.BF16_POSE_TOL and d_j < tolerances.BF16_POSE_TOL


@pytest.mark.gpu
@pytest.mark.parametrize("precision", ["fp32", "fp32_tc"])
def test_fp32_pipeline_topk_set_vs_fp32_oracle(precision):
    out, ref = _run(precision)
    # the pose drift of a 1e-4 generator seen through the evaluator: 4 x FP32_TOL end to end
    rec = pc.check_pipeline(out, ref, S, THRESH, K, 4 * tolerances.FP32_TOL)
    print(precision, "pipeline vs fp32 oracle pipeline:", rec)
    assert rec["bps_bit_exact"] and rec["max_dp"] < 4 * tolerances.FP32_TOL and rec["set_mismatches"] == 0, rec
