"""GPU parity of evaluator-gradient grasp refinement (reference
DexGanGrasp/models/dexgangrasp.py:516-541 improve_grasps_gradient_based, :619-660 refine_grasps) run on libdexgg's GEMMs,
against (1) the golden trajectory produced by executing the reference's own method (tests/golden/make_refine_golden.py)
and (2) the CPU oracle on other inputs.

Tolerance: scores 1e-4 absolute (north_star's fp32 network tolerance); transl / rot_matrix 1e-4 absolute after up to ten
1 cm steps - each step is `grad * 0.01 / |grad|`, i.e. the direction of a gradient whose fp32 GEMM rounding differs from
the CPU's by ~1e-5 relative."""
import numpy as np
import pytest
import torch

from dex_gan_grasp_b200 import synth
from oracle import nets_oracle as no
from test_oracle_refine import golden_inputs
from test_train_gpu import CFG, SEED

pytestmark = pytest.mark.gpu


@pytest.fixture()
def model():
    from dex_gan_grasp_b200.dexgangrasp import DexGanGrasp
    m = DexGanGrasp(dict(CFG, is_train=False))
    m.DexEvaluator.load_state_dict(synth.synth_eva_state_dict(SEED))
    return m


def _check(data, succ, ref_data, ref_succ, tol=1e-4):
    assert len(data) == len(ref_data) and len(succ) == len(ref_succ)
    for i, (a, b) in enumerate(zip(succ, ref_succ)):
        assert a.shape == b.shape and np.abs(a - b).max() < tol, ("success", i, np.abs(a - b).max())
    for i, (a, b) in enumerate(zip(data, ref_data)):
        assert set(a.keys()) == {"rot_matrix", "transl", "joint_conf"}
        for k in a:
            assert a[k].shape == b[k].shape and np.abs(a[k] - b[k]).max() < tol, (k, i, np.abs(a[k] - b[k]).max())


def test_refine_matches_reference_golden(model):
    g, batch = golden_inputs()
    steps = int(g["steps"])
    data, succ = model.refine_grasps({k: batch[k] for k in ("bps_object", "rot_matrix", "transl", "joint_conf")}, "gradient",
                                     num_refine_steps=steps)
    assert len(data) == steps + 1 and len(succ) == steps + 1
    assert np.abs(np.stack(succ) - g["success"]).max() < 1e-4
    for i in (1, 2, steps):
        assert np.abs(data[i]["rot_matrix"] - g["rot_matrix/%d" % i]).max() < 1e-4
        assert np.abs(data[i]["transl"] - g["transl/%d" % i]).max() < 1e-4
    assert np.array_equal(data[-1]["joint_conf"], data[0]["joint_conf"])


@pytest.mark.parametrize("B,steps", [(1, 3), (33, 10), (700, 4)])
def test_refine_matches_oracle(model, B, steps):
    batch = synth.synth_train_batch(B, seed=41 + B)
    ref_data, ref_succ = no.refine_grasps_gradient(synth.synth_eva_state_dict(SEED), batch, steps)
    data, succ = model.refine_grasps({k: batch[k] for k in ("bps_object", "rot_matrix", "transl", "joint_conf")}, "gradient",
                                     num_refine_steps=steps)
    if B == 1:  # .squeeze() of a single grasp is a 0-d array in the reference too
        assert succ[0].shape == ()
    _check(data, succ, ref_data, ref_succ)


def test_refine_is_fp32_in_bf16_mode_and_single_step_api(model):
    """cfg precision 'bf16' must not change refinement (the step is a gradient direction), and
    improve_grasps_gradient_based is callable on its own with the reference's (p_success, None) return."""
    from dex_gan_grasp_b200 import train_ops
    B = 24
    batch = synth.synth_train_batch(B, seed=5)
    ref_data, ref_succ = no.refine_grasps_gradient(synth.synth_eva_state_dict(SEED), batch, 1)
    train_ops.set_precision("bf16")
    try:
        d = {"bps_object": batch["bps_object"].cuda()}
        for k in ("rot_matrix", "transl", "joint_conf"):
            d[k] = batch[k].float().cuda().requires_grad_(True)
        p, last = model.improve_grasps_gradient_based(d, None)
    finally:
        train_ops.set_precision("fp32")
    assert last is None and p.shape == (B,)
    assert np.abs(p.detach().cpu().numpy() - ref_succ[0]).max() < 1e-4
    assert np.abs(d["transl"].detach().cpu().numpy() - ref_data[1]["transl"]).max() < 1e-4
    assert np.abs(d["rot_matrix"].detach().cpu().numpy() - ref_data[1]["rot_matrix"]).max() < 1e-4
    assert d["joint_conf"].grad is None


def test_refine_sampling_variant_not_provided(model):
    with pytest.raises(NotImplementedError):
        model.refine_grasps({}, "sampling")
