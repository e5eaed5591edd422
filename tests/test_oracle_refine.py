"""CPU: the grasp-refinement oracle (oracle/nets_oracle.refine_grasps_gradient) against the golden trajectory produced by
EXECUTING the reference's own `improve_grasps_gradient_based` in refine_grasps' loop
(tests/golden/make_refine_golden.py; reference DexGanGrasp/models/dexgangrasp.py:516-541, 619-660)."""
import os

import numpy as np

from dex_gan_grasp_b200 import synth
from oracle import nets_oracle as no

GOLD = os.path.join(os.path.dirname(__file__), "golden", "refine_golden.npz")


def golden_inputs():
    g = np.load(GOLD)
    B = int(g["B"])
    batch = synth.synth_train_batch(B, seed=int(g["batch_seed"]))
    batch["bps_object"] = batch["bps_object"][:4].repeat_interleave(B // 4, dim=0).contiguous()
    return g, batch


def test_oracle_refine_matches_reference_golden():
    g, batch = golden_inputs()
    steps = int(g["steps"])
    data, succ = no.refine_grasps_gradient(synth.synth_eva_state_dict(int(g["eva_seed"])), batch, steps)
    assert len(data) == steps + 1 and len(succ) == steps + 1  # refine_grasps :643-656
    assert np.abs(np.stack(succ) - g["success"]).max() < 1e-6
    for i in (1, 2, steps):
        assert np.abs(data[i]["rot_matrix"] - g["rot_matrix/%d" % i]).max() < 1e-6
        assert np.abs(data[i]["transl"] - g["transl/%d" % i]).max() < 1e-6
    # every step moves transl by exactly the 1 cm cap here (alpha = 0.01 / |accumulated grad| < 1) ...
    step_len = np.linalg.norm(data[1]["transl"] - data[0]["transl"], axis=1)
    assert np.allclose(step_len, 0.01, atol=1e-6)
    # ... joint_conf is not an evaluator input and stays as given, and the scores improve on average
    assert np.array_equal(data[-1]["joint_conf"], data[0]["joint_conf"])
    assert g["success"][-1].mean() > g["success"][0].mean()
