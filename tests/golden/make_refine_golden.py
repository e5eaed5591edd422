#!/usr/bin/env python
"""Generates tests/golden/refine_golden.npz by EXECUTING the reference's own improve_grasps_gradient_based
(DexGanGrasp/models/dexgangrasp.py:516-541) in the loop refine_grasps runs it in (:645-655).

As in make_wgan_golden.py the reference wrapper is instantiated with `object.__new__` (its __init__ needs cuda:0) and
given the reference's own DexEvaluator module on CPU; the method then runs unmodified.  The reference's refine_grasps
itself cannot run: DexEvaluator never reads joint_conf (networks.py:307-313), so data['joint_conf'].grad is None at :539
and the first step raises TypeError.  Here that .grad is pre-set to zeros - the gradient it mathematically is - which
lets every line of the reference method execute.  No reference source is copied into the repository.  Run in the build
container only:

    python tests/golden/make_refine_golden.py

It also asserts that oracle/nets_oracle.refine_grasps_gradient reproduces the reference's trajectory."""
import importlib.util
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"


def _load(name, path):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


synth = _load("dexgg_synth", os.path.join(REPO, "dex_gan_grasp_b200", "synth.py"))
nets_oracle = _load("dexgg_nets_oracle", os.path.join(REPO, "oracle", "nets_oracle.py"))

for m in ("h5py", "open3d", "transforms3d"):
    sys.modules[m] = types.ModuleType(m)
sys.path.insert(0, REF)
from DexGanGrasp.models import dexgangrasp as ref_wrapper  # noqa: E402
from DexGanGrasp.models import networks as ref_networks  # noqa: E402

assert ref_wrapper.__file__.startswith(REF)

SEED, B, STEPS, BATCH_SEED = 7, 40, 10, 29
eva = ref_networks.DexEvaluator(dict(latentD=5, n_hand_joints=12, is_train=False, gpu_ids=[0]))
eva.load_state_dict(synth.synth_eva_state_dict(SEED))
eva.eval()

w = object.__new__(ref_wrapper.DexGanGrasp)
w.DexEvaluator = eva
w.device = torch.device("cpu")

batch = synth.synth_train_batch(B, seed=BATCH_SEED)
# a few rows per object, as refine_grasps sees them after filter_grasps (bps_object tiled per kept grasp)
batch["bps_object"] = batch["bps_object"][:4].repeat_interleave(B // 4, dim=0).contiguous()
data = {"bps_object": batch["bps_object"].clone(), "label": batch["label"].clone()}
for k in ("rot_matrix", "transl", "joint_conf"):  # refine_grasps :633-638
    data[k] = torch.autograd.Variable(batch[k].float().clone(), requires_grad=True)
data["joint_conf"].grad = torch.zeros_like(data["joint_conf"])

traj = [{k: data[k].detach().numpy().copy() for k in ("rot_matrix", "transl", "joint_conf")}]
succ = []
last = None
for _ in range(STEPS):  # refine_grasps :650-653
    p, last = w.improve_grasps_gradient_based(data, last)
    succ.append(p.detach().numpy().copy())
    traj.append({k: data[k].detach().numpy().copy() for k in ("rot_matrix", "transl", "joint_conf")})
succ.append(eva(data).squeeze().detach().numpy().copy())  # :656

rd, rs = nets_oracle.refine_grasps_gradient(synth.synth_eva_state_dict(SEED), batch, STEPS)
worst = 0.0
for i in range(STEPS + 1):
    for k in ("rot_matrix", "transl", "joint_conf"):
        worst = max(worst, float(np.abs(rd[i][k] - traj[i][k]).max()))
    worst = max(worst, float(np.abs(rs[i] - succ[i]).max()))
assert worst < 1e-6, worst
moved = float(np.linalg.norm(traj[-1]["transl"] - traj[0]["transl"], axis=1).max())
print("oracle refine_grasps_gradient == reference improve_grasps_gradient_based x", STEPS, ": worst", worst,
      "| max transl displacement", moved, "| mean score", float(succ[0].mean()), "->", float(succ[-1].mean()))
assert np.array_equal(traj[-1]["joint_conf"], traj[0]["joint_conf"])

out = {"B": np.int64(B), "steps": np.int64(STEPS), "batch_seed": np.int64(BATCH_SEED), "eva_seed": np.int64(SEED),
       "success": np.stack(succ)}
for i in (1, 2, STEPS):
    out["rot_matrix/%d" % i] = traj[i]["rot_matrix"]
    out["transl/%d" % i] = traj[i]["transl"]
np.savez_compressed(os.path.join(HERE, "refine_golden.npz"), **out)
print("wrote refine_golden.npz")
