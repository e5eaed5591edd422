#!/usr/bin/env python
"""Generates tests/golden/wgan_golden.npz by EXECUTING the reference's own WGAN-GP training step.

The reference wrapper class (DexGanGrasp/models/dexgangrasp.py:65) cannot be constructed on a CPU-only box (its
__init__ selects cuda:0 unconditionally), but its METHODS run anywhere: an instance is made with
`object.__new__`, given the attributes __init__ would have set (the reference's own nn.Modules on CPU, torch.optim.Adam,
the reference's loss functions), and `update_dexgangrasp` (:757-823) is called unmodified with `is_wgan = True`, which
goes through calculate_interp / calculate_grad_interp / calculate_gradient_penalty /
compute_loss_dexgangrasp_discriminator_wass / compute_loss_dexgangrasp_generator_wass (:149-270).  `torch.randn` is
wrapped by a recorder so the latent batch and the three interpolation weights the reference drew are stored with the
results.  No reference source is copied into the repository.  Run in the build container only:

    python tests/golden/make_wgan_golden.py

It also asserts that oracle/nets_oracle.wgan_step reproduces the reference's losses, gradients and updated weights."""
import copy
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
from DexGanGrasp.models import losses as ref_losses  # noqa: E402
from DexGanGrasp.models import networks as ref_networks  # noqa: E402
from DexGanGrasp.utils import utils as ref_utils  # noqa: E402

assert ref_wrapper.__file__.startswith(REF)


def _normalize_vector_cpu(v, return_mag=False):
    # utils.py:323-332 minus the `.cuda()` at :326
    batch = v.shape[0]
    v_mag = torch.sqrt(v.pow(2).sum(1))
    v_mag = torch.max(v_mag, torch.FloatTensor([1e-8]))
    v_mag = v_mag.view(batch, 1).expand(batch, v.shape[1])
    v = v / v_mag
    return (v, v_mag[:, 0]) if return_mag else v


ref_utils.normalize_vector = _normalize_vector_cpu

SEED = 7
B = 48
cfg = dict(latentD=5, n_hand_joints=12, is_train=True, gpu_ids=[0])
gan = ref_networks.DexGANGrasp(cfg)
gan.load_state_dict(synth.synth_gan_state_dict(SEED))
gan.train()

w = object.__new__(ref_wrapper.DexGanGrasp)  # the reference class, __init__ skipped (it needs cuda:0)
w.is_wgan = True
w.DexGANGrasp = gan
w.device = torch.device("cpu")
w.bce_weight = 1.0
w.transl_coef, w.rot_coef, w.conf_coef = 100., 1., 10.  # dexgangrasp.py:92-94
w.L2_loss = torch.nn.MSELoss(reduction="mean")
w.BCE_loss = torch.nn.BCELoss(reduction="mean")
w.rec_pose_loss = ref_losses.transl_rot_6D_l2_loss
w.optim_dexgangrasp_generator = torch.optim.Adam(gan.generator.parameters(), lr=1e-4, betas=(0.9, 0.999), weight_decay=5e-4)
w.optim_dexgangrasp_discriminator = torch.optim.Adam(gan.discriminator.parameters(), lr=1e-4, betas=(0.9, 0.999),
                                                     weight_decay=5e-4)
w.last_loss_dict_gen = {}

batch = synth.synth_train_batch(B, seed=13)
real_data = {k: batch[k].clone() for k in ("bps_object", "rot_matrix", "transl", "joint_conf")}

drawn = []
_randn = torch.randn


def _recording_randn(*a, **k):
    t = _randn(*a, **k)
    drawn.append(t.clone())
    return t


torch.manual_seed(2024)
torch.randn = _recording_randn
try:
    loss_dict = w.update_dexgangrasp(real_data, is_train_gen=True)
finally:
    torch.randn = _randn
assert [tuple(t.shape) for t in drawn] == [(B, 5), (B, 1, 1), (B, 1), (B, 1)], [t.shape for t in drawn]
z, alphas = drawn[0], drawn[1:]
losses_ref = {k: float(v) for k, v in loss_dict.items()}
sd_after = {k: v.detach().clone() for k, v in gan.state_dict().items()}
print("reference WGAN-GP step:", losses_ref)

sd_o = copy.deepcopy(synth.synth_gan_state_dict(SEED))
losses_o, grads_o = nets_oracle.wgan_step(sd_o, {}, {}, batch, z, alphas, True)
for k, v in losses_ref.items():
    assert abs(losses_o[k] - v) < 1e-5 * max(1.0, abs(v)), (k, losses_o[k], v)
worst, bad, tot = 0.0, 0, 0
for k, v in sd_after.items():
    if not v.is_floating_point():
        assert int(sd_o[k]) == int(v), k
        continue
    d = (sd_o[k].double() - v.double()).abs()
    worst = max(worst, float(d.max()))
    bad += int((d > 2e-6).sum())
    tot += d.numel()
# first Adam step: |delta| ~ lr whatever the gradient scale; ill-conditioned only where g + wd*w ~ eps (make_golden.py)
assert worst <= 2.1e-4 and bad <= 1e-4 * tot, (worst, bad, tot)
print("oracle wgan_step == reference update_dexgangrasp(is_wgan): worst", worst, "n>2e-6", bad, "of", tot,
      "GP", losses_o["gradient_penalty"])

out = {"z": z.numpy(), "alpha_rot": alphas[0].numpy(), "alpha_transl": alphas[1].numpy(), "alpha_joint": alphas[2].numpy(),
       "batch_seed": np.int64(13), "B": np.int64(B), "gradient_penalty_oracle": np.float64(losses_o["gradient_penalty"])}
out.update({"loss/" + k: np.float64(v) for k, v in losses_ref.items()})
sd0 = synth.synth_gan_state_dict(SEED)
probes = ["discriminator.disc_rb1.fc1.weight", "discriminator.disc_rb1.fc3.weight", "discriminator.disc_rb2.fc2.weight",
          "discriminator.out_success.weight", "discriminator.disc_bn1.weight", "discriminator.disc_rb1.bn1.bias",
          "discriminator.disc_bn1.running_mean", "discriminator.disc_bn1.running_var", "discriminator.disc_rb2.bn2.running_var",
          "generator.gen_rb1.fc1.weight", "generator.gen_rot.weight", "generator.gen_bn1.running_var"]
for k in probes:
    out["after/" + k] = sd_after[k].reshape(-1)[:4096].numpy()
    out["delta/" + k] = (sd_after[k] - sd0[k]).reshape(-1)[:4096].numpy()
for k in ("discriminator.disc_rb2.fc2.weight", "discriminator.disc_rb1.fc1.weight", "discriminator.out_success.weight"):
    out["grad_oracle/" + k] = grads_o["d"][k].reshape(-1)[:4096].numpy()
out["num_batches_tracked/discriminator.disc_bn1"] = np.int64(int(sd_after["discriminator.disc_bn1.num_batches_tracked"]))
np.savez_compressed(os.path.join(HERE, "wgan_golden.npz"), **out)
print("wrote wgan_golden.npz")
