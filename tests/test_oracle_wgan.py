"""CPU: the WGAN-GP oracle (oracle/nets_oracle.wgan_step) against the golden vectors produced by EXECUTING the
reference's own `DexGanGrasp.update_dexgangrasp` with `is_wgan = True` (tests/golden/make_wgan_golden.py; reference
DexGanGrasp/models/dexgangrasp.py:149-270, 757-823)."""
import copy
import os

import numpy as np
import torch

from dex_gan_grasp_b200 import synth
from oracle import nets_oracle as no

GOLD = os.path.join(os.path.dirname(__file__), "golden", "wgan_golden.npz")


def test_oracle_wgan_step_matches_reference_golden():
    g = np.load(GOLD)
    batch = synth.synth_train_batch(int(g["B"]), seed=int(g["batch_seed"]))
    z = torch.from_numpy(g["z"])
    alphas = tuple(torch.from_numpy(g[k]) for k in ("alpha_rot", "alpha_transl", "alpha_joint"))
    sd = copy.deepcopy(synth.synth_gan_state_dict(7))
    losses, grads = no.wgan_step(sd, {}, {}, batch, z, alphas, True)
    for k in [k[5:] for k in g.files if k.startswith("loss/")]:
        v = float(g["loss/" + k])
        assert abs(losses[k] - v) < 1e-5 * max(1.0, abs(v)), (k, losses[k], v)
    assert losses["gradient_penalty"] > 0.1  # the penalty term is live (it dominates total_loss_disc here)
    for k in [k[6:] for k in g.files if k.startswith("after/")]:
        got = sd[k].reshape(-1)[:4096].double().numpy()
        err = np.abs(got - g["after/" + k].astype(np.float64))
        assert err.max() <= (2.1e-4 if "running_" not in k else 1e-6), (k, err.max())
        assert (err < 2e-6).mean() > 0.999, (k, (err < 2e-6).mean())
    # four train-mode passes through D per step: real, fake, interpolates, generator-step pass
    assert int(sd["discriminator.disc_bn1.num_batches_tracked"]) == int(g["num_batches_tracked/discriminator.disc_bn1"]) \
        == int(synth.synth_gan_state_dict(7)["discriminator.disc_bn1.num_batches_tracked"]) + 4
