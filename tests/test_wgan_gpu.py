"""GPU parity of the WGAN-GP training variant (`is_wgan: true`, reference DexGanGrasp/models/dexgangrasp.py:149-211,
234-270, switch at :761-763) run on libdexgg, against (1) the golden vectors produced by executing
the reference's own `update_dexgangrasp` (tests/golden/make_wgan_golden.py) and (2) the CPU oracle
(oracle/nets_oracle.wgan_step) from identical state, with the same latent batch and interpolation weights.

Tolerances: losses 1e-4 (the total discriminator loss carries 10 x the gradient penalty, so 1e-4 relative), BatchNorm
running statistics 1e-5, weights after one Adam step as in test_train_gpu._check_state, gradients 5e-3 relative L2."""
import copy
import os

import numpy as np
import pytest
import torch

from dex_gan_grasp_b200 import synth
from oracle import nets_oracle as no
from test_train_gpu import CFG, SEED, _check_state, grad_close

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden", "wgan_golden.npz")


@pytest.fixture()
def model():
    from dex_gan_grasp_b200.dexgangrasp import DexGanGrasp
    m = DexGanGrasp(dict(CFG, is_wgan=True))
    m.DexGANGrasp.load_state_dict(synth.synth_gan_state_dict(SEED))
    return m


def _golden_inputs():
    g = np.load(GOLD)
    batch = synth.synth_train_batch(int(g["B"]), seed=int(g["batch_seed"]))
    z = torch.from_numpy(g["z"])
    alphas = tuple(torch.from_numpy(g[k]) for k in ("alpha_rot", "alpha_transl", "alpha_joint"))
    return g, batch, z, alphas


def test_wgan_step_matches_reference_golden(model):
    g, batch, z, alphas = _golden_inputs()
    ours = model.update_dexgangrasp(copy.deepcopy(batch), is_train_gen=True, Zgen=z.cuda(), alphas=alphas)
    torch.cuda.synchronize()
    keys = [k[5:] for k in g.files if k.startswith("loss/")]
    assert set(keys) == set(ours.keys())  # the reference's loss-dict keys, no more, no fewer
    for k in keys:
        v = float(g["loss/" + k])
        assert abs(float(ours[k].detach()) - v) < 1e-4 * max(1.0, abs(v)), (k, float(ours[k].detach()), v)
    sd = model.DexGANGrasp.state_dict()
    for k in [k[6:] for k in g.files if k.startswith("after/")]:
        got = sd[k].detach().cpu().reshape(-1)[:4096].double().numpy()
        err = np.abs(got - g["after/" + k].astype(np.float64))
        if "running_" in k:
            assert err.max() < 1e-5 * max(1.0, np.abs(g["after/" + k]).max()), (k, err.max())
        else:  # one Adam step: |delta| <= lr; nearly all elements at 1e-5 (see test_train_gpu._check_state)
            assert err.max() <= 2.1e-4, (k, err.max())
            assert (err < 1e-5).mean() > 0.98, (k, (err < 1e-5).mean())
    assert int(sd["discriminator.disc_bn1.num_batches_tracked"]) == int(g["num_batches_tracked/discriminator.disc_bn1"])


def test_wgan_step_matches_oracle(model):
    B = 96
    batch = synth.synth_train_batch(B, seed=17)
    z = synth.synth_latents(B, 5, seed=6)
    gen = torch.Generator().manual_seed(8)
    alphas = (torch.randn(B, 1, 1, generator=gen), torch.randn(B, 1, generator=gen), torch.randn(B, 1, generator=gen))
    sd = {k: v.clone() for k, v in synth.synth_gan_state_dict(SEED).items()}
    ref_losses, ref_grads = no.wgan_step(sd, {}, {}, copy.deepcopy(batch), z.clone(), alphas, is_train_gen=False)
    ours = model.update_dexgangrasp(copy.deepcopy(batch), is_train_gen=False, Zgen=z.cuda(), alphas=alphas)
    for k in ("total_loss_disc", "bce_loss_real", "bce_loss_fake"):
        v = ref_losses[k]
        assert abs(float(ours[k].detach()) - v) < 1e-4 * max(1.0, abs(v)), (k, float(ours[k].detach()), v)
    named = dict(model.DexGANGrasp.named_parameters())
    for k, gr in ref_grads["d"].items():  # includes the second-order terms of the gradient penalty
        grad_close(named[k].grad, gr, ("D grad (WGAN-GP)", k))
    _check_state(model.DexGANGrasp, sd, ref_grads["d"], synth.synth_gan_state_dict(SEED))


def test_gradient_penalty_value_and_matmul_double_backward(model):
    """The gradient penalty alone against the oracle, and _MatmulFn's second derivative against torch.matmul's."""
    from dex_gan_grasp_b200 import train_ops as T
    g, batch, z, alphas = _golden_inputs()
    sd = {k: v.clone() for k, v in synth.synth_gan_state_dict(SEED).items()}
    fake = {"rot_matrix": batch["rot_matrix"].float().flip(0).contiguous(), "transl": batch["transl"].float().flip(0).contiguous(),
            "joint_conf": batch["joint_conf"].float().flip(0).contiguous()}
    ref = no.gradient_penalty(sd, batch["bps_object"].float(), batch, fake, alphas)
    model.DexGANGrasp.train()
    ours = model.calculate_gradient_penalty(batch, {k: v.cuda() for k, v in fake.items()}, alphas)
    assert abs(float(ours) - float(ref)) < 1e-4 * max(1.0, abs(float(ref))), (float(ours), float(ref))

    gen = torch.Generator().manual_seed(1)
    for ta, tw in ((0, 0), (0, 1), (1, 0), (1, 1)):
        M, N, K = 37, 29, 53
        A = torch.randn((K, M) if ta else (M, K), generator=gen).cuda().requires_grad_(True)
        W = torch.randn((K, N) if tw else (N, K), generator=gen).cuda().requires_grad_(True)
        def f(mm):
            C = mm(A, W)
            gA, = torch.autograd.grad((C ** 2).sum(), A, create_graph=True)
            return C, torch.autograd.grad((gA ** 2).sum(), [A, W])
        C1, (a1, w1) = f(lambda a, w: T._MatmulFn.apply(a, w, ta, tw))
        C2, (a2, w2) = f(lambda a, w: (a.t() if ta else a) @ (w if tw else w.t()))
        for x, y in ((C1, C2), (a1, a2), (w1, w2)):
            assert float((x - y).abs().max()) < 1e-3 * float(y.abs().max()), (ta, tw)
