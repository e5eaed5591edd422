"""GPU parity of the training step (DexGanGrasp.update_dexgangrasp / update_dexevaluator, reference
DexGanGrasp/models/dexgangrasp.py:739-823) run on the train-mode kernels of libdexgg, against the CPU oracle
(oracle/nets_oracle.py: torch-CPU autograd on the functional restatement + restated Adam) from identical state.

Tolerances: losses 1e-4 absolute, BatchNorm running statistics 1e-5, weights after one Adam step 1e-5
(Adam's first step moves every weight by ~lr = 1e-4 whatever the gradient scale, so the gradients themselves are also
compared, in the relative L2 norm — see grad_close)."""
import copy

import numpy as np
import pytest
import torch

from dex_gan_grasp_b200 import synth
from oracle import nets_oracle as no

pytestmark = pytest.mark.gpu
SEED = 7
CFG = dict(latentD=5, n_hand_joints=12, is_train=True, gpu_ids=[0], model="dexgangrasp", is_wgan=False, bce_weight=1.0,
           train_ffhgenerator=True, train_ffhevaluator=True, beta1=0.9, weight_decay=5e-4, lr=1e-4, lr_gen=1e-4, lr_dis=1e-4,
           lr_policy="plateau", threshold_lr_policy_plateau=0.01, patience_lr_policy_plateau=5, weight_init_type="xavier",
           init_gain=0.02)
B = 192


def md(a, b):
    a = a.detach().cpu().double().reshape(-1)
    b = b.detach().cpu().double().reshape(-1)
    return float((a - b).abs().max())


@pytest.fixture()
def model():
    from dex_gan_grasp_b200.dexgangrasp import DexGanGrasp
    m = DexGanGrasp(dict(CFG))
    m.DexGANGrasp.load_state_dict(synth.synth_gan_state_dict(SEED))
    m.DexEvaluator.load_state_dict(synth.synth_eva_state_dict(SEED))
    return m


def _batch():
    return synth.synth_train_batch(B, seed=11)


def grad_close(got, ref, what):
    """Gradients are compared in the relative L2 norm, 5e-3.  Agreement is ~1e-6 when no LeakyReLU pre-activation of the
    batch lies within fp32 rounding of zero; when one does, its sign (slope 1 vs 0.2) legitimately differs between two
    correct fp32 implementations and moves a few gradient entries by percents of the max (seen torch-CUDA vs torch-CPU
    too), so a max-norm test is flaky by construction.  Any real defect (missing term, wrong scale) is O(0.1-1) here."""
    assert got is not None, what
    ref = ref.double()
    den = float(ref.norm())
    if what[1].endswith(("fc1.bias", "fc2.bias")):  # a bias in front of a train-mode BatchNorm: the true gradient is
        assert den < 1e-5 and float(got.detach().cpu().double().norm()) < 1e-5, what  # zero, both sides hold rounding noise
        return
    err = float((got.detach().cpu().double() - ref).norm()) / den
    assert err < 5e-3, (what, err)


def _check_state(module, sd_ref, grads=None, w0=None, lr=1e-4, steps=1, wd=5e-4):
    """Buffers (BatchNorm running statistics) at 1e-5.  Weights after Adam: an element whose gradient is far from zero
    must agree at 1e-5; where g + wd*w is within rounding noise of zero, Adam's g/(|g|+eps) amplifies that noise up to the
    full step lr (in a 1M-element tensor a few hundred elements always are), so those elements are only bounded by 2*lr
    per step.  This is a property of Adam, equally true torch-vs-torch."""
    sd = module.state_dict()
    params = {k for k, _ in module.named_parameters()}
    for k, v in sd_ref.items():
        if not v.is_floating_point():
            assert int(sd[k]) == int(v), k
            continue
        err = (sd[k].detach().cpu().double() - v.double()).abs()
        if k not in params:
            assert float(err.max()) < (1e-5 if steps == 1 else 2e-4) * max(1.0, float(v.abs().max())), (k, float(err.max()))
            continue
        assert float(err.max()) <= 2.1 * lr * steps, (k, float(err.max()))
        if grads is not None and k in grads:
            g = (grads[k] + wd * w0[k]).abs()  # torch.optim.Adam couples the weight decay into the gradient
            solid = g > 1e-3 * float(g.max())
            if solid.any():  # (a LeakyReLU sign flip, see grad_close, can still turn the sign of a few of them)
                frac = float((err[solid] < 1e-5).double().mean())
                assert frac > 0.995, (k, frac, float(err[solid].max()))
        else:
            assert float((err < 1e-5).double().mean()) > 0.98, (k, float((err < 1e-5).double().mean()))


def test_gan_step_matches_oracle(model):
    batch = _batch()
    z = synth.synth_latents(B, 5, seed=3)
    sd = {k: v.clone() for k, v in synth.synth_gan_state_dict(SEED).items()}
    opt_d, opt_g = {}, {}
    ref_losses, ref_grads = no.gan_step(sd, opt_d, opt_g, copy.deepcopy(batch), z.clone(), is_train_gen=True)

    # capture gradients of our step
    ours = model.update_dexgangrasp(copy.deepcopy(batch), is_train_gen=True, Zgen=z.cuda())
    torch.cuda.synchronize()
    for k, v in ref_losses.items():
        assert abs(float(ours[k]) - v) < 1e-4 * max(1.0, abs(v)), (k, float(ours[k]), v)
    named = dict(model.DexGANGrasp.named_parameters())
    for k, g in ref_grads["g"].items():
        grad_close(named[k].grad, g, ("G grad", k))
    _check_state(model.DexGANGrasp, sd, {**ref_grads["d"], **ref_grads["g"]}, synth.synth_gan_state_dict(SEED))


def test_gan_step_without_generator_update(model):
    batch = _batch()
    z = synth.synth_latents(B, 5, seed=4)
    sd = {k: v.clone() for k, v in synth.synth_gan_state_dict(SEED).items()}
    ref_losses, ref_grads = no.gan_step(sd, {}, {}, copy.deepcopy(batch), z.clone(), is_train_gen=False)
    ours = model.update_dexgangrasp(copy.deepcopy(batch), is_train_gen=False, Zgen=z.cuda())
    for k, v in ref_losses.items():
        assert abs(float(ours[k]) - v) < 1e-4 * max(1.0, abs(v)), k
    named = dict(model.DexGANGrasp.named_parameters())
    for k, g in ref_grads["d"].items():
        grad_close(named[k].grad, g, ("D grad", k))
    _check_state(model.DexGANGrasp, sd, ref_grads["d"], synth.synth_gan_state_dict(SEED))


def test_two_gan_steps_adam_state(model):
    """Second step exercises the Adam moments and the re-packed weights."""
    sd = {k: v.clone() for k, v in synth.synth_gan_state_dict(SEED).items()}
    opt_d, opt_g = {}, {}
    for it in range(2):
        batch = synth.synth_train_batch(B, seed=20 + it)
        z = synth.synth_latents(B, 5, seed=30 + it)
        ref_losses, _ = no.gan_step(sd, opt_d, opt_g, copy.deepcopy(batch), z.clone())
        ours = model.update_dexgangrasp(copy.deepcopy(batch), Zgen=z.cuda())
        for k, v in ref_losses.items():
            assert abs(float(ours[k]) - v) < 2e-4 * max(1.0, abs(v)), (it, k, float(ours[k]), v)
    _check_state(model.DexGANGrasp, sd, steps=2)
    st = model.optim_dexgangrasp_generator.state_dict()["state"]
    assert all(set(s.keys()) == {"step", "exp_avg", "exp_avg_sq"} for s in st.values())  # torch.optim.Adam checkpoint layout


def test_evaluator_step_matches_oracle(model):
    batch = _batch()
    g = torch.Generator().manual_seed(5)
    masks = [(torch.rand(B, 512, generator=g) >= 0.3).float() / 0.7 for _ in range(3)]  # Dropout(0.3) keep masks, pre-scaled
    sd = {k: v.clone() for k, v in synth.synth_eva_state_dict(SEED).items()}
    ref_losses, ref_grads = no.evaluator_step(sd, {}, copy.deepcopy(batch), dropout_masks=masks)
    model.DexEvaluator.dropout_masks = [m.cuda() for m in masks]
    ours = model.update_dexevaluator(copy.deepcopy(batch))
    model.DexEvaluator.dropout_masks = None
    for k, v in ref_losses.items():
        assert abs(float(ours[k]) - v) < 1e-4 * max(1.0, abs(v)), k
    named = dict(model.DexEvaluator.named_parameters())
    for k, gr in ref_grads.items():
        grad_close(named[k].grad, gr, ("E grad", k))
    _check_state(model.DexEvaluator, sd, ref_grads, synth.synth_eva_state_dict(SEED))


def test_evaluator_step_with_torch_dropout_runs(model):
    """Fast mode: torch's own Philox dropout stream; the loss must be finite and the weights must move."""
    before = model.DexEvaluator.out_success.weight.detach().clone()
    out = model.update_dexevaluator(_batch())
    assert np.isfinite(float(out["total_loss_eva"]))
    assert md(before, model.DexEvaluator.out_success.weight) > 0


def test_eval_losses_after_training(model):
    """eval_* helpers (dexgangrasp.py:309-365) run the packed eval-mode forward on the just-updated weights."""
    batch = _batch()
    model.update_dexgangrasp(copy.deepcopy(batch))
    d = model.eval_dexgangrasp_generator_loss(copy.deepcopy(batch))
    assert all(np.isfinite(float(v)) for v in d.values())
    e = model.eval_dexevaluator_loss(copy.deepcopy(batch))
    assert np.isfinite(float(e["total_loss_eva"]))
    pos, neg, pred, gt = model.eval_dexevaluator_accuracy(copy.deepcopy(batch))
    assert pred.shape == gt.shape


# ------------------------------------------------------------------ bf16 tensor-core training mode (BASELINE.json configs[4])
from dex_gan_grasp_b200 import tolerances as _tol  # noqa: E402  (the stated tolerances live in ONE place)

BF16_LOSS_TOL = _tol.BF16_TRAIN_LOSS_TOL
BF16_GRAD_TOL_ALL = _tol.BF16_TRAIN_GRAD_TOL_ALL


def _bf16_grads_close(named, ref_grads):
    num = den = 0.0
    for k, g in ref_grads.items():
        if k.endswith(("fc1.bias", "fc2.bias")) or float(g.norm()) < 1e-6:
            continue  # bias in front of a train-mode BatchNorm: true gradient zero
        rel = _grad_rel(named[k].grad, g)
        tol = _tol.BF16_TRAIN_GRAD_TOL_ROT_HEAD if "gen_rot." in k else (_tol.BF16_TRAIN_GRAD_TOL_SMALL if g.numel() < 64 else _tol.BF16_TRAIN_GRAD_TOL)
        assert rel < tol, (k, rel)
        num += float((named[k].grad.detach().cpu().double() - g.double()).pow(2).sum())
        den += float(g.double().pow(2).sum())
    assert (num / den) ** 0.5 < BF16_GRAD_TOL_ALL, (num / den) ** 0.5


@pytest.fixture()
def model_bf16():
    from dex_gan_grasp_b200 import train_ops
    from dex_gan_grasp_b200.dexgangrasp import DexGanGrasp
    m = DexGanGrasp(dict(CFG, precision="bf16"))
    m.DexGANGrasp.load_state_dict(synth.synth_gan_state_dict(SEED))
    m.DexEvaluator.load_state_dict(synth.synth_eva_state_dict(SEED))
    yield m
    train_ops.set_precision("fp32")


def _grad_rel(got, ref):  # (defined before its users run)
    ref = ref.double()
    return float((got.detach().cpu().double() - ref).norm()) / max(float(ref.norm()), 1e-30)


def test_bf16_gan_step_close_to_oracle(model_bf16):
    """Same step as test_gan_step_matches_oracle with every GEMM (forward, dgrad, wgrad) on the tcgen05 kernel: losses within
    2e-2 relative, gradients within 2e-1 per tensor and 5e-2 over the whole network in the relative L2 norm, weights move by at most Adam's lr per step, BatchNorm
    running statistics (computed in fp32 from bf16-GEMM outputs) within 1e-2."""
    batch = _batch()
    z = synth.synth_latents(B, 5, seed=3)
    sd = {k: v.clone() for k, v in synth.synth_gan_state_dict(SEED).items()}
    ref_losses, ref_grads = no.gan_step(sd, {}, {}, copy.deepcopy(batch), z.clone(), is_train_gen=True)
    ours = model_bf16.update_dexgangrasp(copy.deepcopy(batch), is_train_gen=True, Zgen=z.cuda())
    torch.cuda.synchronize()
    for k, v in ref_losses.items():
        assert abs(float(ours[k]) - v) < BF16_LOSS_TOL * max(1.0, abs(v)), (k, float(ours[k]), v)
    named = dict(model_bf16.DexGANGrasp.named_parameters())
    _bf16_grads_close(named, ref_grads["g"])
    _bf16_grads_close(named, ref_grads["d"])  # the discriminator step's gradients (the generator step leaves them in place)
    w0 = synth.synth_gan_state_dict(SEED)
    for k, p in named.items():
        assert float((p.detach().cpu() - w0[k]).abs().max()) <= 1.05e-4, k  # |Adam step 1| <= lr
    for k, v in model_bf16.DexGANGrasp.state_dict().items():
        if "running_" in k:
            assert md(v, sd[k]) < 1e-2 * max(1.0, float(sd[k].abs().max())), k


def test_bf16_evaluator_step_close_to_oracle(model_bf16):
    batch = _batch()
    g = torch.Generator().manual_seed(5)
    masks = [(torch.rand(B, 512, generator=g) >= 0.3).float() / 0.7 for _ in range(3)]
    sd = {k: v.clone() for k, v in synth.synth_eva_state_dict(SEED).items()}
    ref_losses, ref_grads = no.evaluator_step(sd, {}, copy.deepcopy(batch), dropout_masks=masks)
    model_bf16.DexEvaluator.dropout_masks = [m.cuda() for m in masks]
    ours = model_bf16.update_dexevaluator(copy.deepcopy(batch))
    model_bf16.DexEvaluator.dropout_masks = None
    for k, v in ref_losses.items():
        assert abs(float(ours[k]) - v) < BF16_LOSS_TOL * max(1.0, abs(v)), (k, float(ours[k]), v)
    named = dict(model_bf16.DexEvaluator.named_parameters())
    _bf16_grads_close(named, ref_grads)


def test_bf16_training_reduces_the_loss(model_bf16):
    """20 bf16 steps on one batch: the generator's reconstruction losses must go down (the optimisation works end to end)."""
    batch = _batch()
    first = last = None
    for it in range(20):
        out = model_bf16.update_dexgangrasp(copy.deepcopy(batch))
        v = float(out["transl_loss"]) + float(out["conf_loss"])
        first = v if first is None else first
        last = v
    assert np.isfinite(last) and last < first, (first, last)


def test_captured_step_equals_eager_steps():
    """graph_step.CapturedTrainStep (the whole GAN + evaluator step as one CUDA graph, Adam step count on the device) against
    the eager step from the same state on the same batches: identical kernels in identical order, so weights and BatchNorm
    buffers must agree to fp32 atomics noise; dropout is off (p = 0) and z is supplied so that both sides are deterministic."""
    from dex_gan_grasp_b200.dexgangrasp import DexGanGrasp
    from dex_gan_grasp_b200.graph_step import CapturedTrainStep

    def fresh():
        m = DexGanGrasp(dict(CFG))
        m.DexGANGrasp.load_state_dict(synth.synth_gan_state_dict(SEED))
        m.DexEvaluator.load_state_dict(synth.synth_eva_state_dict(SEED))
        m.DexEvaluator.dout.p = 0.0
        return m

    batches = [{k: v.cuda() for k, v in synth.synth_train_batch(B, seed=40 + i).items()} for i in range(6)]
    zs = [synth.synth_latents(B, 5, seed=60 + i).cuda() for i in range(6)]
    eager = fresh()
    for i in [0] + list(range(6)):  # the captured side spends one eager warm-up step on the example batch first
        b = {k: v.clone() for k, v in batches[i].items()}
        le = dict(eager.update_dexgangrasp(b, Zgen=zs[i]))
        le.update(eager.update_dexevaluator({k: v.clone() for k, v in batches[i].items()}))
    torch.cuda.synchronize()
    cap = fresh()
    step = CapturedTrainStep(cap, batches[0], warmup=1, example_z=zs[0])  # one eager step on batch 0, then six replays
    for i in range(6):
        lc = step(batches[i], zs[i])
    torch.cuda.synchronize()
    # Tolerances: two runs of the SAME seven steps differ by fp32 atomics noise in the BatchNorm column sums, which Adam turns into
    # up to a full step of lr on parameters whose gradient is near zero (below); from there the generator's poses, the losses on
    # them and the discriminator's running statistics of the fake batch drift by the same order (measured run to run, this test
    # alone in a process, 12 runs: losses up to 1.4e-4 relative, disc_bn1.running_mean up to 7.7e-4).  Bounds are ~3x that.
    for k in le:
        assert abs(float(le[k]) - float(lc[k])) < 5e-4 * max(1.0, abs(float(le[k]))), (k, float(le[k]), float(lc[k]))
    for name in ("DexGANGrasp", "DexEvaluator"):
        sa, sb = getattr(eager, name).state_dict(), getattr(cap, name).state_dict()
        params = {k for k, _ in getattr(eager, name).named_parameters()}
        for k in sa:
            if not sa[k].is_floating_point():
                assert int(sa[k]) == int(sb[k]), (name, k)
            elif k in params:
                # Adam turns rounding noise in a near-zero gradient into up to a full step of lr (see _check_state): bounded
                # by 2.1 * lr per step, and a small fraction of lr on average (atomics make two eager runs differ the same way)
                d = (sa[k].double() - sb[k].double()).abs()
                assert float(d.max()) <= 2.1e-4 * 7 and (d.numel() < 1024 or float(d.mean()) < 3e-5), (name, k, float(d.max()), float(d.mean()))
            else:
                assert md(sa[k], sb[k]) < 2.5e-3 * max(1.0, float(sa[k].abs().max())), (name, k, md(sa[k], sb[k]))
    so = cap.optim_dexgangrasp_generator.state_dict()["state"]
    assert all(int(s["step"]) == 7 for s in so.values())
