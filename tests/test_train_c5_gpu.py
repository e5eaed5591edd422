"""Training parity at the C5 batch (BASELINE.json configs[4]: batch 4096) and of the eval_* helpers, against the CPU oracle
(oracle/nets_oracle.py) from identical state.  VERDICT r01 items 1(c), 1(d) and ADVICE r01 (packed weights after graph replays).

Reference: DexGanGrasp/models/dexgangrasp.py:739-823 (update_*), :309-365 (eval_*), models/losses.py:5-16."""
import copy

import numpy as np
import pytest
import torch

from dex_gan_grasp_b200 import synth, tolerances
from oracle import nets_oracle as no

pytestmark = pytest.mark.gpu
SEED = 7
B = 4096
CFG = dict(latentD=5, n_hand_joints=12, is_train=True, gpu_ids=[0], model="dexgangrasp", is_wgan=False, bce_weight=1.0,
           train_ffhgenerator=True, train_ffhevaluator=True, beta1=0.9, weight_decay=5e-4, lr=1e-4, lr_gen=1e-4, lr_dis=1e-4,
           lr_policy="plateau", threshold_lr_policy_plateau=0.01, patience_lr_policy_plateau=5, weight_init_type="xavier",
           init_gain=0.02)


def _model(precision):
    from dex_gan_grasp_b200.dexgangrasp import DexGanGrasp
    m = DexGanGrasp(dict(CFG, precision=precision))
    m.DexGANGrasp.load_state_dict(synth.synth_gan_state_dict(SEED))
    m.DexEvaluator.load_state_dict(synth.synth_eva_state_dict(SEED))
    return m


# fp32 mode at B = 4096: ~1e-6 agreement except where a LeakyReLU pre-activation lies within fp32 rounding of zero (its slope, 1
# vs 0.2, then legitimately differs between two correct fp32 implementations - see tests/test_train_gpu.py::grad_close); with
# 21x more rows than the B = 192 tests a few more such rows exist, worst seen 1.1e-2 in the relative L2 norm of one tensor (generator.gen_rot.weight)
FP32_GRAD_TOL = tolerances.FP32_TRAIN_GRAD_TOL_C5


def _bf16_tol(k, g):
    if "gen_rot." in k:
        return tolerances.BF16_TRAIN_GRAD_TOL_ROT_HEAD
    return tolerances.BF16_TRAIN_GRAD_TOL_SMALL if g.numel() < 64 else tolerances.BF16_TRAIN_GRAD_TOL


def _rel(got, ref):
    ref = ref.double()
    return float((got.detach().cpu().double() - ref).norm()) / max(float(ref.norm()), 1e-30)


@pytest.fixture(scope="module")
def oracle_c5():
    """One oracle GAN step + one evaluator step at B = 4096 (a few seconds of CPU), shared by the fp32 and bf16 tests."""
    batch = synth.synth_train_batch(B, seed=11)
    z = synth.synth_latents(B, 5, seed=3)
    sd = {k: v.clone() for k, v in synth.synth_gan_state_dict(SEED).items()}
    losses, grads = no.gan_step(sd, {}, {}, copy.deepcopy(batch), z.clone(), is_train_gen=True)
    g = torch.Generator().manual_seed(5)
    masks = [(torch.rand(B, 512, generator=g) >= 0.3).float() / 0.7 for _ in range(3)]
    sde = {k: v.clone() for k, v in synth.synth_eva_state_dict(SEED).items()}
    e_losses, e_grads = no.evaluator_step(sde, {}, copy.deepcopy(batch), dropout_masks=masks)
    return dict(batch=batch, z=z, sd=sd, losses=losses, grads=grads, masks=masks, sde=sde, e_losses=e_losses, e_grads=e_grads)


def _skip_grad(k, g):
    return k.endswith(("fc1.bias", "fc2.bias")) or float(g.norm()) < 1e-6  # bias in front of a train-mode BN: true gradient 0


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_c5_gan_step_vs_oracle(oracle_c5, precision):
    from dex_gan_grasp_b200 import train_ops
    o = oracle_c5
    m = _model(precision)
    try:
        ours = m.update_dexgangrasp(copy.deepcopy(o["batch"]), is_train_gen=True, Zgen=o["z"].cuda())
        torch.cuda.synchronize()
    finally:
        train_ops.set_precision("fp32")
    ltol = 1e-4 if precision == "fp32" else tolerances.BF16_TRAIN_LOSS_TOL
    for k, v in o["losses"].items():
        assert abs(float(ours[k]) - v) < ltol * max(1.0, abs(v)), (precision, k, float(ours[k]), v)
    named = dict(m.DexGANGrasp.named_parameters())
    worst = {}
    for net in ("d", "g"):
        num = den = 0.0
        for k, g in o["grads"][net].items():
            if _skip_grad(k, g):
                continue
            r = _rel(named[k].grad, g)
            worst[k] = r
            tol = FP32_GRAD_TOL if precision == "fp32" else _bf16_tol(k, g)
            assert r < tol, (precision, net, k, r)
            num += float((named[k].grad.detach().cpu().double() - g.double()).pow(2).sum())
            den += float(g.double().pow(2).sum())
        if precision == "bf16":
            assert (num / den) ** 0.5 < tolerances.BF16_TRAIN_GRAD_TOL_ALL, (net, (num / den) ** 0.5)
    print(precision, "C5 GAN step: worst per-tensor relative gradient errors:",
          sorted(((round(v, 4), k) for k, v in worst.items()), reverse=True)[:5])
    # BatchNorm running statistics (4 updates of disc_bn1 etc.) and Adam's first step
    w0 = synth.synth_gan_state_dict(SEED)
    for k, v in m.DexGANGrasp.state_dict().items():
        if "running_" in k:
            tol = 5e-5 if precision == "fp32" else 1e-2  # fp32: 1e-5 per update at B = 192, three updates of the discriminator's buffers here
            assert float((v.cpu() - o["sd"][k]).abs().max()) < tol * max(1.0, float(o["sd"][k].abs().max())), k
        elif v.is_floating_point():
            assert float((v.cpu() - w0[k]).abs().max()) <= 1.05e-4, k  # |Adam step 1| <= lr


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_c5_evaluator_step_vs_oracle(oracle_c5, precision):
    from dex_gan_grasp_b200 import train_ops
    o = oracle_c5
    m = _model(precision)
    m.DexEvaluator.dropout_masks = [x.cuda() for x in o["masks"]]
    try:
        ours = m.update_dexevaluator(copy.deepcopy(o["batch"]))
        torch.cuda.synchronize()
    finally:
        train_ops.set_precision("fp32")
        m.DexEvaluator.dropout_masks = None
    ltol = 1e-4 if precision == "fp32" else tolerances.BF16_TRAIN_LOSS_TOL
    for k, v in o["e_losses"].items():
        assert abs(float(ours[k]) - v) < ltol * max(1.0, abs(v)), (precision, k, float(ours[k]), v)
    named = dict(m.DexEvaluator.named_parameters())
    for k, g in o["e_grads"].items():
        if _skip_grad(k, g):
            continue
        r = _rel(named[k].grad, g)
        tol = FP32_GRAD_TOL if precision == "fp32" else _bf16_tol(k, g)
        assert r < tol, (precision, k, r)


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_eval_helpers_vs_oracle(precision):
    """eval_dexevaluator_loss / eval_dexevaluator_accuracy / eval_dexgangrasp_generator_loss (dexgangrasp.py:309-365)
    against the oracle restatement on a 512-row batch (closes SURVEY row a11)."""
    m = _model(precision)
    batch = synth.synth_train_batch(512, seed=13)
    sd_gan, sd_eva = synth.synth_gan_state_dict(SEED), synth.synth_eva_state_dict(SEED)
    stol = tolerances.FP32_TOL if precision == "fp32" else tolerances.BF16_SCORE_TOL
    ref_l, ref_p = no.eval_evaluator_loss(sd_eva, batch)
    ours = m.eval_dexevaluator_loss(copy.deepcopy(batch))
    assert set(ours) == {"total_loss_eva", "bce_loss"}
    assert abs(float(ours["total_loss_eva"]) - ref_l["total_loss_eva"]) < 10 * stol, (float(ours["total_loss_eva"]), ref_l)
    pos, neg, pred, gt = m.eval_dexevaluator_accuracy(copy.deepcopy(batch))  # after eval(): the call order of train.py:75-100
    rpos, rneg, rpred, rgt, rp = no.eval_evaluator_accuracy(sd_eva, batch)
    assert pred.shape == rpred.shape == (512, 1) and np.array_equal(gt, rgt)
    sure = np.abs(rp - 0.5) > stol  # rows whose oracle score is within the tolerance of the 0.5 threshold may flip
    assert np.array_equal(pred[sure], rpred[sure])
    n_flip = int((pred != rpred).sum())
    assert abs(float(pos) - rpos) <= n_flip / max(float(gt.sum()), 1.0) + 1e-6
    assert abs(float(neg) - rneg) <= n_flip / max(float((1 - gt).sum()), 1.0) + 1e-6
    torch.manual_seed(77)
    z = torch.randn((512, 5), device="cuda").cpu()  # the draw eval_dexgangrasp_generator_loss makes (dexgangrasp.py:338)
    torch.manual_seed(77)
    ours_g = m.eval_dexgangrasp_generator_loss(copy.deepcopy(batch))
    ref_g = no.eval_generator_loss(sd_gan, batch, z)
    assert set(ours_g) == set(ref_g)
    gtol = 1e-4 if precision == "fp32" else 5e-2
    for k, v in ref_g.items():
        assert abs(float(ours_g[k]) - v) < gtol * max(1.0, abs(v)), (precision, k, float(ours_g[k]), v)


def test_eval_forward_after_graph_replays_uses_trained_weights():
    """after CapturedTrainStep replays (weights and BatchNorm buffers rewritten by the graph, no host-side
    version bump) the next eval-mode call must see the TRAINED weights: compare with a model trained eagerly on the same
    batches.  Also after is_train_gen=False steps (BatchNorm buffers of the generator change without an optimizer step)."""
    from dex_gan_grasp_b200.graph_step import CapturedTrainStep
    Bs = 256
    batches = [{k: v.cuda() for k, v in synth.synth_train_batch(Bs, seed=80 + i).items()} for i in range(4)]
    zs = [synth.synth_latents(Bs, 5, seed=90 + i).cuda() for i in range(4)]
    probe_bps = torch.from_numpy(synth.synth_train_batch(8, seed=99)["bps_object"].numpy())
    probe_z = synth.synth_latents(8 * 16, 5, seed=98)

    def fresh():
        m = _model("fp32")
        m.DexEvaluator.dout.p = 0.0
        return m

    def probe(m):
        rot, transl, joint, _ = m.DexGANGrasp.generator.native_forward(probe_z, probe_bps, 0.025, False)
        p = m.DexEvaluator.native_forward(probe_bps, rot, transl)
        return torch.cat([rot.reshape(-1), transl.reshape(-1), joint.reshape(-1), p.reshape(-1)]).clone()

    eager = fresh()
    probe(eager)  # packs the initial weights (the stale copy a broken sync would keep using)
    for i in [0] + list(range(4)):
        eager.update_dexgangrasp({k: v.clone() for k, v in batches[i].items()}, Zgen=zs[i])
        eager.update_dexevaluator({k: v.clone() for k, v in batches[i].items()})
    want = probe(eager)
    cap = fresh()
    before = probe(cap)
    step = CapturedTrainStep(cap, batches[0], warmup=1, example_z=zs[0])
    for i in range(4):
        step(batches[i], zs[i])
    got = probe(cap)
    assert float((got - before).abs().max()) > 1e-4, "eval forward still uses the weights packed before the replays"
    assert float((got - want).abs().max()) < 5e-3, float((got - want).abs().max())
    step.close()
    # train-mode forward without a generator step: the generator's BatchNorm running statistics move
    m = fresh()
    b0 = probe(m)
    for i in range(3):
        m.update_dexgangrasp({k: v.clone() for k, v in batches[i].items()}, is_train_gen=False, Zgen=zs[i])
    sd = {k: v.detach().cpu().clone() for k, v in m.DexGANGrasp.state_dict().items()}
    with torch.no_grad():
        o = no.generate_poses(sd, probe_z, probe_bps.repeat_interleave(16, 0), 0.025)
    rot, transl, joint, _ = m.DexGANGrasp.generator.native_forward(probe_z, probe_bps, 0.025, False)
    assert float(np.abs(transl.cpu().numpy() - o["transl"]).max()) < tolerances.FP32_TOL * 5
    assert float(np.abs(joint.cpu().numpy() - o["joint_conf"]).max()) < tolerances.FP32_TOL * 5


def test_engine_combined_step_equals_sequential_calls():
    """train_engine.TrainEngine.combined_step (GAN + evaluator update as one schedule, what CapturedTrainStep replays in bf16
    mode) gives what update_dexgangrasp followed by update_dexevaluator gives: same losses, weights within Adam's noise bound."""
    from dex_gan_grasp_b200.graph_step import CapturedTrainStep
    Bs = 256
    batches = [{k: v.cuda() for k, v in synth.synth_train_batch(Bs, seed=140 + i).items()} for i in range(4)]
    zs = [synth.synth_latents(Bs, 5, seed=160 + i).cuda() for i in range(4)]

    def fresh():
        m = _model("bf16")
        m.DexEvaluator.dout.p = 0.0
        return m
    seq = fresh()
    for i in [0] + list(range(4)):
        ls = dict(seq.update_dexgangrasp({k: v.clone() for k, v in batches[i].items()}, Zgen=zs[i]))
        ls.update(seq.update_dexevaluator({k: v.clone() for k, v in batches[i].items()}))
    cap = fresh()
    step = CapturedTrainStep(cap, batches[0], warmup=1, example_z=zs[0])
    for i in range(4):
        lc = step(batches[i], zs[i])
    torch.cuda.synchronize()
    for k in ls:
        assert abs(float(ls[k]) - float(lc[k])) < 5e-3 * max(1.0, abs(float(ls[k]))), (k, float(ls[k]), float(lc[k]))  # two bf16 runs, atomic summation orders
    for name in ("DexGANGrasp", "DexEvaluator"):
        sa, sb = getattr(seq, name).state_dict(), getattr(cap, name).state_dict()
        params = {k for k, _ in getattr(seq, name).named_parameters()}
        for k in sa:
            if not sa[k].is_floating_point():
                assert int(sa[k]) == int(sb[k]), (name, k)
                continue
            d = (sa[k].double() - sb[k].double()).abs()
            if k in params:  # Adam turns rounding noise in a near-zero gradient into up to lr per step (tests/test_train_gpu.py)
                # (the mean bound only means something on tensors with many elements: one noise-driven step of a 6-element bias moves it)
                assert float(d.max()) <= 2.1e-4 * 5 and (d.numel() < 1024 or float(d.mean()) < 1e-4), (name, k, float(d.max()), float(d.mean()))
            else:  # BatchNorm buffers follow the (slightly different) weights of two bf16 runs with atomic summation orders
                assert float(d.max()) < 1e-2 * max(1.0, float(sa[k].abs().max())), (name, k, float(d.max()))
    step.close()


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_captured_step_follows_learning_rate_changes(precision):
    """VERDICT r01 robustness #10: the learning rate is frozen into the captured graph as a kernel argument, so a scheduler
    step (DexGanGrasp.update_learning_rate, once per epoch in train.py) must trigger a re-capture.  With the rate set to 0 the
    next replay must leave every weight where it was."""
    from dex_gan_grasp_b200 import train_ops
    from dex_gan_grasp_b200.graph_step import CapturedTrainStep
    Bs = 256
    batch = {k: v.cuda() for k, v in synth.synth_train_batch(Bs, seed=170).items()}
    m = _model(precision)
    try:
        step = CapturedTrainStep(m, batch, warmup=1)
        step(batch)
        before = [p.detach().clone() for p in m.DexGANGrasp.parameters()] + [p.detach().clone() for p in m.DexEvaluator.parameters()]
        step(batch)
        now = list(m.DexGANGrasp.parameters()) + list(m.DexEvaluator.parameters())
        assert any(float((a - b).abs().max()) > 0 for a, b in zip(before, now))  # training moves the weights
        for o in step.opts:
            for g in o.param_groups:
                g["lr"] = 0.0  # what ReduceLROnPlateau does, taken to the limit
        before = [p.detach().clone() for p in now]
        step(batch)
        torch.cuda.synchronize()
        assert getattr(step, "recaptures", 0) == 1
        assert all(torch.equal(a, b) for a, b in zip(before, now)), "the replay still used the learning rate frozen at capture time"
        step.close()
    finally:
        train_ops.set_precision("fp32")


@pytest.mark.gpu
def test_engine_is_rebuilt_for_another_batch_size_and_after_close():
    """The fused engine is sized for one batch: a different batch (the short last batch of an epoch) closes it - releasing the
    buffers it allocated outside torch's allocator - and builds another; an explicit close() is followed by a fresh engine too.
    Losses stay finite and the parameters keep moving."""
    import torch
    from dex_gan_grasp_b200 import synth
    from dex_gan_grasp_b200.dexgangrasp import DexGanGrasp
    import bench
    m = DexGanGrasp(dict(bench.TRAIN_CFG, gpu_ids=[0], precision="bf16"))
    m.DexGANGrasp.load_state_dict(synth.synth_gan_state_dict(7))
    m.DexEvaluator.load_state_dict(synth.synth_eva_state_dict(7))
    engines = []
    for B in (256, 128, 128):
        batch = {k: v.cuda() for k, v in synth.synth_train_batch(B, seed=B).items()}
        w0 = m.DexEvaluator.out_success.weight.detach().clone()
        out = m.update_dexevaluator(batch)
        out.update(m.update_dexgangrasp(batch, is_train_gen=True))
        torch.cuda.synchronize()
        assert all(torch.isfinite(torch.as_tensor(float(v))) for v in out.values())
        assert not torch.equal(w0, m.DexEvaluator.out_success.weight.detach())
        engines.append(m._train_engine)
        if len(engines) == 2:
            assert engines[0]._closed and engines[0] is not engines[1]
            engines[1].close()  # explicit: the next update builds a new one
    assert engines[1]._closed and engines[2] is not engines[1] and not engines[2]._closed
