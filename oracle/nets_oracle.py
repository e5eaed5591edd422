"""TEST INFRASTRUCTURE — CPU oracle for the DexGenerator / Discriminator / DexEvaluator stages and the
filter.  Never on the product path: only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
`--impl reference` legs may import this.

A plain torch-CPU fp32 restatement (functional, state-dict driven) of the reference's nn.Modules.  It exists
because the reference's own classes cannot travel to the GPU box (/root/reference is absent there) and cannot
even be constructed on CPU without stubbing h5py/open3d/transforms3d and patching a hard-coded `.cuda()`.
Pinned against the real reference classes by tests/golden/make_golden.py (run in the build container, which
imports /root/reference) -> tests/golden/nets_golden.npz, and by tests/test_oracle_nets.py.

Each function cites the reference lines it follows.
"""
import numpy as np
import torch
import torch.nn.functional as F

LRELU = 0.2  # networks.py:41
BN_EPS = 1e-5  # nn.BatchNorm1d default
BN_MOM = 0.1


def _bn(x, sd, p, training):
    """nn.BatchNorm1d (networks.py:33,36,71,117,289).  In training mode updates running stats in `sd` in place."""
    if training:
        out = F.batch_norm(x, sd[p + ".running_mean"], sd[p + ".running_var"], sd[p + ".weight"], sd[p + ".bias"],
                           True, BN_MOM, BN_EPS)
        sd[p + ".num_batches_tracked"] += 1
        return out
    return F.batch_norm(x, sd[p + ".running_mean"], sd[p + ".running_var"], sd[p + ".weight"], sd[p + ".bias"],
                        False, BN_MOM, BN_EPS)


def _lin(x, sd, p):
    return F.linear(x, sd[p + ".weight"], sd[p + ".bias"])


def resblock(x, sd, p, training=False):
    """ResBlock.forward (networks.py:43-56), final_nl=True."""
    fin, fout = sd[p + ".fc1.weight"].shape[1], sd[p + ".fc2.weight"].shape[0]
    xin = x if fin == fout else F.leaky_relu(_lin(x, sd, p + ".fc3"), LRELU)
    h = F.leaky_relu(_bn(_lin(x, sd, p + ".fc1"), sd, p + ".bn1", training), LRELU)
    y = _bn(_lin(h, sd, p + ".fc2"), sd, p + ".bn2", training)
    return F.leaky_relu(xin + y, LRELU)


def generator_forward(sd, z, bps, training=False, prefix="generator."):
    """Generator.forward (networks.py:88-103): z FIRST in X0, X0 FIRST in the rb2 input."""
    o = _bn(bps.contiguous(), sd, prefix + "gen_bn1", training)
    x0 = torch.cat([z, o], dim=1)
    x = resblock(x0, sd, prefix + "gen_rb1", training)
    x = resblock(torch.cat([x0, x], dim=1), sd, prefix + "gen_rb2", training)
    return {
        "rot_6D": _lin(x, sd, prefix + "gen_rot"),
        "transl": _lin(x, sd, prefix + "gen_transl"),
        "joint_conf": _lin(x, sd, prefix + "gen_joint_conf"),
        "z": z,
    }


def normalize_vector(v):
    """utils.normalize_vector (utils.py:323-332) without its hard-coded `.cuda()` (utils.py:326)."""
    mag = torch.sqrt(v.pow(2).sum(1))
    mag = torch.max(mag, torch.tensor([1e-8], dtype=v.dtype, device=v.device))
    return v / mag.view(-1, 1)


def cross_product(u, v):
    """utils.cross_product (utils.py:68-78)."""
    i = u[:, 1] * v[:, 2] - u[:, 2] * v[:, 1]
    j = u[:, 2] * v[:, 0] - u[:, 0] * v[:, 2]
    k = u[:, 0] * v[:, 1] - u[:, 1] * v[:, 0]
    return torch.stack([i, j, k], dim=1)


def rot_matrix_from_ortho6d(r6):
    """utils.rot_matrix_from_ortho6d (utils.py:380-394): columns (x, y, z)."""
    x = normalize_vector(r6[:, 0:3])
    z = normalize_vector(cross_product(x, r6[:, 3:6]))
    y = cross_product(z, x)
    return torch.stack([x, y, z], dim=2)


def translate_along_axis(transl, rot, axis, offset):
    """utils.translate_along_axis (utils.py:19-48) vectorised: float64 `t - R @ e_axis*offset`, stored back as f32."""
    t = np.asarray(transl, dtype=np.float32)
    r = np.asarray(rot, dtype=np.float32)
    e = np.zeros(3, dtype=np.float64)
    e[axis] = offset
    out = t.astype(np.float64) - np.einsum("nij,j->ni", r.astype(np.float64), e)
    return out.astype(np.float32)


def generate_poses(sd, z, bps, z_offset=0.0):
    """networks.DexGANGrasp.generate_poses (networks.py:223-255) with z supplied by the caller (eval mode)."""
    with torch.no_grad():
        out = generator_forward(sd, z, bps, training=False)
        rot = rot_matrix_from_ortho6d(out.pop("rot_6D"))
    res = {k: v.numpy().copy() for k, v in out.items()}
    res["rot_matrix"] = rot.numpy().copy()
    res["transl"] = translate_along_axis(res["transl"], res["rot_matrix"], 0, z_offset)
    return res


def evaluator_forward(sd, bps, rot, transl, training=False, dropout_masks=None):
    """DexEvaluator.set_input/forward (networks.py:307-337): X FIRST in the rb2/rb3 inputs, joint_conf unused.
    dropout_masks: optional 3 pre-scaled keep masks (train mode); None -> dropout inactive."""
    m = bps.shape[0]
    x = torch.cat([bps, rot.reshape(m, -1), transl], dim=1)
    x0 = _bn(x, sd, "bn1", training)
    x = resblock(x0, sd, "rb1", training)
    for i, rb in enumerate(("rb2", "rb3")):
        if dropout_masks is not None:
            x = x * dropout_masks[i]
        x = resblock(torch.cat([x, x0], dim=1), sd, rb, training)
    if dropout_masks is not None:
        x = x * dropout_masks[2]
    return torch.sigmoid(_lin(x, sd, "out_success"))


def discriminator_forward(sd, bps, rot, transl, joint, training=False, prefix="discriminator."):
    """Discriminator.set_input/forward (networks.py:132-159): X0 FIRST in the rb2 input."""
    m = bps.shape[0]
    x = torch.cat([bps, rot.reshape(m, -1), transl, joint], dim=1)
    x0 = _bn(x, sd, prefix + "disc_bn1", training)
    x = resblock(x0, sd, prefix + "disc_rb1", training)
    x = resblock(torch.cat([x0, x], dim=1), sd, prefix + "disc_rb2", training)
    return torch.sigmoid(_lin(x, sd, prefix + "out_success"))


def filter_order(p, thresh):
    """DexGanGrasp.filter_grasps (dexgangrasp.py:412-419): sort scores descending, keep `> thresh` (strict).
    torch's sort leaves the order of equal scores unspecified; the oracle (and the CUDA path) break ties by the
    lower row index.  Returns (kept_indices int64, raises ValueError like :414-416 when best < thresh)."""
    p = np.asarray(p, dtype=np.float32).reshape(-1)
    order = np.argsort(-p.astype(np.float64), kind="stable")
    if len(p) == 0 or p[order[0]] < thresh:
        raise ValueError("best score lower than thresh")
    return order[p[order] > thresh]


def segmented_topk(p, seg_offsets, thresh, k):
    """Per-object variant used by the batched path: for each segment, indices (global row ids) of the up-to-k best
    rows with p > thresh, best first, ties by lower row index.  Returns (idx [n_seg,k] int32 padded -1, count)."""
    p = np.asarray(p, dtype=np.float32).reshape(-1)
    n_seg = len(seg_offsets) - 1
    idx = np.full((n_seg, k), -1, dtype=np.int32)
    cnt = np.zeros(n_seg, dtype=np.int32)
    for s in range(n_seg):
        a, b = int(seg_offsets[s]), int(seg_offsets[s + 1])
        seg = p[a:b]
        order = np.argsort(-seg.astype(np.float64), kind="stable")
        order = order[seg[order] > thresh][:k]
        cnt[s] = len(order)
        idx[s, :len(order)] = order + a
    return idx, cnt


def filter_grasps(sd_eva, bps_row, grasps, thresh):
    """DexGanGrasp.filter_grasps (dexgangrasp.py:393-442) on numpy inputs; returns dict of numpy like the reference."""
    n = grasps["rot_matrix"].shape[0]
    bps = np.tile(np.asarray(bps_row, dtype=np.float32).reshape(-1), (n, 1))
    g = dict(grasps)
    g["bps_object"] = bps
    with torch.no_grad():
        p = evaluator_forward(sd_eva, torch.from_numpy(bps), torch.from_numpy(np.asarray(g["rot_matrix"], np.float32)),
                              torch.from_numpy(np.asarray(g["transl"], np.float32))).squeeze(-1).numpy()
    keep = filter_order(p, thresh)
    return {k: np.asarray(v)[keep] for k, v in g.items()}, p


# ---------------------------------------------------------------------------------------------------------
# training steps (dexgangrasp.py:739-823) — torch autograd on the functional forward, Adam restated.
# ---------------------------------------------------------------------------------------------------------
def _params(sd, prefix=""):
    return [k for k, v in sd.items() if k.startswith(prefix) and v.is_floating_point()
            and not k.endswith(("running_mean", "running_var"))]


def adam_step(sd, grads, state, lr=1e-4, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=5e-4):
    """torch.optim.Adam semantics (dexgangrasp.py:97-112): coupled L2 (g += wd*w), bias-corrected, eps outside sqrt."""
    state["step"] = state.get("step", 0) + 1
    t = state["step"]
    for k, g in grads.items():
        w = sd[k]
        g = g + weight_decay * w
        m = state.setdefault("m." + k, torch.zeros_like(w))
        v = state.setdefault("v." + k, torch.zeros_like(w))
        m.mul_(beta1).add_(g, alpha=1 - beta1)
        v.mul_(beta2).addcmul_(g, g, value=1 - beta2)
        denom = (v.sqrt() / np.sqrt(1 - beta2 ** t)).add_(eps)
        w.addcdiv_(m, denom, value=-lr / (1 - beta1 ** t))


def bce(p, target_value):
    """nn.BCELoss(reduction='mean') against a constant target (dexgangrasp.py:122,215-223,293); log clamped at -100."""
    if target_value == 1.0:
        return -torch.clamp(torch.log(p), min=-100.0).mean()
    return -torch.clamp(torch.log(1.0 - p), min=-100.0).mean()


def gan_step(sd, opt_d, opt_g, batch, z, is_train_gen=True, bce_weight=1.0, lr=1e-4, beta1=0.9, wd=5e-4):
    """DexGanGrasp.update_dexgangrasp (dexgangrasp.py:757-823), non-WGAN branch.  `sd` (container state dict) and the
    optimizer states are updated in place; returns the loss dict (python floats) and the D/G gradients."""
    bps = batch["bps_object"].float()
    rot = batch["rot_matrix"].float()
    transl = batch["transl"].float()
    joint = batch["joint_conf"].float()
    gk = _params(sd, "generator.")
    dk = _params(sd, "discriminator.")
    for k in gk + dk:
        sd[k].requires_grad_(True)
    try:
        fake = generator_forward(sd, z, bps, training=True)  # :772
        fake_rot = rot_matrix_from_ortho6d(fake["rot_6D"])  # :776
        real_score = discriminator_forward(sd, bps, rot, transl, joint, training=True)  # :788
        fake_score = discriminator_forward(sd, bps, fake_rot.detach(), fake["transl"].detach(),
                                           fake["joint_conf"].detach(), training=True)  # :789
        l_real = bce_weight * bce(real_score, 1.0)
        l_fake = bce_weight * bce(fake_score, 0.0)
        l_d = (l_real + l_fake) / 2  # :225
        gd = torch.autograd.grad(l_d, [sd[k] for k in dk])
        for k in gk + dk:
            sd[k].requires_grad_(False)
        adam_step(sd, dict(zip(dk, gd)), opt_d, lr, beta1, 0.999, 1e-8, wd)  # :792-794
        losses = {"total_loss_disc": l_d.item(), "bce_loss_real": l_real.item(), "bce_loss_fake": l_fake.item()}
        grads = {"d": dict(zip(dk, gd))}
        if is_train_gen:
            for k in gk:
                sd[k].requires_grad_(True)
            score_gen = discriminator_forward(sd, bps, fake_rot, fake["transl"], fake["joint_conf"], training=True)  # :808
            l_gen_fake = bce_weight * bce(score_gen, 1.0)  # :293
            l_t = 100.0 * F.mse_loss(fake["transl"], transl)  # :92, losses.py:74
            l_r = 1.0 * F.mse_loss(fake_rot.reshape(-1, 9), rot.reshape(-1, 9))  # losses.py:67-72
            l_j = 10.0 * F.mse_loss(fake["joint_conf"], joint)  # :284
            l_g = l_gen_fake + (l_t + l_r + l_j)  # :304
            gg = torch.autograd.grad(l_g, [sd[k] for k in gk])
            for k in gk:
                sd[k].requires_grad_(False)
            adam_step(sd, dict(zip(gk, gg)), opt_g, lr, beta1, 0.999, 1e-8, wd)  # :814-816
            losses.update({"gen_loss_fake": l_gen_fake.item(), "transl_loss": l_t.item(), "rot_loss": l_r.item(),
                           "conf_loss": l_j.item(), "total_loss_gen": l_g.item()})
            grads["g"] = dict(zip(gk, gg))
        return losses, grads
    finally:
        for k in gk + dk:
            sd[k].requires_grad_(False)


def gradient_penalty(sd, bps, real, fake, alphas, prefix="discriminator."):
    """DexGanGrasp.calculate_gradient_penalty (dexgangrasp.py:175-202) with calculate_interp (:149-158; alpha is drawn
    with randn, NOT rand — reference quirk, kept) and calculate_grad_interp (:160-173).  `real` holds the batch's own
    dtypes (float64 from the DataLoader), so the interpolates and their gradients are float64 like in the reference;
    the discriminator casts to float32 (networks.py:136-139).  alphas = (a_rot [B,1,1], a_transl [B,1], a_joint [B,1])."""
    names = ("rot_matrix", "transl", "joint_conf")
    interp = {}
    for n, a in zip(names, alphas):
        interp[n] = (a * real[n] + ((1 - a) * fake[n])).requires_grad_(True)  # :157
    score = discriminator_forward(sd, bps, interp["rot_matrix"].float(), interp["transl"].float(),
                                  interp["joint_conf"].float(), training=True, prefix=prefix)  # :194
    pens = []
    for n in names:  # :197-200
        g = torch.autograd.grad(score, interp[n], torch.ones_like(score), create_graph=True, retain_graph=True)[0]
        g = g.view(g.size(0), -1)
        pens.append(torch.mean((g.norm(2, dim=1) - 1) ** 2))
    return torch.mean(torch.stack(pens))  # :202


def wgan_step(sd, opt_d, opt_g, batch, z, alphas, is_train_gen=True, penalty_gain=10.0, lr=1e-4, beta1=0.9, wd=5e-4):
    """DexGanGrasp.update_dexgangrasp (dexgangrasp.py:757-823), `is_wgan` branch: discriminator loss
    -mean(D(real)) + mean(D(fake)) + 10*GP (:204-218; D keeps its sigmoid), generator loss -mean(D(fake)) + pose terms
    (:234-270).  Same in-place conventions as gan_step."""
    bps = batch["bps_object"].float()
    rot = batch["rot_matrix"].float()
    transl = batch["transl"].float()
    joint = batch["joint_conf"].float()
    gk = _params(sd, "generator.")
    dk = _params(sd, "discriminator.")
    for k in gk + dk:
        sd[k].requires_grad_(True)
    try:
        fake = generator_forward(sd, z, bps, training=True)  # :772
        fake_rot = rot_matrix_from_ortho6d(fake["rot_6D"])  # :776
        fake_d = {"rot_matrix": fake_rot.detach(), "transl": fake["transl"].detach(), "joint_conf": fake["joint_conf"].detach()}
        real_score = discriminator_forward(sd, bps, rot, transl, joint, training=True)  # :788
        fake_score = discriminator_forward(sd, bps, fake_d["rot_matrix"], fake_d["transl"], fake_d["joint_conf"], training=True)
        l_real = torch.mean(real_score)  # :206
        l_fake = torch.mean(fake_score)
        gp = gradient_penalty(sd, bps, batch, fake_d, alphas)  # :208
        l_d = -l_real + l_fake + gp * penalty_gain  # :209
        gd = torch.autograd.grad(l_d, [sd[k] for k in dk])
        for k in gk + dk:
            sd[k].requires_grad_(False)
        adam_step(sd, dict(zip(dk, gd)), opt_d, lr, beta1, 0.999, 1e-8, wd)
        losses = {"total_loss_disc": l_d.item(), "bce_loss_real": l_real.item(), "bce_loss_fake": l_fake.item(),
                  "gradient_penalty": gp.item()}
        grads = {"d": dict(zip(dk, gd))}
        if is_train_gen:
            for k in gk:
                sd[k].requires_grad_(True)
            score_gen = discriminator_forward(sd, bps, fake_rot, fake["transl"], fake["joint_conf"], training=True)  # :808
            l_gen_fake = -torch.mean(score_gen)  # :258
            l_t = 100.0 * F.mse_loss(fake["transl"], transl)
            l_r = 1.0 * F.mse_loss(fake_rot.reshape(-1, 9), rot.reshape(-1, 9))
            l_j = 10.0 * F.mse_loss(fake["joint_conf"], joint)
            l_g = l_gen_fake + (l_t + l_r + l_j)  # :266
            gg = torch.autograd.grad(l_g, [sd[k] for k in gk])
            for k in gk:
                sd[k].requires_grad_(False)
            adam_step(sd, dict(zip(gk, gg)), opt_g, lr, beta1, 0.999, 1e-8, wd)
            losses.update({"gen_loss_fake": l_gen_fake.item(), "transl_loss": l_t.item(), "rot_loss": l_r.item(),
                           "conf_loss": l_j.item(), "total_loss_gen": l_g.item()})
            grads["g"] = dict(zip(gk, gg))
        return losses, grads
    finally:
        for k in gk + dk:
            sd[k].requires_grad_(False)


def evaluator_step(sd, opt, batch, dropout_masks=None, bce_weight=1.0, lr=1e-4, beta1=0.9, wd=5e-4):
    """DexGanGrasp.update_dexevaluator (dexgangrasp.py:739-755)."""
    keys = _params(sd)
    for k in keys:
        sd[k].requires_grad_(True)
    try:
        p = evaluator_forward(sd, batch["bps_object"].float(), batch["rot_matrix"].float(), batch["transl"].float(),
                              training=True, dropout_masks=dropout_masks)
        label = batch["label"].float().unsqueeze(-1)  # networks.py:311-312
        loss = bce_weight * F.binary_cross_entropy(p, label)  # :142-147
        g = torch.autograd.grad(loss, [sd[k] for k in keys])
    finally:
        for k in keys:
            sd[k].requires_grad_(False)
    adam_step(sd, dict(zip(keys, g)), opt, lr, beta1, 0.999, 1e-8, wd)
    return {"total_loss_eva": loss.item(), "bce_loss": loss.item()}, dict(zip(keys, g))


# ---------------------------------------------------------------------------------------------------------
# evaluation helpers (dexgangrasp.py:309-365)
# ---------------------------------------------------------------------------------------------------------
def eval_evaluator_loss(sd, batch, bce_weight=1.0):
    """eval_dexevaluator_loss (dexgangrasp.py:325-332): eval-mode forward, BCE against the labels (:142-147)."""
    with torch.no_grad():
        p = evaluator_forward(sd, batch["bps_object"].float(), batch["rot_matrix"].float(), batch["transl"].float(), training=False)
        loss = bce_weight * F.binary_cross_entropy(p, batch["label"].float().unsqueeze(-1))
    return {"total_loss_eva": loss.item(), "bce_loss": loss.item()}, p


def eval_evaluator_accuracy(sd, batch, logit_thresh=0.5):
    """eval_dexevaluator_accuracy (dexgangrasp.py:309-322) after eval(): scores thresholded at 0.5 (utils.py:60-64), then
    losses.accuracy_evaluator (losses.py:5-16): TP / P, TN / N.  Returns (pos_acc, neg_acc, pred [B,1], gt [B,1], scores)."""
    with torch.no_grad():
        p = evaluator_forward(sd, batch["bps_object"].float(), batch["rot_matrix"].float(), batch["transl"].float(), training=False)
    pred = (p >= logit_thresh).float()
    gt = batch["label"].float().unsqueeze(-1)
    correct = torch.eq(pred, gt)
    pos = torch.sum(correct * gt) / torch.sum(gt)
    neg = torch.sum(correct * (1.0 - gt)) / torch.sum(1.0 - gt)
    return pos.item(), neg.item(), pred.numpy(), gt.numpy(), p.numpy()


def eval_generator_loss(sd, batch, z, bce_weight=1.0):
    """eval_dexgangrasp_generator_loss (dexgangrasp.py:334-365), non-WGAN: eval-mode G and D, the generator loss terms of
    compute_loss_dexgangrasp_generator (:272-307) under no_grad."""
    bps = batch["bps_object"].float()
    rot = batch["rot_matrix"].float()
    transl = batch["transl"].float()
    joint = batch["joint_conf"].float()
    with torch.no_grad():
        fake = generator_forward(sd, z, bps, training=False)
        fake_rot = rot_matrix_from_ortho6d(fake["rot_6D"])
        score = discriminator_forward(sd, bps, fake_rot, fake["transl"], fake["joint_conf"], training=False)
        l_fake = bce_weight * bce(score, 1.0)
        l_t = 100.0 * F.mse_loss(fake["transl"], transl)
        l_r = 1.0 * F.mse_loss(fake_rot.reshape(-1, 9), rot.reshape(-1, 9))
        l_j = 10.0 * F.mse_loss(fake["joint_conf"], joint)
    return {"gen_loss_fake": l_fake.item(), "transl_loss": l_t.item(), "rot_loss": l_r.item(), "conf_loss": l_j.item(),
            "total_loss_gen": (l_fake + (l_t + l_r + l_j)).item()}


def refine_grasps_gradient(sd_eva, data, num_refine_steps=10):
    """refine_grasps(refine_method='gradient') (dexgangrasp.py:619-660) -> improve_grasps_gradient_based (:516-541):
    every step evaluates the eval-mode evaluator, back-propagates ones through p_success to the rot_matrix / transl
    leaves WITHOUT zeroing their .grad (the reference never does: step i moves along the sum of the gradients of steps
    0..i), alpha = min(0.01 / ||transl.grad||_2, 1) per grasp, transl += grad*alpha, rot_matrix += grad*alpha.
    joint_conf is not an evaluator input (networks.py:307-313): its gradient is identically zero and it stays as given
    (the reference raises TypeError at :539 because the .grad it reads there is None; tests/golden/make_refine_golden.py
    runs the reference's method with that .grad pre-set to zeros).  Returns (refined_data, refined_success) as
    refine_grasps does: num_refine_steps + 1 entries each."""
    bps = data["bps_object"].float()
    rot = data["rot_matrix"].float().clone().requires_grad_(True)
    transl = data["transl"].float().clone().requires_grad_(True)
    joint = data["joint_conf"].float().clone()

    def snap():
        return {"rot_matrix": rot.detach().numpy().copy(), "transl": transl.detach().numpy().copy(),
                "joint_conf": joint.numpy().copy()}
    refined_data, refined_success = [snap()], []
    for _ in range(num_refine_steps):
        p = evaluator_forward(sd_eva, bps, rot, transl, training=False)
        p.squeeze(-1).backward(torch.ones(p.shape[0]))
        alpha = torch.min(0.01 / torch.norm(transl.grad, p=2, dim=-1), torch.tensor(1.0))
        transl.data += transl.grad * alpha[:, None]
        rot.data += rot.grad * alpha[:, None, None]
        refined_success.append(p.detach().squeeze().numpy().copy())
        refined_data.append(snap())
    with torch.no_grad():
        refined_success.append(evaluator_forward(sd_eva, bps, rot, transl, training=False).squeeze().numpy().copy())
    return refined_data, refined_success
