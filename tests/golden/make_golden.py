#!/usr/bin/env python
"""Generates the committed golden fixtures from the LIVE reference in /root/reference.

Run in the build container only:  python tests/golden/make_golden.py
(`/root/reference` does not exist on the GPU box; nothing at test/bench time reads it.)

What it does
  1. bps_real_objects.npz — for the 7 real clouds (data/real_objects/object/<name>.pcd) a REDUCED cloud that
     still reproduces the stored encoding data/real_objects/bps/<name>.npy bit-exactly: all nearest-neighbour
     winners plus 3000 random other points, original order kept.  Plus the stored encodings themselves.
  2. eval_batch.npz — the tensors of data/eval_batch.pth (the reference's own saved DataLoader batch).
  3. nets_golden.npz — outputs of the reference's OWN nn.Modules (DexGanGrasp/models/networks.py, imported with
     empty stubs for the absent h5py/open3d/transforms3d and `utils.normalize_vector` re-stated without its
     hard-coded `.cuda()`), loaded with the seeded synthetic state dicts of dex_gan_grasp_b200/synth.py:
     generator / rot6d / palm offset / evaluator / discriminator forwards on eval_batch and on a shared-BPS
     object batch; one GAN training step and one evaluator step driven exactly like
     DexGanGrasp.update_dexgangrasp / update_dexevaluator (dexgangrasp.py:739-823) with torch.optim.Adam.
  4. asserts that oracle/nets_oracle.py and oracle/bps_oracle.c agree with all of the above.
"""
import copy
import ctypes
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
pcd = _load("dexgg_pcd", os.path.join(REPO, "oracle", "pcd.py"))

# --- import the reference (its package name is DexGanGrasp; the repo's alias package must not shadow it) ---
for m in ("h5py", "open3d", "transforms3d"):
    sys.modules[m] = types.ModuleType(m)
sys.path.insert(0, REF)
from DexGanGrasp.models import networks as ref_networks  # noqa: E402
from DexGanGrasp.models import losses as ref_losses  # noqa: E402
from DexGanGrasp.utils import utils as ref_utils  # noqa: E402

assert ref_networks.__file__.startswith(REF)


def _normalize_vector_cpu(v, return_mag=False):
    # utils.py:323-332 minus the `.cuda()` at :326
    batch = v.shape[0]
    v_mag = torch.sqrt(v.pow(2).sum(1))
    v_mag = torch.max(v_mag, torch.FloatTensor([1e-8]))
    v_mag = v_mag.view(batch, 1).expand(batch, v.shape[1])
    v = v / v_mag
    return (v, v_mag[:, 0]) if return_mag else v


ref_utils.normalize_vector = _normalize_vector_cpu

CFG = dict(latentD=5, n_hand_joints=12, is_train=False, gpu_ids=[0])
SEED = 7


def vp(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def make_bps():
    lib = ctypes.CDLL(os.path.join(REPO, "oracle", "libdexgg_oracle.so"))
    basis = np.load(os.path.join(REF, "models/basis_point_set.npy")).astype(np.float32)
    out = {"basis": basis}
    names = sorted(f[:-4] for f in os.listdir(os.path.join(REF, "data/real_objects/object")) if f.endswith(".pcd"))
    out["names"] = np.array(names)
    rng = np.random.default_rng(0)
    for name in names:
        pts = pcd.read_pcd_xyz(os.path.join(REF, "data/real_objects/object", name + ".pcd"))
        gold = np.load(os.path.join(REF, "data/real_objects/bps", name + ".npy"))
        assert gold.shape == (1, 4096) and gold.dtype == np.float32
        d = np.empty(4096, np.float32)
        idx = np.empty(4096, np.int32)
        lib.dexgg_oracle_bps_encode_mt(vp(basis), 4096, vp(pts), len(pts), vp(d), vp(idx), 8)
        assert np.array_equal(d, gold[0]), name
        win = np.unique(idx)
        others = np.setdiff1d(np.arange(len(pts)), win)
        keep = np.sort(np.concatenate([win, rng.choice(others, 3000, replace=False)]))
        red = np.ascontiguousarray(pts[keep])
        d2 = np.empty(4096, np.float32)
        idx2 = np.empty(4096, np.int32)
        lib.dexgg_oracle_bps_encode_mt(vp(basis), 4096, vp(red), len(red), vp(d2), vp(idx2), 8)
        assert np.array_equal(d2, gold[0]) and np.array_equal(keep[idx2], idx), name
        out[name + "/cloud"] = red
        out[name + "/bps"] = gold
        out[name + "/idx"] = idx2
        out[name + "/n_full"] = np.int64(len(pts))
        print(f"bps {name}: full {len(pts)} -> reduced {len(red)} points, bit-exact")
    np.savez_compressed(os.path.join(HERE, "bps_real_objects.npz"), **out)


def make_eval_batch():
    b = torch.load(os.path.join(REF, "data/eval_batch.pth"), map_location="cpu")
    np.savez_compressed(os.path.join(HERE, "eval_batch.npz"),
                        bps_object=b["bps_object"].numpy(), rot_matrix=b["rot_matrix"].numpy(),
                        transl=b["transl"].numpy(), joint_conf=b["joint_conf"].numpy(),
                        obj_name=np.array(b["obj_name"]))
    return b


def ref_modules(is_train):
    cfg = dict(CFG, is_train=is_train)
    gan = ref_networks.DexGANGrasp(cfg)
    eva = ref_networks.DexEvaluator(cfg)
    gan.load_state_dict(synth.synth_gan_state_dict(SEED))
    eva.load_state_dict(synth.synth_eva_state_dict(SEED))
    return cfg, gan, eva


def maxdiff(a, b):
    return float(np.max(np.abs(np.asarray(a, np.float64) - np.asarray(b, np.float64))))


def step_close(sd_a, sd_b):
    """Post-Adam-step weights.  The first Adam step moves every weight by ~lr*g/(|g|+eps); where the raw gradient and
    the coupled weight decay cancel (|g + wd*w| ~ eps = 1e-8) the step is ill-conditioned, so: 99.99 % of the
    elements within 2e-6 and none further than one full step (lr = 1e-4)."""
    worst, bad, tot = 0.0, 0, 0
    for k, v in sd_b.items():
        if not v.is_floating_point():
            continue
        d = (sd_a[k].double() - v.double()).abs()
        worst = max(worst, float(d.max()))
        bad += int((d > 2e-6).sum())
        tot += d.numel()
    assert worst < 1.01e-4 and bad <= 1e-4 * tot, (worst, bad, tot)
    return worst, bad, tot


def make_nets(batch):
    out = {}
    cfg, gan, eva = ref_modules(False)
    gan.eval()
    eva.eval()
    sd_gan = synth.synth_gan_state_dict(SEED)
    sd_eva = synth.synth_eva_state_dict(SEED)

    # ---- A. row-wise (every row its own BPS): eval_batch, 64 rows ----
    bps = batch["bps_object"].float()
    z = synth.synth_latents(bps.shape[0], 5, seed=3)
    with torch.no_grad():
        g = gan.generator(z, bps)
        rot = ref_utils.rot_matrix_from_ortho6d(g["rot_6D"])
        res = {"rot_matrix": rot.numpy().copy(), "transl": g["transl"].numpy().copy(),
               "joint_conf": g["joint_conf"].numpy().copy()}
        res = ref_utils.translate_along_axis(res, 0, 0.025)
        p_e = eva({"bps_object": bps, "rot_matrix": torch.from_numpy(res["rot_matrix"]),
                   "transl": torch.from_numpy(res["transl"])})
        p_d_fake = gan.discriminator({"bps_object": bps, "rot_matrix": torch.from_numpy(res["rot_matrix"]),
                                      "transl": torch.from_numpy(res["transl"]),
                                      "joint_conf": torch.from_numpy(res["joint_conf"])})
        p_d_real = gan.discriminator({k: batch[k] for k in ("bps_object", "rot_matrix", "transl", "joint_conf")})
        p_e_real = eva({k: batch[k] for k in ("bps_object", "rot_matrix", "transl")})
    out.update({"A/z": z.numpy(), "A/rot_6D": g["rot_6D"].numpy(), "A/rot_matrix": res["rot_matrix"],
                "A/transl": res["transl"], "A/joint_conf": res["joint_conf"], "A/p_eva": p_e.numpy(),
                "A/p_disc_fake": p_d_fake.numpy(), "A/p_disc_real": p_d_real.numpy(), "A/p_eva_real": p_e_real.numpy()})
    o = nets_oracle.generate_poses(sd_gan, z, bps, 0.025)
    for k in ("rot_matrix", "transl", "joint_conf"):
        assert maxdiff(o[k], res[k]) < 2e-6, (k, maxdiff(o[k], res[k]))
    with torch.no_grad():
        po = nets_oracle.evaluator_forward(sd_eva, bps, torch.from_numpy(o["rot_matrix"]), torch.from_numpy(o["transl"]))
        pd = nets_oracle.discriminator_forward(sd_gan, bps, batch["rot_matrix"].float(), batch["transl"].float(),
                                               batch["joint_conf"].float())
    assert maxdiff(po, p_e) < 2e-6 and maxdiff(pd, p_d_real) < 2e-6
    print("A: oracle == reference modules (row-wise), scores span", float(p_e.min()), float(p_e.max()))

    # ---- B. shared-BPS object batch (dexgangrasp_offline.py:147-162 shape): one real BPS row x 256 z ----
    real = np.load(os.path.join(REF, "data/real_objects/bps/red_mug.npy"))
    n = 256
    zb = synth.synth_latents(n, 5, seed=4)
    bps_t = torch.tensor(np.tile(real.squeeze(), (n, 1)), dtype=torch.float32)
    with torch.no_grad():
        g = gan.generator(zb, bps_t)
        rot = ref_utils.rot_matrix_from_ortho6d(g["rot_6D"])
        res = {"rot_matrix": rot.numpy().copy(), "transl": g["transl"].numpy().copy(),
               "joint_conf": g["joint_conf"].numpy().copy(), "z": zb.numpy().copy()}
        res = ref_utils.translate_along_axis(res, 0, 0.025)
        p = eva({"bps_object": bps_t, "rot_matrix": torch.from_numpy(res["rot_matrix"]),
                 "transl": torch.from_numpy(res["transl"])}).squeeze()
        sorted_score, indices = p.sort(descending=True, stable=True)
        kept = indices[sorted_score > 0.5]
        thr_med = float(np.float32(p.median().item()))  # the synthetic evaluator scores one object in a narrow band
        kept_med = indices[sorted_score > thr_med]
    out.update({"B/bps": real, "B/z": zb.numpy(), "B/rot_matrix": res["rot_matrix"], "B/transl": res["transl"],
                "B/joint_conf": res["joint_conf"], "B/p_eva": p.numpy(), "B/kept_0.5": kept.numpy(),
                "B/thresh_med": np.float64(thr_med), "B/kept_med": kept_med.numpy()})
    o = nets_oracle.generate_poses(sd_gan, zb, bps_t, 0.025)
    filt, po = nets_oracle.filter_grasps(sd_eva, real, {k: o[k] for k in ("rot_matrix", "transl", "joint_conf", "z")}, 0.5)
    assert maxdiff(po, p) < 2e-6
    assert np.array_equal(nets_oracle.filter_order(p.numpy(), 0.5), kept.numpy())
    assert np.array_equal(nets_oracle.filter_order(p.numpy(), thr_med), kept_med.numpy()) and 0 < len(kept_med) < n
    assert all(np.array_equal(filt[k], np.asarray(v)[kept.numpy()]) for k, v in res.items())
    print("B: oracle == reference modules (shared BPS); kept", len(kept), "of", n, "scores", float(p.min()), float(p.max()),
          "transl std over z", res["transl"].std(0))

    # ---- C. one GAN step + one evaluator step, train mode, on eval_batch (dexgangrasp.py:739-823) ----
    cfg, gan, eva = ref_modules(True)
    gan.train()
    eva.train()
    eva.dout.p = 0.0  # dropout's Philox stream cannot be shared with an external implementation; parity runs use p=0
    opt_g = torch.optim.Adam(gan.generator.parameters(), lr=1e-4, betas=(0.9, 0.999), weight_decay=5e-4)
    opt_d = torch.optim.Adam(gan.discriminator.parameters(), lr=1e-4, betas=(0.9, 0.999), weight_decay=5e-4)
    opt_e = torch.optim.Adam(eva.parameters(), lr=1e-4, betas=(0.9, 0.999), weight_decay=5e-4)
    bce_loss = torch.nn.BCELoss(reduction="mean")
    l2 = torch.nn.MSELoss(reduction="mean")
    real_data = {k: batch[k].clone() for k in ("bps_object", "rot_matrix", "transl", "joint_conf")}
    zc = synth.synth_latents(64, 5, seed=5)
    y_fake = gan.generator(zc, real_data["bps_object"])
    y_fake["rot_matrix"] = ref_utils.rot_matrix_from_ortho6d(y_fake["rot_6D"])
    fake_disc = {"bps_object": real_data["bps_object"], "rot_matrix": y_fake["rot_matrix"].detach(),
                 "transl": y_fake["transl"].detach(), "joint_conf": y_fake["joint_conf"].detach()}
    real_score = gan.discriminator(real_data)
    fake_score = gan.discriminator(fake_disc)
    l_real = bce_loss(real_score, torch.ones_like(real_score))
    l_fake = bce_loss(fake_score, torch.zeros_like(fake_score))
    l_d = (l_real + l_fake) / 2
    opt_d.zero_grad()
    l_d.backward()
    gd_probe = gan.discriminator.disc_rb2.fc2.weight.grad.clone()
    opt_d.step()
    fake_gen = {"bps_object": real_data["bps_object"], "rot_matrix": y_fake["rot_matrix"], "rot_6D": y_fake["rot_6D"],
                "transl": y_fake["transl"], "joint_conf": y_fake["joint_conf"]}
    score_gen = gan.discriminator(fake_gen)
    real_data["rot_matrix"] = real_data["rot_matrix"].view(64, -1)
    gt = {"transl": real_data["transl"].float(), "rot_matrix": real_data["rot_matrix"].float()}
    l_t, l_r = ref_losses.transl_rot_6D_l2_loss(fake_gen, gt, l2, "cpu")
    l_j = l2(fake_gen["joint_conf"].float(), real_data["joint_conf"].float())
    l_gf = bce_loss(score_gen, torch.ones_like(score_gen))
    l_g = l_gf + (100.0 * l_t + 1.0 * l_r + 10.0 * l_j)
    opt_g.zero_grad()
    l_g.backward()
    gg_probe = gan.generator.gen_rb1.fc1.weight.grad.clone()
    opt_g.step()
    losses_ref = {"total_loss_disc": l_d.item(), "bce_loss_real": l_real.item(), "bce_loss_fake": l_fake.item(),
                  "gen_loss_fake": l_gf.item(), "transl_loss": 100.0 * l_t.item(), "rot_loss": l_r.item(),
                  "conf_loss": 10.0 * l_j.item(), "total_loss_gen": l_g.item()}
    sd_after = {k: v.detach().clone() for k, v in gan.state_dict().items()}

    sd_o = copy.deepcopy(synth.synth_gan_state_dict(SEED))
    losses_o, grads_o = nets_oracle.gan_step(sd_o, {}, {}, batch, zc, True)
    for k, v in losses_ref.items():
        assert abs(losses_o[k] - v) < 1e-5 * max(1.0, abs(v)), (k, losses_o[k], v)
    assert maxdiff(grads_o["d"]["discriminator.disc_rb2.fc2.weight"], gd_probe) < 1e-6
    assert maxdiff(grads_o["g"]["generator.gen_rb1.fc1.weight"], gg_probe) < 1e-5
    print("C: oracle GAN step == reference modules + torch.optim.Adam; (worst, n>2e-6, n)", step_close(sd_o, sd_after))
    out.update({"C/z": zc.numpy()})
    out.update({"C/loss/" + k: np.float64(v) for k, v in losses_ref.items()})
    probes = ["generator.gen_rb1.fc1.weight", "generator.gen_rb2.fc3.bias", "generator.gen_bn1.running_mean",
              "generator.gen_bn1.running_var", "generator.gen_rb2.bn2.running_var", "generator.gen_rot.weight",
              "discriminator.disc_bn1.running_mean", "discriminator.disc_rb1.bn1.running_var",
              "discriminator.disc_rb2.fc2.weight", "discriminator.out_success.weight",
              "discriminator.disc_rb1.fc3.weight"]
    sd0 = synth.synth_gan_state_dict(SEED)
    for k in probes:
        a = sd_after[k].reshape(-1)[:4096].numpy()
        out["C/after/" + k] = a
        out["C/delta/" + k] = (sd_after[k] - sd0[k]).reshape(-1)[:4096].numpy()
    out["C/grad/discriminator.disc_rb2.fc2.weight"] = gd_probe.reshape(-1)[:4096].numpy()
    out["C/grad/generator.gen_rb1.fc1.weight"] = gg_probe.reshape(-1)[:4096].numpy()

    # evaluator step
    g2 = torch.Generator().manual_seed(9)
    label = (torch.rand(64, generator=g2) < 0.3).float()
    data = {"bps_object": batch["bps_object"], "rot_matrix": batch["rot_matrix"], "transl": batch["transl"], "label": label}
    pe = eva(data)
    l_e = bce_loss(pe, eva.gt_label)
    opt_e.zero_grad()
    l_e.backward()
    ge_probe = eva.rb2.fc1.weight.grad.clone()
    opt_e.step()
    sd_e_after = {k: v.detach().clone() for k, v in eva.state_dict().items()}
    sd_eo = copy.deepcopy(synth.synth_eva_state_dict(SEED))
    le_o, ge_o = nets_oracle.evaluator_step(sd_eo, {}, dict(data))
    assert abs(le_o["total_loss_eva"] - l_e.item()) < 1e-5
    assert maxdiff(ge_o["rb2.fc1.weight"], ge_probe) < 1e-6
    print("C: oracle evaluator step == reference; (worst, n>2e-6, n)", step_close(sd_eo, sd_e_after))
    out["C/eva/label"] = label.numpy()
    out["C/eva/loss"] = np.float64(l_e.item())
    out["C/eva/grad/rb2.fc1.weight"] = ge_probe.reshape(-1)[:4096].numpy()
    sde0 = synth.synth_eva_state_dict(SEED)
    for k in ("rb1.fc1.weight", "rb3.fc3.weight", "bn1.running_var", "rb2.bn1.running_mean", "out_success.weight"):
        out["C/eva/after/" + k] = sd_e_after[k].reshape(-1)[:4096].numpy()
        out["C/eva/delta/" + k] = (sd_e_after[k] - sde0[k]).reshape(-1)[:4096].numpy()
    np.savez_compressed(os.path.join(HERE, "nets_golden.npz"), **out)


if __name__ == "__main__":
    torch.manual_seed(0)
    torch.set_num_threads(8)
    make_bps()
    b = make_eval_batch()
    make_nets(b)
    print("golden fixtures written to", HERE)
