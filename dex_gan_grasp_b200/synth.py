"""Seeded synthetic inputs for the grasp-synthesis path.

No trained weights exist anywhere on these boxes (the reference's checkpoints are git-LFS pointers), so every
parity and throughput run uses state dicts generated here: same key layout as the reference's
`networks.DexGANGrasp` / `networks.DexEvaluator` state dicts (DexGanGrasp/models/networks.py:58-337), weights
at 1/sqrt(fan_in) scale and non-trivial BatchNorm running statistics so activations are O(1) and evaluator
scores span (0,1).  Generation is CPU torch with an explicit Generator, so the same seed gives the same
tensors on every box with this image.

Also: synthetic single-view point clouds and GAN training batches of the shapes named in BASELINE.json.
"""
from collections import OrderedDict

import numpy as np
import torch

IN_BPS = 4096
N_NEURONS = 512
RB_INNER = 256


def _lin(g, out_f, in_f, gain=1.0, bias_std=0.05):
    w = torch.randn(out_f, in_f, generator=g, dtype=torch.float32) * (gain / np.sqrt(in_f))
    b = torch.randn(out_f, generator=g, dtype=torch.float32) * bias_std
    return w, b


def _bn(g, n, mean=None, var=None):
    w = 0.8 + 0.4 * torch.rand(n, generator=g, dtype=torch.float32)
    b = 0.1 * torch.randn(n, generator=g, dtype=torch.float32)
    rm = 0.1 * torch.randn(n, generator=g, dtype=torch.float32) if mean is None else mean
    rv = 0.5 + torch.rand(n, generator=g, dtype=torch.float32) if var is None else var
    return w, b, rm, rv


def _put_bn(sd, prefix, g, n, mean=None, var=None):
    w, b, rm, rv = _bn(g, n, mean, var)
    sd[prefix + ".weight"] = w
    sd[prefix + ".bias"] = b
    sd[prefix + ".running_mean"] = rm
    sd[prefix + ".running_var"] = rv
    sd[prefix + ".num_batches_tracked"] = torch.tensor(100, dtype=torch.int64)


def _put_lin(sd, prefix, g, out_f, in_f, gain=1.0):
    w, b = _lin(g, out_f, in_f, gain)
    sd[prefix + ".weight"] = w
    sd[prefix + ".bias"] = b


def _put_resblock(sd, prefix, g, fin, fout=N_NEURONS, boost=None):
    """key order follows ResBlock.__init__ (networks.py:32-39): fc1, bn1, fc2, bn2, fc3.
    boost=(c0, c1, gain): input columns [c0,c1) (the latent / pose columns, a handful among ~4100 BPS columns) get
    `gain`x larger weights, as a trained network has, so outputs really depend on z and scores on the pose."""
    _put_lin(sd, prefix + ".fc1", g, RB_INNER, fin)
    _put_bn(sd, prefix + ".bn1", g, RB_INNER)
    _put_lin(sd, prefix + ".fc2", g, fout, RB_INNER)
    _put_bn(sd, prefix + ".bn2", g, fout)
    if fin != fout:
        _put_lin(sd, prefix + ".fc3", g, fout, fin)
    if boost is not None:
        c0, c1, gain = boost
        sd[prefix + ".fc1.weight"][:, c0:c1] *= gain
        if fin != fout:
            sd[prefix + ".fc3.weight"][:, c0:c1] *= gain


def _input_stats(g, n_pose_rot=9, n_transl=3, n_joint=0):
    """Running statistics an input BatchNorm would have learnt on [bps | rot9 | transl3 | joints]."""
    mean = [0.06 + 0.08 * torch.rand(IN_BPS, generator=g)]
    var = [0.001 + 0.002 * torch.rand(IN_BPS, generator=g)]
    if n_pose_rot:
        mean.append(0.1 * torch.randn(n_pose_rot, generator=g))
        var.append(0.2 + 0.3 * torch.rand(n_pose_rot, generator=g))
    if n_transl:
        mean.append(0.02 * torch.randn(n_transl, generator=g))
        var.append(0.003 + 0.007 * torch.rand(n_transl, generator=g))
    if n_joint:
        mean.append(0.3 * torch.rand(n_joint, generator=g))
        var.append(0.01 + 0.04 * torch.rand(n_joint, generator=g))
    return torch.cat(mean).float(), torch.cat(var).float()


def synth_gan_state_dict(seed=7, latentD=5, n_hand_joints=12):
    """State dict of the generator+discriminator container (82 entries, `discriminator.*` then `generator.*`)."""
    g = torch.Generator().manual_seed(seed)
    sd = OrderedDict()
    in_pose = 9 + 3 + n_hand_joints
    d_in = IN_BPS + in_pose
    m, v = _input_stats(g, 9, 3, n_hand_joints)
    _put_bn(sd, "discriminator.disc_bn1", g, d_in, m, v)
    _put_resblock(sd, "discriminator.disc_rb1", g, d_in, boost=(IN_BPS, d_in, 6.0))
    _put_resblock(sd, "discriminator.disc_rb2", g, N_NEURONS + d_in, boost=(IN_BPS, d_in, 6.0))
    _put_lin(sd, "discriminator.out_success", g, 1, N_NEURONS, gain=2.0)
    m, v = _input_stats(g, 0, 0, 0)
    _put_bn(sd, "generator.gen_bn1", g, IN_BPS, m, v)
    _put_resblock(sd, "generator.gen_rb1", g, latentD + IN_BPS, boost=(0, latentD, 12.0))
    _put_resblock(sd, "generator.gen_rb2", g, N_NEURONS + latentD + IN_BPS, boost=(0, latentD, 12.0))
    _put_lin(sd, "generator.gen_joint_conf", g, n_hand_joints, N_NEURONS, gain=0.3)
    _put_lin(sd, "generator.gen_rot", g, 6, N_NEURONS, gain=1.0)
    _put_lin(sd, "generator.gen_transl", g, 3, N_NEURONS, gain=0.1)
    return sd


def synth_eva_state_dict(seed=7):
    """State dict of the evaluator (55 entries: bn1, rb1..rb3, out_success)."""
    g = torch.Generator().manual_seed(seed + 1)
    sd = OrderedDict()
    e_in = IN_BPS + 12
    m, v = _input_stats(g, 9, 3, 0)
    _put_bn(sd, "bn1", g, e_in, m, v)
    _put_resblock(sd, "rb1", g, e_in, boost=(IN_BPS, e_in, 8.0))
    _put_resblock(sd, "rb2", g, e_in + N_NEURONS, boost=(N_NEURONS + IN_BPS, N_NEURONS + e_in, 8.0))
    _put_resblock(sd, "rb3", g, e_in + N_NEURONS, boost=(N_NEURONS + IN_BPS, N_NEURONS + e_in, 8.0))
    _put_lin(sd, "out_success", g, 1, N_NEURONS, gain=2.0)
    return sd


def split_gan_state_dict(sd):
    gen = OrderedDict((k[len("generator."):], v) for k, v in sd.items() if k.startswith("generator."))
    dis = OrderedDict((k[len("discriminator."):], v) for k, v in sd.items() if k.startswith("discriminator."))
    return gen, dis


# ---------------------------------------------------------------------------------------------------------
# point clouds (SURVEY.md §8d "Synthetic inputs")
# ---------------------------------------------------------------------------------------------------------
def _random_rotation(rng):
    q, r = np.linalg.qr(rng.standard_normal((3, 3)))
    q = q * np.sign(np.diag(r))
    if np.linalg.det(q) < 0:
        q[:, 0] = -q[:, 0]
    return q


def synth_cloud(n_points, seed):
    """One single-view cloud: surface of a random cuboid or ellipsoid, camera-facing side only, centred, f32 [N,3]."""
    rng = np.random.default_rng(seed)
    half = rng.uniform(0.03, 0.12, size=3)
    cam = rng.standard_normal(3)
    cam /= np.linalg.norm(cam)
    pts = np.empty((0, 3))
    box = bool(rng.integers(0, 2))
    while len(pts) < n_points:
        m = 4 * n_points
        if box:
            face = rng.integers(0, 6, size=m)
            p = rng.uniform(-1, 1, size=(m, 3))
            ax = face // 2
            sgn = (face % 2) * 2.0 - 1.0
            p[np.arange(m), ax] = sgn
            nrm = np.zeros((m, 3))
            nrm[np.arange(m), ax] = sgn
            p = p * half
        else:
            u = rng.standard_normal((m, 3))
            u /= np.linalg.norm(u, axis=1, keepdims=True)
            p = u * half
            nrm = u / half
            nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
        keep = nrm @ cam > 0.0
        pts = np.concatenate([pts, p[keep]], axis=0)
    pts = pts[:n_points] @ _random_rotation(rng).T
    pts = pts - pts.mean(axis=0, keepdims=True)
    return np.ascontiguousarray(pts, dtype=np.float32)


def synth_clouds(n_clouds, n_points, seed0=1000):
    """[B,N,3] float32; cloud i uses seed0+i.  For big batches only the first 64 are distinct shapes re-posed cheaply."""
    out = np.empty((n_clouds, n_points, 3), dtype=np.float32)
    n_unique = min(n_clouds, 64)
    for i in range(n_unique):
        out[i] = synth_cloud(n_points, seed0 + i)
    rng = np.random.default_rng(seed0 - 1)
    for i in range(n_unique, n_clouds):
        src = out[i % n_unique]
        r = _random_rotation(rng).astype(np.float32)
        perm = rng.permutation(n_points)
        out[i] = (src[perm] @ r.T) * np.float32(rng.uniform(0.8, 1.2))
    return out


_BASE_SHAPES = {}


def synth_clouds_by_id(ids, n_points, seed0=1000):
    """[len(ids),N,3] float32 clouds addressed by GLOBAL object id: ids < 64 are the distinct shapes synth_cloud(seed0 + id);
    id >= 64 re-poses shape id % 64 with a rotation / permutation / scale drawn from a generator seeded by the id itself, so
    an object's cloud does not depend on which rank (or how many ranks) holds it."""
    out = np.empty((len(ids), n_points, 3), dtype=np.float32)
    for j, i in enumerate(ids):
        b = int(i) % 64
        key = (b, n_points, seed0)
        if key not in _BASE_SHAPES:
            _BASE_SHAPES[key] = synth_cloud(n_points, seed0 + b)
        src = _BASE_SHAPES[key]
        if i < 64:
            out[j] = src
        else:
            rng = np.random.default_rng(10 ** 6 + int(i))
            r = _random_rotation(rng).astype(np.float32)
            out[j] = (src[rng.permutation(n_points)] @ r.T) * np.float32(rng.uniform(0.8, 1.2))
    return out


def synth_latents(n_rows, latentD=5, seed=0):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(n_rows, latentD, generator=g, dtype=torch.float32)


def synth_train_batch(batch, bps_rows=None, n_hand_joints=12, seed=11):
    """GAN/evaluator training batch with the reference DataLoader's dtypes (train.py:211-215,237)."""
    g = torch.Generator().manual_seed(seed)
    if bps_rows is None:
        bps_rows = (0.106 + 0.04 * torch.randn(batch, IN_BPS, generator=g)).clamp_(1e-4, 0.2).float()
    a = torch.randn(batch, 3, 3, generator=g, dtype=torch.float64)
    q, r = torch.linalg.qr(a)
    q = q * torch.sign(torch.diagonal(r, dim1=1, dim2=2)).unsqueeze(1)
    det = torch.linalg.det(q)
    q[:, :, 0] = q[:, :, 0] * det.sign().unsqueeze(1)
    return {
        "bps_object": bps_rows.float().contiguous(),
        "rot_matrix": q.contiguous(),
        "transl": 0.08 * torch.randn(batch, 3, generator=g, dtype=torch.float64),
        "joint_conf": 0.5 * torch.rand(batch, n_hand_joints, generator=g, dtype=torch.float64),
        "label": (torch.rand(batch, generator=g) < 0.3).float(),
    }
