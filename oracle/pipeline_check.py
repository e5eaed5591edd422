"""TEST INFRASTRUCTURE — the fp32 oracle of the WHOLE batched pipeline and the set comparison used to pin the benchmarked
bf16 path.  Only tests/, __graft_entry__.smoke() and bench.py's `parity_check` (outside every timed region) call this.

oracle pipeline = what the reference does per object (dexgangrasp_real_robot.py:139-150, dexgangrasp_offline.py:147-162):
    bps.encode (oracle/bps_oracle.c)  ->  np.tile  ->  Generator.forward + rot6d + palm offset (networks.py:223-255)
    ->  DexEvaluator.forward (networks.py:315-337)  ->  sort / `p > thresh` (dexgangrasp.py:412-419)  ->  first k rows
"""
import ctypes
import os
import subprocess

import numpy as np
import torch

from . import nets_oracle as no

_HERE = os.path.dirname(os.path.abspath(__file__))


def _lib():
    so = os.path.join(_HERE, "libdexgg_oracle.so")
    if not os.path.exists(so):
        subprocess.check_call(["make", "-C", _HERE])
    return ctypes.CDLL(so)


def oracle_pipeline(sd_gan, sd_eva, basis, clouds, z, samples_per_object, thresh, k, z_offset=0.025, nthreads=None):
    """clouds [B,N,3] f32 numpy, z [B*S,zD] torch/numpy.  Returns dict: bps [B,P], bps_idx [B,P], rot [M,3,3], transl [M,3],
    joint [M,nJ], p [M], topk_idx [B,k] (global rows, -1 padded), topk_count [B]."""
    lib = _lib()
    vp = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    basis = np.ascontiguousarray(basis, np.float32)
    clouds = np.ascontiguousarray(clouds, np.float32)
    B, S, P = clouds.shape[0], int(samples_per_object), basis.shape[0]
    nthreads = nthreads or min(os.cpu_count() or 1, 64)
    enc = np.empty((B, P), np.float32)
    idx = np.empty((B, P), np.int32)
    for b in range(B):
        lib.dexgg_oracle_bps_encode_mt(vp(basis), P, vp(clouds[b]), clouds.shape[1], vp(enc[b]), vp(idx[b]), nthreads)
    z = torch.as_tensor(np.asarray(z), dtype=torch.float32)
    rot, transl, joint, p = [], [], [], []
    with torch.no_grad():
        for b in range(B):  # per object, like the reference (and keeps the tiled [S,4096] matrix small)
            tiled = torch.from_numpy(np.tile(enc[b], (S, 1)))
            g = no.generate_poses(sd_gan, z[b * S:(b + 1) * S], tiled, z_offset)
            pe = no.evaluator_forward(sd_eva, tiled, torch.from_numpy(g["rot_matrix"]), torch.from_numpy(g["transl"]))
            rot.append(g["rot_matrix"]); transl.append(g["transl"]); joint.append(g["joint_conf"])
            p.append(pe.squeeze(1).numpy())
    p = np.concatenate(p)
    tidx, tcnt = no.segmented_topk(p, np.arange(B + 1) * S, thresh, k)
    return {"bps": enc, "bps_idx": idx, "rot": np.concatenate(rot), "transl": np.concatenate(transl),
            "joint": np.concatenate(joint), "p": p, "topk_idx": tidx, "topk_count": tcnt}


def compare_topk(p_oracle, got_idx, got_count, samples_per_object, thresh, k, eps):
    """The filtered-set contract of tolerances.BF16_PIPELINE_EPS.  Per object: cut = max(thresh, k-th best oracle score)
    (thresh when fewer than k oracle rows pass).  A row with oracle score > cut + eps MUST be in the returned set; a row with
    oracle score < cut - eps must NOT be (nor may any row with score <= thresh - eps).  Returns the number of violations and
    how many rows were decided inside the eps band."""
    S = int(samples_per_object)
    p_oracle = np.asarray(p_oracle, np.float32).reshape(-1, S)
    B = p_oracle.shape[0]
    got_idx = np.asarray(got_idx).reshape(B, -1)
    got_count = np.asarray(got_count).reshape(B)
    missing = extra = in_band = 0
    for b in range(B):
        po = p_oracle[b]
        above = np.sort(po[po > thresh])[::-1]
        cut = max(thresh, float(above[k - 1])) if len(above) >= k else thresh
        must = set(np.nonzero(po > cut + eps)[0].tolist())
        never = set(np.nonzero(po < cut - eps)[0].tolist())
        got = set((got_idx[b, :got_count[b]] - b * S).tolist())
        assert all(0 <= g < S for g in got), "top-k row ids must lie inside their object's segment"
        missing += len(must - got)
        extra += len(got & never)
        in_band += len(got - must)
    return {"set_mismatches": int(missing + extra), "missing": int(missing), "extra": int(extra), "in_eps_band": int(in_band)}


def check_pipeline(out, ref, samples_per_object, thresh, k, eps):
    """`out`: GraspPipeline.run_device result (device tensors); `ref`: oracle_pipeline result.  Returns the `parity_check`
    record of bench.py: BPS bit-exactness, max |dp| over every grasp, max pose drift, set mismatches under eps."""
    enc = out["bps"].detach().cpu().numpy()
    p = out["p"].detach().cpu().numpy().reshape(-1)
    rec = {"objects": int(enc.shape[0]), "grasps": int(p.shape[0]), "eps": float(eps),
           "bps_bit_exact": bool(np.array_equal(enc, ref["bps"])),
           "max_dp": float(np.max(np.abs(p.astype(np.float64) - ref["p"].astype(np.float64))))}
    rec.update(compare_topk(ref["p"], out["idx"].detach().cpu().numpy(), out["count"].detach().cpu().numpy(), samples_per_object,
                            thresh, k, eps))
    rec["oracle_kept"] = int(ref["topk_count"].sum())
    rec["kept"] = int(out["count"].sum().item())
    rec["identical_topk_rows"] = int(sum(len(set(a[:ca].tolist()) & set(b[:cb].tolist())) for a, ca, b, cb in
                                         zip(out["idx"].cpu().numpy(), out["count"].cpu().numpy(), ref["topk_idx"], ref["topk_count"])))
    return rec
