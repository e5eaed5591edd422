"""ORACLE (test infrastructure, not product code): numpy restatement of the reference's Mean Absolute Grasp Deviation.

Follows /root/reference/eval.py:
  geodesic_distance_rotmats_pairwise_np      eval.py:48-61   arccos(clip((trace(R1 R2^T) - 1) / 2, -1, 1))
  euclidean_distance_points_pairwise_np      eval.py:64-80   SQUARED distances (the reference never takes the root)
  euclidean_distance_joint_conf_pairwise_np  eval.py:83-99   SQUARED distances
  magd_for_grasp_distribution                eval.py:101-137 per predicted grasp: nearest ground-truth grasp by squared
      translation distance (np.argmin: first index on ties); sums of that translation distance, of the rotation angle and the
      squared joint distance to the SAME ground-truth grasp; coverage = distinct matched ground-truth grasps / M.
Pinned by tests/golden/magd_golden.npz, produced by executing the reference's own function bodies
(tests/golden/make_magd_golden.py).  Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this module.
"""
import numpy as np


def pairwise_sq(a, b):
    a = np.asarray(a)
    b = np.asarray(b)
    d = b[None, :, :] - a[:, None, :]
    return np.einsum("nmj,nmj->nm", d, d)


def geodesic(r1, r2):
    tr = np.einsum("aij,bij->ab", np.asarray(r1), np.asarray(r2))  # trace(R1 R2^T)
    return np.arccos(np.clip((tr - 1.0) / 2.0, -1.0, 1.0))


def magd(grasp1, grasp2, return_match=False):
    """grasp1: predicted set, grasp2: ground truth; dicts with 'transl' [.,3], 'rot_matrix' [.,3,3], 'joint_conf' [.,J]."""
    t = pairwise_sq(grasp1["transl"], grasp2["transl"])
    r = geodesic(grasp1["rot_matrix"], grasp2["rot_matrix"])
    j = pairwise_sq(grasp1["joint_conf"], grasp2["joint_conf"])
    idx = np.argmin(t, axis=1)
    rows = np.arange(t.shape[0])
    out = (float(np.sum(t[rows, idx])), float(np.sum(r[rows, idx])), float(np.sum(j[rows, idx])),
           len(set(idx.tolist())) / len(grasp2["transl"]))
    return out + (idx,) if return_match else out
