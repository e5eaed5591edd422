"""ORACLE (test infrastructure, not product code): numpy restatement of the numeric core of
DexGanGrasp/utils/filter_grasps_given_mask.py — `euclidean_distance_points_pairwise_np` (:12-30, squared distances of |deltas|),
the body of `filter_grasps_given_mask` without its file / Open3D / viewer lines (:100-104 shift and crops, :117 mean,
:123-126 distance, min, argsort) and `sort_grasps` (:140-144).  Pinned in tests/test_oracle_part_filter.py by executing the
reference's own `euclidean_distance_points_pairwise_np` and `sort_grasps` (extracted with ast) when /root/reference exists."""
import numpy as np


def pairwise_sq(pt1, pt2):
    d = np.abs(pt2[None, :, :] - pt1[:, None, :])
    return np.einsum("nmj,nmj->nm", d, d)


def filter_grasps_given_part(grasps, part_pcd_np, pc_center, crop=0.2):
    part = np.array(part_pcd_np, dtype=np.float64) - np.asarray(pc_center, dtype=np.float64)
    part = part[np.abs(part[:, 2]) < crop]
    part = part[np.abs(part[:, 1]) < crop]
    part = part[np.abs(part[:, 0]) < crop]
    mean = part.mean(axis=0).reshape(1, 3)
    dist = pairwise_sq(np.asarray(grasps["transl"]), mean)
    mins = np.min(dist, axis=1)
    return np.argsort(mins), mean, mins


def sort_grasps(grasps, idx, sort_num):
    return {k: (v[idx][:sort_num] if k in ("transl", "joint_conf", "rot_matrix") else v) for k, v in grasps.items()}
