"""Part-conditioned grasp ranking on the GPU — the numeric core of DexGanGrasp/utils/filter_grasps_given_mask.py.

`filter_grasps_given_part(grasps, part_pcd_np, pc_center)` returns what `filter_grasps_given_mask` returns
(filter_grasps_given_mask.py:84-138: `sorted_grasp_indices`, `part_pcd_mean` [1,3]) for an already masked part cloud — the
mask file lookup, Open3D objects and the viewer of the reference function are upstream / out of scope (no graphics stack).
`sort_grasps` is filter_grasps_given_mask.py:140-144.  The arithmetic runs in csrc/part_rank.cu (`dexgg_part_rank`).
"""
import numpy as np
import torch

from . import _native


def filter_grasps_given_part(grasps, part_pcd_np, pc_center, crop=0.2, device=None):
    if not torch.cuda.is_available():
        raise RuntimeError("dexgg: part ranking needs a CUDA (B200) device; there is no CPU path")
    device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    t = grasps["transl"]
    t = (t if torch.is_tensor(t) else torch.from_numpy(np.ascontiguousarray(t))).to(device, torch.float32).reshape(-1, 3).contiguous()
    p = part_pcd_np
    p = (p if torch.is_tensor(p) else torch.from_numpy(np.ascontiguousarray(p))).to(device, torch.float32).reshape(-1, 3).contiguous()
    n = t.shape[0]
    c = np.ascontiguousarray(np.asarray(pc_center, dtype=np.float32).reshape(3))
    L = _native.lib()
    ws = torch.empty(L.dexgg_part_rank_workspace_bytes(n), dtype=torch.uint8, device=device)
    idx = torch.empty(n, dtype=torch.int32, device=device)
    mean = torch.empty(3, dtype=torch.float32, device=device)
    with torch.cuda.device(device):
        _native.check(L.dexgg_part_rank(t.data_ptr(), n, p.data_ptr(), p.shape[0], c.ctypes.data, float(crop), n, idx.data_ptr(), None,
                                        mean.data_ptr(), ws.data_ptr(), ws.numel(), _native.stream_ptr(device)))
    return idx.cpu().numpy().astype(np.int64), mean.cpu().numpy().reshape(1, 3).astype(np.float64)


def sort_grasps(grasps, sorted_grasp_indices, sort_num):
    for k in ("transl", "joint_conf", "rot_matrix"):
        grasps[k] = grasps[k][sorted_grasp_indices][:sort_num]
    return grasps
