"""Mean Absolute Grasp Deviation on the GPU — drop-in for eval.py's `magd_for_grasp_distribution` (eval.py:101-137).

Same call and return convention: `magd_for_grasp_distribution(grasp1, grasp2) -> (transl_sum, rot_sum, joint_sum, coverage)`
with grasp1 the predicted set and grasp2 the ground truth (dicts holding 'transl' [.,3], 'rot_matrix' [.,3,3],
'joint_conf' [.,J] as numpy arrays or tensors).  The arithmetic runs in csrc/magd.cu through the C ABI (`dexgg_magd`);
there is no CPU path.
"""
import numpy as np
import torch

from . import _native


def _dev(x, device):
    if not torch.is_tensor(x):
        x = torch.from_numpy(np.ascontiguousarray(x))
    return x.to(device=device, dtype=torch.float32).contiguous()


def magd_device(grasp1, grasp2, device=None, return_match=False):
    """Returns a device tensor [4] f64 (transl, rot, joint, coverage) [and the matched ground-truth index per grasp]."""
    if not torch.cuda.is_available():
        raise RuntimeError("dexgg: MAGD needs a CUDA (B200) device; there is no CPU path")
    if device is None:
        t = grasp1["transl"]
        device = t.device if torch.is_tensor(t) and t.is_cuda else torch.device("cuda", torch.cuda.current_device())
    pt, gt = _dev(grasp1["transl"], device).reshape(-1, 3), _dev(grasp2["transl"], device).reshape(-1, 3)
    N, M = pt.shape[0], gt.shape[0]
    pr, gr = _dev(grasp1["rot_matrix"], device).reshape(N, 9), _dev(grasp2["rot_matrix"], device).reshape(M, 9)
    pj, gj = _dev(grasp1["joint_conf"], device).reshape(N, -1), _dev(grasp2["joint_conf"], device).reshape(M, -1)
    J = pj.shape[1]
    if gj.shape[1] != J:
        raise ValueError("joint_conf widths differ")
    L = _native.lib()
    ws = torch.empty(L.dexgg_magd_workspace_bytes(M), dtype=torch.uint8, device=device)
    sums = torch.empty(4, dtype=torch.float64, device=device)
    match = torch.empty(N, dtype=torch.int32, device=device) if return_match else None
    with torch.cuda.device(device):
        _native.check(L.dexgg_magd(pt.data_ptr(), pr.data_ptr(), pj.data_ptr(), N, gt.data_ptr(), gr.data_ptr(), gj.data_ptr(), M, J,
                                   _native.ptr(match), None, sums.data_ptr(), ws.data_ptr(), ws.numel(), _native.stream_ptr(device)))
    return (sums, match) if return_match else sums


def magd_for_grasp_distribution(grasp1, grasp2):
    s = magd_device(grasp1, grasp2).cpu().numpy()
    return float(s[0]), float(s[1]), float(s[2]), float(s[3])
