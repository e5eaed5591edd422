"""Host-side mirror of the hot-path functions of DexGanGrasp/utils/utils.py (only those; the dataset / Open3D helpers of
that module are out of scope)."""
import numpy as np
import torch

from . import _native


class _Rot6dFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, r6):
        r6c = r6.contiguous()
        M = r6c.shape[0]
        rot = torch.empty((M, 3, 3), dtype=torch.float32, device=r6c.device)
        with torch.cuda.device(r6c.device):
            _native.check(_native.lib().dexgg_rot6d_forward(r6c.data_ptr(), M, rot.data_ptr(), _native.stream_ptr(r6c.device)))
        ctx.save_for_backward(r6c)
        return rot

    @staticmethod
    def backward(ctx, g):
        (r6,) = ctx.saved_tensors
        g = g.contiguous()
        out = torch.empty_like(r6)
        with torch.cuda.device(r6.device):
            _native.check(_native.lib().dexgg_rot6d_backward(r6.data_ptr(), g.data_ptr(), r6.shape[0], out.data_ptr(),
                                                             _native.stream_ptr(r6.device)))
        return out


def rot_matrix_from_ortho6d(ortho6d):
    """utils.py:380-394 (with normalize_vector :323-332 and cross_product :68-78) — [B,6] cuda f32 -> [B,3,3], columns
    (x, y, z).  Differentiable (hand-written backward kernel)."""
    if not (torch.is_tensor(ortho6d) and ortho6d.is_cuda):
        raise RuntimeError("dexgg: rot_matrix_from_ortho6d runs on a CUDA (B200) tensor only")
    return _Rot6dFn.apply(ortho6d.to(torch.float32))


def translate_along_axis(out, axis, translation):
    """utils.py:19-48: transl_i <- transl_i - R_i @ (e_axis * translation), in place on out['transl'].  The reference loops
    over n in Python with float64 temporaries; this is the same arithmetic vectorised (numpy arrays, as there) or on
    the device (tensors).  The generator path fuses it into the kernel epilogue instead (nets.cu pose_finish_kernel)."""
    t, r = out['transl'], out['rot_matrix']
    if torch.is_tensor(t):
        out['transl'] = (t.double() - r.double()[:, :, axis] * float(translation)).to(t.dtype)
        return out
    e = np.zeros(3)
    e[axis] = translation
    out['transl'][...] = (np.asarray(t, np.float64) - np.einsum("nij,j->ni", np.asarray(r, np.float64), e)).astype(t.dtype)
    return out


def dict_to_tensor(in_dict, device, dtype):
    """utils.py:114-121"""
    out_dict = {}
    for k, v in in_dict.items():
        if torch.is_tensor(v):
            out_dict[k] = v.clone().detach()
        else:
            out_dict[k] = torch.tensor(v, dtype=dtype, device=device)
    return out_dict


def class_labels_from_logits(logits, threshold):
    """utils.py:60-64 (thresholds the sigmoid output IN PLACE, like the reference)"""
    logits[logits >= threshold] = 1.
    logits[logits < threshold] = 0.
    return logits
