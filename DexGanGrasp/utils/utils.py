"""DexGanGrasp.utils.utils: the hot-path functions (reference DexGanGrasp/utils/utils.py:19-48, 60-64, 68-78, 114-121, 323-332,
380-394) on libdexgg.  When the reference checkout is reachable and its module imports (it needs h5py, open3d and
transforms3d at import time, utils.py:3-9), its OTHER helpers (pose / quaternion / dataset utilities the entry scripts
call, e.g. hom_matrix_from_transl_rot_matrix) are re-exported underneath, and only the hot-path names are replaced."""
import importlib.util as _ilu
import os as _os

from .. import _reference_root

_root = _reference_root()
if _root:
    try:
        _spec = _ilu.spec_from_file_location("_dexgangrasp_reference_utils", _os.path.join(_root, "DexGanGrasp", "utils", "utils.py"))
        _ref = _ilu.module_from_spec(_spec)
        _spec.loader.exec_module(_ref)
        globals().update({k: v for k, v in vars(_ref).items() if not k.startswith("__")})
    except Exception:  # a dependency of the reference module is absent: the hot-path functions below are all there is
        pass

import torch as _torch  # noqa: E402

from dex_gan_grasp_b200.utils import (class_labels_from_logits, dict_to_tensor, rot_matrix_from_ortho6d,  # noqa: E402,F401
                                      translate_along_axis)


def normalize_vector(v):
    """utils.py:323-332 (on the tensor's own device; the reference hard-codes `.cuda()` at :326)."""
    v_mag = _torch.sqrt(v.pow(2).sum(1))
    v_mag = _torch.max(v_mag, _torch.tensor([1e-8], dtype=v.dtype, device=v.device))
    return v / v_mag.view(v.shape[0], 1).expand(v.shape[0], v.shape[1])


def cross_product(u, v):
    """utils.py:68-78"""
    i = u[:, 1] * v[:, 2] - u[:, 2] * v[:, 1]
    j = u[:, 2] * v[:, 0] - u[:, 0] * v[:, 2]
    k = u[:, 0] * v[:, 1] - u[:, 1] * v[:, 0]
    return _torch.cat((i.view(-1, 1), j.view(-1, 1), k.view(-1, 1)), 1)
