import torch as _torch

from dex_gan_grasp_b200.bps import to_np  # noqa: F401


def to_tensor(array, dtype=_torch.float32):
    """bps_torch.utils.to_tensor"""
    if not _torch.is_tensor(array):
        array = _torch.tensor(array)
    return array.to(dtype)
