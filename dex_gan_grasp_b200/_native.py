"""ctypes binding of libdexgg.so (include/dexgg.h).  This is the binding INTEGRATION.md describes.

The product path has no fallback: if the library is missing or the device is not a B200 every call raises.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libdexgg.so")

_lib = None

c_void_p, c_int, c_int64, c_size_t, c_float = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_size_t, ctypes.c_float

# name -> (restype, argtypes); every symbol include/dexgg.h declares
SIGNATURES = {
    "dexgg_version": (c_int, []),
    "dexgg_last_error": (ctypes.c_char_p, []),
    "dexgg_device_check": (c_int, []),
    "dexgg_launch_count": (ctypes.c_ulonglong, []),
    "dexgg_bps_workspace_bytes": (c_size_t, [c_int, c_int]),
    "dexgg_bps_encode": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_int, c_int64, c_int64, c_void_p, c_void_p, c_int,
                                 c_void_p, c_size_t, c_void_p]),
    "dexgg_bps_encode_host": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_int]),
    "dexgg_net_create": (c_int, [c_int, c_void_p, c_int, c_int, c_int, c_int, c_void_p, ctypes.POINTER(c_void_p)]),
    "dexgg_net_update": (c_int, [c_void_p, c_void_p, c_int, c_void_p]),
    "dexgg_net_destroy": (None, [c_void_p]),
    "dexgg_net_workspace_bytes": (c_size_t, [c_void_p, c_int64, c_int64]),
    "dexgg_generator_forward": (c_int, [c_void_p, c_void_p, c_int64, c_int64, c_void_p, c_int64, c_float, c_void_p, c_void_p,
                                        c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "dexgg_evaluator_forward": (c_int, [c_void_p, c_void_p, c_int64, c_int64, c_void_p, c_void_p, c_int64, c_void_p, c_void_p,
                                        c_size_t, c_void_p]),
    "dexgg_discriminator_forward": (c_int, [c_void_p, c_void_p, c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_int64,
                                            c_void_p, c_void_p, c_size_t, c_void_p]),
    "dexgg_filter_topk": (c_int, [c_void_p, c_int64, c_int64, c_float, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "dexgg_filter_topk_workspace_bytes": (c_size_t, [c_int64, c_int64]),
    "dexgg_filter_topk_ws": (c_int, [c_void_p, c_int64, c_int64, c_float, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                     c_size_t, c_void_p]),
    "dexgg_filter_topk_gather": (c_int, [c_void_p, c_int64, c_int64, c_float, c_int, c_void_p, c_void_p, c_void_p, c_int, c_void_p,
                                         c_void_p, c_void_p, c_void_p]),
    "dexgg_linear_f32": (c_int, [c_void_p, c_int64, c_int, c_void_p, c_int64, c_int, c_int, c_int, c_int, c_void_p, c_void_p,
                                 c_void_p, c_void_p, c_int64, c_int, c_int, c_void_p, c_int64, c_void_p]),
    "dexgg_linear_bf16": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                  c_void_p, c_int64, c_int, c_void_p, c_void_p, c_int64, c_void_p]),
    "dexgg_gemm_bf16": (c_int, [c_void_p, c_void_p]),
    "dexgg_f32_to_bf16": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_int64, c_int, c_void_p]),
    "dexgg_f32_to_bf16_t": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_int64, c_int, c_void_p]),
    "dexgg_rot6d_forward": (c_int, [c_void_p, c_int64, c_void_p, c_void_p]),
    "dexgg_rot6d_backward": (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_void_p]),
    "dexgg_col_sum": (c_int, [c_void_p, c_int64, c_int64, c_int, c_void_p, c_void_p]),
    "dexgg_col_centered_sumsq": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_int, c_void_p, c_void_p]),
    "dexgg_bn_finalize": (c_int, [c_void_p, c_int, c_float, c_int, c_float, c_float, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "dexgg_bn_apply": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int64, c_int,
                               c_void_p, c_int64, c_void_p]),
    "dexgg_lrelu_backward": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_int64, c_int, c_void_p, c_int64, c_void_p]),
    "dexgg_bn_backward_reduce": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_void_p, c_int64, c_int, c_void_p, c_void_p,
                                         c_void_p]),
    "dexgg_bn_backward_apply": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                        c_float, c_int64, c_int, c_void_p, c_int64, c_void_p]),
    "dexgg_adam_step": (c_int, [c_void_p, c_int, c_float, c_float, c_float, c_float, c_float, c_int, c_void_p]),
    "dexgg_magd_workspace_bytes": (c_size_t, [c_int]),
    "dexgg_magd": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p,
                           c_void_p, c_void_p, c_size_t, c_void_p]),
    "dexgg_part_rank_workspace_bytes": (c_size_t, [c_int64]),
    "dexgg_part_rank": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_float, c_int, c_void_p, c_void_p, c_void_p, c_void_p,
                                c_size_t, c_void_p]),
    "dexgg_adam_step_dev": (c_int, [c_void_p, c_int, c_float, c_float, c_float, c_float, c_float, c_void_p, c_void_p]),
    "dexgg_fp32_pipe_probe": (c_int, [c_int, c_int, ctypes.POINTER(c_float), ctypes.POINTER(ctypes.c_double)]),
}


class GemmSeg(ctypes.Structure):
    _fields_ = [("A", c_void_p), ("lda", c_int64), ("a_mn_major", c_int), ("B", c_void_p), ("ldb", c_int64), ("b_mn_major", c_int),
                ("K", c_int)]


class Gemm(ctypes.Structure):
    """include/dexgg.h: dexgg_gemm"""
    _fields_ = [("M", c_int), ("N", c_int), ("n_seg", c_int), ("seg", GemmSeg * 3), ("bias", c_void_p), ("act_from", c_int),
                ("C16", c_void_p), ("C32", c_void_p), ("ldc", c_int64), ("accumulate", c_int), ("split_k", c_int), ("tile_n", c_int)]


class DexggError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"libdexgg error {code}: {msg}")
        self.code = code


def lib():
    """Loads libdexgg.so (never builds it, never falls back)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing: run `python -m dex_gan_grasp_b200.build` (there is no CPU fallback)")
        l = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(l, name)
            fn.restype = res
            fn.argtypes = args
        _lib = l
    return _lib


def check(rc):
    if rc != 0:
        raise DexggError(rc, lib().dexgg_last_error().decode("utf-8", "replace"))


def ptr(t):
    """Device (or host) address of a torch tensor / numpy array / None."""
    if t is None:
        return None
    if hasattr(t, "data_ptr"):
        return t.data_ptr()
    return t.ctypes.data


def stream_ptr(device=None):
    import torch
    return torch.cuda.current_stream(device).cuda_stream
