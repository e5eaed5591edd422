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
    "dexgg_bps_encode_static_max_keys": (c_size_t, []),
    "dexgg_bps_encode_static": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_int, c_int64, c_int64, c_void_p, c_void_p, c_int, c_void_p]),
    "dexgg_center_clouds": (c_int, [c_void_p, c_void_p, c_int, c_int64, c_int64, c_void_p, c_void_p, c_void_p]),
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
    "dexgg_bn_fwd_slab": (c_int, [c_void_p, c_void_p]),
    "dexgg_bn_bwd_slab": (c_int, [c_void_p, c_void_p]),
    "dexgg_bn_input_stats": (c_int, [c_void_p, c_int64, c_int64, c_int, c_float, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "dexgg_xhat": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_int64, c_int, c_void_p, c_int, c_void_p]),
    "dexgg_tail_pack": (c_int, [c_void_p, c_int64, c_int, c_float, c_float, c_float, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    "dexgg_tail_bn_bwd": (c_int, [c_void_p, c_int64, c_float, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_void_p, c_int64, c_int,
                                  c_void_p, c_int, c_void_p, c_void_p]),
    "dexgg_score_head": (c_int, [c_void_p, c_void_p]),
    "dexgg_gen_pose": (c_int, [c_void_p, c_int64, c_int, c_int64, c_void_p, c_void_p, c_void_p, c_void_p]),
    "dexgg_gen_loss": (c_int, [c_void_p, c_int64, c_int, c_int64, c_float, c_void_p, c_void_p, c_float, c_float, c_float, c_void_p,
                               c_int64, c_void_p, c_void_p, c_void_p]),
    "dexgg_prep_weights": (c_int, [c_void_p, c_void_p]),
    "dexgg_unfold_grads": (c_int, [c_void_p, c_void_p]),
    "dexgg_bn_running_multi": (c_int, [c_void_p, c_int, c_void_p]),
    "dexgg_bn_merge": (c_int, [c_void_p, c_int, c_int, c_int, c_float, c_float, c_void_p, c_void_p, c_void_p, c_void_p]),
    "dexgg_bump_counter": (c_int, [c_void_p, c_void_p]),
    "dexgg_slab_scratch_bytes": (c_size_t, []),
    "dexgg_peer_alloc": (c_int, [c_size_t, ctypes.POINTER(c_void_p), c_void_p]),
    "dexgg_peer_open": (c_int, [c_void_p, ctypes.POINTER(c_void_p)]),
    "dexgg_peer_close": (c_int, [c_void_p]),
    "dexgg_peer_free": (c_int, [c_void_p]),
    "dexgg_slab_step_begin": (c_int, [c_void_p, c_void_p, c_void_p]),
    "dexgg_p2p_small": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p]),
    "dexgg_grad_allreduce": (c_int, [c_void_p, c_void_p, ctypes.c_longlong, ctypes.c_float, c_int, c_void_p]),
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
                ("C16", c_void_p), ("C32", c_void_p), ("ldc", c_int64), ("accumulate", c_int), ("split_k", c_int), ("tile_n", c_int), ("split_slice_stride", c_int64)]


c_uint, c_ull = ctypes.c_uint, ctypes.c_ulonglong


class SlabComm(ctypes.Structure):
    """include/dexgg.h: dexgg_slab_comm"""
    _fields_ = [("peers", c_void_p), ("world", c_int), ("rank", c_int), ("slot", c_int), ("parity", c_void_p)]


class BnFwd(ctypes.Structure):
    """include/dexgg.h: dexgg_bn_fwd"""
    _fields_ = [("x", c_void_p), ("ldx", c_int64), ("C", c_int), ("rows_per_group", c_int64), ("n_groups", c_int), ("gamma", c_void_p),
                ("beta", c_void_p), ("shift", c_void_p), ("mean", c_void_p), ("invstd", c_void_p), ("eps", c_float), ("count_total", c_float),
                ("residual", c_void_p), ("ldr", c_int64), ("act", c_int), ("drop_mask", c_void_p), ("ldm", c_int64), ("drop_p", c_float),
                ("drop_seed", c_void_p), ("drop_stream", c_uint), ("y", c_void_p), ("ldy", c_int64), ("mode", c_int), ("acc", c_void_p), ("comm", SlabComm), ("ctas_per_sm", c_int)]


class BnBwd(ctypes.Structure):
    """include/dexgg.h: dexgg_bn_bwd"""
    _fields_ = [("g", c_void_p), ("ldg", c_int64), ("dl", c_void_p), ("w", c_void_p), ("dw_out", c_void_p), ("out", c_void_p),
                ("ldo", c_int64), ("act", c_int), ("drop_mask", c_void_p), ("ldm", c_int64), ("drop_p", c_float), ("drop_seed", c_void_p),
                ("drop_stream", c_uint), ("x", c_void_p), ("ldx", c_int64), ("mean", c_void_p), ("invstd", c_void_p), ("gamma", c_void_p),
                ("C", c_int), ("rows_per_group", c_int64), ("n_groups", c_int), ("count_total", c_float), ("dgamma", c_void_p),
                ("dbeta", c_void_p), ("gx", c_void_p), ("ldgx", c_int64), ("skip", c_void_p), ("lds", c_int64), ("gskip", c_void_p),
                ("ldgs", c_int64), ("dbias_skip", c_void_p), ("mode", c_int), ("acc", c_void_p), ("comm", SlabComm), ("ctas_per_sm", c_int)]


class TailSrc(ctypes.Structure):
    _fields_ = [("p", c_void_p * 3), ("n", c_int * 3), ("f64", c_int)]


class ScoreHead(ctypes.Structure):
    _fields_ = [("X", c_void_p), ("ldx", c_int64), ("w", c_void_p), ("b", c_void_p), ("rows_per_group", c_int64), ("n_groups", c_int),
                ("labels", c_void_p), ("target", c_float * 2), ("dscale", c_float * 2), ("lscale", c_float * 2), ("p", c_void_p),
                ("dl", c_void_p), ("loss", c_void_p), ("db", c_void_p), ("wass", c_int)]


class FirstLayer(ctypes.Structure):
    _fields_ = [("W1", c_void_p), ("W3", c_void_p), ("b1", c_void_p), ("b3", c_void_p), ("Fin", c_int), ("cmap", c_void_p), ("dmap", c_void_p),
                ("bps_src0", c_int), ("in_gamma", c_void_p), ("in_beta", c_void_p), ("Wb", c_void_p), ("Wxt", c_void_p), ("ldxt", c_int), ("bias13", c_void_p),
                ("dWb", c_void_p), ("dWxt", c_void_p), ("db13", c_void_p), ("gW1", c_void_p), ("gW3", c_void_p), ("gb1", c_void_p),
                ("gb3", c_void_p), ("g_in_gamma", c_void_p), ("g_in_beta", c_void_p)]


class PlainJob(ctypes.Structure):
    _fields_ = [("src", c_void_p), ("lds", c_int64), ("rows", c_int), ("cols", c_int), ("dst", c_void_p), ("ldd", c_int64),
                ("gsrc", c_void_p), ("ldgs", c_int64), ("gdst", c_void_p), ("ldgd", c_int64)]


class WeightJobs(ctypes.Structure):
    _fields_ = [("n_first", c_int), ("n_plain", c_int), ("first", FirstLayer * 3), ("plain", PlainJob * 8)]


class RunEntry(ctypes.Structure):
    _fields_ = [("running_mean", c_void_p), ("running_var", c_void_p), ("mean", c_void_p), ("second", c_void_p),
                ("num_batches_tracked", c_void_p), ("C", c_int), ("n_groups", c_int), ("second_is_var", c_int), ("same_row", c_int),
                ("count", c_float), ("eps", c_float), ("momentum", c_float)]


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
