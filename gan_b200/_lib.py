"""ctypes binding of libganb200.so (the C-ABI declared in include/ganb200.h).

The product has no CPU path: if the shared library is missing, or the device is not sm_100, every
kernel call raises.  Build with ``make -C gan_b200/csrc`` or ``python -c 'import __graft_entry__ as g; g.build()'``.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libganb200.so")

c_void_p, c_int, c_int64, c_float, c_size_t = (ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_float,
                                               ctypes.c_size_t)


class ConvDesc(ctypes.Structure):
    """Mirror of ganb200_conv_desc."""
    _fields_ = [("N", c_int), ("C", c_int), ("H", c_int), ("W", c_int), ("K", c_int), ("R", c_int), ("S", c_int),
                ("stride", c_int), ("pad", c_int), ("pad_mode", c_int), ("P", c_int), ("Q", c_int),
                ("alpha", c_float), ("engine", c_int)]


class AdamEntry(ctypes.Structure):
    """Mirror of ganb200_adam_entry (5 x 8 bytes)."""
    _fields_ = [("p", c_void_p), ("g", c_void_p), ("m", c_void_p), ("v", c_void_p), ("n", c_int64)]


class GlinJobC(ctypes.Structure):
    """Mirror of ganb200_glin_job (96 bytes)."""
    _fields_ = [("a", c_void_p), ("b", c_void_p), ("out", c_void_p), ("bias", c_void_p),
                ("a_off", c_int64), ("b_off", c_int64), ("out_off", c_int64),
                ("ni", c_int), ("nj", c_int), ("nr", c_int), ("lda", c_int), ("ldb", c_int), ("ldo", c_int),
                ("a_red_is_row", c_int), ("b_red_is_row", c_int), ("alpha", c_float), ("first_block", c_int)]


class GbnJobC(ctypes.Structure):
    """Mirror of ganb200_gbn_job (72 bytes)."""
    _fields_ = [("gamma", c_void_p), ("beta", c_void_p), ("slope", c_void_p), ("running_mean", c_void_p),
                ("running_var", c_void_p), ("nbt", c_void_p), ("x_off", c_int64), ("stat_off", c_int64),
                ("C", c_int), ("first_block", c_int)]


P = c_void_p
_DESC_P = ctypes.POINTER(ConvDesc)

# name -> (restype, argtypes).  Every function declared in include/ganb200.h must appear here
# (tests/test_abi.py checks header <-> table <-> exported symbols).
SIGNATURES = {
    "ganb200_abi_version": (c_int, []),
    "ganb200_last_error": (ctypes.c_char_p, []),
    "ganb200_device_check": (c_int, [c_int]),
    "ganb200_conv2d_workspace_bytes": (c_size_t, [_DESC_P, c_int]),
    "ganb200_conv2d_tc_eligible": (c_int, [_DESC_P, c_int]),
    "ganb200_conv_halo_mode": (c_int, [c_int]),
    "ganb200_conv_scatter_mode": (c_int, [c_int]),
    "ganb200_conv_flat_mode": (c_int, [c_int]),
    "ganb200_conv_flat_trace": (c_int, [P]),
    "ganb200_tc_shift_probe": (c_int, [P, P, P, c_int, c_int, c_int, c_int, P]),
    "ganb200_tc_shift_probe_m": (c_int, [P, P, P, c_int, c_int, c_int, c_int, c_int, P]),
    "ganb200_conv2d_route": (c_int, [_DESC_P, c_int]),
    "ganb200_conv2d_fprop": (c_int, [_DESC_P, P, P, P, P, P, P, P, c_size_t, P]),
    "ganb200_conv2d_dgrad": (c_int, [_DESC_P, P, P, P, P, P, P, c_size_t, P]),
    "ganb200_conv2d_wgrad": (c_int, [_DESC_P, P, P, P, P, P, P, c_size_t, P]),
    "ganb200_tc_packed_act_bytes": (c_size_t, [c_int, c_int, c_int, c_int, c_int]),
    "ganb200_tc_pack_act": (c_int, [P, P, P, c_int, c_int, c_int, c_int, c_int, P]),
    "ganb200_conv2d_packed_weight_bytes": (c_size_t, [_DESC_P, c_int]),
    "ganb200_conv2d_pack_weight": (c_int, [_DESC_P, c_int, P, P, P]),
    "ganb200_conv2d_fprop_pk": (c_int, [_DESC_P, P, P, P, P, P, P]),
    "ganb200_conv2d_dgrad_pk": (c_int, [_DESC_P, P, P, P, P, P, P, P, P, c_size_t, P]),
    "ganb200_conv2d_wgrad_pk": (c_int, [_DESC_P, P, P, P, P, P, P, P, P, c_size_t, P]),
    "ganb200_modconv_bwd_pk": (c_int, [_DESC_P, P, P, P, P, P, P, P, P, P, P, c_size_t, P]),
    "ganb200_resample2d": (c_int, [P, P, c_int, c_int, c_int, c_int, c_int, P, P, c_int, P, P, c_int, P]),
    "ganb200_ew_unary": (c_int, [c_int, P, P, c_int64, c_float, c_float, P]),
    "ganb200_ew_binary": (c_int, [c_int, P, P, P, c_int64, c_float, P]),
    "ganb200_ew_bcast": (c_int, [c_int, P, P, P, c_int, c_int, c_int, c_int, P]),
    "ganb200_ew_fma_bcast": (c_int, [P, P, P, P, c_int, c_int, c_int, c_int, P]),
    "ganb200_bmm": (c_int, [P, P, P, c_int, c_int, c_int, c_int, c_int64, c_int64, c_int64, c_int64, c_int64, c_int64,
                            c_float, P]),
    "ganb200_softmax_rows_fwd": (c_int, [P, P, c_int64, c_int, P]),
    "ganb200_softmax_rows_bwd": (c_int, [P, P, P, c_int64, c_int, P]),
    "ganb200_pixnorm_fwd": (c_int, [P, P, P, c_int, c_int, c_int64, c_float, P]),
    "ganb200_pixnorm_bwd": (c_int, [P, P, P, P, c_int, c_int, c_int64, P]),
    "ganb200_pool_sum2_eligible": (c_int, [c_int, c_int, c_int]),
    "ganb200_pool_sum2": (c_int, [P, P, P, c_int64, c_int, c_int, c_int, P]),
    "ganb200_sk_bwd_finish": (c_int, [P, P, P, P, P, c_int, c_int, c_int, c_int, c_int, P]),
    "ganb200_sk_merge2_fwd": (c_int, [P, P, P, P, c_int, c_int, c_int, P]),
    "ganb200_sk_merge2_bwd": (c_int, [P, P, P, P, P, P, P, c_int, c_int, c_int, P]),
    "ganb200_reduce": (c_int, [P, P, P, c_int, c_int, c_int, c_int, c_float, P]),
    "ganb200_prelu_fwd": (c_int, [P, P, P, c_int, c_int, c_int, c_int, P]),
    "ganb200_maskscale": (c_int, [P, P, P, P, c_int, c_int, c_int, c_int, P]),
    "ganb200_prelu_bwd": (c_int, [P, P, P, P, P, c_int, c_int, c_int, c_int, P]),
    "ganb200_masked_dot": (c_int, [P, P, P, P, c_int, c_int, c_int, c_int, P]),
    "ganb200_bn_stats": (c_int, [P, P, P, c_int, c_int, c_int, P]),
    "ganb200_bn_apply": (c_int, [P, P, P, P, P, P, P, P, P, P, c_int, c_int, c_int, c_float, c_float, c_int64, P]),
    "ganb200_bn_bwd_reduce": (c_int, [P, P, P, P, P, P, c_int, c_int, c_int, P]),
    "ganb200_bn_bwd_apply": (c_int, [P, P, P, P, P, P, P, P, P, P, c_int, c_int, c_int, c_int64, P]),
    "ganb200_bn_small_eligible": (c_int, [c_int, c_int, c_int]),
    "ganb200_bn_small_fwd": (c_int, [P, P, P, P, P, P, P, P, P, c_int, c_int, c_int, c_float, c_float, P, P]),
    "ganb200_bn_small_bwd": (c_int, [P, P, P, P, P, P, P, P, c_int, c_int, c_int, P, P, P, P]),
    "ganb200_softmax_m_fwd": (c_int, [P, P, c_int, c_int, c_int, P]),
    "ganb200_softmax_m_bwd": (c_int, [P, P, P, c_int, c_int, c_int, P]),
    "ganb200_mbstd_fwd": (c_int, [P, P, c_int, c_int64, c_float, P]),
    "ganb200_mbstd_bwd": (c_int, [P, P, P, c_int, c_int64, c_float, P]),
    "ganb200_mbstd_bwd_bwd": (c_int, [P, P, P, P, P, c_int, c_int64, c_float, P]),
    "ganb200_linear_group": (c_int, [P, c_int, c_int, P, P, P, P]),
    "ganb200_bn1d_group_stats": (c_int, [P, c_int, c_int, P, c_int, P, P, P]),
    "ganb200_bn1d_group_apply": (c_int, [P, c_int, c_int, P, P, c_int, P, P, P, c_float, c_float, c_int64, P]),
    "ganb200_bn1d_group_bwd_reduce": (c_int, [P, c_int, c_int, P, P, c_int, P, P, P, P, P, P]),
    "ganb200_bn1d_group_bwd_apply": (c_int, [P, c_int, c_int, P, P, P, c_int, P, P, P, P, c_int64, P, P]),
    "ganb200_gather_flip": (c_int, [P, P, P, P, c_int, c_int, c_int, c_int, P]),
    "ganb200_minmax": (c_int, [P, c_int64, P, P]),
    "ganb200_make_grid_u8": (c_int, [P, P, P, c_int, c_int, c_int, c_int, c_int, c_int, P]),
    "ganb200_bce_fwd": (c_int, [P, P, P, P, c_int64, P]),
    "ganb200_bce_fwd_sticky": (c_int, [P, P, P, P, c_int64, P]),
    "ganb200_bce_bwd": (c_int, [P, P, P, P, c_int64, P]),
    "ganb200_multi_tensor_adam_blocks": (c_int, [P, P, c_int, c_int64, c_float, c_float, c_float, c_float, c_float,
                                                 c_int, c_int64, P, c_float, P]),
    "ganb200_multi_tensor_copy": (c_int, [P, P, c_int, c_int64, P]),
    "ganb200_counter_add": (c_int, [P, c_int64, P]),
    "ganb200_peer_bytes": (c_int64, [c_int64]),
    "ganb200_peer_alloc": (c_int, [c_int64, P]),
    "ganb200_peer_free": (c_int, [P]),
    "ganb200_peer_export": (c_int, [P, P]),
    "ganb200_peer_open": (c_int, [P, P]),
    "ganb200_peer_close": (c_int, [P]),
    "ganb200_peer_allreduce": (c_int, [P, c_int, c_int, c_int64, P, P, P, P, c_int64, c_int, P]),
    "ganb200_peer_error": (c_int, [P, P]),
    "ganb200_bn_small_sync_max_channels": (c_int, []),
    "ganb200_bn_small_fwd_sync": (c_int, [P, P, P, P, P, P, P, P, P, c_int, c_int, c_int, c_float, c_float, P, P, c_int,
                                          c_int, P]),
    "ganb200_bn_small_bwd_sync": (c_int, [P, P, P, P, P, P, P, P, c_int, c_int, c_int, P, P, P, P, c_int, c_int, P]),
    "ganb200_chan_copy": (c_int, [P, P, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, P]),
    "ganb200_mixup": (c_int, [P, P, c_float, P, c_int64, P]),
    "ganb200_mixup_wgan_f64": (c_int, [P, P, P, P, c_int, c_int, c_int, c_int, P]),
    "ganb200_clamp_": (c_int, [P, c_int64, c_float, c_float, P]),
    "ganb200_noise_fill": (c_int, [P, c_int64, P, c_int, P]),
    "ganb200_demod": (c_int, [P, P, P, c_int, c_int, c_int, c_float, c_float, P]),
    "ganb200_demod_bwd_s": (c_int, [P, P, P, P, P, P, c_int, c_int, c_int, c_float, P]),
    "ganb200_demod_bwd_w": (c_int, [P, P, P, P, P, c_int, c_int, c_int, c_int, c_float, P]),
    "ganb200_style_tail_fwd": (c_int, [P, P, P, P, P, c_int, c_int, c_int, c_int, P, c_int, P]),
    "ganb200_style_tail_bwd": (c_int, [P, P, P, P, P, c_int, P, P, P, c_int, c_int, c_int, P, c_int, P]),
    "ganb200_modconv_bwd_workspace_bytes": (c_size_t, [_DESC_P]),
    "ganb200_modconv_bwd_eligible": (c_int, [_DESC_P]),
    "ganb200_modconv_bwd": (c_int, [_DESC_P, P, P, P, P, P, P, P, P, P, c_size_t, P]),
    "ganb200_scale_reduce": (c_int, [P, P, P, P, P, c_int, c_int, c_int, P]),
    "ganb200_spectral_power_iter": (c_int, [P, P, P, P, c_int, c_int, c_int, c_float, P]),
    "ganb200_spectral_power_iter_batched": (c_int, [P, c_int, c_int, c_int, c_float, P]),
    "ganb200_spectral_scale": (c_int, [P, P, P, c_int64, P]),
    "ganb200_spectral_scale_bwd": (c_int, [P, P, P, P, P, P, P, c_int, c_int, P]),
    "ganb200_dwconv2d_fprop": (c_int, [P, P, P, P] + [c_int] * 8 + [c_float, P]),
    "ganb200_dwconv2d_dgrad": (c_int, [P, P, P] + [c_int] * 8 + [c_float, P]),
    "ganb200_dwconv2d_wgrad_workspace_bytes": (c_int64, [c_int, c_int]),
    "ganb200_dwconv2d_wgrad_workspace_bytes_rs": (c_int64, [c_int, c_int, c_int, c_int]),
    "ganb200_dwconv2d_wgrad": (c_int, [P, P, P, P, c_int64] + [c_int] * 8 + [c_float, P]),
    "ganb200_dwconv2d_wgrad_static_max_bytes": (c_int64, []),
    "ganb200_dwconv2d_wgrad_static": (c_int, [P, P, P] + [c_int] * 8 + [c_float, P]),
    "ganb200_chan_shuffle": (c_int, [P, P, c_int, c_int, c_int, c_int, P]),
}

_lib = None


class GanB200Error(RuntimeError):
    pass


def load():
    """Load libganb200.so and declare prototypes.  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise GanB200Error(
            f"{LIB_PATH} not found: the CUDA extension is not built (run `make -C gan_b200/csrc`). "
            "gan_b200 has no CPU or PyTorch fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc):
    if rc != 0:
        msg = load().ganb200_last_error()
        raise GanB200Error(f"libganb200 error {rc}: {msg.decode() if msg else ''}")
