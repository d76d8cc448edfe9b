"""ctypes binding of libjumbo_mae_b200.so (the C ABI declared in include/jumbo_mae_b200.h).

There is no fallback: if the library is missing or the device is not a B200 the
import of any op raises.  torch is used only for device memory and streams.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, c_char_p, c_float, c_int, c_int32, c_long, c_size_t, c_void_p

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libjumbo_mae_b200.so")

# Every symbol include/jumbo_mae_b200.h declares; tests check that the library exports them all.
EXPORTS = [
    "jm_last_error", "jm_version", "jm_device_ok", "jm_launch_count", "jm_mask_sort", "jm_index_sequence", "jm_patchify",
    "jm_extract_patches", "jm_patch_loss_workspace", "jm_patch_loss", "jm_patch_loss_static", "jm_layernorm_fwd",
    "jm_layernorm_bwd_workspace", "jm_layernorm_bwd", "jm_gemm", "jm_attention_fwd", "jm_attention_bwd",
    "jm_fill_cls", "jm_unshuffle_fwd", "jm_unshuffle_bwd", "jm_colsum", "jm_adamw", "jm_cast_bf16",
    "jm_add_bias_cast", "jm_clip_by_global_norm", "jm_softmax_ce", "jm_patchify_mix", "jm_patch_mse_loss", "jm_scale_rows_bf16", "jm_rrc_flip_workspace", "jm_rrc_flip", "jm_posemb_grad", "jm_layernorm_bwd_lowp",
    "jm_layernorm_fwd_rows", "jm_layernorm_bwd_rows", "jm_masked_targets", "jm_row_loss_workspace", "jm_row_loss_finalize", "jm_lamb_workspace", "jm_lamb", "jm_zero", "jm_patchify_mix_dev", "jm_softmax_ce_dev", "jm_layernorm_bwd_stream16", "jm_mask_sort_batched", "jm_patchify_pi", "jm_masked_targets_pi",
    "jm_unshuffle_fwd_pi", "jm_unshuffle_bwd_pi", "jm_posemb_grad_pi", "jm_scale_bf16", "jm_layerscale_grad", "jm_layerscale_cls_grad",
]


class JmError(RuntimeError):
    pass


class GemmArgs(ctypes.Structure):
    _fields_ = [
        ("M", c_int), ("N", c_int), ("K", c_int),
        ("A", c_void_p), ("lda", c_long), ("a_mn_major", c_int),
        ("B", c_void_p), ("ldb", c_long), ("b_mn_major", c_int),
        ("out_f32", c_void_p), ("ldo_f32", c_long),
        ("out_bf16", c_void_p), ("ldo_bf16", c_long),
        ("out2_bf16", c_void_p), ("ldo2_bf16", c_long),
        ("bias", c_void_p),
        ("residual", c_void_p), ("ldr", c_long),
        ("aux", c_void_p), ("ldaux", c_long),
        ("pe", c_void_p), ("pe_idx", c_void_p), ("ldpe", c_long),
        ("act", c_int),
        ("col_scale", c_float), ("col_scale_n", c_int),
        ("row_period", c_int), ("row_skip", c_int), ("row_mask_mode", c_int),
        ("accumulate", c_int),
        ("splits", c_int),
        ("block_n", c_int),
        ("colsum", c_void_p),
        ("cta_pair", c_int),
        ("row_scale", c_void_p), ("row_scale_period", c_int),
        ("loss_target", c_void_p), ("ld_target", c_long), ("loss_stats", c_void_p), ("row_loss", c_void_p),
        ("loss_scale", c_float),
        ("col_scale_vec", c_void_p),
        ("pe_idx_stride", c_long),
    ]


_lib = None


def load() -> ctypes.CDLL:
    """Loads the library (once).  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise JmError(
            f"{LIB_PATH} is missing: build it with `python -m jumbo_mae_tpu_b200.build` "
            "(there is no CPU or PyTorch fallback)")
    lib = ctypes.CDLL(LIB_PATH)
    lib.jm_last_error.restype = c_char_p
    lib.jm_launch_count.restype = ctypes.c_longlong
    lib.jm_patch_loss_workspace.restype = c_size_t
    lib.jm_patch_loss_workspace.argtypes = [c_int, c_int]
    lib.jm_rrc_flip_workspace.restype = c_size_t
    lib.jm_rrc_flip_workspace.argtypes = [c_int, c_int, c_int]
    lib.jm_lamb_workspace.restype = c_size_t
    lib.jm_lamb_workspace.argtypes = [c_int]
    lib.jm_row_loss_workspace.restype = c_size_t
    lib.jm_row_loss_workspace.argtypes = []
    lib.jm_layernorm_bwd_workspace.restype = c_size_t
    lib.jm_layernorm_bwd_workspace.argtypes = [c_int, c_int]
    P = c_void_p
    sig = {
        "jm_mask_sort": [P, c_int, c_int, P, P, P, P],
        "jm_index_sequence": [P, P, c_int, c_int, c_int, c_int, P, P],
        "jm_patchify": [P, c_int, c_int, c_int, P, c_int, c_int, c_int, P, c_long, P],
        "jm_extract_patches": [P, c_int, c_int, c_int, P, P],
        "jm_patch_loss": [P, c_long, P, P, c_int, c_int, c_int, c_int, c_int, c_float, P, P, c_long, P, c_size_t, P],
        "jm_layernorm_bwd_lowp": [P, c_long, P, c_long, P, P, P, c_int, c_int, P, c_long, P, c_long, P, c_long, P, P, P, P,
                                  c_size_t, P],
        "jm_layernorm_fwd_rows": [P, c_long, P, c_int, c_int, P, P, c_float, P, c_long, P, P, P],
        "jm_layernorm_bwd_rows": [P, c_long, P, c_long, P, P, P, P, c_int, c_int, c_long, P, c_long, P, c_long, P, P, P, P,
                                  c_size_t, P],
        "jm_lamb": [P, P, P, P, P, P, c_int, P, c_int] + [c_float] * 8 + [P, c_size_t, P],
        "jm_scale_bf16": [P, c_long, c_long, c_int, P, P, c_int, c_int, P, P, P],
        "jm_layerscale_grad": [P, P, c_long, c_int, c_int, P, P, P, P, P],
        "jm_layerscale_cls_grad": [P, c_long, P, c_long, P, c_long, c_int, c_int, P, P, P],
        "jm_mask_sort_batched": [P, c_int, c_int, c_int, P, P, P, P],
        "jm_patchify_pi": [P, c_int, c_int, c_int, P, c_long, c_int, c_int, c_int, P, c_long, P],
        "jm_masked_targets_pi": [P, c_int, c_int, c_int, P, c_long, c_int, c_int, c_int, P, c_long, P, P, P],
        "jm_unshuffle_fwd_pi": [P, P, c_long, P, P, c_int, c_int, c_int, c_int, c_int, P, P],
        "jm_unshuffle_bwd_pi": [P, P, c_long, c_int, c_int, c_int, c_int, c_int, P, P, P],
        "jm_posemb_grad_pi": [P, c_long, c_int, c_int, c_int, c_int, P, c_long, c_int, P, P],
        "jm_zero": [P, c_size_t, P],
        "jm_layernorm_bwd_stream16": [P, c_long, P, c_long, P, P, P, c_int, c_int, P, c_long, P, c_long, c_int, c_int, P,
                                      c_long, c_int, c_int, P, P, P, c_int, c_int, P, c_size_t, P],
        "jm_masked_targets": [P, c_int, c_int, c_int, P, c_int, c_int, c_int, P, c_long, P, P, P],
        "jm_row_loss_finalize": [P, c_long, c_float, P, P, c_size_t, P],
        "jm_posemb_grad": [P, c_long, c_int, c_int, c_int, c_int, P, c_int, P, P],
        "jm_rrc_flip": [P, P, P, P, P, c_int, c_int, c_int, P, P, c_size_t, P],
        "jm_scale_rows_bf16": [P, c_long, c_long, c_int, P, P, c_int, c_int, P],
        "jm_patch_mse_loss": [P, c_long, P, c_long, P, c_int, c_int, c_int, P, P, c_size_t, P],
        "jm_patch_loss_static": [P, c_long, P, P, c_int, c_int, c_int, c_int, c_int, c_float, P, P, c_long, P],
        "jm_layernorm_fwd": [P, c_long, c_int, c_int, P, P, c_float, P, c_long, P, c_long, P, P, P],
        "jm_layernorm_bwd": [P, c_long, P, c_long, P, P, P, c_int, c_int, P, c_long, P, c_long, P, c_long, c_int,
                             c_int, P, P, P, c_int, c_int, P, c_size_t, P],
        "jm_gemm": [POINTER(GemmArgs), P],
        "jm_attention_fwd": [P, c_long, c_int, c_int, c_int, c_int, P, c_long, P, P],
        "jm_attention_bwd": [P, c_long, P, c_long, P, c_long, P, P, c_int, c_int, c_int, c_int, c_float, P, c_long, P, P],
        "jm_fill_cls": [P, c_int, c_int, c_int, c_int, P, P],
        "jm_unshuffle_fwd": [P, P, P, P, c_int, c_int, c_int, c_int, c_int, P, P],
        "jm_unshuffle_bwd": [P, P, c_int, c_int, c_int, c_int, c_int, P, P, P],
        "jm_colsum": [P, c_int, c_long, c_int, c_int, c_int, c_int, c_int, P, P],
        "jm_adamw": [P, P, P, P, P, c_long, c_long, c_float, c_float, c_float, c_float, c_float, c_float, c_float,
                     c_float, P],
        "jm_cast_bf16": [P, P, c_long, P],
        "jm_add_bias_cast": [P, c_long, c_int, c_int, P, P, c_long, P, c_long, P],
        "jm_clip_by_global_norm": [P, c_long, c_float, c_float, P, P],
        "jm_patchify_mix": [P, c_int, c_int, c_int, P, c_int, c_int, c_int, P, c_long, P, c_int, c_float, P, P, P],
        "jm_softmax_ce": [P, P, P, P, c_float, ctypes.c_double, c_int, c_int, c_float, P, P, P, P, P, P],
        "jm_softmax_ce_dev": [P, P, P, P, P, ctypes.c_double, c_int, c_int, c_float, P, P, P, P, P, P],
        "jm_patchify_mix_dev": [P, c_int, c_int, c_int, P, c_int, c_int, c_int, P, c_long, P, c_int, P, P, P, P],
    }
    for name, argtypes in sig.items():
        fn = getattr(lib, name)
        fn.argtypes = argtypes
        fn.restype = c_int
    _lib = lib
    return lib


def check(rc: int) -> None:
    if rc != 0:
        raise JmError(load().jm_last_error().decode("utf-8", "replace"))


def require_device() -> None:
    """The product path needs a B200; fail loudly otherwise."""
    if not torch.cuda.is_available():
        raise JmError("no CUDA device: jumbo_mae_tpu_b200 has no CPU path")
    if not load().jm_device_ok():
        raise JmError("the current CUDA device is not sm_100 (B200); this library is sm_100a-only")


def ptr(t) -> int | None:
    """Device pointer of a tensor (None -> NULL)."""
    if t is None:
        return None
    return t.data_ptr()


def stream_ptr() -> int:
    return torch.cuda.current_stream().cuda_stream
