"""Tensor-level wrappers over the C ABI (include/jumbo_mae_b200.h).

Each function validates dtypes/devices, derives row strides from the tensors and
enqueues the kernel on torch's current stream.  No arithmetic happens in Python or in
PyTorch here: without the CUDA library these functions raise.
"""
from __future__ import annotations

import ctypes
import os
from typing import Optional

import torch

from . import _lib
from ._lib import GemmArgs, check, ptr, stream_ptr

BF16, F32, I32, U8 = torch.bfloat16, torch.float32, torch.int32, torch.uint8


def _chk(t: Optional[torch.Tensor], dtype, name: str) -> None:
    if t is None:
        return
    if not t.is_cuda:
        raise _lib.JmError(f"{name} must be a CUDA tensor")
    if t.dtype != dtype:
        raise _lib.JmError(f"{name} must be {dtype}, got {t.dtype}")


def _expect(cond: bool, msg: str) -> None:
    if not cond:
        raise _lib.JmError(msg)


def _ld(t: Optional[torch.Tensor]) -> int:
    """Row stride (elements) of a 2-D view whose last dim is contiguous."""
    if t is None:
        return 0
    if t.dim() < 2:
        return t.shape[-1] if t.dim() == 1 else 0
    if t.stride(-1) != 1:
        raise _lib.JmError("innermost dimension must be contiguous")
    return t.stride(-2)


# ------------------------------------------------------------------ masking
def mask_sort(noise: torch.Tensor, keep: int):
    """noise f32 [L] -> (ids_shuffle i32 [L], ids_restore i32 [L], mask_row f32 [L])."""
    _chk(noise, F32, "noise")
    L = noise.numel()
    ids_shuffle = torch.empty(L, dtype=I32, device=noise.device)
    ids_restore = torch.empty(L, dtype=I32, device=noise.device)
    mask_row = torch.empty(L, dtype=F32, device=noise.device)
    check(_lib.load().jm_mask_sort(ptr(noise), L, keep, ptr(ids_shuffle), ptr(ids_restore), ptr(mask_row),
                                   stream_ptr()))
    return ids_shuffle, ids_restore, mask_row


def index_sequence(x: torch.Tensor, ids: torch.Tensor) -> torch.Tensor:
    _chk(x, F32, "x"), _chk(ids, I32, "ids")
    x = x.contiguous()
    B, L, D = x.shape
    out = torch.empty(B, ids.numel(), D, dtype=F32, device=x.device)
    check(_lib.load().jm_index_sequence(ptr(x), ptr(ids), B, L, ids.numel(), D, ptr(out), stream_ptr()))
    return out


# ------------------------------------------------------------------ pixels
def patchify(images: torch.Tensor, patch_size: int, ids: Optional[torch.Tensor], n_sel: int, rows_per_image: int,
             row_off: int, out: torch.Tensor) -> torch.Tensor:
    _chk(images, U8, "images"), _chk(ids, I32, "ids"), _chk(out, BF16, "out")
    B, C, H, W = images.shape
    assert C == 3 and H == W and images.is_contiguous()
    if ids is not None and ids.dim() == 2:         # one index row per image (per-image masking)
        check(_lib.load().jm_patchify_pi(ptr(images), B, H, patch_size, ptr(ids), ids.stride(0), n_sel, rows_per_image,
                                         row_off, ptr(out), _ld(out), stream_ptr()))
        return out
    check(_lib.load().jm_patchify(ptr(images), B, H, patch_size, ptr(ids), n_sel, rows_per_image, row_off, ptr(out),
                                  _ld(out), stream_ptr()))
    return out


def patchify_mix(images: torch.Tensor, patch_size: int, ids: Optional[torch.Tensor], n_sel: int, rows_per_image: int,
                 row_off: int, out: torch.Tensor, perm: torch.Tensor, *, ratio: float = 1.0,
                 xmask: Optional[torch.Tensor] = None, ymask: Optional[torch.Tensor] = None,
                 ratio_dev: Optional[torch.Tensor] = None) -> torch.Tensor:
    """patchify with Mixup (``xmask is None``: ratio * a + (1 - ratio) * a[perm]) or CutMix (uint8 ``xmask`` /
    ``ymask`` [image_size]: a[perm] inside the box) on the normalised pixels (jm_patchify_mix; src/utils.py:67-90)."""
    _chk(images, U8, "images"), _chk(ids, I32, "ids"), _chk(out, BF16, "out"), _chk(perm, I32, "perm")
    _chk(xmask, U8, "xmask"), _chk(ymask, U8, "ymask")
    B, C, H, W = images.shape
    _expect(C == 3 and H == W and images.is_contiguous(), "images must be contiguous [B, 3, S, S]")
    _expect(perm.shape == (B,) and perm.is_contiguous(), "perm must be int32 [B]")
    _expect((xmask is None) == (ymask is None), "give both xmask and ymask (cutmix) or neither (mixup)")
    if xmask is not None:
        _expect(xmask.numel() == W and ymask.numel() == H, "xmask / ymask must have image_size entries")
    if ratio_dev is not None:              # the ratio as a device scalar (graph replays with new draws)
        _chk(ratio_dev, F32, "ratio_dev")
        check(_lib.load().jm_patchify_mix_dev(ptr(images), B, H, patch_size, ptr(ids), n_sel, rows_per_image, row_off,
                                              ptr(out), _ld(out), ptr(perm), 1 if xmask is None else 2, ptr(ratio_dev),
                                              ptr(xmask), ptr(ymask), stream_ptr()))
        return out
    check(_lib.load().jm_patchify_mix(ptr(images), B, H, patch_size, ptr(ids), n_sel, rows_per_image, row_off, ptr(out),
                                      _ld(out), ptr(perm), 1 if xmask is None else 2, ratio, ptr(xmask), ptr(ymask),
                                      stream_ptr()))
    return out


def extract_patches_u8(images: torch.Tensor, patch_size: int) -> torch.Tensor:
    _chk(images, U8, "images")
    B, C, H, W = images.shape
    g = H // patch_size
    out = torch.empty(B, g * g, patch_size * patch_size * 3, dtype=F32, device=images.device)
    check(_lib.load().jm_extract_patches(ptr(images.contiguous()), B, H, patch_size, ptr(out), stream_ptr()))
    return out


def patch_loss_workspace(batch: int, num_patches: int) -> int:
    return int(_lib.load().jm_patch_loss_workspace(batch, num_patches))


def patch_loss(pred: torch.Tensor, images: torch.Tensor, mask_row: torch.Tensor, patch_size: int, num_cls: int,
               norm_pix: bool, grad_scale: float, loss: torch.Tensor, dpred: Optional[torch.Tensor],
               workspace: torch.Tensor) -> None:
    _chk(pred, F32, "pred"), _chk(images, U8, "images"), _chk(mask_row, F32, "mask_row")
    _chk(loss, F32, "loss"), _chk(dpred, BF16, "dpred")
    B, _, H, _ = images.shape
    check(_lib.load().jm_patch_loss(ptr(pred), _ld(pred), ptr(images), ptr(mask_row), B, H, patch_size, num_cls,
                                    int(norm_pix), grad_scale, ptr(loss), ptr(dpred), _ld(dpred), ptr(workspace),
                                    workspace.numel() * workspace.element_size(), stream_ptr()))


def posemb_grad(dx: torch.Tensor, batch: int, T: int, J: int, n: int, ids: Optional[torch.Tensor], dim: int,
                dwpe: torch.Tensor) -> None:
    """dwpe[ids[i]] += sum_b dx[b*T + J + i] (learnable position embedding, src/modeling.py:115-118,123)."""
    _chk(dx, F32, "dx"), _chk(ids, I32, "ids"), _chk(dwpe, F32, "dwpe")
    if ids is not None and ids.dim() == 2:
        check(_lib.load().jm_posemb_grad_pi(ptr(dx), _ld(dx), batch, T, J, n, ptr(ids), ids.stride(0), dim, ptr(dwpe),
                                            stream_ptr()))
        return
    check(_lib.load().jm_posemb_grad(ptr(dx), _ld(dx), batch, T, J, n, ptr(ids), dim, ptr(dwpe), stream_ptr()))


def scale_rows_bf16(x: torch.Tensor, rows: int, cols: int, scale: torch.Tensor, period: int,
                    scale_cls: Optional[torch.Tensor] = None, skip: int = 0, ld: Optional[int] = None) -> None:
    """In place x[r] *= (r % period < skip ? scale_cls : scale)[r // period] (DropPath backward)."""
    _chk(x, BF16, "x"), _chk(scale, F32, "scale"), _chk(scale_cls, F32, "scale_cls")
    check(_lib.load().jm_scale_rows_bf16(ptr(x), _ld(x) if ld is None else ld, rows, cols, ptr(scale), ptr(scale_cls),
                                         period, skip, stream_ptr()))


def scale_bf16(x: torch.Tensor, rows: int, cols: int, period: int, scale: Optional[torch.Tensor] = None,
               scale_cls: Optional[torch.Tensor] = None, skip: int = 0, col_scale: Optional[torch.Tensor] = None,
               col_scale_cls: Optional[torch.Tensor] = None, ld: Optional[int] = None) -> None:
    """In place x[r, c] *= row factor (DropPath) * column factor (LayerScale); CLS rows (r % period < skip) have their own."""
    _chk(x, BF16, "x"), _chk(scale, F32, "scale"), _chk(scale_cls, F32, "scale_cls"), _chk(col_scale, F32, "col_scale")
    _chk(col_scale_cls, F32, "col_scale_cls")
    check(_lib.load().jm_scale_bf16(ptr(x), _ld(x) if ld is None else ld, rows, cols, ptr(scale), ptr(scale_cls), period,
                                    skip, ptr(col_scale), ptr(col_scale_cls), stream_ptr()))


def layerscale_grad(W: torch.Tensor, dW: torch.Tensor, b: torch.Tensor, db: torch.Tensor, scale: torch.Tensor,
                    dscale: torch.Tensor, K: int, N: int) -> None:
    """dscale += (colsum(W * dW) + b * db) / scale from the accumulated gradients of the scaled Dense."""
    for t, nm in ((W, "W"), (dW, "dW"), (b, "b"), (db, "db"), (scale, "scale"), (dscale, "dscale")):
        _chk(t, F32, nm)
    check(_lib.load().jm_layerscale_grad(ptr(W), ptr(dW), _ld(W), K, N, ptr(b), ptr(db), ptr(scale), ptr(dscale),
                                         stream_ptr()))


def layerscale_cls_grad(dy: torch.Tensor, c_out: torch.Tensor, c: torch.Tensor, batch: int, N: int, scale: torch.Tensor,
                        dscale: torch.Tensor, lddy: int, ldo: int, ldc: int) -> None:
    """dscale3 += sum_b dy (c_out - c) / scale3 over [batch, N] f32 views (row strides given)."""
    for t, nm in ((dy, "dy"), (c_out, "c_out"), (c, "c"), (scale, "scale"), (dscale, "dscale")):
        _chk(t, F32, nm)
    check(_lib.load().jm_layerscale_cls_grad(ptr(dy), lddy, ptr(c_out), ldo, ptr(c), ldc, batch, N, ptr(scale),
                                             ptr(dscale), stream_ptr()))


def patch_mse_loss(patch_output: torch.Tensor, patch_target: torch.Tensor, valid: Optional[torch.Tensor],
                   loss: torch.Tensor, workspace: torch.Tensor) -> None:
    """src/utils_mae.py:51-64 on explicit f32 targets [B, L, P]; valid [B, L] f32 or None."""
    _chk(patch_output, F32, "patch_output"), _chk(patch_target, F32, "patch_target"), _chk(valid, F32, "valid")
    _chk(loss, F32, "loss")
    B, L, P = patch_output.shape
    check(_lib.load().jm_patch_mse_loss(ptr(patch_output), P, ptr(patch_target), P, ptr(valid), B, L, P, ptr(loss),
                                        ptr(workspace), workspace.numel() * workspace.element_size(), stream_ptr()))


# ------------------------------------------------------------------ LayerNorm
def layernorm_fwd(x: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor, eps: float,
                  y_bf16: Optional[torch.Tensor], y_f32: Optional[torch.Tensor], mean: Optional[torch.Tensor],
                  rstd: Optional[torch.Tensor], rows: Optional[int] = None, dim: Optional[int] = None,
                  ldx: Optional[int] = None, ldy: Optional[int] = None, ldyf: Optional[int] = None) -> None:
    _chk(x, F32, "x"), _chk(gamma, F32, "gamma"), _chk(beta, F32, "beta")
    _chk(y_bf16, BF16, "y_bf16"), _chk(y_f32, F32, "y_f32"), _chk(mean, F32, "mean"), _chk(rstd, F32, "rstd")
    rows = x.shape[0] if rows is None else rows
    dim = x.shape[-1] if dim is None else dim
    check(_lib.load().jm_layernorm_fwd(ptr(x), _ld(x) if ldx is None else ldx, rows, dim, ptr(gamma), ptr(beta), eps,
                                       ptr(y_bf16), (_ld(y_bf16) if ldy is None else ldy), ptr(y_f32),
                                       (_ld(y_f32) if ldyf is None else ldyf), ptr(mean), ptr(rstd), stream_ptr()))


def layernorm_bwd_workspace(rows: int, dim: int) -> int:
    return int(_lib.load().jm_layernorm_bwd_workspace(rows, dim))


def layernorm_bwd(dy: torch.Tensor, x: torch.Tensor, mean: torch.Tensor, rstd: torch.Tensor, gamma: torch.Tensor,
                  dres: Optional[torch.Tensor], dx: Optional[torch.Tensor], dx_bf16: Optional[torch.Tensor],
                  dgamma: torch.Tensor, dbeta: torch.Tensor, workspace: torch.Tensor, row_period: int = 0,
                  row_skip: int = 0, rows: Optional[int] = None, dim: Optional[int] = None,
                  lddy: Optional[int] = None, ldx: Optional[int] = None, lddres: Optional[int] = None,
                  lddx: Optional[int] = None, lddxb: Optional[int] = None, dcol: Optional[torch.Tensor] = None,
                  dcol_period: int = 0, dcol_skip: int = 0) -> None:
    _chk(dcol, F32, "dcol")
    _chk(dy, BF16, "dy"), _chk(x, F32, "x"), _chk(mean, F32, "mean"), _chk(rstd, F32, "rstd")
    _chk(gamma, F32, "gamma"), _chk(dres, F32, "dres"), _chk(dx, F32, "dx"), _chk(dx_bf16, BF16, "dx_bf16")
    _chk(dgamma, F32, "dgamma"), _chk(dbeta, F32, "dbeta")
    rows = x.shape[0] if rows is None else rows
    dim = x.shape[-1] if dim is None else dim
    check(_lib.load().jm_layernorm_bwd(
        ptr(dy), _ld(dy) if lddy is None else lddy, ptr(x), _ld(x) if ldx is None else ldx, ptr(mean), ptr(rstd),
        ptr(gamma), rows, dim, ptr(dres), (_ld(dres) if lddres is None else lddres), ptr(dx),
        (_ld(dx) if lddx is None else lddx), ptr(dx_bf16), (_ld(dx_bf16) if lddxb is None else lddxb), row_period,
        row_skip, ptr(dgamma), ptr(dbeta), ptr(dcol), dcol_period, dcol_skip, ptr(workspace),
        workspace.numel() * workspace.element_size(), stream_ptr()))


def layernorm_bwd_lowp(dy: torch.Tensor, x: torch.Tensor, mean: torch.Tensor, rstd: torch.Tensor, gamma: torch.Tensor,
                       dres_bf16: Optional[torch.Tensor], dx: Optional[torch.Tensor], dx_bf16: torch.Tensor,
                       dgamma: torch.Tensor, dbeta: torch.Tensor, workspace: torch.Tensor,
                       dcol: Optional[torch.Tensor] = None) -> None:
    """LayerNorm backward with the residual gradient read (and written) in bf16 only (dim <= 512)."""
    _chk(dy, BF16, "dy"), _chk(x, F32, "x"), _chk(mean, F32, "mean"), _chk(rstd, F32, "rstd"), _chk(gamma, F32, "gamma")
    _chk(dres_bf16, BF16, "dres_bf16"), _chk(dx, F32, "dx"), _chk(dx_bf16, BF16, "dx_bf16"), _chk(dcol, F32, "dcol")
    _chk(dgamma, F32, "dgamma"), _chk(dbeta, F32, "dbeta")
    rows, dim = x.shape[0], x.shape[-1]
    check(_lib.load().jm_layernorm_bwd_lowp(
        ptr(dy), _ld(dy), ptr(x), _ld(x), ptr(mean), ptr(rstd), ptr(gamma), rows, dim, ptr(dres_bf16), _ld(dres_bf16),
        ptr(dx), _ld(dx), ptr(dx_bf16), _ld(dx_bf16), ptr(dgamma), ptr(dbeta), ptr(dcol), ptr(workspace),
        workspace.numel() * workspace.element_size(), stream_ptr()))


def layernorm_bwd_stream16(dy: torch.Tensor, x: torch.Tensor, mean: torch.Tensor, rstd: torch.Tensor,
                           gamma: torch.Tensor, dres_bf16: Optional[torch.Tensor], dx: Optional[torch.Tensor],
                           dx_bf16: torch.Tensor, dgamma: torch.Tensor, dbeta: torch.Tensor, workspace: torch.Tensor,
                           f32_period: int = 0, f32_skip: int = 0, row_period: int = 0, row_skip: int = 0,
                           dcol: Optional[torch.Tensor] = None, dcol_period: int = 0, dcol_skip: int = 0) -> None:
    """LayerNorm backward on the bf16 gradient stream (dim <= 1280); f32 dx only for rows r % f32_period < f32_skip."""
    _chk(dy, BF16, "dy"), _chk(x, F32, "x"), _chk(mean, F32, "mean"), _chk(rstd, F32, "rstd"), _chk(gamma, F32, "gamma")
    _chk(dres_bf16, BF16, "dres_bf16"), _chk(dx, F32, "dx"), _chk(dx_bf16, BF16, "dx_bf16"), _chk(dcol, F32, "dcol")
    _chk(dgamma, F32, "dgamma"), _chk(dbeta, F32, "dbeta")
    rows, dim = x.shape[0], x.shape[-1]
    check(_lib.load().jm_layernorm_bwd_stream16(
        ptr(dy), _ld(dy), ptr(x), _ld(x), ptr(mean), ptr(rstd), ptr(gamma), rows, dim, ptr(dres_bf16), _ld(dres_bf16),
        ptr(dx), _ld(dx), f32_period, f32_skip, ptr(dx_bf16), _ld(dx_bf16), row_period, row_skip, ptr(dgamma),
        ptr(dbeta), ptr(dcol), dcol_period, dcol_skip, ptr(workspace), workspace.numel() * workspace.element_size(),
        stream_ptr()))


def layernorm_fwd_rows(x: torch.Tensor, row_map: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor, eps: float,
                       y_bf16: torch.Tensor, mean: torch.Tensor, rstd: torch.Tensor) -> None:
    """y[r] = LayerNorm(x[row_map[r]]): the rows that matter, compact (dec_norm in front of the pixel head)."""
    _chk(x, F32, "x"), _chk(row_map, I32, "row_map"), _chk(gamma, F32, "gamma"), _chk(beta, F32, "beta")
    _chk(y_bf16, BF16, "y_bf16"), _chk(mean, F32, "mean"), _chk(rstd, F32, "rstd")
    check(_lib.load().jm_layernorm_fwd_rows(ptr(x), _ld(x), ptr(row_map), row_map.numel(), x.shape[-1], ptr(gamma),
                                            ptr(beta), eps, ptr(y_bf16), _ld(y_bf16), ptr(mean), ptr(rstd),
                                            stream_ptr()))


def layernorm_bwd_rows(dy: torch.Tensor, x: torch.Tensor, row_map: torch.Tensor, mean: torch.Tensor,
                       rstd: torch.Tensor, gamma: torch.Tensor, dx: Optional[torch.Tensor],
                       dx_bf16: Optional[torch.Tensor], dgamma: torch.Tensor, dbeta: torch.Tensor,
                       workspace: torch.Tensor, dcol: Optional[torch.Tensor] = None) -> None:
    """Backward of layernorm_fwd_rows: dx / dx_bf16 [all rows, dim] zeroed, then the mapped rows scattered."""
    _chk(dy, BF16, "dy"), _chk(x, F32, "x"), _chk(row_map, I32, "row_map"), _chk(mean, F32, "mean")
    _chk(rstd, F32, "rstd"), _chk(gamma, F32, "gamma"), _chk(dx, F32, "dx"), _chk(dx_bf16, BF16, "dx_bf16")
    _chk(dgamma, F32, "dgamma"), _chk(dbeta, F32, "dbeta"), _chk(dcol, F32, "dcol")
    check(_lib.load().jm_layernorm_bwd_rows(
        ptr(dy), _ld(dy), ptr(x), _ld(x), ptr(row_map), ptr(mean), ptr(rstd), ptr(gamma), row_map.numel(), x.shape[-1],
        x.shape[0], ptr(dx), _ld(dx), ptr(dx_bf16), _ld(dx_bf16), ptr(dgamma), ptr(dbeta), ptr(dcol), ptr(workspace),
        workspace.numel() * workspace.element_size(), stream_ptr()))


def masked_targets(images: torch.Tensor, patch_size: int, ids: torch.Tensor, num_cls: int, norm_pix: bool,
                   targets: torch.Tensor, stats: Optional[torch.Tensor], row_map: Optional[torch.Tensor]) -> None:
    """Byte targets (+ norm-pix statistics, + decoder rows) of the patches ``ids`` for the loss epilogue of gemm(act=5)."""
    _chk(images, torch.uint8, "images"), _chk(ids, I32, "ids"), _chk(targets, torch.uint8, "targets")
    _chk(stats, F32, "stats"), _chk(row_map, I32, "row_map")
    if ids.dim() == 2:
        check(_lib.load().jm_masked_targets_pi(ptr(images), images.shape[0], images.shape[-1], patch_size, ptr(ids),
                                               ids.stride(0), ids.shape[1], num_cls, int(norm_pix), ptr(targets),
                                               _ld(targets), ptr(stats), ptr(row_map), stream_ptr()))
        return
    check(_lib.load().jm_masked_targets(ptr(images), images.shape[0], images.shape[-1], patch_size, ptr(ids), ids.numel(),
                                        num_cls, int(norm_pix), ptr(targets), _ld(targets), ptr(stats), ptr(row_map),
                                        stream_ptr()))


def row_loss_workspace() -> int:
    return int(_lib.load().jm_row_loss_workspace())


def row_loss_finalize(row_loss: torch.Tensor, scale: float, loss: torch.Tensor, workspace: torch.Tensor) -> None:
    _chk(row_loss, F32, "row_loss"), _chk(loss, F32, "loss")
    check(_lib.load().jm_row_loss_finalize(ptr(row_loss), row_loss.numel(), scale, ptr(loss), ptr(workspace),
                                           workspace.numel() * workspace.element_size(), stream_ptr()))


# ------------------------------------------------------------------ GEMM
def gemm(A: torch.Tensor, B: torch.Tensor, M: int, N: int, K: int, *, a_mn: bool = False, b_mn: bool = False,
         lda: Optional[int] = None, ldb: Optional[int] = None,
         out_f32: Optional[torch.Tensor] = None, ldo_f32: Optional[int] = None,
         out_bf16: Optional[torch.Tensor] = None, ldo_bf16: Optional[int] = None,
         out2_bf16: Optional[torch.Tensor] = None, ldo2_bf16: Optional[int] = None,
         bias: Optional[torch.Tensor] = None,
         residual: Optional[torch.Tensor] = None, ldr: Optional[int] = None,
         aux: Optional[torch.Tensor] = None, ldaux: Optional[int] = None,
         pe: Optional[torch.Tensor] = None, pe_idx: Optional[torch.Tensor] = None,
         act: int = 0, col_scale: float = 1.0, col_scale_n: int = 0,
         row_period: int = 0, row_skip: int = 0, row_mask_mode: int = 0,
         accumulate: bool = False, splits: int = 0, block_n: int = 0,
         colsum: Optional[torch.Tensor] = None, cta_pair: int = 0,
         row_scale: Optional[torch.Tensor] = None, row_scale_period: int = 0,
         loss_target: Optional[torch.Tensor] = None, loss_stats: Optional[torch.Tensor] = None,
         row_loss: Optional[torch.Tensor] = None, loss_scale: float = 0.0,
         col_vec: Optional[torch.Tensor] = None) -> None:
    """D = epilogue(A @ B) through jm_gemm; see include/jumbo_mae_b200.h for the epilogue order."""
    _chk(A, BF16, "A"), _chk(B, BF16, "B"), _chk(out_f32, F32, "out_f32"), _chk(out_bf16, BF16, "out_bf16")
    _chk(out2_bf16, BF16, "out2_bf16"), _chk(bias, F32, "bias"), _chk(residual, F32, "residual")
    _chk(aux, BF16, "aux"), _chk(pe, F32, "pe"), _chk(pe_idx, I32, "pe_idx")
    a = GemmArgs()
    a.M, a.N, a.K = M, N, K
    a.A, a.lda, a.a_mn_major = ptr(A), (_ld(A) if lda is None else lda), int(a_mn)
    a.B, a.ldb, a.b_mn_major = ptr(B), (_ld(B) if ldb is None else ldb), int(b_mn)
    a.out_f32, a.ldo_f32 = ptr(out_f32), (_ld(out_f32) if ldo_f32 is None else ldo_f32)
    a.out_bf16, a.ldo_bf16 = ptr(out_bf16), (_ld(out_bf16) if ldo_bf16 is None else ldo_bf16)
    a.out2_bf16, a.ldo2_bf16 = ptr(out2_bf16), (_ld(out2_bf16) if ldo2_bf16 is None else ldo2_bf16)
    a.bias = ptr(bias)
    a.residual, a.ldr = ptr(residual), (_ld(residual) if ldr is None else ldr)
    a.aux, a.ldaux = ptr(aux), (_ld(aux) if ldaux is None else ldaux)
    a.pe, a.pe_idx, a.ldpe = ptr(pe), ptr(pe_idx), _ld(pe)
    a.pe_idx_stride = pe_idx.stride(0) if (pe_idx is not None and pe_idx.dim() == 2) else 0
    a.act = act
    a.col_scale, a.col_scale_n = col_scale, col_scale_n
    a.row_period, a.row_skip, a.row_mask_mode = row_period, row_skip, row_mask_mode
    a.accumulate, a.splits, a.block_n = int(accumulate), splits, block_n
    _chk(colsum, F32, "colsum")
    a.colsum = ptr(colsum)
    a.cta_pair = int(cta_pair)
    _chk(row_scale, F32, "row_scale")
    a.row_scale, a.row_scale_period = ptr(row_scale), int(row_scale_period)
    _chk(loss_target, torch.uint8, "loss_target"), _chk(loss_stats, F32, "loss_stats"), _chk(row_loss, F32, "row_loss")
    a.loss_target, a.ld_target = ptr(loss_target), _ld(loss_target)
    a.loss_stats, a.row_loss, a.loss_scale = ptr(loss_stats), ptr(row_loss), float(loss_scale)
    _chk(col_vec, F32, "col_vec")
    a.col_scale_vec = ptr(col_vec)
    check(_lib.load().jm_gemm(ctypes.byref(a), stream_ptr()))


def lamb_workspace(n_chunks: int) -> int:
    return int(_lib.load().jm_lamb_workspace(n_chunks))


def lamb(p: torch.Tensor, g: torch.Tensor, m: torch.Tensor, v: torch.Tensor, p_bf16: Optional[torch.Tensor],
         groups: torch.Tensor, chunks: torch.Tensor, lr: float, b1: float, b2: float, eps: float, wd: float, bc1: float,
         bc2: float, grad_scale: float, workspace: torch.Tensor) -> None:
    """modified_lamb (src/utils.py:124-139) on the kernel leaves listed in ``groups`` / ``chunks`` (params.lamb_tables)."""
    _chk(p, F32, "p"), _chk(g, F32, "g"), _chk(m, F32, "m"), _chk(v, F32, "v"), _chk(p_bf16, BF16, "p_bf16")
    _chk(groups, torch.int64, "groups"), _chk(chunks, torch.int64, "chunks")
    check(_lib.load().jm_lamb(ptr(p), ptr(g), ptr(m), ptr(v), ptr(p_bf16), ptr(groups), groups.shape[0], ptr(chunks),
                              chunks.shape[0], lr, b1, b2, eps, wd, bc1, bc2, grad_scale, ptr(workspace),
                              workspace.numel() * workspace.element_size(), stream_ptr()))


def zero_(t: torch.Tensor) -> None:
    """Zeroes a contiguous device tensor with a memset on the current stream (no kernel launch)."""
    if not t.is_cuda or not t.is_contiguous():
        raise _lib.JmError("zero_ takes a contiguous CUDA tensor")
    check(_lib.load().jm_zero(ptr(t), t.numel() * t.element_size(), stream_ptr()))


# ------------------------------------------------------------------ attention
def attention_fwd(qkv: torch.Tensor, batch: int, T: int, heads: int, head_dim: int, out: torch.Tensor,
                  lse: torch.Tensor) -> None:
    _chk(qkv, BF16, "qkv"), _chk(out, BF16, "out"), _chk(lse, F32, "lse")
    check(_lib.load().jm_attention_fwd(ptr(qkv), _ld(qkv), batch, T, heads, head_dim, ptr(out), _ld(out), ptr(lse),
                                       stream_ptr()))


def attention_bwd(qkv: torch.Tensor, out: torch.Tensor, dout: torch.Tensor, lse: torch.Tensor, batch: int, T: int,
                  heads: int, head_dim: int, q_scale: float, dqkv: torch.Tensor,
                  delta_ws: Optional[torch.Tensor] = None, dbias: Optional[torch.Tensor] = None) -> None:
    _chk(qkv, BF16, "qkv"), _chk(out, BF16, "out"), _chk(dout, BF16, "dout"), _chk(lse, F32, "lse")
    _chk(dqkv, BF16, "dqkv"), _chk(delta_ws, F32, "delta_ws"), _chk(dbias, F32, "dbias")
    # (delta_ws is unused since the kernel computes rowsum(dout * out) itself; accepted for compatibility)
    check(_lib.load().jm_attention_bwd(ptr(qkv), _ld(qkv), ptr(out), _ld(out), ptr(dout), _ld(dout), ptr(lse),
                                       ptr(delta_ws), batch, T, heads, head_dim, q_scale, ptr(dqkv), _ld(dqkv),
                                       ptr(dbias), stream_ptr()))


# ------------------------------------------------------------------ token assembly
def fill_cls(x: torch.Tensor, batch: int, T: int, J: int, dim: int, cls: torch.Tensor) -> None:
    _chk(x, F32, "x"), _chk(cls, F32, "cls")
    check(_lib.load().jm_fill_cls(ptr(x), batch, T, J, dim, ptr(cls), stream_ptr()))


def unshuffle_fwd(y: torch.Tensor, ids_restore: torch.Tensor, mask_emb: torch.Tensor, dec_wpe: torch.Tensor,
                  batch: int, N: int, n: int, J: int, dim: int, out: torch.Tensor) -> None:
    _chk(y, F32, "y"), _chk(ids_restore, I32, "ids_restore"), _chk(mask_emb, F32, "mask_emb")
    _chk(dec_wpe, F32, "dec_wpe"), _chk(out, F32, "out")
    if ids_restore.dim() == 2:
        check(_lib.load().jm_unshuffle_fwd_pi(ptr(y), ptr(ids_restore), ids_restore.stride(0), ptr(mask_emb),
                                              ptr(dec_wpe), batch, N, n, J, dim, ptr(out), stream_ptr()))
        return
    check(_lib.load().jm_unshuffle_fwd(ptr(y), ptr(ids_restore), ptr(mask_emb), ptr(dec_wpe), batch, N, n, J, dim,
                                       ptr(out), stream_ptr()))


def unshuffle_bwd(dout: torch.Tensor, ids_restore: torch.Tensor, batch: int, N: int, n: int, J: int, dim: int,
                  dy: torch.Tensor, dmask_emb: torch.Tensor) -> None:
    _chk(dout, F32, "dout"), _chk(ids_restore, I32, "ids_restore"), _chk(dy, BF16, "dy")
    _chk(dmask_emb, F32, "dmask_emb")
    if ids_restore.dim() == 2:
        check(_lib.load().jm_unshuffle_bwd_pi(ptr(dout), ptr(ids_restore), ids_restore.stride(0), batch, N, n, J, dim,
                                              ptr(dy), ptr(dmask_emb), stream_ptr()))
        return
    check(_lib.load().jm_unshuffle_bwd(ptr(dout), ptr(ids_restore), batch, N, n, J, dim, ptr(dy), ptr(dmask_emb),
                                       stream_ptr()))


def colsum(x: torch.Tensor, rows: int, cols: int, out: torch.Tensor, row_period: int = 0, row_skip: int = 0,
           ldx: Optional[int] = None, row_invert: bool = False) -> None:
    if x.dtype not in (BF16, F32):
        raise _lib.JmError("colsum input must be bf16 or f32")
    _chk(out, F32, "out")
    check(_lib.load().jm_colsum(ptr(x), int(x.dtype == F32), _ld(x) if ldx is None else ldx, rows, cols, row_period,
                                row_skip, int(row_invert), ptr(out), stream_ptr()))


# ------------------------------------------------------------------ optimizer
def adamw(p: torch.Tensor, g: torch.Tensor, m: torch.Tensor, v: torch.Tensor, p_bf16: Optional[torch.Tensor],
          n_decay: int, lr: float, b1: float, b2: float, eps: float, wd: float, bc1: float, bc2: float,
          grad_scale: float = 1.0) -> None:
    for t, nm in ((p, "p"), (g, "g"), (m, "m"), (v, "v")):
        _chk(t, F32, nm)
    _chk(p_bf16, BF16, "p_bf16")
    check(_lib.load().jm_adamw(ptr(p), ptr(g), ptr(m), ptr(v), ptr(p_bf16), p.numel(), n_decay, lr, b1, b2, eps, wd,
                               bc1, bc2, grad_scale, stream_ptr()))


def cast_bf16(x: torch.Tensor, y: torch.Tensor) -> None:
    _chk(x, F32, "x"), _chk(y, BF16, "y")
    check(_lib.load().jm_cast_bf16(ptr(x), ptr(y), x.numel(), stream_ptr()))


def add_bias_cast(src: torch.Tensor, rows: int, cols: int, bias: Optional[torch.Tensor] = None,
                  dst_f32: Optional[torch.Tensor] = None, dst_bf16: Optional[torch.Tensor] = None,
                  lds: Optional[int] = None, ldd: Optional[int] = None, ldd16: Optional[int] = None) -> None:
    """dst = src + bias over a strided [rows, cols] f32 view, as f32 and/or bf16 (jm_add_bias_cast)."""
    _chk(src, F32, "src")
    if bias is not None:
        _chk(bias, F32, "bias")
    if dst_f32 is not None:
        _chk(dst_f32, F32, "dst_f32")
    if dst_bf16 is not None:
        _chk(dst_bf16, BF16, "dst_bf16")
    check(_lib.load().jm_add_bias_cast(ptr(src), lds if lds is not None else cols, rows, cols, ptr(bias), ptr(dst_f32),
                                       ldd if ldd is not None else cols, ptr(dst_bf16),
                                       ldd16 if ldd16 is not None else cols, stream_ptr()))


def clip_by_global_norm(g: torch.Tensor, pre_scale: float, max_norm: float, sumsq: torch.Tensor) -> None:
    """In-place optax.clip_by_global_norm of the flat gradient buffer (jm_clip_by_global_norm)."""
    _chk(g, F32, "g"), _chk(sumsq, F32, "sumsq")
    check(_lib.load().jm_clip_by_global_norm(ptr(g), g.numel(), pre_scale, max_norm, ptr(sumsq), stream_ptr()))


def softmax_ce(logits: torch.Tensor, labels: torch.Tensor, *, label_smoothing: float = 0.0,
               perm: Optional[torch.Tensor] = None, ratio: float = 1.0, grad_scale: Optional[float] = None,
               want_grad: bool = True):
    """Classification loss of the fine-tune step (jm_softmax_ce; src/finetuning.py:90-106): label smoothing, the
    Mixup / CutMix label mix ``ratio * labels + (1 - ratio) * labels[perm]``, softmax cross-entropy, top-1 / top-5
    hits and the gradient with respect to the logits, in one launch.

    ``labels`` is int32 ``[B]`` (class indices) or f32 ``[B, C]`` (soft labels).  Returns
    ``(metrics [3] = mean loss, acc1, acc5; loss_rows [B]; acc1 [B]; acc5 [B]; dlogits [B, C] or None)``;
    ``dlogits`` is d(mean loss)/d(logits) unless ``grad_scale`` is given."""
    _chk(logits, F32, "logits")
    if logits.dim() != 2 or not logits.is_contiguous():
        raise _lib.JmError("logits must be a contiguous [B, C] tensor")
    B, C = logits.shape
    soft = idx = None
    if labels.dim() == 1:
        _chk(labels, I32, "labels"), _expect(labels.shape[0] == B and labels.is_contiguous(), "labels must be [B]")
        idx = labels
    else:
        _chk(labels, F32, "labels")
        _expect(tuple(labels.shape) == (B, C) and labels.is_contiguous(), "soft labels must be contiguous [B, C]")
        soft = labels
    if perm is not None:
        _chk(perm, I32, "perm"), _expect(perm.shape == (B,) and perm.is_contiguous(), "perm must be int32 [B]")
    dev = logits.device
    metrics = torch.empty(3, dtype=F32, device=dev)
    rows = torch.empty(3, B, dtype=F32, device=dev)
    dlogits = torch.empty_like(logits) if want_grad else None
    scale = (1.0 / max(B, 1)) if grad_scale is None else grad_scale
    check(_lib.load().jm_softmax_ce(ptr(logits), ptr(soft), ptr(idx), ptr(perm), ratio, float(label_smoothing), B, C,
                                    scale, ptr(rows[0]), ptr(rows[1]), ptr(rows[2]), ptr(metrics), ptr(dlogits),
                                    stream_ptr()))
    return metrics, rows[0], rows[1], rows[2], dlogits
