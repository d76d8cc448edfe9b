"""Device versions of the reference's MAE helpers (src/utils_mae.py), same names and
argument meaning.  Each call runs a kernel from libjumbo_mae_b200.so on torch's current
stream; tensors are CUDA tensors.  ``rng`` of ``random_masking`` is the uniform noise itself
(f32 [length]): the reference's Flax/threefry stream cannot be reproduced outside JAX, so
the boundary takes the noise as data.
"""
from __future__ import annotations

import torch

from . import _lib, ops


def index_sequence(x: torch.Tensor, ids: torch.Tensor) -> torch.Tensor:
    """src/utils_mae.py:84-85: x[:, ids, ...] for f32 x [B, L, D] and int32 ids."""
    return ops.index_sequence(x, ids.to(torch.int32))


def random_masking(x: torch.Tensor, rng: torch.Tensor, keep_len: int, padding_mask=None):
    """src/utils_mae.py:88-99.  Returns (kept [B, keep, D], mask [B, L] f32, ids_restore [L] i32).
    One permutation per call, shared by the whole batch, exactly as the reference."""
    if padding_mask is not None:
        raise NotImplementedError("padding_mask has no caller in the reference (src/utils_mae.py:98-102)")
    batch, length, _ = x.shape
    if rng.numel() != length:
        raise _lib.JmError(f"noise must have {length} elements, got {rng.numel()}")
    ids_shuffle, ids_restore, mask_row = ops.mask_sort(rng.to(torch.float32).contiguous(), keep_len)
    kept = ops.index_sequence(x, ids_shuffle[:keep_len].contiguous())
    return kept, mask_row[None, :].expand(batch, length), ids_restore


def extract_patches(inputs: torch.Tensor, patch_size: int) -> torch.Tensor:
    """src/utils_mae.py:67-73 fused with the pixel normalisation of src/pretraining.py:90-91:
    ``inputs`` is the uint8 NCHW batch; returns normalised patches [B, L, p*p*3] f32 in
    (ph, pw, c) order."""
    if inputs.dtype != torch.uint8:
        raise _lib.JmError("extract_patches takes the uint8 NCHW batch (normalisation is fused)")
    return ops.extract_patches_u8(inputs, patch_size)


def patch_mse_loss(patch_output: torch.Tensor, patch_target: torch.Tensor, valid: torch.Tensor | None = None
                   ) -> torch.Tensor:
    """src/utils_mae.py:51-64, same signature: mean over the batch of the valid patches' mean squared error.
    patch_output / patch_target are f32 [B, L, P], valid f32 [B, L] (``None``: every patch counts)."""
    if patch_output.shape != patch_target.shape or patch_output.dim() != 3:
        raise _lib.JmError("patch_output and patch_target must both be [B, L, P]")
    B, L, _ = patch_output.shape
    if valid is not None:
        valid = valid.to(torch.float32).expand(B, L).contiguous()
    loss = torch.zeros(1, dtype=torch.float32, device=patch_output.device)
    ws = torch.empty(ops.patch_loss_workspace(B, L), dtype=torch.uint8, device=patch_output.device)
    ops.patch_mse_loss(patch_output.to(torch.float32).contiguous(), patch_target.to(torch.float32).contiguous(), valid,
                       loss, ws)
    return loss[0]


def patch_mse_loss_from_images(patch_output: torch.Tensor, images: torch.Tensor, valid: torch.Tensor, patch_size: int,
                               norm_pix: bool = False) -> torch.Tensor:
    """The fused form the training step uses: the target is computed from ``images`` (uint8 NCHW) on the fly
    (src/pretraining.py:113-118 + src/utils_mae.py:51-64), so the f32 target never exists in memory.
    patch_output is [B, L, P] f32; valid is the [L] or [B, L] mask (one row: the mask is shared by the batch)."""
    B, L, P = patch_output.shape
    mask_row = (valid[0] if valid.dim() == 2 else valid).to(torch.float32).contiguous()
    loss = torch.zeros(1, dtype=torch.float32, device=patch_output.device)
    ws = torch.empty(ops.patch_loss_workspace(B, L), dtype=torch.uint8, device=patch_output.device)
    ops.patch_loss(patch_output.reshape(B * L, P).contiguous(), images, mask_row, patch_size, 0, norm_pix, 1.0, loss,
                   None, ws)
    return loss[0]
