"""Module descriptors with the reference's names and constructor fields (src/modeling.py).

In the reference these are Flax modules whose ``__call__`` traces jnp ops; here they carry the
same hyper-parameters and hand the arithmetic to the kernel schedule in ``engine.py``.  Only
the fields the pre-training presets exercise are accepted with non-default values; the rest
raise at construction so that a silently different model cannot be built.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Literal

import torch

from . import _lib


@dataclass
class ViTBase:
    """src/modeling.py:35-70."""
    layers: int = 12
    dim: int = 768
    heads: int = 12
    labels: int | None = 1000
    layerscale: bool = False
    patch_size: int = 16
    image_size: int = 224
    posemb: Literal["learnable", "sincos2d"] = "sincos2d"
    pooling: Literal["cls", "gap"] = "cls"      # accepted and unused, as in the reference (:46)
    dropout: float = 0.0
    droppath: float = 0.0
    grad_ckpt: bool = False
    image_mask_ratio: float = 0.75
    linear_probing: bool = False
    batch_norm: bool = False

    @property
    def head_dim(self) -> int:
        return self.dim // self.heads

    @property
    def hidden_dim(self) -> int:
        return 4 * self.dim

    @property
    def num_patches(self) -> tuple[int, int]:
        return (self.image_size // self.patch_size,) * 2


@dataclass
class MAEDecoderBase:
    """src/modeling.py:72-104."""
    dec_layers: int = 6
    dec_dim: int = 512
    dec_heads: int = 8
    dec_layerscale: bool = False
    dec_posemb: Literal["learnable", "sincos2d"] = "sincos2d"   # ignored by the reference too (:284)
    dec_dropout: float = 0.0
    dec_droppath: float = 0.0
    grad_ckpt: bool = False
    patch_size: int = 16
    image_size: int = 224

    @property
    def head_dim(self) -> int:
        return self.dec_dim // self.dec_heads

    @property
    def hidden_dim(self) -> int:
        return 4 * self.dec_dim

    @property
    def num_patches(self) -> tuple[int, int]:
        return (self.image_size // self.patch_size,) * 2


@dataclass
class ViT(ViTBase):
    """ViT with the Jumbo CLS token, MAE branch (src/modeling.py:221-267).

    ``encode(store, images, noise)`` runs PatchEmbed + CLS + random_masking + JumboLayer x L + norm
    and returns ``(x [B, J+keep, dim] f32, mask [B, L] f32, ids_restore [L] i32)`` like
    ``ViT.__call__`` in MAE mode (``labels <= 0``).  images are the uint8 NCHW batch: the pixel
    normalisation of PretrainModule (src/pretraining.py:90-91) is fused into the patch loader."""
    num_cls_tokens: int = 3

    def __post_init__(self):
        if self.labels is not None and self.labels > 0:
            raise NotImplementedError("classification head (labels > 0) is the fine-tune path, not built yet")
        if self.layerscale or self.dropout or self.droppath or self.posemb != "sincos2d":
            raise NotImplementedError("layerscale / dropout / droppath / learnable posemb are off in the presets")

    def encode(self, engine, images: torch.Tensor, noise: torch.Tensor):
        if engine.cfg.dim != self.dim or engine.cfg.layers != self.layers:
            raise _lib.JmError("engine was built for a different encoder")
        engine.forward(images, noise, want_grad=False)
        B = images.shape[0]
        x = torch.empty(B * engine.Te, self.dim, dtype=torch.float32, device=images.device)
        from . import ops
        ops.layernorm_fwd(engine.xa[self.layers], engine.ps.f("model/norm/scale"), engine.ps.f("model/norm/bias"),
                          1e-6, None, x, None, None)
        mask = engine.mask_row[None, :].expand(B, engine.mask_row.numel())
        return x.view(B, engine.Te, self.dim), mask, engine.ids_restore


@dataclass
class MAEDecoder(MAEDecoderBase):
    """src/modeling.py:276-298."""
    num_cls_tokens: int = 3

    def __post_init__(self):
        if self.dec_layerscale or self.dec_dropout or self.dec_droppath:
            raise NotImplementedError("decoder layerscale / dropout / droppath are off in the presets")
