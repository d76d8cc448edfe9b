"""Shape / recipe configuration of the MAE pre-training step.

Field names follow the reference's dataclasses and CLI flags (src/modeling.py:35-104,
src/main_pretrain.py:98-167); defaults are the ViT-B/16 1600-epoch recipe
(config/pretrain/pretrain-vit-b16-224-in1k-1600ep.sh).
"""
from __future__ import annotations

import argparse
from dataclasses import dataclass, fields


@dataclass
class MAEConfig:
    layers: int = 12
    dim: int = 768
    heads: int = 12
    patch_size: int = 16
    image_size: int = 224
    image_mask_ratio: float = 0.75
    num_cls_tokens: int = 3            # src/modeling.py:171,222,277; pretraining.py:96 hard-codes 3
    dec_layers: int = 8
    dec_dim: int = 512
    dec_heads: int = 16
    norm_pix_loss: bool = False
    # optimizer (src/pretraining.py:223-259)
    learning_rate: float = 1.5e-4
    train_batch_size: int = 4096
    weight_decay: float = 0.05
    adam_b1: float = 0.9
    adam_b2: float = 0.95
    adam_eps: float = 1e-8
    warmup_steps: int = 1281167 * 40 // 4096
    training_steps: int = 1281167 * 1600 // 4096
    grad_accum: int = 1
    posemb: str = "sincos2d"           # "learnable": model/embed/wpe is a trained parameter (src/modeling.py:115-118)
    layerscale: bool = False           # per-channel scale1/2/3 on the encoder's residual branches, init 1e-4 (src/modeling.py:159-162,185-189)
    dec_layerscale: bool = False       # the same for the decoder's ViTLayers (scale1/2)
    per_image_masking: bool = False    # opt-in extension: noise [B, N], one permutation per image (the reference shares one per batch, src/utils_mae.py:88-96)
    grad_ckpt: bool = False            # nn.remat of every layer (src/modeling.py:232,280): activations are recomputed in backward
    labels: int = 0                    # > 0: classification mode (ViT head, no masking, no decoder; src/modeling.py:239-242,252)

    @property
    def grid(self) -> int:
        return self.image_size // self.patch_size

    @property
    def num_patches(self) -> int:
        return self.grid * self.grid

    @property
    def keep(self) -> int:             # src/modeling.py:255 (MAE mode); every patch in classification mode
        if self.labels > 0:
            return self.num_patches
        return int(self.num_patches * (1.0 - self.image_mask_ratio))

    @property
    def num_masked(self) -> int:       # src/pretraining.py:101
        if self.labels > 0:
            return 0
        return int(self.num_patches * self.image_mask_ratio)

    @property
    def patch_dim(self) -> int:
        return self.patch_size * self.patch_size * 3

    @property
    def head_dim(self) -> int:
        return self.dim // self.heads

    @property
    def dec_head_dim(self) -> int:
        return self.dec_dim // self.dec_heads

    def validate(self) -> None:
        if self.num_cls_tokens != 3:
            raise ValueError("the reference hard-codes 3 CLS tokens in PretrainModule (src/pretraining.py:96,110)")
        if self.image_size % self.patch_size:
            raise ValueError("image_size must be a multiple of patch_size")
        if self.head_dim not in (32, 64, 80) or self.dec_head_dim not in (32, 64, 80):
            raise ValueError("head_dim must be 32, 64 or 80 (attention kernel instantiations)")
        if self.dim % 8 or self.dec_dim % 8:
            raise ValueError("dim and dec_dim must be multiples of 8")
        if self.posemb not in ("sincos2d", "learnable"):
            raise ValueError("posemb must be 'sincos2d' or 'learnable'")
        if self.keep + self.num_masked != self.num_patches:
            raise ValueError("keep + masked != patches for this mask ratio (the reference's concat would fail too)")

    @classmethod
    def from_args(cls, args: argparse.Namespace) -> "MAEConfig":
        """Builds the config from the reference's argparse namespace (src/pretraining.py:171-269)."""
        kw = {}
        for f in fields(cls):
            if hasattr(args, f.name):
                kw[f.name] = getattr(args, f.name)
        return cls(**kw)


def vit_b16(**kw) -> MAEConfig:
    return MAEConfig(**kw)


def vit_l16(**kw) -> MAEConfig:
    kw.setdefault("training_steps", 1281167 * 800 // 4096)
    return MAEConfig(layers=24, dim=1024, heads=16, **kw)


def vit_h14(**kw) -> MAEConfig:
    return MAEConfig(layers=32, dim=1280, heads=16, patch_size=14, **kw)


def flops_per_image(cfg: MAEConfig):
    """Algorithmic (forward, full step) FLOPs per image: SURVEY.md section 8d / BASELINE.md section 3.
    Counts the patch embedding on the kept patches and the pixel head on the masked rows only."""
    if cfg.labels > 0:   # classification step (BASELINE.json configs[4]): every patch, no decoder, LinearCLS head
        N, J, d, L, P = cfg.num_patches, cfg.num_cls_tokens, cfg.dim, cfg.layers, cfg.patch_dim
        T = N + J
        fwd = 2.0 * (N * P * d + L * (4 * T * d * d + 2 * T * T * d + 8 * N * d * d + 8 * J * J * d * d)
                     + J * d * cfg.labels)
        return fwd, 3.0 * fwd - 2.0 * N * P * d
    n, m, N = cfg.keep, cfg.num_masked, cfg.num_patches
    J, d, L, P = cfg.num_cls_tokens, cfg.dim, cfg.layers, cfg.patch_dim
    Te, Td, dd, dl = n + J, N + J, cfg.dec_dim, cfg.dec_layers
    fwd = 2.0 * (n * P * d + L * (4 * Te * d * d + 2 * Te * Te * d + 8 * n * d * d + 8 * J * J * d * d)
                 + Te * d * dd + dl * (12 * Td * dd * dd + 2 * Td * Td * dd) + m * dd * P)
    return fwd, 3.0 * fwd - 2.0 * n * P * d
