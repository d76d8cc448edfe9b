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


class _Bound:
    """What ``flax.linen.Module.bind(variables, rngs=...)`` gives the reference's modules: parameters (here the flat
    device ``ParamStore``) and RNG streams (here the uniform mask noise itself, f32 [num_patches], because the
    Flax/threefry stream cannot be reproduced outside JAX).  Engines (activation buffers) are cached per batch size."""

    def bind(self, store, rngs=None):
        import copy
        m = copy.copy(self)
        m._store, m._rngs, m._engines = store, dict(rngs or {}), {}
        return m

    def _need_store(self):
        if getattr(self, "_store", None) is None:
            raise _lib.JmError(f"{type(self).__name__} is unbound: call .bind(store, rngs={{'noise': ...}}) first")
        return self._store

    def _noise(self, device):
        nz = self._rngs.get("noise")
        if nz is None:
            raise _lib.JmError("rngs['noise'] (f32 [num_patches] uniform noise) is required for random masking")
        if not isinstance(nz, torch.Tensor):
            import numpy as np
            nz = torch.from_numpy(np.ascontiguousarray(nz, dtype=np.float32))
        return nz.to(device=device, dtype=torch.float32).contiguous()

    def _engine(self, batch: int):
        store = self._need_store()
        if batch not in self._engines:
            if store.cfg.labels > 0:
                from .finetune_engine import FinetuneEngine
                self._engines[batch] = FinetuneEngine(store.cfg, store, batch)
            else:
                from .engine import MAEEngine
                self._engines[batch] = MAEEngine(store.cfg, store, batch)
        return self._engines[batch]


@dataclass
class ViT(ViTBase, _Bound):
    """ViT with the Jumbo CLS token (src/modeling.py:221-274).

    ``ViT(...).bind(store, rngs={"noise": noise})(x, det=True)`` mirrors ``ViT.__call__(x, det=True)`` (:245):
    MAE mode (``labels`` None or <= 0) returns ``(x [B, J+keep, dim] f32, mask [B, L] f32, ids_restore [L] i32)``,
    classification mode returns the logits ``[B, labels]``.  ``x`` is the uint8 NCHW batch: the pixel normalisation
    of PretrainModule (src/pretraining.py:90-91) is fused into the patch loader, the f32 NHWC image never exists."""
    num_cls_tokens: int = 3

    def __post_init__(self):
        if self.dropout:
            raise NotImplementedError("dropout > 0 is off in every preset and not built")

    def _check(self, cfg):
        if cfg.dim != self.dim or cfg.layers != self.layers or cfg.heads != self.heads or \
                cfg.patch_size != self.patch_size or cfg.image_size != self.image_size:
            raise _lib.JmError("the bound parameter store was built for a different encoder")

    def __call__(self, x: torch.Tensor, det: bool = True):
        store = self._need_store()
        self._check(store.cfg)
        eng = self._engine(x.shape[0])
        x = x.to(store.device).contiguous()
        if store.cfg.labels > 0:                                    # classification branch (:269-274)
            labels = torch.zeros(x.shape[0], dtype=torch.int32, device=store.device)
            eng.forward(x, labels, want_grad=False)
            return eng.logits
        return self.encode(eng, x, self._noise(store.device))

    def encode(self, engine, images: torch.Tensor, noise: torch.Tensor):
        """MAE branch on a given engine: PatchEmbed + CLS + random_masking + JumboLayer x L + norm (:245-267)."""
        self._check(engine.cfg)
        engine.forward_encoder(images, noise)
        B = images.shape[0]
        x = torch.empty(B * engine.Te, self.dim, dtype=torch.float32, device=images.device)
        from . import ops
        ops.layernorm_fwd(engine.xa[self.layers], engine.ps.f("model/norm/scale"), engine.ps.f("model/norm/bias"),
                          1e-6, None, x, None, None)
        mask = engine.mask_row[None, :].expand(B, engine.mask_row.numel())
        return x.view(B, engine.Te, self.dim), mask, engine.ids_restore


@dataclass
class MAEDecoder(MAEDecoderBase, _Bound):
    """src/modeling.py:276-298.  ``MAEDecoder(...).bind(store)(x, det=True)``: ``x`` is the f32 ``[B, J + L, dec_dim]``
    decoder input (CLS tokens first, mask tokens already inserted and unshuffled, src/pretraining.py:95-106); adds the
    fixed sincos position embedding to the patch tokens, runs the ``dec_layers`` ViTLayers and ``dec_norm``."""
    num_cls_tokens: int = 3

    def __post_init__(self):
        if self.dec_dropout:
            raise NotImplementedError("decoder dropout > 0 is off in every preset and not built")

    def __call__(self, x: torch.Tensor, det: bool = True) -> torch.Tensor:
        from . import ops
        store = self._need_store()
        cfg = store.cfg
        if cfg.dec_dim != self.dec_dim or cfg.dec_layers != self.dec_layers or cfg.dec_heads != self.dec_heads:
            raise _lib.JmError("the bound parameter store was built for a different decoder")
        B, Td, dd = x.shape
        eng = self._engine(B)
        if Td != eng.Td or dd != cfg.dec_dim:
            raise _lib.JmError(f"decoder input must be [B, {eng.Td}, {cfg.dec_dim}], got {tuple(x.shape)}")
        J, N = cfg.num_cls_tokens, cfg.num_patches
        xin = x.to(device=store.device, dtype=torch.float32).contiguous().view(B * Td, dd)
        ids = torch.arange(N, dtype=torch.int32, device=store.device)
        # x[:, J:] += dec_wpe (:289-292): the unshuffle kernel with the identity permutation and no mask tokens
        ops.unshuffle_fwd(xin, ids, store.f("image_mask_embedding"), eng.dec_wpe, B, N, N, J, dd, eng.za[0])
        eng.forward_decoder_stack()
        out = torch.empty(B * Td, dd, dtype=torch.float32, device=store.device)
        ops.layernorm_fwd(eng.za[cfg.dec_layers], store.f("decoder_model/dec_norm/scale"),
                          store.f("decoder_model/dec_norm/bias"), 1e-6, None, out, None, None)
        return out.view(B, Td, dd)
