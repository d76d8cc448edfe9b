"""CPU oracle for the Jumbo-CLS MAE pre-training step.

THIS FILE IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only ``tests/``,
``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference``
legs of ``bench.py`` may import it.  The product path
(``jumbo_mae_tpu_b200/``) never does and fails loudly without its CUDA library.

PARITY UNPINNED.  The reference (antofuller/jumbo_mae_tpu) ships no tests,
golden vectors or fixtures, and cannot be imported here (jax / flax / optax /
chex are absent from this image and from the GPU box).  This module is a
line-by-line restatement in PyTorch-CPU (f32, optional f64) of the reference's
own source plus the *documented* behaviour of the third-party calls it makes
(flax.linen LayerNorm/gelu/softmax/Dense, jax.numpy argsort/var, optax adamw /
warmup_cosine_decay_schedule, all un-pinned: scripts/setup.sh:26-28).  Every
assumed third-party default has a named test in tests/test_oracle.py so that a
wrong assumption is a one-line fix once a JAX install exists.

The fine-tune (classification) step is restated the same way further down
(src/finetuning.py:84-124, src/utils.py:67-111,142-147: vit_classify, smooth_labels,
apply_mixup / apply_cutmix / random_bounding_box, softmax_cross_entropy, top_k_indices,
finetune_metrics, get_layer_index / layerwise_scale), with the RNG draws as arguments;
equally unpinned, with its own committed fixture tests/golden/tiny_classify.npz.

Every function cites the reference file:line it follows (paths are relative to
the reference checkout).  The backward pass comes from torch.autograd, the
analogue of ``jax.value_and_grad`` (src/pretraining.py:143).
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Any, Dict, Tuple

import numpy as np
import torch

Tree = Dict[str, Any]

# src/dataset.py:37-38 (float64 numpy constants, demoted to f32 by JAX when they
# meet the f32 image tensor in src/pretraining.py:91).
IMAGENET_DEFAULT_MEAN = np.array([0.485, 0.456, 0.406])
IMAGENET_DEFAULT_STD = np.array([0.229, 0.224, 0.225])


# ----------------------------------------------------------------------------
# configuration (src/modeling.py:35-104, src/main_pretrain.py:98-167)
# ----------------------------------------------------------------------------
@dataclass
class MAEConfig:
    # encoder: ViTBase fields, src/modeling.py:35-56
    layers: int = 12
    dim: int = 768
    heads: int = 12
    patch_size: int = 16
    image_size: int = 224
    image_mask_ratio: float = 0.75
    num_cls_tokens: int = 3          # src/modeling.py:171,222,277 (never overridden)
    # decoder: MAEDecoderBase fields, src/modeling.py:73-87; recipe values from
    # config/pretrain/pretrain-vit-b16-224-in1k-1600ep.sh:28-31
    dec_layers: int = 8
    dec_dim: int = 512
    dec_heads: int = 16
    norm_pix_loss: bool = False      # src/main_pretrain.py:139 (off in both presets)
    layerscale: bool = False         # scale1/2/3 per JumboLayer, init.constant(1e-4) (src/modeling.py:185-189)
    dec_layerscale: bool = False     # scale1/2 per decoder ViTLayer (src/modeling.py:159-162)
    posemb: str = "sincos2d"         # "learnable": embed/wpe is a parameter, truncated_normal(0.02) (src/modeling.py:115-118)
    labels: int = 0                  # > 0: classification mode of ViT (src/modeling.py:236-242): head, no masking, no decoder
    # optimizer: src/pretraining.py:223-259 with the preset's flags
    learning_rate: float = 1.5e-4
    train_batch_size: int = 4096
    weight_decay: float = 0.05
    lr_decay: float = 1.0            # < 1: layer-wise lr scale lr_decay ** (layers - label) behind adamw (src/pretraining.py:235-242)
    clip_grad: float = 0.0           # optax.clip_by_global_norm in front of adamw when > 0 (src/pretraining.py:243-244)
    adam_b1: float = 0.9
    adam_b2: float = 0.95
    adam_eps: float = 1e-8
    warmup_steps: int = 1281167 * 40 // 4096
    training_steps: int = 1281167 * 1600 // 4096

    @property
    def grid(self) -> int:
        return self.image_size // self.patch_size

    @property
    def num_patches(self) -> int:
        return self.grid * self.grid

    @property
    def keep(self) -> int:           # src/modeling.py:255
        return int(self.num_patches * (1.0 - self.image_mask_ratio))

    @property
    def num_masked(self) -> int:     # src/pretraining.py:101
        return int(self.num_patches * self.image_mask_ratio)

    @property
    def patch_dim(self) -> int:
        return self.patch_size * self.patch_size * 3


def vit_b16(**kw) -> MAEConfig:
    return MAEConfig(**kw)


def vit_l16(**kw) -> MAEConfig:
    kw.setdefault("training_steps", 1281167 * 800 // 4096)
    return MAEConfig(layers=24, dim=1024, heads=16, **kw)


def vit_h14(**kw) -> MAEConfig:
    return MAEConfig(layers=32, dim=1280, heads=16, patch_size=14, **kw)


def tiny(**kw) -> MAEConfig:
    """A small shape for fast CPU tests (not a reference preset)."""
    base = dict(layers=2, dim=64, heads=2, patch_size=8, image_size=32, dec_layers=2,
                dec_dim=64, dec_heads=2, warmup_steps=10, training_steps=100,
                train_batch_size=256)
    base.update(kw)
    return MAEConfig(**base)


def golden_tiny(**kw) -> MAEConfig:
    """The shape the committed fixture tests/golden/tiny_step.npz was generated on."""
    base = dict(layers=2, dim=32, heads=1, patch_size=8, image_size=32, dec_layers=1,
                dec_dim=32, dec_heads=1, warmup_steps=10, training_steps=100,
                train_batch_size=256)
    base.update(kw)
    return MAEConfig(**base)


# ----------------------------------------------------------------------------
# src/utils.py:114-121
# ----------------------------------------------------------------------------
def fixed_sincos2d_embeddings(ncols: int, nrows: int, dim: int, dtype=torch.float32) -> torch.Tensor:
    # jnp.linspace(0, 1, dim // 4) includes the endpoint.
    freqs = 1.0 / (10000.0 ** torch.linspace(0, 1, dim // 4, dtype=dtype))
    x = torch.outer(torch.arange(0, nrows, dtype=dtype), freqs)
    y = torch.outer(torch.arange(0, ncols, dtype=dtype), freqs)
    x = x[None, :, :].expand(ncols, nrows, dim // 4)
    y = y[:, None, :].expand(ncols, nrows, dim // 4)
    return torch.cat((torch.sin(x), torch.cos(x), torch.sin(y), torch.cos(y)), dim=2)


# ----------------------------------------------------------------------------
# src/utils_mae.py
# ----------------------------------------------------------------------------
def index_sequence(x: torch.Tensor, ids: torch.Tensor) -> torch.Tensor:
    """src/utils_mae.py:84-85.  ``ids`` [B, k] (per-image masking, an extension: the reference passes one [k] row)
    gathers every image with its own indices."""
    if ids.dim() == 2:
        idx = ids.long()[:, :, None].expand(-1, -1, x.shape[-1])
        return torch.gather(x, 1, idx)
    return x[:, ids.long(), ...]


def masking_from_noise(noise: np.ndarray, keep_len: int) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    """Integer half of random_masking (src/utils_mae.py:90-96) given the noise.

    ``jnp.argsort`` is a stable ascending sort; indices are int32.  Returns
    (ids_shuffle, ids_restore, mask_row) where mask_row[i] = 1.0 iff patch i is
    masked (one row: the reference shares one permutation across the batch).
    """
    noise = np.asarray(noise, dtype=np.float32)
    ids_shuffle = np.argsort(noise, kind="stable").astype(np.int32)
    ids_restore = np.argsort(ids_shuffle, kind="stable").astype(np.int32)
    row = np.ones(noise.shape[0], dtype=np.float32)
    row[:keep_len] = 0.0
    return ids_shuffle, ids_restore, row[ids_restore]


def random_masking(x: torch.Tensor, noise, keep_len: int):
    """src/utils_mae.py:88-99 with the uniform noise supplied as data.

    The reference draws ``noise`` of shape (length,) from the Flax "noise" RNG
    stream (threefry + module-path fold-in), which cannot be reproduced without
    the libraries; the drop-in boundary therefore takes the noise as an input.
    """
    batch, length, _ = x.shape
    noise = np.asarray(noise)
    if noise.ndim == 2:              # per-image masking (extension): one permutation per image
        rows = [masking_from_noise(noise[b], keep_len) for b in range(batch)]
        ids_shuffle = torch.from_numpy(np.stack([r[0] for r in rows]))
        ids_restore = torch.from_numpy(np.stack([r[1] for r in rows]))
        mask = torch.from_numpy(np.stack([r[2] for r in rows])).to(x.dtype)
        return index_sequence(x, ids_shuffle[:, :keep_len]), mask, ids_restore
    ids_shuffle, ids_restore, mask_row = masking_from_noise(noise, keep_len)
    kept = index_sequence(x, torch.from_numpy(ids_shuffle[:keep_len].copy()))
    mask = torch.from_numpy(mask_row)[None, :].expand(batch, length).to(x.dtype)
    return kept, mask, torch.from_numpy(ids_restore)


def extract_patches(inputs: torch.Tensor, patch_size: int) -> torch.Tensor:
    """src/utils_mae.py:67-73 (NHWC in, [B, L, p*p*C] out, (ph, pw, c) order)."""
    batch, height, width, channels = inputs.shape
    height, width = height // patch_size, width // patch_size
    x = inputs.reshape(batch, height, patch_size, width, patch_size, channels)
    x = x.transpose(2, 3)
    return x.reshape(batch, height * width, patch_size ** 2 * channels)


def merge_patches(inputs: torch.Tensor, patch_size: int) -> torch.Tensor:
    """src/utils_mae.py:76-82 (only used by tests: inverse of extract_patches)."""
    batch, length, _ = inputs.shape
    height = width = int(length ** 0.5)
    x = inputs.reshape(batch, height, width, patch_size, patch_size, -1)
    x = x.transpose(2, 3)
    return x.reshape(batch, height * patch_size, width * patch_size, -1)


def patch_mse_loss(patch_output, patch_target, valid=None) -> torch.Tensor:
    """src/utils_mae.py:51-64."""
    if valid is None:
        valid = torch.ones(patch_target.shape[:2], dtype=patch_target.dtype)
    valid_ratio = valid.sum(dim=-1) / valid.shape[-1]
    per_patch = ((patch_target - patch_output) ** 2).mean(dim=-1)
    per_patch = torch.where(valid > 0.0, per_patch, torch.zeros((), dtype=per_patch.dtype))
    return (per_patch.mean(dim=-1) / valid_ratio).mean()


# ----------------------------------------------------------------------------
# third-party primitives (flax.linen), documented defaults
# ----------------------------------------------------------------------------
def layer_norm(x, scale, bias, eps: float = 1e-6):
    """flax.linen.LayerNorm defaults: eps=1e-6, biased "fast" variance
    E[x^2]-E[x]^2 clipped at 0, statistics in f32 (call sites
    src/modeling.py:155-156,177-179,238,282)."""
    mean = x.mean(dim=-1, keepdim=True)
    mean2 = (x * x).mean(dim=-1, keepdim=True)
    var = torch.clamp(mean2 - mean * mean, min=0.0)
    return (x - mean) * torch.rsqrt(var + eps) * scale + bias


def gelu_tanh(x):
    """flax.linen.gelu default approximate=True (src/modeling.py:148)."""
    c = math.sqrt(2.0 / math.pi)
    return 0.5 * x * (1.0 + torch.tanh(c * (x + 0.044715 * x * x * x)))


def dense(x, p: Tree):
    """flax.linen.Dense: x @ kernel + bias, kernel [in, out]."""
    return x @ p["kernel"] + p["bias"]


# ----------------------------------------------------------------------------
# src/modeling.py modules
# ----------------------------------------------------------------------------
def attention(x, p: Tree, heads: int):
    """Attention, src/modeling.py:127-138.  wq/wk/wv kernels [d, h, hd], biases
    [h, hd]; wo kernel [h, hd, d], bias [d].  q is scaled AFTER the bias add."""
    hd = x.shape[-1] // heads
    q = torch.einsum("btd,dhe->bthe", x, p["wq"]["kernel"]) + p["wq"]["bias"]
    k = torch.einsum("btd,dhe->bthe", x, p["wk"]["kernel"]) + p["wk"]["bias"]
    v = torch.einsum("btd,dhe->bthe", x, p["wv"]["kernel"]) + p["wv"]["bias"]
    z = torch.einsum("bqhd,bkhd->bhqk", q / hd ** 0.5, k)
    z = torch.einsum("bhqk,bkhd->bqhd", torch.softmax(z, dim=-1), v)
    return torch.einsum("bqhe,hed->bqd", z, p["wo"]["kernel"]) + p["wo"]["bias"]


def feed_forward(x, p: Tree):
    """FeedForward, src/modeling.py:141-148 (dropout identity at rate 0)."""
    return dense(gelu_tanh(dense(x, p["w1"])), p["w2"])


def drop_path(y, scale):
    """nn.Dropout(rate, broadcast_dims=(1, 2)) / (1,) on a residual branch (src/modeling.py:157,181-183): one draw per
    image, kept branches are divided by (1 - rate).  ``scale`` is that per-image factor, keep / (1 - rate), f32 [B]
    (the Bernoulli draws come from the Flax "dropout" stream and are data at this boundary); None = deterministic."""
    if scale is None:
        return y
    return y * torch.as_tensor(scale, dtype=y.dtype).reshape((-1,) + (1,) * (y.dim() - 1))


def vit_layer(x, p: Tree, heads: int, drop=None):
    """ViTLayer (decoder block), src/modeling.py:150-167; scale1=scale2=1.0.  ``drop`` = (scale1, scale2) per-image
    DropPath factors of the two branches (the SAME nn.Dropout module is applied twice, with two draws), or None."""
    d1, d2 = drop if drop is not None else (None, None)
    s1, s2 = p.get("scale1", 1.0), p.get("scale2", 1.0)        # LayerScale (:159-162,165-166)
    x = x + drop_path(s1 * attention(layer_norm(x, p["norm1"]["scale"], p["norm1"]["bias"]), p["attn"], heads), d1)
    x = x + drop_path(s2 * feed_forward(layer_norm(x, p["norm2"]["scale"], p["norm2"]["bias"]), p["ff"]), d2)
    return x


def jumbo_layer(x, p: Tree, jumbo_mlp: Tree, heads: int, J: int, drop=None):
    """JumboLayer, src/modeling.py:191-206.  Note (:197-199) that the CLS
    residual is taken from the post-norm3 value.  ``drop`` = per-image DropPath factors (drop_path1, drop_path2,
    drop_path3), each f32 [B], or None."""
    d1, d2, d3 = drop if drop is not None else (None, None, None)
    s1, s2, s3 = p.get("scale1", 1.0), p.get("scale2", 1.0), p.get("scale3", 1.0)     # LayerScale (:185-189)
    x = x + drop_path(s1 * attention(layer_norm(x, p["norm1"]["scale"], p["norm1"]["bias"]), p["attn"], heads), d1)
    cls_tokens, patch_tokens = x[:, :J, :], x[:, J:, :]
    bs = cls_tokens.shape[0]
    c = cls_tokens.reshape(bs, -1)
    c = layer_norm(c, p["norm3"]["scale"], p["norm3"]["bias"])
    c = c + drop_path(s3 * feed_forward(c, jumbo_mlp), d3)
    patch_tokens = patch_tokens + drop_path(s2 * feed_forward(
        layer_norm(patch_tokens, p["norm2"]["scale"], p["norm2"]["bias"]), p["ff"]), d2)
    return torch.cat([c.reshape(bs, J, -1), patch_tokens], dim=1)


def patch_embed(images_nhwc, p: Tree, cfg: MAEConfig):
    """PatchEmbed, src/modeling.py:106-124: Conv(dim, p x p, stride p, VALID) with an
    HWIO kernel [p, p, 3, dim] + fixed sincos wpe (posemb == "sincos2d")."""
    patches = extract_patches(images_nhwc, cfg.patch_size)           # (ph, pw, c) order == HWI flatten
    w = p["wte"]["kernel"].reshape(-1, cfg.dim)
    x = patches @ w + p["wte"]["bias"]
    if cfg.posemb == "learnable":
        wpe = p["wpe"]                                                 # [grid, grid, dim] parameter (:115-118)
    else:
        wpe = fixed_sincos2d_embeddings(cfg.grid, cfg.grid, cfg.dim, dtype=x.dtype)
    return x + wpe.reshape(1, -1, cfg.dim)


def vit_forward(images_nhwc, p: Tree, cfg: MAEConfig, noise, collect=None):
    """ViT.__call__ MAE branch, src/modeling.py:245-267."""
    J = cfg.num_cls_tokens
    x = patch_embed(images_nhwc, p["embed"], cfg)
    bs = x.shape[0]
    cls_tokens = p["cls_tokens"].expand(bs, J, cfg.dim)              # jnp.tile (:249)
    patch_tokens, mask, ids_restore = random_masking(x, noise, cfg.keep)
    x = torch.cat((cls_tokens, patch_tokens), dim=1)
    if collect is not None:
        collect["enc_in"] = x
    for i in range(cfg.layers):
        x = jumbo_layer(x, p[f"layer_{i}"], p["jumbo_mlp"], cfg.heads, J)
        if collect is not None:
            collect[f"enc_layer_{i}"] = x
    x = layer_norm(x, p["norm"]["scale"], p["norm"]["bias"])
    return x, mask, ids_restore


def vit_classify(images_nhwc, p: Tree, cfg: MAEConfig, collect=None, drop=None):
    """ViT.__call__ with a head (labels > 0), src/modeling.py:245-274: no masking, every patch, final LayerNorm,
    LinearCLS (no BatchNorm: batch_norm is a linear-probing option) on the concatenated CLS tokens -> logits."""
    J = cfg.num_cls_tokens
    x = patch_embed(images_nhwc, p["embed"], cfg)
    bs = x.shape[0]
    x = torch.cat((p["cls_tokens"].expand(bs, J, cfg.dim), x), dim=1)
    for i in range(cfg.layers):                                       # drop: [layers, 3, B] DropPath factors or None
        x = jumbo_layer(x, p[f"layer_{i}"], p["jumbo_mlp"], cfg.heads, J, None if drop is None else drop[i])
    x = layer_norm(x, p["norm"]["scale"], p["norm"]["bias"])
    if collect is not None:
        collect["enc_out"] = x
    x = x[:, :J, :].reshape(bs, -1)
    return dense(x, p["head"]["Dense_0"])


def finetune_loss_and_grads(params: Tree, images_nhwc, labels_soft, cfg: MAEConfig, drop=None):
    """loss_fn of src/finetuning.py:109-124 on already mixed images / prepared labels: mean softmax cross-entropy of
    FinetuneModule (:98) and its gradient (jax.value_and_grad -> torch.autograd).  Returns (loss, logits, grads)."""
    leaves = tree_map(lambda t: t.detach().clone().requires_grad_(True), params)
    logits = vit_classify(torch.as_tensor(images_nhwc), leaves["model"], cfg, drop=drop)
    lab = torch.as_tensor(labels_soft, dtype=logits.dtype)
    loss = (-(lab * torch.log_softmax(logits, dim=-1)).sum(-1)).mean()
    loss.backward()
    return loss.detach(), logits.detach(), tree_map(lambda t: t.grad if t.grad is not None else torch.zeros_like(t), leaves)


def decoder_forward(x, p: Tree, cfg: MAEConfig, collect=None):
    """MAEDecoder.__call__, src/modeling.py:286-298 (always sincos dec_wpe)."""
    J = cfg.num_cls_tokens
    cls_tokens, patch_tokens = x[:, :J, :], x[:, J:, :]
    dec_wpe = fixed_sincos2d_embeddings(cfg.grid, cfg.grid, cfg.dec_dim, dtype=x.dtype)
    patch_tokens = patch_tokens + dec_wpe.reshape(1, -1, patch_tokens.shape[-1])
    x = torch.cat((cls_tokens, patch_tokens), dim=1)
    if collect is not None:
        collect["dec_in"] = x
    for i in range(cfg.dec_layers):
        x = vit_layer(x, p[f"dec_layer_{i}"], cfg.dec_heads)
        if collect is not None:
            collect[f"dec_layer_{i}"] = x
    return layer_norm(x, p["dec_norm"]["scale"], p["dec_norm"]["bias"])


def preprocess(images_u8: torch.Tensor, dtype=torch.float32) -> torch.Tensor:
    """src/pretraining.py:90-91: uint8 NCHW -> normalised NHWC."""
    x = images_u8.permute(0, 2, 3, 1).to(dtype) / 0xFF
    mean = torch.tensor(IMAGENET_DEFAULT_MEAN, dtype=dtype)
    std = torch.tensor(IMAGENET_DEFAULT_STD, dtype=dtype)
    return (x - mean) / std


def pretrain_forward(params: Tree, images_u8, noise, cfg: MAEConfig, collect=None):
    """PretrainModule.__call__, src/pretraining.py:87-122 -> {"loss": scalar}."""
    dtype = params["image_mask_embedding"].dtype
    J = 3                                                              # hard-coded at :96,:110
    images = preprocess(images_u8, dtype)
    x, image_mask, ids_restore = vit_forward(images, params["model"], cfg, noise, collect)
    if collect is not None:
        collect["enc_out"] = x
    x = dense(x, params["decoder_proj"])
    encoder_cls, image_x = x[:, :J, :], x[:, J:, :]
    bs = image_x.shape[0]
    masked = params["image_mask_embedding"].expand(
        bs, int(ids_restore.shape[-1] * cfg.image_mask_ratio), cfg.dec_dim)
    image_x = index_sequence(torch.cat([image_x, masked], dim=1), ids_restore)
    x = torch.cat([encoder_cls, image_x], dim=1)
    x = decoder_forward(x, params["decoder_model"], cfg, collect)
    if collect is not None:
        collect["dec_out"] = x
    image_output = dense(x[:, J:, :], params["decoder_image_output"])
    image_patches = extract_patches(images, cfg.patch_size)
    if cfg.norm_pix_loss:
        mean = image_patches.mean(dim=-1, keepdim=True)
        var = image_patches.var(dim=-1, keepdim=True, unbiased=False)   # jnp.var ddof=0
        image_patches = (image_patches - mean) / torch.sqrt(var + 1e-6)
    if collect is not None:
        collect["image_output"] = image_output
        collect["mask"] = image_mask
        collect["ids_restore"] = ids_restore
    return {"loss": patch_mse_loss(image_output, image_patches, image_mask)}


# ----------------------------------------------------------------------------
# parameters (Flax naming) and init (src/modeling.py:30-32,243;
# src/pretraining.py:38,83)
# ----------------------------------------------------------------------------
def _tn(gen, shape, dtype):
    """jax.nn.initializers.truncated_normal(0.02): N(0,1) truncated to [-2,2], x 0.02."""
    t = torch.empty(shape, dtype=torch.float32)
    torch.nn.init.trunc_normal_(t, mean=0.0, std=1.0, a=-2.0, b=2.0, generator=gen)
    return (t * 0.02).to(dtype)


def _dense_p(gen, din, dout, dtype):
    return {"kernel": _tn(gen, (din, dout), dtype), "bias": torch.zeros(dout, dtype=dtype)}


def _ln_p(d, dtype):
    return {"scale": torch.ones(d, dtype=dtype), "bias": torch.zeros(d, dtype=dtype)}


def _attn_p(gen, d, h, dtype):
    hd = d // h
    out = {}
    for n in ("wq", "wk", "wv"):
        out[n] = {"kernel": _tn(gen, (d, h, hd), dtype), "bias": torch.zeros(h, hd, dtype=dtype)}
    out["wo"] = {"kernel": _tn(gen, (h, hd, d), dtype), "bias": torch.zeros(d, dtype=dtype)}
    return out


def _ff_p(gen, d, dtype):
    return {"w1": _dense_p(gen, d, 4 * d, dtype), "w2": _dense_p(gen, 4 * d, d, dtype)}


def init_params(cfg: MAEConfig, seed: int = 0, dtype=torch.float32, randomize_all: bool = False) -> Tree:
    """Random-init parameter tree with the reference's init rules.

    ``randomize_all`` additionally perturbs the zero/one-initialised leaves
    (biases, LayerNorm, cls_tokens) so that parity tests exercise every term."""
    gen = torch.Generator().manual_seed(seed)
    d, J = cfg.dim, cfg.num_cls_tokens
    model: Tree = {
        "embed": {"wte": {"kernel": _tn(gen, (cfg.patch_size, cfg.patch_size, 3, d), dtype),
                          "bias": torch.zeros(d, dtype=dtype)}},
        "cls_tokens": torch.zeros(1, J, d, dtype=dtype),
        "jumbo_mlp": _ff_p(gen, J * d, dtype),
    }
    for i in range(cfg.layers):
        model[f"layer_{i}"] = {"attn": _attn_p(gen, d, cfg.heads, dtype), "ff": _ff_p(gen, d, dtype),
                               "norm1": _ln_p(d, dtype), "norm2": _ln_p(d, dtype),
                               "norm3": _ln_p(J * d, dtype)}
        if cfg.layerscale:
            model[f"layer_{i}"].update({"scale1": torch.full((d,), 1e-4, dtype=dtype),
                                        "scale2": torch.full((d,), 1e-4, dtype=dtype),
                                        "scale3": torch.full((J * d,), 1e-4, dtype=dtype)})
    model["norm"] = _ln_p(d, dtype)
    if cfg.posemb == "learnable":    # drawn last, so that the sincos configurations keep their parameter values
        model["embed"]["wpe"] = _tn(torch.Generator().manual_seed(seed + 7919), (cfg.grid, cfg.grid, d), dtype)
    if cfg.labels > 0:               # LinearCLS (src/modeling.py:209-219): nn.compact around one Dense -> "Dense_0"
        model["head"] = {"Dense_0": _dense_p(gen, J * d, cfg.labels, dtype)}
        params = {"model": model}
        if randomize_all:
            for path, leaf in tree_leaves(params):
                if path[-1] in ("bias", "cls_tokens"):
                    leaf.add_(_tn(gen, tuple(leaf.shape), dtype))
                elif path[-1] == "scale":
                    leaf.add_(_tn(gen, tuple(leaf.shape), dtype) * 5.0)
                elif path[-1] in ("scale1", "scale2", "scale3"):
                    leaf.copy_(0.5 + _tn(gen, tuple(leaf.shape), dtype) * 10.0)
        return params
    dd = cfg.dec_dim
    dec: Tree = {}
    for i in range(cfg.dec_layers):
        dec[f"dec_layer_{i}"] = {"attn": _attn_p(gen, dd, cfg.dec_heads, dtype), "ff": _ff_p(gen, dd, dtype),
                                 "norm1": _ln_p(dd, dtype), "norm2": _ln_p(dd, dtype)}
        if cfg.dec_layerscale:
            dec[f"dec_layer_{i}"].update({"scale1": torch.full((dd,), 1e-4, dtype=dtype),
                                          "scale2": torch.full((dd,), 1e-4, dtype=dtype)})
    dec["dec_norm"] = _ln_p(dd, dtype)
    params: Tree = {
        "model": model,
        "decoder_model": dec,
        "image_mask_embedding": _tn(gen, (1, 1, dd), dtype),
        "decoder_proj": _dense_p(gen, d, dd, dtype),
        "decoder_image_output": _dense_p(gen, dd, cfg.patch_dim, dtype),
    }
    if randomize_all:
        for path, leaf in tree_leaves(params):
            if path[-1] in ("bias", "cls_tokens"):
                leaf.add_(_tn(gen, tuple(leaf.shape), dtype))
            elif path[-1] == "scale":
                leaf.add_(_tn(gen, tuple(leaf.shape), dtype) * 5.0)
            elif path[-1] in ("scale1", "scale2", "scale3"):
                leaf.copy_(0.5 + _tn(gen, tuple(leaf.shape), dtype) * 10.0)      # ~0.5 +- 0.2, away from 0
    return params


def tree_leaves(tree: Tree, prefix=()):
    """(path tuple, tensor) pairs in sorted-key order (Flax/JAX dict order)."""
    out = []
    for k in sorted(tree.keys()):
        v = tree[k]
        if isinstance(v, dict):
            out.extend(tree_leaves(v, prefix + (k,)))
        else:
            out.append((prefix + (k,), v))
    return out


def tree_map(fn, tree: Tree, *rest) -> Tree:
    out = {}
    for k, v in tree.items():
        if isinstance(v, dict):
            out[k] = tree_map(fn, v, *[r[k] for r in rest])
        else:
            out[k] = fn(v, *[r[k] for r in rest])
    return out


def count_params(tree: Tree) -> int:
    return sum(int(v.numel()) for _, v in tree_leaves(tree))


# ----------------------------------------------------------------------------
# optimizer: optax.adamw + warmup_cosine_decay_schedule (src/pretraining.py:223-259)
# ----------------------------------------------------------------------------
def lr_schedule(count: int, cfg: MAEConfig) -> float:
    """optax.warmup_cosine_decay_schedule(init_value=1e-6, peak, warmup_steps,
    decay_steps=training_steps, end_value=1e-5) evaluated at ``count``
    (src/pretraining.py:250-259).  peak = lr * batch / 256 (:250)."""
    init_value, end_value = 1e-6, 1e-5
    peak = cfg.learning_rate * cfg.train_batch_size / 256
    warm = cfg.warmup_steps
    if count < warm:
        # linear_schedule(init, peak, warm)
        frac = 1.0 - min(count, warm) / warm
        return (init_value - peak) * frac + peak
    # cosine_decay_schedule(peak, decay_steps - warmup, alpha = end / peak)
    decay_steps = cfg.training_steps - warm
    c = min(count - warm, decay_steps)
    cosine = 0.5 * (1.0 + math.cos(math.pi * c / decay_steps))
    alpha = end_value / peak
    return peak * ((1.0 - alpha) * cosine + alpha)


def get_layer_index(path: Tuple[str, ...], num_layers: int) -> int:
    """get_layer_index_fn, src/utils.py:142-147: ``model/layer_k`` -> k + 1, ``model/embed`` -> 0, else num_layers."""
    if path[0] == "model" and path[1].startswith("layer_"):
        return int(path[1][len("layer_"):]) + 1
    if path[0] == "model" and path[1] == "embed":
        return 0
    return num_layers


def layerwise_scale(path: Tuple[str, ...], cfg: MAEConfig) -> float:
    """optax.multi_transform(layerwise_scales, label_fn) chained behind adamw (src/pretraining.py:235-242): the final
    update of a leaf is multiplied by ``lr_decay ** (layers - label)``; 1.0 when --lr-decay is 1."""
    if cfg.lr_decay >= 1.0:
        return 1.0
    return cfg.lr_decay ** (cfg.layers - get_layer_index(path, cfg.layers))


def adamw_leaf(p, g, m, v, count: int, lr: float, cfg: MAEConfig, decay: bool):
    """One optax.adamw leaf update, ``count`` = number of updates already applied.

    scale_by_adam: m,v EMA; bias correction with t = count + 1;
    u = m_hat / (sqrt(v_hat) + eps); add_decayed_weights (kernel leaves only,
    src/pretraining.py:233); scale by -lr(count); apply_updates."""
    b1, b2 = cfg.adam_b1, cfg.adam_b2
    m = b1 * m + (1.0 - b1) * g
    v = b2 * v + (1.0 - b2) * g * g
    t = count + 1
    m_hat = m / (1.0 - b1 ** t)
    v_hat = v / (1.0 - b2 ** t)
    u = m_hat / (torch.sqrt(v_hat) + cfg.adam_eps)
    if decay:
        u = u + cfg.weight_decay * p
    return p - lr * u, m, v


def lamb_leaf(p, g, m, v, count: int, lr: float, cfg: MAEConfig, decay: bool):
    """One leaf update of the reference's ``modified_lamb`` (src/utils.py:124-139): optax.chain(scale_by_adam,
    add_decayed_weights(mask), masked(scale_by_trust_ratio(), mask), scale_by_learning_rate), mask = kernel leaves.
    scale_by_trust_ratio (optax, defaults min_norm=0, trust_coefficient=1, eps=0, restated from its published
    definition): update * ||param|| / ||update||, and * 1 where either norm is 0."""
    b1, b2 = cfg.adam_b1, cfg.adam_b2
    m = b1 * m + (1.0 - b1) * g
    v = b2 * v + (1.0 - b2) * g * g
    t = count + 1
    u = (m / (1.0 - b1 ** t)) / (torch.sqrt(v / (1.0 - b2 ** t)) + cfg.adam_eps)
    if decay:
        u = u + cfg.weight_decay * p
        pn, un = torch.linalg.vector_norm(p), torch.linalg.vector_norm(u)
        ratio = torch.where((pn == 0) | (un == 0), torch.ones_like(pn), pn / un)
        u = u * ratio
    return p - lr * u, m, v


@dataclass
class OracleState:
    """The numeric part of TrainState (src/pretraining.py:50-57)."""
    params: Tree
    mu: Tree
    nu: Tree
    step: int = 0
    cfg: MAEConfig = field(default_factory=MAEConfig)


def create_state(cfg: MAEConfig, seed: int = 0, dtype=torch.float32, randomize_all=False) -> OracleState:
    params = init_params(cfg, seed, dtype, randomize_all)
    zeros = lambda t: torch.zeros_like(t)
    return OracleState(params, tree_map(zeros, params), tree_map(zeros, params), 0, cfg)


def loss_and_grads(params: Tree, images_u8, noise, cfg: MAEConfig, collect=None):
    """jax.value_and_grad(loss_fn) (src/pretraining.py:127-130,143)."""
    leaves = tree_map(lambda t: t.detach().clone().requires_grad_(True), params)
    loss = pretrain_forward(leaves, images_u8, noise, cfg, collect)["loss"]
    flat = [t for _, t in tree_leaves(leaves)]
    grads_flat = torch.autograd.grad(loss, flat, allow_unused=True)
    it = iter(grads_flat)
    grads = {}

    def fill(tree, out):
        for k in sorted(tree.keys()):
            if isinstance(tree[k], dict):
                out[k] = {}
                fill(tree[k], out[k])
            else:
                g = next(it)
                out[k] = torch.zeros_like(tree[k]) if g is None else g.detach()
    fill(leaves, grads)
    return loss.detach(), grads


def training_step(state: OracleState, images_u8, noise, grads_override: Tree | None = None):
    """training_step without grad accumulation (src/pretraining.py:125-159) on one
    device: loss/grad, adamw, returns (new_state, {"loss", "learning_rate"}).
    ``grads_override`` stands in for ``lax.pmean(grads)`` in data-parallel tests."""
    cfg = state.cfg
    loss, grads = loss_and_grads(state.params, images_u8, noise, cfg)
    if grads_override is not None:
        grads = grads_override
    lr = lr_schedule(state.step, cfg)
    if cfg.clip_grad > 0:
        # optax.clip_by_global_norm: g unchanged if ||g|| < max_norm, else g / ||g|| * max_norm (all leaves together)
        g_norm = torch.sqrt(sum((g.double() ** 2).sum() for _, g in tree_leaves(grads))).float()
        if not bool(g_norm < cfg.clip_grad):
            grads = tree_map(lambda g: (g / g_norm) * cfg.clip_grad, grads)
    new_p, new_m, new_v = {}, {}, {}

    def upd(p, g, m, v, path):
        return adamw_leaf(p, g, m, v, state.step, lr * layerwise_scale(path, cfg), cfg, decay=(path[-1] == "kernel"))

    def walk(pt, gt, mt, vt, np_, nm, nv, prefix=()):
        for k in pt.keys():
            if isinstance(pt[k], dict):
                np_[k], nm[k], nv[k] = {}, {}, {}
                walk(pt[k], gt[k], mt[k], vt[k], np_[k], nm[k], nv[k], prefix + (k,))
            else:
                np_[k], nm[k], nv[k] = upd(pt[k], gt[k], mt[k], vt[k], prefix + (k,))
    walk(state.params, grads, state.mu, state.nu, new_p, new_m, new_v)
    new_state = OracleState(new_p, new_m, new_v, state.step + 1, cfg)
    return new_state, {"loss": float(loss), "learning_rate": lr}, grads

# ----------------------------------------------------------------------------
# fine-tune step, classification loss (src/finetuning.py:84-106; SURVEY.md 8(f) rank 1, first piece)
# ----------------------------------------------------------------------------
def smooth_labels(labels: np.ndarray, alpha: float) -> np.ndarray:
    """optax.smooth_labels (src/finetuning.py:94): ``(1.0 - alpha) * labels + alpha / num_categories`` with
    Python-float constants meeting an f32 array (third-party, restated from its published definition)."""
    labels = np.asarray(labels, dtype=np.float32)
    return np.float32(1.0 - alpha) * labels + np.float32(alpha / labels.shape[-1])


def mix_labels(labels: np.ndarray, perm: np.ndarray, ratio) -> np.ndarray:
    """Label side of Mixup.apply_mixup / apply_cutmix (src/utils.py:70-71,82-83): ``ratio`` is the beta sample
    (mixup) or the mean of the box mask (cutmix), an f32 scalar either way."""
    r = np.float32(ratio)
    return r * labels + (np.float32(1) - r) * labels[perm]


def apply_mixup(images: np.ndarray, labels: np.ndarray, ratio, perm) -> Tuple[np.ndarray, np.ndarray]:
    """Mixup.apply_mixup (src/utils.py:67-72) on normalised NHWC f32 images with the beta sample ``ratio`` and the
    batch permutation given (the reference draws both from the ``mixup`` RNG stream: RNG parity is not attempted)."""
    r, perm = np.float32(ratio), np.asarray(perm)
    images = r * images + (np.float32(1) - r) * images[perm]
    return images, mix_labels(labels, perm, r)


def random_bounding_box(ratio, xstart, ystart, width: int, height: int):
    """Mixup.random_bounding_box (src/utils.py:85-92) with the two uniform draws given: column / row membership of
    the box and the image mask (True OUTSIDE the box) of shape [1, height, width, 1]."""
    size = np.float32((np.float32(1) - np.float32(ratio)) ** np.float32(0.5))
    xs, ys, half = np.float32(xstart), np.float32(ystart), np.float32(0.5)
    xr, yr = np.linspace(0, 1, width, dtype=np.float32), np.linspace(0, 1, height, dtype=np.float32)
    xmask = (xs - half * size <= xr) & (xr < xs + half * size)
    ymask = (ys - half * size <= yr) & (yr < ys + half * size)
    return xmask, ymask, ~(xmask[None, None, :, None] & ymask[None, :, None, None])


def apply_cutmix(images: np.ndarray, labels: np.ndarray, ratio, xstart, ystart, perm):
    """Mixup.apply_cutmix (src/utils.py:74-83): the box of images[perm] pasted in, labels mixed by the box area."""
    perm = np.asarray(perm)
    _, _, image_mask = random_bounding_box(ratio, xstart, ystart, images.shape[2], images.shape[1])
    label_mask = image_mask.mean((1, 2), dtype=np.float32)          # [1, 1]
    m = image_mask.astype(np.float32)
    images = m * images + (np.float32(1) - m) * images[perm]
    labels = label_mask * labels + (np.float32(1) - label_mask) * labels[perm]
    return images, labels


def softmax_cross_entropy(logits: np.ndarray, labels: np.ndarray) -> np.ndarray:
    """optax.softmax_cross_entropy (src/finetuning.py:40,98): ``-sum(labels * log_softmax(logits), -1)`` with
    jax.nn.log_softmax = shifted - log(sum(exp(shifted))), shifted = logits - max(logits)."""
    logits = np.asarray(logits, dtype=np.float32)
    shifted = logits - logits.max(-1, keepdims=True)
    logp = shifted - np.log(np.exp(shifted).sum(-1, keepdims=True, dtype=np.float32))
    return -(labels * logp).sum(-1, dtype=np.float32)


def top_k_indices(logits: np.ndarray, k: int) -> np.ndarray:
    """jax.lax.top_k(...)[1]: descending values, the lower index first among equal values."""
    return np.argsort(-np.asarray(logits, dtype=np.float32), axis=-1, kind="stable")[..., :k]


def finetune_metrics(logits, labels, num_labels: int, label_smoothing: float = 0.0, perm=None, ratio=1.0,
                     det: bool = False):
    """FinetuneModule.__call__ after the model (src/finetuning.py:90-106): per-row loss, acc1, acc5 and the label
    matrix the loss saw.  ``labels``: int [B] class indices or f32 [B, C]."""
    labels = np.asarray(labels)
    if labels.ndim == 1:                                          # nn.one_hot (:90)
        labels = np.eye(num_labels, dtype=np.float32)[labels]
    labels = labels.astype(np.float32)
    if not det:                                                   # :93-95
        labels = smooth_labels(labels, label_smoothing)
        if perm is not None:
            labels = mix_labels(labels, np.asarray(perm), ratio)
    loss = softmax_cross_entropy(logits, labels)                  # :98
    hot = labels == labels.max(-1, keepdims=True)                 # :99
    preds = top_k_indices(logits, min(5, labels.shape[-1]))       # :105
    accs = np.take_along_axis(hot, preds, axis=-1)                # :106
    return {"loss": loss, "acc1": accs[:, 0].astype(np.float32), "acc5": accs.any(-1).astype(np.float32),
            "labels": labels}


def softmax_ce_grad(logits: np.ndarray, labels: np.ndarray) -> np.ndarray:
    """d mean_b(loss) / d logits = (softmax * sum(labels) - labels) / B, the closed form jax.grad arrives at."""
    logits = np.asarray(logits, dtype=np.float64)
    p = np.exp(logits - logits.max(-1, keepdims=True))
    p /= p.sum(-1, keepdims=True)
    return ((p * labels.sum(-1, keepdims=True) - labels) / logits.shape[0]).astype(np.float32)



def synthetic_batch(cfg: MAEConfig, batch: int, seed: int = 0, noise_seed: int = 1000):
    """SURVEY.md section 8d synthetic inputs: uint8 images, f32 uniform noise."""
    g = torch.Generator().manual_seed(seed)
    images = torch.randint(0, 256, (batch, 3, cfg.image_size, cfg.image_size), generator=g, dtype=torch.uint8)
    gn = torch.Generator().manual_seed(noise_seed)
    noise = torch.rand(cfg.num_patches, generator=gn, dtype=torch.float32).numpy()
    return images, noise


def flops_per_image(cfg: MAEConfig) -> Tuple[float, float]:
    """Algorithmic (fwd, step) FLOPs per image, SURVEY.md section 8d formula."""
    n, m, N = cfg.keep, cfg.num_masked, cfg.num_patches
    J, d, L, P = cfg.num_cls_tokens, cfg.dim, cfg.layers, cfg.patch_dim
    Te, Td, dd, dl = n + J, N + J, cfg.dec_dim, cfg.dec_layers
    fwd = 2.0 * (n * P * d + L * (4 * Te * d * d + 2 * Te * Te * d + 8 * n * d * d + 8 * J * J * d * d)
                 + Te * d * dd + dl * (12 * Td * dd * dd + 2 * Td * Td * dd) + m * dd * P)
    return fwd, 3.0 * fwd - 2.0 * n * P * d
