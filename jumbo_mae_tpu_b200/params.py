"""Flat parameter storage with a Flax-shaped view tree.

The reference keeps its parameters in a Flax pytree (names in SURVEY.md section 8a;
src/modeling.py:106-298, src/pretraining.py:82-85).  The kernels want (a) one contiguous
f32 master buffer so that AdamW and the gradient all-reduce are single passes, (b) a bf16
compute copy at the same offsets for the TMA-fed GEMMs, and (c) q/k/v projection weights
fused into one [d, 3d] matrix.  ``ParamStore`` lays the leaves out once and exposes the
reference's tree as (possibly strided) views, so ``state.params["model"]["layer_0"]
["attn"]["wq"]["kernel"]`` still has shape [d, heads, head_dim].

Layout rules
  * leaves whose Flax name is ``kernel`` come first: optax's weight-decay mask
    (src/pretraining.py:233) becomes the single boundary ``n_decay``;
  * every leaf starts at a multiple of 64 elements (TMA needs 16-byte alignment);
  * 2-D kernels get a row stride rounded up to 8 elements (ViT-H/14's 588-wide head).
"""
from __future__ import annotations

import re
from dataclasses import dataclass
from typing import Dict, List, Tuple

import torch

from .config import MAEConfig

ALIGN = 64


def _round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


@dataclass
class Leaf:
    name: str               # internal name, e.g. "model/layer_0/attn/wqkv/kernel"
    rows: int               # 2-D storage shape [rows, cols] with row stride ld (1-D: rows = 1)
    cols: int
    ld: int
    decay: bool
    offset: int = 0

    @property
    def size(self) -> int:
        return self.rows * self.ld


def _leaves(cfg: MAEConfig) -> List[Leaf]:
    d, dd, J, P = cfg.dim, cfg.dec_dim, cfg.num_cls_tokens, cfg.patch_dim
    out: List[Leaf] = []

    def kernel(name, rows, cols):
        out.append(Leaf(name, rows, cols, _round_up(cols, 8), True))

    def vec(name, n):
        out.append(Leaf(name, 1, n, n, False))

    def block(prefix, dim, with_norm3, layerscale):
        kernel(f"{prefix}/attn/wqkv/kernel", dim, 3 * dim)
        vec(f"{prefix}/attn/wqkv/bias", 3 * dim)
        kernel(f"{prefix}/attn/wo/kernel", dim, dim)
        vec(f"{prefix}/attn/wo/bias", dim)
        kernel(f"{prefix}/ff/w1/kernel", dim, 4 * dim)
        vec(f"{prefix}/ff/w1/bias", 4 * dim)
        kernel(f"{prefix}/ff/w2/kernel", 4 * dim, dim)
        vec(f"{prefix}/ff/w2/bias", dim)
        for nm, n in (("norm1", dim), ("norm2", dim)) + ((("norm3", J * dim),) if with_norm3 else ()):
            vec(f"{prefix}/{nm}/scale", n)
            vec(f"{prefix}/{nm}/bias", n)
        if layerscale:                  # LayerScale vectors "scale1/2/3": leaf name is not "kernel" -> no weight decay
            vec(f"{prefix}/scale1", dim)
            vec(f"{prefix}/scale2", dim)
            if with_norm3:
                vec(f"{prefix}/scale3", J * dim)

    kernel("model/embed/wte/kernel", P, d)
    vec("model/embed/wte/bias", d)
    if cfg.posemb == "learnable":       # [grid, grid, d] parameter named "wpe": no weight decay (mask = leaf name "kernel")
        vec("model/embed/wpe", cfg.num_patches * d)
    vec("model/cls_tokens", J * d)
    kernel("model/jumbo_mlp/w1/kernel", J * d, 4 * J * d)
    vec("model/jumbo_mlp/w1/bias", 4 * J * d)
    kernel("model/jumbo_mlp/w2/kernel", 4 * J * d, J * d)
    vec("model/jumbo_mlp/w2/bias", J * d)
    for i in range(cfg.layers):
        block(f"model/layer_{i}", d, True, getattr(cfg, "layerscale", False))
    vec("model/norm/scale", d)
    vec("model/norm/bias", d)
    if cfg.labels > 0:
        # classification mode: LinearCLS on the concatenated CLS tokens (src/modeling.py:209-219,236-240), no decoder
        kernel("model/head/Dense_0/kernel", J * d, cfg.labels)
        vec("model/head/Dense_0/bias", cfg.labels)
        return out
    for i in range(cfg.dec_layers):
        block(f"decoder_model/dec_layer_{i}", dd, False, getattr(cfg, "dec_layerscale", False))
    vec("decoder_model/dec_norm/scale", dd)
    vec("decoder_model/dec_norm/bias", dd)
    vec("image_mask_embedding", dd)
    kernel("decoder_proj/kernel", d, dd)
    vec("decoder_proj/bias", dd)
    kernel("decoder_image_output/kernel", dd, P)
    vec("decoder_image_output/bias", P)
    return out


def get_layer_index(name: str, num_layers: int) -> int:
    """get_layer_index_fn (src/utils.py:142-147) on a '/'-joined leaf path: ``model/layer_k/...`` -> k + 1,
    ``model/embed/...`` -> 0, every other leaf (CLS tokens, shared Jumbo MLP, final norm, decoder, heads) ->
    ``num_layers``.  Label i is updated with the learning rate times ``lr_decay ** (num_layers - i)``
    (src/pretraining.py:235-242, src/finetuning.py:241-248)."""
    parts = name.split("/")
    if parts[0] == "model" and len(parts) > 1 and parts[1].startswith("layer_"):
        return int(re.match(r"layer_(\d+)", parts[1]).group(1)) + 1
    if parts[0] == "model" and len(parts) > 1 and parts[1] == "embed":
        return 0
    return num_layers


class ParamStore:
    """Flat f32 master parameters (+ bf16 compute copy, gradients, Adam moments)."""

    def __init__(self, cfg: MAEConfig, device="cuda"):
        self.cfg = cfg
        leaves = _leaves(cfg)
        ordered = [l for l in leaves if l.decay] + [l for l in leaves if not l.decay]
        off = 0
        for leaf in ordered:
            leaf.offset = off
            off += _round_up(leaf.size, ALIGN)
            if leaf.decay:
                self.n_decay = off
        self.total = off
        self.leaves: Dict[str, Leaf] = {l.name: l for l in ordered}
        self.device = torch.device(device)
        self.p = torch.zeros(self.total, dtype=torch.float32, device=self.device)
        self.p16 = torch.zeros(self.total, dtype=torch.bfloat16, device=self.device)
        self.g = torch.zeros(self.total, dtype=torch.float32, device=self.device)
        self.m = torch.zeros(self.total, dtype=torch.float32, device=self.device)
        self.v = torch.zeros(self.total, dtype=torch.float32, device=self.device)

    # ------------------------------------------------------------- raw views
    def view(self, buf: torch.Tensor, name: str) -> torch.Tensor:
        """[rows, ld] (2-D kernels) or [cols] (vectors) view of a leaf inside ``buf``."""
        l = self.leaves[name]
        flat = buf[l.offset:l.offset + l.size]
        return flat.view(l.rows, l.ld) if l.rows > 1 or l.decay else flat

    def w(self, name: str) -> torch.Tensor:        # bf16 compute copy
        return self.view(self.p16, name)

    def f(self, name: str) -> torch.Tensor:        # f32 master
        return self.view(self.p, name)

    def grad(self, name: str) -> torch.Tensor:
        return self.view(self.g, name)

    def ld(self, name: str) -> int:
        return self.leaves[name].ld

    def ranges(self, select) -> List[Tuple[int, int]]:
        """Merged [start, end) element ranges of the flat buffers covering the leaves whose name satisfies
        ``select(name)`` (padding between consecutive selected leaves is included)."""
        out: List[Tuple[int, int]] = []
        for leaf in sorted(self.leaves.values(), key=lambda l: l.offset):
            if not select(leaf.name):
                continue
            start, end = leaf.offset, leaf.offset + _round_up(leaf.size, ALIGN)
            if out and out[-1][1] == start:
                out[-1] = (out[-1][0], end)
            else:
                out.append((start, end))
        return out

    def lamb_tables(self, select=None, chunk_elems: int = 32768):
        """Group / chunk tables of jm_lamb over the ``kernel`` leaves whose name satisfies ``select`` (all by default).
        A group is one kernel leaf of the FLAX tree -- the trust ratio of modified_lamb (src/utils.py:124-139) is taken
        per leaf -- i.e. a column range of a stored matrix: the fused ``wqkv`` kernel holds wq, wk and wv.  Returns
        (groups int64 [G, 7], chunks int64 [C, 3]) on the store's device, or (None, None) if nothing is selected."""
        groups, chunks = [], []
        for leaf in sorted(self.leaves.values(), key=lambda l: l.offset):
            if not leaf.decay or (select is not None and not select(leaf.name)):
                continue
            parts = 3 if leaf.name.endswith("/wqkv/kernel") else 1
            width = leaf.cols // parts
            for i in range(parts):
                rows_per = max(1, chunk_elems // width)
                first = len(chunks)
                for r0 in range(0, leaf.rows, rows_per):
                    chunks.append([len(groups), r0, min(rows_per, leaf.rows - r0)])
                groups.append([leaf.offset, leaf.rows, leaf.ld, i * width, width, first, len(chunks) - first])
        if not groups:
            return None, None
        return (torch.tensor(groups, dtype=torch.int64, device=self.device),
                torch.tensor(chunks, dtype=torch.int64, device=self.device))

    # ------------------------------------------------------------- Flax-shaped trees
    def tree(self, buf: torch.Tensor | None = None) -> dict:
        """The reference's parameter pytree as views into ``buf`` (default: f32 masters)."""
        buf = self.p if buf is None else buf
        cfg = self.cfg
        J = cfg.num_cls_tokens

        def k(name):                       # 2-D kernel without the stride padding
            l = self.leaves[name]
            return self.view(buf, name)[:, :l.cols]

        def v(name):
            return self.view(buf, name)

        def attn(prefix, dim, heads):
            hd = dim // heads
            wqkv, bqkv = k(f"{prefix}/wqkv/kernel"), v(f"{prefix}/wqkv/bias")
            out = {}
            for i, nm in enumerate(("wq", "wk", "wv")):
                out[nm] = {"kernel": wqkv[:, i * dim:(i + 1) * dim].unflatten(1, (heads, hd)),
                           "bias": bqkv[i * dim:(i + 1) * dim].view(heads, hd)}
            out["wo"] = {"kernel": k(f"{prefix}/wo/kernel").unflatten(0, (heads, hd)), "bias": v(f"{prefix}/wo/bias")}
            return out

        def ff(prefix):
            return {"w1": {"kernel": k(f"{prefix}/w1/kernel"), "bias": v(f"{prefix}/w1/bias")},
                    "w2": {"kernel": k(f"{prefix}/w2/kernel"), "bias": v(f"{prefix}/w2/bias")}}

        def ln(prefix):
            return {"scale": v(f"{prefix}/scale"), "bias": v(f"{prefix}/bias")}

        model = {
            "embed": {"wte": {"kernel": k("model/embed/wte/kernel").view(cfg.patch_size, cfg.patch_size, 3, cfg.dim),
                              "bias": v("model/embed/wte/bias")}},
            "cls_tokens": v("model/cls_tokens").view(1, J, cfg.dim),
            "jumbo_mlp": ff("model/jumbo_mlp"),
            "norm": ln("model/norm"),
        }
        if cfg.posemb == "learnable":
            model["embed"]["wpe"] = v("model/embed/wpe").view(cfg.grid, cfg.grid, cfg.dim)
        for i in range(cfg.layers):
            p = f"model/layer_{i}"
            model[f"layer_{i}"] = {"attn": attn(f"{p}/attn", cfg.dim, cfg.heads), "ff": ff(f"{p}/ff"),
                                   "norm1": ln(f"{p}/norm1"), "norm2": ln(f"{p}/norm2"), "norm3": ln(f"{p}/norm3")}
            if getattr(cfg, "layerscale", False):
                model[f"layer_{i}"].update({nm: v(f"{p}/{nm}") for nm in ("scale1", "scale2", "scale3")})
        if cfg.labels > 0:
            model["head"] = {"Dense_0": {"kernel": k("model/head/Dense_0/kernel"), "bias": v("model/head/Dense_0/bias")}}
            return {"model": model}
        dec = {"dec_norm": ln("decoder_model/dec_norm")}
        for i in range(cfg.dec_layers):
            p = f"decoder_model/dec_layer_{i}"
            dec[f"dec_layer_{i}"] = {"attn": attn(f"{p}/attn", cfg.dec_dim, cfg.dec_heads), "ff": ff(f"{p}/ff"),
                                     "norm1": ln(f"{p}/norm1"), "norm2": ln(f"{p}/norm2")}
            if getattr(cfg, "dec_layerscale", False):
                dec[f"dec_layer_{i}"].update({nm: v(f"{p}/{nm}") for nm in ("scale1", "scale2")})
        return {
            "model": model,
            "decoder_model": dec,
            "image_mask_embedding": v("image_mask_embedding").view(1, 1, cfg.dec_dim),
            "decoder_proj": {"kernel": k("decoder_proj/kernel"), "bias": v("decoder_proj/bias")},
            "decoder_image_output": {"kernel": k("decoder_image_output/kernel"),
                                     "bias": v("decoder_image_output/bias")},
        }

    def load_tree(self, tree: dict) -> None:
        """Copies a Flax-shaped tree of (CPU or CUDA) tensors into the masters and refreshes bf16."""
        mine = self.tree()

        def walk(dst, src, path):
            for key, val in dst.items():
                if key not in src:
                    raise KeyError(f"parameter tree is missing {'/'.join(path + (key,))}")
                if isinstance(val, dict):
                    walk(val, src[key], path + (key,))
                else:
                    s = src[key]
                    if tuple(s.shape) != tuple(val.shape):
                        raise ValueError(f"{'/'.join(path + (key,))}: shape {tuple(s.shape)} != {tuple(val.shape)}")
                    val.copy_(s.to(device=val.device, dtype=val.dtype))
        walk(mine, tree, ())
        self.refresh_bf16()

    def refresh_bf16(self) -> None:
        from . import ops
        ops.cast_bf16(self.p, self.p16)


def tree_leaves(tree: dict, prefix: Tuple[str, ...] = ()):
    out = []
    for key in sorted(tree.keys()):
        val = tree[key]
        if isinstance(val, dict):
            out.extend(tree_leaves(val, prefix + (key,)))
        else:
            out.append((prefix + (key,), val))
    return out
