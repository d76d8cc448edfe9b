"""Host-side mirror of the reference's fine-tune step API (src/finetuning.py) over ``FinetuneEngine``.

Same names and argument meaning: ``create_train_state(args)``, ``training_step(state, batch)`` with
``batch = (images uint8 [B,3,H,W], labels int32 [B])`` returning ``(state, {"loss", "acc1", "acc5",
"learning_rate"})``, ``validation_step(state, batch)`` returning summed ``{"loss", "acc1", "acc5", "num_samples"}``
over the rows whose label is not -1 (src/finetuning.py:157-165).  One process per GPU: under ``torch.distributed``
the gradients are sum-all-reduced over NCCL before the update (1/world folded into AdamW), the step metrics are
averaged and the validation sums are summed over ranks, as ``lax.pmean`` / ``lax.psum`` do (src/finetuning.py:131-165).

RNG: the reference draws the Mixup / CutMix choice, the beta sample, the batch permutation and the box origin from
its ``mixup`` threefry stream (src/utils.py:67-111); those bits cannot be reproduced without JAX, so the draws are
data at this boundary: ``training_step(..., mix=MixDraw(...))`` takes them from the caller, and without ``mix`` they
come from a Philox stream seeded with ``--mixup-seed`` that advances every step.  (In the reference the fine-tune
streams never advance: ``updates`` of ``split_rngs()`` is overwritten by the ``batch_stats`` updates,
src/finetuning.py:136-137,154 -- every step reuses the same key.  That looks unintended and is not reproduced.)

DropPath (``--droppath``, 0.1 in the recipe, config/ft.sh): one Bernoulli draw per image and residual branch
(src/modeling.py:181-183) from this rank's "dropout" Philox stream, or supplied by the caller
(``training_step(..., drop_keep=...)``: bool [layers, 3, B]); the factors keep / (1 - rate) scale the branch outputs in the
GEMM epilogues and the branch gradients in the backward pass.

Not built: Dropout > 0, BatchNorm head and stop-gradient of linear probing (``--mode linear``), ``bce``,
LARS / LAMB / SGD.
"""
from __future__ import annotations

import argparse
import math
from dataclasses import dataclass, field
from typing import Any, Dict, Optional, Tuple

import numpy as np
import torch

from . import _lib, ops
from .config import MAEConfig
from .finetune_engine import FinetuneEngine
from .parallel import GradReducer, process_count, process_index
from .params import ParamStore
from .pretraining import USE_CUDA_GRAPH


@dataclass
class MixDraw:
    """One step's Mixup / CutMix draws (Mixup.__call__, src/utils.py:94-111)."""
    perm: np.ndarray                       # jax.random.permutation(batch)
    ratio: float                           # the beta sample
    box: Optional[Tuple[float, float]] = None   # (xstart, ystart) uniforms: CutMix; None: Mixup


def random_bounding_box(ratio: float, xstart: float, ystart: float, width: int, height: int):
    """Mixup.random_bounding_box (src/utils.py:85-92): column / row membership of the box (uint8) and the label
    weight of the un-pasted image = mean of the image mask (src/utils.py:77-78)."""
    size = np.float32((np.float32(1) - np.float32(ratio)) ** np.float32(0.5))
    xs, ys, half = np.float32(xstart), np.float32(ystart), np.float32(0.5)
    xr, yr = np.linspace(0, 1, width, dtype=np.float32), np.linspace(0, 1, height, dtype=np.float32)
    xmask = (xs - half * size <= xr) & (xr < xs + half * size)
    ymask = (ys - half * size <= yr) & (yr < ys + half * size)
    label_ratio = float(np.float32(1) - np.float32(xmask.sum() * ymask.sum()) / np.float32(width * height))
    return xmask.astype(np.uint8), ymask.astype(np.uint8), label_ratio


def lr_schedule(count: int, state: "TrainState") -> float:
    """optax.warmup_cosine_decay_schedule(1e-6, --learning-rate, warmup, training_steps, 1e-6)
    (src/finetuning.py:262-268; no batch-size scaling for adamw, :256-259)."""
    init_value, end_value, peak, warm = 1e-6, 1e-6, state.learning_rate, state.warmup_steps
    if count < warm:
        return (init_value - peak) * (1.0 - count / warm) + peak
    decay_steps = state.training_steps - warm
    c = min(count - warm, decay_steps)
    cosine = 0.5 * (1.0 + math.cos(math.pi * c / decay_steps))
    alpha = end_value / peak
    return peak * ((1.0 - alpha) * cosine + alpha)


@dataclass
class TrainState:
    """Numeric state of the reference's fine-tune TrainState (src/finetuning.py:51-75)."""
    step: int
    store: ParamStore
    cfg: MAEConfig
    mixup_rng: int
    dropout_rng: int
    noise_rng: int
    learning_rate: float
    warmup_steps: int
    training_steps: int
    label_smoothing: float = 0.1
    mixup: float = 0.8
    cutmix: float = 1.0
    lr_decay: float = 1.0
    clip_grad: float = 0.0
    droppath: float = 0.0
    optimizer: str = "adamw"    # adamw | lamb (modified_lamb, src/utils.py:124-139; src/finetuning.py:45,232-240)
    micro_step: int = 0
    micro_in_mini: int = 1
    batch_stats: Any = None
    engines: Dict[int, FinetuneEngine] = field(default_factory=dict)
    reducer: Optional[GradReducer] = None
    _drop_gen: Optional[np.random.Generator] = None
    _mix_gen: Optional[np.random.Generator] = None

    @property
    def params(self) -> dict:
        return self.store.tree()

    @property
    def opt_state(self) -> dict:
        return {"count": self.step, "mu": self.store.tree(self.store.m), "nu": self.store.tree(self.store.v),
                "hyperparams": {"learning_rate": lr_schedule(max(self.step - 1, 0), self)}}

    def replicate(self) -> "TrainState":
        # one stream per rank, as shard_prng_key gives every device its own key (src/finetuning.py:71-76)
        self._mix_gen = np.random.Generator(np.random.Philox(key=self.mixup_rng + 7919 * process_index()))
        self._drop_gen = np.random.Generator(np.random.Philox(key=self.dropout_rng + 7919 * process_index() + 15485863))
        return self

    def draw_drop_keep(self, batch: int) -> Optional[np.ndarray]:
        """This step's DropPath keep bits, bool [layers, 3, B] (drop_path1 / 2 / 3 of every JumboLayer)."""
        if self.droppath <= 0:
            return None
        if self._drop_gen is None:
            self.replicate()
        return self._drop_gen.random((self.cfg.layers, 3, batch)) >= self.droppath

    def engine(self, batch: int) -> FinetuneEngine:
        if batch not in self.engines:
            self.engines[batch] = FinetuneEngine(self.cfg, self.store, batch)
        return self.engines[batch]

    def draw_mix(self, batch: int) -> Optional[MixDraw]:
        """Mixup.__call__ (src/utils.py:94-111): none, mixup, cutmix, or one of the two by a fair coin."""
        if self.mixup == 0 and self.cutmix == 0:
            return None
        if self._mix_gen is None:
            self.replicate()
        g = self._mix_gen
        use_cutmix = self.cutmix > 0 and (self.mixup == 0 or not (g.random() > 0.5))
        alpha = self.cutmix if use_cutmix else self.mixup
        ratio = float(np.float32(g.beta(alpha, alpha)))
        box = (float(g.random(dtype=np.float32)), float(g.random(dtype=np.float32))) if use_cutmix else None
        return MixDraw(perm=g.permutation(batch).astype(np.int32), ratio=ratio, box=box)


def create_train_state(args: argparse.Namespace) -> TrainState:
    """src/finetuning.py:168-284 for the adamw + cross-entropy fine-tune recipe."""
    _lib.require_device()
    if getattr(args, "optimizer", "adamw") not in ("adamw", "lamb"):
        raise NotImplementedError("--optimizer must be adamw or lamb (sgd / lars are not built)")
    if getattr(args, "criterion", "ce") != "ce":
        raise NotImplementedError("only --criterion ce is built")
    if getattr(args, "mode", "finetune") == "linear":
        raise NotImplementedError("--mode linear (BatchNorm head, stop-gradient) is not built")
    if getattr(args, "dropout", 0.0) != 0.0:
        raise NotImplementedError("--dropout > 0 is not built (0.0 in the fine-tune recipe)")
    if not 0.0 <= float(getattr(args, "droppath", 0.0)) < 1.0:
        raise ValueError("--droppath must be in [0, 1)")
    cfg = MAEConfig(layers=args.layers, dim=args.dim, heads=args.heads, patch_size=args.patch_size,
                    image_size=args.image_size, labels=args.labels, weight_decay=args.weight_decay,
                    adam_b1=args.adam_b1, adam_b2=args.adam_b2, adam_eps=args.adam_eps,
                    grad_accum=getattr(args, "grad_accum", 1), posemb=getattr(args, "posemb", "sincos2d"),
                    grad_ckpt=bool(getattr(args, "grad_ckpt", False)), layerscale=bool(getattr(args, "layerscale", False)))
    store = ParamStore(cfg)
    gen = torch.Generator(device="cpu").manual_seed(int(getattr(args, "init_seed", 0)))
    for name, leaf in store.leaves.items():           # truncated_normal(0.02) kernels, LN scales 1, the rest 0
        view = store.f(name)
        if leaf.decay:
            w = torch.empty(leaf.rows, leaf.cols)
            torch.nn.init.trunc_normal_(w, std=0.02, a=-0.04, b=0.04, generator=gen)
            view[:, :leaf.cols].copy_(w)
        elif name.endswith("/scale"):
            view.fill_(1.0)
        elif name.rsplit("/", 1)[-1] in ("scale1", "scale2", "scale3"):
            view.fill_(1e-4)
    store.refresh_bf16()
    state = _new_state(args, store, cfg)
    if getattr(args, "pretrained_ckpt", None) is not None:         # src/finetuning.py:215-217
        from .checkpoint import load_pretrained_params
        n = load_pretrained_params(state, args.pretrained_ckpt)
        print(f"[*] load pretrained params with overlap of {n}/{len(state.params['model'])}")
    return state


def _new_state(args: argparse.Namespace, store: ParamStore, cfg: MAEConfig) -> TrainState:
    return TrainState(step=0, store=store, cfg=cfg, mixup_rng=int(getattr(args, "mixup_seed", 0)),
                      dropout_rng=int(getattr(args, "dropout_seed", 0)), noise_rng=int(getattr(args, "noise_seed", 0)),
                      learning_rate=float(args.learning_rate), warmup_steps=int(args.warmup_steps),
                      training_steps=int(args.training_steps),
                      label_smoothing=float(getattr(args, "label_smoothing", 0.1)), mixup=float(getattr(args, "mixup", 0.8)),
                      cutmix=float(getattr(args, "cutmix", 1.0)), lr_decay=float(getattr(args, "lr_decay", 1.0)),
                      clip_grad=float(getattr(args, "clip_grad", 0.0)), micro_in_mini=int(getattr(args, "grad_accum", 1)),
                      droppath=float(getattr(args, "droppath", 0.0)), optimizer=getattr(args, "optimizer", "adamw"))


def _to_device(t, dev, dtype):
    t = torch.as_tensor(t)
    if t.dtype != dtype:
        t = t.to(dtype)
    return t.to(dev, non_blocking=True).contiguous()


def _set_droppath(state: TrainState, eng: FinetuneEngine, keep) -> None:
    """keep bits [layers, 3, B] -> the engine's per-image factors keep / (1 - rate) (a persistent device buffer)."""
    if keep is None:
        eng.set_droppath(None)
        return
    fac = torch.from_numpy(np.ascontiguousarray(keep, dtype=np.float32) / np.float32(1.0 - state.droppath))
    buf = getattr(eng, "_drop_buf", None)
    if buf is None:
        buf = eng._drop_buf = torch.empty(fac.shape, dtype=torch.float32, device=state.store.device)
    buf.copy_(fac, non_blocking=True)
    eng.set_droppath(buf)


def training_step(state: TrainState, batch, mix: Optional[MixDraw] = None, drop_keep=None):
    """src/finetuning.py:108-154: forward with label smoothing + Mixup / CutMix + DropPath, backward, (accumulated)
    AdamW.  ``drop_keep`` (bool [layers, 3, B]) overrides this step's DropPath draws (parity tests)."""
    dev = state.store.device
    images, labels = _to_device(batch[0], dev, torch.uint8), _to_device(batch[1], dev, torch.int32)
    B = images.shape[0]
    eng = state.engine(B)
    if mix is None:
        mix = state.draw_mix(B)
    kw: Dict[str, Any] = {}
    if mix is not None:
        kw = dict(perm=_to_device(mix.perm, dev, torch.int32), ratio=float(mix.ratio))
        if mix.box is not None:
            xm, ym, label_ratio = random_bounding_box(mix.ratio, mix.box[0], mix.box[1], state.cfg.image_size,
                                                      state.cfg.image_size)
            kw.update(xmask=torch.from_numpy(xm).to(dev), ymask=torch.from_numpy(ym).to(dev), label_ratio=label_ratio)
    accum = state.micro_in_mini
    world = process_count()
    if world > 1 and getattr(state, "reducer", None) is None:
        state.reducer = GradReducer(state.store)
    _set_droppath(state, eng, drop_keep if drop_keep is not None else state.draw_drop_keep(B))
    mode = "plain" if mix is None else ("cutmix" if mix.box is not None else "mixup")
    first = state.micro_step == 0
    reduced = False
    # Graph replay on a single rank only.  With several ranks the kind of step (captured graph or eager fallback) could
    # differ from rank to rank -- the Mixup / CutMix draws are per rank -- and every rank has to issue the same sequence
    # of collectives, so the data-parallel step keeps one code path: eager launches, one all-reduce after the backward.
    use_graph = USE_CUDA_GRAPH and labels.dim() == 1 and world == 1
    if use_graph and not getattr(eng, "graphs_by_mode", None):
        if first:
            # every kind of step this state can draw is captured once, before any gradient is being accumulated (the
            # capture's warm-up pass overwrites the gradient buffer); the draws themselves -- permutation, ratios, box
            # vectors, DropPath factors -- are data in device buffers the graphs read
            modes = (["mixup"] if state.mixup > 0 else []) + (["cutmix"] if state.cutmix > 0 else [])
            for m in dict.fromkeys(modes + [mode]):
                eng.capture_graphs(label_smoothing=state.label_smoothing, mode=m)
        else:
            use_graph = False
    if use_graph and (mode not in eng.graphs_by_mode or eng._graph_droppath != (eng.drop_masks is not None)):
        use_graph = False                                          # (a kind of step that was not captured: eager)
    if use_graph:
        gkw = dict(mode=mode, perm=kw.get("perm"), xmask=kw.get("xmask"), ymask=kw.get("ymask"),
                   ratios=None if mix is None else (kw["ratio"], kw.get("label_ratio", kw["ratio"])))
        eng.run_graph(images, labels, first, **gkw)
    else:
        eng.forward(images, labels, label_smoothing=state.label_smoothing, **kw)
        eng.backward(zero_grads=first)
    metrics = eng.metrics.clone()
    if world > 1:                                                  # lax.pmean(metrics) (src/finetuning.py:131)
        state.reducer.mean_scalar(metrics)
    state.micro_step += 1
    lr = lr_schedule(state.step, state)
    if state.micro_step == accum:
        if world > 1 and not reduced:                              # lax.pmean(grads) (:138,144); 1/world below
            state.reducer.all_reduce_grads()
        eng.layerscale_finalize()
        if state.clip_grad > 0:
            if not hasattr(state, "_sumsq"):
                state._sumsq = torch.zeros(1, dtype=torch.float32, device=dev)
            ops.clip_by_global_norm(state.store.g, 1.0 / (accum * world), state.clip_grad, state._sumsq)
        step_fn = eng.lamb_step if state.optimizer == "lamb" else eng.adamw_step
        step_fn(state.step, lr, grad_scale=1.0 / (accum * world), lr_decay=state.lr_decay)
        state.step += 1
        state.micro_step = 0
    return state, {"loss": metrics[0], "acc1": metrics[1], "acc5": metrics[2], "learning_rate": lr}


def validation_step(state: TrainState, batch) -> Dict[str, Any]:
    """src/finetuning.py:157-165: det=True forward; rows with label -1 (padding of the last batch) are excluded;
    sums, not means, so that the caller divides by ``num_samples``."""
    dev = state.store.device
    images, labels = _to_device(batch[0], dev, torch.uint8), _to_device(batch[1], dev, torch.int32)
    valid = (labels != -1)
    eng = state.engine(images.shape[0])
    eng.set_droppath(None)                                         # det=True (src/finetuning.py:159)
    eng.forward(images, torch.where(valid, labels, torch.zeros_like(labels)), want_grad=False)
    v = valid.to(torch.float32)
    sums = torch.stack([(eng.loss_rows * v).sum(), (eng.acc1 * v).sum(), (eng.acc5 * v).sum(), v.sum()])
    if process_count() > 1:                                        # lax.psum(metrics) (src/finetuning.py:165)
        import torch.distributed as dist
        dist.all_reduce(sums)
    return {"loss": sums[0], "acc1": sums[1], "acc5": sums[2], "num_samples": sums[3]}
