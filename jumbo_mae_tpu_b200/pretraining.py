"""Train-step entry points with the reference's names and call shapes (src/pretraining.py).

    state = create_train_state(args).replicate()
    state, metrics = training_step(state, batch)       # batch = (images uint8 [B,3,H,W],)
    metrics = validation_step(state, batch)

One process drives one GPU (the reference's ``jax.pmap`` over local devices becomes one
rank per device, ``lax.pmean`` becomes an NCCL all-reduce over ``torch.distributed``), so
``batch`` is this rank's shard and ``replicate()`` only derives the per-rank RNG streams.
"""
from __future__ import annotations

import argparse
import os
from dataclasses import dataclass, field
from typing import Any, Dict, Optional, Tuple

import numpy as np
import torch

from . import _lib, ops
from .config import MAEConfig
from .engine import MAEEngine, lr_schedule
from .modeling import MAEDecoder, ViT, _Bound
from .parallel import GradReducer, process_count, process_index
from .params import ParamStore


# One CUDA-graph launch per micro-step instead of ~1800 Python-issued kernel launches (JM_NO_GRAPH=1: eager).
USE_CUDA_GRAPH = os.environ.get("JM_NO_GRAPH", "0") != "1"


class PretrainModule(_Bound):
    """PretrainModule (src/pretraining.py:76-122): ViT encoder + MAEDecoder + pixel head.

    ``PretrainModule(...).bind(store, rngs={"noise": noise})(images, det=True)`` returns ``{"loss": scalar}`` like
    ``PretrainModule.__call__(images, det=True)`` (:87); images are the uint8 NCHW batch."""

    def __init__(self, model: ViT, decoder_model: MAEDecoder, image_size: int, norm_pix_loss: bool = False):
        self.model, self.decoder_model = model, decoder_model
        self.image_size, self.norm_pix_loss = image_size, norm_pix_loss

    def config(self, **opt) -> MAEConfig:
        m, dm = self.model, self.decoder_model
        return MAEConfig(layers=m.layers, dim=m.dim, heads=m.heads, patch_size=m.patch_size, image_size=m.image_size,
                         image_mask_ratio=m.image_mask_ratio, num_cls_tokens=m.num_cls_tokens,
                         dec_layers=dm.dec_layers, dec_dim=dm.dec_dim, dec_heads=dm.dec_heads,
                         norm_pix_loss=self.norm_pix_loss, **opt)

    def __call__(self, images: torch.Tensor, det: bool = True) -> Dict[str, torch.Tensor]:
        store = self._need_store()
        if store.cfg.norm_pix_loss != self.norm_pix_loss or store.cfg.dim != self.model.dim:
            raise _lib.JmError("the bound parameter store was built for a different module")
        images = _device_images(images).to(store.device)
        eng = self._engine(images.shape[0])
        loss = eng.forward(images, self._noise(store.device), want_grad=False)
        return {"loss": loss[0].clone()}


@dataclass
class TrainState:
    """Numeric state of flax's TrainState + the reference's extra fields (src/pretraining.py:50-57).

    ``params`` / ``opt_state`` are Flax-shaped trees of views into the flat device buffers
    held by ``store`` (the step updates them in place: the reference donates its state)."""
    step: int
    store: ParamStore
    cfg: MAEConfig
    module: PretrainModule
    mixup_rng: int
    dropout_rng: int
    noise_rng: int
    micro_step: int = 0
    micro_in_mini: int = 1
    engines: Dict[int, MAEEngine] = field(default_factory=dict)
    reducer: Optional[GradReducer] = None
    _noise_gen: Optional[np.random.Generator] = None
    lr_decay: float = 1.0       # --lr-decay: layer-wise learning-rate scale table behind adamw (src/pretraining.py:235-242)
    optimizer: str = "adamw"    # --optimizer adamw | lamb (modified_lamb, src/utils.py:124-139)
    clip_grad: float = 0.0      # --clip-grad: optax.clip_by_global_norm in front of adamw (src/pretraining.py:243-244)

    @property
    def params(self) -> dict:
        return self.store.tree()

    @property
    def opt_state(self) -> dict:
        return {"count": self.step, "mu": self.store.tree(self.store.m), "nu": self.store.tree(self.store.v),
                "hyperparams": {"learning_rate": lr_schedule(max(self.step - 1, 0), self.cfg)}}

    @property
    def grad_accum(self):
        return self.store.tree(self.store.g) if self.micro_in_mini > 1 else None

    def replicate(self) -> "TrainState":
        """src/pretraining.py:68-73.  Parameters already live on this rank's device; what
        remains is giving every device its own RNG streams (shard_prng_key)."""
        self._noise_gen = np.random.Generator(np.random.Philox(key=self.noise_rng + 7919 * process_index()))
        return self

    def noise_shape(self, batch: Optional[int]):
        """[num_patches] (one permutation per batch, the reference) or [batch, num_patches] (--per-image-masking)."""
        if getattr(self.cfg, "per_image_masking", False):
            if batch is None:
                raise _lib.JmError("per-image masking: the noise draw needs the batch size")
            return (batch, self.cfg.num_patches)
        return (self.cfg.num_patches,)

    def split_rngs(self, batch: Optional[int] = None) -> Tuple[dict, dict]:
        """src/pretraining.py:59-66: draws this step's "noise" stream (mixup / dropout are
        identity at the recipe's rates).  Returns (rngs, updates) like the reference."""
        if self._noise_gen is None:
            self.replicate()
        noise = self._noise_gen.random(self.noise_shape(batch), dtype=np.float32)
        return {"noise": noise}, {}

    def engine(self, batch: int) -> MAEEngine:
        """Activation buffers (and CUDA graphs) exist per per-rank batch size: keep it fixed, as the reference's loader
        does (it pads / drops the ragged last batch).  At most two engines are kept (training + validation batch size);
        an older one-off shape is released when a third appears."""
        if batch not in self.engines:
            if len(self.engines) >= 2:
                oldest = next(iter(self.engines))
                del self.engines[oldest]
            self.engines[batch] = MAEEngine(self.cfg, self.store, batch)
        return self.engines[batch]

    def prefetch(self, batch):
        """Starts the host-to-device copy of the NEXT batch on a copy stream, so that it overlaps the step that
        is running (the reference's ``shard(batch)`` + XLA's asynchronous dispatch give it
        the same overlap, src/main_pretrain.py:61-64).  ``batch`` is ``(images,)`` with uint8 images in pinned
        host memory; returns the device-side ``(images,)`` to hand to ``training_step``.  Two staging buffers
        alternate; a buffer is rewritten only after the step that consumed it has copied it out."""
        images = batch[0] if isinstance(batch, (tuple, list)) else batch
        if images.is_cuda:
            return (images,)
        dev = self.store.device
        if not hasattr(self, "_stage"):
            self._stage = {"stream": torch.cuda.Stream(device=dev), "bufs": {}, "next": 0, "ready": {}, "freed": {}}
        st = self._stage
        key = (tuple(images.shape), st["next"])
        st["next"] ^= 1
        if key not in st["bufs"]:
            st["bufs"][key] = torch.empty(images.shape, dtype=torch.uint8, device=dev)
        buf = st["bufs"][key]
        with torch.cuda.stream(st["stream"]):
            freed = st["freed"].pop(buf.data_ptr(), None)
            if freed is not None:
                st["stream"].wait_event(freed)          # the step that read this buffer has copied it out
            buf.copy_(images, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(st["stream"])
        st["ready"][buf.data_ptr()] = ev
        return (buf,)

    def _consume_prefetched(self, images: torch.Tensor) -> None:
        """Orders the compute stream after the copy that filled a prefetched buffer."""
        st = getattr(self, "_stage", None)
        if st is None:
            return
        ev = st["ready"].pop(images.data_ptr(), None)
        if ev is not None:
            torch.cuda.current_stream().wait_event(ev)

    def _release_prefetched(self, images: torch.Tensor) -> None:
        st = getattr(self, "_stage", None)
        if st is None or not any(b.data_ptr() == images.data_ptr() for b in st["bufs"].values()):
            return
        ev = torch.cuda.Event()
        ev.record()
        st["freed"][images.data_ptr()] = ev


def create_train_state(args: argparse.Namespace) -> TrainState:
    """src/pretraining.py:170-270: builds the modules from the CLI namespace, initialises the
    parameters (truncated-normal(0.02) kernels, zero biases / CLS, unit LayerNorm scales) and the
    AdamW state, seeds the RNG streams with ``seed + process_index``."""
    _lib.require_device()
    if getattr(args, "optimizer", "adamw") not in ("adamw", "lamb"):
        raise NotImplementedError("--optimizer must be adamw (both pre-training presets) or lamb")
    for nm in ("dropout", "dec_dropout"):
        if getattr(args, nm, 0.0) != 0.0:
            raise NotImplementedError(f"--{nm.replace('_', '-')} > 0 is not built (0.0 in both presets)")
    posemb = getattr(args, "posemb", "sincos2d")
    model = ViT(layers=args.layers, dim=args.dim, heads=args.heads, labels=getattr(args, "labels", 0),
                patch_size=args.patch_size, image_size=args.image_size, posemb=posemb,
                image_mask_ratio=args.image_mask_ratio)
    decoder = MAEDecoder(dec_layers=args.dec_layers, dec_dim=args.dec_dim, dec_heads=args.dec_heads,
                         patch_size=args.patch_size, image_size=args.image_size)
    module = PretrainModule(model, decoder, args.image_size, getattr(args, "norm_pix_loss", False))
    cfg = module.config(learning_rate=args.learning_rate, train_batch_size=args.train_batch_size,
                        weight_decay=args.weight_decay, adam_b1=args.adam_b1, adam_b2=args.adam_b2,
                        adam_eps=args.adam_eps, warmup_steps=args.warmup_steps, training_steps=args.training_steps,
                        grad_accum=getattr(args, "grad_accum", 1), posemb=posemb,
                        grad_ckpt=bool(getattr(args, "grad_ckpt", False)),
                        per_image_masking=bool(getattr(args, "per_image_masking", False)),
                        layerscale=bool(getattr(args, "layerscale", False)),
                        dec_layerscale=bool(getattr(args, "dec_layerscale", False)))
    cfg.validate()
    store = ParamStore(cfg, device=torch.device("cuda", torch.cuda.current_device()))
    init_params_(store, getattr(args, "init_seed", 0))
    return TrainState(step=0, store=store, cfg=cfg, module=module,
                      mixup_rng=getattr(args, "mixup_seed", 0) + process_index(),
                      dropout_rng=getattr(args, "dropout_seed", 0) + process_index(),
                      noise_rng=getattr(args, "noise_seed", 0) + process_index(),
                      micro_step=0, micro_in_mini=cfg.grad_accum, clip_grad=float(getattr(args, "clip_grad", 0.0)),
                      lr_decay=float(getattr(args, "lr_decay", 1.0)), optimizer=getattr(args, "optimizer", "adamw"))


def init_params_(store: ParamStore, seed: int) -> None:
    """Reference init rules (src/modeling.py:30-32,243; src/pretraining.py:38,83), identical on
    every rank: kernels and image_mask_embedding ~ truncated_normal(0.02) on [-2, 2] sigma."""
    gen = torch.Generator(device="cpu").manual_seed(seed)
    tree = store.tree()

    def walk(node, path):
        for key in sorted(node.keys()):
            val = node[key]
            if isinstance(val, dict):
                walk(val, path + (key,))
            elif key in ("kernel", "image_mask_embedding", "wpe"):     # truncated_normal(0.02) (src/modeling.py:116)
                t = torch.empty(tuple(val.shape), dtype=torch.float32)
                torch.nn.init.trunc_normal_(t, mean=0.0, std=1.0, a=-2.0, b=2.0, generator=gen)
                val.copy_((t * 0.02).to(val.device))
            elif key == "scale":
                val.fill_(1.0)
            elif key in ("scale1", "scale2", "scale3"):                 # LayerScale: init.constant(1e-4) (src/modeling.py:160-162)
                val.fill_(1e-4)
            else:
                val.zero_()
    walk(tree, ())
    store.refresh_bf16()


def _device_images(images) -> torch.Tensor:
    if isinstance(images, np.ndarray):
        images = torch.from_numpy(images)
    if images.dtype != torch.uint8:
        raise _lib.JmError("images must be uint8 NCHW (normalisation happens on the device, src/pretraining.py:88-91)")
    if not images.is_cuda and not USE_CUDA_GRAPH:
        images = images.cuda(non_blocking=True)
    return images.contiguous()


def _device_noise(nz, device) -> torch.Tensor:
    """The mask noise (numpy array, host or device tensor) as an f32 device vector."""
    if isinstance(nz, np.ndarray):
        nz = torch.from_numpy(np.ascontiguousarray(nz, dtype=np.float32))
    return nz.to(device=device, dtype=torch.float32, non_blocking=True).contiguous()


def training_step(state: TrainState, batch, noise=None) -> Tuple[TrainState, Dict[str, Any]]:
    """src/pretraining.py:125-159.  ``batch`` is the 1-tuple ``(images,)`` the reference splats
    into ``apply_fn``; images are this rank's uint8 [B, 3, H, W] shard (host or device).
    ``noise`` (f32 [num_patches]) overrides the state's noise stream (parity tests).
    Returns (state, {"loss": device scalar, "learning_rate": float})."""
    images = _device_images(batch[0] if isinstance(batch, (tuple, list)) else batch)
    rngs, _ = state.split_rngs(images.shape[0]) if noise is None else ({"noise": noise}, {})
    nz = _device_noise(rngs["noise"], state.store.device)
    eng = state.engine(images.shape[0])
    state._consume_prefetched(images)
    world = process_count()
    accum = state.micro_in_mini
    if state.reducer is None and world > 1:
        state.reducer = GradReducer(state.store)
    first_micro = state.micro_step == 0
    last_micro = state.micro_step + 1 == accum
    reduced = updated = False
    if USE_CUDA_GRAPH and world > 1 and last_micro:
        # gradient all-reduce overlapped with the backward pass (lax.pmean(grads), :150); without gradient clipping
        # (which needs the global norm) each segment's parameters are also updated as soon as its sum has arrived
        hook = None
        if (state.clip_grad == 0 and state.lr_decay >= 1.0 and state.reducer.stream is not None
                and state.optimizer == "adamw" and not (state.cfg.layerscale or state.cfg.dec_layerscale)):
            # (lamb's trust ratios are per leaf, LayerScale's d scale needs the summed gradients: reduce, then update)
            lr_now, cnt = lr_schedule(state.step, state.cfg), state.step
            hook = lambda ranges: eng.adamw_ranges(ranges, cnt, lr_now, grad_scale=1.0 / (accum * world))
            updated = True
        loss = eng.run_graph_overlapped(images, nz, first_micro, state.reducer, after_reduce=hook)
        reduced = True
    elif USE_CUDA_GRAPH:
        loss = eng.run_graph(images, nz, zero_grads=first_micro)
    else:
        loss = eng.forward(images, nz, want_grad=True)
        eng.backward(zero_grads=first_micro)
    state._release_prefetched(images)
    metrics_loss = loss.clone()
    if world > 1:
        state.reducer.mean_scalar(metrics_loss)                      # lax.pmean(metrics) (:144)
    state.micro_step += 1
    lr = lr_schedule(state.step, state.cfg)                          # hyperparams["learning_rate"] (:159)
    if state.micro_step == accum:                                    # :149-158
        if world > 1 and not reduced:
            state.reducer.all_reduce_grads()                         # lax.pmean(grads) (:150); 1/world below
        eng.layerscale_finalize()                                    # (no-op without LayerScale)
        if state.clip_grad > 0:                                      # chain(clip_by_global_norm, adamw) (:243-244)
            if not hasattr(state, "_sumsq"):
                state._sumsq = torch.zeros(1, dtype=torch.float32, device=state.store.device)
            ops.clip_by_global_norm(state.store.g, 1.0 / (accum * world), state.clip_grad, state._sumsq)
        if not updated:
            step_fn = eng.lamb_step if state.optimizer == "lamb" else eng.adamw_step
            step_fn(state.step, lr, grad_scale=1.0 / (accum * world), lr_decay=state.lr_decay)
        state.step += 1
        state.micro_step = 0
    return state, {"loss": metrics_loss[0], "learning_rate": lr}


def validation_step(state: TrainState, batch, noise=None) -> Dict[str, Any]:
    """src/pretraining.py:162-167: forward only with freshly split (not stored) RNGs;
    returns psum-ed {"loss", "num_samples"}."""
    images = _device_images(batch[0] if isinstance(batch, (tuple, list)) else batch)
    if noise is None:
        # freshly split, not stored (src/pretraining.py:164): one validation stream per rank, advanced per call
        if getattr(state, "_val_gen", None) is None:
            state._val_gen = np.random.Generator(np.random.Philox(key=state.noise_rng + 7919 * process_index() + 104729))
        noise = state._val_gen.random(state.noise_shape(images.shape[0]), dtype=np.float32)
    dev = state.store.device
    images = images.to(dev, non_blocking=True)
    nz = _device_noise(noise, dev)
    eng = state.engine(images.shape[0])
    loss = eng.forward(images, nz, want_grad=False).clone()
    num = torch.tensor([float(images.shape[0])], device=dev)
    if process_count() > 1:
        import torch.distributed as dist
        dist.all_reduce(loss)
        dist.all_reduce(num)
    return {"loss": loss[0], "num_samples": num[0]}
