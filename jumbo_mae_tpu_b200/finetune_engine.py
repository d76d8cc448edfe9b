"""Fine-tune / classification step on the encoder kernels of the pre-training engine.

``FinetuneModule.__call__`` + ``ViT.__call__`` with ``labels > 0`` (src/finetuning.py:84-106, src/modeling.py:245-274):
no masking, all ``N`` patches + ``J`` CLS tokens through the JumboLayers, final LayerNorm, ``LinearCLS`` on the
concatenated CLS tokens, softmax cross-entropy on smoothed / mixed labels, top-1 / top-5 flags.  Everything between
the patch rows and the logits is the schedule of ``engine.MAEEngine`` at ``T = J + N``; the two ends are
``jm_patchify`` / ``jm_patchify_mix`` and ``jm_softmax_ce``; the head is one ``jm_gemm`` over the strided CLS rows.

Built: forward, loss, metrics, backward (all parameter gradients), DropPath (``set_droppath``: per-image factors of
the three branches of every JumboLayer), AdamW with the layer-wise lr table (``adamw_step(..., lr_decay=0.75)``).
Not built (``create``-time options of the fine-tune recipe): BatchNorm in the head (linear probing), LARS / LAMB / SGD;
RNG draws (beta sample, permutation, box, DropPath keep bits) are caller data.
"""
from __future__ import annotations

import os
from typing import List, Optional

import torch

from . import _lib, ops
from .config import MAEConfig
from .engine import BF16, F32, LN_EPS, MAEEngine, _Block, fixed_sincos2d_embeddings, make_blocks
from .params import ParamStore

HEAD = "model/head/Dense_0"     # LinearCLS is nn.compact around one nn.Dense (src/modeling.py:214-219)


class FinetuneEngine(MAEEngine):
    """Buffers + kernel schedule of the classification step for one per-device batch size."""

    def __init__(self, cfg: MAEConfig, store: ParamStore, batch: int):
        _lib.require_device()
        cfg.validate()
        if cfg.labels <= 0:
            raise _lib.JmError("FinetuneEngine needs cfg.labels > 0 (classification mode)")
        if cfg.labels % 8:
            raise _lib.JmError("labels must be a multiple of 8 (16-byte rows of the bf16 logits gradient)")
        self.cfg, self.ps, self.B = cfg, store, batch
        dev = store.device
        self.dev = dev
        B, J, N, d, P, C = batch, cfg.num_cls_tokens, cfg.num_patches, cfg.dim, cfg.patch_dim, cfg.labels
        self.Te = self.Td = J + N
        self.Me = self.Md = B * self.Te
        self.Pp = (P + 7) // 8 * 8
        Me = self.Me
        e = lambda *s, dt=BF16: torch.empty(*s, dtype=dt, device=dev)

        self.wpe = fixed_sincos2d_embeddings(cfg.grid, cfg.grid, d).reshape(N, d).contiguous().to(dev)
        self.ids_all = torch.arange(N, dtype=torch.int32, device=dev)
        self.patches = e(Me, self.Pp)
        self.xa: List[torch.Tensor] = [e(Me, d, dt=F32) for _ in range(cfg.layers + 1)]
        self.remat = bool(getattr(cfg, "grad_ckpt", False))
        self.ls_enc, self.ls_dec = bool(getattr(cfg, "layerscale", False)), False
        self.enc_lowp = (d <= 1280 and cfg.posemb != "learnable" and os.environ.get("JM_ENC_LOWP", "1") != "0")
        self.enc = make_blocks(cfg.layers, Me, d, cfg.heads, B, self.Te, J, True, dev, self.remat)
        self.dec: List[_Block] = []          # no decoder in classification mode (activation_bytes walks both lists)
        L, Jd = cfg.layers, J * d
        self.cn16_all, self.gj_all = e(L * B, Jd), e(L * B, 4 * Jd)
        self.dhj16_all, self.dc16_all = e(L * B, 4 * Jd), e(L * B, Jd)
        for i, blk in enumerate(self.enc):
            blk.cn16, blk.gj = self.cn16_all[i * B:(i + 1) * B], self.gj_all[i * B:(i + 1) * B]
            blk.dhj16, blk.dc16 = self.dhj16_all[i * B:(i + 1) * B], self.dc16_all[i * B:(i + 1) * B]
        self.meanF, self.rstdF = e(Me, dt=F32), e(Me, dt=F32)
        self.enc_n = e(Me, d)
        # head + loss
        self.logits = e(B, C, dt=F32)
        self.dlogits, self.dlog16 = e(B, C, dt=F32), e(B, C)
        self.loss_rows, self.acc1, self.acc5 = e(B, dt=F32), e(B, dt=F32), e(B, dt=F32)
        self.metrics = torch.zeros(3, dtype=F32, device=dev)      # mean loss, acc1, acc5
        self.loss = self.metrics[0:1]
        # backward scratch
        w = Me * d
        self.dxA, self.dxB = e(w, dt=F32), e(w, dt=F32)
        self.dxA16, self.dxB16 = e(w), e(w)
        self.dh16, self.dn16, self.dqkv16, self.datt16 = e(4 * w), e(w), e(3 * w), e(w)
        self.delta_ws = None
        self.dcn16 = e(B, Jd)
        ln_ws = max(ops.layernorm_bwd_workspace(Me, d), ops.layernorm_bwd_workspace(B, Jd))
        self.ln_ws = torch.empty(ln_ws, dtype=torch.uint8, device=dev)
        self.drop_masks = None

    # ------------------------------------------------------------------ forward
    def forward(self, images: torch.Tensor, labels: torch.Tensor, *, label_smoothing: float = 0.0,
                perm: Optional[torch.Tensor] = None, ratio: float = 1.0, label_ratio: Optional[float] = None,
                xmask: Optional[torch.Tensor] = None, ymask: Optional[torch.Tensor] = None,
                want_grad: bool = True, grad_scale: float = 1.0,
                ratio_dev: Optional[torch.Tensor] = None) -> torch.Tensor:
        """images uint8 [B,3,H,W], labels int32 [B] or f32 [B, C] (device) -> mean loss (device tensor [1]);
        ``metrics`` = (loss, acc1, acc5) batch means, ``logits`` f32 [B, C].

        Training (det=False, src/finetuning.py:93-95): ``label_smoothing`` and, with ``perm``, Mixup
        (``ratio`` = the beta sample) or CutMix (``xmask`` / ``ymask`` = the box, ``label_ratio`` = the mean of the
        image mask).  ``ratio_dev`` (f32 [2] on the device: pixel ratio, label ratio) replaces the two host scalars, so
        that a captured graph can be replayed with new draws.  Evaluation: defaults."""
        cfg, ps, B = self.cfg, self.ps, self.B
        J, N, d, P, C, Te, Me = cfg.num_cls_tokens, cfg.num_patches, cfg.dim, cfg.patch_dim, cfg.labels, self.Te, self.Me
        if tuple(images.shape) != (B, 3, cfg.image_size, cfg.image_size):
            raise _lib.JmError(f"engine built for images {(B, 3, cfg.image_size, cfg.image_size)}, got {tuple(images.shape)}")
        self.images = images
        # pre-processing (+ Mixup / CutMix) + patch extraction (src/finetuning.py:87-88,95)
        if perm is None:
            ops.patchify(images, cfg.patch_size, None, N, Te, J, self.patches)
        else:
            ops.patchify_mix(images, cfg.patch_size, None, N, Te, J, self.patches, perm, ratio=ratio, xmask=xmask,
                             ymask=ymask, ratio_dev=None if ratio_dev is None else ratio_dev[0:1])
        x0 = self.xa[0]
        ops.fill_cls(x0, B, Te, J, d, ps.f("model/cls_tokens"))
        self._fwd_linear(self.patches, Me, "model/embed/wte", d, P, lda=self.Pp, out_f32=x0, pe=self._wpe_table(),
                         pe_idx=self.ids_all, row_period=Te, row_skip=J, row_mask_mode=2)
        for i in range(cfg.layers):
            self._block_fwd(f"model/layer_{i}", self.enc[i], self.xa[i], self.xa[i + 1], Me, d, cfg.heads, Te, True,
                            self._dp(i))
        ops.layernorm_fwd(self.xa[cfg.layers], ps.f("model/norm/scale"), ps.f("model/norm/bias"), LN_EPS,
                          self.enc_n, None, self.meanF, self.rstdF)
        # LinearCLS on x[:, :J].reshape(B, J*d) (src/modeling.py:272-274): the CLS rows are the first J of every
        # image's block, i.e. a [B, J*d] view with row stride Te*d
        ops.gemm(self.enc_n, self._W(f"{HEAD}/kernel"), B, C, J * d, b_mn=True, lda=Te * d,
                 ldb=ps.ld(f"{HEAD}/kernel"), bias=self._b(f"{HEAD}/bias"), out_f32=self.logits)
        soft = labels if labels.dim() == 2 else None
        idx = labels if labels.dim() == 1 else None
        ops._chk(soft, F32, "labels"), ops._chk(idx, torch.int32, "labels"), ops._chk(perm, torch.int32, "perm")
        lr = ratio if label_ratio is None else label_ratio
        lib = _lib.load()
        common = (float(label_smoothing), B, C, grad_scale / B, _lib.ptr(self.loss_rows), _lib.ptr(self.acc1),
                  _lib.ptr(self.acc5), _lib.ptr(self.metrics), _lib.ptr(self.dlogits) if want_grad else None,
                  _lib.stream_ptr())
        if ratio_dev is not None and perm is not None:
            _lib.check(lib.jm_softmax_ce_dev(_lib.ptr(self.logits), _lib.ptr(soft), _lib.ptr(idx), _lib.ptr(perm),
                                             _lib.ptr(ratio_dev[1:2]), *common))
        else:
            _lib.check(lib.jm_softmax_ce(_lib.ptr(self.logits), _lib.ptr(soft), _lib.ptr(idx), _lib.ptr(perm), lr,
                                         *common))
        return self.loss

    # ------------------------------------------------------------------ backward
    def segment_grad_ranges(self, k: int):
        if k == 0:
            return self.ps.ranges(lambda n: n.startswith("model/norm/") or n.startswith("model/head/"))
        return super().segment_grad_ranges(k)

    def backward_segment(self, k: int, zero_grads: bool = True) -> None:
        if k != 0:
            return super().backward_segment(k, zero_grads)
        cfg, ps, B = self.cfg, self.ps, self.B
        J, d, C, Te, Me = cfg.num_cls_tokens, cfg.dim, cfg.labels, self.Te, self.Me
        if zero_grads:
            ops.zero_(ps.g)
        # head: dW += cls^T dlogits, db += colsum, d cls = dlogits W^T into the CLS rows; patch rows get no gradient
        ops.add_bias_cast(self.dlogits, B, C, dst_bf16=self.dlog16)
        dn = self.dn16[:Me * d].view(Me, d)
        ops.zero_(dn)
        self._bwd_linear(self.enc_n, self.dlog16, B, HEAD, C, J * d, dX16=dn, x_ld=Te * d, ldo_bf16=Te * d)
        dx, dx16 = self.dxA[:Me * d].view(Me, d), self.dxA16[:Me * d].view(Me, d)
        ops.layernorm_bwd(dn, self.xa[cfg.layers], self.meanF, self.rstdF, ps.f("model/norm/scale"), None, dx, dx16,
                          ps.grad("model/norm/scale"), ps.grad("model/norm/bias"), self.ln_ws,
                          dcol=None if (self.drop_masks is not None or self.ls_enc) else ps.grad(f"model/layer_{cfg.layers - 1}/ff/w2/bias"),
                          dcol_period=Te, dcol_skip=J)

    # ------------------------------------------------------------------ CUDA graphs
    def _ensure_static_inputs(self) -> None:
        if hasattr(self, "images_buf"):
            return
        cfg, B = self.cfg, self.B
        self.images_buf = torch.empty(B, 3, cfg.image_size, cfg.image_size, dtype=torch.uint8, device=self.dev)
        self.labels_buf = torch.zeros(B, dtype=torch.int32, device=self.dev)
        self.perm_buf = torch.arange(B, dtype=torch.int32, device=self.dev).flip(0).contiguous()
        self.ratio_buf = torch.full((2,), 0.5, dtype=F32, device=self.dev)      # pixel ratio, label ratio
        self.xmask_buf = torch.zeros(cfg.image_size, dtype=torch.uint8, device=self.dev)
        self.ymask_buf = torch.zeros(cfg.image_size, dtype=torch.uint8, device=self.dev)
        self.graphs_by_mode, self.seg_graphs_by_mode, self._mode_kw = {}, {}, {}

    def capture_graphs(self, label_smoothing: float = 0.0, mixup: bool = False, mode: Optional[str] = None) -> None:
        """Records forward + backward of one micro-batch into two CUDA graphs (gradients zeroed first / accumulated)
        over static input buffers: images, class indices and, by ``mode`` ("plain", "mixup", "cutmix"), the batch
        permutation, the two Mixup / CutMix ratios (device scalars) and the CutMix box vectors -- everything a step
        draws is data in device memory, so the same graph replays with new draws (src/utils.py:92-111).  DropPath is
        captured as set by ``set_droppath`` (a persistent device buffer, or off).  Call while no gradient is being
        accumulated: the warm-up pass overwrites the gradient buffer."""
        mode = mode or ("mixup" if mixup else "plain")
        self._ensure_static_inputs()
        kw = dict(label_smoothing=label_smoothing)
        if mode != "plain":
            kw.update(perm=self.perm_buf, ratio_dev=self.ratio_buf)
        if mode == "cutmix":
            kw.update(xmask=self.xmask_buf, ymask=self.ymask_buf)
        self._mode_kw[mode] = kw
        self._graph_kw, self._default_mode = kw, mode
        self._graph_droppath = self.drop_masks is not None      # the graphs bake in whether DropPath scaling is on
        side = torch.cuda.Stream(device=self.dev)
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):                  # eager warm-up: one-time function attributes, module load
            self.images_buf.zero_()
            self.forward(self.images_buf, self.labels_buf, **kw)
            self.backward(zero_grads=True)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        graphs = {}
        lib = _lib.load()
        for zero in (True, False):
            g = torch.cuda.CUDAGraph()
            l0 = lib.jm_launch_count()
            with torch.cuda.graph(g):
                self.forward(self.images_buf, self.labels_buf, **kw)
                self.backward(zero_grads=zero)
            self.graph_launches = lib.jm_launch_count() - l0
            graphs[zero] = g
        self.graphs_by_mode[mode] = graphs
        self.graphs = graphs
        ops.zero_(self.ps.g)

    def _stage_inputs(self, images, labels, perm, ratios, xmask, ymask) -> None:
        self.images_buf.copy_(images, non_blocking=True)
        self.labels_buf.copy_(labels, non_blocking=True)
        if perm is not None:
            self.perm_buf.copy_(perm, non_blocking=True)
        if ratios is not None:
            self.ratio_buf.copy_(torch.tensor([float(ratios[0]), float(ratios[1])], dtype=F32), non_blocking=False)
        if xmask is not None:
            self.xmask_buf.copy_(xmask, non_blocking=True)
            self.ymask_buf.copy_(ymask, non_blocking=True)

    def run_graph(self, images: torch.Tensor, labels: torch.Tensor, zero_grads: bool,
                  perm: Optional[torch.Tensor] = None, mode: Optional[str] = None, ratios=None,
                  xmask: Optional[torch.Tensor] = None, ymask: Optional[torch.Tensor] = None) -> torch.Tensor:
        """forward + backward through the captured graph of ``mode``; images may be a pinned host tensor; ``ratios`` =
        (pixel ratio, label ratio) of this step's Mixup / CutMix draw."""
        if not hasattr(self, "graphs"):
            self.capture_graphs()
        mode = mode or self._default_mode
        self._stage_inputs(images, labels, perm, ratios, xmask, ymask)
        self.graphs_by_mode[mode][bool(zero_grads)].replay()
        return self.loss

    def capture_segment_graphs(self, mode: Optional[str] = None) -> None:
        """Graphs for a data-parallel rank's LAST micro-step of an optimizer step: forward + head / final-norm backward,
        then one graph per encoder segment (the pre-training engine's plan), so that the gradient all-reduce of a
        segment runs while the next one computes."""
        if not hasattr(self, "graphs"):
            self.capture_graphs()
        mode = mode or self._default_mode
        kw = self._mode_kw[mode]
        out = {}
        for zero in (True, False):
            graphs = []
            for k in range(self.num_bwd_segments()):
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    if k == 0:
                        self.forward(self.images_buf, self.labels_buf, **kw)
                    self.backward_segment(k, zero_grads=zero)
                graphs.append(g)
            out[zero] = graphs
        self.seg_graphs_by_mode[mode] = out
        self.seg_graphs = out
        ops.zero_(self.ps.g)

    def run_graph_overlapped(self, images: torch.Tensor, labels: torch.Tensor, zero_grads: bool, reducer,
                             perm: Optional[torch.Tensor] = None, after_reduce=None, mode: Optional[str] = None,
                             ratios=None, xmask: Optional[torch.Tensor] = None,
                             ymask: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Last micro-step of a data-parallel optimizer step: every backward segment is followed by the all-reduce of
        the gradients it finalised, on the reducer's stream, overlapping the next segment."""
        mode = mode or getattr(self, "_default_mode", None)
        if mode is None or mode not in getattr(self, "seg_graphs_by_mode", {}):
            self.capture_segment_graphs(mode)
            mode = mode or self._default_mode
        self._stage_inputs(images, labels, perm, ratios, xmask, ymask)
        for k, g in enumerate(self.seg_graphs_by_mode[mode][bool(zero_grads)]):
            g.replay()
            ev = torch.cuda.Event()
            ev.record()
            ranges = self.segment_grad_ranges(k)
            reducer.all_reduce_ranges(ranges, after_event=ev)
            if after_reduce is not None and reducer.stream is not None:
                with torch.cuda.stream(reducer.stream):
                    after_reduce(ranges)
        reducer.join()
        return self.loss
