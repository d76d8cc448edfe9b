"""The MAE pre-training step as an explicit kernel schedule.

This is what XLA executes for the reference's ``training_step`` (src/pretraining.py:125-159):
PretrainModule.__call__ (src/pretraining.py:87-122) forward, its hand-derived backward,
and the optax AdamW update.  Every arithmetic op is one of the sm_100a kernels behind the
C ABI (include/jumbo_mae_b200.h); this module only owns buffers and launch order.

Token layout.  Encoder activations are one [B*Te, d] buffer per layer (Te = J + keep), CLS
rows first in each image block.  The Jumbo CLS view [B, J*d] (src/modeling.py:194-197) is
then the same memory with row stride Te*d, so split / reshape / concat move no data: the
Jumbo FFN's GEMM epilogue stores straight into the CLS rows.  The patch FFN runs on all Te
rows with the CLS rows predicated off in the epilogue (6 % of one GEMM pair).

Precision.  bf16 GEMM / attention operands, f32 accumulation, f32 residual stream,
LayerNorm and softmax statistics, loss and optimizer state (the reference is f32
throughout; BASELINE.json's "bf16" is this compute dtype).
"""
from __future__ import annotations

import copy
import os

import math
from typing import Dict, List, Optional

import torch

from . import _lib, ops
from .config import MAEConfig
from .params import ParamStore, get_layer_index

F32, BF16 = torch.float32, torch.bfloat16
LN_EPS = 1e-6   # flax.linen.LayerNorm default
# jm_gemm activation codes used by the step: the forward FFN GEMM saves gelu'(h) (not h) next to gelu(h), so that
# the backward epilogue is a single multiply
LOSS_EPI = 5   # jm_gemm act: loss epilogue of the pixel head
GELU_FWD, GELU_BWD = (3, 4) if os.environ.get("JM_GELU_SAVED", "1") != "0" else (1, 2)


def fixed_sincos2d_embeddings(ncols: int, nrows: int, dim: int) -> torch.Tensor:
    """src/utils.py:114-121 (host-side constant table, computed once in f32 on the CPU)."""
    freqs = 1.0 / (10000.0 ** torch.linspace(0, 1, dim // 4, dtype=F32))
    x = torch.outer(torch.arange(0, nrows, dtype=F32), freqs)
    y = torch.outer(torch.arange(0, ncols, dtype=F32), freqs)
    x = x[None, :, :].expand(ncols, nrows, dim // 4)
    y = y[:, None, :].expand(ncols, nrows, dim // 4)
    return torch.cat((torch.sin(x), torch.cos(x), torch.sin(y), torch.cos(y)), dim=2)


class _Block:
    """Saved activations of one transformer block (encoder JumboLayer or decoder ViTLayer)."""

    def __init__(self, M: int, dim: int, heads: int, B: int, T: int, J: int, jumbo: bool, dev):
        e = lambda *s, dt=BF16: torch.empty(*s, dtype=dt, device=dev)
        self.mean1, self.rstd1 = e(M, dt=F32), e(M, dt=F32)
        self.mean2, self.rstd2 = e(M, dt=F32), e(M, dt=F32)
        self.xn1, self.xn2 = e(M, dim), e(M, dim)
        self.qkv = e(M, 3 * dim)
        self.lse = e(B, heads, T, dt=F32)
        self.att = e(M, dim)
        self.xb = e(M, dim, dt=F32)            # residual stream after the attention branch
        self.h, self.g = e(M, 4 * dim), e(M, 4 * dim)   # h: gelu'(pre-activation), g: gelu(pre-activation)
        if jumbo:
            Jd = J * dim
            self.mean3, self.rstd3 = e(B, dt=F32), e(B, dt=F32)
            self.cn = e(B, Jd, dt=F32)
            self.hj = e(B, 4 * Jd)              # (cn16, gj, dhj16, dc16: slices of the engine's all-layer buffers)


def make_blocks(count: int, M: int, dim: int, heads: int, B: int, T: int, J: int, jumbo: bool, dev, share: bool):
    """The saved activations of ``count`` blocks.  ``share`` (grad_ckpt, the reference's nn.remat of every layer,
    src/modeling.py:232,280): all blocks use ONE set of buffers -- the forward pass keeps only the block inputs
    (the f32 residual stream) and the backward pass recomputes a block's activations right before it needs them."""
    if not share or count <= 1:
        return [_Block(M, dim, heads, B, T, J, jumbo, dev) for _ in range(count)]
    first = _Block(M, dim, heads, B, T, J, jumbo, dev)
    return [first] + [copy.copy(first) for _ in range(count - 1)]     # (per-layer attributes are set on the copies)


class MAEEngine:
    """Buffers + kernel schedule for one per-device batch size."""

    def __init__(self, cfg: MAEConfig, store: ParamStore, batch: int, fused_head: Optional[bool] = None):
        _lib.require_device()
        cfg.validate()
        self.cfg, self.ps, self.B = cfg, store, batch
        dev = store.device
        self.dev = dev
        B, J, n, N = batch, cfg.num_cls_tokens, cfg.keep, cfg.num_patches
        d, dd, P = cfg.dim, cfg.dec_dim, cfg.patch_dim
        self.Te, self.Td = J + n, J + N
        self.Me, self.Md = B * self.Te, B * self.Td
        self.Pp = (P + 7) // 8 * 8
        e = lambda *s, dt=BF16: torch.empty(*s, dtype=dt, device=dev)

        self.wpe = fixed_sincos2d_embeddings(cfg.grid, cfg.grid, d).reshape(N, d).contiguous().to(dev)
        self.dec_wpe = fixed_sincos2d_embeddings(cfg.grid, cfg.grid, dd).reshape(N, dd).contiguous().to(dev)

        # one permutation for the whole batch (the reference), or one row per image (cfg.per_image_masking)
        self.per_image = bool(getattr(cfg, "per_image_masking", False)) and cfg.labels <= 0
        ishape = (B, N) if self.per_image else (N,)
        self.ids_shuffle = e(*ishape, dt=torch.int32)
        self.ids_restore = e(*ishape, dt=torch.int32)
        self.mask_row = e(*ishape, dt=F32)
        self.patches = e(self.Me, self.Pp)
        # encoder
        self.xa: List[torch.Tensor] = [e(self.Me, d, dt=F32) for _ in range(cfg.layers + 1)]
        self.ls_enc = bool(getattr(cfg, "layerscale", False))        # LayerScale on the encoder / decoder branches
        self.ls_dec = bool(getattr(cfg, "dec_layerscale", False))
        self.remat = bool(getattr(cfg, "grad_ckpt", False))
        self.enc = make_blocks(cfg.layers, self.Me, d, cfg.heads, B, self.Te, J, True, dev, self.remat)
        # The Jumbo MLP is ONE set of weights applied in every layer (src/modeling.py:234): its inputs and
        # output gradients are kept layer after layer in [layers * B, .] buffers, so that each weight gradient
        # is a single GEMM over K = layers * B at the end of the backward pass instead of one read-modify-write
        # of the 85 MB gradient per layer.
        L, Jd = cfg.layers, J * d
        self.cn16_all, self.gj_all = e(L * B, Jd), e(L * B, 4 * Jd)
        self.dhj16_all, self.dc16_all = e(L * B, 4 * Jd), e(L * B, Jd)
        for i, blk in enumerate(self.enc):
            blk.cn16, blk.gj = self.cn16_all[i * B:(i + 1) * B], self.gj_all[i * B:(i + 1) * B]
            blk.dhj16, blk.dc16 = self.dhj16_all[i * B:(i + 1) * B], self.dc16_all[i * B:(i + 1) * B]
        self.meanF, self.rstdF = e(self.Me, dt=F32), e(self.Me, dt=F32)
        self.enc_n = e(self.Me, d)
        self.y = e(self.Me, dd, dt=F32)
        # decoder
        self.za: List[torch.Tensor] = [e(self.Md, dd, dt=F32) for _ in range(cfg.dec_layers + 1)]
        self.dec = make_blocks(cfg.dec_layers, self.Md, dd, cfg.dec_heads, B, self.Td, J, False, dev, self.remat)
        # Pixel head + loss (src/pretraining.py:110-120).  Fused (default for dec_dim <= 512): dec_norm and the head run on
        # the N - keep masked rows of every image only, gathered into compact [B * n_mask, .] buffers, and the head GEMM's
        # epilogue forms the squared error against the byte targets and its gradient -- the f32 prediction never exists.
        # Unfused (fused_head=False / JM_HEAD_FUSED=0): head on all rows -> pred f32 -> jm_patch_loss.
        if fused_head is None:
            fused_head = os.environ.get("JM_HEAD_FUSED", "1") != "0"
        self.fused_head = bool(fused_head) and dd <= 512 and cfg.labels <= 0
        if self.per_image and not self.fused_head:
            raise NotImplementedError("per-image masking is built on the fused pixel head (dec_dim <= 512, fused_head=True)")
        self.n_mask = N - n
        self.Mh = B * self.n_mask if self.fused_head else self.Md
        self.meanD, self.rstdD = e(self.Mh, dt=F32), e(self.Mh, dt=F32)
        self.dec_n = e(self.Mh, dd)
        self.loss = torch.zeros(1, dtype=F32, device=dev)
        if self.fused_head:
            self.pred = None
            self.Pt = (P + 15) // 16 * 16
            self.tgt_u8 = e(self.Mh, self.Pt, dt=torch.uint8)
            self.tgt_stats = e(self.Mh, 2, dt=F32) if cfg.norm_pix_loss else None
            self.row_map = e(self.Mh, dt=torch.int32)
            self.row_loss = e(self.Mh, (P + 15) // 16, dt=F32)
            self.loss_ws = torch.empty(ops.row_loss_workspace(), dtype=torch.uint8, device=dev)
        else:
            self.pred = e(self.Md, P, dt=F32)
            self.loss_ws = torch.empty(ops.patch_loss_workspace(B, N), dtype=torch.uint8, device=dev)
        # backward scratch, shared by both stacks (sized for the larger one)
        Mx = max(self.Me, self.Md)
        wmax = max(self.Me * d, self.Md * dd)
        self.dpred = torch.zeros(self.Mh, self.Pp, dtype=BF16, device=dev)     # pad columns stay zero
        self.dxA, self.dxB = e(wmax, dt=F32), e(wmax, dt=F32)
        self.dxA16, self.dxB16 = e(wmax), e(wmax)
        self.dh16 = e(4 * wmax)
        self.dn16 = e(wmax)
        self.dqkv16 = e(3 * wmax)
        self.datt16 = e(wmax)
        self.delta_ws = e(B * max(cfg.heads * self.Te, cfg.dec_heads * self.Td), dt=F32)
        self.dy16 = e(self.Me, dd)
        self.dcn16 = e(B, J * d)
        ln_ws = max(ops.layernorm_bwd_workspace(Mx, max(d, dd)), ops.layernorm_bwd_workspace(B, J * d))
        self.ln_ws = torch.empty(ln_ws, dtype=torch.uint8, device=dev)
        self.drop_masks = None
        # decoder residual-gradient stream in bf16 alone (LayerNorm backward's dim <= 512 kernel); JM_DEC_LOWP=0: f32
        self.dec_lowp = dd <= 512 and os.environ.get("JM_DEC_LOWP", "1") != "0"
        # the same for the encoder: bf16 stream for every row, plus an f32 copy of the CLS rows only (the Jumbo branch
        # accumulates its split-K gradient into them in f32); a learnable wpe reads f32 patch rows -> f32 stream
        self.enc_lowp = (d <= 1280 and cfg.posemb != "learnable" and os.environ.get("JM_ENC_LOWP", "1") != "0")

    # ------------------------------------------------------------------ DropPath (src/modeling.py:157,181-183)
    def set_droppath(self, masks: Optional[torch.Tensor]) -> None:
        """Per-image DropPath factors keep / (1 - rate) of every residual branch: f32 [blocks, 3, B] on the device
        (encoder layers first, then decoder layers; branch 0 attention, 1 patch FFN, 2 Jumbo MLP), or None for the
        deterministic / rate-0 path.  The buffer is read by the kernels at run time, so a captured CUDA graph keeps
        working when its CONTENTS change; switching between None and a tensor needs a re-capture."""
        if masks is not None:
            nblk = self.cfg.layers + len(self.dec)
            if tuple(masks.shape) != (nblk, 3, self.B) or masks.dtype != F32 or not masks.is_cuda:
                raise _lib.JmError(f"drop masks must be f32 [{nblk}, 3, {self.B}] on the device")
        self.drop_masks = masks

    def _wpe_table(self) -> torch.Tensor:
        """[N, d] position embedding added to the patch tokens: the fixed sincos table, or the trained parameter
        model/embed/wpe with --posemb learnable (src/modeling.py:115-120)."""
        if self.cfg.posemb == "learnable":
            return self.ps.f("model/embed/wpe").view(self.cfg.num_patches, self.cfg.dim)
        return self.wpe

    def _dp(self, block_index: int):
        return None if self.drop_masks is None else self.drop_masks[block_index]

    # ------------------------------------------------------------------ helpers
    def _W(self, name):
        return self.ps.w(name)

    def _b(self, name):
        return self.ps.f(name)

    def _fwd_linear(self, A, M, name, N, K, **epi):
        """out = epilogue(A[M,K] @ kernel[K,N] + bias): a Flax Dense (kernel is [in, out])."""
        ops.gemm(A, self._W(f"{name}/kernel"), M, N, K, b_mn=True, ldb=self.ps.ld(f"{name}/kernel"),
                 bias=self._b(f"{name}/bias"), **epi)

    def _bwd_linear(self, X, dY, M, name, N, K, dX16=None, x_ld=None, dy_ld=None, bias_rows=None, bias=True,
                    **epi):
        """Gradients of Y = X[M,K] @ W[K,N] + b given dY (bf16):
        dW += X^T dY, db += colsum(dY) (bias=False: the producer of dY already accumulated it),
        and optionally dX16 = epilogue(dY @ W^T)."""
        ldw = self.ps.ld(f"{name}/kernel")
        ops.gemm(X, dY, K, N, M, a_mn=True, b_mn=True, lda=x_ld, ldb=dy_ld,
                 out_f32=self.ps.grad(f"{name}/kernel"), ldo_f32=ldw, accumulate=True)
        if bias:
            period, skip = bias_rows if bias_rows else (0, 0)
            ops.colsum(dY, M, N, self.ps.grad(f"{name}/bias"), row_period=period, row_skip=skip, ldx=dy_ld)
        if dX16 is not None or epi:
            ops.gemm(dY, self._W(f"{name}/kernel"), M, K, N, b_mn=False, lda=dy_ld, ldb=ldw, out_bf16=dX16, **epi)

    # ------------------------------------------------------------------ forward
    def forward(self, images: torch.Tensor, noise: torch.Tensor, want_grad: bool = True,
                grad_scale: float = 1.0) -> torch.Tensor:
        """images uint8 [B,3,H,W] (device), noise f32 [N] (device) -> loss (device scalar tensor [1]).
        PretrainModule.__call__ (src/pretraining.py:87-122) = encoder, decoder input assembly, decoder, head + loss."""
        self.forward_encoder(images, noise)
        self.forward_decoder_input()
        self.forward_decoder_stack()
        return self.forward_head_and_loss(want_grad, grad_scale)

    def forward_encoder(self, images: torch.Tensor, noise: torch.Tensor) -> None:
        """ViT.__call__ in MAE mode (src/modeling.py:245-267) incl. the pixel normalisation of
        src/pretraining.py:90-91: fills ids_restore / mask_row, xa[0..L] and the normalised output enc_n."""
        cfg, ps, B = self.cfg, self.ps, self.B
        J, n, N, d, P = cfg.num_cls_tokens, cfg.keep, cfg.num_patches, cfg.dim, cfg.patch_dim
        Te, Me = self.Te, self.Me
        if tuple(images.shape) != (B, 3, cfg.image_size, cfg.image_size):
            raise _lib.JmError(f"engine built for images {(B, 3, cfg.image_size, cfg.image_size)}, got {tuple(images.shape)}")
        self.images = images
        c = _lib.check
        lib = _lib.load()
        st = _lib.stream_ptr()
        # random_masking (src/utils_mae.py:88-96) from host-supplied noise
        if tuple(noise.shape) != tuple(self.mask_row.shape):
            raise _lib.JmError(f"noise must have shape {tuple(self.mask_row.shape)}, got {tuple(noise.shape)}")
        if self.per_image:
            c(lib.jm_mask_sort_batched(noise.data_ptr(), B, N, n, self.ids_shuffle.data_ptr(), self.ids_restore.data_ptr(),
                                       self.mask_row.data_ptr(), st))
        else:
            c(lib.jm_mask_sort(noise.data_ptr(), N, n, self.ids_shuffle.data_ptr(), self.ids_restore.data_ptr(),
                               self.mask_row.data_ptr(), st))
        # pre-processing + extract kept patches (src/pretraining.py:90-91, modeling.py:246,256)
        ops.patchify(images, cfg.patch_size, self.ids_shuffle, n, Te, J, self.patches)
        x0 = self.xa[0]
        ops.fill_cls(x0, B, Te, J, d, ps.f("model/cls_tokens"))
        self._fwd_linear(self.patches, Me, "model/embed/wte", d, P, lda=self.Pp, out_f32=x0, pe=self._wpe_table(),
                         pe_idx=self.ids_shuffle, row_period=Te, row_skip=J, row_mask_mode=2)
        for i in range(cfg.layers):
            self._block_fwd(f"model/layer_{i}", self.enc[i], self.xa[i], self.xa[i + 1], Me, d, cfg.heads, Te, True,
                            self._dp(i))
        ops.layernorm_fwd(self.xa[cfg.layers], ps.f("model/norm/scale"), ps.f("model/norm/bias"), LN_EPS,
                          self.enc_n, None, self.meanF, self.rstdF)

    def forward_decoder_input(self) -> None:
        """decoder_proj + mask tokens + unshuffle + dec_wpe (src/pretraining.py:95-106, modeling.py:289-292) -> za[0]."""
        cfg, ps, B = self.cfg, self.ps, self.B
        J, n, N, d, dd = cfg.num_cls_tokens, cfg.keep, cfg.num_patches, cfg.dim, cfg.dec_dim
        self._fwd_linear(self.enc_n, self.Me, "decoder_proj", dd, d, out_f32=self.y)
        ops.unshuffle_fwd(self.y, self.ids_restore, ps.f("image_mask_embedding"), self.dec_wpe, B, N, n, J, dd,
                          self.za[0])

    def forward_decoder_stack(self) -> None:
        """MAEDecoder's blocks and final norm (src/modeling.py:293-298) on za[0] (which already carries dec_wpe)."""
        cfg, ps = self.cfg, self.ps
        for i in range(cfg.dec_layers):
            self._block_fwd(f"decoder_model/dec_layer_{i}", self.dec[i], self.za[i], self.za[i + 1], self.Md, cfg.dec_dim,
                            cfg.dec_heads, self.Td, False, self._dp(cfg.layers + i))
        if self.fused_head:
            return                       # dec_norm runs on the masked rows only, in forward_head_and_loss
        ops.layernorm_fwd(self.za[cfg.dec_layers], ps.f("decoder_model/dec_norm/scale"),
                          ps.f("decoder_model/dec_norm/bias"), LN_EPS, self.dec_n, None, self.meanD, self.rstdD)

    def forward_head_and_loss(self, want_grad: bool = True, grad_scale: float = 1.0) -> torch.Tensor:
        """decoder_image_output + extract_patches + (norm-pix) + patch_mse_loss (src/pretraining.py:110-120)."""
        cfg, ps = self.cfg, self.ps
        if self.fused_head:
            P, n_mask = cfg.patch_dim, self.n_mask
            ops.masked_targets(self.images, cfg.patch_size, self.ids_shuffle[..., cfg.keep:], cfg.num_cls_tokens,
                               cfg.norm_pix_loss, self.tgt_u8, self.tgt_stats, self.row_map)
            ops.layernorm_fwd_rows(self.za[cfg.dec_layers], self.row_map, ps.f("decoder_model/dec_norm/scale"),
                                   ps.f("decoder_model/dec_norm/bias"), LN_EPS, self.dec_n, self.meanD, self.rstdD)
            denom = float(P) * n_mask * self.B            # mean over images of mean over masked patches of mean_c
            ops.gemm(self.dec_n, self._W("decoder_image_output/kernel"), self.Mh, P, cfg.dec_dim, b_mn=True,
                     ldb=ps.ld("decoder_image_output/kernel"), bias=ps.f("decoder_image_output/bias"), act=LOSS_EPI,
                     out_bf16=self.dpred if want_grad else None, ldo_bf16=self.Pp, loss_target=self.tgt_u8,
                     loss_stats=self.tgt_stats, row_loss=self.row_loss, loss_scale=2.0 * grad_scale / denom)
            ops.row_loss_finalize(self.row_loss, 1.0 / denom, self.loss, self.loss_ws)
            return self.loss
        # pixel head on every token row; CLS / visible rows are ignored by the loss
        self._fwd_linear(self.dec_n, self.Md, "decoder_image_output", cfg.patch_dim, cfg.dec_dim, out_f32=self.pred)
        ops.patch_loss(self.pred, self.images, self.mask_row, cfg.patch_size, cfg.num_cls_tokens, cfg.norm_pix_loss,
                       grad_scale, self.loss, self.dpred if want_grad else None, self.loss_ws)
        return self.loss

    def _block_fwd(self, name, s: _Block, x_in, x_out, M, dim, heads, T, jumbo, dp=None):
        """One JumboLayer / ViTLayer forward.  ``dp``: [3, B] DropPath factors of the block's branches or None."""
        ps, B, J = self.ps, self.B, self.cfg.num_cls_tokens
        hd = dim // heads
        rs = (lambda k: dict(row_scale=dp[k], row_scale_period=T)) if dp is not None else (lambda k: {})
        ls = self.ls_enc if jumbo else self.ls_dec
        # LayerScale (src/modeling.py:159-166,185-201): the branch output times scale_k, per column, in the same epilogue
        cv = (lambda k: dict(col_vec=ps.f(f"{name}/scale{k}"))) if ls else (lambda k: {})
        # x_b = x + drop_path1(attn(LN1(x)))            (src/modeling.py:165 / :192)
        ops.layernorm_fwd(x_in, ps.f(f"{name}/norm1/scale"), ps.f(f"{name}/norm1/bias"), LN_EPS, s.xn1, None,
                          s.mean1, s.rstd1)
        self._fwd_linear(s.xn1, M, f"{name}/attn/wqkv", 3 * dim, dim, out_bf16=s.qkv, col_scale=hd ** -0.5,
                         col_scale_n=dim)
        ops.attention_fwd(s.qkv, B, T, heads, hd, s.att, s.lse)
        self._fwd_linear(s.att, M, f"{name}/attn/wo", dim, dim, out_f32=s.xb, residual=x_in, **rs(0), **cv(1))
        period, skip = (T, J) if jumbo else (0, 0)
        # patch tokens: x = x_b + drop_path2(ff(LN2(x_b)))       (src/modeling.py:166 / :201)
        ops.layernorm_fwd(s.xb, ps.f(f"{name}/norm2/scale"), ps.f(f"{name}/norm2/bias"), LN_EPS, s.xn2, None,
                          s.mean2, s.rstd2)
        self._fwd_linear(s.xn2, M, f"{name}/ff/w1", 4 * dim, dim, out_bf16=s.h, out2_bf16=s.g, act=GELU_FWD,
                         row_period=period, row_skip=skip, row_mask_mode=1 if jumbo else 0)
        # (every row is written, so that the output tiles can leave by TMA; the CLS rows of a Jumbo block get
        #  x_b + bias here -- their h/g rows are zero -- and are overwritten by the Jumbo MLP just below)
        self._fwd_linear(s.g, M, f"{name}/ff/w2", dim, 4 * dim, out_f32=x_out, residual=s.xb, **rs(1), **cv(2))
        if jumbo:
            Jd = J * dim
            # c = LN3(cls concat); c = c + drop_path3(jumbo_mlp(c))      (src/modeling.py:194-199)
            ops.layernorm_fwd(s.xb, ps.f(f"{name}/norm3/scale"), ps.f(f"{name}/norm3/bias"), LN_EPS, s.cn16, s.cn,
                              s.mean3, s.rstd3, rows=B, dim=Jd, ldx=T * dim, ldy=Jd, ldyf=Jd)
            self._fwd_linear(s.cn16, B, "model/jumbo_mlp/w1", 4 * Jd, Jd, out_bf16=s.hj, out2_bf16=s.gj, act=GELU_FWD)
            if dp is None and not ls:
                # w2 has few output tiles and a long K: seed the CLS rows with c + bias, then split-K accumulate
                ops.add_bias_cast(s.cn, B, Jd, bias=ps.f("model/jumbo_mlp/w2/bias"), dst_f32=x_out, ldd=T * dim)
                ops.gemm(s.gj, self._W("model/jumbo_mlp/w2/kernel"), B, Jd, 4 * Jd, b_mn=True,
                         ldb=ps.ld("model/jumbo_mlp/w2/kernel"), out_f32=x_out, ldo_f32=T * dim, accumulate=True)
            else:
                # with DropPath / LayerScale the branch (bias included) is scaled before c is added: one pass over K
                ops.gemm(s.gj, self._W("model/jumbo_mlp/w2/kernel"), B, Jd, 4 * Jd, b_mn=True,
                         ldb=ps.ld("model/jumbo_mlp/w2/kernel"), bias=ps.f("model/jumbo_mlp/w2/bias"), residual=s.cn,
                         ldr=Jd, out_f32=x_out, ldo_f32=T * dim, **(dict(row_scale=dp[2], row_scale_period=1) if dp is not None else {}),
                         **cv(3))

    # ------------------------------------------------------------------ backward
    def backward(self, zero_grads: bool = True) -> None:
        """Accumulates d loss / d params into store.g (after ``forward(want_grad=True)``)."""
        for k in range(self.num_bwd_segments()):
            self.backward_segment(k, zero_grads)

    # The backward pass in segments whose parameter gradients are FINAL when the segment ends, so that a
    # data-parallel caller can start all-reducing them while the next segment computes (the overlap the
    # reference gets from XLA scheduling lax.pmean inside the step, src/pretraining.py:150).
    ENC_LAYERS_PER_SEGMENT = 3

    def _segment_plan(self):
        """Decoder segment, encoder segments of ENC_LAYERS_PER_SEGMENT layers (layers L-1 .. 1), then the shared
        Jumbo MLP on its own and a tail.  The Jumbo gradients (the largest leaves, 28 % of ViT-B) are final as soon
        as layer 0's Jumbo branch has run: that branch and the batched Jumbo weight-gradient GEMMs form one
        segment, so the all-reduce of the layers before it overlaps those GEMMs and the Jumbo all-reduce overlaps
        the tail (layer 0's patch-FFN and attention branches, the patch embedding)."""
        upper = list(reversed(range(1, self.cfg.layers)))
        n = self.ENC_LAYERS_PER_SEGMENT
        return [("dec", [])] + [("enc", upper[i:i + n]) for i in range(0, len(upper), n)] + [("jumbo", [0]), ("tail", [0])]

    def num_bwd_segments(self) -> int:
        return len(self._segment_plan())

    def _enc_segment_layers(self, k: int):
        """Encoder layers (descending) that backward segment k >= 1 works on."""
        return self._segment_plan()[k][1]

    def segment_grad_ranges(self, k: int):
        """[start, end) ranges of store.g that are final once backward segment k has run."""
        kind, layers = self._segment_plan()[k]
        if kind == "dec":
            sel = lambda n: not n.startswith("model/") or n.startswith("model/norm/")
        elif kind == "jumbo":
            sel = lambda n: n.startswith("model/jumbo_mlp/")
        elif kind == "tail":
            sel = lambda n: (n.startswith("model/layer_0/") or n.startswith("model/embed/")
                             or n.startswith("model/cls_tokens"))
        else:
            # a layer's ff/w2 bias is produced by the block ABOVE it (LayerNorm-backward column sums), hence
            # final one block earlier; it is simply reduced together with its own layer's leaves
            sel = lambda n, layers=layers: any(n.startswith(f"model/layer_{i}/") for i in layers)
        return self.ps.ranges(sel)

    def backward_segment(self, k: int, zero_grads: bool = True) -> None:
        cfg, ps, B = self.cfg, self.ps, self.B
        J, n, N, d, dd, P = cfg.num_cls_tokens, cfg.keep, cfg.num_patches, cfg.dim, cfg.dec_dim, cfg.patch_dim
        Te, Td, Me, Md, Pp = self.Te, self.Td, self.Me, self.Md, self.Pp
        view = lambda buf, M, w: buf[:M * w].view(M, w)
        if k == 0:
            if zero_grads:
                ops.zero_(ps.g)
            # ---- pixel head
            Mh = self.Mh
            dn = view(self.dn16, Mh, dd)
            self._bwd_linear(self.dec_n, self.dpred, Mh, "decoder_image_output", P, dd, dX16=dn, dy_ld=Pp)
            dx, dx16 = view(self.dxA, Md, dd), view(self.dxA16, Md, dd)
            # (every LayerNorm backward also emits the column sums of the dx it produces: that dx is the
            #  output gradient of the previous block's ff/w2, i.e. its bias gradient)
            dp_on = self.drop_masks is not None     # then every block sums its ff/w2 bias from its scaled gradient copy
            # The decoder's residual-gradient stream is carried in bf16 alone: its LayerNorm backwards move 10 instead
            # of 16 bytes per element.  (A stated deviation: the reference's gradients are f32 end to end; every
            # gradient leaf stays inside the parity tolerances with it, JM_DEC_LOWP=0 restores the f32 stream.)  Only
            # the gradient that leaves the decoder (layer 0's input, scattered back to the encoder's tokens) is also
            # written in f32.
            lowp = self.dec_lowp and not dp_on and not self.ls_dec
            dcol = None if (dp_on or self.ls_dec) else ps.grad(f"decoder_model/dec_layer_{cfg.dec_layers - 1}/ff/w2/bias")
            if self.fused_head:          # the gradient exists on the masked rows only: scattered into a zeroed dx
                ops.layernorm_bwd_rows(dn, self.za[cfg.dec_layers], self.row_map, self.meanD, self.rstdD,
                                       ps.f("decoder_model/dec_norm/scale"), None if lowp else dx, dx16,
                                       ps.grad("decoder_model/dec_norm/scale"), ps.grad("decoder_model/dec_norm/bias"),
                                       self.ln_ws, dcol=dcol)
            else:
                ops.layernorm_bwd(dn, self.za[cfg.dec_layers], self.meanD, self.rstdD,
                                  ps.f("decoder_model/dec_norm/scale"), None, None if lowp else dx, dx16,
                                  ps.grad("decoder_model/dec_norm/scale"), ps.grad("decoder_model/dec_norm/bias"),
                                  self.ln_ws, dcol=dcol)
            # ---- decoder blocks
            for i in reversed(range(cfg.dec_layers)):
                below = ps.grad(f"decoder_model/dec_layer_{i - 1}/ff/w2/bias") if i > 0 else None
                if self.remat:
                    self._block_fwd(f"decoder_model/dec_layer_{i}", self.dec[i], self.za[i], self.za[i + 1], Md, dd,
                                    cfg.dec_heads, Td, False, self._dp(cfg.layers + i))
                self._block_bwd(f"decoder_model/dec_layer_{i}", self.dec[i], self.za[i], Md, dd, cfg.dec_heads, Td,
                                False, below, dp=self._dp(cfg.layers + i), lowp=(1 + (i == 0)) if lowp else 0)
            # ---- decoder input assembly + decoder_proj
            ops.unshuffle_bwd(dx, self.ids_restore, B, N, n, J, dd, self.dy16, ps.grad("image_mask_embedding"))
            dn = view(self.dn16, Me, d)
            self._bwd_linear(self.enc_n, self.dy16, Me, "decoder_proj", dd, d, dX16=dn)
            dx, dx16 = view(self.dxA, Me, d), view(self.dxA16, Me, d)
            if self.enc_lowp and not dp_on and not self.ls_enc:
                ops.layernorm_bwd_stream16(dn, self.xa[cfg.layers], self.meanF, self.rstdF, ps.f("model/norm/scale"), None,
                                           dx, dx16, ps.grad("model/norm/scale"), ps.grad("model/norm/bias"), self.ln_ws,
                                           f32_period=Te, f32_skip=J,
                                           dcol=ps.grad(f"model/layer_{cfg.layers - 1}/ff/w2/bias"), dcol_period=Te,
                                           dcol_skip=J)
                return
            ops.layernorm_bwd(dn, self.xa[cfg.layers], self.meanF, self.rstdF, ps.f("model/norm/scale"), None, dx,
                              dx16, ps.grad("model/norm/scale"), ps.grad("model/norm/bias"), self.ln_ws,
                              dcol=None if (dp_on or self.ls_enc) else ps.grad(f"model/layer_{cfg.layers - 1}/ff/w2/bias"),
                              dcol_period=Te, dcol_skip=J)
            return
        kind, layers = self._segment_plan()[k]
        if kind == "enc":
            for i in layers:
                if self.remat:
                    self._block_fwd(f"model/layer_{i}", self.enc[i], self.xa[i], self.xa[i + 1], Me, d, cfg.heads, Te,
                                    True, self._dp(i))
                self._block_bwd(f"model/layer_{i}", self.enc[i], self.xa[i], Me, d, cfg.heads, Te, True,
                                ps.grad(f"model/layer_{i - 1}/ff/w2/bias"), dp=self._dp(i), x_out=self.xa[i + 1])
        if kind == "jumbo":
            # layer 0's Jumbo branch alone (its patch-FFN and attention branches do not touch the CLS rows'
            # incoming gradient and run in the tail), then the shared Jumbo MLP's weight gradients over all
            # layers at once (K = layers * B)
            if self.remat:      # (the "rest" part of layer 0 follows in the tail segment and finds the same activations)
                self._block_fwd("model/layer_0", self.enc[0], self.xa[0], self.xa[1], Me, d, cfg.heads, Te, True,
                                self._dp(0))
            self._block_bwd("model/layer_0", self.enc[0], self.xa[0], Me, d, cfg.heads, Te, True, None, part="jumbo",
                            dp=self._dp(0), x_out=self.xa[1])
            L, Jd = cfg.layers, J * d
            for nm, X, dY, rows, cols in (("w1", self.cn16_all, self.dhj16_all, Jd, 4 * Jd),
                                          ("w2", self.gj_all, self.dc16_all, 4 * Jd, Jd)):
                ops.gemm(X, dY, rows, cols, L * B, a_mn=True, b_mn=True,
                         out_f32=ps.grad(f"model/jumbo_mlp/{nm}/kernel"), ldo_f32=ps.ld(f"model/jumbo_mlp/{nm}/kernel"),
                         accumulate=True)
        if kind == "tail":
            self._block_bwd("model/layer_0", self.enc[0], self.xa[0], Me, d, cfg.heads, Te, True,
                            ps.grad("model/embed/wte/bias"), part="rest", dp=self._dp(0), below_is_branch=False)
            # ---- CLS parameter and patch embedding (its bias gradient came from layer 0's LN1 backward)
            dx, dx16 = view(self.dxA, Me, d), view(self.dxA16, Me, d)
            ops.colsum(dx, B, J * d, ps.grad("model/cls_tokens"), ldx=Te * d)
            if cfg.posemb == "learnable":      # d wpe = the patch rows of the encoder-input gradient, summed over the batch
                ops.posemb_grad(dx, B, Te, J, Te - J, self.ids_shuffle if cfg.labels <= 0 else None, d,
                                ps.grad("model/embed/wpe"))
            ops.gemm(self.patches, dx16, P, d, Me, a_mn=True, b_mn=True, lda=Pp,
                     out_f32=ps.grad("model/embed/wte/kernel"), ldo_f32=ps.ld("model/embed/wte/kernel"),
                     accumulate=True)

    def _block_bwd(self, name, s: _Block, x_in, M, dim, heads, T, jumbo, dcol_below, part="all", dp=None,
                   below_is_branch=True, lowp=0, x_out=None):
        """In: d/d(block output) in dxA / dxA16.  Out: d/d(block input) in dxA / dxA16.
        ``part``: "all", or the block in two steps, "jumbo" (the Jumbo CLS branch alone: it reads and writes only
        the CLS rows, which the patch FFN branch leaves alone) followed by "rest".
        Bias gradients are produced where their dY is produced (LayerNorm backward, GELU-backward GEMM
        epilogue, attention backward); ``dcol_below`` is the ff/w2 (or patch-embed) bias gradient of the
        block below, whose output gradient this block's LN1 backward emits.
        ``dp``: [3, B] DropPath factors of this block's branches (or None).  The gradient that enters a branch is the
        block-output gradient times its image's factor, so the bf16 gradient copies are scaled in place before the
        branch's products, and the bias gradients that LayerNorm backward would fuse (column sums of the UNscaled
        gradient) are taken from the scaled copies instead.  ``below_is_branch``: dcol_below is the bias of a
        DropPath-scaled branch of the block below (false for the patch embedding).
        ``lowp`` (plain blocks without DropPath): 1 = the gradient arrives and leaves in dxA16 alone (bf16), 2 = it
        arrives in bf16 and leaves in both dxA (f32) and dxA16."""
        ps, B, J = self.ps, self.B, self.cfg.num_cls_tokens
        hd = dim // heads
        view = lambda buf, w: buf[:M * w].view(M, w)
        dx, dx16 = view(self.dxA, dim), view(self.dxA16, dim)
        dxb, dxb16 = view(self.dxB, dim), view(self.dxB16, dim)
        dh, dn = view(self.dh16, 4 * dim), view(self.dn16, dim)
        period, skip = (T, J) if jumbo else (0, 0)
        mode = 1 if jumbo else 0
        any_dp = self.drop_masks is not None
        # LayerScale: y = scale_k * z, so the gradient that enters the branch is dy * scale_k (per column) -- folded into
        # the same in-place scaling of the bf16 gradient copies as DropPath's per-image factor; the weight / bias
        # gradients taken from the scaled copies are then the true ones, and d scale_k follows from them once per
        # optimizer step (layerscale_finalize); scale3 of the shared Jumbo MLP is done per layer, below.
        ls = self.ls_enc if jumbo else self.ls_dec
        g_ = (lambda k: ps.f(f"{name}/scale{k}")) if ls else (lambda k: None)
        explicit = dp is not None or ls
        # encoder blocks: gradient stream in bf16, f32 only on the CLS rows (see enc_lowp); not with DropPath / LayerScale,
        # whose in-place scaling changes the bf16 copies that would be the residual
        enc16 = jumbo and self.enc_lowp and not explicit and not any_dp
        dpk = (lambda k: dp[k]) if dp is not None else (lambda k: None)
        if explicit and part != "rest":
            # dY of ff/w2 (patch rows, drop_path2 / scale2) and of the Jumbo w2 (CLS rows, drop_path3 / scale3)
            ops.scale_bf16(dx16, M, dim, T, scale=dpk(1), scale_cls=dpk(2) if jumbo else None, skip=J if jumbo else 0,
                           col_scale=g_(2), col_scale_cls=g_(3) if jumbo else None)
        if explicit and part != "jumbo":
            # ff/w2 bias: column sums of the scaled gradient over the rows the branch produced
            ops.colsum(dx16, M, dim, ps.grad(f"{name}/ff/w2/bias"), row_period=period, row_skip=skip)
        if part != "jumbo":
            # ---- patch FFN: y = x_b + w2(gelu(w1 LN2(x_b)))
            self._bwd_linear(s.g, dx16, M, f"{name}/ff/w2", dim, 4 * dim, dX16=dh, bias=False, act=GELU_BWD, aux=s.h,
                             row_period=period, row_skip=skip, row_mask_mode=mode,
                             colsum=ps.grad(f"{name}/ff/w1/bias"))
            self._bwd_linear(s.xn2, dh, M, f"{name}/ff/w1", 4 * dim, dim, dX16=dn, bias=False)
            if lowp:
                ops.layernorm_bwd_lowp(dn, s.xb, s.mean2, s.rstd2, ps.f(f"{name}/norm2/scale"), dx16, None, dxb16,
                                       ps.grad(f"{name}/norm2/scale"), ps.grad(f"{name}/norm2/bias"), self.ln_ws,
                                       dcol=ps.grad(f"{name}/attn/wo/bias"))
            elif enc16:
                ops.layernorm_bwd_stream16(dn, s.xb, s.mean2, s.rstd2, ps.f(f"{name}/norm2/scale"), dx16, None, dxb16,
                                           ps.grad(f"{name}/norm2/scale"), ps.grad(f"{name}/norm2/bias"), self.ln_ws,
                                           row_period=period, row_skip=skip, dcol=ps.grad(f"{name}/attn/wo/bias"))
            else:
                ops.layernorm_bwd(dn, s.xb, s.mean2, s.rstd2, ps.f(f"{name}/norm2/scale"), dx, dxb, dxb16,
                                  ps.grad(f"{name}/norm2/scale"), ps.grad(f"{name}/norm2/bias"), self.ln_ws,
                                  row_period=period, row_skip=skip,
                                  dcol=None if explicit else ps.grad(f"{name}/attn/wo/bias"))
        if jumbo and part != "rest":
            # ---- Jumbo CLS: c_out = c + w2j(gelu(w1j c)), c = LN3(cls concat); strided [B, J*dim] views
            Jd, ldc = J * dim, T * dim
            # output gradient of w2 = the CLS rows of dx: bias gradient now, a bf16 copy for the batched dW
            ops.colsum(dx16, B, Jd, ps.grad("model/jumbo_mlp/w2/bias"), ldx=ldc)
            ops.add_bias_cast(dx, B, Jd, dst_bf16=s.dc16, lds=ldc)
            if explicit:
                ops.scale_bf16(s.dc16, B, Jd, 1, scale=dpk(2), col_scale=g_(3))
            if ls:      # d scale3 += sum_b d c_out * (c_out - c) / scale3, before d c is accumulated into the same rows
                ops.layerscale_cls_grad(dx, x_out, s.cn, B, Jd, ps.f(f"{name}/scale3"), ps.grad(f"{name}/scale3"),
                                        lddy=ldc, ldo=ldc, ldc=Jd)
            ops.gemm(dx16, self._W("model/jumbo_mlp/w2/kernel"), B, 4 * Jd, Jd, b_mn=False, lda=ldc,
                     ldb=ps.ld("model/jumbo_mlp/w2/kernel"), out_bf16=s.dhj16, act=GELU_BWD, aux=s.hj,
                     colsum=ps.grad("model/jumbo_mlp/w1/bias"))
            # d c = d c_out + dh W1^T: split-K accumulate into the CLS rows of dx (f32), bf16 copy for LN3 backward
            ops.gemm(s.dhj16, self._W("model/jumbo_mlp/w1/kernel"), B, Jd, 4 * Jd, b_mn=False,
                     ldb=ps.ld("model/jumbo_mlp/w1/kernel"), out_f32=dx, ldo_f32=ldc, accumulate=True)
            ops.add_bias_cast(dx, B, Jd, dst_bf16=self.dcn16, lds=ldc)
            ops.layernorm_bwd(self.dcn16, s.xb, s.mean3, s.rstd3, ps.f(f"{name}/norm3/scale"), None, dxb, dxb16,
                              ps.grad(f"{name}/norm3/scale"), ps.grad(f"{name}/norm3/bias"), self.ln_ws, rows=B,
                              dim=Jd, ldx=ldc, lddx=ldc, lddxb=ldc)
            if not explicit:
                # wo bias: LN2 backward covered the patch rows, add the CLS rows LN3 backward just wrote
                ops.colsum(dxb16, M, dim, ps.grad(f"{name}/attn/wo/bias"), row_period=T, row_skip=J, row_invert=True)
        if part == "jumbo":
            return
        # ---- attention branch: x_b = x + wo(attn(wqkv LN1(x)))
        if explicit:
            ops.scale_bf16(dxb16, M, dim, T, scale=dpk(0), col_scale=g_(1))
            ops.colsum(dxb16, M, dim, ps.grad(f"{name}/attn/wo/bias"))
        datt, dqkv = view(self.datt16, dim), view(self.dqkv16, 3 * dim)
        self._bwd_linear(s.att, dxb16, M, f"{name}/attn/wo", dim, dim, dX16=datt, bias=False)
        ops.attention_bwd(s.qkv, s.att, datt, s.lse, B, T, heads, hd, hd ** -0.5, dqkv, self.delta_ws,
                          dbias=ps.grad(f"{name}/attn/wqkv/bias"))
        self._bwd_linear(s.xn1, dqkv, M, f"{name}/attn/wqkv", 3 * dim, dim, dX16=dn, bias=False)
        if (any_dp or ls) and below_is_branch:
            dcol_below = None       # the block below takes its ff/w2 bias from its own scaled gradient copy
        if lowp:
            ops.layernorm_bwd_lowp(dn, x_in, s.mean1, s.rstd1, ps.f(f"{name}/norm1/scale"), dxb16,
                                   dx if lowp == 2 else None, dx16, ps.grad(f"{name}/norm1/scale"),
                                   ps.grad(f"{name}/norm1/bias"), self.ln_ws, dcol=dcol_below)
            return
        if enc16:
            ops.layernorm_bwd_stream16(dn, x_in, s.mean1, s.rstd1, ps.f(f"{name}/norm1/scale"), dxb16, dx, dx16,
                                       ps.grad(f"{name}/norm1/scale"), ps.grad(f"{name}/norm1/bias"), self.ln_ws,
                                       f32_period=T, f32_skip=J, dcol=dcol_below, dcol_period=period, dcol_skip=skip)
            return
        ops.layernorm_bwd(dn, x_in, s.mean1, s.rstd1, ps.f(f"{name}/norm1/scale"), dxb, dx, dx16,
                          ps.grad(f"{name}/norm1/scale"), ps.grad(f"{name}/norm1/bias"), self.ln_ws,
                          dcol=dcol_below, dcol_period=period, dcol_skip=skip)

    # ------------------------------------------------------------------ CUDA graphs
    def capture_graphs(self) -> None:
        """Records forward + backward of one micro-batch into two CUDA graphs (gradients zeroed first /
        accumulated) over static input buffers, so that a micro-step is one graph launch instead of
        ~1800 kernel launches issued from Python."""
        cfg, B = self.cfg, self.B
        self.images_buf = torch.empty(B, 3, cfg.image_size, cfg.image_size, dtype=torch.uint8, device=self.dev)
        self.noise_buf = torch.zeros(tuple(self.mask_row.shape), dtype=F32, device=self.dev)
        side = torch.cuda.Stream(device=self.dev)
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):                  # eager warm-up: one-time function attributes, module load
            self.images_buf.zero_()
            self.forward(self.images_buf, self.noise_buf)
            self.backward(zero_grads=True)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        self.graphs = {}
        lib = _lib.load()
        for zero in (True, False):
            g = torch.cuda.CUDAGraph()
            l0 = lib.jm_launch_count()
            with torch.cuda.graph(g):
                self.forward(self.images_buf, self.noise_buf)
                self.backward(zero_grads=zero)
            self.graph_launches = lib.jm_launch_count() - l0
            self.graphs[zero] = g
        ops.zero_(self.ps.g)

    def capture_segment_graphs(self) -> None:
        """Graphs for a data-parallel rank's LAST micro-step of an optimizer step: forward + backward
        segment 0, then one graph per encoder segment, so gradient all-reduces can start between them."""
        if not hasattr(self, "graphs"):
            self.capture_graphs()
        self.seg_graphs = {}
        for zero in (True, False):
            graphs = []
            for k in range(self.num_bwd_segments()):
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    if k == 0:
                        self.forward(self.images_buf, self.noise_buf)
                    self.backward_segment(k, zero_grads=zero)
                graphs.append(g)
            self.seg_graphs[zero] = graphs
        ops.zero_(self.ps.g)

    def run_graph_overlapped(self, images: torch.Tensor, noise: torch.Tensor, zero_grads: bool, reducer,
                             after_reduce=None) -> torch.Tensor:
        """Last micro-step of a data-parallel optimizer step: every backward segment is followed by the
        all-reduce of the gradients it finalised, on the reducer's stream, overlapping the next segment."""
        if not hasattr(self, "seg_graphs"):
            self.capture_segment_graphs()
        self.images_buf.copy_(images, non_blocking=True)
        self.noise_buf.copy_(noise, non_blocking=True)
        for k, g in enumerate(self.seg_graphs[bool(zero_grads)]):
            g.replay()
            ev = torch.cuda.Event()
            ev.record()
            ranges = self.segment_grad_ranges(k)
            reducer.all_reduce_ranges(ranges, after_event=ev)
            if after_reduce is not None and reducer.stream is not None:
                with torch.cuda.stream(reducer.stream):      # e.g. the optimizer update of exactly these ranges
                    after_reduce(ranges)
        reducer.join()
        return self.loss

    def run_graph(self, images: torch.Tensor, noise: torch.Tensor, zero_grads: bool) -> torch.Tensor:
        """forward + backward through the captured graph; images may be a pinned host tensor."""
        if not hasattr(self, "graphs"):
            self.capture_graphs()
        self.images_buf.copy_(images, non_blocking=True)
        self.noise_buf.copy_(noise, non_blocking=True)
        self.graphs[bool(zero_grads)].replay()
        return self.loss

    # ------------------------------------------------------------------ optimizer
    def layerscale_finalize(self) -> None:
        """d scale1 / d scale2 of every LayerScale-d block from the ACCUMULATED gradients of the Dense each one scales
        (attn/wo, ff/w2): call once per optimizer step, after the last micro-step's backward (and the all-reduce: the
        identity is linear in the gradients), before clipping and the optimizer.  No-op without LayerScale."""
        cfg, ps = self.cfg, self.ps
        stacks = []
        if self.ls_enc:
            stacks.append((f"model/layer_{{}}", cfg.layers, cfg.dim))
        if self.ls_dec and cfg.labels <= 0:
            stacks.append((f"decoder_model/dec_layer_{{}}", cfg.dec_layers, cfg.dec_dim))
        for fmt, count, dim in stacks:
            for i in range(count):
                name = fmt.format(i)
                for k, lin, K in ((1, "attn/wo", dim), (2, "ff/w2", 4 * dim)):
                    ops.layerscale_grad(ps.f(f"{name}/{lin}/kernel"), ps.grad(f"{name}/{lin}/kernel"),
                                        ps.f(f"{name}/{lin}/bias"), ps.grad(f"{name}/{lin}/bias"),
                                        ps.f(f"{name}/scale{k}"), ps.grad(f"{name}/scale{k}"), K, dim)

    def adamw_step(self, count: int, lr: float, grad_scale: float = 1.0, lr_decay: float = 1.0) -> None:
        """optax.adamw update number ``count`` (0-based) over the flat buffers (src/pretraining.py:223-234).
        ``lr_decay < 1``: the layer-wise scale table chained behind it (:235-242) — the leaves of label i
        (params.get_layer_index) take the step with ``lr * lr_decay ** (layers - i)``, one range list per label."""
        cfg, ps = self.cfg, self.ps
        if lr_decay < 1.0:
            for i in range(cfg.layers + 1):
                ranges = ps.ranges(lambda n, i=i: get_layer_index(n, cfg.layers) == i)
                self.adamw_ranges(ranges, count, lr * lr_decay ** (cfg.layers - i), grad_scale)
            return
        t = count + 1
        ops.adamw(ps.p, ps.g, ps.m, ps.v, ps.p16, ps.n_decay, lr, cfg.adam_b1, cfg.adam_b2, cfg.adam_eps,
                  cfg.weight_decay, 1.0 - cfg.adam_b1 ** t, 1.0 - cfg.adam_b2 ** t, grad_scale)

    def lamb_step(self, count: int, lr: float, grad_scale: float = 1.0, lr_decay: float = 1.0) -> None:
        """--optimizer lamb: modified_lamb (src/utils.py:124-139) update number ``count`` -- the kernel leaves take the
        trust-ratio step (jm_lamb, one norm pair per Flax leaf), every other leaf a plain Adam step (jm_adamw without
        weight decay); ``lr_decay < 1`` scales the step of label i's leaves as in adamw_step."""
        cfg, ps = self.cfg, self.ps
        t = count + 1
        bc1, bc2 = 1.0 - cfg.adam_b1 ** t, 1.0 - cfg.adam_b2 ** t
        labels = range(cfg.layers + 1) if lr_decay < 1.0 else [None]
        if not hasattr(self, "_lamb_tables"):
            self._lamb_tables = {}
        for i in labels:
            sel = (lambda n, i=i: get_layer_index(n, cfg.layers) == i) if i is not None else (lambda n: True)
            step_lr = lr * (lr_decay ** (cfg.layers - i) if i is not None else 1.0)
            if i not in self._lamb_tables:              # built once per label (host lists -> two small device tensors)
                groups, chunks = ps.lamb_tables(sel)
                ws = None if groups is None else torch.empty(ops.lamb_workspace(chunks.shape[0]), dtype=torch.uint8,
                                                             device=self.dev)
                plain = ps.ranges(lambda n, sel=sel: sel(n) and not ps.leaves[n].decay)
                self._lamb_tables[i] = (groups, chunks, ws, plain)
            groups, chunks, ws, plain = self._lamb_tables[i]
            if groups is not None:
                ops.lamb(ps.p, ps.g, ps.m, ps.v, ps.p16, groups, chunks, step_lr, cfg.adam_b1, cfg.adam_b2, cfg.adam_eps,
                         cfg.weight_decay, bc1, bc2, grad_scale, ws)
            for s_, e_ in plain:
                ops.adamw(ps.p[s_:e_], ps.g[s_:e_], ps.m[s_:e_], ps.v[s_:e_], ps.p16[s_:e_], 0, step_lr, cfg.adam_b1,
                          cfg.adam_b2, cfg.adam_eps, 0.0, bc1, bc2, grad_scale)

    def adamw_ranges(self, ranges, count: int, lr: float, grad_scale: float = 1.0) -> None:
        """The same update restricted to [start, end) ranges of the flat buffers (multiples of 64 elements): lets
        a data-parallel step update the parameters of a backward segment as soon as its gradients are reduced,
        while the remaining segments still compute (they no longer read those parameters)."""
        cfg, ps = self.cfg, self.ps
        t = count + 1
        for s, e in ranges:
            nd = min(max(ps.n_decay - s, 0), e - s)
            ops.adamw(ps.p[s:e], ps.g[s:e], ps.m[s:e], ps.v[s:e], ps.p16[s:e], nd, lr, cfg.adam_b1, cfg.adam_b2,
                      cfg.adam_eps, cfg.weight_decay, 1.0 - cfg.adam_b1 ** t, 1.0 - cfg.adam_b2 ** t, grad_scale)

    # ------------------------------------------------------------------ accounting
    def activation_bytes(self) -> int:
        tot, seen = 0, set()

        def add(t):
            nonlocal tot
            key = (t.data_ptr(), t.numel() * t.element_size())
            if key not in seen:                    # (grad_ckpt: the blocks share one set of buffers)
                seen.add(key)
                tot += key[1]

        for obj in [self] + self.enc + self.dec:
            for v in vars(obj).values():
                if isinstance(v, torch.Tensor):
                    add(v)
                elif isinstance(v, list):
                    for t in v:
                        if isinstance(t, torch.Tensor):
                            add(t)
        return tot


def lr_schedule(count: int, cfg: MAEConfig) -> float:
    """optax.warmup_cosine_decay_schedule(1e-6, lr*batch/256, warmup, training_steps, 1e-5)
    (src/pretraining.py:250-259), evaluated on the host."""
    init_value, end_value = 1e-6, 1e-5
    peak = cfg.learning_rate * cfg.train_batch_size / 256
    warm = cfg.warmup_steps
    if count < warm:
        frac = 1.0 - count / warm
        return (init_value - peak) * frac + peak
    decay_steps = cfg.training_steps - warm
    c = min(count - warm, decay_steps)
    cosine = 0.5 * (1.0 + math.cos(math.pi * c / decay_steps))
    alpha = end_value / peak
    return peak * ((1.0 - alpha) * cosine + alpha)
