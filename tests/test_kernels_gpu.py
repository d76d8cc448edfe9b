"""GPU parity tests of the individual kernels, called through the C ABI (ctypes) and
checked against the CPU oracle (oracle/jumbo_mae_oracle.py) or, for the floating-point
GEMM/attention at sizes the oracle cannot reach in seconds, against a torch fp32
restatement of the same operands.  Tolerances are written next to each assert.
"""
import math

import numpy as np
import pytest
import torch

from oracle import jumbo_mae_oracle as O

pytestmark = pytest.mark.gpu

BF16 = torch.bfloat16


@pytest.fixture(scope="module")
def ops():
    from jumbo_mae_tpu_b200 import _lib, ops as _ops
    _lib.require_device()
    return _ops


def dev(t):
    return t.cuda()


# ------------------------------------------------------------------ masking: bit exact
@pytest.mark.parametrize("L,keep", [(196, 49), (256, 64), (16, 4), (1, 0), (1024, 256), (5, 5)])
def test_mask_sort_bit_exact(ops, L, keep):
    rng = np.random.default_rng(L)
    noise = rng.random(L, dtype=np.float32)
    ids_s, ids_r, mask = ops.mask_sort(dev(torch.from_numpy(noise)), keep)
    ws, wr, wm = O.masking_from_noise(noise, keep)
    assert np.array_equal(ids_s.cpu().numpy(), ws)
    assert np.array_equal(ids_r.cpu().numpy(), wr)
    assert np.array_equal(mask.cpu().numpy(), wm)


def test_mask_sort_ties_stable(ops):
    noise = np.array([0.5, 0.25, 0.5, 0.25, 0.5, 0.0, 0.0, 1.0] * 12, dtype=np.float32)
    ids_s, ids_r, mask = ops.mask_sort(dev(torch.from_numpy(noise)), 24)
    ws, wr, wm = O.masking_from_noise(noise, 24)
    assert np.array_equal(ids_s.cpu().numpy(), ws)
    assert np.array_equal(ids_r.cpu().numpy(), wr)
    assert np.array_equal(mask.cpu().numpy(), wm)
    # permutation property at any size: ids_restore[ids_shuffle] == arange
    assert torch.equal(ids_r[ids_s.long()].cpu(), torch.arange(len(noise), dtype=torch.int32))


def test_index_sequence(ops):
    x = torch.randn(3, 16, 20)
    ids = torch.tensor([5, 0, 15, 7], dtype=torch.int32)
    got = ops.index_sequence(dev(x), dev(ids))
    assert torch.equal(got.cpu(), O.index_sequence(x, ids))          # pure copy: exact


# ------------------------------------------------------------------ pixels
@pytest.mark.parametrize("S,p", [(32, 8), (224, 16), (224, 14)])
def test_extract_patches_bit_exact(ops, S, p):
    g = torch.Generator().manual_seed(0)
    images = torch.randint(0, 256, (2, 3, S, S), generator=g, dtype=torch.uint8)
    got = ops.extract_patches_u8(dev(images), p)
    want = O.extract_patches(O.preprocess(images), p)
    assert torch.equal(got.cpu(), want)                               # same IEEE ops: bit exact


def test_patchify_kept_rows(ops):
    S, p, J = 32, 8, 3
    g = torch.Generator().manual_seed(1)
    images = torch.randint(0, 256, (2, 3, S, S), generator=g, dtype=torch.uint8)
    ids = torch.tensor([7, 2, 11, 0], dtype=torch.int32)
    P, ld = p * p * 3, 200
    out = torch.full((2 * (J + 4), ld), 7.0, dtype=BF16, device="cuda")
    ops.patchify(dev(images), p, dev(ids), 4, J + 4, J, out)
    want = O.extract_patches(O.preprocess(images), p)[:, ids.long()]
    got = out.float().cpu().reshape(2, J + 4, ld)
    assert torch.equal(got[:, :J], torch.zeros(2, J, ld))
    assert torch.equal(got[:, J:, P:], torch.zeros(2, 4, ld - P))
    assert torch.equal(got[:, J:, :P], want.to(BF16).float())          # exact after the same bf16 rounding


@pytest.mark.parametrize("norm_pix", [False, True])
def test_patch_loss_and_grad(ops, norm_pix):
    S, p, J, B = 32, 8, 3, 3
    N, P = (S // p) ** 2, p * p * 3
    g = torch.Generator().manual_seed(2)
    images = torch.randint(0, 256, (B, 3, S, S), generator=g, dtype=torch.uint8)
    noise = torch.rand(N, generator=g).numpy()
    _, _, mask_row = O.masking_from_noise(noise, 4)
    pred = torch.randn(B, J + N, P, generator=g)
    # oracle
    pr = pred[:, J:].clone().requires_grad_(True)
    tgt = O.extract_patches(O.preprocess(images), p)
    if norm_pix:
        tgt = (tgt - tgt.mean(-1, keepdim=True)) / torch.sqrt(tgt.var(-1, keepdim=True, unbiased=False) + 1e-6)
    mask = torch.from_numpy(mask_row)[None].expand(B, N)
    want = O.patch_mse_loss(pr, tgt, mask)
    want.backward()
    # kernel
    loss = torch.zeros(1, device="cuda")
    dpred = torch.full((B * (J + N), P), 9.0, dtype=BF16, device="cuda")
    ws = torch.empty(ops.patch_loss_workspace(B, N), dtype=torch.uint8, device="cuda")
    ops.patch_loss(dev(pred.reshape(-1, P)), dev(images), dev(torch.from_numpy(mask_row)), p, J, norm_pix, 1.0,
                   loss, dpred, ws)
    assert abs(loss.item() - want.item()) < 1e-5 * abs(want.item())   # f32 reduction order only
    got = dpred.float().cpu().reshape(B, J + N, P)
    assert torch.equal(got[:, :J], torch.zeros(B, J, P))
    # bf16 output: 2^-8 relative
    assert torch.allclose(got[:, J:], pr.grad, rtol=2 ** -7, atol=1e-9)


@pytest.mark.parametrize("S,p,keep,B,dd,norm_pix", [(32, 8, 4, 3, 64, False), (32, 8, 4, 3, 64, True),
                                                      (56, 14, 5, 5, 128, True), (224, 16, 49, 2, 512, True)])
def test_pixel_head_loss_epilogue(ops, S, p, keep, B, dd, norm_pix):
    """jm_masked_targets + jm_gemm(act=5) + jm_row_loss_finalize = decoder_image_output + extract_patches + norm-pix +
    patch_mse_loss on the masked rows (src/pretraining.py:110-120, src/utils_mae.py:51-64), and d loss / d prediction;
    P = 192, 588 (ViT-H/14: N not a multiple of 16) and 768."""
    N, P, J = (S // p) ** 2, p * p * 3, 3
    g = torch.Generator().manual_seed(S + p + B)
    images = torch.randint(0, 256, (B, 3, S, S), generator=g, dtype=torch.uint8)
    noise = torch.rand(N, generator=g).numpy()
    ids_shuffle, _, mask_row = O.masking_from_noise(noise, keep)
    ids = torch.from_numpy(np.asarray(ids_shuffle)).to(torch.int32)
    n_mask = N - keep
    Mh = B * n_mask
    X = (torch.randn(Mh, dd, generator=g)).to(BF16)
    W_full = (torch.randn(dd, (P + 7) // 8 * 8, generator=g) * 0.1).to(BF16)        # row stride padded to 16 bytes
    W = W_full[:, :P]
    bias = torch.randn(P, generator=g)
    # oracle: prediction of the masked rows, targets of the same patches
    pred = (X.float() @ W.float() + bias).reshape(B, n_mask, P).clone().requires_grad_(True)
    tgt = O.extract_patches(O.preprocess(images), p)[:, ids[keep:].long()]
    if norm_pix:
        tgt = (tgt - tgt.mean(-1, keepdim=True)) / torch.sqrt(tgt.var(-1, keepdim=True, unbiased=False) + 1e-6)
    want = O.patch_mse_loss(pred, tgt, torch.ones(B, n_mask))
    want.backward()
    # kernels
    Pt, Pp = (P + 15) // 16 * 16, (P + 7) // 8 * 8
    t8 = torch.empty(Mh, Pt, dtype=torch.uint8, device="cuda")
    stats = torch.empty(Mh, 2, device="cuda") if norm_pix else None
    row_map = torch.empty(Mh, dtype=torch.int32, device="cuda")
    ops.masked_targets(dev(images), p, dev(ids)[keep:], J, norm_pix, t8, stats, row_map)
    want_rows = (torch.arange(B)[:, None] * (J + N) + J + ids[keep:][None, :].long()).reshape(-1)
    assert torch.equal(row_map.cpu().long(), want_rows)
    raw = O.extract_patches(images.permute(0, 2, 3, 1).float(), p)[:, ids[keep:].long()].reshape(Mh, P)   # patch-vector order
    assert torch.equal(t8[:, :P].cpu().float(), raw)
    row_loss = torch.full((Mh, (P + 15) // 16), float("nan"), device="cuda")
    dpred = torch.full((Mh, Pp), 9.0, dtype=BF16, device="cuda")
    loss = torch.zeros(1, device="cuda")
    ws = torch.empty(ops.row_loss_workspace(), dtype=torch.uint8, device="cuda")
    denom = float(P) * n_mask * B
    ops.gemm(dev(X), dev(W_full)[:, :P], Mh, P, dd, b_mn=True, bias=dev(bias), act=5, out_bf16=dpred, loss_target=t8,
             loss_stats=stats, row_loss=row_loss, loss_scale=2.0 / denom)
    ops.row_loss_finalize(row_loss, 1.0 / denom, loss, ws)
    assert abs(loss.item() - want.item()) < 2e-5 * abs(want.item())
    got = dpred.float().cpu()
    assert torch.allclose(got[:, :P], pred.grad.reshape(Mh, P), rtol=2 ** -7, atol=1e-9)
    assert torch.equal(got[:, P:], torch.zeros(Mh, Pp - P))          # pad columns written as zeros
    # validation form: no gradient output
    row_loss.fill_(float("nan"))
    ops.gemm(dev(X), dev(W_full)[:, :P], Mh, P, dd, b_mn=True, bias=dev(bias), act=5, loss_target=t8, loss_stats=stats,
             row_loss=row_loss, loss_scale=0.0)
    loss2 = torch.zeros(1, device="cuda")
    ops.row_loss_finalize(row_loss, 1.0 / denom, loss2, ws)
    assert loss2.item() == loss.item()                                # deterministic, same order


@pytest.mark.parametrize("rows,D,total", [(37, 64, 100), (300, 512, 1000)])
def test_layernorm_rows_gather_scatter(ops, rows, D, total):
    """jm_layernorm_fwd_rows / jm_layernorm_bwd_rows: LayerNorm of a subset of rows (compact output), backward scattered
    into a zeroed dx."""
    g = torch.Generator().manual_seed(D + rows)
    x = (torch.randn(total, D, generator=g) * 2 + 0.5)
    gamma, beta = torch.randn(D, generator=g), torch.randn(D, generator=g)
    row_map = torch.randperm(total, generator=g)[:rows].to(torch.int32)
    xs = x[row_map.long()].clone().requires_grad_(True)
    gm, bt = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    y = O.layer_norm(xs, gm, bt)
    dy = torch.randn(rows, D, generator=g).to(BF16)
    (y * dy.float()).sum().backward()
    yb = torch.empty(rows, D, dtype=BF16, device="cuda")
    mean, rstd = torch.empty(rows, device="cuda"), torch.empty(rows, device="cuda")
    ops.layernorm_fwd_rows(dev(x), dev(row_map), dev(gamma), dev(beta), 1e-6, yb, mean, rstd)
    assert torch.allclose(yb.float().cpu(), y.detach(), rtol=2 ** -7, atol=1e-3)
    dx = torch.full((total, D), 5.0, device="cuda")
    dxb = torch.full((total, D), 5.0, dtype=BF16, device="cuda")
    dgamma, dbeta, dcol = (torch.zeros(D, device="cuda") for _ in range(3))
    ws = torch.empty(ops.layernorm_bwd_workspace(total, D), dtype=torch.uint8, device="cuda")
    ops.layernorm_bwd_rows(dev(dy), dev(x), dev(row_map), mean, rstd, dev(gamma), dx, dxb, dgamma, dbeta, ws, dcol=dcol)
    want = torch.zeros(total, D)
    want[row_map.long()] = xs.grad
    assert torch.allclose(dx.cpu(), want, rtol=1e-4, atol=1e-4)
    assert torch.allclose(dxb.float().cpu(), want, rtol=2 ** -7, atol=1e-3)
    scale = max(1.0, math.sqrt(rows))
    assert torch.allclose(dgamma.cpu(), gm.grad, rtol=1e-4, atol=1e-4 * scale)
    assert torch.allclose(dbeta.cpu(), bt.grad, rtol=1e-4, atol=1e-4 * scale)
    assert torch.allclose(dcol.cpu(), want.sum(0), rtol=1e-4, atol=1e-3 * scale)


# ------------------------------------------------------------------ LayerNorm
@pytest.mark.parametrize("rows,D", [(37, 64), (300, 512), (260, 768), (16, 2304), (9, 1280)])
def test_layernorm_fwd(ops, rows, D):
    g = torch.Generator().manual_seed(D)
    x = torch.randn(rows, D, generator=g) * 2 + 0.5
    gamma, beta = torch.randn(D, generator=g), torch.randn(D, generator=g)
    want = O.layer_norm(x, gamma, beta)
    yb = torch.empty(rows, D, dtype=BF16, device="cuda")
    yf = torch.empty(rows, D, device="cuda")
    mean, rstd = torch.empty(rows, device="cuda"), torch.empty(rows, device="cuda")
    ops.layernorm_fwd(dev(x), dev(gamma), dev(beta), 1e-6, yb, yf, mean, rstd)
    assert torch.allclose(yf.cpu(), want, rtol=1e-5, atol=2e-5)       # f32, different summation order
    assert torch.allclose(yb.float().cpu(), want, rtol=2 ** -7, atol=1e-6)
    assert torch.allclose(mean.cpu(), x.mean(-1), atol=1e-5)


def test_layernorm_fwd_strided_rows(ops):
    # the Jumbo view: rows of J*d taken every T*d elements (src/modeling.py:194-197)
    B, T, J, d = 5, 7, 3, 64
    x = torch.randn(B, T, d)
    gamma, beta = torch.randn(J * d), torch.randn(J * d)
    want = O.layer_norm(x[:, :J].reshape(B, J * d), gamma, beta)
    yf = torch.empty(B, J * d, device="cuda")
    xd = dev(x)
    ops.layernorm_fwd(xd, dev(gamma), dev(beta), 1e-6, None, yf, None, None, rows=B, dim=J * d, ldx=T * d)
    assert torch.allclose(yf.cpu(), want, rtol=1e-5, atol=2e-5)


@pytest.mark.parametrize("rows,D,period,skip", [(64, 64, 0, 0), (500, 512, 0, 0), (260, 768, 52, 3),
                                                (12, 2304, 0, 0), (40, 1280, 0, 0),
                                                # several rows per warp: the software-pipelined kernel (dim <= 512)
                                                # with and without skipped rows, and the dim-768 kernel
                                                (5000, 512, 13, 3), (7000, 256, 7, 2), (4001, 128, 0, 0),
                                                (6000, 768, 52, 3)])
def test_layernorm_bwd(ops, rows, D, period, skip):
    g = torch.Generator().manual_seed(D + rows)
    x = (torch.randn(rows, D, generator=g) * 2 + 0.5).requires_grad_(True)
    gamma = torch.randn(D, generator=g).requires_grad_(True)
    beta = torch.randn(D, generator=g).requires_grad_(True)
    dy = torch.randn(rows, D, generator=g).to(BF16)
    dres = torch.randn(rows, D, generator=g)
    keep = torch.ones(rows, dtype=torch.bool)
    if period:
        keep = (torch.arange(rows) % period) >= skip
    y = O.layer_norm(x, gamma, beta)
    (y * dy.float() * keep[:, None]).sum().backward()
    want_dx = x.grad + dres
    # kernel
    mean = x.detach().mean(-1)
    var = torch.clamp((x.detach() ** 2).mean(-1) - mean ** 2, min=0)
    rstd = torch.rsqrt(var + 1e-6)
    dx = torch.full((rows, D), 5.0, device="cuda")
    dxb = torch.full((rows, D), 5.0, dtype=BF16, device="cuda")
    dgamma = torch.ones(D, device="cuda")          # accumulated into
    dbeta = torch.ones(D, device="cuda")
    ws = torch.empty(ops.layernorm_bwd_workspace(rows, D), dtype=torch.uint8, device="cuda")
    dcol = torch.ones(D, device="cuda") if D <= 1280 else None
    ops.layernorm_bwd(dev(dy), dev(x.detach()), dev(mean), dev(rstd), dev(gamma.detach()), dev(dres), dx, dxb,
                      dgamma, dbeta, ws, row_period=period, row_skip=skip, dcol=dcol, dcol_period=7, dcol_skip=2)
    if dcol is not None:      # fused bias gradient: column sums of dx over rows with row % 7 >= 2
        sel = keep & ((torch.arange(rows) % 7) >= 2)
        assert torch.allclose(dcol.cpu() - 1, want_dx[sel].sum(0), rtol=1e-4, atol=1e-3 * max(1.0, math.sqrt(rows)))
    got = dx.cpu()
    assert torch.allclose(got[keep], want_dx[keep], rtol=1e-4, atol=1e-4)
    assert torch.equal(got[~keep], torch.full_like(got[~keep], 5.0))   # skipped rows untouched
    assert torch.allclose(dxb.float().cpu()[keep], want_dx[keep], rtol=2 ** -7, atol=1e-3)
    scale = max(1.0, math.sqrt(rows))
    assert torch.allclose(dgamma.cpu() - 1, gamma.grad, rtol=1e-4, atol=1e-4 * scale)
    assert torch.allclose(dbeta.cpu() - 1, beta.grad, rtol=1e-4, atol=1e-4 * scale)


@pytest.mark.parametrize("rows,D,with_f32", [(257, 512, False), (130, 256, True), (1000, 128, False)])
def test_layernorm_bwd_lowp(ops, rows, D, with_f32):
    """jm_layernorm_bwd_lowp: the residual gradient read from bf16, dx written in bf16 (and f32 on request) --
    same arithmetic as jm_layernorm_bwd fed the same (bf16-rounded) residual gradient."""
    g = torch.Generator().manual_seed(D + rows)
    x = (torch.randn(rows, D, generator=g) * 2 + 0.5).requires_grad_(True)
    gamma = torch.randn(D, generator=g).requires_grad_(True)
    beta = torch.randn(D, generator=g).requires_grad_(True)
    dy = torch.randn(rows, D, generator=g).to(BF16)
    dres = torch.randn(rows, D, generator=g).to(BF16)
    y = O.layer_norm(x, gamma, beta)
    (y * dy.float()).sum().backward()
    want_dx = x.grad + dres.float()
    mean = x.detach().mean(-1)
    var = torch.clamp((x.detach() ** 2).mean(-1) - mean ** 2, min=0)
    rstd = torch.rsqrt(var + 1e-6)
    dx = torch.full((rows, D), 5.0, device="cuda") if with_f32 else None
    dxb = torch.full((rows, D), 5.0, dtype=BF16, device="cuda")
    dgamma, dbeta, dcol = (torch.ones(D, device="cuda") for _ in range(3))
    ws = torch.empty(ops.layernorm_bwd_workspace(rows, D), dtype=torch.uint8, device="cuda")
    ops.layernorm_bwd_lowp(dev(dy), dev(x.detach()), dev(mean), dev(rstd), dev(gamma.detach()), dev(dres), dx, dxb,
                           dgamma, dbeta, ws, dcol=dcol)
    scale = max(1.0, math.sqrt(rows))
    if with_f32:
        assert torch.allclose(dx.cpu(), want_dx, rtol=1e-4, atol=1e-4)
    assert torch.allclose(dxb.float().cpu(), want_dx, rtol=2 ** -7, atol=1e-3)
    assert torch.allclose(dcol.cpu() - 1, want_dx.sum(0), rtol=1e-4, atol=1e-3 * scale)
    assert torch.allclose(dgamma.cpu() - 1, gamma.grad, rtol=1e-4, atol=1e-4 * scale)
    assert torch.allclose(dbeta.cpu() - 1, beta.grad, rtol=1e-4, atol=1e-4 * scale)
    with pytest.raises(RuntimeError):        # rows wider than 1280 do not have the bf16 input
        W = 1536
        ops.layernorm_bwd_lowp(torch.zeros(4, W, dtype=BF16, device="cuda"), torch.zeros(4, W, device="cuda"),
                               torch.zeros(4, device="cuda"), torch.ones(4, device="cuda"),
                               torch.ones(W, device="cuda"), torch.zeros(4, W, dtype=BF16, device="cuda"), None,
                               torch.zeros(4, W, dtype=BF16, device="cuda"), torch.zeros(W, device="cuda"),
                               torch.zeros(W, device="cuda"),
                               torch.empty(ops.layernorm_bwd_workspace(4, W), dtype=torch.uint8, device="cuda"))


@pytest.mark.parametrize("rows,D", [(260, 768), (130, 1024), (65, 1280), (260, 256)])
def test_layernorm_bwd_stream16(ops, rows, D):
    """jm_layernorm_bwd_stream16 (the encoder's bf16 gradient stream): bf16 residual in, rows r % 13 < 3 skipped
    (untouched), bf16 dx for every processed row, the f32 copy only for rows r % 7 < 2, column sums over rows r % 5 >= 1."""
    g = torch.Generator().manual_seed(D + rows)
    x = (torch.randn(rows, D, generator=g) * 2 + 0.5).requires_grad_(True)
    gamma = torch.randn(D, generator=g).requires_grad_(True)
    beta = torch.randn(D, generator=g).requires_grad_(True)
    dy = torch.randn(rows, D, generator=g).to(BF16)
    dres = torch.randn(rows, D, generator=g).to(BF16)
    r = torch.arange(rows)
    keep = (r % 13) >= 3
    y = O.layer_norm(x, gamma, beta)
    (y * dy.float() * keep[:, None]).sum().backward()
    want_dx = x.grad + dres.float()
    mean = x.detach().mean(-1)
    rstd = torch.rsqrt(torch.clamp((x.detach() ** 2).mean(-1) - mean ** 2, min=0) + 1e-6)
    dx = torch.full((rows, D), 5.0, device="cuda")
    dxb = torch.full((rows, D), 5.0, dtype=BF16, device="cuda")
    dgamma, dbeta, dcol = (torch.zeros(D, device="cuda") for _ in range(3))
    ws = torch.empty(ops.layernorm_bwd_workspace(rows, D), dtype=torch.uint8, device="cuda")
    ops.layernorm_bwd_stream16(dev(dy), dev(x.detach()), dev(mean), dev(rstd), dev(gamma.detach()), dev(dres), dx, dxb,
                               dgamma, dbeta, ws, f32_period=7, f32_skip=2, row_period=13, row_skip=3, dcol=dcol,
                               dcol_period=5, dcol_skip=1)
    f32_rows = keep & ((r % 7) < 2)
    got, gotb = dx.cpu(), dxb.float().cpu()
    assert torch.allclose(got[f32_rows], want_dx[f32_rows], rtol=1e-4, atol=1e-4)
    assert torch.equal(got[~f32_rows], torch.full_like(got[~f32_rows], 5.0))          # no f32 copy elsewhere
    assert torch.allclose(gotb[keep], want_dx[keep], rtol=2 ** -7, atol=1e-3)
    assert torch.equal(gotb[~keep], torch.full_like(gotb[~keep], 5.0))                # skipped rows untouched
    scale = max(1.0, math.sqrt(rows))
    sel = keep & ((r % 5) >= 1)
    assert torch.allclose(dcol.cpu(), want_dx[sel].sum(0), rtol=1e-4, atol=1e-3 * scale)
    assert torch.allclose(dgamma.cpu(), gamma.grad, rtol=1e-4, atol=1e-4 * scale)
    assert torch.allclose(dbeta.cpu(), beta.grad, rtol=1e-4, atol=1e-4 * scale)


# ------------------------------------------------------------------ token assembly
def test_fill_cls_and_unshuffle(ops):
    B, N, n, J, d = 3, 16, 4, 3, 32
    g = torch.Generator().manual_seed(3)
    noise = torch.rand(N, generator=g).numpy()
    _, ids_restore, _ = O.masking_from_noise(noise, n)
    y = torch.randn(B, J + n, d, generator=g)
    mask_emb = torch.randn(d, generator=g)
    wpe = torch.randn(N, d, generator=g)
    out = torch.empty(B * (J + N), d, device="cuda")
    ids_d = dev(torch.from_numpy(ids_restore))
    ops.unshuffle_fwd(dev(y.reshape(-1, d)), ids_d, dev(mask_emb), dev(wpe), B, N, n, J, d, out)
    # src/pretraining.py:96-106 + src/modeling.py:289-292
    image_x = torch.cat([y[:, J:], mask_emb.expand(B, N - n, d)], dim=1)
    image_x = O.index_sequence(image_x, torch.from_numpy(ids_restore)) + wpe
    want = torch.cat([y[:, :J], image_x], dim=1)
    assert torch.equal(out.cpu().reshape(B, J + N, d), want)          # copies and one f32 add: exact

    dout = torch.randn(B, J + N, d, generator=g)
    dy = torch.full((B * (J + n), d), 3.0, dtype=BF16, device="cuda")
    dmask = torch.ones(d, device="cuda")
    ops.unshuffle_bwd(dev(dout.reshape(-1, d)), ids_d, B, N, n, J, d, dy, dmask)
    yy = y.clone().requires_grad_(True)
    me = mask_emb.clone().requires_grad_(True)
    ix = torch.cat([yy[:, J:], me.expand(B, N - n, d)], dim=1)
    full = torch.cat([yy[:, :J], O.index_sequence(ix, torch.from_numpy(ids_restore)) + wpe], dim=1)
    (full * dout).sum().backward()
    assert torch.equal(dy.float().cpu().reshape(B, J + n, d), yy.grad.to(BF16).float())
    assert torch.allclose(dmask.cpu() - 1, me.grad, rtol=1e-5, atol=1e-5)

    x = torch.zeros(B * 7, d, device="cuda")
    cls = torch.randn(J, d, generator=g)
    ops.fill_cls(x, B, 7, J, d, dev(cls))
    got = x.cpu().reshape(B, 7, d)
    assert torch.equal(got[:, :J], cls.expand(B, J, d)) and torch.equal(got[:, J:], torch.zeros(B, 4, d))


@pytest.mark.parametrize("f32", [False, True])
def test_colsum(ops, f32):
    rows, cols = 1000, 260
    g = torch.Generator().manual_seed(4)
    x = torch.randn(rows, 264, generator=g)
    x = x if f32 else x.to(BF16)
    out = torch.ones(cols, device="cuda")
    ops.colsum(dev(x), rows, cols, out, row_period=10, row_skip=2)
    keep = (torch.arange(rows) % 10) >= 2
    want = x.float()[keep][:, :cols].sum(0)
    assert torch.allclose(out.cpu() - 1, want, rtol=1e-4, atol=1e-3)
    out2 = torch.zeros(cols, device="cuda")
    ops.colsum(dev(x), rows, cols, out2, row_period=10, row_skip=2, row_invert=True)
    assert torch.allclose(out2.cpu(), x.float()[~keep][:, :cols].sum(0), rtol=1e-4, atol=1e-3)


# ------------------------------------------------------------------ AdamW
def test_adamw_matches_oracle(ops):
    cfg = O.vit_b16()
    n, n_decay = 4096, 1024
    g = torch.Generator().manual_seed(5)
    p, grad = torch.randn(n, generator=g), torch.randn(n, generator=g) * 1e-3
    m, v = torch.randn(n, generator=g) * 1e-3, torch.rand(n, generator=g) * 1e-6
    count, lr = 7, 3e-4
    want_p, want_m, want_v = p.clone(), m.clone(), v.clone()
    a, b, c = O.adamw_leaf(p[:n_decay], grad[:n_decay], m[:n_decay], v[:n_decay], count, lr, cfg, decay=True)
    want_p[:n_decay], want_m[:n_decay], want_v[:n_decay] = a, b, c
    a, b, c = O.adamw_leaf(p[n_decay:], grad[n_decay:], m[n_decay:], v[n_decay:], count, lr, cfg, decay=False)
    want_p[n_decay:], want_m[n_decay:], want_v[n_decay:] = a, b, c
    pd, gd, md, vd = dev(p), dev(grad), dev(m), dev(v)
    pb = torch.empty(n, dtype=BF16, device="cuda")
    t = count + 1
    ops.adamw(pd, gd, md, vd, pb, n_decay, lr, cfg.adam_b1, cfg.adam_b2, cfg.adam_eps, cfg.weight_decay,
              1 - cfg.adam_b1 ** t, 1 - cfg.adam_b2 ** t, 1.0)
    # identical f32 formula; allow a couple of ulp for fma contraction / division rounding
    assert torch.allclose(pd.cpu(), want_p, rtol=3e-7, atol=1e-9)
    # m and v are sums of two rounded terms: an fma keeps one of them exact, so allow 2 ulp of the larger term
    assert torch.allclose(md.cpu(), want_m, rtol=3e-7, atol=2 * 1.2e-7 * want_m.abs().max().item())
    assert torch.allclose(vd.cpu(), want_v, rtol=3e-7, atol=2 * 1.2e-7 * want_v.abs().max().item())
    assert torch.equal(pb.float().cpu(), pd.cpu().to(BF16).float())


# ------------------------------------------------------------------ GEMM
def _gemm_ref(A, B, a_mn, b_mn):
    Af = A.float().t() if a_mn else A.float()
    Bf = B.float() if b_mn else B.float().t()
    return Af @ Bf


GEMM_CASES = [
    # M, N, K, a_mn, b_mn, block_n
    (128, 128, 64, False, False, 128),
    (512, 384, 320, False, False, 128),
    (256, 192, 128, False, False, 64),
    (256, 512, 128, False, False, 256),
    (512, 384, 320, False, True, 128),
    (384, 512, 192, False, True, 256),
    (384, 256, 1000, True, True, 128),
    (384, 256, 328, True, False, 64),
    (200, 588, 72, False, True, 0),          # ragged M, N, K (ViT-H/14 head width)
    (1, 64, 8, False, True, 0),              # single row
    (4160, 2304, 768, False, True, 0),       # persistent: more tiles than SMs
]


@pytest.mark.parametrize("M,N,K,a_mn,b_mn,bn", GEMM_CASES)
def test_gemm_plain(ops, M, N, K, a_mn, b_mn, bn):
    g = torch.Generator(device="cuda").manual_seed(M + N + K)
    A = torch.randn((K, M) if a_mn else (M, K), generator=g, device="cuda").to(BF16)
    # operand row strides must be multiples of 8 elements (16 bytes for TMA): pad, then view
    ldn, ldk = (N + 7) // 8 * 8, (K + 7) // 8 * 8
    B = torch.randn((K, ldn) if b_mn else (N, ldk), generator=g, device="cuda").to(BF16)
    B = B[:, :N] if b_mn else B[:, :K]
    out = torch.full((M, N), float("nan"), device="cuda")
    ops.gemm(A, B, M, N, K, a_mn=a_mn, b_mn=b_mn, out_f32=out, block_n=bn)
    ref = _gemm_ref(A, B, a_mn, b_mn)
    # exact bf16 products, f32 accumulation in a different order: 1e-5 relative to the row scale
    assert (out - ref).abs().max().item() <= 2e-5 * ref.abs().max().item() * math.sqrt(K)


def test_gemm_epilogue_bias_scale_residual_bf16(ops):
    M, N, K = 300, 256, 192
    g = torch.Generator(device="cuda").manual_seed(0)
    A = torch.randn(M, K, generator=g, device="cuda").to(BF16)
    W = (torch.randn(K, N, generator=g, device="cuda") * 0.1).to(BF16)
    bias = torch.randn(N, generator=g, device="cuda")
    res = torch.randn(M, N, generator=g, device="cuda")
    out = torch.empty(M, N, device="cuda")
    outb = torch.empty(M, N, dtype=BF16, device="cuda")
    ops.gemm(A, W, M, N, K, b_mn=True, out_f32=out, out_bf16=outb, bias=bias, residual=res, col_scale=0.125,
             col_scale_n=64)
    ref = A.float() @ W.float() + bias
    ref[:, :64] *= 0.125
    ref = ref + res
    assert torch.allclose(out, ref, rtol=1e-5, atol=1e-4)
    assert torch.allclose(outb.float(), ref, rtol=2 ** -7, atol=1e-3)


def test_gemm_epilogue_gelu_fwd_and_bwd_with_row_mask(ops):
    M, N, K, T, J = 28, 128, 64, 7, 3
    g = torch.Generator(device="cuda").manual_seed(1)
    A = torch.randn(M, K, generator=g, device="cuda").to(BF16)
    W = (torch.randn(K, N, generator=g, device="cuda") * 0.2).to(BF16)
    bias = torch.randn(N, generator=g, device="cuda")
    h = torch.empty(M, N, dtype=BF16, device="cuda")
    gl = torch.empty(M, N, dtype=BF16, device="cuda")
    ops.gemm(A, W, M, N, K, b_mn=True, out_bf16=h, out2_bf16=gl, bias=bias, act=1, row_period=T, row_skip=J,
             row_mask_mode=1)
    keep = ((torch.arange(M, device="cuda") % T) >= J)[:, None]
    pre = (A.float() @ W.float() + bias) * keep
    assert torch.allclose(h.float(), pre, rtol=2 ** -7, atol=1e-3)
    assert torch.allclose(gl.float(), O.gelu_tanh(pre.cpu()).cuda() * keep, rtol=2 ** -7, atol=2e-3)
    # backward: dH = (dY @ W2^T) * gelu'(h), CLS rows zeroed
    dY = torch.randn(M, K, generator=g, device="cuda").to(BF16)
    W2 = (torch.randn(N, K, generator=g, device="cuda") * 0.2).to(BF16)          # Flax kernel [in=N, out=K]
    dH = torch.empty(M, N, dtype=BF16, device="cuda")
    cs = torch.ones(N, device="cuda")
    ops.gemm(dY, W2, M, N, K, b_mn=False, out_bf16=dH, aux=h, act=2, row_period=T, row_skip=J, row_mask_mode=1,
             colsum=cs)
    hh = h.float().cpu().requires_grad_(True)
    O.gelu_tanh(hh).sum().backward()
    ref = (dY.float() @ W2.float().t()) * hh.grad.cuda() * keep
    assert torch.allclose(dH.float(), ref, rtol=2 ** -6, atol=2e-3)
    assert torch.allclose(cs - 1, ref.sum(0), rtol=1e-2, atol=2e-2)           # fused bias gradient


def test_gemm_epilogue_posemb_and_skip_rows(ops):
    M, N, K, T, J, n = 21, 64, 72, 7, 3, 4
    g = torch.Generator(device="cuda").manual_seed(2)
    A = torch.randn(M, K + 8, generator=g, device="cuda").to(BF16)
    W = (torch.randn(K, N, generator=g, device="cuda") * 0.2).to(BF16)
    bias = torch.randn(N, generator=g, device="cuda")
    pe = torch.randn(16, N, generator=g, device="cuda")
    idx = torch.tensor([5, 0, 15, 9], dtype=torch.int32, device="cuda")
    out = torch.full((M, N), -7.0, device="cuda")
    ops.gemm(A, W, M, N, K, b_mn=True, out_f32=out, bias=bias, pe=pe, pe_idx=idx, row_period=T, row_skip=J,
             row_mask_mode=2)
    ref = A[:, :K].float() @ W.float() + bias
    r = torch.arange(M, device="cuda") % T
    keep = r >= J
    ref[keep] += pe[idx[(r[keep] - J).long()].long()]
    assert torch.allclose(out[keep], ref[keep], rtol=1e-5, atol=1e-4)
    assert torch.equal(out[~keep], torch.full_like(out[~keep], -7.0))


def test_gemm_weight_grad_accumulate_and_split_k(ops):
    # dW += X^T dY with both operands token-major (MN-major), split-K over tokens
    tokens, din, dout = 3000, 256, 384
    g = torch.Generator(device="cuda").manual_seed(3)
    X = torch.randn(tokens, din, generator=g, device="cuda").to(BF16)
    dY = torch.randn(tokens, dout, generator=g, device="cuda").to(BF16)
    dW = torch.ones(din, dout, device="cuda")
    ops.gemm(X, dY, din, dout, tokens, a_mn=True, b_mn=True, out_f32=dW, accumulate=True)
    ref = X.float().t() @ dY.float() + 1
    assert torch.allclose(dW, ref, rtol=1e-4, atol=2e-3)
    dW2 = torch.zeros(din, dout, device="cuda")
    ops.gemm(X, dY, din, dout, tokens, a_mn=True, b_mn=True, out_f32=dW2, accumulate=True, splits=5, block_n=128)
    assert torch.allclose(dW2, ref - 1, rtol=1e-4, atol=2e-3)


def test_gemm_strided_jumbo_store(ops):
    # Jumbo FFN output written straight into the CLS rows of x: row stride T*d (src/modeling.py:203-204)
    B, T, J, d = 6, 7, 3, 64
    g = torch.Generator(device="cuda").manual_seed(4)
    A = torch.randn(B, 4 * J * d, generator=g, device="cuda").to(BF16)
    W = (torch.randn(4 * J * d, J * d, generator=g, device="cuda") * 0.05).to(BF16)
    res = torch.randn(B, J * d, generator=g, device="cuda")
    x = torch.zeros(B, T, d, device="cuda")
    ops.gemm(A, W, B, J * d, 4 * J * d, b_mn=True, out_f32=x, ldo_f32=T * d, residual=res)
    ref = A.float() @ W.float() + res
    assert torch.allclose(x[:, :J].reshape(B, J * d), ref, rtol=1e-5, atol=1e-3)
    assert torch.equal(x[:, J:], torch.zeros(B, T - J, d, device="cuda"))


def test_gemm_rejects_bad_arguments(ops):
    from jumbo_mae_tpu_b200._lib import JmError
    A = torch.zeros(8, 12, dtype=BF16, device="cuda")
    B = torch.zeros(12, 8, dtype=BF16, device="cuda")
    with pytest.raises(JmError):
        ops.gemm(A, B, 8, 8, 12, b_mn=True, out_f32=torch.zeros(8, 8, device="cuda"), lda=12)   # ld % 8 != 0
    with pytest.raises(JmError):
        ops.gemm(A, B, 8, 8, 12, b_mn=True)                                                     # no output


# ------------------------------------------------------------------ attention
def _attn_ref(qkv, B, T, H, hd):
    """Attention core of oracle.attention (src/modeling.py:136-137) on the stored (already scaled) q."""
    d = H * hd
    q, k, v = [qkv.float()[:, i * d:(i + 1) * d].reshape(B, T, H, hd) for i in range(3)]
    z = torch.einsum("bqhd,bkhd->bhqk", q, k)
    p = torch.softmax(z, dim=-1)
    return torch.einsum("bhqk,bkhd->bqhd", p, v).reshape(B * T, d), torch.logsumexp(z, dim=-1)


ATTN_CASES = [(3, 52, 12, 64), (1, 199, 2, 64), (2, 199, 16, 32), (2, 67, 4, 80), (1, 259, 4, 32), (2, 7, 2, 32), (2, 19, 2, 32),
              (1, 1, 1, 64), (1, 64, 2, 64)]


@pytest.mark.parametrize("B,T,H,hd", ATTN_CASES)
def test_attention_fwd(ops, B, T, H, hd):
    d = H * hd
    g = torch.Generator(device="cuda").manual_seed(T)
    qkv = torch.randn(B * T, 3 * d, generator=g, device="cuda")
    qkv[:, :d] *= hd ** -0.5
    qkv = qkv.to(BF16)
    out = torch.full((B * T, d), float("nan"), dtype=BF16, device="cuda")
    lse = torch.full((B, H, T), float("nan"), device="cuda")
    ops.attention_fwd(qkv, B, T, H, hd, out, lse)
    ref, ref_lse = _attn_ref(qkv, B, T, H, hd)
    # P is rounded to bf16 before P.V and the output is bf16: 2^-7 of the value scale
    assert (out.float() - ref).abs().max().item() <= 2 ** -6 * ref.abs().max().item()
    assert torch.allclose(lse, ref_lse, rtol=1e-4, atol=1e-4)


@pytest.mark.parametrize("B,T,H,hd", ATTN_CASES)
def test_attention_bwd(ops, B, T, H, hd):
    d = H * hd
    g = torch.Generator(device="cuda").manual_seed(T + 1)
    qkv = torch.randn(B * T, 3 * d, generator=g, device="cuda")
    scale = hd ** -0.5
    qkv[:, :d] *= scale
    qkv = qkv.to(BF16)
    dout = torch.randn(B * T, d, generator=g, device="cuda").to(BF16)
    out = torch.empty(B * T, d, dtype=BF16, device="cuda")
    lse = torch.empty(B, H, T, device="cuda")
    ops.attention_fwd(qkv, B, T, H, hd, out, lse)
    dqkv = torch.full((B * T, 3 * d), float("nan"), dtype=BF16, device="cuda")
    dbias = torch.ones(3 * d, device="cuda")
    ops.attention_bwd(qkv, out, dout, lse, B, T, H, hd, scale, dqkv, dbias=dbias)
    x = qkv.float().requires_grad_(True)
    ref, _ = _attn_ref(x, B, T, H, hd)
    (ref * dout.float()).sum().backward()
    want = x.grad.clone()
    want[:, :d] *= scale           # gradient w.r.t. the un-scaled q projection
    got = dqkv.float()
    for i, name in enumerate("qkv"):
        a, b = got[:, i * d:(i + 1) * d], want[:, i * d:(i + 1) * d]
        rel = ((a - b).norm() / (b.norm() + 1e-6)).item()   # T == 1: dq = dk = 0 exactly
        assert rel < 2e-2, f"d{name}: rel l2 {rel}"          # bf16 P / dS operands, bf16 outputs
    # fused q/k/v bias gradient = column sums (the k part is ~0: softmax shift invariance)
    wb = want.sum(0)
    assert (dbias - 1 - wb).norm().item() <= 2e-2 * wb.norm().item() + 1e-3


@pytest.mark.parametrize("M,N,K,a_mn,b_mn", [(256, 256, 64, False, False), (1280, 512, 512, False, True),
                                             (768, 1024, 3000, True, True), (600, 520, 200, False, True)])
def test_gemm_cta_pair_kernel(ops, M, N, K, a_mn, b_mn):
    """The 2-SM (cta_group::2) 256x256 kernel against the reference product."""
    g = torch.Generator(device="cuda").manual_seed(M + N + K)
    ldn, ldk = (N + 7) // 8 * 8, (K + 7) // 8 * 8
    A = torch.randn((K, M) if a_mn else (M, ldk), generator=g, device="cuda").to(BF16)
    A = A if a_mn else A[:, :K]
    B = torch.randn((K, ldn) if b_mn else (N, ldk), generator=g, device="cuda").to(BF16)
    B = B[:, :N] if b_mn else B[:, :K]
    out = torch.full((M, N), float("nan"), device="cuda")
    ops.gemm(A, B, M, N, K, a_mn=a_mn, b_mn=b_mn, out_f32=out, block_n=256, cta_pair=1)
    ref = _gemm_ref(A, B, a_mn, b_mn)
    assert (out - ref).abs().max().item() <= 2e-5 * ref.abs().max().item() * math.sqrt(K)



@pytest.mark.parametrize("variant", [1, 2, 5])
@pytest.mark.parametrize("M,N,K,b_mn", [(10100, 520, 504, True), (20480, 1024, 512, False), (9000, 768, 136, True)])
def test_gemm_pair_kernel_variants(ops, M, N, K, b_mn, variant):
    """The CTA-pair kernel's variants on problems with more tiles than CTA pairs: plain (2), automatic (1: K <= 512
    keeps B resident in shared memory), resident B with column-tile changes (5) -- bf16, f32 + bias + residual, GELU
    forward / backward epilogues against torch."""
    g = torch.Generator(device="cuda").manual_seed(M + N + K)
    A = torch.randn(M, K, generator=g, device="cuda").to(BF16)
    W = (torch.randn((K, N) if b_mn else (N, K), generator=g, device="cuda") * 0.2).to(BF16)
    bias = torch.randn(N, generator=g, device="cuda")
    pre = A.float() @ (W.float() if b_mn else W.float().t())
    kw = dict(b_mn=b_mn, cta_pair=variant)
    out = torch.empty(M, N, dtype=BF16, device="cuda")
    ops.gemm(A, W, M, N, K, out_bf16=out, bias=bias, **kw)
    assert torch.allclose(out.float(), pre + bias, rtol=2 ** -7, atol=2e-3)
    res = torch.randn(M, N, generator=g, device="cuda")
    o32 = torch.empty(M, N, device="cuda")
    ops.gemm(A, W, M, N, K, out_f32=o32, bias=bias, residual=res, **kw)
    assert torch.allclose(o32, pre + bias + res, rtol=1e-4, atol=2e-3)
    dsave = torch.empty(M, N, dtype=BF16, device="cuda")
    gl = torch.empty(M, N, dtype=BF16, device="cuda")
    ops.gemm(A, W, M, N, K, out_bf16=dsave, out2_bf16=gl, bias=bias, act=3, **kw)
    pp = (pre + bias).cpu().requires_grad_(True)
    O.gelu_tanh(pp).sum().backward()
    assert torch.allclose(dsave.float(), pp.grad.cuda(), rtol=2 ** -7, atol=3e-3)
    assert torch.allclose(gl.float(), O.gelu_tanh(pp.detach()).cuda(), rtol=2 ** -7, atol=2e-3)
    dH = torch.empty(M, N, dtype=BF16, device="cuda")
    cs = torch.zeros(N, device="cuda")
    ops.gemm(A, W, M, N, K, out_bf16=dH, aux=dsave, act=4, colsum=cs, **kw)
    ref = pre * dsave.float()
    assert torch.allclose(dH.float(), ref, rtol=2 ** -6, atol=2e-3)
    assert torch.allclose(cs, ref.sum(0), rtol=1e-2, atol=0.5)
    # twice in a row into the same output: the second launch must not depend on shared-memory leftovers
    ops.gemm(A, W, M, N, K, out_bf16=out, bias=bias, **kw)
    assert torch.allclose(out.float(), pre + bias, rtol=2 ** -7, atol=2e-3)


@pytest.mark.parametrize("bn,pair", [(64, -1), (128, -1), (256, -1), (256, 1)])
@pytest.mark.parametrize("M,N,K", [(700, 328, 136), (1300, 512, 64)])
def test_gemm_bf16_epilogues_every_tile_shape(ops, M, N, K, bn, pair):
    """The row-register epilogues (plain bf16 with bias and q-scale, GELU forward, GELU backward with the fused
    column sums) for every tile width and for the CTA-pair kernel: ragged M, N not a multiple of the tile or
    of 32, zeroed rows (mode 1) and skipped rows (mode 2)."""
    T, J = 13, 3
    g = torch.Generator(device="cuda").manual_seed(M + N + bn)
    A = torch.randn(M, K, generator=g, device="cuda").to(BF16)
    W = (torch.randn(K, N, generator=g, device="cuda") * 0.2).to(BF16)
    bias = torch.randn(N, generator=g, device="cuda")
    rows = torch.arange(M, device="cuda") % T
    keep = (rows >= J)[:, None]
    pre = A.float() @ W.float() + bias
    kw = dict(block_n=bn, cta_pair=pair)
    # plain bf16: (acc + bias) * scale on the first 32*k columns, rows < J of every period left untouched
    out = torch.full((M, N), 3.0, dtype=BF16, device="cuda")
    ops.gemm(A, W, M, N, K, b_mn=True, out_bf16=out, bias=bias, col_scale=0.25, col_scale_n=64, row_period=T,
             row_skip=J, row_mask_mode=2, **kw)
    ref = pre.clone()
    ref[:, :64] *= 0.25
    ref = torch.where(keep, ref, torch.full_like(ref, 3.0))
    assert torch.allclose(out.float(), ref, rtol=2 ** -7, atol=1e-3)
    # GELU forward, masked rows zeroed
    h = torch.empty(M, N, dtype=BF16, device="cuda")
    gl = torch.empty(M, N, dtype=BF16, device="cuda")
    ops.gemm(A, W, M, N, K, b_mn=True, out_bf16=h, out2_bf16=gl, bias=bias, act=1, row_period=T, row_skip=J,
             row_mask_mode=1, **kw)
    assert torch.allclose(h.float(), pre * keep, rtol=2 ** -7, atol=1e-3)
    assert torch.allclose(gl.float(), O.gelu_tanh((pre * keep).cpu()).cuda(), rtol=2 ** -7, atol=2e-3)
    # GELU backward + fused column sums
    dY = torch.randn(M, K, generator=g, device="cuda").to(BF16)
    W2 = (torch.randn(N, K, generator=g, device="cuda") * 0.2).to(BF16)
    dH = torch.empty(M, N, dtype=BF16, device="cuda")
    cs = torch.ones(N, device="cuda")
    ops.gemm(dY, W2, M, N, K, b_mn=False, out_bf16=dH, aux=h, act=2, row_period=T, row_skip=J, row_mask_mode=1,
             colsum=cs, **kw)
    hh = h.float().cpu().requires_grad_(True)
    O.gelu_tanh(hh).sum().backward()
    ref = (dY.float() @ W2.float().t()) * hh.grad.cuda() * keep
    assert torch.allclose(dH.float(), ref, rtol=2 ** -6, atol=2e-3)
    assert torch.allclose(cs - 1, ref.sum(0), rtol=1e-2, atol=0.3)   # sums of bf16-rounded values over M rows
    # acts 3 / 4: the forward saves gelu'(pre) instead of pre, the backward multiplies by it
    dsave = torch.empty(M, N, dtype=BF16, device="cuda")
    gl2 = torch.empty(M, N, dtype=BF16, device="cuda")
    ops.gemm(A, W, M, N, K, b_mn=True, out_bf16=dsave, out2_bf16=gl2, bias=bias, act=3, row_period=T, row_skip=J,
             row_mask_mode=1, **kw)
    pp = (pre * keep).cpu().requires_grad_(True)
    O.gelu_tanh(pp).sum().backward()
    assert torch.allclose(dsave.float(), pp.grad.cuda() * keep, rtol=2 ** -7, atol=3e-3)
    assert torch.allclose(gl2.float(), gl.float(), rtol=2 ** -7, atol=2e-3)
    dH2 = torch.empty(M, N, dtype=BF16, device="cuda")
    cs2 = torch.zeros(N, device="cuda")
    ops.gemm(dY, W2, M, N, K, b_mn=False, out_bf16=dH2, aux=dsave, act=4, row_period=T, row_skip=J, row_mask_mode=1,
             colsum=cs2, **kw)
    ref2 = (dY.float() @ W2.float().t()) * dsave.float() * keep
    assert torch.allclose(dH2.float(), ref2, rtol=2 ** -6, atol=2e-3)
    assert torch.allclose(cs2, ref2.sum(0), rtol=1e-2, atol=0.3)


def test_add_bias_cast_strided(ops):
    """jm_add_bias_cast: dst = src + bias over strided views, f32 and bf16 destinations."""
    g = torch.Generator(device="cuda").manual_seed(5)
    rows, cols, lds, ldd, ldd16 = 37, 24, 40, 56, 32
    src = torch.randn(rows, lds, generator=g, device="cuda")
    bias = torch.randn(cols, generator=g, device="cuda")
    d32 = torch.full((rows, ldd), 9.0, device="cuda")
    d16 = torch.full((rows, ldd16), 9.0, dtype=BF16, device="cuda")
    ops.add_bias_cast(src, rows, cols, bias=bias, dst_f32=d32, dst_bf16=d16, lds=lds, ldd=ldd, ldd16=ldd16)
    want = src[:, :cols] + bias
    assert torch.equal(d32[:, :cols], want) and torch.equal(d32[:, cols:], torch.full_like(d32[:, cols:], 9.0))
    assert torch.equal(d16[:, :cols], want.to(BF16)) and torch.equal(d16[:, cols:], torch.full_like(d16[:, cols:], 9.0))
    ops.add_bias_cast(src, rows, cols, dst_bf16=d16, lds=lds, ldd16=ldd16)
    assert torch.equal(d16[:, :cols], src[:, :cols].to(BF16))


@pytest.mark.parametrize("max_norm", [0.5, 1e6])
def test_clip_by_global_norm(ops, max_norm):
    """jm_clip_by_global_norm against optax's rule: unchanged below the threshold, g / ||g|| * max_norm above."""
    g = torch.Generator(device="cuda").manual_seed(3)
    x = torch.randn(1 << 20, generator=g, device="cuda")
    pre = 0.25
    want_norm = (x.double() * pre).norm().item()
    y = x.clone()
    ss = torch.full((1,), 7.0, device="cuda")
    ops.clip_by_global_norm(y, pre, max_norm, ss)
    assert abs(ss.item() - (x.double() ** 2).sum().item()) <= 1e-5 * ss.item()
    if want_norm < max_norm:
        assert torch.equal(y, x)
    else:
        assert torch.allclose(y, x * (max_norm / want_norm), rtol=1e-5, atol=0)
        assert abs((y.double() * pre).norm().item() - max_norm) <= 1e-5 * max_norm


@pytest.mark.parametrize("M,N,K", [(300, 328, 72), (130, 264, 200)])
def test_gemm_outputs_stay_inside_their_views(ops, M, N, K):
    """Ragged problems through the TMA-store / TMA-reduce epilogues: nothing outside the [M, N] view of a larger,
    sentinel-filled buffer is written (rows beyond M, columns beyond N, the gap between row ends and the stride)."""
    g = torch.Generator(device="cuda").manual_seed(M * N + K)
    A = torch.randn(M, K, generator=g, device="cuda").to(BF16)
    W = (torch.randn(K, N, generator=g, device="cuda") * 0.2).to(BF16)
    bias = torch.randn(N, generator=g, device="cuda")
    ref = A.float() @ W.float() + bias
    pad_r, ld = 5, N + 24
    # bf16 output (TMA store of 32x32 tiles)
    big = torch.full((M + 2 * pad_r, ld), 7.0, dtype=BF16, device="cuda")
    out = big[pad_r:pad_r + M, :N]
    ops.gemm(A, W, M, N, K, b_mn=True, out_bf16=out, bias=bias)
    assert torch.allclose(out.float(), ref, rtol=2 ** -7, atol=1e-3)
    mask = torch.ones_like(big, dtype=torch.bool)
    mask[pad_r:pad_r + M, :N] = False
    assert torch.all(big[mask] == 7.0)
    # f32 output with residual (TMA load + store of 32x16 tiles), then accumulation (TMA reduce-add)
    res = torch.randn(M, N, generator=g, device="cuda")
    big32 = torch.full((M + 2 * pad_r, ld), 7.0, device="cuda")
    out32 = big32[pad_r:pad_r + M, :N]
    ops.gemm(A, W, M, N, K, b_mn=True, out_f32=out32, bias=bias, residual=res)
    assert torch.allclose(out32, ref + res, rtol=1e-5, atol=1e-4)
    assert torch.all(big32[mask] == 7.0)
    ops.gemm(A, W, M, N, K, b_mn=True, out_f32=out32, accumulate=True)
    assert torch.allclose(out32, ref + res + (ref - bias), rtol=1e-5, atol=2e-4)
    assert torch.all(big32[mask] == 7.0)


@pytest.mark.parametrize("T,H,hd", [(199, 2, 32), (52, 2, 64), (67, 1, 80)])
def test_attention_outputs_stay_inside_their_buffers(ops, T, H, hd):
    """Sequence lengths that are not multiples of the 16-row strips: the padded rows of the last strip and the
    last image never reach memory (sentinel rows after the end of out / lse / dqkv stay intact)."""
    B, d, guard = 3, H * hd, 24
    g = torch.Generator(device="cuda").manual_seed(T)
    qkv = (torch.randn(B * T + guard, 3 * d, generator=g, device="cuda") * 0.5).to(BF16)
    dout = torch.randn(B * T + guard, d, generator=g, device="cuda").to(BF16)
    out = torch.full((B * T + guard, d), 5.0, dtype=BF16, device="cuda")
    lse = torch.full((B * H * T + guard,), 5.0, device="cuda")
    dqkv = torch.full((B * T + guard, 3 * d), 5.0, dtype=BF16, device="cuda")
    ops.attention_fwd(qkv[:B * T], B, T, H, hd, out[:B * T], lse[:B * H * T].view(B, H, T))
    ops.attention_bwd(qkv[:B * T], out[:B * T], dout[:B * T], lse[:B * H * T].view(B, H, T), B, T, H, hd, hd ** -0.5,
                      dqkv[:B * T])
    torch.cuda.synchronize()
    assert torch.all(out[B * T:] == 5.0) and torch.all(lse[B * H * T:] == 5.0) and torch.all(dqkv[B * T:] == 5.0)
    assert torch.isfinite(out[:B * T].float()).all() and torch.isfinite(dqkv[:B * T].float()).all()


# ------------------------------------------------------------------ fine-tune classification loss (8(f) rank 1)
@pytest.mark.parametrize("B,C,soft,mix,alpha", [
    (64, 1000, False, False, 0.1), (64, 1000, False, True, 0.1), (33, 1000, True, True, 0.0),
    (8, 3, False, True, 0.1), (5, 8192, False, False, 0.0), (1, 1, False, False, 0.0), (16, 37, True, False, 0.2),
])
def test_softmax_ce_matches_oracle(ops, B, C, soft, mix, alpha):
    """jm_softmax_ce against the numpy restatement of src/finetuning.py:90-106.  Labels and hit flags: bit exact.
    Loss: rel 1e-5 (f32 exp/log of CUDA vs numpy).  Gradient: atol 1e-7 + rel 1e-4 against the f64 closed form."""
    rng = np.random.default_rng(B * 7919 + C)
    logits = (rng.standard_normal((B, C)) * 3).astype(np.float32)
    if C >= 8:
        logits[0, 5] = logits[0, 2] = logits[0].max() + 1       # a tie for the first place: index 2 must win
    idx = rng.integers(0, C, B).astype(np.int32)
    if soft:
        labels = rng.random((B, C), dtype=np.float32)
        labels /= labels.sum(-1, keepdims=True)
    else:
        labels = idx
    perm = rng.permutation(B).astype(np.int32) if mix else None
    ratio = float(np.float32(rng.beta(0.8, 0.8))) if mix else 1.0
    want = O.finetune_metrics(logits, labels, C, alpha, perm=perm, ratio=ratio)
    metrics, loss, acc1, acc5, dlogits = ops.softmax_ce(
        dev(torch.from_numpy(logits)), dev(torch.from_numpy(labels)), label_smoothing=alpha,
        perm=None if perm is None else dev(torch.from_numpy(perm)), ratio=ratio)
    assert np.allclose(loss.cpu().numpy(), want["loss"], rtol=1e-5, atol=1e-6)
    assert np.array_equal(acc1.cpu().numpy(), want["acc1"])
    assert np.array_equal(acc5.cpu().numpy(), want["acc5"])
    got_m = metrics.cpu().numpy()
    assert np.allclose(got_m, [want["loss"].mean(), want["acc1"].mean(), want["acc5"].mean()], rtol=1e-5, atol=1e-6)
    g = O.softmax_ce_grad(logits, want["labels"])
    assert np.allclose(dlogits.cpu().numpy(), g, rtol=1e-4, atol=1e-7)


def test_softmax_ce_eval_mode_and_errors(ops):
    """Evaluation: no smoothing, no mixing, no gradient buffer; bad arguments are refused with a message."""
    from jumbo_mae_tpu_b200._lib import JmError
    rng = np.random.default_rng(11)
    logits = rng.standard_normal((32, 1000)).astype(np.float32)
    idx = rng.integers(0, 1000, 32).astype(np.int32)
    idx[:8] = logits[:8].argmax(-1)                              # eight certain top-1 hits
    want = O.finetune_metrics(logits, idx, 1000, 0.1, det=True)
    metrics, loss, acc1, acc5, dlogits = ops.softmax_ce(dev(torch.from_numpy(logits)), dev(torch.from_numpy(idx)),
                                                        want_grad=False)
    assert dlogits is None and acc1[:8].sum().item() == 8
    assert np.allclose(loss.cpu().numpy(), want["loss"], rtol=1e-5)
    assert np.array_equal(acc5.cpu().numpy(), want["acc5"])
    with pytest.raises(JmError, match="label_smoothing"):
        ops.softmax_ce(dev(torch.from_numpy(logits)), dev(torch.from_numpy(idx)), label_smoothing=1.5)
    with pytest.raises(JmError, match="8192"):
        ops.softmax_ce(torch.zeros(2, 9000, device="cuda"), torch.zeros(2, dtype=torch.int32, device="cuda"))


@pytest.mark.parametrize("mode", ["mixup", "cutmix"])
def test_patchify_mix_matches_oracle(ops, mode):
    """jm_patchify_mix against Mixup.apply_mixup / apply_cutmix of the oracle on the normalised images, followed by
    extract_patches: exact after the same bf16 rounding (the mix is done in f32 with the array ops' roundings)."""
    S, p, J, B = 32, 8, 3, 5
    N, P, ld = (S // p) ** 2, p * p * 3, 200
    g = torch.Generator().manual_seed(4)
    images = torch.randint(0, 256, (B, 3, S, S), generator=g, dtype=torch.uint8)
    perm = np.array([3, 0, 4, 1, 2], dtype=np.int32)
    norm = O.preprocess(images).numpy()
    lab = np.eye(B, dtype=np.float32)
    out = torch.full((B * (J + N), ld), 7.0, dtype=BF16, device="cuda")
    if mode == "mixup":
        ratio = 0.3
        mixed, _ = O.apply_mixup(norm, lab, ratio, perm)
        ops.patchify_mix(dev(images), p, None, N, J + N, J, out, dev(torch.from_numpy(perm)), ratio=ratio)
    else:
        xm, ym, _ = O.random_bounding_box(0.6, 0.4, 0.7, S, S)
        assert 0 < xm.sum() < S and 0 < ym.sum() < S
        mixed, _ = O.apply_cutmix(norm, lab, 0.6, 0.4, 0.7, perm)
        ops.patchify_mix(dev(images), p, None, N, J + N, J, out, dev(torch.from_numpy(perm)),
                         xmask=dev(torch.from_numpy(xm.astype(np.uint8))), ymask=dev(torch.from_numpy(ym.astype(np.uint8))))
    want = O.extract_patches(torch.from_numpy(mixed), p)
    got = out.float().cpu().reshape(B, J + N, ld)
    assert torch.equal(got[:, :J], torch.zeros(B, J, ld))
    assert torch.equal(got[:, J:, P:], torch.zeros(B, N, ld - P))
    assert torch.equal(got[:, J:, :P], want.to(BF16).float())
    assert not torch.equal(got[:, J:, :P], O.extract_patches(torch.from_numpy(norm), p).to(BF16).float())


def test_gemm_row_scale_epilogue(ops):
    """out = row_scale[row // period] * (A W + bias) + residual: the DropPath-scaled residual branch (f32 epilogue),
    also as a split-free accumulate into a strided output (the Jumbo CLS rows)."""
    g = torch.Generator(device="cuda").manual_seed(5)
    M, N, K, T = 12 * 52, 768, 512, 52
    A = torch.randn(M, K, generator=g, device="cuda").to(BF16)
    W = torch.randn(K, N, generator=g, device="cuda").to(BF16)
    bias = torch.randn(N, generator=g, device="cuda")
    res = torch.randn(M, N, generator=g, device="cuda")
    scale = torch.tensor([0.0, 1.25] * 6, device="cuda")
    out = torch.empty(M, N, device="cuda")
    ops.gemm(A, W, M, N, K, b_mn=True, bias=bias, residual=res, out_f32=out, row_scale=scale, row_scale_period=T)
    want = scale.repeat_interleave(T)[:, None] * (A.float() @ W.float() + bias) + res
    assert (out - want).abs().max().item() <= 2e-5 * K ** 0.5 * want.abs().max().item()
    assert torch.equal(out[:T], res[:T])                                       # a dropped image keeps its residual exactly


def test_scale_rows_bf16(ops):
    x = torch.randn(3 * 7, 64, device="cuda").to(BF16)
    a, b = torch.tensor([2.0, 0.0, 0.5], device="cuda"), torch.tensor([0.0, 3.0, 1.0], device="cuda")
    want = x.float().clone().view(3, 7, 64)
    want[:, :2] *= b[:, None, None]
    want[:, 2:] *= a[:, None, None]
    ops.scale_rows_bf16(x, 21, 64, a, 7, scale_cls=b, skip=2)
    assert torch.equal(x.float(), want.view(21, 64).to(BF16).float())
