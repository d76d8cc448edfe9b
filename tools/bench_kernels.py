"""Times the step's kernels in isolation at ViT-B/16 micro-batch shapes (CUDA events, after warm-up).
Development aid: prints per-shape TFLOP/s or GB/s and the per-micro-step total they imply."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from jumbo_mae_tpu_b200 import ops  # noqa: E402

BF16 = torch.bfloat16
B = int(os.environ.get("B", "512"))
ONLY = os.environ.get("ONLY", "")
BN = int(os.environ.get("BN", "0"))
SAVED = os.environ.get("SAVED_GRAD", "1") != "0"   # acts 3/4 (what the engine uses) instead of 1/2


def timeit(fn, reps=20, warm=3):
    for _ in range(warm):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def r(*shape, dt=BF16):
    return (torch.randn(*shape, device="cuda") * 0.1).to(dt)


def gemm_cases():
    d, dd, Te, Td, J = 768, 512, 52, 199, 3
    Me, Md = B * Te, B * Td
    cases = []  # name, count per micro-step, M, N, K, kind
    for nm, M, dim, cnt in (("enc", Me, d, 12), ("dec", Md, dd, 8)):
        cases += [
            (f"{nm} qkv fwd", cnt, M, 3 * dim, dim, "fwd_bf16"),
            (f"{nm} proj fwd", cnt, M, dim, dim, "fwd_res"),
            (f"{nm} ffn1 fwd", cnt, M, 4 * dim, dim, "fwd_gelu"),
            (f"{nm} ffn2 fwd", cnt, M, dim, 4 * dim, "fwd_res"),
            (f"{nm} ffn2 dX", cnt, M, 4 * dim, dim, "dx_gelu"),
            (f"{nm} ffn1 dX", cnt, M, dim, 4 * dim, "dx"),
            (f"{nm} proj dX", cnt, M, dim, dim, "dx"),
            (f"{nm} qkv dX", cnt, M, dim, 3 * dim, "dx"),
            (f"{nm} qkv dW", cnt, dim, 3 * dim, M, "dw"),
            (f"{nm} proj dW", cnt, dim, dim, M, "dw"),
            (f"{nm} ffn1 dW", cnt, dim, 4 * dim, M, "dw"),
            (f"{nm} ffn2 dW", cnt, 4 * dim, dim, M, "dw"),
        ]
    Jd = J * d
    cases += [
        ("jumbo w1 fwd", 12, B, 4 * Jd, Jd, "fwd_gelu"), ("jumbo w2 fwd", 12, B, Jd, 4 * Jd, "fwd_res"),
        ("jumbo w2 dX", 12, B, 4 * Jd, Jd, "dx_gelu"), ("jumbo w1 dX", 12, B, Jd, 4 * Jd, "dx"),
        ("jumbo w1 dW", 12, Jd, 4 * Jd, B, "dw"), ("jumbo w2 dW", 12, 4 * Jd, Jd, B, "dw"),
        ("embed fwd", 1, Me, d, 768, "fwd_res"), ("embed dW", 1, 768, d, Me, "dw"),
        ("dec_proj fwd", 1, Me, dd, d, "fwd_f32"), ("dec_proj dX", 1, Me, d, dd, "dx"), ("dec_proj dW", 1, d, dd, Me, "dw"),
        ("head fwd", 1, Md, 768, dd, "fwd_f32"), ("head dX", 1, Md, dd, 768, "dx"), ("head dW", 1, dd, 768, Md, "dw"),
    ]
    return cases


def run_gemm(name, M, N, K, kind):
    if kind == "dw":
        X, dY = r(K, M), r(K, N)
        out = torch.zeros(M, N, device="cuda")
        return lambda: ops.gemm(X, dY, M, N, K, a_mn=True, b_mn=True, out_f32=out, accumulate=True)
    A = r(M, K)
    bias = torch.zeros(N, device="cuda")
    if kind.startswith("fwd"):
        W = r(K, N)
        if kind == "fwd_bf16":
            o = torch.empty(M, N, dtype=BF16, device="cuda")
            return lambda: ops.gemm(A, W, M, N, K, b_mn=True, out_bf16=o, bias=bias, col_scale=0.125, col_scale_n=N // 3, block_n=BN)
        if kind == "fwd_gelu":
            h, g = torch.empty(M, N, dtype=BF16, device="cuda"), torch.empty(M, N, dtype=BF16, device="cuda")
            return lambda: ops.gemm(A, W, M, N, K, b_mn=True, out_bf16=h, out2_bf16=g, bias=bias, act=3 if SAVED else 1, block_n=BN)
        if kind == "fwd_res":
            res, o = torch.zeros(M, N, device="cuda"), torch.empty(M, N, device="cuda")
            return lambda: ops.gemm(A, W, M, N, K, b_mn=True, out_f32=o, bias=bias, residual=res, block_n=BN)
        o = torch.empty(M, N, device="cuda")
        return lambda: ops.gemm(A, W, M, N, K, b_mn=True, out_f32=o, bias=bias)
    W = r(N, K)
    o = torch.empty(M, N, dtype=BF16, device="cuda")
    if kind == "dx_gelu":
        aux = r(M, N)
        cs = torch.zeros(N, device="cuda") if os.environ.get("NO_CS") is None else None
        return lambda: ops.gemm(A, W, M, N, K, b_mn=False, out_bf16=o, aux=aux, act=4 if SAVED else 2, colsum=cs)
    return lambda: ops.gemm(A, W, M, N, K, b_mn=False, out_bf16=o, block_n=BN)


def main():
    total_ms, total_flop = 0.0, 0.0
    if ONLY in ("", "gemm"):
        for name, cnt, M, N, K, kind in gemm_cases():
            ms = timeit(run_gemm(name, M, N, K, kind))
            fl = 2.0 * M * N * K
            total_ms += ms * cnt
            total_flop += fl * cnt
            print(f"{name:16s} x{cnt:2d} M={M:6d} N={N:5d} K={K:6d} {kind:9s} {ms * 1e3:8.1f} us {fl / ms / 1e9:7.1f} TF/s")
        print(f"GEMM total per micro-step: {total_ms:.2f} ms, {total_flop / total_ms / 1e9:.1f} TF/s average")
    if ONLY in ("", "attn"):
        att_ms = 0.0
        for nm, T, H, hd, cnt in (("enc", 52, 12, 64, 12), ("dec", 199, 16, 32, 8)):
            d = H * hd
            qkv = r(B * T, 3 * d)
            out, dout = torch.empty(B * T, d, dtype=BF16, device="cuda"), r(B * T, d)
            lse = torch.empty(B, H, T, device="cuda")
            dqkv = torch.empty(B * T, 3 * d, dtype=BF16, device="cuda")
            f = timeit(lambda: ops.attention_fwd(qkv, B, T, H, hd, out, lse))
            b = timeit(lambda: ops.attention_bwd(qkv, out, dout, lse, B, T, H, hd, hd ** -0.5, dqkv))
            fl = 4.0 * T * T * hd * B * H
            att_ms += (f + b) * cnt
            print(f"attn {nm} T={T} hd={hd}: fwd {f * 1e3:.0f} us ({fl / f / 1e9:.0f} TF/s)  bwd {b * 1e3:.0f} us "
                  f"({2.5 * fl / b / 1e9:.0f} TF/s algorithmic)")
        print(f"attention total per micro-step: {att_ms:.2f} ms")
        total_ms += att_ms
    if ONLY in ("", "ln"):
        ln_ms = 0.0
        for nm, M, D, cnt in (("enc", B * 52, 768, 24), ("dec", B * 199, 512, 16)):
            x, gm, bt = torch.randn(M, D, device="cuda"), torch.ones(D, device="cuda"), torch.zeros(D, device="cuda")
            y = torch.empty(M, D, dtype=BF16, device="cuda")
            mean, rstd = torch.empty(M, device="cuda"), torch.empty(M, device="cuda")
            f = timeit(lambda: ops.layernorm_fwd(x, gm, bt, 1e-6, y, None, mean, rstd))
            dy, dres = r(M, D), torch.randn(M, D, device="cuda")
            dx, dxb = torch.empty(M, D, device="cuda"), torch.empty(M, D, dtype=BF16, device="cuda")
            dg, db = torch.zeros(D, device="cuda"), torch.zeros(D, device="cuda")
            ws = torch.empty(ops.layernorm_bwd_workspace(M, D), dtype=torch.uint8, device="cuda")
            b = timeit(lambda: ops.layernorm_bwd(dy, x, mean, rstd, gm, dres, dx, dxb, dg, db, ws))
            fb, bb = M * D * 6, M * D * 16
            ln_ms += (f + b) * cnt
            print(f"ln {nm} [{M}x{D}]: fwd {f * 1e3:.0f} us ({fb / f / 1e6:.0f} GB/s)  bwd {b * 1e3:.0f} us ({bb / b / 1e6:.0f} GB/s)")
            cs = timeit(lambda: ops.colsum(dy, M, D, dg))
            print(f"colsum [{M}x{D}] {cs * 1e3:.0f} us ({M * D * 2 / cs / 1e6:.0f} GB/s)")
        print(f"layernorm total per micro-step: {ln_ms:.2f} ms")
        total_ms += ln_ms
    if ONLY in ("", "misc"):
        N, n, J, P, dd = 196, 49, 3, 768, 512
        images = torch.randint(0, 256, (B, 3, 224, 224), dtype=torch.uint8, device="cuda")
        noise = torch.rand(N, device="cuda")
        ids_shuffle = torch.argsort(noise).to(torch.int32)
        ids_restore = torch.argsort(ids_shuffle.long()).to(torch.int32)
        mask_row = (ids_restore >= n).float()
        pred = torch.randn(B * (J + N), P, device="cuda")
        dpred = torch.empty(B * (J + N), P, dtype=BF16, device="cuda")
        loss = torch.zeros(1, device="cuda")
        ws = torch.empty(ops.patch_loss_workspace(B, N), dtype=torch.uint8, device="cuda")
        t = timeit(lambda: ops.patch_loss(pred, images, mask_row, 16, J, False, 1.0, loss, dpred, ws))
        by = pred.numel() * 4 * 0.75 + dpred.numel() * 2 + images.numel() * 0.75
        print(f"patch_loss: {t * 1e3:.0f} us ({by / t / 1e6:.0f} GB/s)")
        patches = torch.empty(B * (J + n), P, dtype=BF16, device="cuda")
        t2 = timeit(lambda: ops.patchify(images, 16, ids_shuffle, n, J + n, J, patches))
        print(f"patchify: {t2 * 1e3:.0f} us ({(patches.numel() * 2 + images.numel() * 0.25) / t2 / 1e6:.0f} GB/s)")
        dout = torch.randn(B * (J + N), dd, device="cuda")
        dy = torch.empty(B * (J + n), dd, dtype=BF16, device="cuda")
        dm = torch.zeros(dd, device="cuda")
        t3 = timeit(lambda: ops.unshuffle_bwd(dout, ids_restore, B, N, n, J, dd, dy, dm))
        print(f"unshuffle_bwd: {t3 * 1e3:.0f} us ({(dout.numel() * 4 + dy.numel() * 2) / t3 / 1e6:.0f} GB/s)")
        total_ms += t + t2 + t3
    print(f"sum of timed kernels per micro-step: {total_ms:.2f} ms -> {B / total_ms * 1e3:.0f} images/s upper bound")


if __name__ == "__main__":
    main()
