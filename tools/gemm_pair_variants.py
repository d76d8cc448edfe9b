"""The plain CTA-pair GEMM (cta_pair=2) against the automatic choice (cta_pair=1: resident B for K <= 512):
correctness against torch and time per launch, at the shapes of the ViT-B step."""
import sys, torch
sys.path.insert(0, ".")
from jumbo_mae_tpu_b200 import ops

dev = "cuda"
BF16 = torch.bfloat16
torch.manual_seed(0)

def timeit(fn, reps=20):
    for _ in range(3 if "cold" in sys.argv else 400):     # sustained clocks (power cap), as inside a training step
        fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); a.record()
    for _ in range(reps):
        fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps

# (M, N, K, a_mn, b_mn, kind)
shapes = [
    (203776, 1536, 512, False, True, "bf16"),
    (203776, 2048, 512, False, True, "gelu"),
    (203776, 512, 2048, False, False, "bf16"),
    (203776, 512, 512, False, True, "bf16"),
    (53248, 3072, 768, False, True, "gelu"),
    (53248, 2304, 768, False, True, "bf16"),
    (53248, 768, 3072, False, False, "bf16"),
    (53248, 768, 768, False, False, "bf16"),
    (512, 2048, 203776, True, True, "dw"),
    (768, 3072, 53248, True, True, "dw"),
    (3072, 768, 53248, True, True, "dw"),
    (768, 2304, 53248, True, True, "dw"),
    (203776, 2048, 512, False, False, "gelubwd"),
    (203776, 512, 512, False, True, "f32res"),
    (203776, 768, 512, False, True, "f32res"),
    (1000, 512, 320, False, True, "bf16"),        # ragged edges
    (70000, 1280, 504, False, True, "gelu"),
]
small = len(sys.argv) > 1 and sys.argv[1] == "small"
for M, N, K, a_mn, b_mn, kind in shapes:
    if small and max(M, K) > 60000:
        M, K = (M if M < 60000 else 8192), (K if K < 60000 else 8192)
    A = (torch.randn((K, M) if a_mn else (M, K), device=dev) * 0.5).to(BF16)
    B = (torch.randn((K, N) if b_mn else (N, K), device=dev) * 0.5).to(BF16)
    Af = (A.t() if a_mn else A).float()
    Bf = (B if b_mn else B.t()).float()
    res = {}
    for mode in (2, 1):
        kw = dict(a_mn=a_mn, b_mn=b_mn, cta_pair=mode)
        if kind == "dw":
            out = torch.zeros(M, N, device=dev)
            run = lambda: ops.gemm(A, B, M, N, K, out_f32=out, accumulate=True, **kw)
            run(); torch.cuda.synchronize()
            got = out.clone(); out.zero_()
        elif kind == "gelu":
            h = torch.empty(M, N, dtype=BF16, device=dev); g = torch.empty_like(h)
            bias = torch.zeros(N, device=dev)
            run = lambda: ops.gemm(A, B, M, N, K, out_bf16=h, out2_bf16=g, bias=bias, act=1, **kw)
            run(); torch.cuda.synchronize()
            got = h.float()
        elif kind == "gelubwd":
            aux = (torch.randn(M, N, device=dev)).to(BF16)
            o = torch.empty(M, N, dtype=BF16, device=dev)
            cs = torch.zeros(N, device=dev)
            run = lambda: ops.gemm(A, B, M, N, K, out_bf16=o, aux=aux, act=4, colsum=cs, **kw)
            run(); torch.cuda.synchronize()
            got = o.float()
            Bf = Bf  # reference below multiplies by aux
        elif kind == "f32res":
            res_in = torch.randn(M, N, device=dev)
            bias = torch.randn(N, device=dev)
            o32 = torch.empty(M, N, device=dev)
            run = lambda: ops.gemm(A, B, M, N, K, out_f32=o32, residual=res_in, bias=bias, **kw)
            run(); torch.cuda.synchronize()
            got = o32.clone()
        else:
            o = torch.empty(M, N, dtype=BF16, device=dev)
            run = lambda: ops.gemm(A, B, M, N, K, out_bf16=o, **kw)
            run(); torch.cuda.synchronize()
            got = o.float()
        rows = slice(0, min(M, 4096))
        post = lambda r, sl: (r * aux[sl].float() if kind == "gelubwd" else (r + bias + res_in[sl] if kind == "f32res" else r))
        ref = post(Af[rows] @ Bf, rows)
        err = ((got[rows] - ref).norm() / ref.norm()).item()
        last = slice(max(0, M - 512), M)
        ref2 = post(Af[last] @ Bf, last)
        err2 = ((got[last] - ref2).norm() / ref2.norm()).item()
        ms = timeit(run)
        res[mode] = (ms, max(err, err2))
    tf = lambda ms: 2.0 * M * N * K / ms / 1e9
    print(f"{M:7d} {N:5d} {K:7d} {kind:5s}  plain {res[2][0]*1e3:8.1f} us {tf(res[2][0]):7.1f} TF/s err {res[2][1]:.2e} | "
          f"auto {res[1][0]*1e3:8.1f} us {tf(res[1][0]):7.1f} TF/s err {res[1][1]:.2e}", flush=True)
