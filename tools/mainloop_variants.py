"""Times long-K weight-gradient GEMMs and a K=512 forward GEMM on an experiment build (tools/_prof/<lib>)."""
import os, sys, torch
sys.path.insert(0, ".")
from jumbo_mae_tpu_b200 import _lib
_lib.LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_prof", sys.argv[1])
from jumbo_mae_tpu_b200 import ops
dev, BF16 = "cuda", torch.bfloat16
cold = "cold" in sys.argv
for (M, N, K, a_mn, kind) in ((512, 2048, 203776, True, "dw"), (203776, 1536, 512, False, "bf16"), (53248, 768, 3072, False, "bf16")):
    A = (torch.randn((K, M) if a_mn else (M, K), device=dev) * 0.5).to(BF16)
    B = (torch.randn(K, N, device=dev) * 0.05).to(BF16)
    if kind == "dw":
        out = torch.zeros(M, N, device=dev)
        fn = lambda: ops.gemm(A, B, M, N, K, a_mn=True, b_mn=True, out_f32=out, accumulate=True, cta_pair=2)
    else:
        o = torch.empty(M, N, dtype=BF16, device=dev)
        fn = lambda: ops.gemm(A, B, M, N, K, b_mn=True, out_bf16=o, cta_pair=2)
    for _ in range(5 if cold else 300):
        fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); a.record()
    for _ in range(20):
        fn()
    b.record(); torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 20
    print(f"{sys.argv[1]:14s} {'cold' if cold else 'hot ':4s} {M} {N} {K} {kind:5s} {ms*1e3:8.1f} us {2.0*M*N*K/ms/1e9:7.1f} TF/s", flush=True)
