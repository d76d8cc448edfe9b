"""Times the decoder / encoder GELU-forward GEMMs on an experiment build (tools/_prof/<lib>) at sustained clocks."""
import os, sys, torch
sys.path.insert(0, ".")
from jumbo_mae_tpu_b200 import _lib
_lib.LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_prof", sys.argv[1])
from jumbo_mae_tpu_b200 import ops
dev, BF16 = "cuda", torch.bfloat16
for (M, N, K) in ((203776, 2048, 512), (53248, 3072, 768)):
    A = (torch.randn(M, K, device=dev) * 0.5).to(BF16)
    B = (torch.randn(K, N, device=dev) * 0.05).to(BF16)
    h = torch.empty(M, N, dtype=BF16, device=dev); g = torch.empty_like(h)
    bias = torch.zeros(N, device=dev)
    for name, fn in (("gelu_fwd", lambda: ops.gemm(A, B, M, N, K, b_mn=True, out_bf16=h, out2_bf16=g, bias=bias, act=3, cta_pair=2)),
                     ("bf16", lambda: ops.gemm(A, B, M, N, K, b_mn=True, out_bf16=h, cta_pair=2))):
        for _ in range(300):
            fn()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); a.record()
        for _ in range(50):
            fn()
        b.record(); torch.cuda.synchronize()
        ms = a.elapsed_time(b) / 50
        print(f"{sys.argv[1]:16s} {M} {N} {K} {name:9s} {ms*1e3:8.1f} us {2.0*M*N*K/ms/1e9:7.1f} TF/s", flush=True)
