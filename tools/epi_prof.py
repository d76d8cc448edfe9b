"""Runs the decoder GELU-forward and plain bf16 GEMMs on the JM_EPI_PROF build (tools/epi_prof_build.sh)."""
import os, sys, torch
sys.path.insert(0, ".")
from jumbo_mae_tpu_b200 import _lib
_lib.LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_prof", "libjm_prof.so")
from jumbo_mae_tpu_b200 import ops

dev, BF16 = "cuda", torch.bfloat16
M, N, K = 203776, 2048, 512
A = (torch.randn(M, K, device=dev) * 0.5).to(BF16)
B = (torch.randn(K, N, device=dev) * 0.5).to(BF16)
h = torch.empty(M, N, dtype=BF16, device=dev); g = torch.empty_like(h)
bias = torch.zeros(N, device=dev)
for name, fn in (("gelu_fwd", lambda: ops.gemm(A, B, M, N, K, b_mn=True, out_bf16=h, out2_bf16=g, bias=bias, act=3)),
                 ("bf16", lambda: ops.gemm(A, B, M, N, K, b_mn=True, out_bf16=h)),
                 ("gelu_bwd", lambda: ops.gemm(A, B, M, N, K, b_mn=True, out_bf16=h, aux=g, act=4))):
    print("==", name, flush=True)
    for _ in range(3):
        fn()
        torch.cuda.synchronize()
