"""Runs one GEMM configuration a few times (for ncu captures). CASE env: name from bench_kernels.gemm_cases()."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench_kernels as bk  # noqa: E402

want = os.environ.get("CASE", "dec proj dX")
for name, cnt, M, N, K, kind in bk.gemm_cases():
    if name == want:
        fn = bk.run_gemm(name, M, N, K, kind)
        for _ in range(int(os.environ.get("REPS", "4"))):
            fn()
        torch.cuda.synchronize()
        print("ran", name, M, N, K, kind)
