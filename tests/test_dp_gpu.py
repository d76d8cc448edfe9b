"""Data-parallel step on real GPUs: 2 NCCL ranks (one process per GPU) run tests/_dp_worker.py.

Skipped on a box with fewer than two GPUs; the host-side logic is covered on CPU by tests/test_parallel_cpu.py."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_two_rank_step_matches_oracle_and_stays_in_lockstep():
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
           "127.0.0.1", "--master-port", "29517", os.path.join(ROOT, "tests", "_dp_worker.py")]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert res.returncode == 0, res.stdout[-4000:] + "\n" + res.stderr[-4000:]
    assert "dp_worker ok" in res.stdout


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_one_process_two_devices_gemm():
    """A host that drives several GPUs from ONE process (the reference's jax.pmap over local devices): the GEMM's
    one-time set-up (dynamic shared-memory opt-in, cluster occupancy) is cached per device, so the same kernel
    instantiation must work on device 1 after it has run on device 0."""
    from jumbo_mae_tpu_b200 import ops
    for dev in (0, 1, 0):
        with torch.cuda.device(dev):
            g = torch.Generator(device="cpu").manual_seed(dev)
            A = torch.randn(512, 256, generator=g).to(torch.bfloat16).cuda(dev)
            B = torch.randn(256, 512, generator=g).to(torch.bfloat16).cuda(dev)
            for pair in (-1, 1):                               # single-CTA and CTA-pair kernels
                out = torch.empty(512, 512, dtype=torch.bfloat16, device=f"cuda:{dev}")
                ops.gemm(A, B, 512, 512, 256, b_mn=True, out_bf16=out, block_n=256, cta_pair=pair)
                torch.cuda.synchronize(dev)
                want = A.float() @ B.float()
                assert (out.float() - want).abs().max().item() <= 2e-2 * want.abs().max().item()
