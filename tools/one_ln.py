"""Runs LayerNorm forward/backward at the decoder shape a few times (for ncu captures)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from jumbo_mae_tpu_b200 import ops
M, D = 512 * 199, 512
x, gm, bt = torch.randn(M, D, device="cuda"), torch.ones(D, device="cuda"), torch.zeros(D, device="cuda")
y = torch.empty(M, D, dtype=torch.bfloat16, device="cuda")
mean, rstd = torch.empty(M, device="cuda"), torch.empty(M, device="cuda")
dy, dres = torch.randn(M, D, device="cuda").to(torch.bfloat16), torch.randn(M, D, device="cuda")
dx, dxb = torch.empty(M, D, device="cuda"), torch.empty(M, D, dtype=torch.bfloat16, device="cuda")
dg, db = torch.zeros(D, device="cuda"), torch.zeros(D, device="cuda")
ws = torch.empty(ops.layernorm_bwd_workspace(M, D), dtype=torch.uint8, device="cuda")
for _ in range(3):
    ops.layernorm_fwd(x, gm, bt, 1e-6, y, None, mean, rstd)
    ops.layernorm_bwd(dy, x, mean, rstd, gm, dres, dx, dxb, dg, db, ws)
torch.cuda.synchronize(); print("ok")
