"""Runs attention fwd+bwd once at decoder shape (for ncu captures)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from jumbo_mae_tpu_b200 import ops
B, T, H, hd = int(os.environ.get("B", "512")), int(os.environ.get("T", "199")), int(os.environ.get("H", "16")), int(os.environ.get("HD", "32"))
d = H * hd
qkv = (torch.randn(B * T, 3 * d, device="cuda") * 0.3).to(torch.bfloat16)
out = torch.empty(B * T, d, dtype=torch.bfloat16, device="cuda"); dout = torch.randn(B * T, d, device="cuda").to(torch.bfloat16)
lse = torch.empty(B, H, T, device="cuda"); dqkv = torch.empty_like(qkv); dl = torch.empty(B * H * T, device="cuda")
for _ in range(3):
    ops.attention_fwd(qkv, B, T, H, hd, out, lse)
    ops.attention_bwd(qkv, out, dout, lse, B, T, H, hd, hd ** -0.5, dqkv, dl)
torch.cuda.synchronize(); print("ok")
