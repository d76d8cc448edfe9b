"""Times LayerNorm backward at the decoder shape: f32 residual-gradient stream against the bf16-only one."""
import sys, torch
sys.path.insert(0, ".")
from jumbo_mae_tpu_b200 import ops

rows, D = 512 * 199, 512
dev = "cuda"
BF16, F32 = torch.bfloat16, torch.float32
mk = lambda dt: torch.randn(rows, D, device=dev).to(dt)
# several buffer sets, larger than L2 together, rotated
sets = []
for _ in range(4):
    sets.append(dict(dy=mk(BF16), x=mk(F32), res=mk(F32), res16=mk(BF16), dx=mk(F32), dxb=mk(BF16)))
mean = torch.zeros(rows, device=dev); rstd = torch.ones(rows, device=dev); gamma = torch.ones(D, device=dev)
dg, db, dc = (torch.zeros(D, device=dev) for _ in range(3))
ws = torch.empty(ops.layernorm_bwd_workspace(rows, D), dtype=torch.uint8, device=dev)

def run(kind, s):
    if kind == "f32":
        ops.layernorm_bwd(s["dy"], s["x"], mean, rstd, gamma, s["res"], s["dx"], s["dxb"], dg, db, ws, dcol=dc)
    elif kind == "lowp":
        ops.layernorm_bwd_lowp(s["dy"], s["x"], mean, rstd, gamma, s["res16"], None, s["dxb"], dg, db, ws, dcol=dc)
    else:
        ops.layernorm_bwd_lowp(s["dy"], s["x"], mean, rstd, gamma, s["res16"], s["dx"], s["dxb"], dg, db, ws, dcol=dc)

for kind, bytes_per in (("f32", 16), ("lowp", 10), ("lowp+f32out", 14)):
    for i in range(4):
        run(kind, sets[i])
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    n = 40
    for i in range(n):
        run(kind, sets[i % 4])
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / n * 1e3
    print(f"{kind:12s} {us:7.1f} us  {rows * D * bytes_per / us / 1e3:7.1f} GB/s")
