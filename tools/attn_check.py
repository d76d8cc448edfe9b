"""Attention forward / backward against an fp32 torch reference of the same bf16 operands, plus CUDA-event timings.
JM_ATTN_TC=0 selects the mma.sync kernels, 1 (default) the tcgen05 kernels where they apply.
usage: python tools/attn_check.py [fwd|bwd|both] [B]"""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from jumbo_mae_tpu_b200 import ops

what = sys.argv[1] if len(sys.argv) > 1 else "both"
Bbig = int(sys.argv[2]) if len(sys.argv) > 2 else 512


def ref(qkv, dout, B, T, H, hd):
    d = H * hd
    x = qkv.float().view(B, T, 3, H, hd).requires_grad_(True)
    q, k, v = x[:, :, 0].transpose(1, 2), x[:, :, 1].transpose(1, 2), x[:, :, 2].transpose(1, 2)
    s = q @ k.transpose(-1, -2)
    lse = torch.logsumexp(s, dim=-1)
    o = (torch.softmax(s, dim=-1) @ v).transpose(1, 2).reshape(B * T, d)
    g = None
    if dout is not None:
        o.backward(dout.float())
        g = x.grad.reshape(B * T, 3 * d)
    return o.detach(), lse.detach(), g


def run(B, T, H, hd, check=True, reps=0):
    d = H * hd
    torch.manual_seed(T * 131 + hd)
    qkv = (torch.randn(B * T, 3 * d, device="cuda") * 0.5).to(torch.bfloat16)
    dout = torch.randn(B * T, d, device="cuda").to(torch.bfloat16)
    out = torch.zeros(B * T, d, dtype=torch.bfloat16, device="cuda")
    lse = torch.zeros(B, H, T, device="cuda")
    dqkv = torch.zeros_like(qkv)
    dl = torch.empty(B * H * T, device="cuda")
    msg = f"B={B} T={T} H={H} hd={hd}:"
    ops.attention_fwd(qkv, B, T, H, hd, out, lse)
    if what != "fwd":
        ops.attention_bwd(qkv, out, dout, lse, B, T, H, hd, 1.0, dqkv, dl)
    torch.cuda.synchronize()
    if check:
        o, l, g = ref(qkv, dout if what != "fwd" else None, B, T, H, hd)
        eo = (out.float() - o).abs().max().item() / o.abs().max().item()
        el = (lse - l).abs().max().item()
        msg += f" out err {eo:.2e} lse err {el:.2e}"
        if g is not None:
            eg = ((dqkv.float() - g).norm() / g.norm()).item()
            msg += f" dqkv rel {eg:.2e}"
        bad = eo > 2e-2 or el > 1e-3 or (g is not None and eg > 2e-2)
        msg += "  BAD" if bad else "  ok"
    if reps:
        for name, fn in (("fwd", lambda: ops.attention_fwd(qkv, B, T, H, hd, out, lse)),
                         ("bwd", lambda: ops.attention_bwd(qkv, out, dout, lse, B, T, H, hd, 1.0, dqkv, dl))):
            if name == "bwd" and what == "fwd":
                continue
            for _ in range(3):
                fn()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                fn()
            e1.record()
            torch.cuda.synchronize()
            msg += f" {name} {e0.elapsed_time(e1) / reps * 1e3:.1f} us"
    print(msg, flush=True)


print("JM_ATTN_TC =", os.environ.get("JM_ATTN_TC", "1"))
for (B, T, H, hd) in [(2, 16, 2, 32), (2, 52, 2, 64), (3, 199, 4, 32), (2, 128, 2, 64), (2, 129, 2, 32), (2, 199, 3, 64),
                      (3, 256, 2, 32), (2, 7, 1, 64), (40, 199, 16, 32), (40, 52, 12, 64)]:
    run(B, T, H, hd)
run(Bbig, 199, 16, 32, check=False, reps=20)
run(Bbig, 52, 12, 64, check=False, reps=20)
run(Bbig // 2, 199, 12, 64, check=False, reps=20)
