"""Timeline of the tcgen05 attention kernels (library built with JM_NVCC_DEFS=-DJM_ATTN_TRACE): cycles between the
hand-over points of CTA 0's first items / tiles.  usage: python tools/attn_trace.py fwd|bwd [B T H hd]"""
import ctypes, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from jumbo_mae_tpu_b200 import _lib, ops
what = sys.argv[1]
B, T, H, hd = [int(x) for x in sys.argv[2:6]] if len(sys.argv) >= 6 else (512, 199, 16, 32)
d = H * hd
qkv = (torch.randn(B * T, 3 * d, device="cuda") * 0.5).to(torch.bfloat16)
out = torch.empty(B * T, d, dtype=torch.bfloat16, device="cuda")
dout = torch.randn(B * T, d, device="cuda").to(torch.bfloat16)
dqkv = torch.empty_like(qkv)
lse = torch.empty(B, H, T, device="cuda")
ops.attention_fwd(qkv, B, T, H, hd, out, lse)
for _ in range(2):
    if what == "fwd":
        ops.attention_fwd(qkv, B, T, H, hd, out, lse)
    else:
        ops.attention_bwd(qkv, out, dout, lse, B, T, H, hd, 1.0, dqkv, None)
torch.cuda.synchronize()
buf = (ctypes.c_longlong * (64 * 16))()
lib = _lib.load()
assert lib.jm_debug_attn_trace(buf) == 0
t = [[buf[i * 16 + k] for k in range(16)] for i in range(64)]
if what == "fwd":
    names = ["sm:enter", "sm:S ready", "sm:pass1 done", "sm:xchg done", "sm:P done", "sm:O ready", "sm:drained", "",
             "mma:P seen", "mma:PV issued", "mma:item top", "mma:slot free", "mma:S issued"]
    t0 = t[0][10]
else:
    names = ["cp:top", "cp:S ready", "", "", "", "cp:P written", "cp:grads done", "", "",
             "mma:S issued", "", "", "mma:P seen", "", "mma:grads issd"]
    t0 = t[0][9]
print("item " + " ".join(f"{n:>13s}" for n in names if n))
for i in range(4, 20):
    print(f"{i:4d} " + " ".join(f"{(t[i][k] - t0) if t[i][k] else 0:13d}" for k, n in enumerate(names) if n))
