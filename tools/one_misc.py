"""Runs the loss kernel a few times (for ncu captures)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from jumbo_mae_tpu_b200 import ops
B, N, n, J, P = 512, 196, 49, 3, 768
images = torch.randint(0, 256, (B, 3, 224, 224), dtype=torch.uint8, device="cuda")
noise = torch.rand(N, device="cuda")
ids_shuffle = torch.argsort(noise).to(torch.int32)
ids_restore = torch.argsort(ids_shuffle.long()).to(torch.int32)
mask_row = (ids_restore >= n).float()
pred = torch.randn(B * (J + N), P, device="cuda")
dpred = torch.empty(B * (J + N), P, dtype=torch.bfloat16, device="cuda")
loss = torch.zeros(1, device="cuda")
ws = torch.empty(ops.patch_loss_workspace(B, N), dtype=torch.uint8, device="cuda")
for _ in range(3):
    ops.patch_loss(pred, images, mask_row, 16, J, False, 1.0, loss, dpred, ws)
torch.cuda.synchronize(); print("ok", loss.item())
