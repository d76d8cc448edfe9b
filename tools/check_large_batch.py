"""One-off consistency check of the micro-batch 2048 step (index arithmetic at 407 552 decoder rows): the loss and
the gradient of one batch of 2048 images equal the mean of the two halves run at micro-batch 1024 (same parameters,
same mask noise)."""
import sys
import numpy as np
import torch
sys.path.insert(0, ".")
from bench import preset_args
from jumbo_mae_tpu_b200.pretraining import create_train_state

ns = preset_args("vit_b16", 4096)
state = create_train_state(ns).replicate()
cfg, store = state.cfg, state.store
gen = torch.Generator().manual_seed(0)
images = torch.randint(0, 256, (2048, 3, 224, 224), generator=gen, dtype=torch.uint8).cuda()
noise = torch.rand(cfg.num_patches, generator=gen).cuda()
big = state.engine(2048)
big.forward(images, noise)
big.backward(zero_grads=True)
torch.cuda.synchronize()
loss_big, g_big = big.loss.item(), store.g.clone()
del big
state.engines.clear()
torch.cuda.empty_cache()
half = state.engine(1024)
losses = []
for i in range(2):
    half.forward(images[i * 1024:(i + 1) * 1024], noise)
    half.backward(zero_grads=(i == 0))
    torch.cuda.synchronize()
    losses.append(half.loss.item())
g_half = store.g * 0.5
rel = ((g_big - g_half).norm() / g_half.norm()).item()
print(f"loss 2048: {loss_big:.7f}  mean of halves: {np.mean(losses):.7f}  gradient rel-L2 difference: {rel:.3e}")
assert abs(loss_big - np.mean(losses)) < 1e-5 * abs(loss_big) and rel < 2e-3
print("ok")
