"""2-rank check (torchrun): training_step with the overlapped all-reduce + per-segment AdamW gives the same
parameters on both ranks and the same result as reduce-then-update."""
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from jumbo_mae_tpu_b200 import parallel
from jumbo_mae_tpu_b200.pretraining import create_train_state, training_step
parallel.init_from_env()
rank = dist.get_rank()
args = bench.preset_args("vit_b16", 16)
args.layers, args.dim, args.heads, args.dec_layers, args.dec_dim, args.dec_heads = 4, 128, 2, 2, 64, 2
g = torch.Generator().manual_seed(100 + rank)
batches = [torch.randint(0, 256, (8, 3, 224, 224), generator=g, dtype=torch.uint8) for _ in range(3)]
outs = []
for mode in ("overlap", "plain"):
    state = create_train_state(args).replicate()
    if mode == "plain":
        state.clip_grad = 1e30          # forces reduce -> (inactive) clip -> one AdamW over the whole buffer
    for b in batches:
        state, m = training_step(state, (b,))
    outs.append(state.store.p.clone())
    other = state.store.p.clone()
    dist.broadcast(other, src=0)
    assert torch.equal(other, state.store.p), "ranks diverged"
rel = ((outs[0] - outs[1]).norm() / outs[1].norm()).item()
if rank == 0:
    print("dp_check ok: overlap vs plain rel diff", rel)
assert rel < 1e-5
