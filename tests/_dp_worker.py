"""Worker of tests/test_dp_gpu.py, launched once per rank by torch.distributed.run (NCCL, one GPU per rank).

Checks of the data-parallel step (src/pretraining.py:125-159 under jax.pmap):
  1. after 3 optimizer steps the parameters are bit-identical on every rank;
  2. the overlapped schedule (per-segment all-reduce + per-segment AdamW while backward continues) gives the
     same parameters as reduce-then-update, rel-L2 <= 1e-5;
  3. with per-rank images and per-rank mask noise, the all-reduced gradient / world equals the mean over ranks of
     the oracle's gradients of each rank's own (images, noise), leaf by leaf (rel-L2 <= 3e-2, cosine >= 0.999).
"""
import argparse
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from jumbo_mae_tpu_b200 import parallel  # noqa: E402
from jumbo_mae_tpu_b200.pretraining import create_train_state, training_step  # noqa: E402
from oracle import jumbo_mae_oracle as O  # noqa: E402


def small_args(**kw):
    base = dict(layers=4, dim=128, heads=2, labels=0, patch_size=16, image_size=64, image_mask_ratio=0.75,
                dec_layers=2, dec_dim=64, dec_heads=2, norm_pix_loss=False, init_seed=0, mixup_seed=0,
                dropout_seed=0, noise_seed=0, optimizer="adamw", learning_rate=1.5e-4, weight_decay=0.05,
                adam_b1=0.9, adam_b2=0.95, adam_eps=1e-8, lr_decay=1.0, clip_grad=0.0, grad_accum=1,
                train_batch_size=256, warmup_steps=10, training_steps=100, posemb="sincos2d", dropout=0.0,
                droppath=0.0, dec_dropout=0.0, dec_droppath=0.0, layerscale=False, dec_layerscale=False)
    base.update(kw)
    return argparse.Namespace(**base)


def main():
    parallel.init_from_env()
    rank, world = dist.get_rank(), dist.get_world_size()
    args = small_args()
    g = torch.Generator().manual_seed(100 + rank)
    batches = [torch.randint(0, 256, (8, 3, 64, 64), generator=g, dtype=torch.uint8) for _ in range(3)]

    # ---- 1 + 2: ranks stay identical; overlapped == reduce-then-update
    outs = []
    for mode in ("overlap", "plain"):
        state = create_train_state(args).replicate()
        if mode == "plain":
            state.clip_grad = 1e30      # forces reduce -> (inactive) clip -> one AdamW over the whole buffer
        for b in batches:
            state, _ = training_step(state, (b,))
        torch.cuda.synchronize()
        outs.append(state.store.p.clone())
        other = state.store.p.clone()
        dist.broadcast(other, src=0)
        assert torch.equal(other, state.store.p), f"rank {rank}: parameters diverged from rank 0 ({mode})"
    rel = ((outs[0] - outs[1]).norm() / outs[1].norm()).item()
    assert rel <= 1e-5, f"overlapped vs reduce-then-update: rel-L2 {rel}"

    # ---- 3: rank-by-rank gradients against the oracle, per-rank noise
    from jumbo_mae_tpu_b200.config import MAEConfig
    from jumbo_mae_tpu_b200.engine import MAEEngine
    from jumbo_mae_tpu_b200.params import ParamStore, tree_leaves
    ocfg = O.MAEConfig(layers=4, dim=128, heads=2, patch_size=16, image_size=64, dec_layers=2, dec_dim=64,
                       dec_heads=2, warmup_steps=10, training_steps=100, train_batch_size=256)
    cfg = MAEConfig(**{f: getattr(ocfg, f) for f in MAEConfig.__dataclass_fields__ if hasattr(ocfg, f)})
    ostate = O.create_state(ocfg, seed=0, randomize_all=True)
    store = ParamStore(cfg)
    store.load_tree(ostate.params)
    eng = MAEEngine(cfg, store, 4)
    want = None
    for r in range(world):                                    # every rank restates every rank's oracle gradient
        images, noise = O.synthetic_batch(ocfg, 4, seed=50 + r, noise_seed=2000 + r)
        _, gr = O.loss_and_grads(ostate.params, images, noise, ocfg)
        want = gr if want is None else O.tree_map(lambda a, b: a + b, want, gr)
        if r == rank:
            mine = (images, noise)
    want = O.tree_map(lambda a: a / world, want)
    eng.forward(mine[0].cuda(), torch.from_numpy(mine[1]).cuda())
    eng.backward()
    red = parallel.GradReducer(store)
    red.all_reduce_grads()
    torch.cuda.synchronize()
    got = dict(tree_leaves(store.tree(store.g)))
    ref = dict(O.tree_leaves(want))
    bad = []
    for path, w in ref.items():
        if path[-2:] == ("wk", "bias"):
            continue                                          # zero true gradient (see tests/test_step_gpu.py)
        gg = got[path].detach().cpu().reshape(w.shape) / world
        r_ = ((gg - w).norm() / (w.norm() + 1e-12)).item()
        cos = (gg.flatten() @ w.flatten() / (gg.norm() * w.norm() + 1e-20)).item()
        if not (r_ <= 3e-2 and cos >= 0.999):
            bad.append(("/".join(path), r_, cos))
    assert not bad, bad
    # ---- 4: classification engine, overlapped per-segment all-reduce == one all-reduce after the backward pass
    from jumbo_mae_tpu_b200.finetune_engine import FinetuneEngine
    ccfg = MAEConfig(layers=4, dim=128, heads=2, patch_size=16, image_size=64, labels=16)
    cstate = O.init_params(O.MAEConfig(layers=4, dim=128, heads=2, patch_size=16, image_size=64, labels=16), seed=0,
                           randomize_all=True)
    cstore = ParamStore(ccfg)
    cstore.load_tree(cstate)
    ceng = FinetuneEngine(ccfg, cstore, 8)
    ceng.capture_graphs(label_smoothing=0.1, mixup=False)
    ceng.capture_segment_graphs()
    cred = parallel.GradReducer(cstore)
    labels = torch.randint(0, 16, (8,), generator=g).to(torch.int32).cuda()
    ceng.run_graph(batches[0].cuda(), labels, True)
    cred.all_reduce_grads()
    torch.cuda.synchronize()
    g_plain = cstore.g.clone()
    ceng.run_graph_overlapped(batches[0].cuda(), labels, True, cred)
    torch.cuda.synchronize()
    crel = ((cstore.g - g_plain).norm() / g_plain.norm()).item()
    assert crel <= 2e-4, f"classification engine, overlapped vs plain all-reduce: rel-L2 {crel}"   # (two passes: f32 atomics order)
    dist.barrier()
    if rank == 0:
        print(f"dp_worker ok: world {world}, overlap vs plain rel {rel:.2e}, classification {crel:.2e}")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
