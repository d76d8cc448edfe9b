"""world_size-2 gloo tests (CPU) of the data-parallel host logic: bucketed all-reduce of the
flat gradient buffer, the 1/world scale, and equivalence with the single-replica step.

pmap semantics being restated (src/pretraining.py:125-150): every replica computes grads on its
shard, grads are pmean-ed, every replica applies the same AdamW update.  With the SAME mask on
every rank the mean of per-rank gradients equals the gradient on the concatenated batch.
"""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import jumbo_mae_oracle as O
from jumbo_mae_tpu_b200.config import MAEConfig
from jumbo_mae_tpu_b200.parallel import GradReducer, bucket_ranges
from jumbo_mae_tpu_b200.params import ParamStore, tree_leaves


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _pc(ocfg):
    return MAEConfig(**{f: getattr(ocfg, f) for f in MAEConfig.__dataclass_fields__ if hasattr(ocfg, f)})


def test_bucket_ranges_cover_buffer_back_to_front():
    r = bucket_ranges(1000, 300)
    assert r == [(700, 1000), (400, 700), (100, 400), (0, 100)]
    assert bucket_ranges(10, 100) == [(0, 10)]


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(1)
    ocfg = O.golden_tiny()
    state = O.create_state(ocfg, seed=0, dtype=torch.float64, randomize_all=True)
    images, noise = O.synthetic_batch(ocfg, 4, seed=3, noise_seed=7)
    shard = images[rank * 2:(rank + 1) * 2]
    _, grads = O.loss_and_grads(state.params, shard, noise, ocfg)
    # product-side flat store on CPU (f32 buffers): write local grads through the Flax-shaped views
    store = ParamStore(_pc(ocfg), device="cpu")
    views = dict(tree_leaves(store.tree(store.g)))
    for path, g in O.tree_leaves(grads):
        views[path].copy_(g.float())
    red = GradReducer(store, bucket_mb=1)
    red.ranges = bucket_ranges(store.total, 4096)          # force many buckets
    # reduce in two pieces through the range API (what the overlapped backward uses), then nothing is left
    half = store.total // 2 // 64 * 64
    red.bucket_elems = 4096
    red.all_reduce_ranges([(half, store.total)])
    red.all_reduce_ranges([(0, half)])
    red.join()
    loss_t = torch.tensor([float(rank + 1)])
    red.mean_scalar(loss_t)
    torch.save({"g": store.g.clone(), "loss": loss_t}, os.path.join(out_dir, f"rank{rank}.pt"))
    dist.destroy_process_group()


def test_two_rank_allreduce_matches_single_replica(tmp_path):
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    r0 = torch.load(tmp_path / "rank0.pt")
    r1 = torch.load(tmp_path / "rank1.pt")
    assert torch.equal(r0["g"], r1["g"])                    # every replica holds the same reduced buffer
    assert r0["loss"].item() == 1.5
    # single replica on the concatenated batch; the reducer sums, AdamW applies 1/world
    ocfg = O.golden_tiny()
    state = O.create_state(ocfg, seed=0, dtype=torch.float64, randomize_all=True)
    images, noise = O.synthetic_batch(ocfg, 4, seed=3, noise_seed=7)
    _, want = O.loss_and_grads(state.params, images, noise, ocfg)
    store = ParamStore(_pc(ocfg), device="cpu")
    got = dict(tree_leaves(store.tree(r0["g"])))
    for path, w in O.tree_leaves(want):
        g = got[path].double() / 2.0
        assert (g - w).norm() <= 1e-5 * w.norm() + 1e-9, path


def test_per_rank_noise_streams_differ(monkeypatch):
    """Each device draws its own mask (shard_prng_key, src/pretraining.py:68-73): TrainState.replicate() derives
    the rank's noise stream from (noise_rng, process_index); split_rngs() then advances it every step."""
    from jumbo_mae_tpu_b200 import pretraining as P
    from jumbo_mae_tpu_b200.config import MAEConfig
    cfg = MAEConfig(layers=1, dim=64, heads=2, patch_size=8, image_size=32, dec_layers=1, dec_dim=64, dec_heads=2)

    def draws(rank, seed=0, n=3):
        monkeypatch.setattr(P, "process_index", lambda: rank)
        st = P.TrainState(step=0, store=None, cfg=cfg, module=None, mixup_rng=seed, dropout_rng=seed,
                          noise_rng=seed).replicate()
        out = [st.split_rngs()[0]["noise"] for _ in range(n)]
        assert all(o.shape == (cfg.num_patches,) and o.dtype == np.float32 for o in out)
        return out

    r0, r1, r0_again = draws(0), draws(1), draws(0)
    assert all(np.array_equal(a, b) for a, b in zip(r0, r0_again))          # reproducible per (seed, rank)
    assert not any(np.array_equal(a, b) for a, b in zip(r0, r1))            # ranks draw different masks
    assert not np.array_equal(r0[0], r0[1])                                 # the stream advances per step
    assert not np.array_equal(draws(0, seed=1)[0], r0[0])                   # and depends on --noise-seed
