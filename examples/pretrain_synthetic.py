"""End-to-end use of the drop-in on one GPU (or under torchrun, one process per GPU): the training loop of the
reference's main_pretrain.py (src/main_pretrain.py:47-94) with synthetic uint8 images --

    create_train_state(args).replicate() -> training_step(state, batch) with the next batch prefetched from pinned
    host memory -> evaluate(state, loader, validation_step) -> save_state / load_state (resume) -> save_params
    (the reference's Flax-msgpack parameter file).

    python examples/pretrain_synthetic.py                       # tiny model, a few steps
    python examples/pretrain_synthetic.py --preset vit_b16 --batch 256 --steps 10
"""
import argparse
import os
import sys
import tempfile

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from jumbo_mae_tpu_b200 import parallel  # noqa: E402
from jumbo_mae_tpu_b200.checkpoint import load_state, save_params, save_state  # noqa: E402
from jumbo_mae_tpu_b200.evaluate import evaluate  # noqa: E402
from jumbo_mae_tpu_b200.pretraining import create_train_state, training_step, validation_step  # noqa: E402


def namespace(preset: str, batch: int, accum: int) -> argparse.Namespace:
    shapes = {"tiny": dict(layers=2, dim=64, heads=2, patch_size=8, image_size=32, dec_layers=2, dec_dim=64, dec_heads=2),
              "vit_b16": dict(layers=12, dim=768, heads=12, patch_size=16, image_size=224, dec_layers=8, dec_dim=512,
                              dec_heads=16)}[preset]
    return argparse.Namespace(image_mask_ratio=0.75, labels=0, posemb="sincos2d", norm_pix_loss=False, init_seed=0,
                              mixup_seed=0, dropout_seed=0, noise_seed=0, optimizer="adamw", learning_rate=1.5e-4,
                              weight_decay=0.05, adam_b1=0.9, adam_b2=0.95, adam_eps=1e-8, lr_decay=1.0, clip_grad=0.0,
                              grad_accum=accum, train_batch_size=batch * accum * parallel.process_count(),
                              warmup_steps=4, training_steps=100, **shapes)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--preset", default="tiny", choices=["tiny", "vit_b16"])
    ap.add_argument("--batch", type=int, default=8, help="per-GPU micro-batch")
    ap.add_argument("--accum", type=int, default=2)
    ap.add_argument("--steps", type=int, default=4, help="optimizer steps")
    args = ap.parse_args()
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    parallel.init_from_env("nccl")
    ns = namespace(args.preset, args.batch, args.accum)
    state = create_train_state(ns).replicate()
    size = ns.image_size
    gen = torch.Generator().manual_seed(parallel.process_index())
    pool = [torch.randint(0, 256, (args.batch, 3, size, size), generator=gen, dtype=torch.uint8).pin_memory()
            for _ in range(4)]

    nxt = state.prefetch((pool[0],))
    for i in range(args.steps * args.accum):
        cur, nxt = nxt, None
        state, metrics = training_step(state, cur)                  # launches the micro-step
        nxt = state.prefetch((pool[(i + 1) % len(pool)],))          # the next batch's copy overlaps it
        if (i + 1) % args.accum == 0 and parallel.process_index() == 0:
            print(f"step {state.step}: loss {metrics['loss'].item():.4f}  lr {metrics['learning_rate']:.3e}")

    # (as in the reference, val/loss = sum of the per-device batch-mean losses / number of samples, src/main_pretrain.py:37-45)
    val = evaluate(state, [(p,) for p in pool[:2]], validation_step)
    if parallel.process_index() == 0:
        print("validation:", {k: round(v, 4) for k, v in val.items()})
        with tempfile.TemporaryDirectory() as tmp:
            save_state(state, os.path.join(tmp, "resume.msgpack"))          # parameters + Adam moments + counters + RNG
            save_params(state, os.path.join(tmp, "params.msgpack"))         # the reference's parameter file
            again = create_train_state(ns).replicate()
            load_state(again, os.path.join(tmp, "resume.msgpack"))
            assert again.step == state.step and torch.equal(again.store.p, state.store.p)
            print(f"resumed at step {again.step}; files: {sorted(os.listdir(tmp))}")


if __name__ == "__main__":
    main()
