"""Generates tests/golden/tiny_step.npz from the CPU oracle (f64 run, stored as f32/f64).

The reference has no golden vectors and cannot be run here (no JAX), so these
fixtures pin the *oracle* against silent drift, and give the GPU tests inputs and
expected outputs that do not depend on torch's RNG stream.  Run from the repo root:
    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from oracle import jumbo_mae_oracle as O  # noqa: E402


def flat(tree):
    return {"/".join(p): t.detach().numpy() for p, t in O.tree_leaves(tree)}


def main():
    out = {}
    for name, norm_pix in (("tiny", False), ("tiny_normpix", True)):
        cfg = O.golden_tiny(norm_pix_loss=norm_pix)
        state = O.create_state(cfg, seed=0, dtype=torch.float64, randomize_all=True)
        images, noise = O.synthetic_batch(cfg, 4, seed=0, noise_seed=1000)
        collect = {}
        loss, grads = O.loss_and_grads(state.params, images, noise, cfg, collect)
        new_state, metrics, _ = O.training_step(state, images, noise)
        out[f"{name}/images"] = images.numpy()
        out[f"{name}/noise"] = noise
        out[f"{name}/loss"] = np.float64(loss)
        out[f"{name}/lr"] = np.float64(metrics["learning_rate"])
        if not norm_pix:                       # same seed -> same params for both variants
            for k, v in flat(state.params).items():
                out[f"{name}/params/{k}"] = v
        if not norm_pix:
            for k, v in flat(grads).items():
                out[f"{name}/grads/{k}"] = v.astype(np.float32)
        old = flat(state.params)
        for k, v in flat(grads).items():
            out[f"{name}/grad_stats/{k}"] = np.array([v.sum(), np.sqrt((v * v).sum())])
        for k, v in flat(new_state.params).items():
            d = v - old[k]
            out[f"{name}/delta_stats/{k}"] = np.array([d.sum(), np.sqrt((d * d).sum())])
        for k in ("enc_in", "enc_out", "dec_in", "dec_out", "image_output"):
            out[f"{name}/act/{k}"] = collect[k].detach().numpy().astype(np.float32)
        out[f"{name}/mask"] = collect["mask"].numpy().astype(np.float32)
        out[f"{name}/ids_restore"] = collect["ids_restore"].numpy().astype(np.int32)
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "tiny_step.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
