"""Generates tests/golden/tiny_classify.npz from the CPU oracle's classification step (f64 run).

Same purpose as make_golden.py: the reference has no golden vectors and cannot be run here, so the fixture pins the
oracle's fine-tune restatement (vit_classify, smooth_labels, apply_mixup, softmax cross-entropy, autograd) against
silent drift and gives the GPU test inputs that do not depend on torch's RNG stream.  Run from the repo root:
    python tests/golden/make_golden_finetune.py
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from oracle import jumbo_mae_oracle as O  # noqa: E402

LABELS, BATCH, SMOOTHING, RATIO = 16, 4, 0.1, 0.35


def inputs(cfg):
    images, _ = O.synthetic_batch(cfg, BATCH, seed=0)
    idx = np.array([3, 0, 15, 8], dtype=np.int32)
    perm = np.array([1, 3, 0, 2], dtype=np.int32)
    return images, idx, perm


def step(params, images, idx, perm, cfg, dtype):
    norm = O.preprocess(images, dtype=dtype).numpy()
    lab = O.smooth_labels(np.eye(LABELS, dtype=np.float32)[idx], SMOOTHING)
    x, lab = O.apply_mixup(norm.astype(np.float32), lab, RATIO, perm) if dtype == torch.float32 else \
        (np.float64(np.float32(RATIO)) * norm + (1.0 - np.float64(np.float32(RATIO))) * norm[perm], O.mix_labels(lab, perm, RATIO))
    return O.finetune_loss_and_grads(params, torch.as_tensor(x, dtype=dtype), lab, cfg) + (lab,)


def main():
    cfg = O.golden_tiny(labels=LABELS)
    params = O.init_params(cfg, seed=0, dtype=torch.float64, randomize_all=True)
    images, idx, perm = inputs(cfg)
    loss, logits, grads, lab = step(params, images, idx, perm, cfg, torch.float64)
    out = {"images": images.numpy(), "idx": idx, "perm": perm, "ratio": np.float32(RATIO), "smoothing": np.float64(SMOOTHING),
           "labels": lab, "loss": np.float64(loss), "logits": logits.numpy()}
    for p, t in O.tree_leaves(params):
        out["params/" + "/".join(p)] = t.numpy()
    for p, g in O.tree_leaves(grads):
        v = g.numpy()
        if p[1] in ("head", "embed", "cls_tokens"):            # full arrays for the two ends, statistics for the rest
            out["grads/" + "/".join(p)] = v.astype(np.float32)
        out["grad_stats/" + "/".join(p)] = np.array([v.sum(), np.sqrt((v * v).sum())])
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "tiny_classify.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
