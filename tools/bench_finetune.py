"""Throughput of the classification (fine-tune) step on one B200: ViT-B/16, 224 px, 196 + 3 tokens, 1000 classes
(BASELINE.json configs[4], SURVEY.md 8(d): 108.306 GFLOP per image per step).  forward + backward through the CUDA
graph with label smoothing 0.1 and Mixup, AdamW with the layer-wise lr table (lr_decay 0.75) every step.

  python tools/bench_finetune.py [--batch 128] [--steps 10] [--warmup 3]
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from jumbo_mae_tpu_b200 import _lib  # noqa: E402
from jumbo_mae_tpu_b200.config import MAEConfig  # noqa: E402
from jumbo_mae_tpu_b200.finetune_engine import FinetuneEngine  # noqa: E402
from jumbo_mae_tpu_b200.params import ParamStore  # noqa: E402


def flops_per_image(cfg):
    N, J, d, L, P = cfg.num_patches, cfg.num_cls_tokens, cfg.dim, cfg.layers, cfg.patch_dim
    T = N + J
    fwd = 2 * (N * P * d + L * (4 * T * d * d + 2 * T * T * d + 8 * N * d * d + 8 * J * J * d * d) + J * d * cfg.labels)
    return fwd, 3 * fwd - 2 * N * P * d


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=128)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--eager", action="store_true", help="no CUDA graph: every kernel is its own launch (for ncu)")
    args = ap.parse_args()
    _lib.require_device()
    cfg = MAEConfig(labels=1000)
    store = ParamStore(cfg)
    g = torch.Generator(device="cuda").manual_seed(0)
    store.p.normal_(0, 0.02, generator=g)
    store.refresh_bf16()
    eng = FinetuneEngine(cfg, store, args.batch)
    if not args.eager:
        eng.capture_graphs(label_smoothing=0.1, mixup=True)
    perm = torch.arange(args.batch, dtype=torch.int32, device="cuda").flip(0).contiguous()
    images = torch.randint(0, 256, (args.batch, 3, 224, 224), dtype=torch.uint8, device="cuda", generator=g)
    labels = torch.randint(0, 1000, (args.batch,), dtype=torch.int32, device="cuda", generator=g)
    lib = _lib.load()

    def step(i):
        if args.eager:
            eng.forward(images, labels, label_smoothing=0.1, perm=perm, ratio=0.5)
            eng.backward()
        else:
            eng.run_graph(images, labels, True)
        eng.adamw_step(i, 1e-4, grad_scale=1.0, lr_decay=0.75)

    for i in range(args.warmup):
        step(i)
    torch.cuda.synchronize()
    l0 = lib.jm_launch_count()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for i in range(args.steps):
        step(args.warmup + i)
    t1.record()
    torch.cuda.synchronize()
    ms = t0.elapsed_time(t1) / args.steps
    _, fstep = flops_per_image(cfg)
    peaks = {}
    pk = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    if os.path.exists(pk):
        peaks = json.load(open(pk))
    ips = args.batch / ms * 1e3
    tf = ips * fstep / 1e12
    peak = peaks.get("bf16_tflops_sustained") or peaks.get("bf16_dense_tflops_sustained")
    print(json.dumps({
        "metric": "finetune_images_per_sec", "value": ips, "unit": "images/s", "n_gpus": 1, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": "ViT-B/16 Jumbo CLS classification step, 224 px, 199 tokens, 1000 classes, label "
                               "smoothing 0.1, Mixup, AdamW with lr_decay 0.75 (BASELINE.json configs[4]); DropPath off",
                   "batch": args.batch},
        "flops_per_image": fstep, "achieved_tflops": tf, "peak_tflops_sustained": peak,
        "frac": (tf / peak) if peak else None,
        "gpu_launches": (lib.jm_launch_count() - l0) + getattr(eng, "graph_launches", 0) * args.steps,
        "loss": eng.loss.item(), "acc1": eng.metrics[1].item()}))


if __name__ == "__main__":
    main()
