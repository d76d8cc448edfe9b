#!/usr/bin/env python
"""Headline benchmark: MAE ViT-B/16 (Jumbo CLS) pre-training step throughput on B200.

    python bench.py --gpus 1 --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus 1 --steps K ...  # the reference algorithm on host cores
    torchrun --nproc-per-node N bench.py --gpus N ...        # one rank per GPU, NCCL

Workload (BASELINE.json configs[1]): ViT-B/16, 224 px, mask 0.75, decoder 8 x 512 (16 heads),
Jumbo CLS J = 3, bf16 compute, GLOBAL batch 4096 synthetic uint8 images per optimizer step,
run as micro-batches of min(1024, 4096 / N) per GPU with the reference's gradient-accumulation semantics
(4096 / (micro * N) micro-steps per rank, one gradient all-reduce and one AdamW per step).
One "step" = one optimizer step = forward + backward of 4096 images + all-reduce + AdamW.

Prints ONE JSON line (rank 0).  `value` is whole-job images/s with inputs resident in HBM;
`e2e` is the same step through the public API (`training_step`) with pinned-host uint8 images
copied in and the loss read back every micro-step inside the timed region.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def preset_args(name: str, global_batch: int) -> argparse.Namespace:
    """The reference's CLI namespace for its pre-training presets (config/pretrain/*.sh)."""
    shapes = {
        "vit_b16": dict(layers=12, dim=768, heads=12, patch_size=16, training_steps=1281167 * 1600 // 4096),
        "vit_l16": dict(layers=24, dim=1024, heads=16, patch_size=16, training_steps=1281167 * 800 // 4096),
        "vit_h14": dict(layers=32, dim=1280, heads=16, patch_size=14, training_steps=1281167 * 800 // 4096),
    }[name]
    return argparse.Namespace(
        image_mask_ratio=0.75, labels=0, image_size=224, posemb="sincos2d", pooling="cls", dropout=0.0, droppath=0.0,
        layerscale=False, grad_ckpt=False, dec_layers=8, dec_dim=512, dec_heads=16, dec_layerscale=False,
        dec_posemb="sincos2d", dec_dropout=0.0, dec_droppath=0.0, norm_pix_loss=False, init_seed=0, mixup_seed=0,
        dropout_seed=0, noise_seed=0, optimizer="adamw", learning_rate=1.5e-4, weight_decay=0.05, adam_b1=0.9,
        adam_b2=0.95, adam_eps=1e-8, lr_decay=1.0, clip_grad=0.0, grad_accum=1, train_batch_size=global_batch,
        warmup_steps=1281167 * 40 // 4096, **shapes)


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return {"bf16_burst": p["bf16_tflops"], "bf16_sustained": p.get("bf16_tflops_sustained", p["bf16_tflops"]),
                "hbm": p["hbm_gbs"], "source": "measured"}
    # /opt/skills/guides/B200_PROFILING.md fallback
    return {"bf16_burst": 1590.0, "bf16_sustained": 1400.0, "hbm": 6650.0, "source": "fallback"}


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons every 200 ms while a timed region runs."""

    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-i", str(self.index),
                 "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.time(), line.strip()))

    def mark(self):
        """Start of a timed region: summary() uses the samples taken from here on (nvidia-smi needs ~0.3 s to
        deliver its first line, so the sampler itself is started once, before the warm-up steps)."""
        self.t_mark = time.time()

    def stop(self):
        if self.proc is None:
            return
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()

    def summary(self):
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        t_mark = getattr(self, "t_mark", 0.0)
        lines = [l for t, l in self.lines if t >= t_mark]
        if not lines:            # region shorter than the sampling period: the latest sample (taken under warm-up load)
            lines = [l for _, l in self.lines[-1:]]
        for line in lines:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
            except ValueError:
                continue
            for nm, val in zip(names, parts[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                "samples": len(sm)}


# ---------------------------------------------------------------------------------------------
def cpu_oracle_rate(config: str, batch: int, steps: int, warmup: int):
    """Times the CPU restatement of the reference step (oracle/) on the host cores."""
    from oracle import jumbo_mae_oracle as O
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    ocfg = {"vit_b16": O.vit_b16, "vit_l16": O.vit_l16, "vit_h14": O.vit_h14}[config]()
    state = O.create_state(ocfg, seed=0)
    images, noise = O.synthetic_batch(ocfg, batch, seed=0, noise_seed=1000)
    for _ in range(warmup):
        state, _, _ = O.training_step(state, images, noise)
    t0 = time.perf_counter()
    for _ in range(steps):
        state, met, _ = O.training_step(state, images, noise)
    dt = time.perf_counter() - t0
    return batch * steps / dt, dt / steps, cores, met["loss"]


def run_reference(args):
    """--impl reference: the reference's own algorithm on the box's host cores.  The JAX
    reference cannot be installed here (no jax/flax/optax wheels, no network), so this arm
    times the CPU port in oracle/ (kind = "port") on a bounded sample: batch 8 per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    batch = 8                                  # BASELINE.json configs[0] / BASELINE.md section 4: batch 8 on the host cores
    ips, sec, cores, loss = cpu_oracle_rate(args.config, batch, args.steps, args.warmup)
    sample = f"{args.config} MAE step, batch {batch} per step, {args.warmup} warm-up + {args.steps} timed, f32 torch-CPU"
    line = {
        "impl": "reference", "metric": "mae_pretrain_images_per_sec", "value": ips, "unit": "images/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"MAE {args.config} 224px mask 0.75 decoder 8x512 Jumbo CLS J=3, batch {batch} synthetic uint8 "
                               f"per step on the host cores (BASELINE.json configs[0]; a bounded sample of the GPU arm's "
                               f"global batch {args.global_batch})",
                   "global_batch": batch, "micro_batch": batch, "grad_accum": 1, "parallelism": "cpu", "norm_pix_loss": False,
                   "l2": "n/a (host)"},
        "cpu_baseline": {"value": ips, "unit": "images/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": ips, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "loss": loss,
    }
    print(json.dumps(line), flush=True)


def workload_config(args, micro, accum, world):
    return {
        "workload": f"MAE {args.config} 224px mask 0.75 decoder 8x512 Jumbo CLS J=3, global batch {args.global_batch} "
                    f"synthetic uint8 (BASELINE.json configs[1])",
        "global_batch": args.global_batch, "micro_batch": micro, "grad_accum": accum, "parallelism": f"dp{world}",
        "norm_pix_loss": bool(getattr(args, "norm_pix_loss", False)),
        "l2": "inputs larger than L2: every micro-batch streams tens of GB of activations through a 126 MB L2",
    }


# ---------------------------------------------------------------------------------------------
def run_ours(args):
    from jumbo_mae_tpu_b200 import _lib, ops
    from jumbo_mae_tpu_b200 import parallel
    from jumbo_mae_tpu_b200.config import flops_per_image
    from jumbo_mae_tpu_b200.engine import lr_schedule
    from jumbo_mae_tpu_b200.pretraining import create_train_state, training_step

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    parallel.init_from_env("nccl")
    import torch.distributed as dist
    _lib.require_device()
    lib = _lib.load()
    peaks = measured_peaks()

    if args.micro_batch is None:
        args.micro_batch = {"vit_b16": 2048, "vit_l16": 512, "vit_h14": 256}[args.config]
    micro = min(args.micro_batch, args.global_batch // world)
    accum = args.global_batch // (micro * world)
    assert micro * accum * world == args.global_batch, "global batch must divide into micro-batches"
    ns = preset_args(args.config, args.global_batch)
    ns.grad_accum = accum
    ns.norm_pix_loss = bool(args.norm_pix_loss)
    state = create_train_state(ns).replicate()
    cfg = state.cfg
    eng = state.engine(micro)
    reducer = parallel.GradReducer(state.store) if world > 1 else None
    state.reducer = reducer
    dev = torch.device("cuda", local)

    gen = torch.Generator().manual_seed(rank)
    host_images = [torch.randint(0, 256, (micro, 3, cfg.image_size, cfg.image_size), generator=gen,
                                 dtype=torch.uint8).pin_memory() for _ in range(accum)]
    dev_images = [h.to(dev) for h in host_images]
    ngen = np.random.Generator(np.random.Philox(key=1000 + rank))
    total_micro = (args.steps + args.warmup) * accum
    dev_noise = torch.from_numpy(ngen.random((total_micro, cfg.num_patches), dtype=np.float32)).to(dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    counter = {"micro": 0, "step": 0}

    use_graph = os.environ.get("JM_NO_GRAPH", "0") != "1"
    if use_graph:
        eng.capture_graphs()

    if use_graph and reducer is not None:
        eng.capture_segment_graphs()

    def resident_step():
        updated = False
        for i in range(accum):
            if use_graph and reducer is not None and i == accum - 1:
                # as training_step does: all-reduce AND AdamW of each backward segment overlap the next segment
                cnt, lr_now = counter["step"], lr_schedule(counter["step"], cfg)
                hook = ((lambda ranges: eng.adamw_ranges(ranges, cnt, lr_now, grad_scale=1.0 / (accum * world)))
                        if reducer.stream is not None else None)
                eng.run_graph_overlapped(dev_images[i], dev_noise[counter["micro"] % total_micro], i == 0, reducer,
                                         after_reduce=hook)
                updated = hook is not None
            elif use_graph:
                eng.run_graph(dev_images[i], dev_noise[counter["micro"] % total_micro], zero_grads=(i == 0))
            else:
                eng.forward(dev_images[i], dev_noise[counter["micro"] % total_micro])
                eng.backward(zero_grads=(i == 0))
            counter["micro"] += 1
        if reducer is not None and not use_graph:
            reducer.all_reduce_grads()
        if not updated:
            eng.adamw_step(counter["step"], lr_schedule(counter["step"], cfg), grad_scale=1.0 / (accum * world))
        counter["step"] += 1

    def timed(fn, steps):
        barrier()
        sampler.mark()
        l0 = lib.jm_launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        launches = lib.jm_launch_count() - l0
        if use_graph:     # kernels inside a replayed graph are not re-counted by the library: add them
            launches += steps * accum * eng.graph_launches
        ms = e0.elapsed_time(e1)
        barrier()
        t = torch.tensor([ms], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item(), launches, sampler.summary()

    sampler = ClockSampler(local)
    sampler.start()
    for _ in range(args.warmup):
        resident_step()
    ms, launches, clocks = timed(resident_step, args.steps)
    ms_per_step = ms / args.steps
    value = args.global_batch / (ms_per_step / 1e3)
    final_loss = eng.loss.item()

    # ---- e2e through the public API: pinned host images in, loss out, every micro-step
    state.step, state.micro_step = counter["step"], 0
    e2e_losses = []

    pending = {"next": None}

    def e2e_step():
        # the input path of a training loop: as soon as a micro-step has been launched, the NEXT micro-batch (of this
        # optimizer step or of the following one) is copied from pinned host memory on a copy stream
        # (TrainState.prefetch), so the copy overlaps the running step also when there is one micro-step per
        # optimizer step; every copy and every loss read is inside the timed region
        if pending["next"] is None:
            pending["next"] = state.prefetch((host_images[0],))
        for i in range(accum):
            cur = pending["next"]
            _, metrics = training_step(state, cur)
            pending["next"] = state.prefetch((host_images[(i + 1) % accum],))
            e2e_losses.append(metrics["loss"].item())          # device -> host read of the step's result

    e2e_step()                                                  # warm-up of the API path (allocations)
    ms_e2e, _, _ = timed(e2e_step, args.steps)
    e2e_value = args.global_batch / (ms_e2e / args.steps / 1e3)
    h2d = accum * (micro * 3 * cfg.image_size * cfg.image_size + cfg.num_patches * 4)
    d2h = accum * 4

    # ---- roofline of the dominant kernel.  By time share (profiles/r1c_launches_summary.md) that is the tcgen05
    # GEMM, and its largest instantiation is the weight-gradient product dW += X^T dY (15 % of the step); it is
    # timed live at the decoder-FFN shape with CUDA events.  The GELU-forward instantiation (12 %), whose epilogue
    # writes two bf16 tensors and is not hidden behind a K = 512 main loop, is reported next to it.
    kmicro = min(micro, 1024)      # the isolated kernel timings keep the shapes the ncu captures in profiles/ were taken at
    Md_, dd_ = kmicro * (cfg.num_patches + cfg.num_cls_tokens), cfg.dec_dim

    def time_kernel(fn, reps=30):
        for _ in range(5):
            fn()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record()
        for _ in range(reps):
            fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / reps

    X = torch.randn(Md_, dd_, device=dev).to(torch.bfloat16)
    dY = torch.randn(Md_, 4 * dd_, device=dev).to(torch.bfloat16)
    dW = torch.zeros(dd_, 4 * dd_, device=dev)
    dw_ms = time_kernel(lambda: ops.gemm(X, dY, dd_, 4 * dd_, Md_, a_mn=True, b_mn=True, out_f32=dW, accumulate=True))
    dw_tflops = 2.0 * dd_ * 4 * dd_ * Md_ / (dw_ms * 1e-3) / 1e12
    M, N, K = Md_, 4 * dd_, dd_
    W = torch.randn(K, N, device=dev).to(torch.bfloat16)
    bias = torch.zeros(N, device=dev)
    h = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
    g = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
    gemm_ms = time_kernel(lambda: ops.gemm(X, W, M, N, K, b_mn=True, out_bf16=h, out2_bf16=g, bias=bias, act=1))
    gemm_tflops = 2.0 * M * N * K / (gemm_ms * 1e-3) / 1e12
    del X, dY, dW, W, h, g
    # ---- every GEMM of one micro-step (forward + backward), each timed alone with CUDA events on its own buffers:
    # sum of algorithmic FLOPs / sum of times = the time-weighted roofline of the kernel family that dominates the step
    recorded = []
    real_gemm = ops.gemm

    def recording_gemm(A, B, M, N, K, **kw):
        recorded.append((A, B, M, N, K, kw))
        return real_gemm(A, B, M, N, K, **kw)

    ops.gemm = recording_gemm
    try:
        eng.forward(dev_images[0], dev_noise[0])
        eng.backward(zero_grads=True)
    finally:
        ops.gemm = real_gemm
    torch.cuda.synchronize()
    gemm_flops = gemm_ms_total = 0.0
    timed_shapes = {}
    for A, B, M_, N_, K_, kw in ([] if args.no_kernel_timings else recorded):
        key = (M_, N_, K_, kw.get("a_mn", False), kw.get("b_mn", False), kw.get("act", 0), kw.get("out_f32") is not None,
               kw.get("out2_bf16") is not None, kw.get("residual") is not None, bool(kw.get("accumulate", False)))
        if key not in timed_shapes:
            timed_shapes[key] = time_kernel(lambda: real_gemm(A, B, M_, N_, K_, **kw), reps=10)
        gemm_ms_total += timed_shapes[key]
        gemm_flops += 2.0 * M_ * N_ * K_
    gemm_weighted_tflops = gemm_flops / (gemm_ms_total * 1e-3) / 1e12 if gemm_ms_total > 0 else 0.0
    if os.environ.get("JM_BENCH_GEMM_TABLE") and rank == 0:     # per-shape table (stderr): where the GEMM time goes
        counts = {}
        for A, B, M_, N_, K_, kw in recorded:
            key = (M_, N_, K_, kw.get("a_mn", False), kw.get("b_mn", False), kw.get("act", 0),
                   kw.get("out_f32") is not None, kw.get("out2_bf16") is not None, kw.get("residual") is not None,
                   bool(kw.get("accumulate", False)))
            counts[key] = counts.get(key, 0) + 1
        print("count  M N K a_mn b_mn act f32 out2 res acc : ms each, TFLOP/s, share of GEMM time", file=sys.stderr)
        for key, n in sorted(counts.items(), key=lambda kv: -kv[1] * timed_shapes.get(kv[0], 0.0)):
            t = timed_shapes.get(key, 0.0)
            tf = 2.0 * key[0] * key[1] * key[2] / (t * 1e-3) / 1e12 if t > 0 else 0.0
            print(f"{n:4d} {key} : {t:.4f} ms, {tf:7.1f}, {100 * n * t / max(gemm_ms_total, 1e-9):5.1f} %",
                  file=sys.stderr)
    fwd_flops, step_flops = flops_per_image(cfg)
    step_tflops_per_gpu = step_flops * args.global_batch / world / (ms_per_step * 1e-3) / 1e12

    line = {
        "metric": "mae_pretrain_images_per_sec", "value": value, "unit": "images/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": workload_config(args, micro, accum, world),
        # the number the target is written on: whole-step algorithmic FLOPs over the step time,
        # against the measured sustained bf16 peak
        "roofline": {
            "bound": "tensor", "kernel": "whole optimizer step (forward + backward + all-reduce + AdamW), algorithmic FLOPs "
                                         f"{step_flops / 1e9:.3f} GFLOP/image",
            "achieved": step_tflops_per_gpu, "peak": peaks["bf16_sustained"], "unit": "TFLOP/s",
            "frac": step_tflops_per_gpu / peaks["bf16_sustained"], "traffic": None,
            "flops_per_image": step_flops, "peak_source": f"{peaks['source']} (sustained: long step)",
        },
        "roofline_gemm_weighted": {
            "bound": "tensor", "kernel": f"all {len(recorded)} jm_gemm launches of one micro-step ({len(timed_shapes)} distinct "
                                         "shapes / epilogues), each timed alone: sum FLOPs / sum time",
            "achieved": gemm_weighted_tflops, "peak": peaks["bf16_burst"], "unit": "TFLOP/s",
            "frac": gemm_weighted_tflops / peaks["bf16_burst"], "traffic": None,
            "peak_source": f"{peaks['source']} (burst: kernels timed alone)", "kernel_ms": gemm_ms_total,
        },
        "roofline_gemm_dw": {
            "bound": "tensor", "kernel": f"gemm2_kernel<MN,MN,f32-accumulate> (CTA pair, 256x256 tile) dW += X^T dY, decoder FFN w1: "
                                         f"M={dd_} N={4 * dd_} K={Md_} (tokens): the best instance",
            "achieved": dw_tflops, "peak": peaks["bf16_burst"], "unit": "TFLOP/s",
            "frac": dw_tflops / peaks["bf16_burst"],
            # dram__bytes_read.sum + dram__bytes_write.sum of one launch of THIS shape under ncu --set full in round 1
            # (profiles/r1f_gemm_b1024_ncu_full.md; algorithmic bytes 2*(M+N)*K + 4*M*N); null for any other shape
            "traffic": 1141.3e6 if (kmicro == 1024 and args.config == "vit_b16") else None,
            "traffic_source": "profiles/r1f_gemm_b1024_ncu_full.md" if (kmicro == 1024 and args.config == "vit_b16") else None,
            "peak_source": f"{peaks['source']} (burst: kernel timed alone)", "kernel_ms": dw_ms,
        },
        "roofline_gelu_fwd": {
            "bound": "tensor", "kernel": f"gemm2_kernel<K,MN,GELU> (CTA pair, 256x256 tile) decoder FFN w1 forward M={M} N={N} K={K}",
            "achieved": gemm_tflops, "peak": peaks["bf16_burst"], "unit": "TFLOP/s",
            "frac": gemm_tflops / peaks["bf16_burst"],
            "traffic": 1841.4e6 if (kmicro == 1024 and args.config == "vit_b16") else None,
            "traffic_source": "profiles/r2i_gemm_gelu_fwd_ncu_full.md" if (kmicro == 1024 and args.config == "vit_b16") else None,
            "kernel_ms": gemm_ms,
            "note": "epilogue-bound: writes 1.6 bytes per FLOP-byte of input (gelu'(h) + gelu(h), bf16, TMA stores); "
                    "see DESIGN.md section 4",
        },
        "step_roofline": {
            "bound": "tensor", "achieved": step_tflops_per_gpu, "peak": peaks["bf16_sustained"], "unit": "TFLOP/s",
            "frac": step_tflops_per_gpu / peaks["bf16_sustained"],
            "flops_per_image": step_flops, "peak_source": f"{peaks['source']} (sustained: long step)",
        },
        "e2e": {"value": e2e_value, "unit": "images/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": ms_e2e / args.steps},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "loss": final_loss,
    }
    if world == 1 and not args.no_cpu_baseline:
        ips, sec, cores, _ = cpu_oracle_rate(args.config, 8, 3, 1)
        line["cpu_baseline"] = {"value": ips, "unit": "images/s", "cores": cores, "kind": "port",
                                "sample": f"{args.config} MAE step, batch 8, 1 warm-up + 3 timed steps of the f32 "
                                          f"torch-CPU port in oracle/ ({sec:.2f} s/step)"}
    sampler.stop()
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


# ---------------------------------------------------------------------------------------------
def run_finetune(args):
    """--workload finetune: BASELINE.json configs[4], the ViT-B/16 Jumbo-CLS classification step at the fine-tune
    recipe (config/ft.sh): global batch 1024, label smoothing 0.1, Mixup, DropPath 0.1, AdamW with lr_decay 0.75.
    One step = forward + backward of 1024 images + gradient all-reduce + AdamW; same JSON contract."""
    from jumbo_mae_tpu_b200 import _lib, parallel
    from jumbo_mae_tpu_b200.config import MAEConfig, flops_per_image
    from jumbo_mae_tpu_b200.finetune_engine import FinetuneEngine
    from jumbo_mae_tpu_b200 import finetuning as FT
    from jumbo_mae_tpu_b200.params import ParamStore

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    parallel.init_from_env("nccl")
    import torch.distributed as dist
    _lib.require_device()
    lib = _lib.load()
    peaks = measured_peaks()
    dev = torch.device("cuda", local)
    gb = 1024 if args.global_batch == 4096 else args.global_batch
    micro = min(args.micro_batch or 1024, gb // world)      # (256 -> 1024: +3.5 % resident, +6 % end to end)
    accum = gb // (micro * world)
    droppath = 0.1
    ns = argparse.Namespace(layers=12, dim=768, heads=12, patch_size=16, image_size=224, labels=1000, weight_decay=0.05,
                            adam_b1=0.9, adam_b2=0.999, adam_eps=1e-8, learning_rate=3.0e-3, lr_decay=0.75, clip_grad=0.0,
                            grad_accum=accum, warmup_steps=1281167 * 10 // 1024, training_steps=1281167 * 110 // 1024,
                            label_smoothing=0.1, mixup=0.8, cutmix=1.0, droppath=droppath, init_seed=1, mixup_seed=1,
                            dropout_seed=1, optimizer="adamw", criterion="ce", mode="finetune")
    state = FT.create_train_state(ns).replicate()
    cfg, store = state.cfg, state.store
    eng = state.engine(micro)
    reducer = parallel.GradReducer(store) if world > 1 else None
    state.reducer = reducer
    gen = torch.Generator().manual_seed(rank)
    host_images = [torch.randint(0, 256, (micro, 3, 224, 224), generator=gen, dtype=torch.uint8).pin_memory()
                   for _ in range(accum)]
    host_labels = [torch.randint(0, 1000, (micro,), generator=gen, dtype=torch.int32).pin_memory() for _ in range(accum)]
    dev_images, dev_labels = [h.to(dev) for h in host_images], [h.to(dev) for h in host_labels]
    # DropPath factors live in a device buffer the captured graph reads; fresh draws are copied in every micro-step
    dgen = np.random.Generator(np.random.Philox(key=77 + rank))
    drop_buf = torch.ones(cfg.layers, 3, micro, device=dev)
    eng.set_droppath(drop_buf)
    eng.capture_graphs(label_smoothing=0.1, mixup=True)
    if reducer is not None:
        eng.capture_segment_graphs()
    perm = torch.arange(micro, dtype=torch.int32, device=dev).flip(0).contiguous()
    counter = {"step": 0}

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def resident_step():
        for i in range(accum):
            keep = dgen.random((cfg.layers, 3, micro)) >= droppath
            drop_buf.copy_(torch.from_numpy(keep.astype(np.float32) / np.float32(1 - droppath)), non_blocking=True)
            if reducer is not None and i == accum - 1:
                # last micro-step: the gradient all-reduce of each backward segment overlaps the next segment
                eng.run_graph_overlapped(dev_images[i], dev_labels[i], i == 0, reducer, perm=perm)
            else:
                eng.run_graph(dev_images[i], dev_labels[i], i == 0, perm=perm)
        eng.adamw_step(counter["step"], FT.lr_schedule(counter["step"], state), grad_scale=1.0 / (accum * world),
                       lr_decay=0.75)
        counter["step"] += 1

    sampler = ClockSampler(local)
    sampler.start()

    def timed(fn, steps):
        barrier()
        sampler.mark()
        l0 = lib.jm_launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        launches = lib.jm_launch_count() - l0
        ms = e0.elapsed_time(e1)
        barrier()
        t = torch.tensor([ms], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item(), launches, sampler.summary()

    for _ in range(args.warmup):
        resident_step()
    ms, launches, clocks = timed(resident_step, args.steps)
    launches += args.steps * accum * eng.graph_launches
    ms_per_step = ms / args.steps
    value = gb / (ms_per_step / 1e3)
    final_loss = eng.loss.item()

    # e2e through the public API (finetuning.training_step: eager launches, fresh Mixup / CutMix / DropPath draws,
    # host images and labels copied in, metrics read back)
    state.step, state.micro_step = counter["step"], 0

    def e2e_step():
        for i in range(accum):
            _, m = FT.training_step(state, (host_images[i], host_labels[i]))
            m["loss"].item()

    e2e_step()
    ms_e2e, _, _ = timed(e2e_step, max(1, args.steps // 2))
    e2e_value = gb / (ms_e2e / max(1, args.steps // 2) / 1e3)
    _, step_flops = flops_per_image(cfg)
    tf = step_flops * gb / world / (ms_per_step * 1e-3) / 1e12
    line = {
        "metric": "finetune_images_per_sec", "value": value, "unit": "images/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": "ViT-B/16 Jumbo CLS fine-tune step, 224 px, 196 + 3 tokens, 1000 classes, label smoothing 0.1, "
                               "Mixup, DropPath 0.1, AdamW lr_decay 0.75, global batch 1024 (BASELINE.json configs[4], "
                               "config/ft.sh)",
                   "global_batch": gb, "micro_batch": micro, "grad_accum": accum, "parallelism": f"dp{world}",
                   "l2": "inputs larger than L2: every micro-batch streams GBs of activations through a 126 MB L2"},
        "roofline": {"bound": "tensor", "kernel": "whole optimizer step, algorithmic FLOPs "
                                                   f"{step_flops / 1e9:.3f} GFLOP/image (SURVEY.md section 8d)",
                     "achieved": tf, "peak": peaks["bf16_sustained"], "unit": "TFLOP/s", "frac": tf / peaks["bf16_sustained"],
                     "traffic": None, "peak_source": f"{peaks['source']} (sustained: long step)"},
        "e2e": {"value": e2e_value, "unit": "images/s",
                "h2d_bytes_per_step": accum * (micro * 3 * 224 * 224 + micro * 4 + cfg.layers * 3 * micro * 4),
                "d2h_bytes_per_step": accum * 4, "note": "public API finetuning.training_step: eager launches"},
        "gpu_launches": int(launches), "clocks": clocks, "loss": final_loss,
    }
    sampler.stop()
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="vit_b16", choices=["vit_b16", "vit_l16", "vit_h14"])
    ap.add_argument("--global-batch", type=int, default=4096)
    ap.add_argument("--micro-batch", type=int, default=None,
                    help="per-GPU micro-batch (default: 2048 for vit_b16 -- ~95 GB of saved activations --, 512 for vit_l16, 256 for vit_h14)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--norm-pix-loss", action="store_true",
                    help="secondary measurement (SURVEY.md 8d): the norm-pix target, off in both reference presets")
    ap.add_argument("--workload", default="pretrain", choices=["pretrain", "finetune"],
                    help="pretrain: BASELINE.json configs[1..3] (the headline); finetune: configs[4]")
    ap.add_argument("--no-kernel-timings", action="store_true", help="skip the isolated GEMM re-timings (profiling runs)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    elif args.workload == "finetune":
        run_finetune(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
