"""Data parallelism: one process per GPU, gradients averaged with NCCL over NVLink.

Replaces ``jax.pmap(axis_name="batch")`` + ``lax.pmean`` (src/pretraining.py:125,144,150).
Gradients live in one flat f32 buffer (params.ParamStore.g), so the all-reduce is a handful of
large bucketed NCCL calls on a side stream; the 1/world factor is folded into AdamW.
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist


def process_index() -> int:
    return dist.get_rank() if dist.is_available() and dist.is_initialized() else 0


def process_count() -> int:
    return dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1


def init_from_env(backend: str | None = None) -> None:
    """Initialises torch.distributed from torchrun's environment (RANK / WORLD_SIZE / MASTER_*)."""
    if dist.is_initialized() or int(os.environ.get("WORLD_SIZE", "1")) <= 1:
        return
    if backend is None:
        backend = "nccl" if torch.cuda.is_available() else "gloo"
    if backend == "nccl":
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    dist.init_process_group(backend=backend)


def bucket_ranges(total: int, bucket_elems: int):
    """[start, end) element ranges covering [0, total) from the END of the buffer backwards.
    The flat layout puts late-layer (decoder) leaves at high offsets, which is the order the
    backward pass finishes them in."""
    out, end = [], total
    while end > 0:
        start = max(0, end - bucket_elems)
        out.append((start, end))
        end = start
    return out


class GradReducer:
    """Sum-all-reduce of the flat gradient buffer in buckets on a dedicated stream."""

    def __init__(self, store, bucket_mb: int = 64):
        self.store = store
        self.bucket_elems = bucket_mb * (1 << 20) // 4
        self.ranges = bucket_ranges(store.total, self.bucket_elems)
        self.stream = torch.cuda.Stream() if store.g.is_cuda else None

    def all_reduce_grads(self) -> None:
        if process_count() == 1:
            return
        g = self.store.g
        if self.stream is None:
            for s, e in self.ranges:
                dist.all_reduce(g[s:e])
            return
        self.stream.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(self.stream):
            for s, e in self.ranges:
                dist.all_reduce(g[s:e])
        torch.cuda.current_stream().wait_stream(self.stream)

    def all_reduce_ranges(self, ranges, after_event=None) -> None:
        """Sum-all-reduces the given [start, end) ranges of the gradient buffer on the communication stream,
        after ``after_event`` (recorded on the compute stream when those gradients became final).  The compute
        stream is NOT blocked: call ``join()`` before the optimizer reads the gradients."""
        if process_count() == 1:
            return
        g = self.store.g
        if self.stream is None:
            for s, e in ranges:
                dist.all_reduce(g[s:e])
            return
        if after_event is not None:
            self.stream.wait_event(after_event)
        else:
            self.stream.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(self.stream):
            for s, e in ranges:
                for b in range(s, e, self.bucket_elems):
                    dist.all_reduce(g[b:min(e, b + self.bucket_elems)])

    def join(self) -> None:
        if self.stream is not None:
            torch.cuda.current_stream().wait_stream(self.stream)

    def mean_scalar(self, t: torch.Tensor) -> None:
        if process_count() == 1:
            return
        dist.all_reduce(t)
        t.div_(process_count())
