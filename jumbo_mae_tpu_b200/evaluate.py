"""Evaluation loop of the reference's entry points (src/main_pretrain.py:37-45, src/main_finetune.py:37-45) and its
metric accumulator (src/utils.py:34-50).  Host-side only: ``validation_step`` (pretraining.py / finetuning.py of this
package) does the device work and returns SUMS over the valid rows of a batch plus ``num_samples``; this module
averages them the way the reference does -- per-key mean over the batches, then divided by the mean ``num_samples``,
i.e. total over total -- so a ragged or padded last batch is weighted by its size.  SURVEY.md section 8(f) rank 2.
"""
from __future__ import annotations

from collections import defaultdict
from typing import Callable, Dict, Iterable, List, Sequence

import numpy as np


def _to_float(v) -> float:
    if hasattr(v, "detach"):                 # a device tensor: one read-back per metric per batch, as device_get does
        v = v.detach().cpu().numpy()
    return float(np.asarray(v).reshape(-1)[0]) if np.asarray(v).size == 1 else np.asarray(v, dtype=np.float64)


class AverageMeter:
    """src/utils.py:34-50: buffers every update; ``summary`` returns the mean (or the latest value for the keys in
    ``use_latest``) under ``prefix`` and clears the buffer."""

    def __init__(self, use_latest: Sequence[str] = ()):
        self.buffer: Dict[str, List] = defaultdict(list)
        self.use_latest = list(use_latest)

    def update(self, **kwargs) -> None:
        for k, v in kwargs.items():
            self.buffer[k].append(_to_float(v))

    def summary(self, prefix: str = "") -> Dict[str, float]:
        buffer = {k: np.array(v) for k, v in self.buffer.items()}
        self.buffer.clear()
        return {f"{prefix}{k}": (v[-1] if k in self.use_latest else np.mean(v)) for k, v in buffer.items()}


def evaluate(state, dataloader: Iterable, validation_step: Callable) -> Dict[str, float]:
    """``evaluate(state, dataloader)`` of both entry points: ``val/<metric>`` = sum over all batches / number of valid
    samples.  ``validation_step`` is passed in because the two steps (pre-training, fine-tune) share this loop."""
    meter = AverageMeter()
    for batch in dataloader:
        meter.update(**validation_step(state, batch))
    metrics = meter.summary("val/")
    if "val/num_samples" not in metrics:
        raise KeyError("validation_step must return 'num_samples'")
    num_samples = metrics.pop("val/num_samples")
    return {k: v / num_samples for k, v in metrics.items()}
