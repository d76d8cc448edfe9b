"""Evaluation loop of the reference's entry points (src/main_pretrain.py:37-45, src/main_finetune.py:37-45) and its
metric accumulator (src/utils.py:34-50).  Host-side only: ``validation_step`` (pretraining.py / finetuning.py of this
package) does the device work and returns SUMS over the valid rows of a batch plus ``num_samples``; this module
averages them the way the reference does -- per-key mean over the batches, then divided by the mean ``num_samples``,
i.e. total over total -- so a ragged or padded last batch is weighted by its size.  SURVEY.md section 8(f) rank 2.
"""
from __future__ import annotations

from collections import defaultdict
from typing import Callable, Dict, Iterable, Sequence

import numpy as np


def _to_float(v) -> float:
    if hasattr(v, "detach"):                 # a device tensor: one read-back per metric per batch, as device_get does
        v = v.detach().cpu().numpy()
    return float(np.asarray(v).reshape(-1)[0]) if np.asarray(v).size == 1 else np.asarray(v, dtype=np.float64)


class AverageMeter:
    """Same contract as the reference's metric accumulator (src/utils.py:34-50): ``update(**metrics)`` once per step,
    ``summary(prefix)`` reports, per key, the mean of the values seen since the last summary -- or the most recent one
    for the keys named in ``use_latest`` (the learning rate) -- and starts over.  Kept as running sums instead of
    lists of samples: a log interval of thousands of steps costs two floats per key."""

    def __init__(self, use_latest: Sequence[str] = ()):
        self.use_latest = set(use_latest)
        self._sum: Dict[str, float] = defaultdict(float)
        self._count: Dict[str, int] = defaultdict(int)
        self._last: Dict[str, float] = {}

    def update(self, **kwargs) -> None:
        for key, value in kwargs.items():
            x = _to_float(value)
            self._sum[key] += x
            self._count[key] += 1
            self._last[key] = x

    def summary(self, prefix: str = "") -> Dict[str, float]:
        out = {}
        for key, last in self._last.items():
            out[f"{prefix}{key}"] = last if key in self.use_latest else self._sum[key] / self._count[key]
        self._sum.clear(), self._count.clear(), self._last.clear()
        return out


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
