"""Device side of the reference's input path (src/dataset.py): what can leave the CPU DataLoader workers.

The reference decodes JPEGs and runs torchvision transforms on PIL images in worker processes, then ships uint8 CHW
tensors (src/dataset.py:56-75).  For the pre-training recipe the only active transforms are
``RandomResizedCrop(image_size, scale=(0.2, 1.0), interpolation=3)`` and ``RandomHorizontalFlip``
(config/pretrain/*.sh:11-18 switch everything else off).  Here the workers only decode; the random crop boxes and flip
bits are drawn on the host (``rrc_params``, the draws are data at this boundary) and the resampling itself runs on the
GPU (``random_resized_crop_flip`` -> ``jm_rrc_flip``), bit-exact with PIL's ``crop(box).resize((S, S), BICUBIC)``.
JPEG decoding stays on the CPU.
"""
from __future__ import annotations

import math
from typing import List, Sequence, Tuple

import numpy as np
import torch

from . import _lib

IMAGENET_DEFAULT_MEAN = np.array([0.485, 0.456, 0.406])     # src/dataset.py:37-38
IMAGENET_DEFAULT_STD = np.array([0.229, 0.224, 0.225])


def rrc_params(rng: np.random.Generator, height: int, width: int, scale=(0.2, 1.0), ratio=(3.0 / 4.0, 4.0 / 3.0)
               ) -> Tuple[int, int, int, int]:
    """torchvision RandomResizedCrop.get_params on a numpy Generator: (top, left, h, w)."""
    area = height * width
    log_ratio = (math.log(ratio[0]), math.log(ratio[1]))
    for _ in range(10):
        target_area = area * rng.uniform(scale[0], scale[1])
        aspect = math.exp(rng.uniform(log_ratio[0], log_ratio[1]))
        w = int(round(math.sqrt(target_area * aspect)))
        h = int(round(math.sqrt(target_area / aspect)))
        if 0 < w <= width and 0 < h <= height:
            return int(rng.integers(0, height - h + 1)), int(rng.integers(0, width - w + 1)), h, w
    in_ratio = float(width) / float(height)
    if in_ratio < min(ratio):
        w, h = width, int(round(width / min(ratio)))
    elif in_ratio > max(ratio):
        h, w = height, int(round(height * max(ratio)))
    else:
        w, h = width, height
    return (height - h) // 2, (width - w) // 2, h, w


def random_resized_crop_flip(images: Sequence[torch.Tensor], boxes: Sequence[Tuple[int, int, int, int]],
                             flips: Sequence[bool], out_size: int, device=None) -> torch.Tensor:
    """``images``: decoded uint8 HWC (RGB) tensors of any sizes (host or device); ``boxes``: (top, left, h, w) per image;
    returns the uint8 ``[B, 3, out_size, out_size]`` device batch the training step takes."""
    _lib.require_device()
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    B = len(images)
    if not (B == len(boxes) == len(flips)) or B == 0:
        raise _lib.JmError("images, boxes and flips must have the same non-zero length")
    offs, hw, total = [], [], 0
    for im, (top, left, h, w) in zip(images, boxes):
        if im.dtype != torch.uint8 or im.dim() != 3 or im.shape[2] != 3:
            raise _lib.JmError("source images must be uint8 [H, W, 3]")
        H, W = int(im.shape[0]), int(im.shape[1])
        if not (0 <= top and 0 <= left and h > 0 and w > 0 and top + h <= H and left + w <= W):
            raise _lib.JmError(f"crop box {(top, left, h, w)} outside an image of {H} x {W}")
        offs.append(total), hw.append((H, W))
        total += (H * W * 3 + 15) // 16 * 16
    packed = torch.empty(total, dtype=torch.uint8, device=dev)
    for im, off in zip(images, offs):
        n = im.numel()
        packed[off:off + n].copy_(im.reshape(-1), non_blocking=True)
    max_crop = max(max(b[2], b[3]) for b in boxes)
    lib = _lib.load()
    ws = torch.empty(lib.jm_rrc_flip_workspace(B, max_crop, out_size), dtype=torch.uint8, device=dev)
    t = lambda a, dt: torch.tensor(a, dtype=dt).to(dev)
    src_off, src_hw = t(offs, torch.int64), t(hw, torch.int32)
    bx, fl = t([list(b) for b in boxes], torch.int32), t([1 if f else 0 for f in flips], torch.uint8)
    out = torch.empty(B, 3, out_size, out_size, dtype=torch.uint8, device=dev)
    _lib.check(lib.jm_rrc_flip(packed.data_ptr(), src_off.data_ptr(), src_hw.data_ptr(), bx.data_ptr(), fl.data_ptr(), B,
                               max_crop, out_size, out.data_ptr(), ws.data_ptr(), ws.numel(), _lib.stream_ptr()))
    return out
