"""CPU restatement of the pre-training recipe's image augmentation (TEST INFRASTRUCTURE: only tests/, smoke() and the
cpu_baseline legs may import this package).

The reference builds its training transform from torchvision (src/dataset.py:56-58,70-71):
``T.RandomResizedCrop(image_size, scale=(0.2, 1.0), interpolation=3)`` + ``T.RandomHorizontalFlip()`` on PIL images,
i.e. ``img.crop(box).resize((S, S), BICUBIC)`` and ``img.transpose(FLIP_LEFT_RIGHT)``.  The arithmetic lives in a
third-party dependency that is not vendored in /root/reference: Pillow's ``ImagingResample`` (src/libImaging/Resample.c,
``precompute_coeffs`` / ``normalize_coeffs_8bpc`` / ``ImagingResampleHorizontal_8bpc`` / ``...Vertical_8bpc``;
Pillow is unpinned in scripts/setup.sh, 12.2.0 in this image).  Its published algorithm is restated below in numpy
(float64 coefficient arithmetic in Pillow's operation order, 22-bit fixed point, two separable passes with a uint8
intermediate).  Pinned: tests/test_augment.py checks this restatement against Pillow itself (when importable) and
against golden vectors that Pillow produced (tests/golden/rrc_golden.npz, made by tests/golden/make_rrc_golden.py).
"""
from __future__ import annotations

import math

import numpy as np

PRECISION_BITS = 32 - 8 - 2          # Resample.c: coefficients are 22-bit fixed point for 8-bit channels
BICUBIC_SUPPORT = 2.0


def bicubic_filter(x: float) -> float:
    """Resample.c bicubic_filter (a = -0.5), same operation order."""
    a = -0.5
    if x < 0.0:
        x = -x
    if x < 1.0:
        return ((a + 2.0) * x - (a + 3.0)) * x * x + 1
    if x < 2.0:
        return (((x - 5) * x + 8) * x - 4) * a
    return 0.0


def precompute_coeffs(in_size: int, out_size: int):
    """Resample.c precompute_coeffs + normalize_coeffs_8bpc for box = (0, in_size): returns (ksize, bounds [out, 2]
    (xmin, count), integer coefficients [out, ksize])."""
    scale = float(in_size) / out_size
    filterscale = scale if scale >= 1.0 else 1.0
    support = BICUBIC_SUPPORT * filterscale
    ksize = int(math.ceil(support)) * 2 + 1
    bounds = np.zeros((out_size, 2), dtype=np.int32)
    kk = np.zeros((out_size, ksize), dtype=np.int32)
    ss = 1.0 / filterscale
    for xx in range(out_size):
        center = 0.0 + (xx + 0.5) * scale
        xmin = int(center - support + 0.5)            # C cast: truncation toward zero
        if xmin < 0:
            xmin = 0
        xmax = int(center + support + 0.5)
        if xmax > in_size:
            xmax = in_size
        xmax -= xmin
        w = [bicubic_filter((x + xmin - center + 0.5) * ss) for x in range(xmax)]
        ww = 0.0
        for v in w:
            ww += v
        if ww != 0.0:
            w = [v / ww for v in w]
        for x, v in enumerate(w):
            kk[xx, x] = int(-0.5 + v * (1 << PRECISION_BITS)) if v < 0 else int(0.5 + v * (1 << PRECISION_BITS))
        bounds[xx] = (xmin, xmax)
    return ksize, bounds, kk


def _clip8(v: np.ndarray) -> np.ndarray:
    return np.clip(v >> PRECISION_BITS, 0, 255).astype(np.uint8)     # arithmetic shift, then the clip8 lookup table


def resize_bicubic(img: np.ndarray, out_size: int) -> np.ndarray:
    """PIL.Image.resize((out_size, out_size), BICUBIC) of a uint8 HWC image: horizontal pass, then vertical pass."""
    h, w, _ = img.shape
    _, bx, kx = precompute_coeffs(w, out_size)
    _, by, ky = precompute_coeffs(h, out_size)
    src = img.astype(np.int64)
    tmp = np.empty((h, out_size, img.shape[2]), dtype=np.uint8)
    for xx in range(out_size):
        x0, n = bx[xx]
        acc = (1 << (PRECISION_BITS - 1)) + (src[:, x0:x0 + n, :] * kx[xx, :n, None].astype(np.int64)).sum(axis=1)
        tmp[:, xx, :] = _clip8(acc)
    out = np.empty((out_size, out_size, img.shape[2]), dtype=np.uint8)
    t = tmp.astype(np.int64)
    for yy in range(out_size):
        y0, n = by[yy]
        acc = (1 << (PRECISION_BITS - 1)) + (t[y0:y0 + n, :, :] * ky[yy, :n, None, None].astype(np.int64)).sum(axis=0)
        out[yy] = _clip8(acc)
    return out


def resized_crop_flip(img: np.ndarray, top: int, left: int, height: int, width: int, out_size: int, flip: bool
                      ) -> np.ndarray:
    """torchvision F.resized_crop(img, top, left, height, width, (S, S), BICUBIC) followed by an optional horizontal
    flip, then PILToTensor: uint8 HWC in, uint8 CHW out (src/dataset.py:57,70,74)."""
    out = resize_bicubic(np.ascontiguousarray(img[top:top + height, left:left + width]), out_size)
    if flip:
        out = out[:, ::-1]
    return np.ascontiguousarray(out.transpose(2, 0, 1))


def rrc_params(rng: np.random.Generator, height: int, width: int, scale=(0.2, 1.0), ratio=(3.0 / 4.0, 4.0 / 3.0)):
    """torchvision RandomResizedCrop.get_params restated on a numpy Generator (the reference draws from torch's global
    RNG in a DataLoader worker; the draws are data at this boundary): returns (top, left, h, w)."""
    area = height * width
    log_ratio = (math.log(ratio[0]), math.log(ratio[1]))
    for _ in range(10):
        target_area = area * rng.uniform(scale[0], scale[1])
        aspect = math.exp(rng.uniform(log_ratio[0], log_ratio[1]))
        w = int(round(math.sqrt(target_area * aspect)))
        h = int(round(math.sqrt(target_area / aspect)))
        if 0 < w <= width and 0 < h <= height:
            return int(rng.integers(0, height - h + 1)), int(rng.integers(0, width - w + 1)), h, w
    in_ratio = float(width) / float(height)          # fallback: central crop with the ratio clamped
    if in_ratio < min(ratio):
        w, h = width, int(round(width / min(ratio)))
    elif in_ratio > max(ratio):
        h, w = height, int(round(height * max(ratio)))
    else:
        w, h = width, height
    return (height - h) // 2, (width - w) // 2, h, w
