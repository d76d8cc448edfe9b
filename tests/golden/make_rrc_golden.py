"""Generates tests/golden/rrc_golden.npz with Pillow itself (the library the reference's torchvision transform calls,
src/dataset.py:57,70): a few synthetic uint8 images, crop boxes that exercise up- and down-sampling and both edges,
``img.crop(box).resize((S, S), Image.BICUBIC)`` (+ FLIP_LEFT_RIGHT), CHW like PILToTensor.
Run in the build container (Pillow 12.2.0): python tests/golden/make_rrc_golden.py"""
import os

import numpy as np
from PIL import Image

HERE = os.path.dirname(os.path.abspath(__file__))
rng = np.random.default_rng(2024)
cases = [  # (H, W, top, left, h, w, S, flip)
    (64, 80, 0, 0, 64, 80, 32, False),       # whole image, down-sampling ~2x
    (64, 80, 5, 7, 20, 30, 32, True),        # up-sampling, flipped
    (96, 72, 10, 3, 77, 61, 48, False),      # non-integer scale both ways
    (50, 50, 49, 49, 1, 1, 16, False),       # a single source pixel
    (120, 90, 0, 30, 120, 45, 28, True),     # tall crop: down 4.3x vertically, up horizontally
    (300, 400, 17, 40, 250, 333, 224, False),  # the recipe's output size
]
out = {}
for i, (H, W, top, left, h, w, S, flip) in enumerate(cases):
    # smooth gradients + noise + hard edges, so that negative bicubic lobes clip at 0 and 255
    yy, xx = np.mgrid[0:H, 0:W]
    base = np.stack([(xx * 255 // max(W - 1, 1)), (yy * 255 // max(H - 1, 1)), ((xx // 8 + yy // 8) % 2) * 255], axis=-1)
    img = np.clip(base + rng.integers(-40, 41, base.shape), 0, 255).astype(np.uint8)
    pil = Image.fromarray(img).crop((left, top, left + w, top + h)).resize((S, S), Image.BICUBIC)
    if flip:
        pil = pil.transpose(Image.FLIP_LEFT_RIGHT)
    out[f"img{i}"] = img
    out[f"box{i}"] = np.array([top, left, h, w, S, int(flip)], dtype=np.int32)
    out[f"out{i}"] = np.asarray(pil).transpose(2, 0, 1).copy()
np.savez_compressed(os.path.join(HERE, "rrc_golden.npz"), **out)
print("wrote", len(cases), "cases")
