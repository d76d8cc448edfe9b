"""RandomResizedCrop (bicubic) + horizontal flip (src/dataset.py:56-58,70-71).

CPU: the numpy restatement (oracle/augment_oracle.py) against golden vectors produced by Pillow itself
(tests/golden/rrc_golden.npz, tests/golden/make_rrc_golden.py) and, when Pillow is importable, against Pillow live on
random boxes -- this is what pins the oracle.  GPU: jm_rrc_flip through the C ABI against the oracle and the golden
vectors, bit for bit."""
import os

import numpy as np
import pytest
import torch

from oracle import augment_oracle as A

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "rrc_golden.npz")


def _cases():
    g = np.load(GOLD)
    n = len([k for k in g.files if k.startswith("img")])
    return [(g[f"img{i}"], [int(v) for v in g[f"box{i}"]], g[f"out{i}"]) for i in range(n)]


def test_oracle_matches_pillow_goldens():
    for img, (top, left, h, w, S, flip), want in _cases():
        got = A.resized_crop_flip(img, top, left, h, w, S, bool(flip))
        assert got.dtype == np.uint8 and np.array_equal(got, want)


def test_oracle_matches_pillow_live():
    Image = pytest.importorskip("PIL.Image")
    rng = np.random.default_rng(7)
    for _ in range(12):
        H, W = int(rng.integers(20, 140)), int(rng.integers(20, 140))
        img = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
        top, left, h, w = A.rrc_params(rng, H, W)
        S = int(rng.choice([16, 24, 56]))
        flip = bool(rng.integers(0, 2))
        pil = Image.fromarray(img).crop((left, top, left + w, top + h)).resize((S, S), Image.BICUBIC)
        if flip:
            pil = pil.transpose(Image.FLIP_LEFT_RIGHT)
        want = np.asarray(pil).transpose(2, 0, 1)
        assert np.array_equal(A.resized_crop_flip(img, top, left, h, w, S, flip), want), (H, W, top, left, h, w, S)


def test_rrc_params_stay_inside_and_follow_the_recipe_scale():
    from jumbo_mae_tpu_b200.dataset import rrc_params
    rng_a, rng_b = np.random.default_rng(3), np.random.default_rng(3)
    for _ in range(200):
        H, W = int(rng_a.integers(32, 600)), int(rng_a.integers(32, 600))
        rng_b.integers(32, 600), rng_b.integers(32, 600)
        box = rrc_params(rng_a, H, W)
        assert box == A.rrc_params(rng_b, H, W)                       # product helper == oracle restatement
        top, left, h, w = box
        assert 0 <= top and 0 <= left and top + h <= H and left + w <= W and h > 0 and w > 0
        assert h * w <= H * W and (h * w >= 0.15 * H * W or (h, w) == (H, W) or h == H or w == W)


@pytest.mark.gpu
def test_rrc_flip_kernel_is_bit_exact():
    from jumbo_mae_tpu_b200.dataset import random_resized_crop_flip
    cases = _cases()
    # all golden cases that share an output size go through ONE launch (ragged sources, mixed flips)
    for S in sorted({c[1][4] for c in cases}):
        grp = [c for c in cases if c[1][4] == S]
        out = random_resized_crop_flip([torch.from_numpy(c[0]) for c in grp], [tuple(c[1][:4]) for c in grp],
                                       [bool(c[1][5]) for c in grp], S).cpu().numpy()
        for o, c in zip(out, grp):
            assert np.array_equal(o, c[2])
    rng = np.random.default_rng(11)
    imgs, boxes, flips = [], [], []
    for _ in range(9):
        H, W = int(rng.integers(40, 400)), int(rng.integers(40, 400))
        imgs.append(rng.integers(0, 256, (H, W, 3), dtype=np.uint8))
        boxes.append(A.rrc_params(rng, H, W))
        flips.append(bool(rng.integers(0, 2)))
    out = random_resized_crop_flip([torch.from_numpy(i) for i in imgs], boxes, flips, 224).cpu().numpy()
    for o, im, bx, fl in zip(out, imgs, boxes, flips):
        assert np.array_equal(o, A.resized_crop_flip(im, *bx, 224, fl))


@pytest.mark.gpu
def test_rrc_flip_rejects_bad_boxes():
    from jumbo_mae_tpu_b200._lib import JmError
    from jumbo_mae_tpu_b200.dataset import random_resized_crop_flip
    img = torch.zeros(10, 10, 3, dtype=torch.uint8)
    with pytest.raises(JmError):
        random_resized_crop_flip([img], [(5, 5, 10, 10)], [False], 8)
