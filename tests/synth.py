"""Seeded synthetic inputs shared by tests and bench.py."""
import numpy as np


def smooth_frame(rng, w, h, bd=8):
    """smooth texture + noise, like the survey's config-1 generator"""
    base = rng.standard_normal((h // 8 + 3, w // 8 + 3))
    img = np.kron(base, np.ones((8, 8)))[:h, :w]
    k = np.ones(5) / 5.0
    for ax in (0, 1):
        img = np.apply_along_axis(lambda v: np.convolve(v, k, mode="same"), ax, img)
    img = (img - img.min()) / max(img.max() - img.min(), 1e-9)
    hi = (1 << bd) - 1
    img = img * 0.8 * hi + 0.1 * hi + rng.normal(0, 2.0 * (1 << (bd - 8)), (h, w))
    return np.clip(np.rint(img), 0, hi).astype(np.uint8 if bd == 8 else np.uint16)


def pred_from(rng, src, bd=8, sigma=3.0):
    hi = (1 << bd) - 1
    p = np.roll(src.astype(np.int32), (1, 1), (0, 1)) + np.rint(
        rng.normal(0, sigma * (1 << (bd - 8)), src.shape)).astype(np.int32)
    return np.clip(p, 0, hi).astype(src.dtype)


def tile_blocks(plane_w, plane_h, stride, tx_w, tx_h, base_offset=0):
    """top-left element offsets of the floor(w/tw)*floor(h/th) blocks tiling a plane"""
    xs = np.arange(plane_w // tx_w) * tx_w
    ys = np.arange(plane_h // tx_h) * tx_h
    return (base_offset + ys[:, None] * stride + xs[None, :]).reshape(-1).astype(np.uint32)
