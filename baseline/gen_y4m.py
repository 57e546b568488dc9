#!/usr/bin/env python3
"""Synthetic 4:2:0 8-bit y4m clip for BASELINE.json configs[0] (352x288, 10 frames).

A textured background translating by a few pixels per frame plus two moving rectangles and
low-amplitude noise, so that the encoder's motion search, transform search and quantiser all
do real work.  Seeded: the same bytes every run.
"""
import sys

import numpy as np


def clip(width=352, height=288, frames=10, seed=7):
    rng = np.random.default_rng(seed)
    pad = 64
    yy, xx = np.mgrid[0:height + 2 * pad, 0:width + 2 * pad]
    base = (128 + 60 * np.sin(xx / 9.0) * np.cos(yy / 13.0) + 30 * np.sin((xx + 2 * yy) / 31.0)
            + rng.normal(0, 6, xx.shape))
    out = []
    for t in range(frames):
        ox, oy = pad + 3 * t, pad + 2 * t
        y = base[oy:oy + height, ox:ox + width].copy()
        bx, by = 40 + 7 * t, 60 + 4 * t
        y[by:by + 48, bx:bx + 64] = 200 - 0.5 * y[by:by + 48, bx:bx + 64]
        cx, cy = 250 - 5 * t, 150 + 3 * t
        y[cy:cy + 32, cx:cx + 32] = 40 + 0.3 * y[cy:cy + 32, cx:cx + 32]
        y += rng.normal(0, 1.5, y.shape)
        yq = np.clip(np.rint(y), 0, 255).astype(np.uint8)
        u = np.clip(128 + (yq[::2, ::2].astype(np.int32) - 128) // 3, 0, 255).astype(np.uint8)
        v = np.clip(128 - (yq[::2, ::2].astype(np.int32) - 128) // 4, 0, 255).astype(np.uint8)
        out.append((yq, u, v))
    return out


def write_y4m(path, width=352, height=288, frames=10, fps=30):
    with open(path, "wb") as f:
        f.write(b"YUV4MPEG2 W%d H%d F%d:1 Ip A1:1 C420jpeg\n" % (width, height, fps))
        for y, u, v in clip(width, height, frames):
            f.write(b"FRAME\n")
            f.write(y.tobytes() + u.tobytes() + v.tobytes())


if __name__ == "__main__":
    write_y4m(sys.argv[1] if len(sys.argv) > 1 else "synthetic_cif.y4m")
