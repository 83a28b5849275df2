#!/usr/bin/env python
"""Synthetic 4:2:0 clip for the drop-in tests (SURVEY 8d recipe: smooth sinusoids + 16x16 block steps + noise, moving per frame).
8-bit: one byte per sample; 10-bit: little-endian 16-bit samples (the layout SvtAv1EncApp reads with --input-depth 10)."""
import sys

import numpy as np


def write_clip(path, w, h, frames, bit_depth, seed=1):
    rng = np.random.default_rng(seed)
    sh = bit_depth - 8
    dt = np.uint8 if bit_depth == 8 else np.dtype("<u2")
    mx = (1 << bit_depth) - 1
    yy, xx = np.mgrid[0:h, 0:w]
    cy, cx = np.mgrid[0:h // 2, 0:w // 2]
    with open(path, "wb") as f:
        for n in range(frames):
            y = 128 + 60 * np.sin((xx + 3 * n) / 23.0) * np.cos((yy - 2 * n) / 17.0) + (((xx // 16 + yy // 16 + n) % 5) * 9) + rng.integers(-6, 7, (h, w))
            u = 128 + 30 * np.sin((cx + n) / 31.0) + (((cx // 8 + n) % 3) * 6) + rng.integers(-3, 4, (h // 2, w // 2))
            v = 128 + 30 * np.cos((cy - n) / 29.0) + (((cy // 8 + n) % 4) * 5) + rng.integers(-3, 4, (h // 2, w // 2))
            for p in (y, u, v):
                f.write(np.clip(np.round(p * (1 << sh)), 0, mx).astype(dt).tobytes())


if __name__ == "__main__":
    write_clip(sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5]))
