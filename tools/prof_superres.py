"""Two super-resolution upscales of a 1920x2160 10-bit picture to 3840x2160 (for ncu)."""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import harness as H  # noqa: E402

S = H.S
filt = np.load(os.path.join(ROOT, "tests", "golden", "superres_filters.npy"))
g, gu = S.make_geom(1920, 2160, 10), S.make_geom(3840, 2160, 10)
pic = H.synth_picture(g, 0xB20051)
with S.FilterContext(gu) as ctx:
    for _ in range(3):
        ctx.begin(g)
        ctx.upload(pic)
        ctx.superres_upscale_frame(3840, 16, filt)
        ctx.sync()
print("ok")
