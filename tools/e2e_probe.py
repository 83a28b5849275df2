"""Where does the e2e time go?  Same submit loop as bench.py's e2e leg, with parts switched off:
  up        upload only                     (H2D ceiling through the C ABI)
  up+down   upload + download, no kernels   (duplex ceiling through the C ABI)
  full      upload + mi + chain + download
Prints ms per picture for each variant."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402

import harness as H  # noqa: E402

S = H.S


def main():
    nin = int(sys.argv[1]) if len(sys.argv) > 1 else 4
    B, reps = 12, 10
    g = S.make_geom(3840, 2160, 10)
    rec = H.synth_picture(g, 1)
    mi = H.pack_mi(H.synth_mode_info(g, 2))
    lf = S.lf_params((24, 24), 12, 12)
    cdef, idx = H.synth_cdef(g, 3, n_strengths=8, damping=5)
    ccso, flags = H.synth_ccso(g, 4)
    ctxs = [S.FilterContext(g) for _ in range(nin)]
    pin_in = [S.PinnedPicture(g) for _ in range(nin)]
    pin_out = [S.PinnedPicture(g) for _ in range(nin)]
    pin_mi = [S.PinnedArray(mi.shape, np.uint32) for _ in range(nin)]
    for k in range(nin):
        for p in range(3):
            pin_in[k].planes[p][...] = rec[p]
        pin_mi[k].array[...] = mi
    L = S._lib.load()
    for c in ctxs:
        c.begin(g)
        c.upload(rec)
        c.set_mi(mi)
        c.sync()

    def submit(k, mode):
        c = ctxs[k]
        if mode in ("up", "up+down", "full"):
            c.begin(g)  # (resets the mi grid: the no-mi variants keep the one set at start-up)
        c.upload(pin_in[k].planes)
        if mode in ("full", "full-nomi"):
            if mode == "full":
                c.set_mi(pin_mi[k].array)
            c.filter_chain(lf, cdef, idx, ccso, flags)
        if mode == "kern":
            c.filter_chain(lf, cdef, idx, ccso, flags)
            return
        if mode != "up":
            c._chk(L.svtf_download_frame_async(c._h, *c._planes_args(pin_out[k].planes)))

    out = {}
    for mode in ("kern", "full-nomi", "up", "up+down", "full"):
        for rep in range(reps + 2):
            if rep == 2:
                t0 = time.perf_counter()
            for j in range(B):
                k = j % nin
                if j >= nin:
                    ctxs[k].sync()
                submit(k, mode)
            for k in range(nin):
                ctxs[k].sync()
        out[mode] = round((time.perf_counter() - t0) / (reps * B) * 1e3, 4)
    # host-side cost of the submit calls alone (time until the calls return, pipeline kept shallow)
    acc = 0.0
    for rep in range(20):
        ts = time.perf_counter()
        submit(0, "full")
        acc += time.perf_counter() - ts
        ctxs[0].sync()
    out["submit_calls_host_ms"] = round(acc * 1e3 / 20, 4)
    print(json.dumps({"inflight": nin, "ms_per_picture": out}))


if __name__ == "__main__":
    main()
