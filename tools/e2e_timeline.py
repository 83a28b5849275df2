"""GPU-side timeline of the e2e submit loop: per picture, when its upload, kernels and download finished
(CUDA events on each context's stream, ms since the first submit)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import harness as H  # noqa: E402

S = H.S


def main():
    nin = int(sys.argv[1]) if len(sys.argv) > 1 else 4
    B = 16
    g = S.make_geom(3840, 2160, 10)
    rec = H.synth_picture(g, 1)
    mi = H.pack_mi(H.synth_mode_info(g, 2))
    lf = S.lf_params((24, 24), 12, 12)
    cdef, idx = H.synth_cdef(g, 3, n_strengths=8, damping=5)
    ccso, flags = H.synth_ccso(g, 4)
    torch.cuda.init()
    ctxs = [S.FilterContext(g) for _ in range(nin)]
    streams = [torch.cuda.ExternalStream(c.cuda_stream) for c in ctxs]
    pin_in = [S.PinnedPicture(g) for _ in range(nin)]
    pin_out = [S.PinnedPicture(g) for _ in range(nin)]
    pin_mi = [S.PinnedArray(mi.shape, np.uint32) for _ in range(nin)]
    for k in range(nin):
        for p in range(3):
            pin_in[k].planes[p][...] = rec[p]
        pin_mi[k].array[...] = mi
    L = S._lib.load()
    ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731
    pipelined = len(sys.argv) > 2 and sys.argv[2] == "pipelined"

    def fetch(j):
        c = ctxs[j % nin]
        c._chk(L.svtf_download_frame_async(c._h, *c._planes_args(pin_out[j % nin].planes)))
        rows[j][3].record(streams[j % nin])

    for rep in range(3):
        rows = []
        start = ev()
        start.record(streams[0])
        for j in range(B):
            k = j % nin
            c = ctxs[k]
            if j >= nin:
                c.sync()
            e = [ev() for _ in range(4)]
            c.begin(g)
            e[0].record(streams[k])
            c.upload(pin_in[k].planes)
            c.set_mi(pin_mi[k].array)
            e[1].record(streams[k])
            c.filter_chain(lf, cdef, idx, ccso, flags)
            e[2].record(streams[k])
            rows.append(e)
            if pipelined and j >= 1:
                fetch(j - 1)
            elif not pipelined:
                fetch(j)
        if pipelined:
            fetch(B - 1)
        torch.cuda.synchronize()
    print("pic  start   up_done  kern_done  down_done   (ms)")
    for j, e in enumerate(rows):
        print(f"{j:3d} " + " ".join(f"{start.elapsed_time(x):8.3f}" for x in e))


if __name__ == "__main__":
    main()
