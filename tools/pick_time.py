"""Latency of finish_cdef_search on the device (svtf_cdef_pick on a resident table) for a 4K and an 8K 10-bit picture."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import harness as H
S = H.S
f1, f2 = H.cdef_level1_lists()
cfg = S.cdef_search_cfg(f1, f2, base_q_idx=120, lambda_=3000)
for (w, h) in ((3840, 2160), (7680, 4320)):
    g = S.make_geom(w, h, 10)
    rec = H.synth_picture(g, 5); src = H.synth_source(g, rec, 6); mi = H.pack_mi(H.synth_mode_info(g, 7))
    with S.FilterContext(g) as c:
        c.begin(g); c.upload(rec); c.upload_source(src); c.set_mi(mi)
        c.cdef_seg_search(cfg, want_outputs=False)
        for _ in range(3): p = c.finish_cdef_search(cfg)
        t0 = time.perf_counter(); n = 20
        for _ in range(n): p = c.finish_cdef_search(cfg)
        dt = (time.perf_counter() - t0) / n
        c.sync(); t0 = time.perf_counter()
        for _ in range(n): c.cdef_seg_search(cfg, want_outputs=False)
        c.sync(); ds = (time.perf_counter() - t0) / n
        print(f"{w}x{h}: pick {dt*1e3:.3f} ms (incl. result download + sync), search {ds*1e3:.3f} ms, bits {p[0].cdef_bits}, launches {c.launch_count}")
