"""One warm + one measured loop-restoration filter search on the 4K 10-bit workload (for ncu launch lists / --set full)."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import harness as H  # noqa: E402

S = H.S
g = S.make_geom(3840, 2160, 10)
rec = H.synth_picture(g, 0xB200C2)
src = H.synth_source(g, rec, 0xB200F1, amp=6)
cfg = H.lr_search_cfg(unit_size=(256, 128, 128))
with S.FilterContext(g) as c:
    c.begin(g)
    c.upload(src)
    c.upload_source(rec)
    for i in range(int(os.environ.get("N", "2"))):
        t0 = time.perf_counter()
        ft, units, search = c.restoration_search(cfg)
        print("search ms", round((time.perf_counter() - t0) * 1e3, 3), ft)
    t0 = time.perf_counter()
    c.compute_stats(0, 256, 7)
    print("luma stats ms", round((time.perf_counter() - t0) * 1e3, 3))
    cfg2 = H.lr_search_cfg(unit_size=(256, 128, 128), sg=1)
    for i in range(int(os.environ.get("N", "2"))):
        t0 = time.perf_counter()
        ft, units, search = c.restoration_search(cfg2)
        print("search (Wiener + self-guided) ms", round((time.perf_counter() - t0) * 1e3, 3), ft, list(c.sg_frame_ep_cnt))
