"""Times the deblocking level search (svtf_lf_search_frame) on one resident 4K 10-bit picture: ms per search, launches per search."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import harness as H
S = H.S
g = S.make_geom(3840, 2160, 10)
rec = H.synth_picture(g, 0xB200C2)
mi = H.pack_mi(H.synth_mode_info(g, 0xB200C3, p_split=0.5, p_skip=0.3, p_inter=0.7))
src = H.synth_source(g, rec, 0xB200D1, amp=6)
lf = S.lf_params((24, 24), 16, 16)
with S.FilterContext(g) as ctx:
    ctx.begin(g); ctx.upload(rec); ctx.upload_source(src); ctx.set_mi(mi)
    for _ in range(3):
        lv = ctx.pick_filter_level(lf)
    l0 = ctx.launch_count
    n = 20
    t0 = time.perf_counter()
    for _ in range(n):
        lv = ctx.pick_filter_level(lf)
    dt = (time.perf_counter() - t0) / n
    print(f"lf search {dt * 1e3:.3f} ms per picture, {(ctx.launch_count - l0) // n} launches, levels {lv}")
