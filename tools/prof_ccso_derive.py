"""One CCSO derivation (ccso_search) of a 4K 10-bit picture, for `ncu --metrics gpu__time_duration.sum` launch lists and wall-clock splits."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import harness as H
import bench
S = H.S
g = S.make_geom(3840, 2160, 10)
wl = bench.workload(H, g, 0xB200C2)
src = H.synth_source(g, wl["rec"], 0xB200CC, amp=6)
with S.FilterContext(g) as c:
    c.begin(g); c.upload(wl["rec"]); c.upload_source(src); c.set_mi(wl["mi"]); c.cdef_frame(wl["cdef"], wl["idx"]); c.sync()
    for _ in range(2):
        l0 = c.launch_count; t0 = time.perf_counter()
        out, _ = c.ccso_search(4000)
        print(f"ccso_search {1e3 * (time.perf_counter() - t0):.1f} ms, {c.launch_count - l0} launches")
