"""Time split of the CDEF strength search at 4K 10-bit, 64 strengths (config C3): search kernel alone,
search + D2H of the table, device pick, host pick."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import harness as H  # noqa: E402

S = H.S


def main():
    w, h = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (3840, 2160)
    g = S.make_geom(w, h, 10)
    rec = H.synth_picture(g, 0xB200C3)
    src = H.synth_source(g, rec, 0xB200C3 + 1)
    mi = H.pack_mi(H.synth_mode_info(g, 0xB200C3 + 2))
    f1, f2 = H.cdef_level1_lists()
    cfg = S.cdef_search_cfg(f1, f2, base_q_idx=120, lambda_=3000)
    out = {}
    with S.FilterContext(g) as c:
        c.begin(g)
        c.upload(rec)
        c.upload_source(src)
        c.set_mi(mi)
        c.loop_filter_frame(S.lf_params((24, 24), 12, 12))
        c.sync()

        def timed(fn, n=5):
            fn()
            c.sync()
            t0 = time.perf_counter()
            for _ in range(n):
                fn()
            c.sync()
            return (time.perf_counter() - t0) / n * 1e3

        out["search_kernel_ms"] = round(timed(lambda: c.cdef_seg_search(cfg, want_outputs=False)), 3)
        out["search_plus_d2h_ms"] = round(timed(lambda: c.cdef_seg_search(cfg)), 3)
        mse, skip, _, _ = c.cdef_seg_search(cfg)
        out["device_pick_ms"] = round(timed(lambda: c.finish_cdef_search(cfg)), 3)
        out["table_pick_ms"] = round(timed(lambda: c.pick_table(cfg, mse, skip)), 3)
        t0 = time.perf_counter()
        S.cdef_pick_host(cfg, mse, skip)
        out["host_pick_ms"] = round((time.perf_counter() - t0) * 1e3, 3)
        nfb = skip.size
        out["nfb"] = int(nfb)
        out["fb_per_s_search_plus_pick"] = round(nfb / ((out["search_kernel_ms"] + out["device_pick_ms"]) * 1e-3), 1)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
