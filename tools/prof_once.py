"""One warm pass + one measured pass of every kernel on the 4K 10-bit workload (for ncu --set full captures)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import harness as H  # noqa: E402

S = H.S
g = S.make_geom(3840, 2160, 10)
rec = H.synth_picture(g, 0xB200C2)
mi = H.pack_mi(H.synth_mode_info(g, 0xB200C2 + 1, p_split=0.5, p_skip=0.3, p_inter=0.7))
lf = S.lf_params((24, 24), 12, 12)
cdef, idx = H.synth_cdef(g, 0xB200C2 + 2, n_strengths=8, damping=5)
ccso, flags = H.synth_ccso(g, 0xB200C2 + 3, p_on=0.8)
src = H.synth_source(g, rec, 0xB200C3)
f1, f2 = H.cdef_level1_lists()
cfg = S.cdef_search_cfg(f1, f2, base_q_idx=120, lambda_=3000)
with S.FilterContext(g) as c:
    c.begin(g)
    c.upload(rec)
    c.upload_source(src)
    c.set_mi(mi)
    for _ in range(2):
        c.rewind()
        c.filter_chain(lf, cdef, idx, ccso, flags)
    c.rewind()
    c.loop_filter_frame(lf)
    c.cdef_seg_search(cfg, want_outputs=False)
    p = c.finish_cdef_search(cfg)
    c.sync()
    # loop restoration after deblock + CDEF, and the side-info packer
    lr, units = H.synth_lr(g, 0xB200F1)
    for _ in range(2):
        c.rewind()
        c.loop_filter_frame(lf)
        c.cdef_frame(cdef, idx)
        c.loop_restoration_filter_frame(lr, units)
    c.set_mi_soa(H.synth_mode_info(g, 0xB200C2 + 1, p_split=0.5, p_skip=0.3, p_inter=0.7))
    c.sync()
print("ok", p[0].cdef_bits)
# recon hand-off: padded reference picture with the 8-bit / 2-bit split (svtf_download_ref_padded)
with S.FilterContext(g) as c:
    c.begin(g)
    c.upload(rec)
    for _ in range(2):
        c.download_ref_padded(3840, 2160, 160, 160, 0, 0, True, True)
print("handoff ok")
