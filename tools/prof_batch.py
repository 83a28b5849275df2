"""The bench's own steady state for ncu: 12 distinct resident 4K 10-bit pictures (12 x 49.8 MB in + 12 x 49.8 MB out > the 126 MB
L2), each through deblock -> CDEF+CCSO, two warm passes and one measured pass on ONE stream, then (with `search`) one
exhaustive CDEF search + pick per picture.  Run under `ncu --set full -k regex:... --launch-skip ...`; the per-launch
dram__bytes of the LAST pass are the > L2 traffic figures of profiles/traffic.json."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import harness as H  # noqa: E402

S = H.S
NP = int(os.environ.get("PROF_PICTURES", "12"))
PASSES = int(os.environ.get("PROF_PASSES", "3"))
g = S.make_geom(3840, 2160, 10)
lf = S.lf_params((24, 24), 12, 12)
cdef, idx = H.synth_cdef(g, 0xB200C2 + 2, n_strengths=8, damping=5)
ccso, flags = H.synth_ccso(g, 0xB200C2 + 3, quant_idx=(0, 0, 0), support=(0, 1, 2), edge_clf=(0, 0, 0), band_log2=(2, 2, 2), p_on=0.8)
mi = H.pack_mi(H.synth_mode_info(g, 0xB200C2 + 1, p_split=0.5, p_skip=0.3, p_inter=0.7))
ctxs = []
for i in range(NP):
    c = S.FilterContext(g)
    c.begin(g)
    c.upload(H.synth_picture(g, 0xB200C2 + 16 * i))
    c.set_mi(mi)
    ctxs.append(c)
for _ in range(PASSES):
    for c in ctxs:
        c.rewind()
        c.filter_chain(lf, cdef, idx, ccso, flags)
for c in ctxs:
    c.sync()
if "search" in sys.argv[1:]:
    f1, f2 = H.cdef_level1_lists()
    cfg = S.cdef_search_cfg(f1, f2, base_q_idx=120, lambda_=3000)
    for i, c in enumerate(ctxs[:3]):
        c.begin(g)
        rec = H.synth_picture(g, 0xB200C2 + 16 * i)
        c.upload(rec)
        c.upload_source(H.synth_source(g, rec, 0xB200C3))
        c.set_mi(mi)
        c.cdef_seg_search(cfg, want_outputs=False)
        c.finish_cdef_search(cfg)
print("ok")
