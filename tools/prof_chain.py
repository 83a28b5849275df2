"""Three warm passes of the 4K 10-bit chain (deblock -> CDEF+CCSO) on one context, for ncu captures of the two chain kernels."""
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
ccso, flags = H.synth_ccso(g, 0xB200C2 + 3, quant_idx=(0, 0, 0), support=(0, 1, 2), edge_clf=(0, 0, 0), band_log2=(2, 2, 2), p_on=0.8)
with S.FilterContext(g) as c:
    c.begin(g)
    c.upload(rec)
    c.set_mi(mi)
    for _ in range(3):
        c.rewind()
        c.filter_chain(lf, cdef, idx, ccso, flags)
    c.sync()
print("ok")
