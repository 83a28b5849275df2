"""One CDEF apply (+ fused CCSO) of a small picture through the C ABI against the oracle; argv: width height bit_depth sample_bytes."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import harness as H
S = H.S
w, h, bd, sb = (int(x) for x in sys.argv[1:5])
g = S.make_geom(w, h, bd, sb)
rec = H.synth_picture(g, 1)
mi = H.pack_mi(H.synth_mode_info(g, 2))
cdef, idx = H.synth_cdef(g, 3)
ccso, flags = H.synth_ccso(g, 4)
orc = H.Oracle()
e = orc.cdef_apply(g, rec, mi, cdef, idx)
e2 = orc.ccso_apply(g, e, rec[0], ccso, flags)
with S.FilterContext(g) as ctx:
    ctx.begin(g); ctx.upload(rec); ctx.set_mi(mi)
    ctx.cdef_frame(cdef, idx)
    got = ctx.download()
    print("cdef", w, h, bd, sb, "equal" if H.planes_equal(got, e) else "DIFF " + H.first_diff(got, e))
    ctx.rewind()
    ctx.filter_chain(None, cdef, idx, ccso, flags)
    got = ctx.download()
    print("cdef+ccso", "equal" if H.planes_equal(got, e2) else "DIFF " + H.first_diff(got, e2))
