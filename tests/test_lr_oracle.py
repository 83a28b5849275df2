"""Loop restoration: the oracle's per-sample restatement against the reference's own svt_av1_loop_restoration_filter_frame +
svt_av1_loop_restoration_save_boundary_lines (Common/Codec/EbRestoration.c, convolve.c compiled into oracle/_ref)."""
import ctypes as C

import numpy as np
import pytest


def lr_case(H, w, h, bd, sb, seed, **kw):
    S = H.S
    g = S.make_geom(w, h, bd, sb)
    dbk = H.synth_picture(g, 12000 + seed, noise=10)
    cur = H.synth_source(g, dbk, 12100 + seed, amp=6)  # "after CDEF": close to, not equal to, the deblocked picture
    lr, units = H.synth_lr(g, 12200 + seed, **kw)
    return g, cur, dbk, lr, units


def vis(H, g, planes):
    return [p[: (g.height + s) >> s, : (g.width + s) >> s] for p, s in zip(planes, (0, 1, 1))]


def test_x_by_xplus1_table(H, orc, ref):
    tab = (C.c_int32 * 256).in_dll(ref.lib, "svt_aom_eb_x_by_xplus1")
    assert [orc.lib.orc_x_by_xplus1(z) for z in range(256)] == list(tab)


@pytest.mark.parametrize("w,h,bd,sb,seed,kw", [
    (136, 72, 8, 1, 1, {}),
    (200, 136, 10, 2, 2, {}),
    (264, 200, 10, 2, 3, {"unit_size": (128, 64, 64)}),
    (72, 136, 12, 2, 4, {}),
    (328, 200, 8, 2, 5, {"unit_size": (256, 128, 128), "p_none": 0.0}),
    (200, 136, 10, 2, 6, {"optimized_lr": 1}),
    (136, 200, 8, 1, 7, {"frame_type": (1, 0, 2)}),
    (134, 70, 10, 2, 8, {}),          # odd chroma size, last stripe one row short
    (200, 121, 8, 1, 9, {}),          # a stripe that ends one row above the crop border
])
def test_oracle_lr_matches_reference(H, orc, ref, w, h, bd, sb, seed, kw):
    g, cur, dbk, lr, units = lr_case(H, w, h, bd, sb, seed, **kw)
    a = orc.lr_apply(g, cur, dbk, lr, units)
    b = ref.lr_apply(g, cur, dbk, lr, units)
    assert H.planes_equal(vis(H, g, a), vis(H, g, b)), H.first_diff(vis(H, g, a), vis(H, g, b))
    assert not H.planes_equal(vis(H, g, a), vis(H, g, cur))


def test_units_match_reference(H, orc, ref):
    S = H.S
    for w, h in [(64, 64), (200, 136), (328, 200), (96, 31), (1920, 1080)]:
        g = S.make_geom(w, h, 8, 1)
        for plane in range(3):
            for us in (32, 64, 128, 256):
                a, b, c, d = C.c_int(), C.c_int(), C.c_int(), C.c_int()
                orc.lib.orc_lr_units(C.byref(g), plane, us, C.byref(a), C.byref(b))
                ref.lib.ref_lr_units(C.byref(g), plane, us, C.byref(c), C.byref(d))
                assert (a.value, b.value) == (c.value, d.value) == H.lr_units_of(g, plane, us)
