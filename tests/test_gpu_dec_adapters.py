"""GPU: the decoder-side Tier-2 adapters (svt_aom_dec_av1_loop_filter_frame, svt_cdef_frame, dec_ccso_frame over the C ABI)
called on a real EbDecHandle and compared with the REAL decoder functions called on an identical handle
(oracle/_ref/libsvtrefdec.so) and with the oracle."""
import os

import numpy as np
import pytest

from test_dec_oracle import synth_sb_delta

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def adaptdec(H):
    path = os.path.join(H.ROOT, "adapters", "_build", "libsvtadapt_dec_test.so")
    if not os.path.exists(path):
        pytest.skip("adapters/_build/libsvtadapt_dec_test.so not built (needs the reference headers)")
    return H.DecDriver(path, "adaptdec")


@pytest.fixture(scope="module")
def refdec(H):
    return H.DecDriver() if H.refdec_available() else None


@pytest.mark.parametrize("w,h,bd,sb,seed", [(136, 72, 8, 1, 1), (200, 136, 10, 2, 2), (328, 200, 10, 2, 3), (264, 136, 8, 2, 4), (72, 200, 12, 2, 5),
                                            (145, 81, 8, 1, 6), (137, 75, 10, 2, 7)])  # odd: super-resolution frames (chroma limit (w + 1) >> 1)
def test_dec_loop_filter_adapter(H, orc, adaptdec, refdec, w, h, bd, sb, seed):
    S = H.S
    g = S.make_geom(w, h, bd, sb)
    rec = H.synth_picture(g, 9300 + seed, noise=8)
    soa = H.synth_dec_mode_info(g, 9400 + seed, p_skip=0.4, segments=seed % 2 == 1)
    lf = S.lf_params((26, 18), 21, 8, sharpness_level=seed % 3, mode_ref_delta_enabled=1, mode_deltas=(0, 2))
    got = adaptdec.deblock(g, rec, soa, lf)
    exp = orc.deblock(g, rec, H.pack_mi_dec(soa), lf)
    assert H.planes_equal(got, exp), H.first_diff(got, exp)
    if refdec:
        assert H.planes_equal(got, refdec.deblock(g, rec, soa, lf))
    # luma only (monochrome call shape: plane_start 0, plane_end 1)
    got1 = adaptdec.deblock(g, rec, soa, lf, 0, 1)
    assert H.planes_equal(got1, orc.deblock(g, rec, H.pack_mi_dec(soa), lf, 0, 1))


@pytest.mark.parametrize("w,h,bd,sb,multi,seed", [(200, 136, 8, 1, 0, 1), (264, 200, 10, 2, 1, 2), (392, 328, 10, 2, 1, 3)])
def test_dec_loop_filter_adapter_sb_delta_lf(H, orc, adaptdec, refdec, w, h, bd, sb, multi, seed):
    S = H.S
    g = S.make_geom(w, h, bd, sb)
    rec = H.synth_picture(g, 9500 + seed, noise=8)
    soa = H.synth_dec_mode_info(g, 9600 + seed, segments=True)
    lf = S.lf_params((24, 17), 20, 11, sharpness_level=1, mode_ref_delta_enabled=1, mode_deltas=(0, 2), delta_lf_present=1, delta_lf_multi=multi)
    d = synth_sb_delta(g, 9700 + seed, multi)
    mi = H.pack_mi_dec(soa)
    exp = orc.deblock(g, rec, mi, lf, sb_delta_lf=d)
    got = adaptdec.deblock(g, rec, soa, lf, sb_delta_lf=d)
    assert H.planes_equal(got, exp), H.first_diff(got, exp)
    if refdec:
        assert H.planes_equal(got, refdec.deblock(g, rec, soa, lf, sb_delta_lf=d))
    # Tier 1 directly
    with S.FilterContext(g) as ctx:
        ctx.begin(g)
        ctx.upload(rec)
        ctx.set_mi(mi)
        ctx.loop_filter_frame(lf, sb_delta_lf=d)
        t1 = ctx.download()
    assert H.planes_equal(t1, exp), H.first_diff(t1, exp)


@pytest.mark.parametrize("w,h,bd,sb,seed", [(136, 72, 8, 1, 1), (200, 136, 10, 2, 2), (328, 200, 10, 2, 3), (264, 136, 8, 2, 4)])
def test_dec_cdef_adapter(H, orc, adaptdec, refdec, w, h, bd, sb, seed):
    S = H.S
    g = S.make_geom(w, h, bd, sb)
    rec = H.synth_picture(g, 9800 + seed, noise=8)
    soa = H.synth_dec_mode_info(g, 9900 + seed, p_skip=0.5)
    cdef, idx = H.synth_cdef(g, 10000 + seed, p_unset=0.2)
    got = adaptdec.cdef_apply(g, rec, soa, cdef, idx)
    exp = orc.cdef_apply(g, rec, H.pack_mi_dec(soa), cdef, idx)
    assert H.planes_equal(got, exp), H.first_diff(got, exp)
    if refdec:
        assert H.planes_equal(got, refdec.cdef_apply(g, rec, soa, cdef, idx))


def test_dec_lf_then_cdef_resident(H, orc, adaptdec):
    """LF -> CDEF with the frame kept on the device between the two calls (svtf_dec_adapter_set_resident)."""
    import ctypes as C

    S = H.S
    g = S.make_geom(264, 200, 10, 2)
    rec = H.synth_picture(g, 10101, noise=8)
    soa = H.synth_dec_mode_info(g, 10102, p_skip=0.4)
    lf = S.lf_params((22, 22), 15, 15)
    cdef, idx = H.synth_cdef(g, 10103)
    out = [p.copy() for p in rec]
    s, keep = H.dec_soa_struct(soa)
    idx8 = np.ascontiguousarray(idx, dtype=np.int8)
    assert adaptdec.lib.adaptdec_lf_cdef_resident(C.byref(g), *H._planes(out), C.byref(s), C.byref(lf), C.byref(cdef), H._p(idx8)) == 0
    mi = H.pack_mi_dec(soa)
    exp = orc.cdef_apply(g, orc.deblock(g, rec, mi, lf), mi, cdef, idx)
    assert H.planes_equal(out, exp), H.first_diff(out, exp)


@pytest.mark.parametrize("w,h,seed,blk,bo,band,clf", [(200, 136, 1, (0, 1, 1), 0, 0, 0), (256, 256, 2, (1, 1, 1), 0, 2, 1), (136, 72, 3, (1, 0, 1), 1, 3, 0),
                                                      (248, 184, 4, (1, 1, 0), 0, 1, 0)])
def test_dec_ccso_adapter_single_block(H, orc, adaptdec, refdec, w, h, seed, blk, bo, band, clf):
    """dec_ccso_frame is defined for 8-bit pictures of one filter block per plane (see svtf_dec_adapters.c); there the
    adapter reproduces it: band index 0 whatever max_band_log2 says, luma switched by the V flag."""
    S = H.S
    g = S.make_geom(w, h, 8, 1)
    pre = H.synth_picture(g, 10200 + seed, noise=8)
    rec = H.synth_source(g, pre, 10300 + seed, amp=5)
    soa = H.synth_dec_mode_info(g, 10400 + seed)
    ccso, _ = H.synth_ccso(g, 10500 + seed, band_log2=(band,) * 3, bo_only=(bo,) * 3, edge_clf=(clf,) * 3)
    got = adaptdec.ccso_apply(g, rec, pre[0], soa, ccso, blk)
    if refdec:
        exp = refdec.ccso_apply(g, rec, pre[0], soa, ccso, blk)
        assert H.planes_equal([a[:h >> s, :w >> s] for a, s in zip(got, (0, 1, 1))], [a[:h >> s, :w >> s] for a, s in zip(exp, (0, 1, 1))]), \
            H.first_diff(got, exp)
    # and it is the common definition with band 0 and the flags the decoder reads
    c0, _ = H.synth_ccso(g, 10500 + seed, band_log2=(0, 0, 0), bo_only=(bo,) * 3, edge_clf=(clf,) * 3)
    flags = [np.array([blk[2]], np.uint8), np.array([blk[1]], np.uint8), np.array([blk[2]], np.uint8)]
    exp2 = orc.ccso_apply(g, rec, pre[0], c0, flags)
    assert H.planes_equal(got, exp2), H.first_diff(got, exp2)


def _vis(g, planes):
    return [p[: (g.height + s) >> s, : (g.width + s) >> s] for p, s in zip(planes, (0, 1, 1))]


@pytest.mark.parametrize("w,h,bd,sb,seed,kw", [
    (136, 72, 8, 1, 1, {}),
    (200, 136, 10, 2, 2, {}),
    (264, 200, 10, 2, 3, {"unit_size": (128, 64, 64)}),
    (72, 136, 12, 2, 4, {}),
    (328, 200, 8, 2, 5, {"unit_size": (256, 128, 128), "p_none": 0.0}),
    (136, 200, 8, 1, 6, {"frame_type": (1, 0, 2)}),
    (200, 120, 8, 1, 7, {}),
])
def test_dec_loop_restoration_adapter(H, orc, adaptdec, refdec, w, h, bd, sb, seed, kw):
    """save_boundary_lines(0) / (1) + svt_aom_dec_av1_loop_restoration_filter_frame, stand-alone, on a bordered picture."""
    S = H.S
    g = S.make_geom(w, h, bd, sb)
    dbk = H.synth_picture(g, 10500 + seed, noise=10)
    cur = H.synth_source(g, dbk, 10550 + seed, amp=6)
    soa = H.synth_dec_mode_info(g, 10600 + seed)
    lr, units = H.synth_lr(g, 10650 + seed, **kw)
    got = adaptdec.lr_apply(g, cur, dbk, soa, lr, units)
    exp = orc.lr_apply(g, cur, dbk, lr, units)
    assert H.planes_equal(_vis(g, got), _vis(g, exp)), H.first_diff(_vis(g, got), _vis(g, exp))
    if refdec:
        ref = refdec.lr_apply(g, cur, dbk, soa, lr, units)
        assert H.planes_equal(_vis(g, got), _vis(g, ref))


@pytest.mark.parametrize("w,h,bd,sb,seed", [(264, 200, 10, 2, 1), (200, 136, 8, 1, 2)])
def test_dec_chain_lf_cdef_lr_resident(H, orc, adaptdec, w, h, bd, sb, seed):
    """The decoder's order LF -> save(0) -> CDEF -> save(1) -> LR with the frame on the device throughout."""
    import ctypes as C

    S = H.S
    g = S.make_geom(w, h, bd, sb)
    rec = H.synth_picture(g, 10700 + seed, noise=8)
    soa = H.synth_dec_mode_info(g, 10710 + seed, p_skip=0.4)
    lf = S.lf_params((22, 20), 15, 13)
    cdef, idx = H.synth_cdef(g, 10720 + seed)
    lr, units = H.synth_lr(g, 10730 + seed)
    out = [p.copy() for p in rec]
    s, keep = H.dec_soa_struct(soa)
    idx8 = np.ascontiguousarray(idx, dtype=np.int8)
    up = (C.c_void_p * 3)(*[C.cast(u, C.c_void_p) for u in units])
    assert adaptdec.lib.adaptdec_chain_lr_resident(C.byref(g), *H._planes(out), C.byref(s), C.byref(lf), C.byref(cdef), H._p(idx8),
                                                   C.byref(lr), up) == 0
    mi = H.pack_mi_dec(soa)
    d = orc.deblock(g, rec, mi, lf)
    c = orc.cdef_apply(g, d, mi, cdef, idx)
    exp = orc.lr_apply(g, c, d, lr, units)
    assert H.planes_equal(_vis(g, out), _vis(g, exp)), H.first_diff(_vis(g, out), _vis(g, exp))
