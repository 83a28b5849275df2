"""The reference DECODER's frame filters (svt_aom_dec_av1_loop_filter_frame, svt_cdef_frame: the real code, compiled into
oracle/_ref/libsvtrefdec.so) against the oracle fed with svtf_pack_mi_dec words: the decoder walks blocks and transform units
(with a transform partition inside inter blocks), the oracle walks the 4x4 grid — same pixels."""
import numpy as np
import pytest


@pytest.fixture(scope="module")
def refdec(H):
    if not H.refdec_available():
        pytest.skip("oracle/_ref/libsvtrefdec.so not built")
    return H.DecDriver()


@pytest.mark.parametrize("w,h,bd,sb,seed", [(136, 72, 8, 1, 1), (200, 136, 10, 2, 2), (64, 64, 8, 1, 3), (328, 200, 10, 2, 4), (72, 200, 12, 2, 5),
                                            (264, 136, 8, 2, 6),
                                            # odd sizes = frames coded with super-resolution / reference scaling; 145 = 8k + 1: the chroma unit at x = 72 has edges
                                            (145, 81, 8, 1, 7), (137, 75, 10, 2, 8)])
def test_decoder_deblock_matches_oracle(H, orc, refdec, w, h, bd, sb, seed):
    S = H.S
    g = S.make_geom(w, h, bd, sb)
    rec = H.synth_picture(g, 8100 + seed, noise=8)
    soa = H.synth_dec_mode_info(g, 8200 + seed, p_skip=0.4, segments=seed % 2 == 0)
    lf = S.lf_params((28, 19), 22, 9, sharpness_level=seed % 4, mode_ref_delta_enabled=seed % 2, mode_deltas=(0, 3))
    exp = orc.deblock(g, rec, H.pack_mi_dec(soa), lf)
    got = refdec.deblock(g, rec, soa, lf)
    assert H.planes_equal(got, exp), H.first_diff(got, exp)
    assert not H.planes_equal(got, rec)


def test_decoder_deblock_plane_ranges_and_zero_levels(H, orc, refdec):
    S = H.S
    g = S.make_geom(136, 104, 10, 2)
    rec = H.synth_picture(g, 8301, noise=8)
    soa = H.synth_dec_mode_info(g, 8302)
    mi = H.pack_mi_dec(soa)
    for lv, uv, ps, pe in [((20, 0), (5, 5), 0, 3), ((0, 0), (12, 0), 0, 3), ((9, 9), (0, 7), 0, 1)]:
        # (the decoder loops plane = 0 .. plane_end - plane_start - 1, EbDecLF.c:612,296: only plane_start = 0 is meaningful there)
        lf = S.lf_params(lv, uv[0], uv[1])
        got = refdec.deblock(g, rec, soa, lf, ps, pe)
        exp = orc.deblock(g, rec, mi, lf, ps, pe)
        assert H.planes_equal(got, exp), (lv, uv, ps, pe, H.first_diff(got, exp))


@pytest.mark.parametrize("w,h,bd,sb,seed", [(136, 72, 8, 1, 1), (200, 136, 10, 2, 2), (328, 200, 10, 2, 3), (264, 136, 8, 2, 4)])
def test_decoder_cdef_matches_oracle(H, orc, refdec, w, h, bd, sb, seed):
    S = H.S
    g = S.make_geom(w, h, bd, sb)
    rec = H.synth_picture(g, 8400 + seed, noise=8)
    soa = H.synth_dec_mode_info(g, 8500 + seed, p_skip=0.5)
    cdef, idx = H.synth_cdef(g, 8600 + seed, p_unset=0.2)
    exp = orc.cdef_apply(g, rec, H.pack_mi_dec(soa), cdef, idx)
    got = refdec.cdef_apply(g, rec, soa, cdef, idx)
    assert H.planes_equal(got, exp), H.first_diff(got, exp)


def synth_sb_delta(g, seed, multi):
    rng = np.random.default_rng(seed)
    d = rng.integers(-24, 25, size=((g.mi_rows + 15) // 16, (g.mi_cols + 15) // 16, 4)).astype(np.int32)
    if not multi:
        d[..., 1:] = 0
    return d


@pytest.mark.parametrize("w,h,bd,sb,multi,seed", [(200, 136, 8, 1, 0, 1), (264, 200, 10, 2, 1, 2), (136, 328, 10, 2, 1, 3)])
def test_decoder_deblock_with_sb_delta_lf(H, orc, refdec, w, h, bd, sb, multi, seed):
    """delta_lf_present: every superblock has its own deltas (SBInfo.sb_delta_lf) and an edge between two superblocks takes
    the current level from one and the neighbour level from the other."""
    S = H.S
    g = S.make_geom(w, h, bd, sb)
    rec = H.synth_picture(g, 8700 + seed, noise=8)
    soa = H.synth_dec_mode_info(g, 8800 + seed, segments=True)
    lf = S.lf_params((24, 17), 20, 11, sharpness_level=1, mode_ref_delta_enabled=1, mode_deltas=(0, 2), delta_lf_present=1, delta_lf_multi=multi)
    d = synth_sb_delta(g, 8900 + seed, multi)
    exp = orc.deblock(g, rec, H.pack_mi_dec(soa), lf, sb_delta_lf=d)
    got = refdec.deblock(g, rec, soa, lf, sb_delta_lf=d)
    assert H.planes_equal(got, exp), H.first_diff(got, exp)
    flat = orc.deblock(g, rec, H.pack_mi_dec(soa), lf)
    assert not H.planes_equal(flat, exp)  # the deltas matter


@pytest.mark.parametrize("w,h,seed,blk,bo,band,clf", [(200, 136, 1, (0, 1, 1), 0, 0, 0), (256, 256, 2, (1, 1, 1), 0, 2, 1), (136, 72, 3, (1, 0, 1), 1, 3, 0),
                                                      (248, 184, 4, (1, 1, 0), 0, 1, 0)])
def test_decoder_ccso_where_defined(H, orc, refdec, w, h, seed, blk, bo, band, clf):
    """dec_ccso_frame (Decoder/Codec/EbDecCcso.c:384) on pictures of one filter block per plane, where its block-flag lookup
    stays inside the SBInfo array: it equals the common definition with band index 0 (shift_bits = bit_depth, :51) and with
    luma switched by the V flag (:66).  blk = the (y, u, v) flags of superblock 0."""
    S = H.S
    g = S.make_geom(w, h, 8, 1)
    pre = H.synth_picture(g, 10200 + seed, noise=8)
    rec = H.synth_source(g, pre, 10300 + seed, amp=5)
    soa = H.synth_dec_mode_info(g, 10400 + seed)
    ccso, _ = H.synth_ccso(g, 10500 + seed, band_log2=(band,) * 3, bo_only=(bo,) * 3, edge_clf=(clf,) * 3)
    got = refdec.ccso_apply(g, rec, pre[0], soa, ccso, blk)
    c0, _ = H.synth_ccso(g, 10500 + seed, band_log2=(0, 0, 0), bo_only=(bo,) * 3, edge_clf=(clf,) * 3)
    flags = [np.array([blk[2]], np.uint8), np.array([blk[1]], np.uint8), np.array([blk[2]], np.uint8)]
    exp = orc.ccso_apply(g, rec, pre[0], c0, flags)
    vis = lambda pl: [a[:h >> s, :w >> s] for a, s in zip(pl, (0, 1, 1))]
    assert H.planes_equal(vis(got), vis(exp)), H.first_diff(vis(got), vis(exp))
    assert not H.planes_equal(vis(got), vis(rec))


# ---- loop restoration as the decoder runs it (Decoder/Codec/EbDecRestoration.c) ----
def _vis(g, planes):
    return [p[: (g.height + s) >> s, : (g.width + s) >> s] for p, s in zip(planes, (0, 1, 1))]


@pytest.mark.parametrize("w,h,bd,sb,seed,kw", [
    (136, 72, 8, 1, 1, {}),
    (200, 136, 10, 2, 2, {}),
    (264, 200, 10, 2, 3, {"unit_size": (128, 64, 64)}),
    (72, 136, 12, 2, 4, {}),
    (328, 200, 8, 2, 5, {"unit_size": (256, 128, 128), "p_none": 0.0}),
    (136, 200, 8, 1, 6, {"frame_type": (1, 0, 2)}),
    (200, 120, 8, 1, 7, {}),          # the last stripe ends on the crop border
    (136, 264, 10, 2, 8, {"unit_size": (64, 64, 64)}),
])
def test_decoder_lr_matches_oracle(H, orc, refdec, w, h, bd, sb, seed, kw):
    """svt_aom_dec_av1_loop_restoration_filter_frame walks SB rows x 64-column blocks with column save/restore; the oracle
    evaluates every sample from its stripe input — same pixels."""
    S = H.S
    g = S.make_geom(w, h, bd, sb)
    dbk = H.synth_picture(g, 8700 + seed, noise=10)
    cur = H.synth_source(g, dbk, 8750 + seed, amp=6)
    soa = H.synth_dec_mode_info(g, 8800 + seed)
    lr, units = H.synth_lr(g, 8850 + seed, **kw)
    exp = orc.lr_apply(g, cur, dbk, lr, units)
    got = refdec.lr_apply(g, cur, dbk, soa, lr, units)
    assert H.planes_equal(_vis(g, got), _vis(g, exp)), H.first_diff(_vis(g, got), _vis(g, exp))
    assert not H.planes_equal(_vis(g, got), _vis(g, cur))
