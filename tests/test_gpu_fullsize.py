"""BASELINE.json's full sizes (4K and 8K 10-bit): the chain against the oracle at 4K, and at 8K through size-independent
properties — the stripe split equals the whole picture, the chain is a pure function of the uploaded picture (rewind
and repeat), filter blocks CDEF and CCSO leave alone carry the deblocked samples through, and the padded reference
picture equals its closed form."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _inputs(H, g, seed):
    S = H.S
    rec = H.synth_picture(g, seed)
    mi = H.pack_mi(H.synth_mode_info(g, seed + 1))
    lf = S.lf_params((24, 24), 12, 12)
    cdef, idx = H.synth_cdef(g, seed + 2)
    ccso, flags = H.synth_ccso(g, seed + 3, support=(1, 3, 5), p_on=0.7)
    return rec, mi, lf, cdef, idx, ccso, flags


def test_chain_4k_10bit_matches_oracle(H, orc):
    S = H.S
    g = S.make_geom(3840, 2160, 10)
    rec, mi, lf, cdef, idx, ccso, flags = _inputs(H, g, 9100)
    with S.FilterContext(g) as ctx:
        ctx.begin(g)
        ctx.upload(rec)
        ctx.set_mi(mi)
        ctx.filter_chain(lf, cdef, idx, ccso, flags)
        got = ctx.download()
    e1 = orc.deblock(g, rec, mi, lf)
    e2 = orc.cdef_apply(g, e1, mi, cdef, idx)
    e3 = orc.ccso_apply(g, e2, e1[0], ccso, flags)
    assert H.planes_equal(got, e3), H.first_diff(got, e3)


def test_chain_8k_10bit_properties(H):
    S = H.S
    g = S.make_geom(7680, 4320, 10)
    rec, mi, lf, cdef, idx, ccso, flags = _inputs(H, g, 9200)
    idx = idx.copy()
    nv, nh = H.nfb_of(g)
    idx.reshape(nv, nh)[::3, ::2] = -1  # CDEF off on these filter blocks
    with S.FilterContext(g) as ctx:
        ctx.begin(g)
        ctx.upload(rec)
        ctx.set_mi(mi)
        ctx.filter_chain(lf, cdef, idx, ccso, flags)
        whole = ctx.download()
        ctx.rewind()
        ctx.filter_chain(lf, cdef, idx, ccso, flags)
        again = ctx.download()
        assert H.planes_equal(whole, again), "chain is not a pure function of the uploaded picture"
        # deblock alone, then CDEF alone: blocks with strength index -1 carry the deblocked samples through
        ctx.rewind()
        ctx.loop_filter_frame(lf)
        dbk = ctx.download()
        ctx.cdef_frame(cdef, idx)
        cd = ctx.download()
        off = idx.reshape(nv, nh) < 0
        for p, s in enumerate((6, 5, 5)):
            a, b = dbk[p], cd[p]
            for r, c in zip(*np.nonzero(off)):
                assert np.array_equal(a[r << s:(r + 1) << s, c << s:(c + 1) << s], b[r << s:(r + 1) << s, c << s:(c + 1) << s])
        # padded reference picture = clamped gather of the current picture
        pad, _, _ = ctx.download_ref_padded(7680, 4320, 160, 160)
        for p, s in enumerate((0, 1, 1)):
            o = 160 >> s
            yy = np.clip(np.arange(pad[p].shape[0]) - o, 0, (4320 >> s) - 1)
            xx = np.clip(np.arange(pad[p].shape[1]) - o, 0, (7680 >> s) - 1)
            assert np.array_equal(pad[p], cd[p][np.ix_(yy, xx)])
    out = [np.zeros_like(p) for p in rec]
    for rank in range(4):
        with S.StripeFilter(g, rank, 4) as sf:
            sf.chain(rec, mi, lf, cdef, idx, ccso, flags, out)
    assert H.planes_equal(out, whole), "4 stripes differ from the whole picture: " + H.first_diff(out, whole)


def _ref_for_timing_free_parity(H):
    """The reference for the full-size comparisons: the SIMD build where the host has AVX2 (pinned equal to the C-only build by
    tests/test_ref_simd.py; several times faster), else the C-only build."""
    if not H.ref_available():
        pytest.skip("oracle/_ref/libsvtref.so not built (reference tree absent)")
    return H.Ref(simd=H.ref_simd_available())


def _sig(p):
    return [p.cdef_bits, p.cdef_damping, *p.cdef_y_strength, *p.cdef_uv_strength]


def test_search_4k_10bit_matches_reference(H):
    """BASELINE configs[2] at its own size: exhaustive CDEF strength search (cdef_level 1 lists: 16 primary + 48 secondary pairs) of a
    3840x2160 10-bit picture — the full mse[2][2040][64] table, skip flags, directions / variances, and the frame-level decision
    (cdef_bits, strength pairs, per-filter-block indices) against the reference's cdef_seg_search + finish_cdef_search
    (Encoder/Codec/EbCdefProcess.c:114-357, Encoder/Codec/EbEncCdef.c:728)."""
    S = H.S
    ref = _ref_for_timing_free_parity(H)
    g = S.make_geom(3840, 2160, 10)
    rec = H.synth_picture(g, 9300)
    src = H.synth_source(g, rec, 9301)
    soa = H.synth_mode_info(g, 9302)
    first, second = H.cdef_level1_lists()
    cfg = S.cdef_search_cfg(first, second, base_q_idx=120, lambda_=3000)
    with S.FilterContext(g) as ctx:
        ctx.begin(g)
        ctx.upload(rec)
        ctx.upload_source(src)
        ctx.set_mi(H.pack_mi(soa))
        got = ctx.cdef_seg_search(cfg)
        params, idx = ctx.finish_cdef_search(cfg)
    exp = ref.cdef_search(g, rec, src, soa, cfg)
    assert got[0].shape == (2, 2040, 64)
    for a, b, name in zip(got, exp, ("mse", "skip", "dir", "var")):
        assert np.array_equal(a, b), name
    ep, eidx = ref.cdef_pick(g, soa, cfg, exp[0], exp[1])
    assert _sig(params) == _sig(ep)
    assert np.array_equal(idx, eidx)


def test_chain_8k_10bit_matches_oracle_and_search_rows_match_reference(H, orc):
    """BASELINE configs[3]'s picture (7680x4320 10-bit) against the CPU checkers, not against itself: the whole chain against the
    oracle, and the search table of the first and the last superblock-row stripes against the reference run on crops (a crop of the
    top / bottom 192 rows gives the exact table of every filter-block row except the one next to the cut: the search reads 2 rows
    beyond a filter block)."""
    S = H.S
    g = S.make_geom(7680, 4320, 10)
    rec, mi, lf, cdef, idx, ccso, flags = _inputs(H, g, 9400)
    src = H.synth_source(g, rec, 9401)
    soa = H.synth_mode_info(g, 9400 + 1)
    assert np.array_equal(H.pack_mi(soa), mi)
    first, second = H.cdef_level1_lists()
    cfg = S.cdef_search_cfg(first, second, base_q_idx=120, lambda_=3000)
    with S.FilterContext(g) as ctx:
        ctx.begin(g)
        ctx.upload(rec)
        ctx.upload_source(src)
        ctx.set_mi(mi)
        mse, skip, dirs, var = ctx.cdef_seg_search(cfg)
        ctx.filter_chain(lf, cdef, idx, ccso, flags)
        got = ctx.download()
    e1 = orc.deblock(g, rec, mi, lf)
    e2 = orc.cdef_apply(g, e1, mi, cdef, idx)
    e3 = orc.ccso_apply(g, e2, e1[0], ccso, flags)
    assert H.planes_equal(got, e3), H.first_diff(got, e3)
    if not H.ref_available():
        return
    ref = H.Ref(simd=H.ref_simd_available())
    nv, nh = H.nfb_of(g)
    rows = 192  # 3 filter-block rows; rows 0 and 1 (top crop) / the last two (bottom crop) are exact
    for top in (True, False):
        gc = S.make_geom(7680, rows, 10)
        sl = (lambda a, s: a[: rows >> s]) if top else (lambda a, s: a[a.shape[0] - (rows >> s):])
        crec = [np.ascontiguousarray(sl(p, s)) for p, s in zip(rec, (0, 1, 1))]
        csrc = [np.ascontiguousarray(sl(p, s)) for p, s in zip(src, (0, 1, 1))]
        # 4320 = 67.5 filter-block rows: the bottom crop must start on a filter-block boundary of the picture
        if not top:
            r0 = (nv - 3) * 64
            crec = [np.ascontiguousarray(p[r0 >> s:]) for p, s in zip(rec, (0, 1, 1))]
            csrc = [np.ascontiguousarray(p[r0 >> s:]) for p, s in zip(src, (0, 1, 1))]
            gc = S.make_geom(7680, 4320 - r0, 10)
        mrows = gc.mi_rows
        csoa = {k: np.ascontiguousarray(v[:mrows] if top else v[v.shape[0] - mrows:]) for k, v in soa.items()}
        emse, eskip, edir, evar = ref.cdef_search(gc, crec, csrc, csoa, cfg)
        cnv = (mrows + 15) // 16
        keep = range(0, 2) if top else range(1, cnv)  # crop fb rows that do not touch the cut
        for cr in keep:
            fr = cr if top else nv - cnv + cr
            a, b = slice(fr * nh, (fr + 1) * nh), slice(cr * nh, (cr + 1) * nh)
            assert np.array_equal(mse[:, a], emse[:, b]), (top, cr)
            assert np.array_equal(skip[a], eskip[b]) and np.array_equal(dirs[a], edir[b]) and np.array_equal(var[a], evar[b]), (top, cr)
