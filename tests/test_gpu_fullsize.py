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
