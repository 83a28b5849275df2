"""Deblocking level search on the GPU (svtf_lf_search_frame) against the oracle restatement of
svt_av1_pick_filter_level / search_filter_level (Encoder/Codec/EbDeblockingFilter.c:840-1240)."""
import pytest

from test_lf_search_oracle import blocky_case

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("w,h,bd,sb,start,kw", [
    (136, 72, 8, 1, (10, 10, 6, 6), {}),
    (200, 136, 10, 2, (30, 30, 20, 12), {"early_exit_convergence": 2}),
    (96, 64, 8, 2, (0, 0, 0, 0), {"only_4x4": 1}),
    (136, 104, 10, 2, (40, 12, 50, 3), {"dlf_avg": 1}),
    (72, 40, 8, 1, (5, 5, 63, 20), {"search_chroma": 0}),
    (648, 360, 10, 2, (20, 20, 10, 10), {}),
    (1000, 568, 8, 2, (12, 12, 30, 30), {"early_exit_convergence": 1}),
])
def test_level_search_matches_oracle(H, orc, w, h, bd, sb, start, kw):
    S = H.S
    g, rec, src, soa = blocky_case(H, w, h, bd, sb, 7000 + w, 5)
    mi = H.pack_mi(soa)
    lf = S.lf_params((start[0], start[1]), start[2], start[3])
    exp = orc.lf_search(g, rec, src, mi, lf, **kw)
    with S.FilterContext(g) as ctx:
        ctx.begin(g)
        ctx.upload(rec)
        ctx.upload_source(src)
        ctx.set_mi(mi)
        got = ctx.pick_filter_level(lf, **kw)
        # the search leaves the uploaded picture as it found it
        back = ctx.download()
    assert got == exp
    assert H.planes_equal([b[: p.shape[0], : p.shape[1]] for b, p in zip(back, rec)], rec)


def test_level_search_needs_source(H):
    S = H.S
    g = S.make_geom(64, 64, 8, 1)
    with S.FilterContext(g) as ctx:
        ctx.begin(g)
        ctx.upload(H.synth_picture(g, 1))
        ctx.set_mi(H.pack_mi(H.synth_mode_info(g, 2)))
        with pytest.raises(S.SvtfError):
            ctx.pick_filter_level(S.lf_params((1, 1), 1, 1))
