"""GPU: svtf_superres_upscale_frame against the oracle (pinned to the reference's svt_av1_upscale_normative_rows in
tests/test_superres_oracle.py), alone and between CDEF and loop restoration."""
import numpy as np
import pytest

from test_superres_oracle import CASES, upscale_cpu

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def filters(H):
    import os

    return np.load(os.path.join(H.ROOT, "tests", "golden", "superres_filters.npy"))


def up_geom(S, g, up_w):
    return S.make_geom_coded(up_w, g.height, 2 * ((up_w + 7) >> 3), g.mi_rows, g.bit_depth, g.sample_bytes)


@pytest.mark.parametrize("w,h,bd,sb,denom,starts", CASES + [(1920, 1080, 10, 2, 12, None)])
def test_superres_matches_oracle(H, orc, filters, w, h, bd, sb, denom, starts):
    S = H.S
    g = S.make_geom(w, h, bd, sb)
    pic = H.synth_picture(g, 15000 + w + denom, noise=12)
    up_w = (w * denom + 4) // 8
    st = list(starts) if starts else [0, g.mi_cols]
    with S.FilterContext(up_geom(S, g, up_w)) as ctx:
        ctx.begin(g)
        ctx.upload(pic)
        ctx.superres_upscale_frame(up_w, denom, filters, st)
        assert ctx.geom.width == up_w
        got = ctx.download()
    for p in range(3):
        exp = upscale_cpu(H, orc.lib.svtf_orc_superres_plane, g, pic[p], p, up_w, denom, st, filters)
        assert np.array_equal(got[p][:, : exp.shape[1]], exp), (p, np.argwhere(got[p][:, : exp.shape[1]] != exp)[:4])


def test_superres_between_cdef_and_restoration(H, orc, filters):
    """deblock -> CDEF -> upscale -> loop restoration, resident: equals the oracle's stages composed (the restoration stripe
    borders come from the UPSCALED deblocked picture, EbRestoration.c:1545-1560)."""
    S = H.S
    g = S.make_geom(264, 136, 10, 2)
    denom, up_w = 12, (264 * 12 + 4) // 8
    rec = H.synth_picture(g, 15100)
    soa = H.synth_mode_info(g, 15101)
    mi = H.pack_mi(soa)
    lf = S.lf_params((24, 24), 12, 12)
    cdef, idx = H.synth_cdef(g, 15102)
    gu = up_geom(S, g, up_w)
    lr, units = H.synth_lr(gu, 15103)
    e1 = orc.deblock(g, rec, mi, lf)
    e2 = orc.cdef_apply(g, e1, mi, cdef, idx)
    st = [0, g.mi_cols]

    def up(planes):
        out = [np.zeros(((gu.mi_rows * 4) >> s, (gu.mi_cols * 4) >> s), planes[0].dtype) for s in (0, 1, 1)]
        for p in range(3):
            u = upscale_cpu(H, orc.lib.svtf_orc_superres_plane, g, planes[p], p, up_w, denom, st, filters)
            out[p][:, : u.shape[1]] = u
            out[p][:, u.shape[1]:] = u[:, -1:]
        return out

    e3 = orc.lr_apply(gu, up(e2), up(e1), lr, units)
    with S.FilterContext(gu) as ctx:
        ctx.begin(g)
        ctx.upload(rec)
        ctx.set_mi(mi)
        ctx.loop_filter_frame(lf)
        ctx.cdef_frame(cdef, idx)
        ctx.superres_upscale_frame(up_w, denom, filters, st)
        ctx.loop_restoration_filter_frame(lr, units)
        got = ctx.download()
    vis = lambda pl: [p[: (gu.height + s) >> s, : (gu.width + s) >> s] for p, s in zip(pl, (0, 1, 1))]
    assert H.planes_equal(vis(got), vis(e3)), H.first_diff(vis(got), vis(e3))
