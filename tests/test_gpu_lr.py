"""Loop restoration on the GPU (svtf_lr_apply_frame) against the oracle and, where it travelled, the reference's own
svt_av1_loop_restoration_filter_frame."""
import numpy as np
import pytest

from test_lr_oracle import vis

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("w,h,bd,sb,seed,kw", [
    (136, 72, 8, 1, 1, {}),
    (200, 136, 10, 2, 2, {}),
    (264, 200, 10, 2, 3, {"unit_size": (128, 64, 64)}),
    (72, 136, 12, 2, 4, {}),
    (328, 200, 8, 2, 5, {"unit_size": (256, 128, 128), "p_none": 0.0}),
    (200, 136, 10, 2, 6, {"optimized_lr": 1}),
    (136, 200, 8, 1, 7, {"frame_type": (1, 0, 2)}),
    (134, 70, 10, 2, 8, {}),
    (200, 121, 8, 1, 9, {}),
    (648, 360, 10, 2, 10, {}),
])
def test_chain_then_lr_matches_oracle(H, orc, w, h, bd, sb, seed, kw):
    """deblock -> CDEF -> loop restoration on one context: the stripe borders come from the deblocked surface."""
    S = H.S
    g = S.make_geom(w, h, bd, sb)
    rec = H.synth_picture(g, 13000 + seed, noise=10)
    mi = H.pack_mi(H.synth_mode_info(g, 13100 + seed, p_skip=0.3))
    lf = S.lf_params((20, 20), 12, 12)
    cdef, idx = H.synth_cdef(g, 13200 + seed)
    lr, units = H.synth_lr(g, 13300 + seed, **kw)
    with S.FilterContext(g) as ctx:
        ctx.begin(g)
        ctx.upload(rec)
        ctx.set_mi(mi)
        ctx.loop_filter_frame(lf)
        ctx.cdef_frame(cdef, idx)
        ctx.loop_restoration_filter_frame(lr, units)
        got = ctx.download()
    dbk = orc.deblock(g, rec, mi, lf)
    post = orc.cdef_apply(g, dbk, mi, cdef, idx)
    exp = orc.lr_apply(g, post, dbk, lr, units)
    assert H.planes_equal(vis(H, g, got), vis(H, g, exp)), H.first_diff(vis(H, g, got), vis(H, g, exp))
    # coded samples outside the visible picture are carried over unchanged
    assert H.planes_equal(got, [np.where(np.pad(np.ones_like(v, bool), ((0, p.shape[0] - v.shape[0]), (0, p.shape[1] - v.shape[1]))), e, p)
                                for v, e, p in zip(vis(H, g, exp), exp, post)])
    if H.ref_available() and w <= 328:
        r = H.Ref().lr_apply(g, post, dbk, lr, units)
        assert H.planes_equal(vis(H, g, got), vis(H, g, r))


def test_lr_without_cdef_uses_the_picture_itself(H, orc):
    S = H.S
    g = S.make_geom(200, 136, 10, 2)
    rec = H.synth_picture(g, 13401, noise=10)
    lr, units = H.synth_lr(g, 13402)
    with S.FilterContext(g) as ctx:
        ctx.begin(g)
        ctx.upload(rec)
        ctx.loop_restoration_filter_frame(lr, units)
        got = ctx.download()
    exp = orc.lr_apply(g, rec, rec, lr, units)
    assert H.planes_equal(vis(H, g, got), vis(H, g, exp)), H.first_diff(vis(H, g, got), vis(H, g, exp))
