"""Loop-restoration filter search on the GPU (svtf_lr_wiener_stats, svtf_lr_search_frame) against the oracle
(oracle/svtf_oracle_lrs.c, pinned against the reference's EbRestorationPick.c in tests/test_lrs_oracle.py)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def search_case(H, w, h, bd, sb, seed, noise=10, amp=6):
    S = H.S
    g = S.make_geom(w, h, bd, sb)
    src = H.synth_picture(g, 13000 + seed, noise=noise)
    cur = H.synth_source(g, src, 13100 + seed, amp=amp)
    return g, cur, src


def vis(g, planes):
    return [np.ascontiguousarray(p[: (g.height + s) >> s, : (g.width + s) >> s]) for p, s in zip(planes, (0, 1, 1))]


def ctx_with(H, g, cur, src):
    ctx = H.S.FilterContext(g)
    ctx.begin(g)
    ctx.upload(cur)
    ctx.upload_source(src)
    return ctx


@pytest.mark.parametrize("w,h,bd,sb,win,plane,unit,seed", [
    (136, 72, 8, 1, 7, 0, 64, 1), (200, 136, 10, 2, 7, 0, 64, 2), (200, 136, 10, 2, 5, 1, 32, 3), (72, 136, 12, 2, 7, 0, 64, 4),
    (134, 70, 8, 2, 3, 2, 32, 5), (328, 200, 10, 2, 5, 0, 128, 6), (520, 300, 12, 2, 7, 0, 256, 7),
])
def test_wiener_stats_match_oracle(H, orc, w, h, bd, sb, win, plane, unit, seed):
    g, cur, src = search_case(H, w, h, bd, sb, seed)
    ctx = ctx_with(H, g, cur, src)
    M, Hm = ctx.compute_stats(plane, unit, win)
    c, s = vis(g, cur), vis(g, src)
    for u, lim in enumerate(H.lr_unit_limits(g, plane, unit)):
        a = orc.wiener_stats(g, plane, c[plane], s[plane], lim, win)
        assert np.array_equal(a[0], M[u]), (u, lim)
        assert np.array_equal(a[1], Hm[u]), (u, lim)


def test_wiener_stats_extreme_content(H, orc):
    """Largest products: a 12-bit checkerboard of 0 / 4095 (32-bit partial sums must be flushed in time)."""
    S = H.S
    g = S.make_geom(200, 136, 12, 2)
    cur = [np.zeros(((g.mi_rows * 4) >> s, (g.mi_cols * 4) >> s), np.uint16) for s in (0, 1, 1)]
    for p in cur:
        yy, xx = np.indices(p.shape)
        p[:] = np.where(((yy // 3) + (xx // 5)) & 1, 4095, 0)
    src = [4095 - p for p in cur]
    ctx = ctx_with(H, g, cur, src)
    M, Hm = ctx.compute_stats(0, 128, 7)
    c, s = vis(g, cur), vis(g, src)
    for u, lim in enumerate(H.lr_unit_limits(g, 0, 128)):
        a = orc.wiener_stats(g, 0, c[0], s[0], lim, 7)
        assert np.array_equal(a[0], M[u]) and np.array_equal(a[1], Hm[u]), u


@pytest.mark.parametrize("w,h,bd,sb,seed,kw", [
    (136, 72, 8, 1, 1, {}),
    (200, 136, 10, 2, 2, {}),
    (136, 104, 10, 2, 3, {"tap_lvl": 2}),
    (104, 72, 12, 2, 4, {"tap_lvl": 3, "one_step": 1}),
    (200, 72, 8, 2, 5, {"unit_size": (128, 64, 64), "refinement": 0}),
    (134, 70, 10, 2, 6, {"chroma": 0, "rdmult": 300000}),
    (136, 72, 8, 1, 7, {"rdmult": 40, "cost": (10, 60)}),
    (328, 200, 10, 2, 8, {"unit_size": (256, 128, 128)}),
    # self-guided tools, alone and with Wiener (-> RESTORE_SWITCHABLE)
    (136, 72, 8, 1, 9, {"wn": 0, "sg": 1}),
    (136, 104, 10, 2, 10, {"sg": 1}),
    (104, 72, 12, 2, 11, {"sg": 1, "sg_ep": ((0, 16, 4, 0), (10, 16, 1, 1))}),
    (200, 72, 8, 2, 12, {"sg": 1, "chroma": 0, "sg_ep": ((2, 12, 3, 1), (8, 16, 2, 1))}),
    (134, 70, 10, 2, 13, {"sg": 1, "sg_chroma": 0, "rdmult": 60000}),
    (136, 72, 10, 2, 14, {"sg": 1, "rdmult": 300, "sw_cost": (20, 40, 30), "sg_cost": (15, 50)}),
    (328, 200, 10, 2, 15, {"sg": 1, "unit_size": (256, 128, 128)}),
    (64, 64, 8, 1, 16, {"sg": 1}),
    (40, 24, 10, 2, 17, {"sg": 1, "sg_ep": ((0, 16, 2, 1), (0, 16, 2, 1))}),
    (16, 16, 8, 1, 18, {"sg": 1, "unit_size": (256, 128, 128)}),
])
def test_lr_search_matches_oracle(H, orc, w, h, bd, sb, seed, kw):
    g, cur, src = search_case(H, w, h, bd, sb, seed)
    cfg = H.lr_search_cfg(**kw)
    ctx = ctx_with(H, g, cur, src)
    ftype, units, search = ctx.restoration_search(cfg)
    a = H.lr_search_sig(g, cfg, ftype, units, search, ctx.sg_frame_ep_cnt)
    b = orc.lr_search(g, vis(g, cur), vis(g, src), cfg)
    for p in range(3):
        assert a[p][0] == b[p][0], ("frame type", p)
        assert a[p][2] == b[p][2], ("unit search", p)
        assert a[p][1] == b[p][1], ("unit info", p)
    assert a[3] == b[3], "sg_frame_ep_cnt"
    if kw.get("tap_lvl", 1) != 3 and w >= 100:  # the 3-tap search legitimately keeps RESTORE_NONE on this content (as the oracle does)
        assert any(t != 0 for t in ftype)


@pytest.mark.parametrize("w,h,bd,sb,seed,kw", [(136, 104, 10, 2, 31, {}), (200, 72, 8, 1, 32, {"sg": 1, "one_step": 1})])
def test_lr_search_prev_frame_coeffs(H, orc, w, h, bd, sb, seed, kw):
    g, cur, src = search_case(H, w, h, bd, sb, seed)
    _, prev = H.synth_lr(g, 900 + seed, p_none=0.3, p_wiener=0.5)
    cfg = H.lr_search_cfg(prev_units=prev, **kw)
    ctx = ctx_with(H, g, cur, src)
    ftype, units, search = ctx.restoration_search(cfg)
    assert H.lr_search_sig(g, cfg, ftype, units, search, ctx.sg_frame_ep_cnt) == orc.lr_search(g, vis(g, cur), vis(g, src), cfg)


def test_lr_search_then_apply_lowers_error(H):
    """Property at a larger size: applying what the search returns must not raise the squared error of any searched plane
    (without stripe boundaries the search's own SSE is what the frame filter would give; with them it stays close)."""
    S = H.S
    g, cur, src = search_case(H, 1920, 1080, 10, 2, 31)
    ctx = ctx_with(H, g, cur, src)
    cfg = H.lr_search_cfg(unit_size=(256, 128, 128))
    ftype, units, search = ctx.restoration_search(cfg)
    assert ftype == [1, 1, 1]
    lr = S._lib.LrParams()
    for p in range(3):
        lr.frame_restoration_type[p], lr.restoration_unit_size[p] = ftype[p], cfg.restoration_unit_size[p]
    lr.optimized_lr = 0
    ctx.loop_restoration_filter_frame(lr, units)
    out = ctx.download()
    c, s, o = vis(g, cur), vis(g, src), vis(g, out)
    for p in range(3):
        before = int(((c[p].astype(np.int64) - s[p]) ** 2).sum())
        after = int(((o[p].astype(np.int64) - s[p]) ** 2).sum())
        assert after < before, (p, before, after)


def test_lr_search_8k_properties(H):
    """BASELINE's largest size (8K 10-bit): the search runs with both tool sets, the per-unit SSEs it reports for RESTORE_NONE sum
    to the picture's squared error, every unit gets a self-guided candidate, and applying the result lowers the error of every
    plane (size-independent properties; the oracle is too slow here)."""
    S = H.S
    g = S.make_geom(7680, 4320, 10, 2)
    rng = np.random.default_rng(8)
    base = H.synth_picture(S.make_geom(960, 544, 10, 2), 13900, noise=10)
    src = [np.ascontiguousarray(np.tile(p, (8, 8))[: (g.mi_rows * 4) >> s, : (g.mi_cols * 4) >> s]) for p, s in zip(base, (0, 1, 1))]
    cur = [np.clip(p.astype(np.int32) + rng.integers(-6, 7, p.shape), 0, 1023).astype(np.uint16) for p in src]
    cfg = H.lr_search_cfg(unit_size=(256, 128, 128), sg=1, sg_ep=((0, 16, 5, 1), (10, 16, 3, 1)))
    with S.FilterContext(g) as ctx:
        ctx.begin(g)
        ctx.upload(cur)
        ctx.upload_source(src)
        ftype, units, search = ctx.restoration_search(cfg)
        assert sum(ctx.sg_frame_ep_cnt) == sum(len(u) for u in units)
        for p in range(3):
            ss = 1 if p else 0
            c, s_ = cur[p][: (g.height + ss) >> ss, : (g.width + ss) >> ss].astype(np.int64), src[p][: (g.height + ss) >> ss, : (g.width + ss) >> ss]
            assert sum(u.sse[0] for u in search[p]) == int(((c - s_) ** 2).sum())
            assert all(u.sse[2] > 0 for u in search[p])
        assert all(t != 0 for t in ftype)
        lr = S._lib.LrParams()
        for p in range(3):
            lr.frame_restoration_type[p], lr.restoration_unit_size[p] = ftype[p], cfg.restoration_unit_size[p]
        lr.optimized_lr = 1  # no deblocked picture on this context: borders from the picture itself
        ctx.loop_restoration_filter_frame(lr, units)
        out = ctx.download()
    for p in range(3):
        ss = 1 if p else 0
        v = lambda a: a[: (g.height + ss) >> ss, : (g.width + ss) >> ss].astype(np.int64)
        assert ((v(out[p]) - v(src[p])) ** 2).sum() < ((v(cur[p]) - v(src[p])) ** 2).sum(), p
