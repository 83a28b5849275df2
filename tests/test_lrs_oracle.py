"""Loop-restoration filter search (Wiener and self-guided tools): the oracle's restatement (oracle/svtf_oracle_lrs.c) against the reference's
own Encoder/Codec/EbRestorationPick.c (restoration_seg_search + rest_finish_search, svt_av1_compute_stats[_highbd]_c,
wiener_decompose_sep_sym / finalize_sym_filter / compute_score) compiled into oracle/_ref (oracle/ref_shim_lrs.c)."""
import numpy as np
import pytest


def search_case(H, w, h, bd, sb, seed, noise=10, amp=6):
    S = H.S
    g = S.make_geom(w, h, bd, sb)
    src = H.synth_picture(g, 13000 + seed, noise=noise)
    cur = H.synth_source(g, src, 13100 + seed, amp=amp)  # the "reconstruction": the source + coding noise
    return g, cur, src


def vis(g, planes):
    return [np.ascontiguousarray(p[: (g.height + s) >> s, : (g.width + s) >> s]) for p, s in zip(planes, (0, 1, 1))]


@pytest.mark.parametrize("w,h,bd,sb,win,plane,seed", [
    (136, 72, 8, 1, 7, 0, 1), (136, 72, 10, 2, 7, 0, 2), (136, 72, 10, 2, 5, 1, 3), (72, 136, 12, 2, 7, 0, 4), (134, 70, 8, 2, 3, 2, 5),
    (200, 100, 8, 1, 5, 0, 6),
])
def test_oracle_wiener_stats_match_reference(H, orc, ref, w, h, bd, sb, win, plane, seed):
    g, cur, src = search_case(H, w, h, bd, sb, seed)
    cur, src = vis(g, cur), vis(g, src)
    for lim in H.lr_unit_limits(g, plane, 64 if plane == 0 else 32):
        a = orc.wiener_stats(g, plane, cur[plane], src[plane], lim, win)
        b = ref.wiener_stats(g, plane, cur[plane], src[plane], lim, win)
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]), lim
        assert orc.wiener_from_stats(win, *a) == ref.wiener_from_stats(win, *b)


def test_oracle_wiener_solve_on_degenerate_stats(H, orc, ref):
    rng = np.random.default_rng(77)
    for win in (7, 5, 3):
        n = win * win
        for k in range(12):
            if k == 0:
                M, Hm = np.zeros(n, np.int64), np.zeros((n, n), np.int64)  # flat unit: every pivot is zero
            else:
                a = rng.integers(-(1 << (8 + k)), 1 << (8 + k), size=(n, n)).astype(np.int64)
                Hm = a @ a.T if k % 2 else a + a.T
                M = rng.integers(-(1 << (12 + 2 * k)), 1 << (12 + 2 * k), size=n).astype(np.int64)
            assert orc.wiener_from_stats(win, M, Hm) == ref.wiener_from_stats(win, M, Hm), (win, k)


@pytest.mark.parametrize("w,h,bd,sb,seed,kw", [
    (136, 72, 8, 1, 1, {}),
    (200, 136, 10, 2, 2, {}),
    (136, 104, 10, 2, 3, {"tap_lvl": 2}),
    (104, 72, 12, 2, 4, {"tap_lvl": 3, "one_step": 1}),
    (200, 72, 8, 2, 5, {"unit_size": (128, 64, 64), "refinement": 0}),
    (134, 70, 10, 2, 6, {"chroma": 0, "rdmult": 300000}),
    (136, 72, 8, 1, 7, {"rdmult": 40, "cost": (10, 60)}),
    # self-guided tools, alone and with Wiener (-> RESTORE_SWITCHABLE)
    (136, 72, 8, 1, 8, {"wn": 0, "sg": 1}),
    (136, 104, 10, 2, 9, {"sg": 1}),
    (104, 72, 12, 2, 10, {"sg": 1, "sg_ep": ((0, 16, 4, 0), (10, 16, 1, 1))}),
    (200, 72, 8, 2, 11, {"sg": 1, "chroma": 0, "sg_ep": ((2, 12, 3, 1), (8, 16, 2, 1))}),
    (134, 70, 10, 2, 12, {"sg": 1, "sg_chroma": 0, "rdmult": 60000}),
    (136, 72, 10, 2, 13, {"sg": 1, "rdmult": 300, "sw_cost": (20, 40, 30), "sg_cost": (15, 50)}),
    # one restoration unit per plane (RESTORE_SWITCHABLE is not a frame type then, :1590), tiny pictures
    (64, 64, 8, 1, 14, {"sg": 1}),
    (40, 24, 10, 2, 15, {"sg": 1, "sg_ep": ((0, 16, 2, 1), (0, 16, 2, 1))}),
    (16, 16, 8, 1, 16, {"sg": 1, "unit_size": (256, 128, 128)}),
])
def test_oracle_lr_search_matches_reference(H, orc, ref, w, h, bd, sb, seed, kw):
    g, cur, src = search_case(H, w, h, bd, sb, seed)
    cfg = H.lr_search_cfg(**kw)
    a = orc.lr_search(g, vis(g, cur), vis(g, src), cfg)
    b = ref.lr_search(g, vis(g, cur), vis(g, src), cfg)
    for p in range(3):
        assert a[p][0] == b[p][0], ("frame type", p)
        assert a[p][2] == b[p][2], ("unit search", p)
        assert a[p][1] == b[p][1], ("unit info", p)
    assert a[3] == b[3], "sg_frame_ep_cnt"


@pytest.mark.parametrize("w,h,bd,sb,seed,kw", [(136, 104, 10, 2, 31, {}), (200, 72, 8, 1, 32, {"sg": 1, "one_step": 1})])
def test_oracle_lr_search_prev_frame_coeffs(H, orc, ref, w, h, bd, sb, seed, kw):
    """wn_filter_ctrls.use_prev_frame_coeffs on an inter frame: units that were RESTORE_WIENER in the reference picture start
    from its taps (no statistics, no score check), the others are solved."""
    g, cur, src = search_case(H, w, h, bd, sb, seed)
    _, prev = H.synth_lr(g, 900 + seed, p_none=0.3, p_wiener=0.5)
    cfg = H.lr_search_cfg(prev_units=prev, **kw)
    a = orc.lr_search(g, vis(g, cur), vis(g, src), cfg)
    b = ref.lr_search(g, vis(g, cur), vis(g, src), cfg)
    assert a == b
    assert a != orc.lr_search(g, vis(g, cur), vis(g, src), H.lr_search_cfg(**kw))


def test_product_host_solve_matches_oracle(H, orc):
    """svtf_lr_wiener_from_stats (the product's host code, no GPU needed) against the oracle on real and on degenerate statistics."""
    S = H.S
    g, cur, src = search_case(H, 136, 104, 10, 2, 21)
    cur, src = vis(g, cur), vis(g, src)
    for plane, win in ((0, 7), (0, 5), (1, 5), (2, 3)):
        for lim in H.lr_unit_limits(g, plane, 64 if plane == 0 else 32):
            M, Hm = orc.wiener_stats(g, plane, cur[plane], src[plane], lim, win)
            assert S.wiener_from_stats(win, M, Hm) == orc.wiener_from_stats(win, M, Hm)
    rng = np.random.default_rng(78)
    for win in (7, 5, 3):
        n = win * win
        for k in range(12):
            a = rng.integers(-(1 << (8 + k)), 1 << (8 + k), size=(n, n)).astype(np.int64)
            Hm = (a @ a.T if k % 2 else a + a.T) if k else np.zeros((n, n), np.int64)
            M = rng.integers(-(1 << (12 + 2 * k)), 1 << (12 + 2 * k), size=n).astype(np.int64) if k else np.zeros(n, np.int64)
            assert S.wiener_from_stats(win, M, Hm) == orc.wiener_from_stats(win, M, Hm), (win, k)
