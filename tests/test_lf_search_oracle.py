"""Deblocking level search (svt_av1_pick_filter_level, LPF_PICK_FROM_FULL_IMAGE): the oracle restatement against the
reference's own static search_filter_level (compiled into oracle/_ref through ref_shim_dlf.c)."""
import pytest


def blocky_case(H, w, h, bd, sb, seed, step):
    """A smooth source and a 'reconstruction' with block-edge steps, so that some deblocking lowers the error."""
    import numpy as np

    S = H.S
    g = S.make_geom(w, h, bd, sb)
    src = H.synth_picture(g, seed, noise=0) if "noise" in H.synth_picture.__code__.co_varnames else H.synth_picture(g, seed)
    rng = np.random.default_rng(seed)
    rec = []
    for p in src:
        hh, ww = p.shape
        dc = rng.integers(-step, step + 1, size=((hh + 7) // 8, (ww + 7) // 8)) << (bd - 8)
        blk = np.kron(dc, np.ones((8, 8), dtype=np.int64))[:hh, :ww]
        rec.append(np.clip(p.astype(np.int64) + blk, 0, (1 << bd) - 1).astype(p.dtype))
    soa = H.synth_mode_info(g, seed + 1)
    return g, rec, src, soa


@pytest.mark.parametrize("w,h,bd,sb,start,kw", [
    (136, 72, 8, 1, (10, 10, 6, 6), {}),
    (200, 136, 10, 2, (30, 30, 20, 12), {"early_exit_convergence": 2}),
    (96, 64, 8, 2, (0, 0, 0, 0), {"only_4x4": 1}),
    (136, 104, 10, 2, (40, 12, 50, 3), {"dlf_avg": 1}),
    (72, 40, 8, 1, (5, 5, 63, 20), {"search_chroma": 0}),
])
def test_oracle_level_search_matches_reference(H, orc, ref, w, h, bd, sb, start, kw):
    S = H.S
    g, rec, src, soa = blocky_case(H, w, h, bd, sb, 7000 + w, 5)
    lf = S.lf_params((start[0], start[1]), start[2], start[3])
    a = orc.lf_search(g, rec, src, H.pack_mi(soa), lf, **kw)
    b = ref.lf_search(g, rec, src, soa, lf, **kw)
    assert a == b
