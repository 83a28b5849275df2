"""CCSO derivation on the GPU (svtf_ccso_derive_frame) against the oracle restatement and, where it travelled, the
reference's own derive_ccso_filter: same enable flag, parameters, LUT bytes and per-block flags for every plane."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("w,h,bd,sb,rdmult,amp", [(96, 64, 8, 1, 40, 6), (264, 136, 10, 2, 300, 8), (136, 264, 10, 2, 5000, 3),
                                                  (328, 200, 8, 2, 90, 10), (600, 392, 10, 2, 700, 6)])
def test_derive_after_cdef_matches_oracle(H, orc, w, h, bd, sb, rdmult, amp):
    S = H.S
    g = S.make_geom(w, h, bd, sb)
    pre = H.synth_picture(g, 9100 + w)
    mi = H.pack_mi(H.synth_mode_info(g, 9101 + w))
    cdef, idx = H.synth_cdef(g, 9102 + w)
    post = orc.cdef_apply(g, pre, mi, cdef, idx)
    org = H.synth_source(g, post, 9103 + w, amp=amp)
    with S.FilterContext(g) as ctx:
        ctx.begin(g)
        ctx.upload(pre)
        ctx.upload_source(org)
        ctx.set_mi(mi)
        ctx.cdef_frame(cdef, idx)
        got, flags = ctx.ccso_search(rdmult)
        # the derived filter applies cleanly through the same context
        ctx.ccso_frame(got, flags)
        filtered = ctx.download()
    ref = H.Ref() if H.ref_available() else None
    for plane in range(3):
        e, ef = orc.ccso_derive_plane(g, plane, pre[0], post[plane], org[plane], rdmult)
        assert H.ccso_plane_sig(got, plane) == H.ccso_plane_sig(e, plane), f"plane {plane}"
        assert np.array_equal(flags[plane], ef), f"plane {plane} flags"
        if ref is not None and w <= 328:
            r, rf = ref.ccso_derive_plane(g, plane, pre[0], post[plane], org[plane], rdmult)
            assert H.ccso_plane_sig(got, plane) == H.ccso_plane_sig(r, plane), f"plane {plane} vs reference"
            assert np.array_equal(flags[plane], rf)
    exp = orc.ccso_apply(g, post, pre[0], got, flags)
    assert H.planes_equal(filtered, exp), H.first_diff(filtered, exp)
    # the derived filter must not increase the distortion of any plane it is enabled on
    for plane in range(3):
        if got.ccso_enable[plane]:
            hh, ww = g.height >> (1 if plane else 0), g.width >> (1 if plane else 0)
            d0 = ((post[plane][:hh, :ww].astype(np.int64) - org[plane][:hh, :ww]) ** 2).sum()
            d1 = ((filtered[plane][:hh, :ww].astype(np.int64) - org[plane][:hh, :ww]) ** 2).sum()
            assert d1 <= d0


def test_derive_extremes_and_single_plane(H, orc):
    """Clipping at 0 / max inside the trial filter, classifier = the picture itself (no CDEF pass), one plane only."""
    S = H.S
    g = S.make_geom(136, 72, 10, 2)
    rng = np.random.default_rng(5)
    rec = [rng.choice(np.array([0, 1, 3, 512, 1020, 1023], dtype=np.uint16), size=((g.mi_rows * 4) >> s, (g.mi_cols * 4) >> s)) for s in (0, 1, 1)]
    org = [np.clip(p.astype(np.int32) + rng.integers(-9, 10, size=p.shape), 0, 1023).astype(np.uint16) for p in rec]
    with S.FilterContext(g) as ctx:
        ctx.begin(g)
        ctx.upload(rec)
        ctx.upload_source(org)
        got, flags = ctx.ccso_search(20, 1, 2)
        with pytest.raises(S.SvtfError):
            ctx.ccso_search(-1)
    e, ef = orc.ccso_derive_plane(g, 1, rec[0], rec[1], org[1], 20)
    assert H.ccso_plane_sig(got, 1) == H.ccso_plane_sig(e, 1)
    assert np.array_equal(flags[1], ef)
    assert not got.ccso_enable[0] and not got.ccso_enable[2]
