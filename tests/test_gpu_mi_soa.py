"""Side info as byte planes packed on the device (svtf_set_mi_soa_enc / _dec) against the host packers
(svtf_pack_mi_enc / _dec, themselves checked against the reference's tables in tests/test_host_logic.py)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("w,h,seed", [(136, 72, 1), (648, 360, 2), (1920, 1080, 3)])
def test_device_pack_equals_host_pack(H, w, h, seed):
    S = H.S
    g = S.make_geom(w, h, 10, 2)
    enc = H.synth_mode_info(g, 11000 + seed, segments=True)
    dec = H.synth_dec_mode_info(g, 11100 + seed, segments=True)
    with S.FilterContext(g) as ctx:
        ctx.begin(g)
        ctx.set_mi_soa(enc)
        assert np.array_equal(ctx.get_mi(), H.pack_mi(enc))
        ctx.set_mi_soa(dec)
        assert np.array_equal(ctx.get_mi(), H.pack_mi_dec(dec))


def test_chain_with_soa_side_info_matches_oracle(H, orc):
    S = H.S
    g = S.make_geom(328, 200, 10, 2)
    rec = H.synth_picture(g, 11201, noise=8)
    soa = H.synth_mode_info(g, 11202, p_skip=0.4)
    lf = S.lf_params((25, 25), 14, 14)
    cdef, idx = H.synth_cdef(g, 11203)
    mi = H.pack_mi(soa)
    with S.FilterContext(g) as ctx:
        ctx.begin(g)
        ctx.upload(rec)
        ctx.set_mi_soa(soa)
        ctx.loop_filter_frame(lf)
        ctx.cdef_frame(cdef, idx)
        got = ctx.download()
    exp = orc.cdef_apply(g, orc.deblock(g, rec, mi, lf), mi, cdef, idx)
    assert H.planes_equal(got, exp), H.first_diff(got, exp)


def test_invalid_enum_is_reported_at_the_next_sync(H):
    S = H.S
    g = S.make_geom(64, 64, 8, 1)
    soa = H.synth_mode_info(g, 11301)
    soa["bsize"] = soa["bsize"].copy()
    soa["bsize"][3, 5] = 200
    with S.FilterContext(g) as ctx:
        ctx.begin(g)
        ctx.set_mi_soa(soa)
        with pytest.raises(S.SvtfError):
            ctx.sync()
