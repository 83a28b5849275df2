"""Recon hand-off on the GPU (svtf_download_ref_padded, svtf_upload_frame_8as16) against the oracle."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

CASES = [  # (width, height, bit_depth, sample_bytes, org_x, org_y, pad_right, pad_bottom)
    (328, 200, 8, 1, 32, 16, 0, 0),
    (328, 200, 10, 2, 96, 96, 0, 0),
    (322, 198, 10, 2, 40, 24, 6, 2),
    (200, 136, 8, 2, 8, 4, 0, 0),
    (64, 64, 10, 2, 160, 160, 0, 0),
    (70, 38, 8, 1, 5, 3, 2, 2),
    (1920, 1080, 10, 2, 160, 160, 0, 0),
]


@pytest.mark.parametrize("w,h,bd,sb,ox,oy,pr,pb", CASES)
def test_download_ref_padded_matches_oracle(H, orc, w, h, bd, sb, ox, oy, pr, pb):
    S = H.S
    g = S.make_geom(w, h, bd, sb)
    pic = H.synth_picture(g, 4200 + w + bd)
    split, two = sb == 2, sb == 2 and bd == 10
    exp = H.handoff_cpu(orc.lib, "svtf_orc_", g, pic, w + pr, h + pb, ox, oy, pr, pb, split, two)
    with S.FilterContext(g) as ctx:
        ctx.begin(g)
        ctx.upload(pic)
        got = ctx.download_ref_padded(w + pr, h + pb, ox, oy, pr, pb, split, two)
    for name, x, y in zip(("padded", "8-bit", "2-bit"), got, exp):
        assert (x is None) == (y is None)
        if x is not None:
            assert H.planes_equal(x, y), name + ": " + H.first_diff(x, y)


def test_handoff_after_chain_and_8as16_upload(H, orc):
    """8-bit planes in, widened on the device, deblock + CDEF in the 16-bit pipeline, padded + narrowed reference out."""
    S = H.S
    g16 = S.make_geom(328, 200, 8, 2)
    g8 = S.make_geom(328, 200, 8, 1)
    pic8 = H.synth_picture(g8, 4300)
    pic16 = H.widen_cpu(orc.lib, "svtf_orc_", pic8)
    mi = H.pack_mi(H.synth_mode_info(g16, 4301))
    lf = S.lf_params((24, 24), 12, 12)
    cdef, idx = H.synth_cdef(g16, 4302)
    e = orc.cdef_apply(g16, orc.deblock(g16, pic16, mi, lf), mi, cdef, idx)
    exp = H.handoff_cpu(orc.lib, "svtf_orc_", g16, e, 328, 200, 64, 48, 0, 0, True, False)
    with S.FilterContext(g16) as ctx:
        ctx.begin(g16)
        ctx.upload_8as16(pic8)
        ctx.set_mi(mi)
        ctx.loop_filter_frame(lf)
        ctx.cdef_frame(cdef, idx)
        got = ctx.download_ref_padded(328, 200, 64, 48, 0, 0, True, False)
    assert H.planes_equal(got[0], exp[0]), H.first_diff(got[0], exp[0])
    assert H.planes_equal(got[1], exp[1]), H.first_diff(got[1], exp[1])


def test_handoff_rejects_bad_arguments(H):
    S = H.S
    g = S.make_geom(64, 64, 8, 1)
    with S.FilterContext(g) as ctx:
        ctx.begin(g)
        ctx.upload(H.synth_picture(g, 1))
        with pytest.raises(S.SvtfError):
            ctx.download_ref_padded(64, 64, 8, 8, split=True)      # split needs 16-bit samples
        with pytest.raises(S.SvtfError):
            ctx.download_ref_padded(640, 64, 8, 8)                  # larger than the picture
        with pytest.raises(S.SvtfError):
            ctx.upload_8as16(H.synth_picture(g, 2))                 # context is not 8-in-16
