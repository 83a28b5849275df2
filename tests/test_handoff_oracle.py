"""Recon hand-off: the oracle restatement against the reference's own pad / unpack / convert functions."""
import numpy as np
import pytest

CASES = [  # (width, height, bit_depth, sample_bytes, org_x, org_y, pad_right, pad_bottom)
    (328, 200, 8, 1, 32, 16, 0, 0),
    (328, 200, 10, 2, 96, 96, 0, 0),
    (322, 198, 10, 2, 40, 24, 6, 2),   # picture padded to a multiple of 8 by replicating its last column / row
    (200, 136, 8, 2, 8, 4, 0, 0),      # 8-bit content in the 16-bit pipeline
    (64, 64, 10, 2, 160, 160, 0, 0),
    (70, 38, 8, 1, 5, 3, 2, 2),
]


@pytest.mark.parametrize("w,h,bd,sb,ox,oy,pr,pb", CASES)
def test_handoff_oracle_matches_reference(H, orc, w, h, bd, sb, ox, oy, pr, pb):
    if not H.ref_available():
        pytest.skip("reference library not built")
    ref = H.Ref()
    S = H.S
    g = S.make_geom(w, h, bd, sb)
    pic = H.synth_picture(g, 4000 + w + bd)
    split, two = sb == 2, sb == 2 and bd == 10
    a = H.handoff_cpu(orc.lib, "svtf_orc_", g, pic, w + pr, h + pb, ox, oy, pr, pb, split, two)
    b = H.handoff_cpu(ref.lib, "ref_", g, pic, w + pr, h + pb, ox, oy, pr, pb, split, two)
    for x, y in zip(a, b):
        assert (x is None) == (y is None)
        if x is not None:
            assert H.planes_equal(x, y), H.first_diff(x, y)
    # every output sample is the clamped source sample (the closed form the CUDA kernel uses)
    for p, (buf, s) in enumerate(zip(a[0], (0, 1, 1))):
        vw, vh = (w + s) >> s, (h + s) >> s
        oxp, oyp = (ox + s) >> s, (oy + s) >> s
        yy = np.clip(np.arange(buf.shape[0]) - oyp, 0, vh - 1)
        xx = np.clip(np.arange(buf.shape[1]) - oxp, 0, vw - 1)
        assert np.array_equal(buf, pic[p][:vh, :vw][np.ix_(yy, xx)])


def test_widen_oracle_matches_reference(H, orc):
    if not H.ref_available():
        pytest.skip("reference library not built")
    ref = H.Ref()
    g = H.S.make_geom(200, 136, 8, 1)
    pic = H.synth_picture(g, 4100)
    a, b = H.widen_cpu(orc.lib, "svtf_orc_", pic), H.widen_cpu(ref.lib, "ref_", pic)
    assert H.planes_equal(a, b) and all(np.array_equal(x, p) for x, p in zip(a, pic))
