"""Super-resolution upscale: the oracle's restatement (svtf_orc_superres_plane) against the reference's own
svt_av1_upscale_normative_rows (Common/Codec/EbSuperRes.c compiled into oracle/_ref)."""
import ctypes as C

import numpy as np
import pytest


def ref_filters(ref):
    ref.lib.ref_superres_filters.restype = C.POINTER(C.c_int16)
    return np.ctypeslib.as_array(ref.lib.ref_superres_filters(), shape=(64, 8)).copy()


def upscale_cpu(H, fn, g, plane_arr, plane, up_w, denom, starts, filters=None):
    ss = 1 if plane else 0
    src = np.ascontiguousarray(plane_arr)
    rows, src_cw = src.shape
    ds_w, upw = (g.width + ss) >> ss, (up_w + ss) >> ss
    dst = np.zeros((rows, upw), src.dtype)
    st = (C.c_int32 * len(starts))(*starts)
    args = [H._p(src), src.strides[0] // src.itemsize, H._p(dst), upw, g.sample_bytes, g.bit_depth, rows, src_cw]
    if filters is None:  # the reference entry takes frame widths and derives the plane widths itself
        assert fn(*args, g.width, up_w, denom, ss, g.mi_cols, len(starts) - 1, st) == 0
    else:
        f = np.ascontiguousarray(filters, np.int16)
        fn.restype = None
        fn(*args, ds_w, upw, denom, ss, g.mi_cols, len(starts) - 1, st, H._p(f))
    return dst


CASES = [(136, 72, 8, 1, 16, None), (200, 64, 10, 2, 9, None), (328, 40, 10, 2, 13, (0, 32, 64, 82)), (134, 70, 12, 2, 11, None),
         (264, 48, 8, 2, 15, (0, 16, 66)), (1920 // 2, 32, 10, 2, 16, None)]


@pytest.mark.parametrize("w,h,bd,sb,denom,starts", CASES)
def test_oracle_superres_matches_reference(H, orc, ref, w, h, bd, sb, denom, starts):
    S = H.S
    g = S.make_geom(w, h, bd, sb)
    pic = H.synth_picture(g, 15000 + w + denom, noise=12)
    up_w = (w * denom + 4) // 8  # a width whose downscale by `denom` is w (calculate_scaled_size_helper rounds to nearest)
    starts = list(starts) if starts else [0, g.mi_cols]
    f = ref_filters(ref)
    for p in range(3):
        a = upscale_cpu(H, orc.lib.svtf_orc_superres_plane, g, pic[p], p, up_w, denom, starts, f)
        b = upscale_cpu(H, ref.lib.ref_superres_plane, g, pic[p], p, up_w, denom, starts)
        assert np.array_equal(a, b), (p, np.argwhere(a != b)[:4])
        assert a.max() > 0


def test_filter_fixture_is_the_reference_table(H, ref):
    """tests/golden/superres_filters.npy (oracle/make_superres_table.py) = av1_resize_filter_normative of the reference."""
    import os

    assert np.array_equal(np.load(os.path.join(H.ROOT, "tests", "golden", "superres_filters.npy")), ref_filters(ref))
