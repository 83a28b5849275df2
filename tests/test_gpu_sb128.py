"""GPU: 128x128 superblocks.  The reference searches a 128-wide / 128-high coding block as ONE filter block: its first 64x64 quadrant's
row of mse_seg carries the distortion of the whole block (the >> 2*coeff_shift comes after the sum over the block), the other
quadrants are neither searched nor picked, and the chosen index is copied to them (Encoder/Codec/EbCdefProcess.c:186-211,
Encoder/Codec/EbEncCdef.c:756-759, :822-826, :884-905).  The oracle restatement has no 128 blocks; the checker here is the reference
itself (oracle/_ref/libsvtref.so)."""
import ctypes as C
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _sig(p):
    return [p.cdef_bits, p.cdef_damping, *p.cdef_y_strength, *p.cdef_uv_strength]


@pytest.mark.parametrize("w,h,bd,sub", [(384, 256, 8, 1), (392, 328, 10, 1), (256, 264, 10, 2), (320, 192, 12, 1)])
def test_search_and_pick_with_128_blocks_match_reference(H, ref, w, h, bd, sub):
    S = H.S
    g = S.make_geom(w, h, bd)
    rec = H.synth_picture(g, 0x128 + w)
    src = H.synth_source(g, rec, 0x129 + w)
    soa = H.synth_mode_info(g, 0x12A + w, p_split=0.35, p_skip=0.45, sb_size=128)
    assert (soa["bsize"] >= 13).any(), "the generator produced no 128-wide block"
    first, second = H.cdef_level1_lists()
    cfg = S.cdef_search_cfg(first, second, subsampling_factor=sub, base_q_idx=140, lambda_=4000)
    with S.FilterContext(g) as ctx:
        ctx.begin(g)
        ctx.upload(rec)
        ctx.upload_source(src)
        ctx.set_mi(H.pack_mi(soa))
        got = ctx.cdef_seg_search(cfg)
    exp = ref.cdef_search(g, rec, src, soa, cfg)
    for a, b, name in zip(got, exp, ("mse", "skip", "dir", "var")):
        assert np.array_equal(a, b), name


@pytest.mark.parametrize("w,h,bd", [(384, 256, 8), (392, 328, 10)])
def test_adapter_search_pick_with_128_blocks_match_reference(H, ref, w, h, bd):
    """svtf_cdef_search_picture + finish_cdef_search (adapters) on the reference's structures: table, frame header strengths and the
    per-block indices written into mi_grid_base (copied to the folded quadrants) equal the reference's."""
    path = os.path.join(H.ROOT, "adapters", "_build", "libsvtadapt_test.so")
    if not os.path.exists(path):
        pytest.skip("adapters/_build/libsvtadapt_test.so not built")
    adapt = C.CDLL(path)
    S = H.S
    g = S.make_geom(w, h, bd, 2)
    rec = H.synth_picture(g, 0x138 + w)
    src = H.synth_source(g, rec, 0x139 + w)
    soa = H.synth_mode_info(g, 0x13A + w, p_split=0.35, p_skip=0.45, sb_size=128)
    s, keep = H.Ref._soa(soa)
    first, second = H.cdef_level1_lists()
    cfg = S.cdef_search_cfg(first, second, base_q_idx=130, zero_fs_cost_bias=0, lambda_=5000)
    nv, nh = H.nfb_of(g)
    nfb = nv * nh
    mse = np.zeros((2, nfb, 64), np.uint64)
    skip = np.zeros(nfb, np.uint8)
    dirs = np.zeros((nfb, 64), np.uint8)
    var = np.zeros((nfb, 64), np.int32)
    out = H.L.CdefParams()
    idx = np.zeros(nfb, np.int8)
    assert adapt.adapt_cdef_search_pick(C.byref(g), *H._planes(rec), *H._planes(src), C.byref(s), C.byref(cfg), H._p(mse), H._p(skip), H._p(dirs),
                                        H._p(var), C.byref(out), H._p(idx)) == 0
    e = ref.cdef_search(g, rec, src, soa, cfg)
    assert np.array_equal(mse, e[0]) and np.array_equal(skip, e[1])
    rp, ridx = ref.cdef_pick(g, soa, cfg, e[0], e[1])
    assert _sig(out) == _sig(rp)
    assert np.array_equal(idx, ridx)
