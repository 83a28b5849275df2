"""CPU: the reference's SIMD build (oracle/_ref/libsvtref_simd.so: the SSE2 loop-filter edges and the AVX2 CDEF / CDEF
distortion / strength pick / CCSO kernels the RTCD tables select on an AVX2 host, Common/Codec/common_dsp_rtcd.c:626-658,
Encoder/Codec/aom_dsp_rtcd.c:187,188,478) produces the same bytes as the C-only build (libsvtref.so) the parity tests use.
That library is the TIMED CPU baseline of bench.py; this file is what lets its numbers stand beside bit-exact parity."""
import numpy as np
import pytest


@pytest.fixture(scope="module")
def simd(H):
    if not H.ref_available() or not H.ref_simd_available():
        pytest.skip("libsvtref_simd.so not built or no AVX2 on this host")
    return H.Ref(simd=True)


@pytest.mark.parametrize("w,h,bd,sb", [(136, 72, 8, 1), (200, 136, 10, 2), (328, 200, 10, 2), (72, 264, 8, 2), (64, 64, 12, 2)])
def test_frame_chain_simd_equals_c(H, ref, simd, w, h, bd, sb):
    S = H.S
    g = S.make_geom(w, h, bd, sb)
    rec = H.synth_picture(g, 0x51D + w)
    soa = H.synth_mode_info(g, 0x51E + w, segments=True)
    lf = S.lf_params((22, 30), 14, 9, sharpness_level=2)
    cdef, idx = H.synth_cdef(g, 0x51F + w, n_strengths=8, damping=3 + (w % 4))
    ccso, flags = H.synth_ccso(g, 0x520 + w, quant_idx=(0, 1, 2), support=(0, 3, 5), edge_clf=(0, 1, 0), band_log2=(2, 0, 1))
    a = [ref.deblock(g, rec, soa, lf)]
    b = [simd.deblock(g, rec, soa, lf)]
    a.append(ref.cdef_apply(g, a[0], soa, cdef, idx))
    b.append(simd.cdef_apply(g, b[0], soa, cdef, idx))
    a.append(ref.ccso_apply(g, a[1], a[0][0], ccso, flags))
    b.append(simd.ccso_apply(g, b[1], b[0][0], ccso, flags))
    for stage, (x, y) in enumerate(zip(a, b)):
        assert H.planes_equal(x, y), (stage, H.first_diff(x, y))


@pytest.mark.parametrize("w,h,bd,sub", [(136, 72, 8, 1), (200, 136, 10, 1), (136, 136, 10, 2), (128, 64, 8, 4)])
def test_search_and_pick_simd_equals_c(H, ref, simd, w, h, bd, sub):
    S = H.S
    g = S.make_geom(w, h, bd)
    rec = H.synth_picture(g, 0x5A0 + w)
    src = H.synth_source(g, rec, 0x5A1 + w)
    soa = H.synth_mode_info(g, 0x5A2 + w)
    f1, f2 = H.cdef_level1_lists()
    cfg = S.cdef_search_cfg(f1, f2, subsampling_factor=sub, base_q_idx=150, lambda_=5000)
    ra = ref.cdef_search(g, rec, src, soa, cfg)
    rb = simd.cdef_search(g, rec, src, soa, cfg)
    for x, y in zip(ra, rb):
        assert np.array_equal(x, y)
    pa = ref.cdef_pick(g, soa, cfg, ra[0], ra[1])
    pb = simd.cdef_pick(g, soa, cfg, rb[0], rb[1])
    sig = lambda p: [p.cdef_bits, p.cdef_damping, *p.cdef_y_strength, *p.cdef_uv_strength]
    assert sig(pa[0]) == sig(pb[0]) and np.array_equal(pa[1], pb[1])


@pytest.mark.parametrize("length", [4, 6, 8, 14])
@pytest.mark.parametrize("bd,dtype", [(8, np.uint8), (10, np.uint16)])
def test_lpf_edge_kernels_simd_equals_c(H, ref, simd, length, bd, dtype):
    rng = np.random.default_rng(77 * length + bd)
    for it in range(200):
        base = int(rng.integers(0, 1 << bd))
        spread = int(rng.choice([1, 2, 4, 16, 64])) << (bd - 8)
        buf = np.clip(base + rng.integers(-spread, spread + 1, (32, 32)), 0, (1 << bd) - 1).astype(dtype)
        blimit, limit, thresh = int(rng.integers(0, 3 * 63 + 5)), int(rng.integers(0, 64)), int(rng.integers(0, 4))
        for vertical in (0, 1):
            a, b = buf.copy(), buf.copy()
            ref.lpf(a, 16 * 32 + 16, 32, vertical, length, bd, blimit, limit, thresh)
            simd.lpf(b, 16 * 32 + 16, 32, vertical, length, bd, blimit, limit, thresh)
            # the SSE2 kernels filter a fixed number of lines (4 for length <= 8 ... ) exactly like the C ones they replace
            assert np.array_equal(a, b), (it, vertical, blimit, limit, thresh)
