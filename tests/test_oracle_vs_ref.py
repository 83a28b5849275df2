"""CPU: pins the oracle restatement (oracle/svtf_oracle.c) against the reference's own C code
(oracle/_ref/libsvtref.so, compiled from /root/reference by oracle/Makefile).

Leaf-level sweeps follow the designs of the reference's unit tests (which compare C with SIMD):
test/DeblockTest.cc:95-309, test/CdefTest.cc:94-411 (block filter), :435-665 (find_dir),
:760-980 (dist).  Frame-level cases cover odd sizes, all bit depths, segmentation, delta-lf,
skipped planes, masked chroma strengths and row subsampling.  Skipped where the reference library
is absent (the golden vectors in tests/golden/ then carry the pin).
"""
import numpy as np
import pytest

LARGE = 0x7F7F


# ---- deblock edge kernels -------------------------------------------------------------------------
@pytest.mark.parametrize("length", [4, 6, 8, 14])
@pytest.mark.parametrize("bd,dtype", [(8, np.uint8), (10, np.uint16), (12, np.uint16), (8, np.uint16)])
def test_lpf_edge_kernels(H, orc, ref, length, bd, dtype):
    rng = np.random.default_rng(1000 * length + bd)
    if dtype == np.uint8 and bd != 8:
        pytest.skip("n/a")
    for it in range(400):
        # test/DeblockTest.cc:131-176: random 32x32 block, smooth-ish so the masks pass often
        base = int(rng.integers(0, 1 << bd))
        spread = int(rng.choice([1, 2, 4, 16, 64])) << (bd - 8)
        buf = np.clip(base + rng.integers(-spread, spread + 1, (32, 32)), 0, (1 << bd) - 1).astype(dtype)
        blimit, limit, thresh = int(rng.integers(0, 3 * 63 + 5)), int(rng.integers(0, 64)), int(rng.integers(0, 4))
        for vertical in (0, 1):
            a, b = buf.copy(), buf.copy()
            off = 16 * 32 + 16
            ref.lpf(a, off, 32, vertical, length, bd, blimit, limit, thresh)
            along, nxt = (1, 32) if vertical else (32, 1)
            orc.lpf_edge(b, off, along, nxt, 4, length, bd, blimit, limit, thresh)
            assert np.array_equal(a, b), (it, vertical, blimit, limit, thresh)


# ---- CDEF block filter -----------------------------------------------------------------------------
@pytest.mark.parametrize("bd", [8, 10, 12])
@pytest.mark.parametrize("bw,bh", [(8, 8), (4, 4), (8, 4), (4, 8)])
def test_cdef_filter_block(H, orc, ref, bd, bw, bh):
    rng = np.random.default_rng(bd * 100 + bw * 10 + bh)
    bs = ref.bstride()
    cs = bd - 8
    for boundary in range(16):  # test/CdefTest.cc:125-146: left/right/above/below = CDEF_VERY_LARGE
        base = int(rng.integers(0, 1 << bd))
        tile = np.clip(base + rng.integers(-(40 << cs), (40 << cs) + 1, (16, bs)), 0, (1 << bd) - 1).astype(np.uint16)
        y0, x0 = 4, 8
        if boundary & 1:
            tile[:, :x0] = LARGE
        if boundary & 2:
            tile[:, x0 + bw:] = LARGE
        if boundary & 4:
            tile[:y0, :] = LARGE
        if boundary & 8:
            tile[y0 + bh:, :] = LARGE
        for direction in range(8):
            for pri in (0, 1, 3, 7, 15, 19):
                for sec in (0, 1, 2, 4):
                    damping = int(rng.integers(3, 7)) + cs
                    sub = int(rng.choice([1, 2])) if bh == 8 else 1
                    a = ref.filter_block(tile, y0, x0, pri << cs, sec << cs, direction, damping, damping, bw, bh, cs, sub)
                    b = orc.filter_block(tile, y0, x0, pri << cs, sec << cs, direction, damping, damping, bw, bh, cs, sub)
                    assert np.array_equal(a[::sub], b[::sub]), (boundary, direction, pri, sec, damping, sub)


# ---- CDEF direction search -------------------------------------------------------------------------
@pytest.mark.parametrize("bd", [8, 10, 12])
def test_cdef_find_dir(H, orc, ref, bd):
    rng = np.random.default_rng(bd)
    for it in range(1500):
        # test/CdefTest.cc:471-480: every DC level x noise bits; plus oriented gradients
        dc = int(rng.integers(0, 1 << bd))
        nb = int(rng.integers(0, bd))
        img = dc + rng.integers(-(1 << nb), (1 << nb) + 1, (8, 8))
        if it % 3 == 0:
            ang = rng.random() * np.pi
            yy, xx = np.mgrid[0:8, 0:8]
            img = img + ((np.cos(ang) * xx + np.sin(ang) * yy) * (rng.integers(1, 40) << (bd - 8))).astype(np.int64)
        img = np.clip(img, 0, (1 << bd) - 1).astype(np.uint16)
        assert orc.find_dir(img, bd - 8) == ref.find_dir(img, bd - 8), it


# ---- CDEF distortion -------------------------------------------------------------------------------
@pytest.mark.parametrize("bd,dtype", [(8, np.uint8), (8, np.uint16), (10, np.uint16), (12, np.uint16)])
@pytest.mark.parametrize("pli,bw,bh", [(0, 8, 8), (1, 4, 4), (2, 4, 4)])
def test_cdef_dist(H, orc, ref, bd, dtype, pli, bw, bh):
    rng = np.random.default_rng(bd * 7 + pli)
    cs = bd - 8
    for it in range(60):
        n = int(rng.integers(1, 65))
        cells = rng.permutation(64)[:n]
        by, bx = (cells // 8).astype(np.uint8), (cells % 8).astype(np.uint8)
        src = rng.integers(0, 1 << bd, (8 * bh, 8 * bw)).astype(dtype)
        noise = int(rng.choice([0, 1, 4, 60])) << cs
        flt = np.zeros((n, bh, bw), np.uint16)
        for i in range(n):
            blk = src[by[i] * bh:(by[i] + 1) * bh, bx[i] * bw:(bx[i] + 1) * bw].astype(np.int64)
            flt[i] = np.clip(blk + rng.integers(-noise, noise + 1, (bh, bw)), 0, (1 << bd) - 1)
        for sub in ((1, 2, 4) if pli == 0 else (1,)):
            a = ref.cdef_dist(src, flt.astype(dtype), by, bx, bw, bh, cs, pli, sub)
            b = orc.cdef_dist(src, flt, by, bx, bw, bh, cs, pli, sub)
            assert a == b, (it, sub, n)


# ---- frame level ---------------------------------------------------------------------------------
FRAME_CASES = [(8, 8, 8, 1), (72, 40, 8, 1), (200, 136, 10, 2), (130, 70, 10, 2), (64, 64, 12, 2), (264, 200, 8, 2), (322, 182, 10, 2)]


@pytest.mark.parametrize("w,h,bd,sb", FRAME_CASES)
def test_deblock_frame(H, orc, ref, w, h, bd, sb):
    S = H.S
    g = S.make_geom(w, h, bd, sb)
    rng = np.random.default_rng(w * h + bd)
    for it in range(6):
        rec = H.synth_picture(g, 10 + it, noise=int(rng.integers(0, 12)))
        soa = H.synth_mode_info(g, 20 + it, p_split=float(rng.uniform(0.2, 0.8)), p_skip=float(rng.uniform(0, 0.9)), segments=True)
        en = rng.integers(0, 2, (8, 8))
        da = rng.integers(-40, 41, (8, 8))
        lf = S.lf_params((int(rng.integers(0, 64)), int(rng.integers(0, 64))), int(rng.integers(0, 64)), int(rng.integers(0, 64)),
                         sharpness_level=int(rng.integers(0, 8)), mode_ref_delta_enabled=int(rng.integers(0, 2)),
                         mode_deltas=(0, int(rng.integers(-4, 5))), segmentation_enabled=int(it % 2), seg_feature_enabled=en,
                         seg_feature_data=da, delta_lf_present=int(it == 3 or it == 4), delta_lf_multi=int(it == 4),
                         curr_delta_lf=tuple(int(v) for v in rng.integers(-20, 21, 4)))
        if it == 5:
            lf.filter_level[0] = lf.filter_level[1] = 0
        ps, pe = (0, 3) if it != 2 else (1, 3)
        a = ref.deblock(g, rec, soa, lf, ps, pe)
        b = orc.deblock(g, rec, H.pack_mi(soa), lf, ps, pe)
        assert H.planes_equal(a, b), f"it {it}: " + H.first_diff(a, b)


@pytest.mark.parametrize("w,h,bd,sb", FRAME_CASES)
def test_cdef_apply_frame(H, orc, ref, w, h, bd, sb):
    S = H.S
    g = S.make_geom(w, h, bd, sb)
    for it in range(4):
        rec = H.synth_picture(g, 30 + it, noise=3 * it)
        soa = H.synth_mode_info(g, 40 + it, p_skip=0.25 * it)
        cdef, idx = H.synth_cdef(g, 50 + it, damping=3 + it)
        a = ref.cdef_apply(g, rec, soa, cdef, idx)
        b = orc.cdef_apply(g, rec, H.pack_mi(soa), cdef, idx)
        assert H.planes_equal(a, b), f"it {it}: " + H.first_diff(a, b)


@pytest.mark.parametrize("w,h,bd,sb", FRAME_CASES)
def test_ccso_apply_frame(H, orc, ref, w, h, bd, sb):
    S = H.S
    g = S.make_geom(w, h, bd, sb)
    rng = np.random.default_rng(w + h)
    for it in range(8):
        rec = H.synth_picture(g, 60 + it)
        pre = H.synth_picture(g, 70 + it)
        ccso, flags = H.synth_ccso(g, 80 + it, tuple(rng.integers(0, 4, 3)), tuple(rng.integers(0, 6, 3)), tuple(rng.integers(0, 2, 3)),
                                   tuple(rng.integers(0, 8, 3)), tuple(rng.integers(0, 2, 3)), tuple(int(v) for v in rng.integers(0, 2, 3) | (it == 0)))
        fl = flags if it % 2 else None
        a = ref.ccso_apply(g, rec, pre[0], ccso, fl)
        b = orc.ccso_apply(g, rec, pre[0], ccso, fl)
        assert H.planes_equal(a, b), f"it {it}: " + H.first_diff(a, b)


@pytest.mark.parametrize("w,h,bd,sb,sub", [(72, 40, 8, 1, 1), (200, 136, 10, 2, 1), (130, 70, 10, 2, 2), (64, 64, 12, 2, 4), (136, 72, 8, 2, 2)])
def test_cdef_search_and_pick(H, orc, ref, w, h, bd, sb, sub):
    S = H.S
    g = S.make_geom(w, h, bd, sb)
    src = H.synth_picture(g, 90, noise=1)
    rec = H.synth_source(g, src, 91, amp=10)
    soa = H.synth_mode_info(g, 92, p_skip=0.55)
    first, second = H.cdef_level1_lists()
    fuv, suv = list(first), list(second)
    fuv[3] = -1
    suv[7] = suv[20] = -1
    cfg = S.cdef_search_cfg(first, second, fuv, suv, subsampling_factor=sub, base_q_idx=150, lambda_=7000)
    a = ref.cdef_search(g, rec, src, soa, cfg)
    b = orc.cdef_search(g, rec, src, H.pack_mi(soa), cfg)
    for x, y, name in zip(a, b, ("mse", "skip", "dir", "var")):
        assert np.array_equal(x, y), name
    for bias in (0, 45):
        cfg.zero_fs_cost_bias = bias
        pa, ia = ref.cdef_pick(g, soa, cfg, a[0], a[1])
        pb, ib = orc.cdef_pick(cfg, b[0], b[1])
        assert (pa.cdef_bits, pa.cdef_damping, list(pa.cdef_y_strength)[:8], list(pa.cdef_uv_strength)[:8]) == \
               (pb.cdef_bits, pb.cdef_damping, list(pb.cdef_y_strength)[:8], list(pb.cdef_uv_strength)[:8])
        assert np.array_equal(ia, ib)
    # a preset-8 style short list (3 + 3 strengths, chroma partly masked)
    cfg2 = S.cdef_search_cfg([0, 8, 20], [1, 10, 22], [0, -1, 20], [1, 10, -1], subsampling_factor=sub, base_q_idx=60, lambda_=900)
    a = ref.cdef_search(g, rec, src, soa, cfg2)
    b = orc.cdef_search(g, rec, src, H.pack_mi(soa), cfg2)
    for x, y, name in zip(a, b, ("mse", "skip", "dir", "var")):
        assert np.array_equal(x, y), "short list " + name
