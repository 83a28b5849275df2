"""GPU parity: the CUDA path (through the C ABI) against the oracle — and against the real reference
library when its prebuilt .so travelled with the repo — on seeded synthetic pictures."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

CASES = [  # (width, height, bit_depth, sample_bytes)
    (328, 200, 8, 1),
    (328, 200, 10, 2),
    (256, 192, 8, 2),   # 8-bit content in the 16-bit pipeline
    (200, 136, 12, 2),
    (64, 64, 10, 2),
    (8, 8, 8, 1),
    (1920, 1080, 8, 1),
    (1920, 1080, 10, 2),
]


def _lf_variants(S):
    return [
        S.lf_params((24, 24), 12, 12),
        S.lf_params((40, 10), 0, 33, sharpness_level=3),
        S.lf_params((63, 63), 63, 63, sharpness_level=7, mode_ref_delta_enabled=1, mode_deltas=(0, 3)),
        S.lf_params((0, 0), 20, 20),  # luma level 0 breaks the plane loop
        S.lf_params((18, 30), 9, 9, delta_lf_present=1, delta_lf_multi=1, curr_delta_lf=(4, -6, 10, -3)),
    ]


@pytest.mark.parametrize("w,h,bd,sb", CASES)
def test_deblock_matches_oracle(H, orc, w, h, bd, sb):
    S = H.S
    g = S.make_geom(w, h, bd, sb)
    rec = H.synth_picture(g, 100 + w + bd)
    soa = H.synth_mode_info(g, 200 + h, segments=True)
    mi = H.pack_mi(soa)
    variants = _lf_variants(S) if w < 1000 else _lf_variants(S)[:1]
    with S.FilterContext(g) as ctx:
        for i, lf in enumerate(variants):
            if i == 2:  # exercise segmentation features too
                en = np.zeros((8, 8), np.int16)
                da = np.zeros((8, 8), np.int16)
                en[:, 1:5] = 1
                da[:, 1:5] = np.arange(-16, 16).reshape(8, 4)
                lf = S.lf_params((50, 45), 30, 28, sharpness_level=1, mode_ref_delta_enabled=1, mode_deltas=(0, 3),
                                 segmentation_enabled=1, seg_feature_enabled=en, seg_feature_data=da)
            ctx.begin(g)
            ctx.upload(rec)
            ctx.set_mi(mi)
            ctx.loop_filter_frame(lf)
            got = ctx.download()
            exp = orc.deblock(g, rec, mi, lf)
            assert H.planes_equal(got, exp), f"variant {i}: " + H.first_diff(got, exp)


@pytest.mark.parametrize("w,h,bd,sb", CASES)
def test_cdef_matches_oracle(H, orc, w, h, bd, sb):
    S = H.S
    g = S.make_geom(w, h, bd, sb)
    rec = H.synth_picture(g, 300 + w + bd)
    soa = H.synth_mode_info(g, 400 + h, p_skip=0.5)
    mi = H.pack_mi(soa)
    with S.FilterContext(g) as ctx:
        for damping in (3, 4, 5, 6) if w < 1000 else (5,):
            cdef, idx = H.synth_cdef(g, 500 + damping, damping=damping)
            idx[::7] = -1
            ctx.begin(g)
            ctx.upload(rec)
            ctx.set_mi(mi)
            ctx.cdef_frame(cdef, idx)
            got = ctx.download()
            exp = orc.cdef_apply(g, rec, mi, cdef, idx)
            assert H.planes_equal(got, exp), f"damping {damping}: " + H.first_diff(got, exp)


@pytest.mark.parametrize("w,h,bd,sb", CASES)
def test_ccso_matches_oracle(H, orc, w, h, bd, sb):
    S = H.S
    g = S.make_geom(w, h, bd, sb)
    rec = H.synth_picture(g, 600 + w + bd)
    pre = H.synth_picture(g, 601 + w + bd)
    soa = H.synth_mode_info(g, 700 + h)
    mi = H.pack_mi(soa)
    combos = [((0, 0, 0), (0, 1, 2), (0, 0, 0), (2, 2, 2), (0, 0, 0)), ((1, 2, 3), (3, 4, 5), (1, 0, 1), (0, 3, 1), (0, 0, 0)),
              ((3, 1, 0), (5, 0, 2), (0, 1, 0), (4, 0, 7), (1, 0, 1))]
    with S.FilterContext(g) as ctx:
        for q, sup, clf, band, bo in combos if w < 1000 else combos[:1]:
            ccso, flags = H.synth_ccso(g, 800 + sum(q), q, sup, clf, band, bo)
            # (a) standalone: luma classifier = the picture itself (out of place on the GPU)
            ctx.begin(g)
            ctx.upload(rec)
            ctx.set_mi(mi)
            ctx.ccso_frame(ccso, flags)
            got = ctx.download()
            exp = orc.ccso_apply(g, rec, rec[0], ccso, flags)
            assert H.planes_equal(got, exp), "standalone: " + H.first_diff(got, exp)
            # (b) after CDEF: classifier = pre-CDEF luma
            cdef, idx = H.synth_cdef(g, 900)
            ctx.begin(g)
            ctx.upload(pre)
            ctx.set_mi(mi)
            ctx.cdef_frame(cdef, idx)
            ctx.ccso_frame(ccso, None)
            got = ctx.download()
            e1 = orc.cdef_apply(g, pre, mi, cdef, idx)
            exp = orc.ccso_apply(g, e1, pre[0], ccso, None)
            assert H.planes_equal(got, exp), "after cdef: " + H.first_diff(got, exp)


@pytest.mark.parametrize("w,h,bd,sb", CASES)
def test_fused_cdef_ccso_matches_oracle(H, orc, w, h, bd, sb):
    """filter_chain(cdef + ccso) is ONE launch (CCSO in the CDEF kernel's epilogue); it must equal the two
    separate passes of the oracle for every tap support / classifier / band / bo_only / block-flag combination,
    including filter blocks CDEF skips (strength index -1, all-skip blocks) where only CCSO touches the samples."""
    S = H.S
    g = S.make_geom(w, h, bd, sb)
    pre = H.synth_picture(g, 650 + w + bd)
    soa = H.synth_mode_info(g, 750 + h)
    mi = H.pack_mi(soa)
    combos = [((0, 0, 0), (0, 1, 2), (0, 0, 0), (2, 2, 2), (0, 0, 0), (1, 1, 1)), ((1, 2, 3), (3, 4, 5), (1, 0, 1), (0, 3, 1), (0, 0, 0), (1, 1, 1)),
              ((3, 1, 0), (5, 0, 2), (0, 1, 0), (4, 0, 7), (1, 0, 1), (1, 1, 1)), ((2, 2, 2), (4, 5, 3), (0, 0, 1), (7, 6, 5), (0, 1, 0), (1, 0, 1)),
              ((0, 3, 1), (1, 2, 0), (1, 1, 0), (1, 1, 1), (0, 0, 0), (0, 1, 0))]
    with S.FilterContext(g) as ctx:
        for n, (q, sup, clf, band, bo, en) in enumerate(combos if w < 1000 else combos[:2]):
            ccso, flags = H.synth_ccso(g, 850 + n, q, sup, clf, band, bo, en, p_on=0.7)
            cdef, idx = H.synth_cdef(g, 950 + n)
            if n == 1:
                idx = idx.copy()
                idx.reshape(-1)[::2] = -1  # half of the filter blocks: CDEF off, CCSO still on
            for fl in (flags, None):
                ctx.begin(g)
                ctx.upload(pre)
                ctx.set_mi(mi)
                before = ctx.launch_count
                ctx.filter_chain(None, cdef, idx, ccso, fl)
                assert ctx.launch_count - before == 1
                got = ctx.download()
                e1 = orc.cdef_apply(g, pre, mi, cdef, idx)
                exp = orc.ccso_apply(g, e1, pre[0], ccso, fl)
                assert H.planes_equal(got, exp), f"combo {n} flags={'y' if fl else 'n'}: " + H.first_diff(got, exp)


@pytest.mark.parametrize("w,h,bd,sb", [(328, 200, 8, 1), (328, 200, 10, 2), (1920, 1080, 10, 2)])
def test_chain_matches_oracle_and_reference(H, orc, w, h, bd, sb):
    S = H.S
    g = S.make_geom(w, h, bd, sb)
    rec = H.synth_picture(g, 1000 + w + bd)
    soa = H.synth_mode_info(g, 1100 + h)
    mi = H.pack_mi(soa)
    lf = S.lf_params((24, 24), 12, 12)
    cdef, idx = H.synth_cdef(g, 1200)
    ccso, flags = H.synth_ccso(g, 1300)
    with S.FilterContext(g) as ctx:
        ctx.begin(g)
        ctx.upload(rec)
        ctx.set_mi(mi)
        ctx.filter_chain(lf, cdef, idx, ccso, flags)
        got = ctx.download()
        # idempotence of the device-resident re-run (rewind does not copy)
        ctx.rewind()
        ctx.filter_chain(lf, cdef, idx, ccso, flags)
        again = ctx.download()
    assert H.planes_equal(got, again)
    e1 = orc.deblock(g, rec, mi, lf)
    e2 = orc.cdef_apply(g, e1, mi, cdef, idx)
    e3 = orc.ccso_apply(g, e2, e1[0], ccso, flags)
    assert H.planes_equal(got, e3), H.first_diff(got, e3)
    if H.ref_available():
        ref = H.Ref()
        r1 = ref.deblock(g, rec, soa, lf)
        r2 = ref.cdef_apply(g, r1, soa, cdef, idx)
        r3 = ref.ccso_apply(g, r2, r1[0], ccso, flags)
        assert H.planes_equal(got, r3), "vs reference: " + H.first_diff(got, r3)


@pytest.mark.parametrize("w,h,bd,sb,sub", [(328, 200, 8, 1, 1), (328, 200, 10, 2, 1), (200, 136, 10, 2, 2), (200, 136, 8, 1, 4),
                                           (200, 136, 12, 2, 1)])
def test_cdef_search_matches_oracle(H, orc, w, h, bd, sb, sub):
    S = H.S
    g = S.make_geom(w, h, bd, sb)
    src = H.synth_picture(g, 1400 + w + bd, noise=0)
    rec = H.synth_source(g, src, 1500, amp=10)
    soa = H.synth_mode_info(g, 1600 + h, p_skip=0.6)
    mi = H.pack_mi(soa)
    first, second = H.cdef_level1_lists()
    suv = list(second)
    suv[5] = -1
    suv[17] = -1
    cfg = S.cdef_search_cfg(first, second, first, suv, subsampling_factor=sub, base_q_idx=120, lambda_=4000)
    with S.FilterContext(g) as ctx:
        ctx.begin(g)
        ctx.upload(rec)
        ctx.upload_source(src)
        ctx.set_mi(mi)
        mse, skip, dirs, var = ctx.cdef_seg_search(cfg)
        params, idx = ctx.finish_cdef_search(cfg)
    e_mse, e_skip, e_dirs, e_var = orc.cdef_search(g, rec, src, mi, cfg)
    assert np.array_equal(skip, e_skip)
    assert np.array_equal(dirs, e_dirs)
    assert np.array_equal(var, e_var)
    bad = np.argwhere(mse != e_mse)
    assert len(bad) == 0, f"{len(bad)} mse entries differ, first {bad[0]}: {mse[tuple(bad[0])]} vs {e_mse[tuple(bad[0])]}"
    e_params, e_idx = orc.cdef_pick(cfg, e_mse, e_skip)
    assert params.cdef_bits == e_params.cdef_bits and params.cdef_damping == e_params.cdef_damping
    assert list(params.cdef_y_strength) == list(e_params.cdef_y_strength)
    assert list(params.cdef_uv_strength) == list(e_params.cdef_uv_strength)
    assert np.array_equal(idx, e_idx)


def _params_sig(p):
    return [p.cdef_bits, p.cdef_damping, *p.cdef_y_strength, *p.cdef_uv_strength]


def test_device_pick_matches_host_and_oracle(H, orc):
    """finish_cdef_search on the GPU (svtf_cdef_pick_table) against the single-thread host implementation and the
    oracle: random tables with ties (first index must win), zero-strength bias, skipped blocks, all blocks skipped,
    short strength lists."""
    S = H.S
    rng = np.random.default_rng(7)
    first, second = H.cdef_level1_lists()
    g = S.make_geom(64, 64, 8)
    with S.FilterContext(g) as ctx:
        for it, (nfb, n1, n2, bias, pskip) in enumerate(((1, 16, 48, 0, 0.0), (7, 3, 3, 0, 0.2), (60, 16, 48, 40, 0.2), (130, 8, 8, 0, 0.3),
                                                        (33, 1, 0, 0, 0.0), (2040, 16, 48, 0, 0.1), (500, 16, 20, 70, 1.0), (64, 2, 1, 0, 0.5))):
            cfg = S.cdef_search_cfg(first[:n1], second[:n2], base_q_idx=int(rng.integers(0, 256)), zero_fs_cost_bias=bias,
                                    lambda_=int(rng.integers(1, 50000)))
            mse = rng.integers(0, 1 << 22, (2, nfb, 64)).astype(np.uint64)
            mse[:, :, 0] += 1000
            if it == 5:
                mse = (mse >> 12) << 12  # coarse values: many exact ties between (j,k) pairs
            skip = (rng.random(nfb) < pskip).astype(np.uint8)
            if nfb > 3:
                mse[:, 1] = mse[:, 0]
            a = ctx.pick_table(cfg, mse, skip)
            b = S.cdef_pick_host(cfg, mse, skip)
            c = orc.cdef_pick(cfg, mse, skip)
            assert _params_sig(a[0]) == _params_sig(b[0]) == _params_sig(c[0]), it
            assert np.array_equal(a[1], b[1]) and np.array_equal(a[1], c[1]), it


def test_device_pick_after_search(H, orc):
    S = H.S
    g = S.make_geom(328, 200, 10)
    rec = H.synth_picture(g, 77)
    src = H.synth_source(g, rec, 78)
    mi = H.pack_mi(H.synth_mode_info(g, 79))
    first, second = H.cdef_level1_lists()
    cfg = S.cdef_search_cfg(first, second, base_q_idx=120, lambda_=4000)
    with S.FilterContext(g) as ctx:
        ctx.begin(g)
        ctx.upload(rec)
        ctx.upload_source(src)
        ctx.set_mi(mi)
        mse, skip, _, _ = ctx.cdef_seg_search(cfg)
        a = ctx.finish_cdef_search(cfg)
    b = orc.cdef_pick(cfg, mse, skip)
    assert _params_sig(a[0]) == _params_sig(b[0])
    assert np.array_equal(a[1], b[1])


def _extreme_pictures(g, seed):
    """Full-range white noise, a 0/max checkerboard with random outliers, a steep wrapping ramp."""
    rng = np.random.default_rng(seed)
    mx = (1 << g.bit_depth) - 1
    dt = np.uint8 if g.sample_bytes == 1 else np.uint16
    pics = []
    for kind in range(3):
        pl = []
        for p in range(3):
            s = 1 if p else 0
            w, h = (g.mi_cols * 4) >> s, (g.mi_rows * 4) >> s
            yy, xx = np.mgrid[0:h, 0:w]
            if kind == 0:
                v = rng.integers(0, mx + 1, size=(h, w))
            elif kind == 1:
                v = (((xx // 3 + yy // 2) & 1) * mx) ^ (rng.integers(0, 4, size=(h, w)) == 0) * rng.integers(0, mx + 1, size=(h, w))
            else:
                v = np.clip((xx * 37 + yy * 91) % (2 * mx + 2) - (mx // 2) + rng.integers(-3, 4, size=(h, w)), 0, mx)
            pl.append(np.ascontiguousarray(v.astype(dt)))
        pics.append(pl)
    return pics


@pytest.mark.parametrize("bd,sb", [(8, 1), (10, 2), (8, 2)])
def test_cdef_fp16_taps_exact_on_extreme_content(H, orc, bd, sb):
    """cdef3_kernel runs the taps of 8/10-bit pictures in packed fp16 (svtf_cdef3.cuh).  Full-range white noise, 0/max
    checkerboards and steep ramps make every |p - x| / strength / shift combination occur; every one of the 64 luma and
    chroma strengths and dampings 3..9 (9 and, at 10 bits, 8 fall back to the packed-integer taps) must stay bit-exact."""
    S = H.S
    g = S.make_geom(512, 256, bd, sb)
    pics = _extreme_pictures(g, 0xF16 + bd)
    soa = H.synth_mode_info(g, 0xF17, p_skip=0.2)
    mi = H.pack_mi(soa)
    nv, nh = H.nfb_of(g)
    with S.FilterContext(g) as ctx:
        for damping in (3, 4, 5, 6, 7, 8, 9):
            for rot in range(8):
                ys = [(8 * rot + k) % 64 for k in range(8)]
                us = [(8 * rot + 5 * k + 3) % 64 for k in range(8)]
                cdef = S.cdef_params(damping, ys, us)
                idx = ((np.arange(nv * nh) + rot) % 8).astype(np.int8)
                rec = pics[(rot + damping) % 3]
                ctx.begin(g)
                ctx.upload(rec)
                ctx.set_mi(mi)
                ctx.cdef_frame(cdef, idx)
                got = ctx.download()
                exp = orc.cdef_apply(g, rec, mi, cdef, idx)
                assert H.planes_equal(got, exp), f"damping {damping} strengths {ys} {us}: " + H.first_diff(got, exp)


@pytest.mark.parametrize("bd,sb,qidx", [(8, 1, 10), (10, 2, 120), (10, 2, 255), (8, 2, 200)])
def test_cdef_search_fp16_exact_on_extreme_content(H, orc, bd, sb, qidx):
    """cdef_search3_kernel (fp16 taps, rounding and clamp; integer moments) on white noise / checkerboard / ramp
    reconstructions, all 64 strengths, search damping 3..6 (= 3 + base_q_idx / 64): the u64 mse table is bit-exact."""
    S = H.S
    g = S.make_geom(256, 128, bd, sb)
    pics = _extreme_pictures(g, 0x5EA + bd + qidx)
    mi = H.pack_mi(H.synth_mode_info(g, 0x5EB, p_skip=0.2))
    first, second = H.cdef_level1_lists()
    cfg = S.cdef_search_cfg(first, second, base_q_idx=qidx, lambda_=4000)
    with S.FilterContext(g) as ctx:
        for k in range(3):
            rec, src = pics[k], pics[(k + 1) % 3]
            ctx.begin(g)
            ctx.upload(rec)
            ctx.upload_source(src)
            ctx.set_mi(mi)
            mse, skip, dirs, var = ctx.cdef_seg_search(cfg)
            e_mse, e_skip, e_dirs, e_var = orc.cdef_search(g, rec, src, mi, cfg)
            assert np.array_equal(skip, e_skip) and np.array_equal(dirs, e_dirs) and np.array_equal(var, e_var)
            bad = np.argwhere(mse != e_mse)
            assert len(bad) == 0, f"picture {k}: {len(bad)} mse entries differ, first {bad[0]}: {mse[tuple(bad[0])]} vs {e_mse[tuple(bad[0])]}"
