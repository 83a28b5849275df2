"""GPU: the Tier-2 adapters — the reference's own entry points (svt_av1_loop_filter_frame(frame, pcs, ..),
svt_av1_cdef_frame(scs, pcs), finish_cdef_search(pcs), ccso_frame(frame, pcs, pd, ext_rec_y)) implemented over
the C ABI — called on real PictureControlSet structures and compared with the real reference functions called
on identical structures (oracle/_ref) and with the oracle."""
import ctypes as C
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def adapt(H):
    path = os.path.join(H.ROOT, "adapters", "_build", "libsvtadapt_test.so")
    if not os.path.exists(path):
        pytest.skip("adapters/_build/libsvtadapt_test.so not built (needs the reference headers)")
    return C.CDLL(path)


def _soa(H, soa):
    return H.Ref._soa(soa)


@pytest.mark.parametrize("w,h,bd,sb", [(200, 136, 8, 1), (328, 200, 10, 2), (136, 72, 12, 2)])
def test_adapters_match_reference(H, orc, adapt, w, h, bd, sb):
    S = H.S
    g = S.make_geom(w, h, bd, sb)
    src = H.synth_picture(g, 2000 + w, noise=1)
    rec = H.synth_source(g, src, 2001, amp=9)
    soa = H.synth_mode_info(g, 2002, p_skip=0.4, segments=True)
    mi = H.pack_mi(soa)
    s, keep = _soa(H, soa)
    ref = H.Ref() if H.ref_available() else None

    # --- svt_av1_loop_filter_frame ---
    lf = S.lf_params((30, 22), 14, 17, sharpness_level=2, mode_ref_delta_enabled=1, mode_deltas=(0, 2))
    got = [p.copy() for p in rec]
    assert adapt.adapt_deblock_frame(C.byref(g), *H._planes(got), C.byref(s), C.byref(lf), 0, 3) == 0
    exp = orc.deblock(g, rec, mi, lf)
    assert H.planes_equal(got, exp), "deblock vs oracle: " + H.first_diff(got, exp)
    if ref:
        assert H.planes_equal(got, ref.deblock(g, rec, soa, lf))
    dbk = got

    # --- cdef search + finish_cdef_search ---
    first, second = H.cdef_level1_lists()
    cfg = S.cdef_search_cfg(first, second, subsampling_factor=1, base_q_idx=130, zero_fs_cost_bias=0, lambda_=5000)
    nv, nh = H.nfb_of(g)
    nfb = nv * nh
    mse = np.zeros((2, nfb, 64), np.uint64)
    skip = np.zeros(nfb, np.uint8)
    dirs = np.zeros((nfb, 64), np.uint8)
    var = np.zeros((nfb, 64), np.int32)
    out = H.L.CdefParams()
    idx = np.zeros(nfb, np.int8)
    assert adapt.adapt_cdef_search_pick(C.byref(g), *H._planes(dbk), *H._planes(src), C.byref(s), C.byref(cfg), H._p(mse), H._p(skip),
                                        H._p(dirs), H._p(var), C.byref(out), H._p(idx)) == 0
    e = orc.cdef_search(g, dbk, src, mi, cfg)
    for a, b, name in zip((mse, skip, dirs, var), e, ("mse", "skip", "dir", "var")):
        assert np.array_equal(a, b), name
    ep, eidx = orc.cdef_pick(cfg, e[0], e[1])
    assert (out.cdef_bits, out.cdef_damping, list(out.cdef_y_strength)[:8], list(out.cdef_uv_strength)[:8]) == \
           (ep.cdef_bits, ep.cdef_damping, list(ep.cdef_y_strength)[:8], list(ep.cdef_uv_strength)[:8])
    assert np.array_equal(idx, eidx)
    if ref:
        rp, ridx = ref.cdef_pick(g, soa, cfg, e[0], e[1])
        assert (out.cdef_bits, list(out.cdef_y_strength)[:8], list(out.cdef_uv_strength)[:8]) == \
               (rp.cdef_bits, list(rp.cdef_y_strength)[:8], list(rp.cdef_uv_strength)[:8])
        assert np.array_equal(idx, ridx)

    # --- svt_av1_cdef_frame with the picked strengths (and with random ones) ---
    for cdef, fidx in ((out, idx), H.synth_cdef(g, 2003)):
        got = [p.copy() for p in dbk]
        fi = np.ascontiguousarray(fidx, np.int8)
        assert adapt.adapt_cdef_apply_frame(C.byref(g), *H._planes(got), C.byref(s), C.byref(cdef), H._p(fi)) == 0
        exp = orc.cdef_apply(g, dbk, mi, cdef, fi)
        assert H.planes_equal(got, exp), "cdef vs oracle: " + H.first_diff(got, exp)
        if ref:
            assert H.planes_equal(got, ref.cdef_apply(g, dbk, soa, cdef, fi))
    post = got

    # --- ccso_frame (the reference wrapper is byte-only; the adapter handles 8- and 16-bit samples) ---
    ccso, flags = H.synth_ccso(g, 2004, (0, 1, 2), (0, 3, 5), (0, 1, 0), (2, 0, 3))
    got = [p.copy() for p in post]
    fl = (C.c_void_p * 3)(*[f.ctypes.data for f in flags])
    assert adapt.adapt_ccso_frame(C.byref(g), *H._planes(got), H._p(dbk[0]), C.c_int(dbk[0].strides[0] // dbk[0].itemsize), C.byref(ccso), fl) == 0
    exp = orc.ccso_apply(g, post, dbk[0], ccso, flags)
    assert H.planes_equal(got, exp), "ccso vs oracle: " + H.first_diff(got, exp)
    if ref:
        assert H.planes_equal(got, ref.ccso_apply(g, post, dbk[0], ccso, flags))


    # --- ccso_search(pcs, pd, rdmult, ext_rec_y, rec_uv, org_uv): uint16 planes at luma stride ---
    W, Hh = g.width, g.height
    luma16 = np.ascontiguousarray(dbk[0][:Hh, :W], dtype=np.uint16)
    rec16, org16 = [], []
    for p in range(3):
        ph, pw = Hh >> (1 if p else 0), W >> (1 if p else 0)
        r = np.zeros((Hh, W), np.uint16)
        o = np.zeros((Hh, W), np.uint16)
        r[:ph, :pw], o[:ph, :pw] = post[p][:ph, :pw], src[p][:ph, :pw]
        rec16.append(r)
        org16.append(o)
    base_q, rdmult = 37, 9
    dout = H.L.CcsoParams()
    dflags = [np.zeros(H.ccso_nblocks(g, p), np.uint8) for p in range(3)]
    vp3 = lambda arrs: (C.c_void_p * 3)(*[a.ctypes.data for a in arrs])  # noqa: E731
    assert adapt.adapt_ccso_search(C.byref(g), base_q, rdmult, H._p(luma16), vp3(rec16), vp3(org16), C.byref(dout), vp3(dflags)) == 0
    for p in range(3):
        e, ef = orc.ccso_derive_plane(g, p, dbk[0], post[p], src[p], rdmult * base_q)
        assert H.ccso_plane_sig(dout, p) == H.ccso_plane_sig(e, p), f"ccso_search plane {p}"
        assert np.array_equal(dflags[p], ef)
        if ref:
            r, rf = ref.ccso_derive_plane(g, p, dbk[0], post[p], src[p], rdmult * base_q)
            assert H.ccso_plane_sig(dout, p) == H.ccso_plane_sig(r, p) and np.array_equal(dflags[p], rf)


@pytest.mark.parametrize("w,h,bd,sb,start,kw,resident", [
    (200, 136, 8, 1, (12, 12, 8, 8), {}, 0),
    (264, 200, 10, 2, (30, 30, 16, 20), {"early_exit_convergence": 2}, 1),
    (136, 104, 10, 2, (8, 8, 40, 12), {"search_chroma": 0, "dlf_avg": 1}, 1),
])
def test_pick_filter_level_adapter_then_deblock(H, orc, adapt, w, h, bd, sb, start, kw, resident):
    """svt_av1_pick_filter_level(enhanced_pic, pcs, LPF_PICK_FROM_FULL_IMAGE) followed by svt_av1_loop_filter_frame(recon,
    pcs, 0, 3), as the DLF kernel calls them (Encoder/Codec/EbDlfProcess.c:101-106)."""
    from test_lf_search_oracle import blocky_case

    S = H.S
    g, rec, src, soa = blocky_case(H, w, h, bd, sb, 7300 + w, 6)
    mi = H.pack_mi(soa)
    s, keep = _soa(H, soa)
    lf = S.lf_params((start[0], start[1]), start[2], start[3], sharpness_level=3)
    cfg = H.L.LfSearchCfg(kw.get("dlf_avg", 0), kw.get("early_exit_convergence", 0), kw.get("only_4x4", 0), kw.get("search_chroma", 1))
    got = [p.copy() for p in rec]
    lv = (C.c_int32 * 4)()
    assert adapt.adapt_pick_and_deblock(C.byref(g), *H._planes(got), *H._planes(src), C.byref(s), C.byref(lf), C.byref(cfg), resident, lv) == 0
    lf0 = S.lf_params((start[0], start[1]), start[2], start[3], sharpness_level=0)  # the pick resets sharpness (:1136)
    exp_lv = orc.lf_search(g, rec, src, mi, lf0, **kw)
    assert list(lv) == exp_lv
    if H.ref_available():
        assert list(lv) == H.Ref().lf_search(g, rec, src, soa, lf0, **kw)
    exp = orc.deblock(g, rec, mi, S.lf_params((exp_lv[0], exp_lv[1]), exp_lv[2], exp_lv[3], sharpness_level=0))
    assert H.planes_equal(got, exp), H.first_diff(got, exp)


@pytest.mark.parametrize("w,h,bd,sb,seed,kw", [(200, 136, 8, 1, 1, {}), (264, 200, 10, 2, 2, {"unit_size": (128, 64, 64)}),
                                               (134, 70, 10, 2, 3, {}), (200, 136, 10, 2, 4, {"optimized_lr": 1})])
def test_loop_restoration_adapter(H, orc, adapt, w, h, bd, sb, seed, kw):
    """svt_av1_loop_restoration_filter_frame(rst_tmpbuf, frame, cm, optimized_lr): stand-alone (stripe borders from
    rst_info[].boundaries) and at the end of a resident DLF -> CDEF -> LR chain, against the oracle and the reference."""
    from test_lr_oracle import vis

    S = H.S
    g = S.make_geom(w, h, bd, sb)
    rec = H.synth_picture(g, 14000 + seed, noise=10)
    soa = H.synth_mode_info(g, 14100 + seed, p_skip=0.3)
    mi = H.pack_mi(soa)
    lf = S.lf_params((20, 20), 12, 12)
    cdef, idx = H.synth_cdef(g, 14200 + seed)
    lr, units = H.synth_lr(g, 14300 + seed, **kw)
    up = (C.c_void_p * 3)(*[C.cast(u, C.c_void_p) for u in units])
    dbk = orc.deblock(g, rec, mi, lf)
    post = orc.cdef_apply(g, dbk, mi, cdef, idx)
    exp = orc.lr_apply(g, post, dbk, lr, units)
    # stand-alone
    got = [p.copy() for p in post]
    assert adapt.adapt_lr_frame(C.byref(g), *H._planes(got), *H._planes(dbk), C.byref(lr), up) == 0
    assert H.planes_equal(vis(H, g, got), vis(H, g, exp)), H.first_diff(vis(H, g, got), vis(H, g, exp))
    if H.ref_available():
        assert H.planes_equal(vis(H, g, got), vis(H, g, H.Ref().lr_apply(g, post, dbk, lr, units)))
    # resident chain
    s, keep = _soa(H, soa)
    got2 = [p.copy() for p in rec]
    idx8 = np.ascontiguousarray(idx, dtype=np.int8)
    assert adapt.adapt_chain_lr(C.byref(g), *H._planes(got2), C.byref(s), C.byref(lf), C.byref(cdef), H._p(idx8), C.byref(lr), up) == 0
    assert H.planes_equal(vis(H, g, got2), vis(H, g, exp)), H.first_diff(vis(H, g, got2), vis(H, g, exp))


@pytest.mark.parametrize("w,h,bd,sb,resident", [(200, 136, 8, 1, 0), (328, 200, 10, 2, 0), (328, 200, 10, 2, 1), (200, 136, 8, 2, 1),
                                                (322, 198, 10, 2, 0)])
def test_pad_ref_pictures_adapter(H, orc, adapt, w, h, bd, sb, resident):
    """svtf_pad_ref_pictures = the pixel work of pad_ref_and_set_flags, through the reference's own structures: padded reference
    picture (+ 8-bit / 2-bit planes in the 16-bit pipeline) from the recon buffer, or from the picture left on the GPU by
    svt_av1_loop_filter_frame (resident)."""
    S = H.S
    pr, pb = (-w) % 8, (-h) % 8  # picture padded to a multiple of 8 by replication
    g = S.make_geom(w, h, bd, sb)
    pic = H.synth_picture(g, 2500 + w + bd)
    soa = H.synth_mode_info(g, 2501)
    mi = H.pack_mi(soa)
    s, keep = _soa(H, soa)
    lf = S.lf_params((26, 20), 11, 13)
    ox, oy = 48, 32
    filt = orc.deblock(g, pic, mi, lf) if resident else pic
    split, two = sb == 2, sb == 2 and bd == 10
    exp = H.handoff_cpu(orc.lib, "svtf_orc_", g, filt, w + pr, h + pb, ox, oy, pr, pb, split, two)
    if H.ref_available():
        ref = H.handoff_cpu(H.Ref().lib, "ref_", g, filt, w + pr, h + pb, ox, oy, pr, pb, split, two)
        assert H.planes_equal(exp[0], ref[0])
    # padded allocations with the picture placed at (org_x, org_y)
    bufs = [np.zeros_like(p) for p in exp[0]]
    for b, p, sft in zip(bufs, pic, (0, 1, 1)):
        vw, vh = (w + sft) >> sft, (h + sft) >> sft
        b[(oy + sft) >> sft:((oy + sft) >> sft) + vh, (ox + sft) >> sft:((ox + sft) >> sft) + vw] = p[:vh, :vw]
    b8 = [np.zeros(p.shape, np.uint8) for p in bufs]
    bn = [np.zeros(p.shape, np.uint8) for p in bufs]

    def arr(lst):
        return (C.c_void_p * 3)(*[a.ctypes.data for a in lst]), (C.c_int * 3)(*[a.strides[0] // a.itemsize for a in lst])

    p16, s16 = arr(bufs)
    p8, s8 = arr(b8)
    pn, sn = arr(bn)
    assert adapt.adapt_pad_ref(C.byref(g), p16, s16, p8, s8, pn if two else None, sn, ox, oy, pr, pb, C.byref(s), C.byref(lf), resident) == 0
    assert H.planes_equal(bufs, exp[0]), "padded: " + H.first_diff(bufs, exp[0])
    if split:
        assert H.planes_equal(b8, exp[1]), "8-bit: " + H.first_diff(b8, exp[1])
    if two:
        assert H.planes_equal(bn, exp[2]), "2-bit: " + H.first_diff(bn, exp[2])


@pytest.mark.parametrize("w,h,bd,sb,kw", [(200, 136, 10, 2, {}), (136, 72, 8, 1, {"tap_lvl": 2, "one_step": 1}), (264, 136, 8, 2, {"chroma": 0}),
                                          (136, 104, 10, 2, {"sg": 1}), (136, 72, 8, 1, {"wn": 0, "sg": 1, "sg_ep": ((0, 16, 3, 1), (8, 16, 1, 0))}),
                                          (200, 136, 10, 2, {"prev": 1})])
def test_restoration_search_adapter(H, orc, adapt, w, h, bd, sb, kw):
    """restoration_seg_search(.., segment 0) + rest_finish_search on real PictureControlSet / Av1Common / Macroblock structures:
    rusi_picture[] and rst_info[] end up as the oracle's (pinned to the reference's own search) say.  The planes handed over
    are the visible picture with arbitrary strides, as the reference's Yv12BufferConfig describes it."""
    S = H.S
    g = S.make_geom(w, h, bd, sb)
    src = H.synth_picture(g, 14000 + w, noise=10)
    cur = H.synth_source(g, src, 14001, amp=6)
    prev = H.synth_lr(g, 14002, p_none=0.3, p_wiener=0.5)[1] if kw.pop("prev", 0) else None
    cfg = H.lr_search_cfg(prev_units=prev, **kw)
    got = H._lr_search_call(adapt.adapt_lr_search, g, cur, src, cfg)
    vis = lambda pl: [np.ascontiguousarray(p[: (g.height + s) >> s, : (g.width + s) >> s]) for p, s in zip(pl, (0, 1, 1))]
    exp = orc.lr_search(g, vis(cur), vis(src), cfg)
    for p in range(3):
        assert got[p][0] == exp[p][0], ("frame type", p)
        assert got[p][2] == exp[p][2], ("unit search", p)
        assert got[p][1] == exp[p][1], ("unit info", p)
    assert got[3] == exp[3], "sg_frame_ep_cnt"


def test_superres_adapter(H, orc, adapt):
    """svt_av1_superres_upscale_frame(cm, pcs, scs) on the reference's own structures: the recon planes are upscaled in place."""
    from test_superres_oracle import upscale_cpu

    S = H.S
    f = np.load(os.path.join(H.ROOT, "tests", "golden", "superres_filters.npy"))
    for (w, h, bd, sb, denom, starts) in [(200, 72, 10, 2, 11, None), (328, 40, 8, 1, 16, (0, 32, 64, 82))]:
        g = S.make_geom(w, h, bd, sb)
        pic = H.synth_picture(g, 15200 + w, noise=12)
        up_w = (w * denom + 4) // 8
        st = list(starts) if starts else [0, g.mi_cols]
        cw = 2 * ((up_w + 7) >> 3) * 4
        bufs = []
        for p in range(3):
            b = np.zeros((pic[p].shape[0], cw >> (1 if p else 0)), pic[p].dtype)
            b[:, : pic[p].shape[1]] = pic[p]
            bufs.append(b)
        arr = (C.c_int32 * len(st))(*st)
        assert adapt.adapt_superres(C.byref(g), *H._planes(bufs), up_w, denom, len(st) - 1, arr) == 0
        for p in range(3):
            exp = upscale_cpu(H, orc.lib.svtf_orc_superres_plane, g, pic[p], p, up_w, denom, st, f)
            assert np.array_equal(bufs[p][:, : exp.shape[1]], exp), p
