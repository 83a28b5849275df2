"""Test/bench harness: seeded synthetic inputs and ctypes bindings of the two CPU checkers.

  Oracle  — oracle/liborc.so   (oracle/svtf_oracle.c, the CPU restatement; always available)
  Ref     — oracle/_ref/libsvtref.so (the reference's own C compiled from /root/reference by
            oracle/Makefile; present wherever `make -C oracle` ran with the reference mounted, and
            it travels to the GPU box as a built file)

Only tests/, __graft_entry__.smoke() and bench.py's CPU legs import this module.
"""
import ctypes as C
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
# SVTF_HARNESS_NO_PRODUCT=1 (set by `bench.py --impl reference`): the CPU checkers alone, in a process that neither imports the
# product's Python package nor maps libsvtfilt_b200.so; the parameter structs then come from tests/ref_abi.py
NO_PRODUCT = os.environ.get("SVTF_HARNESS_NO_PRODUCT") == "1"
if NO_PRODUCT:
    import ref_abi as S  # noqa: E402
    L = S
else:
    import svtfilt_b200 as S  # noqa: E402
    from svtfilt_b200 import _lib as L  # noqa: E402

ORC_PATH = os.path.join(ROOT, "oracle", "liborc.so")
REF_PATH = os.path.join(ROOT, "oracle", "_ref", "libsvtref.so")
REF_SIMD_PATH = os.path.join(ROOT, "oracle", "_ref", "libsvtref_simd.so")  # the same sources + the reference's SSE2/AVX2 kernels
GOLDEN = os.path.join(ROOT, "tests", "golden")


def build_oracle():
    subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "--no-print-directory"], check=True, stdout=subprocess.DEVNULL)


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _planes(pl):
    out = []
    for a in pl:
        assert a.strides[1] == a.itemsize
        out += [_p(a), a.strides[0] // a.itemsize]
    return out


# ------------------------------------------------------------------------------------------------
# synthetic inputs (SURVEY §8d recipe)
# ------------------------------------------------------------------------------------------------
BLOCK_DIMS = [(4, 4), (4, 8), (8, 4), (8, 8), (8, 16), (16, 8), (16, 16), (16, 32), (32, 16), (32, 32), (32, 64), (64, 32),
              (64, 64), (64, 128), (128, 64), (128, 128), (4, 16), (16, 4), (8, 32), (32, 8), (16, 64), (64, 16)]
BSIZE_OF = {d: i for i, d in enumerate(BLOCK_DIMS)}


def synth_picture(geom, seed, frame=0, noise=6, planes_like=None):
    """Smooth sinusoid + 16x16 block DC steps + uniform noise, scaled to the bit depth."""
    rng = np.random.default_rng(seed)
    sh = geom.bit_depth - 8
    dt = np.uint8 if geom.sample_bytes == 1 else np.uint16
    out = []
    for p in range(3):
        s = 1 if p else 0
        w, h = (geom.mi_cols * 4) >> s, (geom.mi_rows * 4) >> s
        x = np.arange(w, dtype=np.float64)[None, :] * (1 << s)
        y = np.arange(h, dtype=np.float64)[:, None] * (1 << s)
        base = 128 + (60 - 20 * p) * np.sin((x + 3 * frame + 11 * p) / 17.0) + (40 - 10 * p) * np.cos((y - 2 * frame) / 11.0)
        bs = 16 >> s
        dc = rng.integers(-6, 7, size=((h + bs - 1) // bs, (w + bs - 1) // bs))
        base = base + np.kron(dc, np.ones((bs, bs)))[:h, :w]
        base = base + rng.integers(-noise, noise + 1, size=(h, w))
        v = np.clip(np.rint(base), 0, 255).astype(np.int64) << sh
        if sh:
            v = v + rng.integers(0, 1 << sh, size=(h, w))
        out.append(np.ascontiguousarray(v.astype(dt)))
    return out


def synth_source(geom, recon, seed, amp=8):
    """Source picture = recon + uniform noise (for the search distortion)."""
    rng = np.random.default_rng(seed ^ 0x5EED)
    sh = geom.bit_depth - 8
    mx = (1 << geom.bit_depth) - 1
    return [np.ascontiguousarray(np.clip(r.astype(np.int64) + (rng.integers(-amp, amp + 1, size=r.shape) << sh), 0, mx).astype(r.dtype))
            for r in recon]


def synth_mode_info(geom, seed, p_split=0.55, p_skip=0.3, p_inter=0.7, segments=False, sb_size=64):
    """Valid AV1 partition of every 64x64 (or, sb_size=128, 128x128) superblock (NONE / SPLIT / HORZ / VERT / HORZ_4 / VERT_4),
    with a uniform tx_depth, skip, ref_frame, mode and segment per block.  Returns a dict of (mi_rows, mi_cols) uint8 arrays."""
    rng = np.random.default_rng(seed)
    R, Cc = geom.mi_rows, geom.mi_cols
    soa = {k: np.zeros((R, Cc), dtype=np.uint8) for k in ("bsize", "tx_depth", "skip", "ref_frame0", "mode", "segment_id")}

    def put(r, c, w, h):
        if r >= R or c >= Cc:
            return
        r1, c1 = min(R, r + h // 4), min(Cc, c + w // 4)
        inter = rng.random() < p_inter
        soa["bsize"][r:r1, c:c1] = BSIZE_OF[(w, h)]
        soa["tx_depth"][r:r1, c:c1] = rng.integers(0, 3)
        soa["skip"][r:r1, c:c1] = rng.random() < p_skip
        soa["ref_frame0"][r:r1, c:c1] = rng.integers(1, 8) if inter else 0
        soa["mode"][r:r1, c:c1] = rng.integers(13, 25) if inter else rng.integers(0, 13)
        soa["segment_id"][r:r1, c:c1] = rng.integers(0, 8) if segments else 0

    def part(r, c, s):
        if r >= R or c >= Cc:
            return
        u = rng.random()
        if s == 4 or (s > 8 and u >= p_split + 0.25) or (s == 8 and u >= p_split + 0.3):
            put(r, c, s, s)
        elif u < p_split:
            if s == 8:
                for dr, dc in ((0, 0), (0, 1), (1, 0), (1, 1)):
                    put(r + dr, c + dc, 4, 4)
            else:
                h = s // 2
                for dr, dc in ((0, 0), (0, h // 4), (h // 4, 0), (h // 4, h // 4)):
                    part(r + dr, c + dc, h)
        else:
            k = rng.integers(0, 4 if 16 <= s <= 64 else 2)  # no 4:1 partitions of 8x8 or 128x128
            if k == 0:  # HORZ
                put(r, c, s, s // 2), put(r + s // 8, c, s, s // 2)
            elif k == 1:  # VERT
                put(r, c, s // 2, s), put(r, c + s // 8, s // 2, s)
            elif k == 2:  # HORZ_4
                for i in range(4):
                    put(r + i * s // 16, c, s, s // 4)
            else:  # VERT_4
                for i in range(4):
                    put(r, c + i * s // 16, s // 4, s)

    for r in range(0, R, sb_size // 4):
        for c in range(0, Cc, sb_size // 4):
            part(r, c, sb_size)
    return soa


# TxSize enum (Common/Codec/EbDefinitions.h:852-880) by (width, height)
TX_DIMS = [(4, 4), (8, 8), (16, 16), (32, 32), (64, 64), (4, 8), (8, 4), (8, 16), (16, 8), (16, 32), (32, 16), (32, 64), (64, 32),
           (4, 16), (16, 4), (8, 32), (32, 8), (16, 64), (64, 16)]
TX_OF = {d: i for i, d in enumerate(TX_DIMS)}


def _sub_tx(w, h):
    """sub_tx_size_map: squares halve both sides, 2:1 and 4:1 rectangles halve the long side"""
    if w == h:
        return (max(w // 2, 4), max(h // 2, 4))
    return (w // 2, h) if w > h else (w, h // 2)


def synth_dec_mode_info(geom, seed, p_split=0.55, p_skip=0.3, p_inter=0.7, segments=False, var_tx=True):
    """The decoder's view of a frame: the blocks of synth_mode_info with explicit per-4x4 transform sizes (TxSize enums) —
    luma: the largest transform of the block split 0..2 times, per transform unit when var_tx (inter blocks in AV1 carry a
    transform partition tree); chroma: av1_get_max_uv_txsize of the block that carries the chroma (the one at the odd
    row/column of an 8x8).  Keys: bsize, tx_size_y, tx_size_uv, skip, ref_frame0, mode, segment_id."""
    soa = synth_mode_info(geom, seed, p_split, p_skip, p_inter, segments)
    rng = np.random.default_rng(seed + 77)
    R, Cc = geom.mi_rows, geom.mi_cols
    txy = np.zeros((R, Cc), np.uint8)

    def fill(r, c, w, h, depth):
        if r >= R or c >= Cc:
            return
        if depth < 2 and (w, h) != (4, 4) and rng.random() < 0.45:
            sw, sh = _sub_tx(w, h)
            for rr in range(r, r + h // 4, sh // 4):
                for cc in range(c, c + w // 4, sw // 4):
                    fill(rr, cc, sw, sh, depth + 1)
        else:
            txy[r:min(R, r + h // 4), c:min(Cc, c + w // 4)] = TX_OF[(w, h)]

    seen = np.zeros((R, Cc), bool)
    for r in range(R):
        for c in range(Cc):
            if seen[r, c]:
                continue
            w, h = BLOCK_DIMS[soa["bsize"][r, c]]
            seen[r:r + h // 4, c:c + w // 4] = True
            tw, th = min(w, 64), min(h, 64)
            inter = soa["ref_frame0"][r, c] > 0
            if var_tx and inter and not soa["skip"][r, c]:
                for rr in range(r, r + h // 4, th // 4):
                    for cc in range(c, c + w // 4, tw // 4):
                        fill(rr, cc, tw, th, 0)
            else:
                d = 0 if (inter and soa["skip"][r, c]) else int(rng.integers(0, 3))
                for _ in range(d):
                    tw, th = _sub_tx(tw, th)
                for rr in range(r, min(R, r + h // 4), th // 4):
                    for cc in range(c, min(Cc, c + w // 4), tw // 4):
                        txy[rr:min(R, rr + th // 4), cc:min(Cc, cc + tw // 4)] = TX_OF[(tw, th)]
    uv_of_bsize = np.zeros(len(BLOCK_DIMS), np.uint8)
    for i, (w, h) in enumerate(BLOCK_DIMS):
        uv_of_bsize[i] = TX_OF[(min(max(w // 2, 4), 32), min(max(h // 2, 4), 32))]
    rr, cc = np.meshgrid(np.arange(R) | 1, np.arange(Cc) | 1, indexing="ij")
    out = {k: soa[k] for k in ("bsize", "skip", "ref_frame0", "mode", "segment_id")}
    out["tx_size_y"] = txy
    out["tx_size_uv"] = uv_of_bsize[soa["bsize"][np.minimum(rr, R - 1), np.minimum(cc, Cc - 1)]]
    return out


def pack_mi_dec(soa):
    return S.pack_mi_dec(soa["bsize"], soa["tx_size_y"], soa["tx_size_uv"], soa["skip"], soa["ref_frame0"], soa["mode"], soa["segment_id"])


def pack_mi(soa):
    return S.pack_mi_enc(soa["bsize"], soa["tx_depth"], soa["skip"], soa["ref_frame0"], soa["mode"], soa["segment_id"])


def nfb_of(geom):
    return (geom.mi_rows + 15) // 16, (geom.mi_cols + 15) // 16


def synth_cdef(geom, seed, n_strengths=8, damping=5, p_unset=0.0):
    rng = np.random.default_rng(seed ^ 0xCDEF)
    ys = [int(v) for v in rng.integers(0, 64, size=n_strengths)]
    us = [int(v) for v in rng.integers(0, 64, size=n_strengths)]
    ys[0] = 0
    us[0] = 0  # one all-zero pair so the "strength (0,0,0,0)" path is exercised
    nv, nh = nfb_of(geom)
    idx = rng.integers(0, n_strengths, size=nv * nh).astype(np.int8)
    return S.cdef_params(damping, ys, us), idx


def synth_ccso(geom, seed, quant_idx=(0, 0, 0), support=(0, 1, 2), edge_clf=(0, 0, 0), band_log2=(2, 2, 2), bo_only=(0, 0, 0),
               enable=(1, 1, 1), p_on=0.8):
    rng = np.random.default_rng(seed ^ 0xCC50)
    lut = rng.choice(np.array([-10, -7, -3, -1, 0, 1, 3, 7], dtype=np.int8), size=(3, L.CCSO_LUT_SIZE))
    flags = []
    for p in range(3):
        pw, ph, bl = geom.width >> (1 if p else 0), geom.height >> (1 if p else 0), 7 if p else 8
        n = ((pw + (1 << bl) - 1) >> bl) * ((ph + (1 << bl) - 1) >> bl)
        flags.append((rng.random(n) < p_on).astype(np.uint8))
    return S.ccso_params(enable, bo_only, quant_idx, support, edge_clf, band_log2, lut), flags


def lr_units_of(geom, plane, unit_size):
    ss = 1 if plane else 0
    pw, ph = (geom.width + ss) >> ss, (geom.height + ss) >> ss
    return max((pw + unit_size // 2) // unit_size, 1), max((ph + unit_size // 2) // unit_size, 1)


def synth_lr(geom, seed, unit_size=(64, 32, 32), frame_type=(3, 3, 3), optimized_lr=0, p_none=0.15, p_wiener=0.45):
    """Restoration parameters for a frame: per-unit RESTORE_NONE / WIENER (symmetric 7-tap, 5-tap for chroma, within the
    coded ranges of Common/Codec/EbRestoration.h:118-160) / SGRPROJ (ep 0..15, xqd within SGRPROJ_PRJ_MIN/MAX).
    Returns (LrParams, [unit arrays])."""
    rng = np.random.default_rng(seed ^ 0x1235)
    lr = L.LrParams()
    units = []
    for p in range(3):
        lr.frame_restoration_type[p] = frame_type[p]
        lr.restoration_unit_size[p] = unit_size[p]
        hu, vu = lr_units_of(geom, p, unit_size[p])
        arr = (L.LrUnit * (hu * vu))()
        for u in arr:
            t = rng.random()
            u.restoration_type = 0 if t < p_none else (1 if t < p_none + p_wiener else 2)
            if frame_type[p] == 1 and u.restoration_type == 2:
                u.restoration_type = 1
            if frame_type[p] == 2 and u.restoration_type == 1:
                u.restoration_type = 2
            for f in (u.wiener_vfilter, u.wiener_hfilter):
                t0 = 0 if p else int(rng.integers(-5, 11))
                t1, t2 = int(rng.integers(-23, 9)), int(rng.integers(-17, 47))
                f[0], f[1], f[2], f[3], f[4], f[5], f[6], f[7] = t0, t1, t2, -2 * (t0 + t1 + t2), t2, t1, t0, 0
            u.sgr_ep = int(rng.integers(0, 16))
            u.sgr_xqd[0], u.sgr_xqd[1] = int(rng.integers(-96, 32)), int(rng.integers(-32, 96))
        units.append(arr)
    lr.optimized_lr = optimized_lr
    return lr, units


def _lr_call(fn, geom, cur, dbk, lr, units):
    out = [p.copy() for p in cur]
    dbk = [np.ascontiguousarray(p) for p in dbk]
    up = (C.c_void_p * 3)(*[C.cast(u, C.c_void_p) for u in units])
    assert fn(C.byref(geom), *_planes(out), *_planes(dbk), C.byref(lr), up) == 0
    return out


def lr_search_cfg(unit_size=(64, 32, 32), tap_lvl=1, refinement=1, one_step=0, chroma=1, rdmult=9000, cost=(120, 900), wn=1, sg=0, sg_chroma=1,
                  sg_ep=((0, 16, 1, 1), (0, 16, 1, 1)), sg_cost=(130, 950), sw_cost=(150, 1000, 1100), prev_units=None):
    """WnFilterCtrls / SgFilterCtrls of a search (Encoder/Codec/EncModeConfig.c set_wn_filter_ctrls / set_sg_filter_ctrls) + the rate
    inputs.  sg_ep: (start_ep, end_ep, ep_inc, refine) for luma and chroma."""
    c = L.LrSearchCfg()
    for p in range(3):
        c.restoration_unit_size[p] = unit_size[p]
    c.wn_enabled, c.wn_filter_tap_lvl, c.wn_use_refinement, c.wn_max_one_refinement_step, c.wn_use_chroma = wn, tap_lvl, refinement, one_step, chroma
    c.sg_enabled, c.sg_use_chroma, c.rdmult = sg, sg_chroma, rdmult
    for t in range(2):
        c.sg_start_ep[t], c.sg_end_ep[t], c.sg_ep_inc[t], c.sg_refine[t] = sg_ep[t]
        c.wiener_restore_cost[t], c.sgrproj_restore_cost[t] = cost[t], sg_cost[t]
    for t in range(3):
        c.switchable_restore_cost[t] = sw_cost[t]
    if prev_units is not None:  # use_prev_frame_coeffs: LrUnit arrays (kept alive on the struct)
        c._keep = prev_units
        for p in range(3):
            c.prev_units[p] = C.cast(prev_units[p], C.c_void_p) if prev_units[p] is not None else None
    return c


def lr_search_sig(geom, cfg, ftype, units, search, ep_cnt):
    """Comparable form of a search result: frame types, per-unit (type, taps, sgr), per-unit search info, ep counters."""
    sig = []
    for p in range(3):
        hu, vu = lr_units_of(geom, p, cfg.restoration_unit_size[p])
        us = [(u.restoration_type, tuple(u.wiener_vfilter), tuple(u.wiener_hfilter), u.sgr_ep, tuple(u.sgr_xqd)) for u in units[p][: hu * vu]]
        ss = [(tuple(s.sse), tuple(s.wiener_vfilter), tuple(s.wiener_hfilter), s.sgr_ep, tuple(s.sgr_xqd)) for s in search[p][: hu * vu]]
        sig.append((ftype[p], us, ss))
    sig.append(list(ep_cnt))
    return sig


def _lr_search_call(fn, geom, cur, src, cfg):
    cur = [np.ascontiguousarray(p) for p in cur]
    src = [np.ascontiguousarray(p) for p in src]
    ftype = (C.c_int32 * 3)()
    cnt = (C.c_int32 * 16)()
    units, search = [], []
    for p in range(3):
        hu, vu = lr_units_of(geom, p, cfg.restoration_unit_size[p])
        units.append((L.LrUnit * (hu * vu))())
        search.append((L.LrUnitSearch * (hu * vu))())
    up = (C.c_void_p * 3)(*[C.cast(u, C.c_void_p) for u in units])
    sp = (C.c_void_p * 3)(*[C.cast(u, C.c_void_p) for u in search])
    assert fn(C.byref(geom), *_planes(cur), *_planes(src), C.byref(cfg), ftype, up, sp, cnt) == 0
    return lr_search_sig(geom, cfg, ftype, units, search, cnt)


def lr_unit_limits(geom, plane, unit_size):
    """[h0, h1, v0, v1] of every restoration unit of a plane, raster order (foreach_rest_unit_in_tile, EbRestoration.c:1250)."""
    ss = 1 if plane else 0
    pw, ph, off = (geom.width + ss) >> ss, (geom.height + ss) >> ss, 8 >> ss
    hu, vu = lr_units_of(geom, plane, unit_size)
    return [(uc * unit_size, pw if uc == hu - 1 else (uc + 1) * unit_size, max(0, ur * unit_size - off),
             ph if ur == vu - 1 else (ur + 1) * unit_size - off) for ur in range(vu) for uc in range(hu)]


def _wiener_stats_call(fn, geom, plane, dgd, src, lim, win):
    ss = 1 if plane else 0
    pw, ph = (geom.width + ss) >> ss, (geom.height + ss) >> ss
    M, Hm = np.zeros(win * win, np.int64), np.zeros((win * win, win * win), np.int64)
    dgd, src = np.ascontiguousarray(dgd), np.ascontiguousarray(src)
    fn(_p(dgd), dgd.strides[0] // dgd.itemsize, _p(src), src.strides[0] // src.itemsize, geom.sample_bytes, geom.bit_depth, pw, ph,
       *[int(v) for v in lim], win, _p(M), _p(Hm))
    return M, Hm


def _wiener_from_stats_call(fn, win, M, Hm):
    vf, hf = (C.c_int16 * 8)(), (C.c_int16 * 8)()
    M, Hm = np.ascontiguousarray(M), np.ascontiguousarray(Hm)
    ok = fn(win, _p(M), _p(Hm), vf, hf)
    return ok, tuple(vf), tuple(hf)


def cdef_level1_lists():
    """cdef_level 1 (exhaustive, presets <= M0): 16 primary {0,4,..,60} then 48 secondary pri+{1,2,3}
    (Encoder/Codec/EncModeConfig.c:871-910)."""
    first = [4 * i for i in range(16)]
    second = [4 * i + s for i in range(16) for s in (1, 2, 3)]
    return first, second



# ------------------------------------------------------------------------------------------------
# CCSO derivation helpers (shared by Oracle and Ref): planes as uint16 at luma stride
# ------------------------------------------------------------------------------------------------
def ccso_nblocks(geom, plane):
    pw, ph, bl = geom.width >> (1 if plane else 0), geom.height >> (1 if plane else 0), 7 if plane else 8
    return ((pw + (1 << bl) - 1) >> bl) * ((ph + (1 << bl) - 1) >> bl)


def _ccso_derive_call(fn, geom, plane, pre_luma, rec_plane, org_plane, rdmult):
    W, H = geom.width, geom.height
    luma = np.ascontiguousarray(pre_luma[:H, :W], dtype=np.uint16)
    ph, pw = H >> (1 if plane else 0), W >> (1 if plane else 0)
    rec = np.zeros((H, W), dtype=np.uint16)
    org = np.zeros((H, W), dtype=np.uint16)
    rec[:ph, :pw] = rec_plane[:ph, :pw]
    org[:ph, :pw] = org_plane[:ph, :pw]
    out = L.CcsoParams()
    flags = np.zeros(ccso_nblocks(geom, plane), dtype=np.uint8)
    assert fn(C.byref(geom), C.c_int(plane), _p(luma), _p(rec), _p(org), C.c_int(rdmult), C.byref(out), _p(flags)) == 0
    return out, flags


def ccso_plane_sig(params, plane):
    """Comparable summary of one plane's derived CCSO parameters."""
    if not params.ccso_enable[plane]:
        return (0,)
    lut = bytes(bytearray(np.ctypeslib.as_array(params.filter_offset[plane]).view(np.uint8)))
    return (1, params.ccso_bo_only[plane], params.quant_idx[plane], params.ext_filter_support[plane], params.edge_clf[plane],
            params.max_band_log2[plane], lut)

# ------------------------------------------------------------------------------------------------
# the CPU restatement
# ------------------------------------------------------------------------------------------------
class Oracle:
    def __init__(self):
        if not os.path.exists(ORC_PATH):
            build_oracle()
        self.lib = C.CDLL(ORC_PATH)
        self.lib.orc_cdef_dist.restype = C.c_uint64

    def deblock(self, geom, planes, mi, lf, plane_start=0, plane_end=3, sb_delta_lf=None, sb_size_log2=6):
        out = [p.copy() for p in planes]
        mi = np.ascontiguousarray(mi, dtype=np.uint32)
        if sb_delta_lf is None:
            r = self.lib.orc_deblock_frame(C.byref(geom), *_planes(out), _p(mi), C.c_int(mi.shape[1]), C.byref(lf), plane_start, plane_end)
        else:
            d = np.ascontiguousarray(sb_delta_lf, dtype=np.int32)
            r = self.lib.orc_deblock_frame_sb_delta(C.byref(geom), *_planes(out), _p(mi), C.c_int(mi.shape[1]), C.byref(lf), plane_start,
                                                    plane_end, _p(d), C.c_int(d.shape[1]), C.c_int(sb_size_log2))
        assert r == 0
        return out

    def cdef_apply(self, geom, planes, mi, cdef, fb_idx):
        out = [p.copy() for p in planes]
        mi = np.ascontiguousarray(mi, dtype=np.uint32)
        idx = np.ascontiguousarray(fb_idx, dtype=np.int8)
        r = self.lib.orc_cdef_apply_frame(C.byref(geom), *_planes(out), _p(mi), C.c_int(mi.shape[1]), C.byref(cdef), _p(idx))
        assert r == 0
        return out

    def ccso_apply(self, geom, planes, pre_luma, ccso, flags):
        out = [p.copy() for p in planes]
        fl = (C.c_void_p * 3)(*[None if f is None else f.ctypes.data for f in (flags or [None] * 3)])
        r = self.lib.orc_ccso_apply_frame(C.byref(geom), *_planes(out), _p(pre_luma), C.c_int(pre_luma.strides[0] // pre_luma.itemsize),
                                          C.byref(ccso), fl if flags is not None else None)
        assert r == 0
        return out

    def ccso_derive_plane(self, geom, plane, pre_luma, rec_plane, org_plane, rdmult):
        return _ccso_derive_call(self.lib.orc_ccso_derive_plane, geom, plane, pre_luma, rec_plane, org_plane, rdmult)

    def lr_apply(self, geom, cur, dbk, lr, units):
        return _lr_call(self.lib.orc_lr_frame, geom, cur, dbk, lr, units)

    def lr_search(self, geom, cur, src, cfg):
        return _lr_search_call(self.lib.orc_lr_search_frame, geom, cur, src, cfg)

    def wiener_stats(self, geom, plane, dgd, src, lim, win):
        self.lib.orc_lr_wiener_stats.restype = None
        return _wiener_stats_call(self.lib.orc_lr_wiener_stats, geom, plane, dgd, src, lim, win)

    def wiener_from_stats(self, win, M, Hm):
        return _wiener_from_stats_call(self.lib.orc_lr_wiener_from_stats, win, M, Hm)

    def lf_search(self, geom, rec, src, mi, lf, dlf_avg=0, early_exit_convergence=0, only_4x4=0, search_chroma=1):
        rec = [p.copy() for p in rec]
        mi = np.ascontiguousarray(mi, dtype=np.uint32)
        cfg = L.LfSearchCfg(dlf_avg, early_exit_convergence, only_4x4, search_chroma)
        out = (C.c_int32 * 4)()
        assert self.lib.orc_lf_search_frame(C.byref(geom), *_planes(rec), *_planes(src), _p(mi), C.c_int(mi.shape[1]), C.byref(lf),
                                            C.byref(cfg), out) == 0
        return list(out)

    def cdef_search(self, geom, rec, src, mi, cfg):
        nv, nh = nfb_of(geom)
        nfb = nv * nh
        mse = np.zeros((2, nfb, 64), dtype=np.uint64)
        skip = np.zeros(nfb, dtype=np.uint8)
        dirs = np.zeros((nfb, 64), dtype=np.uint8)
        var = np.zeros((nfb, 64), dtype=np.int32)
        mi = np.ascontiguousarray(mi, dtype=np.uint32)
        r = self.lib.orc_cdef_search_frame(C.byref(geom), *_planes(rec), *_planes(src), _p(mi), C.c_int(mi.shape[1]), C.byref(cfg),
                                           _p(mse), _p(skip), _p(dirs), _p(var))
        assert r == 0
        return mse, skip, dirs, var

    def cdef_pick(self, cfg, mse, skip):
        mse = np.ascontiguousarray(mse, dtype=np.uint64).copy()
        skip = np.ascontiguousarray(skip, dtype=np.uint8)
        out = L.CdefParams()
        idx = np.empty(skip.size, dtype=np.int8)
        r = self.lib.orc_cdef_pick(C.byref(cfg), _p(mse), _p(skip), C.c_int(skip.size), C.byref(out), _p(idx))
        assert r == 0
        return out, idx

    def lpf_edge(self, buf, offset, along, nxt, count, length, bd, blimit, limit, thresh):
        """buf: 2-D array modified in place; offset: flat index of q0 of line 0."""
        base = buf.ctypes.data + offset * buf.itemsize
        self.lib.orc_lpf_edge(C.c_void_p(base), buf.itemsize, along, nxt, count, length, bd, blimit, limit, thresh)

    def find_dir(self, img, coeff_shift):
        img = np.ascontiguousarray(img, dtype=np.uint16)
        var = C.c_int32()
        d = self.lib.orc_cdef_find_dir(_p(img), C.c_int(img.shape[1]), C.byref(var), coeff_shift)
        return d, var.value

    def filter_block(self, tile, y, x, pri, sec, direction, pri_damping, sec_damping, bw, bh, coeff_shift, sub=1):
        tile = np.ascontiguousarray(tile, dtype=np.uint16)
        dst = np.zeros((bh, bw), dtype=np.uint16)
        base = tile.ctypes.data + (y * tile.shape[1] + x) * 2
        self.lib.orc_cdef_filter_block(_p(dst), bw, C.c_void_p(base), tile.shape[1], pri, sec, direction, pri_damping, sec_damping, bw, bh,
                                       coeff_shift, sub)
        return dst

    def cdef_dist(self, src, flt, by, bx, bw, bh, coeff_shift, pli, sub):
        by = np.ascontiguousarray(by, dtype=np.uint8)
        bx = np.ascontiguousarray(bx, dtype=np.uint8)
        flt = np.ascontiguousarray(flt, dtype=np.uint16)
        return int(self.lib.orc_cdef_dist(_p(src), C.c_int(src.strides[0] // src.itemsize), src.itemsize, _p(flt), _p(by), _p(bx), len(by), bw, bh,
                                          coeff_shift, pli, sub))


# ------------------------------------------------------------------------------------------------
# the real reference (compiled from /root/reference)
# ------------------------------------------------------------------------------------------------
class RefMiSoa(C.Structure):
    _fields_ = [("bsize", C.c_void_p), ("tx_depth", C.c_void_p), ("skip", C.c_void_p), ("ref_frame0", C.c_void_p),
                ("mode", C.c_void_p), ("segment_id", C.c_void_p), ("mi_stride", C.c_int32)]


def ref_available():
    return os.path.exists(REF_PATH)


def cpu_has_avx2():
    try:
        with open("/proc/cpuinfo") as f:
            return " avx2 " in f.read().replace("\n", " ") + " "
    except OSError:
        return False


def ref_simd_available():
    return os.path.exists(REF_SIMD_PATH) and cpu_has_avx2()


class Ref:
    """simd=False: oracle/_ref/libsvtref.so, RTCD pointers -> the `_c` kernels (the parity checker).
    simd=True: oracle/_ref/libsvtref_simd.so, RTCD pointers -> the SSE2 / AVX2 kernels the reference selects on an AVX2 host
    (the timed CPU baseline)."""

    def __init__(self, simd=False):
        self.simd = simd
        self.lib = C.CDLL(REF_SIMD_PATH if simd else REF_PATH)
        self.lib.ref_cdef_dist.restype = C.c_uint64
        assert self.lib.ref_is_simd() == int(simd)

    @staticmethod
    def _soa(soa):
        keep = {k: np.ascontiguousarray(v, dtype=np.uint8) for k, v in soa.items()}
        s = RefMiSoa(*[keep[k].ctypes.data for k in ("bsize", "tx_depth", "skip", "ref_frame0", "mode", "segment_id")],
                     keep["bsize"].shape[1])
        return s, keep

    def deblock(self, geom, planes, soa, lf, plane_start=0, plane_end=3):
        out = [p.copy() for p in planes]
        s, keep = self._soa(soa)
        assert self.lib.ref_deblock_frame(C.byref(geom), *_planes(out), C.byref(s), C.byref(lf), plane_start, plane_end) == 0
        return out

    def cdef_apply(self, geom, planes, soa, cdef, fb_idx):
        out = [p.copy() for p in planes]
        s, keep = self._soa(soa)
        idx = np.ascontiguousarray(fb_idx, dtype=np.int8)
        assert self.lib.ref_cdef_apply_frame(C.byref(geom), *_planes(out), C.byref(s), C.byref(cdef), _p(idx)) == 0
        return out

    def ccso_apply(self, geom, planes, pre_luma, ccso, flags):
        out = [p.copy() for p in planes]
        fl = (C.c_void_p * 3)(*[None if f is None else f.ctypes.data for f in (flags or [None] * 3)])
        assert self.lib.ref_ccso_apply_frame(C.byref(geom), *_planes(out), _p(pre_luma), C.c_int(pre_luma.strides[0] // pre_luma.itemsize),
                                             C.byref(ccso), fl if flags is not None else None) == 0
        return out

    def ccso_derive_plane(self, geom, plane, pre_luma, rec_plane, org_plane, rdmult):
        return _ccso_derive_call(self.lib.ref_ccso_derive_plane, geom, plane, pre_luma, rec_plane, org_plane, rdmult)

    def lr_apply(self, geom, cur, dbk, lr, units):
        return _lr_call(self.lib.ref_lr_frame, geom, cur, dbk, lr, units)

    def lr_search(self, geom, cur, src, cfg):
        return _lr_search_call(self.lib.ref_lr_search, geom, cur, src, cfg)

    def wiener_stats(self, geom, plane, dgd, src, lim, win):
        return _wiener_stats_call(self.lib.ref_lr_wiener_stats, geom, plane, dgd, src, lim, win)

    def wiener_from_stats(self, win, M, Hm):
        return _wiener_from_stats_call(self.lib.ref_lr_wiener_from_stats, win, M, Hm)

    def lf_search(self, geom, rec, src, soa, lf, dlf_avg=0, early_exit_convergence=0, only_4x4=0, search_chroma=1):
        rec = [p.copy() for p in rec]
        s, keep = self._soa(soa)
        cfg = L.LfSearchCfg(dlf_avg, early_exit_convergence, only_4x4, search_chroma)
        out = (C.c_int32 * 4)()
        assert self.lib.ref_lf_search_frame(C.byref(geom), *_planes(rec), *_planes(src), C.byref(s), C.byref(lf), C.byref(cfg), out) == 0
        return list(out)

    def cdef_search(self, geom, rec, src, soa, cfg):
        nv, nh = nfb_of(geom)
        nfb = nv * nh
        mse = np.zeros((2, nfb, 64), dtype=np.uint64)
        skip = np.zeros(nfb, dtype=np.uint8)
        dirs = np.zeros((nfb, 64), dtype=np.uint8)
        var = np.zeros((nfb, 64), dtype=np.int32)
        s, keep = self._soa(soa)
        assert self.lib.ref_cdef_search_frame(C.byref(geom), *_planes(rec), *_planes(src), C.byref(s), C.byref(cfg), _p(mse), _p(skip),
                                              _p(dirs), _p(var)) == 0
        return mse, skip, dirs, var

    def cdef_pick(self, geom, soa, cfg, mse, skip):
        mse = np.ascontiguousarray(mse, dtype=np.uint64)
        skip = np.ascontiguousarray(skip, dtype=np.uint8)
        out = L.CdefParams()
        idx = np.empty(skip.size, dtype=np.int8)
        s, keep = self._soa(soa)
        assert self.lib.ref_cdef_pick(C.byref(geom), C.byref(s), C.byref(cfg), _p(mse), _p(skip), C.byref(out), _p(idx)) == 0
        return out, idx

    def lpf(self, buf, offset, pitch, vertical, length, bd, blimit, limit, thresh):
        base = buf.ctypes.data + offset * buf.itemsize
        self.lib.ref_lpf(C.c_void_p(base), buf.itemsize, pitch, vertical, length, bd, C.c_uint8(blimit), C.c_uint8(limit), C.c_uint8(thresh))

    def find_dir(self, img, coeff_shift):
        img = np.ascontiguousarray(img, dtype=np.uint16)
        var = C.c_int32()
        d = self.lib.ref_cdef_find_dir(_p(img), C.c_int(img.shape[1]), C.byref(var), coeff_shift)
        return d, var.value

    def bstride(self):
        return self.lib.ref_cdef_bstride()

    def filter_block(self, tile, y, x, pri, sec, direction, pri_damping, sec_damping, bw, bh, coeff_shift, sub=1):
        """tile must have row stride CDEF_BSTRIDE."""
        assert tile.shape[1] == self.bstride()
        dst = np.zeros((bh, bw), dtype=np.uint16)
        bsize = {(4, 4): 0, (4, 8): 1, (8, 4): 2, (8, 8): 3}[(bw, bh)]
        base = tile.ctypes.data + (y * tile.shape[1] + x) * 2
        self.lib.ref_cdef_filter_block(_p(dst), bw, C.c_void_p(base), pri, sec, direction, pri_damping, sec_damping, bsize, coeff_shift, sub)
        return dst

    def cdef_dist(self, src, flt, by, bx, bw, bh, coeff_shift, pli, sub):
        by = np.ascontiguousarray(by, dtype=np.uint8)
        bx = np.ascontiguousarray(bx, dtype=np.uint8)
        bsize = {(4, 4): 0, (4, 8): 1, (8, 4): 2, (8, 8): 3}[(bw, bh)]
        is16 = src.itemsize == 2
        f = np.ascontiguousarray(flt, dtype=np.uint16 if is16 else np.uint8)
        return int(self.lib.ref_cdef_dist(_p(src), C.c_int(src.strides[0] // src.itemsize), _p(f), _p(by), _p(bx), len(by), bsize, coeff_shift,
                                          pli, sub, int(is16)))

    def mi_tables(self, bsize, tx_depth):
        v = [C.c_int() for _ in range(6)]
        self.lib.ref_mi_tables(bsize, tx_depth, *[C.byref(x) for x in v])
        return [x.value for x in v]

    def mode_lf(self, mode):
        return self.lib.ref_mode_lf_lut(mode)

    def tx_dims(self, tx):
        w, h = C.c_int(), C.c_int()
        self.lib.ref_tx_size_dims(tx, C.byref(w), C.byref(h))
        return w.value, h.value


REFDEC_PATH = os.path.join(ROOT, "oracle", "_ref", "libsvtrefdec.so")


class DecMiSoa(C.Structure):
    _fields_ = [(k, C.c_void_p) for k in ("bsize", "tx_size_y", "tx_size_uv", "skip", "ref_frame0", "mode", "segment_id")] + [
        ("mi_stride", C.c_int32), ("sb128", C.c_int32), ("sb_delta_lf", C.c_void_p)]


def dec_soa_struct(soa, sb_delta_lf=None):
    keep = {k: np.ascontiguousarray(soa[k], dtype=np.uint8) for k in ("bsize", "tx_size_y", "tx_size_uv", "skip", "ref_frame0", "mode", "segment_id")}
    s = DecMiSoa(*[keep[k].ctypes.data for k in ("bsize", "tx_size_y", "tx_size_uv", "skip", "ref_frame0", "mode", "segment_id")],
                 keep["bsize"].shape[1], 0, None)
    if sb_delta_lf is not None:
        keep["d"] = np.ascontiguousarray(sb_delta_lf, dtype=np.int32)
        s.sb_delta_lf = keep["d"].ctypes.data
    return s, keep


class DecDriver:
    """The three decoder frame functions behind a library that exports <prefix>_lf_frame / _cdef_frame / _ccso_frame:
    oracle/_ref/libsvtrefdec.so (prefix refdec: the REAL decoder code) or the adapter test library (prefix adaptdec)."""

    def __init__(self, path=REFDEC_PATH, prefix="refdec"):
        self.lib = C.CDLL(path)
        self.prefix = prefix

    def _fn(self, name):
        return getattr(self.lib, f"{self.prefix}_{name}")

    def deblock(self, geom, planes, soa, lf, plane_start=0, plane_end=3, sb_delta_lf=None):
        out = [p.copy() for p in planes]
        s, keep = dec_soa_struct(soa, sb_delta_lf)
        assert self._fn("lf_frame")(C.byref(geom), *_planes(out), C.byref(s), C.byref(lf), plane_start, plane_end) == 0
        return out

    def cdef_apply(self, geom, planes, soa, cdef, fb_idx):
        out = [p.copy() for p in planes]
        s, keep = dec_soa_struct(soa)
        idx = np.ascontiguousarray(fb_idx, dtype=np.int8)
        assert self._fn("cdef_frame")(C.byref(geom), *_planes(out), C.byref(s), C.byref(cdef), _p(idx)) == 0
        return out

    def ccso_apply(self, geom, planes, pre_luma, soa, ccso, blk):
        out = [p.copy() for p in planes]
        s, keep = dec_soa_struct(soa)
        pre = np.ascontiguousarray(pre_luma, dtype=np.uint8)
        b = (C.c_uint8 * 3)(*blk)
        assert self._fn("ccso_frame")(C.byref(geom), *_planes(out), _p(pre), C.c_int(pre.shape[1]), C.byref(s), C.byref(ccso), b) == 0
        return out


    def lr_apply(self, geom, cur, dbk, soa, lr, units):
        """svt_aom_dec_av1_loop_restoration_save_boundary_lines(0 / 1) + _filter_frame on the decoder handle; units in
        the svtf layout ([vert][horz] per plane)."""
        out = [p.copy() for p in cur]
        dbk = [np.ascontiguousarray(p) for p in dbk]
        s, keep = dec_soa_struct(soa)
        up = (C.c_void_p * 3)(*[C.cast(u, C.c_void_p) for u in units])
        assert self._fn("lr_frame")(C.byref(geom), *_planes(out), *_planes(dbk), C.byref(s), C.byref(lr), up) == 0
        return out


def refdec_available():
    return os.path.exists(REFDEC_PATH)


def planes_equal(a, b):
    return all(np.array_equal(x, y) for x, y in zip(a, b))


def first_diff(a, b):
    for p, (x, y) in enumerate(zip(a, b)):
        d = np.argwhere(x != y)
        if len(d):
            r, c = d[0]
            return f"plane {p}: {len(d)} diffs, first at (y={r}, x={c}): {x[r, c]} vs {y[r, c]}"
    return "equal"


# ---- recon hand-off (oracle/svtf_oracle_handoff.c, oracle/ref_shim_handoff.c) ----------------------------------
def handoff_cpu(lib, prefix, geom, planes, width, height, org_x, org_y, pad_right=0, pad_bottom=0, split=False, two_bit=False):
    """pad_ref_and_set_flags on the host with `lib` (the oracle: prefix 'svtf_orc_', the reference: 'ref_'): places the valid
    part of every plane at (org_x, org_y) of a padded allocation, pads it, optionally splits it.  Same returns as
    FilterContext.download_ref_padded."""
    vp = C.c_void_p
    out, o8, on = [], [] if split else None, [] if two_bit else None
    for p, s in zip(planes, (0, 1, 1)):
        w, h = (width + s) >> s, (height + s) >> s
        ox, oy = (org_x + s) >> s, (org_y + s) >> s
        pr, pb = pad_right >> s, pad_bottom >> s
        vw, vh = ((width - pad_right) + s) >> s, ((height - pad_bottom) + s) >> s
        pr, pb = w - vw, h - vh
        buf = np.zeros((h + 2 * oy, w + 2 * ox), dtype=p.dtype)
        buf[oy:oy + vh, ox:ox + vw] = p[:vh, :vw]
        getattr(lib, prefix + "pad_plane")(vp(buf.ctypes.data), buf.shape[1], p.dtype.itemsize, w, h, ox, oy, pr, pb)
        out.append(buf)
        if split:
            b8 = np.zeros(buf.shape, np.uint8)
            bn = np.zeros(buf.shape, np.uint8) if two_bit else None
            if geom.bit_depth == 10:
                getattr(lib, prefix + "unpack2d")(vp(buf.ctypes.data), buf.shape[1], vp(b8.ctypes.data), b8.shape[1],
                                                  vp(bn.ctypes.data) if two_bit else None, buf.shape[1], buf.shape[1], buf.shape[0])
            else:
                getattr(lib, prefix + "convert_16_to_8")(vp(buf.ctypes.data), buf.shape[1], vp(b8.ctypes.data), b8.shape[1], buf.shape[1], buf.shape[0])
            o8.append(b8)
            if two_bit:
                on.append(bn)
    return out, o8, on


def widen_cpu(lib, prefix, planes8):
    out = []
    for p in planes8:
        d = np.zeros(p.shape, np.uint16)
        getattr(lib, prefix + "convert_8_to_16")(C.c_void_p(p.ctypes.data), p.shape[1], C.c_void_p(d.ctypes.data), d.shape[1], p.shape[1], p.shape[0])
        out.append(d)
    return out
