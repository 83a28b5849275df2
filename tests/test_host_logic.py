"""CPU: host-side product logic (mi packing, deblock level tables, strength pick) against the
reference's tables/functions and the oracle."""
import ctypes as C

import numpy as np
import pytest


def test_pack_mi_enc_matches_reference_tables(H, ref):
    S = H.S
    one = np.zeros((1, 1), np.uint8)
    for bs in range(22):
        for td in range(3):
            for skip, ref0 in ((0, 0), (1, 0), (1, 3), (0, 5)):
                mi = int(S.pack_mi_enc(one + bs, one + td, one + skip, one + ref0, one, one)[0, 0])
                eff_td = 0 if (skip and ref0 > 0) else td  # get_transform_size: skip&&inter -> depth 0
                exp = ref.mi_tables(bs, eff_td)
                got = [4 << (mi & 7), 4 << ((mi >> 3) & 7), 4 << ((mi >> 6) & 7), 4 << ((mi >> 9) & 7), 4 << ((mi >> 12) & 7), 4 << ((mi >> 15) & 7)]
                assert got == exp, (bs, td, skip, ref0)
                assert (mi >> 18) & 1 == int(skip and ref0 > 0) and (mi >> 19) & 1 == skip and (mi >> 20) & 7 == ref0
    for mode in range(25):
        mi = int(S.pack_mi_enc(one, one, one, one, one + mode, one + 5)[0, 0])
        assert (mi >> 23) & 1 == ref.mode_lf(mode) and (mi >> 24) & 7 == 5


def test_pack_mi_dec_matches_reference_tx_dims(H, ref):
    S = H.S
    one = np.zeros((1, 1), np.uint8)
    for ty in range(19):
        for tuv in (0, 1, 2, 3, 5, 6, 9, 13, 16):
            mi = int(S.pack_mi_dec(one + 12, one + ty, one + tuv, one, one, one, one)[0, 0])
            assert (4 << (mi & 7), 4 << ((mi >> 3) & 7)) == ref.tx_dims(ty)
            assert (4 << ((mi >> 6) & 7), 4 << ((mi >> 9) & 7)) == ref.tx_dims(tuv)


def _tables(fn, lf):
    lvl, lim, mblim, hev = (np.zeros(n, np.uint8) for n in (768, 64, 64, 64))
    fn(C.byref(lf), *[a.ctypes.data_as(C.c_void_p) for a in (lvl, lim, mblim, hev)])
    return lvl, lim, mblim, hev


def test_lf_tables_match_oracle(H, orc):
    S = H.S
    lib = H.L.load()
    rng = np.random.default_rng(7)
    for it in range(200):
        en = rng.integers(0, 2, (8, 8))
        da = rng.integers(-63, 64, (8, 8))
        lf = S.lf_params((int(rng.integers(0, 64)), int(rng.integers(0, 64))), int(rng.integers(0, 64)), int(rng.integers(0, 64)),
                         sharpness_level=int(rng.integers(0, 8)), mode_ref_delta_enabled=int(rng.integers(0, 2)),
                         ref_deltas=tuple(int(v) for v in rng.integers(-20, 21, 8)), mode_deltas=tuple(int(v) for v in rng.integers(-10, 11, 2)),
                         segmentation_enabled=int(rng.integers(0, 2)), seg_feature_enabled=en, seg_feature_data=da,
                         delta_lf_present=int(rng.integers(0, 2)), delta_lf_multi=int(rng.integers(0, 2)),
                         curr_delta_lf=tuple(int(v) for v in rng.integers(-30, 31, 4)))
        a = _tables(lib.svtf_lf_build_tables, lf)
        b = _tables(orc.lib.orc_lf_tables, lf)
        for x, y in zip(a, b):
            assert np.array_equal(x, y), it


def test_cdef_pick_host_matches_oracle_and_reference(H, orc):
    S = H.S
    rng = np.random.default_rng(99)
    first, second = H.cdef_level1_lists()
    for it, (nfb, n1, n2, bias) in enumerate(((1, 16, 48, 0), (7, 3, 3, 0), (60, 16, 48, 40), (130, 8, 8, 0), (33, 1, 0, 0))):
        cfg = S.cdef_search_cfg(first[:n1], second[:n2], base_q_idx=int(rng.integers(0, 256)), zero_fs_cost_bias=bias,
                                lambda_=int(rng.integers(1, 50000)))
        mse = rng.integers(0, 1 << 22, (2, nfb, 64)).astype(np.uint64)
        mse[:, :, 0] += 1000
        skip = (rng.random(nfb) < 0.2).astype(np.uint8)
        if nfb > 3:
            mse[:, 1] = mse[:, 0]  # ties: first index must win
        a = S.cdef_pick_host(cfg, mse, skip)
        b = orc.cdef_pick(cfg, mse, skip)
        assert [a[0].cdef_bits, a[0].cdef_damping, *a[0].cdef_y_strength, *a[0].cdef_uv_strength] == \
               [b[0].cdef_bits, b[0].cdef_damping, *b[0].cdef_y_strength, *b[0].cdef_uv_strength], it
        assert np.array_equal(a[1], b[1]), it
        if H.ref_available():
            g = S.make_geom(64 * nfb, 64, 8)
            soa = {k: np.zeros((g.mi_rows, g.mi_cols), np.uint8) for k in ("bsize", "tx_depth", "skip", "ref_frame0", "mode", "segment_id")}
            soa["bsize"][:] = 12
            r = H.Ref().cdef_pick(g, soa, cfg, mse, skip)
            assert [a[0].cdef_bits, a[0].cdef_damping, *list(a[0].cdef_y_strength)[:8]] == [r[0].cdef_bits, r[0].cdef_damping, *list(r[0].cdef_y_strength)[:8]]
            assert list(a[0].cdef_uv_strength)[:8] == list(r[0].cdef_uv_strength)[:8]
            assert np.array_equal(a[1], r[1]), it
