"""GPU: svtf_cdef_pick_table_noalloc (the pick on the library's static device scratch array, no allocation) against the
existing svtf_cdef_pick_table through the C ABI — bit-exact (integer work), on two calls in a row (the second call finds
the first call's table, totals and state in the shared array), plus the size limit.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _sig(p):
    return [p.cdef_bits, p.cdef_damping, *p.cdef_y_strength, *p.cdef_uv_strength]


def _table(rng, nfb, pskip, coarse):
    mse = rng.integers(0, 1 << 22, (2, nfb, 64)).astype(np.uint64)
    mse[:, :, 0] += 1000
    if coarse:
        mse = (mse >> 12) << 12  # many exact ties between (j, k) pairs
    skip = (rng.random(nfb) < pskip).astype(np.uint8)
    return mse, skip


def test_noalloc_pick_matches_pick_table_twice_in_a_row(H):
    S = H.S
    rng = np.random.default_rng(0xA110C)
    first, second = H.cdef_level1_lists()
    g = S.make_geom(64, 64, 8)
    with S.FilterContext(g) as ctx:
        # a large table first, a small one second: stale rows of the first call stay in the array beyond the second's nfb
        for it, (nfb, n1, n2, bias, pskip) in enumerate(((2040, 16, 48, 0, 0.1), (130, 8, 8, 40, 0.3), (8160, 16, 48, 0, 0.2))):
            cfg = S.cdef_search_cfg(first[:n1], second[:n2], base_q_idx=int(rng.integers(0, 256)), zero_fs_cost_bias=bias,
                                    lambda_=int(rng.integers(1, 50000)))
            mse, skip = _table(rng, nfb, pskip, coarse=(it == 0))
            want = ctx.pick_table(cfg, mse, skip)
            got1 = ctx.pick_table_noalloc(cfg, mse, skip)
            got2 = ctx.pick_table_noalloc(cfg, mse, skip)
            for got in (got1, got2):
                assert _sig(got[0]) == _sig(want[0]), it
                assert np.array_equal(got[1], want[1]), it


def test_noalloc_pick_rejects_tables_beyond_the_static_scratch(H):
    S = H.S
    L = H.L
    first, second = H.cdef_level1_lists()
    cfg = S.cdef_search_cfg(first, second)
    nfb = 8193  # SVTF_NOALLOC_MAX_FB + 1
    mse = np.zeros((2, nfb, 64), np.uint64)
    skip = np.zeros(nfb, np.uint8)
    with S.FilterContext(S.make_geom(64, 64, 8)) as ctx:
        with pytest.raises(S.SvtfError) as e:
            ctx.pick_table_noalloc(cfg, mse, skip)
        assert e.value.code == L.E_ARG
        assert "SVTF_NOALLOC_MAX_FB" in str(e.value)
