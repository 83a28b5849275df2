"""CCSO derivation (Encoder/Codec/EbPickccso.c derive_ccso_filter): the oracle restatement against the reference's
own function compiled into oracle/_ref (pins the oracle for SURVEY rows O4-O7)."""
import numpy as np
import pytest


def _case(H, w, h, bd, seed, amp):
    S = H.S
    g = S.make_geom(w, h, bd, 2)
    pre = H.synth_picture(g, seed)          # pre-CDEF picture (classifier luma)
    rec = H.synth_picture(g, seed + 1)      # what CCSO would modify
    rec = [np.clip(p.astype(np.int32) + ((q.astype(np.int32) - p.astype(np.int32)) >> 2), 0, (1 << bd) - 1).astype(p.dtype)
           for p, q in zip(pre, rec)]
    org = H.synth_source(g, rec, seed + 2, amp=amp)
    return g, pre, rec, org


@pytest.mark.parametrize("w,h,bd,rdmult,amp", [(96, 64, 8, 40, 6), (264, 136, 10, 300, 8), (136, 264, 10, 5000, 3), (72, 40, 8, 1, 12)])
def test_oracle_derive_matches_reference(H, orc, ref, w, h, bd, rdmult, amp):
    g, pre, rec, org = _case(H, w, h, bd, 9000 + w, amp)
    for plane in range(3):
        a = orc.ccso_derive_plane(g, plane, pre[0], rec[plane], org[plane], rdmult)
        b = ref.ccso_derive_plane(g, plane, pre[0], rec[plane], org[plane], rdmult)
        assert H.ccso_plane_sig(a[0], plane) == H.ccso_plane_sig(b[0], plane), f"plane {plane}"
        assert np.array_equal(a[1], b[1]), f"plane {plane} flags"


def test_oracle_derive_with_extremes(H, orc, ref):
    """Samples at 0 and max (the trial filter clips), a constant region (empty classes)."""
    S = H.S
    g = S.make_geom(136, 72, 10, 2)
    rng = np.random.default_rng(5)
    pre = [rng.choice(np.array([0, 1, 3, 512, 1020, 1023], dtype=np.uint16), size=((72 >> s), (136 >> s))) for s in (0, 1, 1)]
    rec = [p.copy() for p in pre]
    org = [np.clip(p.astype(np.int32) + rng.integers(-9, 10, size=p.shape), 0, 1023).astype(np.uint16) for p in rec]
    for plane in range(3):
        a = orc.ccso_derive_plane(g, plane, pre[0], rec[plane], org[plane], 20)
        b = ref.ccso_derive_plane(g, plane, pre[0], rec[plane], org[plane], 20)
        assert H.ccso_plane_sig(a[0], plane) == H.ccso_plane_sig(b[0], plane), f"plane {plane}"
        assert np.array_equal(a[1], b[1])
