"""CPU: the oracle restatement against the golden vectors produced by the real reference
(oracle/make_golden.py).  Runs everywhere, needs neither the reference tree nor a GPU."""
import numpy as np
import pytest

import golden_io as G


@pytest.mark.parametrize("path", G.golden_files(), ids=lambda p: p.split("/")[-1][:-4])
def test_oracle_reproduces_golden(H, orc, path):
    z = G.load(path)
    g, mi = z["geom"], H.pack_mi(z["soa"])
    d1 = orc.deblock(g, z["rec"], mi, z["lf"])
    assert H.planes_equal(d1, z["dbk"]), "deblock: " + H.first_diff(d1, z["dbk"])
    d2 = orc.cdef_apply(g, d1, mi, z["cdef"], z["fb_idx"])
    assert H.planes_equal(d2, z["cdef_out"]), "cdef: " + H.first_diff(d2, z["cdef_out"])
    d3 = orc.ccso_apply(g, d2, d1[0], z["ccso"], None)
    assert H.planes_equal(d3, z["ccso_out"]), "ccso: " + H.first_diff(d3, z["ccso_out"])
    mse, skip, dirs, var = orc.cdef_search(g, d1, z["src"], mi, z["cfg"])
    assert np.array_equal(mse, z["mse"]) and np.array_equal(skip, z["skip"])
    assert np.array_equal(dirs, z["dirs"]) and np.array_equal(var, z["var"])
    params, idx = orc.cdef_pick(z["cfg"], mse, skip)
    assert G.pick_tuple(params) == z["pick"] and np.array_equal(idx, z["pick_idx"])
    # host product logic (no GPU needed): the library's own pick must agree as well
    params2, idx2 = H.S.cdef_pick_host(z["cfg"], mse, skip)
    assert G.pick_tuple(params2) == z["pick"] and np.array_equal(idx2, z["pick_idx"])


def test_golden_present():
    assert len(G.golden_files()) >= 4
