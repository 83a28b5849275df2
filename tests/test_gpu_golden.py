"""GPU: the CUDA path against the reference's golden vectors."""
import numpy as np
import pytest

import golden_io as G

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("path", G.golden_files(), ids=lambda p: p.split("/")[-1][:-4])
def test_cuda_reproduces_golden(H, path):
    S = H.S
    z = G.load(path)
    g, mi = z["geom"], H.pack_mi(z["soa"])
    with S.FilterContext(g) as ctx:
        ctx.begin(g)
        ctx.upload(z["rec"])
        ctx.upload_source(z["src"])
        ctx.set_mi(mi)
        ctx.loop_filter_frame(z["lf"])
        d1 = ctx.download()
        assert H.planes_equal(d1, z["dbk"]), "deblock: " + H.first_diff(d1, z["dbk"])
        mse, skip, dirs, var = ctx.cdef_seg_search(z["cfg"])
        assert np.array_equal(mse, z["mse"]) and np.array_equal(skip, z["skip"])
        assert np.array_equal(dirs, z["dirs"]) and np.array_equal(var, z["var"])
        params, idx = ctx.finish_cdef_search(z["cfg"])
        assert G.pick_tuple(params) == z["pick"] and np.array_equal(idx, z["pick_idx"])
        ctx.cdef_frame(z["cdef"], z["fb_idx"])
        d2 = ctx.download()
        assert H.planes_equal(d2, z["cdef_out"]), "cdef: " + H.first_diff(d2, z["cdef_out"])
        ctx.ccso_frame(z["ccso"], None)
        d3 = ctx.download()
        assert H.planes_equal(d3, z["ccso_out"]), "ccso: " + H.first_diff(d3, z["ccso_out"])
