"""Loads tests/golden/*.npz (written by oracle/make_golden.py from the real reference) back into
the parameter structs."""
import glob
import os

import numpy as np

import harness as H

S = H.S


def golden_files():
    return sorted(glob.glob(os.path.join(H.GOLDEN, "*.npz")))


def load(path):
    z = np.load(path)
    w, h, bd, sb = [int(v) for v in z["geom"]]
    g = S.make_geom(w, h, bd, sb)
    lfv = [int(v) for v in z["lf"]]
    lf = S.lf_params((lfv[0], lfv[1]), lfv[2], lfv[3], lfv[4], lfv[5], tuple(lfv[6:14]), tuple(lfv[14:16]))
    cv = [int(v) for v in z["cdef"]]
    cdef = S.cdef_params(cv[0], cv[1:9], cv[9:17])
    c = [int(v) for v in z["ccso"]]
    ccso = S.ccso_params(c[0:3], c[3:6], c[6:9], c[9:12], c[12:15], c[15:18], z["ccso_lut"])
    sub, bq, lam = [int(v) for v in z["search"]]
    first, second = H.cdef_level1_lists()
    cfg = S.cdef_search_cfg(first, second, subsampling_factor=sub, base_q_idx=bq, lambda_=lam)
    soa = {k[4:]: z[k] for k in z.files if k.startswith("soa_")}
    pl = lambda n: [np.ascontiguousarray(z[f"{n}_{c}"]) for c in "yuv"]  # noqa: E731
    return dict(geom=g, rec=pl("rec"), src=pl("src"), soa=soa, lf=lf, cdef=cdef, fb_idx=z["fb_idx"], ccso=ccso, cfg=cfg,
                dbk=pl("dbk"), cdef_out=pl("cdef"), ccso_out=pl("ccso"), mse=z["mse"], skip=z["skip"], dirs=z["dirs"], var=z["var"],
                pick=[int(v) for v in z["pick"]], pick_idx=z["pick_idx"])


def pick_tuple(params):
    return [params.cdef_bits, params.cdef_damping, *list(params.cdef_y_strength)[:8], *list(params.cdef_uv_strength)[:8]]
