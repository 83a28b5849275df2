#!/usr/bin/env python
"""Generates tests/golden/*.npz from the REAL reference (oracle/_ref/libsvtref.so, i.e. the
reference's own C sources compiled by oracle/Makefile).  Run in the container where /root/reference
is mounted:  `python oracle/make_golden.py`.  Each file stores the inputs and the reference outputs
of one small case, so the pin holds on machines without the reference tree or its library.

The reference ships no golden vectors for this path; these are "outputs of the reference
itself run here".
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import harness as H  # noqa: E402

S = H.S
OUT = os.path.join(ROOT, "tests", "golden")

CASES = [  # name, width, height, bit_depth, sample_bytes, seed
    ("chain_8bit_72x40", 72, 40, 8, 1, 11),
    ("chain_10bit_136x72", 136, 72, 10, 2, 12),
    ("chain_12bit_64x64", 64, 64, 12, 2, 13),
    ("chain_8in16_90x50", 90, 50, 8, 2, 14),
]


def params_for(geom, seed):
    """Plain-value description of the parameters (stored in the npz, rebuilt by tests/golden_io.py)."""
    rng = np.random.default_rng(seed)
    lf = dict(filter_level=(int(rng.integers(1, 64)), int(rng.integers(1, 64))), filter_level_u=int(rng.integers(0, 64)),
              filter_level_v=int(rng.integers(1, 64)), sharpness_level=int(rng.integers(0, 8)), mode_ref_delta_enabled=1,
              ref_deltas=(1, 0, 0, 0, -1, 0, -1, -1), mode_deltas=(0, int(rng.integers(0, 4))))
    n = 8
    cdef = dict(damping=int(rng.integers(3, 7)), y=[0] + [int(v) for v in rng.integers(0, 64, n - 1)],
                uv=[0] + [int(v) for v in rng.integers(0, 64, n - 1)])
    nv, nh = H.nfb_of(geom)
    idx = rng.integers(0, n, nv * nh).astype(np.int8)
    ccso = dict(enable=(1, 1, 1), bo_only=(0, 0, int(rng.integers(0, 2))), quant_idx=tuple(int(v) for v in rng.integers(0, 4, 3)),
                support=tuple(int(v) for v in rng.integers(0, 6, 3)), edge_clf=tuple(int(v) for v in rng.integers(0, 2, 3)),
                band_log2=tuple(int(v) for v in rng.integers(0, 4, 3)))
    lut = rng.choice(np.array([-10, -7, -3, -1, 0, 1, 3, 7], dtype=np.int8), size=(3, 2048))
    first, second = H.cdef_level1_lists()
    search = dict(first=first, second=second, sub=1, base_q_idx=int(rng.integers(0, 256)), lam=int(rng.integers(100, 20000)))
    return lf, cdef, idx, ccso, lut, search


def main():
    os.makedirs(OUT, exist_ok=True)
    ref = H.Ref()
    for name, w, h, bd, sb, seed in CASES:
        g = S.make_geom(w, h, bd, sb)
        src = H.synth_picture(g, seed, noise=1)
        rec = H.synth_source(g, src, seed + 1, amp=9)
        soa = H.synth_mode_info(g, seed + 2, p_skip=0.45, segments=True)
        lfd, cdefd, idx, ccsod, lut, sd = params_for(g, seed + 3)
        lf = S.lf_params(**lfd)
        cdef = S.cdef_params(cdefd["damping"], cdefd["y"], cdefd["uv"])
        ccso = S.ccso_params(ccsod["enable"], ccsod["bo_only"], ccsod["quant_idx"], ccsod["support"], ccsod["edge_clf"], ccsod["band_log2"], lut)
        d1 = ref.deblock(g, rec, soa, lf)
        d2 = ref.cdef_apply(g, d1, soa, cdef, idx)
        d3 = ref.ccso_apply(g, d2, d1[0], ccso, None)
        cfg = S.cdef_search_cfg(sd["first"], sd["second"], subsampling_factor=sd["sub"], base_q_idx=sd["base_q_idx"], lambda_=sd["lam"])
        mse, skip, dirs, var = ref.cdef_search(g, d1, src, soa, cfg)
        pk, pk_idx = ref.cdef_pick(g, soa, cfg, mse, skip)
        np.savez_compressed(
            os.path.join(OUT, name + ".npz"),
            geom=np.array([w, h, bd, sb]), rec_y=rec[0], rec_u=rec[1], rec_v=rec[2], src_y=src[0], src_u=src[1], src_v=src[2],
            **{"soa_" + k: v for k, v in soa.items()},
            lf=np.array([lfd["filter_level"][0], lfd["filter_level"][1], lfd["filter_level_u"], lfd["filter_level_v"], lfd["sharpness_level"],
                         lfd["mode_ref_delta_enabled"], *lfd["ref_deltas"], *lfd["mode_deltas"]]),
            cdef=np.array([cdefd["damping"], *cdefd["y"], *cdefd["uv"]]), fb_idx=idx,
            ccso=np.array([*ccsod["enable"], *ccsod["bo_only"], *ccsod["quant_idx"], *ccsod["support"], *ccsod["edge_clf"], *ccsod["band_log2"]]),
            ccso_lut=lut, search=np.array([sd["sub"], sd["base_q_idx"], sd["lam"]]),
            dbk_y=d1[0], dbk_u=d1[1], dbk_v=d1[2], cdef_y=d2[0], cdef_u=d2[1], cdef_v=d2[2], ccso_y=d3[0], ccso_u=d3[1], ccso_v=d3[2],
            mse=mse, skip=skip, dirs=dirs, var=var,
            pick=np.array([pk.cdef_bits, pk.cdef_damping, *list(pk.cdef_y_strength)[:8], *list(pk.cdef_uv_strength)[:8]]), pick_idx=pk_idx)
        print("wrote", name)


if __name__ == "__main__":
    main()
