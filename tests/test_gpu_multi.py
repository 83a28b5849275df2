"""GPU (needs >= 2 GPUs; skipped on one): ONE picture split across GPUs by superblock-row stripes with the two inter-GPU steps done
device to device through NCCL inside the library — svtf_stripe_exchange (halo rows from the neighbours' memory: every rank uploads ONLY
the rows it owns) and svtf_cdef_search_gather_pick (ragged all-gather of the mse / skip rows + the pick on every rank).  One thread per
GPU in this process (the bench uses one process per GPU; NCCL does not care).  Results must equal the single-GPU whole-picture run."""
import threading

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ngpus():
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


def _run_ranks(world, fn):
    errs, th = [], []

    def wrap(r):
        try:
            fn(r)
        except BaseException as e:  # noqa: BLE001 - reported below
            errs.append((r, e))

    for r in range(world):
        t = threading.Thread(target=wrap, args=(r,))
        t.start()
        th.append(t)
    for t in th:
        t.join(timeout=300)
    assert not errs, errs
    assert all(not t.is_alive() for t in th), "a rank hung"


@pytest.mark.parametrize("world", [2, 4])
@pytest.mark.parametrize("w,h,bd", [(1920, 1080, 10), (840, 520, 8)])
def test_stripes_with_nccl_halo_exchange_equal_whole_picture(H, world, w, h, bd):
    if _ngpus() < world:
        pytest.skip(f"needs {world} GPUs")
    S = H.S
    g = S.make_geom(w, h, bd)
    rec = H.synth_picture(g, 7100 + w)
    mi = H.pack_mi(H.synth_mode_info(g, 7101 + w, segments=True))
    lf = S.lf_params((26, 20), 13, 15, sharpness_level=1)
    cdef, idx = H.synth_cdef(g, 7102 + w)
    ccso, flags = H.synth_ccso(g, 7103 + w, support=(1, 3, 5))
    with S.FilterContext(g) as ctx:
        ctx.begin(g)
        ctx.upload(rec)
        ctx.set_mi(mi)
        ctx.filter_chain(lf, cdef, idx, ccso, flags)
        whole = ctx.download()
    out = [np.zeros_like(p) for p in rec]
    uid = S.Comm.unique_id()
    parts = S.stripes.partition(S.stripes.sb_rows_of(g), world)

    def rank_fn(r):
        first, count = parts[r]
        comm = S.Comm(r, world, r, uid)
        try:
            with S.FilterContext(S.stripes.stripe_geom(g, first, count), device=r) as c:
                c.begin_stripe(g, first, count)
                c.upload_own(rec)      # no halo rows: they must come from the neighbours
                c.set_mi_own(mi)
                c.stripe_exchange(comm, r - 1 if r > 0 else -1, r + 1 if r + 1 < world else -1)
                c.filter_chain(lf, cdef, idx, ccso, flags)
                c.download(out)
        finally:
            comm.close()

    _run_ranks(world, rank_fn)
    assert H.planes_equal(out, whole), "stripes + halo exchange differ from the whole picture: " + H.first_diff(out, whole)


@pytest.mark.parametrize("world", [2, 3])
def test_search_gather_pick_equals_single_gpu(H, world):
    if _ngpus() < world:
        pytest.skip(f"needs {world} GPUs")
    S = H.S
    g = S.make_geom(1280, 720, 10)
    rec = H.synth_picture(g, 7200)
    src = H.synth_source(g, rec, 7201)
    mi = H.pack_mi(H.synth_mode_info(g, 7202))
    first_l, second_l = H.cdef_level1_lists()
    cfg = S.cdef_search_cfg(first_l, second_l, base_q_idx=130, lambda_=4000)
    with S.FilterContext(g) as ctx:
        ctx.begin(g)
        ctx.upload(rec)
        ctx.upload_source(src)
        ctx.set_mi(mi)
        ctx.cdef_seg_search(cfg, want_outputs=False)
        e_params, e_idx = ctx.finish_cdef_search(cfg)
    nv, nh = H.nfb_of(g)
    uid = S.Comm.unique_id()
    parts = S.stripes.partition(S.stripes.sb_rows_of(g), world)
    res = [None] * world

    def rank_fn(r):
        first, count = parts[r]
        comm = S.Comm(r, world, r, uid)
        try:
            with S.FilterContext(S.stripes.stripe_geom(g, first, count), device=r) as c:
                c.begin_stripe(g, first, count)
                c.upload_own(rec)
                c.set_mi_own(mi)
                c.stripe_exchange(comm, r - 1 if r > 0 else -1, r + 1 if r + 1 < world else -1)
                c.upload_source(src)
                c.cdef_seg_search(cfg, want_outputs=False)
                res[r] = c.search_gather_pick(comm, cfg, parts, nv * nh)
        finally:
            comm.close()

    _run_ranks(world, rank_fn)
    sig = lambda p: [p.cdef_bits, p.cdef_damping, *p.cdef_y_strength, *p.cdef_uv_strength]
    for r in range(world):
        assert sig(res[r][0]) == sig(e_params), r
        assert np.array_equal(res[r][1], e_idx), r
