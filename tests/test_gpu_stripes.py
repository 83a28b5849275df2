"""Single-picture stripe split (SURVEY 8e, config C4 at test size): N striped contexts — here all on one GPU, one
after the other, exactly the calls N ranks would make — must reproduce the whole-picture result bit for bit, for the
apply chain (deblock -> CDEF+CCSO) and for the CDEF strength search + pick."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _chain_inputs(H, g, seed):
    S = H.S
    rec = H.synth_picture(g, seed)
    soa = H.synth_mode_info(g, seed + 1)
    mi = H.pack_mi(soa)
    lf = S.lf_params((24, 24), 12, 12)
    cdef, idx = H.synth_cdef(g, seed + 2)
    ccso, flags = H.synth_ccso(g, seed + 3, support=(1, 3, 5), p_on=0.7)
    return rec, mi, lf, cdef, idx, ccso, flags


@pytest.mark.parametrize("w,h,bd,sb", [(328, 456, 10, 2), (200, 520, 8, 1), (136, 330, 12, 2), (90, 200, 8, 2)])
@pytest.mark.parametrize("world", [2, 3, 5])
def test_striped_chain_equals_whole_picture(H, orc, w, h, bd, sb, world):
    S = H.S
    g = S.make_geom(w, h, bd, sb)
    rec, mi, lf, cdef, idx, ccso, flags = _chain_inputs(H, g, 5000 + w + bd)
    with S.FilterContext(g) as ctx:
        ctx.begin(g)
        ctx.upload(rec)
        ctx.set_mi(mi)
        ctx.filter_chain(lf, cdef, idx, ccso, flags)
        whole = ctx.download()
    e1 = orc.deblock(g, rec, mi, lf)
    e2 = orc.cdef_apply(g, e1, mi, cdef, idx)
    e3 = orc.ccso_apply(g, e2, e1[0], ccso, flags)
    assert H.planes_equal(whole, e3), H.first_diff(whole, e3)
    out = [np.zeros_like(p) for p in rec]
    for rank in range(world):
        with S.StripeFilter(g, rank, world) as sf:
            sf.chain(rec, mi, lf, cdef, idx, ccso, flags, out)
    assert H.planes_equal(out, whole), f"world {world}: " + H.first_diff(out, whole)


@pytest.mark.parametrize("world", [2, 4])
def test_striped_stages_separately(H, world):
    """Deblock only, CDEF only, CCSO only (standalone kernel, classifier = the picture) on stripes."""
    S = H.S
    g = S.make_geom(264, 392, 10)
    rec, mi, lf, cdef, idx, ccso, flags = _chain_inputs(H, g, 5100)
    for stage in ("deblock", "cdef", "ccso", "cdef+ccso_unfused"):
        def run(c):
            if stage == "deblock":
                c.loop_filter_frame(lf)
            elif stage == "cdef":
                c.cdef_frame(cdef, idx)
            elif stage == "ccso":
                c.ccso_frame(ccso, flags)
            else:
                c.cdef_frame(cdef, idx)
                c.ccso_frame(ccso, flags)

        with S.FilterContext(g) as ctx:
            ctx.begin(g)
            ctx.upload(rec)
            ctx.set_mi(mi)
            run(ctx)
            whole = ctx.download()
        out = [np.zeros_like(p) for p in rec]
        for rank in range(world):
            with S.StripeFilter(g, rank, world) as sf:
                c = sf.ctx
                c.begin_stripe(g, sf.first, sf.count)
                c.upload(rec)
                c.set_mi(mi)
                run(c)
                c.download(out)
        assert H.planes_equal(out, whole), f"{stage} world {world}: " + H.first_diff(out, whole)


@pytest.mark.parametrize("w,h,bd", [(200, 264, 10), (328, 200, 8)])
@pytest.mark.parametrize("world", [2, 3])
def test_striped_search_and_pick(H, orc, w, h, bd, world):
    S = H.S
    g = S.make_geom(w, h, bd)
    rec = H.synth_picture(g, 5200 + w)
    src = H.synth_source(g, rec, 5201)
    mi = H.pack_mi(H.synth_mode_info(g, 5202))
    lf = S.lf_params((20, 20), 10, 10)
    first, second = H.cdef_level1_lists()
    cfg = S.cdef_search_cfg(first[:8], second[:8], base_q_idx=120, lambda_=2500)
    dbk = orc.deblock(g, rec, mi, lf)
    e_mse, e_skip, e_dir, e_var = orc.cdef_search(g, dbk, src, mi, cfg)
    nfb = e_skip.size
    mse = np.zeros((2, nfb, 64), dtype=np.uint64)
    skip = np.zeros(nfb, dtype=np.uint8)
    dirs = np.zeros((nfb, 64), dtype=np.uint8)
    var = np.zeros((nfb, 64), dtype=np.int32)
    for rank in range(world):
        with S.StripeFilter(g, rank, world) as sf:
            if sf.ctx is None:
                continue
            c = sf.ctx
            c.begin_stripe(g, sf.first, sf.count)
            c.upload(rec)
            c.upload_source(src)
            c.set_mi(mi)
            c.loop_filter_frame(lf)
            c.cdef_seg_search(cfg, out=(mse, skip, dirs, var))
            with pytest.raises(S.SvtfError):
                c.finish_cdef_search(cfg)  # a stripe cannot pick alone
    assert np.array_equal(mse, e_mse) and np.array_equal(skip, e_skip)
    assert np.array_equal(dirs, e_dir) and np.array_equal(var, e_var)
    a = S.cdef_pick_host(cfg, mse, skip)
    b = orc.cdef_pick(cfg, e_mse, e_skip)
    assert a[0].cdef_bits == b[0].cdef_bits and list(a[0].cdef_y_strength) == list(b[0].cdef_y_strength)
    assert np.array_equal(a[1], b[1])
