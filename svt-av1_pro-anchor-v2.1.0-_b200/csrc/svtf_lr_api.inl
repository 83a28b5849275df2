// svtf_lr_api.inl: loop restoration apply (svtf_lr_apply_frame) — textually included by svtf_api.cu (same translation unit: the context struct and the helpers above it are in scope).

// ---- deblocking level search -------------------------------------------------------------------------
// ---- loop restoration (svtf_lr.cuh) --------------------------------------------------------------------
extern "C" int svtf_lr_apply_frame(svtf_ctx *c, const svtf_lr_params *lr, const svtf_lr_unit *const units[3]) {
    if (!c)
        return set_global_error(SVTF_E_ARG, "svtf_lr_apply_frame: NULL context");
    if (!lr || !units)
        return c->fail(SVTF_E_ARG, "svtf_lr_apply_frame: NULL argument");
    if (!c->have_frame)
        return c->fail(SVTF_E_STATE, "svtf_lr_apply_frame: upload a picture first");
    if (c->striped)
        return c->fail(SVTF_E_STATE, "svtf_lr_apply_frame: not available on a striped context");
    bool any = false;
    int  hu[3], vu[3], first[3], total = 0;
    for (int p = 0; p < 3; p++) {
        const int us = lr->restoration_unit_size[p], ft = lr->frame_restoration_type[p];
        hu[p] = vu[p] = 0, first[p] = total;
        if (ft == SVTF_RESTORE_NONE)
            continue;
        if ((us != 32 && us != 64 && us != 128 && us != 256) || (p == 0 && us == 32) || !units[p])
            return c->fail(SVTF_E_ARG, "svtf_lr_apply_frame: bad restoration unit size or NULL unit array");
        svtf_lr_units(&c->geom, p, us, &hu[p], &vu[p]);
        for (int u = 0; u < hu[p] * vu[p]; u++) {
            const svtf_lr_unit &x = units[p][u];
            if (x.restoration_type < 0 || x.restoration_type > SVTF_RESTORE_SGRPROJ ||
                (x.restoration_type == SVTF_RESTORE_SGRPROJ && (x.sgr_ep < 0 || x.sgr_ep > 15)))
                return c->fail(SVTF_E_ARG, "svtf_lr_apply_frame: bad restoration unit");
        }
        total += hu[p] * vu[p];
        any = true;
    }
    if (!any)
        return SVTF_OK; // nothing to restore: the picture stays where it is
    CUDA_TRY(c, cudaSetDevice(c->device));
    if (!c->surf[3].base[0]) {
        cudaError_t e = alloc_surface(c->surf[3], c->max_geom);
        if (e != cudaSuccess)
            return c->cuda_fail(e, "cudaMalloc(restoration surface)");
    }
    if (c->lr_units_cap < (size_t)total) {
        if (c->d_lr_units)
            cudaFree(c->d_lr_units);
        c->d_lr_units = nullptr, c->lr_units_cap = 0;
        CUDA_TRY(c, cudaMalloc(&c->d_lr_units, sizeof(LrUnitDev) * total));
        c->lr_units_cap = total;
    }
    std::vector<LrUnitDev> host(total);
    for (int p = 0; p < 3; p++)
        for (int u = 0; u < hu[p] * vu[p]; u++) {
            const svtf_lr_unit &x = units[p][u];
            LrUnitDev          &d = host[first[p] + u];
            d.type = x.restoration_type, d.ep = x.sgr_ep, d.xqd[0] = x.sgr_xqd[0], d.xqd[1] = x.sgr_xqd[1];
            for (int t = 0; t < 8; t++) d.vf[t] = x.wiener_vfilter[t], d.hf[t] = x.wiener_hfilter[t];
        }
    CUDA_TRY(c, cudaMemcpyAsync(c->d_lr_units, host.data(), sizeof(LrUnitDev) * total, cudaMemcpyHostToDevice, c->stream));
    CUDA_TRY(c, cudaStreamSynchronize(c->stream)); // `host` goes out of scope
    LrArgs a;
    a.cur = c->surf[c->cur].planes();
    a.dbk = c->surf[c->pre_cdef >= 0 ? c->pre_cdef : c->cur].planes();
    a.out = c->surf[3].planes();
    if (c->cur == 3)
        return c->fail(SVTF_E_STATE, "svtf_lr_apply_frame: the current picture is already a restoration output");
    int grid = 0;
    for (int p = 0; p < 3; p++) {
        const int ss = p ? 1 : 0;
        a.pw[p] = (c->geom.width + ss) >> ss, a.ph[p] = (c->geom.height + ss) >> ss;
        a.cw[p] = (c->geom.mi_cols * 4) >> ss, a.chh[p] = (c->geom.mi_rows * 4) >> ss;
        a.unit[p] = lr->restoration_unit_size[p], a.hu[p] = hu[p], a.vu[p] = vu[p];
        a.units[p]      = c->d_lr_units + first[p];
        a.frame_type[p] = lr->frame_restoration_type[p];
        const int sh = 64 >> ss, off = 8 >> ss;
        a.chunks[p]  = (a.cw[p] + (LR_MAXW >> ss) - 1) / (LR_MAXW >> ss);
        a.stripes[p] = (a.ph[p] + off + sh - 1) / sh + 1; // + the tail below the visible picture
        a.base[p]    = grid;
        grid += a.chunks[p] * a.stripes[p];
    }
    a.opt = lr->optimized_lr != 0, a.bd = c->geom.bit_depth;
    if (c->geom.sample_bytes == 1)
        lr_kernel<uint8_t><<<grid, LR_THREADS, 0, c->stream>>>(a);
    else
        lr_kernel<uint16_t><<<grid, LR_THREADS, 0, c->stream>>>(a);
    CUDA_TRY(c, cudaGetLastError());
    c->launches++;
    c->cur = 3;
    return SVTF_OK;
}
