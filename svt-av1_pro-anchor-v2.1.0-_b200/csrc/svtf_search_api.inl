// svtf_search_api.inl: the CDEF strength search and pick entry points (svtf_cdef_search_frame, svtf_cdef_pick, svtf_cdef_pick_table) — textually included by svtf_api.cu (same translation unit: the context struct and the helpers above it are in scope).

// ---- CDEF strength search ------------------------------------------------------------------------
extern "C" int svtf_cdef_search_frame(svtf_ctx *c, const svtf_cdef_search_cfg *cfg, uint64_t *mse, uint8_t *skip, uint8_t *dir,
                                      int32_t *var) {
    if (!c)
        return set_global_error(SVTF_E_ARG, "svtf_cdef_search_frame: NULL context");
    if (!cfg)
        return c->fail(SVTF_E_ARG, "svtf_cdef_search_frame: NULL cfg");
    if (int r = need_frame_mi(c, "svtf_cdef_search_frame"))
        return r;
    if (!c->have_source)
        return c->fail(SVTF_E_STATE, "svtf_cdef_search_frame: upload the source picture first");
    const int n = cfg->first_pass_fs_num + cfg->default_second_pass_fs_num;
    if (n <= 0 || n > SVTF_TOTAL_STRENGTHS || cfg->first_pass_fs_num < 0 || cfg->default_second_pass_fs_num < 0 ||
        (cfg->subsampling_factor != 1 && cfg->subsampling_factor != 2 && cfg->subsampling_factor != 4))
        return c->fail(SVTF_E_ARG, "svtf_cdef_search_frame: bad strength counts or subsampling factor");
    CUDA_TRY(c, cudaSetDevice(c->device));
    const int nvfb = (c->geom.mi_rows + 15) / 16, nhfb = (c->geom.mi_cols + 15) / 16, nfb = nvfb * nhfb;
    SearchArgs a;
    a.rec = c->surf[c->cur].planes();
    a.src = c->source.planes();
    a.mi = c->d_mi, a.mi_stride = c->mi_stride;
    a.d = dims_of(c->geom);
    a.n = n;
    for (int gi = 0; gi < n; gi++) {
        const bool first = gi < cfg->first_pass_fs_num;
        a.fs[gi]    = first ? cfg->default_first_pass_fs[gi] : cfg->default_second_pass_fs[gi - cfg->first_pass_fs_num];
        a.fs_uv[gi] = first ? cfg->default_first_pass_fs_uv[gi] : cfg->default_second_pass_fs_uv[gi - cfg->first_pass_fs_num];
    }
    a.damping = 3 + (cfg->base_q_idx >> 6);
    a.subsampling = cfg->subsampling_factor;
    a.mse = c->d_mse, a.skip = c->d_skip, a.dir = c->d_dir, a.var = c->d_var;
    a.nfb  = nfb;
    a.fbr0 = c->own_lo / 64;
    a.raw  = c->d_mse_raw;
    a.fold = reinterpret_cast<unsigned *>(c->d_mse_raw + (size_t)3 * 64 * nfb);
    a.fbr_par = (c->glob_row0 / 64) & 1;
    CUDA_TRY(c, cudaMemsetAsync(c->d_mse_raw, 0, (sizeof(unsigned long long) * 3 * 64 + sizeof(unsigned)) * (size_t)nfb, c->stream));
    CUDA_TRY(c, cudaMemsetAsync(c->d_skip, 0, (size_t)nfb, c->stream)); // quadrants folded into a 128 block are never written (as in the reference)
    CUDA_TRY(c, cudaMemsetAsync(c->d_dir, 0, 64 * (size_t)nfb, c->stream));
    CUDA_TRY(c, cudaMemsetAsync(c->d_var, 0, sizeof(int32_t) * 64 * (size_t)nfb, c->stream));
    CUDA_TRY(c, cudaMemsetAsync(c->d_mse, 0, sizeof(uint64_t) * 128 * (size_t)nfb, c->stream));
    dim3 grid(nhfb, (c->own_hi + 63) / 64 - a.fbr0);
    // 8/10-bit: separable packed kernel (svtf_search2.cuh); 12-bit (filtered-minus-input does not fit int8) and the
    // SVTF_SEARCH_SCALAR=1 A/B switch: the direct per-strength kernel
    static const bool scalar = getenv("SVTF_SEARCH_SCALAR") != nullptr;
    if (scalar || c->geom.bit_depth > 10) {
        if (c->geom.sample_bytes == 1)
            cdef_search_kernel<uint8_t><<<grid, SEARCH_THREADS, 0, c->stream>>>(a);
        else
            cdef_search_kernel<uint16_t><<<grid, SEARCH_THREADS, 0, c->stream>>>(a);
    } else {
        Search2Args b;
        b.s = a, b.present_y = 0, b.present_uv = 0;
        for (int gi = 0; gi < n; gi++) {
            b.present_y |= 1ull << (a.fs[gi] & 63);
            if (a.fs_uv[gi] != -1)
                b.present_uv |= 1ull << (a.fs[gi] & 63);
        }
#ifdef SVTF_AB_KERNELS
        static const bool use_v2 = getenv("SVTF_SEARCH_V2") != nullptr; // A/B switch: packed-integer taps (svtf_search2.cuh)
        if (use_v2) {
            if (c->geom.sample_bytes == 1)
                cdef_search2_kernel<uint8_t><<<grid, S2_THREADS, 0, c->stream>>>(b);
            else
                cdef_search2_kernel<uint16_t><<<grid, S2_THREADS, 0, c->stream>>>(b);
        } else
#endif
        { // fp16 taps (svtf_search3.cuh); search damping = 3 + (base_q_idx >> 6) <= 6, inside the exactness bound
            constexpr int smem = S3_SMEM_WORDS * 4;
            if (!c->search3_ready) {
                CUDA_TRY(c, cudaFuncSetAttribute(cdef_search3_kernel<uint8_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
                CUDA_TRY(c, cudaFuncSetAttribute(cdef_search3_kernel<uint16_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
                c->search3_ready = true;
            }
            if (c->geom.sample_bytes == 1)
                cdef_search3_kernel<uint8_t><<<grid, S2_THREADS, smem, c->stream>>>(b);
            else
                cdef_search3_kernel<uint16_t><<<grid, S2_THREADS, smem, c->stream>>>(b);
        }
    }
    CUDA_TRY(c, cudaGetLastError());
    c->launches++;
    {
        SearchArgs f = a;
        f.fbr0       = 0;
        search_fold_kernel<<<nfb, 64, 0, c->stream>>>(f); // rows of 128-wide / 128-high coding blocks (returns at once where there are none)
        CUDA_TRY(c, cudaGetLastError());
        c->launches++;
    }
    c->have_search = true;
    // outputs are WHOLE-PICTURE arrays; this context fills the filter-block rows it owns
    const size_t l0 = (size_t)a.fbr0 * nhfb, ln = (size_t)grid.y * nhfb;           // local first fb, count
    const size_t g0 = l0 + (size_t)(c->glob_row0 / 64) * nhfb;                      // same fb in the picture
    const size_t nfb_full = (size_t)((c->full.mi_rows + 15) / 16) * nhfb;
    if (mse)
        for (int pl = 0; pl < 2; pl++)
            CUDA_TRY(c, cudaMemcpyAsync(mse + ((size_t)pl * nfb_full + g0) * 64, c->d_mse + ((size_t)pl * nfb + l0) * 64, sizeof(uint64_t) * 64 * ln,
                                        cudaMemcpyDeviceToHost, c->stream));
    if (skip) CUDA_TRY(c, cudaMemcpyAsync(skip + g0, c->d_skip + l0, ln, cudaMemcpyDeviceToHost, c->stream));
    if (dir) CUDA_TRY(c, cudaMemcpyAsync(dir + g0 * 64, c->d_dir + l0 * 64, 64 * ln, cudaMemcpyDeviceToHost, c->stream));
    if (var) CUDA_TRY(c, cudaMemcpyAsync(var + g0 * 64, c->d_var + l0 * 64, sizeof(int32_t) * 64 * ln, cudaMemcpyDeviceToHost, c->stream));
    if (mse || skip || dir || var)
        CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    return SVTF_OK;
}

// finish_cdef_search on the device (svtf_pick.cuh) over a table that is already in device memory.
static int pick_device(svtf_ctx *c, const svtf_cdef_search_cfg *cfg, const uint64_t *d_mse, const uint8_t *d_skip, int nfb,
                       svtf_cdef_params *out, int8_t *fb_strength_idx, const char *who) {
    const int end_gi = cfg->first_pass_fs_num + cfg->default_second_pass_fs_num;
    if (end_gi <= 0 || end_gi > SVTF_TOTAL_STRENGTHS || cfg->first_pass_fs_num < 0 || cfg->default_second_pass_fs_num < 0)
        return c->fail(SVTF_E_ARG, std::string(who) + ": bad strength counts");
    if (nfb > c->pick_cap) {
        cudaFree(c->d_pick_idx);
        c->d_pick_idx = nullptr, c->pick_cap = 0;
        CUDA_TRY(c, cudaMalloc(&c->d_pick_idx, nfb));
        c->pick_cap = nfb;
    }
    PickArgs a;
    a.mse = d_mse, a.skip = d_skip, a.nfb = nfb, a.end_gi = end_gi, a.bias = cfg->zero_fs_cost_bias;
    a.st = c->d_pick_state, a.tot = c->d_pick_tot;
    // The 75 dependent svt_search_one_dual passes.  Per-pass form: one launch per pass, enqueued back to back (svtf_pick.cuh).  SVTF_PICK_CLUSTER=1: ONE
    // launch, a 16-CTA cluster stays resident for all passes (svtf_pick2.cuh; measured slower in throughput, see its header) when the table
    // fits its shared memory and is narrow (every entry < 2^31; the kernel reports wide == 2 otherwise and the per-pass form runs).
    static const bool per_pass = getenv("SVTF_PICK_CLUSTER") == nullptr;
    if (!c->max_smem_optin) {
        int v = 0;
        CUDA_TRY(c, cudaDeviceGetAttribute(&v, cudaDevAttrMaxSharedMemoryPerBlockOptin, c->device));
        c->max_smem_optin = (size_t)v;
    }
    // Default: ONE cooperative launch over the whole device (svtf_pick3.cuh) when its grid fits the device at once (up to ~9 400 filter
    // blocks: 8K included) and the table is narrow; else, or with SVTF_PICK_GRID=0, the per-pass launches.  Measured, 4K 10-bit,
    // 64 strengths: pick 0.635 -> 0.463 ms; search + pick one picture at a time 1.87 -> 2.21 M filter blocks/s, 8 in flight 2.82 -> 2.99 M.
    static const bool grid_form = getenv("SVTF_PICK_GRID") == nullptr || atoi(getenv("SVTF_PICK_GRID")) != 0;
    PickState st;
    bool      done = false;
    const int rg = per_pass && grid_form ? pick3_fits(nfb, end_gi, c->sm_count) : 0;
    if (rg) {
        const cudaError_t e = pick3_enqueue(a, rg, c->d_pick_idx, (int64_t)cfg->lambda, c->stream, &c->launches);
        if (e == cudaErrorCooperativeLaunchTooLarge || e == cudaErrorNotSupported) {
            (void)cudaGetLastError(); // fewer SMs than the occupancy query promised (a partitioned device): the per-pass form runs
        } else {
            CUDA_TRY(c, e);
            CUDA_TRY(c, cudaMemcpyAsync(&st, c->d_pick_state, sizeof(st), cudaMemcpyDeviceToHost, c->stream));
            CUDA_TRY(c, cudaMemcpyAsync(fb_strength_idx, c->d_pick_idx, nfb, cudaMemcpyDeviceToHost, c->stream));
            CUDA_TRY(c, cudaStreamSynchronize(c->stream));
            done = (st.wide & 3) == 0;
        }
    } else if (!per_pass && pick2_fits(nfb, c->max_smem_optin)) {
        CUDA_TRY(c, pick2_enqueue(a, c->d_pick_idx, (int64_t)cfg->lambda, c->stream, &c->launches));
        CUDA_TRY(c, cudaMemcpyAsync(&st, c->d_pick_state, sizeof(st), cudaMemcpyDeviceToHost, c->stream));
        CUDA_TRY(c, cudaMemcpyAsync(fb_strength_idx, c->d_pick_idx, nfb, cudaMemcpyDeviceToHost, c->stream));
        CUDA_TRY(c, cudaStreamSynchronize(c->stream));
        done = st.wide != 2;
    }
    if (!done) {
        // (Replaying the schedule as one captured graph was measured: 3 % faster one picture at a time, 30 % slower with 8 pictures in
        //  flight — graph launches from several streams do not interleave like plain launches — so the launches stay plain.)
        CUDA_TRY(c, pick_prepare());
        CUDA_TRY(c, pick_enqueue(a, c->d_pick_idx, (int64_t)cfg->lambda, c->sm_count, c->stream, &c->launches));
        CUDA_TRY(c, cudaMemcpyAsync(&st, c->d_pick_state, sizeof(st), cudaMemcpyDeviceToHost, c->stream));
        CUDA_TRY(c, cudaMemcpyAsync(fb_strength_idx, c->d_pick_idx, nfb, cudaMemcpyDeviceToHost, c->stream));
        CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    }
    memset(out, 0, sizeof(*out));
    out->cdef_bits = st.best_bits;
    auto map = [&](int i) { // index -> strength byte through the luma lists (Encoder/Codec/EbEncCdef.c:911-919)
        return i < cfg->first_pass_fs_num ? cfg->default_first_pass_fs[i] : cfg->default_second_pass_fs[i - cfg->first_pass_fs_num];
    };
    for (int g = 0; g < (1 << st.best_bits); g++) {
        out->cdef_y_strength[g]  = map(st.out0[g]);
        out->cdef_uv_strength[g] = map(st.out1[g]);
    }
    out->cdef_damping = 3 + (cfg->base_q_idx >> 6);
    return SVTF_OK;
}

extern "C" int svtf_cdef_pick(svtf_ctx *c, const svtf_cdef_search_cfg *cfg, svtf_cdef_params *out_params, int8_t *fb_strength_idx) {
    if (!c)
        return set_global_error(SVTF_E_ARG, "svtf_cdef_pick: NULL context");
    if (!cfg || !out_params || !fb_strength_idx)
        return c->fail(SVTF_E_ARG, "svtf_cdef_pick: NULL argument");
    if (!c->have_search)
        return c->fail(SVTF_E_STATE, "svtf_cdef_pick: run svtf_cdef_search_frame first");
    if (c->striped)
        return c->fail(SVTF_E_STATE, "svtf_cdef_pick: this context holds one stripe; gather mse/skip and call svtf_cdef_pick_table");
    CUDA_TRY(c, cudaSetDevice(c->device));
    const int nfb = ((c->geom.mi_rows + 15) / 16) * ((c->geom.mi_cols + 15) / 16);
    return pick_device(c, cfg, c->d_mse, c->d_skip, nfb, out_params, fb_strength_idx, "svtf_cdef_pick");
}

extern "C" int svtf_cdef_pick_table(svtf_ctx *c, const svtf_cdef_search_cfg *cfg, const uint64_t *mse, const uint8_t *skip, int32_t nfb,
                                    svtf_cdef_params *out_params, int8_t *fb_strength_idx) {
    if (!c)
        return set_global_error(SVTF_E_ARG, "svtf_cdef_pick_table: NULL context");
    if (!cfg || !mse || !skip || !out_params || !fb_strength_idx || nfb <= 0)
        return c->fail(SVTF_E_ARG, "svtf_cdef_pick_table: bad argument");
    CUDA_TRY(c, cudaSetDevice(c->device));
    if (nfb > c->tab_cap) {
        cudaFree(c->d_tab_mse), cudaFree(c->d_tab_skip);
        c->d_tab_mse = nullptr, c->d_tab_skip = nullptr, c->tab_cap = 0;
        CUDA_TRY(c, cudaMalloc(&c->d_tab_mse, sizeof(uint64_t) * 128 * (size_t)nfb));
        CUDA_TRY(c, cudaMalloc(&c->d_tab_skip, nfb));
        c->tab_cap = nfb;
    }
    CUDA_TRY(c, cudaMemcpyAsync(c->d_tab_mse, mse, sizeof(uint64_t) * 128 * (size_t)nfb, cudaMemcpyHostToDevice, c->stream));
    CUDA_TRY(c, cudaMemcpyAsync(c->d_tab_skip, skip, nfb, cudaMemcpyHostToDevice, c->stream));
    return pick_device(c, cfg, c->d_tab_mse, c->d_tab_skip, nfb, out_params, fb_strength_idx, "svtf_cdef_pick_table");
}
