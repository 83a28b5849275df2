// svtf_lf_search_api.inl: the deblocking level search (svtf_lf_search_frame) — textually included by svtf_api.cu (same translation unit).

// svt_av1_pick_filter_level(LPF_PICK_FROM_FULL_IMAGE) (Encoder/Codec/EbDeblockingFilter.c:1129-1252):
// search_filter_level (:886-985) per plane with try_filter_frame (:841-880) = deblock one plane with a trial level,
// picture_sse_calculations (:716-835) against the source, restore.  On the device a trial is ONE launch: the deblock kernel's
// SSE instantiation filters that plane's tiles in shared memory and measures them against the source instead of writing
// them; nothing is restored because the unfiltered picture is never overwritten.
// search_filter_level (Encoder/Codec/EbDeblockingFilter.c:886-985) for one plane as a state machine (dir 2 = both luma
// directions, as svt_av1_pick_filter_level calls it): `wanted()` names the levels the next step of the reference's loop
// would try (filt_low and filt_high are tried independently of each other's result), `advance()` is the loop body once
// their errors are known.  The three planes' searches are independent of one another (a plane's trial filters that plane
// alone, with that plane's level), so svtf_lf_search_frame runs them in lock-step: one batch of trials = up to six
// deblock+SSE launches and ONE host round trip, instead of one round trip per trial.
struct LfPlaneSearch {
    int     plane, filt_mid, step, filt_best, direction = 0, convergence = 0;
    int64_t best_err = 0, ss_err[64];
    bool    started = false;
    void init(int p, int lvl) {
        plane    = p;
        filt_mid = std::min(std::max(lvl, 0), 63); // MAX_LOOP_FILTER
        step     = filt_mid < 16 ? 4 : filt_mid / 4;
        for (int i = 0; i < 64; i++) ss_err[i] = -1;
        filt_best = filt_mid;
    }
    bool done() const { return started && step <= 0; }
    int  low() const { return std::max(filt_mid - step, 0); }
    int  high() const { return std::min(filt_mid + step, 63); }
    int  wanted(int lv[2]) const {
        int n = 0;
        if (!started) {
            lv[n++] = filt_mid;
            return n;
        }
        if (direction <= 0 && low() != filt_mid && ss_err[low()] < 0)
            lv[n++] = low();
        if (direction >= 0 && high() != filt_mid && ss_err[high()] < 0 && !(n && lv[0] == high()))
            lv[n++] = high();
        return n;
    }
    void advance(const svtf_lf_search_cfg *cfg) {
        if (!started) {
            started  = true;
            best_err = ss_err[filt_mid];
            return;
        }
        const int filt_high = high(), filt_low = low();
        int64_t   bias      = (best_err >> (15 - (filt_mid / 8))) * step; // bias against raising the level
        if (!cfg->only_4x4)
            bias >>= 1;
        if (direction <= 0 && filt_low != filt_mid) {
            if (ss_err[filt_low] < best_err + bias) {
                if (ss_err[filt_low] < best_err)
                    best_err = ss_err[filt_low];
                filt_best = filt_low;
            }
        }
        if (direction >= 0 && filt_high != filt_mid) {
            if (ss_err[filt_high] < best_err - bias) {
                best_err  = ss_err[filt_high];
                filt_best = filt_high;
            }
        }
        if (filt_best == filt_mid) {
            convergence++;
            if (convergence == cfg->early_exit_convergence)
                step = 0;
            else
                step /= 2;
            direction = 0;
        } else {
            direction = filt_best < filt_mid ? -1 : 1;
            filt_mid  = filt_best;
        }
    }
};

extern "C" int svtf_lf_search_frame(svtf_ctx *c, const svtf_lf_params *start, const svtf_lf_search_cfg *cfg, int32_t out_levels[4]) {
    if (!c)
        return set_global_error(SVTF_E_ARG, "svtf_lf_search_frame: NULL context");
    if (!start || !cfg || !out_levels)
        return c->fail(SVTF_E_ARG, "svtf_lf_search_frame: NULL argument");
    if (int r = need_frame_mi(c, "svtf_lf_search_frame"))
        return r;
    if (!c->have_source)
        return c->fail(SVTF_E_STATE, "svtf_lf_search_frame: upload the source picture first");
    if (c->striped)
        return c->fail(SVTF_E_STATE, "svtf_lf_search_frame: not available on a striped context");
    CUDA_TRY(c, cudaSetDevice(c->device));
    const int32_t last[4] = {start->filter_level[0], start->filter_level[1], start->filter_level_u, start->filter_level_v};
    int32_t       y = 0, u = last[2], v = last[3];
    LfPlaneSearch ps[3];
    // luma is searched with dir = 2 and the start level is last_frame_filter_level[dir] unless dlf_avg: index 2 is the U
    // level (:903-909) — kept as the reference has it
    ps[0].init(0, cfg->dlf_avg ? last[0] : last[2]);
    ps[1].init(1, last[2]);
    ps[2].init(2, last[3]);
    const int           nplanes = cfg->search_chroma ? 3 : 1;
    unsigned long long *d_out   = reinterpret_cast<unsigned long long *>(c->d_pick_tot); // 48 bytes of an idle scratch buffer
    for (;;) {
        int trial_plane[6], trial_level[6], nt = 0;
        for (int p = 0; p < nplanes; p++) {
            if (ps[p].done())
                continue;
            int       lv[2];
            const int n = ps[p].wanted(lv);
            for (int k = 0; k < n; k++) trial_plane[nt] = p, trial_level[nt] = lv[k], nt++;
        }
        if (nt) {
            CUDA_TRY(c, cudaMemsetAsync(d_out, 0, sizeof(*d_out) * 6, c->stream));
            for (int t = 0; t < nt; t++) { // try_filter_frame (:841-880): this plane, this level
                svtf_lf_params lf = *start;
                if (trial_plane[t] == 0)
                    lf.filter_level[0] = lf.filter_level[1] = trial_level[t];
                else if (trial_plane[t] == 1)
                    lf.filter_level_u = trial_level[t];
                else
                    lf.filter_level_v = trial_level[t];
                if (int r = deblock_impl(c, &lf, trial_plane[t], trial_plane[t] + 1, true, nullptr, 0, 0, d_out + t))
                    return r;
            }
            unsigned long long res[6] = {0, 0, 0, 0, 0, 0};
            CUDA_TRY(c, cudaMemcpyAsync(res, d_out, sizeof(*d_out) * 6, cudaMemcpyDeviceToHost, c->stream));
            CUDA_TRY(c, cudaStreamSynchronize(c->stream));
            for (int t = 0; t < nt; t++) ps[trial_plane[t]].ss_err[trial_level[t]] = (int64_t)res[t];
        }
        bool all_done = true;
        for (int p = 0; p < nplanes; p++)
            if (!ps[p].done()) {
                ps[p].advance(cfg);
                all_done = all_done && ps[p].done();
            }
        if (all_done)
            break;
    }
    CUDA_TRY(c, cudaMemsetAsync(d_out, 0, sizeof(*d_out) * 6, c->stream)); // the pick expects a zeroed tot
    y = ps[0].filt_best;
    if (cfg->search_chroma)
        u = ps[1].filt_best, v = ps[2].filt_best;
    out_levels[0] = out_levels[1] = y, out_levels[2] = u, out_levels[3] = v;
    return SVTF_OK;
}
