// svtf_ccso_derive_api.inl: the CCSO filter derivation (svtf_ccso_derive_frame) — host driver of svtf_ccso_derive.cuh — textually included by svtf_api.cu (same translation unit: the context struct and the helpers above it are in scope).

// ---- CCSO filter derivation ----------------------------------------------------------------------
// ccso_search / derive_ccso_filter (Encoder/Codec/EbPickccso.c:464-782, :785-812).  Kernels: svtf_ccso_derive.cuh.
// The host keeps the reference's sequential decisions verbatim (float LUT rounding :431-461, double RD cost
// EbPickccso.h:9 + EbRestoration.h:346, per-block on/off :71-121 whose CDF rate never reaches the caller's cost,
// LUT bit count :360-379).  The 4 (band-only: 8) band settings of one (bo, support, quantiser, classifier) group have
// independent training loops, so they run in lock-step: one trial launch per round evaluates every LUT still training.
namespace {

const int     kCcsoOffset[8] = {-10, -7, -3, -1, 0, 1, 3, 7};
const uint8_t kCcsoQuant[4]  = {16, 8, 32, 64};

struct DeriveScratch { // device + pinned host buffers of one context, grown on demand
    int                *d_herr = nullptr, *d_hcnt = nullptr, *h_herr = nullptr, *h_hcnt = nullptr; // [all groups][nfb][nidx]
    unsigned long long *d_dist = nullptr, *h_dist = nullptr;                                       // [slot][CD_MAX_LUT][nfb]
    int8_t             *d_luts = nullptr, *h_luts = nullptr;                                       // [slot][CD_MAX_LUT][2048]
    CdGroup            *d_groups = nullptr, *h_groups = nullptr;                                   // [slot]
    int                 nfb_cap = 0;
};
constexpr int CD_HIST_ENTRIES = 2048 + (CD_MAX_GROUPS - 1) * 128; // histogram entries per filter block over all groups

int ccso_lut_bits(const int8_t *lut, int band_log2, int intervals, int bo) {
    static const int reordered[8] = {0, 1, -1, 3, -3, 7, -7, -10};
    int              bits = 0, n = bo ? 1 : intervals;
    for (int d0 = 0; d0 < n; d0++)
        for (int d1 = 0; d1 < n; d1++)
            for (int b = 0; b < (1 << band_log2); b++)
                for (int i = 0; i < 7; i++) {
                    bits++;
                    if (reordered[i] == lut[(b << 4) + (d0 << 2) + d1])
                        break;
                }
    return bits;
}

void ccso_lut_from_means(int8_t *lut, const int *err, const int *cnt, int band_log2, int intervals, int bo) {
    const int n = bo ? 1 : intervals;
    for (int d0 = 0; d0 < n; d0++)
        for (int d1 = 0; d1 < n; d1++)
            for (int b = 0; b < (1 << band_log2); b++) {
                const int i = (b << 4) + (d0 << 2) + d1;
                if (!cnt[i])
                    continue;
                const float t = (float)err[i] / cnt[i];
                if (t < kCcsoOffset[0] || t >= kCcsoOffset[7]) {
                    lut[i] = (int8_t)std::min(std::max((int)t, kCcsoOffset[0]), kCcsoOffset[7]);
                } else {
                    for (int k = 0; k < 7; k++)
                        if (t >= kCcsoOffset[k] && t <= kCcsoOffset[k + 1]) {
                            lut[i] = (int8_t)(fabs(t - kCcsoOffset[k]) > fabs(t - kCcsoOffset[k + 1]) ? kCcsoOffset[k + 1] : kCcsoOffset[k]);
                            break;
                        }
                }
            }
}

double ccso_rd_cost(int rdmult, int bits, uint64_t dist, int bd) {
    return (((double)bits) * rdmult) / (double)(1 << 9) + ((double)(dist >> (2 * (bd - 8))) * (1 << 7));
}

struct BandState { // one band_log2 training loop
    bool                 enable = true, keep = true;
    double               prev_cost = DBL_MAX, best_cost = DBL_MAX;
    int                  iter = 0;
    int8_t               lut[2048], best_lut[2048];
    std::vector<uint8_t> ctl, best_ctl;
    std::vector<int>     herr, hcnt; // [nfb][entries] at this band count
};

} // namespace

static void derive_scratch_release(DeriveScratch *s) {
    cudaFree(s->d_herr), cudaFree(s->d_hcnt), cudaFree(s->d_dist), cudaFree(s->d_luts), cudaFree(s->d_groups);
    cudaFreeHost(s->h_herr), cudaFreeHost(s->h_hcnt), cudaFreeHost(s->h_dist), cudaFreeHost(s->h_luts), cudaFreeHost(s->h_groups);
}
static void derive_scratch_free(void *p) {
    DeriveScratch *s = static_cast<DeriveScratch *>(p);
    if (!s)
        return;
    derive_scratch_release(s);
    delete s;
}

static DeriveScratch *derive_scratch(svtf_ctx *c, int plane, int nfb) {
    if (!c->derive_scratch[plane])
        c->derive_scratch[plane] = new DeriveScratch();
    DeriveScratch *s = static_cast<DeriveScratch *>(c->derive_scratch[plane]);
    if (nfb > s->nfb_cap) {
        derive_scratch_release(s);
        *s              = DeriveScratch();
        const size_t hb = sizeof(int) * CD_HIST_ENTRIES * (size_t)nfb, db = sizeof(unsigned long long) * CD_MAX_GROUPS * CD_MAX_LUT * (size_t)nfb;
        const size_t lb = (size_t)CD_MAX_GROUPS * CD_MAX_LUT * 2048, gb = sizeof(CdGroup) * CD_MAX_GROUPS;
        bool         ok = cudaMalloc(&s->d_herr, hb) == cudaSuccess && cudaMalloc(&s->d_hcnt, hb) == cudaSuccess &&
            cudaMalloc(&s->d_dist, db) == cudaSuccess && cudaMalloc(&s->d_luts, lb) == cudaSuccess && cudaMalloc(&s->d_groups, gb) == cudaSuccess &&
            cudaHostAlloc(reinterpret_cast<void **>(&s->h_herr), hb, cudaHostAllocDefault) == cudaSuccess &&
            cudaHostAlloc(reinterpret_cast<void **>(&s->h_hcnt), hb, cudaHostAllocDefault) == cudaSuccess &&
            cudaHostAlloc(reinterpret_cast<void **>(&s->h_dist), db, cudaHostAllocDefault) == cudaSuccess &&
            cudaHostAlloc(reinterpret_cast<void **>(&s->h_luts), lb, cudaHostAllocDefault) == cudaSuccess &&
            cudaHostAlloc(reinterpret_cast<void **>(&s->h_groups), gb, cudaHostAllocDefault) == cudaSuccess;
        if (!ok)
            return nullptr;
        s->nfb_cap = nfb;
    }
    return s;
}

template <typename T>
static void launch_hist(const CcsoDeriveArgs &a, dim3 grid, cudaStream_t s) { ccso_hist_kernel<T><<<grid, CD_THREADS, 0, s>>>(a); }
template <typename T>
static void launch_try(const CcsoDeriveArgs &a, dim3 grid, cudaStream_t s) { ccso_try_kernel<T><<<grid, CD_THREADS, 0, s>>>(a); }

// fn(i) for i in [0, n) on a few host threads (the groups of the CCSO derivation are independent of each other)
template <typename F>
static void ccso_parallel_for(int n, F fn) {
    const int nt = std::max(1, std::min({n, 8, (int)std::thread::hardware_concurrency()}));
    if (nt == 1) {
        for (int i = 0; i < n; i++) fn(i);
        return;
    }
    std::atomic<int>         next(0);
    std::vector<std::thread> th;
    for (int t = 0; t < nt; t++)
        th.emplace_back([&] {
            for (int i = next++; i < n; i = next++) fn(i);
        });
    for (auto &t : th) t.join();
}

// derive_ccso_filter for one plane on the context's surfaces: classifier = pre-CDEF luma, rec = current picture,
// org = uploaded source.  The 49 configuration groups (band-only; 6 supports x 4 quantisers x 2 classifiers) are independent of each
// other and so are the band settings inside a group, so they all train in lock-step: ONE launch builds every group's histograms, then
// one launch per training round tries every LUT that is still in training (<= 16 rounds).  The decisions between the rounds (float LUT
// rounding, per-block on/off, double RD cost) stay on the host, per (group, band) exactly as the C code has them, and the final choice
// walks the groups in the reference's order (strict "<": the earlier setting wins ties).
static int ccso_derive_plane(svtf_ctx *c, int plane, int rdmult, svtf_ccso_params *out, uint8_t *flags, cudaStream_t stream, uint64_t *launches) {
    const svtf_geom &g  = c->geom;
    const int        ss = plane ? 1 : 0, bd = g.bit_depth;
    const int        pw = g.width >> ss, ph = g.height >> ss, bl = plane ? 7 : 8;
    const int        nbx = (pw + (1 << bl) - 1) >> bl, nby = (ph + (1 << bl) - 1) >> bl, nfb = nbx * nby;
    DeriveScratch   *sc  = derive_scratch(c, plane, nfb);
    if (!sc)
        return c->fail(SVTF_E_NOMEM, "svtf_ccso_derive_frame: scratch allocation failed");
    const Surface &lum = c->surf[c->pre_cdef >= 0 ? c->pre_cdef : c->cur], &cur = c->surf[c->cur];
    CcsoDeriveArgs a;
    memset(&a, 0, sizeof(a));
    a.luma = lum.base[0], a.luma_pitch = lum.pitch[0];
    a.rec = cur.base[plane], a.rec_pitch = cur.pitch[plane];
    a.org = c->source.base[plane], a.org_pitch = c->source.pitch[plane];
    a.W = g.width, a.H = g.height, a.pw = pw, a.ph = ph, a.ss = ss, a.blk_log2 = bl, a.nbx = nbx, a.max_val = (1 << bd) - 1;
    a.herr = sc->d_herr, a.hcnt = sc->d_hcnt, a.luts = sc->d_luts, a.dist = sc->d_dist, a.groups = sc->d_groups;
    const int  tiles_x = (pw + 63) / 64, tiles_y = (ph + 63) / 64;
    const bool u8 = g.sample_bytes == 1;

    // trial filter + per-block SSE for the `ns` slots described in sc->h_groups / sc->h_luts; results in sc->h_dist[(slot*8 + l)*nfb + fb]
    auto run_try = [&](int ns) -> int {
        CUDA_TRY(c, cudaMemcpyAsync(sc->d_groups, sc->h_groups, sizeof(CdGroup) * (size_t)ns, cudaMemcpyHostToDevice, stream));
        CUDA_TRY(c, cudaMemcpyAsync(sc->d_luts, sc->h_luts, (size_t)ns * CD_MAX_LUT * 2048, cudaMemcpyHostToDevice, stream));
        CUDA_TRY(c, cudaMemsetAsync(sc->d_dist, 0, sizeof(unsigned long long) * (size_t)ns * CD_MAX_LUT * nfb, stream));
        const dim3 grid(tiles_x, tiles_y, ns);
        if (u8)
            launch_try<uint8_t>(a, grid, stream);
        else
            launch_try<uint16_t>(a, grid, stream);
        CUDA_TRY(c, cudaGetLastError());
        (*launches)++;
        CUDA_TRY(c, cudaMemcpyAsync(sc->h_dist, sc->d_dist, sizeof(unsigned long long) * (size_t)ns * CD_MAX_LUT * nfb, cudaMemcpyDeviceToHost, stream));
        CUDA_TRY(c, cudaStreamSynchronize(stream));
        return SVTF_OK;
    };

    // unfiltered distortion: the trial kernel with an all-zero LUT (band-only so no classes are evaluated)
    std::vector<uint64_t> unf(nfb);
    {
        memset(sc->h_luts, 0, 2048);
        CdGroup &z = sc->h_groups[0];
        memset(&z, 0, sizeof(z));
        z.bo = 1, z.nl = 1, z.shift[0] = bd;
        if (int r = run_try(1))
            return r;
        for (int i = 0; i < nfb; i++) unf[i] = sc->h_dist[i];
    }
    uint64_t unf_total = 0;
    for (int i = 0; i < nfb; i++) unf_total += unf[i];
    const double unfiltered_cost = ccso_rd_cost(rdmult, 1, unf_total, bd);
    const int    frame_bits = 10, frame_bits_bo = 5; // EbPickccso.c:531-546 with CONFIG_CCSO_SIGFIX
    static const int taps[6][2] = {{-1, 0}, {-1, -1}, {0, -1}, {1, -1}, {0, -3}, {0, -5}}; // derive_ccso_sample_pos

    // ---- the groups, in the reference's loop order ----
    struct Group {
        int                    bo, sup, q, clf, intervals, nbands, fine;
        CdGroup                dev;
        std::vector<BandState> st;
    };
    std::vector<Group> groups;
    int                hoff = 0;
    for (int bo = 0; bo < 2; bo++)
        for (int sup = 0; sup < (bo ? 1 : 6); sup++)
            for (int q = 0; q < (bo ? 1 : 4); q++)
                for (int clf = 0; clf < (bo ? 1 : 2); clf++) {
                    Group gr;
                    gr.bo = bo, gr.sup = sup, gr.q = q, gr.clf = clf;
                    gr.intervals = clf ? 2 : 3, gr.nbands = bo ? 8 : 4, gr.fine = gr.nbands - 1; // finest band_log2 of the group
                    memset(&gr.dev, 0, sizeof(gr.dev));
                    gr.dev.bo = bo, gr.dev.dy = taps[sup][0], gr.dev.dx = taps[sup][1], gr.dev.thr = kCcsoQuant[q], gr.dev.edge_clf = clf;
                    gr.dev.hist_shift = bd - gr.fine, gr.dev.nidx = bo ? 2048 : 128, gr.dev.hoff = hoff;
                    hoff += gr.dev.nidx * nfb;
                    groups.push_back(std::move(gr));
                }
    const int ng = (int)groups.size();

    static const bool timing = getenv("SVTF_CCSO_TIMING") != nullptr;
    auto              tnow   = [] { return std::chrono::steady_clock::now(); };
    auto              tms    = [](std::chrono::steady_clock::time_point a0, std::chrono::steady_clock::time_point b0) {
        return std::chrono::duration<double, std::milli>(b0 - a0).count();
    };
    double     t_hist = 0, t_fold = 0, t_lut = 0, t_try = 0, t_post = 0;
    int        rounds = 0;
    const auto t0 = tnow();
    // ---- every group's histograms at its finest band split: one launch ----
    {
        for (int k = 0; k < ng; k++) sc->h_groups[k] = groups[k].dev;
        const size_t hbytes = sizeof(int) * (size_t)hoff;
        CUDA_TRY(c, cudaMemcpyAsync(sc->d_groups, sc->h_groups, sizeof(CdGroup) * (size_t)ng, cudaMemcpyHostToDevice, stream));
        CUDA_TRY(c, cudaMemsetAsync(sc->d_herr, 0, hbytes, stream));
        CUDA_TRY(c, cudaMemsetAsync(sc->d_hcnt, 0, hbytes, stream));
        const dim3 grid(tiles_x, tiles_y, ng);
        if (u8)
            launch_hist<uint8_t>(a, grid, stream);
        else
            launch_hist<uint16_t>(a, grid, stream);
        CUDA_TRY(c, cudaGetLastError());
        (*launches)++;
        CUDA_TRY(c, cudaMemcpyAsync(sc->h_herr, sc->d_herr, hbytes, cudaMemcpyDeviceToHost, stream));
        CUDA_TRY(c, cudaMemcpyAsync(sc->h_hcnt, sc->d_hcnt, hbytes, cudaMemcpyDeviceToHost, stream));
        CUDA_TRY(c, cudaStreamSynchronize(stream));
    }
    const auto t1 = tnow();
    t_hist = tms(t0, t1);
    // per band count: histograms at that granularity (band = finest band >> (fine - log2))
    // (finest first, every coarser band count from the next finer one: band j of b = bands 2j and 2j+1 of b+1)
    ccso_parallel_for(ng, [&](int gi) {
        Group &gr = groups[gi];
        gr.st.resize(gr.nbands);
        const int *herr = sc->h_herr + gr.dev.hoff, *hcnt = sc->h_hcnt + gr.dev.hoff;
        for (int b = gr.fine; b >= 0; b--) {
            BandState &s = gr.st[b];
            s.ctl.assign(nfb, 1), s.best_ctl.assign(nfb, 0);
            memset(s.lut, 0, sizeof(s.lut)), memset(s.best_lut, 0, sizeof(s.best_lut));
            const int ne = 16 << b; // LUT indices < (1<<b) << 4
            s.herr.resize((size_t)nfb * ne), s.hcnt.resize((size_t)nfb * ne);
            if (b == gr.fine) { // ne == nidx for the band-only group (2048); the edge groups keep 128 entries per block on the device
                for (int fb = 0; fb < nfb; fb++) {
                    memcpy(&s.herr[(size_t)fb * ne], herr + (size_t)fb * gr.dev.nidx, sizeof(int) * ne);
                    memcpy(&s.hcnt[(size_t)fb * ne], hcnt + (size_t)fb * gr.dev.nidx, sizeof(int) * ne);
                }
                continue;
            }
            const BandState &f  = gr.st[b + 1];
            const int        nf = 16 << (b + 1);
            for (int fb = 0; fb < nfb; fb++)
                for (int j = 0; j < (1 << b); j++)
                    for (int k = 0; k < 16; k++) {
                        s.herr[(size_t)fb * ne + (j << 4) + k] = f.herr[(size_t)fb * nf + ((2 * j) << 4) + k] + f.herr[(size_t)fb * nf + ((2 * j + 1) << 4) + k];
                        s.hcnt[(size_t)fb * ne + (j << 4) + k] = f.hcnt[(size_t)fb * nf + ((2 * j) << 4) + k] + f.hcnt[(size_t)fb * nf + ((2 * j + 1) << 4) + k];
                    }
        }
    });
    t_fold = tms(t1, tnow());
    // ---- lock-step training rounds over every (group, band) still in training ----
    for (;;) {
        const auto r0 = tnow();
        int slot_group[CD_MAX_GROUPS], slot_band[CD_MAX_GROUPS][CD_MAX_LUT], ns = 0;
        int task_g[CD_MAX_GROUPS * CD_MAX_LUT], task_b[CD_MAX_GROUPS * CD_MAX_LUT], ntask = 0;
        for (int k = 0; k < ng; k++)
            for (int b = 0; b < groups[k].nbands; b++)
                if (groups[k].st[b].keep && groups[k].st[b].enable)
                    task_g[ntask] = k, task_b[ntask] = b, ntask++;
        ccso_parallel_for(ntask, [&](int t) { // ccso_compute_class_err + derive_lut_offset of every band still enabled
            Group     &gr = groups[task_g[t]];
            const int  b  = task_b[t], ne = 16 << b;
            BandState &s  = gr.st[b];
            int        err[2048], cnt[2048];
            memset(err, 0, sizeof(int) * ne), memset(cnt, 0, sizeof(int) * ne);
            memset(s.lut, 0, sizeof(s.lut));
            const int n = gr.bo ? 1 : gr.intervals;
            for (int fb = 0; fb < nfb; fb++) {
                if (!s.ctl[fb])
                    continue;
                const int *he = &s.herr[(size_t)fb * ne], *hc = &s.hcnt[(size_t)fb * ne];
                for (int bn = 0; bn < (1 << b); bn++)
                    for (int d0 = 0; d0 < n; d0++)
                        for (int d1 = 0; d1 < n; d1++) {
                            const int i = (bn << 4) + (d0 << 2) + d1;
                            err[i] += he[i], cnt[i] += hc[i];
                        }
            }
            ccso_lut_from_means(s.lut, err, cnt, b, gr.intervals, gr.bo);
        });
        for (int k = 0; k < ng; k++) {
            Group &gr = groups[k];
            int    na = 0;
            for (int b = 0; b < gr.nbands; b++) {
                BandState &s = gr.st[b];
                if (!s.keep)
                    continue;
                memcpy(sc->h_luts + ((size_t)ns * CD_MAX_LUT + na) * 2048, s.lut, 2048);
                gr.dev.shift[na] = bd - b, slot_band[ns][na] = b, na++;
            }
            if (!na)
                continue;
            gr.dev.nl          = na;
            sc->h_groups[ns]   = gr.dev;
            slot_group[ns]     = k, ns++;
        }
        if (!ns)
            break;
        const auto r1 = tnow();
        if (int r = run_try(ns))
            return r;
        const auto r2 = tnow();
        rounds++;
        ccso_parallel_for(ns, [&](int sl) {
            Group &gr = groups[slot_group[sl]];
            for (int k = 0; k < sc->h_groups[sl].nl; k++) {
                BandState      &s   = gr.st[slot_band[sl][k]];
                const uint64_t *trn = reinterpret_cast<const uint64_t *>(sc->h_dist) + ((size_t)sl * CD_MAX_LUT + k) * nfb;
                bool            improvement = false;
                uint64_t        cur_dist = 0;
                if (s.enable) { // derive_blk_md
                    bool any = false;
                    for (int fb = 0; fb < nfb; fb++) {
                        const bool on = trn[fb] < unf[fb];
                        s.ctl[fb]     = on;
                        cur_dist += on ? trn[fb] : unf[fb];
                        any |= on;
                    }
                    s.enable = any;
                }
                if (s.enable) {
                    const int    bits = ccso_lut_bits(s.lut, slot_band[sl][k], gr.intervals, gr.bo) + (gr.bo ? frame_bits_bo : frame_bits) + nfb;
                    const double cost = ccso_rd_cost(rdmult, bits, cur_dist, bd);
                    if (cost < s.prev_cost)
                        s.prev_cost = cost, improvement = true;
                    if (cost < s.best_cost) {
                        s.best_cost = cost;
                        memcpy(s.best_lut, s.lut, sizeof(s.lut));
                        s.best_ctl = s.ctl;
                    }
                }
                s.iter++;
                if (!improvement || s.iter > 15) // CCSO_MAX_ITERATIONS
                    s.keep = false;
            }
        });
        t_lut += tms(r0, r1), t_try += tms(r1, r2), t_post += tms(r2, tnow());
    }
    if (timing)
        fprintf(stderr, "ccso_derive_plane %d: histograms %.2f ms, fold %.2f, %d rounds: luts %.2f, try %.2f, decisions %.2f\n", plane, t_hist, t_fold,
                rounds, t_lut, t_try, t_post);
    // ---- the reference's order: strict "<" keeps the earlier setting on ties ----
    double               final_cost      = DBL_MAX;
    int8_t               final_lut[2048] = {0};
    std::vector<uint8_t> final_ctl(nfb, 0);
    int                  final_clf = 0, final_q = 0, final_sup = 0, final_bo = 0, final_band = 0;
    for (const Group &gr : groups)
        for (int b = 0; b < gr.nbands; b++)
            if (gr.st[b].best_cost < final_cost) {
                final_cost = gr.st[b].best_cost;
                final_q = gr.q, final_sup = gr.sup, final_bo = gr.bo, final_band = b, final_clf = gr.clf;
                memcpy(final_lut, gr.st[b].best_lut, sizeof(final_lut));
                final_ctl = gr.st[b].best_ctl;
            }
    const bool on = !(unfiltered_cost < final_cost);
    out->ccso_enable[plane] = on;
    memset(flags, 0, nfb);
    if (on) {
        memcpy(flags, final_ctl.data(), nfb);
        memcpy(out->filter_offset[plane], final_lut, 2048);
        out->quant_idx[plane] = (uint8_t)final_q, out->ext_filter_support[plane] = (uint8_t)final_sup;
        out->ccso_bo_only[plane] = (uint8_t)final_bo, out->max_band_log2[plane] = (uint8_t)final_band;
        out->edge_clf[plane] = (uint8_t)final_clf;
    }
    return SVTF_OK;
}

extern "C" int svtf_ccso_derive_frame(svtf_ctx *c, int32_t rdmult, int32_t plane_start, int32_t plane_end, svtf_ccso_params *out,
                                      uint8_t *const blk_flags[3]) {
    if (!c)
        return set_global_error(SVTF_E_ARG, "svtf_ccso_derive_frame: NULL context");
    if (!out || !blk_flags || plane_start < 0 || plane_end > 3 || plane_start > plane_end || rdmult < 0)
        return c->fail(SVTF_E_ARG, "svtf_ccso_derive_frame: bad argument");
    if (!c->have_frame || !c->have_source)
        return c->fail(SVTF_E_STATE, "svtf_ccso_derive_frame: upload the picture and the source first");
    if (c->striped)
        return c->fail(SVTF_E_STATE, "svtf_ccso_derive_frame: not available on a striped context");
    CUDA_TRY(c, cudaSetDevice(c->device));
    for (int p = plane_start; p < plane_end; p++)
        if (!blk_flags[p])
            return c->fail(SVTF_E_ARG, "svtf_ccso_derive_frame: NULL flag array for a requested plane");
    // the planes are independent (each writes its own fields of *out): one host thread and one stream per plane, behind what is queued
    // on the context's stream.  SVTF_CCSO_SERIAL=1: one plane after the other on the context's stream.
    static const bool serial = getenv("SVTF_CCSO_SERIAL") != nullptr;
    uint64_t          nl[3]  = {0, 0, 0};
    int               rc[3]  = {SVTF_OK, SVTF_OK, SVTF_OK};
    if (serial || plane_end - plane_start < 2) {
        for (int p = plane_start; p < plane_end; p++) {
            rc[p] = ccso_derive_plane(c, p, rdmult, out, blk_flags[p], c->stream, &nl[p]);
            c->launches += nl[p];
            if (rc[p])
                return rc[p];
        }
        return SVTF_OK;
    }
    if (!c->derive_ev)
        CUDA_TRY(c, cudaEventCreateWithFlags(&c->derive_ev, cudaEventDisableTiming));
    CUDA_TRY(c, cudaEventRecord(c->derive_ev, c->stream));
    for (int p = plane_start; p < plane_end; p++) {
        if (!c->derive_stream[p])
            CUDA_TRY(c, cudaStreamCreateWithFlags(&c->derive_stream[p], cudaStreamNonBlocking));
        CUDA_TRY(c, cudaStreamWaitEvent(c->derive_stream[p], c->derive_ev, 0));
    }
    std::vector<std::thread> th;
    for (int p = plane_start; p < plane_end; p++)
        th.emplace_back([&, p] {
            if (cudaSetDevice(c->device) != cudaSuccess) {
                rc[p] = SVTF_E_CUDA;
                return;
            }
            rc[p] = ccso_derive_plane(c, p, rdmult, out, blk_flags[p], c->derive_stream[p], &nl[p]);
        });
    for (auto &t : th) t.join();
    for (int p = plane_start; p < plane_end; p++) c->launches += nl[p];
    for (int p = plane_start; p < plane_end; p++)
        if (rc[p])
            return rc[p]; // the plane's own message is in svtf_last_error (every round ends with a stream synchronisation)
    return SVTF_OK;
}
