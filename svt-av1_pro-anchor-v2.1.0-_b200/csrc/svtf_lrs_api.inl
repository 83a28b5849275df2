// svtf_lrs_api.inl — part of svtf_api.cu (included there, after svtf_ctx and its helpers): the C-ABI entry points of the stages
// that follow CDEF / CCSO in the encoder besides the restoration filter itself — the super-resolution upscale
// (svtf_superres_upscale_frame) and the loop-restoration filter search (svtf_lr_wiener_stats, svtf_lr_search_frame).
// Kernels: svtf_superres.cuh, svtf_lrs.cuh, svtf_lr.cuh; host arithmetic: svtf_lrs_host.cpp.

// ---- super-resolution upscale (svtf_superres.cuh) --------------------------------------------------------------------
extern "C" int svtf_superres_upscale_frame(svtf_ctx *c, const svtf_superres_params *sr) {
    if (!c)
        return set_global_error(SVTF_E_ARG, "svtf_superres_upscale_frame: NULL context");
    if (!sr || !sr->filters || sr->tile_cols < 1 || sr->tile_cols > 64 || sr->superres_denominator < 8 || sr->superres_denominator > 16)
        return c->fail(SVTF_E_ARG, "svtf_superres_upscale_frame: bad argument");
    if (!c->have_frame)
        return c->fail(SVTF_E_STATE, "svtf_superres_upscale_frame: upload a picture first");
    if (c->striped)
        return c->fail(SVTF_E_STATE, "svtf_superres_upscale_frame: not available on a striped context");
    if (sr->superres_denominator == 8)
        return SVTF_OK; // av1_superres_unscaled
    svtf_geom ng = c->geom;
    ng.width     = sr->upscaled_width;
    ng.mi_cols   = 2 * ((sr->upscaled_width + 7) >> 3);
    if (sr->upscaled_width < c->geom.width || ng.mi_cols > c->max_geom.mi_cols)
        return c->fail(SVTF_E_ARG, "svtf_superres_upscale_frame: the upscaled width must lie between the picture's and the context's width");
    for (int j = 0; j <= sr->tile_cols; j++)
        if (sr->tile_col_start_mi[j] < 0 || (j && sr->tile_col_start_mi[j] <= sr->tile_col_start_mi[j - 1]))
            return c->fail(SVTF_E_ARG, "svtf_superres_upscale_frame: tile_col_start_mi must increase");
    CUDA_TRY(c, cudaSetDevice(c->device));
    if (!c->surf[3].base[0]) {
        cudaError_t e = alloc_surface(c->surf[3], c->max_geom);
        if (e != cudaSuccess)
            return c->cuda_fail(e, "cudaMalloc(restoration surface)");
    }
    if (c->cur == 3)
        return c->fail(SVTF_E_STATE, "svtf_superres_upscale_frame: the current picture is already a restoration output");
    if (int r = handoff_reserve(c, sizeof(SuperresTables)))
        return r;
    // per plane and tile column: where the tile starts in both pictures and the sub-sample offset it starts with (:223-283)
    SuperresTables tab;
    SuperresArgs   a;
    std::memcpy(tab.filt, sr->filters, sizeof(tab.filt));
    a.row_base[0] = 0;
    for (int p = 0; p < 3; p++) {
        const int ss = p ? 1 : 0;
        const int ds_w = (c->geom.width + ss) >> ss, up_w = (sr->upscaled_width + ss) >> ss; // ROUND_POWER_OF_TWO(width, sub_x)
        a.step[p]      = ((ds_w << 14) + up_w / 2) / up_w; // av1_get_upscale_convolve_step
        const int err  = up_w * a.step[p] - (ds_w << 14);
        int       x0   = (-((up_w - ds_w) << 13) + up_w / 2) / up_w + (1 << 7) - err / 2; // get_upscale_convolve_x0
        x0             = (int)((unsigned)x0 & 16383u);
        for (int j = 0; j < sr->tile_cols; j++) {
            const int ds0 = sr->tile_col_start_mi[j] << (2 - ss), ds1 = std::min(sr->tile_col_start_mi[j + 1], c->geom.mi_cols) << (2 - ss);
            const int up0 = ds0 * sr->superres_denominator / 8;
            const int up1 = j == sr->tile_cols - 1 ? up_w : ds1 * sr->superres_denominator / 8;
            tab.up_x0[p][j] = up0, tab.ds_x0[p][j] = ds0, tab.x0_qn[p][j] = x0;
            x0 += (up1 - up0) * a.step[p] - ((ds1 - ds0) << 14);
            tab.up_x0[p][j + 1] = up1;
        }
        a.src_w[p] = (c->geom.mi_cols * 4) >> ss, a.dst_w[p] = up_w, a.dst_cw[p] = (ng.mi_cols * 4) >> ss, a.rows[p] = (c->geom.mi_rows * 4) >> ss;
        a.row_base[p + 1] = a.row_base[p] + a.rows[p];
        a.x0_qn_first[p] = tab.x0_qn[p][0], a.up_x1_first[p] = tab.up_x0[p][1];
    }
    a.tile_cols = sr->tile_cols, a.bd = c->geom.bit_depth;
    a.tab       = reinterpret_cast<const SuperresTables *>(c->d_handoff);
    // pageable source: the runtime stages the 6 KB before the call returns, so `tab` may leave scope
    CUDA_TRY(c, cudaMemcpyAsync(c->d_handoff, &tab, sizeof(tab), cudaMemcpyHostToDevice, c->stream));
    const dim3 blocks((a.dst_cw[0] + 256 * SR_PER - 1) / (256 * SR_PER), a.row_base[3]);
    auto      launch = [&](int from, int to) -> int {
        a.src = c->surf[from].planes(), a.dst = c->surf[to].planes();
        if (c->geom.sample_bytes == 1)
            superres_kernel<uint8_t><<<blocks, 256, 0, c->stream>>>(a);
        else
            superres_kernel<uint16_t><<<blocks, 256, 0, c->stream>>>(a);
        CUDA_TRY(c, cudaGetLastError());
        c->launches++;
        return SVTF_OK;
    };
    // surfaces: the current picture goes to a free one; the picture before CDEF goes through surface 3 (the restoration
    // output) back into its own surface, so that loop restoration still finds surface 3 free
    int free_s = 0;
    while (free_s == c->cur || free_s == c->pre_cdef) free_s++;
    if (int r = launch(c->cur, free_s))
        return r;
    if (c->pre_cdef >= 0 && c->pre_cdef != c->cur) {
        if (int r = launch(c->pre_cdef, 3))
            return r;
        for (int p = 0; p < 3; p++) {
            const size_t bytes = (size_t)c->surf[3].pitch[p] * a.rows[p] * c->geom.sample_bytes;
            CUDA_TRY(c, cudaMemcpyAsync(c->surf[c->pre_cdef].base[p], c->surf[3].base[p], bytes, cudaMemcpyDeviceToDevice, c->stream));
        }
    } else if (c->pre_cdef == c->cur)
        c->pre_cdef = free_s;
    c->cur     = free_s;
    c->geom    = ng;
    c->have_mi = c->have_search = c->have_source = c->edge_valid = false;
    return SVTF_OK;
}

// ---- loop-restoration filter search, Wiener + self-guided tools (svtf_lrs.cuh, svtf_lrs_host.cpp) -----------------------------------
namespace {
struct LrsPlane { // one searched plane
    int plane, win, unit, hu, vu, first; // first: index of its unit 0 in the per-call unit arrays
};
struct LrsPlan {
    std::vector<LrsPlane>           planes;
    std::vector<LrsTile>            tiles;
    std::vector<int>                count;   // samples per unit
    std::vector<unsigned long long> acc_off; // offset of each unit's statistics accumulators
    size_t                          acc_total = 0;
    int                             units = 0, nslots = 1; // nslots: parameter sets the self-guided search measures at once
    // device carve-up
    size_t o_tiles = 0, o_sum = 0, o_sse = 0, o_count = 0, o_off = 0, o_acc = 0, o_trial = 0, o_units = 0, o_proj = 0, o_cand = 0, o_errk = 0, bytes = 0;
};

// Units of a plane (foreach_rest_unit_in_tile, Common/Codec/EbRestoration.c:1250): raster order, the last one of a row /
// column takes the remainder, rows shifted up by 8 luma samples.
void lrs_unit_rect(const svtf_geom &g, const LrsPlane &pl, int u, int lim[4]) {
    const int ss = pl.plane ? 1 : 0, pw = (g.width + ss) >> ss, ph = (g.height + ss) >> ss, off = 8 >> ss;
    const int ur = u / pl.hu, uc = u - ur * pl.hu;
    lim[0] = uc * pl.unit, lim[1] = uc == pl.hu - 1 ? pw : (uc + 1) * pl.unit;
    lim[2] = std::max(0, ur * pl.unit - off), lim[3] = ur == pl.vu - 1 ? ph : (ur + 1) * pl.unit - off;
}

void lrs_add_plane(const svtf_geom &g, LrsPlan &plan, int plane, int unit_size, int win) {
    LrsPlane pl{plane, win, unit_size, 0, 0, plan.units};
    svtf_lr_units(&g, plane, unit_size, &pl.hu, &pl.vu);
    for (int u = 0; u < pl.hu * pl.vu; u++) {
        int lim[4];
        lrs_unit_rect(g, pl, u, lim);
        plan.count.push_back((lim[1] - lim[0]) * (lim[3] - lim[2]));
        plan.acc_off.push_back(plan.acc_total);
        plan.acc_total += (size_t)win * win * win * win + win * win;
        for (int y = lim[2]; y < lim[3]; y += LRS_T)
            for (int x = lim[0]; x < lim[1]; x += LRS_T)
                plan.tiles.push_back(LrsTile{(int16_t)x, (int16_t)y, (int16_t)std::min(LRS_T, lim[1] - x), (int16_t)std::min(LRS_T, lim[3] - y),
                                             plan.units + u, plane});
    }
    plan.units += pl.hu * pl.vu;
    plan.planes.push_back(pl);
}

int lrs_reserve(svtf_ctx *c, LrsPlan &plan) {
    auto carve = [&](size_t &o, size_t n) { o = plan.bytes, plan.bytes += (n + 255) & ~size_t(255); };
    carve(plan.o_tiles, sizeof(LrsTile) * plan.tiles.size());
    carve(plan.o_sum, 8 * (size_t)plan.units), carve(plan.o_sse, 8 * (size_t)plan.units), carve(plan.o_trial, 8 * (size_t)plan.units);
    carve(plan.o_count, 4 * (size_t)plan.units), carve(plan.o_off, 8 * (size_t)plan.units);
    carve(plan.o_units, sizeof(LrUnitDev) * (size_t)plan.units);
    carve(plan.o_proj, 40 * (size_t)plan.units * plan.nslots);
    carve(plan.o_cand, sizeof(LrsCand) * (size_t)plan.units * plan.nslots);
    carve(plan.o_errk, 8 * (size_t)LRS_CAND * plan.units * plan.nslots);
    carve(plan.o_acc, 8 * plan.acc_total);
    if (plan.bytes > c->lrs_cap) {
        if (c->d_lrs) {
            CUDA_TRY(c, cudaStreamSynchronize(c->stream));
            CUDA_TRY(c, cudaFree(c->d_lrs));
            c->d_lrs = nullptr, c->lrs_cap = 0;
        }
        CUDA_TRY(c, cudaMalloc(&c->d_lrs, plan.bytes));
        c->lrs_cap = plan.bytes;
    }
    if (plan.bytes > c->h_lrs_cap) { // same carve-up on the host side, pinned
        if (c->h_lrs) {
            CUDA_TRY(c, cudaStreamSynchronize(c->stream));
            CUDA_TRY(c, cudaFreeHost(c->h_lrs));
            c->h_lrs = nullptr, c->h_lrs_cap = 0;
        }
        CUDA_TRY(c, cudaMallocHost(&c->h_lrs, plan.bytes));
        c->h_lrs_cap = plan.bytes;
    }
    return SVTF_OK;
}

// Uploads the plan, runs the per-unit sums (and, with stats, the Wiener statistics) and brings the results to the host.
int lrs_measure(svtf_ctx *c, LrsPlan &plan, const bool stats_of[3], const unsigned long long *&sse_none, const long long *&acc) {
    const bool stats = stats_of[0] || stats_of[1] || stats_of[2];
    if (int r = lrs_reserve(c, plan))
        return r;
    uint8_t *d = c->d_lrs, *h = c->h_lrs;
    std::memcpy(h + plan.o_tiles, plan.tiles.data(), sizeof(LrsTile) * plan.tiles.size());
    std::memcpy(h + plan.o_count, plan.count.data(), 4 * (size_t)plan.units);
    std::memcpy(h + plan.o_off, plan.acc_off.data(), 8 * (size_t)plan.units);
    CUDA_TRY(c, cudaMemcpyAsync(d + plan.o_tiles, h + plan.o_tiles, sizeof(LrsTile) * plan.tiles.size(), cudaMemcpyHostToDevice, c->stream));
    CUDA_TRY(c, cudaMemcpyAsync(d + plan.o_count, h + plan.o_count, 4 * (size_t)plan.units, cudaMemcpyHostToDevice, c->stream));
    CUDA_TRY(c, cudaMemcpyAsync(d + plan.o_off, h + plan.o_off, 8 * (size_t)plan.units, cudaMemcpyHostToDevice, c->stream));
    CUDA_TRY(c, cudaMemsetAsync(d + plan.o_sum, 0, plan.o_count - plan.o_sum, c->stream)); // sum, sse, trial
    LrsArgs a;
    a.cur = c->surf[c->cur].planes(), a.src = c->source.planes();
    for (int p = 0; p < 3; p++) a.pw[p] = (c->geom.width + (p ? 1 : 0)) >> (p ? 1 : 0), a.ph[p] = (c->geom.height + (p ? 1 : 0)) >> (p ? 1 : 0);
    a.tiles      = reinterpret_cast<const LrsTile *>(d + plan.o_tiles);
    a.unit_sum   = reinterpret_cast<unsigned long long *>(d + plan.o_sum);
    a.unit_sse   = reinterpret_cast<unsigned long long *>(d + plan.o_sse);
    a.unit_count = reinterpret_cast<const int *>(d + plan.o_count);
    a.acc        = reinterpret_cast<long long *>(d + plan.o_acc);
    a.acc_off    = reinterpret_cast<const unsigned long long *>(d + plan.o_off);
    a.bd         = c->geom.bit_depth;
    const int  nt = (int)plan.tiles.size();
    const bool b8 = c->geom.sample_bytes == 1;
    if (b8)
        lrs_sums_kernel<uint8_t><<<nt, LRS_THREADS, 0, c->stream>>>(a);
    else
        lrs_sums_kernel<uint16_t><<<nt, LRS_THREADS, 0, c->stream>>>(a);
    CUDA_TRY(c, cudaGetLastError());
    c->launches++;
    if (stats) {
        CUDA_TRY(c, cudaMemsetAsync(d + plan.o_acc, 0, 8 * plan.acc_total, c->stream));
        // tiles are grouped by plane: one launch per plane with that plane's window
        int t0 = 0;
        for (const LrsPlane &pl : plan.planes) {
            int t1 = t0;
            while (t1 < nt && plan.tiles[t1].plane == pl.plane) t1++;
            if (!stats_of[pl.plane]) {
                t0 = t1;
                continue;
            }
            LrsArgs b = a;
            b.tiles   = a.tiles + t0;
            const int n = t1 - t0;
#ifdef SVTF_AB_KERNELS
            static const bool v1 = std::getenv("SVTF_LRS_STATS_V1") != nullptr;
#define LRS_LAUNCH(T, W) (v1 ? lrs_stats_kernel<T, W><<<n, LRS_THREADS, 0, c->stream>>>(b) : lrs_stats2_kernel<T, W><<<n, LRS_THREADS, 0, c->stream>>>(b))
#else
#define LRS_LAUNCH(T, W) (lrs_stats2_kernel<T, W><<<n, LRS_THREADS, 0, c->stream>>>(b))
#endif
            if (pl.win == 7)
                b8 ? LRS_LAUNCH(uint8_t, 7) : LRS_LAUNCH(uint16_t, 7);
            else if (pl.win == 5)
                b8 ? LRS_LAUNCH(uint8_t, 5) : LRS_LAUNCH(uint16_t, 5);
            else
                b8 ? LRS_LAUNCH(uint8_t, 3) : LRS_LAUNCH(uint16_t, 3);
#undef LRS_LAUNCH
            CUDA_TRY(c, cudaGetLastError());
            c->launches++;
            t0 = t1;
        }
        CUDA_TRY(c, cudaMemcpyAsync(h + plan.o_acc, d + plan.o_acc, 8 * plan.acc_total, cudaMemcpyDeviceToHost, c->stream));
    }
    CUDA_TRY(c, cudaMemcpyAsync(h + plan.o_sse, d + plan.o_sse, 8 * (size_t)plan.units, cudaMemcpyDeviceToHost, c->stream));
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    sse_none = reinterpret_cast<const unsigned long long *>(h + plan.o_sse);
    acc      = reinterpret_cast<const long long *>(h + plan.o_acc);
    return SVTF_OK;
}

// The reference's M / H of one unit from the raw accumulators: bit-depth divider (high-bitdepth path only), H mirrored.
void lrs_finish_stats(const svtf_geom &g, int win, const long long *raw, int64_t *M, int64_t *H) {
    const int win2 = win * win, div = g.sample_bytes == 1 ? 1 : (g.bit_depth == 12 ? 16 : (g.bit_depth == 10 ? 4 : 1));
    for (int k = 0; k < win2; k++) {
        M[k] = raw[win2 * win2 + k] / div;
        for (int l = k; l < win2; l++) H[k * win2 + l] = H[l * win2 + k] = raw[k * win2 + l] / div;
    }
}

int lrs_check(svtf_ctx *c, const char *fn) {
    if (!c->have_frame || !c->have_source)
        return c->fail(SVTF_E_STATE, std::string(fn) + ": upload a picture and a source first");
    if (c->striped)
        return c->fail(SVTF_E_STATE, std::string(fn) + ": not available on a striped context");
    return SVTF_OK;
}
template <class F>
void host_parallel_for(int n, F fn) {
    const int nt = std::max(1, std::min({(int)std::thread::hardware_concurrency(), 16, n / 8}));
    std::atomic<int> next{0};
    auto             worker = [&] {
        for (int i = next.fetch_add(1); i < n; i = next.fetch_add(1)) fn(i);
    };
    std::vector<std::thread> th;
    for (int t = 1; t < nt; t++) th.emplace_back(worker);
    worker();
    for (auto &t : th) t.join();
}
bool lrs_unit_size_ok(int plane, int us) { return (us == 64 || us == 128 || us == 256 || (plane && us == 32)); }
} // namespace

extern "C" int svtf_lr_wiener_stats(svtf_ctx *c, int32_t plane, int32_t unit_size, int32_t wiener_win, int64_t *M, int64_t *H) {
    if (!c)
        return set_global_error(SVTF_E_ARG, "svtf_lr_wiener_stats: NULL context");
    if (plane < 0 || plane > 2 || !lrs_unit_size_ok(plane, unit_size) || (wiener_win != 7 && wiener_win != 5 && wiener_win != 3) || !M || !H)
        return c->fail(SVTF_E_ARG, "svtf_lr_wiener_stats: bad argument");
    if (int r = lrs_check(c, "svtf_lr_wiener_stats"))
        return r;
    CUDA_TRY(c, cudaSetDevice(c->device));
    LrsPlan plan;
    lrs_add_plane(c->geom, plan, plane, unit_size, wiener_win);
    const unsigned long long *sse;
    const long long          *acc;
    const bool all[3] = {true, true, true};
    if (int r = lrs_measure(c, plan, all, sse, acc))
        return r;
    const int win2 = wiener_win * wiener_win;
    for (int u = 0; u < plan.units; u++) lrs_finish_stats(c->geom, wiener_win, acc + plan.acc_off[u], M + (size_t)u * win2, H + (size_t)u * win2 * win2);
    return SVTF_OK;
}

namespace {
// LrArgs of a measuring launch (LR_SSE / LR_FLT) over the planes whose `active` flag is set; returns the grid size
int lrs_trial_args(svtf_ctx *c, const LrsPlan &plan, const bool active[3], int frame_type, LrArgs &a) {
    uint8_t *d = c->d_lrs;
    a.cur = a.dbk = c->surf[c->cur].planes();
    a.out = a.cur; // never written by the measuring instantiations
    a.src = c->source.planes();
    for (int p = 0; p < 3; p++) {
        const int ss = p ? 1 : 0;
        a.pw[p] = (c->geom.width + ss) >> ss, a.ph[p] = (c->geom.height + ss) >> ss;
        a.cw[p] = a.pw[p], a.chh[p] = a.ph[p]; // only the visible picture is measured
        a.unit[p] = 0, a.hu[p] = a.vu[p] = 0, a.units[p] = nullptr, a.sse[p] = nullptr, a.proj[p] = nullptr;
        a.frame_type[p] = SVTF_RESTORE_NONE, a.chunks[p] = a.stripes[p] = 0, a.force_ep[p] = -1, a.flt[p] = nullptr, a.flt_half[p] = 0;
    }
    for (const LrsPlane &pl : plan.planes) {
        const int p = pl.plane, ss = p ? 1 : 0, sh = 64 >> ss, off = 8 >> ss;
        if (!active[p])
            continue;
        a.unit[p] = pl.unit, a.hu[p] = pl.hu, a.vu[p] = pl.vu;
        a.units[p]      = reinterpret_cast<const LrUnitDev *>(d + plan.o_units) + pl.first;
        a.sse[p]        = reinterpret_cast<unsigned long long *>(d + plan.o_trial) + pl.first;
        a.proj[p]       = reinterpret_cast<long long *>(d + plan.o_proj) + 5 * (size_t)pl.first;
        a.frame_type[p] = frame_type;
        a.chunks[p]     = (a.cw[p] + (LR_MAXW >> ss) - 1) / (LR_MAXW >> ss);
        a.stripes[p]    = (a.ph[p] + off + sh - 1) / sh + 1;
    }
    int grid = 0;
    for (int p = 0; p < 3; p++) a.base[p] = grid, grid += a.chunks[p] * a.stripes[p];
    a.opt = 2, a.bd = c->geom.bit_depth;
    return grid;
}

template <int MODE>
int lrs_launch_lr(svtf_ctx *c, const LrArgs &a, int grid) {
    if (grid <= 0)
        return SVTF_OK;
    if (c->geom.sample_bytes == 1)
        lr_kernel<uint8_t, MODE><<<grid, LR_THREADS, 0, c->stream>>>(a);
    else
        lr_kernel<uint16_t, MODE><<<grid, LR_THREADS, 0, c->stream>>>(a);
    CUDA_TRY(c, cudaGetLastError());
    c->launches++;
    return SVTF_OK;
}
} // namespace

extern "C" int svtf_lr_search_frame(svtf_ctx *c, const svtf_lr_search_cfg *cfg, int32_t out_frame_type[3], svtf_lr_unit *const out_units[3],
                                    svtf_lr_unit_search *const out_search[3], int32_t sg_frame_ep_cnt[16]) {
    if (!c)
        return set_global_error(SVTF_E_ARG, "svtf_lr_search_frame: NULL context");
    if (!cfg || !out_frame_type || !out_units)
        return c->fail(SVTF_E_ARG, "svtf_lr_search_frame: NULL argument");
    const bool wn = cfg->wn_enabled != 0, sg = cfg->sg_enabled != 0;
    if (wn && (cfg->wn_filter_tap_lvl < 1 || cfg->wn_filter_tap_lvl > 3))
        return c->fail(SVTF_E_ARG, "svtf_lr_search_frame: wn_filter_tap_lvl must be 1..3");
    if (sg)
        for (int t = 0; t < 2; t++)
            if (cfg->sg_start_ep[t] < 0 || cfg->sg_end_ep[t] > 16 || cfg->sg_ep_inc[t] < 1)
                return c->fail(SVTF_E_ARG, "svtf_lr_search_frame: self-guided ep range must lie in [0, 16) with a positive increment");
    if (int r = lrs_check(c, "svtf_lr_search_frame"))
        return r;
    const int nplanes = (wn && cfg->wn_use_chroma) || (sg && cfg->sg_use_chroma) ? 3 : 1;
    const int wn_luma = cfg->wn_filter_tap_lvl == 1 ? 7 : (cfg->wn_filter_tap_lvl == 2 ? 5 : 3);
    for (int p = 0; p < 3; p++) {
        if (!lrs_unit_size_ok(p, cfg->restoration_unit_size[p]) || !out_units[p])
            return c->fail(SVTF_E_ARG, "svtf_lr_search_frame: bad restoration unit size or NULL unit array");
        int hu, vu;
        svtf_lr_units(&c->geom, p, cfg->restoration_unit_size[p], &hu, &vu);
        out_frame_type[p] = SVTF_RESTORE_NONE;
        std::memset(out_units[p], 0, sizeof(svtf_lr_unit) * hu * vu);
    }
    if (sg_frame_ep_cnt)
        std::memset(sg_frame_ep_cnt, 0, sizeof(int32_t) * 16);
    CUDA_TRY(c, cudaSetDevice(c->device));
    const bool timing = std::getenv("SVTF_LRS_TIMING") != nullptr;
    auto       now    = [] { return std::chrono::steady_clock::now(); };
    auto       ms     = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) {
        return std::chrono::duration<double, std::milli>(b - a).count();
    };
    const auto t_start = now();
    LrsPlan    plan;
    bool       do_wn[3] = {false, false, false}, do_sg[3] = {false, false, false};
    for (int p = 0; p < nplanes; p++) {
        do_wn[p] = wn && (!p || cfg->wn_use_chroma), do_sg[p] = sg && (!p || cfg->sg_use_chroma);
        lrs_add_plane(c->geom, plan, p, cfg->restoration_unit_size[p], p ? std::min(wn_luma, 5) : wn_luma);
    }
    for (int p = 0; p < nplanes; p++)
        if (do_sg[p])
            plan.nslots = std::max(plan.nslots, (cfg->sg_end_ep[p > 0] - cfg->sg_start_ep[p > 0] + cfg->sg_ep_inc[p > 0] - 1) / cfg->sg_ep_inc[p > 0]);
    const unsigned long long *sse_none;
    const long long          *acc;
    if (int r = lrs_measure(c, plan, do_wn, sse_none, acc))
        return r;
    const auto                       t_measured = now();
    std::vector<svtf_lr_unit_search> us(plan.units);
    std::vector<LrsRefine>           ref(plan.units);
    // the solve of a unit (integer alternating least squares, ~20k multiply-divides for a 7x7 window) is independent of every
    // other unit: a few host threads share them (the reference spreads units over its restoration segments the same way)
    std::vector<int> win_of(plan.units), plane_of(plan.units), first_of(plan.units);
    for (const LrsPlane &pl : plan.planes)
        for (int u = 0; u < pl.hu * pl.vu; u++) win_of[pl.first + u] = pl.win, plane_of[pl.first + u] = pl.plane, first_of[pl.first + u] = pl.first;
    host_parallel_for(plan.units, [&](int u) {
        svtf_lr_unit_search &s = us[u];
        std::memset(&s, 0, sizeof(s));
        s.sse[0] = (int64_t)sse_none[u], s.sse[1] = INT64_MAX;
        if (!do_wn[plane_of[u]])
            return;
        const svtf_lr_unit *prev = cfg->prev_units[plane_of[u]] ? cfg->prev_units[plane_of[u]] + (u - first_of[u]) : nullptr;
        if (prev && prev->restoration_type == SVTF_RESTORE_WIENER) { // use_prev_frame_coeffs :1313-1319: no solve, no score check
            ref[u].start(win_of[u], cfg->wn_use_refinement != 0, cfg->wn_max_one_refinement_step != 0, prev->wiener_vfilter, prev->wiener_hfilter);
            return;
        }
        int64_t M[49], H[49 * 49];
        lrs_finish_stats(c->geom, win_of[u], acc + plan.acc_off[u], M, H);
        int16_t vf[8], hf[8];
        if (svtf_lr_wiener_from_stats(win_of[u], M, H, vf, hf) == 1)
            ref[u].start(win_of[u], cfg->wn_use_refinement != 0, cfg->wn_max_one_refinement_step != 0, vf, hf);
    });
    const auto t_solved = now();
    int        rounds = 0, sg_rounds = 0;
    long long  pending_total = 0; // unit trials over all Wiener rounds (timing output)
    uint8_t   *d = c->d_lrs;
    LrUnitDev *host = reinterpret_cast<LrUnitDev *>(c->h_lrs + plan.o_units);
    const unsigned long long *trial = reinterpret_cast<const unsigned long long *>(c->h_lrs + plan.o_trial);
    // one measuring round: upload the candidates in `host`, zero the counters, run `launch`, fetch the counters
    auto round_trip = [&](auto launch) -> int {
        CUDA_TRY(c, cudaMemcpyAsync(d + plan.o_units, host, sizeof(LrUnitDev) * plan.units, cudaMemcpyHostToDevice, c->stream));
        CUDA_TRY(c, cudaMemsetAsync(d + plan.o_trial, 0, 8 * (size_t)plan.units, c->stream));
        if (int r = launch())
            return r;
        CUDA_TRY(c, cudaMemcpyAsync(c->h_lrs + plan.o_trial, d + plan.o_trial, 8 * (size_t)plan.units, cudaMemcpyDeviceToHost, c->stream));
        CUDA_TRY(c, cudaStreamSynchronize(c->stream));
        return SVTF_OK;
    };
    // ---- Wiener refinement in lock-step: one launch filters every unit that has a trial pending and measures it ----
    if (wn) {
        LrArgs    a;
        const int grid = lrs_trial_args(c, plan, do_wn, SVTF_RESTORE_WIENER, a);
        for (;;) {
            int pending = 0;
            for (int u = 0; u < plan.units; u++) {
                LrUnitDev &h = host[u];
                std::memset(&h, 0, sizeof(h));
                if (!ref[u].pending())
                    continue;
                pending++;
                h.type = SVTF_RESTORE_WIENER;
                for (int t = 0; t < 8; t++) h.vf[t] = ref[u].vf[t], h.hf[t] = ref[u].hf[t];
            }
            if (!pending)
                break;
            rounds++, pending_total += pending;
            if (int r = round_trip([&] { return lrs_launch_lr<LR_SSE>(c, a, grid); }))
                return r;
            for (int u = 0; u < plan.units; u++)
                if (ref[u].pending())
                    ref[u].feed((int64_t)trial[u]);
        }
        for (int u = 0; u < plan.units; u++)
            if (ref[u].started) {
                us[u].sse[1] = ref[u].err;
                std::memcpy(us[u].wiener_vfilter, ref[u].vf, 16), std::memcpy(us[u].wiener_hfilter, ref[u].hf, 16);
            }
    }
    const auto t_wiener = now();
    // ---- self-guided search (search_selfguided_restoration :550): the ep candidates one after the other, all units in lock-step ----
    if (sg) {
        // All parameter sets (ep) at once: slot k of a plane = its k-th candidate ep.  Filter outputs: two int32 planes per
        // searched plane and slot.
        const int    nslots = plan.nslots;
        size_t       slot_elems = 0, off_p[3] = {0, 0, 0}, half[3] = {0, 0, 0};
        const Planes cur = c->surf[c->cur].planes();
        for (int p = 0; p < 3; p++)
            if (do_sg[p]) {
                const int ss = p ? 1 : 0;
                half[p]  = (size_t)cur.pitch[p] * (((c->geom.height + ss) >> ss) + 1);
                off_p[p] = slot_elems, slot_elems += 2 * half[p];
            }
        const size_t need = slot_elems * nslots * 4;
        if (need > c->lrs_flt_cap) {
            if (c->d_lrs_flt) {
                CUDA_TRY(c, cudaStreamSynchronize(c->stream));
                CUDA_TRY(c, cudaFree(c->d_lrs_flt));
                c->d_lrs_flt = nullptr, c->lrs_flt_cap = 0;
            }
            CUDA_TRY(c, cudaMalloc(&c->d_lrs_flt, need));
            c->lrs_flt_cap = need;
        }
        auto ep_at = [&](int p, int k) { return cfg->sg_start_ep[p > 0] + k * cfg->sg_ep_inc[p > 0]; };
        auto act_at = [&](int p, int k) { return do_sg[p] && ep_at(p, k) < cfg->sg_end_ep[p > 0]; };
        // the two guided filters of every unit for every ep + the sums of the projection: one launch per slot, no sync between
        for (int u = 0; u < plan.units; u++) {
            std::memset(&host[u], 0, sizeof(LrUnitDev));
            if (do_sg[plane_of[u]])
                host[u].type = SVTF_RESTORE_SGRPROJ;
        }
        CUDA_TRY(c, cudaMemcpyAsync(d + plan.o_units, host, sizeof(LrUnitDev) * plan.units, cudaMemcpyHostToDevice, c->stream));
        CUDA_TRY(c, cudaMemsetAsync(d + plan.o_proj, 0, 40 * (size_t)plan.units * nslots, c->stream));
        for (int k = 0; k < nslots; k++) {
            bool act[3];
            for (int p = 0; p < 3; p++) act[p] = act_at(p, k);
            LrArgs    fa;
            const int grid = lrs_trial_args(c, plan, act, 3 /* any restoring frame type */, fa);
            for (int p = 0; p < 3; p++) {
                fa.flt[p] = c->d_lrs_flt + k * slot_elems + off_p[p], fa.flt_half[p] = half[p], fa.force_ep[p] = ep_at(p, k);
                if (fa.proj[p])
                    fa.proj[p] += 5 * (size_t)plan.units * k;
            }
            if (int r = lrs_launch_lr<LR_FLT>(c, fa, grid))
                return r;
        }
        CUDA_TRY(c, cudaMemcpyAsync(c->h_lrs + plan.o_proj, d + plan.o_proj, 40 * (size_t)plan.units * nslots, cudaMemcpyDeviceToHost, c->stream));
        CUDA_TRY(c, cudaStreamSynchronize(c->stream));
        const long long         *proj = reinterpret_cast<const long long *>(c->h_lrs + plan.o_proj);
        std::vector<LrsSgRefine> sref((size_t)plan.units * nslots);
        for (int k = 0; k < nslots; k++)
            for (int u = 0; u < plan.units; u++)
                if (act_at(plane_of[u], k)) {
                    int       xqd[2];
                    const int ep = ep_at(plane_of[u], k);
                    lrs_sg_initial_xqd(ep, proj + 5 * ((size_t)k * plan.units + u), plan.count[u], xqd);
                    sref[(size_t)k * plan.units + u].start(ep, cfg->sg_refine[plane_of[u] > 0] != 0, xqd);
                }
        // The projection error of the pending candidates, every (unit, ep) in lock-step.  A (unit, ep) brings the candidates its
        // line search asks for next IF each one lowers the error (a copy of its state machine fed with "no worse"); the real
        // errors are then fed in order for as long as the machine asks for exactly the candidate that was measured.
        LrsProjArgs pa;
        pa.cur = cur, pa.src = c->source.planes();
        pa.tiles = reinterpret_cast<const LrsTile *>(d + plan.o_tiles);
        pa.cand  = reinterpret_cast<const LrsCand *>(d + plan.o_cand);
        pa.err   = reinterpret_cast<unsigned long long *>(d + plan.o_errk);
        pa.units = plan.units, pa.slot_elems = slot_elems;
        for (int p = 0; p < 3; p++) pa.flt[p] = c->d_lrs_flt + off_p[p], pa.flt_half[p] = half[p];
        LrsCand                  *cand = reinterpret_cast<LrsCand *>(c->h_lrs + plan.o_cand);
        const unsigned long long *errk = reinterpret_cast<const unsigned long long *>(c->h_lrs + plan.o_errk);
        const int    nt = (int)plan.tiles.size();
        const size_t nu = (size_t)plan.units * nslots;
        for (;;) {
            int pending = 0;
            for (size_t i = 0; i < nu; i++) {
                LrsCand &cd = cand[i];
                cd.n = 0, cd.ep = 0;
                if (!sref[i].pending())
                    continue;
                pending++;
                cd.ep = ep_at(plane_of[i % plan.units], (int)(i / plan.units));
                LrsSgRefine sim = sref[i];
                while (sim.pending() && cd.n < LRS_CAND) {
                    cd.xqd[cd.n][0] = sim.xqd[0], cd.xqd[cd.n][1] = sim.xqd[1];
                    cd.n++;
                    sim.feed(sim.started_err() ? sim.err : 0);
                }
            }
            if (!pending)
                break;
            sg_rounds++;
            CUDA_TRY(c, cudaMemcpyAsync(d + plan.o_cand, cand, sizeof(LrsCand) * nu, cudaMemcpyHostToDevice, c->stream));
            CUDA_TRY(c, cudaMemsetAsync(d + plan.o_errk, 0, 8 * (size_t)LRS_CAND * nu, c->stream));
            if (c->geom.sample_bytes == 1)
                lrs_proj_err_kernel<uint8_t><<<dim3(nt, nslots), LRS_THREADS, 0, c->stream>>>(pa);
            else
                lrs_proj_err_kernel<uint16_t><<<dim3(nt, nslots), LRS_THREADS, 0, c->stream>>>(pa);
            CUDA_TRY(c, cudaGetLastError());
            c->launches++;
            CUDA_TRY(c, cudaMemcpyAsync(c->h_lrs + plan.o_errk, d + plan.o_errk, 8 * (size_t)LRS_CAND * nu, cudaMemcpyDeviceToHost, c->stream));
            CUDA_TRY(c, cudaStreamSynchronize(c->stream));
            for (size_t i = 0; i < nu; i++)
                for (int j = 0; j < cand[i].n && sref[i].pending() && sref[i].xqd[0] == cand[i].xqd[j][0] && sref[i].xqd[1] == cand[i].xqd[j][1]; j++)
                    sref[i].feed((int64_t)errk[i * LRS_CAND + j]);
        }
        // the best ep of every unit: first lowest error in ep order (:637-642)
        for (int u = 0; u < plan.units; u++) {
            int64_t best = -1;
            for (int k = 0; k < nslots; k++) {
                const LrsSgRefine &r = sref[(size_t)k * plan.units + u];
                if (act_at(plane_of[u], k) && (best == -1 || r.err < best))
                    best = r.err, us[u].sgr_ep = ep_at(plane_of[u], k), us[u].sgr_xqd[0] = r.xqd[0], us[u].sgr_xqd[1] = r.xqd[1];
            }
        }
        // the unit filtered for real with the parameters found (clipping included): sse[RESTORE_SGRPROJ]
        LrArgs    a;
        const int grid = lrs_trial_args(c, plan, do_sg, SVTF_RESTORE_SGRPROJ, a);
        for (int u = 0; u < plan.units; u++) {
            std::memset(&host[u], 0, sizeof(LrUnitDev));
            if (!do_sg[plane_of[u]])
                continue;
            host[u].type = SVTF_RESTORE_SGRPROJ, host[u].ep = us[u].sgr_ep, host[u].xqd[0] = us[u].sgr_xqd[0], host[u].xqd[1] = us[u].sgr_xqd[1];
            if (sg_frame_ep_cnt)
                sg_frame_ep_cnt[us[u].sgr_ep]++;
        }
        if (int r = round_trip([&] { return lrs_launch_lr<LR_SSE>(c, a, grid); }))
            return r;
        for (int u = 0; u < plan.units; u++)
            if (do_sg[plane_of[u]])
                us[u].sse[2] = (int64_t)trial[u];
    }
    if (timing)
        std::fprintf(stderr, "svtf_lr_search_frame: plan+stats %.3f ms, host solve %.3f ms, %d Wiener rounds (%lld unit trials of %d units) %.3f ms, "
                     "self-guided (%d rounds) %.3f ms\n",
                     ms(t_start, t_measured), ms(t_measured, t_solved), rounds, pending_total, plan.units, ms(t_solved, t_wiener), sg_rounds,
                     ms(t_wiener, now()));
    // ---- rest_finish_search :1555: the cost of every frame type that applies; per-unit decisions in raster order.  The
    //      reference's per-unit array is allocated and zeroed once BEFORE the plane loop (:1570-1577), so what a plane does not
    //      refresh (e.g. the Wiener entries of a chroma plane searched with the self-guided tools only) still holds the previous
    //      plane's values when search_switchable reads it — kept as is. ----
    struct FinishUnit {
        int64_t sse[3];
        int16_t vf[8], hf[8];
        int     sgr_ep, sgr_xqd[2];
        int     best[3]; // best_rtype[WIENER - 1], [SGRPROJ - 1], [SWITCHABLE - 1]
    };
    size_t most = 0; // the reference sizes the array for luma and asserts that chroma has no more units (:1567); be safe anyway
    for (const LrsPlane &pl : plan.planes) most = std::max(most, (size_t)pl.hu * pl.vu);
    std::vector<FinishUnit> fu(most);
    std::memset(fu.data(), 0, sizeof(FinishUnit) * fu.size());
    const int force = wn ? (sg ? 4 : 1) : (sg ? 2 : 0); // force_restore_type_d; 4 = RESTORE_TYPES: all
    for (const LrsPlane &pl : plan.planes) {
        const int p = pl.plane, n = pl.hu * pl.vu;
        double    best_cost = 0;
        int       best_type = SVTF_RESTORE_NONE;
        for (int r = 0; r < (n > 1 ? 4 : 3); r++) {
            if (force != 4 && r != 0 && r != force)
                continue;
            if (p && ((r == 1 && !cfg->wn_use_chroma) || (r == 2 && !cfg->sg_use_chroma)))
                continue;
            int64_t sse = 0, bits = 0;
            int16_t rvf[8] = {3, -7, 15, -22, 15, -7, 3, 0}, rhf[8] = {3, -7, 15, -22, 15, -7, 3, 0}; // set_default_wiener
            int     rsg[2] = {-32, 31};                                                              // set_default_sgrproj
            for (int u = 0; u < n; u++) {
                FinishUnit                &f = fu[u];
                const svtf_lr_unit_search &s = us[pl.first + u];
                if (r == 0) { // search_norestore_finish
                    f.sse[0] = s.sse[0];
                    sse += f.sse[0];
                } else if (r == 1) { // search_wiener_finish :1380
                    const int64_t bits_none = cfg->wiener_restore_cost[0];
                    f.sse[1]                = s.sse[1];
                    if (f.sse[1] == INT64_MAX) {
                        bits += bits_none, sse += f.sse[0], f.best[0] = SVTF_RESTORE_NONE;
                        continue;
                    }
                    std::memcpy(f.vf, s.wiener_vfilter, 16), std::memcpy(f.hf, s.wiener_hfilter, 16);
                    const int64_t bw = cfg->wiener_restore_cost[1] + ((int64_t)lrs_wiener_bits(pl.win, f.vf, f.hf, rvf, rhf) << 9);
                    const double  cn = lrs_rdcost(cfg->rdmult, bits_none >> 4, f.sse[0]), cw = lrs_rdcost(cfg->rdmult, bw >> 4, f.sse[1]);
                    f.best[0]        = cw < cn ? SVTF_RESTORE_WIENER : SVTF_RESTORE_NONE;
                    sse += f.sse[f.best[0]], bits += cw < cn ? bw : bits_none;
                    if (cw < cn)
                        std::memcpy(rvf, f.vf, 16), std::memcpy(rhf, f.hf, 16);
                } else if (r == 2) { // search_sgrproj_finish :1267
                    f.sse[2] = s.sse[2], f.sgr_ep = s.sgr_ep, f.sgr_xqd[0] = s.sgr_xqd[0], f.sgr_xqd[1] = s.sgr_xqd[1];
                    const int64_t bits_none = cfg->sgrproj_restore_cost[0];
                    const int64_t bs = cfg->sgrproj_restore_cost[1] + ((int64_t)lrs_sgr_bits(f.sgr_ep, f.sgr_xqd, rsg) << 9);
                    const double  cn = lrs_rdcost(cfg->rdmult, bits_none >> 4, f.sse[0]), cs = lrs_rdcost(cfg->rdmult, bs >> 4, f.sse[2]);
                    f.best[1]        = cs < cn ? SVTF_RESTORE_SGRPROJ : SVTF_RESTORE_NONE;
                    sse += f.sse[f.best[1]], bits += cs < cn ? bs : bits_none;
                    if (cs < cn)
                        rsg[0] = f.sgr_xqd[0], rsg[1] = f.sgr_xqd[1];
                } else { // search_switchable :1148 (it counts Wiener bits with the full window of the plane type)
                    double  bc = 0;
                    int64_t bb = 0;
                    int     bt = 0;
                    for (int t = 0; t < 3; t++) {
                        if (t > 0 && f.best[t - 1] == SVTF_RESTORE_NONE)
                            continue;
                        const int64_t coeff = t == 1 ? lrs_wiener_bits(p ? 5 : 7, f.vf, f.hf, rvf, rhf) : (t == 2 ? lrs_sgr_bits(f.sgr_ep, f.sgr_xqd, rsg) : 0);
                        const int64_t b     = cfg->switchable_restore_cost[t] + (coeff << 9);
                        const double  cost  = lrs_rdcost(cfg->rdmult, b >> 4, f.sse[t]);
                        if (t == 0 || cost < bc)
                            bc = cost, bb = b, bt = t;
                    }
                    f.best[2] = bt;
                    sse += f.sse[bt], bits += bb;
                    if (bt == 1)
                        std::memcpy(rvf, f.vf, 16), std::memcpy(rhf, f.hf, 16);
                    if (bt == 2)
                        rsg[0] = f.sgr_xqd[0], rsg[1] = f.sgr_xqd[1];
                }
            }
            const double cost = lrs_rdcost(cfg->rdmult, bits >> 4, sse);
            if (r == 0 || cost < best_cost)
                best_cost = cost, best_type = r;
        }
        out_frame_type[p] = best_type;
        if (best_type != SVTF_RESTORE_NONE)
            for (int u = 0; u < n; u++) { // copy_unit_info :1202
                svtf_lr_unit &o    = out_units[p][u];
                o.restoration_type = fu[u].best[best_type - 1];
                if (o.restoration_type == SVTF_RESTORE_WIENER)
                    std::memcpy(o.wiener_vfilter, fu[u].vf, 16), std::memcpy(o.wiener_hfilter, fu[u].hf, 16);
                else if (o.restoration_type == SVTF_RESTORE_SGRPROJ)
                    o.sgr_ep = fu[u].sgr_ep, o.sgr_xqd[0] = fu[u].sgr_xqd[0], o.sgr_xqd[1] = fu[u].sgr_xqd[1];
            }
        if (out_search && out_search[p])
            std::memcpy(out_search[p], us.data() + pl.first, sizeof(svtf_lr_unit_search) * n);
    }
    return SVTF_OK;
}
