/*
 * svtf_oracle_ccso.c — CPU restatement of the CCSO filter derivation for ONE plane.
 * TEST INFRASTRUCTURE ONLY (tests/, __graft_entry__.smoke(), bench.py's cpu_baseline leg); the product never links it.
 * Parity: PINNED against the reference's own derive_ccso_filter compiled into oracle/_ref/libsvtref.so
 * (tests/test_ccso_derive_oracle.py).
 *
 * Follows Encoder/Codec/EbPickccso.c:
 *   derive_ccso_filter          :464-782   the search over bo_only x support x quant x edge_clf x band_log2
 *   ccso_derive_src_info        :125-160   class planes            (here: classes computed per sample)
 *   ccso_pre_compute_class_err  :382-429   per-block histograms    (and _bo :162-198)
 *   ccso_compute_class_err      :202-235   fold over enabled blocks
 *   derive_lut_offset           :431-461   float mean -> nearest of {-10,-7,-3,-1,0,1,3,7}
 *   ccso_try_{luma,chroma}_filter :238-357 + ccso_filter_block_hbd_with_buf_c (Common/Codec/EbCcso.c:6-38)
 *   compute_distortion          :55-68 + compute_distortion_block_c (Common/Codec/EbCcso.c:64-90)
 *   derive_blk_md               :71-121    per-block on/off (its CDF rate is computed but never used by the caller)
 *   count_lut_bits              :360-379
 * Classifier taps: cal_filter_support / derive_ccso_sample_pos (Common/Codec/EbCcso.c:204-259); the 5-sample
 * replicate border of extend_ccso_border (:185-201) is realised by clamping tap coordinates to the picture.
 */
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../include/svtfilt_b200.h"

#define ORC_API __attribute__((visibility("default")))

static const int     k_offset[8]  = {-10, -7, -3, -1, 0, 1, 3, 7};
static const uint8_t k_quant[4]   = {16, 8, 32, 64};
static const int     k_interval[2] = {3, 2};

static inline int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

typedef struct {
    int             W, H;   /* luma picture */
    int             pw, ph; /* plane */
    int             ss, bd, blk_log2, nbx, nby, nfb;
    const uint16_t *luma;   /* W x H, stride W: pre-CDEF luma */
    const uint16_t *rec, *org; /* plane, stride W (the reference keeps every plane at luma stride) */
} Ctx;

static inline int luma_at(const Ctx *c, int x, int y) { return c->luma[clampi(y, 0, c->H - 1) * c->W + clampi(x, 0, c->W - 1)]; }

static void tap_of(int support, int *dy, int *dx) {
    static const int t[6][2] = {{-1, 0}, {-1, -1}, {0, -1}, {1, -1}, {0, -3}, {0, -5}};
    *dy = t[support][0], *dx = t[support][1];
}

/* cal_filter_support for the luma sample co-located with plane sample (x, y) */
static inline void classes_at(const Ctx *c, int x, int y, int support, int thr, int edge_clf, int *c0, int *c1) {
    int dy, dx;
    tap_of(support, &dy, &dx);
    const int lx = x << c->ss, ly = y << c->ss;
    const int yv = luma_at(c, lx, ly);
    const int d0 = luma_at(c, lx + dx, ly + dy) - yv, d1 = luma_at(c, lx - dx, ly - dy) - yv;
    if (edge_clf == 0) {
        *c0 = d0 > thr ? 2 : (d0 < -thr ? 0 : 1);
        *c1 = d1 > thr ? 2 : (d1 < -thr ? 0 : 1);
    } else {
        *c0 = d0 < -thr ? 0 : 1;
        *c1 = d1 < -thr ? 0 : 1;
    }
}

static void block_sse(const Ctx *c, const uint16_t *rec, uint64_t *blk, uint64_t *total) {
    memset(blk, 0, sizeof(uint64_t) * c->nfb);
    for (int y = 0; y < c->ph; y++)
        for (int x = 0; x < c->pw; x++) {
            const int e = (int)c->org[y * c->W + x] - (int)rec[y * c->W + x];
            blk[(y >> c->blk_log2) * c->nbx + (x >> c->blk_log2)] += (uint64_t)(e * e);
        }
    *total = 0;
    for (int i = 0; i < c->nfb; i++) *total += blk[i];
}

static int lut_bits(const int8_t *lut, int band_log2, int intervals, int bo) {
    static const int reordered[8] = {0, 1, -1, 3, -3, 7, -7, -10};
    int bits = 0, n = bo ? 1 : intervals;
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

static void lut_from_means(int8_t *lut, const int *err, const int *cnt, int band_log2, int intervals, int bo) {
    int n = bo ? 1 : intervals;
    for (int d0 = 0; d0 < n; d0++)
        for (int d1 = 0; d1 < n; d1++)
            for (int b = 0; b < (1 << band_log2); b++) {
                const int i = (b << 4) + (d0 << 2) + d1;
                if (!cnt[i])
                    continue;
                const float t = (float)err[i] / cnt[i];
                if (t < k_offset[0] || t >= k_offset[7]) {
                    lut[i] = (int8_t)clampi((int)t, k_offset[0], k_offset[7]);
                } else {
                    for (int k = 0; k < 7; k++)
                        if (t >= k_offset[k] && t <= k_offset[k + 1]) {
                            lut[i] = (int8_t)(fabs(t - k_offset[k]) > fabs(t - k_offset[k + 1]) ? k_offset[k + 1] : k_offset[k]);
                            break;
                        }
                }
            }
}

static double rd_cost(int rdmult, int bits, uint64_t dist, int bd) { /* RDCOST_DBL_WITH_NATIVE_BD_DIST1 */
    return (((double)bits) * rdmult) / (double)(1 << 9) + ((double)(dist >> (2 * (bd - 8))) * (1 << 7));
}

/* One plane.  luma/rec/org: uint16 samples at stride g->width.  Outputs the plane's entries of `out`
 * (ccso_enable, bo_only, quant_idx, ext_filter_support, edge_clf, max_band_log2, filter_offset) and flags[nfb]. */
ORC_API int orc_ccso_derive_plane(const svtf_geom *g, int plane, const uint16_t *luma, const uint16_t *rec, const uint16_t *org, int rdmult,
                                  svtf_ccso_params *out, uint8_t *flags) {
    Ctx c;
    c.W = g->width, c.H = g->height, c.ss = plane ? 1 : 0;
    c.pw = g->width >> c.ss, c.ph = g->height >> c.ss, c.bd = g->bit_depth;
    c.blk_log2 = plane ? 7 : 8;
    c.nbx = (c.pw + (1 << c.blk_log2) - 1) >> c.blk_log2, c.nby = (c.ph + (1 << c.blk_log2) - 1) >> c.blk_log2;
    c.nfb = c.nbx * c.nby;
    c.luma = luma, c.rec = rec, c.org = org;
    const int  nfb = c.nfb, max_val = (1 << c.bd) - 1;
    uint64_t  *unf = malloc(sizeof(uint64_t) * nfb), *trn = malloc(sizeof(uint64_t) * nfb);
    uint8_t   *ctl = malloc(nfb), *best_ctl = calloc(nfb, 1), *final_ctl = calloc(nfb, 1);
    uint16_t  *tmp = malloc(sizeof(uint16_t) * (size_t)c.W * c.ph);
    int       *herr = malloc(sizeof(int) * 2048 * (size_t)nfb), *hcnt = malloc(sizeof(int) * 2048 * (size_t)nfb);
    uint8_t   *cls = malloc((size_t)c.pw * c.ph * 2);
    uint64_t   unf_total = 0;
    block_sse(&c, rec, unf, &unf_total);
    const double unfiltered_cost = rd_cost(rdmult, 1, unf_total, c.bd);
    double       final_cost = DBL_MAX;
    int8_t       best_lut[2048] = {0}, final_lut[2048] = {0};
    int          best_clf = 0, final_clf = 0, final_q = 0, final_sup = 0, final_bo = 0, final_band = 0;
    const int    frame_bits = 10, frame_bits_bo = 5;

    for (int bo = 0; bo < 2; bo++)
        for (int sup = 0; sup < (bo ? 1 : 6); sup++)
            for (int q = 0; q < (bo ? 1 : 4); q++)
                for (int clf = 0; clf < (bo ? 1 : 2); clf++) {
                    const int intervals = k_interval[clf];
                    if (!bo)
                        for (int y = 0; y < c.ph; y++)
                            for (int x = 0; x < c.pw; x++) {
                                int c0, c1;
                                classes_at(&c, x, y, sup, k_quant[q], clf, &c0, &c1);
                                cls[2 * (y * c.pw + x)] = (uint8_t)c0, cls[2 * (y * c.pw + x) + 1] = (uint8_t)c1;
                            }
                    for (int band_log2 = 0; band_log2 < (bo ? 8 : 4); band_log2++) {
                        const int shift = c.bd - band_log2;
                        double    best_cost = DBL_MAX, prev_cost = DBL_MAX;
                        int       enable = 1, keep = 1, iter = 0;
                        int8_t    lut[2048];
                        memset(lut, 0, sizeof(lut)); /* (the reference's array is uninitialised here; it is zeroed in iteration 0) */
                        memset(ctl, 1, nfb);
                        memset(herr, 0, sizeof(int) * 2048 * (size_t)nfb), memset(hcnt, 0, sizeof(int) * 2048 * (size_t)nfb);
                        for (int y = 0; y < c.ph; y++)
                            for (int x = 0; x < c.pw; x++) {
                                const int band = luma_at(&c, x << c.ss, y << c.ss) >> shift;
                                const int c0 = bo ? 0 : cls[2 * (y * c.pw + x)], c1 = bo ? 0 : cls[2 * (y * c.pw + x) + 1];
                                const int fb = (y >> c.blk_log2) * c.nbx + (x >> c.blk_log2);
                                const int i  = (band << 4) + (c0 << 2) + c1;
                                herr[(size_t)fb * 2048 + i] += (int)org[y * c.W + x] - (int)rec[y * c.W + x];
                                hcnt[(size_t)fb * 2048 + i]++;
                            }
                        while (keep) {
                            int improvement = 0;
                            if (enable) {
                                int err[2048] = {0}, cnt[2048] = {0};
                                memset(lut, 0, sizeof(lut));
                                for (int fb = 0; fb < nfb; fb++) {
                                    if (!ctl[fb])
                                        continue;
                                    const int n = bo ? 1 : intervals;
                                    for (int d0 = 0; d0 < n; d0++)
                                        for (int d1 = 0; d1 < n; d1++)
                                            for (int b = 0; b < (1 << band_log2); b++) {
                                                const int i = (b << 4) + (d0 << 2) + d1;
                                                err[i] += herr[(size_t)fb * 2048 + i], cnt[i] += hcnt[(size_t)fb * 2048 + i];
                                            }
                                }
                                lut_from_means(lut, err, cnt, band_log2, intervals, bo);
                            }
                            /* trial filter of the whole plane (every block, whatever its control flag) */
                            for (int y = 0; y < c.ph; y++)
                                for (int x = 0; x < c.pw; x++) {
                                    const int band = luma_at(&c, x << c.ss, y << c.ss) >> shift;
                                    const int c0 = bo ? 0 : cls[2 * (y * c.pw + x)], c1 = bo ? 0 : cls[2 * (y * c.pw + x) + 1];
                                    tmp[y * c.W + x] = (uint16_t)clampi(lut[(band << 4) + (c0 << 2) + c1] + (int)rec[y * c.W + x], 0, max_val);
                                }
                            uint64_t trn_total = 0;
                            block_sse(&c, tmp, trn, &trn_total);
                            /* derive_blk_md */
                            uint64_t cur_dist = 0;
                            if (enable) {
                                int any = 0;
                                for (int fb = 0; fb < nfb; fb++) {
                                    const int on = trn[fb] < unf[fb];
                                    ctl[fb]      = (uint8_t)on;
                                    cur_dist += on ? trn[fb] : unf[fb];
                                    any |= on;
                                }
                                enable = any;
                            }
                            if (enable) {
                                const int    bits = lut_bits(lut, band_log2, intervals, bo) + (bo ? frame_bits_bo : frame_bits) + nfb;
                                const double cost = rd_cost(rdmult, bits, cur_dist, c.bd);
                                if (cost < prev_cost)
                                    prev_cost = cost, improvement = 1;
                                if (cost < best_cost) {
                                    best_cost = cost;
                                    memcpy(best_lut, lut, sizeof(lut));
                                    best_clf = clf;
                                    memcpy(best_ctl, ctl, nfb);
                                }
                            }
                            iter++;
                            if (!improvement || iter > 15)
                                keep = 0;
                        }
                        if (best_cost < final_cost) {
                            final_cost = best_cost;
                            final_q = q, final_sup = sup, final_bo = bo, final_band = band_log2, final_clf = best_clf;
                            memcpy(final_lut, best_lut, sizeof(best_lut));
                            memcpy(final_ctl, best_ctl, nfb);
                        }
                    }
                }
    const int on = !(unfiltered_cost < final_cost);
    out->ccso_enable[plane] = (uint8_t)on;
    memset(flags, 0, nfb);
    if (on) {
        memcpy(flags, final_ctl, nfb);
        memcpy(out->filter_offset[plane], final_lut, 2048);
        out->quant_idx[plane] = (uint8_t)final_q, out->ext_filter_support[plane] = (uint8_t)final_sup;
        out->ccso_bo_only[plane] = (uint8_t)final_bo, out->max_band_log2[plane] = (uint8_t)final_band;
        out->edge_clf[plane] = (uint8_t)final_clf;
    }
    free(unf), free(trn), free(ctl), free(best_ctl), free(final_ctl), free(tmp), free(herr), free(hcnt), free(cls);
    return 0;
}

/* ---------------------------------------------------------------------------------------------------------
 * Deblocking level search — svt_av1_pick_filter_level(LPF_PICK_FROM_FULL_IMAGE), Encoder/Codec/EbDeblockingFilter.c:
 *   search_filter_level :886-985, try_filter_frame :841-880, picture_sse_calculations :716-835.
 * Pinned against the reference's own static search_filter_level (oracle/ref_shim_dlf.c includes the file).
 * ------------------------------------------------------------------------------------------------------- */
int orc_deblock_frame(const svtf_geom *g, void *y, int sy, void *cb, int scb, void *cr, int scr, const svtf_mi_t *mi, int mi_stride,
                      const svtf_lf_params *lf, int plane_start, int plane_end);

typedef struct {
    const svtf_geom *g;
    void            *rec[3];
    const void      *src[3];
    int              stride[3], sstride[3];
    const svtf_mi_t *mi;
    int              mi_stride;
} LfS;

static int64_t lfs_try(const LfS *s, const svtf_lf_params *lf, int plane) {
    const svtf_geom *g  = s->g;
    const int        sb = g->sample_bytes, ss = plane ? 1 : 0;
    const int        ch = (g->mi_rows * 4) >> ss;
    void            *tmp[3];
    for (int p = 0; p < 3; p++) { /* trial on a copy: the picture itself stays unfiltered */
        const size_t n = (size_t)s->stride[p] * ((g->mi_rows * 4) >> (p ? 1 : 0)) * sb;
        tmp[p]         = malloc(n);
        memcpy(tmp[p], s->rec[p], n);
    }
    orc_deblock_frame(g, tmp[0], s->stride[0], tmp[1], s->stride[1], tmp[2], s->stride[2], s->mi, s->mi_stride, lf, plane, plane + 1);
    int w, h;
    if (sb == 1)
        w = (g->mi_cols * 4) >> ss, h = ch;
    else
        w = (g->width + ss) >> ss, h = (g->height + ss) >> ss;
    int64_t e = 0;
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++) {
            const int64_t a = sb == 1 ? ((const uint8_t *)s->src[plane])[(size_t)y * s->sstride[plane] + x]
                                      : ((const uint16_t *)s->src[plane])[(size_t)y * s->sstride[plane] + x];
            const int64_t b = sb == 1 ? ((const uint8_t *)tmp[plane])[(size_t)y * s->stride[plane] + x]
                                      : ((const uint16_t *)tmp[plane])[(size_t)y * s->stride[plane] + x];
            e += (a - b) * (a - b);
        }
    for (int p = 0; p < 3; p++) free(tmp[p]);
    return e;
}

static int lfs_plane(const LfS *s, svtf_lf_params *lf, const svtf_lf_search_cfg *cfg, const int32_t last[4], int plane) {
    const int lvl      = plane == 0 ? (cfg->dlf_avg ? last[0] : last[2]) : last[plane + 1]; /* dir = 2 for luma, :903-909 */
    int       filt_mid = clampi(lvl, 0, 63), step = filt_mid < 16 ? 4 : filt_mid / 4;
    int64_t   ss_err[64];
    for (int i = 0; i < 64; i++) ss_err[i] = -1;
#define LFS_TRY(level, dst)                                           \
    do {                                                              \
        if (plane == 0)                                               \
            lf->filter_level[0] = lf->filter_level[1] = (level);      \
        else if (plane == 1)                                          \
            lf->filter_level_u = (level);                             \
        else                                                          \
            lf->filter_level_v = (level);                             \
        (dst) = lfs_try(s, lf, plane);                                \
    } while (0)
    int64_t best_err;
    LFS_TRY(filt_mid, best_err);
    int filt_best = filt_mid, direction = 0, convergence = 0;
    ss_err[filt_mid] = best_err;
    while (step > 0) {
        const int filt_high = filt_mid + step < 63 ? filt_mid + step : 63, filt_low = filt_mid - step > 0 ? filt_mid - step : 0;
        int64_t   bias      = (best_err >> (15 - (filt_mid / 8))) * step;
        if (!cfg->only_4x4)
            bias >>= 1;
        if (direction <= 0 && filt_low != filt_mid) {
            if (ss_err[filt_low] < 0)
                LFS_TRY(filt_low, ss_err[filt_low]);
            if (ss_err[filt_low] < best_err + bias) {
                if (ss_err[filt_low] < best_err)
                    best_err = ss_err[filt_low];
                filt_best = filt_low;
            }
        }
        if (direction >= 0 && filt_high != filt_mid) {
            if (ss_err[filt_high] < 0)
                LFS_TRY(filt_high, ss_err[filt_high]);
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
#undef LFS_TRY
    return filt_best;
}

ORC_API int orc_lf_search_frame(const svtf_geom *g, void *y, int sy, void *cb, int scb, void *cr, int scr, const void *src_y, int ssy,
                                const void *src_cb, int sscb, const void *src_cr, int sscr, const svtf_mi_t *mi, int mi_stride,
                                const svtf_lf_params *start, const svtf_lf_search_cfg *cfg, int32_t out_levels[4]) {
    LfS s = {g, {y, cb, cr}, {src_y, src_cb, src_cr}, {sy, scb, scr}, {ssy, sscb, sscr}, mi, mi_stride};
    svtf_lf_params lf = *start;
    const int32_t  last[4] = {start->filter_level[0], start->filter_level[1], start->filter_level_u, start->filter_level_v};
    int            u = last[2], v = last[3];
    const int      yl = lfs_plane(&s, &lf, cfg, last, 0);
    lf.filter_level[0] = lf.filter_level[1] = yl;
    if (cfg->search_chroma) {
        u                 = lfs_plane(&s, &lf, cfg, last, 1);
        lf.filter_level_u = u;
        v                 = lfs_plane(&s, &lf, cfg, last, 2);
    }
    out_levels[0] = out_levels[1] = yl, out_levels[2] = u, out_levels[3] = v;
    return 0;
}
