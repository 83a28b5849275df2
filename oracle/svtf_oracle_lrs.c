/*
 * svtf_oracle_lrs.c — CPU restatement of the reference's loop-restoration filter SEARCH with the Wiener tools.
 * TEST INFRASTRUCTURE ONLY: only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may use it.
 *
 * Follows Encoder/Codec/EbRestorationPick.c:
 *   restoration_seg_search :1471 (one segment = the frame) and rest_finish_search :1555, with
 *   search_norestore_seg :1432, search_wiener_seg :1296, search_wiener_finish :1380, search_rest_type_finish :1456;
 *   svt_av1_compute_stats_c / _highbd_c :671 / :708 (find_average[_highbd], EbRestorationPick.h:24-43);
 *   wiener_decompose_sep_sym :906 (update_a_sep_sym :805, update_b_sep_sym :857, linsolve_wiener :766);
 *   finalize_sym_filter :977, compute_score :937, finer_tile_search_wiener_seg :1042, count_wiener_bits :1008 with
 *   svt_aom_count_primitive_refsubexpfin (Encoder/Codec/EbEntropyCoding.c:2876-3022);
 *   try_restoration_unit_seg :129 = svtf_oracle_lr.c's orc_lr_try_unit (filter without stripe boundaries + SSE).
 * Units: foreach_rest_unit_in_tile (Common/Codec/EbRestoration.c:1250) — raster order, the last unit of a row / column
 * takes the remainder (up to 1.5 units), rows shifted up by 8 luma samples.
 * Pinned against the reference's own code through oracle/ref_shim_lrs.c (tests/test_lrs_oracle.py).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../include/svtfilt_b200.h"

#define ORC_API __attribute__((visibility("default")))
#define IMIN(a, b) ((a) < (b) ? (a) : (b))
#define IMAX(a, b) ((a) > (b) ? (a) : (b))

int     orc_lr_units(const svtf_geom *g, int plane, int unit_size, int *horz, int *vert);
int64_t orc_lr_try_unit(const void *cur, int cst, const void *src, int sst, int sb, int bd, int ss, int pw, int ph, int h0, int h1, int v0,
                        int v1, const svtf_lr_unit *u);

enum { WIN = 7, HALF = 3, WIN2 = 49, STEP = 128 /* WIENER_FILT_STEP */, ITERS = 5 /* NUM_WIENER_ITERS */ };
static const int64_t SCALE = (int64_t)1 << 16; /* WIENER_TAP_SCALE_FACTOR */
/* WIENER_FILT_TAP{0,1,2}_{MINV,MAXV,MIDV}, _SUBEXP_K (Common/Codec/EbRestoration.h:131-153) */
static const int tap_min[3] = {-5, -23, -17}, tap_max[3] = {10, 8, 46}, tap_mid[4] = {3, -7, 15, 106 /* STEP - 2 * (3 - 7 + 15) */}, tap_k[3] = {1, 2, 3};

static int px(const void *p, int st, int sb, int x, int y) {
    return sb == 1 ? ((const uint8_t *)p)[(size_t)y * st + x] : ((const uint16_t *)p)[(size_t)y * st + x];
}
static int pxc(const void *p, int st, int sb, int pw, int ph, int x, int y) { /* the extended frame: edges replicated */
    return px(p, st, sb, IMIN(IMAX(x, 0), pw - 1), IMIN(IMAX(y, 0), ph - 1));
}

/* M[win2], H[win2][win2] of one unit */
ORC_API void orc_lr_wiener_stats(const void *dgd, int dst, const void *src, int sst, int sb, int bd, int pw, int ph, int h0, int h1, int v0,
                                 int v1, int win, int64_t *M, int64_t *H) {
    const int win2 = win * win, hw = win >> 1;
    uint64_t  sum = 0;
    for (int i = v0; i < v1; i++)
        for (int j = h0; j < h1; j++) sum += (uint64_t)px(dgd, dst, sb, j, i);
    const int avg = (int)(sum / (uint64_t)((v1 - v0) * (h1 - h0)));
    memset(M, 0, sizeof(*M) * win2), memset(H, 0, sizeof(*H) * win2 * win2);
    int y[WIN2];
    for (int i = v0; i < v1; i++)
        for (int j = h0; j < h1; j++) {
            const int x = px(src, sst, sb, j, i) - avg;
            for (int c = 0; c < win; c++) /* column offset outer, row offset inner */
                for (int r = 0; r < win; r++) y[c * win + r] = pxc(dgd, dst, sb, pw, ph, j + c - hw, i + r - hw) - avg;
            for (int k = 0; k < win2; k++) {
                M[k] += (int64_t)y[k] * x;
                for (int l = k; l < win2; l++) H[k * win2 + l] += (int64_t)y[k] * y[l];
            }
        }
    const int div = sb == 1 ? 1 : (bd == 12 ? 16 : (bd == 10 ? 4 : 1)); /* bit_depth_divider, high-bitdepth path only */
    for (int k = 0; k < win2; k++) {
        M[k] /= div;
        for (int l = k; l < win2; l++) {
            H[k * win2 + l] /= div;
            H[l * win2 + k] = H[k * win2 + l];
        }
    }
}

static int wrap(int i, int win) { return i >= (win >> 1) + 1 ? win - 1 - i : i; }

/* linsolve_wiener: Gaussian elimination in integers; x scaled by SCALE */
static int linsolve(int n, int64_t *A, int stride, int64_t *b, int32_t *x) {
    for (int k = 0; k < n - 1; k++) {
        for (int i = n - 1; i > k; i--) /* bubble the larger pivot upwards */
            if (llabs(A[(i - 1) * stride + k]) < llabs(A[i * stride + k])) {
                for (int j = 0; j < n; j++) {
                    const int64_t t = A[i * stride + j];
                    A[i * stride + j] = A[(i - 1) * stride + j], A[(i - 1) * stride + j] = t;
                }
                const int64_t t = b[i];
                b[i] = b[i - 1], b[i - 1] = t;
            }
        for (int i = k; i < n - 1; i++) {
            if (A[k * stride + k] == 0)
                return 0;
            const int64_t c = A[(i + 1) * stride + k], cd = A[k * stride + k];
            for (int j = 0; j < n; j++) A[(i + 1) * stride + j] -= c / 256 * A[k * stride + j] / cd * 256;
            b[i + 1] -= c * b[k] / cd;
        }
    }
    for (int i = n - 1; i >= 0; i--) {
        if (A[i * stride + i] == 0)
            return 0;
        int64_t c = 0;
        for (int j = i + 1; j <= n - 1; j++) c += A[i * stride + j] * x[j] / SCALE;
        x[i] = (int32_t)(SCALE * (b[i] - c) / A[i * stride + i]);
    }
    return 1;
}

/* One half-step of the alternating solve.  side 0: b fixed, a solved (update_a_sep_sym); side 1: a fixed, b solved. */
static void sep_update(int win, const int64_t *M, const int64_t *H, const int32_t *fixed, int side, int32_t *out) {
    const int win2 = win * win, h1 = (win >> 1) + 1;
    int64_t   A[HALF + 1], B[(HALF + 1) * (HALF + 1)];
    int32_t   S[WIN];
    memset(A, 0, sizeof(A)), memset(B, 0, sizeof(B));
    for (int i = 0; i < win; i++)
        for (int j = 0; j < win; j++) {
            if (side == 0)
                A[wrap(j, win)] += M[i * win + j] * fixed[i] / SCALE;
            else
                A[wrap(i, win)] += M[i * win + j] * fixed[j] / SCALE;
        }
    for (int i = 0; i < win; i++)
        for (int j = 0; j < win; j++)
            for (int k = 0; k < win; k++)
                for (int l = 0; l < win; l++) {
                    if (side == 0)
                        B[wrap(l, win) * h1 + wrap(k, win)] += H[(j * win + k) * win2 + i * win + l] * fixed[i] / SCALE * fixed[j] / SCALE;
                    else
                        B[wrap(j, win) * h1 + wrap(i, win)] += H[(i * win + k) * win2 + j * win + l] * fixed[k] / SCALE * fixed[l] / SCALE;
                }
    /* the taps sum to one: eliminate the centre tap */
    const int     c  = h1 - 1;
    const int64_t ac = A[c];
    for (int i = 0; i < c; i++) A[i] -= ac * 2 + B[i * h1 + c] - 2 * B[c * h1 + c];
    for (int i = 0; i < c; i++)
        for (int j = 0; j < c; j++) B[i * h1 + j] -= 2 * (B[i * h1 + c] + B[c * h1 + j] - 2 * B[c * h1 + c]);
    if (linsolve(c, B, h1, A, S)) {
        S[c] = (int32_t)SCALE;
        for (int i = h1; i < win; i++) S[i] = S[win - 1 - i], S[c] -= 2 * S[i];
        memcpy(out, S, win * sizeof(*out));
    }
}

static void finalize_taps(int win, const int32_t *f, int16_t *fi) {
    const int hw = win >> 1;
    for (int i = 0; i < hw; i++) {
        const int64_t num = (int64_t)f[i] * STEP;
        fi[i] = (int16_t)(num < 0 ? (num - SCALE / 2) / SCALE : (num + SCALE / 2) / SCALE);
    }
    if (win == WIN) {
        for (int i = 0; i < 3; i++) fi[i] = (int16_t)IMIN(IMAX(fi[i], tap_min[i]), tap_max[i]);
    } else { /* "Specialize for 7-tap filter": any shorter window is stored as 5 taps, from fi[0] and fi[1] */
        fi[2] = (int16_t)IMIN(IMAX(fi[1], tap_min[2]), tap_max[2]);
        fi[1] = (int16_t)IMIN(IMAX(fi[0], tap_min[1]), tap_max[1]);
        fi[0] = 0;
    }
    fi[6] = fi[0], fi[5] = fi[1], fi[4] = fi[2];
    fi[3] = (int16_t)(-2 * (fi[0] + fi[1] + fi[2]));
    fi[7] = 0;
}

static int64_t score(int win, const int64_t *M, const int64_t *H, const int16_t *vf, const int16_t *hf) {
    const int win2 = win * win, off = (WIN - win) >> 1;
    int16_t   a[WIN], b[WIN];
    int32_t   ab[WIN2];
    a[HALF] = b[HALF] = STEP;
    for (int i = 0; i < HALF; i++) {
        a[i] = a[WIN - 1 - i] = vf[i], b[i] = b[WIN - 1 - i] = hf[i];
        a[HALF] -= 2 * a[i], b[HALF] -= 2 * b[i];
    }
    for (int k = 0; k < win; k++)
        for (int l = 0; l < win; l++) ab[k * win + l] = a[l + off] * b[k + off];
    int64_t P = 0, Q = 0;
    for (int k = 0; k < win2; k++) {
        P += ab[k] * M[k] / STEP / STEP;
        for (int l = 0; l < win2; l++) Q += ab[k] * H[k * win2 + l] * ab[l] / STEP / STEP / STEP / STEP;
    }
    const int64_t ip = M[win2 >> 1], iq = H[(win2 >> 1) * win2 + (win2 >> 1)];
    return (Q - 2 * P) - (iq - 2 * ip);
}

/* the host half of search_wiener_seg: 1 = taps usable, 0 = rejected */
ORC_API int orc_lr_wiener_from_stats(int win, const int64_t *M, const int64_t *H, int16_t vf[8], int16_t hf[8]) {
    const int off = (WIN - win) >> 1;
    int32_t   a[WIN], b[WIN];
    for (int i = 0; i < win; i++) a[i] = b[i] = (int32_t)(SCALE / STEP * (i + off <= HALF ? tap_mid[i + off] : tap_mid[WIN - 1 - i - off]));
    for (int it = 1; it < ITERS; it++) {
        sep_update(win, M, H, b, 0, a);
        sep_update(win, M, H, a, 1, b);
    }
    memset(vf, 0, 16), memset(hf, 0, 16);
    finalize_taps(win, a, vf), finalize_taps(win, b, hf);
    return score(win, M, H, vf, hf) > 0 ? 0 : 1;
}

/* ---- bit counting (EbEntropyCoding.c) ---- */
static int recenter(int r, int v) { return v > (r << 1) ? v : (v >= r ? (v - r) << 1 : ((r - v) << 1) - 1); }
static int count_refsubexpfin(int n, int k, int ref, int v) {
    v = (ref << 1) <= n ? recenter(ref, v) : recenter(n - 1 - ref, n - 1 - v);
    v &= 0xffff;
    int count = 0, i = 0, mk = 0;
    for (;;) {
        const int b = i ? k + i - 1 : k, a = 1 << b;
        if (n <= mk + 3 * a) { /* quasi-uniform tail */
            const int nn = n - mk, vv = v - mk;
            if (nn > 1) {
                int l = 0;
                while ((1 << l) < nn) l++; /* get_msb(nn - 1) + 1 */
                count += vv < (1 << l) - nn ? l - 1 : l;
            }
            break;
        }
        count++;
        if (v >= mk + a)
            i++, mk += a;
        else {
            count += b;
            break;
        }
    }
    return count;
}
static int wiener_bits(int win, const int16_t *vf, const int16_t *hf, const int16_t *rvf, const int16_t *rhf) {
    int bits = 0;
    for (int d = 0; d < 2; d++) {
        const int16_t *f = d ? hf : vf, *r = d ? rhf : rvf;
        for (int t = win == WIN ? 0 : 1; t < 3; t++)
            bits += count_refsubexpfin(tap_max[t] - tap_min[t] + 1, tap_k[t], r[t] - tap_min[t], f[t] - tap_min[t]);
    }
    return bits;
}

static double rdcost(int rdmult, int64_t rate, int64_t dist) { return ((double)rate * rdmult) / 512.0 + (double)dist * 128.0; }

typedef struct {
    const void *cur, *src;
    int         cst, sst, sb, bd, ss, pw, ph;
} plane_ctx;

static int64_t try_wiener(const plane_ctx *c, const int lim[4], const int16_t *vf, const int16_t *hf) {
    svtf_lr_unit u;
    memset(&u, 0, sizeof(u));
    u.restoration_type = SVTF_RESTORE_WIENER;
    memcpy(u.wiener_vfilter, vf, 16), memcpy(u.wiener_hfilter, hf, 16);
    return orc_lr_try_unit(c->cur, c->cst, c->src, c->sst, c->sb, c->bd, c->ss, c->pw, c->ph, lim[0], lim[1], lim[2], lim[3], &u);
}

/* finer_tile_search_wiener_seg: coordinate descent on the three outer taps of the horizontal, then of the vertical filter */
static int64_t refine(const plane_ctx *c, const int lim[4], int win, const svtf_lr_search_cfg *cfg, int16_t *vf, int16_t *hf) {
    int64_t err = try_wiener(c, lim, vf, hf);
    if (!cfg->wn_use_refinement)
        return err;
    const int off = (WIN - win) >> 1, one = cfg->wn_max_one_refinement_step;
    for (int s = 4; s >= (one ? 4 : 1); s >>= 1)
        for (int d = 0; d < 2; d++) {
            int16_t *f = d ? vf : hf;
            for (int p = off; p < HALF; p++) {
                int moved_down = 0;
                for (int sign = -1; sign <= 1; sign += 2) {
                    for (;;) {
                        if (sign < 0 ? f[p] - s < tap_min[p] : f[p] + s > tap_max[p])
                            break;
                        f[p] += (int16_t)(sign * s), f[WIN - 1 - p] += (int16_t)(sign * s), f[HALF] -= (int16_t)(2 * sign * s);
                        const int64_t e2 = try_wiener(c, lim, vf, hf);
                        if (e2 > err) {
                            f[p] -= (int16_t)(sign * s), f[WIN - 1 - p] -= (int16_t)(sign * s), f[HALF] += (int16_t)(2 * sign * s);
                            break;
                        }
                        err = e2;
                        if (sign < 0)
                            moved_down = 1;
                        if (!(s == 4 && !one)) /* only the largest step keeps moving in the same direction */
                            break;
                    }
                    if (moved_down)
                        break;
                }
                if (moved_down) /* the reference leaves the tap loop of this filter after a successful downward move */
                    break;
            }
        }
    return err;
}

ORC_API int orc_lr_search_frame(const svtf_geom *g, const void *y, int sy, const void *cb, int scb, const void *cr, int scr, const void *oy,
                                int osy, const void *ocb, int oscb, const void *ocr, int oscr, const svtf_lr_search_cfg *cfg,
                                int32_t out_type[3], svtf_lr_unit *const out_units[3], svtf_lr_unit_search *const out_search[3]) {
    const void *const cur[3] = {y, cb, cr}, *const src[3] = {oy, ocb, ocr};
    const int         cst[3] = {sy, scb, scr}, sst[3] = {osy, oscb, oscr};
    if (cfg->sg_enabled)
        return -1;
    const int planes = cfg->wn_enabled && cfg->wn_use_chroma ? 3 : 1;
    const int wn_luma = cfg->wn_filter_tap_lvl == 1 ? 7 : (cfg->wn_filter_tap_lvl == 2 ? 5 : 3);
    for (int p = 0; p < 3; p++) {
        out_type[p] = SVTF_RESTORE_NONE;
        int hu, vu;
        orc_lr_units(g, p, cfg->restoration_unit_size[p], &hu, &vu);
        memset(out_units[p], 0, sizeof(svtf_lr_unit) * hu * vu);
    }
    for (int p = 0; p < planes; p++) {
        const int ss = p ? 1 : 0, U = cfg->restoration_unit_size[p], off = 8 >> ss, win = p ? IMIN(wn_luma, 5) : wn_luma;
        plane_ctx c = {cur[p], src[p], cst[p], sst[p], g->sample_bytes, g->bit_depth, ss, (g->width + ss) >> ss, (g->height + ss) >> ss};
        int       hu, vu;
        orc_lr_units(g, p, U, &hu, &vu);
        svtf_lr_unit_search *us = calloc((size_t)hu * vu, sizeof(*us));
        int64_t             *M = malloc(sizeof(int64_t) * WIN2), *H = malloc(sizeof(int64_t) * WIN2 * WIN2);
        if (!us || !M || !H)
            return -1;
        /* ---- per unit: SSE unfiltered, Wiener taps + refined SSE ---- */
        for (int u = 0; u < hu * vu; u++) {
            const int ur = u / hu, uc = u % hu;
            const int lim[4] = {uc * U, uc == hu - 1 ? c.pw : (uc + 1) * U, IMAX(0, ur * U - off), ur == vu - 1 ? c.ph : (ur + 1) * U - off};
            svtf_lr_unit none;
            memset(&none, 0, sizeof(none));
            us[u].sse[0] = orc_lr_try_unit(c.cur, c.cst, c.src, c.sst, c.sb, c.bd, ss, c.pw, c.ph, lim[0], lim[1], lim[2], lim[3], &none);
            us[u].sse[1] = INT64_MAX;
            if (!cfg->wn_enabled)
                continue;
            orc_lr_wiener_stats(c.cur, c.cst, c.src, c.sst, c.sb, c.bd, c.pw, c.ph, lim[0], lim[1], lim[2], lim[3], win, M, H);
            int16_t vf[8], hf[8];
            if (!orc_lr_wiener_from_stats(win, M, H, vf, hf))
                continue;
            us[u].sse[1] = refine(&c, lim, win, cfg, vf, hf);
            memcpy(us[u].wiener_vfilter, vf, 16), memcpy(us[u].wiener_hfilter, hf, 16);
        }
        /* ---- frame decision: RESTORE_NONE against RESTORE_WIENER with per-unit on/off (search_wiener_finish) ---- */
        int64_t sse_none = 0, sse_w = 0, bits_w = 0;
        int16_t rvf[8] = {3, -7, 15, -22, 15, -7, 3, 0}, rhf[8] = {3, -7, 15, -22, 15, -7, 3, 0}; /* set_default_wiener */
        for (int u = 0; u < hu * vu; u++) sse_none += us[u].sse[0];
        if (cfg->wn_enabled) {
            for (int u = 0; u < hu * vu; u++) {
                const int64_t bits_none = cfg->wiener_restore_cost[0];
                us[u].best_rtype        = SVTF_RESTORE_NONE;
                if (us[u].sse[1] == INT64_MAX) {
                    bits_w += bits_none, sse_w += us[u].sse[0];
                    continue;
                }
                const int64_t bits_wn = cfg->wiener_restore_cost[1] +
                    ((int64_t)wiener_bits(win, us[u].wiener_vfilter, us[u].wiener_hfilter, rvf, rhf) << 9);
                const double cn = rdcost(cfg->rdmult, bits_none >> 4, us[u].sse[0]), cw = rdcost(cfg->rdmult, bits_wn >> 4, us[u].sse[1]);
                if (cw < cn) {
                    us[u].best_rtype = SVTF_RESTORE_WIENER;
                    sse_w += us[u].sse[1], bits_w += bits_wn;
                    memcpy(rvf, us[u].wiener_vfilter, 16), memcpy(rhf, us[u].wiener_hfilter, 16);
                } else
                    sse_w += us[u].sse[0], bits_w += bits_none;
            }
            const double cost_none = rdcost(cfg->rdmult, 0, sse_none), cost_w = rdcost(cfg->rdmult, bits_w >> 4, sse_w);
            if (cost_w < cost_none) {
                out_type[p] = SVTF_RESTORE_WIENER;
                for (int u = 0; u < hu * vu; u++) { /* copy_unit_info :1202 */
                    out_units[p][u].restoration_type = us[u].best_rtype;
                    if (us[u].best_rtype == SVTF_RESTORE_WIENER) {
                        memcpy(out_units[p][u].wiener_vfilter, us[u].wiener_vfilter, 16);
                        memcpy(out_units[p][u].wiener_hfilter, us[u].wiener_hfilter, 16);
                    }
                }
            }
        }
        if (out_search && out_search[p])
            memcpy(out_search[p], us, sizeof(*us) * hu * vu);
        free(us), free(M), free(H);
    }
    return 0;
}
