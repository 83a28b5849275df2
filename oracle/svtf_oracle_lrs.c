/*
 * svtf_oracle_lrs.c — CPU restatement of the reference's loop-restoration filter SEARCH (Wiener and self-guided tools, switchable).
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
 *   search_sgrproj_seg :1220 / search_selfguided_restoration :550 (apply_sgr :522 = svtf_oracle_lr.c's orc_lr_sgr_unit_flt,
 *   svt_get_proj_subspace_c :417, encode_xq :508, finer_search_pixel_proj_error :320, get_pixel_proj_error :305),
 *   search_sgrproj_finish :1267, count_sgrproj_bits :655, search_switchable :1148, use_prev_frame_coeffs :1313-1319.
 * Units: foreach_rest_unit_in_tile (Common/Codec/EbRestoration.c:1250) — raster order, the last unit of a row / column
 * takes the remainder (up to 1.5 units), rows shifted up by 8 luma samples.
 * Pinned against the reference's own code through oracle/ref_shim_lrs.c (tests/test_lrs_oracle.py).
 */
#include <math.h>
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

/* ---- self-guided tools ---- */
void orc_lr_sgr_unit_flt(const void *cur, int cst, int sb, int bd, int ss, int pw, int ph, int h0, int h1, int v0, int v1, int ep, int32_t *flt0,
                         int32_t *flt1);
/* svt_aom_eb_sgr_params radii (Common/Codec/EbRestoration.c:85-103) */
static const int sg_r[16][2] = {{2, 1}, {2, 1}, {2, 1}, {2, 1}, {2, 1}, {2, 1}, {2, 1}, {2, 1}, {2, 1}, {2, 1}, {0, 1}, {0, 1}, {0, 1}, {0, 1}, {2, 0}, {2, 0}};
static const int sg_min[2] = {-96, -32}, sg_max[2] = {31, 95}; /* SGRPROJ_PRJ_MIN0/1, MAX0/1 */

static int iclamp(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

/* get_pixel_proj_error :305 (svt_decode_xq + svt_av1_[lowbd|highbd]_pixel_proj_error_c :167/:233) */
static int64_t proj_error(const plane_ctx *c, const int lim[4], const int32_t *f0, const int32_t *f1, const int *xqd, int ep) {
    const int r0 = sg_r[ep][0], r1 = sg_r[ep][1], w = lim[1] - lim[0];
    int       xq[2];
    if (r0 == 0)
        xq[0] = 0, xq[1] = 128 - xqd[1];
    else if (r1 == 0)
        xq[0] = xqd[0], xq[1] = 0;
    else
        xq[0] = xqd[0], xq[1] = 128 - xq[0] - xqd[1];
    int64_t err = 0;
    for (int y = lim[2]; y < lim[3]; y++)
        for (int x = lim[0]; x < lim[1]; x++) {
            const size_t i = (size_t)(y - lim[2]) * w + x - lim[0];
            const int    u = px(c->cur, c->cst, c->sb, x, y) << 4;
            int          v = u << 7;
            if (r0 > 0)
                v += xq[0] * (f0[i] - u);
            if (r1 > 0)
                v += xq[1] * (f1[i] - u);
            const int e = ((v + (1 << 10)) >> 11) - px(c->src, c->sst, c->sb, x, y);
            err += (int64_t)e * e;
        }
    return err;
}

/* search_selfguided_restoration :550 for one unit: best (ep, xqd) */
static void search_sgr_unit(const plane_ctx *c, const int lim[4], int start_ep, int end_ep, int ep_inc, int do_refine, int *best_ep, int best_xqd[2]) {
    const int    w = lim[1] - lim[0], h = lim[3] - lim[2], size = w * h;
    int32_t     *f0 = malloc(sizeof(int32_t) * size), *f1 = malloc(sizeof(int32_t) * size);
    int64_t      best_err = -1;
    *best_ep = 0, best_xqd[0] = best_xqd[1] = 0;
    for (int ep = start_ep; ep < end_ep; ep += ep_inc) {
        const int r0 = sg_r[ep][0], r1 = sg_r[ep][1];
        orc_lr_sgr_unit_flt(c->cur, c->cst, c->sb, c->bd, c->ss, c->pw, c->ph, lim[0], lim[1], lim[2], lim[3], ep, f0, f1);
        /* svt_get_proj_subspace_c :417: least squares in double */
        double H[2][2] = {{0, 0}, {0, 0}}, C[2] = {0, 0};
        for (int y = lim[2]; y < lim[3]; y++)
            for (int x = lim[0]; x < lim[1]; x++) {
                const size_t i  = (size_t)(y - lim[2]) * w + x - lim[0];
                const double u  = (double)(px(c->cur, c->cst, c->sb, x, y) << 4);
                const double sv = (double)(px(c->src, c->sst, c->sb, x, y) << 4) - u;
                const double a  = r0 > 0 ? (double)f0[i] - u : 0, b = r1 > 0 ? (double)f1[i] - u : 0;
                H[0][0] += a * a, H[1][1] += b * b, H[0][1] += a * b, C[0] += a * sv, C[1] += b * sv;
            }
        H[0][0] /= size, H[0][1] /= size, H[1][1] /= size, H[1][0] = H[0][1], C[0] /= size, C[1] /= size;
        int xq[2] = {0, 0};
        if (r0 == 0) {
            if (!(H[1][1] < 1e-8))
                xq[1] = (int)rint(C[1] / H[1][1] * 128);
        } else if (r1 == 0) {
            if (!(H[0][0] < 1e-8))
                xq[0] = (int)rint(C[0] / H[0][0] * 128);
        } else {
            const double det = H[0][0] * H[1][1] - H[0][1] * H[1][0];
            if (!(det < 1e-8)) {
                const double x0 = (H[1][1] * C[0] - H[0][1] * C[1]) / det, x1 = (H[0][0] * C[1] - H[1][0] * C[0]) / det;
                xq[0] = (int)rint(x0 * 128), xq[1] = (int)rint(x1 * 128);
            }
        }
        int xqd[2]; /* encode_xq :508 */
        if (r0 == 0)
            xqd[0] = 0, xqd[1] = iclamp(128 - xq[1], sg_min[1], sg_max[1]);
        else if (r1 == 0)
            xqd[0] = iclamp(xq[0], sg_min[0], sg_max[0]), xqd[1] = iclamp(128 - xqd[0], sg_min[1], sg_max[1]);
        else
            xqd[0] = iclamp(xq[0], sg_min[0], sg_max[0]), xqd[1] = iclamp(128 - xqd[0] - xq[1], sg_min[1], sg_max[1]);
        /* finer_search_pixel_proj_error :320, start_step 2 */
        int64_t err = proj_error(c, lim, f0, f1, xqd, ep);
        if (do_refine)
            for (int st = 2; st >= 1; st >>= 1)
                for (int p = 0; p < 2; p++) {
                    if ((r0 == 0 && p == 0) || (r1 == 0 && p == 1))
                        continue;
                    int moved_down = 0;
                    for (int sign = -1; sign <= 1; sign += 2) {
                        for (;;) {
                            if (sign < 0 ? xqd[p] - st < sg_min[p] : xqd[p] + st > sg_max[p])
                                break;
                            xqd[p] += sign * st;
                            const int64_t e2 = proj_error(c, lim, f0, f1, xqd, ep);
                            if (e2 > err) {
                                xqd[p] -= sign * st;
                                break;
                            }
                            err = e2;
                            if (sign < 0)
                                moved_down = 1;
                            if (st != 2)
                                break;
                        }
                        if (moved_down)
                            break;
                    }
                    if (moved_down)
                        break;
                }
        if (best_err == -1 || err < best_err)
            *best_ep = ep, best_err = err, best_xqd[0] = xqd[0], best_xqd[1] = xqd[1];
    }
    free(f0), free(f1);
}

static int sgr_bits(const int *ep_xqd /* ep, xqd0, xqd1 */, const int *ref_xqd) { /* count_sgrproj_bits :655 */
    int bits = 4;
    if (sg_r[ep_xqd[0]][0] > 0)
        bits += count_refsubexpfin(sg_max[0] - sg_min[0] + 1, 4, ref_xqd[0] - sg_min[0], ep_xqd[1] - sg_min[0]);
    if (sg_r[ep_xqd[0]][1] > 0)
        bits += count_refsubexpfin(sg_max[1] - sg_min[1] + 1, 4, ref_xqd[1] - sg_min[1], ep_xqd[2] - sg_min[1]);
    return bits;
}

/* what rest_finish_search keeps per unit across the planes (its `rusi` array is allocated and zeroed ONCE, before the
 * plane loop :1570-1577, so entries a plane does not refresh still hold the previous plane's values) */
typedef struct {
    int64_t sse[3];
    int16_t vf[8], hf[8];
    int     sgr[3]; /* ep, xqd0, xqd1 */
    int     best[3]; /* best_rtype[WIENER - 1], [SGRPROJ - 1], [SWITCHABLE - 1] */
} finish_unit;

ORC_API int orc_lr_search_frame(const svtf_geom *g, const void *y, int sy, const void *cb, int scb, const void *cr, int scr, const void *oy,
                                int osy, const void *ocb, int oscb, const void *ocr, int oscr, const svtf_lr_search_cfg *cfg,
                                int32_t out_type[3], svtf_lr_unit *const out_units[3], svtf_lr_unit_search *const out_search[3],
                                int32_t sg_frame_ep_cnt[16]) {
    const void *const cur[3] = {y, cb, cr}, *const src[3] = {oy, ocb, ocr};
    const int         cst[3] = {sy, scb, scr}, sst[3] = {osy, oscb, oscr};
    const int wn = cfg->wn_enabled != 0, sg = cfg->sg_enabled != 0;
    const int planes  = (wn && cfg->wn_use_chroma) || (sg && cfg->sg_use_chroma) ? 3 : 1;
    const int wn_luma = cfg->wn_filter_tap_lvl == 1 ? 7 : (cfg->wn_filter_tap_lvl == 2 ? 5 : 3);
    int       n0;
    {
        int hu, vu;
        n0 = 0; /* the reference sizes the array for luma and asserts ntiles[1] <= ntiles[0] (:1567) */
        for (int p = 0; p < 3; p++) {
            orc_lr_units(g, p, cfg->restoration_unit_size[p], &hu, &vu);
            n0 = IMAX(n0, hu * vu);
        }
    }
    finish_unit *fu = calloc(n0, sizeof(*fu));
    if (sg_frame_ep_cnt)
        memset(sg_frame_ep_cnt, 0, sizeof(int32_t) * 16);
    for (int p = 0; p < 3; p++) {
        out_type[p] = SVTF_RESTORE_NONE;
        int hu, vu;
        orc_lr_units(g, p, cfg->restoration_unit_size[p], &hu, &vu);
        memset(out_units[p], 0, sizeof(svtf_lr_unit) * hu * vu);
    }
    for (int p = 0; p < planes; p++) {
        const int ss = p ? 1 : 0, U = cfg->restoration_unit_size[p], off = 8 >> ss, win = p ? IMIN(wn_luma, 5) : wn_luma;
        const int do_wn = wn && (!p || cfg->wn_use_chroma), do_sg = sg && (!p || cfg->sg_use_chroma);
        plane_ctx c = {cur[p], src[p], cst[p], sst[p], g->sample_bytes, g->bit_depth, ss, (g->width + ss) >> ss, (g->height + ss) >> ss};
        int       hu, vu;
        orc_lr_units(g, p, U, &hu, &vu);
        const int            n  = hu * vu;
        svtf_lr_unit_search *us = calloc((size_t)n, sizeof(*us));
        int64_t             *M = malloc(sizeof(int64_t) * WIN2), *H = malloc(sizeof(int64_t) * WIN2 * WIN2);
        if (!us || !M || !H)
            return -1;
        /* ---- restoration_seg_search: per unit ---- */
        for (int u = 0; u < n; u++) {
            const int ur = u / hu, uc = u % hu;
            const int lim[4] = {uc * U, uc == hu - 1 ? c.pw : (uc + 1) * U, IMAX(0, ur * U - off), ur == vu - 1 ? c.ph : (ur + 1) * U - off};
            svtf_lr_unit none;
            memset(&none, 0, sizeof(none));
            us[u].sse[0] = orc_lr_try_unit(c.cur, c.cst, c.src, c.sst, c.sb, c.bd, ss, c.pw, c.ph, lim[0], lim[1], lim[2], lim[3], &none);
            us[u].sse[1] = INT64_MAX;
            if (do_wn) {
                int16_t vf[8], hf[8];
                int     ok;
                if (cfg->prev_units[p] && cfg->prev_units[p][u].restoration_type == SVTF_RESTORE_WIENER) { /* :1313-1319 */
                    memcpy(vf, cfg->prev_units[p][u].wiener_vfilter, 16), memcpy(hf, cfg->prev_units[p][u].wiener_hfilter, 16);
                    ok = 1;
                } else {
                    orc_lr_wiener_stats(c.cur, c.cst, c.src, c.sst, c.sb, c.bd, c.pw, c.ph, lim[0], lim[1], lim[2], lim[3], win, M, H);
                    ok = orc_lr_wiener_from_stats(win, M, H, vf, hf);
                }
                if (ok) {
                    us[u].sse[1] = refine(&c, lim, win, cfg, vf, hf);
                    memcpy(us[u].wiener_vfilter, vf, 16), memcpy(us[u].wiener_hfilter, hf, 16);
                }
            }
            if (do_sg) {
                int ep, xqd[2];
                search_sgr_unit(&c, lim, cfg->sg_start_ep[p > 0], cfg->sg_end_ep[p > 0], cfg->sg_ep_inc[p > 0], cfg->sg_refine[p > 0], &ep, xqd);
                us[u].sgr_ep = ep, us[u].sgr_xqd[0] = xqd[0], us[u].sgr_xqd[1] = xqd[1];
                if (sg_frame_ep_cnt)
                    sg_frame_ep_cnt[ep]++;
                svtf_lr_unit t;
                memset(&t, 0, sizeof(t));
                t.restoration_type = SVTF_RESTORE_SGRPROJ, t.sgr_ep = ep, t.sgr_xqd[0] = xqd[0], t.sgr_xqd[1] = xqd[1];
                us[u].sse[2] = orc_lr_try_unit(c.cur, c.cst, c.src, c.sst, c.sb, c.bd, ss, c.pw, c.ph, lim[0], lim[1], lim[2], lim[3], &t);
            }
        }
        /* ---- rest_finish_search: cost of every frame type that applies, per-unit decisions in raster order ---- */
        const int force = wn ? (sg ? 4 : 1) : (sg ? 2 : 0); /* RESTORE_TYPES = 4: all */
        double    best_cost = 0;
        int       best_type = 0;
        for (int r = 0; r < (n > 1 ? 4 : 3); r++) {
            if (force != 4 && r != 0 && r != force)
                continue;
            if (p && ((r == 1 && !cfg->wn_use_chroma) || (r == 2 && !cfg->sg_use_chroma)))
                continue;
            int64_t sse = 0, bits = 0;
            int16_t rvf[8] = {3, -7, 15, -22, 15, -7, 3, 0}, rhf[8] = {3, -7, 15, -22, 15, -7, 3, 0}; /* set_default_wiener */
            int     rsg[2] = {-32, 31};                                                              /* set_default_sgrproj */
            for (int u = 0; u < n; u++) {
                finish_unit *f = &fu[u];
                if (r == 0) {
                    f->sse[0] = us[u].sse[0];
                    sse += f->sse[0];
                } else if (r == 1) { /* search_wiener_finish */
                    const int64_t bits_none = cfg->wiener_restore_cost[0];
                    f->sse[1] = us[u].sse[1];
                    if (f->sse[1] == INT64_MAX) {
                        bits += bits_none, sse += f->sse[0], f->best[0] = 0;
                        continue;
                    }
                    memcpy(f->vf, us[u].wiener_vfilter, 16), memcpy(f->hf, us[u].wiener_hfilter, 16);
                    const int64_t bw = cfg->wiener_restore_cost[1] + ((int64_t)wiener_bits(win, f->vf, f->hf, rvf, rhf) << 9);
                    const double  cn = rdcost(cfg->rdmult, bits_none >> 4, f->sse[0]), cw = rdcost(cfg->rdmult, bw >> 4, f->sse[1]);
                    f->best[0] = cw < cn ? 1 : 0;
                    sse += f->sse[f->best[0]], bits += cw < cn ? bw : bits_none;
                    if (cw < cn)
                        memcpy(rvf, f->vf, 16), memcpy(rhf, f->hf, 16);
                } else if (r == 2) { /* search_sgrproj_finish */
                    f->sse[2] = us[u].sse[2];
                    f->sgr[0] = us[u].sgr_ep, f->sgr[1] = us[u].sgr_xqd[0], f->sgr[2] = us[u].sgr_xqd[1];
                    const int64_t bits_none = cfg->sgrproj_restore_cost[0];
                    const int64_t bs = cfg->sgrproj_restore_cost[1] + ((int64_t)sgr_bits(f->sgr, rsg) << 9);
                    const double  cn = rdcost(cfg->rdmult, bits_none >> 4, f->sse[0]), cs = rdcost(cfg->rdmult, bs >> 4, f->sse[2]);
                    f->best[1] = cs < cn ? 2 : 0;
                    sse += f->sse[f->best[1]], bits += cs < cn ? bs : bits_none;
                    if (cs < cn)
                        rsg[0] = f->sgr[1], rsg[1] = f->sgr[2];
                } else { /* search_switchable */
                    double  bc = 0;
                    int64_t bb = 0;
                    int     bt = 0;
                    for (int t = 0; t < 3; t++) {
                        if (t > 0 && f->best[t - 1] == 0)
                            continue;
                        const int64_t coeff = t == 1 ? wiener_bits(p ? 5 : 7, f->vf, f->hf, rvf, rhf) : (t == 2 ? sgr_bits(f->sgr, rsg) : 0);
                        const int64_t b     = cfg->switchable_restore_cost[t] + (coeff << 9);
                        const double  cost  = rdcost(cfg->rdmult, b >> 4, f->sse[t]);
                        if (t == 0 || cost < bc)
                            bc = cost, bb = b, bt = t;
                    }
                    f->best[2] = bt;
                    sse += f->sse[bt], bits += bb;
                    if (bt == 1)
                        memcpy(rvf, f->vf, 16), memcpy(rhf, f->hf, 16);
                    if (bt == 2)
                        rsg[0] = f->sgr[1], rsg[1] = f->sgr[2];
                }
            }
            const double cost = rdcost(cfg->rdmult, bits >> 4, sse);
            if (r == 0 || cost < best_cost)
                best_cost = cost, best_type = r;
        }
        out_type[p] = best_type;
        if (best_type != 0)
            for (int u = 0; u < n; u++) { /* copy_unit_info :1202 */
                svtf_lr_unit *o = &out_units[p][u];
                o->restoration_type = fu[u].best[best_type - 1];
                if (o->restoration_type == SVTF_RESTORE_WIENER)
                    memcpy(o->wiener_vfilter, fu[u].vf, 16), memcpy(o->wiener_hfilter, fu[u].hf, 16);
                else if (o->restoration_type == SVTF_RESTORE_SGRPROJ)
                    o->sgr_ep = fu[u].sgr[0], o->sgr_xqd[0] = fu[u].sgr[1], o->sgr_xqd[1] = fu[u].sgr[2];
            }
        if (out_search && out_search[p])
            memcpy(out_search[p], us, sizeof(*us) * n);
        free(us), free(M), free(H);
    }
    free(fu);
    return 0;
}
