/*
 * svtf_oracle_lr.c — CPU restatement of the reference's loop restoration for one frame.  TEST INFRASTRUCTURE ONLY: only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may use it; the product never does.
 *
 * Follows svt_av1_loop_restoration_filter_frame (Common/Codec/EbRestoration.c:1179) with the stripe boundaries of
 * svt_av1_loop_restoration_save_boundary_lines (:1682), formulated per sample instead of per buffer:
 *   * the plane is cut into restoration units (foreach_rest_unit_in_tile :1250: unit_size squares, shifted up by 8 luma
 *     rows, the last one up to 1.5x) and into 64-luma-row stripes shifted up by 8 (svt_av1_loop_restoration_filter_unit
 *     :1067); every output sample is a function of a 7x7 window of "stripe input";
 *   * stripe input = the picture inside the stripe; the 3 rows above / below it come from the DEBLOCKED picture, two rows
 *     saved per boundary and the outer one repeated (svt_aom_setup_processing_stripe_boundary :288, save_deblock_boundary_
 *     lines :1522), except at the top / bottom of the frame where the picture's own edge row is repeated (svt_extend_frame
 *     :197); columns outside the plane repeat the edge column (extend_frame, svt_aom_extend_lines :1507); with
 *     optimized_lr the rows come from the picture itself and only the outermost is replaced by its neighbour (:350-377);
 *   * Wiener: svt_av1_[highbd_]wiener_convolve_add_src_c (Common/Codec/convolve.c:100-233), rounding of
 *     get_conv_params_wiener (EbRestoration.c:49);
 *   * self-guided: svt_apply_selfguided_restoration_c (:957), selfguided_restoration_fast_internal (:669, radius 2, every
 *     other row of the stripe) and selfguided_restoration_internal (:803, radius 1), svt_decode_xq (:634).
 * Pinned against the reference's own code through oracle/ref_shim_lr.c (tests/test_lr_oracle.py).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../include/svtfilt_b200.h"

#define ORC_API __attribute__((visibility("default")))
#define IMIN(a, b) ((a) < (b) ? (a) : (b))
#define IMAX(a, b) ((a) > (b) ? (a) : (b))
#define RND(v, n) (((v) + ((1 << (n)) >> 1)) >> (n)) /* ROUND_POWER_OF_TWO */

/* svt_aom_eb_sgr_params (EbRestoration.c:85-103): {r0, r1}, {s0, s1} */
static const int sgr_r[16][2] = {{2, 1}, {2, 1}, {2, 1}, {2, 1}, {2, 1}, {2, 1}, {2, 1}, {2, 1},
                                 {2, 1}, {2, 1}, {0, 1}, {0, 1}, {0, 1}, {0, 1}, {2, 0}, {2, 0}};
static const int sgr_s[16][2] = {{140, 3236}, {112, 2158}, {93, 1618}, {80, 1438}, {70, 1295}, {58, 1177}, {47, 1079}, {37, 996},
                                 {30, 925},   {25, 863},   {-1, 2589}, {-1, 1618}, {-1, 1177}, {-1, 925},  {56, -1},   {22, -1}};
/* svt_aom_eb_one_by_x[n - 1] for n = 9 and n = 25 (:662-665) */
static const int one_by_9 = 455, one_by_25 = 164;

/* svt_aom_eb_x_by_xplus1 (:647-660): round(256 z / (z + 1)) with 0 -> 1 */
static int x_by_xplus1(int z) {
    if (z == 0)
        return 1;
    if (z >= 255)
        return 256;
    return (256 * z + (z + 1) / 2) / (z + 1);
}

typedef struct {
    const void *cur, *dbk;
    int         cur_st, dbk_st, sb;
    int         pw, ph;     /* plane size */
    int         y0, y1;     /* stripe rows [y0, y1) */
    int         above, below; /* the stripe has deblocked context above / below */
    int         opt;
} stripe_in;

static int px(const void *p, int st, int sb, int x, int y) {
    return sb == 1 ? ((const uint8_t *)p)[(size_t)y * st + x] : ((const uint16_t *)p)[(size_t)y * st + x];
}

/* stripe input at (x, r), r in [y0 - 3, y1 + 3) */
static int stripe_px(const stripe_in *s, int x, int r) {
    const int xc = IMIN(IMAX(x, 0), s->pw - 1);
    if (s->opt == 2) /* the encoder's search: no stripe boundaries (need_bounadaries = 0, EbRestoration.c:1112) */
        return px(s->cur, s->cur_st, s->sb, xc, IMIN(IMAX(r, 0), s->ph - 1));
    if (r < s->y0) {
        if (!s->above)
            return px(s->cur, s->cur_st, s->sb, xc, IMAX(r, 0));
        if (s->opt)
            return px(s->cur, s->cur_st, s->sb, xc, r == s->y0 - 3 ? s->y0 - 2 : r);
        return px(s->dbk, s->dbk_st, s->sb, xc, IMAX(r, s->y0 - 2));
    }
    if (r >= s->y1) {
        if (!s->below)
            return px(s->cur, s->cur_st, s->sb, xc, IMIN(r, s->ph - 1));
        if (s->opt)
            return px(s->cur, s->cur_st, s->sb, xc, IMIN(r == s->y1 + 2 ? s->y1 + 1 : r, s->ph - 1));
        return px(s->dbk, s->dbk_st, s->sb, xc, IMIN(IMIN(r, s->y1 + 1), s->ph - 1));
    }
    return px(s->cur, s->cur_st, s->sb, xc, r);
}

/* Wiener: 7 taps (index 0..6, the 8th of the InterpKernel is 0) around the sample, "add_src" = +128 on the centre tap */
static int wiener_px(const stripe_in *s, const svtf_lr_unit *u, int bd, int x, int y) {
    const int r0 = bd == 12 ? 5 : 3, r1 = 14 - r0; /* get_conv_params_wiener: WIENER_ROUND0_BITS 3, +2 when bd + 6 > 16 */
    const int lim = (1 << (bd + 1 + 7 - r0)) - 1;  /* WIENER_CLAMP_LIMIT - 1 */
    int       col[7];
    for (int k = 0; k < 7; k++) {
        int sum = (stripe_px(s, x, y + k - 3) << 7) + (1 << (bd + 7 - 1));
        for (int t = 0; t < 7; t++) sum += stripe_px(s, x + t - 3, y + k - 3) * u->wiener_hfilter[t];
        col[k] = IMIN(IMAX(RND(sum, r0), 0), lim);
    }
    int sum = (col[3] << 7) - (1 << (bd + r1 - 1));
    for (int k = 0; k < 7; k++) sum += col[k] * u->wiener_vfilter[k];
    return IMIN(IMAX(RND(sum, r1), 0), (1 << bd) - 1);
}

/* A and B of the self-guided filter at (x, y): box sums of radius r over the stripe input */
static void sgr_ab(const stripe_in *s, int bd, int r, int sv, int x, int y, int *A, int *B) {
    int64_t sum = 0, sq = 0;
    for (int dy = -r; dy <= r; dy++)
        for (int dx = -r; dx <= r; dx++) {
            const int v = stripe_px(s, x + dx, y + dy);
            sum += v, sq += (int64_t)v * v;
        }
    const int      n = (2 * r + 1) * (2 * r + 1);
    const uint32_t a = (uint32_t)RND((int32_t)sq, 2 * (bd - 8)), b = (uint32_t)RND((int32_t)sum, bd - 8);
    const uint32_t p = (a * n < b * b) ? 0 : a * n - b * b;
    const uint32_t z = (p * (uint32_t)sv + (1u << 19)) >> 20; /* 32-bit arithmetic as in the C code (:739-742) */
    *A               = x_by_xplus1((int)IMIN(z, 255u));
    *B               = (int32_t)(((uint32_t)(256 - *A) * (uint32_t)sum * (uint32_t)(n == 25 ? one_by_25 : one_by_9) + (1u << 11)) >> 12);
}

/* flt0 / flt1 of svt_av1_selfguided_restoration_c (:905) at (x, y): the two guided filters' outputs, 4 fractional bits */
static void sgr_flt(const stripe_in *s, int ep, int bd, int x, int y, int *f0, int *f1) {
    const int r0 = sgr_r[ep][0], r1 = sgr_r[ep][1];
    const int dg = stripe_px(s, x, y);
    *f0 = *f1 = 0;
    if (r0 > 0) { /* fast: A/B exist on the rows y0 - 1, y0 + 1, ... of the stripe only */
        int a = 0, b = 0, A, B, sh;
        if (!((y - s->y0) & 1)) {
            for (int dy = -1; dy <= 1; dy += 2)
                for (int dx = -1; dx <= 1; dx++) {
                    sgr_ab(s, bd, r0, sgr_s[ep][0], x + dx, y + dy, &A, &B);
                    a += A * (dx ? 5 : 6), b += B * (dx ? 5 : 6);
                }
            sh = 8 + 5 - 4;
        } else {
            for (int dx = -1; dx <= 1; dx++) {
                sgr_ab(s, bd, r0, sgr_s[ep][0], x + dx, y, &A, &B);
                a += A * (dx ? 5 : 6), b += B * (dx ? 5 : 6);
            }
            sh = 8 + 4 - 4;
        }
        *f0 = RND(a * dg + b, sh);
    }
    if (r1 > 0) {
        int a = 0, b = 0, A, B;
        for (int dy = -1; dy <= 1; dy++)
            for (int dx = -1; dx <= 1; dx++) {
                sgr_ab(s, bd, r1, sgr_s[ep][1], x + dx, y + dy, &A, &B);
                a += A * ((dx && dy) ? 3 : 4), b += B * ((dx && dy) ? 3 : 4);
            }
        *f1 = RND(a * dg + b, 8 + 5 - 4);
    }
}

static int sgr_px(const stripe_in *s, const svtf_lr_unit *u, int bd, int x, int y) {
    const int ep = u->sgr_ep, r0 = sgr_r[ep][0], r1 = sgr_r[ep][1];
    const int dg = stripe_px(s, x, y), uu = dg << 4;
    int       xq[2], f0, f1; /* svt_decode_xq */
    if (r0 == 0)
        xq[0] = 0, xq[1] = 128 - u->sgr_xqd[1];
    else if (r1 == 0)
        xq[0] = u->sgr_xqd[0], xq[1] = 0;
    else
        xq[0] = u->sgr_xqd[0], xq[1] = 128 - xq[0] - u->sgr_xqd[1];
    sgr_flt(s, ep, bd, x, y, &f0, &f1);
    int v = uu << 7;
    if (r0 > 0)
        v += xq[0] * (f0 - uu);
    if (r1 > 0)
        v += xq[1] * (f1 - uu);
    const int w = (int16_t)RND(v, 7 + 4);
    return IMIN(IMAX(w, 0), (1 << bd) - 1);
}

/* apply_sgr (Encoder/Codec/EbRestorationPick.c:522) for the unit [h0, h1) x [v0, v1): flt0 / flt1 (row pitch h1 - h0) of
 * parameter set ep, no stripe boundaries.  The reference filters 64-row blocks counted from the unit's first row; the only
 * thing that depends on the block is the row parity of the radius-2 filter, and units, blocks and stripes all start on
 * even rows.  Used by svtf_oracle_lrs.c. */
void orc_lr_sgr_unit_flt(const void *cur, int cst, int sb, int bd, int ss, int pw, int ph, int h0, int h1, int v0, int v1, int ep, int32_t *flt0,
                         int32_t *flt1) {
    stripe_in s = {cur, cur, cst, cst, sb, pw, ph, v0, v1, 0, 0, 2};
    for (int yy = v0; yy < v1; yy++)
        for (int xx = h0; xx < h1; xx++) {
            int f0, f1;
            sgr_flt(&s, ep, bd, xx, yy, &f0, &f1);
            flt0[(size_t)(yy - v0) * (h1 - h0) + xx - h0] = f0, flt1[(size_t)(yy - v0) * (h1 - h0) + xx - h0] = f1;
        }
}

ORC_API int orc_lr_units(const svtf_geom *g, int plane, int unit_size, int *horz, int *vert) {
    const int ss = plane ? 1 : 0, pw = (g->width + ss) >> ss, ph = (g->height + ss) >> ss;
    *horz = IMAX((pw + (unit_size >> 1)) / unit_size, 1), *vert = IMAX((ph + (unit_size >> 1)) / unit_size, 1);
    return 0;
}

/* cur: in/out (the picture after CDEF / CCSO); dbk: the deblocked picture */
ORC_API int orc_lr_frame(const svtf_geom *g, void *y, int sy, void *cb, int scb, void *cr, int scr, const void *dy, int dsy, const void *dcb,
                         int dscb, const void *dcr, int dscr, const svtf_lr_params *lr, const svtf_lr_unit *const units[3]) {
    void *const       cur[3] = {y, cb, cr};
    const void *const dbk[3] = {dy, dcb, dcr};
    const int         cst[3] = {sy, scb, scr}, dst[3] = {dsy, dscb, dscr};
    const int         sb = g->sample_bytes, bd = g->bit_depth;
    for (int p = 0; p < 3; p++) {
        if (lr->frame_restoration_type[p] == SVTF_RESTORE_NONE)
            continue;
        const int ss = p ? 1 : 0, pw = (g->width + ss) >> ss, ph = (g->height + ss) >> ss, U = lr->restoration_unit_size[p];
        int       hu, vu;
        orc_lr_units(g, p, U, &hu, &vu);
        const int sh = 64 >> ss, off = 8 >> ss;
        uint8_t  *out = malloc((size_t)pw * ph * sb);
        if (!out)
            return -1;
        for (int k = 0;; k++) {
            stripe_in s = {cur[p], dbk[p], cst[p], dst[p], sb, pw, ph, IMAX(0, k * sh - off), IMIN((k + 1) * sh - off, ph), k > 0, 0, lr->optimized_lr};
            if (s.y0 >= ph)
                break;
            s.below = s.y1 < ph;
            for (int yy = s.y0; yy < s.y1; yy++)
                for (int xx = 0; xx < pw; xx++) {
                    const svtf_lr_unit *u = &units[p][IMIN((yy + off) / U, vu - 1) * hu + IMIN(xx / U, hu - 1)];
                    int                 v;
                    if (u->restoration_type == SVTF_RESTORE_WIENER)
                        v = wiener_px(&s, u, bd, xx, yy);
                    else if (u->restoration_type == SVTF_RESTORE_SGRPROJ)
                        v = sgr_px(&s, u, bd, xx, yy);
                    else
                        v = px(cur[p], cst[p], sb, xx, yy);
                    if (sb == 1)
                        out[(size_t)yy * pw + xx] = (uint8_t)v;
                    else
                        ((uint16_t *)out)[(size_t)yy * pw + xx] = (uint16_t)v;
                }
        }
        for (int yy = 0; yy < ph; yy++) memcpy((uint8_t *)cur[p] + (size_t)yy * cst[p] * sb, out + (size_t)yy * pw * sb, (size_t)pw * sb);
        free(out);
    }
    return 0;
}

/* try_restoration_unit_seg (Encoder/Codec/EbRestorationPick.c:129): the unit [h0, h1) x [v0, v1) of one plane filtered with
 * *u WITHOUT stripe boundaries, squared error against the source (sse_restoration_unit :56).  Used by svtf_oracle_lrs.c. */
int64_t orc_lr_try_unit(const void *cur, int cst, const void *src, int sst, int sb, int bd, int ss, int pw, int ph, int h0, int h1, int v0,
                        int v1, const svtf_lr_unit *u) {
    const int sh = 64 >> ss, off = 8 >> ss;
    int64_t   sse = 0;
    for (int k = 0;; k++) {
        stripe_in s = {cur, cur, cst, cst, sb, pw, ph, IMAX(0, k * sh - off), IMIN((k + 1) * sh - off, ph), k > 0, 0, 2};
        if (s.y0 >= ph)
            break;
        for (int yy = IMAX(s.y0, v0); yy < IMIN(s.y1, v1); yy++)
            for (int xx = h0; xx < h1; xx++) {
                int v;
                if (u->restoration_type == SVTF_RESTORE_WIENER)
                    v = wiener_px(&s, u, bd, xx, yy);
                else if (u->restoration_type == SVTF_RESTORE_SGRPROJ)
                    v = sgr_px(&s, u, bd, xx, yy);
                else
                    v = px(cur, cst, sb, xx, yy);
                const int64_t d = v - px(src, sst, sb, xx, yy);
                sse += d * d;
            }
    }
    return sse;
}

/* the table value, for the test that compares it with the reference's svt_aom_eb_x_by_xplus1 */
ORC_API int orc_x_by_xplus1(int z) { return x_by_xplus1(z); }
