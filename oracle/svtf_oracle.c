/*
 * svtf_oracle.c — CPU restatement of the reference's in-loop filter path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under svt-av1_pro-anchor-v2.1.0-_b200/ may include, link or
 * call this file; it is the checker used by tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline leg.  Parity status: PINNED — tests/test_oracle_vs_ref.py checks every function here
 * against the reference's own C code compiled from /root/reference into oracle/_ref/libsvtref.so
 * (recipe: oracle/Makefile, shim: oracle/ref_shim.c), and tests/golden/ holds vectors produced by
 * that library (generator: oracle/make_golden.py) so the pin also holds where the reference tree is
 * absent.  The reference ships no golden vectors / KATs for this path: its own tests
 * are C-vs-SIMD differential tests, whose parameter sweeps tests/ re-uses.
 *
 * All paths below are relative to the reference tree, Source/Lib/.
 * Plain C99, scalar, single-threaded unless the caller parallelises over frames.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include "../include/svtfilt_b200.h"

#define ORC_API __attribute__((visibility("default")))
#define IMIN(a, b) ((a) < (b) ? (a) : (b))
#define IMAX(a, b) ((a) > (b) ? (a) : (b))
static inline int iclamp(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

/* ------------------------------------------------------------------------------------------
 * sample access: planes are uint8_t or uint16_t depending on sample_bytes
 * ---------------------------------------------------------------------------------------- */
typedef struct {
    void *p;
    int   stride; /* samples */
    int   sb;     /* sample bytes */
} plane_t;
static inline int  pget(const plane_t *pl, int x, int y) {
    ptrdiff_t i = (ptrdiff_t)y * pl->stride + x;
    return pl->sb == 1 ? ((const uint8_t *)pl->p)[i] : ((const uint16_t *)pl->p)[i];
}
static inline void pset(const plane_t *pl, int x, int y, int v) {
    ptrdiff_t i = (ptrdiff_t)y * pl->stride + x;
    if (pl->sb == 1)
        ((uint8_t *)pl->p)[i] = (uint8_t)v;
    else
        ((uint16_t *)pl->p)[i] = (uint16_t)v;
}

/* packed mi word accessors (include/svtfilt_b200.h) */
#define MI_TXW_Y(m) ((int)((m) >> 0) & 7)
#define MI_TXH_Y(m) ((int)((m) >> 3) & 7)
#define MI_TXW_UV(m) ((int)((m) >> 6) & 7)
#define MI_TXH_UV(m) ((int)((m) >> 9) & 7)
#define MI_BW(m) ((int)((m) >> 12) & 7)
#define MI_BH(m) ((int)((m) >> 15) & 7)
#define MI_SKIP_INTER(m) ((int)((m) >> 18) & 1)
#define MI_SKIP(m) ((int)((m) >> 19) & 1)
#define MI_REF0(m) ((int)((m) >> 20) & 7)
#define MI_MODE_LF(m) ((int)((m) >> 23) & 1)
#define MI_SEG(m) ((int)((m) >> 24) & 7)

/* ==========================================================================================
 * DEBLOCK
 * ======================================================================================== */

/* signed_char_clamp_high — Common/Codec/EbDeblockingCommon.c:28-35 (bd 8 == signed_char_clamp :26) */
static inline int sclamp(int t, int bd) {
    const int sh = bd - 8;
    return iclamp(t, -(128 << sh), (128 << sh) - 1);
}

/* One line across an edge.  s[k*st] for k=-7..6 are p6..p0,q0..q6.
 * len 4: svt_aom_[highbd_]lpf_*_4_c  (Common/Codec/EbDeblockingCommon.c:242-272, :460-500)
 * len 6: ..._6_c  (:306-338, highbd :655-700)    len 8: ..._8_c (:340-372, :502-544)
 * len 14: mb_lpf_*_edge_w (:787-865, highbd :590-653, :702-760).
 * The 8-bit kernels are the bd==8 instance of the high-bitdepth formulas (x^0x80 == x-128 in
 * int8, every clamp identical); tests/test_oracle_vs_ref.py checks both families. */
static void lpf_line(int *px /* 14 values p6..q6, index 0..13 */, int len, int bd, int blimit, int limit, int thresh) {
    const int sh   = bd - 8;
    const int lim  = limit << sh, blim = blimit << sh, thr = thresh << sh, one = 1 << sh;
    int *const q   = px + 7; /* q[0]=q0, q[-1]=p0 */
#define P(i) q[-1 - (i)]
#define Q(i) q[(i)]
    const int p0 = P(0), p1 = P(1), q0 = Q(0), q1 = Q(1);
    int mask = 1; /* 1 = filter */
    if (abs(p1 - p0) > lim || abs(q1 - q0) > lim || abs(p0 - q0) * 2 + abs(p1 - q1) / 2 > blim)
        mask = 0;
    int flat = 0, flat2 = 0;
    if (len == 6) {
        const int p2 = P(2), q2 = Q(2);
        if (abs(p2 - p1) > lim || abs(q2 - q1) > lim)
            mask = 0;
        flat = !(abs(p1 - p0) > one || abs(q1 - q0) > one || abs(p2 - p0) > one || abs(q2 - q0) > one);
        if (flat && mask) {
            /* filter6 :274-288 */
            P(1) = (p2 * 3 + p1 * 2 + p0 * 2 + q0 + 4) >> 3;
            P(0) = (p2 + p1 * 2 + p0 * 2 + q0 * 2 + q1 + 4) >> 3;
            Q(0) = (p1 + p0 * 2 + q0 * 2 + q1 * 2 + q2 + 4) >> 3;
            Q(1) = (p0 + q0 * 2 + q1 * 2 + q2 * 3 + 4) >> 3;
            return;
        }
    } else if (len >= 8) {
        const int p2 = P(2), p3 = P(3), q2 = Q(2), q3 = Q(3);
        if (abs(p3 - p2) > lim || abs(p2 - p1) > lim || abs(q2 - q1) > lim || abs(q3 - q2) > lim)
            mask = 0;
        flat = !(abs(p1 - p0) > one || abs(q1 - q0) > one || abs(p2 - p0) > one || abs(q2 - q0) > one ||
                 abs(p3 - p0) > one || abs(q3 - q0) > one);
        if (len == 14) {
            const int p4 = P(4), p5 = P(5), p6 = P(6), q4 = Q(4), q5 = Q(5), q6 = Q(6);
            flat2 = !(abs(p4 - p0) > one || abs(q4 - q0) > one || abs(p5 - p0) > one || abs(q5 - q0) > one ||
                      abs(p6 - p0) > one || abs(q6 - q0) > one);
            if (flat2 && flat && mask) {
                /* filter14 :762-785 */
                P(5) = (p6 * 7 + p5 * 2 + p4 * 2 + p3 + p2 + p1 + p0 + q0 + 8) >> 4;
                P(4) = (p6 * 5 + p5 * 2 + p4 * 2 + p3 * 2 + p2 + p1 + p0 + q0 + q1 + 8) >> 4;
                P(3) = (p6 * 4 + p5 + p4 * 2 + p3 * 2 + p2 * 2 + p1 + p0 + q0 + q1 + q2 + 8) >> 4;
                P(2) = (p6 * 3 + p5 + p4 + p3 * 2 + p2 * 2 + p1 * 2 + p0 + q0 + q1 + q2 + q3 + 8) >> 4;
                P(1) = (p6 * 2 + p5 + p4 + p3 + p2 * 2 + p1 * 2 + p0 * 2 + q0 + q1 + q2 + q3 + q4 + 8) >> 4;
                P(0) = (p6 + p5 + p4 + p3 + p2 + p1 * 2 + p0 * 2 + q0 * 2 + q1 + q2 + q3 + q4 + q5 + 8) >> 4;
                Q(0) = (p5 + p4 + p3 + p2 + p1 + p0 * 2 + q0 * 2 + q1 * 2 + q2 + q3 + q4 + q5 + q6 + 8) >> 4;
                Q(1) = (p4 + p3 + p2 + p1 + p0 + q0 * 2 + q1 * 2 + q2 * 2 + q3 + q4 + q5 + q6 * 2 + 8) >> 4;
                Q(2) = (p3 + p2 + p1 + p0 + q0 + q1 * 2 + q2 * 2 + q3 * 2 + q4 + q5 + q6 * 3 + 8) >> 4;
                Q(3) = (p2 + p1 + p0 + q0 + q1 + q2 * 2 + q3 * 2 + q4 * 2 + q5 + q6 * 4 + 8) >> 4;
                Q(4) = (p1 + p0 + q0 + q1 + q2 + q3 * 2 + q4 * 2 + q5 * 2 + q6 * 5 + 8) >> 4;
                Q(5) = (p0 + q0 + q1 + q2 + q3 + q4 * 2 + q5 * 2 + q6 * 7 + 8) >> 4;
                return;
            }
        }
        if (flat && mask) {
            /* filter8 :290-304 */
            P(2) = (p3 + p3 + p3 + 2 * p2 + p1 + p0 + q0 + 4) >> 3;
            P(1) = (p3 + p3 + p2 + 2 * p1 + p0 + q0 + q1 + 4) >> 3;
            P(0) = (p3 + p2 + p1 + 2 * p0 + q0 + q1 + q2 + 4) >> 3;
            Q(0) = (p2 + p1 + p0 + 2 * q0 + q1 + q2 + q3 + 4) >> 3;
            Q(1) = (p1 + p0 + q0 + 2 * q1 + q2 + q3 + q3 + 4) >> 3;
            Q(2) = (p0 + q0 + q1 + 2 * q2 + q3 + q3 + q3 + 4) >> 3;
            return;
        }
    }
    /* filter4 :214-240 / highbd_filter4 :426-458 */
    {
        const int off = 0x80 << sh;
        const int ps1 = p1 - off, ps0 = p0 - off, qs0 = q0 - off, qs1 = q1 - off;
        const int hev = (abs(p1 - p0) > thr || abs(q1 - q0) > thr) ? -1 : 0;
        const int m   = mask ? -1 : 0;
        int       f   = sclamp(ps1 - qs1, bd) & hev;
        f             = sclamp(f + 3 * (qs0 - ps0), bd) & m;
        const int f1  = sclamp(f + 4, bd) >> 3;
        const int f2  = sclamp(f + 3, bd) >> 3;
        Q(0)          = sclamp(qs0 - f1, bd) + off;
        P(0)          = sclamp(ps0 + f2, bd) + off;
        f             = ((f1 + 1) >> 1) & ~hev;
        Q(1)          = sclamp(qs1 - f, bd) + off;
        P(1)          = sclamp(ps1 + f, bd) + off;
    }
#undef P
#undef Q
}

/* Leaf-level entry used by the kernel sweep tests (design of test/DeblockTest.cc:95-309):
 * filters `count` lines starting at s; `along` = distance between the samples of one line,
 * `next` = distance between consecutive lines. */
ORC_API void orc_lpf_edge(void *s, int sample_bytes, int along, int next, int count, int len, int bd, int blimit,
                          int limit, int thresh) {
    for (int l = 0; l < count; l++) {
        int px[14];
        for (int k = 0; k < 14; k++) {
            ptrdiff_t i = (ptrdiff_t)l * next + (ptrdiff_t)(k - 7) * along;
            int       inside = (k >= 7 - (len == 14 ? 7 : len == 8 ? 4 : len == 6 ? 3 : 2)) &&
                         (k < 7 + (len == 14 ? 7 : len == 8 ? 4 : len == 6 ? 3 : 2));
            px[k] = inside ? (sample_bytes == 1 ? ((uint8_t *)s)[i] : ((uint16_t *)s)[i]) : 0;
        }
        lpf_line(px, len, bd, blimit, limit, thresh);
        const int mod = len == 14 ? 6 : len == 8 ? 3 : 2;
        for (int k = 7 - mod; k < 7 + mod; k++) {
            ptrdiff_t i = (ptrdiff_t)l * next + (ptrdiff_t)(k - 7) * along;
            if (sample_bytes == 1)
                ((uint8_t *)s)[i] = (uint8_t)px[k];
            else
                ((uint16_t *)s)[i] = (uint16_t)px[k];
        }
    }
}

/* Level table lvl[plane][seg][dir][ref][mode_lf] and the sharpness limits.
 * svt_av1_loop_filter_frame_init (Common/Codec/EbDeblockingCommon.c:76-137), or, when
 * delta_lf_present, svt_aom_get_filter_level_delta_lf (:37-71) evaluated for every key (its
 * sb_delta_lf argument is the frame-level ppcs->curr_delta_lf, Encoder/Codec/EbDeblockingFilter.c:213). */
typedef struct {
    uint8_t lvl[3][8][2][8][2];
    uint8_t lim[64], mblim[64], hev[64];
} lf_tables;

static const int seg_lvl_lf_lut[3][2]  = {{1, 2}, {3, 3}, {4, 4}}; /* EbDeblockingCommon.c:17-20 */
static const int delta_lf_id_lut[3][2] = {{0, 1}, {2, 2}, {3, 3}}; /* EbDeblockingCommon.h */

ORC_API void orc_lf_build_tables(const svtf_lf_params *lf, lf_tables *t) {
    memset(t, 0, sizeof(*t));
    /* svt_aom_update_sharpness :554-573 ; hev: Encoder/Codec/EbDeblockingFilter.c:47 */
    for (int lvl = 0; lvl <= 63; lvl++) {
        int bil = lvl >> ((lf->sharpness_level > 0) + (lf->sharpness_level > 4));
        if (lf->sharpness_level > 0 && bil > 9 - lf->sharpness_level)
            bil = 9 - lf->sharpness_level;
        if (bil < 1)
            bil = 1;
        t->lim[lvl]   = (uint8_t)bil;
        t->mblim[lvl] = (uint8_t)(2 * (lvl + 2) + bil);
        t->hev[lvl]   = (uint8_t)(lvl >> 4);
    }
    for (int plane = 0; plane < 3; plane++)
        for (int seg = 0; seg < 8; seg++)
            for (int dir = 0; dir < 2; dir++) {
                int base = plane == 0 ? lf->filter_level[dir] : plane == 1 ? lf->filter_level_u : lf->filter_level_v;
                if (lf->delta_lf_present) {
                    int dlf = lf->delta_lf_multi ? lf->curr_delta_lf[delta_lf_id_lut[plane][dir]] : lf->curr_delta_lf[0];
                    base    = iclamp(dlf + base, 0, 63);
                }
                int       lvl_seg = base;
                const int feat    = seg_lvl_lf_lut[plane][dir];
                if (lf->segmentation_enabled && lf->seg_feature_enabled[seg][feat])
                    lvl_seg = iclamp(lvl_seg + lf->seg_feature_data[seg][feat], 0, 63);
                for (int ref = 0; ref < 8; ref++)
                    for (int m = 0; m < 2; m++) {
                        int v = lvl_seg;
                        if (lf->mode_ref_delta_enabled) {
                            const int scale = 1 << (lvl_seg >> 5);
                            v += lf->ref_deltas[ref] * scale;
                            if (ref > 0)
                                v += lf->mode_deltas[m] * scale;
                            v = iclamp(v, 0, 63);
                        }
                        t->lvl[plane][seg][dir][ref][m] = (uint8_t)v;
                    }
            }
}

/* set_lpf_parameters — Encoder/Codec/EbDeblockingFilter.c:162-285.
 * Returns the filter length (0 = none) and the level for the unit at plane position (x,y);
 * *tx_dim_log2 receives log2 of the current unit's transform size along the walk direction. */
/* per-superblock level tables (decoder: SBInfo.sb_delta_lf, Decoder/Codec/EbDecLF.c:159-193; the table of the superblock
 * containing a unit serves both as "current" and as "left/above" level, :357,548); NULL = one table for the frame */
typedef struct {
    const lf_tables *t;
    int              sb_cols, sb_shift;
} sb_tables;
static const lf_tables *lvl_tab(const lf_tables *t, const sb_tables *sbt, int mi_row, int mi_col) {
    return sbt ? &sbt->t[(mi_row >> sbt->sb_shift) * sbt->sb_cols + (mi_col >> sbt->sb_shift)] : t;
}

static int lpf_params(const svtf_geom *g, const svtf_mi_t *mi, int mi_stride, const lf_tables *t, const sb_tables *sbt, int plane,
                      int dir, int x, int y, int *level, int *tx_dim_log2) {
    const int ss = plane ? 1 : 0;
    /* a chroma unit has edges while its LUMA position is inside the picture: (w + 1) >> 1, not w >> 1, when a super-resolution frame is
     * odd-sized (Encoder/Codec/EbDeblockingFilter.c:305-322, Decoder/Codec/EbDecLF.c:116-121; unscaled encoder pictures are even) */
    const int pw = (g->width + ss) >> ss, ph = (g->height + ss) >> ss;
    *tx_dim_log2 = 2;
    if (x >= pw || y >= ph)
        return 0;
    const int       mi_row = ss | ((y << ss) >> 2);
    const int       mi_col = ss | ((x << ss) >> 2);
    const svtf_mi_t cur    = mi[mi_row * mi_stride + mi_col];
    const int       ts     = dir == 0 ? (plane ? MI_TXW_UV(cur) : MI_TXW_Y(cur)) : (plane ? MI_TXH_UV(cur) : MI_TXH_Y(cur));
    *tx_dim_log2           = ts + 2;
    const int coord        = dir == 0 ? x : y;
    if (coord & ((4 << ts) - 1))
        return 0;
    const int curr_level = lvl_tab(t, sbt, mi_row, mi_col)->lvl[plane][MI_SEG(cur)][dir][MI_REF0(cur)][MI_MODE_LF(cur)];
    if (!coord)
        return 0;
    const svtf_mi_t prv    = dir == 0 ? mi[mi_row * mi_stride + mi_col - (1 << ss)] : mi[(mi_row - (1 << ss)) * mi_stride + mi_col];
    const int       pv_ts  = dir == 0 ? (plane ? MI_TXW_UV(prv) : MI_TXW_Y(prv)) : (plane ? MI_TXH_UV(prv) : MI_TXH_Y(prv));
    const lf_tables *pt    = dir == 0 ? lvl_tab(t, sbt, mi_row, mi_col - (1 << ss)) : lvl_tab(t, sbt, mi_row - (1 << ss), mi_col);
    const int       pv_lvl = pt->lvl[plane][MI_SEG(prv)][dir][MI_REF0(prv)][MI_MODE_LF(prv)];
    /* plane block size along the direction: get_plane_block_size (Common/Codec/EbUtility.h:86-115);
     * for 4:2:0 every entry of ss_size_lookup[.][1][1] is max(4, dim/2) */
    int bdim = 4 << (dir == 0 ? MI_BW(cur) : MI_BH(cur));
    if (ss)
        bdim = IMAX(4, bdim >> 1);
    const int pu_edge = !(coord & (bdim - 1));
    if ((curr_level || pv_lvl) && (!MI_SKIP_INTER(prv) || !MI_SKIP_INTER(cur) || pu_edge)) {
        const int min_ts = IMIN(ts, pv_ts);
        *level           = curr_level ? curr_level : pv_lvl;
        if (min_ts == 0)
            return 4;
        return plane ? 6 : (min_ts == 1 ? 8 : 14);
    }
    return 0;
}

/* svt_av1_loop_filter_frame — Encoder/Codec/EbDeblockingFilter.c:624-651 (+ svt_aom_loop_filter_sb
 * :547-618, svt_av1_filter_block_plane_vert/horz :287-545).  The reference interleaves the passes
 * per superblock (vertical edges of SB x, then horizontal edges of SB x-1); because edges of one
 * pass never overlap this equals "all vertical edges, then all horizontal edges",
 * which is the order used here (and on the GPU).  Checked against the real function. */
static int deblock_frame_impl(const svtf_geom *g, void *y, int sy, void *cb, int scb, void *cr, int scr, const svtf_mi_t *mi,
                              int mi_stride, const svtf_lf_params *lf, int plane_start, int plane_end, const sb_tables *sbt) {
    lf_tables t;
    orc_lf_build_tables(lf, &t);
    plane_t pl[3] = {{y, sy, g->sample_bytes}, {cb, scb, g->sample_bytes}, {cr, scr, g->sample_bytes}};
    for (int plane = plane_start; plane < IMIN(plane_end, 3); plane++) {
        if (plane == 0 && !lf->filter_level[0] && !lf->filter_level[1])
            break;
        if (plane == 1 && !lf->filter_level_u)
            continue;
        if (plane == 2 && !lf->filter_level_v)
            continue;
        const int ss = plane ? 1 : 0;
        /* the SB loop covers aligned_width x aligned_height; units beyond width/height return early */
        const int cw = (g->mi_cols * 4) >> ss, chh = (g->mi_rows * 4) >> ss;
        for (int dir = 0; dir < 2; dir++) {
            for (int a = 0; a < (dir == 0 ? chh : cw); a += 4) {      /* 4-line unit across the walk */
                for (int b = 0; b < (dir == 0 ? cw : chh);) {         /* walk along the direction   */
                    const int x = dir == 0 ? b : a, yy = dir == 0 ? a : b;
                    int       level = 0, tlog2 = 2;
                    const int len = lpf_params(g, mi, mi_stride, &t, sbt, plane, dir, x, yy, &level, &tlog2);
                    if (len) {
                        for (int l = 0; l < 4; l++) {
                            int px[14];
                            const int half = len == 14 ? 7 : len == 8 ? 4 : len == 6 ? 3 : 2;
                            for (int k = -half; k < half; k++)
                                px[7 + k] = dir == 0 ? pget(&pl[plane], x + k, yy + l) : pget(&pl[plane], x + l, yy + k);
                            lpf_line(px, len, g->bit_depth, t.mblim[level], t.lim[level], t.hev[level]);
                            const int mod = len == 14 ? 6 : len == 8 ? 3 : 2;
                            for (int k = -mod; k < mod; k++) {
                                if (dir == 0)
                                    pset(&pl[plane], x + k, yy + l, px[7 + k]);
                                else
                                    pset(&pl[plane], x + l, yy + k, px[7 + k]);
                            }
                        }
                    }
                    b += 1 << tlog2;
                }
            }
        }
    }
    return 0;
}

ORC_API int orc_deblock_frame(const svtf_geom *g, void *y, int sy, void *cb, int scb, void *cr, int scr,
                              const svtf_mi_t *mi, int mi_stride, const svtf_lf_params *lf, int plane_start,
                              int plane_end) {
    return deblock_frame_impl(g, y, sy, cb, scb, cr, scr, mi, mi_stride, lf, plane_start, plane_end, NULL);
}

/* svt_aom_dec_av1_loop_filter_frame with delta_lf_present — Decoder/Codec/EbDecLF.c:697: every superblock carries its
 * own deltas (sb_delta_lf[sb][FRAME_LF_COUNT]); checked against the real decoder function (tests/test_dec_oracle.py). */
ORC_API int orc_deblock_frame_sb_delta(const svtf_geom *g, void *y, int sy, void *cb, int scb, void *cr, int scr, const svtf_mi_t *mi,
                                       int mi_stride, const svtf_lf_params *lf, int plane_start, int plane_end,
                                       const int32_t *sb_delta_lf, int sb_cols, int sb_size_log2) {
    if (!lf->delta_lf_present)
        return deblock_frame_impl(g, y, sy, cb, scb, cr, scr, mi, mi_stride, lf, plane_start, plane_end, NULL);
    const int  sb_mi = 1 << (sb_size_log2 - 2), rows = (g->mi_rows + sb_mi - 1) / sb_mi;
    lf_tables *tabs  = malloc(sizeof(lf_tables) * rows * sb_cols);
    if (!tabs)
        return -1;
    svtf_lf_params l = *lf;
    for (int sb = 0; sb < rows * sb_cols; sb++) {
        for (int i = 0; i < 4; i++) l.curr_delta_lf[i] = sb_delta_lf[sb * 4 + i];
        orc_lf_build_tables(&l, &tabs[sb]);
    }
    const sb_tables sbt = {tabs, sb_cols, sb_size_log2 - 2};
    const int       r   = deblock_frame_impl(g, y, sy, cb, scb, cr, scr, mi, mi_stride, lf, plane_start, plane_end, &sbt);
    free(tabs);
    return r;
}

/* ==========================================================================================
 * CDEF
 * ======================================================================================== */
#define CDEF_VERY_LARGE 0x7F7F /* Common/Codec/EbCdef.h:39 */

static inline int msb(unsigned n) { return 31 - __builtin_clz(n); }

/* constrain — Common/Codec/EbCdef.c:85-91 */
static inline int constrain(int diff, int threshold, int damping) {
    if (!threshold)
        return 0;
    const int shift = IMAX(0, damping - msb((unsigned)threshold));
    const int ad    = abs(diff);
    const int v     = IMIN(ad, IMAX(0, threshold - (ad >> shift)));
    return diff < 0 ? -v : v;
}

/* Cdef_Directions as (dy,dx) for k=0,1 — Common/Codec/EbCdef.c:99-122 */
static const int8_t cdef_dirs[8][2][2] = {{{-1, 1}, {-2, 2}}, {{0, 1}, {-1, 2}}, {{0, 1}, {0, 2}}, {{0, 1}, {1, 2}},
                                          {{1, 1}, {2, 2}},   {{1, 0}, {2, 1}},  {{1, 0}, {2, 0}}, {{1, 0}, {2, -1}}};

/* svt_aom_cdef_find_dir_c — Common/Codec/EbCdef.c:150-210.  img: 8x8, row stride `stride`. */
ORC_API int orc_cdef_find_dir(const uint16_t *img, int stride, int32_t *var, int coeff_shift) {
    int32_t cost[8] = {0}, partial[8][15];
    memset(partial, 0, sizeof(partial));
    static const int div_table[9] = {0, 840, 420, 280, 210, 168, 140, 120, 105};
    for (int i = 0; i < 8; i++)
        for (int j = 0; j < 8; j++) {
            const int x = (img[i * stride + j] >> coeff_shift) - 128;
            partial[0][i + j] += x;
            partial[1][i + j / 2] += x;
            partial[2][i] += x;
            partial[3][3 + i - j / 2] += x;
            partial[4][7 + i - j] += x;
            partial[5][3 - i / 2 + j] += x;
            partial[6][j] += x;
            partial[7][i / 2 + j] += x;
        }
    for (int i = 0; i < 8; i++) {
        cost[2] += partial[2][i] * partial[2][i];
        cost[6] += partial[6][i] * partial[6][i];
    }
    cost[2] *= div_table[8];
    cost[6] *= div_table[8];
    for (int i = 0; i < 7; i++) {
        cost[0] += (partial[0][i] * partial[0][i] + partial[0][14 - i] * partial[0][14 - i]) * div_table[i + 1];
        cost[4] += (partial[4][i] * partial[4][i] + partial[4][14 - i] * partial[4][14 - i]) * div_table[i + 1];
    }
    cost[0] += partial[0][7] * partial[0][7] * div_table[8];
    cost[4] += partial[4][7] * partial[4][7] * div_table[8];
    for (int i = 1; i < 8; i += 2) {
        for (int j = 0; j < 5; j++) cost[i] += partial[i][3 + j] * partial[i][3 + j];
        cost[i] *= div_table[8];
        for (int j = 0; j < 3; j++)
            cost[i] += (partial[i][j] * partial[i][j] + partial[i][10 - j] * partial[i][10 - j]) * div_table[2 * j + 2];
    }
    int best_cost = 0, best_dir = 0;
    for (int i = 0; i < 8; i++)
        if (cost[i] > best_cost) {
            best_cost = cost[i];
            best_dir  = i;
        }
    *var = (best_cost - cost[(best_dir + 4) & 7]) >> 10;
    return best_dir;
}

/* adjust_strength — Common/Codec/EbCdef.c:130-134 */
static inline int adjust_strength(int strength, int var) {
    const int i = (var >> 6) ? IMIN(msb((unsigned)(var >> 6)), 12) : 0;
    return var ? (strength * (4 + i) + 8) >> 4 : 0;
}

/* svt_cdef_filter_block_c for one sample — Common/Codec/EbCdef.c:253-304.
 * `in` points at the sample inside a 16-bit tile with row stride `s`. */
static inline int cdef_filter_px(const uint16_t *in, int s, int pri_strength, int sec_strength, int dir, int pri_damping,
                                 int sec_damping, int coeff_shift) {
    const int     pri_taps[2][2] = {{4, 2}, {3, 3}};
    const int     sec_taps[2]    = {2, 1};
    const int    *pt             = pri_taps[(pri_strength >> coeff_shift) & 1];
    int16_t       sum            = 0;
    const int16_t x              = (int16_t)in[0];
    int           mx = x, mn = x;
    for (int k = 0; k < 2; k++) {
        const int     po = cdef_dirs[dir][k][0] * s + cdef_dirs[dir][k][1];
        const int16_t p0 = (int16_t)in[po], p1 = (int16_t)in[-po];
        sum += (int16_t)(pt[k] * constrain(p0 - x, pri_strength, pri_damping));
        sum += (int16_t)(pt[k] * constrain(p1 - x, pri_strength, pri_damping));
        if (p0 != CDEF_VERY_LARGE) mx = IMAX(p0, mx);
        if (p1 != CDEF_VERY_LARGE) mx = IMAX(p1, mx);
        mn = IMIN(p0, mn);
        mn = IMIN(p1, mn);
        const int     d2 = (dir + 2) & 7, d3 = (dir + 6) & 7; /* padded table == &7 (:93-98) */
        const int     o2 = cdef_dirs[d2][k][0] * s + cdef_dirs[d2][k][1];
        const int     o3 = cdef_dirs[d3][k][0] * s + cdef_dirs[d3][k][1];
        const int16_t s0 = (int16_t)in[o2], s1 = (int16_t)in[-o2], s2 = (int16_t)in[o3], s3 = (int16_t)in[-o3];
        if (s0 != CDEF_VERY_LARGE) mx = IMAX(s0, mx);
        if (s1 != CDEF_VERY_LARGE) mx = IMAX(s1, mx);
        if (s2 != CDEF_VERY_LARGE) mx = IMAX(s2, mx);
        if (s3 != CDEF_VERY_LARGE) mx = IMAX(s3, mx);
        mn = IMIN(s0, mn);
        mn = IMIN(s1, mn);
        mn = IMIN(s2, mn);
        mn = IMIN(s3, mn);
        sum += (int16_t)(sec_taps[k] * constrain(s0 - x, sec_strength, sec_damping));
        sum += (int16_t)(sec_taps[k] * constrain(s1 - x, sec_strength, sec_damping));
        sum += (int16_t)(sec_taps[k] * constrain(s2 - x, sec_strength, sec_damping));
        sum += (int16_t)(sec_taps[k] * constrain(s3 - x, sec_strength, sec_damping));
    }
    return (int16_t)iclamp((int16_t)x + ((8 + sum - (sum < 0)) >> 4), mn, mx);
}

/* Leaf entry for the block sweep test (design of test/CdefTest.cc:94-411): filters a bw x bh
 * block inside a 16-bit tile, writing a packed (stride bw) 16-bit result. */
ORC_API void orc_cdef_filter_block(uint16_t *dst, int dstride, const uint16_t *in, int in_stride, int pri_strength,
                                   int sec_strength, int dir, int pri_damping, int sec_damping, int bw, int bh,
                                   int coeff_shift, int subsampling_factor) {
    for (int i = 0; i < bh; i += subsampling_factor)
        for (int j = 0; j < bw; j++)
            dst[i * dstride + j] = (uint16_t)cdef_filter_px(in + i * in_stride + j, in_stride, pri_strength, sec_strength,
                                                            dir, pri_damping, sec_damping, coeff_shift);
}

/* 16-bit copy of one plane of the coded frame with a 2-sample CDEF_VERY_LARGE apron
 * (what svt_av1_cdef_frame's src[] staging amounts to, Encoder/Codec/EbEncCdef.c:459-568:
 * every tap sees pre-CDEF samples; outside the coded frame -> CDEF_VERY_LARGE). */
typedef struct {
    uint16_t *buf;
    int       w, h, stride; /* origin at buf + 2*stride + 2 */
} tile16;
static int tile16_from_plane(tile16 *t, const plane_t *pl, int w, int h) {
    t->w      = w;
    t->h      = h;
    t->stride = w + 4;
    t->buf    = (uint16_t *)malloc(sizeof(uint16_t) * (size_t)(w + 4) * (h + 4));
    if (!t->buf)
        return -1;
    for (size_t i = 0; i < (size_t)(w + 4) * (h + 4); i++) t->buf[i] = CDEF_VERY_LARGE;
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++) t->buf[(y + 2) * t->stride + x + 2] = (uint16_t)pget(pl, x, y);
    return 0;
}
#define T16(t, x, y) ((t)->buf + ((y) + 2) * (t)->stride + (x) + 2)

/* svt_sb_compute_cdef_list — Encoder/Codec/EbEncCdef.c:238-281 (64x64 filter blocks). */
static int cdef_list(const svtf_geom *g, const svtf_mi_t *mi, int mi_stride, int fbr, int fbc, uint8_t *by, uint8_t *bx) {
    const int mi_row = fbr * 16, mi_col = fbc * 16;
    const int maxr = IMIN(g->mi_rows - mi_row, 16), maxc = IMIN(g->mi_cols - mi_col, 16);
    int       n = 0;
    for (int r = 0; r < maxr; r += 2)
        for (int c = 0; c < maxc; c += 2) {
            const svtf_mi_t *m = mi + (mi_row + r) * mi_stride + mi_col + c;
            if (!MI_SKIP(m[0]) || !MI_SKIP(m[1]) || !MI_SKIP(m[mi_stride]) || !MI_SKIP(m[mi_stride + 1])) {
                by[n] = (uint8_t)(r >> 1);
                bx[n] = (uint8_t)(c >> 1);
                n++;
            }
        }
    return n;
}

/* svt_av1_cdef_frame — Encoder/Codec/EbEncCdef.c:284-612, with svt_cdef_filter_fb
 * (Common/Codec/EbCdef.c:339-430) inlined for dstride != 0.  In place on the planes. */
ORC_API int orc_cdef_apply_frame(const svtf_geom *g, void *y, int sy, void *cb, int scb, void *cr, int scr,
                                 const svtf_mi_t *mi, int mi_stride, const svtf_cdef_params *cp,
                                 const int8_t *fb_strength_idx) {
    plane_t   pl[3] = {{y, sy, g->sample_bytes}, {cb, scb, g->sample_bytes}, {cr, scr, g->sample_bytes}};
    const int cs    = IMAX(g->bit_depth - 8, 0);
    const int nvfb = (g->mi_rows + 15) / 16, nhfb = (g->mi_cols + 15) / 16;
    tile16    pre[3];
    for (int p = 0; p < 3; p++)
        if (tile16_from_plane(&pre[p], &pl[p], (g->mi_cols * 4) >> (p ? 1 : 0), (g->mi_rows * 4) >> (p ? 1 : 0)))
            return -1;
    for (int fbr = 0; fbr < nvfb; fbr++)
        for (int fbc = 0; fbc < nhfb; fbc++) {
            const int idx = fb_strength_idx[fbr * nhfb + fbc];
            if (idx < 0)
                continue; /* never searched == all-skip fb: the reference reads a stale index but
                             then `continue`s on cdef_count==0 (EbEncCdef.c:396-401) */
            int level = cp->cdef_y_strength[idx] / 4, sec = cp->cdef_y_strength[idx] % 4;
            sec += sec == 3;
            int uv_level = cp->cdef_uv_strength[idx] / 4, uv_sec = cp->cdef_uv_strength[idx] % 4;
            uv_sec += uv_sec == 3;
            uint8_t by[64], bx[64];
            int     count;
            if ((level == 0 && sec == 0 && uv_level == 0 && uv_sec == 0) ||
                (count = cdef_list(g, mi, mi_stride, fbr, fbc, by, bx)) == 0)
                continue;
            /* directions from the pre-CDEF luma (cdef_find_dir, EbCdef.c:218-247); identical to the
             * values the search stored in pcs->cdef_dir_data */
            int     dir[64];
            int32_t var[64];
            for (int bi = 0; bi < count; bi++)
                dir[bi] = orc_cdef_find_dir(T16(&pre[0], fbc * 64 + bx[bi] * 8, fbr * 64 + by[bi] * 8), pre[0].stride,
                                            &var[bi], cs);
            for (int p = 0; p < 3; p++) {
                const int lv = p ? uv_level : level, sc = p ? uv_sec : sec;
                if (!lv && !sc)
                    continue; /* plane untouched (or filtered with zero strengths == identity) */
                const int pri = lv << cs, secs = sc << cs;
                const int damp = cp->cdef_damping + cs - (p != 0);
                const int bs   = p ? 4 : 8;
                for (int bi = 0; bi < count; bi++) {
                    const int t  = p ? pri : adjust_strength(pri, var[bi]);
                    const int d  = pri ? dir[bi] : 0;
                    const int x0 = (fbc * 64 >> (p ? 1 : 0)) + bx[bi] * bs, y0 = (fbr * 64 >> (p ? 1 : 0)) + by[bi] * bs;
                    for (int i = 0; i < bs; i++)
                        for (int j = 0; j < bs; j++)
                            pset(&pl[p], x0 + j, y0 + i,
                                 cdef_filter_px(T16(&pre[p], x0 + j, y0 + i), pre[p].stride, t, secs, d, damp, damp, cs));
                }
            }
        }
    for (int p = 0; p < 3; p++) free(pre[p].buf);
    return 0;
}

/* ==========================================================================================
 * CDEF strength search
 * ======================================================================================== */

/* dist_8xn_16bit_c / dist_8xn_8bit_c — Encoder/Codec/EbEncCdef.c:23-48, :76-101.
 * a: source picture block (stride as), b: filtered packed 8x8 block (stride 8). */
static uint64_t dist_8xn(const plane_t *srcpl, int sx, int sy_, const uint16_t *flt, int coeff_shift, int height,
                         int subsampling_factor) {
    uint64_t svar = 0, dvar = 0;
    uint64_t sum_s = 0, sum_d = 0, sum_s2 = 0, sum_d2 = 0, sum_sd = 0;
    for (int i = 0; i < height; i += subsampling_factor)
        for (int j = 0; j < 8; j++) {
            const uint64_t d = (uint64_t)pget(srcpl, sx + j, sy_ + i); /* "dst" in the reference = source picture */
            const uint64_t s = flt[8 * i + j];                         /* "src" = filtered                      */
            sum_s += s;
            sum_d += d;
            sum_s2 += s * s;
            sum_d2 += d * d;
            sum_sd += s * d;
        }
    /* Consider subsampling: variances are normalised as if all rows were summed (:39-44) */
    svar = sum_s2 - ((sum_s * sum_s + 32) >> 6);
    dvar = sum_d2 - ((sum_d * sum_d + 32) >> 6);
    return (uint64_t)floor(.5 + (sum_d2 + sum_s2 - 2 * sum_sd) * .5 * (svar + dvar + (400 << 2 * coeff_shift)) /
                                    (sqrt((20000 << 4 * coeff_shift) + svar * (double)dvar)));
}

/* Leaf entry: svt_aom_compute_cdef_dist_c / _8bit_c — Encoder/Codec/EbEncCdef.c:129-219.
 * src plane origin at the fb; flt: packed list of filtered blocks in dlist order. */
ORC_API uint64_t orc_cdef_dist(const void *src, int sstride, int sample_bytes, const uint16_t *flt, const uint8_t *by,
                               const uint8_t *bx, int count, int bw, int bh, int coeff_shift, int pli,
                               int subsampling_factor) {
    plane_t  sp  = {(void *)src, sstride, sample_bytes};
    uint64_t sum = 0;
    for (int bi = 0; bi < count; bi++) {
        const uint16_t *f = flt + (size_t)bi * bw * bh;
        if (pli == 0 && bw == 8 && bh == 8) {
            sum += dist_8xn(&sp, bx[bi] * 8, by[bi] * 8, f, coeff_shift, 8, subsampling_factor);
        } else {
            /* mse_8xn/4xn — :50-74, :103-126 */
            for (int i = 0; i < bh; i += subsampling_factor)
                for (int j = 0; j < bw; j++) {
                    const int e = pget(&sp, bx[bi] * bw + j, by[bi] * bh + i) - f[i * bw + j];
                    sum += (uint64_t)(e * e);
                }
        }
    }
    return sum >> 2 * coeff_shift;
}

static void search_lists(const svtf_cdef_search_cfg *c, int *n, int fs[64], int fs_uv[64]) {
    *n = c->first_pass_fs_num + c->default_second_pass_fs_num;
    for (int gi = 0; gi < *n; gi++) {
        if (gi < c->first_pass_fs_num) {
            fs[gi]    = c->default_first_pass_fs[gi];
            fs_uv[gi] = c->default_first_pass_fs_uv[gi];
        } else {
            fs[gi]    = c->default_second_pass_fs[gi - c->first_pass_fs_num];
            fs_uv[gi] = c->default_second_pass_fs_uv[gi - c->first_pass_fs_num];
        }
    }
}

/* cdef_seg_search over the whole frame — Encoder/Codec/EbCdefProcess.c:114-357 (64x64 SBs).
 * rec: post-deblock picture, src: source picture.  Outputs as in svtf_cdef_search_frame. */
ORC_API int orc_cdef_search_frame(const svtf_geom *g, const void *ry, int rsy, const void *rcb, int rscb, const void *rcr,
                                  int rscr, const void *oy, int osy, const void *ocb, int oscb, const void *ocr,
                                  int oscr, const svtf_mi_t *mi, int mi_stride, const svtf_cdef_search_cfg *cfg,
                                  uint64_t *mse, uint8_t *skip, uint8_t *dir_out, int32_t *var_out) {
    plane_t   rec[3] = {{(void *)ry, rsy, g->sample_bytes}, {(void *)rcb, rscb, g->sample_bytes}, {(void *)rcr, rscr, g->sample_bytes}};
    plane_t   org[3] = {{(void *)oy, osy, g->sample_bytes}, {(void *)ocb, oscb, g->sample_bytes}, {(void *)ocr, oscr, g->sample_bytes}};
    const int cs     = IMAX(g->bit_depth - 8, 0);
    const int nvfb = (g->mi_rows + 15) / 16, nhfb = (g->mi_cols + 15) / 16, nfb = nvfb * nhfb;
    const int damping = 3 + (cfg->base_q_idx >> 6);
    int       n, fs[64], fs_uv[64];
    search_lists(cfg, &n, fs, fs_uv);
    tile16 pre[3];
    for (int p = 0; p < 3; p++)
        if (tile16_from_plane(&pre[p], &rec[p], (g->mi_cols * 4) >> (p ? 1 : 0), (g->mi_rows * 4) >> (p ? 1 : 0)))
            return -1;
    uint16_t *flt = (uint16_t *)malloc(sizeof(uint16_t) * 64 * 64);
    for (int fbr = 0; fbr < nvfb; fbr++)
        for (int fbc = 0; fbc < nhfb; fbc++) {
            const int fb = fbr * nhfb + fbc;
            uint8_t   by[64], bx[64];
            const int count = cdef_list(g, mi, mi_stride, fbr, fbc, by, bx);
            if (skip)
                skip[fb] = count == 0;
            if (count == 0)
                continue;
            int     dir[64] = {0}, dirinit = 0;
            int32_t var[64] = {0};
            for (int p = 0; p < 3; p++) {
                const int bs = p ? 4 : 8, ssx = p ? 1 : 0;
                int       sub = cfg->subsampling_factor;
                sub           = IMIN(sub, p ? 1 : 4); /* :246-251 (BLOCK_8X8 -> 4, BLOCK_4X4 -> 1) */
                for (int gi = 0; gi < n; gi++) {
                    if (p && fs_uv[gi] == -1) {
                        mse[((size_t)nfb + fb) * 64 + gi] = (uint64_t)1040400 * 64; /* default_mse_uv*64 :86,:259 */
                        continue;
                    }
                    const int pri_l = fs[gi] / 4;
                    int       sec_l = fs[gi] % 4;
                    sec_l += sec_l == 3;
                    const int pri = pri_l << cs, secs = sec_l << cs;
                    const int damp = damping + cs - (p != 0);
                    if (pri == 0 && secs == 0) {
                        /* plain copy, dirinit untouched — Common/Codec/EbCdef.c:356-381 */
                        for (int bi = 0; bi < count; bi++)
                            for (int i = 0; i < bs; i += sub)
                                for (int j = 0; j < bs; j++)
                                    flt[bi * bs * bs + i * bs + j] =
                                        *T16(&pre[p], (fbc * 64 >> ssx) + bx[bi] * bs + j, (fbr * 64 >> ssx) + by[bi] * bs + i);
                    } else {
                        if (p == 0 && !dirinit) {
                            for (int bi = 0; bi < count; bi++)
                                dir[bi] = orc_cdef_find_dir(T16(&pre[0], fbc * 64 + bx[bi] * 8, fbr * 64 + by[bi] * 8),
                                                            pre[0].stride, &var[bi], cs);
                            dirinit = 1;
                        }
                        for (int bi = 0; bi < count; bi++) {
                            const int t = p ? pri : adjust_strength(pri, var[bi]);
                            const int d = pri ? dir[bi] : 0;
                            for (int i = 0; i < bs; i += sub)
                                for (int j = 0; j < bs; j++)
                                    flt[bi * bs * bs + i * bs + j] = (uint16_t)cdef_filter_px(
                                        T16(&pre[p], (fbc * 64 >> ssx) + bx[bi] * bs + j, (fbr * 64 >> ssx) + by[bi] * bs + i),
                                        pre[p].stride, t, secs, d, damp, damp, cs);
                        }
                    }
                    /* compute_cdef_dist :60-107 */
                    plane_t sp = org[p];
                    sp.p       = (uint8_t *)sp.p + ((size_t)(fbr * 64 >> ssx) * sp.stride + (fbc * 64 >> ssx)) * sp.sb;
                    const uint64_t cur = orc_cdef_dist(sp.p, sp.stride, sp.sb, flt, by, bx, count, bs, bs, cs, p, sub);
                    if (p < 2)
                        mse[((size_t)p * nfb + fb) * 64 + gi] = cur * sub;
                    else
                        mse[((size_t)nfb + fb) * 64 + gi] += cur * sub;
                }
            }
            if (dir_out)
                for (int bi = 0; bi < count; bi++) {
                    dir_out[fb * 64 + by[bi] * 8 + bx[bi]] = (uint8_t)dir[bi];
                    var_out[fb * 64 + by[bi] * 8 + bx[bi]] = var[bi];
                }
        }
    free(flt);
    for (int p = 0; p < 3; p++) free(pre[p].buf);
    return 0;
}

/* svt_search_one_dual_c — Encoder/Codec/EbEncCdef.c:627-684 */
static uint64_t search_one_dual(int *lev0, int *lev1, int nb, const uint64_t *const *mse0, const uint64_t *const *mse1,
                                int sb_count, int start_gi, int end_gi) {
    static uint64_t tot[64][64];
    uint64_t        best_tot = (uint64_t)1 << 63;
    int             b0 = 0, b1 = 0;
    memset(tot, 0, sizeof(tot));
    for (int i = 0; i < sb_count; i++) {
        uint64_t best = (uint64_t)1 << 63;
        for (int gi = 0; gi < nb; gi++) {
            const uint64_t c = mse0[i][lev0[gi]] + mse1[i][lev1[gi]];
            if (c < best)
                best = c;
        }
        for (int j = start_gi; j < end_gi; j++)
            for (int k = start_gi; k < end_gi; k++) {
                const uint64_t c = mse0[i][j] + mse1[i][k];
                tot[j][k] += c < best ? c : best;
            }
    }
    for (int j = start_gi; j < end_gi; j++)
        for (int k = start_gi; k < end_gi; k++)
            if (tot[j][k] < best_tot) {
                best_tot = tot[j][k];
                b0       = j;
                b1       = k;
            }
    lev0[nb] = b0;
    lev1[nb] = b1;
    return best_tot;
}

/* finish_cdef_search — Encoder/Codec/EbEncCdef.c:728-925 (search path; 64x64 SBs).
 * mse: [2][nfb][64] (modified in place by zero_fs_cost_bias exactly like pcs->mse_seg). */
ORC_API int orc_cdef_pick(const svtf_cdef_search_cfg *cfg, uint64_t *mse, const uint8_t *skip, int nfb,
                          svtf_cdef_params *out, int8_t *fb_idx) {
    const int        end_gi = cfg->first_pass_fs_num + cfg->default_second_pass_fs_num;
    const uint64_t **m0 = (const uint64_t **)malloc(sizeof(void *) * (nfb + 1)), **m1 = (const uint64_t **)malloc(sizeof(void *) * (nfb + 1));
    int             *sb = (int *)malloc(sizeof(int) * (nfb + 1));
    int              sb_count = 0;
    for (int fb = 0; fb < nfb; fb++) {
        fb_idx[fb] = -1;
        if (skip[fb])
            continue;
        m0[sb_count] = mse + (size_t)fb * 64;
        m1[sb_count] = mse + ((size_t)nfb + fb) * 64;
        sb[sb_count] = fb;
        sb_count++;
    }
    if (cfg->zero_fs_cost_bias) {
        const uint16_t factor = (uint16_t)cfg->zero_fs_cost_bias;
        for (int i = 0; i < sb_count; i++) {
            ((uint64_t *)m0[i])[0] = (factor * m0[i][0]) >> 6;
            ((uint64_t *)m1[i])[0] = (factor * m1[i][0]) >> 6;
        }
    }
    memset(out, 0, sizeof(*out));
    uint64_t best_tot = (uint64_t)1 << 63;
    int      bits     = 0;
    for (int i = 0; i <= 3; i++) {
        int       l0[16] = {0}, l1[16] = {0};
        const int nb  = 1 << i;
        uint64_t  tot = (uint64_t)1 << 63;
        /* joint_strength_search_dual :697-727 */
        for (int k = 0; k < nb; k++) tot = search_one_dual(l0, l1, k, m0, m1, sb_count, 0, end_gi);
        for (int k = 0; k < 4 * nb; k++) {
            for (int j = 0; j < nb - 1; j++) {
                l0[j] = l0[j + 1];
                l1[j] = l1[j + 1];
            }
            tot = search_one_dual(l0, l1, nb - 1, m0, m1, sb_count, 0, end_gi);
        }
        /* av1_cost_literal(n) = n * (1<<AV1_PROB_COST_SHIFT)=n*512 ; RDCOST(RM,R,D) =
         * ROUND_POWER_OF_TWO((int64_t)R*RM, AV1_PROB_COST_SHIFT) + (D << RDDIV_BITS=7)
         * (Encoder/Codec/EbRateDistortionCost.h / EbDefinitions.h) */
        const int      total_bits = sb_count * i + nb * 6 * 2;
        const int64_t  rate       = (int64_t)total_bits * 512;
        const uint64_t dist       = tot * 16;
        tot                       = (uint64_t)(((rate * (int64_t)cfg->lambda + 256) >> 9) + (int64_t)(dist << 7));
        if (tot < best_tot) {
            best_tot = tot;
            bits     = i;
            for (int j = 0; j < nb; j++) {
                out->cdef_y_strength[j]  = (uint8_t)l0[j];
                out->cdef_uv_strength[j] = (uint8_t)l1[j];
            }
        }
    }
    out->cdef_bits = bits;
    const int nb   = 1 << bits;
    for (int i = 0; i < sb_count; i++) {
        uint64_t best = (uint64_t)1 << 63;
        int      bg   = 0;
        for (int gi = 0; gi < nb; gi++) {
            const uint64_t c = m0[i][out->cdef_y_strength[gi]] + m1[i][out->cdef_uv_strength[gi]];
            if (c < best) {
                bg   = gi;
                best = c;
            }
        }
        fb_idx[sb[i]] = (int8_t)bg;
    }
    int map[64] = {0};
    for (int i = 0; i < end_gi; i++)
        map[i] = i < cfg->first_pass_fs_num ? cfg->default_first_pass_fs[i] : cfg->default_second_pass_fs[i - cfg->first_pass_fs_num];
    for (int i = 0; i < nb; i++) {
        out->cdef_y_strength[i]  = (uint8_t)map[out->cdef_y_strength[i]];
        out->cdef_uv_strength[i] = (uint8_t)map[out->cdef_uv_strength[i]];
    }
    out->cdef_damping = 3 + (cfg->base_q_idx >> 6);
    free(m0);
    free(m1);
    free(sb);
    return 0;
}

/* ==========================================================================================
 * CCSO apply
 * ======================================================================================== */

/* ccso_frame — Common/Codec/EbCcso.c:626-678 with ccso_apply_{luma,chroma}_{mb,sb}_filter
 * (:297-622) and ccso_filter_block_hbd_wo_buf_c (:261-294) on 8- or 16-bit samples.
 * pre_y: luma BEFORE CDEF (what ext_rec_y holds, Encoder/Codec/EbCdefProcess.c:480-502); the
 * 5-sample replicate padding of extend_ccso_border (:185-201) is realised by clamping
 * coordinates to the unpadded picture, which yields the same samples (corners included). */
ORC_API int orc_ccso_apply_frame(const svtf_geom *g, void *y, int sy, void *cb, int scb, void *cr, int scr, const void *pre_y,
                                 int pre_sy, const svtf_ccso_params *cc, const uint8_t *const blk_flags[3]) {
    plane_t              pl[3]      = {{y, sy, g->sample_bytes}, {cb, scb, g->sample_bytes}, {cr, scr, g->sample_bytes}};
    const plane_t        lum        = {(void *)pre_y, pre_sy, g->sample_bytes};
    static const uint8_t quant_sz[4] = {16, 8, 32, 64};
    const int            W = g->width, H = g->height;
    const int            max_val = (1 << g->bit_depth) - 1;
    const int            nbx     = (W + 255) / 256;
    for (int p = 0; p < 3; p++) {
        if (!cc->ccso_enable[p])
            continue;
        const int ss = p ? 1 : 0, pw = W >> ss, ph = H >> ss;
        const int thr = quant_sz[cc->quant_idx[p]], neg_thr = -thr;
        const int blk_log2 = p ? 7 : 8;
        int       dy[2], dx[2];
        switch (cc->ext_filter_support[p]) { /* derive_ccso_sample_pos :204-234 */
        case 0: dy[0] = -1, dx[0] = 0, dy[1] = 1, dx[1] = 0; break;
        case 1: dy[0] = -1, dx[0] = -1, dy[1] = 1, dx[1] = 1; break;
        case 2: dy[0] = 0, dx[0] = -1, dy[1] = 0, dx[1] = 1; break;
        case 3: dy[0] = 1, dx[0] = -1, dy[1] = -1, dx[1] = 1; break;
        case 4: dy[0] = 0, dx[0] = -3, dy[1] = 0, dx[1] = 3; break;
        default: dy[0] = 0, dx[0] = -5, dy[1] = 0, dx[1] = 5; break;
        }
        const int single_band = cc->max_band_log2[p] == 0;
        const int shift_bits  = g->bit_depth - cc->max_band_log2[p];
        for (int yy = 0; yy < ph; yy++)
            for (int xx = 0; xx < pw; xx++) {
                if (blk_flags && blk_flags[p] && !blk_flags[p][(yy >> blk_log2) * nbx + (xx >> blk_log2)])
                    continue;
                const int ly = yy << ss, lx = xx << ss;
                const int c  = pget(&lum, lx, ly);
                int       cls[2] = {0, 0};
                if (!cc->ccso_bo_only[p])
                    for (int i = 0; i < 2; i++) {
                        const int d = pget(&lum, iclamp(lx + dx[i], 0, W - 1), iclamp(ly + dy[i], 0, H - 1)) - c;
                        if (cc->edge_clf[p] == 0)
                            cls[i] = d > thr ? 2 : (d < neg_thr ? 0 : 1);
                        else
                            cls[i] = d < neg_thr ? 0 : 1;
                    }
                const int band = single_band ? 0 : c >> shift_bits;
                const int off  = cc->filter_offset[p][(band << 4) + (cls[0] << 2) + cls[1]];
                pset(&pl[p], xx, yy, iclamp(off + pget(&pl[p], xx, yy), 0, max_val));
            }
    }
    return 0;
}

/* table probe for tests: the oracle's own level/threshold tables */
ORC_API void orc_lf_tables(const svtf_lf_params *lf, uint8_t *lvl768, uint8_t *lim, uint8_t *mblim, uint8_t *hev) {
    lf_tables t;
    orc_lf_build_tables(lf, &t);
    memcpy(lvl768, t.lvl, 768);
    memcpy(lim, t.lim, 64);
    memcpy(mblim, t.mblim, 64);
    memcpy(hev, t.hev, 64);
}
