// svtf_cdef5.cuh — CDEF apply (+ optional fused CCSO epilogue) for 8- and 10-bit pictures, fourth formulation: cdef3's packed
// fp16 taps on an instruction diet (profiles/r02c_*: cdef3 spent 37 % of its warp instructions on the taps, 16 % on staging,
// 20 % on per-thread set-up, 18 % on the CCSO epilogue and ran at 3 CTAs per SM).
//
// Same semantics as cdef3_kernel<T, true> (svtf_cdef3.cuh): svt_av1_cdef_frame (Encoder/Codec/EbEncCdef.c:284-612) +
// svt_cdef_filter_fb / svt_cdef_filter_block_c (Common/Codec/EbCdef.c:253-430) and, when `ccso.on`, ccso_frame
// (Common/Codec/EbCcso.c:626-678) on the CDEF result before it is written.  12-bit pictures stay on cdef3_kernel<T, false>.
//
// What changed against cdef3:
//  * ONE representation of a sample in shared memory: the fp16 value 1024 + n, whose bit pattern is 0x6400 | n (n < 1024).
//    Conversion in is one OR per pair and out one AND (cdef3: two instructions each way); differences p - x are still exact;
//    x + delta is exact up to 2047 and is clamped to max <= 2047 beyond; the integer copy of the luma tile (direction search,
//    CCSO classifier) is gone — both read the same words and mask (the direction search folds the offset into its -128) —
//    so a CTA needs 46 KB instead of 57 KB of shared memory and four CTAs fit an SM (registers <= 64);
//  * a tap is 6 instructions instead of 7: sign(d) * min(|d|, v) is one HMNMX2.XORSIGN (min.xorsign.abs.f16x2)
//        d = p - x;  u = fma(|d|, -2^-s, C);  r = u + 1024 (rounds to an integer);  v = relu(r - 1024);
//        c = xorsign_min(d, v);  acc = fma(c, w, acc)
//    (min(a, ceil(z)) = ceil(min(a, z)) for integer a, so the min may follow the rounding; the exactness argument is cdef3's);
//  * staging: a warp owns whole tile rows (3 rows x 10 chunks luma, 5 rows x 6 chunks chroma), so the sample that follows a
//    chunk comes from the next lane by shuffle instead of a second global load, and tiles that lie inside the coded picture
//    (91 % at 4K) skip every bounds test;
//  * per-thread set-up is read, not computed: the strength constants that are uniform over the filter block, the CCSO
//    parameters of each plane, the block flags and the twelve tap offsets per direction are written to shared memory by one
//    thread each before the first barrier; the per-block primary constants (adjusted strength, direction) by the thread that
//    picks the block's direction.
#pragma once
#include "svtf_cdef3.cuh"

namespace svtf {

constexpr uint32_t C5_OFF        = 0x64006400u; // 1024.0 per half
constexpr uint32_t C5_SENT       = 0xF800F800u; // -32768.0 per half: outside the coded plane
constexpr uint32_t C5_MASK       = 0x03FF03FFu;
constexpr int      C5_SMEM_WORDS = 2 * C3_YSZ + 4 * C3_CSZ;

// tap-class constants as they sit in shared memory: x = -2^-s, y = C (svtf_cdef3.cuh: t + 0.5 - 2^-(s+1)), z = first tap weight,
// w = second tap weight (low half) | on << 16 | luma direction << 20 | chroma direction << 24
__device__ __forceinline__ uint4 c5_tap_consts(const int t, const int damp, const uint32_t w0, const uint32_t w1) {
    const int   s = t ? max(0, damp - cdef_msb(t)) : 0;
    const float c = (float)t + (s ? 0.5f - __int_as_float((126 - s) << 23) : 0.f); // 2^-(s+1)
    return make_uint4((0x8000u | (uint32_t)((15 - s) << 10)) * 0x10001u, h2u(__float2half2_rn(c)), w0, (w1 & 0xffffu) | ((t != 0) << 16));
}

struct C5Tap {
    __half2 nsc, C;
    bool    on;
};
__device__ __forceinline__ C5Tap c5_tap(const uint4 k) {
    C5Tap t;
    t.nsc = u2h(k.x), t.C = u2h(k.y), t.on = (k.w >> 16) & 1;
    return t;
}

// NP pairs (2*NP samples) of one row.  ctr: centre words (even-aligned copy); o[k*3+set]: word offsets of the +tap (.x) and
// the -tap (.y) relative to ctr (they already select the even- or odd-aligned copy).  Returns integer sample pairs.
template <int NP>
__device__ __forceinline__ void cdef5_row(const uint32_t *ctr, const int2 (&o)[6], const C5Tap pr, const C5Tap se, const __half2 wp0,
                                          const __half2 wp1, uint32_t *outw) {
    const __half2 k1024 = u2h(C5_OFF), kn1024 = u2h(0xE400E400u), kone = u2h(0x3C003C00u);
    uint32_t      x[NP], mx[NP], mn[NP];
    __half2       acc[NP];
#pragma unroll
    for (int j = 0; j < NP; j++) x[j] = ctr[j], acc[j] = u2h(0u), mx[j] = x[j], mn[j] = x[j];
#pragma unroll
    for (int k = 0; k < 2; k++) {
#pragma unroll
        for (int set = 0; set < 3; set++) {
            const uint32_t *q0 = ctr + o[k * 3 + set].x, *q1 = ctr + o[k * 3 + set].y;
            const C5Tap     t  = set == 0 ? pr : se;
            const __half2   w  = set == 0 ? (k ? wp1 : wp0) : (k ? u2h(0x3C003C00u) : u2h(0x40004000u)); // cdef_sec_taps {2,1}
#pragma unroll
            for (int j = 0; j < NP; j++) {
                const uint32_t p0 = q0[j], p1 = q1[j];
                if (t.on) {
#pragma unroll
                    for (int e = 0; e < 2; e++) {
                        const __half2 d = __hsub2(u2h(e ? p1 : p0), u2h(x[j]));
                        const __half2 r = __hadd2(__hfma2(__habs2(d), t.nsc, t.C), k1024);
                        const __half2 v = __hfma2_relu(r, kone, kn1024);
                        acc[j]          = __hfma2(u2h(pk_xsmin(h2u(d), h2u(v))), w, acc[j]);
                    }
                }
                mx[j] = __vimax3_s16x2(mx[j], p0, p1); // the sentinel is negative as a signed halfword, 0x6400 | n is not
                mn[j] = __vimin3_u16x2(mn[j], p0, p1); // and it is the largest unsigned halfword
            }
        }
    }
    const __half2 k16b = u2h(0x2C012C01u), k1536 = u2h(0x66006600u), kn1536 = u2h(0xE600E600u);
#pragma unroll
    for (int j = 0; j < NP; j++) {
        // (8 + sum - (sum < 0)) >> 4 = sum/16 rounded to the nearest integer, halves away from zero = RN(sum * (1/16 + 2^-14) + 1536) - 1536:
        // the fma is exact before its single rounding to the integer grid of [1024, 2048); the 2^-14 term breaks halves away from zero and
        // is too small (|sum| <= 912) to carry anything else across one (svtf_search3.cuh, s3_err).  x + delta = (x - 1536) + t, exact.
        const __half2 t = __hfma2(acc[j], k16b, k1536);
        const __half2 y = __hmin2(__hmax2(__hadd2(__hadd2(u2h(x[j]), kn1536), t), u2h(mn[j])), u2h(mx[j]));
        outw[j]         = h2u(y) & C5_MASK;
    }
}

// Four directions of one 8x8 block from offset-representation words: cdef2_dir_costs4 with the 0x6400 folded into the -128
// (0x6400 is a multiple of 2^cs, so ((0x6400 | n) >> cs) = (0x6400 >> cs) + (n >> cs)).
template <int DBASE>
__device__ __forceinline__ void cdef5_dir_costs4(const uint32_t *blk, const int pw, const int cs, int *cost) {
    constexpr int NL0 = 15, NL1 = 11, NL2 = 8, NL3 = 11;
    int           p0[NL0], p1[NL1], p2[NL2], p3[NL3];
#pragma unroll
    for (int l = 0; l < NL0; l++) p0[l] = 0;
#pragma unroll
    for (int l = 0; l < NL1; l++) p1[l] = 0;
#pragma unroll
    for (int l = 0; l < NL2; l++) p2[l] = 0;
#pragma unroll
    for (int l = 0; l < NL3; l++) p3[l] = 0;
    const int bias = 128 + (0x6400 >> cs);
#pragma unroll
    for (int i = 0; i < 8; i++)
#pragma unroll
        for (int jp = 0; jp < 4; jp++) {
            const uint32_t w = blk[i * pw + jp];
#pragma unroll
            for (int h = 0; h < 2; h++) {
                const int j = 2 * jp + h;
                const int v = (int)(((h ? (w >> 16) : (w & 0xffff)) >> cs)) - bias;
                p0[cdef_line_of<DBASE>(i, j)] += v;
                p1[cdef_line_of<DBASE + 1>(i, j)] += v;
                p2[cdef_line_of<DBASE + 2>(i, j)] += v;
                p3[cdef_line_of<DBASE + 3>(i, j)] += v;
            }
        }
    int c0 = 0, c1 = 0, c2 = 0, c3 = 0;
#pragma unroll
    for (int l = 0; l < NL0; l++) c0 += p0[l] * p0[l] * (840 / (l < 8 ? l + 1 : 15 - l));
#pragma unroll
    for (int l = 0; l < NL1; l++) c1 += p1[l] * p1[l] * (840 / (2 * min(min(l + 1, 4), 11 - l)));
#pragma unroll
    for (int l = 0; l < NL2; l++) c2 += p2[l] * p2[l] * 105;
#pragma unroll
    for (int l = 0; l < NL3; l++) c3 += p3[l] * p3[l] * (840 / (2 * min(min(l + 1, 4), 11 - l)));
    cost[0] = c0, cost[1] = c1, cost[2] = c2, cost[3] = c3;
}

// CCSO parameters of one plane as they sit in shared memory (written once per CTA): Ccso2Pk without the pointer
struct C5Ccso {
    int      dy0, dx0, shift, pad;
    uint32_t b1, b2, cap1, cap2;
    uint32_t bmask, maxv2, pad2, pad3;
};
__device__ __forceinline__ Ccso2Pk c5_ccso_load(const C5Ccso *s, const uint8_t *lut) {
    const uint4 a = reinterpret_cast<const uint4 *>(s)[0], b = reinterpret_cast<const uint4 *>(s)[1];
    const uint2 c = reinterpret_cast<const uint2 *>(s)[4];
    Ccso2Pk     q;
    q.dy0 = (int)a.x, q.dx0 = (int)a.y, q.shift = (int)a.z;
    q.b1 = b.x, q.b2 = b.y, q.cap1 = b.z, q.cap2 = b.w;
    q.bmask = c.x, q.maxv2 = c.y, q.lut = lut;
    return q;
}
// scalar CCSO for tiles at the left / right picture edge (ccso2_px on offset-representation samples)
__device__ __forceinline__ int c5_ccso_px(const uint16_t *lum16, const int pitch16, const int lr, const int lc, const Ccso2Pk &q, const Ccso2Tile &t,
                                          const int v) {
    const int y  = lum16[lr * pitch16 + lc] & 0x3FF;
    const int n0 = lum16[min(max(lr + q.dy0, t.rmin), t.rmax) * pitch16 + min(max(lc + q.dx0, t.cmin), t.cmax)] & 0x3FF;
    const int n1 = lum16[min(max(lr - q.dy0, t.rmin), t.rmax) * pitch16 + min(max(lc - q.dx0, t.cmin), t.cmax)] & 0x3FF;
    const int d0 = n0 - y, d1 = n1 - y, thr = (int)(q.b1 & 0xffff) - 1;
    const int c0 = (q.cap1 ? (d0 >= -thr) : 0) + (q.cap2 ? (d0 > thr) : 0);
    const int c1 = (q.cap1 ? (d1 >= -thr) : 0) + (q.cap2 ? (d1 > thr) : 0);
    const int band = y >> q.shift;
    return min(max(v + (int)q.lut[(band << 4) + (c0 << 2) + c1] - 128, 0), (int)(q.maxv2 & 0xffff));
}
// NP pairs starting at tile column `col` (any parity) of tile row `row`: the odd-aligned copy serves odd columns
template <int NP>
__device__ __forceinline__ void c5_ld_pairs(const uint32_t *A, const uint32_t *B, const int pw, const int row, const int col, uint32_t *out) {
    const int       odd = col & 1;
    const uint32_t *b   = (odd ? B : A) + row * pw + ((col - odd) >> 1);
#pragma unroll
    for (int j = 0; j < NP; j++) out[j] = b[j];
}

// one 8-sample chunk of a tile row -> even copy A, odd copy B (word i = samples 2i+1, 2i+2); wn = the word that follows
__device__ __forceinline__ void c5_put(uint32_t *A, uint32_t *B, const int o, const uint32_t w0, const uint32_t w1, const uint32_t w2,
                                       const uint32_t w3, const uint32_t wn) {
    A[o] = w0, A[o + 1] = w1, A[o + 2] = w2, A[o + 3] = w3;
    B[o]     = __byte_perm(w0, w1, 0x5432);
    B[o + 1] = __byte_perm(w1, w2, 0x5432);
    B[o + 2] = __byte_perm(w2, w3, 0x5432);
    B[o + 3] = __byte_perm(w3, wn, 0x5432);
}

template <typename T>
__global__ void __launch_bounds__(C2_THREADS, 4) cdef5_kernel(const __grid_constant__ Cdef2Args a) {
    extern __shared__ __align__(16) uint32_t c5_sm[];
    uint32_t *const yA  = c5_sm;              // yB = yA + C3_YSZ
    uint32_t *const cA0 = c5_sm + 2 * C3_YSZ; // plane p: cA0 + p * 2 * C3_CSZ, its B copy + C3_CSZ
    __shared__ int     s_cost[64][9];         // [block][dir], padded
    __shared__ __align__(16) int2 s_off[2][8][6]; // [luma/chroma][dir][k*3+set]: word offsets of the +tap / -tap
    __shared__ __align__(16) uint4  s_blk[64];    // per 8x8 block: primary luma constants + directions (c5_tap_consts)
    __shared__ __align__(16) uint4  s_uni[3];     // filter-block uniform: luma secondary, chroma primary, chroma secondary
    __shared__ __align__(16) C5Ccso s_cq[3];
    __shared__ int     s_flg[3];
    __shared__ uint8_t s_active[64];
    __shared__ __align__(16) uint8_t s_lut[3][SVTF_CCSO_LUT_SIZE]; // CCSO offsets + 128

    const int tid = threadIdx.x, lane = tid & 31, wrp = tid >> 5;
    const int fbc = blockIdx.x, fbr = blockIdx.y + a.fbr0;
    const int nhfb = gridDim.x;
    const int cw = a.d.mi_cols * 4, ch = a.d.mi_rows * 4;
    const int x0 = fbc * 64, y0 = fbr * 64;
    const int cs = a.d.bit_depth - 8;
    const int W = a.d.width, H = a.d.height;

    const T *iny  = static_cast<const T *>(a.in.p[0]);
    T       *outy = static_cast<T *>(a.out.p[0]);

    // ---- every global load of this thread, issued before the first use ----
    const int idx = __ldg(a.fb_idx + fbr * nhfb + fbc);
    svtf_mi_t m4[4] = {~0u, ~0u, ~0u, ~0u};
    bool      mi_in = false;
    int       flg_p = 0;
    if (tid < 64) { // svt_sb_compute_cdef_list — Encoder/Codec/EbEncCdef.c:238-281
        const int r = (tid >> 3) * 2, c = (tid & 7) * 2;
        mi_in = fbr * 16 + r < a.d.mi_rows && fbc * 16 + c < a.d.mi_cols;
        if (mi_in) {
            const svtf_mi_t *m = a.mi + (fbr * 16 + r) * a.mi_stride + fbc * 16 + c;
            m4[0] = __ldg(m), m4[1] = __ldg(m + 1), m4[2] = __ldg(m + a.mi_stride), m4[3] = __ldg(m + a.mi_stride + 1);
        }
    } else if (tid >= 96 && tid < 99 && a.ccso.on) { // CCSO block flag of plane p (one flag per 256x256 luma samples)
        const int p = tid - 96;
        flg_p       = a.ccso.enable[p];
        if (flg_p && a.ccso.blk_flags[p]) {
            const int pwid = W >> (p ? 1 : 0), bl = p ? 7 : 8;
            const int nbx  = (pwid + (1 << bl) - 1) >> bl;
            flg_p          = (x0 >> 8) < nbx && __ldg(a.ccso.blk_flags[p] + ((y0 + a.ccso.y_glob) >> 8) * nbx + (x0 >> 8)) != 0;
        }
    }
    // tiles: a warp owns whole rows, lane = (row in the group, chunk); lanes 30, 31 idle
    const bool interior = x0 >= 16 && x0 + 80 <= cw && y0 >= 4 && y0 + 68 <= ch; // luma 80 x 68 at (x0-8, y0-2) and chroma 48 x 36 at (x0/2-8, y0/2-2) inside the coded planes
    const int  lr3 = lane / 10, lc10 = lane - lr3 * 10;                          // luma: 3 rows x 10 chunks
    const int  lr5 = lane / 6, lc6 = lane - lr5 * 6;                             // chroma: 5 rows x 6 chunks
    uint4      vy[3], vc[2];
    uint32_t   ny[3], nc[2]; // edge tiles only: the sample after the chunk
    int        my[3], mc[2]; // validity (c3_load)
#pragma unroll
    for (int k = 0; k < 3; k++) {
        const int r = 3 * (k * 8 + wrp) + lr3;
        vy[k] = make_uint4(0, 0, 0, 0), ny[k] = 0, my[k] = 0;
        if (lane < 30 && r < C2_YROWS) {
            if (interior)
                vy[k] = ld8<T>(iny + (size_t)(y0 - 2 + r) * a.in.pitch[0] + x0 - 8 + lc10 * 8);
            else {
                const C3Chunk c = c3_load<T>(iny, a.in.pitch[0], x0 - 8 + lc10 * 8, y0 - 2 + r, cw, ch, true);
                vy[k] = make_uint4(c.w0, c.w1, c.w2, c.w3), ny[k] = c.wn, my[k] = c.m;
            }
        }
    }
    const int rc5 = 5 * wrp + lr5;
#pragma unroll
    for (int p = 0; p < 2; p++) {
        vc[p] = make_uint4(0, 0, 0, 0), nc[p] = 0, mc[p] = 0;
        if (lane < 30 && rc5 < C2_CROWS) {
            const T *inc = static_cast<const T *>(a.in.p[1 + p]);
            if (interior)
                vc[p] = ld8<T>(inc + (size_t)(y0 / 2 - 2 + rc5) * a.in.pitch[1 + p] + x0 / 2 - 8 + lc6 * 8);
            else {
                const C3Chunk c = c3_load<T>(inc, a.in.pitch[1 + p], x0 / 2 - 8 + lc6 * 8, y0 / 2 - 2 + rc5, cw / 2, ch / 2, true);
                vc[p] = make_uint4(c.w0, c.w1, c.w2, c.w3), nc[p] = c.wn, mc[p] = c.m;
            }
        }
    }
    // CCSO LUTs (3 x 2048 bytes = 384 x 16 bytes) of the planes that can run
    uint4      l0 = make_uint4(0, 0, 0, 0), l1 = l0;
    const bool lut0 = a.ccso.on && a.ccso.enable[tid >> 7], lut1 = a.ccso.on && a.ccso.enable[2] && tid < 128;
    if (lut0)
        l0 = __ldg(reinterpret_cast<const uint4 *>(a.ccso.lut) + tid);
    if (lut1)
        l1 = __ldg(reinterpret_cast<const uint4 *>(a.ccso.lut) + C2_THREADS + tid);

    int ylev = 0, ysec = 0, uvlev = 0, uvsec = 0;
    if (idx >= 0) {
        const int ys = a.y_strength[idx], us = a.uv_strength[idx];
        ylev = ys >> 2, ysec = ys & 3, ysec += ysec == 3;
        uvlev = us >> 2, uvsec = us & 3, uvsec += uvsec == 3;
    }
    int act = 0;
    if (tid < 64) {
        act           = mi_in && !(mi_skip(m4[0]) & mi_skip(m4[1]) & mi_skip(m4[2]) & mi_skip(m4[3]));
        s_active[tid] = (uint8_t)act;
    } else if (tid < 96) {
        // Cdef_Directions (Common/Codec/EbCdef.c:99-122) as nibble tables (value + 2) indexed by direction:
        // k=0: dy {-1,0,0,0,1,1,1,1} dx {1,1,1,1,1,0,0,0};  k=1: dy {-2,-1,0,1,2,2,2,2} dx {2,2,2,2,2,1,0,-1}
        // entry [pt][dir][k*3+set]: set 0 = dir, 1 = dir + 2, 2 = dir + 6 (the secondary directions)
        const int t = tid - 64, pt = t >> 4, dir = (t >> 1) & 7, k = t & 1;
        const int pw = pt ? C2_CPW : C2_YPW, bd = pt ? C3_CSZ : C3_YSZ;
#pragma unroll
        for (int set = 0; set < 3; set++) {
            const int dd  = (dir + (set == 0 ? 0 : (set == 1 ? 2 : 6))) & 7;
            const int dy  = (int)(((k ? 0x44443210u : 0x33332221u) >> (4 * dd)) & 15) - 2;
            const int dx  = (int)(((k ? 0x12344444u : 0x22233333u) >> (4 * dd)) & 15) - 2;
            const int odd = dx & 1; // centre columns are even
            s_off[pt][dir][k * 3 + set] = make_int2((odd ? bd : 0) + dy * pw + ((dx - odd) >> 1), (odd ? bd : 0) - dy * pw + ((-dx - odd) >> 1));
        }
    } else if (tid < 99) {
        s_flg[tid - 96] = flg_p;
    } else if (tid < 102) {
        const int p = tid - 99;
        if (a.ccso.on) {
            const Ccso2Pk q = ccso2_pk(a.ccso, p, a.d.bit_depth, nullptr);
            C5Ccso        c;
            c.dy0 = q.dy0, c.dx0 = q.dx0, c.shift = q.shift, c.pad = 0;
            c.b1 = q.b1, c.b2 = q.b2, c.cap1 = q.cap1, c.cap2 = q.cap2;
            c.bmask = (((1u << a.d.bit_depth) - 1) >> q.shift) * 0x10001u; // the band of 0x6400 | n is the band of n
            c.maxv2 = q.maxv2, c.pad2 = c.pad3 = 0;
            s_cq[p] = c;
        }
    } else if (tid < 105) {
        const int k = tid - 102; // 0: luma secondary, 1: chroma primary, 2: chroma secondary
        if (k == 0)
            s_uni[0] = c5_tap_consts(ysec << cs, a.damping + cs, 0, 0);
        else if (k == 1) {
            const int odd = uvlev & 1; // cdef_pri_taps {4,2} / {3,3}
            s_uni[1]      = c5_tap_consts(uvlev << cs, a.damping + cs - 1, odd ? 0x42004200u : 0x44004400u, odd ? 0x4200u : 0x4000u);
        } else
            s_uni[2] = c5_tap_consts(uvsec << cs, a.damping + cs - 1, 0, 0);
    }
    const int  any_active = __syncthreads_or(act);
    const bool process    = (ylev | ysec | uvlev | uvsec) && any_active;
    const int  q_on       = a.ccso.on ? (s_flg[0] | s_flg[1] << 1 | s_flg[2] << 2) : 0; // bit p: CCSO (ccso_frame, Common/Codec/EbCcso.c:626) runs on plane p of this tile

    if (!process && !q_on) { // untouched filter block: carry the interior chunks over to the output surface
#pragma unroll
        for (int k = 0; k < 3; k++) {
            const int r = 3 * (k * 8 + wrp) + lr3;
            if (lane < 30 && r >= 2 && r < 66 && lc10 >= 1 && lc10 < 9 && (interior || (my[k] & 3) == 3))
                st8<T>(outy + (size_t)(y0 - 2 + r) * a.out.pitch[0] + x0 - 8 + lc10 * 8, vy[k]);
        }
#pragma unroll
        for (int p = 0; p < 2; p++)
            if (lane < 30 && rc5 >= 2 && rc5 < 34 && lc6 >= 1 && lc6 < 5 && (interior || (mc[p] & 1))) {
                T *o = static_cast<T *>(a.out.p[1 + p]) + (size_t)(y0 / 2 - 2 + rc5) * a.out.pitch[1 + p] + x0 / 2 - 8 + lc6 * 8;
                if (interior || (mc[p] & 3) == 3)
                    st8<T>(o, vc[p]);
                else
                    st4<T>(o, make_uint2(vc[p].x, vc[p].y));
            }
        return;
    }

    // ---- tiles -> shared memory (0x6400 | n; sentinel outside the coded plane) ----
#pragma unroll
    for (int k = 0; k < 3; k++) {
        const int r = 3 * (k * 8 + wrp) + lr3;
        uint32_t  w0 = vy[k].x | C5_OFF, w1 = vy[k].y | C5_OFF, w2 = vy[k].z | C5_OFF, w3 = vy[k].w | C5_OFF, wn;
        if (interior)
            wn = __shfl_down_sync(0xffffffffu, w0, 1); // chunk 9's successor is never read
        else {
            const bool v0 = my[k] & 1, v1 = my[k] & 2;
            w0 = v0 ? w0 : C5_SENT, w1 = v0 ? w1 : C5_SENT, w2 = v1 ? w2 : C5_SENT, w3 = v1 ? w3 : C5_SENT;
            wn = (my[k] & 4) ? (ny[k] | C5_OFF) : C5_SENT;
        }
        if (lane < 30 && r < C2_YROWS)
            c5_put(yA, yA + C3_YSZ, r * C2_YPW + lc10 * 4, w0, w1, w2, w3, wn);
    }
#pragma unroll
    for (int p = 0; p < 2; p++) {
        uint32_t w0 = vc[p].x | C5_OFF, w1 = vc[p].y | C5_OFF, w2 = vc[p].z | C5_OFF, w3 = vc[p].w | C5_OFF, wn;
        if (interior)
            wn = __shfl_down_sync(0xffffffffu, w0, 1);
        else {
            const bool v0 = mc[p] & 1, v1 = mc[p] & 2;
            w0 = v0 ? w0 : C5_SENT, w1 = v0 ? w1 : C5_SENT, w2 = v1 ? w2 : C5_SENT, w3 = v1 ? w3 : C5_SENT;
            wn = (mc[p] & 4) ? (nc[p] | C5_OFF) : C5_SENT;
        }
        if (lane < 30 && rc5 < C2_CROWS)
            c5_put(cA0 + p * 2 * C3_CSZ, cA0 + p * 2 * C3_CSZ + C3_CSZ, rc5 * C2_CPW + lc6 * 4, w0, w1, w2, w3, wn);
    }
    if (lut0) // biased by +128
        reinterpret_cast<uint4 *>(&s_lut[0][0])[tid] = make_uint4(l0.x ^ 0x80808080u, l0.y ^ 0x80808080u, l0.z ^ 0x80808080u, l0.w ^ 0x80808080u);
    if (lut1)
        reinterpret_cast<uint4 *>(&s_lut[0][0])[C2_THREADS + tid] =
            make_uint4(l1.x ^ 0x80808080u, l1.y ^ 0x80808080u, l1.z ^ 0x80808080u, l1.w ^ 0x80808080u);
    __syncthreads();
    // ---- direction search: thread = (block, half of the directions) ----
    if (tid < 128 && process) {
        const int b = tid & 63, half = tid >> 6;
        if (s_active[b]) {
            const uint32_t *blk = yA + (2 + (b >> 3) * 8) * C2_YPW + 4 + (b & 7) * 4;
            int             c4[4];
            if (half == 0)
                cdef5_dir_costs4<0>(blk, C2_YPW, cs, c4);
            else
                cdef5_dir_costs4<4>(blk, C2_YPW, cs, c4);
#pragma unroll
            for (int i = 0; i < 4; i++) s_cost[b][half * 4 + i] = c4[i];
        }
    }
    __syncthreads();
    if (tid < 64 && process && s_active[tid]) {
        int best = 0, bd = 0;
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const int c = s_cost[tid][i];
            if (c > best) { // first maximum wins ties (Common/Codec/EbCdef.c:197-201)
                best = c;
                bd   = i;
            }
        }
        const int var   = (best - s_cost[tid][(bd + 4) & 7]) >> 10;
        const int pri_l = ylev << cs;
        // adjust_strength — Common/Codec/EbCdef.c:130-134
        const int ii  = (var >> 6) ? min(cdef_msb(var >> 6), 12) : 0;
        const int pri = var ? (pri_l * (4 + ii) + 8) >> 4 : 0;
        const int odd = (pri >> cs) & 1; // cdef_pri_taps {4,2} / {3,3}
        uint4     k   = c5_tap_consts(pri, a.damping + cs, odd ? 0x42004200u : 0x44004400u, odd ? 0x4200u : 0x4000u);
        k.w |= (uint32_t)(pri_l ? bd : 0) << 20 | (uint32_t)(uvlev ? bd : 0) << 24;
        s_blk[tid] = k;
    }
    __syncthreads();

    const uint16_t *lum16   = reinterpret_cast<const uint16_t *>(yA);
    const bool      edge_lr = x0 - 5 < 0 || x0 + 69 > W; // CCSO tap columns can leave the picture: per-sample clamping
    Ccso2Tile       ct;
    ct.rmin = 2 - y0, ct.rmax = H - 1 - (y0 - 2), ct.cmin = 8 - x0, ct.cmax = W - 1 - (x0 - 8);
    const int bx = tid & 7;

    // ---- luma: thread = rows r and r+4 of one 8x8 block; a warp = one row of blocks ----
    {
        const int  rp = (tid >> 3) & 3, by = tid >> 5, b = by * 8 + bx;
        const int  gx = x0 + bx * 8;
        const bool inside = gx < cw && y0 + by * 8 < ch; // coded sizes are multiples of 8
        const bool on     = process && (ylev | ysec) && inside && s_active[b];
        int2       o[6];
        C5Tap      pri, sec;
        __half2    w0 = u2h(0u), w1 = u2h(0u);
        pri.on = sec.on = false;
        if (on) {
            const uint4 kb = s_blk[b], ks = s_uni[0];
            pri = c5_tap(kb), sec = c5_tap(ks);
            w0 = u2h(kb.z), w1 = u2h((kb.w & 0xffffu) * 0x10001u);
            const int4 *po = reinterpret_cast<const int4 *>(&s_off[0][(kb.w >> 20) & 7][0]);
            const int4  o0 = po[0], o1 = po[1], o2 = po[2];
            o[0] = make_int2(o0.x, o0.y), o[1] = make_int2(o0.z, o0.w), o[2] = make_int2(o1.x, o1.y);
            o[3] = make_int2(o1.z, o1.w), o[4] = make_int2(o2.x, o2.y), o[5] = make_int2(o2.z, o2.w);
        }
        const bool    cc_on = (q_on & 1) != 0;
        const Ccso2Pk qp    = c5_ccso_load(&s_cq[0], s_lut[0]);
        if (inside) {
#pragma unroll 1
            for (int h = 0; h < 2; h++) {
                const int row = rp + 4 * h, gy = y0 + by * 8 + row;
                const int trow = 2 + by * 8 + row, tcol = 8 + bx * 8;
                uint32_t  o32[4];
                const uint32_t *ctr = yA + trow * C2_YPW + (tcol >> 1);
                if (on && (pri.on || sec.on)) {
                    // two halves of the 8-sample row through one copy of the tap code (measured: same time alone as <4>, 2 % faster
                    // with a second picture's kernels co-resident; 7 KB less code against the 32 KB instruction cache)
#pragma unroll 1
                    for (int hh = 0; hh < 2; hh++) cdef5_row<2>(ctr + 2 * hh, o, pri, sec, w0, w1, o32 + 2 * hh);
                } else {
#pragma unroll
                    for (int j = 0; j < 4; j++) o32[j] = ctr[j] & C5_MASK;
                }
                if (cc_on && gy < H) {
                    if (edge_lr) {
#pragma unroll 1
                        for (int j = 0; j < 8; j++)
                            if (gx + j < W) {
                                const int v = (o32[j >> 1] >> (16 * (j & 1))) & 0xffff;
                                const int r = c5_ccso_px(lum16, 2 * C2_YPW, trow, tcol + j, qp, ct, v);
                                o32[j >> 1] = (j & 1) ? ((o32[j >> 1] & 0xffffu) | ((uint32_t)r << 16)) : ((o32[j >> 1] & 0xffff0000u) | (uint32_t)r);
                            }
                    } else {
                        uint32_t ta[4], tb[4];
                        c5_ld_pairs<4>(yA, yA + C3_YSZ, C2_YPW, min(max(trow + qp.dy0, ct.rmin), ct.rmax), tcol + qp.dx0, ta);
                        c5_ld_pairs<4>(yA, yA + C3_YSZ, C2_YPW, min(max(trow - qp.dy0, ct.rmin), ct.rmax), tcol - qp.dx0, tb);
#pragma unroll
                        for (int j = 0; j < 4; j++) o32[j] = ccso2_pair(ctr[j], ta[j], tb[j], o32[j], qp);
                    }
                }
                st8<T>(outy + (size_t)gy * a.out.pitch[0] + gx, make_uint4(o32[0], o32[1], o32[2], o32[3]));
            }
        }
    }
    // ---- chroma: thread = rows r and r+2 of one 4x4 block of one plane ----
    {
        const int       p = tid >> 7, by = (tid >> 4) & 7, rp = (tid >> 3) & 1, b = by * 8 + bx;
        const int       gx = x0 / 2 + bx * 4;
        const bool      inside = gx < cw / 2 && y0 / 2 + by * 4 < ch / 2;
        const bool      on     = process && (uvlev | uvsec) && inside && s_active[b];
        const uint32_t *cA     = cA0 + p * 2 * C3_CSZ;
        int2            o[6];
        C5Tap           pri, sec;
        __half2         w0 = u2h(0u), w1 = u2h(0u);
        pri.on = sec.on = false;
        if (on) {
            const uint4 kp = s_uni[1], ks = s_uni[2];
            pri = c5_tap(kp), sec = c5_tap(ks);
            w0 = u2h(kp.z), w1 = u2h((kp.w & 0xffffu) * 0x10001u);
            const int4 *po = reinterpret_cast<const int4 *>(&s_off[1][(s_blk[b].w >> 24) & 7][0]);
            const int4  o0 = po[0], o1 = po[1], o2 = po[2];
            o[0] = make_int2(o0.x, o0.y), o[1] = make_int2(o0.z, o0.w), o[2] = make_int2(o1.x, o1.y);
            o[3] = make_int2(o1.z, o1.w), o[4] = make_int2(o2.x, o2.y), o[5] = make_int2(o2.z, o2.w);
        }
        const bool    cc_on = ((q_on >> (1 + p)) & 1) != 0;
        const Ccso2Pk qp    = c5_ccso_load(&s_cq[1 + p], s_lut[1 + p]);
        T            *outc  = static_cast<T *>(a.out.p[1 + p]);
        if (inside) {
#pragma unroll 1
            for (int h = 0; h < 2; h++) {
                const int row = rp + 2 * h, gy = y0 / 2 + by * 4 + row;
                const int trow = 2 + by * 4 + row, tcol = 8 + bx * 4;
                uint32_t  o32[2];
                const uint32_t *ctr = cA + trow * C2_CPW + (tcol >> 1);
                if (on)
                    cdef5_row<2>(ctr, o, pri, sec, w0, w1, o32);
                else
                    o32[0] = ctr[0] & C5_MASK, o32[1] = ctr[1] & C5_MASK;
                if (cc_on && gy < (H >> 1)) {
                    const int lr = 2 + 2 * (by * 4 + row), lc = 8 + 8 * bx; // co-located luma sample of chroma sample 0
                    if (edge_lr) {
#pragma unroll 1
                        for (int j = 0; j < 4; j++)
                            if (gx + j < (W >> 1)) {
                                const int v = (o32[j >> 1] >> (16 * (j & 1))) & 0xffff;
                                const int r = c5_ccso_px(lum16, 2 * C2_YPW, lr, lc + 2 * j, qp, ct, v);
                                o32[j >> 1] = (j & 1) ? ((o32[j >> 1] & 0xffffu) | ((uint32_t)r << 16)) : ((o32[j >> 1] & 0xffff0000u) | (uint32_t)r);
                            }
                    } else {
                        const int ra = min(max(lr + qp.dy0, ct.rmin), ct.rmax), rb = min(max(lr - qp.dy0, ct.rmin), ct.rmax);
#pragma unroll
                        for (int k = 0; k < 2; k++)
                            o32[k] = ccso2_pair(c3_ld_skip(lum16, 2 * C2_YPW, lr, lc + 4 * k), c3_ld_skip(lum16, 2 * C2_YPW, ra, lc + 4 * k + qp.dx0),
                                                c3_ld_skip(lum16, 2 * C2_YPW, rb, lc + 4 * k - qp.dx0), o32[k], qp);
                    }
                }
                st4<T>(outc + (size_t)gy * a.out.pitch[1 + p] + gx, make_uint2(o32[0], o32[1]));
            }
        }
    }
}

} // namespace svtf
