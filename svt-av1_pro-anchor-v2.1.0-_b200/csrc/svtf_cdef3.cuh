// svtf_cdef3.cuh — CDEF apply (+ optional fused CCSO epilogue), third formulation.
//
// Same semantics as cdef2_kernel (svtf_cdef2.cuh): svt_av1_cdef_frame (Encoder/Codec/EbEncCdef.c:284-612) +
// svt_cdef_filter_fb / svt_cdef_filter_block_c (Common/Codec/EbCdef.c:253-430) and, when `ccso.on`, ccso_frame
// (Common/Codec/EbCcso.c:626-678) on the CDEF result before it is written.
//
// What changed against cdef2 (profiles/r01d_ncu_summary.md: the taps were only 26 % of cdef2's instructions):
//  * the odd-aligned tile copy is produced while staging (no second pass over shared memory, one barrier less);
//  * the out-of-frame sentinel is chosen so that no tap needs masking: unsigned min ignores it (it is the largest
//    unsigned halfword), SIGNED max ignores it (it is negative), constrain() of it is 0 — the BORDER code variant
//    is gone and the kernel is half the size (cdef2 stalled 1.2 warps per issue on instruction fetch);
//  * a thread owns two rows (r, r+4) of one 8x8 block (chroma: r, r+2 of a 4x4): strength, direction and the
//    twelve tap offsets (one 64-bit shared load per tap pair from a per-CTA table) are set up once;
//  * the rounding / clamp epilogue stays packed;
//  * 8- and 10-bit pictures run the taps in packed fp16 (FP = true).  Samples (< 1024), differences and the
//    weighted sum (|sum| <= 912 < 2048) are exact in fp16, and with a = |p - x|
//        constrain = sign(d) * min(a, max(0, t - (a >> s)))  =  sign(d) * max(0, RN(min(a, t - a*2^-s + 0.5 - 2^-(s+1))))
//    (a integer, so min(a, ceil(z)) = ceil(min(a, z)); ceil(z) = RN(z + 0.5 - 2^-(s+1)) because z is a multiple of
//    2^-s; wherever the result is positive the argument needs msb(t)+1 integer and s+1 fraction bits =
//    damping+2 <= 11 bits, so the single fma rounding is exact there and cannot change the sign elsewhere).
//    One tap on a sample pair = HADD2, HFMA2, HMNMX2, HADD2 (+1024: round to integer), HFMA2.RELU (-1024), LOP3
//    (sign), HFMA2 (accumulate): 7 instructions, 5 of them on the FMA pipe, against 12 (7 on the ALU pipe).
//    12-bit pictures (samples up to 4095 are not exact in fp16) keep the packed-integer taps (FP = false).
#pragma once
#include <cuda_fp16.h>
#include "svtf_cdef2.cuh"

namespace svtf {

constexpr int      C3_YSZ    = C2_YROWS * C2_YPW; // words per luma tile copy
constexpr int      C3_CSZ    = C2_CROWS * C2_CPW; // words per chroma tile copy
constexpr uint32_t C3_SENT_I = 0xFFFFFFFFu;       // integer tiles: 0xFFFF per half
constexpr uint32_t C3_SENT_F = 0xF800F800u;       // fp16 tiles: -32768.0 per half
// dynamic shared memory, in words: luma A|B, chroma (A|B) x 2, and for FP the integer luma copy
constexpr int C3_SMEM_WORDS_I = 2 * C3_YSZ + 4 * C3_CSZ;
constexpr int C3_SMEM_WORDS_F = 3 * C3_YSZ + 4 * C3_CSZ;

__device__ __forceinline__ uint32_t h2u(const __half2 h) { return *reinterpret_cast<const uint32_t *>(&h); }
__device__ __forceinline__ __half2  u2h(const uint32_t u) { return *reinterpret_cast<const __half2 *>(&u); }
// pair of integers < 1024 -> the same values as fp16 (0x6400 | n is 1024 + n), and back
__device__ __forceinline__ uint32_t pk_i2h(const uint32_t w) { return h2u(__hsub2(u2h(w | 0x64006400u), u2h(0x64006400u))); }
__device__ __forceinline__ uint32_t pk_h2i(const uint32_t f) { return h2u(__hadd2(u2h(f), u2h(0x64006400u))) & 0x03FF03FFu; }

// sign(a) ^ sign(b) applied to min(|a|, |b|), per half (HMNMX2.XORSIGN)
__device__ __forceinline__ uint32_t pk_xsmin(const uint32_t a, const uint32_t b) {
    uint32_t d;
    asm("min.xorsign.abs.f16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
    return d;
}

// Staging is split in two so that every global load of a thread (tile chunks, CCSO LUT, side info) is in flight before
// the first one is consumed: c3_load fetches one 8-sample chunk + the sample that follows it into registers, c3_store
// writes it as A (even pair alignment), B (odd: word i = samples 2i+1, 2i+2) and, FP only, I (the integer pairs; A/B
// then hold fp16 values).  Tile origin = (x0-8, y0-2); outside the coded plane -> sentinel.
struct C3Chunk {
    uint32_t w0, w1, w2, w3, wn; // raw integer pairs; wn.lo = first sample of the next chunk
    int      m;                  // bit 0: w0,w1 valid; bit 1: w2,w3 valid; bit 2: wn valid
};
template <typename T>
__device__ __forceinline__ C3Chunk c3_load(const T *in, const int ipitch, const int gx, const int gy, const int cw, const int ch,
                                           const bool item) {
    C3Chunk k;
    k.w0 = k.w1 = k.w2 = k.w3 = k.wn = 0, k.m = 0;
    if (item && gy >= 0 && gy < ch) {
        const T *row = in + (size_t)gy * ipitch;
        if (gx >= 0 && gx < cw) { // cw is a multiple of 8 (luma) / 4 (chroma)
            if (gx + 8 <= cw) {
                const uint4 v = ld8<T>(row + gx);
                k.w0 = v.x, k.w1 = v.y, k.w2 = v.z, k.w3 = v.w, k.m = 3;
            } else { // chroma planes whose width is 4 mod 8: last half chunk
                const uint2 h = ld4<T>(row + gx);
                k.w0 = h.x, k.w1 = h.y, k.m = 1;
            }
        }
        const int gn = gx + 8;
        if (gn >= 0 && gn < cw)
            k.wn = (uint32_t)__ldg(row + gn), k.m |= 4;
    }
    return k;
}
template <bool FP>
__device__ __forceinline__ void c3_store(uint32_t *A, uint32_t *B, uint32_t *I, const int o, const C3Chunk &k) {
    const uint32_t S  = FP ? C3_SENT_F : C3_SENT_I;
    const bool     v0 = k.m & 1, v1 = k.m & 2, vn = k.m & 4;
    const uint32_t w0 = v0 ? (FP ? pk_i2h(k.w0) : k.w0) : S, w1 = v0 ? (FP ? pk_i2h(k.w1) : k.w1) : S;
    const uint32_t w2 = v1 ? (FP ? pk_i2h(k.w2) : k.w2) : S, w3 = v1 ? (FP ? pk_i2h(k.w3) : k.w3) : S;
    const uint32_t wn = vn ? (FP ? pk_i2h(k.wn) : k.wn) : S;
    A[o] = w0, A[o + 1] = w1, A[o + 2] = w2, A[o + 3] = w3;
    B[o]     = __byte_perm(w0, w1, 0x5432);
    B[o + 1] = __byte_perm(w1, w2, 0x5432);
    B[o + 2] = __byte_perm(w2, w3, 0x5432);
    B[o + 3] = __byte_perm(w3, wn, 0x5432);
    if (FP && I)
        I[o] = k.w0, I[o + 1] = k.w1, I[o + 2] = k.w2, I[o + 3] = k.w3;
}

// Per-thread strength constants of one tap class (primary or secondary).
struct Tap3I {
    uint32_t t2, mask;
    int      s;
};
struct Tap3F {
    __half2 nsc, C; // -2^-s ; t + 0.5 - 2^-(s+1)  (t when s == 0)
    bool    on;
};
__device__ __forceinline__ Tap3I tap3i(const int t, const int damp) {
    Tap3I q;
    q.t2   = (uint32_t)t * 0x10001u;
    q.s    = t ? max(0, damp - cdef_msb(t)) : 0;
    q.mask = (0xFFFFu >> q.s) * 0x10001u;
    return q;
}
__device__ __forceinline__ Tap3F tap3f(const int t, const int damp) {
    Tap3F     q;
    const int s = t ? max(0, damp - cdef_msb(t)) : 0;
    q.on        = t != 0;
    q.nsc       = u2h((0x8000u | (uint32_t)((15 - s) << 10)) * 0x10001u);
    const float c = (float)t + (s ? 0.5f - __int_as_float((126 - s) << 23) : 0.f); // 2^-(s+1)
    q.C           = __float2half2_rn(c);
    return q;
}

// Two rows' worth of work is driven by the caller; this filters NP pairs (2*NP samples) of ONE row.
// ctr: centre words in the even-aligned copy; o[k*3+set] = word offsets of the +tap (.x) and -tap (.y) relative to ctr
// (they already select the even- or odd-aligned copy).  Returns integer sample pairs.
template <int NP>
__device__ __forceinline__ void cdef3_row_i(const uint32_t *ctr, const int2 (&o)[6], const Tap3I pr, const Tap3I se, const uint32_t wp0,
                                            const uint32_t wp1, uint32_t *outw) {
    uint32_t x[NP], P[NP], V[NP], mx[NP], mn[NP];
#pragma unroll
    for (int j = 0; j < NP; j++) x[j] = ctr[j], P[j] = 0, V[j] = 0, mx[j] = x[j], mn[j] = x[j];
#pragma unroll
    for (int k = 0; k < 2; k++) {
#pragma unroll
        for (int set = 0; set < 3; set++) {
            const uint32_t *q0 = ctr + o[k * 3 + set].x, *q1 = ctr + o[k * 3 + set].y;
            const Tap3I     t  = set == 0 ? pr : se;
            const uint32_t  w  = set == 0 ? (k ? wp1 : wp0) : 2u - k; // cdef_sec_taps {2,1}
#pragma unroll
            for (int j = 0; j < NP; j++) {
                const uint32_t p0 = q0[j], p1 = q1[j];
                if (t.t2) {
                    pk_tap(p0, x[j], t.t2, t.s, t.mask, w, P[j], V[j]);
                    pk_tap(p1, x[j], t.t2, t.s, t.mask, w, P[j], V[j]);
                }
                mx[j] = __vimax3_s16x2(mx[j], p0, p1); // the sentinel is negative as a signed halfword
                mn[j] = __vimin3_u16x2(mn[j], p0, p1);
            }
        }
    }
#pragma unroll
    for (int j = 0; j < NP; j++) {
        // y = clamp(x + ((8 + sum - (sum < 0)) >> 4), min, max), sum = 2P - V      (Common/Codec/EbCdef.c:300)
        // packed with a bias of 0x4000 (a multiple of 16) per half: u = sum + 0x4000, bit 14 clear <=> sum < 0
        const uint32_t u   = 2u * P[j] + (0x40004000u - V[j]);
        const uint32_t neg = (~u >> 14) & 0x00010001u;
        const uint32_t R   = ((u + 0x00080008u - neg) >> 4) & 0x0FFF0FFFu; // ((8 + sum - neg) >> 4) + 0x400
        outw[j]            = __vmins2(__viaddmax_s16x2(x[j] + R, 0xFC00FC00u, mn[j]), mx[j]);
    }
}

template <int NP>
__device__ __forceinline__ void cdef3_row_f(const uint32_t *ctr, const int2 (&o)[6], const Tap3F pr, const Tap3F se, const __half2 wp0,
                                            const __half2 wp1, uint32_t *outw) {
    const __half2 k1024 = u2h(0x64006400u), kn1024 = u2h(0xE400E400u), kone = u2h(0x3C003C00u);
    uint32_t      x[NP], mx[NP], mn[NP];
    __half2       acc[NP];
#pragma unroll
    for (int j = 0; j < NP; j++) x[j] = ctr[j], acc[j] = u2h(0u), mx[j] = x[j], mn[j] = x[j];
#pragma unroll
    for (int k = 0; k < 2; k++) {
#pragma unroll
        for (int set = 0; set < 3; set++) {
            const uint32_t *q0 = ctr + o[k * 3 + set].x, *q1 = ctr + o[k * 3 + set].y;
            const Tap3F     t  = set == 0 ? pr : se;
            const __half2   w  = set == 0 ? (k ? wp1 : wp0) : (k ? u2h(0x3C003C00u) : u2h(0x40004000u)); // cdef_sec_taps {2,1}
#pragma unroll
            for (int j = 0; j < NP; j++) {
                const uint32_t p0 = q0[j], p1 = q1[j];
                if (t.on) {
#pragma unroll
                    for (int e = 0; e < 2; e++) {
                        const __half2 d  = __hsub2(u2h(e ? p1 : p0), u2h(x[j]));
                        const __half2 a  = __habs2(d);
                        const __half2 u  = __hmin2(a, __hfma2(a, t.nsc, t.C));
                        const __half2 v  = __hfma2_relu(__hadd2(u, k1024), kone, kn1024);
                        const __half2 c  = u2h(h2u(v) | (h2u(d) & 0x80008000u));
                        acc[j]           = __hfma2(c, w, acc[j]);
                    }
                }
                mx[j] = __vimax3_s16x2(mx[j], p0, p1); // non-negative fp16 values order like their bit patterns
                mn[j] = __vimin3_u16x2(mn[j], p0, p1);
            }
        }
    }
    const __half2 k16 = u2h(0x2C002C00u), k32 = u2h(0x28002800u); // 1/16, 1/32
#pragma unroll
    for (int j = 0; j < NP; j++) {
        // (8 + sum - (sum < 0)) >> 4 = sign(sum) * floor((|sum| + 8) / 16) = sign(sum) * RN(|sum|/16 + 1/32)
        const __half2 q = __hadd2(__hadd2(__hfma2(__habs2(acc[j]), k16, k32), k1024), kn1024);
        const __half2 r = u2h(h2u(q) | (h2u(acc[j]) & 0x80008000u));
        const __half2 y = __hmin2(__hmax2(__hadd2(u2h(x[j]), r), u2h(mn[j])), u2h(mx[j]));
        outw[j]         = pk_h2i(h2u(y));
    }
}

// NP integer pairs starting at tile column `col` (any parity) of tile row `row` from an even-aligned integer tile
template <int NP>
__device__ __forceinline__ void c3_ld_pairs(const uint32_t *I, const int pw, const int row, const int col, uint32_t *out) {
    const int       odd = col & 1;
    const uint32_t *b   = I + row * pw + ((col - odd) >> 1);
    if (!odd) {
#pragma unroll
        for (int j = 0; j < NP; j++) out[j] = b[j];
    } else {
        uint32_t w[NP + 1];
#pragma unroll
        for (int j = 0; j <= NP; j++) w[j] = b[j];
#pragma unroll
        for (int j = 0; j < NP; j++) out[j] = __byte_perm(w[j], w[j + 1], 0x5432);
    }
}
// luma samples (row, col) and (row, col + 2) as one pair — co-located with two neighbouring chroma samples
__device__ __forceinline__ uint32_t c3_ld_skip(const uint16_t *l16, const int pitch16, const int row, const int col) {
    const uint16_t *p = l16 + row * pitch16 + col;
    return (uint32_t)p[0] | ((uint32_t)p[2] << 16);
}

template <typename T, bool FP>
__global__ void __launch_bounds__(C2_THREADS, 3) cdef3_kernel(const __grid_constant__ Cdef2Args a) {
    extern __shared__ __align__(16) uint32_t c3_sm[];
    uint32_t *const yA = c3_sm;                                   // yB = yA + C3_YSZ
    uint32_t *const cA0 = c3_sm + 2 * C3_YSZ;                     // plane p: cA0 + p * 2 * C3_CSZ, its B copy + C3_CSZ
    uint32_t *const yI = FP ? c3_sm + 2 * C3_YSZ + 4 * C3_CSZ : yA; // integer luma (direction search, CCSO classifier)
    __shared__ int     s_cost[64][9];                             // [block][dir], padded
    __shared__ int2    s_off[2][8][2];                            // [luma/chroma][dir][k]: word offsets of the +tap / -tap
    __shared__ uint8_t s_dir[64];
    __shared__ int     s_var[64];
    __shared__ uint8_t s_active[64];
    __shared__ __align__(16) uint8_t s_lut[3][SVTF_CCSO_LUT_SIZE]; // CCSO offsets + 128

    const int tid = threadIdx.x;
    const int fbc = blockIdx.x, fbr = blockIdx.y + a.fbr0;
    const int nhfb = gridDim.x;
    const int cw = a.d.mi_cols * 4, ch = a.d.mi_rows * 4;
    const int x0 = fbc * 64, y0 = fbr * 64;
    const int cs = a.d.bit_depth - 8;

    const T *iny  = static_cast<const T *>(a.in.p[0]);
    T       *outy = static_cast<T *>(a.out.p[0]);
    const int W = a.d.width, H = a.d.height;

    // ---- every global load of this thread, issued before the first use ----
    const int idx = __ldg(a.fb_idx + fbr * nhfb + fbc);
    int       flg = 0; // bit p: CCSO block flag of plane p (one flag per 256x256 luma samples, row = ceil(plane width / block))
    if (a.ccso.on) {
#pragma unroll
        for (int p = 0; p < 3; p++) {
            bool on = a.ccso.enable[p];
            if (on && a.ccso.blk_flags[p]) {
                const int pwid = W >> (p ? 1 : 0), bl = p ? 7 : 8;
                const int nbx  = (pwid + (1 << bl) - 1) >> bl;
                on             = (x0 >> 8) < nbx && __ldg(a.ccso.blk_flags[p] + ((y0 + a.ccso.y_glob) >> 8) * nbx + (x0 >> 8)) != 0;
            }
            flg |= (int)on << p;
        }
    }
    svtf_mi_t m4[4] = {~0u, ~0u, ~0u, ~0u};
    bool      mi_in = false;
    if (tid < 64) { // svt_sb_compute_cdef_list — Encoder/Codec/EbEncCdef.c:238-281
        const int r = (tid >> 3) * 2, c = (tid & 7) * 2;
        mi_in = fbr * 16 + r < a.d.mi_rows && fbc * 16 + c < a.d.mi_cols;
        if (mi_in) {
            const svtf_mi_t *m = a.mi + (fbr * 16 + r) * a.mi_stride + fbc * 16 + c;
            m4[0] = __ldg(m), m4[1] = __ldg(m + 1), m4[2] = __ldg(m + a.mi_stride), m4[3] = __ldg(m + a.mi_stride + 1);
        }
    }
    C3Chunk ky[3], kc[2];
    int     ry[3], cy[3];
#pragma unroll
    for (int k = 0; k < 3; k++) { // luma: 68 rows x 10 chunks
        const int i = tid + k * C2_THREADS;
        ry[k] = i / 10, cy[k] = (i - ry[k] * 10) * 8;
        ky[k] = c3_load<T>(iny, a.in.pitch[0], x0 - 8 + cy[k], y0 - 2 + ry[k], cw, ch, i < C2_YROWS * 10);
    }
    const int rc = tid / 6, cc = (tid - rc * 6) * 8; // chroma: 36 rows x 6 chunks per plane
#pragma unroll
    for (int p = 0; p < 2; p++)
        kc[p] = c3_load<T>(static_cast<const T *>(a.in.p[1 + p]), a.in.pitch[1 + p], x0 / 2 - 8 + cc, y0 / 2 - 2 + rc, cw / 2, ch / 2,
                           tid < C2_CROWS * 6);
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
        const int t = tid - 64, pt = t >> 4, dd = (t >> 1) & 7, k = t & 1;
        const int dy = (int)(((k ? 0x44443210u : 0x33332221u) >> (4 * dd)) & 15) - 2;
        const int dx = (int)(((k ? 0x12344444u : 0x22233333u) >> (4 * dd)) & 15) - 2;
        const int pw = pt ? C2_CPW : C2_YPW, bd = pt ? C3_CSZ : C3_YSZ, odd = dx & 1; // centre columns are even
        s_off[pt][dd][k] = make_int2((odd ? bd : 0) + dy * pw + ((dx - odd) >> 1), (odd ? bd : 0) - dy * pw + ((-dx - odd) >> 1));
    }
    const int  any_active = __syncthreads_or(act);
    const bool process    = (ylev | ysec | uvlev | uvsec) && any_active;
    const int  q_on       = flg; // bit p: CCSO (ccso_frame, Common/Codec/EbCcso.c:626) runs on plane p of this tile

    if (!process && !q_on) { // untouched filter block: carry the interior chunks over to the output surface
#pragma unroll
        for (int k = 0; k < 3; k++)
            if (ry[k] >= 2 && ry[k] < 66 && cy[k] >= 8 && cy[k] < 72 && (ky[k].m & 3) == 3)
                st8<T>(outy + (size_t)(y0 - 2 + ry[k]) * a.out.pitch[0] + x0 - 8 + cy[k], make_uint4(ky[k].w0, ky[k].w1, ky[k].w2, ky[k].w3));
#pragma unroll
        for (int p = 0; p < 2; p++)
            if (rc >= 2 && rc < 34 && cc >= 8 && cc < 40 && (kc[p].m & 1)) {
                T *o = static_cast<T *>(a.out.p[1 + p]) + (size_t)(y0 / 2 - 2 + rc) * a.out.pitch[1 + p] + x0 / 2 - 8 + cc;
                if ((kc[p].m & 3) == 3)
                    st8<T>(o, make_uint4(kc[p].w0, kc[p].w1, kc[p].w2, kc[p].w3));
                else
                    st4<T>(o, make_uint2(kc[p].w0, kc[p].w1));
            }
        return;
    }

#pragma unroll
    for (int k = 0; k < 3; k++)
        if (tid + k * C2_THREADS < C2_YROWS * 10)
            c3_store<FP>(yA, yA + C3_YSZ, FP ? yI : nullptr, ry[k] * C2_YPW + (cy[k] >> 1), ky[k]);
    if (tid < C2_CROWS * 6) {
        c3_store<FP>(cA0, cA0 + C3_CSZ, nullptr, rc * C2_CPW + (cc >> 1), kc[0]);
        c3_store<FP>(cA0 + 2 * C3_CSZ, cA0 + 3 * C3_CSZ, nullptr, rc * C2_CPW + (cc >> 1), kc[1]);
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
            const uint32_t *blk = yI + (2 + (b >> 3) * 8) * C2_YPW + 4 + (b & 7) * 4;
            int             c4[4];
            if (half == 0)
                cdef2_dir_costs4<0>(blk, C2_YPW, cs, c4);
            else
                cdef2_dir_costs4<4>(blk, C2_YPW, cs, c4);
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
        s_dir[tid] = (uint8_t)bd;
        s_var[tid] = (best - s_cost[tid][(bd + 4) & 7]) >> 10;
    }
    __syncthreads();

    const uint16_t *lum16   = reinterpret_cast<const uint16_t *>(yI);
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
        Tap3I      pri_i, sec_i;
        Tap3F      pri_f, sec_f;
        uint32_t   wi0 = 0, wi1 = 0;
        __half2    wf0 = u2h(0u), wf1 = u2h(0u);
        if (on) {
            const int pri_l = ylev << cs, sec = ysec << cs, damp = a.damping + cs;
            const int var = s_var[b]; // adjust_strength — Common/Codec/EbCdef.c:130-134
            const int ii  = (var >> 6) ? min(cdef_msb(var >> 6), 12) : 0;
            const int pri = var ? (pri_l * (4 + ii) + 8) >> 4 : 0;
            const int dir = pri_l ? s_dir[b] : 0;
            const int odd = (pri >> cs) & 1; // cdef_pri_taps {4,2} / {3,3}
#pragma unroll
            for (int k = 0; k < 2; k++) {
                o[k * 3 + 0] = s_off[0][dir][k];
                o[k * 3 + 1] = s_off[0][(dir + 2) & 7][k];
                o[k * 3 + 2] = s_off[0][(dir + 6) & 7][k];
            }
            if (FP) {
                pri_f = tap3f(pri, damp), sec_f = tap3f(sec, damp);
                wf0 = odd ? u2h(0x42004200u) : u2h(0x44004400u), wf1 = odd ? u2h(0x42004200u) : u2h(0x40004000u);
            } else {
                pri_i = tap3i(pri, damp), sec_i = tap3i(sec, damp);
                wi0 = odd ? 3u : 4u, wi1 = odd ? 3u : 2u;
            }
        }
        const bool    cc_on = (q_on & 1) != 0;
        const Ccso2Pk qp    = ccso2_pk(a.ccso, 0, a.d.bit_depth, s_lut[0]);
        if (inside) {
#pragma unroll 1
            for (int h = 0; h < 2; h++) {
                const int row = rp + 4 * h, gy = y0 + by * 8 + row;
                const int trow = 2 + by * 8 + row, tcol = 8 + bx * 8;
                uint32_t  o32[4];
                if (on) {
                    const uint32_t *ctr = yA + trow * C2_YPW + (tcol >> 1);
                    if (FP)
                        cdef3_row_f<4>(ctr, o, pri_f, sec_f, wf0, wf1, o32);
                    else
                        cdef3_row_i<4>(ctr, o, pri_i, sec_i, wi0, wi1, o32);
                } else {
                    const uint32_t *c = yI + trow * C2_YPW + (tcol >> 1);
#pragma unroll
                    for (int j = 0; j < 4; j++) o32[j] = c[j];
                }
                if (cc_on && gy < H) {
                    if (edge_lr) {
#pragma unroll
                        for (int j = 0; j < 8; j++)
                            if (gx + j < W) {
                                const int v = (o32[j >> 1] >> (16 * (j & 1))) & 0xffff;
                                const int r = ccso2_px(lum16, 2 * C2_YPW, trow, tcol + j, qp, ct, v);
                                o32[j >> 1] = (j & 1) ? ((o32[j >> 1] & 0xffffu) | ((uint32_t)r << 16)) : ((o32[j >> 1] & 0xffff0000u) | (uint32_t)r);
                            }
                    } else {
                        const uint32_t *yc = yI + trow * C2_YPW + (tcol >> 1);
                        uint32_t        ta[4], tb[4];
                        c3_ld_pairs<4>(yI, C2_YPW, min(max(trow + qp.dy0, ct.rmin), ct.rmax), tcol + qp.dx0, ta);
                        c3_ld_pairs<4>(yI, C2_YPW, min(max(trow - qp.dy0, ct.rmin), ct.rmax), tcol - qp.dx0, tb);
#pragma unroll
                        for (int j = 0; j < 4; j++) o32[j] = ccso2_pair(yc[j], ta[j], tb[j], o32[j], qp);
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
        Tap3I           pri_i, sec_i;
        Tap3F           pri_f, sec_f;
        uint32_t        wi0 = 0, wi1 = 0;
        __half2         wf0 = u2h(0u), wf1 = u2h(0u);
        if (on) {
            const int pri = uvlev << cs, sec = uvsec << cs, damp = a.damping + cs - 1;
            const int dir = pri ? s_dir[b] : 0;
            const int odd = (pri >> cs) & 1;
#pragma unroll
            for (int k = 0; k < 2; k++) {
                o[k * 3 + 0] = s_off[1][dir][k];
                o[k * 3 + 1] = s_off[1][(dir + 2) & 7][k];
                o[k * 3 + 2] = s_off[1][(dir + 6) & 7][k];
            }
            if (FP) {
                pri_f = tap3f(pri, damp), sec_f = tap3f(sec, damp);
                wf0 = odd ? u2h(0x42004200u) : u2h(0x44004400u), wf1 = odd ? u2h(0x42004200u) : u2h(0x40004000u);
            } else {
                pri_i = tap3i(pri, damp), sec_i = tap3i(sec, damp);
                wi0 = odd ? 3u : 4u, wi1 = odd ? 3u : 2u;
            }
        }
        const bool    cc_on = ((q_on >> (1 + p)) & 1) != 0;
        const Ccso2Pk qp    = ccso2_pk(a.ccso, 1 + p, a.d.bit_depth, s_lut[1 + p]);
        T            *outc  = static_cast<T *>(a.out.p[1 + p]);
        if (inside) {
#pragma unroll 1
            for (int h = 0; h < 2; h++) {
                const int row = rp + 2 * h, gy = y0 / 2 + by * 4 + row;
                const int trow = 2 + by * 4 + row, tcol = 8 + bx * 4;
                uint32_t  o32[2];
                const uint32_t *ctr = cA + trow * C2_CPW + (tcol >> 1);
                if (on) {
                    if (FP)
                        cdef3_row_f<2>(ctr, o, pri_f, sec_f, wf0, wf1, o32);
                    else
                        cdef3_row_i<2>(ctr, o, pri_i, sec_i, wi0, wi1, o32);
                } else {
                    o32[0] = FP ? pk_h2i(ctr[0]) : ctr[0], o32[1] = FP ? pk_h2i(ctr[1]) : ctr[1];
                }
                if (cc_on && gy < (H >> 1)) {
                    const int lr = 2 + 2 * (by * 4 + row), lc = 8 + 8 * bx; // co-located luma sample of chroma sample 0
                    if (edge_lr) {
#pragma unroll
                        for (int j = 0; j < 4; j++)
                            if (gx + j < (W >> 1)) {
                                const int v = (o32[j >> 1] >> (16 * (j & 1))) & 0xffff;
                                const int r = ccso2_px(lum16, 2 * C2_YPW, lr, lc + 2 * j, qp, ct, v);
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
