// svtf_search2.cuh — the encoder's CDEF strength search, separable + packed formulation (8- and 10-bit).
//
// Same results as cdef_seg_search (Encoder/Codec/EbCdefProcess.c:114-357) + compute_cdef_dist (:60-107) +
// svt_aom_compute_cdef_dist_c/_8bit_c (Encoder/Codec/EbEncCdef.c:23-219); the reference filters every filter block
// once per (pri,sec) strength pair (64 pairs at cdef_level 1) with svt_cdef_filter_fb and measures each result.
//
// What is shared between the 64 filterings (Common/Codec/EbCdef.c:253-300):
//   y = clamp(x + round16(sum_pri(pri) + sum_sec(sec)), min, max)
//   * sum_pri depends only on the primary strength (16 values), sum_sec only on the secondary (0,1,2,4);
//   * min/max over the 12 tap positions depend only on the direction, which is the block's direction when
//     pri != 0 and direction 0 when pri == 0 (svt_cdef_filter_fb, :392-396);
// so per sample pair the kernel evaluates 4 primary taps per pri and 8 secondary taps per sec ONCE (108 tap
// evaluations instead of 64*12 = 768) and each of the 64 combinations costs one add, a rounding, a clamp and the
// distortion moments.  Rounding on biased sums: with u = sum + 2048 (never negative, < 4096),
//   (8 + sum - (sum < 0)) >> 4  + 128  =  (u + 7 + bit11(u)) >> 4.
// Distortion: with e = y - x (|e| <= 57 at 10 bits, fits int8) the five moments of dist_8xn are
//   sum_s = Sx + Se,  sum_s2 = Sx2 + 2*Sxe + See,  sum_sd = Sxd + Sde   (sum_d, sum_d2 do not change)
// and Se, See, Sxe, Sde of four samples are DP4A / DP2A instructions on the packed errors; chroma SSE is
// S(g-e)^2 = Sgg - 2*Sge + See with g = source - x.  All integer, so the sums equal the reference's exactly; the
// double-precision luma formula is evaluated on those sums with the same explicit round-to-nearest operations.
#pragma once
#include "svtf_cdef2.cuh"
#include "svtf_search.cuh"

namespace svtf {

constexpr int      S2_THREADS = 256;
constexpr uint32_t S2_B       = 2048; // bias of a full sum
constexpr uint32_t S2_B2      = 256;  // bias of a secondary sum
constexpr uint32_t S2_R0      = 128;  // = S2_B / 16, bias of the rounded offset

struct Search2Args {
    SearchArgs         s;
    unsigned long long present_y, present_uv; // bit f: strength byte f (pri*4 + sec index) is in the list (chroma: unmasked)
};

// Cdef_Directions (Common/Codec/EbCdef.c:99-122) as nibble tables (value + 2), see cdef2_row
__device__ __forceinline__ void s2_dir(const int k, const int dd, int &dy, int &dx) {
    dy = (int)(((k ? 0x44443210u : 0x33332221u) >> (4 * dd)) & 15) - 2;
    dx = (int)(((k ? 0x12344444u : 0x22233333u) >> (4 * dd)) & 15) - 2;
}

// strength-dependent part of one tap on a pair: acc += w * (2*rp - v)   (sum of w*constrain, see svtf_cdef2.cuh)
__device__ __forceinline__ void s2_tap(const uint32_t a, const uint32_t dpos, const uint32_t t2, const int s, const uint32_t smask,
                                       const int w, uint32_t &acc) {
    const uint32_t sh = (a >> s) & smask;
    const uint32_t m  = pk_minu(t2, sh);
    const uint32_t v  = pk_minu(a, t2 - m);
    const uint32_t rp = pk_minu(v, dpos);
    acc += (uint32_t)(2 * w) * rp;
    acc -= (uint32_t)w * v;
}

struct S2Sec { // the three non-zero secondary strengths of one plane type
    uint32_t t2[3], m[3];
    int      s[3];
};
__device__ __forceinline__ S2Sec s2_sec_setup(const int cs, const int damp) {
    S2Sec q;
#pragma unroll
    for (int i = 0; i < 3; i++) {
        const int sec = (i == 2 ? 4 : i + 1) << cs;
        q.t2[i]       = (uint32_t)sec * 0x10001u;
        q.s[i]        = max(0, damp - cdef_msb(sec));
        q.m[i]        = (0xFFFFu >> q.s[i]) * 0x10001u;
    }
    return q;
}

// One row of NP pairs: secondary sums for the three strengths (biased by S2_B2) and min/max over all 12 tap
// positions for direction `dir` (biased by S2_R0).
template <int NP, bool BORDER>
__device__ __forceinline__ void s2_row_setup(const uint32_t *A, const uint32_t *B, const int pw, const int row, const int col, const int dir,
                                             const S2Sec &q, const uint32_t *x, uint32_t (*S)[NP], uint32_t *mnb, uint32_t *mxb) {
    uint32_t mx[NP], mn[NP];
#pragma unroll
    for (int j = 0; j < NP; j++) {
        mx[j] = x[j], mn[j] = x[j];
#pragma unroll
        for (int i = 0; i < 3; i++) S[i][j] = S2_B2 * 0x10001u;
    }
    const int d2 = (dir + 2) & 7, d3 = (dir + 6) & 7;
#pragma unroll
    for (int k = 0; k < 2; k++) {
#pragma unroll
        for (int set = 0; set < 3; set++) {
            int dy, dx;
            s2_dir(k, set == 0 ? dir : (set == 1 ? d2 : d3), dy, dx);
            const uint32_t *q0 = pk_row(A, B, pw, row + dy, col, dx);
            const uint32_t *q1 = pk_row(A, B, pw, row - dy, col, -dx);
#pragma unroll
            for (int j = 0; j < NP; j++) {
#pragma unroll
                for (int side = 0; side < 2; side++) {
                    const uint32_t p = side ? q1[j] : q0[j];
                    if (set != 0) { // secondary taps {2,1} (Common/Codec/EbCdef.c:250)
                        const uint32_t hi = pk_maxu(p, x[j]), lo = pk_minu(p, x[j]);
                        const uint32_t a = hi - lo, dpos = hi - x[j];
#pragma unroll
                        for (int i = 0; i < 3; i++) s2_tap(a, dpos, q.t2[i], q.s[i], q.m[i], 2 - k, S[i][j]);
                    }
                    mx[j] = pk_maxu(mx[j], BORDER ? pk_kill_large(p) : p);
                    mn[j] = pk_minu(mn[j], p);
                }
            }
        }
    }
#pragma unroll
    for (int j = 0; j < NP; j++) mnb[j] = mn[j] + S2_R0 * 0x10001u, mxb[j] = mx[j] + S2_R0 * 0x10001u;
}

// One row of NP pairs: primary sum for strength `pri` (already adjusted), biased by S2_B - S2_B2.
template <int NP>
__device__ __forceinline__ void s2_row_primary(const uint32_t *A, const uint32_t *B, const int pw, const int row, const int col, const int dir,
                                               const int pri, const int damp, const int cs, const uint32_t *x, uint32_t *Ap) {
#pragma unroll
    for (int j = 0; j < NP; j++) Ap[j] = (S2_B - S2_B2) * 0x10001u;
    if (!pri)
        return;
    const uint32_t t2  = (uint32_t)pri * 0x10001u;
    const int      s   = max(0, damp - cdef_msb(pri));
    const uint32_t m   = (0xFFFFu >> s) * 0x10001u;
    const int      odd = (pri >> cs) & 1;
#pragma unroll
    for (int k = 0; k < 2; k++) {
        const int w = k ? (odd ? 3 : 2) : (odd ? 3 : 4); // cdef_pri_taps {4,2} / {3,3}
        int       dy, dx;
        s2_dir(k, dir, dy, dx);
        const uint32_t *q0 = pk_row(A, B, pw, row + dy, col, dx);
        const uint32_t *q1 = pk_row(A, B, pw, row - dy, col, -dx);
#pragma unroll
        for (int j = 0; j < NP; j++) {
#pragma unroll
            for (int side = 0; side < 2; side++) {
                const uint32_t p  = side ? q1[j] : q0[j];
                const uint32_t hi = pk_maxu(p, x[j]), lo = pk_minu(p, x[j]);
                s2_tap(hi - lo, hi - x[j], t2, s, m, w, Ap[j]);
            }
        }
    }
}

// filtered - unfiltered for one pair, as two int16 halves (|e| <= 127)
__device__ __forceinline__ uint32_t s2_err(const uint32_t Ap, const uint32_t S, const uint32_t x, const uint32_t nx, const uint32_t mnb,
                                           const uint32_t mxb) {
    const uint32_t u = Ap + S;                              // sum + 2048 per half
    const uint32_t b = (u >> 11) & 0x00010001u;             // sum >= 0
    const uint32_t r = ((u + 0x00070007u + b) >> 4) & 0x0FFF0FFFu; // round16(sum) + 128
    const uint32_t y = pk_minu(pk_maxu(x + r, mnb), mxb);   // filtered + 128
    return __vadd2(y, nx);                                  // nx = -(x + 128) per half
}

// lane q of every quad ends with the quad total of v[q]
__device__ __forceinline__ int s2_quad_t(const int v0, const int v1, const int v2, const int v3, const int q) {
    const bool hi = q & 2, od = q & 1;
    int        k0 = hi ? v2 : v0, k1 = hi ? v3 : v1;
    k0 += __shfl_xor_sync(0xffffffffu, hi ? v0 : v2, 2);
    k1 += __shfl_xor_sync(0xffffffffu, hi ? v1 : v3, 2);
    int k = od ? k1 : k0;
    k += __shfl_xor_sync(0xffffffffu, od ? k0 : k1, 1);
    return k;
}
__device__ __forceinline__ int s2_quad_sum(int v) {
    v += __shfl_xor_sync(0xffffffffu, v, 1);
    v += __shfl_xor_sync(0xffffffffu, v, 2);
    return v;
}
// sum over the 8 quads of a warp (lanes with equal q), result valid in lanes 0..3
__device__ __forceinline__ unsigned long long s2_warp_quads(unsigned long long v) {
    v += __shfl_xor_sync(0xffffffffu, v, 4);
    v += __shfl_xor_sync(0xffffffffu, v, 8);
    v += __shfl_xor_sync(0xffffffffu, v, 16);
    return v;
}

// ---- luma: thread = (8x8 block, two counted rows); 8 pairs per thread -------------------------------------------
template <typename T, bool BORDER>
__device__ __forceinline__ void s2_luma(const Search2Args &a, const uint32_t *yA, const uint32_t *yB, const uint8_t *s_active, const uint8_t *s_dir,
                                        const int *s_var, unsigned long long *s_acc, const int x0, const int y0, const int cs, const int tid) {
    const int  b = tid >> 2, q = tid & 3, by = b >> 3, bx = b & 7;
    const int  sub = min(a.s.subsampling, 4); // EbCdefProcess.c:246-251; counted rows are 0, sub, 2*sub, ...
    const bool on  = s_active[b] && (2 * q + 1) * sub < 8;
    if (!__any_sync(0xffffffffu, on))
        return;
    const int rows[2] = {on ? 2 * q * sub : 0, on ? (2 * q + 1) * sub : 1};
    const int tcol    = 8 + bx * 8;
    const int damp    = a.s.damping + cs;
    uint32_t  x[2][4], nx[2][4], d[2][4];
    int       X1 = 0, X2 = 0, XD = 0, D1 = 0, D2 = 0;
#pragma unroll
    for (int r = 0; r < 2; r++) {
        const uint32_t *c = yA + (2 + by * 8 + rows[r]) * C2_YPW + (tcol >> 1);
        uint4           v = make_uint4(0, 0, 0, 0);
        if (on)
            v = ld8<T>(static_cast<const T *>(a.s.src.p[0]) + (size_t)(y0 + by * 8 + rows[r]) * a.s.src.pitch[0] + x0 + bx * 8);
        d[r][0] = v.x, d[r][1] = v.y, d[r][2] = v.z, d[r][3] = v.w;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const uint32_t xx = c[j];
            x[r][j]  = xx;
            nx[r][j] = __vadd2(~(xx + S2_R0 * 0x10001u), 0x00010001u);
            const int xl = xx & 0xffff, xh = xx >> 16, dl = d[r][j] & 0xffff, dh = d[r][j] >> 16;
            X1 += xl + xh, X2 += xl * xl + xh * xh, XD += xl * dl + xh * dh, D1 += dl + dh, D2 += dl * dl + dh * dh;
        }
    }
    if (!on)
        X1 = X2 = XD = D1 = D2 = 0;
    X1 = s2_quad_sum(X1), X2 = s2_quad_sum(X2), XD = s2_quad_sum(XD), D1 = s2_quad_sum(D1), D2 = s2_quad_sum(D2);

    const int   var = s_var[b]; // adjust_strength — Common/Codec/EbCdef.c:130-134
    const int   ii  = (var >> 6) ? min(cdef_msb(var >> 6), 12) : 0;
    const S2Sec sq  = s2_sec_setup(cs, damp);
    uint32_t    S[2][3][4], mnb[2][4], mxb[2][4];

    // pass 0: primary strength 0 -> direction 0; pass 1: primary 1..15 -> the block's direction (EbCdef.c:392-396)
#pragma unroll 1
    for (int pass = 0; pass < 2; pass++) {
        const unsigned long long pm = pass ? (a.present_y >> 4) : (a.present_y & 15);
        if (!pm)
            continue;
        const int dir = pass ? s_dir[b] : 0;
#pragma unroll
        for (int r = 0; r < 2; r++) s2_row_setup<4, BORDER>(yA, yB, C2_YPW, 2 + by * 8 + rows[r], tcol, dir, sq, x[r], S[r], mnb[r], mxb[r]);
#pragma unroll 1
        for (int pi = pass; pi < (pass ? 16 : 1); pi++) {
            const unsigned m4 = (unsigned)(a.present_y >> (4 * pi)) & 15u;
            if (!m4)
                continue;
            const int pri_l = pi << cs;
            const int pri   = var ? (pri_l * (4 + ii) + 8) >> 4 : 0;
            uint32_t  Ap[2][4];
#pragma unroll
            for (int r = 0; r < 2; r++) s2_row_primary<4>(yA, yB, C2_YPW, 2 + by * 8 + rows[r], tcol, dir, pri, damp, cs, x[r], Ap[r]);
            int E1[4], Q1[4], ED[4]; // per secondary index: S e, S e^2 + 2 S x e, S d e
#pragma unroll
            for (int si = 0; si < 4; si++) {
                int e1 = 0, ee = 0, xe = 0, de = 0;
                if ((m4 >> si) & 1) {
#pragma unroll
                    for (int r = 0; r < 2; r++)
#pragma unroll
                        for (int h = 0; h < 2; h++) {
                            const uint32_t sa = si ? S[r][si - 1][2 * h] : S2_B2 * 0x10001u, sb = si ? S[r][si - 1][2 * h + 1] : S2_B2 * 0x10001u;
                            const uint32_t ea = s2_err(Ap[r][2 * h], sa, x[r][2 * h], nx[r][2 * h], mnb[r][2 * h], mxb[r][2 * h]);
                            const uint32_t eb = s2_err(Ap[r][2 * h + 1], sb, x[r][2 * h + 1], nx[r][2 * h + 1], mnb[r][2 * h + 1], mxb[r][2 * h + 1]);
                            const int      e4 = (int)__byte_perm(ea, eb, 0x6420);
                            e1 = __dp4a(e4, 0x01010101, e1);
                            ee = __dp4a(e4, e4, ee);
                            xe = __dp2a_hi((int)x[r][2 * h + 1], e4, __dp2a_lo((int)x[r][2 * h], e4, xe));
                            de = __dp2a_hi((int)d[r][2 * h + 1], e4, __dp2a_lo((int)d[r][2 * h], e4, de));
                        }
                }
                E1[si] = on ? e1 : 0, Q1[si] = on ? ee + 2 * xe : 0, ED[si] = on ? de : 0;
            }
            const int e1 = s2_quad_t(E1[0], E1[1], E1[2], E1[3], q);
            const int q1 = s2_quad_t(Q1[0], Q1[1], Q1[2], Q1[3], q);
            const int ed = s2_quad_t(ED[0], ED[1], ED[2], ED[3], q);
            // lane q: distortion of this block for secondary index q
            unsigned long long dist = 0;
            if (s_active[b] && ((m4 >> q) & 1))
                dist = cdef_dist_8x8((unsigned long long)(X1 + e1), (unsigned long long)D1, (unsigned long long)(X2 + q1), (unsigned long long)D2,
                                     (unsigned long long)(XD + ed), cs);
            dist = s2_warp_quads(dist);
            if ((tid & 31) < 4 && ((m4 >> q) & 1))
                atomicAdd(&s_acc[pi * 4 + q], dist);
        }
    }
}

} // namespace svtf
