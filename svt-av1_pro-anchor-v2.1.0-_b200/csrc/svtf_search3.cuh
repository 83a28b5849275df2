// svtf_search3.cuh — the encoder's CDEF strength search: the separable formulation of svtf_search2.cuh with the taps,
// the rounding and the clamp in packed fp16 (8- and 10-bit pictures; the same exactness argument as svtf_cdef3.cuh).
//
// Same results as cdef_seg_search (Encoder/Codec/EbCdefProcess.c:114-357) + compute_cdef_dist (:60-107) +
// svt_aom_compute_cdef_dist_c/_8bit_c (Encoder/Codec/EbEncCdef.c:23-219).  Per sample pair the kernel evaluates the 4
// primary taps per primary strength and the 8 secondary taps per secondary strength once (svtf_search2.cuh explains
// why that is the same as filtering once per (pri,sec) pair); a combination is then
//   s  = Ap + S                                   HADD2   (exact: |s| <= 912)
//   r  = (8 + s - (s < 0)) >> 4 = s/16 rounded half away from zero = RN(s * (1/16 + 2^-14) + 1536) - 1536   (s3_err)
//   e' = clamp(r, min - x, max - x) + 1536        HFMA2 (rounds to the integer), HMNMX2 x2
// and because 1536 + e (|e| <= 57) lies in [1024, 2048), the low byte of each fp16 half IS e as int8: one PRMT packs
// four errors for the DP4A / DP2A moments, which stay integer exactly as in search2.  A primary tap costs 7
// instructions per pair (12 in search2), a secondary tap 1 + 3*6 (4 + 3*8), a combination 6 + the moments (9 + ...).
#pragma once
#include "svtf_cdef3.cuh"
#include "svtf_search2.cuh"

namespace svtf {

constexpr int S3_SMEM_WORDS = C3_SMEM_WORDS_F;

// one tap on a pair, fp16: acc += w * constrain(p - x)   (svtf_cdef3.cuh, cdef3_row_f)
__device__ __forceinline__ __half2 s3_constrain(const __half2 d, const Tap3F &t) {
    const __half2 k1024 = u2h(0x64006400u), kn1024 = u2h(0xE400E400u), kone = u2h(0x3C003C00u);
    // min(a, ceil(z)) = ceil(min(a, z)) for integer a: the min follows the rounding and takes the sign along (svtf_cdef5.cuh)
    const __half2 v = __hfma2_relu(__hadd2(__hfma2(__habs2(d), t.nsc, t.C), k1024), kone, kn1024);
    return u2h(pk_xsmin(h2u(d), h2u(v)));
}

// One row of NP pairs: secondary sums for the three non-zero secondary strengths and the clamp bounds
// (min - x, max - x over all 12 tap positions of direction `dir`, limited to +-100 and biased by 1536).
template <int NP>
__device__ __forceinline__ void s3_row_setup(const uint32_t *ctr, const int2 (&o)[6], const Tap3F (&sq)[3], const uint32_t *x, __half2 (*S)[NP],
                                             __half2 *lo, __half2 *hi) {
    uint32_t mx[NP], mn[NP];
#pragma unroll
    for (int j = 0; j < NP; j++) {
        mx[j] = x[j], mn[j] = x[j];
#pragma unroll
        for (int i = 0; i < 3; i++) S[i][j] = u2h(0u);
    }
#pragma unroll
    for (int k = 0; k < 2; k++) {
#pragma unroll
        for (int set = 0; set < 3; set++) {
            const uint32_t *q0 = ctr + o[k * 3 + set].x, *q1 = ctr + o[k * 3 + set].y;
            const __half2   w  = k ? u2h(0x3C003C00u) : u2h(0x40004000u); // cdef_sec_taps {2,1} (Common/Codec/EbCdef.c:250)
#pragma unroll
            for (int j = 0; j < NP; j++) {
                const uint32_t p0 = q0[j], p1 = q1[j];
                if (set != 0) {
#pragma unroll
                    for (int side = 0; side < 2; side++) {
                        const __half2 d = __hsub2(u2h(side ? p1 : p0), u2h(x[j]));
#pragma unroll
                        for (int i = 0; i < 3; i++) S[i][j] = __hfma2(s3_constrain(d, sq[i]), w, S[i][j]);
                    }
                }
                mx[j] = __vimax3_s16x2(mx[j], p0, p1);
                mn[j] = __vimin3_u16x2(mn[j], p0, p1);
            }
        }
    }
    const __half2 k1536 = u2h(0x66006600u), k100 = u2h(0x56405640u), kn100 = u2h(0xD640D640u);
#pragma unroll
    for (int j = 0; j < NP; j++) {
        lo[j] = __hadd2(__hmax2(__hsub2(u2h(mn[j]), u2h(x[j])), kn100), k1536);
        hi[j] = __hadd2(__hmin2(__hsub2(u2h(mx[j]), u2h(x[j])), k100), k1536);
    }
}

// One row of NP pairs: primary sum for strength pr (already adjusted).  o0 / o1: tap offsets of k = 0 / 1.
template <int NP>
__device__ __forceinline__ void s3_row_primary(const uint32_t *ctr, const int2 o0, const int2 o1, const Tap3F &pr, const __half2 w0,
                                               const __half2 w1, const uint32_t *x, __half2 *Ap) {
#pragma unroll
    for (int j = 0; j < NP; j++) Ap[j] = u2h(0u);
    if (!pr.on)
        return;
#pragma unroll
    for (int k = 0; k < 2; k++) {
        const uint32_t *q0 = ctr + (k ? o1.x : o0.x), *q1 = ctr + (k ? o1.y : o0.y);
        const __half2   w  = k ? w1 : w0;
#pragma unroll
        for (int j = 0; j < NP; j++) {
#pragma unroll
            for (int side = 0; side < 2; side++) {
                const __half2 d = __hsub2(u2h(side ? q1[j] : q0[j]), u2h(x[j]));
                Ap[j]           = __hfma2(s3_constrain(d, pr), w, Ap[j]);
            }
        }
    }
}

// filtered - unfiltered for one pair: fp16 halves 1536 + e, whose low bytes are e as int8
__device__ __forceinline__ uint32_t s3_err(const __half2 s, const __half2 lo, const __half2 hi) {
    // (8 + s - (s < 0)) >> 4 is s/16 rounded to the nearest integer, halves away from zero.  ONE fma does it: s * (1/16 + 2^-14) + 1536
    // is computed exactly and rounded once to the fp16 grid of [1024, 2048), i.e. to an integer; the 2^-14 term pushes a half away from
    // zero (|s| * 2^-14 > 0) and never carries a non-half across one (|s| * 2^-14 <= 912 / 16384 < 1/16, the spacing of s/16).
    const __half2 k16b = u2h(0x2C012C01u), k1536 = u2h(0x66006600u);
    const __half2 t    = __hfma2(s, k16b, k1536);
    return h2u(__hmin2(__hmax2(t, lo), hi));
}

__device__ __forceinline__ void s3_offsets(const int2 (*tab)[2], const int dir, int2 (&o)[6]) {
#pragma unroll
    for (int k = 0; k < 2; k++) {
        o[k * 3 + 0] = tab[dir][k];
        o[k * 3 + 1] = tab[(dir + 2) & 7][k];
        o[k * 3 + 2] = tab[(dir + 6) & 7][k];
    }
}

// ---- luma: thread = (8x8 block, two counted rows); 8 pairs per thread -------------------------------------------
template <typename T>
__device__ __forceinline__ void s3_luma(const Search2Args &a, const uint32_t *yA, const uint32_t *yI, const int2 (*tab)[2], const uint8_t *s_active,
                                        const uint8_t *s_dir, const int *s_var, unsigned long long *s_acc, const int x0, const int y0, const int cs,
                                        const int tid) {
    const int  b = tid >> 2, q = tid & 3, by = b >> 3, bx = b & 7;
    const int  sub = min(a.s.subsampling, 4); // EbCdefProcess.c:246-251; counted rows are 0, sub, 2*sub, ...
    const bool on  = s_active[b] && (2 * q + 1) * sub < 8;
    if (!__any_sync(0xffffffffu, on))
        return;
    const int rows[2] = {on ? 2 * q * sub : 0, on ? (2 * q + 1) * sub : 1};
    const int tcol    = 8 + bx * 8;
    const int damp    = a.s.damping + cs;
    uint32_t  xf[2][4], xi[2][4], d[2][4];
    int       X1 = 0, X2 = 0, XD = 0, D1 = 0, D2 = 0;
#pragma unroll
    for (int r = 0; r < 2; r++) {
        const int       off = (2 + by * 8 + rows[r]) * C2_YPW + (tcol >> 1);
        uint4           v   = make_uint4(0, 0, 0, 0);
        if (on)
            v = ld8<T>(static_cast<const T *>(a.s.src.p[0]) + (size_t)(y0 + by * 8 + rows[r]) * a.s.src.pitch[0] + x0 + bx * 8);
        d[r][0] = v.x, d[r][1] = v.y, d[r][2] = v.z, d[r][3] = v.w;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const uint32_t xx = yI[off + j];
            xi[r][j] = xx, xf[r][j] = yA[off + j];
            const int xl = xx & 0xffff, xh = xx >> 16, dl = d[r][j] & 0xffff, dh = d[r][j] >> 16;
            X1 += xl + xh, X2 += xl * xl + xh * xh, XD += xl * dl + xh * dh, D1 += dl + dh, D2 += dl * dl + dh * dh;
        }
    }
    if (!on)
        X1 = X2 = XD = D1 = D2 = 0;
    X1 = s2_quad_sum(X1), X2 = s2_quad_sum(X2), XD = s2_quad_sum(XD), D1 = s2_quad_sum(D1), D2 = s2_quad_sum(D2);

    const int var = s_var[b]; // adjust_strength — Common/Codec/EbCdef.c:130-134
    const int ii  = (var >> 6) ? min(cdef_msb(var >> 6), 12) : 0;
    Tap3F     sq[3];
#pragma unroll
    for (int i = 0; i < 3; i++) sq[i] = tap3f((i == 2 ? 4 : i + 1) << cs, damp);
    __half2 S[2][3][4], lo[2][4], hi[2][4];

    // pass 0: primary strength 0 -> direction 0; pass 1: primary 1..15 -> the block's direction (EbCdef.c:392-396)
#pragma unroll 1
    for (int pass = 0; pass < 2; pass++) {
        const unsigned long long pm = pass ? (a.present_y >> 4) : (a.present_y & 15);
        if (!pm)
            continue;
        int2 o[6];
        s3_offsets(tab, pass ? s_dir[b] : 0, o);
#pragma unroll
        for (int r = 0; r < 2; r++)
            s3_row_setup<4>(yA + (2 + by * 8 + rows[r]) * C2_YPW + (tcol >> 1), o, sq, xf[r], S[r], lo[r], hi[r]);
#pragma unroll 1
        for (int pi = pass; pi < (pass ? 16 : 1); pi++) {
            const unsigned m4 = (unsigned)(a.present_y >> (4 * pi)) & 15u;
            if (!m4)
                continue;
            const int     pri_l = pi << cs;
            const int     pri   = var ? (pri_l * (4 + ii) + 8) >> 4 : 0;
            const int     odd   = (pri >> cs) & 1; // cdef_pri_taps {4,2} / {3,3}
            const Tap3F   pr    = tap3f(pri, damp);
            const __half2 w0 = odd ? u2h(0x42004200u) : u2h(0x44004400u), w1 = odd ? u2h(0x42004200u) : u2h(0x40004000u);
            __half2       Ap[2][4];
#pragma unroll
            for (int r = 0; r < 2; r++)
                s3_row_primary<4>(yA + (2 + by * 8 + rows[r]) * C2_YPW + (tcol >> 1), o[0], o[3], pr, w0, w1, xf[r], Ap[r]);
            int E1[4], Q1[4], ED[4]; // per secondary index: S e, S e^2 + 2 S x e, S d e
#pragma unroll
            for (int si = 0; si < 4; si++) {
                int e1 = 0, ee = 0, xe = 0, de = 0;
                if ((m4 >> si) & 1) {
#pragma unroll
                    for (int r = 0; r < 2; r++)
#pragma unroll
                        for (int h = 0; h < 2; h++) {
                            const __half2  sa = si ? __hadd2(Ap[r][2 * h], S[r][si - 1][2 * h]) : Ap[r][2 * h];
                            const __half2  sb = si ? __hadd2(Ap[r][2 * h + 1], S[r][si - 1][2 * h + 1]) : Ap[r][2 * h + 1];
                            const uint32_t ea = s3_err(sa, lo[r][2 * h], hi[r][2 * h]);
                            const uint32_t eb = s3_err(sb, lo[r][2 * h + 1], hi[r][2 * h + 1]);
                            const int      e4 = (int)__byte_perm(ea, eb, 0x6420);
                            e1 = __dp4a(e4, 0x01010101, e1);
                            ee = __dp4a(e4, e4, ee);
                            xe = __dp2a_hi((int)xi[r][2 * h + 1], e4, __dp2a_lo((int)xi[r][2 * h], e4, xe));
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

// ---- chroma: thread = (4x4 block, row); 2 pairs per thread and plane; SSE only (EbEncCdef.c:103-126) --------------
template <typename T>
__device__ __forceinline__ void s3_chroma(const Search2Args &a, const int p, const uint32_t *cA, const int2 (*tab)[2], const uint8_t *s_active,
                                          const uint8_t *s_dir, unsigned long long *s_acc, const int x0, const int y0, const int cs, const int tid) {
    const int  b = tid >> 2, q = tid & 3, by = b >> 3, bx = b & 7;
    const bool on = s_active[b];
    if (!__any_sync(0xffffffffu, on))
        return;
    const int       trow = 2 + by * 4 + q, tcol = 8 + bx * 4;
    const int       damp = a.s.damping + cs - 1;
    const uint32_t *ctr  = cA + trow * C2_CPW + (tcol >> 1);
    uint32_t        x[2], g[2];
    int             G2 = 0;
    {
        uint2 v = make_uint2(0, 0);
        if (on)
            v = ld4<T>(static_cast<const T *>(a.s.src.p[1 + p]) + (size_t)(y0 / 2 + by * 4 + q) * a.s.src.pitch[1 + p] + x0 / 2 + bx * 4);
        const uint32_t dd[2] = {v.x, v.y};
#pragma unroll
        for (int j = 0; j < 2; j++) {
            x[j] = ctr[j];
            const uint32_t xi = on ? pk_h2i(x[j]) : 0u;
            g[j]              = __vadd2(dd[j], __vadd2(~xi, 0x00010001u)); // source - x per half (int16)
            const int gl = (int)(short)(g[j] & 0xffff), gh = (int)(short)(g[j] >> 16);
            G2 += gl * gl + gh * gh;
        }
    }
    if (!on)
        G2 = 0;
    Tap3F sq[3];
#pragma unroll
    for (int i = 0; i < 3; i++) sq[i] = tap3f((i == 2 ? 4 : i + 1) << cs, damp);
    __half2 S[3][2], lo[2], hi[2];
#pragma unroll 1
    for (int pass = 0; pass < 2; pass++) {
        const unsigned long long pm = pass ? (a.present_uv >> 4) : (a.present_uv & 15);
        if (!pm)
            continue;
        int2 o[6];
        s3_offsets(tab, pass ? s_dir[b] : 0, o);
        s3_row_setup<2>(ctr, o, sq, x, S, lo, hi);
#pragma unroll 1
        for (int pi = pass; pi < (pass ? 16 : 1); pi++) {
            const unsigned m4 = (unsigned)(a.present_uv >> (4 * pi)) & 15u;
            if (!m4)
                continue;
            const int     pri = pi << cs, odd = pi & 1;
            const Tap3F   pr  = tap3f(pri, damp);
            const __half2 w0 = odd ? u2h(0x42004200u) : u2h(0x44004400u), w1 = odd ? u2h(0x42004200u) : u2h(0x40004000u);
            __half2       Ap[2];
            s3_row_primary<2>(ctr, o[0], o[3], pr, w0, w1, x, Ap);
            int V[4]; // per secondary index: S g^2 - 2 S g e + S e^2 of this thread's 4 samples
#pragma unroll
            for (int si = 0; si < 4; si++) {
                int v = 0;
                if ((m4 >> si) & 1) {
                    const uint32_t ea = s3_err(si ? __hadd2(Ap[0], S[si - 1][0]) : Ap[0], lo[0], hi[0]);
                    const uint32_t eb = s3_err(si ? __hadd2(Ap[1], S[si - 1][1]) : Ap[1], lo[1], hi[1]);
                    const int      e4 = (int)__byte_perm(ea, eb, 0x6420);
                    const int      ge = __dp2a_hi((int)g[1], e4, __dp2a_lo((int)g[0], e4, 0));
                    v                 = G2 + __dp4a(e4, e4, 0) - 2 * ge;
                }
                V[si] = on ? v : 0;
            }
            unsigned long long sse = (unsigned long long)(unsigned)s2_quad_t(V[0], V[1], V[2], V[3], q);
            sse                    = s2_warp_quads(sse);
            if ((tid & 31) < 4 && ((m4 >> q) & 1))
                atomicAdd(&s_acc[pi * 4 + q], sse);
        }
    }
}

template <typename T>
__global__ void __launch_bounds__(S2_THREADS, 2) cdef_search3_kernel(const __grid_constant__ Search2Args a) {
    extern __shared__ __align__(16) uint32_t c3_sm[];
    uint32_t *const yA  = c3_sm;
    uint32_t *const cA0 = c3_sm + 2 * C3_YSZ;
    uint32_t *const yI  = c3_sm + 2 * C3_YSZ + 4 * C3_CSZ;
    __shared__ int                s_cost[64][9];
    __shared__ int2               s_off[2][8][2];
    __shared__ uint8_t            s_dir[64];
    __shared__ int                s_var[64];
    __shared__ uint8_t            s_active[64];
    __shared__ unsigned long long s_acc[3][64]; // indexed by strength byte (pri*4 + secondary index)

    const int tid = threadIdx.x;
    const int fbc = blockIdx.x, fbr = blockIdx.y + a.s.fbr0;
    const int fb  = fbr * gridDim.x + fbc;
    const int cw = a.s.d.mi_cols * 4, ch = a.s.d.mi_rows * 4;
    const int x0 = fbc * 64, y0 = fbr * 64;
    const int cs = a.s.d.bit_depth - 8;

    int act = 0;
    if (tid < 64) { // svt_sb_compute_cdef_list — Encoder/Codec/EbEncCdef.c:238-281
        const int r = (tid >> 3) * 2, c = (tid & 7) * 2;
        if (fbr * 16 + r < a.s.d.mi_rows && fbc * 16 + c < a.s.d.mi_cols) {
            const svtf_mi_t *m = a.s.mi + (fbr * 16 + r) * a.s.mi_stride + fbc * 16 + c;
            act = !(mi_skip(__ldg(m)) & mi_skip(__ldg(m + 1)) & mi_skip(__ldg(m + a.s.mi_stride)) & mi_skip(__ldg(m + a.s.mi_stride + 1)));
        }
        s_active[tid] = (uint8_t)act;
    } else if (tid < 96) { // tap offsets per direction, as in cdef3_kernel
        const int t = tid - 64, pt = t >> 4, dd = (t >> 1) & 7, k = t & 1;
        int       dy, dx;
        s2_dir(k, dd, dy, dx);
        const int pw = pt ? C2_CPW : C2_YPW, bd = pt ? C3_CSZ : C3_YSZ, odd = dx & 1;
        s_off[pt][dd][k] = make_int2((odd ? bd : 0) + dy * pw + ((dx - odd) >> 1), (odd ? bd : 0) - dy * pw + ((-dx - odd) >> 1));
    }
    for (int i = tid; i < 3 * 64; i += S2_THREADS) (&s_acc[0][0])[i] = 0;
    const int        any_active = __syncthreads_or(act);
    const SearchFold fold       = search_fold(a.s, fbr, fbc, gridDim.x);
    if (tid == 0) {
        if (fold.folded) // part of a 128-wide / 128-high coding block: the first quadrant's row carries the whole block
            atomicOr(&a.s.fold[fold.ofb], any_active ? 3u : 1u);
        else
            a.s.skip[fb] = any_active ? 0 : 1;
    }
    if (!any_active) // skip_cdef_seg = 1 (EbCdefProcess.c:203-207); mse rows stay zero
        return;

    {
        C3Chunk ky[3], kc[2];
        int     ry[3], cy[3];
#pragma unroll
        for (int k = 0; k < 3; k++) { // luma: 68 rows x 10 chunks
            const int i = tid + k * S2_THREADS;
            ry[k] = i / 10, cy[k] = (i - ry[k] * 10) * 8;
            ky[k] = c3_load<T>(static_cast<const T *>(a.s.rec.p[0]), a.s.rec.pitch[0], x0 - 8 + cy[k], y0 - 2 + ry[k], cw, ch, i < C2_YROWS * 10);
        }
        const int rc = tid / 6, cc = (tid - rc * 6) * 8; // chroma: 36 rows x 6 chunks per plane
#pragma unroll
        for (int p = 0; p < 2; p++)
            kc[p] = c3_load<T>(static_cast<const T *>(a.s.rec.p[1 + p]), a.s.rec.pitch[1 + p], x0 / 2 - 8 + cc, y0 / 2 - 2 + rc, cw / 2, ch / 2,
                               tid < C2_CROWS * 6);
#pragma unroll
        for (int k = 0; k < 3; k++)
            if (tid + k * S2_THREADS < C2_YROWS * 10)
                c3_store<true>(yA, yA + C3_YSZ, yI, ry[k] * C2_YPW + (cy[k] >> 1), ky[k]);
        if (tid < C2_CROWS * 6) {
            c3_store<true>(cA0, cA0 + C3_CSZ, nullptr, rc * C2_CPW + (cc >> 1), kc[0]);
            c3_store<true>(cA0 + 2 * C3_CSZ, cA0 + 3 * C3_CSZ, nullptr, rc * C2_CPW + (cc >> 1), kc[1]);
        }
    }
    __syncthreads();
    if (tid < 128) { // direction search: thread = (block, half of the directions)
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
    if (tid < 64) {
        int best = 0, bd = 0;
        if (s_active[tid]) {
#pragma unroll
            for (int i = 0; i < 8; i++) {
                const int c = s_cost[tid][i];
                if (c > best) { // first maximum wins ties (Common/Codec/EbCdef.c:197-201)
                    best = c;
                    bd   = i;
                }
            }
            const int v = (best - s_cost[tid][(bd + 4) & 7]) >> 10;
            s_var[tid]  = v;
            a.s.dir[fb * 64 + tid] = (uint8_t)bd;
            a.s.var[fb * 64 + tid] = v;
        } else
            s_var[tid] = 0;
        s_dir[tid] = (uint8_t)bd;
    }
    __syncthreads();

    s3_luma<T>(a, yA, yI, s_off[0], s_active, s_dir, s_var, s_acc[0], x0, y0, cs, tid);
    s3_chroma<T>(a, 0, cA0, s_off[1], s_active, s_dir, s_acc[1], x0, y0, cs, tid);
    s3_chroma<T>(a, 1, cA0 + 2 * C3_CSZ, s_off[1], s_active, s_dir, s_acc[2], x0, y0, cs, tid);
    __syncthreads();
    // mse_seg[0][fb][gi] = (sum >> 2cs) * sub ; mse_seg[1][fb][gi] = U + V (EbCdefProcess.c:297-300),
    // masked chroma strengths -> default_mse_uv * 64 (:86,:258-261)
    if (tid < a.s.n) {
        const int f = a.s.fs[tid];
        search_store(a.s, fb, fold, tid, s_acc[0][f], s_acc[1][f], s_acc[2][f], cs);
    }
}

} // namespace svtf
