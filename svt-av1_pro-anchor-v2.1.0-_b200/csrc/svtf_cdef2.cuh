// svtf_cdef2.cuh — CDEF apply (+ optional fused CCSO epilogue), packed 16-bit formulation.
//
// Same semantics as svt_av1_cdef_frame (Encoder/Codec/EbEncCdef.c:284-612) + svt_cdef_filter_fb /
// svt_cdef_filter_block_c (Common/Codec/EbCdef.c:253-430), and — when `ccso.on` — ccso_frame
// (Common/Codec/EbCcso.c:626-678) applied to the CDEF result before it is written, with the pre-CDEF luma
// tile that is already in shared memory as the classifier input.
//
// Why packed: ncu (profiles/r01_ncu_summary.md) shows the scalar kernel is ALU-issue bound at 142
// thread-instructions per sample.  Here two samples share one 32-bit register and every tap is evaluated with the
// sm_100a packed-halfword instructions (VIMNMX.U16x2, VIMNMX3, 32-bit adds that cannot carry between halves):
//
//   constrain(d) = sign(d) * v,  v = min(|d|, max(0, t - (|d| >> s)))          (Common/Codec/EbCdef.c:85-91)
//   hi = max(p,x)  lo = min(p,x)  a = hi-lo = |d|   dpos = hi-x = max(d,0)
//   m = min(t, a>>s)   v = min(a, t-m)   rp = min(v, dpos)       (rp = v where d>0, else 0; v = 0 when d = 0)
//   sum(w * constrain) = 2*sum(w*rp) - sum(w*v)                    (the d<0 part is v - rp)
//
// so a tap costs 7 ALU-pipe + 5 FMA-pipe instructions per sample PAIR.  Taps at odd sample offsets would need
// a funnel shift per pair; instead each tile is kept twice in shared memory, once with even and once with odd
// pair alignment, so every tap row is four aligned 32-bit shared loads.
#pragma once
#include "svtf_kernels.cuh"

namespace svtf {

constexpr int C2_THREADS = 256;
constexpr int C2_YPW     = 41; // luma pitch in 32-bit words: 82 samples = 8 + 64 + 8 (+2), odd
constexpr int C2_YROWS   = 68; // 2 + 64 + 2
constexpr int C2_CPW     = 25; // chroma pitch: 50 samples = 8 + 32 + 8 (+2), odd
constexpr int C2_CROWS   = 36;

struct CcsoFuse {
    int            on; // 0: plain CDEF
    const int8_t  *lut;          // device [3][2048]
    const uint8_t *blk_flags[3]; // device or NULL
    uint8_t        enable[3], bo_only[3], thr[3], support[3], edge_clf[3], band_log2[3];
    int            y_glob; // picture row of local row 0 (block flags are indexed in picture coordinates)
};

struct Cdef2Args {
    Planes           in, out;
    const svtf_mi_t *mi;
    int              mi_stride;
    Dims             d;
    const int8_t    *fb_idx;
    int              fbr0; // first filter-block row of the launch (stripe window)
    int              damping;
    uint8_t          y_strength[SVTF_CDEF_MAX_STRENGTHS], uv_strength[SVTF_CDEF_MAX_STRENGTHS];
    CcsoFuse         ccso;
};

__device__ __forceinline__ uint32_t pk_maxu(uint32_t a, uint32_t b) { return __vmaxu2(a, b); }
__device__ __forceinline__ uint32_t pk_minu(uint32_t a, uint32_t b) { return __vminu2(a, b); }

// CDEF_VERY_LARGE (0x7F7F) halves -> 0 so they never win a max (Common/Codec/EbCdef.c:273-290).  Valid samples are
// < 2^12, so bit 14 identifies the sentinel.
__device__ __forceinline__ uint32_t pk_kill_large(uint32_t p) {
    const uint32_t m = (p >> 14) & 0x00010001u;
    return p & ~(m * 0xFFFFu);
}

// One tap on a pair of samples; accumulates P += w*rp and V += w*v (see the header comment).
__device__ __forceinline__ void pk_tap(const uint32_t p, const uint32_t x, const uint32_t t2, const int s, const uint32_t smask,
                                       const uint32_t w, uint32_t &P, uint32_t &V) {
    const uint32_t hi = pk_maxu(p, x), lo = pk_minu(p, x);
    const uint32_t a = hi - lo, dpos = hi - x;
    const uint32_t sh = (a >> s) & smask;
    const uint32_t m  = pk_minu(t2, sh);
    const uint32_t v  = pk_minu(a, t2 - m);
    const uint32_t rp = pk_minu(v, dpos);
    P += w * rp;
    V += w * v;
}

// Pointer to the first word of the NP-pair row that starts `dx` samples right of tile column `col` (even) in
// tile row `row`; A = even-aligned copy, B = odd-aligned copy (word i of B holds samples 2i+1, 2i+2).
__device__ __forceinline__ const uint32_t *pk_row(const uint32_t *A, const uint32_t *B, const int pw, const int row, const int col,
                                                  const int dx) {
    const int c = col + dx, odd = c & 1;
    return (odd ? B : A) + row * pw + ((c - odd) >> 1);
}

// Filters NP pairs (2*NP samples) of one row.  xc: centre words.  Returns filtered samples packed like the input.
template <int NP, bool BORDER>
__device__ __forceinline__ void cdef2_row(const uint32_t *A, const uint32_t *B, const int pw, const int row, const int col,
                                          const int dir, const int pri, const int sec, const int damp, const int cs, uint32_t *outw) {
    uint32_t x[NP], P[NP], V[NP], mx[NP], mn[NP];
    {
        const uint32_t *c = A + row * pw + (col >> 1);
#pragma unroll
        for (int j = 0; j < NP; j++) x[j] = c[j], P[j] = 0, V[j] = 0, mx[j] = x[j], mn[j] = x[j];
    }
    const uint32_t pt2 = (uint32_t)pri * 0x10001u, st2 = (uint32_t)sec * 0x10001u;
    const int      ps = pri ? max(0, damp - cdef_msb(pri)) : 0, ss = sec ? max(0, damp - cdef_msb(sec)) : 0;
    const uint32_t pm = (0xFFFFu >> ps) * 0x10001u, sm = (0xFFFFu >> ss) * 0x10001u;
    const int      odd = (pri >> cs) & 1;
    const int      d2 = (dir + 2) & 7, d3 = (dir + 6) & 7;
#pragma unroll
    for (int k = 0; k < 2; k++) {
        const uint32_t wp = k ? (odd ? 3u : 2u) : (odd ? 3u : 4u); // cdef_pri_taps {4,2} / {3,3}
        const uint32_t ws = 2u - k;                                // cdef_sec_taps {2,1}
#pragma unroll
        for (int set = 0; set < 3; set++) {
            const int       dd = set == 0 ? dir : (set == 1 ? d2 : d3);
            // Cdef_Directions (Common/Codec/EbCdef.c:99-122) as nibble tables (value + 2) indexed by direction:
            // k=0: dy {-1,0,0,0,1,1,1,1} dx {1,1,1,1,1,0,0,0};  k=1: dy {-2,-1,0,1,2,2,2,2} dx {2,2,2,2,2,1,0,-1}
            const int dy = (int)(((k ? 0x44443210u : 0x33332221u) >> (4 * dd)) & 15) - 2;
            const int dx = (int)(((k ? 0x12344444u : 0x22233333u) >> (4 * dd)) & 15) - 2;
            const uint32_t *q0 = pk_row(A, B, pw, row + dy, col, dx);
            const uint32_t *q1 = pk_row(A, B, pw, row - dy, col, -dx);
            const uint32_t  t2 = set == 0 ? pt2 : st2;
            const int       s  = set == 0 ? ps : ss;
            const uint32_t  m  = set == 0 ? pm : sm;
            const uint32_t  w  = set == 0 ? wp : ws;
#pragma unroll
            for (int j = 0; j < NP; j++) {
                const uint32_t p0 = q0[j], p1 = q1[j];
                if (t2) {
                    pk_tap(p0, x[j], t2, s, m, w, P[j], V[j]);
                    pk_tap(p1, x[j], t2, s, m, w, P[j], V[j]);
                }
                mx[j] = __vimax3_u16x2(mx[j], BORDER ? pk_kill_large(p0) : p0, BORDER ? pk_kill_large(p1) : p1);
                mn[j] = __vimin3_u16x2(mn[j], p0, p1);
            }
        }
    }
#pragma unroll
    for (int j = 0; j < NP; j++) {
        // y = clamp(x + ((8 + sum - (sum < 0)) >> 4), min, max), sum = 2P - V      (Common/Codec/EbCdef.c:300)
        const int s0 = 2 * (int)(P[j] & 0xffff) - (int)(V[j] & 0xffff), s1 = 2 * (int)(P[j] >> 16) - (int)(V[j] >> 16);
        const int y0 = min(max((int)(x[j] & 0xffff) + ((8 + s0 - (s0 < 0)) >> 4), (int)(mn[j] & 0xffff)), (int)(mx[j] & 0xffff));
        const int y1 = min(max((int)(x[j] >> 16) + ((8 + s1 - (s1 < 0)) >> 4), (int)(mn[j] >> 16)), (int)(mx[j] >> 16));
        outw[j]      = (uint32_t)y0 | ((uint32_t)y1 << 16);
    }
}

// Stage rows x (8*chunks) samples; tile origin = (x0-8, y0-2); outside the coded plane -> CDEF_VERY_LARGE.
template <typename T>
__device__ __forceinline__ void cdef2_stage(uint32_t *A, const int pw, const int rows, const int chunks, const T *in, const int ipitch,
                                            const int x0, const int y0, const int cw, const int ch, const int tid) {
    const uint32_t LL = CDEF_LARGE | (CDEF_LARGE << 16);
    for (int i = tid; i < rows * chunks; i += C2_THREADS) {
        const int r = i / chunks, c = (i - r * chunks) * 8;
        const int gy = y0 - 2 + r, gx = x0 - 8 + c;
        uint4     v  = make_uint4(LL, LL, LL, LL);
        if (gy >= 0 && gy < ch && gx >= 0 && gx < cw) { // cw is a multiple of 8 (luma) / 4 (chroma)
            if (gx + 8 <= cw)
                v = ld8<T>(in + (size_t)gy * ipitch + gx);
            else { // chroma planes whose width is 4 mod 8: last half chunk
                const uint2 h = ld4<T>(in + (size_t)gy * ipitch + gx);
                v.x = h.x, v.y = h.y;
            }
        }
        uint32_t *d = A + r * pw + (c >> 1);
        d[0] = v.x, d[1] = v.y, d[2] = v.z, d[3] = v.w;
    }
}
// Odd-aligned copy: B[i] = (A[i].hi, A[i+1].lo)
__device__ __forceinline__ void cdef2_make_odd(const uint32_t *A, uint32_t *B, const int pw, const int rows, const int words, const int tid) {
    for (int i = tid; i < rows * (words - 1); i += C2_THREADS) {
        const int r = i / (words - 1), c = i - r * (words - 1);
        B[r * pw + c] = __byte_perm(A[r * pw + c], A[r * pw + c + 1], 0x5432);
    }
}

// Four directions (DBASE..DBASE+3) of one 8x8 block in one pass over its samples —
// svt_aom_cdef_find_dir_c, Common/Codec/EbCdef.c:150-210.  Tile samples are read as packed pairs.
template <int DBASE>
__device__ __forceinline__ void cdef2_dir_costs4(const uint32_t *blk, const int pw, const int cs, int *cost) {
    constexpr int NL0 = (DBASE == 0) ? 15 : 15, NL1 = 11, NL2 = 8, NL3 = 11;
    int p0[NL0], p1[NL1], p2[NL2], p3[NL3];
#pragma unroll
    for (int l = 0; l < NL0; l++) p0[l] = 0;
#pragma unroll
    for (int l = 0; l < NL1; l++) p1[l] = 0;
#pragma unroll
    for (int l = 0; l < NL2; l++) p2[l] = 0;
#pragma unroll
    for (int l = 0; l < NL3; l++) p3[l] = 0;
#pragma unroll
    for (int i = 0; i < 8; i++)
#pragma unroll
        for (int jp = 0; jp < 4; jp++) {
            const uint32_t w = blk[i * pw + jp];
#pragma unroll
            for (int h = 0; h < 2; h++) {
                const int j = 2 * jp + h;
                const int v = (int)(((h ? (w >> 16) : (w & 0xffff)) >> cs)) - 128;
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

// ---- fused CCSO ------------------------------------------------------------------------------------
// ccso_filter_block_hbd_wo_buf_c (Common/Codec/EbCcso.c:238-294) on sample PAIRS.  With d = tap - centre:
//   class = [d >= -thr] + [d > thr]            (edge_clf 0;  edge_clf 1 drops the second term, bo_only both)
//   [d >= -thr] = max(min(d + thr + 1, 1), 0)   [d > thr] = max(min(d - thr, 1), 0)
// each one VIADDMNMX.S16x2.RELU; the LUT is kept in shared memory biased by +128 so that the offset add cannot
// carry between the halves and "add, clip to [0, max]" is one more VIADDMNMX with addend -128.
struct Ccso2Pk { // per-plane state of one CTA, packed path
    int            dy0, dx0; // first classifier tap (the second is its mirror; derive_ccso_sample_pos, EbCcso.c:204-234)
    uint32_t       b1, b2, cap1, cap2, bmask, maxv2;
    int            shift;
    const uint8_t *lut; // shared, biased
};
__device__ __forceinline__ void ccso2_tap_of(const int support, int &dy0, int &dx0) {
    // (dy,dx) of tap 0 for ext_filter_support 0..5 as nibbles (value + 5)
    dy0 = (int)((0x556544u >> (4 * support)) & 15) - 5;
    dx0 = (int)((0x024445u >> (4 * support)) & 15) - 5;
}
__device__ __forceinline__ Ccso2Pk ccso2_pk(const CcsoFuse &f, const int p, const int bit_depth, const uint8_t *lut) {
    Ccso2Pk q;
    ccso2_tap_of(f.support[p], q.dy0, q.dx0);
    const uint32_t thr = f.thr[p];
    q.b1    = (thr + 1) * 0x10001u;
    q.b2    = ((0u - thr) & 0xffffu) * 0x10001u;
    q.cap1  = f.bo_only[p] ? 0u : 0x10001u;
    q.cap2  = (f.bo_only[p] || f.edge_clf[p]) ? 0u : 0x10001u;
    q.shift = bit_depth - f.band_log2[p]; // band_log2 0 -> shift = bit depth -> band 0 (the "single band" case)
    q.bmask = (0xFFFFu >> q.shift) * 0x10001u;
    q.maxv2 = ((1u << bit_depth) - 1) * 0x10001u;
    q.lut   = lut;
    return q;
}
// y2: centre luma pair, na/nb: the two tap pairs, v2: the (CDEF-filtered) samples the offset is added to
__device__ __forceinline__ uint32_t ccso2_pair(const uint32_t y2, const uint32_t na, const uint32_t nb, const uint32_t v2, const Ccso2Pk &q) {
    const uint32_t negy = __vadd2(~y2, 0x00010001u);
    const uint32_t da = __vadd2(na, negy), db = __vadd2(nb, negy);
    const uint32_t ca = __viaddmin_s16x2_relu(da, q.b1, q.cap1) + __viaddmin_s16x2_relu(da, q.b2, q.cap2);
    const uint32_t cb = __viaddmin_s16x2_relu(db, q.b1, q.cap1) + __viaddmin_s16x2_relu(db, q.b2, q.cap2);
    const uint32_t band = (y2 >> q.shift) & q.bmask;
    const uint32_t idx  = band * 16u + ca * 4u + cb;
    const uint32_t o    = (uint32_t)q.lut[idx & 0xffffu] | ((uint32_t)q.lut[idx >> 16] << 16);
    return __viaddmin_s16x2_relu(v2 + o, 0xFF80FF80u, q.maxv2);
}
// low halves of two consecutive words -> one pair (luma samples co-located with two neighbouring chroma samples)
__device__ __forceinline__ uint32_t pk_lo2(const uint32_t *w) { return __byte_perm(w[0], w[1], 0x5410); }

// Scalar path for tiles at the left/right picture edge, where the replicate border (EbCcso.c:185-201) clamps
// tap columns per sample.  (lr, lc): tile coordinates of the co-located luma sample.
struct Ccso2Tile {
    int rmin, rmax, cmin, cmax;
};
__device__ __forceinline__ int ccso2_px(const uint16_t *lum16, const int pitch16, const int lr, const int lc, const Ccso2Pk &q,
                                        const Ccso2Tile &t, const int v) {
    const int y  = lum16[lr * pitch16 + lc];
    const int n0 = lum16[min(max(lr + q.dy0, t.rmin), t.rmax) * pitch16 + min(max(lc + q.dx0, t.cmin), t.cmax)];
    const int n1 = lum16[min(max(lr - q.dy0, t.rmin), t.rmax) * pitch16 + min(max(lc - q.dx0, t.cmin), t.cmax)];
    const int d0 = n0 - y, d1 = n1 - y, thr = (int)(q.b1 & 0xffff) - 1;
    const int c0 = (q.cap1 ? (d0 >= -thr) : 0) + (q.cap2 ? (d0 > thr) : 0);
    const int c1 = (q.cap1 ? (d1 >= -thr) : 0) + (q.cap2 ? (d1 > thr) : 0);
    const int band = y >> q.shift;
    return min(max(v + (int)q.lut[(band << 4) + (c0 << 2) + c1] - 128, 0), (int)(q.maxv2 & 0xffff));
}

#ifdef SVTF_AB_KERNELS // superseded kernel, kept for A/B measurements only (not in the product build)
template <typename T>
// (C2_THREADS, 4) = 64 registers with 28 bytes of spills was measured: no gain (106 vs 105 us at 4K), occupancy is not the limiter
__global__ void __launch_bounds__(C2_THREADS, 3) cdef2_kernel(const __grid_constant__ Cdef2Args a) {
    __shared__ __align__(16) uint32_t yA[C2_YROWS * C2_YPW], yB[C2_YROWS * C2_YPW];
    __shared__ __align__(16) uint32_t cA[2][C2_CROWS * C2_CPW], cB[2][C2_CROWS * C2_CPW];
    __shared__ int                    s_cost[64][9]; // [block][dir], padded
    __shared__ uint8_t                s_dir[64];
    __shared__ int                    s_var[64];
    __shared__ uint8_t                s_active[64];
    __shared__ uint8_t                s_lut[3][SVTF_CCSO_LUT_SIZE]; // CCSO offsets + 128

    const int tid = threadIdx.x;
    const int fbc = blockIdx.x, fbr = blockIdx.y + a.fbr0;
    const int nhfb = gridDim.x;
    const int cw = a.d.mi_cols * 4, ch = a.d.mi_rows * 4;
    const int x0 = fbc * 64, y0 = fbr * 64;
    const int cs = a.d.bit_depth - 8;

    const int idx = a.fb_idx[fbr * nhfb + fbc];
    int       ylev = 0, ysec = 0, uvlev = 0, uvsec = 0;
    if (idx >= 0) {
        const int ys = a.y_strength[idx], us = a.uv_strength[idx];
        ylev = ys >> 2, ysec = ys & 3, ysec += ysec == 3;
        uvlev = us >> 2, uvsec = us & 3, uvsec += uvsec == 3;
    }
    int act = 0;
    if (tid < 64) { // svt_sb_compute_cdef_list — Encoder/Codec/EbEncCdef.c:238-281
        const int r = (tid >> 3) * 2, c = (tid & 7) * 2;
        if (fbr * 16 + r < a.d.mi_rows && fbc * 16 + c < a.d.mi_cols) {
            const svtf_mi_t *m = a.mi + (fbr * 16 + r) * a.mi_stride + fbc * 16 + c;
            act = !(mi_skip(__ldg(m)) & mi_skip(__ldg(m + 1)) & mi_skip(__ldg(m + a.mi_stride)) & mi_skip(__ldg(m + a.mi_stride + 1)));
        }
        s_active[tid] = (uint8_t)act;
    }
    const int  any_active = __syncthreads_or(act);
    const bool process    = (ylev | ysec | uvlev | uvsec) && any_active;

    // ---- fused CCSO (ccso_frame, Common/Codec/EbCcso.c:626): per-plane state of this 64x64 luma tile ----
    const int  W = a.d.width, H = a.d.height;
    int        q_on = 0; // bit p: CCSO runs on plane p of this tile
#pragma unroll
    for (int p = 0; p < 3; p++) {
        bool on = a.ccso.on && a.ccso.enable[p];
        if (on && a.ccso.blk_flags[p]) { // one flag per 256x256 luma samples, row length = ceil(plane width / block)
            const int pwid = W >> (p ? 1 : 0), bl = p ? 7 : 8;
            const int nbx  = (pwid + (1 << bl) - 1) >> bl;
            on             = (x0 >> 8) < nbx && a.ccso.blk_flags[p][((y0 + a.ccso.y_glob) >> 8) * nbx + (x0 >> 8)] != 0;
        }
        q_on |= (int)on << p;
    }
    const bool ccso_any = q_on != 0;
    const uint16_t *lum16   = reinterpret_cast<const uint16_t *>(yA);
    const bool      edge_lr = x0 - 5 < 0 || x0 + 69 > W; // tap columns can leave the picture: per-sample clamping
    Ccso2Tile       ct;
    ct.rmin = 2 - y0, ct.rmax = H - 1 - (y0 - 2), ct.cmin = 8 - x0, ct.cmax = W - 1 - (x0 - 8);

    const T *iny  = static_cast<const T *>(a.in.p[0]);
    T       *outy = static_cast<T *>(a.out.p[0]);
    if (!process && !ccso_any) { // untouched filter block: carry the samples over to the output surface
        for (int i = tid; i < 64 * 8; i += C2_THREADS) {
            const int r = i >> 3, c = (i & 7) * 8;
            if (y0 + r < ch && x0 + c < cw)
                st8<T>(outy + (size_t)(y0 + r) * a.out.pitch[0] + x0 + c, ld8<T>(iny + (size_t)(y0 + r) * a.in.pitch[0] + x0 + c));
        }
        for (int i = tid; i < 2 * 32 * 8; i += C2_THREADS) {
            const int p = 1 + (i >> 8), r = (i >> 3) & 31, c = (i & 7) * 4;
            if (y0 / 2 + r < ch / 2 && x0 / 2 + c < cw / 2)
                st4<T>(static_cast<T *>(a.out.p[p]) + (size_t)(y0 / 2 + r) * a.out.pitch[p] + x0 / 2 + c,
                       ld4<T>(static_cast<const T *>(a.in.p[p]) + (size_t)(y0 / 2 + r) * a.in.pitch[p] + x0 / 2 + c));
        }
        return;
    }

    cdef2_stage<T>(yA, C2_YPW, C2_YROWS, 10, iny, a.in.pitch[0], x0, y0, cw, ch, tid);
    cdef2_stage<T>(cA[0], C2_CPW, C2_CROWS, 6, static_cast<const T *>(a.in.p[1]), a.in.pitch[1], x0 / 2, y0 / 2, cw / 2, ch / 2, tid);
    cdef2_stage<T>(cA[1], C2_CPW, C2_CROWS, 6, static_cast<const T *>(a.in.p[2]), a.in.pitch[2], x0 / 2, y0 / 2, cw / 2, ch / 2, tid);
    if (ccso_any)
        for (int i = tid; i < 3 * SVTF_CCSO_LUT_SIZE / 4; i += C2_THREADS)
            reinterpret_cast<uint32_t *>(&s_lut[0][0])[i] = __ldg(reinterpret_cast<const uint32_t *>(a.ccso.lut) + i) ^ 0x80808080u;
    __syncthreads();
    cdef2_make_odd(yA, yB, C2_YPW, C2_YROWS, 40, tid);
    cdef2_make_odd(cA[0], cB[0], C2_CPW, C2_CROWS, 24, tid);
    cdef2_make_odd(cA[1], cB[1], C2_CPW, C2_CROWS, 24, tid);
    // ---- direction search: thread = (block, half of the directions) ----
    if (tid < 128 && process) {
        const int b = tid & 63, half = tid >> 6;
        if (s_active[b]) {
            const uint32_t *blk = yA + (2 + (b >> 3) * 8) * C2_YPW + 4 + (b & 7) * 4;
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

    const bool border = fbc == 0 || fbr == 0 || x0 + 66 > cw || y0 + 66 > ch;
    // ---- luma: work item = one 8-sample row of one 8x8 block ----
    {
        const int  pri_l = ylev << cs, sec = ysec << cs;
        const int  damp  = a.damping + cs;
        const bool on    = process && (ylev | ysec);
        for (int i = tid; i < 512; i += C2_THREADS) {
            const int bx = i & 7, row = (i >> 3) & 7, by = i >> 6;
            const int b  = by * 8 + bx;
            const int gy = y0 + by * 8 + row, gx = x0 + bx * 8;
            if (gy >= ch || gx >= cw)
                continue;
            const int trow = 2 + by * 8 + row, tcol = 8 + bx * 8;
            uint32_t  o32[4];
            if (on && s_active[b]) {
                const int var = s_var[b]; // adjust_strength — Common/Codec/EbCdef.c:130-134
                const int ii  = (var >> 6) ? min(cdef_msb(var >> 6), 12) : 0;
                const int pri = var ? (pri_l * (4 + ii) + 8) >> 4 : 0;
                const int dir = pri_l ? s_dir[b] : 0;
                if (border)
                    cdef2_row<4, true>(yA, yB, C2_YPW, trow, tcol, dir, pri, sec, damp, cs, o32);
                else
                    cdef2_row<4, false>(yA, yB, C2_YPW, trow, tcol, dir, pri, sec, damp, cs, o32);
            } else {
                const uint32_t *c = yA + trow * C2_YPW + (tcol >> 1);
#pragma unroll
                for (int j = 0; j < 4; j++) o32[j] = c[j];
            }
            if ((q_on & 1) && gy < H) {
                const Ccso2Pk qp = ccso2_pk(a.ccso, 0, a.d.bit_depth, s_lut[0]);
                if (edge_lr) {
#pragma unroll
                    for (int j = 0; j < 8; j++)
                        if (gx + j < W) {
                            const int v = (o32[j >> 1] >> (16 * (j & 1))) & 0xffff;
                            const int r = ccso2_px(lum16, 2 * C2_YPW, trow, tcol + j, qp, ct, v);
                            o32[j >> 1] = (j & 1) ? ((o32[j >> 1] & 0xffffu) | ((uint32_t)r << 16)) : ((o32[j >> 1] & 0xffff0000u) | (uint32_t)r);
                        }
                } else {
                    const uint32_t *yc = yA + trow * C2_YPW + (tcol >> 1);
                    const uint32_t *ta = pk_row(yA, yB, C2_YPW, min(max(trow + qp.dy0, ct.rmin), ct.rmax), tcol, qp.dx0);
                    const uint32_t *tb = pk_row(yA, yB, C2_YPW, min(max(trow - qp.dy0, ct.rmin), ct.rmax), tcol, -qp.dx0);
#pragma unroll
                    for (int j = 0; j < 4; j++) o32[j] = ccso2_pair(yc[j], ta[j], tb[j], o32[j], qp);
                }
            }
            st8<T>(outy + (size_t)gy * a.out.pitch[0] + gx, make_uint4(o32[0], o32[1], o32[2], o32[3]));
        }
    }
    // ---- chroma: work item = one 4-sample row of one 4x4 block, both planes ----
    {
        const int  pri = uvlev << cs, sec = uvsec << cs;
        const int  damp = a.damping + cs - 1;
        const bool on   = process && (uvlev | uvsec);
        for (int i = tid; i < 512; i += C2_THREADS) {
            const int p  = i >> 8;
            const int bx = i & 7, row = (i >> 3) & 3, by = (i >> 5) & 7;
            const int b  = by * 8 + bx;
            const int gy = y0 / 2 + by * 4 + row, gx = x0 / 2 + bx * 4;
            if (gy >= ch / 2 || gx >= cw / 2)
                continue;
            const int trow = 2 + by * 4 + row, tcol = 8 + bx * 4;
            uint32_t  o32[2];
            if (on && s_active[b]) {
                const int dir = pri ? s_dir[b] : 0;
                if (border)
                    cdef2_row<2, true>(cA[p], cB[p], C2_CPW, trow, tcol, dir, pri, sec, damp, cs, o32);
                else
                    cdef2_row<2, false>(cA[p], cB[p], C2_CPW, trow, tcol, dir, pri, sec, damp, cs, o32);
            } else {
                const uint32_t *c = cA[p] + trow * C2_CPW + (tcol >> 1);
                o32[0] = c[0], o32[1] = c[1];
            }
            if (((q_on >> (1 + p)) & 1) && gy < (H >> 1)) {
                const Ccso2Pk qp = ccso2_pk(a.ccso, 1 + p, a.d.bit_depth, s_lut[1 + p]);
                const int     lr = 2 + 2 * (by * 4 + row), lc = 8 + 8 * bx; // co-located luma sample of chroma sample 0
                if (edge_lr) {
#pragma unroll
                    for (int j = 0; j < 4; j++)
                        if (gx + j < (W >> 1)) {
                            const int v = (o32[j >> 1] >> (16 * (j & 1))) & 0xffff;
                            const int r = ccso2_px(lum16, 2 * C2_YPW, lr, lc + 2 * j, qp, ct, v);
                            o32[j >> 1] = (j & 1) ? ((o32[j >> 1] & 0xffffu) | ((uint32_t)r << 16)) : ((o32[j >> 1] & 0xffff0000u) | (uint32_t)r);
                        }
                } else {
                    const uint32_t *yc = yA + lr * C2_YPW + (lc >> 1);
                    const uint32_t *ta = pk_row(yA, yB, C2_YPW, min(max(lr + qp.dy0, ct.rmin), ct.rmax), lc, qp.dx0);
                    const uint32_t *tb = pk_row(yA, yB, C2_YPW, min(max(lr - qp.dy0, ct.rmin), ct.rmax), lc, -qp.dx0);
#pragma unroll
                    for (int k = 0; k < 2; k++) o32[k] = ccso2_pair(pk_lo2(yc + 2 * k), pk_lo2(ta + 2 * k), pk_lo2(tb + 2 * k), o32[k], qp);
                }
            }
            st4<T>(static_cast<T *>(a.out.p[1 + p]) + (size_t)gy * a.out.pitch[1 + p] + gx, make_uint2(o32[0], o32[1]));
        }
    }
}
#endif // SVTF_AB_KERNELS

} // namespace svtf
