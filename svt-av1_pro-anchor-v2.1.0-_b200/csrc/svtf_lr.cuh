// svtf_lr.cuh — loop restoration (Wiener / self-guided) on the device: the stage after CDEF / CCSO on the same surfaces.
//
// Reference: svt_av1_loop_restoration_filter_frame (Common/Codec/EbRestoration.c:1179) -> svt_av1_loop_restoration_filter_unit
// (:1067) per restoration unit and 64-row stripe, with the stripe borders prepared by svt_aom_setup_processing_stripe_boundary
// (:288) from the lines svt_av1_loop_restoration_save_boundary_lines (:1682) kept of the DEBLOCKED picture; filters
// svt_av1_[highbd_]wiener_convolve_add_src_c (Common/Codec/convolve.c:100/194) and svt_apply_selfguided_restoration_c (:957).
//
// The reference patches three rows above and below every stripe into the picture, filters, and restores them — a serial
// walk over one buffer.  Here both pictures (current = after CDEF/CCSO, deblocked = what CDEF read) are resident, so a CTA
// = (plane, stripe, 64-column processing unit) simply STAGES its (h + 6) x (w + 6) input in shared memory, choosing the
// source per row: inside the stripe the current picture; the 3 rows above / below from the deblocked picture (two rows
// per boundary, the outer one repeated) or, at the top / bottom of the frame, the current picture's edge row; columns
// outside the plane repeat the edge column.  Every CTA is independent; nothing is patched or restored.
#pragma once
#include "svtf_kernels.cuh"

namespace svtf {

constexpr int LR_THREADS = 256;
constexpr int LR_MAXW    = 64;               // processing-unit width (RESTORATION_PROC_UNIT_SIZE), 32 for 4:2:0 chroma
constexpr int LR_C0      = 8;                // shared-input column of plane column x0 (column 0 <-> x0 - 8: 8-sample vectors line up)
constexpr int LR_IN_P    = LR_MAXW + 16 + 2; // shared input pitch in samples: 41 words, odd -> row and column walks are conflict-free
constexpr int LR_IN_ROWS = 64 + 6;

struct LrUnitDev { // svtf_lr_unit, device copy
    int32_t type;
    int16_t vf[8], hf[8];
    int32_t ep, xqd[2];
};

struct LrArgs {
    Planes           cur, dbk, out;
    int              pw[3], ph[3];   // visible plane size
    int              cw[3], chh[3];  // coded plane size (what the surfaces hold; samples outside the visible size are copied)
    int              unit[3], hu[3], vu[3];
    const LrUnitDev *units[3];
    int              frame_type[3];
    int              opt, bd;                        // opt 2: no stripe boundaries (the encoder's filter search)
    int              chunks[3], stripes[3], base[3]; // linearised (plane, stripe, chunk) grid
    // SSE instantiation (try_restoration_unit_seg, Encoder/Codec/EbRestorationPick.c:129): nothing is written; the squared
    // error of the filtered samples against `src` is added to sse[plane][unit].  Units of type NONE are skipped.
    Planes              src;
    unsigned long long *sse[3];
    // FLT instantiation (apply_sgr + the sums of svt_get_proj_subspace_c, EbRestorationPick.c:522/:417): for units of type
    // SGRPROJ the two guided filters' outputs minus the sample (flt0 - u, flt1 - u, 4 fractional bits) are stored as int32 at
    // flt[plane][y * cur.pitch + x] and flt[plane][flt_half[plane] + ...], and sum f0^2, f1^2, f0 f1, f0 s, f1 s (s = source
    // - sample, same scale) are added to proj[plane][5 * unit + ...].
    int32_t   *flt[3];
    size_t     flt_half[3];
    long long *proj[3];
    int        force_ep[3]; // >= 0: the parameter set of every unit of the plane (one launch per candidate ep)
};
enum { LR_APPLY = 0, LR_SSE = 1, LR_FLT = 2 };

// svt_aom_eb_sgr_params (EbRestoration.c:85-103)
__constant__ int c_sgr_r[16][2] = {{2, 1}, {2, 1}, {2, 1}, {2, 1}, {2, 1}, {2, 1}, {2, 1}, {2, 1},
                                   {2, 1}, {2, 1}, {0, 1}, {0, 1}, {0, 1}, {0, 1}, {2, 0}, {2, 0}};
__constant__ int c_sgr_s[16][2] = {{140, 3236}, {112, 2158}, {93, 1618}, {80, 1438}, {70, 1295}, {58, 1177}, {47, 1079}, {37, 996},
                                   {30, 925},   {25, 863},   {-1, 2589}, {-1, 1618}, {-1, 1177}, {-1, 925},  {56, -1},   {22, -1}};

__device__ __forceinline__ int lr_rnd(int v, int n) { return (v + ((1 << n) >> 1)) >> n; } // ROUND_POWER_OF_TWO

// svt_aom_eb_x_by_xplus1[min(z, 255)] (:647-660): round(256 z / (z + 1)), 0 -> 1, 255 -> 256
__device__ __forceinline__ int lr_x_by_xplus1(unsigned z) {
    if (z == 0)
        return 1;
    if (z >= 255)
        return 256;
    return (int)((256u * z + (z + 1) / 2) / (z + 1));
}

// A and B of the self-guided filter (selfguided_restoration_[fast_]internal, EbRestoration.c:669/803) for LR_SEG consecutive
// positions of one row: the (2R+1)-row column sums of the LR_SEG + 2R columns are formed once in registers and the box
// sums slide over them — (2R+1)(LR_SEG+2R)/LR_SEG shared loads per position instead of (2R+1)^2.  All sums are exact
// 32-bit integers, so the order of the additions is free.  `in` points at the centre sample of the first position.
constexpr int LR_SEG = 6;
template <int R>
__device__ __forceinline__ void lr_sgr_ab_seg(const uint16_t *in, const uint16_t *xb, const int bd, const unsigned s, const int npos, uint16_t *A,
                                              uint32_t *B) {
    constexpr int NC = LR_SEG + 2 * R;
    uint32_t      cs[NC], cq[NC];
#pragma unroll
    for (int j = 0; j < NC; j++) {
        uint32_t a = 0, q = 0;
#pragma unroll
        for (int dy = -R; dy <= R; dy++) {
            const uint32_t v = in[dy * LR_IN_P + j - R];
            a += v, q += v * v;
        }
        cs[j] = a, cq[j] = q;
    }
    constexpr uint32_t n = (2 * R + 1) * (2 * R + 1), one_by = R == 2 ? 164 : 455; // svt_aom_eb_one_by_x[n - 1]
#pragma unroll
    for (int j = 0; j < LR_SEG; j++) {
        if (j >= npos)
            break;
        uint32_t sum = 0, sq = 0;
#pragma unroll
        for (int d = 0; d <= 2 * R; d++) sum += cs[j + d], sq += cq[j + d];
        const uint32_t a = (uint32_t)lr_rnd((int)sq, 2 * (bd - 8)), b = (uint32_t)lr_rnd((int)sum, bd - 8);
        const uint32_t p = (a * n < b * b) ? 0 : a * n - b * b;
        const uint32_t z = (p * s + (1u << 19)) >> 20;
        const int      Av = xb[min(z, 255u)];
        A[j]              = (uint16_t)Av;
        B[j]              = ((uint32_t)(256 - Av) * sum * one_by + (1u << 11)) >> 12;
    }
}

template <typename T, int MODE = LR_APPLY>
__global__ void __launch_bounds__(LR_THREADS) lr_kernel(const __grid_constant__ LrArgs a) {
    __shared__ __align__(16) uint16_t s_in[LR_IN_ROWS * LR_IN_P];
    __shared__ __align__(16) unsigned char s_buf[66 * 66 * 6]; // Wiener: horizontal pass (u16 [70][TW]); SGR: A (u16) and B (u32)
    __shared__ LrUnitDev s_unit;
    __shared__ uint16_t  s_xb[256];
    constexpr bool SSE = MODE == LR_SSE, FLT = MODE == LR_FLT;
    const int tid = threadIdx.x;
    int       plane = 0;
    if ((int)blockIdx.x >= a.base[2])
        plane = 2;
    else if ((int)blockIdx.x >= a.base[1])
        plane = 1;
    const int lin = blockIdx.x - a.base[plane], k = lin / a.chunks[plane], ch = lin - k * a.chunks[plane];
    const int ss = plane ? 1 : 0, TW = LR_MAXW >> ss, SH = 64 >> ss, off = 8 >> ss;
    const int sgs = 3 - ss;    // log2(8-column segments per row)
    const int pw = a.pw[plane], ph = a.ph[plane], cw = a.cw[plane], chh = a.chh[plane];
    // the last "stripe" of the grid is the tail: coded rows below the visible picture, copied
    const bool beyond = k == a.stripes[plane] - 1;
    const int  x0 = ch * TW, y0 = beyond ? ph : max(0, k * SH - off);
    const T  *cur = static_cast<const T *>(a.cur.p[plane]), *dbk = static_cast<const T *>(a.dbk.p[plane]);
    T        *out = static_cast<T *>(a.out.p[plane]);
    const int cp = a.cur.pitch[plane], dp = a.dbk.pitch[plane], op = a.out.pitch[plane];
    const int  y1 = beyond ? chh : min((k + 1) * SH - off, ph), h = y1 - y0, w = min(TW, cw - x0);
    if (h <= 0 || w <= 0)
        return;
    const bool filt = !beyond && a.frame_type[plane] != SVTF_RESTORE_NONE && x0 < pw;
    int        type = SVTF_RESTORE_NONE, unit_idx = 0;
    __shared__ unsigned long long s_sse;
    if (filt) {
        const int U = a.unit[plane];
        unit_idx    = min((y0 + off) / U, a.vu[plane] - 1) * a.hu[plane] + min(x0 / U, a.hu[plane] - 1);
        if (tid == 0)
            s_unit = a.units[plane][unit_idx], s_sse = 0;
        __syncthreads();
        type = s_unit.type;
    }
    if (MODE != LR_APPLY && type == SVTF_RESTORE_NONE)
        return;
    if (FLT && type != SVTF_RESTORE_SGRPROJ)
        return;
    const T  *srcp = MODE != LR_APPLY ? static_cast<const T *>(a.src.p[plane]) : nullptr;
    const int sp   = MODE != LR_APPLY ? a.src.pitch[plane] : 0;
    unsigned long long my_sse = 0;
    if (type == SVTF_RESTORE_NONE) { // copy (also: planes without restoration, coded samples outside the visible picture)
        // coded widths are multiples of 4 (chroma) / 8 (luma) and x0 of 32: 4-sample vectors
        for (int i = tid; i < h * (TW / 4); i += LR_THREADS) {
            const int r = i / (TW / 4), c = (i - r * (TW / 4)) * 4;
            if (c < w)
                st4<T>(out + (size_t)(y0 + r) * op + x0 + c, ld4<T>(cur + (size_t)(y0 + r) * cp + x0 + c));
        }
        return;
    }
    // ---- stage the stripe input: rows y0-3 .. y1+2, columns x0-8 .. x0+TW+7 (setup_processing_stripe_boundary :288) ----
    const bool above = k > 0, below = y1 < ph;
    const int  wv = min(w, pw - x0); // visible columns of this chunk
    s_xb[tid] = (uint16_t)lr_x_by_xplus1((unsigned)tid); // svt_aom_eb_x_by_xplus1 as a table: one entry per thread
    {
        const int nch = TW / 8 + 2;
        for (int i = tid; i < (h + 6) * nch; i += LR_THREADS) {
            const int rr = ss ? i / 6 : i / 10, j = i - rr * nch;
            const int r  = y0 - 3 + rr;
            const T  *src;
            if (a.opt == 2)
                src = cur + (size_t)min(max(r, 0), ph - 1) * cp;
            else if (r < y0) {
                if (!above)
                    src = cur + (size_t)max(r, 0) * cp;
                else if (a.opt)
                    src = cur + (size_t)(r == y0 - 3 ? y0 - 2 : r) * cp;
                else
                    src = dbk + (size_t)max(r, y0 - 2) * dp;
            } else if (r >= y1) {
                if (!below)
                    src = cur + (size_t)min(r, ph - 1) * cp;
                else if (a.opt)
                    src = cur + (size_t)min(r == y1 + 2 ? y1 + 1 : r, ph - 1) * cp;
                else
                    src = dbk + (size_t)min(min(r, y1 + 1), ph - 1) * dp;
            } else
                src = cur + (size_t)r * cp;
            const int gx = x0 - 8 + 8 * j;
            uint4     v;
            if (gx >= 0 && gx + 8 <= pw)
                v = ld8<T>(src + gx);
            else { // columns outside the plane repeat the edge column
                uint32_t e[8];
#pragma unroll
                for (int t = 0; t < 8; t++) e[t] = src[min(max(gx + t, 0), pw - 1)];
                v = make_uint4(e[0] | (e[1] << 16), e[2] | (e[3] << 16), e[4] | (e[5] << 16), e[6] | (e[7] << 16));
            }
            uint32_t *d = reinterpret_cast<uint32_t *>(s_in + rr * LR_IN_P + 8 * j);
            d[0] = v.x, d[1] = v.y, d[2] = v.z, d[3] = v.w;
        }
    }
    __syncthreads();
    const int bd = a.bd, maxv = (1 << bd) - 1;
    if (type == SVTF_RESTORE_WIENER) {
        // svt_av1_[highbd_]wiener_convolve_add_src_c: horizontal pass into 16 bits, then vertical.  A thread produces 8
        // consecutive outputs of a row (horizontal) / of a column (vertical) from 14 inputs held in registers.
        uint16_t *tmp = reinterpret_cast<uint16_t *>(s_buf); // [h + 6][TW]
        const int r0 = bd == 12 ? 5 : 3, r1 = 14 - r0, lim = (1 << (bd + 1 + 7 - r0)) - 1;
        int       hf[7], vf[7];
#pragma unroll
        for (int t = 0; t < 7; t++) hf[t] = s_unit.hf[t], vf[t] = s_unit.vf[t];
        for (int i = tid; i < ((h + 6) << sgs); i += LR_THREADS) {
            const int       rr = i >> sgs, c = (i & ((1 << sgs) - 1)) * 8;
            const uint16_t *p = &s_in[rr * LR_IN_P + LR_C0 + c - 3];
            int             x[14];
#pragma unroll
            for (int t = 0; t < 14; t++) x[t] = p[t];
            uint32_t o[4];
#pragma unroll
            for (int j = 0; j < 8; j++) {
                int sum = (x[j + 3] << 7) + (1 << (bd + 6));
#pragma unroll
                for (int t = 0; t < 7; t++) sum += x[j + t] * hf[t];
                const uint32_t v = (uint32_t)min(max(lr_rnd(sum, r0), 0), lim);
                o[j >> 1]        = (j & 1) ? (o[j >> 1] | (v << 16)) : v;
            }
            uint32_t *d = reinterpret_cast<uint32_t *>(tmp + rr * TW + c);
            d[0] = o[0], d[1] = o[1], d[2] = o[2], d[3] = o[3];
        }
        __syncthreads();
        const int nseg = (h + 7) >> 3;
        for (int i = tid; i < nseg * TW; i += LR_THREADS) {
            const int sg = i / TW, c = i - sg * TW, rb = sg * 8;
            if (c >= w)
                continue;
            T *o = out + (size_t)(y0 + rb) * op + x0 + c;
            if (c < wv) {
                int x[14], sv[8];
                if (SSE) { // the source samples first: their latency hides behind the filter arithmetic
                    const T *sq = srcp + (size_t)(y0 + rb) * sp + x0 + c;
#pragma unroll
                    for (int j = 0; j < 8; j++) sv[j] = rb + j < h ? (int)sq[j * sp] : 0;
                }
                // rows beyond h + 6 of the last segment are never-written shared memory inside s_buf: they only feed outputs
                // rb + j >= h, which are dropped
                const uint16_t *tp = tmp + rb * TW + c;
#pragma unroll
                for (int t = 0; t < 14; t++) x[t] = tp[t * TW];
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    if (rb + j >= h)
                        break;
                    int sum = (x[j + 3] << 7) - (1 << (bd + r1 - 1));
#pragma unroll
                    for (int t = 0; t < 7; t++) sum += x[j + t] * vf[t];
                    const int v = min(max(lr_rnd(sum, r1), 0), maxv);
                    if (SSE) {
                        const int d = v - sv[j];
                        my_sse += (unsigned)(d * d);
                    } else
                        o[(size_t)j * op] = (T)v;
                }
            } else if (!SSE) {
                for (int j = 0; j < 8 && rb + j < h; j++) o[(size_t)j * op] = cur[(size_t)(y0 + rb + j) * cp + x0 + c];
            }
        }
        if (SSE) {
#pragma unroll
            for (int m = 16; m; m >>= 1) my_sse += __shfl_xor_sync(0xffffffffu, my_sse, m);
            if (!(tid & 31))
                atomicAdd(&s_sse, my_sse);
            __syncthreads();
            if (tid == 0)
                atomicAdd(a.sse[plane] + unit_idx, s_sse);
        }
        return;
    }
    // ---- self-guided (svt_apply_selfguided_restoration_c :957) ----
    const int ep = FLT && a.force_ep[plane] >= 0 ? a.force_ep[plane] : s_unit.ep, rad0 = c_sgr_r[ep][0], rad1 = c_sgr_r[ep][1];
    int       xq0, xq1; // svt_decode_xq :634
    if (FLT)
        xq0 = xq1 = 1;
    else if (rad0 == 0)
        xq0 = 0, xq1 = 128 - s_unit.xqd[1];
    else if (rad1 == 0)
        xq0 = s_unit.xqd[0], xq1 = 0;
    else
        xq0 = s_unit.xqd[0], xq1 = 128 - xq0 - s_unit.xqd[1];
    // a thread owns two work items of 8 consecutive samples of a row (item i = tid + n*256: row i >> sgs, segment i & ..)
    int acc[2][8], acc_b[2][8]; // FLT keeps the two filters apart (acc: radius 2, acc_b: radius 1)
#pragma unroll
    for (int n = 0; n < 2; n++)
#pragma unroll
        for (int j = 0; j < 8; j++) acc[n][j] = acc_b[n][j] = 0;
    int(&acc1)[2][8] = FLT ? acc_b : acc;
    constexpr int AP = LR_MAXW + 2; // pitch of A / B: columns -1 .. TW
    uint16_t     *sA = reinterpret_cast<uint16_t *>(s_buf);
    uint32_t     *sB = reinterpret_cast<uint32_t *>(s_buf + 66 * 66 * 2);
    const int     nsg = (TW + 2 + LR_SEG - 1) / LR_SEG; // segments of LR_SEG positions per A/B row
    if (rad0 > 0) { // selfguided_restoration_fast_internal :669 — A/B on the rows -1, 1, 3, .. of the stripe
        const unsigned sv   = (unsigned)c_sgr_s[ep][0];
        const int      rows = (h + 1) / 2 + 1; // i = -1, 1, .., <= h
        for (int i = tid; i < rows * nsg; i += LR_THREADS) {
            const int ri = i / nsg, c = (i - ri * nsg) * LR_SEG; // positions c .. c+5 <-> stripe columns c-1 ..
            lr_sgr_ab_seg<2>(s_in + (3 + 2 * ri - 1) * LR_IN_P + LR_C0 + c - 1, s_xb, bd, sv, TW + 2 - c, sA + ri * AP + c, sB + ri * AP + c);
        }
        __syncthreads();
#pragma unroll
        for (int n = 0; n < 2; n++) {
            const int i = tid + n * LR_THREADS, r = i >> sgs, c = (i & ((1 << sgs) - 1)) * 8;
            if (r >= h)
                continue;
            const uint16_t *dgp = s_in + (r + 3) * LR_IN_P + LR_C0 + c;
            if (!(r & 1)) { // even row: rows r-1 and r+1 -> A rows r/2 and r/2+1
                const int q0 = (r >> 1) * AP + c, q1 = q0 + AP;
                int       as[10], bs[10]; // column sums of the two rows at positions c .. c+9
#pragma unroll
                for (int t = 0; t < 10; t++) as[t] = sA[q0 + t] + sA[q1 + t], bs[t] = (int)(sB[q0 + t] + sB[q1 + t]);
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    const int dg = dgp[j];
                    const int av = as[j + 1] * 6 + (as[j] + as[j + 2]) * 5, bv = bs[j + 1] * 6 + (bs[j] + bs[j + 2]) * 5;
                    acc[n][j] += xq0 * (lr_rnd(av * dg + bv, 9) - (dg << 4));
                }
            } else { // odd row r -> A row (r+1)/2
                const int q = ((r + 1) >> 1) * AP + c;
                int       as[10], bs[10];
#pragma unroll
                for (int t = 0; t < 10; t++) as[t] = sA[q + t], bs[t] = (int)sB[q + t];
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    const int dg = dgp[j];
                    const int av = as[j + 1] * 6 + (as[j] + as[j + 2]) * 5, bv = bs[j + 1] * 6 + (bs[j] + bs[j + 2]) * 5;
                    acc[n][j] += xq0 * (lr_rnd(av * dg + bv, 8) - (dg << 4));
                }
            }
        }
        __syncthreads();
    }
    if (rad1 > 0) { // selfguided_restoration_internal :803 — A/B on rows -1 .. h
        const unsigned sv = (unsigned)c_sgr_s[ep][1];
        for (int i = tid; i < (h + 2) * nsg; i += LR_THREADS) {
            const int ri = i / nsg, c = (i - ri * nsg) * LR_SEG;
            lr_sgr_ab_seg<1>(s_in + (3 + ri - 1) * LR_IN_P + LR_C0 + c - 1, s_xb, bd, sv, TW + 2 - c, sA + ri * AP + c, sB + ri * AP + c);
        }
        __syncthreads();
#pragma unroll
        for (int n = 0; n < 2; n++) {
            const int i = tid + n * LR_THREADS, r = i >> sgs, c = (i & ((1 << sgs) - 1)) * 8;
            if (r >= h)
                continue;
            const uint16_t *dgp = s_in + (r + 3) * LR_IN_P + LR_C0 + c;
            const int       q   = (r + 1) * AP + c; // position c <-> stripe column c-1
            int             a4[10], a3[10], b4[10], b3[10]; // per column: centre row (weight 4 / 3 pattern) and the rows above + below
#pragma unroll
            for (int t = 0; t < 10; t++) {
                const int am = sA[q + t], ao = sA[q - AP + t] + sA[q + AP + t];
                const int bm = (int)sB[q + t], bo = (int)(sB[q - AP + t] + sB[q + AP + t]);
                a4[t] = am, a3[t] = ao, b4[t] = bm, b3[t] = bo;
            }
#pragma unroll
            for (int j = 0; j < 8; j++) {
                const int dg = dgp[j];
                // (centre + left + right + up + down) * 4 + (four diagonals) * 3
                const int av = (a4[j + 1] + a4[j] + a4[j + 2] + a3[j + 1]) * 4 + (a3[j] + a3[j + 2]) * 3;
                const int bv = (b4[j + 1] + b4[j] + b4[j + 2] + b3[j + 1]) * 4 + (b3[j] + b3[j + 2]) * 3;
                acc1[n][j] += xq1 * (lr_rnd(av * dg + bv, 9) - (dg << 4));
            }
        }
    }
    if (FLT) {
        long long  sum[5] = {0, 0, 0, 0, 0};
        int32_t   *f0 = a.flt[plane], *f1 = a.flt[plane] + a.flt_half[plane];
#pragma unroll
        for (int n = 0; n < 2; n++) {
            const int i = tid + n * LR_THREADS, r = i >> sgs, c = (i & ((1 << sgs) - 1)) * 8;
            if (r >= h || c >= wv)
                continue;
            const uint16_t *dgp = s_in + (r + 3) * LR_IN_P + LR_C0 + c;
            const size_t    row = (size_t)(y0 + r) * cp + x0 + c;
#pragma unroll
            for (int j = 0; j < 8; j++) {
                if (c + j < wv) {
                    const int       d0 = acc[n][j], d1 = acc_b[n][j];
                    const long long sv = ((int)srcp[(size_t)(y0 + r) * sp + x0 + c + j] - (int)dgp[j]) << 4;
                    f0[row + j] = d0, f1[row + j] = d1;
                    sum[0] += (long long)d0 * d0, sum[1] += (long long)d1 * d1, sum[2] += (long long)d0 * d1;
                    sum[3] += d0 * sv, sum[4] += d1 * sv;
                }
            }
        }
        __shared__ unsigned long long s_proj[5];
        if (tid < 5)
            s_proj[tid] = 0;
        __syncthreads();
#pragma unroll
        for (int q = 0; q < 5; q++) {
#pragma unroll
            for (int m = 16; m; m >>= 1) sum[q] += __shfl_xor_sync(0xffffffffu, sum[q], m);
            if (!(tid & 31))
                atomicAdd(&s_proj[q], (unsigned long long)sum[q]);
        }
        __syncthreads();
        if (tid < 5)
            atomicAdd(reinterpret_cast<unsigned long long *>(a.proj[plane] + 5 * unit_idx + tid), s_proj[tid]);
        return;
    }
#pragma unroll
    for (int n = 0; n < 2; n++) {
        const int i = tid + n * LR_THREADS, r = i >> sgs, c = (i & ((1 << sgs) - 1)) * 8;
        if (r >= h || c >= w)
            continue;
        const uint16_t *dgp = s_in + (r + 3) * LR_IN_P + LR_C0 + c;
        uint32_t        o[4];
#pragma unroll
        for (int j = 0; j < 8; j++) {
            const int      dg   = dgp[j];
            const int      wv16 = (int)(short)lr_rnd((dg << 11) + acc[n][j], 11);
            const uint32_t v    = (uint32_t)min(max(wv16, 0), maxv);
            o[j >> 1]           = (j & 1) ? (o[j >> 1] | (v << 16)) : v;
            if (SSE && c + j < wv) {
                const int d = (int)v - (int)srcp[(size_t)(y0 + r) * sp + x0 + c + j];
                my_sse += (unsigned)(d * d);
            }
        }
        if (SSE)
            continue;
        T *op8 = out + (size_t)(y0 + r) * op + x0 + c;
        if (c + 8 <= wv)
            st8<T>(op8, make_uint4(o[0], o[1], o[2], o[3]));
        else {
            for (int j = 0; j < 8 && c + j < w; j++)
                op8[j] = c + j < wv ? (T)((o[j >> 1] >> (16 * (j & 1))) & 0xffff) : cur[(size_t)(y0 + r) * cp + x0 + c + j];
        }
    }
    if (SSE) {
#pragma unroll
        for (int m = 16; m; m >>= 1) my_sse += __shfl_xor_sync(0xffffffffu, my_sse, m);
        if (!(tid & 31))
            atomicAdd(&s_sse, my_sse);
        __syncthreads();
        if (tid == 0)
            atomicAdd(a.sse[plane] + unit_idx, s_sse);
    }
}

} // namespace svtf
