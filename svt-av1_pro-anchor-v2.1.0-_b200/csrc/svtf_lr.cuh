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
constexpr int LR_IN_P    = LR_MAXW + 6 + 2;  // shared input pitch
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
    int              opt, bd;
    int              chunks[3], stripes[3], base[3]; // linearised (plane, stripe, chunk) grid
};

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

// A and B of the self-guided filter from the box sums around (r, c) of the staged input
template <int R>
__device__ __forceinline__ void lr_sgr_ab(const uint16_t *in, const uint16_t *xb, const int r, const int c, const int bd, const unsigned s,
                                          uint16_t &A, uint32_t &B) {
    uint32_t sum = 0, sq = 0;
#pragma unroll
    for (int dy = -R; dy <= R; dy++)
#pragma unroll
        for (int dx = -R; dx <= R; dx++) {
            const uint32_t v = in[(r + dy) * LR_IN_P + c + dx];
            sum += v, sq += v * v;
        }
    constexpr uint32_t n = (2 * R + 1) * (2 * R + 1), one_by = R == 2 ? 164 : 455; // svt_aom_eb_one_by_x[n - 1]
    const uint32_t     a = (uint32_t)lr_rnd((int)sq, 2 * (bd - 8)), b = (uint32_t)lr_rnd((int)sum, bd - 8);
    const uint32_t     p = (a * n < b * b) ? 0 : a * n - b * b;
    const uint32_t     z = (p * s + (1u << 19)) >> 20;
    const int          Av = xb[min(z, 255u)];
    A                     = (uint16_t)Av;
    B                     = ((uint32_t)(256 - Av) * sum * one_by + (1u << 11)) >> 12;
}

template <typename T>
__global__ void __launch_bounds__(LR_THREADS) lr_kernel(const __grid_constant__ LrArgs a) {
    __shared__ uint16_t s_in[LR_IN_ROWS * LR_IN_P];
    __shared__ __align__(16) unsigned char s_buf[66 * 66 * 6]; // Wiener: horizontal pass (u16 [70][64]); SGR: A (u16) and B (u32)
    __shared__ LrUnitDev s_unit;
    __shared__ uint16_t  s_xb[256];
    const int tid = threadIdx.x;
    int       plane = 0;
    if ((int)blockIdx.x >= a.base[2])
        plane = 2;
    else if ((int)blockIdx.x >= a.base[1])
        plane = 1;
    const int lin = blockIdx.x - a.base[plane], k = lin / a.chunks[plane], ch = lin - k * a.chunks[plane];
    const int ss = plane ? 1 : 0, TW = LR_MAXW >> ss, tws = 6 - ss, SH = 64 >> ss, off = 8 >> ss;
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
    int        type = SVTF_RESTORE_NONE;
    if (filt) {
        const int U = a.unit[plane];
        if (tid == 0)
            s_unit = a.units[plane][min((y0 + off) / U, a.vu[plane] - 1) * a.hu[plane] + min(x0 / U, a.hu[plane] - 1)];
        __syncthreads();
        type = s_unit.type;
    }
    if (type == SVTF_RESTORE_NONE) { // copy (also: planes without restoration, coded samples outside the visible picture)
        for (int r = tid >> 5; r < h; r += LR_THREADS / 32)
            for (int c = tid & 31; c < w; c += 32) out[(size_t)(y0 + r) * op + x0 + c] = cur[(size_t)(y0 + r) * cp + x0 + c];
        return;
    }
    // ---- stage the stripe input: rows y0-3 .. y1+2, columns x0-3 .. x0+w+2 (setup_processing_stripe_boundary :288) ----
    const bool above = k > 0, below = y1 < ph;
    const int  wv = min(w, pw - x0); // visible columns of this chunk
    const int lane = tid & 31, wrp = tid >> 5;
    if (tid < 256) // svt_aom_eb_x_by_xplus1 as a table: one entry per thread
        s_xb[tid] = (uint16_t)lr_x_by_xplus1((unsigned)tid);
    for (int rr = wrp; rr < h + 6; rr += LR_THREADS / 32) { // one source row per warp pass: the row decision is made once
        const int r = y0 - 3 + rr;
        const T  *src;
        if (r < y0) {
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
        for (int cc = lane; cc < TW + 6; cc += 32) s_in[rr * LR_IN_P + cc] = (uint16_t)src[min(max(x0 - 3 + cc, 0), pw - 1)];
    }
    __syncthreads();
    const int bd = a.bd, maxv = (1 << bd) - 1;
    if (type == SVTF_RESTORE_WIENER) {
        // svt_av1_[highbd_]wiener_convolve_add_src_c: horizontal pass into 16 bits, then vertical
        uint16_t *tmp = reinterpret_cast<uint16_t *>(s_buf); // [h + 6][TW]
        const int r0 = bd == 12 ? 5 : 3, r1 = 14 - r0, lim = (1 << (bd + 1 + 7 - r0)) - 1;
        int       hf[7], vf[7];
#pragma unroll
        for (int t = 0; t < 7; t++) hf[t] = s_unit.hf[t], vf[t] = s_unit.vf[t];
        for (int i = tid; i < (h + 6) * TW; i += LR_THREADS) {
            const int       rr = i >> tws, c = i & (TW - 1);
            const uint16_t *p = &s_in[rr * LR_IN_P + c];
            int             sum = ((int)p[3] << 7) + (1 << (bd + 6));
#pragma unroll
            for (int t = 0; t < 7; t++) sum += (int)p[t] * hf[t];
            tmp[rr * TW + c] = (uint16_t)min(max(lr_rnd(sum, r0), 0), lim);
        }
        __syncthreads();
        for (int i = tid; i < h * TW; i += LR_THREADS) {
            const int r = i >> tws, c = i & (TW - 1);
            if (c >= w)
                continue;
            int v;
            if (c < wv) {
                int sum = ((int)tmp[(r + 3) * TW + c] << 7) - (1 << (bd + r1 - 1));
#pragma unroll
                for (int t = 0; t < 7; t++) sum += (int)tmp[(r + t) * TW + c] * vf[t];
                v = min(max(lr_rnd(sum, r1), 0), maxv);
            } else
                v = cur[(size_t)(y0 + r) * cp + x0 + c];
            out[(size_t)(y0 + r) * op + x0 + c] = (T)v;
        }
        return;
    }
    // ---- self-guided (svt_apply_selfguided_restoration_c :957) ----
    const int ep = s_unit.ep, rad0 = c_sgr_r[ep][0], rad1 = c_sgr_r[ep][1];
    int       xq0, xq1; // svt_decode_xq :634
    if (rad0 == 0)
        xq0 = 0, xq1 = 128 - s_unit.xqd[1];
    else if (rad1 == 0)
        xq0 = s_unit.xqd[0], xq1 = 0;
    else
        xq0 = s_unit.xqd[0], xq1 = 128 - xq0 - s_unit.xqd[1];
    constexpr int PER = 64 * 64 / LR_THREADS;
    int           acc[PER];
#pragma unroll
    for (int n = 0; n < PER; n++) acc[n] = 0;
    constexpr int AP = LR_MAXW + 2; // pitch of A / B: columns -1 .. TW
    uint16_t     *sA = reinterpret_cast<uint16_t *>(s_buf);
    uint32_t     *sB = reinterpret_cast<uint32_t *>(s_buf + 66 * 66 * 2);
    if (rad0 > 0) { // selfguided_restoration_fast_internal :669 — A/B on the rows -1, 1, 3, .. of the stripe
        const unsigned sv   = (unsigned)c_sgr_s[ep][0];
        const int      rows = (h + 1) / 2 + 1; // i = -1, 1, .., <= h
        for (int ri = wrp; ri < rows; ri += LR_THREADS / 32)
            for (int c = lane; c < TW + 2; c += 32)
                lr_sgr_ab<2>(s_in, s_xb, 3 + (2 * ri - 1), 3 + (c - 1), bd, sv, sA[ri * AP + c], sB[ri * AP + c]);
        __syncthreads();
#pragma unroll
        for (int n = 0; n < PER; n++) {
            const int i = tid + n * LR_THREADS, r = i >> tws, c = i & (TW - 1);
            if (i >= h * TW)
                continue;
            const int dg = s_in[(r + 3) * LR_IN_P + c + 3], cc = c + 1;
            int       av, bv, sh;
            if (!(r & 1)) { // even row: rows r-1 and r+1 -> A rows r/2 and r/2+1
                const int q0 = (r >> 1) * AP, q1 = q0 + AP;
                av = (sA[q0 + cc] + sA[q1 + cc]) * 6 + (sA[q0 + cc - 1] + sA[q1 + cc - 1] + sA[q0 + cc + 1] + sA[q1 + cc + 1]) * 5;
                bv = (int)((sB[q0 + cc] + sB[q1 + cc]) * 6 + (sB[q0 + cc - 1] + sB[q1 + cc - 1] + sB[q0 + cc + 1] + sB[q1 + cc + 1]) * 5);
                sh = 9;
            } else { // odd row r -> A row (r+1)/2
                const int q = ((r + 1) >> 1) * AP;
                av = sA[q + cc] * 6 + (sA[q + cc - 1] + sA[q + cc + 1]) * 5;
                bv = (int)(sB[q + cc] * 6 + (sB[q + cc - 1] + sB[q + cc + 1]) * 5);
                sh = 8;
            }
            acc[n] += xq0 * (lr_rnd(av * dg + bv, sh) - (dg << 4));
        }
        __syncthreads();
    }
    if (rad1 > 0) { // selfguided_restoration_internal :803 — A/B on rows -1 .. h
        const unsigned sv = (unsigned)c_sgr_s[ep][1];
        for (int ri = wrp; ri < h + 2; ri += LR_THREADS / 32)
            for (int c = lane; c < TW + 2; c += 32)
                lr_sgr_ab<1>(s_in, s_xb, 3 + (ri - 1), 3 + (c - 1), bd, sv, sA[ri * AP + c], sB[ri * AP + c]);
        __syncthreads();
#pragma unroll
        for (int n = 0; n < PER; n++) {
            const int i = tid + n * LR_THREADS, r = i >> tws, c = i & (TW - 1);
            if (i >= h * TW)
                continue;
            const int dg = s_in[(r + 3) * LR_IN_P + c + 3], q = (r + 1) * AP + c + 1;
            const int av = (sA[q] + sA[q - 1] + sA[q + 1] + sA[q - AP] + sA[q + AP]) * 4 +
                           (sA[q - 1 - AP] + sA[q - 1 + AP] + sA[q + 1 - AP] + sA[q + 1 + AP]) * 3;
            const int bv = (int)((sB[q] + sB[q - 1] + sB[q + 1] + sB[q - AP] + sB[q + AP]) * 4 +
                                 (sB[q - 1 - AP] + sB[q - 1 + AP] + sB[q + 1 - AP] + sB[q + 1 + AP]) * 3);
            acc[n] += xq1 * (lr_rnd(av * dg + bv, 9) - (dg << 4));
        }
    }
#pragma unroll
    for (int n = 0; n < PER; n++) {
        const int i = tid + n * LR_THREADS, r = i >> tws, c = i & (TW - 1);
        if (i >= h * TW || c >= w)
            continue;
        int v;
        if (c < wv) {
            const int dg = s_in[(r + 3) * LR_IN_P + c + 3];
            const int wv16 = (int)(short)lr_rnd((dg << 11) + acc[n], 11);
            v              = min(max(wv16, 0), maxv);
        } else
            v = cur[(size_t)(y0 + r) * cp + x0 + c];
        out[(size_t)(y0 + r) * op + x0 + c] = (T)v;
    }
}

} // namespace svtf
