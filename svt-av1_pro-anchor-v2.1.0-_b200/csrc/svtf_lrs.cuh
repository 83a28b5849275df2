// svtf_lrs.cuh — device half of the loop-restoration filter search: per-unit sums, the Wiener statistics, and the projection
// error of the self-guided coefficient search (the guided-filter outputs themselves: lr_kernel<T, LR_FLT>, svtf_lr.cuh).
//
// Reference: svt_av1_compute_stats_c / _highbd_c (Encoder/Codec/EbRestorationPick.c:671 / :708), find_average[_highbd]
// (EbRestorationPick.h:24-43), search_norestore_seg :1432 (sse_restoration_unit :56).  The trial filtering + SSE of
// finer_tile_search_wiener_seg is lr_kernel<T, true> (svtf_lr.cuh).
//
// Statistics of one restoration unit: with y = the win x win window of (picture - avg) around a sample (index = column offset
// * win + row offset) and x = source - avg,  M[k] = sum y[k] x,  H[k][l] = sum y[k] y[l] over the unit's samples; window
// samples outside the picture repeat its edge (svt_extend_frame), samples outside the unit are the picture's own.
//
// The reference runs win^2 (win^2 + 1) / 2 + win^2 multiply-adds per sample out of a 49-entry array.  Here H is cut into
// win x win blocks (kb, lb) = (column offset of y[k], column offset of y[l]); a lane owns one block of the upper triangle
// (28 / 15 / 6 blocks for win 7 / 5 / 3 -> 1 / 2 / 5 lane groups per warp) and walks 8 samples DOWN a column: the two
// vertical strips of 8 + win - 1 picture samples it needs (columns c + kb and c + lb) are read once into registers, and the
// 8 x win x win products accumulate in 32-bit registers (flushed to 64-bit shared accumulators before they can overflow).
// Lanes of a diagonal block also accumulate M for their column.  One CTA = one 64 x 64 tile of one unit; its totals are
// added to the unit's 64-bit accumulators in global memory (the host divides by the bit-depth divider and mirrors H).
#pragma once
#include "svtf_kernels.cuh"
#include "svtf_lr.cuh"

namespace svtf {

struct LrsTile { // a tile of at most 64 x 64 samples inside one restoration unit (plane coordinates)
    int16_t x0, y0, w, h;
    int32_t unit; // index into the per-call unit arrays (all searched planes concatenated)
    int32_t plane;
};

constexpr int LRS_THREADS = 256;
constexpr int LRS_T       = 64;

struct LrsArgs {
    Planes                    cur, src;
    int                       pw[3], ph[3];
    const LrsTile            *tiles;
    unsigned long long       *unit_sum; // sum of the picture's samples per unit (find_average)
    unsigned long long       *unit_sse; // squared error picture - source per unit (RESTORE_NONE)
    const int                *unit_count; // samples per unit
    long long                *acc;        // statistics kernel: per unit win^4 (H, upper triangle) + win^2 (M) accumulators
    const unsigned long long *acc_off;    // offset of each unit's accumulators in `acc`
    int                       bd;
};

template <typename T>
__global__ void __launch_bounds__(LRS_THREADS) lrs_sums_kernel(const __grid_constant__ LrsArgs a) {
    const LrsTile t   = a.tiles[blockIdx.x];
    const T      *cur = static_cast<const T *>(a.cur.p[t.plane]), *src = static_cast<const T *>(a.src.p[t.plane]);
    const int     cp = a.cur.pitch[t.plane], sp = a.src.pitch[t.plane];
    unsigned long long sum = 0, sse = 0;
    for (int i = threadIdx.x; i < t.w * t.h; i += LRS_THREADS) {
        const int r = i / t.w, c = i - r * t.w;
        const int v = cur[(size_t)(t.y0 + r) * cp + t.x0 + c], d = v - (int)src[(size_t)(t.y0 + r) * sp + t.x0 + c];
        sum += (unsigned)v, sse += (unsigned)(d * d);
    }
#pragma unroll
    for (int m = 16; m; m >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, m), sse += __shfl_xor_sync(0xffffffffu, sse, m);
    if (!(threadIdx.x & 31)) {
        atomicAdd(a.unit_sum + t.unit, sum);
        atomicAdd(a.unit_sse + t.unit, sse);
    }
}

#ifdef SVTF_AB_KERNELS // superseded kernel, kept for A/B measurements only (not in the product build)
template <typename T, int WIN>
__global__ void __launch_bounds__(LRS_THREADS) lrs_stats_kernel(const __grid_constant__ LrsArgs a) {
    constexpr int WIN2 = WIN * WIN, HW = WIN / 2, NB = WIN * (WIN + 1) / 2, G = 32 / NB, NACC = WIN2 * WIN2 + WIN2;
    constexpr int DP = LRS_T + 2 * HW + 2; // pitch of the staged picture window (int16)
    __shared__ int16_t   s_d[(LRS_T + 2 * HW) * DP];
    __shared__ int16_t   s_x[LRS_T * LRS_T];
    __shared__ long long s_acc[NACC];
    const LrsTile t   = a.tiles[blockIdx.x];
    const int     tid = threadIdx.x, plane = t.plane, pw = a.pw[plane], ph = a.ph[plane];
    const T      *cur = static_cast<const T *>(a.cur.p[plane]), *src = static_cast<const T *>(a.src.p[plane]);
    const int     cp = a.cur.pitch[plane], sp = a.src.pitch[plane];
    const int     avg = (int)(a.unit_sum[t.unit] / (unsigned long long)a.unit_count[t.unit]);
    for (int i = tid; i < NACC; i += LRS_THREADS) s_acc[i] = 0;
    for (int i = tid; i < (t.h + 2 * HW) * (t.w + 2 * HW); i += LRS_THREADS) {
        const int r = i / (t.w + 2 * HW), c = i - r * (t.w + 2 * HW);
        const int y = min(max(t.y0 + r - HW, 0), ph - 1), x = min(max(t.x0 + c - HW, 0), pw - 1);
        s_d[r * DP + c] = (int16_t)((int)cur[(size_t)y * cp + x] - avg);
    }
    for (int i = tid; i < t.h * t.w; i += LRS_THREADS) {
        const int r = i / t.w, c = i - r * t.w;
        s_x[r * LRS_T + c] = (int16_t)((int)src[(size_t)(t.y0 + r) * sp + t.x0 + c] - avg);
    }
    __syncthreads();
    const int lane = tid & 31, warp = tid >> 5, g = lane / NB, blk = lane - g * NB;
    if (g < G) {
        int kb = 0, lb = blk; // block (kb, lb), kb <= lb: row kb of the upper triangle holds WIN - kb blocks
        while (lb >= WIN - kb) lb -= WIN - kb, kb++;
        lb += kb;
        const bool diag = kb == lb;
        int        acc[WIN][WIN], accm[WIN];
#pragma unroll
        for (int r1 = 0; r1 < WIN; r1++) {
            accm[r1] = 0;
#pragma unroll
            for (int r2 = 0; r2 < WIN; r2++) acc[r1][r2] = 0;
        }
        auto flush = [&]() {
#pragma unroll
            for (int r1 = 0; r1 < WIN; r1++) {
#pragma unroll
                for (int r2 = 0; r2 < WIN; r2++) {
                    if ((!diag || r1 <= r2) && acc[r1][r2])
                        atomicAdd(reinterpret_cast<unsigned long long *>(&s_acc[(kb * WIN + r1) * WIN2 + lb * WIN + r2]),
                                  (unsigned long long)(long long)acc[r1][r2]);
                    acc[r1][r2] = 0;
                }
                if (diag && accm[r1])
                    atomicAdd(reinterpret_cast<unsigned long long *>(&s_acc[WIN2 * WIN2 + kb * WIN + r1]), (unsigned long long)(long long)accm[r1]);
                accm[r1] = 0;
            }
        };
        // 32-bit partial sums: |y| < 2^bd, so 2^(31 - 2 bd) products fit; a column segment adds 8
        const int flush_every = a.bd > 10 ? 8 : (a.bd > 8 ? 128 : 2048);
        const int nbands = (t.h + 7) >> 3, nitems = nbands * t.w;
        int       since = 0;
        for (int it = warp * G + g; it < nitems; it += (LRS_THREADS / 32) * G) {
            const int band = it / t.w, c = it - band * t.w, r0 = band * 8;
            int       ya[8 + WIN - 1], yb[8 + WIN - 1];
#pragma unroll
            for (int k = 0; k < 8 + WIN - 1; k++) {
                const int rr = min(r0 + k, t.h + 2 * HW - 1);
                ya[k] = s_d[rr * DP + c + kb], yb[k] = s_d[rr * DP + c + lb];
            }
#pragma unroll
            for (int j = 0; j < 8; j++) {
                if (r0 + j < t.h) {
                    const int x = s_x[(r0 + j) * LRS_T + c];
#pragma unroll
                    for (int r1 = 0; r1 < WIN; r1++) {
#pragma unroll
                        for (int r2 = 0; r2 < WIN; r2++) acc[r1][r2] += ya[j + r1] * yb[j + r2];
                        accm[r1] += ya[j + r1] * x;
                    }
                }
            }
            if (++since >= flush_every)
                flush(), since = 0;
        }
        flush();
    }
    __syncthreads();
    long long *out = a.acc + a.acc_off[t.unit];
    for (int i = tid; i < NACC; i += LRS_THREADS)
        if (s_acc[i])
            atomicAdd(reinterpret_cast<unsigned long long *>(out + i), (unsigned long long)s_acc[i]);
}
#endif // SVTF_AB_KERNELS

// ---- statistics, second form: one product image per window-offset DIFFERENCE --------------------------------------------
// H[(c1,r1),(c2,r2)] = sum over the tile of d(p + o1) d(p + o2) depends on o1 and on delta = o2 - o1 only: with
// P_delta(q) = d(q) d(q + delta) it is the sum of P_delta over the tile shifted by o1.  The upper triangle needs the
// (WIN - 1)(2 WIN - 1) + WIN differences with dc > 0, or dc = 0 and dr >= 0 (85 for a 7x7 window), each with (WIN - |dr|)(WIN - dc)
// shifts: 1 225 entries from 85 x 70 x 70 products per 64 x 64 tile instead of 4 096 x 1 225.
//   stage 1, item = (delta, row qy of the extended tile): the row's products once, R(qy, c1) = sum over the w columns starting
//            at c1, slid over c1 (one product in, one out); 32-bit, to shared memory (a batch of LRS_DB differences at a time);
//   stage 2, item = (delta, c1): T(r1, c1) = sum of R over the h rows starting at r1, slid over r1, added to the unit's
//            64-bit accumulator of entry ((c1, r1), (c1 + dc, r1 + dr)).
// M (49 products per sample) is accumulated directly.  Same exact integers as lrs_stats_kernel (kept for reference and
// selectable with SVTF_LRS_STATS_V1=1).
constexpr int LRS_DB = 11;
template <typename T, int WIN>
__global__ void __launch_bounds__(LRS_THREADS, 4) lrs_stats2_kernel(const __grid_constant__ LrsArgs a) {
    constexpr int WIN2 = WIN * WIN, HW = WIN / 2, ND = (WIN - 1) * (2 * WIN - 1) + WIN, ER = LRS_T + 2 * HW;
    constexpr int DP = ((ER / 2) & 1) ? ER : ER + 2; // odd pitch in 32-bit words
    __shared__ int16_t   s_d[ER * DP];
    __shared__ int16_t   s_x[LRS_T * LRS_T];
    __shared__ int       s_R[LRS_DB * ER * WIN];
    __shared__ long long s_M[WIN2];
    const LrsTile t   = a.tiles[blockIdx.x];
    const int     tid = threadIdx.x, plane = t.plane, pw = a.pw[plane], ph = a.ph[plane], w = t.w, h = t.h, eh = h + 2 * HW, ew = w + 2 * HW;
    const T      *cur = static_cast<const T *>(a.cur.p[plane]), *src = static_cast<const T *>(a.src.p[plane]);
    const int     cp = a.cur.pitch[plane], sp = a.src.pitch[plane];
    const int     avg = (int)(a.unit_sum[t.unit] / (unsigned long long)a.unit_count[t.unit]);
    long long    *out = a.acc + a.acc_off[t.unit];
    if (tid < WIN2)
        s_M[tid] = 0;
    for (int i = tid; i < eh * ew; i += LRS_THREADS) {
        const int r = i / ew, c = i - r * ew;
        const int y = min(max(t.y0 + r - HW, 0), ph - 1), x = min(max(t.x0 + c - HW, 0), pw - 1);
        s_d[r * DP + c] = (int16_t)((int)cur[(size_t)y * cp + x] - avg);
    }
    for (int i = tid; i < h * w; i += LRS_THREADS) {
        const int r = i / w, c = i - r * w;
        s_x[r * LRS_T + c] = (int16_t)((int)src[(size_t)(t.y0 + r) * sp + t.x0 + c] - avg);
    }
    __syncthreads();
    { // M[c * WIN + r] = sum d(py + r, px + c) x(py, px): at most 16 samples per thread, |product| < 2^24
        // one window column at a time: WIN accumulators live, so the kernel stays at 4+ CTAs per SM
        for (int cc = 0; cc < WIN; cc++) {
            int m[WIN];
#pragma unroll
            for (int rr = 0; rr < WIN; rr++) m[rr] = 0;
            for (int i = tid; i < h * w; i += LRS_THREADS) {
                const int py = i / w, px = i - py * w, x = s_x[py * LRS_T + px];
#pragma unroll
                for (int rr = 0; rr < WIN; rr++) m[rr] += (int)s_d[(py + rr) * DP + px + cc] * x;
            }
#pragma unroll
            for (int rr = 0; rr < WIN; rr++) {
                int v = m[rr]; // a warp's total: 32 * 16 products < 2^24 * 2^9 overflows 32 bits at 12-bit -> reduce in 64
                long long v64 = v;
#pragma unroll
                for (int q = 16; q; q >>= 1) v64 += __shfl_xor_sync(0xffffffffu, v64, q);
                if (!(tid & 31) && v64)
                    atomicAdd(reinterpret_cast<unsigned long long *>(&s_M[cc * WIN + rr]), (unsigned long long)v64);
            }
        }
    }
    for (int d0 = 0; d0 < ND; d0 += LRS_DB) {
        const int nd = min(LRS_DB, ND - d0);
        __syncthreads(); // s_R of the previous batch consumed (first pass: s_M zeroed, tiles staged)
        for (int it = tid; it < nd * eh; it += LRS_THREADS) {
            const int qy = it / nd, dl = it - qy * nd, di = d0 + dl;
            const int dc = di < WIN ? 0 : 1 + (di - WIN) / (2 * WIN - 1), dr = di < WIN ? di : (di - WIN) % (2 * WIN - 1) - (WIN - 1);
            int      *R = s_R + (dl * ER + qy) * WIN;
            if (qy + dr < 0 || qy + dr >= eh) {
#pragma unroll
                for (int c1 = 0; c1 < WIN; c1++) R[c1] = 0;
                continue;
            }
            const int16_t *pa = s_d + qy * DP, *pb = s_d + (qy + dr) * DP + dc;
            int            sum = 0;
#pragma unroll 8
            for (int qx = 0; qx < w; qx++) sum += (int)pa[qx] * pb[qx];
            R[0] = sum;
            for (int c1 = 1; c1 < WIN - dc; c1++) {
                sum += (int)pa[c1 + w - 1] * pb[c1 + w - 1] - (int)pa[c1 - 1] * pb[c1 - 1];
                R[c1] = sum;
            }
        }
        __syncthreads();
        for (int it = tid; it < nd * WIN; it += LRS_THREADS) {
            const int dl = it / WIN, c1 = it - dl * WIN, di = d0 + dl;
            const int dc = di < WIN ? 0 : 1 + (di - WIN) / (2 * WIN - 1), dr = di < WIN ? di : (di - WIN) % (2 * WIN - 1) - (WIN - 1);
            if (c1 >= WIN - dc)
                continue;
            const int  lo = max(0, -dr), hi = min(WIN - 1, WIN - 1 - dr);
            const int *R = s_R + dl * ER * WIN + c1;
            long long  sum = 0;
#pragma unroll 8
            for (int qy = lo; qy < lo + h; qy++) sum += R[qy * WIN];
            for (int r1 = lo; r1 <= hi; r1++) {
                if (sum)
                    atomicAdd(reinterpret_cast<unsigned long long *>(out + (c1 * WIN + r1) * WIN2 + (c1 + dc) * WIN + r1 + dr), (unsigned long long)sum);
                if (r1 < hi)
                    sum += R[(r1 + h) * WIN] - R[r1 * WIN];
            }
        }
    }
    __syncthreads();
    if (tid < WIN2 && s_M[tid])
        atomicAdd(reinterpret_cast<unsigned long long *>(out + WIN2 * WIN2 + tid), (unsigned long long)s_M[tid]);
}

// get_pixel_proj_error (Encoder/Codec/EbRestorationPick.c:305; svt_av1_[lowbd|highbd]_pixel_proj_error_c :167/:233) for every
// unit that has candidates pending: from the stored filter outputs (lr_kernel<T, LR_FLT>),
// v = (u << 7) + xq0 (flt0 - u) + xq1 (flt1 - u), e = round(v, 11) - source (no clipping, as the reference), sum e^2 per unit.
// A unit brings up to LRS_CAND candidate coefficient pairs per launch (the host speculates along the reference's line
// search: the candidates it will ask for if each one lowers the error); they share the 12 B read per sample.
// One CTA per 64 x 64 tile.
constexpr int LRS_CAND = 8;
struct LrsCand { // per unit
    int32_t n, ep;
    int32_t xqd[LRS_CAND][2];
};
struct LrsProjArgs {
    Planes              cur, src;
    const LrsTile      *tiles;
    const LrsCand      *cand; // [slot][unit]: all searched planes concatenated, indexed by LrsTile::unit
    const int32_t      *flt[3]; // slot 0; slot k at + k * slot_elems
    size_t              flt_half[3], slot_elems;
    int                 units;
    unsigned long long *err; // [slot][unit][LRS_CAND]
};

template <typename T>
__global__ void __launch_bounds__(LRS_THREADS) lrs_proj_err_kernel(const __grid_constant__ LrsProjArgs a) {
    const LrsTile  t = a.tiles[blockIdx.x];
    const size_t   ci = (size_t)blockIdx.y * a.units + t.unit; // (slot = one candidate ep, unit)
    const LrsCand &cd = a.cand[ci];
    const int      n = cd.n;
    if (n <= 0)
        return;
    const int ep = cd.ep, r0 = c_sgr_r[ep][0], r1 = c_sgr_r[ep][1];
    int       xq0[LRS_CAND], xq1[LRS_CAND]; // svt_decode_xq
#pragma unroll
    for (int j = 0; j < LRS_CAND; j++) {
        const int q0 = cd.xqd[j][0], q1 = cd.xqd[j][1];
        if (r0 == 0)
            xq0[j] = 0, xq1[j] = 128 - q1;
        else if (r1 == 0)
            xq0[j] = q0, xq1[j] = 0;
        else
            xq0[j] = q0, xq1[j] = 128 - q0 - q1;
    }
    const T       *cur = static_cast<const T *>(a.cur.p[t.plane]), *src = static_cast<const T *>(a.src.p[t.plane]);
    const int      cp = a.cur.pitch[t.plane], sp = a.src.pitch[t.plane];
    const int32_t *f0 = a.flt[t.plane] + blockIdx.y * a.slot_elems, *f1 = f0 + a.flt_half[t.plane];
    // a thread sees at most 64 * 64 / 256 = 16 samples and |e| < 2^13 (unclipped 12-bit): 32-bit partial sums
    unsigned err32[LRS_CAND];
#pragma unroll
    for (int j = 0; j < LRS_CAND; j++) err32[j] = 0;
    // tiles are at most 64 wide: a thread keeps its column and walks every fourth row (no division per sample).  Issuing the
    // loads of four rows together was measured slower (12.5 against 11.4 ms for the whole self-guided search).
    const int c = threadIdx.x & (LRS_T - 1);
    for (int r = threadIdx.x / LRS_T; r < t.h && c < t.w; r += LRS_THREADS / LRS_T) {
        const size_t o = (size_t)(t.y0 + r) * cp + t.x0 + c;
        const int    uu = (int)cur[o] << 4, d0 = f0[o], d1 = f1[o], sv = (int)src[(size_t)(t.y0 + r) * sp + t.x0 + c];
#pragma unroll
        for (int j = 0; j < LRS_CAND; j++) {
            if (j < n) { // uniform over the CTA
                const int v = (uu << 7) + xq0[j] * d0 + xq1[j] * d1;
                const int e = ((v + (1 << 10)) >> 11) - sv;
                err32[j] += (unsigned)(e * e);
            }
        }
    }
    unsigned long long err[LRS_CAND];
#pragma unroll
    for (int j = 0; j < LRS_CAND; j++) {
        err[j] = err32[j];
#pragma unroll
        for (int m = 16; m; m >>= 1) err[j] += __shfl_xor_sync(0xffffffffu, err[j], m);
        if (!(threadIdx.x & 31) && j < n)
            atomicAdd(a.err + ci * LRS_CAND + j, err[j]);
    }
}

} // namespace svtf
