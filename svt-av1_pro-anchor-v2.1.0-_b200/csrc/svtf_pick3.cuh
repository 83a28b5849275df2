// svtf_pick3.cuh — finish_cdef_search in ONE cooperative launch spread over the whole device.
//
// The same 75 dependent svt_search_one_dual calls as svtf_pick.cuh (Encoder/Codec/EbEncCdef.c:627-925), with that file's split of a call
// (CTA (x, y) = rows [8x, 8x + 8) of tot over the y-th chunk of 128 filter blocks, 4 x 4 register tiles, VIADDMNMX for min(a + b, best)),
// but the CTAs stay resident for the whole schedule:
//   * the chunk's table rows are read ONCE into shared memory (svtf_pick.cuh stages 64 KB per CTA again for every call — most of a call's
//     8 us once launch latency is hidden);
//   * a call = best_so_far from the CTA's own copy of the greedy lists, accumulate, 512 atomics into tot[call % 3], ONE grid barrier,
//     then EVERY CTA reads the 32 KB of totals back from L2 and takes the argmin itself — the copies of the lists evolve identically, so
//     nothing is broadcast and there is no second barrier; tot[(call + 2) % 3] is cleared behind the barrier (its last readers have all
//     arrived), so clearing needs no barrier either;
//   * searched-block count, the width test, the RD bookkeeping and the per-filter-block assignment (pick_init_kernel, pick_argmin_body's
//     tail, pick_assign_kernel) are inside the same launch.
// The grid barrier is a monotonic counter in PickState.ticket (arrive: fence + atomicAdd by thread 0; wait: acquire loads); the launch is
// cooperative (cudaLaunchCooperativeKernel), so all CTAs are resident or the launch fails and the per-call form runs.  The wait is
// bounded: a CTA that polls for 2 s sets bit 1 of PickState.wide and every CTA leaves — no hang, the host runs the per-call form.
// Preconditions as for the cluster form: every entry the pick reads is below 2^31 (else bit 0 of wide: "not handled"), and the grid fits the
// device (4K: 128 CTAs of 97 KB; 8K: 256 CTAs with two row groups per CTA, RG = 2; 148 SMs x 2 CTAs is the limit).
// Partial sums are 32-bit when every entry the pick reads is below 2^26 (two instructions per entry and block instead of three).
// All results are the u64 / i64 integer arithmetic of the C code (first (j, k) in raster order wins ties).
#pragma once
#include "svtf_pick.cuh"

namespace svtf {

constexpr int PG_THREADS = 256;
constexpr int PG_CHUNK   = 128;                                 // filter blocks per CTA
constexpr int PG_ROWS    = 8;                                   // rows of tot per row group; a CTA covers RG of them (template)
constexpr int PG_SMEM    = PG_CHUNK * 128 * 4 + 8 * 512 * 8 + PG_CHUNK * 4; // rows (u32) + the warps' slabs (u64) + best_so_far

__device__ __forceinline__ unsigned pg_ld_acquire(const unsigned *p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// Returns false when the wait gave up (some CTA is missing or another one reported failure).
__device__ __forceinline__ bool pg_grid_barrier(PickState *st, const unsigned target, int *s_ok) {
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(&st->ticket, 1u);
        int                ok = 1;
        unsigned           spins = 0;
        unsigned long long t0 = 0;
        while (pg_ld_acquire(&st->ticket) < target) {
            if ((++spins & 4095) == 0) { // every few milliseconds of polling: never reached in a healthy run
                unsigned long long t;
                asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
                if (!t0)
                    t0 = t;
                if ((__ldcg(&st->wide) & 2) || t - t0 > 2000000000ull) { // another CTA gave up, or 2 s
                    atomicOr(&st->wide, 2);
                    ok = 0;
                    break;
                }
            }
        }
        __threadfence();
        *s_ok = ok;
    }
    __syncthreads();
    return *s_ok != 0;
}

// The accumulate step of one warp: filter blocks [f_lo, f_hi) of the chunk into the warp's 8 x 64 slab (rows jr .., 4 x 4 per thread).
// ACC = uint32_t when every entry the pick reads is below 2^26 (a + b < 2^27, at most 32 blocks per warp: no overflow), else uint64_t.
template <typename ACC>
__device__ __forceinline__ void pg_accumulate(const uint32_t (*srow)[128], const uint32_t *sbest, const int f_lo, const int f_hi, const int jr,
                                              const int tj, const int tk, uint64_t *slab) {
    ACC acc[4][4];
#pragma unroll
    for (int r = 0; r < 4; r++)
#pragma unroll
        for (int c = 0; c < 4; c++) acc[r][c] = 0;
#pragma unroll 2
    for (int f = f_lo; f < f_hi; f++) {
        const uint4    av  = *reinterpret_cast<const uint4 *>(&srow[f][jr + tj * 4]);
        const uint4    bv  = *reinterpret_cast<const uint4 *>(&srow[f][64 + tk * 4]);
        const unsigned bst = sbest[f];
        const unsigned aa[4] = {av.x, av.y, av.z, av.w}, bb[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
        for (int r = 0; r < 4; r++)
#pragma unroll
            for (int c = 0; c < 4; c++) acc[r][c] += __viaddmin_u32(aa[r], bb[c], bst);
    }
#pragma unroll
    for (int r = 0; r < 4; r++) {
        ulonglong2 *d = reinterpret_cast<ulonglong2 *>(&slab[(tj * 4 + r) * 64 + tk * 4]);
        d[0] = make_ulonglong2(acc[r][0], acc[r][1]), d[1] = make_ulonglong2(acc[r][2], acc[r][3]);
    }
}

// RG: row groups (of 8 rows of tot) per CTA.  RG = 1: grid (end_gi / 8) x chunks, a warp takes 16 filter blocks; RG = 2: half as many
// CTAs, a warp takes 32 filter blocks of one of the two row groups — for tables whose RG = 1 grid does not fit the device (8K).
template <int RG>
__global__ void __launch_bounds__(PG_THREADS, 2) pick_grid_kernel(const PickArgs a, const int64_t lambda, int8_t *fb_idx) {
    extern __shared__ __align__(16) unsigned char pg_dyn[];
    uint32_t(*srow)[128] = reinterpret_cast<uint32_t(*)[128]>(pg_dyn);                                      // [fb][luma 64 | chroma 64]
    uint64_t *red        = reinterpret_cast<uint64_t *>(pg_dyn + PG_CHUNK * 128 * 4);                       // [8 warps][8 x 64]
    uint32_t *sbest      = reinterpret_cast<uint32_t *>(pg_dyn + PG_CHUNK * 128 * 4 + 8 * 512 * 8);         // [fb]
    __shared__ uint8_t  s_skip[PG_CHUNK];
    __shared__ int      s_lev0[8], s_lev1[8], s_ok, s_pick[2];
    __shared__ uint64_t s_v[PG_THREADS / 32];
    __shared__ int      s_i[PG_THREADS / 32];
    constexpr int  FGW = 8 / RG; // warps (= subsets of the chunk's filter blocks) per row group
    const int      tid = threadIdx.x, jt = blockIdx.x * PG_ROWS * RG, fb0 = blockIdx.y * PG_CHUNK;
    const unsigned nct = gridDim.x * gridDim.y, bid = blockIdx.y * gridDim.x + blockIdx.x;
    PickState     *st  = a.st;

    // ---- once: the chunk's rows (whole rows: best_so_far reads whichever entries the lists name), width test, searched-block count ----
    unsigned wide = 0; // bit 0: an entry >= 2^31 (not handled), bit 2: an entry >= 2^26 (64-bit partial sums)
#pragma unroll
    for (int n0 = 0; n0 < PG_CHUNK * 64 / PG_THREADS; n0 += 8) {
        ulonglong2 v[8];
#pragma unroll
        for (int n = 0; n < 8; n++) {
            const int i = tid + (n0 + n) * PG_THREADS, f = i >> 6, e = i & 63, fb = min(fb0 + f, a.nfb - 1);
            v[n] = __ldg(reinterpret_cast<const ulonglong2 *>(a.mse + ((size_t)(e >> 5) * a.nfb + fb) * 64) + (e & 31));
        }
#pragma unroll
        for (int n = 0; n < 8; n++) {
            const int i = tid + (n0 + n) * PG_THREADS, f = i >> 6, e = i & 63;
            if (a.bias && (e & 31) == 0)
                v[n].x = ((uint64_t)(uint16_t)a.bias * v[n].x) >> 6; // entry 0 of a row, EbEncCdef.c:776-789
            const int  s0   = 2 * (e & 31); // strength index of v.x within its plane
            const bool live = fb0 + f < a.nfb && !a.skip[min(fb0 + f, a.nfb - 1)];
            const uint64_t m = live ? ((s0 < a.end_gi ? v[n].x : 0) | (s0 + 1 < a.end_gi ? v[n].y : 0)) : 0;
            wide |= (m >> 31 ? 1u : 0u) | (m >> 26 ? 4u : 0u);
            *reinterpret_cast<uint2 *>(&srow[f][2 * e]) = make_uint2((uint32_t)v[n].x, (uint32_t)v[n].y);
        }
    }
    int cnt = 0;
    if (tid < PG_CHUNK) {
        const uint8_t s = fb0 + tid < a.nfb ? a.skip[fb0 + tid] : (uint8_t)1;
        s_skip[tid]     = s;
        cnt             = !s;
    }
    cnt = __syncthreads_count(cnt);
    {
        const int w1 = __syncthreads_or((int)(wide & 1)), w4 = __syncthreads_or((int)(wide & 4));
        if ((w1 || w4) && tid == 0)
            atomicOr(&st->wide, (w1 ? 1 : 0) | (w4 ? 4 : 0));
    }
    if (tid == 0 && blockIdx.x == 0 && cnt)
        atomicAdd(&st->count, cnt);
    for (unsigned i = bid * PG_THREADS + tid; i < 3u * 64 * 64; i += nct * PG_THREADS) a.tot[i] = 0;
    unsigned target = nct;
    if (!pg_grid_barrier(st, target, &s_ok))
        return;
    const int flags = __ldcg(&st->wide); // uniform over the grid: every CTA's flags landed before the barrier, nobody clears them
    if (flags & 1)
        return;
    const bool small = !(flags & 4);
    const int count = __ldcg(&st->count);

    uint64_t tot_mse = PICK_INF, best_cost = PICK_INF; // thread 0's bookkeeping, identical in every CTA
    int      best_bits = 0, out0[8], out1[8];
#pragma unroll
    for (int g = 0; g < 8; g++) out0[g] = out1[g] = 0;
    if (tid < 8)
        s_lev0[tid] = s_lev1[tid] = 0;
    __syncthreads();

    const int lane = tid & 31, warp = tid >> 5, tj = lane >> 4, tk = lane & 15;
    int       call = 0;
    for (int bits = 0; bits <= 3; bits++) {
        const int nbs = 1 << bits;
        for (int step = 0; step < 5 * nbs; step++, call++) {
            // joint_strength_search_dual: nbs greedy picks (list entries 0..step-1 fixed), then 4*nbs refinements (drop the oldest entry,
            // re-pick the last)
            const bool     refine = step >= nbs;
            const int      nb = refine ? nbs - 1 : step, rot = refine ? 1 : 0, pos = refine ? nbs - 1 : step;
            uint64_t      *tot = a.tot + (size_t)(call % 3) * 64 * 64;
            if (tid < PG_CHUNK) {
                uint32_t best = 0;
                if (!s_skip[tid]) {
                    best = 0xffffffffu;
                    for (int g = 0; g < nb; g++) best = min(best, srow[tid][s_lev0[min(g + rot, 7)]] + srow[tid][64 + s_lev1[min(g + rot, 7)]]);
                }
                sbest[tid] = best;
            }
            __syncthreads();
            {
                const int rg = warp / FGW, fg = warp - rg * FGW, jr = jt + rg * PG_ROWS;
                if (small)
                    pg_accumulate<uint32_t>(srow, sbest, fg * (PG_CHUNK / FGW), (fg + 1) * (PG_CHUNK / FGW), jr, tj, tk, red + warp * 512);
                else
                    pg_accumulate<uint64_t>(srow, sbest, fg * (PG_CHUNK / FGW), (fg + 1) * (PG_CHUNK / FGW), jr, tj, tk, red + warp * 512);
            }
            __syncthreads();
#pragma unroll
            for (int h = 0; h < 2 * RG; h++) {
                const int e = tid + h * PG_THREADS, rg = e >> 9, i = e & 511, j = jt + rg * PG_ROWS + (i >> 6), k = i & 63;
                uint64_t  v = 0;
#pragma unroll
                for (int w = 0; w < FGW; w++) v += red[(rg * FGW + w) * 512 + i];
                if (j < a.end_gi && k < a.end_gi && v)
                    atomicAdd(reinterpret_cast<unsigned long long *>(&tot[j * 64 + k]), (unsigned long long)v);
            }
            target += nct;
            if (!pg_grid_barrier(st, target, &s_ok))
                return;
            // ---- every CTA: argmin of the totals (svt_search_one_dual's scan: strict "<" from 1 << 63, first in raster order) ----
            uint64_t      bv = ~(uint64_t)0;
            int           bi = 1 << 30;
            constexpr int PER = 64 * 64 / PG_THREADS;
            {
                uint64_t tv[PER];
#pragma unroll
                for (int n = 0; n < PER; n++) tv[n] = __ldcg(&tot[tid + n * PG_THREADS]);
#pragma unroll
                for (int n = 0; n < PER; n++) {
                    const int i = tid + n * PG_THREADS, j = i >> 6, k = i & 63;
                    if (j < a.end_gi && k < a.end_gi && (tv[n] < bv || (tv[n] == bv && i < bi)))
                        bv = tv[n], bi = i;
                }
            }
            // the buffer of call + 2 was last read behind the barrier of call - 1: everyone is past that
            {
                uint64_t *z = a.tot + (size_t)((call + 2) % 3) * 64 * 64;
                for (unsigned i = bid * PG_THREADS + tid; i < 64u * 64; i += nct * PG_THREADS) z[i] = 0;
            }
#pragma unroll
            for (int m = 16; m > 0; m >>= 1) {
                const uint64_t v = __shfl_xor_sync(0xffffffffu, bv, m);
                const int      i = __shfl_xor_sync(0xffffffffu, bi, m);
                if (v < bv || (v == bv && i < bi))
                    bv = v, bi = i;
            }
            if (lane == 0)
                s_v[warp] = bv, s_i[warp] = bi;
            __syncthreads();
            if (tid == 0) {
                for (int w = 1; w < PG_THREADS / 32; w++)
                    if (s_v[w] < bv || (s_v[w] == bv && s_i[w] < bi))
                        bv = s_v[w], bi = s_i[w];
                int      b0 = 0, b1 = 0;
                uint64_t best = PICK_INF;
                if (bv < PICK_INF)
                    best = bv, b0 = bi >> 6, b1 = bi & 63;
                if (rot)
                    for (int j = 0; j < pos; j++) s_lev0[j] = s_lev0[j + 1], s_lev1[j] = s_lev1[j + 1];
                s_lev0[pos] = b0, s_lev1[pos] = b1;
                tot_mse     = best;
                if (step == 5 * nbs - 1) { // EbEncCdef.c:845-876
                    const int      total_bits = count * bits + nbs * 6 * 2;
                    const int64_t  rate       = (int64_t)total_bits * 512; // av1_cost_literal
                    const uint64_t dist       = tot_mse * 16;
                    const uint64_t cost       = (uint64_t)(((rate * lambda + 256) >> 9) + (int64_t)dist * 128); // RDCOST
                    if (cost < best_cost) {
                        best_cost = cost, best_bits = bits;
                        for (int j = 0; j < nbs; j++) out0[j] = s_lev0[j], out1[j] = s_lev1[j];
                    }
                }
            }
            __syncthreads();
        }
    }
    // ---- per-filter-block index into the chosen set (EbEncCdef.c:884-905), by the CTAs of row group 0 from their own rows ----
    if (tid == 0) {
        s_pick[0] = best_bits;
        for (int g = 0; g < 8; g++) s_lev0[g] = out0[g], s_lev1[g] = out1[g];
        if (bid == 0) {
            st->best_bits = best_bits, st->best_cost = best_cost, st->tot_mse = tot_mse;
            for (int g = 0; g < 8; g++) st->out0[g] = out0[g], st->out1[g] = out1[g];
        }
    }
    __syncthreads();
    if (blockIdx.x == 0 && tid < PG_CHUNK && fb0 + tid < a.nfb) {
        int8_t r = -1;
        if (!s_skip[tid]) {
            const int nb  = 1 << s_pick[0];
            uint32_t  bst = 0xffffffffu;
            int       bg  = 0;
            for (int g = 0; g < nb; g++) {
                const uint32_t c = srow[tid][s_lev0[g]] + srow[tid][64 + s_lev1[g]];
                if (c < bst)
                    bst = c, bg = g;
            }
            r = (int8_t)bg;
        }
        fb_idx[fb0 + tid] = r;
    }
}

// Row groups per CTA for a table of nfb filter blocks: 1, 2 (the whole grid must fit the device at once) or 0 = does not fit.
// (cached per process: one device kind)
inline int pick3_fits(int nfb, int end_gi, int sm_count) {
    static int per_sm = -1;
    if (per_sm < 0) {
        int n1 = 0, n2 = 0;
        if (cudaFuncSetAttribute(pick_grid_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, PG_SMEM) != cudaSuccess ||
            cudaFuncSetAttribute(pick_grid_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, PG_SMEM) != cudaSuccess ||
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n1, pick_grid_kernel<1>, PG_THREADS, PG_SMEM) != cudaSuccess ||
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n2, pick_grid_kernel<2>, PG_THREADS, PG_SMEM) != cudaSuccess)
            n1 = n2 = 0;
        per_sm = n1 < n2 ? n1 : n2;
    }
    if (per_sm <= 0)
        return 0;
    const long chunks = (nfb + PG_CHUNK - 1) / PG_CHUNK, cap = (long)per_sm * sm_count;
    for (int rg = 1; rg <= 2; rg++)
        if ((long)((end_gi + PG_ROWS * rg - 1) / (PG_ROWS * rg)) * chunks <= cap)
            return rg;
    return 0;
}

// One cooperative launch of the whole pick on `stream` (rg from pick3_fits); a.tot must hold 3 x 64 x 64 entries.  PickState.wide
// afterwards: bit 0 = wide table, bit 1 = barrier gave up (either: not handled, run the per-call form); bit 2 is informational
// (64-bit partial sums were used).
inline cudaError_t pick3_enqueue(const PickArgs &a, int rg, int8_t *d_fb_idx, int64_t lambda, cudaStream_t stream, uint64_t *launches) {
    cudaError_t e = cudaMemsetAsync(a.st, 0, sizeof(PickState), stream);
    if (e != cudaSuccess)
        return e;
    PickArgs aa    = a;
    int64_t  lam   = lambda;
    int8_t  *idx   = d_fb_idx;
    void    *args[] = {&aa, &lam, &idx};
    const dim3 grid((a.end_gi + PG_ROWS * rg - 1) / (PG_ROWS * rg), (a.nfb + PG_CHUNK - 1) / PG_CHUNK);
    const void *fn = rg == 1 ? reinterpret_cast<const void *>(pick_grid_kernel<1>) : reinterpret_cast<const void *>(pick_grid_kernel<2>);
    e = cudaLaunchCooperativeKernel(fn, grid, dim3(PG_THREADS), args, PG_SMEM, stream);
    if (e == cudaSuccess && launches)
        *launches += 1;
    return e;
}

} // namespace svtf
