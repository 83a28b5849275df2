// svtf_pick.cuh — frame-level CDEF strength pick on the device.
//
// finish_cdef_search (Encoder/Codec/EbEncCdef.c:728-925) with joint_strength_search_dual (:697-727) and
// svt_search_one_dual_c (:627-684): for cdef_bits = 0..3 pick nb = 1<<bits (luma, chroma) strength pairs greedily,
// refine 4*nb times, keep the bits with the lowest RD cost, then give every filter block the best pair of the set.
// One svt_search_one_dual call is tot[j][k] = sum over searched filter blocks of min(best_so_far, mse0[fb][j] +
// mse1[fb][k]) followed by an argmin: nfb * 64 * 64 u64 min/adds, 75 calls per picture — a serial tail of ~1 s at
// 8K on one host thread (and under pcs->cdef_search_mutex in the reference, EbCdefProcess.c:406).  Here each call is
// one launch (accumulate; the last CTA to finish does the argmin + list update) and the whole schedule is enqueued
// without a host round trip:
// the greedy lists live in device memory.  All arithmetic is u64/i64 integer, identical to the C code.
#pragma once
#include <cstdint>
#include <type_traits>
#include <cuda_runtime.h>

namespace svtf {

constexpr int      PICK_THREADS = 256;
constexpr uint64_t PICK_INF     = (uint64_t)1 << 63;

struct PickState {
    int      lev0[8], lev1[8]; // greedy list being built (indices into the strength lists)
    uint64_t tot_mse;
    int      count; // searched (non-skipped) filter blocks
    uint64_t best_cost;
    int      best_bits;
    int      out0[8], out1[8]; // lists of the best cdef_bits so far
    unsigned ticket;           // CTAs of the running accumulate launch that have finished
    int      wide;             // some table entry is >= 2^31: accumulate with 64-bit sums (else 32-bit, same totals)
};

struct PickArgs {
    const uint64_t *mse; // [2][nfb][64]
    const uint8_t  *skip;
    int             nfb, end_gi;
    uint32_t        bias; // zero_fs_cost_bias: entry 0 of every row is read as (bias * v) >> 6 (EbEncCdef.c:776-789)
    PickState      *st;
    uint64_t       *tot; // [64*64], zero on entry of every accumulate launch
};

__device__ __forceinline__ uint64_t pick_load(const PickArgs &a, int pl, int fb, int i) {
    const uint64_t v = a.mse[((size_t)pl * a.nfb + fb) * 64 + i];
    return (a.bias && i == 0) ? ((uint64_t)(uint16_t)a.bias * v) >> 6 : v;
}

// State reset (CTA 0), count of searched filter blocks and the width test of the table.  The state is zeroed by a memset
// ahead of this launch; the fields below are written by disjoint threads / atomics.
__global__ void __launch_bounds__(PICK_THREADS) pick_init_kernel(const PickArgs a) {
    int      c = 0;
    unsigned w = 0;
    for (int fb = blockIdx.x; fb < a.nfb; fb += gridDim.x) {
        if (a.skip[fb])
            continue;
        c += threadIdx.x == 0;
        if (threadIdx.x < 128) {
            const int pl = threadIdx.x >> 6, i = threadIdx.x & 63;
            if (i < a.end_gi)
                w |= pick_load(a, pl, fb, i) >> 31 ? 1u : 0u;
        }
    }
    if (__syncthreads_or((int)w) && threadIdx.x == 0)
        atomicOr(&a.st->wide, 1);
    if (threadIdx.x == 0 && c)
        atomicAdd(&a.st->count, c);
    if (blockIdx.x == 0) {
        for (int i = threadIdx.x; i < 64 * 64; i += PICK_THREADS) a.tot[i] = 0;
        if (threadIdx.x == 0) {
            PickState &s = *a.st;
            s.tot_mse = PICK_INF, s.best_cost = PICK_INF;
        }
    }
}

// svt_search_one_dual_c, argmin part (first (j,k) in raster order wins ties) + list update + re-zeroing of tot.
// pos: list slot that receives the pick; rot: rotate the list first; bits >= 0: this was the last call for that
// cdef_bits -> RD cost and best-so-far update (EbEncCdef.c:845-876).  Runs in the last CTA of an accumulate launch.
constexpr int PT_THREADS = 256;       // 8 warps = 8 subsets of the CTA's filter blocks; a warp covers the CTA's 8 x 64 slab of tot
constexpr int PT_CHUNK_N = 128;       // filter blocks per CTA, narrow table (32-bit rows in shared memory)
constexpr int PT_CHUNK_W = 64;        // filter blocks per CTA, wide table (64-bit rows): the same 64 KB
constexpr int PT_ROWS    = 8;         // rows of tot (luma strengths) per CTA
constexpr int PT_SMEM    = 64 * 1024 + PT_CHUNK_N * 8; // table rows of the chunk + best_so_far

__device__ __forceinline__ void pick_argmin_body(const PickArgs &a, const int pos, const int rot, const int bits, const int64_t lambda,
                                                 uint64_t *s_v, int *s_i) {
    const int     tid = threadIdx.x;
    uint64_t      bv  = ~(uint64_t)0;
    int           bi  = 1 << 30;
    constexpr int PER = 64 * 64 / PT_THREADS;
#pragma unroll
    for (int h = 0; h < 2; h++) { // two batches of 16 loads in flight (other CTAs' atomics: read at L2)
        uint64_t tv[PER / 2];
#pragma unroll
        for (int n = 0; n < PER / 2; n++) tv[n] = __ldcg(&a.tot[tid + (h * (PER / 2) + n) * PT_THREADS]);
#pragma unroll
        for (int n = 0; n < PER / 2; n++) {
            const int i = tid + (h * (PER / 2) + n) * PT_THREADS, j = i >> 6, k = i & 63;
            a.tot[i] = 0;
            if (j < a.end_gi && k < a.end_gi && (tv[n] < bv || (tv[n] == bv && i < bi)))
                bv = tv[n], bi = i;
        }
    }
#pragma unroll
    for (int m = 16; m > 0; m >>= 1) {
        const uint64_t v = __shfl_xor_sync(0xffffffffu, bv, m);
        const int      i = __shfl_xor_sync(0xffffffffu, bi, m);
        if (v < bv || (v == bv && i < bi))
            bv = v, bi = i;
    }
    if ((tid & 31) == 0)
        s_v[tid >> 5] = bv, s_i[tid >> 5] = bi;
    __syncthreads();
    if (tid == 0) {
        for (int w = 1; w < PT_THREADS / 32; w++)
            if (s_v[w] < bv || (s_v[w] == bv && s_i[w] < bi))
                bv = s_v[w], bi = s_i[w];
        PickState &s = *a.st;
        // the scan in the C code starts from best = 1<<63 with a strict "<": totals >= 1<<63 never win and (0,0) stays
        int      b0 = 0, b1 = 0;
        uint64_t best = PICK_INF;
        if (bv < PICK_INF)
            best = bv, b0 = bi >> 6, b1 = bi & 63;
        if (rot)
            for (int j = 0; j < pos; j++) s.lev0[j] = s.lev0[j + 1], s.lev1[j] = s.lev1[j + 1];
        s.lev0[pos] = b0, s.lev1[pos] = b1;
        s.tot_mse = best;
        if (bits >= 0) {
            const int      nbs        = 1 << bits;
            const int      total_bits = s.count * bits + nbs * 6 * 2;
            const int64_t  rate       = (int64_t)total_bits * 512; // av1_cost_literal
            const uint64_t dist       = best * 16;
            const uint64_t cost       = (uint64_t)(((rate * lambda + 256) >> 9) + (int64_t)dist * 128); // RDCOST
            if (cost < s.best_cost) {
                s.best_cost = cost, s.best_bits = bits;
                for (int j = 0; j < nbs; j++) s.out0[j] = s.lev0[j], s.out1[j] = s.lev1[j];
            }
        }
        s.ticket = 0;
    }
}

// One call for the CTA's chunk of PT_CHUNK_N filter blocks.  NARROW: every table entry is below 2^31 (PickState.wide == 0), rows are
// kept as 32-bit words and min(a + b, best) is one VIADDMNMX; else 64-bit rows, staged and accumulated in two halves of PT_CHUNK_W
// blocks through the same 64 KB (the second half is staged after the dependency: the rare case pays for it).  Totals are u64 either way.
template <bool NARROW>
__device__ __forceinline__ void pick_tot_body(const PickArgs &a, const int nb, const int rot, unsigned char *pt_dyn) {
    typedef typename std::conditional<NARROW, uint32_t, uint64_t>::type E;
    constexpr int CH = NARROW ? PT_CHUNK_N : PT_CHUNK_W;
    E(*srow)[128]    = reinterpret_cast<E(*)[128]>(pt_dyn); // [fb][plane 0: 64 strengths | plane 1: 64 strengths]
    E *sbest         = reinterpret_cast<E *>(pt_dyn + 64 * 1024);
    const int tid = threadIdx.x, jt = blockIdx.x * PT_ROWS;
    uint64_t  acc[4][4];
#pragma unroll
    for (int r = 0; r < 4; r++)
#pragma unroll
        for (int c = 0; c < 4; c++) acc[r][c] = 0;
#pragma unroll 1
    for (int half = 0; half < PT_CHUNK_N / CH; half++) {
        const int fb0 = blockIdx.y * PT_CHUNK_N + half * CH;
        if (half)
            __syncthreads(); // the first half's rows have been read
        // rows of the chunk, skipped or not (a skipped block gets best = 0 below and then adds min(0, .) = 0); the whole luma row is
        // staged, not only this CTA's 8 entries: best_so_far reads whichever entries the greedy lists name
        constexpr int VPT = CH * 64 / PT_THREADS; // 16-byte vectors per thread
#pragma unroll
        for (int n0 = 0; n0 < VPT; n0 += 8) {
            ulonglong2 v[8];
#pragma unroll
            for (int n = 0; n < 8; n++) {
                const int i = tid + (n0 + n) * PT_THREADS, f = i >> 6, e = i & 63, fb = min(fb0 + f, a.nfb - 1);
                v[n] = __ldg(reinterpret_cast<const ulonglong2 *>(a.mse + ((size_t)(e >> 5) * a.nfb + fb) * 64) + (e & 31));
            }
#pragma unroll
            for (int n = 0; n < 8; n++) {
                const int i = tid + (n0 + n) * PT_THREADS, f = i >> 6, e = i & 63;
                if (a.bias && (e & 31) == 0)
                    v[n].x = ((uint64_t)(uint16_t)a.bias * v[n].x) >> 6; // entry 0 of a row, EbEncCdef.c:776-789
                if (NARROW)
                    *reinterpret_cast<uint2 *>(&srow[f][2 * e]) = make_uint2((uint32_t)v[n].x, (uint32_t)v[n].y);
                else
                    *reinterpret_cast<ulonglong2 *>(&srow[f][2 * e]) = v[n];
            }
        }
        uint8_t skp = 1;
        if (tid < CH && fb0 + tid < a.nfb)
            skp = a.skip[fb0 + tid];
        if (!half)
            cudaGridDependencySynchronize(); // the previous call's list update and the re-zeroed tot are visible from here on
        __syncthreads();
        if (tid < CH) {
            E best = 0;
            if (!skp) {
                best = NARROW ? (E)0xffffffffu : (E)PICK_INF;
#pragma unroll
                for (int g = 0; g < 8; g++)
                    if (g < nb)
                        best = min(best, (E)(srow[tid][a.st->lev0[min(g + rot, 7)]] + srow[tid][64 + a.st->lev1[min(g + rot, 7)]]));
            }
            sbest[tid] = best;
        }
        __syncthreads();
        const int lane = tid & 31, f0 = (tid >> 5) * (CH / 8), tj = lane >> 4, tk = lane & 15;
#pragma unroll 2
        for (int f = f0; f < f0 + CH / 8; f++) {
            E av[4], bv[4];
#pragma unroll
            for (int r = 0; r < 4; r++) av[r] = srow[f][jt + tj * 4 + r], bv[r] = srow[f][64 + tk * 4 + r];
            const E bst = sbest[f];
#pragma unroll
            for (int r = 0; r < 4; r++)
#pragma unroll
                for (int c = 0; c < 4; c++) {
                    if (NARROW)
                        acc[r][c] += __viaddmin_u32((unsigned)av[r], (unsigned)bv[c], (unsigned)bst);
                    else
                        acc[r][c] += min((uint64_t)bst, (uint64_t)av[r] + (uint64_t)bv[c]);
                }
        }
    }
    cudaTriggerProgrammaticLaunchCompletion(); // the next call may launch and stage its rows now
    // the eight warps' partial slabs -> one 8 x 64 slab (through the rows' shared memory), then one atomic per entry and CTA
    __syncthreads();
    uint64_t *red = reinterpret_cast<uint64_t *>(pt_dyn); // [8 warps][8 * 64]
    {
        const int w = tid >> 5, lane = tid & 31, tj = lane >> 4, tk = lane & 15;
#pragma unroll
        for (int r = 0; r < 4; r++) {
            ulonglong2 *d = reinterpret_cast<ulonglong2 *>(&red[w * 512 + (tj * 4 + r) * 64 + tk * 4]);
            d[0] = make_ulonglong2(acc[r][0], acc[r][1]), d[1] = make_ulonglong2(acc[r][2], acc[r][3]);
        }
    }
    __syncthreads();
#pragma unroll
    for (int h = 0; h < 2; h++) {
        const int i = tid + h * PT_THREADS, j = jt + (i >> 6), k = i & 63;
        uint64_t  v = 0;
#pragma unroll
        for (int w = 0; w < 8; w++) v += red[w * 512 + i];
        if (j < a.end_gi && k < a.end_gi && v)
            atomicAdd(reinterpret_cast<unsigned long long *>(&a.tot[j * 64 + k]), (unsigned long long)v);
    }
}

// svt_search_one_dual_c, accumulation part: tot[j][k] += min(best_so_far(fb), mse0[fb][j] + mse1[fb][k]) over the searched filter
// blocks.  nb entries of the list are fixed; `rot` = 1 reads them one position further (the refinement step drops entry 0,
// EbEncCdef.c:717-722).  CTA (x, y): rows [8x, 8x + 8) of tot over the y-th chunk of filter blocks; warp w takes an eighth of the chunk
// and covers the 8 x 64 slab with 4 x 4 register tiles (a filter block costs a thread three shared loads and 16 min-adds); the warps'
// slabs are summed through shared memory, so a CTA issues 512 atomics per call (round 1: 2 048 from a 32-block chunk — the atomics
// and the fence behind them were most of a call's 13 us).
//
// The 75 calls of a picture are a dependent chain, so the latency of one call is what counts:
//  * the table rows of the chunk do not depend on the previous call: they are staged (16-byte loads, all in flight) BEFORE
//    cudaGridDependencySynchronize(), and every launch carries the programmatic-stream-serialization attribute, so launch latency and
//    staging overlap the previous call's atomics / argmin tail (64 KB of rows per CTA: the next call's CTA fits beside the running one);
//  * after the dependency only the 16 list entries are read from global memory; best_so_far comes from shared memory.
// The CTA that finishes last (ticket counter, no waiting) runs the argmin + list update of the same call.
__global__ void __launch_bounds__(PT_THREADS) pick_tot_kernel(const PickArgs a, const int nb, const int rot, const int pos, const int bits,
                                                              const int64_t lambda) {
    extern __shared__ __align__(16) unsigned char pt_dyn[];
    __shared__ int s_last;
    const int      tid = threadIdx.x;
    // PickState.wide is written by pick_init_kernel, which has completed before the first call starts (it never triggers early)
    const bool narrow = !a.st->wide;
    const bool active = blockIdx.x * PT_ROWS < a.end_gi;
    if (active) {
        if (narrow)
            pick_tot_body<true>(a, nb, rot, pt_dyn);
        else
            pick_tot_body<false>(a, nb, rot, pt_dyn);
    } else
        cudaGridDependencySynchronize();
    __threadfence();
    __syncthreads();
    if (tid == 0)
        s_last = atomicAdd(&a.st->ticket, 1u) == gridDim.x * gridDim.y - 1;
    __syncthreads();
    if (s_last) {
        __threadfence();
        pick_argmin_body(a, pos, rot, bits, lambda, reinterpret_cast<uint64_t *>(pt_dyn), reinterpret_cast<int *>(pt_dyn + 1024));
    }
}

// per-filter-block index into the chosen set (EbEncCdef.c:884-905); skipped blocks -> -1
__global__ void __launch_bounds__(PICK_THREADS) pick_assign_kernel(const PickArgs a, int8_t *fb_idx) {
    const int fb = blockIdx.x * PICK_THREADS + threadIdx.x;
    if (fb >= a.nfb)
        return;
    if (a.skip[fb]) {
        fb_idx[fb] = -1;
        return;
    }
    const PickState &s  = *a.st;
    const int        nb = 1 << s.best_bits;
    uint64_t         best = PICK_INF;
    int              bg = 0;
    for (int g = 0; g < nb; g++) {
        const uint64_t c = pick_load(a, 0, fb, s.out0[g]) + pick_load(a, 1, fb, s.out1[g]);
        if (c < best)
            best = c, bg = g;
    }
    fb_idx[fb] = (int8_t)bg;
}

// Once per device before the first pick_enqueue (not a stream operation: kept out of pick_enqueue so that the schedule can be captured).
inline cudaError_t pick_prepare() { return cudaFuncSetAttribute(pick_tot_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, PT_SMEM); }

// Enqueues the whole schedule of finish_cdef_search on `stream` (77 small launches, no host synchronisation).
inline cudaError_t pick_enqueue(const PickArgs &a, int8_t *d_fb_idx, int64_t lambda, int sm_count, cudaStream_t stream, uint64_t *launches) {
    cudaError_t e = cudaMemsetAsync(a.st, 0, sizeof(PickState), stream);
    if (e != cudaSuccess)
        return e;
    pick_init_kernel<<<max(1, min(a.nfb, 2 * max(sm_count, 1))), PICK_THREADS, 0, stream>>>(a);
    uint64_t           n = 1;
    cudaLaunchAttribute attr[1];
    attr[0].id                                         = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim            = dim3((a.end_gi + PT_ROWS - 1) / PT_ROWS, (a.nfb + PT_CHUNK_N - 1) / PT_CHUNK_N);
    cfg.blockDim           = dim3(PT_THREADS);
    cfg.dynamicSmemBytes   = PT_SMEM;

    cfg.stream             = stream;
    cfg.attrs              = attr;
    cfg.numAttrs           = 1;
    for (int bits = 0; bits <= 3 && e == cudaSuccess; bits++) {
        const int nb = 1 << bits;
        for (int k = 0; k < nb && e == cudaSuccess; k++) { // joint_strength_search_dual: greedy fill
            e = cudaLaunchKernelEx(&cfg, pick_tot_kernel, a, k, 0, k, -1, lambda);
            n += 1;
        }
        for (int k = 0; k < 4 * nb && e == cudaSuccess; k++) { // refinement: drop the oldest entry, re-pick the last
            e = cudaLaunchKernelEx(&cfg, pick_tot_kernel, a, nb - 1, 1, nb - 1, k == 4 * nb - 1 ? bits : -1, lambda);
            n += 1;
        }
    }
    if (e != cudaSuccess)
        return e;
    pick_assign_kernel<<<(a.nfb + PICK_THREADS - 1) / PICK_THREADS, PICK_THREADS, 0, stream>>>(a, d_fb_idx);
    if (launches)
        *launches += n + 1;
    return cudaGetLastError();
}

} // namespace svtf
