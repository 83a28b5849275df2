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

__global__ void __launch_bounds__(PICK_THREADS) pick_init_kernel(const PickArgs a) {
    __shared__ int s_cnt;
    if (threadIdx.x == 0)
        s_cnt = 0;
    __syncthreads();
    int c = 0;
    for (int i = threadIdx.x; i < a.nfb; i += PICK_THREADS) c += a.skip[i] == 0;
    atomicAdd(&s_cnt, c);
    for (int i = threadIdx.x; i < 64 * 64; i += PICK_THREADS) a.tot[i] = 0;
    __syncthreads();
    if (threadIdx.x == 0) {
        PickState &s = *a.st;
        for (int i = 0; i < 8; i++) s.lev0[i] = s.lev1[i] = s.out0[i] = s.out1[i] = 0;
        s.tot_mse = PICK_INF, s.count = s_cnt, s.best_cost = PICK_INF, s.best_bits = 0, s.ticket = 0;
    }
}

// svt_search_one_dual_c, argmin part (first (j,k) in raster order wins ties) + list update + re-zeroing of tot.
// pos: list slot that receives the pick; rot: rotate the list first; bits >= 0: this was the last call for that
// cdef_bits -> RD cost and best-so-far update (EbEncCdef.c:845-876).
__device__ __forceinline__ void pick_argmin_body(const PickArgs &a, const int pos, const int rot, const int bits, const int64_t lambda,
                                                 uint64_t *s_v, int *s_i) {
    const int tid = threadIdx.x;
    uint64_t  bv = ~(uint64_t)0;
    int       bi = 1 << 30;
    uint64_t tv[64 * 64 / PICK_THREADS];
#pragma unroll
    for (int n = 0; n < 64 * 64 / PICK_THREADS; n++) tv[n] = __ldcg(&a.tot[tid + n * PICK_THREADS]); // other CTAs' atomics: read at L2
#pragma unroll
    for (int n = 0; n < 64 * 64 / PICK_THREADS; n++) {
        const int i = tid + n * PICK_THREADS, j = i >> 6, k = i & 63;
        a.tot[i] = 0;
        if (j < a.end_gi && k < a.end_gi && (tv[n] < bv || (tv[n] == bv && i < bi)))
            bv = tv[n], bi = i;
    }
    s_v[tid] = bv, s_i[tid] = bi;
    __syncthreads();
    for (int s = PICK_THREADS / 2; s > 0; s >>= 1) {
        if (tid < s) {
            const uint64_t v = s_v[tid + s];
            const int      i = s_i[tid + s];
            if (v < s_v[tid] || (v == s_v[tid] && i < s_i[tid]))
                s_v[tid] = v, s_i[tid] = i;
        }
        __syncthreads();
    }
    if (tid == 0) {
        PickState &s = *a.st;
        // the scan in the C code starts from best = 1<<63 with a strict "<": totals >= 1<<63 never win and (0,0) stays
        int      b0 = 0, b1 = 0;
        uint64_t best = PICK_INF;
        if (s_v[0] < PICK_INF)
            best = s_v[0], b0 = s_i[0] >> 6, b1 = s_i[0] & 63;
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

// svt_search_one_dual_c, accumulation part.  nb entries of the list are fixed; `rot` = 1 reads them one position
// further (the refinement step drops entry 0, EbEncCdef.c:717-722).  CTA (tile, chunk): a 16x16 tile of tot over a
// chunk of 64 filter blocks; thread (tj, tk) owns one entry.  One u64 atomic per entry and CTA.
constexpr int PICK_CHUNK = 64;
// The CTA that finishes last (ticket counter, no waiting) runs the argmin + list update of the same call.
__global__ void __launch_bounds__(PICK_THREADS) pick_tot_kernel(const PickArgs a, const int nb, const int rot, const int pos, const int bits,
                                                                const int64_t lambda) {
    __shared__ uint64_t sa[PICK_CHUNK][16], sb[PICK_CHUNK][16], sbest[PICK_CHUNK];
    __shared__ uint8_t  svalid[PICK_CHUNK];
    __shared__ int      s_last;
    const int tid = threadIdx.x;
    const int jt = (blockIdx.x >> 2) * 16, kt = (blockIdx.x & 3) * 16, fb0 = blockIdx.y * PICK_CHUNK;
    if (jt < a.end_gi && kt < a.end_gi) {
    for (int i = tid; i < PICK_CHUNK * 32; i += PICK_THREADS) {
        const int f = i >> 5, e = i & 31, fb = fb0 + f;
        if (fb < a.nfb && !a.skip[fb]) {
            if (e < 16)
                sa[f][e] = pick_load(a, 0, fb, jt + e);
            else
                sb[f][e - 16] = pick_load(a, 1, fb, kt + e - 16);
        }
    }
    if (tid < PICK_CHUNK) {
        const int  fb = fb0 + tid;
        const bool ok = fb < a.nfb && !a.skip[fb];
        // all list entries first, then all table loads: two dependent memory round trips instead of 2*nb
        int l0[8], l1[8];
#pragma unroll
        for (int g = 0; g < 8; g++) l0[g] = a.st->lev0[min(g + rot, 7)], l1[g] = a.st->lev1[min(g + rot, 7)];
        uint64_t va[8], vb[8];
#pragma unroll
        for (int g = 0; g < 8; g++) {
            const bool use = ok && g < nb;
            va[g] = use ? pick_load(a, 0, fb, l0[g]) : PICK_INF, vb[g] = use ? pick_load(a, 1, fb, l1[g]) : 0;
        }
        uint64_t best = PICK_INF;
#pragma unroll
        for (int g = 0; g < 8; g++) best = min(best, va[g] + vb[g]);
        sbest[tid]  = best;
        svalid[tid] = ok;
    }
    __syncthreads();
    const int tj = tid >> 4, tk = tid & 15;
    uint64_t  acc = 0;
#pragma unroll 4
    for (int f = 0; f < PICK_CHUNK; f++)
        if (svalid[f])
            acc += min(sbest[f], sa[f][tj] + sb[f][tk]);
    if (jt + tj < a.end_gi && kt + tk < a.end_gi && acc)
        atomicAdd(reinterpret_cast<unsigned long long *>(&a.tot[(jt + tj) * 64 + kt + tk]), (unsigned long long)acc);
    }
    __threadfence();
    __syncthreads();
    if (tid == 0)
        s_last = atomicAdd(&a.st->ticket, 1u) == gridDim.x * gridDim.y - 1;
    __syncthreads();
    if (s_last) {
        __threadfence();
        pick_argmin_body(a, pos, rot, bits, lambda, &sa[0][0], reinterpret_cast<int *>(&sb[0][0]));
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

// Enqueues the whole schedule of finish_cdef_search on `stream` (77 small launches, no host synchronisation).
inline cudaError_t pick_enqueue(const PickArgs &a, int8_t *d_fb_idx, int64_t lambda, int sm_count, cudaStream_t stream, uint64_t *launches) {
    (void)sm_count;
    const dim3 grid(16, (a.nfb + PICK_CHUNK - 1) / PICK_CHUNK);
    pick_init_kernel<<<1, PICK_THREADS, 0, stream>>>(a);
    uint64_t n = 1;
    for (int bits = 0; bits <= 3; bits++) {
        const int nb = 1 << bits;
        for (int k = 0; k < nb; k++) { // joint_strength_search_dual: greedy fill
            pick_tot_kernel<<<grid, PICK_THREADS, 0, stream>>>(a, k, 0, k, -1, lambda);
            n += 1;
        }
        for (int k = 0; k < 4 * nb; k++) { // refinement: drop the oldest entry, re-pick the last
            pick_tot_kernel<<<grid, PICK_THREADS, 0, stream>>>(a, nb - 1, 1, nb - 1, k == 4 * nb - 1 ? bits : -1, lambda);
            n += 1;
        }
    }
    pick_assign_kernel<<<(a.nfb + PICK_THREADS - 1) / PICK_THREADS, PICK_THREADS, 0, stream>>>(a, d_fb_idx);
    if (launches)
        *launches += n + 1;
    return cudaGetLastError();
}

} // namespace svtf
