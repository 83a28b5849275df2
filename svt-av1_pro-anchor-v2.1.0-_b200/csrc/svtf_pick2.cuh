// svtf_pick2.cuh — finish_cdef_search in ONE launch: a thread-block cluster runs the whole greedy search.
//
// finish_cdef_search (Encoder/Codec/EbEncCdef.c:728-925) calls svt_search_one_dual (:627-684) 75 times per picture (cdef_bits 0..3:
// nb greedy picks + 4*nb refinement picks, joint_strength_search_dual :697-727); every call needs the totals
//     tot[j][k] = sum over searched filter blocks of min(best_so_far(fb), mse0[fb][j] + mse1[fb][k])
// over ALL filter blocks before its argmin, and the next call needs that argmin: a chain of 75 picture-wide reductions.  svtf_pick.cuh
// runs each call as its own launch (77 dependent launches, 13 us each).  Here ONE cluster of PK2_CTAS CTAs stays resident for the whole
// schedule.  MEASURED (4K 10-bit, 2 040 filter blocks, 64 strengths): 0.92 ms per pick against 0.99 ms for the 77 launches — the 16 SMs
// of one cluster are bound by the integer pipe (VIADDMNMX + IADD3 per entry and filter block at one ALU-pipe instruction every other
// cycle per scheduler = 8 us per call), where the per-call launches spread the same work over all 148 SMs — and with 8 pictures in
// flight the resident clusters take SMs from the search kernels (2.05 M against 2.88 M filter blocks/s).  So it is NOT the default:
// SVTF_PICK_CLUSTER=1 selects it (tests/test_gpu_pick_cluster.py keeps it bit-exact).  How it works:
//   * CTA r keeps the table rows of its slice of the filter blocks in shared memory as 32-bit words, [fb][128] (read from global
//     memory once), so a thread's four luma and four chroma entries of a filter block are two conflict-free 128-bit loads;
//   * a call = every CTA accumulates the 64 x 64 totals of its slice — a thread owns a 4 x 4 tile of (j, k): one VIADDMNMX
//     (min(a + b, best)) and one 32-bit add per entry and filter block, flushed into 64-bit totals before 32 bits can overflow —
//     and writes them to its shared memory; cluster barrier; CTA r then sums entries [256 r, 256 r + 256) over the 16 CTAs through
//     distributed shared memory, reduces them to one (value, index) candidate and stores it into every CTA's candidate array; cluster
//     barrier; every CTA takes the argmin of the 16 candidates and updates ITS OWN copy of the greedy lists — the copies evolve
//     identically, so nothing is broadcast;
//   * no global-memory traffic and no kernel boundary between calls: two hardware cluster barriers per call.
// Preconditions (else the kernel sets PickState.wide = 2 and returns at once, and the host runs the per-call launches of svtf_pick.cuh):
// every table entry the pick reads is below 2^31 (then a + b fits 32 bits), and a CTA's slice fits its shared memory (the host checks
// the latter before launching: up to ~5 900 filter blocks, i.e. everything up to 6K pictures).
// All results are the u64 / i64 integer arithmetic of the C code (same tie-breaking: the first (j, k) in raster order wins).
#pragma once
#include <cooperative_groups.h>
#include "svtf_pick.cuh"

namespace svtf {

constexpr int PK2_CTAS    = 16; // cluster size (non-portable size, opt-in; the B200 schedules it)
constexpr int PK2_THREADS = 256;

struct Pick2Shared {
    unsigned long long tot[64 * 64];        // this CTA's partial totals of the running call
    unsigned long long cand_v[2][PK2_CTAS]; // per-CTA argmin candidates of the running call (double-buffered by call parity)
    int                cand_i[2][PK2_CTAS];
    unsigned long long red_v[PK2_THREADS / 32];
    int                red_i[PK2_THREADS / 32];
    int                lev0[8], lev1[8];     // this CTA's copy of the greedy lists
    int                count_part[PK2_CTAS]; // searched filter blocks per CTA
    unsigned           max_part[PK2_CTAS];   // largest table entry per CTA, saturated at 2^31 (= "wide")
};

__global__ void __launch_bounds__(PK2_THREADS, 1) pick_cluster_kernel(const PickArgs a, const int64_t lambda, int8_t *fb_idx) {
    namespace cg = cooperative_groups;
    cg::cluster_group cluster = cg::this_cluster();
    extern __shared__ __align__(16) unsigned char pk2_dyn[];
    Pick2Shared &S    = *reinterpret_cast<Pick2Shared *>(pk2_dyn);
    const int    rank = (int)cluster.block_rank(), tid = threadIdx.x;
    const int    per  = (a.nfb + PK2_CTAS - 1) / PK2_CTAS; // filter blocks per CTA (the last slices may be short or empty)
    const int    f0 = min(rank * per, a.nfb), nloc = min(per, a.nfb - f0);
    uint32_t    *best = reinterpret_cast<uint32_t *>(pk2_dyn + sizeof(Pick2Shared));                                // [per]
    uint32_t    *rows = reinterpret_cast<uint32_t *>(pk2_dyn + sizeof(Pick2Shared) + ((sizeof(uint32_t) * per + 15) & ~(size_t)15)); // [per][128]

    // ---- set-up: searched-block count, the largest entry (wide check + flush interval), then this CTA's rows ----
    int      cnt = 0;
    unsigned mx  = 0;
    for (int i = tid; i < nloc * 128; i += PK2_THREADS) {
        const int f = i >> 7, e = i & 127;
        if ((e & 63) < a.end_gi && !a.skip[f0 + f]) {
            const unsigned long long v = pick_load(a, e >> 6, f0 + f, e & 63);
            mx                         = max(mx, (v >> 31) ? 0x80000000u : (unsigned)v);
        }
    }
    for (int f = tid; f < nloc; f += PK2_THREADS) cnt += a.skip[f0 + f] ? 0 : 1;
#pragma unroll
    for (int m = 16; m > 0; m >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, m), mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, m));
    if ((tid & 31) == 0)
        S.red_i[tid >> 5] = cnt, S.red_v[tid >> 5] = mx;
    __syncthreads();
    if (tid == 0) {
        int      c = 0;
        unsigned m = 0;
        for (int w = 0; w < PK2_THREADS / 32; w++) c += S.red_i[w], m = max(m, (unsigned)S.red_v[w]);
        for (int r = 0; r < PK2_CTAS; r++) {
            Pick2Shared *o      = cluster.map_shared_rank(&S, r);
            o->count_part[rank] = c, o->max_part[rank] = m;
        }
    }
    cluster.sync();
    int      count = 0;
    unsigned vmax  = 0;
    for (int r = 0; r < PK2_CTAS; r++) count += S.count_part[r], vmax = max(vmax, S.max_part[r]);
    if (vmax >= 0x80000000u) { // wide table: not this kernel's case (the host falls back)
        if (rank == 0 && tid == 0)
            a.st->wide = 2;
        cluster.sync();
        return;
    }
    // min(a + b, best) < 2 * (vmax + 1) (best = 2^32 - 1 only bounds sums that are smaller anyway): filter blocks per 32-bit accumulation
    const int flush = (int)min(0xffffffffull / (2ull * vmax + 2ull), 4096ull);
    for (int i = tid; i < nloc * 128; i += PK2_THREADS) rows[i] = (uint32_t)pick_load(a, (i & 127) >> 6, f0 + (i >> 7), i & 63);

    // ---- the schedule of finish_cdef_search; every CTA runs the same bookkeeping on its own copy of the state ----
    uint64_t  best_cost = PICK_INF;
    int       best_bits = 0, out0[8] = {0, 0, 0, 0, 0, 0, 0, 0}, out1[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    int       call = 0;
    const int tj = tid >> 4, tk = tid & 15;
    for (int bits = 0; bits <= 3; bits++) {
        const int nbs = 1 << bits;
        if (tid < 8)
            S.lev0[tid] = 0, S.lev1[tid] = 0;
        uint64_t tot_mse = PICK_INF;
        for (int step = 0; step < 5 * nbs; step++, call++) {
            // greedy fill: entries [0, step) are fixed, pick entry `step`; refinement: drop the oldest entry, re-pick the last one
            const bool refine = step >= nbs;
            const int  nb = refine ? nbs - 1 : step, rot = refine ? 1 : 0, pos = refine ? nbs - 1 : step;
            __syncthreads(); // the lists of this call are in place (and the previous call's reads of `best` are done)
            for (int f = tid; f < nloc; f += PK2_THREADS) { // best_so_far of every local filter block (skipped: 0 -> contributes 0)
                uint32_t b = 0;
                if (!a.skip[f0 + f]) {
                    b = 0xffffffffu;
                    for (int g = 0; g < nb; g++) b = min(b, rows[f * 128 + S.lev0[min(g + rot, 7)]] + rows[f * 128 + 64 + S.lev1[min(g + rot, 7)]]);
                }
                best[f] = b;
            }
            __syncthreads();
            unsigned long long acc[4][4];
#pragma unroll
            for (int r = 0; r < 4; r++)
#pragma unroll
                for (int c = 0; c < 4; c++) acc[r][c] = 0;
            for (int fa = 0; fa < nloc; fa += flush) {
                uint32_t  p[4][4];
                const int fe = min(nloc, fa + flush);
#pragma unroll
                for (int r = 0; r < 4; r++)
#pragma unroll
                    for (int c = 0; c < 4; c++) p[r][c] = 0;
#pragma unroll 4
                for (int f = fa; f < fe; f++) {
                    const uint4    av = *reinterpret_cast<const uint4 *>(rows + f * 128 + tj * 4);
                    const uint4    bv = *reinterpret_cast<const uint4 *>(rows + f * 128 + 64 + tk * 4);
                    const uint32_t bst = best[f];
                    const uint32_t A[4] = {av.x, av.y, av.z, av.w}, B[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
                    for (int r = 0; r < 4; r++)
#pragma unroll
                        for (int c = 0; c < 4; c++) p[r][c] += __viaddmin_u32(A[r], B[c], bst);
                }
#pragma unroll
                for (int r = 0; r < 4; r++)
#pragma unroll
                    for (int c = 0; c < 4; c++) acc[r][c] += p[r][c];
            }
#pragma unroll
            for (int r = 0; r < 4; r++)
#pragma unroll
                for (int c = 0; c < 4; c++) S.tot[(tj * 4 + r) * 64 + tk * 4 + c] = acc[r][c];
            cluster.sync(); // every CTA's partial totals are written
            // entries [256 rank, 256 rank + 256): sum over the CTAs, then this CTA's candidate
            {
                const int          i = rank * PK2_THREADS + tid, j = i >> 6, k = i & 63;
                unsigned long long v = 0;
#pragma unroll
                for (int r = 0; r < PK2_CTAS; r++) v += cluster.map_shared_rank(&S.tot[0], r)[i];
                unsigned long long bv = ~0ull;
                int                bi = 1 << 30;
                if (j < a.end_gi && k < a.end_gi)
                    bv = v, bi = i;
#pragma unroll
                for (int m = 16; m > 0; m >>= 1) {
                    const unsigned long long ov = __shfl_xor_sync(0xffffffffu, bv, m);
                    const int                oi = __shfl_xor_sync(0xffffffffu, bi, m);
                    if (ov < bv || (ov == bv && oi < bi))
                        bv = ov, bi = oi;
                }
                if ((tid & 31) == 0)
                    S.red_v[tid >> 5] = bv, S.red_i[tid >> 5] = bi;
                __syncthreads();
                if (tid < PK2_CTAS) { // thread r reduces the warps' candidates and stores the result into CTA r's array
                    bv = S.red_v[0], bi = S.red_i[0];
                    for (int w = 1; w < PK2_THREADS / 32; w++)
                        if (S.red_v[w] < bv || (S.red_v[w] == bv && S.red_i[w] < bi))
                            bv = S.red_v[w], bi = S.red_i[w];
                    Pick2Shared *o            = cluster.map_shared_rank(&S, tid);
                    o->cand_v[call & 1][rank] = bv, o->cand_i[call & 1][rank] = bi;
                }
            }
            cluster.sync(); // the candidates are everywhere; the partial totals may be overwritten by the next call
            if (tid == 0) { // svt_search_one_dual's argmin + the list update, identically in every CTA
                unsigned long long bv = S.cand_v[call & 1][0];
                int                bi = S.cand_i[call & 1][0];
                for (int r = 1; r < PK2_CTAS; r++)
                    if (S.cand_v[call & 1][r] < bv || (S.cand_v[call & 1][r] == bv && S.cand_i[call & 1][r] < bi))
                        bv = S.cand_v[call & 1][r], bi = S.cand_i[call & 1][r];
                // the scan in the C code starts from best = 1 << 63 with a strict "<": totals >= 1 << 63 never win and (0, 0) stays
                int b0 = 0, b1 = 0;
                tot_mse = PICK_INF;
                if (bv < PICK_INF)
                    tot_mse = bv, b0 = bi >> 6, b1 = bi & 63;
                if (rot)
                    for (int g = 0; g < pos; g++) S.lev0[g] = S.lev0[g + 1], S.lev1[g] = S.lev1[g + 1];
                S.lev0[pos] = b0, S.lev1[pos] = b1;
                if (step == 5 * nbs - 1) { // last call of this cdef_bits: RD cost (EbEncCdef.c:845-876)
                    const int      total_bits = count * bits + nbs * 6 * 2;
                    const int64_t  rate       = (int64_t)total_bits * 512; // av1_cost_literal
                    const uint64_t dist       = tot_mse * 16;
                    const uint64_t cost       = (uint64_t)(((rate * lambda + 256) >> 9) + (int64_t)dist * 128); // RDCOST
                    if (cost < best_cost) {
                        best_cost = cost, best_bits = bits;
                        for (int g = 0; g < nbs; g++) out0[g] = S.lev0[g], out1[g] = S.lev1[g];
                    }
                }
            }
        }
    }
    // ---- per-filter-block index into the chosen set (EbEncCdef.c:884-905); the winner is published through shared memory ----
    __syncthreads();
    if (tid == 0) {
        S.red_i[0] = best_bits;
        for (int g = 0; g < 8; g++) S.lev0[g] = out0[g], S.lev1[g] = out1[g];
        if (rank == 0) {
            PickState &s = *a.st;
            s.best_bits = best_bits, s.best_cost = best_cost, s.count = count;
            for (int g = 0; g < 8; g++) s.out0[g] = out0[g], s.out1[g] = out1[g];
        }
    }
    __syncthreads();
    const int nb = 1 << S.red_i[0];
    for (int f = tid; f < nloc; f += PK2_THREADS) {
        int8_t r = -1;
        if (!a.skip[f0 + f]) {
            uint32_t bst = 0xffffffffu;
            int      bg  = 0;
            for (int g = 0; g < nb; g++) {
                const uint32_t c = rows[f * 128 + S.lev0[g]] + rows[f * 128 + 64 + S.lev1[g]];
                if (c < bst)
                    bst = c, bg = g;
            }
            r = (int8_t)bg;
        }
        fb_idx[f0 + f] = r;
    }
    cluster.sync(); // no CTA leaves while another may still read its shared memory
}

// Bytes of dynamic shared memory for `per` filter blocks per CTA.
inline size_t pick2_smem_bytes(int per) { return sizeof(Pick2Shared) + ((sizeof(uint32_t) * (size_t)per + 15) & ~(size_t)15) + sizeof(uint32_t) * 128 * (size_t)per; }
// Does a table of nfb filter blocks fit the cluster's shared memory (max_smem: the device's opt-in shared memory per block)?
inline bool pick2_fits(int nfb, size_t max_smem) { return pick2_smem_bytes((nfb + PK2_CTAS - 1) / PK2_CTAS) <= max_smem; }

// One launch of the whole pick on `stream` (the PickState block must be zero on entry: wide == 2 afterwards means "not handled").
inline cudaError_t pick2_enqueue(const PickArgs &a, int8_t *d_fb_idx, int64_t lambda, cudaStream_t stream, uint64_t *launches) {
    const size_t smem = pick2_smem_bytes((a.nfb + PK2_CTAS - 1) / PK2_CTAS);
    cudaError_t  e    = cudaMemsetAsync(a.st, 0, sizeof(PickState), stream);
    if (e == cudaSuccess)
        e = cudaFuncSetAttribute(pick_cluster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e == cudaSuccess)
        e = cudaFuncSetAttribute(pick_cluster_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    if (e != cudaSuccess)
        return e;
    cudaLaunchAttribute attr[1];
    attr[0].id               = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = PK2_CTAS, attr[0].val.clusterDim.y = 1, attr[0].val.clusterDim.z = 1;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim            = dim3(PK2_CTAS);
    cfg.blockDim           = dim3(PK2_THREADS);
    cfg.dynamicSmemBytes   = smem;
    cfg.stream             = stream;
    cfg.attrs              = attr;
    cfg.numAttrs           = 1;
    e = cudaLaunchKernelEx(&cfg, pick_cluster_kernel, a, lambda, d_fb_idx);
    if (e == cudaSuccess && launches)
        *launches += 1;
    return e;
}

} // namespace svtf
