// svtf_search.cuh — the encoder's CDEF strength search on the GPU.
//
//   cdef_search_kernel<T>  one CTA per 64x64 filter block: stages the post-deblock tile (+2 halo) of
//                          Y, U, V, finds the 8x8 directions once, then for every (pri,sec) pair of the
//                          level's list filters all non-skip blocks and reduces the distortion against
//                          the source picture with warp shuffles + shared-memory u64 atomics.
//                          Writes mse[2][nfb][64] (the layout of pcs->mse_seg), skip[nfb], dir/var.
//
// Restates cdef_seg_search (Encoder/Codec/EbCdefProcess.c:114-357), compute_cdef_dist (:60-107),
// svt_aom_compute_cdef_dist_c/_8bit_c + dist_8xn/mse_* (Encoder/Codec/EbEncCdef.c:23-219).
// The luma distortion is IEEE double; every operation uses the explicit round-to-nearest intrinsics so
// nvcc cannot contract a*b+c into an FMA (the reference's C is compiled without FMA).
#pragma once
#include "svtf_kernels.cuh"

namespace svtf {

constexpr int SEARCH_THREADS = 256;

struct SearchArgs {
    Planes           rec, src;
    const svtf_mi_t *mi;
    int              mi_stride;
    Dims             d;
    int              n;                            // number of (pri,sec) pairs
    uint8_t          fs[SVTF_TOTAL_STRENGTHS];     // strength byte per index
    int8_t           fs_uv[SVTF_TOTAL_STRENGTHS];  // -1 = chroma masked
    int              damping;
    int              subsampling;
    uint64_t        *mse;  // [2][nfb][64]
    uint8_t         *skip; // [nfb]
    uint8_t         *dir;  // [nfb][64]
    int32_t         *var;  // [nfb][64]
    int              nfb;
    int              fbr0; // first filter-block row of the launch (stripe window)
    // 128-wide / 128-high coding blocks (sb_size 128): the reference searches such a block as ONE filter block whose table row
    // belongs to its first 64x64 quadrant (Encoder/Codec/EbCdefProcess.c:186-203).  Quadrant CTAs add their raw sums here and
    // search_fold_kernel turns them into the first quadrant's row afterwards (the >> 2*coeff_shift comes after the sum over the
    // whole block, so per-quadrant rows cannot simply be added for 10/12-bit pictures).
    unsigned long long *raw;  // [3][nfb][64] raw sums (Y, U, V) per first quadrant, zero on entry
    unsigned           *fold; // [nfb] bit 0: first quadrant of a folded block; bit 1: some 8x8 block of it is not skipped; zero on entry
    int                 fbr_par; // parity of the picture's filter-block row of local row 0 (stripes)
};

// Which filter block carries the table row of filter block (fbr, fbc)?  `folded`: it lies in a 128-wide / 128-high coding block.
struct SearchFold {
    bool folded;
    int  ofb; // the block's first quadrant (== fb when !folded)
};
__device__ __forceinline__ SearchFold search_fold(const SearchArgs &a, const int fbr, const int fbc, const int nhfb) {
    const svtf_mi_t m  = __ldg(a.mi + (size_t)(fbr * 16) * a.mi_stride + fbc * 16);
    const bool      w5 = mi_bdim(m, 0) == 5, h5 = mi_bdim(m, 1) == 5; // log2(128) - 2
    SearchFold      f;
    f.folded      = w5 || h5;
    const int ofr = h5 && ((fbr + a.fbr_par) & 1) ? fbr - 1 : fbr, ofc = w5 && (fbc & 1) ? fbc - 1 : fbc;
    f.ofb         = max(ofr, 0) * nhfb + ofc;
    return f;
}
// One table entry of one filter block: plain store, or the raw sums added to the folded block's first quadrant.
__device__ __forceinline__ void search_store(const SearchArgs &a, const int fb, const SearchFold &f, const int gi, const unsigned long long y,
                                             const unsigned long long u, const unsigned long long v, const int cs) {
    if (f.folded) {
        atomicAdd(&a.raw[((size_t)0 * a.nfb + f.ofb) * 64 + gi], y);
        atomicAdd(&a.raw[((size_t)1 * a.nfb + f.ofb) * 64 + gi], u);
        atomicAdd(&a.raw[((size_t)2 * a.nfb + f.ofb) * 64 + gi], v);
        return;
    }
    const int sub = min(a.subsampling, 4);
    a.mse[(size_t)fb * 64 + gi]             = (y >> (2 * cs)) * (unsigned long long)sub;
    a.mse[((size_t)a.nfb + fb) * 64 + gi] = a.fs_uv[gi] == -1 ? 1040400ull * 64 : (u >> (2 * cs)) + (v >> (2 * cs));
}
// After the search kernel: rows and skip flags of the folded blocks' first quadrants.  One CTA of 64 threads per filter block.
__global__ void __launch_bounds__(64) search_fold_kernel(const __grid_constant__ SearchArgs a) {
    const int      fb = blockIdx.x, gi = threadIdx.x;
    const unsigned fl = a.fold[fb];
    if (!(fl & 1u))
        return;
    if (gi == 0)
        a.skip[fb] = (fl & 2u) ? 0 : 1;
    if (!(fl & 2u) || gi >= a.n) // nothing to filter in the whole block: the row stays zero (EbCdefProcess.c:203-207)
        return;
    const int cs = a.d.bit_depth - 8, sub = min(a.subsampling, 4);
    a.mse[(size_t)fb * 64 + gi] = (a.raw[((size_t)0 * a.nfb + fb) * 64 + gi] >> (2 * cs)) * (unsigned long long)sub;
    a.mse[((size_t)a.nfb + fb) * 64 + gi] =
        a.fs_uv[gi] == -1 ? 1040400ull * 64
                          : (a.raw[((size_t)1 * a.nfb + fb) * 64 + gi] >> (2 * cs)) + (a.raw[((size_t)2 * a.nfb + fb) * 64 + gi] >> (2 * cs));
}

// dist_8xn_16bit_c / dist_8xn_8bit_c — Encoder/Codec/EbEncCdef.c:23-48, :76-101
__device__ __forceinline__ unsigned long long cdef_dist_8x8(unsigned long long sum_s, unsigned long long sum_d,
                                                            unsigned long long sum_s2, unsigned long long sum_d2,
                                                            unsigned long long sum_sd, const int cs) {
    const unsigned long long svar = sum_s2 - ((sum_s * sum_s + 32) >> 6);
    const unsigned long long dvar = sum_d2 - ((sum_d * sum_d + 32) >> 6);
    const double num = __dmul_rn(__dmul_rn((double)(sum_d2 + sum_s2 - 2 * sum_sd), .5), (double)(svar + dvar + (unsigned long long)(400 << 2 * cs)));
    const double den = __dsqrt_rn(__dadd_rn((double)(20000 << 4 * cs), __dmul_rn((double)svar, (double)dvar)));
    return (unsigned long long)floor(__dadd_rn(.5, __ddiv_rn(num, den)));
}

template <typename T>
__global__ void __launch_bounds__(SEARCH_THREADS) cdef_search_kernel(const __grid_constant__ SearchArgs a) {
    __shared__ __align__(16) uint16_t ytile[CDEF_YROWS * CDEF_YP];
    __shared__ __align__(16) uint16_t ctile[2][CDEF_CROWS * CDEF_CP];
    __shared__ int                    s_cost[8][64];
    __shared__ uint8_t                s_dir[64];
    __shared__ int                    s_var[64];
    __shared__ uint8_t                s_active[64];
    __shared__ unsigned long long     s_acc[3][SVTF_TOTAL_STRENGTHS];

    const int tid = threadIdx.x;
    const int fbc = blockIdx.x, fbr = blockIdx.y + a.fbr0;
    const int fb  = fbr * gridDim.x + fbc;
    const int cw = a.d.mi_cols * 4, ch = a.d.mi_rows * 4;
    const int x0 = fbc * 64, y0 = fbr * 64;
    const int cs = a.d.bit_depth - 8;

    int act = 0;
    if (tid < 64) {
        const int r = (tid >> 3) * 2, c = (tid & 7) * 2;
        if (fbr * 16 + r < a.d.mi_rows && fbc * 16 + c < a.d.mi_cols) {
            const svtf_mi_t *m = a.mi + (fbr * 16 + r) * a.mi_stride + fbc * 16 + c;
            act = !(mi_skip(__ldg(m)) & mi_skip(__ldg(m + 1)) & mi_skip(__ldg(m + a.mi_stride)) & mi_skip(__ldg(m + a.mi_stride + 1)));
        }
        s_active[tid] = (uint8_t)act;
    }
    for (int i = tid; i < 3 * SVTF_TOTAL_STRENGTHS; i += SEARCH_THREADS) (&s_acc[0][0])[i] = 0;
    const int        any_active = __syncthreads_or(act);
    const SearchFold fold       = search_fold(a, fbr, fbc, gridDim.x);
    if (tid == 0) {
        if (fold.folded)
            atomicOr(&a.fold[fold.ofb], any_active ? 3u : 1u);
        else
            a.skip[fb] = any_active ? 0 : 1;
    }
    if (!any_active) // skip_cdef_seg = 1 (EbCdefProcess.c:203-207); mse rows stay zero
        return;

    cdef_stage<T>(ytile, CDEF_YP, CDEF_YROWS, 34, static_cast<const T *>(a.rec.p[0]), a.rec.pitch[0], x0, y0, cw, ch, tid);
    cdef_stage<T>(ctile[0], CDEF_CP, CDEF_CROWS, 18, static_cast<const T *>(a.rec.p[1]), a.rec.pitch[1], x0 / 2, y0 / 2, cw / 2, ch / 2, tid);
    cdef_stage<T>(ctile[1], CDEF_CP, CDEF_CROWS, 18, static_cast<const T *>(a.rec.p[2]), a.rec.pitch[2], x0 / 2, y0 / 2, cw / 2, ch / 2, tid);
    __syncthreads();
    {
        const int w = tid >> 5, lane = tid & 31;
#pragma unroll
        for (int h = 0; h < 2; h++) {
            const int       b  = h * 32 + lane;
            const uint16_t *bp = ytile + (2 + (b >> 3) * 8) * CDEF_YP + 2 + (b & 7) * 8;
            int             c  = 0;
            if (s_active[b]) {
                switch (w) {
                case 0: c = cdef_dir_cost<0>(bp, CDEF_YP, cs); break;
                case 1: c = cdef_dir_cost<1>(bp, CDEF_YP, cs); break;
                case 2: c = cdef_dir_cost<2>(bp, CDEF_YP, cs); break;
                case 3: c = cdef_dir_cost<3>(bp, CDEF_YP, cs); break;
                case 4: c = cdef_dir_cost<4>(bp, CDEF_YP, cs); break;
                case 5: c = cdef_dir_cost<5>(bp, CDEF_YP, cs); break;
                case 6: c = cdef_dir_cost<6>(bp, CDEF_YP, cs); break;
                default: c = cdef_dir_cost<7>(bp, CDEF_YP, cs); break;
                }
            }
            s_cost[w][b] = c;
        }
    }
    __syncthreads();
    if (tid < 64) {
        int best = 0, bd = 0;
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const int c = s_cost[i][tid];
            if (c > best) {
                best = c;
                bd   = i;
            }
        }
        const int v = (best - s_cost[(bd + 4) & 7][tid]) >> 10;
        s_dir[tid] = (uint8_t)bd;
        s_var[tid] = v;
        if (s_active[tid]) {
            a.dir[fb * 64 + tid] = (uint8_t)bd;
            a.var[fb * 64 + tid] = v;
        }
    }
    __syncthreads();

    // ---------------- luma: 8 lanes per 8x8 block (one row each), 32 blocks per pass ----------------
    {
        const int sub  = min(a.subsampling, 4); // EbCdefProcess.c:246-251
        const int damp = a.damping + cs;
        const T  *srcy = static_cast<const T *>(a.src.p[0]);
#pragma unroll 1
        for (int h = 0; h < 2; h++) {
            const int  b   = h * 32 + (tid >> 3), row = tid & 7;
            const int  by = b >> 3, bx = b & 7;
            const bool on = s_active[b] && (row % sub) == 0;
            const uint16_t *c = ytile + (2 + by * 8 + row) * CDEF_YP + 2 + bx * 8;
            int             d[8];
            if (on) {
                const uint4 v = ld8<T>(srcy + (size_t)(y0 + by * 8 + row) * a.src.pitch[0] + x0 + bx * 8);
                d[0] = v.x & 0xffff, d[1] = v.x >> 16, d[2] = v.y & 0xffff, d[3] = v.y >> 16;
                d[4] = v.z & 0xffff, d[5] = v.z >> 16, d[6] = v.w & 0xffff, d[7] = v.w >> 16;
            }
            const int var = s_var[b];
            const int ii  = (var >> 6) ? min(cdef_msb(var >> 6), 12) : 0;
#pragma unroll 1
            for (int gi = 0; gi < a.n; gi++) {
                const int lev = a.fs[gi] >> 2;
                int       sl  = a.fs[gi] & 3;
                sl += sl == 3;
                const int pri_l = lev << cs, sec = sl << cs;
                unsigned  ss = 0, sd = 0, ss2 = 0, sd2 = 0, ssd = 0;
                if (on) {
                    const int pri = var ? (pri_l * (4 + ii) + 8) >> 4 : 0;
                    int       o[6];
                    cdef_offsets(o, pri_l ? s_dir[b] : 0, CDEF_YP);
                    const int pri_shift = pri ? max(0, damp - cdef_msb(pri)) : 0;
                    const int sec_shift = sec ? max(0, damp - cdef_msb(sec)) : 0;
                    const int odd = (pri >> cs) & 1, w0 = odd ? 3 : 4, w1 = odd ? 3 : 2;
#pragma unroll
                    for (int j = 0; j < 8; j++) {
                        // (0,0): plain copy, Common/Codec/EbCdef.c:356-381 — identical to filtering with zero strengths
                        const unsigned s = (pri_l | sec) ? (unsigned)cdef_px(c + j, o, pri, sec, pri_shift, sec_shift, w0, w1) : c[j];
                        const unsigned dd = (unsigned)d[j];
                        ss += s, sd += dd, ss2 += s * s, sd2 += dd * dd, ssd += s * dd;
                    }
                }
#pragma unroll
                for (int m = 1; m < 8; m <<= 1) {
                    ss += __shfl_xor_sync(0xffffffffu, ss, m);
                    sd += __shfl_xor_sync(0xffffffffu, sd, m);
                    ss2 += __shfl_xor_sync(0xffffffffu, ss2, m);
                    sd2 += __shfl_xor_sync(0xffffffffu, sd2, m);
                    ssd += __shfl_xor_sync(0xffffffffu, ssd, m);
                }
                if (row == 0 && s_active[b])
                    atomicAdd(&s_acc[0][gi], cdef_dist_8x8(ss, sd, ss2, sd2, ssd, cs));
            }
        }
    }
    // ---------------- chroma: 4 lanes per 4x4 block, 64 blocks per plane, 1 pass per plane ----------------
#pragma unroll 1
    for (int p = 0; p < 2; p++) {
        const int  damp = a.damping + cs - 1;
        const T   *srcc = static_cast<const T *>(a.src.p[1 + p]);
        const int  b = tid >> 2, row = tid & 3;
        const int  by = b >> 3, bx = b & 7;
        const bool on = s_active[b];
        const uint16_t *c = ctile[p] + (2 + by * 4 + row) * CDEF_CP + 2 + bx * 4;
        int             d[4];
        if (on) {
            const uint2 v = ld4<T>(srcc + (size_t)(y0 / 2 + by * 4 + row) * a.src.pitch[1 + p] + x0 / 2 + bx * 4);
            d[0] = v.x & 0xffff, d[1] = v.x >> 16, d[2] = v.y & 0xffff, d[3] = v.y >> 16;
        }
#pragma unroll 1
        for (int gi = 0; gi < a.n; gi++) {
            if (a.fs_uv[gi] == -1)
                continue;
            const int lev = a.fs[gi] >> 2;
            int       sl  = a.fs[gi] & 3;
            sl += sl == 3;
            const int pri = lev << cs, sec = sl << cs;
            unsigned  sse = 0;
            if (on) {
                int o[6];
                cdef_offsets(o, pri ? s_dir[b] : 0, CDEF_CP);
                const int pri_shift = pri ? max(0, damp - cdef_msb(pri)) : 0;
                const int sec_shift = sec ? max(0, damp - cdef_msb(sec)) : 0;
                const int odd = (pri >> cs) & 1, w0 = odd ? 3 : 4, w1 = odd ? 3 : 2;
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const int s = (pri | sec) ? cdef_px(c + j, o, pri, sec, pri_shift, sec_shift, w0, w1) : (int)c[j];
                    const int e = d[j] - s;
                    sse += (unsigned)(e * e);
                }
            }
            sse += __shfl_xor_sync(0xffffffffu, sse, 1);
            sse += __shfl_xor_sync(0xffffffffu, sse, 2);
            if (row == 0 && on)
                atomicAdd(&s_acc[1 + p][gi], (unsigned long long)sse);
        }
    }
    __syncthreads();
    // mse_seg[0][fb][gi] = (sum >> 2cs) * sub ; mse_seg[1][fb][gi] = U + V (EbCdefProcess.c:297-300),
    // masked chroma strengths -> default_mse_uv * 64 (:86,:258-261)
    if (tid < a.n)
        search_store(a, fb, fold, tid, s_acc[0][tid], s_acc[1][tid], s_acc[2][tid], cs);
}

} // namespace svtf
