// svtf_ccso_derive.cuh — device side of the CCSO filter derivation (ccso_search / derive_ccso_filter,
// Encoder/Codec/EbPickccso.c:464-782).
//
// The reference tries, per plane, 200 configurations (bo_only x 6 tap supports x 4 quantisers x 2 classifiers x 4 band
// counts, + 8 band-only settings), each with up to 16 training iterations of: fold per-block histograms into a LUT,
// filter the WHOLE plane with it, per-block SSE against the source, per-block on/off.  The two things that touch every
// sample are kernels here; the small sequential decisions (float LUT rounding, double RD cost, on/off flags) stay on the
// host exactly as the C code has them (svtf_api.cu: ccso_derive_plane).
//
//   ccso_hist_kernel  per-block histograms  sum(org - rec), count  by LUT index, at the FINEST band split of the group
//                     (coarser band counts are sums of these: band = Y >> (bd - log2)), ccso_pre_compute_class_err(_bo)
//   ccso_try_kernel   trial filter + per-block SSE for up to 8 LUTs in one pass (the band settings of a group train
//                     in lock-step, their loops are independent), ccso_try_*_filter + compute_distortion
// Both take a batch of configuration groups in gridDim.z (CdGroup): the 49 groups of a plane are independent of each other, so one
// launch builds all their histograms and one launch per training round tries every LUT still in training (round 2: 18 launches and
// host round trips per plane instead of ~280).
// Classes are recomputed per sample from the luma surface (cal_filter_support, Common/Codec/EbCcso.c:238-259, taps from
// derive_ccso_sample_pos :204-234, replicate border = clamped coordinates) instead of materialising class planes.
#pragma once
#include "svtf_kernels.cuh"

namespace svtf {

constexpr int CD_THREADS    = 256;
constexpr int CD_MAX_LUT    = 8;
constexpr int CD_MAX_GROUPS = 49; // 1 band-only group + 6 supports x 4 quantisers x 2 classifiers

// One configuration group of a batched launch (blockIdx.z): the classifier of the group, where its histograms go, and for the
// trial kernel its LUTs — slot z uses luts[z][CD_MAX_LUT][2048] and dist[z][CD_MAX_LUT][nfb].
struct CdGroup {
    int dy, dx, thr, edge_clf, bo;
    int hist_shift, nidx, hoff; // hoff: offset of the group's [nfb][nidx] histograms in herr / hcnt
    int nl, shift[CD_MAX_LUT];
};

struct CcsoDeriveArgs {
    const void *luma, *rec, *org; // classifier luma plane; plane to modify; source plane
    int         luma_pitch, rec_pitch, org_pitch;
    int         W, H, pw, ph, ss;   // luma picture, plane, subsampling shift
    int         blk_log2, nbx;      // filter-block size (8 luma / 7 chroma) and blocks per row
    int         dy, dx;             // first classifier tap (second is mirrored)
    int         thr, edge_clf, bo;  // quantiser step, classifier, band-only
    int         max_val;
    // hist
    int  hist_shift, nidx; // finest band shift; entries per block (128 edge groups / 2048 band-only)
    int *herr, *hcnt;      // [nfb][nidx]
    // try
    int                 nl;
    int                 shift[CD_MAX_LUT];
    const int8_t       *luts; // [nl][2048]
    unsigned long long *dist; // [nl][nfb]
    // batched form: gridDim.z groups, the fields above that a group owns come from groups[blockIdx.z]
    const CdGroup *groups;
};

// the launch arguments as group blockIdx.z sees them
__device__ __forceinline__ CcsoDeriveArgs cd_group_args(const CcsoDeriveArgs &a) {
    CcsoDeriveArgs b = a;
    if (a.groups) {
        const CdGroup &g   = a.groups[blockIdx.z];
        const int      nfb = a.nbx * ((a.ph + (1 << a.blk_log2) - 1) >> a.blk_log2);
        b.dy = g.dy, b.dx = g.dx, b.thr = g.thr, b.edge_clf = g.edge_clf, b.bo = g.bo;
        b.hist_shift = g.hist_shift, b.nidx = g.nidx, b.herr = a.herr + g.hoff, b.hcnt = a.hcnt + g.hoff;
        b.nl = g.nl;
#pragma unroll
        for (int l = 0; l < CD_MAX_LUT; l++) b.shift[l] = g.shift[l];
        b.luts = a.luts + (size_t)blockIdx.z * CD_MAX_LUT * 2048;
        b.dist = a.dist + (size_t)blockIdx.z * CD_MAX_LUT * nfb;
    }
    return b;
}

template <typename T>
__device__ __forceinline__ void cd_sample(const CcsoDeriveArgs &a, const int x, const int y, int &yv, int &c0, int &c1) {
    const T  *l  = static_cast<const T *>(a.luma);
    const int lx = x << a.ss, ly = y << a.ss;
    yv = l[(size_t)ly * a.luma_pitch + lx];
    c0 = c1 = 0;
    if (!a.bo) {
        const int d0 = (int)l[(size_t)min(max(ly + a.dy, 0), a.H - 1) * a.luma_pitch + min(max(lx + a.dx, 0), a.W - 1)] - yv;
        const int d1 = (int)l[(size_t)min(max(ly - a.dy, 0), a.H - 1) * a.luma_pitch + min(max(lx - a.dx, 0), a.W - 1)] - yv;
        if (a.edge_clf == 0) {
            c0 = d0 > a.thr ? 2 : (d0 < -a.thr ? 0 : 1);
            c1 = d1 > a.thr ? 2 : (d1 < -a.thr ? 0 : 1);
        } else {
            c0 = d0 < -a.thr ? 0 : 1;
            c1 = d1 < -a.thr ? 0 : 1;
        }
    }
}

// CTA = one 64x64 tile of the plane (always inside one filter block: blocks are 128 or 256 samples).
template <typename T>
__global__ void __launch_bounds__(CD_THREADS) ccso_hist_kernel(const __grid_constant__ CcsoDeriveArgs a0) {
    const CcsoDeriveArgs a = cd_group_args(a0);
    __shared__ int s_err[2048], s_cnt[2048];
    const int tid = threadIdx.x, x0 = blockIdx.x * 64, y0 = blockIdx.y * 64;
    for (int i = tid; i < a.nidx; i += CD_THREADS) s_err[i] = 0, s_cnt[i] = 0;
    __syncthreads();
    const T *rec = static_cast<const T *>(a.rec), *org = static_cast<const T *>(a.org);
    for (int i = tid; i < 64 * 64; i += CD_THREADS) {
        const int x = x0 + (i & 63), y = y0 + (i >> 6);
        if (x >= a.pw || y >= a.ph)
            continue;
        int yv, c0, c1;
        cd_sample<T>(a, x, y, yv, c0, c1);
        const int idx = ((yv >> a.hist_shift) << 4) + (c0 << 2) + c1;
        atomicAdd(&s_err[idx], (int)org[(size_t)y * a.org_pitch + x] - (int)rec[(size_t)y * a.rec_pitch + x]);
        atomicAdd(&s_cnt[idx], 1);
    }
    __syncthreads();
    const int fb = (y0 >> a.blk_log2) * a.nbx + (x0 >> a.blk_log2);
    for (int i = tid; i < a.nidx; i += CD_THREADS)
        if (s_cnt[i]) {
            atomicAdd(&a.herr[(size_t)fb * a.nidx + i], s_err[i]);
            atomicAdd(&a.hcnt[(size_t)fb * a.nidx + i], s_cnt[i]);
        }
}

template <typename T>
__global__ void __launch_bounds__(CD_THREADS) ccso_try_kernel(const __grid_constant__ CcsoDeriveArgs a0) {
    const CcsoDeriveArgs a = cd_group_args(a0);
    __shared__ int8_t             s_lut[CD_MAX_LUT][2048];
    __shared__ unsigned long long s_sum[CD_MAX_LUT];
    const int tid = threadIdx.x, x0 = blockIdx.x * 64, y0 = blockIdx.y * 64;
    for (int i = tid; i < a.nl * 512; i += CD_THREADS)
        reinterpret_cast<uint32_t *>(&s_lut[0][0])[i] = __ldg(reinterpret_cast<const uint32_t *>(a.luts) + i);
    if (tid < CD_MAX_LUT)
        s_sum[tid] = 0;
    __syncthreads();
    const T *rec = static_cast<const T *>(a.rec), *org = static_cast<const T *>(a.org);
    unsigned sse[CD_MAX_LUT];
#pragma unroll
    for (int l = 0; l < CD_MAX_LUT; l++) sse[l] = 0;
    for (int i = tid; i < 64 * 64; i += CD_THREADS) { // 16 samples per thread: 16 * 4095^2 < 2^32
        const int x = x0 + (i & 63), y = y0 + (i >> 6);
        if (x >= a.pw || y >= a.ph)
            continue;
        int yv, c0, c1;
        cd_sample<T>(a, x, y, yv, c0, c1);
        const int r = rec[(size_t)y * a.rec_pitch + x], o = org[(size_t)y * a.org_pitch + x], cl = (c0 << 2) + c1;
#pragma unroll
        for (int l = 0; l < CD_MAX_LUT; l++)
            if (l < a.nl) {
                const int v = min(max(r + s_lut[l][((yv >> a.shift[l]) << 4) + cl], 0), a.max_val);
                const int e = o - v;
                sse[l] += (unsigned)(e * e);
            }
    }
#pragma unroll
    for (int l = 0; l < CD_MAX_LUT; l++)
        if (l < a.nl) {
            unsigned long long v = sse[l];
#pragma unroll
            for (int m = 16; m > 0; m >>= 1) v += __shfl_xor_sync(0xffffffffu, v, m);
            if ((tid & 31) == 0)
                atomicAdd(&s_sum[l], v);
        }
    __syncthreads();
    const int fb = (y0 >> a.blk_log2) * a.nbx + (x0 >> a.blk_log2);
    const int nfb = a.nbx * ((a.ph + (1 << a.blk_log2) - 1) >> a.blk_log2);
    if (tid < a.nl && s_sum[tid])
        atomicAdd(&a.dist[(size_t)tid * nfb + fb], s_sum[tid]);
}

} // namespace svtf
