// svtf_superres.cuh — the normative super-resolution upscale, the step between CDEF and loop restoration.
//
// Reference: svt_av1_upscale_normative_rows (Common/Codec/EbSuperRes.c:214), av1_[highbd_]convolve_horiz_rs_c (:60 / :80),
// upscale_normative_rect (:102).  Per plane and tile column j: output column x samples the coded picture at
// x_qn = x0_qn[j] + (x - up_x0[j]) * step (14 fractional bits): 8 taps starting at ds_x0[j] - 4 + (x_qn >> 14), filter phase
// (x_qn & 16383) >> 8; columns outside the coded picture repeat its edge (the reference pads 5 columns in place and restores
// them).  One thread per output sample; bandwidth-bound (read b, write b * denom / 8 bytes per coded sample).
#pragma once
#include "svtf_kernels.cuh"

namespace svtf {

struct SuperresTables { // device memory, built by the host per call
    int16_t filt[64 * 8];
    int32_t up_x0[3][65]; // first output column of tile column j (entry tile_cols = the upscaled plane width)
    int32_t ds_x0[3][64]; // first coded column of tile column j
    int32_t x0_qn[3][64]; // sub-sample offset at the start of tile column j
};
struct SuperresArgs {
    Planes                src, dst;
    const SuperresTables *tab;
    int                   src_w[3], dst_w[3], dst_cw[3], rows[3]; // coded source width, upscaled width, coded upscaled width, rows
    int                   step[3], tile_cols, bd;
    long long             base[4]; // linearised (plane, row, column) items
};

template <typename T>
__global__ void __launch_bounds__(256) superres_kernel(const __grid_constant__ SuperresArgs a) {
    const SuperresTables &tb = *a.tab;
    for (long long i = blockIdx.x * 256ll + threadIdx.x; i < a.base[3]; i += (long long)gridDim.x * 256) {
        const int p  = i >= a.base[2] ? 2 : (i >= a.base[1] ? 1 : 0);
        const int li = (int)(i - a.base[p]), y = li / a.dst_cw[p], xo = li - y * a.dst_cw[p];
        const int x  = min(xo, a.dst_w[p] - 1); // coded columns beyond the upscaled width repeat the last one
        int       j  = 0;
        while (j + 1 < a.tile_cols && x >= tb.up_x0[p][j + 1]) j++;
        const int      x_qn = tb.x0_qn[p][j] + (x - tb.up_x0[p][j]) * a.step[p];
        const int      b0 = tb.ds_x0[p][j] - 4 + (x_qn >> 14);
        const int16_t *f  = tb.filt + ((x_qn & 16383) >> 8) * 8;
        const T       *s  = static_cast<const T *>(a.src.p[p]) + (size_t)y * a.src.pitch[p];
        int            sum = 0;
#pragma unroll
        for (int k = 0; k < 8; k++) sum += (int)s[min(max(b0 + k, 0), a.src_w[p] - 1)] * f[k];
        const int v = min(max((sum + 64) >> 7, 0), (1 << a.bd) - 1);
        static_cast<T *>(a.dst.p[p])[(size_t)y * a.dst.pitch[p] + xo] = (T)v;
    }
}

} // namespace svtf
