// svtf_superres.cuh — the normative super-resolution upscale, the step between CDEF and loop restoration.
//
// Reference: svt_av1_upscale_normative_rows (Common/Codec/EbSuperRes.c:214), av1_[highbd_]convolve_horiz_rs_c (:60 / :80),
// upscale_normative_rect (:102).  Per plane and tile column j: output column x samples the coded picture at
// x_qn = x0_qn[j] + (x - up_x0[j]) * step (14 fractional bits): 8 taps starting at ds_x0[j] - 4 + (x_qn >> 14), filter phase
// (x_qn & 16383) >> 8; columns outside the coded picture repeat its edge (the reference pads 5 columns in place and restores
// them).  One thread per output sample; bandwidth-bound (read b, write b * denom / 8 bytes per coded sample).
#pragma once
#include <climits>
#include "svtf_kernels.cuh"

namespace svtf {

struct alignas(16) SuperresTables { // device memory, built by the host per call
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
    int                   row_base[4]; // blockIdx.y -> (plane, row)
    int                   x0_qn_first[3], up_x1_first[3]; // tile column 0: its sub-sample offset and its end (starts at 0 / 0)
};

// grid: x = 256-column chunks of the widest plane, y = every row of the three planes; one thread per output sample.  The
// start position of the first tile column travels in the kernel arguments (one tile column is the common case: no dependent
// table loads before the taps); the eight source samples come through L1 (neighbouring threads share them).  A version that
// staged the block's source window in shared memory was slower (88 against 61 us for 1920 -> 3840 x 2160 10-bit).
constexpr int SR_PER = 4; // consecutive outputs per thread: their loads are in flight together, one vector store
template <typename T>
__global__ void __launch_bounds__(256) superres_kernel(const __grid_constant__ SuperresArgs a) {
    const SuperresTables &tb = *a.tab;
    const int by = blockIdx.y, p = by >= a.row_base[2] ? 2 : (by >= a.row_base[1] ? 1 : 0), y = by - a.row_base[p];
    const int xo = (blockIdx.x * 256 + threadIdx.x) * SR_PER; // coded widths are multiples of 4
    if (xo >= a.dst_cw[p])
        return;
    const int last = a.src_w[p] - 1, maxv = (1 << a.bd) - 1;
    const T  *s = static_cast<const T *>(a.src.p[p]) + (size_t)y * a.src.pitch[p];
    int       b0[SR_PER];
    uint4     fv[SR_PER];
#pragma unroll
    for (int q = 0; q < SR_PER; q++) {
        const int x = min(xo + q, a.dst_w[p] - 1); // coded columns beyond the upscaled width repeat the last one
        int       x_qn;
        if (a.tile_cols == 1 || x < a.up_x1_first[p]) {
            x_qn  = a.x0_qn_first[p] + x * a.step[p];
            b0[q] = -4 + (x_qn >> 14);
        } else {
            int j = 1;
            while (j + 1 < a.tile_cols && x >= tb.up_x0[p][j + 1]) j++;
            x_qn  = tb.x0_qn[p][j] + (x - tb.up_x0[p][j]) * a.step[p];
            b0[q] = tb.ds_x0[p][j] - 4 + (x_qn >> 14);
        }
        fv[q] = __ldg(reinterpret_cast<const uint4 *>(tb.filt + ((x_qn & 16383) >> 8) * 8)); // the 8 taps of this phase
    }
    uint32_t out[SR_PER];
#pragma unroll
    for (int q = 0; q < SR_PER; q++) {
        const int f[8] = {(short)(fv[q].x & 0xffff), (short)(fv[q].x >> 16), (short)(fv[q].y & 0xffff), (short)(fv[q].y >> 16),
                          (short)(fv[q].z & 0xffff), (short)(fv[q].z >> 16), (short)(fv[q].w & 0xffff), (short)(fv[q].w >> 16)};
        int       sum = 0;
        if (b0[q] >= 0 && b0[q] + 7 <= last) {
#pragma unroll
            for (int k = 0; k < 8; k++) sum += (int)__ldg(s + b0[q] + k) * f[k];
        } else {
#pragma unroll
            for (int k = 0; k < 8; k++) sum += (int)__ldg(s + min(max(b0[q] + k, 0), last)) * f[k];
        }
        out[q] = (uint32_t)min(max((sum + 64) >> 7, 0), maxv);
    }
    T *d = static_cast<T *>(a.dst.p[p]) + (size_t)y * a.dst.pitch[p] + xo;
    if (sizeof(T) == 2)
        *reinterpret_cast<uint2 *>(d) = make_uint2(out[0] | (out[1] << 16), out[2] | (out[3] << 16));
    else
        *reinterpret_cast<uint32_t *>(d) = out[0] | (out[1] << 8) | (out[2] << 16) | (out[3] << 24);
}

} // namespace svtf
