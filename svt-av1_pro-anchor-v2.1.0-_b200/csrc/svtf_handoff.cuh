// svtf_handoff.cuh — recon hand-off: the filtered picture leaves the device once, in the layout the
// reference's later stages read.
//
//   handoff_kernel<T>  pad_ref_and_set_flags (Encoder/Codec/EbEncDecProcess.c:1589-1760): right/bottom replication of the
//                      last valid column/row (svt_aom_pad_input_picture_16bit, Common/Codec/EbMcp.c:296), border
//                      replication by org_x/org_y (svt_aom_generate_padding[16_bit], Common/Codec/EbMcp.c:95/201) and
//                      the split of 10-bit samples into the 8-bit + 2-bit planes mode decision reads
//                      (svt_aom_un_pack2d -> svt_enc_msb_un_pack2_d, Common/C_DEFAULT/EbPackUnPack_C.c:202), or the
//                      narrowing svt_convert_16bit_to_8bit_c (:277) for 8-bit content in the 16-bit pipeline.
//                      All of it is one pure function of the resident surface:
//                          out[y][x] = S[clamp(y - org_y, 0, valid_h-1)][clamp(x - org_x, 0, valid_w-1)]
//                      so one pass writes the padded planes; HBM-bound (read 1 sample, write 1..2 per output sample).
//   widen_kernel       svt_convert_pic_8bit_to_16bit (Encoder/Codec/EbRestProcess.c:235, call site
//                      Encoder/Codec/EbDlfProcess.c:81-95): 8-bit planes uploaded as bytes, widened on the device.
#pragma once
#include "svtf_kernels.cuh"

namespace svtf {

struct HandoffPlane {
    const void *src; // resident surface plane
    int         src_pitch;
    int         valid_w, valid_h; // samples that exist (everything else replicates them)
    int         org_x, org_y;     // border
    int         out_w, out_h;     // padded size = width + 2*org_x, height + 2*org_y
    void       *dst;              // device, pitch dpitch samples of T
    uint8_t    *dst8, *dstn;      // device or NULL, pitch dpitch bytes
    int         dpitch;           // multiple of 8
    int         chunks;           // dpitch / 8
    int         first_item;       // first linear work item of this plane
};
struct HandoffArgs {
    HandoffPlane p[3];
    int          total;  // work items (8-sample chunks) over the three planes
    int          shift8; // dst8 = sample >> shift8 (2 for 10-bit, 0 for 8-bit content); dstn = (uint8_t)(sample << 6)
};

template <typename T>
__global__ void __launch_bounds__(256) handoff_kernel(const __grid_constant__ HandoffArgs a) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < a.total; i += gridDim.x * blockDim.x) {
        const int           pl = i >= a.p[2].first_item ? 2 : (i >= a.p[1].first_item ? 1 : 0);
        const HandoffPlane &q  = a.p[pl];
        const int           li = i - q.first_item;
        const int           y = li / q.chunks, x = (li - y * q.chunks) * 8;
        if (x >= q.out_w)
            continue;
        const int sy = min(max(y - q.org_y, 0), q.valid_h - 1), sx = x - q.org_x;
        const T  *row = static_cast<const T *>(q.src) + (size_t)sy * q.src_pitch;
        uint32_t  w[4];
        if (sx >= 0 && sx + 8 <= q.valid_w && (sx & 7) == 0) {
            const uint4 v = ld8<T>(row + sx);
            w[0] = v.x, w[1] = v.y, w[2] = v.z, w[3] = v.w;
        } else {
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const uint32_t lo = row[min(max(sx + 2 * k, 0), q.valid_w - 1)], hi = row[min(max(sx + 2 * k + 1, 0), q.valid_w - 1)];
                w[k]              = lo | (hi << 16);
            }
        }
        const size_t o = (size_t)y * q.dpitch + x;
        st8<T>(static_cast<T *>(q.dst) + o, make_uint4(w[0], w[1], w[2], w[3]));
        if (q.dst8) {
            uint32_t s[4];
#pragma unroll
            for (int k = 0; k < 4; k++) s[k] = (w[k] >> a.shift8) & 0x00FF00FFu;
            *reinterpret_cast<uint2 *>(q.dst8 + o) = make_uint2(__byte_perm(s[0], s[1], 0x6420), __byte_perm(s[2], s[3], 0x6420));
        }
        if (q.dstn) {
            uint32_t s[4];
#pragma unroll
            for (int k = 0; k < 4; k++) s[k] = (w[k] << 6) & 0x00C000C0u;
            *reinterpret_cast<uint2 *>(q.dstn + o) = make_uint2(__byte_perm(s[0], s[1], 0x6420), __byte_perm(s[2], s[3], 0x6420));
        }
    }
}

struct WidenArgs {
    const uint8_t *src[3]; // device bytes, pitch spitch
    uint16_t      *dst[3];
    int            spitch[3], dpitch[3], w[3], h[3], first_item[3], chunks[3];
    int            total;
};
__global__ void __launch_bounds__(256) widen_kernel(const __grid_constant__ WidenArgs a) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < a.total; i += gridDim.x * blockDim.x) {
        const int pl = i >= a.first_item[2] ? 2 : (i >= a.first_item[1] ? 1 : 0);
        const int li = i - a.first_item[pl];
        const int y = li / a.chunks[pl], x = (li - y * a.chunks[pl]) * 8;
        if (x >= a.w[pl])
            continue;
        const uint4 v = ld8<uint8_t>(a.src[pl] + (size_t)y * a.spitch[pl] + x); // spitch is a multiple of 8
        st8<uint16_t>(a.dst[pl] + (size_t)y * a.dpitch[pl] + x, v);
    }
}

} // namespace svtf
