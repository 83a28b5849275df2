// svtf_mi_pack.cuh — side-info packing on the device.
//
// The filters read one packed word per 4x4 (svtf_mi_t, include/svtfilt_b200.h).  The producers of that information —
// EncDec's per-block mode-info updates (Encoder/Codec/EbCodingLoop.c, BlockModeInfoEnc in EbCodingUnit.h:66-68) and the
// decoder's parser (Decoder/Codec/EbDecParseFrame.c, LfCtxt fill EbDecLF.c:63-90) — know it as a handful of byte-sized
// fields per block.  Shipping those bytes as separate planes (1 B per field, 6 or 7 B per 4x4) and packing here replaces
// the host pass over the picture (svtf_pack_mi_enc/_dec: 2.7 ms per 4K picture on one core, against 0.10 ms for the
// planes' DMA + this kernel) at the price of 6-7 B instead of 4 B per 4x4 over PCIe.  The geometry bits of the word come from tables the HOST builds by calling
// svtf_pack_mi_enc/_dec on every (block size, transform, skip) combination, so the device words are the host words by
// construction (tests/test_gpu_mi_soa.py compares them).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>
#include "../../include/svtfilt_b200.h"

namespace svtf {

constexpr int MP_THREADS = 256;

struct MiPackTables {
    uint32_t enc_geom[2][3][22]; // [skip && inter][tx_depth][bsize] -> bits 0-17
    uint32_t blk_geom[22];       // decoder: block dims only (bits 12-17)
    uint32_t tx_y[19], tx_uv[19]; // decoder: TxSize -> bits 0-5 / 6-11
    uint32_t mode_lf;            // bit m = mode_lf_lut[m]
};

struct MiPackArgs {
    const uint8_t *bsize, *tx0, *tx1, *skip, *ref0, *mode, *seg; // device planes, pitch `in_pitch`; tx1 only for the decoder
    int            in_pitch, rows, cols;
    svtf_mi_t     *out;
    int            out_pitch; // in words
    int            dec;
    int           *bad;       // set to 1 on an out-of-range enum
    MiPackTables   t;
};

__global__ void __launch_bounds__(MP_THREADS) mi_pack_kernel(const __grid_constant__ MiPackArgs a) {
    const int c = blockIdx.x * MP_THREADS + threadIdx.x, r = blockIdx.y;
    if (c >= a.cols)
        return;
    const size_t i  = (size_t)r * a.in_pitch + c;
    const int    bs = a.bsize[i], sk = a.skip[i] != 0, ref0 = a.ref0[i] & 7, md = a.mode[i], sg = a.seg[i] & 7;
    const int    si = sk && ref0 > 0; // skip && is_inter_block (Encoder/Codec/EbDeblockingFilter.c:198, Decoder/Codec/EbDecLF.c:166)
    uint32_t     w  = 0;
    bool         ok = bs < 22;
    if (a.dec) {
        const int ty = a.tx0[i], tuv = a.tx1[i];
        ok           = ok && ty < 19 && tuv < 19;
        if (ok)
            w = a.t.tx_y[ty] | a.t.tx_uv[tuv] | a.t.blk_geom[bs];
    } else if (ok)
        w = a.t.enc_geom[si][min((int)a.tx0[i], 2)][bs];
    if (!ok) {
        *a.bad = 1;
        return;
    }
    w |= (uint32_t)si << 18 | (uint32_t)sk << 19 | (uint32_t)ref0 << 20 | (md < 32 ? (a.t.mode_lf >> md) & 1u : 1u) << 23 | (uint32_t)sg << 24;
    a.out[(size_t)r * a.out_pitch + c] = w;
}

} // namespace svtf
