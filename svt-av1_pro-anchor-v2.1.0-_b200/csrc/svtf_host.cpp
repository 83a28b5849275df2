// svtf_host.cpp — host-side logic of the filter path: mi packing, deblock level/threshold
// tables, and the frame-level CDEF strength pick.  Pure C++ (no CUDA), part of libsvtfilt_b200.so.
//
// Reference paths are relative to Source/Lib/ of the reference tree.
#include <algorithm>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "svtf_internal.h"

namespace {

// Block dimensions as log2(px)-2, indexed by the BlockSize enum
// (Common/Codec/EbDefinitions.h:770-795 order: 4x4,4x8,8x4,8x8,8x16,16x8,16x16,16x32,32x16,32x32,
//  32x64,64x32,64x64,64x128,128x64,128x128,4x16,16x4,8x32,32x8,16x64,64x16; dims at :1611-1615).
constexpr int kNumBlockSizes = 22;
const uint8_t kBlkW[kNumBlockSizes] = {0, 0, 1, 1, 1, 2, 2, 2, 3, 3, 3, 4, 4, 4, 5, 5, 0, 2, 1, 3, 2, 4};
const uint8_t kBlkH[kNumBlockSizes] = {0, 1, 0, 1, 2, 1, 2, 3, 2, 3, 4, 3, 4, 5, 4, 5, 2, 0, 3, 1, 4, 2};

// TxSize enum -> log2(dim)-2 (Common/Codec/EbDefinitions.h tx_size_wide/high :~1670-1690):
// 4x4,8x8,16x16,32x32,64x64,4x8,8x4,8x16,16x8,16x32,32x16,32x64,64x32,4x16,16x4,8x32,32x8,16x64,64x16
constexpr int kNumTxSizes = 19;
const uint8_t kTxW[kNumTxSizes] = {0, 1, 2, 3, 4, 0, 1, 1, 2, 2, 3, 3, 4, 0, 2, 1, 3, 2, 4};
const uint8_t kTxH[kNumTxSizes] = {0, 1, 2, 3, 4, 1, 0, 2, 1, 3, 2, 4, 3, 2, 0, 3, 1, 4, 2};

// Luma transform dims for (tx_depth, bsize): tx_depth_to_tx_size (Common/Codec/EbDefinitions.h:885-909),
// stored as {w,h} log2-2 pairs.  depth 0 = the largest (rectangular) transform, capped at 64.
struct WH {
    uint8_t w, h;
};
const WH kTxDepth[3][kNumBlockSizes] = {
    // depth 0
    {{0, 0}, {0, 1}, {1, 0}, {1, 1}, {1, 2}, {2, 1}, {2, 2}, {2, 3}, {3, 2}, {3, 3}, {3, 4}, {4, 3}, {4, 4}, {4, 4},
     {4, 4}, {4, 4}, {0, 2}, {2, 0}, {1, 3}, {3, 1}, {2, 4}, {4, 2}},
    // depth 1
    {{0, 0}, {0, 1}, {1, 0}, {0, 0}, {1, 1}, {1, 1}, {1, 1}, {2, 2}, {2, 2}, {2, 2}, {3, 3}, {3, 3}, {3, 3}, {4, 4},
     {4, 4}, {4, 4}, {0, 1}, {1, 0}, {1, 2}, {2, 1}, {2, 3}, {3, 2}},
    // depth 2
    {{0, 0}, {0, 1}, {1, 0}, {1, 1}, {0, 0}, {0, 0}, {0, 0}, {1, 1}, {1, 1}, {1, 1}, {2, 2}, {2, 2}, {2, 2}, {4, 4},
     {4, 4}, {4, 4}, {0, 0}, {0, 0}, {1, 1}, {1, 1}, {2, 2}, {2, 2}}};

// mode_lf_lut (Common/Codec/EbDeblockingCommon.h:62-66): 0 for intra modes, GLOBALMV and
// GLOBAL_GLOBALMV, 1 for the other inter modes.
inline int mode_lf(int mode) {
    if (mode < 13)
        return 0;
    if (mode == 15 || mode == 23)
        return 0;
    return 1;
}

// av1_get_max_uv_txsize(bsize, 1, 1) (Common/Codec/EbUtility.h:117-123): the largest transform of the
// 4:2:0 plane block (ss_size_lookup[.][1][1] == max(4, dim/2) on both axes), with 64 -> 32
// (av1_get_adjusted_tx_size, EbDefinitions.h:925-935).  Rectangular 1:4 plane blocks
// keep their 1:4 transform (max_txsize_rect_lookup).
inline WH uv_tx(int bsize) {
    int w = std::max(0, kBlkW[bsize] - 1), h = std::max(0, kBlkH[bsize] - 1);
    w     = std::min(w, 3);
    h     = std::min(h, 3);
    return {(uint8_t)w, (uint8_t)h};
}

inline svtf_mi_t pack(int txw_y, int txh_y, int txw_uv, int txh_uv, int bw, int bh, int skip_inter, int skip, int ref0,
                      int mlf, int seg) {
    return (svtf_mi_t)(txw_y | (txh_y << 3) | (txw_uv << 6) | (txh_uv << 9) | (bw << 12) | (bh << 15) |
                       (skip_inter << 18) | (skip << 19) | (ref0 << 20) | (mlf << 23) | (seg << 24));
}

} // namespace

extern "C" int svtf_pack_mi_enc(size_t n, const uint8_t *bsize, const uint8_t *tx_depth, const uint8_t *skip,
                                const uint8_t *ref_frame0, const uint8_t *mode, const uint8_t *segment_id,
                                svtf_mi_t *out) {
    if (!bsize || !out)
        return svtf::set_global_error(SVTF_E_ARG, "svtf_pack_mi_enc: NULL argument");
    for (size_t i = 0; i < n; i++) {
        const int bs = bsize[i];
        if (bs >= kNumBlockSizes)
            return svtf::set_global_error(SVTF_E_ARG, "svtf_pack_mi_enc: invalid block size");
        const int sk   = skip ? (skip[i] != 0) : 0;
        const int ref0 = ref_frame0 ? (ref_frame0[i] & 7) : 0;
        const int si   = sk && ref0 > 0; // skip && is_inter_block_no_intrabc (EbDeblockingFilter.c:198)
        const int td   = tx_depth ? std::min<int>(tx_depth[i], 2) : 0;
        const WH  ty   = kTxDepth[si ? 0 : td][bs];
        const WH  tuv  = uv_tx(bs);
        out[i] = pack(ty.w, ty.h, tuv.w, tuv.h, kBlkW[bs], kBlkH[bs], si, sk, ref0, mode ? mode_lf(mode[i]) : 0,
                      segment_id ? (segment_id[i] & 7) : 0);
    }
    return SVTF_OK;
}

extern "C" int svtf_pack_mi_dec(size_t n, const uint8_t *bsize, const uint8_t *tx_size_y, const uint8_t *tx_size_uv,
                                const uint8_t *skip, const uint8_t *ref_frame0, const uint8_t *mode,
                                const uint8_t *segment_id, svtf_mi_t *out) {
    if (!bsize || !tx_size_y || !tx_size_uv || !out)
        return svtf::set_global_error(SVTF_E_ARG, "svtf_pack_mi_dec: NULL argument");
    for (size_t i = 0; i < n; i++) {
        const int bs = bsize[i];
        if (bs >= kNumBlockSizes || tx_size_y[i] >= kNumTxSizes || tx_size_uv[i] >= kNumTxSizes)
            return svtf::set_global_error(SVTF_E_ARG, "svtf_pack_mi_dec: invalid enum value");
        const int sk   = skip ? (skip[i] != 0) : 0;
        const int ref0 = ref_frame0 ? (ref_frame0[i] & 7) : 0;
        const int si   = sk && ref0 > 0;
        out[i] = pack(kTxW[tx_size_y[i]], kTxH[tx_size_y[i]], kTxW[tx_size_uv[i]], kTxH[tx_size_uv[i]], kBlkW[bs],
                      kBlkH[bs], si, sk, ref0, mode ? mode_lf(mode[i]) : 0, segment_id ? (segment_id[i] & 7) : 0);
    }
    return SVTF_OK;
}

// count_units_in_tile (Common/Codec/EbRestoration.c:124) over the whole-frame rectangle (svt_aom_whole_frame_rect :104)
extern "C" int svtf_lr_units(const svtf_geom *geom, int32_t plane, int32_t unit_size, int32_t *horz_units, int32_t *vert_units) {
    if (!geom || !horz_units || !vert_units || unit_size <= 0 || plane < 0 || plane > 2)
        return svtf::set_global_error(SVTF_E_ARG, "svtf_lr_units: bad argument");
    const int ss = plane ? 1 : 0, pw = (geom->width + ss) >> ss, ph = (geom->height + ss) >> ss;
    *horz_units  = std::max((pw + (unit_size >> 1)) / unit_size, 1);
    *vert_units  = std::max((ph + (unit_size >> 1)) / unit_size, 1);
    return SVTF_OK;
}

namespace svtf {

// Per-frame deblock tables.  Level table: svt_av1_loop_filter_frame_init
// (Common/Codec/EbDeblockingCommon.c:76-137) — or svt_aom_get_filter_level_delta_lf (:37-71) for every
// (plane, seg, dir, ref, mode_lf) key when delta_lf_present (its sb_delta_lf argument is the
// frame-level ppcs->curr_delta_lf).  Thresholds: svt_aom_update_sharpness (:554-573) and
// hev_thr = lvl >> 4 (Encoder/Codec/EbDeblockingFilter.c:47).
void build_lf_tables(const svtf_lf_params &lf, LfTables &t) {
    static const int seg_feature[3][2] = {{1, 2}, {3, 3}, {4, 4}}; // SEG_LVL_ALT_LF_{Y_V,Y_H,U,V}
    static const int delta_id[3][2]    = {{0, 1}, {2, 2}, {3, 3}};
    std::memset(&t, 0, sizeof(t));
    const int sharp = lf.sharpness_level;
    for (int lvl = 0; lvl < 64; lvl++) {
        int inside = lvl >> ((sharp > 0) + (sharp > 4));
        if (sharp > 0)
            inside = std::min(inside, 9 - sharp);
        inside       = std::max(inside, 1);
        t.lim[lvl]   = (uint8_t)inside;
        t.mblim[lvl] = (uint8_t)(2 * (lvl + 2) + inside);
        t.hev[lvl]   = (uint8_t)(lvl >> 4);
    }
    auto clamp63 = [](int v) { return std::min(std::max(v, 0), 63); };
    for (int plane = 0; plane < 3; plane++)
        for (int dir = 0; dir < 2; dir++) {
            int base = plane == 0 ? lf.filter_level[dir] : (plane == 1 ? lf.filter_level_u : lf.filter_level_v);
            if (lf.delta_lf_present)
                base = clamp63(base + (lf.delta_lf_multi ? lf.curr_delta_lf[delta_id[plane][dir]] : lf.curr_delta_lf[0]));
            for (int seg = 0; seg < 8; seg++) {
                int       lvl_seg = base;
                const int f       = seg_feature[plane][dir];
                if (lf.segmentation_enabled && lf.seg_feature_enabled[seg][f])
                    lvl_seg = clamp63(lvl_seg + lf.seg_feature_data[seg][f]);
                for (int ref = 0; ref < 8; ref++)
                    for (int m = 0; m < 2; m++) {
                        int v = lvl_seg;
                        if (lf.mode_ref_delta_enabled) {
                            const int scale = 1 << (lvl_seg >> 5);
                            v += lf.ref_deltas[ref] * scale;
                            if (ref > 0)
                                v += lf.mode_deltas[m] * scale;
                            v = clamp63(v);
                        }
                        t.lvl[plane][seg][dir][ref][m] = (uint8_t)v;
                    }
            }
        }
}

} // namespace svtf

extern "C" int svtf_lf_build_tables(const svtf_lf_params *lf, uint8_t lvl[768], uint8_t lim[64], uint8_t mblim[64],
                                    uint8_t hev[64]) {
    if (!lf || !lvl || !lim || !mblim || !hev)
        return svtf::set_global_error(SVTF_E_ARG, "svtf_lf_build_tables: NULL argument");
    svtf::LfTables t;
    svtf::build_lf_tables(*lf, t);
    std::memcpy(lvl, t.lvl, 768);
    std::memcpy(lim, t.lim, 64);
    std::memcpy(mblim, t.mblim, 64);
    std::memcpy(hev, t.hev, 64);
    return SVTF_OK;
}

// ---------------------------------------------------------------------------------------------
// Frame-level strength pick: finish_cdef_search (Encoder/Codec/EbEncCdef.c:728-925),
// joint_strength_search_dual (:697-727), svt_search_one_dual_c (:627-684).
// ---------------------------------------------------------------------------------------------
namespace {

struct PickTable {
    std::vector<const uint64_t *> m0, m1; // per searched fb: luma / chroma rows of 64
    std::vector<int>              fb;
};

uint64_t search_one_dual(int *lev0, int *lev1, int nb, const PickTable &t, int end_gi, std::vector<uint64_t> &tot) {
    const uint64_t kInf = (uint64_t)1 << 63;
    std::fill(tot.begin(), tot.end(), 0);
    const int count = (int)t.fb.size();
    for (int i = 0; i < count; i++) {
        const uint64_t *a = t.m0[i], *b = t.m1[i];
        uint64_t        best = kInf;
        for (int g = 0; g < nb; g++) best = std::min(best, a[lev0[g]] + b[lev1[g]]);
        for (int j = 0; j < end_gi; j++) {
            uint64_t *row = &tot[(size_t)j * 64];
            for (int k = 0; k < end_gi; k++) row[k] += std::min(best, a[j] + b[k]);
        }
    }
    uint64_t best_tot = kInf;
    int      b0 = 0, b1 = 0;
    for (int j = 0; j < end_gi; j++)
        for (int k = 0; k < end_gi; k++)
            if (tot[(size_t)j * 64 + k] < best_tot) { // strict: first index wins ties
                best_tot = tot[(size_t)j * 64 + k];
                b0       = j;
                b1       = k;
            }
    lev0[nb] = b0;
    lev1[nb] = b1;
    return best_tot;
}

} // namespace

extern "C" int svtf_cdef_pick_host(const svtf_cdef_search_cfg *cfg, const uint64_t *mse_in, const uint8_t *skip, int32_t nfb,
                                   svtf_cdef_params *out, int8_t *fb_strength_idx) {
    if (!cfg || !mse_in || !skip || !out || !fb_strength_idx || nfb <= 0)
        return svtf::set_global_error(SVTF_E_ARG, "svtf_cdef_pick_host: bad argument");
    const int end_gi = cfg->first_pass_fs_num + cfg->default_second_pass_fs_num;
    if (end_gi <= 0 || end_gi > SVTF_TOTAL_STRENGTHS)
        return svtf::set_global_error(SVTF_E_ARG, "svtf_cdef_pick_host: bad strength counts");
    // the bias rescales entry 0 in place in the reference; work on a private copy of column 0
    std::vector<uint64_t> mse(mse_in, mse_in + (size_t)2 * nfb * 64);
    PickTable             t;
    for (int fb = 0; fb < nfb; fb++) {
        fb_strength_idx[fb] = -1;
        if (skip[fb])
            continue;
        t.m0.push_back(&mse[(size_t)fb * 64]);
        t.m1.push_back(&mse[((size_t)nfb + fb) * 64]);
        t.fb.push_back(fb);
    }
    const int count = (int)t.fb.size();
    if (cfg->zero_fs_cost_bias) {
        const uint16_t factor = (uint16_t)cfg->zero_fs_cost_bias;
        for (int i = 0; i < count; i++) {
            const_cast<uint64_t *>(t.m0[i])[0] = (factor * t.m0[i][0]) >> 6;
            const_cast<uint64_t *>(t.m1[i])[0] = (factor * t.m1[i][0]) >> 6;
        }
    }
    std::memset(out, 0, sizeof(*out));
    std::vector<uint64_t> tot(64 * 64);
    uint64_t              best_cost = (uint64_t)1 << 63;
    int                   best_bits = 0;
    for (int bits = 0; bits <= 3; bits++) {
        int       l0[SVTF_CDEF_MAX_STRENGTHS] = {0}, l1[SVTF_CDEF_MAX_STRENGTHS] = {0};
        const int nb      = 1 << bits;
        uint64_t  tot_mse = (uint64_t)1 << 63;
        for (int k = 0; k < nb; k++) tot_mse = search_one_dual(l0, l1, k, t, end_gi, tot);
        for (int k = 0; k < 4 * nb; k++) { // refinement: rotate the list, re-pick the last entry
            for (int j = 0; j < nb - 1; j++) {
                l0[j] = l0[j + 1];
                l1[j] = l1[j + 1];
            }
            tot_mse = search_one_dual(l0, l1, nb - 1, t, end_gi, tot);
        }
        // RDCOST(lambda, av1_cost_literal(total_bits), tot_mse*16)
        // (Encoder/Codec/EbRateDistortionCost.h:37-39, EbMdRateEstimation.h:33-35)
        const int      total_bits = count * bits + nb * 6 * 2;
        const int64_t  rate       = (int64_t)total_bits * 512;
        const uint64_t dist       = tot_mse * 16;
        const uint64_t cost       = (uint64_t)(((rate * (int64_t)cfg->lambda + 256) >> 9) + (int64_t)dist * 128);
        if (cost < best_cost) {
            best_cost = cost;
            best_bits = bits;
            for (int j = 0; j < nb; j++) {
                out->cdef_y_strength[j]  = (uint8_t)l0[j];
                out->cdef_uv_strength[j] = (uint8_t)l1[j];
            }
        }
    }
    out->cdef_bits = best_bits;
    const int nb   = 1 << best_bits;
    for (int i = 0; i < count; i++) {
        uint64_t best = (uint64_t)1 << 63;
        int      bg   = 0;
        for (int g = 0; g < nb; g++) {
            const uint64_t c = t.m0[i][out->cdef_y_strength[g]] + t.m1[i][out->cdef_uv_strength[g]];
            if (c < best) {
                best = c;
                bg   = g;
            }
        }
        fb_strength_idx[t.fb[i]] = (int8_t)bg;
    }
    for (int g = 0; g < nb; g++) { // index -> strength byte through the luma lists (:911-919)
        auto map = [&](int i) {
            return i < cfg->first_pass_fs_num ? cfg->default_first_pass_fs[i]
                                              : cfg->default_second_pass_fs[i - cfg->first_pass_fs_num];
        };
        out->cdef_y_strength[g]  = map(out->cdef_y_strength[g]);
        out->cdef_uv_strength[g] = map(out->cdef_uv_strength[g]);
    }
    out->cdef_damping = 3 + (cfg->base_q_idx >> 6);
    return SVTF_OK;
}
