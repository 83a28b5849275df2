/*
 * ref_shim.c — drives the REAL reference functions (compiled from /root/reference where they lie,
 * see oracle/Makefile) through a plain-pointer interface, so tests can compare them with
 * oracle/svtf_oracle.c (the restatement) and with the CUDA path.
 *
 * TEST INFRASTRUCTURE ONLY (same rule as svtf_oracle.c).  This file contains no reference source:
 * it includes the reference's headers, fills the few PictureControlSet / SequenceControlSet fields
 * the frame-level functions read, defines the RTCD function pointers as the `_c`
 * kernels (what svt_aom_setup_common_rtcd_internal(0) would do) and stubs the pipeline symbols the
 * hot-path objects reference but never call on these paths.
 *
 * Built into oracle/_ref/libsvtref.so together with the reference objects:
 *   Common/Codec/EbCdef.c EbDeblockingCommon.c EbCcso.c, Encoder/Codec/EbDeblockingFilter.c
 *   EbEncCdef.c EbPickccso.c, and EbCdefProcess.c (included below as a TU to reach the static
 *   cdef_seg_search).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <stdio.h>

#include "EbPictureControlSet.h"
#include "EbSequenceControlSet.h"
#include "EbEncCdef.h"
#include "EbCdef.h"
#include "EbDeblockingFilter.h"
#include "EbDeblockingCommon.h"
#include "EbCcso.h"
#include "EbPickccso.h"
#include "EbReferenceObject.h"
#include "EbModeDecisionProcess.h"

/* pull the static cdef_seg_search() into this TU; the file stays where it lies */
#include "EbCdefProcess.c"

#include "../include/svtfilt_b200.h"

#define REF_API __attribute__((visibility("default")))

/* ---- RTCD pointers (Common/Codec/common_dsp_rtcd.c:626-658, Encoder/Codec/aom_dsp_rtcd.c:187,188,478) ----
 * Default build (libsvtref.so, the parity checker): the `_c` kernels, what svt_aom_setup_common_rtcd_internal(0) selects.
 * -DSVTF_REF_SIMD=1 (libsvtref_simd.so, the TIMED CPU baseline): the variants the same tables select on an AVX2 host —
 * SET_SSE2 for the loop-filter edges, SET_SSE41_AVX2 / SET_AVX2 for CDEF, the CDEF distortion, the strength pick and CCSO.
 * Their sources (Common/ASM_AVX2/cdef_block_avx2.c, ccso_avx2.c, Common/ASM_SSE2/EbDeblockingFilter_Intrinsic_SSE2.c,
 * Encoder/ASM_AVX2/EbCdef_AVX2.c) are C intrinsics and are compiled where they lie by oracle/Makefile. */
#ifndef SVTF_REF_SIMD
#define SVTF_REF_SIMD 0
#endif
#if SVTF_REF_SIMD
#define LPF_SFX(n) n##_sse2
#define AVX2_OR_C(n_c, n_avx2) n_avx2
#else
#define LPF_SFX(n) n##_c
#define AVX2_OR_C(n_c, n_avx2) n_c
#endif
REF_API int ref_is_simd(void) { return SVTF_REF_SIMD; }
#define LPF_DECL(n)                                                                                                   \
    void LPF_SFX(svt_aom_lpf_##n)(uint8_t *, int32_t, const uint8_t *, const uint8_t *, const uint8_t *);             \
    void (*svt_aom_lpf_##n)(uint8_t *, int32_t, const uint8_t *, const uint8_t *, const uint8_t *) = LPF_SFX(svt_aom_lpf_##n); \
    void LPF_SFX(svt_aom_highbd_lpf_##n)(uint16_t *, int32_t, const uint8_t *, const uint8_t *, const uint8_t *, int32_t); \
    void (*svt_aom_highbd_lpf_##n)(uint16_t *, int32_t, const uint8_t *, const uint8_t *, const uint8_t *, int32_t) = \
        LPF_SFX(svt_aom_highbd_lpf_##n);
LPF_DECL(horizontal_4)
LPF_DECL(horizontal_6)
LPF_DECL(horizontal_8)
LPF_DECL(horizontal_14)
LPF_DECL(vertical_4)
LPF_DECL(vertical_6)
LPF_DECL(vertical_8)
LPF_DECL(vertical_14)

#define FIND_DIR AVX2_OR_C(svt_aom_cdef_find_dir_c, svt_aom_cdef_find_dir_avx2)
#define FIND_DIR_DUAL AVX2_OR_C(svt_aom_cdef_find_dir_dual_c, svt_aom_cdef_find_dir_dual_avx2)
#define FILTER_BLOCK AVX2_OR_C(svt_cdef_filter_block_c, svt_cdef_filter_block_avx2)
#define COPY_RECT8 AVX2_OR_C(svt_aom_copy_rect8_8bit_to_16bit_c, svt_aom_copy_rect8_8bit_to_16bit_avx2)
#define CDEF_DIST16 AVX2_OR_C(svt_aom_compute_cdef_dist_c, svt_aom_compute_cdef_dist_16bit_avx2)
#define CDEF_DIST8 AVX2_OR_C(svt_aom_compute_cdef_dist_8bit_c, svt_aom_compute_cdef_dist_8bit_avx2)
#define SEARCH_ONE_DUAL AVX2_OR_C(svt_search_one_dual_c, svt_search_one_dual_avx2)
#define CCSO_WO_BUF AVX2_OR_C(ccso_filter_block_hbd_wo_buf_c, ccso_filter_block_hbd_wo_buf_avx2)

uint8_t FIND_DIR(const uint16_t *img, int32_t stride, int32_t *var, int32_t coeff_shift);
void    FIND_DIR_DUAL(const uint16_t *img1, const uint16_t *img2, int stride, int32_t *var1, int32_t *var2, int32_t coeff_shift,
                      uint8_t *out1, uint8_t *out2);
void    FILTER_BLOCK(uint8_t *dst8, uint16_t *dst16, int32_t dstride, const uint16_t *in, int32_t pri_strength, int32_t sec_strength,
                     int32_t dir, int32_t pri_damping, int32_t sec_damping, int32_t bsize, int32_t coeff_shift,
                     uint8_t subsampling_factor);
void    COPY_RECT8(uint16_t *dst, int32_t dstride, const uint8_t *src, int32_t sstride, int32_t v, int32_t h);
uint64_t CDEF_DIST16(const uint16_t *dst, int32_t dstride, const uint16_t *src, const CdefList *dlist, int32_t cdef_count,
                     BlockSize bsize, int32_t coeff_shift, int32_t pli, uint8_t subsampling_factor);
uint64_t CDEF_DIST8(const uint8_t *dst8, int32_t dstride, const uint8_t *src8, const CdefList *dlist, int32_t cdef_count,
                    BlockSize bsize, int32_t coeff_shift, int32_t pli, uint8_t subsampling_factor);
uint64_t SEARCH_ONE_DUAL(int *lev0, int *lev1, int nb_strengths, uint64_t **mse[2], int sb_count, int start_gi, int end_gi);

uint8_t (*svt_aom_cdef_find_dir)(const uint16_t *, int32_t, int32_t *, int32_t) = FIND_DIR;
void (*svt_aom_cdef_find_dir_dual)(const uint16_t *, const uint16_t *, int, int32_t *, int32_t *, int32_t, uint8_t *,
                                   uint8_t *)                                   = FIND_DIR_DUAL;
void (*svt_cdef_filter_block)(uint8_t *, uint16_t *, int32_t, const uint16_t *, int32_t, int32_t, int32_t, int32_t, int32_t,
                              int32_t, int32_t, uint8_t)                        = FILTER_BLOCK;
void (*svt_aom_copy_rect8_8bit_to_16bit)(uint16_t *, int32_t, const uint8_t *, int32_t, int32_t,
                                         int32_t)                               = COPY_RECT8;
uint64_t (*svt_compute_cdef_dist_16bit)(const uint16_t *, int32_t, const uint16_t *, const CdefList *, int32_t, BlockSize,
                                        int32_t, int32_t, uint8_t)              = CDEF_DIST16;
uint64_t (*svt_compute_cdef_dist_8bit)(const uint8_t *, int32_t, const uint8_t *, const CdefList *, int32_t, BlockSize,
                                       int32_t, int32_t, uint8_t)               = CDEF_DIST8;
uint64_t (*svt_search_one_dual)(int *, int *, int, uint64_t **[2], int, int, int) = SEARCH_ONE_DUAL;

#if SVTF_REF_SIMD
/* svt_cdef_filter_block_avx2 dispatches its 16-bit 8xN case through this pointer (common_dsp_rtcd.c:629-630) */
void svt_cdef_filter_block_8xn_16_avx2(const uint16_t *const in, const int32_t pri_strength, const int32_t sec_strength, const int32_t dir,
                                       int32_t pri_damping, int32_t sec_damping, const int32_t coeff_shift, uint16_t *const dst,
                                       const int32_t dstride, uint8_t height, uint8_t subsampling_factor);
void (*svt_cdef_filter_block_8xn_16)(const uint16_t *const, const int32_t, const int32_t, const int32_t, int32_t, int32_t, const int32_t,
                                     uint16_t *const, const int32_t, uint8_t, uint8_t) = svt_cdef_filter_block_8xn_16_avx2;
#endif

void CCSO_WO_BUF(const uint16_t *src_y, uint16_t *dst_yuv, const int x, const int y, const int pic_width, const int pic_height,
                 int *src_cls, const int8_t *offset_buf, const int src_y_stride, const int dst_stride, const int y_uv_hscale,
                 const int y_uv_vscale, const int thr, const int neg_thr, const int *src_loc, const int max_val,
                 const int blk_size, const bool isSingleBand, const uint8_t shift_bits, const int edge_clf,
                 const uint8_t ccso_bo_only);
void (*ccso_filter_block_hbd_wo_buf)(const uint16_t *, uint16_t *, const int, const int, const int, const int, int *,
                                     const int8_t *, const int, const int, const int, const int, const int, const int,
                                     const int *, const int, const int, const bool, const uint8_t, const int,
                                     const uint8_t) = CCSO_WO_BUF;
/* the per-plane apply functions are global in Common/Codec/EbCcso.c:297,375,455,542 but not declared in EbCcso.h */
#define CCSO_APPLY_DECL(n)                                                                                            \
    void n(PictureControlSet *pcs, MacroblockdPlane *pd, const int plane, const uint16_t *src_y, uint16_t *dst_yuv,   \
           const int dst_stride, const uint8_t thr, const uint8_t filter_sup, const uint8_t max_band_log2,            \
           const int edge_clf);
CCSO_APPLY_DECL(ccso_apply_luma_mb_filter)
CCSO_APPLY_DECL(ccso_apply_luma_sb_filter)
CCSO_APPLY_DECL(ccso_apply_chroma_mb_filter)
CCSO_APPLY_DECL(ccso_apply_chroma_sb_filter)

void svt_aom_get_recon_pic(PictureControlSet *pcs, EbPictureBufferDesc **recon_ptr, Bool is_highbd) {
    /* Encoder/Codec/EbRestProcess.c:135-153 with is_ref == FALSE */
    *recon_ptr = is_highbd ? pcs->ppcs->enc_dec_ptr->recon_pic_16bit : pcs->ppcs->enc_dec_ptr->recon_pic;
}
static uint32_t g_lambda = 0;
static void     shim_lambda(PictureControlSet *pcs, uint32_t *fast_lambda, uint32_t *full_lambda, uint8_t bit_depth,
                            uint16_t qp_index, Bool multiply_lambda) {
    (void)pcs, (void)bit_depth, (void)qp_index, (void)multiply_lambda;
    *fast_lambda = g_lambda;
    *full_lambda = g_lambda;
}
const EbAv1LambdaAssignFunc svt_aom_av1_lambda_assignment_function_table[4] = {shim_lambda, shim_lambda, shim_lambda,
                                                                               shim_lambda};

#include "ref_harness.h"

/* svt_av1_loop_filter_frame — Encoder/Codec/EbDeblockingFilter.c:624 */
REF_API int ref_deblock_frame(const svtf_geom *g, void *y, int sy, void *cb, int scb, void *cr, int scr,
                              const ref_mi_soa *soa, const svtf_lf_params *lf, int plane_start, int plane_end) {
    harness h;
    if (harness_init(&h, g, soa, y, sy, cb, scb, cr, scr))
        return -1;
    FrameHeader       *fh = &h.ppcs->frm_hdr;
    struct LoopFilter *l  = &fh->loop_filter_params;
    l->filter_level[0]    = lf->filter_level[0];
    l->filter_level[1]    = lf->filter_level[1];
    l->filter_level_u     = lf->filter_level_u;
    l->filter_level_v     = lf->filter_level_v;
    l->sharpness_level    = lf->sharpness_level;
    l->mode_ref_delta_enabled = (uint8_t)lf->mode_ref_delta_enabled;
    memcpy(l->ref_deltas, lf->ref_deltas, sizeof(l->ref_deltas));
    memcpy(l->mode_deltas, lf->mode_deltas, sizeof(l->mode_deltas));
    fh->segmentation_params.segmentation_enabled = (uint8_t)lf->segmentation_enabled;
    for (int s = 0; s < 8; s++)
        for (int f = 0; f < 8; f++) {
            fh->segmentation_params.feature_enabled[s][f] = lf->seg_feature_enabled[s][f];
            fh->segmentation_params.feature_data[s][f]    = lf->seg_feature_data[s][f];
        }
    fh->delta_lf_params.delta_lf_present = (uint8_t)lf->delta_lf_present;
    fh->delta_lf_params.delta_lf_multi   = (uint8_t)lf->delta_lf_multi;
    for (int i = 0; i < 4; i++) h.ppcs->curr_delta_lf[i] = lf->curr_delta_lf[i];
    svt_av1_loop_filter_init(h.pcs);
    svt_av1_loop_filter_frame(h.recon, h.pcs, plane_start, plane_end);
    harness_free(&h);
    return 0;
}

/* svt_av1_cdef_frame — Encoder/Codec/EbEncCdef.c:284.  use_reference_cdef_fs=1 makes the apply
 * recompute directions itself (dirinit = 0), i.e. no dependence on a preceding search. */
REF_API int ref_cdef_apply_frame(const svtf_geom *g, void *y, int sy, void *cb, int scb, void *cr, int scr,
                                 const ref_mi_soa *soa, const svtf_cdef_params *cp, const int8_t *fb_strength_idx) {
    harness h;
    if (harness_init(&h, g, soa, y, sy, cb, scb, cr, scr))
        return -1;
    FrameHeader *fh              = &h.ppcs->frm_hdr;
    fh->cdef_params.cdef_damping = (uint8_t)cp->cdef_damping;
    fh->cdef_params.cdef_bits    = (uint8_t)cp->cdef_bits;
    for (int i = 0; i < CDEF_MAX_STRENGTHS && i < SVTF_CDEF_MAX_STRENGTHS; i++) {
        fh->cdef_params.cdef_y_strength[i]  = cp->cdef_y_strength[i];
        fh->cdef_params.cdef_uv_strength[i] = cp->cdef_uv_strength[i];
    }
    h.ppcs->cdef_ctrls.use_reference_cdef_fs = 1;
    const int nvfb = (g->mi_rows + 15) / 16, nhfb = (g->mi_cols + 15) / 16;
    for (int fbr = 0; fbr < nvfb; fbr++)
        for (int fbc = 0; fbc < nhfb; fbc++) {
            int8_t idx = fb_strength_idx[fbr * nhfb + fbc];
            /* all-skip fbs carry a stale index in the reference; any valid value works */
            h.mi[16 * fbr * soa->mi_stride + 16 * fbc].mbmi.cdef_strength = idx < 0 ? 0 : idx;
        }
    svt_av1_cdef_frame(h.scs, h.pcs);
    harness_free(&h);
    return 0;
}

/* cdef_seg_search (static, Encoder/Codec/EbCdefProcess.c:114) over one whole-frame segment. */
REF_API int ref_cdef_search_frame(const svtf_geom *g, void *ry, int rsy, void *rcb, int rscb, void *rcr, int rscr, void *oy,
                                  int osy, void *ocb, int oscb, void *ocr, int oscr, const ref_mi_soa *soa,
                                  const svtf_cdef_search_cfg *cfg, uint64_t *mse, uint8_t *skip, uint8_t *dir,
                                  int32_t *var) {
    harness h;
    if (harness_init(&h, g, soa, ry, rsy, rcb, rscb, rcr, rscr))
        return -1;
    EbPictureBufferDesc *s = h.source;
    *s                     = *h.recon;
    s->buffer_y = oy, s->buffer_cb = ocb, s->buffer_cr = ocr;
    s->stride_y = (uint16_t)osy, s->stride_cb = (uint16_t)oscb, s->stride_cr = (uint16_t)oscr;
    h.pcs->input_frame16bit     = s;
    h.pcs->cdef_input_recon[0]  = ry;
    h.pcs->cdef_input_recon[1]  = rcb;
    h.pcs->cdef_input_recon[2]  = rcr;
    h.pcs->cdef_input_source[0] = oy;
    h.pcs->cdef_input_source[1] = ocb;
    h.pcs->cdef_input_source[2] = ocr;
    CdefControls *c             = &h.ppcs->cdef_ctrls;
    c->first_pass_fs_num          = (uint8_t)cfg->first_pass_fs_num;
    c->default_second_pass_fs_num = (uint8_t)cfg->default_second_pass_fs_num;
    memcpy(c->default_first_pass_fs, cfg->default_first_pass_fs, 64);
    memcpy(c->default_second_pass_fs, cfg->default_second_pass_fs, 64);
    memcpy(c->default_first_pass_fs_uv, cfg->default_first_pass_fs_uv, 64);
    memcpy(c->default_second_pass_fs_uv, cfg->default_second_pass_fs_uv, 64);
    c->subsampling_factor                               = (uint8_t)cfg->subsampling_factor;
    h.ppcs->frm_hdr.quantization_params.base_q_idx      = (uint8_t)cfg->base_q_idx;
    cdef_seg_search(h.pcs, h.scs, 0);
    const int nfb = ((g->mi_rows + 15) / 16) * ((g->mi_cols + 15) / 16);
    memcpy(mse, h.pcs->mse_seg[0], sizeof(uint64_t) * 64 * nfb);
    memcpy(mse + (size_t)64 * nfb, h.pcs->mse_seg[1], sizeof(uint64_t) * 64 * nfb);
    if (skip)
        memcpy(skip, h.pcs->skip_cdef_seg, nfb);
    if (dir) {
        /* per filter block, its own 8x8 grid.  A 128-wide / 128-high coding block is searched as one filter block whose CdefDirData
         * (16x16 entries) belongs to its first quadrant (EbCdefProcess.c:186-211): read the other quadrants' entries from there. */
        const int nhfb = (g->mi_cols + 15) / 16, nvfb = (g->mi_rows + 15) / 16;
        for (int fbr = 0; fbr < nvfb; fbr++)
            for (int fbc = 0; fbc < nhfb; fbc++) {
                const BlockSize bs  = h.mi[16 * fbr * soa->mi_stride + 16 * fbc].mbmi.block_mi.bsize;
                const int       dc  = (fbc & 1) && (bs == BLOCK_128X128 || bs == BLOCK_128X64);
                const int       dr  = (fbr & 1) && (bs == BLOCK_128X128 || bs == BLOCK_64X128);
                const int       fb  = fbr * nhfb + fbc, ofb = (fbr - dr) * nhfb + fbc - dc;
                for (int b = 0; b < 64; b++) {
                    dir[fb * 64 + b] = h.pcs->cdef_dir_data[ofb].dir[(b >> 3) + 8 * dr][(b & 7) + 8 * dc];
                    var[fb * 64 + b] = h.pcs->cdef_dir_data[ofb].var[(b >> 3) + 8 * dr][(b & 7) + 8 * dc];
                }
            }
    }
    harness_free(&h);
    return 0;
}

/* finish_cdef_search — Encoder/Codec/EbEncCdef.c:728, fed with a given mse table. */
REF_API int ref_cdef_pick(const svtf_geom *g, const ref_mi_soa *soa, const svtf_cdef_search_cfg *cfg, const uint64_t *mse,
                          const uint8_t *skip, svtf_cdef_params *out, int8_t *fb_idx) {
    harness h;
    if (harness_init(&h, g, soa, NULL, 0, NULL, 0, NULL, 0))
        return -1;
    const int nvfb = (g->mi_rows + 15) / 16, nhfb = (g->mi_cols + 15) / 16, nfb = nvfb * nhfb;
    memcpy(h.pcs->mse_seg[0], mse, sizeof(uint64_t) * 64 * nfb);
    memcpy(h.pcs->mse_seg[1], mse + (size_t)64 * nfb, sizeof(uint64_t) * 64 * nfb);
    memcpy(h.pcs->skip_cdef_seg, skip, nfb);
    CdefControls *c               = &h.ppcs->cdef_ctrls;
    c->first_pass_fs_num          = (uint8_t)cfg->first_pass_fs_num;
    c->default_second_pass_fs_num = (uint8_t)cfg->default_second_pass_fs_num;
    memcpy(c->default_first_pass_fs, cfg->default_first_pass_fs, 64);
    memcpy(c->default_second_pass_fs, cfg->default_second_pass_fs, 64);
    c->zero_fs_cost_bias                           = (uint16_t)cfg->zero_fs_cost_bias;
    h.ppcs->frm_hdr.quantization_params.base_q_idx = (uint8_t)cfg->base_q_idx;
    h.ppcs->pred_structure                         = 0;
    h.source->bit_depth                            = (EbBitDepth)g->bit_depth;
    g_lambda                                       = cfg->lambda;
    for (int i = 0; i < g->mi_rows * soa->mi_stride; i++) h.mi[i].mbmi.cdef_strength = -1;
    finish_cdef_search(h.pcs);
    memset(out, 0, sizeof(*out));
    out->cdef_damping = h.ppcs->frm_hdr.cdef_params.cdef_damping;
    out->cdef_bits    = h.ppcs->frm_hdr.cdef_params.cdef_bits;
    for (int i = 0; i < CDEF_MAX_STRENGTHS && i < SVTF_CDEF_MAX_STRENGTHS; i++) {
        out->cdef_y_strength[i]  = h.ppcs->frm_hdr.cdef_params.cdef_y_strength[i];
        out->cdef_uv_strength[i] = h.ppcs->frm_hdr.cdef_params.cdef_uv_strength[i];
    }
    for (int fbr = 0; fbr < nvfb; fbr++)
        for (int fbc = 0; fbc < nhfb; fbc++)
            fb_idx[fbr * nhfb + fbc] = h.mi[16 * fbr * soa->mi_stride + 16 * fbc].mbmi.cdef_strength;
    harness_free(&h);
    return 0;
}

/* CCSO apply.  8-bit: ccso_frame itself (Common/Codec/EbCcso.c:626).  16-bit samples: the
 * frame wrapper is byte-only (SURVEY key fact 2), so the per-plane ccso_apply_* functions
 * (:297-622) are called on the uint16 planes directly — the same functions ccso_frame calls.
 * ext_rec_y is built here exactly as Encoder/Codec/EbCdefProcess.c:480-502 + extend_ccso_border. */
REF_API int ref_ccso_apply_frame(const svtf_geom *g, void *y, int sy, void *cb, int scb, void *cr, int scr, const void *pre_y,
                                 int pre_sy, const svtf_ccso_params *cc, const uint8_t *const blk_flags[3]) {
    harness    h;
    const int  mi_stride = g->mi_cols;
    ref_mi_soa soa       = {0};
    soa.mi_stride        = mi_stride;
    if (harness_init(&h, g, &soa, y, sy, cb, scb, cr, scr))
        return -1;
    CcsoInfo *ci = &h.ppcs->frm_hdr.ccso_info;
    for (int p = 0; p < 3; p++) {
        ci->ccso_enable[p]        = cc->ccso_enable[p];
        ci->ccso_bo_only[p]       = cc->ccso_bo_only[p];
        ci->quant_idx[p]          = cc->quant_idx[p];
        ci->ext_filter_support[p] = cc->ext_filter_support[p];
        ci->edge_clf[p]           = cc->edge_clf[p];
        ci->max_band_log2[p]      = cc->max_band_log2[p];
        memcpy(ci->filter_offset[p], cc->filter_offset[p], SVTF_CCSO_LUT_SIZE);
        const int pw = g->width >> (p ? 1 : 0), ph = g->height >> (p ? 1 : 0), bl = p ? 7 : 8;
        const int nbx = (pw + (1 << bl) - 1) >> bl, nby = (ph + (1 << bl) - 1) >> bl;
        for (int by = 0; by < nby; by++)
            for (int bx = 0; bx < nbx; bx++) {
                const int f  = (blk_flags && blk_flags[p]) ? blk_flags[p][by * nbx + bx] : 1;
                MbModeInfo *m = &h.mi[64 * by * mi_stride + 64 * bx].mbmi;
                if (p == 0) m->ccso_blk_y = f;
                if (p == 1) m->ccso_blk_u = f;
                if (p == 2) m->ccso_blk_v = f;
            }
    }
    struct MacroblockdPlane pd[3];
    memset(pd, 0, sizeof(pd));
    const int is16 = g->sample_bytes == 2;
    for (int p = 0; p < 3; p++) {
        pd[p].subsampling_x = pd[p].subsampling_y = p ? 1 : 0;
        pd[p].is_16bit                            = is16;
        pd[p].dst.width                           = g->width >> (p ? 1 : 0);
        pd[p].dst.height                          = g->height >> (p ? 1 : 0);
    }
    const int W = g->width, H = g->height, es = W + 10;
    uint16_t *ext = calloc((size_t)es * (H + 10), sizeof(uint16_t));
    for (int r = 0; r < H; r++)
        for (int c = 0; c < W; c++)
            ext[(r + 5) * es + c + 5] = is16 ? ((const uint16_t *)pre_y)[r * pre_sy + c] : ((const uint8_t *)pre_y)[r * pre_sy + c];
    extend_ccso_border(ext, CCSO_PADDING_SIZE, pd);
    if (!is16) {
        ccso_frame(h.recon, h.pcs, pd, ext);
    } else {
        static const uint8_t quant_sz[4] = {16, 8, 32, 64};
        void                *planes[3]   = {y, cb, cr};
        const int            strides[3]  = {sy, scb, scr};
        for (int p = 0; p < 3; p++) {
            if (!ci->ccso_enable[p])
                continue;
            CCSO_FILTER_FUNC fn = ci->max_band_log2[p] ? (p > 0 ? ccso_apply_chroma_mb_filter : ccso_apply_luma_mb_filter)
                                                       : (p > 0 ? ccso_apply_chroma_sb_filter : ccso_apply_luma_sb_filter);
            fn(h.pcs, pd, p, ext, (uint16_t *)planes[p], strides[p], quant_sz[ci->quant_idx[p]], ci->ext_filter_support[p],
               (uint8_t)ci->max_band_log2[p], ci->edge_clf[p]);
        }
    }
    free(ext);
    harness_free(&h);
    return 0;
}

/* CCSO derivation for one plane: derive_ccso_filter (Encoder/Codec/EbPickccso.c:464).  luma/rec/org are uint16 samples
 * at stride g->width with g->height rows allocated (the reference keeps every plane at luma stride and copies
 * pd[0].dst.height rows).  ccso_search (:785) itself only scales rdmult and sets up pd from the recon descriptor. */
extern int ccso_stride, ccso_stride_ext;
void       derive_ccso_filter(PictureControlSet *pcs, const int plane, MacroblockdPlane *pd, const uint16_t *org_uv, const uint16_t *ext_rec_y,
                              const uint16_t *rec_uv, int rdmult);
REF_API int ref_ccso_derive_plane(const svtf_geom *g, int plane, const uint16_t *luma, const uint16_t *rec, const uint16_t *org, int rdmult,
                                  svtf_ccso_params *out, uint8_t *flags) {
    harness    h;
    const int  mi_stride = g->mi_cols;
    ref_mi_soa soa       = {0};
    soa.mi_stride        = mi_stride;
    if (harness_init(&h, g, &soa, (void *)luma, g->width, (void *)luma, g->width, (void *)luma, g->width))
        return -1;
    h.scs->static_config.encoder_bit_depth = g->bit_depth;
    struct MacroblockdPlane pd[3];
    memset(pd, 0, sizeof(pd));
    for (int p = 0; p < 3; p++) {
        pd[p].subsampling_x = pd[p].subsampling_y = p ? 1 : 0;
        pd[p].is_16bit                            = 1;
        pd[p].dst.width                           = g->width >> (p ? 1 : 0);
        pd[p].dst.height                          = g->height >> (p ? 1 : 0);
    }
    const int W = g->width, H = g->height, es = W + 10;
    uint16_t *ext = calloc((size_t)es * (H + 10), sizeof(uint16_t));
    for (int r = 0; r < H; r++) memcpy(ext + (size_t)(r + 5) * es + 5, luma + (size_t)r * W, sizeof(uint16_t) * W);
    extend_ccso_border(ext, CCSO_PADDING_SIZE, pd);
    ccso_stride     = W;
    ccso_stride_ext = es;
    derive_ccso_filter(h.pcs, plane, pd, org, ext, rec, rdmult);
    const CcsoInfo *ci      = &h.ppcs->frm_hdr.ccso_info;
    out->ccso_enable[plane] = ci->ccso_enable[plane];
    const int pw = g->width >> (plane ? 1 : 0), ph = g->height >> (plane ? 1 : 0), bl = plane ? 7 : 8;
    const int nbx = (pw + (1 << bl) - 1) >> bl, nby = (ph + (1 << bl) - 1) >> bl;
    memset(flags, 0, (size_t)nbx * nby);
    if (ci->ccso_enable[plane]) {
        out->ccso_bo_only[plane]       = ci->ccso_bo_only[plane];
        out->quant_idx[plane]          = ci->quant_idx[plane];
        out->ext_filter_support[plane] = ci->ext_filter_support[plane];
        out->edge_clf[plane]           = ci->edge_clf[plane];
        out->max_band_log2[plane]      = ci->max_band_log2[plane];
        memcpy(out->filter_offset[plane], ci->filter_offset[plane], SVTF_CCSO_LUT_SIZE);
        for (int by = 0; by < nby; by++)
            for (int bx = 0; bx < nbx; bx++) {
                const MbModeInfo *m  = &h.mi[64 * by * mi_stride + 64 * bx].mbmi;
                flags[by * nbx + bx] = plane == 0 ? m->ccso_blk_y : (plane == 1 ? m->ccso_blk_u : m->ccso_blk_v);
            }
    }
    free(ext);
    harness_free(&h);
    return 0;
}

/* =========================================================================================
 * leaf-level wrappers (the functions the reference's own unit tests exercise)
 * ======================================================================================= */
/* svt_aom_[highbd_]lpf_{horizontal,vertical}_{4,6,8,14}_c — test/DeblockTest.cc:95-309 */
/* (all leaf wrappers go through the RTCD pointers above: the `_c` kernels in libsvtref.so, the SSE2/AVX2 ones in libsvtref_simd.so) */
REF_API void ref_lpf(void *s, int sample_bytes, int pitch, int vertical, int len, int bd, uint8_t blimit, uint8_t limit,
                     uint8_t thresh) {
    uint8_t bl[16], li[16], th[16];
    memset(bl, blimit, 16), memset(li, limit, 16), memset(th, thresh, 16);
    if (sample_bytes == 1) {
        void (*f)(uint8_t *, int32_t, const uint8_t *, const uint8_t *, const uint8_t *) =
            vertical ? (len == 4 ? svt_aom_lpf_vertical_4 : len == 6 ? svt_aom_lpf_vertical_6 : len == 8 ? svt_aom_lpf_vertical_8 : svt_aom_lpf_vertical_14)
                     : (len == 4 ? svt_aom_lpf_horizontal_4 : len == 6 ? svt_aom_lpf_horizontal_6 : len == 8 ? svt_aom_lpf_horizontal_8 : svt_aom_lpf_horizontal_14);
        f(s, pitch, bl, li, th);
    } else {
        void (*f)(uint16_t *, int32_t, const uint8_t *, const uint8_t *, const uint8_t *, int32_t) =
            vertical ? (len == 4 ? svt_aom_highbd_lpf_vertical_4 : len == 6 ? svt_aom_highbd_lpf_vertical_6 : len == 8 ? svt_aom_highbd_lpf_vertical_8 : svt_aom_highbd_lpf_vertical_14)
                     : (len == 4 ? svt_aom_highbd_lpf_horizontal_4 : len == 6 ? svt_aom_highbd_lpf_horizontal_6 : len == 8 ? svt_aom_highbd_lpf_horizontal_8 : svt_aom_highbd_lpf_horizontal_14);
        f(s, pitch, bl, li, th, bd);
    }
}
/* svt_cdef_filter_block_c — test/CdefTest.cc:94-411.  `in` is inside a CDEF_BSTRIDE tile. */
REF_API void ref_cdef_filter_block(uint16_t *dst16, int dstride, const uint16_t *in, int pri, int sec, int dir, int pri_damping,
                                   int sec_damping, int bsize, int coeff_shift, int subsampling) {
    svt_cdef_filter_block(NULL, dst16, dstride, in, pri, sec, dir, pri_damping, sec_damping, bsize, coeff_shift,
                            (uint8_t)subsampling);
}
REF_API int ref_cdef_bstride(void) { return CDEF_BSTRIDE; }
/* svt_aom_cdef_find_dir_c — test/CdefTest.cc:435-665 */
REF_API int ref_cdef_find_dir(const uint16_t *img, int stride, int32_t *var, int coeff_shift) {
    return svt_aom_cdef_find_dir(img, stride, var, coeff_shift);
}
/* svt_aom_compute_cdef_dist_c / _8bit_c — test/CdefTest.cc:760-980 */
REF_API uint64_t ref_cdef_dist(const void *dst, int dstride, const void *src, const uint8_t *by, const uint8_t *bx, int count,
                               int bsize, int coeff_shift, int pli, int subsampling, int is16) {
    CdefList dl[256];
    for (int i = 0; i < count; i++) dl[i].by = by[i], dl[i].bx = bx[i];
    return is16 ? svt_compute_cdef_dist_16bit(dst, dstride, src, dl, count, (BlockSize)bsize, coeff_shift, pli, (uint8_t)subsampling)
                : svt_compute_cdef_dist_8bit(dst, dstride, src, dl, count, (BlockSize)bsize, coeff_shift, pli,
                                                   (uint8_t)subsampling);
}
/* table probes used to pin svtf_pack_mi_enc: get_transform_size (Encoder/Codec/EbDeblockingFilter.c:143-160)
 * inputs, block_size_wide/high, mode_lf_lut */
REF_API void ref_mi_tables(int bsize, int tx_depth, int *txw_y, int *txh_y, int *txw_uv, int *txh_uv, int *bw, int *bh) {
    const TxSize ty  = tx_depth_to_tx_size[tx_depth][bsize];
    const TxSize tuv = av1_get_max_uv_txsize((BlockSize)bsize, 1, 1);
    *txw_y = tx_size_wide[ty], *txh_y = tx_size_high[ty];
    *txw_uv = tx_size_wide[tuv], *txh_uv = tx_size_high[tuv];
    *bw = block_size_wide[bsize], *bh = block_size_high[bsize];
}
REF_API int ref_mode_lf_lut(int mode) { return mode_lf_lut[mode]; }
REF_API int ref_tx_size_dims(int tx_size, int *w, int *h) {
    *w = tx_size_wide[tx_size], *h = tx_size_high[tx_size];
    return 0;
}
REF_API size_t ref_sizeof_pcs(void) { return sizeof(PictureControlSet); }
