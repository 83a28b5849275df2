/*
 * adapter_harness.c — TEST INFRASTRUCTURE: builds the same PictureControlSet graph oracle/ref_shim.c builds
 * (oracle/ref_harness.h) and calls the Tier-2 adapters (adapters/svtf_enc_adapters.c) through the reference's own
 * signatures, so tests can compare `svt_av1_loop_filter_frame(frame, pcs, 0, 3)` & co. of this repo with the
 * real reference functions on identical structures.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "EbPictureControlSet.h"
#include "EbSequenceControlSet.h"
#include "EbEncCdef.h"
#include "EbCdef.h"
#include "EbDeblockingFilter.h"
#include "EbDeblockingCommon.h"
#include "EbCcso.h"
#include "EbPickccso.h"
#include "EbModeDecisionProcess.h"

#include "../include/svtfilt_b200.h"
#include "../include/svtfilt_b200_adapters.h"
#include "../oracle/ref_harness.h"

#define API __attribute__((visibility("default")))

void svt_aom_get_recon_pic(PictureControlSet *pcs, EbPictureBufferDesc **recon_ptr, Bool is_highbd) {
    *recon_ptr = is_highbd ? pcs->ppcs->enc_dec_ptr->recon_pic_16bit : pcs->ppcs->enc_dec_ptr->recon_pic;
}
static uint32_t g_lambda = 0;
static void     h_lambda(PictureControlSet *pcs, uint32_t *fast_lambda, uint32_t *full_lambda, uint8_t bit_depth,
                         uint16_t qp_index, Bool multiply_lambda) {
    (void)pcs, (void)bit_depth, (void)qp_index, (void)multiply_lambda;
    *fast_lambda = *full_lambda = g_lambda;
}
const EbAv1LambdaAssignFunc svt_aom_av1_lambda_assignment_function_table[4] = {h_lambda, h_lambda, h_lambda, h_lambda};

static void fill_lf(harness *h, const svtf_lf_params *lf) {
    FrameHeader       *fh = &h->ppcs->frm_hdr;
    struct LoopFilter *l  = &fh->loop_filter_params;
    l->filter_level[0] = lf->filter_level[0], l->filter_level[1] = lf->filter_level[1];
    l->filter_level_u = lf->filter_level_u, l->filter_level_v = lf->filter_level_v;
    l->sharpness_level        = lf->sharpness_level;
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
    for (int i = 0; i < 4; i++) h->ppcs->curr_delta_lf[i] = lf->curr_delta_lf[i];
}

API int adapt_deblock_frame(const svtf_geom *g, void *y, int sy, void *cb, int scb, void *cr, int scr, const ref_mi_soa *soa,
                            const svtf_lf_params *lf, int plane_start, int plane_end) {
    harness h;
    if (harness_init(&h, g, soa, y, sy, cb, scb, cr, scr))
        return -1;
    fill_lf(&h, lf);
    svt_av1_loop_filter_frame(h.recon, h.pcs, plane_start, plane_end);
    svtf_adapter_release(h.pcs);
    harness_free(&h);
    return 0;
}

API int adapt_cdef_apply_frame(const svtf_geom *g, void *y, int sy, void *cb, int scb, void *cr, int scr, const ref_mi_soa *soa,
                               const svtf_cdef_params *cp, const int8_t *fb_strength_idx) {
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
    const int nvfb = (g->mi_rows + 15) / 16, nhfb = (g->mi_cols + 15) / 16;
    for (int fbr = 0; fbr < nvfb; fbr++)
        for (int fbc = 0; fbc < nhfb; fbc++) {
            int8_t idx = fb_strength_idx[fbr * nhfb + fbc];
            h.mi[16 * fbr * soa->mi_stride + 16 * fbc].mbmi.cdef_strength = idx < 0 ? 0 : idx;
        }
    svt_av1_cdef_frame(h.scs, h.pcs);
    svtf_adapter_release(h.pcs);
    harness_free(&h);
    return 0;
}

/* search + pick through svtf_cdef_search_picture() and finish_cdef_search() */
API int adapt_cdef_search_pick(const svtf_geom *g, void *ry, int rsy, void *rcb, int rscb, void *rcr, int rscr, void *oy, int osy,
                               void *ocb, int oscb, void *ocr, int oscr, const ref_mi_soa *soa, const svtf_cdef_search_cfg *cfg,
                               uint64_t *mse, uint8_t *skip, uint8_t *dir, int32_t *var, svtf_cdef_params *out, int8_t *fb_idx) {
    harness h;
    if (harness_init(&h, g, soa, ry, rsy, rcb, rscb, rcr, rscr))
        return -1;
    EbPictureBufferDesc *s = h.source;
    *s                     = *h.recon;
    s->buffer_y = oy, s->buffer_cb = ocb, s->buffer_cr = ocr;
    s->stride_y = (uint16_t)osy, s->stride_cb = (uint16_t)oscb, s->stride_cr = (uint16_t)oscr;
    h.pcs->input_frame16bit = s;
    CdefControls *c           = &h.ppcs->cdef_ctrls;
    c->first_pass_fs_num          = (uint8_t)cfg->first_pass_fs_num;
    c->default_second_pass_fs_num = (uint8_t)cfg->default_second_pass_fs_num;
    memcpy(c->default_first_pass_fs, cfg->default_first_pass_fs, 64);
    memcpy(c->default_second_pass_fs, cfg->default_second_pass_fs, 64);
    memcpy(c->default_first_pass_fs_uv, cfg->default_first_pass_fs_uv, 64);
    memcpy(c->default_second_pass_fs_uv, cfg->default_second_pass_fs_uv, 64);
    c->subsampling_factor                          = (uint8_t)cfg->subsampling_factor;
    c->zero_fs_cost_bias                           = (uint16_t)cfg->zero_fs_cost_bias;
    h.ppcs->frm_hdr.quantization_params.base_q_idx = (uint8_t)cfg->base_q_idx;
    h.source->bit_depth                            = (EbBitDepth)g->bit_depth;
    g_lambda                                       = cfg->lambda;
    for (int i = 0; i < g->mi_rows * soa->mi_stride; i++) h.mi[i].mbmi.cdef_strength = -1;
    svtf_cdef_search_picture(h.pcs, h.scs);
    const int nvfb = (g->mi_rows + 15) / 16, nhfb = (g->mi_cols + 15) / 16, nfb = nvfb * nhfb;
    memcpy(mse, h.pcs->mse_seg[0], sizeof(uint64_t) * 64 * nfb);
    memcpy(mse + (size_t)64 * nfb, h.pcs->mse_seg[1], sizeof(uint64_t) * 64 * nfb);
    memcpy(skip, h.pcs->skip_cdef_seg, nfb);
    for (int fb = 0; fb < nfb; fb++)
        for (int b = 0; b < 64; b++) {
            dir[fb * 64 + b] = h.pcs->cdef_dir_data[fb].dir[b >> 3][b & 7];
            var[fb * 64 + b] = h.pcs->cdef_dir_data[fb].var[b >> 3][b & 7];
        }
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
    svtf_adapter_release(h.pcs);
    harness_free(&h);
    return 0;
}

/* The DLF kernel's sequence (Encoder/Codec/EbDlfProcess.c:101-106): pick the levels against the source, then filter. */
void svt_av1_pick_filter_level_by_q(PictureControlSet *pcs, uint8_t qindex, int32_t *filter_level) {
    (void)pcs, (void)qindex, (void)filter_level;
    fprintf(stderr, "adapter_harness: svt_av1_pick_filter_level_by_q is the reference's own scalar formula; not part of this harness\n");
    abort();
}
API int adapt_pick_and_deblock(const svtf_geom *g, void *y, int sy, void *cb, int scb, void *cr, int scr, void *oy, int osy, void *ocb,
                               int oscb, void *ocr, int oscr, const ref_mi_soa *soa, const svtf_lf_params *lf,
                               const svtf_lf_search_cfg *cfg, int resident, int32_t out_levels[4]) {
    harness h;
    if (harness_init(&h, g, soa, y, sy, cb, scb, cr, scr))
        return -1;
    fill_lf(&h, lf);
    EbPictureBufferDesc *s = h.source;
    *s                     = *h.recon;
    s->buffer_y = oy, s->buffer_cb = ocb, s->buffer_cr = ocr;
    s->stride_y = (uint16_t)osy, s->stride_cb = (uint16_t)oscb, s->stride_cr = (uint16_t)oscr;
    h.pcs->input_frame16bit                  = s;
    h.ppcs->frm_hdr.tx_mode                  = cfg->only_4x4 ? ONLY_4X4 : TX_MODE_SELECT;
    h.ppcs->dlf_ctrls.dlf_avg                = (uint8_t)cfg->dlf_avg;
    h.ppcs->dlf_ctrls.early_exit_convergence = (uint8_t)cfg->early_exit_convergence;
    h.ppcs->dlf_ctrls.dlf_avg_uv             = (uint8_t)!cfg->search_chroma;
    h.pcs->temporal_layer_index              = 1;
    h.ppcs->tot_ref_frame_types              = 0;
    svtf_adapter_set_resident(resident);
    svt_av1_pick_filter_level((EbPictureBufferDesc *)h.ppcs->enhanced_pic, h.pcs, LPF_PICK_FROM_FULL_IMAGE);
    const struct LoopFilter *l = &h.ppcs->frm_hdr.loop_filter_params;
    out_levels[0] = l->filter_level[0], out_levels[1] = l->filter_level[1], out_levels[2] = l->filter_level_u, out_levels[3] = l->filter_level_v;
    svt_av1_loop_filter_frame(h.recon, h.pcs, 0, 3);
    if (resident) { /* bring the picture back as the last stage of a resident chain would */
        svtf_adapter_set_resident(0);
        svtf_adapter_fetch(h.pcs, h.recon);
    }
    svtf_adapter_release(h.pcs);
    harness_free(&h);
    return 0;
}

/* ccso_frame(frame, pcs, pd, ext_rec_y) with ext_rec_y built like Encoder/Codec/EbCdefProcess.c:480-502 */
API int adapt_ccso_frame(const svtf_geom *g, void *y, int sy, void *cb, int scb, void *cr, int scr, const void *pre_y, int pre_sy,
                         const svtf_ccso_params *cc, const uint8_t *const blk_flags[3]) {
    harness    h;
    ref_mi_soa soa = {0};
    soa.mi_stride  = g->mi_cols;
    if (harness_init(&h, g, &soa, y, sy, cb, scb, cr, scr))
        return -1;
    CcsoInfo *ci = &h.ppcs->frm_hdr.ccso_info;
    for (int p = 0; p < 3; p++) {
        ci->ccso_enable[p] = cc->ccso_enable[p], ci->ccso_bo_only[p] = cc->ccso_bo_only[p], ci->quant_idx[p] = cc->quant_idx[p];
        ci->ext_filter_support[p] = cc->ext_filter_support[p], ci->edge_clf[p] = cc->edge_clf[p];
        ci->max_band_log2[p] = cc->max_band_log2[p];
        memcpy(ci->filter_offset[p], cc->filter_offset[p], SVTF_CCSO_LUT_SIZE);
        const int pw = g->width >> (p ? 1 : 0), ph = g->height >> (p ? 1 : 0), bl = p ? 7 : 8;
        const int nbx = (pw + (1 << bl) - 1) >> bl, nby = (ph + (1 << bl) - 1) >> bl;
        for (int by = 0; by < nby; by++)
            for (int bx = 0; bx < nbx; bx++) {
                const int   f = (blk_flags && blk_flags[p]) ? blk_flags[p][by * nbx + bx] : 1;
                MbModeInfo *m = &h.mi[64 * by * soa.mi_stride + 64 * bx].mbmi;
                if (p == 0) m->ccso_blk_y = f;
                if (p == 1) m->ccso_blk_u = f;
                if (p == 2) m->ccso_blk_v = f;
            }
    }
    const int is16 = g->sample_bytes == 2;
    const int W = g->width, H = g->height, es = W + 10;
    uint16_t *ext = calloc((size_t)es * (H + 10), sizeof(uint16_t));
    for (int r = 0; r < H + 10; r++)
        for (int c = 0; c < W + 10; c++) {
            const int rr = r - 5 < 0 ? 0 : (r - 5 >= H ? H - 1 : r - 5), ccn = c - 5 < 0 ? 0 : (c - 5 >= W ? W - 1 : c - 5);
            ext[r * es + c] = is16 ? ((const uint16_t *)pre_y)[rr * pre_sy + ccn] : ((const uint8_t *)pre_y)[rr * pre_sy + ccn];
        }
    struct MacroblockdPlane pd[3];
    memset(pd, 0, sizeof(pd));
    ccso_frame(h.recon, h.pcs, pd, ext);
    free(ext);
    svtf_adapter_release(h.pcs);
    harness_free(&h);
    return 0;
}

/* ccso_search(pcs, pd, rdmult, ext_rec_y, rec_uv, org_uv) on uint16 planes at luma stride; returns the derived
 * parameters and per-block flags as the reference leaves them in frm_hdr.ccso_info / mbmi.ccso_blk_*. */
API int adapt_ccso_search(const svtf_geom *g, int base_q_idx, int rdmult, const uint16_t *luma, uint16_t *rec[3], uint16_t *org[3],
                          svtf_ccso_params *out, uint8_t *const flags[3]) {
    harness    h;
    ref_mi_soa soa = {0};
    soa.mi_stride  = g->mi_cols;
    if (harness_init(&h, g, &soa, (void *)luma, g->width, (void *)luma, g->width, (void *)luma, g->width))
        return -1;
    h.ppcs->frm_hdr.quantization_params.base_q_idx = (uint8_t)base_q_idx;
    const int W = g->width, H = g->height, es = W + 10;
    uint16_t *ext = calloc((size_t)es * (H + 10), sizeof(uint16_t));
    for (int r = 0; r < H + 10; r++)
        for (int c = 0; c < W + 10; c++) {
            const int rr = r - 5 < 0 ? 0 : (r - 5 >= H ? H - 1 : r - 5), cc = c - 5 < 0 ? 0 : (c - 5 >= W ? W - 1 : c - 5);
            ext[r * es + c] = luma[rr * W + cc];
        }
    struct MacroblockdPlane pd[3];
    memset(pd, 0, sizeof(pd));
    ccso_search(h.pcs, pd, rdmult, ext, rec, org);
    const CcsoInfo *ci = &h.ppcs->frm_hdr.ccso_info;
    memset(out, 0, sizeof(*out));
    for (int p = 0; p < 3; p++) {
        out->ccso_enable[p] = ci->ccso_enable[p];
        const int pw = W >> (p ? 1 : 0), ph = H >> (p ? 1 : 0), bl = p ? 7 : 8;
        const int nbx = (pw + (1 << bl) - 1) >> bl, nby = (ph + (1 << bl) - 1) >> bl;
        memset(flags[p], 0, (size_t)nbx * nby);
        if (!ci->ccso_enable[p])
            continue;
        out->ccso_bo_only[p] = ci->ccso_bo_only[p], out->quant_idx[p] = ci->quant_idx[p], out->ext_filter_support[p] = ci->ext_filter_support[p];
        out->edge_clf[p] = ci->edge_clf[p], out->max_band_log2[p] = ci->max_band_log2[p];
        memcpy(out->filter_offset[p], ci->filter_offset[p], SVTF_CCSO_LUT_SIZE);
        for (int by = 0; by < nby; by++)
            for (int bx = 0; bx < nbx; bx++) {
                const MbModeInfo *m = &h.mi[64 * by * soa.mi_stride + 64 * bx].mbmi;
                flags[p][by * nbx + bx] = p == 0 ? m->ccso_blk_y : (p == 1 ? m->ccso_blk_u : m->ccso_blk_v);
            }
    }
    free(ext);
    svtf_adapter_release(h.pcs);
    harness_free(&h);
    return 0;
}

/* ---- loop restoration -------------------------------------------------------------------------------------
 * RestorationInfo of the three planes from plain data; the stripe boundaries (what
 * svt_av1_loop_restoration_save_boundary_lines keeps of the deblocked picture, Common/Codec/EbRestoration.c:1522-1574:
 * rows y0-2, y0-1 above and y1, y1+1 below every stripe, the last one repeated at the crop border) are written here from
 * the deblocked picture the test passes. */
#include "EbRestoration.h"
static void lr_fill(harness *h, Av1Common *cm, const svtf_geom *g, const svtf_lr_params *lr, const svtf_lr_unit *const units[3],
                    void *const dbk[3], const int dst[3]) {
    memset(cm, 0, sizeof(*cm));
    cm->child_pcs = h->pcs;
    cm->mi_rows = g->mi_rows, cm->mi_cols = g->mi_cols;
    cm->frm_size.frame_width = (uint16_t)g->width, cm->frm_size.frame_height = (uint16_t)g->height;
    cm->frm_size.superres_upscaled_width = (uint16_t)g->width, cm->frm_size.superres_upscaled_height = (uint16_t)g->height;
    cm->frm_size.superres_denominator    = SCALE_NUMERATOR;
    cm->subsampling_x = cm->subsampling_y = 1;
    cm->bit_depth = g->bit_depth, cm->use_highbitdepth = g->sample_bytes == 2;
    const int sb = g->sample_bytes;
    for (int p = 0; p < 3; p++) {
        RestorationInfo *rsi = &h->pcs->rst_info[p];
        const int        ss = p ? 1 : 0, pw = (g->width + ss) >> ss, ph = (g->height + ss) >> ss, sh = 64 >> ss, off = 8 >> ss;
        int32_t          hu = 1, vu = 1;
        rsi->frame_restoration_type = (RestorationType)lr->frame_restoration_type[p];
        rsi->restoration_unit_size  = lr->restoration_unit_size[p];
        svtf_lr_units(g, p, lr->restoration_unit_size[p], &hu, &vu);
        rsi->horz_units_per_tile = hu, rsi->vert_units_per_tile = vu, rsi->units_per_tile = hu * vu;
        rsi->unit_info = calloc(hu * vu, sizeof(RestorationUnitInfo));
        for (int u = 0; u < hu * vu && units[p]; u++) {
            RestorationUnitInfo *ri = &rsi->unit_info[u];
            ri->restoration_type    = (RestorationType)units[p][u].restoration_type;
            memcpy(ri->wiener_info.vfilter, units[p][u].wiener_vfilter, sizeof(ri->wiener_info.vfilter));
            memcpy(ri->wiener_info.hfilter, units[p][u].wiener_hfilter, sizeof(ri->wiener_info.hfilter));
            ri->sgrproj_info.ep = units[p][u].sgr_ep, ri->sgrproj_info.xqd[0] = units[p][u].sgr_xqd[0], ri->sgrproj_info.xqd[1] = units[p][u].sgr_xqd[1];
        }
        RestorationStripeBoundaries *b = &rsi->boundaries;
        const int nstripes = (ph + off + sh - 1) / sh, stride = (pw + 2 * RESTORATION_EXTRA_HORZ + 31) & ~31;
        b->stripe_boundary_stride = stride;
        b->stripe_boundary_above  = calloc((size_t)nstripes * RESTORATION_CTX_VERT * stride, sb);
        b->stripe_boundary_below  = calloc((size_t)nstripes * RESTORATION_CTX_VERT * stride, sb);
        if (!dbk)
            continue;
        for (int k = 0; k < nstripes; k++) {
            const int y0 = k * sh - off > 0 ? k * sh - off : 0, y1 = (k + 1) * sh - off < ph ? (k + 1) * sh - off : ph;
            for (int i = 0; i < RESTORATION_CTX_VERT; i++) {
                const size_t row = ((size_t)(RESTORATION_CTX_VERT * k + i) * stride + RESTORATION_EXTRA_HORZ) * sb;
                if (k > 0)
                    memcpy(b->stripe_boundary_above + row, (uint8_t *)dbk[p] + (size_t)(y0 - 2 + i) * dst[p] * sb, (size_t)pw * sb);
                if (y1 < ph)
                    memcpy(b->stripe_boundary_below + row, (uint8_t *)dbk[p] + (size_t)(y1 + i < ph ? y1 + i : ph - 1) * dst[p] * sb, (size_t)pw * sb);
            }
        }
    }
}
static void lr_free(harness *h) {
    for (int p = 0; p < 3; p++) {
        free(h->pcs->rst_info[p].unit_info);
        free(h->pcs->rst_info[p].boundaries.stripe_boundary_above), free(h->pcs->rst_info[p].boundaries.stripe_boundary_below);
    }
}
static void lr_yv12(Yv12BufferConfig *f, const svtf_geom *g, void *const pl[3], const int st[3]) {
    memset(f, 0, sizeof(*f));
    for (int p = 0; p < 3; p++) {
        const int ss = p ? 1 : 0;
        f->buffers[p]      = g->sample_bytes == 2 ? CONVERT_TO_BYTEPTR(pl[p]) : (uint8_t *)pl[p];
        f->strides[ss]     = st[p];
        f->crop_widths[ss] = (g->width + ss) >> ss, f->crop_heights[ss] = (g->height + ss) >> ss;
    }
}

/* stand-alone: `cur` (after CDEF) is restored in place, the stripe borders come from `dbk` through rst_info[].boundaries */
API int adapt_lr_frame(const svtf_geom *g, void *y, int sy, void *cb, int scb, void *cr, int scr, void *dy, int dsy, void *dcb, int dscb,
                       void *dcr, int dscr, const svtf_lr_params *lr, const svtf_lr_unit *const units[3]) {
    static const uint8_t zero = 0;
    harness              h;
    ref_mi_soa           soa = {NULL, NULL, NULL, NULL, NULL, NULL, g->mi_cols};
    (void)zero;
    if (harness_init(&h, g, &soa, y, sy, cb, scb, cr, scr))
        return -1;
    Av1Common   cm;
    void *const dbk[3] = {dy, dcb, dcr}, *const cur[3] = {y, cb, cr};
    const int   dst[3] = {dsy, dscb, dscr}, cst[3] = {sy, scb, scr};
    lr_fill(&h, &cm, g, lr, units, dbk, dst);
    Yv12BufferConfig f;
    lr_yv12(&f, g, cur, cst);
    svtf_adapter_set_resident(0);
    svt_av1_loop_restoration_filter_frame(NULL, &f, &cm, lr->optimized_lr);
    svtf_adapter_release(h.pcs);
    lr_free(&h);
    harness_free(&h);
    return 0;
}

/* svt_av1_superres_upscale_frame on the reference's own structures: the planes (strides wide enough for the upscaled picture)
 * are upscaled in place */
void svt_av1_superres_upscale_frame(struct Av1Common *cm, PictureControlSet *pcs, SequenceControlSet *scs);
API int adapt_superres(const svtf_geom *g, void *y, int sy, void *cb, int scb, void *cr, int scr, int up_w, int denom, int tile_cols,
                       const int32_t *tile_col_start_mi) {
    harness    h;
    ref_mi_soa soa = {NULL, NULL, NULL, NULL, NULL, NULL, g->mi_cols};
    if (harness_init(&h, g, &soa, y, sy, cb, scb, cr, scr))
        return -1;
    h.cm->frm_size.frame_width = (uint16_t)g->width, h.cm->frm_size.frame_height = (uint16_t)g->height;
    h.cm->frm_size.superres_upscaled_width = (uint16_t)up_w, h.cm->frm_size.superres_upscaled_height = (uint16_t)g->height;
    h.cm->frm_size.superres_denominator    = (uint8_t)denom;
    h.cm->tiles_info.tile_cols             = (uint8_t)tile_cols;
    for (int j = 0; j <= tile_cols; j++) h.cm->tiles_info.tile_col_start_mi[j] = (uint16_t)tile_col_start_mi[j];
    svtf_adapter_set_resident(0);
    svt_av1_superres_upscale_frame(h.cm, h.pcs, h.scs);
    svtf_adapter_release(h.pcs);
    harness_free(&h);
    return 0;
}

/* restoration_seg_search (segment 0) + rest_finish_search on the reference's own structures */
void restoration_seg_search(int32_t *rst_tmpbuf, Yv12BufferConfig *org_fts, const Yv12BufferConfig *src, Yv12BufferConfig *trial_frame_rst,
                            PictureControlSet *pcs, uint32_t segment_index);
void rest_finish_search(PictureControlSet *pcs);
API int adapt_lr_search(const svtf_geom *g, void *y, int sy, void *cb, int scb, void *cr, int scr, void *oy, int osy, void *ocb, int oscb, void *ocr,
                        int oscr, const svtf_lr_search_cfg *cfg, int32_t out_type[3], svtf_lr_unit *const out_units[3],
                        svtf_lr_unit_search *const out_search[3], int32_t sg_frame_ep_cnt[16]) {
    harness    h;
    ref_mi_soa soa = {NULL, NULL, NULL, NULL, NULL, NULL, g->mi_cols};
    if (harness_init(&h, g, &soa, y, sy, cb, scb, cr, scr))
        return -1;
    Av1Common      cm;
    Macroblock    *x = calloc(1, sizeof(*x));
    svtf_lr_params lr;
    memset(&lr, 0, sizeof(lr));
    for (int p = 0; p < 3; p++) lr.restoration_unit_size[p] = cfg->restoration_unit_size[p];
    const svtf_lr_unit *none[3] = {NULL, NULL, NULL};
    lr_fill(&h, &cm, g, &lr, none, NULL, NULL);
    h.ppcs->av1_cm = &cm, h.ppcs->av1x = x;
    cm.wn_filter_ctrls.enabled                 = (Bool)cfg->wn_enabled;
    cm.wn_filter_ctrls.filter_tap_lvl          = (uint8_t)cfg->wn_filter_tap_lvl;
    cm.wn_filter_ctrls.use_refinement          = (Bool)cfg->wn_use_refinement;
    cm.wn_filter_ctrls.max_one_refinement_step = (Bool)cfg->wn_max_one_refinement_step;
    cm.wn_filter_ctrls.use_chroma              = (Bool)cfg->wn_use_chroma;
    x->rdmult                                  = cfg->rdmult;
    x->wiener_restore_cost[0] = cfg->wiener_restore_cost[0], x->wiener_restore_cost[1] = cfg->wiener_restore_cost[1];
    x->sgrproj_restore_cost[0] = cfg->sgrproj_restore_cost[0], x->sgrproj_restore_cost[1] = cfg->sgrproj_restore_cost[1];
    for (int t = 0; t < 3; t++) x->switchable_restore_cost[t] = cfg->switchable_restore_cost[t];
    cm.sg_filter_ctrls.enabled = (Bool)cfg->sg_enabled, cm.sg_filter_ctrls.use_chroma = (Bool)cfg->sg_use_chroma;
    cm.sg_filter_ctrls.step_range = 16;
    for (int t = 0; t < 2; t++) {
        cm.sg_filter_ctrls.start_ep[t] = (int8_t)cfg->sg_start_ep[t], cm.sg_filter_ctrls.end_ep[t] = (int8_t)cfg->sg_end_ep[t];
        cm.sg_filter_ctrls.ep_inc[t] = (int8_t)cfg->sg_ep_inc[t], cm.sg_filter_ctrls.refine[t] = (int8_t)cfg->sg_refine[t];
    }
    cm.sg_ref_frame_ep[0] = cm.sg_ref_frame_ep[1] = -1;
    cm.wn_filter_ctrls.use_prev_frame_coeffs = cfg->prev_units[0] || cfg->prev_units[1] || cfg->prev_units[2];
    h.ppcs->frm_hdr.frame_type               = cm.wn_filter_ctrls.use_prev_frame_coeffs ? INTER_FRAME : KEY_FRAME;
    for (int p = 0; p < 3; p++)
        for (int u = 0; u < h.pcs->rst_info[p].units_per_tile && cfg->prev_units[p]; u++) {
            h.pcs->rst_info[p].unit_info[u].restoration_type = (RestorationType)cfg->prev_units[p][u].restoration_type;
            memcpy(h.pcs->rst_info[p].unit_info[u].wiener_info.vfilter, cfg->prev_units[p][u].wiener_vfilter, 16);
            memcpy(h.pcs->rst_info[p].unit_info[u].wiener_info.hfilter, cfg->prev_units[p][u].wiener_hfilter, 16);
        }
    for (int p = 0; p < 3; p++) h.pcs->rusi_picture[p] = calloc(h.pcs->rst_info[p].units_per_tile, sizeof(RestUnitSearchInfo));
    void *const      cur[3] = {y, cb, cr}, *const org[3] = {oy, ocb, ocr};
    const int        cst[3] = {sy, scb, scr}, ost[3] = {osy, oscb, oscr};
    Yv12BufferConfig f, s;
    lr_yv12(&f, g, cur, cst), lr_yv12(&s, g, org, ost);
    svtf_adapter_set_resident(0);
    restoration_seg_search(NULL, &f, &s, NULL, h.pcs, 1); /* a later segment: nothing to do */
    restoration_seg_search(NULL, &f, &s, NULL, h.pcs, 0);
    for (int p = 0; p < 3; p++)
        for (int u = 0; u < h.pcs->rst_info[p].units_per_tile && out_search && out_search[p]; u++) {
            const RestUnitSearchInfo *ri = &h.pcs->rusi_picture[p][u];
            svtf_lr_unit_search      *os = &out_search[p][u];
            memset(os, 0, sizeof(*os));
            if (p && !(cfg->wn_enabled && cfg->wn_use_chroma) && !(cfg->sg_enabled && cfg->sg_use_chroma))
                continue;
            os->sse[0] = ri->sse[RESTORE_NONE];
            os->sse[1] = cfg->wn_enabled && (!p || cfg->wn_use_chroma) ? ri->sse[RESTORE_WIENER] : INT64_MAX;
            if (os->sse[1] != INT64_MAX) {
                memcpy(os->wiener_vfilter, ri->wiener.vfilter, 16);
                memcpy(os->wiener_hfilter, ri->wiener.hfilter, 16);
            }
            if (cfg->sg_enabled && (!p || cfg->sg_use_chroma)) {
                os->sse[2] = ri->sse[RESTORE_SGRPROJ];
                os->sgr_ep = ri->sgrproj.ep, os->sgr_xqd[0] = ri->sgrproj.xqd[0], os->sgr_xqd[1] = ri->sgrproj.xqd[1];
            }
        }
    rest_finish_search(h.pcs);
    if (sg_frame_ep_cnt)
        for (int i = 0; i < 16; i++) sg_frame_ep_cnt[i] = cm.sg_frame_ep_cnt[i];
    for (int p = 0; p < 3; p++) {
        const RestorationInfo *rsi = &h.pcs->rst_info[p];
        out_type[p]                = (int32_t)rsi->frame_restoration_type;
        for (int u = 0; u < rsi->units_per_tile; u++) {
            svtf_lr_unit *o = &out_units[p][u];
            memset(o, 0, sizeof(*o));
            if (rsi->frame_restoration_type == RESTORE_NONE)
                continue;
            o->restoration_type = (int32_t)rsi->unit_info[u].restoration_type;
            if (o->restoration_type == SVTF_RESTORE_WIENER) {
                memcpy(o->wiener_vfilter, rsi->unit_info[u].wiener_info.vfilter, 16);
                memcpy(o->wiener_hfilter, rsi->unit_info[u].wiener_info.hfilter, 16);
            } else if (o->restoration_type == SVTF_RESTORE_SGRPROJ) {
                o->sgr_ep = rsi->unit_info[u].sgrproj_info.ep;
                o->sgr_xqd[0] = rsi->unit_info[u].sgrproj_info.xqd[0], o->sgr_xqd[1] = rsi->unit_info[u].sgrproj_info.xqd[1];
            }
        }
        free(h.pcs->rusi_picture[p]);
    }
    svtf_adapter_release(h.pcs);
    lr_free(&h);
    free(x);
    harness_free(&h);
    return 0;
}

/* resident chain through the adapters: svt_av1_loop_filter_frame -> svt_av1_cdef_frame -> svt_av1_loop_restoration_filter_frame */
API int adapt_chain_lr(const svtf_geom *g, void *y, int sy, void *cb, int scb, void *cr, int scr, const ref_mi_soa *soa,
                       const svtf_lf_params *lf, const svtf_cdef_params *cp, const int8_t *fb_strength_idx, const svtf_lr_params *lr,
                       const svtf_lr_unit *const units[3]) {
    harness h;
    if (harness_init(&h, g, soa, y, sy, cb, scb, cr, scr))
        return -1;
    fill_lf(&h, lf);
    FrameHeader *fh              = &h.ppcs->frm_hdr;
    fh->cdef_params.cdef_damping = (uint8_t)cp->cdef_damping;
    fh->cdef_params.cdef_bits    = (uint8_t)cp->cdef_bits;
    for (int i = 0; i < CDEF_MAX_STRENGTHS && i < SVTF_CDEF_MAX_STRENGTHS; i++) {
        fh->cdef_params.cdef_y_strength[i]  = cp->cdef_y_strength[i];
        fh->cdef_params.cdef_uv_strength[i] = cp->cdef_uv_strength[i];
    }
    const int nvfb = (g->mi_rows + 15) / 16, nhfb = (g->mi_cols + 15) / 16;
    for (int fbr = 0; fbr < nvfb; fbr++)
        for (int fbc = 0; fbc < nhfb; fbc++) {
            int8_t idx = fb_strength_idx[fbr * nhfb + fbc];
            h.mi[16 * fbr * soa->mi_stride + 16 * fbc].mbmi.cdef_strength = idx < 0 ? 0 : idx;
        }
    Av1Common cm;
    lr_fill(&h, &cm, g, lr, units, NULL, NULL);
    void *const      cur[3] = {y, cb, cr};
    const int        cst[3] = {sy, scb, scr};
    Yv12BufferConfig f;
    lr_yv12(&f, g, cur, cst);
    svtf_adapter_set_resident(1);
    svt_av1_loop_filter_frame(h.recon, h.pcs, 0, 3);
    svt_av1_cdef_frame(h.scs, h.pcs);
    svt_av1_loop_restoration_filter_frame(NULL, &f, &cm, lr->optimized_lr);
    svtf_adapter_set_resident(0);
    svtf_adapter_release(h.pcs);
    lr_free(&h);
    harness_free(&h);
    return 0;
}

/* svtf_pad_ref_pictures on padded reference buffers: `pad16` / `pad8` / `padn` point at the first sample of the padded
 * allocations (strides in samples), whose picture part at (org_x, org_y) already holds the filtered picture.  resident:
 * deblock the picture on the GPU first (svt_av1_loop_filter_frame) and pad from the device copy. */
API int adapt_pad_ref(const svtf_geom *g, void *pad[3], const int pst[3], uint8_t *pad8[3], const int st8[3], uint8_t *padn[3], const int stn[3],
                      int org_x, int org_y, int pad_right, int pad_bottom, const ref_mi_soa *soa, const svtf_lf_params *lf, int resident) {
    harness h;
    const int sb = g->sample_bytes;
    /* the recon descriptor points into the padded allocation */
    if (harness_init(&h, g, soa, pad[0], pst[0], pad[1], pst[1], pad[2], pst[2]))
        return -1;
    EbPictureBufferDesc *r = h.recon;
    r->org_x = (uint16_t)org_x, r->org_y = (uint16_t)org_y;
    r->width = (uint16_t)(g->width + pad_right), r->height = (uint16_t)(g->height + pad_bottom);
    h.scs->pad_right = pad_right, h.scs->pad_bottom = pad_bottom;
    EbPictureBufferDesc r8 = *r; /* the 8-bit reference picture of a 16-bit pipeline */
    if (sb == 2) {
        r8.buffer_y = pad8[0], r8.buffer_cb = pad8[1], r8.buffer_cr = pad8[2];
        r8.stride_y = (uint16_t)st8[0], r8.stride_cb = (uint16_t)st8[1], r8.stride_cr = (uint16_t)st8[2];
        r8.buffer_bit_inc_y = padn ? padn[0] : NULL, r8.buffer_bit_inc_cb = padn ? padn[1] : NULL, r8.buffer_bit_inc_cr = padn ? padn[2] : NULL;
        if (padn)
            r8.stride_bit_inc_y = (uint16_t)stn[0], r8.stride_bit_inc_cb = (uint16_t)stn[1], r8.stride_bit_inc_cr = (uint16_t)stn[2];
        h.eds->recon_pic = &r8;
    }
    if (resident) {
        fill_lf(&h, lf);
        svtf_adapter_set_resident(1);
        svt_av1_loop_filter_frame(h.recon, h.pcs, 0, 3);
    }
    svtf_pad_ref_pictures(h.pcs, h.scs);
    svtf_adapter_set_resident(0);
    svtf_adapter_release(h.pcs);
    harness_free(&h);
    return 0;
}

/* ---- the whole chain through the reference's own entry points on ONE long-lived PictureControlSet (bench.py's end-to-end leg) ----
 * adapt_chain_open builds the structures once: the recon picture is an ordinary malloc'ed (pageable) EbPictureBufferDesc allocation of
 * the harness, the mode info a real pointer grid (ModeInfo behind mi_grid_base), the frame header carries the loop-filter levels, the
 * CDEF strengths (+ per-block indices in mbmi.cdef_strength) and the CCSO parameters (+ block flags in mbmi.ccso_blk_*).
 * adapt_chain_run = what the pipeline's DLF and CDEF stages call for one picture:
 *     svt_av1_loop_filter_frame(recon, pcs, 0, 3); svt_av1_cdef_frame(scs, pcs); ccso_frame(recon, pcs, pd, ext_rec_y);
 * with the adapters in resident mode (one upload, one download).  adapt_chain_load copies a fresh reconstruction into the picture
 * (what EncDec would have left there); adapt_chain_planes exposes the picture for checking. */
typedef struct {
    harness   h;
    svtf_geom g;
    uint8_t  *plane[3];
    int       stride[3], rows[3];
} chain_handle;

API void *adapt_chain_open(const svtf_geom *g, const ref_mi_soa *soa, const svtf_lf_params *lf, const svtf_cdef_params *cp, const int8_t *fb_strength_idx,
                           const svtf_ccso_params *cc, const uint8_t *const blk_flags[3]) {
    chain_handle *c = calloc(1, sizeof(*c));
    if (!c)
        return NULL;
    c->g = *g;
    for (int p = 0; p < 3; p++) {
        const int ss = p ? 1 : 0;
        c->stride[p] = ((g->mi_cols * 4) >> ss) + 64; /* a stride that is NOT the device pitch: a 2-D copy, like a padded recon picture */
        c->rows[p]   = (g->mi_rows * 4) >> ss;
        c->plane[p]  = malloc((size_t)c->stride[p] * c->rows[p] * g->sample_bytes);
        if (!c->plane[p])
            return NULL;
    }
    if (harness_init(&c->h, g, soa, c->plane[0], c->stride[0], c->plane[1], c->stride[1], c->plane[2], c->stride[2]))
        return NULL;
    c->h.recon->height = (uint16_t)g->height; /* allocation rows = coded rows (org_y = 0): what pin_picture sizes the lock by */
    fill_lf(&c->h, lf);
    FrameHeader *fh              = &c->h.ppcs->frm_hdr;
    fh->cdef_params.cdef_damping = (uint8_t)cp->cdef_damping;
    fh->cdef_params.cdef_bits    = (uint8_t)cp->cdef_bits;
    for (int i = 0; i < CDEF_MAX_STRENGTHS && i < SVTF_CDEF_MAX_STRENGTHS; i++) {
        fh->cdef_params.cdef_y_strength[i]  = cp->cdef_y_strength[i];
        fh->cdef_params.cdef_uv_strength[i] = cp->cdef_uv_strength[i];
    }
    const int nvfb = (g->mi_rows + 15) / 16, nhfb = (g->mi_cols + 15) / 16;
    for (int fbr = 0; fbr < nvfb; fbr++)
        for (int fbc = 0; fbc < nhfb; fbc++) {
            const int8_t idx = fb_strength_idx[fbr * nhfb + fbc];
            c->h.mi[16 * fbr * soa->mi_stride + 16 * fbc].mbmi.cdef_strength = idx < 0 ? 0 : idx;
        }
    CcsoInfo *ci = &fh->ccso_info;
    for (int p = 0; p < 3; p++) {
        ci->ccso_enable[p] = cc->ccso_enable[p], ci->ccso_bo_only[p] = cc->ccso_bo_only[p], ci->quant_idx[p] = cc->quant_idx[p];
        ci->ext_filter_support[p] = cc->ext_filter_support[p], ci->edge_clf[p] = cc->edge_clf[p];
        ci->max_band_log2[p] = cc->max_band_log2[p];
        memcpy(ci->filter_offset[p], cc->filter_offset[p], SVTF_CCSO_LUT_SIZE);
        const int pw = g->width >> (p ? 1 : 0), ph = g->height >> (p ? 1 : 0), bl = p ? 7 : 8;
        const int nbx = (pw + (1 << bl) - 1) >> bl, nby = (ph + (1 << bl) - 1) >> bl;
        for (int by = 0; by < nby; by++)
            for (int bx = 0; bx < nbx; bx++) {
                const int   f = (blk_flags && blk_flags[p]) ? blk_flags[p][by * nbx + bx] : 1;
                MbModeInfo *m = &c->h.mi[64 * by * soa->mi_stride + 64 * bx].mbmi;
                if (p == 0) m->ccso_blk_y = f;
                if (p == 1) m->ccso_blk_u = f;
                if (p == 2) m->ccso_blk_v = f;
            }
    }
    return c;
}

API void adapt_chain_load(void *handle, const void *y, int sy, const void *cb, int scb, const void *cr, int scr) {
    chain_handle *c      = handle;
    const void   *src[3] = {y, cb, cr};
    const int     st[3]  = {sy, scb, scr};
    const int     sb     = c->g.sample_bytes;
    for (int p = 0; p < 3; p++) {
        const int w = (c->g.mi_cols * 4) >> (p ? 1 : 0);
        for (int r = 0; r < c->rows[p]; r++)
            memcpy(c->plane[p] + (size_t)r * c->stride[p] * sb, (const uint8_t *)src[p] + (size_t)r * st[p] * sb, (size_t)w * sb);
    }
}

API void adapt_chain_run(void *handle) {
    chain_handle           *c = handle;
    struct MacroblockdPlane pd[3];
    memset(pd, 0, sizeof(pd));
    svt_av1_loop_filter_frame(c->h.recon, c->h.pcs, 0, 3);
    svt_av1_cdef_frame(c->h.scs, c->h.pcs);
    ccso_frame(c->h.recon, c->h.pcs, pd, NULL); /* resident chain: the classifier luma is the pre-CDEF surface on the device */
}

API void adapt_chain_planes(void *handle, void *out[3], int stride[3]) {
    chain_handle *c = handle;
    for (int p = 0; p < 3; p++) out[p] = c->plane[p], stride[p] = c->stride[p];
}

API void adapt_chain_close(void *handle) {
    chain_handle *c = handle;
    svtf_adapter_release(c->h.pcs);
    harness_free(&c->h);
    for (int p = 0; p < 3; p++) free(c->plane[p]);
    free(c);
}

API void adapt_set_modes(int resident, int pin, int mi_threads) {
    svtf_adapter_set_resident(resident);
    svtf_adapter_set_pin_pictures(pin);
    svtf_adapter_set_mi_threads(mi_threads);
}
