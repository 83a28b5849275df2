/*
 * ref_shim_lrs.c — drives the reference's loop-restoration SEARCH on synthetic frames.  TEST INFRASTRUCTURE ONLY.
 *
 * Encoder/Codec/EbRestorationPick.c is compiled as part of this translation unit (it lies under $(REF); nothing is copied)
 * so that its static functions are reachable:
 *   restoration_seg_search :1471, rest_finish_search :1555   — the whole search on one picture, one segment
 *   svt_av1_compute_stats_c / _highbd_c :671 / :708           — the statistics of one unit
 *   wiener_decompose_sep_sym :906, finalize_sym_filter :977, compute_score :937 — the host half of search_wiener_seg
 * Bit counting comes from the real Encoder/Codec/EbEntropyCoding.c, SSE from EbPsnr.c + C_DEFAULT/variance.c.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "EbPictureControlSet.h"
#include "EbSequenceControlSet.h"
#include "EbRestoration.h"
#include "EbRestorationPick.c"

#include "../include/svtfilt_b200.h"

#define REF_API __attribute__((visibility("default")))
#define LRS_BORDER 64

void RunEmms(void) {}

typedef struct {
    Yv12BufferConfig f;
    void            *alloc[3];
} lrs_frame;

static void lrs_frame_init(lrs_frame *fr, const svtf_geom *g, const void *const pl[3], const int st[3]) {
    memset(fr, 0, sizeof(*fr));
    const int sb = g->sample_bytes;
    for (int p = 0; p < 3; p++) {
        const int ss = p ? 1 : 0, w = (g->width + ss) >> ss, h = (g->height + ss) >> ss;
        const int stride = (w + 2 * LRS_BORDER + 31) & ~31;
        uint8_t  *buf    = calloc((size_t)stride * (h + 2 * LRS_BORDER), sb);
        uint8_t  *org    = buf + ((size_t)LRS_BORDER * stride + LRS_BORDER) * sb;
        for (int r = 0; r < h; r++) memcpy(org + (size_t)r * stride * sb, (const uint8_t *)pl[p] + (size_t)r * st[p] * sb, (size_t)w * sb);
        fr->alloc[p]          = buf;
        fr->f.buffers[p]      = sb == 2 ? CONVERT_TO_BYTEPTR(org) : org;
        fr->f.strides[ss]     = stride;
        fr->f.crop_widths[ss] = w, fr->f.crop_heights[ss] = h;
        fr->f.widths[ss] = w, fr->f.heights[ss] = h;
    }
    fr->f.border = LRS_BORDER;
    fr->f.flags  = sb == 2 ? YV12_FLAG_HIGHBITDEPTH : 0;
}
static void lrs_frame_free(lrs_frame *fr) {
    for (int p = 0; p < 3; p++) free(fr->alloc[p]);
}

REF_API int ref_lr_search(const svtf_geom *g, const void *y, int sy, const void *cb, int scb, const void *cr, int scr, const void *oy, int osy,
                          const void *ocb, int oscb, const void *ocr, int oscr, const svtf_lr_search_cfg *cfg, int32_t out_type[3],
                          svtf_lr_unit *const out_units[3], svtf_lr_unit_search *const out_search[3], int32_t sg_frame_ep_cnt[16]) {
    PictureControlSet       *pcs  = calloc(1, sizeof(*pcs));
    PictureParentControlSet *ppcs = calloc(1, sizeof(*ppcs));
    Av1Common               *cm   = calloc(1, sizeof(*cm));
    Macroblock              *x    = calloc(1, sizeof(*x));
    int32_t                 *tmp  = malloc(RESTORATION_TMPBUF_SIZE);
    if (!pcs || !ppcs || !cm || !x || !tmp)
        return -1;
    pcs->ppcs = ppcs, ppcs->av1_cm = cm, ppcs->av1x = x, cm->child_pcs = pcs;
    cm->mi_rows = g->mi_rows, cm->mi_cols = g->mi_cols;
    cm->frm_size.frame_width = (uint16_t)g->width, cm->frm_size.frame_height = (uint16_t)g->height;
    cm->frm_size.superres_upscaled_width = (uint16_t)g->width, cm->frm_size.superres_upscaled_height = (uint16_t)g->height;
    cm->frm_size.superres_denominator    = SCALE_NUMERATOR;
    cm->subsampling_x = cm->subsampling_y = 1;
    cm->bit_depth                         = g->bit_depth;
    cm->use_highbitdepth                  = g->sample_bytes == 2;
    cm->use_boundaries_in_rest_search     = 0; /* EbEncHandle.c:4162 */
    cm->wn_filter_ctrls.enabled                 = (Bool)cfg->wn_enabled;
    cm->wn_filter_ctrls.filter_tap_lvl          = (uint8_t)cfg->wn_filter_tap_lvl;
    cm->wn_filter_ctrls.use_refinement          = (Bool)cfg->wn_use_refinement;
    cm->wn_filter_ctrls.max_one_refinement_step = (Bool)cfg->wn_max_one_refinement_step;
    cm->wn_filter_ctrls.use_prev_frame_coeffs   = cfg->prev_units[0] || cfg->prev_units[1] || cfg->prev_units[2];
    cm->wn_filter_ctrls.use_chroma              = (Bool)cfg->wn_use_chroma;
    cm->sg_filter_ctrls.enabled                 = (Bool)cfg->sg_enabled;
    cm->sg_filter_ctrls.use_chroma              = (Bool)cfg->sg_use_chroma;
    cm->sg_filter_ctrls.step_range              = 16; /* the explicit ep range */
    for (int t = 0; t < 2; t++) {
        cm->sg_filter_ctrls.start_ep[t] = (int8_t)cfg->sg_start_ep[t], cm->sg_filter_ctrls.end_ep[t] = (int8_t)cfg->sg_end_ep[t];
        cm->sg_filter_ctrls.ep_inc[t] = (int8_t)cfg->sg_ep_inc[t], cm->sg_filter_ctrls.refine[t] = (int8_t)cfg->sg_refine[t];
    }
    cm->sg_ref_frame_ep[0] = cm->sg_ref_frame_ep[1] = -1;
    x->sgrproj_restore_cost[0] = cfg->sgrproj_restore_cost[0], x->sgrproj_restore_cost[1] = cfg->sgrproj_restore_cost[1];
    for (int t = 0; t < 3; t++) x->switchable_restore_cost[t] = cfg->switchable_restore_cost[t];
    ppcs->frm_hdr.frame_type                    = cm->wn_filter_ctrls.use_prev_frame_coeffs ? INTER_FRAME : KEY_FRAME;
    x->rdmult                                   = cfg->rdmult;
    x->wiener_restore_cost[0] = cfg->wiener_restore_cost[0], x->wiener_restore_cost[1] = cfg->wiener_restore_cost[1];
    for (int p = 0; p < 3; p++) {
        pcs->rst_info[p].frame_restoration_type = RESTORE_NONE;
        pcs->rst_info[p].restoration_unit_size  = cfg->restoration_unit_size[p];
    }
    if (svt_av1_alloc_restoration_buffers(pcs, cm) != EB_ErrorNone)
        return -1;
    for (int p = 0; p < 3; p++) pcs->rusi_picture[p] = calloc(pcs->rst_info[p].units_per_tile, sizeof(RestUnitSearchInfo));
    for (int p = 0; p < 3; p++) /* what EbRestProcess.c:586-597 copies from the reference picture */
        for (int u = 0; u < pcs->rst_info[p].units_per_tile && cfg->prev_units[p]; u++) {
            pcs->rst_info[p].unit_info[u].restoration_type = (RestorationType)cfg->prev_units[p][u].restoration_type;
            memcpy(pcs->rst_info[p].unit_info[u].wiener_info.vfilter, cfg->prev_units[p][u].wiener_vfilter, 16);
            memcpy(pcs->rst_info[p].unit_info[u].wiener_info.hfilter, cfg->prev_units[p][u].wiener_hfilter, 16);
        }
    pcs->rest_segments_column_count = pcs->rest_segments_row_count = 1;
    const void *const cur[3] = {y, cb, cr}, *const src[3] = {oy, ocb, ocr};
    const int         cst[3] = {sy, scb, scr}, sst[3] = {osy, oscb, oscr};
    lrs_frame         fc, fs, ft;
    lrs_frame_init(&fc, g, cur, cst);
    lrs_frame_init(&fs, g, src, sst);
    lrs_frame_init(&ft, g, cur, cst);
    cm->frame_to_show = &fc.f;
    restoration_seg_search(tmp, &fc.f, &fs.f, &ft.f, pcs, 0);
    rest_finish_search(pcs);
    for (int p = 0; p < 3; p++) {
        const RestorationInfo *rsi = &pcs->rst_info[p];
        out_type[p]                = (int32_t)rsi->frame_restoration_type;
        for (int u = 0; u < rsi->units_per_tile; u++) {
            svtf_lr_unit *o = &out_units[p][u];
            memset(o, 0, sizeof(*o));
            if (rsi->frame_restoration_type != RESTORE_NONE) {
                o->restoration_type = (int32_t)rsi->unit_info[u].restoration_type;
                if (o->restoration_type == SVTF_RESTORE_WIENER) {
                    memcpy(o->wiener_vfilter, rsi->unit_info[u].wiener_info.vfilter, 16);
                    memcpy(o->wiener_hfilter, rsi->unit_info[u].wiener_info.hfilter, 16);
                } else if (o->restoration_type == SVTF_RESTORE_SGRPROJ) {
                    o->sgr_ep = rsi->unit_info[u].sgrproj_info.ep;
                    o->sgr_xqd[0] = rsi->unit_info[u].sgrproj_info.xqd[0], o->sgr_xqd[1] = rsi->unit_info[u].sgrproj_info.xqd[1];
                }
            }
            if (out_search && out_search[p]) {
                const RestUnitSearchInfo *ri = &pcs->rusi_picture[p][u];
                svtf_lr_unit_search      *os = &out_search[p][u];
                memset(os, 0, sizeof(*os));
                const int searched = !p || (cfg->wn_enabled && cfg->wn_use_chroma) || (cfg->sg_enabled && cfg->sg_use_chroma);
                if (!searched)
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
        }
    }
    if (sg_frame_ep_cnt)
        for (int i = 0; i < 16; i++) sg_frame_ep_cnt[i] = cm->sg_frame_ep_cnt[i];
    for (int p = 0; p < 3; p++) {
        free(pcs->rusi_picture[p]);
        free(pcs->rst_info[p].unit_info), free(pcs->rst_info[p].boundaries.stripe_boundary_above), free(pcs->rst_info[p].boundaries.stripe_boundary_below);
    }
    lrs_frame_free(&fc), lrs_frame_free(&fs), lrs_frame_free(&ft);
    free(tmp), free(x), free(cm), free(ppcs), free(pcs);
    return 0;
}

/* statistics of the unit [h0, h1) x [v0, v1) of one plane (pw x ph samples, frame edges replicated as svt_extend_frame does) */
REF_API int ref_lr_wiener_stats(const void *dgd, int dst, const void *src, int sst, int sb, int bd, int pw, int ph, int h0, int h1, int v0, int v1,
                                int win, int64_t *M, int64_t *H) {
    const int stride = (pw + 2 * LRS_BORDER + 31) & ~31;
    uint8_t  *bd_ = calloc((size_t)stride * (ph + 2 * LRS_BORDER), sb), *bs = calloc((size_t)stride * (ph + 2 * LRS_BORDER), sb);
    if (!bd_ || !bs)
        return -1;
    uint8_t *od = bd_ + ((size_t)LRS_BORDER * stride + LRS_BORDER) * sb, *os = bs + ((size_t)LRS_BORDER * stride + LRS_BORDER) * sb;
    for (int r = 0; r < ph; r++) {
        memcpy(od + (size_t)r * stride * sb, (const uint8_t *)dgd + (size_t)r * dst * sb, (size_t)pw * sb);
        memcpy(os + (size_t)r * stride * sb, (const uint8_t *)src + (size_t)r * sst * sb, (size_t)pw * sb);
    }
    svt_extend_frame(sb == 2 ? CONVERT_TO_BYTEPTR(od) : od, pw, ph, stride, RESTORATION_BORDER + 1, RESTORATION_BORDER, sb == 2);
    if (sb == 2)
        svt_av1_compute_stats_highbd_c(win, CONVERT_TO_BYTEPTR(od), CONVERT_TO_BYTEPTR(os), h0, h1, v0, v1, stride, stride, M, H, (EbBitDepth)bd);
    else
        svt_av1_compute_stats_c(win, od, os, h0, h1, v0, v1, stride, stride, M, H);
    free(bd_), free(bs);
    return 0;
}

REF_API int ref_lr_wiener_from_stats(int win, const int64_t *M, const int64_t *H, int16_t vf[8], int16_t hf[8]) {
    int64_t Mc[WIENER_WIN2], *Hc = malloc(sizeof(int64_t) * WIENER_WIN2 * WIENER_WIN2);
    int32_t vd[WIENER_WIN], hd[WIENER_WIN];
    memcpy(Mc, M, sizeof(int64_t) * win * win), memcpy(Hc, H, sizeof(int64_t) * win * win * win * win);
    memset(vf, 0, 16), memset(hf, 0, 16);
    int ok = wiener_decompose_sep_sym(win, Mc, Hc, vd, hd);
    if (ok) {
        finalize_sym_filter(win, vd, vf), finalize_sym_filter(win, hd, hf);
        if (compute_score(win, Mc, Hc, vf, hf) > 0)
            ok = 0;
    }
    free(Hc);
    return ok;
}
