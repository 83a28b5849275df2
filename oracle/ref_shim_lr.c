/*
 * ref_shim_lr.c — drives the reference's loop restoration on synthetic frames.  TEST INFRASTRUCTURE ONLY.
 *
 *   svt_av1_alloc_restoration_buffers              Common/Codec/EbRestoration.c:1700
 *   svt_av1_loop_restoration_save_boundary_lines   :1682  (after deblock: after_cdef = 0; after CDEF: after_cdef = 1)
 *   svt_av1_loop_restoration_filter_frame          :1179
 * are the reference's own code (EbRestoration.c, convolve.c, EbPictureBufferDesc.c compiled where they lie,
 * oracle/Makefile).  The shim copies the caller's planes into bordered Yv12BufferConfig frames (the filters read 3 samples
 * and, for Wiener, up to 15 + 3 samples beyond the picture), fills RestorationInfo / RestorationUnitInfo from plain data
 * and copies the result back.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "EbPictureControlSet.h"
#include "EbSequenceControlSet.h"
#include "EbRestoration.h"

#include "../include/svtfilt_b200.h"

#define REF_API __attribute__((visibility("default")))
#define LR_BORDER 64

typedef struct {
    Yv12BufferConfig f;
    void            *alloc[3];
} lr_frame;

static void lr_frame_init(lr_frame *fr, const svtf_geom *g, void *const pl[3], const int st[3]) {
    memset(fr, 0, sizeof(*fr));
    const int sb = g->sample_bytes;
    for (int p = 0; p < 3; p++) {
        const int ss = p ? 1 : 0, w = (g->width + ss) >> ss, h = (g->height + ss) >> ss;
        const int cw = (g->mi_cols * 4) >> ss, chh = (g->mi_rows * 4) >> ss; /* the caller's planes hold the coded size */
        const int stride = (cw + 2 * LR_BORDER + 31) & ~31;
        uint8_t  *buf    = calloc((size_t)stride * (chh + 2 * LR_BORDER), sb);
        uint8_t  *org    = buf + ((size_t)LR_BORDER * stride + LR_BORDER) * sb;
        for (int r = 0; r < chh; r++) memcpy(org + (size_t)r * stride * sb, (uint8_t *)pl[p] + (size_t)r * st[p] * sb, (size_t)cw * sb);
        fr->alloc[p]        = buf;
        fr->f.buffers[p]    = sb == 2 ? CONVERT_TO_BYTEPTR(org) : org;
        fr->f.strides[ss]   = stride;
        fr->f.crop_widths[ss] = w, fr->f.crop_heights[ss] = h;
        fr->f.widths[ss] = cw, fr->f.heights[ss] = chh;
    }
    fr->f.border = LR_BORDER;
    fr->f.flags  = sb == 2 ? YV12_FLAG_HIGHBITDEPTH : 0;
}
static void lr_frame_store(const lr_frame *fr, const svtf_geom *g, void *const pl[3], const int st[3]) {
    const int sb = g->sample_bytes;
    for (int p = 0; p < 3; p++) {
        const int ss = p ? 1 : 0, w = (g->width + ss) >> ss, h = (g->height + ss) >> ss, stride = fr->f.strides[ss];
        const uint8_t *org = (const uint8_t *)fr->alloc[p] + ((size_t)LR_BORDER * stride + LR_BORDER) * sb;
        for (int r = 0; r < h; r++) memcpy((uint8_t *)pl[p] + (size_t)r * st[p] * sb, org + (size_t)r * stride * sb, (size_t)w * sb);
    }
}

/* cur: the picture to restore (post-CDEF), in/out; dbk: the deblocked picture the stripe boundaries are taken from */
REF_API int ref_lr_frame(const svtf_geom *g, void *y, int sy, void *cb, int scb, void *cr, int scr, void *dy, int dsy, void *dcb, int dscb,
                         void *dcr, int dscr, const svtf_lr_params *lr, const svtf_lr_unit *const units[3]) {
    PictureControlSet *pcs = calloc(1, sizeof(*pcs));
    Av1Common         *cm  = calloc(1, sizeof(*cm));
    int32_t           *tmp = malloc(RESTORATION_TMPBUF_SIZE);
    if (!pcs || !cm || !tmp)
        return -1;
    cm->child_pcs                        = pcs;
    cm->mi_rows = g->mi_rows, cm->mi_cols = g->mi_cols;
    cm->frm_size.frame_width = (uint16_t)g->width, cm->frm_size.frame_height = (uint16_t)g->height;
    cm->frm_size.superres_upscaled_width = (uint16_t)g->width, cm->frm_size.superres_upscaled_height = (uint16_t)g->height;
    cm->frm_size.superres_denominator    = SCALE_NUMERATOR;
    cm->subsampling_x = cm->subsampling_y = 1;
    cm->bit_depth                         = g->bit_depth;
    cm->use_highbitdepth                  = g->sample_bytes == 2;
    for (int p = 0; p < 3; p++) {
        pcs->rst_info[p].frame_restoration_type = (RestorationType)lr->frame_restoration_type[p];
        pcs->rst_info[p].restoration_unit_size  = lr->restoration_unit_size[p];
    }
    if (svt_av1_alloc_restoration_buffers(pcs, cm) != EB_ErrorNone)
        return -1;
    for (int p = 0; p < 3; p++) {
        RestorationInfo *rsi = &pcs->rst_info[p];
        for (int u = 0; u < rsi->units_per_tile; u++) {
            RestorationUnitInfo *ri = &rsi->unit_info[u];
            const svtf_lr_unit  *su = &units[p][u];
            memset(ri, 0, sizeof(*ri));
            ri->restoration_type = (RestorationType)su->restoration_type;
            memcpy(ri->wiener_info.vfilter, su->wiener_vfilter, sizeof(ri->wiener_info.vfilter));
            memcpy(ri->wiener_info.hfilter, su->wiener_hfilter, sizeof(ri->wiener_info.hfilter));
            ri->sgrproj_info.ep = su->sgr_ep, ri->sgrproj_info.xqd[0] = su->sgr_xqd[0], ri->sgrproj_info.xqd[1] = su->sgr_xqd[1];
        }
    }
    void *const cur[3] = {y, cb, cr}, *const dbk[3] = {dy, dcb, dcr};
    const int   cst[3] = {sy, scb, scr}, dst[3] = {dsy, dscb, dscr};
    lr_frame    fc, fd;
    lr_frame_init(&fc, g, cur, cst);
    lr_frame_init(&fd, g, dbk, dst);
    svt_av1_loop_restoration_save_boundary_lines(&fd.f, cm, 0); /* EbDlfProcess.c:118 */
    svt_av1_loop_restoration_save_boundary_lines(&fc.f, cm, 1); /* EbCdefProcess.c:668 */
    svt_av1_loop_restoration_filter_frame(tmp, &fc.f, cm, lr->optimized_lr);
    lr_frame_store(&fc, g, cur, cst);
    for (int p = 0; p < 3; p++) {
        free(fc.alloc[p]), free(fd.alloc[p]);
        free(pcs->rst_info[p].unit_info), free(pcs->rst_info[p].boundaries.stripe_boundary_above), free(pcs->rst_info[p].boundaries.stripe_boundary_below);
    }
    free(tmp), free(cm), free(pcs);
    return 0;
}

REF_API int ref_lr_units(const svtf_geom *g, int plane, int unit_size, int *horz, int *vert) {
    Av1Common       cm;
    RestorationInfo rsi;
    memset(&cm, 0, sizeof(cm)), memset(&rsi, 0, sizeof(rsi));
    cm.frm_size.frame_width = (uint16_t)g->width, cm.frm_size.frame_height = (uint16_t)g->height;
    cm.frm_size.superres_upscaled_width = (uint16_t)g->width, cm.frm_size.superres_upscaled_height = (uint16_t)g->height;
    cm.subsampling_x = cm.subsampling_y = 1;
    rsi.restoration_unit_size           = unit_size;
    if (svt_av1_alloc_restoration_struct(&cm, &rsi, plane > 0) != EB_ErrorNone)
        return -1;
    *horz = rsi.horz_units_per_tile, *vert = rsi.vert_units_per_tile;
    free(rsi.unit_info);
    return 0;
}
