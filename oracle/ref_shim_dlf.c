/*
 * ref_shim_dlf.c — drives the reference's deblocking LEVEL SEARCH.  TEST INFRASTRUCTURE ONLY.
 *
 * search_filter_level / try_filter_frame / picture_sse_calculations are static in
 * Encoder/Codec/EbDeblockingFilter.c, so that file is compiled as part of THIS translation unit (it is therefore not in
 * oracle/Makefile's REF_SRCS); its public functions (svt_av1_loop_filter_frame, ...) are exported from here for
 * ref_shim.c as before.  svt_av1_pick_filter_level itself (:1129) also allocates a scratch picture through the encoder's
 * object system (EB_NEW + svt_recon_picture_buffer_desc_ctor) — the shim supplies that buffer and calls the three
 * search_filter_level invocations of :1196-1240 directly.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "EbPictureControlSet.h"
#include "EbSequenceControlSet.h"
#include "EbDeblockingFilter.h"
#include "EbDeblockingCommon.h"
#include "EbDeblockingFilter.c" /* found through -I$(REF)/Source/Lib/Encoder/Codec */

#include "ref_harness.h"

#define REF_API __attribute__((visibility("default")))

REF_API int ref_lf_search_frame(const svtf_geom *g, void *y, int sy, void *cb, int scb, void *cr, int scr, void *src_y, int ssy, void *src_cb,
                                int sscb, void *src_cr, int sscr, const ref_mi_soa *soa, const svtf_lf_params *lf,
                                const svtf_lf_search_cfg *cfg, int32_t out_levels[4]) {
    harness h;
    if (harness_init(&h, g, soa, y, sy, cb, scb, cr, scr))
        return -1;
    FrameHeader       *fh = &h.ppcs->frm_hdr;
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
    for (int i = 0; i < 4; i++) h.ppcs->curr_delta_lf[i] = lf->curr_delta_lf[i];
    fh->tx_mode                                = cfg->only_4x4 ? ONLY_4X4 : TX_MODE_SELECT;
    h.ppcs->dlf_ctrls.dlf_avg                  = (uint8_t)cfg->dlf_avg;
    h.ppcs->dlf_ctrls.early_exit_convergence   = (uint8_t)cfg->early_exit_convergence;
    /* the source picture: enhanced_pic (8-bit) / input_frame16bit (16-bit), picture_sse_calculations :716-835 */
    EbPictureBufferDesc *s = h.source;
    *s                     = *h.recon;
    s->buffer_y = src_y, s->buffer_cb = src_cb, s->buffer_cr = src_cr;
    s->stride_y = (uint16_t)ssy, s->stride_cb = (uint16_t)sscb, s->stride_cr = (uint16_t)sscr;
    h.pcs->input_frame16bit = s;
    /* scratch picture with the recon's layout (svt_copy_buffer :655 copies ALIGN8(width) x ALIGN8(height)) */
    EbPictureBufferDesc tmp = *h.recon;
    const int           sb = g->sample_bytes, ch = g->mi_rows * 4;
    tmp.buffer_y  = malloc((size_t)sy * ch * sb);
    tmp.buffer_cb = malloc((size_t)scb * (ch / 2) * sb);
    tmp.buffer_cr = malloc((size_t)scr * (ch / 2) * sb);
    svt_av1_loop_filter_init(h.pcs);
    const int32_t last[4] = {l->filter_level[0], l->filter_level[1], l->filter_level_u, l->filter_level_v};
    l->filter_level[0] = l->filter_level[1] = search_filter_level(s, &tmp, h.pcs, 0, last, NULL, 0, 2);
    if (cfg->search_chroma) {
        l->filter_level_u = search_filter_level(s, &tmp, h.pcs, 0, last, NULL, 1, 0);
        l->filter_level_v = search_filter_level(s, &tmp, h.pcs, 0, last, NULL, 2, 0);
    } else {
        l->filter_level_u = last[2], l->filter_level_v = last[3];
    }
    out_levels[0] = l->filter_level[0], out_levels[1] = l->filter_level[1], out_levels[2] = l->filter_level_u, out_levels[3] = l->filter_level_v;
    free(tmp.buffer_y), free(tmp.buffer_cb), free(tmp.buffer_cr);
    harness_free(&h);
    return 0;
}
