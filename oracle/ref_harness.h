/*
 * ref_harness.h — builds the minimal PictureControlSet / SequenceControlSet graph the reference's
 * frame-level filter functions read, from plain pointers.  Shared by oracle/ref_shim.c
 * (drives the real reference) and adapters/adapter_harness.c (drives the Tier-2 adapters with the very
 * same structures).  TEST INFRASTRUCTURE ONLY.  Requires the reference headers to be included first.
 */
#ifndef REF_HARNESS_H
#define REF_HARNESS_H
#include "../include/svtfilt_b200.h"

/* =========================================================================================
 * frame-level harness
 * ======================================================================================= */
typedef struct {
    PictureControlSet       *pcs;
    PictureParentControlSet *ppcs;
    SequenceControlSet      *scs;
    Av1Common               *cm;
    EncDecSet               *eds;
    EbPictureBufferDesc     *recon, *source;
    ModeInfo                *mi;
    ModeInfo               **grid;
} harness;

static void harness_free(harness *h) {
    if (h->pcs) {
        free(h->pcs->cdef_dir_data);
        free(h->pcs->skip_cdef_seg);
        free(h->pcs->mse_seg[0]);
        free(h->pcs->mse_seg[1]);
    }
    free(h->pcs), free(h->ppcs), free(h->scs), free(h->cm), free(h->eds), free(h->recon), free(h->source), free(h->mi),
        free(h->grid);
}

/* Mode-info SoA: arrays of mi_rows*mi_stride entries (NULL -> zeros). */
typedef struct ref_mi_soa {
    const uint8_t *bsize, *tx_depth, *skip, *ref_frame0, *mode, *segment_id;
    int32_t        mi_stride;
} ref_mi_soa;

static int harness_init(harness *h, const svtf_geom *g, const ref_mi_soa *soa, void *y, int sy, void *cb, int scb, void *cr,
                        int scr) {
    memset(h, 0, sizeof(*h));
    h->pcs    = calloc(1, sizeof(PictureControlSet));
    h->ppcs   = calloc(1, sizeof(PictureParentControlSet));
    h->scs    = calloc(1, sizeof(SequenceControlSet));
    h->cm     = calloc(1, sizeof(Av1Common));
    h->eds    = calloc(1, sizeof(EncDecSet));
    h->recon  = calloc(1, sizeof(EbPictureBufferDesc));
    h->source = calloc(1, sizeof(EbPictureBufferDesc));
    const int n = g->mi_rows * soa->mi_stride;
    h->mi       = calloc((size_t)n + soa->mi_stride + 2, sizeof(ModeInfo));
    h->grid     = calloc((size_t)n + soa->mi_stride + 2, sizeof(ModeInfo *));
    if (!h->pcs || !h->ppcs || !h->scs || !h->cm || !h->eds || !h->recon || !h->source || !h->mi || !h->grid)
        return -1;
    const int is16 = g->sample_bytes == 2;
    for (int i = 0; i < n + soa->mi_stride + 2; i++) h->grid[i] = &h->mi[i];
    for (int i = 0; i < n; i++) {
        BlockModeInfoEnc *b = &h->mi[i].mbmi.block_mi;
        b->bsize            = soa->bsize ? soa->bsize[i] : BLOCK_4X4;
        b->tx_depth         = soa->tx_depth ? soa->tx_depth[i] : 0;
        b->skip             = soa->skip ? soa->skip[i] : 0;
        b->ref_frame[0]     = soa->ref_frame0 ? (MvReferenceFrame)soa->ref_frame0[i] : INTRA_FRAME;
        b->mode             = soa->mode ? soa->mode[i] : DC_PRED;
        b->segment_id       = soa->segment_id ? soa->segment_id[i] : 0;
    }
    PictureControlSet *pcs        = h->pcs;
    pcs->scs                      = h->scs;
    pcs->ppcs                     = h->ppcs;
    pcs->mi_grid_base             = h->grid;
    pcs->mi_stride                = soa->mi_stride;
    h->ppcs->scs                  = h->scs;
    h->ppcs->av1_cm               = h->cm;
    h->cm->mi_rows                = g->mi_rows;
    h->cm->mi_cols                = g->mi_cols;
    h->cm->mi_stride              = soa->mi_stride;
    h->ppcs->aligned_width        = (uint16_t)(g->mi_cols * 4);
    h->ppcs->aligned_height       = (uint16_t)(g->mi_rows * 4);
    h->ppcs->enc_dec_ptr          = h->eds;
    h->eds->recon_pic             = h->recon;
    h->eds->recon_pic_16bit       = h->recon;
    h->ppcs->is_ref               = FALSE;
    h->ppcs->render_width         = (uint16_t)g->width;
    h->ppcs->render_height        = (uint16_t)g->height;
    h->ppcs->enhanced_pic         = h->source;
    h->scs->is_16bit_pipeline     = is16;
    h->scs->sb_size               = 64;
    h->scs->super_block_size      = 64;
    h->scs->seq_header.sb_size    = BLOCK_64X64;
    h->scs->static_config.encoder_bit_depth = g->bit_depth;
    h->scs->max_input_luma_width  = (uint16_t)g->width;
    h->scs->max_input_luma_height = (uint16_t)g->height;
    h->scs->max_input_pad_right   = 0;
    h->scs->max_input_pad_bottom  = 0;
    h->scs->subsampling_x = h->scs->subsampling_y = 1;
    EbPictureBufferDesc *r = h->recon;
    r->buffer_y = y, r->buffer_cb = cb, r->buffer_cr = cr;
    r->stride_y = (uint16_t)sy, r->stride_cb = (uint16_t)scb, r->stride_cr = (uint16_t)scr;
    r->org_x = r->org_y = 0;
    r->width = (uint16_t)g->width, r->height = (uint16_t)g->height;
    r->max_width = r->width, r->max_height = r->height;
    r->bit_depth    = (EbBitDepth)g->bit_depth;
    r->color_format = EB_YUV420;
    const int nfb      = ((g->mi_rows + 15) / 16) * ((g->mi_cols + 15) / 16);
    pcs->cdef_dir_data = calloc(nfb, sizeof(CdefDirData));
    pcs->skip_cdef_seg = calloc(nfb, 1);
    pcs->mse_seg[0]    = calloc(nfb, sizeof(uint64_t) * TOTAL_STRENGTHS);
    pcs->mse_seg[1]    = calloc(nfb, sizeof(uint64_t) * TOTAL_STRENGTHS);
    pcs->cdef_segments_column_count = 1;
    pcs->cdef_segments_row_count    = 1;
    return 0;
}


#endif
