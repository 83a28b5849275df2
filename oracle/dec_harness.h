/*
 * dec_harness.h — builds the part of an EbDecHandle the decoder's post-recon filters read
 * (svt_aom_dec_av1_loop_filter_frame Decoder/Codec/EbDecLF.c:697, svt_cdef_frame EbDecCdef.c:507, dec_ccso_frame
 * EbDecCcso.c:384) from plain per-4x4 arrays.  Shared by oracle/ref_shim_dec.c (drives the REAL decoder functions) and
 * adapters/adapter_harness_dec.c (drives the Tier-2 decoder adapters on the very same structures).
 * TEST INFRASTRUCTURE ONLY.  Requires the reference's decoder headers to be included first.
 *
 * The decoder keeps, per superblock, the blocks in decode order (SBInfo.sb_mode_info[num_block]) with their transform
 * units (SBInfo.sb_trans_info[0] luma, [1] chroma: the U units of a block followed by its V units), a frame map from
 * 4x4 position to (SB, block index) (FrameMiMap.p_mi_offset) and LfCtxt.tx_size_l / tx_size_uv per 4x4
 * (svt_aom_fill_4x4_lf_param, EbDecLF.c:63).  All of it is derived here from bsize / tx_size per 4x4:
 *   block origins   a 4x4 is the origin of its block iff its row/col are multiples of the block's height/width (AV1
 *                   blocks are aligned to their own size); blocks of an SB are ordered by the Morton code of the
 *                   origin, which keeps the sub-8x8 siblings adjacent as dec_is_8x8_block_skip (EbDecCdef.c:22-34) needs
 *   luma TUs        origins of the tx_size_l map inside the block (and inside the frame)
 *   chroma TUs      only blocks that carry chroma (is_chroma_reference); uniform size tx_size_uv at the block's (odd, odd)
 *                   4x4, tiled over the chroma block
 */
#ifndef DEC_HARNESS_H
#define DEC_HARNESS_H
#include "../include/svtfilt_b200.h"

typedef struct dec_mi_soa {
    const uint8_t *bsize, *tx_size_y, *tx_size_uv, *skip, *ref_frame0, *mode, *segment_id; /* [mi_rows][mi_stride] */
    int32_t        mi_stride;
    int32_t        sb128;            /* 1: 128x128 superblocks */
    const int32_t *sb_delta_lf;      /* [sb_rows*sb_cols][4] or NULL (zeros) */
} dec_mi_soa;

typedef struct {
    EbDecHandle         *dec;
    EbDecPicBuf         *pic_buf;
    EbPictureBufferDesc *recon;
    LfCtxt               lf_ctxt;
    SBInfo              *sb_info;
    SBInfo             **pps;
    uint16_t            *mi_offset;
    BlockModeInfo       *blocks;
    TransformInfo_t     *tus[2];
    int32_t             *delta_lf;
    int8_t              *cdef_strength;
    uint8_t             *ccso_blk[3];
    int                  sb_cols, sb_rows, sb_mi;
} dec_harness;

static void dec_harness_free(dec_harness *h) {
    free(h->dec), free(h->pic_buf), free(h->recon), free(h->lf_ctxt.tx_size_l), free(h->lf_ctxt.tx_size_uv), free(h->sb_info);
    free(h->pps), free(h->mi_offset), free(h->blocks), free(h->tus[0]), free(h->tus[1]), free(h->delta_lf), free(h->cdef_strength);
    for (int p = 0; p < 3; p++) free(h->ccso_blk[p]);
}

static uint32_t dech_morton(uint32_t r, uint32_t c) {
    uint32_t m = 0;
    for (int b = 0; b < 6; b++) m |= ((c >> b) & 1u) << (2 * b) | ((r >> b) & 1u) << (2 * b + 1);
    return m;
}
typedef struct {
    uint32_t key;
    uint16_t r, c;
} dech_org;
static int dech_cmp(const void *a, const void *b) {
    const uint32_t x = ((const dech_org *)a)->key, y = ((const dech_org *)b)->key;
    return x < y ? -1 : x > y;
}

static int dec_harness_init(dec_harness *h, const svtf_geom *g, const dec_mi_soa *soa, void *y, int sy, void *cb, int scb, void *cr,
                            int scr) {
    memset(h, 0, sizeof(*h));
    const int mi_rows = g->mi_rows, mi_cols = g->mi_cols, st = soa->mi_stride;
    const int sb_log2 = soa->sb128 ? 7 : 6, sb_mi = 1 << (sb_log2 - 2);
    const int sb_cols = (mi_cols + sb_mi - 1) / sb_mi, sb_rows = (mi_rows + sb_mi - 1) / sb_mi, nsb = sb_cols * sb_rows;
    const int cols_al = sb_cols * sb_mi, rows_al = sb_rows * sb_mi, factor = soa->sb128 ? 4 : 1;
    h->sb_cols = sb_cols, h->sb_rows = sb_rows, h->sb_mi = sb_mi;
    h->dec                = calloc(1, sizeof(EbDecHandle));
    h->pic_buf            = calloc(1, sizeof(EbDecPicBuf));
    h->recon              = calloc(1, sizeof(EbPictureBufferDesc));
    h->lf_ctxt.tx_size_l  = calloc((size_t)(rows_al + 2) * st, sizeof(TxSize));
    h->lf_ctxt.tx_size_uv = calloc((size_t)(rows_al + 2) * st, sizeof(TxSize));
    h->sb_info            = calloc(nsb, sizeof(SBInfo));
    h->pps                = calloc(nsb, sizeof(SBInfo *));
    h->mi_offset          = calloc((size_t)rows_al * cols_al, sizeof(uint16_t));
    h->blocks             = calloc((size_t)mi_rows * mi_cols + 8, sizeof(BlockModeInfo)); /* +8: dec_is_8x8_block_skip reads mbmi[1..3] */
    h->tus[0]             = calloc((size_t)mi_rows * mi_cols, sizeof(TransformInfo_t));
    h->tus[1]             = calloc((size_t)mi_rows * mi_cols, sizeof(TransformInfo_t));
    h->delta_lf           = calloc((size_t)(nsb + sb_cols + 2) * 4, sizeof(int32_t));
    h->cdef_strength      = calloc((size_t)nsb * factor, 1);
    for (int p = 0; p < 3; p++) h->ccso_blk[p] = calloc((size_t)nsb * factor, 1);
    if (!h->dec || !h->pic_buf || !h->recon || !h->lf_ctxt.tx_size_l || !h->lf_ctxt.tx_size_uv || !h->sb_info || !h->pps || !h->mi_offset ||
        !h->blocks || !h->tus[0] || !h->tus[1] || !h->delta_lf || !h->cdef_strength)
        return -1;
    EbDecHandle *d  = h->dec;
    FrameHeader *fh = &d->frame_header;
    fh->mi_cols = mi_cols, fh->mi_rows = mi_rows, fh->mi_stride = st;
    fh->frame_size.frame_width = (uint16_t)g->width, fh->frame_size.frame_height = (uint16_t)g->height;
    fh->frame_size.superres_upscaled_width = (uint16_t)g->width, fh->frame_size.superres_upscaled_height = (uint16_t)g->height;
    d->seq_header.color_config.subsampling_x = d->seq_header.color_config.subsampling_y = 1;
    d->seq_header.color_config.bit_depth   = (EbBitDepth)g->bit_depth;
    d->seq_header.color_config.mono_chrome = 0;
    d->seq_header.sb_size                  = soa->sb128 ? BLOCK_128X128 : BLOCK_64X64;
    d->seq_header.sb_size_log2             = (uint8_t)sb_log2;
    d->seq_header.use_128x128_superblock   = (uint8_t)soa->sb128;
    d->is_16bit_pipeline                   = g->sample_bytes == 2;
    EbPictureBufferDesc *r = h->recon;
    r->buffer_y = y, r->buffer_cb = cb, r->buffer_cr = cr;
    r->stride_y = (uint16_t)sy, r->stride_cb = (uint16_t)scb, r->stride_cr = (uint16_t)scr;
    r->width = (uint16_t)g->width, r->height = (uint16_t)g->height, r->max_width = r->width, r->max_height = r->height;
    r->bit_depth         = (EbBitDepth)g->bit_depth;
    r->is_16bit_pipeline = g->sample_bytes == 2;
    r->color_format      = EB_YUV420;
    h->pic_buf->ps_pic_buf = r;
    d->cur_pic_buf[0]      = h->pic_buf;
    MainFrameBuf *mf = &d->main_frame_buf;
    mf->sb_cols = sb_cols, mf->sb_rows = sb_rows, mf->num_mis_in_sb = sb_mi * sb_mi;
    FrameMiMap *mm = &mf->frame_mi_map;
    mm->pps_sb_info = h->pps, mm->p_mi_offset = h->mi_offset, mm->sb_size_log2 = (uint8_t)sb_log2;
    mm->mi_cols_algnsb = cols_al, mm->mi_rows_algnsb = rows_al, mm->sb_cols = sb_cols, mm->sb_rows = sb_rows, mm->num_mis_in_sb_wd = sb_mi;
    CurFrameBuf *cf = &mf->cur_frame_bufs[0];
    cf->sb_info = h->sb_info, cf->mode_info = h->blocks, cf->trans_info[0] = h->tus[0], cf->trans_info[1] = h->tus[1];
    cf->cdef_strength = h->cdef_strength, cf->delta_lf = h->delta_lf + (sb_cols + 1) * 4;
    cf->ccso_blk_y = h->ccso_blk[0], cf->ccso_blk_u = h->ccso_blk[1], cf->ccso_blk_v = h->ccso_blk[2];
    for (int r4 = 0; r4 < mi_rows; r4++)
        for (int c4 = 0; c4 < mi_cols; c4++) {
            h->lf_ctxt.tx_size_l[r4 * st + c4]  = (TxSize)soa->tx_size_y[r4 * st + c4];
            h->lf_ctxt.tx_size_uv[r4 * st + c4] = (TxSize)soa->tx_size_uv[r4 * st + c4];
        }
    BlockModeInfo   *blk = h->blocks;
    TransformInfo_t *tl = h->tus[0], *tc = h->tus[1];
    dech_org        *org = malloc(sizeof(dech_org) * sb_mi * sb_mi);
    int16_t         *idx = malloc(sizeof(int16_t) * sb_mi * sb_mi);
    if (!org || !idx)
        return -1;
    for (int sbr = 0; sbr < sb_rows; sbr++)
        for (int sbc = 0; sbc < sb_cols; sbc++) {
            const int sb = sbr * sb_cols + sbc, r0 = sbr * sb_mi, c0 = sbc * sb_mi;
            SBInfo   *si = &h->sb_info[sb];
            h->pps[sb]   = si;
            si->sb_mode_info = blk, si->sb_trans_info[0] = tl, si->sb_trans_info[1] = tc;
            si->sb_cdef_strength = h->cdef_strength + sb * factor;
            si->sb_delta_lf      = cf->delta_lf + sb * 4;
            si->sb_ccso_blk_y = h->ccso_blk[0] + sb * factor, si->sb_ccso_blk_u = h->ccso_blk[1] + sb * factor,
            si->sb_ccso_blk_v = h->ccso_blk[2] + sb * factor;
            if (soa->sb_delta_lf)
                for (int i = 0; i < 4; i++) si->sb_delta_lf[i] = soa->sb_delta_lf[sb * 4 + i];
            int n = 0;
            for (int r4 = r0; r4 < r0 + sb_mi && r4 < mi_rows; r4++)
                for (int c4 = c0; c4 < c0 + sb_mi && c4 < mi_cols; c4++) {
                    const BlockSize bs = (BlockSize)soa->bsize[r4 * st + c4];
                    if (r4 % mi_size_high[bs] == 0 && c4 % mi_size_wide[bs] == 0)
                        org[n].key = dech_morton(r4 - r0, c4 - c0), org[n].r = (uint16_t)r4, org[n].c = (uint16_t)c4, n++;
                }
            qsort(org, n, sizeof(*org), dech_cmp);
            memset(idx, -1, sizeof(int16_t) * sb_mi * sb_mi);
            int ntl = 0, ntc = 0;
            for (int b = 0; b < n; b++) {
                const int       br = org[b].r, bc = org[b].c, i = br * st + bc;
                const BlockSize bs = (BlockSize)soa->bsize[i];
                const int       bw4 = mi_size_wide[bs], bh4 = mi_size_high[bs];
                BlockModeInfo  *m = &blk[b];
                m->bsize = bs, m->mode = (PredictionMode)soa->mode[i], m->skip = (int8_t)soa->skip[i];
                m->segment_id   = (int8_t)soa->segment_id[i];
                m->ref_frame[0] = (MvReferenceFrame)soa->ref_frame0[i], m->ref_frame[1] = NONE_FRAME;
                m->mi_row_in_sb = (int8_t)(br - r0), m->mi_col_in_sb = (int8_t)(bc - c0);
                for (int r4 = br; r4 < br + bh4 && r4 < mi_rows; r4++)
                    for (int c4 = bc; c4 < bc + bw4 && c4 < mi_cols; c4++) {
                        idx[(r4 - r0) * sb_mi + (c4 - c0)] = (int16_t)b;
                        h->mi_offset[r4 * cols_al + c4]    = (uint16_t)b;
                    }
                /* luma transform units */
                m->first_txb_offset[0] = (uint16_t)ntl;
                int cnt                = 0;
                for (int r4 = br; r4 < br + bh4 && r4 < mi_rows; r4++)
                    for (int c4 = bc; c4 < bc + bw4 && c4 < mi_cols; c4++) {
                        const TxSize tx = (TxSize)soa->tx_size_y[r4 * st + c4];
                        if ((r4 - br) % tx_size_high_unit[tx] || (c4 - bc) % tx_size_wide_unit[tx])
                            continue;
                        TransformInfo_t *t = &tl[ntl++];
                        t->tx_size = tx, t->txb_x_offset = (uint8_t)(c4 - bc), t->txb_y_offset = (uint8_t)(r4 - br), cnt++;
                    }
                m->num_tus[0] = (uint8_t)cnt;
                /* chroma transform units (U then V), only on blocks that carry chroma */
                m->first_txb_offset[1] = (uint16_t)ntc;
                cnt                    = 0;
                if (((br & 1) || !(bh4 & 1)) && ((bc & 1) || !(bw4 & 1))) {
                    const int    er = br & ~1, ec = bc & ~1, cw = bw4 > 1 ? bw4 >> 1 : 1, ch = bh4 > 1 ? bh4 >> 1 : 1;
                    const TxSize tx = (TxSize)soa->tx_size_uv[(br | 1) * st + (bc | 1)];
                    for (int pass = 0; pass < 2; pass++)
                        for (int yy = 0; yy < ch; yy += tx_size_high_unit[tx])
                            for (int xx = 0; xx < cw; xx += tx_size_wide_unit[tx]) {
                                if (er + (yy << 1) >= mi_rows || ec + (xx << 1) >= mi_cols)
                                    continue;
                                TransformInfo_t *t = &tc[ntc++];
                                t->tx_size = tx, t->txb_x_offset = (uint8_t)xx, t->txb_y_offset = (uint8_t)yy;
                                cnt += pass == 0;
                            }
                }
                m->num_tus[1] = (uint8_t)cnt;
            }
            si->num_block = n;
            blk += n, tl += ntl, tc += ntc;
        }
    free(org), free(idx);
    return 0;
}

static void dec_harness_fill_lf(dec_harness *h, const svtf_lf_params *lf) {
    FrameHeader       *fh = &h->dec->frame_header;
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
}

/* frame_header.cdef_params + per-64x64 strengths (SBInfo.sb_cdef_strength, index = quadrant inside a 128x128 SB,
 * Decoder/Codec/EbDecCdef.c:104-110) */
static void dec_harness_fill_cdef(dec_harness *h, const svtf_geom *g, const dec_mi_soa *soa, const svtf_cdef_params *cdef, const int8_t *fb_idx) {
    FrameHeader *fh = &h->dec->frame_header;
    fh->cdef_params.cdef_damping = (uint8_t)cdef->cdef_damping;
    fh->cdef_params.cdef_bits    = (uint8_t)cdef->cdef_bits;
    for (int i = 0; i < CDEF_MAX_STRENGTHS && i < SVTF_CDEF_MAX_STRENGTHS; i++) {
        fh->cdef_params.cdef_y_strength[i]  = cdef->cdef_y_strength[i];
        fh->cdef_params.cdef_uv_strength[i] = cdef->cdef_uv_strength[i];
    }
    const int nvfb = (g->mi_rows + 15) / 16, nhfb = (g->mi_cols + 15) / 16;
    for (int fbr = 0; fbr < nvfb; fbr++)
        for (int fbc = 0; fbc < nhfb; fbc++) {
            SBInfo *si = soa->sb128 ? &h->sb_info[(fbr >> 1) * h->sb_cols + (fbc >> 1)] : &h->sb_info[fbr * h->sb_cols + fbc];
            si->sb_cdef_strength[soa->sb128 ? (fbc & 1) + 2 * (fbr & 1) : 0] = fb_idx[fbr * nhfb + fbc];
        }
}

static void dec_harness_fill_ccso(dec_harness *h, const svtf_ccso_params *p, const uint8_t blk[3]) {
    CcsoInfo *ci = &h->dec->frame_header.ccso_info;
    for (int pl = 0; pl < 3; pl++) {
        ci->ccso_enable[pl] = p->ccso_enable[pl] != 0;
        memcpy(ci->filter_offset[pl], p->filter_offset[pl], SVTF_CCSO_LUT_SIZE);
        ci->quant_idx[pl] = p->quant_idx[pl], ci->ext_filter_support[pl] = p->ext_filter_support[pl];
        ci->ccso_bo_only[pl] = p->ccso_bo_only[pl], ci->max_band_log2[pl] = p->max_band_log2[pl], ci->edge_clf[pl] = p->edge_clf[pl];
    }
    h->sb_info[0].sb_ccso_blk_y[0] = blk[0], h->sb_info[0].sb_ccso_blk_u[0] = blk[1], h->sb_info[0].sb_ccso_blk_v[0] = blk[2];
}

/* ---- loop restoration (needs EbDecProcessFrame.h + EbRestoration.h; define DEC_HARNESS_LR before including) ----
 * LrCtxt as init_lr_ctxt builds it for one thread (Decoder/Codec/EbDecMemInit.c:430-503): one RestorationLineBuffers and
 * one scratch per plane, the deblocked-line store per plane, the [SB x 64] output scratch; lr_unit[plane] indexed
 * [unit_row * sb_cols + unit_col] (EbDecMemInit.c:292-303, EbDecRestoration.c:341).  The picture gets a border: lr_pad_pic
 * (EbDecRestoration.c:105) writes 6 samples around it and the Wiener kernels read a little further. */
#ifdef DEC_HARNESS_LR
#define DECH_BORDER 64
typedef struct {
    LrCtxt                  ctx;
    RestorationLineBuffers *rlb[3], **rlbs0, ***rlbs;
    int32_t                *tmp, **tmps;
    void                   *padded[3], *orig[3];
    int                     orig_st[3];
} dec_harness_lr;

static void dec_harness_lr_free(dec_harness_lr *l) {
    for (int p = 0; p < 3; p++) {
        free(l->ctx.lr_unit[p]), free(l->rlb[p]), free(l->padded[p]);
        free(l->ctx.boundaries[p].stripe_boundary_above), free(l->ctx.boundaries[p].stripe_boundary_below);
    }
    free(l->rlbs0), free(l->rlbs), free(l->tmp), free(l->tmps), free(l->ctx.dst);
}

/* units[p]: svtf layout [vert][horz] (svtf_lr_units); pictures: moved into bordered buffers (dec_harness_lr_store copies back) */
static int dec_harness_fill_lr(dec_harness *h, dec_harness_lr *l, const svtf_geom *g, const svtf_lr_params *lr,
                               const svtf_lr_unit *const units[3]) {
    memset(l, 0, sizeof(*l));
    EbDecHandle *d   = h->dec;
    const int    sb  = g->sample_bytes, nsb = h->sb_cols * h->sb_rows;
    d->pv_lr_ctxt    = &l->ctx;
    d->dec_config.threads = 1;
    d->seq_header.max_frame_width = (uint16_t)g->width, d->seq_header.max_frame_height = (uint16_t)g->height;
    Av1Common *cm = &d->cm;
    cm->mi_rows = g->mi_rows, cm->mi_cols = g->mi_cols;
    cm->frm_size.frame_width = (uint16_t)g->width, cm->frm_size.frame_height = (uint16_t)g->height;
    cm->frm_size.superres_upscaled_width = (uint16_t)g->width, cm->frm_size.superres_upscaled_height = (uint16_t)g->height;
    cm->frm_size.superres_denominator = SCALE_NUMERATOR;
    d->frame_header.frame_size.superres_denominator = SCALE_NUMERATOR;
    cm->subsampling_x = cm->subsampling_y = 1;
    cm->bit_depth = g->bit_depth, cm->use_highbitdepth = sb == 2;
    l->ctx.dec_handle_ptr = d;
    l->ctx.is_thread_min  = TRUE;
    l->rlbs = calloc(1, sizeof(*l->rlbs)), l->rlbs0 = calloc(3, sizeof(*l->rlbs0));
    l->tmps = calloc(1, sizeof(*l->tmps)), l->tmp = malloc(RESTORATION_TMPBUF_SIZE);
    l->ctx.dst = calloc((size_t)(MAX_SB_SIZE + 8) * RESTORATION_PROC_UNIT_SIZE, sb);
    if (!l->rlbs || !l->rlbs0 || !l->tmps || !l->tmp || !l->ctx.dst)
        return -1;
    l->rlbs[0] = l->rlbs0, l->tmps[0] = l->tmp;
    l->ctx.rlbs = l->rlbs, l->ctx.rst_tmpbuf = l->tmps;
    const int num_stripes = (RESTORATION_UNIT_OFFSET + g->height + 63) / 64;
    for (int p = 0; p < 3; p++) {
        const int ss = p ? 1 : 0;
        LrParams *fp = &d->frame_header.lr_params[p];
        fp->frame_restoration_type = (RestorationType)lr->frame_restoration_type[p];
        fp->loop_restoration_size  = (uint16_t)lr->restoration_unit_size[p];
        fp->lr_size_log2           = 0;
        while ((1 << fp->lr_size_log2) < lr->restoration_unit_size[p]) fp->lr_size_log2++;
        l->rlb[p] = l->rlbs0[p] = calloc(1, sizeof(RestorationLineBuffers));
        l->ctx.lr_unit[p]   = calloc(nsb, sizeof(RestorationUnitInfo));
        l->ctx.lr_stride[p] = h->sb_cols;
        const int plane_w = ((g->width + ss) >> ss) + 2 * RESTORATION_EXTRA_HORZ, stride = (plane_w + 31) & ~31;
        const int size    = num_stripes * stride * RESTORATION_CTX_VERT * sb;
        RestorationStripeBoundaries *b = &l->ctx.boundaries[p];
        b->stripe_boundary_above = calloc(size, 1), b->stripe_boundary_below = calloc(size, 1);
        b->stripe_boundary_size = size, b->stripe_boundary_stride = stride;
        if (!l->rlb[p] || !l->ctx.lr_unit[p] || !b->stripe_boundary_above || !b->stripe_boundary_below)
            return -1;
        if (lr->frame_restoration_type[p] != SVTF_RESTORE_NONE) {
            /* svt_av1_lr_count_units_in_tile (Common/Codec/EbRestoration.c:1318): max((size + unit / 2) / unit, 1) */
            const int U = lr->restoration_unit_size[p], pw = (g->width + ss) >> ss, ph = (g->height + ss) >> ss;
            const int hu = (pw + U / 2) / U > 1 ? (pw + U / 2) / U : 1, vu = (ph + U / 2) / U > 1 ? (ph + U / 2) / U : 1;
            if (hu > h->sb_cols || vu * h->sb_cols > nsb)
                return -1;
            for (int v = 0; v < vu; v++)
                for (int u = 0; u < hu; u++) {
                    RestorationUnitInfo *ri = &l->ctx.lr_unit[p][v * h->sb_cols + u];
                    const svtf_lr_unit  *su = &units[p][v * hu + u];
                    ri->restoration_type = (RestorationType)su->restoration_type;
                    memcpy(ri->wiener_info.vfilter, su->wiener_vfilter, sizeof(ri->wiener_info.vfilter));
                    memcpy(ri->wiener_info.hfilter, su->wiener_hfilter, sizeof(ri->wiener_info.hfilter));
                    ri->sgrproj_info.ep = su->sgr_ep, ri->sgrproj_info.xqd[0] = su->sgr_xqd[0], ri->sgrproj_info.xqd[1] = su->sgr_xqd[1];
                }
        }
    }
    /* bordered picture */
    EbPictureBufferDesc *r = h->recon;
    uint8_t **bufs[3] = {&r->buffer_y, &r->buffer_cb, &r->buffer_cr};
    uint16_t *sts[3]  = {&r->stride_y, &r->stride_cb, &r->stride_cr};
    for (int p = 0; p < 3; p++) {
        const int ss = p ? 1 : 0, cw = (g->mi_cols * 4) >> ss, chh = (g->mi_rows * 4) >> ss, bd = DECH_BORDER >> ss;
        const int stride = (cw + 2 * bd + 31) & ~31;
        l->orig[p] = *bufs[p], l->orig_st[p] = *sts[p];
        uint8_t *buf = l->padded[p] = calloc((size_t)stride * (chh + 2 * bd), sb);
        if (!buf)
            return -1;
        for (int y = 0; y < chh; y++)
            memcpy(buf + ((size_t)(y + bd) * stride + bd) * sb, (uint8_t *)l->orig[p] + (size_t)y * l->orig_st[p] * sb, (size_t)cw * sb);
        *bufs[p] = buf, *sts[p] = (uint16_t)stride;
    }
    r->org_x = r->org_y = DECH_BORDER;
    return 0;
}

/* bordered picture -> the caller's planes (visible size: lr_pad_pic overwrites coded rows below the picture) */
static void dec_harness_lr_store(dec_harness *h, dec_harness_lr *l, const svtf_geom *g) {
    EbPictureBufferDesc *r = h->recon;
    const uint8_t *bufs[3] = {r->buffer_y, r->buffer_cb, r->buffer_cr};
    const int      sts[3]  = {r->stride_y, r->stride_cb, r->stride_cr}, sb = g->sample_bytes;
    for (int p = 0; p < 3; p++) {
        const int ss = p ? 1 : 0, w = (g->width + ss) >> ss, hh = (g->height + ss) >> ss, bd = DECH_BORDER >> ss;
        for (int y = 0; y < hh; y++)
            memcpy((uint8_t *)l->orig[p] + (size_t)y * l->orig_st[p] * sb, bufs[p] + ((size_t)(y + bd) * sts[p] + bd) * sb, (size_t)w * sb);
    }
}
#endif /* DEC_HARNESS_LR */
#endif /* DEC_HARNESS_H */
