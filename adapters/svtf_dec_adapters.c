/*
 * svtf_dec_adapters.c — Tier-2 of the drop-in boundary, decoder side: the reference decoder's frame-level post-recon
 * filters (same names, same signatures) as thin marshalling over the Tier-1 C ABI (include/svtfilt_b200.h).  Compiled
 * against the reference's decoder headers; linked INSTEAD of the bodies in
 *   Decoder/Codec/EbDecLF.c:697    svt_aom_dec_av1_loop_filter_frame
 *   Decoder/Codec/EbDecCdef.c:507  svt_cdef_frame
 *   Decoder/Codec/EbDecCcso.c:384  dec_ccso_frame
 *   Decoder/Codec/EbDecRestoration.c:429 / :473  svt_aom_dec_av1_loop_restoration_filter_frame / _save_boundary_lines
 *   Decoder/Codec/EbDecProcess.c:668 / :810 / :1022  the three `_mt` stage functions every decoder thread calls when
 *                                   EbSvtAv1DecConfiguration.threads > 1 (end of this file)
 * (decoder section).  The multi-threaded row workers (svt_aom_dec_loop_filter_row EbDecLF.c:650,
 * svt_cdef_sb_row_mt EbDecCdef.c:385) exist to spread the CPU loops over worker threads with spin-wait row counters; with
 * the filters on the GPU the frame-level call is made once, from the thread that called the decoder.
 *
 * What an adapter does: (1) flatten the decoder's SB -> block -> transform-unit side info into one packed word per 4x4
 * (svtf_set_mi_soa_dec: block size / skip / ref / mode / segment from the covering BlockModeInfo through FrameMiMap, the
 * transform sizes from LfCtxt.tx_size_l / tx_size_uv, which the parser already keeps per 4x4), (2) upload the picture,
 * (3) launch, (4) download into the same buffers.  Errors: the reference functions return void; a failing Tier-1 call
 * logs and aborts — there is no CPU fallback.
 */
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "EbDecHandle.h"
#include "EbDecUtils.h"
#include "EbDecLF.h"
#include "EbDecCdef.h"
#include "EbDecCcso.h"
#include "EbCcso.h"
#include "EbDecProcessFrame.h"
#include "EbDecRestoration.h"
#include "EbRestoration.h"

#include "../include/svtfilt_b200.h"
#include "../include/svtfilt_b200_adapters.h"

/* 1: reproduce what dec_ccso_frame does today where that is well defined (see dec_ccso_frame below);
 * 0: the encoder/common definition of CCSO (Common/Codec/EbCcso.c) on the decoder's data */
#ifndef SVTF_DEC_CCSO_AS_REFERENCE
#define SVTF_DEC_CCSO_AS_REFERENCE 1
#endif

#define SVTF_DEC_MAX_CACHED 16
typedef struct {
    const EbDecHandle *dec;
    svtf_ctx          *ctx;
    svtf_geom          geom;     /* of the stage running: the coded (narrow) picture up to CCSO, the upscaled one behind super-resolution */
    svtf_geom          max_geom; /* what the context was created for: the upscaled width */
    int                resident;  /* the frame being decoded is on the device */
    int                have_tx;   /* the mi words on the device carry the transform sizes (set by the LF call) */
    uint8_t           *dbk[3];    /* stand-alone loop restoration: the deblocked rows around the stripe boundaries */
    int                have_dbk;
} DecSlot;
static DecSlot         g_slots[SVTF_DEC_MAX_CACHED];
static pthread_mutex_t g_lock     = PTHREAD_MUTEX_INITIALIZER;
static int             g_device   = 0;
static int             g_resident = 0;

void svtf_dec_adapter_set_device(int device) { g_device = device; }
void svtf_dec_adapter_set_resident(int on) { g_resident = on; }

static void die(const char *where, svtf_ctx *ctx) {
    fprintf(stderr, "Svt[error]: svtfilt_b200 decoder adapter: %s failed: %s (no CPU fallback on this path)\n", where,
            svtf_last_error(ctx));
    abort();
}
#define CHECK(call, ctx)       \
    do {                       \
        if ((call) != SVTF_OK) \
            die(#call, ctx);   \
    } while (0)

/* Call counters, printed at exit when SVTF_ADAPTER_STATS is set (see svtf_enc_adapters.c). */
enum { DS_LF, DS_CDEF, DS_CCSO, DS_LR, DS_SUPERRES, DS_N };
static const char *const g_stat_name[DS_N] = {"dec_loop_filter_frame", "dec_cdef_frame", "dec_ccso_frame", "dec_loop_restoration_filter_frame", "dec_superres_upscale"};
static unsigned long     g_stat[DS_N];
static pthread_once_t    g_stat_once = PTHREAD_ONCE_INIT;
static void              stat_print(void) {
    if (!getenv("SVTF_ADAPTER_STATS"))
        return;
    unsigned long launches = 0;
    for (int i = 0; i < SVTF_DEC_MAX_CACHED; i++)
        if (g_slots[i].ctx)
            launches += (unsigned long)svtf_launch_count(g_slots[i].ctx);
    fprintf(stderr, "svtfilt_b200 adapter stats:");
    for (int i = 0; i < DS_N; i++) fprintf(stderr, " %s=%lu", g_stat_name[i], g_stat[i]);
    fprintf(stderr, " gpu_launches=%lu\n", launches);
}
static void stat_init(void) { atexit(stat_print); }
static void stat_hit(int which) {
    pthread_once(&g_stat_once, stat_init);
    __sync_fetch_and_add(&g_stat[which], 1);
}

/* after_superres: the picture as the stages behind the super-resolution upscale see it (loop restoration, EbDecParseObu.c:2434-2455) */
static svtf_geom geom_of(const EbDecHandle *dec, const EbPictureBufferDesc *pic, int after_superres) {
    const FrameHeader   *fh = &dec->frame_header;
    const EbColorConfig *cc = &dec->seq_header.color_config;
    svtf_geom            g;
    /* the device path is 4:2:0 with three planes (what the reference ENCODER writes, Encoder/Codec/EbEncCdef.c:324-329); the other
     * formats the reference decoder reads (4:0:0, 4:2:2, 4:4:4: Decoder/Codec/EbDecCdef.c:172-173) have no kernels and no fallback */
    if (cc->mono_chrome || cc->subsampling_x != 1 || cc->subsampling_y != 1) {
        fprintf(stderr, "Svt[error]: svtfilt_b200 decoder adapter: only 4:2:0 streams are supported (mono_chrome %d, subsampling %d/%d); "
                        "no CPU fallback on this path\n", (int)cc->mono_chrome, (int)cc->subsampling_x, (int)cc->subsampling_y);
        abort();
    }
    g.width        = fh->frame_size.frame_width; /* the deblock limit, EbDecLF.c:119-124 */
    g.height       = fh->frame_size.frame_height;
    g.mi_cols      = fh->mi_cols;
    g.mi_rows      = fh->mi_rows;
    if (after_superres && fh->frame_size.superres_upscaled_width > fh->frame_size.frame_width) {
        g.width   = fh->frame_size.superres_upscaled_width;
        g.mi_cols = 2 * ((g.width + 7) >> 3);
    }
    g.bit_depth    = pic->bit_depth;
    g.sample_bytes = (pic->bit_depth > EB_EIGHT_BIT || dec->is_16bit_pipeline) ? 2 : 1;
    return g;
}

static DecSlot *slot_for_at(const EbDecHandle *dec, const EbPictureBufferDesc *pic, int after_superres) {
    pthread_mutex_lock(&g_lock);
    DecSlot        *free_slot = NULL, *s = NULL;
    const svtf_geom g = geom_of(dec, pic, 1), now = geom_of(dec, pic, after_superres);
    for (int i = 0; i < SVTF_DEC_MAX_CACHED; i++) {
        if (g_slots[i].dec == dec) {
            s = &g_slots[i];
            break;
        }
        if (!g_slots[i].dec && !free_slot)
            free_slot = &g_slots[i];
    }
    if (s && memcmp(&s->max_geom, &g, sizeof(g))) { /* a new sequence / frame size on the same handle */
        svtf_ctx_destroy(s->ctx);
        for (int p = 0; p < 3; p++) free(s->dbk[p]);
        memset(s, 0, sizeof(*s));
        free_slot = s, s = NULL;
    }
    if (!s) {
        if (!free_slot) {
            fprintf(stderr, "Svt[error]: svtfilt_b200 decoder adapter: more than %d decoder handles\n", SVTF_DEC_MAX_CACHED);
            abort();
        }
        s           = free_slot;
        s->dec      = dec;
        s->max_geom = g;
        if (svtf_ctx_create(g_device, &s->max_geom, &s->ctx) != SVTF_OK)
            die("svtf_ctx_create", NULL);
    }
    if (memcmp(&s->geom, &now, sizeof(now))) /* another picture size: nothing on the device is this picture */
        s->resident = 0;
    s->geom = now;
    pthread_mutex_unlock(&g_lock);
    return s;
}
static DecSlot *slot_for(const EbDecHandle *dec, const EbPictureBufferDesc *pic) { return slot_for_at(dec, pic, 0); }

void svtf_dec_adapter_release(const struct EbDecHandle *dec) {
    pthread_mutex_lock(&g_lock);
    for (int i = 0; i < SVTF_DEC_MAX_CACHED; i++)
        if (g_slots[i].dec == dec) {
            svtf_ctx_destroy(g_slots[i].ctx);
            for (int p = 0; p < 3; p++) free(g_slots[i].dbk[p]);
            memset(&g_slots[i], 0, sizeof(g_slots[i]));
        }
    pthread_mutex_unlock(&g_lock);
}

/* the block covering a 4x4 (what svt_aom_get_cur_mode_info resolves, Decoder/Codec/EbDecNbr.c:39-48) */
static const BlockModeInfo *mode_at(const EbDecHandle *dec, int mi_row, int mi_col) {
    const FrameMiMap *m  = &dec->main_frame_buf.frame_mi_map;
    const int         sh = m->sb_size_log2 - MI_SIZE_LOG2;
    const SBInfo     *sb = m->pps_sb_info[(mi_row >> sh) * m->sb_cols + (mi_col >> sh)];
    return &sb->sb_mode_info[m->p_mi_offset[mi_row * m->mi_cols_algnsb + mi_col]];
}

/* SB/block/TU side info -> packed words.  lf_ctxt == NULL (CDEF-only call): transform sizes are not needed. */
static void send_mi(DecSlot *s, const EbDecHandle *dec, const LfCtxt *lf_ctxt) {
    const int    rows = s->geom.mi_rows, cols = s->geom.mi_cols, st = dec->frame_header.mi_stride;
    const size_t n    = (size_t)rows * cols;
    uint8_t     *soa  = (uint8_t *)calloc(7, n);
    if (!soa)
        die("malloc(mi)", s->ctx);
    uint8_t *bsize = soa, *txy = soa + n, *txuv = soa + 2 * n, *skip = soa + 3 * n, *ref0 = soa + 4 * n, *mode = soa + 5 * n, *seg = soa + 6 * n;
    for (int r = 0; r < rows; r++)
        for (int c = 0; c < cols; c++) {
            const BlockModeInfo *b = mode_at(dec, r, c);
            const size_t         i = (size_t)r * cols + c;
            bsize[i] = (uint8_t)b->bsize, skip[i] = (uint8_t)b->skip, ref0[i] = (uint8_t)b->ref_frame[0];
            mode[i] = (uint8_t)b->mode, seg[i] = (uint8_t)b->segment_id;
            if (lf_ctxt)
                txy[i] = (uint8_t)lf_ctxt->tx_size_l[r * st + c], txuv[i] = (uint8_t)lf_ctxt->tx_size_uv[r * st + c];
        }
    CHECK(svtf_set_mi_soa_dec(s->ctx, bsize, txy, txuv, skip, ref0, mode, seg, cols), s->ctx); /* packed on the device */
    CHECK(svtf_sync(s->ctx), s->ctx);
    s->have_tx = lf_ctxt != NULL;
    free(soa);
}

static void plane_ptrs(const EbPictureBufferDesc *pic, int sb, void *p[3], int32_t st[3]) {
    /* visible origin as svt_aom_derive_blk_pointers computes it (Decoder/Codec/EbDecUtils.c:42-78), 4:2:0 */
    p[0]  = pic->buffer_y + ((size_t)(pic->org_x + pic->org_y * pic->stride_y) * sb);
    p[1]  = pic->buffer_cb + ((size_t)((pic->org_x >> 1) + (pic->org_y >> 1) * pic->stride_cb) * sb);
    p[2]  = pic->buffer_cr + ((size_t)((pic->org_x >> 1) + (pic->org_y >> 1) * pic->stride_cr) * sb);
    st[0] = pic->stride_y, st[1] = pic->stride_cb, st[2] = pic->stride_cr;
}

static void begin_frame(DecSlot *s, const EbDecHandle *dec, const EbPictureBufferDesc *pic, const LfCtxt *lf_ctxt) {
    if (g_resident && s->resident) {
        if (lf_ctxt && !s->have_tx)
            send_mi(s, dec, lf_ctxt);
        return;
    }
    void   *p[3];
    int32_t st[3];
    plane_ptrs(pic, s->geom.sample_bytes, p, st);
    CHECK(svtf_frame_begin(s->ctx, &s->geom), s->ctx);
    CHECK(svtf_upload_frame(s->ctx, p[0], st[0], p[1], st[1], p[2], st[2]), s->ctx);
    send_mi(s, dec, lf_ctxt);
    s->resident = 1;
}

static void end_frame(DecSlot *s, const EbPictureBufferDesc *pic) {
    void   *p[3];
    int32_t st[3];
    plane_ptrs(pic, s->geom.sample_bytes, p, st);
    CHECK(svtf_download_frame(s->ctx, p[0], st[0], p[1], st[1], p[2], st[2]), s->ctx);
}

/* Copy the resident frame back into `pic` (the end of a resident chain). */
void svtf_dec_adapter_fetch(const struct EbDecHandle *dec, void *pic) {
    DecSlot *s = NULL; /* whatever is resident: after a super-resolution upscale that is the wide picture */
    pthread_mutex_lock(&g_lock);
    for (int i = 0; i < SVTF_DEC_MAX_CACHED; i++)
        if (g_slots[i].dec == dec)
            s = &g_slots[i];
    pthread_mutex_unlock(&g_lock);
    if (!s)
        s = slot_for(dec, (const EbPictureBufferDesc *)pic);
    if (!s->resident) {
        fprintf(stderr, "Svt[error]: svtfilt_b200 decoder adapter: svtf_dec_adapter_fetch: no resident frame for this handle\n");
        abort();
    }
    end_frame(s, (const EbPictureBufferDesc *)pic);
    s->resident = 0;
}

static void lf_of(const EbDecHandle *dec, svtf_lf_params *out) {
    const FrameHeader       *fh = &dec->frame_header;
    const struct LoopFilter *l  = &fh->loop_filter_params;
    svtf_lf_params           lf;
    memset(&lf, 0, sizeof(lf));
    lf.filter_level[0] = l->filter_level[0], lf.filter_level[1] = l->filter_level[1];
    lf.filter_level_u = l->filter_level_u, lf.filter_level_v = l->filter_level_v;
    lf.sharpness_level        = l->sharpness_level;
    lf.mode_ref_delta_enabled = l->mode_ref_delta_enabled;
    memcpy(lf.ref_deltas, l->ref_deltas, sizeof(lf.ref_deltas));
    memcpy(lf.mode_deltas, l->mode_deltas, sizeof(lf.mode_deltas));
    lf.segmentation_enabled = fh->segmentation_params.segmentation_enabled;
    for (int i = 0; i < 8; i++)
        for (int f = 0; f < 8; f++) {
            lf.seg_feature_enabled[i][f] = fh->segmentation_params.feature_enabled[i][f];
            lf.seg_feature_data[i][f]    = fh->segmentation_params.feature_data[i][f];
        }
    lf.delta_lf_present = fh->delta_lf_params.delta_lf_present;
    lf.delta_lf_multi   = fh->delta_lf_params.delta_lf_multi;
    *out                = lf;
}

/* ---------------------------------------------------------------------------------------------------- */
void svt_aom_dec_av1_loop_filter_frame(EbDecHandle *dec_handle_ptr, EbPictureBufferDesc *recon_picture_buf, LfCtxt *lf_ctxt,
                                       int32_t plane_start, int32_t plane_end, int32_t is_mt, int enable_flag) {
    (void)is_mt;
    if (!enable_flag)
        return;
    stat_hit(DS_LF);
    DecSlot       *s = slot_for(dec_handle_ptr, recon_picture_buf);
    svtf_lf_params lf;
    lf_of(dec_handle_ptr, &lf);
    begin_frame(s, dec_handle_ptr, recon_picture_buf, lf_ctxt);
    /* the reference filters planes 0 .. plane_end - plane_start - 1 (dec_loop_filter_sb :612, plane loop :296): the
     * decoder only ever calls it with plane_start = 0 (EbDecParseObu.c:2385-2396) */
    const int32_t num_planes = plane_end - plane_start;
    if (lf.delta_lf_present) { /* per-superblock deltas, SBInfo.sb_delta_lf */
        const MainFrameBuf *mf  = &dec_handle_ptr->main_frame_buf;
        const int           nsb = mf->sb_cols * mf->sb_rows;
        int32_t            *d   = (int32_t *)malloc(sizeof(int32_t) * 4 * (size_t)nsb);
        if (!d)
            die("malloc(sb_delta_lf)", s->ctx);
        for (int sb = 0; sb < nsb; sb++)
            for (int i = 0; i < FRAME_LF_COUNT; i++) d[sb * 4 + i] = mf->cur_frame_bufs[0].sb_info[sb].sb_delta_lf[i];
        CHECK(svtf_deblock_frame_sb_delta(s->ctx, &lf, 0, num_planes, d, mf->sb_cols, dec_handle_ptr->seq_header.sb_size_log2), s->ctx);
        free(d);
    } else
        CHECK(svtf_deblock_frame(s->ctx, &lf, 0, num_planes), s->ctx);
    if (!g_resident)
        end_frame(s, recon_picture_buf), s->resident = 0;
}

/* ---------------------------------------------------------------------------------------------------- */
void svt_cdef_frame(EbDecHandle *dec_handle, int enable_flag) {
    if (!enable_flag)
        return;
    stat_hit(DS_CDEF);
    EbPictureBufferDesc *recon = dec_handle->cur_pic_buf[0]->ps_pic_buf;
    DecSlot             *s     = slot_for(dec_handle, recon);
    const FrameHeader   *fh    = &dec_handle->frame_header;
    begin_frame(s, dec_handle, recon, NULL);
    svtf_cdef_params p;
    memset(&p, 0, sizeof(p));
    p.cdef_damping = fh->cdef_params.cdef_damping;
    p.cdef_bits    = fh->cdef_params.cdef_bits;
    for (int i = 0; i < CDEF_MAX_STRENGTHS && i < SVTF_CDEF_MAX_STRENGTHS; i++) {
        p.cdef_y_strength[i]  = (uint8_t)fh->cdef_params.cdef_y_strength[i];
        p.cdef_uv_strength[i] = (uint8_t)fh->cdef_params.cdef_uv_strength[i];
    }
    /* per-64x64 strength index: SBInfo.sb_cdef_strength[quadrant] (svt_cdef_block, EbDecCdef.c:94-116); -1 = not coded */
    const int           nvfb = (s->geom.mi_rows + 15) / 16, nhfb = (s->geom.mi_cols + 15) / 16;
    const MainFrameBuf *mf   = &dec_handle->main_frame_buf;
    const int           b128 = dec_handle->seq_header.sb_size == BLOCK_128X128;
    int8_t             *idx  = (int8_t *)malloc((size_t)nvfb * nhfb);
    if (!idx)
        die("malloc(fb_idx)", s->ctx);
    for (int fbr = 0; fbr < nvfb; fbr++)
        for (int fbc = 0; fbc < nhfb; fbc++) {
            const SBInfo *sb = b128 ? mf->cur_frame_bufs[0].sb_info + (fbr >> 1) * mf->sb_cols + (fbc >> 1)
                                    : mf->cur_frame_bufs[0].sb_info + fbr * mf->sb_cols + fbc;
            idx[fbr * nhfb + fbc] = sb->sb_cdef_strength[b128 ? (fbc & 1) + 2 * (fbr & 1) : 0];
        }
    CHECK(svtf_cdef_apply_frame(s->ctx, &p, idx), s->ctx);
    free(idx);
    if (!g_resident)
        end_frame(s, recon), s->resident = 0;
}

/* ----------------------------------------------------------------------------------------------------
 * dec_ccso_frame (Decoder/Codec/EbDecCcso.c:384).  The decoder's CCSO is reachable only when a stream signals it (the
 * encoder's call sites are commented out) and its apply loop differs from the common definition (Common/Codec/EbCcso.c:
 * 297-622) in ways that look unintended:
 *   (1) band index: shift_bits = bit_depth in every variant (:51,136,222,310), i.e. band 0 regardless of max_band_log2;
 *   (2) block flags: the luma variants read sb_ccso_blk_v (:66,152) — `plane == 1 ? _u : _v`;
 *   (3) block flags are looked up at frame_buf->sb_info + <mi-grid offset of the block> (:63-65), which is inside the SBInfo
 *       array only for the block at (0,0): defined behaviour exists for pictures of one filter block per plane only
 *       (<= 256x256 luma);
 *   (4) samples are addressed from buffer_y/cb/cr without the picture origin (:396-404), i.e. the padding is filtered
 *       when org_x/org_y are not 0; 8-bit pictures only.
 * With SVTF_DEC_CCSO_AS_REFERENCE (default) this adapter reproduces (1) and (2) so that its output equals the reference's
 * wherever the reference is defined (pinned on single-block pictures with a zero origin, tests/test_gpu_dec_adapters.py);
 * for (3) it takes block (by, bx)'s flag from the superblock that contains the block's top-left sample — where the parser
 * stores it (EbDecParseBlock.c:357-377) — and for (4) it honours the origin.  Built with SVTF_DEC_CCSO_AS_REFERENCE=0 it
 * applies the common definition.
 */
void dec_ccso_frame(EbPictureBufferDesc *frame, EbDecHandle *dec_handle, uint16_t *ext_rec_y) {
    stat_hit(DS_CCSO);
    DecSlot           *s  = slot_for(dec_handle, frame);
    const FrameHeader *fh = &dec_handle->frame_header;
    const CcsoInfo    *ci = &fh->ccso_info;
    if (s->geom.sample_bytes != 1) {
        fprintf(stderr, "Svt[error]: svtfilt_b200 decoder adapter: dec_ccso_frame is defined for 8-bit pictures only (EbDecCcso.c:396)\n");
        abort();
    }
    svtf_ccso_params p;
    memset(&p, 0, sizeof(p));
    uint8_t            *flags[3];
    const MainFrameBuf *mf    = &dec_handle->main_frame_buf;
    const int           sb_l2 = dec_handle->seq_header.sb_size_log2;
    const int           W = frame->width, H = frame->height;
    for (int pl = 0; pl < 3; pl++) {
        p.ccso_enable[pl] = ci->ccso_enable[pl], p.ccso_bo_only[pl] = ci->ccso_bo_only[pl], p.quant_idx[pl] = ci->quant_idx[pl];
        p.ext_filter_support[pl] = ci->ext_filter_support[pl], p.edge_clf[pl] = ci->edge_clf[pl];
        p.max_band_log2[pl] = SVTF_DEC_CCSO_AS_REFERENCE ? 0 : (uint8_t)ci->max_band_log2[pl];
        memcpy(p.filter_offset[pl], ci->filter_offset[pl], SVTF_CCSO_LUT_SIZE);
        const int pw = W >> (pl ? 1 : 0), ph = H >> (pl ? 1 : 0), bl = pl ? 7 : 8;
        const int nbx = (pw + (1 << bl) - 1) >> bl, nby = (ph + (1 << bl) - 1) >> bl;
        flags[pl] = (uint8_t *)malloc((size_t)nbx * nby);
        for (int by = 0; by < nby; by++)
            for (int bx = 0; bx < nbx; bx++) {
                const SBInfo *sb = mf->cur_frame_bufs[0].sb_info + ((by << 8) >> sb_l2) * mf->sb_cols + ((bx << 8) >> sb_l2);
                const int     fp = SVTF_DEC_CCSO_AS_REFERENCE ? (pl == 1 ? 1 : 2) : pl;
                flags[pl][by * nbx + bx] = fp == 0 ? sb->sb_ccso_blk_y[0] : (fp == 1 ? sb->sb_ccso_blk_u[0] : sb->sb_ccso_blk_v[0]);
            }
    }
    if (!(g_resident && s->resident)) {
        /* stand-alone call: the classifier input is the caller's ext_rec_y (luma before CDEF, padded by CCSO_PADDING_SIZE,
         * EbDecParseObu.c:2399-2419): stage it as the pre-CDEF picture with an identity CDEF pass, then put the picture
         * to correct on top */
        const int cw = s->geom.mi_cols * 4, chh = s->geom.mi_rows * 4, es = W + 2 * CCSO_PADDING_SIZE;
        void     *pp[3];
        int32_t   st[3];
        plane_ptrs(frame, 1, pp, st);
        uint8_t *ly = (uint8_t *)malloc((size_t)cw * chh);
        for (int r = 0; r < chh; r++)
            for (int c = 0; c < cw; c++)
                ly[(size_t)r * cw + c] = (uint8_t)ext_rec_y[(size_t)(CCSO_PADDING_SIZE + (r < H ? r : H - 1)) * es + CCSO_PADDING_SIZE + (c < W ? c : W - 1)];
        CHECK(svtf_frame_begin(s->ctx, &s->geom), s->ctx);
        CHECK(svtf_upload_frame(s->ctx, ly, cw, pp[1], st[1], pp[2], st[2]), s->ctx);
        send_mi(s, dec_handle, NULL);
        svtf_cdef_params idp;
        memset(&idp, 0, sizeof(idp));
        const int nfb = ((s->geom.mi_rows + 15) / 16) * ((s->geom.mi_cols + 15) / 16);
        int8_t   *idx = (int8_t *)calloc(nfb, 1);
        CHECK(svtf_cdef_apply_frame(s->ctx, &idp, idx), s->ctx);
        CHECK(svtf_sync(s->ctx), s->ctx);
        free(idx), free(ly);
        CHECK(svtf_overwrite_current(s->ctx, pp[0], st[0], pp[1], st[1], pp[2], st[2]), s->ctx);
    }
    const uint8_t *cf[3] = {flags[0], flags[1], flags[2]};
    CHECK(svtf_ccso_apply_frame(s->ctx, &p, cf), s->ctx);
    if (g_resident)
        s->resident = 1; /* loop restoration or svtf_dec_adapter_fetch follows */
    else
        end_frame(s, frame), s->resident = 0;
    for (int pl = 0; pl < 3; pl++) free(flags[pl]);
}

/* ----------------------------------------------------------------------------------------------------
 * Loop restoration (Decoder/Codec/EbDecRestoration.c).  The decoder calls, single-threaded (EbDecParseObu.c:2421-2455):
 *   svt_aom_dec_av1_loop_restoration_save_boundary_lines(dec, 0)   after the loop filter: keep two deblocked rows above and
 *                                                                  below every 64-row stripe
 *   svt_cdef_frame / dec_ccso_frame / superres
 *   svt_aom_dec_av1_loop_restoration_save_boundary_lines(dec, 1)   the picture's own edge rows for the frame top / bottom
 *   svt_aom_dec_av1_loop_restoration_filter_frame(dec, 0, do_lr)   lr_pad_pic + SB rows x 64-column blocks with column
 *                                                                  save / restore around every block
 * On the device the deblocked picture is still resident when the frame went LF -> CDEF here (resident mode), so both
 * saves are no-ops and the filter is one launch (svtf_lr_apply_frame).  Stand-alone, save(0) keeps the boundary rows in
 * an adapter-owned sparse picture which the filter call uploads as the pre-CDEF picture — only those rows are read;
 * save(1) has nothing to keep (the device repeats the picture's edge rows itself, as lr_pad_pic does).
 * Restoration units: LrCtxt.lr_unit[plane][unit_row * lr_stride + unit_col] (:341; written by the parser,
 * EbDecParseBlock.c:2692) -> the dense [vert][horz] array of svtf_lr_apply_frame; the decoder's "use the previous unit when
 * less than half a unit remains" (:333-338) is the same count as svtf_lr_units.
 * Super-resolution: svt_av1_superres_upscale below; loop restoration then runs on the upscaled picture, its stripe borders are the
 * deblocked rows upscaled (lr_stripe_rows stand-alone; on the device in a resident chain).
 */
void svt_av1_upscale_normative_rows(const Av1Common *cm, const uint8_t *src, int src_stride, uint8_t *dst, int dst_stride, int rows,
                                    int sub_x, int bd, Bool is_16bit_pipeline); /* Common/Codec/EbSuperRes.c:214 */

/* Stand-alone loop restoration: the two deblocked rows above and below every 64-row stripe, kept in an adapter-owned sparse picture
 * of the size loop restoration works on.  With super-resolution that is the UPSCALED size and the rows are upscaled on the way in,
 * as svt_aom_save_deblock_boundary_lines does (Common/Codec/EbRestoration.c:1547-1560) — with the reference's own row function: four
 * rows per stripe are not worth a launch. */
static void lr_stripe_rows(DecSlot *s, const EbDecHandle *dec, const EbPictureBufferDesc *pic) {
    const int sb = s->geom.sample_bytes;
    const int sr = s->max_geom.width > s->geom.width; /* s->geom: the coded picture (this runs after the loop filter) */
    void     *pp[3];
    int32_t   st[3];
    plane_ptrs(pic, sb, pp, st);
    for (int p = 0; p < 3; p++) {
        const int ss = p ? 1 : 0, cw = (s->max_geom.mi_cols * 4) >> ss, chh = (s->max_geom.mi_rows * 4) >> ss;
        const int pw = (s->geom.width + ss) >> ss, ph = (s->geom.height + ss) >> ss, sh = 64 >> ss, off = 8 >> ss;
        if (!s->dbk[p] && !(s->dbk[p] = (uint8_t *)calloc((size_t)cw * chh, sb)))
            die("malloc(deblocked rows)", s->ctx);
        for (int k = 0;; k++) {
            const int y0 = k * sh - off > 0 ? k * sh - off : 0, y1 = (k + 1) * sh - off < ph ? (k + 1) * sh - off : ph;
            if (y0 >= ph)
                break;
            for (int i = 0; i < RESTORATION_CTX_VERT; i++) {
                int rows[2], n = 0;
                if (k > 0)
                    rows[n++] = y0 - 2 + i;
                if (y1 + i < ph)
                    rows[n++] = y1 + i;
                for (int j = 0; j < n; j++) {
                    uint8_t       *d   = s->dbk[p] + (size_t)rows[j] * cw * sb;
                    const uint8_t *src = (const uint8_t *)pp[p] + (size_t)rows[j] * st[p] * sb;
                    if (sr)
                        svt_av1_upscale_normative_rows(&dec->cm, src, st[p], d, cw, 1, ss, s->geom.bit_depth, sb == 2);
                    else
                        memcpy(d, src, (size_t)pw * sb);
                }
            }
        }
    }
    s->have_dbk = 1;
}

void svt_aom_dec_av1_loop_restoration_save_boundary_lines(EbDecHandle *dec_handle, int after_cdef) {
    EbPictureBufferDesc *pic = dec_handle->cur_pic_buf[0]->ps_pic_buf;
    if (after_cdef)
        return;
    DecSlot *s = slot_for(dec_handle, pic);
    if (g_resident && s->resident)
        return;
    lr_stripe_rows(s, dec_handle, pic);
}

void svt_aom_dec_av1_loop_restoration_filter_frame(EbDecHandle *dec_handle, int optimized_lr, int enable_flag) {
    if (!enable_flag)
        return;
    stat_hit(DS_LR);
    EbPictureBufferDesc *pic = dec_handle->cur_pic_buf[0]->ps_pic_buf;
    DecSlot             *s   = slot_for_at(dec_handle, pic, 1); /* the upscaled picture when the frame uses super-resolution */
    const LrCtxt        *lc  = (const LrCtxt *)dec_handle->pv_lr_ctxt;
    const FrameHeader   *fh  = &dec_handle->frame_header;
    svtf_lr_params lr;
    svtf_lr_unit  *units[3] = {NULL, NULL, NULL};
    memset(&lr, 0, sizeof(lr));
    lr.optimized_lr = optimized_lr;
    for (int p = 0; p < 3; p++) {
        const LrParams *fp = &fh->lr_params[p];
        lr.frame_restoration_type[p] = fp->frame_restoration_type == RESTORE_NONE ? SVTF_RESTORE_NONE : (int32_t)fp->frame_restoration_type;
        lr.restoration_unit_size[p]  = fp->loop_restoration_size;
        if (fp->frame_restoration_type == RESTORE_NONE)
            continue;
        int32_t hu, vu;
        CHECK(svtf_lr_units(&s->geom, p, fp->loop_restoration_size, &hu, &vu), s->ctx);
        if (!(units[p] = (svtf_lr_unit *)calloc((size_t)hu * vu, sizeof(svtf_lr_unit))))
            die("malloc(units)", s->ctx);
        for (int v = 0; v < vu; v++)
            for (int u = 0; u < hu; u++) {
                const RestorationUnitInfo *ri = lc->lr_unit[p] + v * lc->lr_stride[p] + u;
                svtf_lr_unit              *o  = &units[p][v * hu + u];
                o->restoration_type = ri->restoration_type == RESTORE_WIENER ? SVTF_RESTORE_WIENER
                    : ri->restoration_type == RESTORE_SGRPROJ                 ? SVTF_RESTORE_SGRPROJ
                                                                              : SVTF_RESTORE_NONE;
                memcpy(o->wiener_vfilter, ri->wiener_info.vfilter, sizeof(o->wiener_vfilter));
                memcpy(o->wiener_hfilter, ri->wiener_info.hfilter, sizeof(o->wiener_hfilter));
                o->sgr_ep = ri->sgrproj_info.ep, o->sgr_xqd[0] = ri->sgrproj_info.xqd[0], o->sgr_xqd[1] = ri->sgrproj_info.xqd[1];
            }
    }
    if (!(g_resident && s->resident)) {
        void   *pp[3];
        int32_t st[3];
        plane_ptrs(pic, s->geom.sample_bytes, pp, st);
        CHECK(svtf_frame_begin(s->ctx, &s->geom), s->ctx);
        CHECK(svtf_upload_frame(s->ctx, pp[0], st[0], pp[1], st[1], pp[2], st[2]), s->ctx);
        if (!optimized_lr) {
            if (!s->have_dbk) {
                fprintf(stderr, "Svt[error]: svtfilt_b200 decoder adapter: loop restoration without the deblocked boundary rows "
                                "(svt_aom_dec_av1_loop_restoration_save_boundary_lines(dec, 0) was not called for this frame)\n");
                abort();
            }
            const int cw = s->max_geom.mi_cols * 4;
            CHECK(svtf_upload_pre_cdef(s->ctx, s->dbk[0], cw, s->dbk[1], cw >> 1, s->dbk[2], cw >> 1), s->ctx);
            CHECK(svtf_sync(s->ctx), s->ctx);
        }
        s->resident = 1;
    }
    s->have_dbk = 0;
    const svtf_lr_unit *cu[3] = {units[0], units[1], units[2]};
    CHECK(svtf_lr_apply_frame(s->ctx, &lr, cu), s->ctx);
    if (!g_resident)
        end_frame(s, pic), s->resident = 0;
    for (int p = 0; p < 3; p++) free(units[p]);
}

/* ----------------------------------------------------------------------------------------------------
 * svt_av1_superres_upscale (Decoder/Codec/EbDecSuperRes.c:119), called between CDEF / CCSO and loop restoration
 * (EbDecParseObu.c:2434-2438): the reference copies the picture aside, clears it and runs svt_av1_upscale_normative_rows per plane
 * into the same buffers, then calls the picture `superres_upscaled_width` wide.  Here: one launch per plane on the device
 * (svtf_superres_upscale_frame: the picture, and in a resident chain the deblocked picture loop restoration borrows its stripe
 * borders from); stand-alone the picture is uploaded narrow and written back wide.  The padding the reference zeroes is rewritten
 * by the padding that follows loop restoration (svt_aom_pad_pic / the `_mt` restoration stage).
 * `cm` is EbDecHandle.cm (the only caller passes &dec_handle_ptr->cm), which is how the handle is found.
 * ---------------------------------------------------------------------------------------------------- */
#include <stddef.h>
#include "EbInterPrediction.h" /* RS_SUBPEL_BITS for the table below */
#include "EbSuperRes.h"
void svt_av1_superres_upscale(Av1Common *cm, FrameHeader *frm_hdr, SeqHeader *seq_hdr, EbPictureBufferDesc *recon_picture_src, int enable_flag) {
    (void)seq_hdr;
    if (!enable_flag || av1_superres_unscaled(&frm_hdr->frame_size))
        return;
    stat_hit(DS_SUPERRES);
    const EbDecHandle *dec = (const EbDecHandle *)((const char *)cm - offsetof(EbDecHandle, cm));
    DecSlot           *s   = slot_for(dec, recon_picture_src);
    void              *p[3];
    int32_t            st[3];
    plane_ptrs(recon_picture_src, s->geom.sample_bytes, p, st);
    const int resident = g_resident && s->resident;
    if (!resident) {
        CHECK(svtf_frame_begin(s->ctx, &s->geom), s->ctx);
        CHECK(svtf_upload_frame(s->ctx, p[0], st[0], p[1], st[1], p[2], st[2]), s->ctx);
    }
    svtf_superres_params sr;
    memset(&sr, 0, sizeof(sr));
    sr.upscaled_width       = frm_hdr->frame_size.superres_upscaled_width;
    sr.superres_denominator = frm_hdr->frame_size.superres_denominator;
    sr.tile_cols            = cm->tiles_info.tile_cols;
    for (int j = 0; j <= sr.tile_cols && j < 65; j++) sr.tile_col_start_mi[j] = cm->tiles_info.tile_col_start_mi[j];
    sr.filters = &av1_resize_filter_normative[0][0];
    CHECK(svtf_superres_upscale_frame(s->ctx, &sr), s->ctx);
    s->geom = s->max_geom; /* the picture of this frame is the upscaled one from here on */
    recon_picture_src->width = (uint16_t)sr.upscaled_width;
    if (!resident) {
        CHECK(svtf_download_frame(s->ctx, p[0], st[0], p[1], st[1], p[2], st[2]), s->ctx);
        s->resident = 0;
    }
}

/* ----------------------------------------------------------------------------------------------------
 * Multi-threaded decoding (dec_config.threads > 1).  The reference gives every thread — the caller (thread_ctxt == NULL,
 * EbDecParseObu.c:2382-2453) and the workers (dec_all_stage_kernel, EbDecProcess.c:1167-1203) — the same three calls per
 * frame, and the bodies share superblock rows through job counters and spin-wait row maps:
 *   Decoder/Codec/EbDecProcess.c:668   svt_aom_dec_av1_loop_filter_frame_mt            (+ the LR boundary rows per SB row)
 *   Decoder/Codec/EbDecProcess.c:810   svt_cdef_frame_mt
 *   Decoder/Codec/EbDecProcess.c:1022  svt_aom_dec_av1_loop_restoration_filter_frame_mt (+ the picture padding per SB row)
 * On the device a stage is one launch over the whole picture, so here the CALLING thread runs the frame-level adapter above
 * once the stage's input is complete, and the workers pass straight through to the frame-end barrier in the third call
 * (num_threads_lred, :1150-1165), which is kept as the reference has it — it is what keeps a worker from entering the next
 * frame early and what clears the start_* flags.  The row maps these bodies used among themselves (lf_row_map,
 * cdef_completed_for_row_map, lr_row_map) have no other reader and are left untouched.
 */
#include <sched.h>
#include "EbDecProcess.h"

static int dec_do_lr(const EbDecHandle *dec) {
    const LrParams *lp = dec->frame_header.lr_params;
    return !dec->frame_header.allow_intrabc &&
        (lp[0].frame_restoration_type != RESTORE_NONE || lp[1].frame_restoration_type != RESTORE_NONE ||
         lp[2].frame_restoration_type != RESTORE_NONE);
}

void svt_aom_dec_av1_loop_filter_frame_mt(EbDecHandle *dec_handle, EbPictureBufferDesc *recon_picture_buf, LfCtxt *lf_ctxt,
                                          int32_t plane_start, int32_t plane_end, DecThreadCtxt *thread_ctxt) {
    DecMtFrameData *mt = &dec_handle->main_frame_buf.cur_frame_bufs[0].dec_mt_frame_data;
    volatile Bool  *start = &mt->start_lf_frame;
    while (*start != TRUE) svt_block_on_semaphore(thread_ctxt ? thread_ctxt->thread_semaphore : dec_handle->thread_semaphore);
    if (thread_ctxt)
        return;
    /* every tile's every SB row reconstructed (the per-row wait of :741-763, over the whole picture) */
    const int                n   = mt->sb_rows * dec_handle->frame_header.tiles_info.tile_cols;
    volatile const uint32_t *map = (volatile const uint32_t *)mt->sb_recon_row_map;
    for (int i = 0; i < n; i++)
        while (!map[i] && !mt->end_flag) sched_yield();
    const FrameHeader *fh = &dec_handle->frame_header;
    lf_ctxt->delta_lf_stride = dec_handle->main_frame_buf.sb_cols * FRAME_LF_COUNT;
    svt_aom_dec_av1_loop_filter_frame(dec_handle, recon_picture_buf, lf_ctxt, plane_start, plane_end, /*is_mt*/ 1,
                                      !fh->allow_intrabc && (fh->loop_filter_params.filter_level[0] || fh->loop_filter_params.filter_level[1]));
    /* dec_save_lf_boundary_lines_sb_row (:600-629) per SB row == the frame-level save after the loop filter */
    if (dec_do_lr(dec_handle))
        svt_aom_dec_av1_loop_restoration_save_boundary_lines(dec_handle, 0);
}

void svt_cdef_frame_mt(EbDecHandle *dec_handle_ptr, DecThreadCtxt *thread_ctxt) {
    DecMtFrameData *mt = &dec_handle_ptr->main_frame_buf.cur_frame_bufs[0].dec_mt_frame_data;
    volatile Bool  *start = &mt->start_cdef_frame;
    while (*start != TRUE) svt_block_on_semaphore(thread_ctxt ? thread_ctxt->thread_semaphore : dec_handle_ptr->thread_semaphore);
    const FrameHeader *fh = &dec_handle_ptr->frame_header;
    if (!thread_ctxt)
        svt_cdef_frame(dec_handle_ptr,
                       !fh->allow_intrabc && !fh->coded_lossless &&
                           (fh->cdef_params.cdef_bits || fh->cdef_params.cdef_y_strength[0] || fh->cdef_params.cdef_uv_strength[0]));
    /* :938-945 — with superres every thread waits here until CDEF is over, because the caller upscales next */
    svt_block_on_mutex(mt->temp_mutex);
    mt->num_threads_cdefed++;
    svt_release_mutex(mt->temp_mutex);
    if (!fh->allow_intrabc && !av1_superres_unscaled(&fh->frame_size)) {
        volatile uint32_t *done = &mt->num_threads_cdefed;
        while (*done != dec_handle_ptr->dec_config.threads && !mt->end_flag) sched_yield();
    }
}

void svt_aom_dec_av1_loop_restoration_filter_frame_mt(EbDecHandle *dec_handle, DecThreadCtxt *thread_ctxt) {
    DecMtFrameData *mt = &dec_handle->main_frame_buf.cur_frame_bufs[0].dec_mt_frame_data;
    volatile Bool  *start = &mt->start_lr_frame;
    while (*start != TRUE) svt_block_on_semaphore(thread_ctxt ? thread_ctxt->thread_semaphore : dec_handle->thread_semaphore);
    if (!thread_ctxt && !mt->end_flag) {
        svt_aom_dec_av1_loop_restoration_filter_frame(dec_handle, 0, dec_do_lr(dec_handle));
        svt_aom_pad_pic(dec_handle); /* pad_pre_lr + pad_post_lr (:968-1020) over all rows == the single-threaded padding */
    }
    svt_block_on_mutex(mt->temp_mutex);
    mt->num_threads_lred++;
    if (dec_handle->dec_config.threads == mt->num_threads_lred) {
        mt->start_motion_proj = FALSE, mt->start_parse_frame = FALSE, mt->start_decode_frame = FALSE;
        mt->start_lf_frame = FALSE, mt->start_cdef_frame = FALSE, mt->start_lr_frame = FALSE;
    }
    svt_release_mutex(mt->temp_mutex);
    volatile uint32_t *done = &mt->num_threads_lred;
    while (*done != dec_handle->dec_config.threads && FALSE == mt->end_flag) sched_yield();
}
