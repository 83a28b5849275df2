/*
 * adapter_harness_dec.c — drives the Tier-2 DECODER adapters (svtf_dec_adapters.c) on the same synthetic EbDecHandle the
 * real decoder functions are driven with in oracle/ref_shim_dec.c (oracle/dec_harness.h); same exported signatures with
 * the prefix adaptdec_.  TEST INFRASTRUCTURE ONLY.
 */
#include <stdint.h>
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
#define DEC_HARNESS_LR
#include "../oracle/dec_harness.h"

#define API __attribute__((visibility("default")))

API int adaptdec_lf_frame(const svtf_geom *g, void *y, int sy, void *cb, int scb, void *cr, int scr, const dec_mi_soa *soa,
                          const svtf_lf_params *lf, int plane_start, int plane_end) {
    dec_harness h;
    if (dec_harness_init(&h, g, soa, y, sy, cb, scb, cr, scr))
        return -1;
    dec_harness_fill_lf(&h, lf);
    svt_aom_dec_av1_loop_filter_frame(h.dec, h.recon, &h.lf_ctxt, plane_start, plane_end, 0, 1);
    svtf_dec_adapter_release(h.dec);
    dec_harness_free(&h);
    return 0;
}

API int adaptdec_cdef_frame(const svtf_geom *g, void *y, int sy, void *cb, int scb, void *cr, int scr, const dec_mi_soa *soa,
                            const svtf_cdef_params *cdef, const int8_t *fb_idx) {
    dec_harness h;
    if (dec_harness_init(&h, g, soa, y, sy, cb, scb, cr, scr))
        return -1;
    dec_harness_fill_cdef(&h, g, soa, cdef, fb_idx);
    svt_cdef_frame(h.dec, 1);
    svtf_dec_adapter_release(h.dec);
    dec_harness_free(&h);
    return 0;
}

API int adaptdec_ccso_frame(const svtf_geom *g, void *y, int sy, void *cb, int scb, void *cr, int scr, const uint8_t *pre_y, int pre_sy,
                            const dec_mi_soa *soa, const svtf_ccso_params *p, const uint8_t blk[3]) {
    dec_harness h;
    if (g->sample_bytes != 1 || dec_harness_init(&h, g, soa, y, sy, cb, scb, cr, scr))
        return -1;
    dec_harness_fill_ccso(&h, p, blk);
    const int es  = g->width + 2 * CCSO_PADDING_SIZE;
    uint16_t *ext = calloc((size_t)es * (g->height + 2 * CCSO_PADDING_SIZE), sizeof(uint16_t));
    for (int r = 0; r < g->height; r++)
        for (int c = 0; c < g->width; c++) ext[(r + CCSO_PADDING_SIZE) * es + c + CCSO_PADDING_SIZE] = pre_y[r * pre_sy + c];
    /* (the border of ext_rec_y is not read by the adapter: the device clamps coordinates instead) */
    dec_ccso_frame(h.recon, h.dec, ext);
    free(ext);
    svtf_dec_adapter_release(h.dec);
    dec_harness_free(&h);
    return 0;
}

/* The decoder's frame sequence with the frame kept on the device between the stages: LF -> CDEF, then fetch
 * (Decoder/Codec/EbDecParseObu.c:2382-2428 without CCSO / LR). */
API int adaptdec_lf_cdef_resident(const svtf_geom *g, void *y, int sy, void *cb, int scb, void *cr, int scr, const dec_mi_soa *soa,
                                  const svtf_lf_params *lf, const svtf_cdef_params *cdef, const int8_t *fb_idx) {
    dec_harness h;
    if (dec_harness_init(&h, g, soa, y, sy, cb, scb, cr, scr))
        return -1;
    dec_harness_fill_lf(&h, lf);
    dec_harness_fill_cdef(&h, g, soa, cdef, fb_idx);
    svtf_dec_adapter_set_resident(1);
    svt_aom_dec_av1_loop_filter_frame(h.dec, h.recon, &h.lf_ctxt, 0, 3, 0, 1);
    svt_cdef_frame(h.dec, 1);
    svtf_dec_adapter_fetch(h.dec, h.recon);
    svtf_dec_adapter_set_resident(0);
    svtf_dec_adapter_release(h.dec);
    dec_harness_free(&h);
    return 0;
}

/* Loop restoration, stand-alone calls in the decoder's order (EbDecParseObu.c:2421-2455): save(0) sees the deblocked
 * picture, save(1) and the filter the picture to restore. */
API int adaptdec_lr_frame(const svtf_geom *g, void *y, int sy, void *cb, int scb, void *cr, int scr, void *dy, int dsy, void *dcb, int dscb,
                          void *dcr, int dscr, const dec_mi_soa *soa, const svtf_lr_params *lr, const svtf_lr_unit *const units[3]) {
    dec_harness    h, hd;
    dec_harness_lr l, ld;
    if (dec_harness_init(&h, g, soa, y, sy, cb, scb, cr, scr) || dec_harness_fill_lr(&h, &l, g, lr, units))
        return -1;
    if (dec_harness_init(&hd, g, soa, dy, dsy, dcb, dscb, dcr, dscr) || dec_harness_fill_lr(&hd, &ld, g, lr, units))
        return -1;
    EbPictureBufferDesc *cur = h.pic_buf->ps_pic_buf;
    h.pic_buf->ps_pic_buf    = hd.recon;
    svt_aom_dec_av1_loop_restoration_save_boundary_lines(h.dec, 0);
    h.pic_buf->ps_pic_buf = cur;
    svt_aom_dec_av1_loop_restoration_save_boundary_lines(h.dec, 1);
    svt_aom_dec_av1_loop_restoration_filter_frame(h.dec, 0, 1);
    dec_harness_lr_store(&h, &l, g);
    svtf_dec_adapter_release(h.dec);
    dec_harness_lr_free(&l), dec_harness_lr_free(&ld);
    dec_harness_free(&h), dec_harness_free(&hd);
    return 0;
}

/* The decoder's whole post-recon sequence with the frame kept on the device: LF -> save(0) -> CDEF -> save(1) -> LR -> fetch */
API int adaptdec_chain_lr_resident(const svtf_geom *g, void *y, int sy, void *cb, int scb, void *cr, int scr, const dec_mi_soa *soa,
                                   const svtf_lf_params *lf, const svtf_cdef_params *cdef, const int8_t *fb_idx, const svtf_lr_params *lr,
                                   const svtf_lr_unit *const units[3]) {
    dec_harness    h;
    dec_harness_lr l;
    if (dec_harness_init(&h, g, soa, y, sy, cb, scb, cr, scr) || dec_harness_fill_lr(&h, &l, g, lr, units))
        return -1;
    dec_harness_fill_lf(&h, lf);
    dec_harness_fill_cdef(&h, g, soa, cdef, fb_idx);
    svtf_dec_adapter_set_resident(1);
    svt_aom_dec_av1_loop_filter_frame(h.dec, h.recon, &h.lf_ctxt, 0, 3, 0, 1);
    svt_aom_dec_av1_loop_restoration_save_boundary_lines(h.dec, 0);
    svt_cdef_frame(h.dec, 1);
    svt_aom_dec_av1_loop_restoration_save_boundary_lines(h.dec, 1);
    svt_aom_dec_av1_loop_restoration_filter_frame(h.dec, 0, 1);
    svtf_dec_adapter_fetch(h.dec, h.recon);
    svtf_dec_adapter_set_resident(0);
    dec_harness_lr_store(&h, &l, g);
    svtf_dec_adapter_release(h.dec);
    dec_harness_lr_free(&l);
    dec_harness_free(&h);
    return 0;
}

/* The multi-threaded stage entries of svtf_dec_adapters.c use the decoder library's thread primitives and its picture padding;
 * this harness drives single-threaded frame calls only (the `_mt` entries are exercised by the real decoder library in
 * tests/test_gpu_app_dropin.py), so here they only have to resolve. */
static void harness_unreachable(const char *what) {
    fprintf(stderr, "adapter_harness_dec: %s is not available in the test harness\n", what);
    abort();
}
EbErrorType svt_block_on_semaphore(EbHandle h) { (void)h; harness_unreachable("svt_block_on_semaphore"); return EB_ErrorNone; }
EbErrorType svt_block_on_mutex(EbHandle h) { (void)h; harness_unreachable("svt_block_on_mutex"); return EB_ErrorNone; }
EbErrorType svt_release_mutex(EbHandle h) { (void)h; harness_unreachable("svt_release_mutex"); return EB_ErrorNone; }
void        svt_aom_pad_pic(EbDecHandle *d) { (void)d; harness_unreachable("svt_aom_pad_pic"); }
void        svt_av1_upscale_normative_rows(const Av1Common *cm, const uint8_t *src, int src_stride, uint8_t *dst, int dst_stride, int rows, int sub_x,
                                           int bd, Bool is_16bit_pipeline) { /* super-resolution streams: application test only */
    (void)cm, (void)src, (void)src_stride, (void)dst, (void)dst_stride, (void)rows, (void)sub_x, (void)bd, (void)is_16bit_pipeline;
    harness_unreachable("svt_av1_upscale_normative_rows");
}

/* tests/test_dec_adapter_formats.py: the adapter's answer to a stream format it has no kernels for (expected: message + abort, before
 * any device call) */
void harness_dec_format_probe(int mono, int ssx, int ssy) {
    EbDecHandle         *d = calloc(1, sizeof(*d));
    EbPictureBufferDesc *r = calloc(1, sizeof(*r));
    LfCtxt               lf;
    memset(&lf, 0, sizeof(lf));
    d->seq_header.color_config.mono_chrome   = (Bool)mono;
    d->seq_header.color_config.subsampling_x = (uint8_t)ssx;
    d->seq_header.color_config.subsampling_y = (uint8_t)ssy;
    d->frame_header.frame_size.frame_width = d->frame_header.frame_size.superres_upscaled_width = 64;
    d->frame_header.frame_size.frame_height = d->frame_header.frame_size.superres_upscaled_height = 64;
    d->frame_header.mi_cols = d->frame_header.mi_rows = 16;
    r->bit_depth = EB_EIGHT_BIT;
    svt_aom_dec_av1_loop_filter_frame(d, r, &lf, 0, mono ? 1 : 3, 0, 1);
}
