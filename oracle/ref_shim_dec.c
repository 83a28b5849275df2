/*
 * ref_shim_dec.c — drives the reference DECODER's post-recon filters on synthetic frames.  TEST INFRASTRUCTURE ONLY.
 *
 *   svt_aom_dec_av1_loop_filter_frame  Decoder/Codec/EbDecLF.c:697
 *   svt_cdef_frame                     Decoder/Codec/EbDecCdef.c:507
 *   dec_ccso_frame                     Decoder/Codec/EbDecCcso.c:384  (+ the ext_rec_y preparation of
 *                                      Decoder/Codec/EbDecParseObu.c:2399-2419)
 *   svt_aom_dec_av1_loop_restoration_filter_frame / _save_boundary_lines   Decoder/Codec/EbDecRestoration.c:429 / :473
 * are compiled from the reference tree into oracle/_ref/libsvtrefdec.so (oracle/Makefile) and called on an EbDecHandle
 * built by dec_harness.h.  The decoder walks blocks -> transform units instead of the 4x4 grid, so these entry points
 * pin that the packed-mi formulation (svtf_pack_mi_dec) describes the same edges.
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

#define DEC_HARNESS_LR
#include "dec_harness.h"

#define REF_API __attribute__((visibility("default")))

REF_API int refdec_lf_frame(const svtf_geom *g, void *y, int sy, void *cb, int scb, void *cr, int scr, const dec_mi_soa *soa,
                            const svtf_lf_params *lf, int plane_start, int plane_end) {
    dec_harness h;
    if (dec_harness_init(&h, g, soa, y, sy, cb, scb, cr, scr))
        return -1;
    dec_harness_fill_lf(&h, lf);
    svt_aom_dec_av1_loop_filter_frame(h.dec, h.recon, &h.lf_ctxt, plane_start, plane_end, 0, 1);
    dec_harness_free(&h);
    return 0;
}

REF_API int refdec_cdef_frame(const svtf_geom *g, void *y, int sy, void *cb, int scb, void *cr, int scr, const dec_mi_soa *soa,
                              const svtf_cdef_params *cdef, const int8_t *fb_idx) {
    dec_harness h;
    if (dec_harness_init(&h, g, soa, y, sy, cb, scb, cr, scr))
        return -1;
    dec_harness_fill_cdef(&h, g, soa, cdef, fb_idx);
    svt_cdef_frame(h.dec, 1);
    dec_harness_free(&h);
    return 0;
}

/* 8-bit pictures only (dec_ccso_frame reads bytes); pre_y = luma before CDEF; blk[3] = the flags of SB 0 */
REF_API int refdec_ccso_frame(const svtf_geom *g, void *y, int sy, void *cb, int scb, void *cr, int scr, const uint8_t *pre_y, int pre_sy,
                              const dec_mi_soa *soa, const svtf_ccso_params *p, const uint8_t blk[3]) {
    dec_harness h;
    if (g->sample_bytes != 1 || dec_harness_init(&h, g, soa, y, sy, cb, scb, cr, scr))
        return -1;
    dec_harness_fill_ccso(&h, p, blk);
    const int es  = g->width + 2 * CCSO_PADDING_SIZE;
    uint16_t *ext = malloc(sizeof(uint16_t) * es * (g->height + 2 * CCSO_PADDING_SIZE));
    for (int r = 0; r < g->height; r++)
        for (int c = 0; c < g->width; c++) ext[(r + CCSO_PADDING_SIZE) * es + c + CCSO_PADDING_SIZE] = pre_y[r * pre_sy + c];
    dec_extend_ccso_border(ext, CCSO_PADDING_SIZE, h.recon);
    dec_ccso_frame(h.recon, h.dec, ext);
    free(ext);
    dec_harness_free(&h);
    return 0;
}

/* Loop restoration as the decoder runs it (Decoder/Codec/EbDecParseObu.c:2421-2455, single thread): the boundary lines of
 * the DEBLOCKED picture (after_cdef = 0), of the picture to restore (after_cdef = 1), then the frame filter with
 * optimized_lr = 0.  cur: in/out; dbk: the deblocked picture. */
REF_API int refdec_lr_frame(const svtf_geom *g, void *y, int sy, void *cb, int scb, void *cr, int scr, void *dy, int dsy, void *dcb, int dscb,
                            void *dcr, int dscr, const dec_mi_soa *soa, const svtf_lr_params *lr, const svtf_lr_unit *const units[3]) {
    dec_harness    h, hd;
    dec_harness_lr l, ld;
    if (dec_harness_init(&h, g, soa, y, sy, cb, scb, cr, scr) || dec_harness_fill_lr(&h, &l, g, lr, units))
        return -1;
    /* the deblocked picture goes through the same handle: swap it in for the first save */
    if (dec_harness_init(&hd, g, soa, dy, dsy, dcb, dscb, dcr, dscr) || dec_harness_fill_lr(&hd, &ld, g, lr, units))
        return -1;
    EbPictureBufferDesc *cur = h.pic_buf->ps_pic_buf;
    h.pic_buf->ps_pic_buf    = hd.recon;
    svt_aom_dec_av1_loop_restoration_save_boundary_lines(h.dec, 0);
    h.pic_buf->ps_pic_buf = cur;
    svt_aom_dec_av1_loop_restoration_save_boundary_lines(h.dec, 1);
    svt_aom_dec_av1_loop_restoration_filter_frame(h.dec, 0, 1);
    dec_harness_lr_store(&h, &l, g);
    dec_harness_lr_free(&l), dec_harness_lr_free(&ld);
    dec_harness_free(&h), dec_harness_free(&hd);
    return 0;
}
