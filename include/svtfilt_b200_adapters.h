/*
 * svtfilt_b200_adapters.h — Tier 2 of the drop-in boundary: the reference's own frame-level entry points,
 * re-implemented over the Tier-1 C ABI (adapters/svtf_enc_adapters.c).  Include AFTER the reference headers
 * (EbPictureControlSet.h, EbSequenceControlSet.h, EbCcso.h); the prototypes below are the reference's own and
 * are repeated only for documentation.
 *
 *   void svt_av1_loop_filter_frame(EbPictureBufferDesc *frame_buffer, PictureControlSet *pcs,
 *                                  int32_t plane_start, int32_t plane_end);   Encoder/Codec/EbDeblockingFilter.c:624
 *   void svt_av1_cdef_frame(SequenceControlSet *scs, PictureControlSet *pcs); Encoder/Codec/EbEncCdef.c:284
 *   void finish_cdef_search(PictureControlSet *pcs);                          Encoder/Codec/EbEncCdef.c:728
 *   void ccso_frame(EbPictureBufferDesc *frame, PictureControlSet *pcs, MacroblockdPlane *pd,
 *                   uint16_t *ext_rec_y);                                     Common/Codec/EbCcso.c:626
 *   void ccso_search(PictureControlSet *pcs, MacroblockdPlane *pd, int rdmult, const uint16_t *ext_rec_y,
 *                    uint16_t *rec_uv[3], uint16_t *org_uv[3]);               Encoder/Codec/EbPickccso.c:785
 *   EbErrorType svt_av1_pick_filter_level(EbPictureBufferDesc *srcBuffer, PictureControlSet *pcs,
 *                                         LpfPickMethod method);              Encoder/Codec/EbDeblockingFilter.c:1129
 */
#ifndef SVTFILT_B200_ADAPTERS_H
#define SVTFILT_B200_ADAPTERS_H
#ifdef __cplusplus
extern "C" {
#endif

struct PictureControlSet;
struct SequenceControlSet;

/* Whole-picture replacement of the static cdef_seg_search (Encoder/Codec/EbCdefProcess.c:114): fills
 * pcs->mse_seg, pcs->skip_cdef_seg and pcs->cdef_dir_data.  The first CDEF segment to arrive calls it. */
void svtf_cdef_search_picture(struct PictureControlSet *pcs, struct SequenceControlSet *scs);

/* The pixel work of pad_ref_and_set_flags (Encoder/Codec/EbEncDecProcess.c:1589-1744): right/bottom replication, border
 * padding and the 8-bit / 2-bit split (or narrowing) of the reference picture, produced on the GPU (resident picture, or the
 * recon buffer uploaded first).  pad_ref_and_set_flags calls it in place of its lines 1604-1744 and keeps its bookkeeping. */
void svtf_pad_ref_pictures(struct PictureControlSet *pcs, struct SequenceControlSet *scs);

/* Also provided with the reference's own name and signature (Encoder/Codec/EbRestProcess.c:405):
 *   void svt_av1_superres_upscale_frame(struct Av1Common *cm, PictureControlSet *pcs, SequenceControlSet *scs);
 * and (Encoder/Codec/EbRestorationPick.c:1471 / :1555):
 *   void restoration_seg_search(int32_t *rst_tmpbuf, Yv12BufferConfig *org_fts, const Yv12BufferConfig *src,
 *                               Yv12BufferConfig *trial_frame_rst, PictureControlSet *pcs, uint32_t segment_index);
 *   void rest_finish_search(PictureControlSet *pcs);
 * (Wiener and self-guided tools; segment 0 searches the whole picture on the GPU, the other segments return.) */

/* CUDA device used by contexts the adapters create (default 0). */
void svtf_adapter_set_device(int device);
/* 1: keep the picture resident on the GPU between DLF, CDEF search, CDEF apply and CCSO of one pcs (no
 * intermediate download/upload); ccso_frame (or svtf_adapter_fetch) brings it back.  0 (default): every entry
 * point uploads and downloads, i.e. each is a stand-alone drop-in. */
void svtf_adapter_set_resident(int on);
/* 1: page-lock (svtf_host_register) the plane allocations of the pictures the adapters see, once per allocation — the recon pictures
 * of the encoder live in a long-lived pool — so uploads and downloads are direct DMA.  Only for callers whose picture buffers outlive
 * the adapter's slot for that pcs (svtf_adapter_release unlocks them).  0 (default): plain (staged) copies from pageable memory. */
void svtf_adapter_set_pin_pictures(int on);
/* Host threads that gather mi_grid_base into the side-info byte planes (default 8). */
void svtf_adapter_set_mi_threads(int n);
/* Copy the resident picture of `pcs` back into `pic` (EbPictureBufferDesc *) — the end of a resident chain that does
 * not finish with ccso_frame. */
void svtf_adapter_fetch(const struct PictureControlSet *pcs, void *pic);
/* Destroy the cached GPU context of a picture (call where the pipeline recycles the pcs). */
void svtf_adapter_release(const struct PictureControlSet *pcs);

/* ---- decoder side (adapters/svtf_dec_adapters.c; include after EbDecHandle.h) ----
 *   void svt_aom_dec_av1_loop_filter_frame(EbDecHandle*, EbPictureBufferDesc *recon, LfCtxt*, int32_t plane_start,
 *                                          int32_t plane_end, int32_t is_mt, int enable_flag);   Decoder/Codec/EbDecLF.c:697
 *   void svt_cdef_frame(EbDecHandle*, int enable_flag);                                          Decoder/Codec/EbDecCdef.c:507
 *   void dec_ccso_frame(EbPictureBufferDesc *frame, EbDecHandle*, uint16_t *ext_rec_y);          Decoder/Codec/EbDecCcso.c:384
 *   void svt_aom_dec_av1_loop_restoration_save_boundary_lines(EbDecHandle*, int after_cdef);     Decoder/Codec/EbDecRestoration.c:473
 *   void svt_aom_dec_av1_loop_restoration_filter_frame(EbDecHandle*, int optimized_lr, int enable_flag);           :429
 * and, for EbSvtAv1DecConfiguration.threads > 1 (every decoder thread calls these; the calling thread launches, the workers
 * pass through to the frame-end barrier):
 *   void svt_aom_dec_av1_loop_filter_frame_mt(EbDecHandle*, EbPictureBufferDesc *recon, LfCtxt*, int32_t plane_start,
 *                                             int32_t plane_end, DecThreadCtxt*);                Decoder/Codec/EbDecProcess.c:668
 *   void svt_cdef_frame_mt(EbDecHandle*, DecThreadCtxt*);                                                            :810
 *   void svt_aom_dec_av1_loop_restoration_filter_frame_mt(EbDecHandle*, DecThreadCtxt*);                             :1022
 */
struct EbDecHandle;
void svtf_dec_adapter_set_device(int device);
/* 1: the frame stays on the GPU between the LF, CDEF, CCSO and loop-restoration calls of one decoder handle (the two
 * save_boundary_lines calls become no-ops: the deblocked picture is still on the device); svtf_dec_adapter_fetch after the
 * last enabled stage brings it back.  0 (default): every entry point uploads and downloads. */
void svtf_dec_adapter_set_resident(int on);
void svtf_dec_adapter_fetch(const struct EbDecHandle *dec, void *pic);
void svtf_dec_adapter_release(const struct EbDecHandle *dec);

#ifdef __cplusplus
}
#endif
#endif
