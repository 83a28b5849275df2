/*
 * svtfilt_b200.h — Tier-1 C ABI of the B200-native in-loop filter path
 * (deblock -> CDEF -> CCSO, plus the encoder's CDEF strength search).
 *
 * Plain C, plain pointers and sizes, no SVT-AV1 headers, no torch types.
 * Every entry point names the reference interface it stands in for
 * (paths relative to the reference tree, `Source/Lib/...`).  The Tier-2
 * adapters that keep the reference's own function names/signatures
 * (svt_av1_loop_filter_frame(pcs...), svt_av1_cdef_frame(scs,pcs), ...) are
 * declared in include/svtfilt_b200_adapters.h and described in INTEGRATION.md.
 *
 * Conventions
 *   - all functions return 0 on success and a negative SVTF_E_* code on error;
 *     svtf_last_error() gives a human-readable message (per context).
 *   - strides are in SAMPLES (like EbPictureBufferDesc::stride_y), plane
 *     pointers address the visible origin (buffer + org_y*stride + org_x).
 *   - sample_bytes 1 -> uint8_t samples (8-bit pipeline), 2 -> uint16_t
 *     (10/12-bit or `is_16bit_pipeline`), exactly like the reference picks
 *     recon_pic vs recon_pic_16bit (Encoder/Codec/EbRestProcess.c:135-153).
 *   - 4:2:0 only (the encoder path is 4:2:0 only, Encoder/Codec/EbEncCdef.c:324-329).
 *   - work is stream-ordered on the context's CUDA stream; host-visible
 *     results are valid after svtf_sync() (download calls sync themselves
 *     unless stated otherwise).
 *   - there is NO CPU fallback: without a CUDA device every call fails with
 *     SVTF_E_CUDA.
 */
#ifndef SVTFILT_B200_H
#define SVTFILT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SVTF_OK 0
#define SVTF_E_ARG (-1)      /* bad argument / unsupported configuration            */
#define SVTF_E_CUDA (-2)     /* CUDA runtime error (message in svtf_last_error)     */
#define SVTF_E_STATE (-3)    /* call order violated (e.g. no frame uploaded)        */
#define SVTF_E_NOMEM (-4)    /* host or device allocation failed                    */

#define SVTF_MAX_SEGMENTS 8
#define SVTF_SEG_LVL_MAX 8
#define SVTF_REF_FRAMES 8
#define SVTF_CDEF_MAX_STRENGTHS 16 /* CDEF_MAX_STRENGTHS, Common/Codec/EbAv1Structs.h:359-369 */
#define SVTF_TOTAL_STRENGTHS 64    /* TOTAL_STRENGTHS,   Common/Codec/EbCdef.h:47            */
#define SVTF_CCSO_BAND_NUM 128
#define SVTF_CCSO_LUT_SIZE (SVTF_CCSO_BAND_NUM * 16)

typedef struct svtf_ctx svtf_ctx;

/* Frame geometry.  width/height: unpadded source size (deblock and CCSO limit,
 * Encoder/Codec/EbDeblockingFilter.c:99-101,173-178); mi_cols/mi_rows: coded size in 4x4
 * units (CDEF frame limit, Encoder/Codec/EbEncCdef.c:356-382). */
typedef struct svtf_geom {
    int32_t width, height;
    int32_t mi_cols, mi_rows;
    int32_t bit_depth;    /* 8, 10 or 12 */
    int32_t sample_bytes; /* 1 or 2      */
} svtf_geom;

/* ---- per-4x4 ("mi") side info, one packed 32-bit word per unit ------------------
 * Replaces the pointer chase through pcs->mi_grid_base[..]->mbmi (Encoder/Codec/
 * EbDeblockingFilter.c:187-199, Encoder/Codec/EbEncCdef.c:238-280).
 *   bits  0-2  log2(luma tx width)-2     bits  3-5  log2(luma tx height)-2
 *   bits  6-8  log2(chroma tx width)-2   bits  9-11 log2(chroma tx height)-2
 *   bits 12-14 log2(block width)-2       bits 15-17 log2(block height)-2
 *   bit  18    skip && is_inter (deblock "skipped")   bit 19  skip (CDEF)
 *   bits 20-22 ref_frame[0]              bit  23    mode_lf_lut[mode]
 *   bits 24-26 segment_id                bits 27-31 reserved (0)
 */
typedef uint32_t svtf_mi_t;

/* Encoder-side packer: same derivation as get_transform_size()
 * (Encoder/Codec/EbDeblockingFilter.c:143-160): luma tx = tx_depth_to_tx_size[skip&&inter ? 0 :
 * tx_depth][bsize], chroma tx = av1_get_max_uv_txsize(bsize,1,1).  Arrays hold n entries
 * (BlockSize enum, tx_depth 0..2, skip 0/1, ref_frame[0] 0..7, PredictionMode 0..24, segment id). */
int svtf_pack_mi_enc(size_t n, const uint8_t *bsize, const uint8_t *tx_depth, const uint8_t *skip,
                     const uint8_t *ref_frame0, const uint8_t *mode, const uint8_t *segment_id, svtf_mi_t *out);
/* Decoder-side packer: per-unit transform sizes are given explicitly as TxSize enums
 * (LfCtxt.tx_size_l / tx_size_uv, Decoder/Codec/EbDecLF.c:133-145). */
int svtf_pack_mi_dec(size_t n, const uint8_t *bsize, const uint8_t *tx_size_y, const uint8_t *tx_size_uv,
                     const uint8_t *skip, const uint8_t *ref_frame0, const uint8_t *mode,
                     const uint8_t *segment_id, svtf_mi_t *out);

/* ---- deblock -----------------------------------------------------------------------------
 * struct LoopFilter (Common/Codec/EbDefinitions.h:1903-1920), the LF part of
 * SegmentationParams (Common/Codec/EbSegmentationParams.h) and DeltaLfParams
 * (Common/Codec/EbAv1Structs.h:345-357) + ppcs->curr_delta_lf. */
typedef struct svtf_lf_params {
    int32_t filter_level[2];
    int32_t filter_level_u, filter_level_v;
    int32_t sharpness_level;
    int32_t mode_ref_delta_enabled;
    int8_t  ref_deltas[SVTF_REF_FRAMES];
    int8_t  mode_deltas[2];
    int32_t segmentation_enabled;
    int16_t seg_feature_enabled[SVTF_MAX_SEGMENTS][SVTF_SEG_LVL_MAX];
    int16_t seg_feature_data[SVTF_MAX_SEGMENTS][SVTF_SEG_LVL_MAX];
    int32_t delta_lf_present, delta_lf_multi;
    int32_t curr_delta_lf[4];
} svtf_lf_params;

/* Host-side per-frame tables the deblock kernel consumes: lvl[plane][segment][dir][ref][mode_lf]
 * (3*8*2*8*2 = 768 bytes; svt_av1_loop_filter_frame_init, Common/Codec/EbDeblockingCommon.c:76, or
 * svt_aom_get_filter_level_delta_lf :37 when delta_lf_present) and the per-level thresholds
 * (svt_aom_update_sharpness :554; hev = level >> 4, Encoder/Codec/EbDeblockingFilter.c:47). */
int svtf_lf_build_tables(const svtf_lf_params *lf, uint8_t lvl[768], uint8_t lim[64], uint8_t mblim[64], uint8_t hev[64]);

/* ---- CDEF ---------------------------------------------------------------------------------
 * CdefParams (Common/Codec/EbAv1Structs.h:359-369): strength byte v -> pri=v/4, sec=v%4 (3->4). */
typedef struct svtf_cdef_params {
    int32_t cdef_damping;
    int32_t cdef_bits;
    uint8_t cdef_y_strength[SVTF_CDEF_MAX_STRENGTHS];
    uint8_t cdef_uv_strength[SVTF_CDEF_MAX_STRENGTHS];
} svtf_cdef_params;

/* CdefControls search lists (Encoder/Codec/EbPictureControlSet.h:592-628). */
typedef struct svtf_cdef_search_cfg {
    int32_t first_pass_fs_num;
    int32_t default_second_pass_fs_num;
    uint8_t default_first_pass_fs[SVTF_TOTAL_STRENGTHS];
    uint8_t default_second_pass_fs[SVTF_TOTAL_STRENGTHS];
    int8_t  default_first_pass_fs_uv[SVTF_TOTAL_STRENGTHS];  /* -1 = masked */
    int8_t  default_second_pass_fs_uv[SVTF_TOTAL_STRENGTHS]; /* -1 = masked */
    int32_t subsampling_factor; /* 1, 2 or 4 */
    int32_t base_q_idx;         /* damping = 3 + (base_q_idx >> 6), Encoder/Codec/EbCdefProcess.c:147 */
    int32_t zero_fs_cost_bias;  /* 0 = off; Encoder/Codec/EbEncCdef.c:845-851 */
    uint32_t lambda;            /* full_lambda as produced by svt_aom_av1_lambda_assignment_function_table */
} svtf_cdef_search_cfg;

/* ---- CCSO ---------------------------------------------------------------------------------
 * CcsoInfo (Common/Codec/EbAv1Structs.h:407-427). */
typedef struct svtf_ccso_params {
    uint8_t ccso_enable[3];
    uint8_t ccso_bo_only[3];
    uint8_t quant_idx[3];          /* thr = {16,8,32,64}[quant_idx], Common/Codec/EbCcso.c:636 */
    uint8_t ext_filter_support[3]; /* 0..5, Common/Codec/EbCcso.c:204-234 */
    uint8_t edge_clf[3];
    uint8_t max_band_log2[3];
    int8_t  filter_offset[3][SVTF_CCSO_LUT_SIZE];
} svtf_ccso_params;

/* ---- context ---------------------------------------------------------------------------- */
/* One context = one picture in flight on one GPU: device surfaces for max_geom, side info,
 * pinned staging and one CUDA stream.  Re-entrant per context (the reference is re-entrant per
 * pcs, SURVEY §8b "Threading"). */
int  svtf_ctx_create(int device, const svtf_geom *max_geom, svtf_ctx **out);
void svtf_ctx_destroy(svtf_ctx *ctx);
const char *svtf_last_error(const svtf_ctx *ctx); /* ctx may be NULL: last global error */
int  svtf_sync(svtf_ctx *ctx);
/* Run this context's work on a caller-owned CUDA stream (a cudaStream_t); NULL restores the
 * context's own stream.  The caller keeps the stream alive while the context uses it. */
int svtf_ctx_set_stream(svtf_ctx *ctx, void *cuda_stream);
/* Number of this library's kernel launches issued on the context so far. */
uint64_t svtf_launch_count(const svtf_ctx *ctx);

/* Page-locked host memory for picture buffers: uploads/downloads from such buffers are true
 * asynchronous DMA (a caller may equally cudaHostRegister its own recon buffers). */
void *svtf_host_alloc(size_t bytes);
void  svtf_host_free(void *p);

/* Page-lock a caller-owned buffer in place (cudaHostRegister; already-registered buffers are accepted): the recon pictures of an
 * encoder / decoder live in long-lived pools (EbPictureBufferDesc allocations), so registering each once makes every later
 * svtf_upload_frame / svtf_download_frame on it a direct DMA instead of a staged pageable copy. */
int svtf_host_register(void *p, size_t bytes);
int svtf_host_unregister(void *p);

/* Start a picture: sets the geometry (must fit max_geom) and resets the surface chain. */
int svtf_frame_begin(svtf_ctx *ctx, const svtf_geom *geom);

/* Start ONE HORIZONTAL STRIPE of a picture on this context (single-picture multi-GPU split): the context
 * owns superblock rows [first_sb_row, first_sb_row + num_sb_rows), 64 luma rows each.  After this call every
 * entry point keeps taking WHOLE-PICTURE host arrays (planes with their strides, the mi grid, fb_strength_idx, CCSO
 * block flags, the search outputs) and reads/writes only the stripe's part: upload copies the stripe plus 8 rows
 * beyond each interior boundary, download writes the stripe's rows only, svtf_cdef_search_frame fills the stripe's
 * filter-block rows of mse/skip/dir/var.  N contexts (one per GPU), called with identical arguments and disjoint
 * row ranges, produce the same bytes as one whole-picture context.  The boundary deblock edge is recomputed on both
 * sides (the reference's serial raster order, Encoder/Codec/EbDeblockingFilter.c:624-660, has no such boundary); no
 * inter-GPU exchange is needed when the input comes from the host.  max_geom must cover the stripe plus up to two
 * superblock rows of context.  svtf_cdef_pick is not available on a striped context: gather mse/skip and call
 * svtf_cdef_pick_table (GPU) or svtf_cdef_pick_host. */
int svtf_stripe_begin(svtf_ctx *ctx, const svtf_geom *full_geom, int32_t first_sb_row, int32_t num_sb_rows);

/* Host -> device copy of the reconstructed picture (the `frame_buffer` / recon_pic argument of
 * the reference entry points).  Copies the coded area (mi_cols*4 x mi_rows*4). */
int svtf_upload_frame(svtf_ctx *ctx, const void *y, int32_t stride_y, const void *cb, int32_t stride_cb,
                      const void *cr, int32_t stride_cr);
/* Host -> device copy into the CURRENT end of the chain, keeping the chain state (in particular the
 * remembered pre-CDEF picture).  Lets a caller that ran CDEF elsewhere hand over the post-CDEF picture while
 * the pre-CDEF luma (ccso_frame's ext_rec_y argument) is the one uploaded before. */
int svtf_overwrite_current(svtf_ctx *ctx, const void *y, int32_t stride_y, const void *cb, int32_t stride_cb,
                           const void *cr, int32_t stride_cr);
/* The picture as it was BEFORE CDEF, for a caller that ran deblock / CDEF elsewhere and hands over only the later
 * stages: ccso_frame classifies on its luma (ext_rec_y), loop restoration takes its stripe borders from it
 * (RestorationStripeBoundaries, Common/Codec/EbRestoration.c:1522-1574).  The current picture is not touched. */
int svtf_upload_pre_cdef(svtf_ctx *ctx, const void *y, int32_t stride_y, const void *cb, int32_t stride_cb,
                         const void *cr, int32_t stride_cr);
/* Source picture (pcs->cdef_input_source / input_frame16bit) for the strength search. */
int svtf_upload_source(svtf_ctx *ctx, const void *y, int32_t stride_y, const void *cb, int32_t stride_cb,
                       const void *cr, int32_t stride_cr);
/* Packed mi grid, mi_rows x mi_cols entries with row stride mi_stride. */
int svtf_set_mi(svtf_ctx *ctx, const svtf_mi_t *mi, int32_t mi_stride);

/* The same side info as separate byte planes ([mi_rows][stride] each), packed on the device: what the producers know per
 * block (EncDec's mode-info update, Encoder/Codec/EbCodingUnit.h:66-68 / the decoder's parser and LfCtxt fill,
 * Decoder/Codec/EbDecLF.c:63-90) can be written straight into these planes, 1 B per field, with no host pass over the
 * picture (6-7 B instead of 4 B per 4x4 over PCIe; at 4K: 0.10 ms against 2.7 ms for svtf_pack_mi_enc on one core +
 * svtf_set_mi).  Fields and derivations as svtf_pack_mi_enc / svtf_pack_mi_dec.
 * Asynchronous; an out-of-range enum is reported by the next svtf_sync / svtf_download_frame (SVTF_E_ARG). */
int svtf_set_mi_soa_enc(svtf_ctx *ctx, const uint8_t *bsize, const uint8_t *tx_depth, const uint8_t *skip, const uint8_t *ref_frame0,
                        const uint8_t *mode, const uint8_t *segment_id, int32_t stride);
int svtf_set_mi_soa_dec(svtf_ctx *ctx, const uint8_t *bsize, const uint8_t *tx_size_y, const uint8_t *tx_size_uv, const uint8_t *skip,
                        const uint8_t *ref_frame0, const uint8_t *mode, const uint8_t *segment_id, int32_t stride);
/* The packed words as the device holds them ([mi_rows][stride]); for tests and debugging. */
int svtf_get_mi(svtf_ctx *ctx, svtf_mi_t *out, int32_t stride);
/* ---- recon hand-off -----------------------------------------------------------------------------------
 * svtf_download_ref_padded: the current end of the chain leaves the device as the PADDED reference picture that
 * pad_ref_and_set_flags (Encoder/Codec/EbEncDecProcess.c:1589-1760) builds on the host: the last valid column/row is
 * replicated into pad_right/pad_bottom (svt_aom_pad_input_picture_16bit, Common/Codec/EbMcp.c:296), then the borders
 * org_x/org_y are replicated (svt_aom_generate_padding / _16_bit, Common/Codec/EbMcp.c:95/201); optionally the padded
 * picture is also split into the 8-bit + 2-bit planes (svt_aom_un_pack2d, Common/Codec/EbPictureOperators.c:235) or, for
 * 8-bit content held in 16-bit samples, narrowed (svt_convert_16bit_to_8bit_c, Common/C_DEFAULT/EbPackUnPack_C.c:277).
 * dst[p] points at the FIRST sample of the padded allocation of plane p (EbPictureBufferDesc.buffer_y/cb/cr), strides in
 * samples; the written rectangle is (width + 2*org_x) x (height + 2*org_y) (chroma: every term halved, rounding up).
 * dst8 / dstn may be NULL (no split); they require 16-bit samples, dstn additionally bit_depth 10.  Synchronises.
 * dst == NULL: the kernel only — the padded planes stay in the context's device staging block (non-NULL dst8 / dstn then
 * only say which planes to produce), stream-ordered, not synchronised; used to time the kernel apart from the DMA. */
typedef struct svtf_ref_layout {
    int32_t width, height;         /* EbPictureBufferDesc.width / height of the destination (luma samples) */
    int32_t org_x, org_y;          /* its border (luma samples) */
    int32_t pad_right, pad_bottom; /* scs->pad_right / pad_bottom: trailing columns / rows of width / height that replicate the last valid one */
} svtf_ref_layout;
int svtf_download_ref_padded(svtf_ctx *ctx, const svtf_ref_layout *layout, void *const dst[3], const int32_t dst_stride[3],
                             uint8_t *const dst8[3], const int32_t stride8[3], uint8_t *const dstn[3], const int32_t striden[3]);
/* svt_convert_pic_8bit_to_16bit (Encoder/Codec/EbRestProcess.c:235; call site Encoder/Codec/EbDlfProcess.c:81-95): a context with
 * 16-bit samples takes the picture as 8-bit planes (half the bytes over PCIe) and widens it on the device.  Otherwise as
 * svtf_upload_frame. */
int svtf_upload_frame_8as16(svtf_ctx *ctx, const uint8_t *y, int32_t stride_y, const uint8_t *cb, int32_t stride_cb, const uint8_t *cr,
                            int32_t stride_cr);
/* Device -> host copy of the current end of the filter chain; synchronises. */
int svtf_download_frame(svtf_ctx *ctx, void *y, int32_t stride_y, void *cb, int32_t stride_cb, void *cr,
                        int32_t stride_cr);
/* Same copy, stream-ordered and NOT synchronised: the data is valid after svtf_sync(). */
int svtf_download_frame_async(svtf_ctx *ctx, void *y, int32_t stride_y, void *cb, int32_t stride_cb, void *cr,
                              int32_t stride_cr);
/* Rewind the chain to the uploaded (unfiltered) picture without copying — lets a benchmark
 * re-run the chain on device-resident input. */
int svtf_rewind(svtf_ctx *ctx);

/* ---- filters ------------------------------------------------------------------------------ */
/* svt_av1_loop_filter_frame(frame_buffer, pcs, plane_start, plane_end)
 * (Encoder/Codec/EbDeblockingFilter.c:624) == svt_aom_dec_av1_loop_filter_frame
 * (Decoder/Codec/EbDecLF.c:697) on the same side info. */
int svtf_deblock_frame(svtf_ctx *ctx, const svtf_lf_params *lf, int32_t plane_start, int32_t plane_end);
/* The part of set_lpf_parameters (Encoder/Codec/EbDeblockingFilter.c:162-285) that depends on the mode info alone — transform-edge
 * alignment, the skip / prediction-block rule, the filter length, the two level keys — is evaluated once per mode-info upload into
 * two edge-code planes (one word per 4x4 unit; luma, chroma) that every deblock launch on that mode info reads: svtf_set_mi and
 * svtf_set_mi_soa_* launch that kernel behind their copy (one launch, ~6 us at 4K).  This call rebuilds the planes from the mode
 * info on the device (stream-ordered) — for timing, and for callers that assemble the grid from pieces (svtf_set_mi_own +
 * svtf_stripe_exchange), where otherwise the first svtf_deblock_frame* / svtf_lf_search_frame does it. */
int svtf_deblock_prepare(svtf_ctx *ctx);
/* The same with per-superblock loop-filter deltas, as the DECODER keeps them (SBInfo.sb_delta_lf, read by
 * dec_set_lpf_parameters through svt_aom_get_filter_level_delta_lf, Decoder/Codec/EbDecLF.c:159-193,357,548):
 * sb_delta_lf[sb_rows*sb_cols][4] (FRAME_LF_COUNT entries per superblock, raster order) replaces lf->curr_delta_lf for
 * the 4x4 units of each superblock when lf->delta_lf_present; sb_size_log2 is 6 or 7. */
int svtf_deblock_frame_sb_delta(svtf_ctx *ctx, const svtf_lf_params *lf, int32_t plane_start, int32_t plane_end,
                                const int32_t *sb_delta_lf, int32_t sb_cols, int32_t sb_size_log2);

/* svt_av1_cdef_frame(scs, pcs) (Encoder/Codec/EbEncCdef.c:284) == svt_cdef_frame
 * (Decoder/Codec/EbDecCdef.c:507).  fb_strength_idx: one int8 per 64x64 filter block
 * (row-major, nvfb x nhfb) = mbmi.cdef_strength of its top-left mi. */
int svtf_cdef_apply_frame(svtf_ctx *ctx, const svtf_cdef_params *cdef, const int8_t *fb_strength_idx);

/* ccso_frame(frame, pcs, pd, ext_rec_y) (Common/Codec/EbCcso.c:626) on 8- or 16-bit samples.
 * Luma classification input = the picture as it was before the last svtf_cdef_apply_frame
 * (post-deblock / pre-CDEF, Encoder/Codec/EbCdefProcess.c:490-502); if no CDEF ran in this chain, the
 * current picture.  blk_flags[p]: one byte per CCSO filter block of plane p (256x256 luma
 * samples; row-major ceil(height/256) x ceil(width/256)) = mbmi.ccso_blk_{y,u,v}; NULL = all on. */
int svtf_ccso_apply_frame(svtf_ctx *ctx, const svtf_ccso_params *ccso, const uint8_t *const blk_flags[3]);

/* Whole post-reconstruction chain in one call (decoder order, Decoder/Codec/EbDecParseObu.c:2382-2433):
 * deblock -> CDEF -> CCSO; any of lf / cdef / ccso may be NULL to skip that stage. */
int svtf_filter_chain(svtf_ctx *ctx, const svtf_lf_params *lf, const svtf_cdef_params *cdef,
                      const int8_t *fb_strength_idx, const svtf_ccso_params *ccso,
                      const uint8_t *const blk_flags[3]);

/* ---- loop restoration (the stage after CDEF/CCSO on the same surfaces) ----------------------------------
 * svt_av1_loop_restoration_filter_frame (Common/Codec/EbRestoration.c:1179) with the stripe boundaries of
 * svt_av1_loop_restoration_save_boundary_lines (:1682): Wiener (7-tap separable, svt_av1_[highbd_]wiener_convolve_add_src_c,
 * Common/Codec/convolve.c:100/194) and self-guided (svt_apply_selfguided_restoration_c :957) restoration per unit.
 * RestorationUnitInfo (EbRestoration.h:184-188) / RestorationInfo (:225-242) as plain data: */
#define SVTF_RESTORE_NONE 0
#define SVTF_RESTORE_WIENER 1
#define SVTF_RESTORE_SGRPROJ 2
typedef struct svtf_lr_unit {
    int32_t restoration_type;                     /* SVTF_RESTORE_*                                        */
    int16_t wiener_vfilter[8], wiener_hfilter[8]; /* WienerInfo.vfilter / hfilter (InterpKernel, tap 7 = 0) */
    int32_t sgr_ep, sgr_xqd[2];                   /* SgrprojInfo                                           */
} svtf_lr_unit;
typedef struct svtf_lr_params {
    int32_t frame_restoration_type[3]; /* SVTF_RESTORE_NONE: the plane is left as it is; anything else: per-unit types apply */
    int32_t restoration_unit_size[3];  /* in samples of the plane (64, 128 or 256; chroma may be 32)                         */
    int32_t optimized_lr;              /* rsi->optimized_lr: stripe borders from the picture itself (:350-377)               */
} svtf_lr_params;
/* Units per row / column of a plane (count_units_in_tile, EbRestoration.c:124: round to nearest, at least 1). */
int svtf_lr_units(const svtf_geom *geom, int32_t plane, int32_t unit_size, int32_t *horz_units, int32_t *vert_units);
/* Filters the current picture.  units[p]: vert_units x horz_units entries, raster order.  The rows above/below each
 * 64-row stripe come from the DEBLOCKED picture (the surface CDEF read), as the reference saves them before CDEF; on a
 * context where no CDEF ran they come from the current picture. */
int svtf_lr_apply_frame(svtf_ctx *ctx, const svtf_lr_params *lr, const svtf_lr_unit *const units[3]);

/* svt_av1_pick_filter_level(src, pcs, LPF_PICK_FROM_FULL_IMAGE) (Encoder/Codec/EbDeblockingFilter.c:1129-1252): searches
 * the deblocking level of luma (one level for both directions), then U, then V, each with search_filter_level (:886-985)
 * — trial deblock of that plane, SSE against the uploaded source, keep the unfiltered picture.  `start` carries the
 * levels the search starts from (the reference's last_frame_filter_level = the frame header's levels on entry, after the
 * optional reference-frame averaging) and every other loop-filter parameter; note the reference starts the luma search
 * from last_frame_filter_level[2] (the U level) unless dlf_avg is set (:903-909, dir = 2) and so does this function.
 * The picture on the context is left unfiltered.  out_levels = {luma, luma, u, v}. */
typedef struct svtf_lf_search_cfg {
    int32_t dlf_avg;                /* ppcs->dlf_ctrls.dlf_avg */
    int32_t early_exit_convergence; /* ppcs->dlf_ctrls.early_exit_convergence (0 = never) */
    int32_t only_4x4;               /* frm_hdr->tx_mode == ONLY_4X4 */
    int32_t search_chroma;          /* 0: keep the start chroma levels (the dlf_avg_uv path, :1211-1219) */
} svtf_lf_search_cfg;
int svtf_lf_search_frame(svtf_ctx *ctx, const svtf_lf_params *start, const svtf_lf_search_cfg *cfg, int32_t out_levels[4]);

/* ---- loop-restoration filter search (the caller of the restoration filters) ----------------------
 * restoration_seg_search + rest_finish_search (Encoder/Codec/EbRestorationPick.c:1471/1555):
 *   search_norestore_seg :1432 (SSE of the unfiltered unit), search_wiener_seg :1296 (svt_av1_compute_stats[_highbd]_c
 *   :671/:708 -> wiener_decompose_sep_sym :906 / linsolve_wiener :766 -> finalize_sym_filter :977 -> compute_score :937 ->
 *   finer_tile_search_wiener_seg :1042, every trial = try_restoration_unit_seg :129 = filter the unit + SSE),
 *   search_sgrproj_seg :1220 (search_selfguided_restoration :550: per parameter set ep the two guided filters of the unit,
 *   svt_get_proj_subspace_c :417, encode_xq :508, finer_search_pixel_proj_error :320; then the unit filtered for real), then
 *   search_wiener_finish :1380 / search_sgrproj_finish :1267 / search_switchable :1148 (count_wiener_bits :1008,
 *   count_sgrproj_bits :655, RDCOST_DBL) and the per-plane frame decision (:1596-1624).
 * The picture searched is the context's current picture (after CDEF / CCSO), the source the uploaded source picture.
 * The search filters WITHOUT stripe boundaries (scs->use_boundaries_in_rest_search = 0, EbEncHandle.c:4162: rows
 * above / below a stripe are the picture's own rows; the frame edge is replicated, svt_extend_frame).
 * use_prev_frame_coeffs: see svtf_lr_search_cfg.prev_units. */
typedef struct svtf_lr_search_cfg {
    int32_t restoration_unit_size[3];   /* rst_info[p].restoration_unit_size                                      */
    int32_t wn_enabled;                 /* cm->wn_filter_ctrls.enabled (Common/Codec/Av1Common.h:32-45)           */
    int32_t wn_filter_tap_lvl;          /* 1: Y 7-tap, UV 5-tap; 2: 5 / 5; 3: 3 / 3                               */
    int32_t wn_use_refinement;          /* refinement search around the solved taps                               */
    int32_t wn_max_one_refinement_step; /* only the step of 4                                                     */
    int32_t wn_use_chroma;              /* search U and V as well                                                 */
    int32_t sg_enabled;                 /* cm->sg_filter_ctrls.enabled (Av1Common.h:47-60)                        */
    int32_t sg_use_chroma;
    /* the ep range search_selfguided_restoration ends up with, per plane type [luma, chroma] (:567-589): either
     * ctrls->start_ep / end_ep / ep_inc / refine (step_range 16) or the range around the reference frames' ep with
     * ep_inc 1 and refine 1 (step_range < 16) */
    int32_t sg_start_ep[2], sg_end_ep[2], sg_ep_inc[2], sg_refine[2];
    int32_t rdmult;                     /* x->rdmult                                                              */
    int32_t wiener_restore_cost[2];     /* x->wiener_restore_cost (bits << AV1_PROB_COST_SHIFT)                   */
    int32_t sgrproj_restore_cost[2];    /* x->sgrproj_restore_cost                                                */
    int32_t switchable_restore_cost[3]; /* x->switchable_restore_cost                                             */
    /* wn_filter_ctrls.use_prev_frame_coeffs on a frame that is neither KEY nor INTRA_ONLY (:1313-1319): units whose entry here
     * has restoration_type SVTF_RESTORE_WIENER start the refinement from these taps (what the reference finds in
     * pcs->rst_info[p].unit_info[], copied from the first list-0 reference picture, EbRestProcess.c:586-597) instead of
     * solving for new ones.  NULL pointers: every unit is solved. */
    const svtf_lr_unit *prev_units[3];
} svtf_lr_search_cfg;
/* Per restoration unit, what restoration_seg_search leaves in RestUnitSearchInfo (Common/Codec/EbRestoration.h:349): sse of
 * RESTORE_NONE / RESTORE_WIENER (INT64_MAX: no usable filter) / RESTORE_SGRPROJ, the refined Wiener taps, the self-guided
 * parameters. */
typedef struct svtf_lr_unit_search {
    int64_t sse[3];
    int16_t wiener_vfilter[8], wiener_hfilter[8];
    int32_t sgr_ep, sgr_xqd[2];
    int32_t reserved;
} svtf_lr_unit_search;
/* out_frame_type[p] = rst_info[p].frame_restoration_type; out_units[p] (vert_units x horz_units of svtf_lr_units) = the
 * unit_info the reference leaves in rst_info[p] (ready for svtf_lr_apply_frame); out_search[p] may be NULL;
 * sg_frame_ep_cnt (may be NULL): the 16 counters the search adds to cm->sg_frame_ep_cnt (:1250). */
int svtf_lr_search_frame(svtf_ctx *ctx, const svtf_lr_search_cfg *cfg, int32_t out_frame_type[3], svtf_lr_unit *const out_units[3],
                         svtf_lr_unit_search *const out_search[3], int32_t sg_frame_ep_cnt[16]);
/* svt_av1_compute_stats_c / _highbd_c (:671/:708) of every restoration unit of one plane of the current picture against
 * the source: M[unit][win*win], H[unit][win*win][win*win] (int64, as the reference leaves them, bit-depth scaling and
 * symmetric fill included).  wiener_win = 7, 5 or 3. */
int svtf_lr_wiener_stats(svtf_ctx *ctx, int32_t plane, int32_t unit_size, int32_t wiener_win, int64_t *M, int64_t *H);
/* The host half of search_wiener_seg for one unit, from its statistics: taps (InterpKernel layout, tap 7 = 0) or
 * return 0 when the reference rejects the filter (decomposition failure or compute_score > 0); 1 = usable. */
int svtf_lr_wiener_from_stats(int32_t wiener_win, const int64_t *M, const int64_t *H, int16_t vfilter[8], int16_t hfilter[8]);

/* ccso_search / derive_ccso_filter (Encoder/Codec/EbPickccso.c:785, :464) for planes [plane_start, plane_end) on the
 * context's surfaces: classifier = the luma before the last CDEF pass (ext_rec_y), rec = the current picture (rec_uv),
 * org = the uploaded source (org_uv).  rdmult is the value derive_ccso_filter receives (ccso_search first multiplies
 * the lambda by clamp(base_q_idx, 1, 63), :788-793).  Writes the planes' entries of *out and blk_flags[p]
 * (one byte per 256x256 luma / 128x128 chroma block, the reference's mbmi.ccso_blk_y/u/v).  Same result as the C code:
 * the per-sample work (class histograms, trial filter + per-block SSE) runs on the GPU, the float LUT rounding and the
 * double RD decisions on the host as written in the reference. */
int svtf_ccso_derive_frame(svtf_ctx *ctx, int32_t rdmult, int32_t plane_start, int32_t plane_end, svtf_ccso_params *out,
                           uint8_t *const blk_flags[3]);

/* ---- CDEF strength search ------------------------------------------------------------------ */
/* cdef_seg_search(pcs, scs, segment) over the whole frame (Encoder/Codec/EbCdefProcess.c:114).
 * Input = current picture (post-deblock) + uploaded source.  Outputs (host, caller-owned, may be
 * NULL):  mse[2][nfb][64] u64 laid out like pcs->mse_seg, skip[nfb] like pcs->skip_cdef_seg,
 * dir[nfb][64] / var[nfb][64] = CdefDirData of each fb restricted to its 8x8 grid (by*8+bx). */
int svtf_cdef_search_frame(svtf_ctx *ctx, const svtf_cdef_search_cfg *cfg, uint64_t *mse, uint8_t *skip,
                           uint8_t *dir, int32_t *var);
/* finish_cdef_search(pcs) (Encoder/Codec/EbEncCdef.c:728) on the table left on the device by
 * svtf_cdef_search_frame: picks cdef_bits, the strength pairs and the per-fb index.
 * fb_strength_idx: nfb entries, -1 for fbs the search skipped.  Runs on the GPU: the 75 svt_search_one_dual passes
 * (Encoder/Codec/EbEncCdef.c:627-727) are enqueued back to back, the greedy lists stay in device memory. */
int svtf_cdef_pick(svtf_ctx *ctx, const svtf_cdef_search_cfg *cfg, svtf_cdef_params *out_params,
                   int8_t *fb_strength_idx);
/* Same pick on the GPU from a caller-owned HOST table laid out like pcs->mse_seg / pcs->skip_cdef_seg (the reference's
 * finish_cdef_search input; also the stripe path after the all-gather of the per-stripe rows).  Uploads 1 KiB per
 * filter block, then runs as svtf_cdef_pick. */
int svtf_cdef_pick_table(svtf_ctx *ctx, const svtf_cdef_search_cfg *cfg, const uint64_t *mse, const uint8_t *skip, int32_t nfb,
                         svtf_cdef_params *out_params, int8_t *fb_strength_idx);
/* svtf_cdef_pick_table WITHOUT any allocation, for callers that cannot call cudaMalloc (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid it).  No counterpart in the reference.
 * Same arguments, same results as svtf_cdef_pick_table; every device buffer the pick needs (table copy, skip
 * flags, search state, totals, per-block result) lives in one file-scope `static __device__` array of
 * SVTF_NOALLOC_SCRATCH_BYTES that every GPU loading the library holds until the process exits.  nfb above
 * SVTF_NOALLOC_MAX_FB is rejected with SVTF_E_ARG.  The array is shared by all calls on a GPU and keeps the previous
 * call's contents (the call clears what must start at zero); calls that overlap on one GPU would race on it, so the
 * library serialises them per device inside the process (the call returns only after its stream has drained). */
#define SVTF_NOALLOC_MAX_FB 8192                       /* filter blocks; an 8K picture has 120 x 68 = 8160 */
#define SVTF_NOALLOC_SCRATCH_BYTES (8u * 1024u * 1024u + 64u * 1024u)
int svtf_cdef_pick_table_noalloc(svtf_ctx *ctx, const svtf_cdef_search_cfg *cfg, const uint64_t *mse, const uint8_t *skip, int32_t nfb,
                                 svtf_cdef_params *out_params, int8_t *fb_strength_idx);
/* Pure host implementation of the same pick (single thread, as the reference runs it); no context needed. */
int svtf_cdef_pick_host(const svtf_cdef_search_cfg *cfg, const uint64_t *mse, const uint8_t *skip, int32_t nfb,
                        svtf_cdef_params *out_params, int8_t *fb_strength_idx);

/* ---- super-resolution: the normative upscale between CDEF and loop restoration ----------------
 * svt_av1_superres_upscale_frame (Encoder/Codec/EbRestProcess.c:405; decoder svt_av1_superres_upscale,
 * Decoder/Codec/EbDecSuperRes.c:119) = svt_av1_upscale_normative_rows (Common/Codec/EbSuperRes.c:214) on every plane:
 * an 8-tap horizontal filter with 64 phases, position step ((in << 14) + out / 2) / out, restarted per tile column
 * (:223-283), left / right frame edge replicated (upscale_normative_rect :102).  The current picture and, when one is
 * kept, the picture before CDEF (the restoration stripe borders, EbRestoration.c:1545-1560) are upscaled; the context's
 * geometry becomes upscaled_width x height (it must fit the geometry the context was created for).  The mode info no
 * longer matches the picture afterwards: only loop restoration, the hand-off and downloads follow. */
typedef struct svtf_superres_params {
    int32_t        upscaled_width;        /* cm->frm_size.superres_upscaled_width                                  */
    int32_t        superres_denominator;  /* cm->frm_size.superres_denominator, 9..16 (8 = unscaled: nothing to do) */
    int32_t        tile_cols;             /* cm->tiles_info.tile_cols, 1..64                                       */
    int32_t        tile_col_start_mi[65]; /* cm->tiles_info.tile_col_start_mi[0..tile_cols] of the coded picture   */
    const int16_t *filters;               /* av1_resize_filter_normative[64][8] (Common/Codec/EbSuperRes.h:22)     */
} svtf_superres_params;
int svtf_superres_upscale_frame(svtf_ctx *ctx, const svtf_superres_params *sr);

/* ---- one picture across several GPUs (BASELINE configs[3]: an 8K picture split by superblock-row stripes) ----------------
 * One context per GPU in stripe mode (svtf_stripe_begin), one process or thread per context.  The two inter-GPU steps of the path run
 * device to device over NVLink through NCCL (bound at run time from libnccl.so.2; SVTF_E_STATE when it cannot be loaded) and never touch
 * host memory.  They stand in for what the reference does inside one address space: its deblock / CDEF segment threads read each other's
 * rows of the same recon picture, and finish_cdef_search joins the segments' mse_seg rows under cdef_search_mutex
 * (Encoder/Codec/EbCdefProcess.c:406, :730). */
typedef struct svtf_comm svtf_comm;
/* ncclGetUniqueId on one rank; the caller distributes the 128 bytes to the others (any transport). */
int  svtf_comm_unique_id(void *id128);
/* ncclCommInitRank: collective over the nranks callers.  device = this rank's CUDA device. */
int  svtf_comm_create(int device, int nranks, int rank, const void *id128, svtf_comm **out);
void svtf_comm_destroy(svtf_comm *comm);
/* A picture that reaches the GPUs already split: only the rows the stripe OWNS (no halo) of the whole-picture arrays are copied. */
int svtf_upload_frame_own(svtf_ctx *ctx, const void *y, int32_t stride_y, const void *cb, int32_t stride_cb, const void *cr, int32_t stride_cr);
int svtf_set_mi_own(svtf_ctx *ctx, const svtf_mi_t *mi, int32_t mi_stride);
/* Halo exchange of the current picture with the stripes above / below (rank_up / rank_down; -1 where the stripe touches the picture
 * edge): the 8 luma / 4 chroma input rows and the 2 mode-info rows on either side of each interior boundary go to the neighbour's halo
 * rows and the neighbour's come here — what every later stage of the stripe reads beyond its own rows (the boundary deblock edge reads
 * <= 7 and writes <= 6 rows per side, Common/Codec/EbDeblockingCommon.c:787-823; CDEF taps reach 2 rows, Common/Codec/EbCdef.c:99-120).
 * Stream-ordered on the context's stream; all ranks of the picture must call it (it is a grouped ncclSend / ncclRecv). */
int svtf_stripe_exchange(svtf_ctx *ctx, svtf_comm *comm, int32_t rank_up, int32_t rank_down);
/* finish_cdef_search (Encoder/Codec/EbEncCdef.c:728) of a picture whose stripes were searched on different GPUs (svtf_cdef_search_frame on
 * each striped context, host outputs NULL): every rank's filter-block rows of mse / skip are gathered from device table to device table
 * (ragged all-gather = one NCCL group of broadcasts), then every rank runs the same pick on its device.  first_sb_row / num_sb_rows:
 * the stripe of every rank ([nranks]); fb_strength_idx: WHOLE-picture array (nfb entries); identical results on all ranks. */
int svtf_cdef_search_gather_pick(svtf_ctx *ctx, svtf_comm *comm, const svtf_cdef_search_cfg *cfg, const int32_t *first_sb_row,
                                 const int32_t *num_sb_rows, svtf_cdef_params *out_params, int8_t *fb_strength_idx);

/* ---- device-side access for multi-GPU stripe exchange and device-resident benches ----------- */
typedef struct svtf_dev_plane {
    void   *ptr;   /* device pointer to sample (0,0) */
    int32_t pitch; /* in samples */
    int32_t width, height;
} svtf_dev_plane;
/* which: 0 = uploaded recon, 1 = current end of chain, 2 = source */
int svtf_dev_planes(svtf_ctx *ctx, int which, svtf_dev_plane out[3]);
void *svtf_cuda_stream(svtf_ctx *ctx);

#ifdef __cplusplus
}
#endif
#endif /* SVTFILT_B200_H */
