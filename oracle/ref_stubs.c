/*
 * ref_stubs.c — link-time stand-ins for symbols the reference hot-path objects reference but never
 * reach from the calls ref_shim.c makes (pipeline FIFOs, mutexes, resize/super-res, allocation
 * failure printing), plus the two RTCD utility pointers (svt_memcpy, svt_log2f).  Deliberately
 * includes no reference header so the loose `void f(void)` stubs do not clash with prototypes.
 * TEST INFRASTRUCTURE ONLY.
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define STUB(name) /* weak: a reference object that defines the symbol wins */ \
    __attribute__((weak)) void name(void) {                                                \
        fprintf(stderr, "ref_stubs: unexpected call to " #name "\n"); \
        abort();                                                     \
    }
STUB(set_unscaled_input_16bit)
STUB(svt_aom_ac_quant_qtx)
STUB(svt_aom_assert_err)
STUB(svt_aom_downscaled_source_buffer_desc_ctor)
STUB(svt_aom_resize_frame)
/* referenced only by parts of EbMdRateEstimation.c the oracle never calls (it is compiled in for
 * svt_aom_get_syntax_rate_from_cdf, used by derive_blk_md in EbPickccso.c) */
STUB(av1_get_skip_context)
STUB(av1_get_skip_mode_context)
STUB(svt_aom_allow_intrabc)
STUB(svt_aom_allow_palette)
STUB(svt_aom_blk_geom_mds)
STUB(svt_aom_filter_intra_allowed)
STUB(svt_aom_filter_intra_allowed_bsize)
STUB(svt_aom_get_comp_group_idx_context_enc)
STUB(svt_aom_get_comp_index_context_enc)
STUB(svt_aom_get_comp_reference_type_cdf)
STUB(svt_aom_get_kf_y_mode_ctx)
STUB(svt_aom_get_palette_bsize_ctx)
STUB(svt_aom_get_palette_mode_ctx)
STUB(svt_aom_get_pred_cdf_comp_bwdref_p)
STUB(svt_aom_get_pred_cdf_comp_bwdref_p1)
STUB(svt_aom_get_pred_cdf_comp_ref_p)
STUB(svt_aom_get_pred_cdf_comp_ref_p1)
STUB(svt_aom_get_pred_cdf_comp_ref_p2)
STUB(svt_aom_get_pred_cdf_single_ref_p1)
STUB(svt_aom_get_pred_cdf_single_ref_p2)
STUB(svt_aom_get_pred_cdf_single_ref_p3)
STUB(svt_aom_get_pred_cdf_single_ref_p4)
STUB(svt_aom_get_pred_cdf_single_ref_p5)
STUB(svt_aom_get_pred_cdf_single_ref_p6)
STUB(svt_aom_get_pred_cdf_uni_comp_ref_p)
STUB(svt_aom_get_pred_cdf_uni_comp_ref_p1)
STUB(svt_aom_get_pred_cdf_uni_comp_ref_p2)
STUB(svt_aom_get_pred_context_switchable_interp)
STUB(svt_aom_get_reference_mode_cdf)
STUB(svt_aom_get_wedge_params_bits)
STUB(svt_aom_has_second_ref)
STUB(svt_aom_has_uni_comp_refs)
STUB(svt_aom_have_newmv_in_inter_mode)
STUB(svt_aom_is_interintra_allowed)
STUB(svt_aom_is_interintra_wedge_used)
STUB(svt_aom_is_masked_compound_type)
STUB(svt_aom_is_nontrans_global_motion_ec)
STUB(svt_aom_mode_context_analyzer)
STUB(svt_aom_motion_mode_allowed)
STUB(svt_aom_partition_cdf_length)
STUB(svt_av1_build_nmv_cost_table)
STUB(svt_av1_get_intra_inter_context)
STUB(svt_av1_get_mv_joint)
STUB(svt_av1_superres_upscale_frame)
STUB(svt_get_empty_object)
STUB(svt_get_full_object)
STUB(svt_post_full_object)
STUB(svt_print_alloc_fail_impl)
STUB(svt_release_object)
STUB(svt_system_resource_get_consumer_fifo)
STUB(svt_system_resource_get_producer_fifo)

/* RTCD data pointers that are not on the exercised paths */
uint64_t svt_full_distortion_kernel16_bits_c();
uint64_t svt_spatial_full_distortion_kernel_c();
void    *svt_full_distortion_kernel16_bits  = (void *)svt_full_distortion_kernel16_bits_c;  /* picture_sse_calculations */
void    *svt_spatial_full_distortion_kernel = (void *)svt_spatial_full_distortion_kernel_c;
/* RTCD pointers referenced by parts of EbPictureOperators.c that are never called here */
void svt_enc_msb_un_pack2_d(); /* Common/C_DEFAULT/EbPackUnPack_C.c, compiled in for the hand-off pin */
void *svt_aom_un_pack2d_16_bit_src_mul4 = (void *)svt_enc_msb_un_pack2_d; /* common_dsp_rtcd.c:883 (C only) */
void *svt_compressed_packmsb = 0;
void *svt_full_distortion_kernel32_bits = 0, *svt_full_distortion_kernel_cbf_zero32_bits = 0, *svt_pack2d_16_bit_src_mul4 = 0;
/* RTCD pointers used by derive_ccso_filter (Encoder/Codec/EbPickccso.c): the _c kernels, as common_dsp_rtcd.c:636-658
 * sets them without SIMD.  (Declared without prototypes: this file includes no reference header.) */
void     ccso_filter_block_hbd_with_buf_c();
void     ccso_derive_src_block_c();
uint64_t compute_distortion_block_c();
void *ccso_filter_block_hbd_with_buf = (void *)ccso_filter_block_hbd_with_buf_c;
void *ccso_derive_src_block          = (void *)ccso_derive_src_block_c;
void *compute_distortion_block       = (void *)compute_distortion_block_c;

static void     memcpy_c(void *d, const void *s, size_t n) { memcpy(d, s, n); }
static uint32_t log2f_c(uint32_t x) {
    uint32_t l = 0;
    while (x >>= 1) l++;
    return l;
}
void (*svt_memcpy)(void *, const void *, size_t) = memcpy_c;
uint32_t (*svt_log2f)(uint32_t)                  = log2f_c;

/* loop restoration (Common/Codec/EbRestoration.c, convolve.c): RTCD pointers -> the _c kernels; super-resolution is off */
STUB(svt_av1_upscale_normative_rows)
void svt_apply_selfguided_restoration_c();
void svt_av1_wiener_convolve_add_src_c();
void svt_av1_highbd_wiener_convolve_add_src_c();
void *svt_apply_selfguided_restoration       = (void *)svt_apply_selfguided_restoration_c;
void *svt_av1_wiener_convolve_add_src        = (void *)svt_av1_wiener_convolve_add_src_c;
void *svt_av1_highbd_wiener_convolve_add_src = (void *)svt_av1_highbd_wiener_convolve_add_src_c;
void  svt_log(int level, const char *tag, const char *fmt, ...) { (void)level, (void)tag, (void)fmt; }

/* loop-restoration search (Encoder/Codec/EbRestorationPick.c is part of ref_shim_lrs.c; EbEntropyCoding.c is compiled in
 * for svt_aom_count_primitive_refsubexpfin, EbPsnr.c + C_DEFAULT/variance.c for the SSE of a unit) */
void svt_block_on_mutex(void *m) { (void)m; } /* single-threaded harness */
void svt_release_mutex(void *m) { (void)m; }
void svt_av1_compute_stats_c();
void svt_av1_compute_stats_highbd_c();
void svt_av1_highbd_pixel_proj_error_c();
void svt_av1_lowbd_pixel_proj_error_c();
void svt_av1_selfguided_restoration_c();
void svt_get_proj_subspace_c();
void svt_aom_highbd_8_mse16x16_c();
void svt_aom_mse16x16_c();
void *svt_av1_compute_stats          = (void *)svt_av1_compute_stats_c;
void *svt_av1_compute_stats_highbd   = (void *)svt_av1_compute_stats_highbd_c;
void *svt_av1_highbd_pixel_proj_error = (void *)svt_av1_highbd_pixel_proj_error_c;
void *svt_av1_lowbd_pixel_proj_error = (void *)svt_av1_lowbd_pixel_proj_error_c;
void *svt_av1_selfguided_restoration = (void *)svt_av1_selfguided_restoration_c;
void *svt_get_proj_subspace          = (void *)svt_get_proj_subspace_c;
void *svt_aom_highbd_8_mse16x16      = (void *)svt_aom_highbd_8_mse16x16_c;
void *svt_aom_mse16x16               = (void *)svt_aom_mse16x16_c;
STUB(get_mbmi)
STUB(svt_aom_daala_stop_encode)
STUB(svt_aom_film_grain_params_equal)
STUB(svt_aom_get_block_dimensions)
STUB(svt_aom_get_relative_dist_enc)
STUB(svt_aom_init_mode_probs)
STUB(svt_aom_neighbor_array_unit_mode_write)
STUB(svt_aom_output_bitstream_reset)
STUB(svt_aom_output_bitstream_unit_ctor)
STUB(svt_aom_segmentation_feature_bits)
STUB(svt_aom_segmentation_feature_signed)
STUB(svt_av1_allow_palette)
STUB(svt_av1_default_coef_probs)
STUB(svt_av1_get_nz_map_contexts)
STUB(svt_av1_index_color_cache)
STUB(svt_av1_tokenize_color_map)
STUB(svt_av1_txb_init_levels)
STUB(svt_get_palette_cache_y)
STUB(svt_od_ec_encode_bool_q15)
STUB(svt_od_ec_encode_cdf_q15)
STUB(svt_realloc_output_bitstream_unit)
