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

#define STUB(name)                                                   \
    void name(void) {                                                \
        fprintf(stderr, "ref_stubs: unexpected call to " #name "\n"); \
        abort();                                                     \
    }
STUB(set_unscaled_input_16bit)
STUB(svt_aom_ac_quant_qtx)
STUB(svt_aom_assert_err)
STUB(svt_aom_downscaled_source_buffer_desc_ctor)
STUB(svt_aom_get_syntax_rate_from_cdf)
STUB(svt_aom_resize_frame)
STUB(svt_av1_loop_restoration_save_boundary_lines)
STUB(svt_av1_superres_upscale_frame)
STUB(svt_block_on_mutex)
STUB(svt_get_empty_object)
STUB(svt_get_full_object)
STUB(svt_post_full_object)
STUB(svt_print_alloc_fail_impl)
STUB(svt_recon_picture_buffer_desc_ctor)
STUB(svt_release_mutex)
STUB(svt_release_object)
STUB(svt_system_resource_get_consumer_fifo)
STUB(svt_system_resource_get_producer_fifo)

/* RTCD data pointers that are not on the exercised paths */
void *svt_full_distortion_kernel16_bits  = 0;
void *svt_spatial_full_distortion_kernel = 0;
void *ccso_filter_block_hbd_with_buf     = 0;
void *ccso_derive_src_block              = 0;
void *compute_distortion_block           = 0;

static void     memcpy_c(void *d, const void *s, size_t n) { memcpy(d, s, n); }
static uint32_t log2f_c(uint32_t x) {
    uint32_t l = 0;
    while (x >>= 1) l++;
    return l;
}
void (*svt_memcpy)(void *, const void *, size_t) = memcpy_c;
uint32_t (*svt_log2f)(uint32_t)                  = log2f_c;
