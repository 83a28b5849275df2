/*
 * ref_stubs_dec.c — link-time stand-ins for oracle/_ref/libsvtrefdec.so: the RTCD pointers the decoder's filters call,
 * bound to the reference's own _c kernels (what common_dsp_rtcd.c selects without SIMD), the decoder's memory-map
 * globals, and the picture-padding helpers EbDecUtils.c references but the filter calls never reach.
 * Includes no reference header so the prototype-less declarations do not clash.  TEST INFRASTRUCTURE ONLY.
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define STUB(name)                                                       \
    void name(void) {                                                    \
        fprintf(stderr, "ref_stubs_dec: unexpected call to " #name "\n"); \
        abort();                                                         \
    }
STUB(generate_padding_b)
STUB(generate_padding_l)
STUB(generate_padding_l_hbd)
STUB(generate_padding_r)
STUB(generate_padding_r_hbd)
STUB(generate_padding_t)
STUB(svt_aom_realloc_frame_buffer)
STUB(svt_aom_yv12_copy_y_c)
STUB(svt_aom_yv12_copy_u_c)
STUB(svt_aom_yv12_copy_v_c)
STUB(svt_av1_upscale_normative_rows)
STUB(svt_print_alloc_fail_impl)

#define RTCD(name)   \
    void name##_c(); \
    void *name = (void *)name##_c;
RTCD(svt_aom_lpf_horizontal_4)
RTCD(svt_aom_lpf_horizontal_6)
RTCD(svt_aom_lpf_horizontal_8)
RTCD(svt_aom_lpf_horizontal_14)
RTCD(svt_aom_lpf_vertical_4)
RTCD(svt_aom_lpf_vertical_6)
RTCD(svt_aom_lpf_vertical_8)
RTCD(svt_aom_lpf_vertical_14)
RTCD(svt_aom_highbd_lpf_horizontal_4)
RTCD(svt_aom_highbd_lpf_horizontal_6)
RTCD(svt_aom_highbd_lpf_horizontal_8)
RTCD(svt_aom_highbd_lpf_horizontal_14)
RTCD(svt_aom_highbd_lpf_vertical_4)
RTCD(svt_aom_highbd_lpf_vertical_6)
RTCD(svt_aom_highbd_lpf_vertical_8)
RTCD(svt_aom_highbd_lpf_vertical_14)
RTCD(svt_aom_cdef_find_dir)
RTCD(svt_aom_cdef_find_dir_dual)
RTCD(svt_cdef_filter_block)
RTCD(svt_aom_copy_rect8_8bit_to_16bit)
RTCD(ccso_filter_block_hbd_wo_buf)
RTCD(svt_apply_selfguided_restoration)
RTCD(svt_av1_wiener_convolve_add_src)
RTCD(svt_av1_highbd_wiener_convolve_add_src)
void svt_log(int level, const char *tag, const char *fmt, ...) { (void)level, (void)tag, (void)fmt; }

static void memcpy_c(void *d, const void *s, size_t n) { memcpy(d, s, n); }
void (*svt_memcpy)(void *, const void *, size_t) = memcpy_c;

void    *svt_dec_memory_map       = 0;
uint32_t svt_dec_memory_map_index = 0;
uint64_t svt_dec_total_lib_memory = 0;
uint32_t svt_dec_lib_malloc_count = 0;
