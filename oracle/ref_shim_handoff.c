/* ref_shim_handoff.c — TEST INFRASTRUCTURE ONLY.  Drives the reference's own recon hand-off functions (compiled where they
 * lie: Common/Codec/EbMcp.c, Common/C_DEFAULT/EbPackUnPack_C.c) with the plane-level signatures of svtf_oracle_handoff.c,
 * in the order pad_ref_and_set_flags (Encoder/Codec/EbEncDecProcess.c:1589-1760) calls them. */
#include <stddef.h>
#include <stdint.h>

void pad_input_picture(uint8_t *src_pic, uint32_t src_stride, uint32_t original_src_width, uint32_t original_src_height,
                               uint32_t pad_right, uint32_t pad_bottom);
void svt_aom_pad_input_picture_16bit(uint16_t *src_pic, uint32_t src_stride, uint32_t original_src_width, uint32_t original_src_height,
                                     uint32_t pad_right, uint32_t pad_bottom);
void svt_aom_generate_padding(uint8_t *src_pic, uint32_t src_stride, uint32_t original_src_width, uint32_t original_src_height,
                              uint32_t padding_width, uint32_t padding_height);
void svt_aom_generate_padding16_bit(uint16_t *src_pic, uint32_t src_stride, uint32_t original_src_width, uint32_t original_src_height,
                                    uint32_t padding_width, uint32_t padding_height);
void svt_aom_un_pack2d(uint16_t *in16_bit_buffer, uint32_t in_stride, uint8_t *out8_bit_buffer, uint32_t out8_stride, uint8_t *outn_bit_buffer,
                       uint32_t outn_stride, uint32_t width, uint32_t height); /* Common/Codec/EbPictureOperators.c:235 */
void svt_convert_8bit_to_16bit_c(uint8_t *src, uint32_t src_stride, uint16_t *dst, uint32_t dst_stride, uint32_t width, uint32_t height);
void svt_convert_16bit_to_8bit_c(uint16_t *src, uint32_t src_stride, uint8_t *dst, uint32_t dst_stride, uint32_t width, uint32_t height);

void ref_pad_plane(void *buf, int32_t stride, int32_t sb, int32_t width, int32_t height, int32_t org_x, int32_t org_y, int32_t pad_right,
                   int32_t pad_bottom) {
    if (sb == 1) {
        uint8_t *b = (uint8_t *)buf;
        pad_input_picture(b + org_x + (size_t)org_y * stride, stride, width - pad_right, height - pad_bottom, pad_right, pad_bottom);
        svt_aom_generate_padding(b, stride, width, height, org_x, org_y);
    } else {
        uint16_t *b = (uint16_t *)buf;
        svt_aom_pad_input_picture_16bit(b + org_x + (size_t)org_y * stride, stride, width - pad_right, height - pad_bottom, pad_right, pad_bottom);
        svt_aom_generate_padding16_bit(b, stride, width, height, org_x, org_y);
    }
}
void ref_unpack2d(const uint16_t *in, int32_t in_stride, uint8_t *out8, int32_t stride8, uint8_t *outn, int32_t striden, int32_t width,
                  int32_t height) {
    svt_aom_un_pack2d((uint16_t *)in, in_stride, out8, stride8, outn, striden, width, height);
}
void ref_convert_8_to_16(const uint8_t *src, int32_t src_stride, uint16_t *dst, int32_t dst_stride, int32_t width, int32_t height) {
    svt_convert_8bit_to_16bit_c((uint8_t *)src, src_stride, dst, dst_stride, width, height);
}
void ref_convert_16_to_8(const uint16_t *src, int32_t src_stride, uint8_t *dst, int32_t dst_stride, int32_t width, int32_t height) {
    svt_convert_16bit_to_8bit_c((uint16_t *)src, src_stride, dst, dst_stride, width, height);
}

/* ---- super-resolution upscale of one plane through the reference's own svt_av1_upscale_normative_rows
 * (Common/Codec/EbSuperRes.c, compiled where it lies) ---- */
#include "EbInterPrediction.h"
#include "EbSuperRes.h"
#include "Av1Common.h"
__attribute__((visibility("default"))) const int16_t *ref_superres_filters(void) { return &av1_resize_filter_normative[0][0]; }
__attribute__((visibility("default"))) int ref_superres_plane(const void *src, int sst, void *dst, int dstst, int sb, int bd, int rows, int src_cw,
                                                              int ds_frame_w, int up_frame_w, int denom, int ss, int mi_cols, int tile_cols,
                                                              const int32_t *tile_col_start_mi) {
    Av1Common *cm = calloc(1, sizeof(*cm));
    if (!cm)
        return -1;
    cm->frm_size.frame_width = (uint16_t)ds_frame_w, cm->frm_size.superres_upscaled_width = (uint16_t)up_frame_w;
    cm->frm_size.superres_denominator = (uint8_t)denom;
    cm->mi_cols                       = mi_cols;
    cm->tiles_info.tile_cols          = (uint8_t)tile_cols;
    for (int j = 0; j <= tile_cols; j++) cm->tiles_info.tile_col_start_mi[j] = (uint16_t)tile_col_start_mi[j];
    /* the function pads 5 columns left and right of the coded plane in place: give it room */
    const int stride = src_cw + 32;
    uint8_t  *buf    = calloc((size_t)stride * rows, sb);
    for (int y = 0; y < rows; y++) memcpy(buf + ((size_t)y * stride + 16) * sb, (const uint8_t *)src + (size_t)y * sst * sb, (size_t)src_cw * sb);
    svt_av1_upscale_normative_rows(cm, buf + 16 * sb, stride, (uint8_t *)dst, dstst, rows, ss, bd, sb == 2);
    free(buf), free(cm);
    return 0;
}
