/* svtf_oracle_handoff.c — TEST INFRASTRUCTURE ONLY (see svtf_oracle.c): CPU restatement of the recon hand-off.
 *
 * Follows, step by step and in the reference's order (paths relative to Source/Lib/ of the reference):
 *   svt_aom_pad_input_picture_16bit   Common/Codec/EbMcp.c:296-341   right columns then bottom rows replicate the last valid
 *   svt_aom_generate_padding[16_bit]  Common/Codec/EbMcp.c:95-140, 201-243   left/right borders per row, then whole rows up/down
 *   svt_enc_msb_un_pack2_d            Common/C_DEFAULT/EbPackUnPack_C.c:202-218   out8 = v >> 2, outn = (uint8_t)(v << 6)
 *   svt_convert_8bit_to_16bit_c / svt_convert_16bit_to_8bit_c   Common/C_DEFAULT/EbPackUnPack_C.c:270-283
 * as used by pad_ref_and_set_flags (Encoder/Codec/EbEncDecProcess.c:1589-1760) and svt_convert_pic_8bit_to_16bit
 * (Encoder/Codec/EbRestProcess.c:235).  Pinned against those functions themselves (oracle/ref_shim_handoff.c,
 * tests/test_handoff_oracle.py). */
#include <stdint.h>
#include <string.h>

#define ORC_API __attribute__((visibility("default")))

/* buf: first sample of a padded plane allocation (stride samples of `sb` bytes); the picture proper starts at
 * (org_x, org_y) and already holds valid_w x valid_h samples there.  Fills everything else of the
 * (width + 2*org_x) x (height + 2*org_y) rectangle. */
ORC_API void svtf_orc_pad_plane(void *buf, int32_t stride, int32_t sb, int32_t width, int32_t height, int32_t org_x, int32_t org_y,
                                int32_t pad_right, int32_t pad_bottom) {
    const int32_t vw = width - pad_right, vh = height - pad_bottom;
#define AT(T, y, x) (((T *)buf)[(size_t)(y) * stride + (x)])
#define PAD(T)                                                                                                   \
    do {                                                                                                         \
        /* pad_input_picture: right columns of the valid rows, then bottom rows (each a copy of the last row,  \
           right padding included) */                                                                            \
        for (int32_t y = 0; y < vh; y++)                                                                         \
            for (int32_t x = 0; x < pad_right; x++) AT(T, org_y + y, org_x + vw + x) = AT(T, org_y + y, org_x + vw - 1); \
        for (int32_t y = 0; y < pad_bottom; y++)                                                                 \
            for (int32_t x = 0; x < width; x++) AT(T, org_y + vh + y, org_x + x) = AT(T, org_y + vh - 1, org_x + x);      \
        /* generate_padding: horizontal borders of every picture row, then rows up and down */                   \
        for (int32_t y = 0; y < height; y++)                                                                     \
            for (int32_t x = 0; x < org_x; x++) {                                                                \
                AT(T, org_y + y, x)                 = AT(T, org_y + y, org_x);                                   \
                AT(T, org_y + y, org_x + width + x) = AT(T, org_y + y, org_x + width - 1);                       \
            }                                                                                                    \
        for (int32_t y = 0; y < org_y; y++)                                                                      \
            for (int32_t x = 0; x < width + 2 * org_x; x++) {                                                    \
                AT(T, y, x)                   = AT(T, org_y, x);                                                 \
                AT(T, org_y + height + y, x) = AT(T, org_y + height - 1, x);                                     \
            }                                                                                                    \
    } while (0)
    if (sb == 1)
        PAD(uint8_t);
    else
        PAD(uint16_t);
#undef PAD
#undef AT
}

ORC_API void svtf_orc_unpack2d(const uint16_t *in, int32_t in_stride, uint8_t *out8, int32_t stride8, uint8_t *outn, int32_t striden,
                               int32_t width, int32_t height) {
    for (int32_t y = 0; y < height; y++)
        for (int32_t x = 0; x < width; x++) {
            const uint16_t v = in[(size_t)y * in_stride + x];
            out8[(size_t)y * stride8 + x] = (uint8_t)(v >> 2);
            if (outn)
                outn[(size_t)y * striden + x] = (uint8_t)(v << 6);
        }
}

ORC_API void svtf_orc_convert_8_to_16(const uint8_t *src, int32_t src_stride, uint16_t *dst, int32_t dst_stride, int32_t width, int32_t height) {
    for (int32_t y = 0; y < height; y++)
        for (int32_t x = 0; x < width; x++) dst[(size_t)y * dst_stride + x] = src[(size_t)y * src_stride + x];
}
ORC_API void svtf_orc_convert_16_to_8(const uint16_t *src, int32_t src_stride, uint8_t *dst, int32_t dst_stride, int32_t width, int32_t height) {
    for (int32_t y = 0; y < height; y++)
        for (int32_t x = 0; x < width; x++) dst[(size_t)y * dst_stride + x] = (uint8_t)src[(size_t)y * src_stride + x];
}

/* ---- super-resolution upscale of one plane: svt_av1_upscale_normative_rows (Common/Codec/EbSuperRes.c:214) with
 * av1_[highbd_]convolve_horiz_rs_c (:60 / :80) and the edge replication of upscale_normative_rect (:102).
 * src: coded plane (src_cw = mi_cols * 4 >> ss columns), ds_w / up_w: visible plane widths before / after, filters:
 * av1_resize_filter_normative[64][8] handed in by the caller (Common/Codec/EbSuperRes.h:22). ---- */
ORC_API void svtf_orc_superres_plane(const void *src, int sst, void *dst, int dstst, int sb, int bd, int rows, int src_cw, int ds_w, int up_w,
                                     int denom, int ss, int mi_cols, int tile_cols, const int32_t *tile_col_start_mi, const int16_t *filters) {
    const int step = ((ds_w << 14) + up_w / 2) / up_w;
    const int err  = up_w * step - (ds_w << 14);
    int       x0   = (int)((unsigned)((-((up_w - ds_w) << 13) + up_w / 2) / up_w + (1 << 7) - err / 2) & 16383u);
    for (int j = 0; j < tile_cols; j++) {
        const int end = tile_col_start_mi[j + 1] < mi_cols ? tile_col_start_mi[j + 1] : mi_cols;
        const int ds0 = tile_col_start_mi[j] << (2 - ss), ds1 = end << (2 - ss);
        const int up0 = ds0 * denom / 8, up1 = j == tile_cols - 1 ? up_w : ds1 * denom / 8;
        for (int y = 0; y < rows; y++) {
            int xq = x0;
            for (int x = up0; x < up1; x++, xq += step) {
                const int16_t *f = filters + ((xq & 16383) >> 8) * 8;
                int            sum = 0;
                for (int k = 0; k < 8; k++) {
                    int c = ds0 - 4 + (xq >> 14) + k;
                    c     = c < 0 ? 0 : (c > src_cw - 1 ? src_cw - 1 : c);
                    sum += (sb == 1 ? ((const uint8_t *)src)[(size_t)y * sst + c] : ((const uint16_t *)src)[(size_t)y * sst + c]) * f[k];
                }
                int v = (sum + 64) >> 7;
                v     = v < 0 ? 0 : (v > (1 << bd) - 1 ? (1 << bd) - 1 : v);
                if (sb == 1)
                    ((uint8_t *)dst)[(size_t)y * dstst + x] = (uint8_t)v;
                else
                    ((uint16_t *)dst)[(size_t)y * dstst + x] = (uint16_t)v;
            }
        }
        x0 += (up1 - up0) * step - ((ds1 - ds0) << 14);
    }
}
