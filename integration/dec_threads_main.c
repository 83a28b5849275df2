/*
 * dec_threads_main.c — a minimal IVF decoder over the reference decoder's PUBLIC API (Source/API/EbSvtAv1Dec.h), used only to
 * reach the decoder's multi-threaded path: the reference SvtAv1DecApp ignores `-threads` ("Disabled for now",
 * App/DecApp/EbDecParamParser.c:56-60, :99), so EbSvtAv1DecConfiguration.threads > 1 — and with it
 * svt_aom_dec_av1_loop_filter_frame_mt / svt_cdef_frame_mt / svt_aom_dec_av1_loop_restoration_filter_frame_mt
 * (Decoder/Codec/EbDecProcess.c:668 / :810 / :1022) — is reachable from library users only.
 * Built twice by integration/build.sh, against the unmodified and the patched libSvtAv1Dec.a:
 *     dec_threads_{ref,b200} <in.ivf> <out.yuv> <threads>
 * Follows what App/DecApp/EbDecAppMain.c:160-260 does with the API (init_handle, set_parameter, init, dec_frame,
 * get_picture, write the visible planes); no code is taken from it.
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "EbSvtAv1Dec.h"

static uint32_t le32(const uint8_t *p) { return p[0] | (p[1] << 8) | (p[2] << 16) | ((uint32_t)p[3] << 24); }

static void write_plane(FILE *f, const uint8_t *p, int stride, int w, int h, int bytes) {
    for (int y = 0; y < h; y++) fwrite(p + (size_t)y * stride * bytes, bytes, w, f);
}

int main(int argc, char **argv) {
    if (argc != 4) {
        fprintf(stderr, "usage: %s <in.ivf> <out.yuv> <threads>\n", argv[0]);
        return 2;
    }
    FILE *in = fopen(argv[1], "rb"), *out = fopen(argv[2], "wb");
    if (!in || !out) {
        perror("open");
        return 2;
    }
    uint8_t hdr[32];
    if (fread(hdr, 1, 32, in) != 32 || memcmp(hdr, "DKIF", 4)) {
        fprintf(stderr, "not an IVF file\n");
        return 2;
    }
    const int width = hdr[12] | (hdr[13] << 8), height = hdr[14] | (hdr[15] << 8);

    EbSvtAv1DecConfiguration cfg;
    EbComponentType         *h = NULL;
    if (svt_av1_dec_init_handle(&h, NULL, &cfg) != EB_ErrorNone)
        return 3;
    cfg.max_picture_width  = width;
    cfg.max_picture_height = height;
    cfg.threads            = (uint32_t)atoi(argv[3]);
    if (svt_av1_dec_set_parameter(h, &cfg) != EB_ErrorNone || svt_av1_dec_init(h) != EB_ErrorNone)
        return 3;

    const int          w2 = (width + 1) & ~1, h2 = (height + 1) & ~1;
    EbSvtIOFormat      img;
    EbBufferHeaderType buf;
    memset(&img, 0, sizeof(img));
    memset(&buf, 0, sizeof(buf));
    img.luma = (uint8_t *)malloc((size_t)w2 * h2 * 2), img.cb = (uint8_t *)malloc((size_t)w2 * h2 / 2), img.cr = (uint8_t *)malloc((size_t)w2 * h2 / 2);
    img.y_stride = w2, img.cb_stride = img.cr_stride = width / 2;
    img.width = width, img.height = height, img.bit_depth = cfg.max_bit_depth;
    buf.p_buffer = (uint8_t *)&img;

    EbAV1StreamInfo si;
    EbAV1FrameInfo  fi;
    uint8_t        *data = NULL;
    size_t          cap  = 0;
    int             frames = 0, err = 0;
    uint8_t         fh[12];
    while (fread(fh, 1, 12, in) == 12) {
        const size_t n = le32(fh);
        if (n > cap)
            data = (uint8_t *)realloc(data, cap = n);
        if (fread(data, 1, n, in) != n)
            break;
        err |= svt_av1_dec_frame(h, data, n, 0) != EB_ErrorNone;
        if (svt_av1_dec_get_picture(h, &buf, &si, &fi) != EB_DecNoOutputPicture) {
            const int b = img.bit_depth == EB_EIGHT_BIT ? 1 : 2;
            write_plane(out, img.luma, img.y_stride, img.width, img.height, b);
            write_plane(out, img.cb, img.cb_stride, (img.width + 1) >> 1, (img.height + 1) >> 1, b);
            write_plane(out, img.cr, img.cr_stride, (img.width + 1) >> 1, (img.height + 1) >> 1, b);
            frames++;
        }
    }
    fclose(out);
    fprintf(stderr, "dec_threads: %d pictures, threads=%u%s\n", frames, cfg.threads, err ? " (decode errors)" : "");
    /* threads > 1: the unmodified reference library double-frees in svt_av1_dec_deinit (ASan: a block dec_system_resource_init
     * already freed; glibc aborts with "malloc_consolidate(): invalid chunk size") — the output is complete and closed by now,
     * so the multi-threaded runs leave the teardown to process exit, for both builds alike */
    if (cfg.threads <= 1) {
        svt_av1_dec_deinit(h);
        svt_av1_dec_deinit_handle(h);
    }
    fflush(NULL);
    return err ? 1 : 0;
}
