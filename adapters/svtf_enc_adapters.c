/*
 * svtf_enc_adapters.c — Tier-2 of the drop-in boundary: the reference's OWN frame-level entry points
 * (same names, same signatures) implemented as thin marshalling over the Tier-1 C ABI
 * (include/svtfilt_b200.h).  Compiled against the reference's headers; linked INSTEAD of the bodies in
 *   Encoder/Codec/EbDeblockingFilter.c:624  svt_av1_loop_filter_frame
 *   Encoder/Codec/EbEncCdef.c:284           svt_av1_cdef_frame
 *   Encoder/Codec/EbEncCdef.c:728           finish_cdef_search
 *   Encoder/Codec/EbCdefProcess.c:114       cdef_seg_search      (static there -> svtf_cdef_search_picture here)
 *   Common/Codec/EbCcso.c:626               ccso_frame
 *   Encoder/Codec/EbPickccso.c:785          ccso_search
 *   Encoder/Codec/EbDeblockingFilter.c:1129 svt_av1_pick_filter_level
 * (see INTEGRATION.md for the three-line changes at the call sites).
 *
 * What an adapter does: (1) flatten pcs->mi_grid_base into the packed 4-byte mi words, (2) hand the recon
 * picture's plane pointers/strides to svtf_upload_frame, (3) launch, (4) svtf_download_frame back into the
 * same picture buffers (the reference filters in place).  Errors: the reference functions return void, so a
 * failing Tier-1 call logs through SVT_ERROR-style stderr and aborts — there is no CPU fallback.
 *
 * One GPU context per PictureControlSet, created on first use and cached (a picture's DLF, CDEF search, CDEF
 * apply and CCSO run on the same context, so the picture stays resident between stages:
 * svtf_adapter_set_resident(1) skips the intermediate download/upload pairs).
 */
#include <pthread.h>
#include <time.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "EbPictureControlSet.h"
#include "EbSequenceControlSet.h"
#include "EbEncCdef.h"
#include "EbCdef.h"
#include "EbDeblockingFilter.h"
#include "EbDeblockingCommon.h"
#include "EbCcso.h"
#include "EbPickccso.h"
#include "EbRestoration.h"
#include "EbInterPrediction.h"
#include "EbSuperRes.h"

#include "../include/svtfilt_b200.h"
#include "../include/svtfilt_b200_adapters.h"

void svt_aom_get_recon_pic(PictureControlSet *pcs, EbPictureBufferDesc **recon_ptr, Bool is_highbd);

#define SVTF_MAX_CACHED 64
typedef struct {
    const PictureControlSet *pcs;
    svtf_ctx                *ctx;
    svtf_geom                geom, max_geom;
    int                      resident; /* picture of this pcs currently on the device */
    uint64_t                 launches_seen;
    uint8_t                 *mi_stage;   /* pinned staging of the six side-info byte planes (svtf_host_alloc), reused per picture */
    size_t                   mi_stage_cap;
    int8_t                  *idx_stage;  /* per-filter-block strength indices handed to svtf_cdef_apply_frame */
    size_t                   idx_cap;
    void                    *reg[8];     /* host ranges this slot page-locked (svtf_adapter_set_pin_pictures) */
    int                      nreg;
} CtxSlot;
static CtxSlot         g_slots[SVTF_MAX_CACHED];
static pthread_mutex_t g_lock     = PTHREAD_MUTEX_INITIALIZER;
static int             g_device   = 0;
static int             g_resident = 0;
static int             g_pin      = 0;
static int             g_mi_threads = 8;

void svtf_adapter_set_device(int device) { g_device = device; }
void svtf_adapter_set_resident(int on) { g_resident = on; }
void svtf_adapter_set_pin_pictures(int on) { g_pin = on; }
void svtf_adapter_set_mi_threads(int n) { g_mi_threads = n < 1 ? 1 : (n > 32 ? 32 : n); }

static void die(const char *where, svtf_ctx *ctx) {
    fprintf(stderr, "Svt[error]: svtfilt_b200 adapter: %s failed: %s (no CPU fallback on this path)\n", where,
            svtf_last_error(ctx));
    abort();
}
/* checked allocations: an adapter has no way to report failure (the replaced functions return void) */
static void *xmalloc(size_t n, svtf_ctx *ctx) {
    void *p = malloc(n ? n : 1);
    if (!p)
        die("malloc", ctx);
    return p;
}
static void *xcalloc(size_t n, size_t m, svtf_ctx *ctx) {
    void *p = calloc(n ? n : 1, m ? m : 1);
    if (!p)
        die("calloc", ctx);
    return p;
}
#define CHECK(call, ctx)       \
    do {                       \
        if ((call) != SVTF_OK) \
            die(#call, ctx);   \
    } while (0)

/* Call counters, printed at exit when SVTF_ADAPTER_STATS is set in the environment: lets a test of the real application prove
 * which stages ran on the GPU (`name=calls` pairs on one stderr line), and how many kernel launches they issued. */
enum { ST_DLF, ST_DLF_SB_LAST, ST_LF_SEARCH, ST_CDEF_SEARCH, ST_CDEF_PICK, ST_CDEF_PICK_REFFS, ST_CDEF_APPLY, ST_CCSO, ST_CCSO_SEARCH, ST_LR_APPLY, ST_LR_SEARCH,
       ST_SUPERRES, ST_PAD_REF, ST_N };
static const char *const g_stat_name[ST_N] = {"loop_filter_frame", "loop_filter_sb_last", "pick_filter_level_full", "cdef_search_picture", "finish_cdef_search",
                                              "finish_cdef_search_ref_fs", "cdef_frame", "ccso_frame", "ccso_search", "loop_restoration_filter_frame",
                                              "restoration_search", "superres_upscale_frame", "pad_ref_pictures"};
static unsigned long   g_stat[ST_N];
static unsigned long   g_stat_launches;
/* host wall time spent in the phases of a picture's round trip (ns, summed over threads), same switch: where an end-to-end run waits */
enum { TP_UPLOAD_CALL, TP_GATHER, TP_SET_MI_CALL, TP_DOWNLOAD, TP_N };
static const char *const g_tp_name[TP_N] = {"upload_call", "mi_gather", "set_mi_call", "download_wait"};
static unsigned long long g_tp[TP_N], g_tp_pictures;
static int                g_tp_on = -1;
static inline unsigned long long tp_now(void) {
    if (g_tp_on < 0)
        g_tp_on = getenv("SVTF_ADAPTER_STATS") != NULL;
    if (!g_tp_on)
        return 0;
    struct timespec t;
    clock_gettime(CLOCK_MONOTONIC, &t);
    return (unsigned long long)t.tv_sec * 1000000000ull + (unsigned long long)t.tv_nsec;
}
static inline void tp_add(int which, unsigned long long t0) {
    if (g_tp_on > 0)
        __sync_fetch_and_add(&g_tp[which], tp_now() - t0);
}
static pthread_once_t  g_stat_once = PTHREAD_ONCE_INIT;
static void            stat_print(void) {
    if (!getenv("SVTF_ADAPTER_STATS"))
        return;
    pthread_mutex_lock(&g_lock);
    for (int i = 0; i < SVTF_MAX_CACHED; i++)
        if (g_slots[i].ctx)
            g_stat_launches += (unsigned long)svtf_launch_count(g_slots[i].ctx) - g_slots[i].launches_seen, g_slots[i].launches_seen = svtf_launch_count(g_slots[i].ctx);
    pthread_mutex_unlock(&g_lock);
    fprintf(stderr, "svtfilt_b200 adapter stats:");
    for (int i = 0; i < ST_N; i++) fprintf(stderr, " %s=%lu", g_stat_name[i], g_stat[i]);
    fprintf(stderr, " gpu_launches=%lu\n", g_stat_launches);
    if (g_tp_pictures) {
        fprintf(stderr, "svtfilt_b200 adapter host time per uploaded picture (ms):");
        for (int i = 0; i < TP_N; i++) fprintf(stderr, " %s=%.3f", g_tp_name[i], (double)g_tp[i] / (double)g_tp_pictures * 1e-6);
        fprintf(stderr, " pictures=%llu\n", g_tp_pictures);
    }
}
static void stat_init(void) { atexit(stat_print); }
static void stat_hit(int which) {
    pthread_once(&g_stat_once, stat_init);
    __sync_fetch_and_add(&g_stat[which], 1);
}

static svtf_geom geom_of(const PictureControlSet *pcs) {
    const SequenceControlSet *scs = pcs->scs;
    const Av1Common          *cm  = pcs->ppcs->av1_cm;
    svtf_geom                 g;
    /* unpadded size: the width/height svt_av1_setup_dst_planes gives the deblock (EbDeblockingFilter.c:99-101) */
    g.width        = scs->max_input_luma_width - scs->max_input_pad_right;
    g.height       = scs->max_input_luma_height - scs->max_input_pad_bottom;
    /* a frame coded with super-resolution is narrower than the input, one coded with reference scaling (resize_mode) narrower and
     * lower: the picture the stages work on is ppcs->enhanced_pic (what svt_av1_filter_block_plane_vert/_horz cut their ranges to,
     * EbDeblockingFilter.c:305-322), and cm->mi_cols / mi_rows match it */
    if (pcs->ppcs->frame_superres_enabled || pcs->ppcs->frame_resize_enabled)
        g.width = pcs->ppcs->enhanced_pic->width, g.height = pcs->ppcs->enhanced_pic->height;
    g.mi_cols      = cm->mi_cols;
    g.mi_rows      = cm->mi_rows;
    g.bit_depth    = scs->static_config.encoder_bit_depth;
    g.sample_bytes = scs->is_16bit_pipeline ? 2 : 1;
    return g;
}

/* Frees everything a slot owns (context, staging, page-locks).  g_lock held. */
static void slot_teardown(CtxSlot *s) {
    if (s->ctx) {
        g_stat_launches += (unsigned long)svtf_launch_count(s->ctx) - s->launches_seen;
        svtf_ctx_destroy(s->ctx);
    }
    if (s->mi_stage)
        svtf_host_free(s->mi_stage);
    free(s->idx_stage);
    for (int i = 0; i < s->nreg; i++) svtf_host_unregister(s->reg[i]);
    memset(s, 0, sizeof(*s));
}

/* Geometry a context must hold for this picture (room for the super-resolution upscale included). */
static svtf_geom max_geom_of(const PictureControlSet *pcs, const svtf_geom *g) {
    svtf_geom        mg = *g;
    const Av1Common *cm = pcs->ppcs->av1_cm;
    if (cm->frm_size.superres_denominator > SCALE_NUMERATOR && cm->frm_size.superres_upscaled_width > mg.width) {
        mg.width   = cm->frm_size.superres_upscaled_width;
        mg.mi_cols = 2 * ((mg.width + 7) >> 3);
    }
    return mg;
}

/* The GPU context of a picture.  Contexts are cached per pcs pointer — the pipeline recycles a small pool of PictureControlSets, so
 * a recycled pcs finds its surfaces allocated — but the geometry is recomputed on EVERY entry: a recycled pcs may carry another
 * super-resolution / resize width, and then the context is rebuilt and nothing stale (resident picture included) survives. */
static CtxSlot *slot_for_at(const PictureControlSet *pcs, int after_superres) {
    svtf_geom       g  = geom_of(pcs);
    const svtf_geom mg = max_geom_of(pcs, &g);
    /* the stages behind the super-resolution upscale (EbCdefProcess.c:675-677: restoration search and apply, reference padding) see
     * the upscaled picture; cm->frm_size.frame_width and cm->mi_cols keep describing the coded (narrow) one */
    if (after_superres)
        g.width = mg.width, g.mi_cols = mg.mi_cols;
    pthread_mutex_lock(&g_lock);
    CtxSlot *free_slot = NULL, *s = NULL;
    for (int i = 0; i < SVTF_MAX_CACHED; i++) {
        if (g_slots[i].pcs == pcs) {
            s = &g_slots[i];
            break;
        }
        if (!g_slots[i].pcs && !free_slot)
            free_slot = &g_slots[i];
    }
    if (s && (mg.mi_cols > s->max_geom.mi_cols || mg.mi_rows > s->max_geom.mi_rows || mg.sample_bytes != s->max_geom.sample_bytes)) {
        slot_teardown(s); /* the picture no longer fits: rebuild */
        free_slot = s, s = NULL;
    }
    if (!s) {
        if (!free_slot) {
            fprintf(stderr, "Svt[error]: svtfilt_b200 adapter: more than %d pictures in flight\n", SVTF_MAX_CACHED);
            abort();
        }
        s           = free_slot;
        s->pcs      = pcs;
        s->max_geom = mg;
        s->resident = 0;
        if (svtf_ctx_create(g_device, &mg, &s->ctx) != SVTF_OK)
            die("svtf_ctx_create", NULL);
    }
    if (memcmp(&s->geom, &g, sizeof(g))) /* another picture size on the same pcs: whatever is on the device is not this picture */
        s->resident = 0;
    s->geom = g;
    pthread_mutex_unlock(&g_lock);
    return s;
}

static CtxSlot *slot_for(const PictureControlSet *pcs) { return slot_for_at(pcs, 0); }

/* Release the cached context of a picture (call where the reference recycles the pcs). */
void svtf_adapter_release(const PictureControlSet *pcs) {
    pthread_mutex_lock(&g_lock);
    for (int i = 0; i < SVTF_MAX_CACHED; i++)
        if (g_slots[i].pcs == pcs)
            slot_teardown(&g_slots[i]);
    pthread_mutex_unlock(&g_lock);
}

/* mi_grid_base -> six byte planes (the fields set_lpf_parameters and svt_sb_compute_cdef_list read: EbDeblockingFilter.c:187-260,
 * EbEncCdef.c:267-270), gathered by a few threads into the slot's PINNED staging block and sent with one asynchronous DMA per plane;
 * the words are packed on the device (svtf_set_mi_soa_enc).  No allocation and no synchronisation per picture: the staging block is
 * reused, and it is not rewritten before the picture's final download has drained the stream.  (An encoder that keeps these six
 * planes next to mi_grid_base — written where EncDec updates the mode info of a block — passes them directly and this pass disappears.) */
typedef struct {
    const PictureControlSet *pcs;
    uint8_t                 *soa;
    int                      rows, cols, r0, r1;
} MiGatherJob;
static void *mi_gather_rows(void *arg) {
    const MiGatherJob *j = (const MiGatherJob *)arg;
    const size_t       n = (size_t)j->rows * j->cols;
    uint8_t *bsize = j->soa, *txd = j->soa + n, *skip = j->soa + 2 * n, *ref0 = j->soa + 3 * n, *mode = j->soa + 4 * n, *seg = j->soa + 5 * n;
    for (int r = j->r0; r < j->r1; r++) {
        ModeInfo *const *row = j->pcs->mi_grid_base + (size_t)r * j->pcs->mi_stride;
        const size_t     o   = (size_t)r * j->cols;
        for (int c = 0; c < j->cols; c++) {
            const BlockModeInfoEnc *b = &row[c]->mbmi.block_mi;
            bsize[o + c] = (uint8_t)b->bsize, txd[o + c] = b->tx_depth, skip[o + c] = b->skip;
            ref0[o + c] = (uint8_t)b->ref_frame[0], mode[o + c] = (uint8_t)b->mode, seg[o + c] = (uint8_t)b->segment_id;
        }
    }
    return NULL;
}
static void send_mi(CtxSlot *s, const PictureControlSet *pcs) {
    const int    rows = s->geom.mi_rows, cols = s->geom.mi_cols;
    const size_t n    = (size_t)rows * cols;
    if (s->mi_stage_cap < 6 * n) {
        if (s->mi_stage)
            svtf_host_free(s->mi_stage);
        s->mi_stage     = (uint8_t *)svtf_host_alloc(6 * n);
        s->mi_stage_cap = s->mi_stage ? 6 * n : 0;
        if (!s->mi_stage)
            die("svtf_host_alloc(mi staging)", s->ctx);
    }
    uint8_t *soa = s->mi_stage;
    int      nt  = g_mi_threads;
    if (rows < 64 * nt)
        nt = rows / 64 > 0 ? rows / 64 : 1; /* small pictures: a thread per >= 64 rows */
    MiGatherJob job[32];
    pthread_t   th[32];
    for (int t = 0; t < nt; t++) {
        job[t].pcs = pcs, job[t].soa = soa, job[t].rows = rows, job[t].cols = cols;
        job[t].r0 = (int)((long)rows * t / nt), job[t].r1 = (int)((long)rows * (t + 1) / nt);
    }
    unsigned long long t0 = tp_now();
    int created[32] = {0};
    for (int t = 1; t < nt; t++) created[t] = pthread_create(&th[t], NULL, mi_gather_rows, &job[t]) == 0;
    mi_gather_rows(&job[0]);
    for (int t = 1; t < nt; t++) {
        if (created[t])
            pthread_join(th[t], NULL);
        else
            mi_gather_rows(&job[t]); /* could not start the thread: do its rows here */
    }
    tp_add(TP_GATHER, t0);
    t0 = tp_now();
    CHECK(svtf_set_mi_soa_enc(s->ctx, soa, soa + n, soa + 2 * n, soa + 3 * n, soa + 4 * n, soa + 5 * n, cols), s->ctx);
    tp_add(TP_SET_MI_CALL, t0);
}

/* svtf_adapter_set_pin_pictures(1): page-lock the picture's plane allocations in place on first sight (the recon pictures of the
 * encoder live in a long-lived pool), so that the upload and the download are direct DMA.  Unlocked when the slot is torn down. */
static void pin_picture(CtxSlot *s, const EbPictureBufferDesc *pic, int sb) {
    if (!g_pin)
        return;
    const int vis = pic->height > s->geom.mi_rows * 4 ? pic->height : s->geom.mi_rows * 4; /* the copies cover the coded rows */
    const int ph  = vis + 2 * pic->org_y; /* rows of the allocation; chroma: half */
    void     *base[3] = {pic->buffer_y, pic->buffer_cb, pic->buffer_cr};
    size_t    len[3]  = {(size_t)pic->stride_y * ph * sb, (size_t)pic->stride_cb * ((ph + 1) >> 1) * sb, (size_t)pic->stride_cr * ((ph + 1) >> 1) * sb};
    for (int p = 0; p < 3; p++) {
        int seen = 0;
        for (int i = 0; i < s->nreg; i++) seen |= s->reg[i] == base[p];
        if (seen || !base[p] || s->nreg >= 8)
            continue;
        if (svtf_host_register(base[p], len[p]) == SVTF_OK)
            s->reg[s->nreg++] = base[p];
    }
}

static void plane_ptrs(const EbPictureBufferDesc *pic, int sb, void *p[3], int32_t st[3]) {
    /* visible origin; chroma origin as the reference computes it (EbEncCdef.c:296-299, EbDeblockingFilter.c:116) */
    p[0]  = pic->buffer_y + ((size_t)(pic->org_x + pic->org_y * pic->stride_y) * sb);
    p[1]  = pic->buffer_cb + ((size_t)((pic->org_x + pic->org_y * pic->stride_cb) >> 1) * sb);
    p[2]  = pic->buffer_cr + ((size_t)((pic->org_x + pic->org_y * pic->stride_cr) >> 1) * sb);
    st[0] = pic->stride_y, st[1] = pic->stride_cb, st[2] = pic->stride_cr;
}

static void begin_picture(CtxSlot *s, const PictureControlSet *pcs, const EbPictureBufferDesc *pic) {
    if (g_resident && s->resident)
        return;
    void   *p[3];
    int32_t st[3];
    plane_ptrs(pic, s->geom.sample_bytes, p, st);
    pin_picture(s, pic, s->geom.sample_bytes);
    unsigned long long t0 = tp_now();
    CHECK(svtf_frame_begin(s->ctx, &s->geom), s->ctx);
    CHECK(svtf_upload_frame(s->ctx, p[0], st[0], p[1], st[1], p[2], st[2]), s->ctx); /* asynchronous from page-locked memory */
    tp_add(TP_UPLOAD_CALL, t0);
    send_mi(s, pcs); /* gathers while the picture is on its way */
    if (g_tp_on > 0)
        __sync_fetch_and_add(&g_tp_pictures, 1);
    s->resident = 1;
}

static void end_picture(CtxSlot *s, const EbPictureBufferDesc *pic) {
    void   *p[3];
    int32_t st[3];
    plane_ptrs(pic, s->geom.sample_bytes, p, st);
    const unsigned long long t0 = tp_now();
    CHECK(svtf_download_frame(s->ctx, p[0], st[0], p[1], st[1], p[2], st[2]), s->ctx);
    tp_add(TP_DOWNLOAD, t0);
}

void svtf_adapter_fetch(const PictureControlSet *pcs, void *pic) {
    /* whatever is resident: after a super-resolution upscale that is the wide picture */
    CtxSlot *s = NULL;
    pthread_mutex_lock(&g_lock);
    for (int i = 0; i < SVTF_MAX_CACHED; i++)
        if (g_slots[i].pcs == pcs)
            s = &g_slots[i];
    pthread_mutex_unlock(&g_lock);
    if (!s)
        s = slot_for(pcs);
    if (!s->resident) {
        fprintf(stderr, "Svt[error]: svtfilt_b200 adapter: svtf_adapter_fetch: no resident picture for this pcs\n");
        abort();
    }
    end_picture(s, (const EbPictureBufferDesc *)pic);
}

/* ---------------------------------------------------------------------------------------------------- */
static void lf_of(const PictureControlSet *pcs, svtf_lf_params *out) {
    const FrameHeader       *fh = &pcs->ppcs->frm_hdr;
    const struct LoopFilter *l  = &fh->loop_filter_params;
    svtf_lf_params           lf;
    memset(&lf, 0, sizeof(lf));
    lf.filter_level[0] = l->filter_level[0], lf.filter_level[1] = l->filter_level[1];
    lf.filter_level_u = l->filter_level_u, lf.filter_level_v = l->filter_level_v;
    lf.sharpness_level        = l->sharpness_level;
    lf.mode_ref_delta_enabled = l->mode_ref_delta_enabled;
    memcpy(lf.ref_deltas, l->ref_deltas, sizeof(lf.ref_deltas));
    memcpy(lf.mode_deltas, l->mode_deltas, sizeof(lf.mode_deltas));
    lf.segmentation_enabled = fh->segmentation_params.segmentation_enabled;
    for (int i = 0; i < 8; i++)
        for (int f = 0; f < 8; f++) {
            lf.seg_feature_enabled[i][f] = fh->segmentation_params.feature_enabled[i][f];
            lf.seg_feature_data[i][f]    = fh->segmentation_params.feature_data[i][f];
        }
    lf.delta_lf_present = fh->delta_lf_params.delta_lf_present;
    lf.delta_lf_multi   = fh->delta_lf_params.delta_lf_multi;
    for (int i = 0; i < 4; i++) lf.curr_delta_lf[i] = pcs->ppcs->curr_delta_lf[i];
    *out = lf;
}

void svt_av1_loop_filter_frame(EbPictureBufferDesc *frame_buffer, PictureControlSet *pcs, int32_t plane_start,
                               int32_t plane_end) {
    stat_hit(ST_DLF);
    CtxSlot       *s = slot_for(pcs);
    svtf_lf_params lf;
    lf_of(pcs, &lf);
    begin_picture(s, pcs, frame_buffer);
    CHECK(svtf_deblock_frame(s->ctx, &lf, plane_start, plane_end), s->ctx);
    if (!g_resident)
        end_picture(s, frame_buffer), s->resident = 0;
}

/* svt_av1_loop_filter_init (Encoder/Codec/EbDeblockingFilter.c:35): the limits / thresholds it fills in pcs->ppcs->lf_info are
 * rebuilt on every deblock call from the frame header (svtf_lf_build_tables), so only the flag the per-SB filter reads stays. */
void svt_av1_loop_filter_init(PictureControlSet *pcs) { pcs->ppcs->frm_hdr.loop_filter_params.combine_vert_horz_lf = 1; }

/* svt_aom_loop_filter_sb (Encoder/Codec/EbDeblockingFilter.c:547), the per-superblock deblock EncDec runs right after a superblock is
 * reconstructed (Encoder/Codec/EbCodingLoop.c:2260-2281, dlf_level >= 3: presets M6 and up).  The GPU filters whole pictures: the call
 * is a no-op until the LAST superblock of the picture arrives — EncDec walks superblocks in wavefront order (a superblock starts after
 * its left and top-right neighbours), so the bottom-right one comes last, when every superblock's reconstruction and mode info are
 * final — and then the whole picture is filtered in one launch.  The per-SB and the per-frame order give the same pixels: edges of one
 * pass never overlap, and every horizontal edge reads the column-pass result of all samples it touches in both orders. */
void svt_aom_loop_filter_sb(EbPictureBufferDesc *frame_buffer, PictureControlSet *pcs, int32_t mi_row, int32_t mi_col, int32_t plane_start,
                            int32_t plane_end, uint8_t last_col) {
    const Av1Common *cm    = pcs->ppcs->av1_cm;
    const int        sb_mi = pcs->scs->sb_size >> 2;
    (void)mi_col;
    if (!last_col || mi_row + sb_mi < cm->mi_rows)
        return;
    stat_hit(ST_DLF_SB_LAST);
    svt_av1_loop_filter_frame(frame_buffer, pcs, plane_start, plane_end);
}

/* svt_av1_pick_filter_level (Encoder/Codec/EbDeblockingFilter.c:1129).  LPF_PICK_MINIMAL_LPF and LPF_PICK_FROM_Q are
 * the reference's scalar formulas (svt_av1_pick_filter_level_by_q stays where it is); the full-image search
 * (:1147-1249: up to ~3 x 8 trial deblocks of a whole plane + SSE against the source) is svtf_lf_search_frame.  The
 * reconstruction is uploaded once and stays on the device for the svt_av1_loop_filter_frame call that follows
 * (EbDlfProcess.c:104-106) when svtf_adapter_set_resident(1) is on; the temp_lf_recon_pic allocation is not needed. */
EbErrorType svt_av1_pick_filter_level(EbPictureBufferDesc *srcBuffer, PictureControlSet *pcs, LpfPickMethod method) {
    SequenceControlSet *scs = pcs->scs;
    FrameHeader        *fh  = &pcs->ppcs->frm_hdr;
    struct LoopFilter  *l   = &fh->loop_filter_params;
    l->sharpness_level      = 0;
    if (method == LPF_PICK_MINIMAL_LPF) {
        l->filter_level[0] = l->filter_level[1] = 0;
        return EB_ErrorNone;
    }
    if (method >= LPF_PICK_FROM_Q) {
        int32_t lvl[4];
        svt_av1_pick_filter_level_by_q(pcs, fh->quantization_params.base_q_idx, lvl);
        l->filter_level[0] = lvl[0], l->filter_level[1] = lvl[1], l->filter_level_u = lvl[2], l->filter_level_v = lvl[3];
        return EB_ErrorNone;
    }
    if (pcs->ppcs->dlf_ctrls.dlf_avg && pcs->ppcs->tot_ref_frame_types > 0) { /* :1175-1205, start from the references' mean */
        int32_t tot[4] = {0, 0, 0, 0}, refs = 0;
        for (uint32_t it = 0; it < pcs->ppcs->tot_ref_frame_types; ++it) {
            MvReferenceFrame rf[2];
            av1_set_ref_frame(rf, pcs->ppcs->ref_frame_type_arr[it]);
            if (rf[1] != NONE_FRAME)
                continue;
            const EbReferenceObject *ro = pcs->ref_pic_ptr_array[get_list_idx(rf[0])][get_ref_frame_idx(rf[0])]->object_ptr;
            tot[0] += ro->filter_level[0], tot[1] += ro->filter_level[1], tot[2] += ro->filter_level_u, tot[3] += ro->filter_level_v;
            refs++;
        }
        l->filter_level[0] = tot[0] / refs, l->filter_level[1] = tot[1] / refs;
        l->filter_level_u = tot[2] / refs, l->filter_level_v = tot[3] / refs;
    }
    const Bool           is_16bit = scs->is_16bit_pipeline;
    EbPictureBufferDesc *recon;
    svt_aom_get_recon_pic(pcs, &recon, is_16bit);
    /* the picture the error is measured against (picture_sse_calculations :736, :783) */
    const EbPictureBufferDesc *src = is_16bit ? pcs->input_frame16bit : (EbPictureBufferDesc *)pcs->ppcs->enhanced_pic;
    (void)srcBuffer;
    CtxSlot       *s = slot_for(pcs);
    svtf_lf_params lf;
    lf_of(pcs, &lf);
    begin_picture(s, pcs, recon);
    void   *p[3];
    int32_t st[3];
    plane_ptrs(src, s->geom.sample_bytes, p, st);
    CHECK(svtf_upload_source(s->ctx, p[0], st[0], p[1], st[1], p[2], st[2]), s->ctx);
    svtf_lf_search_cfg cfg;
    cfg.dlf_avg                = pcs->ppcs->dlf_ctrls.dlf_avg;
    cfg.early_exit_convergence = pcs->ppcs->dlf_ctrls.early_exit_convergence;
    cfg.only_4x4               = fh->tx_mode == ONLY_4X4;
    cfg.search_chroma          = !(pcs->ppcs->dlf_ctrls.dlf_avg_uv && pcs->temporal_layer_index > 0); /* :1218 */
    int32_t out[4];
    stat_hit(ST_LF_SEARCH);
    CHECK(svtf_lf_search_frame(s->ctx, &lf, &cfg, out), s->ctx);
    l->filter_level[0] = out[0], l->filter_level[1] = out[1], l->filter_level_u = out[2], l->filter_level_v = out[3];
    if (!g_resident)
        s->resident = 0; /* the picture on the device is unchanged; nothing to copy back */
    return EB_ErrorNone;
}

static void fb_indices(const PictureControlSet *pcs, const svtf_geom *g, int8_t *idx) {
    const int nvfb = (g->mi_rows + 15) / 16, nhfb = (g->mi_cols + 15) / 16;
    for (int fbr = 0; fbr < nvfb; fbr++)
        for (int fbc = 0; fbc < nhfb; fbc++)
            idx[fbr * nhfb + fbc] = pcs->mi_grid_base[16 * fbr * pcs->mi_stride + 16 * fbc]->mbmi.cdef_strength;
}

void svt_av1_cdef_frame(SequenceControlSet *scs, PictureControlSet *pcs) {
    stat_hit(ST_CDEF_APPLY);
    CtxSlot             *s = slot_for(pcs);
    EbPictureBufferDesc *recon;
    svt_aom_get_recon_pic(pcs, &recon, scs->is_16bit_pipeline);
    const CdefParams *cp = &pcs->ppcs->frm_hdr.cdef_params;
    svtf_cdef_params  p;
    memset(&p, 0, sizeof(p));
    p.cdef_damping = cp->cdef_damping, p.cdef_bits = cp->cdef_bits;
    for (int i = 0; i < CDEF_MAX_STRENGTHS && i < SVTF_CDEF_MAX_STRENGTHS; i++)
        p.cdef_y_strength[i] = cp->cdef_y_strength[i], p.cdef_uv_strength[i] = cp->cdef_uv_strength[i];
    const int nfb = ((s->geom.mi_rows + 15) / 16) * ((s->geom.mi_cols + 15) / 16);
    if (s->idx_cap < (size_t)nfb) {
        free(s->idx_stage);
        s->idx_stage = (int8_t *)malloc(nfb);
        s->idx_cap   = s->idx_stage ? (size_t)nfb : 0;
        if (!s->idx_stage)
            die("malloc(cdef indices)", s->ctx);
    }
    int8_t *idx = s->idx_stage;
    fb_indices(pcs, &s->geom, idx);
    /* the reference reads a stale index for all-skip fbs and then skips them; clamp it into range */
    for (int i = 0; i < nfb; i++)
        if (idx[i] >= (1 << p.cdef_bits) || idx[i] < 0)
            idx[i] = 0;
    begin_picture(s, pcs, recon);
    CHECK(svtf_cdef_apply_frame(s->ctx, &p, idx), s->ctx); /* copies the indices into its own staging: nothing to wait for */
    if (!g_resident)
        end_picture(s, recon), s->resident = 0;
}

static void search_cfg_of(const PictureControlSet *pcs, svtf_cdef_search_cfg *c) {
    const CdefControls *cc = &pcs->ppcs->cdef_ctrls;
    memset(c, 0, sizeof(*c));
    c->first_pass_fs_num          = cc->first_pass_fs_num;
    c->default_second_pass_fs_num = cc->default_second_pass_fs_num;
    memcpy(c->default_first_pass_fs, cc->default_first_pass_fs, 64);
    memcpy(c->default_second_pass_fs, cc->default_second_pass_fs, 64);
    memcpy(c->default_first_pass_fs_uv, cc->default_first_pass_fs_uv, 64);
    memcpy(c->default_second_pass_fs_uv, cc->default_second_pass_fs_uv, 64);
    c->subsampling_factor = cc->subsampling_factor;
    c->base_q_idx         = pcs->ppcs->frm_hdr.quantization_params.base_q_idx;
    c->zero_fs_cost_bias  = cc->zero_fs_cost_bias;
}

/* cdef_seg_search over the whole picture (the first CDEF segment to arrive calls this, the others no-op):
 * fills pcs->mse_seg, pcs->skip_cdef_seg and pcs->cdef_dir_data like the reference. */
void svtf_cdef_search_picture(PictureControlSet *pcs, SequenceControlSet *scs) {
    stat_hit(ST_CDEF_SEARCH);
    CtxSlot             *s = slot_for(pcs);
    EbPictureBufferDesc *recon;
    svt_aom_get_recon_pic(pcs, &recon, scs->is_16bit_pipeline);
    const EbPictureBufferDesc *input = scs->is_16bit_pipeline ? pcs->input_frame16bit : pcs->ppcs->enhanced_pic;
    svtf_cdef_search_cfg       cfg;
    search_cfg_of(pcs, &cfg);
    const int nfb = ((s->geom.mi_rows + 15) / 16) * ((s->geom.mi_cols + 15) / 16);
    begin_picture(s, pcs, recon);
    void   *p[3];
    int32_t st[3];
    plane_ptrs(input, s->geom.sample_bytes, p, st);
    CHECK(svtf_upload_source(s->ctx, p[0], st[0], p[1], st[1], p[2], st[2]), s->ctx);
    uint64_t *mse = (uint64_t *)xmalloc(sizeof(uint64_t) * 128 * (size_t)nfb, s->ctx);
    uint8_t  *dir = (uint8_t *)xmalloc(64 * (size_t)nfb, s->ctx);
    int32_t  *var = (int32_t *)xmalloc(sizeof(int32_t) * 64 * (size_t)nfb, s->ctx);
    CHECK(svtf_cdef_search_frame(s->ctx, &cfg, mse, pcs->skip_cdef_seg, dir, var), s->ctx);
    memcpy(pcs->mse_seg[0], mse, sizeof(uint64_t) * 64 * (size_t)nfb);
    memcpy(pcs->mse_seg[1], mse + (size_t)64 * nfb, sizeof(uint64_t) * 64 * (size_t)nfb);
    /* directions / variances: the 16x16 CdefDirData of a 128-wide / 128-high coding block belongs to its first quadrant */
    const int nhfb = (s->geom.mi_cols + 15) / 16, nvfb = (s->geom.mi_rows + 15) / 16;
    for (int fbr = 0; fbr < nvfb; fbr++)
        for (int fbc = 0; fbc < nhfb; fbc++) {
            const BlockSize bs = pcs->mi_grid_base[16 * fbr * pcs->mi_stride + 16 * fbc]->mbmi.block_mi.bsize;
            const int       dc = (fbc & 1) && (bs == BLOCK_128X128 || bs == BLOCK_128X64);
            const int       dr = (fbr & 1) && (bs == BLOCK_128X128 || bs == BLOCK_64X128);
            const int       fb = fbr * nhfb + fbc, ofb = (fbr - dr) * nhfb + fbc - dc;
            for (int b = 0; b < 64; b++) {
                if ((dr || dc) && !dir[fb * 64 + b] && !var[fb * 64 + b])
                    continue; /* a block the search did not list: the first quadrant's array keeps what it had */
                pcs->cdef_dir_data[ofb].dir[(b >> 3) + 8 * dr][(b & 7) + 8 * dc] = dir[fb * 64 + b];
                pcs->cdef_dir_data[ofb].var[(b >> 3) + 8 * dr][(b & 7) + 8 * dc] = var[fb * 64 + b];
            }
        }
    free(mse), free(dir), free(var);
}

/* 64x64 filter blocks that the reference folds into the first filter block of a 128-wide / 128-high coding block
 * (Encoder/Codec/EbCdefProcess.c:186-189, Encoder/Codec/EbEncCdef.c:756-759, :822-826): the right / lower quadrants are neither
 * searched nor picked on their own — the first quadrant carries the distortion of the whole block and its index is copied to them. */
static int fb_folded(const PictureControlSet *pcs, int fbr, int fbc) {
    const BlockSize bs = pcs->mi_grid_base[16 * fbr * pcs->mi_stride + 16 * fbc]->mbmi.block_mi.bsize;
    return ((fbc & 1) && (bs == BLOCK_128X128 || bs == BLOCK_128X64)) || ((fbr & 1) && (bs == BLOCK_128X128 || bs == BLOCK_64X128));
}
/* "we generate param only for the first 64x64 ... we need to duplicate data" (EbEncCdef.c:779-793, :884-905) */
static void fb_set_strength(PictureControlSet *pcs, int fbr, int fbc, int8_t gi) {
    const int  at = 16 * fbr * pcs->mi_stride + 16 * fbc;
    ModeInfo **g  = pcs->mi_grid_base;
    g[at]->mbmi.cdef_strength = gi;
    switch (g[at]->mbmi.block_mi.bsize) {
    case BLOCK_128X128:
        g[at + 16]->mbmi.cdef_strength                       = gi;
        g[at + 16 * pcs->mi_stride]->mbmi.cdef_strength      = gi;
        g[at + 16 * pcs->mi_stride + 16]->mbmi.cdef_strength = gi;
        break;
    case BLOCK_128X64: g[at + 16]->mbmi.cdef_strength = gi; break;
    case BLOCK_64X128: g[at + 16 * pcs->mi_stride]->mbmi.cdef_strength = gi; break;
    default: break;
    }
}

void finish_cdef_search(PictureControlSet *pcs) {
    CtxSlot                 *s    = slot_for(pcs);
    PictureParentControlSet *ppcs = pcs->ppcs;
    FrameHeader             *fh   = &ppcs->frm_hdr;
    const CdefControls      *cc   = &ppcs->cdef_ctrls;
    const int nvfb = (s->geom.mi_rows + 15) / 16, nhfb = (s->geom.mi_cols + 15) / 16, nfb = nvfb * nhfb;
    if (cc->use_reference_cdef_fs) {
        /* EbEncCdef.c:744-793: no search ran; one strength pair predicted from the reference frames, index 0 for every filter block
         * that is not all-skip.  Pure bookkeeping on the host — no pixel work, so nothing for the GPU to do. */
        stat_hit(ST_CDEF_PICK_REFFS);
        for (int fbr = 0; fbr < nvfb; fbr++)
            for (int fbc = 0; fbc < nhfb; fbc++) {
                if (fb_folded(pcs, fbr, fbc))
                    continue;
                int all_skip = 1; /* svt_sb_all_skip (:219-235) */
                const int maxr = s->geom.mi_rows - 16 * fbr < 16 ? s->geom.mi_rows - 16 * fbr : 16;
                const int maxc = s->geom.mi_cols - 16 * fbc < 16 ? s->geom.mi_cols - 16 * fbc : 16;
                for (int r = 0; r < maxr && all_skip; r++)
                    for (int c = 0; c < maxc; c++)
                        if (!pcs->mi_grid_base[(16 * fbr + r) * pcs->mi_stride + 16 * fbc + c]->mbmi.block_mi.skip) {
                            all_skip = 0;
                            break;
                        }
                if (!all_skip)
                    fb_set_strength(pcs, fbr, fbc, 0);
            }
        fh->cdef_params.cdef_bits           = 0;
        ppcs->nb_cdef_strengths             = 1;
        fh->cdef_params.cdef_damping        = (uint8_t)(3 + (fh->quantization_params.base_q_idx >> 6));
        fh->cdef_params.cdef_y_strength[0]  = cc->pred_y_f;
        fh->cdef_params.cdef_uv_strength[0] = cc->pred_uv_f;
        return;
    }
    stat_hit(ST_CDEF_PICK);
    svtf_cdef_search_cfg cfg;
    search_cfg_of(pcs, &cfg);
    uint32_t fast_lambda = 0, full_lambda = 0;
    (*svt_aom_av1_lambda_assignment_function_table[ppcs->pred_structure])(
        pcs, &fast_lambda, &full_lambda, (uint8_t)ppcs->enhanced_pic->bit_depth, fh->quantization_params.base_q_idx, FALSE);
    cfg.lambda = full_lambda;
    /* the table lives in pcs->mse_seg (the reference's bias pass edits it in place; so do we) */
    uint64_t *mse  = (uint64_t *)malloc(sizeof(uint64_t) * 128 * (size_t)nfb);
    uint8_t  *skip = (uint8_t *)malloc(nfb);
    int8_t   *idx  = (int8_t *)malloc(nfb);
    if (!mse || !skip || !idx)
        die("malloc(finish_cdef_search)", s->ctx);
    memcpy(mse, pcs->mse_seg[0], sizeof(uint64_t) * 64 * (size_t)nfb);
    memcpy(mse + (size_t)64 * nfb, pcs->mse_seg[1], sizeof(uint64_t) * 64 * (size_t)nfb);
    for (int fbr = 0; fbr < nvfb; fbr++) /* folded quadrants take no part in the pick (:822-826) */
        for (int fbc = 0; fbc < nhfb; fbc++) skip[fbr * nhfb + fbc] = fb_folded(pcs, fbr, fbc) ? 1 : pcs->skip_cdef_seg[fbr * nhfb + fbc];
    svtf_cdef_params out;
    /* GPU pick from the caller's table: 75 svt_search_one_dual passes without a host round trip */
    CHECK(svtf_cdef_pick_table(s->ctx, &cfg, mse, skip, nfb, &out, idx), s->ctx);
    if (cfg.zero_fs_cost_bias)
        for (int fb = 0; fb < nfb; fb++)
            if (!skip[fb]) {
                pcs->mse_seg[0][fb][0] = ((uint16_t)cfg.zero_fs_cost_bias * pcs->mse_seg[0][fb][0]) >> 6;
                pcs->mse_seg[1][fb][0] = ((uint16_t)cfg.zero_fs_cost_bias * pcs->mse_seg[1][fb][0]) >> 6;
            }
    fh->cdef_params.cdef_bits    = (uint8_t)out.cdef_bits;
    fh->cdef_params.cdef_damping = (uint8_t)out.cdef_damping;
    ppcs->nb_cdef_strengths      = 1 << out.cdef_bits;
    for (int i = 0; i < (1 << out.cdef_bits); i++) {
        fh->cdef_params.cdef_y_strength[i]  = out.cdef_y_strength[i];
        fh->cdef_params.cdef_uv_strength[i] = out.cdef_uv_strength[i];
    }
    for (int fbr = 0; fbr < nvfb; fbr++)
        for (int fbc = 0; fbc < nhfb; fbc++)
            if (idx[fbr * nhfb + fbc] >= 0)
                fb_set_strength(pcs, fbr, fbc, idx[fbr * nhfb + fbc]);
    free(mse), free(skip), free(idx);
}

/* extend_ccso_border (Common/Codec/EbCcso.c:185) replicates the frame edge of the classifier luma into a 5-sample border of the
 * caller's ext_rec_y so that the CPU kernels may read taps beyond the picture.  The GPU kernels clamp tap coordinates to the picture
 * instead (same values), and ccso_frame / ccso_search below read only the interior of ext_rec_y: nothing to materialise. */
void extend_ccso_border(uint16_t *buf, const int d, MacroblockdPlane *pd) { (void)buf, (void)d, (void)pd; }

void ccso_frame(EbPictureBufferDesc *frame, PictureControlSet *pcs, MacroblockdPlane *pd, uint16_t *ext_rec_y) {
    stat_hit(ST_CCSO);
    (void)pd;
    CtxSlot        *s  = slot_for(pcs);
    const CcsoInfo *ci = &pcs->ppcs->frm_hdr.ccso_info;
    svtf_ccso_params p;
    memset(&p, 0, sizeof(p));
    uint8_t *flags[3];
    for (int pl = 0; pl < 3; pl++) {
        p.ccso_enable[pl] = ci->ccso_enable[pl], p.ccso_bo_only[pl] = ci->ccso_bo_only[pl], p.quant_idx[pl] = ci->quant_idx[pl];
        p.ext_filter_support[pl] = ci->ext_filter_support[pl], p.edge_clf[pl] = ci->edge_clf[pl];
        p.max_band_log2[pl] = (uint8_t)ci->max_band_log2[pl];
        memcpy(p.filter_offset[pl], ci->filter_offset[pl], SVTF_CCSO_LUT_SIZE);
        const int pw = s->geom.width >> (pl ? 1 : 0), ph = s->geom.height >> (pl ? 1 : 0), bl = pl ? 7 : 8;
        const int nbx = (pw + (1 << bl) - 1) >> bl, nby = (ph + (1 << bl) - 1) >> bl;
        flags[pl] = (uint8_t *)xmalloc((size_t)nbx * nby, s->ctx);
        for (int by = 0; by < nby; by++)
            for (int bx = 0; bx < nbx; bx++) { /* Common/Codec/EbCcso.c:317-322 */
                const MbModeInfo *m = &pcs->mi_grid_base[64 * by * pcs->mi_stride + 64 * bx]->mbmi;
                flags[pl][by * nbx + bx] = pl == 0 ? m->ccso_blk_y : (pl == 1 ? m->ccso_blk_u : m->ccso_blk_v);
            }
    }
    if (g_resident && s->resident) {
        /* post-CDEF picture and pre-CDEF luma are already on the device */
    } else {
        /* stand-alone call: the classifier input is the caller's ext_rec_y (pre-CDEF luma, 5-sample padded,
         * Encoder/Codec/EbCdefProcess.c:480-502): stage it as the "pre-CDEF" picture by running an identity CDEF */
        const int       W = s->geom.width, H = s->geom.height, cw = s->geom.mi_cols * 4, chh = s->geom.mi_rows * 4, es = W + 10;
        const int       sb = s->geom.sample_bytes;
        void           *pp[3];
        int32_t         st[3];
        plane_ptrs(frame, sb, pp, st);
        uint8_t *ly = (uint8_t *)xmalloc((size_t)cw * chh * sb, s->ctx);
        for (int r = 0; r < chh; r++)
            for (int c = 0; c < cw; c++) {
                const uint16_t v = ext_rec_y[(size_t)(5 + (r < H ? r : H - 1)) * es + 5 + (c < W ? c : W - 1)];
                if (sb == 1)
                    ly[(size_t)r * cw + c] = (uint8_t)v;
                else
                    ((uint16_t *)ly)[(size_t)r * cw + c] = v;
            }
        CHECK(svtf_frame_begin(s->ctx, &s->geom), s->ctx);
        CHECK(svtf_upload_frame(s->ctx, ly, cw, pp[1], st[1], pp[2], st[2]), s->ctx);
        send_mi(s, pcs);
        svtf_cdef_params idp;
        memset(&idp, 0, sizeof(idp));
        const int nfb = ((s->geom.mi_rows + 15) / 16) * ((s->geom.mi_cols + 15) / 16);
        int8_t   *idx = (int8_t *)xcalloc(nfb, 1, s->ctx);
        CHECK(svtf_cdef_apply_frame(s->ctx, &idp, idx), s->ctx); /* strengths 0: copies, marks the pre-CDEF surface */
        CHECK(svtf_sync(s->ctx), s->ctx);
        free(idx), free(ly);
        /* now overwrite the current (post-"CDEF") surface with the picture to be modified */
        svtf_dev_plane dp[3];
        CHECK(svtf_dev_planes(s->ctx, 1, dp), s->ctx);
        CHECK(svtf_overwrite_current(s->ctx, pp[0], st[0], pp[1], st[1], pp[2], st[2]), s->ctx);
    }
    const uint8_t *cf[3] = {flags[0], flags[1], flags[2]};
    CHECK(svtf_ccso_apply_frame(s->ctx, &p, cf), s->ctx);
    end_picture(s, frame);
    s->resident = 0;
    for (int pl = 0; pl < 3; pl++) free(flags[pl]);
}

/* ccso_search (Encoder/Codec/EbPickccso.c:785): ext_rec_y = pre-CDEF luma padded by 5, rec_uv / org_uv = the picture to
 * correct and the source, uint16 samples at luma stride (Encoder/Codec/EbCdefProcess.c:524-661).  The three planes are
 * staged on the picture's context (classifier luma as the "pre-CDEF" picture, rec as the current picture, org as the
 * source), svtf_ccso_derive_frame runs the search, the result goes where the reference writes it. */
void ccso_search(PictureControlSet *pcs, MacroblockdPlane *pd, int rdmult, const uint16_t *ext_rec_y, uint16_t *rec_uv[3], uint16_t *org_uv[3]) {
    stat_hit(ST_CCSO_SEARCH);
    (void)pd;
    FrameHeader  *fh     = &pcs->ppcs->frm_hdr;
    int           weight = fh->quantization_params.base_q_idx;
    weight               = weight < 1 ? 1 : (weight > 63 ? 63 : weight);
    const int64_t rd     = (int64_t)rdmult * weight; /* :788-793 */
    if (rd >= 2147483647)
        return;
    CtxSlot  *s  = slot_for(pcs);
    const int W = s->geom.width, H = s->geom.height, cw = s->geom.mi_cols * 4, chh = s->geom.mi_rows * 4, es = W + 10;
    const int sb = s->geom.sample_bytes;
    /* planes in the context's sample type and coded size (replicated beyond the visible picture; those samples are
     * never read by the derivation, which works on width x height) */
    uint8_t *buf[3][3]; /* [kind: classifier, rec, org][plane] */
    for (int k = 0; k < 3; k++)
        for (int p = 0; p < 3; p++) {
            const int pw = cw >> (p ? 1 : 0), ph = chh >> (p ? 1 : 0), vw = W >> (p ? 1 : 0), vh = H >> (p ? 1 : 0);
            buf[k][p]    = (uint8_t *)xcalloc((size_t)pw * ph, sb, s->ctx);
            for (int r = 0; r < ph; r++)
                for (int c = 0; c < pw; c++) {
                    const int rr = r < vh ? r : vh - 1, cc = c < vw ? c : vw - 1;
                    uint16_t  v;
                    if (k == 0)
                        v = p == 0 ? ext_rec_y[(size_t)(5 + rr) * es + 5 + cc] : 0;
                    else
                        v = (k == 1 ? rec_uv[p] : org_uv[p])[(size_t)rr * W + cc];
                    if (sb == 1)
                        buf[k][p][(size_t)r * pw + c] = (uint8_t)v;
                    else
                        ((uint16_t *)buf[k][p])[(size_t)r * pw + c] = v;
                }
        }
    CHECK(svtf_frame_begin(s->ctx, &s->geom), s->ctx);
    CHECK(svtf_upload_frame(s->ctx, buf[0][0], cw, buf[0][1], cw / 2, buf[0][2], cw / 2), s->ctx);
    send_mi(s, pcs);
    svtf_cdef_params idp;
    memset(&idp, 0, sizeof(idp));
    const int nfb = ((s->geom.mi_rows + 15) / 16) * ((s->geom.mi_cols + 15) / 16);
    int8_t   *idx = (int8_t *)xcalloc(nfb, 1, s->ctx);
    CHECK(svtf_cdef_apply_frame(s->ctx, &idp, idx), s->ctx); /* strengths 0: copies, marks the pre-CDEF surface */
    CHECK(svtf_overwrite_current(s->ctx, buf[1][0], cw, buf[1][1], cw / 2, buf[1][2], cw / 2), s->ctx);
    CHECK(svtf_upload_source(s->ctx, buf[2][0], cw, buf[2][1], cw / 2, buf[2][2], cw / 2), s->ctx);
    svtf_ccso_params out;
    memset(&out, 0, sizeof(out));
    uint8_t *flags[3];
    int      nbx[3], nby[3];
    for (int p = 0; p < 3; p++) {
        const int pw = W >> (p ? 1 : 0), ph = H >> (p ? 1 : 0), bl = p ? 7 : 8;
        nbx[p] = (pw + (1 << bl) - 1) >> bl, nby[p] = (ph + (1 << bl) - 1) >> bl;
        flags[p] = (uint8_t *)xcalloc((size_t)nbx[p] * nby[p], 1, s->ctx);
    }
    CHECK(svtf_ccso_derive_frame(s->ctx, (int)rd, 0, 3, &out, flags), s->ctx);
    CcsoInfo *ci = &fh->ccso_info;
    for (int p = 0; p < 3; p++) {
        ci->ccso_enable[p] = out.ccso_enable[p];
        if (!out.ccso_enable[p])
            continue;
        for (int by = 0; by < nby[p]; by++)
            for (int bx = 0; bx < nbx[p]; bx++) { /* :731-751 */
                MbModeInfo *m = &pcs->mi_grid_base[64 * by * pcs->mi_stride + 64 * bx]->mbmi;
                const uint8_t f = flags[p][by * nbx[p] + bx];
                if (p == 0) m->ccso_blk_y = f;
                if (p == 1) m->ccso_blk_u = f;
                if (p == 2) m->ccso_blk_v = f;
            }
        memcpy(ci->filter_offset[p], out.filter_offset[p], SVTF_CCSO_LUT_SIZE);
        ci->quant_idx[p] = out.quant_idx[p], ci->ext_filter_support[p] = out.ext_filter_support[p];
        ci->ccso_bo_only[p] = out.ccso_bo_only[p], ci->max_band_log2[p] = out.max_band_log2[p], ci->edge_clf[p] = out.edge_clf[p];
    }
#if CONFIG_D143_CCSO_FM_FLAG
    ci->ccso_frame_flag = ci->ccso_enable[0] | ci->ccso_enable[1] | ci->ccso_enable[2];
#endif
    s->resident = 0;
    free(idx);
    for (int p = 0; p < 3; p++) free(flags[p]);
    for (int k = 0; k < 3; k++)
        for (int p = 0; p < 3; p++) free(buf[k][p]);
}

/* ----------------------------------------------------------------------------------------------------
 * svt_av1_loop_restoration_filter_frame (Common/Codec/EbRestoration.c:1179), called by the restoration kernel
 * (Encoder/Codec/EbRestProcess.c) on cm->frame_to_show after the filters were picked.
 * Resident chain (svtf_adapter_set_resident(1), the picture went DLF -> CDEF on this pcs's context): the stripe borders
 * come from the deblocked surface still on the device, svt_av1_loop_restoration_save_boundary_lines has nothing to do,
 * and this call ends the chain (download).  Stand-alone: `frame` is uploaded and the deblocked rows the reference saved
 * in rst_info[p].boundaries (two above and two below every 64-row stripe, :1522-1574) are written into an otherwise
 * empty picture that is uploaded as the pre-CDEF picture — only those rows are ever read. */
void svt_av1_loop_restoration_filter_frame(int32_t *rst_tmpbuf, Yv12BufferConfig *frame, Av1Common *cm, int32_t optimized_lr) {
    stat_hit(ST_LR_APPLY);
    (void)rst_tmpbuf;
    PictureControlSet *pcs = cm->child_pcs;
    CtxSlot           *s   = slot_for_at(pcs, 1);
    const int          sb  = s->geom.sample_bytes;
    svtf_lr_params     lr;
    svtf_lr_unit      *units[3] = {NULL, NULL, NULL};
    memset(&lr, 0, sizeof(lr));
    lr.optimized_lr = optimized_lr;
    for (int p = 0; p < 3; p++) {
        const RestorationInfo *rsi = &pcs->rst_info[p];
        lr.frame_restoration_type[p] = rsi->frame_restoration_type == RESTORE_NONE ? SVTF_RESTORE_NONE : (int32_t)rsi->frame_restoration_type;
        lr.restoration_unit_size[p]  = rsi->restoration_unit_size;
        if (rsi->frame_restoration_type == RESTORE_NONE)
            continue;
        units[p] = (svtf_lr_unit *)xcalloc(rsi->units_per_tile, sizeof(svtf_lr_unit), s->ctx);
        for (int u = 0; u < rsi->units_per_tile; u++) {
            const RestorationUnitInfo *ri = &rsi->unit_info[u];
            units[p][u].restoration_type  = ri->restoration_type == RESTORE_WIENER ? SVTF_RESTORE_WIENER
                 : ri->restoration_type == RESTORE_SGRPROJ                          ? SVTF_RESTORE_SGRPROJ
                                                                                    : SVTF_RESTORE_NONE;
            memcpy(units[p][u].wiener_vfilter, ri->wiener_info.vfilter, sizeof(units[p][u].wiener_vfilter));
            memcpy(units[p][u].wiener_hfilter, ri->wiener_info.hfilter, sizeof(units[p][u].wiener_hfilter));
            units[p][u].sgr_ep = ri->sgrproj_info.ep, units[p][u].sgr_xqd[0] = ri->sgrproj_info.xqd[0], units[p][u].sgr_xqd[1] = ri->sgrproj_info.xqd[1];
        }
    }
    void   *fp[3];
    int32_t fst[3];
    for (int p = 0; p < 3; p++) {
        fp[p]  = cm->use_highbitdepth ? (void *)CONVERT_TO_SHORTPTR(frame->buffers[p]) : (void *)frame->buffers[p];
        fst[p] = frame->strides[p > 0];
    }
    if (!(g_resident && s->resident)) {
        CHECK(svtf_frame_begin(s->ctx, &s->geom), s->ctx);
        CHECK(svtf_upload_frame(s->ctx, fp[0], fst[0], fp[1], fst[1], fp[2], fst[2]), s->ctx);
        if (!optimized_lr) {
            uint8_t *dbk[3];
            int32_t  dst[3];
            for (int p = 0; p < 3; p++) {
                const int ss = p ? 1 : 0, cw = (s->geom.mi_cols * 4) >> ss, chh = (s->geom.mi_rows * 4) >> ss;
                const int pw = (s->geom.width + ss) >> ss, ph = (s->geom.height + ss) >> ss, sh = 64 >> ss, off = 8 >> ss;
                const RestorationStripeBoundaries *b = &pcs->rst_info[p].boundaries;
                dbk[p] = (uint8_t *)xcalloc((size_t)cw * chh, sb, s->ctx), dst[p] = cw;
                for (int k = 0;; k++) {
                    const int y0 = k * sh - off > 0 ? k * sh - off : 0, y1 = (k + 1) * sh - off < ph ? (k + 1) * sh - off : ph;
                    if (y0 >= ph)
                        break;
                    for (int i = 0; i < RESTORATION_CTX_VERT; i++) {
                        const size_t row = (size_t)(RESTORATION_CTX_VERT * k + i) * b->stripe_boundary_stride + RESTORATION_EXTRA_HORZ;
                        if (k > 0)
                            memcpy(dbk[p] + (size_t)(y0 - 2 + i) * cw * sb, b->stripe_boundary_above + row * sb, (size_t)pw * sb);
                        if (y1 < ph && y1 + i < chh)
                            memcpy(dbk[p] + (size_t)(y1 + i) * cw * sb, b->stripe_boundary_below + row * sb, (size_t)pw * sb);
                    }
                }
            }
            CHECK(svtf_upload_pre_cdef(s->ctx, dbk[0], dst[0], dbk[1], dst[1], dbk[2], dst[2]), s->ctx);
            CHECK(svtf_sync(s->ctx), s->ctx);
            for (int p = 0; p < 3; p++) free(dbk[p]);
        }
    }
    const svtf_lr_unit *cu[3] = {units[0], units[1], units[2]};
    CHECK(svtf_lr_apply_frame(s->ctx, &lr, cu), s->ctx);
    CHECK(svtf_download_frame(s->ctx, fp[0], fst[0], fp[1], fst[1], fp[2], fst[2]), s->ctx);
    s->resident = 0;
    for (int p = 0; p < 3; p++) free(units[p]);
}

/* ----------------------------------------------------------------------------------------------------
 * svt_av1_superres_upscale_frame (Encoder/Codec/EbRestProcess.c:405), called by the CDEF kernel after CDEF / CCSO when the
 * frame was coded with super-resolution (Encoder/Codec/EbCdefProcess.c:675-677): the recon picture is upscaled in place
 * (the reference copies it aside and runs svt_av1_upscale_normative_rows per plane).  Resident chain: the picture on the
 * device is upscaled (with the deblocked picture loop restoration takes its stripe borders from) and stays there;
 * stand-alone: recon is uploaded, upscaled and written back into the same buffers.
 * ---------------------------------------------------------------------------------------------------- */
void svt_av1_superres_upscale_frame(struct Av1Common *cm, PictureControlSet *pcs, SequenceControlSet *scs) {
    stat_hit(ST_SUPERRES);
    EbPictureBufferDesc *recon;
    svt_aom_get_recon_pic(pcs, &recon, scs->is_16bit_pipeline);
    CtxSlot *s = slot_for(pcs);
    void    *p[3];
    int32_t  st[3];
    plane_ptrs(recon, s->geom.sample_bytes, p, st);
    const int resident = g_resident && s->resident;
    if (!resident) {
        CHECK(svtf_frame_begin(s->ctx, &s->geom), s->ctx);
        CHECK(svtf_upload_frame(s->ctx, p[0], st[0], p[1], st[1], p[2], st[2]), s->ctx);
    }
    svtf_superres_params sr;
    memset(&sr, 0, sizeof(sr));
    sr.upscaled_width       = cm->frm_size.superres_upscaled_width;
    sr.superres_denominator = cm->frm_size.superres_denominator;
    sr.tile_cols            = cm->tiles_info.tile_cols;
    for (int j = 0; j <= sr.tile_cols && j < 65; j++) sr.tile_col_start_mi[j] = cm->tiles_info.tile_col_start_mi[j];
    sr.filters = &av1_resize_filter_normative[0][0];
    CHECK(svtf_superres_upscale_frame(s->ctx, &sr), s->ctx);
    if (sr.superres_denominator > SCALE_NUMERATOR) { /* the picture of this pcs is the upscaled one from here on */
        s->geom.width   = sr.upscaled_width;
        s->geom.mi_cols = 2 * ((sr.upscaled_width + 7) >> 3);
    }
    if (!resident) {
        CHECK(svtf_download_frame(s->ctx, p[0], st[0], p[1], st[1], p[2], st[2]), s->ctx);
        s->resident = 0;
    }
}

/* ----------------------------------------------------------------------------------------------------
 * restoration_seg_search + rest_finish_search (Encoder/Codec/EbRestorationPick.c:1471 / :1555), called by the restoration
 * kernel (Encoder/Codec/EbRestProcess.c:599-610) once per restoration segment and once per picture.  The GPU searches the
 * whole picture in one call, so the segment that carries segment_index 0 does the work and the others return; the results
 * land where the reference leaves them: pcs->rusi_picture[plane][unit] (sse[RESTORE_NONE], sse[RESTORE_WIENER], wiener) after
 * the segment call, rst_info[plane].frame_restoration_type / unit_info[] after rest_finish_search.
 * Covered: the Wiener and the self-guided tools (and RESTORE_SWITCHABLE between them) without stripe boundaries
 * (scs->use_boundaries_in_rest_search = 0, the reference's setting, Encoder/Globals/EbEncHandle.c:4162; with 1 the adapter
 * aborts rather than return a different decision), wn_filter_ctrls.use_prev_frame_coeffs included.
 * org_fts is the picture after CDEF (uploaded unless the chain is resident), src the source picture.
 * ---------------------------------------------------------------------------------------------------- */
typedef struct {
    const PictureControlSet *pcs;
    int32_t                  frame_type[3];
    svtf_lr_unit            *units[3];
} LrSearchResult;
static LrSearchResult g_lr_res[SVTF_MAX_CACHED];

static LrSearchResult *lr_result_for(const PictureControlSet *pcs, int create) {
    LrSearchResult *r = NULL;
    pthread_mutex_lock(&g_lock);
    for (int i = 0; i < SVTF_MAX_CACHED && !r; i++)
        if (g_lr_res[i].pcs == pcs)
            r = &g_lr_res[i];
    for (int i = 0; i < SVTF_MAX_CACHED && !r && create; i++)
        if (!g_lr_res[i].pcs) {
            r      = &g_lr_res[i];
            r->pcs = pcs;
        }
    pthread_mutex_unlock(&g_lock);
    return r;
}

void restoration_seg_search(int32_t *rst_tmpbuf, Yv12BufferConfig *org_fts, const Yv12BufferConfig *src, Yv12BufferConfig *trial_frame_rst,
                            PictureControlSet *pcs, uint32_t segment_index) {
    (void)rst_tmpbuf, (void)trial_frame_rst;
    if (segment_index != 0)
        return;
    stat_hit(ST_LR_SEARCH);
    Av1Common        *cm = pcs->ppcs->av1_cm;
    const Macroblock *x  = pcs->ppcs->av1x;
    CtxSlot          *s  = slot_for_at(pcs, 1);
    if (cm->use_boundaries_in_rest_search) {
        fprintf(stderr, "Svt[error]: svtfilt_b200 adapter: restoration_seg_search needs use_boundaries_in_rest_search = 0\n");
        abort();
    }
    svtf_lr_search_cfg cfg;
    memset(&cfg, 0, sizeof(cfg));
    for (int p = 0; p < 3; p++) cfg.restoration_unit_size[p] = pcs->rst_info[p].restoration_unit_size;
    cfg.wn_enabled                 = cm->wn_filter_ctrls.enabled;
    cfg.wn_filter_tap_lvl          = cm->wn_filter_ctrls.filter_tap_lvl;
    cfg.wn_use_refinement          = cm->wn_filter_ctrls.use_refinement;
    cfg.wn_max_one_refinement_step = cm->wn_filter_ctrls.max_one_refinement_step;
    cfg.wn_use_chroma              = cm->wn_filter_ctrls.use_chroma;
    cfg.rdmult                     = x->rdmult;
    cfg.wiener_restore_cost[0] = x->wiener_restore_cost[0], cfg.wiener_restore_cost[1] = x->wiener_restore_cost[1];
    cfg.sgrproj_restore_cost[0] = x->sgrproj_restore_cost[0], cfg.sgrproj_restore_cost[1] = x->sgrproj_restore_cost[1];
    for (int t = 0; t < 3; t++) cfg.switchable_restore_cost[t] = x->switchable_restore_cost[t];
    cfg.sg_enabled    = cm->sg_filter_ctrls.enabled;
    cfg.sg_use_chroma = cm->sg_filter_ctrls.use_chroma;
    for (int t = 0; t < 2; t++) { /* the ep range of search_selfguided_restoration (:567-589) */
        const int8_t step = cm->sg_filter_ctrls.step_range, e0 = cm->sg_ref_frame_ep[0], e1 = cm->sg_ref_frame_ep[1];
        if (step < 16) {
            const int mid        = e0 < 0 && e1 < 0 ? 0 : (e1 < 0 ? e0 : (e0 < 0 ? e1 : (e0 + e1) / 2));
            cfg.sg_start_ep[t]   = e0 < 0 && e1 < 0 ? 0 : (mid - step > 0 ? mid - step : 0);
            cfg.sg_end_ep[t]     = e0 < 0 && e1 < 0 ? SGRPROJ_PARAMS : (mid + step < SGRPROJ_PARAMS ? mid + step : SGRPROJ_PARAMS);
            cfg.sg_ep_inc[t] = 1, cfg.sg_refine[t] = 1;
        } else {
            cfg.sg_start_ep[t] = cm->sg_filter_ctrls.start_ep[t], cfg.sg_end_ep[t] = cm->sg_filter_ctrls.end_ep[t];
            cfg.sg_ep_inc[t] = cm->sg_filter_ctrls.ep_inc[t], cfg.sg_refine[t] = cm->sg_filter_ctrls.refine[t];
        }
    }
    void   *fp[3], *sp[3];
    int32_t fst[3], sst[3];
    for (int p = 0; p < 3; p++) {
        fp[p]  = cm->use_highbitdepth ? (void *)CONVERT_TO_SHORTPTR(org_fts->buffers[p]) : (void *)org_fts->buffers[p];
        sp[p]  = cm->use_highbitdepth ? (void *)CONVERT_TO_SHORTPTR(src->buffers[p]) : (void *)src->buffers[p];
        fst[p] = org_fts->strides[p > 0], sst[p] = src->strides[p > 0];
    }
    if (!(g_resident && s->resident)) {
        CHECK(svtf_frame_begin(s->ctx, &s->geom), s->ctx);
        CHECK(svtf_upload_frame(s->ctx, fp[0], fst[0], fp[1], fst[1], fp[2], fst[2]), s->ctx);
    }
    CHECK(svtf_upload_source(s->ctx, sp[0], sst[0], sp[1], sst[1], sp[2], sst[2]), s->ctx);
    /* use_prev_frame_coeffs (:1313-1319): the taps EbRestProcess.c:586-597 left in rst_info[].unit_info[] */
    svtf_lr_unit *prev[3] = {NULL, NULL, NULL};
    if (cm->wn_filter_ctrls.enabled && cm->wn_filter_ctrls.use_prev_frame_coeffs && pcs->ppcs->frm_hdr.frame_type != KEY_FRAME &&
        pcs->ppcs->frm_hdr.frame_type != INTRA_ONLY_FRAME)
        for (int p = 0; p < 3; p++) {
            prev[p] = (svtf_lr_unit *)xcalloc(pcs->rst_info[p].units_per_tile, sizeof(svtf_lr_unit), s->ctx);
            for (int u = 0; u < pcs->rst_info[p].units_per_tile; u++) {
                const RestorationUnitInfo *ri = &pcs->rst_info[p].unit_info[u];
                prev[p][u].restoration_type   = ri->restoration_type == RESTORE_WIENER ? SVTF_RESTORE_WIENER : SVTF_RESTORE_NONE;
                memcpy(prev[p][u].wiener_vfilter, ri->wiener_info.vfilter, sizeof(prev[p][u].wiener_vfilter));
                memcpy(prev[p][u].wiener_hfilter, ri->wiener_info.hfilter, sizeof(prev[p][u].wiener_hfilter));
            }
            cfg.prev_units[p] = prev[p];
        }
    LrSearchResult      *r = lr_result_for(pcs, 1);
    svtf_lr_unit_search *us[3];
    if (!r) {
        fprintf(stderr, "Svt[error]: svtfilt_b200 adapter: more than %d restoration searches in flight\n", SVTF_MAX_CACHED);
        abort();
    }
    for (int p = 0; p < 3; p++) {
        free(r->units[p]);
        r->units[p] = (svtf_lr_unit *)xcalloc(pcs->rst_info[p].units_per_tile, sizeof(svtf_lr_unit), s->ctx);
        us[p]       = (svtf_lr_unit_search *)xcalloc(pcs->rst_info[p].units_per_tile, sizeof(svtf_lr_unit_search), s->ctx);
    }
    int32_t ep_cnt[16];
    CHECK(svtf_lr_search_frame(s->ctx, &cfg, r->frame_type, r->units, us, ep_cnt), s->ctx);
    for (int i = 0; i < 16; i++) cm->sg_frame_ep_cnt[i] += ep_cnt[i]; /* :1250 */
    const int wn = cm->wn_filter_ctrls.enabled, sg = cm->sg_filter_ctrls.enabled;
    const int planes = (wn && cm->wn_filter_ctrls.use_chroma) || (sg && cm->sg_filter_ctrls.use_chroma) ? 3 : 1;
    for (int p = 0; p < planes; p++)
        for (int u = 0; u < pcs->rst_info[p].units_per_tile; u++) {
            RestUnitSearchInfo *ri = &pcs->rusi_picture[p][u];
            ri->sse[RESTORE_NONE]  = us[p][u].sse[0];
            if (wn && (!p || cm->wn_filter_ctrls.use_chroma)) {
                ri->sse[RESTORE_WIENER] = us[p][u].sse[1];
                if (us[p][u].sse[1] != INT64_MAX) {
                    memcpy(ri->wiener.vfilter, us[p][u].wiener_vfilter, sizeof(ri->wiener.vfilter));
                    memcpy(ri->wiener.hfilter, us[p][u].wiener_hfilter, sizeof(ri->wiener.hfilter));
                } else
                    ri->best_rtype[RESTORE_WIENER - 1] = RESTORE_NONE;
            }
            if (sg && (!p || cm->sg_filter_ctrls.use_chroma)) {
                ri->sse[RESTORE_SGRPROJ] = us[p][u].sse[2];
                ri->sgrproj.ep           = us[p][u].sgr_ep;
                ri->sgrproj.xqd[0] = us[p][u].sgr_xqd[0], ri->sgrproj.xqd[1] = us[p][u].sgr_xqd[1];
            }
        }
    for (int p = 0; p < 3; p++) free(us[p]), free(prev[p]);
}

void rest_finish_search(PictureControlSet *pcs) {
    LrSearchResult *r = lr_result_for(pcs, 0);
    if (!r) {
        fprintf(stderr, "Svt[error]: svtfilt_b200 adapter: rest_finish_search without restoration_seg_search\n");
        abort();
    }
    for (int p = 0; p < 3; p++) {
        RestorationInfo *rsi        = &pcs->rst_info[p];
        rsi->frame_restoration_type = (RestorationType)r->frame_type[p];
        if (r->frame_type[p] != SVTF_RESTORE_NONE)
            for (int u = 0; u < rsi->units_per_tile; u++) { /* copy_unit_info :1202 */
                rsi->unit_info[u].restoration_type = (RestorationType)r->units[p][u].restoration_type;
                if (r->units[p][u].restoration_type == SVTF_RESTORE_WIENER) {
                    memcpy(rsi->unit_info[u].wiener_info.vfilter, r->units[p][u].wiener_vfilter, sizeof(rsi->unit_info[u].wiener_info.vfilter));
                    memcpy(rsi->unit_info[u].wiener_info.hfilter, r->units[p][u].wiener_hfilter, sizeof(rsi->unit_info[u].wiener_info.hfilter));
                } else if (r->units[p][u].restoration_type == SVTF_RESTORE_SGRPROJ) {
                    rsi->unit_info[u].sgrproj_info.ep     = r->units[p][u].sgr_ep;
                    rsi->unit_info[u].sgrproj_info.xqd[0] = r->units[p][u].sgr_xqd[0];
                    rsi->unit_info[u].sgrproj_info.xqd[1] = r->units[p][u].sgr_xqd[1];
                }
            }
        free(r->units[p]);
    }
    pthread_mutex_lock(&g_lock);
    memset(r, 0, sizeof(*r));
    pthread_mutex_unlock(&g_lock);
}

/* ----------------------------------------------------------------------------------------------------
 * Recon hand-off: the pixel work of pad_ref_and_set_flags (Encoder/Codec/EbEncDecProcess.c:1589-1744) —
 * right/bottom replication, border padding and the 8-bit / 2-bit split (or narrowing) of the reference picture — from the
 * picture on the GPU.  pad_ref_and_set_flags keeps its bookkeeping (ref_poc, qp, slice_type, ...) and calls this in place of
 * lines 1604-1744.  With a resident chain the picture never was on the host; otherwise it is uploaded from the recon buffer
 * first.  (is_16bit_pipeline with 8-bit content and max_input_pad_right/bottom != 0: the reference narrows the full
 * allocation width including columns its padding did not write; this function writes the padded rectangle only.)
 * ---------------------------------------------------------------------------------------------------- */
void svtf_pad_ref_pictures(PictureControlSet *pcs, SequenceControlSet *scs) {
    stat_hit(ST_PAD_REF);
    EbPictureBufferDesc *ref8, *ref16;
    svt_aom_get_recon_pic(pcs, &ref8, 0);
    svt_aom_get_recon_pic(pcs, &ref16, 1);
    const Bool is_16bit = scs->static_config.encoder_bit_depth > EB_EIGHT_BIT;
    const Bool pipe16   = scs->is_16bit_pipeline;
    CtxSlot   *s        = slot_for_at(pcs, 1);
    if (!(g_resident && s->resident)) { /* stand-alone: the filtered picture is in the recon buffer */
        const EbPictureBufferDesc *pic = (is_16bit || pipe16) ? ref16 : ref8;
        void                      *p[3];
        int32_t                    st[3];
        plane_ptrs(pic, s->geom.sample_bytes, p, st);
        CHECK(svtf_frame_begin(s->ctx, &s->geom), s->ctx);
        CHECK(svtf_upload_frame(s->ctx, p[0], st[0], p[1], st[1], p[2], st[2]), s->ctx);
    }
    const EbPictureBufferDesc *d = (is_16bit || pipe16) ? ref16 : ref8;
    svtf_ref_layout            lay;
    lay.org_x = d->org_x, lay.org_y = d->org_y;
    if (pipe16 && !is_16bit) { /* :1687-1710: borders around the unpadded input size, no right/bottom replication */
        lay.width = d->width - scs->max_input_pad_right, lay.height = d->height - scs->max_input_pad_bottom;
        lay.pad_right = lay.pad_bottom = 0;
    } else {
        lay.width = d->width, lay.height = d->height;
        lay.pad_right = (int32_t)scs->pad_right, lay.pad_bottom = (int32_t)scs->pad_bottom;
    }
    void         *dst[3]  = {d->buffer_y, d->buffer_cb, d->buffer_cr};
    const int32_t dstr[3] = {d->stride_y, d->stride_cb, d->stride_cr};
    if (is_16bit || pipe16) {
        uint8_t      *d8[3]  = {ref8->buffer_y, ref8->buffer_cb, ref8->buffer_cr};
        const int32_t s8[3]  = {ref8->stride_y, ref8->stride_cb, ref8->stride_cr};
        uint8_t      *dn[3]  = {ref8->buffer_bit_inc_y, ref8->buffer_bit_inc_cb, ref8->buffer_bit_inc_cr};
        const int32_t sn[3]  = {ref8->stride_bit_inc_y, ref8->stride_bit_inc_cb, ref8->stride_bit_inc_cr};
        const int     two    = is_16bit && scs->static_config.encoder_bit_depth == EB_TEN_BIT && dn[0] && dn[1] && dn[2];
        CHECK(svtf_download_ref_padded(s->ctx, &lay, dst, dstr, d8, s8, two ? dn : NULL, two ? sn : NULL), s->ctx);
    } else
        CHECK(svtf_download_ref_padded(s->ctx, &lay, dst, dstr, NULL, NULL, NULL, NULL), s->ctx);
}
