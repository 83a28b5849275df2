// svtf_api.cu — the Tier-1 C ABI (include/svtfilt_b200.h): context, surfaces, launches.
// One context = one picture in flight on one CUDA stream.  No CPU fallback: every entry point that
// needs the device fails with SVTF_E_CUDA when there is none.
#include <cuda_runtime.h>
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cfloat>
#include <cmath>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <atomic>
#include <chrono>
#include <thread>
#include <vector>

#include "svtf_kernels.cuh"
#include "svtf_search.cuh"
#include "svtf_search2_kernel.cuh"
#include "svtf_pick.cuh"
#include "svtf_pick2.cuh"
#include "svtf_pick3.cuh"
#include "svtf_ccso_derive.cuh"
#include "svtf_cdef2.cuh"
#include "svtf_cdef3.cuh"
#include "svtf_cdef4.cuh"
#include "svtf_cdef5.cuh"
#include "svtf_search3.cuh"
#include "svtf_mi_pack.cuh"
#include "svtf_lr.cuh"
#include "svtf_lrs.cuh"
#include "svtf_handoff.cuh"
#include "svtf_superres.cuh"

namespace svtf {

static std::mutex  g_err_mutex;
static std::string g_err;

int set_global_error(int code, const char *msg) {
    std::lock_guard<std::mutex> lk(g_err_mutex);
    g_err = msg ? msg : "";
    return code;
}

struct Surface {
    void  *base[3] = {nullptr, nullptr, nullptr};
    int    pitch[3] = {0, 0, 0}; // samples
    Planes planes() const {
        Planes p;
        for (int i = 0; i < 3; i++) p.p[i] = base[i], p.pitch[i] = pitch[i];
        return p;
    }
};

} // namespace svtf

using namespace svtf;

static void derive_scratch_free(void *p);

struct svtf_ctx {
    int          device = 0;
    cudaStream_t stream = nullptr;     // stream in use
    cudaStream_t own_stream = nullptr; // the context's own stream
    svtf_geom    max_geom{};
    svtf_geom    geom{};
    bool         have_geom = false, have_frame = false, have_mi = false, have_source = false, have_search = false;
    Surface      surf[4];  // filter chain ping-pong; [3] = loop-restoration output, allocated on first use
    Surface      source;   // source picture for the search
    // Stripe window (svtf_stripe_begin): the context holds picture rows [glob_row0, glob_row0 + geom.mi_rows*4) of the
    // FULL picture `full`; it owns local rows [own_lo, own_hi).  Whole-picture mode: glob_row0 = 0, own = everything.
    svtf_geom    full{};
    bool         striped = false;
    int          glob_row0 = 0, own_lo = 0, own_hi = 0;
    int          cur = 0;  // surface holding the current end of the chain
    int          pre_cdef = -1; // surface holding the picture before the last CDEF (-1: none)
    svtf_mi_t   *d_mi = nullptr;
    int          mi_stride = 0; // device mi stride == geom.mi_cols rounded
    uint32_t    *d_edge[2] = {nullptr, nullptr}; // deblock edge-code planes of d_mi (dbk_edge_plane_kernel): luma units, chroma units
    bool         edge_valid = false;             // false whenever d_mi was written since the planes were computed
    // Small per-picture parameters (CDEF strength index per filter block, CCSO LUT and block flags) live in ONE device
    // block and are sent with ONE DMA from a pinned staging slot, issued before the first kernel of a chain: a small
    // copy queued between two kernels would stall the H2D engine (and the next picture's upload behind it) until the
    // first kernel finishes.
    uint8_t     *d_params = nullptr;
    int8_t      *d_fb_idx = nullptr;     // = d_params
    int8_t      *d_ccso_lut = nullptr;   // = d_params + poff_lut
    uint8_t     *d_ccso_flags = nullptr; // = d_params + poff_flags, 3 * max_blocks
    int          max_ccso_blocks = 0;
    size_t       poff_lut = 0, poff_flags = 0, psize = 0;
    struct ParamSlot {
        uint8_t    *h = nullptr;
        cudaEvent_t ev = nullptr;
        bool        pending = false;
    } pslot[2];
    // The parameter DMA runs on its own stream: queued on the picture's stream behind a kernel it would sit in the (in-order, shared)
    // host-to-device copy queue until that kernel has run and hold every other context's upload behind it (measured: a 6 us kernel
    // ahead of a small copy cost 0.1 ms per 4K picture end to end).  ev_pfree = the last kernel that reads d_params has been queued.
    cudaStream_t pstream = nullptr;
    cudaEvent_t  ev_pfree = nullptr;
    bool         pfree_rec = false;
    int    pcur = 0;
    bool   popen = false;
    size_t plo = 0, phi = 0;
    uint64_t    *d_mse = nullptr;  // [2][nfb][64]
    unsigned long long *d_mse_raw = nullptr; // [3][nfb][64] raw sums of 128-wide / 128-high coding blocks (svtf_search.cuh), then [nfb] flags
    uint8_t     *d_skip = nullptr; // [nfb]
    uint8_t     *d_dir = nullptr;  // [nfb][64]
    int32_t     *d_var = nullptr;  // [nfb][64]
    int          max_nfb = 0;
    // device-side pick (svtf_pick.cuh)
    PickState   *d_pick_state = nullptr;
    uint64_t    *d_pick_tot = nullptr;   // [3][64*64] (the per-call pick uses the first)
    int8_t      *d_pick_idx = nullptr;   // [pick_cap]
    uint64_t    *d_tab_mse = nullptr;    // table uploaded by svtf_cdef_pick_table, [2][pick_cap][64]
    uint8_t     *d_tab_skip = nullptr;
    int          pick_cap = 0, tab_cap = 0;
    int          sm_count = 0;
    size_t       max_smem_optin = 0;
    LrUnitDev   *d_lr_units = nullptr;   // restoration units of the three planes (svtf_lr_apply_frame)
    size_t       lr_units_cap = 0;
    int         *h_mi_bad = nullptr;     // pinned + mapped: "invalid enum" flag of svtf_set_mi_soa_*, written by the pack kernel itself
    int         *d_mi_bad = nullptr;     // its device alias
    uint8_t     *d_mi_soa = nullptr;     // byte planes of svtf_set_mi_soa_* (7 x mi_rows x mi_stride) + 4 bytes error flag
    size_t       mi_soa_cap = 0;
    uint8_t     *d_sb_lvl = nullptr;     // per-superblock deblock level tables (svtf_deblock_frame_sb_delta)
    size_t       sb_lvl_cap = 0;
    void        *derive_scratch[3] = {nullptr, nullptr, nullptr}; // CCSO derivation buffers per plane (DeriveScratch), created on first use
    cudaStream_t derive_stream[3]  = {nullptr, nullptr, nullptr}; // the planes derive concurrently, each on its own stream
    cudaEvent_t  derive_ev         = nullptr;
    uint64_t     launches = 0;
    uint8_t     *d_lrs = nullptr;          // scratch of the loop-restoration filter search (svtf_lr_search_frame), allocated on first use
    size_t       lrs_cap = 0;
    uint8_t     *h_lrs = nullptr;          // pinned mirror of the parts of d_lrs that cross PCIe every refinement round
    size_t       h_lrs_cap = 0;
    int32_t     *d_lrs_flt = nullptr;      // filter outputs of the self-guided search's current ep (two int32 planes per plane)
    size_t       lrs_flt_cap = 0;
    uint8_t     *d_handoff = nullptr;      // staging of svtf_download_ref_padded / svtf_upload_frame_8as16, allocated on first use
    size_t       handoff_cap = 0;
    bool         search3_ready = false;    // same for cdef_search3_kernel
    bool         cdef5_ready = false;
    bool         cdef3_ready = false;      // dynamic shared-memory opt-in of cdef3_kernel done on this device
    unsigned    *d_sched = nullptr;        // tile scheduler of the persistent kernels: {next tile, CTAs done}, zero between launches
    int          cdef4_ctas_per_sm[4] = {0, 0, 0, 0}; // cdef4_kernel<u8/u16, int/fp16>: resident CTAs per SM (0 = not set up yet)
    std::string  err;

    int fail(int code, const std::string &m) {
        err = m;
        set_global_error(code, m.c_str());
        return code;
    }
    int cuda_fail(cudaError_t e, const char *what) {
        return fail(SVTF_E_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
    }
};

#define CUDA_TRY(ctx, call)                         \
    do {                                            \
        cudaError_t e_ = (call);                    \
        if (e_ != cudaSuccess)                      \
            return (ctx)->cuda_fail(e_, #call);     \
    } while (0)

static bool geom_ok(const svtf_geom *g) {
    return g && g->width > 0 && g->height > 0 && g->mi_cols > 0 && g->mi_rows > 0 && !(g->mi_cols & 1) && !(g->mi_rows & 1) &&
        g->width <= g->mi_cols * 4 && g->height <= g->mi_rows * 4 && (g->bit_depth == 8 || g->bit_depth == 10 || g->bit_depth == 12) &&
        (g->sample_bytes == 1 || g->sample_bytes == 2) && !(g->sample_bytes == 1 && g->bit_depth != 8);
}

static int round_up(int v, int a) { return (v + a - 1) / a * a; }

static cudaError_t alloc_surface(Surface &s, const svtf_geom &g) {
    for (int p = 0; p < 3; p++) {
        const int w = (g.mi_cols * 4) >> (p ? 1 : 0), h = (g.mi_rows * 4) >> (p ? 1 : 0);
        s.pitch[p]  = round_up(w, 128);
        cudaError_t e = cudaMalloc(&s.base[p], (size_t)s.pitch[p] * (h + 8) * g.sample_bytes);
        if (e != cudaSuccess)
            return e;
    }
    return cudaSuccess;
}

extern "C" const char *svtf_last_error(const svtf_ctx *ctx) {
    if (ctx)
        return ctx->err.c_str();
    static thread_local std::string copy;
    std::lock_guard<std::mutex> lk(g_err_mutex);
    copy = g_err;
    return copy.c_str();
}

extern "C" int svtf_ctx_create(int device, const svtf_geom *max_geom, svtf_ctx **out) {
    if (!out || !geom_ok(max_geom))
        return set_global_error(SVTF_E_ARG, "svtf_ctx_create: bad geometry");
    int         ndev = 0;
    cudaError_t e    = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev <= 0 || device < 0 || device >= ndev) {
        std::string m = "svtf_ctx_create: no usable CUDA device (this library has no CPU fallback): ";
        m += e != cudaSuccess ? cudaGetErrorString(e) : "device index out of range";
        return set_global_error(SVTF_E_CUDA, m.c_str());
    }
    svtf_ctx *c = new (std::nothrow) svtf_ctx();
    if (!c)
        return set_global_error(SVTF_E_NOMEM, "svtf_ctx_create: out of host memory");
    c->device   = device;
    c->max_geom = *max_geom;
    auto bail   = [&](cudaError_t ce, const char *what) {
        int r = c->cuda_fail(ce, what);
        svtf_ctx_destroy(c);
        return r;
    };
    if ((e = cudaSetDevice(device)) != cudaSuccess) return bail(e, "cudaSetDevice");
    if ((e = cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking)) != cudaSuccess) return bail(e, "cudaStreamCreate");
    c->stream = c->own_stream;
    for (int i = 0; i < 3; i++)
        if ((e = alloc_surface(c->surf[i], *max_geom)) != cudaSuccess) return bail(e, "cudaMalloc(surface)");
    if ((e = alloc_surface(c->source, *max_geom)) != cudaSuccess) return bail(e, "cudaMalloc(source)");
    const int nvfb = (max_geom->mi_rows + 15) / 16, nhfb = (max_geom->mi_cols + 15) / 16;
    c->max_nfb     = nvfb * nhfb;
    c->mi_stride   = round_up(max_geom->mi_cols, 32);
    if ((e = cudaMalloc(&c->d_mi, sizeof(svtf_mi_t) * (size_t)c->mi_stride * max_geom->mi_rows)) != cudaSuccess) return bail(e, "cudaMalloc(mi)");
    if ((e = cudaMalloc(&c->d_edge[0], sizeof(uint32_t) * (size_t)c->mi_stride * max_geom->mi_rows)) != cudaSuccess) return bail(e, "cudaMalloc(edge)");
    if ((e = cudaMalloc(&c->d_edge[1], sizeof(uint32_t) * (size_t)(c->mi_stride / 2) * ((max_geom->mi_rows + 1) / 2))) != cudaSuccess)
        return bail(e, "cudaMalloc(edge)");
    // flags are indexed in whole-picture coordinates even when the context holds one stripe: size for 16K rows
    c->max_ccso_blocks = ((max_geom->width + 255) / 256) * std::max((max_geom->height + 255) / 256, 64);
    c->poff_lut        = (size_t)round_up(c->max_nfb, 16);
    c->poff_flags      = c->poff_lut + 3 * SVTF_CCSO_LUT_SIZE;
    c->psize           = c->poff_flags + (size_t)round_up(3 * c->max_ccso_blocks, 16);
    if ((e = cudaMalloc(&c->d_params, c->psize)) != cudaSuccess) return bail(e, "cudaMalloc(params)");
    c->d_fb_idx     = reinterpret_cast<int8_t *>(c->d_params);
    c->d_ccso_lut   = reinterpret_cast<int8_t *>(c->d_params + c->poff_lut);
    c->d_ccso_flags = c->d_params + c->poff_flags;
    if ((e = cudaStreamCreateWithFlags(&c->pstream, cudaStreamNonBlocking)) != cudaSuccess) return bail(e, "cudaStreamCreate(params)");
    if ((e = cudaEventCreateWithFlags(&c->ev_pfree, cudaEventDisableTiming)) != cudaSuccess) return bail(e, "cudaEventCreate");
    for (auto &sl : c->pslot) {
        if ((e = cudaHostAlloc(reinterpret_cast<void **>(&sl.h), c->psize, cudaHostAllocDefault)) != cudaSuccess) return bail(e, "cudaHostAlloc(params)");
        if ((e = cudaEventCreateWithFlags(&sl.ev, cudaEventDisableTiming)) != cudaSuccess) return bail(e, "cudaEventCreate");
    }
    if ((e = cudaMalloc(&c->d_mse, sizeof(uint64_t) * 2 * 64 * (size_t)c->max_nfb)) != cudaSuccess) return bail(e, "cudaMalloc(mse)");
    if ((e = cudaMalloc(&c->d_mse_raw, (sizeof(unsigned long long) * 3 * 64 + sizeof(unsigned)) * (size_t)c->max_nfb)) != cudaSuccess)
        return bail(e, "cudaMalloc(mse raw)");
    if ((e = cudaMalloc(&c->d_skip, c->max_nfb)) != cudaSuccess) return bail(e, "cudaMalloc(skip)");
    if ((e = cudaMalloc(&c->d_dir, 64 * (size_t)c->max_nfb)) != cudaSuccess) return bail(e, "cudaMalloc(dir)");
    if ((e = cudaMalloc(&c->d_var, sizeof(int32_t) * 64 * (size_t)c->max_nfb)) != cudaSuccess) return bail(e, "cudaMalloc(var)");
    if ((e = cudaMalloc(&c->d_pick_state, sizeof(PickState))) != cudaSuccess) return bail(e, "cudaMalloc(pick state)");
    if ((e = cudaMalloc(&c->d_pick_tot, sizeof(uint64_t) * 3 * 64 * 64)) != cudaSuccess) return bail(e, "cudaMalloc(pick tot)");
    if ((e = cudaMalloc(&c->d_sched, 64)) != cudaSuccess) return bail(e, "cudaMalloc(sched)");
    if ((e = cudaMemset(c->d_sched, 0, 64)) != cudaSuccess) return bail(e, "cudaMemset(sched)");
    cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, device);
    *out = c;
    return SVTF_OK;
}

extern "C" void svtf_ctx_destroy(svtf_ctx *c) {
    if (!c)
        return;
    cudaSetDevice(c->device);
    if (c->stream)
        cudaStreamSynchronize(c->stream);
    for (int i = 0; i < 4; i++)
        for (int p = 0; p < 3; p++) cudaFree(c->surf[i].base[p]);
    for (int p = 0; p < 3; p++) cudaFree(c->source.base[p]);
    cudaFree(c->d_mi), cudaFree(c->d_edge[0]), cudaFree(c->d_edge[1]), cudaFree(c->d_params), cudaFree(c->d_sched);
    for (auto &sl : c->pslot) {
        if (sl.h) cudaFreeHost(sl.h);
        if (sl.ev) cudaEventDestroy(sl.ev);
    }
    if (c->pstream) cudaStreamDestroy(c->pstream);
    if (c->ev_pfree) cudaEventDestroy(c->ev_pfree);
    cudaFree(c->d_mse), cudaFree(c->d_mse_raw), cudaFree(c->d_skip), cudaFree(c->d_dir), cudaFree(c->d_var);
    for (int p = 0; p < 3; p++) {
        derive_scratch_free(c->derive_scratch[p]);
        if (c->derive_stream[p])
            cudaStreamDestroy(c->derive_stream[p]);
    }
    if (c->derive_ev)
        cudaEventDestroy(c->derive_ev);
    cudaFree(c->d_pick_state), cudaFree(c->d_pick_tot), cudaFree(c->d_pick_idx), cudaFree(c->d_tab_mse), cudaFree(c->d_tab_skip);
    cudaFree(c->d_sb_lvl), cudaFree(c->d_mi_soa), cudaFree(c->d_lr_units), cudaFree(c->d_handoff), cudaFree(c->d_lrs), cudaFree(c->d_lrs_flt);
    if (c->h_lrs) cudaFreeHost(c->h_lrs);
    if (c->h_mi_bad)
        cudaFreeHost(c->h_mi_bad);
    if (c->own_stream)
        cudaStreamDestroy(c->own_stream);
    delete c;
}

static int mi_soa_check(svtf_ctx *c);

extern "C" int svtf_sync(svtf_ctx *c) {
    if (!c)
        return set_global_error(SVTF_E_ARG, "svtf_sync: NULL context");
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    return mi_soa_check(c);
}

extern "C" int svtf_ctx_set_stream(svtf_ctx *c, void *cuda_stream) {
    if (!c)
        return set_global_error(SVTF_E_ARG, "svtf_ctx_set_stream: NULL context");
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    c->stream = cuda_stream ? (cudaStream_t)cuda_stream : c->own_stream;
    return SVTF_OK;
}

extern "C" uint64_t svtf_launch_count(const svtf_ctx *c) { return c ? c->launches : 0; }
extern "C" void    *svtf_cuda_stream(svtf_ctx *c) { return c ? (void *)c->stream : nullptr; }

extern "C" void *svtf_host_alloc(size_t bytes) {
    void *p = nullptr;
    if (cudaHostAlloc(&p, bytes, cudaHostAllocDefault) != cudaSuccess)
        return nullptr;
    return p;
}
extern "C" void svtf_host_free(void *p) {
    if (p)
        cudaFreeHost(p);
}

// Page-lock a caller-owned buffer in place (cudaHostRegister) so that uploads from / downloads into it are asynchronous DMA at full
// PCIe rate; a buffer that is already registered is fine.  Long-lived picture buffers (an encoder's recon pool) are registered once.
extern "C" int svtf_host_register(void *p, size_t bytes) {
    if (!p || !bytes)
        return set_global_error(SVTF_E_ARG, "svtf_host_register: NULL buffer");
    const cudaError_t e = cudaHostRegister(p, bytes, cudaHostRegisterPortable);
    if (e == cudaErrorHostMemoryAlreadyRegistered) {
        cudaGetLastError();
        return SVTF_OK;
    }
    if (e != cudaSuccess)
        return set_global_error(SVTF_E_CUDA, (std::string("cudaHostRegister: ") + cudaGetErrorString(e)).c_str());
    return SVTF_OK;
}
extern "C" int svtf_host_unregister(void *p) {
    if (!p)
        return set_global_error(SVTF_E_ARG, "svtf_host_unregister: NULL buffer");
    const cudaError_t e = cudaHostUnregister(p);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return set_global_error(SVTF_E_CUDA, (std::string("cudaHostUnregister: ") + cudaGetErrorString(e)).c_str());
    }
    return SVTF_OK;
}

extern "C" int svtf_frame_begin(svtf_ctx *c, const svtf_geom *g) {
    if (!c)
        return set_global_error(SVTF_E_ARG, "svtf_frame_begin: NULL context");
    if (!geom_ok(g) || g->mi_cols > c->max_geom.mi_cols || g->mi_rows > c->max_geom.mi_rows ||
        g->sample_bytes != c->max_geom.sample_bytes)
        return c->fail(SVTF_E_ARG, "svtf_frame_begin: geometry invalid or larger than the context's max_geom");
    CUDA_TRY(c, cudaSetDevice(c->device));
    c->geom      = *g;
    c->full      = *g;
    c->striped   = false;
    c->glob_row0 = 0, c->own_lo = 0, c->own_hi = g->mi_rows * 4;
    c->have_geom = true;
    c->have_frame = c->have_mi = c->have_source = c->have_search = false;
    c->cur       = 0;
    c->pre_cdef  = -1;
    return SVTF_OK;
}

// One horizontal stripe of a picture: the context owns superblock rows [first_sb_row, first_sb_row +
// num_sb_rows) (64 luma rows each) and keeps one extra SB row of context above and below where the picture continues.
// Every later call takes WHOLE-PICTURE host arrays (planes, mi grid, fb_strength_idx, CCSO flags, search outputs) and
// touches only this stripe's part of them, so N contexts on N GPUs called with the same arguments filter one picture.
// Bit-exactness: a horizontal edge reads <= 7 rows and writes <= 6 rows on each side and edges of one pass never
// overlap (Common/Codec/EbDeblockingCommon.c:787-823), CDEF taps reach +-2 rows (Common/Codec/EbCdef.c:99-122), CCSO
// +-1 row; with 8 input rows beyond the boundary the boundary edge is recomputed redundantly on both sides and every
// row the stripe's own CDEF/CCSO reads is exact.  No write-back, no exchange when the input comes from the host.
extern "C" int svtf_stripe_begin(svtf_ctx *c, const svtf_geom *full, int32_t first_sb_row, int32_t num_sb_rows) {
    if (!c)
        return set_global_error(SVTF_E_ARG, "svtf_stripe_begin: NULL context");
    if (!geom_ok(full) || full->sample_bytes != c->max_geom.sample_bytes)
        return c->fail(SVTF_E_ARG, "svtf_stripe_begin: bad picture geometry");
    const int rows = full->mi_rows * 4, sb_rows = (rows + 63) / 64;
    if (first_sb_row < 0 || num_sb_rows <= 0 || first_sb_row + num_sb_rows > sb_rows)
        return c->fail(SVTF_E_ARG, "svtf_stripe_begin: superblock rows outside the picture");
    const int ht = first_sb_row > 0, hb = first_sb_row + num_sb_rows < sb_rows;
    const int g0 = 64 * (first_sb_row - ht), g1 = std::min(rows, 64 * (first_sb_row + num_sb_rows + hb));
    svtf_geom l = *full;
    l.mi_rows   = (g1 - g0) / 4;
    l.height    = std::min(full->height, g1) - g0;
    if (l.height <= 0 || l.mi_cols > c->max_geom.mi_cols || l.mi_rows > c->max_geom.mi_rows)
        return c->fail(SVTF_E_ARG, "svtf_stripe_begin: stripe larger than the context's max_geom");
    CUDA_TRY(c, cudaSetDevice(c->device));
    c->geom      = l;
    c->full      = *full;
    c->striped   = true;
    c->glob_row0 = g0;
    c->own_lo    = 64 * ht;
    c->own_hi    = std::min(g1 - g0, 64 * (ht + num_sb_rows));
    c->have_geom = true;
    c->have_frame = c->have_mi = c->have_source = c->have_search = false;
    c->cur       = 0;
    c->pre_cdef  = -1;
    return SVTF_OK;
}

// Copies local luma rows [lo, hi) (and the matching chroma rows) between the caller's WHOLE-PICTURE planes and a surface.
static int copy_planes(svtf_ctx *c, Surface &s, const void *const src[3], const int32_t stride[3], bool to_device, int lo, int hi) {
    const svtf_geom &g = c->geom;
    for (int p = 0; p < 3; p++) {
        const int ss = p ? 1 : 0;
        const int w = (g.mi_cols * 4) >> ss, r0 = lo >> ss, h = (hi >> ss) - r0;
        if (!src[p] || stride[p] < w)
            return c->fail(SVTF_E_ARG, "plane pointer NULL or stride smaller than the coded width");
        if (h <= 0)
            continue;
        const size_t sb = g.sample_bytes;
        uint8_t     *dev = static_cast<uint8_t *>(s.base[p]) + (size_t)r0 * s.pitch[p] * sb;
        uint8_t     *hst = static_cast<uint8_t *>(const_cast<void *>(src[p])) + ((size_t)(c->glob_row0 >> ss) + r0) * stride[p] * sb;
        // same pitch on both sides: one linear DMA instead of a 2-D copy — for uploads, where the bytes between the coded width and the pitch
        // land in device padding; a download only when there are none (the host's columns [w, stride) are the caller's: picture border,
        // a neighbouring region)
        if (stride[p] == s.pitch[p] && (to_device || w == s.pitch[p])) {
            const size_t bytes = ((size_t)(h - 1) * s.pitch[p] + w) * sb;
            if (to_device)
                CUDA_TRY(c, cudaMemcpyAsync(dev, hst, bytes, cudaMemcpyHostToDevice, c->stream));
            else
                CUDA_TRY(c, cudaMemcpyAsync(hst, dev, bytes, cudaMemcpyDeviceToHost, c->stream));
            continue;
        }
        if (to_device)
            CUDA_TRY(c, cudaMemcpy2DAsync(dev, s.pitch[p] * sb, hst, stride[p] * sb, w * sb, h, cudaMemcpyHostToDevice, c->stream));
        else
            CUDA_TRY(c, cudaMemcpy2DAsync(hst, stride[p] * sb, dev, s.pitch[p] * sb, w * sb, h, cudaMemcpyDeviceToHost, c->stream));
    }
    return SVTF_OK;
}
// rows a stripe needs as input: its own plus 8 beyond each interior boundary
static int in_lo(const svtf_ctx *c) { return std::max(0, c->own_lo - 8); }
static int in_hi(const svtf_ctx *c) { return std::min(c->geom.mi_rows * 4, c->own_hi + 8); }

extern "C" int svtf_upload_frame(svtf_ctx *c, const void *y, int32_t sy, const void *cb, int32_t scb, const void *cr, int32_t scr) {
    if (!c)
        return set_global_error(SVTF_E_ARG, "svtf_upload_frame: NULL context");
    if (!c->have_geom)
        return c->fail(SVTF_E_STATE, "svtf_upload_frame: call svtf_frame_begin first");
    const void   *src[3] = {y, cb, cr};
    const int32_t st[3]  = {sy, scb, scr};
    c->cur = 0, c->pre_cdef = -1;
    int r = copy_planes(c, c->surf[0], src, st, true, in_lo(c), in_hi(c));
    if (r == SVTF_OK)
        c->have_frame = true;
    return r;
}

extern "C" int svtf_overwrite_current(svtf_ctx *c, const void *y, int32_t sy, const void *cb, int32_t scb, const void *cr, int32_t scr) {
    if (!c)
        return set_global_error(SVTF_E_ARG, "svtf_overwrite_current: NULL context");
    if (!c->have_frame)
        return c->fail(SVTF_E_STATE, "svtf_overwrite_current: no picture uploaded");
    const void   *src[3] = {y, cb, cr};
    const int32_t st[3]  = {sy, scb, scr};
    return copy_planes(c, c->surf[c->cur], src, st, true, in_lo(c), in_hi(c));
}

// The picture as it was before CDEF, for a caller that ran deblock / CDEF elsewhere: CCSO classifies on its luma, loop
// restoration takes its stripe borders from it.  Goes to a surface of its own; the current picture is not touched.
extern "C" int svtf_upload_pre_cdef(svtf_ctx *c, const void *y, int32_t sy, const void *cb, int32_t scb, const void *cr, int32_t scr) {
    if (!c)
        return set_global_error(SVTF_E_ARG, "svtf_upload_pre_cdef: NULL context");
    if (!c->have_frame)
        return c->fail(SVTF_E_STATE, "svtf_upload_pre_cdef: upload the current picture first");
    if (c->cur == 3)
        return c->fail(SVTF_E_STATE, "svtf_upload_pre_cdef: the current picture is a restoration output");
    const void   *src[3] = {y, cb, cr};
    const int32_t st[3]  = {sy, scb, scr};
    const int     slot   = c->cur == 1 ? 2 : 1;
    int           r      = copy_planes(c, c->surf[slot], src, st, true, in_lo(c), in_hi(c));
    if (r == SVTF_OK)
        c->pre_cdef = slot;
    return r;
}

extern "C" int svtf_upload_source(svtf_ctx *c, const void *y, int32_t sy, const void *cb, int32_t scb, const void *cr, int32_t scr) {
    if (!c)
        return set_global_error(SVTF_E_ARG, "svtf_upload_source: NULL context");
    if (!c->have_geom)
        return c->fail(SVTF_E_STATE, "svtf_upload_source: call svtf_frame_begin first");
    const void   *src[3] = {y, cb, cr};
    const int32_t st[3]  = {sy, scb, scr};
    int r = copy_planes(c, c->source, src, st, true, c->own_lo, c->own_hi);
    if (r == SVTF_OK)
        c->have_source = true;
    return r;
}

static int edge_planes(svtf_ctx *c, bool force);

extern "C" int svtf_set_mi(svtf_ctx *c, const svtf_mi_t *mi, int32_t mi_stride) {
    if (!c)
        return set_global_error(SVTF_E_ARG, "svtf_set_mi: NULL context");
    if (!c->have_geom)
        return c->fail(SVTF_E_STATE, "svtf_set_mi: call svtf_frame_begin first");
    if (!mi || mi_stride < c->geom.mi_cols)
        return c->fail(SVTF_E_ARG, "svtf_set_mi: NULL grid or stride smaller than mi_cols");
    mi += (size_t)(c->glob_row0 / 4) * mi_stride; // whole-picture grid -> this context's rows
    if (mi_stride == c->mi_stride) // same pitch: one linear DMA
        CUDA_TRY(c, cudaMemcpyAsync(c->d_mi, mi, sizeof(svtf_mi_t) * ((size_t)(c->geom.mi_rows - 1) * mi_stride + c->geom.mi_cols),
                                    cudaMemcpyHostToDevice, c->stream));
    else
        CUDA_TRY(c, cudaMemcpy2DAsync(c->d_mi, sizeof(svtf_mi_t) * c->mi_stride, mi, sizeof(svtf_mi_t) * mi_stride,
                                      sizeof(svtf_mi_t) * c->geom.mi_cols, c->geom.mi_rows, cudaMemcpyHostToDevice, c->stream));
    c->have_mi = true;
    return edge_planes(c, true); // the mode info's second half: the deblocking edge-code planes, behind the upload on the stream
}

// ---- side info as byte planes, packed on the device (svtf_mi_pack.cuh) --------------------------------------
static const MiPackTables &mi_pack_tables() {
    static MiPackTables   t;
    static std::once_flag once;
    std::call_once(once, [] {
        // every geometry combination through the host packers: the device words are the host words by construction
        for (int si = 0; si < 2; si++)
            for (int td = 0; td < 3; td++)
                for (int bs = 0; bs < 22; bs++) {
                    const uint8_t b = (uint8_t)bs, d = (uint8_t)td, sk = (uint8_t)si, rf = (uint8_t)si;
                    svtf_mi_t     w = 0;
                    svtf_pack_mi_enc(1, &b, &d, &sk, &rf, nullptr, nullptr, &w);
                    t.enc_geom[si][td][bs] = w & 0x3ffffu;
                }
        for (int bs = 0; bs < 22; bs++) {
            const uint8_t b = (uint8_t)bs, z = 0;
            svtf_mi_t     w = 0;
            svtf_pack_mi_dec(1, &b, &z, &z, nullptr, nullptr, nullptr, nullptr, &w);
            t.blk_geom[bs] = w & (0x3fu << 12);
        }
        for (int tx = 0; tx < 19; tx++) {
            const uint8_t b = 0, x = (uint8_t)tx, z = 0;
            svtf_mi_t     w = 0;
            svtf_pack_mi_dec(1, &b, &x, &z, nullptr, nullptr, nullptr, nullptr, &w);
            t.tx_y[tx] = w & 0x3fu;
            svtf_pack_mi_dec(1, &b, &z, &x, nullptr, nullptr, nullptr, nullptr, &w);
            t.tx_uv[tx] = w & (0x3fu << 6);
        }
        t.mode_lf = 0;
        for (int m = 0; m < 32; m++) {
            const uint8_t b = 0, z = 0, md = (uint8_t)m;
            svtf_mi_t     w = 0;
            svtf_pack_mi_dec(1, &b, &z, &z, nullptr, nullptr, &md, nullptr, &w);
            t.mode_lf |= ((w >> 23) & 1u) << m;
        }
    });
    return t;
}

static int set_mi_soa(svtf_ctx *c, const uint8_t *const planes[7], int nplanes, int32_t stride, bool dec, const char *who) {
    if (!c->have_geom)
        return c->fail(SVTF_E_STATE, std::string(who) + ": call svtf_frame_begin first");
    for (int p = 0; p < nplanes; p++)
        if (!planes[p])
            return c->fail(SVTF_E_ARG, std::string(who) + ": NULL plane");
    if (stride < c->geom.mi_cols)
        return c->fail(SVTF_E_ARG, std::string(who) + ": stride smaller than mi_cols");
    CUDA_TRY(c, cudaSetDevice(c->device));
    const int    rows = c->geom.mi_rows, cols = c->geom.mi_cols;
    const size_t plane_bytes = (size_t)rows * cols, need = 7 * plane_bytes;
    if (c->mi_soa_cap < need) {
        if (c->d_mi_soa)
            cudaFree(c->d_mi_soa);
        c->d_mi_soa = nullptr, c->mi_soa_cap = 0;
        CUDA_TRY(c, cudaMalloc(&c->d_mi_soa, need));
        c->mi_soa_cap = need;
    }
    // The "invalid enum" flag lives in mapped host memory and is stored by the kernel directly (only when something IS invalid): no
    // memset ahead of the copies and no 4-byte read-back behind the kernel — either would sit at the head of the shared, in-order
    // copy queues until this picture's own upload has finished and hold every other context's transfers behind it.
    if (!c->h_mi_bad) {
        CUDA_TRY(c, cudaHostAlloc(reinterpret_cast<void **>(&c->h_mi_bad), sizeof(int), cudaHostAllocMapped));
        *c->h_mi_bad = 0;
        CUDA_TRY(c, cudaHostGetDevicePointer(reinterpret_cast<void **>(&c->d_mi_bad), c->h_mi_bad, 0));
    }
    for (int p = 0; p < nplanes; p++) {
        const uint8_t *src = planes[p] + (size_t)(c->glob_row0 / 4) * stride; // whole-picture planes -> this context's rows
        if (stride == cols)
            CUDA_TRY(c, cudaMemcpyAsync(c->d_mi_soa + p * plane_bytes, src, plane_bytes, cudaMemcpyHostToDevice, c->stream));
        else
            CUDA_TRY(c, cudaMemcpy2DAsync(c->d_mi_soa + p * plane_bytes, cols, src, stride, cols, rows, cudaMemcpyHostToDevice, c->stream));
    }
    MiPackArgs a;
    const uint8_t *d = c->d_mi_soa;
    a.bsize = d;
    if (dec)
        a.tx0 = d + plane_bytes, a.tx1 = d + 2 * plane_bytes, a.skip = d + 3 * plane_bytes, a.ref0 = d + 4 * plane_bytes,
        a.mode = d + 5 * plane_bytes, a.seg = d + 6 * plane_bytes;
    else
        a.tx0 = d + plane_bytes, a.tx1 = nullptr, a.skip = d + 2 * plane_bytes, a.ref0 = d + 3 * plane_bytes, a.mode = d + 4 * plane_bytes,
        a.seg = d + 5 * plane_bytes;
    a.in_pitch = cols, a.rows = rows, a.cols = cols, a.out = c->d_mi, a.out_pitch = c->mi_stride, a.dec = dec, a.bad = c->d_mi_bad;
    a.t = mi_pack_tables();
    mi_pack_kernel<<<dim3((cols + MP_THREADS - 1) / MP_THREADS, rows), MP_THREADS, 0, c->stream>>>(a);
    CUDA_TRY(c, cudaGetLastError());
    c->launches++;
    // no host synchronisation here: the flag is looked at by the next svtf_sync / download
    c->have_mi = true;
    return edge_planes(c, true); // the mode info's second half: the deblocking edge-code planes, behind the upload on the stream
}

static int mi_soa_check(svtf_ctx *c) { // after a stream synchronisation
    if (c->h_mi_bad && *c->h_mi_bad) {
        *c->h_mi_bad = 0;
        c->have_mi   = false;
        return c->fail(SVTF_E_ARG, "svtf_set_mi_soa: invalid enum value in the side info (the picture was filtered with undefined side info)");
    }
    return SVTF_OK;
}

extern "C" int svtf_set_mi_soa_enc(svtf_ctx *c, const uint8_t *bsize, const uint8_t *tx_depth, const uint8_t *skip, const uint8_t *ref_frame0,
                                   const uint8_t *mode, const uint8_t *segment_id, int32_t stride) {
    if (!c)
        return set_global_error(SVTF_E_ARG, "svtf_set_mi_soa_enc: NULL context");
    const uint8_t *pl[7] = {bsize, tx_depth, skip, ref_frame0, mode, segment_id, nullptr};
    return set_mi_soa(c, pl, 6, stride, false, "svtf_set_mi_soa_enc");
}

extern "C" int svtf_set_mi_soa_dec(svtf_ctx *c, const uint8_t *bsize, const uint8_t *tx_size_y, const uint8_t *tx_size_uv, const uint8_t *skip,
                                   const uint8_t *ref_frame0, const uint8_t *mode, const uint8_t *segment_id, int32_t stride) {
    if (!c)
        return set_global_error(SVTF_E_ARG, "svtf_set_mi_soa_dec: NULL context");
    const uint8_t *pl[7] = {bsize, tx_size_y, tx_size_uv, skip, ref_frame0, mode, segment_id};
    return set_mi_soa(c, pl, 7, stride, true, "svtf_set_mi_soa_dec");
}

// The packed words as the device holds them (tests; [mi_rows][stride] words)
extern "C" int svtf_get_mi(svtf_ctx *c, svtf_mi_t *out, int32_t stride) {
    if (!c)
        return set_global_error(SVTF_E_ARG, "svtf_get_mi: NULL context");
    if (!c->have_mi || !out || stride < c->geom.mi_cols)
        return c->fail(SVTF_E_STATE, "svtf_get_mi: no side info set, NULL output or stride smaller than mi_cols");
    CUDA_TRY(c, cudaSetDevice(c->device));
    CUDA_TRY(c, cudaMemcpy2DAsync(out, sizeof(svtf_mi_t) * stride, c->d_mi, sizeof(svtf_mi_t) * c->mi_stride, sizeof(svtf_mi_t) * c->geom.mi_cols,
                                  c->geom.mi_rows, cudaMemcpyDeviceToHost, c->stream));
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    return SVTF_OK;
}

extern "C" int svtf_download_frame_async(svtf_ctx *c, void *y, int32_t sy, void *cb, int32_t scb, void *cr, int32_t scr) {
    if (!c)
        return set_global_error(SVTF_E_ARG, "svtf_download_frame: NULL context");
    if (!c->have_frame)
        return c->fail(SVTF_E_STATE, "svtf_download_frame: no picture uploaded");
    const void   *dst[3] = {y, cb, cr};
    const int32_t st[3]  = {sy, scb, scr};
    return copy_planes(c, c->surf[c->cur], dst, st, false, c->own_lo, c->own_hi);
}

extern "C" int svtf_download_frame(svtf_ctx *c, void *y, int32_t sy, void *cb, int32_t scb, void *cr, int32_t scr) {
    int r = svtf_download_frame_async(c, y, sy, cb, scb, cr, scr);
    if (r != SVTF_OK)
        return r;
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    return mi_soa_check(c);
}

// ---- recon hand-off (svtf_handoff.cuh) ---------------------------------------------------------------------------
static int handoff_reserve(svtf_ctx *c, size_t bytes) {
    if (bytes <= c->handoff_cap)
        return SVTF_OK;
    if (c->d_handoff) {
        CUDA_TRY(c, cudaStreamSynchronize(c->stream));
        CUDA_TRY(c, cudaFree(c->d_handoff));
        c->d_handoff = nullptr, c->handoff_cap = 0;
    }
    CUDA_TRY(c, cudaMalloc(&c->d_handoff, bytes));
    c->handoff_cap = bytes;
    return SVTF_OK;
}

extern "C" int svtf_download_ref_padded(svtf_ctx *c, const svtf_ref_layout *lay, void *const dst[3], const int32_t dst_stride[3],
                                        uint8_t *const dst8[3], const int32_t stride8[3], uint8_t *const dstn[3], const int32_t striden[3]) {
    if (!c)
        return set_global_error(SVTF_E_ARG, "svtf_download_ref_padded: NULL context");
    const bool resident = dst == nullptr; // kernel only: the padded planes stay in the context's staging block
    if (!lay || (!resident && !dst_stride))
        return c->fail(SVTF_E_ARG, "svtf_download_ref_padded: NULL argument");
    if (!c->have_frame)
        return c->fail(SVTF_E_STATE, "svtf_download_ref_padded: no picture uploaded");
    if (c->striped)
        return c->fail(SVTF_E_STATE, "svtf_download_ref_padded: not available on a striped context");
    const svtf_geom &g = c->geom;
    const int vw = lay->width - lay->pad_right, vh = lay->height - lay->pad_bottom;
    if (lay->org_x < 0 || lay->org_y < 0 || lay->pad_right < 0 || lay->pad_bottom < 0 || vw <= 0 || vh <= 0 || vw > g.mi_cols * 4 ||
        vh > g.mi_rows * 4)
        return c->fail(SVTF_E_ARG, "svtf_download_ref_padded: layout does not fit the picture");
    const bool want8 = dst8 != nullptr, wantn = dstn != nullptr;
    if ((want8 || wantn) && g.sample_bytes != 2)
        return c->fail(SVTF_E_ARG, "svtf_download_ref_padded: the 8-bit / 2-bit split needs 16-bit samples");
    if (wantn && g.bit_depth != 10)
        return c->fail(SVTF_E_ARG, "svtf_download_ref_padded: the 2-bit plane exists for 10-bit pictures only");
    if (!resident && ((want8 && !stride8) || (wantn && !striden)))
        return c->fail(SVTF_E_ARG, "svtf_download_ref_padded: missing strides");
    CUDA_TRY(c, cudaSetDevice(c->device));
    HandoffArgs  a;
    const size_t sb = g.sample_bytes;
    size_t       off = 0, o16[3], o8[3], on[3];
    int          items = 0;
    for (int p = 0; p < 3; p++) {
        const int     ss = p ? 1 : 0;
        HandoffPlane &q  = a.p[p];
        q.valid_w = (vw + ss) >> ss, q.valid_h = (vh + ss) >> ss;
        q.org_x = (lay->org_x + ss) >> ss, q.org_y = (lay->org_y + ss) >> ss;
        q.out_w = ((lay->width + ss) >> ss) + 2 * q.org_x, q.out_h = ((lay->height + ss) >> ss) + 2 * q.org_y;
        q.dpitch = (q.out_w + 7) & ~7, q.chunks = q.dpitch / 8;
        q.first_item = items;
        items += q.chunks * q.out_h;
        if (!resident &&
            (!dst[p] || dst_stride[p] < q.out_w || (want8 && (!dst8[p] || stride8[p] < q.out_w)) || (wantn && (!dstn[p] || striden[p] < q.out_w))))
            return c->fail(SVTF_E_ARG, "svtf_download_ref_padded: plane pointer NULL or stride smaller than the padded width");
        const size_t n = (size_t)q.dpitch * q.out_h;
        o16[p] = off, off += (n * sb + 255) & ~(size_t)255;
        o8[p] = off, off += want8 ? (n + 255) & ~(size_t)255 : 0;
        on[p] = off, off += wantn ? (n + 255) & ~(size_t)255 : 0;
    }
    if (int r = handoff_reserve(c, off))
        return r;
    const Planes cur = c->surf[c->cur].planes();
    for (int p = 0; p < 3; p++) {
        a.p[p].src = cur.p[p], a.p[p].src_pitch = cur.pitch[p];
        a.p[p].dst  = c->d_handoff + o16[p];
        a.p[p].dst8 = want8 ? c->d_handoff + o8[p] : nullptr;
        a.p[p].dstn = wantn ? c->d_handoff + on[p] : nullptr;
    }
    a.total  = items;
    a.shift8 = g.bit_depth == 10 ? 2 : 0;
    const int blocks = std::min((items + 255) / 256, std::max(1, c->sm_count) * 8);
    if (sb == 1)
        handoff_kernel<uint8_t><<<blocks, 256, 0, c->stream>>>(a);
    else
        handoff_kernel<uint16_t><<<blocks, 256, 0, c->stream>>>(a);
    CUDA_TRY(c, cudaGetLastError());
    c->launches++;
    if (resident)
        return SVTF_OK;
    for (int p = 0; p < 3; p++) {
        const HandoffPlane &q = a.p[p];
        CUDA_TRY(c, cudaMemcpy2DAsync(dst[p], (size_t)dst_stride[p] * sb, q.dst, (size_t)q.dpitch * sb, (size_t)q.out_w * sb, q.out_h,
                                      cudaMemcpyDeviceToHost, c->stream));
        if (want8)
            CUDA_TRY(c, cudaMemcpy2DAsync(dst8[p], stride8[p], q.dst8, q.dpitch, q.out_w, q.out_h, cudaMemcpyDeviceToHost, c->stream));
        if (wantn)
            CUDA_TRY(c, cudaMemcpy2DAsync(dstn[p], striden[p], q.dstn, q.dpitch, q.out_w, q.out_h, cudaMemcpyDeviceToHost, c->stream));
    }
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    return mi_soa_check(c);
}

extern "C" int svtf_upload_frame_8as16(svtf_ctx *c, const uint8_t *y, int32_t sy, const uint8_t *cb, int32_t scb, const uint8_t *cr, int32_t scr) {
    if (!c)
        return set_global_error(SVTF_E_ARG, "svtf_upload_frame_8as16: NULL context");
    if (!c->have_geom)
        return c->fail(SVTF_E_STATE, "svtf_upload_frame_8as16: call svtf_frame_begin first");
    if (c->geom.sample_bytes != 2 || c->geom.bit_depth != 8)
        return c->fail(SVTF_E_ARG, "svtf_upload_frame_8as16: needs a context with 8-bit content in 16-bit samples");
    if (c->striped)
        return c->fail(SVTF_E_STATE, "svtf_upload_frame_8as16: not available on a striped context");
    const uint8_t *src[3] = {y, cb, cr};
    const int32_t  st[3]  = {sy, scb, scr};
    CUDA_TRY(c, cudaSetDevice(c->device));
    WidenArgs a;
    size_t    off = 0, o[3];
    int       items = 0;
    for (int p = 0; p < 3; p++) {
        const int ss = p ? 1 : 0;
        a.w[p] = (c->geom.mi_cols * 4) >> ss, a.h[p] = (c->geom.mi_rows * 4) >> ss;
        if (!src[p] || st[p] < a.w[p])
            return c->fail(SVTF_E_ARG, "plane pointer NULL or stride smaller than the coded width");
        a.spitch[p] = (a.w[p] + 7) & ~7, a.chunks[p] = a.spitch[p] / 8;
        a.first_item[p] = items;
        items += a.chunks[p] * a.h[p];
        o[p] = off, off += ((size_t)a.spitch[p] * a.h[p] + 255) & ~(size_t)255;
    }
    if (int r = handoff_reserve(c, off))
        return r;
    Planes d = c->surf[0].planes();
    for (int p = 0; p < 3; p++) {
        CUDA_TRY(c, cudaMemcpy2DAsync(c->d_handoff + o[p], a.spitch[p], src[p], st[p], a.w[p], a.h[p], cudaMemcpyHostToDevice, c->stream));
        a.src[p] = c->d_handoff + o[p], a.dst[p] = static_cast<uint16_t *>(d.p[p]), a.dpitch[p] = d.pitch[p];
    }
    a.total = items;
    widen_kernel<<<std::min((items + 255) / 256, std::max(1, c->sm_count) * 8), 256, 0, c->stream>>>(a);
    CUDA_TRY(c, cudaGetLastError());
    c->launches++;
    c->cur = 0, c->pre_cdef = -1;
    c->have_frame = true;
    return SVTF_OK;
}

extern "C" int svtf_rewind(svtf_ctx *c) {
    if (!c)
        return set_global_error(SVTF_E_ARG, "svtf_rewind: NULL context");
    if (!c->have_frame)
        return c->fail(SVTF_E_STATE, "svtf_rewind: no picture uploaded");
    c->cur = 0, c->pre_cdef = -1;
    return SVTF_OK;
}

static Dims dims_of(const svtf_geom &g) {
    Dims d;
    d.width = g.width, d.height = g.height, d.mi_cols = g.mi_cols, d.mi_rows = g.mi_rows, d.bit_depth = g.bit_depth;
    return d;
}

// surface to write next: any surface that is neither the current one nor the pre-CDEF one nor the upload (0)
static int next_surface(const svtf_ctx *c) {
    for (int i = 1; i < 3; i++)
        if (i != c->cur && i != c->pre_cdef)
            return i;
    return c->cur == 1 ? 2 : 1;
}

static int need_frame_mi(svtf_ctx *c, const char *fn) {
    if (!c->have_frame || !c->have_mi)
        return c->fail(SVTF_E_STATE, std::string(fn) + ": upload a picture and its mi grid first");
    return SVTF_OK;
}

// Step 1 of set_lpf_parameters (svtf_kernels.cuh, dbk_edge_entry): once per mode-info upload (svtf_set_mi / svtf_set_mi_soa_* launch it
// behind the copy; where the grid is assembled from pieces — stripes, halo exchange — the first deblock launch does); every deblock
// launch on that mode info (24 level-search trials + the final one) reads the planes.
static int edge_planes(svtf_ctx *c, bool force) {
    if (c->edge_valid && !force)
        return SVTF_OK;
    EdgePlaneArgs ea;
    ea.mi = c->d_mi, ea.mi_stride = c->mi_stride, ea.mi_cols = c->geom.mi_cols, ea.mi_rows = c->geom.mi_rows;
    ea.ey = c->d_edge[0], ea.ec = c->d_edge[1], ea.ey_pitch = c->mi_stride, ea.ec_pitch = c->mi_stride / 2;
    dbk_edge_plane_kernel<<<dim3((c->geom.mi_cols + 63) / 64, (c->geom.mi_rows + 3) / 4), 256, 0, c->stream>>>(ea);
    CUDA_TRY(c, cudaGetLastError());
    c->launches++;
    c->edge_valid = true;
    return SVTF_OK;
}

extern "C" int svtf_deblock_prepare(svtf_ctx *c) {
    if (!c)
        return set_global_error(SVTF_E_ARG, "svtf_deblock_prepare: NULL context");
    if (!c->have_mi)
        return c->fail(SVTF_E_STATE, "svtf_deblock_prepare: upload the mi grid first");
    CUDA_TRY(c, cudaSetDevice(c->device));
    return edge_planes(c, true);
}

// only_selected: launch no tiles for planes outside the mask (their samples in the output surface are then undefined;
// used by the level search, which measures the filtered plane alone) instead of copying them through
// sse_out: level-search trial — ONE plane (plane_start) is filtered (or, at level 0, left alone) and measured against the source
// into *sse_out instead of being written; the chain position does not move.
static int deblock_impl(svtf_ctx *c, const svtf_lf_params *lf, int32_t plane_start, int32_t plane_end, bool only_selected,
                        const uint8_t *d_sb_lvl = nullptr, int sb_cols = 0, int sb_shift = 0, unsigned long long *sse_out = nullptr) {
    if (!lf || plane_start < 0 || plane_end > 3 || plane_start > plane_end)
        return c->fail(SVTF_E_ARG, "svtf_deblock_frame: bad arguments");
    if (int r = need_frame_mi(c, "svtf_deblock_frame"))
        return r;
    // plane selection — svt_aom_loop_filter_sb, Encoder/Codec/EbDeblockingFilter.c:572-578: luma level 0
    // breaks the plane loop (chroma skipped too), a zero chroma level skips that plane.
    unsigned mask = 0;
    for (int p = plane_start; p < plane_end; p++) {
        if (p == 0 && !lf->filter_level[0] && !lf->filter_level[1])
            break;
        if (p == 1 && !lf->filter_level_u)
            continue;
        if (p == 2 && !lf->filter_level_v)
            continue;
        mask |= 1u << p;
    }
    if (!mask && !sse_out)
        return SVTF_OK; // nothing to filter: the picture stays where it is
    CUDA_TRY(c, cudaSetDevice(c->device));
    DeblockArgs a;
    a.sse_src = nullptr, a.sse_pitch = a.sse_w = a.sse_h = 0, a.sse_out = sse_out;
    const int   nxt = next_surface(c);
    a.in        = c->surf[c->cur].planes();
    a.out       = c->surf[nxt].planes();
    a.mi        = c->d_mi;
    a.mi_stride = c->mi_stride;
    a.d         = dims_of(c->geom);
    if (int r = edge_planes(c, false))
        return r;
    a.edge[0] = c->d_edge[0], a.edge[1] = c->d_edge[1], a.edge_pitch[0] = c->mi_stride, a.edge_pitch[1] = c->mi_stride / 2;
    // "no deblocking is required" limit of set_lpf_parameters: a chroma unit has edges while its LUMA position is inside the picture, i.e.
    // up to (w + 1) >> 1 — one 4x4 column more than w >> 1 when a super-resolution frame is 8k + 1 samples wide.  Encoder: the x_range /
    // y_range cut of svt_av1_filter_block_plane_vert/_horz for scaled frames (Encoder/Codec/EbDeblockingFilter.c:305-322; unscaled pictures
    // have even sizes); decoder: dec_set_lpf_parameters tests luma coordinates against frame_width / frame_height (EbDecLF.c:116-121).
    for (int p = 0; p < 3; p++) {
        const int ss = p ? 1 : 0;
        a.lim_w[p] = (c->geom.width + ss) >> ss, a.lim_h[p] = (c->geom.height + ss) >> ss;
    }
    int total   = 0;
    // rows to produce: the stripe's own plus the 8 beyond each interior boundary that its CDEF/CCSO read (tile t
    // of a plane writes rows [64t-8, 64t+56) of that plane)
    const int rows = c->geom.mi_rows * 4;
    const int dlo = c->own_lo > 0 ? c->own_lo - 8 : 0, dhi = c->own_hi < rows ? c->own_hi + 8 : rows;
    for (int p = 0; p < 3; p++) {
        const int ss = p ? 1 : 0;
        const int w = (c->geom.mi_cols * 4) >> ss;
        a.tiles_x[p]   = (w + 8 + DBK_TILE - 1) / DBK_TILE;
        a.ty0[p]       = ((dlo >> ss) + 8) / DBK_TILE;
        a.tiles_y[p]   = ((dhi >> ss) - 1 + 8) / DBK_TILE + 1 - a.ty0[p];
        if (only_selected && !((mask >> p) & 1) && !(sse_out && p == plane_start))
            a.tiles_y[p] = 0;
        a.tile_base[p] = total;
        total += a.tiles_x[p] * a.tiles_y[p];
    }
    a.plane_mask = mask;
    a.sb_lvl = d_sb_lvl, a.sb_cols = sb_cols, a.sb_shift = sb_shift;
    build_lf_tables(*lf, a.t);
    if (sse_out) {
        const int p = plane_start, ss = p ? 1 : 0;
        a.sse_src = c->source.base[p], a.sse_pitch = c->source.pitch[p];
        // picture_sse_calculations (Encoder/Codec/EbDeblockingFilter.c:716-835): 8-bit pictures are measured over the aligned
        // size, 16-bit ones over the visible size
        if (c->geom.sample_bytes == 1)
            a.sse_w = (c->geom.mi_cols * 4) >> ss, a.sse_h = (c->geom.mi_rows * 4) >> ss;
        else
            a.sse_w = (c->geom.width + ss) >> ss, a.sse_h = (c->geom.height + ss) >> ss;
        if (c->geom.sample_bytes == 1)
            deblock_kernel<uint8_t, false, true><<<total, DBK_THREADS, 0, c->stream>>>(a);
        else
            deblock_kernel<uint16_t, false, true><<<total, DBK_THREADS, 0, c->stream>>>(a);
        CUDA_TRY(c, cudaGetLastError());
        c->launches++;
        return SVTF_OK;
    }
    if (d_sb_lvl) {
        if (c->geom.sample_bytes == 1)
            deblock_kernel<uint8_t, true><<<total, DBK_THREADS, 0, c->stream>>>(a);
        else
            deblock_kernel<uint16_t, true><<<total, DBK_THREADS, 0, c->stream>>>(a);
    } else if (c->geom.sample_bytes == 1)
        deblock_kernel<uint8_t><<<total, DBK_THREADS, 0, c->stream>>>(a);
    else
        deblock_kernel<uint16_t><<<total, DBK_THREADS, 0, c->stream>>>(a);
    CUDA_TRY(c, cudaGetLastError());
    c->launches++;
    c->cur = nxt;
    return SVTF_OK;
}

extern "C" int svtf_deblock_frame(svtf_ctx *c, const svtf_lf_params *lf, int32_t plane_start, int32_t plane_end) {
    if (!c)
        return set_global_error(SVTF_E_ARG, "svtf_deblock_frame: NULL context");
    return deblock_impl(c, lf, plane_start, plane_end, false);
}

// Deblock with per-superblock loop-filter deltas (the decoder's SBInfo.sb_delta_lf, Decoder/Codec/EbDecLF.c:159-193): one
// level table per superblock, built on the host exactly like the frame table with that superblock's deltas in place of
// curr_delta_lf, looked up by the superblock containing each 4x4.
extern "C" int svtf_deblock_frame_sb_delta(svtf_ctx *c, const svtf_lf_params *lf, int32_t plane_start, int32_t plane_end,
                                           const int32_t *sb_delta_lf, int32_t sb_cols, int32_t sb_size_log2) {
    if (!c)
        return set_global_error(SVTF_E_ARG, "svtf_deblock_frame_sb_delta: NULL context");
    if (!lf || !sb_delta_lf || (sb_size_log2 != 6 && sb_size_log2 != 7))
        return c->fail(SVTF_E_ARG, "svtf_deblock_frame_sb_delta: bad arguments");
    if (c->striped)
        return c->fail(SVTF_E_STATE, "svtf_deblock_frame_sb_delta: not available on a striped context");
    if (!lf->delta_lf_present)
        return deblock_impl(c, lf, plane_start, plane_end, false);
    const int sb_mi = 1 << (sb_size_log2 - 2);
    const int cols = (c->geom.mi_cols + sb_mi - 1) / sb_mi, rows = (c->geom.mi_rows + sb_mi - 1) / sb_mi;
    if (sb_cols != cols)
        return c->fail(SVTF_E_ARG, "svtf_deblock_frame_sb_delta: sb_cols does not match the frame");
    const size_t bytes = (size_t)rows * cols * 768;
    CUDA_TRY(c, cudaSetDevice(c->device));
    if (c->sb_lvl_cap < bytes) {
        if (c->d_sb_lvl)
            cudaFree(c->d_sb_lvl);
        c->d_sb_lvl = nullptr, c->sb_lvl_cap = 0;
        CUDA_TRY(c, cudaMalloc(&c->d_sb_lvl, bytes));
        c->sb_lvl_cap = bytes;
    }
    std::vector<uint8_t> host(bytes);
    svtf_lf_params       t = *lf;
    uint8_t              lim[64], mblim[64], hev[64];
    for (int sb = 0; sb < rows * cols; sb++) {
        for (int i = 0; i < 4; i++) t.curr_delta_lf[i] = sb_delta_lf[sb * 4 + i];
        if (int r = svtf_lf_build_tables(&t, &host[(size_t)sb * 768], lim, mblim, hev))
            return c->fail(r, "svtf_deblock_frame_sb_delta: bad loop-filter parameters");
    }
    CUDA_TRY(c, cudaMemcpyAsync(c->d_sb_lvl, host.data(), bytes, cudaMemcpyHostToDevice, c->stream));
    CUDA_TRY(c, cudaStreamSynchronize(c->stream)); // `host` goes out of scope
    return deblock_impl(c, lf, plane_start, plane_end, false, c->d_sb_lvl, cols, sb_size_log2 - 2);
}

// ---- CCSO parameters -> device (LUT + block flags) --------------------------------------------------
static int ccso_check(svtf_ctx *c, const svtf_ccso_params *cc, const char *who) {
    for (int p = 0; p < 3; p++)
        if (cc->quant_idx[p] > 3 || cc->ext_filter_support[p] > 5 || cc->max_band_log2[p] > 7 || cc->edge_clf[p] > 1) {
            char msg[96];
            snprintf(msg, sizeof(msg), "%s: CCSO parameter out of range", who);
            return c->fail(SVTF_E_ARG, msg);
        }
    return SVTF_OK;
}

// ---- parameter staging: host slot -> one DMA ---------------------------------------------------------
static int params_open(svtf_ctx *c) {
    if (c->popen)
        return SVTF_OK;
    svtf_ctx::ParamSlot &sl = c->pslot[c->pcur];
    if (sl.pending) { // the DMA that last read this slot (two chains ago) must be done before it is rewritten
        CUDA_TRY(c, cudaEventSynchronize(sl.ev));
        sl.pending = false;
    }
    c->popen = true, c->plo = c->psize, c->phi = 0;
    return SVTF_OK;
}
static void params_put(svtf_ctx *c, size_t off, const void *src, size_t n) {
    memcpy(c->pslot[c->pcur].h + off, src, n);
    c->plo = std::min(c->plo, off), c->phi = std::max(c->phi, off + n);
}
static int params_flush(svtf_ctx *c) {
    if (!c->popen)
        return SVTF_OK;
    svtf_ctx::ParamSlot &sl = c->pslot[c->pcur];
    c->popen = false;
    if (c->phi > c->plo) {
        if (c->pfree_rec) // the kernels that read the previous contents (this context's last CDEF / CCSO launch) come first
            CUDA_TRY(c, cudaStreamWaitEvent(c->pstream, c->ev_pfree, 0));
        CUDA_TRY(c, cudaMemcpyAsync(c->d_params + c->plo, sl.h + c->plo, c->phi - c->plo, cudaMemcpyHostToDevice, c->pstream));
        CUDA_TRY(c, cudaEventRecord(sl.ev, c->pstream));
        CUDA_TRY(c, cudaStreamWaitEvent(c->stream, sl.ev, 0)); // whatever is launched on the picture's stream from here on sees them
        sl.pending = true;
        c->pcur ^= 1;
    }
    return SVTF_OK;
}

// after the launch of a kernel that reads d_params (fb_idx, CCSO LUTs / flags)
static int params_consumed(svtf_ctx *c) {
    CUDA_TRY(c, cudaEventRecord(c->ev_pfree, c->stream));
    c->pfree_rec = true;
    return SVTF_OK;
}

static int ccso_nblocks(const svtf_ctx *c, int p) {
    const int pw = c->full.width >> (p ? 1 : 0), ph = c->full.height >> (p ? 1 : 0), bl = p ? 7 : 8;
    return ((pw + (1 << bl) - 1) >> bl) * ((ph + (1 << bl) - 1) >> bl);
}
static void ccso_put(svtf_ctx *c, const svtf_ccso_params *cc, const uint8_t *const blk_flags[3]) {
    params_put(c, c->poff_lut, cc->filter_offset, 3 * SVTF_CCSO_LUT_SIZE);
    for (int p = 0; p < 3; p++)
        if (blk_flags && blk_flags[p])
            params_put(c, c->poff_flags + (size_t)p * c->max_ccso_blocks, blk_flags[p], ccso_nblocks(c, p));
}
static void ccso_fill(svtf_ctx *c, const svtf_ccso_params *cc, const uint8_t *const blk_flags[3], CcsoFuse &f) {
    static const uint8_t quant_sz[4] = {16, 8, 32, 64}; // Common/Codec/EbCcso.c:636
    for (int p = 0; p < 3; p++) {
        f.enable[p] = cc->ccso_enable[p] != 0, f.bo_only[p] = cc->ccso_bo_only[p] != 0;
        f.thr[p] = quant_sz[cc->quant_idx[p]], f.support[p] = cc->ext_filter_support[p];
        f.edge_clf[p] = cc->edge_clf[p], f.band_log2[p] = cc->max_band_log2[p];
        f.blk_flags[p] = (blk_flags && blk_flags[p]) ? c->d_ccso_flags + (size_t)p * c->max_ccso_blocks : nullptr;
    }
    f.lut    = c->d_ccso_lut;
    f.on     = 1;
    f.y_glob = c->glob_row0;
}
// whole-picture per-filter-block array -> the rows this context holds
static const int8_t *fb_local(const svtf_ctx *c, const int8_t *fb_strength_idx) {
    return fb_strength_idx + (size_t)(c->glob_row0 / 64) * ((c->geom.mi_cols + 15) / 16);
}
static int fb_idx_check(svtf_ctx *c, const int8_t *fb_strength_idx) {
    const int nfb = ((c->geom.mi_rows + 15) / 16) * ((c->geom.mi_cols + 15) / 16);
    fb_strength_idx = fb_local(c, fb_strength_idx);
    for (int i = 0; i < nfb; i++)
        if (fb_strength_idx[i] >= SVTF_CDEF_MAX_STRENGTHS)
            return c->fail(SVTF_E_ARG, "svtf_cdef_apply_frame: strength index out of range");
    return SVTF_OK;
}

// ---- cdef4_kernel: persistent CTAs fed by TMA (svtf_cdef4.cuh) ---------------------------------------------------------------
// cuTensorMapEncodeTiled through the runtime's driver entry point lookup (the library does not link libcuda).
typedef CUresult (*svtf_encode_tiled_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *, const cuuint32_t *,
                                         const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static svtf_encode_tiled_fn encode_tiled_fn() {
    static svtf_encode_tiled_fn fn = [] {
        void                           *p = nullptr;
        cudaDriverEntryPointQueryResult q = cudaDriverEntryPointSymbolNotFound;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
            p = nullptr;
        return reinterpret_cast<svtf_encode_tiled_fn>(p);
    }();
    return fn;
}
// Tensor map of one plane: dims = the coded plane (TMA zero-fills what a box reads beyond it), box = one staged tile.
static int plane_tensor_map(svtf_ctx *c, CUtensorMap *tm, void *base, int pitch, int w, int h, int box_w, int box_h) {
    svtf_encode_tiled_fn enc = encode_tiled_fn();
    if (!enc)
        return c->fail(SVTF_E_CUDA, "cuTensorMapEncodeTiled is not available from this driver");
    const size_t     sb = c->geom.sample_bytes;
    const cuuint64_t dims[2] = {(cuuint64_t)w, (cuuint64_t)h}, strides[1] = {(cuuint64_t)pitch * sb};
    const cuuint32_t box[2] = {(cuuint32_t)box_w, (cuuint32_t)box_h}, es[2] = {1, 1};
    const CUresult   r = enc(tm, sb == 1 ? CU_TENSOR_MAP_DATA_TYPE_UINT8 : CU_TENSOR_MAP_DATA_TYPE_UINT16, 2, base, dims, strides, box, es,
                             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        char msg[96];
        snprintf(msg, sizeof(msg), "cuTensorMapEncodeTiled failed (CUresult %d)", (int)r);
        return c->fail(SVTF_E_CUDA, msg);
    }
    return SVTF_OK;
}
template <typename T, bool FP>
static int cdef4_launch_t(svtf_ctx *c, const Cdef4Args &g, int which) {
    constexpr int smem = c4_smem_bytes<T, FP>();
    if (!c->cdef4_ctas_per_sm[which]) {
        CUDA_TRY(c, cudaFuncSetAttribute(cdef4_kernel<T, FP>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        CUDA_TRY(c, cudaFuncSetAttribute(cdef4_kernel<T, FP>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
        int n = 0;
        CUDA_TRY(c, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, cdef4_kernel<T, FP>, C2_THREADS, smem));
        c->cdef4_ctas_per_sm[which] = std::max(n, 1);
    }
    // one CTA per SM slot; tiles beyond the first of each CTA are drawn from a counter
    const int ctas = std::min(g.ntiles, std::max(1, c->sm_count) * c->cdef4_ctas_per_sm[which]);
    cdef4_kernel<T, FP><<<ctas, C2_THREADS, smem, c->stream>>>(g);
    return SVTF_OK;
}
static int cdef4_launch(svtf_ctx *c, const Cdef2Args &b, bool fp, int nhfb, int nvfb_launch) {
    Cdef4Args g;
    g.c      = b;
    g.nhfb   = nhfb;
    g.ntiles = nhfb * nvfb_launch;
    g.sched  = c->d_sched;
    if (g.ntiles <= 0)
        return SVTF_OK;
    const Planes in = b.in;
    for (int p = 0; p < 3; p++) {
        const int ss = p ? 1 : 0;
        if (int r = plane_tensor_map(c, &g.tm[p], in.p[p], in.pitch[p], (c->geom.mi_cols * 4) >> ss, (c->geom.mi_rows * 4) >> ss,
                                     c->geom.sample_bytes == 1 ? (p ? C4Raw<uint8_t>::C_W : C4Raw<uint8_t>::Y_W) : (p ? C4Raw<uint16_t>::C_W : C4Raw<uint16_t>::Y_W),
                                     p ? C4_CBOX_H : C4_YBOX_H))
            return r;
    }
    if (c->geom.sample_bytes == 1)
        return fp ? cdef4_launch_t<uint8_t, true>(c, g, 0) : cdef4_launch_t<uint8_t, false>(c, g, 1);
    return fp ? cdef4_launch_t<uint16_t, true>(c, g, 2) : cdef4_launch_t<uint16_t, false>(c, g, 3);
}

// CDEF apply; with `cc` the CCSO pass of the same picture runs in the kernel's epilogue (one launch, one write).
static int cdef_apply_impl(svtf_ctx *c, const svtf_cdef_params *cp, const int8_t *fb_strength_idx, const svtf_ccso_params *cc,
                           const uint8_t *const blk_flags[3], const char *who, bool staged) {
    if (int r = need_frame_mi(c, who))
        return r;
    CUDA_TRY(c, cudaSetDevice(c->device));
    const int nvfb = (c->geom.mi_rows + 15) / 16, nhfb = (c->geom.mi_cols + 15) / 16;
    if (!staged) {
        if (int r = fb_idx_check(c, fb_strength_idx))
            return r;
        if (int r = params_open(c))
            return r;
        params_put(c, 0, fb_local(c, fb_strength_idx), (size_t)nvfb * nhfb);
        if (cc)
            ccso_put(c, cc, blk_flags);
        if (int r = params_flush(c))
            return r;
    }
    const int nxt  = next_surface(c);
    const int fbr0 = c->own_lo / 64;
    dim3      grid(nhfb, (c->own_hi + 63) / 64 - fbr0);
    (void)nvfb;
#ifdef SVTF_AB_KERNELS
    static const bool use_v1 = getenv("SVTF_CDEF_SCALAR") != nullptr; // A/B switch: the round-1 scalar kernel
    static const bool use_v2 = getenv("SVTF_CDEF_V2") != nullptr;     // A/B switch: the round-1b kernel
    if (use_v1 && !cc) {
        CdefArgs a;
        a.in = c->surf[c->cur].planes(), a.out = c->surf[nxt].planes();
        a.mi = c->d_mi, a.mi_stride = c->mi_stride, a.d = dims_of(c->geom), a.fb_idx = c->d_fb_idx, a.fbr0 = fbr0, a.damping = cp->cdef_damping;
        memcpy(a.y_strength, cp->cdef_y_strength, sizeof(a.y_strength));
        memcpy(a.uv_strength, cp->cdef_uv_strength, sizeof(a.uv_strength));
        if (c->geom.sample_bytes == 1)
            cdef_kernel<uint8_t><<<grid, CDEF_THREADS, 0, c->stream>>>(a);
        else
            cdef_kernel<uint16_t><<<grid, CDEF_THREADS, 0, c->stream>>>(a);
    } else
#endif
    {
        Cdef2Args b;
        b.in = c->surf[c->cur].planes(), b.out = c->surf[nxt].planes();
        b.mi = c->d_mi, b.mi_stride = c->mi_stride, b.d = dims_of(c->geom), b.fb_idx = c->d_fb_idx, b.fbr0 = fbr0, b.damping = cp->cdef_damping;
        memcpy(b.y_strength, cp->cdef_y_strength, sizeof(b.y_strength));
        memcpy(b.uv_strength, cp->cdef_uv_strength, sizeof(b.uv_strength));
        memset(&b.ccso, 0, sizeof(b.ccso));
        if (cc)
            ccso_fill(c, cc, blk_flags, b.ccso);
#ifdef SVTF_AB_KERNELS
        if (use_v2) {
            if (c->geom.sample_bytes == 1)
                cdef2_kernel<uint8_t><<<grid, C2_THREADS, 0, c->stream>>>(b);
            else
                cdef2_kernel<uint16_t><<<grid, C2_THREADS, 0, c->stream>>>(b);
        } else
#endif
        {
            // fp16 taps are exact for samples < 1024 and luma damping + (bit_depth - 8) <= 9 (svtf_cdef3.cuh); otherwise (12-bit pictures,
            // or SVTF_CDEF_INT=1 to force it) the packed-integer taps
            static const bool force_int = getenv("SVTF_CDEF_INT") != nullptr;
            const bool fp = !force_int && c->geom.bit_depth <= 10 && cp->cdef_damping + (c->geom.bit_depth - 8) <= 9;
            // SVTF_CDEF_TMA=1: cdef4_kernel (svtf_cdef4.cuh) — the same arithmetic in persistent CTAs fed by TMA.  Measured SLOWER than
            // cdef3 on the 4K 10-bit workload (95 vs 86 us alone), so it is not the default; kept selectable and parity-tested.
            static const bool use_tma = getenv("SVTF_CDEF_TMA") != nullptr;
            // fp16 taps: cdef5_kernel (svtf_cdef5.cuh).  SVTF_CDEF_V3=1 (builds with -DSVTF_AB_KERNELS only) selects its predecessor
            // cdef3_kernel<T, true> for A/B runs; cdef3_kernel<T, false> stays the integer path.
#ifdef SVTF_AB_KERNELS
            static const bool use_v5 = getenv("SVTF_CDEF_V3") == nullptr;
#else
            constexpr bool use_v5 = true;
#endif
            if (use_v5 && fp && !use_tma) {
                constexpr int smem5 = C5_SMEM_WORDS * 4;
                if (!c->cdef5_ready) {
                    CUDA_TRY(c, cudaFuncSetAttribute(cdef5_kernel<uint8_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem5));
                    CUDA_TRY(c, cudaFuncSetAttribute(cdef5_kernel<uint16_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem5));
                    CUDA_TRY(c, cudaFuncSetAttribute(cdef5_kernel<uint8_t>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
                    CUDA_TRY(c, cudaFuncSetAttribute(cdef5_kernel<uint16_t>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
                    c->cdef5_ready = true;
                }
                if (c->geom.sample_bytes == 1)
                    cdef5_kernel<uint8_t><<<grid, C2_THREADS, smem5, c->stream>>>(b);
                else
                    cdef5_kernel<uint16_t><<<grid, C2_THREADS, smem5, c->stream>>>(b);
            } else
            if (!use_tma) {
                constexpr int smem_i = C3_SMEM_WORDS_I * 4, smem_f = C3_SMEM_WORDS_F * 4;
                if (!c->cdef3_ready) {
                    CUDA_TRY(c, cudaFuncSetAttribute(cdef3_kernel<uint8_t, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_i));
                    CUDA_TRY(c, cudaFuncSetAttribute(cdef3_kernel<uint16_t, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_i));
#ifdef SVTF_AB_KERNELS
                    CUDA_TRY(c, cudaFuncSetAttribute(cdef3_kernel<uint8_t, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_f));
                    CUDA_TRY(c, cudaFuncSetAttribute(cdef3_kernel<uint16_t, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_f));
#endif
                    c->cdef3_ready = true;
                }
#ifdef SVTF_AB_KERNELS
                if (fp) {
                    if (c->geom.sample_bytes == 1)
                        cdef3_kernel<uint8_t, true><<<grid, C2_THREADS, smem_f, c->stream>>>(b);
                    else
                        cdef3_kernel<uint16_t, true><<<grid, C2_THREADS, smem_f, c->stream>>>(b);
                } else
#endif
                if (c->geom.sample_bytes == 1)
                    cdef3_kernel<uint8_t, false><<<grid, C2_THREADS, smem_i, c->stream>>>(b);
                else
                    cdef3_kernel<uint16_t, false><<<grid, C2_THREADS, smem_i, c->stream>>>(b);
            } else
            if (int r = cdef4_launch(c, b, fp, (int)grid.x, (int)grid.y))
                return r;
        }
    }
    CUDA_TRY(c, cudaGetLastError());
    c->launches++;
    c->pre_cdef = c->cur;
    c->cur      = nxt;
    return params_consumed(c);
}

extern "C" int svtf_cdef_apply_frame(svtf_ctx *c, const svtf_cdef_params *cp, const int8_t *fb_strength_idx) {
    if (!c)
        return set_global_error(SVTF_E_ARG, "svtf_cdef_apply_frame: NULL context");
    if (!cp || !fb_strength_idx)
        return c->fail(SVTF_E_ARG, "svtf_cdef_apply_frame: NULL argument");
    return cdef_apply_impl(c, cp, fb_strength_idx, nullptr, nullptr, "svtf_cdef_apply_frame", false);
}

static int ccso_apply_impl(svtf_ctx *c, const svtf_ccso_params *cc, const uint8_t *const blk_flags[3], bool staged) {
    if (!c->have_frame)
        return c->fail(SVTF_E_STATE, "svtf_ccso_apply_frame: no picture uploaded");
    if (int r = ccso_check(c, cc, "svtf_ccso_apply_frame"))
        return r;
    if (!(cc->ccso_enable[0] | cc->ccso_enable[1] | cc->ccso_enable[2]))
        return SVTF_OK;
    CUDA_TRY(c, cudaSetDevice(c->device));
    if (!staged) {
        if (int r = params_open(c))
            return r;
        ccso_put(c, cc, blk_flags);
        if (int r = params_flush(c))
            return r;
    }
    CcsoFuse f;
    ccso_fill(c, cc, blk_flags, f);
    CcsoArgs a;
    for (int p = 0; p < 3; p++) {
        a.enable[p] = f.enable[p], a.bo_only[p] = f.bo_only[p], a.thr[p] = f.thr[p], a.support[p] = f.support[p];
        a.edge_clf[p] = f.edge_clf[p], a.band_log2[p] = f.band_log2[p], a.blk_flags[p] = f.blk_flags[p];
    }
    a.lut = f.lut;
    a.ty0 = c->own_lo / 64, a.y_glob = c->glob_row0;
    a.d   = dims_of(c->geom);
    a.in  = c->surf[c->cur].planes();
    int nxt = c->cur;
    if (c->pre_cdef >= 0) { // classify on the pre-CDEF luma, modify the post-CDEF picture in place
        a.luma_src = c->surf[c->pre_cdef].planes();
        a.out      = a.in;
    } else { // no CDEF in this chain: luma is both classifier input and output -> go out of place
        nxt        = next_surface(c);
        a.luma_src = a.in;
        a.out      = c->surf[nxt].planes();
    }
    dim3 grid((c->geom.mi_cols * 4 + 63) / 64, (c->own_hi + 63) / 64 - a.ty0);
    if (c->geom.sample_bytes == 1)
        ccso_kernel<uint8_t><<<grid, CCSO_THREADS, 0, c->stream>>>(a);
    else
        ccso_kernel<uint16_t><<<grid, CCSO_THREADS, 0, c->stream>>>(a);
    CUDA_TRY(c, cudaGetLastError());
    c->launches++;
    c->cur = nxt;
    return params_consumed(c);
}

extern "C" int svtf_ccso_apply_frame(svtf_ctx *c, const svtf_ccso_params *cc, const uint8_t *const blk_flags[3]) {
    if (!c)
        return set_global_error(SVTF_E_ARG, "svtf_ccso_apply_frame: NULL context");
    if (!cc)
        return c->fail(SVTF_E_ARG, "svtf_ccso_apply_frame: NULL argument");
    return ccso_apply_impl(c, cc, blk_flags, false);
}

// Deblock -> CDEF -> CCSO on the uploaded picture.  All small parameters go to the device in one DMA ahead of the
// first kernel.  CDEF and CCSO of the same picture run as ONE kernel (the CCSO classifier input is the pre-CDEF luma
// tile the CDEF kernel already holds in shared memory); SVTF_CHAIN_UNFUSED=1 keeps them as separate launches for
// A/B measurements.
extern "C" int svtf_filter_chain(svtf_ctx *c, const svtf_lf_params *lf, const svtf_cdef_params *cdef, const int8_t *fb_strength_idx,
                                 const svtf_ccso_params *ccso, const uint8_t *const blk_flags[3]) {
    if (!c)
        return set_global_error(SVTF_E_ARG, "svtf_filter_chain: NULL context");
    if (cdef && !fb_strength_idx)
        return c->fail(SVTF_E_ARG, "svtf_filter_chain: CDEF needs fb_strength_idx");
    if (!c->have_frame)
        return c->fail(SVTF_E_STATE, "svtf_filter_chain: no picture uploaded");
    int r = SVTF_OK;
    static const bool unfused = getenv("SVTF_CHAIN_UNFUSED") != nullptr;
    const bool        ccso_on = ccso && (ccso->ccso_enable[0] | ccso->ccso_enable[1] | ccso->ccso_enable[2]);
    if (ccso && (r = ccso_check(c, ccso, "svtf_filter_chain")) != SVTF_OK)
        return r;
    if ((lf || cdef) && (r = need_frame_mi(c, "svtf_filter_chain")) != SVTF_OK)
        return r;
    if (cdef && (r = fb_idx_check(c, fb_strength_idx)) != SVTF_OK)
        return r;
    CUDA_TRY(c, cudaSetDevice(c->device));
    if (cdef || ccso_on) {
        if ((r = params_open(c)) != SVTF_OK)
            return r;
        if (cdef)
            params_put(c, 0, fb_local(c, fb_strength_idx), (size_t)((c->geom.mi_rows + 15) / 16) * ((c->geom.mi_cols + 15) / 16));
        if (ccso_on)
            ccso_put(c, ccso, blk_flags);
        if ((r = params_flush(c)) != SVTF_OK)
            return r;
    }
    if (lf && (r = svtf_deblock_frame(c, lf, 0, 3)) != SVTF_OK)
        return r;
    if (cdef && ccso_on && !unfused)
        return cdef_apply_impl(c, cdef, fb_strength_idx, ccso, blk_flags, "svtf_filter_chain", true);
    if (cdef && (r = cdef_apply_impl(c, cdef, fb_strength_idx, nullptr, nullptr, "svtf_filter_chain", true)) != SVTF_OK)
        return r;
    if (ccso_on && (r = ccso_apply_impl(c, ccso, blk_flags, true)) != SVTF_OK)
        return r;
    return r;
}

extern "C" int svtf_dev_planes(svtf_ctx *c, int which, svtf_dev_plane out[3]) {
    if (!c || !out || which < 0 || which > 2)
        return set_global_error(SVTF_E_ARG, "svtf_dev_planes: bad argument");
    if (!c->have_geom)
        return c->fail(SVTF_E_STATE, "svtf_dev_planes: call svtf_frame_begin first");
    const Surface &s = which == 0 ? c->surf[0] : (which == 1 ? c->surf[c->cur] : c->source);
    for (int p = 0; p < 3; p++) {
        out[p].ptr    = s.base[p];
        out[p].pitch  = s.pitch[p];
        out[p].width  = (c->geom.mi_cols * 4) >> (p ? 1 : 0);
        out[p].height = (c->geom.mi_rows * 4) >> (p ? 1 : 0);
    }
    return SVTF_OK;
}

#include "svtf_search_api.inl"
#include "svtf_ccso_derive_api.inl"
#include "svtf_lr_api.inl"
#include "svtf_lrs_api.inl" // super-resolution upscale + loop-restoration filter search entry points
#include "svtf_lf_search_api.inl"
#include "svtf_multi_api.inl"   // one picture across several GPUs: NCCL halo exchange and search-table gather
#include "svtf_noalloc_api.inl" // svtf_cdef_pick_table_noalloc: the pick on a static device scratch array (no allocation)
