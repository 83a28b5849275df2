// svtf_multi_api.inl — textual part of svtf_api.cu: ONE picture across several GPUs (BASELINE configs[3]).
//
// A picture is split by superblock-row stripes (svtf_stripe_begin); one context per GPU, one process (or thread) per context.  Two steps
// of the path exchange data between GPUs, both device to device over NVLink through NCCL, neither touches host memory:
//   * svtf_stripe_exchange — the halo rows: every stripe needs the 8 luma / 4 chroma input rows and the 2 mode-info rows beyond each
//     interior boundary (the boundary deblock edge reads <= 7 and writes <= 6 rows per side, Common/Codec/EbDeblockingCommon.c:787-823; CDEF
//     taps reach 2 rows, Common/Codec/EbCdef.c:99-120; CCSO 1).  A picture that is ALREADY RESIDENT as stripes (uploaded with
//     svtf_upload_frame_own, or left on the devices by an earlier stage) gets them from the neighbours' memory: grouped ncclSend / ncclRecv
//     of contiguous row ranges, stream-ordered on the contexts' streams.
//   * svtf_cdef_search_gather_pick — finish_cdef_search (Encoder/Codec/EbEncCdef.c:728) needs the mse / skip rows of ALL filter blocks;
//     each rank searched its own filter-block rows (svtf_cdef_search_frame on a striped context, outputs NULL).  Every rank broadcasts
//     its rows from its device table into every rank's full-picture device table (one NCCL group: an all-gather with ragged counts), then
//     each rank runs the same deterministic pick on the device.  This replaces the mutex-guarded join of the reference's CDEF segments
//     (Encoder/Codec/EbCdefProcess.c:406, :730).
// NCCL is bound at run time (dlopen of libnccl.so.2: the copy the process already holds — PyTorch's — or the system's); the library has no
// link-time dependency on it, and every entry point here fails with SVTF_E_STATE when NCCL cannot be loaded.
#include <dlfcn.h>

namespace {

typedef struct ncclComm *svtf_nccl_comm_t;
struct svtf_nccl_uid {
    char internal[128];
};
struct NcclApi {
    int (*GetUniqueId)(svtf_nccl_uid *)                                                            = nullptr;
    int (*CommInitRank)(svtf_nccl_comm_t *, int, svtf_nccl_uid, int)                               = nullptr;
    int (*CommDestroy)(svtf_nccl_comm_t)                                                           = nullptr;
    int (*GroupStart)()                                                                            = nullptr;
    int (*GroupEnd)()                                                                              = nullptr;
    int (*Send)(const void *, size_t, int, int, svtf_nccl_comm_t, cudaStream_t)                    = nullptr;
    int (*Recv)(void *, size_t, int, int, svtf_nccl_comm_t, cudaStream_t)                          = nullptr;
    int (*Broadcast)(const void *, void *, size_t, int, int, svtf_nccl_comm_t, cudaStream_t)       = nullptr;
    const char *(*GetErrorString)(int)                                                             = nullptr;
    bool ok                                                                                        = false;
};
constexpr int NCCL_UINT8 = 1; // ncclUint8

const NcclApi &nccl_api() {
    static NcclApi api = [] {
        NcclApi a;
        void   *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h)
            h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!h)
            return a;
#define SVTF_NCCL_SYM(field, name) *reinterpret_cast<void **>(&a.field) = dlsym(h, name)
        SVTF_NCCL_SYM(GetUniqueId, "ncclGetUniqueId");
        SVTF_NCCL_SYM(CommInitRank, "ncclCommInitRank");
        SVTF_NCCL_SYM(CommDestroy, "ncclCommDestroy");
        SVTF_NCCL_SYM(GroupStart, "ncclGroupStart");
        SVTF_NCCL_SYM(GroupEnd, "ncclGroupEnd");
        SVTF_NCCL_SYM(Send, "ncclSend");
        SVTF_NCCL_SYM(Recv, "ncclRecv");
        SVTF_NCCL_SYM(Broadcast, "ncclBroadcast");
        SVTF_NCCL_SYM(GetErrorString, "ncclGetErrorString");
#undef SVTF_NCCL_SYM
        a.ok = a.GetUniqueId && a.CommInitRank && a.CommDestroy && a.GroupStart && a.GroupEnd && a.Send && a.Recv && a.Broadcast;
        return a;
    }();
    return api;
}

} // namespace

struct svtf_comm {
    svtf_nccl_comm_t comm = nullptr;
    int              device = 0, nranks = 1, rank = 0;
};

#define NCCL_TRY(ctx, call)                                                                                                   \
    do {                                                                                                                      \
        const int r_ = (call);                                                                                                \
        if (r_ != 0)                                                                                                          \
            return (ctx)->fail(SVTF_E_CUDA, std::string(#call ": ") + (nccl_api().GetErrorString ? nccl_api().GetErrorString(r_) : "NCCL error")); \
    } while (0)

extern "C" int svtf_comm_unique_id(void *id128) {
    if (!id128)
        return set_global_error(SVTF_E_ARG, "svtf_comm_unique_id: NULL argument");
    if (!nccl_api().ok)
        return set_global_error(SVTF_E_STATE, "svtf_comm_unique_id: NCCL (libnccl.so.2) could not be loaded");
    svtf_nccl_uid id;
    if (nccl_api().GetUniqueId(&id) != 0)
        return set_global_error(SVTF_E_CUDA, "svtf_comm_unique_id: ncclGetUniqueId failed");
    memcpy(id128, &id, sizeof(id));
    return SVTF_OK;
}

extern "C" int svtf_comm_create(int device, int nranks, int rank, const void *id128, svtf_comm **out) {
    if (!out || !id128 || nranks < 1 || rank < 0 || rank >= nranks)
        return set_global_error(SVTF_E_ARG, "svtf_comm_create: bad argument");
    if (!nccl_api().ok)
        return set_global_error(SVTF_E_STATE, "svtf_comm_create: NCCL (libnccl.so.2) could not be loaded");
    if (cudaSetDevice(device) != cudaSuccess)
        return set_global_error(SVTF_E_CUDA, "svtf_comm_create: cudaSetDevice failed");
    svtf_comm *c = new (std::nothrow) svtf_comm();
    if (!c)
        return set_global_error(SVTF_E_NOMEM, "svtf_comm_create: out of host memory");
    svtf_nccl_uid id;
    memcpy(&id, id128, sizeof(id));
    c->device = device, c->nranks = nranks, c->rank = rank;
    const int r = nccl_api().CommInitRank(&c->comm, nranks, id, rank);
    if (r != 0) {
        std::string m = std::string("svtf_comm_create: ncclCommInitRank: ") + (nccl_api().GetErrorString ? nccl_api().GetErrorString(r) : "NCCL error");
        delete c;
        return set_global_error(SVTF_E_CUDA, m.c_str());
    }
    *out = c;
    return SVTF_OK;
}

extern "C" void svtf_comm_destroy(svtf_comm *c) {
    if (!c)
        return;
    cudaSetDevice(c->device);
    if (c->comm && nccl_api().ok)
        nccl_api().CommDestroy(c->comm);
    delete c;
}

// Rows [lo, hi) of the stripe's own rows only (no halo): what a rank holds when the picture reaches the GPUs already split.
extern "C" int svtf_upload_frame_own(svtf_ctx *c, const void *y, int32_t sy, const void *cb, int32_t scb, const void *cr, int32_t scr) {
    if (!c)
        return set_global_error(SVTF_E_ARG, "svtf_upload_frame_own: NULL context");
    if (!c->have_geom)
        return c->fail(SVTF_E_STATE, "svtf_upload_frame_own: call svtf_stripe_begin first");
    const void   *src[3] = {y, cb, cr};
    const int32_t st[3]  = {sy, scb, scr};
    c->cur = 0, c->pre_cdef = -1;
    int r = copy_planes(c, c->surf[0], src, st, true, c->own_lo, c->own_hi);
    if (r == SVTF_OK)
        c->have_frame = true;
    return r;
}

// The packed mi rows of the stripe's own rows only.
extern "C" int svtf_set_mi_own(svtf_ctx *c, const svtf_mi_t *mi, int32_t mi_stride) {
    if (!c)
        return set_global_error(SVTF_E_ARG, "svtf_set_mi_own: NULL context");
    if (!c->have_geom)
        return c->fail(SVTF_E_STATE, "svtf_set_mi_own: call svtf_stripe_begin first");
    if (!mi || mi_stride < c->geom.mi_cols)
        return c->fail(SVTF_E_ARG, "svtf_set_mi_own: NULL grid or stride smaller than mi_cols");
    const int r0 = c->own_lo / 4, r1 = (c->own_hi + 3) / 4;
    mi += (size_t)(c->glob_row0 / 4 + r0) * mi_stride; // whole-picture grid -> this context's own rows
    CUDA_TRY(c, cudaMemcpy2DAsync(c->d_mi + (size_t)r0 * c->mi_stride, sizeof(svtf_mi_t) * c->mi_stride, mi, sizeof(svtf_mi_t) * mi_stride,
                                  sizeof(svtf_mi_t) * c->geom.mi_cols, r1 - r0, cudaMemcpyHostToDevice, c->stream));
    c->have_mi = true, c->edge_valid = false;
    return SVTF_OK;
}

extern "C" int svtf_stripe_exchange(svtf_ctx *c, svtf_comm *comm, int32_t rank_up, int32_t rank_down) {
    if (!c)
        return set_global_error(SVTF_E_ARG, "svtf_stripe_exchange: NULL context");
    if (!comm || rank_up >= comm->nranks || rank_down >= comm->nranks)
        return c->fail(SVTF_E_ARG, "svtf_stripe_exchange: bad communicator or neighbour rank");
    if (!c->striped || !c->have_frame || !c->have_mi)
        return c->fail(SVTF_E_STATE, "svtf_stripe_exchange: needs a striped context with its own rows and mode info on the device");
    const int rows = c->geom.mi_rows * 4;
    const bool up = rank_up >= 0 && c->own_lo >= 8, down = rank_down >= 0 && c->own_hi + 8 <= rows;
    if ((rank_up >= 0) != (c->own_lo > 0) || (rank_down >= 0) != (c->own_hi < rows))
        return c->fail(SVTF_E_ARG, "svtf_stripe_exchange: a neighbour must be named exactly where the stripe has an interior boundary");
    if (!up && !down)
        return SVTF_OK;
    CUDA_TRY(c, cudaSetDevice(c->device));
    const NcclApi &n  = nccl_api();
    const size_t   sb = c->geom.sample_bytes;
    Surface       &s  = c->surf[c->cur];
    NCCL_TRY(c, n.GroupStart());
    for (int p = 0; p < 3; p++) {
        const int    ss = p ? 1 : 0, h = 8 >> ss; // 8 luma / 4 chroma rows
        const size_t rb = (size_t)s.pitch[p] * sb; // bytes per row
        uint8_t     *b  = static_cast<uint8_t *>(s.base[p]);
        if (up) { // my first rows -> the upper neighbour's lower halo; its last rows -> my upper halo
            NCCL_TRY(c, n.Send(b + (size_t)(c->own_lo >> ss) * rb, h * rb, NCCL_UINT8, rank_up, comm->comm, c->stream));
            NCCL_TRY(c, n.Recv(b + (size_t)((c->own_lo >> ss) - h) * rb, h * rb, NCCL_UINT8, rank_up, comm->comm, c->stream));
        }
        if (down) {
            NCCL_TRY(c, n.Send(b + (size_t)((c->own_hi >> ss) - h) * rb, h * rb, NCCL_UINT8, rank_down, comm->comm, c->stream));
            NCCL_TRY(c, n.Recv(b + (size_t)(c->own_hi >> ss) * rb, h * rb, NCCL_UINT8, rank_down, comm->comm, c->stream));
        }
    }
    { // two mode-info rows (8 luma rows) on either side
        const size_t rb = sizeof(svtf_mi_t) * (size_t)c->mi_stride;
        uint8_t     *b  = reinterpret_cast<uint8_t *>(c->d_mi);
        if (up) {
            NCCL_TRY(c, n.Send(b + (size_t)(c->own_lo / 4) * rb, 2 * rb, NCCL_UINT8, rank_up, comm->comm, c->stream));
            NCCL_TRY(c, n.Recv(b + (size_t)(c->own_lo / 4 - 2) * rb, 2 * rb, NCCL_UINT8, rank_up, comm->comm, c->stream));
        }
        if (down) {
            NCCL_TRY(c, n.Send(b + (size_t)(c->own_hi / 4 - 2) * rb, 2 * rb, NCCL_UINT8, rank_down, comm->comm, c->stream));
            NCCL_TRY(c, n.Recv(b + (size_t)(c->own_hi / 4) * rb, 2 * rb, NCCL_UINT8, rank_down, comm->comm, c->stream));
        }
    }
    NCCL_TRY(c, n.GroupEnd());
    c->edge_valid = false; // the halo mode-info rows changed
    return SVTF_OK;
}

extern "C" int svtf_cdef_search_gather_pick(svtf_ctx *c, svtf_comm *comm, const svtf_cdef_search_cfg *cfg, const int32_t *first_sb_row,
                                            const int32_t *num_sb_rows, svtf_cdef_params *out_params, int8_t *fb_strength_idx) {
    if (!c)
        return set_global_error(SVTF_E_ARG, "svtf_cdef_search_gather_pick: NULL context");
    if (!comm || !cfg || !first_sb_row || !num_sb_rows || !out_params || !fb_strength_idx)
        return c->fail(SVTF_E_ARG, "svtf_cdef_search_gather_pick: NULL argument");
    if (!c->striped || !c->have_search)
        return c->fail(SVTF_E_STATE, "svtf_cdef_search_gather_pick: run svtf_cdef_search_frame on the striped context first");
    CUDA_TRY(c, cudaSetDevice(c->device));
    const int nhfb = (c->geom.mi_cols + 15) / 16, nvfb_full = (c->full.mi_rows + 15) / 16, nfb_full = nvfb_full * nhfb;
    const int nfb_loc = ((c->geom.mi_rows + 15) / 16) * nhfb;
    if (first_sb_row[comm->rank] * 64 != c->glob_row0 + c->own_lo)
        return c->fail(SVTF_E_ARG, "svtf_cdef_search_gather_pick: first_sb_row[rank] does not match this context's stripe");
    if (nfb_full > c->tab_cap) {
        cudaFree(c->d_tab_mse), cudaFree(c->d_tab_skip);
        c->d_tab_mse = nullptr, c->d_tab_skip = nullptr, c->tab_cap = 0;
        CUDA_TRY(c, cudaMalloc(&c->d_tab_mse, sizeof(uint64_t) * 128 * (size_t)nfb_full));
        CUDA_TRY(c, cudaMalloc(&c->d_tab_skip, nfb_full));
        c->tab_cap = nfb_full;
    }
    const NcclApi &n  = nccl_api();
    const size_t   l0 = (size_t)(c->own_lo / 64) * nhfb; // this rank's first own filter block in its local table
    NCCL_TRY(c, n.GroupStart());
    for (int r = 0; r < comm->nranks; r++) {
        if (first_sb_row[r] < 0 || num_sb_rows[r] < 0 || first_sb_row[r] + num_sb_rows[r] > nvfb_full)
            return c->fail(SVTF_E_ARG, "svtf_cdef_search_gather_pick: stripe outside the picture");
        const size_t g0 = (size_t)first_sb_row[r] * nhfb, cnt = (size_t)num_sb_rows[r] * nhfb; // rank r's filter blocks in the picture
        if (!cnt)
            continue;
        for (int pl = 0; pl < 2; pl++)
            NCCL_TRY(c, n.Broadcast(c->d_mse + ((size_t)pl * nfb_loc + l0) * 64, c->d_tab_mse + ((size_t)pl * nfb_full + g0) * 64, sizeof(uint64_t) * 64 * cnt,
                                    NCCL_UINT8, r, comm->comm, c->stream));
        NCCL_TRY(c, n.Broadcast(c->d_skip + l0, c->d_tab_skip + g0, cnt, NCCL_UINT8, r, comm->comm, c->stream));
    }
    NCCL_TRY(c, n.GroupEnd());
    return pick_device(c, cfg, c->d_tab_mse, c->d_tab_skip, nfb_full, out_params, fb_strength_idx, "svtf_cdef_search_gather_pick");
}
