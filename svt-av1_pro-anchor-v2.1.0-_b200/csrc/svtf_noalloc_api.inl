// svtf_noalloc_api.inl — textual part of svtf_api.cu (same translation unit as the pick kernels and the scratch array).
//
// svtf_cdef_pick_table_noalloc: svtf_cdef_pick_table for callers that may not allocate (code under CUDA graph capture,
// tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// Same kernels as svtf_cdef_pick_table (svtf_pick.cuh), but every device buffer the call needs — the
// uploaded table copy, the skip flags, the greedy-search state, the 64x64 totals and the per-block result — lives in ONE
// file-scope `static __device__` array instead of cudaMalloc'ed context members.  The array is part of the module image:
// every GPU that loads the library holds SVTF_NOALLOC_SCRATCH_BYTES (8.5 MB) until the process exits.
//
//   * In HOST code the name `g_noalloc_scratch` is only a host-side stand-in (a shadow symbol), not a device address;
//     the address a kernel can use comes from cudaGetSymbolAddress, asked once per device with that device current.
//   * All calls on one GPU share the array and it keeps the previous call's contents: the state block is cleared by a
//     cudaMemsetAsync and the totals by pick_init_kernel at the start of every call, the table / skip / result parts
//     are fully overwritten for the nfb entries in use.
//   * Calls that overlap on one GPU would race on the array.  Inside one process they cannot: a per-device host mutex is
//     held from the first copy to the final stream synchronisation.  (Other processes have their own copy of the array.)
//   * Limit: nfb <= SVTF_NOALLOC_MAX_FB (8192 filter blocks: an 8K picture has 120 x 68 = 8160); larger tables are
//     rejected with SVTF_E_ARG.
//   * The call itself issues no cudaMalloc / cudaFree / cudaHostAlloc and no allocating library routine: copies go
//     straight from / to the caller's buffers, the ~200-byte state is read back into a stack variable.

namespace {

constexpr size_t na_align(size_t v) { return (v + 255) & ~(size_t)255; }
constexpr size_t NA_OFF_MSE   = 0;                                                                  // uint64 [2][MAX_FB][64]
constexpr size_t NA_OFF_TOT   = na_align(NA_OFF_MSE + sizeof(uint64_t) * 128 * SVTF_NOALLOC_MAX_FB); // uint64 [64*64]
constexpr size_t NA_OFF_STATE = na_align(NA_OFF_TOT + sizeof(uint64_t) * 64 * 64);                   // PickState
constexpr size_t NA_OFF_SKIP  = na_align(NA_OFF_STATE + sizeof(PickState));                          // uint8 [MAX_FB]
constexpr size_t NA_OFF_IDX   = na_align(NA_OFF_SKIP + SVTF_NOALLOC_MAX_FB);                         // int8  [MAX_FB]
constexpr size_t NA_BYTES     = na_align(NA_OFF_IDX + SVTF_NOALLOC_MAX_FB);
static_assert(NA_BYTES <= SVTF_NOALLOC_SCRATCH_BYTES, "scratch layout exceeds the size stated in the header");

} // namespace

static __device__ __align__(256) unsigned char g_noalloc_scratch[SVTF_NOALLOC_SCRATCH_BYTES];

namespace {

constexpr int NA_MAX_DEVICES = 64;
struct NoAllocDevice {
    std::mutex     use;           // one call at a time per GPU: the array is shared
    unsigned char *base = nullptr; // device address of g_noalloc_scratch on this GPU
};
NoAllocDevice g_na_dev[NA_MAX_DEVICES];

// Device address of the scratch array on `device` (which must be current).  Called with g_na_dev[device].use held.
cudaError_t noalloc_base(int device, unsigned char **out) {
    NoAllocDevice &d = g_na_dev[device];
    if (!d.base) {
        void       *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess)
            return e;
        d.base = static_cast<unsigned char *>(p);
    }
    *out = d.base;
    return cudaSuccess;
}

} // namespace

extern "C" int svtf_cdef_pick_table_noalloc(svtf_ctx *c, const svtf_cdef_search_cfg *cfg, const uint64_t *mse, const uint8_t *skip, int32_t nfb,
                                            svtf_cdef_params *out_params, int8_t *fb_strength_idx) {
    if (!c)
        return set_global_error(SVTF_E_ARG, "svtf_cdef_pick_table_noalloc: NULL context");
    if (!cfg || !mse || !skip || !out_params || !fb_strength_idx || nfb <= 0)
        return c->fail(SVTF_E_ARG, "svtf_cdef_pick_table_noalloc: bad argument");
    if (nfb > SVTF_NOALLOC_MAX_FB) {
        char msg[160];
        snprintf(msg, sizeof(msg), "svtf_cdef_pick_table_noalloc: nfb = %d exceeds SVTF_NOALLOC_MAX_FB = %d (static scratch); use svtf_cdef_pick_table", nfb,
                 SVTF_NOALLOC_MAX_FB);
        return c->fail(SVTF_E_ARG, msg);
    }
    const int end_gi = cfg->first_pass_fs_num + cfg->default_second_pass_fs_num;
    if (end_gi <= 0 || end_gi > SVTF_TOTAL_STRENGTHS || cfg->first_pass_fs_num < 0 || cfg->default_second_pass_fs_num < 0)
        return c->fail(SVTF_E_ARG, "svtf_cdef_pick_table_noalloc: bad strength counts");
    if (c->device < 0 || c->device >= NA_MAX_DEVICES)
        return c->fail(SVTF_E_ARG, "svtf_cdef_pick_table_noalloc: device index out of range");
    CUDA_TRY(c, cudaSetDevice(c->device));
    std::lock_guard<std::mutex> lk(g_na_dev[c->device].use);
    unsigned char              *base = nullptr;
    CUDA_TRY(c, noalloc_base(c->device, &base));
    uint64_t  *d_mse  = reinterpret_cast<uint64_t *>(base + NA_OFF_MSE);
    uint8_t   *d_skip = base + NA_OFF_SKIP;
    int8_t    *d_idx  = reinterpret_cast<int8_t *>(base + NA_OFF_IDX);
    // the table keeps the caller's [2][nfb][64] layout (plane stride = nfb rows), so one copy per array
    CUDA_TRY(c, cudaMemcpyAsync(d_mse, mse, sizeof(uint64_t) * 128 * (size_t)nfb, cudaMemcpyHostToDevice, c->stream));
    CUDA_TRY(c, cudaMemcpyAsync(d_skip, skip, nfb, cudaMemcpyHostToDevice, c->stream));
    PickArgs a;
    a.mse = d_mse, a.skip = d_skip, a.nfb = nfb, a.end_gi = end_gi, a.bias = cfg->zero_fs_cost_bias;
    a.st  = reinterpret_cast<PickState *>(base + NA_OFF_STATE);
    a.tot = reinterpret_cast<uint64_t *>(base + NA_OFF_TOT);
    // pick_enqueue clears the state block (cudaMemsetAsync); pick_init_kernel zeroes the totals
    CUDA_TRY(c, pick_prepare()); // a function attribute, not an allocation
    CUDA_TRY(c, pick_enqueue(a, d_idx, (int64_t)cfg->lambda, c->sm_count, c->stream, &c->launches));
    PickState st;
    CUDA_TRY(c, cudaMemcpyAsync(&st, a.st, sizeof(st), cudaMemcpyDeviceToHost, c->stream));
    CUDA_TRY(c, cudaMemcpyAsync(fb_strength_idx, d_idx, nfb, cudaMemcpyDeviceToHost, c->stream));
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    memset(out_params, 0, sizeof(*out_params));
    out_params->cdef_bits = st.best_bits;
    for (int g = 0; g < (1 << st.best_bits); g++) { // index -> strength byte through the luma lists (Encoder/Codec/EbEncCdef.c:911-919)
        const int i0 = st.out0[g], i1 = st.out1[g];
        out_params->cdef_y_strength[g] =
            i0 < cfg->first_pass_fs_num ? cfg->default_first_pass_fs[i0] : cfg->default_second_pass_fs[i0 - cfg->first_pass_fs_num];
        out_params->cdef_uv_strength[g] =
            i1 < cfg->first_pass_fs_num ? cfg->default_first_pass_fs[i1] : cfg->default_second_pass_fs[i1 - cfg->first_pass_fs_num];
    }
    out_params->cdef_damping = 3 + (cfg->base_q_idx >> 6);
    return SVTF_OK;
}
