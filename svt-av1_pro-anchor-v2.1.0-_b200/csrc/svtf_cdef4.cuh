// svtf_cdef4.cuh — CDEF apply (+ optional fused CCSO epilogue): persistent CTAs fed by TMA.
//
// Same semantics and the same arithmetic as cdef3_kernel (svtf_cdef3.cuh: packed fp16 taps for 8/10-bit pictures, packed-integer
// taps for 12-bit, CCSO in the epilogue): svt_av1_cdef_frame (Encoder/Codec/EbEncCdef.c:284-612), svt_cdef_filter_fb /
// svt_cdef_filter_block_c (Common/Codec/EbCdef.c:253-430), ccso_frame (Common/Codec/EbCcso.c:626-678).  What changed is how tiles
// reach shared memory and how CTAs live (profiles/r01e_ncu_summary.md: staging + conversion were 16.5 % of cdef3's instructions and
// 35 % of its stall samples, the per-CTA prologue — CCSO LUT, tap-offset table — another 10 %):
//
//  * one PERSISTENT CTA per SM slot (grid = SMs x CTAs/SM) walks the filter blocks tile += gridDim.x; the CCSO LUT and the
//    (direction, tap) -> word-offset table are built ONCE per CTA;
//  * the three tiles of a filter block (68 x 80 luma, 36 x 48 per chroma plane: block + 2-row / 8-column halo; what
//    svt_aom_copy_sb8_16 / svt_aom_fill_rect stage in the reference, Common/Codec/EbCdef.c:308-337, Encoder/Codec/EbEncCdef.c:459-568)
//    arrive by cp.async.bulk.tensor.2d into a raw staging buffer, completion on an mbarrier; samples outside the coded plane are
//    zero-filled by the TMA unit and replaced by the sentinel in the conversion pass (the masks only differ on picture-edge tiles);
//  * the TMA of tile n+1 is issued as soon as tile n's conversion pass (raw -> fp16 even / odd copies + integer luma) has consumed
//    the staging buffer, so the load runs under the direction search, the twelve taps and the epilogue of tile n: no thread ever
//    waits for a global tile load except at the first tile of a CTA.
#pragma once
#include <cuda.h>
#include "svtf_cdef3.cuh"

namespace svtf {

constexpr int C4_YBOX_H = C2_YROWS, C4_CBOX_H = C2_CROWS; // box heights: 2 + 64 + 2 luma rows, 2 + 32 + 2 chroma rows
// Box geometry.  A tile needs the block plus 8 samples on either side (8 + 64 + 8 luma, 8 + 32 + 8 chroma).  The TMA unit faults
// ("illegal instruction") when the first byte of a box row is not 16-byte aligned in the tensor's inner dimension — measured with
// tools/tma_probe/tma_probe.cu on this B200: a box at x = -8 bytes or +56 bytes faults, -16 is fine, whatever the element type —
// so 8-bit planes start 16 samples left of the block (X_OFF) and the conversion pass skips the 8 extra columns.
template <typename T>
struct C4Raw {
    static constexpr int X_OFF = sizeof(T) == 1 ? 16 : 8;                       // samples between the box origin and the block
    static constexpr int Y_W = sizeof(T) == 1 ? 96 : 80, C_W = sizeof(T) == 1 ? 64 : 48; // box widths (row bytes: multiples of 16)
    static constexpr int Y_BYTES = (Y_W * C4_YBOX_H * (int)sizeof(T) + 127) & ~127;
    static constexpr int C_BYTES = (C_W * C4_CBOX_H * (int)sizeof(T) + 127) & ~127;
    static constexpr int BYTES   = Y_BYTES + 2 * C_BYTES;
    static constexpr int TX      = (Y_W * C4_YBOX_H + 2 * C_W * C4_CBOX_H) * (int)sizeof(T); // bytes one tile's three copies deliver
};
template <typename T, bool FP>
constexpr int c4_smem_bytes() {
    return C4Raw<T>::BYTES + 4 * (FP ? C3_SMEM_WORDS_F : C3_SMEM_WORDS_I);
}

struct Cdef4Args {
    Cdef2Args   c;
    CUtensorMap tm[3]; // input planes (Y, Cb, Cr): dims = coded plane size, box = the tile sizes above
    int         nhfb, ntiles;
    unsigned   *sched; // [0] work counter (tile = gridDim.x + draw), [1] CTAs that have left; both zero on entry, rearmed by the last CTA
};

// ---- mbarrier / TMA (PTX; sm_90+) -----------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t c4_smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void     c4_mbar_init(uint64_t *bar, const uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(c4_smem_addr(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void c4_fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
// generic-proxy accesses to shared memory that precede this fence are ordered before async-proxy (TMA) accesses that follow it
__device__ __forceinline__ void c4_fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void c4_mbar_expect_tx(uint64_t *bar, const uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(c4_smem_addr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void c4_mbar_wait(uint64_t *bar, const uint32_t parity) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "C4_WAIT:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra C4_DONE;\n\t"
        "bra C4_WAIT;\n\t"
        "C4_DONE:\n\t"
        "}" ::"r"(c4_smem_addr(bar)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void c4_tma_load_2d(void *dst, const CUtensorMap *tm, const int x, const int y, uint64_t *bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(c4_smem_addr(dst)),
                 "l"(reinterpret_cast<uint64_t>(tm)), "r"(x), "r"(y), "r"(c4_smem_addr(bar))
                 : "memory");
}

// 8 consecutive samples of a raw (dense, TMA-written) tile row as four packed u16 pairs; 4 samples as two
template <typename T>
__device__ __forceinline__ uint4 c4_lds8(const T *p);
template <>
__device__ __forceinline__ uint4 c4_lds8<uint16_t>(const uint16_t *p) {
    return *reinterpret_cast<const uint4 *>(p);
}
template <>
__device__ __forceinline__ uint4 c4_lds8<uint8_t>(const uint8_t *p) {
    const uint2 v = *reinterpret_cast<const uint2 *>(p);
    return make_uint4(__byte_perm(v.x, 0, 0x4140), __byte_perm(v.x, 0, 0x4342), __byte_perm(v.y, 0, 0x4140), __byte_perm(v.y, 0, 0x4342));
}

// One 8-sample chunk of a raw tile as the C3Chunk c3_store consumes: validity masks as c3_load computes them (the TMA unit wrote zeros
// where the plane ends; the masks turn those into the sentinel).  bw: box width in samples; (r, c): tile row / first column of the chunk.
template <typename T>
__device__ __forceinline__ C3Chunk c4_chunk(const T *raw, const int bw, const int r, const int c, const int gx, const int gy, const int cw,
                                            const int ch) {
    C3Chunk       k;
    const T      *src = raw + r * bw + c + (C4Raw<T>::X_OFF - 8); // tile column c = box column c + (X_OFF - 8)
    const uint4   v = c4_lds8<T>(src);
    const bool    row_in = gy >= 0 && gy < ch;
    k.w0 = v.x, k.w1 = v.y, k.w2 = v.z, k.w3 = v.w;
    k.wn = c + (C4Raw<T>::X_OFF - 8) + 8 < bw ? (uint32_t)src[8] : 0u;
    k.m  = 0;
    if (row_in) {
        if (gx >= 0 && gx < cw)
            k.m = gx + 8 <= cw ? 3 : 1; // chroma planes whose width is 4 mod 8: last half chunk
        if (gx + 8 >= 0 && gx + 8 < cw)
            k.m |= 4;
    }
    return k;
}

template <typename T, bool FP>
__global__ void __launch_bounds__(C2_THREADS, 3) cdef4_kernel(const __grid_constant__ Cdef4Args g) {
    extern __shared__ __align__(128) uint8_t c4_dyn[];
    const Cdef2Args &a = g.c;
    T *const         rawY  = reinterpret_cast<T *>(c4_dyn);
    T *const         rawC0 = reinterpret_cast<T *>(c4_dyn + C4Raw<T>::Y_BYTES);
    T *const         rawC1 = reinterpret_cast<T *>(c4_dyn + C4Raw<T>::Y_BYTES + C4Raw<T>::C_BYTES);
    uint32_t *const  c3_sm = reinterpret_cast<uint32_t *>(c4_dyn + C4Raw<T>::BYTES);
    uint32_t *const  yA  = c3_sm;                                     // yB = yA + C3_YSZ
    uint32_t *const  cA0 = c3_sm + 2 * C3_YSZ;                        // plane p: cA0 + p * 2 * C3_CSZ, its B copy + C3_CSZ
    uint32_t *const  yI  = FP ? c3_sm + 2 * C3_YSZ + 4 * C3_CSZ : yA; // integer luma (direction search, CCSO classifier)
    __shared__ __align__(8) uint64_t s_full;                          // mbarrier: the raw staging buffer holds the next tile
    __shared__ int     s_next[2];                                     // tile after the current one (double-buffered broadcast)
    __shared__ int     s_cost[64][9];                                 // [block][dir], padded
    __shared__ int2    s_off[2][8][2];                                // [luma/chroma][dir][k]: word offsets of the +tap / -tap
    __shared__ uint8_t s_dir[64];
    __shared__ int     s_var[64];
    __shared__ uint8_t s_active[64];
    __shared__ __align__(16) uint8_t s_lut[3][SVTF_CCSO_LUT_SIZE];    // CCSO offsets + 128

    const int tid = threadIdx.x;
    const int cw = a.d.mi_cols * 4, ch = a.d.mi_rows * 4;
    const int cs = a.d.bit_depth - 8;
    const int W = a.d.width, H = a.d.height;
    T *const  outy = static_cast<T *>(a.out.p[0]);

    auto issue = [&](const int tile) { // one elected thread: the three copies of `tile` into the raw buffer
        const int fbc = tile % g.nhfb, fbr = tile / g.nhfb + a.fbr0;
        c4_mbar_expect_tx(&s_full, (uint32_t)C4Raw<T>::TX);
        c4_tma_load_2d(rawY, &g.tm[0], fbc * 64 - C4Raw<T>::X_OFF, fbr * 64 - 2, &s_full);
        c4_tma_load_2d(rawC0, &g.tm[1], fbc * 32 - C4Raw<T>::X_OFF, fbr * 32 - 2, &s_full);
        c4_tma_load_2d(rawC1, &g.tm[2], fbc * 32 - C4Raw<T>::X_OFF, fbr * 32 - 2, &s_full);
    };

    // ---- once per CTA ----
    int tile = blockIdx.x; // later tiles are handed out dynamically: a CTA that drew cheap (skipped) filter blocks comes back sooner
    if (tid == 0) {
        c4_mbar_init(&s_full, 1);
        c4_fence_mbar_init();
        if (tile < g.ntiles)
            issue(tile);
        s_next[0] = tile < g.ntiles ? (int)(gridDim.x + atomicAdd(&g.sched[0], 1u)) : g.ntiles;
    }
    if (tid >= 64 && tid < 96) {
        // Cdef_Directions (Common/Codec/EbCdef.c:99-122) as nibble tables (value + 2) indexed by direction:
        // k=0: dy {-1,0,0,0,1,1,1,1} dx {1,1,1,1,1,0,0,0};  k=1: dy {-2,-1,0,1,2,2,2,2} dx {2,2,2,2,2,1,0,-1}
        const int t = tid - 64, pt = t >> 4, dd = (t >> 1) & 7, k = t & 1;
        const int dy = (int)(((k ? 0x44443210u : 0x33332221u) >> (4 * dd)) & 15) - 2;
        const int dx = (int)(((k ? 0x12344444u : 0x22233333u) >> (4 * dd)) & 15) - 2;
        const int pw = pt ? C2_CPW : C2_YPW, bd = pt ? C3_CSZ : C3_YSZ, odd = dx & 1; // centre columns are even
        s_off[pt][dd][k] = make_int2((odd ? bd : 0) + dy * pw + ((dx - odd) >> 1), (odd ? bd : 0) - dy * pw + ((-dx - odd) >> 1));
    }
    if (a.ccso.on) { // CCSO LUTs (3 x 2048 bytes = 384 x 16 bytes) of the planes that can run, biased by +128
        for (int i = tid; i < 3 * SVTF_CCSO_LUT_SIZE / 16; i += C2_THREADS)
            if (a.ccso.enable[i >> 7]) {
                const uint4 l = __ldg(reinterpret_cast<const uint4 *>(a.ccso.lut) + i);
                reinterpret_cast<uint4 *>(&s_lut[0][0])[i] = make_uint4(l.x ^ 0x80808080u, l.y ^ 0x80808080u, l.z ^ 0x80808080u, l.w ^ 0x80808080u);
            }
    }
    __syncthreads(); // the mbarrier is initialised before any thread waits on it
    uint32_t parity = 0;

    for (int it = 0; tile < g.ntiles; it++) {
        const int nxt = s_next[it & 1]; // written before the previous iteration's last barrier (or the set-up barrier)
        const int fbc = tile % g.nhfb, fbr = tile / g.nhfb + a.fbr0;
        const int x0 = fbc * 64, y0 = fbr * 64;

        // ---- side info of this tile: in flight while the tile is converted ----
        const int idx = __ldg(a.fb_idx + fbr * g.nhfb + fbc);
        int       flg = 0; // bit p: CCSO block flag of plane p (one flag per 256x256 luma samples, row = ceil(plane width / block))
        if (a.ccso.on) {
#pragma unroll
            for (int p = 0; p < 3; p++) {
                bool on = a.ccso.enable[p];
                if (on && a.ccso.blk_flags[p]) {
                    const int pwid = W >> (p ? 1 : 0), bl = p ? 7 : 8;
                    const int nbx  = (pwid + (1 << bl) - 1) >> bl;
                    on             = (x0 >> 8) < nbx && __ldg(a.ccso.blk_flags[p] + ((y0 + a.ccso.y_glob) >> 8) * nbx + (x0 >> 8)) != 0;
                }
                flg |= (int)on << p;
            }
        }
        svtf_mi_t m4[4] = {~0u, ~0u, ~0u, ~0u};
        bool      mi_in = false;
        if (tid < 64) { // svt_sb_compute_cdef_list — Encoder/Codec/EbEncCdef.c:238-281
            const int r = (tid >> 3) * 2, c = (tid & 7) * 2;
            mi_in = fbr * 16 + r < a.d.mi_rows && fbc * 16 + c < a.d.mi_cols;
            if (mi_in) {
                const svtf_mi_t *m = a.mi + (fbr * 16 + r) * a.mi_stride + fbc * 16 + c;
                m4[0] = __ldg(m), m4[1] = __ldg(m + 1), m4[2] = __ldg(m + a.mi_stride), m4[3] = __ldg(m + a.mi_stride + 1);
            }
        }

        // ---- the tile has landed: raw -> even / odd (fp16) copies + integer luma ----
        c4_mbar_wait(&s_full, parity);
        parity ^= 1;
#pragma unroll
        for (int k = 0; k < 3; k++) { // luma: 68 rows x 10 chunks
            const int i = tid + k * C2_THREADS;
            if (i < C2_YROWS * 10) {
                const int     r = i / 10, c = (i - r * 10) * 8;
                const C3Chunk q = c4_chunk<T>(rawY, C4Raw<T>::Y_W, r, c, x0 - 8 + c, y0 - 2 + r, cw, ch);
                c3_store<FP>(yA, yA + C3_YSZ, FP ? yI : nullptr, r * C2_YPW + (c >> 1), q);
            }
        }
        if (tid < C2_CROWS * 6) { // chroma: 36 rows x 6 chunks per plane
            const int rc = tid / 6, cc = (tid - rc * 6) * 8;
#pragma unroll
            for (int p = 0; p < 2; p++) {
                const C3Chunk q = c4_chunk<T>(p ? rawC1 : rawC0, C4Raw<T>::C_W, rc, cc, x0 / 2 - 8 + cc, y0 / 2 - 2 + rc, cw / 2, ch / 2);
                c3_store<FP>(cA0 + p * 2 * C3_CSZ, cA0 + p * 2 * C3_CSZ + C3_CSZ, nullptr, rc * C2_CPW + (cc >> 1), q);
            }
        }
        int ylev = 0, ysec = 0, uvlev = 0, uvsec = 0;
        if (idx >= 0) {
            const int ys = a.y_strength[idx], us = a.uv_strength[idx];
            ylev = ys >> 2, ysec = ys & 3, ysec += ysec == 3;
            uvlev = us >> 2, uvsec = us & 3, uvsec += uvsec == 3;
        }
        int act = 0;
        if (tid < 64) {
            act           = mi_in && !(mi_skip(m4[0]) & mi_skip(m4[1]) & mi_skip(m4[2]) & mi_skip(m4[3]));
            s_active[tid] = (uint8_t)act;
        }
        const int  any_active = __syncthreads_or(act); // also: every thread is done with the raw buffer, the copies are visible
        const bool process    = (ylev | ysec | uvlev | uvsec) && any_active;
        const int  q_on       = flg; // bit p: CCSO (ccso_frame, Common/Codec/EbCcso.c:626) runs on plane p of this tile
        if (tid == 0) { // the next tile loads under this tile's arithmetic
            if (nxt < g.ntiles) {
                c4_fence_proxy_async();
                issue(nxt);
                s_next[(it + 1) & 1] = (int)(gridDim.x + atomicAdd(&g.sched[0], 1u));
            } else
                s_next[(it + 1) & 1] = g.ntiles;
        }

        if (!process && !q_on) { // untouched filter block: carry the samples over to the output surface
            for (int i = tid; i < 64 * 8; i += C2_THREADS) {
                const int r = i >> 3, c = (i & 7) * 8;
                if (y0 + r < ch && x0 + c < cw) {
                    const uint32_t *s = yI + (2 + r) * C2_YPW + 4 + (c >> 1);
                    st8<T>(outy + (size_t)(y0 + r) * a.out.pitch[0] + x0 + c, make_uint4(s[0], s[1], s[2], s[3]));
                }
            }
            for (int i = tid; i < 2 * 32 * 8; i += C2_THREADS) {
                const int p = i >> 8, r = (i >> 3) & 31, c = (i & 7) * 4;
                if (y0 / 2 + r < ch / 2 && x0 / 2 + c < cw / 2) {
                    const uint32_t *s  = cA0 + p * 2 * C3_CSZ + (2 + r) * C2_CPW + 4 + (c >> 1);
                    const uint32_t  w0 = FP ? pk_h2i(s[0]) : s[0], w1 = FP ? pk_h2i(s[1]) : s[1];
                    st4<T>(static_cast<T *>(a.out.p[1 + p]) + (size_t)(y0 / 2 + r) * a.out.pitch[1 + p] + x0 / 2 + c, make_uint2(w0, w1));
                }
            }
            __syncthreads(); // the copies are rewritten by the next tile's conversion pass
            tile = nxt;
            continue;
        }

        // ---- direction search: thread = (block, half of the directions) ----
        if (tid < 128 && process) {
            const int b = tid & 63, half = tid >> 6;
            if (s_active[b]) {
                const uint32_t *blk = yI + (2 + (b >> 3) * 8) * C2_YPW + 4 + (b & 7) * 4;
                int             c4[4];
                if (half == 0)
                    cdef2_dir_costs4<0>(blk, C2_YPW, cs, c4);
                else
                    cdef2_dir_costs4<4>(blk, C2_YPW, cs, c4);
#pragma unroll
                for (int i = 0; i < 4; i++) s_cost[b][half * 4 + i] = c4[i];
            }
        }
        __syncthreads();
        if (tid < 64 && process && s_active[tid]) {
            int best = 0, bd = 0;
#pragma unroll
            for (int i = 0; i < 8; i++) {
                const int c = s_cost[tid][i];
                if (c > best) { // first maximum wins ties (Common/Codec/EbCdef.c:197-201)
                    best = c;
                    bd   = i;
                }
            }
            s_dir[tid] = (uint8_t)bd;
            s_var[tid] = (best - s_cost[tid][(bd + 4) & 7]) >> 10;
        }
        __syncthreads();

        const uint16_t *lum16   = reinterpret_cast<const uint16_t *>(yI);
        const bool      edge_lr = x0 - 5 < 0 || x0 + 69 > W; // CCSO tap columns can leave the picture: per-sample clamping
        Ccso2Tile       ct;
        ct.rmin = 2 - y0, ct.rmax = H - 1 - (y0 - 2), ct.cmin = 8 - x0, ct.cmax = W - 1 - (x0 - 8);
        const int bx = tid & 7;

        // ---- luma: thread = rows r and r+4 of one 8x8 block; a warp = one row of blocks ----
        {
            const int  rp = (tid >> 3) & 3, by = tid >> 5, b = by * 8 + bx;
            const int  gx = x0 + bx * 8;
            const bool inside = gx < cw && y0 + by * 8 < ch; // coded sizes are multiples of 8
            const bool on     = process && (ylev | ysec) && inside && s_active[b];
            int2       o[6];
            Tap3I      pri_i, sec_i;
            Tap3F      pri_f, sec_f;
            uint32_t   wi0 = 0, wi1 = 0;
            __half2    wf0 = u2h(0u), wf1 = u2h(0u);
            if (on) {
                const int pri_l = ylev << cs, sec = ysec << cs, damp = a.damping + cs;
                const int var = s_var[b]; // adjust_strength — Common/Codec/EbCdef.c:130-134
                const int ii  = (var >> 6) ? min(cdef_msb(var >> 6), 12) : 0;
                const int pri = var ? (pri_l * (4 + ii) + 8) >> 4 : 0;
                const int dir = pri_l ? s_dir[b] : 0;
                const int odd = (pri >> cs) & 1; // cdef_pri_taps {4,2} / {3,3}
#pragma unroll
                for (int k = 0; k < 2; k++) {
                    o[k * 3 + 0] = s_off[0][dir][k];
                    o[k * 3 + 1] = s_off[0][(dir + 2) & 7][k];
                    o[k * 3 + 2] = s_off[0][(dir + 6) & 7][k];
                }
                if (FP) {
                    pri_f = tap3f(pri, damp), sec_f = tap3f(sec, damp);
                    wf0 = odd ? u2h(0x42004200u) : u2h(0x44004400u), wf1 = odd ? u2h(0x42004200u) : u2h(0x40004000u);
                } else {
                    pri_i = tap3i(pri, damp), sec_i = tap3i(sec, damp);
                    wi0 = odd ? 3u : 4u, wi1 = odd ? 3u : 2u;
                }
            }
            const bool    cc_on = (q_on & 1) != 0;
            const Ccso2Pk qp    = ccso2_pk(a.ccso, 0, a.d.bit_depth, s_lut[0]);
            if (inside) {
#pragma unroll 1
                for (int h = 0; h < 2; h++) {
                    const int row = rp + 4 * h, gy = y0 + by * 8 + row;
                    const int trow = 2 + by * 8 + row, tcol = 8 + bx * 8;
                    uint32_t  o32[4];
                    if (on) {
                        const uint32_t *ctr = yA + trow * C2_YPW + (tcol >> 1);
                        if (FP)
                            cdef3_row_f<4>(ctr, o, pri_f, sec_f, wf0, wf1, o32);
                        else
                            cdef3_row_i<4>(ctr, o, pri_i, sec_i, wi0, wi1, o32);
                    } else {
                        const uint32_t *c = yI + trow * C2_YPW + (tcol >> 1);
#pragma unroll
                        for (int j = 0; j < 4; j++) o32[j] = c[j];
                    }
                    if (cc_on && gy < H) {
                        if (edge_lr) {
#pragma unroll
                            for (int j = 0; j < 8; j++)
                                if (gx + j < W) {
                                    const int v = (o32[j >> 1] >> (16 * (j & 1))) & 0xffff;
                                    const int r = ccso2_px(lum16, 2 * C2_YPW, trow, tcol + j, qp, ct, v);
                                    o32[j >> 1] = (j & 1) ? ((o32[j >> 1] & 0xffffu) | ((uint32_t)r << 16)) : ((o32[j >> 1] & 0xffff0000u) | (uint32_t)r);
                                }
                        } else {
                            const uint32_t *yc = yI + trow * C2_YPW + (tcol >> 1);
                            uint32_t        ta[4], tb[4];
                            c3_ld_pairs<4>(yI, C2_YPW, min(max(trow + qp.dy0, ct.rmin), ct.rmax), tcol + qp.dx0, ta);
                            c3_ld_pairs<4>(yI, C2_YPW, min(max(trow - qp.dy0, ct.rmin), ct.rmax), tcol - qp.dx0, tb);
#pragma unroll
                            for (int j = 0; j < 4; j++) o32[j] = ccso2_pair(yc[j], ta[j], tb[j], o32[j], qp);
                        }
                    }
                    st8<T>(outy + (size_t)gy * a.out.pitch[0] + gx, make_uint4(o32[0], o32[1], o32[2], o32[3]));
                }
            }
        }
        // ---- chroma: thread = rows r and r+2 of one 4x4 block of one plane ----
        {
            const int       p = tid >> 7, by = (tid >> 4) & 7, rp = (tid >> 3) & 1, b = by * 8 + bx;
            const int       gx = x0 / 2 + bx * 4;
            const bool      inside = gx < cw / 2 && y0 / 2 + by * 4 < ch / 2;
            const bool      on     = process && (uvlev | uvsec) && inside && s_active[b];
            const uint32_t *cA     = cA0 + p * 2 * C3_CSZ;
            int2            o[6];
            Tap3I           pri_i, sec_i;
            Tap3F           pri_f, sec_f;
            uint32_t        wi0 = 0, wi1 = 0;
            __half2         wf0 = u2h(0u), wf1 = u2h(0u);
            if (on) {
                const int pri = uvlev << cs, sec = uvsec << cs, damp = a.damping + cs - 1;
                const int dir = pri ? s_dir[b] : 0;
                const int odd = (pri >> cs) & 1;
#pragma unroll
                for (int k = 0; k < 2; k++) {
                    o[k * 3 + 0] = s_off[1][dir][k];
                    o[k * 3 + 1] = s_off[1][(dir + 2) & 7][k];
                    o[k * 3 + 2] = s_off[1][(dir + 6) & 7][k];
                }
                if (FP) {
                    pri_f = tap3f(pri, damp), sec_f = tap3f(sec, damp);
                    wf0 = odd ? u2h(0x42004200u) : u2h(0x44004400u), wf1 = odd ? u2h(0x42004200u) : u2h(0x40004000u);
                } else {
                    pri_i = tap3i(pri, damp), sec_i = tap3i(sec, damp);
                    wi0 = odd ? 3u : 4u, wi1 = odd ? 3u : 2u;
                }
            }
            const bool    cc_on = ((q_on >> (1 + p)) & 1) != 0;
            const Ccso2Pk qp    = ccso2_pk(a.ccso, 1 + p, a.d.bit_depth, s_lut[1 + p]);
            T            *outc  = static_cast<T *>(a.out.p[1 + p]);
            if (inside) {
#pragma unroll 1
                for (int h = 0; h < 2; h++) {
                    const int row = rp + 2 * h, gy = y0 / 2 + by * 4 + row;
                    const int trow = 2 + by * 4 + row, tcol = 8 + bx * 4;
                    uint32_t  o32[2];
                    const uint32_t *ctr = cA + trow * C2_CPW + (tcol >> 1);
                    if (on) {
                        if (FP)
                            cdef3_row_f<2>(ctr, o, pri_f, sec_f, wf0, wf1, o32);
                        else
                            cdef3_row_i<2>(ctr, o, pri_i, sec_i, wi0, wi1, o32);
                    } else {
                        o32[0] = FP ? pk_h2i(ctr[0]) : ctr[0], o32[1] = FP ? pk_h2i(ctr[1]) : ctr[1];
                    }
                    if (cc_on && gy < (H >> 1)) {
                        const int lr = 2 + 2 * (by * 4 + row), lc = 8 + 8 * bx; // co-located luma sample of chroma sample 0
                        if (edge_lr) {
#pragma unroll
                            for (int j = 0; j < 4; j++)
                                if (gx + j < (W >> 1)) {
                                    const int v = (o32[j >> 1] >> (16 * (j & 1))) & 0xffff;
                                    const int r = ccso2_px(lum16, 2 * C2_YPW, lr, lc + 2 * j, qp, ct, v);
                                    o32[j >> 1] = (j & 1) ? ((o32[j >> 1] & 0xffffu) | ((uint32_t)r << 16)) : ((o32[j >> 1] & 0xffff0000u) | (uint32_t)r);
                                }
                        } else {
                            const int ra = min(max(lr + qp.dy0, ct.rmin), ct.rmax), rb = min(max(lr - qp.dy0, ct.rmin), ct.rmax);
#pragma unroll
                            for (int k = 0; k < 2; k++)
                                o32[k] = ccso2_pair(c3_ld_skip(lum16, 2 * C2_YPW, lr, lc + 4 * k), c3_ld_skip(lum16, 2 * C2_YPW, ra, lc + 4 * k + qp.dx0),
                                                    c3_ld_skip(lum16, 2 * C2_YPW, rb, lc + 4 * k - qp.dx0), o32[k], qp);
                        }
                    }
                    st4<T>(outc + (size_t)gy * a.out.pitch[1 + p] + gx, make_uint2(o32[0], o32[1]));
                }
            }
        }
        __syncthreads(); // every thread is done with this tile's copies before the next tile's conversion pass rewrites them
        tile = nxt;
    }
    // the last CTA to leave rearms the scheduler for the next launch on this context
    if (tid == 0 && atomicAdd(&g.sched[1], 1u) == gridDim.x - 1) {
        g.sched[0] = 0;
        g.sched[1] = 0;
    }
}

} // namespace svtf
