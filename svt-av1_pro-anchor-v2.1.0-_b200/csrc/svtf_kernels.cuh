// svtf_kernels.cuh — sm_100a kernels of the in-loop filter chain.
//
//   deblock_kernel<T>   one CTA per 64x64 output tile per plane; stages an 80x80 tile in shared
//                       memory, runs the column pass (vertical edges) then the row pass
//                       (horizontal edges) entirely in shared memory, writes the tile once.
//   cdef_kernel<T>      one CTA per 64x64 filter block (luma + both chroma tiles); 2-sample halo,
//                       8x8 direction search, primary/secondary taps, out of place.
//   ccso_kernel<T>      one CTA per 64x64 luma tile; classifies the pre-CDEF luma neighbourhood,
//                       applies the offset LUT (shared memory) to Y, U and V.
//
// T is the sample type in HBM (uint8_t or uint16_t); shared-memory tiles are always 16-bit.
// All arithmetic is integer and restates the reference's C kernels (cited per function, paths
// relative to Source/Lib/ of the reference).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include "svtf_internal.h"

namespace svtf {

struct Planes {
    void *p[3];
    int   pitch[3]; // samples
};

struct Dims {
    int width, height;    // unpadded luma size
    int mi_cols, mi_rows; // coded size / 4
    int bit_depth;
};

// ---- packed mi word (include/svtfilt_b200.h) -------------------------------------------------
__device__ __forceinline__ int mi_tx(svtf_mi_t m, int plane, int dir) { return (m >> ((plane ? 6 : 0) + (dir ? 3 : 0))) & 7; }
__device__ __forceinline__ int mi_bdim(svtf_mi_t m, int dir) { return (m >> (dir ? 15 : 12)) & 7; }
__device__ __forceinline__ int mi_skip_inter(svtf_mi_t m) { return (m >> 18) & 1; }
__device__ __forceinline__ int mi_skip(svtf_mi_t m) { return (m >> 19) & 1; }
__device__ __forceinline__ int mi_lvl_key(svtf_mi_t m) { return (m >> 20) & 0x7f; } // ref0(3) | mode_lf(1) | seg(3)

// ---- 8-sample vector access -------------------------------------------------------------------
// Loads 8 consecutive samples (16-byte aligned for u16, 8-byte for u8) as four packed u16 pairs.
template <typename T>
__device__ __forceinline__ uint4 ld8(const T *p);
template <>
__device__ __forceinline__ uint4 ld8<uint16_t>(const uint16_t *p) {
    return __ldg(reinterpret_cast<const uint4 *>(p));
}
template <>
__device__ __forceinline__ uint4 ld8<uint8_t>(const uint8_t *p) {
    const uint2 v = __ldg(reinterpret_cast<const uint2 *>(p));
    uint4       r;
    r.x = __byte_perm(v.x, 0, 0x4140);
    r.y = __byte_perm(v.x, 0, 0x4342);
    r.z = __byte_perm(v.y, 0, 0x4140);
    r.w = __byte_perm(v.y, 0, 0x4342);
    return r;
}
template <typename T>
__device__ __forceinline__ void st8(T *p, uint4 v);
template <>
__device__ __forceinline__ void st8<uint16_t>(uint16_t *p, uint4 v) {
    *reinterpret_cast<uint4 *>(p) = v;
}
template <>
__device__ __forceinline__ void st8<uint8_t>(uint8_t *p, uint4 v) {
    uint2 r;
    r.x = __byte_perm(v.x, v.y, 0x6420);
    r.y = __byte_perm(v.z, v.w, 0x6420);
    *reinterpret_cast<uint2 *>(p) = r;
}
// 4 samples
template <typename T>
__device__ __forceinline__ uint2 ld4(const T *p);
template <>
__device__ __forceinline__ uint2 ld4<uint16_t>(const uint16_t *p) {
    return __ldg(reinterpret_cast<const uint2 *>(p));
}
template <>
__device__ __forceinline__ uint2 ld4<uint8_t>(const uint8_t *p) {
    const uint32_t v = __ldg(reinterpret_cast<const uint32_t *>(p));
    return make_uint2(__byte_perm(v, 0, 0x4140), __byte_perm(v, 0, 0x4342));
}
template <typename T>
__device__ __forceinline__ void st4(T *p, uint2 v);
template <>
__device__ __forceinline__ void st4<uint16_t>(uint16_t *p, uint2 v) {
    *reinterpret_cast<uint2 *>(p) = v;
}
template <>
__device__ __forceinline__ void st4<uint8_t>(uint8_t *p, uint2 v) {
    *reinterpret_cast<uint32_t *>(p) = __byte_perm(v.x, v.y, 0x6420);
}

// =================================================================================================
// DEBLOCK
// =================================================================================================
constexpr int DBK_TILE    = 64; // output tile (shifted up-left by 8 so no edge straddles the tile's ownership)
constexpr int DBK_LOAD    = 80; // staged tile
constexpr int DBK_PITCH   = 82; // shared-memory pitch in samples (41 words: odd -> row walks are conflict-free)
constexpr int DBK_THREADS = 256;
constexpr int DBK_EDGES   = 17; // edges at tile-relative 8,12,...,72
constexpr int DBK_VUNITS  = 18; // 4-row units covering staged rows 4..75

struct DeblockArgs {
    Planes           in, out;
    const svtf_mi_t *mi;
    int              mi_stride;
    Dims             d;
    int              tiles_x[3], tiles_y[3], tile_base[3]; // linearised (plane, ty, tx) grid
    int              ty0[3];                               // first tile row of the launch (stripe window)
    unsigned         plane_mask;                           // bit p: plane p is filtered (else copied)
    int              lim_w[3], lim_h[3];                   // units at or beyond these plane coordinates have no edges (set_lpf_parameters :175)
    // edge-code planes (dbk_edge_plane_kernel): one word per 4x4 unit of the luma plane [0] / of a chroma plane [1], vertical edge in the
    // low half, horizontal edge in the high half
    const uint32_t *edge[2];
    int             edge_pitch[2];
    LfTables         t;
    // per-superblock level tables [sb][plane][256] (delta_lf_present with the decoder's per-SB deltas) or nullptr
    const uint8_t *sb_lvl;
    int            sb_cols, sb_shift; // superblocks per row; log2(4x4 units per superblock side)
    // SSE instantiation (deblock level search, try_filter_frame + picture_sse_calculations): the filtered tile is not written but
    // measured against the source plane over [0, sse_w) x [0, sse_h); one launch covers ONE plane
    const void         *sse_src;
    int                 sse_pitch, sse_w, sse_h;
    unsigned long long *sse_out;
};

// One line across an edge; s[k*st], k=-7..6 = p6..p0,q0..q6 (16-bit samples in shared memory).
// len 4: svt_aom_[highbd_]lpf_*_4_c   Common/Codec/EbDeblockingCommon.c:242-272, :460-500 (filter4 :214-240 / :426-458)
// len 6: ..._6_c :306-338 / :655-700 (filter6 :274-288)   len 8: ..._8_c :340-372 / :502-544 (filter8 :290-304)
// len 14: mb_lpf_*_edge_w :787-865 / :590-653 (filter14 :762-785).  sh = bit_depth-8; the 8-bit kernels are
// the sh==0 instance of the high-bitdepth formulas.
__device__ __forceinline__ void lpf_line(uint16_t *s, const int st, const int len, const int sh, const int blimit,
                                         const int limit, const int thresh) {
    const int lim = limit << sh, blim = blimit << sh, one = 1 << sh;
    const int p1 = s[-2 * st], p0 = s[-st], q0 = s[0], q1 = s[st];
    if (abs(p1 - p0) > lim || abs(q1 - q0) > lim || abs(p0 - q0) * 2 + (abs(p1 - q1) >> 1) > blim)
        return; // mask == 0: every branch below degenerates to filter4 with a zero filter value
    if (len == 6) {
        const int p2 = s[-3 * st], q2 = s[2 * st];
        if (abs(p2 - p1) > lim || abs(q2 - q1) > lim)
            return;
        if (!(abs(p1 - p0) > one || abs(q1 - q0) > one || abs(p2 - p0) > one || abs(q2 - q0) > one)) {
            s[-2 * st] = (uint16_t)((p2 * 3 + p1 * 2 + p0 * 2 + q0 + 4) >> 3);
            s[-st]     = (uint16_t)((p2 + p1 * 2 + p0 * 2 + q0 * 2 + q1 + 4) >> 3);
            s[0]       = (uint16_t)((p1 + p0 * 2 + q0 * 2 + q1 * 2 + q2 + 4) >> 3);
            s[st]      = (uint16_t)((p0 + q0 * 2 + q1 * 2 + q2 * 3 + 4) >> 3);
            return;
        }
    } else if (len >= 8) {
        const int p2 = s[-3 * st], p3 = s[-4 * st], q2 = s[2 * st], q3 = s[3 * st];
        if (abs(p3 - p2) > lim || abs(p2 - p1) > lim || abs(q2 - q1) > lim || abs(q3 - q2) > lim)
            return;
        const bool flat = !(abs(p1 - p0) > one || abs(q1 - q0) > one || abs(p2 - p0) > one || abs(q2 - q0) > one ||
                            abs(p3 - p0) > one || abs(q3 - q0) > one);
        if (flat && len == 14) {
            const int  p4 = s[-5 * st], p5 = s[-6 * st], p6 = s[-7 * st], q4 = s[4 * st], q5 = s[5 * st], q6 = s[6 * st];
            const bool flat2 = !(abs(p4 - p0) > one || abs(q4 - q0) > one || abs(p5 - p0) > one || abs(q5 - q0) > one ||
                                 abs(p6 - p0) > one || abs(q6 - q0) > one);
            if (flat2) {
                s[-6 * st] = (uint16_t)((p6 * 7 + p5 * 2 + p4 * 2 + p3 + p2 + p1 + p0 + q0 + 8) >> 4);
                s[-5 * st] = (uint16_t)((p6 * 5 + p5 * 2 + p4 * 2 + p3 * 2 + p2 + p1 + p0 + q0 + q1 + 8) >> 4);
                s[-4 * st] = (uint16_t)((p6 * 4 + p5 + p4 * 2 + p3 * 2 + p2 * 2 + p1 + p0 + q0 + q1 + q2 + 8) >> 4);
                s[-3 * st] = (uint16_t)((p6 * 3 + p5 + p4 + p3 * 2 + p2 * 2 + p1 * 2 + p0 + q0 + q1 + q2 + q3 + 8) >> 4);
                s[-2 * st] = (uint16_t)((p6 * 2 + p5 + p4 + p3 + p2 * 2 + p1 * 2 + p0 * 2 + q0 + q1 + q2 + q3 + q4 + 8) >> 4);
                s[-st]     = (uint16_t)((p6 + p5 + p4 + p3 + p2 + p1 * 2 + p0 * 2 + q0 * 2 + q1 + q2 + q3 + q4 + q5 + 8) >> 4);
                s[0]       = (uint16_t)((p5 + p4 + p3 + p2 + p1 + p0 * 2 + q0 * 2 + q1 * 2 + q2 + q3 + q4 + q5 + q6 + 8) >> 4);
                s[st]      = (uint16_t)((p4 + p3 + p2 + p1 + p0 + q0 * 2 + q1 * 2 + q2 * 2 + q3 + q4 + q5 + q6 * 2 + 8) >> 4);
                s[2 * st]  = (uint16_t)((p3 + p2 + p1 + p0 + q0 + q1 * 2 + q2 * 2 + q3 * 2 + q4 + q5 + q6 * 3 + 8) >> 4);
                s[3 * st]  = (uint16_t)((p2 + p1 + p0 + q0 + q1 + q2 * 2 + q3 * 2 + q4 * 2 + q5 + q6 * 4 + 8) >> 4);
                s[4 * st]  = (uint16_t)((p1 + p0 + q0 + q1 + q2 + q3 * 2 + q4 * 2 + q5 * 2 + q6 * 5 + 8) >> 4);
                s[5 * st]  = (uint16_t)((p0 + q0 + q1 + q2 + q3 + q4 * 2 + q5 * 2 + q6 * 7 + 8) >> 4);
                return;
            }
        }
        if (flat) {
            s[-3 * st] = (uint16_t)((p3 + p3 + p3 + 2 * p2 + p1 + p0 + q0 + 4) >> 3);
            s[-2 * st] = (uint16_t)((p3 + p3 + p2 + 2 * p1 + p0 + q0 + q1 + 4) >> 3);
            s[-st]     = (uint16_t)((p3 + p2 + p1 + 2 * p0 + q0 + q1 + q2 + 4) >> 3);
            s[0]       = (uint16_t)((p2 + p1 + p0 + 2 * q0 + q1 + q2 + q3 + 4) >> 3);
            s[st]      = (uint16_t)((p1 + p0 + q0 + 2 * q1 + q2 + q3 + q3 + 4) >> 3);
            s[2 * st]  = (uint16_t)((p0 + q0 + q1 + 2 * q2 + q3 + q3 + q3 + 4) >> 3);
            return;
        }
    }
    // filter4 with mask == all ones
    const int thr = thresh << sh, off = 0x80 << sh, lo = -(128 << sh), hi = (128 << sh) - 1;
    const int ps1 = p1 - off, ps0 = p0 - off, qs0 = q0 - off, qs1 = q1 - off;
    const int hev = (abs(p1 - p0) > thr || abs(q1 - q0) > thr) ? -1 : 0;
    int       f   = min(max(ps1 - qs1, lo), hi) & hev;
    f             = min(max(f + 3 * (qs0 - ps0), lo), hi);
    const int f1  = min(f + 4, hi) >> 3;
    const int f2  = min(f + 3, hi) >> 3;
    s[0]          = (uint16_t)(min(max(qs0 - f1, lo), hi) + off);
    s[-st]        = (uint16_t)(min(max(ps0 + f2, lo), hi) + off);
    f             = ((f1 + 1) >> 1) & ~hev;
    s[st]         = (uint16_t)(min(max(qs1 - f, lo), hi) + off);
    s[-2 * st]    = (uint16_t)(min(max(ps1 + f, lo), hi) + off);
}

// set_lpf_parameters — Encoder/Codec/EbDeblockingFilter.c:162-285, evaluated independently per 4-sample unit (the reference walks by
// transform size; skipped units are never on a transform edge), in two steps.
// Step 1, once per mode-info upload (dbk_edge_plane_kernel): what depends on the mode info alone — is the unit's left / top boundary
// (coord = its x or y) a transform edge, does the skip / prediction-block rule let it filter, which filter length.  Entry, 16 bits per
// direction:  bits 0-1  0 = no edge, 1 = 4-tap, 2 = 8-tap (luma) / 6-tap (chroma), 3 = 14-tap;  bits 2-8 / 9-15  level key of the unit /
// of its left / top neighbour.  One word per unit (vertical edge low half, horizontal high half); one plane for luma, one for U and V
// (same transform sizes and mode info, only the level tables differ).
// Step 2, in every tile of every deblock launch (the level search runs 24 per picture): the levels of the two keys (frame table or
// per-superblock table) and level != 0.
__device__ __forceinline__ uint32_t dbk_edge_entry(const svtf_mi_t cur, const svtf_mi_t prv, const int plane, const int dir, const int coord) {
    const int ts = mi_tx(cur, plane, dir);
    if (coord == 0 || (coord & ((4 << ts) - 1)))
        return 0;
    int bdim = 4 << mi_bdim(cur, dir);
    if (plane)
        bdim = max(4, bdim >> 1);
    const bool pu_edge = !(coord & (bdim - 1));
    if (mi_skip_inter(prv) && mi_skip_inter(cur) && !pu_edge)
        return 0;
    const int min_ts = min(ts, mi_tx(prv, plane, dir));
    const int code   = min_ts == 0 ? 1 : (plane ? 2 : (min_ts == 1 ? 2 : 3));
    return (uint32_t)code | ((uint32_t)mi_lvl_key(cur) << 2) | ((uint32_t)mi_lvl_key(prv) << 9);
}
struct EdgePlaneArgs {
    const svtf_mi_t *mi;
    int              mi_stride, mi_cols, mi_rows;
    uint32_t        *ey, *ec;
    int              ey_pitch, ec_pitch;
};
// thread = one mode-info word: its luma entry and, at odd (row, col), the entry of the chroma unit it stands for (deblock_kernel's
// mi_row = 1 | 2 * unit row)
__global__ void __launch_bounds__(256) dbk_edge_plane_kernel(const EdgePlaneArgs a) {
    const int c = blockIdx.x * 64 + (threadIdx.x & 63), r = blockIdx.y * 4 + (threadIdx.x >> 6);
    if (r >= a.mi_rows || c >= a.mi_cols)
        return;
    const svtf_mi_t *mp = a.mi + r * a.mi_stride + c;
    const bool       ch = r & c & 1;
    // all loads first (independent), neighbours outside the picture read as the unit itself: their edges have coord 0
    const svtf_mi_t cur = __ldg(mp), lft = __ldg(c ? mp - 1 : mp), top = __ldg(r ? mp - a.mi_stride : mp);
    const svtf_mi_t lft2 = ch && c > 1 ? __ldg(mp - 2) : cur, top2 = ch && r > 1 ? __ldg(mp - 2 * a.mi_stride) : cur;
    a.ey[r * a.ey_pitch + c] = dbk_edge_entry(cur, lft, 0, 0, 4 * c) | (dbk_edge_entry(cur, top, 0, 1, 4 * r) << 16);
    if (ch) {
        const int cr = r >> 1, cc = c >> 1;
        a.ec[cr * a.ec_pitch + cc] = dbk_edge_entry(cur, lft2, 1, 0, 4 * cc) | (dbk_edge_entry(cur, top2, 1, 1, 4 * cr) << 16);
    }
}

// SBLVL: per-superblock level tables (decoder delta_lf); SSE: measure against the source instead of writing (level search).
// Both are separate instantiations so the common path keeps its 32 registers.
template <typename T, bool SBLVL = false, bool SSE = false>
__global__ void __launch_bounds__(DBK_THREADS, 8) deblock_kernel(const __grid_constant__ DeblockArgs a) {
    __shared__ __align__(16) uint16_t tile[DBK_LOAD * DBK_PITCH];
    // work lists of the two passes, one per filter code (1 = 4-tap, 2 = 8/6-tap, 3 = 14-tap): entry = level << 10 | unit << 5 | edge.
    // Lanes then work on edges that exist, four lines of a 4-sample unit next to each other, one filter length per list.
    __shared__ uint16_t s_list[2][3][DBK_VUNITS * DBK_EDGES + 2];
    __shared__ int      s_cnt[2][3];
    __shared__ uint8_t  s_lvl[256];
    __shared__ uint8_t  s_thr[3][64]; // lim, mblim, hev

    const int tid   = threadIdx.x;
    const int bid   = blockIdx.x;
    const int plane = bid >= a.tile_base[2] ? 2 : (bid >= a.tile_base[1] ? 1 : 0);
    const int lb    = bid - a.tile_base[plane];
    const int tyl = lb / a.tiles_x[plane], tx = lb - tyl * a.tiles_x[plane];
    const int ty  = tyl + a.ty0[plane];
    const int ss = plane ? 1 : 0;
    const int cw = (a.d.mi_cols * 4) >> ss, ch = (a.d.mi_rows * 4) >> ss;
    const int pw = a.lim_w[plane], ph = a.lim_h[plane];
    const int X0 = tx * DBK_TILE - 16, Y0 = ty * DBK_TILE - 16;
    const T  *in    = static_cast<const T *>(a.in.p[plane]);
    T        *out   = static_cast<T *>(a.out.p[plane]);
    const int ipitch = a.in.pitch[plane], opitch = a.out.pitch[plane];
    const bool filter = (a.plane_mask >> plane) & 1;

    // ---- stage: 80x80 samples, 8-sample vectors, zero outside the coded plane ----
    for (int i = tid; i < DBK_LOAD * (DBK_LOAD / 8); i += DBK_THREADS) {
        const int r = i / (DBK_LOAD / 8), c = (i - r * (DBK_LOAD / 8)) * 8;
        const int gy = Y0 + r, gx = X0 + c;
        uint4     v  = make_uint4(0, 0, 0, 0);
        if (gy >= 0 && gy < ch && gx >= 0 && gx + 8 <= ipitch)
            v = ld8<T>(in + (size_t)gy * ipitch + gx);
        uint32_t *d = reinterpret_cast<uint32_t *>(tile + r * DBK_PITCH + c);
        d[0] = v.x, d[1] = v.y, d[2] = v.z, d[3] = v.w;
    }
    if (filter) {
        // frame level table re-indexed [dir][key], key = ref | mode_lf << 3 | segment << 4 as the edge entries hold it
        s_lvl[tid] = (&a.t.lvl[plane][0][0][0][0])[((((tid >> 4) & 7) * 2 + (tid >> 7)) * 8 + (tid & 7)) * 2 + ((tid >> 3) & 1)];
        if (tid < 64) {
            s_thr[0][tid] = a.t.lim[tid];
            s_thr[1][tid] = a.t.mblim[tid];
            s_thr[2][tid] = a.t.hev[tid];
        }
        if (tid < 6)
            (&s_cnt[0][0])[tid] = 0;
        __syncthreads();
        // ---- edge parameters -> work lists.  One thread per 4x4 unit of the tile decides its left (vertical) and its top
        //      (horizontal) edge from one read of the unit's mode info: units (ur, uc) = plane position (Y0 + 4ur, X0 + 4uc),
        //      ur 1..18, uc 2..18; vertical edges: row unit ur-1, edge uc-2; horizontal: edge ur-2 (ur >= 2), col unit uc-2 (<= 15)
        constexpr int NUC = DBK_EDGES;
        for (int i = tid; i < DBK_VUNITS * NUC; i += DBK_THREADS) {
            const int ur = 1 + i / NUC, uc = 2 + (i - (i / NUC) * NUC);
            const int x = X0 + 4 * uc, y = Y0 + 4 * ur;
            if (x < 0 || y < 0 || x >= pw || y >= ph)
                continue;
            const int      mi_row = ss | ((y << ss) >> 2), mi_col = ss | ((x << ss) >> 2);
            const uint32_t e      = __ldg(a.edge[ss] + (y >> 2) * a.edge_pitch[ss] + (x >> 2));
            // level of a key: frame table, or the table of the superblock that CONTAINS the unit (for the unit and for its left / top
            // neighbour alike: sb_delta_lf / sb_delta_lf_left / _above, Decoder/Codec/EbDecLF.c:357,548)
            auto level_of = [&](const int key, const int dir, const int r, const int c) {
                if (SBLVL) {
                    const int idx = (((key >> 4) * 2 + dir) * 8 + (key & 7)) * 2 + ((key >> 3) & 1);
                    return (int)__ldg(a.sb_lvl + ((size_t)((r >> a.sb_shift) * a.sb_cols + (c >> a.sb_shift)) * 3 + plane) * 256 + idx);
                }
                return (int)s_lvl[dir * 128 + key];
            };
            const uint32_t ev = e & 0xffffu, eh = e >> 16;
            if (ev & 3) {
                int lv = level_of((ev >> 2) & 127, 0, mi_row, mi_col);
                if (!lv)
                    lv = level_of((ev >> 9) & 127, 0, mi_row, mi_col - (1 << ss));
                if (lv) {
                    const int code = (ev & 3) - 1;
                    s_list[0][code][atomicAdd(&s_cnt[0][code], 1)] = (uint16_t)((lv << 10) | ((ur - 1) << 5) | (uc - 2));
                }
            }
            if (ur >= 2 && uc <= 17 && (eh & 3)) {
                int lv = level_of(eh >> 2 & 127, 1, mi_row, mi_col);
                if (!lv)
                    lv = level_of(eh >> 9, 1, mi_row - (1 << ss), mi_col);
                if (lv) {
                    const int code = (eh & 3) - 1;
                    s_list[1][code][atomicAdd(&s_cnt[1][code], 1)] = (uint16_t)((lv << 10) | ((uc - 2) << 5) | (ur - 2));
                }
            }
        }
        __syncthreads();
        const int sh = a.d.bit_depth - 8;
        // ---- column pass: vertical edges; thread = one line (row) of a listed unit ----
#pragma unroll
        for (int code = 0; code < 3; code++) {
            const int n   = s_cnt[0][code] * 4;
            const int len = code == 0 ? 4 : (code == 2 ? 14 : (plane ? 6 : 8));
            for (int i = tid; i < n; i += DBK_THREADS) {
                const unsigned e  = s_list[0][code][i >> 2];
                const int      lv = e >> 10, r = ((e >> 5) & 31) * 4 + (i & 3), k = e & 31;
                lpf_line(tile + (r + 4) * DBK_PITCH + 8 + 4 * k, 1, len, sh, s_thr[1][lv], s_thr[0][lv], s_thr[2][lv]);
            }
        }
        __syncthreads();
        // ---- row pass: horizontal edges; thread = one line (column) of a listed unit ----
#pragma unroll
        for (int code = 0; code < 3; code++) {
            const int n   = s_cnt[1][code] * 4;
            const int len = code == 0 ? 4 : (code == 2 ? 14 : (plane ? 6 : 8));
            for (int i = tid; i < n; i += DBK_THREADS) {
                const unsigned e  = s_list[1][code][i >> 2];
                const int      lv = e >> 10, c = ((e >> 5) & 31) * 4 + (i & 3), k = e & 31;
                lpf_line(tile + (8 + 4 * k) * DBK_PITCH + 8 + c, DBK_PITCH, len, sh, s_thr[1][lv], s_thr[0][lv], s_thr[2][lv]);
            }
        }
    }
    __syncthreads();
    if (SSE) { // ---- squared error of the tile against the source (svt_spatial_full_distortion_kernel / _16bits) ----
        const T           *src = static_cast<const T *>(a.sse_src);
        unsigned long long acc = 0;
        for (int i = tid; i < DBK_TILE * (DBK_TILE / 8); i += DBK_THREADS) {
            const int r = i >> 3, c = (i & 7) * 8;
            const int gy = Y0 + 8 + r, gx = X0 + 8 + c;
            if (gy < 0 || gy >= a.sse_h || gx < 0 || gx >= a.sse_w)
                continue;
            const uint16_t *s = tile + (r + 8) * DBK_PITCH + 8 + c;
            unsigned        e = 0;
            if (gx + 8 <= a.sse_w) {
                const uint4    v    = ld8<T>(src + (size_t)gy * a.sse_pitch + gx);
                const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    const int d0 = (int)s[2 * k] - (int)(w[k] & 0xffff), d1 = (int)s[2 * k + 1] - (int)(w[k] >> 16);
                    e += (unsigned)(d0 * d0) + (unsigned)(d1 * d1);
                }
            } else {
                for (int k = 0; gx + k < a.sse_w; k++) {
                    const int d = (int)s[k] - (int)src[(size_t)gy * a.sse_pitch + gx + k];
                    e += (unsigned)(d * d);
                }
            }
            acc += e;
        }
#pragma unroll
        for (int m = 16; m > 0; m >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, m);
        __shared__ unsigned long long s_sse[DBK_THREADS / 32];
        if ((tid & 31) == 0)
            s_sse[tid >> 5] = acc;
        __syncthreads();
        if (tid == 0) {
            unsigned long long t = 0;
            for (int i = 0; i < DBK_THREADS / 32; i++) t += s_sse[i];
            if (t)
                atomicAdd(a.sse_out, t);
        }
        return;
    }
    // ---- write the 64x64 tile (staged rows/cols 8..71) ----
    for (int i = tid; i < DBK_TILE * (DBK_TILE / 8); i += DBK_THREADS) {
        const int r = i >> 3, c = (i & 7) * 8;
        const int gy = Y0 + 8 + r, gx = X0 + 8 + c;
        if (gy >= 0 && gy < ch && gx >= 0 && gx < cw && gx + 8 <= opitch) {
            const uint32_t *s = reinterpret_cast<const uint32_t *>(tile + (r + 8) * DBK_PITCH + 8 + c);
            st8<T>(out + (size_t)gy * opitch + gx, make_uint4(s[0], s[1], s[2], s[3]));
        }
    }
}

// =================================================================================================
// CDEF
// =================================================================================================
constexpr int CDEF_THREADS    = 256;
constexpr int CDEF_LARGE      = 0x7F7F; // CDEF_VERY_LARGE, Common/Codec/EbCdef.h:39
constexpr int CDEF_YP         = 70;     // luma tile pitch: 64 + 2*2 samples, padded to 35 words (odd)
constexpr int CDEF_YROWS      = 68;
constexpr int CDEF_CP         = 38; // chroma tile pitch: 32 + 2*2, 19 words (odd)
constexpr int CDEF_CROWS      = 36;

struct CdefArgs {
    Planes           in, out;
    const svtf_mi_t *mi;
    int              mi_stride;
    Dims             d;
    const int8_t    *fb_idx; // [nvfb*nhfb]
    int              fbr0;   // first filter-block row of the launch (stripe window)
    int              damping;
    uint8_t          y_strength[SVTF_CDEF_MAX_STRENGTHS], uv_strength[SVTF_CDEF_MAX_STRENGTHS];
};

// Cdef_Directions as (dy,dx) per tap k — Common/Codec/EbCdef.c:99-122
__constant__ int8_t c_cdef_dirs[8][2][2] = {{{-1, 1}, {-2, 2}}, {{0, 1}, {-1, 2}}, {{0, 1}, {0, 2}}, {{0, 1}, {1, 2}},
                                            {{1, 1}, {2, 2}},   {{1, 0}, {2, 1}},  {{1, 0}, {2, 0}}, {{1, 0}, {2, -1}}};

// Directional cost of one 8x8 block for direction D — svt_aom_cdef_find_dir_c, Common/Codec/EbCdef.c:150-210.
// cost = sum over lines of (line sum)^2 * 840/len(line); the reference's div_table weights are exactly
// 840/len, and no int32 overflow is possible, so the accumulation order is free.
template <int D>
__device__ __forceinline__ int cdef_line_of(int i, int j) {
    return D == 0 ? i + j : D == 1 ? i + j / 2 : D == 2 ? i : D == 3 ? 3 + i - j / 2 : D == 4 ? 7 + i - j : D == 5 ? 3 - i / 2 + j : D == 6 ? j : i / 2 + j;
}
template <int D>
__device__ __forceinline__ int cdef_dir_cost(const uint16_t *b, const int pitch, const int cs) {
    constexpr int NL = (D == 0 || D == 4) ? 15 : ((D & 1) ? 11 : 8);
    int           part[NL];
#pragma unroll
    for (int l = 0; l < NL; l++) part[l] = 0;
#pragma unroll
    for (int i = 0; i < 8; i++)
#pragma unroll
        for (int j = 0; j < 8; j++) part[cdef_line_of<D>(i, j)] += (b[i * pitch + j] >> cs) - 128;
    int cost = 0;
#pragma unroll
    for (int l = 0; l < NL; l++) {
        const int n = (D == 0 || D == 4) ? (l < 8 ? l + 1 : 15 - l) : ((D & 1) ? 2 * min(min(l + 1, 4), 11 - l) : 8);
        cost += part[l] * part[l] * (840 / n);
    }
    return cost;
}

__device__ __forceinline__ int cdef_msb(unsigned v) { return 31 - __clz(v); }

// constrain — Common/Codec/EbCdef.c:85-91 with the shift precomputed (max(0, damping - msb(threshold)))
__device__ __forceinline__ int cdef_constrain(const int diff, const int threshold, const int shift) {
    const int ad = abs(diff);
    const int v  = min(ad, max(0, threshold - (ad >> shift)));
    return diff < 0 ? -v : v;
}

// svt_cdef_filter_block_c for one sample — Common/Codec/EbCdef.c:253-304.  c: centre sample in a 16-bit tile.
// o[6]: offsets of primary k=0,1, secondary (dir+2) k=0,1, secondary (dir-2) k=0,1.
__device__ __forceinline__ int cdef_px(const uint16_t *c, const int *o, const int pri, const int sec, const int pri_shift,
                                       const int sec_shift, const int w0, const int w1) {
    const int x   = c[0];
    int       sum = 0, mx = x, mn = x;
#pragma unroll
    for (int k = 0; k < 2; k++) {
        const int p0 = c[o[k]], p1 = c[-o[k]];
        if (pri) {
            const int w = k ? w1 : w0;
            sum += w * cdef_constrain(p0 - x, pri, pri_shift);
            sum += w * cdef_constrain(p1 - x, pri, pri_shift);
        }
        if (p0 != CDEF_LARGE) mx = max(p0, mx);
        if (p1 != CDEF_LARGE) mx = max(p1, mx);
        mn = min(min(p0, p1), mn);
        const int s0 = c[o[2 + k]], s1 = c[-o[2 + k]], s2 = c[o[4 + k]], s3 = c[-o[4 + k]];
        if (s0 != CDEF_LARGE) mx = max(s0, mx);
        if (s1 != CDEF_LARGE) mx = max(s1, mx);
        if (s2 != CDEF_LARGE) mx = max(s2, mx);
        if (s3 != CDEF_LARGE) mx = max(s3, mx);
        mn = min(min(min(s0, s1), min(s2, s3)), mn);
        if (sec) {
            const int w = 2 - k; // sec taps {2,1}
            sum += w * (cdef_constrain(s0 - x, sec, sec_shift) + cdef_constrain(s1 - x, sec, sec_shift) +
                        cdef_constrain(s2 - x, sec, sec_shift) + cdef_constrain(s3 - x, sec, sec_shift));
        }
    }
    return min(max(x + ((8 + sum - (sum < 0)) >> 4), mn), mx);
}

__device__ __forceinline__ void cdef_offsets(int *o, const int dir, const int pitch) {
    const int d2 = (dir + 2) & 7, d3 = (dir + 6) & 7;
#pragma unroll
    for (int k = 0; k < 2; k++) {
        o[k]     = c_cdef_dirs[dir][k][0] * pitch + c_cdef_dirs[dir][k][1];
        o[2 + k] = c_cdef_dirs[d2][k][0] * pitch + c_cdef_dirs[d2][k][1];
        o[4 + k] = c_cdef_dirs[d3][k][0] * pitch + c_cdef_dirs[d3][k][1];
    }
}

// Stage a (rows x 2*pairs) tile whose origin is (x0-2, y0-2); samples outside the coded plane become
// CDEF_VERY_LARGE (Encoder/Codec/EbEncCdef.c:549-568).  Two samples per thread per step.
template <typename T>
__device__ __forceinline__ void cdef_stage(uint16_t *tile, const int pitch, const int rows, const int pairs, const T *in,
                                           const int ipitch, const int x0, const int y0, const int cw, const int ch,
                                           const int tid) {
    for (int i = tid; i < rows * pairs; i += CDEF_THREADS) {
        const int r = i / pairs, cp = i - r * pairs;
        const int gy = y0 - 2 + r, gx = x0 - 2 + 2 * cp;
        uint32_t  v  = CDEF_LARGE | (CDEF_LARGE << 16);
        if (gy >= 0 && gy < ch && gx >= 0 && gx < cw) {
            const T *p = in + (size_t)gy * ipitch + gx;
            if (sizeof(T) == 2)
                v = __ldg(reinterpret_cast<const uint32_t *>(p));
            else {
                const uint32_t b = __ldg(reinterpret_cast<const uint16_t *>(p));
                v                = (b & 0xff) | ((b & 0xff00) << 8);
            }
        }
        *reinterpret_cast<uint32_t *>(tile + r * pitch + 2 * cp) = v;
    }
}

// svt_av1_cdef_frame — Encoder/Codec/EbEncCdef.c:284-612 + svt_cdef_filter_fb Common/Codec/EbCdef.c:339-430
// (== decoder svt_cdef_frame, Decoder/Codec/EbDecCdef.c:507).  Out of place: every tap reads the
// pre-CDEF picture, which is what the reference's line/column buffers emulate.
#ifdef SVTF_AB_KERNELS // superseded kernel, kept for A/B measurements only (not in the product build)
template <typename T>
__global__ void __launch_bounds__(CDEF_THREADS) cdef_kernel(const __grid_constant__ CdefArgs a) {
    __shared__ __align__(16) uint16_t ytile[CDEF_YROWS * CDEF_YP];
    __shared__ __align__(16) uint16_t ctile[2][CDEF_CROWS * CDEF_CP];
    __shared__ int                    s_cost[8][64];
    __shared__ uint8_t                s_dir[64];
    __shared__ int                    s_var[64];
    __shared__ uint8_t                s_active[64];

    const int tid = threadIdx.x;
    const int fbc = blockIdx.x, fbr = blockIdx.y + a.fbr0;
    const int nhfb = gridDim.x;
    const int cw = a.d.mi_cols * 4, ch = a.d.mi_rows * 4;
    const int x0 = fbc * 64, y0 = fbr * 64;
    const int cs = a.d.bit_depth - 8;

    const int idx = a.fb_idx[fbr * nhfb + fbc];
    int       ylev = 0, ysec = 0, uvlev = 0, uvsec = 0;
    if (idx >= 0) {
        const int ys = a.y_strength[idx], us = a.uv_strength[idx];
        ylev = ys >> 2, ysec = ys & 3, ysec += ysec == 3;
        uvlev = us >> 2, uvsec = us & 3, uvsec += uvsec == 3;
    }
    // svt_sb_compute_cdef_list — Encoder/Codec/EbEncCdef.c:238-281: an 8x8 is filtered iff any of its 4 mi is not skipped
    int act = 0;
    if (tid < 64) {
        const int r = (tid >> 3) * 2, c = (tid & 7) * 2;
        if (fbr * 16 + r < a.d.mi_rows && fbc * 16 + c < a.d.mi_cols) {
            const svtf_mi_t *m = a.mi + (fbr * 16 + r) * a.mi_stride + fbc * 16 + c;
            act = !(mi_skip(__ldg(m)) & mi_skip(__ldg(m + 1)) & mi_skip(__ldg(m + a.mi_stride)) & mi_skip(__ldg(m + a.mi_stride + 1)));
        }
        s_active[tid] = (uint8_t)act;
    }
    const int  any_active = __syncthreads_or(act);
    const bool process    = (ylev | ysec | uvlev | uvsec) && any_active;

    const T *iny = static_cast<const T *>(a.in.p[0]);
    T       *outy = static_cast<T *>(a.out.p[0]);
    if (!process) {
        // untouched filter block: carry the samples over to the output surface
        for (int i = tid; i < 64 * 8; i += CDEF_THREADS) {
            const int r = i >> 3, c = (i & 7) * 8;
            if (y0 + r < ch && x0 + c < cw)
                st8<T>(outy + (size_t)(y0 + r) * a.out.pitch[0] + x0 + c, ld8<T>(iny + (size_t)(y0 + r) * a.in.pitch[0] + x0 + c));
        }
        for (int i = tid; i < 2 * 32 * 8; i += CDEF_THREADS) {
            const int p = 1 + (i >> 8), r = (i >> 3) & 31, c = (i & 7) * 4;
            if (y0 / 2 + r < ch / 2 && x0 / 2 + c < cw / 2)
                st4<T>(static_cast<T *>(a.out.p[p]) + (size_t)(y0 / 2 + r) * a.out.pitch[p] + x0 / 2 + c,
                       ld4<T>(static_cast<const T *>(a.in.p[p]) + (size_t)(y0 / 2 + r) * a.in.pitch[p] + x0 / 2 + c));
        }
        return;
    }

    cdef_stage<T>(ytile, CDEF_YP, CDEF_YROWS, 34, iny, a.in.pitch[0], x0, y0, cw, ch, tid);
    cdef_stage<T>(ctile[0], CDEF_CP, CDEF_CROWS, 18, static_cast<const T *>(a.in.p[1]), a.in.pitch[1], x0 / 2, y0 / 2, cw / 2, ch / 2, tid);
    cdef_stage<T>(ctile[1], CDEF_CP, CDEF_CROWS, 18, static_cast<const T *>(a.in.p[2]), a.in.pitch[2], x0 / 2, y0 / 2, cw / 2, ch / 2, tid);
    __syncthreads();

    // ---- direction search: warp w evaluates direction w for 2x32 blocks ----
    {
        const int w = tid >> 5, lane = tid & 31;
#pragma unroll
        for (int h = 0; h < 2; h++) {
            const int       b  = h * 32 + lane;
            const uint16_t *bp = ytile + (2 + (b >> 3) * 8) * CDEF_YP + 2 + (b & 7) * 8;
            int             c  = 0;
            if (s_active[b]) {
                switch (w) {
                case 0: c = cdef_dir_cost<0>(bp, CDEF_YP, cs); break;
                case 1: c = cdef_dir_cost<1>(bp, CDEF_YP, cs); break;
                case 2: c = cdef_dir_cost<2>(bp, CDEF_YP, cs); break;
                case 3: c = cdef_dir_cost<3>(bp, CDEF_YP, cs); break;
                case 4: c = cdef_dir_cost<4>(bp, CDEF_YP, cs); break;
                case 5: c = cdef_dir_cost<5>(bp, CDEF_YP, cs); break;
                case 6: c = cdef_dir_cost<6>(bp, CDEF_YP, cs); break;
                default: c = cdef_dir_cost<7>(bp, CDEF_YP, cs); break;
                }
            }
            s_cost[w][b] = c;
        }
    }
    __syncthreads();
    if (tid < 64) {
        int best = 0, bd = 0;
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const int c = s_cost[i][tid];
            if (c > best) { // first maximum wins ties (:197-201)
                best = c;
                bd   = i;
            }
        }
        s_dir[tid] = (uint8_t)bd;
        s_var[tid] = (best - s_cost[(bd + 4) & 7][tid]) >> 10;
    }
    __syncthreads();

    // ---- luma: work item = one 8-sample row of one 8x8 block ----
    {
        const int  pri_l = ylev << cs, sec = ysec << cs;
        const int  damp  = a.damping + cs;
        const bool on    = ylev | ysec;
        const int  sec_shift = sec ? max(0, damp - cdef_msb(sec)) : 0;
        for (int i = tid; i < 512; i += CDEF_THREADS) {
            // lanes 0..7 -> 8 blocks side by side, lanes /8 -> 4 consecutive rows
            const int bx = i & 7, row = (i >> 3) & 7, by = i >> 6;
            const int b  = by * 8 + bx;
            const int gy = y0 + by * 8 + row, gx = x0 + bx * 8;
            if (gy >= ch || gx >= cw)
                continue;
            const uint16_t *c = ytile + (2 + by * 8 + row) * CDEF_YP + 2 + bx * 8;
            uint32_t        o32[4];
            if (on && s_active[b]) {
                // adjust_strength — Common/Codec/EbCdef.c:130-134
                const int var = s_var[b];
                const int ii  = (var >> 6) ? min(cdef_msb(var >> 6), 12) : 0;
                const int pri = var ? (pri_l * (4 + ii) + 8) >> 4 : 0;
                int       o[6];
                cdef_offsets(o, pri_l ? s_dir[b] : 0, CDEF_YP);
                const int pri_shift = pri ? max(0, damp - cdef_msb(pri)) : 0;
                const int odd = (pri >> cs) & 1, w0 = odd ? 3 : 4, w1 = odd ? 3 : 2;
#pragma unroll
                for (int j = 0; j < 8; j += 2) {
                    const int v0 = cdef_px(c + j, o, pri, sec, pri_shift, sec_shift, w0, w1);
                    const int v1 = cdef_px(c + j + 1, o, pri, sec, pri_shift, sec_shift, w0, w1);
                    o32[j >> 1]  = (uint32_t)v0 | ((uint32_t)v1 << 16);
                }
            } else {
#pragma unroll
                for (int j = 0; j < 4; j++) o32[j] = *reinterpret_cast<const uint32_t *>(c + 2 * j);
            }
            st8<T>(outy + (size_t)gy * a.out.pitch[0] + gx, make_uint4(o32[0], o32[1], o32[2], o32[3]));
        }
    }
    // ---- chroma: work item = one 4-sample row of one 4x4 block, both planes ----
    {
        const int  pri = uvlev << cs, sec = uvsec << cs;
        const int  damp = a.damping + cs - 1;
        const bool on   = uvlev | uvsec;
        const int  pri_shift = pri ? max(0, damp - cdef_msb(pri)) : 0;
        const int  sec_shift = sec ? max(0, damp - cdef_msb(sec)) : 0;
        const int  odd = (pri >> cs) & 1, w0 = odd ? 3 : 4, w1 = odd ? 3 : 2;
        for (int i = tid; i < 512; i += CDEF_THREADS) {
            const int p  = i >> 8;
            const int bx = i & 7, row = (i >> 3) & 3, by = (i >> 5) & 7;
            const int b  = by * 8 + bx;
            const int gy = y0 / 2 + by * 4 + row, gx = x0 / 2 + bx * 4;
            if (gy >= ch / 2 || gx >= cw / 2)
                continue;
            const uint16_t *c = ctile[p] + (2 + by * 4 + row) * CDEF_CP + 2 + bx * 4;
            uint32_t        o32[2];
            if (on && s_active[b]) {
                int o[6];
                cdef_offsets(o, pri ? s_dir[b] : 0, CDEF_CP);
#pragma unroll
                for (int j = 0; j < 4; j += 2) {
                    const int v0 = cdef_px(c + j, o, pri, sec, pri_shift, sec_shift, w0, w1);
                    const int v1 = cdef_px(c + j + 1, o, pri, sec, pri_shift, sec_shift, w0, w1);
                    o32[j >> 1]  = (uint32_t)v0 | ((uint32_t)v1 << 16);
                }
            } else {
                o32[0] = *reinterpret_cast<const uint32_t *>(c);
                o32[1] = *reinterpret_cast<const uint32_t *>(c + 2);
            }
            st4<T>(static_cast<T *>(a.out.p[1 + p]) + (size_t)gy * a.out.pitch[1 + p] + gx, make_uint2(o32[0], o32[1]));
        }
    }
}
#endif // SVTF_AB_KERNELS

// =================================================================================================
// CCSO apply
// =================================================================================================
constexpr int CCSO_THREADS = 256;
constexpr int CCSO_LP      = 82; // luma tile: 64 + 2*8 columns (41 words), 66 rows
constexpr int CCSO_LROWS   = 66;

struct CcsoArgs {
    Planes        luma_src; // only p[0]/pitch[0] used: pre-CDEF luma
    Planes        in, out;  // picture to modify (may alias)
    Dims          d;
    const int8_t *lut;                 // device: [3][2048]
    const uint8_t *blk_flags[3];       // device or NULL (all on)
    uint8_t       enable[3], bo_only[3], thr[3], support[3], edge_clf[3], band_log2[3];
    int           ty0;    // first 64-row tile row of the launch (stripe window)
    int           y_glob; // picture row of local row 0 (block flags are indexed in picture coordinates)
};

// ccso_filter_block_hbd_wo_buf_c — Common/Codec/EbCcso.c:261-294 (classification cal_filter_support :238-259,
// tap positions derive_ccso_sample_pos :204-234) for one sample; lc: centre luma sample in the staged tile.
__device__ __forceinline__ int ccso_px(const uint16_t *lc, const int o0, const int o1, const int8_t *lut, const int v,
                                       const int thr, const int edge_clf, const int bo_only, const int shift_bits,
                                       const int single_band, const int max_val) {
    const int y = lc[0];
    int       c0 = 0, c1 = 0;
    if (!bo_only) {
        const int d0 = lc[o0] - y, d1 = lc[o1] - y;
        if (edge_clf == 0) {
            c0 = d0 > thr ? 2 : (d0 < -thr ? 0 : 1);
            c1 = d1 > thr ? 2 : (d1 < -thr ? 0 : 1);
        } else {
            c0 = d0 < -thr ? 0 : 1;
            c1 = d1 < -thr ? 0 : 1;
        }
    }
    const int band = single_band ? 0 : (y >> shift_bits);
    return min(max(v + lut[(band << 4) + (c0 << 2) + c1], 0), max_val);
}

// ccso_frame — Common/Codec/EbCcso.c:626-678 with ccso_apply_{luma,chroma}_{mb,sb}_filter :297-622.
// The 5-sample replicate border of extend_ccso_border (:185-201) is realised by clamping the staged
// coordinates to the unpadded picture.
template <typename T>
__global__ void __launch_bounds__(CCSO_THREADS) ccso_kernel(const __grid_constant__ CcsoArgs a) {
    __shared__ __align__(16) uint16_t ltile[CCSO_LROWS * CCSO_LP];
    __shared__ int8_t                 s_lut[3][SVTF_CCSO_LUT_SIZE];

    const int tid = threadIdx.x;
    const int x0 = blockIdx.x * 64, y0 = (blockIdx.y + a.ty0) * 64;
    const int W = a.d.width, H = a.d.height;
    const int cw = a.d.mi_cols * 4, ch = a.d.mi_rows * 4;
    const int max_val = (1 << a.d.bit_depth) - 1;

    // block flags: one per 256x256 luma samples (128x128 chroma), row length = ceil(plane_width / blk)
    bool on[3];
#pragma unroll
    for (int p = 0; p < 3; p++) {
        on[p] = a.enable[p];
        if (on[p] && a.blk_flags[p]) {
            const int pwid = W >> (p ? 1 : 0), bl = p ? 7 : 8;
            const int nbx  = (pwid + (1 << bl) - 1) >> bl;
            on[p]          = (x0 >> 8) < nbx && a.blk_flags[p][((y0 + a.y_glob) >> 8) * nbx + (x0 >> 8)] != 0;
        }
    }
    const bool inplace = a.in.p[0] == a.out.p[0];
    if (!(on[0] | on[1] | on[2]) && inplace)
        return;

    if (on[0] | on[1] | on[2]) {
        const T *ls = static_cast<const T *>(a.luma_src.p[0]);
        for (int i = tid; i < CCSO_LROWS * 80; i += CCSO_THREADS) {
            const int r = i / 80, c = i - r * 80;
            const int gy = min(max(y0 - 1 + r, 0), H - 1), gx = min(max(x0 - 8 + c, 0), W - 1);
            ltile[r * CCSO_LP + c] = ls[(size_t)gy * a.luma_src.pitch[0] + gx];
        }
        for (int i = tid; i < 3 * SVTF_CCSO_LUT_SIZE / 4; i += CCSO_THREADS)
            reinterpret_cast<uint32_t *>(&s_lut[0][0])[i] = __ldg(reinterpret_cast<const uint32_t *>(a.lut) + i);
    }
    __syncthreads();

#pragma unroll
    for (int p = 0; p < 3; p++) {
        const int ss   = p ? 1 : 0;
        const int pwid = W >> ss, phei = H >> ss;      // CCSO region (unpadded)
        const int pcw = cw >> ss, pch = ch >> ss;      // coded plane (copy region when out of place)
        const int tw = 64 >> ss;                       // tile width in this plane
        const T  *in  = static_cast<const T *>(a.in.p[p]);
        T        *out = static_cast<T *>(a.out.p[p]);
        if (!on[p] && inplace)
            continue;
        int o0 = 0, o1 = 0;
        switch (a.support[p]) {
        case 0: o0 = -CCSO_LP, o1 = CCSO_LP; break;
        case 1: o0 = -CCSO_LP - 1, o1 = CCSO_LP + 1; break;
        case 2: o0 = -1, o1 = 1; break;
        case 3: o0 = CCSO_LP - 1, o1 = -CCSO_LP + 1; break;
        case 4: o0 = -3, o1 = 3; break;
        default: o0 = -5, o1 = 5; break;
        }
        const int thr = a.thr[p], clf = a.edge_clf[p], bo = a.bo_only[p];
        const int single = a.band_log2[p] == 0, shift_bits = a.d.bit_depth - a.band_log2[p];
        // work item: 4 consecutive samples
        for (int i = tid; i < tw * (tw / 4); i += CCSO_THREADS) {
            const int r = i / (tw / 4), c = (i - r * (tw / 4)) * 4;
            const int gy = (y0 >> ss) + r, gx = (x0 >> ss) + c;
            if (gy >= pch || gx >= pcw)
                continue;
            uint2 v = ld4<T>(in + (size_t)gy * a.in.pitch[p] + gx);
            if (on[p] && gy < phei) {
                uint32_t px[4] = {v.x & 0xffff, v.x >> 16, v.y & 0xffff, v.y >> 16};
#pragma unroll
                for (int j = 0; j < 4; j++)
                    if (gx + j < pwid) {
                        const uint16_t *lc = ltile + (1 + (r << ss)) * CCSO_LP + 8 + ((c + j) << ss);
                        px[j] = (uint32_t)ccso_px(lc, o0, o1, s_lut[p], (int)px[j], thr, clf, bo, shift_bits, single, max_val);
                    }
                v = make_uint2(px[0] | (px[1] << 16), px[2] | (px[3] << 16));
            } else if (inplace)
                continue;
            st4<T>(out + (size_t)gy * a.out.pitch[p] + gx, v);
        }
    }
}

} // namespace svtf
