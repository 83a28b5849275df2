// svtf_search2_kernel.cuh — chroma part and the kernel of the separable CDEF strength search (see svtf_search2.cuh).
#pragma once
#include "svtf_search2.cuh"

namespace svtf {

// ---- chroma: thread = (4x4 block, row); 2 pairs per thread and plane; SSE only (EbEncCdef.c:103-126) --------------
template <typename T, bool BORDER>
__device__ __forceinline__ void s2_chroma(const Search2Args &a, const int p, const uint32_t *cA, const uint32_t *cB, const uint8_t *s_active,
                                          const uint8_t *s_dir, unsigned long long *s_acc, const int x0, const int y0, const int cs, const int tid) {
    const int  b = tid >> 2, q = tid & 3, by = b >> 3, bx = b & 7;
    const bool on = s_active[b];
    if (!__any_sync(0xffffffffu, on))
        return;
    const int trow = 2 + by * 4 + q, tcol = 8 + bx * 4;
    const int damp = a.s.damping + cs - 1;
    uint32_t  x[2], nx[2], g[2];
    int       G2 = 0;
    {
        const uint32_t *c = cA + trow * C2_CPW + (tcol >> 1);
        uint2           v = make_uint2(0, 0);
        if (on)
            v = ld4<T>(static_cast<const T *>(a.s.src.p[1 + p]) + (size_t)(y0 / 2 + by * 4 + q) * a.s.src.pitch[1 + p] + x0 / 2 + bx * 4);
        const uint32_t dd[2] = {v.x, v.y};
#pragma unroll
        for (int j = 0; j < 2; j++) {
            x[j]  = c[j];
            nx[j] = __vadd2(~(x[j] + S2_R0 * 0x10001u), 0x00010001u);
            g[j]  = __vadd2(dd[j], __vadd2(~x[j], 0x00010001u)); // source - x per half (int16)
            const int gl = (int)(short)(g[j] & 0xffff), gh = (int)(short)(g[j] >> 16);
            G2 += gl * gl + gh * gh;
        }
    }
    if (!on)
        G2 = 0;
    const S2Sec sq = s2_sec_setup(cs, damp);
    uint32_t    S[3][2], mnb[2], mxb[2];
#pragma unroll 1
    for (int pass = 0; pass < 2; pass++) {
        const unsigned long long pm = pass ? (a.present_uv >> 4) : (a.present_uv & 15);
        if (!pm)
            continue;
        const int dir = pass ? s_dir[b] : 0;
        s2_row_setup<2, BORDER>(cA, cB, C2_CPW, trow, tcol, dir, sq, x, S, mnb, mxb);
#pragma unroll 1
        for (int pi = pass; pi < (pass ? 16 : 1); pi++) {
            const unsigned m4 = (unsigned)(a.present_uv >> (4 * pi)) & 15u;
            if (!m4)
                continue;
            uint32_t Ap[2];
            s2_row_primary<2>(cA, cB, C2_CPW, trow, tcol, dir, pi << cs, damp, cs, x, Ap);
            int V[4]; // per secondary index: S g^2 - 2 S g e + S e^2 of this thread's 4 samples
#pragma unroll
            for (int si = 0; si < 4; si++) {
                int v = 0;
                if ((m4 >> si) & 1) {
                    const uint32_t sa = si ? S[si - 1][0] : S2_B2 * 0x10001u, sb = si ? S[si - 1][1] : S2_B2 * 0x10001u;
                    const uint32_t ea = s2_err(Ap[0], sa, x[0], nx[0], mnb[0], mxb[0]);
                    const uint32_t eb = s2_err(Ap[1], sb, x[1], nx[1], mnb[1], mxb[1]);
                    const int      e4 = (int)__byte_perm(ea, eb, 0x6420);
                    const int      ge = __dp2a_hi((int)g[1], e4, __dp2a_lo((int)g[0], e4, 0));
                    v                 = G2 + __dp4a(e4, e4, 0) - 2 * ge;
                }
                V[si] = on ? v : 0;
            }
            unsigned long long sse = (unsigned long long)(unsigned)s2_quad_t(V[0], V[1], V[2], V[3], q);
            sse                    = s2_warp_quads(sse);
            if ((tid & 31) < 4 && ((m4 >> q) & 1))
                atomicAdd(&s_acc[pi * 4 + q], sse);
        }
    }
}

#ifdef SVTF_AB_KERNELS // superseded kernel, kept for A/B measurements only (not in the product build)
template <typename T>
__global__ void __launch_bounds__(S2_THREADS, 2) cdef_search2_kernel(const __grid_constant__ Search2Args a) {
    __shared__ __align__(16) uint32_t yA[C2_YROWS * C2_YPW], yB[C2_YROWS * C2_YPW];
    __shared__ __align__(16) uint32_t cA[2][C2_CROWS * C2_CPW], cB[2][C2_CROWS * C2_CPW];
    __shared__ int                    s_cost[64][9];
    __shared__ uint8_t                s_dir[64];
    __shared__ int                    s_var[64];
    __shared__ uint8_t                s_active[64];
    __shared__ unsigned long long     s_acc[3][64]; // indexed by strength byte (pri*4 + secondary index)

    const int tid = threadIdx.x;
    const int fbc = blockIdx.x, fbr = blockIdx.y + a.s.fbr0;
    const int fb  = fbr * gridDim.x + fbc;
    const int cw = a.s.d.mi_cols * 4, ch = a.s.d.mi_rows * 4;
    const int x0 = fbc * 64, y0 = fbr * 64;
    const int cs = a.s.d.bit_depth - 8;

    int act = 0;
    if (tid < 64) { // svt_sb_compute_cdef_list — Encoder/Codec/EbEncCdef.c:238-281
        const int r = (tid >> 3) * 2, c = (tid & 7) * 2;
        if (fbr * 16 + r < a.s.d.mi_rows && fbc * 16 + c < a.s.d.mi_cols) {
            const svtf_mi_t *m = a.s.mi + (fbr * 16 + r) * a.s.mi_stride + fbc * 16 + c;
            act = !(mi_skip(__ldg(m)) & mi_skip(__ldg(m + 1)) & mi_skip(__ldg(m + a.s.mi_stride)) & mi_skip(__ldg(m + a.s.mi_stride + 1)));
        }
        s_active[tid] = (uint8_t)act;
    }
    for (int i = tid; i < 3 * 64; i += S2_THREADS) (&s_acc[0][0])[i] = 0;
    const int any_active = __syncthreads_or(act);
    if (!any_active) { // skip_cdef_seg = 1 (EbCdefProcess.c:203-207); mse rows stay zero
        if (tid == 0)
            a.s.skip[fb] = 1;
        return;
    }
    if (tid == 0)
        a.s.skip[fb] = 0;

    cdef2_stage<T>(yA, C2_YPW, C2_YROWS, 10, static_cast<const T *>(a.s.rec.p[0]), a.s.rec.pitch[0], x0, y0, cw, ch, tid);
    cdef2_stage<T>(cA[0], C2_CPW, C2_CROWS, 6, static_cast<const T *>(a.s.rec.p[1]), a.s.rec.pitch[1], x0 / 2, y0 / 2, cw / 2, ch / 2, tid);
    cdef2_stage<T>(cA[1], C2_CPW, C2_CROWS, 6, static_cast<const T *>(a.s.rec.p[2]), a.s.rec.pitch[2], x0 / 2, y0 / 2, cw / 2, ch / 2, tid);
    __syncthreads();
    cdef2_make_odd(yA, yB, C2_YPW, C2_YROWS, 40, tid);
    cdef2_make_odd(cA[0], cB[0], C2_CPW, C2_CROWS, 24, tid);
    cdef2_make_odd(cA[1], cB[1], C2_CPW, C2_CROWS, 24, tid);
    if (tid < 128) { // direction search: thread = (block, half of the directions)
        const int b = tid & 63, half = tid >> 6;
        if (s_active[b]) {
            const uint32_t *blk = yA + (2 + (b >> 3) * 8) * C2_YPW + 4 + (b & 7) * 4;
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
    if (tid < 64) {
        int best = 0, bd = 0;
        if (s_active[tid]) {
#pragma unroll
            for (int i = 0; i < 8; i++) {
                const int c = s_cost[tid][i];
                if (c > best) { // first maximum wins ties (Common/Codec/EbCdef.c:197-201)
                    best = c;
                    bd   = i;
                }
            }
            const int v = (best - s_cost[tid][(bd + 4) & 7]) >> 10;
            s_var[tid]  = v;
            a.s.dir[fb * 64 + tid] = (uint8_t)bd;
            a.s.var[fb * 64 + tid] = v;
        } else
            s_var[tid] = 0;
        s_dir[tid] = (uint8_t)bd;
    }
    __syncthreads();

    const bool border = fbc == 0 || fbr == 0 || x0 + 66 > cw || y0 + 66 > ch;
    if (border) {
        s2_luma<T, true>(a, yA, yB, s_active, s_dir, s_var, s_acc[0], x0, y0, cs, tid);
        s2_chroma<T, true>(a, 0, cA[0], cB[0], s_active, s_dir, s_acc[1], x0, y0, cs, tid);
        s2_chroma<T, true>(a, 1, cA[1], cB[1], s_active, s_dir, s_acc[2], x0, y0, cs, tid);
    } else {
        s2_luma<T, false>(a, yA, yB, s_active, s_dir, s_var, s_acc[0], x0, y0, cs, tid);
        s2_chroma<T, false>(a, 0, cA[0], cB[0], s_active, s_dir, s_acc[1], x0, y0, cs, tid);
        s2_chroma<T, false>(a, 1, cA[1], cB[1], s_active, s_dir, s_acc[2], x0, y0, cs, tid);
    }
    __syncthreads();
    // mse_seg[0][fb][gi] = (sum >> 2cs) * sub ; mse_seg[1][fb][gi] = U + V (EbCdefProcess.c:297-300),
    // masked chroma strengths -> default_mse_uv * 64 (:86,:258-261)
    if (tid < a.s.n) {
        const int sub = min(a.s.subsampling, 4);
        const int f   = a.s.fs[tid];
        a.s.mse[(size_t)fb * 64 + tid] = (s_acc[0][f] >> (2 * cs)) * (unsigned long long)sub;
        a.s.mse[((size_t)a.s.nfb + fb) * 64 + tid] =
            a.s.fs_uv[tid] == -1 ? 1040400ull * 64 : (s_acc[1][f] >> (2 * cs)) + (s_acc[2][f] >> (2 * cs));
    }
}
#endif // SVTF_AB_KERNELS

} // namespace svtf
