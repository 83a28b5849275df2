// Stand-alone probe: which 2-D TMA box / element configurations load correctly on this GPU (run on the box: nvcc + run).
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>
__global__ void k(const __grid_constant__ CUtensorMap tm, int x, int y, uint32_t bytes, uint8_t *out, int dst_off) {
    extern __shared__ __align__(128) uint8_t sm[];
    __shared__ __align__(8) uint64_t bar;
    if (threadIdx.x == 0) {
        uint32_t b = (uint32_t)__cvta_generic_to_shared(&bar), d = (uint32_t)__cvta_generic_to_shared(sm + dst_off);
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(b) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(bytes) : "memory");
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(d),
                     "l"(reinterpret_cast<uint64_t>(&tm)), "r"(x), "r"(y), "r"(b)
                     : "memory");
        asm volatile("{\n.reg .pred p;\nW:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n@p bra D;\nbra W;\nD:\n}" ::"r"(b) : "memory");
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < bytes; i += blockDim.x) out[i] = sm[dst_off + i];
}
int main(int argc, char **argv) {
    const int only = argc > 1 ? atoi(argv[1]) : -1; int ci = -1;
    typedef CUresult (*fn_t)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *, const cuuint32_t *, const cuuint32_t *,
                             CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    void *p = nullptr; cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
    fn_t enc = (fn_t)p;
    const int W = 384, H = 200, pitch = 384;
    std::vector<uint8_t> h((size_t)pitch * H);
    for (size_t i = 0; i < h.size(); i++) h[i] = (uint8_t)(i * 7 + (i >> 8));
    uint8_t *d, *out; cudaMalloc(&d, h.size()); cudaMalloc(&out, 1 << 16);
    cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice);
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    struct Cfg { int esz, bw, bh, x, y, off; } cfgs[] = {
        {2, 80, 68, -8, -2, 0}, {1, 80, 68, -8, -2, 0}, {1, 96, 68, -8, -2, 0}, {1, 128, 68, -8, -2, 0}, {2, 40, 68, -4, -2, 0}, {2, 48, 68, -4, -2, 0},
        {1, 80, 68, 56, 62, 0}, {1, 64, 32, 0, 0, 0}, {1, 80, 32, 0, 0, 0}, {1, 80, 68, 0, 0, 0}, {1, 80, 68, -8, 0, 0}, {1, 80, 68, 0, -2, 0}, {2, 24, 36, -4, -2, 6528},
        {1, 48, 36, -8, -2, 5504}, {1, 256, 16, 0, 0, 0}, {1, 16, 16, -8, -2, 0}};
    for (auto &c : cfgs) {
        if (++ci != only && only >= 0) continue;
        CUtensorMap tm;
        cuuint64_t dims[2] = {(cuuint64_t)(W / c.esz), (cuuint64_t)H}, st[1] = {(cuuint64_t)pitch};
        cuuint32_t box[2] = {(cuuint32_t)c.bw, (cuuint32_t)c.bh}, es[2] = {1, 1};
        CUresult r = enc(&tm, c.esz == 1 ? CU_TENSOR_MAP_DATA_TYPE_UINT8 : CU_TENSOR_MAP_DATA_TYPE_UINT16, 2, d, dims, st, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                         CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        const uint32_t bytes = c.bw * c.bh * c.esz;
        cudaMemset(out, 0xEE, 1 << 16);
        k<<<1, 128, 64 * 1024>>>(tm, c.x, c.y, bytes, out, c.off);
        cudaError_t e = cudaDeviceSynchronize();
        int bad = -1;
        if (e == cudaSuccess) {
            std::vector<uint8_t> o(bytes);
            cudaMemcpy(o.data(), out, bytes, cudaMemcpyDeviceToHost);
            bad = 0;
            const int rowb = c.bw * c.esz;
            for (int r2 = 0; r2 < c.bh; r2++)
                for (int b = 0; b < rowb; b++) {
                    const int gy = c.y + r2, gxb = c.x * c.esz + b;
                    const uint8_t exp = (gy >= 0 && gy < H && gxb >= 0 && gxb < W) ? h[(size_t)gy * pitch + gxb] : 0;
                    bad += o[(size_t)r2 * rowb + b] != exp;
                }
        }
        printf("esz %d box %3dx%2d at (%3d,%3d) smem+%5d rowbytes %3d: encode %d, run %s, mismatches %d\n", c.esz, c.bw, c.bh, c.x, c.y, c.off, c.bw * c.esz, (int)r,
               cudaGetErrorString(e), bad);
        if (e != cudaSuccess) { printf("(context lost: stopping)\n"); return 0; }
    }
    return 0;
}
