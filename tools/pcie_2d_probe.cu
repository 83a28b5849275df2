// pcie_2d_probe.cu — linear vs row-wise (cudaMemcpy2DAsync) PCIe copies of one 4K 10-bit luma plane's worth of rows, from cudaHostAlloc'ed
// and from malloc'ed + cudaHostRegister'ed host memory.  nvcc -O2 -o tools/_bin/pcie_2d_probe tools/pcie_2d_probe.cu
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)
int main() {
    const size_t w = 3840 * 2, h = 2160 * 3 / 2, spitch = (3840 + 2 * 160) * 2; // bytes per row, rows (luma + chroma as rows), host stride
    void *hp, *hr_raw, *d, *d2;
    CK(cudaHostAlloc(&hp, spitch * h, cudaHostAllocDefault));
    hr_raw = malloc(spitch * h + 4096);
    memset(hr_raw, 1, spitch * h + 4096);
    CK(cudaHostRegister(hr_raw, spitch * h + 4096, cudaHostRegisterDefault));
    CK(cudaMalloc(&d, spitch * h));
    CK(cudaMalloc(&d2, spitch * h));
    cudaStream_t s;
    CK(cudaStreamCreate(&s));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    const char *names[2] = {"cudaHostAlloc", "malloc+cudaHostRegister"};
    void *hosts[2] = {hp, hr_raw};
    for (int m = 0; m < 2; m++)
        for (int dir = 0; dir < 2; dir++)
            for (int mode = 0; mode < 3; mode++) { // 0 linear payload, 1 2D rows, 2 linear whole stride
                float best = 1e9f;
                for (int it = 0; it < 6; it++) {
                    CK(cudaEventRecord(e0, s));
                    for (int r = 0; r < 4; r++) {
                        if (mode == 0)
                            CK(cudaMemcpyAsync(dir ? hosts[m] : d, dir ? d : hosts[m], w * h, dir ? cudaMemcpyDeviceToHost : cudaMemcpyHostToDevice, s));
                        else if (mode == 1)
                            CK(cudaMemcpy2DAsync(dir ? hosts[m] : d, dir ? spitch : w, dir ? d : hosts[m], dir ? w : spitch, w, h,
                                                 dir ? cudaMemcpyDeviceToHost : cudaMemcpyHostToDevice, s));
                        else
                            CK(cudaMemcpyAsync(dir ? hosts[m] : d, dir ? d : hosts[m], spitch * h, dir ? cudaMemcpyDeviceToHost : cudaMemcpyHostToDevice, s));
                    }
                    CK(cudaEventRecord(e1, s));
                    CK(cudaEventSynchronize(e1));
                    float ms;
                    CK(cudaEventElapsedTime(&ms, e0, e1));
                    if (ms < best) best = ms;
                }
                printf("%-24s %s %-22s %.3f ms per picture  %.1f GB/s payload\n", names[m], dir ? "D2H" : "H2D",
                       mode == 0 ? "linear (payload only)" : (mode == 1 ? "2D rows 7680 B" : "linear (whole stride)"), best / 4, w * h * 4 / best / 1e6);
            }
    // device-side re-pitch cost
    CK(cudaEventRecord(e0, s));
    for (int r = 0; r < 20; r++) CK(cudaMemcpy2DAsync(d2, w, d, spitch, w, h, cudaMemcpyDeviceToDevice, s));
    CK(cudaEventRecord(e1, s));
    CK(cudaEventSynchronize(e1));
    float ms;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    printf("device-to-device 2D re-pitch: %.1f us per picture\n", ms / 20 * 1e3);
    return 0;
}
