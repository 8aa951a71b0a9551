// Host->device copy rate for the tip-state matrix (100 taxa x 1M patterns, taxon-major): contiguous vs pattern-range chunks (tuning tool).
// nvcc -O2 -o tools/h2d_probe tools/h2d_probe.cu
#include <cstdio>
#include <cuda_runtime.h>
int main() {
    const size_t ntaxa = 100, np = 1000000, pitch = 1000064;
    unsigned char *h, *d;
    cudaMallocHost(&h, ntaxa * np); cudaMalloc(&d, ntaxa * pitch);
    cudaStream_t s; cudaStreamCreate(&s);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int mode = 0; mode < 6; mode++) {
        int chunks = mode == 0 ? 1 : (mode == 1 ? 1 : (mode == 2 ? 4 : (mode == 3 ? 8 : (mode == 4 ? 16 : 8))));
        float best = 1e9;
        for (int rep = 0; rep < 4; rep++) {
            cudaEventRecord(e0, s);
            if (mode == 0) cudaMemcpyAsync(d, h, ntaxa * np, cudaMemcpyHostToDevice, s);
            else if (mode == 5) {
                size_t w = np / chunks;
                for (int c = 0; c < chunks; c++)
                    for (size_t t = 0; t < ntaxa; t++) cudaMemcpyAsync(d + t * pitch + c * w, h + t * np + c * w, w, cudaMemcpyHostToDevice, s);
            } else {
                size_t w = np / chunks;
                for (int c = 0; c < chunks; c++) cudaMemcpy2DAsync(d + c * w, pitch, h + c * w, np, w, ntaxa, cudaMemcpyHostToDevice, s);
            }
            cudaEventRecord(e1, s); cudaStreamSynchronize(s);
            float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
        }
        const char *names[] = {"1-D contiguous", "2-D x1", "2-D x4", "2-D x8", "2-D x16", "1-D rows x8 chunks"};
        printf("%-22s %.3f ms %.1f GB/s\n", names[mode], best, ntaxa * np / best / 1e6);
    }
    return 0;
}
