// Ceiling of a WRITE-ONLY HBM stream on this GPU, next to the copy figure MEASURED_PEAKS.json holds: the DNA whole-traversal kernel
// writes 12.7 GB and reads 0.1 GB per launch, so the copy bandwidth (half reads, half writes) is not obviously its roofline.
// Kernels: cudaMemsetAsync; 256-bit streaming stores (st.global.L1::no_allocate.v4.f64, the traversal kernel's store) from a
// persistent grid (148 x k CTAs) and from one CTA per 64 KB; 128-bit stores; and a 256-bit load + store copy for comparison.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/write_peak tools/write_peak.cu      run: tools/write_peak [GB=8]
#include <cuda_runtime.h>
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t rc_ = (x); if (rc_ != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(rc_)); return 1; } } while (0)

__device__ __forceinline__ void st256(double *p, double a) {
    asm volatile("st.global.L1::no_allocate.v4.f64 [%0], {%1,%1,%1,%1};" :: "l"(p), "d"(a) : "memory");
}
__device__ __forceinline__ void st128(double *p, double a) {
    asm volatile("st.global.L1::no_allocate.v2.f64 [%0], {%1,%1};" :: "l"(p), "d"(a) : "memory");
}

__global__ void __launch_bounds__(384) write256_kernel(double *dst, size_t n4, double val) {       // n4: number of 32-byte units
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) st256(dst + 4 * i, val);
}
__global__ void __launch_bounds__(384) write128_kernel(double *dst, size_t n2, double val) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += stride) st128(dst + 2 * i, val);
}
// each CTA owns contiguous 64 KB chunks (the traversal kernel's pattern: a CTA writes whole tiles of every node's vector)
__global__ void __launch_bounds__(384) write256_chunk_kernel(double *dst, size_t nchunk, double val) {
    for (size_t c = blockIdx.x; c < nchunk; c += gridDim.x) {
        double *base = dst + c * 8192;
        for (int i = threadIdx.x; i < 2048; i += blockDim.x) st256(base + 4 * i, val);
    }
}
__global__ void __launch_bounds__(384) copy256_kernel(double *dst, const double *src, size_t n4) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        double a, b, c, d;
        asm volatile("ld.global.L1::no_allocate.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(a), "=d"(b), "=d"(c), "=d"(d) : "l"(src + 4 * i) : "memory");
        asm volatile("st.global.L1::no_allocate.v4.f64 [%0], {%1,%2,%3,%4};" :: "l"(dst + 4 * i), "d"(a), "d"(b), "d"(c), "d"(d) : "memory");
    }
}

template <typename F>
static int timed(const char *name, double bytes, F launch) {
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    std::vector<float> ms;
    for (int r = 0; r < 7; r++) {
        CK(cudaEventRecord(e0));
        launch();
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        CK(cudaGetLastError());
        float t; CK(cudaEventElapsedTime(&t, e0, e1));
        if (r >= 2) ms.push_back(t);
    }
    std::sort(ms.begin(), ms.end());
    printf("{\"kernel\": \"%s\", \"gb\": %.3f, \"best_ms\": %.4f, \"median_ms\": %.4f, \"best_gbps\": %.1f, \"median_gbps\": %.1f}\n", name, bytes / 1e9, ms.front(),
           ms[ms.size() / 2], bytes / 1e6 / ms.front(), bytes / 1e6 / ms[ms.size() / 2]);
    return 0;
}

int main(int argc, char **argv) {
    const double gb = argc > 1 ? atof(argv[1]) : 8.0;
    const size_t bytes = (size_t)(gb * 1e9) / 65536 * 65536;
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    double *a = NULL, *b = NULL;
    CK(cudaMalloc(&a, bytes)); CK(cudaMalloc(&b, bytes));
    CK(cudaMemset(a, 0, bytes)); CK(cudaMemset(b, 0, bytes));
    int rc = 0;
    rc |= timed("cudaMemsetAsync", (double)bytes, [&] { cudaMemsetAsync(a, 1, bytes); });
    for (int k : {1, 2, 4, 8})
        rc |= timed(k == 1 ? "st.v4.f64 grid-stride, 148x1 CTAs of 384" : k == 2 ? "st.v4.f64 grid-stride, 148x2" : k == 4 ? "st.v4.f64 grid-stride, 148x4" : "st.v4.f64 grid-stride, 148x8",
                    (double)bytes, [&] { write256_kernel<<<sms * k, 384>>>(a, bytes / 32, 1.5); });
    rc |= timed("st.v4.f64, 64 KB chunk per CTA, 148x1", (double)bytes, [&] { write256_chunk_kernel<<<sms, 384>>>(a, bytes / 65536, 2.5); });
    rc |= timed("st.v4.f64, 64 KB chunk per CTA, 148x4", (double)bytes, [&] { write256_chunk_kernel<<<sms * 4, 384>>>(a, bytes / 65536, 2.5); });
    rc |= timed("st.v2.f64 grid-stride, 148x4", (double)bytes, [&] { write128_kernel<<<sms * 4, 384>>>(a, bytes / 16, 3.5); });
    rc |= timed("copy ld.v4 + st.v4 (bytes read + written), 148x4", 2.0 * bytes, [&] { copy256_kernel<<<sms * 4, 384>>>(b, a, bytes / 32); });
    rc |= timed("cudaMemcpyAsync d2d (bytes read + written)", 2.0 * bytes, [&] { cudaMemcpyAsync(b, a, bytes, cudaMemcpyDeviceToDevice); });
    cudaFree(a); cudaFree(b);
    return rc;
}
