/*
 * fp64_peak.cu — measures the FP64 roofline denominators of the box (no published B200 number is trusted):
 *   DFMA   vector FP64 FMA throughput (TFLOP/s), ILP 8 per thread
 *   DMMA   mma.sync.aligned.m8n8k4.f64 and m16n8k8.f64 throughput (TFLOP/s)
 *   copy   device-to-device copy GB/s with 256-bit accesses (cross-check of MEASURED_PEAKS.json hbm_gbs)
 *   write  pure streaming write GB/s (the fused traversal kernel is write-dominated)
 * Prints one JSON line.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/fp64_peak tools/fp64_peak.cu
 */
#include <cstdio>
#include <cuda_runtime.h>

__global__ void dfma_kernel(double *out, int iters, double a, double b) {
    double x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
    for (int i = 0; i < iters; i++) {
        x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
        x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
}

__global__ void dmma884_kernel(double *out, int iters) {
    double a = threadIdx.x * 1e-3, b = threadIdx.x * 2e-3;
    double c0[4] = {0, 0, 0, 0}, c1[4] = {0, 0, 0, 0};
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int u = 0; u < 4; u++)
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                         : "+d"(c0[u]), "+d"(c1[u]) : "d"(a), "d"(b));
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = c0[0] + c1[0] + c0[1] + c1[1] + c0[2] + c1[2] + c0[3] + c1[3];
}

__global__ void dmma1688_kernel(double *out, int iters) {
    double a[4], b[2];
    for (int i = 0; i < 4; i++) a[i] = threadIdx.x * 1e-3 + i;
    b[0] = threadIdx.x * 2e-3; b[1] = b[0] + 1;
    double c[4][4];
    for (int u = 0; u < 4; u++) for (int i = 0; i < 4; i++) c[u][i] = 0;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int u = 0; u < 4; u++)
            asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+d"(c[u][0]), "+d"(c[u][1]), "+d"(c[u][2]), "+d"(c[u][3])
                         : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(b[0]), "d"(b[1]));
    }
    double s = 0;
    for (int u = 0; u < 4; u++) for (int i = 0; i < 4; i++) s += c[u][i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

struct __align__(32) d4 { double x, y, z, w; };
__global__ void copy_kernel(const d4 *__restrict__ in, d4 *__restrict__ out, size_t n) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        d4 r;
        asm volatile("ld.global.nc.L1::no_allocate.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(r.x), "=d"(r.y), "=d"(r.z), "=d"(r.w) : "l"(in + i));
        asm volatile("st.global.L1::no_allocate.v4.f64 [%0], {%1,%2,%3,%4};" :: "l"(out + i), "d"(r.x), "d"(r.y), "d"(r.z), "d"(r.w) : "memory");
    }
}
__global__ void write_kernel(d4 *__restrict__ out, size_t n, double v) {
    d4 r = {v, v + 1, v + 2, v + 3};
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        asm volatile("st.global.L1::no_allocate.v4.f64 [%0], {%1,%2,%3,%4};" :: "l"(out + i), "d"(r.x), "d"(r.y), "d"(r.z), "d"(r.w) : "memory");
}
__global__ void read_kernel(const d4 *__restrict__ in, double *out, size_t n) {
    double s = 0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        d4 r;
        asm volatile("ld.global.nc.L1::no_allocate.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(r.x), "=d"(r.y), "=d"(r.z), "=d"(r.w) : "l"(in + i));
        s += r.x + r.y + r.z + r.w;
    }
    if (s == 1.2345e-300) out[0] = s;
}

template <typename F> static float time_best(F f, int reps) {
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    float best = 1e30f;
    for (int r = 0; r < reps + 2; r++) {
        cudaEventRecord(a); f(); cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b);
        if (r >= 2 && ms < best) best = ms;
    }
    return best;
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int sms = p.multiProcessorCount;
    double *out; cudaMalloc(&out, sizeof(double) * sms * 16 * 256);
    int iters = 1 << 14;
    float ms = time_best([&] { dfma_kernel<<<sms * 8, 256>>>(out, iters, 1.0000001, 1e-9); }, 5);
    double dfma = 2.0 * 8 * iters * (double)sms * 8 * 256 / (ms * 1e-3) / 1e12;
    ms = time_best([&] { dmma884_kernel<<<sms * 8, 256>>>(out, iters); }, 5);
    double dmma884 = 2.0 * 8 * 8 * 4 * 4.0 * iters * (double)sms * 8 * 8 / (ms * 1e-3) / 1e12;
    ms = time_best([&] { dmma1688_kernel<<<sms * 8, 256>>>(out, iters); }, 5);
    double dmma1688 = 2.0 * 16 * 8 * 8 * 4.0 * iters * (double)sms * 8 * 8 / (ms * 1e-3) / 1e12;
    size_t n = (size_t)1 << 25;   /* 1 GiB of d4 */
    d4 *a, *b; cudaMalloc(&a, n * 32); cudaMalloc(&b, n * 32); cudaMemset(a, 0, n * 32);
    ms = time_best([&] { copy_kernel<<<sms * 16, 256>>>(a, b, n); }, 5);
    double copy = 2.0 * n * 32 / (ms * 1e-3) / 1e9;
    ms = time_best([&] { write_kernel<<<sms * 16, 256>>>(b, n, 1.0); }, 5);
    double wr = 1.0 * n * 32 / (ms * 1e-3) / 1e9;
    ms = time_best([&] { read_kernel<<<sms * 16, 256>>>(a, out, n); }, 5);
    double rd = 1.0 * n * 32 / (ms * 1e-3) / 1e9;
    ms = time_best([&] { cudaMemcpyAsync(b, a, n * 32, cudaMemcpyDeviceToDevice); }, 5);
    double memcpy_gbs = 2.0 * n * 32 / (ms * 1e-3) / 1e9;
    printf("{\"gpu\": \"%s\", \"sms\": %d, \"dfma_tflops\": %.2f, \"dmma_m8n8k4_tflops\": %.2f, \"dmma_m16n8k8_tflops\": %.2f, "
           "\"copy_gbs\": %.1f, \"write_gbs\": %.1f, \"read_gbs\": %.1f, \"cudamemcpy_d2d_gbs\": %.1f, \"err\": \"%s\"}\n",
           p.name, sms, dfma, dmma884, dmma1688, copy, wr, rd, memcpy_gbs, cudaGetErrorString(cudaGetLastError()));
    return 0;
}
