/*
 * lkcuda_aux.cu — the consumers of the likelihood kernels' per-pattern outputs that SURVEY 8(f) ranks next, on the device:
 *
 *   (f.2) UFBoot RELL: IQTree::saveCurrentTree (iqtree.cpp:2676-2786) computes, for every candidate tree, nsamples (1000) float
 *         dot products of the scale-corrected pattern log-likelihoods (computePatternLikelihood, phylotree.cpp:1213-1259) against
 *         the bootstrap pattern frequencies boot_samples[s][ptn] (dotProductSIMD, phylokernel.h:55-61) - a GEMV whose matrix
 *         (nsamples x patterns) is constant for the whole search.  Here the matrix lives in HBM, stored LOSSLESSLY as uint8 /
 *         uint16 counts when its entries allow it (they are multinomial counts, iqtree.cpp:300-340), so the HBM-bound pass moves
 *         1/4 or 1/2 of the reference's float bytes; products are formed in fp32 like the reference, summed in short fp32 chains
 *         and accumulated in fp64.
 *   (f.1) _pattern_lh_cat consumers: PhyloTree::computePatternLhCat's collapse to mixture classes / rate categories
 *         (phylotree.cpp:1108-1160), the EM step shared by ModelMixture::optimizeWeights (model/modelmixture.cpp:1356-1416) and
 *         RateFree::optimizeWithEM (model/ratefree.cpp:477-560), and RateGamma::computePatternRates (model/rategamma.cpp:235-257).
 *         The reference runs them as host passes over P*C*M doubles; here only ncol sums (or P rates) come back.
 */
#include "lkcuda_traverse.cuh"

using namespace lkc;

#define AUX_THREADS 256

/* ------------------------------------------------------------------ RELL ------------------------------------------------------------------ */

/* x[p] = (float)(_pattern_lh[p] + (max(scale_a[p],0) + max(scale_b[p],0)) * LOG_SCALING_THRESHOLD), zero padded to the row stride */
__global__ void __launch_bounds__(AUX_THREADS) rell_x_kernel(DevView v, int vec_a, int vec_b, float *x, long long stride) {
    const int16_t *sa = vec_a >= 0 ? vec_sc(v, vec_a) : NULL, *sb = vec_b >= 0 ? vec_sc(v, vec_b) : NULL;
    for (long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x; p < stride; p += (long long)gridDim.x * blockDim.x) {
        float r = 0.f;
        if (p < v.P) {
            int s = 0;
            if (sa) { int t = sa[p]; s += t > 0 ? t : 0; }
            if (sb) { int t = sb[p]; s += t > 0 ? t : 0; }
            r = (float)(s ? v.pattern_lh[p] + (double)s * LK_LOG_THR : v.pattern_lh[p]);
        }
        x[p] = r;
    }
}

/* 16 entries of a row as floats */
__device__ __forceinline__ void rell_unpack(const uint8_t *row, long long p, float (&f)[16]) {
    const uint4 w = *reinterpret_cast<const uint4 *>(row + p);
    const unsigned int q[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
    for (int i = 0; i < 4; i++) {
#pragma unroll
        for (int b = 0; b < 4; b++)     /* float(byte) without a conversion instruction: 2^23 + byte is exact, then subtract 2^23 */
            f[4 * i + b] = __uint_as_float(__byte_perm(q[i], 0x4B000000u, 0x7650 + b)) - 8388608.0f;
    }
}
__device__ __forceinline__ void rell_unpack(const uint16_t *row, long long p, float (&f)[16]) {
    const uint4 w0 = *reinterpret_cast<const uint4 *>(row + p), w1 = *reinterpret_cast<const uint4 *>(row + p + 8);
    const unsigned int q[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
    for (int i = 0; i < 8; i++) {
        f[2 * i] = __uint_as_float(__byte_perm(q[i], 0x4B000000u, 0x7610)) - 8388608.0f;
        f[2 * i + 1] = __uint_as_float(__byte_perm(q[i], 0x4B000000u, 0x7632)) - 8388608.0f;
    }
}
__device__ __forceinline__ void rell_unpack(const float *row, long long p, float (&f)[16]) {
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const float4 w = *reinterpret_cast<const float4 *>(row + p + 4 * i);
        f[4 * i] = w.x; f[4 * i + 1] = w.y; f[4 * i + 2] = w.z; f[4 * i + 3] = w.w;
    }
}

/* grid = (pattern chunks, sample groups).  A block owns RELL_SPB consecutive samples and one chunk of patterns: every thread keeps
   its 16 x values in registers and streams the RELL_SPB rows past them, so x is read once per sample GROUP (from L2) while the matrix
   - the only HBM traffic - is read exactly once, 16 bytes or more per lane and request.  Block sums go to part[chunk][sample]; the
   last block of a sample group (ticket counter) adds the chunks in order: deterministic, no atomics on doubles. */
#define RELL_SPB 8
template <typename T>
__global__ void __launch_bounds__(AUX_THREADS) rell_kernel(const T *__restrict__ mat, long long stride, int nsamples, const float *__restrict__ x,
                                                           double *part, unsigned int *tickets, double *out) {
    __shared__ double sh[RELL_SPB][AUX_THREADS / 32];
    __shared__ bool is_last;
    const int s0 = blockIdx.y * RELL_SPB;
    const long long chunk = (long long)AUX_THREADS * 16;
    double acc[RELL_SPB];
#pragma unroll
    for (int s = 0; s < RELL_SPB; s++) acc[s] = 0.0;
    for (long long base = (long long)blockIdx.x * chunk; base < stride; base += (long long)gridDim.x * chunk) {
        const long long p = base + (long long)threadIdx.x * 16;
        if (p >= stride) continue;          /* stride is a multiple of 16 */
        float xv[16];
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const float4 w = *reinterpret_cast<const float4 *>(x + p + 4 * i);
            xv[4 * i] = w.x; xv[4 * i + 1] = w.y; xv[4 * i + 2] = w.z; xv[4 * i + 3] = w.w;
        }
        if constexpr (sizeof(T) == 1) {
            /* all RELL_SPB rows requested before the first is used: 8 x 16 bytes in flight per lane */
            uint4 raw[RELL_SPB];
#pragma unroll
            for (int s = 0; s < RELL_SPB; s++) {
                const int ss = s0 + s < nsamples ? s0 + s : nsamples - 1;
                const uint8_t *src = reinterpret_cast<const uint8_t *>(mat) + (long long)ss * stride + p;
                asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                             : "=r"(raw[s].x), "=r"(raw[s].y), "=r"(raw[s].z), "=r"(raw[s].w) : "l"(src));
            }
#pragma unroll
            for (int s = 0; s < RELL_SPB; s++) {
                const unsigned int q[4] = {raw[s].x, raw[s].y, raw[s].z, raw[s].w};
                float t = 0.f;
#pragma unroll
                for (int i = 0; i < 4; i++) {
#pragma unroll
                    for (int b = 0; b < 4; b++)
                        t = fmaf(__uint_as_float(__byte_perm(q[i], 0x4B000000u, 0x7650 + b)) - 8388608.0f, xv[4 * i + b], t);
                }
                if (s0 + s < nsamples) acc[s] += (double)t;
            }
        } else {
#pragma unroll
            for (int s = 0; s < RELL_SPB; s++) {
                if (s0 + s >= nsamples) break;
                float f[16];
                rell_unpack(mat + (long long)(s0 + s) * stride, p, f);
                float t = 0.f;
#pragma unroll
                for (int i = 0; i < 16; i++) t = fmaf(f[i], xv[i], t);
                acc[s] += (double)t;
            }
        }
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int s = 0; s < RELL_SPB; s++) {
        double a = acc[s];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) a += __shfl_down_sync(0xffffffffu, a, o);
        if (lane == 0) sh[s][warp] = a;
    }
    __syncthreads();
    if (threadIdx.x < RELL_SPB) {
        double a = 0.0;
#pragma unroll
        for (int w = 0; w < AUX_THREADS / 32; w++) a += sh[threadIdx.x][w];
        part[((long long)blockIdx.x * gridDim.y + blockIdx.y) * RELL_SPB + threadIdx.x] = a;
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) is_last = atomicAdd(&tickets[blockIdx.y], 1u) == gridDim.x - 1;
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    if (threadIdx.x < RELL_SPB && s0 + threadIdx.x < nsamples) {
        double a = 0.0;
        for (unsigned int c = 0; c < gridDim.x; c++) a += __ldcg(&part[((long long)c * gridDim.y + blockIdx.y) * RELL_SPB + threadIdx.x]);
        out[s0 + threadIdx.x] = a;
    }
    if (threadIdx.x == 0) tickets[blockIdx.y] = 0u;
}

/* ------------------------------------------------------------------ _pattern_lh_cat consumers ------------------------------------------------------------------ */

/* w[p][j] = sum of _pattern_lh_cat[p][m*ncat + c] over c (mode 1: per mixture class, j = m) or over m (mode 2: per rate category, j = c);
   mode 0: copy (phylotree.cpp:1133-1158) */
__global__ void __launch_bounds__(AUX_THREADS) lhcat_collapse_kernel(DevView v, int mode, int ncat, int nmix, double *w) {
    const int ncol = mode == 1 ? nmix : (mode == 2 ? ncat : ncat * nmix);
    const long long total = v.P * ncol;
    for (long long u = (long long)blockIdx.x * blockDim.x + threadIdx.x; u < total; u += (long long)gridDim.x * blockDim.x) {
        const long long p = u / ncol; const int j = (int)(u - p * ncol);
        const double *src = v.pattern_lh_cat + p * v.K;
        double s;
        if (mode == 1) { s = src[j * ncat]; for (int c = 1; c < ncat; c++) s += src[j * ncat + c]; }
        else if (mode == 2) { s = src[j]; for (int m = 1; m < nmix; m++) s += src[m * ncat + j]; }
        else s = src[j];
        w[u] = s;
    }
}

/* One EM pass over the patterns (modelmixture.cpp:1369-1388, ratefree.cpp:499-520): w[p][c] *= ratio[c] (when given), then
   sums[c] = sum_p w[p][c] * ptn_freq[p] / (ptn_invar[p] + sum_c w[p][c]).
   Tiles of AUX_THREADS patterns are staged in shared memory; thread (seg, c) adds its segment of the tile in pattern order, thread c
   adds the segments in order: the result does not depend on scheduling. */
#define EM_MAXCOL 96      /* a tile of 256 patterns x (ncol + 1) doubles must fit in shared memory; C60 and C20+G4 (80) do */
__global__ void __launch_bounds__(AUX_THREADS) lhcat_em_kernel(DevView v, int ncol, double *w, const double *__restrict__ ratio, int use_ratio, int include_invar,
                                                               double *part, unsigned int *ticket, double *out) {
    extern __shared__ __align__(16) double shm[];
    double *tile = shm;                                   /* [AUX_THREADS][ncol + 1] */
    double *scale = tile + (size_t)AUX_THREADS * (ncol + 1);  /* [AUX_THREADS]: ptn_freq / lk_ptn */
    double *seg = scale + AUX_THREADS;                    /* [nseg][ncol] */
    double *rat = seg + AUX_THREADS;                      /* [ncol] */
    __shared__ bool is_last;
    const int nseg = AUX_THREADS / ncol > 0 ? AUX_THREADS / ncol : 1;
    const int seglen = (AUX_THREADS + nseg - 1) / nseg;
    const int my_c = threadIdx.x % ncol, my_seg = threadIdx.x / ncol;
    for (int c = threadIdx.x; c < ncol; c += blockDim.x) rat[c] = use_ratio ? ratio[c] : 1.0;
    __syncthreads();
    double total = 0.0;          /* thread c < ncol: running sum of column c */
    for (long long base = (long long)blockIdx.x * AUX_THREADS; base < v.P; base += (long long)gridDim.x * AUX_THREADS) {
        const int np = (int)(v.P - base < AUX_THREADS ? v.P - base : AUX_THREADS);
        for (int i = threadIdx.x; i < np * ncol; i += blockDim.x) {      /* coalesced: the tile is contiguous in w */
            const int p = i / ncol, c = i - p * ncol;
            double x = w[base * ncol + i];
            if (use_ratio) { x *= rat[c]; w[base * ncol + i] = x; }
            tile[p * (ncol + 1) + c] = x;
        }
        __syncthreads();
        if (threadIdx.x < np) {
            double lk = include_invar ? v.invar[base + threadIdx.x] : 0.0;
            const double *row = tile + threadIdx.x * (ncol + 1);
            for (int c = 0; c < ncol; c++) lk += row[c];
            scale[threadIdx.x] = v.freq[base + threadIdx.x] / lk;
        }
        __syncthreads();
        if (my_seg < nseg) {
            double s = 0.0;
            const int b = my_seg * seglen, e = b + seglen < np ? b + seglen : np;
            for (int p = b; p < e; p++) s += tile[p * (ncol + 1) + my_c] * scale[p];
            seg[my_seg * ncol + my_c] = s;
        }
        __syncthreads();
        if (threadIdx.x < ncol) for (int g = 0; g < nseg; g++) total += seg[g * ncol + threadIdx.x];
        __syncthreads();
    }
    if (threadIdx.x < ncol) part[(long long)blockIdx.x * ncol + threadIdx.x] = total;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) is_last = atomicAdd(ticket, 1u) == gridDim.x - 1;
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    for (int c = threadIdx.x; c < ncol; c += blockDim.x) {
        double a = 0.0;
        for (unsigned int b = 0; b < gridDim.x; b++) a += __ldcg(&part[(long long)b * ncol + c]);
        out[c] = a;
    }
    if (threadIdx.x == 0) *ticket = 0u;
}

/* RateGamma::computePatternRates (rategamma.cpp:235-257): posterior mean rate and most likely category of every pattern.
   The reference breaks exact ties between categories with random_double(); here the first maximum wins. */
__global__ void __launch_bounds__(AUX_THREADS) lhcat_rates_kernel(DevView v, int ncol, const double *__restrict__ w, const double *__restrict__ rates,
                                                                  int invar_class, double *ptn_rates, int32_t *ptn_cat) {
    for (long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x; p < v.P; p += (long long)gridDim.x * blockDim.x) {
        const double *lh = w + p * ncol;
        /* RateGammaInvar (rategammainvar.cpp:193-221): the invariant class is category 0 with likelihood ptn_invar and rate 0 */
        double sum_rate = 0.0, sum_lh = invar_class ? v.invar[p] : 0.0, bestv = invar_class ? v.invar[p] : lh[0];
        int best = 0;
        for (int c = 0; c < ncol; c++) {
            const double x = lh[c];
            sum_rate += rates[c] * x; sum_lh += x;
            if (x > bestv) { bestv = x; best = c + invar_class; }
        }
        ptn_rates[p] = sum_rate / sum_lh;
        ptn_cat[p] = best;
    }
}

/* ------------------------------------------------------------------ host side ------------------------------------------------------------------ */

struct lkcuda_aux {
    /* RELL */
    void *d_boot; int boot_kind; int nsamples; long long boot_stride; size_t boot_bytes;
    float *d_x; double *d_part; unsigned int *d_tickets; double *d_rell; double *h_rell; size_t part_cap; int groups_cap; int rell_gx;
    /* _pattern_lh_cat consumers */
    double *d_w; int ncol; double *d_em_part; unsigned int *d_em_ticket; double *d_em_io; double *h_em_io;
    double *d_rates; int32_t *d_cat;
};

static lkcuda_aux *aux_of(lkcuda_engine *e) {
    if (!e->aux) {
        e->aux = (lkcuda_aux *)calloc(1, sizeof(lkcuda_aux));
    }
    return e->aux;
}

void lkc::aux_destroy(lkcuda_engine *e) {
    lkcuda_aux *a = e->aux;
    if (!a) return;
    cudaFree(a->d_boot); cudaFree(a->d_x); cudaFree(a->d_part); cudaFree(a->d_tickets); cudaFree(a->d_rell);
    if (a->h_rell) cudaFreeHost(a->h_rell);
    cudaFree(a->d_w); cudaFree(a->d_em_part); cudaFree(a->d_em_ticket); cudaFree(a->d_em_io);
    if (a->h_em_io) cudaFreeHost(a->h_em_io);
    cudaFree(a->d_rates); cudaFree(a->d_cat);
    free(a);
    e->aux = NULL;
}

template <typename T>
static int upload_boot(lkcuda_engine *e, lkcuda_aux *a, const float *samples, int nsamples, long long src_stride, long long P, long long stride) {
    std::vector<T> buf((size_t)nsamples * stride, (T)0);
    for (int s = 0; s < nsamples; s++) {
        const float *src = samples + (size_t)s * src_stride;
        T *dst = &buf[(size_t)s * stride];
        for (long long p = 0; p < P; p++) dst[p] = (T)src[p];
    }
    a->boot_bytes = buf.size() * sizeof(T);
    LK_CUDA(e, cudaMalloc(&a->d_boot, a->boot_bytes ? a->boot_bytes : 16));
    e->total_dev_bytes += a->boot_bytes;
    LK_CUDA(e, cudaMemcpy(a->d_boot, buf.data(), a->boot_bytes, cudaMemcpyHostToDevice));
    return LKCUDA_OK;
}

extern "C" {

int lkcuda_set_boot_samples(lkcuda_engine *e, const float *samples, int32_t nsamples, int64_t src_stride) {
    LK_ENTER(e);
    const DevView &v = e->v;
    if (nsamples < 0 || (nsamples && !samples) || src_stride < v.P) return fail(e, LKCUDA_ERR_ARG, "lkcuda_set_boot_samples: bad arguments");
    lkcuda_aux *a = aux_of(e);
    if (!a) return fail(e, LKCUDA_ERR_NOMEM, "out of host memory");
    LK_CUDA(e, cudaStreamSynchronize(e->stream));
    if (a->d_boot) { cudaFree(a->d_boot); e->total_dev_bytes -= a->boot_bytes; a->d_boot = NULL; a->boot_bytes = 0; }
    a->nsamples = nsamples;
    const long long stride = (v.P + 15) / 16 * 16;
    a->boot_stride = stride;
    if (nsamples == 0) return LKCUDA_OK;
    /* the narrowest LOSSLESS storage: bootstrap frequencies are small non-negative integers (iqtree.cpp:326-338) */
    float mx = 0.f; bool integral = true;
    for (int s = 0; s < nsamples && integral; s++) {
        const float *src = samples + (size_t)s * src_stride;
        for (long long p = 0; p < v.P; p++) {
            const float x = src[p];
            if (!(x >= 0.f) || x != (float)(long long)x || x > 65535.f) { integral = false; break; }
            if (x > mx) mx = x;
        }
    }
    a->boot_kind = !integral ? 2 : (mx <= 255.f ? 0 : 1);
    int rc;
    if (a->boot_kind == 0) rc = upload_boot<uint8_t>(e, a, samples, nsamples, src_stride, v.P, stride);
    else if (a->boot_kind == 1) rc = upload_boot<uint16_t>(e, a, samples, nsamples, src_stride, v.P, stride);
    else rc = upload_boot<float>(e, a, samples, nsamples, src_stride, v.P, stride);
    if (rc) return rc;
    if (!a->d_x) { LK_CUDA(e, cudaMalloc((void **)&a->d_x, (size_t)stride * 4 + 64)); e->total_dev_bytes += (size_t)stride * 4; }
    const int ngroups = (nsamples + RELL_SPB - 1) / RELL_SPB;
    long long chunks = (stride + AUX_THREADS * 16 - 1) / (AUX_THREADS * 16);
    /* gx chunk-blocks per sample group: the grid (gx * ngroups blocks) runs in ceil(blocks / resident) waves of ceil(chunks / gx) iterations
       each; take the gx with the least waves x iterations (a grid of 1.06 waves costs two) */
    int occ = 1;
    if (a->boot_kind == 0) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, rell_kernel<uint8_t>, AUX_THREADS, 0);
    else if (a->boot_kind == 1) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, rell_kernel<uint16_t>, AUX_THREADS, 0);
    else cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, rell_kernel<float>, AUX_THREADS, 0);
    if (occ < 1) occ = 1;
    const long long resident = (long long)e->num_sms * occ;
    long long best_cost = -1; a->rell_gx = 1;
    for (long long gx = 1; gx <= chunks && gx <= 4096; gx++) {
        const long long waves = (gx * ngroups + resident - 1) / resident, iters = (chunks + gx - 1) / gx;
        const long long cost = waves * iters * 64 + waves;          /* a little per wave for launch and reduction */
        if (best_cost < 0 || cost < best_cost) { best_cost = cost; a->rell_gx = (int)gx; }
    }
    if (a->rell_gx < 1) a->rell_gx = 1;
    const size_t need = (size_t)a->rell_gx * ngroups * RELL_SPB;
    if (need > a->part_cap || ngroups > a->groups_cap) {
        cudaFree(a->d_part); cudaFree(a->d_tickets); cudaFree(a->d_rell);
        if (a->h_rell) cudaFreeHost(a->h_rell);
        LK_CUDA(e, cudaMalloc((void **)&a->d_part, need * 8));
        LK_CUDA(e, cudaMalloc((void **)&a->d_tickets, (size_t)ngroups * 4));
        LK_CUDA(e, cudaMemset(a->d_tickets, 0, (size_t)ngroups * 4));
        LK_CUDA(e, cudaMalloc((void **)&a->d_rell, (size_t)ngroups * RELL_SPB * 8));
        LK_CUDA(e, cudaMallocHost((void **)&a->h_rell, (size_t)ngroups * RELL_SPB * 8));
        a->part_cap = need; a->groups_cap = ngroups;
    }
    return LKCUDA_OK;
}

int lkcuda_boot_storage(lkcuda_engine *e, int32_t *bytes_per_entry, uint64_t *bytes) {
    if (!e || !e->aux || !e->aux->d_boot) return LKCUDA_ERR_STATE;
    if (bytes_per_entry) *bytes_per_entry = e->aux->boot_kind == 0 ? 1 : (e->aux->boot_kind == 1 ? 2 : 4);
    if (bytes) *bytes = e->aux->boot_bytes;
    return LKCUDA_OK;
}

int lkcuda_rell(lkcuda_engine *e, int32_t vec_a, int32_t vec_b, double *rell) {
    LK_ENTER(e);
    lkcuda_aux *a = e->aux;
    if (!a || !a->d_boot) return fail(e, LKCUDA_ERR_STATE, "lkcuda_rell: no bootstrap samples (lkcuda_set_boot_samples)");
    if (vec_a >= e->cfg.nvec || vec_b >= e->cfg.nvec) return fail(e, LKCUDA_ERR_ARG, "lkcuda_rell: vector index out of range");
    const DevView &v = e->v;
    const int ngroups = (a->nsamples + RELL_SPB - 1) / RELL_SPB;
    launch_begin(e, KIND_MISC);
    int gx = (int)((a->boot_stride + AUX_THREADS - 1) / AUX_THREADS);
    if (gx > e->num_sms * 8) gx = e->num_sms * 8;
    if (gx < 1) gx = 1;
    rell_x_kernel<<<gx, AUX_THREADS, 0, e->stream>>>(e->v, vec_a, vec_b, a->d_x, a->boot_stride);
    int rc = launch_end(e, KIND_MISC, (uint64_t)v.P * 16, "rell_x_kernel");
    if (rc) return rc;
    launch_begin(e, KIND_RELL);
    dim3 grid((unsigned)a->rell_gx, (unsigned)ngroups);
    if (a->boot_kind == 0) rell_kernel<uint8_t><<<grid, AUX_THREADS, 0, e->stream>>>((const uint8_t *)a->d_boot, a->boot_stride, a->nsamples, a->d_x, a->d_part, a->d_tickets, a->d_rell);
    else if (a->boot_kind == 1) rell_kernel<uint16_t><<<grid, AUX_THREADS, 0, e->stream>>>((const uint16_t *)a->d_boot, a->boot_stride, a->nsamples, a->d_x, a->d_part, a->d_tickets, a->d_rell);
    else rell_kernel<float><<<grid, AUX_THREADS, 0, e->stream>>>((const float *)a->d_boot, a->boot_stride, a->nsamples, a->d_x, a->d_part, a->d_tickets, a->d_rell);
    /* algorithmic bytes: the reference's float matrix + the float pattern log-likelihoods once per sample (phylokernel.h:55-61) */
    rc = launch_end(e, KIND_RELL, (uint64_t)a->nsamples * (uint64_t)v.P * 8, "rell_kernel");
    if (rc) return rc;
    if (!rell) return LKCUDA_OK;
    return lkcuda_rell_fetch(e, rell);
}

int lkcuda_rell_fetch(lkcuda_engine *e, double *rell) {
    LK_ENTER(e);
    lkcuda_aux *a = e->aux;
    if (!a || !a->d_boot || !rell) return fail(e, LKCUDA_ERR_STATE, "lkcuda_rell_fetch: nothing to fetch");
    LK_CUDA(e, cudaMemcpyAsync(a->h_rell, a->d_rell, (size_t)a->nsamples * 8, cudaMemcpyDeviceToHost, e->stream));
    LK_CUDA(e, cudaStreamSynchronize(e->stream));
    memcpy(rell, a->h_rell, (size_t)a->nsamples * 8);
    return LKCUDA_OK;
}

int lkcuda_rell_device_ptr(lkcuda_engine *e, void **p) {
    if (!e || !p || !e->aux || !e->aux->d_rell) return LKCUDA_ERR_STATE;
    *p = e->aux->d_rell;
    return LKCUDA_OK;
}

int lkcuda_lh_cat_collapse(lkcuda_engine *e, int32_t mode, int32_t ncat, int32_t nmix, int32_t *ncol_out) {
    LK_ENTER(e);
    const DevView &v = e->v;
    if (mode < 0 || mode > 2 || ncat < 1 || nmix < 1 || (long long)ncat * nmix != v.K) return fail(e, LKCUDA_ERR_ARG, "lkcuda_lh_cat_collapse: ncat * nmix must equal the block count");
    lkcuda_aux *a = aux_of(e);
    if (!a) return fail(e, LKCUDA_ERR_NOMEM, "out of host memory");
    if (!a->d_w) { LK_CUDA(e, cudaMalloc((void **)&a->d_w, ((size_t)v.P * v.K + 4) * 8)); e->total_dev_bytes += (size_t)v.P * v.K * 8; }
    a->ncol = mode == 1 ? nmix : (mode == 2 ? ncat : ncat * nmix);
    if (ncol_out) *ncol_out = a->ncol;
    if (v.P == 0) return LKCUDA_OK;
    long long total = v.P * a->ncol;
    int grid = (int)((total + AUX_THREADS - 1) / AUX_THREADS < (long long)e->num_sms * 8 ? (total + AUX_THREADS - 1) / AUX_THREADS : (long long)e->num_sms * 8);
    launch_begin(e, KIND_MISC);
    lhcat_collapse_kernel<<<grid, AUX_THREADS, 0, e->stream>>>(e->v, mode, ncat, nmix, a->d_w);
    return launch_end(e, KIND_MISC, (uint64_t)v.P * (v.K + a->ncol) * 8, "lhcat_collapse_kernel");
}

/* E-step posteriors of one class as the pattern weights of ANOTHER engine (ratefree.cpp:487-503, 547-549; modelmixture.cpp:1453-1511):
   dst_freq[p] = w[p][col] * (ptn_freq[p] / ([ptn_invar[p] +] sum_c w[p][c])) - the reference's own order of operations; RateFree's
   E-step leaves ptn_invar out, ModelMixture's starts the sum with it */
__global__ void __launch_bounds__(AUX_THREADS) lhcat_posterior_kernel(DevView v, int ncol, const double *__restrict__ w, int col, int include_invar, double *dst_freq) {
    for (long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x; p < v.P; p += (long long)gridDim.x * blockDim.x) {
        const double *row = w + p * ncol;
        double lk = include_invar ? v.invar[p] : 0.0;
        for (int c = 0; c < ncol; c++) lk += row[c];
        lk = v.freq[p] / lk;
        dst_freq[p] = row[col] * lk;
    }
}

int lkcuda_lh_cat_posterior_weights(lkcuda_engine *e, int32_t col, int32_t include_invar, lkcuda_engine *dst) {
    LK_ENTER(e);
    lkcuda_aux *a = e->aux;
    if (!a || !a->d_w || a->ncol < 1) return fail(e, LKCUDA_ERR_STATE, "lkcuda_lh_cat_posterior_weights: call lkcuda_lh_cat_collapse first");
    if (!dst || dst == e) return fail(e, LKCUDA_ERR_ARG, "lkcuda_lh_cat_posterior_weights: destination engine");
    if (col < 0 || col >= a->ncol) return fail(e, LKCUDA_ERR_ARG, "lkcuda_lh_cat_posterior_weights: column %d of %d", col, a->ncol);
    if (dst->device != e->device || dst->v.P != e->v.P) return fail(e, LKCUDA_ERR_ARG, "lkcuda_lh_cat_posterior_weights: the engines must hold the same patterns on the same device");
    if (!e->weights_set || !dst->weights_set) return fail(e, LKCUDA_ERR_STATE, "lkcuda_lh_cat_posterior_weights: pattern weights not set (the destination keeps its ptn_invar)");
    const DevView &v = e->v;
    if (v.P == 0) return LKCUDA_OK;
    /* whatever the destination still runs with its old weights finishes first; its later work waits for the new ones */
    LK_CUDA(e, cudaStreamSynchronize(dst->stream));
    long long grid = (v.P + AUX_THREADS - 1) / AUX_THREADS;
    if (grid > (long long)e->num_sms * 8) grid = (long long)e->num_sms * 8;
    launch_begin(e, KIND_EM);
    lhcat_posterior_kernel<<<(int)grid, AUX_THREADS, 0, e->stream>>>(e->v, a->ncol, a->d_w, col, include_invar ? 1 : 0, dst->d_freq);
    int rc = launch_end(e, KIND_EM, (uint64_t)v.P * ((uint64_t)a->ncol * 8 + 16), "lhcat_posterior_kernel");
    if (rc) return rc;
    LK_CUDA(e, cudaStreamSynchronize(e->stream));
    dst->theta_ready = false;
    return LKCUDA_OK;
}

int lkcuda_lh_cat_em(lkcuda_engine *e, const double *ratio, double *sums) { return lkcuda_lh_cat_em_ex(e, ratio, 1, sums); }

int lkcuda_lh_cat_em_ex(lkcuda_engine *e, const double *ratio, int32_t include_invar, double *sums) {
    LK_ENTER(e);
    lkcuda_aux *a = e->aux;
    if (!a || !a->d_w || a->ncol < 1) return fail(e, LKCUDA_ERR_STATE, "lkcuda_lh_cat_em: call lkcuda_lh_cat_collapse first");
    if (!sums) return fail(e, LKCUDA_ERR_ARG, "null output");
    if (!e->weights_set) return fail(e, LKCUDA_ERR_STATE, "lkcuda_lh_cat_em: pattern weights not set");
    const int ncol = a->ncol;
    if (ncol > EM_MAXCOL) return fail(e, LKCUDA_ERR_ARG, "lkcuda_lh_cat_em: more than %d columns", EM_MAXCOL);
    const DevView &v = e->v;
    int grid = (int)((v.P + AUX_THREADS - 1) / AUX_THREADS);
    if (grid > e->num_sms * 2) grid = e->num_sms * 2;
    if (grid < 1) grid = 1;
    if (!a->d_em_part) {
        LK_CUDA(e, cudaMalloc((void **)&a->d_em_part, (size_t)e->num_sms * 2 * EM_MAXCOL * 8));
        LK_CUDA(e, cudaMalloc((void **)&a->d_em_ticket, 16));
        LK_CUDA(e, cudaMemset(a->d_em_ticket, 0, 16));
        LK_CUDA(e, cudaMalloc((void **)&a->d_em_io, 2 * EM_MAXCOL * 8));
        LK_CUDA(e, cudaMallocHost((void **)&a->h_em_io, 2 * EM_MAXCOL * 8));
    }
    if (ratio) {
        memcpy(a->h_em_io + EM_MAXCOL, ratio, (size_t)ncol * 8);
        LK_CUDA(e, cudaMemcpyAsync(a->d_em_io + EM_MAXCOL, a->h_em_io + EM_MAXCOL, (size_t)ncol * 8, cudaMemcpyHostToDevice, e->stream));
    }
    size_t sm = ((size_t)AUX_THREADS * (ncol + 1) + AUX_THREADS + AUX_THREADS + ncol) * 8;
    if (sm > 48 * 1024) LK_CUDA(e, cudaFuncSetAttribute(lhcat_em_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, LK_MAX_DYN_SMEM));
    launch_begin(e, KIND_EM);
    lhcat_em_kernel<<<grid, AUX_THREADS, sm, e->stream>>>(e->v, ncol, a->d_w, a->d_em_io + EM_MAXCOL, ratio ? 1 : 0, include_invar ? 1 : 0, a->d_em_part, a->d_em_ticket, a->d_em_io);
    int rc = launch_end(e, KIND_EM, (uint64_t)v.P * ((uint64_t)ncol * (ratio ? 16 : 8) + 16), "lhcat_em_kernel");
    if (rc) return rc;
    LK_CUDA(e, cudaMemcpyAsync(a->h_em_io, a->d_em_io, (size_t)ncol * 8, cudaMemcpyDeviceToHost, e->stream));
    LK_CUDA(e, cudaStreamSynchronize(e->stream));
    memcpy(sums, a->h_em_io, (size_t)ncol * 8);
    return LKCUDA_OK;
}

int lkcuda_lh_cat_pattern_rates(lkcuda_engine *e, const double *rates, int32_t invar_class, double *ptn_rates, int32_t *ptn_cat) {
    LK_ENTER(e);
    lkcuda_aux *a = e->aux;
    if (!a || !a->d_w || a->ncol < 1) return fail(e, LKCUDA_ERR_STATE, "lkcuda_lh_cat_pattern_rates: call lkcuda_lh_cat_collapse first");
    if (!rates || !ptn_rates || !ptn_cat) return fail(e, LKCUDA_ERR_ARG, "null argument");
    const DevView &v = e->v;
    if (a->ncol > EM_MAXCOL) return fail(e, LKCUDA_ERR_ARG, "lkcuda_lh_cat_pattern_rates: more than %d columns", EM_MAXCOL);
    if (v.P == 0) return LKCUDA_OK;
    if (!a->d_rates) {
        LK_CUDA(e, cudaMalloc((void **)&a->d_rates, ((size_t)v.P + EM_MAXCOL) * 8));
        LK_CUDA(e, cudaMalloc((void **)&a->d_cat, (size_t)v.P * 4));
        e->total_dev_bytes += (size_t)v.P * 12;
    }
    double *d_r = a->d_rates + v.P;
    LK_CUDA(e, cudaMemcpyAsync(d_r, rates, (size_t)a->ncol * 8, cudaMemcpyHostToDevice, e->stream));
    int grid = (int)((v.P + AUX_THREADS - 1) / AUX_THREADS < (long long)e->num_sms * 8 ? (v.P + AUX_THREADS - 1) / AUX_THREADS : (long long)e->num_sms * 8);
    launch_begin(e, KIND_MISC);
    lhcat_rates_kernel<<<grid, AUX_THREADS, 0, e->stream>>>(e->v, a->ncol, a->d_w, d_r, invar_class ? 1 : 0, a->d_rates, a->d_cat);
    int rc = launch_end(e, KIND_MISC, (uint64_t)v.P * ((uint64_t)a->ncol * 8 + 12), "lhcat_rates_kernel");
    if (rc) return rc;
    LK_CUDA(e, cudaMemcpyAsync(ptn_rates, a->d_rates, (size_t)v.P * 8, cudaMemcpyDeviceToHost, e->stream));
    LK_CUDA(e, cudaMemcpyAsync(ptn_cat, a->d_cat, (size_t)v.P * 4, cudaMemcpyDeviceToHost, e->stream));
    LK_CUDA(e, cudaStreamSynchronize(e->stream));
    return LKCUDA_OK;
}

int lkcuda_fetch_lh_cat_collapsed(lkcuda_engine *e, double *dst) {
    LK_ENTER(e);
    lkcuda_aux *a = e->aux;
    if (!a || !a->d_w || a->ncol < 1) return fail(e, LKCUDA_ERR_STATE, "lkcuda_fetch_lh_cat_collapsed: call lkcuda_lh_cat_collapse first");
    if (!dst) return fail(e, LKCUDA_ERR_ARG, "null destination");
    if (e->v.P) LK_CUDA(e, cudaMemcpyAsync(dst, a->d_w, (size_t)e->v.P * a->ncol * 8, cudaMemcpyDeviceToHost, e->stream));
    LK_CUDA(e, cudaStreamSynchronize(e->stream));
    return LKCUDA_OK;
}

}  // extern "C"
