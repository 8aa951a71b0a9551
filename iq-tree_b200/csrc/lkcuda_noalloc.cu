/*
 * lkcuda_derv_noalloc (include/lkcuda.h): df / ddf of one branch from a caller-owned theta_all, for callers that must not allocate
 * (code under CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc).  It has no counterpart in the reference; it
 * was added and touches no other code path of the library.
 *
 * The arithmetic is the one lkcuda_derv runs on an engine's theta_all (PhyloTree::computeLikelihoodDervEigenSIMD's buffered loop,
 * phylokernel.h:606-651): val0[c][i] = exp(eval[i] * rate[c] * len) * prop[c], val1 = eval*rate*val0, val2 = (eval*rate)^2 * val0;
 * per pattern lh = |sum theta*val0 + ptn_invar|, df += freq * (sum theta*val1) / lh, ddf += freq * ((sum theta*val2) / lh - (.. / lh)^2).
 *
 * The cross-block scratch (one {df, ddf} pair per block + the ticket that elects the finishing block) does NOT come from cudaMalloc
 * like lkcuda_engine::d_red: it is the file-scope __device__ array below, LKCUDA_NOALLOC_SCRATCH_BYTES on every device that loads
 * the library, until the process exits.  In host code the array's name is only a stand-in: the device address is taken with
 * cudaGetSymbolAddress once per device (that device current) and handed to the kernel.
 * Every call shares the array: calls that overlap on the GPU (two streams, two host threads) race.  The ticket is cleared by a
 * cudaMemsetAsync on the caller's stream in front of every launch; the partial sums are fully rewritten by every launch that reads them.
 */
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <mutex>

#include "../../include/lkcuda.h"
#include "lkcuda_common.cuh"

#define NA_THREADS 256
#define NA_WARPS (NA_THREADS / 32)
#define NA_TILES_PER_WARP 2
#define NA_PTN_PER_BLOCK (NA_WARPS * NA_TILES_PER_WARP * 32)                              /* 512 */
#define NA_MAX_BLOCKS (LKCUDA_NOALLOC_MAX_PATTERNS / NA_PTN_PER_BLOCK)                    /* 4096 */
#define NA_MAX_DIM 64                                                                     /* states, categories */
#define NA_COLS 16                                                                        /* columns of a pattern block staged at a time */

struct NaScratch {
    double part[NA_MAX_BLOCKS][2];      /* {df, ddf} of block b, rewritten by every launch */
    unsigned int ticket;                /* blocks that have published; must be 0 when a launch starts */
    unsigned int pad[63];
};
static_assert(sizeof(NaScratch) == LKCUDA_NOALLOC_SCRATCH_BYTES, "LKCUDA_NOALLOC_SCRATCH_BYTES (include/lkcuda.h) is the size of the static scratch");
static_assert(NA_MAX_BLOCKS * NA_PTN_PER_BLOCK == LKCUDA_NOALLOC_MAX_PATTERNS, "block count of the scratch");

static __device__ __align__(256) NaScratch g_noalloc_scratch;

struct NaModel {
    double eval[NA_MAX_DIM], rate[NA_MAX_DIM], prop[NA_MAX_DIM];
    int S, K;
};

/* block b owns patterns [512 b, 512 b + 512): a fixed assignment and a fixed summation order, so two calls on the same inputs give
   the same bits.  A warp stages 32 patterns x 16 columns of theta_all in shared memory with coalesced reads (a pattern's block is
   contiguous), then lane l sums pattern l. */
__global__ void __launch_bounds__(NA_THREADS) derv_noalloc_kernel(NaScratch *scr, const double *__restrict__ theta, const double *__restrict__ freq,
                                                                  const double *__restrict__ invar, long long nptn, NaModel m, double len,
                                                                  double *out) {
    __shared__ double val[3][LKCUDA_NOALLOC_MAX_BLOCK];
    __shared__ double tile[NA_WARPS][32 * (NA_COLS + 1)];
    __shared__ double red[2][NA_THREADS];
    __shared__ unsigned int is_last;
    const int B = m.S * m.K, tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    for (int j = tid; j < B; j += NA_THREADS) {
        const int k = j / m.S, i = j - k * m.S;
        const double cof = m.eval[i] * m.rate[k];
        const double x = exp(cof * len) * m.prop[k], y = cof * x;
        val[0][j] = x; val[1][j] = y; val[2][j] = cof * y;
    }
    __syncthreads();
    double df = 0.0, ddf = 0.0;
    double *tw = tile[w];
    for (int it = 0; it < NA_TILES_PER_WARP; it++) {
        const long long p0 = (long long)blockIdx.x * NA_PTN_PER_BLOCK + (long long)(it * NA_WARPS + w) * 32;
        if (p0 >= nptn) break;          /* uniform over the warp */
        const int np = (int)min(32ll, nptn - p0);
        double s0 = 0.0, s1 = 0.0, s2 = 0.0;
        for (int c0 = 0; c0 < B; c0 += NA_COLS) {
            const int nc = min(NA_COLS, B - c0);
            for (int idx = lane; idx < 32 * NA_COLS; idx += 32) {
                const int row = idx / NA_COLS, col = idx - row * NA_COLS;
                if (row < np && col < nc) tw[row * (NA_COLS + 1) + col] = __ldg(theta + (p0 + row) * B + c0 + col);
            }
            __syncwarp();
            if (lane < np) {
                for (int col = 0; col < nc; col++) {
                    const double x = tw[lane * (NA_COLS + 1) + col];
                    s0 = fma(x, val[0][c0 + col], s0); s1 = fma(x, val[1][c0 + col], s1); s2 = fma(x, val[2][c0 + col], s2);
                }
            }
            __syncwarp();
        }
        if (lane < np) {
            const long long p = p0 + lane;
            const double lh = fabs(s0 + invar[p]), df_frac = s1 / lh, ddf_frac = s2 / lh, f = freq[p];
            const double t1 = df_frac * f, t2 = ddf_frac * f;
            df += t1;
            ddf += t2 - t1 * df_frac;
        }
    }
    /* block sum in a fixed order */
    red[0][tid] = df; red[1][tid] = ddf;
    __syncthreads();
    for (int off = NA_THREADS / 2; off > 0; off >>= 1) {
        if (tid < off) { red[0][tid] += red[0][tid + off]; red[1][tid] += red[1][tid + off]; }
        __syncthreads();
    }
    if (tid == 0) {
        scr->part[blockIdx.x][0] = red[0][0]; scr->part[blockIdx.x][1] = red[1][0];
        __threadfence();
        is_last = atomicAdd(&scr->ticket, 1u) == gridDim.x - 1 ? 1u : 0u;
    }
    __syncthreads();
    if (!is_last) return;
    /* the block that took the last ticket adds the blocks' pairs, again in a fixed order (whichever block it is) */
    __threadfence();
    double a = 0.0, b = 0.0;
    for (unsigned int j = tid; j < gridDim.x; j += NA_THREADS) { a += __ldcg(&scr->part[j][0]); b += __ldcg(&scr->part[j][1]); }
    red[0][tid] = a; red[1][tid] = b;
    __syncthreads();
    for (int off = NA_THREADS / 2; off > 0; off >>= 1) {
        if (tid < off) { red[0][tid] += red[0][tid + off]; red[1][tid] += red[1][tid + off]; }
        __syncthreads();
    }
    if (tid == 0) {
        double rdf = red[0][0], rddf = red[1][0];
        if (isnan(rdf) || isinf(rdf)) { rdf = 0.0; rddf = 0.0; }       /* phylokernel.h:647-651 */
        out[0] = rdf; out[1] = rddf;
        scr->ticket = 0;        /* left clean as well; the next call clears it anyway */
    }
}

#define NA_MAX_DEVICES 64
static std::mutex g_na_mu;
static NaScratch *g_na_ptr[NA_MAX_DEVICES];      /* device address of g_noalloc_scratch per device (cudaGetSymbolAddress) */

/* once per device, that device current: the device address behind the host-side stand-in `g_noalloc_scratch` */
static int noalloc_scratch(int device, NaScratch **ptr) {
    std::lock_guard<std::mutex> lock(g_na_mu);
    if (!g_na_ptr[device]) {
        void *p = NULL;
        cudaError_t rc = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (rc != cudaSuccess) return lkc::cuda_fail(NULL, rc, "lkcuda_derv_noalloc: cudaGetSymbolAddress(scratch)");
        if (!p) return lkc::fail(NULL, LKCUDA_ERR_CUDA, "lkcuda_derv_noalloc: cudaGetSymbolAddress returned a null address");
        g_na_ptr[device] = static_cast<NaScratch *>(p);
    }
    *ptr = g_na_ptr[device];
    return LKCUDA_OK;
}

extern "C" int lkcuda_derv_noalloc(int32_t device, void *cuda_stream, const double *theta_all, const double *ptn_freq, const double *ptn_invar,
                                   int64_t nptn, int32_t nstates, int32_t ncat, const double *eval, const double *cat_rate, const double *cat_prop,
                                   double len, double *df_ddf) {
    using lkc::fail;
    if (!theta_all || !ptn_freq || !ptn_invar || !eval || !cat_rate || !cat_prop || !df_ddf) return fail(NULL, LKCUDA_ERR_ARG, "lkcuda_derv_noalloc: null argument");
    if (nptn < 0 || nstates <= 0 || ncat <= 0) return fail(NULL, LKCUDA_ERR_ARG, "lkcuda_derv_noalloc: invalid shape");
    if (nptn > LKCUDA_NOALLOC_MAX_PATTERNS)
        return fail(NULL, LKCUDA_ERR_ARG, "lkcuda_derv_noalloc: %lld patterns exceed LKCUDA_NOALLOC_MAX_PATTERNS = %d (the static scratch holds one pair per %d patterns); use lkcuda_derv on an engine",
                    (long long)nptn, LKCUDA_NOALLOC_MAX_PATTERNS, NA_PTN_PER_BLOCK);
    if (nstates > NA_MAX_DIM || ncat > NA_MAX_DIM || nstates * ncat > LKCUDA_NOALLOC_MAX_BLOCK)
        return fail(NULL, LKCUDA_ERR_ARG, "lkcuda_derv_noalloc: %d states x %d categories exceed the limits (%d each, LKCUDA_NOALLOC_MAX_BLOCK = %d together)",
                    nstates, ncat, NA_MAX_DIM, LKCUDA_NOALLOC_MAX_BLOCK);
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) { cudaGetLastError(); return fail(NULL, LKCUDA_ERR_NODEVICE, "no CUDA device available (lkcuda_derv_noalloc has no CPU path)"); }
    if (device < 0 || device >= ndev || device >= NA_MAX_DEVICES) return fail(NULL, LKCUDA_ERR_ARG, "lkcuda_derv_noalloc: device ordinal %d out of range (%d devices)", device, ndev);
    int prev = -1;
    cudaError_t rc = cudaGetDevice(&prev);
    if (rc != cudaSuccess) return lkc::cuda_fail(NULL, rc, "cudaGetDevice");
    if (prev != device) { rc = cudaSetDevice(device); if (rc != cudaSuccess) return lkc::cuda_fail(NULL, rc, "cudaSetDevice"); }
    cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
    NaScratch *scr = NULL;
    int code = noalloc_scratch(device, &scr);
    if (code == LKCUDA_OK) {
        if (nptn == 0) {
            rc = cudaMemsetAsync(df_ddf, 0, 2 * sizeof(double), stream);
        } else {
            rc = cudaMemsetAsync(&scr->ticket, 0, sizeof(unsigned int), stream);        /* what must start at zero */
            if (rc == cudaSuccess) {
                NaModel m;
                for (int i = 0; i < NA_MAX_DIM; i++) { m.eval[i] = i < nstates ? eval[i] : 0.0; m.rate[i] = i < ncat ? cat_rate[i] : 0.0; m.prop[i] = i < ncat ? cat_prop[i] : 0.0; }
                m.S = nstates; m.K = ncat;
                const unsigned int grid = (unsigned int)((nptn + NA_PTN_PER_BLOCK - 1) / NA_PTN_PER_BLOCK);
                derv_noalloc_kernel<<<grid, NA_THREADS, 0, stream>>>(scr, theta_all, ptn_freq, ptn_invar, (long long)nptn, m, len, df_ddf);
                rc = cudaGetLastError();
            }
        }
        if (rc != cudaSuccess) code = lkc::cuda_fail(NULL, rc, "lkcuda_derv_noalloc");
    }
    if (prev != device) cudaSetDevice(prev);
    return code;
}
