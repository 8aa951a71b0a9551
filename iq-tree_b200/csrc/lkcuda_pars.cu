/*
 * lkcuda_pars.cu — SURVEY 8(f).4: Fitch parsimony on the device (include/lkpars.h).
 * Replaces PhyloTree::computePartialParsimonyFast / computeParsimonyBranchFast (phylotreepars.cpp:18-282) and their SIMD twins
 * (phylokernel.h:1265-1572): per 32-site word, z[i] = x[i] & y[i]; w = ~(OR_i z[i]); score += popc(w); z[i] |= w & (x[i] | y[i]).
 *
 * B200 design.  Every 32-site word column is independent of every other one through the WHOLE tree, so - as in the likelihood
 * traversal kernels - one launch runs a complete dependent op list: a thread owns a word column, walks the ops in order and re-reads
 * only what it wrote itself (plain coherent loads, no fences, no grid synchronisation).  The counts are integer sums (REDUX warp
 * reduction -> shared-memory atomics -> one global atomic per op and block): exact and order-independent.  The ops at the END of a list
 * that write nothing (branch scores, and the 2k-3 candidate insertions of one stepwise-addition round, phylotreepars.cpp:366-374) do
 * not depend on each other: they run as a second, WIDE launch with one grid row per op, which is what fills the machine
 * (1M sites are only 31 250 word columns).  Vectors are bits[state][word]: a warp reads 128 contiguous bytes per state plane.
 * Bound: HBM / L2 bandwidth, 3 * S * 4 bytes per word and partial op, 2 * S * 4 (3 * S * 4) per branch (insertion) score.
 */
#include <cuda_runtime.h>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <algorithm>
#include <string>
#include <vector>
#include "../../include/lkcuda.h"
#include "../../include/lkpars.h"

struct lkpars_engine {
    int device, S, ntaxa, nvec;
    long long nsites, nw;           /* informative sites, uint32 words per plane */
    uint32_t *d_bits;               /* [ntaxa + nvec][S][nw] */
    lkpars_op *d_ops;  int ops_cap;
    unsigned int *d_counts;         /* [ops_cap][2] */
    lkpars_op *h_ops;  unsigned int *h_counts;      /* pinned */
    cudaStream_t stream;
    cudaEvent_t ev0, ev1;
    bool profiling;
    uint64_t launches;  double device_ms;
    int num_sms;
    std::string err;
};

static thread_local std::string g_pars_create_error;

static int pfail(lkpars_engine *e, int code, const char *fmt, ...) {
    char buf[512];
    va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
    if (e) e->err = buf; else g_pars_create_error = buf;
    return code;
}
#define PARS_CUDA(e, call) do { cudaError_t _c = (call); if (_c != cudaSuccess) return pfail(e, LKCUDA_ERR_CUDA, "%s: %s", #call, cudaGetErrorString(_c)); } while (0)

#define PARS_THREADS 128

/* tip vectors: thread = (taxon, word); bit s of plane i is set when state i is in the code's state set; the dummy sites that fill the
   last word get state 0 (phylotreepars.cpp:72-74) */
__global__ void pars_tips_kernel(uint32_t *bits, const uint8_t *__restrict__ states, const unsigned long long *__restrict__ mask,
                                 int nmask, int S, int ntaxa, long long nsites, long long nw) {
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)ntaxa * nw) return;
    const long long t = idx / nw, w = idx - t * nw;
    const uint8_t *st = states + t * nsites + w * 32;
    unsigned long long m[32];
#pragma unroll
    for (int s = 0; s < 32; s++) {
        const long long site = w * 32 + s;
        if (site < nsites) { const int code = st[s]; m[s] = code < nmask ? mask[code] : 0ull; }
        else m[s] = 1ull;
    }
    uint32_t *dst = bits + t * S * nw + w;
    for (int i = 0; i < S; i++) {
        uint32_t v = 0;
#pragma unroll
        for (int s = 0; s < 32; s++) v |= (uint32_t)((m[s] >> i) & 1ull) << s;
        dst[(long long)i * nw] = v;
    }
}

/* one op on one word column; S > 0: planes in registers, S == 0: two passes over the planes (any state count <= 64) */
template <int S>
__device__ __forceinline__ void pars_column(uint32_t *bits, long long vstride, long long nw, long long w, int ns, const lkpars_op &op,
                                            unsigned int &c0, unsigned int &c1) {
    const uint32_t *x = bits + op.a * vstride + w, *y = bits + op.b * vstride + w;
    c0 = 0; c1 = 0;
    if (S > 0) {
        uint32_t xv[S > 0 ? S : 1], yv[S > 0 ? S : 1];
#pragma unroll
        for (int i = 0; i < S; i++) { xv[i] = x[i * nw]; yv[i] = y[i * nw]; }
        uint32_t any = 0;
#pragma unroll
        for (int i = 0; i < S; i++) any |= xv[i] & yv[i];
        const uint32_t u = ~any;
        c0 = __popc(u);
        if (op.kind == LKPARS_OP_BRANCH) return;
#pragma unroll
        for (int i = 0; i < S; i++) xv[i] = (xv[i] & yv[i]) | (u & (xv[i] | yv[i]));
        if (op.dst >= 0) {
            uint32_t *z = bits + op.dst * vstride + w;
#pragma unroll
            for (int i = 0; i < S; i++) z[i * nw] = xv[i];
        }
        if (op.kind == LKPARS_OP_INSERT) {
            const uint32_t *t = bits + op.c * vstride + w;
            uint32_t hit = 0;
#pragma unroll
            for (int i = 0; i < S; i++) hit |= xv[i] & t[i * nw];
            c1 = __popc(~hit);
        }
    } else {
        uint32_t any = 0;
        for (int i = 0; i < ns; i++) any |= x[i * nw] & y[i * nw];
        const uint32_t u = ~any;
        c0 = __popc(u);
        if (op.kind == LKPARS_OP_BRANCH) return;
        uint32_t *z = op.dst >= 0 ? bits + op.dst * vstride + w : NULL;
        const uint32_t *t = op.kind == LKPARS_OP_INSERT ? bits + op.c * vstride + w : NULL;
        uint32_t hit = 0;
        for (int i = 0; i < ns; i++) {
            const uint32_t xi = x[i * nw], yi = y[i * nw];
            const uint32_t zi = (xi & yi) | (u & (xi | yi));
            if (z) z[i * nw] = zi;
            if (t) hit |= zi & t[i * nw];
        }
        if (t) c1 = __popc(~hit);
    }
}

/* the dependent part of an op list: a thread owns word columns and walks all ops.  A warp has only 2 * S loads of one op in flight
   and 1M sites are just 977 warps, so a naive walk is bound by the memory round trip per op (about 1 us).  Consecutive PARTIAL ops
   that do not read each other's results (the host orders a queue by tree level, so most do not) are therefore taken up to 8 at a
   time: all their operands are loaded before the first result is stored (which also makes write-after-read inside a group safe). */
#define PARS_GROUP_MAX 8

__device__ __forceinline__ void pars_count(unsigned int *sh, int o, unsigned int c0, unsigned int c1, bool two) {
    c0 = __reduce_add_sync(0xffffffffu, c0);
    if (two) c1 = __reduce_add_sync(0xffffffffu, c1);
    if ((threadIdx.x & 31) == 0) {
        if (c0) atomicAdd(&sh[2 * o], c0);
        if (two && c1) atomicAdd(&sh[2 * o + 1], c1);
    }
}

template <int S>
__global__ void __launch_bounds__(PARS_THREADS) pars_chain_kernel(uint32_t *bits, long long vstride, long long nw, int ns,
                                                                   const lkpars_op *ops, int nops, unsigned int *counts) {
    extern __shared__ unsigned int sh[];                /* [nops][2] */
    for (int i = threadIdx.x; i < 2 * nops; i += blockDim.x) sh[i] = 0;
    __syncthreads();
    constexpr int G = S == 4 ? 8 : (S > 0 ? 4 : 1);     /* 2 * S * G registers of operands */
    const long long nthreads = (long long)gridDim.x * blockDim.x;
    const long long nw_round = (nw + 31) / 32 * 32;
    for (long long w = (long long)blockIdx.x * blockDim.x + threadIdx.x; w < nw_round; w += nthreads) {
        const bool live = w < nw;
        int o = 0;
        while (o < nops) {
            const lkpars_op op = ops[o];
            /* lkpars_run stored, in the unused third operand of a PARTIAL op, how many of the following ops form a hazard-free group */
            const int g = (S > 0 && op.kind == LKPARS_OP_PARTIAL && op.c > 1) ? (op.c < G ? op.c : G) : 1;
            if (S > 0 && g > 1) {
                uint32_t xv[G][S > 0 ? S : 1], yv[G][S > 0 ? S : 1];
                if (live) {
#pragma unroll
                    for (int j = 0; j < G; j++) if (j < g) {
                        const lkpars_op n = ops[o + j];
                        const uint32_t *x = bits + n.a * vstride + w, *y = bits + n.b * vstride + w;
#pragma unroll
                        for (int i = 0; i < S; i++) { xv[j][i] = x[i * nw]; yv[j][i] = y[i * nw]; }
                    }
                }
#pragma unroll
                for (int j = 0; j < G; j++) if (j < g) {
                    unsigned int c0 = 0;
                    if (live) {
                        uint32_t any = 0;
#pragma unroll
                        for (int i = 0; i < S; i++) any |= xv[j][i] & yv[j][i];
                        const uint32_t u = ~any;
                        c0 = __popc(u);
                        uint32_t *z = bits + ops[o + j].dst * vstride + w;
#pragma unroll
                        for (int i = 0; i < S; i++) z[i * nw] = (xv[j][i] & yv[j][i]) | (u & (xv[j][i] | yv[j][i]));
                    }
                    pars_count(sh, o + j, c0, 0, false);
                }
            } else {
                unsigned int c0 = 0, c1 = 0;
                if (live) pars_column<S>(bits, vstride, nw, w, ns, op, c0, c1);
                pars_count(sh, o, c0, c1, op.kind == LKPARS_OP_INSERT);
            }
            o += g;
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 2 * nops; i += blockDim.x) if (sh[i]) atomicAdd(&counts[i], sh[i]);
}

/* the read-only tail of an op list: grid row = op, every op streams its 2 or 3 vectors independently */
template <int S>
__global__ void __launch_bounds__(PARS_THREADS) pars_wide_kernel(uint32_t *bits, long long vstride, long long nw, int ns,
                                                                  const lkpars_op *ops, unsigned int *counts) {
    __shared__ unsigned int sh[2];
    if (threadIdx.x < 2) sh[threadIdx.x] = 0;
    __syncthreads();
    const lkpars_op op = ops[blockIdx.y];
    const long long nthreads = (long long)gridDim.x * blockDim.x;
    unsigned int t0 = 0, t1 = 0;
    for (long long w = (long long)blockIdx.x * blockDim.x + threadIdx.x; w < nw; w += nthreads) {
        unsigned int c0, c1;
        pars_column<S>(bits, vstride, nw, w, ns, op, c0, c1);
        t0 += c0; t1 += c1;
    }
    t0 = __reduce_add_sync(0xffffffffu, t0);
    t1 = __reduce_add_sync(0xffffffffu, t1);
    if ((threadIdx.x & 31) == 0) { if (t0) atomicAdd(&sh[0], t0); if (t1) atomicAdd(&sh[1], t1); }
    __syncthreads();
    if (threadIdx.x < 2 && sh[threadIdx.x]) atomicAdd(&counts[2 * blockIdx.y + threadIdx.x], sh[threadIdx.x]);
}

template <int S>
static void launch_pars(lkpars_engine *e, const lkpars_op *d_ops, unsigned int *d_counts, int nchain, int nwide) {
    const long long vstride = (long long)e->S * e->nw;
    if (nchain > 0) {
        long long blocks = (e->nw + PARS_THREADS - 1) / PARS_THREADS;
        const long long cap = (long long)e->num_sms * 8;
        if (blocks > cap) blocks = cap;
        pars_chain_kernel<S><<<(int)blocks, PARS_THREADS, 2 * nchain * sizeof(unsigned int), e->stream>>>(e->d_bits, vstride, e->nw, e->S, d_ops, nchain, d_counts);
        e->launches++;
    }
    if (nwide > 0) {
        /* enough columns per thread to amortise the block, enough blocks to fill the machine */
        long long bx = (e->nw + PARS_THREADS * 4 - 1) / (PARS_THREADS * 4);
        if (bx < 1) bx = 1;
        const long long cap = ((long long)e->num_sms * 16 + nwide - 1) / nwide;
        if (bx > cap) bx = cap < 1 ? 1 : cap;
        for (int o = 0; o < nwide; o += 65535) {
            const int rows = nwide - o < 65535 ? nwide - o : 65535;
            pars_wide_kernel<S><<<dim3((unsigned)bx, (unsigned)rows), PARS_THREADS, 0, e->stream>>>(e->d_bits, vstride, e->nw, e->S, d_ops + nchain + o, d_counts + 2 * (nchain + o));
            e->launches++;
        }
    }
}

extern "C" int lkpars_create(int32_t device, int32_t nstates, int32_t ntaxa, int64_t nsites, const uint8_t *site_states,
                             const uint64_t *state_mask, int32_t nmask, int32_t nvec, lkpars_engine **out) {
    if (!out) return pfail(NULL, LKCUDA_ERR_ARG, "null out pointer");
    *out = NULL;
    if (nstates < 2 || nstates > 64) return pfail(NULL, LKCUDA_ERR_ARG, "nstates %d outside 2..64", nstates);
    if (ntaxa < 1 || nsites < 0 || nvec < 0 || nmask < nstates || nmask > 256) return pfail(NULL, LKCUDA_ERR_ARG, "bad sizes (ntaxa %d, nsites %lld, nvec %d, nmask %d)", ntaxa, (long long)nsites, nvec, nmask);
    if (!state_mask || (nsites > 0 && !site_states)) return pfail(NULL, LKCUDA_ERR_ARG, "null argument");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) return pfail(NULL, LKCUDA_ERR_CUDA, "no CUDA device: the parsimony engine has no CPU fallback");
    if (device < 0 || device >= ndev) return pfail(NULL, LKCUDA_ERR_ARG, "device %d of %d", device, ndev);
    lkpars_engine *e = new lkpars_engine();
    e->device = device; e->S = nstates; e->ntaxa = ntaxa; e->nvec = nvec; e->nsites = nsites;
    e->nw = (nsites + 31) / 32;
    if (e->nw == 0) e->nw = 1;
    e->d_bits = NULL; e->d_ops = NULL; e->d_counts = NULL; e->h_ops = NULL; e->h_counts = NULL; e->ops_cap = 0;
    e->profiling = false; e->launches = 0; e->device_ms = 0.0;
    cudaError_t c;
    uint8_t *d_states = NULL; unsigned long long *d_mask = NULL;
#define CREATE_CUDA(call) if ((c = (call)) != cudaSuccess) { int rc = pfail(NULL, LKCUDA_ERR_CUDA, "%s: %s", #call, cudaGetErrorString(c)); cudaFree(d_states); cudaFree(d_mask); lkpars_destroy(e); return rc; }
    CREATE_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    CREATE_CUDA(cudaGetDeviceProperties(&prop, device));
    e->num_sms = prop.multiProcessorCount;
    CREATE_CUDA(cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking));
    CREATE_CUDA(cudaEventCreate(&e->ev0));
    CREATE_CUDA(cudaEventCreate(&e->ev1));
    const size_t words = (size_t)(ntaxa + nvec) * nstates * e->nw;
    CREATE_CUDA(cudaMalloc((void **)&e->d_bits, words * 4));
    CREATE_CUDA(cudaMalloc((void **)&d_states, (size_t)ntaxa * (nsites > 0 ? nsites : 1)));
    CREATE_CUDA(cudaMalloc((void **)&d_mask, 256 * 8));
    if (nsites > 0) CREATE_CUDA(cudaMemcpyAsync(d_states, site_states, (size_t)ntaxa * nsites, cudaMemcpyHostToDevice, e->stream));
    CREATE_CUDA(cudaMemcpyAsync(d_mask, state_mask, (size_t)nmask * 8, cudaMemcpyHostToDevice, e->stream));
    const long long n = (long long)ntaxa * e->nw;
    pars_tips_kernel<<<(unsigned)((n + 127) / 128), 128, 0, e->stream>>>(e->d_bits, d_states, d_mask, nmask, nstates, ntaxa, nsites, e->nw);
    e->launches++;
    CREATE_CUDA(cudaGetLastError());
    CREATE_CUDA(cudaStreamSynchronize(e->stream));
    cudaFree(d_states); cudaFree(d_mask);
#undef CREATE_CUDA
    *out = e;
    return LKCUDA_OK;
}

extern "C" int lkpars_destroy(lkpars_engine *e) {
    if (!e) return LKCUDA_OK;
    cudaSetDevice(e->device);
    if (e->d_bits) cudaFree(e->d_bits);
    if (e->d_ops) cudaFree(e->d_ops);
    if (e->d_counts) cudaFree(e->d_counts);
    if (e->h_ops) cudaFreeHost(e->h_ops);
    if (e->h_counts) cudaFreeHost(e->h_counts);
    if (e->ev0) cudaEventDestroy(e->ev0);
    if (e->ev1) cudaEventDestroy(e->ev1);
    if (e->stream) cudaStreamDestroy(e->stream);
    delete e;
    return LKCUDA_OK;
}

extern "C" const char *lkpars_last_error(const lkpars_engine *e) { return e ? e->err.c_str() : g_pars_create_error.c_str(); }

extern "C" int lkpars_run(lkpars_engine *e, const lkpars_op *ops, int32_t nops, uint32_t *counts) {
    if (!e) return LKCUDA_ERR_ARG;
    if (nops < 0 || (nops > 0 && (!ops || !counts))) return pfail(e, LKCUDA_ERR_ARG, "null argument");
    if (nops == 0) return LKCUDA_OK;
    PARS_CUDA(e, cudaSetDevice(e->device));
    const int total = e->ntaxa + e->nvec;
    for (int i = 0; i < nops; i++) {
        const lkpars_op &op = ops[i];
        if (op.kind < 0 || op.kind > 2) return pfail(e, LKCUDA_ERR_ARG, "op %d: kind %d", i, op.kind);
        if (op.a < 0 || op.a >= total || op.b < 0 || op.b >= total) return pfail(e, LKCUDA_ERR_ARG, "op %d: operand out of range", i);
        if (op.kind == LKPARS_OP_INSERT && (op.c < 0 || op.c >= total)) return pfail(e, LKCUDA_ERR_ARG, "op %d: third operand out of range", i);
        if (op.kind == LKPARS_OP_PARTIAL && (op.dst < e->ntaxa || op.dst >= total)) return pfail(e, LKCUDA_ERR_ARG, "op %d: dst %d is not an internal vector", i, op.dst);
        if (op.kind == LKPARS_OP_INSERT && op.dst >= 0 && (op.dst < e->ntaxa || op.dst >= total)) return pfail(e, LKCUDA_ERR_ARG, "op %d: dst %d is not an internal vector", i, op.dst);
        if (op.kind != LKPARS_OP_BRANCH && op.dst >= 0 && (op.dst == op.a || op.dst == op.b)) return pfail(e, LKCUDA_ERR_ARG, "op %d: dst aliases an operand", i);
    }
    if (nops > e->ops_cap) {
        int cap = nops + 256;
        if (e->d_ops) cudaFree(e->d_ops);
        if (e->d_counts) cudaFree(e->d_counts);
        if (e->h_ops) cudaFreeHost(e->h_ops);
        if (e->h_counts) cudaFreeHost(e->h_counts);
        e->d_ops = NULL; e->d_counts = NULL; e->h_ops = NULL; e->h_counts = NULL; e->ops_cap = 0;
        PARS_CUDA(e, cudaMalloc((void **)&e->d_ops, sizeof(lkpars_op) * cap));
        PARS_CUDA(e, cudaMalloc((void **)&e->d_counts, sizeof(unsigned int) * 2 * cap));
        PARS_CUDA(e, cudaMallocHost((void **)&e->h_ops, sizeof(lkpars_op) * cap));
        PARS_CUDA(e, cudaMallocHost((void **)&e->h_counts, sizeof(unsigned int) * 2 * cap));
        e->ops_cap = cap;
    }
    memcpy(e->h_ops, ops, sizeof(lkpars_op) * nops);
    /* the trailing ops that write nothing are independent of each other */
    int nchain = nops;
    while (nchain > 0 && (ops[nchain - 1].kind == LKPARS_OP_BRANCH || (ops[nchain - 1].kind == LKPARS_OP_INSERT && ops[nchain - 1].dst < 0))) nchain--;
    const int nwide = nops - nchain;
    /* hazard-free groups of consecutive PARTIAL ops (no op reads or rewrites what an earlier op of the group writes): the size goes
       into the unused third operand of the group's first op in the device copy */
    const int CHAIN_MAX = 4096;      /* ops per chain launch: 8 bytes of shared memory each */
    const int gmax = e->S == 4 ? 8 : (e->S == 20 ? 4 : 1);     /* = G of the kernel instantiation */
    for (int o = 0; o < nchain;) {
        int g = 1;
        if (ops[o].kind == LKPARS_OP_PARTIAL) {
            const int end = std::min(nchain, (o / CHAIN_MAX + 1) * CHAIN_MAX);
            while (g < gmax && o + g < end && ops[o + g].kind == LKPARS_OP_PARTIAL) {
                const lkpars_op &n = ops[o + g];
                bool hazard = false;
                for (int j = 0; j < g; j++) { const int d = ops[o + j].dst; hazard |= n.a == d || n.b == d || n.dst == d; }
                if (hazard) break;
                g++;
            }
            e->h_ops[o].c = g;
        }
        o += g;
    }
    PARS_CUDA(e, cudaMemcpyAsync(e->d_ops, e->h_ops, sizeof(lkpars_op) * nops, cudaMemcpyHostToDevice, e->stream));
    PARS_CUDA(e, cudaMemsetAsync(e->d_counts, 0, sizeof(unsigned int) * 2 * nops, e->stream));
    if (e->profiling) PARS_CUDA(e, cudaEventRecord(e->ev0, e->stream));
    int o = 0;
    do {
        const int n = nchain - o < CHAIN_MAX ? nchain - o : CHAIN_MAX;
        const int w = o + n >= nchain ? nwide : 0;
        switch (e->S) {
            case 4: launch_pars<4>(e, e->d_ops + o, e->d_counts + 2 * o, n, w); break;
            case 20: launch_pars<20>(e, e->d_ops + o, e->d_counts + 2 * o, n, w); break;
            default: launch_pars<0>(e, e->d_ops + o, e->d_counts + 2 * o, n, w); break;
        }
        o += n;
    } while (o < nchain);
    PARS_CUDA(e, cudaGetLastError());
    if (e->profiling) PARS_CUDA(e, cudaEventRecord(e->ev1, e->stream));
    PARS_CUDA(e, cudaMemcpyAsync(e->h_counts, e->d_counts, sizeof(unsigned int) * 2 * nops, cudaMemcpyDeviceToHost, e->stream));
    PARS_CUDA(e, cudaStreamSynchronize(e->stream));
    if (e->profiling) { float ms = 0; cudaEventElapsedTime(&ms, e->ev0, e->ev1); e->device_ms += ms; }
    memcpy(counts, e->h_counts, sizeof(unsigned int) * 2 * nops);
    return LKCUDA_OK;
}

extern "C" int lkpars_words(const lkpars_engine *e, int64_t *nwords) {
    if (!e || !nwords) return LKCUDA_ERR_ARG;
    *nwords = e->nw;
    return LKCUDA_OK;
}

extern "C" int lkpars_fetch_vector(lkpars_engine *e, int32_t vec, uint32_t *bits) {
    if (!e) return LKCUDA_ERR_ARG;
    if (!bits || vec < 0 || vec >= e->ntaxa + e->nvec) return pfail(e, LKCUDA_ERR_ARG, "bad vector %d", vec);
    PARS_CUDA(e, cudaSetDevice(e->device));
    const size_t n = (size_t)e->S * e->nw;
    PARS_CUDA(e, cudaMemcpyAsync(bits, e->d_bits + (size_t)vec * n, n * 4, cudaMemcpyDeviceToHost, e->stream));
    PARS_CUDA(e, cudaStreamSynchronize(e->stream));
    return LKCUDA_OK;
}

extern "C" int lkpars_launch_count(const lkpars_engine *e, uint64_t *launches, double *device_ms) {
    if (!e) return LKCUDA_ERR_ARG;
    if (launches) *launches = e->launches;
    if (device_ms) *device_ms = e->device_ms;
    return LKCUDA_OK;
}

extern "C" int lkpars_set_profiling(lkpars_engine *e, int32_t on) {
    if (!e) return LKCUDA_ERR_ARG;
    e->profiling = on != 0;
    return LKCUDA_OK;
}
