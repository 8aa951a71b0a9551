/*
 * lkcuda_partial.cu — partial-likelihood kernels with per-pattern underflow scaling (sm_100a, FP64).
 *
 * Replaces PhyloTree::computePartialLikelihoodEigenSIMD (phylokernel.h:71-483), its mixture and fused
 * mix-rate twins (phylokernelmixture.h:21-460, phylokernelmixrate.h:23-449) and the generic kernels
 * (phylotreesse.cpp:581-1010, 1350, 2039) for one node:
 *     L_dad[p][k][i] = sum_x inv_evec_k[i][x] * prod_child ( sum_j echild_k[x][j] * L_child[p][k][j] )
 *     echild_k[x][j] = evec_k[x][j] * exp(eval_k[j] * rate_k * len_child)          (phylokernel.h:163-181)
 * followed by the scaling test on lh_max = max_{k,i} |L_dad[p][k][i]| (phylokernel.h:380-392, 462-474).
 *
 * Launch structure: the host hands over a post-order schedule already grouped into tree levels; one
 * `prep_tables_kernel` launch builds every child's transition table (and, for tip children, the
 * per-state lookup table of phylokernel.h:191-232), then ONE launch per level covers all nodes of the
 * level (gridDim.y = nodes of the level) instead of one launch per node.
 *
 * Kernels
 *   partial_s4_kernel      DNA fast path (S = 4, K a power of two <= 32): one thread per (pattern, category),
 *                          each lane moves exactly one 32-byte sector per child and per result
 *                          (LDG.E.256 / STG.E.256), lh_max via warp shuffles. HBM-bound by design.
 *   partial_generic_kernel any S, any K, any number of children (multifurcations): pattern tiles staged
 *                          in shared memory; correctness fallback for shapes without a tuned kernel.
 */
#include "lkcuda_common.cuh"

using namespace lkc;

namespace lkc {

int ensure_tables(lkcuda_engine *e, long long slots) {
    if (slots <= e->table_slots_cap) return LKCUDA_OK;
    /* growing: wait for in-flight users of the old buffer */
    LK_CUDA(e, cudaStreamSynchronize(e->stream));
    if (e->d_tables) { cudaFree(e->d_tables); e->total_dev_bytes -= (uint64_t)e->table_slots_cap * e->v.table_slot * 8; }
    e->d_tables = NULL; e->table_slots_cap = 0;
    long long cap = slots < 64 ? 64 : slots + slots / 2;
    /* do not over-reserve for big mixture tables */
    if (cap * e->v.table_slot * 8 > (1ll << 30) && slots * e->v.table_slot * 8 <= (1ll << 30)) cap = (1ll << 30) / (e->v.table_slot * 8);
    if (cap < slots) cap = slots;
    LK_CUDA(e, cudaMalloc((void **)&e->d_tables, (size_t)cap * e->v.table_slot * 8));
    e->total_dev_bytes += (uint64_t)cap * e->v.table_slot * 8;
    e->table_slots_cap = cap;
    e->v.tables = e->d_tables;
    return LKCUDA_OK;
}

}  // namespace lkc

/* ------------------------------------------------------------------ table preparation ------------------------------------------------------------------ */

/* one CTA per child slot: ech[k][x][j] and, for a tip child, tab[state][k][x] (unknown-state row = 1) */
__global__ void __launch_bounds__(256) prep_tables_kernel(DevView v, const DevChild *__restrict__ children) {
    extern __shared__ __align__(32) double sh_ex[];            /* exp(eval*rate*len) per (k, j): B doubles */
    const int c = blockIdx.x;
    const DevChild ch = children[c];
    const int S = v.S, K = v.K, B = v.B;
    double *ech = v.tables + (long long)c * v.table_slot;
    double *tab = ech + (long long)K * S * S;
    for (int idx = threadIdx.x; idx < B; idx += blockDim.x) {
        int k = idx / S, j = idx - k * S;
        sh_ex[idx] = exp(v.eval[v.blk_eigen[k] * S + j] * (v.blk_rate[k] * ch.len));
    }
    __syncthreads();
    const int SS = S * S;
    for (int idx = threadIdx.x; idx < K * SS; idx += blockDim.x) {
        int k = idx / SS, r = idx - k * SS, j = r % S;
        ech[idx] = v.evec[(long long)v.blk_eigen[k] * SS + r] * sh_ex[k * S + j];
    }
    if (!LKCUDA_IS_TIP(ch.id)) return;
    __syncthreads();
    const int SU = v.SU;
    for (int idx = threadIdx.x; idx < (SU + 1) * B; idx += blockDim.x) {
        int st = idx / B, r = idx - st * B;
        double val = 1.0;
        if (st < SU) {
            int k = r / S, x = r - k * S;
            const double *row = ech + (long long)k * SS + (long long)x * S;
            const double *tip = v.tip + ((long long)st * v.E + v.blk_eigen[k]) * S;
            val = 0.0;
            for (int j = 0; j < S; j++) val += row[j] * tip[j];
        }
        tab[idx] = val;
    }
}

/* ------------------------------------------------------------------ DNA fast path ------------------------------------------------------------------ */

#define S4_PAD 18   /* doubles per (k) matrix row group in shared memory: 16 + 2 keeps LDS.128 conflict-free across k */

__global__ void __launch_bounds__(256) partial_s4_kernel(DevView v, const DevOp *__restrict__ ops, const DevChild *__restrict__ children,
                                                         int op_begin, int kshift) {
    extern __shared__ __align__(32) double sh[];
    const int K = 1 << kshift;
    const DevOp op = ops[op_begin + blockIdx.y];
    const DevChild cl = children[op.child0], cr = children[op.child0 + 1];
    const bool tipl = LKCUDA_IS_TIP(cl.id), tipr = LKCUDA_IS_TIP(cr.id);
    double *sEL = sh, *sER = sEL + K * S4_PAD, *sINV = sER + K * S4_PAD;
    double *sTL = sINV + K * S4_PAD;                 /* [(SU+1)][K][4] */
    double *sTR = sTL + (v.SU + 1) * K * 4;

    const long long unit = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long p = unit >> kshift;
    const int k = (int)(unit & (K - 1));
    const bool active = p < v.NP;

    /* issue the streaming loads first, stage the tables while they are in flight */
    d4 ll = {0, 0, 0, 0}, lr = {0, 0, 0, 0};
    int stl = 0, str = 0;
    if (active) {
        if (tipl) stl = tip_states(v, cl.id)[p]; else ll = ld_stream256(vec_lh(v, cl.id) + p * v.B + k * 4);
        if (tipr) str = tip_states(v, cr.id)[p]; else lr = ld_stream256(vec_lh(v, cr.id) + p * v.B + k * 4);
    }
    int sc = 0;
    if (active && k == 0) {
        if (!tipl) sc += vec_sc(v, cl.id)[p];
        if (!tipr) sc += vec_sc(v, cr.id)[p];
    }

    const double *gl = v.tables + (long long)op.child0 * v.table_slot;
    const double *gr = gl + v.table_slot;
    for (int idx = threadIdx.x; idx < K * 16; idx += blockDim.x) {
        int kk = idx >> 4, r = idx & 15;
        if (!tipl) sEL[kk * S4_PAD + r] = gl[idx];
        if (!tipr) sER[kk * S4_PAD + r] = gr[idx];
        sINV[kk * S4_PAD + r] = v.inv_evec[v.blk_eigen[kk] * 16 + r];
    }
    const int ntab = (v.SU + 1) * K * 4;
    if (tipl) for (int idx = threadIdx.x; idx < ntab; idx += blockDim.x) sTL[idx] = gl[K * 16 + idx];
    if (tipr) for (int idx = threadIdx.x; idx < ntab; idx += blockDim.x) sTR[idx] = gr[K * 16 + idx];
    __syncthreads();

    double vl[4], vr[4];
    if (tipl) {
        const double *t = sTL + (stl * K + k) * 4;
        vl[0] = t[0]; vl[1] = t[1]; vl[2] = t[2]; vl[3] = t[3];
    } else {
        const double *m = sEL + k * S4_PAD;
#pragma unroll
        for (int x = 0; x < 4; x++) vl[x] = m[x * 4 + 0] * ll.x + m[x * 4 + 1] * ll.y + m[x * 4 + 2] * ll.z + m[x * 4 + 3] * ll.w;
    }
    if (tipr) {
        const double *t = sTR + (str * K + k) * 4;
        vr[0] = t[0]; vr[1] = t[1]; vr[2] = t[2]; vr[3] = t[3];
    } else {
        const double *m = sER + k * S4_PAD;
#pragma unroll
        for (int x = 0; x < 4; x++) vr[x] = m[x * 4 + 0] * lr.x + m[x * 4 + 1] * lr.y + m[x * 4 + 2] * lr.z + m[x * 4 + 3] * lr.w;
    }
#pragma unroll
    for (int x = 0; x < 4; x++) vl[x] *= vr[x];
    const double *inv = sINV + k * S4_PAD;
    double o[4];
#pragma unroll
    for (int i = 0; i < 4; i++) o[i] = inv[i * 4 + 0] * vl[0] + inv[i * 4 + 1] * vl[1] + inv[i * 4 + 2] * vl[2] + inv[i * 4 + 3] * vl[3];

    if (!(tipl && tipr)) {   /* tip-tip never scales (phylokernel.h:246-281) */
        double lmax = fmax(fmax(fabs(o[0]), fabs(o[1])), fmax(fabs(o[2]), fabs(o[3])));
        for (int off = 1; off < K; off <<= 1) lmax = fmax(lmax, __shfl_xor_sync(0xffffffffu, lmax, off));
        if (active && lmax < LK_THR) {
            if ((v.flags & LKCUDA_FLAG_ZERO_GUARD) && lmax == 0.0) {
                const double *t = v.tip + ((long long)v.SU * v.E + v.blk_eigen[k]) * 4;
                o[0] = t[0]; o[1] = t[1]; o[2] = t[2]; o[3] = t[3];
                sc += 4;
            } else if ((v.flags & LKCUDA_FLAG_SCALE_IGNORES_INVAR) || v.invar[p] == 0.0) {
                o[0] *= LK_THR_INV; o[1] *= LK_THR_INV; o[2] *= LK_THR_INV; o[3] *= LK_THR_INV;
                sc += 1;
            }
        }
    }
    if (active) {
        d4 r = {o[0], o[1], o[2], o[3]};
        st_stream256(vec_lh_w(v, op.dst) + p * v.B + k * 4, r);
        if (k == 0) vec_sc_w(v, op.dst)[p] = (int16_t)sc;
    }
}

/* ------------------------------------------------------------------ generic path ------------------------------------------------------------------ */

#define GEN_TP 32        /* patterns per CTA tile (lane <-> pattern) */
#define GEN_THREADS 128

__device__ __forceinline__ void atomic_max_nonneg(unsigned long long *addr, double x) {
    /* the bit patterns of non-negative doubles order like the values; NaN never raises the maximum */
    if (x == x) atomicMax(addr, (unsigned long long)__double_as_longlong(x));
}

__global__ void __launch_bounds__(GEN_THREADS) partial_generic_kernel(DevView v, const DevOp *__restrict__ ops,
                                                                      const DevChild *__restrict__ children, int op_begin) {
    extern __shared__ __align__(32) double sh[];
    const int S = v.S, K = v.K, B = v.B, SS = S * S, LD = S + 1;
    const DevOp op = ops[op_begin + blockIdx.y];
    double *sEC = sh;                        /* child matrix of block k           [S*S]      */
    double *sINV = sEC + SS;                 /* inverse eigenvectors of block k   [S*S]      */
    double *sL = sINV + SS;                  /* child tile / result tile          [TP][S+1]  */
    double *sT = sL + GEN_TP * LD;           /* running product over children     [TP][S+1]  */
    unsigned long long *sMax = (unsigned long long *)(sT + GEN_TP * LD);   /* [TP] */
    int *sFlag = (int *)(sMax + GEN_TP);     /* [TP] */

    const long long p0 = (long long)blockIdx.x * GEN_TP;
    const int np = (int)min((long long)GEN_TP, v.NP - p0);
    const int tid = threadIdx.x, nt = blockDim.x;
    if (tid < GEN_TP) { sMax[tid] = 0ull; sFlag[tid] = 0; }
    bool all_tips = true;
    for (int c = 0; c < op.nchild; c++) all_tips = all_tips && LKCUDA_IS_TIP(children[op.child0 + c].id);
    const bool noscale = all_tips && op.nchild == 2;
    double *dst = vec_lh_w(v, op.dst) + p0 * B;
    int last_inv = -1;

    for (int k = 0; k < K; k++) {
        for (int idx = tid; idx < GEN_TP * S; idx += nt) { int x = idx / GEN_TP, p = idx - x * GEN_TP; sT[p * LD + x] = 1.0; }
        for (int c = 0; c < op.nchild; c++) {
            const DevChild ch = children[op.child0 + c];
            const double *tb = v.tables + (long long)(op.child0 + c) * v.table_slot;
            __syncthreads();
            if (LKCUDA_IS_TIP(ch.id)) {
                const uint8_t *st = tip_states(v, ch.id) + p0;
                const double *tab = tb + (long long)K * SS;
                for (int idx = tid; idx < GEN_TP * S; idx += nt) {
                    int x = idx / GEN_TP, p = idx - x * GEN_TP;
                    if (p < np) sT[p * LD + x] *= tab[(long long)st[p] * B + k * S + x];
                }
            } else {
                const double *src = vec_lh(v, ch.id) + p0 * B + k * S;
                for (int idx = tid; idx < SS; idx += nt) sEC[idx] = tb[(long long)k * SS + idx];
                for (int idx = tid; idx < GEN_TP * S; idx += nt) {
                    int p = idx / S, j = idx - p * S;
                    sL[p * LD + j] = p < np ? src[(long long)p * B + j] : 0.0;
                }
                __syncthreads();
                for (int idx = tid; idx < GEN_TP * S; idx += nt) {
                    int x = idx / GEN_TP, p = idx - x * GEN_TP;
                    const double *m = sEC + x * S, *l = sL + p * LD;
                    double acc = 0.0;
                    for (int j = 0; j < S; j++) acc += m[j] * l[j];
                    sT[p * LD + x] *= acc;
                }
            }
        }
        const int eig = v.blk_eigen[k];
        __syncthreads();
        if (eig != last_inv) {
            for (int idx = tid; idx < SS; idx += nt) sINV[idx] = v.inv_evec[(long long)eig * SS + idx];
            last_inv = eig;
            __syncthreads();
        }
        for (int idx = tid; idx < GEN_TP * S; idx += nt) {
            int i = idx / GEN_TP, p = idx - i * GEN_TP;
            const double *m = sINV + i * S, *t = sT + p * LD;
            double acc = 0.0;
            for (int x = 0; x < S; x++) acc += m[x] * t[x];
            sL[p * LD + i] = acc;
            atomic_max_nonneg(&sMax[p], fabs(acc));
        }
        __syncthreads();
        for (int idx = tid; idx < GEN_TP * S; idx += nt) {
            int p = idx / S, i = idx - p * S;
            if (p < np) dst[(long long)p * B + k * S + i] = sL[p * LD + i];
        }
    }
    __syncthreads();
    if (tid < np) {
        const long long p = p0 + tid;
        int sc = 0;
        for (int c = 0; c < op.nchild; c++) {
            int id = children[op.child0 + c].id;
            if (!LKCUDA_IS_TIP(id)) sc += vec_sc(v, id)[p];
        }
        double lmax = __longlong_as_double((long long)sMax[tid]);
        int flag = 0;
        if (!noscale && lmax < LK_THR) {
            if ((v.flags & LKCUDA_FLAG_ZERO_GUARD) && lmax == 0.0) { flag = 2; sc += 4; }
            else if ((v.flags & LKCUDA_FLAG_SCALE_IGNORES_INVAR) || v.invar[p] == 0.0) { flag = 1; sc += 1; }
        }
        sFlag[tid] = flag;
        vec_sc_w(v, op.dst)[p] = (int16_t)sc;
    }
    __syncthreads();
    bool any = false;
    for (int p = 0; p < np; p++) any = any || sFlag[p] != 0;
    if (any) {
        for (int idx = tid; idx < np * B; idx += nt) {
            int p = idx / B, r = idx - p * B;
            if (sFlag[p] == 1) dst[(long long)p * B + r] *= LK_THR_INV;
            else if (sFlag[p] == 2) { int k = r / S, i = r - k * S; dst[(long long)p * B + r] = v.tip[((long long)v.SU * v.E + v.blk_eigen[k]) * S + i]; }
        }
    }
}

/* ------------------------------------------------------------------ host side ------------------------------------------------------------------ */

static size_t generic_smem(const DevView &v) {
    return ((size_t)2 * v.S * v.S + (size_t)2 * GEN_TP * (v.S + 1)) * 8 + GEN_TP * 8 + GEN_TP * 4;
}

static bool use_s4(const DevView &v, int *kshift) {
    if (v.S != 4 || v.K > 32 || (v.K & (v.K - 1))) return false;
    int s = 0;
    while ((1 << s) < v.K) s++;
    *kshift = s;
    return true;
}

static int ring_acquire(lkcuda_engine *e, int nops, int nchild, int *slot_out) {
    int s = e->ring_next;
    e->ring_next = (s + 1) % lkcuda_engine::RING;
    if (!e->ring_ev[s]) LK_CUDA(e, cudaEventCreateWithFlags(&e->ring_ev[s], cudaEventDisableTiming));
    if (e->ring_used[s]) LK_CUDA(e, cudaEventSynchronize(e->ring_ev[s]));
    if (nops > e->ring_cap_ops[s]) {
        if (e->h_ops[s]) cudaFreeHost(e->h_ops[s]);
        cudaFree(e->d_ops[s]);
        int cap = nops < 256 ? 256 : nops * 2;
        LK_CUDA(e, cudaMallocHost((void **)&e->h_ops[s], sizeof(DevOp) * cap));
        LK_CUDA(e, cudaMalloc((void **)&e->d_ops[s], sizeof(DevOp) * cap));
        e->ring_cap_ops[s] = cap;
    }
    if (nchild > e->ring_cap_child[s]) {
        if (e->h_child[s]) cudaFreeHost(e->h_child[s]);
        cudaFree(e->d_child[s]);
        int cap = nchild < 512 ? 512 : nchild * 2;
        LK_CUDA(e, cudaMallocHost((void **)&e->h_child[s], sizeof(DevChild) * cap));
        LK_CUDA(e, cudaMalloc((void **)&e->d_child[s], sizeof(DevChild) * cap));
        e->ring_cap_child[s] = cap;
    }
    *slot_out = s;
    return LKCUDA_OK;
}

/* algorithmic bytes of one node: internal children are read, the result written, scale_num likewise */
static uint64_t op_bytes(const DevView &v, int ntip, int nint) {
    uint64_t b = (uint64_t)v.B * 8;
    return (uint64_t)v.NP * ((uint64_t)(nint + 1) * b + (uint64_t)ntip * 1 + (uint64_t)(nint + 1) * 2);
}

static int launch_ops(lkcuda_engine *e, int slot, int nops, int nchild_total, const int32_t *level_start, int nlevels, uint64_t *bytes_per_level) {
    const DevView &v = e->v;
    int rc = ensure_tables(e, nchild_total);
    if (rc) return rc;
    LK_CUDA(e, cudaMemcpyAsync(e->d_ops[slot], e->h_ops[slot], sizeof(DevOp) * nops, cudaMemcpyHostToDevice, e->stream));
    LK_CUDA(e, cudaMemcpyAsync(e->d_child[slot], e->h_child[slot], sizeof(DevChild) * nchild_total, cudaMemcpyHostToDevice, e->stream));
    LK_CUDA(e, cudaEventRecord(e->ring_ev[slot], e->stream));
    e->ring_used[slot] = true;
    if (v.NP == 0) return LKCUDA_OK;

    launch_begin(e, KIND_PREP);
    prep_tables_kernel<<<nchild_total, 256, (size_t)v.B * 8, e->stream>>>(e->v, e->d_child[slot]);
    rc = launch_end(e, KIND_PREP, 0, "prep_tables_kernel");
    if (rc) return rc;

    int kshift = 0;
    const bool s4 = use_s4(v, &kshift);
    size_t gsm = generic_smem(v);
    if (!s4 && gsm > 48 * 1024) {
        if (gsm > 227 * 1024) return fail(e, LKCUDA_ERR_ARG, "state space too large for the generic kernel (%d states)", v.S);
        LK_CUDA(e, cudaFuncSetAttribute(partial_generic_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gsm));
    }
    for (int l = 0; l < nlevels; l++) {
        int b = level_start[l], n = level_start[l + 1] - b;
        if (n <= 0) continue;
        if (n > 65535) return fail(e, LKCUDA_ERR_ARG, "more than 65535 nodes in one level");
        launch_begin(e, KIND_PARTIAL);
        if (s4) {
            size_t sm = ((size_t)3 * v.K * S4_PAD + (size_t)2 * (v.SU + 1) * v.K * 4) * 8;
            dim3 grid((unsigned)((v.NP * v.K + 255) / 256), (unsigned)n);
            partial_s4_kernel<<<grid, 256, sm, e->stream>>>(e->v, e->d_ops[slot], e->d_child[slot], b, kshift);
        } else {
            dim3 grid((unsigned)((v.NP + GEN_TP - 1) / GEN_TP), (unsigned)n);
            partial_generic_kernel<<<grid, GEN_THREADS, gsm, e->stream>>>(e->v, e->d_ops[slot], e->d_child[slot], b);
        }
        rc = launch_end(e, KIND_PARTIAL, bytes_per_level[l], "partial kernel");
        if (rc) return rc;
    }
    return LKCUDA_OK;
}

extern "C" {

int lkcuda_partial_batch(lkcuda_engine *e, const lkcuda_partial_op *ops, int32_t nops,
                         const int32_t *level_start, int32_t nlevels) {
    LK_ENTER(e);
    if (nops < 0 || (nops && !ops)) return fail(e, LKCUDA_ERR_ARG, "bad op list");
    if (!e->model_set || !e->states_set || !e->weights_set) return fail(e, LKCUDA_ERR_STATE, "lkcuda_partial_batch: tip states, pattern weights and model must be set first");
    if (nops == 0) return LKCUDA_OK;
    std::vector<int32_t> seq;
    if (level_start) {
        if (nlevels < 0 || level_start[0] != 0 || level_start[nlevels] != nops) return fail(e, LKCUDA_ERR_ARG, "bad level_start");
        for (int l = 0; l < nlevels; l++) if (level_start[l] > level_start[l + 1]) return fail(e, LKCUDA_ERR_ARG, "bad level_start");
    } else {
        seq.resize(nops + 1);
        for (int i = 0; i <= nops; i++) seq[i] = i;
        level_start = &seq[0]; nlevels = nops;
    }
    for (int i = 0; i < nops; i++) {
        const lkcuda_partial_op &op = ops[i];
        if (op.dst < 0 || op.dst >= e->cfg.nvec || !valid_end(e, op.left) || !valid_end(e, op.right))
            return fail(e, LKCUDA_ERR_ARG, "partial op %d: bad vector / taxon id", i);
        if (!LKCUDA_IS_TIP(op.left) && LKCUDA_IS_TIP(op.right))
            return fail(e, LKCUDA_ERR_ARG, "partial op %d: a leaf child must be 'left' (phylokernel.h:116-121)", i);
        if (op.dst == op.left || op.dst == op.right) return fail(e, LKCUDA_ERR_ARG, "partial op %d: destination aliases a child", i);
    }
    int slot;
    int rc = ring_acquire(e, nops, 2 * nops, &slot);
    if (rc) return rc;
    std::vector<uint64_t> lbytes(nlevels, 0);
    int l = 0;
    for (int i = 0; i < nops; i++) {
        while (l + 1 < nlevels && i >= level_start[l + 1]) l++;
        DevOp &d = e->h_ops[slot][i];
        d.dst = ops[i].dst; d.nchild = 2; d.child0 = 2 * i; d.pad = 0;
        DevChild &a = e->h_child[slot][2 * i], &b = e->h_child[slot][2 * i + 1];
        a.id = ops[i].left; a.pad = 0; a.len = ops[i].len_left;
        b.id = ops[i].right; b.pad = 0; b.len = ops[i].len_right;
        int ntip = (LKCUDA_IS_TIP(a.id) ? 1 : 0) + (LKCUDA_IS_TIP(b.id) ? 1 : 0);
        lbytes[l] += op_bytes(e->v, ntip, 2 - ntip);
    }
    e->theta_ready = false;
    return launch_ops(e, slot, nops, 2 * nops, level_start, nlevels, &lbytes[0]);
}

int lkcuda_partial_multi(lkcuda_engine *e, int32_t dst, int32_t nchild, const int32_t *child, const double *child_len) {
    LK_ENTER(e);
    if (!child || !child_len || nchild < 2 || dst < 0 || dst >= e->cfg.nvec) return fail(e, LKCUDA_ERR_ARG, "bad multifurcation arguments");
    if (!e->model_set || !e->states_set || !e->weights_set) return fail(e, LKCUDA_ERR_STATE, "lkcuda_partial_multi: inputs not set");
    for (int c = 0; c < nchild; c++) if (!valid_end(e, child[c]) || child[c] == dst) return fail(e, LKCUDA_ERR_ARG, "bad child");
    int slot;
    int rc = ring_acquire(e, 1, nchild, &slot);
    if (rc) return rc;
    DevOp &d = e->h_ops[slot][0];
    d.dst = dst; d.nchild = nchild; d.child0 = 0; d.pad = 0;
    int ntip = 0;
    for (int c = 0; c < nchild; c++) {
        DevChild &a = e->h_child[slot][c];
        a.id = child[c]; a.pad = 0; a.len = child_len[c];
        ntip += LKCUDA_IS_TIP(child[c]) ? 1 : 0;
    }
    uint64_t bytes = op_bytes(e->v, ntip, nchild - ntip);
    e->theta_ready = false;
    /* multifurcations always take the generic kernel: force it by launching here */
    const DevView &v = e->v;
    rc = ensure_tables(e, nchild);
    if (rc) return rc;
    LK_CUDA(e, cudaMemcpyAsync(e->d_ops[slot], e->h_ops[slot], sizeof(DevOp), cudaMemcpyHostToDevice, e->stream));
    LK_CUDA(e, cudaMemcpyAsync(e->d_child[slot], e->h_child[slot], sizeof(DevChild) * nchild, cudaMemcpyHostToDevice, e->stream));
    LK_CUDA(e, cudaEventRecord(e->ring_ev[slot], e->stream));
    e->ring_used[slot] = true;
    if (v.NP == 0) return LKCUDA_OK;
    launch_begin(e, KIND_PREP);
    prep_tables_kernel<<<nchild, 256, (size_t)v.B * 8, e->stream>>>(e->v, e->d_child[slot]);
    rc = launch_end(e, KIND_PREP, 0, "prep_tables_kernel");
    if (rc) return rc;
    size_t gsm = generic_smem(v);
    if (gsm > 227 * 1024) return fail(e, LKCUDA_ERR_ARG, "state space too large for the generic kernel (%d states)", v.S);
    if (gsm > 48 * 1024) LK_CUDA(e, cudaFuncSetAttribute(partial_generic_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gsm));
    launch_begin(e, KIND_PARTIAL);
    dim3 grid((unsigned)((v.NP + GEN_TP - 1) / GEN_TP), 1);
    partial_generic_kernel<<<grid, GEN_THREADS, gsm, e->stream>>>(e->v, e->d_ops[slot], e->d_child[slot], 0);
    return launch_end(e, KIND_PARTIAL, bytes, "partial_generic_kernel");
}

}  // extern "C"
