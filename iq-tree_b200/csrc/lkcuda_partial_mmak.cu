/*
 * lkcuda_partial_mmak.cu — codon (S = 64, the reference's 64-state layout) and wide-block protein
 * (S = 20 with more than 4 blocks: profile mixtures such as LG+C20+G4, K = 80) partial likelihoods on the FP64
 * tensor cores (sm_100a, mma.sync.aligned.m8n8k4.f64).
 *
 * Replaces PhyloTree::computePartialLikelihoodEigenSIMD<Vec4d,4,64> (phylokernel.h:71-483, instantiated at
 * phylotreeavx.cpp:107-129), computeMixturePartialLikelihoodEigenSIMD<Vec4d,4,20> (phylokernelmixture.h:21-460)
 * and the fused mix-rate twin (phylokernelmixrate.h:23-449):
 *     L_dad[p][k][i] = sum_x inv_evec_k[i][x] * (sum_j eleft_k[x][j] L_left[p][k][j]) * (sum_j eright_k[x][j] L_right[p][k][j])
 *
 * Same whole-traversal structure and the same fragment algebra as lkcuda_partial_mma.cu (the C fragment of one
 * contraction is the A fragment of the next; only the matrices are permuted into B-fragment order), with one
 * difference forced by size: a fragment-ordered 64x64 matrix is 32 KB, so the operands of a node cannot all sit in
 * shared memory at once.  The ring therefore streams one (node, block k) pair per stage: the two children's
 * matrices (or tip lookup tables) of block k, and - when the model has more than one eigensystem - the inverse
 * eigenvector matrix of block k.  The scaling test needs lh_max over all blocks of a pattern, so it runs after the
 * last k; the rare rescale re-reads what the lane itself wrote (as in the protein kernel).
 *
 * These shapes are FP64-pipe bound (codon: 16 flop/B, protein mixtures: 5 flop/B before any on-chip reuse), so
 * children are re-read from their arena vectors (L2 / HBM) unless a shared-memory slot fits (codon, K = 1).
 */
#include "lkcuda_traverse.cuh"

using namespace lkc;

#define MK_MAX_STAGES 8
#define MK_THR_HI 0x2FF00000    /* high word of 2^-256 */

template <int S> struct MkShape {
    static constexpr int NT = (S + 7) / 8;              /* 8-wide tiles */
    static constexpr int ROW = NT * 8;                  /* padded row of a tip table */
    static constexpr int FRAGS = NT * NT * 2;           /* B fragments of one matrix: (rt, ct, c) */
    static constexpr int FRAG_DOUBLES = FRAGS * 32;
    static constexpr int LASTW = S - 8 * (NT - 1);      /* valid entries of the last tile (even) */
    /* HALF (S % 8 == 4, protein): the 4 entries of the last tile sit in its EVEN slots (slot 2q <-> entry 8(NT-1) + q), so that its
       odd slots form one all-padding k-step that every contraction skips: 5 x 3 = 15 DMMAs per 20x20 product instead of 18 */
    static constexpr bool HALF = LASTW == 4;
    /* fragment slot (tile t, column j) -> state / eigen index, -1 = padding */
    __host__ __device__ static constexpr int index_of(int t, int j) {
        return t < NT - 1 ? 8 * t + j : (HALF ? ((j & 1) ? -1 : 8 * t + (j >> 1)) : (j < LASTW ? 8 * t + j : -1));
    }
    /* consumer warps per CTA (one 8-pattern group each): register budget allows 16 for the 3-tile protein shape */
    static constexpr int CW = S <= 24 ? 16 : 8;
    static constexpr int TP = CW * 8;                   /* patterns per round of a CTA */
    static constexpr int THREADS = CW * 32 + 32;        /* + the producer warp */
};

/* ------------------------------------------------------------------ operand preparation ------------------------------------------------------------------ */

/* grid (child slot c, block k): fragment-ordered matrix evec*exp(eval*rate*len) at [k][FRAGS][32] and, for a tip child,
   the lookup table tab[k][state][ROW] = sum_j ech_k[x][j] * tip_partial_lh[state][j] (x >= S: 0; unknown state: 1) */
template <int S>
__global__ void __launch_bounds__(256) prep_fragk_kernel(DevView v, const DevChild *__restrict__ children) {
    typedef MkShape<S> Sh;
    extern __shared__ __align__(16) double sh_m[];      /* S*S */
    const int c = blockIdx.x, k = blockIdx.y, K = v.K;
    const DevChild ch = children[c];
    double *out = v.tables + (long long)c * v.table_slot + (long long)k * Sh::FRAG_DOUBLES;
    double *tab = v.tables + (long long)c * v.table_slot + (long long)K * Sh::FRAG_DOUBLES + (long long)k * (v.SU + 1) * Sh::ROW;
    const double *evec = v.evec + (long long)v.blk_eigen[k] * S * S, *eval = v.eval + v.blk_eigen[k] * S;
    const double rl = v.blk_rate[k] * child_len(v, ch);
    for (int idx = threadIdx.x; idx < S * S; idx += blockDim.x) sh_m[idx] = evec[idx] * exp(eval[idx % S] * rl);
    __syncthreads();
    for (int idx = threadIdx.x; idx < Sh::FRAG_DOUBLES; idx += blockDim.x) {
        int lane = idx & 31, f = idx >> 5, cc = f & 1, ct = (f >> 1) % Sh::NT, rt = (f >> 1) / Sh::NT;
        int r = Sh::index_of(rt, lane >> 2), col = Sh::index_of(ct, 2 * (lane & 3) + cc);
        out[idx] = (r >= 0 && col >= 0) ? sh_m[r * S + col] : 0.0;
    }
    if (!LKCUDA_IS_TIP(ch.id)) return;
    for (int idx = threadIdx.x; idx < (v.SU + 1) * Sh::ROW; idx += blockDim.x) {
        int st = idx / Sh::ROW, pos = idx - st * Sh::ROW, x = Sh::index_of(pos >> 3, pos & 7);   /* rows are stored in fragment-slot order */
        double val = 0.0;
        if (x >= 0) {
            if (st == v.SU) val = 1.0;
            else {
                const double *tip = v.tip + ((long long)st * v.E + v.blk_eigen[k]) * S;
                for (int j = 0; j < S; j++) val += sh_m[x * S + j] * tip[j];
            }
        }
        tab[idx] = val;
    }
}

/* fragment-ordered inverse eigenvectors, one matrix per EIGENSYSTEM: out[e][FRAGS][32] */
template <int S>
__global__ void __launch_bounds__(256) prep_inv_fragk_kernel(DevView v, double *out) {
    typedef MkShape<S> Sh;
    const int eg = blockIdx.x;
    const double *inv = v.inv_evec + (long long)eg * S * S;
    for (int idx = threadIdx.x; idx < Sh::FRAG_DOUBLES; idx += blockDim.x) {
        int lane = idx & 31, f = idx >> 5, cc = f & 1, ct = (f >> 1) % Sh::NT, rt = (f >> 1) / Sh::NT;
        int r = Sh::index_of(rt, lane >> 2), col = Sh::index_of(ct, 2 * (lane & 3) + cc);
        out[(long long)eg * Sh::FRAG_DOUBLES + idx] = (r >= 0 && col >= 0) ? inv[r * S + col] : 0.0;
    }
}

/* ------------------------------------------------------------------ the kernel ------------------------------------------------------------------ */

__device__ __forceinline__ void mk_dmma(double &d0, double &d1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

/* acc[rt][c] = sum over k-steps (ct, cc) of X[ct][cc] * frag[rt][ct][cc]   (frag: this lane's column of a fragment-ordered matrix) */
template <int NT, bool HALF>
__device__ __forceinline__ void mk_contract(double (&acc)[NT][2], const double (&X)[NT][2], const double *frag) {
#pragma unroll
    for (int rt = 0; rt < NT; rt++) { acc[rt][0] = 0.0; acc[rt][1] = 0.0; }
    /* ct outermost: the NT accumulator chains advance together, so consecutive DMMAs are independent */
#pragma unroll
    for (int ct = 0; ct < NT; ct++) {
#pragma unroll
        for (int cc = 0; cc < 2; cc++) {
            if (HALF && ct == NT - 1 && cc == 1) continue;      /* all-padding k-step */
#pragma unroll
            for (int rt = 0; rt < NT; rt++) mk_dmma(acc[rt][0], acc[rt][1], X[ct][cc], frag[((rt * NT + ct) * 2 + cc) * 32]);
        }
    }
}

__device__ __forceinline__ double2 mk_ld128(const double *p) {   /* may have been written earlier by this very lane */
    double2 r;
    asm volatile("ld.global.L1::no_allocate.v2.f64 {%0,%1}, [%2];" : "=d"(r.x), "=d"(r.y) : "l"(p) : "memory");
    return r;
}
__device__ __forceinline__ void mk_st128(double *p, double x, double y) {
    asm volatile("st.global.L1::no_allocate.v2.f64 [%0], {%1,%2};" :: "l"(p), "d"(x), "d"(y) : "memory");
}

/* lane (g, q) <-> the entries of tile nt it holds in a C / A fragment: slots 2q, 2q + 1 (see MkShape::index_of) */
template <int S> __device__ __forceinline__ bool mk_lane_has(int nt, int q) {
    typedef MkShape<S> Sh;
    return nt < Sh::NT - 1 || Sh::HALF || q < Sh::LASTW / 2;
}
__device__ __forceinline__ double mk_ld64(const double *p) {
    double r;
    asm volatile("ld.global.L1::no_allocate.f64 %0, [%1];" : "=d"(r) : "l"(p) : "memory");
    return r;
}
/* blk: start of block k of one pattern in an arena vector (S doubles, reference layout) */
template <int S> __device__ __forceinline__ double2 mk_g_ld(const double *blk, int nt, int q) {
    typedef MkShape<S> Sh;
    if (Sh::HALF && nt == Sh::NT - 1) return make_double2(mk_ld64(blk + 8 * nt + q), 0.0);
    return mk_lane_has<S>(nt, q) ? mk_ld128(blk + 8 * nt + 2 * q) : make_double2(0.0, 0.0);
}
template <int S> __device__ __forceinline__ void mk_g_st(double *blk, int nt, int q, double x, double y) {
    typedef MkShape<S> Sh;
    if (Sh::HALF && nt == Sh::NT - 1) asm volatile("st.global.L1::no_allocate.f64 [%0], %1;" :: "l"(blk + 8 * nt + q), "d"(x) : "memory");
    else if (mk_lane_has<S>(nt, q)) mk_st128(blk + 8 * nt + 2 * q, x, y);
}
/* shared-memory slot (doubles): [k][ tiles 0..NT-2: TP*8 each | last tile: TP*LASTW ] */
template <int S> __device__ __forceinline__ double2 mk_s_ld(const double *slot, int k, int nt, int pat, int q) {
    typedef MkShape<S> Sh;
    const double *b = slot + k * (Sh::TP * S) + nt * (Sh::TP * 8);
    if (nt < Sh::NT - 1) return *reinterpret_cast<const double2 *>(b + pat * 8 + 2 * q);
    if (Sh::HALF) return make_double2(b[pat * 4 + q], 0.0);
    return mk_lane_has<S>(nt, q) ? *reinterpret_cast<const double2 *>(b + pat * Sh::LASTW + 2 * q) : make_double2(0.0, 0.0);
}
template <int S> __device__ __forceinline__ void mk_s_st(double *slot, int k, int nt, int pat, int q, double x, double y) {
    typedef MkShape<S> Sh;
    double *b = slot + k * (Sh::TP * S) + nt * (Sh::TP * 8);
    if (nt < Sh::NT - 1) *reinterpret_cast<double2 *>(b + pat * 8 + 2 * q) = make_double2(x, y);
    else if (Sh::HALF) b[pat * 4 + q] = x;
    else if (mk_lane_has<S>(nt, q)) *reinterpret_cast<double2 *>(b + pat * Sh::LASTW + 2 * q) = make_double2(x, y);
}
template <int S>
__device__ __forceinline__ void mk_load_x(double (&X)[MkShape<S>::NT][2], const double *blk, int q) {
#pragma unroll
    for (int nt = 0; nt < MkShape<S>::NT; nt++) { const double2 a = mk_g_ld<S>(blk, nt, q); X[nt][0] = a.x; X[nt][1] = a.y; }
}

/* ring stage: descriptor (64) | child 0 block | child 1 block | [inverse eigenvectors of block k] (cbs doubles each) | states 0 | states 1 */
__host__ __device__ __forceinline__ int mk_stage_bytes(int cbs, int nblk, int tp) { return 64 + nblk * cbs * 8 + 2 * (tp + 16); }

/* the 8-pattern groups [g0, g1) of CTA b: an even split of all groups over the grid (the last round of a CTA may run with idle warps) */
__device__ __forceinline__ void mk_cta_groups(long long ngroups, long long &g0, long long &g1) {
    g0 = ngroups * blockIdx.x / gridDim.x; g1 = ngroups * (blockIdx.x + 1) / gridDim.x;
}

/* KONE: one block per pattern (codon without rate heterogeneity): the previous node's result stays in registers
   (TRV_SLOT_PREV) like in the DNA kernel.  Otherwise the next block's child entries are loaded while the current
   block's contractions run. */
template <int S, bool KONE>
__global__ void __launch_bounds__(MkShape<S>::THREADS, 1) traverse_mmak_kernel(DevView v, const TravOp *__restrict__ ops, int nops, int nslots, int cbs,
                                                                              int inv_streamed, int nstages, const double *__restrict__ inv_frags, long long ngroups) {
    typedef MkShape<S> Sh;
    constexpr int NT = Sh::NT, FD = Sh::FRAG_DOUBLES, ROW = Sh::ROW, CW = Sh::CW, TP = Sh::TP, THREADS = Sh::THREADS;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int K = KONE ? 1 : v.K, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nblk = inv_streamed ? 3 : 2;
    const int stage_bytes = mk_stage_bytes(cbs, nblk, TP);
    unsigned char *ring = smem_raw;                                                         /* [nstages][stage_bytes] */
    double *sINV = reinterpret_cast<double *>(ring + nstages * stage_bytes);                /* [FD] when resident (one eigensystem) */
    double *sl = sINV + (inv_streamed ? 0 : FD);                                            /* [nslots][K*TP*S] */
    int *ssc = reinterpret_cast<int *>(sl + (size_t)nslots * K * TP * S);                   /* [nslots + 1][TP] */
    uint64_t *full = reinterpret_cast<uint64_t *>(ssc + (nslots + 1) * TP), *empty = full + MK_MAX_STAGES;
    if (!inv_streamed) for (int idx = tid; idx < FD; idx += THREADS) sINV[idx] = inv_frags[idx];
    if (tid == 0) {
        for (int s = 0; s < nstages; s++) { mbar_init(&full[s], 1); mbar_init(&empty[s], CW); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const int states_off = 64 + nblk * cbs * 8;
    long long g0, g1;
    mk_cta_groups(ngroups, g0, g1);

    if (warp == CW) {
        /* ------------------------------ producer: one stage per (node, block k) ------------------------------ */
        if (lane == 0) {
            uint32_t s = 0, ph = 0;
            const uint32_t mat_bytes = (uint32_t)FD * 8, tab_bytes = (uint32_t)(v.SU + 1) * ROW * 8;
            for (long long gr = g0; gr < g1; gr += CW) {
                const long long p0 = gr * 8, a0 = p0 & ~15ll;               /* bulk copies start 16-byte aligned */
                for (int i = 0; i < nops; i++) {
                    const int src0 = ops[i].src0, src1 = ops[i].src1;
                    const bool t0 = LKCUDA_IS_TIP(src0), t1 = LKCUDA_IS_TIP(src1);
                    const uint32_t b0 = t0 ? tab_bytes : mat_bytes, b1 = t1 ? tab_bytes : mat_bytes;
                    const double *s0 = v.tables + (long long)(2 * i) * v.table_slot, *s1 = s0 + v.table_slot;
                    for (int k = 0; k < K; k++) {
                        unsigned char *st = ring + s * stage_bytes;
                        const double *q0 = t0 ? s0 + (long long)K * FD + (long long)k * (v.SU + 1) * ROW : s0 + (long long)k * FD;
                        const double *q1 = t1 ? s1 + (long long)K * FD + (long long)k * (v.SU + 1) * ROW : s1 + (long long)k * FD;
                        mbar_wait_backoff(&empty[s], ph ^ 1);
                        mbar_expect_tx(&full[s], 64 + b0 + b1 + (inv_streamed ? mat_bytes : 0) + (t0 ? TP + 16 : 0) + (t1 ? TP + 16 : 0));
                        bulk_g2s(st, ops + i, 64, &full[s]);
                        bulk_g2s(st + 64, q0, b0, &full[s]);
                        bulk_g2s(st + 64 + cbs * 8, q1, b1, &full[s]);
                        if (inv_streamed) bulk_g2s(st + 64 + 2 * cbs * 8, inv_frags + (long long)v.blk_eigen[k] * FD, mat_bytes, &full[s]);
                        if (t0) bulk_g2s(st + states_off, tip_states(v, src0) + a0, TP + 16, &full[s]);
                        if (t1) bulk_g2s(st + states_off + TP + 16, tip_states(v, src1) + a0, TP + 16, &full[s]);
                        if (++s == (uint32_t)nstages) { s = 0; ph ^= 1; }
                    }
                }
            }
        }
        return;
    }

    /* ------------------------------ consumers: warp w owns group gr + w of the round ------------------------------ */
    const int g = lane >> 2, q = lane & 3, pat = warp * 8 + g;
    const bool zero_guard = (v.flags & LKCUDA_FLAG_ZERO_GUARD) != 0, ignore_invar = (v.flags & LKCUDA_FLAG_SCALE_IGNORES_INVAR) != 0;
    const int B = v.B;
    uint32_t s = 0, ph = 0;

    for (long long gr = g0; gr < g1; gr += CW) {
        long long p = (gr + warp) * 8 + g;
        const bool valid = gr + warp < g1 && p < v.NP;
        if (!valid) p = v.NP - 1;                   /* idle lanes shadow the last pattern; only their stores are masked */
        const int sdelta = (int)((gr * 8) & 15);
        double pv[NT][2];                           /* KONE: result of the previous node */
        int pvsc = 0;
#pragma unroll
        for (int nt = 0; nt < NT; nt++) { pv[nt][0] = 0.0; pv[nt][1] = 0.0; }
        if (gr + warp >= g1) {
            /* this warp has no group in the CTA's last round: keep the ring protocol going, leave the tensor pipe to the others */
            for (int n = nops * K; n > 0; n--) {
                mbar_wait(&full[s], ph);
                if (lane == 0) mbar_arrive(&empty[s]);
                if (++s == (uint32_t)nstages) { s = 0; ph ^= 1; }
            }
            continue;
        }
        for (int i = 0; i < nops; i++) {
            int sc = 0, mx = 0, nz = 0, dst_slot = -1;
            bool tip0 = false, tip1 = false;
            double *dst_lh = NULL, *dslot = NULL;
            int16_t *dst_sc = NULL;
            double X0[NT][2], X1[NT][2];
            for (int k = 0; k < K; k++) {
                mbar_wait(&full[s], ph);
                const unsigned char *stage = ring + s * stage_bytes;
                const TravOp *op = reinterpret_cast<const TravOp *>(stage);
                const int src0 = op->src0, src1 = op->src1, slot0 = op->slot0, slot1 = op->slot1;
                tip0 = LKCUDA_IS_TIP(src0); tip1 = LKCUDA_IS_TIP(src1);
                const bool arena0 = !tip0 && slot0 < 0, arena1 = !tip1 && slot1 < 0 && slot1 != TRV_SLOT_PREV;
                const double *blk0 = reinterpret_cast<const double *>(stage + 64), *blk1 = blk0 + cbs;
                const double *invf = inv_streamed ? blk1 + cbs : sINV;
                const double *gp0 = op->src0_lh + p * B, *gp1 = op->src1_lh + p * B;
                if (k == 0) {
                    if (arena0) mk_load_x<S>(X0, gp0, q);
                    if (arena1) mk_load_x<S>(X1, gp1, q);
                    dst_slot = op->dst_slot; dst_lh = op->dst_lh; dst_sc = op->dst_sc;
                    dslot = dst_slot >= 0 ? sl + (size_t)dst_slot * K * TP * S : NULL;
                    if (!tip0) sc += slot0 >= 0 ? ssc[slot0 * TP + pat] : (int)vec_sc(v, src0)[p];
                    if (!tip1) sc += slot1 == TRV_SLOT_PREV ? pvsc : (slot1 >= 0 ? ssc[slot1 * TP + pat] : (int)vec_sc(v, src1)[p]);
                }
                double t[NT][2];
                /* ---- child 0 ---- */
                if (tip0) {
                    const int st0 = stage[states_off + sdelta + pat];
                    const double *row = blk0 + (size_t)st0 * ROW + 2 * q;
#pragma unroll
                    for (int nt = 0; nt < NT; nt++) { const double2 a = *reinterpret_cast<const double2 *>(row + 8 * nt); t[nt][0] = a.x; t[nt][1] = a.y; }
                } else {
                    if (slot0 >= 0) {
                        const double *ps = sl + (size_t)slot0 * K * TP * S;
#pragma unroll
                        for (int nt = 0; nt < NT; nt++) { const double2 a = mk_s_ld<S>(ps, k, nt, pat, q); X0[nt][0] = a.x; X0[nt][1] = a.y; }
                    }
                    mk_contract<NT, Sh::HALF>(t, X0, blk0 + lane);
                    if (!KONE && arena0 && k + 1 < K) mk_load_x<S>(X0, gp0 + (k + 1) * S, q);     /* next block, in flight during the other contractions */
                }
                /* ---- child 1 ---- */
                if (tip1) {
                    const int st1 = stage[states_off + TP + 16 + sdelta + pat];
                    const double *row = blk1 + (size_t)st1 * ROW + 2 * q;
#pragma unroll
                    for (int nt = 0; nt < NT; nt++) { const double2 a = *reinterpret_cast<const double2 *>(row + 8 * nt); t[nt][0] *= a.x; t[nt][1] *= a.y; }
                } else {
                    double u[NT][2];
                    if (KONE && slot1 == TRV_SLOT_PREV) {
                        mk_contract<NT, Sh::HALF>(u, pv, blk1 + lane);
                    } else {
                        if (slot1 >= 0) {
                            const double *ps = sl + (size_t)slot1 * K * TP * S;
#pragma unroll
                        for (int nt = 0; nt < NT; nt++) { const double2 a = mk_s_ld<S>(ps, k, nt, pat, q); X1[nt][0] = a.x; X1[nt][1] = a.y; }
                        }
                        mk_contract<NT, Sh::HALF>(u, X1, blk1 + lane);
                        if (!KONE && arena1 && k + 1 < K) mk_load_x<S>(X1, gp1 + (k + 1) * S, q);
                    }
#pragma unroll
                    for (int nt = 0; nt < NT; nt++) { t[nt][0] *= u[nt][0]; t[nt][1] *= u[nt][1]; }
                }
                /* ---- back to state space ---- */
                double o[NT][2];
                mk_contract<NT, Sh::HALF>(o, t, invf + lane);
                /* every operand of this (node, k) has been consumed: hand the stage back */
                __syncwarp();
                if (lane == 0) mbar_arrive(&empty[s]);
                if (++s == (uint32_t)nstages) { s = 0; ph ^= 1; }
                double *gd = dst_lh + p * B + k * S;
#pragma unroll
                for (int nt = 0; nt < NT; nt++) {
                    if (KONE) { pv[nt][0] = o[nt][0]; pv[nt][1] = o[nt][1]; }
                    if (mk_lane_has<S>(nt, q)) {
                        const int h0 = __double2hiint(o[nt][0]) & 0x7fffffff, h1 = __double2hiint(o[nt][1]) & 0x7fffffff;
                        mx = max(mx, max(h0, h1));
                        nz |= h0 | h1 | __double2loint(o[nt][0]) | __double2loint(o[nt][1]);
                        if (valid) mk_g_st<S>(gd, nt, q, o[nt][0], o[nt][1]);
                        if (dslot) mk_s_st<S>(dslot, k, nt, pat, q, o[nt][0], o[nt][1]);
                    }
                }
            }
            /* ---- scaling (phylokernel.h:380-392, 462-474): lh_max over all k and states of the pattern = the quad's maximum ---- */
            mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, 1)); mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
            nz |= __shfl_xor_sync(0xffffffffu, nz, 1); nz |= __shfl_xor_sync(0xffffffffu, nz, 2);
            const bool tiny = !(tip0 && tip1) && mx < MK_THR_HI;
            if (__any_sync(0xffffffffu, tiny)) {
                /* rare: re-read what this lane just wrote, rescale (or refill), write again */
                const bool refill = tiny && zero_guard && nz == 0;
                const bool rescale = tiny && !refill && (ignore_invar || v.invar[p] == 0.0);
                if (refill || rescale) {
                    for (int k = 0; k < K; k++) {
                        double *gd = dst_lh + p * B + k * S;
                        const double *tu = v.tip + ((long long)v.SU * v.E + v.blk_eigen[k]) * S;
#pragma unroll
                        for (int nt = 0; nt < NT; nt++) {
                            if (mk_lane_has<S>(nt, q)) {
                                double2 a;
                                if (refill) {
                                    const int i0 = Sh::index_of(nt, 2 * q), i1 = Sh::index_of(nt, 2 * q + 1);
                                    a = make_double2(i0 >= 0 ? tu[i0] : 0.0, i1 >= 0 ? tu[i1] : 0.0);
                                } else {
                                    if (KONE) a = make_double2(pv[nt][0], pv[nt][1]);
                                    else if (dslot) a = mk_s_ld<S>(dslot, k, nt, pat, q);
                                    else a = mk_g_ld<S>(gd, nt, q);
                                    a.x *= LK_THR_INV; a.y *= LK_THR_INV;
                                }
                                if (KONE) { pv[nt][0] = a.x; pv[nt][1] = a.y; }
                                if (valid) mk_g_st<S>(gd, nt, q, a.x, a.y);
                                if (dslot) mk_s_st<S>(dslot, k, nt, pat, q, a.x, a.y);
                            }
                        }
                    }
                    sc += refill ? 4 : 1;
                }
            }
            pvsc = sc;
            if (q == 0) {
                if (valid) dst_sc[p] = (int16_t)sc;
                if (dst_slot >= 0) ssc[dst_slot * TP + pat] = sc;
            }
            /* a parked result is read back by the other lanes of the quad (scale count) and, when spilled, from the arena */
            __syncwarp();
            if (dst_slot == -2) { __threadfence_block(); __syncwarp(); }
        }
    }
}

/* ------------------------------------------------------------------ host side ------------------------------------------------------------------ */

namespace lkc {

static bool mmak_shape(const DevView &v) { return v.S == 64 || (v.S == 20 && v.K > 4); }

/* doubles one child needs in the table area: K fragment-ordered matrices + K tip lookup tables */
long long mmak_table_slot(const DevView &v) {
    if (!mmak_shape(v)) return 0;
    const long long NT = (v.S + 7) / 8, FD = NT * NT * 2 * 32, ROW = NT * 8;
    return (long long)v.K * FD + (long long)v.K * (v.SU + 1) * ROW;
}

bool use_traverse_mmak(const lkcuda_engine *e) { return mmak_shape(e->v) && !e->trav_disable_mma; }

template <int S>
static int launch_mmak(lkcuda_engine *e, const lkcuda_partial_op *ops, int nops) {
    typedef MkShape<S> Sh;
    const DevView &v = e->v;
    const int K = v.K, TP = Sh::TP;
    const bool kone = K == 1;
    const int cbs_mat = Sh::FRAG_DOUBLES, cbs_tab = (v.SU + 1) * Sh::ROW, cbs = cbs_mat > cbs_tab ? cbs_mat : cbs_tab;
    if (mmak_table_slot(v) > v.table_slot) return fail(e, LKCUDA_ERR_STATE, "table slot too small for the MMA layout");
    const int inv_streamed = v.E > 1 ? 1 : 0;
    const size_t stage_b = (size_t)mk_stage_bytes(cbs, inv_streamed ? 3 : 2, TP);
    const size_t slot_bytes = (size_t)K * TP * S * 8 + (size_t)TP * 4;
    const size_t other = (inv_streamed ? 0 : (size_t)Sh::FRAG_DOUBLES * 8) + (size_t)TP * 4 + 2 * MK_MAX_STAGES * 8 + 128;
    /* ring depth: up to 8 stages (small protein stages), at least 2 (codon: 67 KB per stage) */
    int nstages = MK_MAX_STAGES;
    while (nstages > 2 && nstages * stage_b + other > 160 * 1024) nstages--;
    const size_t fixed = nstages * stage_b + other;
    if (fixed > 227 * 1024) return fail(e, LKCUDA_ERR_ARG, "state space too large for the MMA traversal kernel");
    int maxslots = (int)((227 * 1024 - fixed) / slot_bytes);
    if (e->trav_maxslots >= 0 && e->trav_maxslots < maxslots) maxslots = e->trav_maxslots;
    TravSchedule sch;
    int rc = schedule_traverse(e, ops, nops, maxslots, kone, sch);
    if (rc) return rc;
    int slot;
    rc = ring_acquire(e, nops, 2 * nops, &slot);
    if (rc) return rc;
    TravOp *h = reinterpret_cast<TravOp *>(e->h_ops[slot]);
    uint64_t bytes = 0;
    for (int j = 0; j < nops; j++) {
        const TravNode &n = sch.nodes[j];
        TravOp &t = h[j];
        t.dst = ops[n.op].dst; t.dst_slot = n.dst_slot;
        t.dst_lh = e->v.arena + (size_t)t.dst * v.vec_stride; t.dst_sc = e->v.scale + (size_t)t.dst * v.sc_stride;
        t.src0 = n.src0; t.slot0 = n.slot0;
        t.src1 = n.src1; t.slot1 = n.slot1;
        t.src0_lh = n.src0 >= 0 ? e->v.arena + (size_t)n.src0 * v.vec_stride : NULL;
        t.src1_lh = n.src1 >= 0 ? e->v.arena + (size_t)n.src1 * v.vec_stride : NULL;
        t.pad[0] = t.pad[1] = 0;
        DevChild &a = e->h_child[slot][2 * j], &b = e->h_child[slot][2 * j + 1];
        a.id = n.src0; a.pad = n.ref0; a.len = n.len0;
        b.id = n.src1; b.pad = n.ref1; b.len = n.len1;
        int ntip = (LKCUDA_IS_TIP(a.id) ? 1 : 0) + (LKCUDA_IS_TIP(b.id) ? 1 : 0);
        bytes += op_bytes(v, ntip, 2 - ntip);
        e->kind_flops[KIND_PARTIAL] += op_flops(v, ntip, 2 - ntip);
    }
    rc = ensure_tables(e, 2 * nops);
    if (rc) return rc;
    if (!e->d_inv_frags) LK_CUDA(e, cudaMalloc((void **)&e->d_inv_frags, (size_t)v.E * Sh::FRAG_DOUBLES * 8));
    TravOp *d = reinterpret_cast<TravOp *>(e->d_ops[slot]);
    LK_CUDA(e, cudaMemcpyAsync(d, h, sizeof(TravOp) * nops, cudaMemcpyHostToDevice, e->stream));
    LK_CUDA(e, cudaMemcpyAsync(e->d_child[slot], e->h_child[slot], sizeof(DevChild) * 2 * nops, cudaMemcpyHostToDevice, e->stream));
    LK_CUDA(e, cudaEventRecord(e->ring_ev[slot], e->stream));
    e->ring_used[slot] = true;
    if (v.NP == 0) return LKCUDA_OK;
    rc = states_barrier(e);
    if (rc) return rc;

    if ((size_t)S * S * 8 > 48 * 1024) return fail(e, LKCUDA_ERR_ARG, "state space too large for prep_fragk_kernel");
    launch_begin(e, KIND_PREP);
    prep_fragk_kernel<S><<<dim3(2 * nops, K), 256, (size_t)S * S * 8, e->stream>>>(e->v, e->d_child[slot]);
    rc = launch_end(e, KIND_PREP, 0, "prep_fragk_kernel");
    if (rc) return rc;
    launch_begin(e, KIND_PREP);
    prep_inv_fragk_kernel<S><<<v.E, 256, 0, e->stream>>>(e->v, e->d_inv_frags);
    rc = launch_end(e, KIND_PREP, 0, "prep_inv_fragk_kernel");
    if (rc) return rc;

    const long long ngroups = (v.NP + 7) / 8;
    const size_t sm = (size_t)sch.nslots * slot_bytes + fixed;
    if (e->mma_smem_set != LK_MAX_DYN_SMEM) {
        LK_CUDA(e, cudaFuncSetAttribute(traverse_mmak_kernel<S, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, LK_MAX_DYN_SMEM));
        LK_CUDA(e, cudaFuncSetAttribute(traverse_mmak_kernel<S, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, LK_MAX_DYN_SMEM));
        e->mma_smem_set = LK_MAX_DYN_SMEM;
    }
    long long grid = e->num_sms;      /* one persistent CTA per SM; every CTA gets at least one 8-pattern group */
    if (grid > ngroups) grid = ngroups;
    launch_begin(e, KIND_PARTIAL);
    if (kone) traverse_mmak_kernel<S, true><<<(unsigned)grid, Sh::THREADS, sm, e->stream>>>(e->v, d, nops, sch.nslots, cbs, inv_streamed, nstages, e->d_inv_frags, ngroups);
    else traverse_mmak_kernel<S, false><<<(unsigned)grid, Sh::THREADS, sm, e->stream>>>(e->v, d, nops, sch.nslots, cbs, inv_streamed, nstages, e->d_inv_frags, ngroups);
    return launch_end(e, KIND_PARTIAL, bytes, "traverse_mmak_kernel");
}

int launch_traverse_mmak(lkcuda_engine *e, const lkcuda_partial_op *ops, int nops) {
    if (e->v.S == 64) return launch_mmak<64>(e, ops, nops);
    return launch_mmak<20>(e, ops, nops);
}

}  // namespace lkc
