/*
 * lkcuda_partial_mma.cu — protein (S = 20) partial likelihoods on the FP64 tensor cores (sm_100a).
 *
 * Replaces PhyloTree::computePartialLikelihoodEigenSIMD<Vec4d,4,20> (phylokernel.h:71-483, instantiated at
 * phylotreeavx.cpp:85-105) and its fused mix-rate twin (phylokernelmixrate.h:23-449) for K <= 4 blocks:
 *     L_dad[p][k][i] = sum_x inv_evec_k[i][x] * (sum_j eleft_k[x][j] L_left[p][k][j]) * (sum_j eright_k[x][j] L_right[p][k][j])
 * Per node and (pattern, k) that is three 20x20 matrix-vector products: FP64-pipe bound (AI > 15 flop/B once
 * the children stay on chip), so the contraction runs as mma.sync.aligned.m8n8k4.f64 (DMMA; measured
 * 37.1 TFLOP/s vs 33.8 for DFMA on this B200, and one instruction instead of 32 per 256 FMAs).
 *
 * Same whole-traversal structure as the DNA kernel (lkcuda_partial.cu): a persistent CTA executes the entire
 * post-order schedule for a tile of TP patterns; results needed again by a later node of the launch are parked in
 * shared-memory slots (Sethi-Ullman order, spill = re-read the arena vector); one producer warp streams the per-node
 * operands into a 2-stage shared-memory ring with cp.async.bulk + mbarriers; every node is written to its arena
 * vector once.
 *
 * Fragment algebra.  A warp owns 8 patterns (rows m of the MMA) and walks k = 0..K-1.  With 20 padded to 24 = 3
 * tiles of 8, the C fragment of an m8n8 tile gives lane (g = lane/4, q = lane%4) the entries
 *     C[pattern g][8*tile + 2q + c],  c = 0, 1.
 * The A operand of m8n8k4 wants lane (g, q) to hold A[pattern g][k-index q].  The eigen index that a contraction
 * sums over is a dummy, so k-step (tile, c) is DEFINED to carry the index 8*tile + 2q + c in lane q: then the C
 * fragment of one contraction IS the A fragment of the next (no shuffle, no shared-memory transpose) and only the
 * B fragments (the matrices, prepared once per node by prep_frag_kernel) are permuted:
 *     Bfrag[rt][ct][c][lane] = M[8*rt + lane/4][8*ct + 2*(lane%4) + c]      (0 outside 20x20)
 * A partial-likelihood vector is read from / written to memory directly in that fragment layout (16 bytes per
 * lane and tile: entries 8*tile + 2q, 8*tile + 2q + 1).
 */
#include "lkcuda_traverse.cuh"

using namespace lkc;

#define MM_S 20
#define MM_NT 3                 /* 8-wide tiles covering the padded 24 */
#define MM_FRAGS (MM_NT * MM_NT * 2)    /* B fragments of one 20x20 matrix: (rt, ct, c) */
#define MM_FRAG_DOUBLES (MM_FRAGS * 32) /* 576 doubles = 4608 bytes per matrix */
#ifndef MM_CWARPS
#define MM_CWARPS 16     /* 8 -> 16 consumer warps: 14.1 -> 12.9 ms per 200 x 200k LG+G4 traversal although only one slot is left (profiles/r2_s13_mma_variants.txt) */
#endif
#define MM_CTHREADS (MM_CWARPS * 32)
#define MM_THREADS (MM_CTHREADS + 32)
#ifndef MM_G
#define MM_G 1                  /* 8-pattern groups per consumer warp: they share the B fragments of every contraction */
#endif
#define MM_TP (MM_CWARPS * 8 * MM_G)   /* patterns per tile */
#define MM_STAGES 2
#define MM_MAXK 4
#define MM_ROW 24               /* doubles per (state, k) row of a staged tip table */
#define MM_THR_HI 0x2FF00000

/* fragment slot (tile t, column j) -> state / eigen index, -1 = padding.  The 4 entries of the last tile sit in its EVEN slots
   (slot 2q <-> entry 16 + q), so its odd slots form one all-padding k-step that every contraction skips: 5 k-steps x 3 tiles =
   15 DMMAs per 20x20 product instead of 18. */
__host__ __device__ __forceinline__ int mm_index_of(int t, int j) { return t < 2 ? 8 * t + j : ((j & 1) ? -1 : 16 + (j >> 1)); }

/* ------------------------------------------------------------------ operand preparation ------------------------------------------------------------------ */

/* per child slot c (one CTA): K fragment-ordered matrices evec*exp(eval*rate*len) [K][MM_FRAGS][32] followed, for a tip
   child, by the lookup table tab[state][k][24] = sum_j ech_k[x][j] * tip_partial_lh[state][j] (x >= 20: 0; unknown state: 1) */
__global__ void __launch_bounds__(256) prep_frag_kernel(DevView v, const DevChild *__restrict__ children) {
    __shared__ double sh_m[MM_S * MM_S];
    const int c = blockIdx.x, K = v.K;
    const DevChild ch = children[c];
    double *out = v.tables + (long long)c * v.table_slot;
    double *tab = out + (long long)K * MM_FRAG_DOUBLES;
    for (int k = 0; k < K; k++) {
        const double *evec = v.evec + (long long)v.blk_eigen[k] * MM_S * MM_S, *eval = v.eval + v.blk_eigen[k] * MM_S;
        __syncthreads();
        for (int idx = threadIdx.x; idx < MM_S * MM_S; idx += blockDim.x) {
            int j = idx % MM_S;
            sh_m[idx] = evec[idx] * exp(eval[j] * (v.blk_rate[k] * child_len(v, ch)));
        }
        __syncthreads();
        for (int idx = threadIdx.x; idx < MM_FRAG_DOUBLES; idx += blockDim.x) {
            int lane = idx & 31, f = idx >> 5, cc = f & 1, ct = (f >> 1) % MM_NT, rt = (f >> 1) / MM_NT;
            int r = mm_index_of(rt, lane >> 2), col = mm_index_of(ct, 2 * (lane & 3) + cc);
            out[(long long)k * MM_FRAG_DOUBLES + idx] = (r >= 0 && col >= 0) ? sh_m[r * MM_S + col] : 0.0;
        }
        if (LKCUDA_IS_TIP(ch.id)) {
            for (int idx = threadIdx.x; idx < (v.SU + 1) * MM_ROW; idx += blockDim.x) {
                int st = idx / MM_ROW, pos = idx - st * MM_ROW, x = mm_index_of(pos >> 3, pos & 7);   /* rows in fragment-slot order */
                double val = 0.0;
                if (x >= 0) {
                    if (st == v.SU) val = 1.0;
                    else {
                        const double *tip = v.tip + ((long long)st * v.E + v.blk_eigen[k]) * MM_S;
                        for (int j = 0; j < MM_S; j++) val += sh_m[x * MM_S + j] * tip[j];
                    }
                }
                tab[((long long)st * K + k) * MM_ROW + pos] = val;
            }
        }
    }
}

/* fragment-ordered inverse eigenvectors [K][MM_FRAGS][32]: out[i] = sum_x inv_evec[i][x] t[x] */
__global__ void __launch_bounds__(256) prep_inv_frag_kernel(DevView v, double *out) {
    const int k = blockIdx.x;
    const double *inv = v.inv_evec + (long long)v.blk_eigen[k] * MM_S * MM_S;
    for (int idx = threadIdx.x; idx < MM_FRAG_DOUBLES; idx += blockDim.x) {
        int lane = idx & 31, f = idx >> 5, cc = f & 1, ct = (f >> 1) % MM_NT, rt = (f >> 1) / MM_NT;
        int r = mm_index_of(rt, lane >> 2), col = mm_index_of(ct, 2 * (lane & 3) + cc);
        out[(long long)k * MM_FRAG_DOUBLES + idx] = (r >= 0 && col >= 0) ? inv[r * MM_S + col] : 0.0;
    }
}

/* ------------------------------------------------------------------ the kernel ------------------------------------------------------------------ */

__device__ __forceinline__ void dmma(double &d0, double &d1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

/* acc[gi][rt][c] += sum over k-steps (ct, cc) of X[gi][ct][cc] * frag[rt][ct][cc]   (frag: this lane's column of a fragment-ordered matrix;
   gi: the warp's G groups of 8 patterns, which share every B fragment - one LDS.64 feeds G DMMAs).
   The k-steps run in the OUTER loop: the accumulator tiles are independent chains, and the asm statements keep their program
   order, so consecutive DMMAs never wait for each other (with the tiles outside, five dependent DMMAs followed one another) */
template <int G>
__device__ __forceinline__ void contract(double (&acc)[G][MM_NT][2], const double (&X)[G][MM_NT][2], const double *frag) {
#pragma unroll
    for (int gi = 0; gi < G; gi++)
#pragma unroll
        for (int rt = 0; rt < MM_NT; rt++) { acc[gi][rt][0] = 0.0; acc[gi][rt][1] = 0.0; }
#pragma unroll
    for (int ct = 0; ct < MM_NT; ct++) {
#pragma unroll
        for (int cc = 0; cc < 2; cc++) {
            if (ct == MM_NT - 1 && cc == 1) continue;       /* all-padding k-step */
#pragma unroll
            for (int rt = 0; rt < MM_NT; rt++) {
                const double f = frag[((rt * MM_NT + ct) * 2 + cc) * 32];
#pragma unroll
                for (int gi = 0; gi < G; gi++) dmma(acc[gi][rt][0], acc[gi][rt][1], X[gi][ct][cc], f);
            }
        }
    }
}

/* the two children's contractions of an internal-internal node, interleaved: 6 G independent chains */
template <int G>
__device__ __forceinline__ void contract2(double (&a0)[G][MM_NT][2], const double (&X0)[G][MM_NT][2], const double *f0,
                                          double (&a1)[G][MM_NT][2], const double (&X1)[G][MM_NT][2], const double *f1) {
#pragma unroll
    for (int gi = 0; gi < G; gi++)
#pragma unroll
        for (int rt = 0; rt < MM_NT; rt++) { a0[gi][rt][0] = 0.0; a0[gi][rt][1] = 0.0; a1[gi][rt][0] = 0.0; a1[gi][rt][1] = 0.0; }
#pragma unroll
    for (int ct = 0; ct < MM_NT; ct++) {
#pragma unroll
        for (int cc = 0; cc < 2; cc++) {
            if (ct == MM_NT - 1 && cc == 1) continue;
#pragma unroll
            for (int rt = 0; rt < MM_NT; rt++) {
                const double b0 = f0[((rt * MM_NT + ct) * 2 + cc) * 32], b1 = f1[((rt * MM_NT + ct) * 2 + cc) * 32];
#pragma unroll
                for (int gi = 0; gi < G; gi++) {
                    dmma(a0[gi][rt][0], a0[gi][rt][1], X0[gi][ct][cc], b0);
                    dmma(a1[gi][rt][0], a1[gi][rt][1], X1[gi][ct][cc], b1);
                }
            }
        }
    }
}

__device__ __forceinline__ void st_stream128(double *p, double x, double y) {
    asm volatile("st.global.L1::no_allocate.v2.f64 [%0], {%1,%2};" :: "l"(p), "d"(x), "d"(y) : "memory");
}

/* lane (g, q) holds slots 2q, 2q + 1 of every tile: entries 8t + 2q, 8t + 2q + 1 for t < 2, entry 16 + q (and padding) for t = 2.
   blk: start of block k of one pattern in an arena vector (20 doubles, reference layout) */
__device__ __forceinline__ double2 mm_g_ld(const double *blk, int nt, int q) {
    if (nt < 2) return ld_coherent128(blk + 8 * nt + 2 * q);
    double r;
    asm volatile("ld.global.L1::no_allocate.f64 %0, [%1];" : "=d"(r) : "l"(blk + 16 + q) : "memory");
    return make_double2(r, 0.0);
}
__device__ __forceinline__ void mm_g_st(double *blk, int nt, int q, double x, double y) {
    if (nt < 2) st_stream128(blk + 8 * nt + 2 * q, x, y);
    else asm volatile("st.global.L1::no_allocate.f64 [%0], %1;" :: "l"(blk + 16 + q), "d"(x) : "memory");
}
/* slot layout (doubles): [k][ tile 0: TP*8 | tile 1: TP*8 | tile 2: TP*4 ] - conflict-free LDS/STS.128 (tile 2: .64) */
__device__ __forceinline__ double2 mm_s_ld(const double *slot, int k, int nt, int pat, int q) {
    const double *b = slot + k * (MM_TP * MM_S) + nt * (MM_TP * 8);
    if (nt < 2) return *reinterpret_cast<const double2 *>(b + pat * 8 + 2 * q);
    return make_double2(b[pat * 4 + q], 0.0);
}
__device__ __forceinline__ void mm_s_st(double *slot, int k, int nt, int pat, int q, double x, double y) {
    double *b = slot + k * (MM_TP * MM_S) + nt * (MM_TP * 8);
    if (nt < 2) *reinterpret_cast<double2 *>(b + pat * 8 + 2 * q) = make_double2(x, y);
    else b[pat * 4 + q] = x;
}

/* the warp's G pattern groups of block k of a child vector: from its slot, else from the arena */
template <int G>
__device__ __forceinline__ void mm_load(double (&X)[G][MM_NT][2], const double *ps, const double *gv, const long long (&p)[G], const int (&pat)[G],
                                        int B, int k, int q) {
    if (ps) {
#pragma unroll
        for (int gi = 0; gi < G; gi++)
#pragma unroll
            for (int nt = 0; nt < MM_NT; nt++) { const double2 a = mm_s_ld(ps, k, nt, pat[gi], q); X[gi][nt][0] = a.x; X[gi][nt][1] = a.y; }
    } else {
#pragma unroll
        for (int gi = 0; gi < G; gi++)
#pragma unroll
            for (int nt = 0; nt < MM_NT; nt++) { const double2 a = mm_g_ld(gv + p[gi] * B + k * MM_S, nt, q); X[gi][nt][0] = a.x; X[gi][nt][1] = a.y; }
    }
}

/* ring stage: descriptor (64) | child 0 block | child 1 block (cbs doubles each: K fragment matrices or tip table) | states 0 | states 1 */
__host__ __device__ __forceinline__ int mm_stage_bytes(int cbs) { return 64 + 2 * cbs * 8 + 2 * (MM_TP + 16); }

__global__ void __launch_bounds__(MM_THREADS, 1) traverse_mma20_kernel(DevView v, const TravOp *__restrict__ ops, int nops, int nslots, int cbs,
                                                                      const double *__restrict__ inv_frags, long long ntiles) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int K = v.K, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int stage_bytes = mm_stage_bytes(cbs);
    unsigned char *ring = smem_raw;                                                         /* [MM_STAGES][stage_bytes] */
    double *sINV = reinterpret_cast<double *>(ring + MM_STAGES * stage_bytes);              /* [K][MM_FRAG_DOUBLES] */
    double *sl = sINV + K * MM_FRAG_DOUBLES;                                                /* [nslots][K*TP*20] */
    int *ssc = reinterpret_cast<int *>(sl + (size_t)nslots * K * MM_TP * MM_S);             /* [nslots][TP] */
    uint64_t *full = reinterpret_cast<uint64_t *>(ssc + nslots * MM_TP + (nslots * MM_TP & 1)), *empty = full + MM_STAGES;
    for (int idx = tid; idx < K * MM_FRAG_DOUBLES; idx += MM_THREADS) sINV[idx] = inv_frags[idx];
    if (tid == 0) {
        for (int s = 0; s < MM_STAGES; s++) { mbar_init(&full[s], 1); mbar_init(&empty[s], MM_CWARPS); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    if (warp == MM_CWARPS) {
        /* ------------------------------ producer ------------------------------ */
        if (lane == 0) {
            uint32_t c = 0;
            const uint32_t mat_bytes = (uint32_t)K * MM_FRAG_DOUBLES * 8, tab_bytes = (uint32_t)(v.SU + 1) * K * MM_ROW * 8;
            for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
                const long long p0 = tile * MM_TP;          /* MM_TP is a multiple of 16: bulk copies stay 16-byte aligned */
                for (int i = 0; i < nops; i++, c++) {
                    const uint32_t s = c % MM_STAGES, ph = (c / MM_STAGES) & 1;
                    const int src0 = ops[i].src0, src1 = ops[i].src1;
                    unsigned char *st = ring + s * stage_bytes;
                    const bool t0 = LKCUDA_IS_TIP(src0), t1 = LKCUDA_IS_TIP(src1);
                    const uint32_t b0 = t0 ? tab_bytes : mat_bytes, b1 = t1 ? tab_bytes : mat_bytes;
                    const double *g0 = v.tables + (long long)(2 * i) * v.table_slot + (t0 ? K * MM_FRAG_DOUBLES : 0);
                    const double *g1 = v.tables + (long long)(2 * i + 1) * v.table_slot + (t1 ? K * MM_FRAG_DOUBLES : 0);
                    mbar_wait_backoff(&empty[s], ph ^ 1);
                    mbar_expect_tx(&full[s], 64 + b0 + b1 + (t0 ? MM_TP : 0) + (t1 ? MM_TP : 0));
                    bulk_g2s(st, ops + i, 64, &full[s]);
                    bulk_g2s(st + 64, g0, b0, &full[s]);
                    bulk_g2s(st + 64 + cbs * 8, g1, b1, &full[s]);
                    if (t0) bulk_g2s(st + 64 + 2 * cbs * 8, tip_states(v, src0) + p0, MM_TP, &full[s]);
                    if (t1) bulk_g2s(st + 64 + 2 * cbs * 8 + MM_TP + 16, tip_states(v, src1) + p0, MM_TP, &full[s]);
                }
            }
        }
        return;
    }

    /* ------------------------------ consumers: warp w owns MM_G groups of 8 patterns of the tile ------------------------------ */
    const int g = lane >> 2, q = lane & 3;
    int pat[MM_G];
#pragma unroll
    for (int gi = 0; gi < MM_G; gi++) pat[gi] = (warp * MM_G + gi) * 8 + g;
    const bool zero_guard = (v.flags & LKCUDA_FLAG_ZERO_GUARD) != 0, ignore_invar = (v.flags & LKCUDA_FLAG_SCALE_IGNORES_INVAR) != 0;
    const int B = v.B;
    uint32_t c = 0;

    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        long long p[MM_G];
        bool valid[MM_G];
#pragma unroll
        for (int gi = 0; gi < MM_G; gi++) {
            p[gi] = tile * MM_TP + pat[gi];
            valid[gi] = p[gi] < v.NP;
            if (!valid[gi]) p[gi] = v.NP - 1;
        }
        for (int i = 0; i < nops; i++, c++) {
            const uint32_t s = c % MM_STAGES;
            mbar_wait(&full[s], (c / MM_STAGES) & 1);
            const unsigned char *stage = ring + s * stage_bytes;
            const TravOp *op = reinterpret_cast<const TravOp *>(stage);
            const int src0 = op->src0, src1 = op->src1, slot0 = op->slot0, slot1 = op->slot1, dst_slot = op->dst_slot;
            const bool tip0 = LKCUDA_IS_TIP(src0), tip1 = LKCUDA_IS_TIP(src1);
            const double *blk0 = reinterpret_cast<const double *>(stage + 64), *blk1 = blk0 + cbs;
            const double *g0 = op->src0_lh, *g1 = op->src1_lh;
            double *const dst_lh = op->dst_lh;
            int16_t *const dst_sc = op->dst_sc;
            int st0[MM_G], st1[MM_G], sc[MM_G], mx[MM_G], nz[MM_G];
#pragma unroll
            for (int gi = 0; gi < MM_G; gi++) {
                st0[gi] = tip0 ? stage[64 + 2 * cbs * 8 + pat[gi]] : 0;
                st1[gi] = tip1 ? stage[64 + 2 * cbs * 8 + MM_TP + 16 + pat[gi]] : 0;
                sc[gi] = 0; mx[gi] = 0; nz[gi] = 0;
                if (!tip0) sc[gi] += slot0 >= 0 ? ssc[slot0 * MM_TP + pat[gi]] : (int)vec_sc(v, src0)[p[gi]];
                if (!tip1) sc[gi] += slot1 >= 0 ? ssc[slot1 * MM_TP + pat[gi]] : (int)vec_sc(v, src1)[p[gi]];
            }
            double *dslot = dst_slot >= 0 ? sl + (size_t)dst_slot * K * MM_TP * MM_S : NULL;
            const double *ps0 = slot0 >= 0 ? sl + (size_t)slot0 * K * MM_TP * MM_S : NULL;
            const double *ps1 = slot1 >= 0 ? sl + (size_t)slot1 * K * MM_TP * MM_S : NULL;

            for (int k = 0; k < K; k++) {
                double t[MM_G][MM_NT][2];
                if (!tip0 && !tip1) {
                    /* ---- both children are vectors: load both, contract them together ---- */
                    double X0[MM_G][MM_NT][2], X1[MM_G][MM_NT][2], u[MM_G][MM_NT][2];
                    mm_load<MM_G>(X0, ps0, g0, p, pat, B, k, q);
                    mm_load<MM_G>(X1, ps1, g1, p, pat, B, k, q);
                    contract2<MM_G>(t, X0, blk0 + (size_t)k * MM_FRAG_DOUBLES + lane, u, X1, blk1 + (size_t)k * MM_FRAG_DOUBLES + lane);
#pragma unroll
                    for (int gi = 0; gi < MM_G; gi++)
#pragma unroll
                        for (int nt = 0; nt < MM_NT; nt++) { t[gi][nt][0] *= u[gi][nt][0]; t[gi][nt][1] *= u[gi][nt][1]; }
                } else {
                    /* ---- child 0 ---- */
                    if (tip0) {
#pragma unroll
                        for (int gi = 0; gi < MM_G; gi++) {
                            const double *row = blk0 + ((size_t)st0[gi] * K + k) * MM_ROW + 2 * q;
#pragma unroll
                            for (int nt = 0; nt < MM_NT; nt++) { const double2 a = *reinterpret_cast<const double2 *>(row + 8 * nt); t[gi][nt][0] = a.x; t[gi][nt][1] = a.y; }
                        }
                    } else {
                        double X[MM_G][MM_NT][2];
                        mm_load<MM_G>(X, ps0, g0, p, pat, B, k, q);
                        contract<MM_G>(t, X, blk0 + (size_t)k * MM_FRAG_DOUBLES + lane);
                    }
                    /* ---- child 1 ---- */
                    if (tip1) {
#pragma unroll
                        for (int gi = 0; gi < MM_G; gi++) {
                            const double *row = blk1 + ((size_t)st1[gi] * K + k) * MM_ROW + 2 * q;
#pragma unroll
                            for (int nt = 0; nt < MM_NT; nt++) { const double2 a = *reinterpret_cast<const double2 *>(row + 8 * nt); t[gi][nt][0] *= a.x; t[gi][nt][1] *= a.y; }
                        }
                    } else {
                        double X[MM_G][MM_NT][2], u[MM_G][MM_NT][2];
                        mm_load<MM_G>(X, ps1, g1, p, pat, B, k, q);
                        contract<MM_G>(u, X, blk1 + (size_t)k * MM_FRAG_DOUBLES + lane);
#pragma unroll
                        for (int gi = 0; gi < MM_G; gi++)
#pragma unroll
                            for (int nt = 0; nt < MM_NT; nt++) { t[gi][nt][0] *= u[gi][nt][0]; t[gi][nt][1] *= u[gi][nt][1]; }
                    }
                }
                /* ---- back to state space ---- */
                double o[MM_G][MM_NT][2];
                contract<MM_G>(o, t, sINV + (size_t)k * MM_FRAG_DOUBLES + lane);
#pragma unroll
                for (int gi = 0; gi < MM_G; gi++) {
                    double *gd = dst_lh + p[gi] * B + k * MM_S;
#pragma unroll
                    for (int nt = 0; nt < MM_NT; nt++) {
                        const int h0 = __double2hiint(o[gi][nt][0]) & 0x7fffffff, h1 = __double2hiint(o[gi][nt][1]) & 0x7fffffff;
                        mx[gi] = max(mx[gi], max(h0, h1));
                        nz[gi] |= h0 | h1 | __double2loint(o[gi][nt][0]) | __double2loint(o[gi][nt][1]);
                        if (valid[gi]) mm_g_st(gd, nt, q, o[gi][nt][0], o[gi][nt][1]);
                        if (dslot) mm_s_st(dslot, k, nt, pat[gi], q, o[gi][nt][0], o[gi][nt][1]);
                    }
                }
            }
            /* every operand of this node has been consumed: hand the stage back */
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[s]);
            /* ---- scaling (phylokernel.h:380-392, 462-474): lh_max over all k and states of the pattern = the quad's maximum ---- */
#pragma unroll
            for (int gi = 0; gi < MM_G; gi++) {
                int m = mx[gi], z = nz[gi];
                m = max(m, __shfl_xor_sync(0xffffffffu, m, 1)); m = max(m, __shfl_xor_sync(0xffffffffu, m, 2));
                z |= __shfl_xor_sync(0xffffffffu, z, 1); z |= __shfl_xor_sync(0xffffffffu, z, 2);
                const bool tiny = !(tip0 && tip1) && m < MM_THR_HI;
                if (__any_sync(0xffffffffu, tiny)) {
                    /* rare: re-read what this lane just wrote, rescale (or refill), write again */
                    const bool refill = tiny && zero_guard && z == 0;
                    const bool rescale = tiny && !refill && (ignore_invar || v.invar[p[gi]] == 0.0);
                    if (refill || rescale) {
                        for (int k = 0; k < K; k++) {
                            double *gd = dst_lh + p[gi] * B + k * MM_S;
                            const double *tu = v.tip + ((long long)v.SU * v.E + v.blk_eigen[k]) * MM_S;
#pragma unroll
                            for (int nt = 0; nt < MM_NT; nt++) {
                                double2 a;
                                if (refill) a = nt < 2 ? make_double2(tu[8 * nt + 2 * q], tu[8 * nt + 2 * q + 1]) : make_double2(tu[16 + q], 0.0);
                                else { a = dslot ? mm_s_ld(dslot, k, nt, pat[gi], q) : mm_g_ld(gd, nt, q); a.x *= LK_THR_INV; a.y *= LK_THR_INV; }
                                if (valid[gi]) mm_g_st(gd, nt, q, a.x, a.y);
                                if (dslot) mm_s_st(dslot, k, nt, pat[gi], q, a.x, a.y);
                            }
                        }
                        sc[gi] += refill ? 4 : 1;
                    }
                }
                if (q == 0) {
                    if (valid[gi]) dst_sc[p[gi]] = (int16_t)sc[gi];
                    if (dst_slot >= 0) ssc[dst_slot * MM_TP + pat[gi]] = sc[gi];
                }
            }
            /* a parked result is read back by the other lanes of the quad (scale count) and, when spilled, from the arena */
            __syncwarp();
            if (dst_slot == -2) { __threadfence_block(); __syncwarp(); }
        }
    }
}

/* ------------------------------------------------------------------ host side ------------------------------------------------------------------ */

namespace lkc {

bool use_traverse_mma(const lkcuda_engine *e) {
    const DevView &v = e->v;
    if (v.S != MM_S || v.K > MM_MAXK) return false;
    if (e->trav_disable_mma) return false;
    return true;
}

int launch_traverse_mma(lkcuda_engine *e, const lkcuda_partial_op *ops, int nops) {
    const DevView &v = e->v;
    const int K = v.K;
    const int cbs_mat = K * MM_FRAG_DOUBLES, cbs_tab = (v.SU + 1) * K * MM_ROW, cbs = cbs_mat > cbs_tab ? cbs_mat : cbs_tab;
    if ((long long)cbs_mat + cbs_tab > v.table_slot) return fail(e, LKCUDA_ERR_STATE, "table slot too small for the MMA layout");
    const size_t slot_bytes = (size_t)K * MM_TP * MM_S * 8 + (size_t)MM_TP * 4;
    const size_t fixed = (size_t)MM_STAGES * mm_stage_bytes(cbs) + (size_t)K * MM_FRAG_DOUBLES * 8 + 2 * MM_STAGES * 8 + 128;
    if (fixed > 227 * 1024) return fail(e, LKCUDA_ERR_ARG, "too many categories for the protein MMA kernel");
    int maxslots = (int)((227 * 1024 - fixed) / slot_bytes);
    if (e->trav_maxslots >= 0 && e->trav_maxslots < maxslots) maxslots = e->trav_maxslots;
    TravSchedule sch;
    int rc = schedule_traverse(e, ops, nops, maxslots, false, sch);
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
    if (!e->d_inv_frags) LK_CUDA(e, cudaMalloc((void **)&e->d_inv_frags, (size_t)MM_MAXK * MM_FRAG_DOUBLES * 8));
    TravOp *d = reinterpret_cast<TravOp *>(e->d_ops[slot]);
    LK_CUDA(e, cudaMemcpyAsync(d, h, sizeof(TravOp) * nops, cudaMemcpyHostToDevice, e->stream));
    LK_CUDA(e, cudaMemcpyAsync(e->d_child[slot], e->h_child[slot], sizeof(DevChild) * 2 * nops, cudaMemcpyHostToDevice, e->stream));
    LK_CUDA(e, cudaEventRecord(e->ring_ev[slot], e->stream));
    e->ring_used[slot] = true;
    if (v.NP == 0) return LKCUDA_OK;
    rc = states_barrier(e);
    if (rc) return rc;

    launch_begin(e, KIND_PREP);
    prep_frag_kernel<<<2 * nops, 256, 0, e->stream>>>(e->v, e->d_child[slot]);
    rc = launch_end(e, KIND_PREP, 0, "prep_frag_kernel");
    if (rc) return rc;
    launch_begin(e, KIND_PREP);
    prep_inv_frag_kernel<<<K, 256, 0, e->stream>>>(e->v, e->d_inv_frags);
    rc = launch_end(e, KIND_PREP, 0, "prep_inv_frag_kernel");
    if (rc) return rc;

    const long long ntiles = (v.NP + MM_TP - 1) / MM_TP;
    const size_t sm = (size_t)sch.nslots * slot_bytes + fixed;
    if (e->mma_smem_set != LK_MAX_DYN_SMEM) {
        LK_CUDA(e, cudaFuncSetAttribute(traverse_mma20_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, LK_MAX_DYN_SMEM));
        e->mma_smem_set = LK_MAX_DYN_SMEM;
    }
    long long grid = e->num_sms;
    if (grid > ntiles) grid = ntiles;
    launch_begin(e, KIND_PARTIAL);
    traverse_mma20_kernel<<<(unsigned)grid, MM_THREADS, sm, e->stream>>>(e->v, d, nops, sch.nslots, cbs, e->d_inv_frags, ntiles);
    return launch_end(e, KIND_PARTIAL, bytes, "traverse_mma20_kernel");
}

}  // namespace lkc
