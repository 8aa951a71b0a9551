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
 * Launch structure: the host hands over a post-order schedule; one `prep_tables_kernel` launch builds
 * every child's transition table (and, for tip children, the per-state lookup table of
 * phylokernel.h:191-232), then
 *   traverse_s4_kernel     DNA (S = 4, K <= 32): ONE launch executes the whole schedule, pattern tile by
 *                          pattern tile, with the intermediate partials parked in shared memory; see the
 *                          comment above the kernel.
 *   partial_generic_kernel any S, any K, any number of children (multifurcations): one launch per tree
 *                          level (gridDim.y = nodes of the level), pattern tiles staged in shared memory;
 *                          correctness fallback for shapes without a tuned kernel.
 */
#include "lkcuda_traverse.cuh"
#include <type_traits>

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
/* mpad: doubles of padding after each k's S*S matrix (0 for the generic kernel, 2 for the DNA traversal kernel's
   conflict-free shared-memory image); the tip table follows the K padded matrices */
__global__ void __launch_bounds__(256) prep_tables_kernel(DevView v, const DevChild *__restrict__ children, int mpad) {
    extern __shared__ __align__(32) double sh_ex[];            /* exp(eval*rate*len) per (k, j): B doubles */
    const int c = blockIdx.x;
    const DevChild ch = children[c];
    const int S = v.S, K = v.K, B = v.B;
    double *ech = v.tables + (long long)c * v.table_slot;
    const int MS = S * S + mpad;
    double *tab = ech + (long long)K * MS;
    for (int idx = threadIdx.x; idx < B; idx += blockDim.x) {
        int k = idx / S, j = idx - k * S;
        sh_ex[idx] = exp(v.eval[v.blk_eigen[k] * S + j] * (v.blk_rate[k] * child_len(v, ch)));
    }
    __syncthreads();
    const int SS = S * S;
    for (int idx = threadIdx.x; idx < K * SS; idx += blockDim.x) {
        int k = idx / SS, r = idx - k * SS, j = r % S;
        ech[k * MS + r] = v.evec[(long long)v.blk_eigen[k] * SS + r] * sh_ex[k * S + j];
    }
    if (!LKCUDA_IS_TIP(ch.id)) return;
    __syncthreads();
    const int SU = v.SU;
    /* traversal layout (mpad != 0): row `st` is skewed by (st >> 1) 16-byte granules so that two lanes of a quarter-warp
       that hold different states of {0,1} x {2,3} never meet in a bank (TRV_TAB_ROW below) */
    const int skew = mpad ? 2 : 0;
    for (int idx = threadIdx.x; idx < (SU + 1) * B; idx += blockDim.x) {
        int st = idx / B, r = idx - st * B;
        double val = 1.0;
        if (st < SU) {
            int k = r / S, x = r - k * S;
            const double *row = ech + (long long)k * MS + (long long)x * S;
            const double *tip = v.tip + ((long long)st * v.E + v.blk_eigen[k]) * S;
            val = 0.0;
            for (int j = 0; j < S; j++) val += row[j] * tip[j];
        }
        tab[idx + (st >> 1) * skew] = val;
    }
}

/* ------------------------------------------------------------------ DNA: whole-traversal kernel ------------------------------------------------------------------ */

/*
 * Every per-pattern quantity of the pruning recursion is independent of every other pattern, so the
 * tree does not have to be walked level by level across the whole alignment: a CTA takes a TILE of
 * patterns and executes the ENTIRE post-order schedule for it.  A thread owns the same units
 * (pattern, category) for all nodes, therefore
 *   - a node's result is consumed by the very thread that produced it: no inter-thread (let alone
 *     inter-CTA) synchronisation on the data exists in the loop;
 *   - results that a later node of the same schedule needs are parked in shared-memory SLOTS (allocated
 *     by the host with a Sethi-Ullman ordering of the schedule, so ~log2(taxa) slots suffice): internal
 *     children are read from shared memory, not from HBM;
 *   - every node's partial_lh / scale_num is still written to its arena vector (the reference's buffer
 *     semantics: other branches' derivatives read them later), as whole-sector streaming stores.
 * HBM traffic per node drops from "read 2 children + write 1" to "write 1" (+ 1 byte per tip child).
 *
 * Unit mapping: unit u = ptn*K + k is 32-byte sector u of a vector.  A warp covers UPW = (32/K)*K
 * consecutive units (all K units of a pattern in one warp, so lh_max is a shuffle reduction inside the
 * lane group); a thread handles R units spaced by the CTA's unit count, all with the same k, so the
 * 4x4 transition matrices it pulls into registers are used R times.
 *
 * Warp specialisation: TRV_CWARPS consumer warps do the arithmetic; ONE producer warp streams the
 * per-node operands (node descriptor + the two children's transition matrix / tip lookup table, built
 * by prep_tables_kernel) from global memory into a TRV_STAGES-deep shared-memory ring with bulk
 * asynchronous copies (cp.async.bulk -> UBLKCP) that complete on mbarriers; consumers hand a stage
 * back through a second mbarrier.  Nothing in the consumers' inner loop reads a table from global memory.
 */
#ifndef TRV_CWARPS
#define TRV_CWARPS 11
#endif
#define TRV_CTHREADS (TRV_CWARPS * 32)
#define TRV_THREADS (TRV_CTHREADS + 32)
/* Tried on the B200 and not kept (round 2): (a) forming pattern / offset / validity of a unit where they are used instead of
   keeping p[], off[], valid[] per unit - no unit more fits and the kernel gets 2 % slower; (b) 12 consumer warps x 152 registers with the
   producer as a fourth warpgroup that hands its registers over (setmaxnreg): 2.33 vs 2.40 ms per 1M-pattern traversal
   (profiles/r2_s21_traverse_setmaxnreg_variants.txt), but in 3 of 9 runs that build left the 32 patterns of ONE warp of ONE tile of a
   partial traversal unwritten (found with LKCUDA_POISON=1) - the stage-release hole described at TRV_ARRIVE_DEP, which that build's timing
   exposed; not re-measured since the fix. */
#define TRV_MAX_STAGES 8
/* units (pattern, category) per consumer thread.  Measured on one box, A/B (profiles/r2_s50_step_variants.txt): 5 units make the full-tree
   traversal 3 % faster than 4 (2.56 vs 2.63 ms per lnL evaluation at 1M patterns) but a list whose internal children mostly come from
   the arena - the path to the derivative's branch: one streamed child per node - 10 % slower (1.00 vs 0.90 ms); launch_traverse picks
   per launch */
#ifndef TRV_R
#define TRV_R 5             /* lists that read (almost) no child from the arena */
#endif
#ifndef TRV_R_ARENA
#define TRV_R_ARENA 4       /* lists with arena-read children; also the instantiation that can carry the fused root branch */
#endif
#ifndef TRV_INV_REGS
#define TRV_INV_REGS 0      /* 1: the inverse eigenvectors of the lane's category stay in registers for the whole launch */
#endif
#define S4_MPAD 18  /* doubles per k of a staged 4x4 matrix: 16 + 2 keeps LDS.128 conflict-free across k */
#define THR_HI 0x2FF00000   /* high word of 2^-256: |x| < 2^-256  <=>  (hi(x) & 0x7fffffff) < THR_HI */
/* doubles from the start of a staged tip table to the row of state st (K*4 doubles per row + the skew written by prep_tables_kernel) */
#define TRV_TAB_ROW(st, K) ((st) * ((K) * 4) + ((st) >> 1) * 2)
#define TRV_TAB_DOUBLES(SU, K) (((SU) + 1) * (K) * 4 + (((SU) >> 1) + 1) * 2)

/* ring stage: descriptor (64) | child 0 block | child 1 block (cbs doubles each: padded matrices or tip lookup table) |
   child 0 tip states | child 1 tip states (spad bytes each: the tile's patterns of that taxon) */
__host__ __device__ __forceinline__ int trv_stage_bytes(int cbs, int spad) { return 64 + 2 * cbs * 8 + 2 * spad; }

/* one row of a 4x4 row-major matrix (a = (Mi0, Mi1), b = (Mi2, Mi3)) times L.  The order of the operations is spelled out - it is the one
   nvcc's contraction gives `a.x*L.x + a.y*L.y + b.x*L.z + b.y*L.w` (first product fused, second rounded) - so that every form of the
   products below (whole matrix per unit, column halves over all units, row by row over all units) gives the same bits */
#define TRV_DOT(a, b, L) fma((b).y, (L).w, fma((b).x, (L).z, fma((a).x, (L).x, (a).y * (L).y)))
/* dst = M * L for a 4x4 row-major matrix held as 8 double2 */
#define TRV_MATVEC(dst, m, L)                   \
    dst[0] = TRV_DOT(m[0], m[1], L);            \
    dst[1] = TRV_DOT(m[2], m[3], L);            \
    dst[2] = TRV_DOT(m[4], m[5], L);            \
    dst[3] = TRV_DOT(m[6], m[7], L);
/* TRV_FORM 1: the matrix of child 0 is applied in two column halves over all units of the thread (8 of its 16 entries live at a time),
   the matrix of child 1 and the inverse eigenvectors row by row (4 entries live at a time); 0: one whole matrix per unit */
#ifndef TRV_FORM
#define TRV_FORM 0
#endif
/* How a consumer warp hands a ring stage back.  On sm_100a mbarrier.arrive does not wait for an earlier LDS of the same warp that is still
   in flight, and ptxas adds no scoreboard wait: the producer's next bulk copy can replace the stage under the load.  Measured (round 2,
   profiles/r2_s44_traverse_stage_release_race.txt): with the descriptor words dst_lh / dst_sc loaded two instructions before the arrive, whole warps stored a node's
   result through the descriptor of the node that got the stage next (8 patterns x R slices of NaN in the poisoned arena, a few per 1M x 98
   nodes) as soon as the code in front of the arrive was LDS-heavy (TRV_FORM 1); the shipped form had the same instruction pair and was only
   luckier.  1 (shipped): the barrier address depends on the words loaded last, so the arrive cannot issue before the stage reads are in
   (shared loads of a warp complete in order); 2 / 3: membar.cta / fence.proxy.async in front of the arrive (also clean, 2-3 % slower);
   0: nothing.  tools/sass_release_check.py (tests/test_sass_release.py) checks the built SASS for an LDS in flight at a consumer arrive. */
#ifndef TRV_ARRIVE_DEP
#define TRV_ARRIVE_DEP 1
#endif
#ifndef TRV_INV_FORM
#define TRV_INV_FORM 0      /* 1: the inverse eigenvectors are applied row by row over all units */
#endif
/* columns (0,1) of M times (x,y) of L: starts the four sums of a unit; columns (2,3) times (z,w): finishes them */
#define TRV_COLS_LO(t, c0, c1, c2, c3, a)                                       \
    t[0] = fma((c0).x, (a).x, (c0).y * (a).y); t[1] = fma((c1).x, (a).x, (c1).y * (a).y); \
    t[2] = fma((c2).x, (a).x, (c2).y * (a).y); t[3] = fma((c3).x, (a).x, (c3).y * (a).y);
#define TRV_COLS_HI(t, c0, c1, c2, c3, b)                                       \
    t[0] = fma((c0).y, (b).y, fma((c0).x, (b).x, t[0])); t[1] = fma((c1).y, (b).y, fma((c1).x, (b).x, t[1])); \
    t[2] = fma((c2).y, (b).y, fma((c2).x, (b).x, t[2])); t[3] = fma((c3).y, (b).y, fma((c3).x, (b).x, t[3]));

/* root branch fused into the tail of the traversal (lkcuda_partial_batch_lnl): the leaf end, its length and the table parameters */
#define TRV_ROOT_ROW 18
struct TravRoot { int on, tip; double len; B16Tab tab; };     /* tip: LKCUDA_TIP(taxon) */
#define TRV_ROOT_SMEM ((32 * TRV_ROOT_ROW + 4 * TRV_CTHREADS + 4 * 32) * 8)

/* first pattern of this CTA's tile in `round`: full rounds give every CTA R slices (tile_pats patterns), interleaved over the CTAs so
   that the pattern ranges are visited in ascending order (the pipelined upload delivers them in that order); the last round gives
   every CTA only ract_last slices */
__device__ __forceinline__ long long trv_tile_start(int round, int nfull, int ract_last, int tile_pats, int slice_pats) {
    if (round < nfull) return ((long long)round * gridDim.x + blockIdx.x) * tile_pats;
    return (long long)nfull * gridDim.x * tile_pats + (long long)blockIdx.x * ract_last * slice_pats;
}

__device__ __forceinline__ void st_stream256_if(bool pred, double *p, const d4 &v) {
    asm volatile("{\n.reg .pred q;\nsetp.ne.u32 q, %5, 0;\n@q st.global.L1::no_allocate.v4.f64 [%0], {%1,%2,%3,%4};\n}"
                 :: "l"(p), "d"(v.x), "d"(v.y), "d"(v.z), "d"(v.w), "r"((unsigned)pred) : "memory");
}
__device__ __forceinline__ void st_s16_if(bool pred, int16_t *p, int val) {
    asm volatile("{\n.reg .pred q;\nsetp.ne.u32 q, %2, 0;\n@q st.global.u16 [%0], %1;\n}" :: "l"(p), "h"((short)val), "r"((unsigned)pred) : "memory");
}

template <int R, bool FUSE>
__global__ void __launch_bounds__(TRV_THREADS, 1) traverse_s4_kernel(DevView v, const TravOp *__restrict__ ops, int nops, int upw, int nslots,
                                                                     int cbs, int spad, int stage_shift, int nfull, int ract_last,
                                                                     const unsigned int *upload_flag, unsigned int upload_base, long long upload_chunk_pat,
                                                                     TravRoot root) {
    constexpr int CT = TRV_CTHREADS;
    static_assert(R <= 6, "the last round is instantiated for 1..5 slices and R");
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int K = v.K, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int stage_bytes = trv_stage_bytes(cbs, spad);
    const uint32_t nstages = 1u << stage_shift, stage_mask = nstages - 1;
    unsigned char *ring = smem_raw;                                                          /* [nstages][stage_bytes] */
    double2 *sl = reinterpret_cast<double2 *>(ring + nstages * stage_bytes);                 /* [nslots][R][2][CT] */
    int *ssc = reinterpret_cast<int *>(sl + nslots * R * 2 * CT);                            /* [nslots][R][CT] */
    double *sINV = reinterpret_cast<double *>(ssc + nslots * R * CT);                        /* [K][S4_MPAD] */
    uint64_t *full = reinterpret_cast<uint64_t *>(sINV + K * S4_MPAD), *empty = full + TRV_MAX_STAGES;
    /* fused root-branch lnL (root.on; K = 4): weighted tip table of the root leaf [state][TRV_ROOT_ROW] | per-thread sums [4][CT] | [4][32] */
    double *rootB = reinterpret_cast<double *>(empty + TRV_MAX_STAGES), *racc = rootB + 32 * TRV_ROOT_ROW, *rred = racc + 4 * CT;
    constexpr bool fuse_root = FUSE && R <= 4;       /* lane k of a pattern finishes slice k: at most K = 4 slices (launch_traverse) */
    for (int idx = tid; idx < K * 16; idx += TRV_THREADS) {
        int kk = idx >> 4, r = idx & 15;
        sINV[kk * S4_MPAD + r] = v.inv_evec[v.blk_eigen[kk] * 16 + r];
    }
    if (fuse_root) {        /* val[k][i] * tip_partial_lh[state][i] of computeLikelihoodBranchEigenSIMD's tip case (phylokernel.h:767-775, 779-830) */
        for (int idx = tid; idx < (v.SU + 1) * 16; idx += TRV_THREADS) {
            const int st = idx >> 4, j = idx & 15;
            rootB[st * TRV_ROOT_ROW + j] = exp(root.tab.cof[j] * root.len) * root.tab.w[j >> 2] * v.tip[((long long)st * v.E + root.tab.eig[j >> 2]) * 4 + (j & 3)];
        }
        for (int idx = tid; idx < 4 * CT; idx += TRV_THREADS) racc[idx] = 0.0;
    }
    if (tid == 0) {
        for (uint32_t s = 0; s < nstages; s++) { mbar_init(&full[s], 1); mbar_init(&empty[s], TRV_CWARPS); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const int cta_units = TRV_CWARPS * upw;
    const int slice_pats = cta_units / K;        /* upw is a multiple of K: one slice = the patterns of ONE unit per consumer thread */
    const int tile_pats = slice_pats * R;
    const int nrounds = nfull + (ract_last > 0 ? 1 : 0);

    if (warp == TRV_CWARPS) {
        /* ------------------------------ producer: one lane feeds the ring ------------------------------ */
        if (lane == 0) {
            uint32_t c = 0;
            unsigned int seen = upload_base;
            const uint32_t mat_bytes = (uint32_t)K * S4_MPAD * 8, tab_bytes = (uint32_t)TRV_TAB_DOUBLES(v.SU, K) * 8;
            for (int round = 0; round < nrounds; round++) {
                const int ract = round < nfull ? R : ract_last;
                const long long p0 = trv_tile_start(round, nfull, ract_last, tile_pats, slice_pats);
                if (p0 >= v.NP) break;
                const long long a0 = p0 & ~15ll;         /* bulk copies start 16-byte aligned */
                if (upload_flag) {
                    /* pipelined upload (lkcuda_set_tip_states_async): wait until the copy engine has delivered this tile's pattern range */
                    long long last = p0 + (long long)ract * slice_pats - 1;
                    if (last > v.NP - 1) last = v.NP - 1;
                    const unsigned int need = upload_base + (unsigned int)(last / upload_chunk_pat) + 1u;
                    while ((int)(seen - need) < 0) {
                        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(seen) : "l"(upload_flag) : "memory");
                        if ((int)(seen - need) < 0) __nanosleep(500);
                    }
                }
                const uint32_t st_bytes = (uint32_t)((p0 - a0 + ract * slice_pats + 15) & ~15ll);
                for (int i = 0; i < nops; i++, c++) {
                    const uint32_t s = c & stage_mask, ph = (c >> stage_shift) & 1;
                    const int src0 = ops[i].src0, src1 = ops[i].src1;
                    unsigned char *st = ring + s * stage_bytes;
                    const bool t0 = LKCUDA_IS_TIP(src0), t1 = LKCUDA_IS_TIP(src1);
                    const uint32_t b0 = t0 ? tab_bytes : mat_bytes, b1 = t1 ? tab_bytes : mat_bytes;
                    const double *g0 = v.tables + (long long)(2 * i) * v.table_slot + (t0 ? K * S4_MPAD : 0);
                    const double *g1 = v.tables + (long long)(2 * i + 1) * v.table_slot + (t1 ? K * S4_MPAD : 0);
                    mbar_wait_backoff(&empty[s], ph ^ 1);
                    mbar_expect_tx(&full[s], 64 + b0 + b1 + (t0 ? st_bytes : 0) + (t1 ? st_bytes : 0));
                    bulk_g2s(st, ops + i, 64, &full[s]);
                    bulk_g2s(st + 64, g0, b0, &full[s]);
                    bulk_g2s(st + 64 + cbs * 8, g1, b1, &full[s]);
                    if (t0) bulk_g2s(st + 64 + 2 * cbs * 8, tip_states(v, src0) + a0, st_bytes, &full[s]);
                    if (t1) bulk_g2s(st + 64 + 2 * cbs * 8 + spad, tip_states(v, src1) + a0, st_bytes, &full[s]);
                }
            }
        }
        return;
    }

    /* ------------------------------ consumers ------------------------------ */
    const bool lane_on = lane < upw;
    const int k = lane_on ? lane % K : 0;
    const int gbase = lane_on ? lane - k : lane;
    const unsigned gmask = K >= 32 ? 0xffffffffu : ((1u << K) - 1u);
    const double2 *inv = reinterpret_cast<const double2 *>(sINV + k * S4_MPAD);
    const bool zero_guard = (v.flags & LKCUDA_FLAG_ZERO_GUARD) != 0, ignore_invar = (v.flags & LKCUDA_FLAG_SCALE_IGNORES_INVAR) != 0;
    double2 *const sl_t = sl + tid;
    int *const ssc_t = ssc + tid;
    const int B = v.B;
    int q[R];           /* my units' pattern index inside a tile */
#pragma unroll
    for (int r = 0; r < R; r++) q[r] = (r * cta_units + warp * upw + (lane_on ? lane : 0)) / K;
    uint32_t c = 0;     /* ring position of the node being computed, continues across tiles */
#if TRV_INV_REGS
    const double2 iv[8] = {inv[0], inv[1], inv[2], inv[3], inv[4], inv[5], inv[6], inv[7]};
#endif

    /* one tile = RA <= R slices, RA a compile-time constant: the R units of a thread are independent instruction streams the compiler
       interleaves, so the last round (fewer slices) runs its own instantiation instead of predicating slices off */
    auto do_tile = [&](auto ra_tag, const long long p0) {
        constexpr int RA = decltype(ra_tag)::value;
        long long p[RA], off[RA];     /* pattern, element offset of the unit inside a vector */
        bool valid[RA];
        const int sdelta = (int)(p0 & 15);
#pragma unroll
        for (int r = 0; r < RA; r++) {
            p[r] = p0 + q[r];
            valid[r] = lane_on && p[r] < v.NP;
            if (!valid[r]) p[r] = v.NP - 1;          /* out-of-range units shadow the last pattern; only their stores are masked */
            off[r] = p[r] * B + k * 4;
        }
        /* fused root branch: lane k of a pattern's lane group finishes the pattern of slice r = k (one log per lane); everything it needs
           from global memory is requested now and used after the last node */
        int root_st = 0; double root_inv = 0.0, root_f = 0.0; long long root_p = 0; bool root_mine = false;
        d4 pv[RA];       /* result of the previous node (TRV_SLOT_PREV) */
        int pvsc[RA];
#pragma unroll
        for (int r = 0; r < RA; r++) { pv[r].x = pv[r].y = pv[r].z = pv[r].w = 0.0; pvsc[r] = 0; }

        for (int i = 0; i < nops; i++, c++) {
            const uint32_t s = c & stage_mask;
            mbar_wait(&full[s], (c >> stage_shift) & 1);
            if (fuse_root && i == 0) {
                /* requested behind the tile's first stage: with a pipelined upload the producer has seen this pattern range arrive by then
                   (read around L1: the copy engine may have written the bytes while this kernel was already running) */
                const uint8_t *rst = tip_states(v, root.tip);
#pragma unroll
                for (int r = 0; r < RA; r++) {
                    root_st |= min((int)__ldcg(rst + p[r]), v.SU) << (8 * (r & 3));
                    if (r == k) { root_p = p[r]; root_mine = valid[r]; }
                }
                if (root_mine) { root_inv = v.invar[root_p]; root_f = root_p < v.P ? v.freq[root_p] : 0.0; }
            }
            const unsigned char *stage = ring + s * stage_bytes;
            const TravOp *op = reinterpret_cast<const TravOp *>(stage);
            const int src0 = op->src0, src1 = op->src1, slot0 = op->slot0, slot1 = op->slot1, dst_slot = op->dst_slot;
            const bool tip0 = LKCUDA_IS_TIP(src0), tip1 = LKCUDA_IS_TIP(src1);
            const double *blk0 = reinterpret_cast<const double *>(stage + 64), *blk1 = blk0 + cbs;
            const unsigned char *sst0 = stage + 64 + 2 * cbs * 8 + sdelta, *sst1 = sst0 + spad;
            double t[RA][4];
            int sc[RA];
            /* ---- child 0: tip | slot | arena ---- */
            if (tip0) {
                const double *tab = blk0 + k * 4;
#pragma unroll
                for (int r = 0; r < RA; r++) {
                    const int st = min((int)sst0[q[r]], v.SU);      /* clamped: an out-of-range code is reported by the upload's range check */
                    const double2 *row = reinterpret_cast<const double2 *>(tab + TRV_TAB_ROW(st, K));
                    const double2 a = row[0], b = row[1];
                    t[r][0] = a.x; t[r][1] = a.y; t[r][2] = b.x; t[r][3] = b.y; sc[r] = 0;
                }
            } else {
                const double2 *mp = reinterpret_cast<const double2 *>(blk0 + k * S4_MPAD);
#if TRV_FORM == 1
                if (slot0 >= 0) {
                    const double2 *ps = sl_t + slot0 * (R * 2 * CT);
                    const int *pc = ssc_t + slot0 * (R * CT);
                    {
                        const double2 c0 = mp[0], c1 = mp[2], c2 = mp[4], c3 = mp[6];
#pragma unroll
                        for (int r = 0; r < RA; r++) {
                            const double2 a = ps[(r * 2) * CT];
                            sc[r] = pc[r * CT];
                            TRV_COLS_LO(t[r], c0, c1, c2, c3, a)
                        }
                    }
                    {
                        const double2 c0 = mp[1], c1 = mp[3], c2 = mp[5], c3 = mp[7];
#pragma unroll
                        for (int r = 0; r < RA; r++) {
                            const double2 b = ps[(r * 2 + 1) * CT];
                            TRV_COLS_HI(t[r], c0, c1, c2, c3, b)
                        }
                    }
                } else {
                    const double *g = op->src0_lh;
                    const int16_t *gs = vec_sc(v, src0);
                    {
                        const double2 c0 = mp[0], c1 = mp[2], c2 = mp[4], c3 = mp[6];
#pragma unroll
                        for (int r = 0; r < RA; r++) {
                            const double2 a = ld_coherent128(g + off[r]);
                            sc[r] = gs[p[r]];
                            TRV_COLS_LO(t[r], c0, c1, c2, c3, a)
                        }
                    }
                    {
                        const double2 c0 = mp[1], c1 = mp[3], c2 = mp[5], c3 = mp[7];
#pragma unroll
                        for (int r = 0; r < RA; r++) {
                            const double2 b = ld_coherent128(g + off[r] + 2);
                            TRV_COLS_HI(t[r], c0, c1, c2, c3, b)
                        }
                    }
                }
#else
                const double2 m[8] = {mp[0], mp[1], mp[2], mp[3], mp[4], mp[5], mp[6], mp[7]};
                if (slot0 >= 0) {
                    const double2 *ps = sl_t + slot0 * (R * 2 * CT);
                    const int *pc = ssc_t + slot0 * (R * CT);
#pragma unroll
                    for (int r = 0; r < RA; r++) {
                        const double2 a = ps[(r * 2) * CT], b = ps[(r * 2 + 1) * CT];
                        d4 L; L.x = a.x; L.y = a.y; L.z = b.x; L.w = b.y;
                        sc[r] = pc[r * CT];
                        TRV_MATVEC(t[r], m, L)
                    }
                } else {
                    const double *g = op->src0_lh;
                    const int16_t *gs = vec_sc(v, src0);
#pragma unroll
                    for (int r = 0; r < RA; r++) {
                        const d4 L = ld_coherent256(g + off[r]);
                        sc[r] = gs[p[r]];
                        TRV_MATVEC(t[r], m, L)
                    }
                }
#endif
            }
            /* ---- child 1: tip | previous node (registers) | slot | arena ---- */
            if (tip1) {
                const double *tab = blk1 + k * 4;
#pragma unroll
                for (int r = 0; r < RA; r++) {
                    const int st = min((int)sst1[q[r]], v.SU);
                    const double2 *row = reinterpret_cast<const double2 *>(tab + TRV_TAB_ROW(st, K));
                    const double2 a = row[0], b = row[1];
                    t[r][0] *= a.x; t[r][1] *= a.y; t[r][2] *= b.x; t[r][3] *= b.y;
                }
            } else {
                const double2 *mp = reinterpret_cast<const double2 *>(blk1 + k * S4_MPAD);
#if TRV_FORM == 1
                if (slot1 == TRV_SLOT_PREV) {
#pragma unroll
                    for (int r = 0; r < RA; r++) sc[r] += pvsc[r];
#pragma unroll
                    for (int i4 = 0; i4 < 4; i4++) {
                        const double2 ra = mp[2 * i4], rb = mp[2 * i4 + 1];
#pragma unroll
                        for (int r = 0; r < RA; r++) t[r][i4] *= TRV_DOT(ra, rb, pv[r]);
                    }
                } else if (slot1 >= 0) {    /* rare (a vector with two consumers in the list): the slot is re-read for every row */
                    const double2 *ps = sl_t + slot1 * (R * 2 * CT);
                    const int *pc = ssc_t + slot1 * (R * CT);
#pragma unroll
                    for (int r = 0; r < RA; r++) sc[r] += pc[r * CT];
#pragma unroll
                    for (int i4 = 0; i4 < 4; i4++) {
                        const double2 ra = mp[2 * i4], rb = mp[2 * i4 + 1];
#pragma unroll
                        for (int r = 0; r < RA; r++) {
                            const double2 a = ps[(r * 2) * CT], b = ps[(r * 2 + 1) * CT];
                            d4 L; L.x = a.x; L.y = a.y; L.z = b.x; L.w = b.y;
                            t[r][i4] *= TRV_DOT(ra, rb, L);
                        }
                    }
                } else {
                    const double *g = op->src1_lh;
                    const int16_t *gs = vec_sc(v, src1);
#pragma unroll
                    for (int r = 0; r < RA; r++) sc[r] += gs[p[r]];
#pragma unroll
                    for (int i4 = 0; i4 < 4; i4++) {
                        const double2 ra = mp[2 * i4], rb = mp[2 * i4 + 1];
#pragma unroll
                        for (int r = 0; r < RA; r++) {
                            const d4 L = ld_coherent256(g + off[r]);
                            t[r][i4] *= TRV_DOT(ra, rb, L);
                        }
                    }
                }
#else
                const double2 m[8] = {mp[0], mp[1], mp[2], mp[3], mp[4], mp[5], mp[6], mp[7]};
                double u[4];
                if (slot1 == TRV_SLOT_PREV) {
#pragma unroll
                    for (int r = 0; r < RA; r++) {
                        sc[r] += pvsc[r];
                        TRV_MATVEC(u, m, pv[r])
                        t[r][0] *= u[0]; t[r][1] *= u[1]; t[r][2] *= u[2]; t[r][3] *= u[3];
                    }
                } else if (slot1 >= 0) {
                    const double2 *ps = sl_t + slot1 * (R * 2 * CT);
                    const int *pc = ssc_t + slot1 * (R * CT);
#pragma unroll
                    for (int r = 0; r < RA; r++) {
                        const double2 a = ps[(r * 2) * CT], b = ps[(r * 2 + 1) * CT];
                        d4 L; L.x = a.x; L.y = a.y; L.z = b.x; L.w = b.y;
                        sc[r] += pc[r * CT];
                        TRV_MATVEC(u, m, L)
                        t[r][0] *= u[0]; t[r][1] *= u[1]; t[r][2] *= u[2]; t[r][3] *= u[3];
                    }
                } else {
                    const double *g = op->src1_lh;
                    const int16_t *gs = vec_sc(v, src1);
#pragma unroll
                    for (int r = 0; r < RA; r++) {
                        const d4 L = ld_coherent256(g + off[r]);
                        sc[r] += gs[p[r]];
                        TRV_MATVEC(u, m, L)
                        t[r][0] *= u[0]; t[r][1] *= u[1]; t[r][2] *= u[2]; t[r][3] *= u[3];
                    }
                }
#endif
            }
            double *const dst_lh = op->dst_lh;
            int16_t *const dst_sc = op->dst_sc;
            /* everything this node needs from the stage is in registers now: hand the stage back to the producer */
#if TRV_ARRIVE_DEP == 2
            __threadfence_block();
#elif TRV_ARRIVE_DEP == 3
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
#endif
            __syncwarp();
#if TRV_ARRIVE_DEP == 1
            /* the barrier's address depends on the two descriptor words loaded last (always + 0): the arrive cannot issue before they are in */
            if (lane == 0) mbar_arrive(&empty[s] + (((unsigned long long)dst_lh | (unsigned long long)dst_sc) >> 63));
#else
            if (lane == 0) mbar_arrive(&empty[s]);
#endif
            /* ---- back to state space, scaling test, stores ---- */
            const bool may_scale = !(tip0 && tip1);   /* tip-tip never scales (phylokernel.h:246-281) */
            bool tiny = false;
#if TRV_INV_FORM == 1
#define TRV_INV_ROW(i4, comp) do {                                               \
                const double2 ra = inv[2 * (i4)], rb = inv[2 * (i4) + 1];         \
                _Pragma("unroll")                                                 \
                for (int r = 0; r < RA; r++) {                                    \
                    const d4 T = {t[r][0], t[r][1], t[r][2], t[r][3]};            \
                    pv[r].comp = TRV_DOT(ra, rb, T);                              \
                }                                                                 \
            } while (0)
            TRV_INV_ROW(0, x); TRV_INV_ROW(1, y); TRV_INV_ROW(2, z); TRV_INV_ROW(3, w);
#undef TRV_INV_ROW
#pragma unroll
            for (int r = 0; r < RA; r++) {
                const int mx = max(max(__double2hiint(pv[r].x) & 0x7fffffff, __double2hiint(pv[r].y) & 0x7fffffff),
                                   max(__double2hiint(pv[r].z) & 0x7fffffff, __double2hiint(pv[r].w) & 0x7fffffff));
                tiny |= mx < THR_HI;
            }
#else
#if !TRV_INV_REGS
            const double2 iv[8] = {inv[0], inv[1], inv[2], inv[3], inv[4], inv[5], inv[6], inv[7]};
#endif
#pragma unroll
            for (int r = 0; r < RA; r++) {
                const d4 T = {t[r][0], t[r][1], t[r][2], t[r][3]};
                double w[4];
                TRV_MATVEC(w, iv, T)
                pv[r].x = w[0]; pv[r].y = w[1]; pv[r].z = w[2]; pv[r].w = w[3];
                /* lh_max < 2^-256 is decided on the exponent words: integer pipe, not the FP64 pipe */
                const int mx = max(max(__double2hiint(w[0]) & 0x7fffffff, __double2hiint(w[1]) & 0x7fffffff),
                                   max(__double2hiint(w[2]) & 0x7fffffff, __double2hiint(w[3]) & 0x7fffffff));
                tiny |= mx < THR_HI;
            }
#endif
            if (may_scale && __any_sync(0xffffffffu, tiny)) {
                /* rare path: some unit of this warp is below the threshold; agree per pattern (lane group) with ballots */
#pragma unroll
                for (int r = 0; r < RA; r++) {
                    const int h0 = __double2hiint(pv[r].x) & 0x7fffffff, h1 = __double2hiint(pv[r].y) & 0x7fffffff,
                              h2 = __double2hiint(pv[r].z) & 0x7fffffff, h3 = __double2hiint(pv[r].w) & 0x7fffffff;
                    const unsigned small = __ballot_sync(0xffffffffu, max(max(h0, h1), max(h2, h3)) < THR_HI);
                    const bool nz = (h0 | h1 | h2 | h3 | __double2loint(pv[r].x) | __double2loint(pv[r].y) | __double2loint(pv[r].z) | __double2loint(pv[r].w)) != 0;
                    const unsigned anynz = __ballot_sync(0xffffffffu, nz);
                    if (((small >> gbase) & gmask) == gmask) {
                        if (zero_guard && ((anynz >> gbase) & gmask) == 0) {
                            const double *tu = v.tip + ((long long)v.SU * v.E + v.blk_eigen[k]) * 4;
                            pv[r].x = tu[0]; pv[r].y = tu[1]; pv[r].z = tu[2]; pv[r].w = tu[3];
                            sc[r] += 4;
                        } else if (ignore_invar || v.invar[p[r]] == 0.0) {
                            pv[r].x *= LK_THR_INV; pv[r].y *= LK_THR_INV; pv[r].z *= LK_THR_INV; pv[r].w *= LK_THR_INV;
                            sc[r] += 1;
                        }
                    }
                }
            }
#pragma unroll
            for (int r = 0; r < RA; r++) {
                pvsc[r] = sc[r];
                st_stream256_if(valid[r], dst_lh + off[r], pv[r]);
                st_s16_if(valid[r] && k == 0, dst_sc + p[r], sc[r]);
            }
            if (dst_slot >= 0) {   /* parked for a later node of this launch */
                double2 *pd = sl_t + dst_slot * (R * 2 * CT);
                int *pdc = ssc_t + dst_slot * (R * CT);
#pragma unroll
                for (int r = 0; r < RA; r++) {
                    pd[(r * 2) * CT] = make_double2(pv[r].x, pv[r].y);
                    pd[(r * 2 + 1) * CT] = make_double2(pv[r].z, pv[r].w);
                    pdc[r * CT] = sc[r];
                }
            }
            /* spilled result: its consumer re-reads scale_num, which lane k == 0 wrote for the whole lane group */
            if (dst_slot == -2) { __threadfence_block(); __syncwarp(); }
            if (fuse_root && i == nops - 1) {
                /* the last node is the inner end of the root branch and still in registers: lh = sum_k sum_i val[k][i] * tip[state][i] * L[k][i]
                   over the 4 lanes of the pattern (every lane ends up with the total), then lane k runs branch_tip_b16_kernel's tail for slice k */
                double lh_mine = 0.0; int sc_mine = 0;
#pragma unroll
                for (int r = 0; r < RA; r++) {
                    const int st = (root_st >> (8 * r)) & 255;
                    const double2 *row = reinterpret_cast<const double2 *>(rootB + st * TRV_ROOT_ROW + 4 * k);
                    const double2 a = row[0], b = row[1];
                    double lh = a.x * pv[r].x;
                    lh = fma(a.y, pv[r].y, lh); lh = fma(b.x, pv[r].z, lh); lh = fma(b.y, pv[r].w, lh);
                    lh += __shfl_xor_sync(0xffffffffu, lh, 1);
                    lh += __shfl_xor_sync(0xffffffffu, lh, 2);
                    if (r == k) { lh_mine = lh; sc_mine = sc[r]; }       /* scale_num is per pattern: the group's lanes all carry it */
                }
                if (root_mine) {
                    double a4[4] = {0.0, 0.0, 0.0, 0.0};
                    lnl_tail_vals(v, root_p, lh_mine, sc_mine, root_inv, root_f, a4);
#pragma unroll
                    for (int j = 0; j < 4; j++) racc[j * CT + tid] += a4[j];
                }
            }
        }
    };
    for (int round = 0; round < nrounds; round++) {
        const long long p0 = trv_tile_start(round, nfull, ract_last, tile_pats, slice_pats);
        if (p0 >= v.NP) break;
        if (round < nfull) { do_tile(std::integral_constant<int, R>{}, p0); continue; }
        switch (ract_last) {
            case 1: do_tile(std::integral_constant<int, 1>{}, p0); break;
            case 2: do_tile(std::integral_constant<int, (R >= 2 ? 2 : R)>{}, p0); break;
            case 3: do_tile(std::integral_constant<int, (R >= 3 ? 3 : R)>{}, p0); break;
            case 4: do_tile(std::integral_constant<int, (R >= 4 ? 4 : R)>{}, p0); break;
            case 5: do_tile(std::integral_constant<int, (R >= 5 ? 5 : R)>{}, p0); break;
            default: do_tile(std::integral_constant<int, R>{}, p0); break;
        }
    }
    if (fuse_root) {
        double acc[4] = {racc[tid], racc[CT + tid], racc[2 * CT + tid], racc[3 * CT + tid]};
        if (grid_reduce_part<4>(acc, rred, v.red, v.counter, CT)) lnl_finish(v, acc, acc[1]);
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

static_assert(sizeof(TravOp) == 64 && sizeof(TravOp) >= sizeof(DevOp), "the op ring is sized in TravOp units");

static bool use_traverse(const DevView &v) { return v.S == 4 && v.K <= 32; }

/*
 * Order a batch of nodes for the whole-traversal kernel and decide where each node finds its children:
 *   - the child executed immediately before its parent stays in REGISTERS (TRV_SLOT_PREV) - in a
 *     post-order schedule that is one child of every internal-internal node and the only internal child of
 *     every tip-internal node;
 *   - the other internal child of an internal-internal node is parked in a shared-memory SLOT.
 *     Sethi-Ullman: evaluating the child that needs more slots first, a node needs max(a, b + 1) slots
 *     (a >= b the children's needs), so a tree with n leaves never needs more than ~log2(n) - 1;
 *   - when the slots run out the result is not parked (dst_slot = -2) and its consumer reads it back from
 *     its arena vector, like a child that was computed by an earlier launch.
 * Multiplication commutes exactly in IEEE arithmetic, so the two children may be swapped freely.
 */
int lkc::schedule_traverse(lkcuda_engine *e, const lkcuda_partial_op *ops, int nops, int maxslots, bool use_prev, TravSchedule &sch) {
    std::vector<int> producer(e->cfg.nvec, -1);
    for (int i = 0; i < nops; i++) {
        if (producer[ops[i].dst] >= 0) return fail(e, LKCUDA_ERR_ARG, "partial ops %d and %d write the same vector", producer[ops[i].dst], i);
        producer[ops[i].dst] = i;
    }
    std::vector<int> kid0(nops), kid1(nops), nconsumer(nops, 0), need(nops, 0), state(nops, 0), last_kid(nops, 0), pos(nops, -1);
    for (int i = 0; i < nops; i++) {
        kid0[i] = ops[i].left >= 0 ? producer[ops[i].left] : -1;
        kid1[i] = ops[i].right >= 0 ? producer[ops[i].right] : -1;
        if (kid0[i] >= 0) nconsumer[kid0[i]]++;
        if (kid1[i] >= 0) nconsumer[kid1[i]]++;
    }
    /* pass 1: slot need of every node (iterative post-order; state 0 = new, 1 = open, 2 = done) */
    std::vector<int> stack;
    for (int root = 0; root < nops; root++) {
        if (state[root]) continue;
        stack.push_back(root);
        while (!stack.empty()) {
            int i = stack.back();
            if (state[i] == 0) {
                state[i] = 1;
                if (kid0[i] >= 0) { if (state[kid0[i]] == 1) return fail(e, LKCUDA_ERR_ARG, "partial ops form a cycle"); if (!state[kid0[i]]) stack.push_back(kid0[i]); }
                if (kid1[i] >= 0) { if (state[kid1[i]] == 1) return fail(e, LKCUDA_ERR_ARG, "partial ops form a cycle"); if (!state[kid1[i]]) stack.push_back(kid1[i]); }
            } else {
                stack.pop_back();
                if (state[i] == 2) continue;
                state[i] = 2;
                if (kid0[i] >= 0 && kid1[i] >= 0) {
                    int a = need[kid0[i]], b = need[kid1[i]];
                    int hi = a > b ? a : b, lo = a > b ? b : a;
                    need[i] = hi > lo + 1 ? hi : lo + 1;
                } else need[i] = kid0[i] >= 0 ? need[kid0[i]] : (kid1[i] >= 0 ? need[kid1[i]] : 0);
            }
        }
    }
    /* pass 2: emit, needier child first; slots from a free list */
    sch.nodes.clear(); sch.nodes.reserve(nops);
    std::vector<int> free_slots, slot_of(nops, -1);
    for (int sidx = maxslots - 1; sidx >= 0; sidx--) free_slots.push_back(sidx);
    int used = 0;
    std::fill(state.begin(), state.end(), 0);
    for (int root = 0; root < nops; root++) {
        if (nconsumer[root] || state[root]) continue;       /* start from the nodes nobody in the batch consumes */
        stack.push_back(root);
        while (!stack.empty()) {
            int i = stack.back();
            if (state[i] == 0) {
                state[i] = 1;
                int first = kid0[i], second = kid1[i];
                if (first >= 0 && second >= 0 && need[second] > need[first]) { int t = first; first = second; second = t; }
                if (second < 0) { second = first; first = -1; }
                if (second >= 0 && first == second) first = -1;      /* both children are the same vector */
                /* pushed in reverse: `first` is expanded (and finished) before `second`, which ends right before i */
                if (second >= 0 && !state[second]) { stack.push_back(second); last_kid[second] = use_prev && nconsumer[second] == 1; }
                if (first >= 0 && !state[first]) stack.push_back(first);
            } else {
                stack.pop_back();
                if (state[i] == 2) continue;
                state[i] = 2;
                const int j = (int)sch.nodes.size();
                pos[i] = j;
                TravNode n;
                n.op = i; n.dst_slot = -1;
                int k0 = kid0[i], k1 = kid1[i];
                n.src0 = ops[i].left; n.len0 = ops[i].len_left; n.src1 = ops[i].right; n.len1 = ops[i].len_right;
                n.ref0 = ops[i].reserved & 0xffff; n.ref1 = (ops[i].reserved >> 16) & 0xffff;
                /* the child executed right before this node goes to position 1 and is taken from registers */
                if (k0 >= 0 && k0 != k1 && pos[k0] == j - 1 && last_kid[k0] && !(k1 >= 0 && pos[k1] == j - 1)) {
                    std::swap(n.src0, n.src1); std::swap(n.len0, n.len1); std::swap(n.ref0, n.ref1); std::swap(k0, k1);
                }
                n.slot0 = n.slot1 = -1;
                if (k0 >= 0) { n.slot0 = slot_of[k0]; }
                if (k1 >= 0) {
                    if (pos[k1] == j - 1 && last_kid[k1] && k1 != k0) n.slot1 = TRV_SLOT_PREV;
                    else n.slot1 = slot_of[k1];
                }
                if (k0 >= 0 && slot_of[k0] >= 0 && --nconsumer[k0] == 0) { free_slots.push_back(slot_of[k0]); slot_of[k0] = -1; }
                else if (k0 >= 0 && slot_of[k0] < 0) --nconsumer[k0];
                if (k1 >= 0 && k1 != k0) {
                    if (slot_of[k1] >= 0 && --nconsumer[k1] == 0) { free_slots.push_back(slot_of[k1]); slot_of[k1] = -1; }
                    else if (slot_of[k1] < 0) --nconsumer[k1];
                } else if (k1 >= 0 && k1 == k0) {
                    if (slot_of[k1] >= 0 && --nconsumer[k1] == 0) { free_slots.push_back(slot_of[k1]); slot_of[k1] = -1; }
                    else if (slot_of[k1] < 0) --nconsumer[k1];
                }
                if (nconsumer[i] > 0 && !last_kid[i]) {
                    if (!free_slots.empty()) {
                        slot_of[i] = free_slots.back(); free_slots.pop_back();
                        n.dst_slot = slot_of[i];
                        if (slot_of[i] + 1 > used) used = slot_of[i] + 1;
                    } else n.dst_slot = -2;     /* spilled: the consumer reads the arena vector */
                }
                sch.nodes.push_back(n);
            }
        }
    }
    if ((int)sch.nodes.size() != nops) return fail(e, LKCUDA_ERR_ARG, "partial ops form a cycle");
    sch.nslots = used;
    return LKCUDA_OK;
}

int lkc::ring_acquire(lkcuda_engine *e, int nops, int nchild, int *slot_out) {
    int s = e->ring_next;
    e->ring_next = (s + 1) % lkcuda_engine::RING;
    if (!e->ring_ev[s]) LK_CUDA(e, cudaEventCreateWithFlags(&e->ring_ev[s], cudaEventDisableTiming));
    if (e->ring_used[s]) LK_CUDA(e, cudaEventSynchronize(e->ring_ev[s]));
    if (nops > e->ring_cap_ops[s]) {
        if (e->h_ops[s]) cudaFreeHost(e->h_ops[s]);
        cudaFree(e->d_ops[s]);
        int cap = nops < 256 ? 256 : nops * 2;
        LK_CUDA(e, cudaMallocHost((void **)&e->h_ops[s], sizeof(TravOp) * cap));
        LK_CUDA(e, cudaMalloc((void **)&e->d_ops[s], sizeof(TravOp) * cap));
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
uint64_t lkc::op_bytes(const DevView &v, int ntip, int nint) {
    uint64_t b = (uint64_t)v.B * 8;
    return (uint64_t)v.NP * ((uint64_t)(nint + 1) * b + (uint64_t)ntip * 1 + (uint64_t)(nint + 1) * 2);
}

/* algorithmic flops of one node: 2*S^2 per internal child and for the way back to state space, S per product; per pattern and block */
uint64_t lkc::op_flops(const DevView &v, int ntip, int nint) {
    const uint64_t S = (uint64_t)v.S;
    return (uint64_t)v.NP * (uint64_t)v.K * (2 * S * S * (uint64_t)(nint + 1) + S * (uint64_t)(ntip + nint - 1));
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
    rc = states_barrier(e);
    if (rc) return rc;

    launch_begin(e, KIND_PREP);
    prep_tables_kernel<<<nchild_total, 256, (size_t)v.B * 8, e->stream>>>(e->v, e->d_child[slot], 0);
    rc = launch_end(e, KIND_PREP, 0, "prep_tables_kernel");
    if (rc) return rc;

    size_t gsm = generic_smem(v);
    if (gsm > 48 * 1024) {
        if (gsm > 227 * 1024) return fail(e, LKCUDA_ERR_ARG, "state space too large for the generic kernel (%d states)", v.S);
        LK_CUDA(e, cudaFuncSetAttribute(partial_generic_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, LK_MAX_DYN_SMEM));
    }
    for (int l = 0; l < nlevels; l++) {
        int b = level_start[l], n = level_start[l + 1] - b;
        if (n <= 0) continue;
        if (n > 65535) return fail(e, LKCUDA_ERR_ARG, "more than 65535 nodes in one level");
        launch_begin(e, KIND_PARTIAL);
        dim3 grid((unsigned)((v.NP + GEN_TP - 1) / GEN_TP), (unsigned)n);
        partial_generic_kernel<<<grid, GEN_THREADS, gsm, e->stream>>>(e->v, e->d_ops[slot], e->d_child[slot], b);
        rc = launch_end(e, KIND_PARTIAL, bytes_per_level[l], "partial kernel");
        if (rc) return rc;
    }
    return LKCUDA_OK;
}

/* the whole batch in ONE launch of the whole-traversal kernel (+ one table-preparation launch) */
static int launch_traverse(lkcuda_engine *e, const lkcuda_partial_op *ops, int nops) {
    const DevView &v = e->v;
    /* child block of a ring stage: a padded matrix set or a tip lookup table, whichever is larger */
    const int cbs_mat = v.K * S4_MPAD, cbs_tab = TRV_TAB_DOUBLES(v.SU, v.K), cbs = cbs_mat > cbs_tab ? cbs_mat : cbs_tab;
    /* ring depth: 8 stages when they fit beside the slots the schedule is likely to need, else 4, else 2 */
    /* the root branch rides along when the caller asked for it (lkcuda_partial_batch_lnl), this is the list's last piece, the leaf end
       is a tip and the shape is DNA + 4 categories; whether the LAST scheduled node is the branch's inner end is checked below */
    bool fuse = TRV_R_ARENA <= 4 && e->fuse_state == 2 && v.K == 4 && v.SU < 32 && LKCUDA_IS_TIP(e->fuse_dad) && v.NP > 0;
    /* units per thread: children that are vectors of EARLIER launches are streamed from the arena, and a list made of those runs better
       with fewer units per thread (see TRV_R) */
    int narena = 0;
    {
        std::vector<char> &made = e->trav_mark;
        if ((int)made.size() < e->cfg.nvec) made.assign(e->cfg.nvec, 0);
        for (int j = 0; j < nops; j++) made[ops[j].dst] = 1;
        for (int j = 0; j < nops; j++) {
            if (ops[j].left >= 0 && !made[ops[j].left]) narena++;
            if (ops[j].right >= 0 && !made[ops[j].right]) narena++;
        }
        for (int j = 0; j < nops; j++) made[ops[j].dst] = 0;
    }
    const int R = (fuse || TRV_R == TRV_R_ARENA || 8 * narena > nops) ? TRV_R_ARENA : TRV_R;
    const size_t slot_bytes = (size_t)R * TRV_CTHREADS * (32 + 4);
    const int upw = (32 / v.K) * v.K, tile_pats = TRV_CWARPS * upw * R / v.K, spad = (tile_pats + 15 + 15) & ~15;
    const size_t other = (size_t)v.K * S4_MPAD * 8 + 2 * TRV_MAX_STAGES * 8 + 128 + (fuse ? TRV_ROOT_SMEM : 0);
    const size_t stage_b = (size_t)trv_stage_bytes(cbs, spad);
    int stage_shift = 3;
    while (stage_shift > 1 && (size_t)(1 << stage_shift) * stage_b + other + 3 * slot_bytes > 227 * 1024) stage_shift--;
    const size_t fixed = (size_t)(1 << stage_shift) * stage_b + other;
    if (fixed > 227 * 1024) return fail(e, LKCUDA_ERR_ARG, "too many categories for the DNA traversal kernel");
    int maxslots = (int)((227 * 1024 - fixed) / slot_bytes);
    if (e->trav_maxslots >= 0 && e->trav_maxslots < maxslots) maxslots = e->trav_maxslots;   /* test knob: forces the spill path */
    TravSchedule sch;
    int rc = schedule_traverse(e, ops, nops, maxslots, true, sch);
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
    TravOp *d = reinterpret_cast<TravOp *>(e->d_ops[slot]);
    LK_CUDA(e, cudaMemcpyAsync(d, h, sizeof(TravOp) * nops, cudaMemcpyHostToDevice, e->stream));
    LK_CUDA(e, cudaMemcpyAsync(e->d_child[slot], e->h_child[slot], sizeof(DevChild) * 2 * nops, cudaMemcpyHostToDevice, e->stream));
    LK_CUDA(e, cudaEventRecord(e->ring_ev[slot], e->stream));
    e->ring_used[slot] = true;
    if (v.NP == 0) return LKCUDA_OK;

    launch_begin(e, KIND_PREP);
    prep_tables_kernel<<<2 * nops, 256, (size_t)v.B * 8, e->stream>>>(e->v, e->d_child[slot], S4_MPAD - 16);
    rc = launch_end(e, KIND_PREP, 0, "prep_tables_kernel");
    if (rc) return rc;

    /* work is dealt in SLICES (one unit per consumer thread = slice_pats patterns): every CTA runs nfull rounds of TRV_R slices and
       one last round of ract_last <= TRV_R slices, so the cost of the launch is ceil-ed to one slice per CTA, not to one whole tile */
    const int slice_pats = tile_pats / R;
    const long long nslices = (v.NP + slice_pats - 1) / slice_pats;
    long long grid = e->num_sms;      /* one persistent CTA per SM (384 threads, shared memory bound) */
    if (grid > nslices) grid = nslices;
    const int nfull = (int)(nslices / (R * grid));
    const long long rem = nslices - (long long)nfull * R * grid;
    const int ract_last = (int)((rem + grid - 1) / grid);
    const size_t sm = (size_t)sch.nslots * slot_bytes + fixed;
    if (e->trav_smem_set != LK_MAX_DYN_SMEM) {
        LK_CUDA(e, cudaFuncSetAttribute(traverse_s4_kernel<TRV_R, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, LK_MAX_DYN_SMEM));
        LK_CUDA(e, cudaFuncSetAttribute(traverse_s4_kernel<TRV_R_ARENA, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, LK_MAX_DYN_SMEM));
        LK_CUDA(e, cudaFuncSetAttribute(traverse_s4_kernel<TRV_R_ARENA, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, LK_MAX_DYN_SMEM));
        e->trav_smem_set = LK_MAX_DYN_SMEM;
    }
    TravRoot root;
    root.on = 0; root.tip = 0; root.len = 0.0; root.tab = e->b16tab;
    if (fuse && ops[sch.nodes[nops - 1].op].dst == e->fuse_node) {
        root.on = 1; root.tip = e->fuse_dad; root.len = e->fuse_len;
        next_result(e);
        e->fuse_state = 3;
        bytes += (uint64_t)v.NP * 25;       /* states, ptn_freq, ptn_invar read, _pattern_lh written: the vector itself is not re-read */
    }
    launch_begin(e, KIND_PARTIAL);
    /* with a pipelined upload in flight the kernel is NOT ordered after it: its producer polls the per-chunk counter instead */
    const unsigned int *flag = e->upload_pending ? e->d_upload_flag : NULL;
    if (root.on)
        traverse_s4_kernel<TRV_R_ARENA, true><<<(unsigned)grid, TRV_THREADS, sm, e->stream>>>(e->v, d, nops, upw, sch.nslots, cbs, spad, stage_shift, nfull, ract_last,
                                                                                flag, e->upload_base, e->upload_chunk_pat > 0 ? e->upload_chunk_pat : 1, root);
    else if (R == TRV_R_ARENA)
        traverse_s4_kernel<TRV_R_ARENA, false><<<(unsigned)grid, TRV_THREADS, sm, e->stream>>>(e->v, d, nops, upw, sch.nslots, cbs, spad, stage_shift, nfull, ract_last,
                                                                                                flag, e->upload_base, e->upload_chunk_pat > 0 ? e->upload_chunk_pat : 1, root);
    else
        traverse_s4_kernel<TRV_R, false><<<(unsigned)grid, TRV_THREADS, sm, e->stream>>>(e->v, d, nops, upw, sch.nslots, cbs, spad, stage_shift, nfull, ract_last,
                                                                                 flag, e->upload_base, e->upload_chunk_pat > 0 ? e->upload_chunk_pat : 1, root);
    return launch_end(e, KIND_PARTIAL, bytes, "traverse_s4_kernel");
}

/* one hazard-free group of nodes: every vector is written at most once and never written after it was read */
static int partial_batch_segment(lkcuda_engine *e, const lkcuda_partial_op *ops, int32_t nops,
                                 const int32_t *level_start, int32_t nlevels) {
    const bool sitemodel = (e->v.flags & LKCUDA_FLAG_SITEMODEL) != 0;
    if (!sitemodel && use_traverse(e->v)) {
        e->theta_ready = false;
        return launch_traverse(e, ops, nops);
    }
    if (!sitemodel && use_traverse_mma(e)) {
        e->theta_ready = false;
        return launch_traverse_mma(e, ops, nops);
    }
    if (!sitemodel && use_traverse_mmak(e)) {
        e->theta_ready = false;
        return launch_traverse_mmak(e, ops, nops);
    }
    /* level-batched kernels: without caller-supplied levels, level(op) = 1 + max level of the children produced in this group */
    std::vector<lkcuda_partial_op> sorted;
    std::vector<int32_t> starts;
    if (!level_start && !sitemodel) {
        std::vector<int> producer(e->cfg.nvec, -1), lvl(nops, 1);
        int maxlevel = 1;
        for (int i = 0; i < nops; i++) {
            const int a = ops[i].left >= 0 ? producer[ops[i].left] : -1, b = ops[i].right >= 0 ? producer[ops[i].right] : -1;
            if (a >= 0 && lvl[a] + 1 > lvl[i]) lvl[i] = lvl[a] + 1;
            if (b >= 0 && lvl[b] + 1 > lvl[i]) lvl[i] = lvl[b] + 1;
            if (lvl[i] > maxlevel) maxlevel = lvl[i];
            producer[ops[i].dst] = i;
        }
        starts.assign(maxlevel + 1, 0);
        for (int i = 0; i < nops; i++) starts[lvl[i]]++;
        int32_t acc = 0;
        for (int l = 1; l <= maxlevel; l++) { const int32_t c = starts[l]; starts[l] = acc; acc += c; }
        std::vector<int32_t> cursor(starts);
        sorted.resize(nops);
        for (int i = 0; i < nops; i++) sorted[cursor[lvl[i]]++] = ops[i];       /* stable inside a level */
        for (int l = 1; l <= maxlevel; l++) starts[l - 1] = starts[l];
        starts[maxlevel] = nops;
        ops = &sorted[0]; level_start = &starts[0]; nlevels = maxlevel;
    } else if (!level_start) {
        starts.resize(nops + 1);
        for (int i = 0; i <= nops; i++) starts[i] = i;
        level_start = &starts[0]; nlevels = nops;
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
        a.id = ops[i].left; a.pad = ops[i].reserved & 0xffff; a.len = ops[i].len_left;
        b.id = ops[i].right; b.pad = (ops[i].reserved >> 16) & 0xffff; b.len = ops[i].len_right;
        int ntip = (LKCUDA_IS_TIP(a.id) ? 1 : 0) + (LKCUDA_IS_TIP(b.id) ? 1 : 0);
        lbytes[l] += op_bytes(e->v, ntip, 2 - ntip);
        e->kind_flops[KIND_PARTIAL] += op_flops(e->v, ntip, 2 - ntip);
    }
    e->theta_ready = false;
    if (sitemodel) return launch_traverse_sitemodel(e, slot, nops);
    return launch_ops(e, slot, nops, 2 * nops, level_start, nlevels, &lbytes[0]);
}

extern "C" {

int lkcuda_partial_batch(lkcuda_engine *e, const lkcuda_partial_op *ops, int32_t nops,
                         const int32_t *level_start, int32_t nlevels) {
    LK_ENTER(e);
    if (nops < 0 || (nops && !ops)) return fail(e, LKCUDA_ERR_ARG, "bad op list");
    if (!e->model_set || !e->states_set || !e->weights_set) return fail(e, LKCUDA_ERR_STATE, "lkcuda_partial_batch: tip states, pattern weights and model must be set first");
    if (nops == 0) return LKCUDA_OK;
    if (level_start) {
        if (nlevels < 0 || level_start[0] != 0 || level_start[nlevels] != nops) return fail(e, LKCUDA_ERR_ARG, "bad level_start");
        for (int l = 0; l < nlevels; l++) if (level_start[l] > level_start[l + 1]) return fail(e, LKCUDA_ERR_ARG, "bad level_start");
    }
    for (int i = 0; i < nops; i++) {
        const lkcuda_partial_op &op = ops[i];
        if (op.dst < 0 || op.dst >= e->cfg.nvec || !valid_end(e, op.left) || !valid_end(e, op.right))
            return fail(e, LKCUDA_ERR_ARG, "partial op %d: bad vector / taxon id", i);
        if (!LKCUDA_IS_TIP(op.left) && LKCUDA_IS_TIP(op.right))
            return fail(e, LKCUDA_ERR_ARG, "partial op %d: a leaf child must be 'left' (phylokernel.h:116-121)", i);
        if (op.dst == op.left || op.dst == op.right) return fail(e, LKCUDA_ERR_ARG, "partial op %d: destination aliases a child", i);
        if ((op.reserved & 0xffff) > LKCUDA_NLENSLOT || ((op.reserved >> 16) & 0xffff) > LKCUDA_NLENSLOT) return fail(e, LKCUDA_ERR_ARG, "partial op %d: bad length slot", i);
    }
    /* A vector may be the destination of several ops of one list: under the reference's default LM_PER_NODE buffer mode a node's single
       vector is re-oriented on demand (phylonode.cpp:22-39, phylokernel.h:127-143), so a later op may overwrite what an earlier op of
       the same list wrote or read.  With the ops in EXECUTION order (level_start == NULL) the list is cut at every write-after-write /
       write-after-read hazard and the pieces run back to back; a caller-supplied level grouping has no defined order across such a
       hazard and is refused. */
    std::vector<unsigned char> mark(e->cfg.nvec, 0);        /* 1 = written, 2 = read in the current piece */
    int begin = 0;
    for (int i = 0; i <= nops; i++) {
        const bool cut = i == nops || mark[ops[i].dst] != 0;
        if (cut) {
            if (i < nops && level_start) return fail(e, LKCUDA_ERR_ARG, "partial op %d overwrites a vector an earlier op of the list wrote or read: pass the ops in execution order with level_start = NULL", i);
            if (i > begin) {
                if (e->fuse_state == 1 && i == nops) e->fuse_state = 2;     /* the last piece may carry the root branch */
                int rc = level_start ? partial_batch_segment(e, ops, nops, level_start, nlevels) : partial_batch_segment(e, ops + begin, i - begin, NULL, 0);
                if (rc) return rc;
            }
            if (i == nops) break;
            for (int j = begin; j < i; j++) {
                mark[ops[j].dst] = 0;
                if (ops[j].left >= 0) mark[ops[j].left] = 0;
                if (ops[j].right >= 0) mark[ops[j].right] = 0;
            }
            begin = i;
        }
        mark[ops[i].dst] |= 1;
        if (ops[i].left >= 0) mark[ops[i].left] |= 2;
        if (ops[i].right >= 0) mark[ops[i].right] |= 2;
    }
    return LKCUDA_OK;
}

/* lkcuda_partial_batch followed by lkcuda_branch_lnl(dad, node, len): what PhyloTree::computeLikelihood asks for (phylotree.cpp:1031-1071:
   the traversal towards the root branch, then computeLikelihoodBranch on it).  When the list's last node is the branch's inner end, the
   other end is a leaf and the shape is DNA + 4 categories, the branch is evaluated in the TAIL of the traversal launch from the registers
   that still hold the node (no re-read of the vector, no second launch); every other case runs the two calls back to back. */
int lkcuda_partial_batch_lnl(lkcuda_engine *e, const lkcuda_partial_op *ops, int32_t nops, const int32_t *level_start, int32_t nlevels,
                             int32_t dad, int32_t node, double len, double *result) {
    LK_ENTER(e);
    if (node < 0 || node >= e->cfg.nvec || !valid_end(e, dad) || dad == node) return fail(e, LKCUDA_ERR_ARG, "lkcuda_partial_batch_lnl: bad vector / taxon id");
    /* OFF unless LKCUDA_FUSED_LNL=1: measured on B200 (profiles/r2_s16_fused_root_lnl.txt) the traversal instantiation that carries the
       branch is slower than the plain one by more than branch_tip_b16_kernel costs (1M patterns: 2.52 vs 2.44 ms per evaluation) */
    static const bool on = getenv("LKCUDA_FUSED_LNL") && atoi(getenv("LKCUDA_FUSED_LNL")) != 0;
    const bool off = !on;
    e->fuse_state = (!off && nops > 0 && LKCUDA_IS_TIP(dad) && !(e->v.flags & LKCUDA_FLAG_SITEMODEL)) ? 1 : 0;
    e->fuse_dad = dad; e->fuse_node = node; e->fuse_len = len;
    int rc = lkcuda_partial_batch(e, ops, nops, level_start, nlevels);
    const bool done = e->fuse_state == 3;
    e->fuse_state = 0;
    if (rc) return rc;
    if (!done) return lkcuda_branch_lnl(e, dad, node, len, result);
    return lkc::branch_lnl_followups(e, dad, node, len, result);
}

int lkcuda_partial_multi(lkcuda_engine *e, int32_t dst, int32_t nchild, const int32_t *child, const double *child_len) {
    LK_ENTER(e);
    if (!child || !child_len || nchild < 2 || dst < 0 || dst >= e->cfg.nvec) return fail(e, LKCUDA_ERR_ARG, "bad multifurcation arguments");
    if (!e->model_set || !e->states_set || !e->weights_set) return fail(e, LKCUDA_ERR_STATE, "lkcuda_partial_multi: inputs not set");
    for (int c = 0; c < nchild; c++) if (!valid_end(e, child[c]) || child[c] == dst) return fail(e, LKCUDA_ERR_ARG, "bad child");
    if (e->v.flags & LKCUDA_FLAG_SITEMODEL) return fail(e, LKCUDA_ERR_ARG, "site-specific models need a bifurcating tree (phylokernelsitemodel.cpp:86)");
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
    e->kind_flops[KIND_PARTIAL] += op_flops(e->v, ntip, nchild - ntip);
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
    rc = states_barrier(e);
    if (rc) return rc;
    launch_begin(e, KIND_PREP);
    prep_tables_kernel<<<nchild, 256, (size_t)v.B * 8, e->stream>>>(e->v, e->d_child[slot], 0);
    rc = launch_end(e, KIND_PREP, 0, "prep_tables_kernel");
    if (rc) return rc;
    size_t gsm = generic_smem(v);
    if (gsm > 227 * 1024) return fail(e, LKCUDA_ERR_ARG, "state space too large for the generic kernel (%d states)", v.S);
    if (gsm > 48 * 1024) LK_CUDA(e, cudaFuncSetAttribute(partial_generic_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, LK_MAX_DYN_SMEM));
    launch_begin(e, KIND_PARTIAL);
    dim3 grid((unsigned)((v.NP + GEN_TP - 1) / GEN_TP), 1);
    partial_generic_kernel<<<grid, GEN_THREADS, gsm, e->stream>>>(e->v, e->d_ops[slot], e->d_child[slot], 0);
    return launch_end(e, KIND_PARTIAL, bytes, "partial_generic_kernel");
}

}  // extern "C"
