/*
 * lkcuda_sitemodel.cu — SURVEY 8(f).3: partial likelihoods under site-specific frequency models (-fs / PMSF).
 * Replaces PhyloTree::computeSitemodelPartialLikelihoodEigen (phylokernelsitemodel.cpp:20-436) and its SIMD twin
 * (phylokernelsitemodel.h): every pattern has its own eigensystem (ModelSet, models->at(ptn)), so the per-branch matrices
 * evec * exp(eval * rate * len) of the other families do not exist - a node costs, per pattern, 2*K*S exponentials, two S x S
 * products with evec and one with inv_evec, and the reference re-reads the pattern's 2*S*S + S eigen doubles at every node.
 *
 * B200 design: ONE WARP OWNS A PATTERN FOR THE WHOLE SCHEDULE, lane = state.  Lane x keeps row x of evec and row x of inv_evec of its
 * pattern's model in REGISTERS (2*S doubles) for all nodes of the batch, so the eigensystems are read from HBM once per pattern and
 * launch instead of once per pattern and node (6.5 KB per pattern for protein - more than a node's partial_lh block).  The vectors
 * exp(eval*rate*len) * L_child and the products' intermediate go through a 3 x 32-double shared-memory scratch per warp and are read
 * back as broadcasts; a child computed earlier in the same launch is re-read by the very lane that wrote it, so no fence is needed.
 * Tip vectors (tip_partial_lh[taxon][ptn][S] in the reference, phylotreesse.cpp:373-456) are formed on the fly from the inv_evec row.
 * Bound: FP64 pipe (the exponentials and 6*S*S + S flops per pattern, category and node); HBM sees each node's vector written once.
 */
#include "lkcuda_traverse.cuh"

using namespace lkc;

#define SM_WARPS 8
#define SM_THREADS (SM_WARPS * 32)

/* S > 0: rows in registers (static indexing); S == 0: any state count <= 32, rows read through the cache */
template <int S>
__global__ void __launch_bounds__(SM_THREADS) sitemodel_traverse_kernel(DevView v, const DevOp *__restrict__ ops, const DevChild *__restrict__ children, int nops) {
    __shared__ __align__(16) double scratch[SM_WARPS][3][32];
    const int wlane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int ns = S > 0 ? S : v.S;
    /* a warp holds 32 / SP patterns, SP = the state count rounded up to a power of two (DNA: 8 patterns of 4 lanes, protein: 1 of 32) */
    int SP = 1;
    while (SP < ns) SP <<= 1;
    const int gpw = 32 / SP, grp = wlane / SP, lane = wlane - grp * SP;
    double *bl = scratch[warp][0] + grp * SP, *br = scratch[warp][1] + grp * SP, *tm = scratch[warp][2] + grp * SP;
    const long long nwarps = (long long)gridDim.x * SM_WARPS;
    for (long long p0 = ((long long)blockIdx.x * SM_WARPS + warp) * gpw; p0 < v.NP; p0 += nwarps * gpw) {
        const bool live = p0 + grp < v.NP;
        const long long p = live ? p0 + grp : v.NP - 1;        /* idle groups shadow the last pattern and write nothing */
        const bool on = live && lane < ns;
        const long long m = v.ptn_model[p];
        const double *evec_m = v.evec + m * ns * ns, *inv_m = v.inv_evec + m * ns * ns;
        const double ev = on ? v.eval[m * ns + lane] : 0.0;
        double evr[S > 0 ? S : 1], ivr[S > 0 ? S : 1];
        if (S > 0 && lane < ns) {
#pragma unroll
            for (int j = 0; j < S; j++) { evr[j] = evec_m[lane * S + j]; ivr[j] = inv_m[lane * S + j]; }
        }
        const double *evrow = evec_m + (lane < ns ? lane : 0) * ns, *ivrow = inv_m + (lane < ns ? lane : 0) * ns;
        for (int o = 0; o < nops; o++) {
            const DevOp op = ops[o];
            const DevChild c0 = children[op.child0], c1 = children[op.child0 + 1];
            const bool tip0 = LKCUDA_IS_TIP(c0.id), tip1 = LKCUDA_IS_TIP(c1.id);
            double t0 = 0.0, t1 = 0.0;          /* tip vectors, one entry per lane */
            int sc = 0;
#pragma unroll
            for (int side = 0; side < 2; side++) {
                const DevChild &c = side ? c1 : c0;
                if (!LKCUDA_IS_TIP(c.id)) { sc += vec_sc(v, c.id)[p]; continue; }
                const int state = tip_states(v, c.id)[p];
                double t = 0.0;
                if (on) {
                    if (state < ns) t = ivrow[state];
                    else {
                        const unsigned long long mask = v.state_mask[state];
                        if (S > 0) {
#pragma unroll
                            for (int x = 0; x < S; x++) if ((mask >> x) & 1ull) t += ivr[x];
                        } else for (int x = 0; x < ns; x++) if ((mask >> x) & 1ull) t += ivrow[x];
                    }
                }
                if (side) t1 = t; else t0 = t;
            }
            double *dst = vec_lh_w(v, op.dst) + p * v.B;
            const double *l0 = tip0 ? NULL : vec_lh(v, c0.id) + p * v.B, *l1 = tip1 ? NULL : vec_lh(v, c1.id) + p * v.B;
            double lh_max = 0.0;
            for (int c = 0; c < v.K; c++) {
                const double rate = v.blk_rate[c];
                if (on) {
                    const double a = tip0 ? t0 : l0[c * ns + lane], b = tip1 ? t1 : l1[c * ns + lane];
                    bl[lane] = exp(ev * (rate * child_len(v, c0))) * a;
                    br[lane] = exp(ev * (rate * child_len(v, c1))) * b;
                }
                __syncwarp();
                double vl = 0.0, vr = 0.0;
                if (S > 0) {
#pragma unroll
                    for (int i = 0; i < S; i++) { vl += evr[i] * bl[i]; vr += evr[i] * br[i]; }
                } else if (on) for (int i = 0; i < ns; i++) { vl += evrow[i] * bl[i]; vr += evrow[i] * br[i]; }
                if (on) tm[lane] = vl * vr;
                __syncwarp();
                double res = 0.0;
                if (S > 0) {
#pragma unroll
                    for (int x = 0; x < S; x++) res += tm[x] * ivr[x];
                } else if (on) for (int x = 0; x < ns; x++) res += tm[x] * ivrow[x];
                if (on) { dst[c * ns + lane] = res; lh_max = fmax(lh_max, fabs(res)); }
                __syncwarp();
            }
            for (int off = SP >> 1; off > 0; off >>= 1) lh_max = fmax(lh_max, __shfl_xor_sync(0xffffffffu, lh_max, off));
            if (!(tip0 && tip1) && lh_max < LK_THR) {       /* phylokernelsitemodel.cpp:306-335, 404-430; the cherry case never scales */
                if (lh_max == 0.0) {                        /* refill with taxon 0's tip vector of this pattern (&tip_partial_lh[ptn*nstates]) */
                    const int state = v.states[p];
                    double t = 0.0;
                    if (on) {
                        if (state < ns) t = ivrow[state];
                        else { const unsigned long long mask = v.state_mask[state]; for (int x = 0; x < ns; x++) if ((mask >> x) & 1ull) t += ivrow[x]; }
                        for (int c = 0; c < v.K; c++) dst[c * ns + lane] = t;
                    }
                    sc += 4;
                } else {
                    if (on) for (int c = 0; c < v.K; c++) dst[c * ns + lane] *= LK_THR_INV;
                    sc += 1;
                }
            }
            if (lane == 0 && live) vec_sc_w(v, op.dst)[p] = (int16_t)sc;
        }
    }
}

int lkc::launch_traverse_sitemodel(lkcuda_engine *e, int slot, int nops) {
    const DevView &v = e->v;
    if (!v.ptn_model) return fail(e, LKCUDA_ERR_STATE, "site-specific model: lkcuda_set_site_models first");
    LK_CUDA(e, cudaMemcpyAsync(e->d_ops[slot], e->h_ops[slot], sizeof(DevOp) * nops, cudaMemcpyHostToDevice, e->stream));
    LK_CUDA(e, cudaMemcpyAsync(e->d_child[slot], e->h_child[slot], sizeof(DevChild) * 2 * nops, cudaMemcpyHostToDevice, e->stream));
    LK_CUDA(e, cudaEventRecord(e->ring_ev[slot], e->stream));
    e->ring_used[slot] = true;
    if (v.NP == 0) return LKCUDA_OK;
    int rc = states_barrier(e);
    if (rc) return rc;
    int sp = 1;
    while (sp < v.S) sp <<= 1;
    long long need = (v.NP + SM_WARPS * (32 / sp) - 1) / (SM_WARPS * (32 / sp));
    long long cap = (long long)e->num_sms * 2;
    int grid = (int)(need < cap ? need : cap);
    uint64_t bytes = 0;
    for (int i = 0; i < nops; i++) {
        const int ntip = (LKCUDA_IS_TIP(e->h_child[slot][2 * i].id) ? 1 : 0) + (LKCUDA_IS_TIP(e->h_child[slot][2 * i + 1].id) ? 1 : 0);
        bytes += op_bytes(v, ntip, 2 - ntip);
        e->kind_flops[KIND_PARTIAL] += op_flops(v, ntip, 2 - ntip);
    }
    bytes += (uint64_t)v.NP * (uint64_t)(2 * v.S * v.S + v.S) * 8;      /* the pattern's eigensystem, once per launch */
    launch_begin(e, KIND_PARTIAL);
    if (v.S == 4) sitemodel_traverse_kernel<4><<<grid, SM_THREADS, 0, e->stream>>>(e->v, e->d_ops[slot], e->d_child[slot], nops);
    else if (v.S == 20) sitemodel_traverse_kernel<20><<<grid, SM_THREADS, 0, e->stream>>>(e->v, e->d_ops[slot], e->d_child[slot], nops);
    else sitemodel_traverse_kernel<0><<<grid, SM_THREADS, 0, e->stream>>>(e->v, e->d_ops[slot], e->d_child[slot], nops);
    return launch_end(e, KIND_PARTIAL, bytes, "sitemodel_traverse_kernel");
}

extern "C" int lkcuda_set_site_models(lkcuda_engine *e, const int32_t *ptn_model, const uint64_t *state_mask) {
    LK_ENTER(e);
    if (!ptn_model || !state_mask) return fail(e, LKCUDA_ERR_ARG, "null argument");
    DevView &v = e->v;
    if (!(v.flags & LKCUDA_FLAG_SITEMODEL)) return fail(e, LKCUDA_ERR_STATE, "lkcuda_set_site_models: engine created without LKCUDA_FLAG_SITEMODEL");
    for (long long p = 0; p < v.NP; p++)
        if (ptn_model[p] < 0 || ptn_model[p] >= v.E) return fail(e, LKCUDA_ERR_ARG, "ptn_model[%lld] out of range", p);
    if (!e->d_ptn_model) {
        LK_CUDA(e, cudaMalloc((void **)&e->d_ptn_model, ((size_t)v.NP + 4) * 4));
        LK_CUDA(e, cudaMalloc((void **)&e->d_state_mask, ((size_t)v.SU + 1) * 8));
        e->total_dev_bytes += (size_t)v.NP * 4;
    }
    if (v.NP) LK_CUDA(e, cudaMemcpyAsync(e->d_ptn_model, ptn_model, (size_t)v.NP * 4, cudaMemcpyHostToDevice, e->stream));
    LK_CUDA(e, cudaMemcpyAsync(e->d_state_mask, state_mask, ((size_t)v.SU + 1) * 8, cudaMemcpyHostToDevice, e->stream));
    LK_CUDA(e, cudaStreamSynchronize(e->stream));
    v.ptn_model = e->d_ptn_model; v.state_mask = e->d_state_mask;
    e->theta_ready = false;
    return LKCUDA_OK;
}
