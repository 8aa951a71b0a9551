/*
 * lkcuda_reduce.cu — the streaming reductions of the likelihood engine (sm_100a, FP64):
 *   branch log-likelihood   computeLikelihoodBranchEigenSIMD   phylokernel.h:734-1020 (+ mixture :728, mixrate :700,
 *                                                               generic phylotreesse.cpp:1158/1892/2549)
 *   theta_all build         computeLikelihoodDervEigenSIMD      phylokernel.h:535-579
 *   df / ddf                computeLikelihoodDervEigenSIMD      phylokernel.h:521-532, 599-725
 *   lnL from theta_all      computeLikelihoodFromBufferEigenSIMD phylokernel.h:1023-1192
 *   Newton-Raphson          Optimization::minimizeNewton        optimization.cpp:388-465
 *
 * All of them stream `B = K*S` doubles per pattern once and are HBM-bound.  Mapping: G lanes share one
 * pattern (G = largest power of two dividing B/4, <= 32); lane g reads the 32-byte sectors g, g+G, ... of
 * the pattern's block, so every load instruction covers G*32 contiguous bytes per pattern and only whole
 * sectors move.  Per-pattern sums are finished with shuffles inside the G-lane group; the cross-pattern
 * sums use a fixed-order block reduction and a last-block-done pass (deterministic run to run).
 * Shapes with B % 4 != 0 take the scalar path (G = 1).
 */
#include "lkcuda_common.cuh"
#include <math.h>

using namespace lkc;

#define RED_THREADS 256
#define RED_BLOCKS_PER_SM 3     /* __launch_bounds__(256, 4): <= 64 registers, 1024 threads per SM with 4 x 32-byte loads each in flight */

struct GroupShape { int G, gshift, vec; };

static GroupShape group_shape(const DevView &v) {
    GroupShape s;
    if (v.B % 4 != 0) { s.G = 1; s.gshift = 0; s.vec = 0; return s; }
    /* lanes per pattern: a power of two dividing the sector count that leaves every lane at least 4 sectors (4 independent
       32-byte loads in flight per lane); DNA+G4 (4 sectors): one lane per pattern, no shuffles at all */
    int nsec = v.B / 4, g = 1, sh = 0;
    while (g < 32 && nsec % (g * 2) == 0 && nsec / (g * 2) >= 4) { g *= 2; sh++; }
    s.G = g; s.gshift = sh; s.vec = 1;
    return s;
}

static int red_grid(const lkcuda_engine *e, long long NP, int G) {
    (void)G;
    long long per_block = RED_THREADS;      /* a warp takes 32 patterns per outer step whatever G is */
    long long need = (NP + per_block - 1) / per_block;
    long long cap = (long long)e->num_sms * RED_BLOCKS_PER_SM;      /* one resident wave, grid-stride inside */
    if (cap > e->red_blocks) cap = e->red_blocks;
    if (need < 1) need = 1;
    return (int)(need < cap ? need : cap);
}

/* val0[j] = exp(eval*rate*len) * weight, optionally val1 = cof*val0, val2 = cof*val1  (phylokernel.h:521-532, 767-775) */
__device__ __forceinline__ void build_val(const DevView &v, double len, double *val0, double *val1, double *val2) {
    for (int j = threadIdx.x; j < v.B; j += blockDim.x) {
        int k = j / v.S, i = j - k * v.S;
        double cof = v.eval[v.blk_eigen[k] * v.S + i] * v.blk_rate[k];
        double x = exp(cof * len) * v.blk_weight[k];
        val0[j] = x;
        if (val1) { double y = cof * x; val1[j] = y; val2[j] = cof * y; }
    }
}

/* tipB[st][j] = tip_partial_lh[st][eigen(k)][i] (* val[j]): the tip end's partial expanded to block layout */
__device__ __forceinline__ void build_tipB(const DevView &v, double *tipB, const double *val) {
    const int n = (v.SU + 1) * v.B;
    for (int idx = threadIdx.x; idx < n; idx += blockDim.x) {
        int st = idx / v.B, j = idx - st * v.B, k = j / v.S, i = j - k * v.S;
        double t = v.tip[((long long)st * v.E + v.blk_eigen[k]) * v.S + i];
        tipB[idx] = val ? val[j] * t : t;
    }
}

__device__ __forceinline__ double tip_elem(const DevView &v, int st, int j, const double *val) {
    int k = j / v.S, i = j - k * v.S;
    double t = v.tip[((long long)st * v.E + v.blk_eigen[k]) * v.S + i];
    return val ? val[j] * t : t;
}

template <int N>
__device__ __forceinline__ void group_sum(double (&a)[N], int G) {
    for (int off = 1; off < G; off <<= 1) {
#pragma unroll
        for (int i = 0; i < N; i++) a[i] += __shfl_xor_sync(0xffffffffu, a[i], off);
    }
}

/* ------------------------------------------------------------------ branch lnL ------------------------------------------------------------------ */

template <bool TIP, bool VEC>
__global__ void __launch_bounds__(RED_THREADS, RED_BLOCKS_PER_SM) branch_kernel(DevView v, int dad, int node, double len, int gshift, int use_tab) {
    extern __shared__ __align__(32) double sh[];
    __shared__ double sred[4 * 32];
    double *val = sh, *tipB = sh + ((v.B + 3) & ~3);
    build_val(v, len, val, NULL, NULL);
    __syncthreads();
    if (TIP && use_tab) { build_tipB(v, tipB, val); __syncthreads(); }

    const int G = 1 << gshift, g = threadIdx.x & (G - 1);
    const int gpw = 32 >> gshift;                                  /* groups per warp */
    const long long warp_id = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    const int gin = (threadIdx.x & 31) >> gshift;
    const double *lh_node = vec_lh(v, node);
    const double *lh_dad = TIP ? NULL : vec_lh(v, dad);
    const int16_t *sc_node = vec_sc(v, node), *sc_dad = TIP ? NULL : vec_sc(v, dad);
    const uint8_t *st = TIP ? tip_states(v, dad) : NULL;
    const int nsec = v.B >> 2;
    double acc[4] = {0.0, 0.0, 0.0, 0.0};

    /* one outer step = 32 patterns per warp: G sub-steps of 32/G patterns, each summed inside its G-lane group;
       lane (gin, g) keeps the total of sub-step g, so afterwards every lane owns one distinct pattern for the tail */
    for (long long base = warp_id * 32; base < v.NP; base += nwarps * 32) {
        double mine = 0.0;
        for (int j = 0; j < G; j++) {
            const long long p = base + (long long)j * gpw + gin;
            double s0 = 0.0;
            if (p < v.NP) {
                const int state = TIP ? st[p] : 0;
                if (VEC) {
                    const double *a = lh_node + p * v.B;
                    constexpr int NB = TIP ? 4 : 2;     /* sectors per vector in flight before the first is used */
                    for (int q0 = g; q0 < nsec; q0 += NB * G) {
                        d4 xs[NB], ys[NB];
#pragma unroll
                        for (int u = 0; u < NB; u++) if (q0 + u * G < nsec) {
                            xs[u] = ld_stream256(a + 4 * (q0 + u * G));
                            if (!TIP) ys[u] = ld_stream256(lh_dad + p * v.B + 4 * (q0 + u * G));
                        }
#pragma unroll
                        for (int u = 0; u < NB; u++) {
                            const int q = q0 + u * G;
                            if (q >= nsec) break;
                            const d4 x = xs[u];
                            if (TIP) {
                                double t0, t1, t2, t3;
                                if (use_tab) { const double *t = tipB + state * v.B + 4 * q; t0 = t[0]; t1 = t[1]; t2 = t[2]; t3 = t[3]; }
                                else { t0 = tip_elem(v, state, 4 * q, val); t1 = tip_elem(v, state, 4 * q + 1, val); t2 = tip_elem(v, state, 4 * q + 2, val); t3 = tip_elem(v, state, 4 * q + 3, val); }
                                s0 = fma(t0, x.x, s0); s0 = fma(t1, x.y, s0); s0 = fma(t2, x.z, s0); s0 = fma(t3, x.w, s0);
                            } else {
                                const d4 y = ys[u];
                                const double *w = val + 4 * q;
                                s0 = fma(w[0] * x.x, y.x, s0); s0 = fma(w[1] * x.y, y.y, s0);
                                s0 = fma(w[2] * x.z, y.z, s0); s0 = fma(w[3] * x.w, y.w, s0);
                            }
                        }
                    }
                } else {
                    const double *a = lh_node + p * v.B;
                    for (int jj = 0; jj < v.B; jj++) {
                        if (TIP) s0 = fma(use_tab ? tipB[state * v.B + jj] : tip_elem(v, state, jj, val), a[jj], s0);
                        else s0 = fma(val[jj] * a[jj], lh_dad[p * v.B + jj], s0);
                    }
                }
            }
            double s[1] = {s0};
            group_sum<1>(s, G);
            if (j == g) mine = s[0];
        }
        const long long p = base + (long long)g * gpw + gin;
        if (p < v.NP) {
            int sc = sc_node[p] + (TIP ? 0 : sc_dad[p]);
            lnl_tail(v, p, mine, sc, acc);
        }
    }
    if (grid_reduce<4>(acc, sred, v.red, v.counter)) lnl_finish(v, acc, acc[1]);
}

/* ---- DNA + 4 categories (B = 16, 128 bytes per pattern): the mapping of the dedicated streaming kernels ----
   FOUR lanes share a pattern; lane c of the group owns category c = one 32-byte sector, so one warp-wide LDG.E.256 covers 1 KB of
   contiguous memory = 8 whole 128-byte lines (one lane per pattern would touch 32 different lines per instruction).  A warp works on
   tiles of 32 patterns = 4 KB: load u (0..3) of lane l moves bytes [1024u + 32l, +32) of the tile, i.e. category l&3 of pattern
   8u + (l>>2).  The sums over the four categories are finished by a transposed butterfly (b16_gather: 3 shuffles finish 4 patterns);
   afterwards lane l owns pattern 8(l&3) + (l>>2) of the tile for the tail (division / log, frequency weighting, pattern_lh store): all
   32 lanes stay busy and the per-pattern arrays are still read and written as contiguous 256-byte pieces.  The NEXT tile's sectors are
   requested before the current tile is reduced.  Tiles go to warps round-robin with the warp index running over the BLOCKS first, so
   the warps that get one tile more than the others are spread over all SMs. */
struct B16Lane { int lane, c, pg; long long t0, nw, ntile; };

__device__ __forceinline__ B16Lane b16_lane(const DevView &v) {
    B16Lane m;
    m.lane = threadIdx.x & 31; m.c = m.lane & 3; m.pg = m.lane >> 2;
    m.nw = (long long)gridDim.x * (blockDim.x >> 5);
    m.t0 = (long long)(threadIdx.x >> 5) * gridDim.x + blockIdx.x;
    m.ntile = (v.NP + 31) >> 5;
    return m;
}

__device__ __forceinline__ d4 ld_coherent256(const double *p) {      /* data written earlier by this very launch */
    d4 r;
    asm volatile("ld.global.L1::no_allocate.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(r.x), "=d"(r.y), "=d"(r.z), "=d"(r.w) : "l"(p) : "memory");
    return r;
}

/* the four sectors of this lane in tile t of a B = 16 vector (zeros beyond the last pattern) */
template <bool COHERENT>
__device__ __forceinline__ void b16_load(const double *vec, long long t, const B16Lane &m, long long NP, d4 (&x)[4]) {
    const double *src = vec + t * 512 + m.lane * 4;
    const long long p0 = t * 32 + m.pg;
#pragma unroll
    for (int u = 0; u < 4; u++) {
        if (p0 + 8 * u < NP) x[u] = COHERENT ? ld_coherent256(src + 128 * u) : ld_stream256(src + 128 * u);
        else { x[u].x = 0.0; x[u].y = 0.0; x[u].z = 0.0; x[u].w = 0.0; }
    }
}

/* s[u] = this lane's (category c) share of pattern 8u + pg; returns the sum over the group's four lanes of s[c]: the finished value of
   pattern 8c + pg.  Fixed order (s_c0 + s_c1) + (s_c2 + s_c3) in every lane. */
__device__ __forceinline__ double b16_gather(const double (&s)[4], int c) {
    const bool b0 = (c & 1) != 0, b1 = (c & 2) != 0;
    double a0 = b0 ? s[1] : s[0], a1 = b0 ? s[3] : s[2];
    const double g0 = b0 ? s[0] : s[1], g1 = b0 ? s[2] : s[3];
    a0 += __shfl_xor_sync(0xffffffffu, g0, 1);
    a1 += __shfl_xor_sync(0xffffffffu, g1, 1);
    const double r = b1 ? a1 : a0, g = b1 ? a0 : a1;
    return r + __shfl_xor_sync(0xffffffffu, g, 2);
}

__device__ __forceinline__ double dot4(const d4 &x, const double (&w)[4]) {
    double s = x.x * w[0];
    s = fma(x.y, w[1], s); s = fma(x.z, w[2], s); s = fma(x.w, w[3], s);
    return s;
}

/* lane c's table entries of category c straight from the kernel parameters (build_val's operations, no memory traffic, no barrier) */
template <bool DERIV>
__device__ __forceinline__ void b16_lane_tables(const B16Tab &tab, int c, double len, double (&w0)[4], double (&w1)[4], double (&w2)[4]) {
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const double cof = tab.cof[4 * c + i];
        const double x = exp(cof * len) * tab.w[c];
        w0[i] = x;
        if (DERIV) { const double y = cof * x; w1[i] = y; w2[i] = cof * y; }
    }
}

static int b16_grid(const lkcuda_engine *e, long long NP) {
    long long need = (((NP + 31) >> 5) + (RED_THREADS / 32) - 1) / (RED_THREADS / 32);
    long long cap = (long long)e->num_sms * 2;
    if (cap > e->red_blocks) cap = e->red_blocks;
    if (need < 1) need = 1;
    return (int)(need < cap ? need : cap);
}

/* tipB rows are 18 doubles apart: lanes with different states read different banks */
#define B16_TIPROW 18

/* DNA + 4 categories, tip end (the root branch of every full-tree lnL) */
__global__ void __launch_bounds__(RED_THREADS, 2) branch_tip_b16_kernel(DevView v, B16Tab tab, int dad, int node, double len) {
    __shared__ __align__(32) double tipB[32 * B16_TIPROW];
    __shared__ double sred[4 * 32];
    const B16Lane m = b16_lane(v);
    const double *lh_node = vec_lh(v, node);
    const int16_t *sc_node = vec_sc(v, node);
    const uint8_t *st = tip_states(v, dad);
    long long t = m.t0;
    d4 n[4]; double ninv = 0.0, nf = 0.0; int nst = 0, nsc = 0;
#define B16_BRANCH_FETCH() do { \
        b16_load<false>(lh_node, t, m, v.NP, n); \
        const long long pl = t * 32 + m.lane, pt = t * 32 + 8 * m.c + m.pg; \
        nst = pl < v.NP ? (int)st[pl] : 0; \
        if (pt < v.NP) { ninv = v.invar[pt]; nf = pt < v.P ? v.freq[pt] : 0.0; nsc = sc_node[pt]; } \
    } while (0)
    if (t < m.ntile) B16_BRANCH_FETCH();
    for (int idx = threadIdx.x; idx < (v.SU + 1) * 16; idx += blockDim.x) {
        const int s = idx >> 4, j = idx & 15;
        const double tp = v.tip[((long long)s * v.E + tab.eig[j >> 2]) * 4 + (j & 3)];
        tipB[s * B16_TIPROW + j] = exp(tab.cof[j] * len) * tab.w[j >> 2] * tp;
    }
    __syncthreads();
    double acc[4] = {0.0, 0.0, 0.0, 0.0};
    while (t < m.ntile) {
        const d4 x[4] = {n[0], n[1], n[2], n[3]};
        const double inv = ninv, f = nf;
        const int stl = nst, sc = nsc;
        const long long pt = t * 32 + 8 * m.c + m.pg;
        t += m.nw;
        if (t < m.ntile) B16_BRANCH_FETCH();
        double s[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int state = __shfl_sync(0xffffffffu, stl, 8 * u + m.pg);
            const double2 *tb = reinterpret_cast<const double2 *>(tipB + state * B16_TIPROW + 4 * m.c);
            const double2 a = tb[0], b = tb[1];
            double q = a.x * x[u].x;
            q = fma(a.y, x[u].y, q); q = fma(b.x, x[u].z, q); q = fma(b.y, x[u].w, q);
            s[u] = q;
        }
        const double lh = b16_gather(s, m.c);
        if (pt < v.NP) lnl_tail_vals(v, pt, lh, sc, inv, f, acc);
    }
#undef B16_BRANCH_FETCH
    if (grid_reduce<4>(acc, sred, v.red, v.counter)) lnl_finish(v, acc, acc[1]);
}

/* _pattern_lh[p] -= log(1 - prob_const) for the observed patterns (+ASC, phylokernel.h:1009-1016) */
__global__ void asc_adjust_kernel(DevView v) {
    const double lc = v.result[6];
    for (long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x; p < v.P; p += (long long)gridDim.x * blockDim.x)
        v.pattern_lh[p] -= lc;
}

/* _pattern_lh_cat[p][k] = sum_i val[k][i] * L_node * L_dad (weighted by the block weight, without ptn_invar and
   scaling: phylotreesse.cpp:1194-1237, 2588-2635) */
__global__ void __launch_bounds__(RED_THREADS) pattern_lh_cat_kernel(DevView v, int dad, int node, double len) {
    extern __shared__ __align__(32) double sh[];
    double *val = sh;
    build_val(v, len, val, NULL, NULL);
    __syncthreads();
    const bool tip = LKCUDA_IS_TIP(dad);
    const double *lh_node = vec_lh(v, node);
    const long long total = v.P * v.K;
    for (long long u = (long long)blockIdx.x * blockDim.x + threadIdx.x; u < total; u += (long long)gridDim.x * blockDim.x) {
        long long p = u / v.K; int k = (int)(u - p * v.K);
        const double *a = lh_node + p * v.B + k * v.S;
        double s = 0.0;
        if (tip) {
            int state = tip_states(v, dad)[p];
            const double *t = v.tip + ((long long)state * v.E + v.blk_eigen[k]) * v.S;
            for (int i = 0; i < v.S; i++) s += (val[k * v.S + i] * t[i]) * a[i];
        } else {
            const double *b = vec_lh(v, dad) + p * v.B + k * v.S;
            for (int i = 0; i < v.S; i++) s += val[k * v.S + i] * a[i] * b[i];
        }
        v.pattern_lh_cat[u] = s;
    }
}

/* ------------------------------------------------------------------ theta_all ------------------------------------------------------------------ */

template <bool TIP, bool VEC>
__global__ void __launch_bounds__(RED_THREADS) theta_kernel(DevView v, int dad, int node, int use_tab) {
    extern __shared__ __align__(32) double sh[];
    __shared__ double sred[32];
    double *tipB = sh;
    if (TIP && use_tab) { build_tipB(v, tipB, NULL); __syncthreads(); }
    const double *lh_node = vec_lh(v, node);
    const double *lh_dad = TIP ? NULL : vec_lh(v, dad);
    const uint8_t *st = TIP ? tip_states(v, dad) : NULL;
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x, nthr = (long long)gridDim.x * blockDim.x;
    if (VEC) {
        const int nsec = v.B >> 2;
        const long long total = v.NP * nsec;
        for (long long idx = tid; idx < total; idx += nthr) {
            d4 x = ld_stream256(lh_node + 4 * idx), r;
            if (TIP) {
                long long p = idx / nsec; int q = (int)(idx - p * nsec), state = st[p];
                if (use_tab) { const double *t = tipB + state * v.B + 4 * q; r.x = t[0] * x.x; r.y = t[1] * x.y; r.z = t[2] * x.z; r.w = t[3] * x.w; }
                else { r.x = tip_elem(v, state, 4 * q, NULL) * x.x; r.y = tip_elem(v, state, 4 * q + 1, NULL) * x.y;
                       r.z = tip_elem(v, state, 4 * q + 2, NULL) * x.z; r.w = tip_elem(v, state, 4 * q + 3, NULL) * x.w; }
            } else {
                d4 y = ld_stream256(lh_dad + 4 * idx);
                r.x = x.x * y.x; r.y = x.y * y.y; r.z = x.z * y.z; r.w = x.w * y.w;
            }
            st_stream256(v.theta + 4 * idx, r);
        }
    } else {
        const long long total = v.NP * v.B;
        for (long long idx = tid; idx < total; idx += nthr) {
            if (TIP) { long long p = idx / v.B; int j = (int)(idx - p * v.B); v.theta[idx] = tip_elem(v, st[p], j, NULL) * lh_node[idx]; }
            else v.theta[idx] = lh_node[idx] * lh_dad[idx];
        }
    }
    /* scale bookkeeping of the branch: lh_scale_factor(node)+lh_scale_factor(dad) and the unobserved patterns' scale_num */
    const int16_t *sc_node = vec_sc(v, node), *sc_dad = TIP ? NULL : vec_sc(v, dad);
    double acc[1] = {0.0};
    for (long long p = tid; p < v.NP; p += nthr) {
        int sc = sc_node[p] + (TIP ? 0 : sc_dad[p]);
        if (p < v.P) { if (sc) acc[0] += LK_LOG_THR * v.freq[p] * (double)sc; }
        else v.theta_sc[p - v.P] = (int16_t)sc;
    }
    if (grid_reduce<1>(acc, sred, v.red, v.counter)) v.theta_meta[0] = acc[0];
}

/* ------------------------------------------------------------------ df / ddf and lnL from theta_all ------------------------------------------------------------------ */

/* per-pattern sums over theta: NV = 3 (lh, df, ddf) or 1 (lh) */
template <int NV, bool VEC>
__device__ __forceinline__ void theta_dot(const DevView &v, long long p, int g, int G, const double *val0, const double *val1, const double *val2, double (&s)[NV]) {
    const double *th = v.theta + p * v.B;
    if (VEC) {
        const int nsec = v.B >> 2;
        for (int q0 = g; q0 < nsec; q0 += 4 * G) {
            d4 xs[4];       /* up to 4 sectors in flight before the first is used */
#pragma unroll
            for (int u = 0; u < 4; u++) if (q0 + u * G < nsec) xs[u] = ld_stream256(th + 4 * (q0 + u * G));
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const int q = q0 + u * G;
                if (q >= nsec) break;
                const d4 x = xs[u];
                const double *w0 = val0 + 4 * q;
                s[0] = fma(x.x, w0[0], s[0]); s[0] = fma(x.y, w0[1], s[0]); s[0] = fma(x.z, w0[2], s[0]); s[0] = fma(x.w, w0[3], s[0]);
                if (NV == 3) {
                    const double *w1 = val1 + 4 * q, *w2 = val2 + 4 * q;
                    s[1] = fma(x.x, w1[0], s[1]); s[1] = fma(x.y, w1[1], s[1]); s[1] = fma(x.z, w1[2], s[1]); s[1] = fma(x.w, w1[3], s[1]);
                    s[2] = fma(x.x, w2[0], s[2]); s[2] = fma(x.y, w2[1], s[2]); s[2] = fma(x.z, w2[2], s[2]); s[2] = fma(x.w, w2[3], s[2]);
                }
            }
        }
    } else {
        for (int j = 0; j < v.B; j++) {
            double x = th[j];
            s[0] = fma(x, val0[j], s[0]);
            if (NV == 3) { s[1] = fma(x, val1[j], s[1]); s[2] = fma(x, val2[j], s[2]); }
        }
    }
}

struct DervOut { double df, ddf; };

/* one pass over theta_all at branch length len: returns in the LAST block's thread 0 (true) the finished df, ddf */
template <bool VEC>
__device__ __forceinline__ bool derv_pass(const DevView &v, double len, int gshift, double *val0, double *val1, double *val2, double *sred, DervOut *out) {
    build_val(v, len, val0, val1, val2);
    __syncthreads();
    const int G = 1 << gshift, g = threadIdx.x & (G - 1);
    const int gpw = 32 >> gshift;
    const long long warp_id = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    const int gin = (threadIdx.x & 31) >> gshift;
    double acc[5] = {0.0, 0.0, 0.0, 0.0, 0.0};
    /* 32 patterns per warp and outer step; after the G sub-steps every lane owns one distinct pattern (see branch_kernel) */
    for (long long base = warp_id * 32; base < v.NP; base += nwarps * 32) {
        double m0 = 0.0, m1 = 0.0, m2 = 0.0;
        for (int j = 0; j < G; j++) {
            const long long p = base + (long long)j * gpw + gin;
            double s[3] = {0.0, 0.0, 0.0};
            if (p < v.NP) theta_dot<3, VEC>(v, p, g, G, val0, val1, val2, s);
            group_sum<3>(s, G);
            if (j == g) { m0 = s[0]; m1 = s[1]; m2 = s[2]; }
        }
        const long long p = base + (long long)g * gpw + gin;
        if (p < v.NP) {
            double lh = fabs(m0 + v.invar[p]);
            if (p < v.P) {      /* phylotreesse.cpp:1106-1132 */
                double df_frac = m1 / lh, ddf_frac = m2 / lh, f = v.freq[p];
                double t1 = df_frac * f, t2 = ddf_frac * f;
                acc[0] += t1;
                acc[1] += t2 - t1 * df_frac;
            } else { acc[2] += lh; acc[3] += m1; acc[4] += m2; }
        }
    }
    if (!grid_reduce<5>(acc, sred, v.red, v.counter)) return false;
    double df = acc[0], ddf = acc[1];
    if (isnan(df) || isinf(df)) { df = 0.0; ddf = 0.0; }            /* phylokernel.h:647-651 */
    if (v.U > 0) {                                                  /* phylokernel.h:719-724 */
        double pc = 1.0 - acc[2], df_frac = acc[3] / pc, ddf_frac = acc[4] / pc;
        df += v.nsite * df_frac;
        ddf += v.nsite * (ddf_frac + df_frac * df_frac);
    }
    out->df = df; out->ddf = ddf;
    return true;
}

/* ---- DNA + 4 categories (B = 16, 128 bytes per pattern): dedicated streaming kernels ---- */

/* finish df/ddf from the five grid sums (phylokernel.h:647-651, 719-724) */
__device__ __forceinline__ void derv_finish(const DevView &v, const double (&acc)[5]) {
    double df = acc[0], ddf = acc[1];
    if (isnan(df) || isinf(df)) { df = 0.0; ddf = 0.0; }
    if (v.U > 0) {
        double pc = 1.0 - acc[2], df_frac = acc[3] / pc, ddf_frac = acc[4] / pc;
        df += v.nsite * df_frac;
        ddf += v.nsite * (ddf_frac + df_frac * df_frac);
    }
    double r[LKCUDA_NRESULT] = {0.0, df, ddf, 0.0, 0.0, 0.0, 0.0, 0.0};
    publish_result(v, r);
}

/* the streaming pass of derv_b16_kernel / newton_b16_kernel (mapping: see B16Lane): acc = {df, ddf, prob_const, df_const, ddf_const}
   partial sums of this thread.  The first tile's requests go out before the exp() tables are built; lane c copies the 12 table entries
   of its category into registers.  COHERENT: theta_all was written earlier by this very launch (by this very lane). */
template <bool COHERENT, bool WREG>
__device__ __forceinline__ void derv_b16_stream(const DevView &v, const B16Tab &tab, double len, double *val0, double *val1, double *val2, double (&acc)[5]) {
    const B16Lane m = b16_lane(v);
    long long t = m.t0;
    d4 n[4]; double ninv = 0.0, nf = 0.0;
#define B16_DERV_FETCH() do { \
        b16_load<COHERENT>(v.theta, t, m, v.NP, n); \
        const long long pt = t * 32 + 8 * m.c + m.pg; \
        if (pt < v.NP) { ninv = v.invar[pt]; nf = pt < v.P ? v.freq[pt] : 0.0; } \
    } while (0)
    if (t < m.ntile) B16_DERV_FETCH();
    double w0[4], w1[4], w2[4];
    b16_lane_tables<true>(tab, m.c, len, w0, w1, w2);
    if (!WREG) {        /* every lane of category c stores the same values; a lane only ever re-reads after its own store */
#pragma unroll
        for (int i = 0; i < 4; i++) {
            *reinterpret_cast<volatile double *>(val0 + 4 * m.c + i) = w0[i]; *reinterpret_cast<volatile double *>(val1 + 4 * m.c + i) = w1[i];
            *reinterpret_cast<volatile double *>(val2 + 4 * m.c + i) = w2[i];
        }
    }
#pragma unroll
    for (int i = 0; i < 5; i++) acc[i] = 0.0;
    while (t < m.ntile) {
        const d4 x[4] = {n[0], n[1], n[2], n[3]};
        const double inv = ninv, f = nf;
        const long long pt = t * 32 + 8 * m.c + m.pg;
        t += m.nw;
        if (t < m.ntile) B16_DERV_FETCH();
        if (!WREG) {        /* the persistent Newton kernel is short of registers: the tables are re-read from shared memory per tile */
#pragma unroll
            for (int i = 0; i < 4; i++) {
                w0[i] = *reinterpret_cast<volatile double *>(val0 + 4 * m.c + i); w1[i] = *reinterpret_cast<volatile double *>(val1 + 4 * m.c + i);
                w2[i] = *reinterpret_cast<volatile double *>(val2 + 4 * m.c + i);
            }
        }
        double s0[4], s1[4], s2[4];
#pragma unroll
        for (int u = 0; u < 4; u++) { s0[u] = dot4(x[u], w0); s1[u] = dot4(x[u], w1); s2[u] = dot4(x[u], w2); }
        const double m0 = b16_gather(s0, m.c), m1 = b16_gather(s1, m.c), m2 = b16_gather(s2, m.c);
        if (pt < v.NP) {
            const double lh = fabs(m0 + inv);
            if (pt < v.P) {
                const double df_frac = m1 / lh, ddf_frac = m2 / lh, t1 = df_frac * f, t2 = ddf_frac * f;
                acc[0] += t1;
                acc[1] += t2 - t1 * df_frac;
            } else { acc[2] += lh; acc[3] += m1; acc[4] += m2; }
        }
    }
#undef B16_DERV_FETCH
}

__global__ void __launch_bounds__(RED_THREADS, 2) derv_b16_kernel(DevView v, B16Tab tab, double len) {
    __shared__ double sred[5 * 32];
    double acc[5];
    derv_b16_stream<false, true>(v, tab, len, NULL, NULL, NULL, acc);
    if (grid_reduce<5>(acc, sred, v.red, v.counter)) derv_finish(v, acc);
}

/* lnL from theta_all (computeLikelihoodFromBufferEigenSIMD, phylokernel.h:1023-1192), B = 16: the same stream with one dot product */
__global__ void __launch_bounds__(RED_THREADS, 2) frombuf_b16_kernel(DevView v, B16Tab tab, double len, const double *len_slot) {
    __shared__ double sred[4 * 32];
    if (len_slot) len = __ldcg(len_slot);
    const B16Lane m = b16_lane(v);
    long long t = m.t0;
    d4 n[4];
    if (t < m.ntile) b16_load<false>(v.theta, t, m, v.NP, n);
    double w0[4], wu1[4], wu2[4];
    b16_lane_tables<false>(tab, m.c, len, w0, wu1, wu2);
    double acc[4] = {0.0, 0.0, 0.0, 0.0};
    while (t < m.ntile) {
        const d4 x[4] = {n[0], n[1], n[2], n[3]};
        const long long pt = t * 32 + 8 * m.c + m.pg;
        t += m.nw;
        if (t < m.ntile) b16_load<false>(v.theta, t, m, v.NP, n);
        double s0[4];
#pragma unroll
        for (int u = 0; u < 4; u++) s0[u] = dot4(x[u], w0);
        const double lh = b16_gather(s0, m.c);
        /* scale factors were folded into theta_meta[0] when theta_all was built: sc = 0 for observed patterns */
        if (pt < v.NP) lnl_tail(v, pt, lh, pt < v.P ? 0 : (int)v.theta_sc[pt - v.P], acc);
    }
    if (grid_reduce<4>(acc, sred, v.red, v.counter)) lnl_finish(v, acc, v.theta_meta[0]);
}

/* ---- Optimization::minimizeNewton (optimization.cpp:388-465) as ONE persistent launch (DNA + 4 categories) ----
   All blocks are co-resident (cooperative launch).  Per Newton step every block streams theta_all at the current length (the same
   pass, grid and summation order as derv_b16_kernel: bit-identical df/ddf), the block that arrives last at the reduction forms df/ddf
   (+ASC tail, NaN guard, cross-GPU exchange when peers are attached - every rank then takes the same decisions), advances the
   reference's bracketing / Newton state machine by one evaluation and publishes the next length under a generation counter the other
   blocks poll.  No launch, no host hand-over and no stream round trip between steps.  The state machine uses explicit _rn intrinsics:
   no FMA contraction, the same IEEE operations as the host code it mirrors. */
struct NewtonCtl {
    double x;                       /* next length to evaluate */
    double xl, xh, rts, dx, d2l, f, df;
    int j, nev, phase, done;        /* done: 1 = finished, 2 = non-finite derivative, 3 = too many iterations */
    double optx, last_x;
    unsigned int gen;
};

/* a finished run leaves the length optimizeOneBranch would put into the branch (optx; the caller applies the rare "diverged to the upper
   bound" rule, phylotree.cpp:2172-2180, after it fetched the slots) in its device length slot, with {optx, d2l, evaluations, status} beside it */
__device__ __forceinline__ void newton_leave_length(const DevView &v, int len_slot, const NewtonCtl &c) {
    if (len_slot < 0) return;
    v.lens[len_slot] = c.optx;
    double *info = v.lens + LKCUDA_NLENSLOT + 4 * len_slot;
    info[0] = c.optx; info[1] = c.d2l; info[2] = (double)c.nev; info[3] = (double)c.done;
}

__device__ __forceinline__ void newton_advance(NewtonCtl &c, double x1, double x2, double xacc, int maxNRStep) {
    /* (f, df) = -(df, ddf) at c.rts have just been evaluated; mirrors optimization.cpp:396-463 */
    const double f = c.f, df = c.df;
    c.nev++; c.last_x = c.rts;
    if (!isfinite(f) || !isfinite(df)) { c.done = 2; return; }
    if (c.phase == 0) {
        c.d2l = df;
        if (df >= 0.0 && fabs(f) < xacc) { c.optx = c.rts; c.done = 1; return; }
        if (f < 0.0) { c.xl = c.rts; c.xh = x2; } else { c.xh = c.rts; c.xl = x1; }
        c.dx = fabs(__dsub_rn(c.xh, c.xl));
        c.phase = 1; c.j = 1;
    } else {
        if (df > 0.0 && fabs(f) < xacc) { c.d2l = df; c.optx = c.rts; c.done = 1; return; }
        if (f < 0.0) c.xl = c.rts; else c.xh = c.rts;
        c.j++;
        if (c.j > maxNRStep) { c.done = 3; return; }
    }
    const double rts_old = c.rts;
    const double ta = __dsub_rn(__dmul_rn(__dsub_rn(c.rts, c.xh), df), f), tb = __dsub_rn(__dmul_rn(__dsub_rn(c.rts, c.xl), df), f);
    if ((df <= 0.0) || (__dmul_rn(ta, tb) >= 0.0)) {
        c.dx = __dmul_rn(0.5, __dsub_rn(c.xh, c.xl)); c.rts = __dadd_rn(c.xl, c.dx); c.d2l = df;
        if (c.xl == c.rts) { c.optx = c.rts; c.done = 1; return; }
    } else {
        c.dx = __ddiv_rn(f, df); const double temp = c.rts; c.rts = __dsub_rn(c.rts, c.dx); c.d2l = df;
        if (temp == c.rts) { c.optx = c.rts; c.done = 1; return; }
    }
    if (fabs(c.dx) < xacc || c.j == maxNRStep) { c.optx = rts_old; c.done = 1; return; }
    c.x = c.rts;
}

template <bool TIP, bool WREG, bool PF>
__device__ __forceinline__ void theta_derv_b16_stream(const DevView &v, const B16Tab &tab, int dad, int node, double len, double *val0, double *val1,
                                                      double *val2, double *tipT, double (&acc)[6]);

/* BUILD 0: theta_all is there; 1 / 2: the first step builds it from the branch's two ends (dad a tip / an internal vector) while it
   forms the first df/ddf (the pass of theta_derv_b16_kernel) - every lane later re-reads exactly the theta rows it wrote itself */
template <int BUILD>
__global__ void __launch_bounds__(RED_THREADS, 2) newton_b16_kernel(DevView v, B16Tab tab, NewtonCtl *ctl, int dad, int node, double x1, double xguess, double x2, double xacc, int maxNRStep,
                  int xguess_slot, int len_slot) {
    __shared__ __align__(32) double val0[16], val1[16], val2[16];
    __shared__ __align__(32) double tipT[BUILD == 1 ? 32 * B16_TIPROW : 4];
    __shared__ double sred[6 * 32];
    __shared__ double sh_x; __shared__ int sh_done;
    double x = xguess_slot >= 0 ? __ldcg(v.lens + xguess_slot) : xguess;       /* a length an earlier launch left on the device */
    if (x < x1) x = x1;
    if (x > x2) x = x2;
    for (unsigned int it = 0;; it++) {
        double acc[5];
        bool elected;
        if (BUILD != 0 && it == 0) {
            double a6[6];
            theta_derv_b16_stream<BUILD == 1, false, false>(v, tab, dad, node, x, val0, val1, val2, tipT, a6);
            elected = grid_reduce<6>(a6, sred, v.red, v.counter);
            if (elected) v.theta_meta[0] = a6[5];
#pragma unroll
            for (int i = 0; i < 5; i++) acc[i] = a6[i];
        } else {
            derv_b16_stream<BUILD != 0, false>(v, tab, x, val0, val1, val2, acc);
            elected = grid_reduce<5>(acc, sred, v.red, v.counter);
        }
        if (elected) {          /* thread 0 of the block that arrived last */
            double df = acc[0], ddf = acc[1];
            if (isnan(df) || isinf(df)) { df = 0.0; ddf = 0.0; }    /* derv_finish */
            if (v.U > 0) {
                const double pc = 1.0 - acc[2], df_frac = acc[3] / pc, ddf_frac = acc[4] / pc;
                df += v.nsite * df_frac;
                ddf += v.nsite * (ddf_frac + df_frac * df_frac);
            }
            unsigned int err = 0;
            if (v.peer_world > 1) {
                double r[LKCUDA_NRESULT] = {0.0, df, ddf, 0.0, 0.0, 0.0, 0.0, 0.0};
                err = peer_exchange(v, r, v.peer_seq + it);
                df = r[LKCUDA_R_DF]; ddf = r[LKCUDA_R_DDF];
            }
            NewtonCtl c;            /* written by another SM in the previous step: read around L1 */
            c.x = __ldcg(&ctl->x); c.xl = __ldcg(&ctl->xl); c.xh = __ldcg(&ctl->xh); c.rts = __ldcg(&ctl->rts); c.dx = __ldcg(&ctl->dx);
            c.d2l = __ldcg(&ctl->d2l); c.j = __ldcg(&ctl->j); c.nev = __ldcg(&ctl->nev); c.phase = __ldcg(&ctl->phase); c.done = __ldcg(&ctl->done);
            c.optx = __ldcg(&ctl->optx); c.last_x = __ldcg(&ctl->last_x);
            if (it == 0) { c.rts = x; c.phase = 0; c.nev = 0; c.done = 0; c.j = 0; c.d2l = 0.0; c.optx = x; }
            if (isnan(df) || isinf(df)) { df = 0.0; ddf = 0.0; }    /* computeLikelihoodDerv's own guard (phylokernel.h:647-651) */
            c.f = -df; c.df = -ddf;                                 /* computeFuncDerv, phylotree.cpp:2143-2145 */
            newton_advance(c, x1, x2, xacc, maxNRStep);
            if (err) c.done = 4;
            if (c.done) {
                double r[LKCUDA_NRESULT] = {c.optx, c.d2l, (double)c.nev, c.last_x, (double)c.done, 0.0, 0.0, 0.0};
                post_result(v, r, err);
                newton_leave_length(v, len_slot, c);
            }
            ctl->x = c.x; ctl->xl = c.xl; ctl->xh = c.xh; ctl->rts = c.rts; ctl->dx = c.dx; ctl->d2l = c.d2l;
            ctl->j = c.j; ctl->nev = c.nev; ctl->phase = c.phase; ctl->done = c.done; ctl->optx = c.optx; ctl->last_x = c.last_x;
            __threadfence();
            asm volatile("st.release.gpu.global.u32 [%0], %1;" :: "l"(&ctl->gen), "r"(it + 1) : "memory");
        }
        if (threadIdx.x == 0) {
            unsigned int g;
            do { asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(g) : "l"(&ctl->gen) : "memory"); } while (g < it + 1);
            sh_x = *reinterpret_cast<volatile double *>(&ctl->x);
            sh_done = *reinterpret_cast<volatile int *>(&ctl->done);
        }
        __syncthreads();
        if (sh_done) break;
        x = sh_x;
        __syncthreads();
    }
}

/* theta_all build + first df/ddf in one pass (B = 16; mapping: see B16Lane): reads the node vector and the dad vector (or the dad
   tip's state), writes theta_all at the addresses derv_b16_stream reads with the same lane, keeps the three dot products.
   acc[5] carries the scale sum that theta_kernel leaves in theta_meta. */
template <bool TIP, bool WREG, bool PF>
__device__ __forceinline__ void theta_derv_b16_stream(const DevView &v, const B16Tab &tab, int dad, int node, double len, double *val0, double *val1,
                                                      double *val2, double *tipT, double (&acc)[6]) {
    const B16Lane m = b16_lane(v);
    const double *lh_node = vec_lh(v, node), *lh_dad = TIP ? NULL : vec_lh(v, dad);
    const int16_t *sc_node = vec_sc(v, node), *sc_dad = TIP ? NULL : vec_sc(v, dad);
    const uint8_t *st = TIP ? tip_states(v, dad) : NULL;
    long long t = m.t0;
    d4 n[4]; int nst = 0;       /* PF (tip end only): the next tile's node sectors are requested ahead; two vectors: 8 loads per lane go out together */
    if (PF && t < m.ntile) {
        b16_load<false>(lh_node, t, m, v.NP, n);
        const long long pl = t * 32 + m.lane; nst = pl < v.NP ? (int)st[pl] : 0;
    }
    if (TIP) {
        for (int idx = threadIdx.x; idx < (v.SU + 1) * 16; idx += blockDim.x) {
            const int s = idx >> 4, j = idx & 15;
            tipT[s * B16_TIPROW + j] = v.tip[((long long)s * v.E + tab.eig[j >> 2]) * 4 + (j & 3)];
        }
    }
    double w0[4], w1[4], w2[4];
    b16_lane_tables<true>(tab, m.c, len, w0, w1, w2);
    if (!WREG) {
#pragma unroll
        for (int i = 0; i < 4; i++) {
            *reinterpret_cast<volatile double *>(val0 + 4 * m.c + i) = w0[i]; *reinterpret_cast<volatile double *>(val1 + 4 * m.c + i) = w1[i];
            *reinterpret_cast<volatile double *>(val2 + 4 * m.c + i) = w2[i];
        }
    }
    if (TIP) __syncthreads();
#pragma unroll
    for (int i = 0; i < 6; i++) acc[i] = 0.0;
    while (t < m.ntile) {
        const long long tc = t, pt = t * 32 + 8 * m.c + m.pg;
        d4 x[4], y[4];
        if (PF) {
#pragma unroll
            for (int u = 0; u < 4; u++) x[u] = n[u];
        } else {
            b16_load<false>(lh_node, tc, m, v.NP, x);
            if (!TIP) b16_load<false>(lh_dad, tc, m, v.NP, y);
            else { const long long pl = tc * 32 + m.lane; nst = pl < v.NP ? (int)st[pl] : 0; }
        }
        const int stl = nst;
        double inv = 0.0, f = 0.0; int sc = 0;
        if (pt < v.NP) { inv = v.invar[pt]; f = pt < v.P ? v.freq[pt] : 0.0; sc = sc_node[pt] + (TIP ? 0 : sc_dad[pt]); }
        t += m.nw;
        if (PF && t < m.ntile) {
            b16_load<false>(lh_node, t, m, v.NP, n);
            const long long pl = t * 32 + m.lane; nst = pl < v.NP ? (int)st[pl] : 0;
        }
        if (!WREG) {
#pragma unroll
            for (int i = 0; i < 4; i++) {
                w0[i] = *reinterpret_cast<volatile double *>(val0 + 4 * m.c + i); w1[i] = *reinterpret_cast<volatile double *>(val1 + 4 * m.c + i);
                w2[i] = *reinterpret_cast<volatile double *>(val2 + 4 * m.c + i);
            }
        }
        double s0[4], s1[4], s2[4];
        double *dst = v.theta + tc * 512 + m.lane * 4;
#pragma unroll
        for (int u = 0; u < 4; u++) {
            if (TIP) {
                const int state = __shfl_sync(0xffffffffu, stl, 8 * u + m.pg);
                const double2 *tb = reinterpret_cast<const double2 *>(tipT + state * B16_TIPROW + 4 * m.c);
                const double2 a = tb[0], b = tb[1];
                y[u].x = a.x; y[u].y = a.y; y[u].z = b.x; y[u].w = b.y;
            }
            d4 r;
            r.x = x[u].x * y[u].x; r.y = x[u].y * y[u].y; r.z = x[u].z * y[u].z; r.w = x[u].w * y[u].w;
            if (tc * 32 + 8 * u + m.pg < v.NP) st_stream256(dst + 128 * u, r);
            s0[u] = dot4(r, w0); s1[u] = dot4(r, w1); s2[u] = dot4(r, w2);
        }
        const double m0 = b16_gather(s0, m.c), m1 = b16_gather(s1, m.c), m2 = b16_gather(s2, m.c);
        if (pt < v.NP) {
            const double lh = fabs(m0 + inv);
            if (pt < v.P) {
                const double df_frac = m1 / lh, ddf_frac = m2 / lh, t1 = df_frac * f, t2 = ddf_frac * f;
                acc[0] += t1;
                acc[1] += t2 - t1 * df_frac;
                if (sc) acc[5] += LK_LOG_THR * f * (double)sc;
            } else { acc[2] += lh; acc[3] += m1; acc[4] += m2; v.theta_sc[pt - v.P] = (int16_t)sc; }
        }
    }
}

template <bool TIP>
__global__ void __launch_bounds__(RED_THREADS, 2) theta_derv_b16_kernel(DevView v, B16Tab tab, int dad, int node, double len) {
    __shared__ __align__(32) double val0[16], val1[16], val2[16];
    __shared__ __align__(32) double tipT[TIP ? 32 * B16_TIPROW : 4];
    __shared__ double sred[6 * 32];
    double acc[6];
    theta_derv_b16_stream<TIP, !TIP, TIP>(v, tab, dad, node, len, val0, val1, val2, tipT, acc);
    if (grid_reduce<6>(acc, sred, v.red, v.counter)) {
        v.theta_meta[0] = acc[5];
        const double a5[5] = {acc[0], acc[1], acc[2], acc[3], acc[4]};
        derv_finish(v, a5);
    }
}

template <bool VEC>
__global__ void __launch_bounds__(RED_THREADS, RED_BLOCKS_PER_SM) derv_kernel(DevView v, double len, int gshift) {
    extern __shared__ __align__(32) double sh[];
    __shared__ double sred[5 * 32];
    const int Bp = (v.B + 3) & ~3;
    DervOut o;
    if (derv_pass<VEC>(v, len, gshift, sh, sh + Bp, sh + 2 * Bp, sred, &o)) {
        double r[LKCUDA_NRESULT] = {0.0, o.df, o.ddf, 0.0, 0.0, 0.0, 0.0, 0.0};
        publish_result(v, r);
    }
}

/* one evaluation of the persistent Newton loop, run by the thread that finished the grid reduction: the cross-GPU exchange, the
   reference's state machine (newton_advance), the result when the loop ends, and the hand-over of the next length to the other blocks
   (the same statements as in newton_b16_kernel) */
__device__ __noinline__ void newton_elected(const DevView &v, NewtonCtl *ctl, unsigned int it, double x, double df, double ddf,
                                            double x1, double x2, double xacc, int maxNRStep, int len_slot) {
    unsigned int err = 0;
    if (v.peer_world > 1) {
        double r[LKCUDA_NRESULT] = {0.0, df, ddf, 0.0, 0.0, 0.0, 0.0, 0.0};
        err = peer_exchange(v, r, v.peer_seq + it);
        df = r[LKCUDA_R_DF]; ddf = r[LKCUDA_R_DDF];
    }
    NewtonCtl c;
    c.x = __ldcg(&ctl->x); c.xl = __ldcg(&ctl->xl); c.xh = __ldcg(&ctl->xh); c.rts = __ldcg(&ctl->rts); c.dx = __ldcg(&ctl->dx);
    c.d2l = __ldcg(&ctl->d2l); c.j = __ldcg(&ctl->j); c.nev = __ldcg(&ctl->nev); c.phase = __ldcg(&ctl->phase); c.done = __ldcg(&ctl->done);
    c.optx = __ldcg(&ctl->optx); c.last_x = __ldcg(&ctl->last_x);
    if (it == 0) { c.rts = x; c.phase = 0; c.nev = 0; c.done = 0; c.j = 0; c.d2l = 0.0; c.optx = x; }
    if (isnan(df) || isinf(df)) { df = 0.0; ddf = 0.0; }
    c.f = -df; c.df = -ddf;
    newton_advance(c, x1, x2, xacc, maxNRStep);
    if (err) c.done = 4;
    if (c.done) {
        double r[LKCUDA_NRESULT] = {c.optx, c.d2l, (double)c.nev, c.last_x, (double)c.done, 0.0, 0.0, 0.0};
        post_result(v, r, err);
        newton_leave_length(v, len_slot, c);
    }
    ctl->x = c.x; ctl->xl = c.xl; ctl->xh = c.xh; ctl->rts = c.rts; ctl->dx = c.dx; ctl->d2l = c.d2l;
    ctl->j = c.j; ctl->nev = c.nev; ctl->phase = c.phase; ctl->done = c.done; ctl->optx = c.optx; ctl->last_x = c.last_x;
    __threadfence();
    asm volatile("st.release.gpu.global.u32 [%0], %1;" :: "l"(&ctl->gen), "r"(it + 1) : "memory");
}

/* Optimization::minimizeNewton as ONE persistent cooperative launch for ANY block shape (protein, codon, mixtures): the pass, grid and
   summation order of derv_kernel per evaluation (bit-identical df / ddf), theta_all built beforehand by theta_kernel */
template <bool VEC>
__global__ void __launch_bounds__(RED_THREADS, RED_BLOCKS_PER_SM) newton_generic_kernel(DevView v, NewtonCtl *ctl, double x1, double xguess, double x2,
                                                                                       double xacc, int maxNRStep, int gshift, int xguess_slot, int len_slot) {
    extern __shared__ __align__(32) double sh[];
    __shared__ double sred[5 * 32];
    __shared__ double sh_x; __shared__ int sh_done;
    const int Bp = (v.B + 3) & ~3;
    double x = xguess_slot >= 0 ? __ldcg(v.lens + xguess_slot) : xguess;
    if (x < x1) x = x1;
    if (x > x2) x = x2;
    for (unsigned int it = 0;; it++) {
        DervOut o;
        if (derv_pass<VEC>(v, x, gshift, sh, sh + Bp, sh + 2 * Bp, sred, &o)) newton_elected(v, ctl, it, x, o.df, o.ddf, x1, x2, xacc, maxNRStep, len_slot);
        if (threadIdx.x == 0) {
            unsigned int g;
            do { asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(g) : "l"(&ctl->gen) : "memory"); } while (g < it + 1);
            sh_x = *reinterpret_cast<volatile double *>(&ctl->x);
            sh_done = *reinterpret_cast<volatile int *>(&ctl->done);
        }
        __syncthreads();
        if (sh_done) break;
        x = sh_x;
        __syncthreads();
    }
}

template <bool VEC>
__global__ void __launch_bounds__(RED_THREADS, RED_BLOCKS_PER_SM) frombuf_kernel(DevView v, double len, int gshift, const double *len_slot) {
    extern __shared__ __align__(32) double sh[];
    __shared__ double sred[4 * 32];
    if (len_slot) len = __ldcg(len_slot);
    build_val(v, len, sh, NULL, NULL);
    __syncthreads();
    const int G = 1 << gshift, g = threadIdx.x & (G - 1);
    const int gpw = 32 >> gshift;
    const long long warp_id = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    const int gin = (threadIdx.x & 31) >> gshift;
    double acc[4] = {0.0, 0.0, 0.0, 0.0};
    for (long long base = warp_id * 32; base < v.NP; base += nwarps * 32) {
        double mine = 0.0;
        for (int j = 0; j < G; j++) {
            const long long p = base + (long long)j * gpw + gin;
            double s[1] = {0.0};
            if (p < v.NP) theta_dot<1, VEC>(v, p, g, G, sh, NULL, NULL, s);
            group_sum<1>(s, G);
            if (j == g) mine = s[0];
        }
        const long long p = base + (long long)g * gpw + gin;
        /* scale factors were folded into theta_meta[0] when theta_all was built: pass sc = 0 for observed patterns */
        if (p < v.NP) lnl_tail(v, p, mine, p < v.P ? 0 : (int)v.theta_sc[p - v.P], acc);
    }
    if (grid_reduce<4>(acc, sred, v.red, v.counter)) lnl_finish(v, acc, v.theta_meta[0]);
}

/* out[p] = _pattern_lh[p] + max(scale_num[p], 0) * LOG_SCALING_THRESHOLD (phylotree.cpp:1058-1064) */
__global__ void __launch_bounds__(RED_THREADS) pattern_lh_corrected_kernel(DevView v, int vec, double *out) {
    const int16_t *sc = vec_sc(v, vec);
    for (long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x; p < v.P; p += (long long)gridDim.x * blockDim.x) {
        const int s = sc[p];
        out[p] = s > 0 ? v.pattern_lh[p] + (double)s * LK_LOG_THR : v.pattern_lh[p];
    }
}

/* lh_scale_factor of one vector = sum_p LOG_SCALING_THRESHOLD * ptn_freq[p] * scale_num[p] */
__global__ void __launch_bounds__(RED_THREADS) scale_factor_kernel(DevView v, int vec) {
    __shared__ double sred[32];
    const int16_t *sc = vec_sc(v, vec);
    double acc[1] = {0.0};
    for (long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x; p < v.P; p += (long long)gridDim.x * blockDim.x)
        if (sc[p]) acc[0] += LK_LOG_THR * v.freq[p] * (double)sc[p];
    if (grid_reduce<1>(acc, sred, v.red, v.counter)) v.result[7] = acc[0];
}


/* ------------------------------------------------------------------ site-specific frequency models ------------------------------------------------------------------
   SURVEY 8(f).3: computeSitemodelLikelihoodBranchEigen / DervEigen / FromBufferEigen (phylokernelsitemodel.cpp:548-660, 438-546, 663-703).
   The eigenvalues differ per pattern, so there is no shared val[] table: a warp takes a pattern, lane i owns state i and forms
   exp(eval_i * rate_c * len) itself; category sums by warp shuffles.  MODE 0: branch lnL (also fills _pattern_lh_cat, like the
   reference's generic kernel), 1: theta_all, 2: df/ddf from theta_all, 3: lnL from theta_all. */
template <int MODE>
__global__ void __launch_bounds__(RED_THREADS) sitemodel_reduce_kernel(DevView v, int dad, int node, double len) {
    __shared__ double sred[5 * 32];
    const int wlane = threadIdx.x & 31, ns = v.S;
    /* a warp holds 32 / SP patterns, SP = the state count rounded up to a power of two (DNA: 8 patterns of 4 lanes), as in the traversal kernel;
       the butterfly sums stay inside a pattern's lanes and add the same values in the same order as a 32-lane one padded with zeros */
    int SP = 1;
    while (SP < ns) SP <<= 1;
    const int gpw = 32 / SP, grp = wlane / SP, lane = wlane - grp * SP;
    const long long warp0 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    const bool tip = (MODE <= 1) && LKCUDA_IS_TIP(dad);
    const double *lh_node = MODE <= 1 ? vec_lh(v, node) : NULL, *lh_dad = (MODE <= 1 && !tip) ? vec_lh(v, dad) : NULL;
    double acc4[4] = {0.0, 0.0, 0.0, 0.0}, acc5[5] = {0.0, 0.0, 0.0, 0.0, 0.0};
    double theta_sc = 0.0;
    for (long long base = warp0 * gpw; base < v.NP; base += nwarps * gpw) {
        const bool live = base + grp < v.NP;
        const long long p = live ? base + grp : v.NP - 1;       /* idle groups shadow the last pattern, write and add nothing */
        const bool on = live && lane < ns;
        const long long m = v.ptn_model[p];
        const double ev = on ? v.eval[m * ns + lane] : 0.0;
        double t = 0.0;
        if (tip && on) {
            const int state = tip_states(v, dad)[p];
            const double *ivrow = v.inv_evec + m * ns * ns + lane * ns;
            if (state < ns) t = ivrow[state];
            else { const unsigned long long mask = v.state_mask[state]; for (int x = 0; x < ns; x++) if ((mask >> x) & 1ull) t += ivrow[x]; }
        }
        double lh_ptn = 0.0, df_ptn = 0.0, ddf_ptn = 0.0;
        for (int c = 0; c < v.K; c++) {
            const long long j = p * v.B + c * ns + lane;
            double th = 0.0;
            if (on) th = MODE <= 1 ? (tip ? t : lh_dad[j]) * lh_node[j] : v.theta[j];
            if (MODE == 1) { if (on) v.theta[j] = th; continue; }
            const double cof = ev * v.blk_rate[c];
            double val = on ? exp(cof * len) * th : 0.0, val1 = cof * val, val2 = cof * val1;
            for (int off = SP >> 1; off > 0; off >>= 1) {
                val += __shfl_xor_sync(0xffffffffu, val, off);
                if (MODE == 2) { val1 += __shfl_xor_sync(0xffffffffu, val1, off); val2 += __shfl_xor_sync(0xffffffffu, val2, off); }
            }
            const double prop = v.blk_weight[c];
            if (MODE == 0 && lane == 0 && live && p < v.P) v.pattern_lh_cat[p * v.K + c] = val * prop;
            lh_ptn += val * prop; df_ptn += val1 * prop; ddf_ptn += val2 * prop;
        }
        if (lane != 0 || !live) continue;
        if (MODE == 1) {
            const int sc = vec_sc(v, node)[p] + (tip ? 0 : vec_sc(v, dad)[p]);
            if (sc && p < v.P) theta_sc += LK_LOG_THR * v.freq[p] * (double)sc;
        } else if (MODE == 2) {
            const double lh = 1.0 / fabs(lh_ptn + v.invar[p]);
            const double df_frac = df_ptn * lh, ddf_frac = ddf_ptn * lh, f = v.freq[p], t1 = df_frac * f, t2 = ddf_frac * f;
            acc5[0] += t1; acc5[1] += t2 - t1 * df_frac;
        } else {
            const int sc = MODE == 0 ? vec_sc(v, node)[p] + (tip ? 0 : vec_sc(v, dad)[p]) : 0;
            lnl_tail(v, p, lh_ptn, sc, acc4);
        }
    }
    if (MODE == 1) { double a[1] = {theta_sc}; if (grid_reduce<1>(a, sred, v.red, v.counter)) v.theta_meta[0] = a[0]; }
    else if (MODE == 2) { if (grid_reduce<5>(acc5, sred, v.red, v.counter)) derv_finish(v, acc5); }
    else if (grid_reduce<4>(acc4, sred, v.red, v.counter)) lnl_finish(v, acc4, MODE == 0 ? acc4[1] : v.theta_meta[0]);
}

/* With peers attached a reduction kernel waits, inside the kernel, for the other ranks' kernels.  When several engines are driven by ONE
   host thread, nothing the host does between launching engine A's reduction and engine B's may wait for the device - and the FIRST launch
   of a kernel does exactly that under CUDA's lazy module loading (the load synchronises with running kernels: A spins, the host is stuck
   loading the follow-up kernel, B is never launched).  So every kernel of this file is loaded when peers are attached. */
void lkc::preload_reduce_kernels() {
    cudaFuncAttributes fa;
#define PRELOAD(k) cudaFuncGetAttributes(&fa, k)
    PRELOAD((branch_kernel<true, true>)); PRELOAD((branch_kernel<true, false>)); PRELOAD((branch_kernel<false, true>)); PRELOAD((branch_kernel<false, false>));
    PRELOAD(branch_tip_b16_kernel); PRELOAD(asc_adjust_kernel); PRELOAD(pattern_lh_cat_kernel);
    PRELOAD((theta_kernel<true, true>)); PRELOAD((theta_kernel<true, false>)); PRELOAD((theta_kernel<false, true>)); PRELOAD((theta_kernel<false, false>));
    PRELOAD(newton_generic_kernel<true>); PRELOAD(newton_generic_kernel<false>); PRELOAD(newton_b16_kernel<0>); PRELOAD(newton_b16_kernel<1>); PRELOAD(newton_b16_kernel<2>); PRELOAD(derv_b16_kernel); PRELOAD(theta_derv_b16_kernel<true>); PRELOAD(theta_derv_b16_kernel<false>);
    PRELOAD(derv_kernel<true>); PRELOAD(derv_kernel<false>); PRELOAD(frombuf_kernel<true>); PRELOAD(frombuf_kernel<false>);
    PRELOAD(pattern_lh_corrected_kernel); PRELOAD(scale_factor_kernel);
    PRELOAD(sitemodel_reduce_kernel<0>); PRELOAD(sitemodel_reduce_kernel<1>); PRELOAD(sitemodel_reduce_kernel<2>); PRELOAD(sitemodel_reduce_kernel<3>);
#undef PRELOAD
    cudaGetLastError();
}

static int sitemodel_reduce(lkcuda_engine *e, int mode, int dad, int node, double len) {
    const DevView &v = e->v;
    if (!v.ptn_model) return fail(e, LKCUDA_ERR_STATE, "site-specific model: lkcuda_set_site_models first");
    int sp = 1;
    while (sp < v.S) sp <<= 1;
    long long need = (v.NP * sp + RED_THREADS - 1) / RED_THREADS;       /* 32 / sp patterns per warp and step */
    long long cap = (long long)e->num_sms * 4 < e->red_blocks ? (long long)e->num_sms * 4 : e->red_blocks;
    int grid = (int)(need < 1 ? 1 : (need < cap ? need : cap));
    const uint64_t b = (uint64_t)v.B * 8, tipd = LKCUDA_IS_TIP(dad) ? 1 : 0;
    if (mode != 1) next_result(e);
    if (mode == 0) {
        launch_begin(e, KIND_BRANCH);
        sitemodel_reduce_kernel<0><<<grid, RED_THREADS, 0, e->stream>>>(e->v, dad, node, len);
        return launch_end(e, KIND_BRANCH, (uint64_t)v.NP * (tipd ? b + 1 + 24 : 2 * b + 24), "sitemodel_reduce_kernel<branch>");
    }
    if (mode == 1) {
        launch_begin(e, KIND_THETA);
        sitemodel_reduce_kernel<1><<<grid, RED_THREADS, 0, e->stream>>>(e->v, dad, node, len);
        return launch_end(e, KIND_THETA, (uint64_t)v.NP * (tipd ? 2 * b + 1 : 3 * b), "sitemodel_reduce_kernel<theta>");
    }
    if (mode == 2) {
        launch_begin(e, KIND_DERV);
        sitemodel_reduce_kernel<2><<<grid, RED_THREADS, 0, e->stream>>>(e->v, dad, node, len);
        return launch_end(e, KIND_DERV, (uint64_t)v.NP * (b + 16), "sitemodel_reduce_kernel<derv>");
    }
    launch_begin(e, KIND_FROMBUF);
    sitemodel_reduce_kernel<3><<<grid, RED_THREADS, 0, e->stream>>>(e->v, dad, node, len);
    return launch_end(e, KIND_FROMBUF, (uint64_t)v.NP * (b + 24), "sitemodel_reduce_kernel<frombuf>");
}
#define IS_SITEMODEL(e) (((e)->v.flags & LKCUDA_FLAG_SITEMODEL) != 0)

/* ------------------------------------------------------------------ host side ------------------------------------------------------------------ */

#define TAB_SMEM_LIMIT (44 * 1024)

static int check_ready(lkcuda_engine *e, const char *who) {
    if (!e->model_set || !e->states_set || !e->weights_set) return fail(e, LKCUDA_ERR_STATE, "%s: tip states, pattern weights and model must be set first", who);
    return LKCUDA_OK;
}

static int finish_result(lkcuda_engine *e, double *result) {
    if (!result) return LKCUDA_OK;
    return lkcuda_wait_result(e, result);
}

template <bool TIP>
static void launch_branch(lkcuda_engine *e, const GroupShape &gs, int grid, size_t sm, int dad, int node, double len, int use_tab) {
    if (gs.vec) branch_kernel<TIP, true><<<grid, RED_THREADS, sm, e->stream>>>(e->v, dad, node, len, gs.gshift, use_tab);
    else branch_kernel<TIP, false><<<grid, RED_THREADS, sm, e->stream>>>(e->v, dad, node, len, gs.gshift, use_tab);
}

template <bool TIP>
static void launch_theta(lkcuda_engine *e, const GroupShape &gs, int grid, size_t sm, int dad, int node, int use_tab) {
    if (gs.vec) theta_kernel<TIP, true><<<grid, RED_THREADS, sm, e->stream>>>(e->v, dad, node, use_tab);
    else theta_kernel<TIP, false><<<grid, RED_THREADS, sm, e->stream>>>(e->v, dad, node, use_tab);
}

static int enqueue_derv(lkcuda_engine *e, double len) {
    if (IS_SITEMODEL(e)) return sitemodel_reduce(e, 2, 0, 0, len);
    const DevView &v = e->v;
    GroupShape gs = group_shape(v);
    int grid = red_grid(e, v.NP, gs.G);
    size_t sm = (size_t)3 * ((v.B + 3) & ~3) * 8;
    next_result(e);
    launch_begin(e, KIND_DERV);
    if (gs.vec && v.S == 4 && v.K == 4) derv_b16_kernel<<<b16_grid(e, v.NP), RED_THREADS, 0, e->stream>>>(e->v, e->b16tab, len);
    else if (gs.vec) derv_kernel<true><<<grid, RED_THREADS, sm, e->stream>>>(e->v, len, gs.gshift);
    else derv_kernel<false><<<grid, RED_THREADS, sm, e->stream>>>(e->v, len, gs.gshift);
    return launch_end(e, KIND_DERV, (uint64_t)v.NP * ((uint64_t)v.B * 8 + 16), "derv_kernel");
}

/* what follows the lnL reduction of a branch (also when the traversal kernel did the reduction, lkcuda_partial_batch_lnl): the +ASC
   correction of _pattern_lh, _pattern_lh_cat when the engine keeps it, the hand-over of the result block */
int lkc::branch_lnl_followups(lkcuda_engine *e, int32_t dad, int32_t node, double len, double *result) {
    const DevView &v = e->v;
    int rc;
    if (v.U > 0 && v.P > 0) {
        launch_begin(e, KIND_MISC);
        asc_adjust_kernel<<<(int)((v.P + 255) / 256 < 1184 ? (v.P + 255) / 256 : 1184), 256, 0, e->stream>>>(e->v);
        rc = launch_end(e, KIND_MISC, 0, "asc_adjust_kernel");
        if (rc) return rc;
    }
    if ((v.flags & LKCUDA_FLAG_PATTERN_LH_CAT) && v.P > 0) {
        long long total = v.P * v.K;
        int g2 = (int)((total + RED_THREADS - 1) / RED_THREADS < (long long)e->num_sms * 8 ? (total + RED_THREADS - 1) / RED_THREADS : (long long)e->num_sms * 8);
        launch_begin(e, KIND_MISC);
        pattern_lh_cat_kernel<<<g2, RED_THREADS, (size_t)((v.B + 3) & ~3) * 8, e->stream>>>(e->v, dad, node, len);
        rc = launch_end(e, KIND_MISC, 0, "pattern_lh_cat_kernel");
        if (rc) return rc;
    }
    return finish_result(e, result);
}

extern "C" {

int lkcuda_branch_lnl(lkcuda_engine *e, int32_t dad, int32_t node, double len, double *result) {
    LK_ENTER(e);
    if (node < 0 || node >= e->cfg.nvec || !valid_end(e, dad) || dad == node) return fail(e, LKCUDA_ERR_ARG, "lkcuda_branch_lnl: bad vector / taxon id");
    int rc = check_ready(e, "lkcuda_branch_lnl");
    if (rc) return rc;
    rc = states_barrier(e);
    if (rc) return rc;
    if (IS_SITEMODEL(e)) { rc = sitemodel_reduce(e, 0, dad, node, len); return rc ? rc : finish_result(e, result); }
    const DevView &v = e->v;
    GroupShape gs = group_shape(v);
    const bool tip = LKCUDA_IS_TIP(dad);
    int grid = red_grid(e, v.NP, gs.G);
    size_t Bp = (size_t)((v.B + 3) & ~3);
    size_t tabsz = (size_t)(v.SU + 1) * v.B * 8;
    int use_tab = tip && (Bp * 8 + tabsz <= TAB_SMEM_LIMIT);
    size_t sm = Bp * 8 + (use_tab ? tabsz : 0);
    next_result(e);
    launch_begin(e, KIND_BRANCH);
    if (tip && gs.vec && v.S == 4 && v.K == 4 && v.SU < 32) branch_tip_b16_kernel<<<b16_grid(e, v.NP), RED_THREADS, 0, e->stream>>>(e->v, e->b16tab, dad, node, len);
    else if (tip) launch_branch<true>(e, gs, grid, sm, dad, node, len, use_tab);
    else launch_branch<false>(e, gs, grid, sm, dad, node, len, use_tab);
    uint64_t b = (uint64_t)v.B * 8;
    rc = launch_end(e, KIND_BRANCH, (uint64_t)v.NP * (tip ? b + 1 + 24 : 2 * b + 24), "branch_kernel");
    if (rc) return rc;
    return branch_lnl_followups(e, dad, node, len, result);
}

int lkcuda_pattern_lh_cat(lkcuda_engine *e, int32_t dad, int32_t node, double len) {
    LK_ENTER(e);
    if (node < 0 || node >= e->cfg.nvec || !valid_end(e, dad) || dad == node) return fail(e, LKCUDA_ERR_ARG, "lkcuda_pattern_lh_cat: bad vector / taxon id");
    int rc = check_ready(e, "lkcuda_pattern_lh_cat");
    if (rc) return rc;
    rc = states_barrier(e);
    if (rc) return rc;
    if (IS_SITEMODEL(e)) return sitemodel_reduce(e, 0, dad, node, len);     /* the site-model branch kernel is what fills it */
    const DevView &v = e->v;
    if (v.P == 0) return LKCUDA_OK;
    long long total = v.P * v.K;
    int g2 = (int)((total + RED_THREADS - 1) / RED_THREADS < (long long)e->num_sms * 8 ? (total + RED_THREADS - 1) / RED_THREADS : (long long)e->num_sms * 8);
    launch_begin(e, KIND_MISC);
    pattern_lh_cat_kernel<<<g2, RED_THREADS, (size_t)((v.B + 3) & ~3) * 8, e->stream>>>(e->v, dad, node, len);
    return launch_end(e, KIND_MISC, 0, "pattern_lh_cat_kernel");
}

int lkcuda_theta_build(lkcuda_engine *e, int32_t dad, int32_t node) {
    LK_ENTER(e);
    if (node < 0 || node >= e->cfg.nvec || !valid_end(e, dad) || dad == node) return fail(e, LKCUDA_ERR_ARG, "lkcuda_theta_build: bad vector / taxon id");
    int rc = check_ready(e, "lkcuda_theta_build");
    if (rc) return rc;
    rc = states_barrier(e);
    if (rc) return rc;
    if (IS_SITEMODEL(e)) {
        rc = sitemodel_reduce(e, 1, dad, node, 0.0);
        if (rc) return rc;
        e->theta_ready = true; e->theta_dad = dad; e->theta_node = node;
        return LKCUDA_OK;
    }
    const DevView &v = e->v;
    GroupShape gs = group_shape(v);
    const bool tip = LKCUDA_IS_TIP(dad);
    long long work = gs.vec ? v.NP * (v.B / 4) : v.NP * v.B;
    long long need = (work + RED_THREADS - 1) / RED_THREADS;
    long long cap = (long long)e->num_sms * 8 < e->red_blocks ? (long long)e->num_sms * 8 : e->red_blocks;
    int grid = (int)(need < 1 ? 1 : (need < cap ? need : cap));
    size_t tabsz = (size_t)(v.SU + 1) * v.B * 8;
    int use_tab = tip && tabsz <= TAB_SMEM_LIMIT;
    size_t sm = use_tab ? tabsz : 0;
    launch_begin(e, KIND_THETA);
    if (tip) launch_theta<true>(e, gs, grid, sm, dad, node, use_tab);
    else launch_theta<false>(e, gs, grid, sm, dad, node, use_tab);
    uint64_t b = (uint64_t)v.B * 8;
    rc = launch_end(e, KIND_THETA, (uint64_t)v.NP * (tip ? 2 * b + 1 : 3 * b), "theta_kernel");
    if (rc) return rc;
    e->theta_ready = true; e->theta_dad = dad; e->theta_node = node;
    return LKCUDA_OK;
}

int lkcuda_theta_derv(lkcuda_engine *e, int32_t dad, int32_t node, double len, double *result) {
    LK_ENTER(e);
    const DevView &v = e->v;
    GroupShape gs = group_shape(v);
    if (IS_SITEMODEL(e) || !(gs.vec && v.S == 4 && v.K == 4 && v.SU < 32)) {      /* other shapes: the two passes */
        int rc = lkcuda_theta_build(e, dad, node);
        return rc ? rc : lkcuda_derv(e, len, result);
    }
    if (node < 0 || node >= e->cfg.nvec || !valid_end(e, dad) || dad == node) return fail(e, LKCUDA_ERR_ARG, "lkcuda_theta_derv: bad vector / taxon id");
    int rc = check_ready(e, "lkcuda_theta_derv");
    if (rc) return rc;
    rc = states_barrier(e);
    if (rc) return rc;
    const bool tip = LKCUDA_IS_TIP(dad);
    int grid = b16_grid(e, v.NP);
    next_result(e);
    launch_begin(e, KIND_THETA);
    if (tip) theta_derv_b16_kernel<true><<<grid, RED_THREADS, 0, e->stream>>>(e->v, e->b16tab, dad, node, len);
    else theta_derv_b16_kernel<false><<<grid, RED_THREADS, 0, e->stream>>>(e->v, e->b16tab, dad, node, len);
    uint64_t b = (uint64_t)v.B * 8;
    rc = launch_end(e, KIND_THETA, (uint64_t)v.NP * ((tip ? 2 * b + 1 : 3 * b) + 16), "theta_derv_b16_kernel");
    if (rc) return rc;
    e->theta_ready = true; e->theta_dad = dad; e->theta_node = node;
    return finish_result(e, result);
}

int lkcuda_derv(lkcuda_engine *e, double len, double *result) {
    LK_ENTER(e);
    if (!e->theta_ready) return fail(e, LKCUDA_ERR_STATE, "lkcuda_derv: theta_all not built");
    int rc = enqueue_derv(e, len);
    if (rc) return rc;
    return finish_result(e, result);
}

int lkcuda_lnl_from_buffer(lkcuda_engine *e, double len, double *result) {
    LK_ENTER(e);
    if (!e->theta_ready) return fail(e, LKCUDA_ERR_STATE, "lkcuda_lnl_from_buffer: theta_all not built");
    if (IS_SITEMODEL(e)) { int rcs = sitemodel_reduce(e, 3, 0, 0, len); return rcs ? rcs : finish_result(e, result); }
    const DevView &v = e->v;
    GroupShape gs = group_shape(v);
    int grid = red_grid(e, v.NP, gs.G);
    size_t sm = (size_t)((v.B + 3) & ~3) * 8;
    next_result(e);
    launch_begin(e, KIND_FROMBUF);
    if (gs.vec && v.S == 4 && v.K == 4) frombuf_b16_kernel<<<b16_grid(e, v.NP), RED_THREADS, 0, e->stream>>>(e->v, e->b16tab, len, NULL);
    else if (gs.vec) frombuf_kernel<true><<<grid, RED_THREADS, sm, e->stream>>>(e->v, len, gs.gshift, NULL);
    else frombuf_kernel<false><<<grid, RED_THREADS, sm, e->stream>>>(e->v, len, gs.gshift, NULL);
    int rc = launch_end(e, KIND_FROMBUF, (uint64_t)v.NP * ((uint64_t)v.B * 8 + 24), "frombuf_kernel");
    if (rc) return rc;
    if (v.U > 0 && v.P > 0) {
        launch_begin(e, KIND_MISC);
        asc_adjust_kernel<<<(int)((v.P + 255) / 256 < 1184 ? (v.P + 255) / 256 : 1184), 256, 0, e->stream>>>(e->v);
        rc = launch_end(e, KIND_MISC, 0, "asc_adjust_kernel");
        if (rc) return rc;
    }
    return finish_result(e, result);
}

/* Optimization::minimizeNewton (optimization.cpp:388-465) with computeFuncDerv = -(df, ddf) (phylotree.cpp:2135-2146).
   v1: the step sequence is driven from the host, one df/ddf launch + one 64-byte read-back per step. */
/* the persistent launch; build: 0 = theta_all is ready, 1 / 2 = build it from (dad, node) in the first step.  Returns 1 when the shape has
   no persistent kernel (the caller falls back to host-driven steps), 0 on success, < 0 = -(error code) */
static int newton_persistent(lkcuda_engine *e, int build, int32_t dad, int32_t node, double x1, double xguess, double x2, double xacc, int32_t maxNRStep, double *out,
                             int32_t xguess_slot = -1, int32_t len_slot = -1, bool async = false) {
    const DevView &v = e->v;
    GroupShape gs = group_shape(v);
    static const bool host_loop = getenv("LKCUDA_HOST_NEWTON") && atoi(getenv("LKCUDA_HOST_NEWTON")) != 0;      /* test knob */
    if (IS_SITEMODEL(e) || host_loop) return 1;
    const bool b16 = gs.vec && v.S == 4 && v.K == 4 && v.SU < 32;
    if (!b16) {
        /* any other block shape: theta_all must exist (the caller builds it with theta_kernel), then derv_kernel's pass in a loop */
        if (build != 0) return 1;
        static const bool generic_off = getenv("LKCUDA_NEWTON_B16_ONLY") && atoi(getenv("LKCUDA_NEWTON_B16_ONLY")) != 0;   /* test knob */
        if (generic_off) return 1;
        const size_t sm = (size_t)3 * ((v.B + 3) & ~3) * 8;
        const void *fn = gs.vec ? (const void *)newton_generic_kernel<true> : (const void *)newton_generic_kernel<false>;
        if (sm > 48 * 1024) return 1;
        int occ = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fn, RED_THREADS, sm) != cudaSuccess || occ < 1) { cudaGetLastError(); return 1; }
        int grid = red_grid(e, v.NP, gs.G);
        if (grid > occ * e->num_sms) grid = occ * e->num_sms;
        if (!e->d_newton_ctl && cudaMalloc((void **)&e->d_newton_ctl, 256) != cudaSuccess) return -cuda_fail(e, cudaGetLastError(), "cudaMalloc");
        if (cudaMemsetAsync(e->d_newton_ctl, 0, 256, e->stream) != cudaSuccess) return -cuda_fail(e, cudaGetLastError(), "cudaMemsetAsync");
        next_result(e);
        launch_begin(e, KIND_NEWTON);
        lkc::DevView vv = e->v;
        NewtonCtl *ctl = (NewtonCtl *)e->d_newton_ctl;
        int imax = maxNRStep, igs = gs.gshift, ixs = xguess_slot, ils = len_slot;
        void *args[] = {&vv, &ctl, &x1, &xguess, &x2, &xacc, &imax, &igs, &ixs, &ils};
        cudaError_t ce = cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(RED_THREADS), args, sm, e->stream);
        if (ce != cudaSuccess) return -cuda_fail(e, ce, "newton_generic_kernel");
        double r[LKCUDA_NRESULT];
        int rcn = launch_end(e, KIND_NEWTON, 0, "newton_generic_kernel");
        if (rcn) return -rcn;
        if (async) return 0;
        rcn = lkcuda_wait_result(e, r);
        if (rcn) return -rcn;
        const int nev = (int)r[2], status = (int)r[4];
        const uint64_t bytes = (uint64_t)nev * (uint64_t)v.NP * ((uint64_t)v.B * 8 + 16);
        e->kind_bytes[KIND_NEWTON] += bytes; e->bytes += bytes;
        if (v.peer_world > 1 && nev > 1) e->peer_seq += (unsigned int)(nev - 1);
        out[0] = r[0]; out[1] = r[1]; out[2] = r[2]; out[3] = r[3];
        if (status == 2) return -fail(e, LKCUDA_ERR_STATE, "Wrong computeFuncDerv");
        if (status == 3) return -fail(e, LKCUDA_ERR_STATE, "Maximum number of iterations exceeded in minimizeNewton");
        if (status != 1) return -fail(e, LKCUDA_ERR_STATE, "device Newton loop failed (status %d)", status);
        return 0;
    }
    int grid = b16_grid(e, v.NP);
    if (!e->newton_max_grid) {
        if (!e->d_newton_ctl && cudaMalloc((void **)&e->d_newton_ctl, 256) != cudaSuccess) return -cuda_fail(e, cudaGetLastError(), "cudaMalloc");
        int occ = 0, o2 = 0;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, newton_b16_kernel<0>, RED_THREADS, 0);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o2, newton_b16_kernel<1>, RED_THREADS, 0); if (o2 < occ) occ = o2;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o2, newton_b16_kernel<2>, RED_THREADS, 0); if (o2 < occ) occ = o2;
        e->newton_max_grid = occ * e->num_sms;
    }
    if (grid > e->newton_max_grid) grid = e->newton_max_grid;
    if (grid < 1) return 1;
    if (cudaMemsetAsync(e->d_newton_ctl, 0, 256, e->stream) != cudaSuccess) return -cuda_fail(e, cudaGetLastError(), "cudaMemsetAsync");
    next_result(e);
    launch_begin(e, KIND_NEWTON);
    lkc::DevView vv = e->v;
    NewtonCtl *ctl = (NewtonCtl *)e->d_newton_ctl;
    int idad = dad, inode = node, imax = maxNRStep, ixs = xguess_slot, ils = len_slot;
    lkc::B16Tab tab = e->b16tab;
    void *args[] = {&vv, &tab, &ctl, &idad, &inode, &x1, &xguess, &x2, &xacc, &imax, &ixs, &ils};
    const void *fn = build == 0 ? (const void *)newton_b16_kernel<0> : (build == 1 ? (const void *)newton_b16_kernel<1> : (const void *)newton_b16_kernel<2>);
    cudaError_t ce = cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(RED_THREADS), args, 0, e->stream);
    if (ce != cudaSuccess) return -cuda_fail(e, ce, "newton_b16_kernel");
    double r[LKCUDA_NRESULT];
    int rcn = launch_end(e, KIND_NEWTON, 0, "newton_b16_kernel");
    if (rcn) return -rcn;
    if (async) return 0;
    rcn = lkcuda_wait_result(e, r);
    if (rcn) return -rcn;
    const int nev = (int)r[2], status = (int)r[4];
    const uint64_t b = (uint64_t)v.B * 8;
    uint64_t bytes = (uint64_t)nev * (uint64_t)v.NP * (b + 16);
    if (build) bytes += (uint64_t)v.NP * (build == 1 ? 2 * b + 1 : 3 * b) - (uint64_t)v.NP * b;     /* the first step reads the two ends instead of theta_all */
    e->kind_bytes[KIND_NEWTON] += bytes; e->bytes += bytes;
    if (v.peer_world > 1 && nev > 1) e->peer_seq += (unsigned int)(nev - 1);     /* one exchange per evaluation */
    out[0] = r[0]; out[1] = r[1]; out[2] = r[2]; out[3] = r[3];
    if (status == 2) return -fail(e, LKCUDA_ERR_STATE, "Wrong computeFuncDerv");
    if (status == 3) return -fail(e, LKCUDA_ERR_STATE, "Maximum number of iterations exceeded in minimizeNewton");
    if (status != 1) return -fail(e, LKCUDA_ERR_STATE, "device Newton loop failed (status %d)", status);
    return 0;
}

/* theta_all build + minimizeNewton on the branch (dad, node) in one call: optimizeOneBranch's device part (phylotree.cpp:2148-2170).
   DNA + 4 categories: ONE persistent launch; other shapes: lkcuda_theta_build + lkcuda_newton_branch. */
int lkcuda_theta_newton(lkcuda_engine *e, int32_t dad, int32_t node, double x1, double xguess, double x2, double xacc, int32_t maxNRStep, double *out) {
    LK_ENTER(e);
    if (!out) return fail(e, LKCUDA_ERR_ARG, "null output");
    if (node < 0 || node >= e->cfg.nvec || !valid_end(e, dad) || dad == node) return fail(e, LKCUDA_ERR_ARG, "lkcuda_theta_newton: bad vector / taxon id");
    int rc = check_ready(e, "lkcuda_theta_newton");
    if (rc) return rc;
    rc = states_barrier(e);
    if (rc) return rc;
    int pr = newton_persistent(e, LKCUDA_IS_TIP(dad) ? 1 : 2, dad, node, x1, xguess, x2, xacc, maxNRStep, out);
    if (pr <= 0) {
        if (pr == 0) { e->theta_ready = true; e->theta_dad = dad; e->theta_node = node; }
        return -pr;
    }
    rc = lkcuda_theta_build(e, dad, node);
    return rc ? rc : lkcuda_newton_branch(e, x1, xguess, x2, xacc, maxNRStep, out);
}

int lkcuda_newton_branch(lkcuda_engine *e, double x1, double xguess, double x2, double xacc, int32_t maxNRStep, double *out) {
    LK_ENTER(e);
    if (!out) return fail(e, LKCUDA_ERR_ARG, "null output");
    if (!e->theta_ready) return fail(e, LKCUDA_ERR_STATE, "lkcuda_newton_branch: theta_all not built");
    {
        int pr = newton_persistent(e, 0, 0, 0, x1, xguess, x2, xacc, maxNRStep, out);
        if (pr <= 0) return -pr;
    }
    int j, nev = 0, rc;
    double df, dx, f, temp, xh, xl, rts, rts_old, d2l, r[LKCUDA_NRESULT];
#define FUNC_DERV(x) do { rc = enqueue_derv(e, (x)); if (rc) return rc; rc = lkcuda_wait_result(e, r); if (rc) return rc; \
                          f = -r[LKCUDA_R_DF]; df = -r[LKCUDA_R_DDF]; nev++; out[3] = (x); } while (0)
#define NR_RETURN(x) do { out[0] = (x); out[1] = d2l; out[2] = (double)nev; return LKCUDA_OK; } while (0)
    rts = xguess;
    if (rts < x1) rts = x1;
    if (rts > x2) rts = x2;
    FUNC_DERV(rts);
    d2l = df;
    if (!isfinite(f) || !isfinite(df)) return fail(e, LKCUDA_ERR_STATE, "Wrong computeFuncDerv");
    if (df >= 0.0 && fabs(f) < xacc) NR_RETURN(rts);
    if (f < 0.0) { xl = rts; xh = x2; } else { xh = rts; xl = x1; }
    dx = fabs(xh - xl);
    for (j = 1; j <= maxNRStep; j++) {
        rts_old = rts;
        if ((df <= 0.0) || (((rts - xh) * df - f) * ((rts - xl) * df - f) >= 0.0)) {
            dx = 0.5 * (xh - xl); rts = xl + dx; d2l = df;
            if (xl == rts) NR_RETURN(rts);
        } else {
            dx = f / df; temp = rts; rts -= dx; d2l = df;
            if (temp == rts) NR_RETURN(rts);
        }
        if (fabs(dx) < xacc || (j == maxNRStep)) NR_RETURN(rts_old);
        FUNC_DERV(rts);
        if (!isfinite(f) || !isfinite(df)) return fail(e, LKCUDA_ERR_STATE, "Wrong computeFuncDerv");
        if (df > 0.0 && fabs(f) < xacc) { d2l = df; NR_RETURN(rts); }
        if (f < 0.0) xl = rts; else xh = rts;
    }
    return fail(e, LKCUDA_ERR_STATE, "Maximum number of iterations exceeded in minimizeNewton");
#undef FUNC_DERV
#undef NR_RETURN
}

/* ---- device length slots (lkcuda.h): a whole NNI evaluation queued without a host wait ---- */
int lkcuda_theta_newton_async(lkcuda_engine *e, int32_t dad, int32_t node, double x1, double xguess, int32_t xguess_slot, double x2,
                              double xacc, int32_t maxNRStep, int32_t len_slot) {
    LK_ENTER(e);
    if (node < 0 || node >= e->cfg.nvec || !valid_end(e, dad) || dad == node) return fail(e, LKCUDA_ERR_ARG, "lkcuda_theta_newton_async: bad vector / taxon id");
    if (len_slot < 0 || len_slot >= LKCUDA_NLENSLOT || xguess_slot >= LKCUDA_NLENSLOT) return fail(e, LKCUDA_ERR_ARG, "lkcuda_theta_newton_async: bad length slot");
    if (e->v.peer_world > 1) return fail(e, LKCUDA_ERR_STATE, "lkcuda_theta_newton_async: not with peers attached (the exchange counts evaluations on the host)");
    int rc = check_ready(e, "lkcuda_theta_newton_async");
    if (rc) return rc;
    rc = states_barrier(e);
    if (rc) return rc;
    const DevView &v = e->v;
    GroupShape gs = group_shape(v);
    const bool b16 = gs.vec && v.S == 4 && v.K == 4 && v.SU < 32;
    int pr;
    if (b16) pr = newton_persistent(e, LKCUDA_IS_TIP(dad) ? 1 : 2, dad, node, x1, xguess, x2, xacc, maxNRStep, NULL, xguess_slot, len_slot, true);
    else {
        rc = lkcuda_theta_build(e, dad, node);
        if (rc) return rc;
        pr = newton_persistent(e, 0, 0, 0, x1, xguess, x2, xacc, maxNRStep, NULL, xguess_slot, len_slot, true);
    }
    if (pr > 0) return fail(e, LKCUDA_ERR_STATE, "lkcuda_theta_newton_async: this shape has no persistent Newton kernel");
    if (pr < 0) return -pr;
    e->theta_ready = true; e->theta_dad = dad; e->theta_node = node;
    return LKCUDA_OK;
}

/* frombuf kernels take the length by value; a one-thread kernel cannot hand it over, so the B = 16 kernel and the generic one read the
   slot themselves when given one */
int lkcuda_lnl_from_buffer_slot(lkcuda_engine *e, int32_t len_slot, double *result) {
    LK_ENTER(e);
    if (len_slot < 0 || len_slot >= LKCUDA_NLENSLOT) return fail(e, LKCUDA_ERR_ARG, "lkcuda_lnl_from_buffer_slot: bad length slot");
    if (!e->theta_ready) return fail(e, LKCUDA_ERR_STATE, "lkcuda_lnl_from_buffer_slot: theta_all not built");
    if (IS_SITEMODEL(e)) return fail(e, LKCUDA_ERR_STATE, "lkcuda_lnl_from_buffer_slot: not for site-specific models");
    const DevView &v = e->v;
    GroupShape gs = group_shape(v);
    int grid = red_grid(e, v.NP, gs.G);
    size_t sm = (size_t)((v.B + 3) & ~3) * 8;
    next_result(e);
    launch_begin(e, KIND_FROMBUF);
    const double *lp = v.lens + len_slot;
    if (gs.vec && v.S == 4 && v.K == 4) frombuf_b16_kernel<<<b16_grid(e, v.NP), RED_THREADS, 0, e->stream>>>(e->v, e->b16tab, 0.0, lp);
    else if (gs.vec) frombuf_kernel<true><<<grid, RED_THREADS, sm, e->stream>>>(e->v, 0.0, gs.gshift, lp);
    else frombuf_kernel<false><<<grid, RED_THREADS, sm, e->stream>>>(e->v, 0.0, gs.gshift, lp);
    int rc = launch_end(e, KIND_FROMBUF, (uint64_t)v.NP * ((uint64_t)v.B * 8 + 24), "frombuf_kernel");
    if (rc) return rc;
    if (v.U > 0 && v.P > 0) {
        launch_begin(e, KIND_MISC);
        asc_adjust_kernel<<<(int)((v.P + 255) / 256 < 1184 ? (v.P + 255) / 256 : 1184), 256, 0, e->stream>>>(e->v);
        rc = launch_end(e, KIND_MISC, 0, "asc_adjust_kernel");
        if (rc) return rc;
    }
    return finish_result(e, result);
}

int lkcuda_len_slots_set(lkcuda_engine *e, int32_t first, int32_t n, const double *values) {
    LK_ENTER(e);
    if (first < 0 || n < 0 || first + n > LKCUDA_NLENSLOT || (n && !values)) return fail(e, LKCUDA_ERR_ARG, "lkcuda_len_slots_set: bad range");
    if (n == 0) return LKCUDA_OK;
    LK_CUDA(e, cudaStreamSynchronize(e->stream));           /* h_lens may still be the source of an earlier copy */
    memcpy(e->h_lens, values, sizeof(double) * n);
    LK_CUDA(e, cudaMemcpyAsync(e->d_lens + first, e->h_lens, sizeof(double) * n, cudaMemcpyHostToDevice, e->stream));
    return LKCUDA_OK;
}

int lkcuda_len_slots_fetch(lkcuda_engine *e, double *lens, double *info) {
    LK_ENTER(e);
    LK_CUDA(e, cudaMemcpyAsync(e->h_lens, e->d_lens, sizeof(double) * 5 * LKCUDA_NLENSLOT, cudaMemcpyDeviceToHost, e->stream));
    LK_CUDA(e, cudaStreamSynchronize(e->stream));
    if (lens) memcpy(lens, e->h_lens, sizeof(double) * LKCUDA_NLENSLOT);
    if (info) memcpy(info, e->h_lens + LKCUDA_NLENSLOT, sizeof(double) * 4 * LKCUDA_NLENSLOT);
    return upload_verdict(e);
}

/* the correction kernel on the engine's stream, the copy on a stream of its own: later kernels (the next branch's theta_all, df/ddf)
   overlap the transfer; lkcuda_fetch_wait (or the synchronous variant) ends it.  dst should be pinned host memory. */
int lkcuda_fetch_pattern_lh_corrected_async(lkcuda_engine *e, int32_t vec, double *dst) {
    LK_ENTER(e);
    if (!dst || vec < 0 || vec >= e->cfg.nvec) return fail(e, LKCUDA_ERR_ARG, "lkcuda_fetch_pattern_lh_corrected_async: bad argument");
    const DevView &v = e->v;
    if (!v.P) return LKCUDA_OK;
    if (!e->d_plh_out) {
        LK_CUDA(e, cudaMalloc((void **)&e->d_plh_out, (size_t)v.P * 8));
        e->total_dev_bytes += (uint64_t)v.P * 8;
    }
    if (!e->dstream) {
        LK_CUDA(e, cudaStreamCreateWithFlags(&e->dstream, cudaStreamNonBlocking));
        LK_CUDA(e, cudaEventCreateWithFlags(&e->ev_plh_ready, cudaEventDisableTiming));
        LK_CUDA(e, cudaEventCreateWithFlags(&e->ev_plh_done, cudaEventDisableTiming));
    }
    if (e->plh_pending) LK_CUDA(e, cudaStreamWaitEvent(e->stream, e->ev_plh_done, 0));     /* the staging buffer is still on its way out */
    int grid = red_grid(e, v.P, 1);
    launch_begin(e, KIND_MISC);
    pattern_lh_corrected_kernel<<<grid, RED_THREADS, 0, e->stream>>>(e->v, vec, e->d_plh_out);
    int rc = launch_end(e, KIND_MISC, 0, "pattern_lh_corrected_kernel");
    if (rc) return rc;
    LK_CUDA(e, cudaEventRecord(e->ev_plh_ready, e->stream));
    LK_CUDA(e, cudaStreamWaitEvent(e->dstream, e->ev_plh_ready, 0));
    LK_CUDA(e, cudaMemcpyAsync(dst, e->d_plh_out, (size_t)v.P * 8, cudaMemcpyDeviceToHost, e->dstream));
    LK_CUDA(e, cudaEventRecord(e->ev_plh_done, e->dstream));
    e->plh_pending = true;
    return LKCUDA_OK;
}

int lkcuda_fetch_wait(lkcuda_engine *e) {
    LK_ENTER(e);
    if (e->plh_pending) {
        LK_CUDA(e, cudaEventSynchronize(e->ev_plh_done));
        e->plh_pending = false;
    }
    return upload_verdict(e);
}

int lkcuda_fetch_pattern_lh_corrected(lkcuda_engine *e, int32_t vec, double *dst) {
    LK_ENTER(e);
    if (!dst || vec >= e->cfg.nvec) return fail(e, LKCUDA_ERR_ARG, "lkcuda_fetch_pattern_lh_corrected: bad argument");
    if (vec < 0) return lkcuda_fetch_pattern_lh(e, dst);
    const DevView &v = e->v;
    if (v.P) {
        if (!e->d_plh_out) {
            LK_CUDA(e, cudaMalloc((void **)&e->d_plh_out, (size_t)v.P * 8));
            e->total_dev_bytes += (uint64_t)v.P * 8;
        }
        if (e->plh_pending) { LK_CUDA(e, cudaEventSynchronize(e->ev_plh_done)); e->plh_pending = false; }
        int grid = red_grid(e, v.P, 1);
        launch_begin(e, KIND_MISC);
        pattern_lh_corrected_kernel<<<grid, RED_THREADS, 0, e->stream>>>(e->v, vec, e->d_plh_out);
        int rc = launch_end(e, KIND_MISC, 0, "pattern_lh_corrected_kernel");
        if (rc) return rc;
        LK_CUDA(e, cudaMemcpyAsync(dst, e->d_plh_out, (size_t)v.P * 8, cudaMemcpyDeviceToHost, e->stream));
    }
    LK_CUDA(e, cudaStreamSynchronize(e->stream));
    return upload_verdict(e);
}

int lkcuda_vector_scale_factor(lkcuda_engine *e, int32_t vec, double *out) {
    LK_ENTER(e);
    if (!out || vec < 0 || vec >= e->cfg.nvec) return fail(e, LKCUDA_ERR_ARG, "lkcuda_vector_scale_factor: bad argument");
    if (!e->weights_set) return fail(e, LKCUDA_ERR_STATE, "lkcuda_vector_scale_factor: pattern weights not set");
    const DevView &v = e->v;
    int grid = red_grid(e, v.P, 1);
    launch_begin(e, KIND_MISC);
    scale_factor_kernel<<<grid, RED_THREADS, 0, e->stream>>>(e->v, vec);
    int rc = launch_end(e, KIND_MISC, 0, "scale_factor_kernel");
    if (rc) return rc;
    LK_CUDA(e, cudaMemcpyAsync(e->h_newton_out, e->v.result + 7, sizeof(double), cudaMemcpyDeviceToHost, e->stream));
    LK_CUDA(e, cudaStreamSynchronize(e->stream));
    *out = e->h_newton_out[0];
    return LKCUDA_OK;
}

}  // extern "C"
