/*
 * lk_oracle.c — ORACLE: CPU restatement of the reference's likelihood kernels.  TEST INFRASTRUCTURE.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may build, link or call this
 * file.  It is never part of the product path (the product is iq-tree_b200/csrc, CUDA only).
 *
 * It implements the same C ABI as the CUDA engine (include/lkcuda.h) with plain scalar C following
 * the reference's generic kernels statement by statement (same loop nests, same summation order as
 * phylotreesse.cpp), so that the host mirror can be linked against either and the two compared.
 * Parity pin: checked against the golden fixtures produced by the UNMODIFIED reference
 * (oracle/_ref/ref_harness, see oracle/make_golden.py and tests/test_oracle_golden.py).
 *
 * Reference functions restated (IQ-TREE 1.4.3-beta, /root/reference):
 *   partial likelihoods  phylotreesse.cpp:581-1010 (generic), phylokernel.h:71-483 (SIMD semantics of the
 *                        scaling test), phylokernelmixture.h:21-460, phylokernelmixrate.h:23-449
 *   branch lnL           phylotreesse.cpp:1158-1339, phylokernel.h:734-1020
 *   theta + df/ddf       phylotreesse.cpp:1013-1155, phylokernel.h:486-730
 *   lnL from buffer      phylokernel.h:1023-1192
 *   Newton               optimization.cpp:388-465
 */
#include "../include/lkcuda.h"
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* phylotree.h:52-54 */
#define SCALING_THRESHOLD_INVER 115792089237316195423570985008687907853269984665640564039457584007913129639936.0
#define SCALING_THRESHOLD (1.0 / SCALING_THRESHOLD_INVER)
#define LOG_SCALING_THRESHOLD log(SCALING_THRESHOLD)

struct lkcuda_engine {
    lkcuda_config cfg;
    size_t S, K, E, block, nptn_all;
    uint8_t *states;                 /* [ntaxa][P+U] */
    double *ptn_freq, *ptn_invar;    /* [P], [P+U]   */
    double *eval, *evec, *inv_evec;  /* [E*S], [E*S*S], [E*S*S] */
    int32_t *blk_eigen; double *blk_rate, *blk_weight;
    double *tip_lh;                  /* [(SU+1)][E][S] */
    double *arena;                   /* [nvec][vec_stride], own: vec_stride = (P+U)*block */
    int16_t *scale;                  /* [nvec][sc_stride], own: sc_stride = P+U */
    double *own_arena; int16_t *own_scale; size_t vec_stride, sc_stride; int32_t own_nvec;   /* lkcuda_attach_arena borrows */
    double *theta;                   /* [(P+U)*block] */
    double *pattern_lh, *pattern_lh_cat;
    int theta_ready, theta_dad_tip;
    double theta_scale_sum;
    int16_t *theta_scale_num;        /* [U] summed scale_num of the unobserved patterns */
    double result_own[LKCUDA_NRESULT];
    double *result;                  /* result_own or a caller-supplied block (lkcuda_set_result_buffer) */
    uint64_t launches, bytes;
    /* RELL matrix and the collapsed _pattern_lh_cat work matrix */
    float *boot; int32_t nboot; double *rell; double *w; int32_t ncol;
    /* site-specific models */
    int32_t *ptn_model; uint64_t *state_mask;
    double lens[5 * LKCUDA_NLENSLOT];   /* length slots (lkcuda.h): the oracle executes "queued" work at once, so a slot is just a variable */
    char err[256];
};

/* a child's branch length: the value, or the slot named in op->reserved (LKCUDA_LEN_REFS) */
static double op_len(const lkcuda_engine *e, const lkcuda_partial_op *op, int right) {
    int ref = right ? (op->reserved >> 16) & 0xffff : op->reserved & 0xffff;
    return ref ? e->lens[ref - 1] : (right ? op->len_right : op->len_left);
}

static char g_create_err[256];

static int fail(lkcuda_engine *e, int code, const char *msg) {
    snprintf(e ? e->err : g_create_err, 256, "%s", msg);
    return code;
}

int lkcuda_device_count(int32_t *count) { if (count) *count = 0; return LKCUDA_OK; }

const char *lkcuda_last_error(const lkcuda_engine *e) { return e ? e->err : g_create_err; }

int lkcuda_create(const lkcuda_config *cfg, lkcuda_engine **out) {
    if (!cfg || !out) return fail(NULL, LKCUDA_ERR_ARG, "null argument");
    if (cfg->nstates < 1 || cfg->nblocks < 1 || cfg->neigen < 1 || cfg->ntaxa < 1 || cfg->nptn < 0 ||
        cfg->nunobs < 0 || cfg->nvec < 0 || cfg->state_unknown < cfg->nstates)
        return fail(NULL, LKCUDA_ERR_ARG, "invalid configuration");
    lkcuda_engine *e = (lkcuda_engine *)calloc(1, sizeof(*e));
    if (e) e->result = e->result_own;
    if (!e) return fail(NULL, LKCUDA_ERR_NOMEM, "out of memory");
    e->cfg = *cfg;
    e->S = cfg->nstates; e->K = cfg->nblocks; e->E = cfg->neigen;
    e->block = e->S * e->K; e->nptn_all = (size_t)(cfg->nptn + cfg->nunobs);
    size_t n = e->nptn_all ? e->nptn_all : 1;
    e->states = (uint8_t *)calloc((size_t)cfg->ntaxa * n, 1);
    e->ptn_freq = (double *)calloc(n, 8); e->ptn_invar = (double *)calloc(n, 8);
    e->eval = (double *)calloc(e->E * e->S, 8);
    e->evec = (double *)calloc(e->E * e->S * e->S, 8); e->inv_evec = (double *)calloc(e->E * e->S * e->S, 8);
    e->blk_eigen = (int32_t *)calloc(e->K, 4); e->blk_rate = (double *)calloc(e->K, 8); e->blk_weight = (double *)calloc(e->K, 8);
    e->tip_lh = (double *)calloc((cfg->flags & LKCUDA_FLAG_SITEMODEL) ? 1 : (size_t)(cfg->state_unknown + 1) * e->E * e->S, 8);
    e->arena = (double *)calloc((size_t)(cfg->nvec ? cfg->nvec : 1) * n * e->block, 8);
    e->scale = (int16_t *)calloc((size_t)(cfg->nvec ? cfg->nvec : 1) * n, 2);
    e->theta = (double *)calloc(n * e->block, 8);
    e->pattern_lh = (double *)calloc(n, 8); e->pattern_lh_cat = (double *)calloc(n * e->K, 8);
    e->theta_scale_num = (int16_t *)calloc((size_t)cfg->nunobs + 1, 2);
    e->own_arena = e->arena; e->own_scale = e->scale; e->vec_stride = n * e->block; e->sc_stride = n; e->own_nvec = cfg->nvec;
    if (!e->states || !e->arena || !e->scale || !e->theta || !e->pattern_lh_cat) {
        lkcuda_destroy(e);
        return fail(NULL, LKCUDA_ERR_NOMEM, "out of memory");
    }
    *out = e;
    return LKCUDA_OK;
}

int lkcuda_destroy(lkcuda_engine *e) {
    if (!e) return LKCUDA_OK;
    free(e->states); free(e->ptn_freq); free(e->ptn_invar); free(e->eval); free(e->evec); free(e->inv_evec);
    free(e->blk_eigen); free(e->blk_rate); free(e->blk_weight); free(e->tip_lh); free(e->own_arena); free(e->own_scale);
    free(e->theta); free(e->pattern_lh); free(e->pattern_lh_cat); free(e->theta_scale_num);
    free(e->boot); free(e->rell); free(e->w); free(e->ptn_model); free(e->state_mask);
    free(e);
    return LKCUDA_OK;
}

int lkcuda_memory_bytes(const lkcuda_engine *e, uint64_t *bytes) {
    if (!e || !bytes) return LKCUDA_ERR_ARG;
    *bytes = 0;
    return LKCUDA_OK;
}
/* caller-owned arenas: plain host memory here */
int lkcuda_device_malloc(int32_t device, uint64_t bytes, void **ptr) {
    (void)device;
    if (!ptr) return LKCUDA_ERR_ARG;
    *ptr = aligned_alloc(256, ((bytes ? (size_t)bytes : 1) + 255) / 256 * 256);
    if (!*ptr) return LKCUDA_ERR_NOMEM;
    memset(*ptr, 0, bytes ? (size_t)bytes : 1);
    return LKCUDA_OK;
}
int lkcuda_device_free(int32_t device, void *ptr) { (void)device; free(ptr); return LKCUDA_OK; }
int lkcuda_host_register(lkcuda_engine *e, void *ptr, uint64_t bytes) { (void)bytes; return e && ptr ? LKCUDA_OK : LKCUDA_ERR_ARG; }
int lkcuda_host_unregister(lkcuda_engine *e, void *ptr) { (void)ptr; return e ? LKCUDA_OK : LKCUDA_ERR_ARG; }
int lkcuda_arena_layout(const lkcuda_engine *e, int64_t *vec_stride, int64_t *scale_stride) {
    if (!e) return LKCUDA_ERR_ARG;
    if (vec_stride) *vec_stride = (int64_t)e->vec_stride;
    if (scale_stride) *scale_stride = (int64_t)e->sc_stride;
    return LKCUDA_OK;
}
int lkcuda_attach_arena(lkcuda_engine *e, void *partial_lh, int64_t vec_stride, void *scale_num, int64_t scale_stride, int32_t nvec) {
    if (!e) return LKCUDA_ERR_ARG;
    if (!partial_lh && !scale_num) {
        e->arena = e->own_arena; e->scale = e->own_scale; e->vec_stride = e->nptn_all * e->block; e->sc_stride = e->nptn_all;
        e->cfg.nvec = e->own_nvec; e->theta_ready = 0;
        return LKCUDA_OK;
    }
    if (!partial_lh || !scale_num || nvec < 0 || vec_stride < (int64_t)(e->nptn_all * e->block) || scale_stride < (int64_t)e->nptn_all)
        return fail(e, LKCUDA_ERR_ARG, "lkcuda_attach_arena: bad arena");
    /* same contract as the CUDA engine (whole 32-byte sectors) */
    if (((uintptr_t)partial_lh & 31) || (e->block % 4 == 0 && vec_stride % 4 != 0) || ((uintptr_t)scale_num & 1))
        return fail(e, LKCUDA_ERR_ARG, "lkcuda_attach_arena: base must be 32-byte aligned and the stride a multiple of 4 doubles");
    e->arena = (double *)partial_lh; e->scale = (int16_t *)scale_num;
    e->vec_stride = (size_t)vec_stride; e->sc_stride = (size_t)scale_stride; e->cfg.nvec = nvec; e->theta_ready = 0;
    return LKCUDA_OK;
}

int lkcuda_set_stream(lkcuda_engine *e, void *s) { (void)s; return e ? LKCUDA_OK : LKCUDA_ERR_ARG; }
int lkcuda_synchronize(lkcuda_engine *e) { return e ? LKCUDA_OK : LKCUDA_ERR_ARG; }

int lkcuda_set_tip_states(lkcuda_engine *e, const uint8_t *states) {
    if (!e || !states) return LKCUDA_ERR_ARG;
    size_t n = (size_t)e->cfg.ntaxa * e->nptn_all;
    for (size_t i = 0; i < n; i++)
        if (states[i] > e->cfg.state_unknown) return fail(e, LKCUDA_ERR_ARG, "tip state out of range");
    memcpy(e->states, states, n);
    return LKCUDA_OK;
}

/* the oracle has no copy engine: the pipelined upload is the plain one */
int lkcuda_set_tip_states_async(lkcuda_engine *e, const uint8_t *states, int32_t nchunks) { (void)nchunks; return lkcuda_set_tip_states(e, states); }

int lkcuda_set_pattern_weights(lkcuda_engine *e, const double *ptn_freq, const double *ptn_invar) {
    if (!e || !ptn_freq || !ptn_invar) return LKCUDA_ERR_ARG;
    memcpy(e->ptn_freq, ptn_freq, (size_t)e->cfg.nptn * 8);
    memcpy(e->ptn_invar, ptn_invar, e->nptn_all * 8);
    return LKCUDA_OK;
}

int lkcuda_set_model(lkcuda_engine *e, const double *eval, const double *evec, const double *inv_evec,
                     const int32_t *blk_eigen, const double *blk_rate, const double *blk_weight,
                     const double *tip_partial_lh) {
    const int sm = e && (e->cfg.flags & LKCUDA_FLAG_SITEMODEL);
    if (!e || !eval || !evec || !inv_evec || !blk_rate || !blk_weight || (!sm && (!blk_eigen || !tip_partial_lh)))
        return LKCUDA_ERR_ARG;
    if (!sm)
        for (size_t k = 0; k < e->K; k++)
            if (blk_eigen[k] < 0 || (size_t)blk_eigen[k] >= e->E) return fail(e, LKCUDA_ERR_ARG, "blk_eigen out of range");
    memcpy(e->eval, eval, e->E * e->S * 8);
    memcpy(e->evec, evec, e->E * e->S * e->S * 8);
    memcpy(e->inv_evec, inv_evec, e->E * e->S * e->S * 8);
    if (!sm) memcpy(e->blk_eigen, blk_eigen, e->K * 4);
    memcpy(e->blk_rate, blk_rate, e->K * 8);
    memcpy(e->blk_weight, blk_weight, e->K * 8);
    if (!sm) memcpy(e->tip_lh, tip_partial_lh, (size_t)(e->cfg.state_unknown + 1) * e->E * e->S * 8);
    e->theta_ready = 0;
    return LKCUDA_OK;
}

static double *vec_lh(lkcuda_engine *e, int32_t v) { return e->arena + (size_t)v * e->vec_stride; }
static int16_t *vec_sc(lkcuda_engine *e, int32_t v) { return e->scale + (size_t)v * e->sc_stride; }
static int bad_end(lkcuda_engine *e, int32_t id) {
    return id >= 0 ? id >= e->cfg.nvec : LKCUDA_TIP_TAXON(id) >= e->cfg.ntaxa;
}

/* echild[k][x][i] = evec[e(k)][x][i] * exp(eval[e(k)][i] * rate_k * len)   (phylotreesse.cpp:654-663) */
static void build_echild(lkcuda_engine *e, double len, double *echild) {
    size_t S = e->S;
    for (size_t k = 0; k < e->K; k++) {
        const double *eval = e->eval + (size_t)e->blk_eigen[k] * S;
        const double *evec = e->evec + (size_t)e->blk_eigen[k] * S * S;
        double len_child = e->blk_rate[k] * len;
        for (size_t x = 0; x < S; x++)
            for (size_t i = 0; i < S; i++)
                echild[k * S * S + x * S + i] = evec[x * S + i] * exp(eval[i] * len_child);
    }
}

/* partial_lh_leaf[state][k][x] = sum_i echild[k][x][i] * tip_partial_lh[state][e(k)][i]; unknown row = 1
   (phylotreesse.cpp:666-683; mixture indexing phylokernelmixture.h:150) */
static void build_leaf_table(lkcuda_engine *e, const double *echild, double *tab) {
    size_t S = e->S, SU = (size_t)e->cfg.state_unknown;
    for (size_t st = 0; st < SU; st++)
        for (size_t k = 0; k < e->K; k++) {
            const double *tip = e->tip_lh + (st * e->E + (size_t)e->blk_eigen[k]) * S;
            for (size_t x = 0; x < S; x++) {
                double v = 0.0;
                for (size_t i = 0; i < S; i++) v += echild[k * S * S + x * S + i] * tip[i];
                tab[st * e->block + k * S + x] = v;
            }
        }
    for (size_t x = 0; x < e->block; x++) tab[SU * e->block + x] = 1.0;
}

/* the scaling epilogue shared by all cases (phylokernel.h:380-392; generic phylotreesse.cpp:893-919) */
static void scale_pattern(lkcuda_engine *e, double *partial_lh, int16_t *scale_num, size_t ptn, double lh_max) {
    size_t S = e->S;
    if (!(lh_max < SCALING_THRESHOLD)) return;
    if ((e->cfg.flags & LKCUDA_FLAG_ZERO_GUARD) && lh_max == 0.0) {
        for (size_t k = 0; k < e->K; k++)
            memcpy(&partial_lh[k * S], e->tip_lh + ((size_t)e->cfg.state_unknown * e->E + (size_t)e->blk_eigen[k]) * S, S * 8);
        scale_num[ptn] += 4;
        return;
    }
    if ((e->cfg.flags & LKCUDA_FLAG_SCALE_IGNORES_INVAR) || e->ptn_invar[ptn] == 0.0) {
        for (size_t i = 0; i < e->block; i++) partial_lh[i] *= SCALING_THRESHOLD_INVER;
        scale_num[ptn] += 1;
    }
}

static int sm_partial_one(lkcuda_engine *e, const lkcuda_partial_op *op);
static int sm_branch(lkcuda_engine *e, int32_t dad, int32_t node, double len, double *result);
static int sm_theta(lkcuda_engine *e, int32_t dad, int32_t node);
static int sm_derv(lkcuda_engine *e, double len, double *df_out, double *ddf_out);
static int sm_frombuf(lkcuda_engine *e, double len, double *result);
#define IS_SM(e) ((e)->cfg.flags & LKCUDA_FLAG_SITEMODEL)

static int partial_one(lkcuda_engine *e, const lkcuda_partial_op *op) {
    if (IS_SM(e)) return sm_partial_one(e, op);
    size_t S = e->S, K = e->K, block = e->block, nptn = e->nptn_all, SU = (size_t)e->cfg.state_unknown;
    if (op->dst < 0 || op->dst >= e->cfg.nvec || bad_end(e, op->left) || bad_end(e, op->right))
        return fail(e, LKCUDA_ERR_ARG, "partial op: bad vector / taxon id");
    if (!LKCUDA_IS_TIP(op->left) && LKCUDA_IS_TIP(op->right))
        return fail(e, LKCUDA_ERR_ARG, "partial op: a leaf child must be 'left' (phylokernel.h:116-121)");
    double *eleft = (double *)malloc(block * S * 8), *eright = (double *)malloc(block * S * 8);
    double *tab_l = NULL, *tab_r = NULL;
    double *tmp = (double *)malloc(S * 8);
    build_echild(e, op_len(e, op, 0), eleft);
    build_echild(e, op_len(e, op, 1), eright);
    if (LKCUDA_IS_TIP(op->left)) { tab_l = (double *)malloc((SU + 1) * block * 8); build_leaf_table(e, eleft, tab_l); }
    if (LKCUDA_IS_TIP(op->right)) { tab_r = (double *)malloc((SU + 1) * block * 8); build_leaf_table(e, eright, tab_r); }
    double *dst = vec_lh(e, op->dst);
    int16_t *dsc = vec_sc(e, op->dst);
    const uint8_t *st_l = LKCUDA_IS_TIP(op->left) ? e->states + (size_t)LKCUDA_TIP_TAXON(op->left) * nptn : NULL;
    const uint8_t *st_r = LKCUDA_IS_TIP(op->right) ? e->states + (size_t)LKCUDA_TIP_TAXON(op->right) * nptn : NULL;
    const double *lh_l = st_l ? NULL : vec_lh(e, op->left), *lh_r = st_r ? NULL : vec_lh(e, op->right);
    const int16_t *sc_l = st_l ? NULL : vec_sc(e, op->left), *sc_r = st_r ? NULL : vec_sc(e, op->right);

    for (size_t ptn = 0; ptn < nptn; ptn++) {
        double *partial_lh = dst + ptn * block;
        double lh_max = 0.0;
        /* scale_num: 0 for tip-tip (phylokernel.h:247), right's for tip-internal (:290), sum otherwise (:418) */
        dsc[ptn] = (int16_t)((sc_l ? sc_l[ptn] : 0) + (sc_r ? sc_r[ptn] : 0));
        for (size_t k = 0; k < K; k++) {
            const double *inv_evec = e->inv_evec + (size_t)e->blk_eigen[k] * S * S;
            for (size_t x = 0; x < S; x++) {
                double vleft, vright;
                if (st_l) vleft = tab_l[(size_t)st_l[ptn] * block + k * S + x];
                else {
                    vleft = 0.0;
                    for (size_t i = 0; i < S; i++) vleft += eleft[k * S * S + x * S + i] * lh_l[ptn * block + k * S + i];
                }
                if (st_r) vright = tab_r[(size_t)st_r[ptn] * block + k * S + x];
                else {
                    vright = 0.0;
                    for (size_t i = 0; i < S; i++) vright += eright[k * S * S + x * S + i] * lh_r[ptn * block + k * S + i];
                }
                tmp[x] = vleft * vright;
            }
            for (size_t i = 0; i < S; i++) {
                double res = 0.0;
                for (size_t x = 0; x < S; x++) res += tmp[x] * inv_evec[i * S + x];
                partial_lh[k * S + i] = res;
                if (fabs(res) > lh_max) lh_max = fabs(res);
            }
        }
        /* tip-tip never scales (phylokernel.h:246-281) */
        if (!(st_l && st_r)) scale_pattern(e, partial_lh, dsc, ptn, lh_max);
    }
    free(eleft); free(eright); free(tab_l); free(tab_r); free(tmp);
    e->launches++;
    return LKCUDA_OK;
}

int lkcuda_partial_batch(lkcuda_engine *e, const lkcuda_partial_op *ops, int32_t nops,
                         const int32_t *level_start, int32_t nlevels) {
    if (!e || (nops && !ops) || nops < 0) return LKCUDA_ERR_ARG;
    if (level_start) {
        if (nlevels < 0 || level_start[0] != 0 || level_start[nlevels] != nops) return fail(e, LKCUDA_ERR_ARG, "bad level_start");
        for (int32_t l = 0; l < nlevels; l++)
            if (level_start[l] > level_start[l + 1]) return fail(e, LKCUDA_ERR_ARG, "bad level_start");
    }
    for (int32_t i = 0; i < nops; i++) {
        int rc = partial_one(e, &ops[i]);
        if (rc) return rc;
    }
    e->theta_ready = 0;
    return LKCUDA_OK;
}

/* multifurcating node, phylotreesse.cpp:702-806 */
int lkcuda_partial_multi(lkcuda_engine *e, int32_t dst, int32_t nchild, const int32_t *child, const double *child_len) {
    if (!e || !child || !child_len || nchild < 2 || dst < 0 || dst >= e->cfg.nvec) return LKCUDA_ERR_ARG;
    size_t S = e->S, K = e->K, block = e->block, nptn = e->nptn_all, SU = (size_t)e->cfg.state_unknown;
    for (int32_t c = 0; c < nchild; c++) if (bad_end(e, child[c])) return fail(e, LKCUDA_ERR_ARG, "bad child");
    double *ech = (double *)malloc((size_t)nchild * block * S * 8);
    double *tabs = (double *)malloc((size_t)nchild * (SU + 1) * block * 8);
    double *all = (double *)malloc(block * 8);
    for (int32_t c = 0; c < nchild; c++) {
        build_echild(e, child_len[c], ech + (size_t)c * block * S);
        if (LKCUDA_IS_TIP(child[c])) build_leaf_table(e, ech + (size_t)c * block * S, tabs + (size_t)c * (SU + 1) * block);
    }
    double *dlh = vec_lh(e, dst); int16_t *dsc = vec_sc(e, dst);
    for (size_t ptn = 0; ptn < nptn; ptn++) {
        for (size_t i = 0; i < block; i++) all[i] = 1.0;
        dsc[ptn] = 0;
        for (int32_t c = 0; c < nchild; c++) {
            if (LKCUDA_IS_TIP(child[c])) {
                size_t st = e->states[(size_t)LKCUDA_TIP_TAXON(child[c]) * nptn + ptn];
                const double *t = tabs + (size_t)c * (SU + 1) * block + st * block;
                for (size_t i = 0; i < block; i++) all[i] *= t[i];
            } else {
                const double *lh = vec_lh(e, child[c]) + ptn * block;
                const double *ec = ech + (size_t)c * block * S;
                dsc[ptn] = (int16_t)(dsc[ptn] + vec_sc(e, child[c])[ptn]);
                for (size_t k = 0; k < K; k++)
                    for (size_t x = 0; x < S; x++) {
                        double v = 0.0;
                        for (size_t i = 0; i < S; i++) v += ec[k * S * S + x * S + i] * lh[k * S + i];
                        all[k * S + x] *= v;
                    }
            }
        }
        double lh_max = 0.0, *partial_lh = dlh + ptn * block;
        for (size_t k = 0; k < K; k++) {
            const double *inv_evec = e->inv_evec + (size_t)e->blk_eigen[k] * S * S;
            for (size_t i = 0; i < S; i++) {
                double res = 0.0;
                for (size_t x = 0; x < S; x++) res += all[k * S + x] * inv_evec[i * S + x];
                partial_lh[k * S + i] = res;
                if (fabs(res) > lh_max) lh_max = fabs(res);
            }
        }
        scale_pattern(e, partial_lh, dsc, ptn, lh_max);
    }
    free(ech); free(tabs); free(all);
    e->launches++; e->theta_ready = 0;
    return LKCUDA_OK;
}

/* lh_scale_factor of a vector = sum over its patterns of LOG_SCALING_THRESHOLD * ptn_freq * scale_num
   (closed form of the recursion phylokernel.h:157,389,395,471,477) */
static double vec_scale_factor(lkcuda_engine *e, int32_t v) {
    if (v < 0) return 0.0;
    const int16_t *sc = vec_sc(e, v);
    double s = 0.0;
    for (size_t p = 0; p < (size_t)e->cfg.nptn; p++)
        if (sc[p]) s += LOG_SCALING_THRESHOLD * e->ptn_freq[p] * (double)sc[p];
    return s;
}

int lkcuda_vector_scale_factor(lkcuda_engine *e, int32_t vec, double *out) {
    if (!e || !out || vec < 0 || vec >= e->cfg.nvec) return LKCUDA_ERR_ARG;
    *out = vec_scale_factor(e, vec);
    return LKCUDA_OK;
}

static void build_val(lkcuda_engine *e, double len, double *val0, double *val1, double *val2) {
    /* phylokernel.h:521-532, 767-775; mixture weight phylokernelmixture.h:505,766 */
    for (size_t k = 0; k < e->K; k++) {
        const double *eval = e->eval + (size_t)e->blk_eigen[k] * e->S;
        for (size_t i = 0; i < e->S; i++) {
            double cof = eval[i] * e->blk_rate[k];
            double val = exp(cof * len) * e->blk_weight[k];
            val0[k * e->S + i] = val;
            if (val1) { val1[k * e->S + i] = cof * val; val2[k * e->S + i] = cof * (cof * val); }
        }
    }
}

/* common tail of branch lnL and lnL-from-buffer: log, pattern_lh, NaN repair, +ASC (phylokernel.h:828-866, 1009-1016) */
static void finish_lnl(lkcuda_engine *e, const double *lh_ptn_all, const int16_t *sum_scale_unobs, double scale_sum) {
    size_t P = (size_t)e->cfg.nptn, U = (size_t)e->cfg.nunobs;
    double tree_lh = 0.0, nrep = 0.0;
    for (size_t p = 0; p < P; p++) {
        double l = log(fabs(lh_ptn_all[p]));
        if (isnan(l) || isinf(l)) { l = LOG_SCALING_THRESHOLD * 4; nrep += 1.0; }
        e->pattern_lh[p] = l;
        tree_lh += l * e->ptn_freq[p];
    }
    tree_lh += scale_sum;
    double prob_const = 0.0;
    if (U) {
        for (size_t u = 0; u < U; u++) {
            double l = lh_ptn_all[P + u];                              /* sum part only, ptn_invar not yet added */
            if (sum_scale_unobs[u] >= 1) l *= SCALING_THRESHOLD;      /* bugfix 2016-01-21, phylokernel.h:895-900 */
            prob_const += l + e->ptn_invar[P + u];
        }
        double lc = log(1.0 - prob_const);
        for (size_t p = 0; p < P; p++) e->pattern_lh[p] -= lc;
        tree_lh -= (double)e->cfg.nsite * lc;
    }
    memset(e->result, 0, sizeof(e->result_own));
    e->result[LKCUDA_R_LNL] = tree_lh;
    e->result[LKCUDA_R_PROB_CONST] = prob_const;
    e->result[LKCUDA_R_NREPAIRED] = nrep;
    e->result[LKCUDA_R_SCALE_SUM] = scale_sum;
}

int lkcuda_branch_lnl(lkcuda_engine *e, int32_t dad, int32_t node, double len, double *result) {
    if (!e || node < 0 || node >= e->cfg.nvec || bad_end(e, dad)) return LKCUDA_ERR_ARG;
    if (IS_SM(e)) return sm_branch(e, dad, node, len, result);
    size_t S = e->S, K = e->K, block = e->block, nptn = e->nptn_all, P = (size_t)e->cfg.nptn;
    double *val = (double *)malloc(block * 8);
    double *lh_all = (double *)malloc((nptn + 1) * 8);
    int16_t *ssn = (int16_t *)calloc((size_t)e->cfg.nunobs + 1, 2);
    build_val(e, len, val, NULL, NULL);
    const double *lh_node = vec_lh(e, node);
    const uint8_t *st = LKCUDA_IS_TIP(dad) ? e->states + (size_t)LKCUDA_TIP_TAXON(dad) * nptn : NULL;
    const double *lh_dad = st ? NULL : vec_lh(e, dad);
    for (size_t ptn = 0; ptn < nptn; ptn++) {
        double lh_ptn = 0.0;
        for (size_t k = 0; k < K; k++) {
            double lh_cat = 0.0;
            if (st) {   /* phylotreesse.cpp:1206-1237: lh_node = val * tip_partial_lh[state] */
                const double *tip = e->tip_lh + ((size_t)st[ptn] * e->E + (size_t)e->blk_eigen[k]) * S;
                for (size_t i = 0; i < S; i++) lh_cat += (val[k * S + i] * tip[i]) * lh_node[ptn * block + k * S + i];
            } else {    /* phylotreesse.cpp:1263-1272 */
                for (size_t i = 0; i < S; i++) lh_cat += val[k * S + i] * lh_node[ptn * block + k * S + i] * lh_dad[ptn * block + k * S + i];
            }
            if (ptn < P) e->pattern_lh_cat[ptn * K + k] = lh_cat;
            lh_ptn += lh_cat;
        }
        lh_all[ptn] = (ptn < P) ? lh_ptn + e->ptn_invar[ptn] : lh_ptn;   /* unobserved: invar added after rescaling */
        if (ptn >= P) ssn[ptn - P] = (int16_t)(vec_sc(e, node)[ptn] + (st ? 0 : vec_sc(e, dad)[ptn]));
    }
    double scale_sum = vec_scale_factor(e, node) + (st ? 0.0 : vec_scale_factor(e, dad));
    finish_lnl(e, lh_all, ssn, scale_sum);
    free(val); free(lh_all); free(ssn);
    e->launches++;
    if (result) memcpy(result, e->result, sizeof(e->result_own));
    return LKCUDA_OK;
}

/* the fused call of the CUDA engine, restated as its definition: the two calls back to back (phylotree.cpp:1031-1071) */
int lkcuda_partial_batch_lnl(lkcuda_engine *e, const lkcuda_partial_op *ops, int32_t nops, const int32_t *level_start, int32_t nlevels,
                             int32_t dad, int32_t node, double len, double *result) {
    int rc = lkcuda_partial_batch(e, ops, nops, level_start, nlevels);
    return rc ? rc : lkcuda_branch_lnl(e, dad, node, len, result);
}

/* _pattern_lh_cat only (phylotreesse.cpp:1194-1237) */
int lkcuda_pattern_lh_cat(lkcuda_engine *e, int32_t dad, int32_t node, double len) {
    if (!e || node < 0 || node >= e->cfg.nvec || bad_end(e, dad)) return LKCUDA_ERR_ARG;
    if (IS_SM(e)) return sm_branch(e, dad, node, len, NULL);    /* the site-model branch kernel is what fills it (phylokernelsitemodel.cpp:579-603) */
    size_t S = e->S, K = e->K, block = e->block, nptn = e->nptn_all, P = (size_t)e->cfg.nptn;
    double *val = (double *)malloc(block * 8);
    build_val(e, len, val, NULL, NULL);
    const double *lh_node = vec_lh(e, node);
    const uint8_t *st = LKCUDA_IS_TIP(dad) ? e->states + (size_t)LKCUDA_TIP_TAXON(dad) * nptn : NULL;
    const double *lh_dad = st ? NULL : vec_lh(e, dad);
    for (size_t ptn = 0; ptn < P; ptn++)
        for (size_t k = 0; k < K; k++) {
            double lh_cat = 0.0;
            if (st) {
                const double *tip = e->tip_lh + ((size_t)st[ptn] * e->E + (size_t)e->blk_eigen[k]) * S;
                for (size_t i = 0; i < S; i++) lh_cat += (val[k * S + i] * tip[i]) * lh_node[ptn * block + k * S + i];
            } else {
                for (size_t i = 0; i < S; i++) lh_cat += val[k * S + i] * lh_node[ptn * block + k * S + i] * lh_dad[ptn * block + k * S + i];
            }
            e->pattern_lh_cat[ptn * K + k] = lh_cat;
        }
    free(val);
    e->launches++;
    return LKCUDA_OK;
}

int lkcuda_theta_build(lkcuda_engine *e, int32_t dad, int32_t node);
int lkcuda_derv(lkcuda_engine *e, double len, double *result);
int lkcuda_theta_derv(lkcuda_engine *e, int32_t dad, int32_t node, double len, double *result) {
    int rc = lkcuda_theta_build(e, dad, node);
    return rc ? rc : lkcuda_derv(e, len, result);
}

int lkcuda_theta_build(lkcuda_engine *e, int32_t dad, int32_t node) {
    if (!e || node < 0 || node >= e->cfg.nvec || bad_end(e, dad)) return LKCUDA_ERR_ARG;
    if (IS_SM(e)) return sm_theta(e, dad, node);
    size_t S = e->S, K = e->K, block = e->block, nptn = e->nptn_all, P = (size_t)e->cfg.nptn;
    const double *lh_node = vec_lh(e, node);
    if (LKCUDA_IS_TIP(dad)) {   /* phylotreesse.cpp:1046-1063 */
        const uint8_t *st = e->states + (size_t)LKCUDA_TIP_TAXON(dad) * nptn;
        for (size_t ptn = 0; ptn < nptn; ptn++)
            for (size_t k = 0; k < K; k++) {
                const double *tip = e->tip_lh + ((size_t)st[ptn] * e->E + (size_t)e->blk_eigen[k]) * S;
                for (size_t i = 0; i < S; i++) e->theta[ptn * block + k * S + i] = tip[i] * lh_node[ptn * block + k * S + i];
            }
    } else {                    /* phylotreesse.cpp:1072-1079 */
        const double *lh_dad = vec_lh(e, dad);
        for (size_t i = 0; i < nptn * block; i++) e->theta[i] = lh_node[i] * lh_dad[i];
    }
    e->theta_scale_sum = vec_scale_factor(e, node) + (LKCUDA_IS_TIP(dad) ? 0.0 : vec_scale_factor(e, dad));
    for (size_t u = 0; u < (size_t)e->cfg.nunobs; u++)
        e->theta_scale_num[u] = (int16_t)(vec_sc(e, node)[P + u] + (LKCUDA_IS_TIP(dad) ? 0 : vec_sc(e, dad)[P + u]));
    e->theta_ready = 1;
    e->launches++;
    return LKCUDA_OK;
}

static int derv_at(lkcuda_engine *e, double len, double *df_out, double *ddf_out) {
    if (IS_SM(e)) return sm_derv(e, len, df_out, ddf_out);
    size_t block = e->block, nptn = e->nptn_all, P = (size_t)e->cfg.nptn;
    double *val0 = (double *)malloc(block * 8), *val1 = (double *)malloc(block * 8), *val2 = (double *)malloc(block * 8);
    build_val(e, len, val0, val1, val2);
    double my_df = 0.0, my_ddf = 0.0, prob_const = 0.0, df_const = 0.0, ddf_const = 0.0;
    for (size_t ptn = 0; ptn < nptn; ptn++) {   /* phylotreesse.cpp:1106-1132 */
        double lh_ptn = e->ptn_invar[ptn], df_ptn = 0.0, ddf_ptn = 0.0;
        const double *theta = e->theta + ptn * block;
        for (size_t i = 0; i < block; i++) {
            lh_ptn += val0[i] * theta[i];
            df_ptn += val1[i] * theta[i];
            ddf_ptn += val2[i] * theta[i];
        }
        lh_ptn = fabs(lh_ptn);
        if (ptn < P) {
            double df_frac = df_ptn / lh_ptn, ddf_frac = ddf_ptn / lh_ptn, freq = e->ptn_freq[ptn];
            double tmp1 = df_frac * freq, tmp2 = ddf_frac * freq;
            my_df += tmp1;
            my_ddf += tmp2 - tmp1 * df_frac;
        } else {
            prob_const += lh_ptn; df_const += df_ptn; ddf_const += ddf_ptn;
        }
    }
    if (isnan(my_df) || isinf(my_df)) { my_df = 0.0; my_ddf = 0.0; }   /* phylokernel.h:647-651 */
    if (P < nptn) {                                                    /* phylokernel.h:719-724 */
        prob_const = 1.0 - prob_const;
        double df_frac = df_const / prob_const, ddf_frac = ddf_const / prob_const;
        my_df += (double)e->cfg.nsite * df_frac;
        my_ddf += (double)e->cfg.nsite * (ddf_frac + df_frac * df_frac);
    }
    free(val0); free(val1); free(val2);
    *df_out = my_df; *ddf_out = my_ddf;
    e->launches++;
    return LKCUDA_OK;
}

int lkcuda_derv(lkcuda_engine *e, double len, double *result) {
    if (!e) return LKCUDA_ERR_ARG;
    if (!e->theta_ready) return fail(e, LKCUDA_ERR_STATE, "lkcuda_derv: theta_all not built");
    double df, ddf;
    derv_at(e, len, &df, &ddf);
    memset(e->result, 0, sizeof(e->result_own));
    e->result[LKCUDA_R_DF] = df; e->result[LKCUDA_R_DDF] = ddf;
    if (result) memcpy(result, e->result, sizeof(e->result_own));
    return LKCUDA_OK;
}

/* scalar statement of lkcuda_derv_noalloc (no counterpart in the reference; the loop of phylotreesse.cpp:1106-1132 on the caller's
   arrays, which are HOST arrays here); same limits and refusals as the CUDA entry point */
int lkcuda_derv_noalloc(int32_t device, void *cuda_stream, const double *theta_all, const double *ptn_freq, const double *ptn_invar,
                        int64_t nptn, int32_t nstates, int32_t ncat, const double *eval, const double *cat_rate, const double *cat_prop,
                        double len, double *df_ddf) {
    (void)device; (void)cuda_stream;
    if (!theta_all || !ptn_freq || !ptn_invar || !eval || !cat_rate || !cat_prop || !df_ddf) return fail(NULL, LKCUDA_ERR_ARG, "lkcuda_derv_noalloc: null argument");
    if (nptn < 0 || nstates <= 0 || ncat <= 0) return fail(NULL, LKCUDA_ERR_ARG, "lkcuda_derv_noalloc: invalid shape");
    if (nptn > LKCUDA_NOALLOC_MAX_PATTERNS) return fail(NULL, LKCUDA_ERR_ARG, "lkcuda_derv_noalloc: patterns exceed LKCUDA_NOALLOC_MAX_PATTERNS");
    if (nstates > 64 || ncat > 64 || nstates * ncat > LKCUDA_NOALLOC_MAX_BLOCK) return fail(NULL, LKCUDA_ERR_ARG, "lkcuda_derv_noalloc: states x categories exceed the limits");
    double val0[LKCUDA_NOALLOC_MAX_BLOCK], val1[LKCUDA_NOALLOC_MAX_BLOCK], val2[LKCUDA_NOALLOC_MAX_BLOCK];
    size_t block = (size_t)nstates * ncat;
    for (int c = 0; c < ncat; c++)
        for (int i = 0; i < nstates; i++) {
            double cof = eval[i] * cat_rate[c], x = exp(cof * len) * cat_prop[c];
            val0[c * nstates + i] = x; val1[c * nstates + i] = cof * x; val2[c * nstates + i] = cof * cof * x;
        }
    double my_df = 0.0, my_ddf = 0.0;
    for (int64_t ptn = 0; ptn < nptn; ptn++) {
        double lh_ptn = ptn_invar[ptn], df_ptn = 0.0, ddf_ptn = 0.0;
        const double *theta = theta_all + (size_t)ptn * block;
        for (size_t i = 0; i < block; i++) { lh_ptn += val0[i] * theta[i]; df_ptn += val1[i] * theta[i]; ddf_ptn += val2[i] * theta[i]; }
        lh_ptn = fabs(lh_ptn);
        double df_frac = df_ptn / lh_ptn, ddf_frac = ddf_ptn / lh_ptn, tmp1 = df_frac * ptn_freq[ptn], tmp2 = ddf_frac * ptn_freq[ptn];
        my_df += tmp1;
        my_ddf += tmp2 - tmp1 * df_frac;
    }
    if (isnan(my_df) || isinf(my_df)) { my_df = 0.0; my_ddf = 0.0; }
    df_ddf[0] = my_df; df_ddf[1] = my_ddf;
    return LKCUDA_OK;
}

int lkcuda_lnl_from_buffer(lkcuda_engine *e, double len, double *result) {
    if (!e) return LKCUDA_ERR_ARG;
    if (!e->theta_ready) return fail(e, LKCUDA_ERR_STATE, "lkcuda_lnl_from_buffer: theta_all not built");
    if (IS_SM(e)) return sm_frombuf(e, len, result);
    size_t block = e->block, nptn = e->nptn_all;
    double *val0 = (double *)malloc(block * 8);
    double *lh_all = (double *)malloc((nptn + 1) * 8);
    build_val(e, len, val0, NULL, NULL);
    for (size_t ptn = 0; ptn < nptn; ptn++) {   /* phylokernel.h:1067-1090 */
        double lh = 0.0;
        const double *theta = e->theta + ptn * block;
        for (size_t i = 0; i < block; i++) lh += val0[i] * theta[i];
        lh_all[ptn] = (ptn < (size_t)e->cfg.nptn) ? lh + e->ptn_invar[ptn] : lh;
    }
    finish_lnl(e, lh_all, e->theta_scale_num, e->theta_scale_sum);
    free(val0); free(lh_all);
    e->launches++;
    if (result) memcpy(result, e->result, sizeof(e->result_own));
    return LKCUDA_OK;
}

/* Optimization::minimizeNewton (optimization.cpp:388-465) with computeFuncDerv = -(df, ddf) (phylotree.cpp:2135-2146) */
int lkcuda_newton_branch(lkcuda_engine *e, double x1, double xguess, double x2, double xacc, int32_t maxNRStep, double *out) {
    if (!e || !out) return LKCUDA_ERR_ARG;
    if (!e->theta_ready) return fail(e, LKCUDA_ERR_STATE, "lkcuda_newton_branch: theta_all not built");
    int j, nev = 0;
    double df, dx, dxold, f, temp, xh, xl, rts, rts_old, d2l;
    (void)dxold;
#define FUNC_DERV(x) do { double a_, b_; derv_at(e, (x), &a_, &b_); f = -a_; df = -b_; nev++; out[3] = (x); } while (0)
#define RETURN(x) do { out[0] = (x); out[1] = d2l; out[2] = (double)nev; return LKCUDA_OK; } while (0)
    rts = xguess;
    if (rts < x1) rts = x1;
    if (rts > x2) rts = x2;
    FUNC_DERV(rts);
    d2l = df;
    if (!isfinite(f) || !isfinite(df)) return fail(e, LKCUDA_ERR_STATE, "Wrong computeFuncDerv");
    if (df >= 0.0 && fabs(f) < xacc) RETURN(rts);
    if (f < 0.0) { xl = rts; xh = x2; } else { xh = rts; xl = x1; }
    dx = dxold = fabs(xh - xl);
    for (j = 1; j <= maxNRStep; j++) {
        rts_old = rts;
        if ((df <= 0.0) || (((rts - xh) * df - f) * ((rts - xl) * df - f) >= 0.0)) {
            dxold = dx; dx = 0.5 * (xh - xl); rts = xl + dx; d2l = df;
            if (xl == rts) RETURN(rts);
        } else {
            dxold = dx; dx = f / df; temp = rts; rts -= dx; d2l = df;
            if (temp == rts) RETURN(rts);
        }
        if (fabs(dx) < xacc || (j == maxNRStep)) RETURN(rts_old);
        FUNC_DERV(rts);
        if (!isfinite(f) || !isfinite(df)) return fail(e, LKCUDA_ERR_STATE, "Wrong computeFuncDerv");
        if (df > 0.0 && fabs(f) < xacc) { d2l = df; RETURN(rts); }
        if (f < 0.0) xl = rts; else xh = rts;
    }
    return fail(e, LKCUDA_ERR_STATE, "Maximum number of iterations exceeded in minimizeNewton");
#undef FUNC_DERV
#undef RETURN
}

int lkcuda_theta_newton(lkcuda_engine *e, int32_t dad, int32_t node, double x1, double xguess, double x2, double xacc, int32_t maxNRStep, double *out) {
    int rc = lkcuda_theta_build(e, dad, node);
    return rc ? rc : lkcuda_newton_branch(e, x1, xguess, x2, xacc, maxNRStep, out);
}

/* ---- device length slots, executed at once ---- */
int lkcuda_theta_newton_async(lkcuda_engine *e, int32_t dad, int32_t node, double x1, double xguess, int32_t xguess_slot, double x2,
                              double xacc, int32_t maxNRStep, int32_t len_slot) {
    if (!e) return LKCUDA_ERR_ARG;
    if (len_slot < 0 || len_slot >= LKCUDA_NLENSLOT || xguess_slot >= LKCUDA_NLENSLOT) return fail(e, LKCUDA_ERR_ARG, "lkcuda_theta_newton_async: bad length slot");
    double out[4] = {0.0, 0.0, 0.0, 0.0};
    int rc = lkcuda_theta_newton(e, dad, node, x1, xguess_slot >= 0 ? e->lens[xguess_slot] : xguess, x2, xacc, maxNRStep, out);
    double *info = e->lens + LKCUDA_NLENSLOT + 4 * len_slot;
    info[0] = out[0]; info[1] = out[1]; info[2] = out[2]; info[3] = rc ? 2.0 : 1.0;
    e->lens[len_slot] = out[0];
    return LKCUDA_OK;           /* a failed run is reported through the slot's status, as on the device */
}
int lkcuda_lnl_from_buffer_slot(lkcuda_engine *e, int32_t len_slot, double *result) {
    if (!e) return LKCUDA_ERR_ARG;
    if (len_slot < 0 || len_slot >= LKCUDA_NLENSLOT) return fail(e, LKCUDA_ERR_ARG, "lkcuda_lnl_from_buffer_slot: bad length slot");
    return lkcuda_lnl_from_buffer(e, e->lens[len_slot], result);
}
int lkcuda_len_slots_set(lkcuda_engine *e, int32_t first, int32_t n, const double *values) {
    if (!e || first < 0 || n < 0 || first + n > LKCUDA_NLENSLOT || (n && !values)) return LKCUDA_ERR_ARG;
    memcpy(e->lens + first, values, sizeof(double) * (size_t)n);
    return LKCUDA_OK;
}
int lkcuda_len_slots_fetch(lkcuda_engine *e, double *lens, double *info) {
    if (!e) return LKCUDA_ERR_ARG;
    if (lens) memcpy(lens, e->lens, sizeof(double) * LKCUDA_NLENSLOT);
    if (info) memcpy(info, e->lens + LKCUDA_NLENSLOT, sizeof(double) * 4 * LKCUDA_NLENSLOT);
    return LKCUDA_OK;
}

int lkcuda_read_result(lkcuda_engine *e, double *result) {
    if (!e || !result) return LKCUDA_ERR_ARG;
    memcpy(result, e->result, sizeof(e->result_own));
    return LKCUDA_OK;
}
int lkcuda_wait_result(lkcuda_engine *e, double *result) { return lkcuda_read_result(e, result); }
/* the oracle has no devices and no peers: shards are summed by the host (lk_phylotree.cpp reduceResult) */
int lkcuda_peer_handle(lkcuda_engine *e, uint8_t *handle, void **local_ptr) { (void)handle; (void)local_ptr; return e ? fail(e, LKCUDA_ERR_STATE, "the CPU oracle has no peer mailboxes") : LKCUDA_ERR_ARG; }
int lkcuda_attach_peers(lkcuda_engine *e, int32_t rank, int32_t world, const uint8_t *handles, void *const *raw_ptrs) {
    (void)rank; (void)world; (void)handles; (void)raw_ptrs;
    return e ? fail(e, LKCUDA_ERR_STATE, "the CPU oracle has no peer mailboxes") : LKCUDA_ERR_ARG;
}
int lkcuda_detach_peers(lkcuda_engine *e) { return e ? LKCUDA_OK : LKCUDA_ERR_ARG; }
int lkcuda_result_device_ptr(lkcuda_engine *e, void **p) { if (!e || !p) return LKCUDA_ERR_ARG; *p = e->result; return LKCUDA_OK; }
int lkcuda_set_result_buffer(lkcuda_engine *e, void *ptr) { if (!e) return LKCUDA_ERR_ARG; e->result = ptr ? (double*)ptr : e->result_own; return LKCUDA_OK; }

int lkcuda_fetch_pattern_lh(lkcuda_engine *e, double *dst) {
    if (!e || !dst) return LKCUDA_ERR_ARG;
    memcpy(dst, e->pattern_lh, (size_t)e->cfg.nptn * 8);
    return LKCUDA_OK;
}
int lkcuda_fetch_pattern_lh_corrected(lkcuda_engine *e, int32_t vec, double *dst);
int lkcuda_fetch_pattern_lh_corrected_async(lkcuda_engine *e, int32_t vec, double *dst) { return vec < 0 ? LKCUDA_ERR_ARG : lkcuda_fetch_pattern_lh_corrected(e, vec, dst); }
int lkcuda_fetch_wait(lkcuda_engine *e) { return e ? LKCUDA_OK : LKCUDA_ERR_ARG; }
int lkcuda_fetch_pattern_lh_corrected(lkcuda_engine *e, int32_t vec, double *dst) {   /* phylotree.cpp:1055-1069 */
    int rc = lkcuda_fetch_pattern_lh(e, dst);
    if (rc || vec < 0) return rc;
    if (vec >= e->cfg.nvec) return LKCUDA_ERR_ARG;
    {
        const int16_t *sc = e->scale + (size_t)vec * e->sc_stride;
        int64_t p;
        for (p = 0; p < e->cfg.nptn; p++) if (sc[p] > 0) dst[p] += (double)sc[p] * LOG_SCALING_THRESHOLD;
    }
    return LKCUDA_OK;
}

int lkcuda_fetch_pattern_lh_cat(lkcuda_engine *e, double *dst) {
    if (!e || !dst) return LKCUDA_ERR_ARG;
    memcpy(dst, e->pattern_lh_cat, (size_t)e->cfg.nptn * e->K * 8);
    return LKCUDA_OK;
}
/* ---- UFBoot RELL.  IQTree::saveCurrentTree (iqtree.cpp:2676-2786): pattern_lh[i] = (float)pattern_lh_orig[i] with
   pattern_lh_orig from computePatternLikelihood (phylotree.cpp:1213-1259); rell = dotProduct(pattern_lh, boot_sample, nptn)
   (iqtree.cpp:2733).  dotProductSIMD<float, Vec8f, 8> (phylokernel.h:55-61; AVX binding phylotreeavx.cpp:26) keeps 8 float lanes, lane j
   accumulating entries j, j+8, j+16, ... with mul_add (a separate multiply and add without FMA hardware flags, vectorclass), and
   ends with horizontal_add.  Restated lane by lane in float. ---- */
int lkcuda_set_boot_samples(lkcuda_engine *e, const float *samples, int32_t nsamples, int64_t stride) {
    if (!e || nsamples < 0 || (nsamples && !samples) || stride < e->cfg.nptn) return LKCUDA_ERR_ARG;
    size_t P = (size_t)e->cfg.nptn;
    free(e->boot); free(e->rell);
    e->boot = (float *)calloc((size_t)nsamples * P + 1, sizeof(float));
    e->rell = (double *)calloc((size_t)nsamples + 1, sizeof(double));
    e->nboot = nsamples;
    for (int32_t s = 0; s < nsamples; s++) memcpy(e->boot + (size_t)s * P, samples + (size_t)s * (size_t)stride, P * sizeof(float));
    return LKCUDA_OK;
}
int lkcuda_boot_storage(lkcuda_engine *e, int32_t *bytes_per_entry, uint64_t *bytes) {
    if (!e || !e->boot) return LKCUDA_ERR_STATE;
    if (bytes_per_entry) *bytes_per_entry = 4;
    if (bytes) *bytes = (uint64_t)e->nboot * (uint64_t)e->cfg.nptn * 4;
    return LKCUDA_OK;
}
int lkcuda_rell(lkcuda_engine *e, int32_t vec_a, int32_t vec_b, double *rell) {
    if (!e || vec_a >= e->cfg.nvec || vec_b >= e->cfg.nvec) return LKCUDA_ERR_ARG;
    if (!e->boot) return fail(e, LKCUDA_ERR_STATE, "lkcuda_rell: no bootstrap samples");
    size_t P = (size_t)e->cfg.nptn, Ppad = (P + 7) / 8 * 8;
    float *x = (float *)calloc(Ppad + 8, sizeof(float)), *y = (float *)calloc(Ppad + 8, sizeof(float));
    const int16_t *sa = vec_a >= 0 ? vec_sc(e, vec_a) : NULL, *sb = vec_b >= 0 ? vec_sc(e, vec_b) : NULL;
    for (size_t p = 0; p < P; p++) {       /* phylotree.cpp:1216-1233 */
        int sc = (sa && sa[p] > 0 ? sa[p] : 0) + (sb && sb[p] > 0 ? sb[p] : 0);
        x[p] = (float)(sc ? e->pattern_lh[p] + sc * LOG_SCALING_THRESHOLD : e->pattern_lh[p]);
    }
    for (int32_t s = 0; s < e->nboot; s++) {
        memcpy(y, e->boot + (size_t)s * P, P * sizeof(float));
        volatile float lane[8];
        for (int j = 0; j < 8; j++) lane[j] = x[j] * y[j];
        for (size_t i = 8; i < Ppad; i += 8)
            for (int j = 0; j < 8; j++) { volatile float pr = x[i + j] * y[i + j]; lane[j] = pr + lane[j]; }
        /* horizontal_add(Vec8f) (vectorclass/vectorf256.h): hadd within 128-bit halves twice, then low + high */
        volatile float h0 = lane[0] + lane[1], h1 = lane[2] + lane[3], h2 = lane[4] + lane[5], h3 = lane[6] + lane[7];
        volatile float g0 = h0 + h1, g1 = h2 + h3;
        volatile float r = g0 + g1;
        e->rell[s] = (double)r;
    }
    free(x); free(y);
    if (rell) memcpy(rell, e->rell, (size_t)e->nboot * 8);
    return LKCUDA_OK;
}
int lkcuda_rell_fetch(lkcuda_engine *e, double *rell) { if (!e || !rell || !e->rell) return LKCUDA_ERR_STATE; memcpy(rell, e->rell, (size_t)e->nboot * 8); return LKCUDA_OK; }
int lkcuda_rell_device_ptr(lkcuda_engine *e, void **p) { if (!e || !p || !e->rell) return LKCUDA_ERR_STATE; *p = e->rell; return LKCUDA_OK; }

/* ---- consumers of _pattern_lh_cat ---- */
int lkcuda_lh_cat_collapse(lkcuda_engine *e, int32_t mode, int32_t ncat, int32_t nmix, int32_t *ncol_out) {
    if (!e || mode < 0 || mode > 2 || ncat < 1 || nmix < 1 || (size_t)ncat * (size_t)nmix != e->K) return LKCUDA_ERR_ARG;
    size_t P = (size_t)e->cfg.nptn;
    if (!e->w) e->w = (double *)calloc(P * e->K + 1, 8);
    e->ncol = mode == 1 ? nmix : (mode == 2 ? ncat : ncat * nmix);
    if (ncol_out) *ncol_out = e->ncol;
    const double *lh_cat = e->pattern_lh_cat;
    double *lh_res = e->w;
    for (size_t ptn = 0; ptn < P; ptn++) {          /* phylotree.cpp:1133-1158 */
        if (mode == 1) {
            for (int m = 0; m < nmix; m++) {
                double lh = lh_cat[0];
                for (int c = 1; c < ncat; c++) lh += lh_cat[c];
                lh_res[m] = lh;
                lh_cat += ncat;
            }
            lh_res += nmix;
        } else if (mode == 2) {
            memcpy(lh_res, lh_cat, (size_t)ncat * 8);
            lh_cat += ncat;
            for (int m = 1; m < nmix; m++) {
                for (int c = 0; c < ncat; c++) lh_res[c] += lh_cat[c];
                lh_cat += ncat;
            }
            lh_res += ncat;
        } else {
            memcpy(lh_res, lh_cat, e->K * 8);
            lh_cat += e->K; lh_res += e->K;
        }
    }
    return LKCUDA_OK;
}
int lkcuda_lh_cat_em_ex(lkcuda_engine *e, const double *ratio, int32_t include_invar, double *sums);
int lkcuda_lh_cat_em(lkcuda_engine *e, const double *ratio, double *sums) { return lkcuda_lh_cat_em_ex(e, ratio, 1, sums); }
int lkcuda_lh_cat_em_ex(lkcuda_engine *e, const double *ratio, int32_t include_invar, double *sums) {
    if (!e || !sums) return LKCUDA_ERR_ARG;
    if (!e->w || e->ncol < 1) return fail(e, LKCUDA_ERR_STATE, "lkcuda_lh_cat_em: call lkcuda_lh_cat_collapse first");
    size_t P = (size_t)e->cfg.nptn, n = (size_t)e->ncol;
    if (ratio)                                   /* modelmixture.cpp:1369-1377 */
        for (size_t ptn = 0; ptn < P; ptn++)
            for (size_t c = 0; c < n; c++) e->w[ptn * n + c] *= ratio[c];
    memset(sums, 0, n * 8);
    for (size_t ptn = 0; ptn < P; ptn++) {       /* modelmixture.cpp:1379-1388 */
        const double *this_lk_cat = e->w + ptn * n;
        double lk_ptn = include_invar ? e->ptn_invar[ptn] : 0.0;     /* ratefree.cpp:489 starts from 0 */
        for (size_t c = 0; c < n; c++) lk_ptn += this_lk_cat[c];
        lk_ptn = e->ptn_freq[ptn] / lk_ptn;
        for (size_t c = 0; c < n; c++) sums[c] += this_lk_cat[c] * lk_ptn;
    }
    return LKCUDA_OK;
}
/* ratefree.cpp:487-503 (E-step) + :547-549 (the posterior of one category becomes the sub-tree's ptn_freq) */
int lkcuda_lh_cat_posterior_weights(lkcuda_engine *e, int32_t col, int32_t include_invar, lkcuda_engine *dst) {
    if (!e || !dst || dst == e) return LKCUDA_ERR_ARG;
    if (!e->w || e->ncol < 1) return fail(e, LKCUDA_ERR_STATE, "lkcuda_lh_cat_posterior_weights: call lkcuda_lh_cat_collapse first");
    if (col < 0 || col >= e->ncol || dst->cfg.nptn != e->cfg.nptn) return fail(e, LKCUDA_ERR_ARG, "lkcuda_lh_cat_posterior_weights: bad argument");
    size_t P = (size_t)e->cfg.nptn, n = (size_t)e->ncol;
    for (size_t ptn = 0; ptn < P; ptn++) {
        const double *this_lk_cat = e->w + ptn * n;
        double lk_ptn = include_invar ? e->ptn_invar[ptn] : 0.0;       /* modelmixture.cpp:1460 starts with ptn_invar, ratefree.cpp:489 with 0 */
        for (size_t c = 0; c < n; c++) lk_ptn += this_lk_cat[c];
        lk_ptn = e->ptn_freq[ptn] / lk_ptn;
        dst->ptn_freq[ptn] = this_lk_cat[col] * lk_ptn;
    }
    dst->theta_ready = 0;
    return LKCUDA_OK;
}
int lkcuda_lh_cat_pattern_rates(lkcuda_engine *e, const double *rates, int32_t invar_class, double *ptn_rates, int32_t *ptn_cat) {
    if (!e || !rates || !ptn_rates || !ptn_cat) return LKCUDA_ERR_ARG;
    if (!e->w || e->ncol < 1) return fail(e, LKCUDA_ERR_STATE, "lkcuda_lh_cat_pattern_rates: call lkcuda_lh_cat_collapse first");
    size_t P = (size_t)e->cfg.nptn, n = (size_t)e->ncol;
    for (size_t i = 0; i < P; i++) {             /* rategamma.cpp:244-257 / rategammainvar.cpp:202-219, ties: first maximum */
        const double *lh_cat = e->w + i * n;
        double sum_rate = 0.0, sum_lh = invar_class ? e->ptn_invar[i] : 0.0;
        double best_lh = invar_class ? e->ptn_invar[i] : lh_cat[0];
        size_t best = 0;
        for (size_t c = 0; c < n; c++) {
            sum_rate += rates[c] * lh_cat[c];
            sum_lh += lh_cat[c];
            if (lh_cat[c] > best_lh) { best = c + (invar_class ? 1 : 0); best_lh = lh_cat[c]; }
        }
        ptn_rates[i] = sum_rate / sum_lh;
        ptn_cat[i] = (int32_t)best;
    }
    return LKCUDA_OK;
}
int lkcuda_fetch_lh_cat_collapsed(lkcuda_engine *e, double *dst) {
    if (!e || !dst) return LKCUDA_ERR_ARG;
    if (!e->w || e->ncol < 1) return fail(e, LKCUDA_ERR_STATE, "lkcuda_fetch_lh_cat_collapsed: call lkcuda_lh_cat_collapse first");
    memcpy(dst, e->w, (size_t)e->cfg.nptn * (size_t)e->ncol * 8);
    return LKCUDA_OK;
}

int lkcuda_fetch_vector(lkcuda_engine *e, int32_t vec, double *partial_lh, int16_t *scale_num) {
    if (!e || vec < 0 || vec >= e->cfg.nvec) return LKCUDA_ERR_ARG;
    if (partial_lh) memcpy(partial_lh, vec_lh(e, vec), e->nptn_all * e->block * 8);
    if (scale_num) memcpy(scale_num, vec_sc(e, vec), e->nptn_all * 2);
    return LKCUDA_OK;
}
int lkcuda_store_vector(lkcuda_engine *e, int32_t vec, const double *partial_lh, const int16_t *scale_num) {
    if (!e || vec < 0 || vec >= e->cfg.nvec) return LKCUDA_ERR_ARG;
    if (partial_lh) memcpy(vec_lh(e, vec), partial_lh, e->nptn_all * e->block * 8);
    if (scale_num) memcpy(vec_sc(e, vec), scale_num, e->nptn_all * 2);
    return LKCUDA_OK;
}
int lkcuda_fetch_theta(lkcuda_engine *e, double *theta) {
    if (!e || !theta) return LKCUDA_ERR_ARG;
    memcpy(theta, e->theta, e->nptn_all * e->block * 8);
    return LKCUDA_OK;
}
int lkcuda_vector_device_ptr(lkcuda_engine *e, int32_t vec, void **partial_lh, void **scale_num) {
    if (!e || vec < 0 || vec >= e->cfg.nvec) return LKCUDA_ERR_ARG;
    if (partial_lh) *partial_lh = vec_lh(e, vec);
    if (scale_num) *scale_num = vec_sc(e, vec);
    return LKCUDA_OK;
}
int lkcuda_launch_count(const lkcuda_engine *e, uint64_t *c) { if (!e || !c) return LKCUDA_ERR_ARG; *c = e->launches; return LKCUDA_OK; }
int lkcuda_algorithmic_bytes(const lkcuda_engine *e, uint64_t *b) { if (!e || !b) return LKCUDA_ERR_ARG; *b = e->bytes; return LKCUDA_OK; }
int lkcuda_reset_counters(lkcuda_engine *e) { if (!e) return LKCUDA_ERR_ARG; e->launches = 0; e->bytes = 0; return LKCUDA_OK; }
int lkcuda_set_profiling(lkcuda_engine *e, int32_t on) { (void)on; return e ? LKCUDA_OK : LKCUDA_ERR_ARG; }
int lkcuda_kernel_flops(lkcuda_engine *e, int32_t kind, uint64_t *flops) {
    if (!e || !flops) return LKCUDA_ERR_ARG;
    (void)kind; *flops = 0;      /* the CPU restatement keeps no flop statistics */
    return LKCUDA_OK;
}

int lkcuda_kernel_stats(lkcuda_engine *e, int32_t kind, uint64_t *launches, uint64_t *bytes, double *device_ms) {
    if (!e || kind < 0 || kind >= LKCUDA_NKIND) return LKCUDA_ERR_ARG;
    if (launches) *launches = 0; if (bytes) *bytes = 0; if (device_ms) *device_ms = 0.0;
    return LKCUDA_OK;
}


/* ================================================================================================================================
 * site-specific frequency models, restated from phylokernelsitemodel.cpp (the generic, non-SIMD versions):
 * computeSitemodelPartialLikelihoodEigen :20-436, ...DervEigen :438-546, ...BranchEigen :548-660, ...FromBufferEigen :663-703.
 * eval / evec / inv_evec of pattern ptn = those of model ptn_model[ptn] (models->at(ptn)).  No +ASC (the reference never combines them).
 * ================================================================================================================================ */
int lkcuda_set_site_models(lkcuda_engine *e, const int32_t *ptn_model, const uint64_t *state_mask) {
    if (!e || !ptn_model || !state_mask) return LKCUDA_ERR_ARG;
    if (!IS_SM(e)) return fail(e, LKCUDA_ERR_STATE, "lkcuda_set_site_models: engine created without LKCUDA_FLAG_SITEMODEL");
    for (size_t p = 0; p < e->nptn_all; p++)
        if (ptn_model[p] < 0 || (size_t)ptn_model[p] >= e->E) return fail(e, LKCUDA_ERR_ARG, "ptn_model out of range");
    free(e->ptn_model); free(e->state_mask);
    e->ptn_model = (int32_t *)malloc((e->nptn_all + 1) * 4);
    e->state_mask = (uint64_t *)malloc((size_t)(e->cfg.state_unknown + 1) * 8);
    memcpy(e->ptn_model, ptn_model, e->nptn_all * 4);
    memcpy(e->state_mask, state_mask, (size_t)(e->cfg.state_unknown + 1) * 8);
    e->theta_ready = 0;
    return LKCUDA_OK;
}

/* tip_partial_lh[taxon][ptn][i] (phylotreesse.cpp:373-456): sum over the states the code stands for of inv_evec[i][x], x ascending */
static void sm_tip(lkcuda_engine *e, size_t ptn, int state, double *out) {
    size_t S = e->S;
    const double *inv_evec = e->inv_evec + (size_t)e->ptn_model[ptn] * S * S;
    uint64_t mask = e->state_mask[state];
    if (state < (int)S) { for (size_t i = 0; i < S; i++) out[i] = inv_evec[i * S + (size_t)state]; return; }
    for (size_t i = 0; i < S; i++) {
        double lh = 0.0;
        for (size_t x = 0; x < S; x++) if (mask & (1ull << x)) lh += inv_evec[i * S + x];
        out[i] = lh;
    }
}

static int sm_partial_one(lkcuda_engine *e, const lkcuda_partial_op *op) {
    size_t S = e->S, K = e->K, block = e->block, nptn = e->nptn_all;
    if (!e->ptn_model) return fail(e, LKCUDA_ERR_STATE, "site models not set");
    if (op->dst < 0 || op->dst >= e->cfg.nvec || bad_end(e, op->left) || bad_end(e, op->right))
        return fail(e, LKCUDA_ERR_ARG, "partial op: bad vector / taxon id");
    if (!LKCUDA_IS_TIP(op->left) && LKCUDA_IS_TIP(op->right))
        return fail(e, LKCUDA_ERR_ARG, "partial op: a leaf child must be 'left'");
    double *dst = vec_lh(e, op->dst);
    int16_t *dsc = vec_sc(e, op->dst);
    const uint8_t *st_l = LKCUDA_IS_TIP(op->left) ? e->states + (size_t)LKCUDA_TIP_TAXON(op->left) * nptn : NULL;
    const uint8_t *st_r = LKCUDA_IS_TIP(op->right) ? e->states + (size_t)LKCUDA_TIP_TAXON(op->right) * nptn : NULL;
    const double *lh_l = st_l ? NULL : vec_lh(e, op->left), *lh_r = st_r ? NULL : vec_lh(e, op->right);
    const int16_t *sc_l = st_l ? NULL : vec_sc(e, op->left), *sc_r = st_r ? NULL : vec_sc(e, op->right);
    double *tmp = (double *)malloc(S * 8), *expleft = (double *)malloc(S * 8), *expright = (double *)malloc(S * 8);
    double *tipl = (double *)malloc(S * 8), *tipr = (double *)malloc(S * 8), *tip0 = (double *)malloc(S * 8);
    for (size_t ptn = 0; ptn < nptn; ptn++) {
        const size_t m = (size_t)e->ptn_model[ptn];
        const double *eval = e->eval + m * S, *evec = e->evec + m * S * S, *inv_evec = e->inv_evec + m * S * S;
        double *partial_lh = dst + ptn * block;
        double lh_max = 0.0;
        dsc[ptn] = (int16_t)((sc_l ? sc_l[ptn] : 0) + (sc_r ? sc_r[ptn] : 0));      /* :233, :267, :351 */
        if (st_l) sm_tip(e, ptn, st_l[ptn], tipl);
        if (st_r) sm_tip(e, ptn, st_r[ptn], tipr);
        for (size_t c = 0; c < K; c++) {
            const double len_left = e->blk_rate[c] * op_len(e, op, 0), len_right = e->blk_rate[c] * op_len(e, op, 1);
            const double *pl = st_l ? tipl : lh_l + ptn * block + c * S, *pr = st_r ? tipr : lh_r + ptn * block + c * S;
            for (size_t i = 0; i < S; i++) {
                expleft[i] = exp(eval[i] * len_left) * pl[i];
                expright[i] = exp(eval[i] * len_right) * pr[i];
            }
            for (size_t x = 0; x < S; x++) {
                double vleft = 0.0, vright = 0.0;
                const double *this_evec = evec + x * S;
                for (size_t i = 0; i < S; i++) { vleft += this_evec[i] * expleft[i]; vright += this_evec[i] * expright[i]; }
                tmp[x] = vleft * vright;
            }
            for (size_t i = 0; i < S; i++) {
                double res = 0.0;
                for (size_t x = 0; x < S; x++) res += tmp[x] * inv_evec[i * S + x];
                partial_lh[c * S + i] = res;
                if (fabs(res) > lh_max) lh_max = fabs(res);
            }
        }
        if (!(st_l && st_r) && lh_max < SCALING_THRESHOLD) {       /* the cherry case never scales (:222-262) */
            if (lh_max == 0.0) {                                    /* :308-320: &tip_partial_lh[ptn*nstates] is taxon 0's vector */
                sm_tip(e, ptn, e->states[ptn], tip0);
                for (size_t c = 0; c < K; c++) memcpy(&partial_lh[c * S], tip0, S * 8);
                dsc[ptn] += 4;
            } else {
                for (size_t i = 0; i < block; i++) partial_lh[i] *= SCALING_THRESHOLD_INVER;
                dsc[ptn] += 1;
            }
        }
    }
    free(tmp); free(expleft); free(expright); free(tipl); free(tipr); free(tip0);
    e->launches++;
    return LKCUDA_OK;
}

static int sm_branch(lkcuda_engine *e, int32_t dad, int32_t node, double len, double *result) {
    size_t S = e->S, K = e->K, block = e->block, nptn = e->nptn_all, P = (size_t)e->cfg.nptn;
    if (!e->ptn_model) return fail(e, LKCUDA_ERR_STATE, "site models not set");
    double *lh_all = (double *)malloc((nptn + 1) * 8), *tip = (double *)malloc(S * 8);
    int16_t *ssn = (int16_t *)calloc((size_t)e->cfg.nunobs + 1, 2);
    const double *lh_node = vec_lh(e, node);
    const uint8_t *st = LKCUDA_IS_TIP(dad) ? e->states + (size_t)LKCUDA_TIP_TAXON(dad) * nptn : NULL;
    const double *lh_dad = st ? NULL : vec_lh(e, dad);
    for (size_t ptn = 0; ptn < nptn; ptn++) {       /* :583-606 (tip), :612-636 (internal) */
        const double *eval = e->eval + (size_t)e->ptn_model[ptn] * S;
        double lh_ptn = e->ptn_invar[ptn];
        if (st) sm_tip(e, ptn, st[ptn], tip);
        for (size_t c = 0; c < K; c++) {
            double lh_cat = 0.0;
            const double l = e->blk_rate[c] * len, prop = e->blk_weight[c];
            const double *a = st ? tip : lh_dad + ptn * block + c * S, *b = lh_node + ptn * block + c * S;
            for (size_t i = 0; i < S; i++) lh_cat += exp(eval[i] * l) * a[i] * b[i];
            lh_cat *= prop;
            if (ptn < P) e->pattern_lh_cat[ptn * K + c] = lh_cat;
            lh_ptn += lh_cat;
        }
        lh_all[ptn] = lh_ptn;
    }
    double scale_sum = vec_scale_factor(e, node) + (st ? 0.0 : vec_scale_factor(e, dad));
    finish_lnl(e, lh_all, ssn, scale_sum);
    free(lh_all); free(tip); free(ssn);
    e->launches++;
    if (result) memcpy(result, e->result, sizeof(e->result_own));
    return LKCUDA_OK;
}

static int sm_theta(lkcuda_engine *e, int32_t dad, int32_t node) {
    size_t S = e->S, K = e->K, block = e->block, nptn = e->nptn_all;
    if (!e->ptn_model) return fail(e, LKCUDA_ERR_STATE, "site models not set");
    const double *lh_node = vec_lh(e, node);
    if (LKCUDA_IS_TIP(dad)) {       /* :463-482 */
        const uint8_t *st = e->states + (size_t)LKCUDA_TIP_TAXON(dad) * nptn;
        double *tip = (double *)malloc(S * 8);
        for (size_t ptn = 0; ptn < nptn; ptn++) {
            sm_tip(e, ptn, st[ptn], tip);
            for (size_t c = 0; c < K; c++)
                for (size_t i = 0; i < S; i++) e->theta[ptn * block + c * S + i] = tip[i] * lh_node[ptn * block + c * S + i];
        }
        free(tip);
    } else {                        /* :484-497 */
        const double *lh_dad = vec_lh(e, dad);
        for (size_t i = 0; i < nptn * block; i++) e->theta[i] = lh_dad[i] * lh_node[i];
    }
    e->theta_scale_sum = vec_scale_factor(e, node) + (LKCUDA_IS_TIP(dad) ? 0.0 : vec_scale_factor(e, dad));
    e->theta_ready = 1;
    e->launches++;
    return LKCUDA_OK;
}

static int sm_derv(lkcuda_engine *e, double len, double *df_out, double *ddf_out) {
    size_t S = e->S, K = e->K, block = e->block, P = (size_t)e->cfg.nptn;
    double my_df = 0.0, my_ddf = 0.0;
    for (size_t ptn = 0; ptn < P; ptn++) {      /* :505-538 */
        double lh_ptn = e->ptn_invar[ptn], df_ptn = 0.0, ddf_ptn = 0.0;
        const double *theta = e->theta + ptn * block;
        const double *eval = e->eval + (size_t)e->ptn_model[ptn] * S;
        for (size_t c = 0; c < K; c++) {
            double lh_cat = 0.0, df_cat = 0.0, ddf_cat = 0.0;
            for (size_t i = 0; i < S; i++) {
                double cof = eval[i] * e->blk_rate[c];
                double val = exp(cof * len) * theta[i];
                double val1 = cof * val;
                lh_cat += val; df_cat += val1; ddf_cat += cof * val1;
            }
            double prop = e->blk_weight[c];
            lh_ptn += prop * lh_cat; df_ptn += prop * df_cat; ddf_ptn += prop * ddf_cat;
            theta += S;
        }
        lh_ptn = 1.0 / fabs(lh_ptn);
        double df_frac = df_ptn * lh_ptn, ddf_frac = ddf_ptn * lh_ptn, freq = e->ptn_freq[ptn];
        double tmp1 = df_frac * freq, tmp2 = ddf_frac * freq;
        my_df += tmp1;
        my_ddf += tmp2 - tmp1 * df_frac;
    }
    if (isnan(my_df) || isinf(my_df)) { my_df = 0.0; my_ddf = 0.0; }
    *df_out = my_df; *ddf_out = my_ddf;
    e->launches++;
    return LKCUDA_OK;
}

static int sm_frombuf(lkcuda_engine *e, double len, double *result) {
    size_t S = e->S, K = e->K, block = e->block, nptn = e->nptn_all;
    double *lh_all = (double *)malloc((nptn + 1) * 8);
    int16_t *ssn = (int16_t *)calloc((size_t)e->cfg.nunobs + 1, 2);
    for (size_t ptn = 0; ptn < nptn; ptn++) {   /* :680-699 */
        double lh_ptn = e->ptn_invar[ptn];
        const double *theta = e->theta + ptn * block;
        const double *eval = e->eval + (size_t)e->ptn_model[ptn] * S;
        for (size_t c = 0; c < K; c++) {
            double lh_cat = 0.0;
            const double l = e->blk_rate[c] * len;
            for (size_t i = 0; i < S; i++) lh_cat += exp(eval[i] * l) * theta[i];
            lh_ptn += lh_cat * e->blk_weight[c];
            theta += S;
        }
        lh_all[ptn] = lh_ptn;
    }
    finish_lnl(e, lh_all, ssn, e->theta_scale_sum);
    free(lh_all); free(ssn);
    e->launches++;
    if (result) memcpy(result, e->result, sizeof(e->result_own));
    return LKCUDA_OK;
}
