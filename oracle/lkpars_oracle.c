/*
 * lkpars_oracle.c — TEST INFRASTRUCTURE.  CPU restatement of the reference's Fitch parsimony kernels behind the C ABI of
 * include/lkpars.h, so that the host mirror (lk::PhyloTree::computeParsimonyTree etc.) can run without a GPU in tests/.
 * Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may load the library this is linked into; the product never does.
 *
 * Follows the scalar ("Fast") kernels of the reference, in the reference's own layout partial_pars[(site/32)*nstates + state]:
 *   tips          PhyloTree::computePartialParsimonyFast, external node  phylotreepars.cpp:27-148 (dummy sites :72-74)
 *   PARTIAL       the internal-node loop                                 phylotreepars.cpp:150-212
 *   BRANCH        PhyloTree::computeParsimonyBranchFast                  phylotreepars.cpp:218-282
 *   INSERT        = PARTIAL into a scratch block followed by BRANCH against the third vector, which is what
 *                 addTaxonMPFast -> computeParsimonyBranch does          phylotreepars.cpp:428-465
 * Parity pin: tests/golden/pars/ holds scores, per-branch substitution counts and stepwise-addition trees written by the UNMODIFIED
 * reference (oracle/ref_harness.cpp `dumppars`).
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "../include/lkcuda.h"
#include "../include/lkpars.h"

struct lkpars_engine {
    int S, ntaxa, nvec;
    long long nsites, nw;
    uint32_t *bits;             /* [ntaxa + nvec][nw][S] */
    uint32_t *scratch;          /* [nw][S] */
    unsigned long long launches;
    char err[256];
};

static char g_err[256];

static int pfail(lkpars_engine *e, int code, const char *msg) {
    snprintf(e ? e->err : g_err, 256, "%s", msg);
    return code;
}

int lkpars_create(int32_t device, int32_t nstates, int32_t ntaxa, int64_t nsites, const uint8_t *site_states,
                  const uint64_t *state_mask, int32_t nmask, int32_t nvec, lkpars_engine **out) {
    (void)device;
    if (!out) return pfail(NULL, LKCUDA_ERR_ARG, "null out pointer");
    *out = NULL;
    if (nstates < 2 || nstates > 64 || ntaxa < 1 || nsites < 0 || nvec < 0 || nmask < nstates || nmask > 256) return pfail(NULL, LKCUDA_ERR_ARG, "bad sizes");
    if (!state_mask || (nsites > 0 && !site_states)) return pfail(NULL, LKCUDA_ERR_ARG, "null argument");
    lkpars_engine *e = (lkpars_engine *)calloc(1, sizeof *e);
    e->S = nstates; e->ntaxa = ntaxa; e->nvec = nvec; e->nsites = nsites;
    e->nw = (nsites + 31) / 32;
    if (e->nw == 0) e->nw = 1;
    const size_t vs = (size_t)e->nw * nstates;
    e->bits = (uint32_t *)calloc((size_t)(ntaxa + nvec) * vs, 4);
    e->scratch = (uint32_t *)calloc(vs, 4);
    for (int t = 0; t < ntaxa; t++) {
        uint32_t *p = e->bits + (size_t)t * vs;
        for (long long site = 0; site < e->nw * 32; site++) {
            uint64_t m = 1;                                     /* dummy sites: state 0 */
            if (site < nsites) { int code = site_states[(size_t)t * nsites + site]; m = code < nmask ? state_mask[code] : 0; }
            for (int i = 0; i < nstates; i++)
                if ((m >> i) & 1) p[(site / 32) * nstates + i] |= 1u << (site % 32);
        }
    }
    *out = e;
    return LKCUDA_OK;
}

int lkpars_destroy(lkpars_engine *e) {
    if (!e) return LKCUDA_OK;
    free(e->bits); free(e->scratch); free(e);
    return LKCUDA_OK;
}

const char *lkpars_last_error(const lkpars_engine *e) { return e ? e->err : g_err; }

/* phylotreepars.cpp:189-209 */
static uint32_t fitch(const lkpars_engine *e, const uint32_t *left, const uint32_t *right, uint32_t *dad) {
    uint32_t score = 0;
    const int S = e->S;
    for (long long site = 0; site < e->nw; site++) {
        const uint32_t *x = left + site * S, *y = right + site * S;
        uint32_t *z = dad + site * S, w = 0;
        for (int i = 0; i < S; i++) { z[i] = x[i] & y[i]; w |= z[i]; }
        w = ~w;
        score += (uint32_t)__builtin_popcount(w);
        for (int i = 0; i < S; i++) z[i] |= w & (x[i] | y[i]);
    }
    return score;
}

/* phylotreepars.cpp:260-277 (without the early exit: the OpenMP build has none either) */
static uint32_t mismatch(const lkpars_engine *e, const uint32_t *a, const uint32_t *b) {
    uint32_t score = 0;
    const int S = e->S;
    for (long long site = 0; site < e->nw; site++) {
        const uint32_t *x = a + site * S, *y = b + site * S;
        uint32_t w = x[0] & y[0];
        for (int i = 1; i < S; i++) w |= x[i] & y[i];
        score += (uint32_t)__builtin_popcount(~w);
    }
    return score;
}

int lkpars_run(lkpars_engine *e, const lkpars_op *ops, int32_t nops, uint32_t *counts) {
    if (!e) return LKCUDA_ERR_ARG;
    if (nops < 0 || (nops > 0 && (!ops || !counts))) return pfail(e, LKCUDA_ERR_ARG, "null argument");
    const int total = e->ntaxa + e->nvec;
    const size_t vs = (size_t)e->nw * e->S;
    for (int i = 0; i < nops; i++) {
        const lkpars_op *op = &ops[i];
        if (op->kind < 0 || op->kind > 2 || op->a < 0 || op->a >= total || op->b < 0 || op->b >= total) return pfail(e, LKCUDA_ERR_ARG, "bad op");
        if (op->kind == LKPARS_OP_INSERT && (op->c < 0 || op->c >= total)) return pfail(e, LKCUDA_ERR_ARG, "bad op");
        if (op->kind == LKPARS_OP_PARTIAL && (op->dst < e->ntaxa || op->dst >= total)) return pfail(e, LKCUDA_ERR_ARG, "dst is not an internal vector");
        if (op->kind == LKPARS_OP_INSERT && op->dst >= 0 && (op->dst < e->ntaxa || op->dst >= total)) return pfail(e, LKCUDA_ERR_ARG, "dst is not an internal vector");
        if (op->kind != LKPARS_OP_BRANCH && op->dst >= 0 && (op->dst == op->a || op->dst == op->b)) return pfail(e, LKCUDA_ERR_ARG, "dst aliases an operand");
    }
    for (int i = 0; i < nops; i++) {
        const lkpars_op *op = &ops[i];
        const uint32_t *a = e->bits + op->a * vs, *b = e->bits + op->b * vs;
        counts[2 * i] = counts[2 * i + 1] = 0;
        if (op->kind == LKPARS_OP_BRANCH) counts[2 * i] = mismatch(e, a, b);
        else {
            uint32_t *z = op->dst >= 0 ? e->bits + op->dst * vs : e->scratch;
            counts[2 * i] = fitch(e, a, b, z);
            if (op->kind == LKPARS_OP_INSERT) counts[2 * i + 1] = mismatch(e, z, e->bits + op->c * vs);
        }
    }
    e->launches++;
    return LKCUDA_OK;
}

int lkpars_words(const lkpars_engine *e, int64_t *nwords) {
    if (!e || !nwords) return LKCUDA_ERR_ARG;
    *nwords = e->nw;
    return LKCUDA_OK;
}

/* output in the ABI's plane-major layout bits[state][word] */
int lkpars_fetch_vector(lkpars_engine *e, int32_t vec, uint32_t *bits) {
    if (!e) return LKCUDA_ERR_ARG;
    if (!bits || vec < 0 || vec >= e->ntaxa + e->nvec) return pfail(e, LKCUDA_ERR_ARG, "bad vector");
    const uint32_t *p = e->bits + (size_t)vec * e->nw * e->S;
    for (long long w = 0; w < e->nw; w++)
        for (int i = 0; i < e->S; i++) bits[(size_t)i * e->nw + w] = p[w * e->S + i];
    return LKCUDA_OK;
}

int lkpars_launch_count(const lkpars_engine *e, uint64_t *launches, double *device_ms) {
    if (!e) return LKCUDA_ERR_ARG;
    if (launches) *launches = e->launches;
    if (device_ms) *device_ms = 0.0;
    return LKCUDA_OK;
}

int lkpars_set_profiling(lkpars_engine *e, int32_t on) { (void)on; return e ? LKCUDA_OK : LKCUDA_ERR_ARG; }
