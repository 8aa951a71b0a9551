/*
 * lkpars.h — C ABI of the Fitch-parsimony engine: the bit-parallel kernels behind
 *     PhyloTree::computePartialParsimonyFast[SIMD]   phylotreepars.cpp:18-215,  phylokernel.h:1265-1502
 *     PhyloTree::computeParsimonyBranchFast[SIMD]    phylotreepars.cpp:218-282, phylokernel.h:1504-1572
 * which the reference binds through computePartialParsimonyPointer / computeParsimonyBranchPointer (phylotree.h:539, 555;
 * bound in phylotreesse.cpp:42-50, phylotreeavx.cpp:20-21).  Their callers that decide ORDER stay on the host (lk::PhyloTree mirrors
 * them): computeParsimonyTree / addTaxonMPFast (phylotreepars.cpp:309-465, the stepwise-addition start tree), computeParsimony
 * (phylotree.cpp:658), fixNegativeBranch (phylotree.cpp:2654-2694).
 *
 * Data: the reference keeps, per PhyloNeighbor, `partial_pars` = one bit per informative SITE (patterns of
 * Alignment::ordered_pattern repeated `frequency` times, alignment.cpp:731-775) and state, plus the subtree's score in the last word.
 * Here a vector is bits[state][word] (uint32 words, plane-major so that a warp reads 128 contiguous bytes per state), the per-vector
 * scores live on the host (they are sums of the counts lkpars_run returns), and the dummy sites that pad the last word carry state 0 at
 * every tip exactly as in the reference (phylotreepars.cpp:72-74), so they never count.
 *
 * Vectors 0 .. ntaxa-1 are the tips (built on the device at create time); ntaxa .. ntaxa+nvec-1 are free for internal results.
 * Every function returns 0 on success, a LKCUDA_ERR_* code otherwise (text in lkpars_last_error); nothing throws, no CPU fallback.
 */
#ifndef LKPARS_H
#define LKPARS_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct lkpars_engine lkpars_engine;

#define LKPARS_OP_PARTIAL 0   /* dst = Fitch(a, b);                          count = sites where a & b is empty (the union sites)      */
#define LKPARS_OP_BRANCH  1   /* nothing written;                            count = sites where a & b is empty (branch_subst)         */
#define LKPARS_OP_INSERT  2   /* z = Fitch(a, b), written to dst if dst >= 0; count = union sites of (a, b) + sites where z & c is empty:
                                 addTaxonMPFast's score of hanging tip/subtree c onto the branch between a and b, minus the three subtree scores */

typedef struct lkpars_op {
    int32_t kind;
    int32_t dst;              /* vector written (PARTIAL; INSERT: -1 = none) */
    int32_t a, b;             /* vectors read */
    int32_t c;                /* INSERT: the third vector */
} lkpars_op;

/* site_states: [ntaxa][nsites] state codes of the INFORMATIVE sites in ordered_pattern order (pattern p repeated frequency[p] times);
   state_mask[code] = bit set of the states code stands for (code < nstates: 1 << code; STATE_UNKNOWN: all states; DNA ambiguity codes
   4..17: code - 3; protein B/Z/J: phylotreepars.cpp:36, 99-108), nmask entries.  nstates <= 64. */
int lkpars_create(int32_t device, int32_t nstates, int32_t ntaxa, int64_t nsites, const uint8_t *site_states,
                  const uint64_t *state_mask, int32_t nmask, int32_t nvec, lkpars_engine **out);
int lkpars_destroy(lkpars_engine *e);
const char *lkpars_last_error(const lkpars_engine *e);   /* e may be NULL: error of the last failed create */

/* runs the ops IN ORDER (an op may read what an earlier op of the same call wrote) in ONE launch; counts[i] = the count of op i */
int lkpars_run(lkpars_engine *e, const lkpars_op *ops, int32_t nops, uint32_t *counts);

int lkpars_words(const lkpars_engine *e, int64_t *nwords);                    /* uint32 words per state plane */
int lkpars_fetch_vector(lkpars_engine *e, int32_t vec, uint32_t *bits /*[nstates][nwords]*/);
int lkpars_launch_count(const lkpars_engine *e, uint64_t *launches, double *device_ms /* of profiled launches, or NULL */);
int lkpars_set_profiling(lkpars_engine *e, int32_t on);

#ifdef __cplusplus
}
#endif
#endif
