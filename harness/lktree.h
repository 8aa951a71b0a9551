/*
 * lktree.h — flat C entry points of the host-side PhyloTree mirror (iq-tree_b200/csrc/lk_phylotree.*).
 *
 * The mirror keeps the reference's PhyloTree likelihood API (same names, argument meaning and control
 * flow: phylotree.cpp:1031 computeLikelihood, :2148 optimizeOneBranch, :2252 optimizeAllBranches,
 * phylotreesse.cpp:335-357 the four kernel wrappers) and calls the device through include/lkcuda.h.
 * These functions exist so that non-C++ hosts (the ctypes tests, bench.py) drive exactly that code.
 * Every function returns 0 on success, 1 on error (text in lktree_last_error()).
 * Nodes are addressed by their reference ids: leaves = alignment sequence index, internal nodes as
 * numbered by MTree::initializeTree (mtree.cpp:561-583) or as given in an adjacency fixture.
 */
#ifndef LKTREE_H
#define LKTREE_H

#include <stddef.h>
#include <stdint.h>
#include "../include/lkcuda.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct lktree lktree;

/* sum `n` doubles over the ranks of a one-process-per-GPU job: in place in `dev_buf` (device memory,
   host_buf == NULL) or in `host_buf` (dev_buf == NULL) */
typedef void (*lktree_allreduce_fn)(void *user, double *host_buf, void *dev_buf, int n);

lktree *lktree_create(void);
void lktree_destroy(lktree *h);
const char *lktree_last_error(const lktree *h);

/* inputs: the pattern-compressed alignment (Alignment : vector<Pattern>, alignment.h:50), taxon-major */
int lktree_set_alignment(lktree *h, int32_t ntaxa, int64_t nptn, const uint8_t *states, const double *ptn_freq,
                         const char *const *names, int32_t seq_type, int32_t nstates, int32_t state_unknown);
int lktree_set_const_char(lktree *h, const int32_t *const_char);             /* Pattern::const_char */
int lktree_set_unobserved(lktree *h, const uint8_t *unobs, int32_t n);       /* +ASC */
/* family: 0 plain, 1 fused mix-rate, 2 mixture; eigensystems concatenated as ModelMixture does */
int lktree_set_model(lktree *h, int32_t family, int32_t ncat, int32_t nmix, const double *eval, const double *evec,
                     const double *inv_evec, const double *rates, const double *props, const double *mixprops,
                     const double *state_freq, double p_invar);
/* family 3 = site-specific frequency models (-fs / PMSF, phylokernelsitemodel.cpp): nmix = number of distinct site models whose
   eigensystems lktree_set_model receives; ptn_model[nptn] maps every pattern to one (ModelSet::pattern_model_map) */
int lktree_set_site_models(lktree *h, const int32_t *ptn_model);
int lktree_read_newick(lktree *h, const char *newick);
int lktree_read_adjacency(lktree *h, const char *text);
int lktree_get_tree_string(lktree *h, char *buf, size_t buflen);

/* placement */
int lktree_set_devices(lktree *h, const int32_t *devices, int32_t n);        /* in-process shards, one per device */
int lktree_set_process_shard(lktree *h, int32_t rank, int32_t world);        /* one process per GPU */
int lktree_set_allreduce(lktree *h, lktree_allreduce_fn fn, void *user);
/* consumers of the kernels' per-pattern outputs, on the device.
   lktree_set_boot_samples / lktree_compute_rell: IQTree::saveCurrentTree's RELL dot products (iqtree.cpp:2689-2735), samples = boot_samples
   [nsamples][nptn] float; rell[nsamples] for the branch of the last computeLikelihood.
   lktree_optimize_mixture_weights: ModelMixture::optimizeWeights (model/modelmixture.cpp:1356-1416): EM on the class weights, returns
   the new lnL and (optionally) the nmix weights.
   lktree_compute_pattern_rates: RateGamma::computePatternRates (model/rategamma.cpp:235-257), outputs for this process' patterns. */
int lktree_set_boot_samples(lktree *h, const float *samples, int32_t nsamples);
int lktree_compute_rell(lktree *h, double *rell);
int lktree_optimize_mixture_weights(lktree *h, double *lnl, double *weights);
int lktree_compute_pattern_rates(lktree *h, double *pattern_rates, int32_t *pattern_cat);
/* fused cross-GPU all-reduce inside the reduction kernels (lkcuda_peer_handle / lkcuda_attach_peers), one process per GPU:
   every rank exports its handle (LKCUDA_PEER_HANDLE_BYTES), the host exchanges them and passes all `world` of them in rank order */
int lktree_peer_handle(lktree *h, uint8_t *handle);
int lktree_attach_peers(lktree *h, const uint8_t *handles);
int lktree_detach_peers(lktree *h);
int lktree_peers_attached(lktree *h);
int lktree_set_stream(lktree *h, void *cuda_stream);
int lktree_set_params(lktree *h, double min_branch_length, double max_branch_length, int32_t device_newton, int32_t generic_kernel);

/* PhyloTree API */
/* PhyloTree::getBestNNIForBran (phylotree.cpp:2873-3066): both NNIs around the internal branch (node1, node2); out[16] = per NNI
   {newloglh, newLen[5], id of the subtree root taken from node1, id of the one taken from node2}.  The tree is left unchanged. */
int lktree_get_best_nni_for_bran(lktree *h, int32_t node1, int32_t node2, int32_t nni5, double *out);
int lktree_initialize_all_partial_lh(lktree *h);
int lktree_delete_all_partial_lh(lktree *h);
int lktree_clear_all_partial_lh(lktree *h);
int lktree_compute_likelihood(lktree *h, double *lnl, double *pattern_lh /* [local P] or NULL */);
/* computeLikelihood whose per-pattern values (pinned host memory, [local P]) are still being copied when it returns: the caller
   goes on (e.g. with computeLikelihoodDerv) and calls lktree_wait_pattern_lh before reading them */
int lktree_compute_likelihood_async_fetch(lktree *h, double *lnl, double *pattern_lh);
int lktree_wait_pattern_lh(lktree *h);
/* PhyloTree::getBestNNIForBran in its queued form (with device_newton: the ten branch optimisations of a call are queued through device
   length slots, one host wait per NNI): calls that took it / calls that had to be redone step by step (a run hit the upper bound) */
int lktree_nni_queued_stats(lktree *h, uint64_t *calls, uint64_t *fallbacks);
/* bench.py's timed loop: `steps` x { clearAllPartialLH; computeLikelihood; computeLikelihoodDerv(node1 - node2, theta rebuilt) }; out[3] = lnl, df, ddf */
int lktree_bench_steps(lktree *h, int32_t node1, int32_t node2, int32_t steps, double *out);
int lktree_compute_likelihood_branch(lktree *h, int32_t node1, int32_t node2, double *lnl);
int lktree_compute_likelihood_derv(lktree *h, int32_t node1, int32_t node2, int32_t reset_theta, double *df, double *ddf);
int lktree_compute_likelihood_from_buffer(lktree *h, double *lnl);
int lktree_optimize_one_branch(lktree *h, int32_t node1, int32_t node2, int32_t max_nr_step, double *len);
int lktree_optimize_all_branches(lktree *h, int32_t iterations, double tolerance, int32_t max_nr_step, double *lnl);

/* inspection */
int lktree_set_branch_length(lktree *h, int32_t node1, int32_t node2, double len);
int lktree_get_branch_length(lktree *h, int32_t node1, int32_t node2, double *len);
int lktree_get_traversal(lktree *h, int32_t *pairs, int32_t capacity, int32_t *count);
int lktree_branch_vector(lktree *h, int32_t from, int32_t to, int32_t *vec, int32_t *computed);
int lktree_root_branch(lktree *h, int32_t *from, int32_t *to);
int lktree_fetch_tip_partial_lh(lktree *h, double *out, int64_t capacity, int64_t *count);
int lktree_fetch_ptn_invar(lktree *h, double *out);
/* PhyloTree::computePatternLhCat (phylotree.cpp:1108-1160); wsl = 2 rate categories, 3 mixture classes, 4 both;
   out holds [local P][nmix*ncat] doubles at most, *ncol = entries per pattern actually written */
int lktree_compute_pattern_lh_cat(lktree *h, int32_t wsl, double *out, int32_t *ncol);
lkcuda_engine *lktree_engine(lktree *h, int32_t i);
int lktree_local_patterns(lktree *h, int64_t *begin, int64_t *end);
/* global pattern range of in-process shard i (engine i) */
int lktree_shard_patterns(lktree *h, int32_t i, int64_t *begin, int64_t *end);
/* RateFree::optimizeWithEM (model/ratefree.cpp:450-578): EM on the +R proportions and rates; rates / props [ncat] out; copy_precision =
   the reference's MTree::num_precision (6 unless min_branch_length was set before the tree was constructed, mtree.cpp:36-39) */
int lktree_optimize_rate_free_em(lktree *h, int32_t copy_precision, double *lnl, double *rates, double *props);

/* Fitch parsimony on the device (include/lkpars.h).  site_states[ntaxa][nsites] = the informative sites in
   Alignment::ordered_pattern order with every pattern repeated `frequency` times (alignment.cpp:731-775); aln_nsite = getNSite().
   lktree_compute_parsimony        PhyloTree::computeParsimony         phylotree.cpp:658-660
   lktree_compute_parsimony_branch PhyloTree::computeParsimonyBranch   phylotree.cpp:653-655 (score of the tree, substitutions on the branch)
   lktree_fix_negative_branch      PhyloTree::fixNegativeBranch        phylotree.cpp:2654-2694 (branch lengths from parsimony, JC-corrected)
   lktree_compute_parsimony_tree   PhyloTree::computeParsimonyTree     phylotreepars.cpp:309-426: REPLACES the tree by the stepwise-addition
                                   tree for the given addition order (the reference shuffles it with its own RNG), returns best_pars_score
   lktree_get_branches             MTree::getBranches                  mtree.cpp:905-919: (node1, node2) id pairs in visiting order */
int lktree_set_parsimony_sites(lktree *h, int64_t nsites, const uint8_t *site_states, int64_t aln_nsite, int32_t device /* -1: default */);
int lktree_compute_parsimony(lktree *h, int32_t *score);
int lktree_compute_parsimony_branch(lktree *h, int32_t node1, int32_t node2, int32_t *score, int32_t *branch_subst);
int lktree_fix_negative_branch(lktree *h, int32_t force, int32_t *fixed);
int lktree_compute_parsimony_tree(lktree *h, const int32_t *taxon_order /*[ntaxa]*/, int32_t *score);
int lktree_get_branches(lktree *h, int32_t *pairs, int32_t capacity, int32_t *count);
int lktree_get_adjacency(lktree *h, char *buf, size_t buflen);               /* the text format lktree_read_adjacency reads */
struct lkpars_engine *lktree_pars_engine(lktree *h);
uint64_t lktree_num_partial_lh_computations(lktree *h);

#ifdef __cplusplus
}
#endif
#endif
