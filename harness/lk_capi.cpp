/*
 * lk_capi.cpp — flat C entry points over lk::PhyloTree (the host mirror of the reference's PhyloTree
 * likelihood API) so that Python (ctypes) tests and bench.py drive exactly the code a C++ host would.
 * Declared in include/lktree.h.  C++ exceptions (the mirror's outError analogue) become status 1 +
 * lktree_last_error().
 */
#include "lk_phylotree.h"
#include "lktree.h"
#include <cstring>
#include <map>

using lk::PhyloTree;
using lk::Node;
using lk::Neighbor;

struct lktree {
    PhyloTree tree;
    std::string err;
    std::string text;
};

#define LK_TRY(h) try {
#define LK_CATCH(h) } catch (const std::exception &ex) { (h)->err = ex.what(); return 1; } catch (...) { (h)->err = "unknown error"; return 1; } return 0;

static Node *nodeById(lktree *h, int id) {
    for (size_t i = 0; i < h->tree.allNodes.size(); i++) if (h->tree.allNodes[i]->id == id) return h->tree.allNodes[i];
    throw std::runtime_error("no node with id " + std::to_string(id));
}

extern "C" {

lktree *lktree_create(void) { return new (std::nothrow) lktree(); }
void lktree_destroy(lktree *h) { delete h; }
const char *lktree_last_error(const lktree *h) { return h ? h->err.c_str() : "null handle"; }

int lktree_set_alignment(lktree *h, int32_t ntaxa, int64_t nptn, const uint8_t *states, const double *ptn_freq,
                         const char *const *names, int32_t seq_type, int32_t nstates, int32_t state_unknown) {
    LK_TRY(h)
    std::vector<std::string> nm;
    for (int i = 0; i < ntaxa; i++) nm.push_back(names ? names[i] : ("t" + std::to_string(i)));
    h->tree.setAlignment(ntaxa, nptn, states, ptn_freq, nm, seq_type, nstates, state_unknown);
    LK_CATCH(h)
}

int lktree_set_const_char(lktree *h, const int32_t *const_char) { LK_TRY(h) h->tree.setConstChar(const_char); LK_CATCH(h) }

int lktree_set_unobserved(lktree *h, const uint8_t *unobs, int32_t n) {
    LK_TRY(h) h->tree.setUnobservedPatterns(std::vector<uint8_t>(unobs, unobs + n)); LK_CATCH(h)
}

int lktree_set_model(lktree *h, int32_t family, int32_t ncat, int32_t nmix, const double *eval, const double *evec,
                     const double *inv_evec, const double *rates, const double *props, const double *mixprops,
                     const double *state_freq, double p_invar) {
    LK_TRY(h) h->tree.setModel(family, ncat, nmix, eval, evec, inv_evec, rates, props, mixprops, state_freq, p_invar); LK_CATCH(h)
}

int lktree_set_site_models(lktree *h, const int32_t *ptn_model) { LK_TRY(h) h->tree.setSiteModels(ptn_model); LK_CATCH(h) }

int lktree_read_newick(lktree *h, const char *newick) { LK_TRY(h) h->tree.readTreeString(newick); LK_CATCH(h) }
int lktree_read_adjacency(lktree *h, const char *text) { LK_TRY(h) h->tree.readAdjacency(text); LK_CATCH(h) }

int lktree_get_tree_string(lktree *h, char *buf, size_t buflen) {
    LK_TRY(h)
    std::string s = h->tree.getTreeString();
    if (s.size() + 1 > buflen) throw std::runtime_error("buffer too small");
    memcpy(buf, s.c_str(), s.size() + 1);
    LK_CATCH(h)
}

int lktree_set_devices(lktree *h, const int32_t *devices, int32_t n) {
    LK_TRY(h) h->tree.setDevices(std::vector<int>(devices, devices + n)); LK_CATCH(h)
}
int lktree_set_process_shard(lktree *h, int32_t rank, int32_t world) { LK_TRY(h) h->tree.setProcessShard(rank, world); LK_CATCH(h) }
int lktree_set_allreduce(lktree *h, lktree_allreduce_fn fn, void *user) { LK_TRY(h) h->tree.setAllReduce(fn, user); LK_CATCH(h) }
int lktree_set_boot_samples(lktree *h, const float *samples, int32_t nsamples) { LK_TRY(h) h->tree.setBootSamples(samples, nsamples); LK_CATCH(h) }
int lktree_compute_rell(lktree *h, double *rell) { LK_TRY(h) h->tree.computeRell(rell); LK_CATCH(h) }
int lktree_optimize_mixture_weights(lktree *h, double *lnl, double *weights) {
    LK_TRY(h)
    *lnl = h->tree.optimizeMixtureWeights();
    if (weights) memcpy(weights, &h->tree.mixtureWeights()[0], h->tree.mixtureWeights().size() * sizeof(double));
    LK_CATCH(h)
}
int lktree_compute_pattern_rates(lktree *h, double *pattern_rates, int32_t *pattern_cat) {
    LK_TRY(h) h->tree.computePatternRates(pattern_rates, pattern_cat); LK_CATCH(h)
}
int lktree_peer_handle(lktree *h, uint8_t *handle) { LK_TRY(h) h->tree.peerHandle(handle); LK_CATCH(h) }
int lktree_attach_peers(lktree *h, const uint8_t *handles) { LK_TRY(h) h->tree.attachPeers(handles); LK_CATCH(h) }
int lktree_detach_peers(lktree *h) { LK_TRY(h) h->tree.detachPeers(); LK_CATCH(h) }
int lktree_peers_attached(lktree *h) { return h && h->tree.peersAttached() ? 1 : 0; }
int lktree_set_stream(lktree *h, void *stream) { LK_TRY(h) h->tree.setStream(stream); LK_CATCH(h) }

int lktree_set_params(lktree *h, double min_branch_length, double max_branch_length, int32_t device_newton, int32_t generic_kernel) {
    LK_TRY(h)
    h->tree.min_branch_length = min_branch_length;
    h->tree.max_branch_length = max_branch_length;
    h->tree.device_newton = device_newton != 0;
    if ((generic_kernel != 0) != h->tree.generic_kernel) { h->tree.generic_kernel = generic_kernel != 0; h->tree.deleteAllPartialLh(); }
    LK_CATCH(h)
}

int lktree_initialize_all_partial_lh(lktree *h) { LK_TRY(h) h->tree.initializeAllPartialLh(); LK_CATCH(h) }
int lktree_delete_all_partial_lh(lktree *h) { LK_TRY(h) h->tree.deleteAllPartialLh(); LK_CATCH(h) }
int lktree_clear_all_partial_lh(lktree *h) { LK_TRY(h) h->tree.clearAllPartialLH(); LK_CATCH(h) }

int lktree_compute_likelihood(lktree *h, double *lnl, double *pattern_lh) {
    LK_TRY(h) *lnl = h->tree.computeLikelihood(pattern_lh); LK_CATCH(h)
}

/* how often getBestNNIForBran ran in its queued form (device length slots) and how often that had to be redone call by call */
int lktree_nni_queued_stats(lktree *h, uint64_t *calls, uint64_t *fallbacks) {
    LK_TRY(h)
    if (calls) *calls = h->tree.nni_queued_calls;
    if (fallbacks) *fallbacks = h->tree.nni_queued_fallbacks;
    LK_CATCH(h)
}

/* the benchmark step as a host C++ caller issues it - clearAllPartialLH(); computeLikelihood(); computeLikelihoodDerv(node1 - node2, theta
   rebuilt) - repeated `steps` times without returning to the interpreter in between; out = {lnl, df, ddf} of the last step */
int lktree_bench_steps(lktree *h, int32_t node1, int32_t node2, int32_t steps, double *out) {
    LK_TRY(h)
    Node *n1 = nodeById(h, node1), *n2 = nodeById(h, node2);
    Neighbor *nb = n1->findNeighbor(n2);
    if (!nb) throw std::runtime_error("nodes are not adjacent");
    for (int s = 0; s < steps; s++) {
        h->tree.clearAllPartialLH();
        out[0] = h->tree.computeLikelihood(NULL);
        h->tree.theta_computed = false;
        h->tree.current_it = nb;
        h->tree.current_it_back = n2->findNeighbor(n1);
        h->tree.computeLikelihoodDerv(nb, n1, out[1], out[2]);
    }
    LK_CATCH(h)
}

int lktree_compute_likelihood_branch(lktree *h, int32_t node1, int32_t node2, double *lnl) {
    LK_TRY(h)
    Node *n1 = nodeById(h, node1), *n2 = nodeById(h, node2);
    Neighbor *nb = n1->findNeighbor(n2);
    if (!nb) throw std::runtime_error("nodes are not adjacent");
    *lnl = h->tree.computeLikelihoodBranch(nb, n1);
    LK_CATCH(h)
}

int lktree_compute_likelihood_derv(lktree *h, int32_t node1, int32_t node2, int32_t reset_theta, double *df, double *ddf) {
    LK_TRY(h)
    Node *n1 = nodeById(h, node1), *n2 = nodeById(h, node2);
    Neighbor *nb = n1->findNeighbor(n2);
    if (!nb) throw std::runtime_error("nodes are not adjacent");
    if (reset_theta) h->tree.theta_computed = false;
    h->tree.current_it = nb;
    h->tree.current_it_back = n2->findNeighbor(n1);
    h->tree.computeLikelihoodDerv(nb, n1, *df, *ddf);
    LK_CATCH(h)
}

int lktree_compute_likelihood_from_buffer(lktree *h, double *lnl) { LK_TRY(h) *lnl = h->tree.computeLikelihoodFromBuffer(); LK_CATCH(h) }

int lktree_set_branch_length(lktree *h, int32_t node1, int32_t node2, double len) {
    LK_TRY(h)
    Node *n1 = nodeById(h, node1), *n2 = nodeById(h, node2);
    Neighbor *a = n1->findNeighbor(n2), *b = n2->findNeighbor(n1);
    if (!a || !b) throw std::runtime_error("nodes are not adjacent");
    a->length = b->length = len;
    LK_CATCH(h)
}

int lktree_optimize_one_branch(lktree *h, int32_t node1, int32_t node2, int32_t max_nr_step, double *len) {
    LK_TRY(h)
    Node *n1 = nodeById(h, node1), *n2 = nodeById(h, node2);
    h->tree.optimizeOneBranch(n1, n2, true, max_nr_step);
    if (len) *len = n1->findNeighbor(n2)->length;
    LK_CATCH(h)
}

int lktree_get_best_nni_for_bran(lktree *h, int32_t node1, int32_t node2, int32_t nni5, double *out) {
    LK_TRY(h)
    Node *n1 = nodeById(h, node1), *n2 = nodeById(h, node2);
    PhyloTree::NNIMove mv[2];
    h->tree.getBestNNIForBran(n1, n2, mv, nni5 != 0);
    for (int c = 0; c < 2; c++) {
        out[8 * c] = mv[c].newloglh;
        for (int i = 0; i < 5; i++) out[8 * c + 1 + i] = mv[c].newLen[i];
        out[8 * c + 6] = mv[c].node1_nei; out[8 * c + 7] = mv[c].node2_nei;
    }
    LK_CATCH(h)
}

int lktree_optimize_all_branches(lktree *h, int32_t iterations, double tolerance, int32_t max_nr_step, double *lnl) {
    LK_TRY(h) *lnl = h->tree.optimizeAllBranches(iterations, tolerance, max_nr_step); LK_CATCH(h)
}

int lktree_get_traversal(lktree *h, int32_t *pairs, int32_t capacity, int32_t *count) {
    LK_TRY(h)
    std::vector<Node*> a, b;
    h->tree.computeBestTraversal(a, b);
    *count = (int32_t)a.size();
    for (size_t i = 0; i < a.size() && (int32_t)i < capacity; i++) { pairs[2 * i] = a[i]->id; pairs[2 * i + 1] = b[i]->id; }
    LK_CATCH(h)
}

int lktree_get_branch_length(lktree *h, int32_t node1, int32_t node2, double *len) {
    LK_TRY(h)
    Neighbor *a = nodeById(h, node1)->findNeighbor(nodeById(h, node2));
    if (!a) throw std::runtime_error("nodes are not adjacent");
    *len = a->length;
    LK_CATCH(h)
}

int lktree_branch_vector(lktree *h, int32_t from, int32_t to, int32_t *vec, int32_t *computed) {
    LK_TRY(h)
    Neighbor *a = nodeById(h, from)->findNeighbor(nodeById(h, to));
    if (!a) throw std::runtime_error("nodes are not adjacent");
    *vec = a->vec;
    if (computed) *computed = a->partial_lh_computed & 1;
    LK_CATCH(h)
}

int lktree_root_branch(lktree *h, int32_t *from, int32_t *to) {
    LK_TRY(h)
    if (!h->tree.current_it) throw std::runtime_error("no current branch");
    *from = h->tree.current_it_back->node->id;
    *to = h->tree.current_it->node->id;
    LK_CATCH(h)
}

int lktree_fetch_tip_partial_lh(lktree *h, double *out, int64_t capacity, int64_t *count) {
    LK_TRY(h)
    h->tree.computeTipPartialLikelihood();
    *count = (int64_t)h->tree.tip_partial_lh.size();
    if (*count <= capacity) memcpy(out, &h->tree.tip_partial_lh[0], (size_t)*count * 8);
    LK_CATCH(h)
}

int lktree_fetch_ptn_invar(lktree *h, double *out) {
    LK_TRY(h)
    h->tree.computeTipPartialLikelihood();
    memcpy(out, &h->tree.ptn_invar[0], h->tree.ptn_invar.size() * 8);
    LK_CATCH(h)
}

int lktree_compute_pattern_lh_cat(lktree *h, int32_t wsl, double *out, int32_t *ncol) {
    LK_TRY(h) int n = h->tree.computePatternLhCat(out, wsl); if (ncol) *ncol = n; LK_CATCH(h)
}

lkcuda_engine *lktree_engine(lktree *h, int32_t i) { return h ? h->tree.engine((size_t)i) : NULL; }

int lktree_local_patterns(lktree *h, int64_t *begin, int64_t *end) {
    LK_TRY(h) *begin = h->tree.localPatternBegin(); *end = h->tree.localPatternEnd(); LK_CATCH(h)
}

int lktree_shard_patterns(lktree *h, int32_t i, int64_t *begin, int64_t *end) {
    LK_TRY(h)
    if (!h->tree.engine((size_t)i)) throw std::runtime_error("no such shard");
    *begin = h->tree.shardBegin((size_t)i); *end = h->tree.shardBegin((size_t)i + 1);
    LK_CATCH(h)
}

int lktree_compute_likelihood_async_fetch(lktree *h, double *lnl, double *pattern_lh) {
    LK_TRY(h) *lnl = h->tree.computeLikelihoodAsyncFetch(pattern_lh); LK_CATCH(h)
}
int lktree_wait_pattern_lh(lktree *h) { LK_TRY(h) h->tree.waitPatternLh(); LK_CATCH(h) }

int lktree_optimize_rate_free_em(lktree *h, int32_t copy_precision, double *lnl, double *rates, double *props) {
    LK_TRY(h)
    *lnl = h->tree.optimizeRateFreeEM(copy_precision);
    for (int c = 0; c < h->tree.ncat; c++) { if (rates) rates[c] = h->tree.categoryRates()[c]; if (props) props[c] = h->tree.categoryProps()[c]; }
    LK_CATCH(h)
}

/* ---- Fitch parsimony ---- */
int lktree_set_parsimony_sites(lktree *h, int64_t nsites, const uint8_t *site_states, int64_t aln_nsite, int32_t device) {
    LK_TRY(h) h->tree.setParsimonySites(nsites, site_states, aln_nsite); h->tree.pars_device = device; LK_CATCH(h)
}
int lktree_compute_parsimony(lktree *h, int32_t *score) { LK_TRY(h) *score = h->tree.computeParsimony(); LK_CATCH(h) }
int lktree_compute_parsimony_branch(lktree *h, int32_t node1, int32_t node2, int32_t *score, int32_t *branch_subst) {
    LK_TRY(h)
    Node *n1 = nodeById(h, node1), *n2 = nodeById(h, node2);
    Neighbor *nei = n1->findNeighbor(n2);
    if (!nei) throw std::runtime_error("not a branch");
    int subst = 0;
    *score = h->tree.computeParsimonyBranch(nei, n1, &subst);
    if (branch_subst) *branch_subst = subst;
    LK_CATCH(h)
}
int lktree_fix_negative_branch(lktree *h, int32_t force, int32_t *fixed) {
    LK_TRY(h) int n = h->tree.fixNegativeBranch(force != 0); if (fixed) *fixed = n; LK_CATCH(h)
}
int lktree_compute_parsimony_tree(lktree *h, const int32_t *taxon_order, int32_t *score) {
    LK_TRY(h) *score = h->tree.computeParsimonyTree(taxon_order); LK_CATCH(h)
}
int lktree_get_branches(lktree *h, int32_t *pairs, int32_t capacity, int32_t *count) {
    LK_TRY(h)
    std::vector<Node*> a, b;
    h->tree.getBranches(a, b);
    *count = (int32_t)a.size();
    for (size_t i = 0; i < a.size() && (int32_t)i < capacity; i++) { pairs[2 * i] = a[i]->id; pairs[2 * i + 1] = b[i]->id; }
    LK_CATCH(h)
}
static void collectNodes(Node *node, Node *dad, std::vector<Node*> &out) {
    out.push_back(node);
    for (size_t i = 0; i < node->neighbors.size(); i++) if (node->neighbors[i]->node != dad) collectNodes(node->neighbors[i]->node, node, out);
}
/* the text lktree_read_adjacency reads: nodes in pre-order from the root, neighbours in their stored order */
int lktree_get_adjacency(lktree *h, char *buf, size_t buflen) {
    LK_TRY(h)
    if (!h->tree.root) throw std::runtime_error("no tree");
    std::vector<Node*> nodes;
    collectNodes(h->tree.root, NULL, nodes);
    std::string out = "root " + std::to_string(h->tree.root->id) + " nodes " + std::to_string(nodes.size()) + " leaves " + std::to_string(h->tree.leafNum) + "\n";
    char num[64];
    for (size_t i = 0; i < nodes.size(); i++) {
        Node *n = nodes[i];
        out += "node " + std::to_string(n->id) + " " + (n->isLeaf() ? "1" : "0") + " " + (n->name.empty() ? std::string("-") : n->name) + " " + std::to_string(n->neighbors.size());
        for (size_t k = 0; k < n->neighbors.size(); k++) {
            snprintf(num, sizeof num, " %d %.17g", n->neighbors[k]->node->id, n->neighbors[k]->length);
            out += num;
        }
        out += "\n";
    }
    if (out.size() + 1 > buflen) throw std::runtime_error("buffer too small");
    memcpy(buf, out.c_str(), out.size() + 1);
    LK_CATCH(h)
}

lkpars_engine *lktree_pars_engine(lktree *h) { return h ? h->tree.parsEngine() : NULL; }

uint64_t lktree_num_partial_lh_computations(lktree *h) { return h ? h->tree.num_partial_lh_computations : 0; }

}  // extern "C"
