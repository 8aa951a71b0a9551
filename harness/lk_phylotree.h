/*
 * lk_phylotree.h — host-side mirror of the reference's PhyloTree likelihood API, written against the
 * C ABI of include/lkcuda.h.  It keeps the reference's names, argument meaning and control flow for
 * the functions that decide ORDER: which partials are recomputed, left/right
 * orientation, root-branch choice, branch visiting order and the Newton step sequence.  All
 * per-pattern arithmetic happens on the device behind lkcuda_*.
 *
 * Mirrors (reference file:line):
 *   Node / Neighbor / PhyloNeighbor      node.h, phylonode.h:102-127
 *   clearAllPartialLH                    phylotree.cpp:495-502, phylonode.cpp:53-65
 *   clearReversePartialLh                phylonode.cpp:42-51
 *   initializeAllPartialLh               phylotree.cpp:667-716, 834-987 (index assignment :916-974)
 *   computeTipPartialLikelihood          phylotreesse.cpp:359-529
 *   computePtnFreq / computePtnInvar     phylotreesse.cpp:531-569
 *   computePartialLikelihood             phylokernel.h:71-143 (recursion + dirty bits + leaf-left swap)
 *   computeLikelihoodBranch              phylokernel.h:734-751
 *   computeLikelihoodDerv                phylokernel.h:486-503 (theta_computed flag :535)
 *   computeLikelihoodFromBuffer          phylotreesse.cpp:349-357
 *   computeLikelihood                    phylotree.cpp:1031-1071
 *   findFarthestLeaf / getPreOrderBranches   mtree.cpp:2052-2068, 2096-2116
 *   computeBestTraversal                 phylotree.cpp:2241-2250
 *   computeFuncDerv                      phylotree.cpp:2135-2146
 *   minimizeNewton                       optimization.cpp:388-465
 *   optimizeOneBranch / optimizeAllBranches  phylotree.cpp:2148-2192, 2252-2332
 *   getBestNNIForBran                    phylotree.cpp:2873-3066 (NNI evaluation of the tree search)
 *   readTree / parseFile (Newick)        mtree.cpp:503-680
 *   initializeAllPartialPars             phylotree.cpp:596-624            (lk_parsimony.cpp)
 *   computePartialParsimony / computeParsimonyBranch / computeParsimony   phylotreepars.cpp:18-282, phylotree.cpp:638-660
 *   computeParsimonyTree / addTaxonMPFast    phylotreepars.cpp:309-465    (stepwise-addition start tree)
 *   fixNegativeBranch                    phylotree.cpp:2654-2694
 *   getBranches                          mtree.cpp:905-919
 */
#ifndef LK_PHYLOTREE_H
#define LK_PHYLOTREE_H

#include <cstdint>
#include <string>
#include <vector>
#include <stdexcept>
#include "../include/lkcuda.h"
#include "../include/lkpars.h"

namespace lk {

struct Node;

/* PhyloNeighbor (phylonode.h:102-127): partial_lh/scale_num become the arena vector index `vec` */
struct Neighbor {
    Node *node;                 // the node this neighbour points to
    double length;
    int id;
    int vec;                    // arena vector holding the partial of the subtree rooted at `node`; -1 = none (leaf)
    int partial_lh_computed;    // bit 0 = likelihood partial up to date, bit 1 = parsimony partial up to date
    int pars_vec;               // partial_pars: block index in central_partial_pars (a neighbour that points to a leaf uses the tip's vector instead)
    int len_slot;               // >= 0 while an NNI is queued: the branch's current length is in this device length slot (lkcuda.h), not in `length`
    Neighbor(Node *n, double len) : node(n), length(len), id(-1), vec(-1), partial_lh_computed(0), pars_vec(-1), len_slot(-1) {}
};

struct Node {
    int id;
    std::string name;
    std::vector<Neighbor*> neighbors;
    double height;
    Node() : id(-1), height(0.0) {}
    bool isLeaf() const { return neighbors.size() <= 1; }
    int degree() const { return (int)neighbors.size(); }
    Neighbor *findNeighbor(const Node *n) const {
        for (size_t i = 0; i < neighbors.size(); i++) if (neighbors[i]->node == n) return neighbors[i];
        return NULL;
    }
};

enum SeqType { SEQ_DNA = 0, SEQ_PROTEIN = 1, SEQ_BINARY = 2, SEQ_MORPH = 3, SEQ_MULTISTATE = 4, SEQ_CODON = 5 };
enum ModelFamily { FAMILY_PLAIN = 0, FAMILY_MIXRATE = 1, FAMILY_MIXTURE = 2,
                   FAMILY_SITEMODEL = 3 /* -fs / PMSF (phylokernelsitemodel.cpp): nmix = number of distinct site models, setSiteModels maps patterns to them */ };

typedef void (*AllReduceFn)(void *user, double *host_buf /*may be NULL*/, void *dev_buf, int n);

class PhyloTree {
public:
    PhyloTree();
    ~PhyloTree();

    /* ---- tree ---- */
    void readTreeString(const std::string &newick);                 // mtree.cpp:503-680 + setAlignment leaf ids
    void readAdjacency(const std::string &adj_text);                // fixture format of oracle/ref_harness.cpp
    std::string getTreeString(int precision = 17) const;
    void setRootNode(const std::string &name);                      // phylotree.cpp:344-352
    int leafNum, nodeNum;
    Node *root;
    std::vector<Node*> allNodes;

    /* ---- alignment (already pattern-compressed; taxon-major states) ---- */
    void setAlignment(int ntaxa, int64_t nptn, const uint8_t *states, const double *ptn_freq,
                      const std::vector<std::string> &names, int seq_type, int nstates, int state_unknown);
    void setUnobservedPatterns(const std::vector<uint8_t> &unobs);  // +ASC (model_factory->unobserved_ptns)
    void setConstChar(const int32_t *const_char);                   // Pattern::const_char, for ptn_invar

    /* ---- model (host-side contract: SURVEY §2 rows 8, 9) ---- */
    void setModel(int family, int ncat, int nmix, const double *eval, const double *evec, const double *inv_evec,
                  const double *rates, const double *props, const double *mixprops,
                  const double *state_freq, double p_invar);

    void setSiteModels(const int32_t *ptn_model);                   // ModelSet::pattern_model_map (model/modelfactory.cpp:316-320), [nptn]

    /* ---- engine / shards ---- */
    // devices: CUDA ordinals, one pattern shard per entry; shard_rank/shard_world further restrict this
    // process to the patterns of one rank (one process per GPU under torchrun)
    void setDevices(const std::vector<int> &devices) { devices_ = devices; }
    void setProcessShard(int rank, int world) { proc_rank_ = rank; proc_world_ = world; }
    void setAllReduce(AllReduceFn fn, void *user) { allreduce_ = fn; allreduce_user_ = user; }
    void setStream(void *cuda_stream);
    // fused cross-GPU all-reduce inside the reduction kernels (lkcuda_attach_peers), one engine per process: every rank
    // exports peerHandle(), the host exchanges the handles by any transport and gives all of them to attachPeers()
    void peerHandle(uint8_t *handle /*[LKCUDA_PEER_HANDLE_BYTES]*/);
    void attachPeers(const uint8_t *handles /*[world][LKCUDA_PEER_HANDLE_BYTES]*/);
    void detachPeers();
    bool peersAttached() const { return peers_attached_; }
    void initializeAllPartialLh();
    void deleteAllPartialLh();
    int64_t localPatternBegin() const { return ptn_begin_; }
    int64_t localPatternEnd() const { return ptn_end_; }
    int64_t shardBegin(size_t i) const { return i < shard_begin_.size() ? shard_begin_[i] : ptn_end_; }
    size_t numEngines() const { return engines_.size(); }
    lkcuda_engine *engine(size_t i = 0) { return i < engines_.size() ? engines_[i] : NULL; }

    /* ---- the PhyloTree likelihood API ---- */
    void clearAllPartialLH();
    void computeTipPartialLikelihood();
    void computePtnFreq();
    void computePtnInvar();
    void computePartialLikelihood(Neighbor *dad_branch, Node *dad);
    double computeLikelihoodBranch(Neighbor *dad_branch, Node *dad);
    void computeLikelihoodDerv(Neighbor *dad_branch, Node *dad, double &df, double &ddf);
    double computeLikelihoodFromBuffer();
    double computeLikelihood(double *pattern_lh = NULL);
    // the same with the per-pattern values still on their way to (pinned) host memory when it returns; waitPatternLh() ends the transfer
    double computeLikelihoodAsyncFetch(double *pattern_lh);
    void waitPatternLh();
    int computePatternLhCat(double *pattern_lh_cat, int wsl = 4);    // phylotree.cpp:1108-1160; returns entries per pattern
    /* NNI evaluation (phylotree.cpp:2873-3066): both NNIs around the internal branch (node1, node2), each with its 1 (nni5 = false) or 5
       surrounding branches re-optimised by Newton-Raphson (NNI_MAX_NR_STEP = 10, pllnni.cpp:21); the tree is left as it was */
    struct NNIMove { double newloglh; double newLen[5]; int node1_nei, node2_nei; /* ids of the swapped subtrees' roots */ };
    NNIMove getBestNNIForBran(Node *node1, Node *node2, NNIMove moves[2], bool nni5 = true);
    uint64_t nni_queued_calls, nni_queued_fallbacks;    // how often the queued form ran / had to be redone call by call
    /* the consumers of _pattern_lh_cat and _pattern_lh on the device */
    double optimizeMixtureWeights();                                 // ModelMixture::optimizeWeights, model/modelmixture.cpp:1356-1416
    const std::vector<double> &mixtureWeights() const { return mixprops_; }
    int computePatternRates(double *pattern_rates, int32_t *pattern_cat);   // RateGamma::computePatternRates, model/rategamma.cpp:235-257
    void setBootSamples(const float *samples, int nsamples);         // boot_samples, iqtree.cpp:300-340
    void computeRell(double *rell);                                  // IQTree::saveCurrentTree's dot products, iqtree.cpp:2689-2735
    void computeFuncDerv(double value, double &df, double &ddf);
    double minimizeNewton(double x1, double xguess, double x2, double xacc, double &d2l, int maxNRStep = 100);
    void optimizeOneBranch(Node *node1, Node *node2, bool clearLH = true, int maxNRStep = 100);
    double optimizeAllBranches(int my_iterations = 100, double tolerance = 0.001, int maxNRStep = 100);
    Node *findFarthestLeaf(Node *node = NULL, Node *dad = NULL);
    void getPreOrderBranches(std::vector<Node*> &nodes, std::vector<Node*> &nodes2, Node *node, Node *dad = NULL);
    void computeBestTraversal(std::vector<Node*> &nodes, std::vector<Node*> &nodes2);
    void saveBranchLengths(std::vector<double> &lenvec, int startid = 0, Node *node = NULL, Node *dad = NULL);
    void restoreBranchLengths(const std::vector<double> &lenvec, int startid = 0, Node *node = NULL, Node *dad = NULL);
    void clearReversePartialLh(Node *node, Node *dad);

    /* ---- RateFree::optimizeWithEM (model/ratefree.cpp:450-578; lk_rate_em.cpp) ---- */
    // EM over the +R proportions and rates; copy_precision = MTree::num_precision of the reference's tree copy (mtree.cpp:36-39:
    // 6, or ceil(-log10(min_branch_length)) + 1 if that is larger and min_branch_length was set before the tree was made)
    double optimizeRateFreeEM(int copy_precision = 6);
    const std::vector<double> &categoryRates() const { return rates_; }
    const std::vector<double> &categoryProps() const { return props_; }
    std::string printTreeFixed(int precision) const;                 // MTree::printTree(out, WT_BR_LEN), mtree.cpp:245-379
    void scaleLength(double norm);                                   // mtree.cpp:1175-1186
    double optimizeTreeLengthScaling(double min_scaling, double &scaling, double max_scaling, double gradient_epsilon);   // phylotree.cpp:2106-2118
    double minimizeOneDimen(double xmin, double xguess, double xmax, double tolerance, double *fx, double *f2x);            // optimization.cpp:288-339

    /* ---- Fitch parsimony on the device (include/lkpars.h) ---- */
    // the informative sites in Alignment::ordered_pattern order, pattern p repeated frequency[p] times (alignment.cpp:731-775):
    // site_states[ntaxa][nsites]; aln_nsite = Alignment::getNSite() (all sites, for fixNegativeBranch)
    void setParsimonySites(int64_t nsites, const uint8_t *site_states, int64_t aln_nsite);
    void initializeAllPartialPars();                                 // phylotree.cpp:596-624
    void deleteAllPartialPars();
    void computePartialParsimony(Neighbor *dad_branch, Node *dad);   // phylotreepars.cpp:18-215: queues the dirty partials, post-order
    int computeParsimonyBranch(Neighbor *dad_branch, Node *dad, int *branch_subst = NULL);   // phylotreepars.cpp:218-282
    int computeParsimony();                                          // phylotree.cpp:658-660
    int fixNegativeBranch(bool force, Node *node = NULL, Node *dad = NULL);   // phylotree.cpp:2654-2694, all branches in one launch
    // phylotreepars.cpp:309-426; taxon_order = the shuffled addition order (my_random_shuffle stays with the caller's RNG).
    // Replaces the tree; returns best_pars_score.  All 2k-3 insertions of a round are scored by ONE launch pair.
    int computeParsimonyTree(const int32_t *taxon_order);
    void getBranches(std::vector<Node*> &nodes, std::vector<Node*> &nodes2, Node *node = NULL, Node *dad = NULL);   // mtree.cpp:905-919
    lkpars_engine *parsEngine() { return pars_; }
    int pars_device;             // CUDA ordinal of the parsimony engine (default: the first of setDevices, else 0)

    /* parameters (tools.cpp:785-786, phyloanalysis.cpp:2488-2497) */
    double min_branch_length, max_branch_length;
    bool device_newton;          // run minimizeNewton's loop on the device (single shard only)
    bool nni_queued;             // with device_newton: queue a whole NNI (5 branch optimisations + score) without waiting (device length slots)
    bool generic_kernel;         // -fastlk semantics: zero guard + _pattern_lh_cat in every branch call

    Neighbor *current_it, *current_it_back;
    bool theta_computed;
    uint64_t num_partial_lh_computations;

    /* data exposed for tests */
    int nstates, ncat, nmix, family, state_unknown, seq_type, ntaxa;
    int64_t nptn, nunobs, nsite;
    std::vector<double> ptn_freq, ptn_invar, tip_partial_lh;
    std::vector<std::string> seq_names;

private:
    double computeFunctionScaling(double value);                     // PhyloTree::computeFunction with is_opt_scaling, phylotree.cpp:2125-2132
    double brentOpt(double ax, double bx, double cx, double tol, double *foptx, double *f2optx, double fax, double fbx, double fcx);
    double current_scaling_;
    bool keep_engines_;                                              // readTreeString keeps the engines (same taxa, same model shape)

    /* parsimony bookkeeping (lk_parsimony.cpp) */
    int parsVec(const Neighbor *nei) const;                          // the engine vector a neighbour's partial_pars lives in
    int parsQueue(int kind, int dst, int a, int b, int c);
    void parsFlush();                                                // one lkpars_run of everything queued; resolves the subtree scores
    void parsCheck(int rc, const char *what);
    void freeNodes();
    lkpars_engine *pars_;
    std::vector<uint8_t> pars_states_;
    int64_t pars_nsites_, pars_aln_nsite_;
    int pars_nvec_;
    std::vector<uint32_t> pars_score_;                               // per engine vector: partial_pars[scoreid]
    std::vector<lkpars_op> pars_ops_;
    std::vector<uint32_t> pars_counts_;                              // [op][2] of the last flush

    struct BranchRequest { int32_t dad_end, node_vec; double len; };
    bool nni_async_;             // inside a queued NNI: optimizeOneBranch queues, lengths live in slots
    int nni_next_slot_;
    std::vector<Neighbor*> nni_touched_;
    void flushPartialOps(const BranchRequest *branch = NULL);
    void enqueuePatternLhCat();
    int collapsePatternLhCat(int wsl);
    void check(int rc, lkcuda_engine *e, const char *what);
    void reduceResult(double *result);
    Node *newNode();
    void parseNewick(const std::string &s, size_t &pos, Node *&node, double &branch_len);
    void initializeTreeIds(Node *node, Node *dad);
    void printSubtree(std::string &out, const Node *node, const Node *dad, int precision) const;
    int assignVectors(Node *node, Node *dad, int index);
    int opLevel(Neighbor *child);
    void ensureInputs();

    std::vector<uint8_t> states_;          // [ntaxa][nptn] global
    std::vector<uint8_t> unobs_;
    std::vector<int32_t> const_char_, ptn_model_;
    std::vector<double> freq_global_;
    std::vector<double> eval_, evec_, inv_evec_, rates_, props_, mixprops_, state_freq_;
    double p_invar_;
    std::vector<int> devices_;
    int proc_rank_, proc_world_;
    AllReduceFn allreduce_;
    void *allreduce_user_;
    void *stream_;
    int boot_nsamples_;
    bool peers_attached_;                  // the engines' reduction kernels already return sums over all shards / ranks
    std::vector<lkcuda_engine*> engines_;
    std::vector<int64_t> shard_begin_;     // per engine, global pattern offsets (size engines+1)
    int64_t ptn_begin_, ptn_end_;
    bool tip_partial_lh_computed_, weights_uploaded_, model_uploaded_;
    bool freq_dirty_, invar_dirty_, model_dirty_;   // which host inputs changed since they were last derived / uploaded
    int nvec_;                             // vectors of the tree; NNI_VECS more follow them (nni_partial_lh, phylotree.cpp:852-860)
    std::vector<lkcuda_partial_op> pending_ops_;
    std::vector<int> pending_level_;
    std::vector<int> vec_level_;           // level of the pending op writing each vector (0 = none)
};

}  // namespace lk
#endif
