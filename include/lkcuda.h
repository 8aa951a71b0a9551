/*
 * lkcuda.h — C ABI of the B200-native Felsenstein-pruning likelihood engine.
 *
 * This is the drop-in boundary for IQ-TREE 1.4.3's likelihood kernels.  The reference has no plugin
 * ABI; its seam is four PhyloTree member-function pointers plus one allocator routine
 * (phylotree.h:658-659, 697-698, 742-743, 979-980, 587-600), bound in
 * PhyloTree::setLikelihoodKernel (phylotreesse.cpp:60-311) and called through the wrappers at
 * phylotreesse.cpp:335-357.  Each entry point below names the reference function it replaces.
 * INTEGRATION.md shows the shim a maintainer adds on the reference side.
 *
 * Conventions
 *   - plain C, no C++/torch types; all functions return an int status (0 = LKCUDA_OK) and never throw;
 *     lkcuda_last_error() gives the text the caller hands to outError() (tools.cpp:99-106).
 *   - one opaque engine per PhyloTree (per partition under PhyloSuperTree); engines are independent and
 *     may be driven concurrently from different host threads (phylosupertree.cpp:969-976).
 *   - an engine owns ONE pattern shard [its P observed patterns (+ the U unobserved +ASC patterns,
 *     replicated)] on ONE device.  Every returned sum (lnL, df, ddf) is additive over shards: the
 *     caller adds the shards' values (host sum, or an NCCL all-reduce on lkcuda_result_device_ptr()).
 *   - "vector" v = one partial-likelihood buffer of the reference's central_partial_lh arena
 *     (phylotree.cpp:863-974: central_partial_lh + indexlh*block_size) together with its scale_num
 *     row (central_scale_num + indexlh*scale_block_size).  Layout inside a vector is the reference's:
 *     partial_lh[ptn][blk][state] doubles (blk = category, or mixture*ncat + category) and
 *     scale_num[ptn] as int16 (UBYTE = short, phylonode.h:17).
 *   - a child/end of a branch is either a vector index (>= 0) or a tip: LKCUDA_TIP(taxon_id) < 0.
 *   - all arithmetic is FP64.
 */
#ifndef LKCUDA_H
#define LKCUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LKCUDA_ABI_VERSION 1

enum {
    LKCUDA_OK = 0,
    LKCUDA_ERR_ARG = 1,      /* invalid argument                                   */
    LKCUDA_ERR_CUDA = 2,     /* a CUDA runtime call or kernel failed               */
    LKCUDA_ERR_STATE = 3,    /* call order violated (e.g. derv before theta)       */
    LKCUDA_ERR_NOMEM = 4,    /* host or device allocation failed                   */
    LKCUDA_ERR_NODEVICE = 5  /* no CUDA device / extension built without a device  */
};

/* flags of lkcuda_config.flags */
enum {
    /* scaling test is `lh_max < 2^-256` only — the mixture and fused mix-rate families
       (phylokernelmixture.h:356,439; phylokernelmixrate.h:344,428).  Without it the plain family's
       `&& ptn_invar[ptn] == 0.0` condition applies (phylokernel.h:381,463). */
    LKCUDA_FLAG_SCALE_IGNORES_INVAR = 1,
    /* the generic (non-SIMD) kernels' extra `lh_max == 0.0` branch: refill the block with
       tip_partial_lh[STATE_UNKNOWN], scale_num += 4 (phylotreesse.cpp:777-789, 895-907, 976-988) */
    LKCUDA_FLAG_ZERO_GUARD = 2,
    /* also fill _pattern_lh_cat in every branch-lnL call (the generic kernels always do,
       phylotreesse.cpp:1198,1225-1237; the SIMD kernels never do) */
    LKCUDA_FLAG_PATTERN_LH_CAT = 4,
    /* site-specific frequency models (-fs / PMSF; phylokernelsitemodel.cpp): every pattern has its own eigensystem
       (ModelSet, models->at(ptn)).  neigen = number of DISTINCT site models, lkcuda_set_site_models maps patterns to them; blocks are
       the rate categories; the scaling test has no ptn_invar condition and the lh_max == 0 branch refills with taxon 0's tip vector
       (phylokernelsitemodel.cpp:306-335, 404-430).  nstates <= 32. */
    LKCUDA_FLAG_SITEMODEL = 8
};

#define LKCUDA_TIP(taxon) (-1 - (int32_t)(taxon))
#define LKCUDA_IS_TIP(id) ((id) < 0)
#define LKCUDA_TIP_TAXON(id) (-1 - (id))

typedef struct lkcuda_engine lkcuda_engine;

typedef struct lkcuda_config {
    int32_t device;         /* CUDA device ordinal                                                    */
    int32_t nstates;        /* S = aln->num_states (4 DNA, 20 AA, 64 codon, any other -> generic path) */
    int32_t nblocks;        /* K = blocks per pattern: ncat (plain, fused mix-rate) or nmix*ncat       */
    int32_t neigen;         /* E = eigensystems: 1 (plain), ncat (fused), nmix (mixture)               */
    int32_t ntaxa;          /* leaves = rows of the tip-state matrix                                   */
    int32_t state_unknown;  /* aln->STATE_UNKNOWN (18 DNA, 23 AA, S otherwise; alignment.cpp:470-472)  */
    int32_t nvec;           /* vectors in the arena (central + nni buffers, phylotree.cpp:852-894)     */
    int32_t flags;          /* LKCUDA_FLAG_*                                                           */
    int64_t nptn;           /* P = observed patterns of this shard                                     */
    int64_t nunobs;         /* U = model_factory->unobserved_ptns.size() (+ASC), 0 otherwise           */
    int64_t nsite;          /* sites represented by this shard (= sum of its ptn_freq); +ASC term only */
} lkcuda_config;

/* one node of the post-order schedule: replaces one invocation of
   PhyloTree::computePartialLikelihoodEigenSIMD (phylokernel.h:71-483) and its mixture / mix-rate /
   generic twins (phylokernelmixture.h:21, phylokernelmixrate.h:23, phylotreesse.cpp:581/1350/2039)
   for a bifurcating node.  The caller has already ordered a leaf child first (phylokernel.h:116-121). */
typedef struct lkcuda_partial_op {
    int32_t dst;            /* vector receiving L_dad                          */
    int32_t left;           /* vector (>=0) or LKCUDA_TIP(taxon)               */
    int32_t right;          /* vector (>=0) or LKCUDA_TIP(taxon)               */
    int32_t reserved;       /* 0, or LKCUDA_LEN_REFS(left_slot, right_slot): take a child's branch length from a device length slot */
    double  len_left;       /* left->length                                    */
    double  len_right;      /* right->length                                   */
} lkcuda_partial_op;

/* ---- device length slots ----
   PhyloTree::getBestNNIForBran (phylotree.cpp:2873-3066) optimises five branches per NNI, and every optimised length feeds the next node
   update: driven call by call, the host waits for each Newton run before it can state the next op's lengths.  A length slot breaks that
   dependency: lkcuda_theta_newton_async leaves the optimised length in slot s ON THE DEVICE, a later op names the slot instead of a value
   (LKCUDA_LEN_REFS), a later Newton run may start from it, lkcuda_lnl_from_buffer_slot evaluates at it - the whole NNI is queued
   without a host wait and lkcuda_len_slots_fetch returns all lengths at the end.  Slot -1 = none. */
#define LKCUDA_NLENSLOT 64
#define LKCUDA_LEN_REFS(left_slot, right_slot) ((int32_t)((((left_slot) + 1) & 0xffff) | ((((right_slot) + 1) & 0xffff) << 16)))

/* results of the reductions, all additive over shards except [LKCUDA_R_PROB_CONST] */
enum {
    LKCUDA_R_LNL = 0,        /* tree_lh of the shard (scale factors and +ASC term included)  */
    LKCUDA_R_DF = 1,
    LKCUDA_R_DDF = 2,
    LKCUDA_R_PROB_CONST = 3, /* +ASC: sum of unobserved-pattern likelihoods (replicated)     */
    LKCUDA_R_NREPAIRED = 4,  /* patterns whose log-lh was NaN/Inf and got the repair value   */
    LKCUDA_R_SCALE_SUM = 5,  /* lh_scale_factor(node)+lh_scale_factor(dad) of the shard      */
    LKCUDA_NRESULT = 8
};
enum { LKCUDA_MAX_PEERS = 16, LKCUDA_PEER_HANDLE_BYTES = 64 };

/* ---- life cycle : PhyloTree::initializeAllPartialLh / deleteAllPartialLh (phylotree.cpp:667-987, 718) ---- */
int lkcuda_device_count(int32_t *count);
int lkcuda_create(const lkcuda_config *cfg, lkcuda_engine **out);
int lkcuda_destroy(lkcuda_engine *e);
const char *lkcuda_last_error(const lkcuda_engine *e);   /* e may be NULL: error of the last failed create */
/* Caller-owned arenas.  PhyloSuperTreePlen::initializeAllPartialLh (phylosupertreeplen.cpp:1742-1871) sub-allocates all partitions'
   vectors from ONE central_partial_lh / central_scale_num with per-partition strides block_size[part] / scale_block_size[part]
   (:1766-1768, 1834-1855), and ModelMixture::optimizeWithEM aliases a tree's arena into a temporary tree and NULLs it before
   delete (model/modelmixture.cpp:1427-1429, 1524-1526).  lkcuda_device_malloc gives the shim (host C++, no CUDA headers) zeroed
   device memory; lkcuda_attach_arena makes the engine use `nvec` vectors at partial_lh + i*vec_stride (doubles) and
   scale_num + i*scale_stride (int16) instead of its own arena - create the engine with cfg.nvec = 0 then.  The memory is
   BORROWED: lkcuda_destroy never frees it.  Passing NULL, NULL detaches.  Requirements: 32-byte aligned base, vec_stride a
   multiple of 4 doubles and >= (P+U)*K*S, scale_stride >= P+U.  lkcuda_arena_layout returns the strides the engine uses for its
   own arena (what PhyloTree::initializeAllPartialLh's block_size / scale_block_size become, phylotree.cpp:845, 863-914). */
int lkcuda_device_malloc(int32_t device, uint64_t bytes, void **ptr);
int lkcuda_device_free(int32_t device, void *ptr);
/* Page-lock a HOST array the caller keeps for the engine's life time - the reference's ptn_freq / ptn_invar / _pattern_lh, allocated once
   in PhyloTree::initializeAllPartialLh (phylotree.cpp:683-693) and re-sent by lkcuda_set_pattern_weights whenever clearAllPartialLH reset
   tip_partial_lh_computed (phylotree.cpp:499, phylotreesse.cpp:359-372): copies from / to a registered array run at PCIe speed instead of
   through the driver's staging buffer (16 MB at 1M patterns: ~0.3 ms instead of ~1.2 ms per evaluation).  Purely an optimisation: the
   calls that use the array behave the same without it.  Unregister before freeing the array. */
int lkcuda_host_register(lkcuda_engine *e, void *ptr, uint64_t bytes);
int lkcuda_host_unregister(lkcuda_engine *e, void *ptr);
int lkcuda_arena_layout(const lkcuda_engine *e, int64_t *vec_stride, int64_t *scale_stride);
int lkcuda_attach_arena(lkcuda_engine *e, void *partial_lh, int64_t vec_stride, void *scale_num, int64_t scale_stride, int32_t nvec);
/* bytes of device memory the engine holds (getMemoryRequired analogue, phylotree.cpp:995-1024) */
int lkcuda_memory_bytes(const lkcuda_engine *e, uint64_t *bytes);
/* run on the caller's CUDA stream (cudaStream_t as void*); NULL = the engine's own stream */
int lkcuda_set_stream(lkcuda_engine *e, void *cuda_stream);
int lkcuda_synchronize(lkcuda_engine *e);

/* ---- inputs ---- */
/* states[taxon][ptn], ptn < P+U: aln->at(ptn)[taxon] then model_factory->unobserved_ptns (phylokernel.h:237-244) */
int lkcuda_set_tip_states(lkcuda_engine *e, const uint8_t *states);
/* the same upload, pipelined: returns as soon as the copies are queued.  `states` must be page-locked host memory and stay valid
   until the next synchronising call.  The matrix goes up in `nchunks` pattern ranges on a copy stream; the DNA traversal kernel
   starts on the first range while the rest is still in flight (its producer warp polls a per-chunk counter), every other
   kernel that reads tip states is ordered after the whole upload.  A state out of range is reported by the next
   lkcuda_read_result / lkcuda_synchronize / lkcuda_fetch_* (LKCUDA_ERR_ARG), the engine computes on clamped states until then. */
int lkcuda_set_tip_states_async(lkcuda_engine *e, const uint8_t *states, int32_t nchunks);
/* PhyloTree::computePtnFreq / computePtnInvar outputs (phylotreesse.cpp:531-569): ptn_freq[P], ptn_invar[P+U] */
int lkcuda_set_pattern_weights(lkcuda_engine *e, const double *ptn_freq, const double *ptn_invar);
/* model->getEigenvalues/getEigenvectors/getInverseEigenvectors (model/modelsubst.h:248-258; mixtures
   concatenated, model/modelmixture.cpp:1239-1262): eval[E*S], evec[E*S*S], inv_evec[E*S*S];
   per block: its eigensystem, site_rate->getRate(c), and the weight site_rate->getProp(c) (* prop[m]);
   tip_partial_lh[(state_unknown+1)][E][S] from computeTipPartialLikelihood (phylotreesse.cpp:359-529). */
int lkcuda_set_model(lkcuda_engine *e, const double *eval, const double *evec, const double *inv_evec,
                     const int32_t *blk_eigen, const double *blk_rate, const double *blk_weight,
                     const double *tip_partial_lh);

/* ---- partial likelihoods: PhyloTree::computePartialLikelihood (phylotreesse.cpp:335) ---- */
/* ops are in post-order; level_start[nlevels+1] partitions them into groups of mutually independent
   nodes (one launch per level).  Asynchronous: enqueues on the engine stream and returns. */
int lkcuda_partial_batch(lkcuda_engine *e, const lkcuda_partial_op *ops, int32_t nops,
                         const int32_t *level_start, int32_t nlevels);
/* PhyloTree::computeLikelihood (phylotree.cpp:1031-1071) in one call: the traversal towards the root branch (lkcuda_partial_batch) and
   computeLikelihoodBranch on that branch (lkcuda_branch_lnl; dad = LKCUDA_TIP(taxon) or a vector, node = the vector of the branch's
   other end).  Same results and side effects (_pattern_lh, result block) as the two calls; when the list ends in `node`, `dad` is a
   leaf and the data are DNA with 4 categories the branch is evaluated in the tail of the traversal launch, from registers. */
int lkcuda_partial_batch_lnl(lkcuda_engine *e, const lkcuda_partial_op *ops, int32_t nops,
                             const int32_t *level_start, int32_t nlevels,
                             int32_t dad, int32_t node, double len, double *result);
/* multifurcating node (degree > 3) — the generic kernel's path (phylotreesse.cpp:702-806) */
int lkcuda_partial_multi(lkcuda_engine *e, int32_t dst, int32_t nchild, const int32_t *child,
                         const double *child_len);

/* ---- branch log-likelihood: PhyloTree::computeLikelihoodBranch (phylotreesse.cpp:339) ----
   dad = vector or tip, node = vector (caller swapped so that only dad may be a leaf, phylokernel.h:739-746);
   writes _pattern_lh (and _pattern_lh_cat with LKCUDA_FLAG_PATTERN_LH_CAT) on the device.
   result == NULL: enqueue only (read later with lkcuda_read_result); else synchronises and fills
   result[LKCUDA_NRESULT]. */
int lkcuda_branch_lnl(lkcuda_engine *e, int32_t dad, int32_t node, double len, double *result);
/* fill only _pattern_lh_cat[p][k] for that branch: what PhyloTree::computePatternLhCat (phylotree.cpp:1108-1124)
   obtains by calling the generic branch kernels (phylotreesse.cpp:1194-1237, 2588-2635).  Asynchronous. */
int lkcuda_pattern_lh_cat(lkcuda_engine *e, int32_t dad, int32_t node, double len);

/* ---- derivatives: PhyloTree::computeLikelihoodDerv (phylotreesse.cpp:344) split at its theta_computed flag ---- */
/* theta_all = L_node * L_dad (phylokernel.h:535-579) */
int lkcuda_theta_build(lkcuda_engine *e, int32_t dad, int32_t node);
/* both halves of one PhyloTree::computeLikelihoodDerv call with theta_computed == false (phylokernel.h:535-635): builds theta_all
   AND returns df, ddf at `len` in the same pass over the two vectors (one read of theta_all saved per branch) */
int lkcuda_theta_derv(lkcuda_engine *e, int32_t dad, int32_t node, double len, double *result);
/* df, ddf at branch length len from theta_all (phylokernel.h:521-532, 599-725) */
int lkcuda_derv(lkcuda_engine *e, double len, double *result);
/* lkcuda_derv for callers that must not allocate (code under CUDA-graph capture, tight serving loops, harnesses that forbid
   cudaMalloc).  NO COUNTERPART IN THE REFERENCE; the arithmetic is lkcuda_derv's
   (phylokernel.h:606-651) without +ASC and for one eigen system.  It needs no engine and calls no cudaMalloc / cudaFree / allocating
   library routine: the caller passes every input and the output as DEVICE pointers - theta_all[nptn][ncat][nstates], ptn_freq[nptn],
   ptn_invar[nptn], df_ddf[2] = {df, ddf} - and the model as small HOST arrays - eval[nstates], cat_rate[ncat], cat_prop[ncat] (copied
   into the kernel's parameters).  The cross-block scratch is a static device array of LKCUDA_NOALLOC_SCRATCH_BYTES that every GPU
   loading the library holds until the process exits; hence the limits: nptn <= LKCUDA_NOALLOC_MAX_PATTERNS, nstates <= 64,
   ncat <= 64, nstates * ncat <= LKCUDA_NOALLOC_MAX_BLOCK - anything larger is refused with LKCUDA_ERR_ARG (text:
   lkcuda_last_error(NULL)).  Asynchronous on `cuda_stream` (NULL = the default stream): df_ddf is valid once the stream has reached the
   call.  All calls on a device share the one scratch array, so calls that OVERLAP on the GPU (different streams or host threads without
   an ordering between them) race; calls ordered on one stream do not.  Same inputs give the same bits on every call. */
#define LKCUDA_NOALLOC_MAX_PATTERNS (1 << 21)
#define LKCUDA_NOALLOC_MAX_BLOCK 256
#define LKCUDA_NOALLOC_SCRATCH_BYTES (65536 + 256)
int lkcuda_derv_noalloc(int32_t device, void *cuda_stream, const double *theta_all, const double *ptn_freq, const double *ptn_invar,
                        int64_t nptn, int32_t nstates, int32_t ncat, const double *eval, const double *cat_rate, const double *cat_prop,
                        double len, double *df_ddf);
/* PhyloTree::computeLikelihoodFromBuffer (phylotreesse.cpp:349; phylokernel.h:1023-1192) */
int lkcuda_lnl_from_buffer(lkcuda_engine *e, double len, double *result);
/* Optimization::minimizeNewton (optimization.cpp:388-465) driven on the device over theta_all:
   same step sequence, one launch instead of one launch + one read-back per Newton step.
   out[4]: [0] = optx, [1] = d2l, [2] = number of computeFuncDerv evaluations, [3] = the last length
   evaluated (where the reference leaves current_it->length).  One persistent cooperative launch for every block shape
   (site-specific models: host-driven steps); with peers attached every evaluation is exchanged inside the loop. */
int lkcuda_newton_branch(lkcuda_engine *e, double x1, double xguess, double x2, double xacc,
                         int32_t max_nr_step, double *out);
/* lkcuda_theta_newton without the wait: theta_all of (dad, node) and minimizeNewton are queued; the start is `xguess`, or the content of
   slot `xguess_slot` when that is >= 0; the run leaves optx in slot `len_slot` and {optx, d2l, evaluations, status} beside it
   (status 1 = converged, 2 / 3 / 4 = the errors lkcuda_theta_newton reports).  One engine without peers, shapes that have a
   persistent Newton kernel; returns LKCUDA_ERR_STATE otherwise (the caller then uses the synchronous calls). */
int lkcuda_theta_newton_async(lkcuda_engine *e, int32_t dad, int32_t node, double x1, double xguess, int32_t xguess_slot, double x2,
                              double xacc, int32_t maxNRStep, int32_t len_slot);
/* lkcuda_lnl_from_buffer at the length held by a slot */
int lkcuda_lnl_from_buffer_slot(lkcuda_engine *e, int32_t len_slot, double *result);
/* set slots from the host (n values starting at slot `first`), and read all of them back after the queued work has finished:
   lens[LKCUDA_NLENSLOT], info[LKCUDA_NLENSLOT][4] = {optx, d2l, evaluations, status} (either may be NULL) */
int lkcuda_len_slots_set(lkcuda_engine *e, int32_t first, int32_t n, const double *values);
int lkcuda_len_slots_fetch(lkcuda_engine *e, double *lens, double *info);

/* lkcuda_theta_build + lkcuda_newton_branch on the branch (dad, node) in one call - for DNA with 4 categories ONE persistent launch that
   builds theta_all in its first Newton step, for other shapes theta_kernel + one persistent launch (optimizeOneBranch's device part, phylotree.cpp:2148-2170); same out[4] */
int lkcuda_theta_newton(lkcuda_engine *e, int32_t dad, int32_t node, double x1, double xguess, double x2, double xacc,
                        int32_t max_nr_step, double *out);

/* results of the last enqueued reduction (synchronises) / their device address (for NCCL) */
int lkcuda_read_result(lkcuda_engine *e, double *result);
int lkcuda_result_device_ptr(lkcuda_engine *e, void **dev_ptr);
/* keep the result block in caller-owned device memory (LKCUDA_NRESULT doubles, e.g. a buffer registered with NCCL or a
   torch tensor) so that the caller's all-reduce runs in place on it; NULL = back to the engine's own block */
int lkcuda_set_result_buffer(lkcuda_engine *e, void *dev_ptr);
/* Results of the last enqueued reduction WITHOUT a copy or a stream synchronisation: every reduction kernel ends by writing its
   result block and a sequence number into host-mapped memory, which this call polls (the wrappers at phylotreesse.cpp:335-357 return
   a double, so each evaluation ends in exactly one such wait).  With peers attached the block already holds the sums over all
   ranks.  lkcuda_read_result stays the path to use after the caller changed the block behind the engine's back (NCCL in place). */
int lkcuda_wait_result(lkcuda_engine *e, double *result);

/* ---- fused reduction + cross-GPU all-reduce over NVLink peer memory (only the scalar lnL/df/ddf sums are exchanged) ----
   Each engine (= one pattern shard on one GPU, normally one process per GPU) owns a small mailbox in device memory.
   lkcuda_peer_handle exports it: `handle` receives a CUDA IPC handle (LKCUDA_PEER_HANDLE_BYTES) for engines in OTHER processes,
   `local_ptr` the raw device address for engines in the SAME process.  After the ranks exchanged them (any transport: MPI,
   torch.distributed, a pipe) lkcuda_attach_peers installs the table; from then on every reduction kernel (branch lnL, df/ddf, lnL from
   buffer) deposits its result block in all peers' mailboxes with P2P stores and sums the deposits in rank order before it exits, so
   that lkcuda_wait_result returns the GLOBAL sums on every rank (bit-identical across ranks).  All ranks must enqueue the same
   sequence of reductions (they do: tree, model and host control flow are replicated).  `handles`: [world][LKCUDA_PEER_HANDLE_BYTES]
   or NULL; `raw_ptrs`: [world] device addresses or NULL; entry `rank` of either is ignored.  A peer that never shows up makes the
   kernel give up after 10 s and the next lkcuda_wait_result fail with LKCUDA_ERR_STATE.
   CONSTRAINT: a reduction kernel WAITS, on the device, for its peers' kernels - the peers' kernels must be able to run while it
   waits.  One process per GPU satisfies that by construction.  Engines of ONE process on one device (several shards driven by one
   host thread) need the device to run their streams concurrently and the host to queue every shard's preceding work before the
   first reduction is launched (harness/lk_phylotree.cpp does): under CUDA_LAUNCH_BLOCKING=1, compute-sanitizer or ncu (which
   serialise kernels) such a set-up runs into the 10 s time-out; use lkcuda_detach_peers and add the shards' blocks on the host there. */
int lkcuda_peer_handle(lkcuda_engine *e, uint8_t *handle /*[LKCUDA_PEER_HANDLE_BYTES] or NULL*/, void **local_ptr /*or NULL*/);
int lkcuda_attach_peers(lkcuda_engine *e, int32_t rank, int32_t world, const uint8_t *handles, void *const *raw_ptrs);
int lkcuda_detach_peers(lkcuda_engine *e);

/* ---- host-visible mirrors (buffers host code reads or writes) ---- */
int lkcuda_fetch_pattern_lh(lkcuda_engine *e, double *pattern_lh /*[P]*/);
int lkcuda_fetch_pattern_lh_cat(lkcuda_engine *e, double *pattern_lh_cat /*[P][K]*/);
/* the tail of PhyloTree::computeLikelihood(double *pattern_lh) (phylotree.cpp:1055-1069) done on the device:
   dst[p] = _pattern_lh[p] + max(scale_num(vec)[p], 0) * LOG_SCALING_THRESHOLD; vec < 0 (a tip end): plain copy.
   _pattern_lh itself is left untouched. */
/* the same without waiting: the correction runs on the engine's stream, the copy to `pattern_lh` (pinned host memory) on a copy stream,
   so that the next kernels overlap it; lkcuda_fetch_wait returns when the data has arrived (one fetch in flight per engine) */
int lkcuda_fetch_pattern_lh_corrected_async(lkcuda_engine *e, int32_t vec, double *pattern_lh /*[P], pinned*/);
int lkcuda_fetch_wait(lkcuda_engine *e);
int lkcuda_fetch_pattern_lh_corrected(lkcuda_engine *e, int32_t vec, double *pattern_lh /*[P]*/);
int lkcuda_fetch_vector(lkcuda_engine *e, int32_t vec, double *partial_lh /*[(P+U)*K*S] or NULL*/,
                        int16_t *scale_num /*[P+U] or NULL*/);
int lkcuda_store_vector(lkcuda_engine *e, int32_t vec, const double *partial_lh, const int16_t *scale_num);
int lkcuda_fetch_theta(lkcuda_engine *e, double *theta /*[(P+U)*K*S]*/);
/* lh_scale_factor of a vector = sum_p LOG_SCALING_THRESHOLD * ptn_freq[p] * scale_num[p] (phylokernel.h:389,395) */
int lkcuda_vector_scale_factor(lkcuda_engine *e, int32_t vec, double *lh_scale_factor);
/* raw device addresses of a vector (PhyloNeighbor::partial_lh / scale_num become these) */
int lkcuda_vector_device_ptr(lkcuda_engine *e, int32_t vec, void **partial_lh, void **scale_num);

/* ---- site-specific frequency models (LKCUDA_FLAG_SITEMODEL) ----
   ptn_model[p] = index of pattern p's eigensystem among the `neigen` given to lkcuda_set_model (ModelSet::pattern_model_map,
   model/modelfactory.cpp:316-320); state_mask[s] = bit set of the plain states tip code s stands for (bit x set <=> state x
   contributes): 1 << s for s < nstates, all ones for STATE_UNKNOWN, the ambiguity sets of phylotreesse.cpp:398-440 otherwise.  The
   per-taxon, per-pattern tip vectors the reference precomputes (tip_partial_lh[taxon][ptn][S], phylotreesse.cpp:373-456) are
   formed on the fly as sum over the set of inv_evec[i][x].  With this flag lkcuda_set_model ignores blk_eigen and tip_partial_lh. */
int lkcuda_set_site_models(lkcuda_engine *e, const int32_t *ptn_model /*[P+U]*/, const uint64_t *state_mask /*[state_unknown+1]*/);

/* ---- UFBoot RELL (IQTree::saveCurrentTree, iqtree.cpp:2676-2786; dotProductSIMD, phylokernel.h:55-61) ----
   lkcuda_set_boot_samples: boot_samples[s][ptn] (float pattern frequencies of bootstrap replicate s, iqtree.cpp:300-340) of THIS
   shard's patterns, row s at samples + s*stride; kept in HBM for the whole search, stored losslessly as uint8 / uint16 when all
   entries are small non-negative integers (lkcuda_boot_storage tells which).
   lkcuda_rell: rell[s] = sum_ptn boot_samples[s][ptn] * (float)ptn_lh[ptn], ptn_lh = _pattern_lh of the last branch-lnL plus the
   scale_num correction of computePatternLikelihood (phylotree.cpp:1213-1259) taken from vectors vec_a / vec_b (-1 = a tip end).
   Additive over shards; lkcuda_rell_device_ptr gives the device copy (double[nsamples]) for an NCCL sum.  rell may be NULL. */
int lkcuda_set_boot_samples(lkcuda_engine *e, const float *samples, int32_t nsamples, int64_t stride);
int lkcuda_boot_storage(lkcuda_engine *e, int32_t *bytes_per_entry, uint64_t *bytes);
int lkcuda_rell(lkcuda_engine *e, int32_t vec_a, int32_t vec_b, double *rell /*[nsamples]*/);
int lkcuda_rell_fetch(lkcuda_engine *e, double *rell /*[nsamples]*/);      /* the sums of the last lkcuda_rell (synchronises) */
int lkcuda_rell_device_ptr(lkcuda_engine *e, void **dev_ptr);

/* ---- consumers of _pattern_lh_cat (filled by lkcuda_pattern_lh_cat or a generic-kernel branch lnL) ----
   lkcuda_lh_cat_collapse: the tail of PhyloTree::computePatternLhCat (phylotree.cpp:1133-1158) into a work matrix w[P][ncol]:
     mode 0 = as is (ncol = ncat*nmix), 1 = per mixture class (WSL_MIXTURE, ncol = nmix), 2 = per rate category (WSL_RATECAT, ncol = ncat);
     blocks are ordered m*ncat + c (phylokernelmixture.h:121); for plain / fused models pass nmix = 1.
   lkcuda_lh_cat_em: one EM pass (ModelMixture::optimizeWeights, model/modelmixture.cpp:1369-1388; RateFree::optimizeWithEM,
     model/ratefree.cpp:499-520): w[p][c] *= ratio[c] (ratio may be NULL), sums[c] = sum_p w[p][c] * ptn_freq[p] / (ptn_invar[p] + sum_c w[p][c]).
     Additive over shards.  ncol <= 96.
   lkcuda_lh_cat_pattern_rates: RateGamma::computePatternRates (model/rategamma.cpp:235-257): posterior mean rate and most likely
     category of every pattern (exact ties: first maximum; the reference tosses a coin).  invar_class != 0: RateGammaInvar's variant
     (model/rategammainvar.cpp:193-221) - ptn_invar joins the denominator and competes as category 0, the others become c + 1. */
int lkcuda_lh_cat_collapse(lkcuda_engine *e, int32_t mode, int32_t ncat, int32_t nmix, int32_t *ncol);
int lkcuda_lh_cat_em(lkcuda_engine *e, const double *ratio /*[ncol] or NULL*/, double *sums /*[ncol]*/);
/* include_invar = 0: the pattern likelihood of RateFree::optimizeWithEM's E-step, which leaves ptn_invar out (ratefree.cpp:487-503, also under +I+R) */
int lkcuda_lh_cat_em_ex(lkcuda_engine *e, const double *ratio /*[ncol] or NULL*/, int32_t include_invar, double *sums /*[ncol]*/);
/* the E-step posterior of class `col` (w[p][col] * ptn_freq[p] / ([ptn_invar[p] +] sum_c w[p][c]): ratefree.cpp:487-503 without,
   modelmixture.cpp:1458-1471 with ptn_invar) becomes the
   ptn_freq of `dst`, an engine over the same patterns on the same device (the sub-tree of RateFree::optimizeWithEM, ratefree.cpp:533-549,
   and of ModelMixture::optimizeWithEM, modelmixture.cpp:1488-1511); nothing crosses PCIe */
int lkcuda_lh_cat_posterior_weights(lkcuda_engine *e, int32_t col, int32_t include_invar /* ModelMixture: 1, RateFree: 0 */, lkcuda_engine *dst);
int lkcuda_lh_cat_pattern_rates(lkcuda_engine *e, const double *rates /*[ncol]*/, int32_t invar_class, double *ptn_rates /*[P]*/, int32_t *ptn_cat /*[P]*/);
int lkcuda_fetch_lh_cat_collapsed(lkcuda_engine *e, double *w /*[P][ncol]*/);

/* ---- instrumentation ---- */
/* kernels launched by this engine since creation (bench.py's gpu_launches) */
int lkcuda_launch_count(const lkcuda_engine *e, uint64_t *count);
/* algorithmic bytes moved by those launches, for the roofline */
int lkcuda_algorithmic_bytes(const lkcuda_engine *e, uint64_t *bytes);
int lkcuda_reset_counters(lkcuda_engine *e);
/* per-kernel-kind statistics; with profiling on, every launch is bracketed by CUDA events on the engine
   stream and device_ms is their summed elapsed time (the reference's num_partial_lh_computations /
   getRealTime instrumentation, phylokernel.h:85, timeutil.h:134) */
enum {
    LKCUDA_KIND_PREP = 0, LKCUDA_KIND_PARTIAL = 1, LKCUDA_KIND_BRANCH = 2, LKCUDA_KIND_THETA = 3,
    LKCUDA_KIND_DERV = 4, LKCUDA_KIND_FROMBUF = 5, LKCUDA_KIND_NEWTON = 6, LKCUDA_KIND_MISC = 7, LKCUDA_KIND_RELL = 8,
    LKCUDA_KIND_EM = 9, LKCUDA_NKIND = 10
};
int lkcuda_set_profiling(lkcuda_engine *e, int32_t on);
int lkcuda_kernel_stats(lkcuda_engine *e, int32_t kind, uint64_t *launches, uint64_t *bytes, double *device_ms);
/* algorithmic FP64 operations of those launches (per pattern and block 2*S^2 for every internal child and for the way back
   to state space, S per product - 6S^2+S / 4S^2+S / 2S^2+S for the three node cases); counted for the partial-likelihood kernels:
   the numerator of the FP64 roofline */
int lkcuda_kernel_flops(lkcuda_engine *e, int32_t kind, uint64_t *flops);

#ifdef __cplusplus
}
#endif
#endif /* LKCUDA_H */
