/*
 * phylotree_cuda.cpp — the IQ-TREE side of the B200 likelihood engine.
 *
 * This file is added to IQ-TREE 1.4.3's source list next to phylotreesse.cpp (together with integration/iqtree-1.4.3-lkcuda.patch,
 * which declares the members below in phylotree.h and adds the `-cuda [device]` switch).  It binds PhyloTree's four kernel pointers
 * (phylotree.h:658-659, 697-698, 742-743, 979-980) and the buffer allocator (initializeAllPartialLh, phylotree.cpp:667-716, 834-987) to
 * the C ABI of include/lkcuda.h.  Everything else - tree search, model optimisation, ModelFinder, the Newton / Brent drivers, dirty bits -
 * stays the reference's own code and runs unchanged on top.
 *
 * What stays where:
 *   - PhyloNeighbor::partial_lh / scale_num hold DEVICE addresses, assigned by the reference's own recursion
 *     (phylotree.cpp:931-974) with its own block_size / scale_block_size arithmetic; both buffer modes work: LM_ALL_BRANCH and the
 *     default LM_PER_NODE with its on-demand re-orientation (phylonode.cpp:22-39, phylokernel.h:127-143).  The six NNI vectors
 *     (nni_partial_lh, phylotree.cpp:852-860) sit behind the central ones in the same device arena.
 *   - partial_lh_computed, length, the recursion order, the leaf-left swap and the re-orientation are executed here on the host exactly
 *     as the CPU kernels do; instead of computing, a node is RECORDED and the recorded post-order list is handed to the device
 *     in one call before the next reduction.
 *   - _pattern_lh, _pattern_lh_cat, ptn_freq, ptn_invar, tip_partial_lh stay host arrays (host code reads and writes them): inputs
 *     are uploaded when the reference recomputes them (tip_partial_lh_computed), outputs are fetched where host code reads them.
 *
 * `sse` keeps its meaning (LK_EIGEN: generic kernel semantics, i.e. the extra lh_max == 0 branch; LK_EIGEN_SSE: SIMD semantics), the CUDA
 * back end is an orthogonal switch (Params::lk_cuda) - ~40 places of the reference test `sse == LK_EIGEN || sse == LK_EIGEN_SSE`.
 */
#include "phylotree.h"
#include "model/modelfactory.h"
#include "model/modelmixture.h"
#include "model/rateheterogeneity.h"
#include "vectorclass/instrset.h"
#include "lkcuda.h"
#include <vector>
#include <string.h>

extern int instruction_set;     // pda.cpp

/* set around lkCudaBranchEnds by the branch-lnL wrapper: the recorded nodes are not flushed there but handed over together with the branch
   (thread_local: partition models evaluate different PhyloTree objects from several OpenMP threads, phylosupertree.cpp:969-976) */
static thread_local bool lk_defer_flush = false;

struct PhyloTree::LkCuda {
    lkcuda_engine *engine;
    int device;
    void *d_lh, *d_sc;                  // the device arena: central vectors, then the NNI vectors
    size_t block_size, scale_block_size;
    int nvec_central, nvec_total;
    double *tip_partial_lh;             // host (computeTipPartialLikelihood writes it)
    size_t tip_partial_lh_size;
    // shape the engine was created for
    int nstates, nblocks, neigen, ntaxa, flags, nodeNum, leafNum;
    size_t nptn, nunobs;
    LhMemSave mem_save;
    std::vector<lkcuda_partial_op> pending;
    bool weights_uploaded;
    std::vector<double> model_cache;    // what the engine currently holds (eval | evec | inv_evec | rate | weight | tip table)
    uint64_t nflush, nops;
    void *pinned[3];                    // ptn_freq, ptn_invar, _pattern_lh while they are page-locked (lkcuda_host_register)
};

static void lkCheck(int rc, lkcuda_engine *e, const char *what) {
    if (rc == LKCUDA_OK) return;
    string msg = string(what) + ": " + lkcuda_last_error(e);
    outError(msg.c_str());
}

bool PhyloTree::lkCudaWanted() {
    if (!Params::getInstance().lk_cuda || !aln || !model || !site_rate || !model_factory) return false;
    if (model->isSiteSpecificModel()) return false;     // -fs / PMSF stays on the CPU kernels in this binding
    return model->isReversible();                       // the eigen kernels need a reversible model (phylokernel.h:104)
}

/* dispatch: the reference's own choice first (it also fixes `sse` and the parsimony / dot-product kernels), then the four pointers */
void PhyloTree::setLikelihoodKernel(LikelihoodKernel lk) {
    setLikelihoodKernelCPU(lk);
    if (!Params::getInstance().lk_cuda || !aln) return;
    if (model_factory && model_factory->model && model_factory->model->isSiteSpecificModel()) return;
    computePartialLikelihoodPointer = &PhyloTree::computePartialLikelihoodCuda;
    computeLikelihoodBranchPointer = &PhyloTree::computeLikelihoodBranchCuda;
    computeLikelihoodDervPointer = &PhyloTree::computeLikelihoodDervCuda;
    computeLikelihoodFromBufferPointer = &PhyloTree::computeLikelihoodFromBufferCuda;
}

/* ------------------------------------------------------------ buffers ------------------------------------------------------------ */

void PhyloTree::lkCudaRelease() {
    if (!lk_cuda) return;
    // the reference frees ptn_freq / ptn_invar / _pattern_lh right after this call (phylotree.cpp:736-746)
    for (int i = 0; i < 3; i++)
        if (lk_cuda->engine && lk_cuda->pinned[i]) lkcuda_host_unregister(lk_cuda->engine, lk_cuda->pinned[i]);
    if (lk_cuda->engine) lkcuda_destroy(lk_cuda->engine);
    if (lk_cuda->d_lh) lkcuda_device_free(lk_cuda->device, lk_cuda->d_lh);
    if (lk_cuda->d_sc) lkcuda_device_free(lk_cuda->device, lk_cuda->d_sc);
    if (lk_cuda->tip_partial_lh) aligned_free(lk_cuda->tip_partial_lh);
    delete lk_cuda;
    lk_cuda = NULL;
    // the reference's own clean-up (phylotree.cpp:162-180, 720-734) would aligned_free() these
    central_partial_lh = NULL; nni_partial_lh = NULL;
    central_scale_num = NULL; nni_scale_num = NULL;
    tip_partial_lh = NULL;
}

/* called by initializeAllPartialLh() right before the reference's recursion: with the four arena pointers already set the recursion
   skips its own allocations (phylotree.cpp:852, 863, 896) and only hands out addresses */
void PhyloTree::lkCudaPrepareArena() {
    if (!lkCudaWanted()) {
        if (lk_cuda) lkCudaRelease();
        return;
    }
    const bool mixture = model->isMixture(), fused = model_factory->fused_mix_rate;
    const int S = aln->num_states, ncat = site_rate->getNRate(), nmix = model->getNMixtures();
    const int K = (mixture && !fused) ? ncat * nmix : ncat;
    const size_t P = aln->size(), U = model_factory->unobserved_ptns.size();
    const int ntaxa = aln->getNSeq();
    const int flags = (mixture ? LKCUDA_FLAG_SCALE_IGNORES_INVAR : 0)           // phylokernelmixture.h:356, phylokernelmixrate.h:344
                    | (sse == LK_EIGEN ? LKCUDA_FLAG_ZERO_GUARD : 0);           // the generic kernels' lh_max == 0 branch, phylotreesse.cpp:777-789
    const int nvec_central = params->lh_mem_save == LM_PER_NODE ? nodeNum - leafNum : (nodeNum - 1) * 2 - leafNum;    // phylotree.cpp:706-711
    const int nvec_total = nvec_central + (params->nni5 ? 6 : 2);                                                       // phylotree.cpp:857
    if (lk_cuda && (lk_cuda->nstates != S || lk_cuda->nblocks != K || lk_cuda->neigen != nmix || lk_cuda->ntaxa != ntaxa ||
                    lk_cuda->nptn != P || lk_cuda->nunobs != U || lk_cuda->flags != flags || lk_cuda->nodeNum != nodeNum ||
                    lk_cuda->leafNum != leafNum || lk_cuda->mem_save != params->lh_mem_save || lk_cuda->nvec_total != nvec_total))
        lkCudaRelease();        // the caller changed model or tree without deleteAllPartialLh(): start over
    if (!lk_cuda) {
        // block_size / scale_block_size exactly as the recursion computes them (phylotree.cpp:835-848)
        size_t nptn = P + S;
        if (instruction_set >= 7) nptn = ((nptn + 3) / 4) * 4;
        else nptn = ((nptn % 2) == 0) ? nptn : (nptn + 1);
        lk_cuda = new LkCuda;
        lk_cuda->engine = NULL; lk_cuda->d_lh = lk_cuda->d_sc = NULL; lk_cuda->tip_partial_lh = NULL;
        lk_cuda->device = Params::getInstance().cuda_device;
        lk_cuda->scale_block_size = nptn;
        lk_cuda->block_size = nptn * S * ncat * (fused ? 1 : nmix);
        lk_cuda->nvec_central = nvec_central; lk_cuda->nvec_total = nvec_total;
        lk_cuda->nstates = S; lk_cuda->nblocks = K; lk_cuda->neigen = nmix; lk_cuda->ntaxa = ntaxa; lk_cuda->flags = flags;
        lk_cuda->nptn = P; lk_cuda->nunobs = U; lk_cuda->nodeNum = nodeNum; lk_cuda->leafNum = leafNum; lk_cuda->mem_save = params->lh_mem_save;
        lk_cuda->weights_uploaded = false; lk_cuda->nflush = lk_cuda->nops = 0;
        lk_cuda->pinned[0] = lk_cuda->pinned[1] = lk_cuda->pinned[2] = NULL;
        lk_cuda->tip_partial_lh_size = (size_t)S * (aln->STATE_UNKNOWN + 1) * nmix;
        lk_cuda->tip_partial_lh = aligned_alloc<double>(lk_cuda->tip_partial_lh_size);

        lkcuda_config cfg;
        memset(&cfg, 0, sizeof(cfg));
        cfg.device = lk_cuda->device; cfg.nstates = S; cfg.nblocks = K; cfg.neigen = nmix; cfg.ntaxa = ntaxa;
        cfg.state_unknown = aln->STATE_UNKNOWN; cfg.nvec = 0; cfg.flags = flags;
        cfg.nptn = (int64_t)P; cfg.nunobs = (int64_t)U; cfg.nsite = aln->getNSite();
        if (lkcuda_create(&cfg, &lk_cuda->engine)) {
            string msg = string("lkcuda_create: ") + lkcuda_last_error(NULL);
            lkCudaRelease();
            outError(msg.c_str());
        }
        lkcuda_engine *e = lk_cuda->engine;
        lkCheck(lkcuda_device_malloc(cfg.device, (uint64_t)lk_cuda->nvec_total * lk_cuda->block_size * sizeof(double), &lk_cuda->d_lh), NULL, "lkcuda_device_malloc");
        lkCheck(lkcuda_device_malloc(cfg.device, (uint64_t)lk_cuda->nvec_total * lk_cuda->scale_block_size * sizeof(UBYTE), &lk_cuda->d_sc), NULL, "lkcuda_device_malloc");
        lkCheck(lkcuda_attach_arena(e, lk_cuda->d_lh, (int64_t)lk_cuda->block_size, lk_cuda->d_sc, (int64_t)lk_cuda->scale_block_size, lk_cuda->nvec_total),
                e, "lkcuda_attach_arena");
        // taxon-major tip states: aln->at(ptn)[taxon], then the +ASC patterns (phylokernel.h:237-244)
        const size_t NP = P + U;
        std::vector<uint8_t> st((size_t)ntaxa * NP + 1);
        for (size_t p = 0; p < P; p++) {
            const Pattern &pat = aln->at(p);
            for (int t = 0; t < ntaxa; t++) st[(size_t)t * NP + p] = (uint8_t)pat[t];
        }
        for (size_t u = 0; u < U; u++)
            for (int t = 0; t < ntaxa; t++) st[(size_t)t * NP + P + u] = (uint8_t)model_factory->unobserved_ptns[u];
        lkCheck(lkcuda_set_tip_states(e, &st[0]), e, "lkcuda_set_tip_states");
    }
    lk_cuda->pending.clear();
    central_partial_lh = (double*)lk_cuda->d_lh;
    central_scale_num = (UBYTE*)lk_cuda->d_sc;
    nni_partial_lh = central_partial_lh + (size_t)lk_cuda->nvec_central * lk_cuda->block_size;
    nni_scale_num = central_scale_num + (size_t)lk_cuda->nvec_central * lk_cuda->scale_block_size;
}

/* after the recursion: it pointed tip_partial_lh behind the last vector of the (device) arena, phylotree.cpp:886-894 */
void PhyloTree::lkCudaAfterInit() {
    if (!lk_cuda) return;
    tip_partial_lh = lk_cuda->tip_partial_lh;
    // the per-pattern host arrays live until deleteAllPartialLh (phylotree.cpp:683-693): page-locked, the re-upload of ptn_freq / ptn_invar
    // after every clearAllPartialLH (phylotree.cpp:499) and the fetch of _pattern_lh run at PCIe speed.  Best effort: a refusal is ignored.
    void *arr[3] = {ptn_freq, ptn_invar, _pattern_lh};
    const size_t bytes[3] = {lk_cuda->nptn * sizeof(double), (lk_cuda->nptn + lk_cuda->nunobs) * sizeof(double), lk_cuda->nptn * sizeof(double)};
    for (int i = 0; i < 3; i++) {
        if (lk_cuda->pinned[i] == arr[i]) continue;
        if (lk_cuda->pinned[i]) { lkcuda_host_unregister(lk_cuda->engine, lk_cuda->pinned[i]); lk_cuda->pinned[i] = NULL; }
        if (arr[i] && bytes[i] >= (1u << 20) && lkcuda_host_register(lk_cuda->engine, arr[i], bytes[i]) == LKCUDA_OK) lk_cuda->pinned[i] = arr[i];
    }
}

int PhyloTree::lkCudaScaleNumBytes() {
    return (int)(lk_cuda->scale_block_size * sizeof(UBYTE));
}

int PhyloTree::lkCudaVector(PhyloNeighbor *nei) {
    if (!nei->partial_lh || !nei->scale_num) outError("likelihood vector without memory (re-orientation missed?)", __func__);
    const size_t off = nei->partial_lh - (double*)lk_cuda->d_lh, soff = nei->scale_num - (UBYTE*)lk_cuda->d_sc;
    const size_t v = off / lk_cuda->block_size;
    if (off % lk_cuda->block_size || soff != v * lk_cuda->scale_block_size || v >= (size_t)lk_cuda->nvec_total)
        outError("likelihood vector outside the device arena", __func__);
    return (int)v;
}

/* inputs: tip table / ptn_freq / ptn_invar when the reference has just recomputed them, the model's small arrays whenever they differ
   from what the engine holds (the CPU kernels read them from the model objects at every call) */
void PhyloTree::lkCudaSyncInputs() {
    lkcuda_engine *e = lk_cuda->engine;
    if (!tip_partial_lh_computed) {
        computeTipPartialLikelihood();      // also computePtnFreq / computePtnInvar, phylotreesse.cpp:359-372
        lk_cuda->weights_uploaded = false;
    }
    if (!lk_cuda->weights_uploaded) {
        lkCheck(lkcuda_set_pattern_weights(e, ptn_freq, ptn_invar), e, "lkcuda_set_pattern_weights");
        lk_cuda->weights_uploaded = true;
    }
    const bool mixture = model->isMixture(), fused = model_factory->fused_mix_rate;
    const size_t S = aln->num_states, ncat = site_rate->getNRate(), nmix = model->getNMixtures(), K = lk_cuda->nblocks;
    std::vector<double> cur(nmix * S + 2 * nmix * S * S + 2 * K + lk_cuda->tip_partial_lh_size);
    double *p = &cur[0];
    memcpy(p, model->getEigenvalues(), nmix * S * sizeof(double)); p += nmix * S;
    memcpy(p, model->getEigenvectors(), nmix * S * S * sizeof(double)); p += nmix * S * S;
    memcpy(p, model->getInverseEigenvectors(), nmix * S * S * sizeof(double)); p += nmix * S * S;
    double *rate = p, *weight = p + K;
    std::vector<int32_t> eig(K);
    for (size_t k = 0; k < K; k++) {
        if (mixture && !fused) {            // block m*ncat + c (phylokernelmixture.h:121), weight prop_c * prop[m] (:766)
            const size_t m = k / ncat, c = k % ncat;
            eig[k] = (int32_t)m; rate[k] = site_rate->getRate(c); weight[k] = site_rate->getProp(c) * ((ModelMixture*)model)->prop[m];
        } else {                            // fused: category c uses component c (phylokernelmixrate.h:52, 99-130)
            eig[k] = fused ? (int32_t)k : 0; rate[k] = site_rate->getRate(k); weight[k] = site_rate->getProp(k);
        }
    }
    p += 2 * K;
    memcpy(p, tip_partial_lh, lk_cuda->tip_partial_lh_size * sizeof(double));
    if (cur.size() != lk_cuda->model_cache.size() || memcmp(&cur[0], &lk_cuda->model_cache[0], cur.size() * sizeof(double)) != 0) {
        lkCheck(lkcuda_set_model(e, &cur[0], &cur[nmix * S], &cur[nmix * S + nmix * S * S], &eig[0], rate, weight, p), e, "lkcuda_set_model");
        lk_cuda->model_cache.swap(cur);
    }
}

void PhyloTree::lkCudaFlush() {
    lkCudaSyncInputs();
    if (lk_cuda->pending.empty()) return;
    // execution order, level_start = NULL: the engine cuts the list where LM_PER_NODE re-uses a vector
    lkCheck(lkcuda_partial_batch(lk_cuda->engine, &lk_cuda->pending[0], (int32_t)lk_cuda->pending.size(), NULL, 0), lk_cuda->engine, "lkcuda_partial_batch");
    lk_cuda->nflush++; lk_cuda->nops += lk_cuda->pending.size();
    lk_cuda->pending.clear();
}

/* ------------------------------------------------------------ the four kernels ------------------------------------------------------------ */

/* phylokernel.h:71-143 (generic twin phylotreesse.cpp:581-643): same early-outs, same recursion, same re-orientation; records the node */
void PhyloTree::computePartialLikelihoodCuda(PhyloNeighbor *dad_branch, PhyloNode *dad) {
    assert(dad);
    if (dad_branch->partial_lh_computed & 1)
        return;
    dad_branch->partial_lh_computed |= 1;
    num_partial_lh_computations++;
    if (!lk_cuda) outError("CUDA likelihood kernels are bound but the buffers are host buffers: call initializeAllPartialLh()");
    PhyloNode *node = (PhyloNode*)(dad_branch->node);
    dad_branch->lh_scale_factor = 0.0;      // not maintained per node: the reductions derive it from scale_num (lkCudaScaleFactor on demand)
    if (node->isLeaf())
        return;
    const bool multi = node->degree() > 3;
    PhyloNeighbor *left = NULL, *right = NULL;
    if (multi || sse == LK_EIGEN) {
        // generic kernel: children in neighbour order (phylotreesse.cpp:617-624)
        FOR_NEIGHBOR_IT(node, dad, it) {
            PhyloNeighbor *nei = (PhyloNeighbor*)*it;
            if (!left) left = nei; else right = nei;
            if ((nei->partial_lh_computed & 1) == 0)
                computePartialLikelihoodCuda(nei, node);
        }
        if (!multi && !left->node->isLeaf() && right->node->isLeaf()) { PhyloNeighbor *tmp = left; left = right; right = tmp; }
    } else {
        FOR_NEIGHBOR_IT(node, dad, it) {
            if (!left) left = (PhyloNeighbor*)(*it); else right = (PhyloNeighbor*)(*it);
        }
        if (!left->node->isLeaf() && right->node->isLeaf()) { PhyloNeighbor *tmp = left; left = right; right = tmp; }
        if ((left->partial_lh_computed & 1) == 0)
            computePartialLikelihoodCuda(left, node);
        if ((right->partial_lh_computed & 1) == 0)
            computePartialLikelihoodCuda(right, node);
    }
    if (params->lh_mem_save == LM_PER_NODE && !dad_branch->partial_lh) {
        // re-orient partial_lh (phylokernel.h:127-143): take the vector of the first child-side neighbour that still owns one
        bool done = false;
        FOR_NEIGHBOR_IT(node, dad, it2) {
            PhyloNeighbor *backnei = ((PhyloNeighbor*)(*it2)->node->findNeighbor(node));
            if (backnei->partial_lh) {
                dad_branch->partial_lh = backnei->partial_lh;
                dad_branch->scale_num = backnei->scale_num;
                backnei->partial_lh = NULL;
                backnei->scale_num = NULL;
                backnei->partial_lh_computed &= ~1; // clear bit
                done = true;
                break;
            }
        }
        if (!done) outError("partial_lh is not re-oriented", __func__);
    }
    if (multi) {
        // multifurcation: the generic kernel's path (phylotreesse.cpp:702-806), issued right behind its children
        std::vector<int32_t> child; std::vector<double> len;
        FOR_NEIGHBOR_IT(node, dad, it) {
            PhyloNeighbor *nei = (PhyloNeighbor*)*it;
            child.push_back(nei->node->isLeaf() ? LKCUDA_TIP(nei->node->id) : lkCudaVector(nei));
            len.push_back(nei->length);
        }
        lkCudaFlush();
        lkCheck(lkcuda_partial_multi(lk_cuda->engine, lkCudaVector(dad_branch), (int32_t)child.size(), &child[0], &len[0]), lk_cuda->engine, "lkcuda_partial_multi");
        return;
    }
    lkcuda_partial_op op;
    op.dst = lkCudaVector(dad_branch);
    op.left = left->node->isLeaf() ? LKCUDA_TIP(left->node->id) : lkCudaVector(left);
    op.right = right->node->isLeaf() ? LKCUDA_TIP(right->node->id) : lkCudaVector(right);
    op.reserved = 0;
    op.len_left = left->length;
    op.len_right = right->length;
    lk_cuda->pending.push_back(op);
}

/* both directions of the branch, the swap that makes `dad` the leaf end (phylokernel.h:734-751), then everything recorded goes to the device */
void PhyloTree::lkCudaBranchEnds(PhyloNeighbor *&dad_branch, PhyloNode *&dad, int &dad_end, int &node_vec) {
    PhyloNode *node = (PhyloNode*) dad_branch->node;
    PhyloNeighbor *node_branch = (PhyloNeighbor*) node->findNeighbor(dad);
    if (!central_partial_lh)
        initializeAllPartialLh();
    if (!lk_cuda) outError("CUDA likelihood kernels are bound but the buffers are host buffers");
    if (node->isLeaf()) {
        PhyloNode *tmp_node = dad; dad = node; node = tmp_node;
        PhyloNeighbor *tmp_nei = dad_branch; dad_branch = node_branch; node_branch = tmp_nei;
    }
    if ((dad_branch->partial_lh_computed & 1) == 0)
        computePartialLikelihoodCuda(dad_branch, dad);
    if ((node_branch->partial_lh_computed & 1) == 0)
        computePartialLikelihoodCuda(node_branch, node);
    if (!lk_defer_flush)
        lkCudaFlush();
    // dad_branch (stored at dad, pointing to node) holds the subtree at `node`; node_branch the one at `dad`
    dad_end = dad->isLeaf() ? LKCUDA_TIP(dad->id) : lkCudaVector(node_branch);
    node_vec = lkCudaVector(dad_branch);
}

double PhyloTree::computeLikelihoodBranchCuda(PhyloNeighbor *dad_branch, PhyloNode *dad) {
    int dad_end, node_vec;
    lk_defer_flush = true;      // the recorded traversal and this branch go to the device in ONE call (lkcuda_partial_batch_lnl)
    lkCudaBranchEnds(dad_branch, dad, dad_end, node_vec);
    lk_defer_flush = false;
    double r[LKCUDA_NRESULT];
    lkCudaSyncInputs();
    if (!lk_cuda->pending.empty()) {
        lkCheck(lkcuda_partial_batch_lnl(lk_cuda->engine, &lk_cuda->pending[0], (int32_t)lk_cuda->pending.size(), NULL, 0,
                                         dad_end, node_vec, dad_branch->length, r), lk_cuda->engine, "lkcuda_partial_batch_lnl");
        lk_cuda->nflush++; lk_cuda->nops += lk_cuda->pending.size();
        lk_cuda->pending.clear();
    } else
        lkCheck(lkcuda_branch_lnl(lk_cuda->engine, dad_end, node_vec, dad_branch->length, r), lk_cuda->engine, "lkcuda_branch_lnl");
    if (r[LKCUDA_R_NREPAIRED] > 0)      // phylokernel.h:848-866
        outWarning("Numerical underflow for " + convertIntToString((int)r[LKCUDA_R_NREPAIRED]) + " alignment patterns");
    assert(!isnan(r[LKCUDA_R_LNL]) && !isinf(r[LKCUDA_R_LNL]));
    return r[LKCUDA_R_LNL];             // lh_scale_factor of both ends and the +ASC term are inside
}

void PhyloTree::computeLikelihoodDervCuda(PhyloNeighbor *dad_branch, PhyloNode *dad, double &df, double &ddf) {
    int dad_end, node_vec;
    lkCudaBranchEnds(dad_branch, dad, dad_end, node_vec);
    double r[LKCUDA_NRESULT];
    if (!theta_computed) {              // phylokernel.h:535-579: theta_all and the first df/ddf in one pass over the two vectors
        theta_computed = true;
        lkCheck(lkcuda_theta_derv(lk_cuda->engine, dad_end, node_vec, dad_branch->length, r), lk_cuda->engine, "lkcuda_theta_derv");
    } else
        lkCheck(lkcuda_derv(lk_cuda->engine, dad_branch->length, r), lk_cuda->engine, "lkcuda_derv");
    df = r[LKCUDA_R_DF];
    ddf = r[LKCUDA_R_DDF];
    if (isnan(df) || isinf(df)) {       // phylokernel.h:647-651
        df = 0.0;
        ddf = 0.0;
    }
}

double PhyloTree::computeLikelihoodFromBufferCuda() {
    assert(current_it && current_it_back);
    if (!theta_computed || !lk_cuda)    // families without a from-buffer kernel do the same (phylotreesse.cpp:349-357)
        return computeLikelihoodBranchCuda(current_it, (PhyloNode*)current_it_back->node);
    double r[LKCUDA_NRESULT];
    lkCheck(lkcuda_lnl_from_buffer(lk_cuda->engine, current_it->length, r), lk_cuda->engine, "lkcuda_lnl_from_buffer");
    return r[LKCUDA_R_LNL];
}

/* ------------------------------------------------------------ host code that reads buffer contents ------------------------------------------------------------ */

/* lh_scale_factor of a neighbour = sum_p LOG_SCALING_THRESHOLD * ptn_freq[p] * scale_num[p] (phylokernel.h:389-395) */
double PhyloTree::lkCudaScaleFactor(PhyloNeighbor *nei) {
    if (nei->node->isLeaf() || !nei->partial_lh) return 0.0;
    double f = 0.0;
    lkCheck(lkcuda_vector_scale_factor(lk_cuda->engine, lkCudaVector(nei), &f), lk_cuda->engine, "lkcuda_vector_scale_factor");
    return f;
}

/* fetch one vector to host memory (debugging, computeBayesianBranchLength-style readers, the test harness) */
void PhyloTree::lkCudaFetchVector(PhyloNeighbor *nei, double *partial_lh, UBYTE *scale_num) {
    lkCudaFlush();
    lkCheck(lkcuda_fetch_vector(lk_cuda->engine, lkCudaVector(nei), partial_lh, (int16_t*)scale_num), lk_cuda->engine, "lkcuda_fetch_vector");
}

void PhyloTree::lkCudaFetchTheta(double *theta) {
    lkCheck(lkcuda_fetch_theta(lk_cuda->engine, theta), lk_cuda->engine, "lkcuda_fetch_theta");
}

/* bring the host copy of _pattern_lh up to date (the kernels leave it on the device until host code asks) */
void PhyloTree::lkCudaSyncPatternLh() {
    if (lk_cuda) lkCheck(lkcuda_fetch_pattern_lh(lk_cuda->engine, _pattern_lh), lk_cuda->engine, "lkcuda_fetch_pattern_lh");
}

/* the tail of computeLikelihood(double *pattern_lh) (phylotree.cpp:1055-1069): _pattern_lh plus the scale_num correction of current_it */
void PhyloTree::lkCudaPatternLh(double *pattern_lh) {
    lkcuda_engine *e = lk_cuda->engine;
    lkCheck(lkcuda_fetch_pattern_lh(e, _pattern_lh), e, "lkcuda_fetch_pattern_lh");
    current_it->lh_scale_factor = lkCudaScaleFactor(current_it);
    if (current_it->lh_scale_factor < 0.0)
        lkCheck(lkcuda_fetch_pattern_lh_corrected(e, lkCudaVector(current_it), pattern_lh), e, "lkcuda_fetch_pattern_lh_corrected");
    else
        memmove(pattern_lh, _pattern_lh, aln->size() * sizeof(double));
}

/* PhyloTree::computePatternLhCat (phylotree.cpp:1108-1160): the generic branch kernels' _pattern_lh_cat for current_it, then the same
   collapse over mixture classes / rate categories on the host copy */
double PhyloTree::computePatternLhCatCuda(SiteLoglType wsl) {
    PhyloNeighbor *dad_branch = current_it;
    PhyloNode *dad = (PhyloNode*)current_it_back->node;
    int dad_end, node_vec;
    lkCudaBranchEnds(dad_branch, dad, dad_end, node_vec);
    lkcuda_engine *e = lk_cuda->engine;
    double r[LKCUDA_NRESULT];
    lkCheck(lkcuda_branch_lnl(e, dad_end, node_vec, dad_branch->length, r), e, "lkcuda_branch_lnl");
    lkCheck(lkcuda_pattern_lh_cat(e, dad_end, node_vec, dad_branch->length), e, "lkcuda_pattern_lh_cat");
    lkCheck(lkcuda_fetch_pattern_lh_cat(e, _pattern_lh_cat), e, "lkcuda_fetch_pattern_lh_cat");
    const double score = r[LKCUDA_R_LNL];
    if (!getModel()->isMixture() || getModelFactory()->fused_mix_rate || wsl == WSL_MIXTURE_RATECAT)
        return score;
    const size_t nptn = aln->getNPattern(), nmixture = getModel()->getNMixtures(), ncat = getRate()->getNRate();
    double *src = _pattern_lh_cat, *dst = _pattern_lh_cat;
    if (wsl == WSL_MIXTURE && ncat > 1) {               // one value per mixture class: sum the class's categories, first to last
        for (size_t ptn = 0; ptn < nptn; ptn++)
            for (size_t m = 0; m < nmixture; m++, src += ncat) {
                double lh = src[0];
                for (size_t c = 1; c < ncat; c++) lh += src[c];
                *dst++ = lh;
            }
    } else if (wsl == WSL_RATECAT && nmixture > 1) {    // one value per rate category: add the classes, first to last
        for (size_t ptn = 0; ptn < nptn; ptn++, dst += ncat) {
            double acc[64];
            std::vector<double> big;
            double *a = acc;
            if (ncat > 64) { big.resize(ncat); a = &big[0]; }
            for (size_t c = 0; c < ncat; c++) a[c] = src[c];
            src += ncat;
            for (size_t m = 1; m < nmixture; m++, src += ncat)
                for (size_t c = 0; c < ncat; c++) a[c] += src[c];
            for (size_t c = 0; c < ncat; c++) dst[c] = a[c];
        }
    }
    return score;
}

/* PhyloTree::computePatternLikelihood (phylotree.cpp:1200-1262) on host copies of _pattern_lh and of the two ends' scale_num */
void PhyloTree::computePatternLikelihoodCuda(double *ptn_lh, double *cur_logl, double *ptn_lh_cat, SiteLoglType wsl) {
    const size_t nptn = aln->getNPattern();
    const int ncat = ptn_lh_cat ? getNumLhCat(wsl) : 0;
    if (ptn_lh_cat)
        computePatternLhCatCuda(wsl);
    lkcuda_engine *e = lk_cuda->engine;
    lkCudaFlush();
    lkCheck(lkcuda_fetch_pattern_lh(e, _pattern_lh), e, "lkcuda_fetch_pattern_lh");
    // scale_num of both ends; a leaf end (or an end without a vector) never scales
    std::vector<int16_t> sc(nptn, 0), tmp(lk_cuda->nptn + lk_cuda->nunobs);
    PhyloNeighbor *ends[2] = {current_it, current_it_back};
    bool any = false;
    for (int k = 0; k < 2; k++) {
        if (ends[k]->node->isLeaf() || !ends[k]->partial_lh) continue;
        lkCheck(lkcuda_fetch_vector(e, lkCudaVector(ends[k]), NULL, &tmp[0]), e, "lkcuda_fetch_vector");
        for (size_t i = 0; i < nptn; i++) if (tmp[i] > 0) { sc[i] += tmp[i]; any = true; }
    }
    if (any) {
        for (size_t i = 0; i < nptn; i++) ptn_lh[i] = _pattern_lh[i] + sc[i] * LOG_SCALING_THRESHOLD;
    } else
        memmove(ptn_lh, _pattern_lh, nptn * sizeof(double));
    if (ptn_lh_cat) {
        size_t offset = 0;
        for (size_t i = 0; i < nptn; i++) {
            const double scale = sc[i] * LOG_SCALING_THRESHOLD;
            for (int j = 0; j < ncat; j++, offset++)
                ptn_lh_cat[offset] = any ? log(_pattern_lh_cat[offset]) + scale : log(_pattern_lh_cat[offset]);
        }
    }
}

/* per-kernel-kind statistics of the engine (lkcuda_kernel_stats): launches, algorithmic bytes / flops, device time with profiling on */
void PhyloTree::lkCudaProfile(bool on, bool reset) {
    if (!lk_cuda) return;
    if (reset) lkcuda_reset_counters(lk_cuda->engine);
    lkcuda_set_profiling(lk_cuda->engine, on ? 1 : 0);
}

bool PhyloTree::lkCudaKernelStats(int kind, uint64_t &launches, uint64_t &bytes, uint64_t &flops, double &device_ms) {
    launches = bytes = flops = 0; device_ms = 0.0;
    if (!lk_cuda || kind < 0 || kind >= LKCUDA_NKIND) return false;
    lkcuda_kernel_stats(lk_cuda->engine, kind, &launches, &bytes, &device_ms);
    lkcuda_kernel_flops(lk_cuda->engine, kind, &flops);
    return true;
}

/* counters for the harness / log output */
void PhyloTree::lkCudaStats(uint64_t &flushes, uint64_t &ops, uint64_t &launches) {
    flushes = ops = launches = 0;
    if (!lk_cuda) return;
    flushes = lk_cuda->nflush; ops = lk_cuda->nops;
    lkcuda_launch_count(lk_cuda->engine, &launches);
}
