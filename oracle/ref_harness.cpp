/*
 * ref_harness — TEST INFRASTRUCTURE (oracle), not product code.
 *
 * Drives the UNMODIFIED reference PhyloTree (IQ-TREE 1.4.3-beta, compiled by oracle/Makefile.ref
 * from /root/reference) as a library and
 *   (1) dumps golden fixtures: the exact kernel inputs the reference sees (pattern matrix, eigen
 *       system, rates, tip table, tree adjacency) and the reference's own outputs (lnL, _pattern_lh,
 *       _pattern_lh_cat, df/ddf, partial_lh / scale_num of every computed neighbour, branch lengths
 *       after optimizeAllBranches, the branch traversal order);
 *   (2) times the reference's CPU kernels (AVX + OpenMP) for bench.py --impl reference.
 *
 * The set-up sequence mirrors runTreeReconstruction (phyloanalysis.cpp:1611-1726).
 *
 *   (3) dumps the SURVEY 8(f) consumers' golden values (dumpx): UFBoot RELL dot products of IQTree::saveCurrentTree, RateGamma::
 *       computePatternRates and ModelMixture::optimizeWeights, all computed by the reference's own functions.
 *
 * usage: ref_harness dump <outdir> <iqtree args: -s aln -m MODEL -te tree -nt N ...>
 *        ref_harness dumpx <outdir> <nsamples> <iqtree args ...>
 *        ref_harness dumpnni <outdir> <nni5: 0|1> <iqtree args ...>     PhyloTree::getBestNNIForBran on every internal branch
 *        ref_harness dumppars <outdir> <seed> <iqtree args ...>         Fitch parsimony: scores per branch, stepwise-addition tree
 *        ref_harness dumpem <outdir> <iqtree args ...>                  RateFree::optimizeWithEM: rates, proportions, lnL before / after
 *        ref_harness time <lnl_reps> <derv_reps> <oab_reps> <iqtree args ...>
 *        ref_harness dumplite <outdir> <stride> <iqtree args ...>      BASELINE-size parity: model inputs, lnL, every stride-th pattern log-likelihood,
 *                                                                       scale_num of three vectors, df / ddf / lnL-from-buffer on three branches
 *        ref_harness dumprepair <outdir> <iqtree args ...>             the NaN / Inf repair of the branch kernel (phylokernel.h:848-866) on a zeroed vector
 *        ref_harness iqtree <iqtree args ...>                          the complete command line
 */
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <string>
#include <vector>
#include <map>
#include <set>
#include <sstream>
#include <fstream>
#include <iostream>
#include <iomanip>
#include <algorithm>
#include <sys/stat.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define private public
#define protected public
#include "phylotree.h"
#include "iqtree.h"
#include "phyloanalysis.h"
#include "phylosupertree.h"
#include "phylosupertreeplen.h"
#include "tools.h"
#include "alignment.h"
#include "checkpoint.h"
#include "timeutil.h"
#include "model/modelfactory.h"
#include "model/modelmixture.h"
#include "model/modelset.h"
#include "model/rateheterogeneity.h"
#include "model/ratefree.h"
#include "vectorclass/instrset.h"
#undef private
#undef protected

extern int instruction_set;   // pda.cpp:2135

static std::string g_out;

static void wbin(const std::string &name, const void *p, size_t bytes) {
    std::string fn = g_out + "/" + name;
    FILE *f = fopen(fn.c_str(), "wb");
    if (!f) { perror(fn.c_str()); exit(3); }
    if (bytes) fwrite(p, 1, bytes, f);
    fclose(f);
}

/* With -DLK_CUDA_SHIM this file is compiled against the patched reference (integration/): under `-cuda` the likelihood buffers live on
   the device and are fetched through the shim's accessors; without it the helpers below read the reference's host buffers directly. */
#ifdef LK_CUDA_SHIM
#define ON_DEVICE(tree) ((tree)->lk_cuda != NULL)
#else
#define ON_DEVICE(tree) false
#endif

static void syncPatternLh(IQTree *tree) {
#ifdef LK_CUDA_SHIM
    tree->lkCudaSyncPatternLh();
#else
    (void)tree;
#endif
}

static void collectNodes(Node *node, Node *dad, std::vector<Node*> &out) {
    out.push_back(node);
    FOR_NEIGHBOR_IT(node, dad, it) collectNodes((*it)->node, node, out);
}

/* adjacency dump: one line per (node, neighbour) in the reference's own neighbour order */
static void dumpAdjacency(IQTree *tree, const std::string &name) {
    std::vector<Node*> nodes;
    collectNodes(tree->root, NULL, nodes);
    std::ofstream f((g_out + "/" + name).c_str());
    f << std::setprecision(17);
    f << "root " << tree->root->id << " nodes " << nodes.size() << " leaves " << tree->leafNum << "\n";
    for (size_t i = 0; i < nodes.size(); i++) {
        Node *n = nodes[i];
        f << "node " << n->id << " " << (n->isLeaf() ? 1 : 0) << " " << (n->name.empty() ? "-" : n->name)
          << " " << n->neighbors.size();
        for (size_t k = 0; k < n->neighbors.size(); k++)
            f << " " << n->neighbors[k]->node->id << " " << n->neighbors[k]->length;
        f << "\n";
    }
}

static void dumpPartials(IQTree *tree, Node *node, Node *dad, std::ofstream &idx, size_t nptn, size_t block) {
    FOR_NEIGHBOR_IT(node, dad, it) {
        PhyloNeighbor *nei = (PhyloNeighbor*)(*it);
        PhyloNeighbor *back = (PhyloNeighbor*)(*it)->node->findNeighbor(node);
        PhyloNeighbor *both[2] = {nei, back};
        Node *from[2] = {node, (*it)->node};
        for (int k = 0; k < 2; k++) {
            PhyloNeighbor *b = both[k];
            // neighbour b: stored at from[k], pointing to b->node; holds partial of subtree rooted at b->node
            if ((b->partial_lh_computed & 1) && b->partial_lh && !b->node->isLeaf()) {
                std::ostringstream nm;
                nm << "partial_" << from[k]->id << "_" << b->node->id;
                double factor = b->lh_scale_factor;
#ifdef LK_CUDA_SHIM
                if (ON_DEVICE(tree)) {
                    std::vector<double> lh(nptn * block);
                    std::vector<UBYTE> sc(nptn);
                    tree->lkCudaFetchVector(b, &lh[0], &sc[0]);
                    factor = tree->lkCudaScaleFactor(b);
                    wbin(nm.str() + ".f64", &lh[0], nptn * block * sizeof(double));
                    wbin(nm.str() + ".i16", &sc[0], nptn * sizeof(UBYTE));
                } else
#endif
                {
                wbin(nm.str() + ".f64", b->partial_lh, nptn * block * sizeof(double));
                wbin(nm.str() + ".i16", b->scale_num, nptn * sizeof(UBYTE));
                }
                idx << from[k]->id << " " << b->node->id << " " << std::setprecision(17) << factor << "\n";
            }
        }
        dumpPartials(tree, (*it)->node, node, idx, nptn, block);
    }
}

static IQTree *setup(int argc, char **argv, Params &params) {
    parseArg(argc, argv, params);
    instruction_set = instrset_detect();
#ifdef _OPENMP
    if (params.num_threads) omp_set_num_threads(params.num_threads);
    omp_set_nested(false);
#endif
    init_random(params.ran_seed, false);
    IQTree *tree;
    if (params.partition_file) {        // partition models, as runPhyloAnalysis creates them (phyloanalysis.cpp:2461-2476)
        if (params.partition_type) tree = new PhyloSuperTreePlen(params);
        else tree = new PhyloSuperTree(params);
    } else
        tree = new IQTree(new Alignment(params.aln_file, params.sequence_type, params.intype));
    tree->setCheckpoint(new Checkpoint);
    if (params.min_branch_length <= 0.0) {       // phyloanalysis.cpp:2488-2497
        params.min_branch_length = 1e-6;
        if (tree->getAlnNSite() >= 100000)
            params.min_branch_length = 0.1 / (tree->getAlnNSite());
    }
    tree->setParams(&params);
    std::string dist_file;
    tree->computeInitialTree(dist_file, params.SSE);
    tree->setRootNode(params.root);
    ModelsBlock *mb = readModelsDefinition(params);
    tree->initSettings(params);
    tree->initializeModel(params, mb);
    delete mb;
    tree->initializeAllPartialLh();
    return tree;
}

static const char *familyName(IQTree *tree) {
    if (tree->getModel()->isSiteSpecificModel()) return "sitemodel";
    if (tree->getModel()->isMixture())
        return tree->getModelFactory()->fused_mix_rate ? "mixrate" : "mixture";
    return "plain";
}

static int do_dump(int argc, char **argv) {
    g_out = argv[2];
    mkdir(g_out.c_str(), 0755);
    std::vector<char*> args; args.push_back(argv[0]);
    for (int i = 3; i < argc; i++) args.push_back(argv[i]);
    Params &params = Params::getInstance();
    IQTree *tree = setup((int)args.size(), &args[0], params);
    Alignment *aln = tree->aln;
    ModelSubst *model = tree->getModel();
    RateHeterogeneity *rate = tree->getRate();
    ModelFactory *mf = tree->getModelFactory();

    size_t S = aln->num_states, C = rate->getNRate(), M = model->getNMixtures();
    bool fused = mf->fused_mix_rate;
    size_t P = aln->size(), U = mf->unobserved_ptns.size(), nptn = P + U, ntaxa = aln->getNSeq();
    size_t block = tree->getModel()->isMixture() && !fused ? S * C * M : S * C;
    size_t E = M;   // number of eigensystems
    const bool sitemodel = model->isSiteSpecificModel();     // -fs: one model per pattern (ModelSet, phylokernelsitemodel.cpp)
    if (sitemodel) { E = P; M = P; }

    /* ---- kernel inputs ---- */
    std::vector<unsigned char> states(ntaxa * nptn);
    for (size_t t = 0; t < ntaxa; t++) {
        for (size_t p = 0; p < P; p++) states[t * nptn + p] = (unsigned char)(aln->at(p))[t];
        for (size_t u = 0; u < U; u++) states[t * nptn + P + u] = (unsigned char)mf->unobserved_ptns[u];
    }
    wbin("tip_states.u8", &states[0], states.size());
    std::vector<int> site2ptn(aln->getNSite());
    for (int s = 0; s < aln->getNSite(); s++) site2ptn[s] = aln->getPatternID(s);
    wbin("site_pattern.i32", &site2ptn[0], site2ptn.size() * sizeof(int));
    std::vector<int> cc(P);
    for (size_t p = 0; p < P; p++) cc[p] = (int)(signed char)aln->at(p).const_char;
    wbin("const_char.i32", &cc[0], P * sizeof(int));
    {
        std::ofstream f((g_out + "/taxa.txt").c_str());
        for (size_t t = 0; t < ntaxa; t++) f << aln->getSeqName(t) << "\n";
    }
    if (sitemodel) {
        ModelSet *models = (ModelSet*)model;
        std::vector<double> ev(E * S), evc(E * S * S), inv(E * S * S);
        std::vector<int> pm(P);
        for (size_t p = 0; p < P; p++) {
            memcpy(&ev[p * S], models->at(p)->getEigenvalues(), S * 8);
            memcpy(&evc[p * S * S], models->at(p)->getEigenvectors(), S * S * 8);
            memcpy(&inv[p * S * S], models->at(p)->getInverseEigenvectors(), S * S * 8);
            pm[p] = (int)p;
        }
        wbin("eval.f64", &ev[0], ev.size() * 8); wbin("evec.f64", &evc[0], evc.size() * 8); wbin("inv_evec.f64", &inv[0], inv.size() * 8);
        wbin("ptn_model.i32", &pm[0], P * sizeof(int));
    } else {
    wbin("eval.f64", model->getEigenvalues(), E * S * sizeof(double));
    wbin("evec.f64", model->getEigenvectors(), E * S * S * sizeof(double));
    wbin("inv_evec.f64", model->getInverseEigenvectors(), E * S * S * sizeof(double));
    }
    std::vector<double> rates(C), props(C), mixprops(M, 1.0), sfreq(S);
    for (size_t c = 0; c < C; c++) { rates[c] = rate->getRate(c); props[c] = rate->getProp(c); }
    if (model->isMixture() && !sitemodel) for (size_t m = 0; m < M; m++) mixprops[m] = ((ModelMixture*)model)->prop[m];
    model->getStateFrequency(&sfreq[0]);
    wbin("rates.f64", &rates[0], C * 8); wbin("props.f64", &props[0], C * 8);
    wbin("mixprops.f64", &mixprops[0], M * 8); wbin("state_freq.f64", &sfreq[0], S * 8);
    dumpAdjacency(tree, "tree_adj.txt");
    {
        std::ofstream f((g_out + "/tree0.nwk").c_str());
        f << std::setprecision(17);
        tree->printTree(f, WT_BR_LEN);
        f << "\n";
    }

    /* ---- reference outputs ---- */
    tree->clearAllPartialLH();
    double lnl = tree->computeLikelihood();
    // computeTipPartialLikelihood has run now
    if (sitemodel) wbin("tip_partial_lh.f64", tree->tip_partial_lh, 0);     // [taxon][ptn][S] there: formed on the fly by the engine
    else wbin("tip_partial_lh.f64", tree->tip_partial_lh, (aln->STATE_UNKNOWN + 1) * S * E * sizeof(double));
    wbin("ptn_freq.f64", tree->ptn_freq, P * 8);
    wbin("ptn_invar.f64", tree->ptn_invar, nptn * 8);
    syncPatternLh(tree);
    wbin("pattern_lh.f64", tree->_pattern_lh, P * 8);
    int root_from = tree->current_it_back->node->id, root_to = tree->current_it->node->id;
    {
        std::vector<double> plh(P);
        double l2 = tree->computeLikelihood(&plh[0]);       // with scale_num correction
        (void)l2;
        wbin("pattern_lh_scaled.f64", &plh[0], P * 8);
    }
    {
        std::ofstream idx((g_out + "/partials.txt").c_str());
        dumpPartials(tree, tree->root, NULL, idx, nptn, block);
    }
    /* per-category pattern likelihoods via the generic kernel (phylotree.cpp:1108) */
    if (sitemodel) {     // the site-model branch kernel always fills _pattern_lh_cat[ptn][c] (phylokernelsitemodel.cpp:579-603)
        tree->computeSitemodelLikelihoodBranchEigen(tree->current_it, (PhyloNode*)tree->current_it_back->node);   // the SIMD twins never fill it
        wbin("pattern_lh_cat.f64", tree->_pattern_lh_cat, P * C * 8);
        tree->clearAllPartialLH();
        tree->computeLikelihood();
    } else {
        tree->computePatternLhCat(WSL_RATECAT);
        size_t ncat_tot = (model->isMixture() && !fused) ? C * M : C;
        wbin("pattern_lh_cat.f64", tree->_pattern_lh_cat, P * ncat_tot * 8);
        tree->clearAllPartialLH();
        tree->computeLikelihood();
    }

    /* df/ddf on the root-leaf branch (root = tree->root, a leaf) and on every branch */
    PhyloNode *n1 = (PhyloNode*)tree->root, *n2 = (PhyloNode*)tree->root->neighbors[0]->node;
    double df, ddf;
    tree->theta_computed = false;
    tree->computeLikelihoodDerv((PhyloNeighbor*)n1->findNeighbor(n2), n1, df, ddf);
#ifdef LK_CUDA_SHIM
    if (ON_DEVICE(tree)) {
        std::vector<double> th(nptn * block);
        tree->lkCudaFetchTheta(&th[0]);
        wbin("theta_root.f64", &th[0], nptn * block * 8);
    } else
#endif
    wbin("theta_root.f64", tree->theta_all, nptn * block * 8);
    NodeVector nodes, nodes2;
    tree->computeBestTraversal(nodes, nodes2);
    std::vector<int> trav(nodes.size() * 2);
    std::vector<double> dervs(nodes.size() * 3);
    for (size_t j = 0; j < nodes.size(); j++) {
        trav[2 * j] = nodes[j]->id; trav[2 * j + 1] = nodes2[j]->id;
        double a, b;
        tree->theta_computed = false;
        PhyloNeighbor *nb = (PhyloNeighbor*)nodes[j]->findNeighbor(nodes2[j]);
        tree->computeLikelihoodDerv(nb, (PhyloNode*)nodes[j], a, b);
        tree->current_it = nb;
        tree->current_it_back = (PhyloNeighbor*)nodes2[j]->findNeighbor(nodes[j]);
        double fb = tree->computeLikelihoodFromBuffer();
        dervs[3 * j] = a; dervs[3 * j + 1] = b; dervs[3 * j + 2] = fb;
    }
    wbin("traversal.i32", &trav[0], trav.size() * sizeof(int));
    wbin("branch_derv.f64", &dervs[0], dervs.size() * 8);

    /* one pass and full branch-length optimisation */
    tree->clearAllPartialLH();
    tree->current_it = tree->current_it_back = NULL;
    tree->computeLikelihood();
    double oab1 = tree->optimizeAllBranches(1);
    dumpAdjacency(tree, "tree_adj_oab1.txt");
    double oabfull = tree->optimizeAllBranches();
    dumpAdjacency(tree, "tree_adj_oabfull.txt");

    {
        std::ofstream f((g_out + "/meta.json").c_str());
        f << std::setprecision(17);
        f << "{\n"
          << " \"nstates\": " << S << ", \"ncat\": " << C << ", \"nmix\": " << M << ",\n"
          << " \"family\": \"" << familyName(tree) << "\", \"block\": " << block << ",\n"
          << " \"nptn\": " << P << ", \"nunobs\": " << U << ", \"ntaxa\": " << ntaxa << ", \"nsite\": " << aln->getNSite() << ",\n"
          << " \"state_unknown\": " << (int)aln->STATE_UNKNOWN << ", \"seq_type\": " << (int)aln->seq_type << ",\n"
          << " \"model\": \"" << tree->getModelName() << "\", \"p_invar\": " << rate->getPInvar() << ",\n"
          << " \"min_branch_length\": " << params.min_branch_length << ", \"max_branch_length\": " << params.max_branch_length << ",\n"
          << " \"root_branch\": [" << root_from << ", " << root_to << "],\n"
          << " \"lnl\": " << lnl << ", \"df_root\": " << df << ", \"ddf_root\": " << ddf << ",\n"
          << " \"derv_branch\": [" << n1->id << ", " << n2->id << "],\n"
          << " \"oab1_lnl\": " << oab1 << ", \"oabfull_lnl\": " << oabfull << ",\n"
          << " \"kernel\": \"" << (ON_DEVICE(tree) ? "cuda" : (tree->sse == LK_EIGEN_SSE ? (instruction_set >= 7 ? "AVX" : "SSE") : "generic")) << "\",\n"
          << " \"lh_mem_save\": " << (int)params.lh_mem_save << ",\n"
          << " \"threads\": " << params.num_threads << "\n}\n";
    }
    printf("lnl %.10f df %.10g ddf %.10g oab1 %.10f oabfull %.10f\n", lnl, df, ddf, oab1, oabfull);
    return 0;
}

/* golden values of the consumers of _pattern_lh / _pattern_lh_cat from the reference's own functions */
static int do_dumpx(int argc, char **argv) {
    g_out = argv[2];
    mkdir(g_out.c_str(), 0755);
    int nsamples = atoi(argv[3]);
    std::vector<char*> args; args.push_back(argv[0]);
    for (int i = 4; i < argc; i++) args.push_back(argv[i]);
    Params &params = Params::getInstance();
    IQTree *tree = setup((int)args.size(), &args[0], params);
    Alignment *aln = tree->aln;
    ModelSubst *model = tree->getModel();
    RateHeterogeneity *rate = tree->getRate();
    bool fused = tree->getModelFactory()->fused_mix_rate;
    size_t P = aln->size(), C = rate->getNRate(), M = model->getNMixtures();

    tree->clearAllPartialLH();
    double lnl = tree->computeLikelihood();

    /* ---- RELL: the statements of IQTree::saveCurrentTree (iqtree.cpp:2689-2735) and of the boot_samples set-up (iqtree.cpp:300-338) ---- */
    size_t maxnptn = get_safe_upper_limit_float(P);
    BootValType *pattern_lh = aligned_alloc<BootValType>(maxnptn);
    memset(pattern_lh, 0, maxnptn * sizeof(BootValType));
    double *pattern_lh_orig = aligned_alloc<double>(P);
    double cur_logl = lnl;
    tree->computePatternLikelihood(pattern_lh_orig, &cur_logl);
    for (size_t i = 0; i < P; i++) pattern_lh[i] = (float)pattern_lh_orig[i];
    BootValType *mem = aligned_alloc<BootValType>(maxnptn * (size_t)nsamples);
    memset(mem, 0, maxnptn * (size_t)nsamples * sizeof(BootValType));
    std::vector<float> samples((size_t)nsamples * P);
    std::vector<double> rell(nsamples);
    VerboseMode saved_mode = verbose_mode;
    verbose_mode = VB_QUIET;
    for (int s = 0; s < nsamples; s++) {
        IntVector this_sample;
        aln->createBootstrapAlignment(this_sample, params.bootstrap_spec);
        for (size_t j = 0; j < P; j++) { mem[s * maxnptn + j] = this_sample[j]; samples[s * P + j] = (float)this_sample[j]; }
    }
    verbose_mode = saved_mode;
    for (int s = 0; s < nsamples; s++)
        rell[s] = (tree->*(tree->dotProduct))(pattern_lh, mem + s * maxnptn, (int)P);
    wbin("boot_samples.f32", &samples[0], samples.size() * sizeof(float));
    wbin("rell.f64", &rell[0], rell.size() * 8);
    wbin("ptn_lh_full.f64", pattern_lh_orig, P * 8);
    int root_from = tree->current_it_back->node->id, root_to = tree->current_it->node->id;

    /* ---- RateGamma::computePatternRates (model/rategamma.cpp:235-257); other rate models return 0 or their own estimate ---- */
    int nrate_cat = 0;
    {
        DoubleVector pr; IntVector pc;
        nrate_cat = rate->computePatternRates(pr, pc);
        if (nrate_cat > 0 && pr.size() == P) {
            std::vector<int> pci(pc.begin(), pc.end());
            wbin("pattern_rates.f64", &pr[0], P * 8);
            wbin("pattern_cat.i32", &pci[0], P * sizeof(int));
        } else nrate_cat = 0;
    }

    /* ---- ModelMixture::optimizeWeights (model/modelmixture.cpp:1356-1416) ---- */
    double em_lnl = 0.0; int has_em = 0;
    if (model->isMixture() && !fused) {
        ModelMixture *mm = (ModelMixture*)model;
        tree->computePatternLhCat(WSL_MIXTURE);
        wbin("lhcat_mixture.f64", tree->_pattern_lh_cat, P * M * 8);
        std::vector<double> before(mm->prop, mm->prop + M);
        em_lnl = mm->optimizeWeights();
        wbin("em_props_before.f64", &before[0], M * 8);
        wbin("em_props.f64", mm->prop, M * 8);
        has_em = 1;
    }
    {
        std::ofstream f((g_out + "/metax.json").c_str());
        f << std::setprecision(17);
        f << "{ \"nptn\": " << P << ", \"ncat\": " << C << ", \"nmix\": " << M << ", \"nsamples\": " << nsamples
          << ", \"lnl\": " << lnl << ", \"root_branch\": [" << root_from << ", " << root_to << "]"
          << ", \"nrate_cat\": " << nrate_cat << ", \"has_em\": " << has_em << ", \"em_lnl\": " << em_lnl
          << ", \"dot_kernel\": \"" << (tree->sse == LK_EIGEN_SSE ? (instruction_set >= 7 ? "AVX" : "SSE") : "generic") << "\" }\n";
    }
    printf("dumpx lnl %.10f rell0 %.6f nrate_cat %d em %d em_lnl %.10f\n", lnl, rell[0], nrate_cat, has_em, em_lnl);
    return 0;
}

/* golden values of the NNI evaluation (PhyloTree::getBestNNIForBran, phylotree.cpp:2873-3066) on every internal branch, in the order of
   computeBestTraversal; per branch: node1, node2 and, for both NNIs, newloglh, newLen[5] and the ids of the two swapped subtree roots */
static int do_dumpnni(int argc, char **argv) {
    g_out = argv[2];
    mkdir(g_out.c_str(), 0755);
    int nni5 = atoi(argv[3]);
    std::vector<char*> args; args.push_back(argv[0]);
    for (int i = 4; i < argc; i++) args.push_back(argv[i]);
    Params &params = Params::getInstance();
    IQTree *tree = setup((int)args.size(), &args[0], params);
    if ((int)params.nni5 != nni5) {         // the nni_partial_lh buffers are sized by params->nni5 at allocation (phylotree.cpp:857)
        params.nni5 = nni5 != 0;
        tree->deleteAllPartialLh();
        tree->initializeAllPartialLh();
    }
    tree->clearAllPartialLH();
    double lnl = tree->computeLikelihood();
    tree->curScore = lnl;
    NodeVector nodes, nodes2;
    tree->computeBestTraversal(nodes, nodes2);
    std::vector<int> ids;
    std::vector<double> vals;
    for (size_t j = 0; j < nodes.size(); j++) {
        if (nodes[j]->isLeaf() || nodes2[j]->isLeaf()) continue;
        PhyloNode *node1 = (PhyloNode*)nodes[j], *node2 = (PhyloNode*)nodes2[j];
        NNIMove mv[2];
        mv[0].ptnlh = mv[1].ptnlh = NULL;
        mv[0].node1 = NULL;
        tree->getBestNNIForBran(node1, node2, mv);
        ids.push_back(node1->id); ids.push_back(node2->id);
        for (int c = 0; c < 2; c++) {
            ids.push_back((*mv[c].node1Nei_it)->node->id); ids.push_back((*mv[c].node2Nei_it)->node->id);
            vals.push_back(mv[c].newloglh);
            for (int i = 0; i < 5; i++) vals.push_back(i < (nni5 ? 5 : 1) ? mv[c].newLen[i] : 0.0);
        }
    }
    tree->clearAllPartialLH();
    double lnl_after = tree->computeLikelihood();
    wbin("nni_ids.i32", &ids[0], ids.size() * sizeof(int));
    wbin("nni_vals.f64", &vals[0], vals.size() * 8);
    {
        std::ofstream f((g_out + "/metannni.json").c_str());
        f << std::setprecision(17);
        f << "{ \"nni5\": " << nni5 << ", \"nbranch\": " << ids.size() / 6 << ", \"lnl\": " << lnl << ", \"lnl_after\": " << lnl_after << " }\n";
    }
    printf("dumpnni nni5 %d branches %zu lnl %.10f after %.10f\n", nni5, ids.size() / 6, lnl, lnl_after);
    return 0;
}

/* Fitch parsimony: the inputs and outputs of the reference's own computeParsimony / computeParsimonyBranch on the
   fixture tree, and of computeParsimonyTree (stepwise addition) with the addition order its RNG produces for <seed> */
static int do_dumppars(int argc, char **argv) {
    g_out = argv[2];
    mkdir(g_out.c_str(), 0755);
    int seed = atoi(argv[3]);
    std::vector<char*> args; args.push_back(argv[0]);
    for (int i = 4; i < argc; i++) args.push_back(argv[i]);
    Params &params = Params::getInstance();
    IQTree *tree = setup((int)args.size(), &args[0], params);
    Alignment *aln = tree->aln;
    if (aln->ordered_pattern.empty()) aln->orderPatternByNumChars();
    const int ntaxa = aln->getNSeq();
    const size_t nsites = aln->num_informative_sites;
    std::vector<unsigned char> sites((size_t)ntaxa * nsites);
    {
        size_t site = 0;
        for (Alignment::iterator pat = aln->ordered_pattern.begin(); pat != aln->ordered_pattern.end(); pat++)
            for (int j = 0; j < pat->frequency; j++, site++)
                for (int t = 0; t < ntaxa; t++) sites[(size_t)t * nsites + site] = (unsigned char)pat->at(t);
        if (site != nsites) { fprintf(stderr, "informative site count mismatch\n"); return 4; }
    }
    wbin("pars_sites.u8", sites.empty() ? NULL : &sites[0], sites.size());
    dumpAdjacency(tree, "pars_tree.txt");
    int score = tree->computeParsimony();
    NodeVector nodes1, nodes2;
    tree->getBranches(nodes1, nodes2);
    std::vector<int> br;
    for (size_t i = 0; i < nodes1.size(); i++) {
        int subst = -1;
        int sc = tree->computeParsimonyBranch((PhyloNeighbor*)nodes1[i]->findNeighbor(nodes2[i]), (PhyloNode*)nodes1[i], &subst);
        br.push_back(nodes1[i]->id); br.push_back(nodes2[i]->id); br.push_back(sc); br.push_back(subst);
    }
    wbin("pars_branches.i32", &br[0], br.size() * sizeof(int));
    /* the addition order computeParsimonyTree will draw: same generator state, same shuffle (phylotreepars.cpp:318-323) */
    init_random(seed, false);
    IntVector order(ntaxa);
    for (int i = 0; i < ntaxa; i++) order[i] = i;
    my_random_shuffle(order.begin(), order.end());
    wbin("pars_order.i32", &order[0], order.size() * sizeof(int));
    init_random(seed, false);
    IQTree *t2 = new IQTree(aln);
    t2->setParams(&params);
    double t0 = omp_get_wtime();
    int best = t2->computeParsimonyTree(NULL, aln);
    double secs = omp_get_wtime() - t0;
    {   /* the reference fills its private taxon order from the same generator: check that ours is the one it used */
        Node *r = t2->root;
        if (r->id != order[0]) { fprintf(stderr, "taxon order mismatch: root %d vs order[0] %d\n", r->id, order[0]); return 5; }
    }
    dumpAdjacency(t2, "pars_stepwise_tree.txt");
    int check = t2->computeParsimony();
    {
        std::ofstream f((g_out + "/metapars.json").c_str());
        f << std::setprecision(17);
        f << "{ \"ntaxa\": " << ntaxa << ", \"nsites_informative\": " << nsites << ", \"aln_nsite\": " << aln->getNSite()
          << ", \"nstates\": " << aln->num_states << ", \"state_unknown\": " << (int)aln->STATE_UNKNOWN << ", \"seq_type\": " << (int)aln->seq_type
          << ", \"min_branch_length\": " << params.min_branch_length
          << ", \"score\": " << score << ", \"seed\": " << seed << ", \"stepwise_score\": " << best << ", \"stepwise_check\": " << check
          << ", \"stepwise_seconds\": " << secs << ", \"threads\": " << omp_get_max_threads() << " }\n";
    }
    printf("dumppars score %d stepwise %d (check %d) %.3f s, %zu informative sites\n", score, best, check, secs, nsites);
    return 0;
}

/* RateFree::optimizeWithEM: rates / proportions / lnL before and after the reference's own EM */
static int do_dumpem(int argc, char **argv) {
    g_out = argv[2];
    mkdir(g_out.c_str(), 0755);
    std::vector<char*> args; args.push_back(argv[0]);
    for (int i = 3; i < argc; i++) args.push_back(argv[i]);
    Params &params = Params::getInstance();
    IQTree *tree = setup((int)args.size(), &args[0], params);
    RateFree *rf = dynamic_cast<RateFree*>(tree->getRate());
    if (!rf) { fprintf(stderr, "dumpem: not a +R model\n"); return 4; }
    tree->clearAllPartialLH();
    double lnl0 = tree->computeLikelihood();
    int nc = rf->getNRate();
    std::vector<double> v;
    for (int c = 0; c < nc; c++) v.push_back(rf->getRate(c));
    for (int c = 0; c < nc; c++) v.push_back(rf->getProp(c));
    double t_em = omp_get_wtime();
    double lnl1 = rf->optimizeWithEM();
    t_em = omp_get_wtime() - t_em;
    for (int c = 0; c < nc; c++) v.push_back(rf->getRate(c));
    for (int c = 0; c < nc; c++) v.push_back(rf->getProp(c));
    wbin("em_rates_props.f64", &v[0], v.size() * 8);
    {
        std::ofstream f((g_out + "/metaem.json").c_str());
        f << std::setprecision(17);
        f << "{ \"ncat\": " << nc << ", \"lnl_before\": " << lnl0 << ", \"lnl_after\": " << lnl1 << ", \"num_precision\": " << tree->num_precision
          << ", \"em_seconds\": " << t_em << ", \"threads\": " << omp_get_max_threads() << ", \"nptn\": " << tree->aln->getNPattern() << " }\n";
    }
    printf("dumpem ncat %d lnl %.10f -> %.10f\n", nc, lnl0, lnl1);
    return 0;
}


/* the fixtures of the BASELINE-size parity tests: what do_dump writes, without the bulk (no tip states, no partial_lh values) */
static int do_dumplite(int argc, char **argv) {
    g_out = argv[2];
    mkdir(g_out.c_str(), 0755);
    size_t stride = (size_t)atol(argv[3]);
    if (stride < 1) stride = 1;
    std::vector<char*> args; args.push_back(argv[0]);
    for (int i = 4; i < argc; i++) args.push_back(argv[i]);
    Params &params = Params::getInstance();
    IQTree *tree = setup((int)args.size(), &args[0], params);
    Alignment *aln = tree->aln;
    ModelSubst *model = tree->getModel();
    RateHeterogeneity *rate = tree->getRate();
    ModelFactory *mf = tree->getModelFactory();
    size_t S = aln->num_states, C = rate->getNRate(), M = model->getNMixtures();
    bool fused = mf->fused_mix_rate;
    size_t P = aln->size(), U = mf->unobserved_ptns.size(), nptn = P + U, ntaxa = aln->getNSeq();
    size_t block = model->isMixture() && !fused ? S * C * M : S * C;
    size_t E = M;
    wbin("eval.f64", model->getEigenvalues(), E * S * sizeof(double));
    wbin("evec.f64", model->getEigenvectors(), E * S * S * sizeof(double));
    wbin("inv_evec.f64", model->getInverseEigenvectors(), E * S * S * sizeof(double));
    std::vector<double> rates(C), props(C), mixprops(M, 1.0), sfreq(S);
    for (size_t c = 0; c < C; c++) { rates[c] = rate->getRate(c); props[c] = rate->getProp(c); }
    if (model->isMixture()) for (size_t m = 0; m < M; m++) mixprops[m] = ((ModelMixture*)model)->prop[m];
    model->getStateFrequency(&sfreq[0]);
    wbin("rates.f64", &rates[0], C * 8); wbin("props.f64", &props[0], C * 8);
    wbin("mixprops.f64", &mixprops[0], M * 8); wbin("state_freq.f64", &sfreq[0], S * 8);
    dumpAdjacency(tree, "tree_adj.txt");

    tree->clearAllPartialLH();
    double lnl = tree->computeLikelihood();
    int root_from = tree->current_it_back->node->id, root_to = tree->current_it->node->id;
    std::vector<double> plh(P), samp;
    tree->computeLikelihood(&plh[0]);       // with the scale_num correction of current_it
    for (size_t p = 0; p < P; p += stride) samp.push_back(plh[p]);
    wbin("pattern_lh_scaled_sample.f64", &samp[0], samp.size() * 8);
    double freq_sum = 0.0;
    for (size_t p = 0; p < P; p++) freq_sum += tree->ptn_freq[p];

    /* scale_num of three vectors valid after the call above + strided row sums of their partial_lh */
    std::vector<PhyloNeighbor*> vecs; std::vector<std::pair<int,int> > vid;
    {
        std::vector<Node*> nodes;
        collectNodes(tree->root, NULL, nodes);
        for (size_t i = 0; i < nodes.size(); i++)
            for (size_t k = 0; k < nodes[i]->neighbors.size(); k++) {
                PhyloNeighbor *b = (PhyloNeighbor*)nodes[i]->neighbors[k];
                if ((b->partial_lh_computed & 1) && b->partial_lh && !b->node->isLeaf()) { vecs.push_back(b); vid.push_back(std::make_pair(nodes[i]->id, b->node->id)); }
            }
    }
    std::ofstream vf((g_out + "/vectors.txt").c_str());
    size_t pick[3] = {0, vecs.size() / 2, vecs.size() - 1};
    for (int q = 0; q < 3 && !vecs.empty(); q++) {
        PhyloNeighbor *b = vecs[pick[q]];
        std::vector<UBYTE> sc(nptn); std::vector<double> rs;
        std::vector<double> lh;
        const double *src = b->partial_lh;
        if (ON_DEVICE(tree)) {
#ifdef LK_CUDA_SHIM
            lh.resize(nptn * block);
            tree->lkCudaFetchVector(b, &lh[0], &sc[0]);
            src = &lh[0];
#endif
        } else memcpy(&sc[0], b->scale_num, nptn * sizeof(UBYTE));
        for (size_t p = 0; p < nptn; p += stride) { double a = 0.0; for (size_t i = 0; i < block; i++) a += src[p * block + i]; rs.push_back(a); }
        std::ostringstream nm; nm << "vec" << q;
        wbin(nm.str() + "_scale.i16", &sc[0], nptn * sizeof(UBYTE));
        wbin(nm.str() + "_rowsum_sample.f64", &rs[0], rs.size() * 8);
        vf << vid[pick[q]].first << " " << vid[pick[q]].second << "\n";
    }
    vf.close();

    /* df / ddf / lnL-from-buffer on three branches of the reference's traversal */
    NodeVector nodes, nodes2;
    tree->computeBestTraversal(nodes, nodes2);
    size_t bpick[3] = {0, nodes.size() / 2, nodes.size() - 1};
    std::vector<int> bid; std::vector<double> dv;
    for (int q = 0; q < 3; q++) {
        size_t j = bpick[q];
        double a, b;
        tree->theta_computed = false;
        PhyloNeighbor *nb = (PhyloNeighbor*)nodes[j]->findNeighbor(nodes2[j]);
        tree->computeLikelihoodDerv(nb, (PhyloNode*)nodes[j], a, b);
        tree->current_it = nb;
        tree->current_it_back = (PhyloNeighbor*)nodes2[j]->findNeighbor(nodes[j]);
        double fb = tree->computeLikelihoodFromBuffer();
        bid.push_back(nodes[j]->id); bid.push_back(nodes2[j]->id);
        dv.push_back(a); dv.push_back(b); dv.push_back(fb);
    }
    wbin("branch_ids.i32", &bid[0], bid.size() * sizeof(int));
    wbin("branch_derv.f64", &dv[0], dv.size() * 8);
    {
        std::ofstream f((g_out + "/meta.json").c_str());
        f << std::setprecision(17);
        f << "{ \"nstates\": " << S << ", \"ncat\": " << C << ", \"nmix\": " << M << ", \"family\": \"" << familyName(tree) << "\", \"block\": " << block
          << ", \"nptn\": " << P << ", \"nunobs\": " << U << ", \"ntaxa\": " << ntaxa << ", \"nsite\": " << aln->getNSite() << ", \"freq_sum\": " << freq_sum
          << ", \"state_unknown\": " << (int)aln->STATE_UNKNOWN << ", \"seq_type\": " << (int)aln->seq_type << ", \"p_invar\": " << rate->getPInvar()
          << ", \"min_branch_length\": " << params.min_branch_length << ", \"max_branch_length\": " << params.max_branch_length
          << ", \"root_branch\": [" << root_from << ", " << root_to << "], \"lnl\": " << lnl << ", \"stride\": " << stride
          << ", \"kernel\": \"" << (ON_DEVICE(tree) ? "cuda" : (tree->sse == LK_EIGEN_SSE ? (instruction_set >= 7 ? "AVX" : "SSE") : "generic"))
          << "\", \"threads\": " << params.num_threads << " }\n";
    }
    printf("dumplite lnl %.10f, %zu vectors valid, %zu pattern samples\n", lnl, vecs.size(), samp.size());
    return 0;
}

/* the NaN / Inf repair of the branch kernels (phylokernel.h:848-866): zero three patterns' rows of the root branch's vector, so that
   their likelihood is 0 and log() gives -inf, and let the reference's own kernel repair them (CPU buffers only) */
static int do_dumprepair(int argc, char **argv) {
    g_out = argv[2];
    mkdir(g_out.c_str(), 0755);
    std::vector<char*> args; args.push_back(argv[0]);
    for (int i = 3; i < argc; i++) args.push_back(argv[i]);
    Params &params = Params::getInstance();
    IQTree *tree = setup((int)args.size(), &args[0], params);
    if (ON_DEVICE(tree)) { fprintf(stderr, "dumprepair: host buffers only\n"); return 4; }
    size_t P = tree->aln->size();
    size_t block = tree->getPartialLhBytes() / sizeof(double) / ((P + tree->aln->num_states + 3) / 4 * 4);
    tree->clearAllPartialLH();
    double lnl0 = tree->computeLikelihood();
    PhyloNeighbor *it = tree->current_it;
    PhyloNode *dad = (PhyloNode*)tree->current_it_back->node;
    if (!it->partial_lh) { fprintf(stderr, "dumprepair: root branch vector missing\n"); return 5; }
    int zp[3] = {0, (int)(P / 2), (int)P - 1};
    for (int q = 0; q < 3; q++) memset(it->partial_lh + (size_t)zp[q] * block, 0, block * sizeof(double));
    double lnl1 = tree->computeLikelihoodBranch(it, dad);
    wbin("pattern_lh_repaired.f64", tree->_pattern_lh, P * 8);
    {
        std::ofstream f((g_out + "/metarepair.json").c_str());
        f << std::setprecision(17);
        f << "{ \"lnl\": " << lnl0 << ", \"lnl_repaired\": " << lnl1 << ", \"zeroed\": [" << zp[0] << ", " << zp[1] << ", " << zp[2] << "]"
          << ", \"vector\": [" << dad->id << ", " << it->node->id << "], \"block\": " << block << " }\n";
    }
    printf("dumprepair lnl %.10f -> %.10f\n", lnl0, lnl1);
    return 0;
}

static int do_time(int argc, char **argv) {
    int lnl_reps = atoi(argv[2]), derv_reps = atoi(argv[3]), oab_reps = atoi(argv[4]);
    std::vector<char*> args; args.push_back(argv[0]);
    for (int i = 5; i < argc; i++) args.push_back(argv[i]);
    Params &params = Params::getInstance();
    double t0 = getRealTime();
    IQTree *tree = setup((int)args.size(), &args[0], params);
    double t_setup = getRealTime() - t0;
    Alignment *aln = tree->aln;
    size_t C = tree->getRate()->getNRate(), M = tree->getModel()->getNMixtures();
    bool fused = tree->getModelFactory()->fused_mix_rate;
    size_t P = aln->size(), ntaxa = aln->getNSeq();
    size_t blocks = (tree->getModel()->isMixture() && !fused) ? C * M : C;

    tree->clearAllPartialLH();
    double lnl = tree->computeLikelihood();           // warm-up
    PhyloNode *n1 = (PhyloNode*)tree->root, *n2 = (PhyloNode*)tree->root->neighbors[0]->node;
    double df = 0, ddf = 0;
    if (tree->isSuperTree()) {
        /* PartitionModel: full-tree lnL of all partitions (PhyloSuperTree::computeLikelihood, phylosupertree.cpp:955-980) */
        std::vector<double> t;
        for (int r = 0; r < lnl_reps; r++) {
            double a = getRealTime();
            tree->clearAllPartialLH();
            lnl = tree->computeLikelihood();
            t.push_back(getRealTime() - a);
        }
        size_t warm = getenv("LK_REF_WARMUP") ? (size_t)atoi(getenv("LK_REF_WARMUP")) : 0;
        if (warm >= t.size()) warm = 0;
        double sum = 0;
        for (size_t i = warm; i < t.size(); i++) sum += t[i];
        PhyloSuperTree *st = (PhyloSuperTree*)tree;
        size_t units = 0, ptns = 0;
        bool dev = false;
        for (PhyloSuperTree::iterator it = st->begin(); it != st->end(); it++) {
            size_t k = (*it)->getRate()->getNRate() * ((*it)->getModel()->isMixture() && !(*it)->getModelFactory()->fused_mix_rate ? (*it)->getModel()->getNMixtures() : 1);
            units += (size_t)((*it)->leafNum - 2) * (*it)->aln->size() * k;
            ptns += (*it)->aln->size();
            dev = dev || ON_DEVICE(*it);
        }
        printf("{\"ref_time\": true, \"kernel\": \"%s\", \"threads\": %d, \"partitions\": %zu, \"nptn\": %zu, \"ntaxa\": %d, \"updates_per_eval\": %zu, "
               "\"lnl\": %.17g, \"lnl_reps\": %d, \"lnl_s_mean\": %.6g, \"setup_s\": %.4g}\n",
               dev ? "cuda" : "cpu", params.num_threads, st->size(), ptns, tree->leafNum, units, lnl, (int)(t.size() - warm),
               t.size() > warm ? sum / (t.size() - warm) : 0.0, t_setup);
        return 0;
    }
    /* one "step" = full-tree lnL (all partials recomputed) + theta_all rebuild + df/ddf on the root-leaf branch */
    std::vector<double> lnl_t, step_t;
    for (int r = 0; r < lnl_reps; r++) {
        double a = getRealTime();
        tree->clearAllPartialLH();
        lnl = tree->computeLikelihood();
        double b = getRealTime();
        tree->theta_computed = false;
        tree->computeLikelihoodDerv((PhyloNeighbor*)n1->findNeighbor(n2), n1, df, ddf);
        double c = getRealTime();
        lnl_t.push_back(b - a);
        step_t.push_back(c - a);
    }
    tree->theta_computed = false;
    tree->computeLikelihoodDerv((PhyloNeighbor*)n1->findNeighbor(n2), n1, df, ddf);   // builds theta
    double a = getRealTime();
    for (int r = 0; r < derv_reps; r++)
        tree->computeLikelihoodDerv((PhyloNeighbor*)n1->findNeighbor(n2), n1, df, ddf);
    double t_derv = derv_reps ? (getRealTime() - a) / derv_reps : 0.0;
    double oab = 0, t_oab = 0;
    for (int r = 0; r < oab_reps; r++) {
        a = getRealTime();
        oab = tree->optimizeAllBranches(1);
        t_oab += getRealTime() - a;
    }
    if (oab_reps) t_oab /= oab_reps;
    /* LK_REF_NNI=n: time PhyloTree::getBestNNIForBran (both NNIs, 5 branches each) on the first n internal branches of the traversal */
    int nni_n = getenv("LK_REF_NNI") ? atoi(getenv("LK_REF_NNI")) : 0, nni_done = 0;
    double t_nni = 0.0;
    if (nni_n > 0) {
        tree->clearAllPartialLH();
        tree->curScore = tree->computeLikelihood();
        NodeVector nodes, nodes2;
        tree->computeBestTraversal(nodes, nodes2);
        for (size_t j = 0; j < nodes.size() && nni_done < nni_n; j++) {
            if (nodes[j]->isLeaf() || nodes2[j]->isLeaf()) continue;
            NNIMove mv[2];
            mv[0].ptnlh = mv[1].ptnlh = NULL; mv[0].node1 = NULL;
            a = getRealTime();
            tree->getBestNNIForBran((PhyloNode*)nodes[j], (PhyloNode*)nodes2[j], mv);
            t_nni += getRealTime() - a;
            nni_done++;
        }
        if (nni_done) t_nni /= nni_done;
    }
    /* under -cuda: the same steps once more with every launch bracketed by CUDA events -> device time per kernel kind */
    std::ostringstream kstats;
    int profiled_steps = 0;
    kstats << "{";
#ifdef LK_CUDA_SHIM
    if (ON_DEVICE(tree) && lnl_reps > 0) {
        static const char *kind_name[] = {"prep", "partial", "branch", "theta", "derv", "frombuf", "newton", "misc", "rell", "em"};
        tree->lkCudaProfile(true, true);
        profiled_steps = lnl_reps;
        for (int r = 0; r < lnl_reps; r++) {
            tree->clearAllPartialLH();
            tree->computeLikelihood();
            tree->theta_computed = false;
            double df2, ddf2;
            tree->computeLikelihoodDerv((PhyloNeighbor*)n1->findNeighbor(n2), n1, df2, ddf2);
        }
        bool first = true;
        for (int k = 0; k < 10; k++) {
            uint64_t nl, nb, nf; double ms;
            if (!tree->lkCudaKernelStats(k, nl, nb, nf, ms) || !nl) continue;
            kstats << (first ? "" : ", ") << "\"" << kind_name[k] << "\": {\"launches\": " << nl << ", \"bytes\": " << nb << ", \"flops\": " << nf
                   << ", \"device_ms\": " << std::setprecision(8) << ms << "}";
            first = false;
        }
        tree->lkCudaProfile(false, false);
    }
#endif
    kstats << "}";
    /* LK_REF_WARMUP=w: the first w steps are untimed warm-up (bench.py --warmup) */
    size_t warm = getenv("LK_REF_WARMUP") ? (size_t)atoi(getenv("LK_REF_WARMUP")) : 0;
    if (warm >= lnl_t.size()) warm = 0;
    double sum = 0, best = 1e300;
    for (size_t i = warm; i < lnl_t.size(); i++) { sum += lnl_t[i]; best = std::min(best, lnl_t[i]); }
    double mean = lnl_t.empty() ? 0 : sum / (lnl_t.size() - warm);
    double step_sum = 0;
    for (size_t i = warm; i < step_t.size(); i++) step_sum += step_t[i];
    double step_mean = step_t.empty() ? 0 : step_sum / (step_t.size() - warm);
    lnl_reps -= (int)warm;
    printf("{\"ref_time\": true, \"kernel\": \"%s\", \"threads\": %d, \"nptn\": %zu, \"ntaxa\": %zu, \"blocks\": %zu, "
           "\"lnl\": %.17g, \"lnl_reps\": %d, \"lnl_s_mean\": %.6g, \"lnl_s_best\": %.6g, \"step_s_mean\": %.6g, \"step_s_total\": %.6g, "
           "\"df\": %.17g, \"ddf\": %.17g, \"derv_reps\": %d, \"derv_s_mean\": %.6g, "
           "\"oab_lnl\": %.12f, \"oab_reps\": %d, \"oab_s_mean\": %.6g, \"nni_branches\": %d, \"nni_s_mean\": %.6g, \"setup_s\": %.4g, \"profiled_steps\": %d, \"kernels\": %s}\n",
           ON_DEVICE(tree) ? "cuda" : (tree->sse == LK_EIGEN_SSE ? (instruction_set >= 7 ? "AVX" : "SSE") : "generic"),
           params.num_threads, P, ntaxa, blocks, lnl, lnl_reps, mean, lnl_t.empty() ? 0.0 : best, step_mean, step_sum,
           df, ddf, derv_reps, t_derv, oab, oab_reps, t_oab, nni_done, t_nni, t_setup, profiled_steps, kstats.str().c_str());
    return 0;
}

int iqtree_main(int argc, char *argv[]);     /* the reference's own main() (pda.cpp compiled with -Dmain=iqtree_main) */

int main(int argc, char **argv) {
    /* ref_harness iqtree <iqtree args>: the complete IQ-TREE command line (tree search, ModelFinder, ...) of the unmodified reference */
    if (argc >= 2 && !strcmp(argv[1], "iqtree")) { argv[1] = argv[0]; return iqtree_main(argc - 1, argv + 1); }
    if (argc >= 4 && !strcmp(argv[1], "dump")) return do_dump(argc, argv);
    if (argc >= 6 && !strcmp(argv[1], "time")) return do_time(argc, argv);
    if (argc >= 5 && !strcmp(argv[1], "dumpx")) return do_dumpx(argc, argv);
    if (argc >= 5 && !strcmp(argv[1], "dumpnni")) return do_dumpnni(argc, argv);
    if (argc >= 5 && !strcmp(argv[1], "dumppars")) return do_dumppars(argc, argv);
    if (argc >= 4 && !strcmp(argv[1], "dumpem")) return do_dumpem(argc, argv);
    if (argc >= 5 && !strcmp(argv[1], "dumplite")) return do_dumplite(argc, argv);
    if (argc >= 4 && !strcmp(argv[1], "dumprepair")) return do_dumprepair(argc, argv);
    fprintf(stderr, "usage: %s dump <outdir> <iqtree args>\n       %s time <lnl_reps> <derv_reps> <oab_reps> <iqtree args>\n", argv[0], argv[0]);
    return 2;
}
