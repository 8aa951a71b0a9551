/*
 * lk_rate_em.cpp — SURVEY 8(f).1: RateFree::optimizeWithEM (model/ratefree.cpp:450-578) on the engine.
 * The EM of Wang, Li, Susko & Roger (2008) for the +R model: per step one per-category likelihood pass (computePatternLhCat), the E-step
 * over P x C doubles, and per category a copy of the tree with ONE rate category whose pattern weights are that category's posteriors and
 * whose total length is re-optimised by Brent (optimizeTreeLengthScaling) - dozens of full-tree likelihoods per category.  In the
 * reference the P x C matrix and every posterior column live on the host; here the matrix never leaves the device:
 *   E-step sums      lkcuda_lh_cat_em                  (ratefree.cpp:487-503 -> new_prop)
 *   posterior column lkcuda_lh_cat_posterior_weights   (ratefree.cpp:547-549 -> ptn_freq of the one-category tree's engine)
 * Host statements mirrored one for one (they decide the result): the M-step with MIN_PROP regularisation (ratefree.cpp:505-531), the
 * per-category loop (:533-570), MTree::copyTree's round trip through Newick TEXT with `num_precision` digits (mtree.cpp:81-86, 291-379:
 * the copy's branch lengths are the printed ones), MTree::scaleLength (mtree.cpp:1175-1186), optimizeTreeLengthScaling /
 * computeFunction (phylotree.cpp:2106-2133), Optimization::minimizeOneDimen and brent_opt (optimization.cpp:183-339).
 */
#include "lk_phylotree.h"
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <memory>

namespace lk {

/* MTree::printTree(out, WT_BR_LEN) - fixed notation, num_precision digits (mtree.cpp:245-268, 291-379) */
static void printTreeRec(std::string &out, const Node *node, const Node *dad, int precision) {
    char buf[64];
    if (node->isLeaf()) {
        out += node->name;
        snprintf(buf, sizeof buf, ":%.*f", precision, node->neighbors[0]->length);
        out += buf;
        return;
    }
    out += "(";
    bool first = true;
    double length = 0.0;
    for (size_t i = 0; i < node->neighbors.size(); i++) {
        const Neighbor *it = node->neighbors[i];
        if (it->node == dad) { length = it->length; continue; }
        if (!first) out += ",";
        printTreeRec(out, it->node, node, precision);
        first = false;
    }
    out += ")";
    out += node->name;
    if (dad != NULL || length > 0.0) {
        snprintf(buf, sizeof buf, ":%.*f", precision, length);
        out += buf;
    }
}

std::string PhyloTree::printTreeFixed(int precision) const {
    std::string out;
    if (!root) return out;
    if (root->isLeaf()) printTreeRec(out, root->neighbors[0]->node, NULL, precision);
    else printTreeRec(out, root, NULL, precision);
    out += ";";
    return out;
}

void PhyloTree::scaleLength(double norm) {
    for (size_t i = 0; i < allNodes.size(); i++)
        for (size_t k = 0; k < allNodes[i]->neighbors.size(); k++) allNodes[i]->neighbors[k]->length *= norm;
}

/* PhyloTree::computeFunction with is_opt_scaling (phylotree.cpp:2125-2132) */
double PhyloTree::computeFunctionScaling(double value) {
    if (value != current_scaling_) {
        scaleLength(value / current_scaling_);
        current_scaling_ = value;
        clearAllPartialLH();
    }
    return -computeLikelihood();
}

#define LK_ITMAX 100
#define LK_CGOLD 0.3819660
#define LK_ZEPS 1.0e-10
#define LK_SIGN(a, b) ((b) >= 0.0 ? fabs(a) : -fabs(a))

/* Optimization::brent_opt, optimization.cpp:183-274 */
double PhyloTree::brentOpt(double ax, double bx, double cx, double tol, double *foptx, double *f2optx, double fax, double fbx, double fcx) {
    double a, b, d = 0, etemp, fu, fv, fw, fx, p, q, r, tol1, tol2, u, v, w, x, xm;
    double xw, wv, vx;
    double e = 0.0;
    a = (ax < cx ? ax : cx);
    b = (ax > cx ? ax : cx);
    x = bx;
    fx = fbx;
    if (fax < fcx) { w = ax; fw = fax; v = cx; fv = fcx; }
    else { w = cx; fw = fcx; v = ax; fv = fax; }
    for (int iter = 1; iter <= LK_ITMAX; iter++) {
        xm = 0.5 * (a + b);
        tol2 = 2.0 * (tol1 = tol * fabs(x) + LK_ZEPS);
        if (fabs(x - xm) <= (tol2 - 0.5 * (b - a))) {
            *foptx = fx;
            xw = x - w; wv = w - v; vx = v - x;
            *f2optx = 2.0 * (fv * xw + fx * wv + fw * vx) / (v * v * xw + x * x * wv + w * w * vx);
            return x;
        }
        if (fabs(e) > tol1) {
            r = (x - w) * (fx - fv);
            q = (x - v) * (fx - fw);
            p = (x - v) * q - (x - w) * r;
            q = 2.0 * (q - r);
            if (q > 0.0) p = -p;
            q = fabs(q);
            etemp = e;
            e = d;
            if (fabs(p) >= fabs(0.5 * q * etemp) || p <= q * (a - x) || p >= q * (b - x))
                d = LK_CGOLD * (e = (x >= xm ? a - x : b - x));
            else {
                d = p / q;
                u = x + d;
                if (u - a < tol2 || b - u < tol2) d = LK_SIGN(tol1, xm - x);
            }
        } else {
            d = LK_CGOLD * (e = (x >= xm ? a - x : b - x));
        }
        u = (fabs(d) >= tol1 ? x + d : x + LK_SIGN(tol1, d));
        fu = computeFunctionScaling(u);
        if (fu <= fx) {
            if (u >= x) a = x; else b = x;
            v = w; w = x; x = u;
            fv = fw; fw = fx; fx = fu;
        } else {
            if (u < x) a = u; else b = u;
            if (fu <= fw || w == x) { v = w; w = u; fv = fw; fw = fu; }
            else if (fu <= fv || v == x || v == w) { v = u; fv = fu; }
        }
    }
    *foptx = fx;
    xw = x - w; wv = w - v; vx = v - x;
    *f2optx = 2.0 * (fv * xw + fx * wv + fw * vx) / (v * v * xw + x * x * wv + w * w * vx);
    return x;
}

/* Optimization::minimizeOneDimen, optimization.cpp:288-339 */
double PhyloTree::minimizeOneDimen(double xmin, double xguess, double xmax, double tolerance, double *fx, double *f2x) {
    double eps, optx, ax, bx, cx, fa, fb, fc;
    if (xguess < xmin) xguess = xmin;
    if (xguess > xmax) xguess = xmax;
    eps = xguess * tolerance * 50.0;
    ax = xguess - eps;
    if (ax < xmin) ax = xmin;
    bx = xguess;
    cx = xguess + eps;
    if (cx > xmax) cx = xmax;
    fb = computeFunctionScaling(bx);
    fa = computeFunctionScaling(ax);
    fc = computeFunctionScaling(cx);
    if ((fa < fb) || (fc < fb)) {
        if (ax != xmin) fa = computeFunctionScaling(xmin);
        if (cx != xmax) fc = computeFunctionScaling(xmax);
        ax = xmin;
        cx = xmax;
    }
    optx = brentOpt(ax, bx, cx, tolerance, fx, f2x, fa, fb, fc);
    if (*fx > fb) {          // if worse, return initial value
        *fx = computeFunctionScaling(bx);
        return bx;
    }
    return optx;
}

/* phylotree.cpp:2103-2118 */
double PhyloTree::optimizeTreeLengthScaling(double min_scaling, double &scaling, double max_scaling, double gradient_epsilon) {
    const double TOL_TREE_LENGTH_SCALE = 0.001;
    current_scaling_ = scaling;
    double negative_lh, ferror;
    scaling = minimizeOneDimen(std::min(scaling, min_scaling), scaling, std::max(max_scaling, scaling), std::max(TOL_TREE_LENGTH_SCALE, gradient_epsilon), &negative_lh, &ferror);
    if (scaling != current_scaling_) {
        scaleLength(scaling / current_scaling_);
        current_scaling_ = scaling;
        clearAllPartialLH();
    }
    return computeLikelihood();
}

/* RateFree::optimizeWithEM, model/ratefree.cpp:450-578 */
double PhyloTree::optimizeRateFreeEM(int copy_precision) {
    if (family != FAMILY_PLAIN && family != FAMILY_MIXRATE) throw std::runtime_error("optimizeRateFreeEM: a +R model (plain or fused mixture) is needed");
    if (p_invar_ != 0.0) throw std::runtime_error("optimizeRateFreeEM: +I+R is not mirrored (no reference values: the reference's own EM asserts at the default +I+R start)");
    if (nunobs) throw std::runtime_error("optimizeRateFreeEM: the reference turns the EM off for +ASC (ratefree.cpp:211)");
    const int nmixc = ncat;
    const double MIN_PROP = 1e-4;
    if (engines_.empty()) initializeAllPartialLh();
    std::vector<double> new_prop(nmixc), part(nmixc);
    std::vector<double> &prop = props_, &rates = rates_;

    /* the one-category copy: `tree` with RateHeterogeneity (1 category, rate 1, no invariant sites), ratefree.cpp:458-472 */
    std::unique_ptr<PhyloTree> tree(new PhyloTree());
    tree->setAlignment(ntaxa, nptn, &states_[0], &freq_global_[0], seq_names, seq_type, nstates, state_unknown);
    tree->setDevices(devices_);
    tree->setProcessShard(proc_rank_, proc_world_);
    if (allreduce_) tree->setAllReduce(allreduce_, allreduce_user_);
    tree->min_branch_length = min_branch_length; tree->max_branch_length = max_branch_length;
    tree->generic_kernel = generic_kernel;
    tree->keep_engines_ = true;
    const double one = 1.0;
    const size_t S = (size_t)nstates;

    for (int step = 0; step < ncat; step++) {
        enqueuePatternLhCat();                               // phylo_tree->computePatternLhCat(WSL_RATECAT)
        int ncol = collapsePatternLhCat(2);
        if (ncol != nmixc) throw std::runtime_error("optimizeRateFreeEM: unexpected column count");
        std::fill(new_prop.begin(), new_prop.end(), 0.0);
        for (size_t d = 0; d < engines_.size(); d++) {       // E-step; the reference leaves ptn_invar out of lk_ptn, also for +I+R (RateFreeInvar)
            check(lkcuda_lh_cat_em_ex(engines_[d], NULL, 0, &part[0]), engines_[d], "lkcuda_lh_cat_em_ex");
            for (int c = 0; c < nmixc; c++) new_prop[c] += part[c];
        }
        if (allreduce_) allreduce_(allreduce_user_, &new_prop[0], NULL, nmixc);
        // M-step
        int maxpropid = 0;
        for (int c = 0; c < nmixc; c++) {
            new_prop[c] = new_prop[c] / (double)nsite;
            if (new_prop[c] > new_prop[maxpropid]) maxpropid = c;
        }
        bool zero_prop = false;
        for (int c = 0; c < nmixc; c++) {
            if (new_prop[c] < MIN_PROP) {
                new_prop[maxpropid] -= (MIN_PROP - new_prop[c]);
                new_prop[c] = MIN_PROP;
                zero_prop = true;
            }
        }
        if (zero_prop) break;
        bool converged = true;
        for (int c = 0; c < nmixc; c++) {
            converged = converged && (fabs(prop[c] - new_prop[c]) < 1e-4);
            prop[c] = new_prop[c];
        }
        // now optimize rates one by one
        const std::string copy = printTreeFixed(copy_precision);
        for (int c = 0; c < nmixc; c++) {
            tree->readTreeString(copy);                      // tree->copyPhyloTree(phylo_tree)
            const size_t m = family == FAMILY_MIXRATE ? (size_t)c : 0;
            tree->setModel(FAMILY_PLAIN, 1, 1, &eval_[m * S], &evec_[m * S * S], &inv_evec_[m * S * S], &one, &one, NULL, &state_freq_[0], 0.0);
            tree->initializeAllPartialLh();
            tree->ensureInputs();                            // computePtnFreq + uploads; then the posteriors replace ptn_freq on the device
            for (size_t d = 0; d < engines_.size(); d++)
                check(lkcuda_lh_cat_posterior_weights(engines_[d], c, 0, tree->engines_[d]), engines_[d], "lkcuda_lh_cat_posterior_weights");
            double scaling = rates[c];
            tree->scaleLength(scaling);
            tree->optimizeTreeLengthScaling(MIN_PROP, scaling, 1.0 / prop[c], 0.001);
            converged = converged && (fabs(rates[c] - scaling) < 1e-4);
            rates[c] = scaling;
        }
        clearAllPartialLH();
        model_dirty_ = true;                                 // rates and proportions changed: the val tables and eleft/eright are rebuilt
        if (converged) break;
    }
    clearAllPartialLH();
    model_dirty_ = true;
    return computeLikelihood();
}

}  // namespace lk
