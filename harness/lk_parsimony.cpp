/*
 * lk_parsimony.cpp — host mirror of the reference's Fitch-parsimony callers, written against include/lkpars.h.
 * The reference's control flow decides WHICH partials are recomputed and WHICH branch wins a tie; it is kept statement by statement:
 *   initializeAllPartialPars   phylotree.cpp:596-624       block indices in visiting order, two per branch
 *   computePartialParsimony    phylotreepars.cpp:18-26, 150-161   dirty bit 2 of partial_lh_computed, post-order recursion
 *   computeParsimonyBranch     phylotreepars.cpp:218-282
 *   computeParsimonyTree       phylotreepars.cpp:309-426    + addTaxonMPFast :428-465
 *   fixNegativeBranch          phylotree.cpp:2654-2694
 * What changes is WHEN the device is asked: the recursion only queues ops; a whole stepwise-addition round (every dirty partial and
 * the 2k-3 candidate insertions) or a whole fixNegativeBranch pass is ONE lkpars_run.  The reference's per-vector score word
 * (partial_pars[nstates*nsites]) lives in pars_score_ and is resolved from the returned counts in op order.
 */
#include "lk_phylotree.h"
#include <algorithm>
#include <climits>
#include <cmath>

namespace lk {

void PhyloTree::parsCheck(int rc, const char *what) {
    if (rc == LKCUDA_OK) return;
    throw std::runtime_error(std::string(what) + ": " + lkpars_last_error(pars_) + " (status " + std::to_string(rc) + ")");
}

void PhyloTree::freeNodes() {
    for (size_t i = 0; i < allNodes.size(); i++) {
        for (size_t k = 0; k < allNodes[i]->neighbors.size(); k++) delete allNodes[i]->neighbors[k];
        delete allNodes[i];
    }
    allNodes.clear();
    root = NULL;
    current_it = current_it_back = NULL;
}

void PhyloTree::setParsimonySites(int64_t nsites, const uint8_t *site_states, int64_t aln_nsite) {
    deleteAllPartialPars();
    if (ntaxa <= 0) throw std::runtime_error("setParsimonySites: setAlignment first");
    pars_nsites_ = nsites; pars_aln_nsite_ = aln_nsite;
    pars_states_.assign(site_states, site_states + (size_t)ntaxa * (size_t)nsites);
}

void PhyloTree::deleteAllPartialPars() {
    if (pars_) { lkpars_destroy(pars_); pars_ = NULL; }
    pars_ops_.clear();
    pars_score_.clear();
}

/* the state sets of the codes, phylotreepars.cpp:36, 45-68 (DNA), 82-108 (protein), 122-138 (others) */
static void parsStateMasks(int seq_type, int nstates, int state_unknown, std::vector<uint64_t> &mask) {
    mask.assign((size_t)std::max(state_unknown, nstates) + 1, 0);
    for (int s = 0; s < nstates; s++) mask[s] = 1ull << s;
    const uint64_t all = nstates == 64 ? ~0ull : (1ull << nstates) - 1;
    if (seq_type == SEQ_DNA && nstates == 4)
        for (int s = 4; s < state_unknown && s < (int)mask.size(); s++) mask[s] = (uint64_t)(s - 3) & 15ull;
    if (seq_type == SEQ_PROTEIN && nstates == 20) {
        static const int ambi_aa[] = {2, 3, 5, 6, 9, 10};
        for (int s = 20; s < 23 && s < (int)mask.size(); s++) mask[s] = (1ull << ambi_aa[(s - 20) * 2]) | (1ull << ambi_aa[(s - 20) * 2 + 1]);
    }
    mask[state_unknown] = all;
}

/* phylotree.cpp:596-624: (aln->getNSeq() - 1) * 4 blocks, two per branch in root-first visiting order; every flag cleared */
static void assignParsVectors(int &index, Node *node, Node *dad) {
    if (dad) {
        node->findNeighbor(dad)->pars_vec = index;
        dad->findNeighbor(node)->pars_vec = index + 1;
        index += 2;
    }
    for (size_t i = 0; i < node->neighbors.size(); i++)
        if (node->neighbors[i]->node != dad) assignParsVectors(index, node->neighbors[i]->node, node);
}

void PhyloTree::initializeAllPartialPars() {
    if (!root) throw std::runtime_error("initializeAllPartialPars: no tree");
    if (pars_states_.empty() && pars_nsites_ != 0) throw std::runtime_error("initializeAllPartialPars: setParsimonySites first");
    if (!pars_) {
        std::vector<uint64_t> mask;
        parsStateMasks(seq_type, nstates, state_unknown, mask);
        pars_nvec_ = (ntaxa - 1) * 4 + 1;                        /* central_partial_pars (phylotree.cpp:602) + tmp_partial_pars */
        int dev = pars_device >= 0 ? pars_device : (devices_.empty() ? 0 : devices_[0]);
        static const uint8_t none = 0;
        int rc = lkpars_create(dev, nstates, ntaxa, pars_nsites_, pars_states_.empty() ? &none : &pars_states_[0], &mask[0], (int)mask.size(), pars_nvec_, &pars_);
        if (rc != LKCUDA_OK) throw std::runtime_error(std::string("lkpars_create: ") + lkpars_last_error(NULL));
        pars_score_.assign((size_t)ntaxa + pars_nvec_, 0);
    }
    int index = 0;
    assignParsVectors(index, root, NULL);
    clearAllPartialLH();
    pars_ops_.clear();
}

int PhyloTree::parsVec(const Neighbor *nei) const {
    if (nei->node->isLeaf()) return nei->node->id;              /* the tip vectors are built once, at lkpars_create */
    if (nei->pars_vec < 0) throw std::runtime_error("parsimony: neighbour without a partial_pars block");
    return ntaxa + nei->pars_vec;
}

int PhyloTree::parsQueue(int kind, int dst, int a, int b, int c) {
    lkpars_op op; op.kind = kind; op.dst = dst; op.a = a; op.b = b; op.c = c;
    pars_ops_.push_back(op);
    return (int)pars_ops_.size() - 1;
}

/* everything queued goes out as one call: the ops that write (in queue order, they depend on each other), then the read-only ones
   (independent: the engine runs them as one wide launch).  No vector is written twice or after being read within a queue, because a
   queue never spans a change of the tree. */
void PhyloTree::parsFlush() {
    const size_t n = pars_ops_.size();
    pars_counts_.assign(2 * n, 0);
    if (n == 0) return;
    /* level of a writing op = 1 + the highest level among the queued ops that produce its operands; ops of one level are
       independent, so ordering the writers by level lets the engine overlap their memory round trips */
    std::vector<int> level(n, 0), order;
    {
        std::vector<int> producer(pars_score_.size(), -1);
        int maxlevel = 0;
        for (size_t i = 0; i < n; i++) {
            const lkpars_op &op = pars_ops_[i];
            if (!(op.kind == LKPARS_OP_PARTIAL || (op.kind == LKPARS_OP_INSERT && op.dst >= 0))) continue;
            int l = 0;
            if (producer[op.a] >= 0) l = std::max(l, level[producer[op.a]]);
            if (producer[op.b] >= 0) l = std::max(l, level[producer[op.b]]);
            if (op.kind == LKPARS_OP_INSERT && producer[op.c] >= 0) l = std::max(l, level[producer[op.c]]);
            level[i] = l + 1;
            producer[op.dst] = (int)i;
            maxlevel = std::max(maxlevel, l + 1);
        }
        order.reserve(n);
        for (int l = 1; l <= maxlevel; l++) for (size_t i = 0; i < n; i++) if (level[i] == l) order.push_back((int)i);
        for (size_t i = 0; i < n; i++) if (level[i] == 0) order.push_back((int)i);
    }
    std::vector<lkpars_op> sorted(n);
    for (size_t i = 0; i < n; i++) sorted[i] = pars_ops_[order[i]];
    std::vector<uint32_t> counts(2 * n);
    parsCheck(lkpars_run(pars_, &sorted[0], (int)n, &counts[0]), "lkpars_run");
    for (size_t i = 0; i < n; i++) { pars_counts_[2 * order[i]] = counts[2 * i]; pars_counts_[2 * order[i] + 1] = counts[2 * i + 1]; }
    /* partial_pars[nstates*nsites] = score + left + right, phylotreepars.cpp:213 */
    for (size_t i = 0; i < n; i++) {
        const lkpars_op &op = pars_ops_[i];
        if (op.kind != LKPARS_OP_BRANCH && op.dst >= 0) pars_score_[op.dst] = pars_counts_[2 * i] + pars_score_[op.a] + pars_score_[op.b];
    }
    pars_ops_.clear();
}

/* phylotreepars.cpp:18-26, 150-161: a leaf's vector always exists here, so only the flag is set */
void PhyloTree::computePartialParsimony(Neighbor *dad_branch, Node *dad) {
    if (dad_branch->partial_lh_computed & 2) return;
    Node *node = dad_branch->node;
    dad_branch->partial_lh_computed |= 2;
    if (node->isLeaf() && dad) return;
    if (node->degree() != 3) throw std::runtime_error("parsimony: it works only for strictly bifurcating tree");
    Neighbor *left = NULL, *right = NULL;
    for (size_t i = 0; i < node->neighbors.size(); i++) {
        Neighbor *pit = node->neighbors[i];
        if (pit->node == dad) continue;
        if ((pit->partial_lh_computed & 2) == 0) computePartialParsimony(pit, node);
        if (!left) left = pit; else right = pit;
    }
    parsQueue(LKPARS_OP_PARTIAL, parsVec(dad_branch), parsVec(left), parsVec(right), -1);
}

int PhyloTree::computeParsimonyBranch(Neighbor *dad_branch, Node *dad, int *branch_subst) {
    Node *node = dad_branch->node;
    Neighbor *node_branch = node->findNeighbor(dad);
    if (!node_branch) throw std::runtime_error("computeParsimonyBranch: not a branch");
    if (!pars_) initializeAllPartialPars();
    if ((dad_branch->partial_lh_computed & 2) == 0) computePartialParsimony(dad_branch, dad);
    if ((node_branch->partial_lh_computed & 2) == 0) computePartialParsimony(node_branch, node);
    const int a = parsVec(dad_branch), b = parsVec(node_branch);
    const int op = parsQueue(LKPARS_OP_BRANCH, -1, a, b, -1);
    parsFlush();
    const uint32_t sum_end_node = pars_score_[a] + pars_score_[b];
    if (branch_subst) *branch_subst = (int)pars_counts_[2 * op];
    return (int)(sum_end_node + pars_counts_[2 * op]);
}

int PhyloTree::computeParsimony() {
    return computeParsimonyBranch(root->neighbors[0], root);
}

void PhyloTree::getBranches(std::vector<Node*> &nodes, std::vector<Node*> &nodes2, Node *node, Node *dad) {
    if (!node) node = root;
    for (size_t i = 0; i < node->neighbors.size(); i++) {
        Node *child = node->neighbors[i]->node;
        if (child == dad) continue;
        if (node->id < child->id) { nodes.push_back(node); nodes2.push_back(child); }
        else { nodes.push_back(child); nodes2.push_back(node); }
        getBranches(nodes, nodes2, child, node);
    }
}

/* phylotree.cpp:2654-2694.  The recursion only decides the visiting order; the substitution counts of ALL branches come from one
   launch, then the Jukes-Cantor corrected lengths are set in the same order. */
int PhyloTree::fixNegativeBranch(bool force, Node *node, Node *dad) {
    if (node) throw std::runtime_error("fixNegativeBranch: call it for the whole tree");
    if (!pars_) initializeAllPartialPars();
    struct Item { Neighbor *nei; Node *node; int op, a, b; };
    std::vector<Item> items;
    /* pre-order walk, phylotree.cpp:2664-2692 */
    struct Walk {
        static void run(PhyloTree *t, Node *node, Node *dad, bool force, std::vector<Item> &items) {
            for (size_t i = 0; i < node->neighbors.size(); i++) {
                Neighbor *it = node->neighbors[i];
                if (it->node == dad) continue;
                if (it->length < 0.0 || force) {
                    Neighbor *back = it->node->findNeighbor(node);
                    if ((it->partial_lh_computed & 2) == 0) t->computePartialParsimony(it, node);
                    if ((back->partial_lh_computed & 2) == 0) t->computePartialParsimony(back, it->node);
                    Item m; m.nei = it; m.node = node; m.a = t->parsVec(it); m.b = t->parsVec(back);
                    m.op = t->parsQueue(LKPARS_OP_BRANCH, -1, m.a, m.b, -1);
                    items.push_back(m);
                }
                run(t, it->node, node, force, items);
            }
        }
    };
    for (size_t i = 0; i < allNodes.size(); i++) if (!allNodes[i]->isLeaf() && allNodes[i]->degree() != 3)
        throw std::runtime_error("fixNegativeBranch: multifurcating trees take setNegativeBranch(0.1) in the reference, not parsimony");
    Walk::run(this, root, NULL, force, items);
    parsFlush();
    int fixed = 0;
    const double nsite_all = (double)pars_aln_nsite_;
    for (size_t k = 0; k < items.size(); k++) {
        const int branch_subst = (int)pars_counts_[2 * items[k].op];
        double branch_length = branch_subst > 0 ? (double)branch_subst / nsite_all : 1.0 / nsite_all;
        const double z = (double)nstates / (nstates - 1);
        const double x = 1.0 - z * branch_length;
        if (x > 0) branch_length = -log(x) / z;
        if (branch_length < min_branch_length) branch_length = min_branch_length;
        items[k].nei->length = branch_length;
        items[k].nei->node->findNeighbor(items[k].node)->length = branch_length;
        fixed++;
    }
    /* phylotree.cpp:2687-2690 */
    for (size_t i = 0; i < allNodes.size(); i++)
        for (size_t k = 0; k < allNodes[i]->neighbors.size(); k++)
            if (allNodes[i]->neighbors[k]->length <= 0.0) allNodes[i]->neighbors[k]->length = min_branch_length;
    return fixed;
}

static void updateNeighbor(Node *self, Node *old_node, Node *new_node, double newlen) {     /* node.cpp:221-228 */
    for (size_t i = 0; i < self->neighbors.size(); i++)
        if (self->neighbors[i]->node == old_node) { self->neighbors[i]->node = new_node; self->neighbors[i]->length = newlen; break; }
}

static Node *findNodeID(int id, Node *node, Node *dad) {                                    /* mtree.cpp:1164-1172 */
    if (node->id == id) return node;
    for (size_t i = 0; i < node->neighbors.size(); i++) {
        if (node->neighbors[i]->node == dad) continue;
        Node *res = findNodeID(id, node->neighbors[i]->node, node);
        if (res) return res;
    }
    return NULL;
}

int PhyloTree::computeParsimonyTree(const int32_t *taxon_order) {
    const int size = ntaxa;
    if (size < 3) throw std::runtime_error("Number of taxa must be greater than 2.");
    if ((int)seq_names.size() != size) throw std::runtime_error("computeParsimonyTree: setAlignment first");
    {
        std::vector<char> seen(size, 0);
        for (int i = 0; i < size; i++) {
            if (taxon_order[i] < 0 || taxon_order[i] >= size || seen[taxon_order[i]]) throw std::runtime_error("computeParsimonyTree: taxon_order is not a permutation");
            seen[taxon_order[i]] = 1;
        }
    }
    deleteAllPartialLh();
    freeNodes();
    pars_ops_.clear();

    root = newNode(); root->id = size;
    for (leafNum = 0; leafNum < 3; leafNum++) {
        Node *new_taxon = newNode();
        new_taxon->id = taxon_order[leafNum]; new_taxon->name = seq_names[taxon_order[leafNum]];
        root->neighbors.push_back(new Neighbor(new_taxon, -1.0));
        new_taxon->neighbors.push_back(new Neighbor(root, -1.0));
    }
    root = findNodeID(taxon_order[0], root, NULL);
    initializeAllPartialPars();
    int index = 6;
    const int tmp_partial_pars = pars_nvec_ - 1;
    Node dummy1, dummy2;                         /* the reference's place holders (Node*) 1 and (Node*) 2 */
    int best_pars_score = INT_MAX;

    for (leafNum = 3; leafNum < size; leafNum++) {
        std::vector<Node*> nodes1, nodes2;
        getBranches(nodes1, nodes2);
        Node *target_node = NULL, *target_dad = NULL;
        best_pars_score = INT_MAX;
        const int new_taxon_partial_pars = index++;
        Node *new_taxon = newNode();
        new_taxon->id = taxon_order[leafNum]; new_taxon->name = seq_names[taxon_order[leafNum]];
        Node *added_node = newNode();
        added_node->id = size + leafNum - 2;
        added_node->neighbors.push_back(new Neighbor(new_taxon, -1.0));
        new_taxon->neighbors.push_back(new Neighbor(added_node, -1.0));
        added_node->findNeighbor(new_taxon)->pars_vec = index++;
        new_taxon->findNeighbor(added_node)->pars_vec = tmp_partial_pars;
        added_node->neighbors.push_back(new Neighbor(&dummy1, -1.0));
        added_node->neighbors.push_back(new Neighbor(&dummy2, -1.0));

        /* addTaxonMPFast on every branch (phylotreepars.cpp:366-374): hanging added_node into (node, dad) makes the partial of
           new_taxon -> added_node the Fitch set of the two directions of that branch, and the score its subtree score plus the
           mismatches against the new tip.  The dirty partials are queued by the same recursion; nothing else of the tree changes. */
        std::vector<int> ops(nodes1.size()), va(nodes1.size()), vb(nodes1.size());
        for (size_t i = 0; i < nodes1.size(); i++) {
            Node *node = nodes1[i], *dad = nodes2[i];
            Neighbor *dad_nei = dad->findNeighbor(node), *node_nei = node->findNeighbor(dad);
            if ((dad_nei->partial_lh_computed & 2) == 0) computePartialParsimony(dad_nei, dad);
            if ((node_nei->partial_lh_computed & 2) == 0) computePartialParsimony(node_nei, node);
            va[i] = parsVec(dad_nei); vb[i] = parsVec(node_nei);
            ops[i] = parsQueue(LKPARS_OP_INSERT, -1, va[i], vb[i], new_taxon->id);
        }
        parsFlush();
        size_t best = 0;
        for (size_t i = 0; i < nodes1.size(); i++) {
            const long long score = (long long)pars_counts_[2 * ops[i]] + pars_counts_[2 * ops[i] + 1] + pars_score_[va[i]] + pars_score_[vb[i]];
            if (score < best_pars_score) {
                best_pars_score = (int)score;
                target_node = nodes1[i]; target_dad = nodes2[i];
                best = i;
            }
        }
        /* memcpy(new_taxon_partial_pars, tmp_partial_pars): the winner's Fitch set is rebuilt at the head of the next launch */
        parsQueue(LKPARS_OP_PARTIAL, ntaxa + new_taxon_partial_pars, va[best], vb[best], -1);
        pars_score_[ntaxa + new_taxon_partial_pars] = pars_counts_[2 * ops[best]] + pars_score_[va[best]] + pars_score_[vb[best]];

        /* insert the new node in the middle of the branch node-dad (phylotreepars.cpp:378-407) */
        updateNeighbor(target_node, target_dad, added_node, -1.0);
        updateNeighbor(target_dad, target_node, added_node, -1.0);
        updateNeighbor(added_node, &dummy1, target_node, -1.0);
        updateNeighbor(added_node, &dummy2, target_dad, -1.0);
        added_node->findNeighbor(target_node)->pars_vec = target_dad->findNeighbor(added_node)->pars_vec;
        added_node->findNeighbor(target_dad)->pars_vec = target_node->findNeighbor(added_node)->pars_vec;
        added_node->findNeighbor(target_node)->partial_lh_computed = target_dad->findNeighbor(added_node)->partial_lh_computed;
        added_node->findNeighbor(target_dad)->partial_lh_computed = target_node->findNeighbor(added_node)->partial_lh_computed;
        new_taxon->findNeighbor(added_node)->partial_lh_computed |= 2;
        new_taxon->findNeighbor(added_node)->pars_vec = new_taxon_partial_pars;
        target_dad->findNeighbor(added_node)->partial_lh_computed = 0;
        target_dad->findNeighbor(added_node)->pars_vec = index++;
        target_node->findNeighbor(added_node)->partial_lh_computed = 0;
        target_node->findNeighbor(added_node)->pars_vec = index++;
        clearReversePartialLh(target_dad, added_node);
        clearReversePartialLh(target_node, added_node);
    }
    if (index != 4 * leafNum - 6) throw std::runtime_error("computeParsimonyTree: block index out of step");
    parsFlush();                                 /* the last winner's vector */

    nodeNum = leafNum;                           /* initializeTree, mtree.cpp:561-583 */
    initializeTreeIds(root, NULL);
    fixNegativeBranch(true);
    if (size == 3) return computeParsimony();    /* no insertion round ran (the reference returns its stale best_pars_score here) */
    return best_pars_score;
}

}  // namespace lk
