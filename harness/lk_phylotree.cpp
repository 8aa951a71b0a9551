/*
 * lk_phylotree.cpp — host-side mirror of the reference's PhyloTree likelihood API over the lkcuda C ABI.
 * See lk_phylotree.h for the reference file:line of every mirrored function.
 *
 * What is different from the reference by design (B200-first):
 *   - computePartialLikelihood does not compute: its post-order recursion (same dirty-bit and
 *     leaf-left rules) only RECORDS nodes; the recorded schedule is levelised and handed to the device
 *     in one lkcuda_partial_batch call (one launch per tree level instead of one per node).
 *   - lh_scale_factor is not carried on the host: the device derives it from scale_num.
 *   - patterns are split into shards (one engine per device / per process); lnL, df, ddf are sums.
 */
#include "lk_phylotree.h"
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <sstream>

namespace lk {

enum { NNI_VECS = 6 };     // IT_NUM of getBestNNIForBran with params->nni5 (phylotree.cpp:857, 2878)

#define FOR_NEI(nd_, dad_, it) \
    for (std::vector<Neighbor*>::iterator it = (nd_)->neighbors.begin(); it != (nd_)->neighbors.end(); ++it) \
        if ((*it)->node != (dad_))

PhyloTree::PhyloTree()
    : leafNum(0), nodeNum(0), root(NULL), min_branch_length(1e-6), max_branch_length(100.0),
      device_newton(false), nni_queued(true), nni_queued_calls(0), nni_queued_fallbacks(0), nni_async_(false), nni_next_slot_(0),
      generic_kernel(false), current_it(NULL), current_it_back(NULL),
      theta_computed(false), num_partial_lh_computations(0), nstates(0), ncat(1), nmix(1), family(FAMILY_PLAIN),
      state_unknown(0), seq_type(SEQ_DNA), ntaxa(0), nptn(0), nunobs(0), nsite(0), p_invar_(0.0),
      proc_rank_(0), proc_world_(1), allreduce_(NULL), allreduce_user_(NULL), stream_(NULL), peers_attached_(false), boot_nsamples_(0),
      ptn_begin_(0), ptn_end_(0), tip_partial_lh_computed_(false), weights_uploaded_(false), model_uploaded_(false),
      freq_dirty_(true), invar_dirty_(true), model_dirty_(true), nvec_(0),
      pars_device(-1), current_scaling_(1.0), keep_engines_(false), pars_(NULL), pars_nsites_(0), pars_aln_nsite_(0), pars_nvec_(0) {}

PhyloTree::~PhyloTree() {
    deleteAllPartialLh();
    deleteAllPartialPars();
    for (size_t i = 0; i < allNodes.size(); i++) {
        for (size_t k = 0; k < allNodes[i]->neighbors.size(); k++) delete allNodes[i]->neighbors[k];
        delete allNodes[i];
    }
}

void PhyloTree::check(int rc, lkcuda_engine *e, const char *what) {
    if (rc == LKCUDA_OK) return;
    std::string msg = std::string(what) + ": " + lkcuda_last_error(e) + " (status " + std::to_string(rc) + ")";
    throw std::runtime_error(msg);   // the reference's outError(): fatal (tools.cpp:99-106)
}

/* ------------------------------------------------------------------ tree ------------------------------------------------------------------ */

Node *PhyloTree::newNode() {
    Node *n = new Node();
    allNodes.push_back(n);
    return n;
}

static bool isNewickToken(char ch) { return ch == ':' || ch == ';' || ch == ',' || ch == ')' || ch == '(' || ch == '[' || ch == ']'; }
static bool isControl(char ch) { return ch <= 32; }

static char nextChar(const std::string &s, size_t &pos) {
    // skip whitespace and [comments] (mtree.cpp readNextChar)
    while (pos < s.size()) {
        char ch = s[pos++];
        if (isControl(ch)) continue;
        if (ch == '[') {
            while (pos < s.size() && s[pos] != ']') pos++;
            if (pos < s.size()) pos++;
            continue;
        }
        return ch;
    }
    return 0;
}

/* recursive-descent Newick reader with the reference's neighbour order: children in file order, then the
   parent (mtree.cpp:586-680); `ch` convention: s[pos-1] is the current character */
void PhyloTree::parseNewick(const std::string &s, size_t &pos, Node *&node, double &branch_len) {
    branch_len = -1.0;
    node = newNode();
    char ch = s[pos - 1];
    if (ch == '(') {
        ch = nextChar(s, pos);
        while (ch != ')' && ch != 0) {
            Node *child = NULL;
            double brlen;
            parseNewick(s, pos, child, brlen);
            node->neighbors.push_back(new Neighbor(child, brlen));
            child->neighbors.push_back(new Neighbor(node, brlen));
            ch = s[pos - 1];
            if (ch == ',') ch = nextChar(s, pos);
            else if (ch != ')') throw std::runtime_error(std::string("Expecting ')', but found '") + ch + "' instead");
        }
        if (ch == 0) throw std::runtime_error("Expecting ')', but end of file instead");
        ch = nextChar(s, pos);
    }
    std::string name;
    char end_ch = (ch == '\'' || ch == '"') ? ch : 0;
    while (ch != 0) {
        if (end_ch == 0 && (isNewickToken(ch) || isControl(ch))) break;
        name += ch;
        ch = pos < s.size() ? s[pos++] : 0;
        if (end_ch != 0 && ch == end_ch) { name += ch; ch = nextChar(s, pos); break; }
    }
    if (isControl(ch) && ch != 0) ch = nextChar(s, pos);
    if (name.empty() && node->isLeaf()) throw std::runtime_error("A taxon has no name.");
    node->name = name;
    if (node->isLeaf()) { node->id = leafNum; leafNum++; }
    if (ch == ';' || ch == 0) return;
    if (ch == ':') {
        ch = nextChar(s, pos);
        std::string num;
        while (ch != 0 && !isNewickToken(ch) && !isControl(ch)) { num += ch; ch = pos < s.size() ? s[pos++] : 0; }
        if (isControl(ch) && ch != 0) ch = nextChar(s, pos);
        if (ch == 0) throw std::runtime_error("branch length format error.");
        branch_len = strtod(num.c_str(), NULL);
    }
}

void PhyloTree::initializeTreeIds(Node *node, Node *dad) {   // mtree.cpp:561-583
    if (!node->isLeaf()) node->id = nodeNum++;
    FOR_NEI(node, dad, it) {
        initializeTreeIds((*it)->node, node);
    }
}

void PhyloTree::readTreeString(const std::string &newick) {
    if (!keep_engines_) deleteAllPartialLh();
    for (size_t i = 0; i < allNodes.size(); i++) {
        for (size_t k = 0; k < allNodes[i]->neighbors.size(); k++) delete allNodes[i]->neighbors[k];
        delete allNodes[i];
    }
    allNodes.clear();
    leafNum = 0; root = NULL;
    size_t pos = 0;
    char ch = nextChar(newick, pos);
    if (ch != '(') throw std::runtime_error("Tree file does not start with an opening-bracket '('");
    Node *top = NULL;
    double brlen;
    parseNewick(newick, pos, top, brlen);
    if (brlen > 0.0) throw std::runtime_error("rooted trees are not supported by the likelihood engine");
    if (pos == 0 || newick[pos - 1] != ';') throw std::runtime_error("Tree file must be ended with a semi-colon ';'");
    FOR_NEI(top, (Node*)NULL, it) if ((*it)->node->isLeaf()) { root = (*it)->node; break; }
    if (!root) for (size_t i = 0; i < allNodes.size(); i++) if (allNodes[i]->isLeaf() && allNodes[i]->id == 0) root = allNodes[i];
    nodeNum = leafNum;
    initializeTreeIds(root, NULL);
    // setAlignment (phylotree.cpp:315-342): leaf ids = alignment sequence ids; setRootNode: first sequence
    if (!seq_names.empty()) {
        if ((int)seq_names.size() != leafNum) throw std::runtime_error("Tree taxa and alignment sequence do not match");
        for (size_t i = 0; i < allNodes.size(); i++) {
            Node *n = allNodes[i];
            if (!n->isLeaf()) continue;
            std::vector<std::string>::iterator f = std::find(seq_names.begin(), seq_names.end(), n->name);
            if (f == seq_names.end()) throw std::runtime_error("Tree taxon " + n->name + " does not appear in the alignment");
            n->id = (int)(f - seq_names.begin());
        }
        setRootNode(seq_names[0]);
    }
    current_it = current_it_back = NULL;
    if (keep_engines_ && !engines_.empty()) {       // same taxa: the new nodes take over the arena vectors, nothing is computed yet
        if (assignVectors(root, NULL, 0) != nvec_) throw std::runtime_error("readTreeString: the tree no longer fits the engines");
        pending_ops_.clear(); pending_level_.clear();
        std::fill(vec_level_.begin(), vec_level_.end(), 0);
        theta_computed = false;
    }
}

void PhyloTree::readAdjacency(const std::string &text) {
    deleteAllPartialLh();
    for (size_t i = 0; i < allNodes.size(); i++) {
        for (size_t k = 0; k < allNodes[i]->neighbors.size(); k++) delete allNodes[i]->neighbors[k];
        delete allNodes[i];
    }
    allNodes.clear();
    std::istringstream in(text);
    std::string tok;
    int root_id = -1, nnodes = 0, nleaves = 0;
    in >> tok >> root_id >> tok >> nnodes >> tok >> nleaves;
    struct Rec { int id, leaf; std::string name; std::vector<int> nb; std::vector<double> len; };
    std::vector<Rec> recs(nnodes);
    int maxid = -1;
    for (int i = 0; i < nnodes; i++) {
        int deg;
        in >> tok >> recs[i].id >> recs[i].leaf >> recs[i].name >> deg;
        if (!in) throw std::runtime_error("bad adjacency text");
        recs[i].nb.resize(deg); recs[i].len.resize(deg);
        for (int k = 0; k < deg; k++) in >> recs[i].nb[k] >> recs[i].len[k];
        maxid = std::max(maxid, recs[i].id);
    }
    std::vector<Node*> byid(maxid + 1, (Node*)NULL);
    for (int i = 0; i < nnodes; i++) {
        Node *n = newNode();
        n->id = recs[i].id;
        n->name = recs[i].name == "-" ? "" : recs[i].name;
        byid[n->id] = n;
    }
    for (int i = 0; i < nnodes; i++)
        for (size_t k = 0; k < recs[i].nb.size(); k++)
            byid[recs[i].id]->neighbors.push_back(new Neighbor(byid[recs[i].nb[k]], recs[i].len[k]));
    leafNum = nleaves; nodeNum = nnodes;
    root = byid[root_id];
    current_it = current_it_back = NULL;
}

void PhyloTree::setRootNode(const std::string &name) {
    for (size_t i = 0; i < allNodes.size(); i++)
        if (allNodes[i]->name == name) { root = allNodes[i]; return; }
    throw std::runtime_error("setRootNode: no node named " + name);
}

void PhyloTree::printSubtree(std::string &out, const Node *node, const Node *dad, int precision) const {
    char buf[64];
    if (node->isLeaf() && dad) { out += node->name; return; }
    out += "(";
    bool first = true;
    for (size_t i = 0; i < node->neighbors.size(); i++) {
        const Neighbor *nb = node->neighbors[i];
        if (nb->node == dad) continue;
        if (!first) out += ",";
        first = false;
        printSubtree(out, nb->node, node, precision);
        snprintf(buf, sizeof(buf), ":%.*g", precision, nb->length);
        out += buf;
    }
    out += ")";
}

std::string PhyloTree::getTreeString(int precision) const {
    // printed from the root leaf's neighbour, root leaf first (same topology, reference-independent text)
    std::string out;
    if (!root) return out;
    const Neighbor *nb = root->neighbors[0];
    char buf[64];
    out += "(" + root->name;
    snprintf(buf, sizeof(buf), ":%.*g", precision, nb->length);
    out += buf;
    for (size_t i = 0; i < nb->node->neighbors.size(); i++) {
        const Neighbor *c = nb->node->neighbors[i];
        if (c->node == root) continue;
        out += ",";
        printSubtree(out, c->node, nb->node, precision);
        snprintf(buf, sizeof(buf), ":%.*g", precision, c->length);
        out += buf;
    }
    out += ");";
    return out;
}

/* ------------------------------------------------------------ alignment / model ------------------------------------------------------------ */

void PhyloTree::setAlignment(int ntaxa_, int64_t nptn_, const uint8_t *states, const double *freq,
                             const std::vector<std::string> &names, int seq_type_, int nstates_, int state_unknown_) {
    // computeTipPartialLikelihood fills the DNA ambiguity rows 4..17 and the protein rows 20..22 of a (state_unknown + 1)-row table
    if (state_unknown_ < nstates_ || (seq_type_ == SEQ_DNA && nstates_ == 4 && state_unknown_ < 18) ||
        (seq_type_ == SEQ_PROTEIN && nstates_ == 20 && state_unknown_ < 23))
        throw std::runtime_error("setAlignment: state_unknown too small for this sequence type (DNA: 18, protein: 23; alignment.cpp:470-472)");
    deleteAllPartialLh();
    ntaxa = ntaxa_; nptn = nptn_; seq_type = seq_type_; nstates = nstates_; state_unknown = state_unknown_;
    states_.assign(states, states + (size_t)ntaxa * (size_t)nptn);
    freq_global_.assign(freq, freq + nptn);
    nsite = 0;
    for (int64_t p = 0; p < nptn; p++) nsite += (int64_t)freq[p];
    seq_names = names;
    const_char_.clear();
    unobs_.clear(); nunobs = 0;
    freq_dirty_ = invar_dirty_ = true;
}

void PhyloTree::setUnobservedPatterns(const std::vector<uint8_t> &unobs) {
    deleteAllPartialLh();
    unobs_ = unobs; nunobs = (int64_t)unobs.size();
    invar_dirty_ = true;
}

void PhyloTree::setConstChar(const int32_t *cc) { const_char_.assign(cc, cc + nptn); invar_dirty_ = true; tip_partial_lh_computed_ = false; }

void PhyloTree::setModel(int family_, int ncat_, int nmix_, const double *eval, const double *evec, const double *inv_evec,
                         const double *rates, const double *props, const double *mixprops,
                         const double *state_freq, double p_invar) {
    if (!engines_.empty() && (family_ != family || ncat_ != ncat || nmix_ != nmix)) deleteAllPartialLh();
    family = family_; ncat = ncat_; nmix = nmix_;
    size_t E = (size_t)nmix, S = (size_t)nstates;
    eval_.assign(eval, eval + E * S);
    evec_.assign(evec, evec + E * S * S);
    inv_evec_.assign(inv_evec, inv_evec + E * S * S);
    rates_.assign(rates, rates + ncat);
    props_.assign(props, props + ncat);
    if (mixprops) mixprops_.assign(mixprops, mixprops + nmix); else mixprops_.assign(nmix, 1.0);
    std::vector<double> sf;
    if (state_freq) sf.assign(state_freq, state_freq + S); else sf.assign(S, 1.0 / S);
    // ptn_invar (phylotreesse.cpp:543-569) depends on p_invar and the state frequencies only
    if (p_invar != p_invar_ || sf != state_freq_) invar_dirty_ = true;
    state_freq_.swap(sf);
    p_invar_ = p_invar;
    // every parameter change is followed by clearAllPartialLH in the reference; the tip table must be redone
    tip_partial_lh_computed_ = false;
    model_dirty_ = true;
}

void PhyloTree::setSiteModels(const int32_t *ptn_model) {
    ptn_model_.assign(ptn_model, ptn_model + nptn);
    if (!engines_.empty()) deleteAllPartialLh();
}

/* phylotreesse.cpp:459-527 */
void PhyloTree::computeTipPartialLikelihood() {
    if (tip_partial_lh_computed_) return;
    tip_partial_lh_computed_ = true;
    // the reference recomputes all three after every clearAllPartialLH (phylotree.cpp:499); their inputs are tracked
    // here so that an unchanged 8 MB ptn_freq / ptn_invar is neither rebuilt nor sent to the device again
    if (freq_dirty_) { computePtnFreq(); freq_dirty_ = false; weights_uploaded_ = false; }
    if (invar_dirty_) { computePtnInvar(); invar_dirty_ = false; weights_uploaded_ = false; }
    if (!model_dirty_) return;
    model_dirty_ = false; model_uploaded_ = false;
    if (family == FAMILY_SITEMODEL) { tip_partial_lh.clear(); return; }   // per-pattern tip vectors are formed on the device (phylotreesse.cpp:373-456)
    int S = nstates, M = nmix;
    tip_partial_lh.assign((size_t)(state_unknown + 1) * S * M, 0.0);
    const double *all_inv_evec = &inv_evec_[0];
    for (int state = 0; state < S; state++)
        for (int m = 0; m < M; m++) {
            const double *inv_evec = all_inv_evec + (size_t)m * S * S;
            for (int i = 0; i < S; i++) tip_partial_lh[(size_t)state * S * M + m * S + i] = inv_evec[i * S + state];
        }
    for (int i = 0; i < S; i++)
        for (int m = 0; m < M; m++) {
            const double *inv_evec = all_inv_evec + (size_t)m * S * S;
            double lh_unknown = 0.0;
            for (int x = 0; x < S; x++) lh_unknown += inv_evec[i * S + x];
            tip_partial_lh[(size_t)state_unknown * S * M + m * S + i] = lh_unknown;
        }
    static const int ambi_aa[] = {4 + 8, 32 + 64, 512 + 1024};
    if (seq_type == SEQ_DNA && S == 4) {
        for (int state = 4; state < 18; state++) {
            int cstate = state - S + 1;
            for (int m = 0; m < M; m++) {
                const double *inv_evec = all_inv_evec + (size_t)m * S * S;
                for (int i = 0; i < S; i++) {
                    double lh_ambiguous = 0.0;
                    for (int x = 0; x < S; x++) if (cstate & (1 << x)) lh_ambiguous += inv_evec[i * S + x];
                    tip_partial_lh[(size_t)state * S * M + m * S + i] = lh_ambiguous;
                }
            }
        }
    } else if (seq_type == SEQ_PROTEIN && S == 20) {
        for (int state = 0; state < 3; state++)
            for (int m = 0; m < M; m++) {
                const double *inv_evec = all_inv_evec + (size_t)m * S * S;
                for (int i = 0; i < S; i++) {
                    double lh_ambiguous = 0.0;
                    for (int x = 0; x < 11; x++) if (ambi_aa[state] & (1 << x)) lh_ambiguous += inv_evec[i * S + x];
                    tip_partial_lh[(size_t)(state + 20) * S * M + m * S + i] = lh_ambiguous;
                }
            }
    }
}

void PhyloTree::computePtnFreq() {   // phylotreesse.cpp:531-541
    ptn_freq.assign(freq_global_.begin(), freq_global_.end());
}

void PhyloTree::computePtnInvar() {  // phylotreesse.cpp:543-569
    ptn_invar.assign((size_t)(nptn + nunobs), 0.0);
    if (p_invar_ != 0.0) {
        if ((int64_t)const_char_.size() != nptn) throw std::runtime_error("computePtnInvar: +I needs Pattern::const_char (setConstChar)");
        for (int64_t p = 0; p < nptn; p++) {
            if (const_char_[p] == nstates) ptn_invar[p] = p_invar_;
            else if (const_char_[p] >= 0 && const_char_[p] < nstates) ptn_invar[p] = p_invar_ * state_freq_[const_char_[p]];
        }
        for (int64_t u = 0; u < nunobs; u++) ptn_invar[nptn + u] = p_invar_ * state_freq_[unobs_[u]];
    }
}

/* ------------------------------------------------------------ buffers / engines ------------------------------------------------------------ */

int PhyloTree::assignVectors(Node *node, Node *dad, int index) {
    // every neighbour pointing to an internal node owns a vector (LM_ALL_BRANCH, phylotree.cpp:916-974)
    FOR_NEI(node, dad, it) {
        Neighbor *nei = *it, *back = nei->node->findNeighbor(node);
        nei->vec = nei->node->isLeaf() ? -1 : index++;
        back->vec = node->isLeaf() ? -1 : index++;
        nei->partial_lh_computed = back->partial_lh_computed = 0;
        index = assignVectors(nei->node, node, index);
    }
    return index;
}

void PhyloTree::setStream(void *s) {
    stream_ = s;
    for (size_t i = 0; i < engines_.size(); i++) check(lkcuda_set_stream(engines_[i], s), engines_[i], "lkcuda_set_stream");
}

void PhyloTree::initializeAllPartialLh() {
    if (!engines_.empty()) return;
    if (!root) throw std::runtime_error("initializeAllPartialLh: no tree");
    if (states_.empty() || eval_.empty()) throw std::runtime_error("initializeAllPartialLh: alignment and model must be set first");
    nvec_ = assignVectors(root, NULL, 0);
    // this process' pattern range, then one sub-shard per device
    int64_t pb = nptn * proc_rank_ / proc_world_, pe = nptn * (proc_rank_ + 1) / proc_world_;
    ptn_begin_ = pb; ptn_end_ = pe;
    std::vector<int> devs = devices_;
    if (devs.empty()) devs.push_back(0);
    size_t nd = devs.size();
    shard_begin_.resize(nd + 1);
    for (size_t d = 0; d <= nd; d++) shard_begin_[d] = pb + (pe - pb) * (int64_t)d / (int64_t)nd;
    int K = family == FAMILY_MIXTURE ? ncat * nmix : ncat;
    for (size_t d = 0; d < nd; d++) {
        lkcuda_config cfg;
        memset(&cfg, 0, sizeof(cfg));
        cfg.device = devs[d];
        cfg.nstates = nstates; cfg.nblocks = K; cfg.neigen = nmix; cfg.ntaxa = ntaxa;
        cfg.state_unknown = state_unknown; cfg.nvec = nvec_ + NNI_VECS;
        cfg.flags = (family == FAMILY_SITEMODEL ? LKCUDA_FLAG_SITEMODEL : 0) |
                    (family != FAMILY_PLAIN ? LKCUDA_FLAG_SCALE_IGNORES_INVAR : 0) |
                    (generic_kernel ? (LKCUDA_FLAG_ZERO_GUARD | LKCUDA_FLAG_PATTERN_LH_CAT) : 0);
        cfg.nptn = shard_begin_[d + 1] - shard_begin_[d];
        cfg.nunobs = nunobs;
        int64_t ns = 0;
        for (int64_t p = shard_begin_[d]; p < shard_begin_[d + 1]; p++) ns += (int64_t)freq_global_[p];
        cfg.nsite = ns;
        lkcuda_engine *e = NULL;
        int rc = lkcuda_create(&cfg, &e);
        if (rc != LKCUDA_OK) {
            std::string msg = std::string("lkcuda_create: ") + lkcuda_last_error(NULL);
            deleteAllPartialLh();
            throw std::runtime_error(msg);
        }
        engines_.push_back(e);
        if (stream_) check(lkcuda_set_stream(e, stream_), e, "lkcuda_set_stream");
        // tip states of the shard: [taxon][local ptn + unobserved]
        size_t np = (size_t)cfg.nptn, na = np + (size_t)nunobs;
        std::vector<uint8_t> st((size_t)ntaxa * na);
        for (int t = 0; t < ntaxa; t++) {
            memcpy(&st[(size_t)t * na], &states_[(size_t)t * nptn + shard_begin_[d]], np);
            for (int64_t u = 0; u < nunobs; u++) st[(size_t)t * na + np + u] = unobs_[u];
        }
        check(lkcuda_set_tip_states(e, na ? &st[0] : (const uint8_t*)""), e, "lkcuda_set_tip_states");
        if (family == FAMILY_SITEMODEL) {
            if ((int64_t)ptn_model_.size() != nptn) throw std::runtime_error("site-specific model: setSiteModels first");
            if (nunobs) throw std::runtime_error("site-specific models do not combine with +ASC");
            // the states each tip code stands for (phylotreesse.cpp:386-440): plain, unknown, DNA / protein ambiguity codes
            std::vector<uint64_t> mask((size_t)state_unknown + 1, 0);
            const uint64_t all = nstates >= 64 ? ~0ull : ((1ull << nstates) - 1);
            for (int s = 0; s <= state_unknown; s++) {
                if (s < nstates) mask[s] = 1ull << s;
                else if (s == state_unknown) mask[s] = all;
                else if (seq_type == SEQ_DNA && nstates == 4) mask[s] = (uint64_t)(s - nstates + 1);
                else if (seq_type == SEQ_PROTEIN && nstates == 20 && s < 23) { static const uint64_t ambi_aa[] = {4 + 8, 32 + 64, 512 + 1024}; mask[s] = ambi_aa[s - 20]; }
            }
            static const int32_t zero_model = 0;
            check(lkcuda_set_site_models(e, np ? &ptn_model_[shard_begin_[d]] : &zero_model, &mask[0]), e, "lkcuda_set_site_models");
        }
    }
    // several engines in ONE process, each on its own stream: their reduction kernels exchange and add the result blocks among
    // themselves (raw device addresses; peer access across devices), the host no longer sums anything.  With a caller-supplied
    // common stream the kernels would serialise behind each other and could not wait for one another: host sum then.
    peers_attached_ = false;
    if (engines_.size() > 1 && engines_.size() <= (size_t)LKCUDA_MAX_PEERS && proc_world_ == 1 && !stream_ && !getenv("LKTREE_NO_PEERS")) {
        std::vector<void*> ptrs(engines_.size(), (void*)NULL);
        bool ok = true;
        for (size_t d = 0; d < engines_.size() && ok; d++) ok = lkcuda_peer_handle(engines_[d], NULL, &ptrs[d]) == LKCUDA_OK;
        for (size_t d = 0; d < engines_.size() && ok; d++)
            ok = lkcuda_attach_peers(engines_[d], (int32_t)d, (int32_t)engines_.size(), NULL, &ptrs[0]) == LKCUDA_OK;
        if (!ok) for (size_t d = 0; d < engines_.size(); d++) lkcuda_detach_peers(engines_[d]);
        peers_attached_ = ok;
    }
    weights_uploaded_ = model_uploaded_ = false;
    freq_dirty_ = invar_dirty_ = model_dirty_ = true;
    tip_partial_lh_computed_ = false;
    vec_level_.assign(nvec_ + NNI_VECS, 0);
    clearAllPartialLH();
}

void PhyloTree::peerHandle(uint8_t *handle) {
    if (engines_.empty()) initializeAllPartialLh();
    if (engines_.size() != 1) throw std::runtime_error("peerHandle: one engine per process expected");
    check(lkcuda_peer_handle(engines_[0], handle, NULL), engines_[0], "lkcuda_peer_handle");
}

void PhyloTree::detachPeers() {
    for (size_t d = 0; d < engines_.size(); d++) check(lkcuda_detach_peers(engines_[d]), engines_[d], "lkcuda_detach_peers");
    peers_attached_ = false;
}

void PhyloTree::attachPeers(const uint8_t *handles) {
    if (engines_.size() != 1) throw std::runtime_error("attachPeers: one engine per process expected");
    check(lkcuda_attach_peers(engines_[0], proc_rank_, proc_world_, handles, NULL), engines_[0], "lkcuda_attach_peers");
    peers_attached_ = proc_world_ > 1;
}

void PhyloTree::deleteAllPartialLh() {
    for (size_t i = 0; i < engines_.size(); i++) lkcuda_destroy(engines_[i]);
    engines_.clear();
    peers_attached_ = false;
    pending_ops_.clear(); pending_level_.clear();
    theta_computed = false;
}

void PhyloTree::ensureInputs() {
    if (!tip_partial_lh_computed_) computeTipPartialLikelihood();
    if (weights_uploaded_ && model_uploaded_) return;
    int K = family == FAMILY_MIXTURE ? ncat * nmix : ncat;
    std::vector<int32_t> blk_eigen(K);
    std::vector<double> blk_rate(K), blk_weight(K);
    for (int k = 0; k < K; k++) {
        if (family == FAMILY_MIXTURE) {        // block = m*ncat + c (phylokernelmixture.h:121)
            int m = k / ncat, c = k % ncat;
            blk_eigen[k] = m; blk_rate[k] = rates_[c]; blk_weight[k] = props_[c] * mixprops_[m];
        } else if (family == FAMILY_MIXRATE) { // category c uses component c (phylokernelmixrate.h:52,99-130)
            blk_eigen[k] = k; blk_rate[k] = rates_[k]; blk_weight[k] = props_[k];
        } else {
            blk_eigen[k] = 0; blk_rate[k] = rates_[k]; blk_weight[k] = props_[k];
        }
    }
    for (size_t d = 0; d < engines_.size(); d++) {
        lkcuda_engine *e = engines_[d];
        size_t np = (size_t)(shard_begin_[d + 1] - shard_begin_[d]);
        std::vector<double> inv(np + (size_t)nunobs + 1);
        if (np) memcpy(&inv[0], &ptn_invar[shard_begin_[d]], np * 8);
        for (int64_t u = 0; u < nunobs; u++) inv[np + u] = ptn_invar[nptn + u];
        static const double zero = 0.0;
        if (!weights_uploaded_)
            check(lkcuda_set_pattern_weights(e, np ? &ptn_freq[shard_begin_[d]] : &zero, &inv[0]), e, "lkcuda_set_pattern_weights");
        if (!model_uploaded_)
            check(lkcuda_set_model(e, &eval_[0], &evec_[0], &inv_evec_[0], &blk_eigen[0], &blk_rate[0], &blk_weight[0],
                                   tip_partial_lh.empty() ? NULL : &tip_partial_lh[0]), e, "lkcuda_set_model");
    }
    weights_uploaded_ = model_uploaded_ = true;
}

/* ------------------------------------------------------------ dirty bits ------------------------------------------------------------ */

static void clearAllRec(Node *node, Node *dad) {     // phylonode.cpp:53-65
    node->findNeighbor(dad)->partial_lh_computed = 0;
    dad->findNeighbor(node)->partial_lh_computed = 0;
    FOR_NEI(node, dad, it) clearAllRec((*it)->node, node);
}

void PhyloTree::clearAllPartialLH() {                // phylotree.cpp:495-502
    if (!root) return;
    clearAllRec(root->neighbors[0]->node, root);
    tip_partial_lh_computed_ = false;
    current_it = current_it_back = NULL;
    pending_ops_.clear(); pending_level_.clear();
    std::fill(vec_level_.begin(), vec_level_.end(), 0);
}

void PhyloTree::clearReversePartialLh(Node *node, Node *dad) {   // phylonode.cpp:42-51
    FOR_NEI(node, dad, it) {
        (*it)->node->findNeighbor(node)->partial_lh_computed = 0;
        clearReversePartialLh((*it)->node, node);
    }
}

/* ------------------------------------------------------------ kernels ------------------------------------------------------------ */

int PhyloTree::opLevel(Neighbor *child) {
    return (child->vec >= 0) ? vec_level_[child->vec] : 0;
}

/* phylokernel.h:71-143: same early-outs, same recursion order (left then right), same leaf-left swap.
   Records the node instead of computing it. */
void PhyloTree::computePartialLikelihood(Neighbor *dad_branch, Node *dad) {
    if (dad_branch->partial_lh_computed & 1) return;
    dad_branch->partial_lh_computed |= 1;
    num_partial_lh_computations++;
    Node *node = dad_branch->node;
    if (node->isLeaf()) return;
    if (node->degree() > 3) {
        // multifurcating node: generic path, issued immediately after its children (phylotreesse.cpp:702-806)
        std::vector<int32_t> child; std::vector<double> len;
        FOR_NEI(node, dad, it) {
            if (((*it)->partial_lh_computed & 1) == 0) computePartialLikelihood(*it, node);
            child.push_back((*it)->node->isLeaf() ? LKCUDA_TIP((*it)->node->id) : (*it)->vec);
            len.push_back((*it)->length);
        }
        flushPartialOps();
        for (size_t d = 0; d < engines_.size(); d++)
            check(lkcuda_partial_multi(engines_[d], dad_branch->vec, (int32_t)child.size(), &child[0], &len[0]), engines_[d], "lkcuda_partial_multi");
        return;
    }
    Neighbor *left = NULL, *right = NULL;
    FOR_NEI(node, dad, it) { if (!left) left = *it; else right = *it; }
    if (!left->node->isLeaf() && right->node->isLeaf()) std::swap(left, right);
    if ((left->partial_lh_computed & 1) == 0) computePartialLikelihood(left, node);
    if ((right->partial_lh_computed & 1) == 0) computePartialLikelihood(right, node);
    lkcuda_partial_op op;
    op.dst = dad_branch->vec;
    op.left = left->node->isLeaf() ? LKCUDA_TIP(left->node->id) : left->vec;
    op.right = right->node->isLeaf() ? LKCUDA_TIP(right->node->id) : right->vec;
    op.reserved = LKCUDA_LEN_REFS(left->len_slot, right->len_slot);      // 0 unless a queued NNI left a length on the device
    op.len_left = left->length;
    op.len_right = right->length;
    int level = 1 + std::max(opLevel(left), opLevel(right));
    vec_level_[op.dst] = level;
    pending_ops_.push_back(op);
    pending_level_.push_back(level);
}

/* hand the recorded post-order schedule to the device, grouped by tree level; with `branch` the branch lnL of (dad_end, node_vec) is
   asked for in the same call (lkcuda_partial_batch_lnl: PhyloTree::computeLikelihood's traversal + root branch) */
void PhyloTree::flushPartialOps(const BranchRequest *branch) {
    ensureInputs();
    if (pending_ops_.empty()) {
        if (branch)
            for (size_t d = 0; d < engines_.size(); d++)
                check(lkcuda_branch_lnl(engines_[d], branch->dad_end, branch->node_vec, branch->len, NULL), engines_[d], "lkcuda_branch_lnl");
        return;
    }
    size_t n = pending_ops_.size();
    int maxlevel = 0;
    for (size_t i = 0; i < n; i++) maxlevel = std::max(maxlevel, pending_level_[i]);
    std::vector<int32_t> level_start(maxlevel + 1, 0);
    for (size_t i = 0; i < n; i++) level_start[pending_level_[i]]++;       // counts at [level]
    int32_t acc = 0;
    for (int l = 1; l <= maxlevel; l++) { int32_t c = level_start[l]; level_start[l] = acc; acc += c; }
    // level_start[l] = first slot of level l (1-based levels) -> shift to 0-based array of size maxlevel+1
    std::vector<lkcuda_partial_op> sorted(n);
    std::vector<int32_t> cursor(level_start.begin(), level_start.end());
    for (size_t i = 0; i < n; i++) sorted[cursor[pending_level_[i]]++] = pending_ops_[i];   // stable
    std::vector<int32_t> starts(maxlevel + 1);
    for (int l = 1; l <= maxlevel; l++) starts[l - 1] = level_start[l];
    starts[maxlevel] = (int32_t)n;
    if (branch && engines_.size() == 1) {
        check(lkcuda_partial_batch_lnl(engines_[0], &sorted[0], (int32_t)n, &starts[0], maxlevel, branch->dad_end, branch->node_vec, branch->len, NULL),
              engines_[0], "lkcuda_partial_batch_lnl");
    } else {
        // in-process shards exchange their sums inside the reduction kernels: every shard's traversal is queued before the first
        // reduction starts to wait for its peers
        for (size_t d = 0; d < engines_.size(); d++)
            check(lkcuda_partial_batch(engines_[d], &sorted[0], (int32_t)n, &starts[0], maxlevel), engines_[d], "lkcuda_partial_batch");
        if (branch)
            for (size_t d = 0; d < engines_.size(); d++)
                check(lkcuda_branch_lnl(engines_[d], branch->dad_end, branch->node_vec, branch->len, NULL), engines_[d], "lkcuda_branch_lnl");
    }
    for (size_t i = 0; i < n; i++) vec_level_[pending_ops_[i].dst] = 0;
    pending_ops_.clear(); pending_level_.clear();
}

/* sum the shards' results; under one-process-per-GPU the caller-supplied all-reduce sums across ranks */
void PhyloTree::reduceResult(double *result) {
    double tot[LKCUDA_NRESULT];
    memset(tot, 0, sizeof(tot));
    if (peers_attached_) {
        // the kernels already exchanged and added their blocks: every engine holds the same global sums
        for (size_t d = 0; d < engines_.size(); d++)
            check(lkcuda_wait_result(engines_[d], tot), engines_[d], "lkcuda_wait_result");
    } else if (allreduce_ && engines_.size() == 1) {
        void *dev = NULL;
        check(lkcuda_result_device_ptr(engines_[0], &dev), engines_[0], "lkcuda_result_device_ptr");
        allreduce_(allreduce_user_, NULL, dev, LKCUDA_NRESULT);
        check(lkcuda_read_result(engines_[0], tot), engines_[0], "lkcuda_read_result");
    } else {
        for (size_t d = 0; d < engines_.size(); d++) {
            double r[LKCUDA_NRESULT];
            check(lkcuda_wait_result(engines_[d], r), engines_[d], "lkcuda_wait_result");
            for (int i = 0; i < LKCUDA_NRESULT; i++) tot[i] += r[i];
        }
        if (allreduce_) allreduce_(allreduce_user_, tot, NULL, LKCUDA_NRESULT);
    }
    memcpy(result, tot, sizeof(tot));
}

double PhyloTree::computeLikelihoodBranch(Neighbor *dad_branch, Node *dad) {   // phylokernel.h:734-751
    Node *node = dad_branch->node;
    Neighbor *node_branch = node->findNeighbor(dad);
    if (engines_.empty()) initializeAllPartialLh();
    if (node->isLeaf()) { std::swap(dad, node); std::swap(dad_branch, node_branch); }
    if ((dad_branch->partial_lh_computed & 1) == 0) computePartialLikelihood(dad_branch, dad);
    if ((node_branch->partial_lh_computed & 1) == 0) computePartialLikelihood(node_branch, node);
    // dad_branch points from dad to node: its vector is the subtree at `node`; node_branch's is the subtree at `dad`
    BranchRequest br;
    br.dad_end = dad->isLeaf() ? LKCUDA_TIP(dad->id) : node_branch->vec;
    br.node_vec = dad_branch->vec; br.len = dad_branch->length;
    flushPartialOps(&br);
    double r[LKCUDA_NRESULT];
    reduceResult(r);
    if (r[LKCUDA_R_NREPAIRED] > 0)
        fprintf(stderr, "WARNING: Numerical underflow caused by %.0f alignment patterns\n", r[LKCUDA_R_NREPAIRED]);
    return r[LKCUDA_R_LNL];
}

void PhyloTree::computeLikelihoodDerv(Neighbor *dad_branch, Node *dad, double &df, double &ddf) {   // phylokernel.h:486-503
    Node *node = dad_branch->node;
    Neighbor *node_branch = node->findNeighbor(dad);
    if (engines_.empty()) initializeAllPartialLh();
    if (node->isLeaf()) { std::swap(dad, node); std::swap(dad_branch, node_branch); }
    if ((dad_branch->partial_lh_computed & 1) == 0) computePartialLikelihood(dad_branch, dad);
    if ((node_branch->partial_lh_computed & 1) == 0) computePartialLikelihood(node_branch, node);
    flushPartialOps();
    if (!theta_computed) {
        theta_computed = true;
        int32_t dad_end = dad->isLeaf() ? LKCUDA_TIP(dad->id) : node_branch->vec;
        for (size_t d = 0; d < engines_.size(); d++)      // theta_all and the first df/ddf in one pass
            check(lkcuda_theta_derv(engines_[d], dad_end, dad_branch->vec, dad_branch->length, NULL), engines_[d], "lkcuda_theta_derv");
    } else {
        for (size_t d = 0; d < engines_.size(); d++)
            check(lkcuda_derv(engines_[d], dad_branch->length, NULL), engines_[d], "lkcuda_derv");
    }
    double r[LKCUDA_NRESULT];
    reduceResult(r);
    df = r[LKCUDA_R_DF]; ddf = r[LKCUDA_R_DDF];
    if (std::isnan(df) || std::isinf(df)) { df = 0.0; ddf = 0.0; }
}

double PhyloTree::computeLikelihoodFromBuffer() {   // phylotreesse.cpp:349-357, phylokernel.h:1023-1192
    if (!current_it || !current_it_back) throw std::runtime_error("computeLikelihoodFromBuffer: no current branch");
    if (!theta_computed) return computeLikelihoodBranch(current_it, current_it_back->node);
    for (size_t d = 0; d < engines_.size(); d++)
        check(lkcuda_lnl_from_buffer(engines_[d], current_it->length, NULL), engines_[d], "lkcuda_lnl_from_buffer");
    double r[LKCUDA_NRESULT];
    reduceResult(r);
    return r[LKCUDA_R_LNL];
}

Node *PhyloTree::findFarthestLeaf(Node *node, Node *dad) {   // mtree.cpp:2052-2068
    if (!node) node = root;
    if (dad && node->isLeaf()) { node->height = 0.0; return node; }
    Node *res = NULL;
    node->height = 0.0;
    FOR_NEI(node, dad, it) {
        Node *leaf = findFarthestLeaf((*it)->node, node);
        if (node->height < (*it)->node->height + 1) {
            node->height = (*it)->node->height + 1;
            res = leaf;
        }
    }
    return res;
}

void PhyloTree::getPreOrderBranches(std::vector<Node*> &nodes, std::vector<Node*> &nodes2, Node *node, Node *dad) {   // mtree.cpp:2096-2116
    if (dad) { nodes.push_back(node); nodes2.push_back(dad); }
    std::vector<Neighbor*> neivec = node->neighbors;
    for (size_t i1 = 0; i1 < neivec.size(); i1++)
        for (size_t i2 = i1 + 1; i2 < neivec.size(); i2++)
            if (neivec[i1]->node->height > neivec[i2]->node->height) std::swap(neivec[i1], neivec[i2]);
    for (size_t i1 = 0; i1 < neivec.size(); i1++)
        if (neivec[i1]->node != dad) getPreOrderBranches(nodes, nodes2, neivec[i1]->node, node);
}

void PhyloTree::computeBestTraversal(std::vector<Node*> &nodes, std::vector<Node*> &nodes2) {   // phylotree.cpp:2241-2250
    Node *farleaf = findFarthestLeaf();
    findFarthestLeaf(farleaf);
    getPreOrderBranches(nodes, nodes2, farleaf);
}

double PhyloTree::computeLikelihood(double *pattern_lh) {   // phylotree.cpp:1031-1071
    if (!root || !root->isLeaf()) throw std::runtime_error("computeLikelihood: root must be a leaf");
    if (!current_it) {
        Node *leaf = findFarthestLeaf();
        current_it = leaf->neighbors[0];
        current_it_back = current_it->node->findNeighbor(leaf);
    }
    double score = computeLikelihoodBranch(current_it, current_it_back->node);
    if (pattern_lh) {
        // _pattern_lh + max(scale_num,0)*LOG_SCALING_THRESHOLD of current_it (phylotree.cpp:1055-1069)
        for (size_t d = 0; d < engines_.size(); d++) {
            int64_t b = shard_begin_[d] - ptn_begin_, np = shard_begin_[d + 1] - shard_begin_[d];
            if (!np) continue;
            check(lkcuda_fetch_pattern_lh_corrected(engines_[d], current_it->vec, pattern_lh + b), engines_[d], "lkcuda_fetch_pattern_lh_corrected");
        }
    }
    return score;
}

double PhyloTree::computeLikelihoodAsyncFetch(double *pattern_lh) {
    double score = computeLikelihood(NULL);
    for (size_t d = 0; d < engines_.size(); d++) {
        int64_t b = shard_begin_[d] - ptn_begin_, np = shard_begin_[d + 1] - shard_begin_[d];
        if (!np) continue;
        if (current_it->vec < 0) throw std::runtime_error("computeLikelihoodAsyncFetch: the root branch has no vector");
        check(lkcuda_fetch_pattern_lh_corrected_async(engines_[d], current_it->vec, pattern_lh + b), engines_[d], "lkcuda_fetch_pattern_lh_corrected_async");
    }
    return score;
}

void PhyloTree::waitPatternLh() {
    for (size_t d = 0; d < engines_.size(); d++) check(lkcuda_fetch_wait(engines_[d]), engines_[d], "lkcuda_fetch_wait");
}

/* phylotree.cpp:1108-1160: per-category pattern likelihoods on current_it (picked like the reference when unset);
   wsl: 2 = WSL_RATECAT, 3 = WSL_MIXTURE, 4 = WSL_MIXTURE_RATECAT (tools.h:406). Returns the entries per pattern. */
void PhyloTree::enqueuePatternLhCat() {
    if (engines_.empty()) initializeAllPartialLh();
    if (!current_it) {
        Node *leaf = findFarthestLeaf();   // findFirstFarLeaf (phylotree.cpp:1018) picks the same leaf: first maximal height
        current_it = leaf->neighbors[0];
        current_it_back = current_it->node->findNeighbor(leaf);
    }
    Neighbor *dad_branch = current_it; Node *dad = current_it_back->node;
    Node *node = dad_branch->node;
    Neighbor *node_branch = current_it_back;
    if (node->isLeaf()) { std::swap(dad, node); std::swap(dad_branch, node_branch); }
    if ((dad_branch->partial_lh_computed & 1) == 0) computePartialLikelihood(dad_branch, dad);
    if ((node_branch->partial_lh_computed & 1) == 0) computePartialLikelihood(node_branch, node);
    flushPartialOps();
    int32_t dad_end = dad->isLeaf() ? LKCUDA_TIP(dad->id) : node_branch->vec;
    for (size_t d = 0; d < engines_.size(); d++)
        check(lkcuda_pattern_lh_cat(engines_[d], dad_end, dad_branch->vec, dad_branch->length), engines_[d], "lkcuda_pattern_lh_cat");
}

int PhyloTree::computePatternLhCat(double *out, int wsl) {
    enqueuePatternLhCat();
    int K = family == FAMILY_MIXTURE ? ncat * nmix : ncat;
    std::vector<double> raw((size_t)(ptn_end_ - ptn_begin_) * K + 1);
    for (size_t d = 0; d < engines_.size(); d++) {
        int64_t b = shard_begin_[d] - ptn_begin_;
        check(lkcuda_fetch_pattern_lh_cat(engines_[d], &raw[b * K]), engines_[d], "lkcuda_fetch_pattern_lh_cat");
    }
    int64_t np = ptn_end_ - ptn_begin_;
    if (family == FAMILY_MIXTURE && wsl == 3 && ncat > 1) {          // per mixture class
        for (int64_t p = 0; p < np; p++)
            for (int m = 0; m < nmix; m++) {
                const double *lc = &raw[(size_t)p * K + (size_t)m * ncat];
                double lh = lc[0];
                for (int c = 1; c < ncat; c++) lh += lc[c];
                out[p * nmix + m] = lh;
            }
        return nmix;
    }
    if (family == FAMILY_MIXTURE && wsl == 2 && nmix > 1) {          // per rate category
        for (int64_t p = 0; p < np; p++)
            for (int c = 0; c < ncat; c++) {
                double lh = raw[(size_t)p * K + c];
                for (int m = 1; m < nmix; m++) lh += raw[(size_t)p * K + (size_t)m * ncat + c];
                out[p * ncat + c] = lh;
            }
        return ncat;
    }
    memcpy(out, &raw[0], (size_t)np * K * 8);
    return K;
}

/* ---- the consumers of _pattern_lh_cat, run on the device (only ncol sums / P rates come back) ---- */

/* the collapse of computePatternLhCat (phylotree.cpp:1133-1158) on every engine; returns the columns per pattern */
int PhyloTree::collapsePatternLhCat(int wsl) {
    int mode = 0, inner = ncat, outer = 1;
    if (family == FAMILY_MIXTURE) {
        outer = nmix;
        if (wsl == 3 && ncat > 1) mode = 1; else if (wsl == 2 && nmix > 1) mode = 2;
        if (wsl == 3 && ncat == 1) mode = 0;
    }
    int32_t ncol = 0;
    for (size_t d = 0; d < engines_.size(); d++)
        check(lkcuda_lh_cat_collapse(engines_[d], mode, inner, outer, &ncol), engines_[d], "lkcuda_lh_cat_collapse");
    return ncol;
}

/* ModelMixture::optimizeWeights (model/modelmixture.cpp:1356-1416), statement for statement; the two passes over the patterns
   (:1369-1388) are one kernel per EM step */
double PhyloTree::optimizeMixtureWeights() {
    if (family != FAMILY_MIXTURE) throw std::runtime_error("optimizeMixtureWeights: not a mixture model");
    enqueuePatternLhCat();                               // phylo_tree->computePatternLhCat(WSL_MIXTURE)
    int ncol = collapsePatternLhCat(3);
    if (ncol != nmix) throw std::runtime_error("optimizeMixtureWeights: unexpected column count");
    std::vector<double> new_prop(nmix), ratio_prop(nmix), part(nmix);
    std::vector<double> &prop = mixprops_;
    const bool has_invar = p_invar_ != 0.0;             // RateHeterogeneity::setPInvar is a no-op without an invariant class
    for (int step = 0; step < nmix; step++) {
        std::fill(new_prop.begin(), new_prop.end(), 0.0);
        for (size_t d = 0; d < engines_.size(); d++) {
            check(lkcuda_lh_cat_em(engines_[d], step > 0 ? &ratio_prop[0] : NULL, &part[0]), engines_[d], "lkcuda_lh_cat_em");
            for (int c = 0; c < nmix; c++) new_prop[c] += part[c];
        }
        if (allreduce_) allreduce_(allreduce_user_, &new_prop[0], NULL, nmix);
        bool converged = true;
        double new_pinvar = 0.0;
        for (int c = 0; c < nmix; c++) {
            new_prop[c] /= (double)nsite;
            converged = converged && (fabs(prop[c] - new_prop[c]) < 1e-4);
            ratio_prop[c] = new_prop[c] / prop[c];
            prop[c] = new_prop[c];
            new_pinvar += prop[c];
        }
        new_pinvar = 1.0 - new_pinvar;
        if (new_pinvar != 0.0) {
            converged = converged && (fabs((has_invar ? p_invar_ : 0.0) - new_pinvar) < 1e-4);
            if (has_invar) {
                p_invar_ = new_pinvar;
                computePtnInvar();
                weights_uploaded_ = false;
                ensureInputs();
            }
        }
        if (converged) break;
    }
    // the weights only enter the branch reductions (val = exp(eval*rate*t)*prop): partials stay valid, the model goes up again
    tip_partial_lh_computed_ = false; model_dirty_ = true;
    return computeLikelihood();
}

/* RateGamma::computePatternRates (model/rategamma.cpp:235-257); with an invariant class (+I+G, p_invar != 0) RateGammaInvar's
   variant (model/rategammainvar.cpp:193-221), whose categories are 0 = invariant, c + 1 = Gamma category c */
int PhyloTree::computePatternRates(double *pattern_rates, int32_t *pattern_cat) {
    enqueuePatternLhCat();                               // phylo_tree->computePatternLhCat(WSL_RATECAT)
    int ncol = collapsePatternLhCat(2);
    if (ncol != ncat) throw std::runtime_error("computePatternRates: unexpected column count");
    for (size_t d = 0; d < engines_.size(); d++) {
        int64_t b = shard_begin_[d] - ptn_begin_;
        check(lkcuda_lh_cat_pattern_rates(engines_[d], &rates_[0], p_invar_ != 0.0 ? 1 : 0, pattern_rates + b, pattern_cat + b), engines_[d], "lkcuda_lh_cat_pattern_rates");
    }
    return p_invar_ != 0.0 ? ncat + 1 : ncat;
}

/* ---- UFBoot RELL (IQTree::saveCurrentTree, iqtree.cpp:2676-2786) ---- */
void PhyloTree::setBootSamples(const float *samples, int nsamples) {     // boot_samples (iqtree.cpp:300-340), [nsamples][nptn]
    if (engines_.empty()) initializeAllPartialLh();
    boot_nsamples_ = nsamples;
    for (size_t d = 0; d < engines_.size(); d++)
        check(lkcuda_set_boot_samples(engines_[d], samples + shard_begin_[d], nsamples, nptn), engines_[d], "lkcuda_set_boot_samples");
}

/* rell[s] = dotProduct((float)computePatternLikelihood(), boot_samples[s]) (iqtree.cpp:2689-2735); the branch is current_it, whose
   _pattern_lh the last computeLikelihood / computeLikelihoodBranch left behind */
void PhyloTree::computeRell(double *rell) {
    if (!current_it || !current_it_back) throw std::runtime_error("computeRell: no current branch (computeLikelihood first)");
    if (boot_nsamples_ <= 0) throw std::runtime_error("computeRell: no bootstrap samples");
    for (size_t d = 0; d < engines_.size(); d++)
        check(lkcuda_rell(engines_[d], current_it->vec, current_it_back->vec, NULL), engines_[d], "lkcuda_rell");
    std::vector<double> part(boot_nsamples_);
    std::fill(rell, rell + boot_nsamples_, 0.0);
    for (size_t d = 0; d < engines_.size(); d++) {
        check(lkcuda_rell_fetch(engines_[d], &part[0]), engines_[d], "lkcuda_rell_fetch");
        for (int s = 0; s < boot_nsamples_; s++) rell[s] += part[s];
    }
    if (allreduce_) allreduce_(allreduce_user_, rell, NULL, boot_nsamples_);
}

void PhyloTree::computeFuncDerv(double value, double &df, double &ddf) {   // phylotree.cpp:2135-2146
    current_it->length = value;
    current_it_back->length = value;
    computeLikelihoodDerv(current_it, current_it_back->node, df, ddf);
    df = -df;
    ddf = -ddf;
}

/* optimization.cpp:388-465, statement for statement */
double PhyloTree::minimizeNewton(double x1, double xguess, double x2, double xacc, double &d2l, int maxNRStep) {
    int j;
    double df, dx, dxold, f;
    double temp, xh, xl, rts, rts_old;
    rts = xguess;
    if (rts < x1) rts = x1;
    if (rts > x2) rts = x2;
    computeFuncDerv(rts, f, df);
    d2l = df;
    if (!std::isfinite(f) || !std::isfinite(df)) throw std::runtime_error("Wrong computeFuncDerv");
    if (df >= 0.0 && fabs(f) < xacc) return rts;
    if (f < 0.0) { xl = rts; xh = x2; } else { xh = rts; xl = x1; }
    dx = dxold = fabs(xh - xl);
    (void)dxold;
    for (j = 1; j <= maxNRStep; j++) {
        rts_old = rts;
        if ((df <= 0.0) || (((rts - xh) * df - f) * ((rts - xl) * df - f) >= 0.0)) {
            dxold = dx; dx = 0.5 * (xh - xl); rts = xl + dx; d2l = df;
            if (xl == rts) return rts;
        } else {
            dxold = dx; dx = f / df; temp = rts; rts -= dx; d2l = df;
            if (temp == rts) return rts;
        }
        if (fabs(dx) < xacc || (j == maxNRStep)) return rts_old;
        computeFuncDerv(rts, f, df);
        if (!std::isfinite(f) || !std::isfinite(df)) throw std::runtime_error("Wrong computeFuncDerv");
        if (df > 0.0 && fabs(f) < xacc) { d2l = df; return rts; }
        if (f < 0.0) xl = rts; else xh = rts;
    }
    throw std::runtime_error("Maximum number of iterations exceeded in minimizeNewton");
}

void PhyloTree::optimizeOneBranch(Node *node1, Node *node2, bool clearLH, int maxNRStep) {   // phylotree.cpp:2148-2192
    current_it = node1->findNeighbor(node2);
    current_it_back = node2->findNeighbor(node1);
    if (!current_it || !current_it_back) throw std::runtime_error("optimizeOneBranch: nodes are not adjacent");
    double current_len = current_it->length;
    theta_computed = false;
    double optx, d2l;
    if (device_newton && engines_.size() == 1 && !allreduce_) {
        // same Newton sequence, driven on the device: one launch instead of one per step
        Node *dad = current_it_back->node;
        Neighbor *dad_branch = current_it; Node *node = dad_branch->node; Neighbor *node_branch = current_it_back;
        if (node->isLeaf()) { std::swap(dad, node); std::swap(dad_branch, node_branch); }
        if ((dad_branch->partial_lh_computed & 1) == 0) computePartialLikelihood(dad_branch, dad);
        if ((node_branch->partial_lh_computed & 1) == 0) computePartialLikelihood(node_branch, node);
        flushPartialOps();
        theta_computed = true;
        int32_t dad_end = dad->isLeaf() ? LKCUDA_TIP(dad->id) : node_branch->vec;
        if (nni_async_) {
            // queued NNI: no wait.  The run starts from the branch's current length (a host value, or the slot an earlier run of this
            // NNI pair left it in) and leaves optx in a fresh slot, which later node updates, Newton starts and the score refer to
            if (nni_next_slot_ >= LKCUDA_NLENSLOT) throw std::runtime_error("queued NNI: out of length slots");
            const int slot = nni_next_slot_++;
            check(lkcuda_theta_newton_async(engines_[0], dad_end, dad_branch->vec, min_branch_length, current_len, current_it->len_slot, max_branch_length,
                                            min_branch_length, maxNRStep, slot), engines_[0], "lkcuda_theta_newton_async");
            current_it->len_slot = slot; current_it_back->len_slot = slot;
            nni_touched_.push_back(current_it); nni_touched_.push_back(current_it_back);
            return;     // the "diverged to the upper bound" rule below is applied by getBestNNIForBran once the slots are back
        }
        double out[4];
        check(lkcuda_theta_newton(engines_[0], dad_end, dad_branch->vec, min_branch_length, current_len, max_branch_length, min_branch_length, maxNRStep, out),
              engines_[0], "lkcuda_theta_newton");
        optx = out[0]; d2l = out[1];
        // minimizeNewton leaves the branch at the last length it evaluated (computeFuncDerv, phylotree.cpp:2136-2137)
        current_it->length = out[3]; current_it_back->length = out[3];
    } else {
        optx = minimizeNewton(min_branch_length, current_len, max_branch_length, min_branch_length, d2l, maxNRStep);
    }
    if (optx > max_branch_length * 0.95) {
        // newton raphson diverged, reset
        double opt_lh = computeLikelihoodFromBuffer();
        current_it->length = current_len;
        current_it_back->length = current_len;
        double orig_lh = computeLikelihoodFromBuffer();
        if (orig_lh > opt_lh) optx = current_len;
    }
    current_it->length = optx;
    current_it_back->length = optx;
    if (clearLH && current_len != optx) {
        clearReversePartialLh(node1, node2);
        clearReversePartialLh(node2, node1);
    }
}

/* phylotree.cpp:2873-3066, statement for statement (without the upper-bound shortcut, params->upper_bound_NNI, and the bootstrap hooks):
   six (nni5) or two neighbour slots around the branch get fresh Neighbor objects whose partials live in the nni_partial_lh vectors, so the
   tree's own partials survive; per NNI: swap, clear, optimise the branches in the reference's order with at most 10 Newton steps,
   computeLikelihoodFromBuffer, swap back. */
PhyloTree::NNIMove PhyloTree::getBestNNIForBran(Node *node1, Node *node2, NNIMove moves[2], bool nni5) {
    if (node1->isLeaf() || node2->isLeaf() || node1->degree() != 3 || node2->degree() != 3)
        throw std::runtime_error("getBestNNIForBran: both ends must be internal nodes of degree 3");
    if (engines_.empty()) initializeAllPartialLh();
    const int NNI_MAX_NR_STEP = 10;                 // pllnni.cpp:21
    const int IT_NUM = nni5 ? 6 : 2;
    /* queued form: the ten branch optimisations and the two scores of the call are queued with ONE host wait per NNI;
       optimised lengths stay in device length slots until the end.  Same statements otherwise; if a run did not converge or a length
       ran into the upper bound (the rule of optimizeOneBranch that needs a host decision), the call is redone one step at a time. */
    static const bool env_off = getenv("LKTREE_NNI_QUEUED") && atoi(getenv("LKTREE_NNI_QUEUED")) == 0;      // measurement knob
    const bool queued = nni_queued && !env_off && device_newton && engines_.size() == 1 && !allreduce_ && !peers_attached_;
    int len_slots[2][5];
    for (int a = 0; a < 2; a++) for (int b = 0; b < 5; b++) len_slots[a][b] = -1;
    if (queued) { nni_async_ = true; nni_next_slot_ = 0; nni_touched_.clear(); nni_queued_calls++; }
    Neighbor **saved_it[6];
    int id = 0;
    auto slot = [](Node *at, Node *to) -> Neighbor** {
        for (size_t i = 0; i < at->neighbors.size(); i++) if (at->neighbors[i]->node == to) return &at->neighbors[i];
        throw std::runtime_error("getBestNNIForBran: nodes are not adjacent");
    };
    saved_it[id++] = slot(node1, node2);
    saved_it[id++] = slot(node2, node1);
    if (nni5) {
        FOR_NEI(node1, node2, it) saved_it[id++] = slot((*it)->node, node1);
        FOR_NEI(node2, node1, it) saved_it[id++] = slot((*it)->node, node2);
    }
    Neighbor *saved_nei[6];
    for (id = 0; id < IT_NUM; id++) {
        saved_nei[id] = *saved_it[id];
        *saved_it[id] = new Neighbor(saved_nei[id]->node, saved_nei[id]->length);
        (*saved_it[id])->id = saved_nei[id]->id;
        (*saved_it[id])->vec = nvec_ + id;          // nni_partial_lh + id*partial_lh_size
    }
    Neighbor *node12_it = node1->findNeighbor(node2), *node21_it = node2->findNeighbor(node1);
    // the two NNIs: the first neighbour of node1 (other than node2) swaps with each of node2's two (phylotree.cpp:2951-2962)
    Neighbor **node1_it = NULL, **node2_its[2] = {NULL, NULL};
    for (size_t i = 0; i < node1->neighbors.size() && !node1_it; i++) if (node1->neighbors[i]->node != node2) node1_it = &node1->neighbors[i];
    { int cnt = 0; for (size_t i = 0; i < node2->neighbors.size(); i++) if (node2->neighbors[i]->node != node1) node2_its[cnt++] = &node2->neighbors[i]; }
    for (int cnt = 0; cnt < 2; cnt++) {
        Neighbor **node2_it = node2_its[cnt];
        Neighbor *node1_nei = *node1_it, *node2_nei = *node2_it;
        moves[cnt].node1_nei = node1_nei->node->id; moves[cnt].node2_nei = node2_nei->node->id;
        // do the NNI swap
        *node1_it = node2_nei;
        node2_nei->node->findNeighbor(node2)->node = node1;
        *node2_it = node1_nei;
        node1_nei->node->findNeighbor(node1)->node = node2;
        node12_it->partial_lh_computed = 0;
        node21_it->partial_lh_computed = 0;
        int i = 1;
        if (nni5) {
            FOR_NEI(node1, node2, it) {
                (*it)->node->findNeighbor(node1)->partial_lh_computed = 0;
                optimizeOneBranch(node1, (*it)->node, false, NNI_MAX_NR_STEP);
                moves[cnt].newLen[i] = node1->findNeighbor((*it)->node)->length;
                len_slots[cnt][i] = node1->findNeighbor((*it)->node)->len_slot;
                i++;
            }
            node21_it->partial_lh_computed = 0;
        }
        optimizeOneBranch(node1, node2, false, NNI_MAX_NR_STEP);
        moves[cnt].newLen[0] = node1->findNeighbor(node2)->length;
        len_slots[cnt][0] = node1->findNeighbor(node2)->len_slot;
        if (nni5) {
            FOR_NEI(node2, node1, it) {
                (*it)->node->findNeighbor(node2)->partial_lh_computed = 0;
                optimizeOneBranch(node2, (*it)->node, false, NNI_MAX_NR_STEP);
                moves[cnt].newLen[i] = node2->findNeighbor((*it)->node)->length;
                len_slots[cnt][i] = node2->findNeighbor((*it)->node)->len_slot;
                i++;
            }
            node12_it->partial_lh_computed = 0;
        }
        for (; i < 5; i++) moves[cnt].newLen[i] = 0.0;
        if (queued) {
            // the score at the length the last run left on the device: the only host wait of this NNI
            check(lkcuda_lnl_from_buffer_slot(engines_[0], current_it->len_slot, NULL), engines_[0], "lkcuda_lnl_from_buffer_slot");
            double r[LKCUDA_NRESULT];
            reduceResult(r);
            moves[cnt].newloglh = r[LKCUDA_R_LNL];
        } else
            moves[cnt].newloglh = computeLikelihoodFromBuffer();
        // swap back
        *node1_it = node1_nei;
        node1_nei->node->findNeighbor(node2)->node = node1;
        *node2_it = node2_nei;
        node2_nei->node->findNeighbor(node1)->node = node2;
    }
    // restore the Neighbor*
    for (id = IT_NUM - 1; id >= 0; id--) {
        if (*saved_it[id] == current_it) current_it = saved_nei[id];
        if (*saved_it[id] == current_it_back) current_it_back = saved_nei[id];
        delete *saved_it[id];
        *saved_it[id] = saved_nei[id];
    }
    // restore the length of 4 branches around node1, node2
    FOR_NEI(node1, node2, it) (*it)->length = (*it)->node->findNeighbor(node1)->length;
    FOR_NEI(node2, node1, it) (*it)->length = (*it)->node->findNeighbor(node2)->length;
    if (queued) {
        nni_async_ = false;
        // the tree's own neighbours that took part only ever carried a slot number (their lengths were never overwritten); the temporary
        // ones are gone: every survivor goes back to "length is the value"
        FOR_NEI(node1, node2, it) { (*it)->len_slot = -1; (*it)->node->findNeighbor(node1)->len_slot = -1; }
        FOR_NEI(node2, node1, it) { (*it)->len_slot = -1; (*it)->node->findNeighbor(node2)->len_slot = -1; }
        node1->findNeighbor(node2)->len_slot = -1; node2->findNeighbor(node1)->len_slot = -1;
        nni_touched_.clear();
        double lens[LKCUDA_NLENSLOT], info[4 * LKCUDA_NLENSLOT];
        check(lkcuda_len_slots_fetch(engines_[0], lens, info), engines_[0], "lkcuda_len_slots_fetch");
        bool redo = false;
        for (int sl = 0; sl < nni_next_slot_; sl++)
            if ((int)info[4 * sl + 3] != 1 || lens[sl] > max_branch_length * 0.95) redo = true;
        if (redo) {
            nni_queued_fallbacks++;
            const bool keep = nni_queued;
            nni_queued = false;
            NNIMove best = getBestNNIForBran(node1, node2, moves, nni5);
            nni_queued = keep;
            return best;
        }
        for (int cnt = 0; cnt < 2; cnt++)
            for (int b = 0; b < 5; b++) if (len_slots[cnt][b] >= 0) moves[cnt].newLen[b] = lens[len_slots[cnt][b]];
    }
    return moves[0].newloglh > moves[1].newloglh ? moves[0] : moves[1];
}

void PhyloTree::saveBranchLengths(std::vector<double> &lenvec, int startid, Node *node, Node *dad) {
    if (!node) { node = root; lenvec.clear(); }
    FOR_NEI(node, dad, it) {
        lenvec.push_back((*it)->length);
        saveBranchLengths(lenvec, startid, (*it)->node, node);
    }
}

static void restoreRec(const std::vector<double> &lenvec, size_t &pos, Node *node, Node *dad) {
    FOR_NEI(node, dad, it) {
        (*it)->length = (*it)->node->findNeighbor(node)->length = lenvec[pos++];
        restoreRec(lenvec, pos, (*it)->node, node);
    }
}

void PhyloTree::restoreBranchLengths(const std::vector<double> &lenvec, int, Node *, Node *) {
    size_t pos = 0;
    restoreRec(lenvec, pos, root, NULL);
}

double PhyloTree::optimizeAllBranches(int my_iterations, double tolerance, int maxNRStep) {   // phylotree.cpp:2252-2332
    std::vector<Node*> nodes, nodes2;
    computeBestTraversal(nodes, nodes2);
    double tree_lh = computeLikelihoodBranch(nodes[0]->findNeighbor(nodes2[0]), nodes[0]);
    for (int i = 0; i < my_iterations; i++) {
        std::vector<double> lenvec;
        saveBranchLengths(lenvec);
        // the reference calls optimizeOneBranch with its defaults here (clearLH = true, maxNRStep = 100) whatever maxNRStep the caller
        // passed (phylotree.cpp:2252-2332; IQTree::optimizeNNI passes PLL_NEWZPERCYCLE = 10, iqtree.cpp:2215, without effect)
        (void)maxNRStep;
        for (size_t j = 0; j < nodes.size(); j++) optimizeOneBranch(nodes[j], nodes2[j], true, 100);
        double new_tree_lh = computeLikelihoodFromBuffer();
        if (new_tree_lh < tree_lh) {
            // IN RARE CASE: tree log-likelihood decreases, revert the branch length and stop
            clearAllPartialLH();
            restoreBranchLengths(lenvec);
            new_tree_lh = computeLikelihood();
            return new_tree_lh;
        }
        if (tree_lh <= new_tree_lh && new_tree_lh <= tree_lh + tolerance) return new_tree_lh;
        tree_lh = new_tree_lh;
    }
    return tree_lh;
}

}  // namespace lk
