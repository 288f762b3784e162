// host_capi.cpp -- flat C entry points over the C++ host classes, for drivers that cannot include C++ headers
// (the ctypes test and bench drivers).  Every function returns 0 on success and -1 on error with the message in
// b200host_last_error(); nothing throws across this boundary.
#include <string.h>

#include <stdexcept>

#include "b200host.h"

using namespace b200;

static thread_local std::string g_err;
#define GUARD(...)                        \
    try {                                 \
        __VA_ARGS__;                         \
        return 0;                         \
    } catch (const std::exception &e) {   \
        g_err = e.what();                 \
        return -1;                        \
    }

struct HostTree {
    PhyloTree *tree = NULL;
    PhyloSuperTree *super = NULL;
    Params params;
    std::vector<PhyloNode *> b1, b2;  // branch id -> its two nodes
};

extern "C" {

const char *b200host_last_error(void) { return g_err.c_str(); }

#ifdef B200_TESTING
int b200host_set_kernel_library(const char *path, const char *prefix) {
    GUARD(setKernelLibrary(path ? path : "", prefix ? prefix : ""))
}
#endif

// ---- alignment ---------------------------------------------------------------------------------------------
int b200host_alignment_from_states(void **out, int ntaxa, int64_t nsite, const uint8_t *codes, int seq_type, int nstates,
                                   int compress) {
    GUARD({
        Alignment *a = new Alignment();
        std::vector<std::string> names;
        for (int t = 0; t < ntaxa; t++) names.push_back("T" + std::to_string(t));
        a->buildFromStates(names, codes, nsite, (SeqType)seq_type, nstates, compress != 0);
        *out = a;
    })
}
int b200host_alignment_from_phylip(void **out, const char *path, int seq_type) {
    GUARD({
        Alignment *a = new Alignment();
        a->readPhylip(path, (SeqType)seq_type);
        *out = a;
    })
}
int64_t b200host_alignment_npattern(void *a) { return (int64_t)((Alignment *)a)->getNPattern(); }
int b200host_alignment_ntaxa(void *a) { return (int)((Alignment *)a)->getNSeq(); }
int b200host_alignment_get(void *av, uint8_t *states, double *freq, int32_t *const_char) {
    Alignment *a = (Alignment *)av;
    if (states) memcpy(states, a->states.data(), a->states.size());
    if (freq) memcpy(freq, a->ptn_freq.data(), sizeof(double) * a->ptn_freq.size());
    if (const_char)
        for (size_t i = 0; i < a->const_char.size(); i++) const_char[i] = a->const_char[i];
    return 0;
}
const char *b200host_alignment_name(void *a, int i) { return ((Alignment *)a)->seq_names[i].c_str(); }
void b200host_alignment_free(void *a) { delete (Alignment *)a; }

// ---- model, rates --------------------------------------------------------------------------------------------
int b200host_model_gtr(void **out, const double *rates6, const double *freqs4) {
    GUARD(*out = new ModelSubst(ModelSubst::GTR(rates6, freqs4)))
}
int b200host_model_named(void **out, const char *name) {  // "LG", "LG+C60", "JC"
    GUARD({
        const std::string nm(name ? name : "");
        if (nm == "LG") *out = new ModelSubst(ModelSubst::LG());
        else if (nm == "LG+C60") *out = new ModelSubst(ModelSubst::LGC60());
        else if (nm == "JC") *out = new ModelSubst(ModelSubst::JC());
        else outError("unknown model name " + nm);
    })
}
int b200host_model_reversible(void **out, int n, const double *exch, const double *freqs) {
    GUARD(*out = new ModelSubst(ModelSubst::reversible(n, exch, freqs)))
}
int b200host_model_mixture(void **out, void **classes, const double *weights, int k) {
    GUARD({
        std::vector<ModelSubst> cl;
        for (int i = 0; i < k; i++) cl.push_back(*(ModelSubst *)classes[i]);
        *out = new ModelSubst(ModelSubst::mixture(cl, std::vector<double>(weights, weights + k)));
    })
}
int b200host_model_from_eigen(void **out, int n, int nmix, const double *eval, const double *evec, const double *inv_evec,
                              const double *freqs, const double *mixw) {
    GUARD(*out = new ModelSubst(ModelSubst::fromEigen(n, nmix, eval, evec, inv_evec, freqs, mixw)))
}
int b200host_model_get(void *mv, double *eval, double *evec, double *inv_evec, double *freqs, double *mixw) {
    ModelSubst *m = (ModelSubst *)mv;
    if (eval) memcpy(eval, m->eval.data(), sizeof(double) * m->eval.size());
    if (evec) memcpy(evec, m->evec.data(), sizeof(double) * m->evec.size());
    if (inv_evec) memcpy(inv_evec, m->inv_evec.data(), sizeof(double) * m->inv_evec.size());
    if (freqs) memcpy(freqs, m->state_freq.data(), sizeof(double) * m->state_freq.size());
    if (mixw) memcpy(mixw, m->mixw.data(), sizeof(double) * m->mixw.size());
    return 0;
}
void b200host_model_free(void *m) { delete (ModelSubst *)m; }

int b200host_rate_gamma(void **out, double alpha, int ncat, double p_invar) {
    GUARD(*out = new RateHeterogeneity(ncat > 1 ? RateHeterogeneity::gamma(alpha, ncat, p_invar)
                                                : RateHeterogeneity::gamma(alpha, 1, p_invar)))
}
int b200host_rate_categories(void **out, const double *rates, const double *props, int ncat, double p_invar) {
    GUARD(*out = new RateHeterogeneity(RateHeterogeneity::fromCategories(rates, props, ncat, p_invar)))
}
int b200host_rate_get(void *rv, double *rates, double *props) {
    RateHeterogeneity *r = (RateHeterogeneity *)rv;
    memcpy(rates, r->rates.data(), sizeof(double) * r->rates.size());
    memcpy(props, r->props.data(), sizeof(double) * r->props.size());
    return 0;
}
void b200host_rate_free(void *r) { delete (RateHeterogeneity *)r; }

// ---- tree ------------------------------------------------------------------------------------------------------
// the reference's set-up order: readTree, setAlignment, setModelFactory, setLikelihoodKernel, initializeAllPartialLh
static double g_robust_keep = 1.0;  // Params::robust_phy_keep / robust_median of the next tree
static int g_robust_median = 0;
void b200host_set_robust(double keep, int median) { g_robust_keep = keep; g_robust_median = median; }
static std::vector<double> g_mixlen_mult;  // non-empty: the next tree is a PhyloTreeMixlen with these class multipliers
void b200host_set_next_mixlen(const double *mult, int nclass) { g_mixlen_mult.assign(mult, mult + nclass); }
static double g_mem_fraction = 0.0;  // Params::max_mem_size of the next -mem tree (0: the minimum number of slots)
void b200host_set_mem_fraction(double f) { g_mem_fraction = f; }

static std::vector<int> g_next_devices;  // Params::devices of the next tree (more than one: pattern shards over these GPUs)
void b200host_set_next_devices(const int32_t *devices, int n) { g_next_devices.assign(devices, devices + n); }

int b200host_tree_new(void **out, const char *newick, void *aln, void *model, void *rate, int safe, int device,
                      int device_newton) {
    HostTree *h = new HostTree();
    try {
        h->params.lk_safe_scaling = (safe & 1) != 0;
        h->params.asc_lewis = (safe & 2) != 0;  // bit 1 of `safe`: +ASC
        h->params.asc_holder = (safe & 8) != 0; // bit 3: +ASC with missing data (Holder)
        h->params.robust_phy_keep = g_robust_keep;
        h->params.robust_median = g_robust_median != 0;
        g_robust_keep = 1.0;
        g_robust_median = 0;
        if (safe & 4) {                         // bit 2: -mem (LM_MEM_SAVE)
            h->params.lh_mem_save = LM_MEM_SAVE;
            h->params.max_mem_size = g_mem_fraction;
        }
        h->params.device = device;
        h->params.devices.swap(g_next_devices);
        g_next_devices.clear();
        h->params.device_newton = device_newton != 0;
        std::vector<double> mult;
        mult.swap(g_mixlen_mult);
        if (mult.empty()) {
            h->tree = new PhyloTree();
        } else {
            PhyloTreeMixlen *mt = new PhyloTreeMixlen();
            mt->setMixlen((int)mult.size());
            h->tree = mt;
        }
        h->tree->setParams(&h->params);
        h->tree->readTreeString(newick);
        h->tree->setAlignment((Alignment *)aln);
        h->tree->setModel((ModelSubst *)model);
        h->tree->setRate((RateHeterogeneity *)rate);
        h->tree->setLikelihoodKernel(LK_CUDA);
        h->tree->initializeAllPartialLh();
        if (!mult.empty()) ((PhyloTreeMixlen *)h->tree)->initializeMixBranches(mult);
        h->tree->getBranches(h->b1, h->b2);
        *out = h;
        return 0;
    } catch (const std::exception &e) {
        g_err = e.what();
        delete h->tree;
        delete h;
        return -1;
    }
}
void b200host_tree_free(void *hv) {
    HostTree *h = (HostTree *)hv;
    delete h->tree;
    delete h->super;
    delete h;
}
static PhyloTree *T(void *hv) {
    HostTree *h = (HostTree *)hv;
    return h->super ? (PhyloTree *)h->super : h->tree;
}
int b200host_tree_nbranches(void *hv) { return T(hv)->branchNum; }
int b200host_tree_max_lh_slots(void *hv) { return T(hv)->max_lh_slots; }
int b200host_tree_set_min_branch(void *hv, double lo, double hi) {
    HostTree *h = (HostTree *)hv;
    h->params.min_branch_length = lo;
    h->params.max_branch_length = hi;
    return 0;
}
int b200host_tree_compute_likelihood(void *hv, double *pattern_lh, double *lnl) {
    GUARD(*lnl = T(hv)->computeLikelihood(pattern_lh))
}
int b200host_tree_clear_partials(void *hv) { GUARD(T(hv)->clearAllPartialLH()) }
int b200host_tree_model_changed(void *hv) { GUARD(T(hv)->modelChanged()) }
// lnL taken at branch `bid` (any branch gives the same value: the pulley principle)
int b200host_tree_branch_lnl(void *hv, int bid, double *lnl) {
    HostTree *h = (HostTree *)hv;
    GUARD(*lnl = h->tree->computeLikelihoodBranch(h->b1[bid]->findNeighbor(h->b2[bid]), h->b1[bid]))
}
int b200host_tree_derv(void *hv, int bid, double length, double *df, double *ddf) {
    HostTree *h = (HostTree *)hv;
    GUARD({
        PhyloTree *t = h->tree;
        t->current_it = h->b1[bid]->findNeighbor(h->b2[bid]);
        t->current_it_back = h->b2[bid]->findNeighbor(h->b1[bid]);
        t->theta_computed = false;
        double f, g;
        t->computeFuncDerv(length, f, g);
        *df = -f;
        *ddf = -g;
    })
}
int b200host_tree_optimize_one_branch(void *hv, int bid, int max_steps, double *length) {
    HostTree *h = (HostTree *)hv;
    GUARD({
        h->tree->optimizeOneBranch(h->b1[bid], h->b2[bid], true, max_steps);
        *length = h->b1[bid]->findNeighbor(h->b2[bid])->length;
    })
}
// computePatternLhCat(wsl): lnL and the per-pattern category likelihoods; *ncols = columns per pattern after folding
int b200host_tree_pattern_lh_cat(void *hv, int wsl, double *out, int32_t *ncols, double *lnl) {
    GUARD({
        PhyloTree *t = T(hv);
        *lnl = t->computePatternLhCat((SiteLoglType)wsl);
        const size_t np = t->aln->getNPattern();
        *ncols = (int32_t)(t->_pattern_lh_cat.size() / (np ? np : 1));
        if (out) memcpy(out, t->_pattern_lh_cat.data(), sizeof(double) * t->_pattern_lh_cat.size());
    })
}
// heterotachy: computeFuncDervMulti at branch `bid` (gradient [mixlen] + lnL, Hessian [mixlen][mixlen]; signs as the
// kernel returns them)
int b200host_tree_derv_multi(void *hv, int bid, const double *lens, double *df, double *ddf, double *lnl) {
    HostTree *h = (HostTree *)hv;
    GUARD({
        PhyloTreeMixlen *t = dynamic_cast<PhyloTreeMixlen *>(h->tree);
        if (!t) outError("not a mixed-branch-length tree");
        t->current_it = h->b1[bid]->findNeighbor(h->b2[bid]);
        t->current_it_back = h->b2[bid]->findNeighbor(h->b1[bid]);
        t->theta_computed = false;
        const int m = t->getMixlen();
        std::vector<double> g(m + 1), hs((size_t)m * m);
        t->computeFuncDervMulti(lens, g.data(), hs.data());
        for (int i = 0; i < m; i++) df[i] = -g[i];
        for (int i = 0; i < m * m; i++) ddf[i] = -hs[i];
        *lnl = -g[m];
    })
}
int b200host_tree_from_buffer(void *hv, double *lnl) { GUARD(*lnl = T(hv)->computeLikelihoodFromBuffer()) }
int b200host_tree_optimize_all_branches(void *hv, int iterations, double tolerance, double *lnl) {
    GUARD(*lnl = T(hv)->optimizeAllBranches(iterations, tolerance))
}
int b200host_tree_get_lengths(void *hv, double *out) {
    std::vector<double> v;
    T(hv)->saveBranchLengths(v);
    memcpy(out, v.data(), sizeof(double) * T(hv)->branchNum);
    return 0;
}
// heterotachy: class lengths [branchNum][mixlen] by branch id
int b200host_tree_get_mix_lengths(void *hv, double *out) {
    PhyloTree *t = T(hv);
    std::vector<double> v;
    t->saveBranchLengths(v);
    memcpy(out, v.data() + t->branchNum, sizeof(double) * (v.size() - t->branchNum));
    return 0;
}
int b200host_tree_set_lengths(void *hv, const double *in) {
    GUARD({
        PhyloTree *t = T(hv);
        t->restoreBranchLengths(std::vector<double>(in, in + t->branchNum));
        t->clearAllPartialLH();
    })
}
// leaf sets on the b1 side of every branch as alignment rows: lets a driver match branch ids to its own tree
int b200host_tree_branch_split(void *hv, int bid, int32_t *rows, int32_t *nrows) {
    HostTree *h = (HostTree *)hv;
    GUARD({
        std::vector<std::pair<PhyloNode *, PhyloNode *>> st;
        st.push_back(std::make_pair(h->b1[bid], h->b2[bid]));
        int k = 0;
        while (!st.empty()) {
            PhyloNode *n = st.back().first, *d = st.back().second;
            st.pop_back();
            if (n->isLeaf()) rows[k++] = h->tree->aln->getSeqID(n->name);
            for (PhyloNeighbor *nb : n->neighbors)
                if (nb->node != d) st.push_back(std::make_pair(nb->node, n));
        }
        *nrows = k;
    })
}
// Params::device_sweep of an existing tree, and how many optimizeAllBranches sweeps ran with device-resident lengths
int b200host_tree_set_device_sweep(void *hv, int on) {
    ((HostTree *)hv)->params.device_sweep = on != 0;
    return 0;
}
int64_t b200host_tree_device_sweeps(void *hv) { return T(hv)->device_sweeps; }

int b200host_tree_newick(void *hv, char *buf, int cap) {
    GUARD({
        std::string s = T(hv)->getTreeString(17);
        if ((int)s.size() + 1 > cap) outError("newick buffer too small");
        memcpy(buf, s.c_str(), s.size() + 1);
    })
}
int64_t b200host_tree_kernel_calls(void *hv) {
    HostTree *h = (HostTree *)hv;
    if (!h->super) return h->tree->kernel_calls;
    int64_t n = 0;
    for (PhyloTree *t : h->super->parts) n += t->kernel_calls;
    return n;
}
// kernels launched by the tree's context(s) since creation (b200lk_get_counters); -1 when the library has no counters
int64_t b200host_tree_gpu_launches(void *hv) {
    HostTree *h = (HostTree *)hv;
    const KernelABI &k = kernelABI();
    if (!k.get_counters) return -1;
    int64_t total = 0, upd = 0, n = 0;
    if (!h->super) {
        if (h->tree->ctx && k.get_counters(h->tree->ctx, &upd, &n) == 0) total = n;
        return total;
    }
    for (PhyloTree *t : h->super->parts)
        if (t->ctx && k.get_counters(t->ctx, &upd, &n) == 0) total += n;
    return total;
}
int b200host_tree_safe_numeric(void *hv) { return T(hv)->safe_numeric ? 1 : 0; }
int b200host_tree_set_kernel(void *hv, int lk) { GUARD(T(hv)->setLikelihoodKernel((LikelihoodKernel)lk)) }

// ---- partitions --------------------------------------------------------------------------------------------------
int b200host_super_new(void **out, const char *newick, int edge_linked, int device) {
    HostTree *h = new HostTree();
    try {
        h->params.device = device;
        h->super = new PhyloSuperTree();
        h->super->setParams(&h->params);
        h->super->edge_linked = edge_linked != 0;
        h->super->readTreeString(newick);
        *out = h;
        return 0;
    } catch (const std::exception &e) {
        g_err = e.what();
        delete h->super;
        delete h;
        return -1;
    }
}
int b200host_super_add(void *hv, const char *name, void *aln, void *model, void *rate, double part_rate) {
    HostTree *h = (HostTree *)hv;
    GUARD(h->super->addPartition(name, (Alignment *)aln, (ModelSubst *)model, (RateHeterogeneity *)rate, part_rate))
}
int b200host_super_init(void *hv) {
    HostTree *h = (HostTree *)hv;
    GUARD({
        h->super->setLikelihoodKernel(LK_CUDA);
        h->super->initializeAllPartialLh();
        h->super->getBranches(h->b1, h->b2);
    })
}
// batched launches over partitions: switch (before init) and what was formed
int b200host_super_set_batching(void *hv, int on) {
    ((HostTree *)hv)->super->batch_partitions = on != 0;
    return 0;
}
int b200host_super_batch_info(void *hv, int32_t *nbatches, int32_t *nbatched_parts, int64_t *batch_calls) {
    PhyloSuperTree *s = ((HostTree *)hv)->super;
    int32_t members = 0;
    for (const PhyloSuperTree::PartitionBatch &b : s->batches) members += (int32_t)b.members.size();
    *nbatches = (int32_t)s->batches.size();
    *nbatched_parts = members;
    *batch_calls = s->batch_calls;
    return 0;
}
int b200host_super_part_scores(void *hv, double *out) {
    HostTree *h = (HostTree *)hv;
    for (size_t p = 0; p < h->super->part_info.size(); p++) out[p] = h->super->part_info[p].cur_score;
    return 0;
}
int b200host_super_part_order(void *hv, int32_t *out) {
    HostTree *h = (HostTree *)hv;
    for (size_t p = 0; p < h->super->part_order.size(); p++) out[p] = h->super->part_order[p];
    return 0;
}

}  // extern "C"
