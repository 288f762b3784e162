// mixlen.cpp -- b200::PhyloTreeMixlen: the heterotachy callers of the dispatch API (tree/phylotreemixlen.cpp),
// restricted to what reaches the likelihood kernel: per-class branch lengths, the gradient/Hessian form of
// computeLikelihoodDerv, and the EM update of one branch (the reference's default `optimize_alg_mixlen`).
#include <float.h>
#include <math.h>

#include <algorithm>

#include "b200host.h"

namespace b200 {

void PhyloTreeMixlen::initializeMixBranches(const std::vector<double> &mult) {
    if ((int)mult.size() != mixlen) outError("initializeMixBranches: one multiplier per class expected");
    for (PhyloNode *n : all_nodes)
        for (PhyloNeighbor *nb : n->neighbors) {
            nb->lengths.resize(mixlen);
            for (int c = 0; c < mixlen; c++) nb->lengths[c] = nb->length * mult[c];
        }
    clearAllPartialLH();
}

// tree/phylotreemixlen.cpp:728-737: inside the EM loop the function being minimised is the class's own
void PhyloTreeMixlen::computeFuncDerv(double value, double &df, double &ddf) {
    if (cur_mixture < 0) return PhyloTree::computeFuncDerv(value, df, ddf);
    current_it->setLength(cur_mixture, value);
    current_it_back->setLength(cur_mixture, value);
    (this->*computeLikelihoodDervMixlenPointer)(current_it, current_it_back->node, df, ddf);
    df = -df;
    ddf = -ddf;
}

// tree/phylotreemixlen.cpp:544-559: gradient (+ lnL in df[mixlen]) and Hessian at the class lengths `value`, negated
void PhyloTreeMixlen::computeFuncDervMulti(const double *value, double *df, double *ddf) {
    for (int i = 0; i < mixlen; i++) {
        current_it->setLength(i, value[i]);
        current_it_back->setLength(i, value[i]);
    }
    computeLikelihoodDerv(current_it, current_it_back->node, df, ddf);
    for (int i = 0; i <= mixlen; i++) df[i] = -df[i];
    for (int i = 0; i < mixlen * mixlen; i++) ddf[i] = -ddf[i];
}

// The EM branch of PhyloTreeMixlen::optimizeOneBranch (tree/phylotreemixlen.cpp:425-495): two EM steps; each takes the
// per-class pattern likelihoods (computePatternLhCat), turns them into posterior class weights, loads these as pattern
// frequencies and runs the safeguarded Newton-Raphson on the class's own length.
void PhyloTreeMixlen::optimizeOneBranch(PhyloNode *node1, PhyloNode *node2, bool clearLH, int maxNRStep) {
    if (model->isMixture()) outError("Please use option -optlen BFGS to disable EM algorithm");  // :427-428
    const KernelABI &k = kernelABI();
    if (!k.set_ptn_freq) outError("the kernel library has no b200lk_set_ptn_freq");
    current_it = node1->findNeighbor(node2);
    current_it_back = node2->findNeighbor(node1);
    if (!current_it || !current_it_back) outError("optimizeOneBranch: nodes are not adjacent");
    theta_computed = false;
    const size_t nptn = aln->getNPattern();
    const size_t nmix = site_rate->getNRate();
    if ((int)nmix != mixlen) outError("optimizeOneBranch: the rate model must have one category per length class");
    const double lo = params ? params->min_branch_length : 1e-6, hi = params ? params->max_branch_length : 10.0;
    double tree_lh = -DBL_MAX;
    std::vector<double> ptn_invar_host(nptn, 0.0), freq(nptn);
    if (site_rate->getPInvar() != 0.0) outError("heterotachy EM step with +I is not built in the host mirror");
    for (int EM_step = 0; EM_step < 2; EM_step++) {
        const double new_tree_lh = computePatternLhCat(WSL_RATECAT);
        if (new_tree_lh - lo < tree_lh) break;
        tree_lh = new_tree_lh;
        // E-step: posterior probability of every class, times the pattern frequency (:449-462)
        std::vector<double> post(_pattern_lh_cat);
        for (size_t ptn = 0; ptn < nptn; ptn++) {
            double *this_lk_cat = &post[ptn * nmix];
            double lk_ptn = ptn_invar_host[ptn];
            for (size_t c = 0; c < nmix; c++) lk_ptn += this_lk_cat[c];
            if (lk_ptn == 0.0) outError("optimizeOneBranch: zero pattern likelihood");
            lk_ptn = aln->ptn_freq[ptn] / lk_ptn;
            for (size_t c = 0; c < nmix; c++) this_lk_cat[c] *= lk_ptn;
        }
        theta_computed = false;
        double negative_lh;
        for (cur_mixture = 0; cur_mixture < mixlen; cur_mixture++) {  // M-step, one class at a time (:470-483)
            for (size_t ptn = 0; ptn < nptn; ptn++) freq[ptn] = post[ptn * nmix + cur_mixture];
            check(k.set_ptn_freq(ctx, freq.data()), "b200lk_set_ptn_freq");
            kernel_calls++;
            const double current_len = current_it->getLength(cur_mixture);
            const double optx = minimizeNewton(lo, current_len, hi, lo, negative_lh, maxNRStep);
            current_it->setLength(cur_mixture, optx);
            current_it_back->setLength(cur_mixture, optx);
        }
        cur_mixture = -1;
        check(k.set_ptn_freq(ctx, aln->ptn_freq.data()), "b200lk_set_ptn_freq");  // computePtnFreq again (:486-488)
        kernel_calls++;
    }
    double mean = 0.0;
    for (int c = 0; c < mixlen; c++) mean += current_it->getLength(c);
    current_it->length = current_it_back->length = mean / mixlen;
    if (clearLH) {
        node1->clearReversePartialLh(node2);
        node2->clearReversePartialLh(node1);
    }
}

}  // namespace b200
