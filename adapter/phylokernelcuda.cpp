// tree/phylokernelcuda.cpp -- LK_CUDA for IQ-TREE 2's PhyloTree: the four likelihood-kernel pointers bound to
// libb200lk.so (include/b200lk.h).  This is the translation unit INTEGRATION.md section 2 describes, compiled against
// the reference's REAL tree/phylotree.h (plus the selector patch adapter/lk_cuda.patch).  Everything above the four
// pointers -- computeTraversalInfo, reorientPartialLh, MemSlotVector, optimizeOneBranch / minimizeNewton,
// optimizeAllBranches, getBestNNIForBran -- is the reference's own code and runs unchanged.
//
// Test infrastructure in this repository (built by adapter/Makefile into a git-ignored directory and driven by
// tests/test_adapter.py through oracle/ref_harness.cpp); for a maintainer it is the file to drop into tree/.
//
// Host buffers: the reference still sizes and allocates central_partial_lh / nni_partial_lh (tree/phylotree.cpp:
// 1054-1182).  Their CONTENTS are never touched here: a PhyloNeighbor's partial_lh pointer only names a device slot,
//     slot = (partial_lh - central_partial_lh) / block_size            (the two nni_partial_lh vectors: max_lh_slots + k)
// so reorientPartialLh (:5887-5901), mem_slots.takeover and the NNI re-pointing (:4119-4128) keep working on pointers
// that are never dereferenced.  (A production build would reserve that address range without committing memory.)
#include "phylotree.h"
#include "phylonode.h"
#include "model/modelfactory.h"
#include "model/modelsubst.h"
#include "model/rateheterogeneity.h"
#include "alignment/alignment.h"

#include <stdlib.h>
#include <string.h>

#include <string>
#include <vector>

#include "b200lk.h"

struct PhyloTreeCudaState {
    b200lk_ctx *ctx = NULL;
    size_t block_size = 0;        // doubles per partial vector in the host arena: what slot ids are derived from
    int nslots = 0;               // max_lh_slots + 2
    std::vector<double> uploaded; // signature of the model last uploaded (eigenvalues, rates, weights, tip table)
    bool eager_fetch = true;      // copy _pattern_lh / _pattern_lh_cat back after every lnL (see fetchPatternLh)
    long calls = 0;               // C-ABI calls made (statistics)
};

static void cudaCheck(PhyloTree *tree, PhyloTreeCudaState *st, int rc, const char *what, bool safe_numeric) {
    (void)tree;
    if (rc >= 0) {  // positive: numerical warning mask -- the reference's own messages (tree/phylokernelnew.h:2573-2578, 3077-3094)
        if (rc & B200LK_WARN_LNL_NONFINITE) outWarning("Numerical underflow for lh-branch");
        if (rc & B200LK_WARN_DERV_NONFINITE) {
            if (safe_numeric) outWarning("Numerical underflow for lh-derivative");
            else outError("Numerical underflow (lh-derivative). Run again with the safe likelihood kernel via `-safe` option");
        }
        return;
    }
    std::string msg = std::string(what) + " failed with status " + convertIntToString(rc);
    if (st && st->ctx) msg += std::string(": ") + b200lk_last_error(st->ctx);
    outError(msg);
}

void PhyloTree::setLikelihoodKernelCUDA() {
    vector_size = 1;                // host-visible buffers (_pattern_lh, _pattern_lh_cat) are plain pattern-major
    num_threads = 1;                // the device owns the pattern loop
    num_packets = 1;
    computePartialLikelihoodPointer = &PhyloTree::computePartialLikelihoodCUDA;
    computeLikelihoodBranchPointer = &PhyloTree::computeLikelihoodBranchCUDA;
    computeLikelihoodDervPointer = &PhyloTree::computeLikelihoodDervCUDA;
    computeLikelihoodFromBufferPointer = &PhyloTree::computeLikelihoodFromBufferCUDA;
    computeLikelihoodDervMixlenPointer = NULL;
}

void PhyloTree::releaseCUDA() {
    if (!cuda_state) return;
    if (cuda_state->ctx) b200lk_destroy(cuda_state->ctx);
    delete cuda_state;
    cuda_state = NULL;
}

// ---- helpers with access to the tree's protected fields (friend struct PhyloTreeCudaOps, lk_cuda.patch) --------------
struct PhyloTreeCudaOps {
    static int slotOf(PhyloTree *t, PhyloNeighbor *nei) {
        const double *p = nei->get_partial_lh();
        const PhyloTreeCudaState *st = t->cuda_state;
        const size_t bs = st->block_size;
        const double *central = t->central_partial_lh, *nni = t->nni_partial_lh;
        if (!p) outError("LK_CUDA: a directed branch of the traversal has no partial_lh slot");
        if (p >= central && p < central + (size_t)t->max_lh_slots * bs) {
            const size_t off = (size_t)(p - central);
            if (off % bs) outError("LK_CUDA: partial_lh does not point at a slot boundary");
            return (int)(off / bs);
        }
        if (nni && p >= nni && p < nni + 2 * bs) return (int)t->max_lh_slots + (int)((p - nni) / bs);
        outError("LK_CUDA: partial_lh points outside central_partial_lh / nni_partial_lh");
        return -1;
    }

    // device context sized like initializeAllPartialLh sized the host buffers (tree/phylotree.cpp:1054-1110)
    static void ensureContext(PhyloTree *t) {
        if (t->cuda_state && t->cuda_state->ctx) return;
        if (!t->central_partial_lh) t->initializeAllPartialLh();
        ModelSubst *model = t->getModel();
        if (model->isSiteSpecificModel() || !model->useRevKernel() || t->params->kernel_nonrev || t->isMixlen())
            outError("LK_CUDA adapter: site-specific, non-reversible and heterotachy trees use the b200lk_set_site_models / "
                     "b200lk_set_model_nonrev / b200lk_*_mixlen entry points, which this adapter does not bind yet");
        if (!t->cuda_state) t->cuda_state = new PhyloTreeCudaState();
        PhyloTreeCudaState *st = t->cuda_state;
        Alignment *aln = t->aln;
        const size_t nptn = aln->size(), nun = t->model_factory->unobserved_ptns.size();
        b200lk_dims d;
        memset(&d, 0, sizeof(d));
        d.ntaxa = (int32_t)t->leafNum;
        d.nstates = aln->num_states;
        d.ncat = t->site_rate->getNRate();
        d.nmix = model->getNMixtures();
        d.fused_mix_rate = t->model_factory->fused_mix_rate ? 1 : 0;
        d.safe_numeric = t->safe_numeric ? 1 : 0;
        d.n_state_codes = aln->STATE_UNKNOWN + 1;
        d.nslots = (int32_t)t->max_lh_slots + 2;  // + nni_partial_lh (:1066-1070)
        d.nptn = (int64_t)nptn;
        d.n_unobserved = (int64_t)nun;
        d.device = getenv("IQTREE_CUDA_DEVICE") ? atoi(getenv("IQTREE_CUDA_DEVICE")) : 0;
        st->block_size = t->getPartialLhSize();
        st->nslots = d.nslots;
        int rc = b200lk_create(&st->ctx, &d);
        if (rc != B200LK_OK || !st->ctx) {
            st->ctx = NULL;
            outError("b200lk_create failed with status " + convertIntToString(rc) +
                     (rc == B200LK_ERR_CUDA ? " (no usable CUDA device; LK_CUDA has no CPU fallback)" : ""));
        }
        // tip states, taxon-major: aln->at(ptn)[taxon] (alignment/pattern.h:29); leaf node id == alignment row
        std::vector<uint8_t> states((size_t)d.ntaxa * nptn), unobs((size_t)d.ntaxa * nun);
        for (size_t p = 0; p < nptn; p++)
            for (int s = 0; s < d.ntaxa; s++) states[(size_t)s * nptn + p] = (uint8_t)(*aln)[p][s];
        for (size_t u = 0; u < nun; u++)
            for (int s = 0; s < d.ntaxa; s++) unobs[(size_t)s * nun + u] = (uint8_t)t->model_factory->unobserved_ptns[u][s];
        if ((t->tip_partial_lh_computed & 1) == 0) t->computeTipPartialLikelihood();  // also computePtnFreq / computePtnInvar
        cudaCheck(t, st, b200lk_upload_tips(st->ctx, states.data(), nun ? unobs.data() : NULL, t->ptn_freq), "b200lk_upload_tips",
                  t->safe_numeric);
        if (nun && t->model_factory->getASC() == ASC_VARIANT_MISSING)
            cudaCheck(t, st, b200lk_set_asc(st->ctx, B200LK_ASC_HOLDER), "b200lk_set_asc", t->safe_numeric);
        if (t->params->robust_phy_keep < 1.0 || t->params->robust_median)
            cudaCheck(t, st, b200lk_set_robust(st->ctx, t->params->robust_phy_keep, t->params->robust_median ? 1 : 0),
                      "b200lk_set_robust", t->safe_numeric);
        st->uploaded.clear();
        st->calls += 2;
    }

    // the eigen-system, rates and tip table the reference's kernels would read (model/modelmarkov.cpp:1904-1916,
    // computeTipPartialLikelihood tree/phylotreesse.cpp:243-375), re-uploaded whenever they changed
    static void syncModel(PhyloTree *t) {
        ensureContext(t);
        PhyloTreeCudaState *st = t->cuda_state;
        if ((t->tip_partial_lh_computed & 1) == 0) t->computeTipPartialLikelihood();
        ModelSubst *model = t->getModel();
        const int n = t->aln->num_states, nmix = model->getNMixtures(), ncat = t->site_rate->getNRate();
        const size_t nptn = t->aln->size(), nun = t->model_factory->unobserved_ptns.size();
        std::vector<double> rates(ncat), props(ncat), mixw(nmix);
        for (int c = 0; c < ncat; c++) { rates[c] = t->site_rate->getRate(c); props[c] = t->site_rate->getProp(c); }
        for (int m = 0; m < nmix; m++) mixw[m] = model->getMixtureWeight(m);
        const double *eval = model->getEigenvalues(), *evec = model->getEigenvectors(), *inv = model->getInverseEigenvectors();
        const size_t tipsz = (size_t)(t->aln->STATE_UNKNOWN + 1) * nmix * n;
        std::vector<double> sig;
        sig.insert(sig.end(), eval, eval + (size_t)nmix * n);
        sig.insert(sig.end(), evec, evec + (size_t)nmix * n * n);
        sig.insert(sig.end(), rates.begin(), rates.end());
        sig.insert(sig.end(), props.begin(), props.end());
        sig.insert(sig.end(), mixw.begin(), mixw.end());
        sig.insert(sig.end(), t->tip_partial_lh, t->tip_partial_lh + tipsz);
        sig.push_back(t->site_rate->getPInvar());
        if (sig.size() == st->uploaded.size() && memcmp(sig.data(), st->uploaded.data(), sizeof(double) * sig.size()) == 0) return;
        cudaCheck(t, st, b200lk_set_model(st->ctx, eval, evec, inv, rates.data(), props.data(), mixw.data(), t->tip_partial_lh),
                  "b200lk_set_model", t->safe_numeric);
        // ptn_invar: observed patterns as computePtnInvar left them; the unobserved (+ASC) ones carry p_invar (:607-609)
        std::vector<double> invar(nptn + nun, t->site_rate->getPInvar());
        memcpy(invar.data(), t->ptn_invar, sizeof(double) * nptn);
        cudaCheck(t, st, b200lk_set_ptn_invar(st->ctx, invar.data()), "b200lk_set_ptn_invar", t->safe_numeric);
        st->uploaded.swap(sig);
        st->calls += 2;
    }

    // The reference's OWN traversal builder (tree/phylotree.cpp:5907-6015; driver tree/phylokernelnew.h:1150-1200):
    // flags, reorientPartialLh, memory slots and the post-order list are its work; the list then goes to the device.
    static void traverse(PhyloTree *t, PhyloNode *node, PhyloNode *dad) {
        syncModel(t);
        t->traversal_info.clear();
        if (t->params->lh_mem_save == LM_MEM_SAVE) {  // larger side first (:1172-1186)
            if (node->computeSize(dad) < dad->computeSize(node)) std::swap(node, dad);
        }
        PhyloNeighbor *dad_branch = (PhyloNeighbor *)dad->findNeighbor(node), *node_branch = (PhyloNeighbor *)node->findNeighbor(dad);
        double *buffer = t->buffer_partial_lh;  // the builder only does pointer arithmetic with it (echildren, partial_lh_leaves)
        const bool dad_locked = t->computeTraversalInfo(dad_branch, dad, buffer);
        const bool node_locked = t->computeTraversalInfo(node_branch, node, buffer);
        if (t->params->lh_mem_save == LM_MEM_SAVE) {
            if (dad_locked) t->mem_slots.unlock(dad_branch);
            if (node_locked) t->mem_slots.unlock(node_branch);
        }
        flush(t);
    }

    static void flush(PhyloTree *t) {
        if (t->traversal_info.empty()) return;
        PhyloTreeCudaState *st = t->cuda_state;
        std::vector<b200lk_node> nodes;
        std::vector<b200lk_child> kids;
        for (auto &info : t->traversal_info) {
            PhyloNode *n = (PhyloNode *)info.dad_branch->node;
            b200lk_node nd = {slotOf(t, info.dad_branch), (int32_t)kids.size(), 0, 0};
            FOR_NEIGHBOR_IT(n, info.dad, it) {
                PhyloNeighbor *c = (PhyloNeighbor *)*it;
                b200lk_child ch;
                if (c->node->isLeaf()) { ch.kind = B200LK_CHILD_LEAF; ch.id = c->node->id; }
                else { ch.kind = B200LK_CHILD_INTERNAL; ch.id = slotOf(t, c); }
                ch.length = c->length;
                kids.push_back(ch);
                nd.nchild++;
            }
            nodes.push_back(nd);
        }
        cudaCheck(t, st, b200lk_update_partials(st->ctx, nodes.data(), (int32_t)nodes.size(), kids.data(), (int32_t)kids.size()),
                  "b200lk_update_partials", t->safe_numeric);
        st->calls++;
        t->traversal_info.clear();
    }

    // b200lk wants `node` internal; the reference's kernels swap likewise when the node end is a leaf (:2660-2668)
    static void ends(PhyloTree *t, PhyloNeighbor *dad_branch, PhyloNode *dad, b200lk_end &de, b200lk_end &ne) {
        PhyloNode *node = (PhyloNode *)dad_branch->node;
        PhyloNeighbor *node_branch = (PhyloNeighbor *)node->findNeighbor(dad);
        if (node->isLeaf()) {
            std::swap(node, dad);
            std::swap(dad_branch, node_branch);
        }
        if (node->isLeaf()) outError("LK_CUDA: a two-taxon tree has no internal partial to evaluate");
        ne.kind = B200LK_CHILD_INTERNAL;
        ne.id = slotOf(t, dad_branch);
        if (dad->isLeaf()) { de.kind = B200LK_CHILD_LEAF; de.id = dad->id; }
        else { de.kind = B200LK_CHILD_INTERNAL; de.id = slotOf(t, node_branch); }
    }

    // Host readers of device data (tree/phylotree.cpp:1252-1259, 1463-1581) read _pattern_lh / _pattern_lh_cat right
    // after a likelihood call.  This adapter copies them back eagerly so that every such reader works unpatched; a
    // production build would instead fetch at those readers ("Host readers of device data").
    static void fetchPatternLh(PhyloTree *t) {
        PhyloTreeCudaState *st = t->cuda_state;
        if (!st->eager_fetch) return;
        cudaCheck(t, st, b200lk_fetch_pattern_lh(st->ctx, t->_pattern_lh), "b200lk_fetch_pattern_lh", t->safe_numeric);
        cudaCheck(t, st, b200lk_fetch_pattern_lh_cat(st->ctx, t->_pattern_lh_cat), "b200lk_fetch_pattern_lh_cat", t->safe_numeric);
        st->calls += 2;
    }
};

// ---- the four members -------------------------------------------------------------------------------------------------
void PhyloTree::computePartialLikelihoodCUDA(TraversalInfo &info, size_t, size_t, int) {
    // one directed branch; the pattern range / packet arguments of the SIMD kernels have no meaning on the device
    PhyloTreeCudaOps::syncModel(this);
    std::vector<TraversalInfo> saved;
    saved.swap(traversal_info);
    traversal_info.push_back(info);
    PhyloTreeCudaOps::flush(this);
    traversal_info.swap(saved);
}

double PhyloTree::computeLikelihoodBranchCUDA(PhyloNeighbor *dad_branch, PhyloNode *dad) {
    PhyloTreeCudaOps::traverse(this, (PhyloNode *)dad_branch->node, dad);
    b200lk_end de, ne;
    PhyloTreeCudaOps::ends(this, dad_branch, dad, de, ne);
    double lnl = 0.0;
    cudaCheck(this, cuda_state, b200lk_branch_lnl(cuda_state->ctx, de, ne, dad_branch->length, &lnl), "b200lk_branch_lnl", safe_numeric);
    cuda_state->calls++;
    PhyloTreeCudaOps::fetchPatternLh(this);
    return lnl;
}

void PhyloTree::computeLikelihoodDervCUDA(PhyloNeighbor *dad_branch, PhyloNode *dad, double *df, double *ddf) {
    if (!theta_computed) {  // tree/phylokernelnew.h:2263-2266: theta is built once per branch, callers reset the flag
        PhyloTreeCudaOps::traverse(this, (PhyloNode *)dad_branch->node, dad);
        b200lk_end de, ne;
        PhyloTreeCudaOps::ends(this, dad_branch, dad, de, ne);
        cudaCheck(this, cuda_state, b200lk_build_theta(cuda_state->ctx, de, ne), "b200lk_build_theta", safe_numeric);
        theta_computed = true;
        cuda_state->calls++;
    }
    cudaCheck(this, cuda_state, b200lk_derv(cuda_state->ctx, dad_branch->length, df, ddf), "b200lk_derv", safe_numeric);
    cuda_state->calls++;
}

double PhyloTree::computeLikelihoodFromBufferCUDA() {
    ASSERT(theta_computed && cuda_state && cuda_state->ctx);
    double lnl = 0.0;
    cudaCheck(this, cuda_state, b200lk_lnl_from_buffer(cuda_state->ctx, current_it->length, &lnl), "b200lk_lnl_from_buffer", safe_numeric);
    cuda_state->calls++;
    PhyloTreeCudaOps::fetchPatternLh(this);
    return lnl;
}
