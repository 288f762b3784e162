// dispatch.cpp -- setLikelihoodKernel, device residency, traversal building and the four dispatched members of
// b200::PhyloTree, plus the Newton-Raphson glue that calls them (see b200host.h for the reference citations).
#include <math.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>

#include "b200host.h"

namespace b200 {

void PhyloTree::check(int rc, const char *what) {
    if (rc >= 0) {
        // positive = numerical warning mask; the reference prints and carries on (tree/phylokernelnew.h:2573-2578, 3077-3094)
        if (rc & B200LK_WARN_LNL_NONFINITE) outWarning("Numerical underflow for lh-branch");
        if (rc & B200LK_WARN_DERV_NONFINITE) {
            if (safe_numeric) outWarning("Numerical underflow for lh-derivative");
            else outError("Numerical underflow (lh-derivative). Run again with the safe likelihood kernel via `-safe` option");
        }
        return;
    }
    const KernelABI &k = kernelABI();
    std::string msg = std::string(what) + " failed with status " + std::to_string(rc);
    if (k.last_error && ctx) msg += std::string(": ") + k.last_error(ctx);
    outError(msg);
}

// tree/phylotreesse.cpp:91-199.  Only LK_CUDA is built; every other value is refused (no CPU fallback).
void PhyloTree::setLikelihoodKernel(LikelihoodKernel lk) {
    sse = lk;
    vector_size = 1;
    // SAFE_NUMERIC exactly as the reference decides it (tree/phylotreesse.cpp:93-96)
    const int nstates = aln ? aln->num_states : 4;
    safe_numeric = (params && params->lk_safe_scaling) ||
                   leafNum >= (params ? params->numseq_safe_scaling : 2000) || (nstates != 4 && nstates != 20);
    computePartialLikelihoodPointer = NULL;
    computeLikelihoodBranchPointer = NULL;
    computeLikelihoodDervPointer = NULL;
    computeLikelihoodFromBufferPointer = NULL;
    if (lk != LK_CUDA)
        outError("this build provides the LK_CUDA likelihood kernel only (no SSE/AVX variants, no CPU fallback)");
    setLikelihoodKernelCUDA();
}

void PhyloTree::setLikelihoodKernelCUDA() {
    kernelABI();  // fail here, loudly, if the library cannot be loaded
    num_threads = num_packets = 1;  // the device kernels own the pattern loop (SURVEY.md 8b "Threading")
    computePartialLikelihoodPointer = &PhyloTree::computePartialLikelihoodCUDA;
    computeLikelihoodBranchPointer = &PhyloTree::computeLikelihoodBranchCUDA;
    computeLikelihoodDervPointer = &PhyloTree::computeLikelihoodDervCUDA;
    computeLikelihoodFromBufferPointer = &PhyloTree::computeLikelihoodFromBufferCUDA;
    computeLikelihoodDervMixlenPointer = &PhyloTree::computeLikelihoodDervMixlenCUDA;
}

// tree/phylotreesse.cpp:212-236
void PhyloTree::computePartialLikelihood(TraversalInfo &info, size_t ptn_lower, size_t ptn_upper, int packet_id) {
    if (!computePartialLikelihoodPointer) outError("setLikelihoodKernel has not been called");
    (this->*computePartialLikelihoodPointer)(info, ptn_lower, ptn_upper, packet_id);
}
double PhyloTree::computeLikelihoodBranch(PhyloNeighbor *dad_branch, PhyloNode *dad) {
    if (!computeLikelihoodBranchPointer) outError("setLikelihoodKernel has not been called");
    return (this->*computeLikelihoodBranchPointer)(dad_branch, dad);
}
void PhyloTree::computeLikelihoodDerv(PhyloNeighbor *dad_branch, PhyloNode *dad, double *df, double *ddf) {
    if (!computeLikelihoodDervPointer) outError("setLikelihoodKernel has not been called");
    (this->*computeLikelihoodDervPointer)(dad_branch, dad, df, ddf);
}
double PhyloTree::computeLikelihoodFromBuffer() {
    if (computeLikelihoodFromBufferPointer && optimize_by_newton)
        return (this->*computeLikelihoodFromBufferPointer)();
    return computeLikelihoodBranch(current_it, current_it_back->node);  // :230-234
}

// ---------------------------------------------------------------------------------------------------------
// device residency
// ---------------------------------------------------------------------------------------------------------
void PhyloTree::initializeAllPartialLh() {
    if (!aln || !model || !site_rate || !root) outError("initializeAllPartialLh: tree, alignment, model and rates must be set");
    if (!computeLikelihoodBranchPointer) outError("initializeAllPartialLh: call setLikelihoodKernel first");
    deleteAllPartialLh();
    const KernelABI &k = kernelABI();
    // one slot per internal node, handed to the direction that was computed last (LM_PER_NODE,
    // tree/phylotree.cpp:1131-1150); +2 scratch slots where the reference keeps nni_partial_lh (:1066-1070)
    int nint = 0;
    for (PhyloNode *n : all_nodes)
        for (PhyloNeighbor *nb : n->neighbors) { nb->slot = -1; nb->partial_lh_computed = 0; }
    const bool mem_save = params && params->lh_mem_save == LM_MEM_SAVE;
    if (!mem_save) {
        std::vector<std::pair<PhyloNode *, PhyloNode *>> stack;
        stack.push_back(std::make_pair(root, (PhyloNode *)NULL));
        while (!stack.empty()) {
            PhyloNode *node = stack.back().first, *dad = stack.back().second;
            stack.pop_back();
            if (dad && !node->isLeaf()) dad->findNeighbor(node)->slot = nint++;
            for (PhyloNeighbor *nb : node->neighbors)
                if (nb->node != dad) stack.push_back(std::make_pair(nb->node, node));
        }
        mem_slots.enabled = false;
    } else {
        // getMemoryRequired, tree/phylotree.cpp:993-1015: no slot is bound up front (:1153-1158); the traversal
        // builder allocates them
        const int full = leafNum - 2, min_slots = (int)log2((double)leafNum) + 1;  // LH_MIN_CONST = 1 (:42)
        int want = params->max_mem_size == 0.0 ? min_slots
                   : (params->max_mem_size <= 1.0 ? (int)floor(params->max_mem_size * full) : full);
        nint = std::max(min_slots, std::min(want, full));
        mem_slots.enabled = true;
        mem_slots.init(nint);
    }
    max_lh_slots = nint;
    computeSizes(root, NULL);
    for (PhyloNode *n : all_nodes)
        for (PhyloNeighbor *nb : n->neighbors)
            if (nb->size == 0) {
                PhyloNeighbor *back = nb->node->findNeighbor(n);
                nb->size = leafNum - back->size;
            }
    b200lk_dims d;
    memset(&d, 0, sizeof(d));
    d.ntaxa = leafNum;
    d.nstates = aln->num_states;
    d.ncat = site_rate->getNRate();
    d.nmix = model->getNMixtures();
    d.fused_mix_rate = 0;
    d.safe_numeric = safe_numeric ? 1 : 0;
    d.n_state_codes = aln->STATE_UNKNOWN + 1;
    d.nslots = nint + 2;
    d.nptn = (int64_t)aln->getNPattern();
    // +ASC (Lewis): the nstates constant patterns are the unobserved ones (ModelFactory, model/modelfactory.cpp:
    // "unobserved_ptns" built by Alignment::getUnobservedConstPatterns for ASC_VARIANT)
    // +ASC with missing data (Holder, ASC_VARIANT_MISSING): for every state x and observed pattern p the constant-x
    // column that keeps p's missing entries (Alignment::getUnobservedConstPatterns, alignment/alignment.cpp:4150-4172)
    const bool holder = params && params->asc_holder;
    const bool asc = params && (params->asc_lewis || holder);
    if (holder && !k.set_asc) outError("the kernel library has no Holder +ASC (b200lk_set_asc)");
    d.n_unobserved = holder ? (int64_t)aln->num_states * (int64_t)aln->getNPattern() : (asc ? aln->num_states : 0);
    d.device = params ? params->device : 0;
    int rc;
    if (params && params->devices.size() > 1) {
        if (!k.create_sharded) outError("the kernel library has no b200lk_create_sharded (Params::devices)");
        std::vector<int32_t> devs(params->devices.begin(), params->devices.end());
        rc = k.create_sharded(&ctx, &d, devs.data(), (int32_t)devs.size());
        if (rc == B200LK_ERR_UNSUPPORTED)
            outError("this model / +ASC setting cannot be split over several devices (b200lk_create_sharded refused it)");
    } else {
        if (params && params->devices.size() == 1) d.device = params->devices[0];
        rc = k.create(&ctx, &d);
    }
    if (rc != B200LK_OK || !ctx) {
        ctx = NULL;
        outError("b200lk_create failed with status " + std::to_string(rc) +
                 (rc == B200LK_ERR_CUDA ? " (no usable CUDA device; there is no CPU fallback)" : ""));
    }
    kernel_calls++;
    // tip states in alignment-row order == taxon id used in traversal descriptors
    std::vector<uint8_t> unobs;
    if (holder) {
        const size_t np = aln->getNPattern(), nu = (size_t)d.n_unobserved;
        unobs.resize((size_t)leafNum * nu);
        for (int t = 0; t < leafNum; t++)
            for (int x = 0; x < aln->num_states; x++)
                for (size_t p = 0; p < np; p++)
                    unobs[(size_t)t * nu + (size_t)x * np + p] =
                        aln->states[(size_t)t * np + p] < aln->num_states ? (uint8_t)x : (uint8_t)aln->STATE_UNKNOWN;
    } else if (asc) {
        unobs.resize((size_t)leafNum * aln->num_states);
        for (int t = 0; t < leafNum; t++)
            for (int x = 0; x < aln->num_states; x++) unobs[(size_t)t * aln->num_states + x] = (uint8_t)x;
    }
    check(k.upload_tips(ctx, aln->states.data(), asc ? unobs.data() : NULL, aln->ptn_freq.data()), "b200lk_upload_tips");
    kernel_calls++;
    if (holder) check(k.set_asc(ctx, B200LK_ASC_HOLDER), "b200lk_set_asc");
    if (params && (params->robust_phy_keep < 1.0 || params->robust_median)) {  // tree/phylokernelnew.h:3096-3135
        if (!k.set_robust) outError("the kernel library has no robust-phylogeny scoring (b200lk_set_robust)");
        check(k.set_robust(ctx, params->robust_phy_keep, params->robust_median ? 1 : 0), "b200lk_set_robust");
    }
    if (isMixlen() && (!k.update_partials_mixlen || !k.branch_lnl_mixlen || !k.derv_mixlen))
        outError("the kernel library has no mixed-branch-length entry points (b200lk_*_mixlen)");
    model_changed = true;
    buffers_ready = true;
    _pattern_lh.assign(aln->getNPattern(), 0.0);
    _pattern_lh_cat.clear();
    clearAllPartialLH();
}

// computeTipPartialLikelihood + computePtnInvar (tree/phylotreesse.cpp:243-375, 549-612) and the eigen-system upload
void PhyloTree::syncModel() {
    if (!buffers_ready) outError("initializeAllPartialLh has not been called");
    if (!model_changed) return;
    const KernelABI &k = kernelABI();
    const int n = aln->num_states, nmix = model->getNMixtures(), codes = aln->STATE_UNKNOWN + 1;
    std::vector<double> tip((size_t)codes * nmix * n), lk(n);
    for (int s = 0; s < codes; s++) {
        ModelSubst::computeTipLikelihood(aln->seq_type, n, aln->STATE_UNKNOWN, s, lk.data());
        for (int m = 0; m < nmix; m++) {
            const double *inv = model->inv_evec.data() + (size_t)m * n * n;
            for (int i = 0; i < n; i++) {
                double v = 0.0;
                for (int x = 0; x < n; x++) v += inv[i * n + x] * lk[x];
                tip[((size_t)s * nmix + m) * n + i] = v;
            }
        }
    }
    check(k.set_model(ctx, model->eval.data(), model->evec.data(), model->inv_evec.data(), site_rate->rates.data(),
                      site_rate->props.data(), model->mixw.data(), tip.data()),
          "b200lk_set_model");
    // ptn_invar
    const size_t nptn = aln->getNPattern();
    const size_t nunobs = (params && params->asc_holder) ? (size_t)n * nptn : ((params && params->asc_lewis) ? (size_t)n : 0);
    const double p_invar = site_rate->getPInvar();
    // entries of the unobserved (+ASC) patterns are the reference's "dummy values" p_invar (:607-609)
    std::vector<double> inv(nptn + nunobs, p_invar), fr(n);
    std::fill(inv.begin(), inv.begin() + nptn, 0.0);
    if (p_invar != 0.0) {
        model->getStateFrequency(fr.data(), -1);
        for (size_t p = 0; p < nptn; p++) {
            const int cc = aln->const_char[p];
            if (cc > aln->STATE_UNKNOWN) continue;
            if (cc == aln->STATE_UNKNOWN) { inv[p] = p_invar; continue; }
            ModelSubst::computeTipLikelihood(aln->seq_type, n, aln->STATE_UNKNOWN, cc, lk.data());
            double s = 0.0;
            for (int x = 0; x < n; x++) s += lk[x] * fr[x];
            inv[p] = p_invar * s;
        }
    }
    check(k.set_ptn_invar(ctx, inv.data()), "b200lk_set_ptn_invar");
    kernel_calls += 2;
    model_changed = false;
    theta_computed = false;
}

// ---------------------------------------------------------------------------------------------------------
// traversal
// ---------------------------------------------------------------------------------------------------------
void PhyloTree::reorientPartialLh(PhyloNeighbor *dad_branch, PhyloNode *dad) {
    if (dad_branch->slot >= 0) return;
    PhyloNode *node = dad_branch->node;
    for (PhyloNeighbor *nb : node->neighbors) {
        if (nb->node == dad) continue;
        PhyloNeighbor *back = nb->node->findNeighbor(node);
        if (back->slot >= 0) {
            mem_slots.takeover(dad_branch, back);  // tree/memslot.cpp:206-222
            break;
        }
    }
    if (!mem_slots.enabled && dad_branch->slot < 0) outError("partial_lh is not re-oriented");
}

// tree/phylotree.cpp:5907-6015.  Returns whether the branch's slot was locked here (the caller unlocks it once the
// node that reads it has been added).
bool PhyloTree::computeTraversalInfo(PhyloNeighbor *dad_branch, PhyloNode *dad) {
    PhyloNode *node = dad_branch->node;
    if ((dad_branch->partial_lh_computed & 1) || node->isLeaf()) return mem_slots.lock(dad_branch);
    // children in descending subtree size (:5920-5931), same exchange sort
    std::vector<PhyloNeighbor *> nv = node->neighbors;
    for (size_t i = 0; i < nv.size(); i++)
        for (size_t j = i + 1; j < nv.size(); j++)
            if (nv[i]->size < nv[j]->size) std::swap(nv[i], nv[j]);
    std::vector<char> locked(nv.size(), 0);
    for (size_t i = 0; i < nv.size(); i++)
        if (nv[i]->node != dad) locked[i] = computeTraversalInfo(nv[i], node);
    dad_branch->partial_lh_computed |= 1;
    reorientPartialLh(dad_branch, dad);
    if (dad_branch->slot < 0 || mem_slots.locked(dad_branch)) {
        // still no entry: memory-saving mode hands out a free or the smallest unlocked slot (:5951-5983)
        if (mem_slots.allocate(dad_branch) < 0) outError("No free/unlocked mem slot found!");
    } else {
        mem_slots.update(dad_branch);
    }
    if (mem_slots.enabled)
        for (size_t i = 0; i < nv.size(); i++)
            if (nv[i]->node != dad && !nv[i]->node->isLeaf() && locked[i]) mem_slots.unlock(nv[i]);
    TraversalInfo info = {dad_branch, dad};
    traversal_info.push_back(info);
    return mem_slots.lock(dad_branch);
}

// tree/phylokernelnew.h:1150-1200
void PhyloTree::computeTraversalInfo(PhyloNode *node, PhyloNode *dad) {
    traversal_info.clear();
    if (mem_slots.enabled) {  // larger side first (:1172-1186)
        PhyloNeighbor *db = dad->findNeighbor(node), *nb = node->findNeighbor(dad);
        if (db->size < nb->size) std::swap(node, dad);
    }
    PhyloNeighbor *dad_branch = dad->findNeighbor(node), *node_branch = node->findNeighbor(dad);
    const bool dad_locked = computeTraversalInfo(dad_branch, dad);
    const bool node_locked = computeTraversalInfo(node_branch, node);
    if (mem_slots.enabled) {
        if (dad_locked) mem_slots.unlock(dad_branch);
        if (node_locked) mem_slots.unlock(node_branch);
    }
}

// ---- MemSlotVector (tree/memslot.cpp) ---------------------------------------------------------------------
static const int MEM_LOCKED = 1;

void MemSlotVector::init(int num_slot) {
    assign(num_slot, MemSlot());
    reset();
}
void MemSlotVector::reset() {
    for (MemSlot &m : *this) { m.status = 0; m.nei = NULL; }
    nei_id_map.clear();
    free_count = 0;
}
int MemSlotVector::findNei(PhyloNeighbor *nei) {
    std::map<PhyloNeighbor *, int>::iterator it = nei_id_map.find(nei);
    if (it == nei_id_map.end()) outError("MemSlotVector: neighbour has no slot");
    return it->second;
}
void MemSlotVector::addNei(PhyloNeighbor *nei, int id) {
    nei->slot = id;
    (*this)[id].nei = nei;
    nei_id_map[nei] = id;
}
bool MemSlotVector::lock(PhyloNeighbor *nei) {
    if (!enabled || nei->node->isLeaf()) return false;
    MemSlot &m = (*this)[findNei(nei)];
    if (m.status & MEM_LOCKED) outError("MemSlotVector: slot is already locked");
    m.status |= MEM_LOCKED;
    return true;
}
void MemSlotVector::unlock(PhyloNeighbor *nei) {
    if (!enabled || nei->node->isLeaf()) return;
    MemSlot &m = (*this)[findNei(nei)];
    if (!(m.status & MEM_LOCKED)) outError("MemSlotVector: slot is not locked");
    m.status &= ~MEM_LOCKED;
}
bool MemSlotVector::locked(PhyloNeighbor *nei) {
    if (!enabled || nei->node->isLeaf()) return false;
    return ((*this)[findNei(nei)].status & MEM_LOCKED) != 0;
}
int MemSlotVector::allocate(PhyloNeighbor *nei) {
    if (!enabled) return -1;
    if (free_count < (int)size()) {  // a slot never used so far
        addNei(nei, free_count);
        return free_count++;
    }
    int best = -1, min_size = INT32_MAX;
    for (int i = 0; i < (int)size(); i++) {  // the unlocked slot whose subtree is cheapest to recompute
        const MemSlot &m = (*this)[i];
        if (!(m.status & MEM_LOCKED) && min_size > m.nei->size) {
            best = i;
            min_size = m.nei->size;
            if (min_size == 2) break;
        }
    }
    if (best < 0) return -1;
    (*this)[best].nei->clearPartialLh();
    addNei(nei, best);
    return best;
}
void MemSlotVector::update(PhyloNeighbor *nei) {
    if (!enabled) return;
    const int id = findNei(nei);
    if ((*this)[id].nei != nei) {  // the slot was lent to someone else meanwhile: take it back
        (*this)[id].nei->clearPartialLh();
        addNei(nei, id);
    }
}
void MemSlotVector::takeover(PhyloNeighbor *nei, PhyloNeighbor *taken_nei) {
    nei->slot = taken_nei->slot;
    taken_nei->slot = -1;
    taken_nei->partial_lh_computed &= ~1;
    if (!enabled) return;
    const int id = findNei(taken_nei);
    nei_id_map.erase(taken_nei);
    nei_id_map[nei] = id;
    if ((*this)[id].nei == taken_nei) (*this)[id].nei = nei;
}

void PhyloTree::packTraversal(std::vector<b200lk_node> &nodes, std::vector<b200lk_child> &children, std::vector<int32_t> *slots) {
    nodes.clear();
    children.clear();
    if (slots) slots->clear();
    for (const TraversalInfo &info : traversal_info) {
        PhyloNode *node = info.dad_branch->node;
        b200lk_node nd;
        nd.out_slot = info.dad_branch->slot;
        nd.child_begin = (int32_t)children.size();
        nd.nchild = 0;
        nd.reserved = 0;
        for (PhyloNeighbor *nb : node->neighbors) {
            if (nb->node == info.dad) continue;
            b200lk_child ch;
            if (nb->node->isLeaf()) {
                ch.kind = B200LK_CHILD_LEAF;
                ch.id = leaf_row[nb->node->id];
            } else {
                ch.kind = B200LK_CHILD_INTERNAL;
                ch.id = nb->slot;
                if (ch.id < 0) outError("traversal: child partial has no slot");
            }
            ch.length = nb->length;
            children.push_back(ch);
            if (slots) slots->push_back(nb->id);  // device-resident lengths are indexed by branch id
            nd.nchild++;
        }
        nodes.push_back(nd);
    }
}

void PhyloTree::flushTraversal() {
    if (traversal_info.empty()) return;
    std::vector<b200lk_node> nodes;
    std::vector<b200lk_child> children;
    packTraversal(nodes, children);
    if (isMixlen()) {  // child->getLength(c), tree/phylokernelnew.h:1017, 1069
        const int m = getMixlen();
        std::vector<double> child_len;
        child_len.reserve(children.size() * m);
        for (const TraversalInfo &info : traversal_info)
            for (PhyloNeighbor *nb : info.dad_branch->node->neighbors)
                if (nb->node != info.dad)
                    for (int c = 0; c < m; c++) child_len.push_back(nb->getLength(c));
        check(kernelABI().update_partials_mixlen(ctx, nodes.data(), (int32_t)nodes.size(), children.data(),
                                                 (int32_t)children.size(), child_len.data()),
              "b200lk_update_partials_mixlen");
    } else
    check(kernelABI().update_partials(ctx, nodes.data(), (int32_t)nodes.size(), children.data(), (int32_t)children.size()),
          "b200lk_update_partials");
    kernel_calls++;
    traversal_info.clear();
}

void PhyloTree::branchEnds(PhyloNeighbor *dad_branch, PhyloNode *dad, b200lk_end &dad_end, b200lk_end &node_end) {
    // b200lk wants `node` internal; the reference swaps likewise when the node end is a leaf
    // (tree/phylokernelnew.h:2660-2668)
    PhyloNode *node = dad_branch->node;
    PhyloNeighbor *node_branch = node->findNeighbor(dad);
    if (node->isLeaf()) {
        std::swap(node, dad);
        std::swap(dad_branch, node_branch);
    }
    if (node->isLeaf()) outError("a two-taxon tree has no internal partial to evaluate");
    node_end.kind = B200LK_CHILD_INTERNAL;
    node_end.id = dad_branch->slot;
    if (dad->isLeaf()) {
        dad_end.kind = B200LK_CHILD_LEAF;
        dad_end.id = leaf_row[dad->id];
    } else {
        dad_end.kind = B200LK_CHILD_INTERNAL;
        dad_end.id = node_branch->slot;
    }
}

// ---------------------------------------------------------------------------------------------------------
// the four LK_CUDA members
// ---------------------------------------------------------------------------------------------------------
void PhyloTree::computePartialLikelihoodCUDA(TraversalInfo &info, size_t, size_t, int) {
    // one directed branch; the pattern range / packet arguments of the SIMD kernels have no meaning on the device
    syncModel();
    std::vector<TraversalInfo> saved;
    saved.swap(traversal_info);
    traversal_info.push_back(info);
    flushTraversal();
    traversal_info.swap(saved);
}

double PhyloTree::computeLikelihoodBranchCUDA(PhyloNeighbor *dad_branch, PhyloNode *dad) {
    syncModel();
    PhyloNode *node = dad_branch->node;
    computeTraversalInfo(node, dad);
    flushTraversal();
    b200lk_end de, ne;
    branchEnds(dad_branch, dad, de, ne);
    double lnl = 0.0;
    if (isMixlen()) {  // dad_branch->getLength(mycat), :2711, 2725
        std::vector<double> lens;
        branchLengths(dad_branch, lens);
        check(kernelABI().branch_lnl_mixlen(ctx, de, ne, lens.data(), &lnl), "b200lk_branch_lnl_mixlen");
    } else
    check(kernelABI().branch_lnl(ctx, de, ne, dad_branch->length, &lnl), "b200lk_branch_lnl");
    kernel_calls++;
    return lnl;
}

void PhyloTree::branchLengths(const PhyloNeighbor *nb, std::vector<double> &out) const {
    const int m = site_rate->getNRate();
    out.resize(m);
    for (int c = 0; c < m; c++) out[c] = nb->getLength(c);
}

void PhyloTree::ensureTheta(PhyloNeighbor *dad_branch, PhyloNode *dad) {
    if (theta_computed) return;  // tree/phylokernelnew.h:2263-2266, 2554
    PhyloNode *node = dad_branch->node;
    computeTraversalInfo(node, dad);
    flushTraversal();
    b200lk_end de, ne;
    branchEnds(dad_branch, dad, de, ne);
    check(kernelABI().build_theta(ctx, de, ne), "b200lk_build_theta");
    kernel_calls++;
    theta_computed = true;
}

void PhyloTree::computeLikelihoodDervCUDA(PhyloNeighbor *dad_branch, PhyloNode *dad, double *df, double *ddf) {
    syncModel();
    const KernelABI &k = kernelABI();
    ensureTheta(dad_branch, dad);
    if (isMixlen()) {
        // the reference's isMixlen() form (:2365-2462, 2556-2568): df[mixlen] gradient + the lnL in df[mixlen],
        // ddf[mixlen][mixlen] Hessian
        const int m = getMixlen();
        std::vector<double> lens;
        branchLengths(dad_branch, lens);
        check(k.derv_mixlen(ctx, lens.data(), df, ddf, df + m), "b200lk_derv_mixlen");
    } else
    check(k.derv(ctx, dad_branch->length, df, ddf), "b200lk_derv");
    kernel_calls++;
}

// computeLikelihoodDervMixlen*SIMD (tree/phylokernelnew.h:3397-3580): the class getCurMixture() alone
void PhyloTree::computeLikelihoodDervMixlenCUDA(PhyloNeighbor *dad_branch, PhyloNode *dad, double &df, double &ddf) {
    syncModel();
    const KernelABI &k = kernelABI();
    if (!k.derv_mixlen_class) outError("the kernel library has no b200lk_derv_mixlen_class");
    ensureTheta(dad_branch, dad);
    const int cls = getCurMixture();
    check(k.derv_mixlen_class(ctx, cls, dad_branch->getLength(cls), &df, &ddf), "b200lk_derv_mixlen_class");
    kernel_calls++;
}

double PhyloTree::computeLikelihoodFromBufferCUDA() {
    if (!theta_computed) outError("computeLikelihoodFromBuffer: theta has not been computed");
    double lnl = 0.0;
    if (isMixlen()) {  // current_it->getLength(mycat), :3248, 3263
        std::vector<double> lens;
        branchLengths(current_it, lens);
        check(kernelABI().lnl_from_buffer_mixlen(ctx, lens.data(), &lnl), "b200lk_lnl_from_buffer_mixlen");
    } else
    check(kernelABI().lnl_from_buffer(ctx, current_it->length, &lnl), "b200lk_lnl_from_buffer");
    kernel_calls++;
    return lnl;
}

// ---------------------------------------------------------------------------------------------------------
// callers
// ---------------------------------------------------------------------------------------------------------
double PhyloTree::computeLikelihood(double *pattern_lh) {
    if (!current_it) {  // tree/phylotree.cpp:1231-1235
        PhyloNode *leaf = findFarthestLeaf();
        current_it = leaf->neighbors[0];
        current_it_back = current_it->node->findNeighbor(leaf);
    }
    const double score = computeLikelihoodBranch(current_it, current_it_back->node);
    if (pattern_lh) {
        check(kernelABI().fetch_pattern_lh(ctx, _pattern_lh.data()), "b200lk_fetch_pattern_lh");
        memcpy(pattern_lh, _pattern_lh.data(), sizeof(double) * aln->getNPattern());
    }
    curScore = score;
    return score;
}

void PhyloTree::computePatternLikelihood(double *pattern_lh, double *cur_logl, double *pattern_lh_cat) {
    // tree/phylotree.cpp:1450-1595 for the reversible kernel: per-pattern log-likelihoods of the current root branch
    const double score = computeLikelihood(pattern_lh);
    if (cur_logl) *cur_logl = score;
    if (pattern_lh_cat) {
        const size_t ncm = (size_t)site_rate->getNRate() * model->getNMixtures();
        _pattern_lh_cat.resize(aln->getNPattern() * ncm);
        check(kernelABI().fetch_pattern_lh_cat(ctx, _pattern_lh_cat.data()), "b200lk_fetch_pattern_lh_cat");
        memcpy(pattern_lh_cat, _pattern_lh_cat.data(), sizeof(double) * _pattern_lh_cat.size());
    }
}

// tree/phylotree.cpp:1330-1388 (transformPatternLhCat is the identity: vector_size == 1)
double PhyloTree::computePatternLhCat(SiteLoglType wsl) {
    if (!current_it) {
        PhyloNode *leaf = findFarthestLeaf();
        current_it = leaf->neighbors[0];
        current_it_back = current_it->node->findNeighbor(leaf);
    }
    const double score = computeLikelihoodBranch(current_it, current_it_back->node);
    const size_t nptn = aln->getNPattern(), ncat = site_rate->getNRate(), nmixture = model->getNMixtures();
    _pattern_lh_cat.resize(nptn * ncat * nmixture);
    check(kernelABI().fetch_pattern_lh_cat(ctx, _pattern_lh_cat.data()), "b200lk_fetch_pattern_lh_cat");
    kernel_calls++;
    if (model->isMixture() && (wsl == WSL_MIXTURE || wsl == WSL_RATECAT)) {  // unfused mixtures only exist here
        double *lh_cat = _pattern_lh_cat.data(), *lh_res = _pattern_lh_cat.data();
        if (wsl == WSL_MIXTURE && ncat > 1) {  // likelihood per mixture class (:1356-1366)
            for (size_t ptn = 0; ptn < nptn; ptn++) {
                for (size_t m = 0; m < nmixture; m++) {
                    double lh = lh_cat[0];
                    for (size_t c = 1; c < ncat; c++) lh += lh_cat[c];
                    lh_res[m] = lh;
                    lh_cat += ncat;
                }
                lh_res += nmixture;
            }
            _pattern_lh_cat.resize(nptn * nmixture);
        } else if (wsl == WSL_RATECAT && nmixture > 1) {  // likelihood per rate category (:1367-1381)
            for (size_t ptn = 0; ptn < nptn; ptn++) {
                if (lh_res != lh_cat) memmove(lh_res, lh_cat, ncat * sizeof(double));
                lh_cat += ncat;
                for (size_t m = 1; m < nmixture; m++) {
                    for (size_t c = 0; c < ncat; c++) lh_res[c] += lh_cat[c];
                    lh_cat += ncat;
                }
                lh_res += ncat;
            }
            _pattern_lh_cat.resize(nptn * ncat);
        }
    }
    curScore = score;
    return score;
}

double PhyloTree::computeFunction(double value) {
    current_it->length = value;
    current_it_back->length = value;
    return -computeLikelihoodBranch(current_it, current_it_back->node);
}

void PhyloTree::computeFuncDerv(double value, double &df, double &ddf) {
    current_it->length = value;
    current_it_back->length = value;
    computeLikelihoodDerv(current_it, current_it_back->node, &df, &ddf);
    df = -df;
    ddf = -ddf;
}

// Safeguarded Newton-Raphson on f = -dlnL/dt (utils/optimization.cpp:422-499): bisect when the function is concave
// or the Newton step leaves the bracket, otherwise step; stop when the step is below xacc or the gradient vanishes.
double PhyloTree::minimizeNewton(double x1, double xguess, double x2, double xacc, double &d2l, int maxNRStep) {
    double rts = std::min(std::max(xguess, x1), x2);
    double f, df;
    computeFuncDerv(rts, f, df);
    d2l = df;
    if (!std::isfinite(f) || !std::isfinite(df)) outError("Wrong computeFuncDerv");
    if (df >= 0.0 && fabs(f) < xacc) return rts;
    double xl, xh;
    if (f < 0.0) { xl = rts; xh = x2; } else { xh = rts; xl = x1; }
    double dx = fabs(xh - xl);
    for (int j = 1; j <= maxNRStep; j++) {
        const double rts_old = rts;
        if (df <= 0.0 || (((rts - xh) * df - f) * ((rts - xl) * df - f) >= 0.0)) {
            dx = 0.5 * (xh - xl);
            rts = xl + dx;
            d2l = df;
            if (xl == rts) return rts;
        } else {
            dx = f / df;
            const double temp = rts;
            rts -= dx;
            d2l = df;
            if (temp == rts) return rts;
        }
        if (fabs(dx) < xacc || j == maxNRStep) return rts_old;
        computeFuncDerv(rts, f, df);
        if (!std::isfinite(f) || !std::isfinite(df)) outError("Wrong computeFuncDerv");
        if (df > 0.0 && fabs(f) < xacc) { d2l = df; return rts; }
        if (f < 0.0) xl = rts;
        else if (f > 0.0) xh = rts;
    }
    outError("Maximum number of iterations exceeded in minimizeNewton");
}

void PhyloTree::optimizeOneBranch(PhyloNode *node1, PhyloNode *node2, bool clearLH, int maxNRStep) {
    current_it = node1->findNeighbor(node2);
    current_it_back = node2->findNeighbor(node1);
    if (!current_it || !current_it_back) outError("optimizeOneBranch: nodes are not adjacent");
    const double current_len = current_it->length;
    theta_computed = false;
    const double lo = params ? params->min_branch_length : 1e-6, hi = params ? params->max_branch_length : 10.0;
    double negative_lh, optx;
    const KernelABI &k = kernelABI();
    if (k.optimize_branch && (!params || params->device_newton) && !isMixlen()) {
        // the same iteration, run on the device in one call: partial updates still pending for the two directions
        // are flushed first, then theta + every derivative evaluation + the safeguarded update stay in HBM
        syncModel();
        computeTraversalInfo(current_it->node, current_it_back->node);
        flushTraversal();
        b200lk_end de, ne;
        branchEnds(current_it, current_it_back->node, de, ne);
        b200lk_newton nr;
        nr.min_len = lo; nr.max_len = hi; nr.xacc = lo; nr.guess = current_len;
        nr.max_steps = maxNRStep; nr.reserved = 0;
        int32_t steps = 0;
        const int rc = k.optimize_branch(ctx, de, ne, &nr, &optx, &negative_lh, &steps);
        kernel_calls++;
        if (rc == B200LK_ERR_UNSUPPORTED) {  // e.g. +ASC: the iteration is driven from the host instead
            optx = minimizeNewton(lo, current_len, hi, lo, negative_lh, maxNRStep);
        } else {
            check(rc, "b200lk_optimize_branch");
            theta_computed = true;
            current_it->length = current_it_back->length = optx;
        }
    } else {
        optx = minimizeNewton(lo, current_len, hi, lo, negative_lh, maxNRStep);
    }
    if (optx > hi * 0.95) {  // Newton-Raphson diverged: keep the better of the two (tree/phylotree.cpp:2607-2616)
        current_it->length = current_it_back->length = optx;
        const double opt_lh = computeLikelihoodFromBuffer();
        current_it->length = current_it_back->length = current_len;
        const double orig_lh = computeLikelihoodFromBuffer();
        if (orig_lh > opt_lh) optx = current_len;
    }
    current_it->length = optx;
    current_it_back->length = optx;
    if (clearLH && current_len != optx) {
        node1->clearReversePartialLh(node2);
        node2->clearReversePartialLh(node1);
    }
}

// One sweep over the branches with the lengths resident on the device (b200lk_*_slots, include/b200lk.h): every branch is
// enqueued -- the partial update across the branch just optimised, theta, the whole Newton-Raphson iteration -- and the
// lengths come back in ONE copy at the end.  The per-branch host logic of optimizeOneBranch that needs the optimum
// (tree/phylotree.cpp:2607-2616: an optimum above 0.95 max_branch_length is compared with the old length) cannot run
// inside such a sweep; it is checked afterwards, and when it would have fired the caller repeats the sweep branch by branch.
int PhyloTree::sweepOnDevice(const std::vector<PhyloNode *> &nodes, const std::vector<PhyloNode *> &nodes2, int maxNRStep) {
    const KernelABI &k = kernelABI();
    if (!k.set_lengths || !k.get_lengths || !k.update_partials_slots || !k.optimize_branch_slot || !k.optimize_branch) return 0;
    if (!params || !params->device_sweep || !params->device_newton || isMixlen() || params->devices.size() > 1) return 0;
    if (params->asc_lewis || params->asc_holder || aln->num_states > 64) return 0;
    syncModel();
    std::vector<double> len, aux;
    saveBranchLengths(len);
    len.resize(branchNum);
    if (k.set_lengths(ctx, len.data(), branchNum) != B200LK_OK) return 0;
    kernel_calls++;
    const double lo = params->min_branch_length, hi = params->max_branch_length;
    std::vector<b200lk_node> tn;
    std::vector<b200lk_child> tc;
    std::vector<int32_t> slots;
    for (size_t j = 0; j < nodes.size(); j++) {
        current_it = nodes[j]->findNeighbor(nodes2[j]);
        current_it_back = nodes2[j]->findNeighbor(nodes[j]);
        theta_computed = false;
        computeTraversalInfo(current_it->node, current_it_back->node);
        if (!traversal_info.empty()) {
            packTraversal(tn, tc, &slots);
            traversal_info.clear();
            const int rc = k.update_partials_slots(ctx, tn.data(), (int32_t)tn.size(), tc.data(), (int32_t)tc.size(), slots.data());
            kernel_calls++;
            if (rc == B200LK_ERR_UNSUPPORTED) return -1;  // e.g. a multifurcation on the 4-state path
            check(rc, "b200lk_update_partials_slots");
        }
        b200lk_end de, ne;
        branchEnds(current_it, current_it_back->node, de, ne);
        b200lk_newton nr;
        nr.min_len = lo; nr.max_len = hi; nr.xacc = lo; nr.guess = current_it->length;  // the kernel starts from lengths[id]
        nr.max_steps = maxNRStep; nr.reserved = 0;
        const int rc = k.optimize_branch_slot(ctx, de, ne, &nr, current_it->id);
        kernel_calls++;
        if (rc == B200LK_ERR_UNSUPPORTED) return -1;
        check(rc, "b200lk_optimize_branch_slot");
        theta_computed = true;
        // the new length is not known here: treat it as changed (optimizeOneBranch clears only when it did change)
        nodes[j]->clearReversePartialLh(nodes2[j]);
        nodes2[j]->clearReversePartialLh(nodes[j]);
    }
    aux.assign((size_t)2 * branchNum, 0.0);
    check(k.get_lengths(ctx, len.data(), aux.data(), branchNum), "b200lk_get_lengths");
    kernel_calls++;
    for (int b = 0; b < branchNum; b++) {
        if (aux[2 * b + 1] != 0.0) check(B200LK_WARN_DERV_NONFINITE, "b200lk_optimize_branch_slot");
        if (!(len[b] <= hi * 0.95)) return -1;  // optimizeOneBranch would have compared with the old length here
    }
    restoreBranchLengths(len);
    device_sweeps++;
    return 1;
}

double PhyloTree::optimizeAllBranches(int my_iterations, double tolerance, int maxNRStep) {
    std::vector<PhyloNode *> nodes, nodes2;
    PhyloNode *farleaf = findFarthestLeaf();  // computeBestTraversal, tree/phylotree.cpp:2680-2687
    findFarthestLeaf(farleaf);
    getPreOrderBranches(nodes, nodes2, farleaf);
    double tree_lh = computeLikelihoodBranch(nodes[0]->findNeighbor(nodes2[0]), nodes[0]);
    std::vector<double> lenvec;
    for (int i = 0; i < my_iterations; i++) {
        saveBranchLengths(lenvec);
        const int swept = sweepOnDevice(nodes, nodes2, maxNRStep);
        if (swept < 0) {  // given up midway (a per-branch check of optimizeOneBranch fired): back to where the sweep started
            restoreBranchLengths(lenvec);
            clearAllPartialLH();
        }
        if (swept <= 0)  // the reference's loop, branch by branch
            for (size_t j = 0; j < nodes.size(); j++) optimizeOneBranch(nodes[j], nodes2[j], true, maxNRStep);
        double new_tree_lh = computeLikelihoodFromBuffer();
        if (new_tree_lh < tree_lh - tolerance * 0.1) {
            // rare: the likelihood decreased; restore and stop (tree/phylotree.cpp:2724-2752)
            clearAllPartialLH();
            restoreBranchLengths(lenvec);
            new_tree_lh = computeLikelihood();
            if (fabs(new_tree_lh - tree_lh) > 1.0) outError("optimizeAllBranches: likelihood not restored");
            return new_tree_lh;
        }
        if (tree_lh <= new_tree_lh && new_tree_lh <= tree_lh + tolerance) {
            curScore = new_tree_lh;
            return new_tree_lh;
        }
        tree_lh = new_tree_lh;
    }
    curScore = tree_lh;
    return tree_lh;
}

}  // namespace b200
