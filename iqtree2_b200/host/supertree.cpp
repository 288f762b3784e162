// supertree.cpp -- b200::PhyloSuperTree: the partition dispatch of the reference over LK_CUDA contexts.
//
// The reference runs one PhyloTree per partition and fans every likelihood call out over them under OpenMP
// (tree/phylosupertree.cpp:701-725), ordered by decreasing cost (:654-700); with edge-linked proportional
// lengths (-p, PhyloSuperTreePlen) the Newton-Raphson derivatives are sum_p rate_p df_p and
// sum_p rate_p^2 ddf_p (tree/phylosupertreeplen.cpp:400-402).  Here every partition is one b200lk context with
// its own stream; a likelihood call enqueues all partitions asynchronously (b200lk_evaluate_async) and reads the
// scalars back afterwards, so the partitions overlap on the device instead of on host threads.
#include <math.h>

#include <stdio.h>
#include <stdlib.h>

#include <algorithm>
#include <chrono>
#include <stdexcept>
#include <thread>

#include "b200host.h"

namespace b200 {

PhyloSuperTree::~PhyloSuperTree() {
    destroyBatches();
    for (PhyloTree *t : parts) delete t;
}

void PhyloSuperTree::destroyBatches() {
    for (PartitionBatch &b : batches)
        if (b.handle) kernelABI().batch_destroy(b.handle);
    batches.clear();
    unbatched.clear();
}

// Partitions whose contexts have equal dimensions go into one batch; what the library cannot batch (other state
// counts, large mixtures, +ASC) and groups of one stay on their own streams.
void PhyloSuperTree::buildBatches() {
    destroyBatches();
    const KernelABI &k = kernelABI();
    const bool can = batch_partitions && k.batch_create && k.batch_destroy && k.batch_evaluate_async && k.batch_read_results && k.batch_derv &&
                     k.batch_lnl_from_buffer;
    std::vector<char> taken(parts.size(), 0);
    for (int p : part_order) {
        if (taken[p]) continue;
        taken[p] = 1;
        std::vector<int> group(1, p);
        if (can)
            for (int q : part_order) {
                if (taken[q]) continue;
                const PhyloTree *a = parts[p], *b = parts[q];
                if (a->aln->num_states == b->aln->num_states && a->aln->STATE_UNKNOWN == b->aln->STATE_UNKNOWN &&
                    a->site_rate->getNRate() == b->site_rate->getNRate() &&
                    a->model->getNMixtures() == b->model->getNMixtures() && a->leafNum == b->leafNum) {
                    taken[q] = 1;
                    group.push_back(q);
                }
            }
        if (!can || group.size() < 2) {
            unbatched.insert(unbatched.end(), group.begin(), group.end());
            continue;
        }
        std::vector<b200lk_ctx *> ctxs;
        for (int q : group) ctxs.push_back(parts[q]->ctx);
        PartitionBatch b;
        const int rc = k.batch_create(&b.handle, ctxs.data(), (int32_t)ctxs.size());
        if (rc == B200LK_ERR_UNSUPPORTED) {
            unbatched.insert(unbatched.end(), group.begin(), group.end());
            continue;
        }
        if (rc < 0) {
            std::string why = (b.handle && k.batch_last_error) ? k.batch_last_error(b.handle) : "";
            if (b.handle) k.batch_destroy(b.handle);
            outError("b200lk_batch_create failed with status " + std::to_string(rc) + " " + why);
        }
        b.members = group;
        batches.push_back(b);
    }
}

void PhyloSuperTree::addPartition(const std::string &name, Alignment *a, ModelSubst *m, RateHeterogeneity *r,
                                  double part_rate) {
    if (!root) outError("PhyloSuperTree: read the tree before adding partitions");
    // the reference maps a gene with missing taxa onto an induced subtree (SuperNeighbor::link_neighbors,
    // tree/phylosupertree.cpp:528-650); that mapping is not built here (scope): every partition must hold
    // all taxa of the super-tree -- refuse anything else instead of evaluating the wrong tree
    if (!a || (int)a->getNSeq() != leafNum)
        outError("PhyloSuperTree::addPartition(" + name + "): the alignment has " + std::to_string(a ? a->getNSeq() : 0) +
                 " sequences but the tree has " + std::to_string(leafNum) +
                 " taxa; partitions with missing taxa are not supported by this build");
    PhyloTree *t = new PhyloTree();
    t->setParams(params);
    t->copyTree(*this);
    t->setAlignment(a);
    t->setModel(m);
    t->setRate(r);
    parts.push_back(t);
    PartitionInfo info;
    info.name = name;
    info.part_rate = part_rate;
    part_info.push_back(info);
}

// tree/phylosupertree.cpp:654-700: decreasing T*N*n (here every partition has all taxa)
void PhyloSuperTree::computePartitionOrder() {
    part_order.resize(parts.size());
    for (size_t i = 0; i < parts.size(); i++) part_order[i] = (int)i;
    std::vector<double> cost(parts.size());
    for (size_t i = 0; i < parts.size(); i++)
        cost[i] = (double)parts[i]->leafNum * parts[i]->aln->getNPattern() * parts[i]->aln->num_states;
    std::stable_sort(part_order.begin(), part_order.end(), [&](int a, int b) { return cost[a] > cost[b]; });
}

void PhyloSuperTree::clearAllPartialLH() {
    PhyloTree::clearAllPartialLH();
    for (PhyloTree *t : parts) t->clearAllPartialLH();
}

void PhyloSuperTree::setLikelihoodKernel(LikelihoodKernel lk) {
    sse = lk;
    if (lk != LK_CUDA)
        outError("this build provides the LK_CUDA likelihood kernel only (no SSE/AVX variants, no CPU fallback)");
    kernelABI();
    for (PhyloTree *t : parts) t->setLikelihoodKernel(lk);
}

void PhyloSuperTree::initializeAllPartialLh() {
    computePartitionOrder();
    if (num_threads_partitions <= 0) {
        const unsigned hw = std::thread::hardware_concurrency();
        num_threads_partitions = (int)std::max(1u, std::min(hw, 16u));
    }
    for (PhyloTree *t : parts) t->initializeAllPartialLh();
    buildBatches();
    mapBranchLengths();
}

// partition trees are copies of the super-tree (copyTree): same node ids, same branch ids
void PhyloSuperTree::mapBranchLengths() {
    std::vector<double> len;
    saveBranchLengths(len);
    for (size_t p = 0; p < parts.size(); p++) {
        if (!edge_linked) continue;
        std::vector<double> scaled(len);
        for (double &x : scaled) x *= part_info[p].part_rate;
        parts[p]->restoreBranchLengths(scaled);
        parts[p]->clearAllPartialLH();
    }
}

static void rootBranch(PhyloTree *t) {  // tree/phylotree.cpp:1231-1235
    if (t->current_it) return;
    PhyloNode *leaf = t->findFarthestLeaf();
    t->current_it = leaf->neighbors[0];
    t->current_it_back = t->current_it->node->findNeighbor(leaf);
}

// enqueue the listed partitions (own stream each), then collect: the partitions overlap on the device.  Like the
// reference (tree/phylosupertree.cpp:716, OpenMP over partitions) the fan-out runs on host threads: every context is
// independently thread-safe, and many small genes are limited by how fast the host can enqueue.
void PhyloSuperTree::evaluatePartitionsAsync(const std::vector<int> &which) {
    const KernelABI &k = kernelABI();
    const int np = (int)which.size();
    std::vector<std::string> errors(np);
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic, 4) num_threads(num_threads_partitions)
#endif
    for (int i = 0; i < np; i++) {
        const int p = which[i];
        try {
            PhyloTree *t = parts[p];
            std::vector<b200lk_node> nodes;
            std::vector<b200lk_child> children;
            t->syncModel();
            rootBranch(t);
            PhyloNode *dad = t->current_it_back->node;
            t->computeTraversalInfo(t->current_it->node, dad);
            t->packTraversal(nodes, children);
            t->traversal_info.clear();
            b200lk_end de, ne;
            t->branchEnds(t->current_it, dad, de, ne);
            t->check(k.evaluate_async(t->ctx, nodes.data(), (int32_t)nodes.size(), children.data(),
                                      (int32_t)children.size(), de, ne, t->current_it->length, 1, 0),
                     "b200lk_evaluate_async");
            t->kernel_calls++;
        } catch (const std::exception &e) {
            errors[i] = e.what();
        }
    }
    for (const std::string &e : errors)
        if (!e.empty()) outError(e);
    for (int p : which) {
        double lnl = 0.0;
        parts[p]->check(k.read_results(parts[p]->ctx, &lnl, NULL, NULL), "b200lk_read_results");
        part_info[p].cur_score = lnl;
    }
}

bool PhyloSuperTree::enqueueBatch(PartitionBatch &b, bool want_derv) {
    const KernelABI &k = kernelABI();
    const int n = (int)b.members.size();
    const auto t_begin = std::chrono::steady_clock::now();
    // host part per member: model upload if it changed, traversal flags, packing -- independent per tree, so it is
    // spread over the partition threads like the reference's loop (tree/phylosupertree.cpp:716).  The traversal flags
    // of every member are advanced here, so on a mismatch the members' partials are brought up to date one by one.
    std::vector<std::vector<b200lk_node>> all_nodes(n);
    std::vector<std::vector<b200lk_child>> all_children(n);
    std::vector<b200lk_end> de(n), ne(n);
    std::vector<double> branch_len(n);
    std::vector<std::string> errors(n);
#ifdef _OPENMP
#pragma omp parallel for schedule(static) num_threads(num_threads_partitions) if (n >= 64)
#endif
    for (int i = 0; i < n; i++) {
        try {
            PhyloTree *t = parts[b.members[i]];
            t->syncModel();
            rootBranch(t);
            PhyloNode *dad = t->current_it_back->node;
            t->computeTraversalInfo(t->current_it->node, dad);
            t->packTraversal(all_nodes[i], all_children[i]);
            t->traversal_info.clear();
            t->branchEnds(t->current_it, dad, de[i], ne[i]);
            branch_len[i] = t->current_it->length;
        } catch (const std::exception &e) {
            errors[i] = e.what();
        }
    }
    for (const std::string &e : errors)
        if (!e.empty()) outError(e);
    const std::vector<b200lk_node> &nodes0 = all_nodes[0];
    const std::vector<b200lk_child> &children0 = all_children[0];
    std::vector<double> child_len((size_t)n * children0.size());
    bool same = true;
    for (int i = 0; i < n && same; i++) {
        const std::vector<b200lk_node> &nodes = all_nodes[i];
        const std::vector<b200lk_child> &children = all_children[i];
        same = nodes.size() == nodes0.size() && children.size() == children0.size() && de[i].kind == de[0].kind &&
               de[i].id == de[0].id && ne[i].kind == ne[0].kind && ne[i].id == ne[0].id;
        for (size_t j = 0; same && j < nodes.size(); j++)
            same = nodes[j].out_slot == nodes0[j].out_slot && nodes[j].child_begin == nodes0[j].child_begin &&
                   nodes[j].nchild == nodes0[j].nchild;
        for (size_t j = 0; same && j < children.size(); j++) {
            same = children[j].kind == children0[j].kind && children[j].id == children0[j].id;
            child_len[(size_t)i * children0.size() + j] = children[j].length;
        }
    }
    int rc = B200LK_ERR_UNSUPPORTED;
    static const bool timing = getenv("B200_HOST_TIMING") != NULL;
    const auto t_prep = std::chrono::steady_clock::now();
    if (same) {
        rc = k.batch_evaluate_async(b.handle, nodes0.data(), (int32_t)nodes0.size(), children0.data(),
                                    (int32_t)children0.size(), child_len.data(), de[0], ne[0], branch_len.data(),
                                    want_derv ? 1 : 0);
        batch_calls++;
        if (rc < 0 && rc != B200LK_ERR_UNSUPPORTED)
            outError(std::string("b200lk_batch_evaluate_async failed with status ") + std::to_string(rc) + " " +
                     (k.batch_last_error ? k.batch_last_error(b.handle) : ""));
    }
    if (timing)
        fprintf(stderr, "[b200host] batch of %d: host prep %.2f ms, enqueue %.2f ms (%zu nodes)\n", n,
                std::chrono::duration<double, std::milli>(t_prep - t_begin).count(),
                std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_prep).count(),
                nodes0.size());
    if (rc >= 0) {
        for (int i = 0; i < n; i++) parts[b.members[i]]->theta_computed = want_derv;
        return true;
    }
    // members out of step (or a traversal the batch cannot express): update each one's partials on its own context
    for (int i = 0; i < n; i++) {
        PhyloTree *t = parts[b.members[i]];
        t->check(k.update_partials(t->ctx, all_nodes[i].data(), (int32_t)all_nodes[i].size(), all_children[i].data(),
                                   (int32_t)all_children[i].size()),
                 "b200lk_update_partials");
        t->kernel_calls++;
        t->theta_computed = false;
    }
    return false;
}

void PhyloSuperTree::readBatch(PartitionBatch &b, std::vector<double> &lnl, std::vector<double> &df,
                               std::vector<double> &ddf) {
    const KernelABI &k = kernelABI();
    const size_t n = b.members.size();
    lnl.assign(n, 0.0);
    df.assign(n, 0.0);
    ddf.assign(n, 0.0);
    const int rc = k.batch_read_results(b.handle, lnl.data(), df.data(), ddf.data());
    if (rc < 0)
        outError(std::string("b200lk_batch_read_results failed with status ") + std::to_string(rc) + " " +
                 (k.batch_last_error ? k.batch_last_error(b.handle) : ""));
    // positive = numerical warning mask, as PhyloTree::check treats it for the per-partition calls; the library has
    // already repaired the members concerned (tree/phylokernelnew.h:3077-3094)
    if (rc & B200LK_WARN_LNL_NONFINITE) outWarning("Numerical underflow for lh-branch");
}

double PhyloSuperTree::computeLikelihood(double *pattern_lh) {
    const KernelABI &k = kernelABI();
    double total = 0.0;
    const bool async = k.evaluate_async && k.read_results && !pattern_lh;
    if (!async) {
        for (int p : part_order) {
            part_info[p].cur_score = parts[p]->computeLikelihood();
            total += part_info[p].cur_score;
        }
        curScore = total;
        return total;
    }
    // every batch is enqueued before any result is waited for: the host part of the next batch overlaps the device
    // part of the previous one
    std::vector<int> rest(unbatched);
    std::vector<double> lnl, df, ddf;
    std::vector<char> enqueued(batches.size(), 0);
    for (size_t i = 0; i < batches.size(); i++) {
        enqueued[i] = enqueueBatch(batches[i], false);
        if (!enqueued[i]) rest.insert(rest.end(), batches[i].members.begin(), batches[i].members.end());
    }
    if (!rest.empty()) evaluatePartitionsAsync(rest);
    for (size_t i = 0; i < batches.size(); i++) {
        if (!enqueued[i]) continue;
        readBatch(batches[i], lnl, df, ddf);
        for (size_t j = 0; j < batches[i].members.size(); j++) part_info[batches[i].members[j]].cur_score = lnl[j];
    }
    for (int p : part_order) total += part_info[p].cur_score;  // fixed summation order
    curScore = total;
    return total;
}

void PhyloSuperTree::computeFuncDervLinked(PhyloNode *node1, PhyloNode *node2, double value, double &df, double &ddf) {
    const KernelABI &k = kernelABI();
    df = ddf = 0.0;
    std::vector<double> pdf(parts.size(), 0.0), pddf(parts.size(), 0.0);
    std::vector<int> rest(unbatched);
    for (int p : part_order) {
        PhyloTree *t = parts[p];
        PhyloNode *a = t->all_nodes[node1->id], *b = t->all_nodes[node2->id];
        t->current_it = a->findNeighbor(b);
        t->current_it_back = b->findNeighbor(a);
        t->current_it->length = t->current_it_back->length = value * part_info[p].part_rate;
    }
    std::vector<double> lnl, d1, d2, len;
    std::vector<char> pending(batches.size(), 0);
    for (size_t bi = 0; bi < batches.size(); bi++) {  // first derivative call on this branch: traversal + theta
        bool have_theta = true;
        for (int p : batches[bi].members) have_theta &= parts[p]->theta_computed;
        if (have_theta) continue;
        if (enqueueBatch(batches[bi], true)) pending[bi] = 1;
        else pending[bi] = 2;
    }
    for (size_t bi = 0; bi < batches.size(); bi++) {
        PartitionBatch &b = batches[bi];
        const size_t n = b.members.size();
        bool done;
        if (pending[bi] == 1) {
            readBatch(b, lnl, d1, d2);
            done = true;
        } else if (pending[bi] == 2) {
            done = false;
        } else {
            len.resize(n);
            d1.assign(n, 0.0);
            d2.assign(n, 0.0);
            for (size_t i = 0; i < n; i++) len[i] = parts[b.members[i]]->current_it->length;
            const int rc = k.batch_derv(b.handle, len.data(), d1.data(), d2.data());
            batch_calls++;
            if (rc < 0)
                outError(std::string("b200lk_batch_derv failed with status ") + std::to_string(rc) + " " +
                         (k.batch_last_error ? k.batch_last_error(b.handle) : ""));
            done = true;
        }
        if (done)
            for (size_t i = 0; i < n; i++) {  // computeFuncDerv negates (tree/phylotree.cpp:2664-2672)
                pdf[b.members[i]] = -d1[i];
                pddf[b.members[i]] = -d2[i];
            }
        else
            rest.insert(rest.end(), b.members.begin(), b.members.end());
    }
    for (int p : rest) {
        PhyloTree *t = parts[p];
        t->computeFuncDerv(t->current_it->length, pdf[p], pddf[p]);
    }
    for (int p : part_order) {
        const double r = part_info[p].part_rate;
        df += r * pdf[p];
        ddf += r * r * pddf[p];
    }
}

// One pass structure of PhyloTree::optimizeAllBranches with the linked derivative in place of the single-tree one.
double PhyloSuperTree::optimizeAllBranches(int my_iterations, double tolerance, int maxNRStep) {
    if (!edge_linked) {
        double total = 0.0;
        for (int p : part_order) {
            part_info[p].cur_score = parts[p]->optimizeAllBranches(my_iterations, tolerance, maxNRStep);
            total += part_info[p].cur_score;
        }
        curScore = total;
        return total;
    }
    std::vector<PhyloNode *> nodes, nodes2;
    PhyloNode *farleaf = findFarthestLeaf();
    findFarthestLeaf(farleaf);
    getPreOrderBranches(nodes, nodes2, farleaf);
    mapBranchLengths();
    double tree_lh = computeLikelihood();
    const double lo = params ? params->min_branch_length : 1e-6, hi = params ? params->max_branch_length : 10.0;
    for (int it = 0; it < my_iterations; it++) {
        for (size_t j = 0; j < nodes.size(); j++) {
            PhyloNeighbor *nb = nodes[j]->findNeighbor(nodes2[j]), *back = nodes2[j]->findNeighbor(nodes[j]);
            for (PhyloTree *t : parts) t->theta_computed = false;
            // safeguarded Newton-Raphson on the summed derivatives (same iteration as PhyloTree::minimizeNewton)
            double rts = std::min(std::max(nb->length, lo), hi), f, df;
            computeFuncDervLinked(nodes[j], nodes2[j], rts, f, df);
            double xl, xh, best = rts;
            if (!(df >= 0.0 && fabs(f) < lo)) {
                if (f < 0.0) { xl = rts; xh = hi; } else { xh = rts; xl = lo; }
                for (int s = 1; s <= maxNRStep; s++) {
                    const double old = rts;
                    double dx;
                    if (df <= 0.0 || (((rts - xh) * df - f) * ((rts - xl) * df - f) >= 0.0)) {
                        dx = 0.5 * (xh - xl);
                        rts = xl + dx;
                        if (xl == rts) { best = rts; break; }
                    } else {
                        dx = f / df;
                        rts -= dx;
                        if (old == rts) { best = rts; break; }
                    }
                    if (fabs(dx) < lo || s == maxNRStep) { best = old; break; }
                    computeFuncDervLinked(nodes[j], nodes2[j], rts, f, df);
                    if (df > 0.0 && fabs(f) < lo) { best = rts; break; }
                    if (f < 0.0) xl = rts; else if (f > 0.0) xh = rts;
                }
            }
            const bool changed = best != nb->length;
            nb->length = back->length = best;
            for (size_t p = 0; p < parts.size(); p++) {
                PhyloTree *t = parts[p];
                PhyloNode *a = t->all_nodes[nodes[j]->id], *b = t->all_nodes[nodes2[j]->id];
                a->findNeighbor(b)->length = b->findNeighbor(a)->length = best * part_info[p].part_rate;
                if (changed) {
                    a->clearReversePartialLh(b);
                    b->clearReversePartialLh(a);
                }
            }
        }
        double new_lh = 0.0;
        {
            const KernelABI &k = kernelABI();
            std::vector<int> rest(unbatched);
            std::vector<double> len, lnl;
            for (PartitionBatch &b : batches) {
                bool have_theta = true;
                for (int p : b.members) have_theta &= parts[p]->theta_computed;
                if (!have_theta) {
                    rest.insert(rest.end(), b.members.begin(), b.members.end());
                    continue;
                }
                len.clear();
                for (int p : b.members) len.push_back(parts[p]->current_it->length);
                lnl.assign(len.size(), 0.0);
                const int rc = k.batch_lnl_from_buffer(b.handle, len.data(), lnl.data());
                batch_calls++;
                if (rc < 0) outError("b200lk_batch_lnl_from_buffer failed with status " + std::to_string(rc));
                for (size_t i = 0; i < b.members.size(); i++) part_info[b.members[i]].cur_score = lnl[i];
            }
            for (int p : rest) part_info[p].cur_score = parts[p]->computeLikelihoodFromBuffer();
            for (int p : part_order) new_lh += part_info[p].cur_score;
        }
        if (tree_lh <= new_lh && new_lh <= tree_lh + tolerance) { curScore = new_lh; return new_lh; }
        if (new_lh < tree_lh - tolerance * 0.1) { curScore = new_lh; return new_lh; }
        tree_lh = new_lh;
    }
    curScore = tree_lh;
    return tree_lh;
}

}  // namespace b200
