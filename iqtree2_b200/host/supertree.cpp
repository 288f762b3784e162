// supertree.cpp -- b200::PhyloSuperTree: the partition dispatch of the reference over LK_CUDA contexts.
//
// The reference runs one PhyloTree per partition and fans every likelihood call out over them under OpenMP
// (tree/phylosupertree.cpp:701-725), ordered by decreasing cost (:654-700); with edge-linked proportional
// lengths (-p, PhyloSuperTreePlen) the Newton-Raphson derivatives are sum_p rate_p df_p and
// sum_p rate_p^2 ddf_p (tree/phylosupertreeplen.cpp:400-402).  Here every partition is one b200lk context with
// its own stream; a likelihood call enqueues all partitions asynchronously (b200lk_evaluate_async) and reads the
// scalars back afterwards, so the partitions overlap on the device instead of on host threads.
#include <math.h>

#include <algorithm>
#include <stdexcept>
#include <thread>

#include "b200host.h"

namespace b200 {

PhyloSuperTree::~PhyloSuperTree() {
    for (PhyloTree *t : parts) delete t;
}

void PhyloSuperTree::addPartition(const std::string &name, Alignment *a, ModelSubst *m, RateHeterogeneity *r,
                                  double part_rate) {
    if (!root) outError("PhyloSuperTree: read the tree before adding partitions");
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
    // enqueue every partition (own stream each), then collect: the partitions overlap on the device.  Like the
    // reference (tree/phylosupertree.cpp:716, OpenMP over partitions) the fan-out runs on host threads: every
    // context is independently thread-safe, and a launch-bound supermatrix (1000 small genes) is limited by how fast
    // the host can enqueue.
    const int np = (int)part_order.size();
    std::vector<std::string> errors(np);
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic, 4) num_threads(num_threads_partitions)
#endif
    for (int i = 0; i < np; i++) {
        const int p = part_order[i];
        try {
            PhyloTree *t = parts[p];
            std::vector<b200lk_node> nodes;
            std::vector<b200lk_child> children;
            t->syncModel();
            if (!t->current_it) {
                PhyloNode *leaf = t->findFarthestLeaf();
                t->current_it = leaf->neighbors[0];
                t->current_it_back = t->current_it->node->findNeighbor(leaf);
            }
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
    for (int p : part_order) {
        double lnl = 0.0;
        parts[p]->check(k.read_results(parts[p]->ctx, &lnl, NULL, NULL), "b200lk_read_results");
        part_info[p].cur_score = lnl;
        total += lnl;
    }
    curScore = total;
    return total;
}

void PhyloSuperTree::computeFuncDervLinked(PhyloNode *node1, PhyloNode *node2, double value, double &df, double &ddf) {
    df = ddf = 0.0;
    for (int p : part_order) {
        PhyloTree *t = parts[p];
        PhyloNode *a = t->all_nodes[node1->id], *b = t->all_nodes[node2->id];
        t->current_it = a->findNeighbor(b);
        t->current_it_back = b->findNeighbor(a);
        const double r = part_info[p].part_rate;
        double d1, d2;
        t->computeFuncDerv(value * r, d1, d2);
        df += r * d1;
        ddf += r * r * d2;
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
        for (int p : part_order) {
            part_info[p].cur_score = parts[p]->computeLikelihoodFromBuffer();
            new_lh += part_info[p].cur_score;
        }
        if (tree_lh <= new_lh && new_lh <= tree_lh + tolerance) { curScore = new_lh; return new_lh; }
        if (new_lh < tree_lh - tolerance * 0.1) { curScore = new_lh; return new_lh; }
        tree_lh = new_lh;
    }
    curScore = tree_lh;
    return tree_lh;
}

}  // namespace b200
