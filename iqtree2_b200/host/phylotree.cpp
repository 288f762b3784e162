// phylotree.cpp -- b200::PhyloTree: the reference's kernel-dispatch API over the C ABI (see b200host.h).
#include <dlfcn.h>
#include <math.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <sstream>

#include "b200host.h"

namespace b200 {

// ---------------------------------------------------------------------------------------------------------
// kernel library
// ---------------------------------------------------------------------------------------------------------
static KernelABI g_abi = {};
static std::string g_lib_path, g_lib_prefix = "b200lk_";

static std::string default_library_path() {
    Dl_info info;
    if (dladdr((void *)&default_library_path, &info) && info.dli_fname) {
        std::string p(info.dli_fname);
        size_t k = p.find_last_of('/');
        return (k == std::string::npos ? std::string(".") : p.substr(0, k)) + "/libb200lk.so";
    }
    return "libb200lk.so";
}

#ifdef B200_TESTING
// Test builds only (libb200host_testing.so, -DB200_TESTING): the product library binds libb200lk.so beside it and
// nothing else -- there is no way to re-point it at another backend.
void setKernelLibrary(const std::string &path, const std::string &prefix) {
    if (g_abi.dl) { dlclose(g_abi.dl); g_abi = KernelABI(); }
    g_lib_path = path;
    g_lib_prefix = prefix.empty() ? "b200lk_" : prefix;
}
#endif

template <class F>
static void bind(F &fn, const char *name, bool required) {
    const std::string sym = g_lib_prefix + name;
    fn = (F)dlsym(g_abi.dl, sym.c_str());
    if (!fn && required) outError("kernel library " + g_abi.path + " does not export " + sym);
}

const KernelABI &kernelABI() {
    if (g_abi.dl) return g_abi;
    g_abi.path = g_lib_path.empty() ? default_library_path() : g_lib_path;
    g_abi.prefix = g_lib_prefix;
    g_abi.dl = dlopen(g_abi.path.c_str(), RTLD_NOW | RTLD_LOCAL);
    if (!g_abi.dl)
        outError("cannot load the CUDA likelihood kernel " + g_abi.path + " (" + dlerror() +
                 "); build it first -- there is no CPU fallback on this path");
    bind(g_abi.create, "create", true);
    bind(g_abi.create_sharded, "create_sharded", false);
    bind(g_abi.destroy, "destroy", true);
    bind(g_abi.upload_tips, "upload_tips", true);
    bind(g_abi.set_ptn_invar, "set_ptn_invar", true);
    bind(g_abi.set_model, "set_model", true);
    bind(g_abi.update_partials, "update_partials", true);
    bind(g_abi.branch_lnl, "branch_lnl", true);
    bind(g_abi.build_theta, "build_theta", true);
    bind(g_abi.derv, "derv", true);
    bind(g_abi.lnl_from_buffer, "lnl_from_buffer", true);
    bind(g_abi.fetch_pattern_lh, "fetch_pattern_lh", true);
    bind(g_abi.fetch_pattern_lh_cat, "fetch_pattern_lh_cat", true);
    bind(g_abi.fetch_scale_num, "fetch_scale_num", true);
    bind(g_abi.evaluate_async, "evaluate_async", false);
    bind(g_abi.read_results, "read_results", false);
    bind(g_abi.optimize_branch, "optimize_branch", false);
    bind(g_abi.set_lengths, "set_lengths", false);
    bind(g_abi.get_lengths, "get_lengths", false);
    bind(g_abi.update_partials_slots, "update_partials_slots", false);
    bind(g_abi.optimize_branch_slot, "optimize_branch_slot", false);
    bind(g_abi.set_asc, "set_asc", false);
    bind(g_abi.set_robust, "set_robust", false);
    bind(g_abi.set_ptn_freq, "set_ptn_freq", false);
    bind(g_abi.update_partials_mixlen, "update_partials_mixlen", false);
    bind(g_abi.branch_lnl_mixlen, "branch_lnl_mixlen", false);
    bind(g_abi.lnl_from_buffer_mixlen, "lnl_from_buffer_mixlen", false);
    bind(g_abi.derv_mixlen, "derv_mixlen", false);
    bind(g_abi.derv_mixlen_class, "derv_mixlen_class", false);
    bind(g_abi.last_error, "last_error", false);
    bind(g_abi.get_counters, "get_counters", false);
    bind(g_abi.batch_create, "batch_create", false);
    bind(g_abi.batch_destroy, "batch_destroy", false);
    bind(g_abi.batch_last_error, "batch_last_error", false);
    bind(g_abi.batch_evaluate, "batch_evaluate", false);
    bind(g_abi.batch_evaluate_async, "batch_evaluate_async", false);
    bind(g_abi.batch_read_results, "batch_read_results", false);
    bind(g_abi.batch_derv, "batch_derv", false);
    bind(g_abi.batch_lnl_from_buffer, "batch_lnl_from_buffer", false);
    return g_abi;
}

// ---------------------------------------------------------------------------------------------------------
// nodes
// ---------------------------------------------------------------------------------------------------------
PhyloNeighbor *PhyloNode::findNeighbor(const PhyloNode *other) const {
    for (PhyloNeighbor *nb : neighbors)
        if (nb->node == other) return nb;
    return NULL;
}

void PhyloNode::clearReversePartialLh(PhyloNode *dad) {  // tree/phylonode.cpp:22-33
    for (PhyloNeighbor *nb : neighbors)
        if (nb->node != dad) {
            PhyloNeighbor *back = nb->node->findNeighbor(this);
            back->partial_lh_computed = 0;
            nb->node->clearReversePartialLh(this);
        }
}

void PhyloNode::clearAllPartialLh(PhyloNode *dad) {
    for (PhyloNeighbor *nb : neighbors) {
        nb->partial_lh_computed = 0;
        nb->node->findNeighbor(this)->partial_lh_computed = 0;
        if (nb->node != dad) nb->node->clearAllPartialLh(this);
    }
}

// ---------------------------------------------------------------------------------------------------------
// construction, Newick
// ---------------------------------------------------------------------------------------------------------
PhyloTree::PhyloTree() {}

void PhyloTree::freeNodes() {
    for (PhyloNode *n : all_nodes) {
        for (PhyloNeighbor *nb : n->neighbors) delete nb;
        delete n;
    }
    all_nodes.clear();
    root = NULL;
    current_it = current_it_back = NULL;
}

PhyloTree::~PhyloTree() {
    deleteAllPartialLh();
    freeNodes();
}

void PhyloTree::deleteAllPartialLh() {
    if (ctx) {
        kernelABI().destroy(ctx);
        ctx = NULL;
    }
    buffers_ready = false;
}

namespace {
struct NewickParser {
    const std::string &s;
    size_t pos = 0;
    PhyloTree &tree;
    int next_branch = 0;
    NewickParser(const std::string &text, PhyloTree &t) : s(text), tree(t) {}
    void skip() { while (pos < s.size() && isspace((unsigned char)s[pos])) pos++; }
    PhyloNode *newNode(const std::string &name) {
        PhyloNode *n = new PhyloNode((int)tree.all_nodes.size(), name);
        tree.all_nodes.push_back(n);
        return n;
    }
    void link(PhyloNode *a, PhyloNode *b, double len) {
        const int id = next_branch++;
        a->neighbors.push_back(new PhyloNeighbor(b, len, id));
        b->neighbors.push_back(new PhyloNeighbor(a, len, id));
    }
    std::string label() {
        skip();
        std::string out;
        if (pos < s.size() && (s[pos] == '\'' || s[pos] == '"')) {
            const char q = s[pos++];
            while (pos < s.size() && s[pos] != q) out.push_back(s[pos++]);
            pos++;
            return out;
        }
        while (pos < s.size() && !strchr("():,;[", s[pos]) && !isspace((unsigned char)s[pos])) out.push_back(s[pos++]);
        return out;
    }
    double length() {
        skip();
        if (pos < s.size() && s[pos] == ':') {
            pos++;
            skip();
            size_t used = 0;
            double v = std::stod(s.substr(pos), &used);
            pos += used;
            return v;
        }
        return -1.0;
    }
    PhyloNode *subtree() {
        skip();
        if (pos >= s.size()) outError("Unexpected end of tree string");
        if (s[pos] == '(') {
            pos++;
            PhyloNode *n = newNode("");
            while (true) {
                PhyloNode *child = subtree();
                double len = length();
                link(n, child, len);
                skip();
                if (pos < s.size() && s[pos] == ',') { pos++; continue; }
                if (pos < s.size() && s[pos] == ')') { pos++; break; }
                outError("Expected , or ) in tree string");
            }
            n->name = label();  // internal label / support value, kept as the name
            return n;
        }
        return newNode(label());
    }
};
}  // namespace

void PhyloTree::readTreeString(const std::string &newick) {
    deleteAllPartialLh();
    freeNodes();
    NewickParser ps(newick, *this);
    PhyloNode *top = ps.subtree();
    ps.length();
    // an unrooted tree is stored with a leaf as `root` (ASSERT(root->isLeaf()), tree/phylotree.cpp:1230);
    // a degree-2 top node (rooted input) is dissolved into one branch
    if (top->degree() == 2) {
        PhyloNode *a = top->neighbors[0]->node, *b = top->neighbors[1]->node;
        const double la = top->neighbors[0]->length, lb = top->neighbors[1]->length;
        PhyloNeighbor *ab = a->findNeighbor(top), *ba = b->findNeighbor(top);
        ab->node = b;
        ba->node = a;
        ab->length = ba->length = (la < 0 ? 0 : la) + (lb < 0 ? 0 : lb);
        ba->id = ab->id;
        for (PhyloNeighbor *nb : top->neighbors) delete nb;
        all_nodes.erase(std::find(all_nodes.begin(), all_nodes.end(), top));
        delete top;
    }
    // leaves first (ids 0..leafNum-1 in order of appearance), internal nodes after; branch ids compacted
    std::vector<PhyloNode *> leaves, internals;
    for (PhyloNode *n : all_nodes) (n->isLeaf() ? leaves : internals).push_back(n);
    all_nodes = leaves;
    all_nodes.insert(all_nodes.end(), internals.begin(), internals.end());
    for (size_t i = 0; i < all_nodes.size(); i++) all_nodes[i]->id = (int)i;
    leafNum = (int)leaves.size();
    nodeNum = (int)all_nodes.size();
    std::map<int, int> bid;
    for (PhyloNode *n : all_nodes)
        for (PhyloNeighbor *nb : n->neighbors) {
            if (!bid.count(nb->id)) { int k = (int)bid.size(); bid[nb->id] = k; }
            if (nb->length < 0) nb->length = 0.1;  // missing length: the reference's default fixed length
        }
    for (PhyloNode *n : all_nodes)
        for (PhyloNeighbor *nb : n->neighbors) nb->id = bid[nb->id];
    branchNum = (int)bid.size();
    if (leafNum < 2) outError("Tree must have at least 2 taxa");
    root = leaves[0];
    if (aln) setAlignment(aln);
}

void PhyloTree::copyTree(const PhyloTree &other) {
    deleteAllPartialLh();
    freeNodes();
    for (const PhyloNode *n : other.all_nodes) all_nodes.push_back(new PhyloNode(n->id, n->name));
    for (const PhyloNode *n : other.all_nodes)
        for (const PhyloNeighbor *nb : n->neighbors)
            all_nodes[n->id]->neighbors.push_back(new PhyloNeighbor(all_nodes[nb->node->id], nb->length, nb->id));
    leafNum = other.leafNum;
    nodeNum = other.nodeNum;
    branchNum = other.branchNum;
    root = all_nodes[other.root->id];
    if (aln) setAlignment(aln);
}

void PhyloTree::setAlignment(Alignment *alignment) {
    aln = alignment;
    leaf_row.clear();
    if (!root) return;
    for (PhyloNode *n : all_nodes)
        if (n->isLeaf()) {
            const int row = aln->getSeqID(n->name);
            if (row < 0) outError("Alignment has no sequence name " + n->name);
            leaf_row[n->id] = row;
        }
    if ((int)aln->getNSeq() != leafNum) outError("Tree and alignment have different numbers of taxa");
}

PhyloNode *PhyloTree::findNodeName(const std::string &name) const {
    for (PhyloNode *n : all_nodes)
        if (n->name == name) return n;
    return NULL;
}

PhyloNode *PhyloTree::findFarthestLeaf(PhyloNode *node, PhyloNode *dad) {
    if (!node) node = root;
    if (dad && node->isLeaf()) {
        node->height = 0.0;
        return node;
    }
    PhyloNode *res = NULL;
    node->height = 0.0;
    for (PhyloNeighbor *nb : node->neighbors) {
        if (nb->node == dad) continue;
        PhyloNode *leaf = findFarthestLeaf(nb->node, node);
        if (node->height < nb->node->height + 1) {
            node->height = nb->node->height + 1;
            res = leaf;
        }
    }
    return res;
}

void PhyloTree::getPreOrderBranches(std::vector<PhyloNode *> &nodes, std::vector<PhyloNode *> &nodes2, PhyloNode *node,
                                    PhyloNode *dad) {
    if (dad) {
        nodes.push_back(node);
        nodes2.push_back(dad);
    }
    // children by increasing height, with the reference's exchange sort so that ties resolve identically
    std::vector<PhyloNeighbor *> nv = node->neighbors;
    for (size_t i = 0; i < nv.size(); i++)
        for (size_t j = i + 1; j < nv.size(); j++)
            if (nv[i]->node->height > nv[j]->node->height) std::swap(nv[i], nv[j]);
    for (PhyloNeighbor *nb : nv)
        if (nb->node != dad) getPreOrderBranches(nodes, nodes2, nb->node, node);
}

void PhyloTree::getBranches(std::vector<PhyloNode *> &nodes, std::vector<PhyloNode *> &nodes2) const {
    nodes.assign(branchNum, NULL);
    nodes2.assign(branchNum, NULL);
    for (PhyloNode *n : all_nodes)
        for (PhyloNeighbor *nb : n->neighbors)
            if (n->id < nb->node->id) { nodes[nb->id] = n; nodes2[nb->id] = nb->node; }
}

static void write_subtree(std::ostringstream &os, const PhyloNode *node, const PhyloNode *dad) {
    if (node->isLeaf()) { os << node->name; return; }
    os << "(";
    bool first = true;
    for (PhyloNeighbor *nb : node->neighbors) {
        if (nb->node == dad) continue;
        if (!first) os << ",";
        first = false;
        write_subtree(os, nb->node, node);
        os << ":" << nb->length;
    }
    os << ")";
}

std::string PhyloTree::getTreeString(int digits) const {
    std::ostringstream os;
    os.precision(digits);
    os << "(" << root->name << ":" << root->neighbors[0]->length;
    PhyloNode *in = root->neighbors[0]->node;
    for (PhyloNeighbor *nb : in->neighbors) {
        if (nb->node == root) continue;
        os << ",";
        write_subtree(os, nb->node, in);
        os << ":" << nb->length;
    }
    os << ");";
    return os.str();
}

void PhyloTree::scaleLength(double factor) {
    for (PhyloNode *n : all_nodes)
        for (PhyloNeighbor *nb : n->neighbors) nb->length *= factor;
}

// with heterotachy the class lengths follow the plain ones: v[branchNum + id*mixlen + c]
void PhyloTree::saveBranchLengths(std::vector<double> &v) const {
    const int m = isMixlen() ? getMixlen() : 0;
    v.assign((size_t)branchNum * (1 + m), 0.0);
    for (PhyloNode *n : all_nodes)
        for (PhyloNeighbor *nb : n->neighbors) {
            v[nb->id] = nb->length;
            for (int c = 0; c < m && c < (int)nb->lengths.size(); c++) v[branchNum + (size_t)nb->id * m + c] = nb->lengths[c];
        }
}

void PhyloTree::restoreBranchLengths(const std::vector<double> &v) {
    const int m = isMixlen() ? getMixlen() : 0;
    for (PhyloNode *n : all_nodes)
        for (PhyloNeighbor *nb : n->neighbors) {
            nb->length = v[nb->id];
            if (m && v.size() >= (size_t)branchNum * (1 + m)) {
                nb->lengths.resize(m);
                for (int c = 0; c < m; c++) nb->lengths[c] = v[branchNum + (size_t)nb->id * m + c];
            }
        }
}

void PhyloTree::clearAllPartialLH() {
    for (PhyloNode *n : all_nodes)
        for (PhyloNeighbor *nb : n->neighbors) nb->partial_lh_computed = 0;
    theta_computed = false;
}

int PhyloTree::computeSizes(PhyloNode *node, PhyloNode *dad) {
    if (node->isLeaf() && dad) return 1;
    int total = 0;
    for (PhyloNeighbor *nb : node->neighbors) {
        if (nb->node == dad) continue;
        nb->size = computeSizes(nb->node, node);
        total += nb->size;
    }
    return total;
}

}  // namespace b200
