// b200host.h -- C++ host side above the C ABI (include/b200lk.h).
//
// The reference is compiled code, so the host layer is C++: it mirrors, by name and argument meaning, the part of
// PhyloTree that reaches the likelihood kernel -- setLikelihoodKernel and the four member-function pointers
// (tree/phylotree.h:895-896, 942-943, 995-996, 1336-1337; tree/phylotreesse.cpp:91-236), the traversal builder
// (tree/phylotree.cpp:5907-6015), computeLikelihood (tree/phylotree.cpp:1227-1268), the Newton-Raphson glue
// (tree/phylotree.cpp:2505-2527, 2585-2635, 2689-2765; utils/optimization.cpp:422-499) and the PhyloSuperTree
// partition dispatch (tree/phylosupertree.cpp:211-231, 701-725; tree/phylosupertreeplen.cpp:254-412) -- so that a
// caller of the reference finds the same entry points, and the parity tests read like the reference's own use.
// It is NOT a rebuild of IQ-TREE: no tree search, no model selection, no I/O beyond PHYLIP/Newick for fixtures.
//
// The only kernel implemented is LK_CUDA (libb200lk.so, sm_100a).  There is no CPU fallback: selecting any other
// kernel, or LK_CUDA without the library / a device, fails loudly.
#pragma once
#include <stddef.h>
#include <stdint.h>

#include <map>
#include <string>
#include <vector>

#include "../../include/b200lk.h"

namespace b200 {

// utils/tools.h:526-528, with LK_CUDA appended (keeps `>=` comparisons such as get_safe_upper_limit meaningful)
enum LikelihoodKernel { LK_386, LK_SSE, LK_SSE2, LK_AVX, LK_AVX_FMA, LK_AVX512, LK_CUDA };
enum SeqType { SEQ_DNA, SEQ_PROTEIN, SEQ_MORPH };
enum SiteLoglType { WSL_NONE, WSL_SITE, WSL_RATECAT, WSL_MIXTURE, WSL_MIXTURE_RATECAT };  // utils/tools.h:534-536
enum ASCType { ASC_NONE, ASC_VARIANT, ASC_VARIANT_MISSING };                               // utils/tools.h:549-555

const double TOL_BRANCH_LEN = 0.000001;  // tree/phylotree.h:64
const double TOL_LIKELIHOOD = 0.001;     // tree/phylotree.h:65

[[noreturn]] void outError(const std::string &msg);  // utils/tools.cpp:70-77: print, then throw (no exit in a library)
void outWarning(const std::string &msg);

// ---- kernel library binding ------------------------------------------------------------------------------
// The C ABI resolved at run time from libb200lk.so (dlopen next to this library).  Only a TEST build of this layer
// (-DB200_TESTING -> libb200host_testing.so, never loaded by bench.py or smoke()) can be pointed at another library
// exporting the same ABI under another prefix (tests/ do this with the CPU oracle to check the host logic where no
// GPU exists); the product library has no such switch.
struct KernelABI {
    void *dl;
    std::string path, prefix;
    int (*create)(b200lk_ctx **, const b200lk_dims *);
    int (*create_sharded)(b200lk_ctx **, const b200lk_dims *, const int32_t *, int32_t);  // optional
    int (*destroy)(b200lk_ctx *);
    int (*upload_tips)(b200lk_ctx *, const uint8_t *, const uint8_t *, const double *);
    int (*set_ptn_invar)(b200lk_ctx *, const double *);
    int (*set_model)(b200lk_ctx *, const double *, const double *, const double *, const double *, const double *,
                     const double *, const double *);
    int (*update_partials)(b200lk_ctx *, const b200lk_node *, int32_t, const b200lk_child *, int32_t);
    int (*branch_lnl)(b200lk_ctx *, b200lk_end, b200lk_end, double, double *);
    int (*build_theta)(b200lk_ctx *, b200lk_end, b200lk_end);
    int (*derv)(b200lk_ctx *, double, double *, double *);
    int (*lnl_from_buffer)(b200lk_ctx *, double, double *);
    int (*evaluate_async)(b200lk_ctx *, const b200lk_node *, int32_t, const b200lk_child *, int32_t, b200lk_end,
                          b200lk_end, double, int, int);                  // optional (NULL when absent)
    int (*read_results)(b200lk_ctx *, double *, double *, double *);      // optional
    int (*fetch_pattern_lh)(b200lk_ctx *, double *);
    int (*fetch_pattern_lh_cat)(b200lk_ctx *, double *);
    int (*fetch_scale_num)(b200lk_ctx *, int32_t, uint8_t *);
    int (*optimize_branch)(b200lk_ctx *, b200lk_end, b200lk_end, const b200lk_newton *, double *, double *,
                           int32_t *);                                    // optional
    // optional: branch lengths resident on the device (a whole sweep without a read-back)
    int (*set_lengths)(b200lk_ctx *, const double *, int32_t);
    int (*get_lengths)(b200lk_ctx *, double *, double *, int32_t);
    int (*update_partials_slots)(b200lk_ctx *, const b200lk_node *, int32_t, const b200lk_child *, int32_t, const int32_t *);
    int (*optimize_branch_slot)(b200lk_ctx *, b200lk_end, b200lk_end, const b200lk_newton *, int32_t);
    // optional: +ASC flavour, robust scoring, pattern weights, heterotachy (NULL when the library has none)
    int (*set_asc)(b200lk_ctx *, int32_t);
    int (*set_robust)(b200lk_ctx *, double, int32_t);
    int (*set_ptn_freq)(b200lk_ctx *, const double *);
    int (*update_partials_mixlen)(b200lk_ctx *, const b200lk_node *, int32_t, const b200lk_child *, int32_t,
                                  const double *);
    int (*branch_lnl_mixlen)(b200lk_ctx *, b200lk_end, b200lk_end, const double *, double *);
    int (*lnl_from_buffer_mixlen)(b200lk_ctx *, const double *, double *);
    int (*derv_mixlen)(b200lk_ctx *, const double *, double *, double *, double *);
    int (*derv_mixlen_class)(b200lk_ctx *, int32_t, double, double *, double *);
    const char *(*last_error)(const b200lk_ctx *);                        // optional
    int (*get_counters)(b200lk_ctx *, int64_t *, int64_t *);              // optional
    // batched launches over the partitions of a supermatrix (optional; NULL when the library has none)
    int (*batch_create)(b200lk_batch **, b200lk_ctx **, int32_t);
    int (*batch_destroy)(b200lk_batch *);
    const char *(*batch_last_error)(const b200lk_batch *);
    int (*batch_evaluate)(b200lk_batch *, const b200lk_node *, int32_t, const b200lk_child *, int32_t, const double *,
                          b200lk_end, b200lk_end, const double *, int, double *, double *, double *);
    int (*batch_evaluate_async)(b200lk_batch *, const b200lk_node *, int32_t, const b200lk_child *, int32_t,
                                const double *, b200lk_end, b200lk_end, const double *, int);
    int (*batch_read_results)(b200lk_batch *, double *, double *, double *);
    int (*batch_derv)(b200lk_batch *, const double *, double *, double *);
    int (*batch_lnl_from_buffer)(b200lk_batch *, const double *, double *);
};
// libb200lk.so beside this library, prefix "b200lk_"
#ifdef B200_TESTING
void setKernelLibrary(const std::string &path, const std::string &prefix);
#endif
const KernelABI &kernelABI();  // loads on first use; outError if the library or a required symbol is missing

// ---- alignment (alignment/alignment.h, alignment/pattern.h) ----------------------------------------------
class Alignment {
public:
    SeqType seq_type = SEQ_DNA;
    int num_states = 4;
    int STATE_UNKNOWN = 18;  // DNA 18, protein 23
    std::vector<std::string> seq_names;
    int64_t nsite = 0;
    std::vector<uint8_t> states;    // [ntaxa][nptn] state codes, taxon-major: aln->at(ptn)[taxon]
    std::vector<double> ptn_freq;   // Pattern::frequency
    std::vector<int> const_char;    // Pattern::const_char: the constant state, STATE_UNKNOWN (all gaps) or > STATE_UNKNOWN
    std::vector<int32_t> site_pattern;

    size_t getNSeq() const { return seq_names.size(); }
    size_t getNPattern() const { return ptn_freq.size(); }
    size_t size() const { return ptn_freq.size(); }
    int getSeqID(const std::string &name) const;
    uint8_t at(size_t ptn, size_t taxon) const { return states[taxon * getNPattern() + ptn]; }

    static int encodeState(SeqType t, char c);  // alignment/alignment.cpp:738 (DNA), model/modelprotein.cpp:1147-1164
    // columns[taxon][site] characters -> compressed site patterns (Alignment::buildPattern)
    void buildFromSequences(const std::vector<std::string> &names, const std::vector<std::string> &seqs, SeqType t,
                            bool compress = true);
    // already-encoded state codes [ntaxa][nsite]; compress merges identical columns
    void buildFromStates(const std::vector<std::string> &names, const uint8_t *codes, int64_t nsite, SeqType t,
                         int num_states, bool compress);
    void readPhylip(const std::string &path, SeqType t);  // sequential or interleaved PHYLIP
private:
    void computeConst();
};

// ---- substitution model: the eigen-system the kernels read (model/modelmarkov.cpp:1386-1632, 1904-1916) ----
class ModelSubst {
public:
    int num_states = 4;
    int nmix = 1;
    std::string name;
    std::vector<double> eval, evec, inv_evec;  // [nmix][n], [nmix][n][n] row = state, [nmix][n][n] row = eigen index
    std::vector<double> state_freq;            // [nmix][n]
    std::vector<double> mixw;                  // [nmix]

    int getNMixtures() const { return nmix; }
    bool isMixture() const { return nmix > 1; }
    double getMixtureWeight(int m) const { return mixw[m]; }
    double *getEigenvalues() { return eval.data(); }
    double *getEigenvectors() { return evec.data(); }
    double *getInverseEigenvectors() { return inv_evec.data(); }
    void getStateFrequency(double *f, int mixture = 0) const;  // -1: weighted average over classes

    // reversible class from symmetric exchangeabilities r[x][y] and frequencies pi: Q_xy = r_xy pi_y, normalised to
    // one expected substitution per unit time; eigen-system by a symmetric solve (modelmarkov.cpp:1555-1559)
    static ModelSubst reversible(int n, const double *exch, const double *freqs, const std::string &name = "REV");
    static ModelSubst GTR(const double rates[6], const double freqs[4]);  // AC AG AT CG CT GT (model/modeldna.cpp)
    static ModelSubst JC();
    static ModelSubst LG();     // model/modelprotein.cpp:287-310 (Le & Gascuel 2008), as data
    static ModelSubst LGC60();  // LG + the C60 profile mixture, model/modelmixture.cpp:917-977, as data
    static ModelSubst mixture(const std::vector<ModelSubst> &classes, const std::vector<double> &weights);
    static ModelSubst fromEigen(int n, int nmix, const double *eval, const double *evec, const double *inv_evec,
                                const double *freqs, const double *mixw);
    // 0/1 indicator of a state code, ambiguity codes included (model/modelsubst.cpp:182-207, modeldna.cpp:462-475,
    // modelprotein.cpp:1147-1164)
    static void computeTipLikelihood(SeqType t, int nstates, int STATE_UNKNOWN, int state, double *state_lk);
};

// symmetric eigen-solve (cyclic Jacobi): a[n][n] -> eigenvalues w[n] ascending, eigenvectors as COLUMNS of v[n][n]
void eigenSymmetric(int n, const double *a, double *w, double *v);

// ---- rates across sites (model/rategamma.cpp:98-170, model/rateinvar.cpp) ---------------------------------
class RateHeterogeneity {
public:
    std::vector<double> rates, props;
    double p_invar = 0.0;
    double gamma_shape = 0.0;
    int getNRate() const { return (int)rates.size(); }
    double getRate(int c) const { return rates[c]; }
    double getProp(int c) const { return props[c]; }
    double getPInvar() const { return p_invar; }
    static RateHeterogeneity uniform();
    // mean-of-category discrete Gamma rescaled to mean 1; with p_invar the category proportions are (1-p_invar)/ncat
    static RateHeterogeneity gamma(double alpha, int ncat, double p_invar = 0.0);
    static RateHeterogeneity fromCategories(const double *rates, const double *props, int ncat, double p_invar);
};
double gammaIncP(double a, double x);      // regularised lower incomplete gamma
double gammaIncPInv(double a, double p);   // its inverse in x

// ---- tree (tree/node.h, tree/phylonode.h) ----------------------------------------------------------------
class PhyloNode;
class PhyloNeighbor {
public:
    PhyloNode *node;        // the node this record points TO
    double length;
    int id;                 // branch id
    int partial_lh_computed = 0;  // bit 0: likelihood partial valid (tree/phylonode.h:99-101)
    int slot = -1;          // device slot holding the partial of the subtree at `node` seen from the owner; -1 none
                            // == (partial_lh - central_partial_lh) / block_size in the reference
    int size = 0;           // leaves in the subtree at `node` (computeSizes)
    // heterotachy (PhyloNeighborMixlen, tree/phylonodemixlen.h:20-140): one length per class; `length` is their mean
    std::vector<double> lengths;
    PhyloNeighbor(PhyloNode *n, double len, int bid) : node(n), length(len), id(bid) {}
    void clearPartialLh() { partial_lh_computed = 0; }
    double getLength(int c) const { return lengths.empty() ? length : lengths[c]; }
    void setLength(int c, double len) {
        if (lengths.empty()) length = len;
        else lengths[c] = len;
    }
};
class PhyloNode {
public:
    int id;
    std::string name;
    std::vector<PhyloNeighbor *> neighbors;
    double height = 0.0;
    PhyloNode(int i, const std::string &nm) : id(i), name(nm) {}
    bool isLeaf() const { return neighbors.size() <= 1; }
    int degree() const { return (int)neighbors.size(); }
    PhyloNeighbor *findNeighbor(const PhyloNode *other) const;
    void clearReversePartialLh(PhyloNode *dad);  // tree/phylonode.cpp: invalidate partials pointing away from dad
    void clearAllPartialLh(PhyloNode *dad);
};

struct TraversalInfo {  // tree/phylotree.h:320-331
    PhyloNeighbor *dad_branch;
    PhyloNode *dad;
};

enum LhMemSave { LM_PER_NODE, LM_MEM_SAVE };  // utils/tools.h: -mem

struct Params {  // the few fields of utils/tools.h Params the path reads
    // -mem: LM_MEM_SAVE recycles max_lh_slots < leafNum-2 partial vectors (tree/memslot.cpp); max_mem_size as the
    // reference reads it for values <= 1: 0 = the minimum log2(leafNum)+1, else that fraction of leafNum-2
    // (tree/phylotree.cpp:993-1015)
    LhMemSave lh_mem_save = LM_PER_NODE;
    double max_mem_size = 0.0;
    double min_branch_length = 1e-6;  // utils/tools.cpp:1209, main/phyloanalysis.cpp:4171-4175
    double max_branch_length = 10.0;  // utils/tools.cpp:1212
    bool lk_safe_scaling = false;     // -safe
    bool asc_lewis = false;           // +ASC: Lewis' correction for variable-sites-only data (ASC_VARIANT)
    bool asc_holder = false;          // +ASC with missing data: Holder's correction (ASC_VARIANT_MISSING)
    double robust_phy_keep = 1.0;     // --robust-phy: fraction of best sites summed (utils/tools.cpp:1294)
    bool robust_median = false;       // --robust-median (utils/tools.cpp:1295)
    int numseq_safe_scaling = 2000;   // utils/tools.cpp:1235-1236
    int device = 0;
    // more than one entry: the alignment is split by site pattern over these GPUs, all driven from this process
    // (b200lk_create_sharded; the device form of the reference's packets, tree/phylokernelnew.h:1114-1142)
    std::vector<int> devices;
    bool device_newton = true;        // run minimizeNewton's iteration on the device when the library offers it
    bool device_sweep = true;         // optimizeAllBranches: keep the lengths on the device for a whole sweep (one read-back)
};

// ---- memory slots of the -mem mode (tree/memslot.h:38-109) ------------------------------------------------
// A slot is a device partial vector + its scale counters (index into the context's arena); with LM_MEM_SAVE the
// traversal builder hands the few slots to the directed branches it is working on, locks the ones a pending node
// still reads, and recycles the unlocked slot with the smallest subtree.
struct MemSlot {
    int status = 0;
    PhyloNeighbor *nei = NULL;
};
class MemSlotVector : public std::vector<MemSlot> {
public:
    bool enabled = false;
    void init(int num_slot);
    void reset();
    bool lock(PhyloNeighbor *nei);
    void unlock(PhyloNeighbor *nei);
    bool locked(PhyloNeighbor *nei);
    int allocate(PhyloNeighbor *nei);  // -1: every slot is locked
    void update(PhyloNeighbor *nei);
    void takeover(PhyloNeighbor *nei, PhyloNeighbor *taken_nei);

private:
    std::map<PhyloNeighbor *, int> nei_id_map;
    int free_count = 0;
    int findNei(PhyloNeighbor *nei);
    void addNei(PhyloNeighbor *nei, int id);
};

class PhyloTree {
public:
    typedef void (PhyloTree::*ComputePartialLikelihoodType)(TraversalInfo &, size_t, size_t, int);
    typedef double (PhyloTree::*ComputeLikelihoodBranchType)(PhyloNeighbor *, PhyloNode *);
    typedef void (PhyloTree::*ComputeLikelihoodDervType)(PhyloNeighbor *, PhyloNode *, double *, double *);
    typedef double (PhyloTree::*ComputeLikelihoodFromBufferType)();
    typedef void (PhyloTree::*ComputeLikelihoodDervMixlenType)(PhyloNeighbor *, PhyloNode *, double &, double &);

    PhyloTree();
    virtual ~PhyloTree();
    PhyloTree(const PhyloTree &) = delete;
    PhyloTree &operator=(const PhyloTree &) = delete;

    // -- set-up, in the reference's order: readTree, setAlignment, setModel/Rate, setLikelihoodKernel,
    //    initializeAllPartialLh
    void readTreeString(const std::string &newick);  // leaf names must match the alignment once it is set
    void copyTree(const PhyloTree &other);            // same node ids, neighbour order, branch ids and lengths
    void setAlignment(Alignment *alignment);          // maps leaf ids to alignment rows
    void setModel(ModelSubst *m) { model = m; model_changed = true; }
    void setRate(RateHeterogeneity *r) { site_rate = r; model_changed = true; }
    void setParams(Params *p) { params = p; }
    virtual void setLikelihoodKernel(LikelihoodKernel lk);  // tree/phylotreesse.cpp:91-199
    virtual void initializeAllPartialLh();                   // tree/phylotree.cpp:890-934, 1054-1182 -> b200lk_create + uploads
    void deleteAllPartialLh();
    // a model/rate parameter changed (ModelFactory::optimizeParameters ... -> clearAllPartialLH): re-upload the
    // eigen-system, tip table and ptn_invar (computeTipPartialLikelihood, computePtnInvar) before the next kernel call
    void modelChanged() { model_changed = true; clearAllPartialLH(); }

    // -- the dispatch API (tree/phylotreesse.cpp:212-236)
    void computePartialLikelihood(TraversalInfo &info, size_t ptn_lower = 0, size_t ptn_upper = 0, int packet_id = 0);
    double computeLikelihoodBranch(PhyloNeighbor *dad_branch, PhyloNode *dad);
    void computeLikelihoodDerv(PhyloNeighbor *dad_branch, PhyloNode *dad, double *df, double *ddf);
    double computeLikelihoodFromBuffer();

    // -- callers of the dispatch API that the parity tests exercise
    virtual double computeLikelihood(double *pattern_lh = NULL);             // tree/phylotree.cpp:1227-1268
    void computePatternLikelihood(double *pattern_lh, double *cur_logl = NULL, double *pattern_lh_cat = NULL);
    // tree/phylotree.cpp:1330-1388: lnL at the root branch; _pattern_lh_cat then holds the per-pattern likelihood of
    // every category, folded per mixture class (WSL_MIXTURE) or per rate category (WSL_RATECAT) for unfused mixtures
    double computePatternLhCat(SiteLoglType wsl);
    double computeFunction(double value);                                    // tree/phylotree.cpp:2505-2518
    virtual void computeFuncDerv(double value, double &df, double &ddf);     // tree/phylotree.cpp:2520-2527
    virtual void optimizeOneBranch(PhyloNode *node1, PhyloNode *node2, bool clearLH = true, int maxNRStep = 100);
    // heterotachy hooks (tree/phylotree.h:546, 556, 1306)
    virtual bool isMixlen() const { return false; }
    virtual int getMixlen() const { return 1; }
    virtual int getCurMixture() const { return 0; }
    virtual double optimizeAllBranches(int my_iterations = 100, double tolerance = TOL_LIKELIHOOD, int maxNRStep = 100);
    virtual void clearAllPartialLH();
    // utils/optimization.cpp:422-499
    double minimizeNewton(double x1, double xguess, double x2, double xacc, double &d2l, int maxNRStep = 100);

    // -- traversal (tree/phylotree.cpp:5907-6015; tree/phylokernelnew.h:1150-1281)
    void computeTraversalInfo(PhyloNode *node, PhyloNode *dad);  // both directions of branch (dad, node)
    std::vector<TraversalInfo> traversal_info;

    // -- topology helpers
    PhyloNode *findNodeName(const std::string &name) const;
    PhyloNode *findFarthestLeaf(PhyloNode *node = NULL, PhyloNode *dad = NULL);  // tree/mtree.cpp:2988-3008
    void getPreOrderBranches(std::vector<PhyloNode *> &nodes, std::vector<PhyloNode *> &nodes2, PhyloNode *node,
                             PhyloNode *dad = NULL);                              // tree/mtree.cpp:3010-3030
    void getBranches(std::vector<PhyloNode *> &nodes, std::vector<PhyloNode *> &nodes2) const;  // by branch id
    std::string getTreeString(int digits = 10) const;
    void scaleLength(double factor);
    void saveBranchLengths(std::vector<double> &v) const;
    void restoreBranchLengths(const std::vector<double> &v);

    // -- state the reference keeps on PhyloTree
    Alignment *aln = NULL;
    ModelSubst *model = NULL;
    RateHeterogeneity *site_rate = NULL;
    Params *params = NULL;
    PhyloNode *root = NULL;
    std::vector<PhyloNode *> all_nodes;
    int leafNum = 0, nodeNum = 0, branchNum = 0;
    LikelihoodKernel sse = LK_386;
    size_t vector_size = 1;      // setLikelihoodKernelCUDA sets 1: host-visible buffers are plain pattern-major
    bool safe_numeric = false;
    int num_threads = 1, num_packets = 1;
    bool optimize_by_newton = true;
    bool theta_computed = false;
    PhyloNeighbor *current_it = NULL, *current_it_back = NULL;
    double curScore = 0.0;
    std::vector<double> _pattern_lh, _pattern_lh_cat;  // filled lazily (computePatternLikelihood)
    int64_t kernel_calls = 0;  // C-ABI calls made (host-side statistic)
    int64_t device_sweeps = 0; // optimizeAllBranches sweeps that ran with device-resident lengths (sweepOnDevice)
    b200lk_ctx *ctx = NULL;    // device residency of this tree
    MemSlotVector mem_slots;   // -mem (tree/phylotree.h: mem_slots)
    int max_lh_slots = 0;      // partial-vector slots handed out (leafNum-2, or fewer with LM_MEM_SAVE)

protected:
    ComputePartialLikelihoodType computePartialLikelihoodPointer = NULL;
    ComputeLikelihoodBranchType computeLikelihoodBranchPointer = NULL;
    ComputeLikelihoodDervType computeLikelihoodDervPointer = NULL;
    ComputeLikelihoodFromBufferType computeLikelihoodFromBufferPointer = NULL;
    ComputeLikelihoodDervMixlenType computeLikelihoodDervMixlenPointer = NULL;  // tree/phylotree.h:1339
    void computeLikelihoodDervMixlenCUDA(PhyloNeighbor *dad_branch, PhyloNode *dad, double &df, double &ddf);
    void ensureTheta(PhyloNeighbor *dad_branch, PhyloNode *dad);
    void branchLengths(const PhyloNeighbor *nb, std::vector<double> &out) const;  // [ncat] class lengths of a branch
    void setLikelihoodKernelCUDA();
    // the LK_CUDA members bound by setLikelihoodKernelCUDA
    void computePartialLikelihoodCUDA(TraversalInfo &info, size_t ptn_lower, size_t ptn_upper, int packet_id);
    double computeLikelihoodBranchCUDA(PhyloNeighbor *dad_branch, PhyloNode *dad);
    void computeLikelihoodDervCUDA(PhyloNeighbor *dad_branch, PhyloNode *dad, double *df, double *ddf);
    double computeLikelihoodFromBufferCUDA();

    bool computeTraversalInfo(PhyloNeighbor *dad_branch, PhyloNode *dad);
    void reorientPartialLh(PhyloNeighbor *dad_branch, PhyloNode *dad);  // tree/phylotree.cpp:5887-5901
    void packTraversal(std::vector<b200lk_node> &nodes, std::vector<b200lk_child> &children, std::vector<int32_t> *slots = NULL);
    // one sweep of optimizeAllBranches with the lengths resident on the device: 1 done; 0 not applicable (nothing touched);
    // -1 given up midway (a per-branch check of optimizeOneBranch fired) -- the caller restores the lengths, invalidates
    // and runs the sweep branch by branch
    int sweepOnDevice(const std::vector<PhyloNode *> &nodes, const std::vector<PhyloNode *> &nodes2, int maxNRStep);

    void flushTraversal();  // traversal_info -> b200lk_update_partials
    void syncModel();
    void check(int rc, const char *what);
    void branchEnds(PhyloNeighbor *dad_branch, PhyloNode *dad, b200lk_end &dad_end, b200lk_end &node_end);
    int computeSizes(PhyloNode *node, PhyloNode *dad);
    void freeNodes();
    bool model_changed = true;
    bool buffers_ready = false;
    std::map<int, int> leaf_row;  // leaf node id -> alignment row
    friend class PhyloSuperTree;
};

// ---- heterotachy (tree/phylotreemixlen.h, tree/phylotreemixlen.cpp) ---------------------------------------
// A tree whose branches carry one length per rate category of a RateHeterotachy-like rate model (all rates 1, class
// weights as proportions): the dispatch members send per-class lengths through the *_mixlen entry points of the C ABI,
// computeLikelihoodDerv returns the gradient / Hessian / lnL form, and optimizeOneBranch runs the reference's EM step.
class PhyloTreeMixlen : public PhyloTree {
public:
    int mixlen = 1;
    int cur_mixture = -1;
    void setMixlen(int m) { mixlen = m; }
    bool isMixlen() const override { return true; }
    int getMixlen() const override { return mixlen; }
    int getCurMixture() const override { return cur_mixture; }
    // class c of every branch = its current length x mult[c] (the relative class rates initializeMixBranches
    // starts from, tree/phylotreemixlen.cpp:137-190)
    void initializeMixBranches(const std::vector<double> &mult);
    void computeFuncDerv(double value, double &df, double &ddf) override;         // :728-737
    void computeFuncDervMulti(const double *value, double *df, double *ddf);       // :544-559
    void optimizeOneBranch(PhyloNode *node1, PhyloNode *node2, bool clearLH = true, int maxNRStep = 100) override;  // EM, :425-495
};

// ---- partitions (tree/phylosupertree.cpp, tree/phylosupertreeplen.cpp) -----------------------------------
// Edge-linked partition model: every partition tree has the super-tree's topology (all taxa present in every gene)
// and branch lengths = super-tree lengths x part_rate[p] (-p, PhyloSuperTreePlen); or edge-unlinked (-Q):
// every partition keeps its own lengths (PhyloSuperTreeUnlinked).
struct PartitionInfo {  // tree/phylosupertree.h part_info
    std::string name;
    double part_rate = 1.0;
    double cur_score = 0.0;
};
class PhyloSuperTree : public PhyloTree {
public:
    std::vector<PhyloTree *> parts;  // owned
    std::vector<PartitionInfo> part_info;
    std::vector<int> part_order;     // by decreasing cost (computePartitionOrder, tree/phylosupertree.cpp:654-700)
    bool edge_linked = true;
    int num_threads_partitions = 0;  // host threads that enqueue partitions (0: all cores, at most 16)
    // Partitions with equal dimensions form a batch: every step of an evaluation is ONE launch over all of them
    // (b200lk_batch_*, include/b200lk.h) instead of one launch per partition.  Off: per-partition streams only.
    bool batch_partitions = true;
    struct PartitionBatch {
        b200lk_batch *handle = NULL;
        std::vector<int> members;  // partition ids
    };
    std::vector<PartitionBatch> batches;
    std::vector<int> unbatched;     // partitions evaluated on their own streams
    int64_t batch_calls = 0;
    ~PhyloSuperTree();
    // partition p: its alignment, model and rates (ownership stays with the caller); the topology is the super-tree's
    void addPartition(const std::string &name, Alignment *a, ModelSubst *m, RateHeterogeneity *r, double part_rate = 1.0);
    void computePartitionOrder();
    void setLikelihoodKernel(LikelihoodKernel lk) override;  // tree/phylosupertree.cpp:211-215: propagate
    void initializeAllPartialLh() override;
    void mapBranchLengths();  // super-tree lengths x part_rate -> partition trees (edge-linked)
    void clearAllPartialLH() override;
    double computeLikelihood(double *pattern_lh = NULL) override;  // tree/phylosupertree.cpp:701-725: sum over partitions
    // edge-linked Newton-Raphson: df = sum_p rate_p df_p, ddf = sum_p rate_p^2 ddf_p (phylosupertreeplen.cpp:400-402)
    void computeFuncDervLinked(PhyloNode *node1, PhyloNode *node2, double value, double &df, double &ddf);
    double optimizeAllBranches(int my_iterations = 100, double tolerance = TOL_LIKELIHOOD, int maxNRStep = 100) override;

private:
    void buildBatches();
    void destroyBatches();
    // enqueue the evaluation of one batch's members at their current_it branch; false: the members' traversals
    // differ (or the library refused) and nothing was enqueued -- the caller takes the per-partition path for them
    bool enqueueBatch(PartitionBatch &b, bool want_derv);
    void readBatch(PartitionBatch &b, std::vector<double> &lnl, std::vector<double> &df, std::vector<double> &ddf);
    void evaluatePartitionsAsync(const std::vector<int> &which);
};

// ---- multi-GPU sharding (shard.cpp) ------------------------------------------------
// contiguous pattern range [begin, end) of `rank`, whole 32-pattern tiles
void shardPatterns(int64_t nptn, int world, int rank, int tile, int64_t *begin, int64_t *end);
// longest-processing-time-first placement of partitions on GPUs by cost (tree/phylosupertree.cpp:664-667)
void placePartitions(const double *cost, int nparts, int world, int32_t *gpu_of_part);

}  // namespace b200
