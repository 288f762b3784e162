// ref_harness.cpp -- drives the REFERENCE's own likelihood kernels (TEST INFRASTRUCTURE ONLY).
//
// oracle/Makefile.ref compiles, unmodified and in place from /root/reference, the translation units
//   tree/phylokernelsse.cpp, tree/phylotreeavx.cpp, tree/phylokernelfma.cpp     (the SIMD kernel instantiations
//                                                                               of tree/phylokernelnew.h)
//   tree/phylotreesse.cpp (setLikelihoodKernel, computeTipPartialLikelihood, computePtnInvar/Freq, wrappers)
//   tree/phylotree.cpp (buffers, computeTraversalInfo, optimizeOneBranch ...), tree/mtree.cpp, tree/node.cpp,
//   tree/phylonode.cpp, tree/memslot.cpp, utils/optimization.cpp, utils/tools.cpp, utils/checkpoint.cpp,
//   utils/MPIHelper.cpp, utils/progress.cpp, model/modelsubst.cpp, model/rateheterogeneity.cpp,
//   alignment/pattern.cpp
// and links them with THIS file into oracle/_ref/libiqtree_ref.so.  Nothing of the reference is copied.
//
// What this file supplies is the scaffolding the kernels read their inputs through, because the reference's
// own producers cannot be built here (alignment/alignment.cpp and model/modelmarkov.cpp need Eigen3/Boost,
// which are not installed): minimal bodies for the few Alignment / ModelFactory / StateSpace members the
// path touches, and ModelSubst / RateHeterogeneity subclasses that hand the kernels a caller-supplied
// eigen-system and rate categories through the reference's virtual interface
// (model/modelsubst.h:360-372, model/rateheterogeneity.h:107-172).  Every other unresolved symbol of the
// reference objects is an auto-generated trap (oracle/gen_ref_stubs.py): reaching one aborts loudly.
#include <sstream>
#include <string>
#include <vector>
#include <algorithm>
#include <cstring>
#include <complex>

// the harness reads kernel-private buffers (_pattern_lh, partial_lh, scale_num, current_it): open the
// access specifiers for THIS translation unit only (object layout is unaffected)
#define private public
#define protected public
#include "tree/phylotree.h"
#include "tree/phylonode.h"
#include "tree/phylonodemixlen.h"
#include "model/modelsubst.h"
#include "model/rateheterogeneity.h"
#include "model/modelfactory.h"
#include "alignment/alignment.h"
#include "utils/tools.h"
#include "utils/eigendecomposition.h"
#include "utils/starttree.h"
#undef private
#undef protected

int NNI_MAX_NR_STEP = 10;  // tree/phylotree.cpp global defined in main/phyloanalysis.cpp (off the path)

// parseArg() asks for this default (utils/tools.cpp:1362); the distance-tree builders are off the path
const std::string &StartTree::Factory::getNameOfDefaultTreeBuilder() {
    static const std::string name = "none";
    return name;
}

// ---- minimal bodies for members whose translation units cannot be built here ------------------------------
PML::StateSpace::StateSpace() : space_name(), num_states(0), num_all_states(0) {}
PML::StateSpace::~StateSpace() {}

Alignment::Alignment() : vector<Pattern>() {
    num_states = 0;
    frac_const_sites = 0.0;
    frac_invariant_sites = 0.0;
    codon_table = NULL;
    genetic_code = NULL;
    non_stop_codon = NULL;
    seq_type = SEQ_UNKNOWN;
    STATE_UNKNOWN = 126;
    pars_lower_bound = NULL;
}
Alignment::~Alignment() {}
string &Alignment::getSeqName(int i) { return seq_names[i]; }
vector<string> &Alignment::getSeqNames() { return seq_names; }
int Alignment::getSeqID(string &seq_name) {
    for (size_t i = 0; i < seq_names.size(); i++)
        if (seq_names[i] == seq_name) return (int)i;
    return -1;
}

ModelFactory::ModelFactory() : CheckpointFactory() {
    model = NULL;
    site_rate = NULL;
    store_trans_matrix = false;
    is_storing = false;
    joint_optimize = false;
    fused_mix_rate = false;
    ASC_type = ASC_NONE;
}
ModelFactory::~ModelFactory() {}
void ModelFactory::setCheckpoint(Checkpoint *checkpoint) { CheckpointFactory::setCheckpoint(checkpoint); }
// model/modelfactory.cpp:1517-1556 with store_trans_matrix off (the default): straight to the model
void ModelFactory::computeTransMatrix(double time, double *trans_matrix, int mixture, int selected_row) {
    model->computeTransMatrix(time, trans_matrix, mixture, selected_row);
}
void ModelFactory::computeTransDerv(double time, double *trans_matrix, double *trans_derv1, double *trans_derv2,
                                    int mixture) {
    model->computeTransDerv(time, trans_matrix, trans_derv1, trans_derv2, mixture);
}

// ---- caller-supplied model and rates behind the reference's virtual interface -----------------------------
class HarnessModel : public ModelSubst {
public:
    int nmix;
    // the kernels read the eigen-system with aligned SIMD loads (tree/phylokernelnew.h:1019-1026): 64-byte aligned
    double *eval, *evec, *inv_evec;
    std::vector<double> mixw, freqs, tipvec;  // tipvec[code][x]
    int ncodes;
    static double *aligned_copy(const double *src, size_t count) {
        void *p = NULL;
        if (posix_memalign(&p, 64, sizeof(double) * (count + 8))) return NULL;
        memcpy(p, src, sizeof(double) * count);
        return (double *)p;
    }
    HarnessModel(int n, int nmix_) : ModelSubst(n), nmix(nmix_), eval(NULL), evec(NULL), inv_evec(NULL), ncodes(0), site(false), nonrev(false) {}
    virtual ~HarnessModel() { free(eval); free(evec); free(inv_evec); }
    bool site;  // ModelSet-like: per-pattern eigen-systems in the layout of ModelSet::decomposeRateMatrix
    virtual bool isSiteSpecificModel() { return site; }
    // Non-reversible kernel family: what ModelMarkov hands the kernels through computeTransMatrix / computeTransDerv
    // (model/modelmarkov.cpp:460-488, 729-745; that translation unit needs Eigen3 and cannot be built here): the
    // caller supplies Q and its complex eigen-system, P(t) = Re(V exp(ev t) V^-1), dP = Q P, d2P = Q Q P.
    // `nonrev` makes isReversible() false; with params->kernel_nonrev a reversible model takes the same route.
    bool nonrev;
    std::vector<double> Q;                       // [nmix][n][n]
    std::vector<std::complex<double>> cev, cV, cVi;  // [nmix][n], [nmix][n][n] x2
    virtual bool isReversible() { return !nonrev; }
    virtual void computeTransMatrix(double time, double *trans_matrix, int mixture = 0, int selected_row = -1) {
        const int n = num_states;
        const std::complex<double> *ev = &cev[(size_t)mixture * n], *V = &cV[(size_t)mixture * n * n];
        const std::complex<double> *Vi = &cVi[(size_t)mixture * n * n];
        std::vector<std::complex<double>> e(n);
        for (int k = 0; k < n; k++) e[k] = std::exp(ev[k] * time);
        for (int i = 0; i < n; i++)
            for (int j = 0; j < n; j++) {
                std::complex<double> s(0.0, 0.0);
                for (int k = 0; k < n; k++) s += V[(size_t)i * n + k] * e[k] * Vi[(size_t)k * n + j];
                trans_matrix[(size_t)i * n + j] = s.real();
            }
    }
    virtual void computeTransDerv(double time, double *trans_matrix, double *trans_derv1, double *trans_derv2,
                                  int mixture = 0) {
        const int n = num_states;
        const double *q = &Q[(size_t)mixture * n * n];
        computeTransMatrix(time, trans_matrix, mixture);
        for (int pass = 0; pass < 2; pass++) {
            const double *src = pass ? trans_derv1 : trans_matrix;
            double *dst = pass ? trans_derv2 : trans_derv1;
            for (int i = 0; i < n; i++)
                for (int j = 0; j < n; j++) {
                    double v = 0.0;
                    for (int k = 0; k < n; k++) v += q[(size_t)i * n + k] * src[(size_t)k * n + j];
                    dst[(size_t)i * n + j] = v;
                }
        }
    }
    virtual bool isMixture() { return nmix > 1; }
    virtual int getNMixtures() { return nmix; }
    virtual double getMixtureWeight(int cat) { return mixw[cat]; }
    virtual double *getEigenvalues() const { return eval; }
    virtual double *getEigenvectors() const { return evec; }
    virtual double *getInverseEigenvectors() const { return inv_evec; }
    virtual void getStateFrequency(double *f, int mixture = 0) {
        for (int i = 0; i < num_states; i++) {
            if (mixture >= 0) {
                f[i] = freqs[(size_t)mixture * num_states + i];
            } else {  // -1: the weighted average over classes (computePtnInvar, tree/phylotreesse.cpp:570)
                f[i] = 0.0;
                for (int m = 0; m < nmix; m++) f[i] += mixw[m] * freqs[(size_t)m * num_states + i];
            }
        }
    }
    // the 0/1 indicator of a state code, ambiguity codes included (input data)
    virtual void computeTipLikelihood(PML::StateType state, double *state_lk) {
        for (int i = 0; i < num_states; i++) state_lk[i] = tipvec[(size_t)state * num_states + i];
    }
};

class HarnessRate : public RateHeterogeneity {
public:
    std::vector<double> rates, props;
    double pinvar;
    HarnessRate() : RateHeterogeneity(), pinvar(0.0) {}
    virtual int getNRate() { return (int)rates.size(); }
    virtual int getNDiscreteRate() { return (int)rates.size(); }
    virtual double getRate(int c) { return rates[c]; }
    virtual double getProp(int c) { return props[c]; }
    virtual double getPInvar() { return pinvar; }
    virtual int isGammaRate() { return 0; }
};

// Heterotachy (PhyloTreeMixlen, tree/phylotreemixlen.h:89-98, 218; that class derives from IQTree and cannot be built
// here): the three virtuals the kernels consult, and nodes whose neighbours carry one length per class
// (PhyloNeighborMixlen::getLength(c), tree/phylonodemixlen.h:64-69)
class HarnessMixlenTree : public PhyloTree {
public:
    int mixlen, cur_mixture;
    explicit HarnessMixlenTree(int m) : PhyloTree(), mixlen(m), cur_mixture(-1) {}
    virtual Node *newNode(int node_id = -1, const char *node_name = NULL) { return new PhyloNodeMixlen(node_id, node_name); }
    virtual Node *newNode(int node_id, int node_name) { return new PhyloNodeMixlen(node_id, node_name); }
    virtual bool isMixlen() { return true; }
    virtual int getMixlen() { return mixlen; }
    virtual int getCurMixture() { return cur_mixture; }
};

struct RefHandle {
    PhyloTree *tree;
    Alignment *aln;
    HarnessModel *model;
    HarnessRate *rate;
    ModelFactory *factory;
    PhyloNeighbor *cur_dad_branch;
    PhyloNode *cur_dad;
    int ncm, block;
};

extern "C" {

struct ref_problem {
    int32_t ntaxa, nstates, ncat, nmix, n_state_codes, kernel, safe, threads;
    int64_t nptn;
    const uint8_t *states;      // [ntaxa][nptn]
    const double *ptn_freq;     // [nptn] (integers)
    const char *newick;         // leaves named T<taxon>
    const double *tipvec;       // [n_state_codes][nstates]
    const double *rates, *props;
    double pinvar;
    const double *eval, *evec, *inv_evec, *mixw, *freqs;
};

static bool params_ready = false;

// +ASC (Lewis): unobserved patterns of the NEXT ref_create, [ntaxa][n] state codes (model_factory->unobserved_ptns,
// model/modelfactory.h:241; read by the kernels at tree/phylokernelnew.h:1500-1512, 2912-2930, 3137-3183)
static std::vector<uint8_t> next_unobserved;
static int next_n_unobserved = 0;
void ref_set_next_unobserved(int n_unobserved, const uint8_t *codes, int ntaxa) {
    next_n_unobserved = n_unobserved;
    next_unobserved.assign(codes, codes + (size_t)(n_unobserved > 0 ? n_unobserved : 0) * ntaxa);
}

// the reference's own reversible eigen-solver (utils/eigendecomposition.cpp:165-317) on caller-supplied
// exchangeabilities (upper triangle) and frequencies, as ModelMarkov::decomposeRateMatrixRev calls it
// (model/modelmarkov.cpp:1602-1632)
void ref_eigensystem_sym(int n, const double *exch, const double *freqs, double total_num_subst, double *eval,
                         double *evec, double *inv_evec) {
    EigenDecomposition ed;
    ed.total_num_subst = total_num_subst;
    std::vector<std::vector<double>> rm(n, std::vector<double>(n, 0.0));
    std::vector<double *> rows(n);
    std::vector<double> fr(freqs, freqs + n);
    for (int i = 0, k = 0; i < n; i++) {
        rows[i] = rm[i].data();
        for (int j = i + 1; j < n; j++, k++) {
            const double r = (freqs[i] <= ZERO_FREQ || freqs[j] <= ZERO_FREQ) ? 0 : exch[k];
            rm[i][j] = rm[j][i] = r;
        }
    }
    ed.eigensystem_sym(rows.data(), fr.data(), eval, evec, inv_evec, n);
}

// Site-specific models of the NEXT ref_create (what a ModelSet holds after decomposeRateMatrix): eval [nptn][n],
// evec / inv_evec [nptn][n][n], pattern-major; ref_create re-arranges them to the kernel's vector width exactly like
// ModelSet::decomposeRateMatrix (model/modelset.cpp:139-200): blocks of V patterns, element-major inside a block,
// the last model repeated into the padding
static std::vector<double> next_site_eval, next_site_evec, next_site_inv;
void ref_set_next_site_models(int64_t nptn, int n, const double *eval, const double *evec, const double *inv_evec) {
    next_site_eval.assign(eval, eval + (size_t)nptn * n);
    next_site_evec.assign(evec, evec + (size_t)nptn * n * n);
    next_site_inv.assign(inv_evec, inv_evec + (size_t)nptn * n * n);
}
static double *interleave_models(const std::vector<double> &src, size_t nptn, size_t per, size_t V) {
    const size_t padded = (nptn + V - 1) / V * V;
    void *p = NULL;
    if (posix_memalign(&p, 64, sizeof(double) * (padded * per + 8))) return NULL;
    double *out = (double *)p;
    for (size_t q = 0; q < padded; q++) {
        const double *m = &src[std::min(q, nptn - 1) * per];
        for (size_t x = 0; x < per; x++) out[(q / V) * V * per + x * V + q % V] = m[x];
    }
    return out;
}

// model_factory->fused_mix_rate of the NEXT ref_create (model/modelfactory.h; category c = mixture class c with rate
// category c, tree/phylokernelnew.h:1322-1331)
// Heterotachy for the NEXT ref_create: class c of every branch has length  newick length * mult[c]
static std::vector<double> next_mixlen_mult;
void ref_set_next_mixlen(int nclass, const double *mult) { next_mixlen_mult.assign(mult, mult + nclass); }
// +ASC flavour of the NEXT ref_create's unobserved patterns: 0 Lewis (ASC_VARIANT), 1 Holder (ASC_VARIANT_MISSING)
static int next_asc_holder = 0;
void ref_set_next_asc_holder(int on) { next_asc_holder = on; }
static void set_mix_lengths(Node *node, Node *dad, const std::vector<double> &mult) {
    for (size_t i = 0; i < node->neighbors.size(); i++) {
        PhyloNeighborMixlen *nb = (PhyloNeighborMixlen *)node->neighbors[i];
        nb->lengths.resize(mult.size());
        for (size_t c = 0; c < mult.size(); c++) nb->lengths[c] = nb->length * mult[c];
        if (nb->node != dad) set_mix_lengths(nb->node, node, mult);
    }
}

// robust phylogeny (tree/phylokernelnew.h:3096-3135) for the NEXT ref_create: params->robust_phy_keep / robust_median.
// Params is the reference's process-wide singleton, so every ref_create sets both (default: off).
static double next_robust_keep = 1.0;
static int next_robust_median = 0;
void ref_set_next_robust(double keep, int median) { next_robust_keep = keep; next_robust_median = median; }

static int next_fused = 0;
void ref_set_next_fused(int fused) { next_fused = fused; }

// Non-reversible kernel family for the NEXT ref_create.  mode 1: the model reports isReversible() == false and the
// newick is read as a ROOTED tree (tree/mtree.cpp:739-751 adds the root leaf, tree/phylotree.cpp:5710 sets the branch
// directions); mode 2: params->kernel_nonrev on an unrooted tree with a reversible model (what -asr and the marginal
// ancestral state code switch on, tree/phylotreesse.cpp:1441-1446).  Q [nmix][n][n]; ceval [nmix][n], cevec / cinv
// [nmix][n][n] as (re, im) pairs.
static int next_nonrev_mode = 0;
static std::vector<double> next_Q;
static std::vector<std::complex<double>> next_cev, next_cV, next_cVi;
void ref_set_next_nonrev(int mode, int n, int nmix, const double *Q, const double *ceval, const double *cevec,
                         const double *cinv) {
    next_nonrev_mode = mode;
    next_Q.assign(Q, Q + (size_t)nmix * n * n);
    const std::complex<double> *e = (const std::complex<double> *)ceval, *v = (const std::complex<double> *)cevec;
    const std::complex<double> *w = (const std::complex<double> *)cinv;
    next_cev.assign(e, e + (size_t)nmix * n);
    next_cV.assign(v, v + (size_t)nmix * n * n);
    next_cVi.assign(w, w + (size_t)nmix * n * n);
}

#include <execinfo.h>
#include <signal.h>
#include <unistd.h>
static void segv_backtrace(int sig) {  // debugging aid: REF_HARNESS_BACKTRACE=1
    void *frames[64];
    int n = backtrace(frames, 64);
    backtrace_symbols_fd(frames, n, 2);
    _exit(139);
}

void *ref_create(const ref_problem *pb) {
    if (getenv("REF_HARNESS_BACKTRACE")) signal(SIGSEGV, segv_backtrace);
    if (!params_ready) {
        const char *argv[] = {"iqtree2", "-s", "harness", NULL};
        parseArg(3, const_cast<char **>(argv), Params::getInstance());  // the reference's own defaults
        params_ready = true;
        // main/phyloanalysis.cpp:4171-4175 (off the path, not compiled here) turns the parseArg default 0.0 into the
        // working lower bound before any branch is optimised
        if (Params::getInstance().min_branch_length <= 0.0) Params::getInstance().min_branch_length = 1e-6;
    }
    Params &params = Params::getInstance();
    params.lk_safe_scaling = pb->safe != 0;
    params.robust_phy_keep = next_robust_keep;
    params.robust_median = next_robust_median != 0;
    next_robust_keep = 1.0;
    next_robust_median = 0;
    RefHandle *h = new RefHandle();
    const int n = pb->nstates;
    // alignment: pattern-major vector<Pattern>, aln->at(ptn)[taxon] (alignment/pattern.h:29)
    Alignment *aln = new Alignment();
    aln->num_states = n;
    aln->STATE_UNKNOWN = pb->n_state_codes - 1;
    aln->seq_type = n == 4 ? SEQ_DNA : (n == 20 ? SEQ_PROTEIN : SEQ_MORPH);
    for (int t = 0; t < pb->ntaxa; t++) aln->getSeqNames().push_back("T" + std::to_string(t));
    aln->reserve(pb->nptn);
    int site = 0;
    for (int64_t p = 0; p < pb->nptn; p++) {
        Pattern pat;
        pat.resize(pb->ntaxa);
        for (int t = 0; t < pb->ntaxa; t++) pat[t] = pb->states[(size_t)t * pb->nptn + p];
        pat.frequency = (int)pb->ptn_freq[p];
        // constant-pattern bookkeeping of Alignment::computeConst for unambiguous data: a pattern is
        // constant when all non-gap characters agree (gaps alone count as constant UNKNOWN)
        int first = -1;
        bool is_const = true, ambiguous = false;
        for (int t = 0; t < pb->ntaxa; t++) {
            int s = pat[t];
            if (s == (int)aln->STATE_UNKNOWN) continue;
            if (s >= n) ambiguous = true;
            if (first < 0) first = s;
            else if (s != first) is_const = false;
        }
        pat.flag = 0;
        pat.const_char = aln->STATE_UNKNOWN + 1;
        if (is_const && !ambiguous) {
            pat.flag |= PAT_CONST;
            pat.const_char = first < 0 ? (int)aln->STATE_UNKNOWN : first;
        }
        aln->push_back(pat);
        for (int f = 0; f < pat.frequency; f++) aln->site_pattern.push_back((int)p), site++;
    }
    h->aln = aln;
    // model, rates, factory
    HarnessModel *m = new HarnessModel(n, pb->nmix);
    m->eval = HarnessModel::aligned_copy(pb->eval, (size_t)pb->nmix * n);
    m->evec = HarnessModel::aligned_copy(pb->evec, (size_t)pb->nmix * n * n);
    m->inv_evec = HarnessModel::aligned_copy(pb->inv_evec, (size_t)pb->nmix * n * n);
    m->mixw.assign(pb->mixw, pb->mixw + pb->nmix);
    m->freqs.assign(pb->freqs, pb->freqs + (size_t)pb->nmix * n);
    m->ncodes = pb->n_state_codes;
    m->tipvec.assign(pb->tipvec, pb->tipvec + (size_t)pb->n_state_codes * n);
    const int nonrev_mode = next_nonrev_mode;
    next_nonrev_mode = 0;
    params.kernel_nonrev = nonrev_mode == 2;
    if (nonrev_mode) {
        m->nonrev = nonrev_mode == 1;
        m->Q.swap(next_Q);
        m->cev.swap(next_cev);
        m->cV.swap(next_cV);
        m->cVi.swap(next_cVi);
    }
    HarnessRate *r = new HarnessRate();
    r->rates.assign(pb->rates, pb->rates + pb->ncat);
    r->props.assign(pb->props, pb->props + pb->ncat);
    r->pinvar = pb->pinvar;
    ModelFactory *f = new ModelFactory();
    f->model = m;
    f->site_rate = r;
    f->fused_mix_rate = next_fused != 0;
    next_fused = 0;
    if (next_n_unobserved > 0) {
        f->ASC_type = next_asc_holder ? ASC_VARIANT_MISSING : ASC_VARIANT;  // Holder / Lewis (utils/tools.h:549-555)
        next_asc_holder = 0;
        for (int u = 0; u < next_n_unobserved; u++) {
            Pattern pat;
            pat.resize(pb->ntaxa);
            for (int t = 0; t < pb->ntaxa; t++) pat[t] = next_unobserved[(size_t)t * next_n_unobserved + u];
            pat.frequency = 0;
            f->unobserved_ptns.push_back(pat);
        }
        next_n_unobserved = 0;
    }
    h->model = m;
    h->rate = r;
    h->factory = f;
    // tree
    std::vector<double> mixlen_mult;
    mixlen_mult.swap(next_mixlen_mult);
    PhyloTree *tree = mixlen_mult.empty() ? new PhyloTree() : new HarnessMixlenTree((int)mixlen_mult.size());
    tree->params = &params;
    std::stringstream ss;
    ss << pb->newick;
    bool rooted = nonrev_mode == 1;
    tree->readTree(ss, rooted);
    if (!mixlen_mult.empty()) set_mix_lengths(tree->root, NULL, mixlen_mult);
    tree->setAlignment(aln);
    tree->setModelFactory(f);
    m->site = !next_site_eval.empty();
    tree->setLikelihoodKernel((LikelihoodKernel)pb->kernel);
    if (m->site) {
        const size_t V = tree->vector_size;
        free(m->eval); free(m->evec); free(m->inv_evec);
        m->eval = interleave_models(next_site_eval, pb->nptn, n, V);
        m->evec = interleave_models(next_site_evec, pb->nptn, (size_t)n * n, V);
        m->inv_evec = interleave_models(next_site_inv, pb->nptn, (size_t)n * n, V);
        next_site_eval.clear(); next_site_evec.clear(); next_site_inv.clear();
    }
    tree->setNumThreads(pb->threads);
    tree->initializeAllPartialLh();
    h->tree = tree;
    h->ncm = f->fused_mix_rate ? pb->ncat : pb->ncat * pb->nmix;  // tree/phylokernelnew.h:1323
    h->block = h->ncm * n;
    h->cur_dad_branch = NULL;
    h->cur_dad = NULL;
    return h;
}

int ref_vector_size(void *hv) { return (int)((RefHandle *)hv)->tree->vector_size; }
int ref_safe_numeric(void *hv) { return ((RefHandle *)hv)->tree->safe_numeric ? 1 : 0; }

static PhyloNode *find_leaf(PhyloTree *tree, int taxon) {
    NodeVector leaves;
    tree->getTaxa(leaves);
    for (size_t i = 0; i < leaves.size(); i++)
        if (leaves[i]->id == taxon) return (PhyloNode *)leaves[i];
    return NULL;
}

// All partials stale -> computeLikelihoodBranch at the branch (leaf, its neighbour), the root choice of
// PhyloTree::computeLikelihood (tree/phylotree.cpp:1227-1268) made explicit.
double ref_lnl_at_leaf(void *hv, int taxon) {
    RefHandle *h = (RefHandle *)hv;
    PhyloTree *tree = h->tree;
    PhyloNode *leaf = find_leaf(tree, taxon);
    if (!leaf) return 1.0;
    tree->clearAllPartialLH();
    PhyloNeighbor *dad_branch = (PhyloNeighbor *)leaf->neighbors[0];
    h->cur_dad_branch = dad_branch;
    h->cur_dad = leaf;
    tree->current_it = dad_branch;
    tree->current_it_back = (PhyloNeighbor *)dad_branch->node->findNeighbor(leaf);
    return tree->computeLikelihoodBranch(dad_branch, leaf);
}

// lnL at the same branch again WITHOUT invalidating partials (after a length change)
double ref_lnl_again(void *hv) {
    RefHandle *h = (RefHandle *)hv;
    return h->tree->computeLikelihoodBranch(h->cur_dad_branch, h->cur_dad);
}

void ref_set_current_length(void *hv, double len) {
    RefHandle *h = (RefHandle *)hv;
    h->tree->current_it->length = len;
    h->tree->current_it_back->length = len;
}

void ref_derv(void *hv, double *df, double *ddf) {
    RefHandle *h = (RefHandle *)hv;
    h->tree->theta_computed = false;
    h->tree->computeLikelihoodDerv(h->cur_dad_branch, h->cur_dad, df, ddf);
}
// further derivative evaluations reuse theta (what minimizeNewton's later steps do)
void ref_derv_again(void *hv, double *df, double *ddf) {
    RefHandle *h = (RefHandle *)hv;
    h->tree->computeLikelihoodDerv(h->cur_dad_branch, h->cur_dad, df, ddf);
}
// ---- heterotachy ----------------------------------------------------------------------------------------
// class lengths of the current branch (both directions)
void ref_set_current_mixlen(void *hv, const double *len, int nclass) {
    RefHandle *h = (RefHandle *)hv;
    DoubleVector v(len, len + nclass);
    h->tree->current_it->setLength(v);
    h->tree->current_it_back->setLength(v);
}
// computeLikelihoodDerv with isMixlen(): df[nclass], ddf[nclass][nclass] and lnL (df[nclass], tree/phylokernelnew.h:2566)
void ref_derv_mixlen(void *hv, double *df, double *ddf, double *lnl, int again) {
    RefHandle *h = (RefHandle *)hv;
    const int m = h->tree->getMixlen();
    std::vector<double> g(m + 1), hs((size_t)m * m);
    if (!again) h->tree->theta_computed = false;
    h->tree->computeLikelihoodDerv(h->cur_dad_branch, h->cur_dad, g.data(), hs.data());
    memcpy(df, g.data(), sizeof(double) * m);
    memcpy(ddf, hs.data(), sizeof(double) * m * m);
    *lnl = g[m];
}
// computeLikelihoodDervMixlen (tree/phylokernelnew.h:3397-3580) for class `cls` at length `len`, as
// PhyloTreeMixlen::computeFuncDerv calls it (tree/phylotreemixlen.cpp:728-737)
void ref_derv_mixlen_class(void *hv, int cls, double len, double *df, double *ddf, int again) {
    RefHandle *h = (RefHandle *)hv;
    HarnessMixlenTree *t = (HarnessMixlenTree *)h->tree;
    t->cur_mixture = cls;
    t->current_it->setLength(cls, len);
    t->current_it_back->setLength(cls, len);
    if (!again) t->theta_computed = false;
    (t->*(t->computeLikelihoodDervMixlenPointer))(h->cur_dad_branch, h->cur_dad, *df, *ddf);
    t->cur_mixture = -1;
}
// pattern weights as the EM step rewrites them (tree/phylotreemixlen.cpp:470-474)
void ref_set_ptn_freq(void *hv, const double *freq) {
    RefHandle *h = (RefHandle *)hv;
    memcpy(h->tree->ptn_freq, freq, sizeof(double) * h->aln->size());
}

double ref_from_buffer(void *hv) { return ((RefHandle *)hv)->tree->computeLikelihoodFromBuffer(); }

// Newton-Raphson on the current branch exactly as the reference runs it (tree/phylotree.cpp:2585-2635)
double ref_optimize_current_branch(void *hv, int max_steps) {
    RefHandle *h = (RefHandle *)hv;
    PhyloNode *n1 = h->cur_dad, *n2 = (PhyloNode *)h->cur_dad_branch->node;
    h->tree->optimizeOneBranch(n1, n2, false, max_steps);
    return h->tree->current_it->length;
}

// PhyloTree::optimizeAllBranches as the reference runs it (tree/phylotree.cpp:2689-2765)
double ref_optimize_all_branches(void *hv, int iterations, double tolerance) {
    RefHandle *h = (RefHandle *)hv;
    h->tree->clearAllPartialLH();
    h->tree->current_it = h->tree->current_it_back = NULL;
    return h->tree->optimizeAllBranches(iterations, tolerance);
}
// PhyloTree::computeLikelihood (tree/phylotree.cpp:1227-1268): the reference's own root choice
double ref_compute_likelihood(void *hv) {
    RefHandle *h = (RefHandle *)hv;
    h->tree->clearAllPartialLH();
    h->tree->current_it = h->tree->current_it_back = NULL;
    return h->tree->computeLikelihood();
}
// length of the branch that separates the (sorted) leaf set from the rest
double ref_branch_length(void *hv, const int32_t *leaves, int nleaves, int from_taxon);

void ref_pattern_lh(void *hv, double *out) {
    RefHandle *h = (RefHandle *)hv;
    memcpy(out, h->tree->_pattern_lh, sizeof(double) * h->aln->size());
}

// _pattern_lh_cat is V-interleaved: (p/V)*V*ncm + c*V + p%V (tree/phylokernelnew.h:2801, 2857-2867)
void ref_pattern_lh_cat(void *hv, double *out) {
    RefHandle *h = (RefHandle *)hv;
    const size_t V = h->tree->vector_size, ncm = h->ncm, nptn = h->aln->size();
    for (size_t p = 0; p < nptn; p++)
        for (size_t c = 0; c < ncm; c++) out[p * ncm + c] = h->tree->_pattern_lh_cat[(p / V) * V * ncm + c * V + p % V];
}

static void leaf_set(Node *node, Node *dad, std::vector<int> &out) {
    if (node->isLeaf()) {
        out.push_back(node->id);
        return;
    }
    for (size_t i = 0; i < node->neighbors.size(); i++)
        if (node->neighbors[i]->node != dad) leaf_set(node->neighbors[i]->node, node, out);
}

static PhyloNeighbor *find_branch(Node *node, Node *dad, const std::vector<int> &want) {
    // the neighbour record (dad -> node) whose subtree at `node` has exactly the leaf set `want`
    for (size_t i = 0; i < node->neighbors.size(); i++) {
        Node *nx = node->neighbors[i]->node;
        if (nx == dad) continue;
        std::vector<int> ls;
        leaf_set(nx, node, ls);
        std::sort(ls.begin(), ls.end());
        if (ls == want) return (PhyloNeighbor *)node->neighbors[i];
        if (ls.size() > want.size()) {
            PhyloNeighbor *r = find_branch(nx, node, want);
            if (r) return r;
        }
    }
    return NULL;
}

// partial_lh and scale_num of the directed branch whose subtree carries the given (sorted) leaf set,
// de-interleaved to [nptn][ncm][n] and [nptn] (NORM) / [nptn][ncm] (SAFE)
// (layout tree/phylokernelnew.h:1914-1931).  Returns 0 on success.
int ref_fetch_branch(void *hv, const int32_t *leaves, int nleaves, int from_taxon, double *partial, uint8_t *scale) {
    RefHandle *h = (RefHandle *)hv;
    std::vector<int> want(leaves, leaves + nleaves);
    PhyloNode *start = find_leaf(h->tree, from_taxon);
    if (!start) return -1;
    PhyloNeighbor *nb = find_branch(start, NULL, want);
    if (!nb || !nb->partial_lh) return -2;
    const size_t V = h->tree->vector_size, ncm = h->ncm, n = h->aln->num_states, block = h->block;
    const size_t nptn = h->aln->size();
    const bool safe = h->tree->safe_numeric;
    for (size_t p = 0; p < nptn; p++) {
        for (size_t e = 0; e < block; e++) partial[p * block + e] = nb->partial_lh[(p / V) * V * block + e * V + p % V];
        if (safe)
            for (size_t c = 0; c < ncm; c++) scale[p * ncm + c] = nb->scale_num[p * ncm + c];
        else
            scale[p] = nb->scale_num[p];
    }
    (void)n;
    return 0;
}

// lnL at the INTERNAL branch whose far side (seen from leaf `from_taxon`) carries exactly the given leaf set, all
// partials stale: dad = the near end, dad_branch = (dad -> far end).  The kernels choose their own orientation
// (tree/phylokernelnonrev.h:1060-1067 for the non-reversible family).
static PhyloNode *owner_of(PhyloTree *tree, PhyloNeighbor *nb) {
    PhyloNode *far = (PhyloNode *)nb->node;
    for (size_t i = 0; i < far->neighbors.size(); i++) {
        PhyloNode *cand = (PhyloNode *)far->neighbors[i]->node;
        if (cand->findNeighbor(far) == nb) return cand;
    }
    return NULL;
}
double ref_lnl_at_branch(void *hv, const int32_t *leaves, int nleaves, int from_taxon) {
    RefHandle *h = (RefHandle *)hv;
    std::vector<int> want(leaves, leaves + nleaves);
    PhyloNode *start = find_leaf(h->tree, from_taxon);
    if (!start) return 1.0;
    PhyloNeighbor *nb = find_branch(start, NULL, want);
    if (!nb) return 2.0;
    PhyloNode *dad = owner_of(h->tree, nb);
    if (!dad) return 3.0;
    h->tree->clearAllPartialLH();
    h->cur_dad_branch = nb;
    h->cur_dad = dad;
    h->tree->current_it = nb;
    h->tree->current_it_back = (PhyloNeighbor *)nb->node->findNeighbor(dad);
    return h->tree->computeLikelihoodBranch(nb, dad);
}

double ref_branch_length(void *hv, const int32_t *leaves, int nleaves, int from_taxon) {
    RefHandle *h = (RefHandle *)hv;
    std::vector<int> want(leaves, leaves + nleaves);
    PhyloNode *start = find_leaf(h->tree, from_taxon);
    if (!start) return -1.0;
    if (nleaves == 1 && leaves[0] == from_taxon) return start->neighbors[0]->length;
    PhyloNeighbor *nb = find_branch(start, NULL, want);
    return nb ? nb->length : -2.0;
}

// PhyloTree::getBestNNIForBran (tree/phylotree.cpp:4058-4288), the dominant caller of a tree search (IQTree::evaluateNNIs):
// on the internal branch whose far side (seen from leaf `from_taxon`) carries exactly `leaves`, both NNI topologies are
// scored -- nni_partial_lh takes the place of the two central directions (:4119-4128), optimizeOneBranch runs on the
// central branch (and with nni5 on the four around it), computeLikelihoodFromBuffer gives the score -- and the tree is
// restored.  Out: newloglh[2], the re-optimised central length of each, and (return value) the index of the better one.
int ref_nni(void *hv, const int32_t *leaves, int nleaves, int from_taxon, int nni5, double *newloglh, double *newlen,
            double *lnl_before, double *lnl_after) {
    RefHandle *h = (RefHandle *)hv;
    PhyloTree *tree = h->tree;
    std::vector<int> want(leaves, leaves + nleaves);
    PhyloNode *start = find_leaf(tree, from_taxon);
    if (!start) return -1;
    PhyloNeighbor *nb = find_branch(start, NULL, want);
    if (!nb) return -2;
    PhyloNode *node1 = owner_of(tree, nb), *node2 = (PhyloNode *)nb->node;
    if (!node1 || node1->isLeaf() || node2->isLeaf()) return -3;
    tree->params->nni5 = nni5 != 0;
    tree->params->nni5_num_eval = 1;
    tree->clearAllPartialLH();
    tree->curScore = tree->computeLikelihood();
    if (lnl_before) *lnl_before = tree->curScore;
    NNIMove moves[2];
    moves[0].ptnlh = moves[1].ptnlh = NULL;
    moves[0].node1 = NULL;
    tree->getBestNNIForBran(node1, node2, moves);
    for (int k = 0; k < 2; k++) {
        newloglh[k] = moves[k].newloglh;
        newlen[k] = moves[k].newLen[0].empty() ? -1.0 : moves[k].newLen[0][0];
    }
    // the tree (topology, lengths, validity flags, slot ownership) must be as before: same likelihood without clearing
    if (lnl_after) *lnl_after = tree->computeLikelihood();
    return moves[0].newloglh >= moves[1].newloglh ? 0 : 1;
}

void ref_destroy(void *hv) {
    RefHandle *h = (RefHandle *)hv;
    // the tree owns its buffers; model/rate/factory are ours
    h->tree->setModelFactory(NULL);
    delete h->tree;
    delete h->factory;
    delete h->model;
    delete h->rate;
    delete h->aln;
    delete h;
}

}  // extern "C"

// target of the generated traps (oracle/gen_ref_stubs.py)
extern "C" void ref_stub_trap(const char *symbol) {
    fprintf(stderr, "oracle/_ref: reference symbol outside the likelihood path was reached: %s\n", symbol);
    abort();
}
