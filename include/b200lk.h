/*
 * b200lk.h -- C ABI of the B200 tree-likelihood engine (drop-in boundary).
 *
 * The reference (IQ-TREE 2) has no FFI on this path: kernel variants are
 * PhyloTree member-function templates bound to member-function pointers by
 * PhyloTree::setLikelihoodKernel (tree/phylotreesse.cpp:91-199).  This header is
 * what a `setLikelihoodKernelCUDA()` translation unit binds instead of the
 * templates of tree/phylokernelnew.h.  Each entry point names the reference
 * function it replaces.  INTEGRATION.md shows the adapter a maintainer adds.
 *
 * Conventions
 *  - plain C types only; every call returns an int status (B200LK_OK == 0,
 *    negative = hard error, positive = numerical-warning bit mask).  Nothing
 *    throws and nothing calls exit() across this boundary.
 *  - partial-likelihood vectors, scale counters, tip states, theta and the
 *    per-pattern buffers live in HBM for the life of the context.  A *slot* is
 *    one partial vector + its scale counters; the host addresses slots with
 *    the same index it uses for its own storage:
 *        slot = (PhyloNeighbor::partial_lh - central_partial_lh) / block_size
 *    (tree/phylotree.cpp:1145-1163); the two NNI scratch vectors
 *    (nni_partial_lh, tree/phylotree.cpp:1066-1070) are slots nslots-2, nslots-1
 *    when the host reserves them.
 *  - category index is mixture-major c = m*ncat + k, block = nstates*ncat_mix
 *    (tree/phylokernelnew.h:1322-1331).
 *  - a context is bound to one device and one stream; distinct contexts may be
 *    driven from distinct host threads concurrently (PhyloSuperTree runs
 *    partitions under OpenMP, tree/phylosupertree.cpp:716).
 */
#ifndef B200LK_H_
#define B200LK_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200LK_OK 0
#define B200LK_ERR_ARG (-1)      /* bad argument / dimension */
#define B200LK_ERR_CUDA (-2)     /* CUDA runtime error, see b200lk_last_error */
#define B200LK_ERR_STATE (-3)    /* call order violated (e.g. derv before theta) */
#define B200LK_ERR_NOMEM (-4)
#define B200LK_ERR_UNSUPPORTED (-5)
/* numerical warnings (positive bit mask), mirroring the reference's outWarning /
 * outError texts at tree/phylokernelnew.h:2573-2578, 3077-3094, 3333-3336 */
#define B200LK_WARN_LNL_NONFINITE 1  /* "Numerical underflow for lh-branch" */
#define B200LK_WARN_DERV_NONFINITE 2 /* "Numerical underflow (lh-derivative)" */
/* informational: df/ddf deliberately differ from what the reference's kernels return for this context.  With an odd
 * number of states and an even number of categories the reference's dotProductTriple tests the parity of nstates while
 * it loops over nstates*ncat_mix entries (tree/phylokernelnew.h:2491, 700-724) and counts the last (category, eigen
 * index) entry of the block twice; this library returns the mathematically right sums.  4- and 20-state data are
 * unaffected.  Set by b200lk_derv, b200lk_evaluate / b200lk_read_results (when derivatives are read) and
 * b200lk_optimize_branch. */
#define B200LK_INFO_DERV_NOT_AS_REFERENCE 4

#define B200LK_CHILD_LEAF 0
#define B200LK_CHILD_INTERNAL 1
/* non-reversible contexts only (dims.site_specific == B200LK_MODEL_NONREV) */
#define B200LK_CHILD_ROOT 2            /* the root leaf of a rooted tree (isRootLeaf, tree/mtree.h:135); id ignored */
#define B200LK_CHILD_TOWARD_ROOT 0x100 /* OR-ed into b200lk_child.kind: child->direction == TOWARD_ROOT */
#define B200LK_CHILD_KIND_MASK 0xff

/* values of b200lk_dims.site_specific: which kernel family setLikelihoodKernel selects
 * (tree/phylokernelfma.cpp:41-104) */
#define B200LK_MODEL_REV 0    /* reversible kernels, tree/phylokernelnew.h */
#define B200LK_MODEL_SITE 1   /* SITE_MODEL variants: model->isSiteSpecificModel() */
#define B200LK_MODEL_NONREV 2 /* computeNonrev*: !model->isReversible() || params->kernel_nonrev, tree/phylokernelnonrev.h */

typedef struct b200lk_ctx b200lk_ctx;

/* Problem dimensions; fixed for the life of a context
 * (what PhyloTree::initializeAllPartialLh sizes its buffers from,
 * tree/phylotree.cpp:890-934, 1054-1110). */
typedef struct b200lk_dims {
    int32_t ntaxa;          /* leafNum */
    int32_t nstates;        /* aln->num_states */
    int32_t ncat;           /* site_rate->getNRate() */
    int32_t nmix;           /* model->getNMixtures() */
    int32_t fused_mix_rate; /* model_factory->fused_mix_rate (0/1) */
    int32_t safe_numeric;   /* 0 = NORM_LH (scale per pattern), 1 = SAFE_LH (per pattern and category) */
    int32_t n_state_codes;  /* aln->STATE_UNKNOWN + 1 */
    int32_t nslots;         /* number of partial-vector slots (max_lh_slots [+2 NNI]) */
    int64_t nptn;           /* aln->size(): observed site patterns */
    int64_t n_unobserved;   /* model_factory->unobserved_ptns.size() (+ASC), 0 otherwise */
    int32_t device;         /* CUDA device ordinal */
    int32_t site_specific;  /* B200LK_MODEL_REV / _SITE (model->isSiteSpecificModel(): ModelSet, PMSF; per-pattern
                             * eigen-systems, nmix = 1) / _NONREV (non-reversible kernel family) */
} b200lk_dims;

/* One directed branch of a traversal (TraversalInfo, tree/phylotree.h:320-331):
 * the partial vector at `out_slot` is computed from children
 * [child_begin, child_begin + nchild) of the child array. */
typedef struct b200lk_node {
    int32_t out_slot;
    int32_t child_begin;
    int32_t nchild; /* 2 for a bifurcating node, >2 multifurcating */
    int32_t reserved;
} b200lk_node;

typedef struct b200lk_child {
    int32_t kind; /* B200LK_CHILD_LEAF: id = taxon (alignment row); B200LK_CHILD_INTERNAL: id = slot */
    int32_t id;
    double length; /* branch length to the child (child->length) */
} b200lk_child;

/* One end of the branch an lnL / theta is taken at.  kind LEAF: id = taxon;
 * INTERNAL: id = slot holding the partial that points toward that end. */
typedef struct b200lk_end {
    int32_t kind;
    int32_t id;
} b200lk_end;

/* ---- life cycle ------------------------------------------------------- */
int b200lk_create(b200lk_ctx **out, const b200lk_dims *dims);
/* One problem over several GPUs of THIS process (SURVEY.md 8b "device_mask", 8e): the reference is one process whose
 * kernels split the pattern range into packets for its OpenMP threads and sum the per-packet lnL / df / ddf
 * (tree/phylokernelnew.h:1114-1142, 2377, 2792, 3273); this is the device form of that.  `devices` lists ndev CUDA
 * ordinals (dims->device is ignored; an ordinal may repeat -- several shards on one GPU -- which only tests use).  The context owns one ordinary context per device over a contiguous range of
 * whole 32-pattern tiles of EVERY partial vector (b200lk_shard_info); model tables are computed redundantly per GPU,
 * nothing is exchanged during a traversal.  Every entry point below works unchanged on the returned context: it
 * enqueues on all shards first (they run concurrently, driven from the one calling thread), then adds the per-shard
 * sums on the host in device-index order -- bit-reproducible from run to run -- and the b200lk_fetch_* readers gather
 * the shards' ranges.  b200lk_optimize_branch runs the reference's safeguarded Newton iteration with one fan-out and
 * one ordered sum per step.  Not available on a sharded context (B200LK_ERR_UNSUPPORTED, at create or at the call):
 * +ASC (its sums must be combined before the non-linear correction), site-specific and non-reversible contexts,
 * heterotachy, robust scoring, batches, b200lk_set_stream / b200lk_result_device_ptr, b200lk_lnl_from_buffer_noalloc. */
int b200lk_create_sharded(b200lk_ctx **out, const b200lk_dims *dims, const int32_t *devices, int32_t ndev);
/* Shards of a context: their count, devices and observed-pattern ranges [begin, end) (any pointer may be NULL; arrays of
 * at least *nshards entries).  An ordinary context reports one shard covering [0, nptn). */
int b200lk_shard_info(b200lk_ctx *ctx, int32_t *nshards, int32_t *devices, int64_t *begin, int64_t *end);
int b200lk_destroy(b200lk_ctx *ctx);
const char *b200lk_last_error(const b200lk_ctx *ctx);
const char *b200lk_version(void);

/* Optional: run on a caller-owned stream (e.g. the stream NCCL is enqueued on).
 * `stream` is a cudaStream_t passed as void*; NULL restores the context's own -- so the default stream, whose handle
 * is 0, cannot be selected this way: share a dedicated stream, or pass cudaStreamLegacy / cudaStreamPerThread. */
int b200lk_set_stream(b200lk_ctx *ctx, void *stream);

/* ---- inputs (one-time / per model change) ----------------------------- */
/* Tip data + pattern weights.  states: [ntaxa][nptn] uint8 state codes
 * (aln->at(ptn)[taxon], tree/phylotree.cpp:3242-3247); unobserved: [ntaxa][n_unobserved]
 * or NULL; ptn_freq: [nptn] (computePtnFreq, tree/phylotreesse.cpp:537-547). */
int b200lk_upload_tips(b200lk_ctx *ctx, const uint8_t *states, const uint8_t *unobserved,
                       const double *ptn_freq);
/* ptn_invar[nptn + n_unobserved] (computePtnInvar, tree/phylotreesse.cpp:549-612). */
int b200lk_set_ptn_invar(b200lk_ctx *ctx, const double *ptn_invar);
/* Eigen-system and rate model: what the kernels read through
 * model->getEigenvalues()/getEigenvectors()/getInverseEigenvectors()
 * (model/modelmarkov.cpp:1904-1916), site_rate->getRate/getProp, getMixtureWeight and
 * tip_partial_lh (computeTipPartialLikelihood, tree/phylotreesse.cpp:243-375).
 * eval[nmix*n], evec[nmix*n*n] (row = state), inv_evec[nmix*n*n] (row = eigen index),
 * rates[ncat], props[ncat], mixw[nmix], tip_partial_lh[n_state_codes*nmix*n]. */
int b200lk_set_model(b200lk_ctx *ctx, const double *eval, const double *evec, const double *inv_evec,
                     const double *rates, const double *props, const double *mixw,
                     const double *tip_partial_lh);

/* The same from the model's free parameters, eigen-decomposition on the device:
 * ModelMarkov::decomposeRateMatrix for reversible models (model/modelmarkov.cpp:1386-1632 ->
 * EigenDecomposition::eigensystem_sym, utils/eigendecomposition.cpp:165-317) + computeTipPartialLikelihood.
 * exch[nmix][n(n-1)/2]: upper triangle of the exchangeabilities, row-major (ModelMarkov::rates, half_matrix);
 * freqs[nmix][n]; total_num_subst[nmix] or NULL (1.0 each; the class rate of a mixture);
 * tipvec[n_state_codes][n]: 0/1 indicator (ambiguity) vectors of the state codes (computeTipLikelihood,
 * model/modelsubst.cpp:182-207).  nstates <= 64.  Eigenpairs come out in the solver's own order. */
int b200lk_set_model_rates(b200lk_ctx *ctx, const double *exch, const double *freqs, const double *total_num_subst,
                           const double *tipvec, const double *rates, const double *props, const double *mixw);
/* the eigen-system the context currently holds (eval[nmix*n], evec[nmix*n*n], inv_evec[nmix*n*n]) */
int b200lk_fetch_model(b200lk_ctx *ctx, double *eval, double *evec, double *inv_evec);

/* Site-specific models (the SITE_MODEL = true kernel variants selected when
 * model->isSiteSpecificModel(), tree/phylokernelfma.cpp:46-83): every pattern has its own eigen-system, as a ModelSet
 * holds them after decomposeRateMatrix (model/modelset.cpp:129-200; PMSF frequency profiles).  The context must have
 * been created with dims.site_specific = 1 (nmix = 1, no +ASC, nstates <= 32) and the tips uploaded; the per-taxon,
 * per-pattern tip vectors of computeTipPartialLikelihood (tree/phylotreesse.cpp:257-340) are built on the device.
 * eval[nptn][n], evec[nptn][n][n] (row = state), inv_evec[nptn][n][n] (row = eigen index), pattern-major (no vector
 * interleaving), tipvec[n_state_codes][n] as in b200lk_set_model_rates. */
int b200lk_set_site_models(b200lk_ctx *ctx, const double *eval, const double *evec, const double *inv_evec,
                           const double *rates, const double *props, const double *tipvec);

/* Non-reversible kernel family (computeNonrevPartialLikelihood / LikelihoodBranch /
 * LikelihoodDerv, tree/phylokernelnonrev.h:43-1517, selected when !model->isReversible() || params->kernel_nonrev,
 * tree/phylokernelfma.cpp:92-100).  The context must have been created with dims.site_specific = B200LK_MODEL_NONREV
 * (no +ASC: the reference's own tip-internal case indexes the unobserved patterns out of bounds,
 * tree/phylokernelnonrev.h:399).  Partial vectors hold state likelihoods (no eigen-space), children are combined
 * through transition matrices P(rate_c * length), transposed for a child that lies TOWARD_ROOT
 * (tree/phylokernelnew.h:932-956), and the root leaf contributes the state frequencies (:959-964).
 *   rate_matrix[nmix][n][n]  Q, row-major (ModelMarkov::rate_matrix): dP/dt = Q P, d2P/dt2 = Q Q P
 *                            (model/modelmarkov.cpp:729-745)
 *   ceval[nmix][n][2], cevec[nmix][n][n][2], cinv_evec[nmix][n][n][2]
 *                            complex eigen-system (re, im interleaved): P(t) = Re(cevec exp(ceval t) cinv_evec)
 *                            (ModelMarkov::computeTransMatrixNonrev, model/modelmarkov.cpp:481-488)
 *   freqs[nmix][n]           state frequencies (root leaf; unrooted trees weight the branch by them)
 *   tipvec[n_state_codes][n] as in b200lk_set_model_rates; the last code is the unknown state (all ones)
 *   rooted                   tree->rooted
 * Call sequence for a branch: the caller orders the two ends as the kernels do after their swap
 * (tree/phylokernelnonrev.h:596-603, 1060-1067): `dad` is the leaf end if there is one (the ROOT leaf only when it
 * is the original dad), otherwise the end farther from the root; b200lk_build_theta only records the ends,
 * b200lk_derv recomputes from the two partial vectors as the reference does, b200lk_lnl_from_buffer is
 * b200lk_branch_lnl at the recorded ends (the reference binds no from-buffer kernel, tree/phylokernelfma.cpp:99).
 * b200lk_derv at a ROOT dad returns B200LK_ERR_UNSUPPORTED (the virtual root branch has no length,
 * tree/phylokernelnonrev.h:667-670), and so do b200lk_optimize_branch and batches. */
int b200lk_set_model_nonrev(b200lk_ctx *ctx, const double *rate_matrix, const double *ceval, const double *cevec,
                            const double *cinv_evec, const double *freqs, const double *tipvec, const double *rates,
                            const double *props, const double *mixw, int rooted);

/* ---- the four kernel entry points -------------------------------------- */
/* computePartialLikelihood over a whole traversal list
 * (computeTraversalInfo + computePartialInfo + computePartialLikelihood*SIMD,
 * tree/phylokernelnew.h:900-1110, 1150-1281, 1291-2039).  `nodes` is in post-order
 * (children before parents); independent nodes are batched into one launch. */
int b200lk_update_partials(b200lk_ctx *ctx, const b200lk_node *nodes, int32_t nnodes,
                           const b200lk_child *children, int32_t nchildren);

/* computeLikelihoodBranch*SIMD (tree/phylokernelnew.h:2643-3185) on branch (dad, node) of
 * length `length`: writes _pattern_lh / _pattern_lh_cat on the device, returns lnL.
 * `dad` may be a leaf (tip-internal case); `node` must be INTERNAL. */
int b200lk_branch_lnl(b200lk_ctx *ctx, b200lk_end dad, b200lk_end node, double length, double *lnl);

/* computeLikelihoodBuffer*SIMD (tree/phylokernelnew.h:2050-2231): theta_all and buffer_scale_all. */
int b200lk_build_theta(b200lk_ctx *ctx, b200lk_end dad, b200lk_end node);

/* computeLikelihoodDerv*SIMD (tree/phylokernelnew.h:2235-2633) from theta_all. */
int b200lk_derv(b200lk_ctx *ctx, double length, double *df, double *ddf);

/* computeLikelihoodFromBuffer*SIMD (tree/phylokernelnew.h:3196-3392) from theta_all. */
int b200lk_lnl_from_buffer(b200lk_ctx *ctx, double length, double *lnl);

/* b200lk_lnl_from_buffer for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  NO COUNTERPART IN THE REFERENCE.  Same
 * inputs (theta_all of the last b200lk_build_theta / b200lk_evaluate / b200lk_optimize_branch, the model, `length`),
 * same outputs (*lnl, _pattern_lh on the device) as b200lk_lnl_from_buffer, but the scratch of the computation -- the
 * branch coefficient table val[3][nstates*ncat_mix] and the per-CTA partial sums -- lives in one file-scope
 * `static __device__` array of the library (B200LK_NOALLOC_SCRATCH_BYTES on every GPU that loads it, until exit)
 * whose device address is looked up once per device at b200lk_create.  The call allocates and frees nothing.
 * Limits: nstates * ncat_mix <= B200LK_NOALLOC_MAX_BLOCK, otherwise B200LK_ERR_ARG with a message in
 * b200lk_last_error; reversible, non-site-specific contexts without Holder's +ASC and without mixed branch lengths
 * (B200LK_ERR_UNSUPPORTED otherwise).  The array is SHARED by every call on a device and keeps the last call's
 * contents (each call overwrites all it reads, nothing needs clearing): calls that overlap on one GPU -- two streams,
 * two host threads, two contexts -- race on it; serialise them. */
#define B200LK_NOALLOC_MAX_BLOCK 4800 /* doubles per pattern: LG+C60+G4 (20 states x 240 categories) still fits */
#define B200LK_NOALLOC_MAX_CTAS 2048  /* per-CTA partial sums kept: 2 per CTA */
#define B200LK_NOALLOC_SCRATCH_BYTES (8 * (3 * B200LK_NOALLOC_MAX_BLOCK + 2 * B200LK_NOALLOC_MAX_CTAS)) /* 147 968 */
int b200lk_lnl_from_buffer_noalloc(b200lk_ctx *ctx, double length, double *lnl);

/* ---- +ASC flavour -------------------------------------------------------- */
/* Which ascertainment-bias correction the unobserved patterns stand for (model_factory->getASC(), utils/tools.h:549-555).
 * B200LK_ASC_LEWIS (the default when dims.n_unobserved > 0): the sum of their likelihoods is the probability of an
 * unobservable constant site (tree/phylokernelnew.h:2620-2628, 3160-3182).  B200LK_ASC_HOLDER (ASC_VARIANT_MISSING):
 * dims.n_unobserved == nstates * nptn, state-major -- unobserved pattern x*nptn + p is the constant-x column that keeps
 * the missing entries of observed pattern p (Alignment::getUnobservedConstPatterns, alignment/alignment.cpp:4150-4172) --
 * and every observed pattern is corrected by its own 1 - sum_x L(x*nptn + p) (:2522-2525, 2580-2619, 3137-3159,
 * 3349-3368).  Reversible contexts only; b200lk_optimize_branch and batches refuse +ASC as before. */
#define B200LK_ASC_LEWIS 0
#define B200LK_ASC_HOLDER 1
int b200lk_set_asc(b200lk_ctx *ctx, int32_t type);

/* "Robust phylogeny" scoring of computeLikelihoodBranch (tree/phylokernelnew.h:3096-3135): with keep < 1 (params->
 * robust_phy_keep) b200lk_branch_lnl returns the sum of the per-site log-likelihoods without the (int)((1-keep)*nsite)
 * worst sites; with median != 0 (params->robust_median) the median per-site log-likelihood.  keep = 1, median = 0
 * (the default) is the plain sum.  The reference selects with nth_element over a per-site array on the host; here the
 * per-pattern values are sorted with their frequencies on the device.  Like the reference: b200lk_lnl_from_buffer is
 * unaffected, and +ASC contexts refuse it (B200LK_ERR_UNSUPPORTED from b200lk_branch_lnl). */
int b200lk_set_robust(b200lk_ctx *ctx, double keep, int32_t median);

/* New pattern weights without re-uploading the tips: PhyloTree::computePtnFreq (tree/phylotreesse.cpp:537-547), or the
 * posterior class weights the heterotachy EM step writes into ptn_freq (tree/phylotreemixlen.cpp:470-474). */
int b200lk_set_ptn_freq(b200lk_ctx *ctx, const double *ptn_freq /* [nptn] */);

/* ---- heterotachy: mixed branch lengths ------------------------------------------ */
/* PhyloTreeMixlen (GHOST, +H models): every branch carries one length per rate category, and the kernels read
 * child->getLength(c % ncat) / dad_branch->getLength(c % ncat) (tree/phylokernelnew.h:1017, 1069, 2325, 2348, 2711,
 * 2725, 3248, 3263; PhyloNeighborMixlen::getLength, tree/phylonodemixlen.h:64-69).  These are the four kernel calls with
 * that array in place of the single length, plus the two derivative forms the reference binds for isMixlen() trees.
 * Reversible, non-site-specific contexts; not available through batches or b200lk_optimize_branch.
 *   child_len[j*ncat + k]  length of class k on the branch to children[j]  (b200lk_child::length is ignored)
 *   len[k]                 length of class k on the branch (dad, node) */
int b200lk_update_partials_mixlen(b200lk_ctx *ctx, const b200lk_node *nodes, int32_t nnodes,
                                  const b200lk_child *children, int32_t nchildren, const double *child_len);
int b200lk_branch_lnl_mixlen(b200lk_ctx *ctx, b200lk_end dad, b200lk_end node, const double *len, double *lnl);
int b200lk_lnl_from_buffer_mixlen(b200lk_ctx *ctx, const double *len, double *lnl);
/* computeLikelihoodDerv*SIMD with isMixlen() (tree/phylokernelnew.h:2365-2462, 2556-2568): the gradient df[ncat] over
 * the class lengths, the Hessian ddf[ncat][ncat], and the log-likelihood (the reference returns it in df[ncat]).
 * ncat <= 8.  +ASC is refused like the reference does (ASSERT(0) at :2449). */
int b200lk_derv_mixlen(b200lk_ctx *ctx, const double *len, double *df, double *ddf, double *lnl);
/* computeLikelihoodDervMixlen*SIMD (tree/phylokernelnew.h:3397-3580; computeLikelihoodDervMixlenPointer,
 * tree/phylotree.h:1339): df, ddf of the EM sub-problem of class cur_mixlen (getCurMixture()) at length `len`, under the
 * pattern weights currently loaded (b200lk_set_ptn_freq).  Lewis +ASC terms included (:3560-3568).  Like the EM step
 * itself (tree/phylotreemixlen.cpp:427-428) it requires fused_mix_rate or nmix == 1. */
int b200lk_derv_mixlen_class(b200lk_ctx *ctx, int32_t cur_mixlen, double len, double *df, double *ddf);

/* Newton-Raphson on one branch, whole iteration on the device: what PhyloTree::optimizeOneBranch asks of
 * Optimization::minimizeNewton (tree/phylotree.cpp:2585-2635, utils/optimization.cpp:422-499) with
 * computeFuncDerv = computeLikelihoodDerv (tree/phylotree.cpp:2520-2527).  Builds theta for branch (dad, node),
 * then runs the bisection-safeguarded iteration t <- t - f/f' (f = -df, f' = -ddf) inside [min_len, max_len]
 * without a host round trip per step.  Returns the optimised length, d2l (= -ddf at the last evaluation, the
 * reference's `negative_lh` out-parameter) and the number of derivative evaluations.  theta stays valid for a
 * following b200lk_lnl_from_buffer. */
typedef struct b200lk_newton {
    double min_len, max_len; /* params->min_branch_length, max_branch_length */
    double xacc;             /* accuracy: the reference passes min_branch_length */
    double guess;            /* current_it->length */
    int32_t max_steps;       /* maxNRStep (default 100) */
    int32_t reserved;
} b200lk_newton;
int b200lk_optimize_branch(b200lk_ctx *ctx, b200lk_end dad, b200lk_end node, const b200lk_newton *nr,
                           double *optx, double *d2l, int32_t *nsteps);

/* ---- branch lengths resident on the device: a whole optimizeAllBranches sweep without a read-back --------------------
 * PhyloTree::optimizeAllBranches (tree/phylotree.cpp:2689-2765) re-roots at every branch in turn: one partial update across
 * the branch just optimised, theta, Newton-Raphson.  The next update needs the length just found; when the host forms the
 * exp(lambda r t) factors, that is a device->host round trip per branch.  With these entry points the lengths live in a
 * device table indexed by the caller's own branch id: the children of an update name slots (len_slot[j] >= 0; -1 keeps
 * children[j].length), the Newton-Raphson launch starts from lengths[len_slot] and writes the optimum back, and the host
 * reads the table ONCE per sweep -- with each slot's evaluation count and status (aux[slot][0..1]; status 1 = non-finite
 * derivative), from which it applies the checks optimizeOneBranch makes per branch (:2607-2616).  Reversible,
 * non-site-specific, non-sharded contexts without +ASC; bifurcating nodes on the 4-state path. */
int b200lk_set_lengths(b200lk_ctx *ctx, const double *lengths, int32_t n);
int b200lk_get_lengths(b200lk_ctx *ctx, double *lengths, double *aux /* [n][2] or NULL */, int32_t n);
int b200lk_update_partials_slots(b200lk_ctx *ctx, const b200lk_node *nodes, int32_t nnodes, const b200lk_child *children,
                                 int32_t nchildren, const int32_t *len_slot /* [nchildren] */);
int b200lk_optimize_branch_slot(b200lk_ctx *ctx, b200lk_end dad, b200lk_end node, const b200lk_newton *nr, int32_t len_slot);

/* One round trip for a whole evaluation: update_partials + branch_lnl + build_theta + derv
 * (what computeLikelihood followed by the first computeFuncDerv of optimizeOneBranch costs,
 * tree/phylotree.cpp:1227-1268, 2520-2527).  Any of lnl/df/ddf may be NULL. */
int b200lk_evaluate(b200lk_ctx *ctx, const b200lk_node *nodes, int32_t nnodes,
                    const b200lk_child *children, int32_t nchildren, b200lk_end dad, b200lk_end node,
                    double length, double *lnl, double *df, double *ddf);

/* Asynchronous form of b200lk_evaluate: everything is enqueued on the context's stream and the sums
 * stay in the device result buffer (b200lk_result_device_ptr) -- e.g. to be all-reduced across
 * pattern shards by NCCL on the same stream -- until b200lk_read_results copies them back
 * (the only device->host transfer of an evaluation: 3 doubles). */
int b200lk_evaluate_async(b200lk_ctx *ctx, const b200lk_node *nodes, int32_t nnodes,
                          const b200lk_child *children, int32_t nchildren, b200lk_end dad, b200lk_end node,
                          double length, int want_lnl, int want_derv);
int b200lk_read_results(b200lk_ctx *ctx, double *lnl, double *df, double *ddf);

/* ---- pattern shards in several PROCESSES: the sums exchanged inside the finishing kernel ------------------------------
 * One process per GPU (MPI ranks, torchrun), each with a context over its pattern range.  The reference adds the
 * per-packet lnL / df / ddf of its OpenMP threads with reduction(+) (tree/phylokernelnew.h:2377, 2792, 3273); across
 * GPUs that is an all-reduce of 3 doubles.  Instead of a collective library call per evaluation, the kernel that
 * finishes b200lk_evaluate_async stores the rank's sums into every peer's exchange buffer over NVLink (CUDA IPC peer
 * memory), waits for the peers' and adds them in rank order: every rank's result buffer then holds bit-identical
 * totals and b200lk_read_results returns them.  Set-up, once: each rank calls b200lk_peer_handle, the 64-byte handles
 * are gathered by whatever the host has (MPI_Allgather, torch.distributed), each rank calls b200lk_peer_connect.
 * Every rank must then issue the same sequence of b200lk_evaluate* calls (a rank waits in the kernel for the others,
 * like in any collective).  Only b200lk_evaluate_async / b200lk_evaluate exchange; the other entry points stay local. */
int b200lk_peer_handle(b200lk_ctx *ctx, void *handle64 /* 64 bytes out */);
int b200lk_peer_connect(b200lk_ctx *ctx, int32_t rank, int32_t world, const void *handles /* [world][64] */);
int b200lk_peer_disconnect(b200lk_ctx *ctx); /* back to purely local sums */

/* ---- device-resident results on demand --------------------------------- */
/* Unreduced per-shard sums for multi-GPU runs: out[0] = lnL (if requested), out[1] = df,
 * out[2] = ddf are left in a device buffer; this returns its device pointer (3 doubles)
 * so the caller can all-reduce it in place (NCCL) on the context's stream. */
int b200lk_result_device_ptr(b200lk_ctx *ctx, void **dptr);

/* Host readers of device data (tree/phylotree.cpp:1252-1259, 1463-1581). */
int b200lk_fetch_pattern_lh(b200lk_ctx *ctx, double *pattern_lh /* [nptn] */);
int b200lk_fetch_pattern_lh_cat(b200lk_ctx *ctx, double *pattern_lh_cat /* [nptn][ncat_mix] */);
int b200lk_fetch_scale_num(b200lk_ctx *ctx, int32_t slot, uint8_t *scale /* [nptn] or [nptn][ncat_mix] */);
int b200lk_fetch_partial(b200lk_ctx *ctx, int32_t slot, double *partial /* [nptn][ncat_mix][nstates] */);
int b200lk_fetch_theta(b200lk_ctx *ctx, double *theta /* [nptn][ncat_mix][nstates] */,
                       double *buffer_scale_all /* [nptn] or NULL */);

/* ---- UFBoot RELL ------------------------------------ */
/* IQTree::saveCurrentTree (tree/iqtree.cpp:3553-3650): for every bootstrap resample b,
 *     rell[b] = dotProduct(pattern_lh, boot_samples[b], nptn)      (tree/phylokernel.h:64-77, BootValType = float)
 * with pattern_lh the per-pattern log-likelihoods of the tree just evaluated.  The resampled pattern frequencies
 * are uploaded once and stay resident ([nsamples][nptn] float, the reference's own BootValType); _pattern_lh never
 * leaves the device (it is what the last b200lk_branch_lnl / b200lk_lnl_from_buffer / b200lk_evaluate wrote), which
 * removes the largest recurring device->host copy of a tree search (nptn doubles per candidate tree).  Products are
 * formed in float like the reference's, sums are accumulated in double in a fixed order. */
int b200lk_upload_boot_samples(b200lk_ctx *ctx, const float *samples /* [nsamples][nptn] */, int32_t nsamples);
int b200lk_rell(b200lk_ctx *ctx, double *rell /* [nsamples] */);

/* ---- partition dispatch as batched launches --------- */
/* PhyloSuperTree::computeLikelihood (tree/phylosupertree.cpp:701-725) evaluates every partition's own tree with an
 * OpenMP loop over partitions; PhyloSuperTreePlen::computeFuncDerv (tree/phylosupertreeplen.cpp:250-330) does the
 * same for the derivatives of one linked edge.  A batch is the device form of that loop: contexts with the same
 * dimensions (ntaxa, nstates, ncat, nmix, safe_numeric, n_state_codes, nslots; no +ASC), evaluated on the SAME
 * topology and traversal list, run each step as one launch whose extra grid dimension is the partition.  Models,
 * alignments, pattern counts and every branch length stay per partition.
 *   nodes/children: the traversal of the first context (b200lk_child::length is ignored);
 *   child_len[i*nchildren + k]: length of the branch to children[k] in context i;  branch_len[i]: the dad-node branch.
 * Configurations a batched launch cannot express (other state counts, mixtures of more than 4 category-classes,
 * multifurcations, +ASC) return B200LK_ERR_UNSUPPORTED and the caller uses the per-context calls.
 * Results per context: lnl[n] (and df[n], ddf[n] when want_derv; theta is then left built in every context). */
typedef struct b200lk_batch b200lk_batch;
int b200lk_batch_create(b200lk_batch **out, b200lk_ctx **ctxs, int32_t n);
int b200lk_batch_destroy(b200lk_batch *batch); /* the contexts stay owned by the caller */
const char *b200lk_batch_last_error(const b200lk_batch *batch);
int b200lk_batch_evaluate(b200lk_batch *batch, const b200lk_node *nodes, int32_t nnodes, const b200lk_child *children,
                          int32_t nchildren, const double *child_len /* [n][nchildren] */, b200lk_end dad,
                          b200lk_end node, const double *branch_len /* [n] */, int want_derv, double *lnl /* [n] */,
                          double *df /* [n] or NULL */, double *ddf /* [n] or NULL */);
/* The same in two halves, so that several batches (say the DNA genes and the protein genes of one supermatrix) are
 * all enqueued before any result is waited for; read_results blocks and fills what the evaluation asked for. */
int b200lk_batch_evaluate_async(b200lk_batch *batch, const b200lk_node *nodes, int32_t nnodes,
                                const b200lk_child *children, int32_t nchildren, const double *child_len,
                                b200lk_end dad, b200lk_end node, const double *branch_len, int want_derv);
int b200lk_batch_read_results(b200lk_batch *batch, double *lnl, double *df, double *ddf);
int b200lk_batch_build_theta(b200lk_batch *batch, b200lk_end dad, b200lk_end node);
int b200lk_batch_derv(b200lk_batch *batch, const double *branch_len /* [n] */, double *df, double *ddf);
int b200lk_batch_lnl_from_buffer(b200lk_batch *batch, const double *branch_len /* [n] */, double *lnl);

/* Counters for the metric: pattern-node updates and kernel launches since create/reset. */
int b200lk_get_counters(b200lk_ctx *ctx, int64_t *pattern_node_updates, int64_t *kernel_launches);
int b200lk_reset_counters(b200lk_ctx *ctx);
/* Minimum DRAM bytes of the partial-update kernels launched since create/reset, counted from the node programs the
 * library actually launched (not from the reference's loop nest): every partial vector written once with its scale
 * counters, every child vector a launch must fetch from HBM (all internal children for the level-batched kernels;
 * for the 4-state pattern-stationary traversal only children produced by an earlier launch -- a vector of the same
 * launch that no longer fits the per-thread stack is re-read by the CTA that just wrote it, from L2), one byte per
 * tip state.  The denominator-free half of bench.py's roofline fraction.  Replaces
 * nothing in the reference. */
int b200lk_get_min_traffic(b200lk_ctx *ctx, int64_t *min_dram_bytes);
int b200lk_synchronize(b200lk_ctx *ctx);
/* Optional per-launch CUDA-event timing of the partial-update kernels (the dominant kernel):
 * accumulated milliseconds and launch count since the last reset.  Replaces nothing in the
 * reference (its only timers are getRealTime/getCPUTime, utils/timeutil.h:95-134). */
int b200lk_set_profiling(b200lk_ctx *ctx, int on);
int b200lk_get_profile(b200lk_ctx *ctx, double *partial_ms, int64_t *partial_launches, int reset);

#ifdef __cplusplus
}
#endif
#endif /* B200LK_H_ */
