/*
 * phylokernel_oracle.c -- CPU restatement of IQ-TREE 2's reversible likelihood
 * kernel (tree/phylokernelnew.h), scalar (V = 1) semantics, plain C.
 *
 * THIS IS TEST INFRASTRUCTURE, NOT THE PRODUCT.  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may call it; the product
 * path (iqtree2_b200/csrc) never links or loads it.
 *
 * Parity status: PINNED against the reference's own kernels -- oracle/_ref builds
 * tree/phylokernel{sse,fma}.cpp, phylotreeavx.cpp, phylotree.cpp ... unmodified from
 * /root/reference and tests/test_oracle_vs_ref.py + tests/golden/ hold its outputs
 * (the reference ships no golden vectors of its own).
 *
 * It exposes the same call sequence as include/b200lk.h with prefix orc_ so a test
 * can drive both with identical arguments.  Each function cites the reference
 * lines it follows.  Memory layout: element (pattern p, category c, state x) of a
 * partial vector at p*block + c*nstates + x (the reference's layout for
 * VectorClass::size() == 1, tree/phylokernelnew.h:1914-1916); scale_num[p] (NORM)
 * or scale_num[p*ncat_mix + c] (SAFE) (:1920-1931).
 *
 * Floating-point order follows the FMA build (mul_add == fused multiply-add,
 * vectorclass) with the reduction trees of the helper templates at :43-842.
 * Compile with -ffp-contract=off so that only the explicit fma() calls fuse.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "../include/b200lk.h"

#define SCALING_THRESHOLD 8.636168555094444625386e-78 /* 2^-256, tree/phylotree.h:77 */
#define SCALING_THRESHOLD_EXP 256                      /* tree/phylotree.h:73 */
#define LOG_SCALING_THRESHOLD (-177.4456782233459932741) /* tree/phylotree.h:80 */

typedef struct orc_ctx {
    b200lk_dims d;
    int ncat_mix, block, nthreads;
    int64_t ntot; /* nptn + n_unobserved */
    uint8_t *states;     /* [ntaxa][ntot] */
    double *ptn_freq;    /* [ntot] (0 beyond nptn) */
    double *ptn_invar;   /* [ntot] */
    double *eval, *evec, *inv_evec, *rates, *props, *mixw, *tip_partial_lh;
    double **partial;    /* [nslots] -> [ntot][block] */
    uint8_t **scale;     /* [nslots] -> [ntot] or [ntot][ncat_mix] */
    double *theta_all, *buffer_scale_all, *pattern_lh, *pattern_lh_cat;
    int theta_computed;
    int64_t updates;
    /* site-specific models (SITE_MODEL = true variants; ModelSet / PMSF): per-pattern eigen-systems
     * s_eval[nptn][n], s_evec[nptn][n][n], s_inv[nptn][n][n]; s_tip[ntaxa][nptn][n] = the reference's
     * tip_partial_lh of site models (tree/phylotreesse.cpp:257-340) */
    int site;
    double *s_eval, *s_evec, *s_inv, *s_tip;
    /* non-reversible kernel family (nonrev_oracle.inc): Q[nmix][n][n], complex eigen-system (re, im interleaved),
     * state frequencies, tip indicator vectors; the ends recorded by orc_build_theta */
    int nonrev, rooted;
    double *nr_Q, *nr_ev, *nr_V, *nr_Vi, *nr_freq, *nr_tipvec;
    b200lk_end nr_dad, nr_node;
    /* Holder's +ASC for missing data (ASC_VARIANT_MISSING): n_unobserved = nstates * nptn patterns, state-major
     * (Alignment::getUnobservedConstPatterns, alignment/alignment.cpp:4150-4172); per-pattern derivative terms */
    int asc_holder;
    double robust_keep; /* params->robust_phy_keep (1.0 = off) */
    int robust_median;  /* params->robust_median */
    double nsite; /* aln->getNSite(): sum of the pattern frequencies uploaded with the tips */
    double *const_df, *const_ddf; /* [n_unobserved] */
} orc_ctx;

/* ------------------------------------------------------------------ helpers */

/* dotProductVec, tree/phylokernelnew.h:82-126 (JAMES_VERSION branch) */
static inline double dot_vec(const double *A, const double *B, int N) {
    if (N == 4) return (A[0] * B[0] + A[1] * B[1]) + (A[2] * B[2] + A[3] * B[3]);
    int cruft = N & 3;
    if (N == cruft) {
        double X = 0;
        for (int i = 0; i < cruft; i++) X = fma(A[i], B[i], X);
        return X;
    }
    double X = A[0] * B[0], v1 = A[1] * B[1], v2 = A[2] * B[2], v3 = A[3] * B[3];
    int i;
    for (i = 4; i < N - cruft; i += 4) {
        X = fma(A[i], B[i], X);
        v1 = fma(A[i + 1], B[i + 1], v1);
        v2 = fma(A[i + 2], B[i + 2], v2);
        v3 = fma(A[i + 3], B[i + 3], v3);
    }
    if (0 < cruft) {
        X = fma(A[i], B[i], X);
        if (1 < cruft) {
            v1 = fma(A[i + 1], B[i + 1], v1);
            if (2 < cruft) v2 = fma(A[i + 2], B[i + 2], v2);
        }
    }
    return (X + v1) + (v2 + v3);
}

/* one row of productVecMat / one factor of dotProductDualVec: the reduction tree at
 * tree/phylokernelnew.h:223-313 and :333-435 (identical shapes) */
static inline double dot_tree(const double *A, const double *B, int N) {
    switch (N) {
    case 1: return A[0] * B[0];
    case 2: return fma(A[1], B[1], A[0] * B[0]);
    case 3: return fma(A[2], B[2], fma(A[1], B[1], A[0] * B[0]));
    case 4: return fma(A[1], B[1], A[0] * B[0]) + fma(A[3], B[3], A[2] * B[2]);
    default: break;
    }
    switch (N % 4) {
    case 0: {
        double V[4];
        for (int j = 0; j < 4; j++) V[j] = A[j] * B[j];
        for (int i = 4; i < N; i += 4)
            for (int j = 0; j < 4; j++) V[j] = fma(A[i + j], B[i + j], V[j]);
        return (V[0] + V[1]) + (V[2] + V[3]);
    }
    case 1: {
        double V[4];
        for (int j = 0; j < 4; j++) V[j] = A[j] * B[j];
        for (int i = 4; i < N - 1; i += 4)
            for (int j = 0; j < 4; j++) V[j] = fma(A[i + j], B[i + j], V[j]);
        return fma(A[N - 1], B[N - 1], (V[0] + V[1]) + (V[2] + V[3]));
    }
    case 2: {
        double V[2];
        for (int j = 0; j < 2; j++) V[j] = A[j] * B[j];
        for (int i = 2; i < N; i += 2)
            for (int j = 0; j < 2; j++) V[j] = fma(A[i + j], B[i + j], V[j]);
        return V[0] + V[1];
    }
    default: {
        double V[2];
        for (int j = 0; j < 2; j++) V[j] = A[j] * B[j];
        for (int i = 2; i < N - 1; i += 2)
            for (int j = 0; j < 2; j++) V[j] = fma(A[i + j], B[i + j], V[j]);
        return fma(A[N - 1], B[N - 1], V[0] + V[1]);
    }
    }
}

/* second factor of dotProductDualVec for N%4 in {0,1}: note the reference sums CD as
 * ((CD0+CD1)+CD2)+CD3, tree/phylokernelnew.h:257, 275 */
static inline double dot_tree_cd(const double *A, const double *B, int N) {
    if (N <= 4 || (N % 4) >= 2) return dot_tree(A, B, N);
    double V[4];
    for (int j = 0; j < 4; j++) V[j] = A[j] * B[j];
    if (N % 4 == 0) {
        for (int i = 4; i < N; i += 4)
            for (int j = 0; j < 4; j++) V[j] = fma(A[i + j], B[i + j], V[j]);
        return (V[0] + V[1]) + V[2] + V[3];
    }
    for (int i = 4; i < N - 1; i += 4)
        for (int j = 0; j < 4; j++) V[j] = fma(A[i + j], B[i + j], V[j]);
    return fma(A[N - 1], B[N - 1], (V[0] + V[1]) + V[2] + V[3]);
}

/* productVecMat with max tracking, tree/phylokernelnew.h:451-567:
 * X[i] = sum_x A[x]*M[i][x]; *xmax = max(*xmax, |X[i]|) */
static inline void prod_vec_mat(const double *A, const double *M, double *X, double *xmax, int N) {
    for (int i = 0; i < N; i++) {
        X[i] = dot_tree(A, M + (size_t)i * N, N);
        double a = fabs(X[i]);
        if (a > *xmax) *xmax = a;
    }
}

/* dotProduct3Vec, tree/phylokernelnew.h:737-793 */
static inline double dot3(const double *A, const double *B, const double *C, int N) {
    switch (N % 4) {
    case 0: {
        double V[4];
        for (int j = 0; j < 4; j++) V[j] = A[j] * B[j] * C[j];
        for (int i = 4; i < N; i += 4)
            for (int j = 0; j < 4; j++) V[j] = fma(A[i + j] * B[i + j], C[i + j], V[j]);
        return (V[0] + V[1]) + (V[2] + V[3]);
    }
    case 1: {
        double V[4];
        for (int j = 0; j < 4; j++) V[j] = A[j] * B[j] * C[j];
        for (int i = 4; i < N - 1; i += 4)
            for (int j = 0; j < 4; j++) V[j] = fma(A[i + j] * B[i + j], C[i + j], V[j]);
        return fma(A[N - 1] * B[N - 1], C[N - 1], (V[0] + V[1]) + (V[2] + V[3]));
    }
    case 2: {
        double V[2];
        for (int j = 0; j < 2; j++) V[j] = A[j] * B[j] * C[j];
        for (int i = 2; i < N; i += 2)
            for (int j = 0; j < 2; j++) V[j] = fma(A[i + j] * B[i + j], C[i + j], V[j]);
        return V[0] + V[1];
    }
    default: {
        double V[2];
        for (int j = 0; j < 2; j++) V[j] = A[j] * B[j] * C[j];
        for (int i = 2; i < N - 1; i += 2)
            for (int j = 0; j < 2; j++) V[j] = fma(A[i + j] * B[i + j], C[i + j], V[j]);
        return fma(A[N - 1] * B[N - 1], C[N - 1], V[0] + V[1]);
    }
    }
}

/* dotProductTriple<ADD=false>, tree/phylokernelnew.h:666-725: two interleaved
 * accumulators over the flat block of N elements */
/* REFERENCE QUIRK (kept by default, orc_set_ref_quirks(0) turns it off): the parity test is on
 * nstates while the loop runs over N = block (:2491).  For odd nstates with even ncat_mix the
 * "odd" branch visits every element in the loop AND adds element N-1 once more, so the reference's
 * lh/df/ddf count the last (category, eigen index) entry twice.  nstates 4 and 20 are unaffected. */
static int orc_ref_quirks = 1;
int orc_set_ref_quirks(int on) {
    orc_ref_quirks = on;
    return 0;
}
static inline void dot_triple(const double *A, const double *B, const double *C, const double *D,
                              double *X, double *Y, double *Z, int N, int nstates) {
    if (!orc_ref_quirks) nstates = N;
    double AD[2], BD[2], CD[2];
    for (int j = 0; j < 2; j++) {
        AD[j] = A[j] * D[j];
        BD[j] = B[j] * D[j];
        CD[j] = C[j] * D[j];
    }
    if ((nstates & 1) == 0) {
        for (int i = 2; i < N; i += 2)
            for (int j = 0; j < 2; j++) {
                AD[j] = fma(A[i + j], D[i + j], AD[j]);
                BD[j] = fma(B[i + j], D[i + j], BD[j]);
                CD[j] = fma(C[i + j], D[i + j], CD[j]);
            }
        *X = AD[0] + AD[1];
        *Y = BD[0] + BD[1];
        *Z = CD[0] + CD[1];
    } else {
        for (int i = 2; i < N - 1; i += 2)
            for (int j = 0; j < 2; j++) {
                AD[j] = fma(A[i + j], D[i + j], AD[j]);
                BD[j] = fma(B[i + j], D[i + j], BD[j]);
                CD[j] = fma(C[i + j], D[i + j], CD[j]);
            }
        *X = fma(A[N - 1], D[N - 1], AD[0] + AD[1]);
        *Y = fma(B[N - 1], D[N - 1], BD[0] + BD[1]);
        *Z = fma(C[N - 1], D[N - 1], CD[0] + CD[1]);
    }
}

/* sumVec<append=true>, tree/phylokernelnew.h:43-69 */
static inline double sum_vec_append(const double *A, double X, int N) {
    int cruft = N & 3;
    if (cruft < N) {
        double V[4] = {0, 0, 0, 0};
        int i;
        for (i = 0; i < N - cruft; i += 4) {
            V[0] += A[i];
            V[1] += A[i + 1];
            V[2] += A[i + 2];
            V[3] += A[i + 3];
        }
        V[0] += V[2];
        V[1] += V[3];
        V[0] += V[1];
        X += V[0];
        A += N - cruft;
    }
    for (int i = 0; i < cruft; i++) X += A[i];
    return X;
}

/* ------------------------------------------------------------------ context */

int orc_create(orc_ctx **out, const b200lk_dims *dims) {
    if (!out || !dims || dims->ntaxa < 2 || dims->nstates < 1 || dims->ncat < 1 || dims->nmix < 1 ||
        dims->nptn < 0 || dims->nslots < 0)
        return B200LK_ERR_ARG;
    orc_ctx *c = (orc_ctx *)calloc(1, sizeof(orc_ctx));
    c->d = *dims;
    c->ncat_mix = dims->fused_mix_rate ? dims->ncat : dims->ncat * dims->nmix; /* :1323 */
    c->block = dims->nstates * c->ncat_mix;
    c->ntot = dims->nptn + dims->n_unobserved;
    c->nthreads = 1;
    int n = dims->nstates, nmix = dims->nmix;
    size_t ntot = (size_t)c->ntot;
    c->states = (uint8_t *)calloc((size_t)dims->ntaxa * (ntot ? ntot : 1), 1);
    c->ptn_freq = (double *)calloc(ntot + 1, sizeof(double));
    c->ptn_invar = (double *)calloc(ntot + 1, sizeof(double));
    c->eval = (double *)calloc((size_t)nmix * n, sizeof(double));
    c->evec = (double *)calloc((size_t)nmix * n * n, sizeof(double));
    c->inv_evec = (double *)calloc((size_t)nmix * n * n, sizeof(double));
    c->rates = (double *)calloc(dims->ncat, sizeof(double));
    c->props = (double *)calloc(dims->ncat, sizeof(double));
    c->mixw = (double *)calloc(nmix, sizeof(double));
    c->tip_partial_lh = (double *)calloc((size_t)dims->n_state_codes * nmix * n, sizeof(double));
    c->partial = (double **)calloc(dims->nslots + 1, sizeof(double *));
    c->scale = (uint8_t **)calloc(dims->nslots + 1, sizeof(uint8_t *));
    c->theta_all = (double *)calloc(ntot * c->block + 1, sizeof(double));
    c->buffer_scale_all = (double *)calloc(ntot + 1, sizeof(double));
    c->pattern_lh = (double *)calloc(ntot + 1, sizeof(double));
    c->pattern_lh_cat = (double *)calloc(ntot * c->ncat_mix + 1, sizeof(double));
    c->robust_keep = 1.0;
    *out = c;
    return B200LK_OK;
}

int orc_destroy(orc_ctx *c) {
    if (!c) return B200LK_OK;
    for (int s = 0; s < c->d.nslots; s++) {
        free(c->partial[s]);
        free(c->scale[s]);
    }
    free(c->partial); free(c->scale); free(c->states); free(c->ptn_freq); free(c->ptn_invar);
    free(c->eval); free(c->evec); free(c->inv_evec); free(c->rates); free(c->props); free(c->mixw);
    free(c->tip_partial_lh); free(c->theta_all); free(c->buffer_scale_all); free(c->pattern_lh);
    free(c->pattern_lh_cat); free(c->s_eval); free(c->s_evec); free(c->s_inv); free(c->s_tip);
    free(c->nr_Q); free(c->nr_ev); free(c->nr_V); free(c->nr_Vi); free(c->nr_freq); free(c->nr_tipvec);
    free(c->const_df); free(c->const_ddf);
    free(c);
    return B200LK_OK;
}

int orc_set_threads(orc_ctx *c, int nthreads) {
    c->nthreads = nthreads > 0 ? nthreads : 1;
    return B200LK_OK;
}

static int ensure_slot(orc_ctx *c, int s) {
    if (s < 0 || s >= c->d.nslots) return B200LK_ERR_ARG;
    if (!c->partial[s]) {
        c->partial[s] = (double *)calloc((size_t)c->ntot * c->block + 1, sizeof(double));
        c->scale[s] = (uint8_t *)calloc((size_t)c->ntot * (c->d.safe_numeric ? c->ncat_mix : 1) + 1, 1);
        if (!c->partial[s] || !c->scale[s]) return B200LK_ERR_NOMEM;
    }
    return B200LK_OK;
}

int orc_upload_tips(orc_ctx *c, const uint8_t *states, const uint8_t *unobserved, const double *ptn_freq) {
    int64_t nptn = c->d.nptn, nun = c->d.n_unobserved;
    for (int t = 0; t < c->d.ntaxa; t++) {
        memcpy(c->states + (size_t)t * c->ntot, states + (size_t)t * nptn, (size_t)nptn);
        if (nun) memcpy(c->states + (size_t)t * c->ntot + nptn, unobserved + (size_t)t * nun, (size_t)nun);
    }
    memcpy(c->ptn_freq, ptn_freq, (size_t)nptn * sizeof(double));
    c->nsite = 0.0;
    for (int64_t p = 0; p < nptn; p++) c->nsite += ptn_freq[p];
    return B200LK_OK;
}

int orc_set_ptn_invar(orc_ctx *c, const double *ptn_invar) {
    memcpy(c->ptn_invar, ptn_invar, (size_t)c->ntot * sizeof(double));
    return B200LK_OK;
}

int orc_set_model(orc_ctx *c, const double *eval, const double *evec, const double *inv_evec,
                  const double *rates, const double *props, const double *mixw, const double *tip_partial_lh) {
    if (c->d.site_specific != B200LK_MODEL_REV) return B200LK_ERR_STATE;
    int n = c->d.nstates, nmix = c->d.nmix;
    memcpy(c->eval, eval, sizeof(double) * nmix * n);
    memcpy(c->evec, evec, sizeof(double) * nmix * n * n);
    memcpy(c->inv_evec, inv_evec, sizeof(double) * nmix * n * n);
    memcpy(c->rates, rates, sizeof(double) * c->d.ncat);
    memcpy(c->props, props, sizeof(double) * c->d.ncat);
    memcpy(c->mixw, mixw, sizeof(double) * nmix);
    memcpy(c->tip_partial_lh, tip_partial_lh, sizeof(double) * c->d.n_state_codes * nmix * n);
    c->theta_computed = 0;
    return B200LK_OK;
}

/* Model from exchangeabilities and frequencies: ModelMarkov::decomposeRateMatrix for reversible models
 * (model/modelmarkov.cpp:1386-1632; eigen_oracle.c) followed by computeTipPartialLikelihood
 * (tree/phylotreesse.cpp:243-375): tip_partial_lh[s][m][i] = sum_x inv_evec[m][i][x] * tipvec[s][x].
 * exch [nmix][n(n-1)/2], freqs [nmix][n], total_num_subst [nmix] or NULL (1.0), tipvec [n_state_codes][n]. */
int orc_eigen_rev(int n, const double *exch, const double *freqs, double total_num_subst, double *eval, double *evec,
                  double *inv_evec);
int orc_set_model_rates(orc_ctx *c, const double *exch, const double *freqs, const double *total_num_subst,
                        const double *tipvec, const double *rates, const double *props, const double *mixw) {
    int n = c->d.nstates, nmix = c->d.nmix, codes = c->d.n_state_codes, ne = n * (n - 1) / 2;
    double *eval = malloc(sizeof(double) * nmix * n), *evec = malloc(sizeof(double) * nmix * n * n);
    double *inv = malloc(sizeof(double) * nmix * n * n), *tip = malloc(sizeof(double) * codes * nmix * n);
    int rc = B200LK_OK;
    for (int m = 0; m < nmix; m++)
        if (orc_eigen_rev(n, exch + (size_t)m * ne, freqs + (size_t)m * n, total_num_subst ? total_num_subst[m] : 1.0,
                          eval + m * n, evec + (size_t)m * n * n, inv + (size_t)m * n * n))
            rc = B200LK_ERR_ARG;
    for (int s = 0; s < codes; s++)
        for (int m = 0; m < nmix; m++)
            for (int i = 0; i < n; i++) {
                double v = 0.0;
                for (int x = 0; x < n; x++) v += inv[(size_t)m * n * n + i * n + x] * tipvec[s * n + x];
                tip[((size_t)s * nmix + m) * n + i] = v;
            }
    if (rc == B200LK_OK) rc = orc_set_model(c, eval, evec, inv, rates, props, mixw, tip);
    free(eval); free(evec); free(inv); free(tip);
    return rc;
}

/* the eigen-system a context currently holds (tests) */
int orc_fetch_model(orc_ctx *c, double *eval, double *evec, double *inv_evec) {
    int n = c->d.nstates, nmix = c->d.nmix;
    memcpy(eval, c->eval, sizeof(double) * nmix * n);
    memcpy(evec, c->evec, sizeof(double) * nmix * n * n);
    memcpy(inv_evec, c->inv_evec, sizeof(double) * nmix * n * n);
    return B200LK_OK;
}

/* Site-specific models: what a ModelSet hands the SITE_MODEL kernels (model/modelset.cpp:129-200, de-interleaved)
 * and computeTipPartialLikelihood's site-specific branch (tree/phylotreesse.cpp:257-340):
 *   tip[t][p][i] = inv_p[i][state]            for a plain state,
 *                = sum_x inv_p[i][x]          for STATE_UNKNOWN,
 *                = sum over the ambiguity set for the other codes (ascending x)
 * -- all three are sum_x inv_p[i][x] * tipvec[state][x] over the non-zero indicator entries. */
int orc_set_site_models(orc_ctx *c, const double *eval, const double *evec, const double *inv_evec, const double *rates,
                        const double *props, const double *tipvec) {
    const int n = c->d.nstates;
    const int64_t nptn = c->d.nptn;
    if (c->d.site_specific != B200LK_MODEL_SITE || c->d.nmix != 1 || c->d.n_unobserved != 0) return B200LK_ERR_ARG;
    free(c->s_eval); free(c->s_evec); free(c->s_inv); free(c->s_tip);
    c->s_eval = malloc(sizeof(double) * nptn * n);
    c->s_evec = malloc(sizeof(double) * nptn * n * n);
    c->s_inv = malloc(sizeof(double) * nptn * n * n);
    c->s_tip = malloc(sizeof(double) * (size_t)c->d.ntaxa * nptn * n);
    memcpy(c->s_eval, eval, sizeof(double) * nptn * n);
    memcpy(c->s_evec, evec, sizeof(double) * nptn * n * n);
    memcpy(c->s_inv, inv_evec, sizeof(double) * nptn * n * n);
    memcpy(c->rates, rates, sizeof(double) * c->d.ncat);
    memcpy(c->props, props, sizeof(double) * c->d.ncat);
    c->mixw[0] = 1.0;
    for (int t = 0; t < c->d.ntaxa; t++)
        for (int64_t p = 0; p < nptn; p++) {
            const double *tv = tipvec + (size_t)c->states[(size_t)t * c->ntot + p] * n;
            const double *inv = c->s_inv + (size_t)p * n * n;
            double *out = c->s_tip + ((size_t)t * nptn + p) * n;
            for (int i = 0; i < n; i++) {
                double v = 0.0;
                for (int x = 0; x < n; x++)
                    if (tv[x] != 0.0) v += inv[i * n + x];
                out[i] = v;
            }
        }
    c->site = 1;
    c->theta_computed = 0;
    return B200LK_OK;
}

/* dotProductExp, tree/phylokernelnew.h:808-842 */
static inline double dot_exp3(const double *A, const double *B, const double *C, double D, int N) {
    double X = exp(A[0] * D) * B[0] * C[0];
    for (int i = 1; i < N; i++) X = fma(exp(A[i] * D), B[i] * C[i], X);
    return X;
}
static inline double dot_exp2(const double *A, const double *B, double D, int N) {
    double X = exp(A[0] * D) * B[0];
    for (int i = 1; i < N; i++) X = fma(exp(A[i] * D), B[i], X);
    return X;
}

static inline void rescale(double *v, int count);

/* SITE_MODEL branches of computePartialLikelihoodGenericSIMD: tip-tip :1653-1680, tip-internal :1772-1800,
 * internal-internal :1952-1968 (scaling as in the ordinary branches).  Bifurcating nodes only. */
static void update_node_site(orc_ctx *c, const b200lk_node *nd, const b200lk_child *ch, int64_t lo, int64_t hi) {
    const int n = c->d.nstates, ncat = c->d.ncat, block = c->block, safe = c->d.safe_numeric;
    const int64_t nptn = c->d.nptn;
    double *out = c->partial[nd->out_slot];
    uint8_t *oscale = c->scale[nd->out_slot];
    int il = 0, ir = 1;
    if (ch[0].kind != B200LK_CHILD_LEAF && ch[1].kind == B200LK_CHILD_LEAF) { il = 1; ir = 0; }
    const b200lk_child *L = &ch[il], *R = &ch[ir];
    const int lleaf = L->kind == B200LK_CHILD_LEAF, rleaf = R->kind == B200LK_CHILD_LEAF;
    double len_left[ncat], len_right[ncat], expl[n], expr[n], tmp[n];
    for (int k = 0; k < ncat; k++) { /* :1362-1374 */
        len_left[k] = c->rates[k] * L->length;
        len_right[k] = c->rates[k] * R->length;
    }
    for (int64_t p = lo; p < hi; p++) {
        const double *eval = c->s_eval + (size_t)p * n, *evec = c->s_evec + (size_t)p * n * n;
        const double *inv = c->s_inv + (size_t)p * n * n;
        uint8_t *sd = safe ? oscale + p * ncat : oscale + p;
        double *po = out + (size_t)p * block;
        const double *vl = lleaf ? c->s_tip + ((size_t)L->id * nptn + p) * n : c->partial[L->id] + (size_t)p * block;
        const double *vr = rleaf ? c->s_tip + ((size_t)R->id * nptn + p) * n : c->partial[R->id] + (size_t)p * block;
        const uint8_t *sl = lleaf ? NULL : c->scale[L->id], *sr = rleaf ? NULL : c->scale[R->id];
        if (lleaf && rleaf) memset(sd, 0, safe ? ncat : 1);
        else if (lleaf) memcpy(sd, safe ? sr + p * ncat : sr + p, safe ? ncat : 1);
        else if (!safe) sd[0] = sl[p] + sr[p];
        double lh_max = 0.0;
        for (int k = 0; k < ncat; k++) {
            if (safe) {
                lh_max = 0.0;
                if (!lleaf) sd[k] = sl[p * ncat + k] + sr[p * ncat + k];
            }
            for (int i = 0; i < n; i++) {
                expl[i] = exp(eval[i] * len_left[k]) * (lleaf ? vl[i] : vl[k * n + i]);
                expr[i] = exp(eval[i] * len_right[k]) * (rleaf ? vr[i] : vr[k * n + i]);
            }
            for (int x = 0; x < n; x++) /* dotProductDualVec */
                tmp[x] = dot_tree(evec + (size_t)x * n, expl, n) * dot_tree_cd(evec + (size_t)x * n, expr, n);
            prod_vec_mat(tmp, inv, po + k * n, &lh_max, n);
            if (safe && !(lleaf && rleaf) && lh_max < SCALING_THRESHOLD && c->ptn_invar[p] == 0.0) {
                rescale(po + k * n, n);
                sd[k] += 1;
            }
        }
        if (!safe && !(lleaf && rleaf) && lh_max < SCALING_THRESHOLD && c->ptn_invar[p] == 0.0) {
            rescale(po, block);
            sd[0] += 1;
        }
    }
}

/* --------------------------------------------------------- computePartialInfo */
/* tree/phylokernelnew.h:1062-1108 (reversible, non-vectorised branch):
 * echild[c][x][i] = evec[m][x][i] * exp(eval[m][i] * rate[c%ncat] * len)
 * leaf[s][c][x]   = sum_i echild[c][x][i] * tip_partial_lh[s][m][i]; row s = STATE_UNKNOWN is 1.0 */
static void partial_info(const orc_ctx *c, double len, const double *lens, int is_leaf, double *echild, double *leaf) {
    int n = c->d.nstates, ncat = c->d.ncat, ncm = c->ncat_mix, block = c->block;
    int denom = c->d.fused_mix_rate ? 1 : ncat;
    int tip_block = n * c->d.nmix;
    double expchild[n];
    for (int cc = 0; cc < ncm; cc++) {
        int m = cc / denom;
        /* child->getLength(cat_id[c]): one length per category under heterotachy (:1017, 1069) */
        double len_child = c->rates[cc % ncat] * (lens ? lens[cc % ncat] : len);
        const double *eval = c->eval + m * n, *evec = c->evec + (size_t)m * n * n;
        for (int i = 0; i < n; i++) expchild[i] = exp(eval[i] * len_child);
        double *e = echild + (size_t)cc * n * n;
        for (int x = 0; x < n; x++)
            for (int i = 0; i < n; i++) e[x * n + i] = evec[x * n + i] * expchild[i];
    }
    if (!is_leaf) return;
    int su = c->d.n_state_codes - 1;
    for (int s = 0; s <= su; s++) {
        for (int cc = 0; cc < ncm; cc++) {
            int m = cc / denom;
            const double *tip = c->tip_partial_lh + (size_t)s * tip_block + m * n;
            const double *e = echild + (size_t)cc * n * n;
            for (int x = 0; x < n; x++) {
                double v = e[x * n] * tip[0];
                for (int i = 1; i < n; i++) v += e[x * n + i] * tip[i];
                leaf[(size_t)s * block + cc * n + x] = v;
            }
        }
    }
    for (int x = 0; x < block; x++) leaf[(size_t)su * block + x] = 1.0;
}

/* ------------------------------------------------- computePartialLikelihood */

static inline int tip_state(const orc_ctx *c, int taxon, int64_t p) {
    return c->states[(size_t)taxon * c->ntot + p]; /* :1498-1512 with V = 1 */
}

static inline void rescale(double *v, int count) { /* ldexp(x, 256), :864, 878 */
    for (int i = 0; i < count; i++) v[i] = ldexp(v[i], SCALING_THRESHOLD_EXP);
}

/* one node, patterns [lo, hi).  Follows computePartialLikelihoodGenericSIMD,
 * tree/phylokernelnew.h:1291-2039 (non-SITE_MODEL branches). */
static void update_node(orc_ctx *c, const b200lk_node *nd, const b200lk_child *ch, double **echild,
                        double **leaf, int64_t lo, int64_t hi) {
    const int n = c->d.nstates, ncm = c->ncat_mix, block = c->block, safe = c->d.safe_numeric;
    const int denom = c->d.fused_mix_rate ? 1 : c->d.ncat;
    double *out = c->partial[nd->out_slot];
    uint8_t *oscale = c->scale[nd->out_slot];
    double tmp[n];

    if (nd->nchild > 2) {
        /* multifurcating node, :1407-1631 */
        double all[block];
        for (int64_t p = lo; p < hi; p++) {
            for (int i = 0; i < block; i++) all[i] = 1.0;
            uint8_t *sd = safe ? oscale + p * ncm : oscale + p;
            memset(sd, 0, safe ? ncm : 1);
            for (int k = 0; k < nd->nchild; k++) {
                const b200lk_child *cd = &ch[k];
                if (cd->kind == B200LK_CHILD_LEAF) {
                    const double *lf = leaf[k] + (size_t)block * tip_state(c, cd->id, p);
                    for (int i = 0; i < block; i++) all[i] *= lf[i]; /* :1520-1524 */
                } else {
                    const double *pc = c->partial[cd->id] + (size_t)p * block;
                    const uint8_t *sc = c->scale[cd->id];
                    if (!safe) sd[0] += sc[p]; /* :1530-1533 */
                    const double *e = echild[k];
                    for (int cc = 0; cc < ncm; cc++) {
                        if (safe) sd[cc] += sc[p * ncm + cc]; /* :1537-1540 */
                        for (int x = 0; x < n; x++) { /* :1542-1550 */
                            double v = e[0] * pc[0];
                            for (int i = 1; i < n; i++) v = fma(e[i], pc[i], v);
                            e += n;
                            all[cc * n + x] *= v;
                        }
                        pc += n;
                    }
                }
                /* rescale the running product after every child, :1558-1599 */
                if (safe) {
                    for (int cc = 0; cc < ncm; cc++) {
                        double mx = 0.0;
                        for (int x = 0; x < n; x++) mx = fmax(mx, fabs(all[cc * n + x]));
                        if (mx < SCALING_THRESHOLD && c->ptn_invar[p] == 0.0) {
                            rescale(all + cc * n, n);
                            sd[cc] += 1;
                        }
                    }
                } else {
                    double mx = 0.0;
                    for (int x = 0; x < block; x++) mx = fmax(mx, fabs(all[x]));
                    if (mx < SCALING_THRESHOLD && c->ptn_invar[p] == 0.0) {
                        rescale(all, block);
                        sd[0] += 1;
                    }
                }
            }
            double lh_max = 0.0; /* :1605-1627, max is computed but not acted upon */
            for (int cc = 0; cc < ncm; cc++)
                prod_vec_mat(all + cc * n, c->inv_evec + (size_t)(cc / denom) * n * n,
                             out + (size_t)p * block + cc * n, &lh_max, n);
        }
        return;
    }

    /* bifurcating: order so that a leaf child is "left", :1395-1405 */
    int il = 0, ir = 1;
    if (ch[0].kind != B200LK_CHILD_LEAF && ch[1].kind == B200LK_CHILD_LEAF) {
        il = 1;
        ir = 0;
    }
    const b200lk_child *L = &ch[il], *R = &ch[ir];

    if (L->kind == B200LK_CHILD_LEAF && R->kind == B200LK_CHILD_LEAF) {
        /* tip-tip, :1632-1746: scale_num zeroed, never rescaled */
        for (int64_t p = lo; p < hi; p++) {
            memset(safe ? oscale + p * ncm : oscale + p, 0, safe ? ncm : 1);
            const double *vl = leaf[il] + (size_t)block * tip_state(c, L->id, p);
            const double *vr = leaf[ir] + (size_t)block * tip_state(c, R->id, p);
            double lh_max = 0.0;
            for (int cc = 0; cc < ncm; cc++) {
                for (int x = 0; x < n; x++) tmp[x] = vl[cc * n + x] * vr[cc * n + x];
                prod_vec_mat(tmp, c->inv_evec + (size_t)(cc / denom) * n * n,
                             out + (size_t)p * block + cc * n, &lh_max, n);
            }
        }
    } else if (L->kind == B200LK_CHILD_LEAF) {
        /* tip-internal, :1748-1905 */
        const double *pr_all = c->partial[R->id];
        const uint8_t *sr = c->scale[R->id];
        for (int64_t p = lo; p < hi; p++) {
            uint8_t *sd = safe ? oscale + p * ncm : oscale + p;
            memcpy(sd, safe ? sr + p * ncm : sr + p, safe ? ncm : 1); /* :1753-1756 */
            const double *vl = leaf[il] + (size_t)block * tip_state(c, L->id, p);
            const double *pr = pr_all + (size_t)p * block;
            double *po = out + (size_t)p * block;
            const double *er = echild[ir];
            double lh_max = 0.0;
            for (int cc = 0; cc < ncm; cc++) {
                if (safe) lh_max = 0.0; /* :1846-1847 */
                for (int x = 0; x < n; x++) {
                    double vright = dot_vec(er, pr + cc * n, n); /* :1853-1858 */
                    er += n;
                    tmp[x] = vl[cc * n + x] * vright;
                }
                prod_vec_mat(tmp, c->inv_evec + (size_t)(cc / denom) * n * n, po + cc * n, &lh_max, n);
                if (safe && lh_max < SCALING_THRESHOLD && c->ptn_invar[p] == 0.0) { /* :1868-1882 */
                    rescale(po + cc * n, n);
                    sd[cc] += 1;
                }
            }
            if (!safe && lh_max < SCALING_THRESHOLD && c->ptn_invar[p] == 0.0) { /* :1889-1903 */
                rescale(po, block);
                sd[0] += 1;
            }
        }
    } else {
        /* internal-internal, :1907-2032 */
        const double *pl_all = c->partial[L->id], *pr_all = c->partial[R->id];
        const uint8_t *sl = c->scale[L->id], *sr = c->scale[R->id];
        for (int64_t p = lo; p < hi; p++) {
            uint8_t *sd = safe ? oscale + p * ncm : oscale + p;
            if (!safe) sd[0] = sl[p] + sr[p]; /* :1929-1931 */
            const double *pl = pl_all + (size_t)p * block, *pr = pr_all + (size_t)p * block;
            double *po = out + (size_t)p * block;
            const double *el = echild[il], *er = echild[ir];
            double lh_max = 0.0;
            for (int cc = 0; cc < ncm; cc++) {
                if (safe) {
                    lh_max = 0.0;
                    sd[cc] = sl[p * ncm + cc] + sr[p * ncm + cc]; /* :1946-1950 */
                }
                for (int x = 0; x < n; x++) { /* dotProductDualVec, :1975-1983 */
                    tmp[x] = dot_tree(el, pl + cc * n, n) * dot_tree_cd(er, pr + cc * n, n);
                    el += n;
                    er += n;
                }
                prod_vec_mat(tmp, c->inv_evec + (size_t)(cc / denom) * n * n, po + cc * n, &lh_max, n);
                if (safe && lh_max < SCALING_THRESHOLD && c->ptn_invar[p] == 0.0) { /* :1994-2005 */
                    rescale(po + cc * n, n);
                    sd[cc] += 1;
                }
            }
            if (!safe && lh_max < SCALING_THRESHOLD && c->ptn_invar[p] == 0.0) { /* :2015-2029 */
                rescale(po, block);
                sd[0] += 1;
            }
        }
    }
}

/* computeTraversalInfo(compute_partial_lh = true), tree/phylokernelnew.h:1150-1281:
 * tables for every entry, then every entry over every pattern range. */
#include "nonrev_oracle.inc"

static int update_partials_impl(orc_ctx *c, const b200lk_node *nodes, int32_t nnodes, const b200lk_child *children,
                                int32_t nchildren, const double *child_len) {
    const int n = c->d.nstates, block = c->block;
    if (nnodes <= 0) return B200LK_OK;
    if (child_len && (c->d.site_specific != B200LK_MODEL_REV)) return B200LK_ERR_UNSUPPORTED;
    if (c->d.site_specific == B200LK_MODEL_NONREV)
        return c->nonrev ? nr_update_partials(c, nodes, nnodes, children, nchildren) : B200LK_ERR_STATE;
    if (c->site) { /* no P-factor tables: every pattern has its own eigen-system */
        for (int i = 0; i < nnodes; i++) {
            if (nodes[i].nchild != 2) return B200LK_ERR_UNSUPPORTED;
            int rc = ensure_slot(c, nodes[i].out_slot);
            if (rc) return rc;
        }
        int64_t ntot_s = c->ntot;
        int nthr_s = c->nthreads;
#ifdef _OPENMP
#pragma omp parallel for schedule(static) num_threads(nthr_s)
#endif
        for (int t = 0; t < nthr_s; t++) {
            int64_t lo = ntot_s * t / nthr_s, hi = ntot_s * (t + 1) / nthr_s;
            for (int i = 0; i < nnodes; i++) update_node_site(c, &nodes[i], children + nodes[i].child_begin, lo, hi);
        }
        c->updates += (int64_t)nnodes * ntot_s;
        c->theta_computed = 0;
        return B200LK_OK;
    }
    double **echild = (double **)calloc(nchildren, sizeof(double *));
    double **leaf = (double **)calloc(nchildren, sizeof(double *));
    for (int i = 0; i < nnodes; i++) {
        int rc = ensure_slot(c, nodes[i].out_slot);
        if (rc) return rc;
        for (int k = 0; k < nodes[i].nchild; k++) {
            int j = nodes[i].child_begin + k;
            echild[j] = (double *)malloc(sizeof(double) * block * n);
            int is_leaf = children[j].kind == B200LK_CHILD_LEAF;
            if (is_leaf) leaf[j] = (double *)malloc(sizeof(double) * block * c->d.n_state_codes);
            partial_info(c, children[j].length, child_len ? child_len + (size_t)j * c->d.ncat : NULL, is_leaf,
                         echild[j], leaf[j]);
        }
    }
    int64_t ntot = c->ntot;
    int nthr = c->nthreads;
#ifdef _OPENMP
#pragma omp parallel for schedule(static) num_threads(nthr)
#endif
    for (int t = 0; t < nthr; t++) {
        int64_t lo = ntot * t / nthr, hi = ntot * (t + 1) / nthr;
        for (int i = 0; i < nnodes; i++) /* nodes inner, pattern range outer, :1273-1277 */
            update_node(c, &nodes[i], children + nodes[i].child_begin, echild + nodes[i].child_begin,
                        leaf + nodes[i].child_begin, lo, hi);
    }
    for (int j = 0; j < nchildren; j++) {
        free(echild[j]);
        free(leaf[j]);
    }
    free(echild);
    free(leaf);
    c->updates += (int64_t)nnodes * ntot;
    c->theta_computed = 0;
    return B200LK_OK;
}

int orc_update_partials(orc_ctx *c, const b200lk_node *nodes, int32_t nnodes, const b200lk_child *children,
                        int32_t nchildren) {
    return update_partials_impl(c, nodes, nnodes, children, nchildren, NULL);
}
/* heterotachy: child_len[j][k] = children[j]->getLength(k), k < ncat (tree/phylokernelnew.h:1017, 1069) */
int orc_update_partials_mixlen(orc_ctx *c, const b200lk_node *nodes, int32_t nnodes, const b200lk_child *children,
                               int32_t nchildren, const double *child_len) {
    if (!child_len) return B200LK_ERR_ARG;
    return update_partials_impl(c, nodes, nnodes, children, nchildren, child_len);
}

/* ------------------------------------------------------- branch-level steps */

/* val[c][i] = exp(eval[m][i]*rate*len) * prop*mixw; optionally val1 = cof*val, val2 = cof*val1
 * (tree/phylokernelnew.h:2342-2357, 2719-2730, 3255-3266) */
static void branch_vals_l(const orc_ctx *c, double len_, const double *lens, double *val0, double *val1, double *val2) {
    int n = c->d.nstates, ncat = c->d.ncat, ncm = c->ncat_mix;
    int denom = c->d.fused_mix_rate ? 1 : ncat;
    for (int cc = 0; cc < ncm; cc++) {
        int m = cc / denom, k = cc % ncat;
        double prop = c->props[k] * c->mixw[m];
        double len = lens ? lens[k] : len_; /* dad_branch->getLength(mycat), :2325, 2348, 2711, 2725, 3248 */
        for (int i = 0; i < n; i++) {
            double cof = c->eval[m * n + i] * c->rates[k];
            double val = exp(cof * len) * prop;
            val0[cc * n + i] = val;
            if (val1) {
                double v1 = cof * val;
                val1[cc * n + i] = v1;
                val2[cc * n + i] = cof * v1;
            }
        }
    }
}

static void branch_vals(const orc_ctx *c, double len, double *val0, double *val1, double *val2) {
    branch_vals_l(c, len, NULL, val0, val1, val2);
}

static int check_ends(const orc_ctx *c, b200lk_end dad, b200lk_end node) {
    if (node.kind != B200LK_CHILD_INTERNAL || node.id < 0 || node.id >= c->d.nslots || !c->partial[node.id])
        return B200LK_ERR_ARG;
    if (dad.kind == B200LK_CHILD_LEAF) return (dad.id >= 0 && dad.id < c->d.ntaxa) ? 0 : B200LK_ERR_ARG;
    if (dad.id < 0 || dad.id >= c->d.nslots || !c->partial[dad.id]) return B200LK_ERR_ARG;
    return 0;
}

/* computeLikelihoodBranchGenericSIMD, tree/phylokernelnew.h:2643-3185.
 * `node` = slot of dad_branch->partial_lh (subtree on the node side);
 * `dad`  = the leaf dad (tip-internal, :2740-2941) or the slot of node_branch->partial_lh (:2942-3074). */
/* Mark Holder's correction for missing data applied to the per-pattern and total lnL (:3137-3159, 3349-3368):
 * const_lh[p] = sum over the n constant states of the unobserved pattern that copies p's gaps */
static double holder_fix_lnl(orc_ctx *c) {
    const int n = c->d.nstates;
    const int64_t nptn = c->d.nptn, step = c->d.n_unobserved / n;
    const double *const_lh = c->pattern_lh + nptn;
    double sum_corr = 0.0;
    for (int64_t p = 0; p < nptn; p++) {
        double cl = const_lh[p];
        for (int s = 1; s < n; s++) cl += const_lh[(size_t)s * step + p];
        double prob_variant = log(1.0 - cl);
        c->pattern_lh[p] -= prob_variant;
        sum_corr += prob_variant * c->ptn_freq[p];
    }
    return sum_corr;
}

static int cmp_double(const void *a, const void *b) {
    double x = *(const double *)a, y = *(const double *)b;
    return (x > y) - (x < y);
}
/* params->robust_phy_keep (fraction of best sites summed; 1.0 = off) and params->robust_median: read by
 * computeLikelihoodBranch only (tree/phylokernelnew.h:3096-3135) */
int orc_set_robust(orc_ctx *c, double keep, int32_t median) {
    if (!(keep > 0.0) || keep > 1.0) return B200LK_ERR_ARG;
    c->robust_keep = keep;
    c->robust_median = median != 0;
    return B200LK_OK;
}

static int branch_lnl_impl(orc_ctx *c, b200lk_end dad, b200lk_end node, double length, const double *lens,
                           double *lnl) {
    if (c->d.site_specific == B200LK_MODEL_NONREV)
        return c->nonrev ? nr_branch_lnl(c, dad, node, length, lnl) : B200LK_ERR_STATE;
    if (lens && c->site) return B200LK_ERR_UNSUPPORTED;
    int rc = check_ends(c, dad, node);
    if (rc) return rc;
    const int n = c->d.nstates, ncm = c->ncat_mix, block = c->block, safe = c->d.safe_numeric;
    const int denom = c->d.fused_mix_rate ? 1 : c->d.ncat;
    const int64_t nptn = c->d.nptn, ntot = c->ntot;
    double val[block];
    if (!c->site) branch_vals_l(c, length, lens, val, NULL, NULL);
    const double *pdad_all = c->partial[node.id]; /* "partial_lh_dad" = dad_branch->partial_lh */
    const uint8_t *sdad = c->scale[node.id];
    double *lh_node_tab = NULL;
    if (dad.kind == B200LK_CHILD_LEAF && !c->site) { /* :2757-2785: partial_lh_node[s][c][i] = val[c][i]*tip[s][m][i] */
        int tip_block = n * c->d.nmix;
        lh_node_tab = (double *)malloc(sizeof(double) * block * c->d.n_state_codes);
        for (int s = 0; s < c->d.n_state_codes; s++)
            for (int cc = 0; cc < ncm; cc++)
                for (int i = 0; i < n; i++)
                    lh_node_tab[(size_t)s * block + cc * n + i] =
                        val[cc * n + i] * c->tip_partial_lh[(size_t)s * tip_block + (cc / denom) * n + i];
    }
    double tree_lh = 0.0, prob_const = 0.0;
    int nthr = c->nthreads;
#ifdef _OPENMP
#pragma omp parallel for schedule(static) num_threads(nthr) reduction(+ : tree_lh, prob_const)
#endif
    for (int t = 0; t < nthr; t++) {
        int64_t lo = ntot * t / nthr, hi = ntot * (t + 1) / nthr;
        for (int64_t p = lo; p < hi; p++) {
            double *lh_cat = c->pattern_lh_cat + (size_t)p * ncm;
            const double *pdad = pdad_all + (size_t)p * block;
            double lh_ptn = 0.0, min_scale;
            uint8_t sum_scale[ncm];
            if (c->site) { /* :2815-2831, 2969-2985: dotProductExp per category, then * cat_prop */
                const double *ev = c->s_eval + (size_t)p * n;
                const int leafdad = dad.kind == B200LK_CHILD_LEAF;
                const double *other = leafdad ? c->s_tip + ((size_t)dad.id * nptn + p) * n
                                              : c->partial[dad.id] + (size_t)p * block;
                const uint8_t *snode = leafdad ? NULL : c->scale[dad.id];
                for (int cc = 0; cc < ncm; cc++) {
                    lh_cat[cc] = dot_exp3(ev, leafdad ? other : other + cc * n, pdad + cc * n, c->rates[cc] * length, n);
                    lh_cat[cc] *= c->props[cc];
                    if (!safe) lh_ptn += lh_cat[cc];
                }
                if (safe)
                    for (int cc = 0; cc < ncm; cc++)
                        sum_scale[cc] = (uint8_t)(sdad[p * ncm + cc] + (leafdad ? 0 : snode[p * ncm + cc]));
                else
                    min_scale = (double)((int)sdad[p] + (leafdad ? 0 : (int)snode[p]));
            } else if (dad.kind == B200LK_CHILD_LEAF) {
                const double *lh_node = lh_node_tab + (size_t)block * tip_state(c, dad.id, p);
                for (int cc = 0; cc < ncm; cc++) { /* :2857-2867 */
                    lh_cat[cc] = dot_vec(lh_node + cc * n, pdad + cc * n, n);
                    if (!safe) lh_ptn += lh_cat[cc];
                }
                if (safe)
                    for (int cc = 0; cc < ncm; cc++) sum_scale[cc] = sdad[p * ncm + cc];
                else
                    min_scale = sdad[p];
            } else {
                const double *pnode = c->partial[dad.id] + (size_t)p * block;
                const uint8_t *snode = c->scale[dad.id];
                for (int cc = 0; cc < ncm; cc++) { /* :2987-2998 */
                    lh_cat[cc] = dot3(val + cc * n, pnode + cc * n, pdad + cc * n, n);
                    if (!safe) lh_ptn += lh_cat[cc];
                }
                if (safe)
                    for (int cc = 0; cc < ncm; cc++)
                        sum_scale[cc] = (uint8_t)(sdad[p * ncm + cc] + snode[p * ncm + cc]);
                else
                    min_scale = (double)((int)sdad[p] + (int)snode[p]); /* :3031-3033, int promotion */
            }
            if (safe) { /* :2873-2897, 3004-3029 */
                uint8_t ms = sum_scale[0];
                for (int cc = 1; cc < ncm; cc++)
                    if (sum_scale[cc] < ms) ms = sum_scale[cc];
                min_scale = ms;
                for (int cc = 0; cc < ncm; cc++) {
                    if (sum_scale[cc] == ms + 1)
                        lh_cat[cc] *= SCALING_THRESHOLD;
                    else if (sum_scale[cc] > ms + 1)
                        lh_cat[cc] = 0.0;
                }
                lh_ptn = sum_vec_append(lh_cat, lh_ptn, ncm);
            }
            min_scale *= LOG_SCALING_THRESHOLD;
            lh_ptn = fabs(lh_ptn) + c->ptn_invar[p]; /* :2907, 3038 */
            if (p < nptn) {
                lh_ptn = log(lh_ptn) + min_scale;
                c->pattern_lh[p] = lh_ptn;
                tree_lh = fma(lh_ptn, c->ptn_freq[p], tree_lh);
            } else { /* +ASC unobserved patterns, :2912-2930 (Lewis) */
                if (min_scale != 0.0) lh_ptn *= SCALING_THRESHOLD;
                c->pattern_lh[p] = lh_ptn;
                prob_const += lh_ptn;
            }
        }
    }
    free(lh_node_tab);
    int warn = 0;
    if (!isfinite(tree_lh)) { /* :3077-3094 */
        warn |= B200LK_WARN_LNL_NONFINITE;
        tree_lh = 0.0;
        for (int64_t p = 0; p < nptn; p++) {
            if (!isfinite(c->pattern_lh[p])) c->pattern_lh[p] = LOG_SCALING_THRESHOLD * 4;
            tree_lh += c->pattern_lh[p] * c->ptn_freq[p];
        }
    }
    if (c->robust_keep < 1.0 || c->robust_median) { /* :3096-3135: robust phylogeny, the best sites only / the median */
        if (c->d.n_unobserved > 0) return B200LK_ERR_UNSUPPORTED; /* outError("+ASC not supported with robust phylo") */
        size_t sites = (size_t)c->nsite, site = 0;
        double *site_lh = (double *)malloc(sizeof(double) * (sites + 1));
        for (int64_t p = 0; p < nptn; p++)
            for (int64_t i = 0; i < (int64_t)c->ptn_freq[p]; i++) site_lh[site++] = c->pattern_lh[p];
        qsort(site_lh, sites, sizeof(double), cmp_double);
        if (c->robust_keep < 1.0) {
            size_t sites_drop = (size_t)(int)((1.0 - c->robust_keep) * sites);
            tree_lh = 0.0;
            for (size_t i = sites_drop; i < sites; i++) tree_lh += site_lh[i];
        } else if (sites & 1) {
            tree_lh = site_lh[sites / 2];
        } else {
            tree_lh = 0.5 * (site_lh[sites / 2] + site_lh[sites / 2 - 1]);
        }
        free(site_lh);
    }
    if (c->asc_holder) { /* :3137-3159 */
        tree_lh -= holder_fix_lnl(c);
    } else if (c->d.n_unobserved > 0) { /* ASC_Lewis, :3160-3182 */
        double nsite = c->nsite; /* aln->getNSite() */
        prob_const = log(1.0 - prob_const);
        for (int64_t p = 0; p < nptn; p++) c->pattern_lh[p] -= prob_const;
        tree_lh -= nsite * prob_const;
    }
    *lnl = tree_lh;
    return warn;
}
int orc_branch_lnl(orc_ctx *c, b200lk_end dad, b200lk_end node, double length, double *lnl) {
    return branch_lnl_impl(c, dad, node, length, NULL, lnl);
}
/* heterotachy: len[k] = dad_branch->getLength(k) (:2711, 2725) */
int orc_branch_lnl_mixlen(orc_ctx *c, b200lk_end dad, b200lk_end node, const double *len, double *lnl) {
    if (!len) return B200LK_ERR_ARG;
    return branch_lnl_impl(c, dad, node, len[0], len, lnl);
}
/* Which +ASC correction the unobserved patterns stand for (model_factory->getASC(), utils/tools.h:549-555):
 * 0 = Lewis (ASC_VARIANT, the default when n_unobserved > 0), 1 = Holder for missing data (ASC_VARIANT_MISSING:
 * n_unobserved = nstates * nptn, state-major) */
int orc_set_asc(orc_ctx *c, int32_t type) {
    if (type != 0 && type != 1) return B200LK_ERR_ARG;
    if (type == 1) {
        if (c->d.n_unobserved != (int64_t)c->d.nstates * c->d.nptn || c->d.n_unobserved == 0) return B200LK_ERR_ARG;
        if (c->d.site_specific != B200LK_MODEL_REV) return B200LK_ERR_UNSUPPORTED;
        if (!c->const_df) {
            c->const_df = (double *)calloc(c->d.n_unobserved, sizeof(double));
            c->const_ddf = (double *)calloc(c->d.n_unobserved, sizeof(double));
        }
    }
    c->asc_holder = type;
    return B200LK_OK;
}

/* computeLikelihoodBufferGenericSIMD, tree/phylokernelnew.h:2050-2231 */
int orc_build_theta(orc_ctx *c, b200lk_end dad, b200lk_end node) {
    if (c->d.site_specific == B200LK_MODEL_NONREV) { /* no theta in this family: the ends are recorded */
        int rn = c->nonrev ? nr_check_ends(c, dad, node) : B200LK_ERR_STATE;
        if (rn) return rn;
        c->nr_dad = dad;
        c->nr_node = node;
        c->theta_computed = 1;
        return B200LK_OK;
    }
    int rc = check_ends(c, dad, node);
    if (rc) return rc;
    const int n = c->d.nstates, ncm = c->ncat_mix, block = c->block, safe = c->d.safe_numeric;
    const int denom = c->d.fused_mix_rate ? 1 : c->d.ncat;
    const int tip_block = n * c->d.nmix;
    const int64_t ntot = c->ntot;
    const double *pdad_all = c->partial[node.id];
    const uint8_t *sdad = c->scale[node.id];
    int nthr = c->nthreads;
#ifdef _OPENMP
#pragma omp parallel for schedule(static) num_threads(nthr)
#endif
    for (int t = 0; t < nthr; t++) {
        int64_t lo = ntot * t / nthr, hi = ntot * (t + 1) / nthr;
        for (int64_t p = lo; p < hi; p++) {
            double *theta = c->theta_all + (size_t)p * block;
            const double *pdad = pdad_all + (size_t)p * block;
            uint8_t sum_scale[ncm];
            int norm_scale;
            if (dad.kind == B200LK_CHILD_LEAF) { /* :2091-2141 */
                const double *tip = c->site ? c->s_tip + ((size_t)dad.id * c->d.nptn + p) * n
                                            : c->tip_partial_lh + (size_t)tip_block * tip_state(c, dad.id, p);
                for (int cc = 0; cc < ncm; cc++)
                    for (int i = 0; i < n; i++)
                        theta[cc * n + i] = tip[c->site ? i : (cc / denom) * n + i] * pdad[cc * n + i];
                if (safe)
                    for (int cc = 0; cc < ncm; cc++) sum_scale[cc] = sdad[p * ncm + cc];
                else
                    norm_scale = sdad[p];
            } else { /* :2183-2189 */
                const double *pnode = c->partial[dad.id] + (size_t)p * block;
                const uint8_t *snode = c->scale[dad.id];
                for (int i = 0; i < block; i++) theta[i] = pnode[i] * pdad[i];
                if (safe)
                    for (int cc = 0; cc < ncm; cc++)
                        sum_scale[cc] = (uint8_t)(sdad[p * ncm + cc] + snode[p * ncm + cc]);
                else
                    norm_scale = (int)sdad[p] + (int)snode[p]; /* :2224 */
            }
            if (safe) { /* :2142-2167, 2190-2221 */
                uint8_t ms = sum_scale[0];
                for (int cc = 1; cc < ncm; cc++)
                    if (sum_scale[cc] < ms) ms = sum_scale[cc];
                for (int cc = 0; cc < ncm; cc++) {
                    if (sum_scale[cc] == ms + 1)
                        for (int x = 0; x < n; x++) theta[cc * n + x] *= SCALING_THRESHOLD;
                    else if (sum_scale[cc] > ms + 1)
                        for (int x = 0; x < n; x++) theta[cc * n + x] = 0.0;
                }
                norm_scale = ms;
            }
            c->buffer_scale_all[p] = (double)norm_scale * LOG_SCALING_THRESHOLD; /* :2174-2175 */
        }
    }
    c->theta_computed = 1;
    return B200LK_OK;
}

/* computeLikelihoodDervGenericSIMD, tree/phylokernelnew.h:2235-2633 (joint branch-length model) */
int orc_derv(orc_ctx *c, double length, double *df, double *ddf) {
    if (!c->theta_computed) return B200LK_ERR_STATE;
    if (c->nonrev) return nr_derv(c, length, df, ddf);
    const int n = c->d.nstates, block = c->block;
    const int64_t nptn = c->d.nptn, ntot = c->ntot;
    double val0[block], val1[block], val2[block];
    if (!c->site) branch_vals(c, length, val0, val1, val2);
    double all_df = 0.0, all_ddf = 0.0, prob_const = 0.0, df_const = 0.0, ddf_const = 0.0;
    int nthr = c->nthreads;
#ifdef _OPENMP
#pragma omp parallel for schedule(static) num_threads(nthr) reduction(+ : all_df, all_ddf, prob_const, df_const, ddf_const)
#endif
    for (int t = 0; t < nthr; t++) {
        int64_t lo = ntot * t / nthr, hi = ntot * (t + 1) / nthr;
        double my_df = 0.0, my_ddf = 0.0;
        for (int64_t p = lo; p < hi; p++) {
            double lh, dfp, ddfp;
            if (c->site) { /* :2469-2487 */
                const double *ev = c->s_eval + (size_t)p * n, *theta = c->theta_all + (size_t)p * block;
                lh = dfp = ddfp = 0.0;
                for (int cc = 0; cc < c->d.ncat; cc++) {
                    double lh_cat = 0.0, df_cat = 0.0, ddf_cat = 0.0;
                    for (int i = 0; i < n; i++) {
                        double cof = ev[i] * c->rates[cc];
                        double val = exp(cof * length) * theta[cc * n + i];
                        double v1 = cof * val;
                        lh_cat += val;
                        df_cat += v1;
                        ddf_cat = fma(cof, v1, ddf_cat);
                    }
                    lh = fma(c->props[cc], lh_cat, lh);
                    dfp = fma(c->props[cc], df_cat, dfp);
                    ddfp = fma(c->props[cc], ddf_cat, ddfp);
                }
            } else
            dot_triple(val0, val1, val2, c->theta_all + (size_t)p * block, &lh, &dfp, &ddfp, block, n); /* :2489-2491 */
            lh = fabs(lh) + c->ptn_invar[p];                                                            /* :2494 */
            if (p < nptn) { /* :2496-2505 */
                lh = 1.0 / lh;
                double df_frac = dfp * lh, ddf_frac = ddfp * lh, freq = c->ptn_freq[p];
                double tmp1 = df_frac * freq, tmp2 = ddf_frac * freq;
                my_df += tmp1;
                my_ddf += fma(-tmp1, df_frac, tmp2); /* nmul_add(tmp1, df_frac, tmp2) */
            } else { /* :2506-2534 (Lewis) */
                if (c->buffer_scale_all[p] != 0.0) {
                    lh *= SCALING_THRESHOLD;
                    dfp *= SCALING_THRESHOLD;
                    ddfp *= SCALING_THRESHOLD;
                }
                if (c->asc_holder) { /* :2522-2525 */
                    c->pattern_lh[p] = lh;
                    c->const_df[p - nptn] = dfp;
                    c->const_ddf[p - nptn] = ddfp;
                } else {
                    prob_const += lh;
                    df_const += dfp;
                    ddf_const += ddfp;
                }
            }
        }
        all_df += my_df;
        all_ddf += my_ddf;
    }
    *df = all_df;
    *ddf = all_ddf;
    if (c->asc_holder) { /* :2580-2619; like the reference, applied whether or not df is finite */
        const int64_t step = c->d.n_unobserved / n;
        const double *const_lh = c->pattern_lh + nptn;
        double sum_df = 0.0, sum_ddf = 0.0;
        for (int64_t p = 0; p < nptn; p++) {
            double cl = const_lh[p], cd = c->const_df[p], cdd = c->const_ddf[p];
            for (int s = 1; s < n; s++) {
                cl += const_lh[(size_t)s * step + p];
                cd += c->const_df[(size_t)s * step + p];
                cdd += c->const_ddf[(size_t)s * step + p];
            }
            double pc = 1.0 - cl;
            double df_frac = cd / pc, ddf_frac = cdd / pc;
            sum_df += c->ptn_freq[p] * df_frac;
            sum_ddf += c->ptn_freq[p] * (ddf_frac + df_frac * df_frac);
        }
        *df += sum_df;
        *ddf += sum_ddf;
    } else if (c->d.n_unobserved > 0 && isfinite(*df)) { /* :2620-2628 */
        double nsite = c->nsite; /* aln->getNSite() */
        prob_const = 1.0 - prob_const;
        double df_frac = df_const / prob_const, ddf_frac = ddf_const / prob_const;
        *df += nsite * df_frac;
        *ddf += nsite * (ddf_frac + df_frac * df_frac);
    }
    if (!isfinite(*df)) { /* :2573-2578, 2629-2632 */
        if (c->d.safe_numeric) *df = *ddf = 0.0;
        return B200LK_WARN_DERV_NONFINITE;
    }
    return B200LK_OK;
}

/* computeLikelihoodFromBufferGenericSIMD, tree/phylokernelnew.h:3196-3392 */
static int from_buffer_impl(orc_ctx *c, double length, const double *lens, double *lnl) {
    if (!c->theta_computed) return B200LK_ERR_STATE;
    if (c->nonrev) return nr_branch_lnl(c, c->nr_dad, c->nr_node, length, lnl); /* tree/phylokernelfma.cpp:99 */
    if (lens && c->site) return B200LK_ERR_UNSUPPORTED;
    const int block = c->block;
    const int64_t nptn = c->d.nptn, ntot = c->ntot;
    double val0[block];
    if (!c->site) branch_vals_l(c, length, lens, val0, NULL, NULL);
    double tree_lh = 0.0, prob_const = 0.0;
    int nthr = c->nthreads;
#ifdef _OPENMP
#pragma omp parallel for schedule(static) num_threads(nthr) reduction(+ : tree_lh, prob_const)
#endif
    for (int t = 0; t < nthr; t++) {
        int64_t lo = ntot * t / nthr, hi = ntot * (t + 1) / nthr;
        for (int64_t p = lo; p < hi; p++) {
            double lh;
            if (c->site) { /* :3278-3289 */
                const int ns = c->d.nstates;
                lh = 0.0;
                for (int cc = 0; cc < c->d.ncat; cc++)
                    lh = fma(dot_exp2(c->s_eval + (size_t)p * ns, c->theta_all + (size_t)p * block + cc * ns,
                                      c->rates[cc] * length, ns), c->props[cc], lh);
            } else
            lh = dot_vec(val0, c->theta_all + (size_t)p * block, block); /* :3291 */
            lh = fabs(lh) + c->ptn_invar[p];                                     /* :3295 */
            if (p < nptn) {
                lh = log(fabs(lh)) + c->buffer_scale_all[p]; /* :3300 */
                c->pattern_lh[p] = lh;
                tree_lh += fma(lh, c->ptn_freq[p], 0.0);
            } else {
                if (c->buffer_scale_all[p] != 0.0) lh *= SCALING_THRESHOLD;
                c->pattern_lh[p] = lh;
                prob_const += lh;
            }
        }
    }
    int warn = 0;
    if (!isfinite(tree_lh)) { /* :3333-3347 */
        warn |= B200LK_WARN_LNL_NONFINITE;
        tree_lh = 0.0;
        for (int64_t p = 0; p < nptn; p++) {
            if (!isfinite(c->pattern_lh[p])) c->pattern_lh[p] = LOG_SCALING_THRESHOLD * 4;
            tree_lh += c->pattern_lh[p] * c->ptn_freq[p];
        }
    }
    if (c->asc_holder) { /* :3349-3368 */
        tree_lh -= holder_fix_lnl(c);
    } else if (c->d.n_unobserved > 0) { /* :3369-3390 */
        double nsite = c->nsite; /* aln->getNSite() */
        prob_const = log(1.0 - prob_const);
        for (int64_t p = 0; p < nptn; p++) c->pattern_lh[p] -= prob_const;
        tree_lh -= nsite * prob_const;
    }
    *lnl = tree_lh;
    return warn;
}
int orc_lnl_from_buffer(orc_ctx *c, double length, double *lnl) { return from_buffer_impl(c, length, NULL, lnl); }
int orc_lnl_from_buffer_mixlen(orc_ctx *c, const double *len, double *lnl) {
    if (!len) return B200LK_ERR_ARG;
    return from_buffer_impl(c, len[0], len, lnl);
}

/* computeLikelihoodDerv*SIMD with isMixlen() (tree/phylokernelnew.h:2365-2462, 2556-2568): the gradient over the ncat
 * class lengths of one branch, the Hessian, and the log-likelihood (the reference returns it in df[nmixlen]).
 * +ASC is not supported there (ASSERT(0) at :2449). */
int orc_derv_mixlen(orc_ctx *c, const double *len, double *df, double *ddf, double *lnl) {
    if (!c->theta_computed) return B200LK_ERR_STATE;
    if (!len || !df || !ddf) return B200LK_ERR_ARG;
    if (c->site || c->nonrev || c->d.n_unobserved > 0) return B200LK_ERR_UNSUPPORTED;
    const int n = c->d.nstates, block = c->block, ncat = c->d.ncat, ncm = c->ncat_mix;
    const int64_t nptn = c->d.nptn;
    double val0[block], val1[block], val2[block];
    branch_vals_l(c, len[0], len, val0, val1, val2);
    double all_lh = 0.0;
    for (int i = 0; i < ncat; i++) df[i] = 0.0;
    for (int i = 0; i < ncat * ncat; i++) ddf[i] = 0.0;
    for (int64_t p = 0; p < nptn; p++) {
        double df_ptn[ncat], ddf_ptn[ncat], lh_ptn = 0.0;
        for (int i = 0; i < ncat; i++) df_ptn[i] = ddf_ptn[i] = 0.0;
        const double *theta = c->theta_all + (size_t)p * block;
        for (int cc = 0; cc < ncm; cc++) { /* dotProductTriple<..., ADD = true> per category, :2410-2423 */
            double l, d1, d2;
            int i = cc % ncat;
            dot_triple(val0 + cc * n, val1 + cc * n, val2 + cc * n, theta + cc * n, &l, &d1, &d2, n, n);
            lh_ptn += l;
            df_ptn[i] += d1;
            ddf_ptn[i] += d2;
        }
        lh_ptn = fabs(lh_ptn) + c->ptn_invar[p];
        double freq = c->ptn_freq[p], inv = 1.0 / lh_ptn;
        for (int i = 0; i < ncat; i++) {
            df_ptn[i] *= inv;
            ddf_ptn[i] *= inv;
            df[i] = fma(df_ptn[i], freq, df[i]);
        }
        for (int i = 0; i < ncat; i++) {
            ddf[i * ncat + i] += fma(-df_ptn[i], df_ptn[i], ddf_ptn[i]) * freq;
            for (int k = 0; k < ncat; k++)
                if (k != i) ddf[i * ncat + k] -= df_ptn[i] * df_ptn[k] * freq;
        }
        all_lh = fma(log(lh_ptn) + c->buffer_scale_all[p], freq, all_lh);
    }
    if (lnl) *lnl = all_lh;
    return B200LK_OK;
}

/* computeLikelihoodDervMixlen*SIMD (tree/phylokernelnew.h:3397-3580): df, ddf of the EM sub-problem of class
 * `cur_mixlen` (getCurMixture()): only that class's slice of theta enters, at length `len` =
 * dad_branch->getLength(cur_mixlen); the caller has put the posterior weights into ptn_freq
 * (tree/phylotreemixlen.cpp:449-481).  The reference indexes val with the mixture class alone (:3473-3485) and theta at
 * cur_mixlen*nstates*nmix (:3512): with nmix > 1 and unfused rates that is not the category layout of the other
 * kernels, so the entry is restricted to what PhyloTreeMixlen allows on this path (fused, or nmix == 1; :427-428). */
int orc_derv_mixlen_class(orc_ctx *c, int32_t cur_mixlen, double len, double *df, double *ddf) {
    if (!c->theta_computed) return B200LK_ERR_STATE;
    if (c->site || c->nonrev || c->asc_holder) return B200LK_ERR_UNSUPPORTED;
    if (!c->d.fused_mix_rate && c->d.nmix > 1) return B200LK_ERR_UNSUPPORTED;
    if (cur_mixlen < 0 || cur_mixlen >= c->d.ncat) return B200LK_ERR_ARG;
    const int n = c->d.nstates, block = c->block;
    const int64_t nptn = c->d.nptn, ntot = c->ntot;
    const int cur_mix = c->d.fused_mix_rate ? cur_mixlen : 0;
    double val0[n], val1[n], val2[n];
    double prop = c->mixw[cur_mix];
    for (int i = 0; i < n; i++) { /* :3472-3486: no rate, no category proportion */
        double cof = c->eval[cur_mix * n + i];
        double val = exp(cof * len) * prop;
        val0[i] = val;
        val1[i] = cof * val;
        val2[i] = cof * val1[i];
    }
    double all_df = 0.0, all_ddf = 0.0, prob_const = 0.0, df_const = 0.0, ddf_const = 0.0;
    for (int64_t p = 0; p < ntot; p++) {
        double lh, d1, d2;
        dot_triple(val0, val1, val2, c->theta_all + (size_t)p * block + (size_t)cur_mixlen * n, &lh, &d1, &d2, n, n);
        lh = fabs(lh) + c->ptn_invar[p];
        if (p < nptn) {
            double inv = 1.0 / lh, freq = c->ptn_freq[p];
            d1 *= inv;
            d2 *= inv;
            all_df = fma(d1, freq, all_df);
            all_ddf += fma(-d1, d1, d2) * freq;
        } else {
            prob_const += lh;
            df_const += d1;
            ddf_const += d2;
        }
    }
    *df = all_df;
    *ddf = all_ddf;
    if (c->d.n_unobserved > 0) { /* :3560-3568 (Lewis) */
        double nsite = c->nsite; /* aln->getNSite() */
        prob_const = 1.0 / (1.0 - prob_const);
        df_const *= prob_const;
        ddf_const *= prob_const;
        *df += nsite * df_const;
        *ddf += nsite * (ddf_const + df_const * df_const);
    }
    if (!isfinite(*df)) {
        *df = *ddf = 0.0;
        return B200LK_WARN_DERV_NONFINITE;
    }
    return B200LK_OK;
}
/* new pattern weights (computePtnFreq, or the EM posteriors of tree/phylotreemixlen.cpp:470-474) */
int orc_set_ptn_freq(orc_ctx *c, const double *ptn_freq) {
    memcpy(c->ptn_freq, ptn_freq, sizeof(double) * c->d.nptn);
    return B200LK_OK;
}

int orc_evaluate(orc_ctx *c, const b200lk_node *nodes, int32_t nnodes, const b200lk_child *children,
                 int32_t nchildren, b200lk_end dad, b200lk_end node, double length, double *lnl, double *df,
                 double *ddf) {
    int rc = orc_update_partials(c, nodes, nnodes, children, nchildren), w = 0;
    if (rc < 0) return rc;
    double tmp;
    if (lnl) {
        rc = orc_branch_lnl(c, dad, node, length, lnl);
        if (rc < 0) return rc;
        w |= rc;
    }
    if (df || ddf) {
        rc = orc_build_theta(c, dad, node);
        if (rc < 0) return rc;
        rc = orc_derv(c, length, df ? df : &tmp, ddf ? ddf : &tmp);
        if (rc < 0) return rc;
        w |= rc;
    }
    return w;
}

/* ------------------------------------------------------------------ readers */
int orc_fetch_pattern_lh(orc_ctx *c, double *out) {
    memcpy(out, c->pattern_lh, sizeof(double) * c->d.nptn);
    return B200LK_OK;
}
int orc_fetch_pattern_lh_cat(orc_ctx *c, double *out) {
    memcpy(out, c->pattern_lh_cat, sizeof(double) * c->d.nptn * c->ncat_mix);
    return B200LK_OK;
}
int orc_fetch_scale_num(orc_ctx *c, int32_t slot, uint8_t *out) {
    if (slot < 0 || slot >= c->d.nslots || !c->scale[slot]) return B200LK_ERR_ARG;
    size_t per = c->d.safe_numeric ? c->ncat_mix : 1;
    memcpy(out, c->scale[slot], (size_t)c->d.nptn * per);
    return B200LK_OK;
}
int orc_fetch_partial(orc_ctx *c, int32_t slot, double *out) {
    if (slot < 0 || slot >= c->d.nslots || !c->partial[slot]) return B200LK_ERR_ARG;
    memcpy(out, c->partial[slot], sizeof(double) * c->d.nptn * c->block);
    return B200LK_OK;
}
int orc_fetch_theta(orc_ctx *c, double *theta, double *bscale) {
    memcpy(theta, c->theta_all, sizeof(double) * c->d.nptn * c->block);
    if (bscale) memcpy(bscale, c->buffer_scale_all, sizeof(double) * c->d.nptn);
    return B200LK_OK;
}
int orc_get_counters(orc_ctx *c, int64_t *updates, int64_t *launches) {
    if (updates) *updates = c->updates;
    if (launches) *launches = 0;
    return B200LK_OK;
}
