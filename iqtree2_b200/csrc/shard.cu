// shard.cu -- one problem split by site pattern over several GPUs of THIS process (SURVEY.md 8b `device_mask`, 8e).
//
// The reference is one process whose kernels split the pattern range into "packets" handed to OpenMP threads
// (tree/phylokernelnew.h:1114-1142) and sum the per-packet lnL / df / ddf (:2377, 2792, 3273).  b200lk_create_sharded is
// the device form of that: the parent context owns no device memory; it owns one ordinary context per device over a
// contiguous range of whole 32-pattern tiles (the arithmetic of b200::shardPatterns), each with its own stream.  Every
// entry point first ENQUEUES its work on all shards (nothing is exchanged during a traversal, so the GPUs run
// concurrently from one host thread) and only then reads the per-shard sums back, adding them on the host in
// device-index order: the total is bit-reproducible from run to run, which an NCCL all-reduce is not required to be.
// Per-pattern readers (b200lk_fetch_*) gather the shards' ranges into the caller's pattern-major arrays.
#include <math.h>

#include <algorithm>

#include "common.cuh"

#define NSHARD(c) ((int)(c)->shards->ctx.size())
#define SHARD(c, i) ((c)->shards->ctx[i])

static int shard_fail(b200lk_ctx *c, int i, int rc) {  // keep the message of the shard that failed
    if (rc < 0) snprintf(c->err, sizeof(c->err), "shard %d (device %d): %.400s", i, SHARD(c, i)->device, SHARD(c, i)->err);
    return rc;
}

int shard_refuse(b200lk_ctx *c, const char *what) {
    snprintf(c->err, sizeof(c->err), "%s is not available on a sharded context (b200lk_create_sharded)", what);
    return B200LK_ERR_UNSUPPORTED;
}

static void shard_range(int64_t nptn, int world, int rank, int tile, int64_t *begin, int64_t *end) {
    const int64_t tiles = (nptn + tile - 1) / tile;
    const int64_t base = tiles / world, extra = tiles % world;  // the first `extra` shards take one more tile
    const int64_t t0 = rank * base + std::min<int64_t>(rank, extra);
    const int64_t t1 = t0 + base + (rank < extra ? 1 : 0);
    *begin = std::min(nptn, t0 * tile);
    *end = std::min(nptn, t1 * tile);
}

extern "C" int b200lk_create_sharded(b200lk_ctx **out, const b200lk_dims *dims, const int32_t *devices, int32_t ndev) {
    if (!out || !dims || !devices || ndev < 1 || ndev > 64) return B200LK_ERR_ARG;
    *out = nullptr;
    // what a pattern shard cannot express: +ASC sums must be combined BEFORE the non-linear correction
    // (tree/phylokernelnew.h:2620-2628, 3160-3181), site-specific and non-reversible contexts keep per-pattern /
    // rooted state the per-shard calls below do not split
    if (dims->n_unobserved != 0 || dims->site_specific != B200LK_MODEL_REV) return B200LK_ERR_UNSUPPORTED;
    b200lk_ctx *c = new b200lk_ctx();
    memset(c, 0, sizeof(*c));
    c->d = *dims;
    c->device = devices[0];
    c->ncat_mix = dims->fused_mix_rate ? dims->ncat : dims->ncat * dims->nmix;
    c->block = dims->nstates * c->ncat_mix;
    c->scale_per_ptn = dims->safe_numeric ? c->ncat_mix : 1;
    c->ntot = dims->nptn;
    c->robust_keep = 1.0;
    c->cur_newton_slot = -1;
    c->shards = new ShardSet();
    for (int i = 0; i < ndev; i++) {
        int64_t b, e;
        shard_range(dims->nptn, ndev, i, 32, &b, &e);
        b200lk_dims di = *dims;
        di.nptn = e - b;
        di.device = devices[i];
        b200lk_ctx *sc = nullptr;
        const int rc = b200lk_create(&sc, &di);
        if (rc != B200LK_OK) {
            for (b200lk_ctx *x : c->shards->ctx) b200lk_destroy(x);
            delete c->shards;
            delete c;
            return rc;
        }
        c->shards->ctx.push_back(sc);
        c->shards->begin.push_back(b);
        c->shards->end.push_back(e);
    }
    *out = c;
    return B200LK_OK;
}

extern "C" int b200lk_shard_info(b200lk_ctx *c, int32_t *nshards, int32_t *devices, int64_t *begin, int64_t *end) {
    if (!c) return B200LK_ERR_ARG;
    const int n = c->shards ? NSHARD(c) : 1;
    if (nshards) *nshards = n;
    for (int i = 0; i < n; i++) {
        if (devices) devices[i] = c->shards ? SHARD(c, i)->device : c->device;
        if (begin) begin[i] = c->shards ? c->shards->begin[i] : 0;
        if (end) end[i] = c->shards ? c->shards->end[i] : c->d.nptn;
    }
    return B200LK_OK;
}

int shard_destroy(b200lk_ctx *c) {
    for (b200lk_ctx *sc : c->shards->ctx) b200lk_destroy(sc);
    delete c->shards;
    delete c;
    return B200LK_OK;
}

// ---- inputs --------------------------------------------------------------------------------------------------------
int shard_upload_tips(b200lk_ctx *c, const uint8_t *states, const uint8_t *unobserved, const double *ptn_freq) {
    if ((!states && c->d.nptn) || !ptn_freq || unobserved) return B200LK_ERR_ARG;
    for (int i = 0; i < NSHARD(c); i++) {
        const int64_t b = c->shards->begin[i];
        int rc = upload_tips_pitched(SHARD(c, i), states ? states + b : nullptr, (size_t)c->d.nptn, nullptr, 0, ptn_freq + b);
        if (rc) return shard_fail(c, i, rc);
    }
    c->tips_set = true;
    return B200LK_OK;
}

#define EACH_SHARD(expr)                            \
    for (int i = 0; i < NSHARD(c); i++) {           \
        b200lk_ctx *sc = SHARD(c, i);               \
        const int64_t b = c->shards->begin[i];      \
        (void)b;                                    \
        int rc = (expr);                            \
        if (rc < 0) return shard_fail(c, i, rc);    \
    }

// The same on one host thread per shard: what a call does on the host before it launches (levels, node programs, the
// exp(lambda r t) factors of every child -- ~0.2 ms for a cfg2 traversal) would otherwise be paid once per shard in
// turn.  Contexts are independently thread-safe (own stream, own staging, thread-local launch parameters).
#define EACH_SHARD_PARALLEL(expr)                                                   \
    {                                                                               \
        const int nsh__ = NSHARD(c);                                                \
        std::vector<int> rcs__(nsh__, 0);                                           \
        _Pragma("omp parallel for schedule(static, 1) num_threads(nsh__)")          \
        for (int i = 0; i < nsh__; i++) {                                           \
            b200lk_ctx *sc = SHARD(c, i);                                           \
            rcs__[i] = (expr);                                                      \
        }                                                                           \
        for (int i = 0; i < nsh__; i++)                                             \
            if (rcs__[i] < 0) return shard_fail(c, i, rcs__[i]);                    \
    }

int shard_set_ptn_invar(b200lk_ctx *c, const double *v) {
    if (!v) return B200LK_ERR_ARG;
    EACH_SHARD(b200lk_set_ptn_invar(sc, v + b));
    return B200LK_OK;
}
int shard_set_ptn_freq(b200lk_ctx *c, const double *v) {
    if (!v) return B200LK_ERR_ARG;
    EACH_SHARD(b200lk_set_ptn_freq(sc, v + b));
    return B200LK_OK;
}
int shard_set_model(b200lk_ctx *c, const double *eval, const double *evec, const double *inv_evec, const double *rates,
                    const double *props, const double *mixw, const double *tip) {
    EACH_SHARD(b200lk_set_model(sc, eval, evec, inv_evec, rates, props, mixw, tip));
    c->model_set = true;
    c->theta_computed = false;
    return B200LK_OK;
}
int shard_set_model_rates(b200lk_ctx *c, const double *exch, const double *freqs, const double *tot, const double *tipvec,
                          const double *rates, const double *props, const double *mixw) {
    EACH_SHARD(b200lk_set_model_rates(sc, exch, freqs, tot, tipvec, rates, props, mixw));  // solved redundantly per GPU
    c->model_set = true;
    c->theta_computed = false;
    return B200LK_OK;
}
int shard_set_robust(b200lk_ctx *c, double keep, int32_t median) {
    if (keep < 1.0 || median) return shard_refuse(c, "robust-phylogeny scoring (a selection over ALL sites)");
    return B200LK_OK;
}

// ---- the four kernel entry points ------------------------------------------------------------------------------------
int shard_update_partials(b200lk_ctx *c, const b200lk_node *nodes, int32_t nnodes, const b200lk_child *children, int32_t nchildren) {
    EACH_SHARD_PARALLEL(b200lk_update_partials(sc, nodes, nnodes, children, nchildren));  // asynchronous on every shard
    c->theta_computed = false;
    return B200LK_OK;
}

int shard_branch_lnl(b200lk_ctx *c, b200lk_end dad, b200lk_end node, double length, double *lnl) {
    if (!lnl) return B200LK_ERR_ARG;
    for (int i = 0; i < NSHARD(c); i++) {  // enqueue everywhere first ...
        b200lk_ctx *sc = SHARD(c, i);
        if (!sc->model_set || !sc->tips_set) return B200LK_ERR_STATE;
        if (check_ends(sc, dad, node)) return B200LK_ERR_ARG;
        if (sc->d.nptn == 0) continue;
        cudaSetDevice(sc->device);
        int rc = launch_branch_lnl(sc, dad, node, length);
        if (rc) return shard_fail(c, i, rc);
    }
    double sum = 0.0;
    int warn = 0;
    for (int i = 0; i < NSHARD(c); i++) {  // ... then read back, summing in device-index order
        b200lk_ctx *sc = SHARD(c, i);
        if (sc->d.nptn == 0) continue;
        cudaSetDevice(sc->device);
        double v = 0.0;
        int rc = finish_lnl(sc, &v, true);
        if (rc < 0) return shard_fail(c, i, rc);
        warn |= rc;
        sum += v;
    }
    *lnl = sum;
    return warn;
}

int shard_build_theta(b200lk_ctx *c, b200lk_end dad, b200lk_end node) {
    EACH_SHARD(b200lk_build_theta(sc, dad, node));
    c->theta_computed = true;
    return B200LK_OK;
}

// sum of the shards' derivative terms; a non-finite first derivative in ANY shard is the reference's "Numerical
// underflow (lh-derivative)" for the whole alignment (tree/phylokernelnew.h:2573-2578, 2629-2632)
static int collect_derv(b200lk_ctx *c, double *df, double *ddf) {
    double a = 0.0, b2 = 0.0;
    int warn = 0;
    for (int i = 0; i < NSHARD(c); i++) {
        b200lk_ctx *sc = SHARD(c, i);
        if (sc->d.nptn == 0) continue;
        cudaSetDevice(sc->device);
        int rc = read_results(sc, 1, 2);
        if (rc < 0) return shard_fail(c, i, rc);
        a += sc->h_result[1];
        b2 += sc->h_result[2];
    }
    if (!isfinite(a)) {
        warn = B200LK_WARN_DERV_NONFINITE;
        if (c->d.safe_numeric) a = b2 = 0.0;
    }
    if (df) *df = a;
    if (ddf) *ddf = b2;
    return warn;
}

int shard_derv(b200lk_ctx *c, double length, double *df, double *ddf) {
    if (!c->theta_computed) return B200LK_ERR_STATE;
    for (int i = 0; i < NSHARD(c); i++) {
        b200lk_ctx *sc = SHARD(c, i);
        if (sc->d.nptn == 0) continue;
        if (!sc->theta_computed) return B200LK_ERR_STATE;
        cudaSetDevice(sc->device);
        int rc = launch_derv(sc, length);
        if (rc) return shard_fail(c, i, rc);
    }
    return collect_derv(c, df, ddf);
}

int shard_lnl_from_buffer(b200lk_ctx *c, double length, double *lnl) {
    if (!lnl) return B200LK_ERR_ARG;
    if (!c->theta_computed) return B200LK_ERR_STATE;
    for (int i = 0; i < NSHARD(c); i++) {
        b200lk_ctx *sc = SHARD(c, i);
        if (sc->d.nptn == 0) continue;
        cudaSetDevice(sc->device);
        int rc = launch_from_buffer(sc, length);
        if (rc) return shard_fail(c, i, rc);
    }
    double sum = 0.0;
    int warn = 0;
    for (int i = 0; i < NSHARD(c); i++) {
        b200lk_ctx *sc = SHARD(c, i);
        if (sc->d.nptn == 0) continue;
        cudaSetDevice(sc->device);
        double v = 0.0;
        int rc = finish_lnl(sc, &v, false);
        if (rc < 0) return shard_fail(c, i, rc);
        warn |= rc;
        sum += v;
    }
    *lnl = sum;
    return warn;
}

int shard_evaluate_async(b200lk_ctx *c, const b200lk_node *nodes, int32_t nnodes, const b200lk_child *children,
                         int32_t nchildren, b200lk_end dad, b200lk_end node, double length, int want_lnl, int want_derv) {
    EACH_SHARD_PARALLEL(b200lk_evaluate_async(sc, nodes, nnodes, children, nchildren, dad, node, length, want_lnl, want_derv));
    if (want_derv) c->theta_computed = true;
    return B200LK_OK;
}

int shard_read_results(b200lk_ctx *c, double *lnl, double *df, double *ddf) {
    double s[3] = {0.0, 0.0, 0.0};
    int warn = 0;
    bool derv_bad = false;
    for (int i = 0; i < NSHARD(c); i++) {
        b200lk_ctx *sc = SHARD(c, i);
        if (sc->d.nptn == 0) continue;
        double v[3] = {0.0, 0.0, 0.0};
        int rc = b200lk_read_results(sc, lnl ? &v[0] : nullptr, df ? &v[1] : nullptr, ddf ? &v[2] : nullptr);
        if (rc < 0) return shard_fail(c, i, rc);
        warn |= rc;
        derv_bad |= (rc & B200LK_WARN_DERV_NONFINITE) != 0;
        for (int k = 0; k < 3; k++) s[k] += v[k];
    }
    if (derv_bad && c->d.safe_numeric) s[1] = s[2] = 0.0;
    if (lnl) *lnl = s[0];
    if (df) *df = s[1];
    if (ddf) *ddf = s[2];
    return warn;
}

// Optimization::minimizeNewton as PhyloTree::optimizeOneBranch calls it (utils/optimization.cpp:422-499,
// tree/phylotree.cpp:2585-2635), f = -df, f' = -ddf summed over the shards per step.  The single-device contexts run
// this loop in one cooperative launch (newton_kernel); across devices every step is one fan-out + one ordered sum.
int shard_optimize_branch(b200lk_ctx *c, b200lk_end dad, b200lk_end node, const b200lk_newton *nr, double *optx, double *d2l,
                          int32_t *nsteps) {
    if (!nr || !optx || !(nr->min_len <= nr->max_len) || nr->max_steps < 1) return B200LK_ERR_ARG;
    int rc = shard_build_theta(c, dad, node);
    if (rc) return rc;
    const double x1 = nr->min_len, x2 = nr->max_len, xacc = nr->xacc;
    int evals = 0, warn = 0;
    auto fd = [&](double x, double &f, double &df) {
        double a = 0.0, b = 0.0;
        int r = shard_derv(c, x, &a, &b);
        evals++;
        if (r > 0) warn |= r;
        f = -a;
        df = -b;
        return r < 0 ? r : 0;
    };
    double rts = std::min(std::max(nr->guess, x1), x2), f, df, d2 = 0.0;
    auto done = [&](double x) {
        *optx = x;
        if (d2l) *d2l = d2;
        if (nsteps) *nsteps = evals;
        return warn;
    };
    if ((rc = fd(rts, f, df)) < 0) return rc;
    d2 = df;
    if (!isfinite(f) || !isfinite(df)) { done(rts); return B200LK_WARN_DERV_NONFINITE; }
    if (df >= 0.0 && fabs(f) < xacc) return done(rts);
    double xl, xh;
    if (f < 0.0) { xl = rts; xh = x2; } else { xh = rts; xl = x1; }
    double dx = fabs(xh - xl);
    for (int j = 1; j <= nr->max_steps; j++) {
        const double rts_old = rts;
        if (df <= 0.0 || (((rts - xh) * df - f) * ((rts - xl) * df - f) >= 0.0)) {
            dx = 0.5 * (xh - xl);
            rts = xl + dx;
            d2 = df;
            if (xl == rts) return done(rts);
        } else {
            dx = f / df;
            const double temp = rts;
            rts -= dx;
            d2 = df;
            if (temp == rts) return done(rts);
        }
        if (fabs(dx) < xacc || j == nr->max_steps) return done(rts_old);
        if ((rc = fd(rts, f, df)) < 0) return rc;
        if (!isfinite(f) || !isfinite(df)) { done(rts); return B200LK_WARN_DERV_NONFINITE; }
        if (df > 0.0 && fabs(f) < xacc) { d2 = df; return done(rts); }
        if (f < 0.0) xl = rts;
        else if (f > 0.0) xh = rts;
    }
    return done(rts);
}

// ---- readers: every shard's range into the caller's pattern-major array ----------------------------------------------
int shard_fetch_pattern_lh(b200lk_ctx *c, double *out) {
    if (!out) return B200LK_ERR_ARG;
    EACH_SHARD(sc->d.nptn ? b200lk_fetch_pattern_lh(sc, out + b) : 0);
    return B200LK_OK;
}
int shard_fetch_pattern_lh_cat(b200lk_ctx *c, double *out) {
    if (!out) return B200LK_ERR_ARG;
    EACH_SHARD(sc->d.nptn ? b200lk_fetch_pattern_lh_cat(sc, out + (size_t)b * c->ncat_mix) : 0);
    return B200LK_OK;
}
int shard_fetch_scale_num(b200lk_ctx *c, int32_t slot, uint8_t *out) {
    if (!out) return B200LK_ERR_ARG;
    EACH_SHARD(sc->d.nptn ? b200lk_fetch_scale_num(sc, slot, out + (size_t)b * c->scale_per_ptn) : 0);
    return B200LK_OK;
}
int shard_fetch_partial(b200lk_ctx *c, int32_t slot, double *out) {
    if (!out) return B200LK_ERR_ARG;
    EACH_SHARD(sc->d.nptn ? b200lk_fetch_partial(sc, slot, out + (size_t)b * c->block) : 0);
    return B200LK_OK;
}
int shard_fetch_theta(b200lk_ctx *c, double *theta, double *bscale) {
    if (!theta) return B200LK_ERR_ARG;
    EACH_SHARD(sc->d.nptn ? b200lk_fetch_theta(sc, theta + (size_t)b * c->block, bscale ? bscale + b : nullptr) : 0);
    return B200LK_OK;
}
int shard_fetch_model(b200lk_ctx *c, double *eval, double *evec, double *inv_evec) {
    return b200lk_fetch_model(SHARD(c, 0), eval, evec, inv_evec);  // replicated: every shard holds the same model
}

// ---- bookkeeping -------------------------------------------------------------------------------------------------------
int shard_get_counters(b200lk_ctx *c, int64_t *updates, int64_t *launches) {
    int64_t u = 0, l = 0;
    for (b200lk_ctx *sc : c->shards->ctx) { u += sc->n_updates; l += sc->n_launches; }
    if (updates) *updates = u;
    if (launches) *launches = l;
    return B200LK_OK;
}
int shard_reset_counters(b200lk_ctx *c) {
    EACH_SHARD(b200lk_reset_counters(sc));
    return B200LK_OK;
}
int shard_get_min_traffic(b200lk_ctx *c, int64_t *bytes) {
    int64_t s = 0;
    for (b200lk_ctx *sc : c->shards->ctx) s += sc->n_min_bytes;
    *bytes = s;
    return B200LK_OK;
}
int shard_synchronize(b200lk_ctx *c) {
    EACH_SHARD(b200lk_synchronize(sc));
    return B200LK_OK;
}
int shard_set_profiling(b200lk_ctx *c, int on) {
    EACH_SHARD(b200lk_set_profiling(sc, on));
    return B200LK_OK;
}
// the shards run concurrently: the partial-update time of the problem is the slowest shard's; launches are summed
int shard_get_profile(b200lk_ctx *c, double *partial_ms, int64_t *partial_launches, int reset) {
    double mx = 0.0;
    int64_t n = 0;
    for (int i = 0; i < NSHARD(c); i++) {
        double ms = 0.0;
        int64_t l = 0;
        int rc = b200lk_get_profile(SHARD(c, i), &ms, &l, reset);
        if (rc < 0) return shard_fail(c, i, rc);
        mx = std::max(mx, ms);
        n += l;
    }
    if (partial_ms) *partial_ms = mx;
    if (partial_launches) *partial_launches = n;
    return B200LK_OK;
}

// ---- UFBoot RELL over shards: rell[b] = sum over shards of their part of the dot product ---------------------------------
int shard_upload_boot_samples(b200lk_ctx *c, const float *samples, int32_t nsamples) {
    if (!samples || nsamples < 1) return B200LK_ERR_ARG;
    EACH_SHARD(upload_boot_samples_pitched(sc, samples + b, (size_t)c->d.nptn, nsamples));
    c->nboot = nsamples;
    return B200LK_OK;
}
int shard_rell(b200lk_ctx *c, double *rell) {
    if (!rell) return B200LK_ERR_ARG;
    if (c->nboot < 1) return B200LK_ERR_STATE;
    std::vector<double> part(c->nboot);
    for (int k = 0; k < c->nboot; k++) rell[k] = 0.0;
    for (int i = 0; i < NSHARD(c); i++) {
        if (SHARD(c, i)->d.nptn == 0) continue;
        int rc = b200lk_rell(SHARD(c, i), part.data());
        if (rc < 0) return shard_fail(c, i, rc);
        for (int k = 0; k < c->nboot; k++) rell[k] += part[k];
    }
    return B200LK_OK;
}
