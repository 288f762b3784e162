// batch.cu -- the PhyloSuperTree partition dispatch as batched launches (7.2 step 8).
//
// The reference fans a likelihood call out over its partitions on host threads, one small kernel invocation per
// gene (tree/phylosupertree.cpp:701-725).  On the device a supermatrix of 1000 genes x ~1k patterns is bound by the
// NUMBER of launches (about 20 per gene), not by their work.  A batch groups contexts that share dimensions and
// tree topology (an edge-linked or unlinked partition model: same tree, per-gene lengths, models and data) and
// runs every step of an evaluation as ONE launch whose extra grid dimension is the context: the traversal structure
// (node list, levels, slots, table offsets, tip-state staging order) is built once from the first context; each CTA
// substitutes its own context's buffers, model and branch lengths (BatchCtxDev, common.cuh).
#include <algorithm>

#include "common.cuh"

struct b200lk_batch {
    std::vector<b200lk_ctx *> ctx;
    int n, device;
    cudaStream_t stream;
    BatchCtxDev *d_ctx;
    double *d_child_len, *h_child_len;  // [n][child_cap]
    int child_cap;
    double *d_branch_len, *h_branch_len;  // [n]
    double *d_results, *h_results;        // [n][4]
    BatchLaunch bl;
    bool pending, pending_derv;  // an asynchronous evaluation whose results have not been read yet
    char err[512];
};

#define BATCH_TRY(b, expr)                                                                       \
    do {                                                                                         \
        cudaError_t e__ = (expr);                                                                \
        if (e__ != cudaSuccess) {                                                                \
            snprintf((b)->err, sizeof((b)->err), "%s:%d %s: %s", __FILE__, __LINE__, #expr,      \
                     cudaGetErrorString(e__));                                                   \
            return B200LK_ERR_CUDA;                                                              \
        }                                                                                        \
    } while (0)

extern "C" const char *b200lk_batch_last_error(const b200lk_batch *b) { return b ? b->err : "null batch"; }

extern "C" int b200lk_batch_destroy(b200lk_batch *b) {
    if (!b) return B200LK_OK;
    cudaSetDevice(b->device);
    if (b->stream) cudaStreamSynchronize(b->stream);
    if (b->d_ctx) cudaFree(b->d_ctx);
    if (b->d_child_len) cudaFree(b->d_child_len);
    if (b->h_child_len) cudaFreeHost(b->h_child_len);
    if (b->d_branch_len) cudaFree(b->d_branch_len);
    if (b->h_branch_len) cudaFreeHost(b->h_branch_len);
    if (b->d_results) cudaFree(b->d_results);
    if (b->h_results) cudaFreeHost(b->h_results);
    if (b->stream) cudaStreamDestroy(b->stream);
    delete b;
    return B200LK_OK;
}

extern "C" int b200lk_batch_create(b200lk_batch **out, b200lk_ctx **ctxs, int32_t n) {
    if (!out || !ctxs || n < 1) return B200LK_ERR_ARG;
    *out = nullptr;
    const b200lk_ctx *c0 = ctxs[0];
    if (!c0) return B200LK_ERR_ARG;
    for (int i = 0; i < n; i++)
        if (ctxs[i] && ctxs[i]->shards) return B200LK_ERR_UNSUPPORTED;  // a sharded parent owns no device buffers
    // what a batched launch can express: the 4-state and the 20-state tensor path, one category group, no +ASC
    if (!(c0->tile == 32 || (c0->dense20 && c0->ncat_mix <= 4)) || c0->d.n_unobserved != 0) return B200LK_ERR_UNSUPPORTED;
    for (int i = 0; i < n; i++) {
        const b200lk_ctx *c = ctxs[i];
        if (!c) return B200LK_ERR_ARG;
        const b200lk_dims &a = c->d, &z = c0->d;
        if (c->device != c0->device || a.ntaxa != z.ntaxa || a.nstates != z.nstates || a.ncat != z.ncat ||
            a.nmix != z.nmix || a.fused_mix_rate != z.fused_mix_rate || a.safe_numeric != z.safe_numeric ||
            a.n_state_codes != z.n_state_codes || a.nslots != z.nslots || a.n_unobserved != 0 || c->tile != c0->tile ||
            c->dense20 != c0->dense20)
            return B200LK_ERR_UNSUPPORTED;
    }
    b200lk_batch *b = new b200lk_batch();  // value-initialised: every scalar member starts at zero
    b->ctx.assign(ctxs, ctxs + n);
    b->n = n;
    b->device = c0->device;
    *out = b;
    BATCH_TRY(b, cudaSetDevice(b->device));
    BATCH_TRY(b, cudaStreamCreateWithFlags(&b->stream, cudaStreamNonBlocking));
    std::vector<BatchCtxDev> h(n);
    int64_t max_ntot = 0;
    for (int i = 0; i < n; i++) {
        const b200lk_ctx *c = ctxs[i];
        BatchCtxDev &d = h[i];
        d.partial = c->partial;
        d.scale = c->scale;
        d.states = c->states;
        d.ptn_invar = c->ptn_invar;
        d.ptn_freq = c->ptn_freq;
        d.tables = c->tables;
        d.m = c->m;
        d.theta = c->theta_all;
        d.buffer_scale_all = c->buffer_scale_all;
        d.pattern_lh = c->pattern_lh;
        d.pattern_lh_cat = c->pattern_lh_cat;
        d.val = c->val;
        d.tip_val = c->tip_val;
        d.red_partial = c->red_partial;
        d.d_result = c->d_result;
        d.slot_doubles = c->slot_doubles;
        d.slot_scale_bytes = c->slot_scale_bytes;
        d.ntot = c->ntot;
        d.nptn = c->d.nptn;
        d.ntot_pad = c->ntot_pad;
        max_ntot = std::max(max_ntot, c->ntot);
    }
    BATCH_TRY(b, cudaMalloc(&b->d_ctx, sizeof(BatchCtxDev) * n));
    BATCH_TRY(b, cudaMemcpy(b->d_ctx, h.data(), sizeof(BatchCtxDev) * n, cudaMemcpyHostToDevice));
    b->child_cap = c0->children_cap;
    BATCH_TRY(b, cudaMalloc(&b->d_child_len, sizeof(double) * (size_t)n * b->child_cap));
    BATCH_TRY(b, cudaMallocHost(&b->h_child_len, sizeof(double) * (size_t)n * b->child_cap));
    BATCH_TRY(b, cudaMalloc(&b->d_branch_len, sizeof(double) * n));
    BATCH_TRY(b, cudaMallocHost(&b->h_branch_len, sizeof(double) * n));
    BATCH_TRY(b, cudaMalloc(&b->d_results, sizeof(double) * 4 * n));
    BATCH_TRY(b, cudaMallocHost(&b->h_results, sizeof(double) * 4 * n));
    b->bl.d_ctx = b->d_ctx;
    b->bl.n = n;
    b->bl.max_ntot = max_ntot;
    b->bl.d_child_len = b->d_child_len;
    b->bl.d_branch_len = b->d_branch_len;
    b->bl.nchildren = 0;
    return B200LK_OK;
}

__global__ void batch_gather_kernel(const BatchCtxDev *ctx, int n, double *out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    for (int k = 0; k < 3; k++) out[4 * i + k] = ctx[i].d_result[k];
}

// run `body` with the first context temporarily bound to the batch (its stream, the per-context substitution table)
template <class F>
static int with_batch(b200lk_batch *b, F body) {
    b200lk_ctx *c0 = b->ctx[0];
    cudaStream_t saved = c0->stream;
    c0->stream = b->stream;
    c0->batch = &b->bl;
    int rc = body(c0);
    c0->batch = nullptr;
    c0->stream = saved;
    if (rc < 0 && !b->err[0]) snprintf(b->err, sizeof(b->err), "%s", c0->err);
    return rc;
}

// a member may have work of its own calls still in flight on its own stream (model upload, a per-context update)
static int wait_for_members(b200lk_batch *b) {
    if (b->pending) return B200LK_ERR_STATE;  // read the results of the evaluation in flight first
    for (b200lk_ctx *c : b->ctx) {
        // vectors a member's own 4-state traversal left virtual (s4.cu): the batched kernels read slots from HBM
        if (int rc = s4_realize(c)) {
            snprintf(b->err, sizeof(b->err), "%s", c->err);
            return rc;
        }
        if (c->stream != b->stream) BATCH_TRY(b, cudaStreamSynchronize(c->stream));
    }
    return B200LK_OK;
}

static int upload_branch_len(b200lk_batch *b, const double *branch_len) {
    int rc = wait_for_members(b);
    if (rc) return rc;
    memcpy(b->h_branch_len, branch_len, sizeof(double) * b->n);
    BATCH_TRY(b, cudaMemcpyAsync(b->d_branch_len, b->h_branch_len, sizeof(double) * b->n, cudaMemcpyHostToDevice, b->stream));
    return B200LK_OK;
}

static int unpack(b200lk_batch *b, double *lnl, double *df, double *ddf) {
    int warn = 0;
    for (int i = 0; i < b->n; i++) {
        if (lnl) {
            lnl[i] = b->h_results[4 * i];
            if (!isfinite(lnl[i])) warn |= B200LK_WARN_LNL_NONFINITE;
        }
        if (df) {
            df[i] = b->h_results[4 * i + 1];
            if (!isfinite(df[i])) warn |= B200LK_WARN_DERV_NONFINITE;
        }
        if (ddf) ddf[i] = b->h_results[4 * i + 2];
    }
    return warn;
}

// The reference repairs a non-finite lnL per partition tree (tree/phylokernelnew.h:3077-3094, 3338-3347: non-finite
// per-pattern values are replaced by 4 * LOG_SCALING_THRESHOLD and the sum retaken); the per-context calls do the same in
// finish_lnl.  A batch does it here for the members concerned, each on its own context, so that results do not depend
// on whether batching is on.  The warning bit stays set.
static int repair_members(b200lk_batch *b, double *lnl) {
    if (!lnl) return B200LK_OK;
    for (int i = 0; i < b->n; i++) {
        if (isfinite(lnl[i])) continue;
        b200lk_ctx *c = b->ctx[i];
        if (c->d.nptn == 0) continue;
        cudaStream_t saved = c->stream;
        c->stream = b->stream;
        int rc = launch_repair(c);
        if (rc == B200LK_OK) {
            cudaError_t e = cudaMemcpyAsync(c->h_result, c->d_result, sizeof(double), cudaMemcpyDeviceToHost, b->stream);
            if (e == cudaSuccess) e = cudaStreamSynchronize(b->stream);
            if (e != cudaSuccess) {
                snprintf(b->err, sizeof(b->err), "repair of batch member %d: %s", i, cudaGetErrorString(e));
                rc = B200LK_ERR_CUDA;
            }
        } else {
            snprintf(b->err, sizeof(b->err), "%s", c->err);
        }
        c->stream = saved;
        if (rc) return rc;
        lnl[i] = c->h_result[0];
    }
    return B200LK_OK;
}

static int collect(b200lk_batch *b, double *lnl, double *df, double *ddf) {
    batch_gather_kernel<<<(b->n + 127) / 128, 128, 0, b->stream>>>(b->d_ctx, b->n, b->d_results);
    BATCH_TRY(b, cudaGetLastError());
    BATCH_TRY(b, cudaMemcpyAsync(b->h_results, b->d_results, sizeof(double) * 4 * b->n, cudaMemcpyDeviceToHost, b->stream));
    BATCH_TRY(b, cudaStreamSynchronize(b->stream));
    const int warn = unpack(b, lnl, df, ddf);
    if (warn & B200LK_WARN_LNL_NONFINITE) {
        int rc = repair_members(b, lnl);
        if (rc) return rc;
    }
    return warn;
}

extern "C" int b200lk_batch_evaluate_async(b200lk_batch *b, const b200lk_node *nodes, int32_t nnodes,
                                           const b200lk_child *children, int32_t nchildren, const double *child_len,
                                           b200lk_end dad, b200lk_end node, const double *branch_len, int want_derv) {
    if (!b || !branch_len || (nnodes > 0 && (!nodes || !children || !child_len)) || nchildren > b->child_cap)
        return B200LK_ERR_ARG;
    b->err[0] = 0;
    BATCH_TRY(b, cudaSetDevice(b->device));
    for (int i = 0; i < nnodes; i++)
        if (nodes[i].nchild != 2) return B200LK_ERR_UNSUPPORTED;  // multifurcations take the per-context path
    for (b200lk_ctx *c : b->ctx)
        if (!c->model_set || !c->tips_set) return B200LK_ERR_STATE;
    if (b->pending) BATCH_TRY(b, cudaStreamSynchronize(b->stream));  // the pinned staging below is still being read
    b->pending = false;
    if (nchildren > 0) {
        memcpy(b->h_child_len, child_len, sizeof(double) * (size_t)b->n * nchildren);
        BATCH_TRY(b, cudaMemcpyAsync(b->d_child_len, b->h_child_len, sizeof(double) * (size_t)b->n * nchildren,
                                     cudaMemcpyHostToDevice, b->stream));
    }
    b->bl.nchildren = nchildren;
    int rc = upload_branch_len(b, branch_len);
    if (rc) return rc;
    rc = with_batch(b, [&](b200lk_ctx *c0) {
        if (check_ends(c0, dad, node)) return B200LK_ERR_ARG;
        int r = enqueue_update(c0, nodes, nnodes, children, nchildren);
        if (r) return r;
        r = launch_branch_lnl(c0, dad, node, branch_len[0]);
        if (r) return r;
        if (want_derv) {
            r = launch_theta(c0, dad, node);
            if (r) return r;
            r = launch_derv(c0, branch_len[0]);
        }
        return r;
    });
    if (rc) return rc;
    for (b200lk_ctx *c : b->ctx) {
        c->theta_computed = want_derv != 0;
        c->n_updates += (int64_t)nnodes * c->ntot;
    }
    batch_gather_kernel<<<(b->n + 127) / 128, 128, 0, b->stream>>>(b->d_ctx, b->n, b->d_results);
    BATCH_TRY(b, cudaGetLastError());
    BATCH_TRY(b, cudaMemcpyAsync(b->h_results, b->d_results, sizeof(double) * 4 * b->n, cudaMemcpyDeviceToHost, b->stream));
    b->pending = true;
    b->pending_derv = want_derv != 0;
    return B200LK_OK;
}

extern "C" int b200lk_batch_read_results(b200lk_batch *b, double *lnl, double *df, double *ddf) {
    if (!b) return B200LK_ERR_ARG;
    if (!b->pending) return B200LK_ERR_STATE;
    BATCH_TRY(b, cudaSetDevice(b->device));
    BATCH_TRY(b, cudaStreamSynchronize(b->stream));
    b->pending = false;
    const int warn = unpack(b, lnl, b->pending_derv ? df : nullptr, b->pending_derv ? ddf : nullptr);
    if (warn & B200LK_WARN_LNL_NONFINITE) {
        int rc = repair_members(b, lnl);
        if (rc) return rc;
    }
    return warn;
}

extern "C" int b200lk_batch_evaluate(b200lk_batch *b, const b200lk_node *nodes, int32_t nnodes, const b200lk_child *children,
                                     int32_t nchildren, const double *child_len, b200lk_end dad, b200lk_end node,
                                     const double *branch_len, int want_derv, double *lnl, double *df, double *ddf) {
    int rc = b200lk_batch_evaluate_async(b, nodes, nnodes, children, nchildren, child_len, dad, node, branch_len, want_derv);
    if (rc) return rc;
    return b200lk_batch_read_results(b, lnl, df, ddf);
}

extern "C" int b200lk_batch_build_theta(b200lk_batch *b, b200lk_end dad, b200lk_end node) {
    if (!b) return B200LK_ERR_ARG;
    b->err[0] = 0;
    BATCH_TRY(b, cudaSetDevice(b->device));
    int rc = wait_for_members(b);
    if (rc) return rc;
    rc = with_batch(b, [&](b200lk_ctx *c0) {
        if (check_ends(c0, dad, node)) return B200LK_ERR_ARG;
        return launch_theta(c0, dad, node);
    });
    if (rc) return rc;
    BATCH_TRY(b, cudaStreamSynchronize(b->stream));
    for (b200lk_ctx *c : b->ctx) c->theta_computed = true;
    return B200LK_OK;
}

extern "C" int b200lk_batch_derv(b200lk_batch *b, const double *branch_len, double *df, double *ddf) {
    if (!b || !branch_len || !df || !ddf) return B200LK_ERR_ARG;
    b->err[0] = 0;
    for (b200lk_ctx *c : b->ctx)
        if (!c->theta_computed) return B200LK_ERR_STATE;
    BATCH_TRY(b, cudaSetDevice(b->device));
    int rc = upload_branch_len(b, branch_len);
    if (rc) return rc;
    rc = with_batch(b, [&](b200lk_ctx *c0) { return launch_derv(c0, branch_len[0]); });
    if (rc) return rc;
    return collect(b, nullptr, df, ddf);
}

extern "C" int b200lk_batch_lnl_from_buffer(b200lk_batch *b, const double *branch_len, double *lnl) {
    if (!b || !branch_len || !lnl) return B200LK_ERR_ARG;
    b->err[0] = 0;
    for (b200lk_ctx *c : b->ctx)
        if (!c->theta_computed) return B200LK_ERR_STATE;
    BATCH_TRY(b, cudaSetDevice(b->device));
    int rc = upload_branch_len(b, branch_len);
    if (rc) return rc;
    rc = with_batch(b, [&](b200lk_ctx *c0) { return launch_from_buffer(c0, branch_len[0]); });
    if (rc) return rc;
    return collect(b, lnl, nullptr, nullptr);
}
