// sweep.cu -- branch lengths resident on the device, so that a whole PhyloTree::optimizeAllBranches sweep
// (tree/phylotree.cpp:2689-2765 -> optimizeOneBranch :2585-2635 -> Optimization::minimizeNewton, utils/optimization.cpp:
// 422-499) is ENQUEUED without reading an optimised length back (second half).
//
// The reference re-roots at every branch in turn: one partial update across the branch just optimised, theta, the
// Newton-Raphson iteration.  The next partial update needs the length just found -- on the host, if the host forms the
// exp(lambda r t) factors.  Here the lengths live in a device table indexed by the host's own branch id:
//   * b200lk_update_partials_slots: children name a slot of the table instead of carrying a length; tables_kernel looks the
//     length up on the device (the 4-state traversal then takes its exp factors from the tables, like the batched kernel);
//   * b200lk_optimize_branch_slot: the Newton-Raphson launch starts from lengths[slot] and writes the optimum back to it;
//   * b200lk_get_lengths: ONE device->host copy of the table (and of each slot's evaluation count / status) per sweep.
#include "common.cuh"

static int ensure_lens(b200lk_ctx *c, int n) {
    if (n <= c->lens_cap) return B200LK_OK;
    double *nl = nullptr, *na = nullptr;
    CUDA_TRY(c, cudaMalloc(&nl, sizeof(double) * n));
    CUDA_TRY(c, cudaMalloc(&na, sizeof(double) * 2 * n));
    CUDA_TRY(c, cudaMemsetAsync(nl, 0, sizeof(double) * n, c->stream));
    CUDA_TRY(c, cudaMemsetAsync(na, 0, sizeof(double) * 2 * n, c->stream));
    if (c->d_lens) {
        CUDA_TRY(c, cudaMemcpyAsync(nl, c->d_lens, sizeof(double) * c->lens_cap, cudaMemcpyDeviceToDevice, c->stream));
        CUDA_TRY(c, cudaStreamSynchronize(c->stream));
        CUDA_TRY(c, cudaFree(c->d_lens));
        CUDA_TRY(c, cudaFree(c->d_lens_aux));
    }
    c->d_lens = nl;
    c->d_lens_aux = na;
    c->lens_cap = n;
    return B200LK_OK;
}

extern "C" int b200lk_set_lengths(b200lk_ctx *c, const double *len, int32_t n) {
    if (!c || !len || n < 1) return B200LK_ERR_ARG;
    SHARD_REFUSE(c, "device-resident branch lengths");
    CUDA_TRY(c, cudaSetDevice(c->device));
    int rc = ensure_lens(c, n);
    if (rc) return rc;
    CUDA_TRY(c, cudaMemcpyAsync(c->d_lens, len, sizeof(double) * n, cudaMemcpyHostToDevice, c->stream));
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));  // pageable source: the caller may reuse it
    return B200LK_OK;
}

// lengths[n]; aux (optional) [n][2]: derivative evaluations and status (1: non-finite derivative) of the last
// b200lk_optimize_branch_slot of every slot.  Blocks until everything enqueued so far has finished.
extern "C" int b200lk_get_lengths(b200lk_ctx *c, double *len, double *aux, int32_t n) {
    if (!c || !len || n < 1 || n > c->lens_cap) return B200LK_ERR_ARG;
    SHARD_REFUSE(c, "device-resident branch lengths");
    CUDA_TRY(c, cudaSetDevice(c->device));
    CUDA_TRY(c, cudaMemcpyAsync(len, c->d_lens, sizeof(double) * n, cudaMemcpyDeviceToHost, c->stream));
    if (aux) CUDA_TRY(c, cudaMemcpyAsync(aux, c->d_lens_aux, sizeof(double) * 2 * n, cudaMemcpyDeviceToHost, c->stream));
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    return B200LK_OK;
}

static int ensure_self_batch(b200lk_ctx *c) {
    if (c->self_dev) return B200LK_OK;
    BatchCtxDev d;
    memset(&d, 0, sizeof(d));
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
    CUDA_TRY(c, cudaMalloc(&c->self_dev, sizeof(BatchCtxDev)));
    CUDA_TRY(c, cudaMemcpy(c->self_dev, &d, sizeof(BatchCtxDev), cudaMemcpyHostToDevice));
    memset(&c->self_bl, 0, sizeof(c->self_bl));
    c->self_bl.d_ctx = c->self_dev;
    c->self_bl.n = 1;
    c->self_bl.max_ntot = c->ntot;
    return B200LK_OK;
}

// b200lk_update_partials with the children's branch lengths taken from the device table: len_slot[j] >= 0 names the slot of
// children[j] (its `length` field is ignored), -1 keeps children[j].length.
extern "C" int b200lk_update_partials_slots(b200lk_ctx *c, const b200lk_node *nodes, int32_t nnodes, const b200lk_child *children,
                                            int32_t nchildren, const int32_t *len_slot) {
    if (!c || !len_slot) return B200LK_ERR_ARG;
    SHARD_REFUSE(c, "device-resident branch lengths");
    if (!c->model_set || !c->tips_set) return B200LK_ERR_STATE;
    if (c->site || c->nonrev || c->batch || c->h_child_mixlen) {
        snprintf(c->err, sizeof(c->err), "device-resident branch lengths: reversible, non-site-specific contexts, one length per branch");
        return B200LK_ERR_UNSUPPORTED;
    }
    if (nnodes == 0) return B200LK_OK;
    for (int j = 0; j < nchildren; j++)
        if (len_slot[j] >= c->lens_cap) return B200LK_ERR_ARG;
    CUDA_TRY(c, cudaSetDevice(c->device));
    const bool s4 = c->tile == 32;
    if (s4) {
        // the 4-state traversal forms exp(lambda r t) on the HOST as kernel parameters; with device-resident lengths it runs
        // as a batch of one instead, which reads the factors tables_kernel computed (s4_traverse_batch_kernel)
        for (int i = 0; i < nnodes; i++)
            if (nodes[i].nchild != 2) {
                snprintf(c->err, sizeof(c->err), "device-resident branch lengths: multifurcating nodes are not supported on the 4-state path");
                return B200LK_ERR_UNSUPPORTED;
            }
        int rc = ensure_self_batch(c);
        if (rc) return rc;
        c->batch = &c->self_bl;
    }
    c->cur_len_slots = len_slot;
    int rc = enqueue_update(c, nodes, nnodes, children, nchildren);
    c->cur_len_slots = nullptr;
    if (s4) c->batch = nullptr;
    return rc;
}

// b200lk_optimize_branch without a read-back: the iteration starts from lengths[len_slot] (nr->guess is ignored) and leaves
// the optimum there; evaluations and status go beside it (b200lk_get_lengths).  theta stays valid as after
// b200lk_optimize_branch.
extern "C" int b200lk_optimize_branch_slot(b200lk_ctx *c, b200lk_end dad, b200lk_end node, const b200lk_newton *nr, int32_t len_slot) {
    if (!c || !nr || len_slot < 0 || len_slot >= c->lens_cap || !(nr->min_len <= nr->max_len) || nr->max_steps < 1) return B200LK_ERR_ARG;
    SHARD_REFUSE(c, "device-resident branch lengths");
    if (!c->model_set || !c->tips_set) return B200LK_ERR_STATE;
    if (check_ends(c, dad, node)) return B200LK_ERR_ARG;
    if (c->nonrev || c->site || c->d.n_unobserved > 0 || c->d.nptn == 0) return B200LK_ERR_UNSUPPORTED;
    CUDA_TRY(c, cudaSetDevice(c->device));
    const b200lk_end ends[2] = {dad, node};
    c->cur_newton_slot = len_slot;
    int rc = launch_newton(c, nr, ends);
    c->cur_newton_slot = -1;
    if (rc == B200LK_OK || rc == B200LK_ERR_UNSUPPORTED) c->theta_computed = true;
    return rc;
}
