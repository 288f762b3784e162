// context.cu -- device residency: creation, one-time uploads, host readers.
// Replaces the allocation side of PhyloTree::initializeAllPartialLh
// (tree/phylotree.cpp:890-934, 1054-1182): one arena of nslots partial vectors and
// scale counters, theta_all, buffer_scale_all, _pattern_lh, _pattern_lh_cat, ptn_freq,
// ptn_invar and the tip-state matrix all stay in HBM for the life of the context.
#include <stdlib.h>

#include "common.cuh"

void peer_release(b200lk_ctx *c);  // peer.cu

extern "C" const char *b200lk_version(void) { return "b200lk 0.2 (sm_100a)"; }

// The A/B and test hooks: read when a context is created (b200lk_create*), never on a launch path.
static B200Env read_env() {
    {
        B200Env e;
        const char *v = getenv("B200LK_S4_MINB");
        e.s4_minb = v ? atoi(v) : 0;
        v = getenv("B200LK_S4_SYM");
        e.s4_sym = !(v && *v == '0');
        v = getenv("B200LK_STACK_DEPTH");  // test hook: force the stack-overflow path
        e.stack_depth = v ? atoi(v) : -1;
        if (v && e.stack_depth < 0) e.stack_depth = 0;
        if (e.stack_depth > 12) e.stack_depth = 12;
        v = getenv("B200LK_LEVEL_BATCHED");
        e.level_batched = v && *v == '1';
        v = getenv("B200LK_NO_DMMA");
        e.no_dmma = v && *v == '1';
        v = getenv("B200LK_DENSE_PIPE");
        e.dense_pipe = v ? atoi(v) : 2;
        v = getenv("B200LK_DENSE_OCC");
        e.dense_occ = (v && atoi(v) == 3) ? 3 : 4;
        v = getenv("B200LK_SITE_DMMA");
        e.site_dmma = !(v && *v == '0');
        v = getenv("B200LK_NO_FUSED_TAIL");
        e.no_fused_tail = v && *v == '1';
        v = getenv("B200LK_BATCH_MINB");
        e.batch_minb = v ? atoi(v) : 0;  // 0: chosen per use
        v = getenv("B200LK_S4_TT");
        e.s4_tt = (v && *v == '0') ? 0 : ((v && *v == '2') ? 2 : 1);
        v = getenv("B200LK_S4_BULK");
        e.s4_bulk = v && *v == '1';
        v = getenv("B200LK_S4_PREFETCH");
        e.s4_prefetch = v && *v == '1';
        v = getenv("B200LK_DENSE_SPLIT_VIRT");
        e.dense_split_virt = !(v && *v == '0');
        v = getenv("B200LK_S4_BUDGET");
        e.s4_budget = v ? atoi(v) : 0;
        v = getenv("B200LK_S4_VIRTUAL");
        e.s4_virtual = !(v && *v == '0');
        v = getenv("B200LK_S4_EVICT");
        e.s4_evict = !(v && *v == '0');
        v = getenv("B200LK_S4_RING");
        e.s4_ring = v ? atoi(v) : 0;
        if (e.s4_ring > 12) e.s4_ring = 12;
        return e;
    }
}
static B200Env g_env = read_env();
const B200Env &b200_env() { return g_env; }
void b200_env_reload() { g_env = read_env(); }

extern "C" const char *b200lk_last_error(const b200lk_ctx *c) { return c ? c->err : "null context"; }

static int alloc_all(b200lk_ctx *c) {
    const b200lk_dims &d = c->d;
    const int n = d.nstates, nmix = d.nmix;
    CUDA_TRY(c, cudaSetDevice(c->device));
    CUDA_TRY(c, cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking));
    c->stream = c->own_stream;
    CUDA_TRY(c, cudaMalloc(&c->states, (size_t)d.ntaxa * c->ntot_pad));
    CUDA_TRY(c, cudaMemset(c->states, d.n_state_codes - 1, (size_t)d.ntaxa * c->ntot_pad));
    CUDA_TRY(c, cudaMalloc(&c->ptn_freq, sizeof(double) * c->ntot_pad));
    CUDA_TRY(c, cudaMalloc(&c->ptn_invar, sizeof(double) * c->ntot_pad));
    CUDA_TRY(c, cudaMemset(c->ptn_freq, 0, sizeof(double) * c->ntot_pad));
    CUDA_TRY(c, cudaMemset(c->ptn_invar, 0, sizeof(double) * c->ntot_pad));
    c->slot_doubles = (size_t)(c->tile == 32 ? ((c->ntot + 31) / 32) * 32 : c->ntot) * c->block;
    c->slot_scale_bytes = (((size_t)c->ntot * c->scale_per_ptn) + 15) & ~(size_t)15;
    // keep every slot 16-byte aligned for vector loads
    if (c->slot_doubles & 1) c->slot_doubles++;
    if (d.nslots > 0) {
        cudaError_t e = cudaMalloc(&c->partial, sizeof(double) * c->slot_doubles * d.nslots);
        if (e != cudaSuccess) {
            snprintf(c->err, sizeof(c->err), "cannot allocate %zu bytes for %d partial-likelihood slots: %s",
                     sizeof(double) * c->slot_doubles * d.nslots, d.nslots, cudaGetErrorString(e));
            cudaGetLastError();
            return B200LK_ERR_NOMEM;
        }
        CUDA_TRY(c, cudaMalloc(&c->scale, c->slot_scale_bytes * d.nslots));
        CUDA_TRY(c, cudaMemset(c->scale, 0, c->slot_scale_bytes * d.nslots));
    }
    {
        // the model arrays are ONE device block (each array 256-byte aligned) with a pinned staging mirror, so that
        // b200lk_set_model is a memcpy into pinned memory + ONE asynchronous H2D copy per model probe
        const size_t sz[7] = {(size_t)nmix * n, (size_t)nmix * n * n, (size_t)nmix * n * n, (size_t)d.ncat, (size_t)d.ncat,
                              (size_t)nmix, (size_t)d.n_state_codes * nmix * n};
        size_t off = 0;
        for (int k = 0; k < 7; k++) {
            c->m_off[k] = off;
            off += (sz[k] + 31) & ~(size_t)31;
        }
        c->m_off[7] = c->m_block_doubles = off;
        CUDA_TRY(c, cudaMalloc(&c->m_block, sizeof(double) * off));
        CUDA_TRY(c, cudaMemset(c->m_block, 0, sizeof(double) * off));
        CUDA_TRY(c, cudaMallocHost(&c->h_model_stage, sizeof(double) * off));
        memset(c->h_model_stage, 0, sizeof(double) * off);
        CUDA_TRY(c, cudaEventCreateWithFlags(&c->model_ev, cudaEventDisableTiming));
        c->m.eval = c->m_block + c->m_off[0];
        c->m.evec = c->m_block + c->m_off[1];
        c->m.inv_evec = c->m_block + c->m_off[2];
        c->m.rates = c->m_block + c->m_off[3];
        c->m.props = c->m_block + c->m_off[4];
        c->m.mixw = c->m_block + c->m_off[5];
        c->m.tip_partial_lh = c->m_block + c->m_off[6];
    }
    CUDA_TRY(c, cudaMalloc(&c->theta_all, sizeof(double) * (c->slot_doubles + 2)));
    CUDA_TRY(c, cudaMalloc(&c->buffer_scale_all, sizeof(double) * c->ntot_pad));
    CUDA_TRY(c, cudaMalloc(&c->pattern_lh, sizeof(double) * c->ntot_pad));
    CUDA_TRY(c, cudaMalloc(&c->pattern_lh_cat, sizeof(double) * c->ntot_pad * c->ncat_mix));
    CUDA_TRY(c, cudaMalloc(&c->val, sizeof(double) * 3 * c->block));
    CUDA_TRY(c, cudaMalloc(&c->tip_val, sizeof(double) * d.n_state_codes * c->block));
    // tables for a full traversal: at most 2*(ntaxa-2)+ntaxa children
    size_t per_child = (size_t)c->block * n + c->block + (size_t)d.n_state_codes * c->block;  // echild, expv, leaf
    per_child += (size_t)c->ncat_mix * c->frag_cat;
    CUDA_TRY(c, cudaMalloc(&c->m.inv_frag, sizeof(double) * nmix * std::max(c->frag_cat, B200_FRAG20)));
    c->h_evec = (double *)malloc(sizeof(double) * nmix * n * n);
    c->h_inv_evec = (double *)malloc(sizeof(double) * nmix * n * n);
    c->h_eval = (double *)malloc(sizeof(double) * nmix * n);
    c->h_rates = (double *)malloc(sizeof(double) * d.ncat);
    c->children_cap = 3 * d.ntaxa + 8;
    c->nodes_cap = 2 * d.ntaxa + 8;
    c->tables_cap = per_child * c->children_cap;
    CUDA_TRY(c, cudaMalloc(&c->tables, sizeof(double) * c->tables_cap));
    CUDA_TRY(c, cudaMalloc(&c->d_nodes, sizeof(DevNode) * c->nodes_cap));
    CUDA_TRY(c, cudaMalloc(&c->d_children, sizeof(DevChild) * c->children_cap));
    CUDA_TRY(c, cudaMallocHost(&c->h_nodes_ring, sizeof(DevNode) * c->nodes_cap * B200_STAGE_RING));
    CUDA_TRY(c, cudaMallocHost(&c->h_children_ring, sizeof(DevChild) * c->children_cap * B200_STAGE_RING));
    // tip-state row sequence of the pattern-stationary traversal (s4.cu: launch_segment): per traversal list at most
    // 3 entries per node (two leaf children and one pad row) + one chunk of rows read past the end by the last refill,
    // whatever the number of launches the list is cut into
    // rows region (taxon ids only) + 4 ints per node for the descriptors of the tip-tip tables, out of the refills' reach
    c->prog_stride = sizeof(int32_t) * ((3 * (size_t)c->nodes_cap + 512) + 4 * (size_t)c->nodes_cap);
    CUDA_TRY(c, cudaMallocHost(&c->h_prog_ring, c->prog_stride * B200_STAGE_RING));
    memset(c->h_prog_ring, 0, c->prog_stride * B200_STAGE_RING);
    CUDA_TRY(c, cudaMalloc(&c->d_prog, c->prog_stride));
    CUDA_TRY(c, cudaMemset(c->d_prog, 0, c->prog_stride));
    c->h_prog = c->h_prog_ring;
    c->max_stack_depth = b200_env().stack_depth;  // -1: chosen per kernel at launch (s4.cu, measured on cfg2)
    c->h_nodes = c->h_nodes_ring;
    c->h_children = c->h_children_ring;
    for (int i = 0; i < B200_STAGE_RING; i++)
        CUDA_TRY(c, cudaEventCreateWithFlags(&c->stage_ev[i], cudaEventDisableTiming));
    c->prof_events = new std::vector<cudaEvent_t>();
    c->sms = 148;
    cudaDeviceGetAttribute(&c->sms, cudaDevAttrMultiProcessorCount, c->device);
    c->red_blocks = c->sms * 8;
    CUDA_TRY(c, cudaMalloc(&c->red_partial, sizeof(double) * 8 * c->red_blocks));
    CUDA_TRY(c, cudaMalloc(&c->d_result, sizeof(double) * 16));
    CUDA_TRY(c, cudaMemset(c->d_result, 0, sizeof(double) * 16));
    CUDA_TRY(c, cudaMallocHost(&c->h_result, sizeof(double) * 8));
    if (c->tile == 32 || c->dense20) {  // virtual tip-tip vectors of the 4-state traversal (s4.cu) / the 20-state path (dense.cu)
        c->virt = (uint8_t *)calloc(d.nslots > 0 ? d.nslots : 1, 1);
        c->virt_desc = (S4Virt *)calloc(d.nslots > 0 ? d.nslots : 1, sizeof(S4Virt));
        if (!c->virt || !c->virt_desc) return B200LK_ERR_NOMEM;
    }
    return noalloc_setup(c);  // once per device: where the static scratch of b200lk_lnl_from_buffer_noalloc lives
}

extern "C" int b200lk_create(b200lk_ctx **out, const b200lk_dims *dims) {
    if (!out || !dims) return B200LK_ERR_ARG;
    *out = nullptr;
    if (dims->ntaxa < 2 || dims->nstates < 1 || dims->nstates > 64 || dims->ncat < 1 || dims->nmix < 1 ||
        dims->nptn < 0 || dims->nslots < 0 || dims->n_state_codes < dims->nstates + 1 ||
        dims->n_state_codes > 256 || dims->n_unobserved < 0)
        return B200LK_ERR_ARG;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0 || dims->device < 0 || dims->device >= ndev) {
        cudaGetLastError();
        return B200LK_ERR_CUDA;  // no CPU fallback on this path
    }
    b200_env_reload();
    b200lk_ctx *c = new b200lk_ctx();
    memset(c, 0, sizeof(*c));
    c->d = *dims;
    c->robust_keep = 1.0;
    c->cur_newton_slot = -1;
    c->device = dims->device;
    c->ncat_mix = dims->fused_mix_rate ? dims->ncat : dims->ncat * dims->nmix;  // tree/phylokernelnew.h:1323
    c->block = dims->nstates * c->ncat_mix;
    c->scale_per_ptn = dims->safe_numeric ? c->ncat_mix : 1;
    c->ntot = dims->nptn + dims->n_unobserved;
    c->ntot_pad = ((c->ntot + 255) / 256) * 256;  // a CTA tile of tip states never leaves its row
    if (c->ntot_pad == 0) c->ntot_pad = 256;
    c->force_level_batched = b200_env().level_batched;  // A/B switch: generic level-batched kernels only
    // 4-state models with <= 4 categories run the pattern-stationary path on the T32 layout (s4.cu)
    {
        const bool dmma = !b200_env().no_dmma && !c->force_level_batched;  // A/B switch: generic kernels for 20-state models
        c->dense20 = dims->nstates == 20 && dmma;
        c->dense_nt = (dmma && !c->dense20 && dims->nstates >= 9 && dims->nstates <= 64) ? (dims->nstates + 7) / 8 : 0;
        c->frag_cat = c->dense20 ? B200_FRAG20 : c->dense_nt * 2 * c->dense_nt * 32;
    }
    c->tile = (dims->nstates == 4 && c->ncat_mix <= 4 && !c->force_level_batched) ? 32 : 1;
    c->site = dims->site_specific == B200LK_MODEL_SITE;
    c->nonrev = dims->site_specific == B200LK_MODEL_NONREV;
    if (dims->site_specific < 0 || dims->site_specific > B200LK_MODEL_NONREV) {
        delete c;
        return B200LK_ERR_ARG;
    }
    if (c->nonrev) {  // non-reversible family: own kernels (nonrev.cu) on the pattern-major layout, no +ASC
        if (dims->n_unobserved != 0) {
            delete c;
            return B200LK_ERR_UNSUPPORTED;
        }
        c->tile = 1;
        c->dense20 = false;
        c->dense_nt = 0;
        c->frag_cat = 0;
    }
    if (c->site) {  // site-specific models: own kernels (site.cu) on the pattern-major layout
        if (dims->nmix != 1 || dims->n_unobserved != 0 || dims->nstates > 32 || dims->fused_mix_rate) {
            delete c;
            return B200LK_ERR_UNSUPPORTED;
        }
        c->tile = 1;
        c->dense20 = false;
        c->dense_nt = 0;
        c->frag_cat = 0;
    }
    int rc = alloc_all(c);
    if (rc != B200LK_OK) {
        // keep the context alive long enough for the caller to read the message? no: report and free
        fprintf(stderr, "b200lk_create: %s\n", c->err);
        b200lk_destroy(c);
        return rc;
    }
    *out = c;
    return B200LK_OK;
}

extern "C" int b200lk_destroy(b200lk_ctx *c) {
    SHARD_FORWARD(c, shard_destroy(c));
    if (!c) return B200LK_OK;
    cudaSetDevice(c->device);
    if (c->own_stream) cudaStreamSynchronize(c->own_stream);
    void *dev[] = {c->states, c->ptn_freq, c->ptn_invar, c->partial, c->scale, c->m_block, c->m.inv_frag, c->theta_all,
                   c->buffer_scale_all, c->pattern_lh, c->pattern_lh_cat, c->val, c->tip_val, c->tables,
                   c->d_nodes, c->d_children, c->red_partial, c->d_result};
    for (void *p : dev)
        if (p) cudaFree(p);
    peer_release(c);
    for (void *p2 : {(void *)c->d_lens, (void *)c->d_lens_aux, (void *)c->self_dev})
        if (p2) cudaFree(p2);
    if (c->tail_counter) cudaFree(c->tail_counter);
    if (c->h_model_stage) cudaFreeHost(c->h_model_stage);
    if (c->model_ev) cudaEventDestroy(c->model_ev);
    if (c->h_nodes_ring) cudaFreeHost(c->h_nodes_ring);
    if (c->h_children_ring) cudaFreeHost(c->h_children_ring);
    if (c->h_prog_ring) cudaFreeHost(c->h_prog_ring);
    if (c->d_prog) cudaFree(c->d_prog);
    if (c->boot_samples) cudaFree(c->boot_samples);
    if (c->d_rell) cudaFree(c->d_rell);
    if (c->h_rell) cudaFreeHost(c->h_rell);
    if (c->pmax) cudaFree(c->pmax);
    if (c->tt_tables) cudaFree(c->tt_tables);
    if (c->pair_idx) cudaFree(c->pair_idx);
    free(c->virt);
    free(c->virt_desc);
    if (c->s_eval) cudaFree(c->s_eval);
    if (c->s_evec) cudaFree(c->s_evec);
    if (c->s_inv) cudaFree(c->s_inv);
    if (c->s_tip) cudaFree(c->s_tip);
    if (c->s_tipvec) cudaFree(c->s_tipvec);
    for (double *p : {c->nr_Q, c->nr_ev, c->nr_V, c->nr_Vi, c->nr_freq, c->nr_tipvec, c->nr_T})
        if (p) cudaFree(p);
    for (double *p : {c->d_child_mixlen, c->d_branch_mixlen, c->mix_red, c->mix_result, c->const_df, c->const_ddf})
        if (p) cudaFree(p);
    if (c->rob_buf) cudaFree(c->rob_buf);
    if (c->rob_temp) cudaFree(c->rob_temp);
    if (c->eig_in) cudaFree(c->eig_in);
    if (c->eig_host) cudaFreeHost(c->eig_host);
    free(c->h_evec);
    free(c->h_inv_evec);
    free(c->h_eval);
    free(c->h_rates);
    for (int i = 0; i < B200_STAGE_RING; i++)
        if (c->stage_ev[i]) cudaEventDestroy(c->stage_ev[i]);
    if (c->prof_events) {
        for (cudaEvent_t e : *c->prof_events) cudaEventDestroy(e);
        delete c->prof_events;
    }
    if (c->h_result) cudaFreeHost(c->h_result);
    if (c->own_stream) cudaStreamDestroy(c->own_stream);
    delete c;
    return B200LK_OK;
}

extern "C" int b200lk_set_stream(b200lk_ctx *c, void *stream) {
    SHARD_REFUSE(c, "a caller-owned stream (every shard runs on its own device and stream)");
    if (!c) return B200LK_ERR_ARG;
    c->stream = stream ? (cudaStream_t)stream : c->own_stream;
    return B200LK_OK;
}

extern "C" int b200lk_synchronize(b200lk_ctx *c) {
    SHARD_FORWARD(c, shard_synchronize(c));
    if (!c) return B200LK_ERR_ARG;
    CUDA_TRY(c, cudaSetDevice(c->device));
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    return B200LK_OK;
}

int upload_tips_pitched(b200lk_ctx *c, const uint8_t *states, size_t spitch, const uint8_t *unobserved, size_t upitch,
                        const double *ptn_freq) {
    if (!c || (!states && c->d.nptn) || !ptn_freq || (c->d.n_unobserved && !unobserved)) return B200LK_ERR_ARG;
    CUDA_TRY(c, cudaSetDevice(c->device));
    const int64_t nptn = c->d.nptn, nun = c->d.n_unobserved;
    if (nptn)
        CUDA_TRY(c, cudaMemcpy2DAsync(c->states, c->ntot_pad, states, spitch, nptn, c->d.ntaxa,
                                      cudaMemcpyHostToDevice, c->stream));
    if (nun)
        CUDA_TRY(c, cudaMemcpy2DAsync(c->states + nptn, c->ntot_pad, unobserved, upitch, nun, c->d.ntaxa,
                                      cudaMemcpyHostToDevice, c->stream));
    CUDA_TRY(c, cudaMemcpyAsync(c->ptn_freq, ptn_freq, sizeof(double) * nptn, cudaMemcpyHostToDevice, c->stream));
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    c->nsite = 0.0;
    for (int64_t p = 0; p < nptn; p++) c->nsite += ptn_freq[p];
    c->tips_set = true;
    return B200LK_OK;
}

extern "C" int b200lk_upload_tips(b200lk_ctx *c, const uint8_t *states, const uint8_t *unobserved,
                                  const double *ptn_freq) {
    SHARD_FORWARD(c, shard_upload_tips(c, states, unobserved, ptn_freq));
    if (!c) return B200LK_ERR_ARG;
    return upload_tips_pitched(c, states, (size_t)c->d.nptn, unobserved, (size_t)c->d.n_unobserved, ptn_freq);
}

extern "C" int b200lk_set_ptn_invar(b200lk_ctx *c, const double *ptn_invar) {
    SHARD_FORWARD(c, shard_set_ptn_invar(c, ptn_invar));
    if (!c || !ptn_invar) return B200LK_ERR_ARG;
    CUDA_TRY(c, cudaSetDevice(c->device));
    CUDA_TRY(c, cudaMemcpyAsync(c->ptn_invar, ptn_invar, sizeof(double) * c->ntot, cudaMemcpyHostToDevice, c->stream));
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    c->theta_computed = false;
    return B200LK_OK;
}

extern "C" int b200lk_set_model(b200lk_ctx *c, const double *eval, const double *evec, const double *inv_evec,
                                const double *rates, const double *props, const double *mixw,
                                const double *tip_partial_lh) {
    SHARD_FORWARD(c, shard_set_model(c, eval, evec, inv_evec, rates, props, mixw, tip_partial_lh));
    if (!c || !eval || !evec || !inv_evec || !rates || !props || !mixw || !tip_partial_lh) return B200LK_ERR_ARG;
    if (c->site || c->nonrev) return B200LK_ERR_STATE;  // these contexts take b200lk_set_site_models / _set_model_nonrev
    CUDA_TRY(c, cudaSetDevice(c->device));
    const int n = c->d.nstates, nmix = c->d.nmix;
    cudaStream_t s = c->stream;
    // one pinned staging block, one H2D copy; the previous probe's copy must have left the staging block first
    if (c->model_ev_used) CUDA_TRY(c, cudaEventSynchronize(c->model_ev));
    double *st = c->h_model_stage;
    memcpy(st + c->m_off[0], eval, sizeof(double) * nmix * n);
    memcpy(st + c->m_off[1], evec, sizeof(double) * nmix * n * n);
    memcpy(st + c->m_off[2], inv_evec, sizeof(double) * nmix * n * n);
    memcpy(st + c->m_off[3], rates, sizeof(double) * c->d.ncat);
    memcpy(st + c->m_off[4], props, sizeof(double) * c->d.ncat);
    memcpy(st + c->m_off[5], mixw, sizeof(double) * nmix);
    memcpy(st + c->m_off[6], tip_partial_lh, sizeof(double) * c->d.n_state_codes * nmix * n);
    CUDA_TRY(c, cudaMemcpyAsync(c->m_block, st, sizeof(double) * c->m_block_doubles, cudaMemcpyHostToDevice, s));
    CUDA_TRY(c, cudaEventRecord(c->model_ev, s));
    c->model_ev_used = true;
    if (c->dense20 || c->dense_nt) {
        int rc = launch_inv_frag(c);
        if (rc) return rc;
    }
    // no stream sync: the caller's arrays were copied into the staging block and may be reused at once
    memcpy(c->h_evec, evec, sizeof(double) * nmix * n * n);
    memcpy(c->h_inv_evec, inv_evec, sizeof(double) * nmix * n * n);
    memcpy(c->h_eval, eval, sizeof(double) * nmix * n);
    memcpy(c->h_rates, rates, sizeof(double) * c->d.ncat);
    c->s4_sym = c->tile == 32 && b200_env().s4_sym && s4_detect_sym(c, c->s4_pi);
    c->model_set = true;
    c->theta_computed = false;
    return B200LK_OK;
}

// ---- host readers (lazy D2H of device-resident results) --------------------
static int d2h(b200lk_ctx *c, void *dst, const void *src, size_t bytes) {
    CUDA_TRY(c, cudaSetDevice(c->device));
    CUDA_TRY(c, cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, c->stream));
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    return B200LK_OK;
}

extern "C" int b200lk_fetch_pattern_lh(b200lk_ctx *c, double *out) {
    SHARD_FORWARD(c, shard_fetch_pattern_lh(c, out));
    if (!c || !out) return B200LK_ERR_ARG;
    return d2h(c, out, c->pattern_lh, sizeof(double) * c->d.nptn);
}
extern "C" int b200lk_fetch_pattern_lh_cat(b200lk_ctx *c, double *out) {
    SHARD_FORWARD(c, shard_fetch_pattern_lh_cat(c, out));
    if (!c || !out) return B200LK_ERR_ARG;
    return d2h(c, out, c->pattern_lh_cat, sizeof(double) * c->d.nptn * c->ncat_mix);
}
extern "C" int b200lk_fetch_scale_num(b200lk_ctx *c, int32_t slot, uint8_t *out) {
    SHARD_FORWARD(c, shard_fetch_scale_num(c, slot, out));
    if (!c || !out || slot < 0 || slot >= c->d.nslots) return B200LK_ERR_ARG;
    if (int rc = s4_realize_slot(c, slot)) return rc;
    return d2h(c, out, slot_scale(c, slot), (size_t)c->d.nptn * c->scale_per_ptn);
}
// device layout -> the host's pattern-major [nptn][ncat_mix][nstates]
static int d2h_vector(b200lk_ctx *c, double *out, const double *dev) {
    if (c->tile == 1) return d2h(c, out, dev, sizeof(double) * c->d.nptn * c->block);
    std::vector<double> tmp(c->slot_doubles);
    int rc = d2h(c, tmp.data(), dev, sizeof(double) * c->slot_doubles);
    if (rc) return rc;
    for (int64_t p = 0; p < c->d.nptn; p++)
        for (int e = 0; e < c->block; e++) out[(size_t)p * c->block + e] = tmp[elem_off(p, e, c->block, c->tile)];
    return B200LK_OK;
}
extern "C" int b200lk_fetch_partial(b200lk_ctx *c, int32_t slot, double *out) {
    SHARD_FORWARD(c, shard_fetch_partial(c, slot, out));
    if (!c || !out || slot < 0 || slot >= c->d.nslots) return B200LK_ERR_ARG;
    if (int rc = s4_realize_slot(c, slot)) return rc;
    return d2h_vector(c, out, slot_partial(c, slot));
}
extern "C" int b200lk_fetch_theta(b200lk_ctx *c, double *theta, double *bscale) {
    SHARD_FORWARD(c, shard_fetch_theta(c, theta, bscale));
    if (!c || !theta) return B200LK_ERR_ARG;
    int rc = d2h_vector(c, theta, c->theta_all);
    if (rc == B200LK_OK && bscale) rc = d2h(c, bscale, c->buffer_scale_all, sizeof(double) * c->d.nptn);
    return rc;
}
extern "C" int b200lk_get_counters(b200lk_ctx *c, int64_t *updates, int64_t *launches) {
    SHARD_FORWARD(c, shard_get_counters(c, updates, launches));
    if (!c) return B200LK_ERR_ARG;
    if (updates) *updates = c->n_updates;
    if (launches) *launches = c->n_launches;
    return B200LK_OK;
}
extern "C" int b200lk_reset_counters(b200lk_ctx *c) {
    SHARD_FORWARD(c, shard_reset_counters(c));
    if (!c) return B200LK_ERR_ARG;
    c->n_updates = c->n_launches = c->n_min_bytes = 0;
    return B200LK_OK;
}
extern "C" int b200lk_get_min_traffic(b200lk_ctx *c, int64_t *bytes) {
    if (c && c->shards && bytes) return shard_get_min_traffic(c, bytes);
    if (!c || !bytes) return B200LK_ERR_ARG;
    *bytes = c->n_min_bytes;
    return B200LK_OK;
}
extern "C" int b200lk_result_device_ptr(b200lk_ctx *c, void **dptr) {
    SHARD_REFUSE(c, "a single device result buffer (the shards' sums are added on the host)");
    if (!c || !dptr) return B200LK_ERR_ARG;
    *dptr = c->d_result;
    return B200LK_OK;
}

// ---- per-launch timing of the partial-update kernels -------------------------
extern "C" int b200lk_set_profiling(b200lk_ctx *c, int on) {
    SHARD_FORWARD(c, shard_set_profiling(c, on));
    if (!c) return B200LK_ERR_ARG;
    c->profiling = on != 0;
    return B200LK_OK;
}

extern "C" int b200lk_get_profile(b200lk_ctx *c, double *partial_ms, int64_t *partial_launches, int reset) {
    SHARD_FORWARD(c, shard_get_profile(c, partial_ms, partial_launches, reset));
    if (!c) return B200LK_ERR_ARG;
    CUDA_TRY(c, cudaSetDevice(c->device));
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    std::vector<cudaEvent_t> &ev = *c->prof_events;
    for (size_t i = 0; i + 1 < ev.size(); i += 2) {
        float ms = 0.f;
        CUDA_TRY(c, cudaEventElapsedTime(&ms, ev[i], ev[i + 1]));
        c->prof_ms += ms;
        c->prof_launches++;
    }
    for (cudaEvent_t e : ev) cudaEventDestroy(e);
    ev.clear();
    if (partial_ms) *partial_ms = c->prof_ms;
    if (partial_launches) *partial_launches = c->prof_launches;
    if (reset) { c->prof_ms = 0.0; c->prof_launches = 0; }
    return B200LK_OK;
}

// ---- UFBoot RELL ---------------------------------------------------------------------------------------------
int upload_boot_samples_pitched(b200lk_ctx *c, const float *samples, size_t pitch_floats, int32_t nsamples) {
    if (!c || !samples || nsamples < 1) return B200LK_ERR_ARG;
    CUDA_TRY(c, cudaSetDevice(c->device));
    if (c->boot_samples) CUDA_TRY(c, cudaFree(c->boot_samples));
    if (c->d_rell) CUDA_TRY(c, cudaFree(c->d_rell));
    if (c->h_rell) CUDA_TRY(c, cudaFreeHost(c->h_rell));
    c->boot_samples = nullptr;
    c->d_rell = c->h_rell = nullptr;
    c->nboot = 0;
    const size_t row = sizeof(float) * (size_t)c->d.nptn, bytes = row * nsamples;
    if (cudaMalloc(&c->boot_samples, bytes ? bytes : 4) != cudaSuccess) {
        cudaGetLastError();
        snprintf(c->err, sizeof(c->err), "cannot allocate %zu bytes for %d bootstrap resamples", bytes, nsamples);
        return B200LK_ERR_NOMEM;
    }
    CUDA_TRY(c, cudaMalloc(&c->d_rell, sizeof(double) * nsamples));
    CUDA_TRY(c, cudaMallocHost(&c->h_rell, sizeof(double) * nsamples));
    if (row)
        CUDA_TRY(c, cudaMemcpy2DAsync(c->boot_samples, row, samples, sizeof(float) * pitch_floats, row, nsamples,
                                      cudaMemcpyHostToDevice, c->stream));
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    c->nboot = nsamples;
    return B200LK_OK;
}

extern "C" int b200lk_upload_boot_samples(b200lk_ctx *c, const float *samples, int32_t nsamples) {
    SHARD_FORWARD(c, shard_upload_boot_samples(c, samples, nsamples));
    if (!c) return B200LK_ERR_ARG;
    return upload_boot_samples_pitched(c, samples, (size_t)c->d.nptn, nsamples);
}

extern "C" int b200lk_rell(b200lk_ctx *c, double *rell) {
    SHARD_FORWARD(c, shard_rell(c, rell));
    if (!c || !rell) return B200LK_ERR_ARG;
    if (c->nboot < 1) {
        snprintf(c->err, sizeof(c->err), "b200lk_rell: upload the bootstrap resamples first");
        return B200LK_ERR_STATE;
    }
    CUDA_TRY(c, cudaSetDevice(c->device));
    int rc = launch_rell(c);
    if (rc) return rc;
    CUDA_TRY(c, cudaMemcpyAsync(c->h_rell, c->d_rell, sizeof(double) * c->nboot, cudaMemcpyDeviceToHost, c->stream));
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    memcpy(rell, c->h_rell, sizeof(double) * c->nboot);
    return B200LK_OK;
}
