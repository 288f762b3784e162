// common.cuh -- context layout and helpers shared by the CUDA translation units.
//
// HBM layout (all private to the engine; the host sees only the C ABI):
//   states      uint8  [ntaxa][ntot_pad]             tip state codes, taxon-major (coalesced per taxon)
//   ptn_freq    double [ntot_pad]   ptn_invar double [ntot_pad]
//   partial     double [nslots][ntot][block]         element (p,c,x) at elem_off(p, c*nstates + x): pattern-major
//                                                    (tile == 1), or "T32" for the 4-state path (tile == 32):
//                                                    32-pattern tiles, pair j of pattern l at j*64 + l*2
//   scale       uint8  [nslots][ntot] (NORM) | [nslots][ntot][ncat_mix] (SAFE)
//   theta_all   double [ntot][block]   buffer_scale_all double [ntot]
//   pattern_lh  double [ntot]          pattern_lh_cat double [ntot][ncat_mix]
//   tables      per child of the current traversal: echild[ncat_mix][n][n] (+ leaf[codes][ncat_mix][n])
// ntot = nptn + n_unobserved (+ASC patterns appended after the observed ones,
// tree/phylokernelnew.h:1314-1316 with VectorClass::size() == 1).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <vector>

#include "../../include/b200lk.h"

#define B200_SCALING_THRESHOLD 8.636168555094444625386e-78  // 2^-256, tree/phylotree.h:77
#define B200_LOG_SCALING_THRESHOLD (-177.4456782233459932741)  // tree/phylotree.h:80
#define B200_TWO_256 1.157920892373161954235709850086879e77    // 2^256: ldexp(x,256) as an exact multiply

#define B200_STAGE_RING 4

struct DevModel {
    double *eval, *evec, *inv_evec, *rates, *props, *mixw, *tip_partial_lh;
    double *inv_frag;  // 20-state tensor path: inv_evec per class in mma fragment order [nmix][15][32]
};
#define B200_FRAG20 480  // doubles of one 20x20 matrix in fragment order: 3 n-tiles x 5 k-steps x 32 lanes

// One child of one traversal entry, device view.
struct DevChild {
    int32_t kind;        // B200LK_CHILD_LEAF / INTERNAL
    int32_t id;          // taxon or slot
    double length;
    int64_t echild_off;  // offset (doubles) into tables
    int64_t leaf_off;    // offset (doubles) into tables, -1 if internal
    int64_t frag_off;    // 20-state tensor path: echild in mma fragment order (dense.cu), -1 if unused
    int64_t aux_off;     // left child of a tip-tip node (20-state path): offset of the node's state-pair table in tt_tables
    int64_t src;         // position in the caller's children array (entries are reordered leaf-first): indexes
                         // BatchLaunch::d_child_len
    int64_t len_slot;    // >= 0: the branch length is lengths[len_slot] of the context's device-resident table (sweep.cu)
    // 20-state path: >= 0 when this INTERNAL child is a tip-tip node left virtual (its vector is not in its slot): offset
    // (doubles) of its state-pair table in tt_tables, and offset (elements) of its per-pattern pair indices in pair_idx
    int64_t vtt_off, vpair_off;
};

struct DevNode {
    int32_t out_slot, child_begin, nchild, kind;  // kind: 0 TT, 1 TI, 2 II, 3 multifurcating
};

// ---- batched launches over the contexts of a supermatrix (batch.cu) ----------------------------------------------
// Device view of one context of a batch: what a kernel of a batched launch substitutes for the launch's own
// pointers and sizes (the traversal STRUCTURE -- nodes, child kinds, slots, table offsets -- is shared by all
// contexts of a batch and comes from the first one).
struct BatchCtxDev {
    double *partial;
    uint8_t *scale;
    const uint8_t *states;
    const double *ptn_invar, *ptn_freq;
    double *tables;
    DevModel m;
    double *theta, *buffer_scale_all, *pattern_lh, *pattern_lh_cat, *val, *tip_val, *red_partial, *d_result;
    size_t slot_doubles, slot_scale_bytes;
    int64_t ntot, nptn, ntot_pad;
};
struct BatchLaunch {           // non-null in b200lk_ctx::batch while a batched call is being enqueued
    const BatchCtxDev *d_ctx;  // [n]
    int n;
    int64_t max_ntot;          // grids are sized for the largest context
    const double *d_child_len; // [n][nchildren] branch length to every child, per context
    int nchildren;
    const double *d_branch_len;  // [n] length of the branch the likelihood is taken at
    const double *d_lens;        // device-resident branch-length table of a single context (sweep.cu), or null
};

// b200lk_create_sharded (shard.cu): one problem split by site pattern over several devices of THIS process.  The parent
// context owns no device memory; every entry point fans out to the per-device contexts and sums on the host in
// device-index order.
struct ShardSet {
    std::vector<b200lk_ctx *> ctx;    // one ordinary context per device, in the order the caller listed the devices
    std::vector<int64_t> begin, end;  // observed-pattern range [begin, end) of each shard (whole 32-pattern tiles)
};

struct S4Virt { int32_t slot, tt_off, ta, tb; };  // a virtual tip-tip vector: rows (state of taxon ta, state of taxon tb) of the table at tt_off

// process-wide A/B and test hooks, read from the environment when a context is created (context.cu) -- never on a launch path
struct B200Env {
    int s4_minb;       // B200LK_S4_MINB: resident-CTA instantiation of s4_traverse_kernel (8 / 10 / 12); 0 = default
    bool s4_sym;       // B200LK_S4_SYM=0 keeps the general 4-state kernel
    int stack_depth;   // B200LK_STACK_DEPTH: shared-memory stack entries per thread (-1 = chosen at launch)
    bool level_batched, no_dmma;  // B200LK_LEVEL_BATCHED=1, B200LK_NO_DMMA=1: generic kernels only
    int dense_pipe, dense_occ;    // B200LK_DENSE_PIPE (0 direct, 1 cp.async, 2 register prefetch), B200LK_DENSE_OCC (3 / 4)
    bool site_dmma;               // B200LK_SITE_DMMA=0 keeps the shuffle kernel for 20-state site models
    bool no_fused_tail;           // B200LK_NO_FUSED_TAIL=1: separate lnl / theta / derv launches in b200lk_evaluate*
    int batch_minb;               // B200LK_BATCH_MINB=2|8: force one build of s4_traverse_batch_kernel (A/B); 0 = per use
    int s4_tt;                    // B200LK_S4_TT=0: tip-tip nodes (4- and 20-state paths) computed, not looked up in state-pair tables; 2: looked up at any size (tests)
    bool s4_bulk;                 // B200LK_S4_BULK=1: node vectors leave through cp.async.bulk instead of the threads' STG (slower: A/B log)
    bool dense_split_virt;        // B200LK_DENSE_SPLIT_VIRT=0: a level's nodes with and without a virtual child in ONE launch of the larger build (A/B)
    int s4_budget;                // B200LK_S4_BUDGET: nodes per launch of the 4-state traversal (A/B: more, shorter launches); 0 = what the window holds
    bool s4_virtual;              // B200LK_S4_VIRTUAL=0: every tip-tip vector is written by the traversal that computes it
    bool s4_evict;                // B200LK_S4_EVICT=0: a full stack keeps what it holds (first come) instead of the soonest-wanted vectors
    bool s4_prefetch;             // B200LK_S4_PREFETCH=1: L1 prefetch of the next node's tip-child table rows (slower: A/B log)
    int s4_ring;                  // B200LK_S4_RING: shared-memory slots of the bulk-store traversal (0 = default 3)
};
const B200Env &b200_env();
void b200_env_reload();  // b200lk_create*: the hooks in force are those of the latest context creation

struct b200lk_ctx {
    ShardSet *shards;  // non-null: a sharded parent
    int sms;                          // multiprocessors of the device
    double *m_block, *h_model_stage;  // the DevModel arrays as ONE device block + its pinned staging mirror
    size_t m_block_doubles, m_off[8];
    cudaEvent_t model_ev;             // recorded after the H2D copy of the staging mirror
    bool model_ev_used;
    const BatchLaunch *batch;
    b200lk_dims d;
    int ncat_mix, block, scale_per_ptn;
    bool dense20;  // nstates == 20: FP64 tensor-core (DMMA) partial update, dense.cu
    int dense_nt;  // 9 <= nstates <= 64 (not 20): 8-state tiles of the generic tensor-core path (densen.cu); else 0
    int frag_cat;  // doubles of one category's P-factor table in mma fragment order (0: no tensor path)
    int tile;  // HBM layout of partial vectors and theta: 1 = pattern-major, 32 = T32 (s4.cu)
    double *h_evec, *h_inv_evec, *h_eval, *h_rates;  // host copies of the model (kernel parameters of the 4-state path)
    int64_t ntot, ntot_pad;
    int device;
    cudaStream_t stream, own_stream;
    char err[512];

    uint8_t *states;
    double *ptn_freq, *ptn_invar;
    double *partial;   // arena
    uint8_t *scale;    // arena
    size_t slot_doubles, slot_scale_bytes;
    DevModel m;
    double *theta_all, *buffer_scale_all, *pattern_lh, *pattern_lh_cat;
    double *val;        // [3][block] branch coefficient tables
    double *tip_val;    // [codes][block] val * tip_partial_lh (tip-internal branch lnL)
    double *tables;     // per-traversal P-factor tables
    size_t tables_cap;  // doubles
    DevNode *d_nodes;
    DevChild *d_children;
    int nodes_cap, children_cap;
    DevNode *h_nodes;      // pinned staging: slice of the ring used by the call in flight
    DevChild *h_children;  // pinned staging
    DevNode *h_nodes_ring;      // B200_STAGE_RING slices of nodes_cap / children_cap entries, so that
    DevChild *h_children_ring;  // asynchronous evaluations never overwrite a copy still in flight
    size_t prog_stride;
    void *h_prog, *h_prog_ring, *d_prog;  // device program of the pattern-stationary traversal (traverse.cu)
    bool trav_attr_set, dense_attr_set, force_level_batched;
    int max_stack_depth;  // shared-memory stack entries per thread (B200LK_STACK_DEPTH; -1 = chosen at launch)
    cudaEvent_t stage_ev[B200_STAGE_RING];
    bool stage_used[B200_STAGE_RING];
    int stage_idx;
    // optional per-launch timing of the partial-update kernels (b200lk_set_profiling)
    bool profiling;
    std::vector<cudaEvent_t> *prof_events;  // start/stop pairs
    double prof_ms;
    int64_t prof_launches;
    float *boot_samples;  // UFBoot: [nboot][nptn] resampled pattern frequencies (b200lk_upload_boot_samples)
    double *d_rell, *h_rell;
    int nboot;
    double *tt_tables;  // 20-state path: per tip-tip node, the partial vector of every (left state, right state) pair
    size_t tt_cap;      // doubles
    size_t s4_ttdesc_used;  // 4-state path: tip-tip table descriptors staged for the traversal being enqueued
    uint16_t *pair_idx;     // 20-state path: per virtual tip-tip node, sa * codes + sb of every pattern (dense.cu)
    size_t pair_cap;        // elements
    uint8_t *virt;          // host [nslots]: 1 = the slot's vector is virtual (4-state: s4.cu, 20-state: dense.cu; s4_realize)
    S4Virt *virt_desc;      // host [nslots]: where a virtual slot's rows are
    int n_virtual;
    const b200lk_node *cur_nodes;  // non-null only inside enqueue_update's 4-state branch: the caller's list ...
    const int *cur_nreaders;       // ... and how many of its nodes read each node's output
    size_t s4_tt_used;  // 4-state path: doubles of tt_tables handed to the tip-tip nodes of the traversal being enqueued
    unsigned long long *pmax;  // mixture launches of dense.cu: per (node of the level, pattern) max|x| bits
    size_t pmax_bytes;
    bool site;  // dims.site_specific: per-pattern eigen-systems (site.cu); no P-factor tables, pattern-major layout
    double *s_eval, *s_evec, *s_inv, *s_tip, *s_tipvec;  // [nptn][n], [nptn][n][n] x2, [ntaxa][nptn][n], [codes][n]
    // non-reversible kernel family (nonrev.cu): dims.site_specific == B200LK_MODEL_NONREV
    bool nonrev, rooted;
    double *nr_Q, *nr_ev, *nr_V, *nr_Vi, *nr_freq, *nr_tipvec;  // Q, complex eigen-system, frequencies, tip vectors
    double *nr_T;                                               // branch tables T0/T1/T2 + tip-dad tables
    b200lk_end nr_dad, nr_node;                                 // ends recorded by b200lk_build_theta
    double *eig_in, *eig_host;  // staging of b200lk_set_model_rates (eigen.cu): device inputs, pinned host mirror
    double *red_partial;   // [3][red_blocks] block partial sums
    int red_blocks;
    double nsite;          // sum of ptn_freq (aln->getNSite()), for the +ASC correction
    double *d_result;      // [16] lnL, df, ddf, -, optx, d2l, evals, status, prob_const, df_const, ddf_const
    double *h_result;      // pinned
    // heterotachy (b200lk_*_mixlen, branch.cu): per-class lengths of the call being enqueued
    const double *h_child_mixlen;     // host [nchildren][ncat]; non-null only inside b200lk_update_partials_mixlen
    double *d_child_mixlen;           // its device copy [children_cap][ncat]
    double *d_branch_mixlen;          // device [ncat]: class lengths of the branch of a *_mixlen branch-level call
    const double *cur_branch_mixlen;  // == d_branch_mixlen while such a call is being enqueued, else null
    double *mix_red, *mix_result;     // derv_mixlen: [K][red_blocks] partial sums, [K] sums (K = 1 + ncat + ncat^2)
    // Holder's +ASC (b200lk_set_asc): per-pattern derivative terms of the unobserved patterns
    bool asc_holder;
    double *const_df, *const_ddf;  // [n_unobserved]
    // robust phylogeny (b200lk_set_robust, robust.cu)
    double robust_keep;  // params->robust_phy_keep; 1.0 = off
    bool robust_median;  // params->robust_median
    double *rob_buf;     // sorted per-pattern lnL, sorted frequencies, their inclusive scan, 2 picks
    void *rob_temp;
    size_t rob_temp_bytes;
    bool s4_sym;       // 4-state path: inv_evec[i][x] == evec[x][i] * s4_pi[x] (s4_detect_sym at set_model)
    double s4_pi[4];
    bool model_set, tips_set, theta_computed;
    int theta_dad_kind;
    // device-resident branch lengths (sweep.cu): a whole optimizeAllBranches sweep is enqueued without reading a length back
    double *d_lens, *d_lens_aux;       // [lens_cap] lengths, [lens_cap][2] evaluations and status of the last optimisation
    int lens_cap;
    const int32_t *cur_len_slots;      // host [nchildren]; non-null only inside b200lk_update_partials_slots
    int cur_newton_slot;               // slot the Newton-Raphson launch being enqueued reads its guess from / writes to; -1 none
    BatchCtxDev *self_dev;             // this context as a batch of one (4-state path: exp factors from the tables)
    BatchLaunch self_bl;
    unsigned *tail_counter;  // eval_tail_kernel: CTAs finished (device)
    // exchange of [lnL, df, ddf] between the pattern shards of several PROCESSES over NVLink peer memory (peer.cu):
    // every rank's exchange buffer is mapped into every other rank (CUDA IPC); the kernel that finishes an evaluation
    // stores its sums into all peers' buffers and adds what the peers stored, in rank order
    double *peer_buf;              // this rank's buffer: vals[2][B200_PEER_MAX][4], then flags[2][B200_PEER_MAX]
    double **d_peer_bufs;          // device array [world]: every rank's buffer as mapped here
    void *peer_mapped[64];         // what cudaIpcOpenMemHandle returned (to close)
    int peer_rank, peer_world;
    unsigned long long peer_seq;   // evaluations exchanged so far
    int64_t n_updates, n_launches;
    // minimum DRAM bytes of the partial-update launches enqueued since the last reset, counted from the node programs
    // actually launched: every vector written once (+ scale bytes), every child vector that a launch has to fetch
    // from HBM (level-batched kernels: all internal children; 4-state traversal: only children produced by an
    // earlier launch; a vector evicted from the per-thread stack is re-read from L2), one byte per tip state read
    int64_t n_min_bytes;
};

#define CUDA_TRY(ctx, expr)                                                                      \
    do {                                                                                         \
        cudaError_t e__ = (expr);                                                                \
        if (e__ != cudaSuccess) {                                                                \
            snprintf((ctx)->err, sizeof((ctx)->err), "%s:%d %s: %s", __FILE__, __LINE__, #expr,  \
                     cudaGetErrorString(e__));                                                   \
            return B200LK_ERR_CUDA;                                                              \
        }                                                                                        \
    } while (0)

// offset (doubles) of element e of pattern p inside one partial vector / theta
__host__ __device__ __forceinline__ size_t elem_off(int64_t p, int e, int block, int tile) {
    if (tile == 1) return (size_t)p * block + e;
    return (size_t)(p >> 5) * ((size_t)32 * block) + (size_t)(e >> 1) * 64 + (size_t)(p & 31) * 2 + (e & 1);
}

// B200LK_INFO_DERV_NOT_AS_REFERENCE: odd nstates with an even number of categories (include/b200lk.h)
static inline int derv_info_bits(const b200lk_ctx *c) {
    return ((c->d.nstates & 1) && !(c->ncat_mix & 1) && !c->nonrev) ? B200LK_INFO_DERV_NOT_AS_REFERENCE : 0;
}
static inline double *slot_partial(const b200lk_ctx *c, int s) { return c->partial + (size_t)s * c->slot_doubles; }
static inline uint8_t *slot_scale(const b200lk_ctx *c, int s) { return c->scale + (size_t)s * c->slot_scale_bytes; }

// record one end of a timed partial-update launch (pairs: before, after)
static inline void prof_mark(b200lk_ctx *c) {
    if (!c->profiling) return;
    cudaEvent_t e;
    if (cudaEventCreate(&e) != cudaSuccess) return;
    cudaEventRecord(e, c->stream);
    c->prof_events->push_back(e);
}

// launchers implemented in tables.cu / partial.cu / branch.cu
// shard.cu: what the entry points forward to when the context is a sharded parent
#define SHARD_FORWARD(c, call) \
    if ((c) && (c)->shards) return call
#define SHARD_REFUSE(c, what) \
    if ((c) && (c)->shards) return shard_refuse(c, what)
int shard_refuse(b200lk_ctx *c, const char *what);
int shard_destroy(b200lk_ctx *c);
int shard_upload_tips(b200lk_ctx *c, const uint8_t *states, const uint8_t *unobserved, const double *ptn_freq);
int shard_set_ptn_invar(b200lk_ctx *c, const double *v);
int shard_set_ptn_freq(b200lk_ctx *c, const double *v);
int shard_set_model(b200lk_ctx *c, const double *eval, const double *evec, const double *inv_evec, const double *rates,
                    const double *props, const double *mixw, const double *tip);
int shard_set_model_rates(b200lk_ctx *c, const double *exch, const double *freqs, const double *tot, const double *tipvec,
                          const double *rates, const double *props, const double *mixw);
int shard_set_robust(b200lk_ctx *c, double keep, int32_t median);
int shard_update_partials(b200lk_ctx *c, const b200lk_node *nodes, int32_t nnodes, const b200lk_child *children, int32_t nchildren);
int shard_branch_lnl(b200lk_ctx *c, b200lk_end dad, b200lk_end node, double length, double *lnl);
int shard_build_theta(b200lk_ctx *c, b200lk_end dad, b200lk_end node);
int shard_derv(b200lk_ctx *c, double length, double *df, double *ddf);
int shard_lnl_from_buffer(b200lk_ctx *c, double length, double *lnl);
int shard_evaluate_async(b200lk_ctx *c, const b200lk_node *nodes, int32_t nnodes, const b200lk_child *children,
                         int32_t nchildren, b200lk_end dad, b200lk_end node, double length, int want_lnl, int want_derv);
int shard_read_results(b200lk_ctx *c, double *lnl, double *df, double *ddf);
int shard_optimize_branch(b200lk_ctx *c, b200lk_end dad, b200lk_end node, const b200lk_newton *nr, double *optx, double *d2l,
                          int32_t *nsteps);
int shard_fetch_pattern_lh(b200lk_ctx *c, double *out);
int shard_fetch_pattern_lh_cat(b200lk_ctx *c, double *out);
int shard_fetch_scale_num(b200lk_ctx *c, int32_t slot, uint8_t *out);
int shard_fetch_partial(b200lk_ctx *c, int32_t slot, double *out);
int shard_fetch_theta(b200lk_ctx *c, double *theta, double *bscale);
int shard_fetch_model(b200lk_ctx *c, double *eval, double *evec, double *inv_evec);
int shard_get_counters(b200lk_ctx *c, int64_t *updates, int64_t *launches);
int shard_reset_counters(b200lk_ctx *c);
int shard_get_min_traffic(b200lk_ctx *c, int64_t *bytes);
int shard_synchronize(b200lk_ctx *c);
int shard_set_profiling(b200lk_ctx *c, int on);
int shard_get_profile(b200lk_ctx *c, double *partial_ms, int64_t *partial_launches, int reset);
int shard_upload_boot_samples(b200lk_ctx *c, const float *samples, int32_t nsamples);
int shard_rell(b200lk_ctx *c, double *rell);
// api.cu: the blocking halves of the entry points (one synchronising D2H of the result doubles)
int read_results(b200lk_ctx *c, int first, int count);
int finish_lnl(b200lk_ctx *c, double *lnl, bool from_branch);
int finish_derv(b200lk_ctx *c, double *df, double *ddf);
// context.cu: the uploads with an explicit source pitch (a shard takes a column range of the caller's matrix)
int upload_tips_pitched(b200lk_ctx *c, const uint8_t *states, size_t spitch, const uint8_t *unobserved, size_t upitch,
                        const double *ptn_freq);
int upload_boot_samples_pitched(b200lk_ctx *c, const float *samples, size_t pitch_floats, int32_t nsamples);
int launch_tables(b200lk_ctx *c, int nchildren);
int enqueue_update(b200lk_ctx *c, const b200lk_node *nodes, int32_t nnodes, const b200lk_child *children, int32_t nchildren);
int check_ends(b200lk_ctx *c, b200lk_end dad, b200lk_end node);
int launch_partials(b200lk_ctx *c, const std::vector<std::vector<int>> &levels);
int launch_branch_lnl(b200lk_ctx *c, b200lk_end dad, b200lk_end node, double length);
int launch_theta(b200lk_ctx *c, b200lk_end dad, b200lk_end node);
int launch_derv(b200lk_ctx *c, double length);
int launch_from_buffer(b200lk_ctx *c, double length);
int launch_derv_mixlen(b200lk_ctx *c, int K);
int launch_derv_mixlen_class(b200lk_ctx *c, int cls, double length);
int launch_repair(b200lk_ctx *c);
int launch_robust(b200lk_ctx *c);  // robust.cu: replaces result[0] by the robust score of the per-pattern lnL
int launch_newton(b200lk_ctx *c, const b200lk_newton *nr, const b200lk_end *ends);  // ends = {dad, node}: builds theta too
int launch_rell(b200lk_ctx *c);
#define B200_PEER_MAX 64
struct PeerArgs {            // kernel-side view of the exchange; world <= 1: nothing to exchange
    double *const *bufs;     // [world]
    int rank, world;
    unsigned long long seq;  // sequence number of THIS evaluation (>= 1)
};
PeerArgs peer_args_next(b200lk_ctx *c);  // peer.cu: the exchange of the evaluation being enqueued (advances the sequence)
int launch_peer_exchange(b200lk_ctx *c, int count);  // peer.cu: exchange d_result[0..count) in a kernel of its own
bool can_fuse_tail(const b200lk_ctx *c);  // branch.cu: lnL + theta + df/ddf of one evaluation as ONE launch (T32 layout)
int launch_eval_tail(b200lk_ctx *c, b200lk_end dad, b200lk_end node, double length);
int noalloc_setup(b200lk_ctx *c);  // branch.cu: device address of the static scratch of b200lk_lnl_from_buffer_noalloc
int launch_s4_traverse(b200lk_ctx *c, const b200lk_node *nodes, int first, int count, int *prog_used, int *seq_used);
int launch_partial_multi(b200lk_ctx *c, int node_index);
bool s4_detect_sym(const b200lk_ctx *c, double pi[4]);
int launch_dense20(b200lk_ctx *c, int node_base, int nnodes);
int launch_inv_frag(b200lk_ctx *c);
int launch_tt_tables(b200lk_ctx *c, int nnodes);
int s4_realize(b200lk_ctx *c);  // s4.cu: gather every virtual tip-tip vector into its slot (no-op when there is none; either layout)
int dense_realize(b200lk_ctx *c);                               // dense.cu: the pattern-major form of the same
int launch_tt_pairs(b200lk_ctx *c, const S4Virt *v, int n, const int64_t *pair_off);  // dense.cu: pair indices of virtual nodes
int s4_before_update(b200lk_ctx *c, const b200lk_node *nodes, int nnodes, const b200lk_child *children);
static inline int s4_realize_slot(b200lk_ctx *c, int slot) { return (c->n_virtual && c->virt[slot]) ? s4_realize(c) : 0; }
int launch_dense20_tt(b200lk_ctx *c, int node_base, int nnodes);
int launch_denseN(b200lk_ctx *c, int node_base, int nnodes);
int launch_site_partials(b200lk_ctx *c, const std::vector<std::vector<int>> &levels);
// nonrev.cu
int enqueue_update_nonrev(b200lk_ctx *c, const b200lk_node *nodes, int32_t nnodes, const b200lk_child *children, int32_t nchildren);
int nr_check_ends(b200lk_ctx *c, b200lk_end dad, b200lk_end node);
int nr_launch_branch_lnl(b200lk_ctx *c, b200lk_end dad, b200lk_end node, double length);
int nr_record_ends(b200lk_ctx *c, b200lk_end dad, b200lk_end node);
int nr_launch_derv(b200lk_ctx *c, double length);
int nr_launch_from_buffer(b200lk_ctx *c, double length);
