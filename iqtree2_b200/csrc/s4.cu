// s4.cu -- the 4-state (DNA) path: pattern-stationary traversal and the branch-level streams.
//
// Replaces, for nstates == 4 and ncat_mix <= 4 (GTR+G4 and friends):
//   computeTraversalInfo(compute_partial_lh) + computePartialLikelihood*SIMD   tree/phylokernelnew.h:1270-1277, 1291-2039
//   computeLikelihoodBuffer*SIMD (theta)                                      :2050-2231
//   computeLikelihoodBranch*SIMD                                              :2643-3185
//   computeLikelihoodDerv*SIMD                                                :2235-2633
//   computeLikelihoodFromBuffer*SIMD                                          :3196-3392
//
// Design (B200-first, not the reference's loop nest):
//  * One THREAD owns one site pattern and walks the WHOLE post-order node list itself: patterns are
//    independent, so no inter-thread synchronisation exists and every level of the tree is in the same
//    launch.  A child produced earlier in the list comes from registers (the previous node) or from a
//    per-thread shared-memory stack; HBM sees each partial vector once -- the write that keeps it
//    resident for later branch work.  HBM traffic per pattern and II node: 128 B + 1 B instead of the
//    algorithmic 387 B.
//  * HBM layout "T32": a partial vector is tiled by 32 patterns, inside a tile element pair j of
//    pattern l sits at j*64 + l*2 doubles, so a warp's 128-bit access is one contiguous 512 B
//    segment for every load and store (common.cuh: elem_off).
//  * Everything that is the same for all patterns is kept OFF the L1/shared-memory data path, which is
//    what bounds this kernel after HBM (measured: the first version sat at 94 % L1TEX throughput):
//    eigenvectors live in kernel parameters (constant bank, read as instruction operands), and per child
//    only expv[c][i] = exp(lambda_i r_c t) is fetched (warp-uniform loads):
//        sum_i (U[x][i] expv[c][i]) v[i]  ==  sum_i U[x][i] (expv[c][i] v[i])
//    The reference multiplies U*expv first (echild); the association differs by rounding only (parity is
//    judged at 1e-9; a scale count could flip only if max|x| were within ~1e-15 relative of 2^-256).
//  * NORM_LH takes the max over the whole block in-thread; SAFE_LH keeps one counter per category, packed
//    four to a word.
#include "common.cuh"

#define SRC_LEAF 0    // id = row of the staged tip-state chunk
#define SRC_STACK 1   // id = stack entry written earlier by this same thread
#define SRC_GLOBAL 2  // id = slot: resident partial vector
#define SRC_PREV 3    // the node computed immediately before (still in registers)

struct FNode {
    int32_t out_slot;
    int32_t out_stack;   // stack entry to push to, -1 = none
    int32_t lk, lid, rk, rid;
    int32_t ltab, rtab;  // table offset (doubles): leaf table for SRC_LEAF, expv otherwise
    int32_t refill;      // >= 0: before this node, stage chunk `refill` of the tip-state rows
    int32_t pad[3];
};
static_assert(sizeof(FNode) == B200_PROG_BYTES, "B200_PROG_BYTES must equal sizeof(FNode)");

template <int NCM>
struct S4Model {        // per category: eigenvectors of its mixture class, passed by value (constant bank)
    double U[NCM][16];  // evec[x][i]
    double M[NCM][16];  // inv_evec[i][x]
};

struct TravArgs {
    const FNode *prog;
    const int32_t *leaf_seq;
    const double *tables, *ptn_invar;
    const uint8_t *states;
    double *partial;
    uint8_t *scale;
    size_t slot_doubles, slot_scale_bytes;
    int64_t ntot, ntot_pad;
    int nnodes, depth, chunk;
};

#define S4_THREADS 128
#define S4_STATE_BYTES 4096  // staged tip states per CTA: chunk rows x S4_THREADS patterns

// reference reduction trees for N == 4 (tree/phylokernelnew.h:90-93, 235-237, 335-338)
__device__ __forceinline__ double s4_tree(const double *a, const double *b) {
    return __dadd_rn(__fma_rn(a[1], b[1], __dmul_rn(a[0], b[0])), __fma_rn(a[3], b[3], __dmul_rn(a[2], b[2])));
}
__device__ __forceinline__ double s4_plain(const double *a, const double *b) {
    return __dadd_rn(__dadd_rn(__dmul_rn(a[0], b[0]), __dmul_rn(a[1], b[1])),
                     __dadd_rn(__dmul_rn(a[2], b[2]), __dmul_rn(a[3], b[3])));
}
__device__ __forceinline__ unsigned long long abs_bits(double x) {
    return (unsigned long long)__double_as_longlong(x) & 0x7fffffffffffffffull;
}

// One child of one node -> its factor f[c*4+x] for every category.  `sc` accumulates the child's scale
// counter(s): NORM one integer, SAFE four bytes packed (UBYTE arithmetic per category).
template <int NCM, bool SAFE, bool PLAIN>
__device__ __forceinline__ void child_factor(const TravArgs &a, const S4Model<NCM> &mdl, int kind, int id, int tab,
                                             const double *prev, unsigned prev_sc, const double2 *stk,
                                             const unsigned *stk_sc, const uint8_t *st_buf, int tid, size_t poff,
                                             int64_t p, double *f, unsigned &sc) {
    constexpr int BLOCK = NCM * 4;
    if (kind == SRC_LEAF) {
        const int st = st_buf[id * S4_THREADS + tid];
        const double2 *lf = reinterpret_cast<const double2 *>(a.tables + tab + (size_t)st * BLOCK);
#pragma unroll
        for (int j = 0; j < BLOCK / 2; j++) {
            const double2 u = __ldg(lf + j);
            f[2 * j] = u.x;
            f[2 * j + 1] = u.y;
        }
        return;
    }
    double v[BLOCK];
    unsigned s;
    if (kind == SRC_PREV) {
#pragma unroll
        for (int j = 0; j < BLOCK; j++) v[j] = prev[j];
        s = prev_sc;
    } else if (kind == SRC_STACK) {
#pragma unroll
        for (int j = 0; j < BLOCK / 2; j++) {
            const double2 u = stk[(id * (BLOCK / 2) + j) * S4_THREADS + tid];
            v[2 * j] = u.x;
            v[2 * j + 1] = u.y;
        }
        s = stk_sc[id * S4_THREADS + tid];
    } else {
        const double2 *g = reinterpret_cast<const double2 *>(a.partial + (size_t)id * a.slot_doubles + poff);
#pragma unroll
        for (int j = 0; j < BLOCK / 2; j++) {
            const double2 u = __ldcg(g + j * 32);
            v[2 * j] = u.x;
            v[2 * j + 1] = u.y;
        }
        const uint8_t *gs = a.scale + (size_t)id * a.slot_scale_bytes;
        if (SAFE) {
            s = 0;
#pragma unroll
            for (int c = 0; c < NCM; c++) s |= (unsigned)__ldcg(gs + p * NCM + c) << (8 * c);
        } else {
            s = __ldcg(gs + p);
        }
    }
    sc = SAFE ? __vadd4(sc, s) : sc + s;
    const double2 *ev = reinterpret_cast<const double2 *>(a.tables + tab);  // expv[c][i], same address in every lane
#pragma unroll
    for (int c = 0; c < NCM; c++) {
        const double2 e0 = __ldg(ev + 2 * c), e1 = __ldg(ev + 2 * c + 1);
        double w[4];
        w[0] = __dmul_rn(e0.x, v[c * 4]);
        w[1] = __dmul_rn(e0.y, v[c * 4 + 1]);
        w[2] = __dmul_rn(e1.x, v[c * 4 + 2]);
        w[3] = __dmul_rn(e1.y, v[c * 4 + 3]);
#pragma unroll
        for (int x = 0; x < 4; x++)
            f[c * 4 + x] = PLAIN ? s4_plain(&mdl.U[c][4 * x], w) : s4_tree(&mdl.U[c][4 * x], w);
    }
}

template <int NCM, bool SAFE>
__global__ void __launch_bounds__(S4_THREADS, 3) s4_traverse_kernel(const TravArgs a, const S4Model<NCM> mdl) {
    constexpr int BLOCK = NCM * 4, PAIRS = BLOCK / 2;
    // [depth][PAIRS][S4_THREADS] double2 | unsigned [depth][S4_THREADS] scale | uint8 [chunk][S4_THREADS] tip states
    extern __shared__ double2 stk[];
    unsigned *stk_sc = reinterpret_cast<unsigned *>(stk + (size_t)a.depth * PAIRS * S4_THREADS);
    uint8_t *st_buf = reinterpret_cast<uint8_t *>(stk_sc + (size_t)a.depth * S4_THREADS);
    const int tid = threadIdx.x;
    const int64_t p0 = (int64_t)blockIdx.x * S4_THREADS;
    int64_t p = p0 + tid;
    const bool valid = p < a.ntot;
    if (!valid) p = a.ntot - 1;  // the lane stays alive for the barriers; nothing is stored
    const size_t poff = (size_t)(p >> 5) * (32 * BLOCK) + (size_t)(p & 31) * 2;  // T32 layout
    const bool invar0 = a.ptn_invar[p] == 0.0;
    double prev[BLOCK];
#pragma unroll
    for (int j = 0; j < BLOCK; j++) prev[j] = 0.0;
    unsigned prev_sc = 0;

    for (int i = 0; i < a.nnodes; i++) {
        const FNode nd = a.prog[i];  // the same address in every thread
        if (nd.refill >= 0) {
            // stage the next chunk of tip-state rows of this CTA's pattern tile: all loads in flight at once
            __syncthreads();
            constexpr int VPR = S4_THREADS / 16;
            const int nvec = a.chunk * VPR;
            const int32_t *seq = a.leaf_seq + (size_t)nd.refill * a.chunk;
            for (int v = tid; v < nvec; v += S4_THREADS) {
                const int row = v / VPR, q = v - row * VPR;
                const uint4 *src = reinterpret_cast<const uint4 *>(a.states + (size_t)seq[row] * a.ntot_pad + p0) + q;
                reinterpret_cast<uint4 *>(st_buf)[v] = __ldg(src);
            }
            __syncthreads();
        }
        double l[BLOCK], r[BLOCK];
        unsigned sc = 0;
        child_factor<NCM, SAFE, false>(a, mdl, nd.lk, nd.lid, nd.ltab, prev, prev_sc, stk, stk_sc, st_buf, tid, poff,
                                       p, l, sc);
        if (nd.lk == SRC_LEAF)  // tip-internal: dotProductVec for the internal child (:1853)
            child_factor<NCM, SAFE, true>(a, mdl, nd.rk, nd.rid, nd.rtab, prev, prev_sc, stk, stk_sc, st_buf, tid,
                                          poff, p, r, sc);
        else                    // internal-internal: dotProductDualVec (:1975-1983)
            child_factor<NCM, SAFE, false>(a, mdl, nd.rk, nd.rid, nd.rtab, prev, prev_sc, stk, stk_sc, st_buf, tid,
                                           poff, p, r, sc);
        // ---- product, back to eigen-space (productVecMat), rescale --------------------------------
        const bool tiptip = nd.lk == SRC_LEAF && nd.rk == SRC_LEAF;
        unsigned long long mxall = 0ull;
#pragma unroll
        for (int c = 0; c < NCM; c++) {
            double tmp[4];
#pragma unroll
            for (int x = 0; x < 4; x++) tmp[x] = __dmul_rn(l[c * 4 + x], r[c * 4 + x]);
            unsigned long long mx = 0ull;
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const double o = s4_tree(&mdl.M[c][4 * k], tmp);
                prev[c * 4 + k] = o;
                const unsigned long long b = abs_bits(o);  // |x| order == unsigned order of the bit pattern
                mx = b > mx ? b : mx;
            }
            if (SAFE) {
                if (!tiptip && __longlong_as_double((long long)mx) < B200_SCALING_THRESHOLD && invar0) {
#pragma unroll
                    for (int k = 0; k < 4; k++) prev[c * 4 + k] *= B200_TWO_256;
                    sc = __vadd4(sc, 1u << (8 * c));
                }
            } else {
                mxall = mx > mxall ? mx : mxall;
            }
        }
        if (!SAFE) {
            if (!tiptip && __longlong_as_double((long long)mxall) < B200_SCALING_THRESHOLD && invar0) {
#pragma unroll
                for (int j = 0; j < BLOCK; j++) prev[j] *= B200_TWO_256;
                sc += 1;
            }
            sc &= 0xff;  // UBYTE arithmetic of the reference
        }
        prev_sc = sc;
        if (valid) {
            double2 *g = reinterpret_cast<double2 *>(a.partial + (size_t)nd.out_slot * a.slot_doubles + poff);
#pragma unroll
            for (int j = 0; j < PAIRS; j++) g[j * 32] = make_double2(prev[2 * j], prev[2 * j + 1]);
            uint8_t *gs = a.scale + (size_t)nd.out_slot * a.slot_scale_bytes;
            if (SAFE) {
#pragma unroll
                for (int c = 0; c < NCM; c++) gs[p * NCM + c] = (uint8_t)(sc >> (8 * c));
            } else {
                gs[p] = (uint8_t)sc;
            }
        }
        if (nd.out_stack >= 0) {
#pragma unroll
            for (int j = 0; j < PAIRS; j++)
                stk[(nd.out_stack * PAIRS + j) * S4_THREADS + tid] = make_double2(prev[2 * j], prev[2 * j + 1]);
            stk_sc[nd.out_stack * S4_THREADS + tid] = sc;
        }
    }
}

// ---- host side -------------------------------------------------------------------------------------------
bool s4_layout(const b200lk_ctx *c) { return c->tile == 32; }

template <int NCM>
static void fill_model(const b200lk_ctx *c, S4Model<NCM> &m) {
    const int denom = c->d.fused_mix_rate ? 1 : c->d.ncat;
    for (int cc = 0; cc < NCM; cc++) {
        const int mm = cc / denom;
        memcpy(m.U[cc], c->h_evec + (size_t)mm * 16, sizeof(double) * 16);
        memcpy(m.M[cc], c->h_inv_evec + (size_t)mm * 16, sizeof(double) * 16);
    }
}

template <int NCM>
static cudaError_t launch_trav(b200lk_ctx *c, const TravArgs &a, unsigned grid, size_t smem) {
    S4Model<NCM> m;
    fill_model<NCM>(c, m);
    if (!c->trav_attr_set) {
        cudaFuncSetAttribute(s4_traverse_kernel<NCM, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
        cudaFuncSetAttribute(s4_traverse_kernel<NCM, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
        c->trav_attr_set = true;
    }
    if (c->d.safe_numeric) s4_traverse_kernel<NCM, true><<<grid, S4_THREADS, smem, c->stream>>>(a, m);
    else s4_traverse_kernel<NCM, false><<<grid, S4_THREADS, smem, c->stream>>>(a, m);
    return cudaGetLastError();
}

// Build the device program for nodes [first, first+count) of the host's post-order list (all bifurcating;
// children already carry table offsets in c->h_children) and launch.  Stack entries are handed out on the
// host: a value goes to the stack when a later node of this run consumes it and it is not simply the next
// node's operand; when the stack is full the consumer re-reads the resident copy instead.
int launch_s4_traverse(b200lk_ctx *c, const b200lk_node *nodes, int first, int count, int *prog_used, int *seq_used) {
    const int nslots = c->d.nslots;
    const int max_depth = c->max_stack_depth;
    const b200lk_node *nl = nodes + first;
    std::vector<int> producer(nslots, -1);  // slot -> index of the node of this run that last wrote it
    std::vector<int> consumer(count, -1);   // node -> index of the node that consumes its output
    for (int i = 0; i < count; i++) {
        for (int k = 0; k < 2; k++) {
            const DevChild &ch = c->h_children[nl[i].child_begin + k];
            if (ch.kind != B200LK_CHILD_LEAF && producer[ch.id] >= 0 && consumer[producer[ch.id]] < 0)
                consumer[producer[ch.id]] = i;
        }
        producer[nl[i].out_slot] = i;
    }
    std::fill(producer.begin(), producer.end(), -1);
    std::vector<int> entry_of(count, -1);
    std::vector<int> free_list;
    for (int e = max_depth - 1; e >= 0; e--) free_list.push_back(e);
    int depth_used = 0;
    FNode *prog = reinterpret_cast<FNode *>(c->h_prog) + *prog_used;
    int32_t *leaf_seq = reinterpret_cast<int32_t *>((char *)c->h_prog + (size_t)B200_PROG_BYTES * c->nodes_cap) + *seq_used;
    const int chunk = S4_STATE_BYTES / S4_THREADS;
    const int block = c->block, n = 4;
    int nseq = 0;
    for (int i = 0; i < count; i++) {
        FNode f;
        memset(&f, 0, sizeof(f));
        f.out_slot = nl[i].out_slot;
        int kind[2], id[2], tab[2];
        int order[2] = {0, 1};  // a single leaf child goes first (tree/phylokernelnew.h:1395-1405)
        {
            const DevChild &a0 = c->h_children[nl[i].child_begin], &a1 = c->h_children[nl[i].child_begin + 1];
            if (a0.kind != B200LK_CHILD_LEAF && a1.kind == B200LK_CHILD_LEAF) { order[0] = 1; order[1] = 0; }
        }
        int nleaf = 0;
        for (int k = 0; k < 2; k++) nleaf += c->h_children[nl[i].child_begin + k].kind == B200LK_CHILD_LEAF;
        f.refill = -1;
        if (nleaf) {  // tip-state rows are staged chunk by chunk in order of use; a node never straddles
            if (nseq % chunk + nleaf > chunk)
                while (nseq % chunk) leaf_seq[nseq++] = 0;
            if (nseq % chunk == 0) f.refill = nseq / chunk;
        }
        for (int k = 0; k < 2; k++) {
            const DevChild &ch = c->h_children[nl[i].child_begin + order[k]];
            if (ch.kind == B200LK_CHILD_LEAF) {
                kind[k] = SRC_LEAF; id[k] = nseq % chunk; tab[k] = (int)ch.leaf_off;
                leaf_seq[nseq++] = ch.id;
            } else {
                tab[k] = (int)ch.echild_off + block * n;  // expv follows echild
                const int pr = producer[ch.id];
                if (pr >= 0 && consumer[pr] == i && pr == i - 1) {
                    kind[k] = SRC_PREV; id[k] = 0;
                } else if (pr >= 0 && consumer[pr] == i && entry_of[pr] >= 0) {
                    kind[k] = SRC_STACK; id[k] = entry_of[pr];
                    free_list.push_back(entry_of[pr]);
                    entry_of[pr] = -1;
                } else {
                    kind[k] = SRC_GLOBAL; id[k] = ch.id;
                }
            }
        }
        f.lk = kind[0]; f.lid = id[0]; f.ltab = tab[0];
        f.rk = kind[1]; f.rid = id[1]; f.rtab = tab[1];
        f.out_stack = -1;
        if (consumer[i] >= 0 && consumer[i] != i + 1 && !free_list.empty()) {
            f.out_stack = free_list.back();
            free_list.pop_back();
            entry_of[i] = f.out_stack;
            depth_used = std::max(depth_used, f.out_stack + 1);
        }
        producer[nl[i].out_slot] = i;
        prog[i] = f;
    }
    while (nseq % chunk) leaf_seq[nseq++] = 0;  // a refill always reads a whole chunk of rows
    char *d_prog = (char *)c->d_prog + sizeof(FNode) * (size_t)*prog_used;
    char *d_seq = (char *)c->d_prog + (size_t)B200_PROG_BYTES * c->nodes_cap + sizeof(int32_t) * (size_t)*seq_used;
    CUDA_TRY(c, cudaMemcpyAsync(d_prog, prog, sizeof(FNode) * count, cudaMemcpyHostToDevice, c->stream));
    if (nseq)
        CUDA_TRY(c, cudaMemcpyAsync(d_seq, leaf_seq, sizeof(int32_t) * nseq, cudaMemcpyHostToDevice, c->stream));
    *prog_used += count;
    *seq_used += nseq;
    TravArgs a;
    a.prog = reinterpret_cast<const FNode *>(d_prog);
    a.leaf_seq = reinterpret_cast<const int32_t *>(d_seq);
    a.tables = c->tables;
    a.ptn_invar = c->ptn_invar;
    a.states = c->states;
    a.partial = c->partial;
    a.scale = c->scale;
    a.slot_doubles = c->slot_doubles;
    a.slot_scale_bytes = c->slot_scale_bytes;
    a.ntot = c->ntot;
    a.ntot_pad = c->ntot_pad;
    a.nnodes = count;
    a.depth = depth_used;
    a.chunk = chunk;
    const int ncm = c->ncat_mix;
    const size_t smem = (size_t)depth_used * S4_THREADS * (2 * ncm * sizeof(double2) + sizeof(unsigned)) + S4_STATE_BYTES;
    const unsigned grid = (unsigned)((c->ntot + S4_THREADS - 1) / S4_THREADS);
    prof_mark(c);
    cudaError_t e;
    switch (ncm) {
        case 1: e = launch_trav<1>(c, a, grid, smem); break;
        case 2: e = launch_trav<2>(c, a, grid, smem); break;
        case 3: e = launch_trav<3>(c, a, grid, smem); break;
        default: e = launch_trav<4>(c, a, grid, smem); break;
    }
    prof_mark(c);
    CUDA_TRY(c, e);
    c->n_launches++;
    return B200LK_OK;
}
