// traverse.cu -- pattern-stationary traversal for 4-state models: the WHOLE post-order node list in
// ONE launch (replaces the per-packet node loop of computeTraversalInfo + computePartialLikelihood*SIMD,
// tree/phylokernelnew.h:1270-1277, 1291-2039, for nstates == 4).
//
// Site patterns are independent, and for n = 4 one (pattern, category) pair of a node depends only on the
// same pair of its children.  So one thread owns a (pattern, category) pair and walks the node list
// itself; level batching degenerates into "every level in the same launch" and no inter-thread
// synchronisation is needed at all (NORM_LH takes one max over the ncat_mix lanes of a pattern by warp
// shuffle).  Children produced earlier in the same list are taken from registers (the previous node) or a
// per-thread shared-memory stack, so HBM sees each partial vector exactly once -- the write that keeps it
// resident for later branch work -- instead of one write plus one read per child:
//
//   algorithmic bytes / pattern (II node, block = 16):  128 (write) + 256 (child reads) + 3
//   HBM traffic of this kernel:                          128 (write) + 1 (scale)        [+ tip states]
//
// Arithmetic order per element is the reference's (dotProductDualVec / dotProductVec / productVecMat
// reduction trees for N = 4, :90-93, 235-237, 335-338) so that scale counts are bit-exact.
#include "common.cuh"

// child source kinds of the device program
#define SRC_LEAF 0    // id = position of the tip-state row in the staged chunk: leaf lookup table by tip state
#define SRC_STACK 1   // id = stack entry written earlier by this same thread
#define SRC_GLOBAL 2  // id = slot: resident partial vector (computed by an earlier call, or stack overflow)
#define SRC_PREV 3    // the node computed immediately before (still in registers)

struct FNode {
    int32_t out_slot;   // slot to store (always >= 0 here: every partial stays resident)
    int32_t out_stack;  // stack entry to push to, -1 = none
    int32_t lk, lid, rk, rid;
    int32_t ltab, rtab;  // table offset (doubles): leaf table for SRC_LEAF, echild otherwise
    int32_t sync;        // 1: a later node re-reads this output from global memory (stack overflow); in
                         //    NORM mode its scale byte is written by lane c == 0 for the other lanes
    int32_t refill;      // >= 0: before this node, stage chunk `refill` of the tip-state rows
    int32_t pad[2];
};
static_assert(sizeof(FNode) == B200_PROG_BYTES, "B200_PROG_BYTES must equal sizeof(FNode)");

struct TravArgs {
    const FNode *prog;
    const int32_t *leaf_seq;  // taxon of every staged tip-state row, in order of use, chunk-padded
    const double *tables, *inv_evec, *ptn_invar;
    const uint8_t *states;
    double *partial;
    uint8_t *scale;
    size_t slot_doubles, slot_scale_bytes;
    int64_t ntot, ntot_pad;
    int nnodes, ncm, denom, depth, chunk, PT;
};

#define TRAV_THREADS 256
#define TRAV_STATE_BYTES 8192  // staged tip states per CTA: chunk rows x PT patterns

__device__ __forceinline__ double t_dot4_tree(const double2 e0, const double2 e1, const double *b) {
    return __dadd_rn(__fma_rn(e0.y, b[1], __dmul_rn(e0.x, b[0])), __fma_rn(e1.y, b[3], __dmul_rn(e1.x, b[2])));
}
__device__ __forceinline__ double t_dot4_plain(const double2 e0, const double2 e1, const double *b) {
    return __dadd_rn(__dadd_rn(__dmul_rn(e0.x, b[0]), __dmul_rn(e0.y, b[1])),
                     __dadd_rn(__dmul_rn(e1.x, b[2]), __dmul_rn(e1.y, b[3])));
}
__device__ __forceinline__ double t_dot4_tree_r(const double *a, const double *b) {
    return __dadd_rn(__fma_rn(a[1], b[1], __dmul_rn(a[0], b[0])), __fma_rn(a[3], b[3], __dmul_rn(a[2], b[2])));
}

template <bool SAFE>
__global__ void __launch_bounds__(TRAV_THREADS, 3) traverse4_kernel(TravArgs a) {
    // [depth][2][TRAV_THREADS] double2 values | int [depth][TRAV_THREADS] scale | uint8 [chunk][PT] tip states
    extern __shared__ double2 stk[];
    int *stk_sc = reinterpret_cast<int *>(stk + (size_t)a.depth * 2 * TRAV_THREADS);
    uint8_t *st_buf = reinterpret_cast<uint8_t *>(stk_sc + (size_t)a.depth * TRAV_THREADS);
    const int tid = threadIdx.x;
    const int ncm = a.ncm, block = ncm * 4, PT = a.PT;
    const int pl = tid / ncm, c = tid - pl * ncm;
    const int64_t p0 = (int64_t)blockIdx.x * PT;
    int64_t p = p0 + pl;
    const bool valid = p < a.ntot;
    if (!valid) p = a.ntot - 1;  // lanes stay alive for shuffles and barriers; nothing is stored
    double M[16];
    {
        const double *m = a.inv_evec + (size_t)(c / a.denom) * 16;
#pragma unroll
        for (int i = 0; i < 16; i++) M[i] = m[i];
    }
    const bool invar0 = a.ptn_invar[p] == 0.0;
    const size_t off = (size_t)p * block + (size_t)c * 4;
    const size_t sc_off = SAFE ? (size_t)p * ncm + c : (size_t)p;
    double prev[4] = {0.0, 0.0, 0.0, 0.0};
    int prev_sc = 0;

    FNode nd = a.prog[0];  // same address in every thread: broadcast loads
    for (int i = 0; i < a.nnodes; i++) {
        FNode nxt = nd;
        if (i + 1 < a.nnodes) nxt = a.prog[i + 1];  // one node ahead of use
        if (nd.refill >= 0) {
            // stage the next `chunk` tip-state rows of this CTA's pattern tile: all loads in flight at once
            __syncthreads();
            const int vec_per_row = PT / 16, nvec = a.chunk * vec_per_row;
            const int32_t *seq = a.leaf_seq + (size_t)nd.refill * a.chunk;
            for (int v = tid; v < nvec; v += TRAV_THREADS) {
                const int row = v / vec_per_row, q = v - row * vec_per_row;
                const uint4 *src = reinterpret_cast<const uint4 *>(a.states + (size_t)seq[row] * a.ntot_pad + p0) + q;
                reinterpret_cast<uint4 *>(st_buf)[v] = __ldg(src);
            }
            __syncthreads();
        }
        double l[4], r[4];
        int sc = 0;
        // ---- left factor --------------------------------------------------------------------------
        if (nd.lk == SRC_LEAF) {
            const int st = st_buf[nd.lid * PT + pl];
            const double2 *lf = reinterpret_cast<const double2 *>(a.tables + nd.ltab + (size_t)st * block + c * 4);
            const double2 u = __ldg(lf), w = __ldg(lf + 1);
            l[0] = u.x; l[1] = u.y; l[2] = w.x; l[3] = w.y;
        } else {
            double v[4];
            if (nd.lk == SRC_PREV) {
                v[0] = prev[0]; v[1] = prev[1]; v[2] = prev[2]; v[3] = prev[3];
                sc += prev_sc;
            } else if (nd.lk == SRC_STACK) {
                const double2 u = stk[(nd.lid * 2) * TRAV_THREADS + tid], w = stk[(nd.lid * 2 + 1) * TRAV_THREADS + tid];
                v[0] = u.x; v[1] = u.y; v[2] = w.x; v[3] = w.y;
                sc += stk_sc[nd.lid * TRAV_THREADS + tid];
            } else {
                const double2 *g = reinterpret_cast<const double2 *>(a.partial + (size_t)nd.lid * a.slot_doubles + off);
                const double2 u = __ldcg(g), w = __ldcg(g + 1);
                v[0] = u.x; v[1] = u.y; v[2] = w.x; v[3] = w.y;
                sc += __ldcg(a.scale + (size_t)nd.lid * a.slot_scale_bytes + sc_off);
            }
            const double2 *e = reinterpret_cast<const double2 *>(a.tables + nd.ltab + c * 16);
#pragma unroll
            for (int x = 0; x < 4; x++) l[x] = t_dot4_tree(__ldg(e + 2 * x), __ldg(e + 2 * x + 1), v);
        }
        // ---- right factor -------------------------------------------------------------------------
        if (nd.rk == SRC_LEAF) {
            const int st = st_buf[nd.rid * PT + pl];
            const double2 *lf = reinterpret_cast<const double2 *>(a.tables + nd.rtab + (size_t)st * block + c * 4);
            const double2 u = __ldg(lf), w = __ldg(lf + 1);
            r[0] = u.x; r[1] = u.y; r[2] = w.x; r[3] = w.y;
        } else {
            double v[4];
            if (nd.rk == SRC_PREV) {
                v[0] = prev[0]; v[1] = prev[1]; v[2] = prev[2]; v[3] = prev[3];
                sc += prev_sc;
            } else if (nd.rk == SRC_STACK) {
                const double2 u = stk[(nd.rid * 2) * TRAV_THREADS + tid], w = stk[(nd.rid * 2 + 1) * TRAV_THREADS + tid];
                v[0] = u.x; v[1] = u.y; v[2] = w.x; v[3] = w.y;
                sc += stk_sc[nd.rid * TRAV_THREADS + tid];
            } else {
                const double2 *g = reinterpret_cast<const double2 *>(a.partial + (size_t)nd.rid * a.slot_doubles + off);
                const double2 u = __ldcg(g), w = __ldcg(g + 1);
                v[0] = u.x; v[1] = u.y; v[2] = w.x; v[3] = w.y;
                sc += __ldcg(a.scale + (size_t)nd.rid * a.slot_scale_bytes + sc_off);
            }
            const double2 *e = reinterpret_cast<const double2 *>(a.tables + nd.rtab + c * 16);
            if (nd.lk == SRC_LEAF) {  // tip-internal: dotProductVec (:1853)
#pragma unroll
                for (int x = 0; x < 4; x++) r[x] = t_dot4_plain(__ldg(e + 2 * x), __ldg(e + 2 * x + 1), v);
            } else {                  // internal-internal: dotProductDualVec (:1975-1983)
#pragma unroll
                for (int x = 0; x < 4; x++) r[x] = t_dot4_tree(__ldg(e + 2 * x), __ldg(e + 2 * x + 1), v);
            }
        }
        // ---- product, back to eigen-space, rescale --------------------------------------------------
        double tmp[4], o[4];
#pragma unroll
        for (int x = 0; x < 4; x++) tmp[x] = __dmul_rn(l[x], r[x]);
        unsigned long long mxb = 0ull;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            o[k] = t_dot4_tree_r(M + 4 * k, tmp);
            const unsigned long long b = (unsigned long long)__double_as_longlong(o[k]) & 0x7fffffffffffffffull;
            mxb = b > mxb ? b : mxb;  // |x| ordering == unsigned ordering of the bit pattern (NaN sorts above)
        }
        if (!SAFE) {
            for (int o2 = 1; o2 < ncm; o2 <<= 1) {
                const unsigned long long other = __shfl_xor_sync(0xffffffffu, mxb, o2);
                mxb = other > mxb ? other : mxb;
            }
        }
        const bool tiptip = nd.lk == SRC_LEAF && nd.rk == SRC_LEAF;
        const bool under = !tiptip && __longlong_as_double((long long)mxb) < B200_SCALING_THRESHOLD && invar0;
        if (under) {
#pragma unroll
            for (int k = 0; k < 4; k++) o[k] *= B200_TWO_256;
            sc += 1;
        }
        sc &= 0xff;  // UBYTE arithmetic of the reference
        if (valid) {
            double2 *g = reinterpret_cast<double2 *>(a.partial + (size_t)nd.out_slot * a.slot_doubles + off);
            g[0] = make_double2(o[0], o[1]);
            g[1] = make_double2(o[2], o[3]);
            if (SAFE || c == 0) a.scale[(size_t)nd.out_slot * a.slot_scale_bytes + sc_off] = (uint8_t)sc;
        }
        if (nd.out_stack >= 0) {
            stk[(nd.out_stack * 2) * TRAV_THREADS + tid] = make_double2(o[0], o[1]);
            stk[(nd.out_stack * 2 + 1) * TRAV_THREADS + tid] = make_double2(o[2], o[3]);
            stk_sc[nd.out_stack * TRAV_THREADS + tid] = sc;
        }
        if (nd.sync) __syncwarp();  // orders the scale store of lane c == 0 before the other lanes' re-read
        prev[0] = o[0]; prev[1] = o[1]; prev[2] = o[2]; prev[3] = o[3];
        prev_sc = sc;
        nd = nxt;
    }
}

// Can this list run as one pattern-stationary launch?  n == 4, bifurcating nodes only, and (NORM) the
// ncat_mix lanes of a pattern inside one warp.
bool traverse4_supported(const b200lk_ctx *c, const b200lk_node *nodes, int nnodes) {
    if (c->d.nstates != 4 || c->force_level_batched) return false;
    const int ncm = c->ncat_mix;
    if (ncm > 16 || (ncm & (ncm - 1)) != 0) return false;  // the lanes of a pattern share a warp; PT % 16 == 0
    for (int i = 0; i < nnodes; i++)
        if (nodes[i].nchild != 2) return false;
    return true;
}

// Build the device program from the host's post-order list (children already carry table offsets in
// c->h_children) and launch.  Stack entries are handed out on the host: a value goes to the stack when a
// later node of this same list consumes it and it is not simply the next node's operand.
int launch_traverse4(b200lk_ctx *c, const b200lk_node *nodes, int nnodes) {
    const int nslots = c->d.nslots;
    const int max_depth = c->max_stack_depth;
    std::vector<int> producer(nslots, -1);      // slot -> index of the node of this list that last wrote it
    std::vector<int> consumer(nnodes, -1);      // node -> index of the node that consumes its output
    for (int i = 0; i < nnodes; i++) {
        for (int k = 0; k < 2; k++) {
            const DevChild &ch = c->h_children[nodes[i].child_begin + k];
            if (ch.kind != B200LK_CHILD_LEAF && producer[ch.id] >= 0 && consumer[producer[ch.id]] < 0)
                consumer[producer[ch.id]] = i;
        }
        producer[nodes[i].out_slot] = i;
    }
    std::fill(producer.begin(), producer.end(), -1);
    std::vector<int> entry_of(nnodes, -1);  // node -> stack entry holding its output
    std::vector<int> free_list;
    for (int e = max_depth - 1; e >= 0; e--) free_list.push_back(e);
    int depth_used = 0;
    FNode *prog = reinterpret_cast<FNode *>(c->h_prog);
    int32_t *leaf_seq = reinterpret_cast<int32_t *>((char *)c->h_prog + (size_t)B200_PROG_BYTES * c->nodes_cap);
    const int PT = TRAV_THREADS / c->ncat_mix, chunk = TRAV_STATE_BYTES / PT;
    int nseq = 0;  // tip-state rows handed out so far (chunk-padded)
    for (int i = 0; i < nnodes; i++) {
        FNode f;
        f.out_slot = nodes[i].out_slot;
        int kind[2], id[2], tab[2];
        // a single leaf child goes first (tree/phylokernelnew.h:1395-1405)
        int order[2] = {0, 1};
        {
            const DevChild &a0 = c->h_children[nodes[i].child_begin], &a1 = c->h_children[nodes[i].child_begin + 1];
            if (a0.kind != B200LK_CHILD_LEAF && a1.kind == B200LK_CHILD_LEAF) { order[0] = 1; order[1] = 0; }
        }
        // tip-state rows are staged chunk by chunk in order of use; a node never straddles two chunks
        int nleaf = 0;
        for (int k = 0; k < 2; k++) nleaf += c->h_children[nodes[i].child_begin + k].kind == B200LK_CHILD_LEAF;
        f.refill = -1;
        if (nleaf) {
            if (nseq % chunk + nleaf > chunk)
                while (nseq % chunk) leaf_seq[nseq++] = 0;
            if (nseq % chunk == 0) f.refill = nseq / chunk;
        }
        for (int k = 0; k < 2; k++) {
            const DevChild &ch = c->h_children[nodes[i].child_begin + order[k]];
            if (ch.kind == B200LK_CHILD_LEAF) {
                kind[k] = SRC_LEAF; id[k] = nseq % chunk; tab[k] = (int)ch.leaf_off;
                leaf_seq[nseq++] = ch.id;
            } else {
                tab[k] = (int)ch.echild_off;
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
        f.sync = 0;
        f.pad[0] = f.pad[1] = 0;
        if (consumer[i] >= 0 && consumer[i] != i + 1) {
            if (!free_list.empty()) {
                f.out_stack = free_list.back();
                free_list.pop_back();
                entry_of[i] = f.out_stack;
                depth_used = std::max(depth_used, f.out_stack + 1);
            } else {
                f.sync = 1;
            }
        }
        producer[nodes[i].out_slot] = i;
        prog[i] = f;
    }
    while (nseq % chunk) leaf_seq[nseq++] = 0;  // a refill always reads a whole chunk of rows
    char *d_seq = (char *)c->d_prog + (size_t)B200_PROG_BYTES * c->nodes_cap;
    CUDA_TRY(c, cudaMemcpyAsync(c->d_prog, prog, sizeof(FNode) * nnodes, cudaMemcpyHostToDevice, c->stream));
    if (nseq)
        CUDA_TRY(c, cudaMemcpyAsync(d_seq, leaf_seq, sizeof(int32_t) * nseq, cudaMemcpyHostToDevice, c->stream));
    TravArgs a;
    a.prog = reinterpret_cast<const FNode *>(c->d_prog);
    a.tables = c->tables;
    a.inv_evec = c->m.inv_evec;
    a.ptn_invar = c->ptn_invar;
    a.states = c->states;
    a.partial = c->partial;
    a.scale = c->scale;
    a.slot_doubles = c->slot_doubles;
    a.slot_scale_bytes = c->slot_scale_bytes;
    a.ntot = c->ntot;
    a.ntot_pad = c->ntot_pad;
    a.nnodes = nnodes;
    a.ncm = c->ncat_mix;
    a.denom = c->d.fused_mix_rate ? 1 : c->d.ncat;
    a.depth = depth_used;
    a.leaf_seq = reinterpret_cast<const int32_t *>(d_seq);
    a.chunk = chunk;
    a.PT = PT;
    const size_t smem = (size_t)depth_used * TRAV_THREADS * (2 * sizeof(double2) + sizeof(int)) + TRAV_STATE_BYTES;
    const unsigned grid = (unsigned)((c->ntot + PT - 1) / PT);
    if (!c->trav_attr_set) {
        CUDA_TRY(c, cudaFuncSetAttribute(traverse4_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
        CUDA_TRY(c, cudaFuncSetAttribute(traverse4_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
        c->trav_attr_set = true;
    }
    prof_mark(c);
    if (c->d.safe_numeric) traverse4_kernel<true><<<grid, TRAV_THREADS, smem, c->stream>>>(a);
    else traverse4_kernel<false><<<grid, TRAV_THREADS, smem, c->stream>>>(a);
    prof_mark(c);
    CUDA_TRY(c, cudaGetLastError());
    c->n_launches++;
    return B200LK_OK;
}
