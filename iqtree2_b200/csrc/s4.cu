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
//    what bounds this kernel after HBM (measured: the first version sat at 94 % L1TEX throughput, the
//    second waited on dependent global loads): the node program, the eigenvectors and, per child, only
//    expv[c][i] = exp(lambda_i r_c t) are KERNEL PARAMETERS (constant bank -> uniform registers -> operands):
//        sum_i (U[x][i] expv[c][i]) v[i]  ==  sum_i U[x][i] (expv[c][i] v[i])
//    The reference multiplies U*expv first (echild); the association differs by rounding only (parity is
//    judged at 1e-9; a scale count could flip only if max|x| were within ~1e-15 relative of 2^-256).
//  * NORM_LH takes the max over the whole block in-thread; SAFE_LH keeps one counter per category, packed
//    four to a word.
#include <omp.h>

#include "common.cuh"

#define SRC_LEAF 0    // id = row of the staged tip-state chunk
#define SRC_STACK 1   // id = stack entry written earlier by this same thread
#define SRC_GLOBAL 2  // id = slot: resident partial vector
#define SRC_PREV 3    // the node computed immediately before (still in registers)

// One node of the device program: 32 bytes.  The program, the per-child factors expv[c][i] = exp(lambda_i r_c t) and
// the eigenvectors travel as KERNEL PARAMETERS: they sit in the constant bank, are read through the uniform datapath
// (LDCU -> uniform registers -> operands of DFMA/DMUL) and never touch L1, shared memory or a vector register.
// The factors are a separate packed array, one block of NCM * 4 doubles per INTERNAL child (a tip child needs none), so a
// launch carries as many nodes as 32 * nodes + 8 * NCM * 4 * internal children + the model fit the 32 KB parameter window:
// all 198 nodes of cfg2 (31.8 KB), where a fixed 288-byte node took two launches of <= 107.
struct PNode {
    int32_t out_slot;
    int32_t info;        // lk | rk << 2 | (out_stack + 1) << 4 | tiptip_table << 12 | virtual << 13 | (refill + 1) << 14
    int32_t lid, rid;    // SRC_LEAF: row in the staged chunk; SRC_STACK: entry; SRC_GLOBAL: slot
    int32_t ltab, rtab;  // SRC_LEAF: offset (doubles) of the child's leaf lookup table in `tables`
    int32_t lexp, rexp;  // internal child: index of its factor block in the launch's factor array; batched launches: offset
                         // (doubles) of the same factors in the context's tables
};

struct TravArgs {
    const int32_t *leaf_seq;  // taxon of every staged tip-state row, in order of use, chunk-padded
    const double *tables, *ptn_invar;
    const double *tt;  // state-pair tables of the tip-tip nodes (s4_tt_tables_kernel): [node][sa * codes + sb][block]
    const uint8_t *states;
    double *partial;
    uint8_t *scale;
    size_t slot_doubles, slot_scale_bytes;
    int64_t ntot, ntot_pad;
    int nnodes, depth, chunk, denom, codes;
    int nchunks;   // staged tip-state chunks of this launch (s4_traverse_kernel fetches chunk k + 1 while chunk k is in use)
    int prefetch;  // 1: the table rows of the NEXT node's tip children are pulled into L1 while this node computes
};

#define S4_PARAM_BYTES 32640  // of the 32764 a kernel may take
// blob: [ model | PNode nodes[NMAX] | double factors[FMAX][NCM * 4] ]
//   model, SYM kernels:   U[16] (evec[x][i]), pi[4]                               160 bytes
//   model, general:       U[NCM][16], M[NCM][16] (inv_evec[i][x]) per category    256 * NCM bytes
// Both arrays start at COMPILE-TIME offsets and grow upwards (element address = constant + counter * stride: ptxas
// keeps such loads on the uniform datapath); the window is split so that as many factor blocks fit as nodes.
template <int NCM>
struct S4Launch {
    static constexpr int BLOB = (S4_PARAM_BYTES - (int)sizeof(TravArgs)) / 32 * 32;
    static constexpr int FAC = NCM * 4 * (int)sizeof(double);  // bytes of one factor block
    __host__ __device__ static constexpr int nmax(int model_bytes) { return (BLOB - model_bytes) / (32 + FAC); }  // nodes AND factor blocks
    __host__ __device__ static constexpr int fac_off(int model_bytes) { return model_bytes + 32 * nmax(model_bytes); }
    TravArgs a;
    alignas(16) unsigned char blob[BLOB];
};
template <int NCM, int MODEL_BYTES>
__device__ __forceinline__ const PNode &s4_node(const S4Launch<NCM> &P, int i) {
    return reinterpret_cast<const PNode *>(P.blob + MODEL_BYTES)[i];
}
template <int NCM, int MODEL_BYTES>
__device__ __forceinline__ const double *s4_factors(const S4Launch<NCM> &P, int idx) {
    constexpr int FAC_OFF = S4Launch<NCM>::fac_off(MODEL_BYTES);
    return reinterpret_cast<const double *>(P.blob + FAC_OFF) + (size_t)idx * (NCM * 4);
}

#ifndef S4_THREADS
#define S4_THREADS 64  // two warps per CTA: 4.71 ms per cfg2 traversal against 5.03 ms with four (profiles/r1t_s4_cta64_ab.log)
#endif
#define S4_THRESHOLD_HI 0x2ff00000  // high word of 2^-256 = B200_SCALING_THRESHOLD (exponent 1023 - 256, mantissa 0)
#ifndef S4_MINB
#define S4_MINB 10  // resident CTAs per SM the register allocation is sized for (SYM kernels: 96 registers; the general ones use 8)
#endif
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

// Shared-memory slot of a node vector.  Classic layout: [slot][pair][thread].  Warp-major layout (WL, the bulk-store
// kernel): [slot][warp][pair][lane] -- a warp's part of a slot is then the byte image of its 32-pattern tile in HBM
// (T32: pair j of lane l at j*32 + l double2), which is what one cp.async.bulk copies out.
template <bool WL, int PAIRS>
__device__ __forceinline__ int stk_at(int slot, int j, int tid) {
    return WL ? ((slot * (S4_THREADS / 32) + (tid >> 5)) * PAIRS + j) * 32 + (tid & 31) : (slot * PAIRS + j) * S4_THREADS + tid;
}

// raw operand of one child: the leaf-table row (SRC_LEAF, already the factor) or the child's partial vector
template <int NCM, bool SAFE, bool WL = false>
__device__ __forceinline__ void child_load(const TravArgs &a, int kind, int id, int tab, const double *prev,
                                           unsigned prev_sc, const double2 *stk, const unsigned *stk_sc,
                                           const uint8_t *st_buf, int tid, size_t poff, int64_t p, double *v,
                                           unsigned &sc) {
    constexpr int BLOCK = NCM * 4;
    if (kind == SRC_LEAF) {
        const int st = st_buf[id * S4_THREADS + tid];
        const double2 *lf = reinterpret_cast<const double2 *>(a.tables + tab + (size_t)st * BLOCK);
#pragma unroll
        for (int j = 0; j < BLOCK / 2; j++) {
            const double2 u = __ldg(lf + j);
            v[2 * j] = u.x;
            v[2 * j + 1] = u.y;
        }
        return;
    }
    unsigned s;
    if (kind == SRC_PREV) {
#pragma unroll
        for (int j = 0; j < BLOCK; j++) v[j] = prev[j];
        s = prev_sc;
    } else if (kind == SRC_STACK) {
#pragma unroll
        for (int j = 0; j < BLOCK / 2; j++) {
            const double2 u = stk[stk_at<WL, BLOCK / 2>(id, j, tid)];
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
}

// internal child: v <- U (expv (.) v) per category, in place
template <int NCM, bool PLAIN, bool SYM = false>
__device__ __forceinline__ void child_apply(const double *U, const double *ev, double *v, const double *Ur = nullptr) {
#pragma unroll
    for (int c = 0; c < NCM; c++) {
        double w[4], f[4];
#pragma unroll
        for (int i = 0; i < 4; i++) w[i] = __dmul_rn(ev[c * 4 + i], v[c * 4 + i]);
#pragma unroll
        for (int x = 0; x < 4; x++) {
            const double *u = SYM ? Ur + 4 * x : U + c * 16 + 4 * x;  // SYM: register copy of the shared eigenvectors
            f[x] = PLAIN ? s4_plain(u, w) : s4_tree(u, w);
        }
#pragma unroll
        for (int x = 0; x < 4; x++) v[c * 4 + x] = f[x];
    }
}

// SYM: every category shares the eigenvectors and the inverse is M[i][x] = U[x][i] * pi[x] (what a reversible model's
// eigen-solver produces: evec = Pi^-1/2 V, inv_evec = V^T Pi^1/2, utils/eigendecomposition.cpp:165-317), checked on the
// host at set_model.  The back-transformation then reads U again -- out[i] = sum_x U[x][i] (pi[x] tmp[x]), same
// reduction tree over x -- so only U and pi (40 uniform registers) are live across the node loop instead of U and M
// (64: more than the uniform register file holds, which made the kernel re-load both for every category of every
// node; that constant traffic, not HBM, was what bounded it -- profiles/r1s_s4_experiments.log).  The SYM model block is U[16], pi[4].
// The 20 values are copied into vector registers once per thread (through shared memory, so that the compiler cannot
// turn them back into constant-bank reads); the kernel then fits 96 registers: five resident CTAs per SM with two
// stack entries (profiles/r1t_s4_ureg_ab.log), later as ten CTAs of 64 threads.
//
// BULK: the node vector leaves through the TMA engine instead of the LSU.  Measured (profiles/r2h_s4_traverse_cfg2.md,
// source page): the traversal is a pure 12.7 GB write stream per launch, and a warp's 512-byte STG.E.128 costs the L1
// data pipe one wavefront per 32-byte sector (16 per instruction) -- 87 % of that pipe, the limiter ahead of HBM.  An
// STS.128 costs 4.  So every node writes its vector to a shared-memory slot laid out as the byte image of the warp's
// T32 tile (stk_at<WL>), and one elected lane hands the 32*BLOCK*8 bytes to cp.async.bulk.global.shared::cta.  The
// slot IS the stack entry when a later node wants the vector (the host keeps it held), else it is free again at the
// next node: the only wait is cp.async.bulk.wait_group.read at the top of the next store, a whole node later.
// A consumer that re-reads a vector of this launch from HBM (stack full) first waits for the warp's copies to complete.
template <int NCM, bool SAFE, bool SYM, int MINB = S4_MINB, bool BULK = false>
__global__ void __launch_bounds__(S4_THREADS, MINB) s4_traverse_kernel(const __grid_constant__ S4Launch<NCM> P) {
    constexpr int BLOCK = NCM * 4, PAIRS = BLOCK / 2;
    constexpr int MODEL_BYTES = SYM ? 160 : 2 * NCM * 16 * (int)sizeof(double);  // == model_bytes() on the host
    const TravArgs &a = P.a;
    // [depth][PAIRS][S4_THREADS] double2 | unsigned [depth][S4_THREADS] scale | uint8 [chunk][S4_THREADS] tip states
    extern __shared__ double2 stk[];
    unsigned *stk_sc = reinterpret_cast<unsigned *>(stk + (size_t)a.depth * PAIRS * S4_THREADS);
    uint8_t *st_base = reinterpret_cast<uint8_t *>(stk_sc + (size_t)a.depth * S4_THREADS);
    const uint8_t *st_buf = st_base;  // the half of the staging buffer the current chunk lives in
    const int tid = threadIdx.x;
    const int64_t p0 = (int64_t)blockIdx.x * S4_THREADS;
    // Tip states are staged chunk by chunk (a.chunk rows x S4_THREADS patterns) into the two halves of the buffer: chunk
    // k + 1 is fetched with cp.async while the nodes of chunk k run, so a refill is one barrier with nothing to wait for.
    // (Fetched at the refill itself it was two dependent global round trips -- taxon id, then its states -- taken by the
    // whole CTA between two barriers: 12 % of the stall samples, profiles/r2n_s4_traverse_cfg2.md.)
    auto fetch_chunk = [&](int k) {
        constexpr int VPR = S4_THREADS / 16;
        const int nvec = a.chunk * VPR;
        const int32_t *seq = a.leaf_seq + (size_t)k * a.chunk;
        uint8_t *dst = st_base + (size_t)(k & 1) * a.chunk * S4_THREADS;
        for (int v = tid; v < nvec; v += S4_THREADS) {
            const int row = v / VPR, q = v - row * VPR;
            const uint8_t *src = a.states + (size_t)__ldg(seq + row) * a.ntot_pad + p0 + 16 * q;
            const uint32_t d = (uint32_t)__cvta_generic_to_shared(dst + 16 * v);
            asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(d), "l"(src) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    if (a.nchunks > 0) fetch_chunk(0);
    int64_t p = p0 + tid;
    const bool valid = p < a.ntot;
    if (!valid) p = a.ntot - 1;  // the lane stays alive for the barriers; nothing is stored
    const size_t poff = (size_t)(p >> 5) * (32 * BLOCK) + (size_t)(p & 31) * 2;  // T32 layout
    // BULK: does this warp own a whole 32-pattern tile?  (p0 is a multiple of 32; the one ragged warp of the grid stores
    // lane by lane, a warp past the end stores nothing)
    const bool warp_full = p0 + (tid | 31) < a.ntot;
    const bool invar0 = a.ptn_invar[p] == 0.0;
    double prev[BLOCK];
#pragma unroll
    for (int j = 0; j < BLOCK; j++) prev[j] = 0.0;
    unsigned prev_sc = 0;
    double Ur[16], pir[4];
    if (SYM) {
        // through shared memory so that the values live in vector registers for the whole loop instead of being
        // re-read from the constant bank into uniform registers by every code path of every node
        __shared__ double u_sh[20];
        if (tid < 20) u_sh[tid] = reinterpret_cast<const double *>(P.blob)[tid];  // U[16], pi[4]
        __syncthreads();
#pragma unroll
        for (int j = 0; j < 16; j++) Ur[j] = reinterpret_cast<volatile double *>(u_sh)[j];
#pragma unroll
        for (int j = 0; j < 4; j++) pir[j] = reinterpret_cast<volatile double *>(u_sh)[16 + j];
    }

    int nfac = 0;  // internal children met so far = index of the next factor block
    for (int i = 0; i < a.nnodes; i++) {
        const PNode &nd = s4_node<NCM, MODEL_BYTES>(P, i);  // constant bank, uniform index
        const int info = nd.info;
        const int lk = info & 3, rk = (info >> 2) & 3, out_stack = ((info >> 4) & 0xff) - 1, refill = (info >> 14) - 1;
        const bool tt = (info >> 12) & 1;
        // a tip-tip vector that only this launch's next consumer needs (from registers or the stack) is not written at
        // all: it stays "virtual" -- (table, two taxa) on the host -- until something else asks for the slot (s4_realize)
        const bool store = valid && !((info >> 13) & 1);
        if (refill >= 0) {
            // chunk `refill` was requested a chunk ago: complete this thread's copies, then the barrier makes everybody's
            // visible and says that nobody reads the other half any more -- which chunk refill + 1 may now overwrite
            asm volatile("cp.async.wait_all;" ::: "memory");
            __syncthreads();
            st_buf = st_base + (size_t)(refill & 1) * a.chunk * S4_THREADS;
            if (refill + 1 < a.nchunks) fetch_chunk(refill + 1);
        }
        unsigned sc = 0;
        if (a.prefetch && i + 1 < a.nnodes) {
            // The tip-child rows (one 128-byte line each, leaf table or state-pair table) miss L1 about every other time --
            // shared memory leaves it ~40 KB and the CTAs of an SM are spread over the whole node list -- and the wait for
            // them is the largest stall of the kernel (long scoreboard at the first use).  The next node's rows depend only
            // on the staged tip states, so they are requested now and arrive while this node computes.  No register is held.
            const PNode &nx = s4_node<NCM, MODEL_BYTES>(P, i + 1);
            const int ni = nx.info;
            if ((ni >> 14) == 0) {  // same staged chunk
                const double *row = nullptr, *row2 = nullptr;
                if ((ni >> 12) & 1) {
                    const int sa = st_buf[nx.lid * S4_THREADS + tid], sb = st_buf[nx.rid * S4_THREADS + tid];
                    row = a.tt + (size_t)nx.ltab + (size_t)(sa * a.codes + sb) * BLOCK;
                } else {
                    if ((ni & 3) == SRC_LEAF) row = a.tables + nx.ltab + (size_t)st_buf[nx.lid * S4_THREADS + tid] * BLOCK;
                    if (((ni >> 2) & 3) == SRC_LEAF) row2 = a.tables + nx.rtab + (size_t)st_buf[nx.rid * S4_THREADS + tid] * BLOCK;
                }
                if (row) asm volatile("prefetch.global.L1 [%0];" ::"l"(row));
                if (row2) asm volatile("prefetch.global.L1 [%0];" ::"l"(row2));
            }
        }
        if (tt) {
            // tip-tip node: its vector depends on the two tip states only, and was computed for every state pair by
            // s4_tt_tables_kernel with this kernel's own arithmetic.  One row load instead of two, no FP64 work, never
            // rescaled (tree/phylokernelnew.h:1632-1746)
            const int sa = st_buf[nd.lid * S4_THREADS + tid], sb = st_buf[nd.rid * S4_THREADS + tid];
            const double2 *row = reinterpret_cast<const double2 *>(a.tt + (size_t)nd.ltab + (size_t)(sa * a.codes + sb) * BLOCK);
#pragma unroll
            for (int j = 0; j < PAIRS; j++) {
                const double2 u = __ldg(row + j);
                prev[2 * j] = u.x;
                prev[2 * j + 1] = u.y;
            }
        } else {
        // both children's operands are requested before either is used: one memory latency per node
        double l[BLOCK], r[BLOCK];
        if (BULK && (lk == SRC_GLOBAL || rk == SRC_GLOBAL)) {
            // the vector may be one this warp handed to the copy engine earlier in this launch: complete, then read
            if ((tid & 31) == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
            __syncwarp();
        }
        child_load<NCM, SAFE, BULK>(a, lk, nd.lid, nd.ltab, prev, prev_sc, stk, stk_sc, st_buf, tid, poff, p, l, sc);
        child_load<NCM, SAFE, BULK>(a, rk, nd.rid, nd.rtab, prev, prev_sc, stk, stk_sc, st_buf, tid, poff, p, r, sc);
        const double *mU = reinterpret_cast<const double *>(P.blob);  // general kernel: U[NCM][16], then M[NCM][16]
        // The factor blocks lie in the order the node loop meets the internal children, so their position is a running
        // count (nd.lexp / nd.rexp hold the same numbers).  An address formed from a LOADED index makes ptxas fetch the
        // factors with register-indexed LDC instead of uniform LDCU: 4.09 against 3.78 ms (profiles/r2r_s4_compact_ab.log).
        if (lk != SRC_LEAF) child_apply<NCM, false, SYM>(mU, s4_factors<NCM, MODEL_BYTES>(P, nfac++), l, Ur);
        if (rk != SRC_LEAF) {
            if (lk == SRC_LEAF) child_apply<NCM, true, SYM>(mU, s4_factors<NCM, MODEL_BYTES>(P, nfac++), r, Ur);   // tip-internal: dotProductVec (:1853)
            else child_apply<NCM, false, SYM>(mU, s4_factors<NCM, MODEL_BYTES>(P, nfac++), r, Ur);                 // internal-internal (:1975-1983)
        }
        // ---- product, back to eigen-space (productVecMat), rescale --------------------------------
        const bool tiptip = lk == SRC_LEAF && rk == SRC_LEAF;
        int mxall = 0;
#pragma unroll
        for (int c = 0; c < NCM; c++) {
            double tmp[4];
#pragma unroll
            for (int x = 0; x < 4; x++) tmp[x] = __dmul_rn(l[c * 4 + x], r[c * 4 + x]);
            if (SYM) {
#pragma unroll
                for (int x = 0; x < 4; x++) tmp[x] = __dmul_rn(tmp[x], pir[x]);
            }
            int mx = 0;
#pragma unroll
            for (int k = 0; k < 4; k++) {
                double o;
                if (SYM) {
                    const double ut[4] = {Ur[k], Ur[4 + k], Ur[8 + k], Ur[12 + k]};
                    o = s4_tree(ut, tmp);
                } else {
                    o = s4_tree(mU + NCM * 16 + c * 16 + 4 * k, tmp);
                }
                prev[c * 4 + k] = o;
                // |x| < 2^-256  <=>  high word of |x| < high word of 2^-256 (whose low word is 0): the test needs only
                // the maximum of the high words -- one LOP3 + one integer max per value
                mx = max(mx, __double2hiint(o) & 0x7fffffff);
            }
            if (SAFE) {
                if (!tiptip && mx < S4_THRESHOLD_HI && invar0) {
#pragma unroll
                    for (int k = 0; k < 4; k++) prev[c * 4 + k] *= B200_TWO_256;
                    sc = __vadd4(sc, 1u << (8 * c));
                }
            } else {
                mxall = max(mxall, mx);
            }
        }
        if (!SAFE) {
            if (!tiptip && mxall < S4_THRESHOLD_HI && invar0) {
#pragma unroll
                for (int j = 0; j < BLOCK; j++) prev[j] *= B200_TWO_256;
                sc += 1;
            }
            sc &= 0xff;  // UBYTE arithmetic of the reference
        }
        }  // !tt
        prev_sc = sc;
        if (BULK) {
            // every node of a BULK program has a slot (held as a stack entry or free again at the next node)
            if ((tid & 31) == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");  // earlier copies have left shared memory
            __syncwarp();
            double2 *s = stk + stk_at<true, PAIRS>(out_stack, 0, tid);
#pragma unroll
            for (int j = 0; j < PAIRS; j++) s[j * 32] = make_double2(prev[2 * j], prev[2 * j + 1]);
            stk_sc[out_stack * S4_THREADS + tid] = sc;
            if (!store) {
            } else if (warp_full) {
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes -> visible to the copy engine
                __syncwarp();
                if ((tid & 31) == 0) {
                    const double *g = a.partial + (size_t)nd.out_slot * a.slot_doubles + (size_t)((p0 + tid) >> 5) * (32 * BLOCK);
                    const uint32_t src = (uint32_t)__cvta_generic_to_shared(s);
                    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(g), "r"(src),
                                 "r"(32 * BLOCK * 8)
                                 : "memory");
                    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                }
            } else {
                double2 *g = reinterpret_cast<double2 *>(a.partial + (size_t)nd.out_slot * a.slot_doubles + poff);
#pragma unroll
                for (int j = 0; j < PAIRS; j++) g[j * 32] = make_double2(prev[2 * j], prev[2 * j + 1]);
            }
        } else if (store) {
            double2 *g = reinterpret_cast<double2 *>(a.partial + (size_t)nd.out_slot * a.slot_doubles + poff);
#pragma unroll
            for (int j = 0; j < PAIRS; j++) g[j * 32] = make_double2(prev[2 * j], prev[2 * j + 1]);
        }
        if (store) {
            uint8_t *gs = a.scale + (size_t)nd.out_slot * a.slot_scale_bytes;
            if (SAFE) {
#pragma unroll
                for (int c = 0; c < NCM; c++) gs[p * NCM + c] = (uint8_t)(sc >> (8 * c));
            } else {
                gs[p] = (uint8_t)sc;
            }
        }
        if (!BULK && out_stack >= 0) {
#pragma unroll
            for (int j = 0; j < PAIRS; j++)
                stk[(out_stack * PAIRS + j) * S4_THREADS + tid] = make_double2(prev[2 * j], prev[2 * j + 1]);
            stk_sc[out_stack * S4_THREADS + tid] = sc;
        }
    }
    if (BULK) {  // shared memory must outlive the copies that read it
        if ((tid & 31) == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
    }
}

// State-pair tables of tip-tip nodes: tt[node][sa * codes + sb][c * 4 + k] = the partial vector s4_traverse_kernel computes
// for a pattern whose two tips carry states (sa, sb) -- the same operations in the same order (product of the two leaf
// table rows, SYM: times pi, then the back-transformation on the reference's reduction tree), so using the table changes
// no bit.  One thread per (node, state pair, category).
struct S4TTNode { int32_t ltab, rtab, out_off, pad; };  // leaf-table offsets of the two children (doubles into tables), table offset
template <int NCM, bool SYM>
__global__ void s4_tt_tables_kernel(const S4TTNode *__restrict__ nodes, const double *__restrict__ tables, double *__restrict__ tt,
                                    DevModel m, int denom, int codes, double pi0, double pi1, double pi2, double pi3) {
    constexpr int BLOCK = NCM * 4;
    const S4TTNode nd = nodes[blockIdx.y];
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= codes * codes * NCM) return;
    const int c = idx % NCM, pair = idx / NCM, sa = pair / codes, sb = pair - sa * codes;
    const double *la = tables + nd.ltab + (size_t)sa * BLOCK + c * 4, *lb = tables + nd.rtab + (size_t)sb * BLOCK + c * 4;
    const double pir[4] = {pi0, pi1, pi2, pi3};
    double tmp[4];
#pragma unroll
    for (int x = 0; x < 4; x++) {
        tmp[x] = __dmul_rn(la[x], lb[x]);
        if (SYM) tmp[x] = __dmul_rn(tmp[x], pir[x]);
    }
    const double *U = m.evec + (size_t)(c / denom) * 16, *M = m.inv_evec + (size_t)(c / denom) * 16;
    double *out = tt + nd.out_off + (size_t)pair * BLOCK + c * 4;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        if (SYM) {
            const double ut[4] = {U[k], U[4 + k], U[8 + k], U[12 + k]};
            out[k] = s4_tree(ut, tmp);
        } else {
            out[k] = s4_tree(M + 4 * k, tmp);
        }
    }
}

// ---- virtual tip-tip vectors ---------------------------------------------------------------------------------------
// A tip-tip node's vector is row (state a, state b) of its state-pair table.  When the only reader inside the launch takes
// it from registers or the stack, s4_traverse_kernel does not write it (a third of the HBM traffic of a cfg2 traversal:
// 65 of 198 nodes); the slot is then "virtual": the host keeps (table offset, taxon a, taxon b).  Whoever else needs the
// slot -- a branch-level kernel, a later traversal that reads it or rebuilds the tables, a host reader, the batched and
// sweep paths -- calls s4_realize first, which gathers ALL virtual slots from the tables that are still in place: the same
// rows the traversal would have stored, so no result changes by a bit.
#define S4_VIRT_PER_LAUNCH 120
struct S4RealizeArgs {
    const double *tt;
    const uint8_t *states;
    double *partial;
    uint8_t *scale;
    size_t slot_doubles, slot_scale_bytes;
    int64_t ntot, ntot_pad;
    int codes;
    S4Virt v[S4_VIRT_PER_LAUNCH];
};
template <int NCM, bool SAFE>
__global__ void __launch_bounds__(256) s4_realize_kernel(const __grid_constant__ S4RealizeArgs a) {
    constexpr int BLOCK = NCM * 4, PAIRS = BLOCK / 2;
    const int64_t p = (int64_t)blockIdx.x * 256 + threadIdx.x;
    if (p >= a.ntot) return;
    const S4Virt &v = a.v[blockIdx.y];
    const int sa = a.states[(size_t)v.ta * a.ntot_pad + p], sb = a.states[(size_t)v.tb * a.ntot_pad + p];
    const double2 *row = reinterpret_cast<const double2 *>(a.tt + (size_t)v.tt_off + (size_t)(sa * a.codes + sb) * BLOCK);
    const size_t poff = (size_t)(p >> 5) * (32 * BLOCK) + (size_t)(p & 31) * 2;  // T32 layout
    double2 *g = reinterpret_cast<double2 *>(a.partial + (size_t)v.slot * a.slot_doubles + poff);
#pragma unroll
    for (int j = 0; j < PAIRS; j++) g[j * 32] = __ldg(row + j);
    uint8_t *gs = a.scale + (size_t)v.slot * a.slot_scale_bytes;  // a tip-tip node is never rescaled (:1640)
    if (SAFE) {
#pragma unroll
        for (int c = 0; c < NCM; c++) gs[p * NCM + c] = 0;
    } else {
        gs[p] = 0;
    }
}

template <int NCM>
static int realize_run(b200lk_ctx *c) {
    static thread_local S4RealizeArgs *Ap = nullptr;
    if (!Ap) Ap = new S4RealizeArgs();
    S4RealizeArgs &a = *Ap;
    a.tt = c->tt_tables;
    a.states = c->states;
    a.partial = c->partial;
    a.scale = c->scale;
    a.slot_doubles = c->slot_doubles;
    a.slot_scale_bytes = c->slot_scale_bytes;
    a.ntot = c->ntot;
    a.ntot_pad = c->ntot_pad;
    a.codes = c->d.n_state_codes;
    const unsigned gx = (unsigned)((c->ntot + 255) / 256);
    int n = 0, total = 0;
    auto flush = [&]() -> int {
        if (!n) return B200LK_OK;
        if (c->d.safe_numeric) s4_realize_kernel<NCM, true><<<dim3(gx, n), 256, 0, c->stream>>>(a);
        else s4_realize_kernel<NCM, false><<<dim3(gx, n), 256, 0, c->stream>>>(a);
        CUDA_TRY(c, cudaGetLastError());
        c->n_launches++;
        total += n;
        n = 0;
        return B200LK_OK;
    };
    for (int s = 0; s < c->d.nslots; s++) {
        if (!c->virt[s]) continue;
        a.v[n++] = c->virt_desc[s];
        c->virt[s] = 0;
        if (n == S4_VIRT_PER_LAUNCH) {
            int rc = flush();
            if (rc) return rc;
        }
    }
    int rc = flush();
    if (rc) return rc;
    c->n_virtual = 0;
    c->n_min_bytes += c->ntot * (8 * (int64_t)c->block + c->scale_per_ptn) * total;  // what the traversal had left out
    return B200LK_OK;
}

int s4_realize(b200lk_ctx *c) {
    if (!c->n_virtual) return B200LK_OK;
    CUDA_TRY(c, cudaSetDevice(c->device));
    if (c->tile != 32) return dense_realize(c);
    switch (c->ncat_mix) {
        case 1: return realize_run<1>(c);
        case 2: return realize_run<2>(c);
        case 3: return realize_run<3>(c);
        default: return realize_run<4>(c);
    }
}

// Before a traversal is enqueued (it rebuilds the state-pair tables): a virtual slot that the new list overwrites without
// reading it first is simply forgotten; every other one is gathered now, while its table is still there.
int s4_before_update(b200lk_ctx *c, const b200lk_node *nodes, int nnodes, const b200lk_child *children) {
    if (!c->n_virtual) return B200LK_OK;
    std::vector<char> written(c->d.nslots, 0), read_first(c->d.nslots, 0);
    for (int i = 0; i < nnodes; i++) {
        for (int k = 0; k < nodes[i].nchild; k++) {
            const b200lk_child &ch = children[nodes[i].child_begin + k];
            if (ch.kind != B200LK_CHILD_LEAF && !written[ch.id]) read_first[ch.id] = 1;
        }
        written[nodes[i].out_slot] = 1;
    }
    for (int s = 0; s < c->d.nslots; s++)
        if (c->virt[s] && written[s] && !read_first[s]) {
            c->virt[s] = 0;
            c->n_virtual--;
        }
    return s4_realize(c);
}

// ---- batched variant: one launch for all 4-state contexts of a supermatrix (grid.y = context) -----------------------
// Same node program (the contexts share topology and traversal), but eigenvectors and exp(lambda r t) factors differ
// per context (own model, own branch lengths x partition rate): they are read from the context's device tables
// (warp-uniform loads through L1) instead of the launch parameters.  Used for many small alignments, where launch
// count, not per-launch speed, is what matters.
template <int NCM, bool PLAIN, bool SHARED>
__device__ __forceinline__ void child_apply_batch(const double *U, const double *Ur, int denom, const double *ev, double *v) {
#pragma unroll
    for (int c = 0; c < NCM; c++) {
        const double *Uc = U + (c / denom) * 16;
        double w[4], f[4];
        const double2 e0 = __ldg(reinterpret_cast<const double2 *>(ev + c * 4)), e1 = __ldg(reinterpret_cast<const double2 *>(ev + c * 4) + 1);
        w[0] = __dmul_rn(e0.x, v[c * 4]);
        w[1] = __dmul_rn(e0.y, v[c * 4 + 1]);
        w[2] = __dmul_rn(e1.x, v[c * 4 + 2]);
        w[3] = __dmul_rn(e1.y, v[c * 4 + 3]);
#pragma unroll
        for (int x = 0; x < 4; x++) {
            if (SHARED) {  // one model class: eigenvectors held in registers for the whole traversal
                f[x] = PLAIN ? s4_plain(Ur + 4 * x, w) : s4_tree(Ur + 4 * x, w);
                continue;
            }
            const double u[4] = {__ldg(Uc + 4 * x), __ldg(Uc + 4 * x + 1), __ldg(Uc + 4 * x + 2), __ldg(Uc + 4 * x + 3)};
            f[x] = PLAIN ? s4_plain(u, w) : s4_tree(u, w);
        }
#pragma unroll
        for (int x = 0; x < 4; x++) v[c * 4 + x] = f[x];
    }
}

// MINB: resident CTAs per SM the registers are sized for.  2 (220 registers, no spills) for the batches of many small genes
// it was written for -- cfg5 33.6 ms against 34.0 ms with 8 (profiles/r2j_batch_minb_cfg5.log); 8 (128 registers) when ONE
// large context runs as a batch of one (device-resident branch lengths, sweep.cu): the one-node updates of a branch sweep
// are HBM-bound streams that need the bytes in flight (cfg2 sweep 88.8 -> 81.9 ms).  B200LK_BATCH_MINB overrides for A/B.
template <int NCM, bool SAFE, bool SHARED, int MINB = 2>
__global__ void __launch_bounds__(S4_THREADS, MINB) s4_traverse_batch_kernel(const __grid_constant__ S4Launch<NCM> P, BatchLaunch bl) {
    constexpr int BLOCK = NCM * 4, PAIRS = BLOCK / 2;
    TravArgs a = P.a;
    const BatchCtxDev &bc = bl.d_ctx[blockIdx.y];
    a.tables = bc.tables;
    a.ptn_invar = bc.ptn_invar;
    a.states = bc.states;
    a.partial = bc.partial;
    a.scale = bc.scale;
    a.slot_doubles = bc.slot_doubles;
    a.slot_scale_bytes = bc.slot_scale_bytes;
    a.ntot = bc.ntot;
    a.ntot_pad = bc.ntot_pad;
    const double *Ug = bc.m.evec, *Mg = bc.m.inv_evec;
    double Ur[16], Mr[16];  // SHARED (every category uses model class 0): the context's eigenvectors, loaded once
    if (SHARED) {
#pragma unroll
        for (int k = 0; k < 16; k++) {
            Ur[k] = __ldg(Ug + k);
            Mr[k] = __ldg(Mg + k);
        }
    }
    const int64_t p0 = (int64_t)blockIdx.x * S4_THREADS;
    if (p0 >= a.ntot) return;  // smaller context than the grid
    extern __shared__ double2 stk[];
    unsigned *stk_sc = reinterpret_cast<unsigned *>(stk + (size_t)a.depth * PAIRS * S4_THREADS);
    uint8_t *st_buf = reinterpret_cast<uint8_t *>(stk_sc + (size_t)a.depth * S4_THREADS);
    const int tid = threadIdx.x;
    int64_t p = p0 + tid;
    const bool valid = p < a.ntot;
    if (!valid) p = a.ntot - 1;
    const size_t poff = (size_t)(p >> 5) * (32 * BLOCK) + (size_t)(p & 31) * 2;
    const bool invar0 = a.ptn_invar[p] == 0.0;
    double prev[BLOCK];
#pragma unroll
    for (int j = 0; j < BLOCK; j++) prev[j] = 0.0;
    unsigned prev_sc = 0;
    for (int i = 0; i < a.nnodes; i++) {
        const PNode &nd = s4_node<NCM, 2 * NCM * 16 * (int)sizeof(double)>(P, i);
        const int info = nd.info;
        const int lk = info & 3, rk = (info >> 2) & 3, out_stack = ((info >> 4) & 0xff) - 1, refill = (info >> 14) - 1;
        if (refill >= 0) {
            __syncthreads();
            constexpr int VPR = S4_THREADS / 16;
            const int nvec = a.chunk * VPR;
            const int32_t *seq = a.leaf_seq + (size_t)refill * a.chunk;
            for (int v = tid; v < nvec; v += S4_THREADS) {
                const int row = v / VPR, q = v - row * VPR;
                const uint4 *src = reinterpret_cast<const uint4 *>(a.states + (size_t)seq[row] * a.ntot_pad + p0) + q;
                reinterpret_cast<uint4 *>(st_buf)[v] = __ldg(src);
            }
            __syncthreads();
        }
        double l[BLOCK], r[BLOCK];
        unsigned sc = 0;
        child_load<NCM, SAFE>(a, lk, nd.lid, nd.ltab, prev, prev_sc, stk, stk_sc, st_buf, tid, poff, p, l, sc);
        child_load<NCM, SAFE>(a, rk, nd.rid, nd.rtab, prev, prev_sc, stk, stk_sc, st_buf, tid, poff, p, r, sc);
        if (lk != SRC_LEAF) child_apply_batch<NCM, false, SHARED>(Ug, Ur, a.denom, a.tables + nd.lexp, l);
        if (rk != SRC_LEAF) {
            if (lk == SRC_LEAF) child_apply_batch<NCM, true, SHARED>(Ug, Ur, a.denom, a.tables + nd.rexp, r);
            else child_apply_batch<NCM, false, SHARED>(Ug, Ur, a.denom, a.tables + nd.rexp, r);
        }
        const bool tiptip = lk == SRC_LEAF && rk == SRC_LEAF;
        unsigned long long mxall = 0ull;
#pragma unroll
        for (int c = 0; c < NCM; c++) {
            const double *Mc = Mg + (c / a.denom) * 16;
            double tmp[4];
#pragma unroll
            for (int x = 0; x < 4; x++) tmp[x] = __dmul_rn(l[c * 4 + x], r[c * 4 + x]);
            unsigned long long mx = 0ull;
#pragma unroll
            for (int k = 0; k < 4; k++) {
                double o;
                if (SHARED) {
                    o = s4_tree(Mr + 4 * k, tmp);
                } else {
                    const double m4[4] = {__ldg(Mc + 4 * k), __ldg(Mc + 4 * k + 1), __ldg(Mc + 4 * k + 2), __ldg(Mc + 4 * k + 3)};
                    o = s4_tree(m4, tmp);
                }
                prev[c * 4 + k] = o;
                const unsigned long long b = abs_bits(o);
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
            sc &= 0xff;
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
        if (out_stack >= 0) {
#pragma unroll
            for (int j = 0; j < PAIRS; j++)
                stk[(out_stack * PAIRS + j) * S4_THREADS + tid] = make_double2(prev[2 * j], prev[2 * j + 1]);
            stk_sc[out_stack * S4_THREADS + tid] = sc;
        }
    }
}

// ---- host side -------------------------------------------------------------------------------------------
// the model block at the start of S4Launch::blob; returns its size in bytes (a multiple of 32)
template <int NCM>
static int model_bytes(const b200lk_ctx *c, bool batch) {
    return (c->s4_sym && !batch) ? 160 : 2 * NCM * 16 * (int)sizeof(double);
}
template <int NCM>
static void fill_model(const b200lk_ctx *c, double *m, bool batch) {
    if (c->s4_sym && !batch) {  // s4_traverse_kernel<.., SYM = true>: U[16], pi[4]
        memcpy(m, c->h_evec, sizeof(double) * 16);
        for (int x = 0; x < 4; x++) m[16 + x] = c->s4_pi[x];
        return;
    }
    const int denom = c->d.fused_mix_rate ? 1 : c->d.ncat;
    for (int cc = 0; cc < NCM; cc++) {
        const int mm = cc / denom;
        memcpy(m + cc * 16, c->h_evec + (size_t)mm * 16, sizeof(double) * 16);
        memcpy(m + NCM * 16 + cc * 16, c->h_inv_evec + (size_t)mm * 16, sizeof(double) * 16);
    }
}

// inv_evec[i][x] == evec[x][i] * pi[x] to rounding for one shared eigen-system?  (reversible models: evec = Pi^-1/2 V,
// inv_evec = V^T Pi^1/2.)  pi[x] is read off the eigenvector whose entries are farthest from zero (the constant one of
// eigenvalue 0); every entry is then checked, so an eigen-system normalised any other way keeps the general kernel.
bool s4_detect_sym(const b200lk_ctx *c, double pi[4]) {
    if (c->d.nstates != 4 || c->d.nmix != 1) return false;
    const double *U = c->h_evec, *M = c->h_inv_evec;
    int i0 = 0;
    double best = -1.0;
    for (int i = 0; i < 4; i++) {
        double mn = 1e300;
        for (int x = 0; x < 4; x++) mn = std::min(mn, fabs(U[x * 4 + i]));
        if (mn > best) { best = mn; i0 = i; }
    }
    if (!(best > 1e-8)) return false;
    for (int x = 0; x < 4; x++) pi[x] = M[i0 * 4 + x] / U[x * 4 + i0];
    double scale = 0.0;
    for (int k = 0; k < 16; k++) scale = std::max(scale, fabs(M[k]));
    for (int i = 0; i < 4; i++)
        for (int x = 0; x < 4; x++)
            if (!(fabs(M[i * 4 + x] - U[x * 4 + i] * pi[x]) <= 8e-16 * scale)) return false;
    return true;
}

// expv[c][i] = exp(eval[m][i] * rate[c % ncat] * len), the factor of computePartialInfo (:1016-1031)
template <int NCM>
static void fill_expv(const b200lk_ctx *c, const DevChild &ch, double *ev) {
    const int denom = c->d.fused_mix_rate ? 1 : c->d.ncat;
    // heterotachy: one length per category (child->getLength(cat_id[c]), :1017, 1069)
    const double *lens = c->h_child_mixlen ? c->h_child_mixlen + (size_t)ch.src * c->d.ncat : nullptr;
    for (int cc = 0; cc < NCM; cc++) {
        const double len_child = c->h_rates[cc % c->d.ncat] * (lens ? lens[cc % c->d.ncat] : ch.length);
        for (int i = 0; i < 4; i++) ev[cc * 4 + i] = exp(c->h_eval[(cc / denom) * 4 + i] * len_child);
    }
}

// One launch: nodes [first, first + count) of the host's post-order list, all bifurcating, count <= MAXN.
// Stack entries are handed out here: a value goes to the stack when a later node of this launch consumes
// it and it is not simply the next node's operand; when the stack is full (or the consumer is in a later
// launch) the consumer re-reads the resident copy instead.
template <int NCM>
static int launch_segment(b200lk_ctx *c, const b200lk_node *nl, int count, int *seq_used) {
    static thread_local S4Launch<NCM> *Lp = nullptr;  // 32 KB of launch parameters, one per host thread
    if (!Lp) Lp = new S4Launch<NCM>();
    S4Launch<NCM> &P = *Lp;
    const int nslots = c->d.nslots;
    const int mdl_bytes = model_bytes<NCM>(c, c->batch != nullptr);
    PNode *const pn = reinterpret_cast<PNode *>(P.blob + mdl_bytes);                 // nodes[count]
    double *const pev = reinterpret_cast<double *>(P.blob + S4Launch<NCM>::fac_off(mdl_bytes));  // factor blocks, one per internal child
    int nev = 0;
    // unset: two entries for the SYM kernel (ten resident CTAs of 64 threads per SM), three otherwise (eight)
    const int max_depth = c->max_stack_depth >= 0 ? c->max_stack_depth : (c->s4_sym && !c->batch ? 2 : 3);
    std::vector<int> producer(nslots, -1);  // slot -> index of the node of this launch that last wrote it
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
    // bulk-store kernel (SYM, one context): a ring of shared-memory slots, one per node output; a slot stays held as a
    // stack entry only while another one is left to stage the next outputs (ring 3 = two stack entries + staging)
    const bool bulk = !c->batch && c->s4_sym && b200_env().s4_bulk;
    const int ring = b200_env().s4_ring > 0 ? b200_env().s4_ring : (c->max_stack_depth >= 0 ? c->max_stack_depth + 1 : 3);
    for (int e = (bulk ? ring : max_depth) - 1; e >= 0; e--) free_list.push_back(e);
    int depth_used = 0;
    int32_t *leaf_seq = reinterpret_cast<int32_t *>(c->h_prog) + *seq_used;
    // rows per staged chunk: the whole buffer for the batched kernel, half of it for s4_traverse_kernel (two halves in turn)
    const int chunk = c->batch ? S4_STATE_BYTES / S4_THREADS : S4_STATE_BYTES / S4_THREADS / 2;
    int nseq = 0;
    int64_t cross_reads = 0;
    // staged tip-state rows of this segment: its leaf children plus at most one pad row per node (a node never
    // straddles two chunks).  Nothing is padded at the END of a segment: a refill always reads a whole chunk of rows,
    // and the rows behind the last one used are the next segment's, or left-overs of an earlier call -- valid taxon ids
    // either way (both copies of the buffer start zeroed and only ever receive taxon ids), loaded and never looked at.
    // `prog_stride` holds children_cap + nodes_cap entries + one chunk of slack, which bounds every traversal list.
    // The staging buffer has two regions: [0, rows_cap) holds ONLY taxon ids (so whatever a refill reads is valid), and
    // [rows_cap, ...) the descriptors of the tip-tip tables, 4 ints per node, which no refill can reach.
    const size_t rows_cap = 3 * (size_t)c->nodes_cap + 512;
    if ((size_t)*seq_used + 3 * (size_t)count + chunk > rows_cap || c->s4_ttdesc_used + (size_t)count > (size_t)c->nodes_cap) {
        snprintf(c->err, sizeof(c->err), "tip-state staging sequence overflow: %d + %d entries (capacity %zu)", *seq_used,
                 3 * count + chunk, rows_cap);
        return B200LK_ERR_NOMEM;
    }
    // tip-tip nodes answered from a per-node state-pair table (s4_tt_tables_kernel): worth it when there are clearly
    // more patterns than state pairs
    const int codes = c->d.n_state_codes;
    const bool use_tt = !c->batch && b200_env().s4_tt && (b200_env().s4_tt == 2 || (int64_t)codes * codes * 4 <= c->ntot);
    const size_t tt_doubles = (size_t)codes * codes * NCM * 4;
    const bool virt_ok = use_tt && b200_env().s4_virtual && c->virt && c->cur_nreaders;
    const int *nreaders = virt_ok ? c->cur_nreaders + (nl - c->cur_nodes) : nullptr;  // of this segment's nodes, over the whole list
    std::vector<int> tip_a(count, -1), tip_b(count, -1);  // taxa of the two tips of a tip-tip node
    std::vector<S4TTNode> tt_nodes;
    std::vector<const DevChild *> exp_src(2 * (size_t)count, nullptr);
    for (int i = 0; i < count; i++) {
        PNode &f = pn[i];
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
        int refill = -1;
        if (nleaf) {  // tip-state rows are staged chunk by chunk in order of use; a node never straddles
            if (nseq % chunk + nleaf > chunk)
                while (nseq % chunk) leaf_seq[nseq++] = 0;
            if (nseq % chunk == 0) refill = nseq / chunk;
        }
        for (int k = 0; k < 2; k++) {
            const DevChild &ch = c->h_children[nl[i].child_begin + order[k]];
            tab[k] = 0;
            if (ch.kind == B200LK_CHILD_LEAF) {
                kind[k] = SRC_LEAF; id[k] = nseq % chunk; tab[k] = (int)ch.leaf_off;
                leaf_seq[nseq++] = ch.id;
            } else {
                // exp(lambda r t) of this child: its own block of the factor array, filled below (all children in parallel);
                // batched launches read the same factors from the context's tables instead
                if (c->batch) {
                    (k == 0 ? f.lexp : f.rexp) = (int32_t)(ch.echild_off + (int64_t)c->block * c->d.nstates);
                } else {
                    (k == 0 ? f.lexp : f.rexp) = nev;
                    exp_src[nev++] = &ch;
                }
                const int pr = producer[ch.id];
                if (pr >= 0 && consumer[pr] == i && pr == i - 1) {
                    kind[k] = SRC_PREV; id[k] = 0;
                    if (virt_ok && nreaders[pr] == 1 && ((pn[pr].info >> 12) & 1)) pn[pr].info |= 1 << 13;  // tip-tip, read from registers: not stored
                } else if (pr >= 0 && consumer[pr] == i && entry_of[pr] >= 0) {
                    kind[k] = SRC_STACK; id[k] = entry_of[pr];
                    free_list.push_back(entry_of[pr]);
                    entry_of[pr] = -1;
                    if (virt_ok && nreaders[pr] == 1 && ((pn[pr].info >> 12) & 1)) pn[pr].info |= 1 << 13;  // ... or from the stack
                } else {
                    kind[k] = SRC_GLOBAL; id[k] = ch.id;
                    // produced by an EARLIER launch (or call): a DRAM read.  A vector of this launch that no longer
                    // fits the stack is re-read a few nodes after the same CTA wrote it: an L2 hit, not DRAM traffic
                    if (pr < 0) cross_reads++;
                }
            }
        }
        f.lid = id[0]; f.ltab = tab[0];
        f.rid = id[1]; f.rtab = tab[1];
        int out_stack = -1;
        const bool wanted = consumer[i] >= 0 && consumer[i] != i + 1;
        if (wanted && b200_env().s4_evict && free_list.size() <= (bulk ? 1u : 0u) && (bulk ? ring > 1 : max_depth > 0)) {
            // Stack full.  Every vector is in HBM anyway, so dropping an entry costs nothing now and one re-read at its
            // consumer: keep the vectors that are wanted soonest (Belady) -- the held one whose consumer is farthest away
            // gives its entry up if this node's consumer comes first.  cfg2: 14 -> 5 re-reads per traversal with two
            // entries, 43 -> 18 with one.
            int far = -1;
            for (int h = 0; h < i; h++)
                if (entry_of[h] >= 0 && (far < 0 || consumer[h] > consumer[far])) far = h;
            if (far >= 0 && consumer[far] > consumer[i]) {
                free_list.push_back(entry_of[far]);
                entry_of[far] = -1;
                if (!bulk) pn[far].info &= ~(0xff << 4);  // its push is no longer needed
            }
        }
        if (bulk) {
            out_stack = free_list.back();  // never empty: at most ring - 1 slots are held
            free_list.pop_back();
            if (wanted && !free_list.empty()) entry_of[i] = out_stack;
            else free_list.insert(free_list.begin(), out_stack);  // staging only: free again, reused last
            depth_used = std::max(depth_used, out_stack + 1);
        } else if (wanted && !free_list.empty()) {
            out_stack = free_list.back();
            free_list.pop_back();
            entry_of[i] = out_stack;
            depth_used = std::max(depth_used, out_stack + 1);
        }
        int tt = 0;
        if (use_tt && kind[0] == SRC_LEAF && kind[1] == SRC_LEAF && c->s4_tt_used + tt_doubles <= (size_t)INT32_MAX) {
            S4TTNode tn = {tab[0], tab[1], (int32_t)c->s4_tt_used, 0};
            tt_nodes.push_back(tn);
            tip_a[i] = c->h_children[nl[i].child_begin + order[0]].id;
            tip_b[i] = c->h_children[nl[i].child_begin + order[1]].id;
            f.ltab = (int32_t)c->s4_tt_used;
            c->s4_tt_used += tt_doubles;
            tt = 1;
        }
        f.info = kind[0] | kind[1] << 2 | (out_stack + 1) << 4 | tt << 12 | (refill + 1) << 14;  // bit 13 (virtual) is set by the consumer
        producer[nl[i].out_slot] = i;
    }
    // launch_run sized the segment for the window; nothing has been written behind the node array yet
    if (c->batch ? mdl_bytes + 32 * count > S4Launch<NCM>::BLOB
                 : (count > S4Launch<NCM>::nmax(mdl_bytes) || nev > S4Launch<NCM>::nmax(mdl_bytes))) {
        snprintf(c->err, sizeof(c->err), "node program of %d nodes / %d internal children exceeds the parameter window", count, nev);
        return B200LK_ERR_ARG;
    }
    // which slots this launch leaves virtual (the last writer of a slot decides)
    int nvirtual = 0;
    if (c->virt)
        for (int i = 0; i < count; i++) {
            const int slot = nl[i].out_slot;
            const bool v = (pn[i].info >> 13) & 1;
            if (c->virt[slot] && !v) c->n_virtual--;
            if (!c->virt[slot] && v) c->n_virtual++;
            c->virt[slot] = v;
            if (v) {
                const S4Virt d = {slot, pn[i].ltab, tip_a[i], tip_b[i]};
                c->virt_desc[slot] = d;
                nvirtual++;
            }
        }
    // the per-child factors: 16 exp() per internal child, ~6000 for a cfg2 traversal -- what a blocking evaluation spends
    // most of its host time on, so a few threads share it (the calling thread alone for short lists)
    {
        const int nexp = nev;
        // at most four threads, and never more than OMP_NUM_THREADS allows: torchrun sets it to 1 for every rank, and eight
        // ranks forking four workers each onto the same pinned cores cost 70 ms per step instead of saving 0.03
        const int nthr = std::min(4, omp_get_max_threads());
#pragma omp parallel for schedule(static) num_threads(nthr) if (nexp >= 64 && nthr > 1)
        for (int j = 0; j < nexp; j++)
            fill_expv<NCM>(c, *exp_src[j], pev + (size_t)j * (NCM * 4));
    }
    int32_t *d_seq = reinterpret_cast<int32_t *>(c->d_prog) + *seq_used;
    if (nseq)
        CUDA_TRY(c, cudaMemcpyAsync(d_seq, leaf_seq, sizeof(int32_t) * nseq, cudaMemcpyHostToDevice, c->stream));
    *seq_used += nseq;
    // the descriptors of the tip-tip tables: their own region of the staging buffer, their own copy
    const int ntt = (int)tt_nodes.size();
    const size_t desc_at = rows_cap + 4 * c->s4_ttdesc_used;  // ints from the start of the buffer (16-byte aligned)
    if (ntt) {
        int32_t *h_desc = reinterpret_cast<int32_t *>(c->h_prog) + desc_at;
        memcpy(h_desc, tt_nodes.data(), sizeof(S4TTNode) * ntt);
        CUDA_TRY(c, cudaMemcpyAsync(reinterpret_cast<int32_t *>(c->d_prog) + desc_at, h_desc, sizeof(S4TTNode) * ntt,
                                    cudaMemcpyHostToDevice, c->stream));
        c->s4_ttdesc_used += ntt;
    }
    if (ntt) {
        if (c->s4_tt_used > c->tt_cap) {  // grows only: a larger traversal than any before (cudaFree waits for the device)
            const size_t want = std::max<size_t>(c->s4_tt_used, 2 * tt_doubles * (size_t)(c->d.ntaxa / 2 + 1));
            if (c->tt_tables) CUDA_TRY(c, cudaFree(c->tt_tables));
            c->tt_tables = nullptr;
            c->tt_cap = 0;
            CUDA_TRY(c, cudaMalloc(&c->tt_tables, sizeof(double) * want));
            c->tt_cap = want;
        }
        const S4TTNode *d_tt = reinterpret_cast<const S4TTNode *>(reinterpret_cast<int32_t *>(c->d_prog) + desc_at);
        const dim3 g((codes * codes * NCM + 127) / 128, ntt);
        const int denom2 = c->d.fused_mix_rate ? 1 : c->d.ncat;
        if (c->s4_sym)
            s4_tt_tables_kernel<NCM, true><<<g, 128, 0, c->stream>>>(d_tt, c->tables, c->tt_tables, c->m, denom2, codes, c->s4_pi[0],
                                                                   c->s4_pi[1], c->s4_pi[2], c->s4_pi[3]);
        else
            s4_tt_tables_kernel<NCM, false><<<g, 128, 0, c->stream>>>(d_tt, c->tables, c->tt_tables, c->m, denom2, codes, 0, 0, 0, 0);
        CUDA_TRY(c, cudaGetLastError());
        c->n_launches++;
    }
    TravArgs &a = P.a;
    a.leaf_seq = d_seq;
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
    a.nchunks = (nseq + chunk - 1) / chunk;
    a.denom = c->d.fused_mix_rate ? 1 : c->d.ncat;
    a.tt = c->tt_tables;
    a.codes = codes;
    a.prefetch = b200_env().s4_prefetch && !c->batch;
    fill_model<NCM>(c, reinterpret_cast<double *>(P.blob), c->batch != nullptr);

    const size_t smem = (size_t)depth_used * S4_THREADS * (2 * NCM * sizeof(double2) + sizeof(unsigned)) + S4_STATE_BYTES;
    const unsigned grid = (unsigned)(((c->batch ? c->batch->max_ntot : c->ntot) + S4_THREADS - 1) / S4_THREADS);
    if (!c->trav_attr_set) {
        CUDA_TRY(c, cudaFuncSetAttribute(s4_traverse_batch_kernel<NCM, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
        CUDA_TRY(c, cudaFuncSetAttribute(s4_traverse_batch_kernel<NCM, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
        CUDA_TRY(c, cudaFuncSetAttribute(s4_traverse_batch_kernel<NCM, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
        CUDA_TRY(c, cudaFuncSetAttribute(s4_traverse_batch_kernel<NCM, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
        CUDA_TRY(c, cudaFuncSetAttribute(s4_traverse_kernel<NCM, false, false, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
        CUDA_TRY(c, cudaFuncSetAttribute(s4_traverse_kernel<NCM, true, false, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
        CUDA_TRY(c, cudaFuncSetAttribute(s4_traverse_kernel<NCM, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
        CUDA_TRY(c, cudaFuncSetAttribute(s4_traverse_kernel<NCM, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
        c->trav_attr_set = true;
    }
    if (!c->batch) {
        const int64_t vec_bytes = 8 * (int64_t)c->block + c->scale_per_ptn;
        c->n_min_bytes += c->ntot * (vec_bytes * (count - nvirtual + cross_reads) + nseq);
    }
    prof_mark(c);
    if (c->batch) {
        const dim3 g2(grid, c->batch->n);
        const bool safe = c->d.safe_numeric != 0, one_class = c->d.nmix == 1;
        auto kern = one_class ? (safe ? s4_traverse_batch_kernel<NCM, true, true> : s4_traverse_batch_kernel<NCM, false, true>)
                              : (safe ? s4_traverse_batch_kernel<NCM, true, false> : s4_traverse_batch_kernel<NCM, false, false>);
        const int minb = b200_env().batch_minb ? b200_env().batch_minb : (c->batch == &c->self_bl ? 8 : 2);
        if (minb == 8)
            kern = one_class ? (safe ? s4_traverse_batch_kernel<NCM, true, true, 8> : s4_traverse_batch_kernel<NCM, false, true, 8>)
                             : (safe ? s4_traverse_batch_kernel<NCM, true, false, 8> : s4_traverse_batch_kernel<NCM, false, false, 8>);
        kern<<<g2, S4_THREADS, smem, c->stream>>>(P, *c->batch);
    } else if (bulk) {
        // ring 3 (29 KB per CTA): seven resident CTAs, the 128-register build; ring 2 (21 KB): ten, the 96-register one
        const int minb = b200_env().s4_minb ? b200_env().s4_minb : (depth_used >= 3 ? 8 : S4_MINB);
        const bool safe = c->d.safe_numeric != 0;
        auto kern = safe ? s4_traverse_kernel<NCM, true, true, 8, true> : s4_traverse_kernel<NCM, false, true, 8, true>;
        if (minb >= 10) kern = safe ? s4_traverse_kernel<NCM, true, true, 10, true> : s4_traverse_kernel<NCM, false, true, 10, true>;
        if (minb <= 6) kern = safe ? s4_traverse_kernel<NCM, true, true, 6, true> : s4_traverse_kernel<NCM, false, true, 6, true>;
        CUDA_TRY(c, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
        kern<<<grid, S4_THREADS, smem, c->stream>>>(P);
    } else if (c->s4_sym) {
        // resident CTAs per SM the registers are sized for: 10 x 64 threads (96 registers, two stack entries) unless overridden
        const int minb = b200_env().s4_minb ? b200_env().s4_minb : S4_MINB;  // test / sweep hook: the 8- and 12-CTA instantiations
        const bool safe = c->d.safe_numeric != 0;
        auto kern = safe ? s4_traverse_kernel<NCM, true, true, S4_MINB> : s4_traverse_kernel<NCM, false, true, S4_MINB>;
        if (minb == 8) kern = safe ? s4_traverse_kernel<NCM, true, true, 8> : s4_traverse_kernel<NCM, false, true, 8>;
        if (minb == 12) kern = safe ? s4_traverse_kernel<NCM, true, true, 12> : s4_traverse_kernel<NCM, false, true, 12>;
        if (minb == 14) kern = safe ? s4_traverse_kernel<NCM, true, true, 14> : s4_traverse_kernel<NCM, false, true, 14>;
        if (minb == 16) kern = safe ? s4_traverse_kernel<NCM, true, true, 16> : s4_traverse_kernel<NCM, false, true, 16>;
        if (minb != S4_MINB) CUDA_TRY(c, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
        kern<<<grid, S4_THREADS, smem, c->stream>>>(P);
    } else if (c->d.safe_numeric) s4_traverse_kernel<NCM, true, false, 8><<<grid, S4_THREADS, smem, c->stream>>>(P);
    else s4_traverse_kernel<NCM, false, false, 8><<<grid, S4_THREADS, smem, c->stream>>>(P);
    prof_mark(c);
    CUDA_TRY(c, cudaGetLastError());
    c->n_launches++;
    return B200LK_OK;
}

template <int NCM>
static int launch_run(b200lk_ctx *c, const b200lk_node *nl, int count, int *seq_used) {
    // What a launch can carry: `cap` nodes and, unless it is a batched launch (which reads the factors from the tables),
    // `cap` factor blocks -- one per internal child.  pre[i] = internal children of nodes [0, i).
    const int mdl_bytes = model_bytes<NCM>(c, c->batch != nullptr);
    int cap = c->batch ? (S4Launch<NCM>::BLOB - mdl_bytes) / 32 : S4Launch<NCM>::nmax(mdl_bytes);
    if (b200_env().s4_budget > 0) cap = std::max(2, std::min(cap, b200_env().s4_budget));  // A/B hook: shorter launches
    std::vector<int> pre(count + 1, 0);
    for (int i = 0; i < count; i++) {
        int internal = 0;
        if (!c->batch)
            for (int k = 0; k < 2; k++) internal += c->h_children[nl[i].child_begin + k].kind != B200LK_CHILD_LEAF;
        pre[i + 1] = pre[i] + internal;
    }
    auto fits = [&](int j, int e) { return e - j <= cap && pre[e] - pre[j] <= cap; };  // nodes [j, e) in one launch?
    // A list longer than one launch's parameter space is cut where the fewest vectors are live across the cut (each
    // of them is re-read from HBM by the next launch instead of coming from registers or the stack):
    // live[s] = #{i < s : the node consuming i's output has index >= s}.
    std::vector<int> live, maxfit, need;
    if (!fits(0, count)) {
        std::vector<int> producer(c->d.nslots, -1), diff(count + 2, 0);
        for (int i = 0; i < count; i++) {
            for (int k = 0; k < 2; k++) {
                const DevChild &ch = c->h_children[nl[i].child_begin + k];
                if (ch.kind == B200LK_CHILD_LEAF || producer[ch.id] < 0) continue;
                diff[producer[ch.id] + 1]++;  // cuts s in (producer, i] separate the two
                diff[i + 1]--;
                producer[ch.id] = -1;
            }
            producer[nl[i].out_slot] = i;
        }
        live.assign(count + 1, 0);
        for (int s2 = 1; s2 <= count; s2++) live[s2] = live[s2 - 1] + diff[s2];
        // maxfit[j] = longest segment starting at j that fits; need[j] = fewest launches for [j, count)
        maxfit.assign(count + 1, 0);
        need.assign(count + 1, 0);
        for (int j = 0, e = 0; j < count; j++) {
            if (e < j) e = j;
            while (e < count && fits(j, e + 1)) e++;
            maxfit[j] = e - j;
        }
        for (int j = count - 1; j >= 0; j--) need[j] = 1 + need[j + std::max(1, maxfit[j])];
    }
    for (int done = 0; done < count;) {
        int seg = count - done;
        if (!fits(done, count)) {
            seg = std::max(1, maxfit[done]);
            // any shorter segment that keeps the launch count minimal and leaves fewer vectors live across the cut
            for (int t = 1; t <= maxfit[done]; t++)
                if (need[done + t] == need[done] - 1 && live[done + t] <= live[done + seg]) seg = t;
        }
        int rc = launch_segment<NCM>(c, nl + done, seg, seq_used);
        if (rc) return rc;
        done += seg;
    }
    return B200LK_OK;
}

// nodes [first, first + count) of the host's post-order list, all bifurcating (children already carry their
// leaf-table offsets in c->h_children)
int launch_s4_traverse(b200lk_ctx *c, const b200lk_node *nodes, int first, int count, int *prog_used, int *seq_used) {
    (void)prog_used;
    switch (c->ncat_mix) {
        case 1: return launch_run<1>(c, nodes + first, count, seq_used);
        case 2: return launch_run<2>(c, nodes + first, count, seq_used);
        case 3: return launch_run<3>(c, nodes + first, count, seq_used);
        default: return launch_run<4>(c, nodes + first, count, seq_used);
    }
}
