// branch.cu -- branch-level steps of the likelihood path, all HBM-bound streams:
//   theta build       computeLikelihoodBuffer*SIMD     tree/phylokernelnew.h:2050-2231
//   lnL at a branch   computeLikelihoodBranch*SIMD     tree/phylokernelnew.h:2643-3185
//   df, ddf           computeLikelihoodDerv*SIMD       tree/phylokernelnew.h:2235-2633
//   lnL from theta    computeLikelihoodFromBuffer*SIMD tree/phylokernelnew.h:3196-3392
//
// A group of G lanes (power of two <= 32) owns one pattern and strides over its flat
// [block] vector so that consecutive lanes read consecutive doubles; per-pattern sums
// finish with warp shuffles.  Sums over patterns are deterministic: fixed grid, fixed
// per-thread order, shuffle tree, then a fixed-order pass over the per-CTA partials.
#include <cooperative_groups.h>

#include <algorithm>

#include "common.cuh"
#include "peer.cuh"

#define RED_THREADS 256
// element e of the current pattern p under the context's layout (common.cuh)
#define E(e) elem_off(p, (e), block, a.tile)

struct BranchArgs {
    const double *pdad;    // dad_branch->partial_lh  (subtree on the node side)
    const double *pnode;   // node_branch->partial_lh (subtree on the dad side), null if dad is a leaf
    const uint8_t *sdad, *snode;
    const uint8_t *dad_states;  // tip states of the leaf dad
    const double *tip_partial_lh, *val, *tip_val;
    const double *ptn_freq, *ptn_invar;
    double *theta, *buffer_scale_all, *pattern_lh, *pattern_lh_cat;
    double *red_partial;
    int64_t nptn, ntot;
    int n, ncm, block, denom, nmix, G, safe, tile;
    int slot_pdad, slot_pnode, dad_taxon;  // batched launches rebuild the pointers above per context from these
    // site-specific models (SITE_MODEL variants): the branch coefficients depend on the pattern,
    //   val_p[c][i] = exp(eval_p[i] * rate_c * len) * prop_c   (tree/phylokernelnew.h:2469-2487, 2815-2831, 3278-3289)
    // and a leaf dad contributes its per-pattern tip vector site_tip[taxon][p][i]
    const double *site_eval, *site_tip, *rates, *props;
    double len;
    // Holder's +ASC: the unobserved patterns keep their own lh (in pattern_lh), df and ddf (:2522-2525)
    double *const_df, *const_ddf;
};

// coefficient of flat element j = c*n + i of pattern p: val0 (which = 0), val1 = cof*val0, val2 = cof*val1
__device__ __forceinline__ void branch_coef(const BranchArgs &a, int64_t p, int j, double &v0, double &v1, double &v2) {
    if (!a.site_eval) {
        v0 = a.val[j];
        v1 = a.val[a.block + j];
        v2 = a.val[2 * a.block + j];
        return;
    }
    const int c = j / a.n, i = j - c * a.n;
    const double cof = a.site_eval[(size_t)p * a.n + i] * a.rates[c];
    v0 = exp(cof * a.len) * a.props[c];
    v1 = cof * v0;
    v2 = cof * v1;
}

// batched launch (grid.y = context of the supermatrix): same branch, same slots, the context's own buffers
__device__ __forceinline__ void rebind_branch(BranchArgs &a, const BatchLaunch &bl) {
    if (!bl.d_ctx) return;
    const BatchCtxDev &b = bl.d_ctx[blockIdx.y];
    if (a.slot_pdad >= 0) {
        a.pdad = b.partial + (size_t)a.slot_pdad * b.slot_doubles;
        a.sdad = b.scale + (size_t)a.slot_pdad * b.slot_scale_bytes;
        if (a.slot_pnode >= 0) {
            a.pnode = b.partial + (size_t)a.slot_pnode * b.slot_doubles;
            a.snode = b.scale + (size_t)a.slot_pnode * b.slot_scale_bytes;
        } else {
            a.dad_states = b.states + (size_t)a.dad_taxon * b.ntot_pad;
        }
    }
    a.tip_partial_lh = b.m.tip_partial_lh;
    a.val = b.val;
    a.tip_val = b.tip_val;
    a.ptn_freq = b.ptn_freq;
    a.ptn_invar = b.ptn_invar;
    a.theta = b.theta;
    a.buffer_scale_all = b.buffer_scale_all;
    a.pattern_lh = b.pattern_lh;
    a.pattern_lh_cat = b.pattern_lh_cat;
    a.red_partial = b.red_partial;
    a.nptn = b.nptn;
    a.ntot = b.ntot;
}

// val0[c][i] = exp(eval*rate*len)*prop*mixw, val1 = cof*val0, val2 = cof*val1 (:2342-2357);
// tip_val[s][c][i] = val0[c][i] * tip_partial_lh[s][m][i] (:2757-2785)
__global__ void val_kernel(DevModel m, double len_, const double *__restrict__ lens, double *val, double *tip_val,
                           int n, int ncat, int ncm, int denom, int nmix, int codes, int want_tip, BatchLaunch bl) {
    if (bl.d_ctx) {
        const BatchCtxDev &b = bl.d_ctx[blockIdx.y];
        m = b.m;
        val = b.val;
        tip_val = b.tip_val;
        len_ = bl.d_branch_len[blockIdx.y];
    }
    const int block = n * ncm;
    // every CTA owns a slice of the block (mixtures: block = 4800); tip_val recomputes its val0 entry instead of
    // waiting for another CTA's
    const int gtid = blockIdx.x * blockDim.x + threadIdx.x, gsz = gridDim.x * blockDim.x;
    for (int idx = gtid; idx < block; idx += gsz) {
        int c = idx / n, i = idx - c * n, mm = c / denom, k = c % ncat;
        double prop = m.props[k] * m.mixw[mm];
        double cof = m.eval[mm * n + i] * m.rates[k];
        const double len = lens ? lens[k] : len_;  // heterotachy: dad_branch->getLength(mycat), :2325, 2711, 3248
        double v = exp(cof * len) * prop;
        double v1 = cof * v;
        val[idx] = v;
        val[block + idx] = v1;
        val[2 * block + idx] = cof * v1;
        if (want_tip) {
            const int tip_block = n * nmix;
            for (int s = 0; s < codes; s++)
                tip_val[(size_t)s * block + idx] = v * m.tip_partial_lh[(size_t)s * tip_block + mm * n + i];
        }
    }
}

template <int K>
__device__ __forceinline__ void block_reduce_store(double (&acc)[K], double *red_partial, int red_blocks) {
    __shared__ double sh[K][RED_THREADS / 32];
#pragma unroll
    for (int k = 0; k < K; k++) {
        double v = acc[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0) sh[k][threadIdx.x >> 5] = v;
    }
    __syncthreads();
    if (threadIdx.x < K) {
        double v = 0.0;
        for (int w = 0; w < RED_THREADS / 32; w++) v += sh[threadIdx.x][w];
        red_partial[(size_t)threadIdx.x * red_blocks + blockIdx.x] = v;
    }
}

// fixed-order final pass: result[k] = sum_b red_partial[k][b]
// (sums k < K0 go to result[dst0 + k], the +ASC sums k >= K0 to result[dst1 + k - K0])
__global__ void final_reduce_kernel(const double *red_partial, int red_blocks, int K, double *result, int dst0,
                                    int K0, int dst1, BatchLaunch bl) {
    __shared__ double sh[RED_THREADS];
    if (bl.d_ctx) {
        red_partial = bl.d_ctx[blockIdx.y].red_partial;
        result = bl.d_ctx[blockIdx.y].d_result;
    }
    for (int k = 0; k < K; k++) {
        double v = 0.0;
        for (int b = threadIdx.x; b < red_blocks; b += RED_THREADS) v += red_partial[(size_t)k * red_blocks + b];
        sh[threadIdx.x] = v;
        __syncthreads();
        for (int s = RED_THREADS / 2; s > 0; s >>= 1) {
            if (threadIdx.x < s) sh[threadIdx.x] += sh[threadIdx.x + s];
            __syncthreads();
        }
        if (threadIdx.x == 0) result[k < K0 ? dst0 + k : dst1 + k - K0] = sh[0];
        __syncthreads();
    }
}

__device__ __forceinline__ double group_sum(double v, int G) {
    for (int o = G >> 1; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ int group_min(int v, int G) {
    for (int o = G >> 1; o > 0; o >>= 1) v = min(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// ---- T32 fast paths (4-state contexts, tile == 32) ---------------------------------------------------------------
// The generic kernels below give a pattern to a group of lanes that stride over its block; on the T32 layout that turns
// every warp access into eight scattered 32-byte sectors (measured on cfg2: derv 0.56 TB/s, theta 2.4 TB/s).  Here one
// THREAD owns a pattern and reads its PAIRS = block/2 element pairs with 128-bit loads, so a warp access is one
// contiguous 512-byte segment, exactly like the traversal kernel that wrote the vectors (s4.cu).
__device__ __forceinline__ const double2 *t32_ptr(const double *base, int64_t p, int pairs) {
    return reinterpret_cast<const double2 *>(base + (size_t)(p >> 5) * (64 * (size_t)pairs) + (size_t)(p & 31) * 2);
}
template <int PAIRS>
__device__ __forceinline__ void t32_load(const double *base, int64_t p, double *v) {
    const double2 *g = t32_ptr(base, p, PAIRS);
#pragma unroll
    for (int j = 0; j < PAIRS; j++) {
        const double2 u = __ldcs(g + j * 32);
        v[2 * j] = u.x;
        v[2 * j + 1] = u.y;
    }
}
// lh, df, ddf = sum_e val{0,1,2}[e] * theta[p][e]; sv = [3][block] in shared memory
template <int PAIRS>
__device__ __forceinline__ void t32_triple_regs(const double *t, const double *sv, double &lh, double &df, double &ddf) {
    constexpr int B = 2 * PAIRS;
    double l0 = 0.0, l1 = 0.0, d0 = 0.0, d1 = 0.0, e0 = 0.0, e1 = 0.0;  // two interleaved accumulators, as dotProductTriple
#pragma unroll
    for (int j = 0; j < PAIRS; j++) {
        l0 = fma(sv[2 * j], t[2 * j], l0);
        l1 = fma(sv[2 * j + 1], t[2 * j + 1], l1);
        d0 = fma(sv[B + 2 * j], t[2 * j], d0);
        d1 = fma(sv[B + 2 * j + 1], t[2 * j + 1], d1);
        e0 = fma(sv[2 * B + 2 * j], t[2 * j], e0);
        e1 = fma(sv[2 * B + 2 * j + 1], t[2 * j + 1], e1);
    }
    lh = l0 + l1;
    df = d0 + d1;
    ddf = e0 + e1;
}
template <int PAIRS>
__device__ __forceinline__ void t32_triple(const double *theta, int64_t p, const double *sv, double &lh, double &df,
                                           double &ddf) {
    double t[2 * PAIRS];
    t32_load<PAIRS>(theta, p, t);
    t32_triple_regs<PAIRS>(t, sv, lh, df, ddf);
}
// per-pattern epilogue of computeLikelihoodDerv (:2494-2534), shared by the generic and the T32 kernels
__device__ __forceinline__ void derv_accumulate(const BranchArgs &a, int64_t p, double lh, double df, double ddf,
                                                double (&acc)[5]) {
    lh = fabs(lh) + a.ptn_invar[p];  // :2494
    if (p < a.nptn) {
        double inv = 1.0 / lh;           // :2497-2505
        double df_frac = df * inv, ddf_frac = ddf * inv, freq = a.ptn_freq[p];
        double tmp1 = df_frac * freq, tmp2 = ddf_frac * freq;
        acc[0] += tmp1;
        acc[1] += fma(-tmp1, df_frac, tmp2);
    } else {  // +ASC unobserved pattern, :2506-2534
        if (a.buffer_scale_all[p] != 0.0) {
            lh *= B200_SCALING_THRESHOLD;
            df *= B200_SCALING_THRESHOLD;
            ddf *= B200_SCALING_THRESHOLD;
        }
        if (a.const_df) {  // Holder: kept per pattern, combined by holder_kernel
            a.pattern_lh[p] = lh;
            a.const_df[p - a.nptn] = df;
            a.const_ddf[p - a.nptn] = ddf;
        }
        acc[2] += lh;
        acc[3] += df;
        acc[4] += ddf;
    }
}
template <int PAIRS>
__device__ __forceinline__ void derv_t32(const BranchArgs &a, int red_blocks) {
    __shared__ double sv[48];
    if (threadIdx.x < 6 * PAIRS) sv[threadIdx.x] = a.val[threadIdx.x];
    __syncthreads();
    double acc[5] = {0.0, 0.0, 0.0, 0.0, 0.0};
    for (int64_t p = (int64_t)blockIdx.x * RED_THREADS + threadIdx.x; p < a.ntot; p += (int64_t)gridDim.x * RED_THREADS) {
        double lh, df, ddf;
        t32_triple<PAIRS>(a.theta, p, sv, lh, df, ddf);
        derv_accumulate(a, p, lh, df, ddf, acc);
    }
    if (a.ntot > a.nptn) block_reduce_store<5>(acc, a.red_partial, red_blocks);
    else {
        double acc2[2] = {acc[0], acc[1]};
        block_reduce_store<2>(acc2, a.red_partial, red_blocks);
    }
}
template <int PAIRS>
__device__ __forceinline__ void frombuf_t32(const BranchArgs &a, int red_blocks) {
    __shared__ double sv[16];
    if (threadIdx.x < 2 * PAIRS) sv[threadIdx.x] = a.val[threadIdx.x];
    __syncthreads();
    double acc[2] = {0.0, 0.0};
    for (int64_t p = (int64_t)blockIdx.x * RED_THREADS + threadIdx.x; p < a.ntot; p += (int64_t)gridDim.x * RED_THREADS) {
        double t[2 * PAIRS];
        t32_load<PAIRS>(a.theta, p, t);
        double l0 = 0.0, l1 = 0.0;
#pragma unroll
        for (int j = 0; j < PAIRS; j++) {
            l0 = fma(sv[2 * j], t[2 * j], l0);
            l1 = fma(sv[2 * j + 1], t[2 * j + 1], l1);
        }
        double lh = fabs(l0 + l1) + a.ptn_invar[p];  // :3295
        if (p < a.nptn) {
            lh = log(fabs(lh)) + a.buffer_scale_all[p];  // :3300
            a.pattern_lh[p] = lh;
            acc[0] = fma(lh, a.ptn_freq[p], acc[0]);
        } else {  // +ASC unobserved pattern
            if (a.buffer_scale_all[p] != 0.0) lh *= B200_SCALING_THRESHOLD;
            a.pattern_lh[p] = lh;
            acc[1] += lh;
        }
    }
    block_reduce_store<2>(acc, a.red_partial, red_blocks);
}
// theta (computeLikelihoodBuffer) and the branch lnL on T32: NCM = categories = PAIRS / 2
// one pattern: theta[p] is stored AND left in th_out (the fused first Newton-Raphson evaluation reads it from there)
template <int PAIRS>
__device__ __forceinline__ void theta_t32_pattern(const BranchArgs &a, int64_t p, bool tip_mode, double *th_out) {
    constexpr int B = 2 * PAIRS, NCM = PAIRS / 2;
    {
        double d[B], o[B];
        t32_load<PAIRS>(a.pdad, p, d);
        int sc[NCM], min_scale;
        if (tip_mode) {
            const double2 *tp = reinterpret_cast<const double2 *>(a.tip_partial_lh + (size_t)a.dad_states[p] * (4 * a.nmix));
#pragma unroll
            for (int c = 0; c < NCM; c++) {  // tip vector of the category's mixture class
                const double2 u0 = __ldg(tp + (c / a.denom) * 2), u1 = __ldg(tp + (c / a.denom) * 2 + 1);
                o[4 * c] = u0.x; o[4 * c + 1] = u0.y; o[4 * c + 2] = u1.x; o[4 * c + 3] = u1.y;
            }
        } else {
            t32_load<PAIRS>(a.pnode, p, o);
        }
        if (a.safe) {
            min_scale = 0x7fffffff;
#pragma unroll
            for (int c = 0; c < NCM; c++) {
                sc[c] = (a.sdad[p * NCM + c] + (tip_mode ? 0 : a.snode[p * NCM + c])) & 0xff;
                min_scale = min(min_scale, sc[c]);
            }
        } else {
            min_scale = a.sdad[p] + (tip_mode ? 0 : a.snode[p]);
        }
        double2 *th = const_cast<double2 *>(t32_ptr(a.theta, p, PAIRS));
#pragma unroll
        for (int j = 0; j < PAIRS; j++) {
            double t0 = o[2 * j] * d[2 * j], t1 = o[2 * j + 1] * d[2 * j + 1];
            if (a.safe) {
                const int s = sc[j / 2];
                if (s == min_scale + 1) { t0 *= B200_SCALING_THRESHOLD; t1 *= B200_SCALING_THRESHOLD; }
                else if (s > min_scale + 1) { t0 = 0.0; t1 = 0.0; }
            }
            th[j * 32] = make_double2(t0, t1);
            th_out[2 * j] = t0;
            th_out[2 * j + 1] = t1;
        }
        a.buffer_scale_all[p] = (double)min_scale * B200_LOG_SCALING_THRESHOLD;
    }
}
template <int PAIRS>
__device__ __forceinline__ void theta_t32(const BranchArgs &a) {
    const bool tip_mode = a.pnode == nullptr;
    for (int64_t p = (int64_t)blockIdx.x * RED_THREADS + threadIdx.x; p < a.ntot; p += (int64_t)gridDim.x * RED_THREADS) {
        double t[2 * PAIRS];
        theta_t32_pattern<PAIRS>(a, p, tip_mode, t);
    }
}
// first evaluation of the device Newton-Raphson loop fused with the theta pass: theta is formed from the two partial
// vectors, stored for the later iterations, and its triple product taken from registers
template <int PAIRS>
__device__ __forceinline__ void theta_triple_t32(const BranchArgs &a, int64_t p, bool tip_mode, const double *sv, double &lh,
                                                 double &df, double &ddf) {
    double t[2 * PAIRS];
    theta_t32_pattern<PAIRS>(a, p, tip_mode, t);
    t32_triple_regs<PAIRS>(t, sv, lh, df, ddf);
}
template <int PAIRS>
__device__ __forceinline__ void lnl_t32(const BranchArgs &a, int red_blocks) {
    constexpr int B = 2 * PAIRS, NCM = PAIRS / 2;
    __shared__ double sv[16];
    if (threadIdx.x < B) sv[threadIdx.x] = a.val[threadIdx.x];
    __syncthreads();
    const bool tip_mode = a.pnode == nullptr;
    double acc[2] = {0.0, 0.0};
    for (int64_t p = (int64_t)blockIdx.x * RED_THREADS + threadIdx.x; p < a.ntot; p += (int64_t)gridDim.x * RED_THREADS) {
        double d[B], w[B];
        t32_load<PAIRS>(a.pdad, p, d);
        if (tip_mode) {  // tip_val[state][c][i] = val0 * tip_partial_lh (:2757-2785)
            const double2 *tv = reinterpret_cast<const double2 *>(a.tip_val + (size_t)a.dad_states[p] * B);
#pragma unroll
            for (int j = 0; j < PAIRS; j++) {
                const double2 u = __ldg(tv + j);
                w[2 * j] = u.x;
                w[2 * j + 1] = u.y;
            }
        } else {
            t32_load<PAIRS>(a.pnode, p, w);
#pragma unroll
            for (int e = 0; e < B; e++) w[e] *= sv[e];
        }
        int min_scale, sc[NCM];
        if (a.safe) {
            min_scale = 0x7fffffff;
#pragma unroll
            for (int c = 0; c < NCM; c++) {
                sc[c] = (a.sdad[p * NCM + c] + (tip_mode ? 0 : a.snode[p * NCM + c])) & 0xff;
                min_scale = min(min_scale, sc[c]);
            }
        } else {
            min_scale = a.sdad[p] + (tip_mode ? 0 : a.snode[p]);
        }
        double lh = 0.0;
#pragma unroll
        for (int c = 0; c < NCM; c++) {
            const double s0 = fma(w[4 * c + 2], d[4 * c + 2], w[4 * c] * d[4 * c]);
            const double s1 = fma(w[4 * c + 3], d[4 * c + 3], w[4 * c + 1] * d[4 * c + 1]);
            double lc = s0 + s1;
            if (a.safe) {  // :2886-2893, 3018-3025
                if (sc[c] == min_scale + 1) lc *= B200_SCALING_THRESHOLD;
                else if (sc[c] > min_scale + 1) lc = 0.0;
            }
            a.pattern_lh_cat[(size_t)p * NCM + c] = lc;
            lh += lc;
        }
        lh = fabs(lh) + a.ptn_invar[p];  // :2907, 3038
        if (p < a.nptn) {
            lh = log(lh) + (double)min_scale * B200_LOG_SCALING_THRESHOLD;  // :2909, 3041
            a.pattern_lh[p] = lh;
            acc[0] = fma(lh, a.ptn_freq[p], acc[0]);
        } else {  // +ASC unobserved pattern (:2912-2930)
            if (min_scale != 0) lh *= B200_SCALING_THRESHOLD;
            a.pattern_lh[p] = lh;
            acc[1] += lh;
        }
    }
    block_reduce_store<2>(acc, a.red_partial, red_blocks);
}
#define T32_DISPATCH(fn, ...)                                  \
    switch (a.block) {                                         \
        case 4: fn<2>(__VA_ARGS__); break;                     \
        case 8: fn<4>(__VA_ARGS__); break;                     \
        case 12: fn<6>(__VA_ARGS__); break;                    \
        default: fn<8>(__VA_ARGS__); break;                    \
    }

// ---- theta ---------------------------------------------------------------
// theta[p][c][i] = partial_dad * partial_node (or * tip_partial_lh[state][m][i] when dad is a leaf);
// buffer_scale_all[p] = LOG_SCALING_THRESHOLD * (NORM: summed scale; SAFE: min over categories, with
// categories scaled once more multiplied by 2^-256 and those scaled more set to 0) (:2142-2175, 2190-2228)
__global__ void __launch_bounds__(RED_THREADS) theta_kernel(BranchArgs a, BatchLaunch bl) {
    rebind_branch(a, bl);
    if (a.tile == 32) {
        T32_DISPATCH(theta_t32, a)
        return;
    }
    const int G = a.G, g = threadIdx.x & (G - 1);
    const int64_t groups_per_grid = (int64_t)gridDim.x * (RED_THREADS / G);
    const int n = a.n, ncm = a.ncm, block = a.block;
    for (int64_t pb = (int64_t)blockIdx.x * (RED_THREADS / G); pb < a.ntot; pb += groups_per_grid) {
        int64_t p = pb + threadIdx.x / G;  // loop bound is CTA-uniform so every lane reaches the shuffles
        const bool valid = p < a.ntot;
        if (!valid) p = a.ntot - 1;
        const double *pd = a.pdad;
        double *th = a.theta;
        int min_scale = 0;
        if (a.safe) {
            int ms = 0x7fffffff;
            for (int c = g; c < ncm; c += G) {
                int s = a.sdad[p * ncm + c] + (a.pnode ? a.snode[p * ncm + c] : 0);
                ms = min(ms, s & 0xff);
            }
            min_scale = group_min(ms, G);
        } else {
            min_scale = a.sdad[p] + (a.pnode ? a.snode[p] : 0);
        }
        const double *other;
        int tip_mode = a.pnode == nullptr;
        if (tip_mode)
            other = a.site_eval ? a.site_tip + ((size_t)a.dad_taxon * a.nptn + p) * n
                                : a.tip_partial_lh + (size_t)a.dad_states[p] * (n * a.nmix);
        else other = a.pnode;
        for (int j = g; j < block; j += G) {
            int c = j / n;
            double o = tip_mode ? other[(a.site_eval ? 0 : (c / a.denom) * n) + (j - c * n)] : other[E(j)];
            double t = o * pd[E(j)];
            if (a.safe) {
                int s = (a.sdad[p * ncm + c] + (a.pnode ? a.snode[p * ncm + c] : 0)) & 0xff;
                if (s == min_scale + 1) t *= B200_SCALING_THRESHOLD;
                else if (s > min_scale + 1) t = 0.0;
            }
            if (valid) th[E(j)] = t;
        }
        if (g == 0 && valid) a.buffer_scale_all[p] = (double)min_scale * B200_LOG_SCALING_THRESHOLD;
    }
}

// ---- df, ddf -------------------------------------------------------------
__global__ void __launch_bounds__(RED_THREADS) derv_kernel(BranchArgs a, int red_blocks, BatchLaunch bl) {
    rebind_branch(a, bl);
    if (a.tile == 32) {
        T32_DISPATCH(derv_t32, a, red_blocks)
        return;
    }
    const int G = a.G, g = threadIdx.x & (G - 1);
    const int64_t groups_per_grid = (int64_t)gridDim.x * (RED_THREADS / G);
    const int block = a.block;
    const double *v0 = a.val, *v1 = a.val + block, *v2 = a.val + 2 * block;
    double acc[5] = {0.0, 0.0, 0.0, 0.0, 0.0};  // df, ddf | +ASC: prob_const, df_const, ddf_const
    if (!a.site_eval && block >= 512 && G == 32 && a.tile == 1) {
        // mixtures (block = 4800 for 60 classes x 4 rates): the three coefficient tables are 115 kB, so a warp that
        // owns one pattern re-reads them from L2 for every pattern -- three table loads per theta load.  A warp takes
        // four patterns at a time instead and uses each coefficient four times.
        constexpr int PP = 4;
        const int64_t warps_per_grid = (int64_t)gridDim.x * (RED_THREADS / 32);
        for (int64_t pb = ((int64_t)blockIdx.x * (RED_THREADS / 32) + (threadIdx.x >> 5)) * PP; pb < a.ntot;
             pb += warps_per_grid * PP) {
            const double *th[PP];
            double lh[PP], df[PP], ddf[PP];
#pragma unroll
            for (int q = 0; q < PP; q++) {
                th[q] = a.theta + (size_t)min(pb + q, a.ntot - 1) * block;
                lh[q] = df[q] = ddf[q] = 0.0;
            }
            for (int j = g; j < block; j += 32) {
                const double w0 = v0[j], w1 = v1[j], w2 = v2[j];
#pragma unroll
                for (int q = 0; q < PP; q++) {
                    const double t = __ldcs(th[q] + j);
                    lh[q] = fma(w0, t, lh[q]);
                    df[q] = fma(w1, t, df[q]);
                    ddf[q] = fma(w2, t, ddf[q]);
                }
            }
#pragma unroll
            for (int q = 0; q < PP; q++) {
                const double l = group_sum(lh[q], 32), d1 = group_sum(df[q], 32), d2 = group_sum(ddf[q], 32);
                if (g == 0 && pb + q < a.ntot) derv_accumulate(a, pb + q, l, d1, d2, acc);
            }
        }
        if (a.ntot > a.nptn) block_reduce_store<5>(acc, a.red_partial, red_blocks);
        else {
            double acc2[2] = {acc[0], acc[1]};
            block_reduce_store<2>(acc2, a.red_partial, red_blocks);
        }
        return;
    }
    for (int64_t pb = (int64_t)blockIdx.x * (RED_THREADS / G); pb < a.ntot; pb += groups_per_grid) {
        int64_t p = pb + threadIdx.x / G;  // loop bound is CTA-uniform so every lane reaches the shuffles
        const bool valid = p < a.ntot;
        if (!valid) p = a.ntot - 1;
        const double *th = a.theta;
        double lh = 0.0, df = 0.0, ddf = 0.0;
        if (a.site_eval) {
            for (int j = g; j < block; j += G) {
                double t = th[E(j)], w0, w1, w2;
                branch_coef(a, p, j, w0, w1, w2);
                lh = fma(w0, t, lh);
                df = fma(w1, t, df);
                ddf = fma(w2, t, ddf);
            }
        } else {
            for (int j = g; j < block; j += G) {
                double t = th[E(j)];
                lh = fma(v0[j], t, lh);
                df = fma(v1[j], t, df);
                ddf = fma(v2[j], t, ddf);
            }
        }
        lh = group_sum(lh, G);
        df = group_sum(df, G);
        ddf = group_sum(ddf, G);
        if (g == 0 && valid) derv_accumulate(a, p, lh, df, ddf, acc);
    }
    if (a.ntot > a.nptn) block_reduce_store<5>(acc, a.red_partial, red_blocks);
    else {
        double acc2[2] = {acc[0], acc[1]};
        block_reduce_store<2>(acc2, a.red_partial, red_blocks);
    }
}

// +ASC (Lewis) corrections applied to the reduced sums (:2620-2628, 3160-3182).  result: [0] lnL, [1] df, [2] ddf,
// [8] prob_const, [9] df_const, [10] ddf_const.  mode 0: lnL (+ per-pattern lnL), mode 1: derivatives.
__global__ void asc_fix_kernel(double *result, double *pattern_lh, int64_t nptn, double nsite, int mode) {
    if (mode == 1) {
        if (blockIdx.x == 0 && threadIdx.x == 0 && isfinite(result[1])) {
            const double prob = 1.0 - result[8];
            const double df_frac = result[9] / prob, ddf_frac = result[10] / prob;
            result[1] += nsite * df_frac;
            result[2] += nsite * (ddf_frac + df_frac * df_frac);
        }
        return;
    }
    const double pc = log(1.0 - result[8]);
    for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < nptn; p += (int64_t)gridDim.x * blockDim.x)
        pattern_lh[p] -= pc;
    if (blockIdx.x == 0 && threadIdx.x == 0) result[0] -= nsite * pc;
}

// ---- Newton-Raphson on the device ---------------------------------------------
// Optimization::minimizeNewton (utils/optimization.cpp:422-499) with computeFuncDerv = the derv pass above, as ONE
// cooperative launch: every iteration all CTAs stream theta once, leave their partial sums in a parity buffer,
// cross a grid barrier, and then EVERY CTA reduces the partials in the same fixed order and advances the same
// safeguarded-Newton state machine, so all CTAs agree on the next branch length without a second barrier or a
// host round trip.  val0/1/2 for the current length are rebuilt per CTA in shared memory.
struct NewtonArgs {
    DevModel m;
    double x1, x2, xacc, guess;
    int max_steps, ncat, safe_zero;
    double *out;  // [0] optx, [1] d2l, [2] derivative evaluations, [3] status (0 ok, 1 non-finite derivative)
    const double *guess_dev;  // non-null: the starting length is read from the device (device-resident lengths, sweep.cu)
    double *len_out, *aux_out;  // non-null: the optimum goes into the length table, evaluations and status beside it
    double *red;  // [2 parities][2][red_blocks]
};

__global__ void __launch_bounds__(RED_THREADS) newton_kernel(BranchArgs a, NewtonArgs nr, int red_blocks, int fuse_theta) {
    cooperative_groups::grid_group grid = cooperative_groups::this_grid();
    extern __shared__ double sval[];  // [3][block]
    __shared__ double sh_red[RED_THREADS];
    __shared__ double sh_state[4];     // rts, stop flag
    const int G = a.G, g = threadIdx.x & (G - 1);
    const int64_t groups_per_grid = (int64_t)gridDim.x * (RED_THREADS / G);
    const int block = a.block, n = a.n;
    const double *v0 = sval, *v1 = sval + block, *v2 = sval + 2 * block;
    // Newton state, replicated in thread 0 of every CTA
    const double guess0 = nr.guess_dev ? *nr.guess_dev : nr.guess;  // read by every CTA before the first grid barrier
    double rts = fmin(fmax(guess0, nr.x1), nr.x2), xl = 0.0, xh = 0.0, dx = 0.0, rts_old = rts, d2l = 0.0, result = rts;
    int j = 0, evals = 0, status = 0;
    bool started = false;
    for (int iter = 0;; iter++) {
        // val tables for the current length (val_kernel)
        for (int idx = threadIdx.x; idx < block; idx += RED_THREADS) {
            const int c = idx / n, i = idx - c * n, mm = c / a.denom, k = c % nr.ncat;
            const double prop = nr.m.props[k] * nr.m.mixw[mm];
            const double cof = nr.m.eval[mm * n + i] * nr.m.rates[k];
            const double v = exp(cof * rts) * prop, w = cof * v;
            sval[idx] = v;
            sval[block + idx] = w;
            sval[2 * block + idx] = cof * w;
        }
        __syncthreads();
        double acc[2] = {0.0, 0.0};
        if (a.tile == 32) {  // thread per pattern, 128-bit loads (see the T32 fast paths above)
            for (int64_t p = (int64_t)blockIdx.x * RED_THREADS + threadIdx.x; p < a.nptn; p += (int64_t)gridDim.x * RED_THREADS) {
                double lh, df, ddf;
                if (fuse_theta && iter == 0) {  // theta does not exist yet: build it on the way (b200lk_optimize_branch)
                    const bool tip_mode = a.pnode == nullptr;
                    switch (block) {
                        case 4: theta_triple_t32<2>(a, p, tip_mode, sval, lh, df, ddf); break;
                        case 8: theta_triple_t32<4>(a, p, tip_mode, sval, lh, df, ddf); break;
                        case 12: theta_triple_t32<6>(a, p, tip_mode, sval, lh, df, ddf); break;
                        default: theta_triple_t32<8>(a, p, tip_mode, sval, lh, df, ddf); break;
                    }
                } else
                switch (block) {
                    case 4: t32_triple<2>(a.theta, p, sval, lh, df, ddf); break;
                    case 8: t32_triple<4>(a.theta, p, sval, lh, df, ddf); break;
                    case 12: t32_triple<6>(a.theta, p, sval, lh, df, ddf); break;
                    default: t32_triple<8>(a.theta, p, sval, lh, df, ddf); break;
                }
                lh = fabs(lh) + a.ptn_invar[p];
                const double inv = 1.0 / lh;
                const double df_frac = df * inv, ddf_frac = ddf * inv, freq = a.ptn_freq[p];
                const double tmp1 = df_frac * freq, tmp2 = ddf_frac * freq;
                acc[0] += tmp1;
                acc[1] += fma(-tmp1, df_frac, tmp2);
            }
        } else
        for (int64_t pb = (int64_t)blockIdx.x * (RED_THREADS / G); pb < a.nptn; pb += groups_per_grid) {
            int64_t p = pb + threadIdx.x / G;
            const bool valid = p < a.nptn;
            if (!valid) p = a.nptn - 1;
            const double *th = a.theta;
            double lh = 0.0, df = 0.0, ddf = 0.0;
            for (int e = g; e < block; e += G) {
                const double t = th[E(e)];
                lh = fma(v0[e], t, lh);
                df = fma(v1[e], t, df);
                ddf = fma(v2[e], t, ddf);
            }
            lh = group_sum(lh, G);
            df = group_sum(df, G);
            ddf = group_sum(ddf, G);
            if (g == 0 && valid) {
                lh = fabs(lh) + a.ptn_invar[p];
                const double inv = 1.0 / lh;
                const double df_frac = df * inv, ddf_frac = ddf * inv, freq = a.ptn_freq[p];
                const double tmp1 = df_frac * freq, tmp2 = ddf_frac * freq;
                acc[0] += tmp1;
                acc[1] += fma(-tmp1, df_frac, tmp2);
            }
        }
        double *red = nr.red + (size_t)(iter & 1) * 2 * red_blocks;
        block_reduce_store<2>(acc, red, red_blocks);
        __threadfence();
        grid.sync();
        // every CTA: the fixed-order final reduction of final_reduce_kernel, on its own
        double sums[2];
        for (int k = 0; k < 2; k++) {
            double v = 0.0;
            for (int b = threadIdx.x; b < (int)gridDim.x; b += RED_THREADS) v += __ldcg(red + (size_t)k * red_blocks + b);
            sh_red[threadIdx.x] = v;
            __syncthreads();
            for (int s = RED_THREADS / 2; s > 0; s >>= 1) {
                if (threadIdx.x < s) sh_red[threadIdx.x] += sh_red[threadIdx.x + s];
                __syncthreads();
            }
            sums[k] = sh_red[0];
            __syncthreads();
        }
        if (threadIdx.x == 0) {
            double f = -sums[0], df = -sums[1];  // computeFuncDerv negates (tree/phylotree.cpp:2520-2527)
            evals++;
            bool stop = false;
            if (!isfinite(f) || !isfinite(df)) {
                if (nr.safe_zero) { f = 0.0; df = 0.0; }  // SAFE_LH: df = ddf = 0 and carry on (:2629-2632)
                else { status = 1; stop = true; result = rts; }
            }
            if (!stop) {
                if (!started) {
                    started = true;
                    d2l = df;
                    if (df >= 0.0 && fabs(f) < nr.xacc) { stop = true; result = rts; }
                    else {
                        if (f < 0.0) { xl = rts; xh = nr.x2; } else { xh = rts; xl = nr.x1; }
                        dx = fabs(xh - xl);
                        j = 1;
                    }
                } else {
                    if (df > 0.0 && fabs(f) < nr.xacc) { d2l = df; stop = true; result = rts; }
                    else {
                        if (f < 0.0) xl = rts;
                        else if (f > 0.0) xh = rts;
                        j++;
                    }
                }
            }
            if (!stop) {  // one step of the safeguarded iteration
                rts_old = rts;
                if (df <= 0.0 || (((rts - xh) * df - f) * ((rts - xl) * df - f) >= 0.0)) {
                    dx = 0.5 * (xh - xl);
                    rts = xl + dx;
                    d2l = df;
                    if (xl == rts) { stop = true; result = rts; }
                } else {
                    dx = f / df;
                    const double temp = rts;
                    rts -= dx;
                    d2l = df;
                    if (temp == rts) { stop = true; result = rts; }
                }
                if (!stop && (fabs(dx) < nr.xacc || j == nr.max_steps)) { stop = true; result = rts_old; }
            }
            sh_state[0] = rts;
            sh_state[1] = stop ? 1.0 : 0.0;
            if (stop && blockIdx.x == 0) {
                nr.out[0] = result;
                nr.out[1] = d2l;
                nr.out[2] = (double)evals;
                nr.out[3] = (double)status;
                if (nr.len_out) {  // written after at least one grid barrier: no CTA still needs the old value
                    *nr.len_out = result;
                    nr.aux_out[0] = (double)evals;
                    nr.aux_out[1] = (double)status;
                }
            }
        }
        __syncthreads();
        rts = sh_state[0];
        const bool stop = sh_state[1] != 0.0;
        __syncthreads();
        if (stop) return;
    }
}

// ---- lnL from theta --------------------------------------------------------
__global__ void __launch_bounds__(RED_THREADS) frombuf_kernel(BranchArgs a, int red_blocks, BatchLaunch bl) {
    rebind_branch(a, bl);
    if (a.tile == 32) {
        T32_DISPATCH(frombuf_t32, a, red_blocks)
        return;
    }
    const int G = a.G, g = threadIdx.x & (G - 1);
    const int64_t groups_per_grid = (int64_t)gridDim.x * (RED_THREADS / G);
    const int block = a.block;
    double acc[2] = {0.0, 0.0};  // lnL | +ASC prob_const
    for (int64_t pb = (int64_t)blockIdx.x * (RED_THREADS / G); pb < a.ntot; pb += groups_per_grid) {
        int64_t p = pb + threadIdx.x / G;  // loop bound is CTA-uniform so every lane reaches the shuffles
        const bool valid = p < a.ntot;
        if (!valid) p = a.ntot - 1;
        const double *th = a.theta;
        double lh = 0.0;
        if (a.site_eval) {
            for (int j = g; j < block; j += G) {
                double w0, w1, w2;
                branch_coef(a, p, j, w0, w1, w2);
                lh = fma(w0, th[E(j)], lh);
            }
        } else {
            for (int j = g; j < block; j += G) lh = fma(a.val[j], th[E(j)], lh);
        }
        lh = group_sum(lh, G);
        if (g == 0 && valid) {
            lh = fabs(lh) + a.ptn_invar[p];               // :3295
            if (p < a.nptn) {
                lh = log(fabs(lh)) + a.buffer_scale_all[p];   // :3300
                a.pattern_lh[p] = lh;
                acc[0] = fma(lh, a.ptn_freq[p], acc[0]);
            } else {  // +ASC unobserved pattern
                if (a.buffer_scale_all[p] != 0.0) lh *= B200_SCALING_THRESHOLD;
                a.pattern_lh[p] = lh;
                acc[1] += lh;
            }
        }
    }
    block_reduce_store<2>(acc, a.red_partial, red_blocks);
}

// ---- lnL at a branch -------------------------------------------------------
// lane g of a group owns categories g, g+G, ...; lh_cat[c] = sum_i val[c][i]*node[c][i]*dad[c][i]
// (or tip_val[state][c][i]*dad[c][i] when dad is a leaf).
__global__ void __launch_bounds__(RED_THREADS) lnl_kernel(BranchArgs a, int red_blocks, BatchLaunch bl) {
    rebind_branch(a, bl);
    if (a.tile == 32) {
        T32_DISPATCH(lnl_t32, a, red_blocks)
        return;
    }
    const int G = a.G, g = threadIdx.x & (G - 1);
    const int64_t groups_per_grid = (int64_t)gridDim.x * (RED_THREADS / G);
    const int n = a.n, ncm = a.ncm, block = a.block;
    const bool tip_mode = a.pnode == nullptr;
    double acc[2] = {0.0, 0.0};  // lnL | +ASC prob_const
    for (int64_t pb = (int64_t)blockIdx.x * (RED_THREADS / G); pb < a.ntot; pb += groups_per_grid) {
        int64_t p = pb + threadIdx.x / G;  // loop bound is CTA-uniform so every lane reaches the shuffles
        const bool valid = p < a.ntot;
        if (!valid) p = a.ntot - 1;
        const double *pd = a.pdad;
        const double *w = (tip_mode && !a.site_eval) ? a.tip_val + (size_t)a.dad_states[p] * block : a.val;
        const double *pn = a.pnode;
        double lh = 0.0;
        int min_scale;
        if (a.safe) {
            int ms = 0x7fffffff;
            for (int c = g; c < ncm; c += G) {
                int s = (a.sdad[p * ncm + c] + (tip_mode ? 0 : a.snode[p * ncm + c])) & 0xff;
                ms = min(ms, s);
            }
            min_scale = group_min(ms, G);
        } else {
            min_scale = a.sdad[p] + (tip_mode ? 0 : a.snode[p]);
        }
        for (int c = g; c < ncm; c += G) {
            double s0 = 0.0, s1 = 0.0;
            const double *wd = w + c * n;
            const int cb = c * n;
            if (a.site_eval) {
                const double *tipv = tip_mode ? a.site_tip + ((size_t)a.dad_taxon * a.nptn + p) * n : nullptr;
                for (int i = 0; i < n; i++) {
                    double w0, w1, w2;
                    branch_coef(a, p, cb + i, w0, w1, w2);
                    s0 = fma(w0 * (tip_mode ? tipv[i] : pn[E(cb + i)]), pd[E(cb + i)], s0);
                }
            } else if (tip_mode) {
                for (int i = 0; i + 1 < n; i += 2) {
                    s0 = fma(wd[i], pd[E(cb + i)], s0);
                    s1 = fma(wd[i + 1], pd[E(cb + i + 1)], s1);
                }
                if (n & 1) s0 = fma(wd[n - 1], pd[E(cb + n - 1)], s0);
            } else {
                for (int i = 0; i + 1 < n; i += 2) {
                    s0 = fma(wd[i] * pn[E(cb + i)], pd[E(cb + i)], s0);
                    s1 = fma(wd[i + 1] * pn[E(cb + i + 1)], pd[E(cb + i + 1)], s1);
                }
                if (n & 1) s0 = fma(wd[n - 1] * pn[E(cb + n - 1)], pd[E(cb + n - 1)], s0);
            }
            double lc = s0 + s1;
            if (a.safe) {  // :2886-2893, 3018-3025
                int s = (a.sdad[p * ncm + c] + (tip_mode ? 0 : a.snode[p * ncm + c])) & 0xff;
                if (s == min_scale + 1) lc *= B200_SCALING_THRESHOLD;
                else if (s > min_scale + 1) lc = 0.0;
            }
            if (valid) a.pattern_lh_cat[(size_t)p * ncm + c] = lc;
            lh += lc;
        }
        lh = group_sum(lh, G);
        if (g == 0 && valid) {
            lh = fabs(lh) + a.ptn_invar[p];                                        // :2907, 3038
            if (p < a.nptn) {
                lh = log(lh) + (double)min_scale * B200_LOG_SCALING_THRESHOLD;    // :2909, 3041
                a.pattern_lh[p] = lh;
                acc[0] = fma(lh, a.ptn_freq[p], acc[0]);
            } else {  // +ASC unobserved pattern (:2912-2930)
                if (min_scale != 0) lh *= B200_SCALING_THRESHOLD;
                a.pattern_lh[p] = lh;
                acc[1] += lh;
            }
        }
    }
    block_reduce_store<2>(acc, a.red_partial, red_blocks);
}

// repair pass for non-finite per-pattern lnL (:3085-3094, 3338-3347)
__global__ void __launch_bounds__(RED_THREADS) repair_kernel(double *pattern_lh, const double *ptn_freq, int64_t nptn,
                                                            double *red_partial, int red_blocks) {
    double acc[1] = {0.0};
    for (int64_t p = (int64_t)blockIdx.x * RED_THREADS + threadIdx.x; p < nptn; p += (int64_t)gridDim.x * RED_THREADS) {
        double v = pattern_lh[p];
        if (!isfinite(v)) {
            v = B200_LOG_SCALING_THRESHOLD * 4;
            pattern_lh[p] = v;
        }
        acc[0] += v * ptn_freq[p];
    }
    block_reduce_store<1>(acc, red_partial, red_blocks);
}

// ---- host-side launchers -----------------------------------------------------
static int pick_group(int count) {  // largest power of two <= min(32, count)
    int G = 1;
    while (G * 2 <= count && G * 2 <= 32) G *= 2;
    return G;
}

static int fill_args(b200lk_ctx *c, BranchArgs &a, b200lk_end dad, b200lk_end node) {
    // a slot the 4-state traversal left virtual (tip-tip vector, s4.cu) is gathered before a branch kernel reads it
    if (int rc = s4_realize_slot(c, node.id)) return rc;
    if (dad.kind != B200LK_CHILD_LEAF)
        if (int rc = s4_realize_slot(c, dad.id)) return rc;
    memset(&a, 0, sizeof(a));
    a.pdad = slot_partial(c, node.id);
    a.sdad = slot_scale(c, node.id);
    a.slot_pdad = node.id;
    a.slot_pnode = -1;
    a.dad_taxon = -1;
    if (dad.kind == B200LK_CHILD_LEAF) {
        a.dad_states = c->states + (size_t)dad.id * c->ntot_pad;
        a.dad_taxon = dad.id;
    } else {
        a.pnode = slot_partial(c, dad.id);
        a.snode = slot_scale(c, dad.id);
        a.slot_pnode = dad.id;
    }
    return B200LK_OK;
}

static void fill_common(b200lk_ctx *c, BranchArgs &a) {
    a.tip_partial_lh = c->m.tip_partial_lh;
    a.val = c->val;
    a.tip_val = c->tip_val;
    a.ptn_freq = c->ptn_freq;
    a.ptn_invar = c->ptn_invar;
    a.theta = c->theta_all;
    a.buffer_scale_all = c->buffer_scale_all;
    a.pattern_lh = c->pattern_lh;
    a.pattern_lh_cat = c->pattern_lh_cat;
    a.red_partial = c->red_partial;
    a.nptn = c->d.nptn;
    a.ntot = c->ntot;
    a.n = c->d.nstates;
    a.ncm = c->ncat_mix;
    a.block = c->block;
    a.denom = c->d.fused_mix_rate ? 1 : c->d.ncat;
    a.nmix = c->d.nmix;
    a.safe = c->d.safe_numeric;
    a.tile = c->tile;
    a.site_eval = c->site ? c->s_eval : nullptr;
    a.site_tip = c->s_tip;
    a.rates = c->m.rates;
    a.props = c->m.props;
    a.len = 0.0;
    a.const_df = c->asc_holder ? c->const_df : nullptr;
    a.const_ddf = c->asc_holder ? c->const_ddf : nullptr;
}

static BatchLaunch batch_of(const b200lk_ctx *c) {
    BatchLaunch bl;
    memset(&bl, 0, sizeof(bl));
    if (c->batch) bl = *c->batch;
    return bl;
}
static dim3 grid2(const b200lk_ctx *c, int gx) { return dim3(gx, c->batch ? c->batch->n : 1); }

static int grid_for(const b200lk_ctx *c, int G, int64_t npat) {
    if (c->batch) npat = c->batch->max_ntot;
    int64_t per_block = RED_THREADS / G;
    int64_t need = (npat + per_block - 1) / per_block;
    if (need < 1) need = 1;
    return (int)(need < c->red_blocks ? need : c->red_blocks);
}

static int launch_holder_fix(b200lk_ctx *c, int mode);
static int launch_asc_fix(b200lk_ctx *c, int mode) {
    if (c->d.n_unobserved == 0) return B200LK_OK;
    if (c->asc_holder) return launch_holder_fix(c, mode);
    const int grid = mode == 1 ? 1 : (int)std::min<int64_t>(c->red_blocks, (c->d.nptn + 255) / 256);
    asc_fix_kernel<<<grid, 256, 0, c->stream>>>(c->d_result, c->pattern_lh, c->d.nptn, c->nsite, mode);
    CUDA_TRY(c, cudaGetLastError());
    c->n_launches++;
    return B200LK_OK;
}

static int launch_val(b200lk_ctx *c, double length, int want_tip) {
    if (c->site) return B200LK_OK;  // per-pattern coefficients are formed inside the kernels
    const int denom = c->d.fused_mix_rate ? 1 : c->d.ncat;
    val_kernel<<<grid2(c, (c->block + 255) / 256), 256, 0, c->stream>>>(c->m, length, c->cur_branch_mixlen, c->val, c->tip_val, c->d.nstates,
                                                                        c->d.ncat, c->ncat_mix, denom, c->d.nmix,
                                                                        c->d.n_state_codes, want_tip, batch_of(c));
    CUDA_TRY(c, cudaGetLastError());
    c->n_launches++;
    return B200LK_OK;
}

int launch_branch_lnl(b200lk_ctx *c, b200lk_end dad, b200lk_end node, double length) {
    if (c->nonrev) return nr_launch_branch_lnl(c, dad, node, length);
    int rc = launch_val(c, length, dad.kind == B200LK_CHILD_LEAF);
    if (rc) return rc;
    BranchArgs a;
    if (int rc = fill_args(c, a, dad, node)) return rc;
    fill_common(c, a);
    a.len = length;
    a.G = pick_group(c->ncat_mix);
    int grid = grid_for(c, a.G, c->ntot);
    lnl_kernel<<<grid2(c, grid), RED_THREADS, 0, c->stream>>>(a, grid, batch_of(c));
    CUDA_TRY(c, cudaGetLastError());
    final_reduce_kernel<<<grid2(c, 1), RED_THREADS, 0, c->stream>>>(c->red_partial, grid, 2, c->d_result, 0, 1, 8, batch_of(c));
    CUDA_TRY(c, cudaGetLastError());
    c->n_launches += 2;
    if (c->robust_keep < 1.0 || c->robust_median) {  // :3096-3135; "+ASC not supported with robust phylo"
        if (c->d.n_unobserved > 0 || c->batch) return B200LK_ERR_UNSUPPORTED;
        return launch_robust(c);
    }
    return launch_asc_fix(c, 0);
}

int launch_repair(b200lk_ctx *c) {
    int grid = grid_for(c, 1, c->d.nptn);
    repair_kernel<<<grid, RED_THREADS, 0, c->stream>>>(c->pattern_lh, c->ptn_freq, c->d.nptn, c->red_partial, grid);
    CUDA_TRY(c, cudaGetLastError());
    final_reduce_kernel<<<1, RED_THREADS, 0, c->stream>>>(c->red_partial, grid, 1, c->d_result, 0, 8, 0, batch_of(c));
    CUDA_TRY(c, cudaGetLastError());
    c->n_launches += 2;
    return B200LK_OK;
}

int launch_theta(b200lk_ctx *c, b200lk_end dad, b200lk_end node) {
    if (c->nonrev) return nr_record_ends(c, dad, node);
    BranchArgs a;
    if (int rc = fill_args(c, a, dad, node)) return rc;
    fill_common(c, a);
    a.G = pick_group(c->block);
    int grid = grid_for(c, a.G, c->ntot);
    theta_kernel<<<grid2(c, grid), RED_THREADS, 0, c->stream>>>(a, batch_of(c));
    CUDA_TRY(c, cudaGetLastError());
    c->n_launches++;
    return B200LK_OK;
}

int launch_derv(b200lk_ctx *c, double length) {
    if (c->nonrev) return nr_launch_derv(c, length);
    int rc = launch_val(c, length, 0);
    if (rc) return rc;
    BranchArgs a;
    memset(&a, 0, sizeof(a));
    fill_common(c, a);
    a.len = length;
    a.G = pick_group(c->block);
    int grid = grid_for(c, a.G, c->ntot);
    derv_kernel<<<grid2(c, grid), RED_THREADS, 0, c->stream>>>(a, grid, batch_of(c));
    CUDA_TRY(c, cudaGetLastError());
    final_reduce_kernel<<<grid2(c, 1), RED_THREADS, 0, c->stream>>>(c->red_partial, grid, c->d.n_unobserved ? 5 : 2, c->d_result, 1, 2, 8, batch_of(c));
    CUDA_TRY(c, cudaGetLastError());
    c->n_launches += 2;
    return launch_asc_fix(c, 1);
}

int launch_from_buffer(b200lk_ctx *c, double length) {
    if (c->nonrev) return nr_launch_from_buffer(c, length);
    int rc = launch_val(c, length, 0);
    if (rc) return rc;
    BranchArgs a;
    memset(&a, 0, sizeof(a));
    fill_common(c, a);
    a.len = length;
    a.G = pick_group(c->block);
    int grid = grid_for(c, a.G, c->ntot);
    frombuf_kernel<<<grid2(c, grid), RED_THREADS, 0, c->stream>>>(a, grid, batch_of(c));
    CUDA_TRY(c, cudaGetLastError());
    final_reduce_kernel<<<grid2(c, 1), RED_THREADS, 0, c->stream>>>(c->red_partial, grid, 2, c->d_result, 0, 1, 8, batch_of(c));
    CUDA_TRY(c, cudaGetLastError());
    c->n_launches += 2;
    return launch_asc_fix(c, 0);
}

// b200lk_optimize_branch.  ends == nullptr: theta for the branch is already built.  Otherwise ends = {dad, node} and theta is
// built here: inside the first iteration on the T32 layout (one launch and one pass over theta less per branch), by
// theta_kernel first everywhere else.
bool newton_fuses_theta(const b200lk_ctx *c) {
    return c->tile == 32 && !c->site && !c->nonrev && !c->batch && c->d.n_unobserved == 0 && c->ntot == c->d.nptn;
}
int launch_newton(b200lk_ctx *c, const b200lk_newton *nr, const b200lk_end *ends) {
    int fuse_theta = ends && newton_fuses_theta(c);
    if (ends && !fuse_theta) {
        int rc = launch_theta(c, ends[0], ends[1]);
        if (rc) return rc;
    }
    if (c->site) return B200LK_ERR_UNSUPPORTED;  // per-pattern coefficients: the host iteration drives b200lk_derv
    BranchArgs a;
    if (fuse_theta) {
        if (int rc = fill_args(c, a, ends[0], ends[1])) return rc;
    } else {
        memset(&a, 0, sizeof(a));
    }
    fill_common(c, a);
    a.G = pick_group(c->block);
    NewtonArgs na;
    na.m = c->m;
    na.x1 = nr->min_len;
    na.x2 = nr->max_len;
    na.xacc = nr->xacc;
    na.guess = nr->guess;
    na.max_steps = nr->max_steps;
    na.ncat = c->d.ncat;
    na.safe_zero = c->d.safe_numeric;
    na.out = c->d_result + 4;
    const int slot = c->cur_newton_slot;
    na.guess_dev = slot >= 0 ? c->d_lens + slot : nullptr;
    na.len_out = slot >= 0 ? c->d_lens + slot : nullptr;
    na.aux_out = slot >= 0 ? c->d_lens_aux + 2 * (size_t)slot : nullptr;
    na.red = c->red_partial;  // [2 parities][2][red_blocks] of its 8 * red_blocks doubles
    const size_t smem = sizeof(double) * 3 * c->block;
    if (smem > 200 * 1024 || c->d.n_unobserved > 0) return B200LK_ERR_UNSUPPORTED;  // +ASC: host-driven iteration
    auto unsupported = [&]() {  // the caller falls back to the host-driven iteration, which expects theta to exist
        if (fuse_theta) {
            int rc = launch_theta(c, ends[0], ends[1]);
            if (rc) return rc;
        }
        return (int)B200LK_ERR_UNSUPPORTED;
    };
    if (smem > 48 * 1024)
        CUDA_TRY(c, cudaFuncSetAttribute(newton_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 0, sms = c->sms, coop = 0;
    CUDA_TRY(c, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, newton_kernel, RED_THREADS, smem));
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->device);
    cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, c->device);
    if (!coop || per_sm < 1) return unsupported();
    int grid = grid_for(c, a.G, c->d.nptn);
    if (grid > per_sm * sms) grid = per_sm * sms;
    int red_blocks = c->red_blocks;
    void *args[] = {&a, &na, &red_blocks, &fuse_theta};
    CUDA_TRY(c, cudaLaunchCooperativeKernel((void *)newton_kernel, dim3(grid), dim3(RED_THREADS), args, smem, c->stream));
    c->n_launches++;
    return B200LK_OK;
}

// ---- UFBoot RELL: rell[b] = sum_p (float)pattern_lh[p] * boot_samples[b][p] ------------------------------------
// One CTA per resample streams its row once (HBM-bound: 4 bytes per pattern and resample); _pattern_lh stays in L2.
__global__ void __launch_bounds__(RED_THREADS) rell_kernel(const float *__restrict__ samples, const double *__restrict__ pattern_lh,
                                                            int64_t nptn, double *__restrict__ rell) {
    const float *row = samples + (size_t)blockIdx.x * nptn;
    double acc = 0.0;
    for (int64_t p = threadIdx.x; p < nptn; p += RED_THREADS) acc += (double)((float)pattern_lh[p] * __ldg(row + p));
    __shared__ double sh[RED_THREADS];
    sh[threadIdx.x] = acc;
    __syncthreads();
    for (int s = RED_THREADS / 2; s > 0; s >>= 1) {
        if (threadIdx.x < s) sh[threadIdx.x] += sh[threadIdx.x + s];
        __syncthreads();
    }
    if (threadIdx.x == 0) rell[blockIdx.x] = sh[0];
}

int launch_rell(b200lk_ctx *c) {
    rell_kernel<<<c->nboot, RED_THREADS, 0, c->stream>>>(c->boot_samples, c->pattern_lh, c->d.nptn, c->d_rell);
    CUDA_TRY(c, cudaGetLastError());
    c->n_launches++;
    return B200LK_OK;
}

// ---- Holder's +ASC for missing data (ASC_VARIANT_MISSING) ---------------------------------------------------------
// Unobserved pattern x*nptn + p is the constant-x column with the gaps of observed pattern p; the kernels above left
// its likelihood in pattern_lh[nptn + ...] (and const_df / const_ddf).  Per observed pattern:
//   mode 0 (:3137-3159, 3349-3368)  pv = log(1 - sum_x L);  pattern_lh[p] -= pv;  lnL -= sum_p f_p pv
//   mode 1 (:2580-2619)             df += sum_p f_p df_frac;  ddf += sum_p f_p (ddf_frac + df_frac^2)
__global__ void __launch_bounds__(RED_THREADS) holder_kernel(double *pattern_lh, const double *const_df,
                                                            const double *const_ddf, const double *ptn_freq,
                                                            int64_t nptn, int n, int mode, double *red_partial,
                                                            int red_blocks) {
    double acc[2] = {0.0, 0.0};
    const double *const_lh = pattern_lh + nptn;
    for (int64_t p = (int64_t)blockIdx.x * RED_THREADS + threadIdx.x; p < nptn; p += (int64_t)gridDim.x * RED_THREADS) {
        double cl = const_lh[p];
        for (int s = 1; s < n; s++) cl += const_lh[(size_t)s * nptn + p];
        if (mode == 0) {
            const double pv = log(1.0 - cl);
            pattern_lh[p] -= pv;
            acc[0] += pv * ptn_freq[p];
        } else {
            double cd = const_df[p], cdd = const_ddf[p];
            for (int s = 1; s < n; s++) {
                cd += const_df[(size_t)s * nptn + p];
                cdd += const_ddf[(size_t)s * nptn + p];
            }
            const double pc = 1.0 - cl, df_frac = cd / pc, ddf_frac = cdd / pc;
            acc[0] += ptn_freq[p] * df_frac;
            acc[1] += ptn_freq[p] * (ddf_frac + df_frac * df_frac);
        }
    }
    block_reduce_store<2>(acc, red_partial, red_blocks);
}
// result[dst + k] += sign * sum_b red_partial[k][b], fixed order
__global__ void holder_final_kernel(const double *red_partial, int red_blocks, int K, double *result, int dst,
                                    double sign) {
    __shared__ double sh[RED_THREADS];
    for (int k = 0; k < K; k++) {
        double v = 0.0;
        for (int b = threadIdx.x; b < red_blocks; b += RED_THREADS) v += red_partial[(size_t)k * red_blocks + b];
        sh[threadIdx.x] = v;
        __syncthreads();
        for (int s = RED_THREADS / 2; s > 0; s >>= 1) {
            if (threadIdx.x < s) sh[threadIdx.x] += sh[threadIdx.x + s];
            __syncthreads();
        }
        if (threadIdx.x == 0) result[dst + k] += sign * sh[0];
        __syncthreads();
    }
}
static int launch_holder_fix(b200lk_ctx *c, int mode) {
    const int grid = grid_for(c, 1, c->d.nptn);
    holder_kernel<<<grid, RED_THREADS, 0, c->stream>>>(c->pattern_lh, c->const_df, c->const_ddf, c->ptn_freq, c->d.nptn,
                                                      c->d.nstates, mode, c->red_partial, grid);
    CUDA_TRY(c, cudaGetLastError());
    holder_final_kernel<<<1, RED_THREADS, 0, c->stream>>>(c->red_partial, grid, mode == 0 ? 1 : 2, c->d_result,
                                                         mode == 0 ? 0 : 1, mode == 0 ? -1.0 : 1.0);
    CUDA_TRY(c, cudaGetLastError());
    c->n_launches += 2;
    return B200LK_OK;
}

// ---- heterotachy: the two derivative forms of isMixlen() trees -------------------------------------------------------
#define MIX_MAXCAT 8
// computeLikelihoodDerv*SIMD with isMixlen() (:2365-2462): per pattern the class sums df_i, ddf_i (i = c % ncat) over
// the flat block, then  g_i += f df_i/L,  H_ii += f (ddf_i/L - (df_i/L)^2),  H_ik -= f df_i df_k / L^2,
// lnL += f (log L + buffer_scale).  Sums: [0] lnL, [1 + i] g_i, [1 + ncat + i*ncat + k] H_ik.
__global__ void __launch_bounds__(RED_THREADS) derv_mixlen_kernel(BranchArgs a, int ncat, double *red, int red_blocks) {
    const int G = a.G, g = threadIdx.x & (G - 1);
    const int64_t groups_per_grid = (int64_t)gridDim.x * (RED_THREADS / G);
    const int block = a.block, n = a.n;
    const double *v0 = a.val, *v1 = a.val + block, *v2 = a.val + 2 * block;
    const int K = 1 + ncat + ncat * ncat;
    double acc[1 + MIX_MAXCAT + MIX_MAXCAT * MIX_MAXCAT];
    for (int k = 0; k < K; k++) acc[k] = 0.0;
    for (int64_t pb = (int64_t)blockIdx.x * (RED_THREADS / G); pb < a.nptn; pb += groups_per_grid) {
        int64_t p = pb + threadIdx.x / G;
        const bool valid = p < a.nptn;
        if (!valid) p = a.nptn - 1;
        const double *th = a.theta;
        double lh = 0.0, d1[MIX_MAXCAT], d2[MIX_MAXCAT];
#pragma unroll
        for (int k = 0; k < MIX_MAXCAT; k++) d1[k] = d2[k] = 0.0;
        for (int j = g; j < block; j += G) {
            const double t = th[E(j)];
            const int i = (j / n) % ncat;
            lh = fma(v0[j], t, lh);
            const double x1 = v1[j] * t, x2 = v2[j] * t;
#pragma unroll
            for (int k = 0; k < MIX_MAXCAT; k++) {
                d1[k] += (k == i) ? x1 : 0.0;
                d2[k] += (k == i) ? x2 : 0.0;
            }
        }
        lh = group_sum(lh, G);
#pragma unroll
        for (int k = 0; k < MIX_MAXCAT; k++) {
            if (k < ncat) {  // ncat is launch-uniform
                d1[k] = group_sum(d1[k], G);
                d2[k] = group_sum(d2[k], G);
            }
        }
        if (g == 0 && valid) {
            lh = fabs(lh) + a.ptn_invar[p];
            const double freq = a.ptn_freq[p], inv = 1.0 / lh;
#pragma unroll
            for (int i = 0; i < MIX_MAXCAT; i++) {
                if (i < ncat) {
                    d1[i] *= inv;
                    d2[i] *= inv;
                    acc[1 + i] = fma(d1[i], freq, acc[1 + i]);
                }
            }
#pragma unroll
            for (int i = 0; i < MIX_MAXCAT; i++) {
                if (i >= ncat) continue;
#pragma unroll
                for (int k = 0; k < MIX_MAXCAT; k++) {
                    if (k >= ncat) continue;
                    const int idx = 1 + ncat + i * ncat + k;
                    if (k == i) acc[idx] += fma(-d1[i], d1[i], d2[i]) * freq;
                    else acc[idx] -= d1[i] * d1[k] * freq;
                }
            }
            acc[0] = fma(log(lh) + a.buffer_scale_all[p], freq, acc[0]);
        }
    }
    __shared__ double sh[RED_THREADS / 32];
    for (int k = 0; k < K; k++) {
        double v = acc[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
        __syncthreads();
        if (threadIdx.x == 0) {
            double s = 0.0;
            for (int w = 0; w < RED_THREADS / 32; w++) s += sh[w];
            red[(size_t)k * red_blocks + blockIdx.x] = s;
        }
        __syncthreads();
    }
}

int launch_derv_mixlen(b200lk_ctx *c, int K) {
    int rc = launch_val(c, 0.0, 0);  // c->cur_branch_mixlen carries the class lengths
    if (rc) return rc;
    BranchArgs a;
    memset(&a, 0, sizeof(a));
    fill_common(c, a);
    a.G = pick_group(c->block);
    const int grid = grid_for(c, a.G, c->d.nptn);
    derv_mixlen_kernel<<<grid, RED_THREADS, 0, c->stream>>>(a, c->d.ncat, c->mix_red, grid);
    CUDA_TRY(c, cudaGetLastError());
    BatchLaunch none;
    memset(&none, 0, sizeof(none));
    final_reduce_kernel<<<1, RED_THREADS, 0, c->stream>>>(c->mix_red, grid, K, c->mix_result, 0, K, 0, none);
    CUDA_TRY(c, cudaGetLastError());
    c->n_launches += 2;
    return B200LK_OK;
}

// computeLikelihoodDervMixlen*SIMD (:3397-3580): only the slice of class `cls` of theta enters, with
// val[i] = exp(eval[m][i] * len) * mixw[m], m = cls (fused) or 0 -- no rate, no category proportion (:3472-3486).
// Sums: df, ddf | +ASC (Lewis) prob_const, df_const, ddf_const (not rescaled here, :3536-3540).
__global__ void __launch_bounds__(RED_THREADS) derv_mixlen_class_kernel(BranchArgs a, DevModel m, int cls, int cur_mix,
                                                                       double len, int red_blocks) {
    __shared__ double sv[3][64];
    const int n = a.n, block = a.block;
    if (threadIdx.x < n) {
        const double cof = m.eval[cur_mix * n + threadIdx.x];
        const double v = exp(cof * len) * m.mixw[cur_mix];
        sv[0][threadIdx.x] = v;
        sv[1][threadIdx.x] = cof * v;
        sv[2][threadIdx.x] = cof * (cof * v);
    }
    __syncthreads();
    double acc[5] = {0.0, 0.0, 0.0, 0.0, 0.0};
    for (int64_t p = (int64_t)blockIdx.x * RED_THREADS + threadIdx.x; p < a.ntot; p += (int64_t)gridDim.x * RED_THREADS) {
        const double *th = a.theta;
        double lh = 0.0, df = 0.0, ddf = 0.0;
        for (int i = 0; i < n; i++) {
            const double t = th[E(cls * n + i)];
            lh = fma(sv[0][i], t, lh);
            df = fma(sv[1][i], t, df);
            ddf = fma(sv[2][i], t, ddf);
        }
        lh = fabs(lh) + a.ptn_invar[p];
        if (p < a.nptn) {
            const double inv = 1.0 / lh, freq = a.ptn_freq[p];
            df *= inv;
            ddf *= inv;
            acc[0] = fma(df, freq, acc[0]);
            acc[1] += fma(-df, df, ddf) * freq;
        } else {
            acc[2] += lh;
            acc[3] += df;
            acc[4] += ddf;
        }
    }
    if (a.ntot > a.nptn) block_reduce_store<5>(acc, a.red_partial, red_blocks);
    else {
        double acc2[2] = {acc[0], acc[1]};
        block_reduce_store<2>(acc2, a.red_partial, red_blocks);
    }
}

int launch_derv_mixlen_class(b200lk_ctx *c, int cls, double length) {
    BranchArgs a;
    memset(&a, 0, sizeof(a));
    fill_common(c, a);
    const int grid = grid_for(c, 1, c->ntot);
    derv_mixlen_class_kernel<<<grid, RED_THREADS, 0, c->stream>>>(a, c->m, cls, c->d.fused_mix_rate ? cls : 0, length, grid);
    CUDA_TRY(c, cudaGetLastError());
    BatchLaunch none;
    memset(&none, 0, sizeof(none));
    final_reduce_kernel<<<1, RED_THREADS, 0, c->stream>>>(c->red_partial, grid, c->d.n_unobserved ? 5 : 2, c->d_result, 1, 2, 8, none);
    CUDA_TRY(c, cudaGetLastError());
    c->n_launches += 2;
    return launch_asc_fix(c, 1);
}

// ---- fused tail of a whole evaluation on the T32 layout ------------------------------------------------------------------
// b200lk_evaluate* needs, at ONE branch and ONE length, the branch lnL (computeLikelihoodBranch), theta
// (computeLikelihoodBuffer) and df/ddf (computeLikelihoodDerv).  As separate launches that is val -> lnl -> reduce ->
// theta -> val -> derv -> reduce: seven launches that read the same two partial vectors three times.  Here one thread
// owns a pattern, loads both vectors once, writes theta / _pattern_lh / _pattern_lh_cat, and accumulates lnL, df and ddf
// with exactly the per-pattern arithmetic of lnl_t32, theta_t32_pattern and t32_triple_regs.  The branch coefficient
// table (val0/1/2, 3 * block <= 48 doubles) is rebuilt per CTA in shared memory with val_kernel's formula, and the CTA
// that finishes last adds the per-CTA partial sums in the fixed order of final_reduce_kernel (deterministic).
struct TailArgs {
    DevModel m;
    double len;
    int ncat;
    unsigned *counter;  // CTAs finished so far; the last one resets it
    double *result;     // [0] lnL, [1] df, [2] ddf
    PeerArgs peer;      // pattern shards in other processes: the last CTA exchanges the sums with them (peer.cuh)
};

template <int PAIRS>
__device__ __forceinline__ void eval_tail_t32(const BranchArgs &a, const TailArgs &ta) {
    constexpr int B = 2 * PAIRS, NCM = PAIRS / 2;
    __shared__ double sv[48];
    __shared__ unsigned s_ticket;
    if (threadIdx.x < B) {  // val_kernel's formula (:2342-2357)
        const int idx = threadIdx.x, c = idx >> 2, i = idx & 3, mm = c / a.denom, k = c % ta.ncat;
        const double prop = ta.m.props[k] * ta.m.mixw[mm];
        const double cof = ta.m.eval[mm * 4 + i] * ta.m.rates[k];
        const double v = exp(cof * ta.len) * prop, v1 = cof * v;
        sv[idx] = v;
        sv[B + idx] = v1;
        sv[2 * B + idx] = cof * v1;
    }
    __syncthreads();
    const bool tip_mode = a.pnode == nullptr;
    double acc[3] = {0.0, 0.0, 0.0};
    for (int64_t p = (int64_t)blockIdx.x * RED_THREADS + threadIdx.x; p < a.ntot; p += (int64_t)gridDim.x * RED_THREADS) {
        double d[B], o[B];
        t32_load<PAIRS>(a.pdad, p, d);
        if (tip_mode) {
            const double2 *tp = reinterpret_cast<const double2 *>(a.tip_partial_lh + (size_t)a.dad_states[p] * (4 * a.nmix));
#pragma unroll
            for (int c = 0; c < NCM; c++) {
                const double2 u0 = __ldg(tp + (c / a.denom) * 2), u1 = __ldg(tp + (c / a.denom) * 2 + 1);
                o[4 * c] = u0.x; o[4 * c + 1] = u0.y; o[4 * c + 2] = u1.x; o[4 * c + 3] = u1.y;
            }
        } else {
            t32_load<PAIRS>(a.pnode, p, o);
        }
        int min_scale, sc[NCM];
        if (a.safe) {
            min_scale = 0x7fffffff;
#pragma unroll
            for (int c = 0; c < NCM; c++) {
                sc[c] = (a.sdad[p * NCM + c] + (tip_mode ? 0 : a.snode[p * NCM + c])) & 0xff;
                min_scale = min(min_scale, sc[c]);
            }
        } else {
            min_scale = a.sdad[p] + (tip_mode ? 0 : a.snode[p]);
        }
        // ---- branch lnL, as lnl_t32 (w = val0 * node side; the tip table tip_val is val0 * tip_partial_lh) ----
        double lhb = 0.0;
#pragma unroll
        for (int c = 0; c < NCM; c++) {
            const double w0 = o[4 * c] * sv[4 * c], w1 = o[4 * c + 1] * sv[4 * c + 1], w2 = o[4 * c + 2] * sv[4 * c + 2],
                         w3 = o[4 * c + 3] * sv[4 * c + 3];
            const double s0 = fma(w2, d[4 * c + 2], w0 * d[4 * c]);
            const double s1 = fma(w3, d[4 * c + 3], w1 * d[4 * c + 1]);
            double lc = s0 + s1;
            if (a.safe) {
                if (sc[c] == min_scale + 1) lc *= B200_SCALING_THRESHOLD;
                else if (sc[c] > min_scale + 1) lc = 0.0;
            }
            a.pattern_lh_cat[(size_t)p * NCM + c] = lc;
            lhb += lc;
        }
        lhb = log(fabs(lhb) + a.ptn_invar[p]) + (double)min_scale * B200_LOG_SCALING_THRESHOLD;
        a.pattern_lh[p] = lhb;
        const double freq = a.ptn_freq[p];
        acc[0] = fma(lhb, freq, acc[0]);
        // ---- theta, as theta_t32_pattern ----
        double t[B];
        double2 *th = const_cast<double2 *>(t32_ptr(a.theta, p, PAIRS));
#pragma unroll
        for (int j = 0; j < PAIRS; j++) {
            double t0 = o[2 * j] * d[2 * j], t1 = o[2 * j + 1] * d[2 * j + 1];
            if (a.safe) {
                const int s2 = sc[j / 2];
                if (s2 == min_scale + 1) { t0 *= B200_SCALING_THRESHOLD; t1 *= B200_SCALING_THRESHOLD; }
                else if (s2 > min_scale + 1) { t0 = 0.0; t1 = 0.0; }
            }
            th[j * 32] = make_double2(t0, t1);
            t[2 * j] = t0;
            t[2 * j + 1] = t1;
        }
        a.buffer_scale_all[p] = (double)min_scale * B200_LOG_SCALING_THRESHOLD;
        // ---- df, ddf, as derv_t32 (observed patterns only: the caller rules +ASC out) ----
        double lh, df, ddf;
        t32_triple_regs<PAIRS>(t, sv, lh, df, ddf);
        lh = fabs(lh) + a.ptn_invar[p];
        const double inv = 1.0 / lh, df_frac = df * inv, ddf_frac = ddf * inv;
        const double tmp1 = df_frac * freq, tmp2 = ddf_frac * freq;
        acc[1] += tmp1;
        acc[2] += fma(-tmp1, df_frac, tmp2);
    }
    block_reduce_store<3>(acc, a.red_partial, (int)gridDim.x);
    // the CTA that finishes last adds the partial sums in a fixed order (final_reduce_kernel's)
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_ticket = atomicAdd(ta.counter, 1u);
    __syncthreads();
    if (s_ticket != gridDim.x - 1) return;
    __threadfence();
    __shared__ double sh[RED_THREADS];
    for (int k = 0; k < 3; k++) {
        double v = 0.0;
        for (int b = threadIdx.x; b < (int)gridDim.x; b += RED_THREADS) v += __ldcg(a.red_partial + (size_t)k * gridDim.x + b);
        sh[threadIdx.x] = v;
        __syncthreads();
        for (int s2 = RED_THREADS / 2; s2 > 0; s2 >>= 1) {
            if (threadIdx.x < s2) sh[threadIdx.x] += sh[threadIdx.x + s2];
            __syncthreads();
        }
        if (threadIdx.x == 0) ta.result[k] = sh[0];
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        *ta.counter = 0u;
        peer_exchange(ta.result, 3, ta.peer);
    }
}

__global__ void __launch_bounds__(RED_THREADS) eval_tail_kernel(BranchArgs a, TailArgs ta) {
    T32_DISPATCH(eval_tail_t32, a, ta)
}

bool can_fuse_tail(const b200lk_ctx *c) {
    return c->tile == 32 && !c->site && !c->nonrev && !c->batch && c->d.n_unobserved == 0 && c->ntot == c->d.nptn &&
           c->robust_keep >= 1.0 && !c->robust_median && !c->cur_branch_mixlen && !b200_env().no_fused_tail;
}

int launch_eval_tail(b200lk_ctx *c, b200lk_end dad, b200lk_end node, double length) {
    if (!c->tail_counter) {
        CUDA_TRY(c, cudaMalloc(&c->tail_counter, sizeof(unsigned)));
        CUDA_TRY(c, cudaMemsetAsync(c->tail_counter, 0, sizeof(unsigned), c->stream));
    }
    BranchArgs a;
    if (int rc = fill_args(c, a, dad, node)) return rc;
    fill_common(c, a);
    a.len = length;
    a.G = 1;
    TailArgs ta;
    ta.m = c->m;
    ta.len = length;
    ta.ncat = c->d.ncat;
    ta.counter = c->tail_counter;
    ta.result = c->d_result;
    ta.peer = peer_args_next(c);
    const int grid = grid_for(c, 1, c->ntot);
    eval_tail_kernel<<<grid, RED_THREADS, 0, c->stream>>>(a, ta);
    CUDA_TRY(c, cudaGetLastError());
    c->n_launches++;
    return B200LK_OK;
}

// ---- b200lk_lnl_from_buffer_noalloc: from-buffer lnL whose scratch lives in a file-scope device array ----------------
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The scratch that
// b200lk_lnl_from_buffer keeps in per-context cudaMalloc'ed buffers -- the branch coefficient table val[3][block] and
// the per-CTA partial sums [2][grid] -- lives here in ONE static array per GPU that loads the library.  Every call on
// a device shares it and it keeps the last call's contents: the kernels overwrite every entry they later read (val for
// all 3*block entries, the partial sums for every launched CTA), so nothing has to start at zero; calls that overlap
// on the same GPU (two streams, two host threads) race on it.
#define NOALLOC_VAL_DOUBLES (3 * B200LK_NOALLOC_MAX_BLOCK)
#define NOALLOC_RED_DOUBLES (2 * B200LK_NOALLOC_MAX_CTAS)
static __device__ __align__(256) double g_noalloc_scratch[NOALLOC_VAL_DOUBLES + NOALLOC_RED_DOUBLES];

#define NOALLOC_MAX_DEVICES 64
static double *g_noalloc_ptr[NOALLOC_MAX_DEVICES];  // device address of the array on each device (the symbol's name is
                                                    // only a host-side stand-in, never a pointer a kernel may be given)

// set-up, once per device, with that device current (called from b200lk_create; the entry point repeats it only for
// a device it has never seen)
int noalloc_setup(b200lk_ctx *c) {
    if (c->device < 0 || c->device >= NOALLOC_MAX_DEVICES) return B200LK_ERR_ARG;
    if (g_noalloc_ptr[c->device]) return B200LK_OK;
    CUDA_TRY(c, cudaSetDevice(c->device));
    double *ptr = nullptr;
    CUDA_TRY(c, cudaGetSymbolAddress((void **)&ptr, g_noalloc_scratch));
    g_noalloc_ptr[c->device] = ptr;
    return B200LK_OK;
}

extern "C" int b200lk_lnl_from_buffer_noalloc(b200lk_ctx *c, double length, double *lnl) {
    SHARD_REFUSE(c, "b200lk_lnl_from_buffer_noalloc");
    if (!c || !lnl) return B200LK_ERR_ARG;
    if (!c->model_set || !c->tips_set || !c->theta_computed) {
        snprintf(c->err, sizeof(c->err), "lnl_from_buffer_noalloc: set the model, upload the tips and build theta first");
        return B200LK_ERR_STATE;
    }
    if (c->nonrev || c->site || c->batch || c->cur_branch_mixlen || c->asc_holder) {
        snprintf(c->err, sizeof(c->err), "lnl_from_buffer_noalloc: reversible, non-site-specific contexts without Holder +ASC only");
        return B200LK_ERR_UNSUPPORTED;
    }
    if (c->block > B200LK_NOALLOC_MAX_BLOCK) {
        snprintf(c->err, sizeof(c->err),
                 "lnl_from_buffer_noalloc: nstates * ncat_mix = %d exceeds the static scratch (B200LK_NOALLOC_MAX_BLOCK = %d); "
                 "use b200lk_lnl_from_buffer",
                 c->block, B200LK_NOALLOC_MAX_BLOCK);
        return B200LK_ERR_ARG;
    }
    CUDA_TRY(c, cudaSetDevice(c->device));
    if (c->d.nptn == 0) { *lnl = 0.0; return B200LK_OK; }
    int rc = noalloc_setup(c);
    if (rc) return rc;
    double *sval = g_noalloc_ptr[c->device], *sred = sval + NOALLOC_VAL_DOUBLES;
    BatchLaunch nobatch;
    memset(&nobatch, 0, sizeof(nobatch));
    const int denom = c->d.fused_mix_rate ? 1 : c->d.ncat;
    val_kernel<<<(c->block + 255) / 256, 256, 0, c->stream>>>(c->m, length, nullptr, sval, c->tip_val, c->d.nstates, c->d.ncat,
                                                             c->ncat_mix, denom, c->d.nmix, c->d.n_state_codes, 0, nobatch);
    CUDA_TRY(c, cudaGetLastError());
    BranchArgs a;
    memset(&a, 0, sizeof(a));
    fill_common(c, a);
    a.val = sval;
    a.red_partial = sred;
    a.len = length;
    a.G = pick_group(c->block);
    int grid = std::min(grid_for(c, a.G, c->ntot), B200LK_NOALLOC_MAX_CTAS);  // grid-stride loops: any grid covers ntot
    frombuf_kernel<<<grid, RED_THREADS, 0, c->stream>>>(a, grid, nobatch);
    CUDA_TRY(c, cudaGetLastError());
    final_reduce_kernel<<<1, RED_THREADS, 0, c->stream>>>(sred, grid, 2, c->d_result, 0, 1, 8, nobatch);
    CUDA_TRY(c, cudaGetLastError());
    c->n_launches += 3;
    rc = launch_asc_fix(c, 0);  // works on d_result and pattern_lh only
    if (rc) return rc;
    CUDA_TRY(c, cudaMemcpyAsync(c->h_result, c->d_result, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    int warn = 0;
    if (!isfinite(c->h_result[0])) {  // tree/phylokernelnew.h:3338-3347, with the scratch for the repair pass's sums too
        warn = B200LK_WARN_LNL_NONFINITE;
        grid = std::min(grid_for(c, 1, c->d.nptn), B200LK_NOALLOC_MAX_CTAS);
        repair_kernel<<<grid, RED_THREADS, 0, c->stream>>>(c->pattern_lh, c->ptn_freq, c->d.nptn, sred, grid);
        CUDA_TRY(c, cudaGetLastError());
        final_reduce_kernel<<<1, RED_THREADS, 0, c->stream>>>(sred, grid, 1, c->d_result, 0, 8, 0, nobatch);
        CUDA_TRY(c, cudaGetLastError());
        c->n_launches += 2;
        CUDA_TRY(c, cudaMemcpyAsync(c->h_result, c->d_result, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
        CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    }
    *lnl = c->h_result[0];
    return warn;
}
