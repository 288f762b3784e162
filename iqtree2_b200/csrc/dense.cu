// dense.cu -- 20-state partial update on the FP64 tensor cores (replaces computePartialLikelihoodSIMD
// <.., nstates = 20, ..>, tree/phylokernelnew.h:1291-2039, for LG/WAG/... +G and the C10..C60 mixtures).
//
// Per node, pattern p and category c the update is three 20x20 products chained through registers:
//     L[p][x]   = sum_i A_left [p][c][i] * echild_L[c][x][i]          (leaf child: table lookup by tip state)
//     R[p][x]   = sum_i A_right[p][c][i] * echild_R[c][x][i]
//     out[p][i] = sum_x (L[p][x] * R[p][x]) * inv_evec[m][i][x]
// A warp owns 16 patterns (two m8n8k4 row tiles) and loops over the categories of a group; the [patterns x 20]
// operand is the A fragment, the 20x20 tables are B fragments pre-arranged in fragment order by tables.cu
// (one conflict-free 8-byte shared-memory load per lane and mma, reused by both row tiles), the 8x24
// accumulator of the first two products is multiplied element-wise in registers and fed back as the A
// operand of the third product by permuting its contraction index (no register shuffles except the last
// k-step).  Arithmetic per (pattern, category): 15 + 15 + 15 DMMA of 256 FMA.
//
// One CTA = one node x TILES_PER_CTA tiles of 64 patterns, so the tables (<= 45 KB per category group) are
// staged once per 64*TILES patterns; grid.y runs over the independent nodes of a level (level batching).
// NORM_LH needs max|out| over ALL categories of a pattern before it can decide to rescale: results are
// stored unscaled, the running max is kept per pattern, and the rare underflowing pattern is rescaled in
// place afterwards (SAFE_LH decides per category on the spot).
#include <stdlib.h>

#include "common.cuh"

#define DN_THREADS 128
#ifndef DN_MT
#define DN_MT 2                                   // row tiles (8 patterns) per warp
#endif
#ifndef DN_RP_MINB
#define DN_RP_MINB 3                              // resident CTAs per SM dense20_rp_kernel's registers are sized for
#endif
#define DN_PT (DN_THREADS / 32 * 8 * DN_MT)       // 64 patterns per CTA tile
#define DN_CG 4                                   // categories per staged group
#ifndef DN_MAX_TILES
#define DN_MAX_TILES 8
#endif

struct DenseArgs {
    const DevNode *nodes;
    const DevChild *children;
    const double *tables, *inv_frag, *ptn_invar;
    const double *tt;            // state-pair tables (virtual tip-tip children are read from them, dense20_rp_kernel)
    const uint16_t *pair_idx;    // their per-pattern pair indices
    const uint8_t *states;
    double *partial;
    uint8_t *scale;
    size_t slot_doubles, slot_scale_bytes;
    int64_t ntot, ntot_pad;
    int node_base, ncm, denom, codes, tiles_per_cta;
    // mixtures (ncm > DN_CG): grid.z runs over the category groups, so a CTA stages ONE group's tables for all its
    // tiles and the upper, narrow levels of the tree still fill the machine; the NORM_LH per-pattern maximum is then
    // combined across CTAs in `pmax` [node of the level][pattern] and applied by dense20_fix_kernel
    unsigned long long *pmax;
    int split;
};

__device__ __forceinline__ void dmma(double &d0, double &d1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(d0), "+d"(d1)
                 : "d"(a), "d"(b));
}

// factor of one child for category `cl` of the staged group, both row tiles, accumulator layout:
// f[mt][2j + r] = F[p = 8mt + g][x = 8j + 2q + r]
__device__ __forceinline__ void dense_child(const DenseArgs &a, const DevChild &ch, const double *sT, int cg, int cl,
                                            int c, const int64_t (&pc)[DN_MT], int lane, double (&f)[DN_MT][6]) {
    const int g = lane >> 2, q = lane & 3;
    (void)g;
    if (ch.kind == B200LK_CHILD_LEAF) {
#pragma unroll
        for (int mt = 0; mt < DN_MT; mt++) {
            const int st = a.states[(size_t)ch.id * a.ntot_pad + pc[mt]];
            const double *row = sT + ((size_t)st * cg + cl) * 20;
#pragma unroll
            for (int j = 0; j < 3; j++) {
                if (8 * j + 2 * q < 20) {
                    const double2 v = *reinterpret_cast<const double2 *>(row + 8 * j + 2 * q);
                    f[mt][2 * j] = v.x;
                    f[mt][2 * j + 1] = v.y;
                } else {
                    f[mt][2 * j] = 0.0;
                    f[mt][2 * j + 1] = 0.0;
                }
            }
        }
        return;
    }
    double A[DN_MT][5];
#pragma unroll
    for (int mt = 0; mt < DN_MT; mt++) {
        const double *src = a.partial + (size_t)ch.id * a.slot_doubles + (size_t)pc[mt] * (a.ncm * 20) + c * 20 + q;
#pragma unroll
        for (int s = 0; s < 5; s++) A[mt][s] = __ldg(src + 4 * s);
#pragma unroll
        for (int k = 0; k < 6; k++) f[mt][k] = 0.0;
    }
    const double *B = sT + (size_t)cl * B200_FRAG20 + lane;
#pragma unroll
    for (int s = 0; s < 5; s++) {
#pragma unroll
        for (int j = 0; j < 3; j++) {
            const double b = B[(j * 5 + s) * 32];
#pragma unroll
            for (int mt = 0; mt < DN_MT; mt++) dmma(f[mt][2 * j], f[mt][2 * j + 1], A[mt][s], b);
        }
    }
}

template <bool SAFE, int MINB>
__global__ void __launch_bounds__(DN_THREADS, MINB) dense20_kernel(DenseArgs a) {
    extern __shared__ double sm[];
    const int ncm = a.ncm, block = ncm * 20;
    const DevNode nd = a.nodes[a.node_base + blockIdx.y];
    const DevChild L = a.children[nd.child_begin];
    const DevChild R = a.children[nd.child_begin + 1];
    const bool Lleaf = L.kind == B200LK_CHILD_LEAF, Rleaf = R.kind == B200LK_CHILD_LEAF;
    const bool tiptip = Lleaf && Rleaf;
    const int leaf_doubles = a.codes * DN_CG * 20, frag_doubles = DN_CG * B200_FRAG20;
    const int child_doubles = leaf_doubles > frag_doubles ? leaf_doubles : frag_doubles;
    double *sL = sm, *sR = sL + child_doubles, *sI = sR + child_doubles;
    unsigned long long *smax = reinterpret_cast<unsigned long long *>(sI + frag_doubles);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, q = lane & 3;
    const int64_t tile0 = (int64_t)blockIdx.x * a.tiles_per_cta;
    double *out_base = a.partial + (size_t)nd.out_slot * a.slot_doubles;
    uint8_t *sc_out = a.scale + (size_t)nd.out_slot * a.slot_scale_bytes;
    const uint8_t *scL = Lleaf ? nullptr : a.scale + (size_t)L.id * a.slot_scale_bytes;
    const uint8_t *scR = Rleaf ? nullptr : a.scale + (size_t)R.id * a.slot_scale_bytes;
    if (!SAFE)
        for (int i = tid; i < a.tiles_per_cta * DN_PT; i += DN_THREADS) smax[i] = 0ull;

    const int cbeg = a.split ? (int)blockIdx.z * DN_CG : 0;
    const int cend = a.split ? min(ncm, cbeg + DN_CG) : ncm;
    for (int c0 = cbeg; c0 < cend; c0 += DN_CG) {
        const int cg = min(DN_CG, ncm - c0);
        __syncthreads();
        // ---- stage this group's tables --------------------------------------------------------------
        if (Lleaf) {
            const double *src = a.tables + L.leaf_off;
            for (int i = tid; i < a.codes * cg * 20; i += DN_THREADS) {
                const int s = i / (cg * 20), r = i - s * (cg * 20);
                sL[i] = src[(size_t)s * block + c0 * 20 + r];
            }
        } else {
            const double *src = a.tables + L.frag_off + (size_t)c0 * B200_FRAG20;
            for (int i = tid; i < cg * B200_FRAG20; i += DN_THREADS) sL[i] = src[i];
        }
        if (Rleaf) {
            const double *src = a.tables + R.leaf_off;
            for (int i = tid; i < a.codes * cg * 20; i += DN_THREADS) {
                const int s = i / (cg * 20), r = i - s * (cg * 20);
                sR[i] = src[(size_t)s * block + c0 * 20 + r];
            }
        } else {
            const double *src = a.tables + R.frag_off + (size_t)c0 * B200_FRAG20;
            for (int i = tid; i < cg * B200_FRAG20; i += DN_THREADS) sR[i] = src[i];
        }
        for (int i = tid; i < cg * B200_FRAG20; i += DN_THREADS) {
            const int cl = i / B200_FRAG20;
            sI[i] = a.inv_frag[(size_t)((c0 + cl) / a.denom) * B200_FRAG20 + (i - cl * B200_FRAG20)];
        }
        __syncthreads();
        // ---- tiles of 64 patterns ---------------------------------------------------------------------
        for (int t = 0; t < a.tiles_per_cta; t++) {
            const int64_t pbase = (tile0 + t) * DN_PT + warp * (8 * DN_MT);
            if (pbase >= a.ntot) break;  // warp-uniform
            int64_t p[DN_MT], pc[DN_MT];
            bool invar0[DN_MT];
#pragma unroll
            for (int mt = 0; mt < DN_MT; mt++) {
                p[mt] = pbase + mt * 8 + g;
                pc[mt] = p[mt] < a.ntot ? p[mt] : a.ntot - 1;
                invar0[mt] = a.ptn_invar[pc[mt]] == 0.0;
            }
            unsigned long long rmax[DN_MT] = {};
            for (int cl = 0; cl < cg; cl++) {
                const int c = c0 + cl;
                double fl[DN_MT][6], fr[DN_MT][6];
                dense_child(a, L, sL, cg, cl, c, pc, lane, fl);
                dense_child(a, R, sR, cg, cl, c, pc, lane, fr);
                const double *B = sI + (size_t)cl * B200_FRAG20 + lane;
#pragma unroll
                for (int mt = 0; mt < DN_MT; mt++) {
                    double tmp[6];
#pragma unroll
                    for (int k = 0; k < 6; k++) tmp[k] = fl[mt][k] * fr[mt][k];
                    // x = 16..19 sit in lanes q = 0, 1 of the quad: bring x = 16 + q to lane q
                    const int srcl = (lane & ~3) | (q >> 1);
                    const double v0 = __shfl_sync(0xffffffffu, tmp[4], srcl), v1 = __shfl_sync(0xffffffffu, tmp[5], srcl);
                    const double a4 = (q & 1) ? v1 : v0;
                    double o[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
#pragma unroll
                    for (int jo = 0; jo < 3; jo++) {
#pragma unroll
                        for (int kk = 0; kk < 4; kk++) dmma(o[2 * jo], o[2 * jo + 1], tmp[kk], B[(jo * 5 + kk) * 32]);
                        dmma(o[2 * jo], o[2 * jo + 1], a4, B[(jo * 5 + 4) * 32]);
                    }
                    unsigned long long mx = 0ull;
#pragma unroll
                    for (int k = 0; k < 6; k++) {
                        const unsigned long long b = (unsigned long long)__double_as_longlong(o[k]) & 0x7fffffffffffffffull;
                        mx = b > mx ? b : mx;
                    }
                    int sc = 0;
                    if (SAFE) {
                        mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
                        mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
                        sc = (Lleaf ? 0 : scL[pc[mt] * ncm + c]) + (Rleaf ? 0 : scR[pc[mt] * ncm + c]);
                        if (!tiptip && __longlong_as_double((long long)mx) < B200_SCALING_THRESHOLD && invar0[mt]) {
#pragma unroll
                            for (int k = 0; k < 6; k++) o[k] *= B200_TWO_256;
                            sc += 1;
                        }
                    } else {
                        rmax[mt] = mx > rmax[mt] ? mx : rmax[mt];
                    }
                    if (p[mt] < a.ntot) {
                        double *po = out_base + (size_t)p[mt] * block + c * 20 + 2 * q;
                        *reinterpret_cast<double2 *>(po) = make_double2(o[0], o[1]);
                        *reinterpret_cast<double2 *>(po + 8) = make_double2(o[2], o[3]);
                        if (q < 2) *reinterpret_cast<double2 *>(po + 16) = make_double2(o[4], o[5]);
                        if (SAFE && q == 0) sc_out[p[mt] * ncm + c] = (uint8_t)sc;
                    }
                }
            }
            if (!SAFE) {
#pragma unroll
                for (int mt = 0; mt < DN_MT; mt++) {
                    unsigned long long mx = rmax[mt];
                    mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
                    mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
                    if (q == 0) {
                        unsigned long long *slot = &smax[t * DN_PT + warp * (8 * DN_MT) + mt * 8 + g];
                        if (mx > *slot) *slot = mx;
                    }
                }
            }
        }
    }
    if (SAFE) return;
    // ---- NORM_LH: one decision per pattern over the whole block (:1889-1903, 2015-2029) ---------------------
    __syncwarp();  // smax entries are written and read by lanes of the same warp
    for (int t = 0; t < a.tiles_per_cta; t++) {
        const int64_t pbase = (tile0 + t) * DN_PT + warp * (8 * DN_MT);
        if (pbase >= a.ntot) break;
#pragma unroll
        for (int mt = 0; mt < DN_MT; mt++) {
            const int64_t p = pbase + mt * 8 + g;
            if (p >= a.ntot) continue;
            if (a.split) {  // this CTA saw one category group only: combine across groups, decide in the fix kernel
                if (q == 0) atomicMax(&a.pmax[(size_t)blockIdx.y * a.ntot + p], smax[t * DN_PT + warp * (8 * DN_MT) + mt * 8 + g]);
                continue;
            }
            const double mx = __longlong_as_double((long long)smax[t * DN_PT + warp * (8 * DN_MT) + mt * 8 + g]);
            const bool under = !tiptip && mx < B200_SCALING_THRESHOLD && a.ptn_invar[p] == 0.0;
            if (under) {  // rare: rescale the stored block in place, the quad shares the work
                double *po = out_base + (size_t)p * block;
                for (int e = q; e < block; e += 4) po[e] *= B200_TWO_256;
            }
            if (q == 0) sc_out[p] = (uint8_t)((Lleaf ? 0 : scL[p]) + (Rleaf ? 0 : scR[p]) + (under ? 1 : 0));
        }
    }
}

// ---- pipelined variant --------------------------------------------------------------------------------------------
// Same arithmetic, different operand path: every warp prefetches the [16 patterns x 20] operand tiles of the NEXT
// (tile, category) step of its internal children into a private shared-memory double buffer with cp.async (16-byte
// chunks, L2 only) while its tensor-core chain works on the current step, so that HBM latency is no longer exposed
// between the load and the first DMMA (the level-batched kernel above sat on long_scoreboard at 12-16 warps per SM).
#define DN_STAGE_DOUBLES (2 * 16 * 20)  // one stage of one warp: two children x 16 patterns x 20

__device__ __forceinline__ void cp_async16(void *smem_dst, const void *gsrc) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gsrc));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N)); }

// factor of an internal child from the staged operand tile `sa` ([16][20] doubles), fragment tables `sT`
__device__ __forceinline__ void dense_child_staged(const double *sa, const double *sT, int cl, int lane,
                                                   double (&f)[DN_MT][6]) {
    const int g = lane >> 2, q = lane & 3;
    double A[DN_MT][5];
#pragma unroll
    for (int mt = 0; mt < DN_MT; mt++) {
#pragma unroll
        for (int s = 0; s < 5; s++) A[mt][s] = sa[(mt * 8 + g) * 20 + 4 * s + q];
#pragma unroll
        for (int k = 0; k < 6; k++) f[mt][k] = 0.0;
    }
    const double *B = sT + (size_t)cl * B200_FRAG20 + lane;
#pragma unroll
    for (int s = 0; s < 5; s++) {
#pragma unroll
        for (int j = 0; j < 3; j++) {
            const double b = B[(j * 5 + s) * 32];
#pragma unroll
            for (int mt = 0; mt < DN_MT; mt++) dmma(f[mt][2 * j], f[mt][2 * j + 1], A[mt][s], b);
        }
    }
}

__device__ __forceinline__ void dense_leaf_lookup(const double *sT, int cg, int cl, const int (&st)[DN_MT], int lane,
                                                  double (&f)[DN_MT][6]) {
    const int q = lane & 3;
#pragma unroll
    for (int mt = 0; mt < DN_MT; mt++) {
        const double *row = sT + ((size_t)st[mt] * cg + cl) * 20;
#pragma unroll
        for (int j = 0; j < 3; j++) {
            if (8 * j + 2 * q < 20) {
                const double2 v = *reinterpret_cast<const double2 *>(row + 8 * j + 2 * q);
                f[mt][2 * j] = v.x;
                f[mt][2 * j + 1] = v.y;
            } else {
                f[mt][2 * j] = 0.0;
                f[mt][2 * j + 1] = 0.0;
            }
        }
    }
}

template <bool SAFE>
__global__ void __launch_bounds__(DN_THREADS, 2) dense20_pipe_kernel(DenseArgs a) {
    extern __shared__ double sm[];
    const int ncm = a.ncm, block = ncm * 20;
    const DevNode nd = a.nodes[a.node_base + blockIdx.y];
    const DevChild L = a.children[nd.child_begin];
    const DevChild R = a.children[nd.child_begin + 1];
    const bool Lleaf = L.kind == B200LK_CHILD_LEAF, Rleaf = R.kind == B200LK_CHILD_LEAF;
    const bool tiptip = Lleaf && Rleaf;
    const int leaf_doubles = a.codes * DN_CG * 20, frag_doubles = DN_CG * B200_FRAG20;
    const int child_doubles = leaf_doubles > frag_doubles ? leaf_doubles : frag_doubles;
    double *sL = sm, *sR = sL + child_doubles, *sI = sR + child_doubles;
    unsigned long long *smax = reinterpret_cast<unsigned long long *>(sI + frag_doubles);
    double *stage_all = reinterpret_cast<double *>(smax + a.tiles_per_cta * DN_PT);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, q = lane & 3;
    double *wbuf = stage_all + (size_t)warp * 2 * DN_STAGE_DOUBLES;  // [2 stages][2 children][16][20]
    const int64_t tile0 = (int64_t)blockIdx.x * a.tiles_per_cta;
    double *out_base = a.partial + (size_t)nd.out_slot * a.slot_doubles;
    uint8_t *sc_out = a.scale + (size_t)nd.out_slot * a.slot_scale_bytes;
    const uint8_t *scL = Lleaf ? nullptr : a.scale + (size_t)L.id * a.slot_scale_bytes;
    const uint8_t *scR = Rleaf ? nullptr : a.scale + (size_t)R.id * a.slot_scale_bytes;
    const double *pL = Lleaf ? nullptr : a.partial + (size_t)L.id * a.slot_doubles;
    const double *pR = Rleaf ? nullptr : a.partial + (size_t)R.id * a.slot_doubles;
    if (!SAFE)
        for (int i = tid; i < a.tiles_per_cta * DN_PT; i += DN_THREADS) smax[i] = 0ull;
    // tiles this warp really has (its 16-pattern strip of each 64-pattern tile)
    int my_tiles = 0;
    for (int t = 0; t < a.tiles_per_cta; t++)
        if ((tile0 + t) * DN_PT + warp * (8 * DN_MT) < a.ntot) my_tiles = t + 1;

    const int cbeg = a.split ? (int)blockIdx.z * DN_CG : 0;
    const int cend = a.split ? min(ncm, cbeg + DN_CG) : ncm;
    for (int c0 = cbeg; c0 < cend; c0 += DN_CG) {
        const int cg = min(DN_CG, ncm - c0);
        __syncthreads();
        if (Lleaf) {
            const double *src = a.tables + L.leaf_off;
            for (int i = tid; i < a.codes * cg * 20; i += DN_THREADS) {
                const int s = i / (cg * 20), r = i - s * (cg * 20);
                sL[i] = src[(size_t)s * block + c0 * 20 + r];
            }
        } else {
            const double *src = a.tables + L.frag_off + (size_t)c0 * B200_FRAG20;
            for (int i = tid; i < cg * B200_FRAG20; i += DN_THREADS) sL[i] = src[i];
        }
        if (Rleaf) {
            const double *src = a.tables + R.leaf_off;
            for (int i = tid; i < a.codes * cg * 20; i += DN_THREADS) {
                const int s = i / (cg * 20), r = i - s * (cg * 20);
                sR[i] = src[(size_t)s * block + c0 * 20 + r];
            }
        } else {
            const double *src = a.tables + R.frag_off + (size_t)c0 * B200_FRAG20;
            for (int i = tid; i < cg * B200_FRAG20; i += DN_THREADS) sR[i] = src[i];
        }
        for (int i = tid; i < cg * B200_FRAG20; i += DN_THREADS) {
            const int cl = i / B200_FRAG20;
            sI[i] = a.inv_frag[(size_t)((c0 + cl) / a.denom) * B200_FRAG20 + (i - cl * B200_FRAG20)];
        }
        __syncthreads();

        const int nsteps = my_tiles * cg;
        // operand tiles of step `step` -> stage `stage` of this warp's buffer
        auto prefetch = [&](int step, int stage) {
            const int t = step / cg, cl = step - t * cg;
            const int64_t pbase = (tile0 + t) * DN_PT + warp * (8 * DN_MT);
            double *dst = wbuf + (size_t)stage * DN_STAGE_DOUBLES;
#pragma unroll
            for (int j = 0; j < 5; j++) {
                const int k = lane + 32 * j, row = k / 10, col = k - row * 10;
                int64_t p = pbase + row;
                if (p >= a.ntot) p = a.ntot - 1;
                const size_t off = (size_t)p * block + (size_t)(c0 + cl) * 20 + col * 2;
                if (!Lleaf) cp_async16(dst + row * 20 + col * 2, pL + off);
                if (!Rleaf) cp_async16(dst + 320 + row * 20 + col * 2, pR + off);
            }
        };
        if (nsteps > 0 && !tiptip) prefetch(0, 0);
        cp_async_commit();
        int stL[DN_MT] = {}, stR[DN_MT] = {};
        int64_t p[DN_MT] = {}, pc[DN_MT] = {};
        bool invar0[DN_MT] = {};
        unsigned long long rmax[DN_MT] = {};
        for (int step = 0; step < nsteps; step++) {
            const int t = step / cg, cl = step - t * cg, c = c0 + cl;
            if (step + 1 < nsteps && !tiptip) prefetch(step + 1, (step + 1) & 1);
            cp_async_commit();
            if (cl == 0) {  // new tile: rows, tip states
                const int64_t pbase = (tile0 + t) * DN_PT + warp * (8 * DN_MT);
#pragma unroll
                for (int mt = 0; mt < DN_MT; mt++) {
                    p[mt] = pbase + mt * 8 + g;
                    pc[mt] = p[mt] < a.ntot ? p[mt] : a.ntot - 1;
                    invar0[mt] = a.ptn_invar[pc[mt]] == 0.0;
                    if (Lleaf) stL[mt] = a.states[(size_t)L.id * a.ntot_pad + pc[mt]];
                    if (Rleaf) stR[mt] = a.states[(size_t)R.id * a.ntot_pad + pc[mt]];
                    rmax[mt] = 0ull;
                }
            }
            cp_async_wait<1>();
            __syncwarp();
            const double *sa = wbuf + (size_t)(step & 1) * DN_STAGE_DOUBLES;
            double fl[DN_MT][6], fr[DN_MT][6];
            if (Lleaf) dense_leaf_lookup(sL, cg, cl, stL, lane, fl);
            else dense_child_staged(sa, sL, cl, lane, fl);
            if (Rleaf) dense_leaf_lookup(sR, cg, cl, stR, lane, fr);
            else dense_child_staged(sa + 320, sR, cl, lane, fr);
            __syncwarp();  // the stage may be overwritten by the prefetch issued in the next iteration
            const double *B = sI + (size_t)cl * B200_FRAG20 + lane;
#pragma unroll
            for (int mt = 0; mt < DN_MT; mt++) {
                double tmp[6];
#pragma unroll
                for (int k = 0; k < 6; k++) tmp[k] = fl[mt][k] * fr[mt][k];
                const int srcl = (lane & ~3) | (q >> 1);
                const double v0 = __shfl_sync(0xffffffffu, tmp[4], srcl), v1 = __shfl_sync(0xffffffffu, tmp[5], srcl);
                const double a4 = (q & 1) ? v1 : v0;
                double o[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
#pragma unroll
                for (int jo = 0; jo < 3; jo++) {
#pragma unroll
                    for (int kk = 0; kk < 4; kk++) dmma(o[2 * jo], o[2 * jo + 1], tmp[kk], B[(jo * 5 + kk) * 32]);
                    dmma(o[2 * jo], o[2 * jo + 1], a4, B[(jo * 5 + 4) * 32]);
                }
                unsigned long long mx = 0ull;
#pragma unroll
                for (int k = 0; k < 6; k++) {
                    const unsigned long long b = (unsigned long long)__double_as_longlong(o[k]) & 0x7fffffffffffffffull;
                    mx = b > mx ? b : mx;
                }
                int sc = 0;
                if (SAFE) {
                    mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
                    mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
                    sc = (Lleaf ? 0 : scL[pc[mt] * ncm + c]) + (Rleaf ? 0 : scR[pc[mt] * ncm + c]);
                    if (!tiptip && __longlong_as_double((long long)mx) < B200_SCALING_THRESHOLD && invar0[mt]) {
#pragma unroll
                        for (int k = 0; k < 6; k++) o[k] *= B200_TWO_256;
                        sc += 1;
                    }
                } else {
                    rmax[mt] = mx > rmax[mt] ? mx : rmax[mt];
                }
                if (p[mt] < a.ntot) {
                    double *po = out_base + (size_t)p[mt] * block + c * 20 + 2 * q;
                    *reinterpret_cast<double2 *>(po) = make_double2(o[0], o[1]);
                    *reinterpret_cast<double2 *>(po + 8) = make_double2(o[2], o[3]);
                    if (q < 2) *reinterpret_cast<double2 *>(po + 16) = make_double2(o[4], o[5]);
                    if (SAFE && q == 0) sc_out[p[mt] * ncm + c] = (uint8_t)sc;
                }
            }
            if (!SAFE && cl == cg - 1) {  // last category of this group for the tile: fold the running maximum
#pragma unroll
                for (int mt = 0; mt < DN_MT; mt++) {
                    unsigned long long mx = rmax[mt];
                    mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
                    mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
                    if (q == 0) {
                        unsigned long long *slot = &smax[t * DN_PT + warp * (8 * DN_MT) + mt * 8 + g];
                        if (mx > *slot) *slot = mx;
                    }
                }
            }
        }
        cp_async_wait<0>();
    }
    if (SAFE) return;
    __syncwarp();
    for (int t = 0; t < a.tiles_per_cta; t++) {
        const int64_t pbase = (tile0 + t) * DN_PT + warp * (8 * DN_MT);
        if (pbase >= a.ntot) break;
#pragma unroll
        for (int mt = 0; mt < DN_MT; mt++) {
            const int64_t pp = pbase + mt * 8 + g;
            if (pp >= a.ntot) continue;
            if (a.split) {
                if (q == 0) atomicMax(&a.pmax[(size_t)blockIdx.y * a.ntot + pp], smax[t * DN_PT + warp * (8 * DN_MT) + mt * 8 + g]);
                continue;
            }
            const double mx = __longlong_as_double((long long)smax[t * DN_PT + warp * (8 * DN_MT) + mt * 8 + g]);
            const bool under = !tiptip && mx < B200_SCALING_THRESHOLD && a.ptn_invar[pp] == 0.0;
            if (under) {
                double *po = out_base + (size_t)pp * block;
                for (int e = q; e < block; e += 4) po[e] *= B200_TWO_256;
            }
            if (q == 0) sc_out[pp] = (uint8_t)((Lleaf ? 0 : scL[pp]) + (Rleaf ? 0 : scR[pp]) + (under ? 1 : 0));
        }
    }
}

// ---- register-prefetch variant ------------------------------------------------------------------------------------
// The operand fragments of BOTH internal children of step s+1 are requested while the tensor-core chain of step s
// is still running: the left child's fragments right after its product has consumed them, the right child's after
// its product; HBM latency then hides behind >= 60 DMMAs of the same warp instead of stalling the first DMMA of
// every product (ncu: ~40 % of all stall samples sat on those DMMAs).
__device__ __forceinline__ void dense_load_A(const double *base, size_t block, int c, const int64_t (&pc)[DN_MT], int q,
                                             double (&A)[DN_MT][5]) {
#pragma unroll
    for (int mt = 0; mt < DN_MT; mt++) {
        const double *src = base + (size_t)pc[mt] * block + c * 20 + q;
#pragma unroll
        for (int s = 0; s < 5; s++) A[mt][s] = __ldg(src + 4 * s);
    }
}
// the same with the rows' element offsets (row * block + q) formed once per tile instead of once per step
__device__ __forceinline__ void dense_load_A_at(const double *base, int c, const int64_t (&off)[DN_MT], double (&A)[DN_MT][5]) {
#pragma unroll
    for (int mt = 0; mt < DN_MT; mt++) {
        const double *src = base + off[mt] + c * 20;
#pragma unroll
        for (int s = 0; s < 5; s++) A[mt][s] = __ldg(src + 4 * s);
    }
}
// ... and for a virtual tip-tip child: the rows are those of its state-pair table (32-bit offsets: api.cu keeps them small)
__device__ __forceinline__ void dense_load_A_at(const double *base, int c, const int (&off)[DN_MT], double (&A)[DN_MT][5]) {
#pragma unroll
    for (int mt = 0; mt < DN_MT; mt++) {
        const double *src = base + off[mt] + c * 20;
#pragma unroll
        for (int s = 0; s < 5; s++) A[mt][s] = __ldg(src + 4 * s);
    }
}
__device__ __forceinline__ void dense_mma_child(const double (&A)[DN_MT][5], const double *sT, int cl, int lane,
                                                double (&f)[DN_MT][6]) {
#pragma unroll
    for (int mt = 0; mt < DN_MT; mt++)
#pragma unroll
        for (int k = 0; k < 6; k++) f[mt][k] = 0.0;
    const double *B = sT + (size_t)cl * B200_FRAG20 + lane;
#pragma unroll
    for (int s = 0; s < 5; s++) {
#pragma unroll
        for (int j = 0; j < 3; j++) {
            const double b = B[(j * 5 + s) * 32];
#pragma unroll
            for (int mt = 0; mt < DN_MT; mt++) dmma(f[mt][2 * j], f[mt][2 * j + 1], A[mt][s], b);
        }
    }
}

// VIRT: the build that can read a virtual tip-tip child from its state-pair table (168 registers); launches without such a
// child -- batched ones never have any -- take the lean build (160 registers: cfg5 31.1 -> 30.1 ms).
template <bool SAFE, bool VIRT = true>
__global__ void __launch_bounds__(DN_THREADS, DN_RP_MINB) dense20_rp_kernel(DenseArgs a, BatchLaunch bl) {
    extern __shared__ double sm[];
    if (bl.d_ctx) {  // batched launch: grid.z = context of the supermatrix, same node structure, own data
        const BatchCtxDev &b = bl.d_ctx[blockIdx.z];
        a.tables = b.tables;
        a.inv_frag = b.m.inv_frag;
        a.ptn_invar = b.ptn_invar;
        a.states = b.states;
        a.partial = b.partial;
        a.scale = b.scale;
        a.slot_doubles = b.slot_doubles;
        a.slot_scale_bytes = b.slot_scale_bytes;
        a.ntot = b.ntot;
        a.ntot_pad = b.ntot_pad;
        if ((int64_t)blockIdx.x * a.tiles_per_cta * DN_PT >= a.ntot) return;  // smaller context than the grid
    }
    const int ncm = a.ncm, block = ncm * 20;
    const DevNode nd = a.nodes[a.node_base + blockIdx.y];
    const DevChild L = a.children[nd.child_begin];
    const DevChild R = a.children[nd.child_begin + 1];
    const bool Lleaf = L.kind == B200LK_CHILD_LEAF, Rleaf = R.kind == B200LK_CHILD_LEAF;
    const bool tiptip = Lleaf && Rleaf;
    // a tip-tip child left virtual (api.cu): its rows are read from the node's state-pair table by per-pattern pair index,
    // not from a resident vector; its scale counters are zero
    const bool Lvirt = VIRT && !Lleaf && L.vtt_off >= 0, Rvirt = VIRT && !Rleaf && R.vtt_off >= 0;
    const int leaf_doubles = a.codes * DN_CG * 20, frag_doubles = DN_CG * B200_FRAG20;
    const int child_doubles = leaf_doubles > frag_doubles ? leaf_doubles : frag_doubles;
    double *sL = sm, *sR = sL + child_doubles, *sI = sR + child_doubles;
    unsigned long long *smax = reinterpret_cast<unsigned long long *>(sI + frag_doubles);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, q = lane & 3;
    const int64_t tile0 = (int64_t)blockIdx.x * a.tiles_per_cta;
    double *out_base = a.partial + (size_t)nd.out_slot * a.slot_doubles;
    uint8_t *sc_out = a.scale + (size_t)nd.out_slot * a.slot_scale_bytes;
    const bool Lsc0 = Lleaf || Lvirt, Rsc0 = Rleaf || Rvirt;  // children that contribute no scale count
    const uint8_t *scL = Lsc0 ? nullptr : a.scale + (size_t)L.id * a.slot_scale_bytes;
    const uint8_t *scR = Rsc0 ? nullptr : a.scale + (size_t)R.id * a.slot_scale_bytes;
    const double *pL = Lleaf ? nullptr : (Lvirt ? a.tt + L.vtt_off : a.partial + (size_t)L.id * a.slot_doubles);
    const double *pR = Rleaf ? nullptr : (Rvirt ? a.tt + R.vtt_off : a.partial + (size_t)R.id * a.slot_doubles);
    const uint16_t *vpL = Lvirt ? a.pair_idx + L.vpair_off : nullptr, *vpR = Rvirt ? a.pair_idx + R.vpair_off : nullptr;
    if (!SAFE)
        for (int i = tid; i < a.tiles_per_cta * DN_PT; i += DN_THREADS) smax[i] = 0ull;
    int my_tiles = 0;
    for (int t = 0; t < a.tiles_per_cta; t++)
        if ((tile0 + t) * DN_PT + warp * (8 * DN_MT) < a.ntot) my_tiles = t + 1;

    const int cbeg = a.split ? (int)blockIdx.z * DN_CG : 0;
    const int cend = a.split ? min(ncm, cbeg + DN_CG) : ncm;
    for (int c0 = cbeg; c0 < cend; c0 += DN_CG) {
        const int cg = min(DN_CG, ncm - c0);
        __syncthreads();
        // The three tables of the group (<= 45 KB) go to shared memory as 16-byte cp.async copies: every copy of a
        // thread is in flight at once and no register holds the data.  (The element-by-element loop this replaces had
        // ONE load in flight per thread and iteration: 14 % of the kernel's stall samples sat on its stores,
        // profiles/r2k_dense20_rp_cfg3_50k.md.)  All offsets are even numbers of doubles: rows of 20 * cg, fragments of 480.
        auto stage_child = [&](const DevChild &ch, bool leaf, double *dst) {
            if (leaf) {
                const double *src = a.tables + ch.leaf_off + c0 * 20;
                const int rowc = cg * 10;  // 16-byte chunks per state row
                for (int i = tid; i < a.codes * rowc; i += DN_THREADS) {
                    const int s = i / rowc, r = i - s * rowc;
                    cp_async16(dst + (size_t)s * (cg * 20) + 2 * r, src + (size_t)s * block + 2 * r);
                }
            } else {
                const double *src = a.tables + ch.frag_off + (size_t)c0 * B200_FRAG20;
                for (int i = tid; i < cg * (B200_FRAG20 / 2); i += DN_THREADS) cp_async16(dst + 2 * i, src + 2 * i);
            }
        };
        stage_child(L, Lleaf, sL);
        stage_child(R, Rleaf, sR);
        for (int i = tid; i < cg * (B200_FRAG20 / 2); i += DN_THREADS) {
            const int cl = i / (B200_FRAG20 / 2), r = i - cl * (B200_FRAG20 / 2);
            cp_async16(sI + (size_t)cl * B200_FRAG20 + 2 * r, a.inv_frag + (size_t)((c0 + cl) / a.denom) * B200_FRAG20 + 2 * r);
        }
        cp_async_commit();

        const int nsteps = my_tiles * cg;
        // Loop bookkeeping is incremental: (t, cl) count up instead of step / cg and step % cg, and everything that depends
        // on the tile only -- its rows (clamped so that out-of-range rows read the last pattern), their element offsets
        // row * block for the operand loads and the output -- is formed when a tile opens, not at every step (on the source
        // page of the r2l capture this arithmetic was 22 % of the instructions and 16 % of the stall samples).
        double AL[DN_MT][5], AR[DN_MT][5];
        int64_t pc[DN_MT] = {}, pcn[DN_MT] = {};          // clamped rows: current tile, tile of the NEXT step
        int64_t offn[DN_MT] = {}, offo[DN_MT] = {};       // element offsets: next step's operands, this tile's output
        bool live[DN_MT] = {}, liven[DN_MT] = {};         // row < ntot
        auto rows_of_tile = [&](int t) {
            const int64_t pbase = (tile0 + t) * DN_PT + warp * (8 * DN_MT);
#pragma unroll
            for (int mt = 0; mt < DN_MT; mt++) {
                const int64_t pp = pbase + mt * 8 + g;
                liven[mt] = pp < a.ntot;
                pcn[mt] = liven[mt] ? pp : a.ntot - 1;
                offn[mt] = pcn[mt] * block + q;
            }
        };
        // virtual children: pair indices of the tile AFTER the current one (requested when a tile opens, a whole tile before
        // the boundary step turns them into the element offsets of the next tile's operand loads)
        int vpnL[DN_MT] = {}, vpnR[DN_MT] = {}, voffL[DN_MT] = {}, voffR[DN_MT] = {};
        auto pairs_of_tile = [&](int t) {
            const int64_t pbase = (tile0 + t) * DN_PT + warp * (8 * DN_MT);
#pragma unroll
            for (int mt = 0; mt < DN_MT; mt++) {
                int64_t pp = pbase + mt * 8 + g;
                pp = pp < a.ntot ? pp : a.ntot - 1;
                if (Lvirt) vpnL[mt] = __ldg(vpL + pp);
                if (Rvirt) vpnR[mt] = __ldg(vpR + pp);
            }
        };
        auto pairs_to_offsets = [&]() {
#pragma unroll
            for (int mt = 0; mt < DN_MT; mt++) {
                voffL[mt] = vpnL[mt] * block + q;
                voffR[mt] = vpnR[mt] * block + q;
            }
        };
        // per-tile inputs (tip states, invariant-site flag): requested one step before the tile starts, so that the
        // first table lookup of a tile does not wait for them
        int stL[DN_MT] = {}, stR[DN_MT] = {}, stLn[DN_MT] = {}, stRn[DN_MT] = {};
        double invn[DN_MT] = {};
        bool invar0[DN_MT] = {};
        auto tile_inputs = [&](const int64_t (&rows)[DN_MT]) {
#pragma unroll
            for (int mt = 0; mt < DN_MT; mt++) {
                invn[mt] = __ldg(a.ptn_invar + rows[mt]);
                if (Lleaf) stLn[mt] = __ldg(a.states + (size_t)L.id * a.ntot_pad + rows[mt]);
                if (Rleaf) stRn[mt] = __ldg(a.states + (size_t)R.id * a.ntot_pad + rows[mt]);
            }
        };
        if (nsteps > 0) {
            rows_of_tile(0);
            if (Lvirt || Rvirt) {
                pairs_of_tile(0);
                pairs_to_offsets();
            }
            if (Lvirt) dense_load_A_at(pL, c0, voffL, AL);
            else if (!Lleaf) dense_load_A_at(pL, c0, offn, AL);
            if (Rvirt) dense_load_A_at(pR, c0, voffR, AR);
            else if (!Rleaf) dense_load_A_at(pR, c0, offn, AR);
            tile_inputs(pcn);
        }
        cp_async_wait<0>();  // the tables; the first operands above stay in flight across the barrier
        __syncthreads();
        unsigned long long rmax[DN_MT] = {};
        int t = 0, cl = 0;
        for (int step = 0; step < nsteps; step++) {
            const int c = c0 + cl;
            const bool more = step + 1 < nsteps, last_cl = cl == cg - 1;
            const int cn = c0 + ((more && !last_cl) ? cl + 1 : 0);
            if (cl == 0) {  // a tile opens: what was "next" is current
#pragma unroll
                for (int mt = 0; mt < DN_MT; mt++) {
                    pc[mt] = pcn[mt];
                    live[mt] = liven[mt];
                    offo[mt] = offn[mt] + q;  // row * block + 2 q
                    invar0[mt] = invn[mt] == 0.0;
                    stL[mt] = stLn[mt];
                    stR[mt] = stRn[mt];
                    rmax[mt] = 0ull;
                }
                if ((Lvirt || Rvirt) && t + 1 < my_tiles) pairs_of_tile(t + 1);
            }
            if (more && last_cl) {  // the next step opens a tile
                rows_of_tile(t + 1);
                tile_inputs(pcn);
                if (Lvirt || Rvirt) pairs_to_offsets();
            }
            double fl[DN_MT][6], fr[DN_MT][6];
            if (Lleaf) dense_leaf_lookup(sL, cg, cl, stL, lane, fl);
            else {
                dense_mma_child(AL, sL, cl, lane, fl);
                if (more) {  // next step's left operands, in flight from here
                    if (Lvirt) dense_load_A_at(pL, cn, voffL, AL);
                    else dense_load_A_at(pL, cn, offn, AL);
                }
            }
            if (Rleaf) dense_leaf_lookup(sR, cg, cl, stR, lane, fr);
            else {
                dense_mma_child(AR, sR, cl, lane, fr);
                if (more) {
                    if (Rvirt) dense_load_A_at(pR, cn, voffR, AR);
                    else dense_load_A_at(pR, cn, offn, AR);
                }
            }
            const double *B = sI + (size_t)cl * B200_FRAG20 + lane;
            // back-transformation of BOTH row tiles together: every B fragment is fetched once and feeds two DMMAs, and six
            // accumulator chains (3 n-tiles x 2 row tiles) are in flight instead of three
            double tmpa[DN_MT][6], a4a[DN_MT], oa[DN_MT][6];
#pragma unroll
            for (int mt = 0; mt < DN_MT; mt++) {
#pragma unroll
                for (int k = 0; k < 6; k++) {
                    tmpa[mt][k] = fl[mt][k] * fr[mt][k];
                    oa[mt][k] = 0.0;
                }
                const int srcl = (lane & ~3) | (q >> 1);
                const double v0 = __shfl_sync(0xffffffffu, tmpa[mt][4], srcl), v1 = __shfl_sync(0xffffffffu, tmpa[mt][5], srcl);
                a4a[mt] = (q & 1) ? v1 : v0;
            }
#pragma unroll
            for (int kk = 0; kk < 5; kk++) {
#pragma unroll
                for (int jo = 0; jo < 3; jo++) {
                    const double b = B[(jo * 5 + kk) * 32];
#pragma unroll
                    for (int mt = 0; mt < DN_MT; mt++)
                        dmma(oa[mt][2 * jo], oa[mt][2 * jo + 1], kk < 4 ? tmpa[mt][kk] : a4a[mt], b);
                }
            }
#pragma unroll
            for (int mt = 0; mt < DN_MT; mt++) {
                double (&o)[6] = oa[mt];
                // |x| < 2^-256  <=>  high word of |x| < 0x2FF00000: the running maximum needs the high words only
                unsigned hmx = 0u;
#pragma unroll
                for (int k = 0; k < 6; k++) {
                    const unsigned h = (unsigned)__double2hiint(o[k]) & 0x7fffffffu;
                    hmx = h > hmx ? h : hmx;
                }
                unsigned long long mx = (unsigned long long)hmx << 32;
                int sc = 0;
                if (SAFE) {
                    mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
                    mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
                    sc = (Lsc0 ? 0 : scL[pc[mt] * ncm + c]) + (Rsc0 ? 0 : scR[pc[mt] * ncm + c]);
                    if (!tiptip && __longlong_as_double((long long)mx) < B200_SCALING_THRESHOLD && invar0[mt]) {
#pragma unroll
                        for (int k = 0; k < 6; k++) o[k] *= B200_TWO_256;
                        sc += 1;
                    }
                } else {
                    rmax[mt] = mx > rmax[mt] ? mx : rmax[mt];
                }
                if (live[mt]) {
                    double *po = out_base + offo[mt] + c * 20;
                    // streaming stores: the vector is next read a level later, after more than an L2 of other writes
                    // (cfg3 33.6 -> 32.7 ms, profiles/r1t_dense20_stream_ab.log; streaming LOADS of the children cost 5 %)
                    __stcs(reinterpret_cast<double2 *>(po), make_double2(o[0], o[1]));
                    __stcs(reinterpret_cast<double2 *>(po + 8), make_double2(o[2], o[3]));
                    if (q < 2) __stcs(reinterpret_cast<double2 *>(po + 16), make_double2(o[4], o[5]));
                    if (SAFE && q == 0) sc_out[pc[mt] * ncm + c] = (uint8_t)sc;
                }
            }
            if (!SAFE && cl == cg - 1) {
#pragma unroll
                for (int mt = 0; mt < DN_MT; mt++) {
                    unsigned long long mx = rmax[mt];
                    mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
                    mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
                    if (q == 0) {
                        unsigned long long *slot = &smax[t * DN_PT + warp * (8 * DN_MT) + mt * 8 + g];
                        if (mx > *slot) *slot = mx;
                    }
                }
            }
            if (last_cl) {
                cl = 0;
                t++;
            } else {
                cl++;
            }
        }
    }
    if (SAFE) return;
    __syncwarp();
    if (a.split) {
        for (int t = 0; t < a.tiles_per_cta; t++) {
            const int64_t pbase = (tile0 + t) * DN_PT + warp * (8 * DN_MT);
            if (pbase >= a.ntot) break;
#pragma unroll
            for (int mt = 0; mt < DN_MT; mt++) {
                const int64_t pp = pbase + mt * 8 + g;
                if (pp < a.ntot && q == 0)
                    atomicMax(&a.pmax[(size_t)blockIdx.y * a.ntot + pp], smax[t * DN_PT + warp * (8 * DN_MT) + mt * 8 + g]);
            }
        }
        return;
    }
    // NORM_LH decision per pattern.  Two phases, so that the scale counters of the children and the invariant-site flags
    // of ALL the warp's patterns are in flight together (one dependent load chain per pattern left 5 % of the stall
    // samples here); the children's counters are other slots than the one written.
    int scsum[DN_MAX_TILES][DN_MT];
    double inv[DN_MAX_TILES][DN_MT];
#pragma unroll
    for (int t = 0; t < DN_MAX_TILES; t++) {
#pragma unroll
        for (int mt = 0; mt < DN_MT; mt++) {
            const int64_t pp = (tile0 + t) * DN_PT + warp * (8 * DN_MT) + mt * 8 + g;
            const bool on = t < a.tiles_per_cta && pp < a.ntot;
            scsum[t][mt] = 0;
            inv[t][mt] = 1.0;
            if (on) {
                scsum[t][mt] = (Lsc0 ? 0 : (int)__ldg(scL + pp)) + (Rsc0 ? 0 : (int)__ldg(scR + pp));
                inv[t][mt] = __ldg(a.ptn_invar + pp);
            }
        }
    }
#pragma unroll
    for (int t = 0; t < DN_MAX_TILES; t++) {
#pragma unroll
        for (int mt = 0; mt < DN_MT; mt++) {
            const int64_t pp = (tile0 + t) * DN_PT + warp * (8 * DN_MT) + mt * 8 + g;
            if (!(t < a.tiles_per_cta && pp < a.ntot)) continue;
            const double mx = __longlong_as_double((long long)smax[t * DN_PT + warp * (8 * DN_MT) + mt * 8 + g]);
            const bool under = !tiptip && mx < B200_SCALING_THRESHOLD && inv[t][mt] == 0.0;
            if (under) {
                double *po = out_base + (size_t)pp * block;
                for (int e = q; e < block; e += 4) po[e] *= B200_TWO_256;
            }
            if (q == 0) sc_out[pp] = (uint8_t)(scsum[t][mt] + (under ? 1 : 0));
        }
    }
}

// NORM_LH decision of the split (mixture) launch: one thread per (node of the level, pattern)
__global__ void __launch_bounds__(256) dense20_fix_kernel(DenseArgs a) {
    const int64_t p = (int64_t)blockIdx.x * 256 + threadIdx.x;  // whole warps stay alive for the ballot below
    const bool valid = p < a.ntot;
    const DevNode nd = a.nodes[a.node_base + blockIdx.y];
    const DevChild L = a.children[nd.child_begin];
    const DevChild R = a.children[nd.child_begin + 1];
    const bool Lleaf = L.kind == B200LK_CHILD_LEAF, Rleaf = R.kind == B200LK_CHILD_LEAF;
    bool under = false;
    if (valid) {
        const double mx = __longlong_as_double((long long)a.pmax[(size_t)blockIdx.y * a.ntot + p]);
        under = !(Lleaf && Rleaf) && mx < B200_SCALING_THRESHOLD && a.ptn_invar[p] == 0.0;
        // a tip child and a virtual tip-tip child (vtt_off >= 0: nothing is in its slot) contribute no scale count
        const int sl = (Lleaf || L.vtt_off >= 0) ? 0 : a.scale[(size_t)L.id * a.slot_scale_bytes + p];
        const int sr = (Rleaf || R.vtt_off >= 0) ? 0 : a.scale[(size_t)R.id * a.slot_scale_bytes + p];
        a.scale[(size_t)nd.out_slot * a.slot_scale_bytes + p] = (uint8_t)(sl + sr + (under ? 1 : 0));
    }
    // underflowing patterns are rescaled by the whole warp, one pattern at a time (coalesced over the block)
    const int block = a.ncm * 20, lane = threadIdx.x & 31;
    unsigned todo = __ballot_sync(0xffffffffu, under);
    const int64_t p0 = p - lane;
    while (todo) {
        const int l = __ffs(todo) - 1;
        todo &= todo - 1;
        double *po = a.partial + (size_t)nd.out_slot * a.slot_doubles + (size_t)(p0 + l) * block;
        for (int e = lane; e < block; e += 32) po[e] *= B200_TWO_256;
    }
}

// Tip-tip nodes: gather the node's partial vector from its state-pair table (tables.cu).  Four lanes copy one
// pattern's block; scale counters of a tip-tip node are zero (:1640).
struct DenseTTArgs {
    const DevNode *nodes;
    const DevChild *children;
    const double *tt;
    const uint8_t *states;
    double *partial;
    uint8_t *scale;
    size_t slot_doubles, slot_scale_bytes;
    int64_t ntot, ntot_pad;
    int node_base, block, codes, scale_per_ptn;
};

__global__ void __launch_bounds__(256) dense20_tt_kernel(DenseTTArgs a) {
    const DevNode nd = a.nodes[a.node_base + blockIdx.y];
    const DevChild L = a.children[nd.child_begin], R = a.children[nd.child_begin + 1];
    const int q = threadIdx.x & 3;
    const int64_t p = ((int64_t)blockIdx.x * 256 + threadIdx.x) >> 2;
    if (p >= a.ntot) return;
    const int sl = a.states[(size_t)L.id * a.ntot_pad + p], sr = a.states[(size_t)R.id * a.ntot_pad + p];
    const double2 *src = reinterpret_cast<const double2 *>(a.tt + L.aux_off + (size_t)(sl * a.codes + sr) * a.block);
    double2 *dst = reinterpret_cast<double2 *>(a.partial + (size_t)nd.out_slot * a.slot_doubles + (size_t)p * a.block);
    const int pairs = a.block >> 1;
#pragma unroll 5
    for (int e = q; e < pairs; e += 4) __stcs(dst + e, __ldg(src + e));
    uint8_t *sc = a.scale + (size_t)nd.out_slot * a.slot_scale_bytes + (size_t)p * a.scale_per_ptn;
    for (int k = q; k < a.scale_per_ptn; k += 4) sc[k] = 0;
}

// ---- virtual tip-tip nodes (api.cu decides which) ------------------------------------------------------------------
// pair_idx[v][p] = state(ta, p) * codes + state(tb, p): what the reader of a virtual node loads per row instead of the row's
// position in a resident vector (4 bytes of traffic per pattern and node instead of the 2 * 8 * block of gather + re-read)
#define DN_VIRT_PER_LAUNCH 120
struct TTPairArgs {
    const uint8_t *states;
    uint16_t *pair_idx;
    int64_t ntot, ntot_pad;
    int codes;
    S4Virt v[DN_VIRT_PER_LAUNCH];
    int64_t off[DN_VIRT_PER_LAUNCH];
};
__global__ void __launch_bounds__(256) tt_pair_kernel(const __grid_constant__ TTPairArgs a) {
    const int64_t p = (int64_t)blockIdx.x * 256 + threadIdx.x;
    if (p >= a.ntot) return;
    const S4Virt &v = a.v[blockIdx.y];
    a.pair_idx[a.off[blockIdx.y] + p] =
        (uint16_t)(a.states[(size_t)v.ta * a.ntot_pad + p] * a.codes + a.states[(size_t)v.tb * a.ntot_pad + p]);
}

int launch_tt_pairs(b200lk_ctx *c, const S4Virt *v, int n, const int64_t *pair_off) {
    const size_t need = (size_t)n * c->ntot_pad;
    if (need > c->pair_cap) {  // grows only (cudaFree waits for the device)
        const size_t want = std::max(need, (size_t)(c->d.ntaxa / 2 + 1) * c->ntot_pad);
        if (c->pair_idx) CUDA_TRY(c, cudaFree(c->pair_idx));
        c->pair_idx = nullptr;
        c->pair_cap = 0;
        CUDA_TRY(c, cudaMalloc(&c->pair_idx, sizeof(uint16_t) * want));
        c->pair_cap = want;
    }
    static thread_local TTPairArgs *Ap = nullptr;
    if (!Ap) Ap = new TTPairArgs();
    TTPairArgs &a = *Ap;
    a.states = c->states;
    a.pair_idx = c->pair_idx;
    a.ntot = c->ntot;
    a.ntot_pad = c->ntot_pad;
    a.codes = c->d.n_state_codes;
    for (int done = 0; done < n; done += DN_VIRT_PER_LAUNCH) {
        const int m = std::min(DN_VIRT_PER_LAUNCH, n - done);
        for (int k = 0; k < m; k++) {
            a.v[k] = v[done + k];
            a.off[k] = pair_off[done + k];
        }
        tt_pair_kernel<<<dim3((unsigned)((c->ntot + 255) / 256), m), 256, 0, c->stream>>>(a);
        CUDA_TRY(c, cudaGetLastError());
        c->n_launches++;
    }
    return B200LK_OK;
}

// s4_realize for the pattern-major layout: every virtual slot gathered from its table (the rows the update would have stored)
struct DenseRealizeArgs {
    const double *tt;
    const uint8_t *states;
    double *partial;
    uint8_t *scale;
    size_t slot_doubles, slot_scale_bytes;
    int64_t ntot, ntot_pad;
    int block, codes, scale_per_ptn;
    S4Virt v[DN_VIRT_PER_LAUNCH];
};
__global__ void __launch_bounds__(256) dense_realize_kernel(const __grid_constant__ DenseRealizeArgs a) {
    const S4Virt &v = a.v[blockIdx.y];
    const int q = threadIdx.x & 3;
    const int64_t p = ((int64_t)blockIdx.x * 256 + threadIdx.x) >> 2;  // four lanes copy one pattern's block
    if (p >= a.ntot) return;
    const int sl = a.states[(size_t)v.ta * a.ntot_pad + p], sr = a.states[(size_t)v.tb * a.ntot_pad + p];
    const double2 *src = reinterpret_cast<const double2 *>(a.tt + (size_t)v.tt_off + (size_t)(sl * a.codes + sr) * a.block);
    double2 *dst = reinterpret_cast<double2 *>(a.partial + (size_t)v.slot * a.slot_doubles + (size_t)p * a.block);
    const int pairs = a.block >> 1;
    for (int e = q; e < pairs; e += 4) __stcs(dst + e, __ldg(src + e));
    uint8_t *sc = a.scale + (size_t)v.slot * a.slot_scale_bytes + (size_t)p * a.scale_per_ptn;
    for (int k = q; k < a.scale_per_ptn; k += 4) sc[k] = 0;
}

int dense_realize(b200lk_ctx *c) {
    if (!c->n_virtual) return B200LK_OK;
    static thread_local DenseRealizeArgs *Ap = nullptr;
    if (!Ap) Ap = new DenseRealizeArgs();
    DenseRealizeArgs &a = *Ap;
    a.tt = c->tt_tables;
    a.states = c->states;
    a.partial = c->partial;
    a.scale = c->scale;
    a.slot_doubles = c->slot_doubles;
    a.slot_scale_bytes = c->slot_scale_bytes;
    a.ntot = c->ntot;
    a.ntot_pad = c->ntot_pad;
    a.block = c->block;
    a.codes = c->d.n_state_codes;
    a.scale_per_ptn = c->scale_per_ptn;
    int n = 0, total = 0;
    auto flush = [&]() -> int {
        if (!n) return B200LK_OK;
        dense_realize_kernel<<<dim3((unsigned)((c->ntot * 4 + 255) / 256), n), 256, 0, c->stream>>>(a);
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
        if (n == DN_VIRT_PER_LAUNCH) {
            int rc = flush();
            if (rc) return rc;
        }
    }
    int rc = flush();
    if (rc) return rc;
    c->n_virtual = 0;
    c->n_min_bytes += c->ntot * (8 * (int64_t)c->block + c->scale_per_ptn) * total;
    return B200LK_OK;
}

int launch_dense20_tt(b200lk_ctx *c, int node_base, int nnodes) {
    DenseTTArgs a;
    a.nodes = c->d_nodes;
    a.children = c->d_children;
    a.tt = c->tt_tables;
    a.states = c->states;
    a.partial = c->partial;
    a.scale = c->scale;
    a.slot_doubles = c->slot_doubles;
    a.slot_scale_bytes = c->slot_scale_bytes;
    a.ntot = c->ntot;
    a.ntot_pad = c->ntot_pad;
    a.node_base = node_base;
    a.block = c->block;
    a.codes = c->d.n_state_codes;
    a.scale_per_ptn = c->scale_per_ptn;
    prof_mark(c);
    dense20_tt_kernel<<<dim3((unsigned)((c->ntot * 4 + 255) / 256), nnodes), 256, 0, c->stream>>>(a);
    prof_mark(c);
    CUDA_TRY(c, cudaGetLastError());
    c->n_launches++;
    return B200LK_OK;
}

// bifurcating nodes [node_base, node_base + nnodes) of the level-sorted device node array
int launch_dense20(b200lk_ctx *c, int node_base, int nnodes) {
    DenseArgs a;
    a.nodes = c->d_nodes;
    a.children = c->d_children;
    a.tables = c->tables;
    a.inv_frag = c->m.inv_frag;
    a.ptn_invar = c->ptn_invar;
    a.tt = c->tt_tables;
    a.pair_idx = c->pair_idx;
    a.states = c->states;
    a.partial = c->partial;
    a.scale = c->scale;
    a.slot_doubles = c->slot_doubles;
    a.slot_scale_bytes = c->slot_scale_bytes;
    a.ntot = c->ntot;
    a.ntot_pad = c->ntot_pad;
    a.node_base = node_base;
    a.ncm = c->ncat_mix;
    a.denom = c->d.fused_mix_rate ? 1 : c->d.ncat;
    a.codes = c->d.n_state_codes;
    const int64_t tiles = ((c->batch ? c->batch->max_ntot : c->ntot) + DN_PT - 1) / DN_PT;
    // enough CTAs to fill the machine a few times over, but amortise the table staging over several tiles
    const int ngroups = (a.ncm + DN_CG - 1) / DN_CG;
    a.split = ngroups > 1;
    if (c->batch && a.split) return B200LK_ERR_UNSUPPORTED;  // grid.z is taken by the category groups
    a.pmax = nullptr;
    int tpc = DN_MAX_TILES;
    while (tpc > 1 && (tiles + tpc - 1) / tpc * nnodes * (a.split ? ngroups : 1) * (c->batch ? c->batch->n : 1) < c->sms * 6) tpc >>= 1;
    a.tiles_per_cta = tpc;
    if (a.split && !c->d.safe_numeric) {
        const size_t need = sizeof(unsigned long long) * (size_t)nnodes * c->ntot;
        if (need > c->pmax_bytes) {
            if (c->pmax) CUDA_TRY(c, cudaFree(c->pmax));
            c->pmax = nullptr;
            c->pmax_bytes = 0;
            CUDA_TRY(c, cudaMalloc(&c->pmax, need));
            c->pmax_bytes = need;
        }
        CUDA_TRY(c, cudaMemsetAsync(c->pmax, 0, need, c->stream));
        a.pmax = c->pmax;
    }
    const int leaf_doubles = a.codes * DN_CG * 20, frag_doubles = DN_CG * B200_FRAG20;
    const size_t smem = sizeof(double) * (2 * (size_t)std::max(leaf_doubles, frag_doubles) + frag_doubles) +
                        sizeof(unsigned long long) * tpc * DN_PT;
    const int pipe = b200_env().dense_pipe;  // A/B switch: 0 direct loads, 1 cp.async double buffer, 2 register prefetch (default)
    const size_t smem_pipe = smem + sizeof(double) * (DN_THREADS / 32) * 2 * DN_STAGE_DOUBLES;
    const int occ = b200_env().dense_occ;  // A/B switch: resident CTAs per SM the kernel is compiled for (registers 160 / 128)
    if (!c->dense_attr_set) {
        CUDA_TRY(c, cudaFuncSetAttribute(dense20_kernel<false, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
        CUDA_TRY(c, cudaFuncSetAttribute(dense20_kernel<true, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
        CUDA_TRY(c, cudaFuncSetAttribute(dense20_kernel<false, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
        CUDA_TRY(c, cudaFuncSetAttribute(dense20_kernel<true, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
        CUDA_TRY(c, cudaFuncSetAttribute(dense20_pipe_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 110 * 1024));
        CUDA_TRY(c, cudaFuncSetAttribute(dense20_pipe_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 110 * 1024));
        CUDA_TRY(c, cudaFuncSetAttribute(dense20_rp_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
        CUDA_TRY(c, cudaFuncSetAttribute(dense20_rp_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
        CUDA_TRY(c, cudaFuncSetAttribute(dense20_rp_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
        CUDA_TRY(c, cudaFuncSetAttribute(dense20_rp_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
        c->dense_attr_set = true;
    }
    dim3 grid((unsigned)((tiles + tpc - 1) / tpc), nnodes, a.split ? ngroups : (c->batch ? c->batch->n : 1));
    BatchLaunch bl;
    memset(&bl, 0, sizeof(bl));
    if (c->batch) bl = *c->batch;
    prof_mark(c);
    if (pipe == 2 || c->batch) {
        bool any_virt = false;  // does a node of this launch read a virtual tip-tip child?
        for (int i = node_base; i < node_base + nnodes && !any_virt && !c->batch; i++)
            for (int k = 0; k < 2; k++) any_virt |= c->h_children[c->h_nodes[i].child_begin + k].vtt_off >= 0;
        if (any_virt) {
            if (c->d.safe_numeric) dense20_rp_kernel<true, true><<<grid, DN_THREADS, smem, c->stream>>>(a, bl);
            else dense20_rp_kernel<false, true><<<grid, DN_THREADS, smem, c->stream>>>(a, bl);
        } else {
            if (c->d.safe_numeric) dense20_rp_kernel<true, false><<<grid, DN_THREADS, smem, c->stream>>>(a, bl);
            else dense20_rp_kernel<false, false><<<grid, DN_THREADS, smem, c->stream>>>(a, bl);
        }
    } else if (pipe == 1) {
        if (c->d.safe_numeric) dense20_pipe_kernel<true><<<grid, DN_THREADS, smem_pipe, c->stream>>>(a);
        else dense20_pipe_kernel<false><<<grid, DN_THREADS, smem_pipe, c->stream>>>(a);
    } else if (occ == 3) {
        if (c->d.safe_numeric) dense20_kernel<true, 3><<<grid, DN_THREADS, smem, c->stream>>>(a);
        else dense20_kernel<false, 3><<<grid, DN_THREADS, smem, c->stream>>>(a);
    } else {
        if (c->d.safe_numeric) dense20_kernel<true, 4><<<grid, DN_THREADS, smem, c->stream>>>(a);
        else dense20_kernel<false, 4><<<grid, DN_THREADS, smem, c->stream>>>(a);
    }
    if (a.split && !c->d.safe_numeric) {
        dense20_fix_kernel<<<dim3((unsigned)((c->ntot + 255) / 256), nnodes), 256, 0, c->stream>>>(a);
        c->n_launches++;
    }
    prof_mark(c);
    CUDA_TRY(c, cudaGetLastError());
    c->n_launches++;
    return B200LK_OK;
}
