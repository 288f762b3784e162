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
#define DN_MT 2                                   // row tiles (8 patterns) per warp
#define DN_PT (DN_THREADS / 32 * 8 * DN_MT)       // 64 patterns per CTA tile
#define DN_CG 4                                   // categories per staged group
#define DN_MAX_TILES 8

struct DenseArgs {
    const DevNode *nodes;
    const DevChild *children;
    const double *tables, *inv_frag, *ptn_invar;
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
            unsigned long long rmax[DN_MT] = {0ull, 0ull};
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

// NORM_LH decision of the split (mixture) launch: one thread per (node of the level, pattern)
__global__ void __launch_bounds__(256) dense20_fix_kernel(DenseArgs a) {
    const int64_t p = (int64_t)blockIdx.x * 256 + threadIdx.x;
    if (p >= a.ntot) return;
    const DevNode nd = a.nodes[a.node_base + blockIdx.y];
    const DevChild L = a.children[nd.child_begin];
    const DevChild R = a.children[nd.child_begin + 1];
    const bool Lleaf = L.kind == B200LK_CHILD_LEAF, Rleaf = R.kind == B200LK_CHILD_LEAF;
    const double mx = __longlong_as_double((long long)a.pmax[(size_t)blockIdx.y * a.ntot + p]);
    const bool under = !(Lleaf && Rleaf) && mx < B200_SCALING_THRESHOLD && a.ptn_invar[p] == 0.0;
    const int block = a.ncm * 20;
    if (under) {
        double *po = a.partial + (size_t)nd.out_slot * a.slot_doubles + (size_t)p * block;
        for (int e = 0; e < block; e++) po[e] *= B200_TWO_256;
    }
    const int sl = Lleaf ? 0 : a.scale[(size_t)L.id * a.slot_scale_bytes + p];
    const int sr = Rleaf ? 0 : a.scale[(size_t)R.id * a.slot_scale_bytes + p];
    a.scale[(size_t)nd.out_slot * a.slot_scale_bytes + p] = (uint8_t)(sl + sr + (under ? 1 : 0));
}

// bifurcating nodes [node_base, node_base + nnodes) of the level-sorted device node array
int launch_dense20(b200lk_ctx *c, int node_base, int nnodes) {
    DenseArgs a;
    a.nodes = c->d_nodes;
    a.children = c->d_children;
    a.tables = c->tables;
    a.inv_frag = c->m.inv_frag;
    a.ptn_invar = c->ptn_invar;
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
    const int64_t tiles = (c->ntot + DN_PT - 1) / DN_PT;
    // enough CTAs to fill the machine a few times over, but amortise the table staging over several tiles
    const int ngroups = (a.ncm + DN_CG - 1) / DN_CG;
    a.split = ngroups > 1;
    a.pmax = nullptr;
    int tpc = DN_MAX_TILES;
    while (tpc > 1 && (tiles + tpc - 1) / tpc * nnodes * (a.split ? ngroups : 1) < 148 * 6) tpc >>= 1;
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
    static int occ = -1;  // A/B switch: resident CTAs per SM the kernel is compiled for (registers 160 / 128)
    if (occ < 0) {
        const char *e = getenv("B200LK_DENSE_OCC");
        occ = (e && atoi(e) == 3) ? 3 : 4;
    }
    if (!c->dense_attr_set) {
        CUDA_TRY(c, cudaFuncSetAttribute(dense20_kernel<false, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
        CUDA_TRY(c, cudaFuncSetAttribute(dense20_kernel<true, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
        CUDA_TRY(c, cudaFuncSetAttribute(dense20_kernel<false, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
        CUDA_TRY(c, cudaFuncSetAttribute(dense20_kernel<true, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
        c->dense_attr_set = true;
    }
    dim3 grid((unsigned)((tiles + tpc - 1) / tpc), nnodes, a.split ? ngroups : 1);
    prof_mark(c);
    if (occ == 3) {
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
