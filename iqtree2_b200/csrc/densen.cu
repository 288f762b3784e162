// densen.cu -- partial-likelihood update for dense state spaces other than 20 (codon 61, morphological, ...) on the
// FP64 tensor cores (mma.sync m8n8k4), level-batched.  Replaces computePartialLikelihoodGenericSIMD
// (tree/phylokernelnew.h:1291-2039, bifurcating cases :1632-2032) for 9 <= nstates <= 64.
//
// States are padded to NP = 8*NT.  Per (pattern tile, category) a warp forms three products on the tensor cores:
//   F_ch[p][x] = sum_i A_ch[p][c][i] * E_ch[c][x][i]      (each internal child; a leaf child is a table lookup by state)
//   out[p][i]  = sum_x (F_L * F_R)[p][x] * Uinv[i][x]
// The accumulator fragment of the child products (lane (g,q) holds x = 8j+2q, 8j+2q+1 of row g) is used directly as
// the A operand of the last product: k-step kk = 2j+b contracts over x = 8j + 2q + b, and the inverse eigenvectors
// are laid out in that permuted order (tables.cu, inv_fragN_kernel), so nothing is shuffled between the products.
// Tables of ONE category are staged in shared memory at a time (3 x NT*2NT*32 doubles: 96 KB for codons), categories
// outermost, so the NORM_LH per-pattern maximum over categories is kept in shared memory per pattern of the CTA.
#include <algorithm>

#include "common.cuh"

#define DX_THREADS 128
#define DX_MAX_TILES 16

struct DenseNArgs {
    const DevNode *nodes;
    const DevChild *children;
    const double *tables, *inv_frag, *ptn_invar;
    const uint8_t *states;
    double *partial;
    uint8_t *scale;
    size_t slot_doubles, slot_scale_bytes;
    int64_t ntot, ntot_pad;
    int node_base, n, ncm, denom, codes, tiles_per_cta, child_doubles;
};

__device__ __forceinline__ void dmma_n(double &d0, double &d1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(d0), "+d"(d1)
                 : "d"(a), "d"(b));
}

template <int NT, int MT, bool SAFE>
__global__ void __launch_bounds__(DX_THREADS) denseN_kernel(DenseNArgs a) {
    constexpr int NP = 8 * NT, KS = 2 * NT, FRAG = NT * KS * 32, LP = NP + 2, PT = (DX_THREADS / 32) * 8 * MT;
    constexpr int JG = NT < 4 ? NT : 4;
    extern __shared__ __align__(16) double sm[];
    const int n = a.n, ncm = a.ncm, block = ncm * n;
    const DevNode nd = a.nodes[a.node_base + blockIdx.y];
    const DevChild L = a.children[nd.child_begin];
    const DevChild R = a.children[nd.child_begin + 1];
    const bool Lleaf = L.kind == B200LK_CHILD_LEAF, Rleaf = R.kind == B200LK_CHILD_LEAF;
    const bool tiptip = Lleaf && Rleaf;
    double *sL = sm, *sR = sL + a.child_doubles, *sI = sR + a.child_doubles;
    unsigned long long *smax = reinterpret_cast<unsigned long long *>(sI + FRAG);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, q = lane & 3;
    const int64_t tile0 = (int64_t)blockIdx.x * a.tiles_per_cta;
    double *out_base = a.partial + (size_t)nd.out_slot * a.slot_doubles;
    uint8_t *sc_out = a.scale + (size_t)nd.out_slot * a.slot_scale_bytes;
    const uint8_t *scL = Lleaf ? nullptr : a.scale + (size_t)L.id * a.slot_scale_bytes;
    const uint8_t *scR = Rleaf ? nullptr : a.scale + (size_t)R.id * a.slot_scale_bytes;
    const double *pL = Lleaf ? nullptr : a.partial + (size_t)L.id * a.slot_doubles;
    const double *pR = Rleaf ? nullptr : a.partial + (size_t)R.id * a.slot_doubles;
    if (!SAFE)
        for (int i = tid; i < a.tiles_per_cta * PT; i += DX_THREADS) smax[i] = 0ull;
    int my_tiles = 0;
    for (int t = 0; t < a.tiles_per_cta; t++)
        if ((tile0 + t) * PT + warp * (8 * MT) < a.ntot) my_tiles = t + 1;

    for (int c = 0; c < ncm; c++) {
        __syncthreads();
        // stage this category's tables: a leaf child's lookup rows [code][x] (padded with zeros, row stride LP keeps
        // rows of different codes on different banks), an internal child's P-factor fragments
        if (Lleaf) {
            const double *src = a.tables + L.leaf_off + (size_t)c * n;
            for (int i = tid; i < a.codes * LP; i += DX_THREADS) {
                const int s = i / LP, x = i - s * LP;
                sL[i] = x < n ? src[(size_t)s * block + x] : 0.0;
            }
        } else {
            const double *src = a.tables + L.frag_off + (size_t)c * FRAG;
            for (int i = tid; i < FRAG; i += DX_THREADS) sL[i] = src[i];
        }
        if (Rleaf) {
            const double *src = a.tables + R.leaf_off + (size_t)c * n;
            for (int i = tid; i < a.codes * LP; i += DX_THREADS) {
                const int s = i / LP, x = i - s * LP;
                sR[i] = x < n ? src[(size_t)s * block + x] : 0.0;
            }
        } else {
            const double *src = a.tables + R.frag_off + (size_t)c * FRAG;
            for (int i = tid; i < FRAG; i += DX_THREADS) sR[i] = src[i];
        }
        {
            const double *src = a.inv_frag + (size_t)(c / a.denom) * FRAG;
            for (int i = tid; i < FRAG; i += DX_THREADS) sI[i] = src[i];
        }
        __syncthreads();

        for (int t = 0; t < my_tiles; t++) {
            int64_t p[MT], pc[MT];
#pragma unroll
            for (int mt = 0; mt < MT; mt++) {
                p[mt] = (tile0 + t) * PT + warp * (8 * MT) + mt * 8 + g;
                pc[mt] = p[mt] < a.ntot ? p[mt] : a.ntot - 1;  // rows past the end repeat the last pattern, never stored
            }
            double f[MT][2 * NT];
            // ---- left child
            if (Lleaf) {
#pragma unroll
                for (int mt = 0; mt < MT; mt++) {
                    const double *row = sL + (size_t)a.states[(size_t)L.id * a.ntot_pad + pc[mt]] * LP + 2 * q;
#pragma unroll
                    for (int j = 0; j < NT; j++) {
                        const double2 v = *reinterpret_cast<const double2 *>(row + 8 * j);
                        f[mt][2 * j] = v.x;
                        f[mt][2 * j + 1] = v.y;
                    }
                }
            } else {
                double A[MT][KS];
#pragma unroll
                for (int mt = 0; mt < MT; mt++) {
                    const double *src = pL + (size_t)pc[mt] * block + (size_t)c * n + q;
#pragma unroll
                    for (int s = 0; s < KS; s++) A[mt][s] = (4 * s + q < n) ? __ldg(src + 4 * s) : 0.0;
#pragma unroll
                    for (int k = 0; k < 2 * NT; k++) f[mt][k] = 0.0;
                }
                const double *B = sL + lane;
                // k-step outermost: the NT accumulator chains of a row tile are independent of each other
#pragma unroll
                for (int s = 0; s < KS; s++)
#pragma unroll
                    for (int j = 0; j < NT; j++) {
                        const double b = B[(j * KS + s) * 32];
#pragma unroll
                        for (int mt = 0; mt < MT; mt++) dmma_n(f[mt][2 * j], f[mt][2 * j + 1], A[mt][s], b);
                    }
            }
            // ---- right child, multiplied into the left factor n-tile by n-tile
            if (Rleaf) {
#pragma unroll
                for (int mt = 0; mt < MT; mt++) {
                    const double *row = sR + (size_t)a.states[(size_t)R.id * a.ntot_pad + pc[mt]] * LP + 2 * q;
#pragma unroll
                    for (int j = 0; j < NT; j++) {
                        const double2 v = *reinterpret_cast<const double2 *>(row + 8 * j);
                        f[mt][2 * j] *= v.x;
                        f[mt][2 * j + 1] *= v.y;
                    }
                }
            } else {
                double A[MT][KS];
#pragma unroll
                for (int mt = 0; mt < MT; mt++) {
                    const double *src = pR + (size_t)pc[mt] * block + (size_t)c * n + q;
#pragma unroll
                    for (int s = 0; s < KS; s++) A[mt][s] = (4 * s + q < n) ? __ldg(src + 4 * s) : 0.0;
                }
                const double *B = sR + lane;
                // n-tiles in groups of JG: JG independent chains in flight, JG*MT*2 extra accumulators live
#pragma unroll
                for (int j0 = 0; j0 < NT; j0 += JG) {
                    double r[MT][2 * JG];
#pragma unroll
                    for (int mt = 0; mt < MT; mt++)
#pragma unroll
                        for (int k = 0; k < 2 * JG; k++) r[mt][k] = 0.0;
#pragma unroll
                    for (int s = 0; s < KS; s++)
#pragma unroll
                        for (int jj = 0; jj < JG; jj++) {
                            if (j0 + jj < NT) {
                                const double b = B[((j0 + jj) * KS + s) * 32];
#pragma unroll
                                for (int mt = 0; mt < MT; mt++) dmma_n(r[mt][2 * jj], r[mt][2 * jj + 1], A[mt][s], b);
                            }
                        }
#pragma unroll
                    for (int mt = 0; mt < MT; mt++)
#pragma unroll
                        for (int jj = 0; jj < JG; jj++)
                            if (j0 + jj < NT) {
                                f[mt][2 * (j0 + jj)] *= r[mt][2 * jj];
                                f[mt][2 * (j0 + jj) + 1] *= r[mt][2 * jj + 1];
                            }
                }
            }
            // ---- back to eigen-space, scaling test, store
            const double *BI = sI + lane;
#pragma unroll
            for (int mt = 0; mt < MT; mt++) {
                double o[2 * NT];
                unsigned hmx = 0u;
#pragma unroll
                for (int k = 0; k < 2 * NT; k++) o[k] = 0.0;
#pragma unroll
                for (int kk = 0; kk < KS; kk++)
#pragma unroll
                    for (int jo = 0; jo < NT; jo++) dmma_n(o[2 * jo], o[2 * jo + 1], f[mt][kk], BI[(jo * KS + kk) * 32]);
                // |x| < 2^-256  <=>  high word of |x| < 0x2FF00000: the maximum needs the high words only
#pragma unroll
                for (int k = 0; k < 2 * NT; k++) hmx = max(hmx, (unsigned)__double2hiint(o[k]) & 0x7fffffffu);
                hmx = max(hmx, __shfl_xor_sync(0xffffffffu, hmx, 1));
                hmx = max(hmx, __shfl_xor_sync(0xffffffffu, hmx, 2));
                const bool live = p[mt] < a.ntot;
                if (SAFE) {
                    int sc = (Lleaf ? 0 : scL[pc[mt] * ncm + c]) + (Rleaf ? 0 : scR[pc[mt] * ncm + c]);
                    if (!tiptip && hmx < 0x2FF00000u && a.ptn_invar[pc[mt]] == 0.0) {
#pragma unroll
                        for (int k = 0; k < 2 * NT; k++) o[k] *= B200_TWO_256;
                        sc += 1;
                    }
                    if (live && q == 0) sc_out[p[mt] * ncm + c] = (uint8_t)sc;
                } else if (q == 0) {
                    unsigned long long *slot = &smax[t * PT + warp * (8 * MT) + mt * 8 + g];
                    const unsigned long long mx = (unsigned long long)hmx << 32;
                    if (mx > *slot) *slot = mx;
                }
                if (live) {
                    double *po = out_base + (size_t)p[mt] * block + (size_t)c * n + 2 * q;
#pragma unroll
                    for (int jo = 0; jo < NT; jo++) {
                        if (8 * jo + 2 * q < n) po[8 * jo] = o[2 * jo];
                        if (8 * jo + 2 * q + 1 < n) po[8 * jo + 1] = o[2 * jo + 1];
                    }
                }
            }
        }
    }
    if (SAFE) return;
    // NORM_LH: one decision per pattern over all categories (:1960-2010); the warp that wrote a pattern's block
    // rescales it (the block is still in L2), four lanes per pattern
    __syncwarp();
    for (int t = 0; t < my_tiles; t++) {
#pragma unroll
        for (int mt = 0; mt < MT; mt++) {
            const int64_t pp = (tile0 + t) * PT + warp * (8 * MT) + mt * 8 + g;
            if (pp >= a.ntot) continue;
            const unsigned hmx = (unsigned)(smax[t * PT + warp * (8 * MT) + mt * 8 + g] >> 32);
            const bool under = !tiptip && hmx < 0x2FF00000u && a.ptn_invar[pp] == 0.0;
            if (under) {
                double *po = out_base + (size_t)pp * block;
                for (int e = q; e < block; e += 4) po[e] *= B200_TWO_256;
            }
            if (q == 0) sc_out[pp] = (uint8_t)((Lleaf ? 0 : scL[pp]) + (Rleaf ? 0 : scR[pp]) + (under ? 1 : 0));
        }
    }
}

template <int NT, int MT>
static int launch_nt(b200lk_ctx *c, DenseNArgs &a, int nnodes) {
    constexpr int NP = 8 * NT, FRAG = NT * 2 * NT * 32, LP = NP + 2, PT = (DX_THREADS / 32) * 8 * MT;
    a.child_doubles = std::max(a.codes * LP, FRAG);
    const int64_t tiles = (c->ntot + PT - 1) / PT;
    int tpc = DX_MAX_TILES;  // amortise the table staging, but keep enough CTAs to fill the machine a few times
    while (tpc > 1 && (tiles + tpc - 1) / tpc * nnodes < c->sms * 4) tpc >>= 1;
    a.tiles_per_cta = tpc;
    const size_t smem = sizeof(double) * (2 * (size_t)a.child_doubles + FRAG) + sizeof(unsigned long long) * tpc * PT;
    if (smem > 200 * 1024) return B200LK_ERR_UNSUPPORTED;
    auto kern = c->d.safe_numeric ? denseN_kernel<NT, MT, true> : denseN_kernel<NT, MT, false>;
    CUDA_TRY(c, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    dim3 grid((unsigned)((tiles + tpc - 1) / tpc), nnodes);
    prof_mark(c);
    kern<<<grid, DX_THREADS, smem, c->stream>>>(a);
    prof_mark(c);
    CUDA_TRY(c, cudaGetLastError());
    c->n_launches++;
    return B200LK_OK;
}

int launch_denseN(b200lk_ctx *c, int node_base, int nnodes) {
    DenseNArgs a;
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
    a.n = c->d.nstates;
    a.ncm = c->ncat_mix;
    a.denom = c->d.fused_mix_rate ? 1 : c->d.ncat;
    a.codes = c->d.n_state_codes;
    switch (c->dense_nt) {
        case 2: return launch_nt<2, 2>(c, a, nnodes);
        case 3: return launch_nt<3, 2>(c, a, nnodes);
        case 4: return launch_nt<4, 2>(c, a, nnodes);
        case 5: return launch_nt<5, 1>(c, a, nnodes);
        case 6: return launch_nt<6, 1>(c, a, nnodes);
        case 7: return launch_nt<7, 1>(c, a, nnodes);
        case 8: return launch_nt<8, 1>(c, a, nnodes);
    }
    return B200LK_ERR_UNSUPPORTED;
}
