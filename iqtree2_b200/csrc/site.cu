// site.cu -- site-specific models (ModelSet / PMSF): every pattern has its own eigen-system.  Replaces the
// SITE_MODEL = true branches of computePartialLikelihoodGenericSIMD (tree/phylokernelnew.h: tip-tip :1653-1680,
// tip-internal :1772-1800, internal-internal :1952-1968) and the site-specific part of computeTipPartialLikelihood
// (tree/phylotreesse.cpp:257-340).  SURVEY.md section 8f, rank 3.
//
//   out[p][c][i] = sum_x Uinv_p[i][x] * ( sum_j U_p[x][j] e^{lambda_p[j] r_c t_L} A_L[p][c][j] )
//                                     * ( sum_j U_p[x][j] e^{lambda_p[j] r_c t_R} A_R[p][c][j] )
// with a leaf child's A = tip[taxon][p][.] (the same vector for every category).  There are no P-factor tables to
// share between patterns, so the update is a batch of per-pattern mat-vecs: one warp per pattern, lane = state;
// the pattern's U and Uinv (2 n^2 doubles, contiguous) are read once per node and kept in registers across the
// categories.  Bound: HBM (6.4 KB of eigen-system per pattern-node for 20 states against 1.9 KB of partials).
#include <stdlib.h>

#include "common.cuh"

#define SITE_THREADS 128

struct SiteArgs {
    const DevNode *nodes;
    const DevChild *children;
    const double *eval, *evec, *inv, *tip, *rates, *ptn_invar;
    double *partial;
    uint8_t *scale;
    size_t slot_doubles, slot_scale_bytes;
    int64_t nptn;
    int node_base, n, ncat;
};

// tip[t][p][i] = sum_x inv_p[i][x] * tipvec[state(t,p)][x]: the plain state's column, the row sum for UNKNOWN, the
// sum over the ambiguity set otherwise
__global__ void __launch_bounds__(256) site_tip_kernel(const uint8_t *__restrict__ states, const double *__restrict__ inv,
                                                        const double *__restrict__ tipvec, double *__restrict__ tip,
                                                        int64_t nptn, int64_t ntot_pad, int n, int ntaxa) {
    const int64_t total = (int64_t)ntaxa * nptn * n;
    for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
        const int i = (int)(idx % n);
        const int64_t tp = idx / n, p = tp % nptn, t = tp / nptn;
        const double *tv = tipvec + (size_t)states[(size_t)t * ntot_pad + p] * n;
        const double *row = inv + ((size_t)p * n + i) * n;
        double v = 0.0;
        for (int x = 0; x < n; x++)
            if (tv[x] != 0.0) v += row[x];
        tip[idx] = v;
    }
}

template <int NMAX, bool SAFE>
__global__ void __launch_bounds__(SITE_THREADS) site_partial_kernel(SiteArgs a) {
    const int n = a.n, ncat = a.ncat, block = n * ncat;
    const DevNode nd = a.nodes[a.node_base + blockIdx.y];
    const DevChild L = a.children[nd.child_begin], R = a.children[nd.child_begin + 1];  // a single leaf comes first
    const bool Lleaf = L.kind == B200LK_CHILD_LEAF, Rleaf = R.kind == B200LK_CHILD_LEAF, tiptip = Lleaf && Rleaf;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const bool act = lane < n;
    double *out = a.partial + (size_t)nd.out_slot * a.slot_doubles;
    uint8_t *sc_out = a.scale + (size_t)nd.out_slot * a.slot_scale_bytes;
    const double *pL = Lleaf ? a.tip + (size_t)L.id * a.nptn * n : a.partial + (size_t)L.id * a.slot_doubles;
    const double *pR = Rleaf ? a.tip + (size_t)R.id * a.nptn * n : a.partial + (size_t)R.id * a.slot_doubles;
    const uint8_t *scL = Lleaf ? nullptr : a.scale + (size_t)L.id * a.slot_scale_bytes;
    const uint8_t *scR = Rleaf ? nullptr : a.scale + (size_t)R.id * a.slot_scale_bytes;
    for (int64_t p = (int64_t)blockIdx.x * (SITE_THREADS / 32) + warp; p < a.nptn; p += (int64_t)gridDim.x * (SITE_THREADS / 32)) {
        // lane x: row x of U_p; lane i: row i of Uinv_p and eigenvalue i
        double U[NMAX], M[NMAX];
        const double *Up = a.evec + (size_t)p * n * n + (size_t)lane * n, *Mp = a.inv + (size_t)p * n * n + (size_t)lane * n;
#pragma unroll
        for (int j = 0; j < NMAX; j++) {
            U[j] = (act && j < n) ? __ldg(Up + j) : 0.0;
            M[j] = (act && j < n) ? __ldg(Mp + j) : 0.0;
        }
        const double ev = act ? a.eval[(size_t)p * n + lane] : 0.0;
        const bool invar0 = a.ptn_invar[p] == 0.0;
        double o[8];  // ncat <= 8 results of this lane (NORM_LH decides after the last category)
        unsigned long long mxall = 0ull;
        for (int c = 0; c < ncat; c++) {
            const double rc = a.rates[c];
            double eL = 0.0, eR = 0.0;
            if (act) {
                eL = exp(ev * (rc * L.length)) * (Lleaf ? pL[(size_t)p * n + lane] : pL[(size_t)p * block + c * n + lane]);
                eR = exp(ev * (rc * R.length)) * (Rleaf ? pR[(size_t)p * n + lane] : pR[(size_t)p * block + c * n + lane]);
            }
            double fl = 0.0, fr = 0.0;
#pragma unroll
            for (int j = 0; j < NMAX; j++) {
                if (j < n) {
                    fl = fma(U[j], __shfl_sync(0xffffffffu, eL, j), fl);
                    fr = fma(U[j], __shfl_sync(0xffffffffu, eR, j), fr);
                }
            }
            const double tmp = fl * fr;
            double v = 0.0;
#pragma unroll
            for (int x = 0; x < NMAX; x++)
                if (x < n) v = fma(M[x], __shfl_sync(0xffffffffu, tmp, x), v);
            unsigned long long mx = act ? ((unsigned long long)__double_as_longlong(v) & 0x7fffffffffffffffull) : 0ull;
#pragma unroll
            for (int s = 16; s > 0; s >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, s));
            if (SAFE) {
                int sc = (Lleaf ? 0 : scL[p * ncat + c]) + (Rleaf ? 0 : scR[p * ncat + c]);
                if (!tiptip && __longlong_as_double((long long)mx) < B200_SCALING_THRESHOLD && invar0) {
                    v *= B200_TWO_256;
                    sc += 1;
                }
                if (act) out[(size_t)p * block + c * n + lane] = v;
                if (lane == 0) sc_out[p * ncat + c] = (uint8_t)sc;
            } else {
                mxall = max(mxall, mx);
                if (c < 8) o[c] = v;
                else if (act) out[(size_t)p * block + c * n + lane] = v;  // more than 8 categories: fixed up below
            }
        }
        if (!SAFE) {
            const bool under = !tiptip && __longlong_as_double((long long)mxall) < B200_SCALING_THRESHOLD && invar0;
            const double f = under ? B200_TWO_256 : 1.0;
            if (act) {
                for (int c = 0; c < ncat && c < 8; c++) out[(size_t)p * block + c * n + lane] = o[c] * f;
                if (under)
                    for (int c = 8; c < ncat; c++) out[(size_t)p * block + c * n + lane] *= f;
            }
            if (lane == 0) sc_out[p] = (uint8_t)((Lleaf ? 0 : scL[p]) + (Rleaf ? 0 : scR[p]) + (under ? 1 : 0));
        }
    }
}

// ---- 20 states on the FP64 tensor cores ---------------------------------------------------------------------------
// One warp per pattern, one mma.m8n8k4 row per (child, category): rows 0-3 = left child x 4 categories, rows 4-7 =
// right child.  The pattern's own U and Uinv are the B operands, gathered straight from global memory in fragment
// order (every byte of the 2 x 3.2 KB is used exactly once per node); as in dense.cu the accumulator fragment of the
// child products is the A operand of the back-transformation (contraction index permuted in the Uinv gather).
// 30 DMMAs per pattern-node and four categories instead of 240 shuffles + 240 FMAs per lane: the update becomes
// HBM-bound on the eigen-systems.
__device__ __forceinline__ void dmma_s(double &d0, double &d1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(d0), "+d"(d1)
                 : "d"(a), "d"(b));
}

template <bool SAFE>
__global__ void __launch_bounds__(SITE_THREADS) site_partial20_kernel(SiteArgs a) {
    constexpr int n = 20;
    const int ncat = a.ncat, block = n * ncat;
    const DevNode nd = a.nodes[a.node_base + blockIdx.y];
    const DevChild L = a.children[nd.child_begin], R = a.children[nd.child_begin + 1];
    const bool Lleaf = L.kind == B200LK_CHILD_LEAF, Rleaf = R.kind == B200LK_CHILD_LEAF, tiptip = Lleaf && Rleaf;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, q = lane & 3;
    const int child = g >> 2, cl = g & 3;  // this lane's A row: child (0 left, 1 right), category within the group
    double *out = a.partial + (size_t)nd.out_slot * a.slot_doubles;
    uint8_t *sc_out = a.scale + (size_t)nd.out_slot * a.slot_scale_bytes;
    const double *pL = Lleaf ? a.tip + (size_t)L.id * a.nptn * n : a.partial + (size_t)L.id * a.slot_doubles;
    const double *pR = Rleaf ? a.tip + (size_t)R.id * a.nptn * n : a.partial + (size_t)R.id * a.slot_doubles;
    const uint8_t *scL = Lleaf ? nullptr : a.scale + (size_t)L.id * a.slot_scale_bytes;
    const uint8_t *scR = Rleaf ? nullptr : a.scale + (size_t)R.id * a.slot_scale_bytes;
    const double *src = child ? pR : pL;
    const bool src_leaf = child ? Rleaf : Lleaf;
    const double len = child ? R.length : L.length;
    for (int64_t p = (int64_t)blockIdx.x * (SITE_THREADS / 32) + warp; p < a.nptn; p += (int64_t)gridDim.x * (SITE_THREADS / 32)) {
        const double *Up = a.evec + (size_t)p * 400, *Mp = a.inv + (size_t)p * 400, *evp = a.eval + (size_t)p * n;
        double bU[3][5], bM[3][5], ev[5];
#pragma unroll
        for (int s = 0; s < 5; s++) ev[s] = __ldg(evp + 4 * s + q);
#pragma unroll
        for (int j = 0; j < 3; j++) {
            const int x = 8 * j + g;  // state (child products) / eigen index (back-transformation) of this lane's column
#pragma unroll
            for (int s = 0; s < 5; s++) {
                bU[j][s] = x < n ? __ldg(Up + x * n + 4 * s + q) : 0.0;
                const int xx = s < 4 ? 8 * (s >> 1) + 2 * q + (s & 1) : 16 + q;
                bM[j][s] = x < n ? __ldg(Mp + x * n + xx) : 0.0;
            }
        }
        const bool invar0 = a.ptn_invar[p] == 0.0;
        unsigned long long mxall = 0ull;
        for (int c0 = 0; c0 < ncat; c0 += 4) {
            const int c = c0 + cl;
            const bool row_on = c < ncat;
            double A[5];
            const double rl = row_on ? a.rates[c] * len : 0.0;
            const double *sp = src + (src_leaf ? (size_t)p * n : (size_t)p * block + (size_t)c * n) + q;
#pragma unroll
            for (int s = 0; s < 5; s++) A[s] = row_on ? exp(ev[s] * rl) * __ldg(sp + 4 * s) : 0.0;
            double F[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
#pragma unroll
            for (int s = 0; s < 5; s++)
#pragma unroll
                for (int j = 0; j < 3; j++) dmma_s(F[2 * j], F[2 * j + 1], A[s], bU[j][s]);
            // rows 0-3 (left) x rows 4-7 (right): the partner row lives 16 lanes up
            double tmp[6];
#pragma unroll
            for (int k = 0; k < 6; k++) {
                const double r = __shfl_down_sync(0xffffffffu, F[k], 16);
                tmp[k] = g < 4 ? F[k] * r : 0.0;
            }
            const int srcl = (lane & ~3) | (q >> 1);
            const double v0 = __shfl_sync(0xffffffffu, tmp[4], srcl), v1 = __shfl_sync(0xffffffffu, tmp[5], srcl);
            const double a4 = (q & 1) ? v1 : v0;
            double o[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
#pragma unroll
            for (int jo = 0; jo < 3; jo++) {
#pragma unroll
                for (int kk = 0; kk < 4; kk++) dmma_s(o[2 * jo], o[2 * jo + 1], tmp[kk], bM[jo][kk]);
                dmma_s(o[2 * jo], o[2 * jo + 1], a4, bM[jo][4]);
            }
            const bool mine = g < 4 && row_on;  // lanes (g < 4, q) hold out[c = c0 + g][i = 8jo + 2q, +1]
            unsigned long long mx = 0ull;
            if (mine) {
#pragma unroll
                for (int k = 0; k < 6; k++) {
                    const bool live = (8 * (k >> 1) + 2 * q) < n;
                    const unsigned long long b = live ? ((unsigned long long)__double_as_longlong(o[k]) & 0x7fffffffffffffffull) : 0ull;
                    mx = max(mx, b);
                }
            }
            if (SAFE) {
                mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
                mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
                if (mine) {
                    int sc = (Lleaf ? 0 : scL[p * ncat + c]) + (Rleaf ? 0 : scR[p * ncat + c]);
                    if (!tiptip && __longlong_as_double((long long)mx) < B200_SCALING_THRESHOLD && invar0) {
#pragma unroll
                        for (int k = 0; k < 6; k++) o[k] *= B200_TWO_256;
                        sc += 1;
                    }
                    if (q == 0) sc_out[p * ncat + c] = (uint8_t)sc;
                }
            } else {
                mxall = max(mxall, mx);
            }
            if (mine) {
                double *po = out + (size_t)p * block + (size_t)c * n + 2 * q;
                *reinterpret_cast<double2 *>(po) = make_double2(o[0], o[1]);
                *reinterpret_cast<double2 *>(po + 8) = make_double2(o[2], o[3]);
                if (q < 2) *reinterpret_cast<double2 *>(po + 16) = make_double2(o[4], o[5]);
            }
        }
        if (!SAFE) {
#pragma unroll
            for (int s = 16; s > 0; s >>= 1) mxall = max(mxall, __shfl_xor_sync(0xffffffffu, mxall, s));
            const bool under = !tiptip && __longlong_as_double((long long)mxall) < B200_SCALING_THRESHOLD && invar0;
            if (under) {  // rare: the warp rescales the block it has just written
                __syncwarp();
                for (int e = lane; e < block; e += 32) out[(size_t)p * block + e] *= B200_TWO_256;
            }
            if (lane == 0) sc_out[p] = (uint8_t)((Lleaf ? 0 : scL[p]) + (Rleaf ? 0 : scR[p]) + (under ? 1 : 0));
        }
    }
}

int launch_site_partials(b200lk_ctx *c, const std::vector<std::vector<int>> &levels) {
    SiteArgs a;
    a.nodes = c->d_nodes;
    a.children = c->d_children;
    a.eval = c->s_eval;
    a.evec = c->s_evec;
    a.inv = c->s_inv;
    a.tip = c->s_tip;
    a.rates = c->m.rates;
    a.ptn_invar = c->ptn_invar;
    a.partial = c->partial;
    a.scale = c->scale;
    a.slot_doubles = c->slot_doubles;
    a.slot_scale_bytes = c->slot_scale_bytes;
    a.nptn = c->d.nptn;
    a.n = c->d.nstates;
    a.ncat = c->d.ncat;
    const bool safe = c->d.safe_numeric != 0;
    int base = 0;
    for (size_t l = 0; l < levels.size(); l++) {
        const int nb = (int)levels[l].size();
        for (int idx : levels[l])
            if (c->h_nodes[idx].kind == 3) return B200LK_ERR_UNSUPPORTED;  // multifurcating nodes (:1437-1556) not built
        if (!nb) continue;
        a.node_base = base;
        const int64_t per = SITE_THREADS / 32;
        const unsigned gx = (unsigned)std::min<int64_t>((c->d.nptn + per - 1) / per, (int64_t)c->sms * 16);
        dim3 grid(gx, nb);
        prof_mark(c);
        const bool use_dmma = b200_env().site_dmma;  // A/B switch: B200LK_SITE_DMMA=0 keeps the shuffle kernel for 20 states
        if (a.n == 20 && use_dmma) {
            if (safe) site_partial20_kernel<true><<<grid, SITE_THREADS, 0, c->stream>>>(a);
            else site_partial20_kernel<false><<<grid, SITE_THREADS, 0, c->stream>>>(a);
        } else if (a.n <= 4) {
            if (safe) site_partial_kernel<4, true><<<grid, SITE_THREADS, 0, c->stream>>>(a);
            else site_partial_kernel<4, false><<<grid, SITE_THREADS, 0, c->stream>>>(a);
        } else if (a.n <= 20) {
            if (safe) site_partial_kernel<20, true><<<grid, SITE_THREADS, 0, c->stream>>>(a);
            else site_partial_kernel<20, false><<<grid, SITE_THREADS, 0, c->stream>>>(a);
        } else {
            if (safe) site_partial_kernel<32, true><<<grid, SITE_THREADS, 0, c->stream>>>(a);
            else site_partial_kernel<32, false><<<grid, SITE_THREADS, 0, c->stream>>>(a);
        }
        prof_mark(c);
        CUDA_TRY(c, cudaGetLastError());
        c->n_launches++;
        base += nb;
    }
    return B200LK_OK;
}

extern "C" int b200lk_set_site_models(b200lk_ctx *c, const double *eval, const double *evec, const double *inv_evec,
                                      const double *rates, const double *props, const double *tipvec) {
    SHARD_REFUSE(c, "site-specific models");
    if (!c || !eval || !evec || !inv_evec || !rates || !props || !tipvec) return B200LK_ERR_ARG;
    if (!c->site) return B200LK_ERR_STATE;  // the context was not created with dims.site_specific
    if (!c->tips_set) return B200LK_ERR_STATE;  // the tip vectors are derived from the uploaded states
    CUDA_TRY(c, cudaSetDevice(c->device));
    const int n = c->d.nstates, codes = c->d.n_state_codes;
    const int64_t nptn = c->d.nptn;
    cudaStream_t s = c->stream;
    if (!c->s_eval) {
        CUDA_TRY(c, cudaMalloc(&c->s_eval, sizeof(double) * std::max<int64_t>(1, nptn) * n));
        CUDA_TRY(c, cudaMalloc(&c->s_evec, sizeof(double) * std::max<int64_t>(1, nptn) * n * n));
        CUDA_TRY(c, cudaMalloc(&c->s_inv, sizeof(double) * std::max<int64_t>(1, nptn) * n * n));
        CUDA_TRY(c, cudaMalloc(&c->s_tip, sizeof(double) * std::max<int64_t>(1, nptn) * n * c->d.ntaxa));
        CUDA_TRY(c, cudaMalloc(&c->s_tipvec, sizeof(double) * codes * n));
    }
    CUDA_TRY(c, cudaMemcpyAsync(c->s_eval, eval, sizeof(double) * nptn * n, cudaMemcpyHostToDevice, s));
    CUDA_TRY(c, cudaMemcpyAsync(c->s_evec, evec, sizeof(double) * nptn * n * n, cudaMemcpyHostToDevice, s));
    CUDA_TRY(c, cudaMemcpyAsync(c->s_inv, inv_evec, sizeof(double) * nptn * n * n, cudaMemcpyHostToDevice, s));
    CUDA_TRY(c, cudaMemcpyAsync(c->s_tipvec, tipvec, sizeof(double) * codes * n, cudaMemcpyHostToDevice, s));
    CUDA_TRY(c, cudaMemcpyAsync(c->m.rates, rates, sizeof(double) * c->d.ncat, cudaMemcpyHostToDevice, s));
    CUDA_TRY(c, cudaMemcpyAsync(c->m.props, props, sizeof(double) * c->d.ncat, cudaMemcpyHostToDevice, s));
    if (nptn > 0) {
        const int64_t total = (int64_t)c->d.ntaxa * nptn * n;
        site_tip_kernel<<<(unsigned)std::min<int64_t>((total + 255) / 256, (int64_t)c->sms * 32), 256, 0, s>>>(
            c->states, c->s_inv, c->s_tipvec, c->s_tip, nptn, c->ntot_pad, n, c->d.ntaxa);
        CUDA_TRY(c, cudaGetLastError());
        c->n_launches++;
    }
    CUDA_TRY(c, cudaStreamSynchronize(s));  // the host arrays may be reused by the caller
    memcpy(c->h_rates, rates, sizeof(double) * c->d.ncat);
    c->model_set = true;
    c->theta_computed = false;
    return B200LK_OK;
}
