// tables.cu -- P-factor tables on device (replaces PhyloTree::computePartialInfo,
// tree/phylokernelnew.h:900-1110, reversible branch :1008-1109).
//
//   echild[c][x][i] = evec[m][x][i] * exp(eval[m][i] * rate[c % ncat] * len)        (:1016-1031)
//   leaf[s][c][x]   = sum_i echild[c][x][i] * tip_partial_lh[s][m][i], row s = UNKNOWN := 1   (:1036-1056)
//
// One CTA per child of the traversal; only branch lengths travel host -> device.
#include <algorithm>

#include "common.cuh"

__global__ void __launch_bounds__(256) tables_kernel(const DevChild *__restrict__ children, DevModel m,
                                                      double *__restrict__ tables, int n, int ncat, int ncm,
                                                      int denom, int nmix, int codes, int frag_nt, int frag_ks,
                                                      const double *__restrict__ mixlen, BatchLaunch bl) {
    DevChild ch = children[blockIdx.x];
    if (bl.d_ctx) {  // batched: context blockIdx.y, its own model, tables and branch lengths
        const BatchCtxDev &b = bl.d_ctx[blockIdx.y];
        m = b.m;
        tables = b.tables;
        if (bl.d_child_len) ch.length = bl.d_child_len[(size_t)blockIdx.y * bl.nchildren + ch.src];
    }
    if (bl.d_lens && ch.len_slot >= 0) ch.length = bl.d_lens[ch.len_slot];  // device-resident lengths (sweep.cu)
    const int block = n * ncm, nn = n * n;
    double *echild = tables + ch.echild_off;
    // grid.z splits the categories of a child over CTAs (mixtures: 240 categories per child); every phase below only
    // reads what the same category produced, so the CTA-local barriers stay sufficient
    const int cpc = (ncm + gridDim.z - 1) / gridDim.z, c0 = blockIdx.z * cpc, c1 = min(ncm, c0 + cpc);
    if (c0 >= c1) return;
    // expv[c][i] = exp(eval[m][i] * rate * len): ONE exp() per (category, eigenvalue) -- n of them per category, not the
    // n * n of echild's elements (supermatrix batches spent 11 % of a step on those, profiles/launches_cfg5_r1t.csv).
    // Also what the 4-state path reads per child (s4.cu).  <= 16 categories per CTA (launch_tables), n <= 64.
    __shared__ double s_expv[16 * 64];
    double *expv = echild + (size_t)block * n;
    for (int idx = c0 * n + threadIdx.x; idx < c1 * n; idx += blockDim.x) {
        int c = idx / n, i = idx - c * n;
        // heterotachy: child->getLength(cat_id[c]) (:1017, 1069), one length per category
        const double e = exp(m.eval[(c / denom) * n + i] *
                             (m.rates[c % ncat] * (mixlen ? mixlen[(size_t)ch.src * ncat + c % ncat] : ch.length)));
        expv[idx] = e;
        s_expv[idx - c0 * n] = e;
    }
    __syncthreads();
    // phase 1: echild
    for (int idx = c0 * nn + threadIdx.x; idx < c1 * nn; idx += blockDim.x) {
        int c = idx / nn, r = idx - c * nn, i = r % n;
        int mm = c / denom;
        echild[idx] = m.evec[(size_t)mm * nn + r] * s_expv[(c - c0) * n + i];
    }
    // The two derived tables below form echild's elements again as evec * expv (the same product, the same bits) instead of
    // reading them back: no barrier, no global round trip between the phases (supermatrix batches: 99 k such CTAs per launch).
    if (ch.frag_off >= 0) {
        // tensor paths (dense.cu: 3 n-tiles x 5 k-steps for 20 states; densen.cu: NT x 2NT): echild as the B operand
        // of mma.m8n8k4, fragment order
        //   frag[c][j][s][lane] = echild[c][x = 8j + lane/4][i = 4s + lane%4]   (x >= n or i >= n -> 0)
        double *frag = tables + ch.frag_off;
        const int fc = frag_nt * frag_ks * 32;
        for (int idx = c0 * fc + threadIdx.x; idx < c1 * fc; idx += blockDim.x) {
            int c = idx / fc, r = idx - c * fc;
            int j = r / (frag_ks * 32), s = (r >> 5) % frag_ks, lane = r & 31;
            int x = 8 * j + (lane >> 2), i = 4 * s + (lane & 3);
            frag[idx] = (x < n && i < n) ? m.evec[(size_t)(c / denom) * nn + x * n + i] * s_expv[(c - c0) * n + i] : 0.0;
        }
    }
    if (ch.leaf_off < 0) return;
    double *leaf = tables + ch.leaf_off;
    const int tip_block = n * nmix;
    const int span = (c1 - c0) * n;
    // One thread per (category, state row x) and group of tip states: the row's products evec * expv are the same for every
    // tip state, so they are formed once (registers for n <= 20) and the loop over the states only reads the tip vectors --
    // the same address for a whole warp, i.e. broadcasts.  One thread per OUTPUT element re-read evec and expv for each of
    // the `codes` states: 60 L1 loads per element, which made this phase -- not exp(), not the barriers -- what a
    // supermatrix batch spent 3 ms per evaluation on (profiles/launches_cfg5_r3b.csv).  Same terms, same order.
    const bool fits = span <= (int)blockDim.x;  // else (many categories of many states) the threads stride over the rows
    const int ngroups = fits ? (int)blockDim.x / span : 1;
    const int grp = fits ? threadIdx.x / span : 0;
    if (grp >= ngroups) return;
    for (int k = fits ? threadIdx.x - grp * span : threadIdx.x; k < span; k += fits ? span : (int)blockDim.x) {
    const int r = c0 * n + k, c = r / n, x = r - c * n;  // r = c*n + x
    const double *ev = m.evec + (size_t)(c / denom) * nn + (size_t)x * n;
    const double *ex = s_expv + (c - c0) * n;
    constexpr int NREG = 20;
    double w[NREG];
    if (n <= NREG) {
#pragma unroll
        for (int i = 0; i < NREG; i++) w[i] = i < n ? ev[i] * ex[i] : 0.0;
    }
    for (int s = grp; s < codes; s += ngroups) {
        double v;
        if (s == codes - 1) {
            v = 1.0;
        } else {
            const double *tip = m.tip_partial_lh + (size_t)s * tip_block + (c / denom) * n;
            if (n <= NREG) {
                v = w[0] * tip[0];
#pragma unroll
                for (int i = 1; i < NREG; i++)
                    if (i < n) v += w[i] * tip[i];
            } else {
                v = (ev[0] * ex[0]) * tip[0];
                for (int i = 1; i < n; i++) v += (ev[i] * ex[i]) * tip[i];
            }
        }
        leaf[s * block + r] = v;
    }
    }
}

int launch_tables(b200lk_ctx *c, int nchildren) {
    if (nchildren <= 0) return B200LK_OK;
    const int denom = c->d.fused_mix_rate ? 1 : c->d.ncat;
    BatchLaunch bl;
    memset(&bl, 0, sizeof(bl));
    if (c->batch) bl = *c->batch;
    bl.d_lens = c->cur_len_slots ? c->d_lens : nullptr;
    const int zsplit = (c->ncat_mix + 15) / 16;  // <= 16 categories per CTA (the size of its shared exp() table)
    tables_kernel<<<dim3(nchildren, c->batch ? c->batch->n : 1, zsplit), 256, 0, c->stream>>>(
        c->d_children, c->m, c->tables, c->d.nstates, c->d.ncat, c->ncat_mix, denom, c->d.nmix, c->d.n_state_codes,
        c->dense20 ? 3 : c->dense_nt, c->dense20 ? 5 : 2 * c->dense_nt,
        c->h_child_mixlen ? c->d_child_mixlen : nullptr, bl);
    CUDA_TRY(c, cudaGetLastError());
    c->n_launches++;
    return B200LK_OK;
}

// Tip-tip nodes of the 20-state path: both children are leaves, so the node's partial vector is a function of the
// (left state, right state) pair only -- 24 x 24 pairs for amino acids against hundreds of thousands of patterns.
//   tt[sl][sr][c][i] = sum_x inv_evec[m][i][x] * leafL[sl][c][x] * leafR[sr][c][x]       (:1632-1746 per pattern)
// The update of such a node is then a gather from this table (dense20_tt_kernel), bounded by the HBM write only.
__global__ void __launch_bounds__(256) tt_tables_kernel(const DevNode *__restrict__ nodes, const DevChild *__restrict__ children,
                                                         const double *__restrict__ tables, const double *__restrict__ inv_evec,
                                                         double *__restrict__ tt, int n, int ncm, int denom, int codes) {
    const DevNode nd = nodes[blockIdx.x];
    if (nd.kind != 0) return;
    const DevChild L = children[nd.child_begin], R = children[nd.child_begin + 1];
    if (L.aux_off < 0) return;
    const int block = n * ncm;
    const double *leafL = tables + L.leaf_off, *leafR = tables + R.leaf_off;
    double *out = tt + L.aux_off;
    // one thread per (category, state pair): pairs vary fastest, so a warp shares the category and reads the same
    // row of inv_evec (broadcast); the n products leafL * leafR stay in registers for all n outputs
    const int npairs = codes * codes;
    const size_t total = (size_t)ncm * npairs;
    for (size_t idx = (size_t)blockIdx.y * blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.y * blockDim.x) {
        const int cc = (int)(idx / npairs), pair = (int)(idx - (size_t)cc * npairs);
        const int sl = pair / codes, sr = pair - sl * codes;
        const double *a = leafL + (size_t)sl * block + cc * n, *b = leafR + (size_t)sr * block + cc * n;
        const double *M = inv_evec + (size_t)(cc / denom) * n * n;
        constexpr int N = 20;  // the tensor path is the 20-state path: registers, fully unrolled
        double tmp[N];
#pragma unroll
        for (int x = 0; x < N; x++) tmp[x] = a[x] * b[x];
        double *o = out + (size_t)pair * block + cc * N;
        for (int i = 0; i < N; i++) {
            double v = 0.0;
#pragma unroll
            for (int x = 0; x < N; x++) v = fma(__ldg(M + i * N + x), tmp[x], v);
            o[i] = v;
        }
    }
}

int launch_tt_tables(b200lk_ctx *c, int nnodes) {
    const int denom = c->d.fused_mix_rate ? 1 : c->d.ncat;
    const size_t total = (size_t)c->d.n_state_codes * c->d.n_state_codes * c->ncat_mix;
    const int gy = (int)std::min<size_t>(4096, (total + 255) / 256);
    tt_tables_kernel<<<dim3(nnodes, gy), 256, 0, c->stream>>>(c->d_nodes, c->d_children, c->tables, c->m.inv_evec,
                                                             c->tt_tables, c->d.nstates, c->ncat_mix, denom,
                                                             c->d.n_state_codes);
    CUDA_TRY(c, cudaGetLastError());
    c->n_launches++;
    return B200LK_OK;
}

// inv_evec per mixture class as the B operand of the final product out[p][i] = sum_x tmp[p][x] inv[i][x]:
// the A operand is the accumulator fragment of the previous mma (lane (g, q) holds x = 8j + 2q + r), so the
// contraction index is permuted instead of shuffling registers:
//   k-step kk = 2j + r (j < 2):  frag[jo][kk][lane] = inv[i = 8jo + lane/4][x = 8j + 2(lane%4) + r]
//   k-step 4 (x = 16..19):       frag[jo][4][lane]  = inv[i = 8jo + lane/4][x = 16 + lane%4]
__global__ void inv_frag_kernel(const double *__restrict__ inv_evec, double *__restrict__ frag, int nmix) {
    const int n = 20;
    for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < nmix * B200_FRAG20; idx += gridDim.x * blockDim.x) {
        int m = idx / B200_FRAG20, r = idx - m * B200_FRAG20;
        int jo = r / 160, kk = (r / 32) % 5, lane = r & 31;
        int i = 8 * jo + (lane >> 2);
        int x = kk < 4 ? 8 * (kk >> 1) + 2 * (lane & 3) + (kk & 1) : 16 + (lane & 3);
        frag[idx] = i < n ? inv_evec[(size_t)m * n * n + i * n + x] : 0.0;
    }
}

// generic tensor path (densen.cu): every k-step pairs with an accumulator register of the child products,
//   k-step kk = 2j + b:  frag[jo][kk][lane] = inv[i = 8jo + lane/4][x = 8j + 2(lane%4) + b]   (i >= n or x >= n -> 0)
__global__ void inv_fragN_kernel(const double *__restrict__ inv_evec, double *__restrict__ frag, int nmix, int n, int nt) {
    const int ks = 2 * nt, fc = nt * ks * 32;
    for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < nmix * fc; idx += gridDim.x * blockDim.x) {
        int m = idx / fc, r = idx - m * fc;
        int jo = r / (ks * 32), kk = (r >> 5) % ks, lane = r & 31;
        int i = 8 * jo + (lane >> 2), x = 8 * (kk >> 1) + 2 * (lane & 3) + (kk & 1);
        frag[idx] = (i < n && x < n) ? inv_evec[(size_t)m * n * n + i * n + x] : 0.0;
    }
}

int launch_inv_frag(b200lk_ctx *c) {
    if (c->dense_nt) {
        const int total = c->d.nmix * c->frag_cat;
        inv_fragN_kernel<<<(total + 255) / 256, 256, 0, c->stream>>>(c->m.inv_evec, c->m.inv_frag, c->d.nmix, c->d.nstates,
                                                                     c->dense_nt);
        CUDA_TRY(c, cudaGetLastError());
        return B200LK_OK;
    }
    inv_frag_kernel<<<(c->d.nmix * B200_FRAG20 + 255) / 256, 256, 0, c->stream>>>(c->m.inv_evec, c->m.inv_frag, c->d.nmix);
    CUDA_TRY(c, cudaGetLastError());
    return B200LK_OK;
}
