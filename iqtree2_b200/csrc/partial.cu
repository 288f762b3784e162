// partial.cu -- the pruning update (replaces computePartialLikelihood{,Generic}SIMD,
// tree/phylokernelnew.h:1291-2039) with the SCALING_THRESHOLD rescale fused in.
//
// Work decomposition: one thread per (pattern, category); a CTA owns a tile of PT
// consecutive patterns of ONE node, grid.y runs over the independent nodes of a level
// (level batching), so a whole level is one launch.  P-factor tables (echild / leaf
// lookup / inv_evec) are staged in shared memory once per CTA.
//
//   tmp[x]  = prod_children ( internal: sum_i echild[c][x][i] * child[p][c][i] | leaf: leaf[state][c][x] )
//   out[i]  = sum_x inv_evec[m][i][x] * tmp[x]
//   NORM: if max_{c,i}|out| < 2^-256 and ptn_invar[p] == 0: out *= 2^256, scale[p] = sum child scale + 1
//   SAFE: the same test per (p, c)
#include "common.cuh"

struct PartialArgs {
    const DevNode *nodes;
    const DevChild *children;
    const double *tables;
    const double *inv_evec;
    const uint8_t *states;
    const double *ptn_invar;
    double *partial;
    uint8_t *scale;
    size_t slot_doubles, slot_scale_bytes;
    int64_t ntot, ntot_pad;
    int node_base, n, ncm, denom, nmix, codes, PT, tables_in_smem, tile;
};

__device__ __forceinline__ double dabs(double x) { return fabs(x); }

// reference reduction tree for N == 4 (dotProductDualVec / productVecMat, :235-237, :335-338)
__device__ __forceinline__ double dot4_tree(const double *a, const double *b) {
    return __dadd_rn(__fma_rn(a[1], b[1], __dmul_rn(a[0], b[0])), __fma_rn(a[3], b[3], __dmul_rn(a[2], b[2])));
}
// dotProductVec N == 4 (:90-93)
__device__ __forceinline__ double dot4_plain(const double *a, const double *b) {
    return __dadd_rn(__dadd_rn(__dmul_rn(a[0], b[0]), __dmul_rn(a[1], b[1])),
                     __dadd_rn(__dmul_rn(a[2], b[2]), __dmul_rn(a[3], b[3])));
}

template <int N>
__device__ __forceinline__ double dotN(const double *__restrict__ e, const double *__restrict__ v, int n) {
    if (N == 4) return dot4_tree(e, v);
    const int nn = N ? N : n;
    double a0 = e[0] * v[0], a1 = 0.0;
    int i = 1;
#pragma unroll 4
    for (; i + 1 < nn; i += 2) {
        a1 = fma(e[i], v[i], a1);
        a0 = fma(e[i + 1], v[i + 1], a0);
    }
    if (i < nn) a1 = fma(e[i], v[i], a1);
    return a0 + a1;
}

template <int N, bool SAFE>
__global__ void __launch_bounds__(256) partial_kernel(PartialArgs a) {
    extern __shared__ double sm[];
    const int n = N ? N : a.n;
    const int ncm = a.ncm, block = n * ncm, nn = n * n;
    const DevNode nd = a.nodes[a.node_base + blockIdx.y];
    const DevChild L = a.children[nd.child_begin];
    const DevChild R = a.children[nd.child_begin + 1];
    const bool Lleaf = L.kind == B200LK_CHILD_LEAF, Rleaf = R.kind == B200LK_CHILD_LEAF;

    // ---- stage tables ------------------------------------------------------
    const double *tL = a.tables + (Lleaf ? L.leaf_off : L.echild_off);
    const double *tR = a.tables + (Rleaf ? R.leaf_off : R.echild_off);
    const double *inv = a.inv_evec;
    unsigned long long *s_max;
    if (a.tables_in_smem) {
        const int szL = Lleaf ? a.codes * block : block * n, szR = Rleaf ? a.codes * block : block * n;
        const int szI = a.nmix * nn;
        double *sI = sm, *sL = sI + szI, *sR = sL + szL;
        for (int i = threadIdx.x; i < szI; i += blockDim.x) sI[i] = inv[i];
        for (int i = threadIdx.x; i < szL; i += blockDim.x) sL[i] = tL[i];
        for (int i = threadIdx.x; i < szR; i += blockDim.x) sR[i] = tR[i];
        inv = sI; tL = sL; tR = sR;
        s_max = (unsigned long long *)(sR + szR);
    } else {
        s_max = (unsigned long long *)sm;
    }
    if (!SAFE)
        for (int i = threadIdx.x; i < a.PT; i += blockDim.x) s_max[i] = 0ull;
    __syncthreads();

    const int64_t p0 = (int64_t)blockIdx.x * a.PT;
    double *out_base = a.partial + (size_t)nd.out_slot * a.slot_doubles;
    uint8_t *sc_out = a.scale + (size_t)nd.out_slot * a.slot_scale_bytes;
    const double *PL = Lleaf ? nullptr : a.partial + (size_t)L.id * a.slot_doubles;
    const double *PR = Rleaf ? nullptr : a.partial + (size_t)R.id * a.slot_doubles;
    const uint8_t *scL = Lleaf ? nullptr : a.scale + (size_t)L.id * a.slot_scale_bytes;
    const uint8_t *scR = Rleaf ? nullptr : a.scale + (size_t)R.id * a.slot_scale_bytes;
    const uint8_t *stL = Lleaf ? a.states + (size_t)L.id * a.ntot_pad : nullptr;
    const uint8_t *stR = Rleaf ? a.states + (size_t)R.id * a.ntot_pad : nullptr;
    const int items = a.PT * ncm;

    for (int it = threadIdx.x; it < items; it += blockDim.x) {
        const int pl_ = it / ncm, c = it - pl_ * ncm;
        const int64_t p = p0 + pl_;
        if (p >= a.ntot) continue;
        const size_t off = (size_t)p * block + (size_t)c * n;
        double vl[N ? N : 1], vr[N ? N : 1];
        const double *pvl = nullptr, *pvr = nullptr;
        if (!Lleaf) {
            pvl = PL + off;
            if (N) {
#pragma unroll
                for (int i = 0; i < N; i += 2) {
                    double2 t = *reinterpret_cast<const double2 *>(pvl + i);
                    vl[i] = t.x; vl[i + 1] = t.y;
                }
                pvl = vl;
            }
        }
        if (!Rleaf) {
            pvr = PR + off;
            if (N) {
#pragma unroll
                for (int i = 0; i < N; i += 2) {
                    double2 t = *reinterpret_cast<const double2 *>(pvr + i);
                    vr[i] = t.x; vr[i + 1] = t.y;
                }
                pvr = vr;
            }
        }
        const double *lfL = Lleaf ? tL + (size_t)stL[p] * block + c * n : nullptr;
        const double *lfR = Rleaf ? tR + (size_t)stR[p] * block + c * n : nullptr;
        const double *eL = Lleaf ? nullptr : tL + (size_t)c * nn;
        const double *eR = Rleaf ? nullptr : tR + (size_t)c * nn;

        double tmp[N ? N : 64];
#pragma unroll
        for (int x = 0; x < (N ? N : n); x++) {
            double l, r;
            if (Lleaf) l = lfL[x];
            else l = dotN<N>(eL + x * n, pvl, n);
            if (Rleaf) r = lfR[x];
            else if (N == 4 && Lleaf) r = dot4_plain(eR + x * n, pvr);  // tip-internal uses dotProductVec (:1853)
            else r = dotN<N>(eR + x * n, pvr, n);
            tmp[x] = __dmul_rn(l, r);
        }
        const double *M = inv + (size_t)(c / a.denom) * nn;
        double mx = 0.0;
        double o[N ? N : 64];
#pragma unroll
        for (int i = 0; i < (N ? N : n); i++) {
            o[i] = dotN<N>(M + i * n, tmp, n);  // productVecMat: X[i] = sum_x A[x]*M[i][x]
            mx = fmax(mx, dabs(o[i]));
        }
        double *po = out_base + off;
        if (SAFE) {
            int s = (Lleaf ? 0 : scL[p * ncm + c]) + (Rleaf ? 0 : scR[p * ncm + c]);
            // tip-tip never rescales (:1632-1746)
            if (!(Lleaf && Rleaf) && mx < B200_SCALING_THRESHOLD && a.ptn_invar[p] == 0.0) {
#pragma unroll
                for (int i = 0; i < (N ? N : n); i++) o[i] *= B200_TWO_256;
                s += 1;
            }
            sc_out[p * ncm + c] = (uint8_t)s;
        } else {
            atomicMax(&s_max[pl_], (unsigned long long)__double_as_longlong(mx));
        }
        if (N) {
#pragma unroll
            for (int i = 0; i < N; i += 2) *reinterpret_cast<double2 *>(po + i) = make_double2(o[i], o[i + 1]);
        } else {
            for (int i = 0; i < n; i++) po[i] = o[i];
        }
    }
    if (SAFE) return;
    __syncthreads();
    // NORM: one decision per pattern over the whole block (:1889-1903, 2015-2029)
    for (int it = threadIdx.x; it < items; it += blockDim.x) {
        const int pl_ = it / ncm, c = it - pl_ * ncm;
        const int64_t p = p0 + pl_;
        if (p >= a.ntot) continue;
        const double mx = __longlong_as_double((long long)s_max[pl_]);
        const bool under = !(Lleaf && Rleaf) && mx < B200_SCALING_THRESHOLD && a.ptn_invar[p] == 0.0;
        if (under) {
            double *po = out_base + (size_t)p * block + (size_t)c * n;
            for (int i = 0; i < n; i++) po[i] *= B200_TWO_256;
        }
        if (c == 0) {
            int s = (Lleaf ? 0 : scL[p]) + (Rleaf ? 0 : scR[p]) + (under ? 1 : 0);
            sc_out[p] = (uint8_t)s;
        }
    }
}

// Multifurcating node (degree > 3), tree/phylokernelnew.h:1407-1631: running product over the
// children with a rescale test after EVERY child, then inv_evec.  Rare (star trees, polytomies):
// one thread per pattern, the output vector doubles as scratch for the running product.
template <bool SAFE>
__global__ void __launch_bounds__(128) partial_multi_kernel(PartialArgs a) {
    const int n = a.n, ncm = a.ncm, block = n * ncm, nn = n * n, tile = a.tile;
    const DevNode nd = a.nodes[a.node_base + blockIdx.y];
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= a.ntot) return;
    double *all = a.partial + (size_t)nd.out_slot * a.slot_doubles;  // element e of this pattern at E(e)
#define E(e) elem_off(p, (e), block, tile)
    uint8_t *sd = a.scale + (size_t)nd.out_slot * a.slot_scale_bytes + (SAFE ? p * ncm : p);
    for (int i = 0; i < block; i++) all[E(i)] = 1.0;
    for (int i = 0; i < (SAFE ? ncm : 1); i++) sd[i] = 0;
    const bool invar0 = a.ptn_invar[p] == 0.0;
    for (int k = 0; k < nd.nchild; k++) {
        const DevChild ch = a.children[nd.child_begin + k];
        if (ch.kind == B200LK_CHILD_LEAF) {
            const double *lf = a.tables + ch.leaf_off + (size_t)a.states[(size_t)ch.id * a.ntot_pad + p] * block;
            for (int i = 0; i < block; i++) all[E(i)] *= lf[i];
        } else {
            const double *pc = a.partial + (size_t)ch.id * a.slot_doubles;
            const uint8_t *sc = a.scale + (size_t)ch.id * a.slot_scale_bytes;
            if (!SAFE) sd[0] += sc[p];
            const double *e = a.tables + ch.echild_off;
            for (int c = 0; c < ncm; c++) {
                if (SAFE) sd[c] += sc[p * ncm + c];
                for (int x = 0; x < n; x++) {
                    double v = e[0] * pc[E(c * n)];
                    for (int i = 1; i < n; i++) v = fma(e[i], pc[E(c * n + i)], v);
                    e += n;
                    all[E(c * n + x)] *= v;
                }
            }
        }
        if (SAFE) {
            for (int c = 0; c < ncm; c++) {
                double mx = 0.0;
                for (int x = 0; x < n; x++) mx = fmax(mx, fabs(all[E(c * n + x)]));
                if (mx < B200_SCALING_THRESHOLD && invar0) {
                    for (int x = 0; x < n; x++) all[E(c * n + x)] *= B200_TWO_256;
                    sd[c] += 1;
                }
            }
        } else {
            double mx = 0.0;
            for (int x = 0; x < block; x++) mx = fmax(mx, fabs(all[E(x)]));
            if (mx < B200_SCALING_THRESHOLD && invar0) {
                for (int x = 0; x < block; x++) all[E(x)] *= B200_TWO_256;
                sd[0] += 1;
            }
        }
    }
    double tmp[64];
    for (int c = 0; c < ncm; c++) {
        const double *M = a.inv_evec + (size_t)(c / a.denom) * nn;
        for (int x = 0; x < n; x++) tmp[x] = all[E(c * n + x)];
        for (int i = 0; i < n; i++) all[E(c * n + i)] = dotN<0>(M + i * n, tmp, n);
    }
#undef E
}

template <int N>
static void launch_one(PartialArgs &a, bool safe, dim3 grid, int threads, size_t smem, cudaStream_t s) {
    if (safe) {
        if (smem > 48 * 1024)
            cudaFuncSetAttribute(partial_kernel<N, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        partial_kernel<N, true><<<grid, threads, smem, s>>>(a);
    } else {
        if (smem > 48 * 1024)
            cudaFuncSetAttribute(partial_kernel<N, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        partial_kernel<N, false><<<grid, threads, smem, s>>>(a);
    }
}

static void fill_partial_args(b200lk_ctx *c, PartialArgs &a) {
    a.nodes = c->d_nodes;
    a.children = c->d_children;
    a.tables = c->tables;
    a.inv_evec = c->m.inv_evec;
    a.states = c->states;
    a.ptn_invar = c->ptn_invar;
    a.partial = c->partial;
    a.scale = c->scale;
    a.slot_doubles = c->slot_doubles;
    a.slot_scale_bytes = c->slot_scale_bytes;
    a.ntot = c->ntot;
    a.ntot_pad = c->ntot_pad;
    a.n = c->d.nstates;
    a.ncm = c->ncat_mix;
    a.denom = c->d.fused_mix_rate ? 1 : c->d.ncat;
    a.nmix = c->d.nmix;
    a.codes = c->d.n_state_codes;
    a.tile = c->tile;
    a.PT = 0;
    a.tables_in_smem = 0;
    a.node_base = 0;
}

// one multifurcating node (index into the device node array); works on either layout
int launch_partial_multi(b200lk_ctx *c, int node_index) {
    PartialArgs a;
    fill_partial_args(c, a);
    a.node_base = node_index;
    dim3 grid((unsigned)((c->ntot + 127) / 128), 1);
    prof_mark(c);
    if (c->d.safe_numeric) partial_multi_kernel<true><<<grid, 128, 0, c->stream>>>(a);
    else partial_multi_kernel<false><<<grid, 128, 0, c->stream>>>(a);
    prof_mark(c);
    CUDA_TRY(c, cudaGetLastError());
    c->n_launches++;
    return B200LK_OK;
}

// levels[l] = indices (into the level-sorted device node array) of the nodes of level l;
// nodes of one level are contiguous: [base, base + count).  Pattern-major layout only (tile == 1).
int launch_partials(b200lk_ctx *c, const std::vector<std::vector<int>> &levels) {
    PartialArgs a;
    fill_partial_args(c, a);
    const int n = a.n, block = c->block;
    const int threads = 256;
    a.PT = threads / a.ncm > 0 ? threads / a.ncm : 1;
    // worst-case table footprint: two leaf tables
    size_t tbl = ((size_t)a.nmix * n * n + 2 * (size_t)std::max(a.codes * block, block * n)) * sizeof(double);
    a.tables_in_smem = tbl + a.PT * 8 <= 200 * 1024;
    size_t smem = (a.tables_in_smem ? tbl : 0) + (size_t)a.PT * 8;
    const bool safe = c->d.safe_numeric != 0;
    int base = 0;
    for (size_t l = 0; l < levels.size(); l++) {
        // within a level: bifurcating nodes first [base, base+nb), multifurcating after
        int nb = 0, nm = 0;
        for (int idx : levels[l]) (c->h_nodes[idx].kind == 3 ? nm : nb)++;
        if (nb && c->dense20) {
            int ntt = 0;  // tip-tip nodes come first inside a level (api.cu)
            for (int idx : levels[l])
                ntt += c->h_nodes[idx].kind == 0 && c->h_children[c->h_nodes[idx].child_begin].aux_off >= 0;
            if (ntt) {
                int rc = launch_dense20_tt(c, base, ntt);  // gather from the state-pair tables
                if (rc) return rc;
            }
            if (nb > ntt) {
                // nodes that read a virtual tip-tip child come last among the bifurcating ones (api.cu): their own launch,
                // so that the others run the lean build of the tensor kernel
                int nv = 0;
                if (b200_env().dense_split_virt)
                    for (int idx : levels[l]) {
                        const DevNode &dn = c->h_nodes[idx];
                        if (dn.kind == 3 || (dn.kind == 0 && c->h_children[dn.child_begin].aux_off >= 0)) continue;
                        bool v = false;
                        for (int k = 0; k < 2; k++) v |= c->h_children[dn.child_begin + k].vtt_off >= 0;
                        nv += v;
                    }
                const int nlean = nb - ntt - nv;
                if (nlean > 0) {
                    int rc = launch_dense20(c, base + ntt, nlean);  // FP64 tensor-core path (dense.cu)
                    if (rc) return rc;
                }
                if (nv > 0) {
                    int rc = launch_dense20(c, base + ntt + nlean, nv);
                    if (rc) return rc;
                }
            }
        } else if (nb && c->dense_nt) {
            int rc = launch_denseN(c, base, nb);  // generic FP64 tensor-core path (densen.cu)
            if (rc) return rc;
        } else if (nb) {
            a.node_base = base;
            dim3 grid((unsigned)((c->ntot + a.PT - 1) / a.PT), nb);
            prof_mark(c);
            if (n == 4) launch_one<4>(a, safe, grid, threads, smem, c->stream);
            else if (n == 20) launch_one<20>(a, safe, grid, threads, smem, c->stream);
            else launch_one<0>(a, safe, grid, threads, smem, c->stream);
            prof_mark(c);
            CUDA_TRY(c, cudaGetLastError());
            c->n_launches++;
        }
        if (nm) {
            a.node_base = base + nb;
            dim3 grid((unsigned)((c->ntot + 127) / 128), nm);
            prof_mark(c);
            if (safe) partial_multi_kernel<true><<<grid, 128, 0, c->stream>>>(a);
            else partial_multi_kernel<false><<<grid, 128, 0, c->stream>>>(a);
            prof_mark(c);
            CUDA_TRY(c, cudaGetLastError());
            c->n_launches++;
        }
        base += nb + nm;
    }
    return B200LK_OK;
}
