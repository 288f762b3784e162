// nonrev.cu -- the NON-REVERSIBLE kernel family: replaces
// computeNonrevPartialLikelihood{,Generic}SIMD (tree/phylokernelnonrev.h:43-579), the non-reversible branch of
// computePartialInfo (tree/phylokernelnew.h:932-1005), computeNonrevLikelihoodBranch (:1051-1517) and
// computeNonrevLikelihoodDerv (:585-1045), selected by setLikelihoodKernel when !model->isReversible() or
// params->kernel_nonrev (tree/phylokernelfma.cpp:92-100).
//
// Differences from the reversible family that shape the kernels:
//   * partial vectors hold state likelihoods (no eigen-space, no U^-1 back-transformation): a node update is
//     out[c][x] = prod_children (sum_y P_c[x][y] child[c][y]), with P transposed for a child that lies TOWARD_ROOT;
//   * P(t) = Re(V exp(ev t) V^-1) comes from a COMPLEX eigen-system (model/modelmarkov.cpp:481-488), built on the
//     device per (child, category); dP/dt = Q P, d2P/dt2 = Q Q P (:729-745);
//   * rooted trees: the root leaf contributes the state frequencies; no theta buffer exists, the derivative kernel
//     reads both partial vectors at every call.
// Work decomposition as partial.cu: one thread per (pattern, category), a CTA owns PT consecutive patterns of one
// node, grid.y = the independent nodes of a level; tables staged in shared memory once per CTA.
#include <algorithm>

#include "common.cuh"

#define NR_THREADS 256

static inline int nr_kind(int kind) { return kind & B200LK_CHILD_KIND_MASK; }
__device__ __forceinline__ int d_kind(int kind) { return kind & B200LK_CHILD_KIND_MASK; }

// ---- model upload --------------------------------------------------------------------------------------------
extern "C" int b200lk_set_model_nonrev(b200lk_ctx *c, const double *rate_matrix, const double *ceval, const double *cevec,
                                       const double *cinv_evec, const double *freqs, const double *tipvec,
                                       const double *rates, const double *props, const double *mixw, int rooted) {
    SHARD_REFUSE(c, "the non-reversible kernel family");
    if (!c || !rate_matrix || !ceval || !cevec || !cinv_evec || !freqs || !tipvec || !rates || !props || !mixw)
        return B200LK_ERR_ARG;
    if (!c->nonrev) return B200LK_ERR_STATE;
    CUDA_TRY(c, cudaSetDevice(c->device));
    const size_t n = c->d.nstates, nmix = c->d.nmix, codes = c->d.n_state_codes, block = c->block;
    cudaStream_t s = c->stream;
    if (!c->nr_Q) {
        CUDA_TRY(c, cudaMalloc(&c->nr_Q, sizeof(double) * nmix * n * n));
        CUDA_TRY(c, cudaMalloc(&c->nr_ev, sizeof(double) * nmix * n * 2));
        CUDA_TRY(c, cudaMalloc(&c->nr_V, sizeof(double) * nmix * n * n * 2));
        CUDA_TRY(c, cudaMalloc(&c->nr_Vi, sizeof(double) * nmix * n * n * 2));
        CUDA_TRY(c, cudaMalloc(&c->nr_freq, sizeof(double) * nmix * n));
        CUDA_TRY(c, cudaMalloc(&c->nr_tipvec, sizeof(double) * codes * n));
        // branch tables: T0, T1, T2 [ncat_mix][n][n], then the three tip tables [codes][block]
        CUDA_TRY(c, cudaMalloc(&c->nr_T, sizeof(double) * 3 * (block * n + codes * block)));
    }
    CUDA_TRY(c, cudaMemcpyAsync(c->nr_Q, rate_matrix, sizeof(double) * nmix * n * n, cudaMemcpyHostToDevice, s));
    CUDA_TRY(c, cudaMemcpyAsync(c->nr_ev, ceval, sizeof(double) * nmix * n * 2, cudaMemcpyHostToDevice, s));
    CUDA_TRY(c, cudaMemcpyAsync(c->nr_V, cevec, sizeof(double) * nmix * n * n * 2, cudaMemcpyHostToDevice, s));
    CUDA_TRY(c, cudaMemcpyAsync(c->nr_Vi, cinv_evec, sizeof(double) * nmix * n * n * 2, cudaMemcpyHostToDevice, s));
    CUDA_TRY(c, cudaMemcpyAsync(c->nr_freq, freqs, sizeof(double) * nmix * n, cudaMemcpyHostToDevice, s));
    CUDA_TRY(c, cudaMemcpyAsync(c->nr_tipvec, tipvec, sizeof(double) * codes * n, cudaMemcpyHostToDevice, s));
    CUDA_TRY(c, cudaMemcpyAsync(c->m.rates, rates, sizeof(double) * c->d.ncat, cudaMemcpyHostToDevice, s));
    CUDA_TRY(c, cudaMemcpyAsync(c->m.props, props, sizeof(double) * c->d.ncat, cudaMemcpyHostToDevice, s));
    CUDA_TRY(c, cudaMemcpyAsync(c->m.mixw, mixw, sizeof(double) * nmix, cudaMemcpyHostToDevice, s));
    CUDA_TRY(c, cudaStreamSynchronize(s));  // the host arrays may be reused by the caller
    c->rooted = rooted != 0;
    c->model_set = true;
    c->theta_computed = false;
    return B200LK_OK;
}

struct NrModel {
    const double *Q, *ev, *V, *Vi, *freq, *tipvec, *rates, *props, *mixw;
};
static NrModel nr_model(const b200lk_ctx *c) {
    NrModel m;
    m.Q = c->nr_Q; m.ev = c->nr_ev; m.V = c->nr_V; m.Vi = c->nr_Vi; m.freq = c->nr_freq; m.tipvec = c->nr_tipvec;
    m.rates = c->m.rates; m.props = c->m.props; m.mixw = c->m.mixw;
    return m;
}

// P[i][j] = Re(sum_k V[i][k] exp(ev_k t) Vi[k][j]) for mixture class mm, by the whole CTA; `sh` holds 2n doubles.
// dst[i*n + j] (or dst[j*n + i] when `transpose`) *= nothing: plain store.  Ends with __syncthreads().
__device__ void nr_trans_matrix(const NrModel &m, int mm, double t, int n, double *sh, double *dst, bool transpose) {
    const double *ev = m.ev + (size_t)mm * n * 2;
    for (int k = threadIdx.x; k < n; k += blockDim.x) {
        double mag = exp(ev[2 * k] * t), sn, cs;
        sincos(ev[2 * k + 1] * t, &sn, &cs);
        sh[2 * k] = mag * cs;
        sh[2 * k + 1] = mag * sn;
    }
    __syncthreads();
    const double *V = m.V + (size_t)mm * n * n * 2, *Vi = m.Vi + (size_t)mm * n * n * 2;
    for (int idx = threadIdx.x; idx < n * n; idx += blockDim.x) {
        const int i = idx / n, j = idx - i * n;
        double s = 0.0;
        for (int k = 0; k < n; k++) {
            const double vr = V[((size_t)i * n + k) * 2], vi = V[((size_t)i * n + k) * 2 + 1];
            const double wr = Vi[((size_t)k * n + j) * 2], wi = Vi[((size_t)k * n + j) * 2 + 1];
            const double ar = vr * sh[2 * k] - vi * sh[2 * k + 1], ai = vr * sh[2 * k + 1] + vi * sh[2 * k];
            s += ar * wr - ai * wi;
        }
        dst[transpose ? j * n + i : idx] = s;
    }
    __syncthreads();
}

// ---- computePartialInfo, non-reversible branch (tree/phylokernelnew.h:932-1005) -------------------------------
// grid (nchildren, ncat_mix): echild[c][x][y] = P_c[x][y] (P_c[y][x] if the child lies TOWARD_ROOT);
// leaf[s][c][x] = sum_y echild[c][x][y] tipvec[s][y], all ones for the unknown state; the root leaf's table holds
// the state frequencies in every row.
__global__ void __launch_bounds__(NR_THREADS) nr_tables_kernel(const DevChild *__restrict__ children, NrModel m,
                                                               double *__restrict__ tables, int n, int ncat, int denom,
                                                               int ncm, int codes) {
    extern __shared__ double sm[];
    const DevChild ch = children[blockIdx.x];
    const int c = blockIdx.y, mm = c / denom, nn = n * n, block = n * ncm;
    double *e = tables + ch.echild_off + (size_t)c * nn;
    nr_trans_matrix(m, mm, m.rates[c % ncat] * ch.length, n, sm, e, (ch.kind & B200LK_CHILD_TOWARD_ROOT) != 0);
    if (ch.leaf_off < 0) return;
    double *leaf = tables + ch.leaf_off;
    const bool root = d_kind(ch.kind) == B200LK_CHILD_ROOT;
    for (int idx = threadIdx.x; idx < codes * n; idx += blockDim.x) {
        const int s = idx / n, x = idx - s * n;
        double v;
        if (root) {
            v = m.freq[(size_t)mm * n + x];
        } else if (s == codes - 1) {
            v = 1.0;
        } else {
            v = 0.0;
            for (int y = 0; y < n; y++) v += e[x * n + y] * m.tipvec[(size_t)s * n + y];
        }
        leaf[(size_t)s * block + c * n + x] = v;
    }
}

// ---- node update -----------------------------------------------------------------------------------------------
struct NrArgs {
    const DevNode *nodes;
    const DevChild *children;
    const double *tables;
    const uint8_t *states;
    const double *ptn_invar;
    double *partial;
    uint8_t *scale;
    size_t slot_doubles, slot_scale_bytes;
    int64_t ntot, ntot_pad;
    int node_base, n, ncm, codes, PT, tables_in_smem;
};

template <int N>
__device__ __forceinline__ double nr_dot(const double *__restrict__ e, const double *__restrict__ v, int n) {
    const int nn = N ? N : n;
    double a0 = 0.0, a1 = 0.0;
    int i = 0;
#pragma unroll
    for (; i + 1 < nn; i += 2) {
        a0 = fma(e[i], v[i], a0);
        a1 = fma(e[i + 1], v[i + 1], a1);
    }
    if (i < nn) a0 = fma(e[i], v[i], a0);
    return a0 + a1;
}

// Bifurcating node (device kinds: 0 tip-tip incl. root-tip, 1 tip-internal, 2 internal-internal, 4 root-internal);
// the host put the single leaf (or the root leaf) first.
template <int N, bool SAFE>
__global__ void __launch_bounds__(NR_THREADS) nr_partial_kernel(NrArgs a) {
    extern __shared__ double sm[];
    const int n = N ? N : a.n;
    const int ncm = a.ncm, block = n * ncm, nn = n * n;
    const DevNode nd = a.nodes[a.node_base + blockIdx.y];
    const DevChild L = a.children[nd.child_begin];
    const DevChild R = a.children[nd.child_begin + 1];
    const int kL = d_kind(L.kind), kR = d_kind(R.kind);
    const bool Lleaf = kL != B200LK_CHILD_INTERNAL, Rleaf = kR != B200LK_CHILD_INTERNAL;
    const bool Lroot = kL == B200LK_CHILD_ROOT;

    const double *tL = a.tables + (Lleaf ? L.leaf_off : L.echild_off);
    const double *tR = a.tables + (Rleaf ? R.leaf_off : R.echild_off);
    unsigned long long *s_max;
    if (a.tables_in_smem) {
        const int szL = Lleaf ? (Lroot ? block : a.codes * block) : block * n;
        const int szR = Rleaf ? a.codes * block : block * n;
        double *sL = sm, *sR = sL + szL;
        for (int i = threadIdx.x; i < szL; i += blockDim.x) sL[i] = tL[i];
        for (int i = threadIdx.x; i < szR; i += blockDim.x) sR[i] = tR[i];
        tL = sL; tR = sR;
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
    const uint8_t *stL = (Lleaf && !Lroot) ? a.states + (size_t)L.id * a.ntot_pad : nullptr;
    const uint8_t *stR = Rleaf ? a.states + (size_t)R.id * a.ntot_pad : nullptr;
    const int items = a.PT * ncm;
    // tip-tip and root-internal nodes never rescale (tree/phylokernelnonrev.h:247-388)
    const bool may_rescale = !(Lleaf && Rleaf) && !Lroot;

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
        const double *lfL = Lleaf ? tL + (size_t)(Lroot ? 0 : stL[p]) * block + c * n : nullptr;
        const double *lfR = Rleaf ? tR + (size_t)stR[p] * block + c * n : nullptr;
        const double *eL = Lleaf ? nullptr : tL + (size_t)c * nn;
        const double *eR = Rleaf ? nullptr : tR + (size_t)c * nn;
        double mx = 0.0;
        double o[N ? N : 64];
#pragma unroll
        for (int x = 0; x < (N ? N : n); x++) {
            const double l = Lleaf ? lfL[x] : nr_dot<N>(eL + x * n, pvl, n);
            const double r = Rleaf ? lfR[x] : nr_dot<N>(eR + x * n, pvr, n);
            o[x] = l * r;
            mx = fmax(mx, o[x]);
        }
        double *po = out_base + off;
        if (SAFE) {
            int s = (Lleaf ? 0 : scL[p * ncm + c]) + (Rleaf ? 0 : scR[p * ncm + c]);
            if (may_rescale && mx < B200_SCALING_THRESHOLD && a.ptn_invar[p] == 0.0) {
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
    for (int it = threadIdx.x; it < items; it += blockDim.x) {
        const int pl_ = it / ncm, c = it - pl_ * ncm;
        const int64_t p = p0 + pl_;
        if (p >= a.ntot) continue;
        const double mx = __longlong_as_double((long long)s_max[pl_]);
        const bool under = may_rescale && mx < B200_SCALING_THRESHOLD && a.ptn_invar[p] == 0.0;
        if (under) {
            double *po = out_base + (size_t)p * block + (size_t)c * n;
            for (int i = 0; i < n; i++) po[i] *= B200_TWO_256;
        }
        if (c == 0) sc_out[p] = (uint8_t)((Lleaf ? 0 : scL[p]) + (Rleaf ? 0 : scR[p]) + (under ? 1 : 0));
    }
}

// Multifurcating node (tree/phylokernelnonrev.h:104-245): running product over the children with a rescale test
// after every child.  One thread per pattern; the output vector holds the running product.
template <bool SAFE>
__global__ void __launch_bounds__(128) nr_partial_multi_kernel(NrArgs a) {
    const int n = a.n, ncm = a.ncm, block = n * ncm, nn = n * n;
    const DevNode nd = a.nodes[a.node_base + blockIdx.y];
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= a.ntot) return;
    double *all = a.partial + (size_t)nd.out_slot * a.slot_doubles + (size_t)p * block;
    uint8_t *sd = a.scale + (size_t)nd.out_slot * a.slot_scale_bytes + (SAFE ? p * ncm : p);
    for (int i = 0; i < block; i++) all[i] = 1.0;
    for (int i = 0; i < (SAFE ? ncm : 1); i++) sd[i] = 0;
    const bool invar0 = a.ptn_invar[p] == 0.0;
    for (int k = 0; k < nd.nchild; k++) {
        const DevChild ch = a.children[nd.child_begin + k];
        const int kind = d_kind(ch.kind);
        if (kind != B200LK_CHILD_INTERNAL) {
            const int st = kind == B200LK_CHILD_ROOT ? 0 : a.states[(size_t)ch.id * a.ntot_pad + p];
            const double *lf = a.tables + ch.leaf_off + (size_t)st * block;
            for (int i = 0; i < block; i++) all[i] *= lf[i];
        } else {
            const double *pc = a.partial + (size_t)ch.id * a.slot_doubles + (size_t)p * block;
            const uint8_t *sc = a.scale + (size_t)ch.id * a.slot_scale_bytes;
            if (!SAFE) sd[0] += sc[p];
            const double *e = a.tables + ch.echild_off;
            for (int c = 0; c < ncm; c++) {
                if (SAFE) sd[c] += sc[p * ncm + c];
                for (int x = 0; x < n; x++) all[c * n + x] *= nr_dot<0>(e + (size_t)c * nn + x * n, pc + c * n, n);
            }
        }
        if (SAFE) {
            for (int c = 0; c < ncm; c++) {
                double mx = all[c * n];
                for (int x = 1; x < n; x++) mx = fmax(mx, all[c * n + x]);
                if (mx < B200_SCALING_THRESHOLD && invar0) {
                    for (int x = 0; x < n; x++) all[c * n + x] *= B200_TWO_256;
                    sd[c] += 1;
                }
            }
        } else {
            double mx = all[0];
            for (int x = 1; x < block; x++) mx = fmax(mx, all[x]);
            if (mx < B200_SCALING_THRESHOLD && invar0) {
                for (int x = 0; x < block; x++) all[x] *= B200_TWO_256;
                sd[0] += 1;
            }
        }
    }
}

template <int N>
static void nr_launch_one(NrArgs &a, bool safe, dim3 grid, size_t smem, cudaStream_t s) {
    if (safe) {
        if (smem > 48 * 1024)
            cudaFuncSetAttribute(nr_partial_kernel<N, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        nr_partial_kernel<N, true><<<grid, NR_THREADS, smem, s>>>(a);
    } else {
        if (smem > 48 * 1024)
            cudaFuncSetAttribute(nr_partial_kernel<N, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        nr_partial_kernel<N, false><<<grid, NR_THREADS, smem, s>>>(a);
    }
}

// b200lk_update_partials of a non-reversible context: validation, level batching (as api.cu), tables, updates.
int enqueue_update_nonrev(b200lk_ctx *c, const b200lk_node *nodes, int32_t nnodes, const b200lk_child *children,
                          int32_t nchildren) {
    const int n = c->d.nstates, block = c->block, nslots = c->d.nslots, codes = c->d.n_state_codes;
    const int ring = c->stage_idx++ % B200_STAGE_RING;
    if (c->stage_used[ring]) CUDA_TRY(c, cudaEventSynchronize(c->stage_ev[ring]));
    c->h_nodes = c->h_nodes_ring + (size_t)ring * c->nodes_cap;
    c->h_children = c->h_children_ring + (size_t)ring * c->children_cap;
    std::vector<int> level_of_slot(nslots, -1), reader_level(nslots, -1), level(nnodes, 0);
    int nlevels = 0;
    for (int i = 0; i < nnodes; i++) {
        const b200lk_node &nd = nodes[i];
        if (nd.out_slot < 0 || nd.out_slot >= nslots || nd.nchild < 2 || nd.child_begin < 0 ||
            nd.child_begin + nd.nchild > nchildren) {
            snprintf(c->err, sizeof(c->err), "bad traversal entry %d", i);
            return B200LK_ERR_ARG;
        }
        int lv = 0, nroot = 0;
        for (int k = 0; k < nd.nchild; k++) {
            const b200lk_child &ch = children[nd.child_begin + k];
            const int kind = nr_kind(ch.kind);
            if (kind == B200LK_CHILD_ROOT) {
                if (!c->rooted || ++nroot > 1) return B200LK_ERR_ARG;
            } else if (kind == B200LK_CHILD_LEAF) {
                if (ch.id < 0 || ch.id >= c->d.ntaxa) return B200LK_ERR_ARG;
            } else if (kind == B200LK_CHILD_INTERNAL) {
                if (ch.id < 0 || ch.id >= nslots || ch.id == nd.out_slot) return B200LK_ERR_ARG;
                lv = std::max(lv, level_of_slot[ch.id] + 1);
            } else {
                return B200LK_ERR_ARG;
            }
        }
        lv = std::max(lv, std::max(reader_level[nd.out_slot], level_of_slot[nd.out_slot]) + 1);
        for (int k = 0; k < nd.nchild; k++) {
            const b200lk_child &ch = children[nd.child_begin + k];
            if (nr_kind(ch.kind) == B200LK_CHILD_INTERNAL) reader_level[ch.id] = std::max(reader_level[ch.id], lv);
        }
        level[i] = lv;
        level_of_slot[nd.out_slot] = lv;
        nlevels = std::max(nlevels, lv + 1);
    }
    size_t off = 0;
    for (int j = 0; j < nchildren; j++) {
        DevChild &dc = c->h_children[j];
        dc.kind = children[j].kind;
        dc.id = children[j].id;
        dc.length = children[j].length;
        dc.echild_off = (int64_t)off;
        off += (size_t)block * n;
        dc.frag_off = dc.aux_off = -1;
        dc.vtt_off = dc.vpair_off = -1;
        dc.src = j;
        if (nr_kind(dc.kind) != B200LK_CHILD_INTERNAL) {
            dc.leaf_off = (int64_t)off;
            off += (size_t)codes * block;
        } else {
            dc.leaf_off = -1;
        }
    }
    if (off > c->tables_cap) return B200LK_ERR_NOMEM;
    // a single leaf goes first (tree/phylokernelnonrev.h:95-102); of two leaves the root leaf goes first (:256-262)
    std::vector<int> kind_of(nnodes, 3);
    for (int i = 0; i < nnodes; i++) {
        if (nodes[i].nchild > 2) continue;
        DevChild &a = c->h_children[nodes[i].child_begin], &b = c->h_children[nodes[i].child_begin + 1];
        const bool al = nr_kind(a.kind) != B200LK_CHILD_INTERNAL, bl = nr_kind(b.kind) != B200LK_CHILD_INTERNAL;
        if ((!al && bl) || (al && bl && nr_kind(b.kind) == B200LK_CHILD_ROOT)) std::swap(a, b);
        const bool l0 = nr_kind(a.kind) != B200LK_CHILD_INTERNAL, l1 = nr_kind(b.kind) != B200LK_CHILD_INTERNAL;
        kind_of[i] = l0 ? (l1 ? 0 : (nr_kind(a.kind) == B200LK_CHILD_ROOT ? 4 : 1)) : 2;
    }
    std::vector<std::vector<int>> levels(nlevels);
    int pos = 0;
    for (int lv = 0; lv < nlevels; lv++)
        for (int pass = 0; pass < 2; pass++)
            for (int i = 0; i < nnodes; i++) {
                if (level[i] != lv || (kind_of[i] == 3) != (pass == 1)) continue;
                DevNode &dn = c->h_nodes[pos];
                dn.out_slot = nodes[i].out_slot;
                dn.child_begin = nodes[i].child_begin;
                dn.nchild = nodes[i].nchild;
                dn.kind = kind_of[i];
                levels[lv].push_back(pos++);
            }
    CUDA_TRY(c, cudaMemcpyAsync(c->d_children, c->h_children, sizeof(DevChild) * nchildren, cudaMemcpyHostToDevice,
                                c->stream));
    CUDA_TRY(c, cudaMemcpyAsync(c->d_nodes, c->h_nodes, sizeof(DevNode) * nnodes, cudaMemcpyHostToDevice, c->stream));
    CUDA_TRY(c, cudaEventRecord(c->stage_ev[ring], c->stream));
    c->stage_used[ring] = true;

    const int denom = c->d.fused_mix_rate ? 1 : c->d.ncat, ncm = c->ncat_mix;
    nr_tables_kernel<<<dim3(nchildren, ncm), NR_THREADS, sizeof(double) * 2 * n, c->stream>>>(
        c->d_children, nr_model(c), c->tables, n, c->d.ncat, denom, ncm, codes);
    CUDA_TRY(c, cudaGetLastError());
    c->n_launches++;

    NrArgs a;
    a.nodes = c->d_nodes; a.children = c->d_children; a.tables = c->tables; a.states = c->states;
    a.ptn_invar = c->ptn_invar; a.partial = c->partial; a.scale = c->scale;
    a.slot_doubles = c->slot_doubles; a.slot_scale_bytes = c->slot_scale_bytes;
    a.ntot = c->ntot; a.ntot_pad = c->ntot_pad; a.n = n; a.ncm = ncm; a.codes = codes;
    a.PT = NR_THREADS / ncm > 0 ? NR_THREADS / ncm : 1;
    const size_t tbl = 2 * (size_t)std::max(codes * block, block * n) * sizeof(double);
    a.tables_in_smem = tbl + a.PT * 8 <= 200 * 1024;
    const size_t smem = (a.tables_in_smem ? tbl : 0) + (size_t)a.PT * 8;
    const bool safe = c->d.safe_numeric != 0;
    int base = 0;
    for (size_t l = 0; l < levels.size(); l++) {
        int nb = 0, nm = 0;
        for (int idx : levels[l]) (c->h_nodes[idx].kind == 3 ? nm : nb)++;
        if (nb) {
            a.node_base = base;
            dim3 grid((unsigned)((c->ntot + a.PT - 1) / a.PT), nb);
            prof_mark(c);
            if (n == 4) nr_launch_one<4>(a, safe, grid, smem, c->stream);
            else if (n == 20) nr_launch_one<20>(a, safe, grid, smem, c->stream);
            else nr_launch_one<0>(a, safe, grid, smem, c->stream);
            prof_mark(c);
            CUDA_TRY(c, cudaGetLastError());
            c->n_launches++;
        }
        if (nm) {
            a.node_base = base + nb;
            dim3 grid((unsigned)((c->ntot + 127) / 128), nm);
            prof_mark(c);
            if (safe) nr_partial_multi_kernel<true><<<grid, 128, 0, c->stream>>>(a);
            else nr_partial_multi_kernel<false><<<grid, 128, 0, c->stream>>>(a);
            prof_mark(c);
            CUDA_TRY(c, cudaGetLastError());
            c->n_launches++;
        }
        base += nb + nm;
    }
    c->n_updates += (int64_t)nnodes * c->ntot;
    c->theta_computed = false;
    return B200LK_OK;
}

// ---- branch level ------------------------------------------------------------------------------------------------
int nr_check_ends(b200lk_ctx *c, b200lk_end dad, b200lk_end node) {
    if (node.kind != B200LK_CHILD_INTERNAL || node.id < 0 || node.id >= c->d.nslots) return B200LK_ERR_ARG;
    if (dad.kind == B200LK_CHILD_ROOT) return c->rooted ? B200LK_OK : B200LK_ERR_ARG;
    if (dad.kind == B200LK_CHILD_LEAF) return (dad.id >= 0 && dad.id < c->d.ntaxa) ? B200LK_OK : B200LK_ERR_ARG;
    if (dad.kind != B200LK_CHILD_INTERNAL || dad.id < 0 || dad.id >= c->d.nslots || dad.id == node.id)
        return B200LK_ERR_ARG;
    return B200LK_OK;
}

// One CTA per category c: T0[c][i][x] = prop_c P_c[i][x] (* freq_m[i] on an unrooted tree), and with `derv`
// T1 = prop rate Q P, T2 = prop rate^2 Q Q P (tree/phylokernelnonrev.h:1101-1120, 628-659); then the tables of a
// tip dad, tab[k][s][c][i] = sum_x Tk[c][i][x] tipvec[s][x] (:1138-1162, 693-720), or of the root leaf,
// tab[0][0][c][i] = freq_m[i] prop_c.  dad_kind: INTERNAL (no table), LEAF, ROOT.
__global__ void __launch_bounds__(NR_THREADS) nr_branch_tables_kernel(NrModel m, double *__restrict__ T, double length,
                                                                      int n, int ncat, int denom, int ncm, int codes,
                                                                      int rooted, int derv, int dad_kind) {
    extern __shared__ double sm[];  // 2n (exp) + 3 n*n
    const int c = blockIdx.x, mm = c / denom, k = c % ncat, nn = n * n, block = n * ncm;
    const double rate = m.rates[k], prop = m.props[k] * m.mixw[mm];
    double *P = sm + 2 * n, *D1 = P + nn, *D2 = D1 + nn;
    nr_trans_matrix(m, mm, rate * length, n, sm, P, false);
    const double *Q = m.Q + (size_t)mm * nn;
    if (derv) {
        for (int pass = 0; pass < 2; pass++) {
            const double *src = pass ? D1 : P;
            double *dst = pass ? D2 : D1;
            for (int idx = threadIdx.x; idx < nn; idx += blockDim.x) {
                const int i = idx / n, j = idx - i * n;
                double s = 0.0;
                for (int q = 0; q < n; q++) s += Q[i * n + q] * src[q * n + j];
                dst[idx] = s;
            }
            __syncthreads();
        }
    }
    const size_t tsz = (size_t)block * n;
    const double sc[3] = {prop, prop * rate, prop * rate * rate};
    for (int idx = threadIdx.x; idx < nn; idx += blockDim.x) {
        const double f = rooted ? 1.0 : m.freq[(size_t)mm * n + idx / n];
        P[idx] *= sc[0] * f;
        T[(size_t)c * nn + idx] = P[idx];
        if (derv) {
            D1[idx] *= sc[1] * f;
            D2[idx] *= sc[2] * f;
            T[tsz + (size_t)c * nn + idx] = D1[idx];
            T[2 * tsz + (size_t)c * nn + idx] = D2[idx];
        }
    }
    if (dad_kind == B200LK_CHILD_INTERNAL) return;
    __syncthreads();
    double *tab = T + 3 * tsz;
    const size_t tabsz = (size_t)codes * block;
    if (dad_kind == B200LK_CHILD_ROOT) {
        for (int i = threadIdx.x; i < n; i += blockDim.x) tab[c * n + i] = m.freq[(size_t)mm * n + i] * prop;
        return;
    }
    for (int idx = threadIdx.x; idx < codes * n; idx += blockDim.x) {
        const int s = idx / n, i = idx - s * n;
        const double *tv = m.tipvec + (size_t)s * n;
        for (int q = 0; q < (derv ? 3 : 1); q++) {
            const double *M = P + (size_t)q * nn + i * n;
            double v = 0.0;
            for (int x = 0; x < n; x++) v += M[x] * tv[x];
            tab[q * tabsz + (size_t)s * block + c * n + i] = v;
        }
    }
}

struct NrBranchArgs {
    const double *T;           // T0, T1, T2, tab0, tab1, tab2
    const double *pdad, *pnode;  // dad_branch->partial_lh (row side), node_branch->partial_lh (column side) or null
    const uint8_t *sdad, *snode, *dad_states;
    const double *ptn_freq, *ptn_invar;
    double *pattern_lh, *pattern_lh_cat, *red_partial;
    int64_t nptn;
    int n, ncm, codes, safe, dad_kind, PT;
};

__device__ __forceinline__ void nr_block_reduce(double *acc, int K, double *red_partial, int red_blocks) {
    __shared__ double sh[2][NR_THREADS / 32];
    for (int k = 0; k < K; k++) {
        double v = acc[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0) sh[k][threadIdx.x >> 5] = v;
    }
    __syncthreads();
    if (threadIdx.x < K) {
        double v = 0.0;
        for (int w = 0; w < NR_THREADS / 32; w++) v += sh[threadIdx.x][w];
        red_partial[(size_t)threadIdx.x * red_blocks + blockIdx.x] = v;
    }
}

__global__ void nr_final_reduce_kernel(const double *red_partial, int red_blocks, int K, double *result, int dst0) {
    __shared__ double sh[NR_THREADS];
    for (int k = 0; k < K; k++) {
        double v = 0.0;
        for (int b = threadIdx.x; b < red_blocks; b += NR_THREADS) v += red_partial[(size_t)k * red_blocks + b];
        sh[threadIdx.x] = v;
        __syncthreads();
        for (int s = NR_THREADS / 2; s > 0; s >>= 1) {
            if (threadIdx.x < s) sh[threadIdx.x] += sh[threadIdx.x + s];
            __syncthreads();
        }
        if (threadIdx.x == 0) result[dst0 + k] = sh[0];
        __syncthreads();
    }
}

// Branch lnL (DERV == false) or df/ddf (DERV == true).  A CTA walks tiles of PT patterns (fixed grid, fixed order:
// the sums are bit-reproducible); phase 1: one thread per (pattern, category) forms the category's likelihood (and
// derivatives); phase 2: one thread per pattern combines the categories (SAFE: min-scale re-weighting).
template <bool DERV>
__global__ void __launch_bounds__(NR_THREADS) nr_branch_kernel(NrBranchArgs a, int red_blocks) {
    extern __shared__ double sm[];  // [DERV ? 3 : 1][PT * ncm]
    const int n = a.n, ncm = a.ncm, block = n * ncm, nn = n * n, items = a.PT * ncm;
    const size_t tsz = (size_t)block * n, tabsz = (size_t)a.codes * block;
    const double *tab = a.T + 3 * tsz;
    double acc[2] = {0.0, 0.0};
    const int64_t ntiles = (a.nptn + a.PT - 1) / a.PT;
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int64_t p0 = tile * a.PT;
        for (int it = threadIdx.x; it < items; it += blockDim.x) {
            const int pl_ = it / ncm, c = it - pl_ * ncm;
            const int64_t p = p0 + pl_;
            if (p >= a.nptn) continue;
            const double *pd = a.pdad + (size_t)p * block + c * n;
            double v0 = 0.0, v1 = 0.0, v2 = 0.0;
            if (a.dad_kind != B200LK_CHILD_INTERNAL) {
                const int st = a.dad_kind == B200LK_CHILD_ROOT ? 0 : a.dad_states[p];
                const double *row = tab + (size_t)st * block + c * n;
                for (int i = 0; i < n; i++) {
                    const double x = pd[i];
                    v0 = fma(row[i], x, v0);
                    if (DERV) {
                        v1 = fma(row[tabsz + i], x, v1);
                        v2 = fma(row[2 * tabsz + i], x, v2);
                    }
                }
            } else {
                const double *pn = a.pnode + (size_t)p * block + c * n;
                const double *T0 = a.T + (size_t)c * nn;
                for (int i = 0; i < n; i++) {
                    double s0 = 0.0, s1 = 0.0, s2 = 0.0;
                    for (int x = 0; x < n; x++) {
                        const double y = pn[x];
                        s0 = fma(T0[i * n + x], y, s0);
                        if (DERV) {
                            s1 = fma(T0[tsz + i * n + x], y, s1);
                            s2 = fma(T0[2 * tsz + i * n + x], y, s2);
                        }
                    }
                    const double x = pd[i];
                    v0 = fma(x, s0, v0);
                    if (DERV) {
                        v1 = fma(x, s1, v1);
                        v2 = fma(x, s2, v2);
                    }
                }
            }
            if (DERV && a.safe) {
                // each side is brought to its own smallest scale count (:750-778, 896-949); its category minimum
                // is recomputed by every thread of the pattern (ncat_mix byte reads from L1)
                double f = 1.0;
                for (int side = 0; side < (a.dad_kind == B200LK_CHILD_INTERNAL ? 2 : 1); side++) {
                    const uint8_t *sv = (side ? a.snode : a.sdad) + (size_t)p * ncm;
                    int mn = sv[0];
                    for (int cc = 1; cc < ncm; cc++) mn = min(mn, (int)sv[cc]);
                    const int d = sv[c] - mn;
                    f *= d == 0 ? 1.0 : (d == 1 ? B200_SCALING_THRESHOLD : 0.0);
                }
                v0 *= f; v1 *= f; v2 *= f;
            }
            sm[it] = v0;
            if (DERV) {
                sm[items + it] = v1;
                sm[2 * items + it] = v2;
            } else {
                a.pattern_lh_cat[(size_t)p * ncm + c] = v0;
            }
        }
        __syncthreads();
        for (int pl_ = threadIdx.x; pl_ < a.PT; pl_ += blockDim.x) {
            const int64_t p = p0 + pl_;
            if (p >= a.nptn) continue;
            const double *l0 = sm + (size_t)pl_ * ncm;
            if (DERV) {
                double lh = 0.0, d1 = 0.0, d2 = 0.0;
                for (int c = 0; c < ncm; c++) {
                    lh += l0[c];
                    d1 += l0[items + c];
                    d2 += l0[2 * items + c];
                }
                lh += a.ptn_invar[p];
                const double inv = 1.0 / lh, df_frac = d1 * inv, ddf_frac = d2 * inv, freq = a.ptn_freq[p];
                const double tmp1 = df_frac * freq, tmp2 = ddf_frac * freq;
                acc[0] += tmp1;
                acc[1] += tmp2 - tmp1 * df_frac;
            } else {
                double lh = 0.0, min_scale;
                if (a.safe) {
                    const uint8_t *sd = a.sdad + (size_t)p * ncm, *sn = a.snode ? a.snode + (size_t)p * ncm : nullptr;
                    int mn = 0;
                    for (int c = 0; c < ncm; c++) {
                        const int s = (uint8_t)(sd[c] + (sn ? sn[c] : 0));
                        if (c == 0 || s < mn) mn = s;
                    }
                    for (int c = 0; c < ncm; c++) {
                        const int s = (uint8_t)(sd[c] + (sn ? sn[c] : 0));
                        double v = l0[c];
                        if (s == mn + 1) v *= B200_SCALING_THRESHOLD;
                        else if (s > mn + 1) v = 0.0;
                        if (s != mn) a.pattern_lh_cat[(size_t)p * ncm + c] = v;
                        lh += v;
                    }
                    min_scale = mn;
                } else {
                    for (int c = 0; c < ncm; c++) lh += l0[c];
                    min_scale = a.sdad[p] + (a.snode ? a.snode[p] : 0);
                }
                lh += a.ptn_invar[p];
                const double v = log(lh) + min_scale * B200_LOG_SCALING_THRESHOLD;
                a.pattern_lh[p] = v;
                acc[0] = fma(v, a.ptn_freq[p], acc[0]);
            }
        }
        __syncthreads();
    }
    nr_block_reduce(acc, DERV ? 2 : 1, a.red_partial, red_blocks);
}

static int nr_launch_branch(b200lk_ctx *c, b200lk_end dad, b200lk_end node, double length, bool derv) {
    const int n = c->d.nstates, ncm = c->ncat_mix, codes = c->d.n_state_codes;
    const int denom = c->d.fused_mix_rate ? 1 : c->d.ncat;
    const size_t tsm = sizeof(double) * (2 * n + 3 * n * n);
    if (tsm > 48 * 1024)
        cudaFuncSetAttribute(nr_branch_tables_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tsm);
    nr_branch_tables_kernel<<<ncm, NR_THREADS, tsm, c->stream>>>(nr_model(c), c->nr_T, length, n, c->d.ncat, denom, ncm,
                                                                 codes, c->rooted, derv, dad.kind);
    CUDA_TRY(c, cudaGetLastError());
    NrBranchArgs a;
    memset(&a, 0, sizeof(a));
    a.T = c->nr_T;
    a.pdad = slot_partial(c, node.id);
    a.sdad = slot_scale(c, node.id);
    if (dad.kind == B200LK_CHILD_INTERNAL) {
        a.pnode = slot_partial(c, dad.id);
        a.snode = slot_scale(c, dad.id);
    } else if (dad.kind == B200LK_CHILD_LEAF) {
        a.dad_states = c->states + (size_t)dad.id * c->ntot_pad;
    }
    a.ptn_freq = c->ptn_freq; a.ptn_invar = c->ptn_invar;
    a.pattern_lh = c->pattern_lh; a.pattern_lh_cat = c->pattern_lh_cat; a.red_partial = c->red_partial;
    a.nptn = c->d.nptn; a.n = n; a.ncm = ncm; a.codes = codes; a.safe = c->d.safe_numeric; a.dad_kind = dad.kind;
    a.PT = NR_THREADS / ncm > 0 ? NR_THREADS / ncm : 1;
    const int64_t ntiles = (a.nptn + a.PT - 1) / a.PT;
    const int grid = (int)std::max<int64_t>(1, std::min<int64_t>(ntiles, c->red_blocks));
    const size_t smem = sizeof(double) * (derv ? 3 : 1) * (size_t)a.PT * ncm;
    if (derv) {
        if (smem > 48 * 1024)
            cudaFuncSetAttribute(nr_branch_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        nr_branch_kernel<true><<<grid, NR_THREADS, smem, c->stream>>>(a, grid);
    } else {
        if (smem > 48 * 1024)
            cudaFuncSetAttribute(nr_branch_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        nr_branch_kernel<false><<<grid, NR_THREADS, smem, c->stream>>>(a, grid);
    }
    CUDA_TRY(c, cudaGetLastError());
    nr_final_reduce_kernel<<<1, NR_THREADS, 0, c->stream>>>(c->red_partial, grid, derv ? 2 : 1, c->d_result, derv ? 1 : 0);
    CUDA_TRY(c, cudaGetLastError());
    c->n_launches += 3;
    return B200LK_OK;
}

int nr_launch_branch_lnl(b200lk_ctx *c, b200lk_end dad, b200lk_end node, double length) {
    return nr_launch_branch(c, dad, node, length, false);
}
// b200lk_build_theta only records the ends: the non-reversible family keeps no theta buffer
int nr_record_ends(b200lk_ctx *c, b200lk_end dad, b200lk_end node) {
    c->nr_dad = dad;
    c->nr_node = node;
    return B200LK_OK;
}
int nr_launch_derv(b200lk_ctx *c, double length) {
    // the virtual root branch has no length to optimise (tree/phylokernelnonrev.h:667-670)
    if (c->nr_dad.kind == B200LK_CHILD_ROOT) return B200LK_ERR_UNSUPPORTED;
    return nr_launch_branch(c, c->nr_dad, c->nr_node, length, true);
}
// the reference binds no from-buffer kernel for this family (tree/phylokernelfma.cpp:99): the wrapper falls back to
// computeLikelihoodBranch at the current branch (tree/phylotreesse.cpp:230-234)
int nr_launch_from_buffer(b200lk_ctx *c, double length) {
    return nr_launch_branch(c, c->nr_dad, c->nr_node, length, false);
}
