// eigen.cu -- the reversible eigen-decomposition on the device.
//
// Replaces, for reversible models, ModelMarkov::decomposeRateMatrix -> EigenDecomposition::eigensystem_sym
// (model/modelmarkov.cpp:1386-1632, utils/eigendecomposition.cpp:165-317) and computeTipPartialLikelihood
// (tree/phylotreesse.cpp:243-375): only the free parameters of a model -- n(n-1)/2 exchangeabilities and n
// frequencies per mixture class -- travel to the device; eval / evec / inv_evec / tip_partial_lh are produced where
// the kernels read them.  ModelFinder-style probes (thousands of small evaluations with fresh eigen-systems,
// main/phylotesting.cpp:1563-1680) then need neither a host eigen-solve nor an n^2 upload per candidate.
//
// Same construction as the reference: frequencies normalised, a_ij = pi_j r_ij scaled to total_num_subst expected
// substitutions, states with pi <= 1e-10 taken out (identity rows), S = Pi^1/2 Q Pi^-1/2 symmetric.  The solver is a
// parallel cyclic Jacobi (one CTA per mixture class, n/2 disjoint rotations per step in round-robin order) instead
// of the reference's tred2 + tqli: it is the parallel algorithm for matrices this small and at least as accurate;
// eigenpairs come out in a different order, which no consumer depends on.
#include "common.cuh"

#define EIG_THREADS 256
#define EIG_ZERO_FREQ 1e-10  // utils/eigendecomposition.h:24

__global__ void __launch_bounds__(EIG_THREADS) jacobi_model_kernel(const double *__restrict__ exch, const double *__restrict__ freqs,
                                                                    const double *__restrict__ total_subst,
                                                                    const double *__restrict__ tipvec, DevModel m, int n,
                                                                    int nmix, int codes, int *status) {
    extern __shared__ double sh[];
    const int np = (n + 1) & ~1, K = np / 2, cls = blockIdx.x, tid = threadIdx.x;
    double *S = sh, *V = S + np * np, *pi = V + np * np, *rt = pi + np, *cs = rt + np, *red = cs + 2 * K;
    __shared__ double s_scalar[2];
    __shared__ int s_done;
    const double *ex = exch + (size_t)cls * (n * (n - 1) / 2), *fr = freqs + (size_t)cls * n;
    // frequencies, normalised (:188-196)
    if (tid == 0) {
        double sum = 0.0;
        for (int i = 0; i < n; i++) sum += fr[i];
        s_scalar[0] = 1.0 / sum;
    }
    __syncthreads();
    for (int i = tid; i < np; i += EIG_THREADS) {
        const double p = i < n ? fr[i] * s_scalar[0] : 0.0;
        pi[i] = p;
        rt[i] = sqrt(p);
    }
    __syncthreads();
    // a_ij = pi_j r_ij (a rate touching a zero-frequency state is dropped, modelmarkov.cpp:1613), kept in S
    for (int idx = tid; idx < np * np; idx += EIG_THREADS) {
        const int i = idx / np, j = idx - i * np;
        double r = 0.0;
        if (i < n && j < n && i != j && fr[i] > EIG_ZERO_FREQ && fr[j] > EIG_ZERO_FREQ) {
            const int a = i < j ? i : j, b = i < j ? j : i;
            r = ex[a * n - a * (a + 1) / 2 + (b - a - 1)];
        }
        S[idx] = pi[j] * r;
        V[idx] = i == j ? 1.0 : 0.0;
    }
    __syncthreads();
    // row sums, expected rate, normalisation (computeRateMatrix :478-519): fixed summation order
    for (int i = tid; i < np; i += EIG_THREADS) {
        double t = 0.0;
        for (int j = 0; j < np; j++) t += S[i * np + j];
        red[i] = t;
    }
    __syncthreads();
    if (tid == 0) {
        double sum = 0.0;
        for (int i = 0; i < np; i++) sum += red[i] * pi[i];
        s_scalar[1] = (total_subst ? total_subst[cls] : 1.0) / sum;
    }
    __syncthreads();
    // S = Pi^1/2 Q Pi^-1/2: S_ij = delta r_ij sqrt(pi_i pi_j); states below ZERO_FREQ are left out (zero row/column)
    for (int idx = tid; idx < np * np; idx += EIG_THREADS) {
        const int i = idx / np, j = idx - i * np;
        const bool act = pi[i] > EIG_ZERO_FREQ && pi[j] > EIG_ZERO_FREQ;
        double v;
        if (i == j) v = act ? -s_scalar[1] * red[i] : 0.0;
        else v = act ? s_scalar[1] * S[idx] * rt[i] / rt[j] : 0.0;
        S[idx] = v;  // each entry depends on itself only
    }
    __syncthreads();
    // symmetrise exactly (the two triangles differ by rounding)
    for (int idx = tid; idx < np * np; idx += EIG_THREADS) {
        const int i = idx / np, j = idx - i * np;
        if (i < j) {
            const double v = 0.5 * (S[i * np + j] + S[j * np + i]);
            S[i * np + j] = v;
            S[j * np + i] = v;
        }
    }
    __syncthreads();

    int sweep = 0;
    for (; sweep < 60; sweep++) {
        // convergence: off-diagonal mass against the diagonal's
        double off = 0.0, dia = 0.0;
        for (int idx = tid; idx < np * np; idx += EIG_THREADS) {
            const int i = idx / np, j = idx - i * np;
            const double v = S[idx];
            if (i == j) dia += v * v; else off += v * v;
        }
        for (int o = 16; o > 0; o >>= 1) {
            off += __shfl_xor_sync(0xffffffffu, off, o);
            dia += __shfl_xor_sync(0xffffffffu, dia, o);
        }
        if ((tid & 31) == 0) {
            red[2 * (tid >> 5)] = off;
            red[2 * (tid >> 5) + 1] = dia;
        }
        __syncthreads();
        if (tid == 0) {
            double o2 = 0.0, d2 = 0.0;
            for (int w = 0; w < EIG_THREADS / 32; w++) {
                o2 += red[2 * w];
                d2 += red[2 * w + 1];
            }
            s_done = o2 <= 1e-32 * d2 || o2 == 0.0;
        }
        __syncthreads();
        if (s_done) break;
        for (int round = 0; round < np - 1; round++) {
            // rotation parameters of the K disjoint pairs of this round (circle method)
            if (tid < K) {
                const int a = tid == 0 ? np - 1 : (round + tid) % (np - 1), b = (round + np - 1 - tid) % (np - 1);
                const int p = a < b ? a : b, q = a < b ? b : a;
                const double apq = S[p * np + q], app = S[p * np + p], aqq = S[q * np + q];
                double c = 1.0, s = 0.0;
                if (apq != 0.0 && fabs(apq) > 1e-300) {
                    const double theta = (aqq - app) / (2.0 * apq);
                    const double t = (theta >= 0.0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
                    c = 1.0 / sqrt(t * t + 1.0);
                    s = t * c;
                }
                cs[2 * tid] = c;
                cs[2 * tid + 1] = s;
            }
            __syncthreads();
            // columns: S <- S J, V <- V J
            for (int idx = tid; idx < K * np; idx += EIG_THREADS) {
                const int k = idx / np, i = idx - k * np;
                const int a = k == 0 ? np - 1 : (round + k) % (np - 1), b = (round + np - 1 - k) % (np - 1);
                const int p = a < b ? a : b, q = a < b ? b : a;
                const double c = cs[2 * k], s = cs[2 * k + 1];
                const double sp = S[i * np + p], sq = S[i * np + q];
                S[i * np + p] = c * sp - s * sq;
                S[i * np + q] = s * sp + c * sq;
                const double vp = V[i * np + p], vq = V[i * np + q];
                V[i * np + p] = c * vp - s * vq;
                V[i * np + q] = s * vp + c * vq;
            }
            __syncthreads();
            // rows: S <- J^T S
            for (int idx = tid; idx < K * np; idx += EIG_THREADS) {
                const int k = idx / np, j = idx - k * np;
                const int a = k == 0 ? np - 1 : (round + k) % (np - 1), b = (round + np - 1 - k) % (np - 1);
                const int p = a < b ? a : b, q = a < b ? b : a;
                const double c = cs[2 * k], s = cs[2 * k + 1];
                const double sp = S[p * np + j], sq = S[q * np + j];
                S[p * np + j] = c * sp - s * sq;
                S[q * np + j] = s * sp + c * sq;
            }
            __syncthreads();
            // the annihilated pair is zero by construction, not merely up to rounding
            if (tid < K) {
                const int a = tid == 0 ? np - 1 : (round + tid) % (np - 1), b = (round + np - 1 - tid) % (np - 1);
                S[a * np + b] = 0.0;
                S[b * np + a] = 0.0;
            }
            __syncthreads();
        }
    }
    if (tid == 0 && sweep >= 60) atomicExch(status, 1);
    // eval, evec = Pi^-1/2 V, inv_evec = V^T Pi^1/2 (:227-251); left-out states: identity
    double *eval = m.eval + (size_t)cls * n, *evec = m.evec + (size_t)cls * n * n, *inv = m.inv_evec + (size_t)cls * n * n;
    for (int i = tid; i < n; i += EIG_THREADS) eval[i] = pi[i] > EIG_ZERO_FREQ ? S[i * np + i] : 0.0;
    for (int idx = tid; idx < n * n; idx += EIG_THREADS) {
        const int i = idx / n, j = idx - i * n;
        const bool act = pi[i] > EIG_ZERO_FREQ && pi[j] > EIG_ZERO_FREQ;
        evec[idx] = act ? V[i * np + j] / rt[i] : (i == j ? 1.0 : 0.0);
        inv[idx] = act ? V[j * np + i] * rt[j] : (i == j ? 1.0 : 0.0);
    }
    __syncthreads();
    // tip_partial_lh[s][cls][i] = sum_x inv[i][x] tipvec[s][x]  (tree/phylotreesse.cpp:368-375)
    for (int idx = tid; idx < codes * n; idx += EIG_THREADS) {
        const int s = idx / n, i = idx - s * n;
        double v = 0.0;
        for (int x = 0; x < n; x++) {
            const bool act = pi[i] > EIG_ZERO_FREQ && pi[x] > EIG_ZERO_FREQ;
            const double e = act ? V[x * np + i] * rt[x] : (i == x ? 1.0 : 0.0);
            v += e * tipvec[s * n + x];
        }
        m.tip_partial_lh[((size_t)s * nmix + cls) * n + i] = v;
    }
}

extern "C" int b200lk_set_model_rates(b200lk_ctx *c, const double *exch, const double *freqs, const double *total_num_subst,
                                      const double *tipvec, const double *rates, const double *props,
                                      const double *mixw) {
    SHARD_FORWARD(c, shard_set_model_rates(c, exch, freqs, total_num_subst, tipvec, rates, props, mixw));
    if (!c || !exch || !freqs || !tipvec || !rates || !props || !mixw) return B200LK_ERR_ARG;
    CUDA_TRY(c, cudaSetDevice(c->device));
    const int n = c->d.nstates, nmix = c->d.nmix, codes = c->d.n_state_codes, ne = n * (n - 1) / 2;
    if (n > 64) return B200LK_ERR_UNSUPPORTED;
    if (c->site || c->nonrev) return B200LK_ERR_STATE;
    cudaStream_t s = c->stream;
    const size_t in_doubles = (size_t)nmix * (ne + n + 1) + (size_t)codes * n;
    if (!c->eig_in) {
        CUDA_TRY(c, cudaMalloc(&c->eig_in, sizeof(double) * in_doubles + sizeof(int)));
        CUDA_TRY(c, cudaMallocHost(&c->eig_host, sizeof(double) * (in_doubles + (size_t)nmix * (n + 2 * n * n)) + sizeof(int)));
    }
    double *h = c->eig_host, *d_exch = c->eig_in, *d_freqs = d_exch + (size_t)nmix * ne, *d_total = d_freqs + (size_t)nmix * n;
    double *d_tip = d_total + nmix;
    int *d_status = reinterpret_cast<int *>(d_tip + (size_t)codes * n);
    memcpy(h, exch, sizeof(double) * nmix * ne);
    memcpy(h + (size_t)nmix * ne, freqs, sizeof(double) * nmix * n);
    for (int k = 0; k < nmix; k++) h[(size_t)nmix * (ne + n) + k] = total_num_subst ? total_num_subst[k] : 1.0;
    memcpy(h + (size_t)nmix * (ne + n + 1), tipvec, sizeof(double) * codes * n);
    CUDA_TRY(c, cudaMemcpyAsync(c->eig_in, h, sizeof(double) * in_doubles, cudaMemcpyHostToDevice, s));
    CUDA_TRY(c, cudaMemsetAsync(d_status, 0, sizeof(int), s));
    CUDA_TRY(c, cudaMemcpyAsync(c->m.rates, rates, sizeof(double) * c->d.ncat, cudaMemcpyHostToDevice, s));
    CUDA_TRY(c, cudaMemcpyAsync(c->m.props, props, sizeof(double) * c->d.ncat, cudaMemcpyHostToDevice, s));
    CUDA_TRY(c, cudaMemcpyAsync(c->m.mixw, mixw, sizeof(double) * nmix, cudaMemcpyHostToDevice, s));
    const int np = (n + 1) & ~1;
    const size_t smem = sizeof(double) * ((size_t)2 * np * np + 2 * np + np + 2 * np + 2 * (EIG_THREADS / 32) + 8);
    CUDA_TRY(c, cudaFuncSetAttribute(jacobi_model_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
    jacobi_model_kernel<<<nmix, EIG_THREADS, smem, s>>>(d_exch, d_freqs, d_total, d_tip, c->m, n, nmix, codes, d_status);
    CUDA_TRY(c, cudaGetLastError());
    c->n_launches++;
    if (c->dense20 || c->dense_nt) {
        int rc = launch_inv_frag(c);
        if (rc) return rc;
    }
    // host copies of the eigen-system: the 4-state path passes it as kernel parameters (s4.cu)
    double *h_out = h + in_doubles;
    int *h_status = reinterpret_cast<int *>(h_out + (size_t)nmix * (n + 2 * n * n));
    CUDA_TRY(c, cudaMemcpyAsync(h_out, c->m.eval, sizeof(double) * nmix * n, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(c, cudaMemcpyAsync(h_out + (size_t)nmix * n, c->m.evec, sizeof(double) * nmix * n * n, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(c, cudaMemcpyAsync(h_out + (size_t)nmix * (n + n * n), c->m.inv_evec, sizeof(double) * nmix * n * n,
                                cudaMemcpyDeviceToHost, s));
    CUDA_TRY(c, cudaMemcpyAsync(h_status, d_status, sizeof(int), cudaMemcpyDeviceToHost, s));
    CUDA_TRY(c, cudaStreamSynchronize(s));
    if (*h_status) {
        snprintf(c->err, sizeof(c->err), "Jacobi eigen-solve did not converge in 60 sweeps");
        return B200LK_ERR_ARG;
    }
    memcpy(c->h_eval, h_out, sizeof(double) * nmix * n);
    memcpy(c->h_evec, h_out + (size_t)nmix * n, sizeof(double) * nmix * n * n);
    memcpy(c->h_inv_evec, h_out + (size_t)nmix * (n + n * n), sizeof(double) * nmix * n * n);
    memcpy(c->h_rates, rates, sizeof(double) * c->d.ncat);
    c->s4_sym = c->tile == 32 && b200_env().s4_sym && s4_detect_sym(c, c->s4_pi);
    c->model_set = true;
    c->theta_computed = false;
    return B200LK_OK;
}

extern "C" int b200lk_fetch_model(b200lk_ctx *c, double *eval, double *evec, double *inv_evec) {
    SHARD_FORWARD(c, shard_fetch_model(c, eval, evec, inv_evec));
    if (!c || !eval || !evec || !inv_evec) return B200LK_ERR_ARG;
    if (!c->model_set) return B200LK_ERR_STATE;
    CUDA_TRY(c, cudaSetDevice(c->device));
    const int n = c->d.nstates, nmix = c->d.nmix;
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    CUDA_TRY(c, cudaMemcpy(eval, c->m.eval, sizeof(double) * nmix * n, cudaMemcpyDeviceToHost));
    CUDA_TRY(c, cudaMemcpy(evec, c->m.evec, sizeof(double) * nmix * n * n, cudaMemcpyDeviceToHost));
    CUDA_TRY(c, cudaMemcpy(inv_evec, c->m.inv_evec, sizeof(double) * nmix * n * n, cudaMemcpyDeviceToHost));
    return B200LK_OK;
}
