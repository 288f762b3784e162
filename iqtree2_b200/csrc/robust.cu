// robust.cu -- "robust phylogeny" scoring of computeLikelihoodBranch (tree/phylokernelnew.h:3096-3135):
//   params->robust_phy_keep < 1   lnL = sum of the per-SITE log-likelihoods without the sites_drop worst ones
//   params->robust_median         lnL = median per-site log-likelihood
// The reference expands _pattern_lh to one entry per site on the host and runs nth_element.  Here the per-pattern
// values never leave the device: patterns are sorted by log-likelihood together with their frequencies
// (cub radix sort, keys = doubles), an inclusive scan of the sorted frequencies gives every pattern its range of
// site ranks, and one pass clips that range against the ranks kept (or picks the median ranks).
#include <cub/cub.cuh>

#include "common.cuh"

#define ROB_THREADS 256

// kept sites have ranks [drop, sites); pattern i covers ranks [cum[i] - f[i], cum[i])
__global__ void __launch_bounds__(ROB_THREADS) robust_keep_kernel(const double *__restrict__ lh, const double *__restrict__ f,
                                                                   const double *__restrict__ cum, int64_t nptn, double drop,
                                                                   double *__restrict__ red, int red_blocks) {
    double acc = 0.0;
    for (int64_t i = (int64_t)blockIdx.x * ROB_THREADS + threadIdx.x; i < nptn; i += (int64_t)gridDim.x * ROB_THREADS) {
        const double kept = fmin(f[i], fmax(cum[i] - drop, 0.0));
        if (kept > 0.0) acc += lh[i] * kept;
    }
    __shared__ double sh[ROB_THREADS];
    sh[threadIdx.x] = acc;
    __syncthreads();
    for (int s = ROB_THREADS / 2; s > 0; s >>= 1) {
        if (threadIdx.x < s) sh[threadIdx.x] += sh[threadIdx.x + s];
        __syncthreads();
    }
    if (threadIdx.x == 0) red[blockIdx.x] = sh[0];
}
__global__ void __launch_bounds__(ROB_THREADS) robust_final_kernel(const double *__restrict__ red, int red_blocks, double *result) {
    __shared__ double sh[ROB_THREADS];
    double v = 0.0;
    for (int b = threadIdx.x; b < red_blocks; b += ROB_THREADS) v += red[b];
    sh[threadIdx.x] = v;
    __syncthreads();
    for (int s = ROB_THREADS / 2; s > 0; s >>= 1) {
        if (threadIdx.x < s) sh[threadIdx.x] += sh[threadIdx.x + s];
        __syncthreads();
    }
    if (threadIdx.x == 0) result[0] = sh[0];
}
// site ranks r_hi = sites/2 and (even site counts) r_lo = sites/2 - 1: the patterns whose rank range holds them
__global__ void __launch_bounds__(ROB_THREADS) robust_median_kernel(const double *__restrict__ lh, const double *__restrict__ f,
                                                                     const double *__restrict__ cum, int64_t nptn, double r_hi,
                                                                     double r_lo, double *__restrict__ pick) {
    for (int64_t i = (int64_t)blockIdx.x * ROB_THREADS + threadIdx.x; i < nptn; i += (int64_t)gridDim.x * ROB_THREADS) {
        const double lo = cum[i] - f[i], hi = cum[i];
        if (f[i] > 0.0 && lo <= r_hi && r_hi < hi) pick[0] = lh[i];
        if (f[i] > 0.0 && lo <= r_lo && r_lo < hi) pick[1] = lh[i];
    }
}
__global__ void robust_median_final_kernel(const double *pick, int odd, double *result) {
    result[0] = odd ? pick[0] : 0.5 * (pick[0] + pick[1]);
}

int launch_robust(b200lk_ctx *c) {
    const int64_t nptn = c->d.nptn;
    if (nptn == 0) return B200LK_OK;
    if (nptn > 0x7fffffff) return B200LK_ERR_UNSUPPORTED;
    const int num = (int)nptn;
    double *keys_out, *vals_out, *cum;
    if (!c->rob_buf) {
        size_t t1 = 0, t2 = 0;
        cub::DeviceRadixSort::SortPairs(nullptr, t1, (const double *)nullptr, (double *)nullptr, (const double *)nullptr,
                                        (double *)nullptr, num);
        cub::DeviceScan::InclusiveSum(nullptr, t2, (const double *)nullptr, (double *)nullptr, num);
        c->rob_temp_bytes = t1 > t2 ? t1 : t2;
        CUDA_TRY(c, cudaMalloc(&c->rob_buf, sizeof(double) * (3 * (size_t)nptn + 2)));
        CUDA_TRY(c, cudaMalloc(&c->rob_temp, c->rob_temp_bytes));
    }
    keys_out = c->rob_buf;
    vals_out = keys_out + nptn;
    cum = vals_out + nptn;
    double *pick = cum + nptn;
    size_t tb = c->rob_temp_bytes;
    CUDA_TRY(c, cub::DeviceRadixSort::SortPairs(c->rob_temp, tb, (const double *)c->pattern_lh, keys_out,
                                                (const double *)c->ptn_freq, vals_out, num, 0, 64, c->stream));
    tb = c->rob_temp_bytes;
    CUDA_TRY(c, cub::DeviceScan::InclusiveSum(c->rob_temp, tb, (const double *)vals_out, cum, num, c->stream));
    const int grid = (int)std::min<int64_t>(c->red_blocks, (nptn + ROB_THREADS - 1) / ROB_THREADS);
    const double sites = c->nsite;
    if (c->robust_keep < 1.0) {
        const double drop = (double)(int)((1.0 - c->robust_keep) * sites);  // :3102, the reference's own truncation
        robust_keep_kernel<<<grid, ROB_THREADS, 0, c->stream>>>(keys_out, vals_out, cum, nptn, drop, c->red_partial, grid);
        CUDA_TRY(c, cudaGetLastError());
        robust_final_kernel<<<1, ROB_THREADS, 0, c->stream>>>(c->red_partial, grid, c->d_result);
        CUDA_TRY(c, cudaGetLastError());
    } else {
        const int64_t ns = (int64_t)sites;
        robust_median_kernel<<<grid, ROB_THREADS, 0, c->stream>>>(keys_out, vals_out, cum, nptn, (double)(ns / 2),
                                                                 (double)(ns / 2 - 1), pick);
        CUDA_TRY(c, cudaGetLastError());
        robust_median_final_kernel<<<1, 1, 0, c->stream>>>(pick, (int)(ns & 1), c->d_result);
        CUDA_TRY(c, cudaGetLastError());
    }
    c->n_launches += 4;
    return B200LK_OK;
}
