// hbm_streams.cu -- measured HBM ceilings of this GPU for the three stream shapes the likelihood kernels have:
//   write-only (the 4-state traversal: 98 % of its DRAM traffic is the partial vectors it leaves resident),
//   read-only  (theta / derv / from-buffer), and copy (the driver's MEASURED_PEAKS.json figure).
// The roofline denominator in MEASURED_PEAKS.json is a COPY bandwidth (read + write bytes); a kernel that only writes
// cannot be judged against it without knowing what a pure write stream reaches.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o hbm_streams tools/hbm_streams.cu && ./hbm_streams
// Prints one JSON line.  4 GiB buffers (>> 126 MB L2), 128-bit accesses, best of 5 after a warm-up.
#include <cuda_runtime.h>
#include <stdio.h>

__global__ void write_kernel(double2 *__restrict__ dst, size_t n, double v) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) dst[i] = make_double2(v, v);
}
// the store shape of s4_traverse_kernel: a thread writes 8 x 16 B at a 512 B stride (T32 tiles), warp = 512 B segments
__global__ void write_t32_kernel(double2 *__restrict__ dst, size_t ntiles, double v) {
    const size_t warp = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    const size_t nwarps = ((size_t)gridDim.x * blockDim.x) >> 5;
    for (size_t t = warp; t < ntiles; t += nwarps) {
        double2 *g = dst + t * 256 + lane;
#pragma unroll
        for (int j = 0; j < 8; j++) g[j * 32] = make_double2(v, v + j);
    }
}
__global__ void read_kernel(const double2 *__restrict__ src, size_t n, double *out) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    double s = 0.0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const double2 u = __ldcg(src + i);
        s += u.x + u.y;
    }
    if (s == 123.456) out[0] = s;
}
__global__ void copy_kernel(const double2 *__restrict__ src, double2 *__restrict__ dst, size_t n) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) dst[i] = __ldcg(src + i);
}

template <class F>
static float best_of(F launch) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 6; rep++) {
        cudaEventRecord(e0);
        launch();
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (rep && ms < best) best = ms;
    }
    return best;
}

int main() {
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const size_t bytes = (size_t)4 << 30, n = bytes / sizeof(double2);
    double2 *a, *b;
    double *out;
    if (cudaMalloc(&a, bytes) != cudaSuccess || cudaMalloc(&b, bytes) != cudaSuccess) return 1;
    cudaMalloc(&out, 8);
    cudaMemset(a, 0, bytes);
    cudaMemset(b, 0, bytes);
    const int threads = 256, blocks = sms * 8;
    const float w = best_of([&] { write_kernel<<<blocks, threads>>>(a, n, 1.0); });
    const float wt = best_of([&] { write_t32_kernel<<<sms * 16, 128>>>(a, n / 256, 1.0); });
    const float ms = best_of([&] { cudaMemsetAsync(a, 1, bytes); });
    const float r = best_of([&] { read_kernel<<<blocks, threads>>>(a, n, out); });
    const float c = best_of([&] { copy_kernel<<<blocks, threads>>>(a, b, n); });
    const float mc = best_of([&] { cudaMemcpyAsync(b, a, bytes, cudaMemcpyDeviceToDevice); });
    const double gb = bytes / 1e9;
    printf("{\"buffer_gib\": 4, \"write_gbs\": %.1f, \"write_t32_gbs\": %.1f, \"memset_gbs\": %.1f, \"read_gbs\": %.1f, "
           "\"copy_gbs\": %.1f, \"memcpy_d2d_gbs\": %.1f, \"err\": \"%s\"}\n",
           gb / w * 1e3, gb / wt * 1e3, gb / ms * 1e3, gb / r * 1e3, 2 * gb / c * 1e3, 2 * gb / mc * 1e3,
           cudaGetErrorString(cudaGetLastError()));
    return 0;
}
