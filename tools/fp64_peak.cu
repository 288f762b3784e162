// fp64_peak.cu -- measured FP64 ceilings of this GPU: the DMMA (mma.sync m8n8k4 f64) pipe and the DFMA pipe.
// SURVEY.md 7.1 asks for the dense-path fraction against a MEASURED peak, not the data-sheet number.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_peak tools/fp64_peak.cu && ./fp64_peak
// Prints one JSON line: {"dmma_tflops": .., "dfma_tflops": .., "sm_mhz": ..}.  Register-resident operands, 8
// independent accumulator chains per warp, enough warps to fill every SM sub-partition.
#include <cuda_runtime.h>
#include <stdio.h>

__global__ void dmma_loop(double *out, int iters) {
    double a = 1.0 + threadIdx.x * 1e-9, b = 1.0 - threadIdx.x * 1e-9;
    double d[16];
#pragma unroll
    for (int k = 0; k < 16; k++) d[k] = 0.0;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int k = 0; k < 8; k++)
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                         : "+d"(d[2 * k]), "+d"(d[2 * k + 1])
                         : "d"(a), "d"(b));
    }
    double s = 0.0;
#pragma unroll
    for (int k = 0; k < 16; k++) s += d[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void dfma_loop(double *out, int iters) {
    double a = 1.0 + threadIdx.x * 1e-9, b = 1e-9;
    double d[16];
#pragma unroll
    for (int k = 0; k < 16; k++) d[k] = k;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int k = 0; k < 16; k++) d[k] = fma(d[k], a, b);
    }
    double s = 0.0;
#pragma unroll
    for (int k = 0; k < 16; k++) s += d[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

int main() {
    int sms = 0, mhz = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    cudaDeviceGetAttribute(&mhz, cudaDevAttrClockRate, 0);
    const int threads = 256, blocks = sms * 8, iters = 20000;
    double *out;
    cudaMalloc(&out, sizeof(double) * threads * blocks);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    float best_mma = 1e30f, best_fma = 1e30f;
    for (int rep = 0; rep < 5; rep++) {
        float ms;
        cudaEventRecord(e0);
        dmma_loop<<<blocks, threads>>>(out, iters);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms, e0, e1);
        if (rep && ms < best_mma) best_mma = ms;
        cudaEventRecord(e0);
        dfma_loop<<<blocks, threads>>>(out, iters);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms, e0, e1);
        if (rep && ms < best_fma) best_fma = ms;
    }
    // one m8n8k4 = 8*8*4 FMA = 512 flop per warp; 8 per iteration
    const double warps = (double)blocks * threads / 32;
    const double mma_flops = warps * iters * 8.0 * 512.0, fma_flops = (double)blocks * threads * iters * 16.0 * 2.0;
    printf("{\"dmma_tflops\": %.3f, \"dfma_tflops\": %.3f, \"sms\": %d, \"sm_khz_attr\": %d, \"err\": \"%s\"}\n",
           mma_flops / (best_mma * 1e-3) / 1e12, fma_flops / (best_fma * 1e-3) / 1e12, sms, mhz,
           cudaGetErrorString(cudaGetLastError()));
    return 0;
}
