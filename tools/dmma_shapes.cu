// dmma_shapes.cu -- throughput of the FP64 mma.sync shapes on this GPU as a function of independent accumulator
// chains per warp and resident warps per SM (what the dense kernels can count on).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o dmma_shapes tools/dmma_shapes.cu && ./dmma_shapes
#include <cuda_runtime.h>
#include <stdio.h>

template <int SHAPE, int CHAINS>
__global__ void loop(double *out, int iters) {
    double a[8], b[4], d[CHAINS][4];
#pragma unroll
    for (int k = 0; k < 8; k++) a[k] = 1.0 + (threadIdx.x + k) * 1e-9;
#pragma unroll
    for (int k = 0; k < 4; k++) b[k] = 1.0 - (threadIdx.x + k) * 1e-9;
#pragma unroll
    for (int c = 0; c < CHAINS; c++)
#pragma unroll
        for (int k = 0; k < 4; k++) d[c][k] = 0.0;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int c = 0; c < CHAINS; c++) {
            if (SHAPE == 0)
                asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                             : "+d"(d[c][0]), "+d"(d[c][1]) : "d"(a[0]), "d"(b[0]));
            else if (SHAPE == 1)
                asm volatile("mma.sync.aligned.m16n8k4.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
                             : "+d"(d[c][0]), "+d"(d[c][1]), "+d"(d[c][2]), "+d"(d[c][3]) : "d"(a[0]), "d"(a[1]), "d"(b[0]));
            else if (SHAPE == 2)
                asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+d"(d[c][0]), "+d"(d[c][1]), "+d"(d[c][2]), "+d"(d[c][3])
                             : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(b[0]), "d"(b[1]));
            else
                asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, {%12,%13,%14,%15}, {%0,%1,%2,%3};"
                             : "+d"(d[c][0]), "+d"(d[c][1]), "+d"(d[c][2]), "+d"(d[c][3])
                             : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(a[4]), "d"(a[5]), "d"(a[6]), "d"(a[7]),
                               "d"(b[0]), "d"(b[1]), "d"(b[2]), "d"(b[3]));
        }
    }
    double s = 0.0;
#pragma unroll
    for (int c = 0; c < CHAINS; c++)
#pragma unroll
        for (int k = 0; k < 4; k++) s += d[c][k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int SHAPE, int CHAINS>
static double run(double *out, int sms, int warps_per_sm) {
    const int threads = 128, blocks = sms * warps_per_sm / 4, iters = 4000;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 4; rep++) {
        float ms;
        cudaEventRecord(e0);
        loop<SHAPE, CHAINS><<<blocks, threads>>>(out, iters);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms, e0, e1);
        if (rep && ms < best) best = ms;
    }
    const double flop_per = SHAPE == 0 ? 512.0 : SHAPE == 1 ? 1024.0 : SHAPE == 2 ? 2048.0 : 4096.0;
    return (double)blocks * (threads / 32) * iters * CHAINS * flop_per / (best * 1e-3) / 1e12;
}

template <int SHAPE>
static void shape(const char *name, double *out, int sms) {
    const int wps[] = {4, 8, 12, 16, 32};
    for (int w : wps)
        printf("{\"shape\": \"%s\", \"warps_per_sm\": %d, \"tflops_by_chains\": {\"1\": %.2f, \"2\": %.2f, \"4\": %.2f, \"8\": %.2f}}\n", name, w,
               run<SHAPE, 1>(out, sms, w), run<SHAPE, 2>(out, sms, w), run<SHAPE, 4>(out, sms, w), run<SHAPE, 8>(out, sms, w));
}

int main() {
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    double *out;
    cudaMalloc(&out, sizeof(double) * 128 * sms * 8);
    shape<0>("m8n8k4", out, sms);
    shape<1>("m16n8k4", out, sms);
    shape<2>("m16n8k8", out, sms);
    shape<3>("m16n8k16", out, sms);
    printf("{\"err\": \"%s\"}\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
