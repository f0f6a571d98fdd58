// Microbenchmark: peak FP64 rate of this GPU through (a) DFMA chains and (b) DMMA
// mma.sync.m8n8k4.f64 tiles.  Gives the compute roof used in DESIGN.md for K6.
#include <cuda_runtime.h>
#include <stdio.h>

template <int ILP>
__global__ void k_dfma(double *out, int iters, double a, double b) {
    double acc[ILP];
#pragma unroll
    for (int i = 0; i < ILP; i++) acc[i] = threadIdx.x * 1e-9 + i;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < ILP; i++) acc[i] = fma(acc[i], a, b);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < ILP; i++) s += acc[i];
    if (s == 12345.678) out[0] = s;
}

__global__ void k_dmma(double *out, int iters, double a0, double b0) {
    double c[4][2];
#pragma unroll
    for (int i = 0; i < 4; i++) c[i][0] = c[i][1] = threadIdx.x * 1e-9;
    double a = a0 + threadIdx.x * 1e-12, b = b0;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 4; i++)
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                         : "+d"(c[i][0]), "+d"(c[i][1])
                         : "d"(a), "d"(b));
    }
    double s = 0;
    for (int i = 0; i < 4; i++) s += c[i][0] + c[i][1];
    if (s == 12345.678) out[0] = s;
}

int main() {
    double *out;
    cudaMalloc(&out, 8);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const int iters = 20000;
    for (int threads : {256, 512, 1024}) {
        int blocks = sms * (2048 / threads);
        for (int rep = 0; rep < 2; rep++) {
            cudaEventRecord(e0);
            k_dfma<8><<<blocks, threads>>>(out, iters, 1.0000001, 1e-9);
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            double flops = 2.0 * 8 * iters * (double)blocks * threads;
            if (rep) printf("DFMA  threads=%4d blocks=%d: %.3f ms  %.2f TFLOP/s\n", threads, blocks, ms, flops / ms / 1e9);
        }
    }
    for (int threads : {256, 512, 1024}) {
        int blocks = sms * (2048 / threads);
        for (int rep = 0; rep < 2; rep++) {
            cudaEventRecord(e0);
            k_dmma<<<blocks, threads>>>(out, iters, 1.0000001, 1e-9);
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            double flops = 2.0 * 8 * 8 * 4 * 4 * iters * (double)blocks * (threads / 32);
            if (rep) printf("DMMA  threads=%4d blocks=%d: %.3f ms  %.2f TFLOP/s\n", threads, blocks, ms, flops / ms / 1e9);
        }
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
