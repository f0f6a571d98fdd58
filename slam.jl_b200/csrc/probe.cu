// Measured FP64 roofs of the device this library runs on: peak rate of (a) DFMA chains and (b)
// mma.sync.m8n8k4.f64 tiles (SASS DMMA).  bench.py quotes K6 above 32 ranks and the ensemble kernel
// against these, measured in the same run, instead of a datasheet number.
#include "common.cuh"
#include "kernels.cuh"

namespace {

template <int ILP>
__global__ void k_probe_dfma(double *out, int iters, double a, double b) {
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

__global__ void k_probe_dmma(double *out, int iters, double a0, double b0) {
    double c[4][2];
#pragma unroll
    for (int i = 0; i < 4; i++) c[i][0] = c[i][1] = threadIdx.x * 1e-9;
    const double a = a0 + threadIdx.x * 1e-12, b = b0;
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

}  // namespace

extern "C" int slam_fp64_peak(int device, double *dfma_tflops, double *dmma_tflops) {
    int rc = slam_require_device(device);
    if (rc) return rc;
    double *out = nullptr;
    cudaEvent_t e0, e1;
    CUDA_TRY(cudaMalloc(&out, 8));
    CUDA_TRY(cudaEventCreate(&e0));
    CUDA_TRY(cudaEventCreate(&e1));
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    const int iters = 20000, threads = 512, blocks = sms * (2048 / threads);
    double best[2] = {0, 0};
    for (int which = 0; which < 2; which++)
        for (int rep = 0; rep < 3; rep++) {
            cudaEventRecord(e0);
            if (which == 0) k_probe_dfma<8><<<blocks, threads>>>(out, iters, 1.0000001, 1e-9);
            else k_probe_dmma<<<blocks, threads>>>(out, iters, 1.0000001, 1e-9);
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float ms = 0;
            cudaEventElapsedTime(&ms, e0, e1);
            const double flops = which == 0 ? 2.0 * 8 * iters * (double)blocks * threads
                                            : 2.0 * 8 * 8 * 4 * 4 * iters * (double)blocks * (threads / 32);
            if (rep && ms > 0) best[which] = fmax(best[which], flops / ms / 1e9);
        }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(out);
    CUDA_TRY(cudaGetLastError());
    if (dfma_tflops) *dfma_tflops = best[0];
    if (dmma_tflops) *dmma_tflops = best[1];
    return SLAM_OK;
}
