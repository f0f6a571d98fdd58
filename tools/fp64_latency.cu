// Dependent-chain latencies (cycles per operation, one warp, one lane's chain) of the FP64
// operations the ensemble kernel's serial phases are made of.  nvcc -arch=sm_100a -O3 -fmad=false
#include <cstdio>
#include <cuda_runtime.h>
#define N 256
template <int OP>
__global__ void k(double *out, double a, double b, long long *cyc) {
    double x = a;
    long long t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N; i++) {
        if (OP == 0) x = x + b;
        if (OP == 1) x = x * b;
        if (OP == 2) x = fma(x, b, a);
        if (OP == 3) x = sqrt(x + 2.0);
        if (OP == 4) x = a / (x + 3.0);
        if (OP == 5) x = sin(x + 0.3);
        if (OP == 6) x = log(x + 2.0);
        if (OP == 7) x = atan2(x, b);
        if (OP == 8) x = __shfl_sync(0xffffffffu, x, (threadIdx.x + 1) & 31);
        if (OP == 9) { __shared__ double s[64]; s[threadIdx.x] = x; __syncwarp(); x = s[(threadIdx.x + 1) & 31]; __syncwarp(); }
        if (OP == 10) { __syncthreads(); }
        if (OP == 11) x = (double)(((int)x + 7) / ((int)b + 3));
    }
    long long t1 = clock64();
    out[threadIdx.x] = x;
    if (threadIdx.x == 0) *cyc = t1 - t0;
}
int main() {
    double *o; long long *c, h;
    cudaMalloc(&o, 8 * 256); cudaMalloc(&c, 8);
    const char *nm[] = {"DADD", "DMUL", "DFMA", "sqrt(+add)", "div(+add)", "sin(+add)", "log(+add)", "atan2", "SHFL", "STS+LDS+2 syncwarp", "__syncthreads (128 thr)", "int div (+cvt)"};
#define RUN(OP, T) k<OP><<<1, T>>>(o, 1.0001, 1.0000001, c); k<OP><<<1, T>>>(o, 1.0001, 1.0000001, c); cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost); printf("%-26s %7.1f cycles\n", nm[OP], (double)h / N);
    RUN(0, 32) RUN(1, 32) RUN(2, 32) RUN(3, 32) RUN(4, 32) RUN(5, 32) RUN(6, 32) RUN(7, 32) RUN(8, 32) RUN(9, 32) RUN(10, 128) RUN(11, 32)
    return 0;
}
