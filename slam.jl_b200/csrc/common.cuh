// Shared host/device helpers of libslamcuda (sm_100a only).
// Reference citations are file:line under /root/reference.
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/slamcuda.h"

// ---------------------------------------------------------------------------------------
// error plumbing
// ---------------------------------------------------------------------------------------
void slam_set_error(const char *fmt, ...);

#define CUDA_TRY(expr)                                                                   \
    do {                                                                                 \
        cudaError_t e__ = (expr);                                                        \
        if (e__ != cudaSuccess) {                                                        \
            slam_set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__),      \
                           __FILE__, __LINE__);                                          \
            return SLAM_ERR_CUDA;                                                        \
        }                                                                                \
    } while (0)

#define ARG_CHECK(cond, msg)                                                             \
    do {                                                                                 \
        if (!(cond)) {                                                                   \
            slam_set_error("%s: %s", __func__, msg);                                     \
            return SLAM_ERR_ARG;                                                         \
        }                                                                                \
    } while (0)

static inline int round_up(int a, int b) { return (a + b - 1) / b * b; }
static inline size_t round_up_sz(size_t a, size_t b) { return (a + b - 1) / b * b; }

// ---------------------------------------------------------------------------------------
// NVTX ranges around the C entry points (tracing): header-only nvtx3, a no-op unless a profiler
// (nsys / ncu --nvtx) injects itself.  One range per reference-level operation, named after the entry point.
// ---------------------------------------------------------------------------------------
#include <nvtx3/nvToolsExt.h>
struct SlamNvtxRange {
    explicit SlamNvtxRange(const char *name) { nvtxRangePushA(name); }
    ~SlamNvtxRange() { nvtxRangePop(); }
};
#define SLAM_NVTX() SlamNvtxRange nvtx_range__(__func__)

// ---------------------------------------------------------------------------------------
// device math shared by the single-filter, ensemble and sharded paths
// ---------------------------------------------------------------------------------------
#define SLAM_PI 3.14159265358979323846

// mpi_to_pi, src/common.jl:102-110: a SINGLE wrap, not a modulo.
__host__ __device__ __forceinline__ double mpi_to_pi(double phi) {
    if (phi > SLAM_PI) return phi - 2 * SLAM_PI;
    if (phi < -SLAM_PI) return phi + 2 * SLAM_PI;
    return phi;
}

// Nonzeros of the observation Jacobian of predict_observation (src/common.jl:139-165):
// h[0..2] = row 1, cols 1:3; h[3..5] = row 2, cols 1:3; h[6..7] = row 1, cols fpos:fpos+1;
// h[8..9] = row 2, cols fpos:fpos+1.  zp = [range; bearing - phi] (bearing NOT wrapped).
__host__ __device__ __forceinline__ void observe_model(double xv, double yv, double phi, double lx,
                                                       double ly, double *zp, double *h) {
    double dx = lx - xv;
    double dy = ly - yv;
    double d2 = dx * dx + dy * dy;
    double d = sqrt(d2);
    zp[0] = d;
    zp[1] = atan2(dy, dx) - phi;
    double xd = dx / d, yd = dy / d, xd2 = dx / d2, yd2 = dy / d2;
    h[0] = -xd;  h[1] = -yd;  h[2] = 0.0;
    h[3] = yd2;  h[4] = -xd2; h[5] = -1.0;
    h[6] = xd;   h[7] = yd;
    h[8] = -yd2; h[9] = xd2;
}

// Upper-triangle-authoritative access to a column-major symmetric matrix.
__device__ __forceinline__ double sym_at(const double *P, size_t ld, int i, int j) {
    return i <= j ? P[(size_t)i + (size_t)j * ld] : P[(size_t)j + (size_t)i * ld];
}

// ---------------------------------------------------------------------------------------
// Philox4x32-10 counter-based generator.  The reference draws from Julia's global RNG
// (sim/sim-utils.jl:5,36-37,68), which cannot be reproduced without Julia; "same seeds"
// is defined at this generator.  Counter = (idx, stream, run, 0), key = seed.
// ---------------------------------------------------------------------------------------
#define SLAM_PRAGMA_(x) _Pragma(#x)
#define SLAM_UNROLL(n) SLAM_PRAGMA_(unroll n)
#ifndef PHILOX_UNROLL
#define PHILOX_UNROLL 2    // rolled: the ensemble kernel is bound by instruction fetch, not by these 70 instructions
#endif
__host__ __device__ __forceinline__ void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
    SLAM_UNROLL(PHILOX_UNROLL)
    for (int r = 0; r < 10; r++) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c[0];
        uint64_t p1 = (uint64_t)0xCD9E8D57u * c[2];
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
}

// Two 53-bit uniforms in [0,1) from one Philox block.
__host__ __device__ __forceinline__ void philox_uniform2(uint64_t seed, uint32_t run, uint32_t stream,
                                                         uint32_t idx, double *u1, double *u2) {
    uint32_t c[4] = {idx, stream, run, 0u};
    philox4x32_10(c, (uint32_t)seed, (uint32_t)(seed >> 32));
    *u1 = ((double)(c[0] >> 5) * 67108864.0 + (double)(c[1] >> 6)) / 9007199254740992.0;
    *u2 = ((double)(c[2] >> 5) * 67108864.0 + (double)(c[3] >> 6)) / 9007199254740992.0;
}

// Standard normal by Box-Muller (cosine branch only, one normal per counter).
__host__ __device__ __forceinline__ double philox_normal(uint64_t seed, uint32_t run, uint32_t stream,
                                                         uint32_t idx) {
    double u1, u2;
    philox_uniform2(seed, run, stream, idx, &u1, &u2);
    return sqrt(-2.0 * log(1.0 - u1)) * cos(6.283185307179586 * u2);
}

// ---------------------------------------------------------------------------------------
// mbarrier + 1-D bulk async copy (TMA engine, SASS UBLKCP) helpers
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra.uni WAIT_DONE;\n"
        "bra.uni WAIT_LOOP;\n"
        "WAIT_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// global -> shared bulk copy; bytes % 16 == 0, both addresses 16-byte aligned.
__device__ __forceinline__ void bulk_g2s(void *smem_dst, const void *gmem_src, uint32_t bytes,
                                         uint64_t *bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
            smem_u32(smem_dst)),
        "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}

// ---------------------------------------------------------------------------------------
// Programmatic dependent launch (PDL).  The kernels of one update form a chain on one stream
// (K3+K4 -> K5 -> K6); launched with the programmatic-stream-serialization attribute, a kernel is
// scheduled while its predecessor drains, runs its own prologue (index arithmetic, mbarrier
// init) and blocks in pdl_wait() until the predecessor has completed and flushed its writes.
// pdl_wait() is a no-op for a kernel launched the ordinary way.
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

#ifdef __CUDACC__
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}
#endif

