// K6 -- the symmetric rank-2m downdate  P -= W1*W1'  of update() (src/ekf.jl:75), in place
// over the authoritative upper triangle of the column-major covariance.
//
// Roofline: HBM.  Algorithmic bytes per launch = 8*n*(n+1) (each of the n(n+1)/2 upper
// doubles read once and written once); flops = 2*R*n(n+1)/2 with R = 2m.
//
// FMA kernel (rank chunk RC <= 32):
//   * a CTA owns a tile of BI = 128*ROWS rows x BJ columns; a thread owns ROWS adjacent
//     rows i and keeps W1[i, 0:RC] in registers for the whole tile;
//   * the W1 rows of the tile's BJ columns are one contiguous block of the row-major W1
//     and are staged into shared memory by a single 1-D bulk async copy (TMA engine,
//     cp.async.bulk + mbarrier); the inner loop reads them as warp-wide broadcasts;
//   * P is streamed column by column: for a fixed column the 128 threads read 128*ROWS
//     consecutive doubles (16-byte vector loads for ROWS = 2), U columns in flight;
//   * each element is a chain  a = fma(-W1[i,k], W1[j,k], a), k ascending -- the same
//     order as the CPU oracle's twin, so results agree to the last bit for equal W1.
//   Entries below the diagonal are never read by any consumer, so diagonal-crossing tiles
//   need no per-element predicate: threads simply start at their own diagonal column.
//
// Ranks above 32 run as ceil(R/32) passes of the FMA kernel or through the DMMA kernel
// (mma.sync.m8n8k4.f64), chosen by launch_downdate().
#include "common.cuh"
#include "kernels.cuh"

namespace {

constexpr int DD_THREADS = 128;
constexpr int DD_BJ = 64;

template <int RC, int ROWS, int U>
__global__ void __launch_bounds__(DD_THREADS)
k_downdate_fma(double *__restrict__ P, size_t ld, int n, const double *__restrict__ W1) {
    constexpr int BI = DD_THREADS * ROWS;
    const int i0 = blockIdx.x * BI;
    const int j0 = blockIdx.y * DD_BJ;
    const int jend = min(j0 + DD_BJ, n);
    if (i0 >= jend) return;  // tile entirely below the diagonal

    __shared__ __align__(128) double sW[DD_BJ * RC];
    __shared__ __align__(8) uint64_t bar;
    if (threadIdx.x == 0) {
        mbar_init(&bar, 1);
        mbar_fence_init();
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        mbar_expect_tx(&bar, DD_BJ * RC * 8);
        bulk_g2s(sW, W1 + (size_t)j0 * RC, DD_BJ * RC * 8, &bar);
    }

    const int i = i0 + threadIdx.x * ROWS;
    if (i >= jend) return;  // this thread's rows lie below every column of the tile

    double wi[ROWS][RC];
#pragma unroll
    for (int r = 0; r < ROWS; r++) {
        const double2 *src = reinterpret_cast<const double2 *>(W1 + (size_t)(i + r) * RC);
#pragma unroll
        for (int k = 0; k < RC / 2; k++) {
            double2 t = __ldg(src + k);
            wi[r][2 * k] = -t.x;
            wi[r][2 * k + 1] = -t.y;
        }
    }
    mbar_wait(&bar, 0);

    // first column this thread touches: its own diagonal, rounded down to the unroll
    int jstart = max(j0, i);
    jstart = j0 + ((jstart - j0) / U) * U;
    double *col = P + (size_t)i + (size_t)jstart * ld;

    for (int j = jstart; j < jend; j += U) {
        if (j + U <= jend) {
            double a[U][ROWS];
#pragma unroll
            for (int u = 0; u < U; u++) {
                if (ROWS == 2) {
                    double2 t = __ldcs(reinterpret_cast<const double2 *>(col + (size_t)u * ld));
                    a[u][0] = t.x;
                    a[u][ROWS - 1] = t.y;
                } else {
                    a[u][0] = __ldcs(col + (size_t)u * ld);
                }
            }
#pragma unroll
            for (int u = 0; u < U; u++) {
                const double2 *wj = reinterpret_cast<const double2 *>(sW + (size_t)(j - j0 + u) * RC);
#pragma unroll
                for (int k = 0; k < RC / 2; k++) {
                    double2 w = wj[k];
#pragma unroll
                    for (int r = 0; r < ROWS; r++) {
                        a[u][r] = fma(wi[r][2 * k], w.x, a[u][r]);
                        a[u][r] = fma(wi[r][2 * k + 1], w.y, a[u][r]);
                    }
                }
            }
#pragma unroll
            for (int u = 0; u < U; u++) {
                if (ROWS == 2) {
                    __stcs(reinterpret_cast<double2 *>(col + (size_t)u * ld),
                           make_double2(a[u][0], a[u][ROWS - 1]));
                } else {
                    __stcs(col + (size_t)u * ld, a[u][0]);
                }
            }
        } else {
            for (int u = 0; j + u < jend; u++) {
                const double *wj = sW + (size_t)(j - j0 + u) * RC;
#pragma unroll
                for (int r = 0; r < ROWS; r++) {
                    double acc = col[(size_t)u * ld + r];
#pragma unroll
                    for (int k = 0; k < RC; k++) acc = fma(wi[r][k], wj[k], acc);
                    col[(size_t)u * ld + r] = acc;
                }
            }
        }
        col += (size_t)U * ld;
    }
}

template <int RC, int ROWS, int U>
void launch_fma(double *P, size_t ld, int n, const double *W1c, cudaStream_t s) {
    constexpr int BI = DD_THREADS * ROWS;
    dim3 grid((n + BI - 1) / BI, (n + DD_BJ - 1) / DD_BJ);
    k_downdate_fma<RC, ROWS, U><<<grid, DD_THREADS, 0, s>>>(P, ld, n, W1c);
}

}  // namespace

int downdate_chunk_width(int R) {
    if (R <= 2) return 2;
    if (R <= 4) return 4;
    if (R <= 8) return 8;
    if (R <= 12) return 12;
    if (R <= 16) return 16;
    if (R <= 24) return 24;
    return 32;
}

// W1c: nchunks blocks of [n_pad][RC] row-major (chunk c holds rank columns c*RC .. c*RC+RC-1,
// zero padded); n_pad is a multiple of DD_ROW_PAD and rows >= n are zero.
int launch_downdate(double *P, size_t ld, int n, const double *W1c, int n_pad, int R, int variant,
                    cudaStream_t s, long long *launches) {
    (void)variant;
    const int RC = downdate_chunk_width(R);
    const int nchunks = (R + RC - 1) / RC;
    for (int c = 0; c < nchunks; c++) {
        const double *Wc = W1c + (size_t)c * n_pad * RC;
        switch (RC) {
            case 2: launch_fma<2, 2, 8>(P, ld, n, Wc, s); break;
            case 4: launch_fma<4, 2, 8>(P, ld, n, Wc, s); break;
            case 8: launch_fma<8, 2, 8>(P, ld, n, Wc, s); break;
            case 12: launch_fma<12, 2, 8>(P, ld, n, Wc, s); break;
            case 16: launch_fma<16, 2, 8>(P, ld, n, Wc, s); break;
            case 24: launch_fma<24, 1, 8>(P, ld, n, Wc, s); break;
            default: launch_fma<32, 1, 8>(P, ld, n, Wc, s); break;
        }
        if (launches) (*launches)++;
    }
    CUDA_TRY(cudaGetLastError());
    return SLAM_OK;
}
