// K6' -- the symmetric rank-2m downdate  P -= W1*W1'  (src/ekf.jl:75) on the FP64 tensor pipe,
// for ranks where the update is a real dense contraction (R/8 flop per byte of P; the ridge of the
// B200 is ~5.7 flop/B, so R >~ 48).  mma.sync.aligned.m8n8k4.f64 (SASS DMMA).
//
// CTA = 128 threads = 2 x 2 warps on a 64 (rows i) x 64 (columns j) tile of the upper triangle;
// a warp owns 32 x 32 = 4 x 4 MMA tiles, 32 FP64 accumulators per thread, initialised from P
// and written back once: P is read and written exactly once for ALL rank chunks (the FMA kernel
// makes one pass per 32 ranks).  The MMA is issued on the TRANSPOSED tile, D'[j][i]: the
// accumulator fragment of a thread is then two consecutive ROWS of one column of the
// column-major P, i.e. one 16-byte load / store, and the four threads of a group cover 64
// contiguous bytes.  A[j][k] = -W1[j][k] (negated once when staged) and B[k][i] = W1[i][k] come
// from the row-major copy of W1 (both fragments have the same (row = lane/4, k = lane%4) shape),
// staged per 32-rank chunk in shared memory with a row stride of 36 doubles (bank-conflict free
// for the fragment pattern).
// Tiles entirely below the diagonal exit; a tile that straddles it is computed whole -- entries
// below the diagonal are not authoritative and no consumer reads them.
// On the B200 the four products of an m8n8k4 MMA are accumulated as a k-ascending fma chain, so
// the results are bit-identical to the FMA kernel and the oracle's twin (tests assert 1e-12, the
// measured difference is 0); PTX does not promise that order.
// Measured (n = 20003, profiles/r1_k6_dmma.md): 19.0 / 25.2 / 28.2 / 29.6 TFLOP/s at R = 48 / 64 /
// 128 / 256 against a measured DMMA roof of 37.1 TFLOP/s; the FMA kernel stays at 17.7.
#include "common.cuh"
#include "kernels.cuh"

namespace {

constexpr int MM_THREADS = 128;
constexpr int MM_TILE = 64;
constexpr int MM_KC = 32;        // rank chunk = chunk width of the W1 workspace for R > 24
constexpr int MM_LDK = 36;       // shared-memory row stride (doubles)
#ifndef MM_MIN_CTAS
#define MM_MIN_CTAS 4            // <= 128 registers: 4 CTAs (16 warps) per SM hide the tile load / store
#endif

__device__ __forceinline__ void dmma(double &d0, double &d1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};"
                 : "+d"(d0), "+d"(d1)
                 : "d"(a), "d"(b));
}

// W1c: per 32-rank chunk c, [n_pad][32 + W1_PAD] row-major at W1c + c*w1_chunk_doubles(n_pad, 32) (rows >= n are zero).
__global__ void __launch_bounds__(MM_THREADS, MM_MIN_CTAS)
k_downdate_dmma(double *__restrict__ Pb, size_t ldb, RowMap map, int n, int n_pad,
                const double *__restrict__ W1c, int nchunks) {
    static_assert(SH_PH % MM_TILE == 0, "a tile's rows must not straddle two row panels");
    int i0, loc0, r1;
    if (!rowmap_tile(map, blockIdx.x, MM_TILE, n, &i0, &loc0, &r1)) return;   // past the rows of the map
    const int r0 = i0 - loc0;                     // local row of global row i is i - r0 inside this tile
    const int j0 = blockIdx.y * MM_TILE;
    if (i0 >= j0 + MM_TILE) return;               // below the diagonal

    __shared__ __align__(16) double sJ[MM_TILE * MM_LDK];   // -W1 rows of the tile's columns
    __shared__ __align__(16) double sI[MM_TILE * MM_LDK];   //  W1 rows of the tile's rows

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, t = lane & 3;
    const int iw = i0 + (warp & 1) * 32, jw = j0 + (warp >> 1) * 32;

    // accumulators <- P (two consecutive rows of one column per fragment)
    double acc[4][4][2];
#pragma unroll
    for (int ja = 0; ja < 4; ja++) {
        const int j = jw + 8 * ja + g;
#pragma unroll
        for (int ib = 0; ib < 4; ib++) {
            const int i = iw + 8 * ib + 2 * t;
            double2 v = make_double2(0.0, 0.0);
            if (j < n && i + 1 < r1) {
                v = *reinterpret_cast<const double2 *>(Pb + (size_t)(i - r0) + (size_t)j * ldb);
            } else if (j < n && i < r1) {
                v.x = Pb[(size_t)(i - r0) + (size_t)j * ldb];
            }
            acc[ja][ib][0] = v.x;
            acc[ja][ib][1] = v.y;
        }
    }

    for (int c = 0; c < nchunks; c++) {
        const double *Wr = W1c + (size_t)c * w1_chunk_doubles(n_pad, MM_KC);
        __syncthreads();
        // stage the two 64 x 32 operand tiles (16-byte loads; W1 is small and L2 resident)
#pragma unroll
        for (int q = 0; q < (MM_TILE * MM_KC / 2) / MM_THREADS; q++) {
            const int e = tid + q * MM_THREADS;          // double2 index
            const int row = e >> 4, k2 = (e & 15) * 2;
            const int rj = min(j0 + row, n_pad - 1), ri = min(i0 + row, n_pad - 1);
            const double2 wj = __ldg(reinterpret_cast<const double2 *>(Wr + (size_t)rj * (MM_KC + W1_PAD) + k2));
            const double2 wi = __ldg(reinterpret_cast<const double2 *>(Wr + (size_t)ri * (MM_KC + W1_PAD) + k2));
            *reinterpret_cast<double2 *>(sJ + row * MM_LDK + k2) = make_double2(-wj.x, -wj.y);
            *reinterpret_cast<double2 *>(sI + row * MM_LDK + k2) = wi;
        }
        __syncthreads();
        const double *aJ = sJ + ((warp >> 1) * 32 + g) * MM_LDK + t;
        const double *bI = sI + ((warp & 1) * 32 + g) * MM_LDK + t;
#pragma unroll 2
        for (int k4 = 0; k4 < MM_KC / 4; k4++) {
            double a[4], b[4];
#pragma unroll
            for (int q = 0; q < 4; q++) {
                a[q] = aJ[q * 8 * MM_LDK + 4 * k4];
                b[q] = bI[q * 8 * MM_LDK + 4 * k4];
            }
#pragma unroll
            for (int ja = 0; ja < 4; ja++)
#pragma unroll
                for (int ib = 0; ib < 4; ib++) dmma(acc[ja][ib][0], acc[ja][ib][1], a[ja], b[ib]);
        }
    }

#pragma unroll
    for (int ja = 0; ja < 4; ja++) {
        const int j = jw + 8 * ja + g;
#pragma unroll
        for (int ib = 0; ib < 4; ib++) {
            const int i = iw + 8 * ib + 2 * t;
            if (j < n && i + 1 < r1) {
                *reinterpret_cast<double2 *>(Pb + (size_t)(i - r0) + (size_t)j * ldb) =
                    make_double2(acc[ja][ib][0], acc[ja][ib][1]);
            } else if (j < n && i < r1) {
                Pb[(size_t)(i - r0) + (size_t)j * ldb] = acc[ja][ib][0];
            }
        }
    }
}

}  // namespace

// Same contract as launch_downdate_rows (downdate.cu); needs the 32-wide chunk layout (R > 24) and an
// even local row origin / leading dimension for the 16-byte accesses.
bool downdate_dmma_usable(int R, size_t ldb) { return downdate_chunk_width(R) == MM_KC && (ldb % 2) == 0; }

int launch_downdate_dmma_rows(double *Pb, size_t ldb, RowMap map, int n, const double *W1c, int n_pad, int R,
                              cudaStream_t s, long long *launches) {
    if (rowmap_local_rows(map) <= 0) return SLAM_OK;
    const int nchunks = (R + MM_KC - 1) / MM_KC;
    dim3 grid((rowmap_local_rows(map) + MM_TILE - 1) / MM_TILE, (n + MM_TILE - 1) / MM_TILE);
    k_downdate_dmma<<<grid, MM_THREADS, 0, s>>>(Pb, ldb, map, n, n_pad, W1c, nchunks);
    if (launches) (*launches)++;
    CUDA_TRY(cudaGetLastError());
    return SLAM_OK;
}
