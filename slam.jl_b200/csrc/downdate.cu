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
//     consecutive doubles (16-byte vector loads for ROWS = 2).  Columns move through a
//     two-deep register pipeline of U columns each: the loads of batch b+1 are issued
//     before the FMAs of batch b, so every warp keeps U*ROWS*8*32 bytes in flight while
//     it computes;
//   * each element is a chain  a = fma(-W1[i,k], W1[j,k], a), k ascending -- the same
//     order as the CPU oracle's twin, so results agree to the last bit for equal W1.
//   Entries below the diagonal are never read by any consumer, so diagonal-crossing tiles
//   need no per-element predicate: threads simply start at their own diagonal column.
//
// Ranks above 32 run as ceil(R/32) passes of the FMA kernel or through the DMMA kernel
// (mma.sync.m8n8k4.f64), chosen by launch_downdate().
#include <stdlib.h>

#include "common.cuh"
#include "kernels.cuh"

namespace {

constexpr int DD_THREADS = 128;

// streaming (evict-first) vector accesses; volatile asm keeps the issue order: all loads of
// the next batch before the FMAs of the current one
template <int ROWS>
__device__ __forceinline__ void ld_rows(double (&a)[ROWS], const double *p) {
    if (ROWS == 2) {
        asm volatile("ld.global.cs.v2.f64 {%0, %1}, [%2];" : "=d"(a[0]), "=d"(a[ROWS - 1]) : "l"(p));
    } else {
        asm volatile("ld.global.cs.f64 %0, [%1];" : "=d"(a[0]) : "l"(p));
    }
}
template <int ROWS>
__device__ __forceinline__ void st_rows(const double (&a)[ROWS], double *p) {
    if (ROWS == 2) {
        asm volatile("st.global.cs.v2.f64 [%0], {%1, %2};" ::"l"(p), "d"(a[0]), "d"(a[ROWS - 1]) : "memory");
    } else {
        asm volatile("st.global.cs.f64 [%0], %1;" ::"l"(p), "d"(a[0]) : "memory");
    }
}

template <int RC, int ROWS, int U>
__device__ __forceinline__ void load_batch(double (&a)[U][ROWS], const double *col, size_t ld) {
#pragma unroll
    for (int u = 0; u < U; u++) ld_rows<ROWS>(a[u], col + (size_t)u * ld);
}

template <int RC, int ROWS, int U>
__device__ __forceinline__ void fma_store_batch(double (&a)[U][ROWS], const double (&wi)[ROWS][RC],
                                                const double *sWj, double *col, size_t ld) {
#pragma unroll
    for (int u = 0; u < U; u++) {
        const double2 *wj = reinterpret_cast<const double2 *>(sWj + (size_t)u * (RC + W1_PAD));
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
    for (int u = 0; u < U; u++) st_rows<ROWS>(a[u], col + (size_t)u * ld);
}

// columns per tile (W1 column-side rows staged in shared memory): the row-side W1 values a thread keeps in
// registers are fetched once per tile, so a wider tile amortises them further; bounded by 48 KB of static
// shared memory (BJ * RC * 8 bytes)
#ifndef DD_BJ_WIDE
#define DD_BJ_WIDE 128
#endif
template <int RC>
struct DdBj { static constexpr int value = RC <= 16 ? DD_BJ_WIDE : 128; };

// One warp-wide L2 prefetch covers a whole batch: lane l asks for 128-byte line (l % LINES) of
// column (l / LINES) of the batch, LINES = lines of this warp's row segment per column.
template <int ROWS, int U>
__device__ __forceinline__ void prefetch_batch(const double *warp_col, size_t ld, int lane, int cols_left) {
    constexpr int LINES = ROWS * 2;            // 32 lanes * ROWS * 8 B = ROWS*256 B per column
    constexpr int COLS_PER_INSTR = 32 / LINES;
#pragma unroll
    for (int u0 = 0; u0 < U; u0 += COLS_PER_INSTR) {
        const int u = u0 + lane / LINES;
        const double *p = warp_col + (size_t)u * ld + (lane % LINES) * 16;
        if (u < cols_left) asm volatile("prefetch.global.L2::evict_last [%0];" ::"l"(p));
    }
}

// Rows [r0, r1) of the upper triangle; element (i, j) lives at Pb[(i - r0) + j*ldb].
// The single-filter path passes r0 = 0, r1 = n, Pb = P.
// W1r: [n_pad][RC + W1_PAD] row-major (column-side rows, staged by one bulk copy per tile);
// W1t: [RC][n_pad] rank-major (row-side values, coalesced register loads).
// PIPE = 1 adds a two-deep register pipeline on top of the L2 prefetch: the loads of batch b+1 are issued before
// the FMAs of batch b.  It pays where the kernel is held at two CTAs per SM by its W1 rows anyway (RC >= 24:
// 2 x RC doubles per thread), so the second buffer costs no occupancy.
template <int RC, int ROWS, int U, int PD, int PIPE>
__global__ void __launch_bounds__(DD_THREADS)
k_downdate_fma(double *__restrict__ Pb, size_t ldb, RowMap map, int n, int n_pad,
               const double *__restrict__ W1r, const double *__restrict__ W1t) {
    constexpr int BI = DD_THREADS * ROWS;
    static_assert(SH_PH % BI == 0, "a CTA's rows must not straddle two row panels");
    int i0, loc0, r1;
    if (!rowmap_tile(map, blockIdx.x, BI, n, &i0, &loc0, &r1)) return;
    const int r0 = i0 - loc0;                         // local row of global row i is i - r0 inside this tile
    constexpr int DD_BJ = DdBj<RC>::value;
    const int j0 = blockIdx.y * DD_BJ;
    const int jend = min(j0 + DD_BJ, n);
    if (i0 >= jend) return;  // tile entirely below the diagonal

    constexpr int LDK = RC + W1_PAD;
    __shared__ __align__(128) double sW[DD_BJ * LDK];
    __shared__ __align__(8) uint64_t bar;
    if (threadIdx.x == 0) {
        mbar_init(&bar, 1);
        mbar_fence_init();
    }
    __syncthreads();
    pdl_wait();                                       // W1 (K5) and P (the previous update) are complete
    pdl_launch_dependents();
    if (threadIdx.x == 0) {
        mbar_expect_tx(&bar, DD_BJ * LDK * 8);
        bulk_g2s(sW, W1r + (size_t)j0 * LDK, DD_BJ * LDK * 8, &bar);
    }

    const int lane = threadIdx.x & 31;
    const int iw = i0 + (threadIdx.x & ~31) * ROWS;   // first row of this warp
    if (iw >= jend || iw >= r1) return;               // whole warp below the tile / outside the block
    const int i = i0 + threadIdx.x * ROWS;

    // the warp walks the columns from its own diagonal (rounded down to a batch) to the tile end
    int jstart = max(j0, iw);
    jstart = j0 + ((jstart - j0) / U) * U;
    const int nfull = (jend - jstart) / U;
    const double *wcol = Pb + (size_t)(iw - r0) + (size_t)jstart * ldb;
    if (PD > 0) {
#pragma unroll
        for (int b = 0; b < PD; b++)
            if (b < nfull) prefetch_batch<ROWS, U>(wcol + (size_t)(b * U) * ldb, ldb, lane, jend - jstart - b * U);
    }

    double wi[ROWS][RC];
#pragma unroll
    for (int k = 0; k < RC; k++) {
        if (ROWS == 2 && (i & 1) == 0) {              // (uniform over the CTA: i0 and the tile height are fixed)
            double2 t = __ldg(reinterpret_cast<const double2 *>(W1t + (size_t)k * n_pad + i));
            wi[0][k] = -t.x;
            wi[ROWS - 1][k] = -t.y;
        } else if (ROWS == 2) {   // row panels of the sharded path start at odd rows
            wi[0][k] = -__ldg(W1t + (size_t)k * n_pad + i);
            // a ragged last warp may stand on row n_pad - 1: its second row does not exist (and is never used)
            wi[ROWS - 1][k] = i + 1 < n_pad ? -__ldg(W1t + (size_t)k * n_pad + i + 1) : 0.0;
        } else {
            wi[0][k] = -__ldg(W1t + (size_t)k * n_pad + i);
        }
    }
    mbar_wait(&bar, 0);

    double *col = Pb + (size_t)(i - r0) + (size_t)jstart * ldb;
    const double *sWj = sW + (size_t)(jstart - j0) * LDK;
    const bool row_ok = i < r1;   // rows past the block end (ragged last warp) do nothing
    if (PIPE) {
        double a0[U][ROWS], a1[U][ROWS];
        if (row_ok && nfull > 0) load_batch<RC, ROWS, U>(a0, col, ldb);
        for (int b = 0; b < nfull; b += 2) {
            if (PD > 0) {
                if (b + PD < nfull) prefetch_batch<ROWS, U>(wcol + (size_t)((b + PD) * U) * ldb, ldb, lane, jend - jstart - (b + PD) * U);
                if (b + 1 + PD < nfull) prefetch_batch<ROWS, U>(wcol + (size_t)((b + 1 + PD) * U) * ldb, ldb, lane, jend - jstart - (b + 1 + PD) * U);
            }
            if (row_ok) {
                if (b + 1 < nfull) load_batch<RC, ROWS, U>(a1, col + (size_t)U * ldb, ldb);
                fma_store_batch<RC, ROWS, U>(a0, wi, sWj, col, ldb);
                if (b + 2 < nfull) load_batch<RC, ROWS, U>(a0, col + (size_t)(2 * U) * ldb, ldb);
                if (b + 1 < nfull) fma_store_batch<RC, ROWS, U>(a1, wi, sWj + (size_t)U * LDK, col + (size_t)U * ldb, ldb);
            }
            col += (size_t)(2 * U) * ldb;
            sWj += (size_t)(2 * U) * LDK;
        }
        if (nfull & 1) {   // the loop stepped one batch past the end
            col -= (size_t)U * ldb;
            sWj -= (size_t)U * LDK;
        }
    } else {
        for (int b = 0; b < nfull; b++) {
            if (PD > 0 && b + PD < nfull) prefetch_batch<ROWS, U>(wcol + (size_t)((b + PD) * U) * ldb, ldb, lane, jend - jstart - (b + PD) * U);
            if (row_ok) {
                double a[U][ROWS];
                load_batch<RC, ROWS, U>(a, col, ldb);
                fma_store_batch<RC, ROWS, U>(a, wi, sWj, col, ldb);
            }
            col += (size_t)U * ldb;
            sWj += (size_t)U * LDK;
        }
    }
    if (row_ok) {
        for (int j = jstart + nfull * U; j < jend; j++) {   // leftover columns of the last tile
#pragma unroll
            for (int r = 0; r < ROWS; r++) {
                double acc = col[r];
#pragma unroll
                for (int k = 0; k < RC; k++) acc = fma(wi[r][k], sWj[k], acc);
                col[r] = acc;
            }
            col += ldb;
            sWj += LDK;
        }
    }
}

struct DdCfg { int rows, u, pd, pipe; };

// experiment hook: SLAM_K6_CFG="rows,u,pd" overrides the tuned choice
DdCfg pick_cfg(int RC) {
    // measured on B200 at n = 20003 (profiles/r1_k6_sweep.md): the L2 prefetch pays once the FMA phase
    // is long enough to starve the register-staged loads (rank >= 8)
    // RC = 32 (254 registers, two CTAs per SM whatever it does): + a two-deep register pipeline, 0.724 -> 0.678 ms
    DdCfg c = RC <= 4 ? DdCfg{2, 8, 0, 0} : RC <= 16 ? DdCfg{2, 8, 1, 0} : RC <= 24 ? DdCfg{2, 4, 4, 0} : DdCfg{2, 8, 2, 1};
    if (const char *e = getenv("SLAM_K6_CFG")) {
        int r, u, pd, pipe = 0;
        if (sscanf(e, "%d,%d,%d,%d", &r, &u, &pd, &pipe) >= 3) c = DdCfg{r, u, pd, pipe};
    }
    return c;
}

template <int RC, int ROWS, int U, int PD, int PIPE>
void launch_one(double *Pb, size_t ldb, const RowMap &map, int n, int n_pad, const double *W1r,
                const double *W1t, cudaStream_t s) {
    constexpr int BI = DD_THREADS * ROWS;
    constexpr int DD_BJ = DdBj<RC>::value;
    dim3 grid((rowmap_local_rows(map) + BI - 1) / BI, (n + DD_BJ - 1) / DD_BJ);
    launch_pdl(k_downdate_fma<RC, ROWS, U, PD, PIPE>, grid, dim3(DD_THREADS), 0, s, Pb, ldb, map, n, n_pad, W1r, W1t);
}

template <int RC>
bool launch_cfg(const DdCfg &c, double *Pb, size_t ldb, const RowMap &map, int n, int n_pad, const double *W1r,
                const double *W1t, cudaStream_t s) {
#define TRY_CFG(R_, U_, PD_, PIPE_)                                                   \
    if (c.rows == R_ && c.u == U_ && c.pd == PD_ && c.pipe == PIPE_) {                  \
        launch_one<RC, R_, U_, PD_, PIPE_>(Pb, ldb, map, n, n_pad, W1r, W1t, s);        \
        return true;                                                                    \
    }
    TRY_CFG(2, 8, 0, 0) TRY_CFG(2, 8, 1, 0) TRY_CFG(2, 8, 2, 0)
    TRY_CFG(2, 4, 1, 0) TRY_CFG(2, 4, 2, 0) TRY_CFG(2, 4, 4, 0)
    TRY_CFG(1, 8, 2, 0)
    if (RC >= 24) {   // register-pipelined variants (two CTAs per SM either way)
        TRY_CFG(2, 4, 2, 1) TRY_CFG(2, 8, 2, 1)
    }
#undef TRY_CFG
    return false;
}

}  // namespace

int downdate_chunk_width(int R) {
    if (R <= 2) return 2;
    if (R <= 4) return 4;
    if (R <= 8) return 8;
    if (R <= 12) return 12;
    if (R <= 16) return 16;
    if (R <= 24) return 24;
    if (R <= 32) return 32;
    if (R <= 40) return 40;    // 33..64 ranks: one chunk, one pass of the streaming DMMA kernel
    if (R <= 48) return 48;
    if (R <= 56) return 56;
    if (R <= 64) return 64;
    return 32;                 // above: 32-wide chunks for the tile DMMA kernel (all chunks in one pass)
}

// W1c: per rank chunk c (RC columns, zero padded) two copies of W1's chunk (layout: kernels.cuh):
//   [n_pad][RC + W1_PAD] row-major, then [RC][n_pad] rank-major.
// n_pad is a multiple of DD_ROW_PAD and rows >= n are zero.
// Which kernel runs the downdate of R ranks: 1 = FMA, 2 = tiled DMMA, 3 = streaming DMMA (0 = none applies).
// variant 0 = automatic (measured cross-overs, profiles/r2_k6_stream.md): FMA up to 4 ranks, streaming DMMA for 8..64,
// tiled DMMA above; 1 / 2 / 3 force a kernel where it applies.  K5 asks too: only the FMA kernel reads the
// rank-major W1 copy.
int downdate_kernel_choice(int R, int variant, size_t ldb) {
    int min_rank = 33;     // above one FMA pass (32 ranks): 19-30 TFLOP/s DMMA vs 13-18 TFLOP/s FMA
    if (const char *e = getenv("SLAM_K6_DMMA_MIN_RANK")) min_rank = atoi(e);
    int stream_min = 8;    // R = 8: 0.505 vs 0.534 ms (FMA) at n = 20003; at 4 ranks the FMA kernel is level or ahead
    if (const char *e = getenv("SLAM_K6_STREAM_MIN_RANK")) stream_min = atoi(e);
    const int RC = downdate_chunk_width(R);
    // (33..64 ranks are laid out as one 40..64-wide chunk, which only the streaming kernel reads)
    if ((variant == 3 || (variant == 0 && R >= stream_min) || RC > 32) && downdate_stream_usable(R, ldb)) return 3;
    if ((variant == 2 || (variant == 0 && R >= min_rank)) && downdate_dmma_usable(R, ldb)) return 2;
    return RC <= 32 ? 1 : 0;
}

int launch_downdate_rows(double *Pb, size_t ldb, int r0, int r1, int n, const double *W1c, int n_pad,
                         int R, int variant, cudaStream_t s, long long *launches) {
    return launch_downdate_map(Pb, ldb, rowmap_rows(r0, r1), n, W1c, n_pad, R, variant, s, launches);
}

int launch_downdate_map(double *Pb, size_t ldb, RowMap map, int n, const double *W1c, int n_pad, int R, int variant,
                        cudaStream_t s, long long *launches) {
    if (rowmap_local_rows(map) <= 0) return SLAM_OK;
    const int RC = downdate_chunk_width(R);
    const int which = downdate_kernel_choice(R, variant, ldb);
    if (which == 3) return launch_downdate_stream_rows(Pb, ldb, map, n, W1c, n_pad, R, s, launches);
    if (which == 2) return launch_downdate_dmma_rows(Pb, ldb, map, n, W1c, n_pad, R, s, launches);
    if (which == 0) {
        slam_set_error("launch_downdate: no kernel for a %d-wide rank chunk", RC);
        return SLAM_ERR_ARG;
    }
    const int nchunks = (R + RC - 1) / RC;
    const DdCfg cfg = pick_cfg(RC);
    for (int c = 0; c < nchunks; c++) {
        const double *Wr = W1c + (size_t)c * w1_chunk_doubles(n_pad, RC);
        const double *Wt = Wr + w1_rank_major_offset(n_pad, RC);
        bool ok = false;
        switch (RC) {
            case 2: ok = launch_cfg<2>(cfg, Pb, ldb, map, n, n_pad, Wr, Wt, s); break;
            case 4: ok = launch_cfg<4>(cfg, Pb, ldb, map, n, n_pad, Wr, Wt, s); break;
            case 8: ok = launch_cfg<8>(cfg, Pb, ldb, map, n, n_pad, Wr, Wt, s); break;
            case 12: ok = launch_cfg<12>(cfg, Pb, ldb, map, n, n_pad, Wr, Wt, s); break;
            case 16: ok = launch_cfg<16>(cfg, Pb, ldb, map, n, n_pad, Wr, Wt, s); break;
            case 24: ok = launch_cfg<24>(cfg, Pb, ldb, map, n, n_pad, Wr, Wt, s); break;
            default: ok = launch_cfg<32>(cfg, Pb, ldb, map, n, n_pad, Wr, Wt, s); break;
        }
        if (!ok) {
            slam_set_error("launch_downdate: no kernel for config rows=%d u=%d pd=%d", cfg.rows, cfg.u, cfg.pd);
            return SLAM_ERR_ARG;
        }
        if (launches) (*launches)++;
    }
    CUDA_TRY(cudaGetLastError());
    return SLAM_OK;
}

int launch_downdate(double *P, size_t ld, int n, const double *W1c, int n_pad, int R, int variant,
                    cudaStream_t s, long long *launches) {
    return launch_downdate_rows(P, ld, 0, n, n, W1c, n_pad, R, variant, s, launches);
}
