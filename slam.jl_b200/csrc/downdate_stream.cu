// K6s -- the symmetric rank-2m downdate  P -= W1*W1'  (src/ekf.jl:75) as a STREAMING FP64 tensor-core kernel for
// 12..64 ranks: one pass over the authoritative upper triangle of P for all ranks, P moved at the HBM rate while
// mma.sync.aligned.m8n8k4.f64 (SASS DMMA) does the arithmetic.
//
// Why (profiles/r1_k6_sweep.md, r1_k6_dmma.md): between 24 and 64 ranks the FMA kernel is held by its operands --
// 2 x RC row values per thread in registers (254 registers at 32 ranks) and one shared-memory broadcast per four
// DFMA -- and reaches 0.72 of the HBM roof at 32 ranks, two passes above; the 64 x 64-tile DMMA kernel re-stages
// both operand tiles from L2 for every tile and every 32-rank chunk and reaches 0.51 / 0.68 of the DMMA roof at
// 48 / 64 ranks.  Here
//   * a warp owns 8*IB consecutive rows for a whole BJ-column tile and keeps its row-side W1 fragments -- the B
//     operand of every MMA it will issue, IB x RC/4 doubles per lane -- in registers, negated once;
//   * the tile's BJ column-side W1 rows are ONE contiguous block of the row-major W1 copy, whose rows K5 pads to
//     RC + 4 doubles: a single 1-D bulk async copy (TMA engine, SASS UBLKCP) lands them in shared memory with a
//     bank-conflict-free stride for the (row = lane/4, rank = lane%4) fragment reads -- one LDS.64 per IB DMMA;
//   * P streams 8 columns at a time.  The MMA is issued on the TRANSPOSED tile, D'[j][i] += W1[j][k] * (-W1[i][k]),
//     so a lane's accumulator pair is two consecutive rows of one column of the column-major P: one 16-byte
//     ld.global.cs / st.global.cs, the four lanes of a group covering 64 contiguous bytes.  The next 8 columns
//     are loaded into a second accumulator set before the MMAs of the current ones are issued, and one warp-wide
//     L2 prefetch per 8 columns runs PD blocks further ahead (as in the FMA kernel);
//   * every element is updated by the chain a = fma(-W1[i,k], W1[j,k], a) with k ascending: on the B200 the four
//     products of an m8n8k4 MMA accumulate in k order, so the result is bit-identical to the FMA kernel and to the
//     oracle's C twin (measured difference 0; tests assert 1e-12 because PTX does not promise the order).
// A warp starts at its own diagonal (rounded down to 8 columns); the entries of that first block that lie below
// the diagonal are updated blindly -- they are not authoritative and nothing reads them.
//
// Roofline: HBM up to ~44 ranks (8 n (n+1) bytes), the FP64 tensor pipe above (R n (n+1) flops at the measured
// 37.1 TFLOP/s DMMA rate).  Measurements: profiles/r2_k6_stream.md.
#include <stdlib.h>

#include "common.cuh"
#include "kernels.cuh"

namespace {

__device__ __forceinline__ void dmma(double2 &d, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};"
                 : "+d"(d.x), "+d"(d.y)
                 : "d"(a), "d"(b));
}
__device__ __forceinline__ double2 ld_cs2(const double *p) {
    double2 v;
    asm volatile("ld.global.cs.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p));
    return v;
}
__device__ __forceinline__ double ld_cs1(const double *p) {
    double v;
    asm volatile("ld.global.cs.f64 %0, [%1];" : "=d"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ void st_cs2(double *p, double2 v) {
    asm volatile("st.global.cs.v2.f64 [%0], {%1, %2};" ::"l"(p), "d"(v.x), "d"(v.y) : "memory");
}
__device__ __forceinline__ void st_cs1(double *p, double v) {
    asm volatile("st.global.cs.f64 [%0], %1;" ::"l"(p), "d"(v) : "memory");
}

// the lane's IB accumulator pairs of one 8-column block: rows (i, i+1), i = i_first + 8 ib, of the lane's column
template <int IB>
__device__ __forceinline__ void ld_block(double2 (&v)[IB], const double *p, bool fast, bool colok, int i_first, int r1) {
    if (fast) {
#pragma unroll
        for (int ib = 0; ib < IB; ib++) v[ib] = ld_cs2(p + 8 * ib);
    } else {
#pragma unroll
        for (int ib = 0; ib < IB; ib++) {
            const int i = i_first + 8 * ib;
            v[ib] = make_double2(0.0, 0.0);
            if (colok && i + 1 < r1) v[ib] = ld_cs2(p + 8 * ib);
            else if (colok && i < r1) v[ib].x = ld_cs1(p + 8 * ib);
        }
    }
}
template <int IB>
__device__ __forceinline__ void st_block(const double2 (&v)[IB], double *p, bool fast, bool colok, int i_first, int r1) {
    if (fast) {
#pragma unroll
        for (int ib = 0; ib < IB; ib++) st_cs2(p + 8 * ib, v[ib]);
    } else {
#pragma unroll
        for (int ib = 0; ib < IB; ib++) {
            const int i = i_first + 8 * ib;
            if (colok && i + 1 < r1) st_cs2(p + 8 * ib, v[ib]);
            else if (colok && i < r1) st_cs1(p + 8 * ib, v[ib].x);
        }
    }
}

// The rows of `map` of the upper triangle; element (i, j) lives at Pb[local row of i + j*ldb] (the single filter
// passes rows [0, n) of P itself; the sharded filter its stack of row panels).  W1r: [n_pad][RC + W1_PAD] row-major,
// rows >= n zero.
template <int RC, int IB, int WARPS, int BJ, int MINB>
__global__ void __launch_bounds__(WARPS * 32, MINB)
k_downdate_stream(double *__restrict__ Pb, size_t ldb, RowMap map, int n, int n_pad, const double *__restrict__ W1r,
                  int pd) {
    constexpr int LDK = RC + W1_PAD, KS = RC / 4, ROWS_W = 8 * IB, BI = WARPS * ROWS_W;
    static_assert(RC % 4 == 0 && BJ % 8 == 0 && 256 % BJ == 0 && SH_PH % BI == 0, "tile shape");
    extern __shared__ __align__(128) double sA[];               // [BJ][LDK], then the mbarrier
    uint64_t *bar = reinterpret_cast<uint64_t *>(sA + BJ * LDK);
    int i0, loc0, r1;
    if (!rowmap_tile(map, blockIdx.x, BI, n, &i0, &loc0, &r1)) return;
    const int r0 = i0 - loc0;                                    // local row of global row i is i - r0 inside this tile
    const int j0 = blockIdx.y * BJ;
    const int jend = min(j0 + BJ, n);
    if (i0 >= jend) return;                                      // tile entirely below the diagonal
    if (threadIdx.x == 0) {
        mbar_init(bar, 1);
        mbar_fence_init();
    }
    __syncthreads();
    pdl_wait();                                                  // W1 (K5) and P (the previous update) are complete
    pdl_launch_dependents();
    if (threadIdx.x == 0) {                                      // j0 + BJ <= n_pad: n_pad is a multiple of 256
        mbar_expect_tx(bar, BJ * LDK * 8);
        bulk_g2s(sA, W1r + (size_t)j0 * LDK, BJ * LDK * 8, bar);
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
    const int iw = i0 + warp * ROWS_W;                           // first row of this warp
    if (iw >= jend || iw >= r1) return;                          // whole warp below the tile / past the row block
    const int jrel = max(iw - j0, 0) & ~7;                       // the warp starts at its own diagonal, 8-column blocks
    const int jb0 = j0 + jrel;
    const int nblk = (jend - jb0 + 7) >> 3;
    const int i_first = iw + 2 * t;                              // the lane's row pair (i_first + 8 ib, + 1)
    const double *pcol = Pb + (size_t)(i_first - r0) + (size_t)(jb0 + g) * ldb;
    // one warp-wide L2 prefetch covers an 8-column block: ROWS_W * 8 bytes = LPC 128-byte lines per column
    constexpr int LPC = (ROWS_W * 8 + 127) / 128;
    const double *pf = Pb + (size_t)(iw - r0) + (size_t)(jb0 + lane / LPC) * ldb + (lane % LPC) * 16;
    const bool pf_lane = lane < 8 * LPC;
    for (int q = 1; q <= pd; q++)
        if (pf_lane && q < nblk && jb0 + 8 * q + lane / LPC < jend)
            asm volatile("prefetch.global.L2::evict_last [%0];" ::"l"(pf + (size_t)(8 * q) * ldb));

    // row-side fragments: B[k = t][i = g] = -W1[iw + 8 ib + g][4 ks + t]
    double b[IB][KS];
#pragma unroll
    for (int ib = 0; ib < IB; ib++) {
        const int row = min(iw + 8 * ib + g, n_pad - 1);         // rows past the block end are never stored
        const double *wr = W1r + (size_t)row * LDK + t;
#pragma unroll
        for (int ks = 0; ks < KS; ks++) b[ib][ks] = -__ldg(wr + 4 * ks);
    }
    const bool rows_full = iw + ROWS_W <= r1;
    double2 acc[IB], nxt[IB];
    ld_block<IB>(nxt, pcol, rows_full && jb0 + 8 <= jend, jb0 + g < jend, i_first, r1);
    mbar_wait(bar, 0);
    const double *aj = sA + (size_t)(jrel + g) * LDK + t;        // A[j = g][k = t] = W1[jb + g][4 ks + t]
#pragma unroll 1
    for (int blk = 0; blk < nblk; blk++) {
        const int jb = jb0 + 8 * blk;
        const bool fast = rows_full && jb + 8 <= jend, colok = jb + g < jend;
#pragma unroll
        for (int ib = 0; ib < IB; ib++) acc[ib] = nxt[ib];
        if (blk + 1 < nblk)
            ld_block<IB>(nxt, pcol + (size_t)8 * ldb, rows_full && jb + 16 <= jend, jb + 8 + g < jend, i_first, r1);
        if (pf_lane && blk + 1 + pd < nblk && jb + 8 * (1 + pd) + lane / LPC < jend)
            asm volatile("prefetch.global.L2::evict_last [%0];" ::"l"(pf + (size_t)(8 * (blk + 1 + pd)) * ldb));
#pragma unroll
        for (int ks = 0; ks < KS; ks++) {
            const double a = aj[4 * ks];
#pragma unroll
            for (int ib = 0; ib < IB; ib++) dmma(acc[ib], a, b[ib][ks]);
        }
        st_block<IB>(acc, const_cast<double *>(pcol), fast, colok, i_first, r1);
        pcol += (size_t)8 * ldb;
        aj += 8 * LDK;
    }
}

struct SCfg { int ib, warps, bj, pd, minb; };

// tuned on the B200 (profiles/r2_k6_stream.md); SLAM_K6S_CFG="ib,warps,bj,pd,minb" overrides
SCfg pick_scfg(int RC) {
    SCfg c = RC <= 24   ? SCfg{4, 8, 64, 2, 2}
             : RC <= 32 ? SCfg{4, 8, 128, 2, 2}
             : RC <= 40 ? SCfg{2, 4, 128, 2, 4}
             : RC <= 48 ? SCfg{2, 4, 128, 3, 4}
                        : SCfg{2, 4, 64, 2, 4};
    if (const char *e = getenv("SLAM_K6S_CFG")) {
        int a, b, d, f, g = 0;
        const int got = sscanf(e, "%d,%d,%d,%d,%d", &a, &b, &d, &f, &g);
        if (got >= 4) c = SCfg{a, b, d, f, got >= 5 ? g : 0};
    }
    return c;
}

template <int RC, int IB, int WARPS, int BJ, int MINB>
int launch_one(double *Pb, size_t ldb, const RowMap &map, int n, int n_pad, const double *W1r, int pd, cudaStream_t s) {
    constexpr int BI = WARPS * 8 * IB;
    const size_t smem = sizeof(double) * BJ * (RC + W1_PAD) + 16;
    auto kern = k_downdate_stream<RC, IB, WARPS, BJ, MINB>;
    if (smem > 48 * 1024) CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid((rowmap_local_rows(map) + BI - 1) / BI, (n + BJ - 1) / BJ);
    CUDA_TRY(launch_pdl(kern, grid, dim3(WARPS * 32), smem, s, Pb, ldb, map, n, n_pad, W1r, pd));
    return SLAM_OK;
}

template <int RC>
int launch_rc(const SCfg &c, double *Pb, size_t ldb, const RowMap &map, int n, int n_pad, const double *W1r, cudaStream_t s) {
    // minb = 0 in the configuration takes the first entry with matching (ib, warps, bj)
#define TRY_S(IB_, W_, BJ_, MINB_)                                                                       \
    if (c.ib == IB_ && c.warps == W_ && c.bj == BJ_ && (c.minb == 0 || c.minb == MINB_))                 \
        return launch_one<RC, IB_, W_, BJ_, MINB_>(Pb, ldb, map, n, n_pad, W1r, c.pd, s);
    if constexpr (RC <= 32) {
        TRY_S(4, 8, 64, 2) TRY_S(4, 8, 128, 2) TRY_S(4, 4, 128, 4) TRY_S(4, 4, 64, 4)
    }
    if constexpr (RC > 16) {
        TRY_S(2, 8, 128, 2) TRY_S(2, 4, 128, 4) TRY_S(2, 8, 64, 2) TRY_S(2, 4, 64, 4)
    }
#undef TRY_S
    slam_set_error("launch_downdate_stream: no kernel for ib=%d warps=%d bj=%d minb=%d at a %d-wide chunk", c.ib, c.warps,
                   c.bj, c.minb, RC);
    return SLAM_ERR_ARG;
}

}  // namespace

// one chunk of 4..64 ranks, 16-byte accesses (even leading dimension; local row origins are even by construction)
bool downdate_stream_usable(int R, size_t ldb) {
    const int RC = downdate_chunk_width(R);
    return R <= 64 && RC >= 4 && RC % 4 == 0 && RC >= R && (ldb % 2) == 0;
}

int launch_downdate_stream_rows(double *Pb, size_t ldb, RowMap map, int n, const double *W1c, int n_pad, int R,
                                cudaStream_t s, long long *launches) {
    if (rowmap_local_rows(map) <= 0) return SLAM_OK;
    const int RC = downdate_chunk_width(R);
    const SCfg c = pick_scfg(RC);
    int rc;
    switch (RC) {
        case 4: rc = launch_rc<4>(c, Pb, ldb, map, n, n_pad, W1c, s); break;
        case 8: rc = launch_rc<8>(c, Pb, ldb, map, n, n_pad, W1c, s); break;
        case 12: rc = launch_rc<12>(c, Pb, ldb, map, n, n_pad, W1c, s); break;
        case 16: rc = launch_rc<16>(c, Pb, ldb, map, n, n_pad, W1c, s); break;
        case 24: rc = launch_rc<24>(c, Pb, ldb, map, n, n_pad, W1c, s); break;
        case 32: rc = launch_rc<32>(c, Pb, ldb, map, n, n_pad, W1c, s); break;
        case 40: rc = launch_rc<40>(c, Pb, ldb, map, n, n_pad, W1c, s); break;
        case 48: rc = launch_rc<48>(c, Pb, ldb, map, n, n_pad, W1c, s); break;
        case 56: rc = launch_rc<56>(c, Pb, ldb, map, n, n_pad, W1c, s); break;
        case 64: rc = launch_rc<64>(c, Pb, ldb, map, n, n_pad, W1c, s); break;
        default:
            slam_set_error("launch_downdate_stream: %d ranks are not a single chunk of 4..64", R);
            return SLAM_ERR_ARG;
    }
    if (rc) return rc;
    if (launches) (*launches)++;
    CUDA_TRY(cudaGetLastError());
    return SLAM_OK;
}
