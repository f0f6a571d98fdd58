// Internal launch interfaces shared by the translation units of libslamcuda.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

constexpr int DD_ROW_PAD = 256;  // W1 row padding: covers every K6 tile shape

// W1 workspace layout (written by K5, read by K6).  Per rank chunk c of width RC (R zero-padded to a multiple):
//   row-major copy   [n_pad][RC + W1_PAD]   at W1c + c * w1_chunk_doubles(n_pad, RC)
//   rank-major copy  [RC][n_pad]            at ... + w1_rank_major_offset(n_pad, RC)
// The row-major rows carry W1_PAD unused (zero) doubles so that a tile of rows fetched by ONE 1-D bulk copy lands
// in shared memory with a row stride of RC + 4: the (row = lane / 4, rank = lane % 4) fragment reads of the
// FP64 MMA are then bank-conflict free (stride 4 mod 16 eight-byte words), with no per-row copies and no swizzle.
constexpr int W1_PAD = 4;
__host__ __device__ inline int w1_ldk(int RC) { return RC + W1_PAD; }
__host__ __device__ inline size_t w1_chunk_doubles(int n_pad, int RC) { return (size_t)n_pad * (2 * RC + W1_PAD); }
__host__ __device__ inline size_t w1_rank_major_offset(int n_pad, int RC) { return (size_t)n_pad * (RC + W1_PAD); }
// upper bound of the whole workspace for any R <= Rw (chunks of width >= 4 need 2 Rpad + 4 nch <= 3 Rpad doubles per row)
inline size_t w1_workspace_doubles(int n_pad, int Rw) { return (size_t)n_pad * (3 * (size_t)Rw + 8); }

// ---- row maps: which global rows of P a launch covers and where they live in the local array -------------
// Contiguous (single filter: r0 = 0, r1 = n): global row i at local row i - r0.
// Panel-cyclic (large-map mode, shard.cu): P is cut into panels of SH_PH = 256 rows -- panel 0 = rows [0, 255),
// panel p >= 1 = rows [256 p - 1, 256 p + 255): every boundary is odd, so a landmark's two rows (3 + 2k, 4 + 2k)
// never straddle two panels -- dealt to the ranks boustrophedon, b = p mod 2W -> rank b (b < W) or 2W-1-b:
// every group of 2W panels gives every rank one "early" and one "late" panel, so the upper-triangle work is
// balanced at ANY state size n, not only at capacity.  A rank stacks its panels in the order they come: local
// panel lp = 2 (p div 2W) + (b >= W), local row = 256 lp + (i - first row of p).
constexpr int SH_PH = 256;
__host__ __device__ inline int panel_first(int p) { return p > 0 ? SH_PH * p - 1 : 0; }
__host__ __device__ inline int panel_of_row(int i) { return (i + 1) / SH_PH; }
__host__ __device__ inline int panel_owner(int p, int W) {
    const int b = p % (2 * W);
    return b < W ? b : 2 * W - 1 - b;
}
__host__ __device__ inline int panel_local(int p, int W) { return 2 * (p / (2 * W)) + ((p % (2 * W)) < W ? 0 : 1); }
__host__ __device__ inline int panel_global(int lp, int rank, int W) {
    return (lp >> 1) * 2 * W + ((lp & 1) ? 2 * W - 1 - rank : rank);
}
struct RowMap {
    int r0, r1;        // contiguous mode (world == 0): rows [r0, r1)
    int rank, world;   // panel-cyclic mode (world >= 1)
    int nlp;           // local panels in use
};
inline RowMap rowmap_rows(int r0, int r1) { return RowMap{r0, r1, 0, 0, 0}; }
inline RowMap rowmap_panels(int rank, int world, int nlp) { return RowMap{0, 0, rank, world, nlp}; }
// local rows a launch has to cover (grid.x = ceil(rowmap_local_rows / rows per CTA))
inline int rowmap_local_rows(const RowMap &m) { return m.world ? m.nlp * SH_PH : m.r1 - m.r0; }
// CTA row tile bx of BI rows (BI divides SH_PH in cyclic mode): first global row *i0, its local row *loc0, and the
// end *r1 of the rows the tile may touch.  False when the tile is empty.
__device__ __forceinline__ bool rowmap_tile(const RowMap &m, int bx, int BI, int n, int *i0, int *loc0, int *r1) {
    if (m.world == 0) {
        *i0 = m.r0 + bx * BI;
        *loc0 = bx * BI;
        *r1 = m.r1;
    } else {
        const int lrow = bx * BI, lp = lrow / SH_PH;
        const int p = panel_global(lp, m.rank, m.world);
        *i0 = panel_first(p) + (lrow % SH_PH);
        *loc0 = lrow;
        *r1 = min(panel_first(p + 1), n);
    }
    return *i0 < *r1;
}

// ---- K6 (downdate.cu / downdate_dmma.cu) -------------------------------------------------
int downdate_chunk_width(int R);
// 1 = FMA kernel, 2 = tiled DMMA, 3 = streaming DMMA, 0 = none; only the FMA kernel reads the rank-major W1 copy
int downdate_kernel_choice(int R, int variant, size_t ldb);
int launch_downdate(double *P, size_t ld, int n, const double *W1c, int n_pad, int R, int variant,
                    cudaStream_t s, long long *launches);
// Rectangular form used by the row-block-sharded path: rows [r0, r1) of the upper triangle
// stored as a local block Pb (element (i, j) at Pb[(i - r0) + j*ldb]), W1 rows global.
int launch_downdate_rows(double *Pb, size_t ldb, int r0, int r1, int n, const double *W1c, int n_pad,
                         int R, int variant, cudaStream_t s, long long *launches);
// General form: the rows of `map` (element (i, j) at Pb[local row + j*ldb]).
int launch_downdate_map(double *Pb, size_t ldb, RowMap map, int n, const double *W1c, int n_pad, int R, int variant,
                        cudaStream_t s, long long *launches);

// streaming FP64 tensor-core variant (downdate_stream.cu): ranks 12..64 in ONE pass over P, the row-side W1
// fragments of a warp in registers, the column-side tile staged once per tile by a bulk copy
bool downdate_stream_usable(int R, size_t ldb);
int launch_downdate_stream_rows(double *Pb, size_t ldb, RowMap map, int n, const double *W1c, int n_pad, int R,
                                cudaStream_t s, long long *launches);

// FP64 tensor-core tile variant (downdate_dmma.cu): same contract; usable for the 32-wide chunk layout
bool downdate_dmma_usable(int R, size_t ldb);
int launch_downdate_dmma_rows(double *Pb, size_t ldb, RowMap map, int n, const double *W1c, int n_pad, int R,
                              cudaStream_t s, long long *launches);

// ---- K2 (gate.cu, compiled with -fmad=false) ---------------------------------------------
// z: device 2 x nz. Results: d_zf (2 x nz), d_idf (nz), d_zn (2 x nz), d_counts[2] = {m, k}; P is the full
// column-major matrix with the upper triangle authoritative.
int launch_gate(const double *d_x, const double *d_P, size_t ld, int Nf, const double *d_z, int nz,
                const double R[4], double gate1, double gate2, void *d_scratch, size_t scratch_bytes,
                double *d_zf, int64_t *d_idf, double *d_zn, int *d_counts,
                const unsigned long long *d_sig2bits, unsigned long long *d_stats, cudaStream_t s,
                long long *launches);
size_t gate_scratch_bytes(int nz);
// Spatial pre-gate: d_sig2bits holds the bit pattern of an upper bound of the landmarks'
// diagonal entries of P (nullptr = exhaustive scan); launch_sig2_refresh recomputes it exactly.
int launch_sig2_refresh(const double *d_P, size_t ld, int n, unsigned long long *d_sig2bits, cudaStream_t s,
                        long long *launches);
int launch_compute_association(const double *d_x, const double *d_P, size_t ld, int n, const double z[2],
                               const double R[4], int64_t idf, double *d_out2, int *d_err,
                               cudaStream_t s, long long *launches);

// ---- synthetic state (synth.cu) -----------------------------------------------------------
int synth_fill(double *d_x, double *d_Pb, size_t ldb, int r0, int r1, int n_landmarks, double side,
               uint64_t seed, cudaStream_t s, long long *launches);
int synth_fill_map(double *d_x, double *d_Pl, size_t ldl, RowMap map, int n_landmarks, double side, uint64_t seed,
                   cudaStream_t s, long long *launches);
int synth_observations_host(int n_landmarks, double side, uint64_t seed, int m, int spurious, double *z,
                            int64_t *ids);
int slam_require_device(int device);

// ---- K4 (filter.cu): S = H*PHt + RR, C = inv(chol(S)), u = C'v; PHt is rank-major with stride npad
int launch_innov_chol(const double *PHt, int npad, const double *Hn, const int *fp, const double *v, int m,
                      const double R[4], double *S, double *C, double *u, int *err, cudaStream_t s);

// ---- K2 on a row-block-sharded filter (gate.cu) ------------------------------------------
constexpr int GATE_RES_BYTES = 24;
int launch_gate_shard(const double *d_x, const double *d_P3, int npad, const double *d_Pl, size_t ldl, int rank, int world,
                      int nlp, int n, const double *d_z, int nz, const double R[4], double gate1,
                      double gate2, void *d_scratch, size_t scratch_bytes, void *d_rank_res,
                      const unsigned long long *d_sig2bits, cudaStream_t s, long long *launches);
int launch_gate_finalize(const void *d_all_res, int world, const double *d_z, int nz, double gate2, double *d_zf,
                         int64_t *d_idf, double *d_zn, int *d_counts, cudaStream_t s, long long *launches);
