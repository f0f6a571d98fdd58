// Internal launch interfaces shared by the translation units of libslamcuda.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

constexpr int DD_ROW_PAD = 256;  // W1 row padding: covers every K6 tile shape

// ---- K6 (downdate.cu / downdate_dmma.cu) -------------------------------------------------
int downdate_chunk_width(int R);
int launch_downdate(double *P, size_t ld, int n, const double *W1c, int n_pad, int R, int variant,
                    cudaStream_t s, long long *launches);
// Rectangular form used by the row-block-sharded path: rows [r0, r1) of the upper triangle
// stored as a local block Pb (element (i, j) at Pb[(i - r0) + j*ldb]), W1 rows global.
int launch_downdate_rows(double *Pb, size_t ldb, int r0, int r1, int n, const double *W1c, int n_pad,
                         int R, int variant, cudaStream_t s, long long *launches);

// FP64 tensor-core variant (downdate_dmma.cu): same contract; usable for the 32-wide chunk layout
bool downdate_dmma_usable(int R, size_t ldb);
int launch_downdate_dmma_rows(double *Pb, size_t ldb, int r0, int r1, int n, const double *W1c, int n_pad, int R,
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
int synth_observations_host(int n_landmarks, double side, uint64_t seed, int m, int spurious, double *z,
                            int64_t *ids);
int slam_require_device(int device);

// ---- K4 (filter.cu): S = H*PHt + RR, C = inv(chol(S)), u = C'v; PHt is rank-major with stride npad
int launch_innov_chol(const double *PHt, int npad, const double *Hn, const int *fp, const double *v, int m,
                      const double R[4], double *S, double *C, double *u, int *err, cudaStream_t s);

// ---- K2 on a row-block-sharded filter (gate.cu) ------------------------------------------
constexpr int GATE_RES_BYTES = 24;
int launch_gate_shard(const double *d_x, const double *d_P3, int npad, const double *const d_Pb[2], size_t ldb,
                      const int lo[2], const int hi[2], const double *d_z, int nz, const double R[4], double gate1,
                      double gate2, void *d_scratch, size_t scratch_bytes, void *d_rank_res,
                      const unsigned long long *d_sig2bits, cudaStream_t s, long long *launches);
int launch_gate_finalize(const void *d_all_res, int world, const double *d_z, int nz, double gate2, double *d_zf,
                         int64_t *d_idf, double *d_zn, int *d_counts, cudaStream_t s, long long *launches);
