/* libslamcuda.so -- C ABI of the B200-native EKF-SLAM hot path.
 *
 * Drop-in boundary for the filter half of andrewadare/SLAM.jl's `module SLAM`
 * (reference paths are under /root/reference):
 *
 *   slam_predict             replaces  predict(state, vehicle, Q, dt)            src/ekf.jl:8-43
 *   slam_associate           replaces  associate(state, z, R, gate1, gate2)      src/data-association.jl:1-51
 *   slam_compute_association replaces  compute_association(x, P, z, R, idf)      src/data-association.jl:53-63
 *   slam_predict_observation replaces  predict_observation(x, idf)               src/common.jl:139-165
 *   slam_update              replaces  update(state, z, R, idf)                  src/ekf.jl:46-77
 *   slam_add_features        replaces  add_features(state, z, R)                 src/ekf.jl:84-122
 *   slam_set_state/get_state carry     EKFSlamState{T}.x / .cov                  src/common.jl:25-28
 *
 * The reference has no FFI of its own: its caller is sim! (sim/ekfslam-sim.jl:100,114,
 * 117,120), which re-binds `state.x, state.cov = f(state, ...)`.  A Julia `SLAM` module
 * keeping those names ccall()s these entry points.
 *
 * Conventions
 *  - extern "C", plain pointers and sizes; every pointer is HOST memory unless the
 *    parameter name starts with d_.
 *  - Matrices are column-major (Julia layout).  z/zf/zn are 2 x count, column i at
 *    z[2*i], z[2*i+1] = (range, bearing).  R and Q are 2 x 2 column-major.
 *  - Landmark ids `idf` are 1-based int64 (Julia Int), exactly as the reference passes
 *    them; state index of landmark id is 3 + 2*(id-1) (0-based).
 *  - Every function returns 0 on success or a negative SLAM_ERR_* code;
 *    slam_last_error() gives the message of the last failure on the calling thread.
 *  - There is no CPU fallback: every call fails with SLAM_ERR_CUDA when no sm_100
 *    device is usable.
 *  - A handle is not re-entrant; distinct handles may be driven from distinct threads.
 *
 * Covariance storage: P lives on the device as a column-major n x n array with a fixed
 * leading dimension chosen at slam_create (capacity n_max), so augmentation is an
 * in-place append.  The upper triangle (i <= j) is authoritative; slam_get_state
 * mirrors it into the lower triangle on export.  slam_set_state reads the upper
 * triangle of the matrix it is given (the reference keeps P symmetric up to rounding in
 * its 3x3 / 2x2 diagonal blocks, src/ekf.jl:32,112).
 */
#ifndef SLAMCUDA_H
#define SLAMCUDA_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SLAM_OK 0
#define SLAM_ERR_NOT_PD    (-1) /* innovation covariance S not positive definite (Julia: chol throws) */
#define SLAM_ERR_BAD_INDEX (-2) /* idf outside 1..Nf (Julia: BoundsError)                            */
#define SLAM_ERR_CAPACITY  (-3) /* n + 2k exceeds n_max, or a count exceeds a fixed limit            */
#define SLAM_ERR_CUDA      (-4) /* CUDA runtime failure, or no usable device                         */
#define SLAM_ERR_NCCL      (-5) /* NCCL failure / libnccl not loadable                               */
#define SLAM_ERR_ARG       (-6) /* NULL handle, negative size, ...                                   */

#define SLAM_MAX_OBS 1024       /* upper limit on nz / m / k per call                                */

typedef struct slam_filter slam_t;

const char *slam_last_error(void);
/* "libslamcuda <version> sm_100a"; never fails, needs no device. */
const char *slam_version(void);
/* Number of CUDA devices visible (0 when none / no driver); never fails. */
int slam_device_count(void);

/* Measured FP64 roofs of `device`, TFLOP/s: DFMA chains and mma.sync.m8n8k4.f64 tiles (SASS DMMA).
 * bench.py quotes K6 above 32 ranks and the ensemble kernel against them.  ~0.1 s. */
int slam_fp64_peak(int device, double *dfma_tflops, double *dmma_tflops);

/* ---- single filter, state resident on one device ------------------------------------ */

/* EKFSlamState(x, cov) -- src/common.jl:25-28, as built by sim/ekfslam-sim.jl:42.
 * Allocates device state with capacity n_max (>= 3).  State starts as n = 3, x = 0, P = 0. */
int slam_create(int device, int n_max, slam_t **out);
int slam_destroy(slam_t *h);

/* Capacity growth: the reference's state grows without bound (hcat / vcat per new landmark,
 * src/ekf.jl:108-109); a handle's capacity is fixed so that slam_add_features appends in place.
 * slam_reserve raises it to at least n_max_new states (new device arrays, one device-to-device copy
 * of x and the live part of P; no-op when already that large; the old state survives a failure).
 * The host mirrors call it when slam_add_features reports SLAM_ERR_CAPACITY. */
int slam_reserve(slam_t *h, int n_max_new);
/* The current capacity n_max; negative on error. */
int slam_capacity(slam_t *h);

/* state.x / state.cov <- host arrays.  P is n x n column-major with leading dim ldp. */
int slam_set_state(slam_t *h, const double *x, const double *P, int n, int ldp);
/* host arrays <- state.x / state.cov (full symmetric matrix).  x or P may be NULL. */
int slam_get_state(slam_t *h, double *x, double *P, int ldp);
/* state.x[1:3], read by sim! every step (sim/ekfslam-sim.jl:128). */
int slam_get_pose(slam_t *h, double pose[3]);
/* length(state.x); negative on error. */
int slam_n(slam_t *h);
/* A sub-block cov[i0:i0+ni, j0:j0+nj] (0-based, full symmetric view) for front ends that
 * must not pull all of P (resident mode). */
int slam_get_cov_block(slam_t *h, int i0, int ni, int j0, int nj, double *out, int ldo);
/* Diagnostic, no counterpart in the reference (its arrays have no capacity padding): arm != 0 paints every element of
 * the device slab (ld x n_max) and of x outside the live state [0, n) with a NaN payload; arm == 0 writes to *changed
 * how many painted elements have been overwritten since.  predict / associate / update must leave it at 0;
 * add_features moves n and owns the rows and columns it adds.  Used by the tests in place of a memory checker. */
int slam_debug_guard(slam_t *h, int arm, long long *changed);

/* slam_get_cov_block without any allocation.  This entry point has no counterpart in the reference and was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  Same arguments, checks and result as slam_get_cov_block: `out` receives
 * cov[i0:i0+ni, j0:j0+nj] (0-based, full symmetric view, column-major, leading dimension ldo >= ni).
 *  - The gather kernel's scratch is a static device array of SLAM_COV_BLOCK_STATIC_MAX doubles (512 KiB) that every
 *    GPU loading the library holds until the process exits; the call issues no cudaMalloc / cudaFree and calls no
 *    allocating library routine.
 *  - ni * nj > SLAM_COV_BLOCK_STATIC_MAX is rejected with SLAM_ERR_CAPACITY and a message naming the limit (`out` is
 *    not written): read a larger window in pieces, or use slam_get_cov_block.
 *  - The array is shared by EVERY call on the device, whatever the handle, and keeps the last call's contents.  The
 *    kernel writes the whole window before it is read back, so nothing needs clearing; but calls that overlap on
 *    the GPU (two handles, two host threads) race on it -- serialise them. */
#define SLAM_COV_BLOCK_STATIC_MAX 65536
int slam_get_cov_block_static(slam_t *h, int i0, int ni, int j0, int nj, double *out, int ldo);

/* compute_landmark_ellipses -- src/common.jl:236-262 (with ellipse, :236-244), computed on the
 * device from the 2x2 diagonal blocks only: out is (2*Nf) x (nsegs+1) column-major with leading
 * dimension ldo >= 2*Nf; rows 2i-1:2i hold nsigma*sqrtm(P_ii)*[cos(phi)'; sin(phi)'] .+ x_i for
 * phi = 0:2pi/nsegs:2pi.  The reference uses nsigma = 2, nsegs = 12 (:248-249). */
int slam_compute_landmark_ellipses(slam_t *h, double nsigma, int nsegs, double *out, int ldo);
/* The vehicle's position ellipse, ellipse(x[1:2], P[1:2,1:2], nsigma, nsegs) as drawn by
 * sim/gr-draw.jl:112-117: out is 2 x (nsegs+1) column-major. */
int slam_vehicle_ellipse(slam_t *h, double nsigma, int nsegs, double *out);

/* Front-end payloads of the browser monitor (sim/browser/wsserver.jl:31-85) as headless buffers computed on the
 * device, so that a monitor of a large resident map never pulls P:
 *  slam_feature_ellipses       "feature-ellipses": out is 5 x Nf column-major, [cx; cy; sqrt(l1); sqrt(l2); phi] with
 *                              l, u = eig(cov[j, j]) ascending and phi = atan2(u[2,1], u[1,1]) normalised to
 *                              (-pi/2, pi/2] (wsserver.jl:72-85; an eigenvector's sign is LAPACK's choice there:
 *                              same ellipse).  Returns Nf.
 *  slam_vehicle_ellipse_params "vehicle-ellipse": [cx, cy, vehicle_phi, rx, ry, phi] (wsserver.jl:58-63).
 *  slam_laser_lines            "lidar": laser_lines(z, state.x[1:3]) of src/common.jl:265-283: lines is 4 x nz,
 *                              [vx; vy; fx; fy] per observation (range, bearing) of z (2 x nz, host).
 * ("tracks" is slam_get_pose per step, "state" is slam_get_state / slam_get_cov_block.) */
int slam_feature_ellipses(slam_t *h, double *out);
int slam_vehicle_ellipse_params(slam_t *h, double out[6]);
int slam_laser_lines(slam_t *h, const double *z, int nz, double *lines);

/* predict -- src/ekf.jl:8-43.  g, v = vehicle.measured_gamma / measured_speed (:14-15),
 * wheelbase = vehicle.wheelbase (:16), Q 2x2 control covariance, dt timestep. In place. */
int slam_predict(slam_t *h, double g, double v, double wheelbase, const double Q[4], double dt);

/* associate -- src/data-association.jl:1-51.  z is 2 x nz.  Outputs (caller buffers of
 * capacity nz columns): zf 2 x *m, idf[*m] (1-based), zn 2 x *k; observation order kept. */
int slam_associate(slam_t *h, const double *z, int nz, const double R[4], double gate1,
                   double gate2, double *zf, int64_t *idf, int *m, double *zn, int *k);

/* Spatial pre-gate of slam_associate (the reference's TODO at
 * src/data-association.jl:18-20).  On (the default), a landmark is scored against an
 * observation only if it lies inside the annulus around the vehicle outside which
 * nis > max(gate1, gate2) is certain (Cauchy-Schwarz bound on the range innovation with a
 * device-resident upper bound of the landmark variances, x 4 head-room); results are identical
 * to the exhaustive scan for any positive semidefinite P.  on = 0 restores the exhaustive scan. */
int slam_set_pregate(slam_t *h, int on);
/* Cumulative counters of slam_associate since slam_create: landmarks scanned (annulus test on
 * x only) and landmarks scored (5 x 5 gather + Mahalanobis distances), and the current bound
 * on the landmark variances.  Any pointer may be NULL.  Synchronises the handle's stream. */
int slam_gate_stats(slam_t *h, long long *scanned, long long *scored, double *sig2max);

/* compute_association -- src/data-association.jl:53-63 for one (z, landmark) pair. */
int slam_compute_association(slam_t *h, const double z[2], const double R[4], int64_t idf,
                             double *nis, double *nd);

/* predict_observation -- src/common.jl:139-165.  zp = [range; bearing] (bearing not wrapped).
 * H is the dense 2 x n Jacobian, column-major with leading dimension 2 (may be NULL). */
int slam_predict_observation(slam_t *h, int64_t idf, double zp[2], double *H);

/* update -- src/ekf.jl:46-77: batch Cholesky-form update with all m associated
 * observations.  m == 0 is a no-op (first observation step of every run).  Duplicate ids
 * in idf are legal.  In place on the device state.
 * Completion: the call returns once the step's result is final -- the status (not-PD / bad id, decided by the
 * innovation Cholesky) and x, which slam_get_pose then serves without another wait -- while the downdate of P may
 * still be running on the handle's stream.  Every later call on the handle, and every read-out (slam_get_state,
 * slam_get_cov_block, ...), is ordered behind it, and slam_sync waits for everything; SLAM_UPDATE_EARLY=0 in the
 * environment makes the call itself wait for the downdate as well. */
int slam_update(slam_t *h, const double *zf, int m, const double R[4], const int64_t *idf);

/* add_features -- src/ekf.jl:84-122: append k new landmarks in place (n += 2k). */
int slam_add_features(slam_t *h, const double *zn, int k, const double R[4]);

/* Resident-mode variants: inputs already on the device, nothing is copied back and the
 * call does not synchronise.  Errors (not-PD, bad index) are latched on the device and
 * returned by the next slam_sync().  Used by bench.py's device-resident timing. */
int slam_update_resident(slam_t *h, const double *d_zf, int m, const double R[4],
                         const int64_t *d_idf);
int slam_sync(slam_t *h);
/* The CUDA stream (cudaStream_t) all work of this handle is issued on. */
void *slam_stream(slam_t *h);
/* Downdate kernel choice for update: 0 = auto (FMA kernel up to 8 ranks, streaming DMMA kernel for 12..64
 * ranks, tiled DMMA kernel above), 1 = force the FP64 FMA kernel (up to 32 ranks per pass), 2 = force the tiled
 * FP64 DMMA tensor-core kernel (where its 32-wide chunk layout applies), 3 = force the streaming DMMA kernel
 * (12..64 ranks).  A forced variant that does not apply to the rank falls through to the automatic choice. */
int slam_set_downdate_variant(slam_t *h, int variant);
/* Which kernel the last update's downdate ran: 0 none yet, 1 k_downdate_fma, 2 k_downdate_dmma (tiled), 3 k_downdate_stream. */
int slam_downdate_kernel(slam_t *h);
/* Number of kernels this handle has launched since creation (bench.py's gpu_launches). */
long long slam_launch_count(slam_t *h);
/* Kernel timing for roofline reports: while on, every update records a CUDA event pair
 * around its downdate (K6) launches on the handle's stream.  slam_profile_read synchronises
 * and returns the summed K6 milliseconds and the number of timed updates since it was
 * switched on (at most 4096 are kept), then clears them. */
int slam_profile(slam_t *h, int on);
int slam_profile_read(slam_t *h, double *k6_ms, int *k6_updates);

/* Synthetic large-map state generated ON the device:
 * landmarks uniform on a side x side square, x_landmark = truth + N(0, 0.1^2),
 * P = D + L L' with D = diag U(0.01, 0.1), L n x 8 N(0, 0.05^2), vehicle at the centre
 * with heading 0.  All draws are Philox4x32-10 keyed by (seed, stream, index) so a CPU
 * mirror reproduces any entry.  Sets n = 3 + 2*n_landmarks (<= n_max). */
int slam_synth_state(slam_t *h, int n_landmarks, double side, uint64_t seed);
/* Observations of the m nearest landmarks ahead of the vehicle (truth position, no
 * noise added) followed by `spurious` observations far beyond gate2; writes 2 x (m+spurious)
 * to z (host) and the true 1-based ids of the first m to ids (may be NULL). */
int slam_synth_observations(slam_t *h, int n_landmarks, double side, uint64_t seed, int m,
                            int spurious, double *z, int64_t *ids);

/* ---- ensemble of independent course runs (one CTA per filter) ------------------------ */

typedef struct slam_ensemble slam_ens_t;

/* Precomputes the noise-free truth track and visibility tables of sim! on the host
 * (sim/ekfslam-sim.jl:80-142 with src/common.jl:172-230, sim/sim-utils.jl:12-28,53-75) and
 * uploads them.  waypoints 2 x nwp, landmarks 2 x nl, both column-major.  Members are
 * run ids [run0, run0 + n_runs); member r draws Philox normals keyed by (seed, r). */
int slam_ens_create(int device, const double *waypoints, int nwp, const double *landmarks,
                    int nl, int nlaps, uint64_t seed, uint32_t run0, int n_runs,
                    slam_ens_t **out);
int slam_ens_destroy(slam_ens_t *e);
/* Member capacity.  The reference's state grows without bound (hcat / vcat, src/ekf.jl:108-109);
 * a member holds 3 + 2*(nl + extra) states: `extra` (default 5) leaves room for the spurious
 * landmarks the algorithm itself creates when an observation of a mapped landmark fails both gates
 * (src/data-association.jl:45-47).  A member that outgrows it is re-run with more room (slam_ens_set_growth)
 * or, with growth = 0, stops with SLAM_ERR_CAPACITY in status[].  Call before slam_ens_run. */
int slam_ens_set_extra_landmarks(slam_ens_t *e, int extra);
/* 3 + 2*(nl + extra); negative on error. */
int slam_ens_capacity(slam_ens_t *e);
/* How a member whose map outgrows that capacity is treated.  The reference's state grows without bound
 * (hcat / vcat, src/ekf.jl:108-109), and so does a member by default:
 *   2 (default)  the members a launch cannot hold are re-run -- from their first step; the noise is keyed by the
 *                run id, so the result is the single-launch result -- by further launches with more
 *                shared memory per member (two CTAs per SM, then one; when a first pass with less head-room runs
 *                in front, the launch at slam_ens_capacity() itself is skipped), and finally with x, P and the
 *                workspace in global memory, the capacity doubling per round up to 3 + 2*(landmarks +
 *                observations of the whole run), which no run can exceed: no member ever stops for capacity;
 *   1            the shared-memory launches only;
 *   0            fixed capacity: a member that outgrows slam_ens_capacity() stops with SLAM_ERR_CAPACITY. */
int slam_ens_set_growth(slam_ens_t *e, int mode);
/* Caps the head-room (landmarks beyond nl) of the shared-memory launches of growth >= 1; -1 (default) = what
 * one CTA's shared memory holds.  (Tests use it to send members to the global-memory rounds.) */
int slam_ens_set_max_shared_extra(slam_ens_t *e, int extra);
/* Algorithmic FP64 operations of a run, tallied by the kernel with SURVEY.md 8d's figures and every member's own
 * n, nz, m at every step (30 n per predict; 80 nz Nf per observation step, plus 2m n(n+1) + 20 n m + 8 n m^2 +
 * 8 m^3 with m matches).  on = 1 before slam_ens_run; *total (may be NULL) = the last run's sum; on = 0 frees. */
int slam_ens_count_flops(slam_ens_t *e, int on, double *total);
/* Every launch of the last run in order (shared-memory tiers, then global-memory rounds): capacity in states
 * and members run, at most `max` entries each; returns the number of launches.  Waits for the run. */
int slam_ens_tier_info(slam_ens_t *e, int *capacity, int *members, int max);
/* Two-pass capacity.  The kernel's throughput follows the member CTAs resident per SM, which the
 * member's shared memory (i.e. its capacity) decides, and few members ever use the head-room.  A run
 * may therefore take two launches: a first pass with `first_extra` < extra landmarks of head-room,
 * and a second one that re-runs, from their first step and with the full capacity, only the members
 * whose map outgrew the first.  A member's noise is keyed by its run id, so every result is the same
 * as in a single pass.  first_extra = -1 (default): chosen from the occupancy; >= extra: single pass. */
int slam_ens_set_first_pass_extra(slam_ens_t *e, int first_extra);
/* After slam_ens_run: *first_extra the first pass used (-1 = single pass), *redone = members re-run. */
int slam_ens_pass_info(slam_ens_t *e, int *first_extra, int *redone);
/* Threads per member CTA: 64, 128 (default) or 256. */
int slam_ens_set_threads(slam_ens_t *e, int threads);
/* Steps of one run and number of observation steps (2802 / 311 for 2 laps of course1). */
int slam_ens_steps(slam_ens_t *e, int *nsteps, int *nupdates);
/* Runs every member from its initial state to the end of the course in one launch.
 * Asynchronous; slam_ens_sync() waits and returns the number of failed members (< 0 on
 * CUDA error). */
int slam_ens_run(slam_ens_t *e);
int slam_ens_sync(slam_ens_t *e);
/* Filter-steps actually completed by the last run: *total over all members (a member that stopped
 * early counts only the steps it ran) and per_member[n_runs]; either may be NULL. */
int slam_ens_steps_done(slam_ens_t *e, long long *total, int *per_member);
/* Results: final_pose 3 x n_runs; n_final[n_runs] = final state length per member;
 * status[n_runs] (0 ok, SLAM_ERR_*).  Any pointer may be NULL. */
int slam_ens_get_results(slam_ens_t *e, double *final_pose, int *n_final, int *status);
/* slam_track (3 x nsteps) of one member (sim/ekfslam-sim.jl:128); requires tracks to have
 * been enabled before the run with slam_ens_keep_tracks(e, 1). */
int slam_ens_keep_tracks(slam_ens_t *e, int on);
int slam_ens_get_track(slam_ens_t *e, int member, double *track);
/* Final x (n_final) and P (n_final^2, ldp, full symmetric) of one member; returns n_final.  x and P may both
 * be NULL to ask for n_final first (with growth > 0 it is not bounded by slam_ens_capacity()). */
int slam_ens_get_state(slam_ens_t *e, int member, double *x, double *P, int ldp);
/* Host-supplied noise instead of Philox (parity runs): ctrl 2 x nsteps per member,
 * obs: pool of `obs_per_member` normals per member consumed in draw order. */
int slam_ens_set_noise(slam_ens_t *e, const double *ctrl, const double *obs, int obs_per_member);
void *slam_ens_stream(slam_ens_t *e);
long long slam_ens_launch_count(slam_ens_t *e);

/* ---- large map, P row-block-sharded over the ranks of one NCCL communicator ---------- */

typedef struct slam_shard slam_shard_t;

/* 128-byte NCCL unique id created on rank 0 and handed to every rank by the host
 * launcher (torch.distributed / MPI / a file). */
int slam_shard_unique_id(char id[128]);
/* P is sharded by row panels of 256 rows (panel 0 = rows [0, 255), panel p >= 1 = rows [256 p - 1, 256 p + 255): odd
 * boundaries, a landmark's two rows never straddle), dealt boustrophedon: panel p goes to rank b = p mod 2*world when
 * b < world, else to 2*world-1-b, so the upper-triangle work is balanced at any state size, not only at capacity
 * (the map grows one landmark at a time, src/ekf.jl:84-122).  world == 1 needs no NCCL. */
int slam_shard_create(int device, int rank, int world, const char id[128], int n_max,
                      slam_shard_t **out);
int slam_shard_destroy(slam_shard_t *s);
/* Pure partition arithmetic (no device needed): rows [lo[k], hi[k]) of the k-th panel `rank` owns at state size n,
 * in the order the rank stacks them (at most `max` entries are written; lo / hi may be NULL); returns the number of
 * panels the rank owns. */
int slam_shard_plan(int n, int world, int rank, int *lo, int *hi, int max);
int slam_shard_synth_state(slam_shard_t *s, int n_landmarks, double side, uint64_t seed);
int slam_shard_n(slam_shard_t *s);
int slam_shard_get_pose(slam_shard_t *s, double pose[3]);
int slam_shard_get_x(slam_shard_t *s, double *x);
/* Rows [i0, i0+ni) x cols [j0, j0+nj) of the authoritative upper triangle owned by this
 * rank: rows this rank does not own are left untouched in `out`; of the rows it owns the whole
 * window is copied as stored, so entries below the diagonal are overwritten with unspecified
 * values (the lower triangle is never maintained).  Returns the count copied. */
int slam_shard_get_cov_block(slam_shard_t *s, int i0, int ni, int j0, int nj, double *out, int ldo);
int slam_shard_predict(slam_shard_t *s, double g, double v, double wheelbase, const double Q[4],
                       double dt);
int slam_shard_associate(slam_shard_t *s, const double *z, int nz, const double R[4], double gate1,
                         double gate2, double *zf, int64_t *idf, int *m, double *zn, int *k);
int slam_shard_update(slam_shard_t *s, const double *zf, int m, const double R[4],
                      const int64_t *idf);
/* 1 when the update exchanges the H rows and W1 through peer memory (one kernel computes and stores
 * into every rank's buffers over NVLink, CUDA IPC mappings, flag barrier), 0 when it uses NCCL
 * collectives (SLAM_SHARD_P2P=0, world = 1, or IPC refused).  Both give bit-identical results. */
int slam_shard_p2p(slam_shard_t *s);
/* As slam_downdate_kernel, for the sharded filter's last update. */
int slam_shard_downdate_kernel(slam_shard_t *s);
/* As slam_debug_guard, for this rank's row-panel stack (ldl x n_max) and its copy of x: a slot is live when its local
 * row holds a global row < n of the current state and its column is < n.  Rank-local, no collective. */
int slam_shard_debug_guard(slam_shard_t *s, int arm, long long *changed);
/* add_features -- src/ekf.jl:84-122 on the sharded covariance: rows 0..2 of P are broadcast from
 * rank 0, every rank fills its own rows of the 2k new columns, the owner of the new rows' panel the diagonal
 * blocks.  Collective: call on every rank. */
int slam_shard_add_features(slam_shard_t *s, const double *zn, int k, const double R[4]);
int slam_shard_update_resident(slam_shard_t *s, const double *d_zf, int m, const double R[4],
                               const int64_t *d_idf);
int slam_shard_sync(slam_shard_t *s);
void *slam_shard_stream(slam_shard_t *s);
long long slam_shard_launch_count(slam_shard_t *s);
/* As slam_profile / slam_profile_read: CUDA events around this rank's K6 launches. */
int slam_shard_profile(slam_shard_t *s, int on);
int slam_shard_profile_read(slam_shard_t *s, double *k6_ms, int *k6_updates);
/* Same window, split: [0] H-row exchange, [1] PHt / S / Cholesky, [2] gain + W1 exchange, [3] K6 (ms). */
int slam_shard_profile_phases(slam_shard_t *s, double phases_ms[4], int *updates);

#ifdef __cplusplus
}
#endif
#endif /* SLAMCUDA_H */
