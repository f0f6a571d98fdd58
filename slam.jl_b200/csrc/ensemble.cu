// Ensemble mode: thousands of independent EKF-SLAM runs of the reference simulator loop
// (sim/ekfslam-sim.jl:80-142), one CTA per filter, the whole run in ONE kernel launch.
//
// The truth side of sim! does not depend on the noise (step_vehicle! uses the target speed and
// steering angle, src/common.jl:175-176), so the truth track, the steering angle and the set of
// visible landmarks per observation step are identical for every member that shares a map.
// They are computed once on the host (steer!, step_vehicle!, nearby_landmark_indices,
// get_observations; src/common.jl:172-230, sim/sim-utils.jl:12-28,53-66) and uploaded as tables.
// Each member then runs predict / associate / update / add_features for all steps with its
// state (x, the packed upper triangle of P, update workspace) resident in shared memory; only the
// noise (Philox4x32-10 keyed by (seed, run), or host-supplied arrays) and the optional track cross
// HBM.  The reference's state grows without bound (hcat / vcat, src/ekf.jl:108-109); a member's
// capacity is 3 + 2*(landmarks + extra): `extra` leaves room for the spurious landmarks the
// algorithm itself creates when an observation fails both gates (src/data-association.jl:45-47).
//
// The device functions follow the operation order of the oracle's plain-C twin.  Unlike filter.cu this file is
// compiled with -fmad=true (build.py): the twin rounds products and sums separately, this kernel fuses them, and the
// device libm's sin / cos / atan2 / log differ from glibc's anyway -- the two agree to ~1e-14, with every map size
// equal, not bit for bit (+4.6 % throughput against -fmad=false, profiles/r2_ensemble.md).
// Members are sharded by run across GPUs by the caller (one handle per GPU, no collective).
#include <stdlib.h>

#include <algorithm>
#include <vector>

#include "common.cuh"
#include "kernels.cuh"

namespace {

// literals of sim/ekfslam-sim.jl:29-37,62-76,114
constexpr double WHEELBASE = 4.0;
constexpr double MAX_GAMMA = 60 * SLAM_PI / 180;
constexpr double STEER_RATE = 60 * SLAM_PI / 180;
constexpr double SENSOR_RANGE = 30.0;
constexpr double TARGET_SPEED = 8.0;
constexpr double D_MIN = 1.0;
constexpr double SIGMA_SPEED = 0.5;
constexpr double SIGMA_STEER = 3.0 * SLAM_PI / 180;
constexpr double SIGMA_R = 0.1;
constexpr double SIGMA_B = 1.0 * SLAM_PI / 180;
constexpr double DT = 0.025;
constexpr double DT_OBS = 8 * DT;
constexpr double GATE1 = 4.0, GATE2 = 25.0;
constexpr int ENS_DEFAULT_EXTRA = 5;   // spare landmarks per member beyond the true map
constexpr int MAX_STEPS = 10000;   // the reference's track buffers (sim/ekfslam-sim.jl:24-25)

double sgn(double a) { return (double)(a > 0) - (double)(a < 0); }

struct TruthTables {
    int nsteps = 0, nupd = 0, max_nz = 0;
    std::vector<double> gamma;       // target steering angle after steer!, per step
    std::vector<int> upd_at;         // per step: observation-step index or -1
    std::vector<int> obs_start;      // per observation step: offset into z_true (nupd + 1 entries)
    std::vector<double> z_true;      // 2 x total noise-free observations
    std::vector<double> true_track;  // 3 x nsteps
    std::vector<int> blocks;         // per block {first step, one-past-last step, first observation, nz}
};

// Noise-free half of sim! (sleep, monitor and the filter calls removed).
int build_truth(const double *wp, int nwp, const double *lm, int nl, int nlaps, int max_block, TruthTables &t) {
    double pose[3] = {wp[0], wp[1], atan2(wp[3] - wp[1], wp[2] - wp[0])};   // initial_pose, src/common.jl:93-96
    double g = 0.0;
    int iwp = 1;
    double dtsum = 0;
    while (iwp != 0) {
        if (t.nsteps >= MAX_STEPS) {
            slam_set_error("slam_ens_create: the course needs more than %d steps", MAX_STEPS);
            return SLAM_ERR_CAPACITY;
        }
        {   // steer!, src/common.jl:189-230
            double cx = wp[2 * (iwp - 1)], cy = wp[2 * (iwp - 1) + 1];
            double d2 = (cx - pose[0]) * (cx - pose[0]) + (cy - pose[1]) * (cy - pose[1]);
            bool done = false;
            if (d2 < D_MIN * D_MIN) {
                iwp += 1;
                if (iwp > nwp) {
                    iwp = 0;
                    done = true;
                } else {
                    cx = wp[2 * (iwp - 1)];
                    cy = wp[2 * (iwp - 1) + 1];
                }
            }
            if (!done) {
                double dg = mpi_to_pi(atan2(cy - pose[1], cx - pose[0]) - pose[2] - g);
                double dgmax = STEER_RATE * DT;
                if (fabs(dg) > dgmax) dg = sgn(dg) * dgmax;
                g += dg;
                if (fabs(g) > MAX_GAMMA) g = sgn(g) * MAX_GAMMA;
            }
        }
        if (iwp == 0 && nlaps > 1) {   // sim/ekfslam-sim.jl:88-91
            iwp = 1;
            nlaps -= 1;
        }
        {   // step_vehicle!, src/common.jl:172-181
            double px = pose[0], py = pose[1], phi = pose[2];
            pose[0] = px + TARGET_SPEED * DT * cos(g + phi);
            pose[1] = py + TARGET_SPEED * DT * sin(g + phi);
            pose[2] = mpi_to_pi(phi + TARGET_SPEED * DT * sin(g) / WHEELBASE);
        }
        t.gamma.push_back(g);
        dtsum += DT;
        int u = -1;
        if (dtsum > DT_OBS) {   // strict '>' on the accumulated sum: fires on every 9th step
            dtsum = 0;
            u = t.nupd++;
            t.obs_start.push_back((int)t.z_true.size() / 2);
            int nz = 0;
            for (int i = 0; i < nl; i++) {   // nearby_landmark_indices + get_observations (noise-free)
                double dx = lm[2 * i] - pose[0], dy = lm[2 * i + 1] - pose[1];
                if ((dx * cos(pose[2]) + dy * sin(pose[2])) > 0 &&
                    (dx * dx + dy * dy) < SENSOR_RANGE * SENSOR_RANGE) {
                    t.z_true.push_back(sqrt(dx * dx + dy * dy));
                    t.z_true.push_back(atan2(dy, dx) - pose[2]);
                    nz++;
                }
            }
            t.max_nz = std::max(t.max_nz, nz);
        }
        t.upd_at.push_back(u);
        t.true_track.insert(t.true_track.end(), pose, pose + 3);
        t.nsteps++;
    }
    t.obs_start.push_back((int)t.z_true.size() / 2);
    // blocks of consecutive steps ending with an observation step, at most max_block steps each
    int s0 = 0;
    for (int s = 0; s < t.nsteps; s++) {
        const bool last = s + 1 == t.nsteps;
        if (t.upd_at[s] >= 0 || s + 1 - s0 == max_block || last) {
            const int u = t.upd_at[s];
            t.blocks.push_back(s0);
            t.blocks.push_back(s + 1);
            t.blocks.push_back(u >= 0 ? t.obs_start[u] : 0);
            t.blocks.push_back(u >= 0 ? t.obs_start[u + 1] - t.obs_start[u] : 0);
            s0 = s + 1;
        }
    }
    return SLAM_OK;
}

}  // namespace

// ---------------------------------------------------------------------------------------
// device side
// ---------------------------------------------------------------------------------------
// A run is cut into BLOCKS of consecutive steps that end with an observation step (every 9th
// step for the reference's dt / dt_observe; at most ENS_MAXB steps).  Inside a block the
// predict steps decouple into three chains that depend on each other only through per-step
// scalars, so they are evaluated without a CTA barrier per step and with bit-identical results
// to the step-by-step loop (same operations on the same operands in the same order):
//   (a) heading  phi_{s+1} = mpi_to_pi(phi_s + v_s*dt*sin(g_s)/wheelbase)   -- needs no sin(g+phi)
//       then all sin/cos(g_s + phi_s) of the block in parallel lanes, then x, y prefix sums;
//   (b) the 3x3 pose block  Pvv <- Gv*Pvv*Gv' + Gu*Q*Gu'  on 9 lanes of warp 0 (register chain,
//       shuffles); the noise term Gu*Q*Gu' of every step is precomputed in parallel lanes;
//   (c) rows 0..2 of every map column: thread j applies the block's Gv_s one after the other
//       to its own three values in registers.
// Multiplications by the literal 1 / 0 entries of Gv (src/ekf.jl:24-26) are exact and are
// skipped: a + 0*b == a and 1*a == a for finite operands.
struct EnsArgs {
    int n_runs, nsteps, nl, ncap, mmax, nblocks;
    int ldw;                   // row stride of the rank-major W1 / PHt workspace (multiple of 4)
    int npk;                   // doubles of the packed upper triangle, rounded up to 4
    int wdoubles;              // doubles of the workspace shared by W1 and the association tables
    uint64_t seed;
    uint32_t run0;
    const double *gamma;       // [nsteps]   target steering angle after steer!
    const int4 *blk;           // [nblocks] = {first step, one-past-last step, first observation, nz}
    const double *z_true;      // [2 * total]; the three doubles in front hold the initial pose
    const double *ctrl_noise;  // optional [n_runs][2*nsteps] (speed, steering per step)
    const double *obs_noise;   // optional [n_runs][obs_per_member]
    int obs_per_member;
    int ocap;                  // capacity the outputs are laid out for (>= ncap; the second pass's ncap)
    double *final_x;           // [n_runs][ocap]
    int *n_final;              // [n_runs]
    int *status;               // [n_runs]
    int *steps_done;           // [n_runs]
    double *tracks;            // optional [n_runs][3*nsteps]
    double *final_P;           // optional [n_runs][ocap*ocap], full symmetric
    // Two-pass capacity: the first pass runs every member with a small map capacity (more CTAs per
    // SM); members whose map outgrows it are appended to over_list and run again, from the start, by
    // a second launch with the full capacity (redo_list / redo_count = the first pass's over_*).
    int *over_list, *over_count;
    const int *redo_list, *redo_count;
    // Global-memory tier (members that outgrow every shared-memory capacity): x, P, the W1 / association workspace
    // and the triangle table of redo member i live at gstate + i*gstride; final_x / final_P are then indexed by i too.
    double *gstate;
    size_t gstride;
    // Checkpoints: a member that outgrows this launch's capacity leaves its state (taken after the update of the block
    // in which the append no longer fits) in slot = its index in over_list, and the next launch -- whose redo_list is
    // that over_list -- resumes it there instead of re-running it from its first step (the median diverged filter
    // outgrows the first pass at step ~1540 of 2802).  Same arithmetic on the same operands: bit-identical to a
    // re-run.  Members beyond the slots available are re-run from the start.
    double *ckpt_out;          // [ckpt_out_slots][ckpt_out_stride] or nullptr
    int ckpt_out_slots, ckpt_out_stride;
    const double *ckpt_in;     // the previous launch's checkpoints (taken at capacity ckpt_in_ncap) or nullptr
    int ckpt_in_slots, ckpt_in_stride, ckpt_in_ncap;
    // optional [n_runs]: algorithmic FP64 operations of the member's run by the figures of SURVEY.md 8d -- per
    // predict 30 n; per observation step 80 nz Nf, and with m matches 2m n(n+1) + 20 n m + 8 n m^2 + 8 m^3
    double *flops;
};

constexpr int ENS_MAXB = 9;       // steps per block (dt_observe = 8 dt: an observation every 9th step)
#ifndef ENS_UNROLL_INV
#define ENS_UNROLL_INV 1
#endif
#ifndef ENS_UNROLL_W1
#define ENS_UNROLL_W1 2
#endif
constexpr int ENS_PAR = 8;        // scalars handed between phases: [0] m, [1] k, [2] Cholesky failed
constexpr int BS = 24;            // doubles of block scratch per step
// block scratch slots (per step of the block)
constexpr int B_V = 0, B_G = 1, B_SG = 2, B_CG = 3, B_TPHI = 4, B_PHI0 = 5, B_VTS = 7, B_VTC = 8, B_N = 9,
              B_X0 = 18, B_X1 = 19, B_PHI = 20, B_GT = 21;
#ifdef ENS_TIMING   // developer build (SLAMCUDA_DEFS=-DENS_TIMING): cycles per phase of member 0, thread 0
__device__ long long g_ens_phase[24];
__device__ __forceinline__ void ens_mark(int k) {
    __shared__ long long tlast, acc[24];
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        const long long t = clock64();
        if (k == -1) { for (int q = 0; q < 24; q++) acc[q] = 0; }
        else if (k == -2) { for (int q = 0; q < 24; q++) g_ens_phase[q] += acc[q]; }
        else acc[k] += t - tlast;
        tlast = clock64();
    }
}
#define ENS_MARK(k) ens_mark(k)
#else
#define ENS_MARK(k)
#endif
constexpr int LM_STRIDE = 17;     // association table per landmark: zp[2], inv(S)[4], log det, h[10]
// checkpoint layout (doubles): [1] n, [2] k, [3] block, [4] flops, [5] steps of the block; [8, 35) the block's poses;
// [CK_ZN, ..) zn (2 mmax); then x (ncap of the writer, rounded to 4); then the packed P (ens_pdoubles(n) of them)
constexpr int CK_POSE = 8, CK_ZN = 40;

struct EnsSmem {
    double *x, *P, *W, *S, *cd, *u, *v, *Hn, *z, *zf, *zn, *bs, *par;   // cd: diagonal of inv(U)
    int *fp, *idf, *cls;
    unsigned short *tri;   // (i', j' << 8), j' <= i' < 32, row by row: the Cholesky's element order
};

__host__ __device__ inline int ens_ru4(int a) { return (a + 3) & ~3; }
// doubles of the packed upper triangle of a member's P (layout: see cofs)
__host__ __device__ inline int ens_pdoubles(int ncap) {
    const int nbc = (ncap + 3) >> 2;
    return 8 * nbc * (nbc + 1);
}
__host__ __device__ inline int ens_wdoubles(int ncap, int mmax) {
    const int a = 2 * mmax * ens_ru4(ncap), b = ((ncap - 3) / 2) * LM_STRIDE;
    return ens_ru4(a > b ? a : b);
}
// S (then U in its upper triangle, inv(U) transposed in its strict lower triangle) has the leading
// dimension R + 2: even, so that two inverse entries of a row load as one 16-byte word, and with two
// columns to spare for the four-column groups of the W1 product.
__host__ __device__ inline int ens_lds(int R) { return R + 2; }
// The triangle table: entry e = (row a, column b <= a) of a lower triangle numbered row by row.  The
// Cholesky reads it as its trailing elements (R <= 32), the downdate as its 16 x 16 warp regions (bj = a, bi = b).
__host__ __device__ inline int ens_ntri(int ncap, int mmax) {
    const int R = 2 * mmax < 32 ? 2 * mmax : 32, nb = (ncap + 15) >> 4;
    const int a = R * (R - 1) / 2, b = nb * (nb + 1) / 2;
    return ens_ru4(a > b ? a : b);
}
// A member's storage has two parts.  The STATE part grows with the capacity: x, the packed triangle of P, the
// W1 / association workspace and the triangle table; it lives in shared memory for the shared tiers and in global
// memory for the last-resort tier.  The FIXED part (the 2m x 2m square and the per-observation vectors) depends on
// the largest number of observations per step only and is always in shared memory.
__host__ __device__ inline size_t ens_state_doubles(int ncap, int mmax) {
    return (size_t)ens_ru4(ncap) + ens_pdoubles(ncap) + ens_wdoubles(ncap, mmax) + ens_ru4((ens_ntri(ncap, mmax) + 3) / 4);
}
__host__ __device__ inline size_t ens_fixed_bytes(int mmax) {
    const int R = 2 * mmax;
    size_t d = ens_ru4(R * ens_lds(R)) + 3 * ens_ru4(R) + ens_ru4(10 * mmax) + 3 * ens_ru4(2 * mmax) + ENS_MAXB * BS + ENS_PAR;
    return d * 8 + ens_ru4(3 * mmax) * 4 + 32;
}
__host__ __device__ inline int ens_ckpt_doubles(int ncap, int mmax) {
    return CK_ZN + ens_ru4(2 * mmax) + ens_ru4(ncap) + ens_pdoubles(ncap);
}
__host__ __device__ inline size_t ens_smem_bytes(int ncap, int mmax) {
    return ens_state_doubles(ncap, mmax) * 8 + ens_fixed_bytes(mmax);
}

__device__ inline EnsSmem ens_carve(double *state, double *fixed, int ncap, int mmax) {
    EnsSmem s;
    const int R = 2 * mmax;
    double *p = state;
    s.x = p;   p += ens_ru4(ncap);
    s.P = p;   p += ens_pdoubles(ncap);
    s.W = p;   p += ens_wdoubles(ncap, mmax);
    s.tri = reinterpret_cast<unsigned short *>(p);
    p = fixed;
    s.S = p;   p += ens_ru4(R * ens_lds(R));
    s.cd = p;  p += ens_ru4(R);
    s.u = p;   p += ens_ru4(R);
    s.v = p;   p += ens_ru4(R);
    s.Hn = p;  p += ens_ru4(10 * mmax);
    s.z = p;   p += ens_ru4(2 * mmax);
    s.zf = p;  p += ens_ru4(2 * mmax);
    s.zn = p;  p += ens_ru4(2 * mmax);
    s.bs = p;  p += ENS_MAXB * BS;
    s.par = p; p += ENS_PAR;
    int *q = reinterpret_cast<int *>(p);
    s.fp = q;  q += mmax;
    s.idf = q; q += mmax;
    s.cls = q;
    return s;
}

// One copy of the FP64 sin / cos / normal code for all call sites: the kernel streams through its
// whole instruction footprint once per block, so code size is fetch latency.
__device__ __noinline__ void ens_sincos(double t, double *s, double *c) {
    sincos(t, s, c);   // one argument reduction for both
}
__device__ __noinline__ double ens_normal(uint64_t seed, uint32_t run, uint32_t stream, uint32_t idx) {
    return philox_normal(seed, run, stream, idx);
}

// idx / d for the small index ranges of this kernel (idx < 2^17, quotient < 2^8) without the
// ~70-cycle integer division: (idx + 0.5) / d is at least 0.5 / d away from an integer, far more
// than the float rounding error.
__device__ __forceinline__ int ens_div(int idx, float rinv) { return (int)(((float)idx + 0.5f) * rinv); }

// D (8 x 8, a lane holds d.x = D[lane/4][2(lane%4)], d.y the next column) += A (8 x 4, a = A[lane/4][lane%4])
// * B (4 x 8, b = B[lane%4][lane/4])
__device__ __forceinline__ void ens_dmma(double2 &d, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};"
                 : "+d"(d.x), "+d"(d.y)
                 : "d"(a), "d"(b));
}

// packed upper triangle, column-major: P(i, j), i <= j, at j(j+1)/2 + i
// Columns are grouped in fours; the four columns of group b = j/4 all hold rows 0 .. 4b+3, so every
// 4 x 4 tile of the upper triangle -- the diagonal ones too -- is a full, 32-byte-aligned block:
// column j starts at 8b(b+1) + (j%4)(4b+4) = 4(b+1)(2b + j%4).  The entries below the diagonal inside
// a diagonal tile are scratch (the downdate updates them blindly; nothing reads them).
__device__ __forceinline__ int cofs(int j) {
    const int b = j >> 2;
    return 4 * (b + 1) * (2 * b + (j & 3));
}
__device__ __forceinline__ double psym(const double *P, int i, int j) {
    return i <= j ? P[cofs(j) + i] : P[cofs(i) + j];
}

// ---- the part of a block that does not depend on the filter state ----------------------------
// add_control_noise! (sim/sim-utils.jl:35-38: speed first, then steering) and the noisy
// observations of get_observations (sim/sim-utils.jl:53-75: noise-free table + this member's
// noise, draw order = all range noises of the step, then all bearing noises), plus sin / cos of
// the measured steering angle.  Run by the HELPER warps (tid >= 32) while warp 0 is busy with a
// serial phase of the previous block (the innovation Cholesky), so its latency is hidden.
template <int T>
__device__ __noinline__ void ens_stage_block(const EnsSmem &sm, const EnsArgs &a, int b, uint32_t run,
                                             const double *ctrl, const double *obsn) {
    const int ht = (int)threadIdx.x - 32;
    const int4 d = a.blk[b];
    const int s0 = d.x, len = d.y - d.x, o0 = d.z;
    int nz = d.w;
    if (obsn && 2 * o0 + 2 * nz > a.obs_per_member) nz = 0;   // the block loop reports the overflow
    const int cursor = 2 * o0;
    // one standard normal per item: 2*len control items (step, speed | steering), then 2*nz observation items
#pragma unroll 1
    for (int item = ht; item < 2 * len + 2 * nz; item += T - 32) {
        const bool is_ctrl = item < 2 * len;
        const int c = is_ctrl ? item : item - 2 * len;
        double e;
        if (is_ctrl ? ctrl != nullptr : obsn != nullptr) {
            e = is_ctrl ? ctrl[2 * s0 + c] : obsn[cursor + c];
        } else {
            e = is_ctrl ? ens_normal(a.seed, run, (uint32_t)(c & 1), (uint32_t)(s0 + (c >> 1)))
                        : ens_normal(a.seed, run, 2u, (uint32_t)(cursor + c));
        }
        if (is_ctrl) {
            const int which = c & 1;
            sm.bs[(c >> 1) * BS + which] = which == 0 ? TARGET_SPEED + e * sqrt(SIGMA_SPEED * SIGMA_SPEED)
                                                      : a.gamma[s0 + (c >> 1)] + e * sqrt(SIGMA_STEER * SIGMA_STEER);
        } else {
            const int ob = c < nz ? c : c - nz, comp = c < nz ? 0 : 1;
            sm.z[2 * ob + comp] = a.z_true[2 * (o0 + ob) + comp] +
                                  e * (comp == 0 ? sqrt(SIGMA_R * SIGMA_R) : sqrt(SIGMA_B * SIGMA_B));
        }
    }
    asm volatile("bar.sync 1, %0;" ::"r"(T - 32) : "memory");   // helper warps only
#pragma unroll 1
    for (int st = ht; st < len; st += T - 32) {
        double *bq = sm.bs + st * BS;
        const double v = bq[B_V], g = bq[B_G];
        double sg, cg;
        ens_sincos(g, &sg, &cg);
        bq[B_SG] = sg;
        bq[B_CG] = cg;
        bq[B_TPHI] = v * DT * sg / WHEELBASE;
    }
}

// ---- predict for all steps [s0, s0 + len) of a staged block, src/ekf.jl:8-43 ------------------
// On return x, P hold the state after the block's last predict; the pose after every step is in
// bs[..][B_X0..B_PHI].  If nz > 0 the helper warps also fill zp / H of every mapped landmark
// (predict_observation, src/common.jl:139-165) into the association table while warp 0 runs the
// 3x3 chain.
template <int T>
__device__ void ens_predict_block(const EnsSmem &sm, int n, int len, int nz) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    double *bs = sm.bs;
    if (warp == 0) {
        if (lane == 0) {   // chain (a): heading; one lane, branch-free wrap, increments loaded a step ahead
            double phi = sm.x[2], tn = bs[B_TPHI];
#pragma unroll 1
            for (int s = 0; s < len; s++) {
                const double t = phi + tn;
                tn = bs[(s + 1 < len ? s + 1 : s) * BS + B_TPHI];
                const double tm = t - 2 * SLAM_PI, tp = t + 2 * SLAM_PI;      // mpi_to_pi, src/common.jl:102-110
                const double w = t > SLAM_PI ? tm : (t < -SLAM_PI ? tp : t);
                bs[s * BS + B_PHI0] = phi;
                bs[s * BS + B_PHI] = w;
                phi = w;
            }
            sm.x[2] = phi;
        }
        __syncwarp();
        ENS_MARK(18);   // heading chain done
        if (lane < len) {
            double *b = bs + lane * BS;
            const double v = b[B_V], g = b[B_G], phi = b[B_PHI0];
            double s, c;
            ens_sincos(g + phi, &s, &c);
            const double vts = v * DT * s, vtc = v * DT * c;
            b[B_VTS] = vts;
            b[B_VTC] = vtc;
            b[B_GT] = -vts;       // Gv[0][2], Gv[1][2], Gv[2][2] - 1 of this step for the 3x3 chain
            b[B_GT + 1] = vtc;
            b[B_GT + 2] = 0.0;
            // Gu*Q*Gu' (src/ekf.jl:27-29,32)
            const double Q[4] = {SIGMA_SPEED * SIGMA_SPEED, 0, 0, SIGMA_STEER * SIGMA_STEER};
            const double Gu[3][2] = {{DT * c, -vts}, {DT * s, vtc}, {DT * b[B_SG] / WHEELBASE, v * DT * b[B_CG] / WHEELBASE}};
            double GQ[3][2];
#pragma unroll
            for (int i = 0; i < 3; i++)
#pragma unroll
                for (int j = 0; j < 2; j++) GQ[i][j] = Gu[i][0] * Q[0 + 2 * j] + Gu[i][1] * Q[1 + 2 * j];
#pragma unroll
            for (int i = 0; i < 3; i++)
#pragma unroll
                for (int j = 0; j < 3; j++) b[B_N + 3 * i + j] = GQ[i][0] * Gu[j][0] + GQ[i][1] * Gu[j][1];
        }
        ENS_MARK(15);   // heading chain + sin/cos + Gu*Q*Gu' done
    }
    __syncthreads();
    ENS_MARK(1);
    if (warp == 0) {
        // chain (b): lane 3*i + j holds Pvv[i][j]; only the upper lanes (i <= j) matter and they read
        // upper lanes only: A[i][j] = P[i][j] + gi*P[j][2], A[i][2] = P[i][2] + gi*P[2][2] with
        // gi = Gv[i][2] = (-vts, vtc, 0); B[i][j] = A[i][j] + A[i][2]*gj; then + (Gu*Q*Gu')[i][j].
        const int l9 = lane < 9 ? lane : 0;
        const int i = l9 / 3, j = l9 % 3;
        double p = psym(sm.P, i, j);
        // the step's scalars (gi, gj from the table phase (i) left, the noise term) are loaded one
        // iteration ahead so that only shuffle + 5 FP64 operations sit on the loop-carried chain
        double gi = bs[B_GT + i], gj = bs[B_GT + j], nn = bs[B_N + l9];
#pragma unroll 1
        for (int s = 0; s < len; s++) {
            const double *bn = bs + (s + 1 < len ? s + 1 : s) * BS;
            const double gi_n = bn[B_GT + i], gj_n = bn[B_GT + j], nn_n = bn[B_N + l9];
            const double pj2 = __shfl_sync(0xffffffffu, p, 3 * j + 2), pi2 = __shfl_sync(0xffffffffu, p, 3 * i + 2),
                         p22 = __shfl_sync(0xffffffffu, p, 8);
            const double aj = p + gi * pj2;
            const double a2 = pi2 + gi * p22;
            p = (aj + a2 * gj) + nn;
            gi = gi_n; gj = gj_n; nn = nn_n;
        }
        if (lane < 9 && i <= j) sm.P[cofs(j) + i] = p;
        ENS_MARK(16);   // 3x3 chain done (barrier wait for the helper warps follows)
    } else {
        // x += vtc_s (lane 0), y += vts_s (lane 1) of the first helper warp, one step after the other
        if (tid - 32 < 2) {
            const int slot = tid == 32 ? B_VTC : B_VTS;
            double acc = sm.x[tid - 32], vn = bs[slot];
#pragma unroll 1
            for (int s = 0; s < len; s++) {
                acc = acc + vn;
                vn = bs[(s + 1 < len ? s + 1 : s) * BS + slot];
                bs[s * BS + B_X0 + (tid - 32)] = acc;
            }
            sm.x[tid - 32] = acc;
        }
        // chain (c): rows 0..2 of map column jc
#pragma unroll 1
        for (int jc = 3 + tid - 32; jc < n; jc += T - 32) {
            double *col = sm.P + cofs(jc);
            double p0 = col[0], p1 = col[1];
            const double p2 = col[2];
#pragma unroll 1
            for (int s = 0; s < len; s++) {
                p0 = p0 + (-bs[s * BS + B_VTS]) * p2;
                p1 = p1 + bs[s * BS + B_VTC] * p2;
            }
            col[0] = p0;
            col[1] = p1;
        }
        if (nz > 0) {
            asm volatile("bar.sync 1, %0;" ::"r"(T - 32) : "memory");   // helper warps: the pose is final
            const int Nf = (n - 3) / 2;
#pragma unroll 1
            for (int jl = tid - 32; jl < Nf; jl += T - 32) {
                const int f = 3 + 2 * jl;
                double *t = sm.W + jl * LM_STRIDE;
                observe_model(sm.x[0], sm.x[1], sm.x[2], sm.x[f], sm.x[f + 1], t, t + 7);
            }
        }
    }
    __syncthreads();
    ENS_MARK(2);
}

// ---- associate, src/data-association.jl:1-63 ---------------------------------------------
// Phase A: one thread per landmark builds everything of compute_association that does not
// depend on the observation (S = H*P*H' + R from the 5x5 sub-block, inv(S), log det S; zp and H
// come from ens_predict_block);
// phase B: one warp per observation scores all landmarks and reduces (nd, j) / min nis with
// shuffles -- strict '<' everywhere, lowest j wins ties, NaN never wins, which is the result of
// the reference's sequential scan (`outer` is only consulted when no landmark was ever gated).
// Leaves m, k in par[0], par[1], zf / idf / zn compacted in observation order, and Hn / fp / v
// of the matched observations ready for update (the same values update recomputes, src/ekf.jl:55-59).
template <int T>
__device__ void ens_associate(const EnsSmem &sm, int n, int nz) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int Nf = (n - 3) / 2;
    double *LM = sm.W;
    const double Rm[4] = {SIGMA_R * SIGMA_R, 0, 0, SIGMA_B * SIGMA_B};
#pragma unroll 1
    for (int jl = tid; jl < Nf; jl += T) {
        const int f = 3 + 2 * jl;
        double *t = LM + jl * LM_STRIDE;     // zp and h were filled by ens_predict_block
        double h[10];
#pragma unroll
        for (int q = 0; q < 10; q++) h[q] = t[7 + q];
        const int c[5] = {0, 1, 2, f, f + 1};
        const double Hr[2][5] = {{h[0], h[1], h[2], h[6], h[7]}, {h[3], h[4], h[5], h[8], h[9]}};
        double P5[5][5];
#pragma unroll
        for (int p = 0; p < 5; p++)
#pragma unroll
            for (int q = p; q < 5; q++) {
                const double val = sm.P[cofs(c[q]) + c[p]];
                P5[p][q] = val;
                P5[q][p] = val;
            }
        double HP[2][5];
#pragma unroll
        for (int a = 0; a < 2; a++)
#pragma unroll
            for (int q = 0; q < 5; q++) {
                double acc = 0;
#pragma unroll
                for (int p = 0; p < 5; p++) acc += Hr[a][p] * P5[p][q];
                HP[a][q] = acc;
            }
        double S[4];
#pragma unroll
        for (int a = 0; a < 2; a++)
#pragma unroll
            for (int b = 0; b < 2; b++) {
                double acc = 0;
#pragma unroll
                for (int q = 0; q < 5; q++) acc += HP[a][q] * Hr[b][q];
                S[2 * a + b] = acc + Rm[a + 2 * b];
            }
        const double det = S[0] * S[3] - S[1] * S[2];
        t[2] = S[3] / det;
        t[3] = -S[1] / det;
        t[4] = -S[2] / det;
        t[5] = S[0] / det;
        t[6] = log(det);
    }
    __syncthreads();
    ENS_MARK(3);
#pragma unroll 1
    for (int i = warp; i < nz; i += T / 32) {
        const double z0 = sm.z[2 * i], z1 = sm.z[2 * i + 1];
        double nbest = INFINITY;
        int jbest = 0;
        bool near = false;   // some landmark has nis <= gate2  <=>  !(outer > gate2) for outer = min nis
#pragma unroll 1
        for (int jl = lane; jl < Nf; jl += 32) {
            const double *t = LM + jl * LM_STRIDE;
            const double v0 = z0 - t[0];
            const double v1 = mpi_to_pi(z1 - t[1]);
            const double t0 = t[2] * v0 + t[3] * v1;
            const double t1 = t[4] * v0 + t[5] * v1;
            const double nis = v0 * t0 + v1 * t1;
            const double nd = nis + t[6];
            if (nis < GATE1 && nd < nbest) {
                nbest = nd;
                jbest = jl + 1;
            }
            near = near || nis <= GATE2;
        }
        // warp argmin of (nd, j) with three 32-bit min reductions (REDUX): nd as an order-preserving
        // 64-bit key (hi word, then lo word among the hi winners), then the lowest j among equal keys
        const long long nb = __double_as_longlong(nbest + 0.0);   // -0 -> +0
        const unsigned long long key =
            jbest ? (nb < 0 ? ~(unsigned long long)nb : ((unsigned long long)nb | 0x8000000000000000ull)) : ~0ull;
        const unsigned hi = (unsigned)(key >> 32), lo = (unsigned)key;
        const unsigned mhi = __reduce_min_sync(0xffffffffu, hi);
        const unsigned mlo = __reduce_min_sync(0xffffffffu, hi == mhi ? lo : 0xffffffffu);
        const unsigned jm = __reduce_min_sync(0xffffffffu, (jbest && hi == mhi && lo == mlo) ? (unsigned)jbest : 0xffffffffu);
        jbest = jm == 0xffffffffu ? 0 : (int)jm;
        const double outer = __any_sync(0xffffffffu, near) ? 0.0 : INFINITY;
        if (lane == 0) sm.cls[i] = jbest != 0 ? jbest : (outer > GATE2 ? -1 : 0);
    }
    __syncthreads();
    ENS_MARK(4);
    if (warp == 0) {
        // stable compaction in observation order (src/data-association.jl:43-48) by ballot / popc, and
        // what update needs of the matched observations: fpos, innovation, the Jacobian's nonzeros
        int m = 0, k = 0;
#pragma unroll 1
        for (int base = 0; base < nz; base += 32) {
            const int i = base + lane;
            const int c = i < nz ? sm.cls[i] : 0;
            const unsigned mf = __ballot_sync(0xffffffffu, c > 0), mn = __ballot_sync(0xffffffffu, c < 0);
            const unsigned below = (1u << lane) - 1u;
            if (c > 0) {
                const int o = m + __popc(mf & below), jl = c - 1;
                const double z0 = sm.z[2 * i], z1 = sm.z[2 * i + 1];
                const double *t = LM + jl * LM_STRIDE;
                sm.zf[2 * o] = z0;
                sm.zf[2 * o + 1] = z1;
                sm.idf[o] = c;
                sm.fp[o] = 3 + 2 * jl;
                sm.v[2 * o] = z0 - t[0];
                sm.v[2 * o + 1] = mpi_to_pi(z1 - t[1]);
#pragma unroll
                for (int q = 0; q < 10; q++) sm.Hn[10 * o + q] = t[7 + q];
            } else if (c < 0) {
                const int o = k + __popc(mn & below);
                sm.zn[2 * o] = sm.z[2 * i];
                sm.zn[2 * o + 1] = sm.z[2 * i + 1];
            }
            m += __popc(mf);
            k += __popc(mn);
        }
        if (lane == 0) {
            sm.par[0] = m;
            sm.par[1] = k;
        }
    }
    __syncthreads();
    ENS_MARK(6);
}

// ---- update, src/ekf.jl:46-77, in the operation order of the oracle's C twin ---------------
// PHt / W1 are rank-major (W[k*ldw + i]).  The 2m x 2m Cholesky, the triangular inverse and
// u = C'v run in ONE warp (lane = column, __syncwarp between the column steps).
template <int T>
__device__ bool ens_update(const EnsSmem &sm, int n, int ldw, int m, const EnsArgs &a, int stage_b, uint32_t run,
                           const double *ctrl, const double *obsn) {
    const int tid = threadIdx.x;
    const int R = 2 * m, lds = ens_lds(R);
    const double Rm[4] = {SIGMA_R * SIGMA_R, 0, 0, SIGMA_B * SIGMA_B};
    double *W = sm.W;
    // PHt = P*H' (five terms per entry, columns 1,2,3,f,f+1); rows n..ldw-1 are zeroed for K6's tiles
    const float rinv_ldw = 1.0f / (float)ldw, rinv_R = 1.0f / (float)R;
#pragma unroll 1
    for (int idx = tid; idx < ldw * m; idx += T) {
        const int o = ens_div(idx, rinv_ldw), i = idx - o * ldw;
        double w0 = 0.0, w1 = 0.0;
        if (i < n) {
            const double *h = sm.Hn + 10 * o;
            const int f = sm.fp[o];
            const double p0 = psym(sm.P, i, 0), p1 = psym(sm.P, i, 1), p2 = psym(sm.P, i, 2), p3 = psym(sm.P, i, f),
                         p4 = psym(sm.P, i, f + 1);
            w0 = p0 * h[0] + p1 * h[1] + p2 * h[2] + p3 * h[6] + p4 * h[7];
            w1 = p0 * h[3] + p1 * h[4] + p2 * h[5] + p3 * h[8] + p4 * h[9];
        }
        W[(2 * o) * ldw + i] = w0;
        W[(2 * o + 1) * ldw + i] = w1;
    }
    __syncthreads();
    ENS_MARK(7);
    // S = H*PHt + RR
#pragma unroll 1
    for (int idx = tid; idx < m * R; idx += T) {
        const int o = ens_div(idx, rinv_R), k = idx - o * R;
        const double *h = sm.Hn + 10 * o;
        const int f = sm.fp[o];
        const double *wk = W + k * ldw;
        const double a0 = wk[0], a1 = wk[1], a2 = wk[2], a3 = wk[f], a4 = wk[f + 1];
        double s0 = h[0] * a0 + h[1] * a1 + h[2] * a2 + h[6] * a3 + h[7] * a4;
        double s1 = h[3] * a0 + h[4] * a1 + h[5] * a2 + h[8] * a3 + h[9] * a4;
        if (k == 2 * o) { s0 += Rm[0]; s1 += Rm[1]; }
        if (k == 2 * o + 1) { s0 += Rm[2]; s1 += Rm[3]; }
        sm.S[(2 * o) + k * lds] = s0;
        sm.S[(2 * o + 1) + k * lds] = s1;
    }
    __syncthreads();
    ENS_MARK(8);
#pragma unroll 1
    for (int idx = tid; idx < R * R; idx += T) {   // S = (S + S')*0.5; only the upper triangle is used
        const int b = ens_div(idx, rinv_R), a = idx - b * R;
        if (a < b) sm.S[a + b * lds] = (sm.S[a + b * lds] + sm.S[b + a * lds]) * 0.5;
    }
    __syncthreads();
    ENS_MARK(9);
    if (tid < 32) {   // one warp: right-looking upper Cholesky, then C = inv(U) and u = C'v
        const int lane = tid;
        int fail = 0;
#pragma unroll 1
        for (int k = 0; k < R; k++) {
            const double d = sm.S[k + k * lds];
            if (!(d > 0)) { fail = 1; break; }    // uniform: every lane reads the same d
            const double ukk = sqrt(d);
            __syncwarp();
            if (lane == 0) sm.S[k + k * lds] = ukk;
#pragma unroll 1
            for (int j = k + 1 + lane; j < R; j += 32) {
                const double ukj = sm.S[k + j * lds] / ukk;
                sm.S[k + j * lds] = ukj;
            }
            __syncwarp();
            if (R <= 32) {
                // trailing update, one ELEMENT per lane: entry e of the triangle table is (i', j') with
                // j' <= i', rows counted from the bottom (i = R-1-i'), so the elements below row k are the
                // first r(r+1)/2 entries, r = R-1-k.  Same subtraction per element as the column loop.
                // (Keeping a lane's table entries in registers was tried: fewer cycles for a lone CTA,
                // but more instructions and registers, and 6 % less throughput at full occupancy.)
                const int r = R - 1 - k;
#pragma unroll 1
                for (int e = lane; e < r * (r + 1) / 2; e += 32) {
                    const int t = sm.tri[e];
                    const int i = R - 1 - (t & 0xff), j = R - 1 - (t >> 8);
                    sm.S[i + j * lds] -= sm.S[k + i * lds] * sm.S[k + j * lds];
                }
            } else {
#pragma unroll 1
                for (int j = k + 1 + lane; j < R; j += 32) {
                    const double ukj = sm.S[k + j * lds];
#pragma unroll 1
                    for (int i = k + 1; i <= j; i++) sm.S[i + j * lds] -= sm.S[k + i * lds] * ukj;
                }
            }
            __syncwarp();
        }
        if (lane == 0) sm.par[2] = fail;
        ENS_MARK(17);   // Cholesky done (inverse + u follow)
        if (!fail) {
            // C = inv(U) in the order of LAPACK's unblocked dtrti2: C[i][j] = (U[i][j]*C[i][i] +
            // sum_{i<jj<j} U[jj][j]*C[i][jj]) * (-1/U[j][j]).  Row i of C depends on row i alone, so
            // lane i builds its own row with no exchange; the reciprocals are taken up front.
            // Storage: C[i][j], i < j, goes to the strict LOWER triangle of S at (j, i) -- row i of C is
            // the contiguous column i there -- and the diagonal to cd; U stays in the upper triangle.
#pragma unroll 1
            for (int j = lane; j < R; j += 32) {
                const double cjj = 1.0 / sm.S[j + j * lds];
                sm.cd[j] = cjj;
                sm.u[j] = -cjj;   // ajj, parked in u until u is formed below
            }
            __syncwarp();
#pragma unroll 1
            for (int i = lane; i < R; i += 32) {
                const double cii = sm.cd[i];
                double *ci = sm.S + i * lds;     // ci[jj] = C[i][jj]
#pragma unroll 1
                for (int j = i + 1; j < R; j++) {
                    double y = sm.S[i + j * lds] * cii;
                    SLAM_UNROLL(ENS_UNROLL_INV)
                    for (int jj = i + 1; jj < j; jj++) y += sm.S[jj + j * lds] * ci[jj];
                    ci[j] = y * sm.u[j];
                }
            }
            __syncwarp();
#pragma unroll 1
            for (int j = lane; j < R; j += 32) {
                double a = 0;
                    SLAM_UNROLL(ENS_UNROLL_INV)
                for (int l = 0; l < j; l++) a += sm.S[j + l * lds] * sm.v[l];
                a += sm.cd[j] * sm.v[j];
                sm.u[j] = a;          // ajj is dead after the barrier above
            }
        }
    } else if (stage_b < a.nblocks) {
        ens_stage_block<T>(sm, a, stage_b, run, ctrl, obsn);   // hidden behind the Cholesky
    }
    __syncthreads();
    ENS_MARK(10);
    if (sm.par[2] != 0.0) return false;
    // W1 = PHt*C in place, x += W1*u.  Four columns k at a time, groups descending (a group needs
    // PHt[i, l <= its last column] only, which later groups leave intact); the columns of a group share
    // the loads of PHt[i, l], and each keeps its own sum (l ascending, the diagonal term last), i.e.
    // the result of one column at a time.  R is even, so a group holds four or two columns; the
    // operands a two-column group reads past column R-1 stay inside the workspace and are not used.
#pragma unroll 1
    for (int i = tid; i < n; i += T) {
        double *wi = W + i;
#pragma unroll 1
        for (int kg = (R - 1) & ~3; kg >= 0; kg -= 4) {
            double a0 = 0, a1 = 0, a2 = 0, a3 = 0;
            const double *cl = sm.S + kg;          // C[l][kg + q] at cl[l*lds + q]
            SLAM_UNROLL(ENS_UNROLL_W1)
            for (int l = 0; l < kg; l++) {
                const double w = wi[l * ldw];
                const double2 c01 = *reinterpret_cast<const double2 *>(cl + l * lds);
                const double2 c23 = *reinterpret_cast<const double2 *>(cl + l * lds + 2);
                a0 += w * c01.x;
                a1 += w * c01.y;
                a2 += w * c23.x;
                a3 += w * c23.y;
            }
            const double w0 = wi[kg * ldw], w1 = wi[(kg + 1) * ldw], w2 = wi[(kg + 2) * ldw], w3 = wi[(kg + 3) * ldw];
            const double *c0 = cl + kg * lds, *c1 = c0 + lds, *c2 = c1 + lds;
            a0 += w0 * sm.cd[kg];
            a1 += w0 * c0[1];
            a1 += w1 * sm.cd[kg + 1];
            a2 += w0 * c0[2];
            a2 += w1 * c1[2];
            a2 += w2 * sm.cd[kg + 2];
            a3 += w0 * c0[3];
            a3 += w1 * c1[3];
            a3 += w2 * c2[3];
            a3 += w3 * sm.cd[kg + 3];
            wi[kg * ldw] = a0;
            wi[(kg + 1) * ldw] = a1;
            if (kg + 2 < R) {
                wi[(kg + 2) * ldw] = a2;
                wi[(kg + 3) * ldw] = a3;
            }
        }
        double dxi = 0;
#pragma unroll 1
        for (int k = 0; k < R; k++) dxi += wi[k * ldw] * sm.u[k];
        sm.x[i] += dxi;
    }
    __syncthreads();
    ENS_MARK(11);
    // P -= W1*W1' over the packed upper triangle on the FP64 tensor pipe (mma.sync m8n8k4, SASS DMMA): a
    // warp owns a 16 x 16 region = 2 x 2 MMA tiles, four ranks per instruction.  The MMA is issued on the
    // TRANSPOSED tile, D'[j][i] -= W1[j][k] W1[i][k], so a lane's accumulator pair is two consecutive rows
    // (i, i+1) of one column j of P: one 16-byte access of the packed layout.  A = -W1[j][k], B = W1[i][k]
    // both come from the rank-major workspace (one 8-byte load per lane each); ranks past R are fed
    // as zeros.  On this hardware the four products of an MMA accumulate as a k-ascending fma chain (the
    // measured difference to the FMA kernel K6 is 0, downdate_dmma.cu), so the result is the chain
    // a = fma(-W1[i,k], W1[j,k], a), k ascending, of the twin.  An MMA replaces 8 warp-wide DFMAs: the
    // kernel is bound by instruction issue and fetch, not by the FP64 pipe.
    // Entries that do not exist in the packed layout (the part of a diagonal region below its 4 x 4 diagonal
    // blocks, columns past the allocation) are neither loaded nor stored; columns n .. ldw-1 have W1 = 0.
    {
        const int lane = tid & 31, warp = tid >> 5, g = lane >> 2, t = lane & 3;
        const int nb16 = (n + 15) >> 4;
#pragma unroll 1
        for (int tl = warp; tl < nb16 * (nb16 + 1) / 2; tl += T / 32) {
            const int tt = sm.tri[tl];
            const int bj = tt & 0xff, bi = tt >> 8;
            const int i0 = 16 * bi, j0 = 16 * bj;
            const bool diag = bi == bj;
            double2 c[2][2];
            bool ok[2][2];
            int off[2][2];
#pragma unroll
            for (int sj = 0; sj < 2; sj++)
#pragma unroll
                for (int si = 0; si < 2; si++) {
                    const int j = j0 + 8 * sj + g, i = i0 + 8 * si + 2 * t;
                    ok[sj][si] = j < ldw && i + 1 <= (j | 3) && !(diag && si > sj);
                    off[sj][si] = cofs(j) + i;
                    c[sj][si] = ok[sj][si] ? *reinterpret_cast<const double2 *>(sm.P + off[sj][si]) : make_double2(0.0, 0.0);
                }
#pragma unroll 1
            for (int k0 = 0; k0 < R; k0 += 4) {
                const int k = k0 + t;
                const double *wk = W + k * ldw;
                double a[2], b[2];
#pragma unroll
                for (int q = 0; q < 2; q++) {
                    const int j = j0 + 8 * q + g, i = i0 + 8 * q + g;
                    a[q] = (k < R && j < ldw) ? -wk[j] : 0.0;
                    b[q] = (k < R && i < ldw) ? wk[i] : 0.0;
                }
                ens_dmma(c[0][0], a[0], b[0]);
                ens_dmma(c[1][0], a[1], b[0]);
                ens_dmma(c[1][1], a[1], b[1]);
                if (!diag) ens_dmma(c[0][1], a[0], b[1]);
            }
#pragma unroll
            for (int sj = 0; sj < 2; sj++)
#pragma unroll
                for (int si = 0; si < 2; si++)
                    if (ok[sj][si]) *reinterpret_cast<double2 *>(sm.P + off[sj][si]) = c[sj][si];
        }
    }
    __syncthreads();
    ENS_MARK(12);
    return true;
}

// ---- add_features, src/ekf.jl:84-122, in place (the member's capacity is fixed) --------------
template <int T>
__device__ int ens_add_features(const EnsSmem &sm, int n, int k) {
    const int tid = threadIdx.x;
    const double phi = sm.x[2];
    const double Rm[4] = {SIGMA_R * SIGMA_R, 0, 0, SIGMA_B * SIGMA_B};
#pragma unroll 1
    for (int q = 0; q < k; q++) {
        const int len = n;
        const double r = sm.zn[2 * q], b = sm.zn[2 * q + 1];
        double s, c;
        ens_sincos(phi + b, &s, &c);
        const double Gv[2][3] = {{1, 0, -r * s}, {0, 1, r * c}};
        if (tid == 0) {
            sm.x[len] = sm.x[0] + r * c;
            sm.x[len + 1] = sm.x[1] + r * s;
            const double Gz[2][2] = {{c, -r * s}, {s, r * c}};
            double GP[2][3], GR[2][2];
            for (int a = 0; a < 2; a++)
                for (int j = 0; j < 3; j++)
                    GP[a][j] = Gv[a][0] * psym(sm.P, 0, j) + Gv[a][1] * psym(sm.P, 1, j) + Gv[a][2] * psym(sm.P, 2, j);
            for (int a = 0; a < 2; a++)
                for (int j = 0; j < 2; j++) GR[a][j] = Gz[a][0] * Rm[0 + 2 * j] + Gz[a][1] * Rm[1 + 2 * j];
            for (int a = 0; a < 2; a++)
                for (int bb = a; bb < 2; bb++) {
                    const double t1 = GP[a][0] * Gv[bb][0] + GP[a][1] * Gv[bb][1] + GP[a][2] * Gv[bb][2];
                    const double t2 = GR[a][0] * Gz[bb][0] + GR[a][1] * Gz[bb][1];
                    sm.P[cofs(len + bb) + len + a] = t1 + t2;
                }
            for (int a = 0; a < 2; a++)
                for (int j = 0; j < 3; j++) sm.P[cofs(len + a) + j] = GP[a][j];
        }
#pragma unroll 1
        for (int j = 3 + tid; j < len; j += T) {
            const double c0 = sm.P[cofs(j)], c1 = sm.P[cofs(j) + 1], c2 = sm.P[cofs(j) + 2];
            for (int a = 0; a < 2; a++)
                sm.P[cofs(len + a) + j] = Gv[a][0] * c0 + Gv[a][1] * c1 + Gv[a][2] * c2;
        }
        __syncthreads();
        n += 2;
    }
    return n;
}

// One CTA = one member = one whole run of sim!.
template <int T, bool GLOBAL>
__global__ void __launch_bounds__(T, T == 128 ? 5 : 1)
k_ensemble(EnsArgs a) {
    extern __shared__ __align__(32) double ens_smem[];
    int member = blockIdx.x;
    if (a.redo_list) {
        if (member >= *a.redo_count) return;
        member = a.redo_list[member];
    } else if (member >= a.n_runs) {
        return;
    }
    const int tid = threadIdx.x;
    // where this member's results go: the global tier lays its outputs out by redo slot
    const size_t orow = GLOBAL ? (size_t)blockIdx.x : (size_t)member;
    const EnsSmem sm = GLOBAL ? ens_carve(a.gstate + (size_t)blockIdx.x * a.gstride, ens_smem, a.ncap, a.mmax)
                              : ens_carve(ens_smem, ens_smem + ens_state_doubles(a.ncap, a.mmax), a.ncap, a.mmax);
    const uint32_t run = a.run0 + (uint32_t)member;
    for (int idx = tid; idx < a.npk; idx += T) sm.P[idx] = 0.0;   // zeros(3,3); spare capacity zeroed
    for (int idx = tid; idx < a.ncap; idx += T) sm.x[idx] = 0.0;
#pragma unroll 1
    for (int e = tid; e < ens_ntri(a.ncap, a.mmax); e += T) {
        int ip = (int)((sqrtf(8.0f * (float)e + 1.0f) - 1.0f) * 0.5f);
        while (ip * (ip + 1) / 2 > e) ip--;
        while ((ip + 1) * (ip + 2) / 2 <= e) ip++;
        sm.tri[e] = (unsigned short)(ip | ((e - ip * (ip + 1) / 2) << 8));
    }
    __syncthreads();
    int n = 3;
    int status = SLAM_OK;
    const double *ctrl = a.ctrl_noise ? a.ctrl_noise + (size_t)member * 2 * a.nsteps : nullptr;
    const double *obsn = a.obs_noise ? a.obs_noise + (size_t)member * a.obs_per_member : nullptr;
    double *track = a.tracks ? a.tracks + (size_t)member * 3 * a.nsteps : nullptr;
    int done = 0, staged = -1, b0 = 0, rk = 0, over_idx = -1;
    bool outgrown = false, resume = false;
    double flops = 0.0;   // thread 0's tally (bench.py's FP64 roofline entry)
    const int ck_xo = CK_ZN + ens_ru4(2 * a.mmax);
    if (a.ckpt_in && (int)blockIdx.x < a.ckpt_in_slots) {
        // resume where the previous launch had to stop: state after that block's update, its new observations pending
        const double *c = a.ckpt_in + (size_t)blockIdx.x * a.ckpt_in_stride;
        n = (int)c[1]; rk = (int)c[2]; b0 = (int)c[3]; flops = c[4];
        const int rlen = (int)c[5], po = ck_xo + ens_ru4(a.ckpt_in_ncap);
        for (int idx = tid; idx < n; idx += T) sm.x[idx] = c[ck_xo + idx];
        for (int idx = tid; idx < ens_pdoubles(n); idx += T) sm.P[idx] = c[po + idx];
        for (int idx = tid; idx < 2 * rk; idx += T) sm.zn[idx] = c[CK_ZN + idx];
        if (tid < 3 * rlen) sm.bs[(tid / 3) * BS + B_X0 + tid % 3] = c[CK_POSE + tid];
        resume = true;
        done = a.blk[b0].x;
    } else if (tid < 3) {
        // state.x = vehicle.pose = initial_pose (sim/ekfslam-sim.jl:37,42); the host stores it in the
        // three doubles in front of z_true
        sm.x[tid] = a.z_true[-3 + tid];
    }
    __syncthreads();
    ENS_MARK(-1);
    int4 dnext = a.blk[b0];
    for (int b = b0; b < a.nblocks; b++) {
        ENS_MARK(0);
        const int4 d = dnext;      // the next block's descriptor is fetched a whole block ahead
        if (b + 1 < a.nblocks) dnext = a.blk[b + 1];
        const int s0 = d.x, s1 = d.y, len = d.y - d.x, nz = d.w;
        int k = rk;
        if (!resume) {
            if (obsn && 2 * d.z + 2 * nz > a.obs_per_member) { status = SLAM_ERR_CAPACITY; break; }
            if (staged != b) {   // not staged during the previous block's update (first block, blocks after
                                 // an observation step without matches): do it now
                if (tid >= 32) ens_stage_block<T>(sm, a, b, run, ctrl, obsn);
                __syncthreads();
            }
            ens_predict_block<T>(sm, n, len, nz);
            if (a.flops) flops += 30.0 * n * len;
            if (nz > 0) {
                ens_associate<T>(sm, n, nz);
                const int m = (int)sm.par[0];
                k = (int)sm.par[1];
                if (a.flops)
                    flops += 40.0 * nz * (n - 3) + (m > 0 ? (double)m * (2.0 * n * (n + 1) + 20.0 * n + 8.0 * n * m + 8.0 * m * m) : 0.0);
                if (m > 0) {
                    if (!ens_update<T>(sm, n, a.ldw, m, a, b + 1, run, ctrl, obsn)) { status = SLAM_ERR_NOT_PD; break; }
                    staged = b + 1;
                }
            }
        }
        if (nz > 0) {
            if (n + 2 * k > a.ncap) {
                status = SLAM_ERR_CAPACITY;
                outgrown = true;
                if (a.over_list) {
                    if (tid == 0) sm.par[3] = (double)atomicAdd(a.over_count, 1);
                    __syncthreads();
                    over_idx = (int)sm.par[3];
                    if (a.ckpt_out && over_idx < a.ckpt_out_slots) {
                        double *c = a.ckpt_out + (size_t)over_idx * a.ckpt_out_stride;
                        const int po = ck_xo + ens_ru4(a.ncap);
                        for (int idx = tid; idx < n; idx += T) c[ck_xo + idx] = sm.x[idx];
                        for (int idx = tid; idx < ens_pdoubles(n); idx += T) c[po + idx] = sm.P[idx];
                        for (int idx = tid; idx < 2 * k; idx += T) c[CK_ZN + idx] = sm.zn[idx];
                        if (tid < 3 * len) {
                            const int st = tid / 3, comp = tid - 3 * st;
                            c[CK_POSE + tid] = st == len - 1 ? sm.x[comp] : sm.bs[st * BS + B_X0 + comp];
                        }
                        if (tid == 0) { c[1] = n; c[2] = k; c[3] = b; c[4] = flops; c[5] = len; }
                    }
                }
                break;
            }
            n = ens_add_features<T>(sm, n, k);
            ENS_MARK(13);
        }
        resume = false;
        // slam_track[:, step] = state.x[1:3] (sim/ekfslam-sim.jl:128): after predict for the inner
        // steps, after update / add_features for the block's last step
        if (track && tid < 3 * len) {
            const int st = tid / 3, comp = tid - 3 * st;
            track[3 * (s0 + st) + comp] = st == len - 1 ? sm.x[comp] : sm.bs[st * BS + B_X0 + comp];
        }
        done = s1;
        __syncthreads();
        ENS_MARK(14);
    }
    __syncthreads();
    for (int idx = tid; idx < a.ocap; idx += T) a.final_x[orow * a.ocap + idx] = idx < a.ncap ? sm.x[idx] : 0.0;
    if (a.final_P)   // exported as the full symmetric matrix
        for (int idx = tid; idx < a.ocap * a.ocap; idx += T) {
            const int i = idx % a.ocap, j = idx / a.ocap;
            a.final_P[orow * a.ocap * a.ocap + idx] = (i < a.ncap && j < a.ncap) ? psym(sm.P, i, j) : 0.0;
        }
    ENS_MARK(-2);
    if (tid == 0) {
        if (outgrown && a.over_list) a.over_list[over_idx] = member;
        a.n_final[member] = n;
        a.status[member] = status;
        a.steps_done[member] = done;
        if (a.flops) a.flops[member] = flops;
        if (track)   // steps not run (stopped member) are marked NaN
            for (int s2 = done; s2 < a.nsteps; s2++) track[3 * s2] = track[3 * s2 + 1] = track[3 * s2 + 2] = nan("");
    }
}

// ---------------------------------------------------------------------------------------
// host API
// ---------------------------------------------------------------------------------------
// Capacity tiers.  The reference's state grows without bound (hcat / vcat, src/ekf.jl:108-109; new landmarks
// from src/data-association.jl:45-47); a member CTA has a fixed capacity, and the kernel's throughput follows
// the CTAs resident per SM, i.e. the member's shared memory.  A run is therefore a sequence of launches with
// growing capacity, each re-running -- from their first step; noise is keyed by run id, so the result is
// the single-launch result -- only the members whose map outgrew the previous one:
//   tier 0   first pass: the largest head-room that still gives the best occupancy (5 CTAs per SM for course1)
//   tier 1   `extra` landmarks of head-room (slam_ens_set_extra_landmarks, default 5)
//   tier 2   the largest head-room with two CTAs per SM          \ growth >= 1
//   tier 3   the largest head-room one CTA's shared memory holds /
//   global   x, P and the workspace in global memory, capacity doubled per round up to the bound
//            3 + 2*(landmarks + observations of the whole run) that no member can exceed (growth == 2)
// growth = 0 keeps the fixed capacity of tier 1: a member that outgrows it stops with SLAM_ERR_CAPACITY.
constexpr int ENS_MAX_LISTS = 6;

struct EnsTier {
    int extra = 0, ncap = 0, occ = 0;
    size_t smem = 0;
};
struct EnsGlobalRound {
    int ncap = 0;
    std::vector<int> members;
    double *d_x = nullptr, *d_P = nullptr;
};

struct slam_ensemble {
    int device = 0, n_runs = 0, nl = 0, ncap = 0, extra = 0, threads = 128;
    uint64_t seed = 0;
    uint32_t run0 = 0;
    TruthTables t;
    cudaStream_t stream = nullptr;
    double *d_gamma = nullptr, *d_ztrue = nullptr;   // d_ztrue[0..2] = initial pose, tables from +3
    int *d_blk = nullptr, *d_obs_start = nullptr, *d_steps_done = nullptr;
    double *d_ctrl = nullptr, *d_obs = nullptr;
    int obs_per_member = 0;
    double *d_final_x = nullptr, *d_tracks = nullptr, *d_final_P = nullptr;
    int *d_n_final = nullptr, *d_status = nullptr;
    double *d_flops = nullptr;     // per member, allocated by slam_ens_count_flops
    bool keep = false, ran = false;
    int first_extra = -1;          // first-pass head-room: -1 = chosen by occupancy, >= extra = no first pass
    int growth = 2;                // 0 fixed capacity, 1 + larger shared-memory tiers, 2 + global-memory tier
    int max_shared_extra = -1;     // cap on the shared tiers' head-room (-1: what the device holds)
    std::vector<EnsTier> tiers;
    bool dirty = true;             // tiers / output layout must be rebuilt before the next run
    int ocap = 0;                  // capacity d_final_x / d_final_P are laid out for (the largest shared tier)
    int *d_over = nullptr;         // ENS_MAX_LISTS lists of n_runs + 1 ints: [0] = count, [1 + i] = member
    double *d_ckpt[ENS_MAX_LISTS] = {nullptr};   // checkpoints written by tier i for tier i + 1 (or the first global round)
    int ckpt_slots[ENS_MAX_LISTS] = {0}, ckpt_stride[ENS_MAX_LISTS] = {0};
    bool global_pending = false;
    std::vector<EnsGlobalRound> grounds;
    std::vector<int> tier_members; // members each launch of the last run ran (shared tiers, then global rounds)
    long long launches = 0;
};

#define E_CHECK(e)                                                        \
    do {                                                                  \
        ARG_CHECK((e) != nullptr, "NULL handle");                         \
        CUDA_TRY(cudaSetDevice((e)->device));                             \
    } while (0)

static void ens_free_rounds(slam_ens_t *e) {
    for (auto &r : e->grounds) {
        cudaFree(r.d_x);
        cudaFree(r.d_P);
    }
    e->grounds.clear();
}

extern "C" int slam_ens_destroy(slam_ens_t *e) {
    if (!e) return SLAM_OK;
    cudaSetDevice(e->device);
    if (e->stream) cudaStreamSynchronize(e->stream);
    ens_free_rounds(e);
    for (auto &c : e->d_ckpt) cudaFree(c);
    cudaFree(e->d_gamma); cudaFree(e->d_ztrue); cudaFree(e->d_blk); cudaFree(e->d_steps_done); cudaFree(e->d_obs_start);
    cudaFree(e->d_ctrl); cudaFree(e->d_obs); cudaFree(e->d_final_x); cudaFree(e->d_tracks);
    cudaFree(e->d_final_P); cudaFree(e->d_n_final); cudaFree(e->d_status); cudaFree(e->d_over); cudaFree(e->d_flops);
    if (e->stream) cudaStreamDestroy(e->stream);
    cudaGetLastError();
    delete e;
    return SLAM_OK;
}

static int ens_mmax(const slam_ens_t *e) { return std::max(e->t.max_nz, 1); }

// the base capacity 3 + 2*(nl + extra): must fit one CTA's shared memory
static int ens_resize(slam_ens_t *e, int extra) {
    const int ncap = 3 + 2 * (e->nl + extra);
    const size_t smem = ens_smem_bytes(ncap, ens_mmax(e));
    int max_optin = 0;
    cudaDeviceGetAttribute(&max_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, e->device);
    if (smem > (size_t)max_optin) {
        slam_set_error("ensemble: a member needs %zu bytes of shared memory (%d + %d landmarks, up to %d "
                       "observations per step); the device offers %d", smem, e->nl, extra, e->t.max_nz, max_optin);
        return SLAM_ERR_CAPACITY;
    }
    e->extra = extra;
    e->ncap = ncap;
    e->dirty = true;
    e->ran = false;
    return SLAM_OK;
}

extern "C" int slam_ens_create(int device, const double *waypoints, int nwp, const double *landmarks, int nl,
                               int nlaps, uint64_t seed, uint32_t run0, int n_runs, slam_ens_t **out) {
    ARG_CHECK(out != nullptr, "out is NULL");
    *out = nullptr;
    ARG_CHECK(waypoints && nwp >= 2 && nl >= 0 && (nl == 0 || landmarks) && nlaps >= 1 && n_runs >= 1, "bad argument");
    int rc = slam_require_device(device);
    if (rc) return rc;
    slam_ens_t *e = new slam_ensemble();
    e->device = device; e->n_runs = n_runs; e->nl = nl; e->extra = ENS_DEFAULT_EXTRA; e->ncap = 3 + 2 * (nl + e->extra); e->seed = seed; e->run0 = run0;
    rc = build_truth(waypoints, nwp, landmarks, nl, nlaps, ENS_MAXB, e->t);
    if (rc) { delete e; return rc; }
    if (e->t.max_nz > 128) {
        slam_set_error("slam_ens_create: %d landmarks visible in one step; the ensemble kernel handles 128", e->t.max_nz);
        delete e;
        return SLAM_ERR_CAPACITY;
    }
#define EN_TRY(expr)                                                                             \
    do {                                                                                         \
        cudaError_t e__ = (expr);                                                                \
        if (e__ != cudaSuccess) {                                                                \
            slam_set_error("%s failed: %s", #expr, cudaGetErrorString(e__));                     \
            slam_ens_destroy(e);                                                                 \
            return SLAM_ERR_CUDA;                                                                \
        }                                                                                        \
    } while (0)
    const TruthTables &t = e->t;
    EN_TRY(cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking));
    EN_TRY(cudaMalloc(&e->d_gamma, sizeof(double) * t.nsteps));
    EN_TRY(cudaMalloc(&e->d_blk, sizeof(int) * t.blocks.size()));
    EN_TRY(cudaMalloc(&e->d_obs_start, sizeof(int) * (t.nupd + 1)));
    EN_TRY(cudaMalloc(&e->d_ztrue, sizeof(double) * (3 + t.z_true.size() + 2)));
    EN_TRY(cudaMalloc(&e->d_n_final, sizeof(int) * n_runs));
    EN_TRY(cudaMalloc(&e->d_steps_done, sizeof(int) * n_runs));
    EN_TRY(cudaMalloc(&e->d_status, sizeof(int) * n_runs));
    EN_TRY(cudaMalloc(&e->d_over, sizeof(int) * (size_t)ENS_MAX_LISTS * (n_runs + 1)));
    const double pose0[3] = {waypoints[0], waypoints[1],
                             atan2(waypoints[3] - waypoints[1], waypoints[2] - waypoints[0])};
    EN_TRY(cudaMemcpy(e->d_gamma, t.gamma.data(), sizeof(double) * t.nsteps, cudaMemcpyHostToDevice));
    EN_TRY(cudaMemcpy(e->d_blk, t.blocks.data(), sizeof(int) * t.blocks.size(), cudaMemcpyHostToDevice));
    EN_TRY(cudaMemcpy(e->d_obs_start, t.obs_start.data(), sizeof(int) * (t.nupd + 1), cudaMemcpyHostToDevice));
    EN_TRY(cudaMemcpy(e->d_ztrue, pose0, sizeof(pose0), cudaMemcpyHostToDevice));
    if (!t.z_true.empty())
        EN_TRY(cudaMemcpy(e->d_ztrue + 3, t.z_true.data(), sizeof(double) * t.z_true.size(), cudaMemcpyHostToDevice));
#undef EN_TRY
    rc = ens_resize(e, ENS_DEFAULT_EXTRA);
    if (rc) { slam_ens_destroy(e); return rc; }
    *out = e;
    return SLAM_OK;
}

extern "C" int slam_ens_set_extra_landmarks(slam_ens_t *e, int extra) {
    E_CHECK(e);
    ARG_CHECK(extra >= 0, "extra < 0");
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    return ens_resize(e, extra);
}

extern "C" int slam_ens_set_threads(slam_ens_t *e, int threads) {
    ARG_CHECK(e != nullptr, "NULL handle");
    ARG_CHECK(threads == 64 || threads == 128 || threads == 256, "threads must be 64, 128 or 256");
    e->threads = threads;
    e->dirty = true;
    return SLAM_OK;
}

extern "C" int slam_ens_set_growth(slam_ens_t *e, int mode) {
    ARG_CHECK(e != nullptr, "NULL handle");
    ARG_CHECK(mode >= 0 && mode <= 2, "mode must be 0, 1 or 2");
    e->growth = mode;
    e->dirty = true;
    return SLAM_OK;
}

extern "C" int slam_ens_set_max_shared_extra(slam_ens_t *e, int extra) {
    ARG_CHECK(e != nullptr, "NULL handle");
    ARG_CHECK(extra >= -1, "extra < -1");
    e->max_shared_extra = extra;
    e->dirty = true;
    return SLAM_OK;
}

extern "C" int slam_ens_capacity(slam_ens_t *e) {
    if (!e) { slam_set_error("slam_ens_capacity: NULL handle"); return SLAM_ERR_ARG; }
    return e->ncap;
}

extern "C" int slam_ens_steps(slam_ens_t *e, int *nsteps, int *nupdates) {
    ARG_CHECK(e != nullptr, "NULL handle");
    if (nsteps) *nsteps = e->t.nsteps;
    if (nupdates) *nupdates = e->t.nupd;
    return SLAM_OK;
}

extern "C" int slam_ens_keep_tracks(slam_ens_t *e, int on) {
    E_CHECK(e);
    e->keep = on != 0;
    if (e->keep && !e->d_tracks)
        CUDA_TRY(cudaMalloc(&e->d_tracks, sizeof(double) * (size_t)e->n_runs * 3 * e->t.nsteps));
    e->dirty = true;   // the covariance export is sized with the output layout
    return SLAM_OK;
}

// Algorithmic FP64 operations of the last run (SURVEY.md 8d's per-unit figures evaluated with every member's own
// n, nz and m at every step; a member re-run by a later launch counts once).  on = 1 before the run; *total after.
extern "C" int slam_ens_count_flops(slam_ens_t *e, int on, double *total) {
    E_CHECK(e);
    if (on && !e->d_flops) {
        CUDA_TRY(cudaMalloc(&e->d_flops, sizeof(double) * e->n_runs));
        CUDA_TRY(cudaMemset(e->d_flops, 0, sizeof(double) * e->n_runs));
    }
    if (total) {
        *total = 0;
        if (e->d_flops && e->ran) {
            CUDA_TRY(cudaStreamSynchronize(e->stream));
            std::vector<double> f(e->n_runs);
            CUDA_TRY(cudaMemcpy(f.data(), e->d_flops, sizeof(double) * e->n_runs, cudaMemcpyDeviceToHost));
            for (double v : f) *total += v;
        }
    }
    if (!on && e->d_flops) {
        CUDA_TRY(cudaStreamSynchronize(e->stream));
        cudaFree(e->d_flops);
        e->d_flops = nullptr;
    }
    return SLAM_OK;
}

extern "C" int slam_ens_set_noise(slam_ens_t *e, const double *ctrl, const double *obs, int obs_per_member) {
    E_CHECK(e);
    ARG_CHECK(ctrl && obs && obs_per_member >= 0, "bad argument");
    cudaFree(e->d_ctrl); cudaFree(e->d_obs);
    e->d_ctrl = e->d_obs = nullptr;
    const size_t nc = (size_t)e->n_runs * 2 * e->t.nsteps, no = (size_t)e->n_runs * obs_per_member;
    CUDA_TRY(cudaMalloc(&e->d_ctrl, sizeof(double) * nc));
    CUDA_TRY(cudaMalloc(&e->d_obs, sizeof(double) * std::max<size_t>(no, 1)));
    CUDA_TRY(cudaMemcpy(e->d_ctrl, ctrl, sizeof(double) * nc, cudaMemcpyHostToDevice));
    if (no) CUDA_TRY(cudaMemcpy(e->d_obs, obs, sizeof(double) * no, cudaMemcpyHostToDevice));
    e->obs_per_member = obs_per_member;
    return SLAM_OK;
}

template <int T>
static int ens_occupancy(size_t smem) {
    int nb = 0;
    if (cudaFuncSetAttribute(k_ensemble<T, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess ||
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k_ensemble<T, false>, T, smem) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return nb;
}
static int ens_occ(const slam_ens_t *e, size_t smem) {
    switch (e->threads) {
        case 64: return ens_occupancy<64>(smem);
        case 256: return ens_occupancy<256>(smem);
        default: return ens_occupancy<128>(smem);
    }
}
static EnsTier ens_tier(const slam_ens_t *e, int extra) {
    EnsTier t;
    t.extra = extra;
    t.ncap = 3 + 2 * (e->nl + extra);
    t.smem = ens_smem_bytes(t.ncap, ens_mmax(e));
    t.occ = ens_occ(e, t.smem);
    return t;
}

// First-pass head-room.  The kernel is latency-bound and its throughput follows the resident CTAs per
// SM, which the member's shared memory decides; few members ever need the full head-room.  Auto: the
// largest head-room that still gives the best occupancy, if that beats the base capacity's and the
// ensemble is longer than one wave (another pass costs about one member's run time).
static int ens_first_extra(const slam_ens_t *e) {
    if (e->first_extra >= 0) return std::min(e->first_extra, e->extra);
    const int full = ens_tier(e, e->extra).occ;
    int best = e->extra, best_occ = full;
    for (int x = e->extra - 1; x >= 0; x--) {
        const int occ = ens_tier(e, x).occ;
        if (occ > best_occ) { best = x; best_occ = occ; }
    }
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, e->device);
    if (best_occ == full || e->n_runs <= full * sms) return e->extra;
    return best;
}

// builds the tier list and (re)allocates what depends on the output layout
static int ens_prepare(slam_ens_t *e) {
    if (!e->dirty) return SLAM_OK;
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    int max_optin = 0;
    cudaDeviceGetAttribute(&max_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, e->device);
    e->tiers.clear();
    const int x1 = ens_first_extra(e);
    if (x1 < e->extra) e->tiers.push_back(ens_tier(e, x1));
    e->tiers.push_back(ens_tier(e, e->extra));
    const bool have_first = x1 < e->extra;
    if (e->growth >= 1) {
        const int limit = e->max_shared_extra >= 0 ? e->max_shared_extra : 1 << 20;
        int x2 = -1, x3 = -1;   // largest head-room with two CTAs per SM / that fits at all
        bool two = true;        // occupancy only falls as x grows: stop asking once it is below two
        for (int x = e->extra + 1; x <= limit; x++) {
            const size_t smem = ens_smem_bytes(3 + 2 * (e->nl + x), ens_mmax(e));
            if (smem > (size_t)max_optin || 3 + 2 * (e->nl + x) > 4080) break;
            x3 = x;
            if (two) {
                if (ens_occ(e, smem) >= 2) x2 = x;
                else two = false;
            }
        }
        // With a first pass in front and a two-CTA tier behind, the base tier is dropped: on course1 four in five of
        // the members that outgrow the first pass outgrow the base capacity too (diverged filters roughly double
        // their map), so that launch was a whole wave -- 5 ms of a 67 ms batch -- for 54 of 8192 members.
        if (x2 > e->extra && have_first) e->tiers.pop_back();
        if (x2 > e->extra) e->tiers.push_back(ens_tier(e, x2));
        if (x3 > e->extra && x3 > x2) e->tiers.push_back(ens_tier(e, x3));
    }
    for (const EnsTier &t : e->tiers)
        if (t.occ < 1) {
            slam_set_error("ensemble: no CTA of %zu bytes of shared memory fits an SM", t.smem);
            return SLAM_ERR_CAPACITY;
        }
    // the opt-in limit of the kernels is the largest tier's
    const int top = (int)e->tiers.back().smem;
    CUDA_TRY(cudaFuncSetAttribute(k_ensemble<64, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, top));
    CUDA_TRY(cudaFuncSetAttribute(k_ensemble<128, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, top));
    CUDA_TRY(cudaFuncSetAttribute(k_ensemble<256, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, top));
    CUDA_TRY(cudaFuncSetAttribute(k_ensemble<128, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)ens_fixed_bytes(ens_mmax(e))));
    // checkpoint slots of every tier that has a successor: an eighth of the members behind the first pass, a
    // sixty-fourth behind the others (course1: 3.6 % / 0.4 % outgrow them); members past the slots are re-run
    for (int i = 0; i < ENS_MAX_LISTS; i++) {
        cudaFree(e->d_ckpt[i]);
        e->d_ckpt[i] = nullptr;
        e->ckpt_slots[i] = e->ckpt_stride[i] = 0;
    }
    const char *ck_env = getenv("SLAM_ENS_CKPT");          // SLAM_ENS_CKPT=0: re-run outgrown members from their first step
    const bool ck_on = !(ck_env && atoi(ck_env) == 0);
    for (int i = 0; i < (int)e->tiers.size() && ck_on; i++) {
        if (!(i + 1 < (int)e->tiers.size() || e->growth == 2)) continue;
        const int slots = std::min(e->n_runs, std::max(64, e->n_runs / (i == 0 ? 8 : 64)));
        const int stride = ens_ckpt_doubles(e->tiers[i].ncap, ens_mmax(e));
        if (cudaMalloc(&e->d_ckpt[i], sizeof(double) * (size_t)slots * stride) != cudaSuccess) {
            cudaGetLastError();      // no room for checkpoints: those members are re-run from their first step
            e->d_ckpt[i] = nullptr;
            continue;
        }
        e->ckpt_slots[i] = slots;
        e->ckpt_stride[i] = stride;
    }
    e->ocap = e->tiers.back().ncap;
    cudaFree(e->d_final_x);
    cudaFree(e->d_final_P);
    e->d_final_x = e->d_final_P = nullptr;
    CUDA_TRY(cudaMalloc(&e->d_final_x, sizeof(double) * (size_t)e->n_runs * e->ocap));
    if (e->keep) CUDA_TRY(cudaMalloc(&e->d_final_P, sizeof(double) * (size_t)e->n_runs * e->ocap * e->ocap));
    e->dirty = false;
    e->ran = false;
    return SLAM_OK;
}

static void ens_launch(const slam_ens_t *e, const EnsArgs &a, int grid, size_t smem) {
    switch (e->threads) {
        case 64: k_ensemble<64, false><<<grid, 64, smem, e->stream>>>(a); break;
        case 256: k_ensemble<256, false><<<grid, 256, smem, e->stream>>>(a); break;
        default: k_ensemble<128, false><<<grid, 128, smem, e->stream>>>(a); break;
    }
}

static EnsArgs ens_args(const slam_ens_t *e) {
    EnsArgs a;
    a.n_runs = e->n_runs; a.nsteps = e->t.nsteps; a.nl = e->nl;
    a.mmax = ens_mmax(e);
    a.seed = e->seed; a.run0 = e->run0;
    a.nblocks = (int)e->t.blocks.size() / 4;
    a.gamma = e->d_gamma; a.blk = reinterpret_cast<const int4 *>(e->d_blk); a.z_true = e->d_ztrue + 3;
    a.ctrl_noise = e->d_ctrl; a.obs_noise = e->d_obs; a.obs_per_member = e->obs_per_member;
    a.ocap = e->ocap;
    a.final_x = e->d_final_x; a.n_final = e->d_n_final; a.status = e->d_status; a.steps_done = e->d_steps_done;
    a.tracks = e->keep ? e->d_tracks : nullptr;
    a.final_P = e->keep ? e->d_final_P : nullptr;
    a.over_list = a.over_count = nullptr;
    a.redo_list = a.redo_count = nullptr;
    a.gstate = nullptr;
    a.gstride = 0;
    a.ckpt_out = nullptr; a.ckpt_out_slots = a.ckpt_out_stride = 0;
    a.ckpt_in = nullptr; a.ckpt_in_slots = a.ckpt_in_stride = a.ckpt_in_ncap = 0;
    a.flops = e->d_flops;
    return a;
}
static void ens_set_capacity(EnsArgs &a, int ncap) {
    a.ncap = ncap;
    a.ldw = ens_ru4(ncap);
    a.npk = ens_pdoubles(ncap);
    a.wdoubles = ens_wdoubles(ncap, a.mmax);
}

extern "C" int slam_ens_run(slam_ens_t *e) {
    SLAM_NVTX();
    E_CHECK(e);
    int rc = ens_prepare(e);
    if (rc) return rc;
    ens_free_rounds(e);
    EnsArgs a = ens_args(e);
    const int nt = (int)e->tiers.size();
    const size_t lw = (size_t)e->n_runs + 1;
    CUDA_TRY(cudaMemsetAsync(e->d_over, 0, sizeof(int) * ENS_MAX_LISTS * lw, e->stream));
    for (int i = 0; i < nt; i++) {
        ens_set_capacity(a, e->tiers[i].ncap);
        int *prev = i ? e->d_over + (size_t)(i - 1) * lw : nullptr, *next = e->d_over + (size_t)i * lw;
        a.redo_count = prev; a.redo_list = prev ? prev + 1 : nullptr;
        const bool more = i + 1 < nt || e->growth == 2;    // someone will re-run what this launch cannot hold
        a.over_count = more ? next : nullptr; a.over_list = more ? next + 1 : nullptr;
        a.ckpt_out = more ? e->d_ckpt[i] : nullptr;
        a.ckpt_out_slots = e->ckpt_slots[i]; a.ckpt_out_stride = e->ckpt_stride[i];
        a.ckpt_in = i ? e->d_ckpt[i - 1] : nullptr;
        a.ckpt_in_slots = i ? e->ckpt_slots[i - 1] : 0; a.ckpt_in_stride = i ? e->ckpt_stride[i - 1] : 0;
        a.ckpt_in_ncap = i ? e->tiers[i - 1].ncap : 0;
        ens_launch(e, a, e->n_runs, e->tiers[i].smem);     // CTAs past the redo count exit at once
        e->launches++;
    }
    e->global_pending = true;   // the first sync reads the tiers' counts and, for growth == 2, runs the global rounds
    e->ran = true;
    CUDA_TRY(cudaGetLastError());
    return SLAM_OK;
}

// Last resort of growth == 2: the members the largest shared-memory tier could not hold run with x, P and the
// workspace in global memory, the capacity doubling per round up to the bound no run can exceed (every
// observation of the run a new landmark).  Synchronous; called by the first sync after a run.
static int ens_global_rounds(slam_ens_t *e) {
    e->global_pending = false;
    const int nt = (int)e->tiers.size();
    const size_t lw = (size_t)e->n_runs + 1;
    e->tier_members.assign(nt, 0);
    std::vector<int> cnts(ENS_MAX_LISTS);
    for (int i = 0; i < nt; i++) CUDA_TRY(cudaMemcpy(&cnts[i], e->d_over + (size_t)i * lw, sizeof(int), cudaMemcpyDeviceToHost));
    e->tier_members[0] = e->n_runs;
    for (int i = 1; i < nt; i++) e->tier_members[i] = cnts[i - 1];
    if (e->growth != 2) return SLAM_OK;
    int cur = nt - 1, spare = nt;   // ping-pong lists (at most four tiers of ENS_MAX_LISTS lists are in use)
    int count = cnts[cur], cap = e->tiers.back().ncap;
    const long long bound_ll = 3 + 2 * ((long long)e->nl + (long long)e->t.z_true.size() / 2);
    const int bound = (int)std::min<long long>(bound_ll, 4080 - 1);   // the triangle table holds 255 blocks of 16 rows
    const int mmax = ens_mmax(e);
    while (count > 0 && cap < bound) {
        const int ncap = (int)std::min<long long>(bound, 3 + 4 * (long long)((cap - 3) / 2));
        EnsGlobalRound r;
        r.ncap = ncap;
        r.members.resize(count);
        CUDA_TRY(cudaMemcpy(r.members.data(), e->d_over + (size_t)cur * lw + 1, sizeof(int) * count, cudaMemcpyDeviceToHost));
        const size_t gstride = ens_state_doubles(ncap, mmax);
        double *d_state = nullptr;
        if (cudaMalloc(&d_state, sizeof(double) * gstride * count) != cudaSuccess ||
            cudaMalloc(&r.d_x, sizeof(double) * (size_t)ncap * count) != cudaSuccess ||
            (e->keep && cudaMalloc(&r.d_P, sizeof(double) * (size_t)ncap * ncap * count) != cudaSuccess)) {
            cudaGetLastError();
            cudaFree(d_state); cudaFree(r.d_x); cudaFree(r.d_P);
            slam_set_error("ensemble: no device memory to re-run %d members with a capacity of %d states", count, ncap);
            return SLAM_ERR_CUDA;
        }
        EnsArgs a = ens_args(e);
        ens_set_capacity(a, ncap);
        a.ocap = ncap;
        a.final_x = r.d_x;
        a.final_P = e->keep ? r.d_P : nullptr;
        a.gstate = d_state;
        a.gstride = gstride;
        int *prev = e->d_over + (size_t)cur * lw, *next = e->d_over + (size_t)spare * lw;
        CUDA_TRY(cudaMemsetAsync(next, 0, sizeof(int), e->stream));
        a.redo_count = prev; a.redo_list = prev + 1;
        a.over_count = next; a.over_list = next + 1;
        if (e->grounds.empty() && e->d_ckpt[nt - 1]) {    // the first round resumes from the last shared tier's checkpoints
            a.ckpt_in = e->d_ckpt[nt - 1];
            a.ckpt_in_slots = e->ckpt_slots[nt - 1]; a.ckpt_in_stride = e->ckpt_stride[nt - 1];
            a.ckpt_in_ncap = e->tiers.back().ncap;
        }
        k_ensemble<128, true><<<count, 128, ens_fixed_bytes(mmax), e->stream>>>(a);
        e->launches++;
        CUDA_TRY(cudaGetLastError());
        CUDA_TRY(cudaStreamSynchronize(e->stream));
        cudaFree(d_state);
        e->tier_members.push_back(count);
        e->grounds.push_back(std::move(r));
        CUDA_TRY(cudaMemcpy(&count, next, sizeof(int), cudaMemcpyDeviceToHost));
        std::swap(cur, spare);
        cap = ncap;
    }
    return SLAM_OK;
}

// Head-room of the first pass: -1 = automatic, >= the base head-room = no separate first pass.
extern "C" int slam_ens_set_first_pass_extra(slam_ens_t *e, int extra) {
    ARG_CHECK(e != nullptr, "NULL handle");
    ARG_CHECK(extra >= -1, "extra < -1");
    e->first_extra = extra;
    e->dirty = true;
    return SLAM_OK;
}

static int ens_finish(slam_ens_t *e) {
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    if (e->ran && e->global_pending) return ens_global_rounds(e);
    return SLAM_OK;
}

// What the last slam_ens_run did: *first_extra = the first pass's head-room (-1: no separate first pass),
// *redone = members the pass after it ran again.
extern "C" int slam_ens_pass_info(slam_ens_t *e, int *first_extra, int *redone) {
    E_CHECK(e);
    int rc = ens_finish(e);
    if (rc) return rc;
    const bool two = e->ran && e->tiers.size() > 1 && e->tiers[0].extra < e->extra;
    if (first_extra) *first_extra = two ? e->tiers[0].extra : -1;
    if (redone) *redone = two && e->tier_members.size() > 1 ? e->tier_members[1] : 0;
    return SLAM_OK;
}

// Every launch of the last run, in order (shared-memory tiers, then global-memory rounds): its capacity in
// states and the members it ran.  Returns the number of launches; fills at most `max` entries.
extern "C" int slam_ens_tier_info(slam_ens_t *e, int *capacity, int *members, int max) {
    E_CHECK(e);
    int rc = ens_finish(e);
    if (rc) return rc;
    if (!e->ran) return 0;
    const int nt = (int)e->tiers.size(), total = (int)e->tier_members.size();
    for (int i = 0; i < total && i < max; i++) {
        if (capacity) capacity[i] = i < nt ? e->tiers[i].ncap : e->grounds[i - nt].ncap;
        if (members) members[i] = e->tier_members[i];
    }
    return total;
}

extern "C" int slam_ens_sync(slam_ens_t *e) {
    SLAM_NVTX();
    E_CHECK(e);
    int rc = ens_finish(e);
    if (rc) return rc;
    if (!e->ran) return 0;
    std::vector<int> st(e->n_runs);
    CUDA_TRY(cudaMemcpy(st.data(), e->d_status, sizeof(int) * e->n_runs, cudaMemcpyDeviceToHost));
    int bad = 0;
    for (int v : st) bad += v != 0;
    return bad;
}

// Filter-steps actually completed: *total over all members (members that stopped early count the
// steps they ran), per_member[n_runs] (may be NULL).
extern "C" int slam_ens_steps_done(slam_ens_t *e, long long *total, int *per_member) {
    E_CHECK(e);
    int rc = ens_finish(e);
    if (rc) return rc;
    std::vector<int> sd(e->n_runs, 0);
    if (e->ran) CUDA_TRY(cudaMemcpy(sd.data(), e->d_steps_done, sizeof(int) * e->n_runs, cudaMemcpyDeviceToHost));
    long long tot = 0;
    for (int v : sd) tot += v;
    if (total) *total = tot;
    if (per_member) memcpy(per_member, sd.data(), sizeof(int) * e->n_runs);
    return SLAM_OK;
}

// the global round (and the slot in it) that holds member's final state, or nullptr: the last round wins
static const EnsGlobalRound *ens_find_round(const slam_ens_t *e, int member, int *slot) {
    for (int r = (int)e->grounds.size() - 1; r >= 0; r--) {
        const auto &mv = e->grounds[r].members;
        for (size_t i = 0; i < mv.size(); i++)
            if (mv[i] == member) {
                *slot = (int)i;
                return &e->grounds[r];
            }
    }
    return nullptr;
}

extern "C" int slam_ens_get_results(slam_ens_t *e, double *final_pose, int *n_final, int *status) {
    E_CHECK(e);
    int rc = ens_finish(e);
    if (rc) return rc;
    if (final_pose) {
        ARG_CHECK(e->d_final_x != nullptr, "no run yet");
        CUDA_TRY(cudaMemcpy2D(final_pose, sizeof(double) * 3, e->d_final_x, sizeof(double) * e->ocap,
                              sizeof(double) * 3, e->n_runs, cudaMemcpyDeviceToHost));
        for (const auto &r : e->grounds)   // later rounds overwrite earlier ones
            for (size_t i = 0; i < r.members.size(); i++)
                CUDA_TRY(cudaMemcpy(final_pose + 3 * (size_t)r.members[i], r.d_x + i * (size_t)r.ncap, sizeof(double) * 3,
                                    cudaMemcpyDeviceToHost));
    }
    if (n_final) CUDA_TRY(cudaMemcpy(n_final, e->d_n_final, sizeof(int) * e->n_runs, cudaMemcpyDeviceToHost));
    if (status) CUDA_TRY(cudaMemcpy(status, e->d_status, sizeof(int) * e->n_runs, cudaMemcpyDeviceToHost));
    return SLAM_OK;
}

extern "C" int slam_ens_get_track(slam_ens_t *e, int member, double *track) {
    E_CHECK(e);
    ARG_CHECK(track && member >= 0 && member < e->n_runs, "bad argument");
    ARG_CHECK(e->keep && e->d_tracks, "tracks were not kept: call slam_ens_keep_tracks(e, 1) before the run");
    int rc = ens_finish(e);
    if (rc) return rc;
    CUDA_TRY(cudaMemcpy(track, e->d_tracks + (size_t)member * 3 * e->t.nsteps, sizeof(double) * 3 * e->t.nsteps,
                        cudaMemcpyDeviceToHost));
    return SLAM_OK;
}

// Final x (n_final) and P (n_final^2, ldp, full symmetric) of one member; returns n_final (x and P may both be
// NULL to ask for the size first: the state of a member is not bounded by slam_ens_capacity() when growth > 0).
extern "C" int slam_ens_get_state(slam_ens_t *e, int member, double *x, double *P, int ldp) {
    E_CHECK(e);
    ARG_CHECK(member >= 0 && member < e->n_runs, "bad member");
    int rc = ens_finish(e);
    if (rc) return rc;
    ARG_CHECK(e->ran, "no run yet");
    int n = 0;
    CUDA_TRY(cudaMemcpy(&n, e->d_n_final + member, sizeof(int), cudaMemcpyDeviceToHost));
    int slot = 0;
    const EnsGlobalRound *r = ens_find_round(e, member, &slot);
    const int cap = r ? r->ncap : e->ocap;
    const double *sx = r ? r->d_x + (size_t)slot * cap : e->d_final_x + (size_t)member * cap;
    const double *sP = r ? r->d_P : e->d_final_P;
    if (x) CUDA_TRY(cudaMemcpy(x, sx, sizeof(double) * n, cudaMemcpyDeviceToHost));
    if (P) {
        ARG_CHECK(e->keep && sP, "covariances were not kept: call slam_ens_keep_tracks(e, 1) before the run");
        ARG_CHECK(ldp >= n, "ldp < n");
        sP += (size_t)(r ? slot : member) * cap * cap;
        CUDA_TRY(cudaMemcpy2D(P, sizeof(double) * (size_t)ldp, sP, sizeof(double) * cap, sizeof(double) * n, n,
                              cudaMemcpyDeviceToHost));
    }
    return n;
}

extern "C" void *slam_ens_stream(slam_ens_t *e) { return e ? (void *)e->stream : nullptr; }
extern "C" long long slam_ens_launch_count(slam_ens_t *e) { return e ? e->launches : 0; }

#ifdef ENS_TIMING
extern "C" int slam_ens_debug_phases(long long *out, int reset) {
    if (out) cudaMemcpyFromSymbol(out, g_ens_phase, sizeof(long long) * 24);
    if (reset) {
        long long z[24] = {0};
        cudaMemcpyToSymbol(g_ens_phase, z, sizeof(z));
    }
    return SLAM_OK;
}
#endif
