// Ensemble mode: thousands of independent EKF-SLAM runs of the reference simulator loop
// (sim/ekfslam-sim.jl:80-142), one CTA per filter, the whole run in ONE kernel launch.
//
// The truth side of sim! does not depend on the noise (step_vehicle! uses the target speed and
// steering angle, src/common.jl:175-176), so the truth track, the steering angle and the set of
// visible landmarks per observation step are identical for every member that shares a map.
// They are computed once on the host (steer!, step_vehicle!, nearby_landmark_indices,
// get_observations; src/common.jl:172-230, sim/sim-utils.jl:12-28,53-66) and uploaded as tables.
// Each member then runs predict / associate / update / add_features for all steps with its
// state (x, full symmetric P, update workspace) resident in shared memory; only the noise
// (Philox4x32-10 keyed by (seed, run), or host-supplied arrays) and the optional track cross HBM.
//
// The device functions follow the operation order of the oracle's plain-C twin; compiled with
// -fmad=false, so the only differences from the twin are the device libm's sin/cos/atan2/log.
// Members are sharded by run across GPUs by the caller (one handle per GPU, no collective).
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
constexpr int MAX_STEPS = 10000;   // the reference's track buffers (sim/ekfslam-sim.jl:24-25)

double sgn(double a) { return (double)(a > 0) - (double)(a < 0); }

struct TruthTables {
    int nsteps = 0, nupd = 0, max_nz = 0;
    std::vector<double> gamma;       // target steering angle after steer!, per step
    std::vector<int> upd_at;         // per step: observation-step index or -1
    std::vector<int> obs_start;      // per observation step: offset into z_true (nupd + 1 entries)
    std::vector<double> z_true;      // 2 x total noise-free observations
    std::vector<double> true_track;  // 3 x nsteps
};

// Noise-free half of sim! (sleep, monitor and the filter calls removed).
int build_truth(const double *wp, int nwp, const double *lm, int nl, int nlaps, TruthTables &t) {
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
    return SLAM_OK;
}

}  // namespace

// ---------------------------------------------------------------------------------------
// device side
// ---------------------------------------------------------------------------------------
struct EnsArgs {
    int n_runs, nsteps, nl, ncap, mmax;      // ncap = 3 + 2*nl (the twin's fixed capacity), mmax = max nz
    uint64_t seed;
    uint32_t run0;
    const double *gamma;       // [nsteps]
    const int *upd_at;         // [nsteps]
    const int *obs_start;      // [nupd + 1]
    const double *z_true;      // [2 * total]
    const double *ctrl_noise;  // optional [n_runs][2*nsteps] (speed, steering per step)
    const double *obs_noise;   // optional [n_runs][obs_per_member]
    int obs_per_member;
    double *final_x;           // [n_runs][ncap]
    int *n_final;              // [n_runs]
    int *status;               // [n_runs]
    double *tracks;            // optional [n_runs][3*nsteps]
    double *final_P;           // optional [n_runs][ncap*ncap]
};

constexpr int ENS_THREADS = 128;

struct EnsSmem {   // carve-up of the dynamic shared memory of one member
    double *x, *P, *PHt, *S, *C, *u, *v, *Hn, *z, *zf, *zn, *nis, *nd, *par;
    int *fp, *idf, *cls;
};

__host__ __device__ inline size_t ens_smem_doubles(int ncap, int mmax, int nl) {
    const size_t R = 2 * (size_t)mmax;
    return (size_t)ncap + (size_t)ncap * ncap + (size_t)ncap * R + 2 * R * R + 2 * R + 10 * (size_t)mmax +
           6 * (size_t)mmax + 2 * (size_t)mmax * (nl > 0 ? nl : 1) + 32;
}
__host__ __device__ inline size_t ens_smem_bytes(int ncap, int mmax, int nl) {
    return ens_smem_doubles(ncap, mmax, nl) * 8 + 3 * (size_t)(mmax > 0 ? mmax : 1) * 4 + 64;
}

__device__ inline EnsSmem ens_carve(double *base, int ncap, int mmax, int nl) {
    EnsSmem s;
    const size_t R = 2 * (size_t)mmax, nlm = nl > 0 ? nl : 1;
    double *p = base;
    s.x = p;   p += ncap;
    s.P = p;   p += (size_t)ncap * ncap;
    s.PHt = p; p += (size_t)ncap * R;
    s.S = p;   p += R * R;
    s.C = p;   p += R * R;
    s.u = p;   p += R;
    s.v = p;   p += R;
    s.Hn = p;  p += 10 * (size_t)mmax;
    s.z = p;   p += 2 * (size_t)mmax;
    s.zf = p;  p += 2 * (size_t)mmax;
    s.zn = p;  p += 2 * (size_t)mmax;
    s.nis = p; p += (size_t)mmax * nlm;
    s.nd = p;  p += (size_t)mmax * nlm;
    s.par = p; p += 32;
    int *q = reinterpret_cast<int *>(p);
    s.fp = q;  q += mmax;
    s.idf = q; q += mmax;
    s.cls = q;
    return s;
}

#define EP(i, j) sm.P[(size_t)(i) + (size_t)(j) * ld]

// predict, src/ekf.jl:8-43, on the member's shared-memory state (both triangles kept, as the
// reference does with P[4:end,1:3] = P[1:3,4:end]')
__device__ void ens_predict(const EnsSmem &sm, int n, int ld, double g, double v) {
    const int tid = threadIdx.x;
    double *par = sm.par;
    if (tid == 0) {
        const double phi = sm.x[2];
        const double s = sin(g + phi), c = cos(g + phi);
        const double vts = v * DT * s, vtc = v * DT * c;
        par[0] = vts;
        par[1] = vtc;
        const double Q[4] = {SIGMA_SPEED * SIGMA_SPEED, 0, 0, SIGMA_STEER * SIGMA_STEER};
        const double Gv[3][3] = {{1, 0, -vts}, {0, 1, vtc}, {0, 0, 1}};
        const double Gu[3][2] = {{DT * c, -vts}, {DT * s, vtc}, {DT * sin(g) / WHEELBASE, v * DT * cos(g) / WHEELBASE}};
        double A[3][3], B[3][3], GQ[3][2];
        for (int i = 0; i < 3; i++)
            for (int j = 0; j < 3; j++) {
                double a = 0;
                for (int k = 0; k < 3; k++) a += Gv[i][k] * EP(k, j);
                A[i][j] = a;
            }
        for (int i = 0; i < 3; i++)
            for (int j = 0; j < 3; j++) {
                double a = 0;
                for (int k = 0; k < 3; k++) a += A[i][k] * Gv[j][k];
                B[i][j] = a;
            }
        for (int i = 0; i < 3; i++)
            for (int j = 0; j < 2; j++) GQ[i][j] = Gu[i][0] * Q[0 + 2 * j] + Gu[i][1] * Q[1 + 2 * j];
        for (int i = 0; i < 3; i++)
            for (int j = 0; j < 3; j++) B[i][j] += GQ[i][0] * Gu[j][0] + GQ[i][1] * Gu[j][1];
        for (int i = 0; i < 3; i++)
            for (int j = 0; j < 3; j++) EP(i, j) = B[i][j];
        double x0 = sm.x[0] + vtc, x1 = sm.x[1] + vts;
        sm.x[2] = mpi_to_pi(phi + v * DT * sin(g) / WHEELBASE);
        sm.x[0] = x0;
        sm.x[1] = x1;
    }
    __syncthreads();
    const double vts = par[0], vtc = par[1];
    for (int j = 3 + tid; j < n; j += ENS_THREADS) {
        double p0 = EP(0, j), p1 = EP(1, j), p2 = EP(2, j);
        double q0 = p0 + (-vts) * p2;
        double q1 = p1 + vtc * p2;
        EP(0, j) = q0; EP(1, j) = q1;
        EP(j, 0) = q0; EP(j, 1) = q1; EP(j, 2) = p2;
    }
    __syncthreads();
}

// 2x2 innovation statistics of (observation, landmark), src/data-association.jl:53-63, on the 5x5
// sub-block of the member's P (same operation order as twin_compute_association)
__device__ void ens_pair(const EnsSmem &sm, int ld, double z0, double z1, int j, double *nis, double *nd) {
    double zp[2], h[10];
    const int f = 3 + 2 * (j - 1);
    observe_model(sm.x[0], sm.x[1], sm.x[2], sm.x[f], sm.x[f + 1], zp, h);
    const int c[5] = {0, 1, 2, f, f + 1};
    const double Hr[2][5] = {{h[0], h[1], h[2], h[6], h[7]}, {h[3], h[4], h[5], h[8], h[9]}};
    double HP[2][5];
    for (int a = 0; a < 2; a++)
        for (int q = 0; q < 5; q++) {
            double acc = 0;
            for (int p = 0; p < 5; p++) acc += Hr[a][p] * EP(c[p], c[q]);
            HP[a][q] = acc;
        }
    const double R[4] = {SIGMA_R * SIGMA_R, 0, 0, SIGMA_B * SIGMA_B};
    double S[4];
    for (int a = 0; a < 2; a++)
        for (int b = 0; b < 2; b++) {
            double acc = 0;
            for (int q = 0; q < 5; q++) acc += HP[a][q] * Hr[b][q];
            S[2 * a + b] = acc + R[a + 2 * b];
        }
    double v0 = z0 - zp[0];
    double v1 = mpi_to_pi(z1 - zp[1]);
    double det = S[0] * S[3] - S[1] * S[2];
    double i11 = S[3] / det, i12 = -S[1] / det, i21 = -S[2] / det, i22 = S[0] / det;
    double t0 = i11 * v0 + i12 * v1;
    double t1 = i21 * v0 + i22 * v1;
    *nis = v0 * t0 + v1 * t1;
    *nd = *nis + log(det);
}

// associate, src/data-association.jl:1-51.  All (observation, landmark) pairs in parallel, then
// one thread per observation replays the reference's sequential scan over j literally.
// Returns m (matched) and k (new) through par[8], par[9].
__device__ void ens_associate(const EnsSmem &sm, int n, int ld, int nz, int nl) {
    const int tid = threadIdx.x;
    const int Nf = (n - 3) / 2;
    const int nlm = nl > 0 ? nl : 1;
    for (int idx = tid; idx < nz * Nf; idx += ENS_THREADS) {
        int i = idx / Nf, j = idx % Nf + 1;
        double nis, nd;
        ens_pair(sm, ld, sm.z[2 * i], sm.z[2 * i + 1], j, &nis, &nd);
        sm.nis[(size_t)i * nlm + (j - 1)] = nis;
        sm.nd[(size_t)i * nlm + (j - 1)] = nd;
    }
    __syncthreads();
    for (int i = tid; i < nz; i += ENS_THREADS) {
        int jbest = 0;
        double nbest = INFINITY, outer = INFINITY;
        for (int j = 1; j <= Nf; j++) {
            double nis = sm.nis[(size_t)i * nlm + (j - 1)], nd = sm.nd[(size_t)i * nlm + (j - 1)];
            bool ingate = false;
            if (nis < GATE1) {
                if (nd < nbest) ingate = true;
            }
            if (ingate) {
                nbest = nd;
                jbest = j;
            } else if (nis < outer) {
                outer = nis;
            }
        }
        sm.cls[i] = jbest != 0 ? jbest : (outer > GATE2 ? -1 : 0);
    }
    __syncthreads();
    if (tid == 0) {
        int m = 0, k = 0;
        for (int i = 0; i < nz; i++) {
            int c = sm.cls[i];
            if (c > 0) {
                sm.zf[2 * m] = sm.z[2 * i];
                sm.zf[2 * m + 1] = sm.z[2 * i + 1];
                sm.idf[m] = c;
                m++;
            } else if (c < 0) {
                sm.zn[2 * k] = sm.z[2 * i];
                sm.zn[2 * k + 1] = sm.z[2 * i + 1];
                k++;
            }
        }
        sm.par[8] = m;
        sm.par[9] = k;
    }
    __syncthreads();
}

// update, src/ekf.jl:46-77, in the operation order of twin_update.  The 2m x 2m innovation
// Cholesky and the triangular inverse run in ONE warp (lane = column, __syncwarp between the
// column steps) while the other warps wait at a single barrier.  Returns false if S is not PD.
__device__ bool ens_update(const EnsSmem &sm, int n, int ld, int m) {
    const int tid = threadIdx.x;
    const int R = 2 * m;
    const double Rm[4] = {SIGMA_R * SIGMA_R, 0, 0, SIGMA_B * SIGMA_B};
    if (tid < m) {
        double zp[2];
        const int f = 3 + 2 * (sm.idf[tid] - 1);
        observe_model(sm.x[0], sm.x[1], sm.x[2], sm.x[f], sm.x[f + 1], zp, sm.Hn + 10 * tid);
        sm.fp[tid] = f;
        sm.v[2 * tid] = sm.zf[2 * tid] - zp[0];
        sm.v[2 * tid + 1] = mpi_to_pi(sm.zf[2 * tid + 1] - zp[1]);
    }
    __syncthreads();
    // PHt = P*H' (five terms per entry, columns 1,2,3,f,f+1)
    for (int idx = tid; idx < n * m; idx += ENS_THREADS) {
        const int i = idx % n, o = idx / n;
        const double *h = sm.Hn + 10 * o;
        const int f = sm.fp[o];
        const double p0 = EP(i, 0), p1 = EP(i, 1), p2 = EP(i, 2), p3 = EP(i, f), p4 = EP(i, f + 1);
        sm.PHt[(size_t)i * R + 2 * o] = p0 * h[0] + p1 * h[1] + p2 * h[2] + p3 * h[6] + p4 * h[7];
        sm.PHt[(size_t)i * R + 2 * o + 1] = p0 * h[3] + p1 * h[4] + p2 * h[5] + p3 * h[8] + p4 * h[9];
    }
    __syncthreads();
    // S = H*PHt + RR
    for (int idx = tid; idx < m * R; idx += ENS_THREADS) {
        const int o = idx / R, k = idx % R;
        const double *h = sm.Hn + 10 * o;
        const int f = sm.fp[o];
        const double a0 = sm.PHt[k], a1 = sm.PHt[(size_t)R + k], a2 = sm.PHt[(size_t)2 * R + k];
        const double a3 = sm.PHt[(size_t)f * R + k], a4 = sm.PHt[(size_t)(f + 1) * R + k];
        double s0 = h[0] * a0 + h[1] * a1 + h[2] * a2 + h[6] * a3 + h[7] * a4;
        double s1 = h[3] * a0 + h[4] * a1 + h[5] * a2 + h[8] * a3 + h[9] * a4;
        if (k == 2 * o) { s0 += Rm[0]; s1 += Rm[1]; }
        if (k == 2 * o + 1) { s0 += Rm[2]; s1 += Rm[3]; }
        sm.S[(size_t)(2 * o) + (size_t)k * R] = s0;
        sm.S[(size_t)(2 * o + 1) + (size_t)k * R] = s1;
    }
    __syncthreads();
    for (int idx = tid; idx < R * R; idx += ENS_THREADS) {   // S = (S + S')*0.5
        const int a = idx % R, b = idx / R;
        if (a < b) {
            double s = (sm.S[(size_t)a + (size_t)b * R] + sm.S[(size_t)b + (size_t)a * R]) * 0.5;
            sm.S[(size_t)a + (size_t)b * R] = s;
            sm.S[(size_t)b + (size_t)a * R] = s;
        }
    }
    __syncthreads();
    if (tid < 32) {   // one warp: right-looking upper Cholesky, then C = inv(U) and u = C'v
        const int lane = tid;
        int fail = 0;
        for (int k = 0; k < R; k++) {
            double d = sm.S[(size_t)k + (size_t)k * R];
            if (!(d > 0)) { fail = 1; break; }    // uniform: every lane reads the same d
            double ukk = sqrt(d);
            __syncwarp();
            if (lane == 0) sm.S[(size_t)k + (size_t)k * R] = ukk;
            for (int j = k + 1 + lane; j < R; j += 32) sm.S[(size_t)k + (size_t)j * R] /= ukk;
            __syncwarp();
            for (int j = k + 1 + lane; j < R; j += 32) {
                const double ukj = sm.S[(size_t)k + (size_t)j * R];
                for (int i = k + 1; i <= j; i++) sm.S[(size_t)i + (size_t)j * R] -= sm.S[(size_t)k + (size_t)i * R] * ukj;
            }
            __syncwarp();
        }
        if (lane == 0) sm.par[10] = fail;
        if (!fail) {
            for (int j = lane; j < R; j += 32) {
                sm.C[(size_t)j + (size_t)j * R] = 1.0 / sm.S[(size_t)j + (size_t)j * R];
                for (int i = j - 1; i >= 0; i--) {
                    double a = 0;
                    for (int k = i + 1; k <= j; k++) a += sm.S[(size_t)i + (size_t)k * R] * sm.C[(size_t)k + (size_t)j * R];
                    sm.C[(size_t)i + (size_t)j * R] = -a / sm.S[(size_t)i + (size_t)i * R];
                }
                double a = 0;
                for (int l = 0; l <= j; l++) a += sm.C[(size_t)l + (size_t)j * R] * sm.v[l];
                sm.u[j] = a;
            }
        }
    }
    __syncthreads();
    if (sm.par[10] != 0.0) return false;
    // W1 = PHt*C in place (k descending keeps PHt[i, l <= k] intact), x += W1*u
    for (int i = tid; i < n; i += ENS_THREADS) {
        double *row = sm.PHt + (size_t)i * R;
        for (int k = R - 1; k >= 0; k--) {
            double a = 0;
            for (int l = 0; l <= k; l++) a += row[l] * sm.C[(size_t)l + (size_t)k * R];
            row[k] = a;
        }
        double dxi = 0;
        for (int k = 0; k < R; k++) dxi += row[k] * sm.u[k];
        sm.x[i] += dxi;
    }
    __syncthreads();
    // P -= W1*W1' over the full matrix (the reference subtracts the full symmetric product)
    for (int idx = tid; idx < n * n; idx += ENS_THREADS) {
        const int i = idx % n, j = idx / n;
        const double *wi = sm.PHt + (size_t)i * R, *wj = sm.PHt + (size_t)j * R;
        double a = EP(i, j);
        for (int k = 0; k < R; k++) a = fma(-wi[k], wj[k], a);
        EP(i, j) = a;
    }
    __syncthreads();
    return true;
}

// add_features, src/ekf.jl:84-122, in place (the member's capacity is fixed); returns the new n
__device__ int ens_add_features(const EnsSmem &sm, int n, int ld, int k) {
    const int tid = threadIdx.x;
    const double phi = sm.x[2];
    const double Rm[4] = {SIGMA_R * SIGMA_R, 0, 0, SIGMA_B * SIGMA_B};
    for (int q = 0; q < k; q++) {
        const int len = n;
        const double r = sm.zn[2 * q], b = sm.zn[2 * q + 1];
        const double s = sin(phi + b), c = cos(phi + b);
        const double Gv[2][3] = {{1, 0, -r * s}, {0, 1, r * c}};
        double *GP = sm.par + 16;   // 2 x 3
        if (tid == 0) {
            sm.x[len] = sm.x[0] + r * c;
            sm.x[len + 1] = sm.x[1] + r * s;
            const double Gz[2][2] = {{c, -r * s}, {s, r * c}};
            double GR[2][2];
            for (int a = 0; a < 2; a++)
                for (int j = 0; j < 3; j++)
                    GP[a * 3 + j] = Gv[a][0] * EP(0, j) + Gv[a][1] * EP(1, j) + Gv[a][2] * EP(2, j);
            for (int a = 0; a < 2; a++)
                for (int j = 0; j < 2; j++) GR[a][j] = Gz[a][0] * Rm[0 + 2 * j] + Gz[a][1] * Rm[1 + 2 * j];
            for (int a = 0; a < 2; a++)
                for (int bb = 0; bb < 2; bb++) {
                    double t1 = GP[a * 3 + 0] * Gv[bb][0] + GP[a * 3 + 1] * Gv[bb][1] + GP[a * 3 + 2] * Gv[bb][2];
                    double t2 = GR[a][0] * Gz[bb][0] + GR[a][1] * Gz[bb][1];
                    EP(len + a, len + bb) = t1 + t2;
                }
            for (int a = 0; a < 2; a++)
                for (int j = 0; j < 3; j++) {
                    EP(len + a, j) = GP[a * 3 + j];
                    EP(j, len + a) = GP[a * 3 + j];
                }
        }
        for (int j = 3 + tid; j < len; j += ENS_THREADS)
            for (int a = 0; a < 2; a++) {
                double t = Gv[a][0] * EP(0, j) + Gv[a][1] * EP(1, j) + Gv[a][2] * EP(2, j);
                EP(len + a, j) = t;
                EP(j, len + a) = t;
            }
        __syncthreads();
        n += 2;
    }
    return n;
}

// One CTA = one member = one whole run of sim!.
__global__ void __launch_bounds__(ENS_THREADS)
k_ensemble(EnsArgs a) {
    extern __shared__ __align__(16) double ens_smem[];
    const int member = blockIdx.x;
    if (member >= a.n_runs) return;
    const int tid = threadIdx.x;
    const int ld = a.ncap;
    const EnsSmem sm = ens_carve(ens_smem, a.ncap, a.mmax, a.nl);
    const uint32_t run = a.run0 + (uint32_t)member;
    for (int idx = tid; idx < a.ncap * a.ncap; idx += ENS_THREADS) sm.P[idx] = 0.0;   // zeros(3,3), capacity zeroed
    for (int idx = tid; idx < a.ncap; idx += ENS_THREADS) sm.x[idx] = 0.0;
    __syncthreads();
    int n = 3;
    int status = SLAM_OK;
    const double *ctrl = a.ctrl_noise ? a.ctrl_noise + (size_t)member * 2 * a.nsteps : nullptr;
    const double *obsn = a.obs_noise ? a.obs_noise + (size_t)member * a.obs_per_member : nullptr;
    double *track = a.tracks ? a.tracks + (size_t)member * 3 * a.nsteps : nullptr;
    // state.x = vehicle.pose = initial_pose (sim/ekfslam-sim.jl:37,42); the host stores it in the
    // three doubles in front of z_true
    if (tid < 3) sm.x[tid] = a.z_true[-3 + tid];
    __syncthreads();
    int step = 0;
    for (; step < a.nsteps; step++) {
        // add_control_noise!, sim/sim-utils.jl:35-38 (speed first, then steering)
        const double n0 = ctrl ? ctrl[2 * step] : philox_normal(a.seed, run, 0u, (uint32_t)step);
        const double n1 = ctrl ? ctrl[2 * step + 1] : philox_normal(a.seed, run, 1u, (uint32_t)step);
        const double mspeed = TARGET_SPEED + n0 * sqrt(SIGMA_SPEED * SIGMA_SPEED);
        const double mgamma = a.gamma[step] + n1 * sqrt(SIGMA_STEER * SIGMA_STEER);
        ens_predict(sm, n, ld, mgamma, mspeed);
        const int u = a.upd_at[step];
        if (u >= 0) {
            const int o0 = a.obs_start[u], nz = a.obs_start[u + 1] - o0;
            const int cursor = 2 * o0;   // draw order: all range noises of the step, then all bearing noises
            if (obsn && cursor + 2 * nz > a.obs_per_member) { status = SLAM_ERR_CAPACITY; break; }
            for (int c = tid; c < nz; c += ENS_THREADS) {
                const double e0 = obsn ? obsn[cursor + c] : philox_normal(a.seed, run, 2u, (uint32_t)(cursor + c));
                const double e1 = obsn ? obsn[cursor + nz + c] : philox_normal(a.seed, run, 2u, (uint32_t)(cursor + nz + c));
                sm.z[2 * c] = a.z_true[2 * (o0 + c)] + e0 * sqrt(SIGMA_R * SIGMA_R);
                sm.z[2 * c + 1] = a.z_true[2 * (o0 + c) + 1] + e1 * sqrt(SIGMA_B * SIGMA_B);
            }
            __syncthreads();
            ens_associate(sm, n, ld, nz, a.nl);
            const int m = (int)sm.par[8], k = (int)sm.par[9];
            if (m > 0 && !ens_update(sm, n, ld, m)) { status = SLAM_ERR_NOT_PD; break; }
            if (n + 2 * k > a.ncap) { status = SLAM_ERR_CAPACITY; break; }   // as the twin: rc = -2
            n = ens_add_features(sm, n, ld, k);
        }
        if (track && tid < 3) track[3 * step + tid] = sm.x[tid];
    }
    __syncthreads();
    for (int idx = tid; idx < a.ncap; idx += ENS_THREADS) a.final_x[(size_t)member * a.ncap + idx] = sm.x[idx];
    if (a.final_P)
        for (int idx = tid; idx < a.ncap * a.ncap; idx += ENS_THREADS)
            a.final_P[(size_t)member * a.ncap * a.ncap + idx] = sm.P[idx];
    if (tid == 0) {
        a.n_final[member] = n;
        a.status[member] = status;
        if (track)   // steps not run (failed member) are marked NaN
            for (int s2 = step; s2 < a.nsteps; s2++) track[3 * s2] = track[3 * s2 + 1] = track[3 * s2 + 2] = nan("");
    }
}

// ---------------------------------------------------------------------------------------
// host API
// ---------------------------------------------------------------------------------------
struct slam_ensemble {
    int device = 0, n_runs = 0, nl = 0, ncap = 0;
    uint64_t seed = 0;
    uint32_t run0 = 0;
    TruthTables t;
    cudaStream_t stream = nullptr;
    double *d_gamma = nullptr, *d_ztrue = nullptr;   // d_ztrue[0..2] = initial pose, tables from +3
    int *d_upd_at = nullptr, *d_obs_start = nullptr;
    double *d_ctrl = nullptr, *d_obs = nullptr;
    int obs_per_member = 0;
    double *d_final_x = nullptr, *d_tracks = nullptr, *d_final_P = nullptr;
    int *d_n_final = nullptr, *d_status = nullptr;
    bool keep = false, ran = false;
    size_t smem = 0;
    long long launches = 0;
};

#define E_CHECK(e)                                                        \
    do {                                                                  \
        ARG_CHECK((e) != nullptr, "NULL handle");                         \
        CUDA_TRY(cudaSetDevice((e)->device));                             \
    } while (0)

extern "C" int slam_ens_destroy(slam_ens_t *e) {
    if (!e) return SLAM_OK;
    cudaSetDevice(e->device);
    if (e->stream) cudaStreamSynchronize(e->stream);
    cudaFree(e->d_gamma); cudaFree(e->d_ztrue); cudaFree(e->d_upd_at); cudaFree(e->d_obs_start);
    cudaFree(e->d_ctrl); cudaFree(e->d_obs); cudaFree(e->d_final_x); cudaFree(e->d_tracks);
    cudaFree(e->d_final_P); cudaFree(e->d_n_final); cudaFree(e->d_status);
    if (e->stream) cudaStreamDestroy(e->stream);
    cudaGetLastError();
    delete e;
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
    e->device = device; e->n_runs = n_runs; e->nl = nl; e->ncap = 3 + 2 * nl; e->seed = seed; e->run0 = run0;
    rc = build_truth(waypoints, nwp, landmarks, nl, nlaps, e->t);
    if (rc) { delete e; return rc; }
    const int mmax = std::max(e->t.max_nz, 1);
    e->smem = ens_smem_bytes(e->ncap, mmax, nl);
    int max_optin = 0;
    cudaDeviceGetAttribute(&max_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, device);
    if (e->smem > (size_t)max_optin) {
        slam_set_error("slam_ens_create: a member needs %zu bytes of shared memory (%d landmarks, up to %d "
                       "observations per step); the device offers %d", e->smem, nl, e->t.max_nz, max_optin);
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
    EN_TRY(cudaFuncSetAttribute(k_ensemble, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)e->smem));
    EN_TRY(cudaMalloc(&e->d_gamma, sizeof(double) * t.nsteps));
    EN_TRY(cudaMalloc(&e->d_upd_at, sizeof(int) * t.nsteps));
    EN_TRY(cudaMalloc(&e->d_obs_start, sizeof(int) * (t.nupd + 1)));
    EN_TRY(cudaMalloc(&e->d_ztrue, sizeof(double) * (3 + t.z_true.size() + 2)));
    EN_TRY(cudaMalloc(&e->d_final_x, sizeof(double) * (size_t)n_runs * e->ncap));
    EN_TRY(cudaMalloc(&e->d_n_final, sizeof(int) * n_runs));
    EN_TRY(cudaMalloc(&e->d_status, sizeof(int) * n_runs));
    const double pose0[3] = {waypoints[0], waypoints[1],
                             atan2(waypoints[3] - waypoints[1], waypoints[2] - waypoints[0])};
    EN_TRY(cudaMemcpy(e->d_gamma, t.gamma.data(), sizeof(double) * t.nsteps, cudaMemcpyHostToDevice));
    EN_TRY(cudaMemcpy(e->d_upd_at, t.upd_at.data(), sizeof(int) * t.nsteps, cudaMemcpyHostToDevice));
    EN_TRY(cudaMemcpy(e->d_obs_start, t.obs_start.data(), sizeof(int) * (t.nupd + 1), cudaMemcpyHostToDevice));
    EN_TRY(cudaMemcpy(e->d_ztrue, pose0, sizeof(pose0), cudaMemcpyHostToDevice));
    if (!t.z_true.empty())
        EN_TRY(cudaMemcpy(e->d_ztrue + 3, t.z_true.data(), sizeof(double) * t.z_true.size(), cudaMemcpyHostToDevice));
#undef EN_TRY
    *out = e;
    return SLAM_OK;
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
    if (e->keep && !e->d_tracks) {
        CUDA_TRY(cudaMalloc(&e->d_tracks, sizeof(double) * (size_t)e->n_runs * 3 * e->t.nsteps));
        CUDA_TRY(cudaMalloc(&e->d_final_P, sizeof(double) * (size_t)e->n_runs * e->ncap * e->ncap));
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

extern "C" int slam_ens_run(slam_ens_t *e) {
    E_CHECK(e);
    EnsArgs a;
    a.n_runs = e->n_runs; a.nsteps = e->t.nsteps; a.nl = e->nl; a.ncap = e->ncap;
    a.mmax = std::max(e->t.max_nz, 1);
    a.seed = e->seed; a.run0 = e->run0;
    a.gamma = e->d_gamma; a.upd_at = e->d_upd_at; a.obs_start = e->d_obs_start; a.z_true = e->d_ztrue + 3;
    a.ctrl_noise = e->d_ctrl; a.obs_noise = e->d_obs; a.obs_per_member = e->obs_per_member;
    a.final_x = e->d_final_x; a.n_final = e->d_n_final; a.status = e->d_status;
    a.tracks = e->keep ? e->d_tracks : nullptr;
    a.final_P = e->keep ? e->d_final_P : nullptr;
    k_ensemble<<<e->n_runs, ENS_THREADS, e->smem, e->stream>>>(a);
    e->launches++;
    e->ran = true;
    CUDA_TRY(cudaGetLastError());
    return SLAM_OK;
}

extern "C" int slam_ens_sync(slam_ens_t *e) {
    E_CHECK(e);
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    if (!e->ran) return 0;
    std::vector<int> st(e->n_runs);
    CUDA_TRY(cudaMemcpy(st.data(), e->d_status, sizeof(int) * e->n_runs, cudaMemcpyDeviceToHost));
    int bad = 0;
    for (int v : st) bad += v != 0;
    return bad;
}

extern "C" int slam_ens_get_results(slam_ens_t *e, double *final_pose, int *n_final, int *status) {
    E_CHECK(e);
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    if (final_pose)
        CUDA_TRY(cudaMemcpy2D(final_pose, sizeof(double) * 3, e->d_final_x, sizeof(double) * e->ncap,
                              sizeof(double) * 3, e->n_runs, cudaMemcpyDeviceToHost));
    if (n_final) CUDA_TRY(cudaMemcpy(n_final, e->d_n_final, sizeof(int) * e->n_runs, cudaMemcpyDeviceToHost));
    if (status) CUDA_TRY(cudaMemcpy(status, e->d_status, sizeof(int) * e->n_runs, cudaMemcpyDeviceToHost));
    return SLAM_OK;
}

extern "C" int slam_ens_get_track(slam_ens_t *e, int member, double *track) {
    E_CHECK(e);
    ARG_CHECK(track && member >= 0 && member < e->n_runs, "bad argument");
    ARG_CHECK(e->keep && e->d_tracks, "tracks were not kept: call slam_ens_keep_tracks(e, 1) before the run");
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    CUDA_TRY(cudaMemcpy(track, e->d_tracks + (size_t)member * 3 * e->t.nsteps, sizeof(double) * 3 * e->t.nsteps,
                        cudaMemcpyDeviceToHost));
    return SLAM_OK;
}

extern "C" int slam_ens_get_state(slam_ens_t *e, int member, double *x, double *P, int ldp) {
    E_CHECK(e);
    ARG_CHECK(member >= 0 && member < e->n_runs, "bad member");
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    int n = 0;
    CUDA_TRY(cudaMemcpy(&n, e->d_n_final + member, sizeof(int), cudaMemcpyDeviceToHost));
    if (x) CUDA_TRY(cudaMemcpy(x, e->d_final_x + (size_t)member * e->ncap, sizeof(double) * n, cudaMemcpyDeviceToHost));
    if (P) {
        ARG_CHECK(e->keep && e->d_final_P, "covariances were not kept: call slam_ens_keep_tracks(e, 1) before the run");
        ARG_CHECK(ldp >= n, "ldp < n");
        CUDA_TRY(cudaMemcpy2D(P, sizeof(double) * (size_t)ldp, e->d_final_P + (size_t)member * e->ncap * e->ncap,
                              sizeof(double) * e->ncap, sizeof(double) * n, n, cudaMemcpyDeviceToHost));
    }
    return n;
}

extern "C" void *slam_ens_stream(slam_ens_t *e) { return e ? (void *)e->stream : nullptr; }
extern "C" long long slam_ens_launch_count(slam_ens_t *e) { return e ? e->launches : 0; }
