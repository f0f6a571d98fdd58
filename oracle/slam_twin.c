/* ORACLE (test infrastructure, not product code) -- plain-C "sparse" restatement of the
 * EKF-SLAM hot path of andrewadare/SLAM.jl, independent of oracle/slam_ref.py.
 *
 * PARITY UNPINNED: the reference has no tests or golden vectors and Julia is absent here;
 * this file is pinned only against the numpy literal restatement (tests/test_oracle.py).
 *
 * Same math as the reference but only on the entries that the sparse Jacobian touches
 * (rows/cols {1,2,3,fpos,fpos+1}), closed-form 2x2 inverse/determinant, no BLAS.
 * Build: gcc -O2 -ffp-contract=off -fopenmp -shared -fPIC (oracle/Makefile).
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load it.  Citations are file:line under /root/reference.
 *
 * Matrices are column-major with leading dimension ld; idf is 1-based (as in Julia).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define P_(i, j) P[(size_t)(i) + (size_t)(j) * (size_t)ld]

/* src/common.jl:102-110 */
double twin_mpi_to_pi(double phi) {
    if (phi > M_PI) return phi - 2 * M_PI;
    if (phi < -M_PI) return phi + 2 * M_PI;
    return phi;
}

/* src/common.jl:139-165.  zp[2]; h[10] = nonzeros of the 2 x n Jacobian:
 * h[0..2] = row 1 cols 1:3, h[3..5] = row 2 cols 1:3, h[6..7] = row 1 cols fpos:fpos+1,
 * h[8..9] = row 2 cols fpos:fpos+1.  Returns the 0-based fpos. */
int twin_predict_observation(const double *x, int64_t idf, double *zp, double *h) {
    int f = 3 + 2 * ((int)idf - 1);
    double dx = x[f] - x[0];
    double dy = x[f + 1] - x[1];
    double d2 = dx * dx + dy * dy;
    double d = sqrt(d2);
    zp[0] = d;
    zp[1] = atan2(dy, dx) - x[2];
    double xd = dx / d, yd = dy / d, xd2 = dx / d2, yd2 = dy / d2;
    h[0] = -xd; h[1] = -yd; h[2] = 0.0;
    h[3] = yd2; h[4] = -xd2; h[5] = -1.0;
    h[6] = xd;  h[7] = yd;
    h[8] = -yd2; h[9] = xd2;
    return f;
}

/* src/ekf.jl:8-43 -- in place on x and the full (both triangles) P. */
void twin_predict(double *x, double *P, int n, int ld, double g, double v, double w,
                  const double *Q /* 2x2 col-major */, double dt) {
    double phi = x[2];
    double s = sin(g + phi), c = cos(g + phi);
    double vts = v * dt * s, vtc = v * dt * c;
    double Gv[3][3] = {{1, 0, -vts}, {0, 1, vtc}, {0, 0, 1}};
    double Gu[3][2] = {{dt * c, -vts}, {dt * s, vtc}, {dt * sin(g) / w, v * dt * cos(g) / w}};
    double A[3][3], B[3][3], GQ[3][2];
    /* (Gv*Pvv)*Gv' */
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
            double a = 0;
            for (int k = 0; k < 3; k++) a += Gv[i][k] * P_(k, j);
            A[i][j] = a;
        }
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
            double a = 0;
            for (int k = 0; k < 3; k++) a += A[i][k] * Gv[j][k];
            B[i][j] = a;
        }
    /* (Gu*Q)*Gu' */
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 2; j++) GQ[i][j] = Gu[i][0] * Q[0 + 2 * j] + Gu[i][1] * Q[1 + 2 * j];
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) B[i][j] += GQ[i][0] * Gu[j][0] + GQ[i][1] * Gu[j][1];
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) P_(i, j) = B[i][j];
    for (int j = 3; j < n; j++) {
        double p0 = P_(0, j), p1 = P_(1, j), p2 = P_(2, j);
        double q0 = p0 + Gv[0][2] * p2;   /* 1*p0 + 0*p1 + (-vts)*p2 */
        double q1 = p1 + Gv[1][2] * p2;
        P_(0, j) = q0; P_(1, j) = q1; P_(2, j) = p2;
        P_(j, 0) = q0; P_(j, 1) = q1; P_(j, 2) = p2;
    }
    double x0 = x[0] + vtc, x1 = x[1] + vts;
    x[2] = twin_mpi_to_pi(phi + v * dt * sin(g) / w);
    x[0] = x0; x[1] = x1;
}

/* 2x2 innovation covariance S = H*P*H' + R for landmark idf, from the 5x5 sub-block. */
static void innov_cov(const double *x, const double *P, int ld, const double *R, int64_t idf,
                      double *zp, double *S /* s11 s12 s21 s22 */) {
    double h[10];
    int f = twin_predict_observation(x, idf, zp, h);
    int c[5] = {0, 1, 2, f, f + 1};
    double Hr[2][5] = {{h[0], h[1], h[2], h[6], h[7]}, {h[3], h[4], h[5], h[8], h[9]}};
    double HP[2][5];
    for (int a = 0; a < 2; a++)
        for (int q = 0; q < 5; q++) {
            double acc = 0;
            for (int p = 0; p < 5; p++) acc += Hr[a][p] * P_(c[p], c[q]);
            HP[a][q] = acc;
        }
    for (int a = 0; a < 2; a++)
        for (int b = 0; b < 2; b++) {
            double acc = 0;
            for (int q = 0; q < 5; q++) acc += HP[a][q] * Hr[b][q];
            S[2 * a + b] = acc + R[a + 2 * b];
        }
}

/* src/data-association.jl:53-63 */
void twin_compute_association(const double *x, const double *P, int n, int ld, const double *z,
                              const double *R, int64_t idf, double *nis, double *nd) {
    (void)n;
    double zp[2], S[4];
    innov_cov(x, P, ld, R, idf, zp, S);
    double v0 = z[0] - zp[0];
    double v1 = twin_mpi_to_pi(z[1] - zp[1]);
    double det = S[0] * S[3] - S[1] * S[2];
    double i11 = S[3] / det, i12 = -S[1] / det, i21 = -S[2] / det, i22 = S[0] / det;
    double t0 = i11 * v0 + i12 * v1;
    double t1 = i21 * v0 + i22 * v1;
    *nis = v0 * t0 + v1 * t1;
    *nd = *nis + log(det);
}

/* src/data-association.jl:1-51.  zf (2 x m), idf (m, 1-based), zn (2 x k) are caller
 * buffers of capacity nz columns.  *min_margin (may be NULL) receives the smallest distance
 * of a decisive comparison from its threshold over this call. */
void twin_associate(const double *x, const double *P, int n, int ld, const double *z, int nz,
                    const double *R, double gate1, double gate2, double *zf, int64_t *idf,
                    int *m_out, double *zn, int *k_out, double *min_margin) {
    int Nf = (n - 3) / 2;
    int m = 0, k = 0;
    double mm = INFINITY;
    for (int i = 0; i < nz; i++) {
        int jbest = 0;
        double nbest = INFINITY, outer = INFINITY;
        double second = INFINITY, gm = INFINITY;
        for (int j = 1; j <= Nf; j++) {
            double nis, nd;
            twin_compute_association(x, P, n, ld, z + 2 * i, R, j, &nis, &nd);
            if (fabs(nis - gate1) < gm) gm = fabs(nis - gate1);
            int ingate = 0;
            if (nis < gate1) {
                if (nd < nbest) ingate = 1;
                else if (nd < second) second = nd;
            }
            if (ingate) {
                if (nbest < second) second = nbest;
                nbest = nd;
                jbest = j;
            } else if (nis < outer) {
                outer = nis;
            }
        }
        if (Nf > 0) {
            if (gm < mm) mm = gm;
            if (jbest != 0) { if (second - nbest < mm) mm = second - nbest; }
            else if (fabs(outer - gate2) < mm) mm = fabs(outer - gate2);
        }
        if (jbest != 0) {
            zf[2 * m] = z[2 * i]; zf[2 * m + 1] = z[2 * i + 1];
            idf[m] = jbest; m++;
        } else if (outer > gate2) {
            zn[2 * k] = z[2 * i]; zn[2 * k + 1] = z[2 * i + 1]; k++;
        }
    }
    *m_out = m; *k_out = k;
    if (min_margin) *min_margin = mm;
}

/* src/ekf.jl:46-77 -- in place on x and the full P.  Returns 0, or -1 if S is not PD. */
int twin_update(double *x, double *P, int n, int ld, const double *zf, int m, const double *R,
                const int64_t *idf) {
    if (m == 0) return 0;
    int r = 2 * m;
    double *Hn = malloc(sizeof(double) * 10 * m);   /* nonzeros per observation */
    int *fp = malloc(sizeof(int) * m);
    double *v = malloc(sizeof(double) * r);
    double *PHt = malloc(sizeof(double) * (size_t)n * r);   /* n x r, row-major [i*r+k] */
    double *W1 = malloc(sizeof(double) * (size_t)n * r);
    double *S = calloc((size_t)r * r, sizeof(double));      /* col-major r x r */
    double *C = calloc((size_t)r * r, sizeof(double));
    double *u = malloc(sizeof(double) * r);
    int rc = 0;
    for (int o = 0; o < m; o++) {
        double zp[2];
        fp[o] = twin_predict_observation(x, idf[o], zp, Hn + 10 * o);
        v[2 * o] = zf[2 * o] - zp[0];
        v[2 * o + 1] = twin_mpi_to_pi(zf[2 * o + 1] - zp[1]);
    }
    /* PHt = P*H' : five terms per entry, columns in ascending order 1,2,3,f,f+1 */
    for (int i = 0; i < n; i++)
        for (int o = 0; o < m; o++) {
            const double *h = Hn + 10 * o;
            int f = fp[o];
            double p0 = P_(i, 0), p1 = P_(i, 1), p2 = P_(i, 2), p3 = P_(i, f), p4 = P_(i, f + 1);
            PHt[(size_t)i * r + 2 * o] = p0 * h[0] + p1 * h[1] + p2 * h[2] + p3 * h[6] + p4 * h[7];
            PHt[(size_t)i * r + 2 * o + 1] = p0 * h[3] + p1 * h[4] + p2 * h[5] + p3 * h[8] + p4 * h[9];
        }
    /* S = H*PHt + RR ; S = (S+S')*0.5 */
    for (int o = 0; o < m; o++) {
        const double *h = Hn + 10 * o;
        int f = fp[o];
        for (int k = 0; k < r; k++) {
            double a0 = PHt[(size_t)0 * r + k], a1 = PHt[(size_t)1 * r + k], a2 = PHt[(size_t)2 * r + k];
            double a3 = PHt[(size_t)f * r + k], a4 = PHt[(size_t)(f + 1) * r + k];
            S[(2 * o) + (size_t)k * r] = h[0] * a0 + h[1] * a1 + h[2] * a2 + h[6] * a3 + h[7] * a4;
            S[(2 * o + 1) + (size_t)k * r] = h[3] * a0 + h[4] * a1 + h[5] * a2 + h[8] * a3 + h[9] * a4;
        }
        for (int a = 0; a < 2; a++)
            for (int b = 0; b < 2; b++) S[(2 * o + a) + (size_t)(2 * o + b) * r] += R[a + 2 * b];
    }
    for (int a = 0; a < r; a++)
        for (int b = a; b < r; b++) {
            double s = (S[a + (size_t)b * r] + S[b + (size_t)a * r]) * 0.5;
            S[a + (size_t)b * r] = s; S[b + (size_t)a * r] = s;
        }
    /* chol: S = U'U, U upper, stored over the upper triangle of S (column by column) */
    for (int j = 0; j < r && rc == 0; j++) {
        for (int i = 0; i < j; i++) {
            double a = S[i + (size_t)j * r];
            for (int k = 0; k < i; k++) a -= S[k + (size_t)i * r] * S[k + (size_t)j * r];
            S[i + (size_t)j * r] = a / S[i + (size_t)i * r];
        }
        double d = S[j + (size_t)j * r];
        for (int k = 0; k < j; k++) d -= S[k + (size_t)j * r] * S[k + (size_t)j * r];
        if (!(d > 0)) { rc = -1; break; }
        S[j + (size_t)j * r] = sqrt(d);
    }
    if (rc == 0) {
        /* C = inv(U), upper triangular, in the order of LAPACK's unblocked dtrti2 ('U', 'N'), which
         * is what inv(::UpperTriangular) reaches for these sizes: column j = the already inverted
         * leading block times U[0:j, j] (reference dtrmv: y_i = x_i*T_ii + sum_{jj>i} x_jj*T_i,jj,
         * jj ascending), scaled by -1/U[j][j]. */
        for (int j = 0; j < r; j++) {
            double cjj = 1.0 / S[j + (size_t)j * r];
            C[j + (size_t)j * r] = cjj;
            double ajj = -cjj;
            for (int i = 0; i < j; i++) {
                double y = S[i + (size_t)j * r] * C[i + (size_t)i * r];
                for (int jj = i + 1; jj < j; jj++) y += S[jj + (size_t)j * r] * C[i + (size_t)jj * r];
                C[i + (size_t)j * r] = y * ajj;
            }
        }
        /* W1 = PHt*C ; u = C'*v ; x += W1*u  ( = W*v with W = W1*C' ) */
        for (int k = 0; k < r; k++) {
            double a = 0;
            for (int l = 0; l <= k; l++) a += C[l + (size_t)k * r] * v[l];
            u[k] = a;
        }
        for (int i = 0; i < n; i++) {
            double dxi = 0;
            for (int k = 0; k < r; k++) {
                double a = 0;
                for (int l = 0; l <= k; l++) a += PHt[(size_t)i * r + l] * C[l + (size_t)k * r];
                W1[(size_t)i * r + k] = a;
                dxi += a * u[k];
            }
            x[i] += dxi;
        }
        /* P -= W1*W1' */
        for (int j = 0; j < n; j++)
            for (int i = 0; i < n; i++) {
                double a = P_(i, j);
                for (int k = 0; k < r; k++) a = fma(-W1[(size_t)i * r + k], W1[(size_t)j * r + k], a);
                P_(i, j) = a;
            }
    }
    free(Hn); free(fp); free(v); free(PHt); free(W1); free(S); free(C); free(u);
    return rc;
}

/* src/ekf.jl:84-122 -- in place; needs ld >= n + 2k.  Returns the new n. */
int twin_add_features(double *x, double *P, int n, int ld, const double *zn, int k,
                      const double *R) {
    double phi = x[2];
    for (int q = 0; q < k; q++) {
        int len = n;
        double r = zn[2 * q], b = zn[2 * q + 1];
        double s = sin(phi + b), c = cos(phi + b);
        x[len] = x[0] + r * c;
        x[len + 1] = x[1] + r * s;
        double Gv[2][3] = {{1, 0, -r * s}, {0, 1, r * c}};
        double Gz[2][2] = {{c, -r * s}, {s, r * c}};
        double GP[2][3], GR[2][2];
        for (int a = 0; a < 2; a++)
            for (int j = 0; j < 3; j++)
                GP[a][j] = Gv[a][0] * P_(0, j) + Gv[a][1] * P_(1, j) + Gv[a][2] * P_(2, j);
        for (int a = 0; a < 2; a++)
            for (int j = 0; j < 2; j++) GR[a][j] = Gz[a][0] * R[0 + 2 * j] + Gz[a][1] * R[1 + 2 * j];
        for (int a = 0; a < 2; a++)
            for (int bb = 0; bb < 2; bb++) {
                double t1 = GP[a][0] * Gv[bb][0] + GP[a][1] * Gv[bb][1] + GP[a][2] * Gv[bb][2];
                double t2 = GR[a][0] * Gz[bb][0] + GR[a][1] * Gz[bb][1];
                P_(len + a, len + bb) = t1 + t2;
            }
        for (int a = 0; a < 2; a++)
            for (int j = 0; j < 3; j++) { P_(len + a, j) = GP[a][j]; P_(j, len + a) = GP[a][j]; }
        for (int j = 3; j < len; j++)
            for (int a = 0; a < 2; a++) {
                double t = Gv[a][0] * P_(0, j) + Gv[a][1] * P_(1, j) + Gv[a][2] * P_(2, j);
                P_(len + a, j) = t; P_(j, len + a) = t;
            }
        n += 2;
    }
    return n;
}

/* ------------------------------------------------------------------------------------ */
/* Counter-based standard normals (Philox4x32-10 + Box-Muller).  Not in the reference:   */
/* the reference draws from Julia's global RNG (sim/sim-utils.jl:36-37,68), which cannot */
/* be reproduced without Julia.  "Same seeds" is defined at this generator instead.      */
/* normal(seed, run, stream, idx): stream 0 = speed noise, 1 = steering noise, 2 = obs.  */
/* ------------------------------------------------------------------------------------ */
static inline void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
    for (int r = 0; r < 10; r++) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c[0];
        uint64_t p1 = (uint64_t)0xCD9E8D57u * c[2];
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
}

double twin_normal(uint64_t seed, uint32_t run, uint32_t stream, uint32_t idx) {
    uint32_t c[4] = {idx, stream, run, 0u};
    philox4x32_10(c, (uint32_t)seed, (uint32_t)(seed >> 32));
    double u1 = ((double)(c[0] >> 5) * 67108864.0 + (double)(c[1] >> 6)) / 9007199254740992.0;
    double u2 = ((double)(c[2] >> 5) * 67108864.0 + (double)(c[3] >> 6)) / 9007199254740992.0;
    return sqrt(-2.0 * log(1.0 - u1)) * cos(6.283185307179586 * u2);
}

/* Truth-side vehicle state (src/common.jl:36-55 with sim/ekfslam-sim.jl:29-37 values). */
typedef struct {
    double pose[3], target_gamma;
    int waypoint_id;
} veh_t;

static double sgn(double a) { return (double)(a > 0) - (double)(a < 0); }

/* src/common.jl:189-230 */
static void steer(veh_t *v, const double *wp, int nwp, double d_min, double dt,
                  double steer_rate, double max_gamma) {
    double g = v->target_gamma;
    int iwp = v->waypoint_id;
    double x = v->pose[0], y = v->pose[1], phi = v->pose[2];
    double cx = wp[2 * (iwp - 1)], cy = wp[2 * (iwp - 1) + 1];
    double d2 = (cx - x) * (cx - x) + (cy - y) * (cy - y);
    if (d2 < d_min * d_min) {
        iwp += 1;
        if (iwp > nwp) { v->waypoint_id = 0; return; }
        cx = wp[2 * (iwp - 1)]; cy = wp[2 * (iwp - 1) + 1];
    }
    double dg = twin_mpi_to_pi(atan2(cy - y, cx - x) - phi - g);
    double dgmax = steer_rate * dt;
    if (fabs(dg) > dgmax) dg = sgn(dg) * dgmax;
    g += dg;
    if (fabs(g) > max_gamma) g = sgn(g) * max_gamma;
    v->target_gamma = g;
    v->waypoint_id = iwp;
}

/* Headless sim!(simdata) of sim/ekfslam-sim.jl:54-143 (sleep and monitor removed).
 * wp: 2 x nwp col-major; lm: 2 x nl col-major.
 * Noise: if ctrl_noise != NULL it is 2 x max_steps col-major and obs_noise is a flat pool of
 * nobs normals consumed in draw order; otherwise Philox normals keyed by (seed, run).
 * Outputs (any may be NULL): true_track / slam_track 3 x max_steps col-major; x_out (3+2nl),
 * P_out ((3+2nl)^2 col-major, ld = 3+2nl); counts[4] = {nsteps, nupdates, n_final, obs_used};
 * *min_margin = smallest association margin of the run.  Returns 0, -1 not PD, -2 overflow. */
int twin_sim_cap(const double *wp, int nwp, const double *lm, int nl, const double *ctrl_noise,
                 const double *obs_noise, int nobs, uint64_t seed, uint32_t run, int nlaps,
                 int max_steps, double *true_track, double *slam_track, double *x_out, double *P_out,
                 int *counts, double *min_margin, int extra) {
    const double wheelbase = 4.0, max_gamma = 60 * M_PI / 180, steer_rate = 60 * M_PI / 180;
    const double sensor_range = 30.0, target_speed = 8.0, d_min = 1.0;
    const double sigma_speed = 0.5, sigma_steer = 3.0 * M_PI / 180;
    const double sigmaR = 0.1, sigmaB = 1.0 * M_PI / 180;
    const double Q[4] = {sigma_speed * sigma_speed, 0, 0, sigma_steer * sigma_steer};
    const double R[4] = {sigmaR * sigmaR, 0, 0, sigmaB * sigmaB};
    const double dt = 0.025, dt_obs = 8 * dt;
    /* the reference's state simply grows (hcat/vcat, src/ekf.jl:108-109); here the capacity is
     * fixed at 3 + 2*(nl + extra): `extra` leaves room for spurious landmarks (an observation of a
     * mapped landmark that fails both gates becomes a new one, src/data-association.jl:45-47) */
    int ld = 3 + 2 * (nl + extra), n = 3, rc = 0;
    double *x = calloc(ld, sizeof(double));
    double *P = calloc((size_t)ld * ld, sizeof(double));
    double *z = malloc(sizeof(double) * 2 * (nl + 1)), *zf = malloc(sizeof(double) * 2 * (nl + 1));
    double *zn = malloc(sizeof(double) * 2 * (nl + 1));
    int64_t *idf = malloc(sizeof(int64_t) * (nl + 1));
    veh_t veh;
    veh.pose[0] = wp[0]; veh.pose[1] = wp[1];
    veh.pose[2] = atan2(wp[3] - wp[1], wp[2] - wp[0]);      /* src/common.jl:93-96 */
    veh.target_gamma = 0; veh.waypoint_id = 1;
    x[0] = veh.pose[0]; x[1] = veh.pose[1]; x[2] = veh.pose[2];
    int nsteps = 0, nupd = 0, cursor = 0;
    double dtsum = 0, mm = INFINITY;
    while (veh.waypoint_id != 0) {
        if (nsteps >= max_steps) { rc = -2; break; }
        steer(&veh, wp, nwp, d_min, dt, steer_rate, max_gamma);
        if (veh.waypoint_id == 0 && nlaps > 1) { veh.waypoint_id = 1; nlaps -= 1; }
        {   /* step_vehicle! src/common.jl:172-181 */
            double px = veh.pose[0], py = veh.pose[1], phi = veh.pose[2], g = veh.target_gamma;
            veh.pose[0] = px + target_speed * dt * cos(g + phi);
            veh.pose[1] = py + target_speed * dt * sin(g + phi);
            veh.pose[2] = twin_mpi_to_pi(phi + target_speed * dt * sin(g) / wheelbase);
        }
        double n0 = ctrl_noise ? ctrl_noise[2 * nsteps] : twin_normal(seed, run, 0, nsteps);
        double n1 = ctrl_noise ? ctrl_noise[2 * nsteps + 1] : twin_normal(seed, run, 1, nsteps);
        double mspeed = target_speed + n0 * sqrt(Q[0]);      /* sim/sim-utils.jl:35-38 */
        double mgamma = veh.target_gamma + n1 * sqrt(Q[3]);
        twin_predict(x, P, n, ld, mgamma, mspeed, wheelbase, Q, dt);
        dtsum += dt;
        if (dtsum > dt_obs) {
            dtsum = 0; nupd++;
            /* get_observations sim/sim-utils.jl:53-75 with nearby_landmark_indices :12-28 */
            int nz = 0;
            double px = veh.pose[0], py = veh.pose[1], phi = veh.pose[2];
            for (int i = 0; i < nl; i++) {
                double dx = lm[2 * i] - px, dy = lm[2 * i + 1] - py;
                if ((dx * cos(phi) + dy * sin(phi)) > 0 && (dx * dx + dy * dy) < sensor_range * sensor_range) {
                    z[2 * nz] = sqrt(dx * dx + dy * dy);
                    z[2 * nz + 1] = atan2(dy, dx) - phi;
                    nz++;
                }
            }
            if (ctrl_noise && cursor + 2 * nz > nobs) { rc = -2; break; }
            for (int c = 0; c < nz; c++)
                z[2 * c] += (ctrl_noise ? obs_noise[cursor + c] : twin_normal(seed, run, 2, cursor + c)) * sqrt(R[0]);
            for (int c = 0; c < nz; c++)
                z[2 * c + 1] += (ctrl_noise ? obs_noise[cursor + nz + c] : twin_normal(seed, run, 2, cursor + nz + c)) * sqrt(R[3]);
            cursor += 2 * nz;
            int m, k; double mg;
            twin_associate(x, P, n, ld, z, nz, R, 4.0, 25.0, zf, idf, &m, zn, &k, &mg);
            if (mg < mm) mm = mg;
            if (twin_update(x, P, n, ld, zf, m, R, idf) != 0) { rc = -1; break; }
            if (n + 2 * k > ld) { rc = -2; break; }
            n = twin_add_features(x, P, n, ld, zn, k, R);
        }
        if (true_track) memcpy(true_track + 3 * nsteps, veh.pose, 3 * sizeof(double));
        if (slam_track) memcpy(slam_track + 3 * nsteps, x, 3 * sizeof(double));
        nsteps++;
    }
    if (x_out) memcpy(x_out, x, sizeof(double) * ld);
    if (P_out) memcpy(P_out, P, sizeof(double) * (size_t)ld * ld);
    if (counts) { counts[0] = nsteps; counts[1] = nupd; counts[2] = n; counts[3] = cursor; }
    if (min_margin) *min_margin = mm;
    free(x); free(P); free(z); free(zf); free(zn); free(idf);
    return rc;
}

/* capacity 3 + 2*nl (no head-room): the original entry point, x_out / P_out sized by that */
int twin_sim(const double *wp, int nwp, const double *lm, int nl, const double *ctrl_noise,
             const double *obs_noise, int nobs, uint64_t seed, uint32_t run, int nlaps,
             int max_steps, double *true_track, double *slam_track, double *x_out, double *P_out,
             int *counts, double *min_margin) {
    return twin_sim_cap(wp, nwp, lm, nl, ctrl_noise, obs_noise, nobs, seed, run, nlaps, max_steps,
                        true_track, slam_track, x_out, P_out, counts, min_margin, 0);
}

/* Ensemble on the host cores: runs [run0, run0+nruns) with Philox noise, one run per OpenMP
 * thread.  final_pose: 3 x nruns; slam_tracks (may be NULL): 3 x max_steps x nruns;
 * n_final (may be NULL): nruns.  Returns the number of failed runs. */
int twin_ensemble_cap(const double *wp, int nwp, const double *lm, int nl, uint64_t seed,
                      uint32_t run0, int nruns, int nlaps, int max_steps, double *final_pose,
                      double *slam_tracks, int *n_final, int nthreads, int extra) {
    int bad = 0;
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic, 1) num_threads(nthreads) reduction(+ : bad)
#endif
    for (int r = 0; r < nruns; r++) {
        double *xo = malloc(sizeof(double) * (3 + 2 * (nl + extra)));
        int counts[4];
        double *trk = slam_tracks ? slam_tracks + (size_t)3 * max_steps * r : NULL;
        int rc = twin_sim_cap(wp, nwp, lm, nl, NULL, NULL, 0, seed, run0 + r, nlaps, max_steps, NULL,
                              trk, xo, NULL, counts, NULL, extra);
        if (rc != 0) bad++;
        memcpy(final_pose + 3 * r, xo, 3 * sizeof(double));
        if (n_final) n_final[r] = counts[2];
        free(xo);
    }
    return bad;
}

int twin_ensemble(const double *wp, int nwp, const double *lm, int nl, uint64_t seed,
                  uint32_t run0, int nruns, int nlaps, int max_steps, double *final_pose,
                  double *slam_tracks, int *n_final, int nthreads) {
    return twin_ensemble_cap(wp, nwp, lm, nl, seed, run0, nruns, nlaps, max_steps, final_pose,
                             slam_tracks, n_final, nthreads, 0);
}
