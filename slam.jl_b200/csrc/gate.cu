// K2 -- Mahalanobis nearest-neighbour gating: associate() and compute_association()
// (src/data-association.jl:1-63).  This translation unit is compiled with -fmad=false so that
// every product/sum rounds exactly as the plain-C oracle twin does (gate tests are strict
// FP64 comparisons, :30-31,38,46).
//
// The reference forms S = H*P*H' + R with a dense 2 x n Jacobian; H has nonzeros only in
// columns {1,2,3,fpos,fpos+1} (src/common.jl:160-162), so S depends on a 5 x 5 sub-block of P.
// One thread owns a landmark j: it gathers that sub-block once, forms S, det(S) and inv(S)
// in closed form, then scores up to OB observations held in registers.  Per observation
// the running (nd, j) minimum over gated landmarks and the minimum nis are reduced with warp
// shuffles (lowest j wins ties, NaN never wins -- the strict `<` of the reference), then
// across warps in shared memory, then across CTAs by the last CTA to finish, which also
// classifies each observation {matched, new, dropped} and compacts zf / idf / zn in
// observation order.
//
// Equivalence of the parallel reduction with the reference's sequential scan: `outer` is
// only consulted when jbest == 0 (:43-46), i.e. when no landmark ever passed
// `nis < gate1 && nd < nbest`; then every landmark fell to the `elseif nis < outer` branch,
// so outer = min_j nis_j independent of scan order.
//
// Spatial pre-gate (the reference's own TODO at src/data-association.jl:18-20).
// A pair (observation i, landmark j) may be skipped without changing any output iff its
// nis exceeds G = max(gate1, gate2): it can then neither pass `nis < gate1` (:30) nor pull
// `outer` to or below gate2 (:38,46).  For a positive definite S, Cauchy-Schwarz gives
//     nis = v' inv(S) v >= v_r^2 / S_rr                       (v_r = range innovation)
// and S_rr = h P5 h' + R_rr with h = [-u, 0, u], |u| = 1, is the variance of (landmark -
// vehicle) along u plus R_rr, so for a positive semidefinite 5 x 5 sub-block
//     S_rr <= 2 (P_xx + P_yy) + 4 sig2max + R_rr =: S_rr_max,
// sig2max being an upper bound of every landmark's diagonal entries of P.  The kernel skips
// landmark j for observation i when (z_r - d_j)^2 > PREGATE_SAFETY * G * S_rr_max (a factor 4
// of head-room over the exact bound covers every rounding of the FP64 evaluation): the
// landmark lies outside an annulus around the vehicle.  That test needs x only (16 coalesced
// bytes per landmark); the scattered 5 x 5 gather, atan2, log and the scoring are done for the
// survivors, compacted per CTA so that they fill whole warps.  sig2max lives on the device: an
// exact diagonal reduction at slam_set_state / slam_synth_state, raised by K7 for new
// landmarks, and never lowered by update (P_ii - sum W1_ik^2 <= P_ii also in floating point).
// NaN anywhere makes the test false, i.e. the landmark is kept.
#include "common.cuh"
#include "kernels.cuh"

namespace {

constexpr int GATE_THREADS = 256;
constexpr int OB = 8;            // observations scored per thread per pass
constexpr int GATE_MAX_BLOCKS = 592;
constexpr int GATE_SMS = 148;

struct Innov {     // per-landmark quantities that do not depend on the observation
    double zp0, zp1, i11, i12, i21, i22, logdet;
};

// Where a landmark's 5x5 sub-block comes from.  Single filter: everything from the one
// column-major P.  Sharded filter: rows 0..2 of every column from P3 (broadcast from rank 0,
// P3[p*npad + j]) and the landmark's 2x2 diagonal block from the owning rank's panel stack.
struct GateSrc {
    const double *P;      // single: full matrix; sharded: nullptr
    size_t ld;
    const double *P3;     // sharded: rows 0..2, stride npad
    int npad;
    const double *Pl;     // sharded: the rank's row panels (kernels.cuh), leading dimension ldl
    size_t ldl;
    int rank, world, n;   // sharded: scan slot t = 128 * local panel + k is the k-th landmark pair of that panel
    int Nf;               // single: landmarks 1..Nf, slot t = landmark t + 1
};

// landmark id (1-based) of scan slot t, or 0 when the slot holds no landmark of the current state
__device__ __forceinline__ int gate_slot_landmark(const GateSrc &g, int t) {
    if (g.P) return t < g.Nf ? t + 1 : 0;
    const int p = panel_global(t / (SH_PH / 2), g.rank, g.world);
    const int f = (p == 0 ? 3 : panel_first(p)) + 2 * (t % (SH_PH / 2));      // panels start on a landmark's first row
    if (f + 1 >= min(panel_first(p + 1), g.n)) return 0;
    return (f - 3) / 2 + 1;
}

__device__ __forceinline__ double gate_pvv(const GateSrc &g, int p, int q) {
    int a = p <= q ? p : q, b = p <= q ? q : p;
    return g.P ? g.P[(size_t)a + (size_t)b * g.ld] : g.P3[(size_t)a * g.npad + b];
}

// 2x2 innovation covariance from the 5x5 sub-block; Pvv is the symmetric 3x3 pose block.
__device__ __forceinline__ Innov landmark_innov(const double *__restrict__ x, const GateSrc &g, int f,
                                                const double *Pvv, double xv, double yv, double phi,
                                                const double *R) {
    double zp[2], h[10];
    observe_model(xv, yv, phi, x[f], x[f + 1], zp, h);
    double P5[5][5];
#pragma unroll
    for (int p = 0; p < 3; p++)
#pragma unroll
        for (int q = 0; q < 3; q++) P5[p][q] = Pvv[p * 3 + q];
    if (g.P) {
        const double *c0 = g.P + (size_t)f * g.ld;
        const double *c1 = c0 + g.ld;
#pragma unroll
        for (int p = 0; p < 3; p++) {
            P5[p][3] = c0[p];
            P5[p][4] = c1[p];
        }
        P5[3][3] = c0[f];
        P5[3][4] = c1[f];
        P5[4][4] = c1[f + 1];
    } else {
#pragma unroll
        for (int p = 0; p < 3; p++) {
            P5[p][3] = g.P3[(size_t)p * g.npad + f];
            P5[p][4] = g.P3[(size_t)p * g.npad + f + 1];
        }
        const int p = panel_of_row(f);
        const double *d0 = g.Pl + (size_t)panel_local(p, g.world) * SH_PH + (f - panel_first(p)) + (size_t)f * g.ldl;
        P5[3][3] = d0[0];
        P5[3][4] = d0[g.ldl];
        P5[4][4] = d0[g.ldl + 1];
    }
#pragma unroll
    for (int p = 0; p < 3; p++) {
        P5[3][p] = P5[p][3];
        P5[4][p] = P5[p][4];
    }
    P5[4][3] = P5[3][4];
    const double Hr[2][5] = {{h[0], h[1], h[2], h[6], h[7]}, {h[3], h[4], h[5], h[8], h[9]}};
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
            S[2 * a + b] = acc + R[a + 2 * b];
        }
    double det = S[0] * S[3] - S[1] * S[2];
    Innov o;
    o.zp0 = zp[0];
    o.zp1 = zp[1];
    o.i11 = S[3] / det;
    o.i12 = -S[1] / det;
    o.i21 = -S[2] / det;
    o.i22 = S[0] / det;
    o.logdet = log(det);
    return o;
}

__device__ __forceinline__ void score(const Innov &L, double z0, double z1, double *nis, double *nd) {
    double v0 = z0 - L.zp0;
    double v1 = mpi_to_pi(z1 - L.zp1);
    double t0 = L.i11 * v0 + L.i12 * v1;
    double t1 = L.i21 * v0 + L.i22 * v1;
    *nis = v0 * t0 + v1 * t1;
    *nd = *nis + L.logdet;
}

constexpr double PREGATE_SAFETY = 4.0;

struct GateParams {
    double R[4];
    double gate1, gate2;
    const unsigned long long *sig2bits;   // bit pattern of sig2max (non-negative double); nullptr: no pre-gate
    unsigned long long *stats;            // [0] landmarks scanned, [1] landmarks scored (cumulative); may be nullptr
};

struct GateRes {   // one rank's reduction for one observation (24 bytes, allgathered as bytes)
    double nd, outer;
    long long j;
};

// Stable compaction of the classified observations into zf / idf / zn (observation order kept, as the
// reference appends in scan order, src/data-association.jl:43-48).  Called by every thread of the CTA:
// ballot ranks inside a warp, warp totals through shared memory, chunks of blockDim observations.
__device__ void gate_compact(const int *cls, const double *__restrict__ z, int nz, double *zf, int64_t *idf,
                             double *zn, int *counts) {
    __shared__ int s_wm[32], s_wk[32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    int m0 = 0, k0 = 0;   // matched / new before this chunk
    for (int c0 = 0; c0 < nz; c0 += blockDim.x) {
        const int i = c0 + threadIdx.x;
        const int c = i < nz ? cls[i] : 0;
        double z0 = 0, z1 = 0;
        if (c != 0) {
            z0 = z[2 * i];
            z1 = z[2 * i + 1];
        }
        const unsigned bm = __ballot_sync(0xffffffffu, c > 0), bk = __ballot_sync(0xffffffffu, c < 0);
        if (lane == 0) {
            s_wm[warp] = __popc(bm);
            s_wk[warp] = __popc(bk);
        }
        __syncthreads();
        int mo = m0, ko = k0, mt = 0, kt = 0;
        for (int w = 0; w < nwarps; w++) {
            if (w < warp) {
                mo += s_wm[w];
                ko += s_wk[w];
            }
            mt += s_wm[w];
            kt += s_wk[w];
        }
        const unsigned lt = (1u << lane) - 1u;
        if (c > 0) {
            int r = mo + __popc(bm & lt);
            zf[2 * r] = z0;
            zf[2 * r + 1] = z1;
            idf[r] = c;
        } else if (c < 0) {
            int r = ko + __popc(bk & lt);
            zn[2 * r] = z0;
            zn[2 * r + 1] = z1;
        }
        m0 += mt;
        k0 += kt;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        counts[0] = m0;
        counts[1] = k0;
    }
}

// sharded path: every rank reduces the allgathered per-rank results identically
__global__ void k_gate_finalize(const GateRes *__restrict__ all, int world, const double *__restrict__ z, int nz,
                                GateParams prm, double *zf, int64_t *idf, double *zn, int *counts) {
    __shared__ int s_cls[SLAM_MAX_OBS];
    for (int i = threadIdx.x; i < nz; i += blockDim.x) {
        double b = INFINITY, ou = INFINITY;
        long long jj = 0;
        for (int r = 0; r < world; r++) {
            GateRes g = all[(size_t)r * nz + i];
            if (g.j != 0 && (jj == 0 || g.nd < b || (g.nd == b && g.j < jj))) {
                b = g.nd;
                jj = g.j;
            }
            if (g.outer < ou) ou = g.outer;
        }
        s_cls[i] = jj != 0 ? (int)jj : (ou > prm.gate2 ? -1 : 0);
    }
    __syncthreads();
    gate_compact(s_cls, z, nz, zf, idf, zn, counts);
}

// scratch: [0] unsigned counter, then part_nd / part_outer (double) and part_j (int) of
// nblocks*nz entries each
__global__ void __launch_bounds__(GATE_THREADS, 2)
k_gate(const double *__restrict__ x, GateSrc src, int Ntot, const double *__restrict__ z, int nz, GateParams prm,
       unsigned *counter, double *part_nd, double *part_outer, int *part_j, double *zf, int64_t *idf, double *zn,
       int *counts, GateRes *rank_res) {
    __shared__ double sPvv[9];
    __shared__ double s_nd[GATE_THREADS / 32][OB], s_outer[GATE_THREADS / 32][OB];
    __shared__ int s_j[GATE_THREADS / 32][OB];
    __shared__ bool s_last;
    __shared__ int s_list[GATE_THREADS];
    __shared__ int s_cnt;
    const int nblocks = gridDim.x;
    const int ob0 = blockIdx.y * OB;
    // every global load that does not depend on another is issued ahead of the first barrier
    const double xv = x[0], yv = x[1], phi = x[2];
    const double sig2 = prm.sig2bits ? __longlong_as_double((long long)*prm.sig2bits) : 0.0;
    if (threadIdx.x < 9) {
        int p = threadIdx.x / 3, q = threadIdx.x % 3;
        sPvv[threadIdx.x] = gate_pvv(src, p, q);
    }
    __syncthreads();
    // pre-gate radius^2; +inf (keep everything) when there is no bound or no positive gate
    double rho2 = INFINITY;
    if (prm.sig2bits) {
        double G = prm.gate1 > prm.gate2 ? prm.gate1 : prm.gate2;
        double srr = 2.0 * (fabs(sPvv[0]) + fabs(sPvv[4])) + 4.0 * sig2 + fabs(prm.R[0]);
        double r2 = PREGATE_SAFETY * G * srr;
        if (G > 0.0 && r2 >= 0.0) rho2 = r2;   // NaN fails the test and leaves +inf
    }
    double zr[OB], zb[OB], best[OB], outer[OB];
    int jb[OB];
#pragma unroll
    for (int o = 0; o < OB; o++) {
        int io = ob0 + o;
        zr[o] = io < nz ? z[2 * io] : INFINITY;   // a padded slot never keeps a landmark; its scores are not stored
        zb[o] = io < nz ? z[2 * io + 1] : 0.0;
        best[o] = INFINITY;
        outer[o] = INFINITY;
        jb[o] = 0;
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned scored = 0;
    for (int base = blockIdx.x * GATE_THREADS; base < Ntot; base += nblocks * GATE_THREADS) {
        if (threadIdx.x == 0) s_cnt = 0;
        __syncthreads();
        // phase A: the annulus test on x alone; survivors are compacted (warp-aggregated)
        const int t = base + threadIdx.x;
        bool keep = false;
        const int j = t < Ntot ? gate_slot_landmark(src, t) : 0;
        if (j > 0) {
            const int f = 3 + 2 * (j - 1);
            const double dx = x[f] - xv, dy = x[f + 1] - yv;
            const double d = sqrt(dx * dx + dy * dy);
#pragma unroll
            for (int o = 0; o < OB; o++) {
                double vr = zr[o] - d;
                keep = keep || !(vr * vr > rho2);
            }
        }
        const unsigned mask = __ballot_sync(0xffffffffu, keep);
        int wbase = 0;
        if (lane == 0 && mask) wbase = atomicAdd(&s_cnt, __popc(mask));
        wbase = __shfl_sync(0xffffffffu, wbase, 0);
        if (keep) s_list[wbase + __popc(mask & ((1u << lane) - 1u))] = t;
        __syncthreads();
        // phase B: full scoring of the survivors, one per thread
        const int cnt = s_cnt;
        if (threadIdx.x < cnt) {
            const int j = gate_slot_landmark(src, s_list[threadIdx.x]);
            Innov L = landmark_innov(x, src, 3 + 2 * (j - 1), sPvv, xv, yv, phi, prm.R);
#pragma unroll
            for (int o = 0; o < OB; o++) {
                double nis, nd;
                score(L, zr[o], zb[o], &nis, &nd);
                if (nis < prm.gate1 && nd < best[o]) {
                    best[o] = nd;
                    jb[o] = j;
                }
                if (nis < outer[o]) outer[o] = nis;
            }
        }
        scored += cnt;
        __syncthreads();
    }
    if (prm.stats && threadIdx.x == 0) {
        if (blockIdx.y == 0) {
            int mine = 0;
            for (int base = blockIdx.x * GATE_THREADS; base < Ntot; base += nblocks * GATE_THREADS)
                mine += min(GATE_THREADS, Ntot - base);
            atomicAdd(&prm.stats[0], (unsigned long long)mine);
        }
        atomicAdd(&prm.stats[1], (unsigned long long)scored);
    }
    if (scored == 0) {
        // no landmark of this CTA survived the pre-gate (the usual case): an empty partial, no reduction
        if (threadIdx.x < OB && ob0 + threadIdx.x < nz) {
            size_t slot = (size_t)blockIdx.x * nz + ob0 + threadIdx.x;
            part_nd[slot] = INFINITY;
            part_j[slot] = 0;
            part_outer[slot] = INFINITY;
        }
    } else {
        // warp reduction: (nd, j) lexicographic minimum, and min outer
#pragma unroll
        for (int o = 0; o < OB; o++) {
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) {
                double nd2 = __shfl_down_sync(0xffffffffu, best[o], off);
                int j2 = __shfl_down_sync(0xffffffffu, jb[o], off);
                double ou2 = __shfl_down_sync(0xffffffffu, outer[o], off);
                if (j2 != 0 && (jb[o] == 0 || nd2 < best[o] || (nd2 == best[o] && j2 < jb[o]))) {
                    best[o] = nd2;
                    jb[o] = j2;
                }
                if (ou2 < outer[o]) outer[o] = ou2;
            }
            if (lane == 0) {
                s_nd[warp][o] = best[o];
                s_j[warp][o] = jb[o];
                s_outer[warp][o] = outer[o];
            }
        }
        __syncthreads();
        if (threadIdx.x < OB && ob0 + threadIdx.x < nz) {
            int o = threadIdx.x;
            double b = s_nd[0][o], ou = s_outer[0][o];
            int jj = s_j[0][o];
            for (int w = 1; w < GATE_THREADS / 32; w++) {
                double nd2 = s_nd[w][o];
                int j2 = s_j[w][o];
                if (j2 != 0 && (jj == 0 || nd2 < b || (nd2 == b && j2 < jj))) {
                    b = nd2;
                    jj = j2;
                }
                if (s_outer[w][o] < ou) ou = s_outer[w][o];
            }
            size_t slot = (size_t)blockIdx.x * nz + ob0 + o;
            part_nd[slot] = b;
            part_j[slot] = jj;
            part_outer[slot] = ou;
        }
    }
    // last CTA to finish reduces across CTAs, classifies and compacts
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned total = gridDim.x * gridDim.y;
        unsigned t = atomicAdd(counter, 1u);
        s_last = (t == total - 1);
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    __shared__ int s_cls[SLAM_MAX_OBS];   // >0: matched landmark id, 0: dropped, -1: new
    for (int i = warp; i < nz; i += GATE_THREADS / 32) {   // a warp per observation, lanes over the CTAs
        double b = INFINITY, ou = INFINITY;
        int jj = 0;
        for (int blk = lane; blk < nblocks; blk += 32) {
            size_t slot = (size_t)blk * nz + i;
            // other CTAs wrote these: read them at L2 (the threadFenceReduction pattern), never from a stale L1 line
            double nd2 = __ldcg(part_nd + slot);
            int j2 = __ldcg(part_j + slot);
            const double ou1 = __ldcg(part_outer + slot);
            if (j2 != 0 && (jj == 0 || nd2 < b || (nd2 == b && j2 < jj))) {
                b = nd2;
                jj = j2;
            }
            if (ou1 < ou) ou = ou1;
        }
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            double nd2 = __shfl_down_sync(0xffffffffu, b, off);
            int j2 = __shfl_down_sync(0xffffffffu, jj, off);
            double ou2 = __shfl_down_sync(0xffffffffu, ou, off);
            if (j2 != 0 && (jj == 0 || nd2 < b || (nd2 == b && j2 < jj))) {
                b = nd2;
                jj = j2;
            }
            if (ou2 < ou) ou = ou2;
        }
        if (lane == 0) {
            s_cls[i] = jj != 0 ? jj : (ou > prm.gate2 ? -1 : 0);
            if (rank_res) {   // sharded: this rank's result; classification happens after the allgather
                rank_res[i].nd = b;
                rank_res[i].outer = ou;
                rank_res[i].j = jj;
            }
        }
    }
    __syncthreads();
    if (rank_res) {
        if (threadIdx.x == 0) *counter = 0;
        return;
    }
    gate_compact(s_cls, z, nz, zf, idf, zn, counts);
    if (threadIdx.x == 0) *counter = 0;
}

__global__ void k_compute_association(const double *__restrict__ x, const double *__restrict__ P,
                                      size_t ld, int n, double z0, double z1, GateParams prm, int64_t idf,
                                      double *out, int *err) {
    int Nf = (n - 3) / 2;
    if (idf < 1 || idf > Nf) {
        atomicCAS(err, 0, SLAM_ERR_BAD_INDEX);
        out[0] = out[1] = nan("");
        return;
    }
    GateSrc src = {};
    src.P = P;
    src.ld = ld;
    double Pvv[9];
    for (int p = 0; p < 3; p++)
        for (int q = 0; q < 3; q++) Pvv[p * 3 + q] = gate_pvv(src, p, q);
    Innov L = landmark_innov(x, src, 3 + 2 * ((int)idf - 1), Pvv, x[0], x[1], x[2], prm.R);
    score(L, z0, z1, &out[0], &out[1]);
}

}  // namespace

static int gate_nblocks(int Nf) {
    int nb = (Nf + GATE_THREADS - 1) / GATE_THREADS;
    if (nb < 1) nb = 1;
    if (nb > GATE_MAX_BLOCKS) nb = GATE_MAX_BLOCKS;
    return nb;
}

size_t gate_scratch_bytes(int nz) {
    size_t per = (size_t)GATE_MAX_BLOCKS * (size_t)(nz > 0 ? nz : 1);
    return 256 + per * (8 + 8 + 4);
}

// sig2max = max(0, max_f P[f, f], f >= 3) as the bit pattern of a non-negative double (ordered like
// an unsigned integer; a NaN entry gives a pattern above +inf, which switches the pre-gate off).
__global__ void k_sig2_diag(const double *__restrict__ P, size_t ld, int n, unsigned long long *sig2bits) {
    unsigned long long mine = 0ull;
    for (int f = 3 + blockIdx.x * blockDim.x + threadIdx.x; f < n; f += gridDim.x * blockDim.x) {
        double v = P[(size_t)f * (ld + 1)];
        unsigned long long b = (unsigned long long)__double_as_longlong(v);
        if (v != v) b = 0x7ff8000000000000ull;
        else if (v < 0.0) b = 0ull;
        mine = b > mine ? b : mine;
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        unsigned long long o = __shfl_down_sync(0xffffffffu, mine, off);
        mine = o > mine ? o : mine;
    }
    if ((threadIdx.x & 31) == 0 && mine) atomicMax(sig2bits, mine);
}

int launch_sig2_refresh(const double *d_P, size_t ld, int n, unsigned long long *d_sig2bits, cudaStream_t s,
                        long long *launches) {
    CUDA_TRY(cudaMemsetAsync(d_sig2bits, 0, sizeof(unsigned long long), s));
    if (n <= 3) return SLAM_OK;
    int nb = (n - 3 + 255) / 256;
    if (nb > 296) nb = 296;
    k_sig2_diag<<<nb, 256, 0, s>>>(d_P, ld, n, d_sig2bits);
    if (launches) (*launches)++;
    CUDA_TRY(cudaGetLastError());
    return SLAM_OK;
}

int launch_gate(const double *d_x, const double *d_P, size_t ld, int Nf, const double *d_z, int nz,
                const double R[4], double gate1, double gate2, void *d_scratch, size_t scratch_bytes,
                double *d_zf, int64_t *d_idf, double *d_zn, int *d_counts,
                const unsigned long long *d_sig2bits, unsigned long long *d_stats, cudaStream_t s,
                long long *launches) {
    if (scratch_bytes < gate_scratch_bytes(nz)) {
        slam_set_error("launch_gate: scratch too small");
        return SLAM_ERR_ARG;
    }
    GateParams prm;
    memcpy(prm.R, R, sizeof(prm.R));
    prm.gate1 = gate1;
    prm.gate2 = gate2;
    prm.sig2bits = d_sig2bits;
    prm.stats = d_stats;
    size_t per = (size_t)GATE_MAX_BLOCKS * (size_t)nz;
    unsigned *counter = (unsigned *)d_scratch;
    double *part_nd = (double *)((char *)d_scratch + 256);
    double *part_outer = part_nd + per;
    int *part_j = (int *)(part_outer + per);
    GateSrc src = {};
    src.P = d_P;
    src.ld = ld;
    src.Nf = Nf;
    const int ny = (nz + OB - 1) / OB;
    int nb = gate_nblocks(Nf);
    if (d_sig2bits) {
        // pre-gated: a landmark costs one coalesced 16-byte load and a sqrt, so the whole grid is kept within
        // one wave (two CTAs per SM) and a CTA strides over several landmarks per thread
        int cap = 2 * GATE_SMS / ny;
        if (cap < 16) cap = 16;
        if (nb > cap) nb = cap;
    }
    dim3 grid(nb, ny);
    k_gate<<<grid, GATE_THREADS, 0, s>>>(d_x, src, Nf, d_z, nz, prm, counter, part_nd, part_outer, part_j, d_zf,
                                        d_idf, d_zn, d_counts, nullptr);
    if (launches) (*launches)++;
    CUDA_TRY(cudaGetLastError());
    return SLAM_OK;
}

int launch_compute_association(const double *d_x, const double *d_P, size_t ld, int n, const double z[2],
                               const double R[4], int64_t idf, double *d_out2, int *d_err,
                               cudaStream_t s, long long *launches) {
    GateParams prm;
    memcpy(prm.R, R, sizeof(prm.R));
    prm.gate1 = prm.gate2 = 0;
    prm.sig2bits = nullptr;
    prm.stats = nullptr;
    k_compute_association<<<1, 1, 0, s>>>(d_x, d_P, ld, n, z[0], z[1], prm, idf, d_out2, d_err);
    if (launches) (*launches)++;
    CUDA_TRY(cudaGetLastError());
    return SLAM_OK;
}

// Sharded gating, phase 1: this rank's landmarks -> d_rank_res[nz] (GateRes, 24 bytes each).
int launch_gate_shard(const double *d_x, const double *d_P3, int npad, const double *d_Pl, size_t ldl, int rank, int world,
                      int nlp, int n, const double *d_z, int nz, const double R[4], double gate1,
                      double gate2, void *d_scratch, size_t scratch_bytes, void *d_rank_res,
                      const unsigned long long *d_sig2bits, cudaStream_t s, long long *launches) {
    if (scratch_bytes < gate_scratch_bytes(nz)) {
        slam_set_error("launch_gate_shard: scratch too small");
        return SLAM_ERR_ARG;
    }
    GateParams prm;
    memcpy(prm.R, R, sizeof(prm.R));
    prm.gate1 = gate1;
    prm.gate2 = gate2;
    prm.sig2bits = d_sig2bits;   // rank-local bound over the landmarks this rank scans (shard.cu); nullptr = exhaustive
    prm.stats = nullptr;
    GateSrc src = {};
    src.P3 = d_P3;
    src.npad = npad;
    src.Pl = d_Pl;
    src.ldl = ldl;
    src.rank = rank;
    src.world = world;
    src.n = n;
    const int total = nlp * (SH_PH / 2);   // scan slots: 128 landmark pairs per local panel (some empty)
    size_t per = (size_t)GATE_MAX_BLOCKS * (size_t)nz;
    unsigned *counter = (unsigned *)d_scratch;
    double *part_nd = (double *)((char *)d_scratch + 256);
    double *part_outer = part_nd + per;
    int *part_j = (int *)(part_outer + per);
    dim3 grid(gate_nblocks(total), (nz + OB - 1) / OB);
    k_gate<<<grid, GATE_THREADS, 0, s>>>(d_x, src, total, d_z, nz, prm, counter, part_nd, part_outer, part_j, nullptr,
                                        nullptr, nullptr, nullptr, (GateRes *)d_rank_res);
    if (launches) (*launches)++;
    CUDA_TRY(cudaGetLastError());
    return SLAM_OK;
}

// phase 2 (after the allgather of the per-rank results): classify and compact, identically on every rank
int launch_gate_finalize(const void *d_all_res, int world, const double *d_z, int nz, double gate2, double *d_zf,
                         int64_t *d_idf, double *d_zn, int *d_counts, cudaStream_t s, long long *launches) {
    GateParams prm = {};
    prm.gate2 = gate2;
    k_gate_finalize<<<1, 256, 0, s>>>((const GateRes *)d_all_res, world, d_z, nz, prm, d_zf, d_idf, d_zn, d_counts);
    if (launches) (*launches)++;
    CUDA_TRY(cudaGetLastError());
    return SLAM_OK;
}
