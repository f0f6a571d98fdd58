// Synthetic large-map state generated on the device: the
// 100k-landmark covariance (320 GB) cannot exist on a host, so each entry is a pure function
// of (seed, row, col) through Philox4x32-10 and a CPU mirror (tests/synth_mirror.py) can
// reproduce any sampled block.
//
//   landmark l truth   = side * (U1, U2),           (U1, U2) = uniform2(seed, 0, stream 10, l)
//   x[3+2l], x[4+2l]   = truth + 0.1 * N(seed, 0, 11, 2l), 0.1 * N(seed, 0, 11, 2l+1)
//   pose               = (side/2, side/2, 0)
//   D[i]               = 0.01 + 0.09 * U1(seed, 0, 12, i)
//   L[i][c]            = 0.05 * N(seed, 0, 13, 8i + c),  c = 0..7
//   P[i][j]            = sum_c L[i][c]*L[j][c]  (+ D[i] if i == j)
#include <algorithm>
#include <vector>

#include "common.cuh"
#include "kernels.cuh"

__host__ __device__ inline void synth_landmark(uint64_t seed, double side, int l, double *lx, double *ly) {
    double u1, u2;
    philox_uniform2(seed, 0u, 10u, (uint32_t)l, &u1, &u2);
    *lx = side * u1;
    *ly = side * u2;
}

__global__ void k_synth_x(double *x, int n_landmarks, double side, uint64_t seed) {
    int l = blockIdx.x * blockDim.x + threadIdx.x;
    if (l == 0) {
        x[0] = side * 0.5;
        x[1] = side * 0.5;
        x[2] = 0.0;
    }
    if (l >= n_landmarks) return;
    double lx, ly;
    synth_landmark(seed, side, l, &lx, &ly);
    x[3 + 2 * l] = lx + 0.1 * philox_normal(seed, 0u, 11u, (uint32_t)(2 * l));
    x[4 + 2 * l] = ly + 0.1 * philox_normal(seed, 0u, 11u, (uint32_t)(2 * l + 1));
}

__global__ void k_synth_L(double *L, double *D, int n, uint64_t seed) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double u1, u2;
    philox_uniform2(seed, 0u, 12u, (uint32_t)i, &u1, &u2);
    D[i] = 0.01 + 0.09 * u1;
    for (int c = 0; c < 8; c++) L[(size_t)i * 8 + c] = 0.05 * philox_normal(seed, 0u, 13u, (uint32_t)(8 * i + c));
}

// Rows [r0, r1) x all columns of P into a block with leading dimension ldb
// (element (i, j) at Pb[(i - r0) + j*ldb]).
__global__ void k_synth_P(double *Pb, size_t ldb, int r0, int r1, int n, const double *__restrict__ L,
                          const double *__restrict__ D) {
    int i = r0 + blockIdx.y * blockDim.x + threadIdx.x;
    int j = blockIdx.x;
    if (i >= r1 || j >= n) return;
    const double *li = L + (size_t)i * 8, *lj = L + (size_t)j * 8;
    double a = 0;
#pragma unroll
    for (int c = 0; c < 8; c++) a += li[c] * lj[c];
    if (i == j) a += D[i];
    Pb[(size_t)(i - r0) + (size_t)j * ldb] = a;
}

int synth_fill(double *d_x, double *d_Pb, size_t ldb, int r0, int r1, int n_landmarks, double side,
               uint64_t seed, cudaStream_t s, long long *launches) {
    int n = 3 + 2 * n_landmarks;
    double *L = nullptr, *D = nullptr;
    CUDA_TRY(cudaMalloc(&L, sizeof(double) * (size_t)n * 8));
    CUDA_TRY(cudaMalloc(&D, sizeof(double) * (size_t)n));
    k_synth_x<<<(std::max(n_landmarks, 1) + 255) / 256, 256, 0, s>>>(d_x, n_landmarks, side, seed);
    k_synth_L<<<(n + 255) / 256, 256, 0, s>>>(L, D, n, seed);
    if (r1 > r0) {
        dim3 grid(n, (r1 - r0 + 255) / 256);
        k_synth_P<<<grid, 256, 0, s>>>(d_Pb, ldb, r0, r1, n, L, D);
        if (launches) (*launches)++;
    }
    if (launches) (*launches) += 2;
    cudaError_t e = cudaStreamSynchronize(s);
    cudaFree(L);
    cudaFree(D);
    CUDA_TRY(e);
    CUDA_TRY(cudaGetLastError());
    return SLAM_OK;
}

// The rows of a row map (kernels.cuh) into a local array: a CTA row of 256 threads = one 256-row tile of the map.
__global__ void k_synth_P_map(double *Pl, size_t ldl, RowMap map, int n, const double *__restrict__ L,
                              const double *__restrict__ D) {
    int i0, loc0, r1;
    if (!rowmap_tile(map, blockIdx.y, 256, n, &i0, &loc0, &r1)) return;
    const int i = i0 + threadIdx.x, j = blockIdx.x;
    if (i >= r1 || j >= n) return;
    const double *li = L + (size_t)i * 8, *lj = L + (size_t)j * 8;
    double a = 0;
#pragma unroll
    for (int c = 0; c < 8; c++) a += li[c] * lj[c];
    if (i == j) a += D[i];
    Pl[(size_t)(loc0 + threadIdx.x) + (size_t)j * ldl] = a;
}

int synth_fill_map(double *d_x, double *d_Pl, size_t ldl, RowMap map, int n_landmarks, double side, uint64_t seed,
                   cudaStream_t s, long long *launches) {
    int n = 3 + 2 * n_landmarks;
    double *L = nullptr, *D = nullptr;
    CUDA_TRY(cudaMalloc(&L, sizeof(double) * (size_t)n * 8));
    CUDA_TRY(cudaMalloc(&D, sizeof(double) * (size_t)n));
    k_synth_x<<<(std::max(n_landmarks, 1) + 255) / 256, 256, 0, s>>>(d_x, n_landmarks, side, seed);
    k_synth_L<<<(n + 255) / 256, 256, 0, s>>>(L, D, n, seed);
    const int rows = rowmap_local_rows(map);
    if (rows > 0) {
        dim3 grid(n, (rows + 255) / 256);
        k_synth_P_map<<<grid, 256, 0, s>>>(d_Pl, ldl, map, n, L, D);
        if (launches) (*launches)++;
    }
    if (launches) (*launches) += 2;
    cudaError_t e = cudaStreamSynchronize(s);
    cudaFree(L);
    cudaFree(D);
    CUDA_TRY(e);
    CUDA_TRY(cudaGetLastError());
    return SLAM_OK;
}

// Host side: the m nearest landmarks ahead of the vehicle (truth positions) followed by
// `spurious` observations of points at least 6 m from every landmark.
int synth_observations_host(int n_landmarks, double side, uint64_t seed, int m, int spurious, double *z,
                            int64_t *ids) {
    if (m < 0 || spurious < 0 || m > n_landmarks) {
        slam_set_error("slam_synth_observations: bad m / spurious");
        return SLAM_ERR_ARG;
    }
    const double xv = side * 0.5, yv = side * 0.5;
    std::vector<double> lx(n_landmarks), ly(n_landmarks);
    std::vector<std::pair<double, int>> ahead;
    for (int l = 0; l < n_landmarks; l++) {
        synth_landmark(seed, side, l, &lx[l], &ly[l]);
        double dx = lx[l] - xv, dy = ly[l] - yv;
        if (dx > 0) ahead.push_back({dx * dx + dy * dy, l});
    }
    if ((int)ahead.size() < m) {
        slam_set_error("slam_synth_observations: only %d landmarks ahead, need %d", (int)ahead.size(), m);
        return SLAM_ERR_ARG;
    }
    std::partial_sort(ahead.begin(), ahead.begin() + m, ahead.end());
    for (int o = 0; o < m; o++) {
        int l = ahead[o].second;
        double dx = lx[l] - xv, dy = ly[l] - yv;
        z[2 * o] = sqrt(dx * dx + dy * dy);
        z[2 * o + 1] = atan2(dy, dx) - 0.0;
        if (ids) ids[o] = l + 1;
    }
    int found = 0;
    for (int t = 0; t < 100000 && found < spurious; t++) {
        double u1, u2;
        philox_uniform2(seed, 0u, 14u, (uint32_t)t, &u1, &u2);
        double r = 5.0 + 40.0 * u1, b = (u2 - 0.5) * 3.0;
        double px = xv + r * cos(b), py = yv + r * sin(b);
        bool ok = true;
        for (int l = 0; l < n_landmarks && ok; l++) {
            double dx = lx[l] - px, dy = ly[l] - py;
            if (dx * dx + dy * dy < 36.0) ok = false;
        }
        if (!ok) continue;
        z[2 * (m + found)] = r;
        z[2 * (m + found) + 1] = b;
        found++;
    }
    if (found < spurious) {
        slam_set_error("slam_synth_observations: could not place %d spurious observations", spurious);
        return SLAM_ERR_ARG;
    }
    return SLAM_OK;
}
