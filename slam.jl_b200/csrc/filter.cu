// Single-filter device-resident EKF-SLAM: handle lifecycle, state I/O and the kernels
// K1 (predict), K3 (P*H' gather), K4 (innovation Cholesky), K5 (gain, x update),
// K7 (in-place augmentation).  K2 lives in gate.cu, K6 in downdate.cu.
// Compiled with -fmad=false: products and sums round separately, in the operation order of
// the CPU oracle's plain-C twin.  Reference citations: file:line under /root/reference.
#include <stdarg.h>

#include <algorithm>
#include <mutex>
#include <vector>

#include "common.cuh"
#include "kernels.cuh"

// ---------------------------------------------------------------------------------------
// errors
// ---------------------------------------------------------------------------------------
static thread_local char g_err[512] = "";

void slam_set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" const char *slam_last_error(void) { return g_err; }
extern "C" const char *slam_version(void) { return "libslamcuda 0.1 sm_100a"; }
extern "C" int slam_device_count(void) {
    int c = 0;
    if (cudaGetDeviceCount(&c) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return c;
}

int slam_require_device(int device) {
    int c = slam_device_count();
    if (device < 0 || device >= c) {
        slam_set_error("no usable CUDA device %d (%d visible); libslamcuda has no CPU fallback", device, c);
        return SLAM_ERR_CUDA;
    }
    cudaDeviceProp prop;
    CUDA_TRY(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) {
        slam_set_error("device %d is sm_%d%d; libslamcuda is built for sm_100a only", device, prop.major,
                       prop.minor);
        return SLAM_ERR_CUDA;
    }
    CUDA_TRY(cudaSetDevice(device));
    return SLAM_OK;
}

// ---------------------------------------------------------------------------------------
// handle
// ---------------------------------------------------------------------------------------
struct slam_filter {
    int device = 0;
    int n_max = 0, n = 3;
    size_t ld = 0;
    cudaStream_t stream = nullptr;
    double *d_x = nullptr;
    double *d_P = nullptr;
    double *d_pose = nullptr;  // 2 x 4 doubles: pose snapshot double buffer read by K1
    int pose_slot = 0;
    int *d_err = nullptr;      // latched device-side error code
    // observation staging (device) and pinned host mirror
    double *d_z = nullptr, *d_zn = nullptr;
    char *d_assoc = nullptr;    // associate's result, packed per call: counts[4] | zf 2 x nz | zn 2 x nz | idf nz (one D2H)
    char *d_upd = nullptr;      // update's host inputs, packed per call: zf 2 x m | idf m (one H2D)
    double *d_small = nullptr;  // 16 doubles of misc outputs
    void *h_pin = nullptr;      // pinned host staging
    size_t h_pin_bytes = 0;
    void *d_gate_scratch = nullptr;
    size_t gate_scratch = 0;
    // spatial pre-gate of associate (gate.cu): [0] bit pattern of sig2max, [1] landmarks scanned,
    // [2] landmarks scored (cumulative)
    unsigned long long *d_sig2 = nullptr;
    bool pregate = true;
    // update workspace, grown on demand
    int ws_R = 0, ws_npad = 0;
    double *d_PHt = nullptr, *d_W1c = nullptr, *d_S = nullptr, *d_C = nullptr, *d_u = nullptr,
           *d_v = nullptr, *d_Hn = nullptr;
    int *d_fp = nullptr;
    double *d_block = nullptr;  // cov-block staging
    size_t block_cap = 0;
    int variant = 0;
    int last_k6 = 0;               // downdate kernel of the last update: 1 FMA, 2 tiled DMMA, 3 streaming DMMA
    // host-buffer slam_update returns when the step's RESULT is final (status after K3 + K4, x after K5), not when the
    // downdate is: ev_x is recorded behind K5 and read back on copy_stream, so the next call's copy and launches queue
    // up behind a K6 that is still running.  ev_stage: the pinned staging buffer's last H2D.  SLAM_UPDATE_EARLY=0 disables.
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t ev_x = nullptr, ev_stage = nullptr;
    bool early_return = true;
    long long launches = 0;
    // x[0:3] as read back with the last host-path update's status (one synchronisation for both): sim! reads
    // state.x[1:3] right after every call (sim/ekfslam-sim.jl:128); valid until the next call that changes the state
    double pose_cache[3] = {0, 0, 0};
    bool pose_valid = false;
    // K6 event timing (slam_profile)
    bool profiling = false;
    std::vector<cudaEvent_t> ev;   // pairs
    int ev_used = 0;
};

#define H_CHECK(h)                                                        \
    do {                                                                  \
        ARG_CHECK((h) != nullptr, "NULL handle");                         \
        CUDA_TRY(cudaSetDevice((h)->device));                             \
    } while (0)

static int ensure_update_ws(slam_t *h, int R) {
    int npad = round_up(h->n_max, DD_ROW_PAD);
    int RC = downdate_chunk_width(R);
    int Rpad = round_up(R, RC);
    if (h->ws_R >= Rpad && h->ws_npad == npad) return SLAM_OK;
    cudaFree(h->d_PHt); cudaFree(h->d_W1c); cudaFree(h->d_S); cudaFree(h->d_C);
    cudaFree(h->d_u); cudaFree(h->d_v); cudaFree(h->d_Hn); cudaFree(h->d_fp);
    h->d_PHt = h->d_W1c = h->d_S = h->d_C = h->d_u = h->d_v = h->d_Hn = nullptr;
    h->d_fp = nullptr;
    h->ws_R = 0;
    int Rw = std::max(Rpad, 32);
    CUDA_TRY(cudaMalloc(&h->d_PHt, sizeof(double) * (size_t)npad * Rw));
    CUDA_TRY(cudaMalloc(&h->d_W1c, sizeof(double) * w1_workspace_doubles(npad, Rw)));   // both layouts, every chunk
    CUDA_TRY(cudaMalloc(&h->d_S, sizeof(double) * (size_t)Rw * Rw));
    CUDA_TRY(cudaMalloc(&h->d_C, sizeof(double) * (size_t)Rw * Rw));
    CUDA_TRY(cudaMalloc(&h->d_u, sizeof(double) * Rw));
    CUDA_TRY(cudaMalloc(&h->d_v, sizeof(double) * Rw));
    CUDA_TRY(cudaMalloc(&h->d_Hn, sizeof(double) * 10 * (Rw / 2)));
    CUDA_TRY(cudaMalloc(&h->d_fp, sizeof(int) * (Rw / 2)));
    h->ws_R = Rw;
    h->ws_npad = npad;
    return SLAM_OK;
}

extern "C" int slam_create(int device, int n_max, slam_t **out) {
    ARG_CHECK(out != nullptr, "out is NULL");
    ARG_CHECK(n_max >= 3, "n_max must be >= 3");
    *out = nullptr;
    int rc = slam_require_device(device);
    if (rc) return rc;
    slam_t *h = new slam_t();
    h->device = device;
    h->n_max = n_max;
    h->ld = (size_t)round_up(n_max, 16);
    h->n = 3;
#define CREATE_TRY(expr)                                                                         \
    do {                                                                                         \
        cudaError_t e__ = (expr);                                                                \
        if (e__ != cudaSuccess) {                                                                \
            slam_set_error("%s failed: %s", #expr, cudaGetErrorString(e__));                     \
            slam_destroy(h);                                                                     \
            return SLAM_ERR_CUDA;                                                                \
        }                                                                                        \
    } while (0)
    CREATE_TRY(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
    CREATE_TRY(cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
    CREATE_TRY(cudaEventCreateWithFlags(&h->ev_x, cudaEventDisableTiming));
    CREATE_TRY(cudaEventCreateWithFlags(&h->ev_stage, cudaEventDisableTiming));
    if (const char *e = getenv("SLAM_UPDATE_EARLY")) h->early_return = atoi(e) != 0;
    CREATE_TRY(cudaMalloc(&h->d_x, sizeof(double) * h->ld));
    CREATE_TRY(cudaMalloc(&h->d_P, sizeof(double) * h->ld * (size_t)n_max));
    CREATE_TRY(cudaMalloc(&h->d_pose, sizeof(double) * 8));
    CREATE_TRY(cudaMalloc(&h->d_err, sizeof(int)));
    const size_t no = SLAM_MAX_OBS;
    CREATE_TRY(cudaMalloc(&h->d_z, sizeof(double) * 2 * no));
    CREATE_TRY(cudaMalloc(&h->d_zn, sizeof(double) * 2 * no));
    CREATE_TRY(cudaMalloc(&h->d_assoc, 16 + 40 * no));
    CREATE_TRY(cudaMalloc(&h->d_upd, 24 * no));
    CREATE_TRY(cudaMalloc(&h->d_small, sizeof(double) * 16));
    h->h_pin_bytes = sizeof(double) * 8 * no + 256;
    CREATE_TRY(cudaMallocHost(&h->h_pin, h->h_pin_bytes));
    h->gate_scratch = gate_scratch_bytes(SLAM_MAX_OBS);
    CREATE_TRY(cudaMalloc(&h->d_gate_scratch, h->gate_scratch));
    CREATE_TRY(cudaMemsetAsync(h->d_gate_scratch, 0, 256, h->stream));
    CREATE_TRY(cudaMalloc(&h->d_sig2, sizeof(unsigned long long) * 4));
    CREATE_TRY(cudaMemsetAsync(h->d_sig2, 0, sizeof(unsigned long long) * 4, h->stream));
    CREATE_TRY(cudaMemsetAsync(h->d_x, 0, sizeof(double) * h->ld, h->stream));
    CREATE_TRY(cudaMemsetAsync(h->d_P, 0, sizeof(double) * h->ld * (size_t)std::min(n_max, 64), h->stream));
    CREATE_TRY(cudaMemsetAsync(h->d_pose, 0, sizeof(double) * 8, h->stream));
    CREATE_TRY(cudaMemsetAsync(h->d_err, 0, sizeof(int), h->stream));
    CREATE_TRY(cudaStreamSynchronize(h->stream));
#undef CREATE_TRY
    *out = h;
    return SLAM_OK;
}

extern "C" int slam_destroy(slam_t *h) {
    if (!h) return SLAM_OK;
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    cudaFree(h->d_x); cudaFree(h->d_P); cudaFree(h->d_pose); cudaFree(h->d_err);
    cudaFree(h->d_z); cudaFree(h->d_zn); cudaFree(h->d_assoc); cudaFree(h->d_upd);
    cudaFree(h->d_small); cudaFree(h->d_gate_scratch); cudaFree(h->d_sig2);
    cudaFree(h->d_PHt); cudaFree(h->d_W1c); cudaFree(h->d_S); cudaFree(h->d_C);
    cudaFree(h->d_u); cudaFree(h->d_v); cudaFree(h->d_Hn); cudaFree(h->d_fp);
    cudaFree(h->d_block);
    if (h->h_pin) cudaFreeHost(h->h_pin);
    if (h->ev_x) cudaEventDestroy(h->ev_x);
    if (h->ev_stage) cudaEventDestroy(h->ev_stage);
    if (h->copy_stream) cudaStreamDestroy(h->copy_stream);
    for (auto &e : h->ev) cudaEventDestroy(e);
    if (h->stream) cudaStreamDestroy(h->stream);
    cudaGetLastError();
    delete h;
    return SLAM_OK;
}

// Capacity growth.  The reference's state grows without bound (hcat / vcat per new landmark,
// src/ekf.jl:108-109); a handle has a fixed capacity so that add_features is an in-place append.  A
// host that cannot size the map up front calls this when slam_add_features reports SLAM_ERR_CAPACITY
// (slamcuda/slam.py and julia/ekf-cuda.jl do): new arrays with the new leading dimension, one
// device-to-device copy of x and of the live columns of P.  On failure the old state is untouched.
extern "C" int slam_reserve(slam_t *h, int n_max_new) {
    H_CHECK(h);
    ARG_CHECK(n_max_new >= 3, "n_max must be >= 3");
    if (n_max_new <= h->n_max) return SLAM_OK;
    const size_t ld_new = (size_t)round_up(n_max_new, 16);
    double *nx = nullptr, *nP = nullptr;
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    if (cudaMalloc(&nx, sizeof(double) * ld_new) != cudaSuccess ||
        cudaMalloc(&nP, sizeof(double) * ld_new * (size_t)n_max_new) != cudaSuccess) {
        cudaGetLastError();
        cudaFree(nx);
        slam_set_error("slam_reserve: no device memory for a capacity of %d states", n_max_new);
        return SLAM_ERR_CUDA;
    }
    cudaError_t e = cudaMemsetAsync(nx, 0, sizeof(double) * ld_new, h->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(nx, h->d_x, sizeof(double) * h->n, cudaMemcpyDeviceToDevice, h->stream);
    if (e == cudaSuccess)
        e = cudaMemcpy2DAsync(nP, sizeof(double) * ld_new, h->d_P, sizeof(double) * h->ld, sizeof(double) * h->n, h->n,
                              cudaMemcpyDeviceToDevice, h->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
    if (e != cudaSuccess) {
        cudaFree(nx);
        cudaFree(nP);
        slam_set_error("slam_reserve: copy failed: %s", cudaGetErrorString(e));
        return SLAM_ERR_CUDA;
    }
    cudaFree(h->d_x);
    cudaFree(h->d_P);
    h->d_x = nx;
    h->d_P = nP;
    h->ld = ld_new;
    h->n_max = n_max_new;   // the update workspace follows on the next update (ensure_update_ws keys on n_max)
    return SLAM_OK;
}

extern "C" int slam_capacity(slam_t *h) {
    if (!h) {
        slam_set_error("slam_capacity: NULL handle");
        return SLAM_ERR_ARG;
    }
    return h->n_max;
}

extern "C" int slam_n(slam_t *h) {
    if (!h) {
        slam_set_error("slam_n: NULL handle");
        return SLAM_ERR_ARG;
    }
    return h->n;
}
extern "C" void *slam_stream(slam_t *h) { return h ? (void *)h->stream : nullptr; }
extern "C" long long slam_launch_count(slam_t *h) { return h ? h->launches : 0; }
extern "C" int slam_set_downdate_variant(slam_t *h, int variant) {
    ARG_CHECK(h != nullptr, "NULL handle");
    ARG_CHECK(variant >= 0 && variant <= 3, "variant must be 0, 1, 2 or 3");
    h->variant = variant;
    return SLAM_OK;
}
extern "C" int slam_downdate_kernel(slam_t *h) {
    ARG_CHECK(h != nullptr, "NULL handle");
    return h->last_k6;
}

static int check_latched(slam_t *h) {
    int *he = (int *)h->h_pin;
    double *hp = (double *)h->h_pin + 8;
    CUDA_TRY(cudaMemcpyAsync(he, h->d_err, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(cudaMemcpyAsync(hp, h->d_x, sizeof(double) * 3, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    h->pose_cache[0] = hp[0]; h->pose_cache[1] = hp[1]; h->pose_cache[2] = hp[2];
    h->pose_valid = true;      // everything issued so far is complete: the pose is current
    int e = *he;
    if (e != 0) {
        CUDA_TRY(cudaMemsetAsync(h->d_err, 0, sizeof(int), h->stream));
        slam_set_error(e == SLAM_ERR_NOT_PD      ? "innovation covariance not positive definite"
                       : e == SLAM_ERR_BAD_INDEX ? "landmark id out of range"
                                                 : "device-side error %d",
                       e);
        return e;
    }
    return SLAM_OK;
}

// check_latched for the host-buffer update: waits for ev_x (behind K5) instead of the whole stream.  The status was
// decided by K3 + K4 and x by K5; K6 neither fails nor touches x, and every later call is ordered behind it on the
// handle's stream, so nothing the caller can observe differs from a full synchronisation -- except that it returns
// while the downdate runs.  slam_sync() and every read-out still wait for everything.
static int check_after_x(slam_t *h) {
    int *he = (int *)h->h_pin;
    double *hp = (double *)h->h_pin + 8;
    CUDA_TRY(cudaStreamWaitEvent(h->copy_stream, h->ev_x, 0));
    CUDA_TRY(cudaMemcpyAsync(he, h->d_err, sizeof(int), cudaMemcpyDeviceToHost, h->copy_stream));
    CUDA_TRY(cudaMemcpyAsync(hp, h->d_x, sizeof(double) * 3, cudaMemcpyDeviceToHost, h->copy_stream));
    CUDA_TRY(cudaStreamSynchronize(h->copy_stream));
    h->pose_cache[0] = hp[0]; h->pose_cache[1] = hp[1]; h->pose_cache[2] = hp[2];
    h->pose_valid = true;      // x is final behind K5 and nothing queued behind it changes x
    int e = *he;
    if (e != 0) {
        CUDA_TRY(cudaMemsetAsync(h->d_err, 0, sizeof(int), h->stream));
        slam_set_error(e == SLAM_ERR_NOT_PD      ? "innovation covariance not positive definite"
                       : e == SLAM_ERR_BAD_INDEX ? "landmark id out of range"
                                                 : "device-side error %d",
                       e);
        return e;
    }
    return SLAM_OK;
}

extern "C" int slam_sync(slam_t *h) {
    H_CHECK(h);
    return check_latched(h);
}

// ---------------------------------------------------------------------------------------
// state I/O
// ---------------------------------------------------------------------------------------
// lower := upper' on the device copy (only for export; the lower triangle is not authoritative)
__global__ void k_mirror_lower(double *P, size_t ld, int n) {
    __shared__ double t[32][33];
    int bi = blockIdx.x, bj = blockIdx.y;
    if (bi > bj) return;
    int i = bi * 32 + threadIdx.y, j = bj * 32 + threadIdx.x;
    // read upper tile (rows bi, cols bj) transposed access pattern: coalesce along i
    int ii = bi * 32 + threadIdx.x, jj = bj * 32 + threadIdx.y;
    t[threadIdx.y][threadIdx.x] = (ii < n && jj < n && ii <= jj) ? P[(size_t)ii + (size_t)jj * ld] : 0.0;
    __syncthreads();
    // element (row = j, col = i) of the lower tile equals upper (i, j) = t[jloc... ]
    // t[a][b] holds upper(row = bi*32 + b, col = bj*32 + a); lower target (row = bj*32 + a', col = bi*32 + b')
    int r = bj * 32 + threadIdx.x;   // row of the lower element (contiguous across threads)
    int c = bi * 32 + threadIdx.y;   // its column
    (void)i; (void)j;
    if (r < n && c < n && r > c) P[(size_t)r + (size_t)c * ld] = t[threadIdx.x][threadIdx.y];
}

__global__ void k_set_pose_snapshot(const double *x, double *pose_slot) {
    if (threadIdx.x < 3) pose_slot[threadIdx.x] = x[threadIdx.x];
}

extern "C" int slam_set_state(slam_t *h, const double *x, const double *P, int n, int ldp) {
    SLAM_NVTX();
    H_CHECK(h);
    ARG_CHECK(x && P, "NULL x or P");
    ARG_CHECK(n >= 3 && (n - 3) % 2 == 0, "n must be 3 + 2*Nf");
    ARG_CHECK(ldp >= n, "ldp < n");
    if (n > h->n_max) {
        slam_set_error("slam_set_state: n = %d exceeds capacity n_max = %d", n, h->n_max);
        return SLAM_ERR_CAPACITY;
    }
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    h->pose_valid = false;
    CUDA_TRY(cudaMemsetAsync(h->d_gate_scratch, 0, 256, h->stream));   // K2's last-CTA ticket (robust to an aborted launch)
    CUDA_TRY(cudaMemcpyAsync(h->d_x, x, sizeof(double) * n, cudaMemcpyHostToDevice, h->stream));
    CUDA_TRY(cudaMemcpy2DAsync(h->d_P, sizeof(double) * h->ld, P, sizeof(double) * (size_t)ldp,
                               sizeof(double) * n, n, cudaMemcpyHostToDevice, h->stream));
    h->n = n;
    k_set_pose_snapshot<<<1, 32, 0, h->stream>>>(h->d_x, h->d_pose + 4 * h->pose_slot);
    h->launches++;
    CUDA_TRY(cudaGetLastError());
    int rc = launch_sig2_refresh(h->d_P, h->ld, n, h->d_sig2, h->stream, &h->launches);
    if (rc) return rc;
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    return SLAM_OK;
}

extern "C" int slam_get_state(slam_t *h, double *x, double *P, int ldp) {
    SLAM_NVTX();
    H_CHECK(h);
    int n = h->n;
    if (P) {
        ARG_CHECK(ldp >= n, "ldp < n");
        dim3 grid((n + 31) / 32, (n + 31) / 32), blk(32, 32);
        k_mirror_lower<<<grid, blk, 0, h->stream>>>(h->d_P, h->ld, n);
        h->launches++;
        CUDA_TRY(cudaGetLastError());
        CUDA_TRY(cudaMemcpy2DAsync(P, sizeof(double) * (size_t)ldp, h->d_P, sizeof(double) * h->ld,
                                   sizeof(double) * n, n, cudaMemcpyDeviceToHost, h->stream));
    }
    if (x) CUDA_TRY(cudaMemcpyAsync(x, h->d_x, sizeof(double) * n, cudaMemcpyDeviceToHost, h->stream));
    return check_latched(h);
}

extern "C" int slam_get_pose(slam_t *h, double pose[3]) {
    H_CHECK(h);
    ARG_CHECK(pose != nullptr, "NULL pose");
    if (h->pose_valid) {       // read back together with the status of the call that produced it
        pose[0] = h->pose_cache[0]; pose[1] = h->pose_cache[1]; pose[2] = h->pose_cache[2];
        return SLAM_OK;
    }
    double *hp = (double *)h->h_pin + 8;
    CUDA_TRY(cudaMemcpyAsync(hp, h->d_x, sizeof(double) * 3, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    pose[0] = hp[0]; pose[1] = hp[1]; pose[2] = hp[2];
    return SLAM_OK;
}

__global__ void k_cov_block(const double *P, size_t ld, int i0, int ni, int j0, int nj, double *out) {
    int a = blockIdx.y * blockDim.x + threadIdx.x;
    int b = blockIdx.x;
    if (a < ni && b < nj) out[(size_t)a + (size_t)b * ni] = sym_at(P, ld, i0 + a, j0 + b);
}

extern "C" int slam_get_cov_block(slam_t *h, int i0, int ni, int j0, int nj, double *out, int ldo) {
    H_CHECK(h);
    ARG_CHECK(out != nullptr, "NULL out");
    ARG_CHECK(i0 >= 0 && j0 >= 0 && ni >= 0 && nj >= 0 && i0 + ni <= h->n && j0 + nj <= h->n,
              "block outside the state");
    ARG_CHECK(ldo >= ni, "ldo < ni");
    if (ni == 0 || nj == 0) return SLAM_OK;
    size_t need = (size_t)ni * nj;
    if (need > h->block_cap) {
        cudaFree(h->d_block);
        h->d_block = nullptr;
        h->block_cap = 0;
        CUDA_TRY(cudaMalloc(&h->d_block, sizeof(double) * need));
        h->block_cap = need;
    }
    dim3 grid(nj, (ni + 127) / 128);
    k_cov_block<<<grid, 128, 0, h->stream>>>(h->d_P, h->ld, i0, ni, j0, nj, h->d_block);
    h->launches++;
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaMemcpy2DAsync(out, sizeof(double) * (size_t)ldo, h->d_block, sizeof(double) * ni,
                               sizeof(double) * ni, nj, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    return SLAM_OK;
}

// ---------------------------------------------------------------------------------------
// slam_get_cov_block_static -- slam_get_cov_block for callers that must not allocate (code under CUDA graph capture,
// tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The gather kernel is k_cov_block, unchanged; its scratch is the file-scope array below
// instead of the handle's cudaMalloc'ed d_block.
//   * SLAM_COV_BLOCK_STATIC_MAX doubles = 512 KiB, held by every GPU that loads the library until the process exits.
//   * In HOST code the array's name is a host-side stand-in, not a device address: the address comes from
//     cudaGetSymbolAddress, once per device with that device current (cov_block_static_scratch), and only that
//     pointer is ever handed to a kernel.
//   * Every call on a device shares the array, whatever the handle, and the array keeps the last call's contents.
//     k_cov_block writes all ni * nj elements of the window before the copy reads them, so nothing has to start at
//     zero and nothing is cleared.  Calls that overlap on the GPU (two handles' streams, two host threads) race on it.
// ---------------------------------------------------------------------------------------
static __device__ __align__(256) double g_cov_block_static[SLAM_COV_BLOCK_STATIC_MAX];

constexpr int COV_STATIC_MAX_DEVICES = 64;

// the device address of g_cov_block_static on `device`, which must be the current device
static int cov_block_static_scratch(int device, double **ptr) {
    static std::mutex mu;
    static double *addr[COV_STATIC_MAX_DEVICES] = {nullptr};
    ARG_CHECK(device >= 0 && device < COV_STATIC_MAX_DEVICES, "device ordinal outside 0..63");
    std::lock_guard<std::mutex> lock(mu);
    if (addr[device] == nullptr) {
        double *p = nullptr;
        CUDA_TRY(cudaGetSymbolAddress((void **)&p, g_cov_block_static));
        if (p == nullptr) {
            slam_set_error("slam_get_cov_block_static: cudaGetSymbolAddress gave a NULL address on device %d", device);
            return SLAM_ERR_CUDA;
        }
        addr[device] = p;
    }
    *ptr = addr[device];
    return SLAM_OK;
}

extern "C" int slam_get_cov_block_static(slam_t *h, int i0, int ni, int j0, int nj, double *out, int ldo) {
    H_CHECK(h);                                                  // makes h->device current
    ARG_CHECK(out != nullptr, "NULL out");
    ARG_CHECK(i0 >= 0 && j0 >= 0 && ni >= 0 && nj >= 0 && i0 + ni <= h->n && j0 + nj <= h->n,
              "block outside the state");
    ARG_CHECK(ldo >= ni, "ldo < ni");
    if ((long long)ni * nj > (long long)SLAM_COV_BLOCK_STATIC_MAX) {
        slam_set_error("slam_get_cov_block_static: a %d x %d window is %lld elements, more than the static scratch of "
                       "SLAM_COV_BLOCK_STATIC_MAX = %d; read it in pieces or use slam_get_cov_block",
                       ni, nj, (long long)ni * nj, SLAM_COV_BLOCK_STATIC_MAX);
        return SLAM_ERR_CAPACITY;
    }
    if (ni == 0 || nj == 0) return SLAM_OK;
    double *scratch = nullptr;
    int rc = cov_block_static_scratch(h->device, &scratch);
    if (rc) return rc;
    dim3 grid(nj, (ni + 127) / 128);
    k_cov_block<<<grid, 128, 0, h->stream>>>(h->d_P, h->ld, i0, ni, j0, nj, scratch);
    h->launches++;
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaMemcpy2DAsync(out, sizeof(double) * (size_t)ldo, scratch, sizeof(double) * ni,
                               sizeof(double) * ni, nj, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    return SLAM_OK;
}

// Padding guard (diagnostic; stands in for compute-sanitizer, which this pool does not offer).  The slab is ld x n_max
// doubles and x is ld doubles; predict / associate / update own only [0, n) x [0, n) and x[0, n).  arm = 1 paints every
// other element with a quiet-NaN payload, arm = 0 counts the painted elements whose bits have changed since.
static constexpr unsigned long long GUARD_BITS = 0x7FF8C0DEC0DEC0DEull;

__global__ void k_guard(double *x, double *P, size_t ld, int n, int n_max, int arm, unsigned long long *changed) {
    const size_t total = ld * (size_t)n_max + ld;                // the slab, then x
    unsigned long long bad = 0;
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        double *p;
        if (e < ld * (size_t)n_max) {
            const size_t i = e % ld, j = e / ld;
            if (i < (size_t)n && j < (size_t)n) continue;
            p = P + e;
        } else {
            const size_t i = e - ld * (size_t)n_max;
            if (i < (size_t)n) continue;
            p = x + i;
        }
        if (arm) *p = __longlong_as_double((long long)GUARD_BITS);
        else if ((unsigned long long)__double_as_longlong(*p) != GUARD_BITS) bad++;
    }
    if (!arm && bad) atomicAdd(changed, bad);
}

extern "C" int slam_debug_guard(slam_t *h, int arm, long long *changed) {
    H_CHECK(h);
    ARG_CHECK(arm || changed != nullptr, "NULL changed");
    unsigned long long *d_cnt = nullptr;
    CUDA_TRY(cudaMalloc(&d_cnt, sizeof(unsigned long long)));
    cudaError_t e = cudaMemsetAsync(d_cnt, 0, sizeof(unsigned long long), h->stream);
    if (e == cudaSuccess) {
        k_guard<<<148 * 8, 256, 0, h->stream>>>(h->d_x, h->d_P, h->ld, h->n, h->n_max, arm, d_cnt);
        h->launches++;
        e = cudaGetLastError();
    }
    unsigned long long cnt = 0;
    if (e == cudaSuccess) e = cudaMemcpyAsync(&cnt, d_cnt, sizeof(cnt), cudaMemcpyDeviceToHost, h->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
    cudaFree(d_cnt);
    CUDA_TRY(e);
    if (changed) *changed = (long long)cnt;
    return SLAM_OK;
}

// ---------------------------------------------------------------------------------------
// K1 predict -- src/ekf.jl:8-43.  O(n): rows 1:3 of every column.
// Bytes: 24 read + 16 written per map column (+ the 3x3 block and x[1:3]).
// ---------------------------------------------------------------------------------------
struct PredictParams {
    double g, v, w, dt, Q[4];
};

__global__ void __launch_bounds__(256)
k_predict(double *x, double *P, size_t ld, int n, PredictParams p, const double *pose_in, double *pose_out) {
    const double phi = pose_in[2];
    const double s = sin(p.g + phi), c = cos(p.g + phi);
    const double vts = p.v * p.dt * s, vtc = p.v * p.dt * c;
    int j = 3 + blockIdx.x * blockDim.x + threadIdx.x;
    if (j < n) {
        double *col = P + (size_t)j * ld;
        double p0 = col[0], p1 = col[1], p2 = col[2];
        col[0] = p0 + (-vts) * p2;   // Gv*P[1:3,4:end], Gv = [1 0 -vts; 0 1 vtc; 0 0 1]   (:24-26,34)
        col[1] = p1 + vtc * p2;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        const double Gv[3][3] = {{1, 0, -vts}, {0, 1, vtc}, {0, 0, 1}};
        const double Gu[3][2] = {{p.dt * c, -vts}, {p.dt * s, vtc},
                                 {p.dt * sin(p.g) / p.w, p.v * p.dt * cos(p.g) / p.w}};
        double Pv[3][3], A[3][3], B[3][3], GQ[3][2];
        for (int i = 0; i < 3; i++)
            for (int q = 0; q < 3; q++) Pv[i][q] = sym_at(P, ld, i, q);
        for (int i = 0; i < 3; i++)
            for (int q = 0; q < 3; q++) {
                double a = 0;
                for (int k = 0; k < 3; k++) a += Gv[i][k] * Pv[k][q];
                A[i][q] = a;
            }
        for (int i = 0; i < 3; i++)
            for (int q = 0; q < 3; q++) {
                double a = 0;
                for (int k = 0; k < 3; k++) a += A[i][k] * Gv[q][k];
                B[i][q] = a;
            }
        for (int i = 0; i < 3; i++)
            for (int q = 0; q < 2; q++) GQ[i][q] = Gu[i][0] * p.Q[0 + 2 * q] + Gu[i][1] * p.Q[1 + 2 * q];
        for (int i = 0; i < 3; i++)
            for (int q = 0; q < 3; q++) B[i][q] += GQ[i][0] * Gu[q][0] + GQ[i][1] * Gu[q][1];
        for (int i = 0; i < 3; i++)
            for (int q = i; q < 3; q++) P[(size_t)i + (size_t)q * ld] = B[i][q];
        double x0 = pose_in[0] + vtc, x1 = pose_in[1] + vts;
        double x2 = mpi_to_pi(phi + p.v * p.dt * sin(p.g) / p.w);
        x[0] = x0; x[1] = x1; x[2] = x2;
        pose_out[0] = x0; pose_out[1] = x1; pose_out[2] = x2;
    }
}

extern "C" int slam_predict(slam_t *h, double g, double v, double wheelbase, const double Q[4], double dt) {
    SLAM_NVTX();
    H_CHECK(h);
    ARG_CHECK(Q != nullptr, "NULL Q");
    h->pose_valid = false;
    PredictParams p;
    p.g = g; p.v = v; p.w = wheelbase; p.dt = dt;
    memcpy(p.Q, Q, sizeof(p.Q));
    int cols = h->n - 3;
    int blocks = std::max(1, (cols + 255) / 256);
    k_predict<<<blocks, 256, 0, h->stream>>>(h->d_x, h->d_P, h->ld, h->n, p, h->d_pose + 4 * h->pose_slot,
                                             h->d_pose + 4 * (h->pose_slot ^ 1));
    h->pose_slot ^= 1;
    h->launches++;
    CUDA_TRY(cudaGetLastError());
    return SLAM_OK;
}

// ---------------------------------------------------------------------------------------
// associate / compute_association / predict_observation host wrappers (kernels in gate.cu)
// ---------------------------------------------------------------------------------------
extern "C" int slam_associate(slam_t *h, const double *z, int nz, const double R[4], double gate1,
                              double gate2, double *zf, int64_t *idf, int *m, double *zn, int *k) {
    SLAM_NVTX();
    H_CHECK(h);
    ARG_CHECK(R && m && k, "NULL argument");
    ARG_CHECK(nz >= 0, "nz < 0");
    *m = 0;
    *k = 0;
    if (nz == 0) return SLAM_OK;
    ARG_CHECK(z && zf && idf && zn, "NULL observation buffer");
    if (nz > SLAM_MAX_OBS) {
        slam_set_error("slam_associate: nz = %d exceeds SLAM_MAX_OBS = %d", nz, SLAM_MAX_OBS);
        return SLAM_ERR_CAPACITY;
    }
    // pinned staging: [0, 16 + 40 nz) the packed result, the observations behind it
    const size_t out_bytes = 16 + 40 * (size_t)nz;
    double *hz = (double *)((char *)h->h_pin + 16 + 40 * (size_t)SLAM_MAX_OBS + 240);
    memcpy(hz, z, sizeof(double) * 2 * nz);
    CUDA_TRY(cudaMemcpyAsync(h->d_z, hz, sizeof(double) * 2 * nz, cudaMemcpyHostToDevice, h->stream));
    int Nf = (h->n - 3) / 2;
    int *d_counts = (int *)h->d_assoc;
    double *d_zf = (double *)(h->d_assoc + 16), *d_zn = d_zf + 2 * nz;
    int64_t *d_idf = (int64_t *)(d_zn + 2 * nz);
    int rc = launch_gate(h->d_x, h->d_P, h->ld, Nf, h->d_z, nz, R, gate1, gate2, h->d_gate_scratch,
                         h->gate_scratch, d_zf, d_idf, d_zn, d_counts,
                         h->pregate ? h->d_sig2 : nullptr, h->d_sig2 + 1, h->stream, &h->launches);
    if (rc) return rc;
    // counts, zf, zn, idf come back in one copy
    CUDA_TRY(cudaMemcpyAsync(h->h_pin, h->d_assoc, out_bytes, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    const int *hc = (const int *)h->h_pin;
    const double *hzf = (const double *)((char *)h->h_pin + 16), *hzn = hzf + 2 * nz;
    const int64_t *hid = (const int64_t *)(hzn + 2 * nz);
    *m = hc[0];
    *k = hc[1];
    memcpy(zf, hzf, sizeof(double) * 2 * (*m));
    memcpy(idf, hid, sizeof(int64_t) * (*m));
    memcpy(zn, hzn, sizeof(double) * 2 * (*k));
    return SLAM_OK;
}

extern "C" int slam_set_pregate(slam_t *h, int on) {
    ARG_CHECK(h != nullptr, "NULL handle");
    h->pregate = on != 0;
    return SLAM_OK;
}

extern "C" int slam_gate_stats(slam_t *h, long long *scanned, long long *scored, double *sig2max) {
    H_CHECK(h);
    unsigned long long v[3];
    CUDA_TRY(cudaMemcpyAsync(v, h->d_sig2, sizeof(v), cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    if (sig2max) memcpy(sig2max, &v[0], sizeof(double));
    if (scanned) *scanned = (long long)v[1];
    if (scored) *scored = (long long)v[2];
    return SLAM_OK;
}

extern "C" int slam_compute_association(slam_t *h, const double z[2], const double R[4], int64_t idf,
                                        double *nis, double *nd) {
    H_CHECK(h);
    ARG_CHECK(z && R && nis && nd, "NULL argument");
    int rc = launch_compute_association(h->d_x, h->d_P, h->ld, h->n, z, R, idf, h->d_small, h->d_err,
                                        h->stream, &h->launches);
    if (rc) return rc;
    double *ho = (double *)h->h_pin + 16;
    CUDA_TRY(cudaMemcpyAsync(ho, h->d_small, sizeof(double) * 2, cudaMemcpyDeviceToHost, h->stream));
    rc = check_latched(h);
    if (rc) return rc;
    *nis = ho[0];
    *nd = ho[1];
    return SLAM_OK;
}

__global__ void k_predict_observation(const double *x, int n, int64_t idf, double *out, int *err) {
    int Nf = (n - 3) / 2;
    if (idf < 1 || idf > Nf) {
        atomicCAS(err, 0, SLAM_ERR_BAD_INDEX);
        return;
    }
    int f = 3 + 2 * ((int)idf - 1);
    observe_model(x[0], x[1], x[2], x[f], x[f + 1], out, out + 2);
}

extern "C" int slam_predict_observation(slam_t *h, int64_t idf, double zp[2], double *H) {
    H_CHECK(h);
    ARG_CHECK(zp != nullptr, "NULL zp");
    k_predict_observation<<<1, 1, 0, h->stream>>>(h->d_x, h->n, idf, h->d_small, h->d_err);
    h->launches++;
    CUDA_TRY(cudaGetLastError());
    double *ho = (double *)h->h_pin + 16;
    CUDA_TRY(cudaMemcpyAsync(ho, h->d_small, sizeof(double) * 12, cudaMemcpyDeviceToHost, h->stream));
    int rc = check_latched(h);
    if (rc) return rc;
    zp[0] = ho[0];
    zp[1] = ho[1];
    if (H) {   // dense 2 x n, column-major (leading dimension 2), zeros elsewhere (src/common.jl:160-162)
        const double *hn = ho + 2;
        int n = h->n, f = 3 + 2 * ((int)idf - 1);
        memset(H, 0, sizeof(double) * 2 * n);
        for (int c = 0; c < 3; c++) {
            H[2 * c] = hn[c];
            H[2 * c + 1] = hn[3 + c];
        }
        H[2 * f] = hn[6];     H[2 * f + 1] = hn[8];
        H[2 * (f + 1)] = hn[7]; H[2 * (f + 1) + 1] = hn[9];
    }
    return SLAM_OK;
}

// ---------------------------------------------------------------------------------------
// K3 -- PHt = P*H' (src/ekf.jl:67) as a gather of 5 columns per observation, plus the
// innovation v (:59).  Thread = row i.  PHt is stored k-major: PHt[k*npad + i].
// Bytes: 8n(3 + 2m) read (5 columns per observation, the 3 pose columns shared), 16nm written.
// ---------------------------------------------------------------------------------------
constexpr int K3_THREADS = 256;
constexpr int K3_OT = 64;   // observations staged per shared-memory tile

__global__ void __launch_bounds__(K3_THREADS)
k_pht(const double *__restrict__ x, const double *__restrict__ P, size_t ld, int n, int npad,
      const double *__restrict__ zf, const int64_t *__restrict__ idf, int m, double *__restrict__ PHt,
      double *__restrict__ v_out, double *__restrict__ Hn_out, int *__restrict__ fp_out, int *err) {
    __shared__ double sH[K3_OT][10];
    __shared__ int sF[K3_OT];
    const int i = blockIdx.x * K3_THREADS + threadIdx.x;
    const int Nf = (n - 3) / 2;
    const double xv = x[0], yv = x[1], phi = x[2];
    double p0 = 0, p1 = 0, p2 = 0;
    if (i < n) {
        p0 = sym_at(P, ld, i, 0);
        p1 = sym_at(P, ld, i, 1);
        p2 = sym_at(P, ld, i, 2);
    }
    for (int ot = 0; ot < m; ot += K3_OT) {
        int cnt = min(K3_OT, m - ot);
        __syncthreads();
        if (threadIdx.x < cnt) {
            int o = ot + threadIdx.x;
            int64_t id = idf[o];
            if (id < 1 || id > Nf) {
                atomicCAS(err, 0, SLAM_ERR_BAD_INDEX);
                id = 1;
            }
            int f = 3 + 2 * ((int)id - 1);
            double zp[2];
            observe_model(xv, yv, phi, x[f], x[f + 1], zp, sH[threadIdx.x]);
            sF[threadIdx.x] = f;
            if (blockIdx.x == 0) {
                v_out[2 * o] = zf[2 * o] - zp[0];
                v_out[2 * o + 1] = mpi_to_pi(zf[2 * o + 1] - zp[1]);
                fp_out[o] = f;
                for (int q = 0; q < 10; q++) Hn_out[10 * o + q] = sH[threadIdx.x][q];
            }
        }
        __syncthreads();
        if (i < n) {
            for (int t = 0; t < cnt; t++) {
                const double *hh = sH[t];
                int f = sF[t];
                double p3, p4;
                if (i <= f) {
                    p3 = P[(size_t)i + (size_t)f * ld];
                    p4 = P[(size_t)i + (size_t)(f + 1) * ld];
                } else {
                    const double *ci = P + (size_t)i * ld;
                    p3 = ci[f];
                    p4 = (i == f + 1) ? P[(size_t)i + (size_t)(f + 1) * ld] : ci[f + 1];
                }
                int o = ot + t;
                PHt[(size_t)(2 * o) * npad + i] = p0 * hh[0] + p1 * hh[1] + p2 * hh[2] + p3 * hh[6] + p4 * hh[7];
                PHt[(size_t)(2 * o + 1) * npad + i] = p0 * hh[3] + p1 * hh[4] + p2 * hh[5] + p3 * hh[8] + p4 * hh[9];
            }
        }
    }
}

// ---------------------------------------------------------------------------------------
// K4 -- S = H*PHt + RR, S = (S+S')/2, C = inv(chol(S)) (src/ekf.jl:68-70), u = C'v, for up to 32 ranks, in
// shared memory by one CTA.  k4_factor is shared by the stand-alone kernel (sharded filter) and the fused
// K3 + K4 launch (single filter), so both produce the same C bit for bit.
//   * right-looking upper Cholesky, the same subtraction order per entry as the left-looking twin; two barriers
//     per step (every thread takes sqrt(d) of the pivot itself; the pivots' roots go to sd, so the step that
//     reads the diagonal never races with one that writes it);
//   * C = inv(U) in the order of LAPACK's unblocked dtrti2 with the reciprocals taken up front, thread i building
//     row i: C[i][j] = (U[i][j] C[i][i] + sum_{i<jj<j} U[jj][j] C[i][jj]) * (-1/U[j][j]) -- the order of the
//     oracle's twin and of the ensemble kernel;  u[j] = sum_{l<j} C[l][j] v[l] + C[j][j] v[j].
// ---------------------------------------------------------------------------------------
constexpr int K4_RMAX = 32;
constexpr int K4_THREADS = 256;

#ifdef K4_TIMING   // developer build (SLAMCUDA_DEFS=-DK4_TIMING): cycles per phase of the CTA that runs K4, printed by its thread 0
__device__ long long g_k4_mark[16];
#define K4_MARK(k) do { if (threadIdx.x == 0) g_k4_mark[k] = clock64(); } while (0)
#else
#define K4_MARK(k)
#endif

// sS: R x R column-major (leading dimension R) holding S; sC: R x R; sv: R; sd: 2R scratch.  Returns false (all
// threads alike) when S is not positive definite.  Ends with a barrier; C and u are written to global memory.
// nt is a multiple of 32 and at least 32 (both callers launch 256 threads); R <= 32.
//
// Warp-synchronous form (r2): the block-wide right-looking loop spent two __syncthreads, an integer division and
// stride-R (16- to 32-way bank-conflicted) accesses per step, and the inverse was one thread per row walking
// R^2/2 dependent shared-memory round trips: +8.6 / +19.7 / +36.5 us per update at R = 16 / 24 / 32
// (tools/k4_probe.py).  Now
//   * Cholesky: warp 0 alone, row by row (the twin's left-looking order: U[i][j] = (S[i][j] - sum_{k<i}
//     U[k][i] U[k][j]) / U[i][i], k ascending -- per entry the same multiply / subtract sequence as the right-looking
//     loop; this file is built with -fmad=false, so no explicit fma() here: it would round differently), lane j
//     owns column j.  U[i][j] overwrites S at sS[j + i*R], the transposed position (S is symmetric there), so every
//     lane access is contiguous; one __syncwarp per row, no block barrier;
//   * C = inv(U): one WARP per row i, lane j accumulates C[i][j].  As soon as C[i][jj] is final it is broadcast by
//     shuffle and every lane j > jj adds U[jj][j] C[i][jj] -- jj ascending per entry, dtrti2's order exactly -- so a
//     row costs R - i dependent steps instead of (R - i)^2 / 2, and the rows run side by side on the warps.
// Measured on one box against the -DK4_LEGACY build (which keeps the block-wide body for exactly that), per update:
// 8.4 -> 7.0 us at R = 16, 19.9 -> 17.5 at R = 24, 37.6 -> 31.4 at R = 32 on a tiny map; 26.9 -> 25.6 us (m = 8) and
// 58.5 -> 52.8 us (m = 16) of exposed small-kernel time at n = 20003 -- a modest gain: most of the rank-dependent
// latency of an update is outside this function.  Same bits as the legacy body (digests of x and P after updates at
// R = 2..32, tools/k4_hash.py) and as the twin (test_update_P_bitwise_vs_twin).
#ifndef K4_LEGACY
__device__ __forceinline__ bool k4_factor(double *sS, double *sC, const double *sv, double *sd, int R, int tid, int nt,
                                          double *__restrict__ C, double *__restrict__ u) {
    constexpr unsigned FULL = 0xffffffffu;
    for (int idx = tid; idx < R * R; idx += nt) {                  // S = (S + S')*0.5
        const int a = idx % R, b = idx / R;
        if (a < b) {
            const double s2 = (sS[a + b * R] + sS[b + a * R]) * 0.5;
            sS[a + b * R] = s2;
            sS[b + a * R] = s2;
        } else if (a == b) {
            sS[a + a * R] = (sS[a + a * R] + sS[a + a * R]) * 0.5;
        }
    }
    __syncthreads();
    K4_MARK(4);   // symmetrised
    double *cd = sd + R;                                           // cd[j] = 1 / U[j][j]; sd[0] = "positive definite"
    const int lane = tid & 31, warp = tid >> 5, nw = nt >> 5;
#ifdef K4_RIGHT_LOOKING   // measured, NOT adopted: off by default (profiles/r2_k4_warp.md)
    if (warp == 0) {
        // Right-looking, the running sums in shared memory: entry (i, j), i <= j, lives at sS[j + i*R] (the transposed
        // position: S is symmetric there, lanes j are contiguous) and holds S[i][j] - sum_{k finished} U[k][i] U[k][j].
        // When row k is final every remaining row takes its term at once; eight rows are loaded before any is stored,
        // so the updates pipeline instead of forming one dependent chain per row (measured with the left-looking
        // loop: ~700 cycles per row fixed plus ~55 per term of its sum, tools/k4_phases.py).  Per entry the terms
        // still arrive k ascending, product and difference rounded separately: the same bits.  (The same scheme with
        // the column in REGISTERS and everything unrolled was 3x slower: ~2000 instructions executed once each.)
        // Result: the Cholesky phase itself -14 % at R = 16, -27 % at R = 32 in the -DK4_TIMING build, but whole updates
        // on one box: +0.4 / +1.2 us on a tiny map (R = 16 / 32), no change at n = 20003 with events, -4..-12 us without
        // (inside that loop's scatter).  Not a demonstrable gain, so the row-by-row loop below stays the default.
        const int j = lane;
        bool ok = true;
        for (int k = 0; k < R; k++) {
            const double d = sS[k + k * R];                        // the pivot: every lane reads it (broadcast)
            if (!(d > 0)) {                                        // uniform across the warp
                ok = false;
                break;
            }
            const double ukk = sqrt(d);                            // every lane takes the root itself
            const bool mine = j > k && j < R;
            const double ukj = mine ? sS[j + k * R] / ukk : 0.0;   // U[k][j]
            __syncwarp();                                          // the pivot has been read by all before lane k replaces it
            if (j == k) {
                sS[k + k * R] = ukk;
                cd[k] = 1.0 / ukk;
            } else if (mine) {
                sS[j + k * R] = ukj;
            }
            __syncwarp();
            for (int i0 = k + 1; i0 < R; i0 += 8) {                // rows i0 .. i0 + 7 of the trailing block
                double uki[8], a[8];
#pragma unroll
                for (int q = 0; q < 8; q++) {
                    const int ii = i0 + q;
                    const bool on = ii < R && mine && ii <= j;
                    uki[q] = ii < R ? sS[ii + k * R] : 0.0;        // U[k][ii], a broadcast read
                    a[q] = on ? sS[j + ii * R] : 0.0;
                }
#pragma unroll
                for (int q = 0; q < 8; q++) {
                    const int ii = i0 + q;
                    if (ii < R && mine && ii <= j) sS[j + ii * R] = a[q] - uki[q] * ukj;
                }
            }
            __syncwarp();
        }
#else
    if (warp == 0) {
        const int j = lane;
        bool ok = true;
        for (int i = 0; i < R; i++) {
            double s = 0.0;
            if (j >= i && j < R) {
                s = sS[j + i * R];                                 // S[i][j] (= S[j][i])
#pragma unroll 4
                for (int k = 0; k < i; k++) s -= sS[i + k * R] * sS[j + k * R];   // U[k][i] (broadcast), U[k][j]
            }
            const double d = __shfl_sync(FULL, s, i);              // the pivot, from lane i
            if (!(d > 0)) {                                        // uniform across the warp
                ok = false;
                break;
            }
            const double ukk = sqrt(d);                            // every lane takes the root itself
            if (j == i) {
                sS[i + i * R] = ukk;
                cd[i] = 1.0 / ukk;
            } else if (j > i && j < R) {
                sS[j + i * R] = s / ukk;                           // U[i][j]
            }
            __syncwarp();
        }
#endif
        if (lane == 0) sd[0] = ok ? 1.0 : 0.0;
    }
    __syncthreads();
    K4_MARK(5);   // Cholesky done
    if (sd[0] == 0.0) return false;
    for (int i = warp; i < R; i += nw) {                           // C = inv(U), row i on this warp
        const int j = lane;
        const double cii = cd[i];
        const double ncdj = j < R ? -cd[j] : 0.0;
        double y = (j > i && j < R) ? sS[j + i * R] * cii : 0.0;   // U[i][j] C[i][i]
        if (j < i) sC[i + j * R] = 0.0;                            // below the diagonal
        if (j == i) sC[i + i * R] = cii;
        for (int jj = i + 1; jj < R; jj++) {
            const double c = __shfl_sync(FULL, y * ncdj, jj);      // C[i][jj], final in lane jj
            if (j == jj) sC[i + jj * R] = c;
            if (j > jj && j < R) y += sS[j + jj * R] * c;          // + U[jj][j] C[i][jj] (separate multiply and add)
        }
    }
    __syncthreads();
    K4_MARK(6);   // inverse done
    if (tid < R) {
        const int j = tid;
        double a = 0;
        for (int l = 0; l < j; l++) a += sC[l + j * R] * sv[l];
        a += cd[j] * sv[j];
        u[j] = a;
    }
    for (int idx = tid; idx < R * R; idx += nt) C[idx] = sC[idx];
    K4_MARK(7);   // u and C written
    return true;
}
#else
__device__ __forceinline__ bool k4_factor(double *sS, double *sC, const double *sv, double *sd, int R, int tid, int nt,
                                          double *__restrict__ C, double *__restrict__ u) {
    for (int idx = tid; idx < R * R; idx += nt) {                  // S = (S + S')*0.5
        const int a = idx % R, b = idx / R;
        if (a < b) {
            const double s2 = (sS[a + b * R] + sS[b + a * R]) * 0.5;
            sS[a + b * R] = s2;
            sS[b + a * R] = s2;
        } else if (a == b) {
            sS[a + a * R] = (sS[a + a * R] + sS[a + a * R]) * 0.5;
        }
    }
    __syncthreads();
    bool ok = true;
    for (int k = 0; k < R; k++) {
        const double d = sS[k + k * R];
        if (!(d > 0)) {          // uniform: every thread reads the same pivot
            ok = false;
            break;
        }
        const double ukk = sqrt(d);
        if (tid == 0) sd[k] = ukk;
        for (int j = k + 1 + tid; j < R; j += nt) sS[k + j * R] /= ukk;
        __syncthreads();
        const int rem = R - k - 1;
        for (int idx = tid; idx < rem * rem; idx += nt) {
            const int i = k + 1 + idx % rem, j = k + 1 + idx / rem;
            if (i <= j) sS[i + j * R] -= sS[k + i * R] * sS[k + j * R];
        }
        __syncthreads();
    }
    if (!ok) return false;
    double *cd = sd + R;
    if (tid < R) cd[tid] = 1.0 / sd[tid];
    for (int idx = tid; idx < R * R; idx += nt) sC[idx] = 0.0;
    __syncthreads();
    if (tid < R) {
        const int i = tid;
        const double cii = cd[i];
        sC[i + i * R] = cii;
        for (int j = i + 1; j < R; j++) {
            double y = sS[i + j * R] * cii;
            for (int jj = i + 1; jj < j; jj++) y += sS[jj + j * R] * sC[i + jj * R];
            sC[i + j * R] = y * (-cd[j]);
        }
    }
    __syncthreads();
    if (tid < R) {
        const int j = tid;
        double a = 0;
        for (int l = 0; l < j; l++) a += sC[l + j * R] * sv[l];
        a += cd[j] * sv[j];
        u[j] = a;
    }
    for (int idx = tid; idx < R * R; idx += nt) C[idx] = sC[idx];
    return true;
}
#endif

__device__ __forceinline__ void k4_noop(int R, int tid, int nt, double *__restrict__ C, double *__restrict__ u) {
    for (int idx = tid; idx < R * R; idx += nt) C[idx] = 0.0;     // C = 0: W1 = 0, x += 0, P -= 0
    for (int k = tid; k < R; k += nt) u[k] = 0.0;
}

__global__ void __launch_bounds__(K4_THREADS)
k_innov_chol(const double *__restrict__ PHt, int npad, const double *__restrict__ Hn,
             const int *__restrict__ fp, const double *__restrict__ v, int m, double R0, double R1,
             double R2, double R3, double *__restrict__ C, double *__restrict__ u, int *err) {
    __shared__ double sS[K4_RMAX * K4_RMAX], sC[K4_RMAX * K4_RMAX], sv[K4_RMAX], sd[2 * K4_RMAX];
    const int R = 2 * m;
    const int tid = threadIdx.x;
    // A latched error (a landmark id out of range found earlier in THIS update, or an earlier failure nobody has
    // read yet) turns the update into a no-op before anything is changed.  The reference throws before any
    // mutation (BoundsError at src/ekf.jl:57, chol at :70).
    if (*reinterpret_cast<volatile int *>(err) != 0) {
        k4_noop(R, tid, K4_THREADS, C, u);
        return;
    }
    for (int idx = tid; idx < m * R; idx += K4_THREADS) {
        const int o = idx / R, k = idx % R;
        const double *h = Hn + 10 * o;
        const int f = fp[o];
        const double *col = PHt + (size_t)k * npad;
        const double a0 = col[0], a1 = col[1], a2 = col[2], a3 = col[f], a4 = col[f + 1];
        double s0 = h[0] * a0 + h[1] * a1 + h[2] * a2 + h[6] * a3 + h[7] * a4;
        double s1 = h[3] * a0 + h[4] * a1 + h[5] * a2 + h[8] * a3 + h[9] * a4;
        if (k == 2 * o) { s0 += R0; s1 += R1; }          // RR[ii,ii] = R (:60), R column-major
        if (k == 2 * o + 1) { s0 += R2; s1 += R3; }
        sS[(2 * o) + k * R] = s0;
        sS[(2 * o + 1) + k * R] = s1;
    }
    if (tid < R) sv[tid] = v[tid];
    __syncthreads();
    if (!k4_factor(sS, sC, sv, sd, R, tid, K4_THREADS, C, u)) {
        if (tid == 0) atomicCAS(err, 0, SLAM_ERR_NOT_PD);
        k4_noop(R, tid, K4_THREADS, C, u);
    }
}

// ---------------------------------------------------------------------------------------
// K5 -- W1 = PHt*C (src/ekf.jl:71) and x += W*v = W1*(C'v) (:72-74).  Thread = row i.
// W1 is written per rank chunk in both layouts K6 reads: row-major [npad][RC] (staged with
// bulk copies for the column side) and rank-major [RC][npad] (coalesced row-side loads).
// ---------------------------------------------------------------------------------------
constexpr int K5_THREADS = 128;

__global__ void __launch_bounds__(K5_THREADS)
k_gain(const double *__restrict__ PHt, int npad, int n, int R, const double *__restrict__ C,
       const double *__restrict__ u, int RC, double *__restrict__ W1c, double *__restrict__ x,
       double *pose_snapshot) {
    extern __shared__ double sC[];   // R*R when it fits, else unused
    const bool use_s = (size_t)R * R * sizeof(double) <= 32768;
    if (use_s) {
        for (int idx = threadIdx.x; idx < R * R; idx += K5_THREADS) sC[idx] = C[idx];
        __syncthreads();
    }
    const double *Cm = use_s ? sC : C;
    const int i = blockIdx.x * K5_THREADS + threadIdx.x;
    if (i >= npad) return;
    const int nch = (R + RC - 1) / RC, ldk = w1_ldk(RC);
    const size_t csz = w1_chunk_doubles(npad, RC), rmo = w1_rank_major_offset(npad, RC);
    for (int c = 0; c < nch; c++)      // the row-major rows' padding (never used by the arithmetic; kept defined)
        for (int kk = RC; kk < ldk; kk++) W1c[(size_t)c * csz + (size_t)i * ldk + kk] = 0.0;
    if (i >= n) {
        for (int c = 0; c < nch; c++)
            for (int kk = 0; kk < RC; kk++) {
                double *chunk = W1c + (size_t)c * csz;
                chunk[(size_t)i * ldk + kk] = 0.0;                  // [npad][RC + W1_PAD] row-major
                chunk[rmo + (size_t)kk * npad + i] = 0.0;           // [RC][npad] rank-major
            }
        return;
    }
    double dxi = 0;
    for (int k = 0; k < nch * RC; k++) {
        double a = 0;
        if (k < R)
            for (int l = 0; l <= k; l++) a += PHt[(size_t)l * npad + i] * Cm[(size_t)l + (size_t)k * R];
        double *chunk = W1c + (size_t)(k / RC) * csz;
        chunk[(size_t)i * ldk + (k % RC)] = a;
        chunk[rmo + (size_t)(k % RC) * npad + i] = a;
        if (k < R) dxi += a * u[k];
    }
    double xi = x[i] + dxi;
    x[i] = xi;
    if (i < 3) pose_snapshot[i] = xi;
}

// K4 for ranks above 32.  Same S, same right-looking Cholesky element by element (so U is bit-identical to
// k_innov_chol's), but laid out for a 2m x 2m problem that is no longer tiny:
//  * 1024 threads as a 32 x 32 grid over the trailing block (no index division); the square lives in
//    shared memory with an odd leading dimension, so walking a row is as conflict-free as a column;
//  * inv(U) by rows from the bottom with every (i, j) sum advancing in parallel -- row k is finished,
//    then every entry above it takes its term U[i][k]*C[k][j].  That is k_innov_chol's sum taken k
//    DESCENDING instead of ascending: R barrier-separated steps of O(R^2) parallel work instead of one
//    thread per column walking j^2/2 dependent terms.  The running sums sit transposed in the lower
//    triangle of the same square (entry (i, j), i < j, at row j, column i), the diagonal in cd.
// Rounding of C differs from the small-rank kernel at the 1e-16 level; the oracle tolerance is 1e-9.
constexpr int K4W_THREADS = 1024;

__global__ void __launch_bounds__(K4W_THREADS)
k_innov_chol_wide(const double *__restrict__ PHt, int npad, const double *__restrict__ Hn,
                  const int *__restrict__ fp, const double *__restrict__ v, int m, double R0, double R1,
                  double R2, double R3, double *Sg, double *C, double *u, int *err, int s_in_smem) {
    extern __shared__ __align__(16) double k4_smem[];
    const int R = 2 * m;
    const int ld = s_in_smem ? (R | 1) : R;          // the global workspace is R x R
    double *S = s_in_smem ? k4_smem : Sg;
    double *cd = s_in_smem ? k4_smem + (size_t)ld * R : u;   // 1/U[k][k]; u is free until the very end
    const int tid = threadIdx.x, tx = tid & 31, ty = tid >> 5;
    __shared__ int s_fail;
    if (tid == 0) s_fail = 0;
    if (*reinterpret_cast<volatile int *>(err) != 0) {   // latched error: no-op (see k_innov_chol)
        for (int idx = tid; idx < R * R; idx += K4W_THREADS) C[idx] = 0.0;
        for (int k = tid; k < R; k += K4W_THREADS) u[k] = 0.0;
        return;
    }
    for (int idx = tid; idx < m * R; idx += K4W_THREADS) {
        int o = idx / R, k = idx % R;
        const double *h = Hn + 10 * o;
        int f = fp[o];
        const double *col = PHt + (size_t)k * npad;
        double a0 = col[0], a1 = col[1], a2 = col[2], a3 = col[f], a4 = col[f + 1];
        double s0 = h[0] * a0 + h[1] * a1 + h[2] * a2 + h[6] * a3 + h[7] * a4;
        double s1 = h[3] * a0 + h[4] * a1 + h[5] * a2 + h[8] * a3 + h[9] * a4;
        if (k == 2 * o) { s0 += R0; s1 += R1; }
        if (k == 2 * o + 1) { s0 += R2; s1 += R3; }
        S[(size_t)(2 * o) + (size_t)k * ld] = s0;
        S[(size_t)(2 * o + 1) + (size_t)k * ld] = s1;
    }
    __syncthreads();
    for (int b = ty; b < R; b += 32)
        for (int a = tx; a <= b; a += 32) {
            const double sv = (S[(size_t)a + (size_t)b * ld] + S[(size_t)b + (size_t)a * ld]) * 0.5;
            S[(size_t)a + (size_t)b * ld] = sv;      // only the upper triangle is used below
        }
    __syncthreads();
    for (int k = 0; k < R; k++) {
        const double d = S[(size_t)k + (size_t)k * ld];
        if (!(d > 0)) {              // uniform: every thread reads the same d
            if (tid == 0) s_fail = 1;
            break;
        }
        const double ukk = sqrt(d);
        __syncthreads();             // everyone has read d
        if (tid == 0) S[(size_t)k + (size_t)k * ld] = ukk;
        for (int j = k + 1 + tid; j < R; j += K4W_THREADS) S[(size_t)k + (size_t)j * ld] /= ukk;
        __syncthreads();
        for (int j = k + 1 + ty; j < R; j += 32) {
            const double ukj = S[(size_t)k + (size_t)j * ld];
            for (int i = k + 1 + tx; i <= j; i += 32)
                S[(size_t)i + (size_t)j * ld] -= S[(size_t)k + (size_t)i * ld] * ukj;
        }
        __syncthreads();
    }
    __syncthreads();
    if (s_fail) {
        if (tid == 0) atomicCAS(err, 0, SLAM_ERR_NOT_PD);
        for (int idx = tid; idx < R * R; idx += K4W_THREADS) C[idx] = 0.0;   // update becomes a no-op
        for (int k = tid; k < R; k += K4W_THREADS) u[k] = 0.0;
        return;
    }
    // running sums of inv(U) start at zero: entry (i, j), i < j, at S[j + i*ld]
    for (int i = ty; i < R; i += 32)
        for (int j = i + 1 + tx; j < R; j += 32) S[(size_t)j + (size_t)i * ld] = 0.0;
    __syncthreads();
    for (int k = R - 1; k >= 0; k--) {
        const double ukk = S[(size_t)k + (size_t)k * ld];
        for (int j = k + tid; j < R; j += K4W_THREADS) {
            if (j == k) {
                cd[k] = 1.0 / ukk;
            } else {
                const size_t e = (size_t)j + (size_t)k * ld;
                S[e] = -S[e] / ukk;
            }
        }
        __syncthreads();
        const double ckk = cd[k];
        for (int i = ty; i < k; i += 32) {
            const double uik = S[(size_t)i + (size_t)k * ld];
            for (int j = k + tx; j < R; j += 32) {
                const double ckj = j == k ? ckk : S[(size_t)j + (size_t)k * ld];
                S[(size_t)j + (size_t)i * ld] += uik * ckj;
            }
        }
        __syncthreads();
    }
    // u = C'v (l ascending, as k_innov_chol), then C out in the layout K5 reads: upper, zeros below
    double uj = 0;
    const int j1 = tid;
    if (j1 < R) {
        for (int l = 0; l < j1; l++) uj += S[(size_t)j1 + (size_t)l * ld] * v[l];
        uj += cd[j1] * v[j1];
    }
    for (int j = ty; j < R; j += 32)
        for (int i = tx; i < R; i += 32)
            C[(size_t)i + (size_t)j * R] = i < j ? S[(size_t)j + (size_t)i * ld] : (i == j ? cd[i] : 0.0);
    __syncthreads();                 // cd may alias u: everyone has read it
    if (j1 < R) u[j1] = uj;
}

int launch_innov_chol(const double *PHt, int npad, const double *Hn, const int *fp, const double *v, int m,
                      const double R[4], double *S, double *C, double *u, int *err, cudaStream_t s) {
    if (2 * m > K4_RMAX) {
        const int Rr = 2 * m;
        const size_t needw = sizeof(double) * ((size_t)(Rr | 1) * Rr + Rr);
        int in_smem = needw <= 48 * 1024;
        if (!in_smem && needw <= 200 * 1024) {   // opt in per call: the attribute belongs to the current device
            if (cudaFuncSetAttribute(k_innov_chol_wide, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)needw) ==
                cudaSuccess)
                in_smem = 1;
            else
                cudaGetLastError();
        }
        k_innov_chol_wide<<<1, K4W_THREADS, in_smem ? needw : 0, s>>>(PHt, npad, Hn, fp, v, m, R[0], R[1], R[2], R[3], S, C,
                                                                      u, err, in_smem);
    } else {
        k_innov_chol<<<1, K4_THREADS, 0, s>>>(PHt, npad, Hn, fp, v, m, R[0], R[1], R[2], R[3], C, u, err);
    }
    CUDA_TRY(cudaGetLastError());
    return SLAM_OK;
}

// K5 for ranks above 32 (the DMMA range): the one-row-per-thread loop above re-reads a row of PHt
// R/2 times and C no longer fits shared memory.  Here a CTA owns 128 rows x one 32-column chunk of W1
// (= one K6 rank chunk): thread = row, 32 accumulators, C staged 32 x 32 at a time in shared memory
// and read as warp-wide broadcasts, PHt read once per chunk.  Each W1[i, k] is still the sum over
// l = 0..k in ascending order with separate multiply and add, i.e. bit-identical to k_gain.
// x += W1*u needs all chunks of a row in order and follows in k_gain_x.
constexpr int K5T_ROWS = 128;
constexpr int K5T_KC = 32;

__global__ void __launch_bounds__(K5T_ROWS)
k_gain_tiled(const double *__restrict__ PHt, int npad, int n, int R, int RC, const double *__restrict__ C,
             double *__restrict__ W1c) {
    __shared__ __align__(16) double sC[K5T_KC * K5T_KC];   // sC[ll][kk] = C[l0 + ll][k0 + kk], 0 where l > k
    const int i = blockIdx.x * K5T_ROWS + threadIdx.x;
    const int k0 = blockIdx.y * K5T_KC;
    const bool live = i < n;
    double acc[K5T_KC];
#pragma unroll
    for (int kk = 0; kk < K5T_KC; kk++) acc[kk] = 0.0;
    const int lend = min(k0 + K5T_KC, R);          // l runs to the chunk's last column
    for (int l0 = 0; l0 < lend; l0 += K5T_KC) {
        __syncthreads();
        for (int e = threadIdx.x; e < K5T_KC * K5T_KC; e += K5T_ROWS) {
            const int ll = e % K5T_KC, kk = e / K5T_KC;      // consecutive threads walk l: C is column-major
            const int l = l0 + ll, k = k0 + kk;
            sC[ll * K5T_KC + kk] = (l <= k && k < R) ? C[(size_t)l + (size_t)k * R] : 0.0;
        }
        __syncthreads();
        if (!live) continue;
        const int lcnt = min(K5T_KC, lend - l0);
        if (l0 < k0) {
            // every l of this block precedes every column of the chunk: no guards
#pragma unroll 2
            for (int ll = 0; ll < lcnt; ll++) {
                const double p = PHt[(size_t)(l0 + ll) * npad + i];
                const double4 *c4 = reinterpret_cast<const double4 *>(sC + ll * K5T_KC);
#pragma unroll
                for (int q = 0; q < K5T_KC / 4; q++) {
                    const double4 c = c4[q];
                    acc[4 * q + 0] += p * c.x;
                    acc[4 * q + 1] += p * c.y;
                    acc[4 * q + 2] += p * c.z;
                    acc[4 * q + 3] += p * c.w;
                }
            }
        } else {
            // the diagonal block: column kk takes l = l0 .. l0 + kk only (a skipped term is not the
            // same as adding p*0 when p is not finite)
            for (int ll = 0; ll < lcnt; ll++) {
                const double p = PHt[(size_t)(l0 + ll) * npad + i];
#pragma unroll
                for (int kk = 0; kk < K5T_KC; kk++)
                    if (kk >= ll) acc[kk] += p * sC[ll * K5T_KC + kk];
            }
        }
    }
    if (i >= npad) return;
    // the CTA's 32 columns k0 .. k0 + 31 inside the RC-wide chunk layout (RC a multiple of 4: a group of four
    // columns never straddles two chunks); columns past the padded rank nch*RC do not exist
    const int ldk = w1_ldk(RC), kend = ((R + RC - 1) / RC) * RC;
    const size_t csz = w1_chunk_doubles(npad, RC), rmo = w1_rank_major_offset(npad, RC);
#pragma unroll
    for (int q = 0; q < K5T_KC / 4; q++) {
        const int k = k0 + 4 * q;
        if (k >= kend) break;
        double *chunk = W1c + (size_t)(k / RC) * csz;
        *reinterpret_cast<double4 *>(chunk + (size_t)i * ldk + (k % RC)) =          // [npad][RC + W1_PAD] row-major
            live ? make_double4(acc[4 * q], acc[4 * q + 1], acc[4 * q + 2], acc[4 * q + 3]) : make_double4(0.0, 0.0, 0.0, 0.0);
#pragma unroll
        for (int e = 0; e < 4; e++)                                                 // [RC][npad] rank-major
            chunk[rmo + (size_t)(k % RC + e) * npad + i] = live ? acc[4 * q + e] : 0.0;
        if ((k % RC) + 4 == RC)                                                     // the row's padding, kept defined
            *reinterpret_cast<double4 *>(chunk + (size_t)i * ldk + RC) = make_double4(0.0, 0.0, 0.0, 0.0);
    }
}

// x += W1*u (src/ekf.jl:72-74), k ascending over all chunks as k_gain does
__global__ void __launch_bounds__(128)
k_gain_x(const double *__restrict__ W1c, int npad, int n, int R, int RC, const double *__restrict__ u,
         double *__restrict__ x, double *pose_snapshot) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const size_t csz = w1_chunk_doubles(npad, RC), rmo = w1_rank_major_offset(npad, RC);
    double dxi = 0;
    for (int k = 0; k < R; k++) {
        const double *chunk = W1c + (size_t)(k / RC) * csz;
        dxi += chunk[rmo + (size_t)(k % RC) * npad + i] * u[k];
    }
    const double xi = x[i] + dxi;
    x[i] = xi;
    if (i < 3) pose_snapshot[i] = xi;
}

// ---------------------------------------------------------------------------------------
// K3 + K4 in one launch, for up to 32 ranks (the HBM-bound regime where the small kernels' latency shows: at
// n = 20003 they were 54 us of a 571 us step).  S = H*P*H' + RR depends on P only through the (3 + 2m)^2
// sub-block of the rows and columns H touches, so the innovation Cholesky does not have to wait for P*H':
// the LAST CTA of the launch gathers that sub-block itself -- PHt at the 3 + 2m rows named by H, with K3's own
// expression, so S is bit-identical to what K4 formed from K3's output -- and runs K4 (S, symmetrise,
// right-looking Cholesky, C = inv(U), u = C'v) entirely in shared memory while the other CTAs stream P*H'.
// inv(U) column j is built by thread j from shared memory (k_innov_chol kept C in global memory: ~120
// dependent L2 round trips for the last column, 20 us for a 16 x 16 problem).
// ---------------------------------------------------------------------------------------
constexpr int K34_RMAX = 32;
constexpr int K34_MMAX = K34_RMAX / 2;

__global__ void __launch_bounds__(K3_THREADS)
k_pht_chol(const double *__restrict__ x, const double *__restrict__ P, size_t ld, int n, int npad,
           const double *__restrict__ zf, const int64_t *__restrict__ idf, int m, double *__restrict__ PHt,
           double *__restrict__ v_out, double *__restrict__ Hn_out, int *__restrict__ fp_out, int *err, double R0,
           double R1, double R2, double R3, double *__restrict__ C, double *__restrict__ u) {
    __shared__ double sH[K34_MMAX][10];
    __shared__ int sF[K34_MMAX];
    __shared__ double sPH[K34_RMAX][K34_RMAX + 3];     // chol CTA: PHt[k][s] at the rows cs[s] named by H
    __shared__ double sS[K34_RMAX * K34_RMAX], sC[K34_RMAX * K34_RMAX], sv[K34_RMAX], sd[2 * K34_RMAX];
    __shared__ int s_fail;
#ifdef NO_PDL_PREFETCH
    pdl_wait();
    pdl_launch_dependents();
#endif
    // Before the wait: x, the ids and the observation model.  Launched behind a K6 that is still running, the only
    // thing this kernel must wait for is P; x was finalised by the previous update's K5 (complete before that K6
    // passed its own wait), the status word by its K3 + K4, and any other predecessor (predict, add_features, the
    // gate, a copy) serialises fully.  The dependent global round trips x -> id -> x[f] and the sqrt / atan2 of the
    // model so overlap K6's tail: -1.5 us per update on a tiny map (13.4 -> 11.9), -2.0 us at n = 20003 (531.7 -> 529.7,
    // same box against -DNO_PDL_PREFETCH, profiles/r2_k4_warp.md), same bits.  (Moving the downdate's first P loads ahead
    // of ITS wait was tried too and reverted: it put every CTA's P loads in front of its W1 copy, K6 +2..15 us.)
    const bool chol = blockIdx.x == gridDim.x - 1;
#ifdef K4_TIMING
    if (chol) K4_MARK(0);   // entry
#endif
    const int tid = threadIdx.x;
    const int Nf = (n - 3) / 2;
    const int R = 2 * m;
    const double xv = x[0], yv = x[1], phi = x[2];
    const int latched = *reinterpret_cast<volatile int *>(err);   // an earlier failure nobody has read yet: no-op
    if (tid == 0) s_fail = 0;
    __syncthreads();
    if (tid < m) {
        int64_t id = idf[tid];
        if (id < 1 || id > Nf) {
            atomicCAS(err, 0, SLAM_ERR_BAD_INDEX);
            s_fail = 2;
            id = 1;
        }
        const int f = 3 + 2 * ((int)id - 1);
        double zp[2];
        observe_model(xv, yv, phi, x[f], x[f + 1], zp, sH[tid]);
        sF[tid] = f;
        if (chol) {
            const double v0 = zf[2 * tid] - zp[0], v1 = mpi_to_pi(zf[2 * tid + 1] - zp[1]);
            sv[2 * tid] = v0;
            sv[2 * tid + 1] = v1;
            v_out[2 * tid] = v0;
            v_out[2 * tid + 1] = v1;
            fp_out[tid] = f;
            for (int q = 0; q < 10; q++) Hn_out[10 * tid + q] = sH[tid][q];
        }
    }
    __syncthreads();
#if defined(GATHER_PREFETCH) && !defined(NO_PDL_PREFETCH)   // measured, NOT adopted: off by default (profiles/r2_k4_warp.md)
    // The K4 CTA's gather touches P at rows / columns {0, 1, 2, f_o, f_o + 1}: 3 + 2m columns on as many 2 MB pages
    // that the downdate has just streamed past -- measured 7.8 k cycles for ONE batch of five loads per thread at
    // R = 16 (-DK4_TIMING, tools/k4_phases.py), i.e. serialised TLB / HBM misses.  The addresses are known here, ~2 us
    // before the wait ends: prefetch them into L2.  L2 is the point of coherence, so a line the running downdate still
    // writes is current when the loads behind the wait read it (nothing is brought into L1).  The gather phase fell to
    // 3-6 k cycles, but same-box A/B of whole updates showed no gain (+0.1 us on a tiny map, nothing at n = 20003).
    if (chol && latched == 0 && s_fail != 2) {
        const int SCp = 3 + R;
        for (int idx = tid; idx < m * SCp; idx += K3_THREADS) {
            const int o = idx / SCp, sx = idx % SCp;
            const int i = sx < 3 ? sx : sF[(sx - 3) >> 1] + ((sx - 3) & 1);
            const int f = sF[o];
            const int cols[5] = {0, 1, 2, f, f + 1};
#pragma unroll
            for (int q = 0; q < 5; q++) {
                const int c = cols[q];
                const double *pa = i <= c ? P + (size_t)i + (size_t)c * ld : P + (size_t)c + (size_t)i * ld;
                asm volatile("prefetch.global.L2 [%0];" ::"l"(pa));
            }
        }
    }
#endif
#ifdef K4_TIMING
    if (chol) K4_MARK(1);   // ids, x, observation model done, gather prefetched
#endif
#ifndef NO_PDL_PREFETCH
    pdl_wait();                          // P (the previous update's downdate) is complete
    pdl_launch_dependents();
#endif
#ifdef K4_TIMING
    if (chol) K4_MARK(2);   // wait for the previous downdate over
#endif
    if (!chol) {
        const int i = blockIdx.x * K3_THREADS + tid;
        if (i >= n) return;
        const double p0 = sym_at(P, ld, i, 0), p1 = sym_at(P, ld, i, 1), p2 = sym_at(P, ld, i, 2);
        for (int o = 0; o < m; o++) {
            const double *hh = sH[o];
            const int f = sF[o];
            const double p3 = sym_at(P, ld, i, f), p4 = sym_at(P, ld, i, f + 1);
            PHt[(size_t)(2 * o) * npad + i] = p0 * hh[0] + p1 * hh[1] + p2 * hh[2] + p3 * hh[6] + p4 * hh[7];
            PHt[(size_t)(2 * o + 1) * npad + i] = p0 * hh[3] + p1 * hh[4] + p2 * hh[5] + p3 * hh[8] + p4 * hh[9];
        }
        return;
    }
    // ---- the Cholesky CTA ----
    if (latched != 0 || s_fail == 2) {   // bad id in this update, or a latched error: C = 0 makes K5 / K6 a no-op
        k4_noop(R, tid, K3_THREADS, C, u);
        return;
    }
    const int SC = 3 + R;                // rows named by H: pose rows, then (f_o, f_o + 1) per observation
    for (int idx = tid; idx < m * SC; idx += K3_THREADS) {
        const int o = idx / SC, s = idx % SC;
        const int i = s < 3 ? s : sF[(s - 3) >> 1] + ((s - 3) & 1);
        const double *hh = sH[o];
        const int f = sF[o];
        const double p0 = sym_at(P, ld, i, 0), p1 = sym_at(P, ld, i, 1), p2 = sym_at(P, ld, i, 2),
                     p3 = sym_at(P, ld, i, f), p4 = sym_at(P, ld, i, f + 1);
        sPH[2 * o][s] = p0 * hh[0] + p1 * hh[1] + p2 * hh[2] + p3 * hh[6] + p4 * hh[7];
        sPH[2 * o + 1][s] = p0 * hh[3] + p1 * hh[4] + p2 * hh[5] + p3 * hh[8] + p4 * hh[9];
    }
    __syncthreads();
    K4_MARK(3);   // sub-block of P gathered
    for (int idx = tid; idx < m * R; idx += K3_THREADS) {          // S = H*PHt + RR, as k_innov_chol
        const int o = idx / R, k = idx % R;
        const double *h = sH[o];
        const double *col = sPH[k];
        const double a0 = col[0], a1 = col[1], a2 = col[2], a3 = col[3 + 2 * o], a4 = col[3 + 2 * o + 1];
        double s0 = h[0] * a0 + h[1] * a1 + h[2] * a2 + h[6] * a3 + h[7] * a4;
        double s1 = h[3] * a0 + h[4] * a1 + h[5] * a2 + h[8] * a3 + h[9] * a4;
        if (k == 2 * o) { s0 += R0; s1 += R1; }
        if (k == 2 * o + 1) { s0 += R2; s1 += R3; }
        sS[(2 * o) + k * R] = s0;
        sS[(2 * o + 1) + k * R] = s1;
    }
    __syncthreads();
    if (!k4_factor(sS, sC, sv, sd, R, tid, K3_THREADS, C, u)) {
        if (tid == 0) atomicCAS(err, 0, SLAM_ERR_NOT_PD);
        k4_noop(R, tid, K3_THREADS, C, u);
    }
#ifdef K4_TIMING
    if (tid == 0) {
        const long long *t = g_k4_mark;
        printf("K4_TIMING R=%d n=%d cycles: model %lld | wait %lld | gather %lld | S+symm %lld | chol %lld | inverse %lld | u+C out %lld | total %lld\n",
               R, n, t[1] - t[0], t[2] - t[1], t[3] - t[2], t[4] - t[3], t[5] - t[4], t[6] - t[5], t[7] - t[6], t[7] - t[0]);
    }
#endif
}

// K5 for one rank chunk (R <= RC <= 32): a thread owns row i, keeps PHt[i, 0:R] in registers (the loop above
// re-read it R/2 times) and forms W1[i, k] = sum_{l <= k} PHt[i, l] C[l, k], l ascending, separate multiply and add
// -- the sums of k_gain, bit for bit -- then writes both K6 layouts and x += W1*u.
template <int RC>
__global__ void __launch_bounds__(K5_THREADS)
k_gain_reg(const double *__restrict__ PHt, int npad, int n, int R, const double *__restrict__ C,
           const double *__restrict__ u, double *__restrict__ W1c, double *__restrict__ x, double *pose_snapshot) {
    __shared__ double sCm[RC * RC], su[RC];           // sCm[l + k*RC] = C[l][k]
    constexpr int LDK = RC + W1_PAD;
    pdl_wait();
    pdl_launch_dependents();
    for (int idx = threadIdx.x; idx < RC * RC; idx += K5_THREADS) {
        const int l = idx % RC, k = idx / RC;
        sCm[idx] = (l < R && k < R) ? C[(size_t)l + (size_t)k * R] : 0.0;
    }
    if (threadIdx.x < RC) su[threadIdx.x] = threadIdx.x < R ? u[threadIdx.x] : 0.0;
    __syncthreads();
    const int i = blockIdx.x * K5_THREADS + threadIdx.x;
    if (i >= npad) return;
    double *row = W1c + (size_t)i * LDK;                            // [npad][RC + W1_PAD] row-major
    double *rk = W1c + w1_rank_major_offset(npad, RC) + i;          // [RC][npad] rank-major
    const bool live = i < n;
    double p[RC];
#pragma unroll
    for (int l = 0; l < RC; l++) p[l] = (live && l < R) ? PHt[(size_t)l * npad + i] : 0.0;
    double dxi = 0;
#pragma unroll
    for (int k = 0; k < RC; k += 2) {
        double a[2];
#pragma unroll
        for (int e = 0; e < 2; e++) {
            double acc = 0;
            if (k + e < R) {
#pragma unroll
                for (int l = 0; l <= k + e; l++) acc += p[l] * sCm[l + (k + e) * RC];
                dxi += acc * su[k + e];
            }
            a[e] = acc;
            rk[(size_t)(k + e) * npad] = acc;
        }
        *reinterpret_cast<double2 *>(row + k) = make_double2(a[0], a[1]);
    }
#pragma unroll
    for (int k = RC; k < LDK; k += 2) *reinterpret_cast<double2 *>(row + k) = make_double2(0.0, 0.0);
    if (live) {
        const double xi = x[i] + dxi;
        x[i] = xi;
        if (i < 3) pose_snapshot[i] = xi;
    }
}

template <int RC>
static cudaError_t launch_gain_reg(int blocks, cudaStream_t s, const double *PHt, int npad, int n, int R, const double *C,
                                   const double *u, double *W1c, double *x, double *pose) {
    return launch_pdl(k_gain_reg<RC>, dim3(blocks), dim3(K5_THREADS), 0, s, PHt, npad, n, R, C, u, W1c, x, pose);
}

static int update_device(slam_t *h, const double *d_zf, int m, const double R[4], const int64_t *d_idf,
                         bool mark_x = false) {
    if (m == 0) return SLAM_OK;   // first observation step of every run (sim/ekfslam-sim.jl:114-117)
    if (m > SLAM_MAX_OBS) {
        slam_set_error("slam_update: m = %d exceeds SLAM_MAX_OBS = %d", m, SLAM_MAX_OBS);
        return SLAM_ERR_CAPACITY;
    }
    if (h->n <= 3) {
        slam_set_error("slam_update: %d observations but the map is empty", m);
        return SLAM_ERR_BAD_INDEX;
    }
    h->pose_valid = false;
    const int Rk = 2 * m;
    int rc = ensure_update_ws(h, Rk);
    if (rc) return rc;
    const int n = h->n, npad = h->ws_npad;
    const int RC = downdate_chunk_width(Rk);
    const int npad_used = round_up(n, DD_ROW_PAD);
    if (Rk <= K34_RMAX) {
        // up to 32 ranks: K3 + K4 in one launch, K5 with the row of PHt in registers, chained by programmatic
        // dependent launch (each kernel is scheduled while its predecessor drains)
        CUDA_TRY(launch_pdl(k_pht_chol, dim3((n + K3_THREADS - 1) / K3_THREADS + 1), dim3(K3_THREADS), 0, h->stream,
                            (const double *)h->d_x, (const double *)h->d_P, h->ld, n, npad, d_zf, d_idf, m, h->d_PHt,
                            h->d_v, h->d_Hn, h->d_fp, h->d_err, R[0], R[1], R[2], R[3], h->d_C, h->d_u));
        const int blocks = (npad_used + K5_THREADS - 1) / K5_THREADS;
        double *pose = h->d_pose + 4 * h->pose_slot;
        cudaError_t e;
        switch (RC) {
            case 2: e = launch_gain_reg<2>(blocks, h->stream, h->d_PHt, npad, n, Rk, h->d_C, h->d_u, h->d_W1c, h->d_x, pose); break;
            case 4: e = launch_gain_reg<4>(blocks, h->stream, h->d_PHt, npad, n, Rk, h->d_C, h->d_u, h->d_W1c, h->d_x, pose); break;
            case 8: e = launch_gain_reg<8>(blocks, h->stream, h->d_PHt, npad, n, Rk, h->d_C, h->d_u, h->d_W1c, h->d_x, pose); break;
            case 12: e = launch_gain_reg<12>(blocks, h->stream, h->d_PHt, npad, n, Rk, h->d_C, h->d_u, h->d_W1c, h->d_x, pose); break;
            case 16: e = launch_gain_reg<16>(blocks, h->stream, h->d_PHt, npad, n, Rk, h->d_C, h->d_u, h->d_W1c, h->d_x, pose); break;
            case 24: e = launch_gain_reg<24>(blocks, h->stream, h->d_PHt, npad, n, Rk, h->d_C, h->d_u, h->d_W1c, h->d_x, pose); break;
            default: e = launch_gain_reg<32>(blocks, h->stream, h->d_PHt, npad, n, Rk, h->d_C, h->d_u, h->d_W1c, h->d_x, pose); break;
        }
        CUDA_TRY(e);
        h->launches += 2;
    } else {
        k_pht<<<(n + K3_THREADS - 1) / K3_THREADS, K3_THREADS, 0, h->stream>>>(
            h->d_x, h->d_P, h->ld, n, npad, d_zf, d_idf, m, h->d_PHt, h->d_v, h->d_Hn, h->d_fp, h->d_err);
        rc = launch_innov_chol(h->d_PHt, npad, h->d_Hn, h->d_fp, h->d_v, m, R, h->d_S, h->d_C, h->d_u, h->d_err, h->stream);
        if (rc) return rc;
        const int Rpad = round_up(Rk, RC);   // RC is 40..64 or (above 64 ranks) 32
        k_gain_tiled<<<dim3((npad_used + K5T_ROWS - 1) / K5T_ROWS, (Rpad + K5T_KC - 1) / K5T_KC), K5T_ROWS, 0, h->stream>>>(
            h->d_PHt, npad, n, Rk, RC, h->d_C, h->d_W1c);
        k_gain_x<<<(n + 127) / 128, 128, 0, h->stream>>>(h->d_W1c, npad, n, Rk, RC, h->d_u, h->d_x,
                                                        h->d_pose + 4 * h->pose_slot);
        h->launches += 4;
    }
    CUDA_TRY(cudaGetLastError());
    if (mark_x) CUDA_TRY(cudaEventRecord(h->ev_x, h->stream));   // status and x are final here; only P is still to come
    const bool timed = h->profiling && h->ev_used + 2 <= (int)h->ev.size();
    if (timed) CUDA_TRY(cudaEventRecord(h->ev[h->ev_used], h->stream));
    h->last_k6 = downdate_kernel_choice(Rk, h->variant, h->ld);
    rc = launch_downdate(h->d_P, h->ld, n, h->d_W1c, npad, Rk, h->variant, h->stream, &h->launches);
    if (timed) {
        CUDA_TRY(cudaEventRecord(h->ev[h->ev_used + 1], h->stream));
        h->ev_used += 2;
    }
    return rc;
}

extern "C" int slam_profile(slam_t *h, int on) {
    H_CHECK(h);
    if (on && h->ev.empty()) {
        h->ev.resize(2 * 4096);
        for (auto &e : h->ev) CUDA_TRY(cudaEventCreate(&e));
    }
    h->profiling = on != 0;
    h->ev_used = 0;
    return SLAM_OK;
}

extern "C" int slam_profile_read(slam_t *h, double *k6_ms, int *k6_updates) {
    H_CHECK(h);
    ARG_CHECK(k6_ms && k6_updates, "NULL argument");
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    double tot = 0;
    for (int i = 0; i + 1 < h->ev_used; i += 2) {
        float ms = 0;
        CUDA_TRY(cudaEventElapsedTime(&ms, h->ev[i], h->ev[i + 1]));
        tot += ms;
    }
    *k6_ms = tot;
    *k6_updates = h->ev_used / 2;
    h->ev_used = 0;
    return SLAM_OK;
}

extern "C" int slam_update_resident(slam_t *h, const double *d_zf, int m, const double R[4],
                                    const int64_t *d_idf) {
    SLAM_NVTX();
    H_CHECK(h);
    ARG_CHECK(m >= 0 && R, "bad argument");
    ARG_CHECK(m == 0 || (d_zf && d_idf), "NULL device buffer");
    return update_device(h, d_zf, m, R, d_idf);
}

extern "C" int slam_update(slam_t *h, const double *zf, int m, const double R[4], const int64_t *idf) {
    SLAM_NVTX();
    H_CHECK(h);
    ARG_CHECK(m >= 0 && R, "bad argument");
    if (m == 0) return SLAM_OK;
    ARG_CHECK(zf && idf, "NULL observation buffer");
    if (m > SLAM_MAX_OBS) {
        slam_set_error("slam_update: m = %d exceeds SLAM_MAX_OBS = %d", m, SLAM_MAX_OBS);
        return SLAM_ERR_CAPACITY;
    }
    // the ids are host data here: refuse a bad one before anything is launched (the reference's
    // BoundsError at src/ekf.jl:57 precedes every mutation)
    const int Nf = (h->n - 3) / 2;
    for (int o = 0; o < m; o++)
        if (idf[o] < 1 || idf[o] > Nf) {
            slam_set_error("slam_update: idf[%d] = %lld outside 1..%d", o, (long long)idf[o], Nf);
            return SLAM_ERR_BAD_INDEX;
        }
    double *hz = (double *)((char *)h->h_pin + 256);
    int64_t *hid = (int64_t *)(hz + 2 * m);       // packed behind zf: one H2D
    // the staging buffer is free once its last H2D is done; d_upd is only read by kernels ordered before this copy
    if (h->early_return) CUDA_TRY(cudaEventSynchronize(h->ev_stage));
    else CUDA_TRY(cudaStreamSynchronize(h->stream));
    memcpy(hz, zf, sizeof(double) * 2 * m);
    memcpy(hid, idf, sizeof(int64_t) * m);
    CUDA_TRY(cudaMemcpyAsync(h->d_upd, hz, 24 * (size_t)m, cudaMemcpyHostToDevice, h->stream));
    CUDA_TRY(cudaEventRecord(h->ev_stage, h->stream));
    int rc = update_device(h, (const double *)h->d_upd, m, R, (const int64_t *)(h->d_upd + 16 * (size_t)m),
                           h->early_return);
    if (rc) return rc;
    return h->early_return ? check_after_x(h) : check_latched(h);
}

// ---------------------------------------------------------------------------------------
// K7 -- add_features (src/ekf.jl:84-122) as an in-place strided append into spare capacity.
// A launch handles up to K7_Q new landmarks; thread (j, q) writes the two new entries of
// existing column/row j for new landmark q.  Bytes per landmark: 24n read, 16n written.
// ---------------------------------------------------------------------------------------
constexpr int K7_Q = 32;
constexpr int K7_THREADS = 256;

__global__ void __launch_bounds__(K7_THREADS)
k_augment(double *x, double *P, size_t ld, int n0, const double *__restrict__ zn, int kq, double R0,
          double R1, double R2, double R3, unsigned long long *sig2bits) {
    // n0: state length before this launch's first landmark; kq <= K7_Q landmarks
    __shared__ double sGv[K7_Q][2];     // Gv = [1 0 g0; 0 1 g1]
    __shared__ double sGP[K7_Q][2][3];  // Gv*Pvv
    __shared__ double sPvv[9];
    if (threadIdx.x < 9) sPvv[threadIdx.x] = sym_at(P, ld, threadIdx.x / 3, threadIdx.x % 3);
    __syncthreads();
    const double xv = x[0], yv = x[1], phi = x[2];
    if (threadIdx.x < kq) {
        int q = threadIdx.x;
        double r = zn[2 * q], b = zn[2 * q + 1];
        double s = sin(phi + b), c = cos(phi + b);
        double g0 = -r * s, g1 = r * c;
        sGv[q][0] = g0;
        sGv[q][1] = g1;
        const double Gv[2][3] = {{1, 0, g0}, {0, 1, g1}};
        for (int a = 0; a < 2; a++)
            for (int j = 0; j < 3; j++)
                sGP[q][a][j] = Gv[a][0] * sPvv[0 * 3 + j] + Gv[a][1] * sPvv[1 * 3 + j] + Gv[a][2] * sPvv[2 * 3 + j];
        if (blockIdx.x == 0 && blockIdx.y == q) {
            // x and the new diagonal block, written once
            int len = n0 + 2 * q;
            x[len] = xv + r * c;
            x[len + 1] = yv + r * s;
            const double Gz[2][2] = {{c, -r * s}, {s, r * c}};
            const double Rm[4] = {R0, R1, R2, R3};
            double GR[2][2];
            for (int a = 0; a < 2; a++)
                for (int j = 0; j < 2; j++) GR[a][j] = Gz[a][0] * Rm[0 + 2 * j] + Gz[a][1] * Rm[1 + 2 * j];
            for (int a = 0; a < 2; a++)
                for (int bb = a; bb < 2; bb++) {
                    double t1 = sGP[q][a][0] * Gv[bb][0] + sGP[q][a][1] * Gv[bb][1] + sGP[q][a][2] * Gv[bb][2];
                    double t2 = GR[a][0] * Gz[bb][0] + GR[a][1] * Gz[bb][1];
                    double pv = t1 + t2;
                    P[(size_t)(len + a) + (size_t)(len + bb) * ld] = pv;
                    if (a == bb) {   // keep the pre-gate's bound on the landmark variances valid (gate.cu)
                        unsigned long long bits = (unsigned long long)__double_as_longlong(pv);
                        if (pv != pv) bits = 0x7ff8000000000000ull;
                        if (!(pv < 0.0)) atomicMax(sig2bits, bits);
                    }
                }
        }
    }
    __syncthreads();
    const int q = blockIdx.y;
    if (q >= kq) return;
    const int len = n0 + 2 * q;
    const int j = blockIdx.x * K7_THREADS + threadIdx.x;
    if (j >= len) return;
    double c0, c1, c2;   // P[1:3, j]
    if (j < 3) {
        c0 = sPvv[0 * 3 + j]; c1 = sPvv[1 * 3 + j]; c2 = sPvv[2 * 3 + j];
    } else if (j < n0) {
        const double *col = P + (size_t)j * ld;
        c0 = col[0]; c1 = col[1]; c2 = col[2];
    } else {   // a landmark added earlier in this launch: P[1:3, j] = (Gv'*Pvv)' of that landmark
        int qq = (j - n0) / 2, a = (j - n0) % 2;
        c0 = sGP[qq][a][0]; c1 = sGP[qq][a][1]; c2 = sGP[qq][a][2];
    }
    double g0 = sGv[q][0], g1 = sGv[q][1];
    // Gv = [1 0 g0; 0 1 g1]: t_a = Gv[a][0]*c0 + Gv[a][1]*c1 + Gv[a][2]*c2
    double t0 = 1.0 * c0 + 0.0 * c1 + g0 * c2;
    double t1 = 0.0 * c0 + 1.0 * c1 + g1 * c2;
    P[(size_t)j + (size_t)len * ld] = t0;
    P[(size_t)j + (size_t)(len + 1) * ld] = t1;
}

extern "C" int slam_add_features(slam_t *h, const double *zn, int k, const double R[4]) {
    SLAM_NVTX();
    H_CHECK(h);
    ARG_CHECK(k >= 0 && R, "bad argument");
    if (k == 0) return SLAM_OK;
    ARG_CHECK(zn != nullptr, "NULL zn");
    if (k > SLAM_MAX_OBS || h->n + 2 * k > h->n_max) {
        slam_set_error("slam_add_features: n + 2k = %d exceeds capacity n_max = %d", h->n + 2 * k, h->n_max);
        return SLAM_ERR_CAPACITY;
    }
    double *hz = (double *)((char *)h->h_pin + 256);
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    memcpy(hz, zn, sizeof(double) * 2 * k);
    CUDA_TRY(cudaMemcpyAsync(h->d_zn, hz, sizeof(double) * 2 * k, cudaMemcpyHostToDevice, h->stream));
    for (int q0 = 0; q0 < k; q0 += K7_Q) {
        int kq = std::min(K7_Q, k - q0);
        int n0 = h->n;
        dim3 grid((n0 + 2 * kq + K7_THREADS - 1) / K7_THREADS, kq);
        k_augment<<<grid, K7_THREADS, 0, h->stream>>>(h->d_x, h->d_P, h->ld, n0, h->d_zn + 2 * q0, kq, R[0],
                                                      R[1], R[2], R[3], h->d_sig2);
        h->launches++;
        h->n = n0 + 2 * kq;
    }
    CUDA_TRY(cudaGetLastError());
    return SLAM_OK;
}

// ---------------------------------------------------------------------------------------
// State export for front ends: compute_landmark_ellipses / ellipse,
// src/common.jl:236-262 -- nsigma*sqrtm(P_jj)*[cos(phi)'; sin(phi)'] .+ x_j with
// phi = 0:2pi/nsegs:2pi, per landmark.  Only the 2x2 diagonal blocks are read (24 B per
// landmark) instead of the whole P the reference's monitor pulls (sim/run-ekfsim.jl:22).
// sqrtm of a symmetric positive-definite 2x2 in closed form: (A + sqrt(det) I) / sqrt(tr + 2 sqrt(det)).
// Block 0 is the vehicle (x[1:2], P[1:2,1:2], sim/gr-draw.jl:112-117), block b >= 1 landmark b.
// ---------------------------------------------------------------------------------------
__global__ void k_ellipses(const double *__restrict__ x, const double *__restrict__ P, size_t ld, int b0, int nb,
                           double nsigma, int nsegs, double *__restrict__ out, int ldo) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= nb) return;
    const int blk = b0 + b;
    const int r = blk == 0 ? 0 : 3 + 2 * (blk - 1);
    const double a = P[(size_t)r + (size_t)r * ld], c = P[(size_t)r + (size_t)(r + 1) * ld],
                 d = P[(size_t)(r + 1) + (size_t)(r + 1) * ld];
    const double s = sqrt(a * d - c * c);
    const double t = sqrt(a + d + 2.0 * s);
    const double m00 = (a + s) / t, m01 = c / t, m11 = (d + s) / t;
    const double cx = x[r], cy = x[r + 1];
    const double step = 2.0 * SLAM_PI / nsegs;
    for (int p = 0; p <= nsegs; p++) {
        const double ph = p * step;
        const double cp = cos(ph), sp = sin(ph);
        out[(size_t)(2 * b) + (size_t)p * ldo] = nsigma * (m00 * cp + m01 * sp) + cx;
        out[(size_t)(2 * b + 1) + (size_t)p * ldo] = nsigma * (m01 * cp + m11 * sp) + cy;
    }
}

static int ellipses_common(slam_t *h, int b0, int nb, double nsigma, int nsegs, double *out, int ldo) {
    if (nb == 0) return SLAM_OK;
    const size_t need = (size_t)2 * nb * (nsegs + 1);
    if (need > h->block_cap) {
        cudaFree(h->d_block);
        h->d_block = nullptr;
        h->block_cap = 0;
        CUDA_TRY(cudaMalloc(&h->d_block, sizeof(double) * need));
        h->block_cap = need;
    }
    k_ellipses<<<(nb + 127) / 128, 128, 0, h->stream>>>(h->d_x, h->d_P, h->ld, b0, nb, nsigma, nsegs, h->d_block,
                                                        2 * nb);
    h->launches++;
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaMemcpy2DAsync(out, sizeof(double) * (size_t)ldo, h->d_block, sizeof(double) * 2 * nb,
                               sizeof(double) * 2 * nb, nsegs + 1, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    return SLAM_OK;
}

extern "C" int slam_compute_landmark_ellipses(slam_t *h, double nsigma, int nsegs, double *out, int ldo) {
    H_CHECK(h);
    const int nf = (h->n - 3) / 2;
    ARG_CHECK(nsegs >= 1 && (nf == 0 || out != nullptr), "bad argument");
    ARG_CHECK(ldo >= 2 * nf, "ldo < 2*nf");
    return ellipses_common(h, 1, nf, nsigma, nsegs, out, ldo);
}

extern "C" int slam_vehicle_ellipse(slam_t *h, double nsigma, int nsegs, double *out) {
    H_CHECK(h);
    ARG_CHECK(nsegs >= 1 && out != nullptr, "bad argument");
    return ellipses_common(h, 0, 1, nsigma, nsegs, out, 2);
}

// ---------------------------------------------------------------------------------------
// Front-end payloads of the browser monitor (sim/browser/wsserver.jl:31-85) as headless
// buffers, computed on the device from the resident state so that a monitor never pulls P:
//   "feature-ellipses"  5 x Nf: [cx; cy; sqrt(l1); sqrt(l2); phi], l, u = eig(cov[j, j]) ascending,
//                       phi = atan2(u[2,1], u[1,1]) (wsserver.jl:72-85); block 0 = the vehicle's position block
//                       ("vehicle-ellipse", :58-63).  An eigenvector's sign is LAPACK's choice in the reference;
//                       here phi is normalised to (-pi/2, pi/2] -- the same ellipse.
//   "lidar"             4 x nz: [vx; vy; fx; fy] per observation, laser_lines (src/common.jl:265-283) with
//                       local_to_global (:112-128) on the device pose.
// ("tracks" is state.x[1:3] per step = slam_get_pose; "state" is pose + cov = slam_get_state / slam_get_cov_block.)
// ---------------------------------------------------------------------------------------
__global__ void k_ellipse_params(const double *__restrict__ x, const double *__restrict__ P, size_t ld, int b0, int nb,
                                 double *__restrict__ out) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= nb) return;
    const int blk = b0 + b;
    const int r = blk == 0 ? 0 : 3 + 2 * (blk - 1);
    const double a = P[(size_t)r + (size_t)r * ld], c = P[(size_t)r + (size_t)(r + 1) * ld],
                 d = P[(size_t)(r + 1) + (size_t)(r + 1) * ld];
    const double mean = 0.5 * (a + d), diff = 0.5 * (a - d);
    const double rad = sqrt(diff * diff + c * c);
    const double l1 = mean - rad, l2 = mean + rad;
    // eigenvector of l1: (c, l1 - a) or (l1 - d, c), whichever is longer
    double vx = c, vy = l1 - a;
    if (fabs(l1 - d) > fabs(vy)) { vx = l1 - d; vy = c; }
    if (vx == 0.0 && vy == 0.0) { vx = a <= d ? 1.0 : 0.0; vy = a <= d ? 0.0 : 1.0; }   // already diagonal
    double phi = atan2(vy, vx);
    if (phi > SLAM_PI / 2) phi -= SLAM_PI;
    if (phi <= -SLAM_PI / 2) phi += SLAM_PI;
    double *o = out + 5 * (size_t)b;
    o[0] = x[r]; o[1] = x[r + 1]; o[2] = sqrt(l1); o[3] = sqrt(l2); o[4] = phi;
}

__global__ void k_laser_lines(const double *__restrict__ x, const double *__restrict__ z, int nz, double *__restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nz) return;
    const double xv = x[0], yv = x[1], phi = x[2];
    const double r = z[2 * i], b = z[2 * i + 1];
    const double lx = r * cos(b), ly = r * sin(b);          // observation in the vehicle frame (:277)
    const double cp = cos(phi), sp = sin(phi);
    out[4 * i] = xv;
    out[4 * i + 1] = yv;
    out[4 * i + 2] = (cp * lx + (-sp) * ly) + xv;           // R*l .+ g (:121)
    out[4 * i + 3] = (sp * lx + cp * ly) + yv;
}

static int front_buffer(slam_t *h, size_t need) {
    if (need > h->block_cap) {
        cudaFree(h->d_block);
        h->d_block = nullptr;
        h->block_cap = 0;
        CUDA_TRY(cudaMalloc(&h->d_block, sizeof(double) * need));
        h->block_cap = need;
    }
    return SLAM_OK;
}

extern "C" int slam_feature_ellipses(slam_t *h, double *out) {
    H_CHECK(h);
    const int nf = (h->n - 3) / 2;
    if (nf == 0) return 0;
    ARG_CHECK(out != nullptr, "NULL out");
    int rc = front_buffer(h, (size_t)5 * nf);
    if (rc) return rc;
    k_ellipse_params<<<(nf + 127) / 128, 128, 0, h->stream>>>(h->d_x, h->d_P, h->ld, 1, nf, h->d_block);
    h->launches++;
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaMemcpyAsync(out, h->d_block, sizeof(double) * 5 * nf, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    return nf;
}

extern "C" int slam_vehicle_ellipse_params(slam_t *h, double out[6]) {
    H_CHECK(h);
    ARG_CHECK(out != nullptr, "NULL out");
    int rc = front_buffer(h, 8);
    if (rc) return rc;
    k_ellipse_params<<<1, 32, 0, h->stream>>>(h->d_x, h->d_P, h->ld, 0, 1, h->d_block);
    h->launches++;
    CUDA_TRY(cudaGetLastError());
    double e[5], pose[3];
    CUDA_TRY(cudaMemcpyAsync(e, h->d_block, sizeof(e), cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(cudaMemcpyAsync(pose, h->d_x, sizeof(pose), cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    out[0] = pose[0]; out[1] = pose[1]; out[2] = pose[2];   // [cx, cy, vehicle_phi, rx, ry, phi] (wsserver.jl:60-61)
    out[3] = e[2]; out[4] = e[3]; out[5] = e[4];
    return SLAM_OK;
}

extern "C" int slam_laser_lines(slam_t *h, const double *z, int nz, double *lines) {
    H_CHECK(h);
    ARG_CHECK(nz >= 0 && nz <= SLAM_MAX_OBS, "bad nz");
    if (nz == 0) return SLAM_OK;
    ARG_CHECK(z && lines, "NULL argument");
    int rc = front_buffer(h, (size_t)4 * nz);
    if (rc) return rc;
    double *hz = (double *)((char *)h->h_pin + 256);
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    memcpy(hz, z, sizeof(double) * 2 * nz);
    CUDA_TRY(cudaMemcpyAsync(h->d_z, hz, sizeof(double) * 2 * nz, cudaMemcpyHostToDevice, h->stream));
    k_laser_lines<<<(nz + 127) / 128, 128, 0, h->stream>>>(h->d_x, h->d_z, nz, h->d_block);
    h->launches++;
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaMemcpyAsync(lines, h->d_block, sizeof(double) * 4 * nz, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    return SLAM_OK;
}

// ---------------------------------------------------------------------------------------
// synthetic large-map state (synth.cu)
// ---------------------------------------------------------------------------------------
extern "C" int slam_synth_state(slam_t *h, int n_landmarks, double side, uint64_t seed) {
    H_CHECK(h);
    ARG_CHECK(n_landmarks >= 0 && side > 0, "bad argument");
    int n = 3 + 2 * n_landmarks;
    if (n > h->n_max) {
        slam_set_error("slam_synth_state: n = %d exceeds capacity n_max = %d", n, h->n_max);
        return SLAM_ERR_CAPACITY;
    }
    h->pose_valid = false;
    CUDA_TRY(cudaMemsetAsync(h->d_gate_scratch, 0, 256, h->stream));
    int rc = synth_fill(h->d_x, h->d_P, h->ld, 0, n, n_landmarks, side, seed, h->stream, &h->launches);
    if (rc) return rc;
    h->n = n;
    k_set_pose_snapshot<<<1, 32, 0, h->stream>>>(h->d_x, h->d_pose + 4 * h->pose_slot);
    h->launches++;
    CUDA_TRY(cudaGetLastError());
    rc = launch_sig2_refresh(h->d_P, h->ld, n, h->d_sig2, h->stream, &h->launches);
    if (rc) return rc;
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    return SLAM_OK;
}

extern "C" int slam_synth_observations(slam_t *h, int n_landmarks, double side, uint64_t seed, int m,
                                       int spurious, double *z, int64_t *ids) {
    (void)h;
    ARG_CHECK(z != nullptr, "NULL z");
    return synth_observations_host(n_landmarks, side, seed, m, spurious, z, ids);
}
