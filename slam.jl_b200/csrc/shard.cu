// Large-map mode: one EKF-SLAM filter whose covariance is sharded by ROW PANELS over the ranks of one NCCL
// communicator (one process per GPU).  P is cut into panels of 256 rows (panel 0 = rows [0, 255), panel p >= 1 =
// rows [256 p - 1, 256 p + 255): odd boundaries, so a landmark's two rows never straddle) dealt boustrophedon --
// panel p goes to rank b = p mod 2W if b < W, else 2W-1-b -- so every group of 2W panels gives each rank one early
// and one late panel and the upper-triangle work of K6 is balanced at ANY state size (the map grows one landmark
// at a time, src/ekf.jl:84-122), not only at capacity.  A rank stacks its panels in one array: local panel
// lp = 2 (p div 2W) + (b >= W), element (i, j) at Pl[256 lp + (i - first row of p) + j*ldl]; only i <= j is live.
//
// update (src/ekf.jl:46-77) per rank:
//   1. prep      Hn / fp / v from the replicated x (every rank, identical)
//   2. extract   the S = 3 + 2m rows of P named by H (pose rows + observed landmark rows): the
//                owner of row c writes P[c, c:n], others write zeros            -> rowbuf[S][n]
//   3. ncclAllReduce(sum) of rowbuf  (x + 0 is exact; needs no host knowledge of idf)
//   4. PHt for the rank's own rows (5 columns per observation; P[i,c] = local for c >= i, else
//      rowbuf), and PHt at the S rows from rowbuf alone -> S, Cholesky, C, u on every rank
//   5. W1 = PHt*C for the rank's rows                                           -> send buffer
//   6. ncclAllGather of W1 (n x 2m doubles in total)
//   7. unpermute into K6's two W1 layouts, x += W1*u for every row (x stays replicated)
//   8. K6 on the rank's panels (one launch, kernels.cuh RowMap).
//
// Peer-memory exchange (default when the ranks can map each other's memory: CUDA IPC over NVLink /
// NVSwitch).  Steps 2+3 and 5+6+7 become one kernel each that STORES its result into every rank's
// buffer, followed by a flag barrier in peer memory:
//   2'. extract_p2p   the owner of row c writes P[c, c:n] into rowbuf[s] of every rank (1.6 MB per row at
//                     n = 200k, instead of an allreduce of S x n mostly-zero doubles)
//   5'. gain_p2p      W1 for the rank's own rows, written straight into every rank's K6 layouts, and
//                     x += W1*u into every rank's x (no send buffer, no gathered copy, no unpermute pass)
// The arithmetic per element is the NCCL path's, so both give bit-identical x and P; SLAM_SHARD_P2P=0
// keeps the NCCL path (also the fallback when IPC is refused).
// Reference citations are file:line under /root/reference.
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <vector>

#include "common.cuh"
#include "kernels.cuh"

// ---------------------------------------------------------------------------------------
// NCCL through dlopen: the library loads (and the single-GPU paths run) without libnccl
// ---------------------------------------------------------------------------------------
namespace {
struct NcclApi {
    void *handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t,
                              cudaStream_t) = nullptr;
    ncclResult_t (*Broadcast)(const void *, void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
};
NcclApi g_nccl;

int load_nccl() {
    if (g_nccl.handle) return SLAM_OK;
    const char *names[] = {"libnccl.so.2", "libnccl.so"};
    void *h = nullptr;
    for (const char *nm : names) {
        h = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
        if (h) break;
    }
    if (!h) {
        slam_set_error("cannot dlopen libnccl.so.2: %s", dlerror());
        return SLAM_ERR_NCCL;
    }
#define LOAD(field, sym)                                                \
    *(void **)(&g_nccl.field) = dlsym(h, sym);                          \
    if (!g_nccl.field) {                                                \
        slam_set_error("libnccl lacks %s", sym);                        \
        return SLAM_ERR_NCCL;                                           \
    }
    LOAD(GetUniqueId, "ncclGetUniqueId")
    LOAD(CommInitRank, "ncclCommInitRank")
    LOAD(CommDestroy, "ncclCommDestroy")
    LOAD(AllGather, "ncclAllGather")
    LOAD(AllReduce, "ncclAllReduce")
    LOAD(Broadcast, "ncclBroadcast")
    LOAD(GetErrorString, "ncclGetErrorString")
#undef LOAD
    g_nccl.handle = h;
    return SLAM_OK;
}
}  // namespace

#define NCCL_TRY(expr)                                                                      \
    do {                                                                                    \
        ncclResult_t r__ = (expr);                                                          \
        if (r__ != ncclSuccess) {                                                           \
            slam_set_error("%s failed: %s", #expr, g_nccl.GetErrorString(r__));             \
            return SLAM_ERR_NCCL;                                                           \
        }                                                                                   \
    } while (0)

// ---------------------------------------------------------------------------------------
// partition arithmetic (no device needed); the panel functions themselves are in kernels.cuh
// ---------------------------------------------------------------------------------------
// local panels rank `rank` of `world` holds when the state has n rows
static int local_panels(int n, int world, int rank) {
    if (n <= 0) return 0;
    const int np = panel_of_row(n - 1) + 1;           // global panels in use
    int cnt = 0;                                      // panel_global grows with the local index
    while (panel_global(cnt, rank, world) < np) cnt++;
    return cnt;
}

// Rows [lo[k], hi[k]) of every panel rank owns at state size n, in local order (empty panels included as lo == hi
// so that index k is the local panel number); returns the number of local panels, fills at most `max` entries.
extern "C" int slam_shard_plan(int n, int world, int rank, int *lo, int *hi, int max) {
    ARG_CHECK(n >= 3 && world >= 1 && rank >= 0 && rank < world, "bad argument");
    const int nlp = local_panels(n, world, rank);
    for (int lp = 0; lp < nlp && lp < max; lp++) {
        const int p = panel_global(lp, rank, world);
        const int a = std::min(panel_first(p), n), b = std::min(panel_first(p + 1), n);
        if (lo) lo[lp] = a;
        if (hi) hi[lp] = b;
    }
    return nlp;
}

extern "C" int slam_shard_unique_id(char id[128]) {
    ARG_CHECK(id != nullptr, "NULL id");
    int rc = load_nccl();
    if (rc) return rc;
    ncclUniqueId u;
    NCCL_TRY(g_nccl.GetUniqueId(&u));
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    memcpy(id, &u, 128);
    return SLAM_OK;
}

constexpr int SH_MAXW = 16;   // ranks the peer-memory exchange handles (one NVSwitch domain)

struct slam_shard {
    int device = 0, rank = 0, world = 1;
    int n_max = 0, n = 3;
    int nloc = 0, nlp = 0;      // local row capacity (256 x local panels at n_max) and local panels in use at n
    // peer-memory exchange
    bool p2p = false;
    int last_k6 = 0;               // downdate kernel of the last update (slam_shard_downdate_kernel)
    void *arenaA = nullptr, *arenaB = nullptr;          // peer-visible: [x | flags], [rowbuf | W1c]
    void *peerA[SH_MAXW] = {nullptr}, *peerB[SH_MAXW] = {nullptr};
    unsigned long long *flags = nullptr;                // [SH_MAXW] in arena A: flags[r] = rank r's last epoch
    unsigned long long epoch = 0;
    char *d_ipc = nullptr;
    int *d_agree = nullptr;
    size_t ldl = 0;             // leading dimension of the local panel stack
    ncclComm_t comm = nullptr;
    cudaStream_t stream = nullptr;
    double *d_x = nullptr;
    double *d_Pl = nullptr;     // the rank's row panels, stacked: [ldl][n_max] column-major
    int *d_err = nullptr;
    void *h_pin = nullptr;
    double *d_zf = nullptr, *d_z = nullptr, *d_zn = nullptr;
    int64_t *d_idf = nullptr;
    // spatial pre-gate of associate (gate.cu): bit pattern of an upper bound of the diagonal entries of the
    // landmarks THIS rank scans (those whose 2x2 diagonal block it owns) -- rank-local, no collective
    unsigned long long *d_sig2 = nullptr;
    bool pregate = true;
    // update workspace
    int ws_R = 0, npad = 0;
    double *d_rowbuf = nullptr, *d_PHtloc = nullptr, *d_PHtH = nullptr, *d_S = nullptr, *d_C = nullptr,
           *d_u = nullptr, *d_v = nullptr, *d_Hn = nullptr, *d_W1send = nullptr, *d_W1gath = nullptr,
           *d_W1c = nullptr;
    int *d_fp = nullptr, *d_fpH = nullptr, *d_cs = nullptr;
    // gating workspace
    double *d_P3 = nullptr;      // rows 0..2 of every column, broadcast from rank 0
    void *d_gate = nullptr;
    size_t gate_bytes = 0;
    long long launches = 0;
    bool profiling = false;
    std::vector<cudaEvent_t> ev;
    int ev_used = 0;
};

#define S_CHECK(s)                                                        \
    do {                                                                  \
        ARG_CHECK((s) != nullptr, "NULL handle");                         \
        CUDA_TRY(cudaSetDevice((s)->device));                             \
    } while (0)

// ---------------------------------------------------------------------------------------
// peer memory: every rank maps `base` of every other rank (CUDA IPC; the 64-byte handles travel by
// ncclAllGather, so the host launcher needs nothing new).  *ok is the ranks' common verdict (an
// allreduce(min) of the local outcomes): either everyone uses the peer path or no one does.
// ---------------------------------------------------------------------------------------
static int shard_share(slam_shard_t *s, void *base, void **peers, bool want, bool *ok) {
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
    cudaStream_t st = s->stream;
    const size_t hb = sizeof(cudaIpcMemHandle_t);
    cudaIpcMemHandle_t mine;
    memset(&mine, 0, sizeof(mine));
    int good = want ? 1 : 0;
    if (good && cudaIpcGetMemHandle(&mine, base) != cudaSuccess) {
        cudaGetLastError();
        good = 0;
    }
    CUDA_TRY(cudaMemcpyAsync(s->d_ipc + hb * s->rank, &mine, hb, cudaMemcpyHostToDevice, st));
    NCCL_TRY(g_nccl.AllGather(s->d_ipc + hb * s->rank, s->d_ipc, hb, ncclChar, s->comm, st));
    std::vector<cudaIpcMemHandle_t> all(s->world);
    CUDA_TRY(cudaMemcpyAsync(all.data(), s->d_ipc, hb * s->world, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    for (int r = 0; r < s->world; r++) peers[r] = nullptr;
    peers[s->rank] = base;
    for (int r = 0; r < s->world && good; r++) {
        if (r == s->rank) continue;
        if (cudaIpcOpenMemHandle(&peers[r], all[r], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
            cudaGetLastError();
            peers[r] = nullptr;
            good = 0;
        }
    }
    CUDA_TRY(cudaMemcpyAsync(s->d_agree, &good, sizeof(int), cudaMemcpyHostToDevice, st));
    NCCL_TRY(g_nccl.AllReduce(s->d_agree, s->d_agree, 1, ncclInt32, ncclMin, s->comm, st));
    int all_good = 0;
    CUDA_TRY(cudaMemcpyAsync(&all_good, s->d_agree, sizeof(int), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    if (!all_good) {
        for (int r = 0; r < s->world; r++) {
            if (r != s->rank && peers[r]) cudaIpcCloseMemHandle(peers[r]);
            peers[r] = nullptr;
        }
        peers[s->rank] = base;
    }
    *ok = all_good != 0;
    return SLAM_OK;
}

// ---------------------------------------------------------------------------------------
// lifecycle
// ---------------------------------------------------------------------------------------
extern "C" int slam_shard_destroy(slam_shard_t *s) {
    if (!s) return SLAM_OK;
    cudaSetDevice(s->device);
    if (s->stream) cudaStreamSynchronize(s->stream);
    for (int r = 0; r < s->world && r < SH_MAXW; r++) {
        if (r == s->rank) continue;
        if (s->peerA[r]) cudaIpcCloseMemHandle(s->peerA[r]);
        if (s->peerB[r]) cudaIpcCloseMemHandle(s->peerB[r]);
    }
    if (s->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(s->comm);   // the ranks meet here before anything is freed
    cudaFree(s->arenaA); cudaFree(s->arenaB); cudaFree(s->d_ipc); cudaFree(s->d_agree);
    cudaFree(s->d_Pl); cudaFree(s->d_err);
    cudaFree(s->d_zf); cudaFree(s->d_z); cudaFree(s->d_zn); cudaFree(s->d_idf);
    cudaFree(s->d_PHtloc); cudaFree(s->d_PHtH); cudaFree(s->d_S); cudaFree(s->d_C);
    cudaFree(s->d_u); cudaFree(s->d_v); cudaFree(s->d_Hn); cudaFree(s->d_W1send); cudaFree(s->d_W1gath);
    cudaFree(s->d_fp); cudaFree(s->d_fpH); cudaFree(s->d_cs);
    cudaFree(s->d_P3); cudaFree(s->d_gate); cudaFree(s->d_sig2);
    if (s->h_pin) cudaFreeHost(s->h_pin);
    for (auto &e : s->ev) cudaEventDestroy(e);
    if (s->stream) cudaStreamDestroy(s->stream);
    cudaGetLastError();
    delete s;
    return SLAM_OK;
}

extern "C" int slam_shard_create(int device, int rank, int world, const char id[128], int n_max,
                                 slam_shard_t **out) {
    ARG_CHECK(out != nullptr, "out is NULL");
    *out = nullptr;
    ARG_CHECK(world >= 1 && rank >= 0 && rank < world && n_max >= 3, "bad rank / world / n_max");
    ARG_CHECK(world == 1 || id != nullptr, "NULL unique id");
    int rc = slam_require_device(device);
    if (rc) return rc;
    slam_shard_t *s = new slam_shard();
    s->device = device; s->rank = rank; s->world = world; s->n_max = n_max;
    s->nloc = SH_PH * local_panels(n_max, world, rank);
    {   // the same local capacity on every rank (buffers indexed by local row are exchanged): the largest count
        int mx = 0;
        for (int r = 0; r < world; r++) mx = std::max(mx, local_panels(n_max, world, r));
        s->nloc = SH_PH * mx;
    }
    s->ldl = (size_t)s->nloc + 16;
    s->npad = round_up(n_max, DD_ROW_PAD);
#define SH_TRY(expr)                                                                             \
    do {                                                                                         \
        cudaError_t e__ = (expr);                                                                \
        if (e__ != cudaSuccess) {                                                                \
            slam_set_error("%s failed: %s", #expr, cudaGetErrorString(e__));                     \
            slam_shard_destroy(s);                                                               \
            return SLAM_ERR_CUDA;                                                                \
        }                                                                                        \
    } while (0)
    SH_TRY(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
    SH_TRY(cudaMalloc(&s->arenaA, sizeof(double) * (size_t)s->npad + sizeof(unsigned long long) * SH_MAXW));
    s->d_x = (double *)s->arenaA;
    s->flags = (unsigned long long *)(s->d_x + s->npad);
    SH_TRY(cudaMemsetAsync(s->flags, 0, sizeof(unsigned long long) * SH_MAXW, s->stream));
    SH_TRY(cudaMalloc(&s->d_ipc, (size_t)SH_MAXW * sizeof(cudaIpcMemHandle_t)));
    SH_TRY(cudaMalloc(&s->d_agree, sizeof(int)));
    SH_TRY(cudaMalloc(&s->d_Pl, sizeof(double) * s->ldl * (size_t)n_max));
    SH_TRY(cudaMalloc(&s->d_err, sizeof(int)));
    SH_TRY(cudaMemsetAsync(s->d_err, 0, sizeof(int), s->stream));
    SH_TRY(cudaMalloc(&s->d_sig2, sizeof(unsigned long long) * 4));
    SH_TRY(cudaMemsetAsync(s->d_sig2, 0, sizeof(unsigned long long) * 4, s->stream));
    if (const char *e = getenv("SLAM_SHARD_PREGATE")) s->pregate = atoi(e) != 0;
    SH_TRY(cudaMemsetAsync(s->d_x, 0, sizeof(double) * (size_t)s->npad, s->stream));
    const size_t no = SLAM_MAX_OBS;
    SH_TRY(cudaMalloc(&s->d_z, sizeof(double) * 2 * no));
    SH_TRY(cudaMalloc(&s->d_zf, sizeof(double) * 2 * no));
    SH_TRY(cudaMalloc(&s->d_zn, sizeof(double) * 2 * no));
    SH_TRY(cudaMalloc(&s->d_idf, sizeof(int64_t) * no));
    SH_TRY(cudaMallocHost(&s->h_pin, sizeof(double) * 8 * no + 256));
    SH_TRY(cudaMalloc(&s->d_P3, sizeof(double) * 4 * (size_t)s->npad));
    SH_TRY(cudaStreamSynchronize(s->stream));
#undef SH_TRY
    if (world > 1) {
        rc = load_nccl();
        if (rc) { slam_shard_destroy(s); return rc; }
        ncclUniqueId u;
        memcpy(&u, id, 128);
        ncclResult_t r = g_nccl.CommInitRank(&s->comm, world, u, rank);
        if (r != ncclSuccess) {
            slam_set_error("ncclCommInitRank failed: %s", g_nccl.GetErrorString(r));
            s->comm = nullptr;
            slam_shard_destroy(s);
            return SLAM_ERR_NCCL;
        }
    }
    if (world > 1) {
        const char *env = getenv("SLAM_SHARD_P2P");
        const bool want = world <= SH_MAXW && !(env && atoi(env) == 0);
        rc = shard_share(s, s->arenaA, s->peerA, want, &s->p2p);
        if (rc) { slam_shard_destroy(s); return rc; }
    }
    s->n = 3;
    s->nlp = local_panels(3, world, rank);
    *out = s;
    return SLAM_OK;
}

// 1 when the update exchanges rows and W1 through peer memory, 0 when it uses NCCL collectives.
extern "C" int slam_shard_p2p(slam_shard_t *s) {
    if (!s) { slam_set_error("slam_shard_p2p: NULL handle"); return SLAM_ERR_ARG; }
    return s->p2p ? 1 : 0;
}

extern "C" int slam_shard_downdate_kernel(slam_shard_t *s) {
    if (!s) { slam_set_error("slam_shard_downdate_kernel: NULL handle"); return SLAM_ERR_ARG; }
    return s->last_k6;
}

extern "C" int slam_shard_n(slam_shard_t *s) {
    if (!s) { slam_set_error("slam_shard_n: NULL handle"); return SLAM_ERR_ARG; }
    return s->n;
}
extern "C" void *slam_shard_stream(slam_shard_t *s) { return s ? (void *)s->stream : nullptr; }
extern "C" long long slam_shard_launch_count(slam_shard_t *s) { return s ? s->launches : 0; }

static int shard_check_latched(slam_shard_t *s) {
    int *he = (int *)s->h_pin;
    CUDA_TRY(cudaMemcpyAsync(he, s->d_err, sizeof(int), cudaMemcpyDeviceToHost, s->stream));
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    int e = *he;
    if (e != 0) {
        CUDA_TRY(cudaMemsetAsync(s->d_err, 0, sizeof(int), s->stream));
        slam_set_error(e == SLAM_ERR_NOT_PD      ? "innovation covariance not positive definite"
                       : e == SLAM_ERR_BAD_INDEX ? "landmark id out of range"
                                                 : "device-side error %d", e);
        return e;
    }
    return SLAM_OK;
}

extern "C" int slam_shard_sync(slam_shard_t *s) {
    S_CHECK(s);
    return shard_check_latched(s);
}

// max over P[f, f] of the landmark rows f this rank owns, as the bit pattern of a non-negative double (same
// convention as k_sig2_diag in gate.cu: NaN -> a pattern above +inf, negatives ignored)
__global__ void k_sh_sig2_diag(const double *__restrict__ Pl, size_t ldl, int rank, int world, int nlp, int n,
                               unsigned long long *sig2bits) {
    unsigned long long mine = 0ull;
    for (int li = blockIdx.x * blockDim.x + threadIdx.x; li < nlp * SH_PH; li += gridDim.x * blockDim.x) {
        const int p = panel_global(li / SH_PH, rank, world);
        const int f = panel_first(p) + li % SH_PH;
        if (f < 3 || f >= min(panel_first(p + 1), n)) continue;
        const double v = Pl[(size_t)li + (size_t)f * ldl];
        unsigned long long b = (unsigned long long)__double_as_longlong(v);
        if (v != v) b = 0x7ff8000000000000ull;
        else if (v < 0.0) b = 0ull;
        mine = b > mine ? b : mine;
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        const unsigned long long o = __shfl_down_sync(0xffffffffu, mine, off);
        mine = o > mine ? o : mine;
    }
    if ((threadIdx.x & 31) == 0 && mine) atomicMax(sig2bits, mine);
}

extern "C" int slam_shard_synth_state(slam_shard_t *s, int n_landmarks, double side, uint64_t seed) {
    S_CHECK(s);
    ARG_CHECK(n_landmarks >= 0 && side > 0, "bad argument");
    const int n = 3 + 2 * n_landmarks;
    if (n > s->n_max) {
        slam_set_error("slam_shard_synth_state: n = %d exceeds capacity n_max = %d", n, s->n_max);
        return SLAM_ERR_CAPACITY;
    }
    s->n = n;
    s->nlp = local_panels(n, s->world, s->rank);
    CUDA_TRY(cudaMemsetAsync(s->d_sig2, 0, sizeof(unsigned long long), s->stream));
    int rc = synth_fill_map(s->d_x, s->d_Pl, s->ldl, rowmap_panels(s->rank, s->world, s->nlp), n_landmarks, side, seed,
                            s->stream, &s->launches);
    if (rc) return rc;
    if (s->nlp > 0) {   // the pre-gate's bound: exact maximum over the diagonal entries this rank owns
        k_sh_sig2_diag<<<std::min((s->nlp * SH_PH + 255) / 256, 296), 256, 0, s->stream>>>(
            s->d_Pl, s->ldl, s->rank, s->world, s->nlp, n, s->d_sig2);
        s->launches++;
    }
    CUDA_TRY(cudaGetLastError());
    return SLAM_OK;
}

extern "C" int slam_shard_get_pose(slam_shard_t *s, double pose[3]) {
    S_CHECK(s);
    ARG_CHECK(pose != nullptr, "NULL pose");
    double *hp = (double *)s->h_pin + 8;
    CUDA_TRY(cudaMemcpyAsync(hp, s->d_x, sizeof(double) * 3, cudaMemcpyDeviceToHost, s->stream));
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    pose[0] = hp[0]; pose[1] = hp[1]; pose[2] = hp[2];
    return SLAM_OK;
}

extern "C" int slam_shard_get_x(slam_shard_t *s, double *x) {
    S_CHECK(s);
    ARG_CHECK(x != nullptr, "NULL x");
    CUDA_TRY(cudaMemcpyAsync(x, s->d_x, sizeof(double) * s->n, cudaMemcpyDeviceToHost, s->stream));
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    return SLAM_OK;
}

// owned entries (row i in one of this rank's panels) of the requested window; others untouched
extern "C" int slam_shard_get_cov_block(slam_shard_t *s, int i0, int ni, int j0, int nj, double *out, int ldo) {
    S_CHECK(s);
    ARG_CHECK(out && i0 >= 0 && j0 >= 0 && ni >= 0 && nj >= 0 && i0 + ni <= s->n && j0 + nj <= s->n && ldo >= ni,
              "bad window");
    int copied = 0;
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    if (ni == 0 || nj == 0) return 0;
    for (int p = panel_of_row(i0); p <= panel_of_row(i0 + ni - 1); p++) {
        if (panel_owner(p, s->world) != s->rank) continue;
        const int a = std::max(i0, panel_first(p)), b = std::min(i0 + ni, std::min(panel_first(p + 1), s->n));
        if (a >= b) continue;
        const size_t li = (size_t)panel_local(p, s->world) * SH_PH + (a - panel_first(p));
        // rows [a, b) of columns [j0, j0+nj): strided 2-D copy out of the panel stack
        CUDA_TRY(cudaMemcpy2D(out + (a - i0), sizeof(double) * (size_t)ldo, s->d_Pl + li + (size_t)j0 * s->ldl,
                              sizeof(double) * s->ldl, sizeof(double) * (b - a), nj, cudaMemcpyDeviceToHost));
        copied += (b - a) * nj;
    }
    return copied;
}

// ---------------------------------------------------------------------------------------
// update kernels
// ---------------------------------------------------------------------------------------
struct ShardGeom {
    int rank, world, n, npad;
    int nloc;          // local row capacity = rows reserved per rank in rank-local buffers (PHtloc, W1send)
    int nlp;           // local panels in use at this n
    size_t ldl;
    double *Pl;        // the rank's panel stack: element (i, j) at Pl[local row of i + j*ldl]
};

// owner and local row of global row i
__device__ __forceinline__ void row_owner(const ShardGeom &g, int i, int *owner, int *li) {
    const int p = panel_of_row(i);
    *owner = panel_owner(p, g.world);
    *li = panel_local(p, g.world) * SH_PH + (i - panel_first(p));
}
// global row of local row li of this rank, or -1 when that slot holds no row of the current state
__device__ __forceinline__ int local_row_global(const ShardGeom &g, int li) {
    const int p = panel_global(li / SH_PH, g.rank, g.world);
    const int i = panel_first(p) + li % SH_PH;
    return i < min(panel_first(p + 1), g.n) ? i : -1;
}

// step 1: observation model for every associated observation, innovation v, and the list cs of
// the S = 3 + 2m state rows that H touches (src/ekf.jl:55-61, src/common.jl:139-165)
__global__ void k_sh_prep(const double *__restrict__ x, int n, const double *__restrict__ zf,
                          const int64_t *__restrict__ idf, int m, double *Hn, int *fp, int *fpH, int *cs,
                          double *v, int *err) {
    const int Nf = (n - 3) / 2;
    if (threadIdx.x < 3 && blockIdx.x == 0) cs[threadIdx.x] = threadIdx.x;
    for (int o = blockIdx.x * blockDim.x + threadIdx.x; o < m; o += gridDim.x * blockDim.x) {
        int64_t id = idf[o];
        if (id < 1 || id > Nf) {
            atomicCAS(err, 0, SLAM_ERR_BAD_INDEX);
            id = 1;
        }
        int f = 3 + 2 * ((int)id - 1);
        double zp[2];
        observe_model(x[0], x[1], x[2], x[f], x[f + 1], zp, Hn + 10 * o);
        fp[o] = f;
        fpH[o] = 3 + 2 * o;
        cs[3 + 2 * o] = f;
        cs[3 + 2 * o + 1] = f + 1;
        v[2 * o] = zf[2 * o] - zp[0];
        v[2 * o + 1] = mpi_to_pi(zf[2 * o + 1] - zp[1]);
    }
}

// step 2: rowbuf[s][j] = P[cs[s], j] for j >= cs[s] if this rank owns row cs[s], else 0
__global__ void k_sh_extract(ShardGeom g, const int *__restrict__ cs, double *__restrict__ rowbuf) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    const int s = blockIdx.y;
    if (j >= g.npad) return;
    const int c = cs[s];
    int owner, li;
    row_owner(g, c, &owner, &li);
    double val = 0.0;
    if (owner == g.rank && j >= c && j < g.n) val = g.Pl[(size_t)li + (size_t)j * g.ldl];
    rowbuf[(size_t)s * g.npad + j] = val;
}

// P[i, cs[s]] for a row i this rank owns (local row li): local when cs[s] >= i, else from the gathered row
__device__ __forceinline__ double sh_get(const ShardGeom &g, const double *rowbuf, int li, int i, int s, int c) {
    return c >= i ? g.Pl[(size_t)li + (size_t)c * g.ldl] : rowbuf[(size_t)s * g.npad + i];
}

// step 4a: PHt rows of the rank's own rows (src/ekf.jl:67), rank-major over the local row index
__global__ void __launch_bounds__(256)
k_sh_pht(ShardGeom g, const double *__restrict__ rowbuf, const double *__restrict__ Hn, const int *__restrict__ fp,
         int m, double *__restrict__ PHtloc) {
    pdl_wait();   // launched behind the exchange barrier with programmatic serialization
    const int li = blockIdx.x * blockDim.x + threadIdx.x;
    const int nloc = g.nloc;
    if (li >= nloc) return;
    const int i = li < g.nlp * SH_PH ? local_row_global(g, li) : -1;
    const bool live = i >= 0;
    double p0 = 0, p1 = 0, p2 = 0;
    if (live) {
        p0 = sh_get(g, rowbuf, li, i, 0, 0);
        p1 = sh_get(g, rowbuf, li, i, 1, 1);
        p2 = sh_get(g, rowbuf, li, i, 2, 2);
    }
    for (int o = 0; o < m; o++) {
        double a0 = 0, a1 = 0;
        if (live) {
            const double *hh = Hn + 10 * o;
            const int f = fp[o];
            double p3 = sh_get(g, rowbuf, li, i, 3 + 2 * o, f);
            double p4 = sh_get(g, rowbuf, li, i, 3 + 2 * o + 1, f + 1);
            a0 = p0 * hh[0] + p1 * hh[1] + p2 * hh[2] + p3 * hh[6] + p4 * hh[7];
            a1 = p0 * hh[3] + p1 * hh[4] + p2 * hh[5] + p3 * hh[8] + p4 * hh[9];
        }
        PHtloc[(size_t)(2 * o) * nloc + li] = a0;
        PHtloc[(size_t)(2 * o + 1) * nloc + li] = a1;
    }
}

// step 4b: PHt at the S rows named by H, from rowbuf alone (identical on every rank)
__global__ void k_sh_pht_hset(int npad, int S, const int *__restrict__ cs, const double *__restrict__ rowbuf,
                              const double *__restrict__ Hn, int m, double *__restrict__ PHtH) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= S * m) return;
    const int s = idx % S, o = idx / S;
    const int i = cs[s];
    auto at = [&](int s2) {   // P[i, cs[s2]]
        int c = cs[s2];
        return c >= i ? rowbuf[(size_t)s * npad + c] : rowbuf[(size_t)s2 * npad + i];
    };
    const double *hh = Hn + 10 * o;
    double p0 = at(0), p1 = at(1), p2 = at(2), p3 = at(3 + 2 * o), p4 = at(3 + 2 * o + 1);
    PHtH[(size_t)(2 * o) * S + s] = p0 * hh[0] + p1 * hh[1] + p2 * hh[2] + p3 * hh[6] + p4 * hh[7];
    PHtH[(size_t)(2 * o + 1) * S + s] = p0 * hh[3] + p1 * hh[4] + p2 * hh[5] + p3 * hh[8] + p4 * hh[9];
}

// step 5: W1 rows of the rank's own rows (src/ekf.jl:71) into the allgather send buffer
// [chunk][local row][RC]
__global__ void __launch_bounds__(128)
k_sh_gain(int nloc, int R, int RC, const double *__restrict__ PHtloc, const double *__restrict__ Cm,
          double *__restrict__ W1send) {
    const int li = blockIdx.x * blockDim.x + threadIdx.x;
    if (li >= nloc) return;
    const int nch = (R + RC - 1) / RC;
    for (int k = 0; k < nch * RC; k++) {
        double a = 0;
        if (k < R)
            for (int l = 0; l <= k; l++) a += PHtloc[(size_t)l * nloc + li] * Cm[(size_t)l + (size_t)k * R];
        W1send[((size_t)(k / RC) * nloc + li) * RC + (k % RC)] = a;
    }
}

// step 7: gathered W1 -> K6's two layouts in global row order; x += W1*u (src/ekf.jl:72-74)
__global__ void __launch_bounds__(128)
k_sh_unpermute(ShardGeom g, int R, int RC, const double *__restrict__ W1gath, const double *__restrict__ u,
               double *__restrict__ W1c, double *__restrict__ x) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= g.npad) return;
    const int nch = (R + RC - 1) / RC, nloc = g.nloc;
    int owner = 0, li = 0;
    const bool live = i < g.n;
    if (live) row_owner(g, i, &owner, &li);
    double dx = 0;
    const int ldk = w1_ldk(RC);
    const size_t csz = w1_chunk_doubles(g.npad, RC), rmo = w1_rank_major_offset(g.npad, RC);
    for (int c = 0; c < nch; c++)
        for (int kk = RC; kk < ldk; kk++) W1c[(size_t)c * csz + (size_t)i * ldk + kk] = 0.0;   // row padding, kept defined
    for (int k = 0; k < nch * RC; k++) {
        const int c = k / RC, kk = k % RC;
        double a = live ? W1gath[(((size_t)owner * nch + c) * nloc + li) * RC + kk] : 0.0;
        double *chunk = W1c + (size_t)c * csz;
        chunk[(size_t)i * ldk + kk] = a;
        chunk[rmo + (size_t)kk * g.npad + i] = a;
        if (k < R) dx += a * u[k];
    }
    if (live) x[i] += dx;
}

// ---------------------------------------------------------------------------------------
// peer-memory variants of steps 2+3 and 5+6+7
// ---------------------------------------------------------------------------------------
struct ShardPeers {
    double *x[SH_MAXW];
    double *rowbuf[SH_MAXW];
    double *W1c[SH_MAXW];
    unsigned long long *flags[SH_MAXW];
};

// step 2': the owner of row cs[s] stores P[cs[s], j], j >= cs[s], into rowbuf[s] of the ranks that read it
// (entries j < cs[s] are never read: P[i, c] with c >= i comes from the reader's own block).  Entry j has two
// kinds of reader: the owner of row j (k_sh_pht forms PHt for its own rows only) and, when j is itself one of
// the S rows named by H, every rank (k_sh_pht_hset forms the S x 2m block redundantly).  So the x-blocks over
// j store to ONE rank each, and gridDim.x - hx extra x-blocks broadcast the S entries of the H set: n + 8 S
// remote stores per row instead of 8 n.
__global__ void __launch_bounds__(256)
k_sh_extract_p2p(ShardGeom g, const int *__restrict__ cs, int S, int hx, ShardPeers pr) {
    const int s = blockIdx.y;
    const int c = cs[s];
    int owner, li;
    row_owner(g, c, &owner, &li);
    if (owner != g.rank) return;
    const double *row = g.Pl + (size_t)li;
    if ((int)blockIdx.x < hx) {
        const int j = blockIdx.x * blockDim.x + threadIdx.x;
        if (j < c || j >= g.npad) return;
        const double val = j < g.n ? row[(size_t)j * g.ldl] : 0.0;
        int oj, lj;
        row_owner(g, j, &oj, &lj);
        pr.rowbuf[oj][(size_t)s * g.npad + j] = val;
    } else {
        const int t = (blockIdx.x - hx) * blockDim.x + threadIdx.x;
        if (t >= S) return;
        const int j = cs[t];
        if (j < c) return;
        const double val = row[(size_t)j * g.ldl];
        for (int r = 0; r < g.world; r++) pr.rowbuf[r][(size_t)s * g.npad + j] = val;
    }
}

// Barrier over peer memory: rank `rank` publishes `epoch` in flags[rank] of every rank, then waits
// until its own flags show `epoch` from everyone.  Stream order puts the kernel after this rank's
// stores to its peers; the release / acquire pair at system scope makes them visible to whoever
// sees the flag.  A rank that never arrives turns into SLAM_ERR_NCCL after ~60 s instead of a hang.
__global__ void k_sh_barrier(ShardPeers pr, int rank, int world, unsigned long long epoch, int *err) {
    pdl_launch_dependents();   // the consumer (k_sh_pht / K6) is scheduled while this kernel spins, and waits in pdl_wait()
    const int t = threadIdx.x;
    if (t >= world) return;
    __threadfence_system();
    unsigned long long *dst = pr.flags[t] + rank;
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(dst), "l"(epoch) : "memory");
    const unsigned long long *src = pr.flags[rank] + t;
    const long long t0 = clock64();
    unsigned long long seen = 0;
    while (true) {
        asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(seen) : "l"(src) : "memory");
        if (seen >= epoch) break;
        if (clock64() - t0 > 120000000000LL) {
            atomicCAS(err, 0, SLAM_ERR_NCCL);
            break;
        }
        __nanosleep(200);
    }
}

// step 5': W1 rows of the rank's own rows (src/ekf.jl:71), stored into both K6 layouts of EVERY rank in
// global row order, and x += W1*u (src/ekf.jl:72-74) into every rank's x.  One CTA = 128 consecutive
// rows of one block; a chunk's row-major tile is staged in shared memory so that the remote stores are
// whole contiguous lines.  Same sums as k_sh_gain / k_sh_unpermute.
constexpr int SHG_ROWS = 128;
__global__ void __launch_bounds__(SHG_ROWS)
k_sh_gain_p2p(ShardGeom g, int R, int RC, const double *__restrict__ PHtloc, const double *__restrict__ Cm,
              const double *__restrict__ u, ShardPeers pr, int rank_major) {
    extern __shared__ double sh_tile[];          // [SHG_ROWS][RC + W1_PAD]: the row-major layout's rows, padding included
    const int ldk = w1_ldk(RC);
    const size_t csz = w1_chunk_doubles(g.npad, RC), rmo = w1_rank_major_offset(g.npad, RC);
    // a CTA = SHG_ROWS consecutive local rows, all inside one panel (SHG_ROWS divides the panel height)
    const int l0 = blockIdx.x * SHG_ROWS;        // first local row of the CTA
    const int p = panel_global(l0 / SH_PH, g.rank, g.world);
    const int i0 = panel_first(p) + l0 % SH_PH;
    const int rows = min(SHG_ROWS, min(panel_first(p + 1), g.n) - i0);
    if (rows <= 0) return;
    const int t = threadIdx.x;
    const int li = l0 + t, nloc = g.nloc;
    const int i = i0 + t;
    const bool live = t < rows;
    const int nch = (R + RC - 1) / RC;
    double dx = 0;
    for (int kk = RC; kk < ldk; kk++) sh_tile[t * ldk + kk] = 0.0;
    for (int c = 0; c < nch; c++) {
        for (int kk = 0; kk < RC; kk++) {
            const int k = c * RC + kk;
            double a = 0;
            if (live && k < R)
                for (int l = 0; l <= k; l++) a += PHtloc[(size_t)l * nloc + li] * Cm[(size_t)l + (size_t)k * R];
            sh_tile[t * ldk + kk] = a;
            if (live) {
                if (rank_major) {   // only the FMA downdate kernel (up to 4 ranks) reads the rank-major copy
                    const size_t o2 = (size_t)c * csz + rmo + (size_t)kk * g.npad + i;
                    for (int r = 0; r < g.world; r++) pr.W1c[r][o2] = a;      // lanes = rows
                }
                if (k < R) dx += a * u[k];
            }
        }
        __syncthreads();
        const size_t o1 = (size_t)c * csz + (size_t)i0 * ldk;   // row-major copy: the tile is contiguous
        for (int e = t; e < rows * ldk; e += SHG_ROWS) {
            const double a = sh_tile[e];
            for (int r = 0; r < g.world; r++) pr.W1c[r][o1 + e] = a;
        }
        __syncthreads();
    }
    if (live) {
        const double xn = pr.x[g.rank][i] + dx;
        for (int r = 0; r < g.world; r++) pr.x[r][i] = xn;
    }
}

// rows [n, npad_used) of the local K6 layouts are zero (the tiles of the last row block read them)
__global__ void k_sh_zero_tail(int n, int npad_used, int npad, int RC, int nch, double *W1c) {
    const int i = n + blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= npad_used) return;
    const int ldk = w1_ldk(RC);
    for (int c = 0; c < nch; c++) {
        double *chunk = W1c + (size_t)c * w1_chunk_doubles(npad, RC);
        for (int kk = 0; kk < ldk; kk++) chunk[(size_t)i * ldk + kk] = 0.0;
        for (int kk = 0; kk < RC; kk++) chunk[w1_rank_major_offset(npad, RC) + (size_t)kk * npad + i] = 0.0;
    }
}

// ---------------------------------------------------------------------------------------
// update driver
// ---------------------------------------------------------------------------------------
static ShardPeers shard_peers(const slam_shard_t *s, size_t rowbuf_doubles) {
    ShardPeers pr;
    for (int r = 0; r < SH_MAXW; r++) {
        const bool on = r < s->world;
        pr.x[r] = on ? (double *)s->peerA[r] : nullptr;
        pr.flags[r] = on ? (unsigned long long *)((double *)s->peerA[r] + s->npad) : nullptr;
        pr.rowbuf[r] = on ? (double *)s->peerB[r] : nullptr;
        pr.W1c[r] = on ? (double *)s->peerB[r] + rowbuf_doubles : nullptr;
    }
    return pr;
}
static ShardGeom geom(const slam_shard_t *s) {
    ShardGeom g;
    g.rank = s->rank; g.world = s->world; g.n = s->n; g.npad = s->npad;
    g.nloc = s->nloc; g.nlp = s->nlp; g.ldl = s->ldl; g.Pl = s->d_Pl;
    return g;
}

// Padding guard of the rank's panel stack and of x (diagnostic, as slam_debug_guard; rank-local, no collective).  A
// slot (li, j) of the ldl x n_max stack is live when local row li holds a global row i < n of the current state and
// j < n; arm = 1 paints every other slot and x[n, npad) with a NaN payload, arm = 0 counts the painted ones that changed.
static constexpr unsigned long long SH_GUARD_BITS = 0x7FF8C0DEC0DEC0DEull;

__global__ void k_sh_guard(ShardGeom g, double *x, int n_max, int arm, unsigned long long *changed) {
    const size_t slab = g.ldl * (size_t)n_max, total = slab + (size_t)g.npad;
    unsigned long long bad = 0;
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        double *p;
        if (e < slab) {
            const int li = (int)(e % g.ldl), j = (int)(e / g.ldl);
            if (j < g.n && li < g.nlp * SH_PH && local_row_global(g, li) >= 0) continue;
            p = g.Pl + e;
        } else {
            const size_t i = e - slab;
            if (i < (size_t)g.n) continue;
            p = x + i;
        }
        if (arm) *p = __longlong_as_double((long long)SH_GUARD_BITS);
        else if ((unsigned long long)__double_as_longlong(*p) != SH_GUARD_BITS) bad++;
    }
    if (!arm && bad) atomicAdd(changed, bad);
}

extern "C" int slam_shard_debug_guard(slam_shard_t *s, int arm, long long *changed) {
    S_CHECK(s);
    ARG_CHECK(arm || changed != nullptr, "NULL changed");
    unsigned long long *d_cnt = nullptr;
    CUDA_TRY(cudaMalloc(&d_cnt, sizeof(unsigned long long)));
    cudaError_t e = cudaMemsetAsync(d_cnt, 0, sizeof(unsigned long long), s->stream);
    if (e == cudaSuccess) {
        k_sh_guard<<<148 * 8, 256, 0, s->stream>>>(geom(s), s->d_x, s->n_max, arm, d_cnt);
        s->launches++;
        e = cudaGetLastError();
    }
    unsigned long long cnt = 0;
    if (e == cudaSuccess) e = cudaMemcpyAsync(&cnt, d_cnt, sizeof(cnt), cudaMemcpyDeviceToHost, s->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s->stream);
    cudaFree(d_cnt);
    CUDA_TRY(e);
    if (changed) *changed = (long long)cnt;
    return SLAM_OK;
}

static int shard_ensure_ws(slam_shard_t *s, int R) {
    const int RC = downdate_chunk_width(R);
    const int Rw = std::max(round_up(R, RC), 32);
    if (s->ws_R >= Rw) return SLAM_OK;
    double **bufs[] = {&s->d_PHtloc, &s->d_PHtH, &s->d_S, &s->d_C, &s->d_u, &s->d_v, &s->d_Hn, &s->d_W1send, &s->d_W1gath};
    for (auto b : bufs) { cudaFree(*b); *b = nullptr; }
    // rowbuf and W1c live in the peer-visible arena B.  Growing it is collective (every rank sees the
    // same m): unmap the peers' old arenas, meet, free, allocate, share again.
    if (s->arenaB) {
        CUDA_TRY(cudaStreamSynchronize(s->stream));
        for (int r = 0; r < s->world && r < SH_MAXW; r++) {
            if (r != s->rank && s->peerB[r]) cudaIpcCloseMemHandle(s->peerB[r]);
            s->peerB[r] = nullptr;
        }
        if (s->world > 1) {
            NCCL_TRY(g_nccl.AllReduce(s->d_agree, s->d_agree, 1, ncclInt32, ncclMin, s->comm, s->stream));
            CUDA_TRY(cudaStreamSynchronize(s->stream));
        }
        cudaFree(s->arenaB);
        s->arenaB = nullptr;
    }
    s->d_rowbuf = s->d_W1c = nullptr;
    cudaFree(s->d_fp); cudaFree(s->d_fpH); cudaFree(s->d_cs);
    s->d_fp = s->d_fpH = s->d_cs = nullptr;
    s->ws_R = 0;
    const size_t S = 3 + Rw, nloc = (size_t)s->nloc, npad = s->npad;
    CUDA_TRY(cudaMalloc(&s->arenaB, sizeof(double) * (S * npad + w1_workspace_doubles((int)npad, Rw))));
    s->d_rowbuf = (double *)s->arenaB;
    s->d_W1c = s->d_rowbuf + S * npad;
    if (s->world > 1 && s->p2p) {
        bool ok = false;
        int rc = shard_share(s, s->arenaB, s->peerB, true, &ok);
        if (rc) return rc;
        s->p2p = ok;   // refused for this arena: every rank falls back to the NCCL path together
    }
    CUDA_TRY(cudaMalloc(&s->d_PHtloc, sizeof(double) * Rw * nloc));
    CUDA_TRY(cudaMalloc(&s->d_PHtH, sizeof(double) * Rw * S));
    CUDA_TRY(cudaMalloc(&s->d_S, sizeof(double) * Rw * Rw));
    CUDA_TRY(cudaMalloc(&s->d_C, sizeof(double) * Rw * Rw));
    CUDA_TRY(cudaMalloc(&s->d_u, sizeof(double) * Rw));
    CUDA_TRY(cudaMalloc(&s->d_v, sizeof(double) * Rw));
    CUDA_TRY(cudaMalloc(&s->d_Hn, sizeof(double) * 10 * (Rw / 2)));
    CUDA_TRY(cudaMalloc(&s->d_W1send, sizeof(double) * Rw * nloc));
    if (s->world > 1) CUDA_TRY(cudaMalloc(&s->d_W1gath, sizeof(double) * Rw * nloc * s->world));
    CUDA_TRY(cudaMalloc(&s->d_fp, sizeof(int) * (Rw / 2)));
    CUDA_TRY(cudaMalloc(&s->d_fpH, sizeof(int) * (Rw / 2)));
    CUDA_TRY(cudaMalloc(&s->d_cs, sizeof(int) * S));
    s->ws_R = Rw;
    return SLAM_OK;
}

static int shard_update_device(slam_shard_t *s, const double *d_zf, int m, const double R[4],
                               const int64_t *d_idf) {
    if (m == 0) return SLAM_OK;
    if (m > SLAM_MAX_OBS) {
        slam_set_error("slam_shard_update: m = %d exceeds SLAM_MAX_OBS", m);
        return SLAM_ERR_CAPACITY;
    }
    if (s->n <= 3) {
        slam_set_error("slam_shard_update: %d observations but the map is empty", m);
        return SLAM_ERR_BAD_INDEX;
    }
    const int Rk = 2 * m, S = 3 + Rk;
    int rc = shard_ensure_ws(s, Rk);
    if (rc) return rc;
    const int RC = downdate_chunk_width(Rk), nch = (Rk + RC - 1) / RC;
    const int nloc = s->nloc, npad = s->npad;
    const ShardGeom g = geom(s);
    cudaStream_t st = s->stream;
    // profiling: five events per update -- start, H rows exchanged, before the gain, K6 start, end
    const bool timed = s->profiling && s->ev_used + 5 <= (int)s->ev.size();
    cudaEvent_t *ev = timed ? &s->ev[s->ev_used] : nullptr;
    if (timed) CUDA_TRY(cudaEventRecord(ev[0], st));
    k_sh_prep<<<(m + 255) / 256, 256, 0, st>>>(s->d_x, s->n, d_zf, d_idf, m, s->d_Hn, s->d_fp, s->d_fpH, s->d_cs,
                                              s->d_v, s->d_err);
    const bool p2p = s->world > 1 && s->p2p;
    const int npad_used = round_up(s->n, DD_ROW_PAD);
    ShardPeers pr;
    if (p2p) {
        pr = shard_peers(s, (size_t)(s->d_W1c - s->d_rowbuf));
        if (npad_used > s->n) {
            k_sh_zero_tail<<<(npad_used - s->n + 127) / 128, 128, 0, st>>>(s->n, npad_used, npad, RC, nch, s->d_W1c);
            s->launches++;
        }
        const int hx = (npad + 255) / 256;
        k_sh_extract_p2p<<<dim3(hx + (S + 255) / 256, S), 256, 0, st>>>(g, s->d_cs, S, hx, pr);
        k_sh_barrier<<<1, 32, 0, st>>>(pr, s->rank, s->world, ++s->epoch, s->d_err);
        s->launches += 3;
    } else {
        k_sh_extract<<<dim3((npad + 255) / 256, S), 256, 0, st>>>(g, s->d_cs, s->d_rowbuf);
        s->launches += 2;
        if (s->world > 1)
            NCCL_TRY(g_nccl.AllReduce(s->d_rowbuf, s->d_rowbuf, (size_t)S * npad, ncclFloat64, ncclSum, s->comm, st));
    }
    if (timed) CUDA_TRY(cudaEventRecord(ev[1], st));
    CUDA_TRY(launch_pdl(k_sh_pht, dim3((nloc + 255) / 256), dim3(256), 0, st, g, (const double *)s->d_rowbuf,
                        (const double *)s->d_Hn, (const int *)s->d_fp, m, s->d_PHtloc));
    k_sh_pht_hset<<<(S * m + 127) / 128, 128, 0, st>>>(npad, S, s->d_cs, s->d_rowbuf, s->d_Hn, m, s->d_PHtH);
    rc = launch_innov_chol(s->d_PHtH, S, s->d_Hn, s->d_fpH, s->d_v, m, R, s->d_S, s->d_C, s->d_u, s->d_err, st);
    if (rc) return rc;
    s->launches += 3;
    if (timed) CUDA_TRY(cudaEventRecord(ev[2], st));
    if (p2p) {
        const size_t gsm = sizeof(double) * SHG_ROWS * w1_ldk(RC);
        if (gsm > 48 * 1024) CUDA_TRY(cudaFuncSetAttribute(k_sh_gain_p2p, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gsm));
        k_sh_gain_p2p<<<dim3(std::max(s->nlp * SH_PH / SHG_ROWS, 1)), SHG_ROWS, gsm, st>>>(
            g, Rk, RC, s->d_PHtloc, s->d_C, s->d_u, pr, downdate_kernel_choice(Rk, 0, s->ldl) == 1);
        k_sh_barrier<<<1, 32, 0, st>>>(pr, s->rank, s->world, ++s->epoch, s->d_err);
        s->launches += 2;
    } else {
        k_sh_gain<<<(nloc + 127) / 128, 128, 0, st>>>(nloc, Rk, RC, s->d_PHtloc, s->d_C, s->d_W1send);
        const double *gathered = s->d_W1send;
        if (s->world > 1) {
            NCCL_TRY(g_nccl.AllGather(s->d_W1send, s->d_W1gath, (size_t)nch * nloc * RC, ncclFloat64, s->comm, st));
            gathered = s->d_W1gath;
        }
        k_sh_unpermute<<<(npad_used + 127) / 128, 128, 0, st>>>(g, Rk, RC, gathered, s->d_u, s->d_W1c, s->d_x);
        s->launches += 2;
    }
    CUDA_TRY(cudaGetLastError());
    if (timed) CUDA_TRY(cudaEventRecord(ev[3], st));
    s->last_k6 = downdate_kernel_choice(Rk, 0, s->ldl);
    rc = launch_downdate_map(s->d_Pl, s->ldl, rowmap_panels(s->rank, s->world, s->nlp), s->n, s->d_W1c, npad, Rk, 0, st,
                             &s->launches);
    if (timed) {
        CUDA_TRY(cudaEventRecord(ev[4], st));
        s->ev_used += 5;
    }
    return rc;
}

extern "C" int slam_shard_update_resident(slam_shard_t *s, const double *d_zf, int m, const double R[4],
                                          const int64_t *d_idf) {
    SLAM_NVTX();
    S_CHECK(s);
    ARG_CHECK(m >= 0 && R, "bad argument");
    ARG_CHECK(m == 0 || (d_zf && d_idf), "NULL device buffer");
    return shard_update_device(s, d_zf, m, R, d_idf);
}

extern "C" int slam_shard_update(slam_shard_t *s, const double *zf, int m, const double R[4], const int64_t *idf) {
    SLAM_NVTX();
    S_CHECK(s);
    ARG_CHECK(m >= 0 && R, "bad argument");
    if (m == 0) return SLAM_OK;
    ARG_CHECK(zf && idf, "NULL observation buffer");
    if (m > SLAM_MAX_OBS) {
        slam_set_error("slam_shard_update: m = %d exceeds SLAM_MAX_OBS", m);
        return SLAM_ERR_CAPACITY;
    }
    const int Nf = (s->n - 3) / 2;   // host ids: refuse a bad one before anything is launched (every rank alike)
    for (int o = 0; o < m; o++)
        if (idf[o] < 1 || idf[o] > Nf) {
            slam_set_error("slam_shard_update: idf[%d] = %lld outside 1..%d", o, (long long)idf[o], Nf);
            return SLAM_ERR_BAD_INDEX;
        }
    double *hz = (double *)((char *)s->h_pin + 256);
    int64_t *hid = (int64_t *)(hz + 6 * SLAM_MAX_OBS);
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    memcpy(hz, zf, sizeof(double) * 2 * m);
    memcpy(hid, idf, sizeof(int64_t) * m);
    CUDA_TRY(cudaMemcpyAsync(s->d_zf, hz, sizeof(double) * 2 * m, cudaMemcpyHostToDevice, s->stream));
    CUDA_TRY(cudaMemcpyAsync(s->d_idf, hid, sizeof(int64_t) * m, cudaMemcpyHostToDevice, s->stream));
    int rc = shard_update_device(s, s->d_zf, m, R, s->d_idf);
    if (rc) return rc;
    return shard_check_latched(s);
}

extern "C" int slam_shard_profile(slam_shard_t *s, int on) {
    S_CHECK(s);
    if (on && s->ev.empty()) {
        s->ev.resize(5 * 2048);
        for (auto &e : s->ev) CUDA_TRY(cudaEventCreate(&e));
    }
    s->profiling = on != 0;
    s->ev_used = 0;
    return SLAM_OK;
}

// Sums over the updates since the last read: phases[0] H-row exchange (prep, extract, allreduce or
// peer stores + barrier), [1] PHt / S / Cholesky, [2] gain + W1 exchange, [3] K6.  Also clears.
extern "C" int slam_shard_profile_phases(slam_shard_t *s, double phases_ms[4], int *updates) {
    S_CHECK(s);
    ARG_CHECK(phases_ms && updates, "NULL argument");
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    for (int p = 0; p < 4; p++) phases_ms[p] = 0;
    for (int i = 0; i + 4 < s->ev_used; i += 5)
        for (int p = 0; p < 4; p++) {
            float ms = 0;
            CUDA_TRY(cudaEventElapsedTime(&ms, s->ev[i + p], s->ev[i + p + 1]));
            phases_ms[p] += ms;
        }
    *updates = s->ev_used / 5;
    s->ev_used = 0;
    return SLAM_OK;
}

extern "C" int slam_shard_profile_read(slam_shard_t *s, double *k6_ms, int *k6_updates) {
    ARG_CHECK(k6_ms && k6_updates, "NULL argument");
    double ph[4];
    int rc = slam_shard_profile_phases(s, ph, k6_updates);
    if (rc) return rc;
    *k6_ms = ph[3];
    return SLAM_OK;
}

// ---------------------------------------------------------------------------------------
// predict (src/ekf.jl:8-43): rank 0 owns rows 0..2 of P; x is replicated
// ---------------------------------------------------------------------------------------
struct ShPredict { double g, v, w, dt, Q[4]; };

__global__ void __launch_bounds__(256)
k_sh_predict_P(const double *__restrict__ x, double *Pb0, size_t ldb, int n, ShPredict p) {
    const double phi = x[2];
    const double s = sin(p.g + phi), c = cos(p.g + phi);
    const double vts = p.v * p.dt * s, vtc = p.v * p.dt * c;
    int j = 3 + blockIdx.x * blockDim.x + threadIdx.x;
    if (j < n) {
        double *col = Pb0 + (size_t)j * ldb;
        double p0 = col[0], p1 = col[1], p2 = col[2];
        col[0] = p0 + (-vts) * p2;
        col[1] = p1 + vtc * p2;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        const double Gv[3][3] = {{1, 0, -vts}, {0, 1, vtc}, {0, 0, 1}};
        const double Gu[3][2] = {{p.dt * c, -vts}, {p.dt * s, vtc},
                                 {p.dt * sin(p.g) / p.w, p.v * p.dt * cos(p.g) / p.w}};
        double Pv[3][3], A[3][3], B[3][3], GQ[3][2];
        for (int i = 0; i < 3; i++)
            for (int q = 0; q < 3; q++) Pv[i][q] = i <= q ? Pb0[i + q * ldb] : Pb0[q + i * ldb];
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
            for (int q = i; q < 3; q++) Pb0[i + q * ldb] = B[i][q];
    }
}

__global__ void k_sh_predict_x(double *x, ShPredict p) {
    const double phi = x[2];
    const double s = sin(p.g + phi), c = cos(p.g + phi);
    double x0 = x[0] + p.v * p.dt * c, x1 = x[1] + p.v * p.dt * s;
    double x2 = mpi_to_pi(phi + p.v * p.dt * sin(p.g) / p.w);
    x[0] = x0; x[1] = x1; x[2] = x2;
}

extern "C" int slam_shard_predict(slam_shard_t *s, double g, double v, double wheelbase, const double Q[4],
                                  double dt) {
    SLAM_NVTX();
    S_CHECK(s);
    ARG_CHECK(Q != nullptr, "NULL Q");
    ShPredict p;
    p.g = g; p.v = v; p.w = wheelbase; p.dt = dt;
    memcpy(p.Q, Q, sizeof(p.Q));
    if (s->rank == 0) {
        int blocks = std::max(1, (s->n - 3 + 255) / 256);
        k_sh_predict_P<<<blocks, 256, 0, s->stream>>>(s->d_x, s->d_Pl, s->ldl, s->n, p);   // panel 0 = rank 0's local rows 0..
        s->launches++;
    }
    k_sh_predict_x<<<1, 1, 0, s->stream>>>(s->d_x, p);   // stream order: after the columns read x[2]
    s->launches++;
    CUDA_TRY(cudaGetLastError());
    return SLAM_OK;
}

// ---------------------------------------------------------------------------------------
// associate (src/data-association.jl:1-51): rows 0..2 broadcast from rank 0, every rank gates the
// landmarks whose 2x2 diagonal block it owns, (nd, j, min nis) per observation allgathered
// ---------------------------------------------------------------------------------------
__global__ void k_sh_rows012(const double *__restrict__ Pb0, size_t ldb, int n, int npad, double *P3) {
    int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= npad) return;
    for (int p = 0; p < 3; p++) P3[(size_t)p * npad + j] = (j < n && j >= p) ? Pb0[p + (size_t)j * ldb] : 0.0;
}

extern "C" int slam_shard_associate(slam_shard_t *s, const double *z, int nz, const double R[4], double gate1,
                                    double gate2, double *zf, int64_t *idf, int *m, double *zn, int *k) {
    SLAM_NVTX();
    S_CHECK(s);
    ARG_CHECK(R && m && k && nz >= 0, "bad argument");
    *m = 0;
    *k = 0;
    if (nz == 0) return SLAM_OK;
    ARG_CHECK(z && zf && idf && zn, "NULL observation buffer");
    if (nz > SLAM_MAX_OBS) {
        slam_set_error("slam_shard_associate: nz = %d exceeds SLAM_MAX_OBS", nz);
        return SLAM_ERR_CAPACITY;
    }
    if (!s->d_gate) {
        s->gate_bytes = gate_scratch_bytes(SLAM_MAX_OBS) + (size_t)GATE_RES_BYTES * SLAM_MAX_OBS * (s->world + 1) + 64;
        CUDA_TRY(cudaMalloc(&s->d_gate, s->gate_bytes));
        CUDA_TRY(cudaMemsetAsync(s->d_gate, 0, 256, s->stream));
    }
    char *res_mine = (char *)s->d_gate + gate_scratch_bytes(SLAM_MAX_OBS);
    char *res_all = res_mine + (size_t)GATE_RES_BYTES * SLAM_MAX_OBS;
    int *d_counts = (int *)(res_all + (size_t)GATE_RES_BYTES * SLAM_MAX_OBS * s->world);
    cudaStream_t st = s->stream;
    double *hz = (double *)((char *)s->h_pin + 256);
    CUDA_TRY(cudaStreamSynchronize(st));
    memcpy(hz, z, sizeof(double) * 2 * nz);
    CUDA_TRY(cudaMemcpyAsync(s->d_z, hz, sizeof(double) * 2 * nz, cudaMemcpyHostToDevice, st));
    if (s->rank == 0) {
        k_sh_rows012<<<(s->npad + 255) / 256, 256, 0, st>>>(s->d_Pl, s->ldl, s->n, s->npad, s->d_P3);
        s->launches++;
    }
    if (s->world > 1) NCCL_TRY(g_nccl.Broadcast(s->d_P3, s->d_P3, (size_t)3 * s->npad, ncclFloat64, 0, s->comm, st));
    int rc = launch_gate_shard(s->d_x, s->d_P3, s->npad, s->d_Pl, s->ldl, s->rank, s->world, s->nlp, s->n, s->d_z, nz, R, gate1, gate2,
                               s->d_gate, gate_scratch_bytes(SLAM_MAX_OBS), res_mine,
                               s->pregate ? s->d_sig2 : nullptr, st, &s->launches);
    if (rc) return rc;
    const void *all = res_mine;
    if (s->world > 1) {
        NCCL_TRY(g_nccl.AllGather(res_mine, res_all, (size_t)GATE_RES_BYTES * nz, ncclChar, s->comm, st));
        all = res_all;
    }
    rc = launch_gate_finalize(all, s->world, s->d_z, nz, gate2, s->d_zf, s->d_idf, s->d_zn, d_counts, st,
                              &s->launches);
    if (rc) return rc;
    int *hc = (int *)s->h_pin;
    double *hzf = hz + 2 * SLAM_MAX_OBS, *hzn = hzf + 2 * SLAM_MAX_OBS;
    int64_t *hid = (int64_t *)(hzn + 2 * SLAM_MAX_OBS);
    CUDA_TRY(cudaMemcpyAsync(hc, d_counts, sizeof(int) * 2, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaMemcpyAsync(hzf, s->d_zf, sizeof(double) * 2 * nz, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaMemcpyAsync(hzn, s->d_zn, sizeof(double) * 2 * nz, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaMemcpyAsync(hid, s->d_idf, sizeof(int64_t) * nz, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    *m = hc[0];
    *k = hc[1];
    memcpy(zf, hzf, sizeof(double) * 2 * (*m));
    memcpy(idf, hid, sizeof(int64_t) * (*m));
    memcpy(zn, hzn, sizeof(double) * 2 * (*k));
    return SLAM_OK;
}

// ---------------------------------------------------------------------------------------
// add_features (src/ekf.jl:84-122) on the sharded covariance.  The new columns are local on every
// rank (a block stores its rows of EVERY column); what a rank lacks are rows 0..2 of the existing
// columns, which rank 0 broadcasts (the same 24 n bytes associate uses).  Every rank then fills its
// own rows of the 2k new columns; the owner of the new rows writes the 2x2 diagonal blocks.  x is
// replicated and every rank appends the same values.  Same expressions as the single-filter K7
// (filter.cu k_augment), so x and P stay bit-identical to the single-filter path.
// ---------------------------------------------------------------------------------------
constexpr int SHA_Q = 32;        // new landmarks per launch
constexpr int SHA_THREADS = 256;

// grid.x = local row tiles of SHA_THREADS rows (one panel each), grid.y = new landmark q of this launch
__global__ void __launch_bounds__(SHA_THREADS)
k_sh_augment(double *x, const double *__restrict__ P3, int npad, double *Pl, size_t ldl, int rank, int world, int n0,
             const double *__restrict__ zn, int kq, double R0, double R1, double R2, double R3,
             unsigned long long *sig2bits) {
    __shared__ double sGv[SHA_Q][2];
    __shared__ double sGP[SHA_Q][2][3];
    __shared__ double sPvv[9];
    if (threadIdx.x < 9) {
        const int r = threadIdx.x / 3, c = threadIdx.x % 3;
        sPvv[threadIdx.x] = r <= c ? P3[(size_t)r * npad + c] : P3[(size_t)c * npad + r];
    }
    __syncthreads();
    const double xv = x[0], yv = x[1], phi = x[2];
    if (threadIdx.x < kq) {
        const int q = threadIdx.x;
        const double r = zn[2 * q], b = zn[2 * q + 1];
        const double s = sin(phi + b), c = cos(phi + b);
        const double g0 = -r * s, g1 = r * c;
        sGv[q][0] = g0;
        sGv[q][1] = g1;
        const double Gv[2][3] = {{1, 0, g0}, {0, 1, g1}};
        for (int a = 0; a < 2; a++)
            for (int j = 0; j < 3; j++)
                sGP[q][a][j] = Gv[a][0] * sPvv[0 * 3 + j] + Gv[a][1] * sPvv[1 * 3 + j] + Gv[a][2] * sPvv[2 * 3 + j];
        const int len = n0 + 2 * q;
        if (blockIdx.x == 0 && blockIdx.y == q) {
            x[len] = xv + r * c;        // x is replicated: every rank appends the same values
            x[len + 1] = yv + r * s;
            // the 2x2 diagonal block, by the owner of row len (a landmark's two rows never straddle panels)
            const int p = panel_of_row(len);
            if (panel_owner(p, world) == rank) {
                const size_t li = (size_t)panel_local(p, world) * SH_PH + (len - panel_first(p));
                const double Gz[2][2] = {{c, -r * s}, {s, r * c}};
                const double Rm[4] = {R0, R1, R2, R3};
                double GR[2][2];
                for (int a = 0; a < 2; a++)
                    for (int j = 0; j < 2; j++) GR[a][j] = Gz[a][0] * Rm[0 + 2 * j] + Gz[a][1] * Rm[1 + 2 * j];
                for (int a = 0; a < 2; a++)
                    for (int bb = a; bb < 2; bb++) {
                        const double t1 = sGP[q][a][0] * Gv[bb][0] + sGP[q][a][1] * Gv[bb][1] + sGP[q][a][2] * Gv[bb][2];
                        const double t2 = GR[a][0] * Gz[bb][0] + GR[a][1] * Gz[bb][1];
                        const double pv = t1 + t2;
                        Pl[li + a + (size_t)(len + bb) * ldl] = pv;
                        if (a == bb) {   // keep this rank's pre-gate bound valid for the landmark it now owns
                            unsigned long long bits = (unsigned long long)__double_as_longlong(pv);
                            if (pv != pv) bits = 0x7ff8000000000000ull;
                            if (!(pv < 0.0)) atomicMax(sig2bits, bits);
                        }
                    }
            }
        }
    }
    __syncthreads();
    const int q = blockIdx.y;
    if (q >= kq) return;
    const int len = n0 + 2 * q;
    const int li = blockIdx.x * SHA_THREADS + threadIdx.x;      // a local row of this rank
    const int p = panel_global(li / SH_PH, rank, world);
    const int j = panel_first(p) + li % SH_PH;                   // its global row
    if (j >= len || j >= panel_first(p + 1)) return;             // rows of the state before landmark q only
    double c0, c1, c2;   // P[1:3, j]
    if (j < 3) {
        c0 = sPvv[0 * 3 + j]; c1 = sPvv[1 * 3 + j]; c2 = sPvv[2 * 3 + j];
    } else if (j < n0) {
        c0 = P3[j]; c1 = P3[(size_t)npad + j]; c2 = P3[(size_t)2 * npad + j];
    } else {   // a landmark added earlier in this launch
        const int qq = (j - n0) / 2, a = (j - n0) % 2;
        c0 = sGP[qq][a][0]; c1 = sGP[qq][a][1]; c2 = sGP[qq][a][2];
    }
    const double g0 = sGv[q][0], g1 = sGv[q][1];
    const double t0 = 1.0 * c0 + 0.0 * c1 + g0 * c2;
    const double t1 = 0.0 * c0 + 1.0 * c1 + g1 * c2;
    Pl[(size_t)li + (size_t)len * ldl] = t0;
    Pl[(size_t)li + (size_t)(len + 1) * ldl] = t1;
}

extern "C" int slam_shard_add_features(slam_shard_t *s, const double *zn, int k, const double R[4]) {
    SLAM_NVTX();
    S_CHECK(s);
    ARG_CHECK(k >= 0 && R, "bad argument");
    if (k == 0) return SLAM_OK;
    ARG_CHECK(zn != nullptr, "NULL zn");
    if (k > SLAM_MAX_OBS || s->n + 2 * k > s->n_max) {
        slam_set_error("slam_shard_add_features: n + 2k = %d exceeds capacity n_max = %d", s->n + 2 * k, s->n_max);
        return SLAM_ERR_CAPACITY;
    }
    cudaStream_t st = s->stream;
    double *hz = (double *)((char *)s->h_pin + 256);
    CUDA_TRY(cudaStreamSynchronize(st));
    memcpy(hz, zn, sizeof(double) * 2 * k);
    CUDA_TRY(cudaMemcpyAsync(s->d_zn, hz, sizeof(double) * 2 * k, cudaMemcpyHostToDevice, st));
    for (int q0 = 0; q0 < k; q0 += SHA_Q) {
        const int kq = std::min(SHA_Q, k - q0);
        const int n0 = s->n, n1 = n0 + 2 * kq;
        if (s->rank == 0) {
            k_sh_rows012<<<(s->npad + 255) / 256, 256, 0, st>>>(s->d_Pl, s->ldl, n0, s->npad, s->d_P3);
            s->launches++;
        }
        if (s->world > 1) NCCL_TRY(g_nccl.Broadcast(s->d_P3, s->d_P3, (size_t)3 * s->npad, ncclFloat64, 0, s->comm, st));
        const int nlp1 = local_panels(n1, s->world, s->rank);    // panels once the new rows exist
        dim3 grid(std::max(nlp1 * SH_PH / SHA_THREADS, 1), kq);
        k_sh_augment<<<grid, SHA_THREADS, 0, st>>>(s->d_x, s->d_P3, s->npad, s->d_Pl, s->ldl, s->rank, s->world, n0,
                                                   s->d_zn + 2 * q0, kq, R[0], R[1], R[2], R[3], s->d_sig2);
        s->launches++;
        s->n = n1;
        s->nlp = nlp1;
    }
    CUDA_TRY(cudaGetLastError());
    return SLAM_OK;
}
