"""Large-map mode: one filter whose covariance is row-block-sharded over the ranks of a
torch.distributed job (one process per GPU).  torch.distributed is only the launcher-side
plumbing (it hands the NCCL unique id to every rank); the collectives on the data path are
issued by libslamcuda itself on its own stream."""
from __future__ import annotations

import ctypes as C

import numpy as np

from ._ffi import lib, check, dptr, iptr


def shard_plan(n: int, world: int, rank: int):
    """Rows [lo, hi) of every 256-row panel rank owns at state size n, in the order the rank stacks them (pure
    arithmetic, no device)."""
    cnt = check(lib.slam_shard_plan(n, world, rank, None, None, 0))
    lo = (C.c_int * max(cnt, 1))()
    hi = (C.c_int * max(cnt, 1))()
    check(lib.slam_shard_plan(n, world, rank, lo, hi, cnt))
    return [(lo[k], hi[k]) for k in range(cnt)]


def exchange_unique_id(rank: int, world: int) -> bytes:
    """Rank 0 creates the NCCL unique id; torch.distributed (any backend) broadcasts it."""
    if world == 1:
        return b"\0" * 128
    import torch.distributed as dist
    box = [None]
    if rank == 0:
        buf = C.create_string_buffer(128)
        check(lib.slam_shard_unique_id(buf))
        box[0] = buf.raw
    dist.broadcast_object_list(box, src=0)
    return box[0]


class ShardedFilter:
    def __init__(self, device: int, rank: int, world: int, n_max: int, unique_id: bytes | None = None):
        self.rank, self.world, self.device = rank, world, device
        uid = unique_id if unique_id is not None else exchange_unique_id(rank, world)
        self.ptr = C.c_void_p()
        check(lib.slam_shard_create(device, rank, world, uid, n_max, C.byref(self.ptr)))

    def close(self):
        if self.ptr:
            lib.slam_shard_destroy(self.ptr)
            self.ptr = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def n(self) -> int:
        return check(lib.slam_shard_n(self.ptr))

    @property
    def p2p(self) -> bool:
        """True when the update exchanges rows / W1 through peer memory instead of NCCL collectives."""
        return check(lib.slam_shard_p2p(self.ptr)) == 1

    def guard_arm(self):
        """Paint this rank's panel stack and x outside the live state (slam_shard_debug_guard)."""
        check(lib.slam_shard_debug_guard(self.ptr, 1, None))

    def guard_changed(self) -> int:
        """How many painted padding slots of this rank have been overwritten since guard_arm()."""
        c = C.c_longlong(-1)
        check(lib.slam_shard_debug_guard(self.ptr, 0, C.byref(c)))
        return int(c.value)

    def synth_state(self, n_landmarks: int, side: float, seed: int):
        check(lib.slam_shard_synth_state(self.ptr, n_landmarks, float(side), seed))

    def get_x(self):
        x = np.empty(self.n)
        check(lib.slam_shard_get_x(self.ptr, dptr(x)))
        return x

    def get_pose(self):
        p = np.empty(3)
        check(lib.slam_shard_get_pose(self.ptr, dptr(p)))
        return p

    def cov_block(self, i0, ni, j0, nj):
        """Window of P with NaN where this rank does not own the row; returns (block, count)."""
        out = np.full((ni, nj), np.nan, order="F")
        cnt = check(lib.slam_shard_get_cov_block(self.ptr, i0, ni, j0, nj, dptr(out), max(ni, 1)))
        return np.ascontiguousarray(out), cnt

    def predict(self, g, v, w, Q, dt):
        q = np.ascontiguousarray(np.asarray(Q, dtype=float).reshape(2, 2).T).reshape(-1)
        check(lib.slam_shard_predict(self.ptr, float(g), float(v), float(w), dptr(q), float(dt)))

    def associate(self, z, R, gate1, gate2):
        z = np.asarray(z, dtype=float).reshape(2, -1)
        nz = z.shape[1]
        zc = np.ascontiguousarray(z.T).reshape(-1)
        r = np.ascontiguousarray(np.asarray(R, dtype=float).reshape(2, 2).T).reshape(-1)
        zf, zn = np.zeros(2 * max(nz, 1)), np.zeros(2 * max(nz, 1))
        idf = np.zeros(max(nz, 1), dtype=np.int64)
        m, k = C.c_int(), C.c_int()
        check(lib.slam_shard_associate(self.ptr, dptr(zc), nz, dptr(r), float(gate1), float(gate2), dptr(zf),
                                       iptr(idf), C.byref(m), dptr(zn), C.byref(k)))
        m, k = m.value, k.value
        return zf[:2 * m].reshape(m, 2).T.copy(), idf[:m].copy(), zn[:2 * k].reshape(k, 2).T.copy()

    def update(self, zf, R, idf):
        zf = np.asarray(zf, dtype=float).reshape(2, -1)
        m = zf.shape[1]
        zc = np.ascontiguousarray(zf.T).reshape(-1)
        ids = np.ascontiguousarray(np.asarray(idf, dtype=np.int64).reshape(-1))
        r = np.ascontiguousarray(np.asarray(R, dtype=float).reshape(2, 2).T).reshape(-1)
        check(lib.slam_shard_update(self.ptr, dptr(zc), m, dptr(r), iptr(ids)))

    def add_features(self, zn, R):
        """src/ekf.jl:84-122 on the sharded covariance (collective: every rank calls it)."""
        zn = np.asarray(zn, dtype=float).reshape(2, -1)
        k = zn.shape[1]
        zc = np.ascontiguousarray(zn.T).reshape(-1)
        r = np.ascontiguousarray(np.asarray(R, dtype=float).reshape(2, 2).T).reshape(-1)
        check(lib.slam_shard_add_features(self.ptr, dptr(zc), k, dptr(r)))

    def sync(self):
        check(lib.slam_shard_sync(self.ptr))

    def launch_count(self) -> int:
        return int(lib.slam_shard_launch_count(self.ptr))
