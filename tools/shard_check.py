"""torchrun check of the sharded path on W GPUs against the single-filter path on rank 0.
usage: python -m torch.distributed.run --nproc-per-node W tools/shard_check.py [nl] [m]"""
import math
import os
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "slam.jl_b200"))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
from slamcuda._ffi import lib, check, dptr, iptr  # noqa: E402
from slamcuda.shard import ShardedFilter, shard_plan  # noqa: E402
from slamcuda.slam import _Handle  # noqa: E402

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("gloo")
nl = int(sys.argv[1]) if len(sys.argv) > 1 else 3000
m = int(sys.argv[2]) if len(sys.argv) > 2 else 8
n = 3 + 2 * nl
side = math.sqrt(nl / 0.0035)
R = np.array([[0.01, 0.0], [0.0, (math.pi / 180) ** 2]])
Q = np.array([[0.25, 0.0], [0.0, (3 * math.pi / 180) ** 2]])
knew = 40
ncap = n + 2 * knew
sf = ShardedFilter(lr, rank, world, ncap)
sf.synth_state(nl, side, 7)
sf.guard_arm()              # padding guard of this rank's panel stack and x (slam_shard_debug_guard)
guard = [sf.guard_changed()]
h = _Handle(lr, ncap)
check(lib.slam_synth_state(h.ptr, nl, side, 7))
nz = m + knew
z = np.zeros(2 * nz)
ids = np.zeros(m, dtype=np.int64)
check(lib.slam_synth_observations(h.ptr, nl, side, 7, m, knew, dptr(z), iptr(ids)))
z2 = z.reshape(nz, 2).T
zf_a, idf_a, zn_a = h.associate(z2, R, 4.0, 25.0)
zf_b, idf_b, zn_b = sf.associate(z2, R, 4.0, 25.0)
ok = np.array_equal(idf_a, idf_b) and np.array_equal(zf_a, zf_b) and np.array_equal(zn_a, zn_b)
for _ in range(2):
    h.predict(0.03, 7.7, 4.0, Q, 0.025)
    sf.predict(0.03, 7.7, 4.0, Q, 0.025)
    h.update(zf_a, R, idf_a)
    sf.update(zf_a, R, idf_a)
guard.append(sf.guard_changed())     # after associate + 2 x (predict, update), 2 * knew states of padding still spare
zn_all = z2[:, m:].copy()
h.add_features(zn_all, R)
sf.add_features(zn_all, R)
kk = knew
ok = ok and sf.n == h.n == n + 2 * kk
idn = np.array([nl + 1, nl + kk, 1], dtype=np.int64)
zq = np.stack([h.predict_observation(int(i))[0] for i in idn], axis=1) + 0.01
h.predict(0.03, 7.7, 4.0, Q, 0.025)
sf.predict(0.03, 7.7, 4.0, Q, 0.025)
h.update(zq, R, idn)
sf.update(zq, R, idn)
# a high-rank update (R = 80): the workspace grows, so the peer-visible arena is re-allocated and
# re-shared collectively, and K4 takes its wide variant
idw = np.array([1 + (7 * c) % nl for c in range(40)], dtype=np.int64)
zw = np.stack([h.predict_observation(int(i))[0] for i in idw], axis=1) + 0.02
h.update(zw, R, idw)
sf.update(zw, R, idw)
# association again on the grown, updated map: the rank-local bound of the spatial pre-gate has followed
# add_features on whichever rank owns the new landmarks
idq = np.array([nl + 1, nl + 3, nl + kk, 2, 5], dtype=np.int64)
zq2 = np.stack([h.predict_observation(int(i))[0] for i in idq], axis=1) + 0.005
zq2 = np.concatenate([zq2, np.array([[3 * side, 0.3], [2.5 * side, -1.0]]).T], axis=1)
a2 = h.associate(zq2, R, 4.0, 25.0)
b2 = sf.associate(zq2, R, 4.0, 25.0)
ok = ok and all(np.array_equal(p, q) for p, q in zip(a2, b2)) and len(a2[1]) >= 3 and a2[2].shape[1] == 2
guard.append(sf.guard_changed())     # after add_features, the rank-80 update (arena re-shared) and the second associate
n = sf.n
x1, P1 = h.get_state()
ex = float(np.max(np.abs(sf.get_x() - x1)) / np.max(np.abs(x1)))
eP = 0.0
cnt_total = 0
for lo, hi in shard_plan(n, world, rank):
    if hi > lo:
        blk, cnt = sf.cov_block(lo, hi - lo, 0, n)
        cnt_total += cnt
        ref = P1[lo:hi, :]
        mask = np.arange(n)[None, :] >= np.arange(lo, hi)[:, None]
        eP = max(eP, float(np.max(np.abs(blk - ref)[mask]) / np.max(np.abs(P1))))
res = [None] * world
dist.all_gather_object(res, (rank, ok, ex, eP, cnt_total, len(shard_plan(n, world, rank)), sum(guard)))
if rank == 0:
    for r in res:
        print("rank %d assoc_ok=%s x_rel=%.2e P_rel=%.2e rows*cols=%d panels=%d padding_slots_overwritten=%d" % r)
    good = all(r[1] and r[2] < 1e-12 and r[3] < 1e-12 and r[6] == 0 for r in res)
    print("SHARD_CHECK", "PASS" if good else "FAIL", f"world={world} n={n} m={m} p2p={sf.p2p}")
sf.close()
h.close()
dist.destroy_process_group()
