"""Full-size parity of the sharded update (torchrun, W GPUs): bench.py's workload (40 GB of P per GPU,
n = 199999 at W = 8) is too large for a single-filter comparison, but H is zero outside the vehicle and
the observed landmarks, so the update restricted to a principal submatrix that contains those columns IS
the update of that sub-state.  The ranks assemble a sample (vehicle + observed pairs + four 150-row
windows) of P before and after one sharded update, and rank 0 runs the numpy literal oracle on it.
usage: python -m torch.distributed.run --nproc-per-node W tools/shard_full_check.py [gb_per_gpu] [m]"""
import math
import os
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "slam.jl_b200"))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
from oracle import slam_ref as ref  # noqa: E402  (checker only: this is a test tool)
from slamcuda._ffi import lib, check, dptr, iptr  # noqa: E402
from slamcuda.shard import ShardedFilter  # noqa: E402

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("gloo")
gb = float(sys.argv[1]) if len(sys.argv) > 1 else 40.0
m = int(sys.argv[2]) if len(sys.argv) > 2 else 8
n = int(math.sqrt(gb * world * 1e9 / 8.0))
n -= (n - 3) % 2
nl = (n - 3) // 2
side = math.sqrt(nl / 0.0035)
R = np.array([[0.01, 0.0], [0.0, (math.pi / 180) ** 2]])
sf = ShardedFilter(lr, rank, world, n)
sf.synth_state(nl, side, 7)
z = np.zeros(2 * (m + 4))
ids = np.zeros(m, dtype=np.int64)
check(lib.slam_synth_observations(None, nl, side, 7, m, 4, dptr(z), iptr(ids)))
z2 = z.reshape(m + 4, 2).T.copy()
z2[0, :m] += 0.05      # the synthetic observations are noise-free: give the update an innovation
z2[1, :m] += 0.003
zf, idf, zn = sf.associate(z2, R, 4.0, 25.0)
seen = sorted({int(i) for i in idf})
segs = [(0, 3)] + [(3 + 2 * (i - 1), 2) for i in seen]
taken = {3 + 2 * (i - 1) for i in seen}
for w0 in (3, n // 3, 2 * n // 3, n - 150):
    w0 = w0 if (w0 - 3) % 2 == 0 else w0 + 1
    while any(w0 - 2 <= t < w0 + 150 for t in taken):
        w0 = w0 + 200 if w0 + 400 < n else w0 - 200
    segs.append((w0, 150))
index = np.concatenate([np.arange(s, s + ln) for s, ln in segs])


def gather():
    """Principal submatrix over the segments: every rank contributes the rows it owns."""
    tot = len(index)
    out = np.zeros((tot, tot))
    owned = np.zeros((tot, tot))
    oi = 0
    for i0, ni in segs:
        oj = 0
        for j0, nj in segs:
            blk, _cnt = sf.cov_block(i0, ni, j0, nj)
            mask = ~np.isnan(blk)
            out[oi:oi + ni, oj:oj + nj] = np.where(mask, blk, 0.0)
            owned[oi:oi + ni, oj:oj + nj] = mask
            oj += nj
        oi += ni
    t = torch.from_numpy(np.stack([out, owned]))
    dist.all_reduce(t)
    out, owned = t[0].numpy(), t[1].numpy()
    assert np.all(owned == 1.0), "every entry must be owned by exactly one rank"
    # only entries with (global) row <= column are authoritative: mirror them
    upper = index[:, None] <= index[None, :]
    return np.where(upper, out, out.T)


x0 = sf.get_x()
P0 = gather()
sf.update(zf, R, idf)
x1 = sf.get_x()
P1 = gather()
if rank == 0:
    pos = {i: 1 + c for c, i in enumerate(seen)}
    ids_sub = np.array([pos[int(i)] for i in idf])
    xo, Po = ref.update(x0[index], P0, zf, R, ids_sub)
    ex = float(np.max(np.abs(x1[index] - xo)) / np.max(np.abs(xo)))
    eP = float(np.max(np.abs(P1 - Po)) / np.max(np.abs(Po)))
    ok = ex < 1e-9 and eP < 1e-9 and len(idf) == m
    print(f"SHARD_FULL_CHECK {'PASS' if ok else 'FAIL'} world={world} n={n} P={8.0 * n * n / 1e9:.0f} GB m={len(idf)} "
          f"p2p={sf.p2p} sampled {len(index)} states: x rel {ex:.2e}  P rel {eP:.2e}")
sf.close()
dist.destroy_process_group()
