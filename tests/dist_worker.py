"""Worker of tests/test_dist_cpu.py: one of W gloo ranks on CPU (launched by torch.distributed.run).
Checks the host-side logic of the two multi-GPU modes without a device:
  plan      the row-panel partition of the sharded covariance (slam_shard_plan) over the ranks
  uid       the NCCL unique id handed from rank 0 to every rank through torch.distributed
  members   the ensemble's run-id split, with the oracle's C twin standing in for the kernel:
            the ranks' members together equal one process running all of them
Rank 0 prints one line `DIST_CHECK <json>`."""
import ctypes as C
import json
import os
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "slam.jl_b200"))
sys.path.insert(0, str(ROOT / "tests"))
import torch.distributed as dist  # noqa: E402
from slamcuda.ensemble import member_range  # noqa: E402
from slamcuda.shard import exchange_unique_id, shard_plan  # noqa: E402
from twin_ffi import d, load_twin  # noqa: E402

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo")
out = {}

# -- plan ---------------------------------------------------------------------------------
plans = {}
for n in (3, 5, 255, 257, 259, 1003, 6083, 20003, 60001, 120001, 200003):   # incl. 0.1 / 0.3 / 0.6 / 1.0 of 200003
    mine = shard_plan(n, world, rank)
    got = [None] * world
    dist.all_gather_object(got, mine)
    plans[n] = got
out["plans"] = plans

# -- uid ----------------------------------------------------------------------------------
uid = exchange_unique_id(rank, world)
got = [None] * world
dist.all_gather_object(got, uid.hex())
out["uids"] = got

# -- members ------------------------------------------------------------------------------
twin = load_twin()
twin.twin_ensemble_cap.restype = C.c_int
twin.twin_ensemble_cap.argtypes = twin.twin_ensemble.argtypes + [C.c_int]
wp2 = np.loadtxt(ROOT / "tests" / "golden" / "course1.txt", skiprows=1).T.copy()
nl, total, seed = 12, 7, 99
lm = np.random.Generator(np.random.PCG64(5)).integers(5, 96, size=(2, nl)).astype(float)
wp = np.ascontiguousarray(wp2.T).reshape(-1)
lmf = np.ascontiguousarray(lm.T).reshape(-1)


def run_members(run0, count):
    fp = np.zeros(3 * max(count, 1))
    nfin = np.zeros(max(count, 1), dtype=np.int32)
    if count:
        twin.twin_ensemble_cap(d(wp), wp2.shape[1], d(lmf), nl, seed, run0, count, 1, 1000000, d(fp), None,
                               nfin.ctypes.data_as(C.POINTER(C.c_int)), 2, 5)
    return fp[:3 * count].tolist(), nfin[:count].tolist()


r0, cnt = member_range(total, world, rank)
got = [None] * world
dist.all_gather_object(got, (r0, cnt) + run_members(r0, cnt))
out["members"] = got
if rank == 0:
    out["members_one_process"] = run_members(0, total)
    print("DIST_CHECK " + json.dumps(out), flush=True)
dist.barrier()
dist.destroy_process_group()
