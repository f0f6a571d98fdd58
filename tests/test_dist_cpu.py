"""world_size-2 (and 3) `gloo` tests of the host logic behind the two multi-GPU modes, on CPU:
the row-block partition of the sharded covariance, the NCCL-unique-id hand-off, the ensemble's
run-id split, and the bench contract's reference arm under torch.distributed.run (rank 0 alone
works and prints).  The device side of the same paths is tests/test_gpu_shard.py and
tools/shard_check.py (torchrun on 2+ GPUs)."""
import json
import os
import socket
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent


def free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def torchrun(world: int, script: str, *args: str, timeout: int = 600):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(free_port()), script, *args]
    env = dict(os.environ, OMP_NUM_THREADS="2")
    p = subprocess.run(cmd, cwd=ROOT, env=env, capture_output=True, text=True, timeout=timeout)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-4000:]
    return p.stdout


@pytest.fixture(scope="module", params=[2, 3])
def dist_result(request):
    world = request.param
    out = torchrun(world, "tests/dist_worker.py")
    lines = [ln for ln in out.splitlines() if ln.startswith("DIST_CHECK ")]
    assert len(lines) == 1, out[-2000:]
    return world, json.loads(lines[0][len("DIST_CHECK "):])


def test_row_panel_partition(dist_result):
    """256-row panels dealt boustrophedon (panel p -> rank b = p mod 2W if b < W else 2W-1-b): rows covered exactly
    once, panel edges never split a landmark's two rows, and the upper-triangle work per rank is balanced at ANY
    state size -- 10 %, 30 %, 60 % and 100 % of a 100k-landmark capacity -- not only when the map is full (the
    reference's state grows one landmark at a time, src/ekf.jl:84-122)."""
    world, res = dist_result
    for n_str, plans in res["plans"].items():
        n = int(n_str)
        owner = np.full(n, -1)
        work = np.zeros(world)
        for r, panels in enumerate(plans):
            for lo, hi in panels:
                if hi <= lo:
                    continue
                assert 0 <= lo < hi <= n and hi - lo <= 256
                assert np.all(owner[lo:hi] == -1), (n, r, lo, hi)
                owner[lo:hi] = r
                if lo > 0:
                    assert lo >= 3 and (lo - 3) % 2 == 0, (n, lo)      # lo is a landmark's first row
                work[r] += sum(n - i for i in (lo, hi - 1)) * (hi - lo) / 2.0   # entries j >= i
        assert np.all(owner >= 0), n
        if n >= 20003:
            assert work.max() / work.mean() < 1.05, (n, work)


def test_unique_id_handoff(dist_result):
    world, res = dist_result
    uids = res["uids"]
    assert len(uids) == world and len(set(uids)) == 1
    assert len(bytes.fromhex(uids[0])) == 128 and any(bytes.fromhex(uids[0]))


def test_member_split_is_result_neutral(dist_result):
    world, res = dist_result
    poses, nfin, cover = [], [], []
    for r0, cnt, fp, nf in res["members"]:
        cover += list(range(r0, r0 + cnt))
        poses += fp
        nfin += nf
    one_fp, one_nf = res["members_one_process"]
    assert cover == list(range(len(one_nf)))
    assert poses == one_fp and nfin == one_nf     # bit-equal: JSON round-trips doubles exactly


def test_member_range_edges():
    sys.path.insert(0, str(ROOT / "slam.jl_b200"))
    from slamcuda.ensemble import member_range
    for total in (0, 1, 7, 8, 8192, 8193):
        for world in (1, 2, 3, 8):
            spans = [member_range(total, world, r) for r in range(world)]
            assert spans[0][0] == 0 and sum(c for _, c in spans) == total
            assert all(spans[i][0] + spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            assert max(c for _, c in spans) - min(c for _, c in spans) <= 1
    with pytest.raises(ValueError):
        member_range(8, 2, 2)


def test_reference_arm_prints_once_under_torchrun():
    """bench.py --impl reference at N = 2: rank 0 alone runs and prints, the other rank exits 0."""
    out = torchrun(2, "bench.py", "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "1",
                   "--sample-landmarks", "300")
    lines = [json.loads(ln) for ln in out.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    line = lines[0]
    assert line["impl"] == "reference" and line["n_gpus"] == 2 and line["value"] > 0
    # torch.distributed.run exports OMP_NUM_THREADS=1 to every rank and numpy's OpenBLAS obeys it at import: the arm must
    # put the BLAS back on every host core (and report the count it really has), or it times a one-thread CPU
    import os
    assert line["cpu_baseline"]["cores"] == (os.cpu_count() or 1), line["cpu_baseline"]["cores"]
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["cpu_baseline"]["kind"] == "port"


def test_cpu_legs_get_every_core_despite_omp_num_threads():
    """The behaviour behind the assertion above: in a process that inherits OMP_NUM_THREADS=1 (what torch.distributed.run
    exports) numpy's BLAS has ONE thread; bench.cpu_use_all_cores() puts it on every host core and returns the count the
    BLAS then really has.  (Before this helper existed the reference arm at N > 1 timed a one-thread CPU and reported
    os.cpu_count() cores.)"""
    import os
    import subprocess
    import sys
    code = ("import json, os, numpy, bench\n"
            "from threadpoolctl import threadpool_info\n"
            "blas = lambda: [i['num_threads'] for i in threadpool_info() if i.get('user_api') == 'blas']\n"
            "before = blas(); n = bench.cpu_use_all_cores(); after = blas()\n"
            "print(json.dumps({'before': before, 'returned': n, 'after': after, 'cpus': os.cpu_count()}))\n")
    env = dict(os.environ, OMP_NUM_THREADS="1")
    env.pop("OPENBLAS_NUM_THREADS", None)
    out = subprocess.run([sys.executable, "-c", code], cwd=ROOT, env=env, capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stderr[-800:]
    r = json.loads(out.stdout.strip().splitlines()[-1])
    cpus = r["cpus"] or 1
    assert r["before"] and all(t == 1 for t in r["before"]), r          # the inherited setting does throttle the BLAS
    assert r["returned"] == cpus and r["after"] and all(t == cpus for t in r["after"]), r
