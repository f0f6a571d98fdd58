"""GPU parity of the row-panel-sharded path.  world = 1 exercises the whole sharded code path
(256-row panels, rank-local buffers, unpermute, panel-mapped K6, sharded gating) on one device
without NCCL; tools/shard_check.py runs the same comparison under torchrun on 2+ GPUs."""
import ctypes as C
import math

import numpy as np
import pytest

from conftest import rel_err

pytestmark = pytest.mark.gpu

R = np.array([[0.01, 0.0], [0.0, (math.pi / 180) ** 2]])
Q = np.array([[0.25, 0.0], [0.0, (3 * math.pi / 180) ** 2]])


def upper(P):
    return np.triu(P)


@pytest.mark.parametrize("nl,m", [(700, 5), (1500, 8), (1500, 20)])
def test_sharded_world1_matches_single_filter(nl, m):
    import slamcuda
    from slamcuda._ffi import lib, check, dptr, iptr
    from slamcuda.shard import ShardedFilter, shard_plan
    from slamcuda.slam import _Handle
    n = 3 + 2 * nl
    side = math.sqrt(nl / 0.0035)
    h = _Handle(0, n)
    check(lib.slam_synth_state(h.ptr, nl, side, 7))
    sf = ShardedFilter(0, 0, 1, n)
    sf.synth_state(nl, side, 7)
    x0, P0 = h.get_state()
    assert np.array_equal(sf.get_x(), x0)
    blk, cnt = sf.cov_block(0, n, 0, n)
    assert cnt == n * n
    assert np.array_equal(np.triu(blk), np.triu(P0))
    # observations
    nz = m + 4
    z = np.zeros(2 * nz)
    ids = np.zeros(m, dtype=np.int64)
    check(lib.slam_synth_observations(h.ptr, nl, side, 7, m, 4, dptr(z), iptr(ids)))
    z2 = z.reshape(nz, 2).T
    zf_a, idf_a, zn_a = h.associate(z2, R, 4.0, 25.0)
    zf_b, idf_b, zn_b = sf.associate(z2, R, 4.0, 25.0)
    assert np.array_equal(idf_a, idf_b) and np.array_equal(zf_a, zf_b) and np.array_equal(zn_a, zn_b)
    assert list(idf_a) == list(ids)
    for _ in range(2):
        h.predict(0.03, 7.7, 4.0, Q, 0.025)
        sf.predict(0.03, 7.7, 4.0, Q, 0.025)
        h.update(zf_a, R, idf_a)
        sf.update(zf_a, R, idf_a)
    x1, P1 = h.get_state()
    blk, _ = sf.cov_block(0, n, 0, n)
    ex, eP = rel_err(sf.get_x(), x1), rel_err(np.triu(blk), np.triu(P1))
    print(f"\n[shard world=1 n={n} m={m}] x rel {ex:.2e}  P rel {eP:.2e}; {len(shard_plan(n, 1, 0))} panels")
    assert ex < 1e-12 and eP < 1e-12
    sf.close()
    h.close()


@pytest.mark.parametrize("nl,knew", [(300, 4), (700, 40)])
def test_sharded_add_features_matches_single_filter(nl, knew):
    """associate -> update -> add_features -> predict -> update on both paths; the new rows and
    columns (and everything the later update makes of them) must agree.  knew = 40 takes two
    augment launches (32 landmarks per launch) and crosses the first block boundary at nl = 700."""
    from slamcuda._ffi import lib, check, dptr, iptr
    from slamcuda.shard import ShardedFilter, shard_plan
    from slamcuda.slam import _Handle
    n0 = 3 + 2 * nl
    ncap = n0 + 2 * knew
    side = math.sqrt(nl / 0.0035)
    h = _Handle(0, ncap)
    check(lib.slam_synth_state(h.ptr, nl, side, 11))
    sf = ShardedFilter(0, 0, 1, ncap)
    sf.synth_state(nl, side, 11)
    m = 6
    z = np.zeros(2 * (m + knew))
    ids = np.zeros(m, dtype=np.int64)
    check(lib.slam_synth_observations(h.ptr, nl, side, 11, m, knew, dptr(z), iptr(ids)))
    z2 = z.reshape(m + knew, 2).T
    zf_a, idf_a, zn_a = h.associate(z2, R, 4.0, 25.0)
    zf_b, idf_b, zn_b = sf.associate(z2, R, 4.0, 25.0)
    assert np.array_equal(idf_a, idf_b) and np.array_equal(zn_a, zn_b)
    assert 0 < zn_a.shape[1] <= knew
    # add_features takes whatever the caller hands it: use all knew spurious observations (not only
    # the ones outside the outer gate) so that the count is fixed
    zn_a = zn_b = z2[:, m:].copy()
    k = knew
    h.update(zf_a, R, idf_a)
    sf.update(zf_a, R, idf_a)
    h.add_features(zn_a, R)
    sf.add_features(zn_b, R)
    assert sf.n == h.n == n0 + 2 * k
    xa, Pa = h.get_state()
    nn = sf.n
    blk, cnt = sf.cov_block(0, nn, 0, nn)
    assert cnt == nn * nn
    assert np.array_equal(sf.get_x(), xa)
    assert np.array_equal(np.triu(blk), np.triu(Pa))
    # the new landmarks take part in the next cycle
    h.predict(0.03, 7.7, 4.0, Q, 0.025)
    sf.predict(0.03, 7.7, 4.0, Q, 0.025)
    idn = np.array([nl + 1, nl + k, 1], dtype=np.int64)
    zq = np.stack([h.predict_observation(int(i))[0] for i in idn], axis=1) + 0.01
    h.update(zq, R, idn)
    sf.update(zq, R, idn)
    xb, Pb = h.get_state()
    blk, _ = sf.cov_block(0, nn, 0, nn)
    ex, eP = rel_err(sf.get_x(), xb), rel_err(np.triu(blk), np.triu(Pb))
    print(f"\n[shard add_features n0={n0} k={k}] x rel {ex:.2e}  P rel {eP:.2e}; {len(shard_plan(ncap, 1, 0))} panels")
    assert ex < 1e-12 and eP < 1e-12
    with pytest.raises(Exception):
        sf.add_features(np.tile(zn_b[:, :1], (1, knew + 1)), R)   # more than the capacity left
    sf.close()
    h.close()


@pytest.mark.parametrize("p2p", ["1", "0"])
def test_sharded_two_ranks_match_single_filter(p2p):
    """tools/shard_check.py under torch.distributed.run on two GPUs: predict / associate / update /
    add_features sharded over two ranks vs the single filter on rank 0, bit-equal -- once with the
    peer-memory exchange (fused compute + store into every rank's buffers over NVLink) and once with
    the NCCL collectives.  Skipped on a one-GPU box (the driver's `-m gpu` run); run by `gpurun --gpus 2`."""
    import os
    import subprocess
    import sys
    from pathlib import Path
    import slamcuda
    if slamcuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    root = Path(__file__).resolve().parent.parent
    env = dict(os.environ, SLAM_SHARD_P2P=p2p)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
           "127.0.0.1", "--master-port", "29533" if p2p == "1" else "29534", str(root / "tools" / "shard_check.py"),
           "3000", "8"]
    out = subprocess.run(cmd, cwd=root, env=env, capture_output=True, text=True, timeout=300)
    tail = out.stdout[-1500:] + out.stderr[-1500:]
    assert out.returncode == 0, tail
    assert "SHARD_CHECK PASS" in out.stdout, tail
    assert f"p2p={'True' if p2p == '1' else 'False'}" in out.stdout, tail
    # the padding guard of each rank's panel stack (slam_shard_debug_guard) ran and counted nothing
    assert out.stdout.count("padding_slots_overwritten=0") == 2, tail


@pytest.mark.parametrize("nl,m,spare", [(300, 5, 7), (700, 8, 40), (700, 20, 1), (1500, 36, 130)])
def test_sharded_padding_guard_no_stray_writes(nl, m, spare):
    """Own memory check of the panel-sharded path (compute-sanitizer is closed on this pool), world = 1: the whole
    sharded code path (256-row panels, rank-local buffers, panel-mapped K6, sharded gating, augment) on one device.
    slam_shard_debug_guard paints every slot of the ldl x n_max panel stack that holds no element of the live state,
    and x[n, npad), with a NaN payload; predict, associate, update and add_features must leave every painted slot as it
    was (add_features moves n and owns what it adds), and x / P must stay equal to the single filter's and finite."""
    from slamcuda._ffi import lib, check, dptr, iptr
    from slamcuda.shard import ShardedFilter
    from slamcuda.slam import _Handle
    n = 3 + 2 * nl
    side = math.sqrt(nl / 0.0035)
    h = _Handle(0, n + spare, grow=False)
    check(lib.slam_synth_state(h.ptr, nl, side, 9))
    sf = ShardedFilter(0, 0, 1, n + spare)
    sf.synth_state(nl, side, 9)
    sf.guard_arm()
    assert sf.guard_changed() == 0
    nz = m + 3
    z = np.zeros(2 * nz)
    ids = np.zeros(m, dtype=np.int64)
    check(lib.slam_synth_observations(h.ptr, nl, side, 9, m, 3, dptr(z), iptr(ids)))
    z2 = z.reshape(nz, 2).T
    zf, idf, zn = sf.associate(z2, R, 4.0, 25.0)
    # (a synthetic "new" observation may fall inside an existing landmark's gate at this density: the single filter
    # on the same state is the judge of the association, not the generator's intent)
    zf_a, idf_a, zn_a = h.associate(z2, R, 4.0, 25.0)
    assert np.array_equal(idf, idf_a) and np.array_equal(zf, zf_a) and np.array_equal(zn, zn_a)
    assert set(ids) <= set(idf) and len(idf) >= m
    assert sf.guard_changed() == 0, "sharded associate wrote outside the live state"
    for _ in range(2):
        sf.predict(0.03, 7.7, 4.0, Q, 0.025)
        h.predict(0.03, 7.7, 4.0, Q, 0.025)
        sf.update(zf, R, idf)
        h.update(zf, R, idf)
    changed = sf.guard_changed()
    assert changed == 0, f"sharded predict / update (rank {2 * len(idf)}) overwrote {changed} padding slots"
    knew = min(spare // 2, zn.shape[1])
    if knew:
        sf.add_features(zn[:, :knew], R)
        h.add_features(zn[:, :knew], R)
        assert sf.n == n + 2 * knew
        assert sf.guard_changed() == 0, "sharded add_features wrote beyond the rows and columns it adds"
        sf.update(zf, R, idf)
        h.update(zf, R, idf)
        assert sf.guard_changed() == 0
    x1, P1 = h.get_state()
    blk, _ = sf.cov_block(0, sf.n, 0, sf.n)
    # only the upper triangle is authoritative (include/slamcuda.h): the rows add_features appended keep the NaN paint
    # below the diagonal, so a kernel that read below the diagonal in the update above would show up here
    assert np.isfinite(np.triu(blk)).all() and np.isfinite(sf.get_x()).all()
    assert rel_err(sf.get_x(), x1) < 1e-12 and rel_err(np.triu(blk), np.triu(P1)) < 1e-12
    sf.close()
    h.close()


def test_sharded_padding_guard_detects_a_write():
    """Positive control for slam_shard_debug_guard: after arming at n0, a smaller synthetic state (n1) leaves the band
    that used to be live outside the new live state.  Its upper triangle was written by the first synth_state (never
    painted), so at least those slots + x[n1, n0) must be counted, and at most the whole band; re-arming clears it."""
    from slamcuda.shard import ShardedFilter
    nl0, nl1 = 400, 390
    n0, n1 = 3 + 2 * nl0, 3 + 2 * nl1
    sf = ShardedFilter(0, 0, 1, n0 + 11)
    sf.synth_state(nl0, math.sqrt(nl0 / 0.0035), 4)
    sf.guard_arm()
    assert sf.guard_changed() == 0
    sf.synth_state(nl1, math.sqrt(nl1 / 0.0035), 4)
    assert sf.n == n1
    lo = (n0 * (n0 + 1) - n1 * (n1 + 1)) // 2 + (n0 - n1)
    hi = n0 * n0 - n1 * n1 + (n0 - n1)
    changed = sf.guard_changed()
    assert lo <= changed <= hi, (lo, changed, hi)
    sf.guard_arm()
    assert sf.guard_changed() == 0
    sf.close()
