"""GPU parity of the row-block-sharded path.  world = 1 exercises the whole sharded code path
(two row blocks, rank-local buffers, unpermute, block-wise K6, sharded gating) on one device
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
    print(f"\n[shard world=1 n={n} m={m}] x rel {ex:.2e}  P rel {eP:.2e}; blocks {shard_plan(n, 1, 0)}")
    assert ex < 1e-12 and eP < 1e-12
    sf.close()
    h.close()
