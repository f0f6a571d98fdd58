"""Small end-to-end exercise of every kernel family for compute-sanitizer (memcheck / racecheck): single filter
(predict, associate with and without the pre-gate, update at ranks that take the FMA, streaming DMMA and tiled DMMA
downdate kernels on a ragged n, add_features), the panel-sharded filter with world = 1, and a tiny ensemble with growth.
usage: compute-sanitizer --tool memcheck python tools/sanitize_small.py"""
import math
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "slam.jl_b200"))
from slamcuda._ffi import lib, check, dptr, iptr  # noqa: E402
from slamcuda.ensemble import Ensemble  # noqa: E402
from slamcuda.shard import ShardedFilter  # noqa: E402
from slamcuda.slam import _Handle  # noqa: E402

R = np.array([[0.01, 0.0], [0.0, (math.pi / 180) ** 2]])
Q = np.array([[0.25, 0.0], [0.0, (3 * math.pi / 180) ** 2]])
nl = 333
n = 3 + 2 * nl
side = math.sqrt(nl / 0.0035)
for m in (1, 2, 4, 8, 14, 20, 30, 36):
    h = _Handle(0, n + 8)
    check(lib.slam_synth_state(h.ptr, nl, side, 5))
    z = np.zeros(2 * (m + 3))
    ids = np.zeros(m, dtype=np.int64)
    check(lib.slam_synth_observations(h.ptr, nl, side, 5, m, 3, dptr(z), iptr(ids)))
    z2 = z.reshape(m + 3, 2).T
    h.predict(0.02, 7.9, 4.0, Q, 0.025)
    for pg in (True, False):
        h.set_pregate(pg)
        zf, idf, zn = h.associate(z2, R, 4.0, 25.0)
    h.update(zf, R, idf)
    if zn.shape[1]:
        h.add_features(zn[:, :2], R)
    h.feature_ellipses()
    h.landmark_ellipses()
    h.get_state()
    h.close()
    print("single filter m =", m, "matched", len(idf), flush=True)
sf = ShardedFilter(0, 0, 1, n + 8)
sf.synth_state(nl, side, 5)
for m in (4, 8, 20):
    z = np.zeros(2 * (m + 3))
    ids = np.zeros(m, dtype=np.int64)
    check(lib.slam_synth_observations(None, nl, side, 5, m, 3, dptr(z), iptr(ids)))
    zf, idf, zn = sf.associate(z.reshape(m + 3, 2).T, R, 4.0, 25.0)
    sf.predict(0.02, 7.9, 4.0, Q, 0.025)
    sf.update(zf, R, idf)
sf.add_features(np.array([[30.0, 40.0], [0.3, -0.4]]), R)
sf.cov_block(0, sf.n, 0, sf.n)
sf.close()
print("sharded world = 1 ok", flush=True)
wp = np.loadtxt(ROOT / "tests" / "golden" / "course1.txt", skiprows=1).T.copy()
lm = np.random.Generator(np.random.PCG64(0)).integers(5, 96, size=(2, 35)).astype(float)
ens = Ensemble(wp, lm, 6, seed=2, run0=18, nlaps=1)
ens.run()
print("ensemble stopped:", ens.sync(), ens.tier_info(), flush=True)
ens.close()
print("SANITIZE_SMALL done")
