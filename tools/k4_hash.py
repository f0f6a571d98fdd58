"""Digests of x and P after updates at ranks 2..32 (and the not-positive-definite error path), for comparing two
builds of the library bit for bit on one box: the warp-synchronous k4_factor against the -DK4_LEGACY body.
usage: python tools/k4_hash.py"""
import hashlib
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "slam.jl_b200"))
from slamcuda._ffi import lib  # noqa: E402
from slamcuda.slam import _Handle  # noqa: E402

R = np.diag([0.01, (np.pi / 180) ** 2])
for nf, m in [(60, 1), (60, 2), (60, 4), (130, 6), (200, 8), (200, 10), (200, 12), (300, 13), (300, 16)]:
    rng = np.random.default_rng(100 * nf + m)
    n = 3 + 2 * nf
    x = np.concatenate([[50.0, 50.0, 0.3], rng.uniform(20, 80, size=2 * nf)])
    L = rng.normal(0, 0.05, size=(n, 8))
    P = L @ L.T + np.diag(rng.uniform(0.01, 0.1, size=n))
    ids = rng.choice(nf, size=m, replace=False) + 1
    z = np.zeros((2, m))
    for c, i in enumerate(ids):
        dx, dy = x[3 + 2 * (i - 1)] - x[0], x[4 + 2 * (i - 1)] - x[1]
        z[:, c] = [np.hypot(dx, dy) + rng.normal(0, 0.1), np.arctan2(dy, dx) - x[2] + rng.normal(0, 0.01)]
    h = _Handle(0, n)
    h.set_state(x, P)
    for _ in range(3):
        h.update(z, R, ids)
    xo, Po = h.get_state()
    h.close()
    print(f"nf={nf} m={m} R={2 * m}: {hashlib.sha256(xo.tobytes() + Po.tobytes()).hexdigest()[:32]}")
# S not positive definite: P = -I on the observed landmark
x = np.array([0.0, 0.0, 0.0, 10.0, 5.0])
P = -np.eye(5)
h = _Handle(0, 5)
h.set_state(x, P)
rc = lib.slam_update(h.ptr, np.array([11.0, 0.4]).ctypes.data_as(lib.slam_update.argtypes[1]), 1,
                     np.asfortranarray(R).ctypes.data_as(lib.slam_update.argtypes[3]),
                     np.array([1], dtype=np.int64).ctypes.data_as(lib.slam_update.argtypes[4]))
xo, Po = h.get_state()
print(f"not PD: rc={rc} state untouched={np.array_equal(xo, x) and np.array_equal(Po, P)}")
h.close()
