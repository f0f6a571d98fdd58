"""GPU probe of K2 (k_gate, nearest-neighbour gating): `probe_gate.py <landmarks> <nz> [reps]` times
slam_associate (host call: H2D of z, one k_gate launch, D2H of the compacted result) on a synthetic
map and prints the call time and the algorithmic bytes (88 B per landmark + 16 B per observation).
Run under `ncu -k regex:k_gate` for the kernel alone: the first reps + 3 launches run with the spatial pre-gate,
the next reps + 3 without."""
import ctypes as C
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "slam.jl_b200"))
from slamcuda._ffi import lib, check, dptr, iptr  # noqa: E402

nl = int(sys.argv[1]) if len(sys.argv) > 1 else 35353
nz = int(sys.argv[2]) if len(sys.argv) > 2 else 12
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 50
n = 3 + 2 * nl
side = (nl / 0.0035) ** 0.5
R = np.array([0.01, 0, 0, (np.pi / 180) ** 2])
h = C.c_void_p()
check(lib.slam_create(0, n, C.byref(h)))
check(lib.slam_synth_state(h, nl, side, 2))
m = max(nz - 4, 1)
z = np.zeros(2 * nz)
ids = np.zeros(m, dtype=np.int64)
check(lib.slam_synth_observations(h, nl, side, 2, m, nz - m, dptr(z), iptr(ids)))
zf = np.zeros_like(z); zn = np.zeros_like(z); idf = np.zeros(nz, dtype=np.int64)
mm, kk = C.c_int(), C.c_int()
alg = 88.0 * nl + 72 + 32.0 * nz
res = {}
for on in (1, 0):          # spatial pre-gate on (default), then the exhaustive scan; reps + 3 launches each
    check(lib.slam_set_pregate(h, on))
    a0, b0 = C.c_longlong(), C.c_longlong()
    check(lib.slam_gate_stats(h, C.byref(a0), C.byref(b0), None))
    ts = []
    for _ in range(reps + 3):
        t0 = time.perf_counter()
        check(lib.slam_associate(h, dptr(z), nz, dptr(R), 4.0, 25.0, dptr(zf), iptr(idf), C.byref(mm), dptr(zn),
                                 C.byref(kk)))
        ts.append(time.perf_counter() - t0)
    a1, b1 = C.c_longlong(), C.c_longlong()
    check(lib.slam_gate_stats(h, C.byref(a1), C.byref(b1), None))
    ts = sorted(ts[3:])
    res[on] = (idf[:mm.value].copy(), zf[:2 * mm.value].copy(), zn[:2 * kk.value].copy())
    print(f"associate Nf={nl} nz={nz} pregate={on}: matched {mm.value} (expected {m}: {np.array_equal(idf[:m], ids)}), "
          f"new {kk.value}; scored {(b1.value - b0.value) // (reps + 3)} of {(a1.value - a0.value) // (reps + 3)} landmarks "
          f"per pass; host call median {ts[len(ts) // 2] * 1e6:.1f} us, min {ts[0] * 1e6:.1f} us; "
          f"algorithmic bytes (exhaustive) {alg / 1e6:.2f} MB")
print("identical:", all(np.array_equal(a, b) for a, b in zip(res[1], res[0])))
lib.slam_destroy(h)
