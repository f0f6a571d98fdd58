"""Quick GPU probe: update() at a synthetic n-landmark map for several m; prints step time,
K6 time and achieved GB/s against 8n(n+1) algorithmic bytes."""
import ctypes as C
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "slam.jl_b200"))
import torch  # noqa: E402
from slamcuda._ffi import lib, check, dptr, iptr  # noqa: E402

nl = int(sys.argv[1]) if len(sys.argv) > 1 else 10000
ms_list = [int(a) for a in sys.argv[2].split(",")] if len(sys.argv) > 2 else [1, 4, 8, 16]
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 20
variant = int(sys.argv[4]) if len(sys.argv) > 4 else 0     # 0 auto, 1 FMA kernel, 2 DMMA kernel
n = 3 + 2 * nl
side = (nl / 0.0035) ** 0.5
R = np.array([0.01, 0, 0, (np.pi / 180) ** 2])
h = C.c_void_p()
check(lib.slam_create(0, n, C.byref(h)))
check(lib.slam_set_downdate_variant(h, variant))
stream = torch.cuda.ExternalStream(lib.slam_stream(h))
for m in ms_list:
    check(lib.slam_synth_state(h, nl, side, 2))
    z = np.zeros(2 * (m + 4))
    ids = np.zeros(m, dtype=np.int64)
    check(lib.slam_synth_observations(h, nl, side, 2, m, 4, dptr(z), iptr(ids)))
    zf = np.zeros_like(z); zn = np.zeros_like(z); idf = np.zeros(m + 4, dtype=np.int64)
    mm, kk = C.c_int(), C.c_int()
    t0 = time.perf_counter()
    check(lib.slam_associate(h, dptr(z), m + 4, dptr(R), 4.0, 25.0, dptr(zf), iptr(idf), C.byref(mm), dptr(zn), C.byref(kk)))
    t_assoc = time.perf_counter() - t0
    ok = mm.value == m and np.array_equal(idf[:m], ids)
    d_zf = torch.tensor(zf[:2 * m], device="cuda")
    d_idf = torch.tensor(idf[:m], device="cuda")
    torch.cuda.synchronize()
    for _ in range(3):
        check(lib.slam_update_resident(h, d_zf.data_ptr(), m, dptr(R), d_idf.data_ptr()))
    check(lib.slam_sync(h))
    check(lib.slam_profile(h, 1))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(steps):
        check(lib.slam_update_resident(h, d_zf.data_ptr(), m, dptr(R), d_idf.data_ptr()))
    e1.record(stream)
    check(lib.slam_sync(h))
    ms = e0.elapsed_time(e1) / steps
    k6 = C.c_double(); cnt = C.c_int()
    check(lib.slam_profile_read(h, C.byref(k6), C.byref(cnt)))
    check(lib.slam_profile(h, 0))
    k6ms = k6.value / max(cnt.value, 1)
    gb = 8.0 * n * (n + 1) / 1e9
    tf = 2.0 * m * n * (n + 1) / (k6ms * 1e-3) / 1e12
    print(f"variant={variant} K6 {tf:.1f} TFLOP/s  ", end="")
    print(f"nl={nl} n={n} m={m}: assoc_ok={ok} (m={mm.value},k={kk.value}, {t_assoc*1e3:.2f} ms)  "
          f"step {ms:.3f} ms  K6 {k6ms:.3f} ms  K6 {gb / (k6ms * 1e-3):.0f} GB/s  step {1e3/ms:.1f} upd/s", flush=True)
lib.slam_destroy(h)
