"""GPU probe of the small kernels in front of the downdate (K3 + K4 fused, K5): where do the microseconds of an update
go when K6 is not the cost?  Two measurements through the C ABI only:
  * a tiny map (40 landmarks, n = 83): K3 / K5 / K6 work is nothing, so the time of back-to-back resident updates is
    the latency chain of the three launches; its growth with m is the rank-dependent part (the innovation Cholesky,
    its inverse, u) of the single CTA that runs K4;
  * config 3's map (10000 landmarks, n = 20003): whole update minus the K6 time from the library's own events = what
    the small kernels leave exposed in front of the downdate.
usage: python tools/k4_probe.py [steps]"""
import ctypes as C
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "slam.jl_b200"))
import torch  # noqa: E402
from slamcuda._ffi import lib, check, dptr, iptr  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 400
R = np.array([0.01, 0, 0, (np.pi / 180) ** 2])


def probe(nl, ms_list, steps, with_k6):
    n = 3 + 2 * nl
    side = (nl / 0.0035) ** 0.5
    h = C.c_void_p()
    check(lib.slam_create(0, n, C.byref(h)))
    stream = torch.cuda.ExternalStream(lib.slam_stream(h))
    rows = []
    for m in ms_list:
        check(lib.slam_synth_state(h, nl, side, 2))
        z = np.zeros(2 * m)
        ids = np.zeros(m, dtype=np.int64)
        check(lib.slam_synth_observations(h, nl, side, 2, m, 0, dptr(z), iptr(ids)))
        d_zf = torch.tensor(z, device="cuda")
        d_idf = torch.tensor(ids, device="cuda")
        torch.cuda.synchronize()
        for _ in range(5):
            check(lib.slam_update_resident(h, d_zf.data_ptr(), m, dptr(R), d_idf.data_ptr()))
        check(lib.slam_sync(h))
        best = None
        for rep in range(3):
            if with_k6:
                check(lib.slam_profile(h, 1))
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(steps):
                check(lib.slam_update_resident(h, d_zf.data_ptr(), m, dptr(R), d_idf.data_ptr()))
            e1.record(stream)
            check(lib.slam_sync(h))
            us = e0.elapsed_time(e1) / steps * 1e3
            k6us = 0.0
            if with_k6:
                k6, cnt = C.c_double(), C.c_int()
                check(lib.slam_profile_read(h, C.byref(k6), C.byref(cnt)))
                check(lib.slam_profile(h, 0))
                k6us = k6.value / max(cnt.value, 1) * 1e3
            if best is None or us < best[0]:
                best = (us, k6us)
        rows.append((m, best[0], best[1]))
    check(lib.slam_destroy(h))
    return rows


print(f"tiny map (n = 83), {steps} back-to-back resident updates, best of 3: us per update")
base = None
for m, us, _ in probe(40, [1, 2, 4, 6, 8, 10, 12, 16], steps, False):
    base = us if base is None else base
    print(f"  m = {m:2d} (R = {2 * m:2d}): {us:7.2f} us   (+{us - base:6.2f} over m = 1)")
print("config 3 map (n = 20003), 50 updates: whole step, K6 (library events), exposed small-kernel time")
for m, us, k6 in probe(10000, [1, 4, 8, 12, 16], 50, True):
    print(f"  m = {m:2d}: step {us:7.1f} us  K6 {k6:7.1f} us  exposed {us - k6:6.1f} us")
print("config 3 map (n = 20003), 200 updates WITHOUT the library's K6 events (they sit between K5 and K6 and behind K6, so the"
      " loop above runs without programmatic-launch overlap; this one is what a caller gets): us per update")
for m, us, _ in probe(10000, [1, 4, 8, 12, 16], 200, False):
    print(f"  m = {m:2d}: step {us:7.1f} us")
print("K4_PROBE done")
