"""GPU probe of the streaming DMMA downdate (downdate_stream.cu): K6 time for several m and tile configurations in one
process (the library reads SLAM_K6S_CFG = "ib,warps,bj,pd" at every launch), next to the FMA / tiled DMMA kernels.
usage: python tools/probe_k6s.py <landmarks> <m list> <steps> [cfg;cfg;...]"""
import ctypes as C
import os
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "slam.jl_b200"))
import torch  # noqa: E402
from slamcuda._ffi import lib, check, dptr, iptr  # noqa: E402

nl = int(sys.argv[1]) if len(sys.argv) > 1 else 10000
ms_list = [int(a) for a in sys.argv[2].split(",")] if len(sys.argv) > 2 else [6, 8, 12, 16, 24, 32]
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 20
cfgs = sys.argv[4].split(";") if len(sys.argv) > 4 else [""]     # "" = the library's tuned default
n = 3 + 2 * nl
side = (nl / 0.0035) ** 0.5
R = np.array([0.01, 0, 0, (np.pi / 180) ** 2])
h = C.c_void_p()
check(lib.slam_create(0, n, C.byref(h)))
stream = torch.cuda.ExternalStream(lib.slam_stream(h))
gb = 8.0 * n * (n + 1) / 1e9


def run(m, variant, cfg=None):
    if cfg:
        os.environ["SLAM_K6S_CFG"] = cfg
    else:
        os.environ.pop("SLAM_K6S_CFG", None)
    check(lib.slam_set_downdate_variant(h, variant))
    z = np.zeros(2 * m)
    ids = np.zeros(m, dtype=np.int64)
    check(lib.slam_synth_observations(h, nl, side, 2, m, 0, dptr(z), iptr(ids)))
    d_zf = torch.tensor(z, device="cuda")
    d_idf = torch.tensor(ids, device="cuda")
    torch.cuda.synchronize()
    for _ in range(3):
        rc = lib.slam_update_resident(h, d_zf.data_ptr(), m, dptr(R), d_idf.data_ptr())
        if rc:
            return None
    check(lib.slam_sync(h))
    check(lib.slam_profile(h, 1))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(steps):
        check(lib.slam_update_resident(h, d_zf.data_ptr(), m, dptr(R), d_idf.data_ptr()))
    e1.record(stream)
    check(lib.slam_sync(h))
    ms = e0.elapsed_time(e1) / steps
    k6, cnt = C.c_double(), C.c_int()
    check(lib.slam_profile_read(h, C.byref(k6), C.byref(cnt)))
    check(lib.slam_profile(h, 0))
    return ms, k6.value / max(cnt.value, 1)


check(lib.slam_synth_state(h, nl, side, 2))
for m in ms_list:
    rank = 2 * m
    rows = []
    if rank <= 32 or rank > 64:
        rows.append(("FMA", 1, None))
    if 24 < rank <= 32 or rank > 64:
        rows.append(("tiled DMMA", 2, None))
    if 3 <= rank <= 64:
        rows += [(f"stream {c}", 3, c) for c in cfgs]
    for name, variant, cfg in rows:
        r = run(m, variant, cfg)
        if r is None:
            print(f"n={n} m={m:3d} R={rank:3d} {name:22s}: not available ({lib.slam_last_error().decode()})", flush=True)
            continue
        ms, k6 = r
        print(f"n={n} m={m:3d} R={rank:3d} {name:22s}: step {ms:7.3f} ms  K6 {k6:7.3f} ms  {gb / (k6 * 1e-3):6.0f} GB/s  "
              f"{2.0 * m * n * (n + 1) / (k6 * 1e-3) / 1e12:5.1f} TFLOP/s", flush=True)
lib.slam_destroy(h)
