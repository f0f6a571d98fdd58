"""Phase split of the CTA that runs K4 inside k_pht_chol, for a library built with SLAMCUDA_DEFS=-DK4_TIMING (the kernel's
thread 0 prints clock64 deltas): model | wait for the previous downdate | gather of the sub-block of P | S + symmetrise |
Cholesky | inverse | u and C written.  Back-to-back resident updates, so the later lines of each m are the steady state.
usage: SLAMCUDA_DEFS=-DK4_TIMING python slam.jl_b200/build.py (after touching csrc/filter.cu); python tools/k4_phases.py [landmarks]"""
import ctypes as C
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "slam.jl_b200"))
import torch  # noqa: E402
from slamcuda._ffi import lib, check, dptr, iptr  # noqa: E402

nl = int(sys.argv[1]) if len(sys.argv) > 1 else 10000
n = 3 + 2 * nl
side = (nl / 0.0035) ** 0.5
R = np.array([0.01, 0, 0, (np.pi / 180) ** 2])
h = C.c_void_p()
check(lib.slam_create(0, n, C.byref(h)))
for m in (1, 4, 8, 12, 16):
    check(lib.slam_synth_state(h, nl, side, 2))
    z = np.zeros(2 * m)
    ids = np.zeros(m, dtype=np.int64)
    check(lib.slam_synth_observations(h, nl, side, 2, m, 0, dptr(z), iptr(ids)))
    d_zf = torch.tensor(z, device="cuda")
    d_idf = torch.tensor(ids, device="cuda")
    torch.cuda.synchronize()
    print(f"--- m = {m} (R = {2 * m}), n = {n}", flush=True)
    for _ in range(6):
        check(lib.slam_update_resident(h, d_zf.data_ptr(), m, dptr(R), d_idf.data_ptr()))
    check(lib.slam_sync(h))
    torch.cuda.synchronize()
check(lib.slam_destroy(h))
print("K4_PHASES done")
