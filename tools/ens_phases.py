"""Developer probe: per-phase cycle counts of ensemble member 0 (needs a library built with
SLAMCUDA_DEFS=-DENS_TIMING).  usage: ens_phases.py <n_runs> [threads]"""
import ctypes as C
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "slam.jl_b200"))
from slamcuda._ffi import lib  # noqa: E402
from slamcuda.ensemble import Ensemble  # noqa: E402

n_runs = int(sys.argv[1]) if len(sys.argv) > 1 else 592
threads = int(sys.argv[2]) if len(sys.argv) > 2 else None
wp = np.loadtxt(ROOT / "tests" / "golden" / "course1.txt", skiprows=1).T.copy()
lm = np.random.Generator(np.random.PCG64(0)).integers(5, 96, size=(2, 35)).astype(float)
ens = Ensemble(wp, lm, n_runs, seed=2, run0=0, threads=threads)
ens.run(); ens.sync()
out = (C.c_longlong * 24)()
lib.slam_ens_debug_phases(out, 1)
ens.run(); ens.sync()
lib.slam_ens_debug_phases(out, 0)
names = ["gap", "predict(i) noise+heading+sincos", "predict(ii) Pvv chain | columns", "assoc A per-landmark", "assoc B score+argmin",
         "classify", "Hn/v copy", "PHt", "S", "symmetrise", "inverse+u (after chol)", "W1, x", "downdate", "add_features", "track/end",
         "  (i): sincos + GuQGu'", "  (ii): 3x3 chain", "  chol", "  (i): heading chain"]
names[1] = "predict(i) tail: prefix sums"
names[2] = "predict(ii) tail: wait helpers"
tot = sum(out[:19])
nb = ens.nupdates + 1
print(f"members {n_runs}: {tot} cycles for member 0, {tot / nb:.0f} per block")
for k, nm in enumerate(names):
    print(f"  {nm:36s} {out[k] / nb:9.0f} cyc/block  {100.0 * out[k] / tot:5.1f}%")
ens.close()
