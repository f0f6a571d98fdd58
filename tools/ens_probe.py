"""Quick GPU probe of the ensemble kernel: `ens_probe.py <n_runs> [reps] [landmarks] [threads] [extra] [first_pass_extra]` runs n_runs
course1 members in one launch and prints the launch time, filter-steps/s and how many members
stopped early.  Used under ncu for the k_ensemble captures in profiles/."""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "slam.jl_b200"))
import torch  # noqa: E402
from slamcuda._ffi import lib  # noqa: E402
from slamcuda.ensemble import Ensemble  # noqa: E402

n_runs = int(sys.argv[1]) if len(sys.argv) > 1 else 444
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
nl = int(sys.argv[3]) if len(sys.argv) > 3 else 35
threads = int(sys.argv[4]) if len(sys.argv) > 4 and sys.argv[4] != "-" else None
extra = int(sys.argv[5]) if len(sys.argv) > 5 and sys.argv[5] != "-" else None
first = int(sys.argv[6]) if len(sys.argv) > 6 and sys.argv[6] != "-" else None
growth = int(sys.argv[7]) if len(sys.argv) > 7 and sys.argv[7] != "-" else None
max_shared = int(sys.argv[8]) if len(sys.argv) > 8 and sys.argv[8] != "-" else None
wp = np.loadtxt(ROOT / "tests" / "golden" / "course1.txt", skiprows=1).T.copy()
lm = np.random.Generator(np.random.PCG64(0)).integers(5, 96, size=(2, nl)).astype(float)
ens = Ensemble(wp, lm, n_runs, seed=2, run0=0, threads=threads, extra_landmarks=extra, first_pass_extra=first,
               growth=growth, max_shared_extra=max_shared)
stream = torch.cuda.ExternalStream(lib.slam_ens_stream(ens.ptr))
ens.run()
bad = ens.sync()
ts = []
for _ in range(reps):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    ens.run()
    ens.sync()
    e1.record(stream)
    e1.synchronize()
    ts.append(e0.elapsed_time(e1))
ms = sorted(ts)[len(ts) // 2]
pose, nf, st = ens.results()
total, per = ens.steps_done()
print(f"runs {n_runs} nl {nl} steps {ens.nsteps} threads {threads} cap {ens.capacity}: {ms:.3f} ms/launch, "
      f"{total / ms / 1e3:.2f} M filter-steps/s ({total} steps done of {n_runs * ens.nsteps}), "
      f"growth {growth}, launches (capacity, members) {ens.tier_info()}, stopped early {bad}, n_final max {nf.max()}, status counts {dict(zip(*np.unique(st, return_counts=True)))}")
ens.close()
