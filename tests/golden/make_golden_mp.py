"""Generates tests/golden/kat_mp.npz: the known-answer states of kat.npz (n = 3, 5, 23) pushed through the
50-digit mpmath restatement of the reference's text (oracle/slam_mp.py) and rounded to float64.  These are the
values every float64 implementation of the same formulas -- the numpy oracle, the C twin, the CUDA path -- must
sit within 1e-12 of; unlike kat.npz they do not come from the float64 oracle.

Run from the repo root:  python tests/golden/make_golden_mp.py   (about a minute)
"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent.parent
sys.path.insert(0, str(ROOT))
from oracle import slam_mp as smp  # noqa: E402
from oracle import slam_ref as ref  # noqa: E402  (constants only)

OUT = Path(__file__).resolve().parent


def compute(kat, n):
    """All mp results for the KAT state of size n, as float64 arrays."""
    k = ref.sim_constants()
    Q, R, dt = smp.M(k["Q"]), smp.M(k["R"]), k["dt"]
    t = f"n{n}"
    x, P, z = smp.M(kat[f"{t}_x"]), smp.M(kat[f"{t}_P"]), smp.M(kat[f"{t}_z"])
    out = {}
    xp, Pp = smp.predict(x, P, 0.07, 7.6, 4.0, Q, dt)
    out["predict_x"], out["predict_P"] = smp.to_np(xp), smp.to_np(Pp)
    zf_i, idf, zn_i = smp.associate(x, P, z, R, k["gate1"], k["gate2"])
    out["idf"] = np.array(idf, dtype=np.int64)
    out["zf_cols"], out["zn_cols"] = np.array(zf_i, dtype=np.int64), np.array(zn_i, dtype=np.int64)
    def cols(idx):
        m = smp.mp.matrix(2, len(idx))
        for c, i in enumerate(idx):
            m[0, c], m[1, c] = z[0, i], z[1, i]
        return m

    zf, zn = cols(zf_i), cols(zn_i)
    xu, Pu = smp.update(x, P, zf, R, idf)
    out["update_x"], out["update_P"] = smp.to_np(xu), smp.to_np(Pu)
    # add_features starts from the float64 update result of kat.npz (what the float64 implementations are fed)
    xa, Pa = smp.add_features(smp.M(kat[f"{t}_update_x"]), smp.M(kat[f"{t}_update_P"]), zn, R)
    out["add_x"], out["add_P"] = smp.to_np(xa), smp.to_np(Pa)
    nf = (n - 3) // 2
    if nf:
        z0 = smp.mp.matrix([z[0, 0], z[1, 0]])
        out["nis_nd"] = np.array([[float(v) for v in smp.compute_association(x, P, z0, R, j)] for j in range(1, nf + 1)])
        zp, H = smp.predict_observation(x, nf)
        out["obs_z"], out["obs_H"] = smp.to_np(zp), smp.to_np(H)
    return out


def main():
    kat = np.load(OUT / "kat.npz")
    res = {}
    for n in (3, 5, 23):
        for key, val in compute(kat, n).items():
            res[f"n{n}_{key}"] = val
    np.savez_compressed(OUT / "kat_mp.npz", **res)
    print("kat_mp.npz", (OUT / "kat_mp.npz").stat().st_size, "bytes")


if __name__ == "__main__":
    main()
