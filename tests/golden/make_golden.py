"""Generates tests/golden/kat.npz and tests/golden/course1_runs.npz from the numpy literal
restatement (oracle/slam_ref.py).  The reference (Julia <= 0.6) cannot run here and ships no
golden vectors, so these vectors pin the oracle against itself over time and
against its independent C twin -- "parity unpinned" with respect to the reference runtime.

Run from the repo root:  python tests/golden/make_golden.py
"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent.parent
sys.path.insert(0, str(ROOT))
from oracle import slam_ref as ref  # noqa: E402

OUT = Path(__file__).resolve().parent


def state(rng, nf):
    n = 3 + 2 * nf
    x = np.concatenate([[50.0, 50.0, 0.3], rng.uniform(20, 80, size=2 * nf)])
    L = rng.normal(0, 0.05, size=(n, 8))
    P = L @ L.T + np.diag(rng.uniform(0.01, 0.1, size=n))
    return x, P


def main():
    k = ref.sim_constants()
    Q, R, dt = k["Q"], k["R"], k["dt"]
    kat = {}
    for nf in (0, 1, 10, 35):          # n = 3, 5, 23, 73
        rng = np.random.default_rng(100 + nf)
        x, P = state(rng, nf)
        tag = f"n{3 + 2 * nf}"
        kat[f"{tag}_x"], kat[f"{tag}_P"] = x, P
        xp, Pp = ref.predict(x.copy(), P.copy(), 0.07, 7.6, 4.0, Q, dt)
        kat[f"{tag}_predict_x"], kat[f"{tag}_predict_P"] = xp, Pp
        # observations: some close to landmarks, one in the drop zone, two far (new), one across +-pi
        zs = []
        for idf in range(1, min(nf, 4) + 1):
            zp, _ = ref.predict_observation(x, idf)
            zs.append(zp + rng.normal(0, [0.05, 0.005]))
        zs.append(np.array([150.0, 0.5]))
        zs.append(np.array([170.0, -2.0]))
        z = np.array(zs).T
        zf, idf, zn = ref.associate(x, P, z, R, k["gate1"], k["gate2"])
        kat[f"{tag}_z"], kat[f"{tag}_zf"], kat[f"{tag}_idf"], kat[f"{tag}_zn"] = z, zf, idf, zn
        xu, Pu = ref.update(x, P, zf, R, idf)
        kat[f"{tag}_update_x"], kat[f"{tag}_update_P"] = xu, Pu
        xa, Pa = ref.add_features(xu, Pu, zn, R)
        kat[f"{tag}_add_x"], kat[f"{tag}_add_P"] = xa, Pa
        if nf:
            nis_nd = np.array([ref.compute_association(x, P, z[:, 0], R, j) for j in range(1, nf + 1)])
            kat[f"{tag}_nis_nd"] = nis_nd
            zp, H = ref.predict_observation(x, nf)
            kat[f"{tag}_obs_z"], kat[f"{tag}_obs_H"] = zp, H
    np.savez_compressed(OUT / "kat.npz", **kat)

    wp = ref.load_waypoints(str(OUT / "course1.txt"))
    runs = {}
    for nl in (10, 35):
        for seed in range(8):
            lm, ctrl, obs = ref.make_inputs(nl, 0, seed)
            rec = []
            r = ref.sim(wp, lm, ctrl, obs, record=rec)
            tag = f"nl{nl}_s{seed}"
            runs[f"{tag}_x"] = r["x"]
            runs[f"{tag}_Pdiag"] = np.diag(r["P"]).copy()
            runs[f"{tag}_track_every50"] = r["slam_track"][:, ::50].copy()
            runs[f"{tag}_counts"] = np.array([r["nsteps"], r["nupdates"], len(r["x"]), r["obs_used"]])
            runs[f"{tag}_idf_concat"] = np.concatenate([e["idf"] for e in rec if e["op"] == "observe"] or [np.zeros(0)])
            runs[f"{tag}_min_margin"] = np.array(min(e["margin"] for e in rec if e["op"] == "observe"))
    np.savez_compressed(OUT / "course1_runs.npz", **runs)
    for f in ("kat.npz", "course1_runs.npz"):
        print(f, (OUT / f).stat().st_size, "bytes")


if __name__ == "__main__":
    main()
