"""Arbitrary-precision restatement of the reference's EKF-SLAM filter functions (TEST INFRASTRUCTURE:
only tests/ imports this file; nothing shipped or timed does).

Why it exists: the reference (Julia <= 0.6) cannot run in this image and ships no golden vectors, so
"parity" is pinned to the reference's TEXT.  The float64 oracle (oracle/slam_ref.py), its C twin
(oracle/slam_twin.c) and the CUDA path could all share a transcription error and agree with each other;
this file restates the same seven functions a third time, straight from the Julia source, in 50-digit
mpmath arithmetic with none of the float64 code's structure (no sparse shortcuts, no closed forms, library
Cholesky / inverse / determinant on mp matrices).  At 50 digits rounding is invisible, so any float64
implementation of the same formulas must sit within a few ulps times the condition number of it:
tests/test_oracle.py requires the numpy oracle and the C twin, tests/test_gpu_single.py the CUDA path, to be
within 1e-12 of these values on the KATs at n = 3, 5, 23.

Each function cites the reference file:line it follows (paths under /root/reference).  Indices are 1-based
where the reference's are (idf), matrices are mp.matrix, inputs are exact conversions of float64 arrays.
"""
from __future__ import annotations

import mpmath as mp
import numpy as np

mp.mp.dps = 50


def M(a) -> mp.matrix:
    """float64 array (1-D -> column vector, 2-D) -> exact mp.matrix."""
    a = np.asarray(a, dtype=np.float64)
    if a.ndim == 1:
        a = a.reshape(-1, 1)
    out = mp.matrix(a.shape[0], a.shape[1])
    for i in range(a.shape[0]):
        for j in range(a.shape[1]):
            out[i, j] = mp.mpf(float(a[i, j]))
    return out


def to_np(m: mp.matrix) -> np.ndarray:
    """mp.matrix -> float64 (correctly rounded); a column vector comes back 1-D."""
    a = np.array([[float(m[i, j]) for j in range(m.cols)] for i in range(m.rows)], dtype=np.float64).reshape(m.rows, m.cols)
    return a[:, 0] if m.cols == 1 else a


def mpi_to_pi(phi):
    """src/common.jl:102-110 -- a single wrap."""
    if phi > mp.pi:
        return phi - 2 * mp.pi
    if phi < -mp.pi:
        return phi + 2 * mp.pi
    return phi


def predict_observation(x: mp.matrix, idf: int):
    """src/common.jl:139-165.  x: n x 1.  Returns z (2 x 1) and the dense H (2 x n)."""
    n = x.rows
    fpos = 3 + idf * 2 - 1          # 1-based position of the landmark's x coordinate   (:143)
    dx = x[fpos - 1] - x[0]         # (:146)
    dy = x[fpos] - x[1]
    d2 = dx ** 2 + dy ** 2
    d = mp.sqrt(d2)
    z = mp.matrix([d, mp.atan2(dy, dx) - x[2]])       # (:152) bearing not wrapped
    xd, yd, xd2, yd2 = dx / d, dy / d, dx / d2, dy / d2
    H = mp.zeros(2, n)                                  # (:160)
    H[0, 0], H[0, 1], H[0, 2] = -xd, -yd, 0             # (:161)
    H[1, 0], H[1, 1], H[1, 2] = yd2, -xd2, -1
    H[0, fpos - 1], H[0, fpos] = xd, yd                 # (:162)
    H[1, fpos - 1], H[1, fpos] = -yd2, xd2
    return z, H


def compute_association(x, P, z, R, idf: int):
    """src/data-association.jl:53-63.  z: 2 x 1."""
    zp, H = predict_observation(x, idf)
    v = z - zp
    v[1] = mpi_to_pi(v[1])
    S = H * P * H.T + R
    nis = (v.T * (mp.inverse(S) * v))[0]
    nd = nis + mp.log(mp.det(S))
    return nis, nd


def associate(x, P, z, R, gate1, gate2):
    """src/data-association.jl:1-51.  z: 2 x nz.  Returns (columns of z matched, idf list, columns new)."""
    Nf = (x.rows - 3) // 2
    zf, idf, zn = [], [], []
    for i in range(z.cols):
        jbest, nbest, outer = 0, mp.inf, mp.inf
        zi = mp.matrix([z[0, i], z[1, i]])
        for j in range(1, Nf + 1):
            nis, nd = compute_association(x, P, zi, R, j)
            ingate = 0
            if nis < gate1:
                if nd < nbest:
                    ingate = 1
            if ingate == 1:
                nbest, jbest = nd, j
            elif nis < outer:
                outer = nis
        if jbest != 0:
            zf.append(i)
            idf.append(jbest)
        elif outer > gate2:
            zn.append(i)
    return zf, idf, zn


def predict(x, P, g, v, w, Q, dt):
    """src/ekf.jl:8-43.  Returns new (x, P) (the reference mutates in place)."""
    x, P = x.copy(), P.copy()
    n = x.rows
    g, v, w, dt = mp.mpf(g), mp.mpf(v), mp.mpf(w), mp.mpf(dt)
    phi = x[2]
    s, c = mp.sin(g + phi), mp.cos(g + phi)
    vts, vtc = v * dt * s, v * dt * c
    Gv = mp.matrix([[1, 0, -vts], [0, 1, vtc], [0, 0, 1]])
    Gu = mp.matrix([[dt * c, -vts], [dt * s, vtc], [dt * mp.sin(g) / w, v * dt * mp.cos(g) / w]])
    Pvv = P[0:3, 0:3]
    P[0:3, 0:3] = Gv * Pvv * Gv.T + Gu * Q * Gu.T          # (:32)
    if n > 3:
        Pvm = Gv * P[0:3, 3:n]                               # (:34)
        P[0:3, 3:n] = Pvm
        P[3:n, 0:3] = Pvm.T                                  # (:35)
    x0, x1 = x[0] + vtc, x[1] + vts
    x[0], x[1], x[2] = x0, x1, mpi_to_pi(phi + v * dt * mp.sin(g) / w)     # (:39-41)
    return x, P


def update(x, P, z, R, idf):
    """src/ekf.jl:46-77.  z: 2 x m, idf: list of 1-based ids."""
    lenz, lenx = z.cols, x.rows
    if lenz == 0:
        return x.copy(), P.copy()
    H = mp.zeros(2 * lenz, lenx)
    v = mp.zeros(2 * lenz, 1)
    RR = mp.zeros(2 * lenz, 2 * lenz)
    for i in range(lenz):
        zp, Hi = predict_observation(x, int(idf[i]))
        H[2 * i:2 * i + 2, :] = Hi
        v[2 * i] = z[0, i] - zp[0]
        v[2 * i + 1] = mpi_to_pi(z[1, i] - zp[1])
        RR[2 * i:2 * i + 2, 2 * i:2 * i + 2] = R
    PHt = P * H.T
    S = H * PHt + RR
    S = (S + S.T) * mp.mpf("0.5")
    U = mp.cholesky(S).T            # chol(S) is the upper factor, S = U'U   (:70)
    Cm = mp.inverse(U)
    W1 = PHt * Cm
    W = W1 * Cm.T
    return x + W * v, P - W1 * W1.T


def add_features(x, P, z, R):
    """src/ekf.jl:84-122.  z: 2 x k."""
    x, P = x.copy(), P.copy()
    phi = x[2]                                               # read once (:88)
    for i in range(z.cols):
        ln = x.rows
        r, b = z[0, i], z[1, i]
        s, c = mp.sin(phi + b), mp.cos(phi + b)
        xn = mp.zeros(ln + 2, 1)
        xn[0:ln, 0] = x
        xn[ln], xn[ln + 1] = x[0] + r * c, x[1] + r * s      # (:99)
        Gv = mp.matrix([[1, 0, -r * s], [0, 1, r * c]])
        Gz = mp.matrix([[c, -r * s], [s, r * c]])
        Pn = mp.zeros(ln + 2, ln + 2)                        # hcat / vcat with zeros (:108-109)
        Pn[0:ln, 0:ln] = P
        Pvv = P[0:3, 0:3]
        Pn[ln:ln + 2, ln:ln + 2] = Gv * Pvv * Gv.T + Gz * R * Gz.T     # (:112)
        GP = Gv * Pvv
        Pn[ln:ln + 2, 0:3] = GP                              # (:113)
        Pn[0:3, ln:ln + 2] = GP.T
        if ln > 3:
            GM = Gv * P[0:3, 3:ln]                           # (:117)
            Pn[ln:ln + 2, 3:ln] = GM
            Pn[3:ln, ln:ln + 2] = GM.T
        x, P = xn, Pn
    return x, P
