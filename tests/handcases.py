"""Hand-derived known-answer cases for the filter functions, independent of every implementation in this repo:
the expected values are closed-form expressions worked out on paper from the reference's formulas
(src/ekf.jl:8-122, src/data-association.jl:1-63, src/common.jl:102-165) for states simple enough to do so
(vehicle at the origin with heading 0, landmarks on the x axis, diagonal P, so H and every Jacobian have
entries 0, +-1, +-1/d and S is diagonal).  They exercise the matched / dropped / new branches of associate and
the strict-< tie (lowest landmark id wins).  tests/test_oracle.py runs the numpy oracle, the C twin and the
mpmath restatement through them, tests/test_gpu_single.py the CUDA path."""
import math

import numpy as np

R = np.array([[0.01, 0.0], [0.0, 0.0004]])          # sigma_r = 0.1, sigma_b = 0.02
Q = np.array([[0.25, 0.0], [0.0, 0.0025]])


def state_two_landmarks():
    """vehicle (0, 0, 0); landmark 1 at (10, 0), landmark 2 at (20, 0); P diagonal."""
    x = np.array([0.0, 0.0, 0.0, 10.0, 0.0, 20.0, 0.0])
    p = np.array([0.04, 0.09, 0.0001, 0.25, 0.16, 0.36, 0.49])
    return x, np.diag(p), p


def case_predict():
    """g = 0, phi = 0: s = 0, c = 1, vts = 0, vtc = v dt.  Gv = [1 0 0; 0 1 vtc; 0 0 1], Gu = [dt 0; 0 vtc; 0 v dt / w]."""
    x, P, p = state_two_landmarks()
    v, dt, w = 8.0, 0.025, 4.0
    vtc = v * dt
    Pe = P.copy()
    Pe[0, 0] = p[0] + dt * dt * Q[0, 0]
    Pe[1, 1] = p[1] + vtc * vtc * p[2] + vtc * vtc * Q[1, 1]
    Pe[1, 2] = Pe[2, 1] = vtc * p[2] + vtc * (v * dt / w) * Q[1, 1]
    Pe[2, 2] = p[2] + (v * dt / w) ** 2 * Q[1, 1]
    xe = x.copy()
    xe[0] = vtc                        # x + vtc, y + vts = 0, phi + v dt sin(0) / w = 0
    return dict(x=x, P=P, g=0.0, v=v, w=w, dt=dt, Q=Q, x_out=xe, P_out=Pe)


def case_observe():
    """predict_observation / compute_association of landmark 1 (d = 10 on the x axis):
    z = [10, 0]; H = [-1 0 0 1 0 | 0 -1/d -1 0 1/d] on columns (1, 2, 3, 4, 5);
    S = diag(p1 + p4 + R11, (p2 + p5) / d^2 + p3 + R22)."""
    x, P, p = state_two_landmarks()
    d = 10.0
    H = np.zeros((2, 7))
    H[0, 0], H[0, 3] = -1.0, 1.0
    H[1, 1], H[1, 2], H[1, 4] = -1.0 / d, -1.0, 1.0 / d
    s11 = p[0] + p[3] + R[0, 0]
    s22 = (p[1] + p[4]) / d ** 2 + p[2] + R[1, 1]
    z = np.array([10.3, 0.01])
    v = z - np.array([10.0, 0.0])
    nis = v[0] ** 2 / s11 + v[1] ** 2 / s22
    return dict(x=x, P=P, idf=1, zp=np.array([10.0, 0.0]), H=H, z=z, R=R, nis=nis, nd=nis + math.log(s11 * s22))


def case_associate():
    """Four observations against the two landmarks (gates 4 / 25):
      A = landmark 1 + (0.3, 0.01)          -> matched to 1 (nis = 0.09/0.3 + 1e-4/0.003 = 0.333)
      B = range 12.0 (2 m beyond landmark 1) -> nis_1 = 4/0.3 = 13.3: fails gate1, below gate2 -> DROPPED
      C = range 60, bearing 1.0              -> nis huge for both -> NEW
      D = landmark 2 + (-0.2, 0)             -> matched to 2
    Observation order is kept in zf / idf / zn."""
    x, P, _p = state_two_landmarks()
    z = np.array([[10.3, 12.0, 60.0, 19.8], [0.01, 0.0, 1.0, 0.0]])
    return dict(x=x, P=P, z=z, R=R, gate1=4.0, gate2=25.0, idf=[1, 2], zf=z[:, [0, 3]], zn=z[:, [2]])


def case_tie():
    """Two landmarks at the SAME position with identical covariance: nis and nd are bit-equal for both, and the
    reference's strict `nd < nbest` (src/data-association.jl:31) keeps the first: id 1."""
    x = np.array([0.0, 0.0, 0.0, 10.0, 0.0, 10.0, 0.0])
    P = np.diag([0.04, 0.09, 0.0001, 0.25, 0.16, 0.25, 0.16])
    z = np.array([[10.1], [0.005]])
    return dict(x=x, P=P, z=z, R=R, gate1=4.0, gate2=25.0, idf=[1], zf=z, zn=np.zeros((2, 0)))


def case_update():
    """One observation of landmark 1 of the two-landmark state.  S is diagonal, so with K = P H' inv(S):
    x+ = x + K v, P+ = P - K S K'.  Only (x1, l1x) move with the range innovation and (y, phi, l1y) with the bearing
    innovation; landmark 2 is uncorrelated and must stay bit-unchanged."""
    c = case_observe()
    x, P, H = c["x"], c["P"], c["H"]
    p = np.diag(P)
    d = 10.0
    s11 = p[0] + p[3] + R[0, 0]
    s22 = (p[1] + p[4]) / d ** 2 + p[2] + R[1, 1]
    v = c["z"] - c["zp"]
    # columns of P H': range row h1 = (-1, 0, 0, 1, 0), bearing row h2 = (0, -1/d, -1, 0, 1/d)
    k1 = np.array([-p[0], 0, 0, p[3], 0, 0, 0]) / s11
    k2 = np.array([0, -p[1] / d, -p[2], 0, p[4] / d, 0, 0]) / s22
    xe = x + k1 * v[0] + k2 * v[1]
    Pe = P - s11 * np.outer(k1, k1) - s22 * np.outer(k2, k2)
    return dict(x=x, P=P, z=c["z"].reshape(2, 1), R=R, idf=[1], x_out=xe, P_out=Pe)


def case_add_features():
    """New landmark from (r, b = 0) with the vehicle at the origin, heading 0: s = 0, c = 1, landmark at (r, 0),
    Gv = [1 0 0; 0 1 r], Gz = [1 0; 0 r].  Second new landmark in the same call sees the first through 4:len."""
    x = np.array([0.0, 0.0, 0.0, 10.0, 0.0])
    p = np.array([0.04, 0.09, 0.0001, 0.25, 0.16])
    P = np.diag(p)
    P[1, 4] = P[4, 1] = 0.02          # a vehicle-y / landmark-y correlation to carry into the new rows
    r1, r2 = 5.0, 8.0
    z = np.array([[r1, r2], [0.0, 0.0]])
    xe = np.concatenate([x, [r1, 0.0, r2, 0.0]])
    Pe = np.zeros((9, 9))
    Pe[:5, :5] = P
    for a, r in ((5, r1), (7, r2)):
        # rows of Gv * [Pvv | Pvm]: row a = P[0, :], row a+1 = P[1, :] + r * P[2, :]
        Pe[a, a] = p[0] + R[0, 0]
        Pe[a + 1, a + 1] = p[1] + r * r * p[2] + r * r * R[1, 1]
        Pe[a, 0] = Pe[0, a] = p[0]
        Pe[a + 1, 1] = Pe[1, a + 1] = p[1]
        Pe[a + 1, 2] = Pe[2, a + 1] = r * p[2]
        Pe[a + 1, 4] = Pe[4, a + 1] = P[1, 4]
    # the second landmark against the first (map-to-feature term Gv * P[1:3, 6:7]):
    # P[0:3, 5] = (p0, 0, 0), P[0:3, 6] = (0, p1, r1 p2)
    Pe[7, 5] = Pe[5, 7] = p[0]
    Pe[8, 6] = Pe[6, 8] = p[1] + r2 * r1 * p[2]
    return dict(x=x, P=P, z=z, R=R, x_out=xe, P_out=Pe)
