"""ORACLE (test infrastructure, not product code) -- "reference-literal" numpy restatement
of the EKF-SLAM hot path of andrewadare/SLAM.jl.

PARITY UNPINNED: the reference ships no tests, golden vectors or fixtures for this path
 and Julia (<= 0.6) is not available in this image, so this
restatement cannot be checked against the reference's own outputs.  It is pinned only
against its independent twin (oracle/slam_twin.c) and follows the Julia text line by line.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import this module.  The product path (slam.jl_b200/) never does.

Every function cites the reference file:line it restates (paths under /root/reference).
Arrays are column-major in the reference; here they are numpy arrays with the same
mathematical indices (z is 2 x nz, P is n x n); landmark ids `idf` stay 1-based.
Dense products go through numpy's OpenBLAS exactly where Julia's `*` would reach its BLAS.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import numpy as np

PI = math.pi


# --------------------------------------------------------------------------------------
# src/common.jl
# --------------------------------------------------------------------------------------
def mpi_to_pi(phi: float) -> float:
    """src/common.jl:102-110 -- single wrap, not a modulo."""
    if phi > PI:
        return phi - 2 * PI
    if phi < -PI:
        return phi + 2 * PI
    return phi


def predict_observation(x: np.ndarray, idf: int):
    """src/common.jl:139-165 -- expected [range; bearing] and the dense 2 x n Jacobian."""
    pose_length = 3
    fpos = pose_length + idf * 2 - 1            # 1-based position of the landmark's x
    f0 = fpos - 1                               # 0-based
    dx = x[f0] - x[0]
    dy = x[f0 + 1] - x[1]
    d2 = dx * dx + dy * dy
    d = math.sqrt(d2)
    z = np.array([d, math.atan2(dy, dx) - x[2]])
    xd = dx / d
    yd = dy / d
    xd2 = dx / d2
    yd2 = dy / d2
    H = np.zeros((2, len(x)))
    H[:, 0:3] = [[-xd, -yd, 0.0], [yd2, -xd2, -1.0]]
    H[:, f0:f0 + 2] = [[xd, yd], [-yd2, xd2]]
    return z, H


def initial_pose(waypoints: np.ndarray) -> np.ndarray:
    """src/common.jl:93-96."""
    wp = waypoints
    return np.array([wp[0, 0], wp[1, 0],
                     math.atan2(wp[1, 1] - wp[1, 0], wp[0, 1] - wp[0, 0])])


@dataclass
class Vehicle:
    """src/common.jl:36-55 with the values of sim/ekfslam-sim.jl:29-37."""
    wheelbase: float = 4.0
    max_gamma: float = 60 * PI / 180
    steer_rate: float = 60 * PI / 180
    sensor_range: float = 30.0
    pose: np.ndarray = field(default_factory=lambda: np.zeros(3))
    target_speed: float = 8.0
    measured_speed: float = 0.0
    target_gamma: float = 0.0
    measured_gamma: float = 0.0
    waypoint_id: int = 1


def step_vehicle(vehicle: Vehicle, dt: float) -> None:
    """src/common.jl:172-181 -- truth motion uses the TARGET speed and steering angle."""
    x, y, phi = vehicle.pose
    v = vehicle.target_speed
    g = vehicle.target_gamma
    vehicle.pose = np.array([x + v * dt * math.cos(g + phi),
                             y + v * dt * math.sin(g + phi),
                             mpi_to_pi(phi + v * dt * math.sin(g) / vehicle.wheelbase)])


def _sign(a: float) -> float:
    return float(a > 0) - float(a < 0)


def steer(vehicle: Vehicle, waypoints: np.ndarray, d_min: float, dt: float) -> None:
    """src/common.jl:189-230."""
    g = vehicle.target_gamma
    iwp = vehicle.waypoint_id
    x, y, phi = vehicle.pose
    cwp = waypoints[:, iwp - 1]
    d2 = (cwp[0] - x) ** 2 + (cwp[1] - y) ** 2
    if d2 < d_min ** 2:
        iwp += 1
        if iwp > waypoints.shape[1]:
            vehicle.waypoint_id = 0
            return
        cwp = waypoints[:, iwp - 1]
    dg = mpi_to_pi(math.atan2(cwp[1] - y, cwp[0] - x) - phi - g)
    dgmax = vehicle.steer_rate * dt
    if abs(dg) > dgmax:
        dg = _sign(dg) * dgmax
    g += dg
    if abs(g) > vehicle.max_gamma:
        g = _sign(g) * vehicle.max_gamma
    vehicle.target_gamma = g
    vehicle.waypoint_id = iwp


# --------------------------------------------------------------------------------------
# src/ekf.jl
# --------------------------------------------------------------------------------------
def predict(x: np.ndarray, P: np.ndarray, g: float, v: float, w: float,
            Q: np.ndarray, dt: float):
    """src/ekf.jl:8-43.  g, v are the MEASURED steering angle and speed, w the wheelbase.
    Mutates x and P in place (as the reference does) and returns them."""
    phi = x[2]
    s = math.sin(g + phi)
    c = math.cos(g + phi)
    vts = v * dt * s
    vtc = v * dt * c
    Gv = np.array([[1.0, 0.0, -vts],
                   [0.0, 1.0, vtc],
                   [0.0, 0.0, 1.0]])
    Gu = np.array([[dt * c, -vts],
                   [dt * s, vtc],
                   [dt * math.sin(g) / w, v * dt * math.cos(g) / w]])
    P[0:3, 0:3] = Gv @ P[0:3, 0:3] @ Gv.T + Gu @ Q @ Gu.T
    if P.shape[0] > 3:
        P[0:3, 3:] = Gv @ P[0:3, 3:]
        P[3:, 0:3] = P[0:3, 3:].T
    x[0:3] = [x[0] + vtc, x[1] + vts, mpi_to_pi(phi + v * dt * math.sin(g) / w)]
    return x, P


def _chol_upper_inv(S: np.ndarray) -> np.ndarray:
    """inv(chol(S)) of src/ekf.jl:70: S = U'U (dpotrf 'U'), C = U^-1 (dtrtri)."""
    if S.shape[0] == 0:
        return np.zeros((0, 0))
    from scipy.linalg import lapack
    U, info = lapack.dpotrf(S, lower=0, clean=1)
    if info != 0:
        raise np.linalg.LinAlgError("innovation covariance not positive definite")
    C, info = lapack.dtrtri(U, lower=0)
    if info != 0:
        raise np.linalg.LinAlgError("singular Cholesky factor")
    return np.triu(C)


def update(x: np.ndarray, P: np.ndarray, z: np.ndarray, R: np.ndarray, idf):
    """src/ekf.jl:46-77 -- batch Cholesky-form update; returns NEW arrays."""
    z = np.asarray(z, dtype=float).reshape(2, -1)
    idf = np.asarray(idf, dtype=np.int64).reshape(-1)
    lenz = z.shape[1]
    lenx = len(x)
    H = np.zeros((2 * lenz, lenx))
    v = np.zeros(2 * lenz)
    RR = np.zeros((2 * lenz, 2 * lenz))
    for i in range(lenz):
        zp, Hi = predict_observation(x, int(idf[i]))
        H[2 * i:2 * i + 2, :] = Hi
        v[2 * i] = z[0, i] - zp[0]
        v[2 * i + 1] = mpi_to_pi(z[1, i] - zp[1])
        RR[2 * i:2 * i + 2, 2 * i:2 * i + 2] = R
    PHt = P @ H.T
    S = H @ PHt + RR
    S = (S + S.T) * 0.5
    C = _chol_upper_inv(S)
    W1 = PHt @ C
    W = W1 @ C.T
    x_new = x + W @ v
    P_new = P - W1 @ W1.T
    return x_new, P_new


def add_features(x: np.ndarray, P: np.ndarray, z: np.ndarray, R: np.ndarray):
    """src/ekf.jl:84-122 -- augment x and P with new landmarks; returns NEW arrays
    (or the inputs themselves when z has no columns)."""
    z = np.asarray(z, dtype=float).reshape(2, -1)
    phi = x[2]
    for i in range(z.shape[1]):
        ln = len(x)
        r, b = z[0, i], z[1, i]
        s, c = math.sin(phi + b), math.cos(phi + b)
        x = np.concatenate([x, [x[0] + r * c, x[1] + r * s]])
        Gv = np.array([[1.0, 0.0, -r * s], [0.0, 1.0, r * c]])
        Gz = np.array([[c, -r * s], [s, r * c]])
        P = np.hstack([P, np.zeros((P.shape[0], 2))])
        P = np.vstack([P, np.zeros((2, P.shape[1]))])
        rng = slice(ln, ln + 2)
        P[rng, rng] = Gv @ P[0:3, 0:3] @ Gv.T + Gz @ R @ Gz.T
        P[rng, 0:3] = Gv @ P[0:3, 0:3]
        P[0:3, rng] = P[rng, 0:3].T
        if ln > 3:
            rnm = slice(3, ln)
            P[rng, rnm] = Gv @ P[0:3, rnm]
            P[rnm, rng] = P[rng, rnm].T
    return x, P


# --------------------------------------------------------------------------------------
# src/data-association.jl
# --------------------------------------------------------------------------------------
def compute_association(x, P, z, R, idf: int):
    """src/data-association.jl:53-63 -- (nis, nd) with the dense H*P*H' + R."""
    zp, H = predict_observation(x, idf)
    v = np.array([z[0] - zp[0], z[1] - zp[1]])
    v[1] = mpi_to_pi(v[1])
    S = H @ P @ H.T + R
    nis = float(np.dot(v, np.linalg.inv(S) @ v))
    det = float(np.linalg.det(S))
    nd = nis + (math.log(det) if det > 0 else (float("-inf") if det == 0 else float("nan")))
    return nis, nd


def associate(x, P, z, R, gate1: float, gate2: float, margins: list | None = None):
    """src/data-association.jl:1-51 -- gated nearest neighbour.
    Returns (zf 2 x m, idf (m,) int64 1-based, zn 2 x k).  If `margins` is a list, the
    smallest distance of any decisive comparison from its threshold is appended to it
    (SURVEY.md section 7 'bit-exact associations')."""
    z = np.asarray(z, dtype=float).reshape(2, -1)
    zf_cols, zn_cols, idf = [], [], []
    Nf = int(round((len(x) - 3) / 2))
    for i in range(z.shape[1]):
        jbest = 0
        nbest = math.inf
        outer = math.inf
        nds, niss = [], []
        for j in range(1, Nf + 1):
            nis, nd = compute_association(x, P, z[:, i], R, j)
            niss.append(nis)
            nds.append(nd)
            ingate = 0
            if nis < gate1:
                if nd < nbest:
                    ingate = 1
            if ingate == 1:
                nbest = nd
                jbest = j
            elif nis < outer:
                outer = nis
        if jbest != 0:
            zf_cols.append(z[:, i])
            idf.append(jbest)
        elif outer > gate2:
            zn_cols.append(z[:, i])
        if margins is not None and Nf > 0:
            niss_a, nds_a = np.array(niss), np.array(nds)
            mg = float(np.min(np.abs(niss_a - gate1)))
            if jbest != 0:
                ing = niss_a < gate1
                others = np.delete(nds_a[ing], np.where(np.flatnonzero(ing) == jbest - 1)[0])
                if others.size:
                    mg = min(mg, float(np.min(np.abs(others - nbest))))
            else:
                mg = min(mg, abs(outer - gate2))
            margins.append(mg)
    zf = np.array(zf_cols, dtype=float).T.reshape(2, -1) if zf_cols else np.zeros((2, 0))
    zn = np.array(zn_cols, dtype=float).T.reshape(2, -1) if zn_cols else np.zeros((2, 0))
    return zf, np.array(idf, dtype=np.int64), zn


# --------------------------------------------------------------------------------------
# sim/sim-utils.jl and sim/ekfslam-sim.jl -- headless loop with injected noise
# --------------------------------------------------------------------------------------
def ellipse(x, P, nsigma, nsegs):
    """src/common.jl:236-244 -- nsigma*sqrtm(P)*[cos(phi)'; sin(phi)'] .+ x[1:2],
    phi = 0:2*pi/nsegs:2*pi (nsegs + 1 points).  sqrtm of the symmetric 2 x 2 through its
    eigen-decomposition, as Julia's sqrtm does for a symmetric argument."""
    phi = np.arange(nsegs + 1) * (2 * PI / nsegs)
    A = np.asarray(P, dtype=float)
    w, V = np.linalg.eigh((A + A.T) * 0.5)
    sq = (V * np.sqrt(w)) @ V.T
    return nsigma * sq @ np.vstack([np.cos(phi), np.sin(phi)]) + np.asarray(x, dtype=float)[0:2].reshape(2, 1)


def compute_landmark_ellipses(x, P):
    """src/common.jl:247-262 -- 2nf x (nsegs+1); rows 2i-1:2i are landmark i's contour."""
    nsigma, nsegs = 2, 12
    nf = (len(x) - 3) // 2
    out = np.zeros((2 * nf, nsegs + 1))
    for i in range(1, nf + 1):
        j0 = 2 * i + 1          # 0-based start of (2i+2):(2i+3)
        out[2 * i - 2:2 * i, :] = ellipse(x[j0:j0 + 2], P[j0:j0 + 2, j0:j0 + 2], nsigma, nsegs)
    return out


def local_to_global(l, g):
    """src/common.jl:112-128: rotate the 2 x n local positions by the pose heading and translate."""
    l = np.asarray(l, dtype=float)
    phi = g[2]
    Rm = np.array([[math.cos(phi), -math.sin(phi)], [math.sin(phi), math.cos(phi)]])
    q = np.zeros_like(l)
    q[0:2, :] = Rm @ l[0:2, :] + np.asarray(g[0:2], dtype=float).reshape(2, 1)
    return q


def laser_lines(z, vehicle_pose):
    """src/common.jl:265-283: 4 x nz, columns [vx; vy; fx; fy]."""
    z = np.asarray(z, dtype=float).reshape(2, -1)
    nl = z.shape[1]
    lines = np.zeros((4, nl))
    lines[0, :] = vehicle_pose[0]
    lines[1, :] = vehicle_pose[1]
    r, b = z[0, :], z[1, :]
    xy_local = np.vstack([r * np.cos(b), r * np.sin(b)])
    lines[2:4, :] = local_to_global(xy_local, vehicle_pose)
    return lines


def feature_ellipses(x, cov):
    """sim/browser/wsserver.jl:72-85: 5 x nf, [x_j; sqrt(l); atan2(u[2,1], u[1,1])] with l, u = eig(cov[j, j])
    (ascending eigenvalues; the eigenvector's sign -- hence phi modulo pi -- is LAPACK's choice)."""
    nf = (len(x) - 3) // 2
    out = np.zeros((5, nf))
    for i in range(1, nf + 1):
        j0 = 2 * i + 1          # 0-based start of 1-based (2i+2):(2i+3)
        lam, u = np.linalg.eigh(cov[j0:j0 + 2, j0:j0 + 2])
        out[:, i - 1] = [x[j0], x[j0 + 1], math.sqrt(lam[0]), math.sqrt(lam[1]), math.atan2(u[1, 0], u[0, 0])]
    return out


def nearby_landmark_indices(pose, lm, rmax):
    """sim/sim-utils.jl:12-28 (returns 0-based indices)."""
    phi = pose[2]
    out = []
    for i in range(lm.shape[1]):
        dx = lm[0, i] - pose[0]
        dy = lm[1, i] - pose[1]
        if (dx * math.cos(phi) + dy * math.sin(phi)) > 0:
            if (dx * dx + dy * dy) < rmax ** 2:
                out.append(i)
    return out


def get_observations(vehicle: Vehicle, landmarks, R, normals):
    """sim/sim-utils.jl:53-75.  `normals` is a callable n -> array of n standard normals,
    called for the range noises first and then for the bearing noises (the reference's
    draw order, :68)."""
    x, y, phi = vehicle.pose
    inearby = nearby_landmark_indices(vehicle.pose, landmarks, vehicle.sensor_range)
    ncols = len(inearby)
    z = np.zeros((2, ncols))
    for c, i in enumerate(inearby):
        dx = landmarks[0, i] - x
        dy = landmarks[1, i] - y
        z[0, c] = math.sqrt(dx * dx + dy * dy)
        z[1, c] = math.atan2(dy, dx) - phi
    if ncols > 0:
        z[0, :] = z[0, :] + normals(ncols) * math.sqrt(R[0, 0])
        z[1, :] = z[1, :] + normals(ncols) * math.sqrt(R[1, 1])
    return z, [i + 1 for i in inearby]


def load_waypoints(path: str) -> np.ndarray:
    """src/common.jl:84-87 -- header line then tab-separated x,y; returned 2 x N."""
    return np.loadtxt(path, skiprows=1).T.copy()


def sim_constants():
    """Literals of sim/ekfslam-sim.jl:62-76,114."""
    sigma_speed = 0.5
    sigma_steer = 3.0 * PI / 180
    Q = np.array([[sigma_speed ** 2, 0.0], [0.0, sigma_steer ** 2]])
    sigmaR = 0.1
    sigmaB = 1.0 * PI / 180
    R = np.array([[sigmaR ** 2, 0.0], [0.0, sigmaB ** 2]])
    return dict(d_min=1.0, Q=Q, R=R, dt=0.025, dt_obs=8 * 0.025, gate1=4.0, gate2=25.0)


def sim(waypoints, landmarks, ctrl_noise, obs_noise, nlaps=2, record=None, max_steps=10000,
        fns=None):
    """sim/ekfslam-sim.jl:54-143 without the real-time sleep and the monitor.

    ctrl_noise: (2, >= nsteps) standard normals: row 0 speed, row 1 steering (sim-utils.jl:36-37).
    obs_noise : flat array of standard normals consumed in draw order (sim-utils.jl:68).
    record    : optional list; receives one dict per filter call (golden trace).
    fns       : optional dict overriding predict/associate/update/add_features with callables
                of the same (x, P, ...) signatures (used by tests to drive other back ends).
    Returns dict(true_track 3 x nsteps, slam_track 3 x nsteps, x, P, nsteps, nupdates)."""
    k = sim_constants()
    Q, R, dt, dt_obs = k["Q"], k["R"], k["dt"], k["dt_obs"]
    f_predict = (fns or {}).get("predict", predict)
    f_associate = (fns or {}).get("associate", associate)
    f_update = (fns or {}).get("update", update)
    f_add = (fns or {}).get("add_features", add_features)
    vehicle = Vehicle()
    vehicle.pose = initial_pose(waypoints)
    x = vehicle.pose.copy()
    P = np.zeros((3, 3))
    true_track = np.zeros((3, max_steps))
    slam_track = np.zeros((3, max_steps))
    nsteps = 0
    nupd = 0
    cursor = [0]

    def normals(n):
        a = obs_noise[cursor[0]:cursor[0] + n]
        if len(a) != n:
            raise ValueError("obs_noise exhausted")
        cursor[0] += n
        return a

    dtsum = 0
    while vehicle.waypoint_id != 0:
        steer(vehicle, waypoints, k["d_min"], dt)
        if vehicle.waypoint_id == 0 and nlaps > 1:
            vehicle.waypoint_id = 1
            nlaps -= 1
        step_vehicle(vehicle, dt)
        # add_control_noise! (sim/sim-utils.jl:35-38)
        vehicle.measured_speed = vehicle.target_speed + ctrl_noise[0, nsteps] * math.sqrt(Q[0, 0])
        vehicle.measured_gamma = vehicle.target_gamma + ctrl_noise[1, nsteps] * math.sqrt(Q[1, 1])
        if record is not None:
            record.append(dict(op="predict", x=x.copy(), P=P.copy(), g=vehicle.measured_gamma,
                               v=vehicle.measured_speed))
        x, P = f_predict(x, P, vehicle.measured_gamma, vehicle.measured_speed,
                         vehicle.wheelbase, Q, dt)
        if record is not None:
            record[-1].update(x_out=x.copy(), P_out=P.copy())
        dtsum += dt
        if dtsum > dt_obs:
            dtsum = 0
            nupd += 1
            z, _tags = get_observations(vehicle, landmarks, R, normals)
            margins = []
            if f_associate is associate:
                zf, idf, zn = associate(x, P, z, R, k["gate1"], k["gate2"], margins)
            else:
                zf, idf, zn = f_associate(x, P, z, R, k["gate1"], k["gate2"])
            if record is not None:
                record.append(dict(op="observe", x=x.copy(), P=P.copy(), z=z.copy(), zf=zf.copy(),
                                   idf=idf.copy(), zn=zn.copy(),
                                   margin=min(margins) if margins else math.inf))
            x, P = f_update(x, P, zf, R, idf)
            if record is not None:
                record[-1].update(x_upd=x.copy(), P_upd=P.copy())
            x, P = f_add(x, P, zn, R)
            if record is not None:
                record[-1].update(x_out=x.copy(), P_out=P.copy())
        nsteps += 1
        true_track[:, nsteps - 1] = vehicle.pose
        slam_track[:, nsteps - 1] = x[0:3]
    return dict(true_track=true_track[:, :nsteps], slam_track=slam_track[:, :nsteps],
                x=x, P=P, nsteps=nsteps, nupdates=nupd, obs_used=cursor[0])


def make_inputs(n_landmarks: int, map_seed: int, noise_seed: int, max_steps=10000,
                obs_pool=None):
    """Synthetic inputs of SURVEY.md section 8(d) config 1: landmarks are integers uniform in
    5..95 for both axes (the make_landmarks quirk, sim/sim-utils.jl:5); noise arrays are
    standard normals from numpy PCG64."""
    lm = np.random.Generator(np.random.PCG64(map_seed)).integers(5, 96, size=(2, n_landmarks))
    lm = lm.astype(float)
    g = np.random.Generator(np.random.PCG64(1000003 + noise_seed))
    ctrl = g.standard_normal((2, max_steps))
    obs = g.standard_normal(obs_pool if obs_pool is not None else 2 * max(n_landmarks, 1) * 400)
    return lm, ctrl, obs
