"""CPU tests of the oracle (test infrastructure): the numpy literal restatement against the
committed golden vectors and against its independent plain-C twin.  Parity with the reference
runtime itself is unpinned (no Julia here, no tests in the reference)."""
import math
from pathlib import Path

import numpy as np
import pytest

from conftest import rel_err
from oracle import slam_ref as ref
from twin_ffi import Twin, d

GOLD = Path(__file__).resolve().parent / "golden"


@pytest.fixture(scope="module")
def kat():
    return np.load(GOLD / "kat.npz")


@pytest.fixture(scope="module")
def tw(twin):
    return Twin(twin)


def test_mpi_to_pi_single_wrap(twin):
    for phi in (0.0, 3.0, -3.0, math.pi, -math.pi, 3.5, -3.5, 7.0, -7.0, 10.0, -10.0):
        assert ref.mpi_to_pi(phi) == twin.twin_mpi_to_pi(phi)
    assert ref.mpi_to_pi(7.0) == 7.0 - 2 * math.pi          # single wrap: still > pi
    assert ref.mpi_to_pi(3 * math.pi + 0.5) > math.pi


@pytest.mark.parametrize("n", [3, 5, 23, 73])
def test_oracle_matches_golden_and_twin(kat, tw, n):
    k = ref.sim_constants()
    Q, R, dt = k["Q"], k["R"], k["dt"]
    t = f"n{n}"
    x, P = kat[f"{t}_x"], kat[f"{t}_P"]
    xp, Pp = ref.predict(x.copy(), P.copy(), 0.07, 7.6, 4.0, Q, dt)
    assert rel_err(xp, kat[f"{t}_predict_x"]) < 1e-14 and rel_err(Pp, kat[f"{t}_predict_P"]) < 1e-13
    xt, Pt = tw.predict(x, P, 0.07, 7.6, 4.0, Q, dt)
    assert rel_err(xt, xp) < 1e-12 and rel_err(Pt, Pp) < 1e-12
    z = kat[f"{t}_z"]
    zf, idf, zn = ref.associate(x, P, z, R, k["gate1"], k["gate2"])
    assert np.array_equal(idf, kat[f"{t}_idf"]) and np.array_equal(zf, kat[f"{t}_zf"])
    assert np.array_equal(zn, kat[f"{t}_zn"])
    zf_t, idf_t, zn_t = tw.associate(x, P, z, R, k["gate1"], k["gate2"])
    assert np.array_equal(idf_t, idf) and np.array_equal(zf_t, zf) and np.array_equal(zn_t, zn)
    xu, Pu = ref.update(x, P, zf, R, idf)
    assert rel_err(xu, kat[f"{t}_update_x"]) < 1e-13 and rel_err(Pu, kat[f"{t}_update_P"]) < 1e-12
    xt, Pt = tw.update(x, P, zf, R, idf)
    assert rel_err(xt, xu) < 1e-12 and rel_err(Pt, Pu) < 1e-12
    xa, Pa = ref.add_features(xu, Pu, zn, R)
    assert rel_err(xa, kat[f"{t}_add_x"]) < 1e-13 and rel_err(Pa, kat[f"{t}_add_P"]) < 1e-12
    xt, Pt = tw.add_features(xu, Pu, zn, R)
    assert rel_err(xt, xa) < 1e-12 and rel_err(Pt, Pa) < 1e-12
    if n > 3:
        nf = (n - 3) // 2
        g = kat[f"{t}_nis_nd"]
        for j in range(1, nf + 1):
            nis, nd = ref.compute_association(x, P, z[:, 0], R, j)
            assert abs(nis - g[j - 1, 0]) <= 1e-10 * max(1, abs(nis))
            nt, ndt = tw.compute_association(x, P, z[:, 0], R, j)
            assert abs(nt - nis) <= 1e-9 * max(1, abs(nis)) and abs(ndt - nd) <= 1e-9 * max(1, abs(nd))
        zp, H = ref.predict_observation(x, nf)
        assert rel_err(zp, kat[f"{t}_obs_z"]) < 1e-15 and rel_err(H, kat[f"{t}_obs_H"]) < 1e-15
        assert np.count_nonzero(H) <= 9            # 5 columns x 2 rows, H[0,2] = 0


def test_properties():
    k = ref.sim_constants()
    R = k["R"]
    rng = np.random.default_rng(1)
    n = 13
    x = np.concatenate([[10.0, 10.0, 0.1], rng.uniform(0, 40, n - 3)])
    L = rng.normal(0, 0.1, (n, 5))
    P = L @ L.T + 0.05 * np.eye(n)
    x1, P1 = ref.update(x, P, np.zeros((2, 0)), R, np.zeros(0, dtype=int))      # m = 0: identity
    assert np.array_equal(x1, x) and np.array_equal(P1, P)
    x1, P1 = ref.add_features(x, P, np.zeros((2, 0)), R)                        # k = 0: identity
    assert x1 is x and P1 is P
    z = np.array([[5.0, 9.0], [0.3, -0.4]])
    zf, idf, zn = ref.associate(x[:3], P[:3, :3], z, R, 4.0, 25.0)               # Nf = 0: all new
    assert zf.shape == (2, 0) and len(idf) == 0 and np.array_equal(zn, z)
    zp, _ = ref.predict_observation(x, 2)
    xu, Pu = ref.update(x, P, (zp + [0.05, 0.01]).reshape(2, 1), R, [2])
    assert np.allclose(Pu, Pu.T, atol=1e-15) and np.min(np.linalg.eigvalsh(Pu)) > 0
    assert np.trace(Pu) < np.trace(P)                                           # information gained


@pytest.mark.parametrize("nl", [10, 35])
def test_course1_runs_match_golden_and_twin(course1, twin, nl):
    g = np.load(GOLD / "course1_runs.npz")
    for seed in range(4):     # the golden file holds seeds 0..7; four of them are re-run here (time)
        lm, ctrl, obs = ref.make_inputs(nl, 0, seed)
        rec = []
        r = ref.sim(course1, lm, ctrl, obs, record=rec)
        t = f"nl{nl}_s{seed}"
        assert list(g[f"{t}_counts"]) == [r["nsteps"], r["nupdates"], len(r["x"]), r["obs_used"]]
        assert rel_err(r["x"], g[f"{t}_x"]) < 1e-9
        assert rel_err(np.diag(r["P"]), g[f"{t}_Pdiag"]) < 1e-9
        assert rel_err(r["slam_track"][:, ::50], g[f"{t}_track_every50"]) < 1e-9
        idfs = np.concatenate([e["idf"] for e in rec if e["op"] == "observe"])
        assert np.array_equal(idfs, g[f"{t}_idf_concat"])
        # independent C twin, same injected noise
        ms = 10000
        wp = np.ascontiguousarray(course1.T).reshape(-1)
        lmf = np.ascontiguousarray(lm.T).reshape(-1)
        cn = np.ascontiguousarray(ctrl.T).reshape(-1)
        tt, st = np.zeros(3 * ms), np.zeros(3 * ms)
        ld = 3 + 2 * nl
        xo, Po = np.zeros(ld), np.zeros(ld * ld)
        import ctypes as C
        counts = (C.c_int * 4)()
        mg = C.c_double()
        rc = twin.twin_sim(d(wp), course1.shape[1], d(lmf), nl, d(cn), d(obs), len(obs), 0, 0, 2, ms, d(tt),
                           d(st), d(xo), d(Po), counts, C.byref(mg))
        if len(r["x"]) > ld:
            # the filter created spurious landmarks (a data-association failure of the algorithm
            # itself, e.g. 35 landmarks / seed 1 ends with 63): the fixed-capacity twin stops with
            # -2 where the map first outgrows 3 + 2*nl; tracks agree up to that step
            assert rc == -2 and counts[2] == ld
            ns = counts[0]
            assert 0 < ns < r["nsteps"]
        else:
            assert rc == 0 and counts[0] == r["nsteps"] and counts[1] == r["nupdates"] and counts[2] == len(r["x"])
            ns = r["nsteps"]
            assert rel_err(xo[:counts[2]], r["x"]) < 1e-6
        assert np.array_equal(tt[:3 * ns].reshape(ns, 3).T, r["true_track"][:, :ns])
        assert rel_err(st[:3 * ns].reshape(ns, 3).T, r["slam_track"][:, :ns]) < 1e-6
        assert mg.value > 1e-9        # no association decision was within rounding of a threshold


def test_update_cadence_every_ninth_step(course1):
    """dt_obs = 8*dt with a strict '>' on the accumulated float sum fires on every 9th step."""
    lm, ctrl, obs = ref.make_inputs(10, 0, 0)
    r = ref.sim(course1, lm, ctrl, obs)
    assert r["nupdates"] == r["nsteps"] // 9


def _fd_jacobian(f, x0, eps=1e-6):
    """Central finite differences of f: R^n -> R^k at x0 (angles in f's output are compared unwrapped)."""
    f0 = np.asarray(f(x0), dtype=float)
    J = np.zeros((f0.size, x0.size))
    for q in range(x0.size):
        dx = np.zeros_like(x0)
        dx[q] = eps
        J[:, q] = (np.asarray(f(x0 + dx)) - np.asarray(f(x0 - dx))) / (2 * eps)
    return J


def test_oracle_against_textbook_identities():
    """The reference ships no known answers (parity unpinned), so besides following its text line by
    line the oracle is checked against what an EKF must satisfy, derived independently of the text:
    Jacobians against finite differences of the models, the update against the information form, the
    augmentation against first-order propagation through the inverse observation model."""
    k = ref.sim_constants()
    Q, R, dt = k["Q"], k["R"], k["dt"]
    rng = np.random.default_rng(42)
    nf = 6
    n = 3 + 2 * nf
    x = np.concatenate([[20.0, 30.0, 0.4], rng.uniform(0, 60, 2 * nf)])
    L = rng.normal(0, 0.2, (n, n))
    P = L @ L.T + 0.1 * np.eye(n)
    g, v, w = 0.12, 7.3, 4.0

    # observation model: H = dh/dx
    for idf in (1, 4, 6):
        zp, H = ref.predict_observation(x, idf)
        J = _fd_jacobian(lambda xx: ref.predict_observation(xx, idf)[0], x)
        assert np.max(np.abs(H - J)) < 1e-7
        f = 3 + 2 * (idf - 1)
        assert abs(zp[0] - np.hypot(x[f] - x[0], x[f + 1] - x[1])) < 1e-12

    # predict: x' = f(x, u), P' = F P F' + Gu Q Gu' with F, Gu the Jacobians of f
    def motion(xu):
        xx, gg, vv = xu[:3], xu[3], xu[4]
        return np.array([xx[0] + vv * dt * math.cos(gg + xx[2]), xx[1] + vv * dt * math.sin(gg + xx[2]),
                         xx[2] + vv * dt * math.sin(gg) / w])
    Jm = _fd_jacobian(motion, np.array([x[0], x[1], x[2], g, v]))
    Fv, Gu = Jm[:, :3], Jm[:, [4, 3]]                    # Q is ordered (speed, steering): src/ekf.jl:27-29
    F = np.eye(n)
    F[:3, :3] = Fv
    Pexp = F @ P @ F.T
    Pexp[:3, :3] += Gu @ Q @ Gu.T
    xp, Pp = ref.predict(x.copy(), P.copy(), g, v, w, Q, dt)
    assert np.max(np.abs(Pp - Pexp)) < 1e-6 * np.max(np.abs(Pexp))
    assert np.max(np.abs(xp[:3] - motion(np.array([x[0], x[1], x[2], g, v])))) < 1e-12
    assert np.array_equal(xp[3:], x[3:])

    # update: information form  inv(P+) = inv(P) + H' inv(RR) H,  x+ = x + P+ H' inv(RR) v
    ids = np.array([2, 5, 3])
    z = np.zeros((2, len(ids)))
    Hs, vs = [], []
    for c, idf in enumerate(ids):
        zp, H = ref.predict_observation(x, int(idf))
        z[:, c] = zp + rng.normal(0, [0.1, 0.01])
        Hs.append(H)
        vs.append([z[0, c] - zp[0], ref.mpi_to_pi(z[1, c] - zp[1])])
    Hall, vall = np.vstack(Hs), np.concatenate(vs)
    RRi = np.kron(np.eye(len(ids)), np.linalg.inv(R))
    xu, Pu = ref.update(x, P, z, R, ids)
    info = np.linalg.inv(P) + Hall.T @ RRi @ Hall
    assert np.max(np.abs(np.linalg.inv(Pu) - info)) < 1e-7 * np.max(np.abs(info))
    assert np.max(np.abs(xu - (x + Pu @ Hall.T @ RRi @ vall))) < 1e-9

    # add_features: [x; g(x_v, z)] with covariance J blkdiag(P, R) J'
    zn = np.array([[12.0], [0.7]])
    def augment(xz):
        xx, r, b = xz[:n], xz[n], xz[n + 1]
        return np.concatenate([xx, [xx[0] + r * math.cos(xx[2] + b), xx[1] + r * math.sin(xx[2] + b)]])
    Ja = _fd_jacobian(augment, np.concatenate([x, zn[:, 0]]))
    big = np.zeros((n + 2, n + 2))
    big[:n, :n] = P
    big[n:, n:] = R
    xa, Pa = ref.add_features(x, P, zn, R)
    assert np.max(np.abs(xa - augment(np.concatenate([x, zn[:, 0]])))) < 1e-12
    assert np.max(np.abs(Pa - Ja @ big @ Ja.T)) < 1e-6 * np.max(np.abs(Pa))

    # association distances: nis = v' inv(S) v, nd = nis + log det S with S = H P H' + R
    zq = z[:, 0]
    for j in (1, 2, 6):
        zp, H = ref.predict_observation(x, j)
        vv = np.array([zq[0] - zp[0], ref.mpi_to_pi(zq[1] - zp[1])])
        S = H @ P @ H.T + R
        nis, nd = ref.compute_association(x, P, zq, R, j)
        assert abs(nis - vv @ np.linalg.solve(S, vv)) < 1e-9 * max(1.0, nis)
        assert abs(nd - (nis + math.log(np.linalg.det(S)))) < 1e-9 * max(1.0, abs(nd))


def test_pregate_bound_against_the_literal_oracle(course1):
    """The inequality chain behind the CUDA spatial pre-gate (slam.jl_b200/csrc/gate.cu), checked
    against the literal restatement of compute_association (src/data-association.jl:53-63) with no GPU involved:
        nis >= v_r^2 / S_rr,   S_rr <= 2 (P_xx + P_yy) + 4 sig2max + R_rr,
    so a pair with (z_r - d)^2 > 4 G S_rr_max (G = max(gate1, gate2)) has nis > G with a factor 4 to spare and can
    change neither `nis < gate1` (:30) nor `outer > gate2` (:46).  Random dense states and every observation step
    of a course1 run (the states an EKF actually produces)."""
    k = ref.sim_constants()
    R, g1, g2 = k["R"], k["gate1"], k["gate2"]
    G = max(g1, g2)

    def check(x, P, z):
        nf = (len(x) - 3) // 2
        if nf == 0 or z.shape[1] == 0:
            return 0, 0
        sig2 = max(0.0, float(np.max(np.diag(P)[3:])))
        srr_max = 2.0 * (abs(P[0, 0]) + abs(P[1, 1])) + 4.0 * sig2 + abs(R[0, 0])
        skipped = total = 0
        for i in range(z.shape[1]):
            for j in range(1, nf + 1):
                zp, H = ref.predict_observation(x, j)
                S = H @ P @ H.T + R
                nis, _ = ref.compute_association(x, P, z[:, i], R, j)
                vr = z[0, i] - zp[0]
                assert S[0, 0] <= srr_max * (1 + 1e-12)
                assert nis >= vr * vr / S[0, 0] * (1 - 1e-9)
                total += 1
                if vr * vr > 4.0 * G * srr_max:          # what k_gate skips
                    skipped += 1
                    assert nis > 3.9 * G                 # far above both gates
        return skipped, total

    rng = np.random.default_rng(5)
    sk = tot = 0
    for nf, side in ((40, 60.0), (120, 400.0)):
        n = 3 + 2 * nf
        x = np.concatenate([[side / 2, side / 2, 0.3], rng.uniform(0, side, size=2 * nf)])
        L = rng.normal(0, 0.05, size=(n, 8))
        P = L @ L.T + np.diag(rng.uniform(0.01, 0.1, size=n))
        z = np.array([rng.uniform(1, side / 2, 10), rng.uniform(-3, 3, 10)])
        a, b = check(x, P, z)
        sk, tot = sk + a, tot + b
    lm, ctrl, obs = ref.make_inputs(35, 0, 1)
    rec = []
    ref.sim(course1, lm, ctrl, obs, nlaps=1, record=rec)
    steps = [r for r in rec if r["op"] == "observe"]
    for r in steps[::6]:
        a, b = check(r["x"], r["P"], r["z"])
        sk, tot = sk + a, tot + b
    print(f"\n[pre-gate bound] {sk} of {tot} pairs skipped")
    assert tot > 3000 and sk > tot // 4      # the bound is not vacuous: it discards a large share of the pairs


def test_pregate_bound_fuzz_against_the_twin(twin):
    """The same implication -- outside the annulus => nis > max(gate1, gate2) -- fuzzed over 300 random
    positive definite states with the plain-C twin's compute_association: strong vehicle / landmark
    correlations, heading variances up to 1 rad^2, observation covariances from 1e-8 to 1, gates from 0.1
    to 1e3, landmarks from 0.1 m to 2 km away."""
    from twin_ffi import Twin
    tw = Twin(twin)
    rng = np.random.default_rng(123)
    pairs = skipped = 0
    worst = np.inf
    for trial in range(300):
        nf = int(rng.integers(1, 40))
        n = 3 + 2 * nf
        side = 10.0 ** rng.uniform(0, 3.3)
        x = np.concatenate([[0.0, 0.0, rng.uniform(-3, 3)], rng.uniform(-side, side, size=2 * nf)])
        rank = int(rng.integers(1, n + 1))
        A = rng.normal(0, 10.0 ** rng.uniform(-3, 0), size=(n, rank))
        A[2, :] *= 10.0 ** rng.uniform(-1, 1)                       # heading
        P = A @ A.T + np.diag(10.0 ** rng.uniform(-8, -1, size=n))
        R = np.diag(10.0 ** rng.uniform(-8, 0, size=2))
        if rng.random() < 0.3:
            c = rng.uniform(-0.9, 0.9) * math.sqrt(R[0, 0] * R[1, 1])
            R[0, 1] = R[1, 0] = c
        g1, g2 = 10.0 ** rng.uniform(-1, 3, size=2)
        G = max(g1, g2)
        sig2 = float(np.max(np.diag(P)[3:]))
        srr_max = 2.0 * (P[0, 0] + P[1, 1]) + 4.0 * sig2 + R[0, 0]
        rho2 = 4.0 * G * srr_max
        for _ in range(6):
            z = np.array([rng.uniform(0, 2 * side), rng.uniform(-3.2, 3.2)])
            for j in range(1, nf + 1):
                d = math.hypot(x[3 + 2 * (j - 1)] - x[0], x[4 + 2 * (j - 1)] - x[1])
                pairs += 1
                if (z[0] - d) ** 2 > rho2:
                    skipped += 1
                    nis, _ = tw.compute_association(x, P, z, R, j)
                    assert nis > G, (trial, j, nis, G)
                    worst = min(worst, nis / G)
    print(f"\n[pre-gate fuzz] {skipped} of {pairs} pairs outside the annulus; smallest nis / G among them {worst:.2f}")
    assert skipped > pairs // 10 and worst > 3.9
