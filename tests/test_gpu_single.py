"""GPU parity of the single-filter path (K1-K7) through the C ABI against the CPU oracle.

Tolerances (BASELINE.json north_star): association indices bit-exact; x and P within 1e-9
relative (max-norm) per step when every step is re-seeded from the oracle; 1e-6 on
end-of-run trajectories in closed loop.
"""
import math

import numpy as np
import pytest

from conftest import rel_err, state_err
from oracle import slam_ref as ref

pytestmark = pytest.mark.gpu

STEP_TOL = 1e-9
TRACK_TOL = 1e-6


@pytest.fixture(scope="module")
def sc():
    import slamcuda
    if slamcuda.device_count() == 0:
        pytest.fail("no CUDA device: the gpu tests must run on the B200 box")
    return slamcuda


@pytest.fixture(scope="module")
def traces(course1):
    out = {}
    for nl in (10, 35):
        lm, ctrl, obs = ref.make_inputs(nl, 0, 0)
        rec = []
        res = ref.sim(course1, lm, ctrl, obs, record=rec)
        out[nl] = (lm, ctrl, obs, rec, res)
    return out


def test_mpi_to_pi(sc):
    for phi in (0.0, 3.0, -3.0, math.pi, -math.pi, 3.5, -3.5, 7.0, -7.0, 10.0):
        assert sc.mpi_to_pi(phi) == ref.mpi_to_pi(phi)


@pytest.mark.parametrize("nl", [10, 35])
def test_step_reseeded_parity(sc, traces, nl):
    """Every recorded call of a course1 run: seed the device from the oracle's input state,
    run the CUDA op, compare with the oracle's output."""
    _lm, _ctrl, _obs, rec, _res = traces[nl]
    k = ref.sim_constants()
    Q, R, dt = k["Q"], k["R"], k["dt"]
    veh = sc.Vehicle()
    worst = dict(predict=0.0, update=0.0, add=0.0)
    min_margin = math.inf
    n_obs = 0
    for i, e in enumerate(rec):
        if e["op"] == "predict":
            if i % 7 and len(e["x"]) > 3:   # sample the 2800 predict steps, keep all n = 3 ones
                continue
            st = sc.EKFSlamState(e["x"], e["P"])
            veh.measured_gamma, veh.measured_speed = e["g"], e["v"]
            x, P = sc.predict(st, veh, Q, dt)
            worst["predict"] = max(worst["predict"], state_err(x, P, e["x_out"], e["P_out"]))
        else:
            n_obs += 1
            st = sc.EKFSlamState(e["x"], e["P"])
            zf, idf, zn = sc.associate(st, e["z"], R, k["gate1"], k["gate2"])
            assert np.array_equal(idf.reshape(-1), e["idf"]), f"association differs at record {i}"
            assert np.array_equal(zf, e["zf"]) and np.array_equal(zn, e["zn"])
            min_margin = min(min_margin, e["margin"])
            x, P = sc.update(st, e["zf"], R, e["idf"])
            worst["update"] = max(worst["update"], state_err(x, P, e["x_upd"], e["P_upd"]))
            st2 = sc.EKFSlamState(e["x_upd"], e["P_upd"])
            x, P = sc.add_features(st2, e["zn"], R)
            assert len(x) == len(e["x_out"])
            worst["add"] = max(worst["add"], state_err(x, P, e["x_out"], e["P_out"]))
    print(f"\n[parity nl={nl}] worst rel err {worst}; {n_obs} observation steps; "
          f"smallest association margin {min_margin:.3e}")
    assert n_obs > 300
    assert max(worst.values()) < STEP_TOL


@pytest.mark.parametrize("nl", [10, 35])
def test_closed_loop_course1(sc, traces, course1, nl):
    """sim! driven end to end through the CUDA library, state resident on the device."""
    lm, ctrl, obs, _rec, res = traces[nl]
    from slamcuda import sim as gsim
    out = gsim.run(course1, lm, ctrl, obs, n_max=3 + 2 * nl)
    assert out["nsteps"] == res["nsteps"] and out["nupdates"] == res["nupdates"]
    assert len(out["x"]) == len(res["x"])
    e_track = float(np.max(np.abs(out["slam_track"] - res["slam_track"])))
    scale = float(np.max(np.abs(res["slam_track"])))
    print(f"\n[closed loop nl={nl}] track max abs err {e_track:.3e} (scale {scale:.1f}), "
          f"x rel {rel_err(out['x'], res['x']):.3e}, P rel {rel_err(out['P'], res['P']):.3e}")
    assert e_track / scale < TRACK_TOL
    assert np.array_equal(out["true_track"], res["true_track"])
    assert rel_err(out["x"], res["x"]) < TRACK_TOL
    assert rel_err(out["P"], res["P"]) < TRACK_TOL
    # P stays symmetric positive semi-definite
    assert np.array_equal(out["P"], out["P"].T)
    assert np.min(np.linalg.eigvalsh(out["P"])) > -1e-12


def _random_state(rng, nf):
    n = 3 + 2 * nf
    x = np.concatenate([[50.0, 50.0, 0.3], rng.uniform(20, 80, size=2 * nf)])
    L = rng.normal(0, 0.05, size=(n, 8))
    P = L @ L.T + np.diag(rng.uniform(0.01, 0.1, size=n))
    return x, P


def test_edge_cases(sc):
    k = ref.sim_constants()
    R = k["R"]
    rng = np.random.default_rng(5)
    x, P = _random_state(rng, 4)
    st = sc.EKFSlamState(x, P)
    # update with m = 0 is the identity (first observation step of every run)
    x1, P1 = sc.update(st, np.zeros((2, 0)), R, np.zeros(0, dtype=np.int64))
    assert np.array_equal(x1, x) and rel_err(P1, P) == 0.0
    # add_features with zero columns is the identity
    x1, P1 = sc.add_features(st, np.zeros((2, 0)), R)
    assert np.array_equal(x1, x) and rel_err(P1, P) == 0.0
    # associate with Nf = 0 sends every observation to zn (outer = Inf > gate2)
    st0 = sc.EKFSlamState(x[:3], P[:3, :3])
    z = np.array([[10.0, 20.0, 5.0], [0.1, -0.2, 3.0]])
    zf, idf, zn = sc.associate(st0, z, R, 4.0, 25.0)
    assert zf.shape == (2, 0) and idf.size == 0 and np.array_equal(zn, z)
    # associate with no observations
    zf, idf, zn = sc.associate(st, np.zeros((2, 0)), R, 4.0, 25.0)
    assert zf.shape == (2, 0) and zn.shape == (2, 0)
    # duplicate landmark ids in one update are legal
    zp, _ = ref.predict_observation(x, 2)
    zdup = np.array([[zp[0] + 0.05, zp[0] - 0.03], [zp[1] + 0.01, zp[1] - 0.02]])
    xg, Pg = sc.update(st, zdup, R, np.array([2, 2]))
    xo, Po = ref.update(x, P, zdup, R, np.array([2, 2]))
    assert rel_err(xg, xo) < STEP_TOL and rel_err(Pg, Po) < STEP_TOL
    # bearing innovation that crosses +-pi
    x2 = x.copy()
    x2[2] = 3.1
    x2[3:5] = [x2[0] - 10.0, x2[1] - 0.5]
    zp, _ = ref.predict_observation(x2, 1)
    zc = np.array([[zp[0]], [zp[1] + 2 * math.pi - 0.01]])
    st2 = sc.EKFSlamState(x2, P)
    xg, Pg = sc.update(st2, zc, R, np.array([1]))
    xo, Po = ref.update(x2, P, zc, R, np.array([1]))
    assert rel_err(xg, xo) < STEP_TOL and rel_err(Pg, Po) < STEP_TOL
    # out-of-range landmark id -> BadIndex (Julia: BoundsError)
    with pytest.raises(sc.BadIndex):
        sc.update(st, zdup[:, :1], R, np.array([99]))
    # non-PD innovation covariance -> NotPositiveDefinite (Julia: chol throws)
    with pytest.raises(sc.NotPositiveDefinite):
        sc.update(sc.EKFSlamState(x, -P), zdup[:, :1], -R, np.array([2]))
    # capacity
    from slamcuda.slam import _Handle
    h = _Handle(0, 5, grow=False)
    h.set_state(x[:5], P[:5, :5])
    with pytest.raises(sc.CapacityError):
        h.add_features(np.array([[5.0], [0.1]]), R)
    h.close()


def test_cuda_hand_derived_cases(sc):
    """The CUDA path against closed-form expected values worked out on paper (tests/handcases.py), independent of
    the oracle: predict, predict_observation, compute_association, the matched / dropped / new branches of
    associate, the strict-< tie (lowest id wins), update, add_features (second new landmark sees the first)."""
    import handcases as hc
    c = hc.case_predict()
    veh = sc.Vehicle(wheelbase=c["w"], measured_gamma=c["g"], measured_speed=c["v"])
    x, P = sc.predict(sc.EKFSlamState(c["x"], c["P"]), veh, c["Q"], c["dt"])
    assert np.allclose(x, c["x_out"], rtol=0, atol=1e-15) and np.allclose(P, c["P_out"], rtol=1e-14, atol=1e-18)
    c = hc.case_observe()
    zp, H = sc.predict_observation(c["x"], c["idf"])
    assert np.allclose(zp, c["zp"], rtol=0, atol=1e-15) and np.allclose(H, c["H"], rtol=0, atol=1e-16)
    nis, nd = sc.compute_association(c["x"], c["P"], c["z"], c["R"], c["idf"])
    assert abs(nis - c["nis"]) < 1e-13 and abs(nd - c["nd"]) < 1e-13
    for c in (hc.case_associate(), hc.case_tie()):
        for pregate in (True, False):
            st = sc.DeviceEKFSlamState(c["x"], c["P"], n_max=len(c["x"]))
            st.h.set_pregate(pregate)
            zf, idf, zn = sc.associate(st, c["z"], c["R"], c["gate1"], c["gate2"])
            st.h.close()
            assert list(idf.reshape(-1)) == c["idf"] and np.array_equal(zf, c["zf"]) and np.array_equal(zn, c["zn"])
    c = hc.case_update()
    x, P = sc.update(sc.EKFSlamState(c["x"], c["P"]), c["z"], c["R"], np.array(c["idf"]))
    assert np.allclose(x, c["x_out"], rtol=0, atol=1e-15) and np.allclose(P, c["P_out"], rtol=0, atol=1e-16)
    assert np.array_equal(P[5:, 5:], c["P"][5:, 5:]) and np.array_equal(x[5:], c["x"][5:])   # landmark 2 is uncorrelated
    c = hc.case_add_features()
    x, P = sc.add_features(sc.EKFSlamState(c["x"], c["P"]), c["z"], c["R"])
    assert np.allclose(x, c["x_out"], rtol=0, atol=1e-15) and np.allclose(P, c["P_out"], rtol=0, atol=1e-16)


@pytest.mark.parametrize("n", [3, 5, 23])
def test_cuda_within_1e12_of_the_mpmath_restatement(sc, n):
    """The 50-digit mpmath restatement of the reference's text (oracle/slam_mp.py, fixture tests/golden/kat_mp.npz
    generated by tests/golden/make_golden_mp.py) pins the CUDA path to that text, not to the float64 oracle: every
    function within 1e-12 (max-norm relative and element-wise), association ids equal."""
    from pathlib import Path
    gold = Path(__file__).resolve().parent / "golden"
    kat, mpk = np.load(gold / "kat.npz"), np.load(gold / "kat_mp.npz")
    k = ref.sim_constants()
    Q, R, dt = k["Q"], k["R"], k["dt"]
    t = f"n{n}"
    x, P, z = kat[f"{t}_x"], kat[f"{t}_P"], kat[f"{t}_z"]
    TOL = 1e-12
    veh = sc.Vehicle(wheelbase=4.0, measured_gamma=0.07, measured_speed=7.6)
    xp, Pp = sc.predict(sc.EKFSlamState(x, P), veh, Q, dt)
    e_pred = state_err(xp, Pp, mpk[f"{t}_predict_x"], mpk[f"{t}_predict_P"])
    zf, idf, zn = sc.associate(sc.EKFSlamState(x, P), z, R, k["gate1"], k["gate2"])
    assert np.array_equal(idf.reshape(-1), mpk[f"{t}_idf"])
    assert np.array_equal(zf, z[:, mpk[f"{t}_zf_cols"]]) and np.array_equal(zn, z[:, mpk[f"{t}_zn_cols"]])
    xu, Pu = sc.update(sc.EKFSlamState(x, P), zf, R, idf)
    e_upd = state_err(xu, Pu, mpk[f"{t}_update_x"].reshape(-1), mpk[f"{t}_update_P"])
    xa, Pa = sc.add_features(sc.EKFSlamState(kat[f"{t}_update_x"], kat[f"{t}_update_P"]), zn, R)
    e_add = state_err(xa, Pa, mpk[f"{t}_add_x"].reshape(-1), mpk[f"{t}_add_P"])
    print(f"\n[cuda vs mpmath n={n}] predict {e_pred:.2e}  update {e_upd:.2e}  add_features {e_add:.2e}")
    assert max(e_pred, e_upd, e_add) < TOL
    if n > 3:
        nf = (n - 3) // 2
        for j in range(1, nf + 1):
            nis, nd = sc.compute_association(x, P, z[:, 0], R, j)
            g = mpk[f"{t}_nis_nd"][j - 1]
            assert abs(nis - g[0]) <= 1e-11 * max(1.0, abs(g[0])) and abs(nd - g[1]) <= 1e-11 * max(1.0, abs(g[1]))
        zp, H = sc.predict_observation(x, nf)
        assert rel_err(zp, mpk[f"{t}_obs_z"]) < 1e-14 and rel_err(H, mpk[f"{t}_obs_H"]) < 1e-14


def test_bad_landmark_id_leaves_the_state_untouched(sc):
    """The reference throws BoundsError (src/ekf.jl:57) before any mutation.  Host ids are refused before any
    launch; device-resident ids are caught by K3 and turn K4 / K5 / K6 into a bit-exact no-op, the error is
    returned by the next slam_sync(), and a latched error keeps later resident updates no-ops until it is read."""
    import torch
    from slamcuda._ffi import lib, dptr
    from slamcuda.slam import _Handle, _mat2
    k = ref.sim_constants()
    R = k["R"]
    rng = np.random.default_rng(8)
    for nf, m in ((40, 3), (900, 20)):          # both K4 kernels (rank <= 32 and wide)
        x, P = _random_state(rng, nf)
        h = _Handle(0, 3 + 2 * nf)
        h.set_state(x, P)
        x0, P0 = h.get_state()
        z = np.zeros((2, m))
        ids = rng.choice(nf, m, replace=False) + 1
        for c, idf in enumerate(ids):
            zp, _ = ref.predict_observation(x, int(idf))
            z[:, c] = zp + rng.normal(0, [0.05, 0.005])
        bad = ids.copy()
        bad[m // 2] = nf + 7
        with pytest.raises(sc.BadIndex):
            h.update(z, R, bad)
        x1, P1 = h.get_state()
        assert np.array_equal(x1, x0) and np.array_equal(P1, P0)
        # resident entry point: ids live on the device
        d_zf = torch.tensor(np.ascontiguousarray(z.T).reshape(-1), device="cuda:0")
        d_bad = torch.tensor(bad, device="cuda:0")
        d_ok = torch.tensor(ids, device="cuda:0")
        r = _mat2(R)
        assert lib.slam_update_resident(h.ptr, d_zf.data_ptr(), m, dptr(r), d_bad.data_ptr()) == 0
        assert lib.slam_update_resident(h.ptr, d_zf.data_ptr(), m, dptr(r), d_ok.data_ptr()) == 0   # still latched: no-op
        with pytest.raises(sc.BadIndex):
            h.sync()
        x1, P1 = h.get_state()
        assert np.array_equal(x1, x0) and np.array_equal(P1, P0)
        # the error is cleared: the same update now runs and matches the oracle
        assert lib.slam_update_resident(h.ptr, d_zf.data_ptr(), m, dptr(r), d_ok.data_ptr()) == 0
        h.sync()
        x2, P2 = h.get_state()
        xo, Po = ref.update(x, P, z, R, ids)
        assert rel_err(x2, xo) < STEP_TOL and rel_err(P2, Po) < STEP_TOL
        h.close()


def test_capacity_growth_matches_the_reference(sc):
    """The reference's state grows without bound (hcat / vcat, src/ekf.jl:108-109).  A resident state created
    with too little capacity grows through slam_reserve (device-to-device copy, new leading dimension) and the
    result equals the oracle's; the pre-gate bound and later updates survive the move."""
    from slamcuda.slam import _Handle
    k = ref.sim_constants()
    R, Q = k["R"], k["Q"]
    rng = np.random.default_rng(9)
    x, P = _random_state(rng, 6)
    st = sc.DeviceEKFSlamState(x, P, n_max=len(x))          # no head-room at all
    zn = np.array([rng.uniform(5, 25, 7), rng.uniform(-1, 1, 7)])
    st.x, st.cov = sc.add_features(st, zn, R)
    xo, Po = ref.add_features(x, P, zn, R)
    assert st.h.n == len(xo) and st.h.n_max >= len(xo)
    xg, Pg = st.h.get_state()
    assert rel_err(xg, xo) < STEP_TOL and rel_err(Pg, Po) < STEP_TOL
    assert st.h.gate_stats()[2] == np.max(np.diag(Pg)[3:])
    zp, _ = ref.predict_observation(xo, 9)
    z = (zp + [0.03, 0.004]).reshape(2, 1)
    zf, idf, zn2 = sc.associate(st, z, R, 4.0, 25.0)
    zf_o, idf_o, zn_o = ref.associate(xo, Po, z, R, 4.0, 25.0)
    assert np.array_equal(idf.reshape(-1), idf_o)
    st.x, st.cov = sc.update(st, zf, R, idf)
    xo, Po = ref.update(xo, Po, zf_o, R, idf_o)
    xg, Pg = st.h.get_state()
    assert rel_err(xg, xo) < STEP_TOL and rel_err(Pg, Po) < STEP_TOL
    st.h.close()


def test_incremental_host_mirror(sc):
    """Host-mirror mode as sim! uses it (`state.x, state.cov = f(state, ...)` on host arrays, sim/ekfslam-sim.jl:100,
    117,120): the device keeps a shadow of what the previous call returned, so a call handed those very arrays uploads
    nothing and predict brings back O(n) doubles.  Results are bit-equal to full upload / download on every call;
    arrays edited in place or replaced are uploaded again."""
    from slamcuda import slam as S
    k = ref.sim_constants()
    Q, R, dt = k["Q"], k["R"], k["dt"]
    rng = np.random.default_rng(12)
    x0, P0 = _random_state(rng, 300)
    n = len(x0)
    full = 8 * n * n + 8 * n
    veh = sc.Vehicle(wheelbase=4.0, measured_gamma=0.04, measured_speed=7.8)
    ids = rng.choice(300, 5, replace=False) + 1

    def obs(x):
        z = np.zeros((2, 5))
        for c, idf in enumerate(ids):
            zp, _ = ref.predict_observation(x, int(idf))
            z[:, c] = zp + rng.normal(0, [0.05, 0.005])
        return z

    z1, z2 = obs(x0), obs(x0)
    zn = np.array([[40.0], [0.3]])
    # A: the incremental mirror, arrays re-bound like sim!
    st = sc.EKFSlamState(x0, P0)
    log = []
    for _ in range(3):
        st.x, st.cov = sc.predict(st, veh, Q, dt); log.append(("predict",) + st.last_transfer)
    zf, idf, _zn = sc.associate(st, z1, R, 4.0, 25.0)
    st.x, st.cov = sc.update(st, zf, R, idf); log.append(("update",) + st.last_transfer)
    st.x, st.cov = sc.add_features(st, np.zeros((2, 0)), R); log.append(("add0",) + st.last_transfer)
    st.x, st.cov = sc.predict(st, veh, Q, dt); log.append(("predict",) + st.last_transfer)
    st.x, st.cov = sc.add_features(st, zn, R); log.append(("add1",) + st.last_transfer)
    st.x, st.cov = sc.predict(st, veh, Q, dt); log.append(("predict",) + st.last_transfer)
    # B: the same calls with the mirror dropped before each one (full upload and download every time)
    sb = sc.EKFSlamState(x0, P0)
    for _ in range(3):
        S.invalidate_mirror(); sb.x, sb.cov = sc.predict(sb, veh, Q, dt)
    S.invalidate_mirror(); zfb, idfb, _ = sc.associate(sb, z1, R, 4.0, 25.0)
    S.invalidate_mirror(); sb.x, sb.cov = sc.update(sb, zfb, R, idfb)
    S.invalidate_mirror(); sb.x, sb.cov = sc.predict(sb, veh, Q, dt)
    S.invalidate_mirror(); sb.x, sb.cov = sc.add_features(sb, zn, R)
    S.invalidate_mirror(); sb.x, sb.cov = sc.predict(sb, veh, Q, dt)
    assert np.array_equal(idf, idfb) and np.array_equal(st.x, sb.x) and np.array_equal(st.cov, sb.cov)
    # ... and equal to the oracle's run of the same calls
    xo, Po = x0.copy(), P0.copy()
    for _ in range(3):
        xo, Po = ref.predict(xo, Po, 0.04, 7.8, 4.0, Q, dt)
    zfo, idfo, _ = ref.associate(xo, Po, z1, R, 4.0, 25.0)
    xo, Po = ref.update(xo, Po, zfo, R, idfo)
    xo, Po = ref.predict(xo, Po, 0.04, 7.8, 4.0, Q, dt)
    xo, Po = ref.add_features(xo, Po, zn, R)
    xo, Po = ref.predict(xo, Po, 0.04, 7.8, 4.0, Q, dt)
    assert state_err(st.x, st.cov, xo, Po) < STEP_TOL
    print(f"\n[host mirror n={n}] (call, bytes up, bytes down): {log}")
    small = 8 * (3 + 6 * n)
    assert log[0] == ("predict", full, small)                 # first call: the state is new to the device
    assert log[1] == ("predict", 0, small) and log[2] == ("predict", 0, small)
    assert log[3] == ("update", 0, full) and log[4] == ("add0", 0, 0) and log[5] == ("predict", 0, small)
    assert log[6][0] == "add1" and log[7][1] == 0             # (add_features moves to a larger handle: one upload)
    # an in-place edit of the pose is seen (digest), a replaced array is seen (identity)
    st.x[0] += 1.0
    st.x, st.cov = sc.predict(st, veh, Q, dt)
    assert st.last_transfer[0] > 0
    st.cov = st.cov.copy()
    st.x, st.cov = sc.predict(st, veh, Q, dt)
    assert st.last_transfer[0] > 0
    st.x, st.cov = sc.predict(st, veh, Q, dt)
    assert st.last_transfer[0] == 0


def test_observation_model_and_association_values(sc):
    k = ref.sim_constants()
    R = k["R"]
    rng = np.random.default_rng(11)
    x, P = _random_state(rng, 6)
    for idf in (1, 3, 6):
        zg, Hg = sc.predict_observation(x, idf)
        zo, Ho = ref.predict_observation(x, idf)
        assert rel_err(zg, zo) < 1e-14 and rel_err(Hg, Ho) < 1e-14
        z = zo + np.array([0.1, -0.01])
        nis_g, nd_g = sc.compute_association(x, P, z, R, idf)
        nis_o, nd_o = ref.compute_association(x, P, z, R, idf)
        assert abs(nis_g - nis_o) < 1e-9 * max(1.0, abs(nis_o))
        assert abs(nd_g - nd_o) < 1e-9 * max(1.0, abs(nd_o))
    with pytest.raises(sc.BadIndex):
        sc.predict_observation(x, 7)


def test_association_outcomes_all_three(sc):
    """One call in which a matched, a dropped (gate1 <= nis, outer <= gate2) and a new
    observation all occur, and two observations pick the same landmark."""
    k = ref.sim_constants()
    R = k["R"]
    rng = np.random.default_rng(3)
    x, P = _random_state(rng, 5)
    P = P * 0.2
    zp2, _ = ref.predict_observation(x, 2)
    zp4, _ = ref.predict_observation(x, 4)
    cand = []
    for dr in np.linspace(0.0, 3.0, 61):
        z = np.array([zp4[0] + dr, zp4[1]])
        nis, _ = ref.compute_association(x, P, z, R, 4)
        cand.append((dr, nis))
    drop = [dr for dr, nis in cand if 6.0 < nis < 20.0]
    assert drop, "no drop-zone candidate"
    z = np.array([[zp2[0] + 0.02, zp2[0] - 0.02, zp4[0] + drop[0], 200.0],
                  [zp2[1], zp2[1] + 0.002, zp4[1], 1.0]])
    zf_o, idf_o, zn_o = ref.associate(x, P, z, R, 4.0, 25.0)
    assert list(idf_o) == [2, 2] and zn_o.shape[1] == 1
    zf, idf, zn = sc.associate(sc.EKFSlamState(x, P), z, R, 4.0, 25.0)
    assert np.array_equal(idf.reshape(-1), idf_o)
    assert np.array_equal(zf, zf_o) and np.array_equal(zn, zn_o)


@pytest.mark.parametrize("nf,m", [(1000, 1), (1000, 8), (1000, 12), (1000, 16), (1000, 40), (300, 3)])
def test_update_large_parity(sc, nf, m):
    """update at n = 2003 (all K6 rank-chunk widths incl. the multi-pass path) vs the numpy
    literal oracle."""
    k = ref.sim_constants()
    R = k["R"]
    rng = np.random.default_rng(nf + m)
    x, P = _random_state(rng, nf)
    ids = rng.choice(nf, size=m, replace=False) + 1
    z = np.zeros((2, m))
    for c, idf in enumerate(ids):
        zp, _ = ref.predict_observation(x, int(idf))
        z[:, c] = zp + rng.normal(0, [0.1, 0.01])
    xo, Po = ref.update(x, P, z, R, ids)
    from slamcuda.slam import _Handle
    h = _Handle(0, len(x))
    h.set_state(x, P)
    h.update(z, R, ids)
    xg, Pg = h.get_state()
    h.close()
    print(f"\n[update n={len(x)} m={m}] x rel {rel_err(xg, xo):.2e}  P rel {rel_err(Pg, Po):.2e}  "
          f"element-wise {state_err(xg, Pg, xo, Po):.2e}")
    assert state_err(xg, Pg, xo, Po) < STEP_TOL
    assert np.array_equal(Pg, Pg.T)


def test_associate_large_parity(sc, twin):
    """Gating over 1000 landmarks x 20 observations vs the twin (the literal oracle is
    O(nz*Nf*n^2)); indices bit-exact."""
    from twin_ffi import Twin
    tw = Twin(twin)
    k = ref.sim_constants()
    R = k["R"]
    rng = np.random.default_rng(77)
    nf = 1000
    x, P = _random_state(rng, nf)
    ids = rng.choice(nf, size=14, replace=False) + 1
    z = np.zeros((2, 20))
    for c, idf in enumerate(ids):
        zp, _ = ref.predict_observation(x, int(idf))
        z[:, c] = zp + rng.normal(0, [0.1, 0.01])
    z[:, 14:] = np.array([rng.uniform(100, 200, 6), rng.uniform(-1, 1, 6)])
    zf_o, idf_o, zn_o = tw.associate(x, P, z, R, 4.0, 25.0)
    zf, idf, zn = sc.associate(sc.EKFSlamState(x, P), z, R, 4.0, 25.0)
    print(f"\n[associate Nf={nf}] matched {len(idf_o)}, new {zn_o.shape[1]}, twin margin {tw.last_margin:.3e}")
    assert np.array_equal(idf.reshape(-1), idf_o)
    assert np.array_equal(zf, zf_o) and np.array_equal(zn, zn_o)
    assert len(idf_o) >= 10


@pytest.mark.parametrize("nf", [0, 1, 35, 700])
def test_landmark_ellipses(sc, nf):
    """compute_landmark_ellipses / ellipse (src/common.jl:236-262) from the 2x2 diagonal blocks on
    the device vs the oracle's eigen-decomposition sqrtm; FP64, 1e-9 relative."""
    rng = np.random.default_rng(5 + nf)
    x, P = _random_state(rng, nf)
    ell_o = ref.compute_landmark_ellipses(x, P)
    ell = sc.compute_landmark_ellipses(x, P)
    assert ell.shape == (2 * nf, 13)
    if nf:
        assert rel_err(ell, ell_o) < STEP_TOL
    veh_o = ref.ellipse(x[0:2], P[0:2, 0:2], 2, 12)
    veh = sc.ellipse(x[0:2], P[0:2, 0:2], 2, 12)
    assert rel_err(veh, veh_o) < STEP_TOL
    # resident state: same numbers without pulling P
    st = sc.DeviceEKFSlamState(x, P, n_max=len(x))
    assert np.array_equal(sc.compute_landmark_ellipses(st), ell)
    st.h.close()


def test_front_end_payloads(sc):
    """The browser monitor's payloads as headless device buffers (sim/browser/wsserver.jl:31-85): the five-parameter
    feature ellipses (eig of the 2 x 2 diagonal blocks) and the lidar line segments (laser_lines,
    src/common.jl:265-283) of a resident state, against the oracle; phi modulo pi (an eigenvector's sign is LAPACK's)."""
    rng = np.random.default_rng(31)
    x, P = _random_state(rng, 300)
    P[3, 3], P[3, 4], P[4, 3], P[4, 4] = 0.3, 0.0, 0.0, 0.1        # a diagonal block, larger variance first
    P[5, 5], P[5, 6], P[6, 5], P[6, 6] = 0.1, 0.0, 0.0, 0.3        # ... and smaller first
    st = sc.DeviceEKFSlamState(x, P, n_max=len(x))
    e = st.h.feature_ellipses()
    eo = ref.feature_ellipses(x, P)
    assert e.shape == eo.shape == (5, 300)
    assert rel_err(e[:4], eo[:4]) < STEP_TOL
    dphi = np.abs(e[4] - eo[4])
    dphi = np.minimum(dphi, np.abs(dphi - math.pi))
    assert np.max(dphi) < 1e-9 and np.all(e[4] > -math.pi / 2 - 1e-15) and np.all(e[4] <= math.pi / 2 + 1e-15)
    assert np.array_equal(sc.feature_ellipses(x, P), e)
    v = st.h.vehicle_ellipse_params()
    lam, u = np.linalg.eigh(P[0:2, 0:2])
    assert np.array_equal(v[:3], x[:3]) and rel_err(v[3:5], np.sqrt(lam)) < STEP_TOL
    dv = abs(v[5] - math.atan2(u[1, 0], u[0, 0]))
    assert min(dv, abs(dv - math.pi)) < 1e-9
    z = np.array([rng.uniform(1, 30, 9), rng.uniform(-1.5, 1.5, 9)])
    lines = st.h.laser_lines(z)
    assert rel_err(lines, ref.laser_lines(z, x[:3])) < 1e-14
    assert np.array_equal(sc.laser_lines(z, x[:3]), lines)
    assert st.h.laser_lines(np.zeros((2, 0))).shape == (4, 0)
    st.h.close()


@pytest.mark.parametrize("nf,m", [(1000, 6), (1000, 8), (1000, 10), (1000, 12), (1000, 13), (1000, 16), (1000, 20),
                                  (1000, 24), (1000, 28), (1000, 32), (1000, 40), (1000, 64), (333, 16), (517, 24),
                                  (130, 6), (61, 30)])
def test_update_rank_sweep_dmma(sc, nf, m):
    """The three downdate kernels against each other and the numpy literal oracle at ranks 12..128: FMA (variant 1,
    up to 32 ranks per pass), tiled DMMA (variant 2, 32-wide chunks), streaming DMMA (variant 3, 12..64 ranks in one
    pass; the only kernel for the 48- / 64-wide single chunk of 33..64 ranks).  On the B200 the four products of an
    m8n8k4 MMA accumulate in k order, so the kernels agree bit for bit (asserted to 1e-12: PTX does not promise the
    order); every one within 1e-9 of the oracle, max-norm and element-wise.  Ragged sizes (n not a multiple of any
    tile) included."""
    from slamcuda._ffi import lib, check
    from slamcuda.slam import _Handle
    k = ref.sim_constants()
    R = k["R"]
    rng = np.random.default_rng(7 * nf + m)
    x, P = _random_state(rng, nf)
    ids = rng.choice(nf, size=m, replace=False) + 1
    z = np.zeros((2, m))
    for c, idf in enumerate(ids):
        zp, _ = ref.predict_observation(x, int(idf))
        z[:, c] = zp + rng.normal(0, [0.1, 0.01])
    xo, Po = ref.update(x, P, z, R, ids)
    rank = 2 * m
    variants = [0]
    if rank <= 32 or rank > 64:
        variants.append(1)
    if 24 < rank <= 32 or rank > 64:
        variants.append(2)
    if 12 <= rank <= 64:
        variants.append(3)
    out = {}
    for variant in variants:
        h = _Handle(0, len(x))
        check(lib.slam_set_downdate_variant(h.ptr, variant))
        h.set_state(x, P)
        h.update(z, R, ids)
        out[variant] = h.get_state()
        h.close()
    diffs = {v: rel_err(out[v][1], out[0][1]) for v in variants[1:]}
    errs = {v: state_err(out[v][0], out[v][1], xo, Po) for v in variants}
    print(f"\n[downdate kernels n={len(x)} R={rank}] P rel diff to the automatic choice {diffs}; vs oracle {errs}")
    for v in variants:
        assert np.array_equal(out[v][0], out[0][0])          # x does not depend on the downdate kernel
        assert np.array_equal(out[v][1], out[v][1].T)
        assert errs[v] < STEP_TOL
    assert all(d < 1e-12 for d in diffs.values())
    assert len(variants) >= 2


def _gather(h, segs):
    """Principal submatrix of the resident P over the index segments [(start, len), ...]."""
    tot = sum(ln for _, ln in segs)
    out = np.empty((tot, tot))
    oi = 0
    for i0, ni in segs:
        oj = 0
        for j0, nj in segs:
            out[oi:oi + ni, oj:oj + nj] = h.cov_block(i0, ni, j0, nj)
            oj += nj
        oi += ni
    return out


@pytest.mark.parametrize("nl", [10000, 35353])
def test_full_size_properties(sc, nl):
    """n = 20003 (the 10k-landmark config, P = 3.2 GB) and n = 70709 (bench.py's N = 1 workload, P = 40 GB);
    P stays on the device.  H is zero outside the
    vehicle and the observed landmarks, so the update restricted to any principal submatrix that
    contains those columns IS the update of that sub-state: the literal oracle runs on a sample
    (vehicle + observed pairs + four 150-row windows) and must agree with the same windows of the
    full-size result.  Plus the size-independent properties: symmetric export, m = 0 identity,
    idempotent association, non-increasing diagonal, PD observed block."""
    from slamcuda._ffi import lib, check, dptr, iptr
    from slamcuda.slam import _Handle
    k = ref.sim_constants()
    R = k["R"]
    m, seed = 8, 3
    n = 3 + 2 * nl
    side = math.sqrt(nl / 0.0035)
    h = _Handle(0, n)
    check(lib.slam_synth_state(h.ptr, nl, side, seed))
    z = np.zeros(2 * (m + 4))
    ids = np.zeros(m, dtype=np.int64)
    check(lib.slam_synth_observations(h.ptr, nl, side, seed, m, 4, dptr(z), iptr(ids)))
    z2 = z.reshape(m + 4, 2).T
    # association: exact ids, idempotent
    zf, idf, zn = h.associate(z2, R, 4.0, 25.0)
    zf2, idf2, zn2 = h.associate(z2, R, 4.0, 25.0)
    # (a noise-free observation has nis = 0 for its own landmark, but the normalised distance adds
    # log det S, so another landmark with a tighter S may legitimately win: most, not all, ids match)
    assert len(idf) == m and sum(int(a) == int(b) for a, b in zip(idf, ids)) >= m - 2
    assert np.array_equal(zf, zf2) and np.array_equal(idf, idf2) and np.array_equal(zn, zn2)
    # the spatial pre-gate (on by default) against the exhaustive scan at full size
    both = _assoc_both(h, z2, R, 4.0, 25.0)
    for a, b in zip(both[1][0], both[0][0]):
        assert np.array_equal(a, b)
    print(f"\n[full size n={n}] pre-gate scored {both[1][2]} of {both[0][2]} landmark passes")
    assert both[0][2] == 2 * nl and both[1][2] < both[0][2] // 50
    # sample: vehicle, observed landmark pairs, four windows
    rng = np.random.default_rng(1)
    seen = sorted({int(i) for i in idf})
    segs = [(0, 3)] + [(3 + 2 * (i - 1), 2) for i in seen]
    taken = {3 + 2 * (i - 1) for i in seen}
    for w0 in (3, 5001, 12345, n - 150):
        w0 = w0 if (w0 - 3) % 2 == 0 else w0 + 1
        while any(w0 - 2 <= t < w0 + 150 for t in taken):    # keep the windows clear of the observed pairs
            w0 = w0 + 200 if w0 + 400 < n else w0 - 200
        segs.append((w0, 150))
    index = np.concatenate([np.arange(s, s + ln) for s, ln in segs])
    x0 = h.get_x()
    P0 = _gather(h, segs)
    assert np.array_equal(P0, P0.T)                         # export mirrors the upper triangle
    # m = 0 is the identity at full size too
    h.update(np.zeros((2, 0)), R, np.zeros(0, dtype=np.int64))
    assert np.array_equal(h.get_x(), x0) and np.array_equal(_gather(h, segs), P0)
    # the oracle on the sub-state
    pos = {i: 1 + c for c, i in enumerate(seen)}               # landmark id inside the sample
    ids_sub = np.array([pos[int(i)] for i in idf])
    xo, Po = ref.update(x0[index], P0, zf, R, ids_sub)
    h.update(zf, R, idf)
    x1 = h.get_x()
    P1 = _gather(h, segs)
    ex, eP, ee = rel_err(x1[index], xo), rel_err(P1, Po), state_err(x1[index], P1, xo, Po)
    print(f"\n[full size n={n} m={m}] sampled {len(index)} states: x rel {ex:.2e}  P rel {eP:.2e}  element-wise {ee:.2e}")
    assert ex < STEP_TOL and eP < STEP_TOL and ee < STEP_TOL
    assert np.array_equal(P1, P1.T)
    d0, d1 = np.diag(P0), np.diag(P1)
    assert np.all(d1 <= d0 * (1 + 1e-12)) and d1.sum() < d0.sum()   # P - W S W' never grows a variance
    nobs = 3 + 2 * len(seen)
    np.linalg.cholesky(P1[:nobs, :nobs])                              # observed block stays PD
    # states outside the sample moved too (the update is dense): spot-check that x changed there
    assert np.any(x1 != x0)
    h.close()


def _assoc_both(h, z, R, g1, g2):
    """associate with the spatial pre-gate on and off on the same resident state -> both results and
    the landmarks (scanned, scored) of each call."""
    out = {}
    for on in (1, 0):
        h.set_pregate(on)
        a0, b0, _ = h.gate_stats()
        res = h.associate(z, R, g1, g2)
        a1, b1, _ = h.gate_stats()
        out[on] = (res, a1 - a0, b1 - b0)
    h.set_pregate(1)
    return out


@pytest.mark.parametrize("nf,nz,side,gates", [
    (5000, 12, 1200.0, (4.0, 25.0)),
    (5000, 64, 1200.0, (4.0, 25.0)),       # eight passes of eight observations
    (1000, 20, 60.0, (4.0, 25.0)),         # dense field: most landmarks inside some annulus
    (3000, 9, 900.0, (25.0, 4.0)),         # gate2 < gate1: the radius follows the larger gate
    (3000, 9, 900.0, (1e6, 1e7)),          # huge gates: radius beyond the field, nothing may be skipped wrongly
    (3000, 9, 900.0, (-1.0, -2.0)),        # no positive gate: pre-gate switches itself off
    (2000, 16, 700.0, (4.0, float("inf"))),
    (2000, 300, 700.0, (4.0, 25.0)),       # 38 passes; the compaction walks two chunks of 256 observations
])
def test_pregate_identical_to_exhaustive(sc, twin, nf, nz, side, gates):
    """SURVEY 8 f4 (the reference's TODO at src/data-association.jl:18-20): the annulus pre-gate may only
    skip pairs whose nis is certainly above both gates, so zf / idf / zn must be bit-identical to the
    exhaustive scan (and to the C twin's sequential scan) -- on observations of mapped landmarks,
    observations placed next to the gate2 boundary of a landmark (drop zone vs new feature) and far
    spurious ones."""
    from slamcuda.slam import _Handle
    from twin_ffi import Twin
    k = ref.sim_constants()
    R = k["R"]
    rng = np.random.default_rng(nf + 13 * nz)
    n = 3 + 2 * nf
    x = np.concatenate([[side / 2, side / 2, 0.3], rng.uniform(0, side, size=2 * nf)])
    L = rng.normal(0, 0.05, size=(n, 8))
    P = L @ L.T + np.diag(rng.uniform(0.01, 0.1, size=n))
    d = np.hypot(x[3::2] - x[0], x[4::2] - x[1])
    near = np.argsort(d)[:max(nz, 8)] + 1
    z = np.zeros((2, nz))
    for c in range(nz):
        idf = int(near[c % len(near)])
        zp, H = ref.predict_observation(x, idf)
        S = H @ P @ H.T + R
        if c % 3 == 0:      # a genuine observation
            z[:, c] = zp + rng.normal(0, [0.1, 0.01])
        elif c % 3 == 1:    # on the gate2 ellipse of that landmark, +-2 %
            u = rng.normal(size=2)
            u /= np.linalg.norm(u)
            z[:, c] = zp + np.linalg.cholesky(S) @ u * math.sqrt(25.0) * rng.choice([0.98, 1.02])
        else:               # spurious: inside the field (may match, drop or be new) or beyond it (always new)
            z[:, c] = [rng.uniform(0.3, 0.6) * side * (1 if c % 2 else 4), rng.uniform(-3, 3)]
    h = _Handle(0, n)
    h.set_state(x, P)
    out = _assoc_both(h, z, R, *gates)
    (zf1, idf1, zn1), scan1, score1 = out[1]
    (zf0, idf0, zn0), scan0, score0 = out[0]
    h.close()
    passes = (nz + 7) // 8
    print(f"\n[pre-gate Nf={nf} nz={nz} gates={gates}] matched {len(idf0)} new {zn0.shape[1]}; scored "
          f"{score1} of {scan1 * passes} landmark passes with the pre-gate, {score0} without")
    assert np.array_equal(idf1, idf0) and np.array_equal(zf1, zf0) and np.array_equal(zn1, zn0)
    assert scan1 == nf and scan0 == nf and score0 == nf * passes
    if gates == (4.0, 25.0) and side > 100:
        if nz <= 64:      # (with hundreds of observations the union of eight annuli per pass covers much of the field)
            assert score1 < score0 // 5, "the pre-gate skipped too little"
        assert len(idf0) >= nz // 3 and 0 < zn0.shape[1]
    if gates[0] < 0:
        assert score1 == score0
    if np.isfinite(gates[1]):
        zf_t, idf_t, zn_t = Twin(twin).associate(x, P, z, R, *gates)
        assert np.array_equal(idf1.reshape(-1), idf_t) and np.array_equal(zf1, zf_t) and np.array_equal(zn1, zn_t)


def test_pregate_bound_follows_the_state(sc):
    """The pre-gate's device-resident bound on the landmark variances: exact at set_state, raised by
    add_features (K7) for far, uncertain new landmarks, never below the diagonal after update; and the
    association stays identical to the exhaustive scan through such a sequence."""
    from slamcuda.slam import _Handle
    k = ref.sim_constants()
    R = k["R"]
    rng = np.random.default_rng(21)
    x, P = _random_state(rng, 400)
    h = _Handle(0, 3 + 2 * 440)
    h.set_state(x, P)
    assert h.gate_stats()[2] == np.max(np.diag(P)[3:])
    zn = np.array([rng.uniform(300, 900, 30), rng.uniform(-3, 3, 30)])     # variance ~ (r * 1 deg)^2, up to ~250
    h.add_features(zn, R)
    x1, P1 = h.get_state()
    s2 = h.gate_stats()[2]
    assert s2 == np.max(np.diag(P1)[3:]) and s2 > 10 * np.max(np.diag(P)[3:])
    # observe some of the far landmarks again (noisy) plus near ones and spurious ones
    ids = np.concatenate([401 + rng.choice(30, 6, replace=False), rng.choice(400, 4, replace=False) + 1])
    z = np.zeros((2, 14))
    for c, idf in enumerate(ids):
        zp, _ = ref.predict_observation(x1, int(idf))
        z[:, c] = zp + rng.normal(0, [0.1, 0.01])
    z[:, 10:] = np.array([rng.uniform(100, 1500, 4), rng.uniform(-3, 3, 4)])
    out = _assoc_both(h, z, R, 4.0, 25.0)
    for a, b in zip(out[1][0], out[0][0]):
        assert np.array_equal(a, b)
    zf, idf, _ = out[1][0]
    assert len(idf) >= 8
    h.update(zf, R, idf)
    _, P2 = h.get_state()
    assert h.gate_stats()[2] == s2 >= np.max(np.diag(P2)[3:])           # update only shrinks variances
    out = _assoc_both(h, z, R, 4.0, 25.0)
    for a, b in zip(out[1][0], out[0][0]):
        assert np.array_equal(a, b)
    # NaN in the state: the bound turns NaN, the annulus test is false everywhere, nothing is skipped
    P2[7, 7] = float("nan")
    h.set_state(x1, P2)
    assert math.isnan(h.gate_stats()[2])
    out = _assoc_both(h, z, R, 4.0, 25.0)
    for a, b in zip(out[1][0], out[0][0]):
        assert np.array_equal(a, b)
    assert out[1][2] == out[0][2]
    h.close()


def test_closed_loop_large_course(sc, twin, course1):
    """sim! closed loop on a course six times the size of course1 (600 m, 725 landmarks in the field, the
    map grows to ~320 landmarks over 886 observation steps), state resident on the device, against the C
    twin's run of the same inputs: every association (pre-gated scan), update and in-place append of a growing
    map has to agree for the tracks to agree.  End-of-run tolerance 1e-6 as north_star states."""
    import ctypes as C
    from slamcuda import sim as gsim
    from twin_ffi import d
    scale, extra, ms = 6, 40, 12000
    wp = course1 * scale
    rng = np.random.default_rng(11)
    lo, hi = wp.min(axis=1) - 20, wp.max(axis=1) + 20
    nl = int(0.0035 * np.prod(hi - lo))
    lm = np.vstack([rng.uniform(lo[0], hi[0], nl), rng.uniform(lo[1], hi[1], nl)])
    ctrl = rng.standard_normal((2, ms))
    obs = rng.standard_normal(200000)
    ld = 3 + 2 * (nl + extra)
    tt, st, xo, Po = np.zeros(3 * ms), np.zeros(3 * ms), np.zeros(ld), np.zeros(ld * ld)
    counts = (C.c_int * 4)()
    mg = C.c_double()
    twin.twin_sim_cap.restype = C.c_int
    rc = twin.twin_sim_cap(d(np.ascontiguousarray(wp.T).reshape(-1)), wp.shape[1],
                           d(np.ascontiguousarray(lm.T).reshape(-1)), nl, d(np.ascontiguousarray(ctrl.T).reshape(-1)),
                           d(obs), len(obs), C.c_uint64(0), C.c_uint32(0), 1, ms, d(tt), d(st), d(xo), d(Po), counts,
                           C.byref(mg), extra)
    assert rc == 0
    ns, nupd, nfin = counts[0], counts[1], counts[2]
    out = gsim.run(wp, lm, ctrl, obs, nlaps=1, n_max=ld, max_steps=ms)
    assert out["nsteps"] == ns and out["nupdates"] == nupd and len(out["x"]) == nfin
    trk = st[:3 * ns].reshape(ns, 3).T
    e_track = float(np.max(np.abs(out["slam_track"] - trk))) / float(np.max(np.abs(trk)))
    Pt = Po.reshape(ld, ld).T[:nfin, :nfin]
    print(f"\n[closed loop x{scale}] {ns} steps, {nupd} observation steps, n = {nfin}, twin margin {mg.value:.2e}; "
          f"track rel {e_track:.2e}, x rel {rel_err(out['x'], xo[:nfin]):.2e}, P rel {rel_err(out['P'], Pt):.2e}, "
          f"{out['launches']} launches")
    assert e_track < TRACK_TOL and rel_err(out["x"], xo[:nfin]) < TRACK_TOL and rel_err(out["P"], Pt) < TRACK_TOL
    assert nfin > 3 + 2 * 250 and mg.value > 1e-9


@pytest.mark.parametrize("nf,m,spare", [(333, 1, 5), (333, 4, 8), (333, 7, 21), (333, 10, 1), (517, 16, 37), (517, 24, 3),
                                        (130, 30, 77), (200, 36, 10), (61, 30, 2)])
def test_padding_guard_no_stray_writes(sc, nf, m, spare):
    """Own memory check (compute-sanitizer is closed on this pool).  The slab has capacity n + spare (ragged, so tiles
    of every kernel straddle the edge of the live state); slam_debug_guard paints the slab and x outside [0, n) with a
    NaN payload, then predict, associate (pre-gate on and off), update with every downdate kernel the rank admits, the
    ellipse / block read-outs and add_features run.  No painted element may change (add_features moves n and owns what
    it adds), and the results still match the oracle to STEP_TOL: a kernel that READ the padding into a result would
    carry the NaN into it."""
    from slamcuda._ffi import lib, check
    from slamcuda.slam import _Handle
    k = ref.sim_constants()
    R, Q = k["R"], k["Q"]
    rng = np.random.default_rng(1000 * nf + 10 * m + spare)
    x, P = _random_state(rng, nf)
    n = len(x)
    ids = rng.choice(nf, size=m, replace=False) + 1
    z = np.zeros((2, m))
    for c, idf in enumerate(ids):
        zp, _ = ref.predict_observation(x, int(idf))
        z[:, c] = zp + rng.normal(0, [0.1, 0.01])
    rank = 2 * m
    variants = [0] + ([1] if rank <= 32 or rank > 64 else []) + ([2] if 24 < rank <= 32 or rank > 64 else []) \
        + ([3] if 12 <= rank <= 64 else [])
    for variant in variants:
        h = _Handle(0, n + spare, grow=False)
        check(lib.slam_set_downdate_variant(h.ptr, variant))
        h.set_state(x, P)
        h.guard_arm()
        assert h.guard_changed() == 0
        h.predict(0.02, 3.0, 4.0, Q, 0.025)
        xr, Pr = ref.predict(x.copy(), P.copy(), 0.02, 3.0, 4.0, Q, 0.025)     # the oracle's predict works in place
        for pg in (True, False):
            h.set_pregate(pg)
            h.associate(z, R, 4.0, 25.0)
        assert h.guard_changed() == 0, "predict / associate wrote outside the live state"
        h.update(z, R, ids)
        xr, Pr = ref.update(xr, Pr, z, R, ids)
        h.cov_block(0, n, max(n - 5, 0), min(5, n))
        h.landmark_ellipses()
        h.feature_ellipses()
        changed = h.guard_changed()
        assert changed == 0, f"update (variant {variant}, rank {rank}) overwrote {changed} padding elements"
        xg, Pg = h.get_state()
        assert np.isfinite(xg).all() and np.isfinite(Pg).all()
        assert state_err(xg, Pg, xr, Pr) < STEP_TOL
        knew = min(spare // 2, 3)
        if knew:
            zn = np.array([rng.uniform(5, 25, knew), rng.uniform(-1, 1, knew)])
            h.add_features(zn, R)
            xr, Pr = ref.add_features(xr, Pr, zn, R)
            assert h.n == n + 2 * knew
            assert h.guard_changed() == 0, "add_features wrote beyond the rows and columns it adds"
            h.update(z, R, ids)                                  # the grown state through the same kernels
            xr, Pr = ref.update(xr, Pr, z, R, ids)
            assert h.guard_changed() == 0
            xg, Pg = h.get_state()
            assert np.isfinite(Pg).all() and state_err(xg, Pg, xr, Pr) < STEP_TOL
        h.close()


def test_padding_guard_detects_a_write(sc):
    """Positive control for slam_debug_guard: after arming at n, a state of n - 2 set into the same handle leaves the
    band that used to be live (never painted) outside the new live state, so the guard must count exactly
    n^2 - (n - 2)^2 elements of the slab + 2 of x; re-arming clears it."""
    from slamcuda.slam import _Handle
    rng = np.random.default_rng(3)
    x, P = _random_state(rng, 20)
    n = len(x)
    h = _Handle(0, n + 9, grow=False)
    h.set_state(x, P)
    h.guard_arm()
    assert h.guard_changed() == 0
    h.set_state(x[:n - 2], np.ascontiguousarray(P[:n - 2, :n - 2]))
    assert h.guard_changed() == n * n - (n - 2) * (n - 2) + 2
    h.guard_arm()
    assert h.guard_changed() == 0
    h.close()


def test_downdate_kernel_choice_is_reported(sc):
    """slam_downdate_kernel names the kernel the dispatcher ran for the last update (bench.py labels its roofline with
    it): FMA up to 4 ranks, streaming DMMA for 8..64, tiled DMMA above; a forced variant is reported as forced."""
    from slamcuda._ffi import lib, check
    from slamcuda.slam import _Handle
    R = ref.sim_constants()["R"]
    rng = np.random.default_rng(11)
    x, P = _random_state(rng, 120)
    h = _Handle(0, len(x))
    assert lib.slam_downdate_kernel(h.ptr) == 0
    for m, want in [(1, 1), (2, 1), (4, 3), (8, 3), (16, 3), (32, 3), (40, 2)]:
        h.set_state(x, P)
        ids = rng.choice(120, size=m, replace=False) + 1
        z = np.stack([ref.predict_observation(x, int(i))[0] for i in ids], axis=1)
        h.update(z, R, ids)
        assert lib.slam_downdate_kernel(h.ptr) == want, (m, lib.slam_downdate_kernel(h.ptr))
    check(lib.slam_set_downdate_variant(h.ptr, 1))
    h.set_state(x, P)
    h.update(z[:, :8], R, ids[:8])
    assert lib.slam_downdate_kernel(h.ptr) == 1
    h.close()


@pytest.mark.parametrize("nf,m", [(60, 1), (60, 2), (60, 4), (130, 6), (200, 8), (200, 10), (200, 12), (300, 13),
                                  (300, 16)])
def test_update_P_bitwise_vs_twin(sc, twin, nf, m):
    """Pins the ARITHMETIC of K3..K6, not just its accuracy.  oracle/slam_twin.c restates the update with the CUDA
    path's operation order: products and sums rounded separately (the twin is built with -ffp-contract=off, filter.cu
    with -fmad=false), the innovation Cholesky per entry in k-ascending order, inv(U) in dtrti2's order, W1 with l
    ascending, the downdate one k-ascending fma chain per element.  H is built from IEEE sqrt and divisions only, and P
    depends on the observation through H alone, so P must come out BIT-EQUAL to the twin's at every rank K4 handles
    (R = 2..32; FMA and streaming-DMMA downdate kernels).  x also takes atan2, which CUDA's libm and glibc round
    differently, so x is held to 1e-12.  (An explicit fma() in K4 passes every tolerance test and fails this one.)
    Scope: R <= 32 only.  Above that the update takes k_innov_chol_wide, the tiled gain and (past 64 ranks) the chunked
    downdate, whose sums run in another order: measured against the twin at R = 34..128, nearly every element of P
    differs in the last place (max-norm relative 4e-16 .. 5e-15), which test_update_rank_sweep_dmma holds to 1e-9."""
    from twin_ffi import Twin
    from slamcuda.slam import _Handle
    R = ref.sim_constants()["R"]
    rng = np.random.default_rng(100 * nf + m)
    x, P = _random_state(rng, nf)
    P = (P + P.T) * 0.5                                         # exactly symmetric: the twin updates both triangles
    ids = rng.choice(nf, size=m, replace=False) + 1
    z = np.zeros((2, m))
    for c, idf in enumerate(ids):
        zp, _ = ref.predict_observation(x, int(idf))
        z[:, c] = zp + rng.normal(0, [0.1, 0.01])
    h = _Handle(0, len(x))
    h.set_state(x, P)
    xg, Pg = x, P
    for step in range(2):
        # both sides start from the same bits: the twin takes the state the GPU exported (after the first update
        # the two x differ in the last place through atan2, and H is built from x)
        xt, Pt = Twin(twin).update(xg, Pg, z, R, ids)
        h.update(z, R, ids)
        xg, Pg = h.get_state()
        ndiff = int((Pg != Pt).sum())
        print(f"\n[update vs twin nf={nf} R={2 * m} step {step}] P elements differing {ndiff} of {Pg.size}, "
              f"max rel {rel_err(Pg, Pt):.1e}; x rel {rel_err(xg, xt):.1e}")
        assert ndiff == 0
        assert rel_err(xg, xt) < 1e-12
    h.close()


def test_update_early_return_is_unobservable(sc, monkeypatch):
    """Host-buffer slam_update returns once the status and x are final, while the downdate of P is still running
    (n = 10003: K6 is ~0.13 ms, the call returns ~30 us after its launches).  Nothing a caller can observe may differ
    from waiting for everything: a handle created with SLAM_UPDATE_EARLY=0 (full wait) and a default one go through the
    same calls -- back-to-back updates with different observations (the pinned staging buffer and d_upd are reused while
    the previous K6 runs), the pose, a block read-out, predict, add_features, the not-PD path -- and every result is
    bit-equal."""
    from slamcuda._ffi import lib, check, dptr, iptr
    from slamcuda.slam import _Handle
    k = ref.sim_constants()
    R, Q = k["R"], k["Q"]
    nl, m = 5000, 8
    side = math.sqrt(nl / 0.0035)

    def drive(early):
        if early:
            monkeypatch.delenv("SLAM_UPDATE_EARLY", raising=False)
        else:
            monkeypatch.setenv("SLAM_UPDATE_EARLY", "0")
        h = _Handle(0, 3 + 2 * nl + 8, grow=False)               # the switch is read when the handle is created
        monkeypatch.delenv("SLAM_UPDATE_EARLY", raising=False)
        check(lib.slam_synth_state(h.ptr, nl, side, 3))
        obs = []
        for seed in (3, 4, 5):
            z = np.zeros(2 * (m + 2))
            ids = np.zeros(m, dtype=np.int64)
            check(lib.slam_synth_observations(h.ptr, nl, side, seed, m, 2, dptr(z), iptr(ids)))
            obs.append((z.reshape(m + 2, 2).T[:, :m].copy(), ids.copy(), z.reshape(m + 2, 2).T[:, m:].copy()))
        out = []
        for rnd in range(2):
            for zf, ids, _ in obs:                              # three updates with no read-out in between
                h.update(zf, R, ids)
                out.append(h.get_pose())
            out.append(h.cov_block(0, 40, 3 + 2 * (int(obs[0][1][0]) - 1), 2))   # ordered behind the running K6
            h.predict(0.01, 6.0, 4.0, Q, 0.025)
            out.append(h.get_pose())
        h.add_features(obs[0][2], R)
        h.update(obs[1][0], R, obs[1][1])
        x, P = h.get_state()
        out += [x, P]
        h.close()
        return out

    a, b = drive(True), drive(False)
    assert len(a) == len(b)
    for i, (u, v) in enumerate(zip(a, b)):
        assert np.array_equal(u, v), f"result {i} differs between early return and full wait"
    assert np.isfinite(a[-1]).all()
    # the error path through the early check: S not positive definite is reported by the call itself, state untouched
    x0 = np.array([0.0, 0.0, 0.0, 10.0, 5.0])
    P0 = -np.eye(5)
    h = _Handle(0, 5)
    h.set_state(x0, P0)
    with pytest.raises(sc.NotPositiveDefinite):
        h.update(np.array([[11.0], [0.4]]), R, np.array([1]))
    xe, Pe = h.get_state()
    assert np.array_equal(xe, x0) and np.array_equal(Pe, P0)
    h.update(np.zeros((2, 0)), R, np.zeros(0, dtype=np.int64))  # and the handle is usable afterwards
    h.close()
