"""Host-side mirror of the reference's `module SLAM` filter surface (src/SLAM.jl:5-30) over
the C ABI of libslamcuda.so.  Same names, positional arguments and return conventions as
the Julia functions so the parity tests read like calls into the reference:

    state.x, state.cov = predict(state, vehicle, Q, dt)          # src/ekf.jl:8
    zf, idf, zn        = associate(state, z, R, gate1, gate2)    # src/data-association.jl:1
    state.x, state.cov = update(state, zf, R, idf)               # src/ekf.jl:46
    state.x, state.cov = add_features(state, zn, R)              # src/ekf.jl:84

Two state types:
  * EKFSlamState        -- host-mirror mode: x / cov are numpy arrays; every call uploads
                           them, runs the CUDA kernels and downloads the result.
  * DeviceEKFSlamState  -- resident mode: x / cov stay on the device behind a handle;
                           calls return the state object's lazy views.
All arithmetic of the filter runs in the CUDA library; there is no CPU fallback.
"""
from __future__ import annotations

import ctypes as C
import math
from dataclasses import dataclass, field

import numpy as np

from . import _ffi
from ._ffi import lib, check, dptr, iptr

PI = math.pi


def mpi_to_pi(phi: float) -> float:
    """src/common.jl:102-110 (scalar host helper; the device has its own copy)."""
    if phi > PI:
        return phi - 2 * PI
    if phi < -PI:
        return phi + 2 * PI
    return phi


@dataclass
class Vehicle:
    """The fields of src/common.jl:36-55 that `predict` reads (src/ekf.jl:14-16)."""
    wheelbase: float = 4.0
    measured_gamma: float = 0.0
    measured_speed: float = 0.0
    pose: np.ndarray = field(default_factory=lambda: np.zeros(3))


def _f64(a, shape=None):
    a = np.ascontiguousarray(np.asarray(a, dtype=np.float64))
    return a if shape is None else a.reshape(shape)


def _cols(z):
    """2 x k Julia matrix -> flat column-major buffer and k."""
    z = np.asarray(z, dtype=np.float64)
    if z.size == 0:
        return np.zeros(0), 0
    z = z.reshape(2, -1)
    return np.ascontiguousarray(z.T).reshape(-1), z.shape[1]


def _mat2(M):
    """2 x 2 matrix -> column-major 4-vector."""
    return np.ascontiguousarray(np.asarray(M, dtype=np.float64).reshape(2, 2).T).reshape(-1)


class _Handle:
    """Owns a slam_t*.  The reference's state grows without bound (hcat / vcat, src/ekf.jl:108-109);
    with grow=True (the default) add_features doubles the device capacity through slam_reserve when
    the append would not fit, so a spurious landmark never aborts a resident run.  grow=False keeps
    the capacity fixed (CapacityError), e.g. for maps sized to the GPU's memory."""

    def __init__(self, device: int = 0, n_max: int = 256, grow: bool = True):
        self.device = device
        self.grow = grow
        self.mirror = None      # host arrays this handle's device state is known to equal (incremental host mirror)
        self.ptr = C.c_void_p()
        check(lib.slam_create(device, n_max, C.byref(self.ptr)))

    @property
    def n_max(self) -> int:
        return check(lib.slam_capacity(self.ptr))

    def reserve(self, n_max: int):
        check(lib.slam_reserve(self.ptr, int(n_max)))

    def close(self):
        if self.ptr:
            lib.slam_destroy(self.ptr)
            self.ptr = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def n(self) -> int:
        return check(lib.slam_n(self.ptr))

    def set_state(self, x, P):
        x = _f64(x)
        n = x.shape[0]
        Pf = np.asfortranarray(np.asarray(P, dtype=np.float64))
        assert Pf.shape == (n, n)
        self.mirror = None
        check(lib.slam_set_state(self.ptr, dptr(x), dptr(Pf), n, n))

    def get_state(self, order: str = "C"):
        """(x, P) as fresh host arrays.  order="F" hands P back in the library's own column-major layout (no host-side
        re-layout: what the incremental host mirror uses for large maps); "C" (default) as a C-contiguous array."""
        n = self.n
        x = np.empty(n)
        P = np.empty((n, n), order="F")
        check(lib.slam_get_state(self.ptr, dptr(x), dptr(P), n))
        return x, (P if order == "F" else np.ascontiguousarray(P))

    def pose_rows(self):
        """What predict changes (src/ekf.jl:32-41): x[0:3], rows 0..2 and columns 0..2 of P -- O(n) doubles instead of n^2."""
        n = self.n
        return self.get_pose(), self.cov_block(0, 3, 0, n), self.cov_block(0, n, 0, 3)

    def get_x(self):
        x = np.empty(self.n)
        check(lib.slam_get_state(self.ptr, dptr(x), None, 0))
        return x

    def get_pose(self):
        p = np.empty(3)
        check(lib.slam_get_pose(self.ptr, dptr(p)))
        return p

    def cov_block(self, i0, ni, j0, nj):
        out = np.empty((ni, nj), order="F")
        check(lib.slam_get_cov_block(self.ptr, i0, ni, j0, nj, dptr(out), max(ni, 1)))
        return np.ascontiguousarray(out)

    def guard_arm(self):
        """Paint the device slab and x outside the live state (slam_debug_guard): see guard_changed."""
        check(lib.slam_debug_guard(self.ptr, 1, None))

    def guard_changed(self) -> int:
        """How many painted padding elements have been overwritten since guard_arm()."""
        c = C.c_longlong(-1)
        check(lib.slam_debug_guard(self.ptr, 0, C.byref(c)))
        return int(c.value)

    # -- raw operations ------------------------------------------------------------------
    def predict(self, g, v, w, Q, dt):
        q = _mat2(Q)
        check(lib.slam_predict(self.ptr, float(g), float(v), float(w), dptr(q), float(dt)))

    def associate(self, z, R, gate1, gate2):
        zc, nz = _cols(z)
        r = _mat2(R)
        zf = np.zeros(2 * max(nz, 1))
        zn = np.zeros(2 * max(nz, 1))
        idf = np.zeros(max(nz, 1), dtype=np.int64)
        m, k = C.c_int(0), C.c_int(0)
        check(lib.slam_associate(self.ptr, dptr(zc), nz, dptr(r), float(gate1), float(gate2), dptr(zf),
                                 iptr(idf), C.byref(m), dptr(zn), C.byref(k)))
        m, k = m.value, k.value
        return (zf[:2 * m].reshape(m, 2).T.copy(), idf[:m].copy(), zn[:2 * k].reshape(k, 2).T.copy())

    def set_pregate(self, on):
        """Spatial pre-gate of associate; on by default, results identical either way."""
        check(lib.slam_set_pregate(self.ptr, int(bool(on))))

    def gate_stats(self):
        """(landmarks scanned, landmarks scored, bound on the landmark variances), cumulative."""
        a, b, s2 = C.c_longlong(0), C.c_longlong(0), C.c_double(0.0)
        check(lib.slam_gate_stats(self.ptr, C.byref(a), C.byref(b), C.byref(s2)))
        return a.value, b.value, s2.value

    def update(self, zf, R, idf):
        zc, m = _cols(zf)
        ids = np.ascontiguousarray(np.asarray(idf, dtype=np.int64).reshape(-1))
        assert ids.shape[0] == m, "idf must have one id per column of z"
        r = _mat2(R)
        check(lib.slam_update(self.ptr, dptr(zc), m, dptr(r), iptr(ids)))

    def add_features(self, zn, R):
        zc, k = _cols(zn)
        r = _mat2(R)
        need = self.n + 2 * k
        if self.grow and need > self.n_max:
            self.reserve(max(need, 2 * self.n_max))
        check(lib.slam_add_features(self.ptr, dptr(zc), k, dptr(r)))

    def compute_association(self, z, R, idf):
        zc = _f64(z).reshape(-1)
        r = _mat2(R)
        nis, nd = C.c_double(), C.c_double()
        check(lib.slam_compute_association(self.ptr, dptr(zc), dptr(r), int(idf), C.byref(nis), C.byref(nd)))
        return nis.value, nd.value

    def predict_observation(self, idf):
        n = self.n
        zp = np.empty(2)
        H = np.empty((2, n), order="F")
        check(lib.slam_predict_observation(self.ptr, int(idf), dptr(zp), dptr(H)))
        return zp, np.ascontiguousarray(H)

    def landmark_ellipses(self, nsigma=2.0, nsegs=12):
        nf = (self.n - 3) // 2
        out = np.zeros((2 * nf, nsegs + 1), order="F")
        check(lib.slam_compute_landmark_ellipses(self.ptr, float(nsigma), int(nsegs), dptr(out), max(2 * nf, 1)))
        return np.ascontiguousarray(out)

    def vehicle_ellipse(self, nsigma=2.0, nsegs=12):
        out = np.zeros((2, nsegs + 1), order="F")
        check(lib.slam_vehicle_ellipse(self.ptr, float(nsigma), int(nsegs), dptr(out)))
        return np.ascontiguousarray(out)

    def feature_ellipses(self):
        """5 x Nf [cx; cy; rx; ry; phi] of sim/browser/wsserver.jl:72-85, from the 2 x 2 diagonal blocks on the device."""
        nf = (self.n - 3) // 2
        out = np.zeros((5, nf), order="F")
        check(lib.slam_feature_ellipses(self.ptr, dptr(out)))
        return np.ascontiguousarray(out)

    def vehicle_ellipse_params(self):
        """[cx, cy, vehicle_phi, rx, ry, phi] of sim/browser/wsserver.jl:58-63."""
        out = np.zeros(6)
        check(lib.slam_vehicle_ellipse_params(self.ptr, dptr(out)))
        return out

    def laser_lines(self, z):
        """laser_lines(z, state.x[1:3]) -- src/common.jl:265-283 -> 4 x nz [vx; vy; fx; fy]."""
        zc, nz = _cols(z)
        out = np.zeros((4, max(nz, 1)), order="F")
        check(lib.slam_laser_lines(self.ptr, dptr(zc), nz, dptr(out)))
        return np.ascontiguousarray(out[:, :nz])

    def sync(self):
        check(lib.slam_sync(self.ptr))

    def launch_count(self) -> int:
        return int(lib.slam_launch_count(self.ptr))


_SCRATCH: dict[tuple[int, int], _Handle] = {}


def _scratch_handle(n_needed: int, device: int = 0) -> _Handle:
    """A cached handle with capacity >= n_needed for host-mirror calls."""
    cap = 128
    while cap < n_needed:
        cap *= 2
    key = (device, cap)
    h = _SCRATCH.get(key)
    if h is None:
        h = _Handle(device, cap)
        _SCRATCH[key] = h
    return h


def _fingerprint(x, P):
    """Cheap digest of a host state: all of x, the 3 x 3 vehicle block of P and up to 512 evenly spaced diagonal
    entries (a strided gather: it must stay far below the O(n) transfer a predict makes).  The incremental host mirror
    skips the upload when the arrays handed in ARE the ones the previous call returned (object identity) and this digest
    still matches what was recorded then -- which catches the usual ways a caller re-seeds a state in place (new pose
    or landmarks, reset vehicle covariance), though not an arbitrary in-place edit of P: call invalidate_mirror()
    after one."""
    n = P.shape[0]
    step = max(1, n // 512)
    d = np.ascontiguousarray(np.diagonal(P)[::step])
    return hash((x.tobytes(), np.ascontiguousarray(P[:3, :3]).tobytes(), d.tobytes()))


def invalidate_mirror():
    """Forget every device shadow: the next host-mirror call uploads its state in full."""
    for h in _SCRATCH.values():
        h.mirror = None


class SlamState:
    """abstract SlamState (src/common.jl:22)."""


class EKFSlamState(SlamState):
    """EKFSlamState{T}(x, cov) -- src/common.jl:25-28.  Host-mirror mode: x / cov are host arrays, as sim! builds and
    re-binds them (sim/ekfslam-sim.jl:42,100,117,120).

    Incremental mirror.  The device keeps a shadow of the state the previous call returned; when the next call is
    handed those very arrays (`state.x, state.cov = f(state, ...)` re-binds them), nothing is uploaded, and only what
    the call changed comes back: predict x[0:3] and rows / columns 0..2 of P (O(n)), update all of P, add_features
    with k = 0 nothing.  Any other arrays (a re-seeded state) are uploaded in full, as before."""

    def __init__(self, x, cov, device: int = 0):
        self.x = np.array(x, dtype=np.float64).reshape(-1)
        self.cov = np.array(cov, dtype=np.float64)
        self.device = device
        self.last_transfer = (0, 0)     # (bytes uploaded, bytes downloaded) by the last call

    def _upload(self, extra: int = 0) -> _Handle:
        n = len(self.x)
        h = _scratch_handle(n + extra, self.device)
        m = getattr(h, "mirror", None)
        if (m is not None and m[0] is self.x and m[1] is self.cov and h.n == n and self.cov.shape == (n, n)
                and m[2] == _fingerprint(self.x, self.cov)):
            self._up = 0
            return h
        h.mirror = None
        h.set_state(self.x, self.cov)
        self._up = 8 * n * n + 8 * n
        return h

    def _adopt(self, h: _Handle, x, P, down: int):
        """The device now mirrors the host arrays (x, P)."""
        h.mirror = (x, P, _fingerprint(x, P))
        self.last_transfer = (getattr(self, "_up", 0), down)


class DeviceEKFSlamState(SlamState):
    """Resident mode: the state lives on the device; `.x` / `.cov` fetch on access."""

    def __init__(self, x, cov, n_max: int, device: int = 0):
        self.h = _Handle(device, n_max)
        self.h.set_state(np.asarray(x, dtype=np.float64), np.asarray(cov, dtype=np.float64))

    @property
    def x(self):
        return self.h.get_x()

    @x.setter
    def x(self, value):   # `state.x, state.cov = f(state, ...)` re-binds; the data is already in place
        pass

    @property
    def cov(self):
        return self.h.get_state()[1]

    @cov.setter
    def cov(self, value):
        pass

    def _upload(self, extra: int = 0) -> _Handle:
        return self.h


def _result(state, h: _Handle):
    """update / add_features return fresh arrays (src/ekf.jl:76,121); resident states return themselves."""
    if isinstance(state, DeviceEKFSlamState):
        return state, state       # placeholders accepted by the setters above
    n = h.n
    x, P = h.get_state(order="F" if n > 512 else "C")
    state._adopt(h, x, P, 8 * n * n + 8 * n)
    return x, P


def predict(state, vehicle, Q, dt):
    """src/ekf.jl:8-43.  Mutates the state in place (as the reference does) and returns (x, P)."""
    h = state._upload()
    h.predict(vehicle.measured_gamma, vehicle.measured_speed, vehicle.wheelbase, Q, dt)
    if isinstance(state, DeviceEKFSlamState):
        return state, state
    n = h.n
    pose, rows, cols = h.pose_rows()          # all predict touches: O(n), not n^2
    state.x[0:3] = pose
    state.cov[0:3, :] = rows
    state.cov[:, 0:3] = cols
    state._adopt(h, state.x, state.cov, 8 * (3 + 6 * n))
    return state.x, state.cov


def associate(state, z, R, gate1, gate2):
    """src/data-association.jl:1-51 -> (zf 2 x m, idf 1 x m Int (1-based), zn 2 x k)."""
    h = state._upload()
    zf, idf, zn = h.associate(z, R, gate1, gate2)
    if isinstance(state, EKFSlamState):
        state._adopt(h, state.x, state.cov, 0)
    return zf, idf.reshape(1, -1), zn


def update(state, z, R, idf):
    """src/ekf.jl:46-77 -> new (x, P); with no associated observation (the first observation step of every run) the
    state as it is."""
    if np.size(z) == 0 and isinstance(state, EKFSlamState):
        state.last_transfer = (0, 0)
        return state.x, state.cov
    h = state._upload()
    h.update(z, R, idf)
    return _result(state, h)


def add_features(state, z, R):
    """src/ekf.jl:84-122 -> augmented (x, P); with no new observation the inputs themselves (:91 never runs)."""
    k = 0 if np.size(z) == 0 else np.asarray(z).reshape(2, -1).shape[1]
    if k == 0 and isinstance(state, EKFSlamState):
        state.last_transfer = (0, 0)
        return state.x, state.cov
    h = state._upload(extra=2 * k)
    h.add_features(z, R)
    return _result(state, h)


def predict_observation(x, idf: int):
    """src/common.jl:139-165 -> (z [range; bearing], H 2 x n)."""
    x = np.asarray(x, dtype=np.float64).reshape(-1)
    h = _scratch_handle(len(x))
    h.set_state(x, np.zeros((len(x), len(x))))
    return h.predict_observation(idf)


def compute_association(x, P, z, R, idf: int):
    """src/data-association.jl:53-63 -> (nis, nd)."""
    x = np.asarray(x, dtype=np.float64).reshape(-1)
    h = _scratch_handle(len(x))
    h.set_state(x, P)
    return h.compute_association(z, R, idf)


def compute_landmark_ellipses(x, P=None):
    """src/common.jl:247-262 -> 2nf x 13 contour points (nsigma = 2, nsegs = 12).  Accepts
    (x, P) arrays like the reference, or a state object (resident states read only the 2 x 2
    diagonal blocks on the device)."""
    if isinstance(x, SlamState):
        return x._upload().landmark_ellipses()
    x = np.asarray(x, dtype=np.float64).reshape(-1)
    h = _scratch_handle(len(x))
    h.set_state(x, P)
    return h.landmark_ellipses()


def feature_ellipses(x, P=None):
    """sim/browser/wsserver.jl:72-85 -> 5 x nf [cx; cy; sqrt(l1); sqrt(l2); phi]; accepts (x, P) or a state object."""
    if isinstance(x, SlamState):
        return x._upload().feature_ellipses()
    x = np.asarray(x, dtype=np.float64).reshape(-1)
    h = _scratch_handle(len(x))
    h.set_state(x, P)
    return h.feature_ellipses()


def laser_lines(z, vehicle_pose):
    """src/common.jl:265-283 -> 4 x nz line segments vehicle -> observed point."""
    xs = np.zeros(3)
    xs[:] = np.asarray(vehicle_pose, dtype=np.float64).reshape(-1)[:3]
    h = _scratch_handle(3)
    h.set_state(xs, np.zeros((3, 3)))
    return h.laser_lines(z)


def ellipse(x, P, nsigma, nsegs):
    """src/common.jl:236-244 for the vehicle block: x[1:2], P[1:2,1:2] -> 2 x (nsegs+1)."""
    xs = np.zeros(3)
    xs[:2] = np.asarray(x, dtype=np.float64).reshape(-1)[:2]
    Ps = np.zeros((3, 3))
    Ps[:2, :2] = np.asarray(P, dtype=np.float64)[:2, :2]
    h = _scratch_handle(3)
    h.set_state(xs, Ps)
    return h.vehicle_ellipse(nsigma, nsegs)
