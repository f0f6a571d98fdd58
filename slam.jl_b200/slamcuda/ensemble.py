"""Ensemble mode: many independent runs of the reference's simulator loop in one kernel launch
(one CTA per filter); members are sharded by run across GPUs by giving each rank its own
Ensemble with a different run0 (no collective)."""
from __future__ import annotations

import ctypes as C

import numpy as np

from ._ffi import lib, check, dptr


def member_range(total_runs: int, world: int, rank: int):
    """(run0, count) of the members rank runs when total_runs are split over world ranks: contiguous
    run ids, the first total_runs % world ranks take one more.  A member's Philox streams are keyed
    by its run id alone, so the split does not change any member's result (no collective)."""
    if world < 1 or not 0 <= rank < world or total_runs < 0:
        raise ValueError("member_range: need 0 <= rank < world and total_runs >= 0")
    base, rem = divmod(total_runs, world)
    return rank * base + min(rank, rem), base + (1 if rank < rem else 0)


class Ensemble:
    def __init__(self, waypoints, landmarks, n_runs: int, seed: int = 0, run0: int = 0, nlaps: int = 2,
                 device: int = 0, extra_landmarks: int | None = None, threads: int | None = None,
                 first_pass_extra: int | None = None, growth: int | None = None, max_shared_extra: int | None = None):
        wp = np.ascontiguousarray(np.asarray(waypoints, dtype=np.float64).T).reshape(-1)   # 2 x nwp column-major
        lm = np.asarray(landmarks, dtype=np.float64).reshape(2, -1)
        lmf = np.ascontiguousarray(lm.T).reshape(-1)
        self.n_runs, self.nl = n_runs, lm.shape[1]
        self.ptr = C.c_void_p()
        check(lib.slam_ens_create(device, dptr(wp), len(wp) // 2, dptr(lmf) if lm.shape[1] else None, lm.shape[1],
                                  nlaps, seed, run0, n_runs, C.byref(self.ptr)))
        ns, nu = C.c_int(), C.c_int()
        check(lib.slam_ens_steps(self.ptr, C.byref(ns), C.byref(nu)))
        self.nsteps, self.nupdates = ns.value, nu.value
        if extra_landmarks is not None:
            check(lib.slam_ens_set_extra_landmarks(self.ptr, int(extra_landmarks)))
        if threads is not None:
            check(lib.slam_ens_set_threads(self.ptr, int(threads)))
        if first_pass_extra is not None:
            check(lib.slam_ens_set_first_pass_extra(self.ptr, int(first_pass_extra)))
        if growth is not None:     # 2 (default): unbounded like the reference; 1: shared-memory tiers; 0: fixed capacity
            check(lib.slam_ens_set_growth(self.ptr, int(growth)))
        if max_shared_extra is not None:
            check(lib.slam_ens_set_max_shared_extra(self.ptr, int(max_shared_extra)))

    @property
    def capacity(self) -> int:
        """3 + 2*(landmarks + extra_landmarks): the base capacity of a member (with growth > 0 a member that
        outgrows it is re-run with more room, as the reference's state simply grows)."""
        return check(lib.slam_ens_capacity(self.ptr))

    def count_flops(self, on: bool = True) -> float:
        """Switches the kernel's tally of algorithmic FP64 operations on / off; returns the last run's total."""
        tot = C.c_double(0.0)
        check(lib.slam_ens_count_flops(self.ptr, int(on), C.byref(tot)))
        return tot.value

    def tier_info(self):
        """[(capacity in states, members run), ...] for every launch of the last run, in order."""
        caps, mem = (C.c_int * 32)(), (C.c_int * 32)()
        n = check(lib.slam_ens_tier_info(self.ptr, caps, mem, 32))
        return [(caps[i], mem[i]) for i in range(min(n, 32))]

    def pass_info(self):
        """(head-room of the last run's first pass or -1 for a single pass, members the second pass re-ran)."""
        fe, redone = C.c_int(), C.c_int()
        check(lib.slam_ens_pass_info(self.ptr, C.byref(fe), C.byref(redone)))
        return fe.value, redone.value

    def steps_done(self):
        """(total filter-steps completed by the last run, per-member array)."""
        tot = C.c_longlong()
        per = np.zeros(self.n_runs, dtype=np.int32)
        check(lib.slam_ens_steps_done(self.ptr, C.byref(tot), per.ctypes.data_as(C.POINTER(C.c_int))))
        return int(tot.value), per

    def close(self):
        if self.ptr:
            lib.slam_ens_destroy(self.ptr)
            self.ptr = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def keep_tracks(self, on: bool = True):
        check(lib.slam_ens_keep_tracks(self.ptr, int(on)))

    def set_noise(self, ctrl, obs):
        """ctrl: (n_runs, 2, nsteps) standard normals (speed, steering); obs: (n_runs, pool)."""
        ctrl = np.asarray(ctrl, dtype=np.float64)
        assert ctrl.shape == (self.n_runs, 2, self.nsteps)
        c = np.ascontiguousarray(ctrl.transpose(0, 2, 1)).reshape(-1)      # per member: (speed, steer) per step
        o = np.ascontiguousarray(np.asarray(obs, dtype=np.float64).reshape(self.n_runs, -1))
        check(lib.slam_ens_set_noise(self.ptr, dptr(c), dptr(o.reshape(-1)), o.shape[1]))

    def run(self):
        check(lib.slam_ens_run(self.ptr))

    def sync(self) -> int:
        """Waits for the run; returns the number of members that stopped early."""
        return check(lib.slam_ens_sync(self.ptr))

    def results(self):
        pose = np.empty((self.n_runs, 3))
        nf = np.empty(self.n_runs, dtype=np.int32)
        st = np.empty(self.n_runs, dtype=np.int32)
        check(lib.slam_ens_get_results(self.ptr, dptr(pose), nf.ctypes.data_as(C.POINTER(C.c_int)),
                                       st.ctypes.data_as(C.POINTER(C.c_int))))
        return pose, nf, st

    def track(self, member: int):
        t = np.empty((self.nsteps, 3))
        check(lib.slam_ens_get_track(self.ptr, member, dptr(t)))
        return t.T.copy()

    def state(self, member: int, with_cov: bool = True):
        cap = check(lib.slam_ens_get_state(self.ptr, member, None, None, 0))   # n_final of this member
        x = np.empty(cap)
        P = np.empty((cap, cap), order="F") if with_cov else None
        n = check(lib.slam_ens_get_state(self.ptr, member, dptr(x), dptr(P) if with_cov else None, cap))
        return x[:n].copy(), (np.ascontiguousarray(P[:n, :n]) if with_cov else None)

    def launch_count(self) -> int:
        return int(lib.slam_ens_launch_count(self.ptr))
