"""Headless mirror of the reference's simulator loop `sim!(simdata)`
(sim/ekfslam-sim.jl:54-143, sleep and monitor removed) driving the CUDA filter.

The truth side (steer!, step_vehicle!, get_observations; src/common.jl:172-230,
sim/sim-utils.jl:12-75) is the caller's host code in the reference and stays host code
here; the four filter calls go to libslamcuda with the state resident on the device.
Noise is injected (the reference draws from Julia's global RNG, which cannot be
reproduced without Julia): ctrl_noise[0/1, step] are the standard normals of
add_control_noise! (speed, then steering), obs_noise is consumed in get_observations'
draw order (all range noises of a step, then all bearing noises).
"""
from __future__ import annotations

import math

import numpy as np

from .slam import DeviceEKFSlamState, Vehicle, predict, associate, update, add_features, mpi_to_pi

# literals of sim/ekfslam-sim.jl:29-37,62-76,114
WHEELBASE = 4.0
MAX_GAMMA = 60 * math.pi / 180
STEER_RATE = 60 * math.pi / 180
SENSOR_RANGE = 30.0
TARGET_SPEED = 8.0
D_MIN = 1.0
SIGMA_SPEED = 0.5
SIGMA_STEER = 3.0 * math.pi / 180
SIGMA_R = 0.1
SIGMA_B = 1.0 * math.pi / 180
DT = 0.025
DT_OBS = 8 * DT
GATE1, GATE2 = 4.0, 25.0


def control_cov():
    return np.array([[SIGMA_SPEED ** 2, 0.0], [0.0, SIGMA_STEER ** 2]])


def observation_cov():
    return np.array([[SIGMA_R ** 2, 0.0], [0.0, SIGMA_B ** 2]])


class Truth:
    """Noise-free vehicle (src/common.jl:172-230)."""

    def __init__(self, waypoints):
        wp = np.asarray(waypoints, dtype=float)
        self.wp = wp
        self.pose = np.array([wp[0, 0], wp[1, 0],
                              math.atan2(wp[1, 1] - wp[1, 0], wp[0, 1] - wp[0, 0])])   # initial_pose :93-96
        self.gamma = 0.0
        self.waypoint_id = 1

    def steer(self):
        """steer! -- src/common.jl:189-230."""
        x, y, phi = self.pose
        iwp = self.waypoint_id
        cx, cy = self.wp[0, iwp - 1], self.wp[1, iwp - 1]
        if (cx - x) ** 2 + (cy - y) ** 2 < D_MIN ** 2:
            iwp += 1
            if iwp > self.wp.shape[1]:
                self.waypoint_id = 0
                return
            cx, cy = self.wp[0, iwp - 1], self.wp[1, iwp - 1]
        dg = mpi_to_pi(math.atan2(cy - y, cx - x) - phi - self.gamma)
        dgmax = STEER_RATE * DT
        if abs(dg) > dgmax:
            dg = math.copysign(dgmax, dg)
        g = self.gamma + dg
        if abs(g) > MAX_GAMMA:
            g = math.copysign(MAX_GAMMA, g)
        self.gamma = g
        self.waypoint_id = iwp

    def step(self):
        """step_vehicle! -- src/common.jl:172-181 (uses the TARGET speed and steering)."""
        x, y, phi = self.pose
        g = self.gamma
        self.pose = np.array([x + TARGET_SPEED * DT * math.cos(g + phi),
                              y + TARGET_SPEED * DT * math.sin(g + phi),
                              mpi_to_pi(phi + TARGET_SPEED * DT * math.sin(g) / WHEELBASE)])

    def observe(self, landmarks):
        """Noise-free get_observations (sim/sim-utils.jl:12-28,53-66): 2 x nz and 1-based tags."""
        x, y, phi = self.pose
        zs, tags = [], []
        cphi, sphi = math.cos(phi), math.sin(phi)
        for i in range(landmarks.shape[1]):
            dx = landmarks[0, i] - x
            dy = landmarks[1, i] - y
            if (dx * cphi + dy * sphi) > 0 and (dx * dx + dy * dy) < SENSOR_RANGE ** 2:
                zs.append((math.sqrt(dx * dx + dy * dy), math.atan2(dy, dx) - phi))
                tags.append(i + 1)
        z = np.array(zs, dtype=float).T.reshape(2, -1) if zs else np.zeros((2, 0))
        return z, tags


def run(waypoints, landmarks, ctrl_noise, obs_noise, nlaps: int = 2, n_max: int | None = None,
        max_steps: int = 10000, device: int = 0):
    """One headless run; returns dict(true_track, slam_track, x, P, nsteps, nupdates)."""
    landmarks = np.asarray(landmarks, dtype=float)
    Q, R = control_cov(), observation_cov()
    truth = Truth(waypoints)
    n_max = n_max or 3 + 2 * landmarks.shape[1]
    state = DeviceEKFSlamState(truth.pose, np.zeros((3, 3)), n_max=n_max, device=device)
    vehicle = Vehicle(wheelbase=WHEELBASE)
    true_track = np.zeros((3, max_steps))
    slam_track = np.zeros((3, max_steps))
    nsteps = nupd = cursor = 0
    dtsum = 0
    while truth.waypoint_id != 0:
        truth.steer()
        if truth.waypoint_id == 0 and nlaps > 1:
            truth.waypoint_id = 1
            nlaps -= 1
        truth.step()
        vehicle.measured_speed = TARGET_SPEED + ctrl_noise[0, nsteps] * math.sqrt(Q[0, 0])
        vehicle.measured_gamma = truth.gamma + ctrl_noise[1, nsteps] * math.sqrt(Q[1, 1])
        state.x, state.cov = predict(state, vehicle, Q, DT)
        dtsum += DT
        if dtsum > DT_OBS:
            dtsum = 0
            nupd += 1
            z, _tags = truth.observe(landmarks)
            nz = z.shape[1]
            if nz > 0:
                z[0, :] = z[0, :] + obs_noise[cursor:cursor + nz] * math.sqrt(R[0, 0])
                z[1, :] = z[1, :] + obs_noise[cursor + nz:cursor + 2 * nz] * math.sqrt(R[1, 1])
                cursor += 2 * nz
            zf, idf, zn = associate(state, z, R, GATE1, GATE2)
            state.x, state.cov = update(state, zf, R, idf)
            state.x, state.cov = add_features(state, zn, R)
        nsteps += 1
        true_track[:, nsteps - 1] = truth.pose
        slam_track[:, nsteps - 1] = state.h.get_pose()
    x, P = state.h.get_state()
    launches = state.h.launch_count()
    state.h.close()
    return dict(true_track=true_track[:, :nsteps], slam_track=slam_track[:, :nsteps], x=x, P=P,
                nsteps=nsteps, nupdates=nupd, launches=launches)
