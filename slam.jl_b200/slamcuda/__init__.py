"""slamcuda -- Python host side of libslamcuda.so, the B200-native EKF-SLAM hot path.

Mirrors the filter surface of andrewadare/SLAM.jl's `module SLAM` (predict, associate,
update, add_features, predict_observation, compute_association, mpi_to_pi,
                   compute_landmark_ellipses, ellipse) over a C ABI;
every filter operation runs in hand-written CUDA kernels for sm_100a.
"""
from ._ffi import (LIB_PATH, SlamError, NotPositiveDefinite, BadIndex, CapacityError, lib)  # noqa: F401
from .slam import (EKFSlamState, DeviceEKFSlamState, SlamState, Vehicle, predict, associate, update,  # noqa: F401
                   add_features, predict_observation, compute_association, mpi_to_pi,
                   compute_landmark_ellipses, ellipse, feature_ellipses, laser_lines, invalidate_mirror)


def version() -> str:
    return lib.slam_version().decode()


def device_count() -> int:
    return int(lib.slam_device_count())
