"""Test configuration: registers the `gpu` marker, puts the package and the oracle on the
path and builds the native pieces (libslamcuda.so by nvcc cross-compilation, the oracle's C
twin by gcc).  The oracle is test infrastructure: only tests/ (and smoke / bench baselines)
import it."""
import ctypes
import os
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "slam.jl_b200"))
sys.path.insert(0, str(ROOT / "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _build_native():
    lib = ROOT / "slam.jl_b200" / "lib" / "libslamcuda.so"
    if not lib.exists() or os.environ.get("SLAMCUDA_REBUILD"):
        subprocess.check_call([sys.executable, str(ROOT / "slam.jl_b200" / "build.py")])
    twin = ROOT / "oracle" / "libslamtwin.so"
    if not twin.exists():
        subprocess.check_call(["make", "-C", str(ROOT / "oracle")])


_build_native()


def has_gpu() -> bool:
    import slamcuda
    return slamcuda.device_count() > 0


@pytest.fixture(scope="session")
def twin():
    """ctypes handle of the oracle's plain-C twin."""
    from twin_ffi import load_twin
    return load_twin()


@pytest.fixture(scope="session")
def course1():
    """sim/course1.txt of the reference, committed as a golden fixture (2 x 19 waypoints)."""
    return np.loadtxt(ROOT / "tests" / "golden" / "course1.txt", skiprows=1).T.copy()


def rel_err(a, b):
    """max-norm relative error of a against reference b."""
    a, b = np.asarray(a, dtype=float), np.asarray(b, dtype=float)
    if a.shape != b.shape:
        return float("inf")
    if a.size == 0:
        return 0.0
    scale = max(float(np.max(np.abs(b))), 1e-300)
    return float(np.max(np.abs(a - b))) / scale


def elem_err(x, P, xo, Po):
    """Element-wise relative error with an absolute floor, the stricter reading of "within 1e-9 relative":
    |a - b| <= tol * (|b| + sqrt(P_ii P_jj)) for every entry of P and tol * (|b| + sqrt(P_ii)) for x -- a small
    cross-covariance is judged against the standard deviations it couples, not against the largest entry of
    the matrix (which is what rel_err's max-norm does: with x ~ 5000 m it cannot see an error in a small entry)."""
    x, P, xo, Po = (np.asarray(a, dtype=float) for a in (x, P, xo, Po))
    if x.shape != xo.shape or P.shape != Po.shape:
        return float("inf")
    if x.size == 0:
        return 0.0
    dg = np.sqrt(np.abs(np.diag(Po)))
    fx = np.abs(xo) + dg
    fP = np.abs(Po) + np.outer(dg, dg)
    ex = np.max(np.where(fx > 0, np.abs(x - xo) / np.where(fx > 0, fx, 1.0), np.abs(x - xo) * 1e300))
    eP = np.max(np.where(fP > 0, np.abs(P - Po) / np.where(fP > 0, fP, 1.0), np.abs(P - Po) * 1e300))
    return float(max(ex, eP))


def state_err(x, P, xo, Po):
    """The figure the parity gates use: the larger of the max-norm relative errors of x and P and elem_err."""
    return max(rel_err(x, xo), rel_err(P, Po), elem_err(x, P, xo, Po))

