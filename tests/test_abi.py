"""CPU tests of the C ABI: libslamcuda.so loads without a GPU, exports every symbol the header
declares, fails loudly (no CPU fallback), and its device-free helpers work."""
import ctypes as C
import re
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent


def header_symbols():
    txt = (ROOT / "include" / "slamcuda.h").read_text()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(slam_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_header_symbol():
    import slamcuda
    from slamcuda._ffi import lib
    syms = header_symbols()
    assert len(syms) >= 50
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/slamcuda.h but not exported"
    assert "sm_100a" in slamcuda.version()


def test_every_exported_symbol_is_declared():
    import subprocess
    from slamcuda._ffi import LIB_PATH
    out = subprocess.check_output(["nm", "-D", "--defined-only", str(LIB_PATH)], text=True)
    exported = sorted(set(re.findall(r" T (slam_[a-z0-9_]+)$", out, flags=re.M)))
    internal = {"slam_set_error", "slam_require_device"}
    assert [s for s in exported if s not in internal and s not in header_symbols()] == []


def test_no_cpu_fallback():
    import slamcuda
    if slamcuda.device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(slamcuda.SlamError) as e:
        slamcuda.predict(slamcuda.EKFSlamState(np.zeros(3), np.zeros((3, 3))), slamcuda.Vehicle(), np.eye(2), 0.025)
    assert "no CPU fallback" in str(e.value) and e.value.code == -4


def test_product_never_imports_oracle():
    pkg = ROOT / "slam.jl_b200"
    for p in list(pkg.rglob("*.py")) + list(pkg.rglob("*.cu")) + list(pkg.rglob("*.cuh")):
        t = p.read_text()
        assert "slam_ref" not in t and "slamtwin" not in t and "import oracle" not in t, p


def test_host_mpi_to_pi():
    import math
    import slamcuda
    assert slamcuda.mpi_to_pi(3.5) == 3.5 - 2 * math.pi
    assert slamcuda.mpi_to_pi(-3.5) == -3.5 + 2 * math.pi
    assert slamcuda.mpi_to_pi(1.0) == 1.0


def test_julia_wrapper_matches_header():
    """julia/ekf-cuda.jl cannot be executed here (no Julia): check that every ccall names a symbol
    the header declares and passes as many arguments as the C prototype has."""
    hdr = (ROOT / "include" / "slamcuda.h").read_text()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    protos = {}
    for m in re.finditer(r"\b(slam_[a-z0-9_]+)\s*\(([^;{]*?)\)\s*;", hdr, flags=re.S):
        args = m.group(2).strip()
        protos[m.group(1)] = 0 if args in ("", "void") else args.count(",") + 1
    jl = (ROOT / "julia" / "ekf-cuda.jl").read_text()
    calls = re.findall(r"ccall\(\(:(slam_[a-z0-9_]+), libslamcuda\),\s*\w+,\s*\(([^)]*)\)", jl, flags=re.S)
    assert len(calls) >= 12
    for name, argt in calls:
        assert name in protos, f"{name} is not declared in include/slamcuda.h"
        n = len([a for a in argt.split(",") if a.strip()])
        assert n == protos[name], f"{name}: ccall passes {n} arguments, the header declares {protos[name]}"
    for f in ("predict", "associate", "update", "add_features", "compute_association"):
        assert re.search(rf"^function {f}\(", jl, flags=re.M), f
