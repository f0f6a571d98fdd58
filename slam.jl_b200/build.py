"""Builds libslamcuda.so (sm_100a only) in-tree: slam.jl_b200/lib/libslamcuda.so.

nvcc cross-compiles without a GPU.  Each .cu is compiled to an object (in parallel, skipped
when newer than its sources) and linked with the static CUDA runtime; NCCL is dlopen()ed at
run time by shard.cu, so the library loads on boxes without it.
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

HERE = Path(__file__).resolve().parent
CSRC = HERE / "csrc"
LIB = HERE / "lib"
OBJ = HERE / "build"
INCLUDE = HERE.parent / "include"
TARGET = LIB / "libslamcuda.so"

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-lineinfo", "-std=c++17", "-Xcompiler", "-fPIC", "-Xcompiler", "-Wall",
          "-Xcompiler", "-fno-strict-aliasing", f"-I{INCLUDE}"]
# -fmad=false: products and sums round separately, matching the plain-C oracle twin
# (oracle/slam_twin.c is built with -ffp-contract=off): the single filter's update is BIT-EQUAL to the twin's for up to
# 32 ranks (tests/test_gpu_single.py::test_update_P_bitwise_vs_twin).  downdate*.cu use explicit fma().
PER_FILE = {
    "filter.cu": ["-fmad=false"],
    "gate.cu": ["-fmad=false"],
    "synth.cu": ["-fmad=false"],
    "shard.cu": ["-fmad=false"],
    # The ensemble kernel is the exception: contraction ON.  It never was bit-equal to the twin (sincos / atan2 of CUDA's
    # libm vs glibc already separate them at 1e-15, and its downdate is an fma / DMMA chain like the twin's), so separate
    # rounding bought nothing there and cost instructions in an instruction-bound kernel: same box, 8192 runs,
    # 64.18 -> 61.37 ms (+4.6 %), every map size still equal to the twin's, errors unchanged or smaller
    # (profiles/r2_ensemble.md).  SLAMCUDA_ENS_FMAD=false restores separate rounding.
    "ensemble.cu": ["-fmad=" + os.environ.get("SLAMCUDA_ENS_FMAD", "true")],
    "downdate.cu": [],
    "downdate_dmma.cu": [],
    "probe.cu": [],
    "downdate_stream.cu": [],
}


def _nvcc() -> str:
    for c in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if c and (os.path.isabs(c) and os.path.exists(c) or not os.path.isabs(c)):
            return c
    return "nvcc"


def _host_compiler() -> list[str]:
    # the image exports CC=/opt/gcc/bin/gcc; nvcc wants the system g++
    return ["-ccbin", "/usr/bin/g++"] if os.path.exists("/usr/bin/g++") else []


def _stale(obj: Path, deps: list[Path]) -> bool:
    if not obj.exists():
        return True
    t = obj.stat().st_mtime
    return any(d.stat().st_mtime > t for d in deps)


def build(verbose: bool = False, force: bool = False) -> Path:
    LIB.mkdir(exist_ok=True)
    OBJ.mkdir(exist_ok=True)
    headers = sorted(CSRC.glob("*.cuh")) + sorted(INCLUDE.glob("*.h")) + [Path(__file__)]
    sources = sorted(CSRC.glob("*.cu"))
    nvcc = _nvcc()
    jobs = []
    for src in sources:
        obj = OBJ / (src.stem + ".o")
        if force or _stale(obj, [src] + headers):
            cmd = [nvcc, *ARCH, *COMMON, *_host_compiler(), *PER_FILE.get(src.name, []),
                   *os.environ.get("SLAMCUDA_DEFS", "").split(),
                   "-Xptxas", "-v", "-c", str(src), "-o", str(obj)]
            jobs.append((src, cmd))

    def run(job):
        src, cmd = job
        r = subprocess.run(cmd, capture_output=True, text=True)
        (OBJ / (src.stem + ".ptxas.log")).write_text(r.stderr)
        return src, r

    with ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        for src, r in ex.map(run, jobs):
            if r.returncode != 0:
                sys.stderr.write("\n".join(l for l in (r.stdout + r.stderr).splitlines()
                                           if "ptxas info" not in l and "bytes stack frame" not in l) + "\n")
                raise RuntimeError(f"nvcc failed on {src.name}")
            if verbose:
                sys.stderr.write(r.stderr)
    objs = [OBJ / (s.stem + ".o") for s in sources]
    if force or jobs or _stale(TARGET, objs):
        cmd = [nvcc, *ARCH, *_host_compiler(), "-shared", "-o", str(TARGET), *map(str, objs),
               "-cudart", "static", "-ldl", "-lpthread"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            sys.stderr.write(r.stdout + r.stderr)
            raise RuntimeError("link failed")
    return TARGET


if __name__ == "__main__":
    p = build(verbose="-v" in sys.argv, force="-f" in sys.argv)
    print(p)
