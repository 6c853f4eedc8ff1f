"""Build the sm_100a shared library in-tree (doapp_b200/lib/libdoapp_b200.so).

nvcc cross-compiles without a GPU; the built .so is git-ignored but travels to
the GPU box with the gpurun snapshot.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(PKG, "csrc")
LIB_DIR = os.path.join(PKG, "lib")
LIB = os.path.join(LIB_DIR, "libdoapp_b200.so")
HOST_LIB = os.path.join(LIB_DIR, "libdoapp_planning_state.so")
HOST_TEST = os.path.join(LIB_DIR, "planning_state_test")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
    "-fmad=false",  # every FMA in the kernels is written explicitly (parity with the oracle)
    "-Xcompiler", "-fPIC,-fvisibility=hidden,-ffp-contract=off",
]


def _nvcc() -> str:
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        raise RuntimeError("nvcc not found: the CUDA library cannot be built")
    return nvcc


def _digest(sources: list[str], flags: list[str]) -> str:
    import hashlib
    h = hashlib.sha256(" ".join(flags).encode())
    for s in sources:
        h.update(os.path.basename(s).encode())
        h.update(open(s, "rb").read())
    return h.hexdigest()


def _stale(target: str, sources: list[str], flags: list[str] | None = None) -> bool:
    """Content-hash staleness (a stamp file beside the target): file mtimes do not survive the
    copy to the GPU box, and a spurious rebuild there must not happen."""
    stamp = target + ".stamp"
    if not os.path.exists(target) or not os.path.exists(stamp):
        return True
    return open(stamp).read().strip() != _digest([s for s in sources if not s.endswith(".so")],
                                                 flags or [])


def _stamp(target: str, sources: list[str], flags: list[str] | None = None) -> None:
    with open(target + ".stamp", "w") as f:
        f.write(_digest([s for s in sources if not s.endswith(".so")], flags or []))


OBJ_DIR = os.path.join(LIB_DIR, "obj")
# translation units: (source, object suffix, extra defines).  The cost / planner kernels are
# compiled once per joint count so that the eight jobs run in parallel.
UNITS = [
    ("kernels_cost.cu", "_d4", ["-DDOAPP_D=4"]), ("kernels_cost.cu", "_d5", ["-DDOAPP_D=5"]),
    ("kernels_pk.cu", "_d4", ["-DDOAPP_D=4"]), ("kernels_pk.cu", "_d5", ["-DDOAPP_D=5"]),
    ("kernels_persist.cu", "_d4", ["-DDOAPP_D=4"]), ("kernels_persist.cu", "_d5", ["-DDOAPP_D=5"]),
    ("kernels_persist2.cu", "_d4", ["-DDOAPP_D=4"]), ("kernels_persist2.cu", "_d5", ["-DDOAPP_D=5"]),
    ("kernels_aux.cu", "", []), ("doapp_capi.cu", "", []),
]
HEADERS = ["kernels.cuh", "device_math.cuh", "cost_device.cuh"]


def build(force: bool = False, verbose: bool = False) -> str:
    from concurrent.futures import ThreadPoolExecutor
    os.makedirs(OBJ_DIR, exist_ok=True)
    hdrs = [os.path.join(CSRC, f) for f in HEADERS] + [os.path.join(ROOT, "include", "doapp", "doapp.h")]
    nvcc = _nvcc()

    def compile_unit(unit):
        src, suffix, defs = unit
        path = os.path.join(CSRC, src)
        obj = os.path.join(OBJ_DIR, os.path.splitext(src)[0] + suffix + ".o")
        flags = NVCC_FLAGS + defs
        deps = [path] + hdrs
        if not (force or _stale(obj, deps, flags)):
            return obj, False
        cmd = [nvcc, *flags, "-c", "-I", os.path.join(ROOT, "include"), "-I", CSRC, path, "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed ({r.returncode}) on {src}{suffix}:\n{r.stdout}\n{r.stderr}")
        if verbose:
            sys.stderr.write(r.stderr)
        _stamp(obj, deps, flags)
        return obj, True

    with ThreadPoolExecutor(max_workers=min(len(UNITS), os.cpu_count() or 1)) as ex:
        results = list(ex.map(compile_unit, UNITS))
    objs = [o for o, _ in results]
    if force or any(changed for _, changed in results) or not os.path.exists(LIB):
        r = subprocess.run([nvcc, *NVCC_FLAGS, "-shared", *objs, "-o", LIB, "-ldl"],
                           capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed ({r.returncode}):\n{r.stdout}\n{r.stderr}")
    host_dir = os.path.join(PKG, "host")
    host_src = os.path.join(host_dir, "planning_state.cpp")
    if os.path.exists(host_src):
        hdeps = [host_src, os.path.join(host_dir, "planning_state.hpp"),
                 os.path.join(host_dir, "ros_mirror.hpp"),
                 os.path.join(host_dir, "follower_check.hpp"), LIB]
        cxx = shutil.which("g++") or "g++"
        common = ["-O2", "-std=c++17", "-fPIC", "-I", os.path.join(ROOT, "include"), "-I", host_dir]
        # the reference vendors nanoflann (header-only, third party): with it on the include path
        # PlanningState::updateDistanceGrid(const KDTree&) exists as in planning_state.cuh:25-28
        nanoflann = os.environ.get("DOAPP_NANOFLANN_INCLUDE",
                                   "/root/reference/perturbator/nanoflann/include")
        if os.path.exists(os.path.join(nanoflann, "nanoflann.hpp")):
            common += ["-DDOAPP_WITH_NANOFLANN", "-isystem", nanoflann]
        if force or _stale(HOST_LIB, hdeps):
            subprocess.run([cxx, *common, "-shared", host_src, "-o", HOST_LIB, "-L", LIB_DIR,
                            "-ldoapp_b200", "-Wl,-rpath,$ORIGIN"], check=True)
            _stamp(HOST_LIB, hdeps)
        test_src = os.path.join(host_dir, "planning_state_test.cpp")
        if os.path.exists(test_src) and (force or _stale(HOST_TEST, hdeps + [test_src, HOST_LIB])):
            subprocess.run([cxx, *common, test_src, "-o", HOST_TEST, "-L", LIB_DIR,
                            "-ldoapp_planning_state", "-ldoapp_b200", "-Wl,-rpath,$ORIGIN"],
                           check=True)
            _stamp(HOST_TEST, hdeps + [test_src, HOST_LIB])
        node_src = os.path.join(host_dir, "motion_planner_main.cpp")
        node_bin = os.path.join(LIB_DIR, "motion_planner_main")
        if os.path.exists(node_src) and (force or _stale(node_bin, hdeps + [node_src, HOST_LIB])):
            subprocess.run([cxx, *common, node_src, "-o", node_bin, "-L", LIB_DIR,
                            "-ldoapp_planning_state", "-ldoapp_b200", "-Wl,-rpath,$ORIGIN"],
                           check=True)
            _stamp(node_bin, hdeps + [node_src, HOST_LIB])
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
