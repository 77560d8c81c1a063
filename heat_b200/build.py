"""In-tree build of the CUDA library (C ABI) and the `cf_c` pybind11 module.

`python -m heat_b200.build` or `__graft_entry__.build()`.  nvcc cross-compiles for sm_100a
without a GPU; the resulting .so files live next to the sources (git-ignored, but they travel
to the GPU box with the repo snapshot).
"""
from __future__ import annotations

import os
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libheat_b200.so")
CUDA_SOURCES = ["heat_abi.cu", "train_launch.cu", "train_launch_tile.cu", "train_bigk.cu", "probe.cu", "probe_tf32.cu", "eval_kernel.cu", "comm.cu", "ingest.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr"]


def _newer(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def _deps(directory):
    out = []
    for root, _dirs, files in os.walk(directory):
        out += [os.path.join(root, f) for f in files if f.endswith((".cu", ".cuh", ".h", ".hpp", ".cpp"))]
    out.append(os.path.join(HERE, "..", "include", "heat_b200.h"))
    return out


def _includes(path, seen=None):
    """The file and every quoted #include it reaches (per-object dependency list)."""
    seen = set() if seen is None else seen
    path = os.path.normpath(path)
    if path in seen or not os.path.exists(path):
        return seen
    seen.add(path)
    with open(path) as f:
        for line in f:
            line = line.strip()
            if line.startswith('#include "'):
                _includes(os.path.join(os.path.dirname(path), line.split('"')[1]), seen)
    return seen


def build_cuda(force=False, verbose=False):
    if not force and not _newer(LIB, _deps(CSRC)):
        return LIB
    objs = []
    procs = []
    os.makedirs(os.path.join(HERE, "build"), exist_ok=True)
    for src in CUDA_SOURCES:
        obj = os.path.join(HERE, "build", src + ".o")
        objs.append(obj)
        if not force and not verbose and not os.environ.get("HEAT_NVCC_DEFS") and \
                not _newer(obj, list(_includes(os.path.join(CSRC, src)))):
            continue
        cmd = ["nvcc", *NVCC_FLAGS, *os.environ.get("HEAT_NVCC_DEFS", "").split(), "-c", os.path.join(CSRC, src), "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        procs.append((cmd, subprocess.Popen(cmd, cwd=CSRC)))
    for cmd, p in procs:
        if p.wait() != 0:
            raise RuntimeError("nvcc failed: " + " ".join(cmd))
    subprocess.check_call(["nvcc", "-shared", "-o", LIB, *objs, "-lcudart", "-lcuda", "-ldl"])
    return LIB


def cf_c_path():
    return os.path.join(HERE, "cf", "cf_c" + sysconfig.get_config_var("EXT_SUFFIX"))


def build_cf_c(force=False):
    """pybind11 module `cf_c` (reference surface: cf_cpu/src/pybind/init_modules.cpp:13-152)."""
    import pybind11
    host = os.path.join(CSRC, "host")
    if not os.path.isdir(host):
        return None
    out = cf_c_path()
    srcs = [os.path.join(host, f) for f in sorted(os.listdir(host)) if f.endswith(".cpp")]
    if not force and not _newer(out, _deps(host) + [LIB]):
        return out
    cmd = ["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-fvisibility=hidden",
           "-I" + pybind11.get_include(), "-I" + sysconfig.get_paths()["include"],
           "-I" + os.path.join(HERE, "..", "include"), *srcs, "-o", out,
           "-L" + HERE, "-lheat_b200", "-Wl,-rpath,$ORIGIN/.."]
    subprocess.check_call(cmd)
    return out


def build_all(force=False):
    build_cuda(force=force)
    build_cf_c(force=force)


if __name__ == "__main__":
    build_all(force="--force" in sys.argv)
    print("built", LIB)
