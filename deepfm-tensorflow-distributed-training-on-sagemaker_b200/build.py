"""In-tree native build: `python -m deepfm_b200.build [--force]`.

Produces two shared libraries next to the Python bindings (git-ignored, shipped to GPU boxes as-is):

  ops/_lib/libdeepfm_b200.so   every CUDA kernel in csrc/*.cu, compiled for sm_100a only
                               (`-gencode arch=compute_100a,code=sm_100a -lineinfo`), cudart linked statically,
                               C ABI declared in include/deepfm_b200.h, loaded with ctypes (no torch headers,
                               so a full rebuild takes seconds)
  data/_lib/libdeepfm_io.so    host-only C++17 (csrc/io/*.cc): TFRecord reader/writer, CRC32C, Example batch
                               parser, pipe-mode reader, prefetch loader, TF tensor-bundle checkpoint codec
"""
from __future__ import annotations

import concurrent.futures as cf
import glob
import os
import shutil
import subprocess
import sys

_HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(_HERE)
CSRC = os.path.join(REPO, "csrc")
INC = os.path.join(REPO, "include")
CUDA_LIB = os.path.join(_HERE, "ops", "_lib", "libdeepfm_b200.so")
TRACE_LIB = os.path.join(_HERE, "ops", "_lib", "libdeepfm_b200_trace.so")  # -DDFM_KTRACE: kernel timeline stamps
IO_LIB = os.path.join(_HERE, "data", "_lib", "libdeepfm_io.so")
OBJ_DIR = os.path.join(REPO, "build", "obj")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr", "-I", INC, "-I", CSRC]
CXX_FLAGS = ["-O3", "-fno-math-errno", "-std=c++17", "-fPIC", "-msse4.2", "-pthread", "-I", INC, "-I", CSRC, "-Wall"]


def _nvcc() -> str:
    return shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"


def _newer(target: str, deps) -> bool:
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.getmtime(d) <= t for d in deps)


def _run(cmd):
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("build command failed:\n%s\n%s\n%s" % (" ".join(cmd), r.stdout, r.stderr))
    return r.stdout + r.stderr


def build_cuda(force: bool = False, verbose: bool = False, trace: bool = False) -> str:
    """trace=True builds the timeline variant (csrc/common.cuh: %globaltimer stamps per kernel; relocatable device code so
    every translation unit shares one trace buffer) next to the normal library; DFM_KTRACE=1 makes ops/lib.py load it."""
    srcs = sorted(glob.glob(os.path.join(CSRC, "*.cu")))
    hdrs = sorted(glob.glob(os.path.join(CSRC, "*.cuh")) + glob.glob(os.path.join(INC, "*.h")))
    out_lib = TRACE_LIB if trace else CUDA_LIB
    obj_dir = OBJ_DIR + ("_trace" if trace else "")
    extra = ["-DDFM_KTRACE", "-rdc=true"] if trace else []
    if not force and _newer(out_lib, srcs + hdrs):
        return out_lib
    os.makedirs(obj_dir, exist_ok=True)
    os.makedirs(os.path.dirname(out_lib), exist_ok=True)
    objs = []

    def compile_one(src):
        obj = os.path.join(obj_dir, os.path.basename(src) + ".o")
        if force or not _newer(obj, [src] + hdrs):
            out = _run([_nvcc()] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + ["-c", src, "-o", obj])
            if verbose:
                print(out)
        return obj

    with cf.ThreadPoolExecutor(max_workers=min(8, len(srcs) or 1)) as ex:
        objs = list(ex.map(compile_one, srcs))
    _run([_nvcc(), "-shared", "-gencode", "arch=compute_100a,code=sm_100a"] + (["-rdc=true"] if trace else []) +
         ["-o", out_lib] + objs)
    return out_lib


def build_io(force: bool = False) -> str:
    srcs = sorted(f for f in glob.glob(os.path.join(CSRC, "io", "*.cc")) if not f.endswith("fuzz_main.cc"))
    hdrs = sorted(glob.glob(os.path.join(CSRC, "io", "*.h")) + glob.glob(os.path.join(INC, "*.h")))
    if not srcs:
        return ""
    if not force and _newer(IO_LIB, srcs + hdrs):
        return IO_LIB
    os.makedirs(os.path.dirname(IO_LIB), exist_ok=True)
    _run(["g++"] + CXX_FLAGS + ["-shared", "-o", IO_LIB] + srcs)
    return IO_LIB


def build_all(force: bool = False, verbose: bool = False):
    return build_cuda(force, verbose), build_io(force)


if __name__ == "__main__":
    libs = build_all(force="--force" in sys.argv, verbose="-v" in sys.argv)
    if "--trace" in sys.argv:
        libs = libs + (build_cuda(force="--force" in sys.argv, trace=True),)
    print("built:", libs)
