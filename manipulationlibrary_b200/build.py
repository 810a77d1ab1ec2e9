"""Builds libe2sec.so (CUDA kernels + C ABI + host resolver/readers) in-tree for sm_100a.

    python -m manipulationlibrary_b200.build [--force]

nvcc cross-compiles without a GPU. The .so is git-ignored but travels to the GPU box with the
repository snapshot. --fmad=false: float64 expressions keep numpy's unfused operation order
(see csrc/common.cuh); the host sources get -ffp-contract=off for the same reason.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libe2sec.so")
SOURCES = ["e2sec_api.cu", "resolver.cpp", "pcd_reader.cpp"]
HEADERS = ["common.cuh", "tma.cuh", "seg_fused.cuh", "seg_stream.cuh", "seg.cuh", "select.cuh", "radius.cuh", "filter.cuh", "pairs.cuh", "cells.cuh", "contact.cuh", "dissim.cuh", "relations.h", "phase2.cuh", "lzf.cuh", "frame0.cuh",
           os.path.join("..", "..", "include", "e2sec.h")]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")


def _stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, s) for s in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def build(force=False, verbose=False):
    if not force and not _stale():
        return LIB
    cmd = [NVCC, "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "--fmad=false", "-std=c++17",
           "-Xcompiler", "-fPIC,-O2,-ffp-contract=off,-fvisibility=default", "-shared", "-o", LIB]
    if verbose:
        cmd += ["-Xptxas", "-v"]
    cmd += os.environ.get("E2SEC_NVCC_FLAGS", "").split()          # tuning experiments (-DE2_CELL_THREADS=256 ...)
    cmd += [os.path.join(CSRC, s) for s in SOURCES]
    cmd += ["-lcudart"]
    subprocess.check_call(cmd)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
