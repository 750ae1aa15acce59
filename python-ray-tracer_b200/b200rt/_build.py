"""Compile csrc/b200rt.cu into libb200rt.so next to this file (in-tree, sm_100a only)."""
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
SRC_DIR = os.path.join(os.path.dirname(_HERE), "csrc")
LIB_PATH = os.path.join(_HERE, "libb200rt.so")

# Arithmetic flags pinned to the reference kernel's (numba default: no fast-math).
NVCC_FLAGS = ["-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo",
              "--fmad=true", "--prec-div=true", "--prec-sqrt=true", "--ftz=false",
              "-shared", "-Xcompiler", "-fPIC,-fvisibility=hidden"]


SOURCES = ["b200rt.cu", "isect_shadow.cu"]       # two CUDA modules on purpose: two constant banks (csrc/isect3.cuh)


def needs_build() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    deps = [os.path.join(SRC_DIR, f) for f in os.listdir(SRC_DIR)] + \
           [os.path.join(os.path.dirname(os.path.dirname(_HERE)), "include", "b200rt.h")]
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def build(force: bool = False, verbose: bool = False) -> str:
    if force or needs_build():
        cmd = ["nvcc"] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + \
              ["--threads", "2", "-o", LIB_PATH] + [os.path.join(SRC_DIR, f) for f in SOURCES]
        subprocess.check_call(cmd)
    return LIB_PATH
