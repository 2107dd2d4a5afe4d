"""Builds csrc/ -> libptcore.so (in-tree) with nvcc for sm_100a.

nvcc cross-compiles without a GPU; the built .so is git-ignored but travels to the
GPU box with the gpurun snapshot.
"""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libptcore.so")
CUDA = os.environ.get("CUDA_HOME", "/usr/local/cuda")
SOURCES = ["ptcore.cu", "kernels.cu", "group.cu", "bvh_builder.cpp", "flatten.cpp", "texpack.cpp", "png.cpp"]
HEADERS = ["bvh.h", "device_types.cuh", "kernels.cuh", "vecmath.cuh", "traverse.cuh", "shade.cuh",
           "ptcore_internal.h", "../../include/ptcore.h"]


def up_to_date():
    if not os.path.exists(OUT):
        return False
    t = os.path.getmtime(OUT)
    deps = [os.path.join(CSRC, f) for f in SOURCES + HEADERS] + [__file__]
    return all(os.path.getmtime(d) <= t for d in deps if os.path.exists(d))


def build(force=False, verbose=False):
    if not force and up_to_date():
        return OUT
    cmd = [os.path.join(CUDA, "bin", "nvcc"), "-std=c++17", "-O3", "-lineinfo",
           "-gencode", "arch=compute_100a,code=sm_100a",
           "-Xcompiler", "-fPIC,-O2,-Wall,-pthread,-ffp-contract=off", "-shared",
           "-I", os.path.join(REPO, "include"), "-I", CSRC]
    if verbose:
        cmd += ["-Xptxas", "-v"]
    cmd += [os.path.join(CSRC, s) for s in SOURCES] + ["-o", OUT]
    subprocess.check_call(cmd)
    return OUT


if __name__ == "__main__":
    import sys
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
