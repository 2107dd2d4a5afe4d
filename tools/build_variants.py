#!/usr/bin/env python3
"""Builds experiment variants of libptcore.so (same ABI, different -D switches) into
cuda-pathtracer_b200/_variants/<name>.so; select one with PTC_LIBRARY=<path>."""
import os, subprocess, sys
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(REPO, "cuda-pathtracer_b200"); CSRC = os.path.join(PKG, "csrc")
OUT = os.path.join(PKG, "_variants"); os.makedirs(OUT, exist_ok=True)
procs = []
for spec in sys.argv[1:]:
    name, _, defs = spec.partition("=")
    cmd = ["/usr/local/cuda/bin/nvcc", "-std=c++17", "-O3", "-lineinfo", "-gencode", "arch=compute_100a,code=sm_100a",
           "-Xcompiler", "-fPIC,-O2,-pthread,-ffp-contract=off", "-shared", "-I", os.path.join(REPO, "include"), "-I", CSRC]
    cmd += ["-D" + d for d in defs.split(",") if d]
    cmd += [os.path.join(CSRC, f) for f in ("ptcore.cu", "kernels.cu", "group.cu", "bvh_builder.cpp", "flatten.cpp", "texpack.cpp")] + ["-o", os.path.join(OUT, name + ".so")]
    procs.append((name, subprocess.Popen(cmd)))
for name, p in procs:
    print(name, "rc", p.wait())
