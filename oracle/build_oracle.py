#!/usr/bin/env python3
"""TEST INFRASTRUCTURE: compiles the CPU oracle (oracle/pt_oracle.cpp -> libpt_oracle.so)."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
SRC = os.path.join(HERE, "pt_oracle.cpp")
OUT = os.path.join(HERE, "libpt_oracle.so")


def build(force=False):
    if (not force and os.path.exists(OUT) and os.path.getmtime(OUT) >= os.path.getmtime(SRC)
            and os.path.getmtime(OUT) >= os.path.getmtime(os.path.join(REPO, "include", "ptcore.h"))):
        return OUT
    # -ffp-contract=off: one rounding per fp32 operation, like the reference's C++ on host cores
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-Wall", "-fPIC", "-shared", "-pthread",
                           "-ffp-contract=off", "-I", os.path.join(REPO, "include"), SRC, "-o", OUT])
    return OUT


if __name__ == "__main__":
    print(build(force=True))
