#!/usr/bin/env python3
"""The reference's shipped cube crosses (cuda_opengl/assets/cubemap/*.jpg, 4096x3072) as 8-bit
texels + stb_image's 256-entry 8-bit -> float table, for the GPU box (where /root/reference does
not exist).

    python tests/golden/make_cubemap_fixture.py        # needs /root/reference and oracle/_ref/ref_loader

oracle/_ref/ref_loader --cross-u8 decodes the JPEG with the reference's vendored stb_image, checks
that table[u8] equals what stbi_loadf (the call uploadCubemap makes, gpu_processor.cpp:96) returns
for EVERY texel, and dumps (w, h, nb_chan, table, texels).  This script packs that into
tests/golden/_large/<name>.cross_u8.npz (11-25 MB; git-ignored like oracle/_ref, travels with the
gpurun snapshot).  pt_testlib.load_cube_cross() rebuilds the float cross as table[u8]: bit-identical
to the reference's stbi_loadf result by the check above.
"""
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference/cuda_opengl/assets/cubemap"
OUT = os.path.join(HERE, "_large")
LOADER = os.path.join(REPO, "oracle", "_ref", "ref_loader")
NAMES = ["field_with_house", "water"]


def build(force=False):
    if not os.path.isdir(REF) or not os.path.exists(LOADER):
        return []
    os.makedirs(OUT, exist_ok=True)
    made = []
    for name in NAMES:
        src, dst = os.path.join(REF, name + ".jpg"), os.path.join(OUT, name + ".cross_u8.npz")
        if not os.path.exists(src):
            continue
        if os.path.exists(dst) and not force and os.path.getmtime(dst) >= os.path.getmtime(LOADER):
            made.append(dst)
            continue
        with tempfile.TemporaryDirectory() as tmp:
            raw = os.path.join(tmp, "cross.bin")
            subprocess.check_call([LOADER, "--cross-u8", src, raw])
            buf = np.fromfile(raw, dtype=np.uint8)
        w, h, nc = (int(x) for x in buf[:12].view(np.uint32))
        table = buf[12:12 + 1024].view(np.float32).copy()
        texels = buf[12 + 1024:].reshape(h, w, nc)
        np.savez_compressed(dst, texels=texels, table=table)
        made.append(dst)
    return made


if __name__ == "__main__":
    for p in build(force="--force" in sys.argv):
        print(p, os.path.getsize(p))
