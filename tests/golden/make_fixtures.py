#!/usr/bin/env python3
"""Generates the committed golden fixtures from the REFERENCE ITSELF.

Runs only in the build container (needs /root/reference).  It builds oracle/_ref
(the reference's loader and the reference's raytrace.cu compiled for host cores,
see oracle/build_ref.py), then freezes:

  *.ptscene[.xz]        output of the reference's own loader for each asset scene
                        ("plain" = as it loads on Linux, F6; "tex" = MTL back-slashes fixed)
  rng_kat.json          WangHash values and cuRAND XORWOW uniforms (toolkit header,
                        called exactly as raytrace.cu:234-235 does)
  ref_indoor_*.npz      small renders by the reference's radiance()/kernel() on host
                        cores: un-clamped radiance, accumulator, RGBA8, primary hits

Everything the -m "not gpu" and -m gpu tests compare against comes from these
files; none of the tests reads /root/reference.
"""
import ctypes as C
import json
import lzma
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "oracle"))
sys.path.insert(0, os.path.join(REPO, "tests"))

import build_ref  # noqa: E402
from pt_testlib import load_scene, RefCpu  # noqa: E402


def dump_scenes():
    loader = os.path.join(build_ref.OUT, "ref_loader")
    plain_assets = os.path.join(build_ref.RC, "assets")
    fixed_assets = build_ref.fixed_assets()
    jobs = [("indoor", plain_assets, "plain"), ("indoor", fixed_assets, "tex"),
            ("color_sample", plain_assets, "plain"), ("island", plain_assets, "plain"),
            ("crate_land", plain_assets, "plain"), ("sss_crate", plain_assets, "plain")]
    for scene, assets, tag in jobs:
        out = os.path.join(HERE, f"{scene}_{tag}.ptscene")
        subprocess.check_call([loader, assets, scene + ".scene", out],
                              stderr=subprocess.DEVNULL)
        raw = open(out, "rb").read()
        if len(raw) > 256 * 1024:
            if scene != "indoor":
                os.remove(out)  # only the headline scene ships its texels
                print(f"  {scene}_{tag}: {len(raw)} B -- too large, not committed")
                continue
            open(out + ".xz", "wb").write(lzma.compress(raw, preset=6))
            os.remove(out)


def dump_indexed():
    """indoor_indexed.ptix: the reference's own tinyobj output for indoor.obj (attribute arrays,
    index records, per-face material ids) BEFORE upload_meshes flattens it (scene.cpp:202-283);
    ptc_flatten_faces on it must reproduce the Face[] of indoor_plain.ptscene bit for bit."""
    loader = os.path.join(build_ref.OUT, "ref_loader")
    obj = os.path.join(build_ref.RC, "assets", "obj")
    subprocess.check_call([loader, "--indexed", os.path.join(obj, "indoor.obj"), obj + "/",
                           os.path.join(HERE, "indoor_indexed.ptix")], stderr=subprocess.DEVNULL)


def rng_kat(ref: RefCpu):
    frames = [1, 2, 3, 64, 1000]
    kat = {"wang_hash": {str(k): int(ref.wang_hash(k)) for k in frames + [0, 0xFFFFFFFF]},
           "uniforms": {}}
    # seeds of the form WangHash(frame) + tid, the tids SURVEY 8c names
    for tid in [0, 1, 255, 256, 8227 * 256 + 255]:
        seed = (ref.wang_hash(1) + tid) & 0xFFFFFFFF
        kat["uniforms"][str(seed)] = [float.hex(float(v)) for v in ref.uniforms(seed, 8)]
    for seed in [0, 1, 0xFFFFFFFF, 0xDEADBEEF]:
        kat["uniforms"][str(seed)] = [float.hex(float(v)) for v in ref.uniforms(seed, 8)]
    json.dump(kat, open(os.path.join(HERE, "rng_kat.json"), "w"), indent=1)


def ref_renders(ref: RefCpu):
    for tag in ["plain", "tex"]:
        sc = load_scene(f"indoor_{tag}")
        w, h = 160, 90
        out = {}
        for bounces in (3, 4):
            out[f"radiance_b{bounces}_f1"] = ref.radiance(sc, w, h, 1, bounces)
        out["radiance_b3_f7"] = ref.radiance(sc, w, h, 7, 3)
        out["radiance_preview"] = ref.radiance(sc, w, h, 1, 3, moved=True)
        accum = np.zeros((h, w, 3), np.float32)
        for k in range(1, 5):
            rgba = ref.kernel(sc, w, h, k, 3, accum)
        out["accum_b3_f1to4"] = accum
        out["rgba_b3_f1to4"] = rgba
        o, d = ref.primary_rays(sc, w, h, 1)
        ids, t = ref.intersect(sc, o, d)
        out["primary_ids_f1"] = ids.reshape(h, w)
        out["primary_t_f1"] = t.reshape(h, w)
        np.savez_compressed(os.path.join(HERE, f"ref_indoor_{tag}_160x90.npz"), **out)


if __name__ == "__main__":
    if not build_ref.build_all():
        sys.exit("needs /root/reference")
    dump_scenes()
    dump_indexed()
    ref = RefCpu()
    rng_kat(ref)
    ref_renders(ref)
    print("fixtures written to", HERE)
