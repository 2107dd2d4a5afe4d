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


MTL_CASES = """
newmtl m0
Kd 0.1 0.2 0.3
Ks 0.3 0.6 0.9
Ni 1.0
newmtl m1
Kd 0.5 0.5 0.5
Ks 0.2 0.2 0.2
map_Kd rgb_a.ppm
newmtl m2
Kd 0.25 0.5 0.75
map_Ks a_a.pgm
newmtl m3
map_Kd rgb_a.ppm
map_Ks a_a.pgm
map_bump n_a.ppm
Ni 1.5
newmtl m4
map_Kd rgb_a.ppm
map_Ks a_a.pgm
norm n_a.ppm
newmtl m5
Kd 0.9 0.8 0.7
map_Kd rgb_a.ppm
map_bump missing_n.ppm
newmtl m6
Kd 0.4 0.3 0.2
Ks 0.6 0.3 0.0
map_Kd missing.ppm
newmtl m7
Kd 0.7 0.1 0.1
map_Kd a_a.pgm
newmtl m8
Kd 0.2 0.9 0.4
map_Kd a_a.pgm
map_Ks a_a.pgm
newmtl m9
Kd 0.0 0.0 1.0
map_Kd a_a.pgm
map_Ks a_a.pgm
newmtl m10
Ks 0.3 0.3 0.9
map_Kd rgb_a.ppm
map_Ks rgb_a.ppm
newmtl m11
Kd 0.11 0.12 0.13
map_Kd missing.ppm
map_Ks missing2.pgm
newmtl m12
Kd 0.21 0.22 0.23
map_Kd missing.ppm
map_Ks missing2.pgm
newmtl m13
Kd 0.6 0.6 0.1
Ks 1 1 1
map_Ks rgb_b.ppm
newmtl m14
Kd 0.3 0.3 0.3
norm a_a.pgm
newmtl m15
map_Kd rgb_b.ppm
map_Ks a_b.pgm
map_bump n_a.ppm
"""


def dump_mtl_cases():
    """mtl_cases.ptscene / .ptmt: a synthetic OBJ whose 16 materials walk every branch of
    MaterialLoader::getTextureId (material_loader.cpp:243-415) -- no map, one map, both maps (cache
    hit on the second use), a normal map through map_bump and through norm, files that do not exist,
    maps with the wrong channel count -- loaded by the REFERENCE's loader (.ptscene = its answer) and
    dumped before packing (.ptmt = tinyobj materials + stb-decoded images = ptc_texpack's input)."""
    import tempfile
    loader = os.path.join(build_ref.OUT, "ref_loader")
    rng = np.random.default_rng(11)
    with tempfile.TemporaryDirectory() as tmp:
        obj = os.path.join(tmp, "obj")
        os.makedirs(obj)

        def pnm(name, w, h, chan):
            px = rng.integers(0, 256, (h, w, chan), dtype=np.uint8)
            with open(os.path.join(obj, name), "wb") as f:
                f.write(b"P6\n" if chan == 3 else b"P5\n")
                f.write(f"{w} {h}\n255\n".encode())
                f.write(px.tobytes())
        pnm("rgb_a.ppm", 8, 8, 3), pnm("a_a.pgm", 8, 8, 1), pnm("n_a.ppm", 8, 8, 3)
        pnm("rgb_b.ppm", 4, 2, 3), pnm("a_b.pgm", 4, 2, 1)
        open(os.path.join(obj, "cases.mtl"), "w").write(MTL_CASES)
        n_mat = MTL_CASES.count("newmtl")
        lines = ["mtllib cases.mtl", "v 0 0 0", "v 1 0 0", "v 0 1 0", "vn 0 0 1", "vt 0 0", "vt 1 0", "vt 0 1", "o tri"]
        for k in range(n_mat):
            lines += [f"usemtl m{k}", "f 1/1/1 2/2/1 3/3/1"]
        open(os.path.join(obj, "cases.obj"), "w").write("\n".join(lines) + "\n")
        open(os.path.join(tmp, "cases.scene"), "w").write(
            "camera 0 0 3 0 0 -1 90.0 3.0 0.01\np_light 0 2 0 1 1 1 2.0 0.3\nscene obj/cases.obj\n")
        subprocess.check_call([loader, tmp, "cases.scene", os.path.join(HERE, "mtl_cases.ptscene")],
                              stderr=subprocess.DEVNULL)
        subprocess.check_call([loader, "--materials", os.path.join(obj, "cases.obj"), obj + "/",
                               os.path.join(HERE, "mtl_cases.ptmt")], stderr=subprocess.DEVNULL)


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
    dump_mtl_cases()
    ref = RefCpu()
    rng_kat(ref)
    ref_renders(ref)
    print("fixtures written to", HERE)
