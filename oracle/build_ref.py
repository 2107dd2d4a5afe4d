#!/usr/bin/env python3
"""TEST INFRASTRUCTURE -- recipe that builds the *reference itself* under oracle/_ref/.

Nothing here is product code and nothing is copied into the repository: the
reference sources are compiled where they lie under /root/reference, outputs go
only into oracle/_ref/ (git-ignored, but shipped to the GPU box by gpurun).

Artefacts
---------
oracle/_ref/ref_loader        the reference's host loader (scene.cpp, material_loader.cpp,
                              utils.cpp, texture_utils.cpp) linked against a malloc-backed
                              CUDA shim; dumps Face[]/Material[]/... to a .ptscene fixture
oracle/_ref/libref_cpu.so     the reference's raytrace.cu + .cuh headers compiled by g++
                              as plain C++ (device built-ins shimmed) -> reference radiance
                              on host cores, real cuRAND XORWOW from the toolkit header
oracle/_ref/libref_gpu.so     the reference's raytrace.cu mechanically patched for CUDA 12
                              (texture object instead of the removed texture/surface
                              references, linear RGBA8 output, runtime bounce count),
                              built -O2 for sm_100a: the GPU baseline "G0"

The patches are pure text substitutions listed in PATCHES_* below; each one is
asserted to apply exactly once so a drift of the reference is noticed.
"""
import os
import re
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
REF = os.environ.get("PT_REFERENCE", "/root/reference")
RC = os.path.join(REF, "cuda_opengl")
OUT = os.path.join(HERE, "_ref")
GEN = os.path.join(OUT, "gen")
CUDA = os.environ.get("CUDA_HOME", "/usr/local/cuda")

INC = ["-I", os.path.join(RC, "include"), "-I", os.path.join(RC, "3rd_party", "include"),
       "-I", os.path.join(CUDA, "include"), "-I", os.path.join(REPO, "include")]


def have_reference():
    return os.path.isfile(os.path.join(RC, "src", "shaders", "raytrace.cu"))


def run(cmd):
    print("+", " ".join(cmd), flush=True)
    subprocess.check_call(cmd)


def newer(target, *sources):
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.getmtime(s) <= t for s in sources if os.path.exists(s))


def sub_once(text, pattern, repl, count=1, flags=0):
    new, n = re.subn(pattern, repl, text, flags=flags)
    if n != count:
        raise RuntimeError(f"patch {pattern!r}: expected {count} match(es), got {n}")
    return new


# --- patches common to the CPU and the GPU build of raytrace.cu -----------------
def patch_common(src):
    # F7: bounce count becomes a runtime value (raytrace.cu:243)
    src = sub_once(src, r"int static_samples = 1;", "int static_samples = g_ref_static_samples;")
    # Q3: the reference reads inter.specular_col before writing it on a light hit
    # (intersection.cuh:199-212, raytrace.cu:129): define it as zero-initialised.
    src = sub_once(src, r"IntersectionData inter;", "IntersectionData inter = IntersectionData();")
    return src


def gen_cpu_source():
    src = open(os.path.join(RC, "src", "shaders", "raytrace.cu")).read()
    src = patch_common(src)
    # kernel<<<grid, block, 0, stream>>>(args) -> ref_launch(kernel, grid, block, args)
    src = sub_once(src, r"kernel<<<nb_blocks, threads_per_block, 0, stream>>>\(\s*",
                   "ref_launch(kernel, nb_blocks, threads_per_block, ")
    os.makedirs(GEN, exist_ok=True)
    open(os.path.join(GEN, "raytrace_cpu.cpp"), "w").write(src)
    # math_functions.h refuses to be included outside nvcc; an empty stand-in is enough
    open(os.path.join(GEN, "math_functions.h"), "w").write("#pragma once\n")


def gen_gpu_source():
    src = open(os.path.join(RC, "src", "shaders", "raytrace.cu")).read()
    src = patch_common(src)
    # F4: legacy texture / surface references were removed in CUDA 12
    src = sub_once(src, r"surface<void, cudaSurfaceType2D> surf;\n",
                   "__constant__ unsigned int* ref_out_rgba;\n"
                   "__constant__ int g_ref_static_samples;\n")
    src = sub_once(src, r"texture<float4, cudaTextureTypeCubemap> cubemap_ref;\n",
                   "__constant__ cudaTextureObject_t cubemap_ref;\n")
    src = sub_once(src, r"texCubemap\(cubemap_ref,", "texCubemap<float4>(cubemap_ref,", count=2)
    src = sub_once(src, r"surf2Dwrite\(rgbx\.b32, surf, x \* sizeof\(rgbx\), y, cudaBoundaryModeZero\);",
                   "ref_out_rgba[y * width + x] = rgbx.b32.x;")
    src = sub_once(src, r"  cudaBindSurfaceToArray\(surf, array\);\n", "")
    src = sub_once(src, r"  const scene::Cubemap& cubemap = cubemaps\[cubemap_id\];\n"
                        r"  cubemap_ref\.addressMode\[0\] = cudaAddressModeWrap;\n"
                        r"  cubemap_ref\.addressMode\[1\] = cudaAddressModeWrap;\n"
                        r"  cubemap_ref\.filterMode = cudaFilterModeLinear;\n"
                        r"  cubemap_ref\.normalized = true;\n"
                        r"  cudaBindTextureToArray\(cubemap_ref, cubemap\.cubemap, cubemap\.cubemap_desc\);\n",
                   "")
    os.makedirs(GEN, exist_ok=True)
    open(os.path.join(GEN, "raytrace_gpu.cu"), "w").write(src)


def build_loader():
    out = os.path.join(OUT, "ref_loader")
    srcs = [os.path.join(HERE, "ref_driver", "loader_dump.cpp"),
            os.path.join(HERE, "ref_shim", "cuda_host_shim.cpp"),
            os.path.join(RC, "src", "scene", "scene.cpp"),
            os.path.join(RC, "src", "scene", "material_loader.cpp"),
            os.path.join(RC, "src", "utils", "utils.cpp"),
            os.path.join(RC, "src", "utils", "texture_utils.cpp")]
    if newer(out, *srcs):
        return out
    run(["g++", "-std=c++14", "-O2", "-w", *INC, *srcs, "-o", out])
    return out


def build_cpu():
    out = os.path.join(OUT, "libref_cpu.so")
    drv = os.path.join(HERE, "ref_driver", "ref_cpu_driver.cpp")
    shim = os.path.join(HERE, "ref_shim", "ref_cpu_shim.h")
    if newer(out, drv, shim, __file__):
        return out
    gen_cpu_source()
    # -ffp-contract=off: the reference arithmetic is evaluated operation by
    # operation in IEEE fp32; the from-scratch oracle is compiled the same way.
    # -Bsymbolic: the CUDA-runtime stand-ins of cuda_host_shim.cpp must win inside this
    # library even when a real libcudart is already loaded globally (torch does that).
    run(["g++", "-std=c++14", "-O2", "-w", "-fPIC", "-shared", "-pthread", "-ffp-contract=off",
         "-Wl,-Bsymbolic", "-I", GEN, *INC, "-I", os.path.join(HERE, "ref_shim"), "-include", shim,
         drv, os.path.join(HERE, "ref_shim", "cuda_host_shim.cpp"), "-o", out])
    return out


def build_gpu():
    out = os.path.join(OUT, "libref_gpu.so")
    drv = os.path.join(HERE, "ref_driver", "ref_gpu_driver.cu")
    if not os.path.exists(drv):
        return None
    pkg = os.path.join(os.path.dirname(HERE), "cuda-pathtracer_b200")
    hdr = os.path.join(os.path.dirname(HERE), "include")
    if newer(out, drv, __file__, os.path.join(hdr, "ptcore.h"), os.path.join(hdr, "ptcore_raytrace.hpp")):
        return out
    if not os.path.exists(os.path.join(pkg, "libptcore.so")):
        raise SystemExit("build_ref: build cuda-pathtracer_b200/libptcore.so first (the compat entry links it)")
    gen_gpu_source()
    # the reference's own flags: -O2, no -arch, no fast-math (CMakeLists.txt:23-25);
    # the arch is the only addition.
    run([os.path.join(CUDA, "bin", "nvcc"), "-std=c++14", "-O2", "-w", "-lineinfo",
         "-gencode", "arch=compute_100a,code=sm_100a", "-Xcompiler", "-fPIC", "-shared",
         "-I", GEN, *INC, drv, "-o", out,
         # ref_gpu_raytrace_compat calls the product through include/ptcore_raytrace.hpp
         "-L", pkg, "-lptcore", "-Xlinker", "-rpath", "-Xlinker", "$ORIGIN/../../cuda-pathtracer_b200"])
    return out


def fixed_assets(tmp="/tmp/pt_assets_fixed"):
    """F6: the MTL files use back-slash texture paths that do not open on Linux.
    Builds a scratch asset tree whose MTL files have '\\' -> '/', everything else
    symlinked, so the UNMODIFIED reference loader takes its textured path."""
    src = os.path.join(RC, "assets")
    shutil.rmtree(tmp, ignore_errors=True)
    os.makedirs(os.path.join(tmp, "obj"))
    for f in os.listdir(src):
        if f.endswith(".scene"):
            shutil.copy(os.path.join(src, f), os.path.join(tmp, f))
    for f in os.listdir(os.path.join(src, "obj")):
        p = os.path.join(src, "obj", f)
        if f.endswith(".mtl"):
            open(os.path.join(tmp, "obj", f), "w").write(open(p).read().replace("\\", "/"))
        else:
            os.symlink(p, os.path.join(tmp, "obj", f))
    os.symlink(os.path.join(src, "cubemap"), os.path.join(tmp, "cubemap"))
    return tmp


def build_all(verbose=True):
    if not have_reference():
        if verbose:
            print("oracle/build_ref.py: /root/reference absent -- using prebuilt oracle/_ref as is")
        return False
    os.makedirs(OUT, exist_ok=True)
    build_loader()
    build_cpu()
    build_gpu()
    return True


if __name__ == "__main__":
    ok = build_all()
    sys.exit(0 if ok else 1)
