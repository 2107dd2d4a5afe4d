"""Shared helpers for tests / bench / fixtures: package import, fixture loading and
ctypes wrappers around the two CHECKERS (oracle/_ref = the reference itself on host
cores, oracle/libpt_oracle.so = the from-scratch restatement).  The product is only
ever reached through ``pkg().Core`` (the C ABI)."""
from __future__ import annotations

import ctypes as C
import importlib.util
import os
import sys

import numpy as np

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(REPO, "tests", "golden")
ORACLE_DIR = os.path.join(REPO, "oracle")
PKG_DIR = os.path.join(REPO, "cuda-pathtracer_b200")
PKG_NAME = "cuda_pathtracer_b200"


def pkg():
    """Imports the package (its directory name has a hyphen, so by path)."""
    if PKG_NAME in sys.modules:
        return sys.modules[PKG_NAME]
    spec = importlib.util.spec_from_file_location(
        PKG_NAME, os.path.join(PKG_DIR, "__init__.py"), submodule_search_locations=[PKG_DIR])
    mod = importlib.util.module_from_spec(spec)
    sys.modules[PKG_NAME] = mod
    spec.loader.exec_module(mod)
    return mod


_scene_cache = {}


def load_scene(name: str):
    """Loads tests/golden/<name>.ptscene[.xz] (cached)."""
    if name not in _scene_cache:
        ptscene = pkg().ptscene
        p = os.path.join(GOLDEN, name + ".ptscene")
        if not os.path.exists(p):
            p += ".xz"
        _scene_cache[name] = ptscene.load_ptscene(p)
    return _scene_cache[name]


def _f32p(a):
    return a.ctypes.data_as(C.c_void_p)


def load_cube_cross(name):
    """One of the reference's shipped cube crosses (assets/cubemap/<name>.jpg) as the float image
    stbi_loadf hands to uploadCubemap, rebuilt from tests/golden/_large/<name>.cross_u8.npz
    (tests/golden/make_cubemap_fixture.py); None when that file was not generated."""
    p = os.path.join(GOLDEN, "_large", name + ".cross_u8.npz")
    if not os.path.exists(p):
        return None
    z = np.load(p)
    return z["table"][z["texels"]]


def cube_of(sc):
    if sc.cubemap is None:
        raise ValueError("fixture carries no cubemap faces")
    return np.ascontiguousarray(sc.cubemap, dtype=np.float32), int(sc.cubemap_size)


class _RenderLib:
    """Common ctypes surface of libref_cpu.so (prefix ref_cpu_) and libpt_oracle.so
    (prefix oracle_): same argument lists on purpose."""

    def __init__(self, path, prefix):
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        self.lib = C.CDLL(path)
        self.p = prefix
        self.threads = os.cpu_count() or 1

    def _fn(self, name):
        return getattr(self.lib, self.p + name)

    def radiance(self, sc, w, h, frame, bounces, moved=False):
        cube, cs = cube_of(sc)
        desc = sc.to_desc()
        cam = sc.camera_struct()
        out = np.zeros((h, w, 3), np.float32)
        self._fn("radiance")(C.byref(desc), _f32p(cube), C.c_uint(cs), C.byref(cam),
                             C.c_uint(w), C.c_uint(h), C.c_uint(frame), C.c_uint(bounces),
                             C.c_int(int(moved)), _f32p(out), C.c_int(self.threads))
        return out

    def kernel(self, sc, w, h, frame, bounces, accum, moved=False, post_id=0):
        """One accumulate-into-framebuffer launch; returns RGBA8 (h,w,4) uint8."""
        cube, cs = cube_of(sc)
        desc = sc.to_desc()
        cam = sc.camera_struct()
        rgba = np.zeros((h, w), np.uint32)
        assert accum.dtype == np.float32 and accum.flags.c_contiguous
        self._fn("kernel")(C.byref(desc), _f32p(cube), C.c_uint(cs), C.byref(cam),
                           C.c_uint(w), C.c_uint(h), C.c_uint(frame), C.c_uint(bounces),
                           C.c_int(int(moved)), C.c_uint(post_id), _f32p(accum),
                           _f32p(rgba), C.c_int(self.threads))
        return rgba.view(np.uint8).reshape(h, w, 4)

    def intersect(self, sc, origins, dirs):
        desc = sc.to_desc()
        o = np.ascontiguousarray(origins, np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, np.float32).reshape(-1, 3)
        n = o.shape[0]
        ids = np.zeros(n, np.int32)
        t = np.zeros(n, np.float32)
        self._fn("intersect")(C.byref(desc), _f32p(o), _f32p(d), C.c_uint64(n),
                              _f32p(ids), _f32p(t), C.c_int(self.threads))
        return ids, t

    def primary_rays(self, sc, w, h, frame):
        cam = sc.camera_struct()
        o = np.zeros((h * w, 3), np.float32)
        d = np.zeros((h * w, 3), np.float32)
        self._fn("primary_rays")(C.byref(cam), C.c_uint(w), C.c_uint(h), C.c_uint(frame),
                                 _f32p(o), _f32p(d))
        return o, d


class RefCpu(_RenderLib):
    """The reference's own raytrace.cu on host cores (oracle/_ref/libref_cpu.so)."""

    def __init__(self):
        super().__init__(os.path.join(ORACLE_DIR, "_ref", "libref_cpu.so"), "ref_cpu_")
        self.lib.ref_wang_hash.restype = C.c_uint

    def wang_hash(self, a):
        return self.lib.ref_wang_hash(C.c_uint(a & 0xFFFFFFFF))

    def uniforms(self, seed, n):
        out = np.zeros(n, np.float32)
        self.lib.ref_curand_uniforms(C.c_uint32(seed), C.c_uint(n), _f32p(out))
        return out


class Oracle(_RenderLib):
    """The from-scratch CPU restatement (oracle/libpt_oracle.so)."""

    def __init__(self):
        super().__init__(os.path.join(ORACLE_DIR, "libpt_oracle.so"), "oracle_")
        self.lib.oracle_wang_hash.restype = C.c_uint

    def wang_hash(self, a):
        return self.lib.oracle_wang_hash(C.c_uint(a & 0xFFFFFFFF))

    def uniforms(self, seed, n):
        out = np.zeros(n, np.float32)
        self.lib.oracle_uniforms(C.c_uint32(seed), C.c_uint(n), _f32p(out))
        return out


class RefGpu:
    """GPU baseline G0: the reference's own kernel patched for CUDA 12 and built for
    sm_100a (oracle/_ref/libref_gpu.so, see oracle/build_ref.py).  Needs a GPU."""

    def __init__(self, sc):
        path = os.path.join(ORACLE_DIR, "_ref", "libref_gpu.so")
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        self.lib = C.CDLL(path)
        self.sc = sc
        self.h = C.c_void_p()
        cube, cs = cube_of(sc)
        desc = sc.to_desc()
        rc = self.lib.ref_gpu_create(C.byref(desc), _f32p(cube), C.c_uint(cs), C.byref(self.h))
        if rc:
            raise RuntimeError(f"ref_gpu_create failed: {rc}")

    def close(self):
        if self.h:
            self.lib.ref_gpu_destroy(self.h)
            self.h = C.c_void_p()

    def kernel(self, w, h, frame, bounces, d_accum, d_out, moved=False, post_id=0, stream=0):
        """Explicit-frame launch; d_accum / d_out are device pointers (ints)."""
        cam = self.sc.camera_struct()
        rc = self.lib.ref_gpu_kernel(self.h, C.byref(cam), C.c_uint(w), C.c_uint(h), C.c_uint(frame),
                                     C.c_uint(bounces), C.c_int(int(moved)), C.c_uint(post_id),
                                     C.c_void_p(d_accum), C.c_void_p(d_out), C.c_void_p(stream))
        if rc:
            raise RuntimeError(f"ref_gpu_kernel failed: {rc}")

    def raytrace(self, w, h, bounces, d_accum, d_out, moved=False, post_id=0, stream=0):
        """The reference's public raytrace() with its own frame counter."""
        cam = self.sc.camera_struct()
        rc = self.lib.ref_gpu_raytrace(self.h, C.byref(cam), C.c_uint(w), C.c_uint(h),
                                       C.c_uint(bounces), C.c_int(int(moved)), C.c_uint(post_id),
                                       C.c_void_p(d_accum), C.c_void_p(d_out), C.c_void_p(stream))
        if rc:
            raise RuntimeError(f"ref_gpu_raytrace failed: {rc}")

    def raytrace_compat(self, w, h, d_accum, d_out, moved=False, post_id=0, stream=0):
        """The PRODUCT through include/ptcore_raytrace.hpp's raytrace(uchar4*, Scenes, ...)
        overload, fed with this handle's reference-layout device tree and cubemap array."""
        cam = self.sc.camera_struct()
        rc = self.lib.ref_gpu_raytrace_compat(self.h, C.byref(cam), C.c_uint(w), C.c_uint(h),
                                              C.c_int(int(moved)), C.c_uint(post_id),
                                              C.c_void_p(d_accum), C.c_void_p(d_out), C.c_void_p(stream))
        if rc:
            raise RuntimeError(f"raytrace() compat overload failed: cudaError {rc}")

    def compat_reset(self):
        self.lib.ref_gpu_compat_reset()

    def compat_use_devices(self, devices):
        """ptc_compat::use_devices: later raytrace_compat calls render on all of `devices`."""
        arr = (C.c_int * len(devices))(*devices)
        rc = self.lib.ref_gpu_compat_use_devices(arr, C.c_uint(len(devices)))
        if rc:
            raise RuntimeError(f"ptc_compat::use_devices failed: {rc}")

    def compat_gather_accum(self, w, h):
        out = np.zeros((h, w, 3), np.float32)
        rc = self.lib.ref_gpu_compat_gather_accum(_f32p(out))
        if rc:
            raise RuntimeError(f"ptc_group_gather_accum failed: {rc}")
        return out

    def primary(self, w, h, frame):
        cam = self.sc.camera_struct()
        o = np.zeros((h * w, 3), np.float32)
        d = np.zeros((h * w, 3), np.float32)
        ids = np.zeros(h * w, np.int32)
        t = np.zeros(h * w, np.float32)
        rc = self.lib.ref_gpu_primary(self.h, C.byref(cam), C.c_uint(w), C.c_uint(h), C.c_uint(frame),
                                      _f32p(o), _f32p(d), _f32p(ids), _f32p(t))
        if rc:
            raise RuntimeError(f"ref_gpu_primary failed: {rc}")
        return o, d, ids.reshape(h, w), t.reshape(h, w)

    def intersect(self, origins, dirs):
        o = np.ascontiguousarray(origins, np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, np.float32).reshape(-1, 3)
        ids = np.zeros(o.shape[0], np.int32)
        t = np.zeros(o.shape[0], np.float32)
        rc = self.lib.ref_gpu_intersect(self.h, _f32p(o), _f32p(d), C.c_uint64(o.shape[0]),
                                        _f32p(ids), _f32p(t))
        if rc:
            raise RuntimeError(f"ref_gpu_intersect failed: {rc}")
        return ids, t

    def intersect_tiled(self, origins, dirs):
        """Same answer through the tiled brute-force kernels (blocks of rays x chunks of faces,
        the reference's own intersectTriangle; winner = min (t, face index)): affordable on a
        10 M-face mesh, pinned to intersect() on the small scenes."""
        o = np.ascontiguousarray(origins, np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, np.float32).reshape(-1, 3)
        ids = np.zeros(o.shape[0], np.int32)
        t = np.zeros(o.shape[0], np.float32)
        rc = self.lib.ref_gpu_intersect_tiled(self.h, _f32p(o), _f32p(d), C.c_uint64(o.shape[0]),
                                              _f32p(ids), _f32p(t))
        if rc:
            raise RuntimeError(f"ref_gpu_intersect_tiled failed: {rc}")
        return ids, t

    def rng(self, seeds, n_draws):
        s = np.ascontiguousarray(seeds, np.uint32)
        out = np.zeros((s.size, n_draws), np.float32)
        rc = self.lib.ref_gpu_rng(_f32p(s), C.c_uint32(s.size), C.c_uint32(n_draws), _f32p(out))
        if rc:
            raise RuntimeError(f"ref_gpu_rng failed: {rc}")
        return out


# ---------------------------------------------------------------------------
# shared input generators
# ---------------------------------------------------------------------------
def random_rays(sc, n, seed=1, pad=0.05):
    """Seeded rays with origins inside the scene's bounding box and uniform directions."""
    rng = np.random.default_rng(seed)
    v = sc.faces_f32()[:, :9].reshape(-1, 3)
    if v.shape[0] == 0:
        lo, hi = np.array([-1.0, -1.0, -1.0]), np.array([1.0, 1.0, 1.0])
    else:
        lo, hi = v.min(0), v.max(0)
    ext = (hi - lo) * pad
    o = (lo - ext + (hi - lo + 2 * ext) * rng.random((n, 3))).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    return o, d


# triangle soups that stress the builder and the quantised boxes (CPU: O1 == O2; GPU: device walk ==
# the reference's brute force)
def soup_scene(v):
    """Scene from bare triangles (F, 3, 3): flat normals, trivial uvs, one material, no lights."""
    synth = pkg().synth
    v = np.asarray(v, np.float32)
    f = v.shape[0]
    n = np.tile(np.array([0, 0, 1], np.float32), (f, 3, 1))
    uv = np.tile(np.array([[0, 0], [1, 0], [0, 1]], np.float32), (f, 1, 1))
    faces = synth.faces_from_arrays(v, n, uv, np.zeros(f, np.uint32))
    return synth.make_blob(faces, [f], [(0, -1, 1.0)], [np.ones((1, 1, 4), np.float32)], np.zeros((0, 8)),
                           synth.camera_array((0, 0, 3), (0, 0, -1)))


def aimed_rays(sc, n, rng):
    """Rays from random origins THROUGH vertices, edge midpoints and centroids of random faces: hits on
    shared edges and vertices are where two faces tie and where a box that is one ulp short loses a hit."""
    v = sc.faces_f32()[:, :9].reshape(-1, 3, 3).astype(np.float64)
    f = rng.integers(0, v.shape[0], n)
    w = rng.integers(0, 7, n)
    bary = np.array([[1, 0, 0], [0, 1, 0], [0, 0, 1], [.5, .5, 0], [0, .5, .5], [.5, 0, .5], [1 / 3, 1 / 3, 1 / 3]])[w]
    target = (v[f] * bary[:, :, None]).sum(1)
    lo, hi = v.reshape(-1, 3).min(0), v.reshape(-1, 3).max(0)
    o = lo - 0.3 * (hi - lo + 1e-3) + (1.6 * (hi - lo + 1e-3)) * rng.random((n, 3))
    d = target - o
    d /= np.maximum(np.linalg.norm(d, axis=1, keepdims=True), 1e-30)
    return o.astype(np.float32), d.astype(np.float32)


def triangle_soups():
    rng = np.random.default_rng(20240611)
    def tri(centres, size):
        return centres[:, None, :] + size * rng.normal(size=(centres.shape[0], 3, 3))
    yield "uniform small triangles", tri(rng.random((2000, 3)), 0.02)
    yield "long thin triangles across the scene", tri(rng.random((400, 3)), np.array([1.0, 0.01, 0.01]))
    flat = tri(rng.random((800, 3)), 0.05)
    flat[..., 2] = 0.25
    yield "all in one axis-aligned plane (zero extent on z)", flat
    c = rng.random((12, 3)) * 1000.0
    yield "tight clusters in a large scene", tri(np.repeat(c, 150, axis=0), 1e-3)
    yield "far from the origin", tri(rng.random((1500, 3)) + 1.0e4, 0.03)
    # (smaller still and the reference's own test rejects every face: det < 1e-7, intersection.cuh:110)
    yield "small scene near the reference's determinant cut-off", tri(rng.random((1500, 3)) * 0.05, 1.5e-3)
    mixed = tri(rng.random((1000, 3)), 0.03)
    mixed[::5, 1] = mixed[::5, 0]          # zero-area (two equal vertices)
    mixed[1::5, 2] = 0.5 * (mixed[1::5, 0] + mixed[1::5, 1])  # zero-area (collinear)
    yield "zero-area triangles mixed in", mixed
    g = 24  # a mesh with exactly shared vertices and edges, every cell twice (duplicates)
    xs = np.linspace(0, 1, g + 1)
    X, Y = np.meshgrid(xs, xs)
    P = np.stack([X, Y, 0.1 * np.sin(5 * X) * np.cos(4 * Y)], -1)
    a, b, c2, d = P[:-1, :-1], P[:-1, 1:], P[1:, 1:], P[1:, :-1]
    cells = np.concatenate([np.stack([a, b, c2], 2).reshape(-1, 3, 3), np.stack([a, c2, d], 2).reshape(-1, 3, 3)])
    yield "shared-edge mesh, every face twice", np.concatenate([cells, cells])
    yield "one triangle", tri(rng.random((1, 3)), 0.3)
    yield "four triangles", tri(rng.random((4, 3)), 0.3)


def oracle_bvh_intersect(orc, sc, hb, origins, dirs):
    """O2: the oracle's scalar traversal of the PRODUCT's BVH arrays (host_bvh())."""
    desc = sc.to_desc()
    o = np.ascontiguousarray(origins, np.float32).reshape(-1, 3)
    d = np.ascontiguousarray(dirs, np.float32).reshape(-1, 3)
    n = o.shape[0]
    ids = np.zeros(n, np.int32)
    t = np.zeros(n, np.float32)
    nn, nt = C.c_uint64(0), C.c_uint64(0)
    nodes = np.ascontiguousarray(hb["nodes"])
    tris = np.ascontiguousarray(hb["tris"])
    orc.lib.oracle_intersect_bvh(C.byref(desc), _f32p(nodes), _f32p(tris), _f32p(o), _f32p(d),
                                 C.c_uint64(n), _f32p(ids), _f32p(t), C.byref(nn), C.byref(nt),
                                 C.c_int(orc.threads))
    return ids, t, nn.value, nt.value


def oracle_render_stats(orc, sc, w, h, frame, bounces):
    """Oracle radiance + primary hit ids + (segments, traced) ray counts."""
    cube, cs = cube_of(sc)
    desc = sc.to_desc()
    cam = sc.camera_struct()
    rad = np.zeros((h, w, 3), np.float32)
    first = np.zeros((h, w), np.int32)
    seg, tr = C.c_uint64(0), C.c_uint64(0)
    orc.lib.oracle_render_stats(C.byref(desc), _f32p(cube), C.c_uint(cs), C.byref(cam), C.c_uint(w),
                                C.c_uint(h), C.c_uint(frame), C.c_uint(bounces), _f32p(rad),
                                _f32p(first), C.byref(seg), C.byref(tr), C.c_int(orc.threads))
    return rad, first, seg.value, tr.value
