"""ctypes binding of the C ABI (include/ptcore.h) -- plumbing, not the product.

``Core`` mirrors the reference's call surface for the hot path: upload once
(``Scene::upload`` + ``uploadTextures``/``uploadCubemaps``, scene.cpp:304-391,
gpu_processor.cpp:163-238), then one ``render`` per frame
(``raytrace(...)``, raytrace.h:9-15).  The library is loaded from the in-tree
``libptcore.so``; a missing library or a machine without a CUDA device raises --
there is no CPU fallback.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import ptscene
from .ptscene import PtcCamera, PtcSceneDesc

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libptcore.so")

PTC_ERR_NO_DEVICE = -2


class PtcRenderArgs(C.Structure):
    _fields_ = [("scene_id", C.c_int32), ("cubemap_id", C.c_int32),
                ("camera", C.POINTER(PtcCamera)),
                ("width", C.c_uint32), ("height", C.c_uint32),
                ("row_begin", C.c_uint32), ("row_end", C.c_uint32),
                ("bounces", C.c_uint32), ("moved", C.c_uint32),
                ("post_id", C.c_uint32), ("frame_nb", C.c_uint32),
                ("accum", C.c_void_p), ("out_rgba", C.c_void_p),
                ("radiance_out", C.c_void_p), ("stream", C.c_void_p),
                ("strip_rows", C.c_uint32), ("strip_stride", C.c_uint32),
                ("strip_phase", C.c_uint32), ("flags", C.c_uint32)]


class PtcCounters(C.Structure):
    _fields_ = [("paths", C.c_uint64), ("segments", C.c_uint64), ("traced", C.c_uint64),
                ("rays_per_bounce", C.c_uint64 * 16),
                ("nodes_visited", C.c_uint64), ("tris_tested", C.c_uint64)]


class PtcSceneInfo(C.Structure):
    _fields_ = [("n_faces", C.c_uint32), ("n_lights", C.c_uint32),
                ("n_materials", C.c_uint32), ("n_textures", C.c_uint32),
                ("bvh_nodes", C.c_uint32), ("bvh_node_bytes", C.c_uint32),
                ("tri_bytes", C.c_uint32), ("shade_bytes", C.c_uint32),
                ("texture_bytes", C.c_uint64), ("smem_resident", C.c_uint32)]


class PtcProfile(C.Structure):
    _fields_ = [("n_launches", C.c_uint32), ("launch_ms", C.c_float * 130),
                ("launch_kind", C.c_uint32 * 130),
                ("launch_bounce0", C.c_uint32 * 130), ("launch_iters", C.c_uint32 * 130),
                ("frame_ms", C.c_float), ("traced", C.c_uint64), ("segments", C.c_uint64),
                ("nodes_visited", C.c_uint64), ("tris_tested", C.c_uint64),
                ("rays_per_bounce", C.c_uint64 * 16), ("queue_rays", C.c_uint64),
                ("queue_bytes_per_ray", C.c_uint32), ("hit_bytes_per_ray", C.c_uint32),
                ("frame_ms_shipped", C.c_float), ("pipelines", C.c_uint32),
                ("extend_lanes", C.c_uint64 * 8)]


class PtcHostBvh(C.Structure):
    _fields_ = [("nodes", C.c_void_p), ("n_nodes", C.c_uint32), ("node_bytes", C.c_uint32),
                ("tris", C.c_void_p), ("n_tris", C.c_uint32), ("tri_bytes", C.c_uint32),
                ("tri_face_ids", C.c_void_p), ("sah_cost", C.c_float),
                ("max_depth", C.c_uint32), ("_owner", C.c_void_p)]


EXPORTS = ["ptc_abi_version", "ptc_render_static", "ptc_group_upload_scene_from_device",
           "ptc_group_set_cubemap_from_array", "ptc_host_image_open", "ptc_host_image_close",
           "ptc_memcpy_strips_to_host", "ptc_memcpy_strips_to_device", "ptc_create", "ptc_destroy", "ptc_last_error", "ptc_upload_scene",
           "ptc_scene_info_get", "ptc_set_cubemap", "ptc_render", "ptc_render_host",
           "ptc_reset_accumulation", "ptc_counters_get", "ptc_counters_reset",
           "ptc_trace_primary", "ptc_intersect_rays", "ptc_debug_rng", "ptc_host_bvh_build",
           "ptc_host_bvh_free", "ptc_band_rows", "ptc_launch_count", "ptc_set_wavefront",
           "ptc_profile_frame", "ptc_render_host_async", "ptc_wait",
           "ptc_upload_scene_from_device", "ptc_set_cubemap_from_array",
           "ptc_set_cubemap_from_cross", "ptc_cubemap_faces_from_cross",
           "ptc_write_ppm", "ptc_write_png", "ptc_accum_save", "ptc_accum_load", "ptc_debug_read_bandwidth",
           "ptc_group_create", "ptc_group_destroy", "ptc_group_last_error", "ptc_group_size",
           "ptc_group_context", "ptc_group_upload_scene", "ptc_group_set_cubemap", "ptc_group_render",
           "ptc_group_sync", "ptc_group_image", "ptc_group_read_image", "ptc_group_gather_accum",
           "ptc_group_reset_accumulation", "ptc_shared_alloc", "ptc_shared_open", "ptc_shared_close",
           "ptc_shared_free", "ptc_flag_signal", "ptc_flag_wait", "ptc_memcpy_to_host",
           "ptc_indexed_face_count", "ptc_flatten_faces", "ptc_upload_scene_indexed",
           "ptc_texpack_create", "ptc_texpack_destroy", "ptc_texpack_last_error", "ptc_texpack_add_materials",
           "ptc_texpack_texture_count", "ptc_texpack_textures"]

_lib = None


class PtcError(RuntimeError):
    pass


STATIC_MAX_PIXELS = 262144  # PTC_STATIC_MAX_PIXELS (include/ptcore.h)


def load_library() -> C.CDLL:
    """Loads libptcore.so; raises (loudly) when the CUDA extension was not built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise PtcError(f"{LIB_PATH} is missing: run `python __graft_entry__.py` (build()) first; "
                           "this package has no CPU fallback")
        # PTC_LIBRARY: experiment builds of the same ABI (tools/build_variants.py)
        lib = C.CDLL(os.environ.get("PTC_LIBRARY") or LIB_PATH)
        lib.ptc_last_error.restype = C.c_char_p
        lib.ptc_last_error.argtypes = [C.c_void_p]
        lib.ptc_destroy.restype = None
        lib.ptc_destroy.argtypes = [C.c_void_p]
        lib.ptc_host_bvh_free.restype = None
        for name in ["ptc_upload_scene", "ptc_scene_info_get", "ptc_set_cubemap", "ptc_render",
                     "ptc_render_host", "ptc_reset_accumulation", "ptc_counters_get",
                     "ptc_counters_reset", "ptc_trace_primary", "ptc_intersect_rays", "ptc_debug_rng",
                     "ptc_launch_count", "ptc_set_wavefront", "ptc_profile_frame",
                     "ptc_render_host_async", "ptc_wait", "ptc_set_cubemap_from_cross",
                     "ptc_cubemap_faces_from_cross", "ptc_upload_scene_from_device",
                     "ptc_set_cubemap_from_array", "ptc_write_ppm", "ptc_write_png", "ptc_accum_save", "ptc_accum_load",
                     "ptc_debug_read_bandwidth", "ptc_group_create", "ptc_group_upload_scene",
                     "ptc_group_set_cubemap", "ptc_group_render", "ptc_group_sync", "ptc_group_read_image",
                     "ptc_group_gather_accum", "ptc_group_reset_accumulation", "ptc_shared_alloc",
                     "ptc_shared_open", "ptc_shared_close", "ptc_shared_free", "ptc_flag_signal",
                     "ptc_flag_wait", "ptc_memcpy_to_host", "ptc_render_static", "ptc_host_image_open",
                     "ptc_host_image_close", "ptc_memcpy_strips_to_host", "ptc_memcpy_strips_to_device"]:
            getattr(lib, name).restype = C.c_int
        lib.ptc_indexed_face_count.restype = C.c_uint64
        lib.ptc_flatten_faces.restype = C.c_int
        lib.ptc_upload_scene_indexed.restype = C.c_int
        lib.ptc_texpack_create.restype = C.c_int
        lib.ptc_texpack_add_materials.restype = C.c_int
        lib.ptc_texpack_textures.restype = C.c_int
        lib.ptc_texpack_texture_count.restype = C.c_uint32
        lib.ptc_texpack_texture_count.argtypes = [C.c_void_p]
        lib.ptc_texpack_last_error.restype = C.c_char_p
        lib.ptc_texpack_last_error.argtypes = [C.c_void_p]
        lib.ptc_texpack_destroy.restype = None
        lib.ptc_texpack_destroy.argtypes = [C.c_void_p]
        lib.ptc_group_last_error.restype = C.c_char_p
        lib.ptc_group_last_error.argtypes = [C.c_void_p]
        lib.ptc_group_destroy.restype = None
        lib.ptc_group_destroy.argtypes = [C.c_void_p]
        lib.ptc_group_size.restype = C.c_uint32
        lib.ptc_group_size.argtypes = [C.c_void_p]
        lib.ptc_group_context.restype = C.c_void_p
        lib.ptc_group_context.argtypes = [C.c_void_p, C.c_uint32]
        lib.ptc_group_image.restype = C.c_void_p
        lib.ptc_group_image.argtypes = [C.c_void_p]
        lib.ptc_band_rows.restype = C.c_uint32
        _lib = lib
    return _lib


def _ptr(a):
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return C.c_void_p(a.ctypes.data)
    if hasattr(a, "data_ptr"):  # torch tensor
        return C.c_void_p(a.data_ptr())
    return C.c_void_p(int(a))


def cubemap_faces_from_cross(cross, rgb_hex: int = 0xFFFFFFFF):
    """Host-only: the six faces (6, size, size, 4) the reference cuts out of a cube cross
    (texture_utils.cpp:5-52), or its constant-colour fallback.  Returns (faces, used_fallback)."""
    lib = load_library()
    size, fb = C.c_uint32(0), C.c_int32(0)
    if cross is None:
        out = np.zeros(24, np.float32)
        rc = lib.ptc_cubemap_faces_from_cross(None, 0, 0, 0, C.c_uint32(rgb_hex), _ptr(out), C.c_uint64(24),
                                              C.byref(size), C.byref(fb))
    else:
        cross = np.ascontiguousarray(cross, dtype=np.float32)
        h, w, nc = cross.shape
        out = np.zeros(max(24, (w // 4) ** 2 * 24), np.float32)
        rc = lib.ptc_cubemap_faces_from_cross(_ptr(cross), C.c_uint32(w), C.c_uint32(h), C.c_uint32(nc),
                                              C.c_uint32(rgb_hex), _ptr(out), C.c_uint64(out.size),
                                              C.byref(size), C.byref(fb))
    if rc:
        raise PtcError(f"ptc_cubemap_faces_from_cross failed: {rc}")
    s = size.value
    return out[: s * s * 24].reshape(6, s, s, 4).copy(), bool(fb.value)


def write_ppm(path: str, rgba: np.ndarray) -> None:
    """RGBA8 frame ((h, w) uint32 or (h, w, 4) uint8, row 0 = top) -> binary PPM."""
    rgba = np.ascontiguousarray(rgba)
    h, w = rgba.shape[:2]
    if load_library().ptc_write_ppm(path.encode(), _ptr(rgba), C.c_uint32(w), C.c_uint32(h)):
        raise PtcError(f"ptc_write_ppm({path}) failed")


def write_png(path: str, rgba: np.ndarray, keep_alpha: bool = False) -> None:
    """RGBA8 frame ((h, w) uint32 or (h, w, 4) uint8, row 0 = top) -> PNG (RGB, or RGBA)."""
    rgba = np.ascontiguousarray(rgba)
    h, w = rgba.shape[:2]
    if load_library().ptc_write_png(path.encode(), _ptr(rgba), C.c_uint32(w), C.c_uint32(h), C.c_int(int(keep_alpha))):
        raise PtcError(f"ptc_write_png({path}) failed")


def accum_save(path: str, accum: np.ndarray, samples: int) -> None:
    accum = np.ascontiguousarray(accum, dtype=np.float32)
    rows, w = accum.shape[:2]
    if load_library().ptc_accum_save(path.encode(), _ptr(accum), C.c_uint32(w), C.c_uint32(rows),
                                     C.c_uint32(samples)):
        raise PtcError(f"ptc_accum_save({path}) failed")


def accum_load(path: str):
    """Returns (accum (rows, w, 3) float32, samples)."""
    lib = load_library()
    w, rows, n = C.c_uint32(0), C.c_uint32(0), C.c_uint32(0)
    if lib.ptc_accum_load(path.encode(), None, C.c_uint64(0), C.byref(w), C.byref(rows), C.byref(n)):
        raise PtcError(f"ptc_accum_load({path}): not an accumulator file")
    out = np.zeros((rows.value, w.value, 3), np.float32)
    if lib.ptc_accum_load(path.encode(), _ptr(out), C.c_uint64(out.size), C.byref(w), C.byref(rows), C.byref(n)):
        raise PtcError(f"ptc_accum_load({path}) failed")
    return out, n.value


def host_bvh(scene: ptscene.SceneBlob) -> dict:
    """Host-only: builds the wide BVH and returns copies of its arrays (no GPU needed)."""
    lib = load_library()
    desc = scene.to_desc()
    hb = PtcHostBvh()
    rc = lib.ptc_host_bvh_build(C.byref(desc), C.byref(hb))
    if rc:
        raise PtcError(f"ptc_host_bvh_build failed: {rc}")
    try:
        nodes = np.ctypeslib.as_array((C.c_uint8 * (hb.n_nodes * hb.node_bytes)).from_address(hb.nodes)).copy()
        n_t = hb.n_tris * hb.tri_bytes
        tris = (np.ctypeslib.as_array((C.c_uint8 * n_t).from_address(hb.tris)).copy()
                if n_t else np.zeros(0, np.uint8))
        ids = (np.ctypeslib.as_array((C.c_uint32 * hb.n_tris).from_address(hb.tri_face_ids)).copy()
               if hb.n_tris else np.zeros(0, np.uint32))
        return {"nodes": nodes.reshape(hb.n_nodes, hb.node_bytes),
                "tris": tris.reshape(hb.n_tris, hb.tri_bytes) if hb.n_tris else tris.reshape(0, 48),
                "face_ids": ids, "sah_cost": float(hb.sah_cost), "max_depth": int(hb.max_depth)}
    finally:
        lib.ptc_host_bvh_free(C.byref(hb))


def flatten_faces(mesh: "ptscene.IndexedMesh"):
    """Host-only: indexed mesh -> (Face records uint8 (F, 112), faces per mesh uint32), ptc_flatten_faces."""
    lib = load_library()
    d = mesh.to_desc()
    n = int(lib.ptc_indexed_face_count(C.byref(d)))
    faces = np.zeros((n, 112), np.uint8)
    counts = np.zeros(len(mesh.shapes), np.uint32)
    rc = lib.ptc_flatten_faces(C.byref(d), _ptr(faces), C.c_uint64(n), _ptr(counts))
    if rc:
        raise PtcError(f"ptc_flatten_faces failed ({rc}): an index outside its attribute array?")
    return faces, counts


class TexPacker:
    """ptc_texpack_*: the material / texture table of MaterialLoader::load, ids continuing across
    scenes.  Host only."""

    def __init__(self):
        self.lib = load_library()
        self.p = C.c_void_p()
        if self.lib.ptc_texpack_create(C.byref(self.p)):
            raise PtcError("ptc_texpack_create failed")
        self._keep = []

    def close(self):
        if self.p:
            self.lib.ptc_texpack_destroy(self.p)
            self.p = C.c_void_p()

    def add_materials(self, mtl: "ptscene.MtlSet") -> np.ndarray:
        """Returns the scene's Material records (structured array, ptscene.MATERIAL_DTYPE)."""
        imgs = {}
        for name, a in mtl.images.items():
            im = ptscene.PtcImage()
            im.name = name.encode()
            if a is not None:
                a = np.ascontiguousarray(a, np.float32)
                im.h, im.w, im.nb_chan = a.shape
                im.data = a.ctypes.data
            self._keep.append(a)
            imgs[name] = im
        n = len(mtl.materials)
        ent = (ptscene.PtcMtlEntry * max(1, n))()
        for k, m in enumerate(mtl.materials):
            ent[k].diffuse[:] = [float(x) for x in m["diffuse"]]
            ent[k].specular[:] = [float(x) for x in m["specular"]]
            ent[k].ior = m["ior"]
            for field, name in zip(("diffuse_map", "specular_map", "normal_map"), m["names"]):
                if name:
                    if name not in imgs:   # named but not handed over: treated like a failed load
                        imgs[name] = ptscene.PtcImage()
                        imgs[name].name = name.encode()
                    setattr(ent[k], field, C.pointer(imgs[name]))
        self._keep.append(imgs)
        out = np.zeros(n, ptscene.MATERIAL_DTYPE)
        rc = self.lib.ptc_texpack_add_materials(self.p, ent, C.c_uint32(n), _ptr(out))
        if rc:
            raise PtcError(f"ptc_texpack_add_materials failed ({rc}): {self.lib.ptc_texpack_last_error(self.p).decode()}")
        return out

    def textures(self) -> list:
        """Copies of the global texture table so far: float32 arrays (h, w, nb_chan)."""
        n = int(self.lib.ptc_texpack_texture_count(self.p))
        arr = (ptscene.PtcTexture * max(1, n))()
        if self.lib.ptc_texpack_textures(self.p, arr, C.c_uint32(n)):
            raise PtcError("ptc_texpack_textures failed")
        out = []
        for i in range(n):
            t = arr[i]
            cnt = t.w * t.h * t.nb_chan
            out.append(np.ctypeslib.as_array((C.c_float * cnt).from_address(t.data)).copy().reshape(t.h, t.w, t.nb_chan))
        return out


class Core:
    """One context per GPU (gpu_processor.h:20-98 holds the same state in GPUProcessor)."""

    def __init__(self, device: int = 0):
        self.lib = load_library()
        self.ctx = C.c_void_p()
        rc = self.lib.ptc_create(C.c_int(device), C.byref(self.ctx))
        if rc == PTC_ERR_NO_DEVICE:
            raise PtcError("no usable CUDA device: the path-tracing core has no CPU fallback")
        if rc:
            raise PtcError(f"ptc_create failed: {rc}")
        self.device = device
        self._keep = []

    def close(self):
        if self.ctx:
            self.lib.ptc_destroy(self.ctx)
            self.ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc, what):
        if rc:
            msg = self.lib.ptc_last_error(self.ctx)
            raise PtcError(f"{what} failed ({rc}): {msg.decode() if msg else ''}")

    # -- upload ----------------------------------------------------------------
    def upload_scene(self, scene: ptscene.SceneBlob) -> int:
        desc = scene.to_desc()
        sid = C.c_int32(-1)
        self._check(self.lib.ptc_upload_scene(self.ctx, C.byref(desc), C.byref(sid)), "ptc_upload_scene")
        return sid.value

    def set_cubemap(self, faces: np.ndarray) -> int:
        faces = np.ascontiguousarray(faces, dtype=np.float32)
        assert faces.ndim == 4 and faces.shape[0] == 6 and faces.shape[3] == 4
        cid = C.c_int32(-1)
        self._check(self.lib.ptc_set_cubemap(self.ctx, _ptr(faces), C.c_uint32(faces.shape[1]),
                                             C.byref(cid)), "ptc_set_cubemap")
        return cid.value

    def set_cubemap_from_cross(self, cross, rgb_hex: int = 0xFFFFFFFF):
        """uploadCubemap (gpu_processor.cpp:68-161) for a (h, w, nb_chan) float cube cross, or --
        ``cross=None`` -- for the constant-colour path.  Returns (cubemap_id, used_fallback)."""
        cid, fb = C.c_int32(-1), C.c_int32(0)
        if cross is None:
            rc = self.lib.ptc_set_cubemap_from_cross(self.ctx, None, 0, 0, 0, C.c_uint32(rgb_hex),
                                                     C.byref(cid), C.byref(fb))
        else:
            cross = np.ascontiguousarray(cross, dtype=np.float32)
            h, w, nc = cross.shape
            rc = self.lib.ptc_set_cubemap_from_cross(self.ctx, _ptr(cross), C.c_uint32(w), C.c_uint32(h),
                                                     C.c_uint32(nc), C.c_uint32(rgb_hex), C.byref(cid),
                                                     C.byref(fb))
        self._check(rc, "ptc_set_cubemap_from_cross")
        return cid.value, bool(fb.value)

    def read_bandwidth(self, working_set_bytes: int, total_bytes: int) -> float:
        """Measured GB/s of 16-byte read-only loads over a working set of that size."""
        out = C.c_float(0)
        self._check(self.lib.ptc_debug_read_bandwidth(self.ctx, C.c_uint64(working_set_bytes),
                                                      C.c_uint64(total_bytes), C.byref(out)),
                    "ptc_debug_read_bandwidth")
        return float(out.value)

    def scene_info(self, scene_id: int) -> PtcSceneInfo:
        info = PtcSceneInfo()
        self._check(self.lib.ptc_scene_info_get(self.ctx, C.c_int32(scene_id), C.byref(info)),
                    "ptc_scene_info_get")
        return info

    # -- render ----------------------------------------------------------------
    def _args(self, scene_id, cubemap_id, camera, width, height, accum, out_rgba=None,
              radiance_out=None, bounces=3, moved=False, post_id=0, frame_nb=0, rows=None,
              stream=0, strips=None, out_full=False):
        cam = camera if isinstance(camera, PtcCamera) else camera.camera_struct()
        self._keep = [cam]
        a = PtcRenderArgs()
        a.scene_id, a.cubemap_id = scene_id, cubemap_id
        a.camera = C.pointer(cam)
        a.width, a.height = width, height
        a.row_begin, a.row_end = rows if rows else (0, 0)
        a.bounces, a.moved, a.post_id, a.frame_nb = bounces, int(bool(moved)), post_id, frame_nb
        a.accum = _ptr(accum)
        a.out_rgba = _ptr(out_rgba)
        a.radiance_out = _ptr(radiance_out)
        a.stream = C.c_void_p(int(stream)) if stream else None
        if strips:  # (strip_rows, strip_stride, strip_phase): interleaved image tiles
            a.strip_rows, a.strip_stride, a.strip_phase = strips
        a.flags = 1 if out_full else 0  # PTC_OUT_FULL_IMAGE
        return a

    def band_rows(self, width, height, rows=None, strips=None) -> int:
        a = PtcRenderArgs()
        a.width, a.height = width, height
        a.row_begin, a.row_end = rows if rows else (0, 0)
        if strips:
            a.strip_rows, a.strip_stride, a.strip_phase = strips
        return int(self.lib.ptc_band_rows(C.byref(a)))

    def launch_count(self) -> int:
        n = C.c_uint64(0)
        self._check(self.lib.ptc_launch_count(self.ctx, C.byref(n)), "ptc_launch_count")
        return n.value

    def set_wavefront(self, first_iters=0, next_iters=0, split_extend=-1, refill=0):
        self._check(self.lib.ptc_set_wavefront(self.ctx, C.c_uint32(first_iters), C.c_uint32(next_iters),
                                               C.c_int32(split_extend), C.c_uint32(refill)),
                    "ptc_set_wavefront")

    def set_wavefront_auto(self):
        """Back to the default schedule (first stage depth chosen per scene size)."""
        self._check(self.lib.ptc_set_wavefront(self.ctx, C.c_uint32(0xFFFFFFFF), C.c_uint32(0), C.c_int32(-1),
                                               C.c_uint32(0)), "ptc_set_wavefront")

    def profile_frame(self, *args, repeats=5, **kw) -> dict:
        """Per-launch CUDA-event times of one frame + traversal statistics (ptc_profile_frame)."""
        a = self._args(*args, **kw)
        pr = PtcProfile()
        self._check(self.lib.ptc_profile_frame(self.ctx, C.byref(a), C.c_uint32(repeats), C.byref(pr)),
                    "ptc_profile_frame")
        n = pr.n_launches
        kinds = ["first", "fused", "extend", "shade", "generate+intersect", "shade+intersect"]
        return {"launch_ms": [pr.launch_ms[i] for i in range(n)],
                "launch_kind": [kinds[pr.launch_kind[i]] for i in range(n)],
                "launch_bounce0": [pr.launch_bounce0[i] for i in range(n)],
                "launch_iters": [pr.launch_iters[i] for i in range(n)],
                "frame_ms": pr.frame_ms, "traced": pr.traced, "segments": pr.segments,
                "nodes_visited": pr.nodes_visited, "tris_tested": pr.tris_tested,
                "rays_per_bounce": [pr.rays_per_bounce[i] for i in range(16)],
                "queue_rays": pr.queue_rays, "queue_bytes_per_ray": pr.queue_bytes_per_ray,
                "hit_bytes_per_ray": pr.hit_bytes_per_ray,
                "frame_ms_shipped": pr.frame_ms_shipped, "pipelines": pr.pipelines,
                "extend_lanes": [pr.extend_lanes[i] for i in range(5)]}

    def render(self, *args, **kw):
        """Device-pointer render call (accum / out_rgba / radiance_out are device buffers)."""
        a = self._args(*args, **kw)
        self._check(self.lib.ptc_render(self.ctx, C.byref(a)), "ptc_render")

    def render_static(self, *args, **kw):
        """ptc_render without any allocation: scratch in the library's file-scope device array
        (at most STATIC_MAX_PIXELS pixels per call; calls on one device must not overlap)."""
        a = self._args(*args, **kw)
        self._check(self.lib.ptc_render_static(self.ctx, C.byref(a)), "ptc_render_static")

    def upload_scene_indexed(self, mesh, scene) -> int:
        """ptc_upload_scene_indexed: `mesh` = ptscene.IndexedMesh, materials / lights / textures of `scene`."""
        d = mesh.to_desc(scene)
        sid = C.c_int32(-1)
        self._check(self.lib.ptc_upload_scene_indexed(self.ctx, C.byref(d), C.byref(sid)), "ptc_upload_scene_indexed")
        return sid.value

    def render_host(self, *args, upload_accum=False, download_accum=False, **kw):
        """Host-buffer render call: copies in/out and synchronises."""
        a = self._args(*args, **kw)
        self._check(self.lib.ptc_render_host(self.ctx, C.byref(a), C.c_int(int(upload_accum)),
                                             C.c_int(int(download_accum))), "ptc_render_host")

    def render_host_async(self, *args, upload_accum=False, **kw):
        """Pipelined host-buffer call (returns before the frame is on the host); see wait()."""
        a = self._args(*args, **kw)
        self._check(self.lib.ptc_render_host_async(self.ctx, C.byref(a), C.c_int(int(upload_accum))),
                    "ptc_render_host_async")

    def wait(self, accum_out=None):
        nbytes = accum_out.nbytes if accum_out is not None else 0
        self._check(self.lib.ptc_wait(self.ctx, _ptr(accum_out), C.c_uint64(nbytes)), "ptc_wait")

    def reset_accumulation(self):
        self._check(self.lib.ptc_reset_accumulation(self.ctx), "ptc_reset_accumulation")

    def counters(self) -> PtcCounters:
        c = PtcCounters()
        self._check(self.lib.ptc_counters_get(self.ctx, C.byref(c)), "ptc_counters_get")
        return c

    def reset_counters(self):
        self._check(self.lib.ptc_counters_reset(self.ctx), "ptc_counters_reset")

    # -- parity entry points -----------------------------------------------------
    def trace_primary(self, scene_id, camera, width, height, frame_nb=1):
        cam = camera if isinstance(camera, PtcCamera) else camera.camera_struct()
        ids = np.zeros((height, width), np.int32)
        t = np.zeros((height, width), np.float32)
        self._check(self.lib.ptc_trace_primary(self.ctx, C.c_int32(scene_id), C.byref(cam),
                                               C.c_uint32(width), C.c_uint32(height),
                                               C.c_uint32(frame_nb), _ptr(ids), _ptr(t)),
                    "ptc_trace_primary")
        return ids, t

    def intersect_rays(self, scene_id, origins, dirs):
        o = np.ascontiguousarray(origins, np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, np.float32).reshape(-1, 3)
        ids = np.zeros(o.shape[0], np.int32)
        t = np.zeros(o.shape[0], np.float32)
        self._check(self.lib.ptc_intersect_rays(self.ctx, C.c_int32(scene_id), _ptr(o), _ptr(d),
                                                C.c_uint64(o.shape[0]), _ptr(ids), _ptr(t)),
                    "ptc_intersect_rays")
        return ids, t

    def debug_rng(self, seeds, n_draws, reference_impl=False):
        s = np.ascontiguousarray(seeds, np.uint32)
        out = np.zeros((s.size, n_draws), np.float32)
        self._check(self.lib.ptc_debug_rng(self.ctx, _ptr(s), C.c_uint32(s.size), C.c_uint32(n_draws),
                                           C.c_int(int(reference_impl)), _ptr(out)), "ptc_debug_rng")
        return out


class Group:
    """Several GPUs behind one call (ptc_group_*): one upload, one render per frame, one image on
    devices[0], the same bits as one GPU.  Single process; `devices` may repeat a device."""

    def __init__(self, devices):
        self.lib = load_library()
        self.devices = list(devices)
        arr = (C.c_int * len(self.devices))(*self.devices)
        g = C.c_void_p()
        rc = self.lib.ptc_group_create(arr, C.c_uint32(len(self.devices)), C.byref(g))
        if rc == PTC_ERR_NO_DEVICE:
            raise PtcError("ptc_group_create: a device is missing or has no peer access to devices[0] "
                           "(the core has no CPU fallback)")
        if rc:
            raise PtcError(f"ptc_group_create failed: {rc}")
        self.g = g
        self._keep = None

    def close(self):
        if self.g:
            self.lib.ptc_group_destroy(self.g)
            self.g = C.c_void_p()

    def _check(self, rc, what):
        if rc:
            raise PtcError(f"{what} failed ({rc}): {self.lib.ptc_group_last_error(self.g).decode()}")

    @property
    def size(self):
        return int(self.lib.ptc_group_size(self.g))

    def upload_scene(self, scene) -> int:
        desc = scene.to_desc()
        sid = C.c_int32(-1)
        self._check(self.lib.ptc_group_upload_scene(self.g, C.byref(desc), C.byref(sid)), "ptc_group_upload_scene")
        return sid.value

    def set_cubemap(self, faces) -> int:
        faces = np.ascontiguousarray(faces, dtype=np.float32)
        cid = C.c_int32(-1)
        self._check(self.lib.ptc_group_set_cubemap(self.g, _ptr(faces), C.c_uint32(faces.shape[1]), C.byref(cid)),
                    "ptc_group_set_cubemap")
        return cid.value

    def render(self, scene_id, cubemap_id, camera, width, height, out_rgba=None, bounces=3, moved=False,
               post_id=0, frame_nb=0, stream=0):
        cam = camera if isinstance(camera, PtcCamera) else camera.camera_struct()
        self._keep = cam
        a = PtcRenderArgs()
        a.scene_id, a.cubemap_id = scene_id, cubemap_id
        a.camera = C.pointer(cam)
        a.width, a.height = width, height
        a.bounces, a.moved, a.post_id, a.frame_nb = bounces, int(bool(moved)), post_id, frame_nb
        a.out_rgba = _ptr(out_rgba)
        a.stream = C.c_void_p(int(stream)) if stream else None
        self._check(self.lib.ptc_group_render(self.g, C.byref(a)), "ptc_group_render")

    def sync(self):
        self._check(self.lib.ptc_group_sync(self.g), "ptc_group_sync")

    def read_image(self, width, height, stream=0) -> np.ndarray:
        out = np.zeros((height, width), np.uint32)
        self._check(self.lib.ptc_group_read_image(self.g, _ptr(out), C.c_void_p(int(stream)) if stream else None),
                    "ptc_group_read_image")
        return out

    def gather_accum(self, width, height) -> np.ndarray:
        out = np.zeros((height, width, 3), np.float32)
        self._check(self.lib.ptc_group_gather_accum(self.g, _ptr(out)), "ptc_group_gather_accum")
        return out

    def reset_accumulation(self):
        self._check(self.lib.ptc_group_reset_accumulation(self.g), "ptc_group_reset_accumulation")


class SharedBuffer:
    """Device memory of one rank that the other ranks of a one-process-per-GPU job map through CUDA
    IPC (ptc_shared_*): the full image the ranks' kernels store their tiles into, and the flags."""

    def __init__(self, device, nbytes=0, handle=None):
        self.lib = load_library()
        self.device = device
        self.ptr = C.c_void_p()
        self.owner = handle is None
        if self.owner:
            h = (C.c_uint8 * 64)()
            rc = self.lib.ptc_shared_alloc(C.c_int(device), C.c_uint64(nbytes), C.byref(self.ptr), h)
            if rc:
                raise PtcError(f"ptc_shared_alloc failed: {rc}")
            self.handle = bytes(h)
        else:
            h = (C.c_uint8 * 64).from_buffer_copy(handle)
            rc = self.lib.ptc_shared_open(C.c_int(device), h, C.byref(self.ptr))
            if rc:
                raise PtcError(f"ptc_shared_open failed: {rc} (no peer access between the GPUs?)")
            self.handle = bytes(handle)

    @property
    def address(self) -> int:
        return int(self.ptr.value)

    def close(self):
        if self.ptr:
            (self.lib.ptc_shared_free if self.owner else self.lib.ptc_shared_close)(C.c_int(self.device), self.ptr)
            self.ptr = C.c_void_p()

    def to_host(self, out: np.ndarray, offset=0, stream=0, sync=True):
        rc = self.lib.ptc_memcpy_to_host(C.c_int(self.device), _ptr(out), C.c_void_p(self.address + offset),
                                         C.c_uint64(out.nbytes), C.c_void_p(int(stream)) if stream else None,
                                         C.c_int(int(sync)))
        if rc:
            raise PtcError(f"ptc_memcpy_to_host failed: {rc}")


class HostImage:
    """A page-locked host image in POSIX shared memory that every rank's process maps
    (ptc_host_image_open); ``slots`` full RGBA8 images of ``height`` x ``width``."""

    def __init__(self, name, height, width, slots=1, create=False):
        self.lib = load_library()
        self.name, self.h, self.w, self.slots, self.created = name, height, width, slots, create
        self.nbytes = slots * height * width * 4
        self.ptr = C.c_void_p()
        rc = self.lib.ptc_host_image_open(name.encode(), C.c_uint64(self.nbytes), C.c_int(int(create)), C.byref(self.ptr))
        if rc:
            raise PtcError(f"ptc_host_image_open({name}) failed: {rc}")

    def array(self) -> np.ndarray:
        buf = (C.c_uint32 * (self.nbytes // 4)).from_address(self.ptr.value)
        return np.frombuffer(buf, dtype=np.uint32).reshape(self.slots, self.h, self.w)

    def put_strips(self, device, slot, dev_band, strips, stream=0):
        """This rank's band (device) -> its rows of image ``slot``; asynchronous on ``stream``."""
        rows, stride, phase = strips
        dst = C.c_void_p(self.ptr.value + slot * self.h * self.w * 4)
        rc = self.lib.ptc_memcpy_strips_to_host(C.c_int(device), dst, _ptr(dev_band), C.c_uint32(self.w), C.c_uint32(self.h),
                                                C.c_uint32(rows), C.c_uint32(stride), C.c_uint32(phase),
                                                C.c_void_p(int(stream)) if stream else None)
        if rc:
            raise PtcError(f"ptc_memcpy_strips_to_host failed: {rc}")

    def close(self):
        if self.ptr:
            self.lib.ptc_host_image_close(self.ptr, C.c_uint64(self.nbytes), self.name.encode() if self.created else None)
            self.ptr = C.c_void_p()


def strips_to_device(device, image_address, dev_band, width, height, strips, stream=0):
    """ptc_memcpy_strips_to_device: this rank's band -> its rows of a (possibly peer-mapped) device image."""
    rows, stride, phase = strips
    rc = load_library().ptc_memcpy_strips_to_device(C.c_int(device), C.c_void_p(int(image_address)), _ptr(dev_band),
                                                    C.c_uint32(width), C.c_uint32(height), C.c_uint32(rows),
                                                    C.c_uint32(stride), C.c_uint32(phase),
                                                    C.c_void_p(int(stream)) if stream else None)
    if rc:
        raise PtcError(f"ptc_memcpy_strips_to_device failed: {rc}")


def flag_signal(device, address, value, stream=0):
    rc = load_library().ptc_flag_signal(C.c_int(device), C.c_void_p(address), C.c_uint32(value),
                                        C.c_void_p(int(stream)) if stream else None)
    if rc:
        raise PtcError(f"ptc_flag_signal failed: {rc}")


def flag_wait(device, address, n, stride_words, value, stream=0):
    rc = load_library().ptc_flag_wait(C.c_int(device), C.c_void_p(address), C.c_uint32(n), C.c_uint32(stride_words),
                                      C.c_uint32(value), C.c_void_p(int(stream)) if stream else None)
    if rc:
        raise PtcError(f"ptc_flag_wait failed: {rc}")
