"""Reader for ``.ptscene`` containers and the ctypes view of the C ABI records.

A ``.ptscene`` file is the flat form of what the reference's loader hands to its
kernel: ``Camera``, ``LightProp[]``, ``Material[]``, per-mesh ``Face[]``, the global
``Texture[]`` table and the cubemap faces (reference:
cuda_opengl/include/scene/scene_data.h:31-133).  Fixtures under ``tests/golden`` are
written by ``oracle/ref_driver/loader_dump.cpp`` from the reference's own loader;
``SceneBlob.to_desc()`` turns one into the ``ptc_scene_desc`` of ``include/ptcore.h``.
"""
from __future__ import annotations

import ctypes as C
import lzma
import struct
from dataclasses import dataclass, field

import numpy as np

FACE_BYTES = 112
LIGHT_BYTES = 32
CAMERA_BYTES = 64


class PtcFace(C.Structure):
    _fields_ = [("vertices", C.c_float * 9), ("normals", C.c_float * 9),
                ("texcoords", C.c_float * 6), ("tangent", C.c_float * 3),
                ("material_id", C.c_uint32)]


class PtcMaterial(C.Structure):
    _fields_ = [("diffuse_spec_map", C.c_int32), ("normal_map", C.c_int32),
                ("ior", C.c_float), ("_pad", C.c_uint32)]


class PtcLight(C.Structure):
    _fields_ = [("color", C.c_float * 3), ("vec", C.c_float * 3),
                ("emission", C.c_float), ("radius", C.c_float)]


class PtcTexture(C.Structure):
    _fields_ = [("w", C.c_int32), ("h", C.c_int32), ("nb_chan", C.c_int32),
                ("_pad0", C.c_int32), ("data", C.c_void_p), ("_pad1", C.c_uint64)]


class PtcCamera(C.Structure):
    _fields_ = [("position", C.c_float * 3), ("dir", C.c_float * 3),
                ("u", C.c_float * 3), ("v", C.c_float * 3), ("fov_x", C.c_float),
                ("speed", C.c_float), ("aperture", C.c_float), ("focus_dist", C.c_float)]


class PtcSceneDesc(C.Structure):
    _fields_ = [("faces", C.c_void_p), ("mesh_face_counts", C.c_void_p),
                ("n_meshes", C.c_uint32), ("materials", C.c_void_p),
                ("n_materials", C.c_uint32), ("lights", C.c_void_p),
                ("n_lights", C.c_uint32), ("textures", C.c_void_p),
                ("n_textures", C.c_uint32)]


assert C.sizeof(PtcFace) == 112 and C.sizeof(PtcMaterial) == 16
assert C.sizeof(PtcLight) == 32 and C.sizeof(PtcTexture) == 32
assert C.sizeof(PtcCamera) == 64


def stb_gamma_lut() -> np.ndarray:
    """stb_image.h:1565-1582: ``(float)pow(u8 / 255.0f, 2.2f)`` (float divide, double pow)."""
    v = (np.arange(256, dtype=np.float32) / np.float32(255.0)).astype(np.float64)
    return np.power(v, np.float64(np.float32(2.2))).astype(np.float32)


@dataclass
class SceneBlob:
    camera: np.ndarray            # float32[16] == scene::Camera
    lights: np.ndarray            # float32[n,8]
    materials: np.ndarray         # structured (i4,i4,f4,u4)[n]
    mesh_face_counts: np.ndarray  # uint32[n_meshes]
    faces: np.ndarray             # uint8[n_faces,112] == scene::Face[]
    textures: list                # list of float32 arrays (h,w,nb_chan)
    cubemap_path: str = ""
    cubemap_size: int = 1
    cubemap: np.ndarray | None = None  # float32[6,size,size,4]
    _keep: list = field(default_factory=list, repr=False)

    @property
    def n_faces(self) -> int:
        return int(self.faces.shape[0])

    @property
    def n_lights(self) -> int:
        return int(self.lights.shape[0])

    def faces_f32(self) -> np.ndarray:
        """float32 view [n_faces, 28]; column 27 holds material_id bits."""
        return self.faces.view(np.float32).reshape(-1, 28)

    def material_ids(self) -> np.ndarray:
        return self.faces.view(np.uint32).reshape(-1, 28)[:, 27].copy()

    def camera_struct(self) -> PtcCamera:
        cam = PtcCamera()
        C.memmove(C.byref(cam), self.camera.ctypes.data, CAMERA_BYTES)
        return cam

    def to_desc(self) -> PtcSceneDesc:
        """ctypes ``ptc_scene_desc`` over this blob's arrays (kept alive by the blob)."""
        tex = (PtcTexture * max(1, len(self.textures)))()
        for i, t in enumerate(self.textures):
            tex[i].w, tex[i].h, tex[i].nb_chan = t.shape[1], t.shape[0], t.shape[2]
            tex[i].data = t.ctypes.data
        self._keep.append(tex)
        d = PtcSceneDesc()
        d.faces = self.faces.ctypes.data
        d.mesh_face_counts = self.mesh_face_counts.ctypes.data
        d.n_meshes = len(self.mesh_face_counts)
        d.materials = self.materials.ctypes.data
        d.n_materials = len(self.materials)
        d.lights = self.lights.ctypes.data
        d.n_lights = len(self.lights)
        d.textures = C.addressof(tex)
        d.n_textures = len(self.textures)
        return d


MATERIAL_DTYPE = np.dtype([("diffuse_spec_map", "<i4"), ("normal_map", "<i4"),
                           ("ior", "<f4"), ("_pad", "<u4")])


def load_ptscene(path: str) -> SceneBlob:
    raw = open(path, "rb").read()
    if path.endswith(".xz"):
        raw = lzma.decompress(raw)
    return parse_ptscene(raw)


def parse_ptscene(raw: bytes) -> SceneBlob:
    if raw[:4] != b"PTSC":
        raise ValueError("not a .ptscene container")
    (version,) = struct.unpack_from("<I", raw, 4)
    if version != 1:
        raise ValueError(f"unsupported .ptscene version {version}")
    off = 8

    def take(n):
        nonlocal off
        b = raw[off:off + n]
        if len(b) != n:
            raise ValueError("truncated .ptscene")
        off += n
        return b

    def u32():
        return struct.unpack("<I", take(4))[0]

    camera = np.frombuffer(take(CAMERA_BYTES), dtype=np.float32).copy()
    n_l = u32()
    lights = np.frombuffer(take(n_l * LIGHT_BYTES), dtype=np.float32).reshape(n_l, 8).copy()
    n_m = u32()
    mats = np.zeros(n_m, dtype=MATERIAL_DTYPE)
    for i in range(n_m):
        d, n, ior = struct.unpack("<iif", take(12))
        mats[i] = (d, n, ior, 0)
    n_meshes = u32()
    counts = np.frombuffer(take(4 * n_meshes), dtype=np.uint32).copy()
    total = int(counts.sum())
    faces = np.frombuffer(take(total * FACE_BYTES), dtype=np.uint8).reshape(total, FACE_BYTES).copy()
    n_t = u32()
    lut = None
    textures = []
    for _ in range(n_t):
        w, h, nc = struct.unpack("<iii", take(12))
        kind = u32()
        if kind == 0:
            t = np.frombuffer(take(w * h * nc * 4), dtype=np.float32).reshape(h, w, nc).copy()
        elif kind == 1:
            alpha_const = u32()
            (alpha_v,) = struct.unpack("<f", take(4))
            stored = 3 if (nc == 4 and alpha_const) else nc
            b = np.frombuffer(take(w * h * stored), dtype=np.uint8).reshape(h, w, stored)
            if lut is None:
                lut = stb_gamma_lut()
            t = np.empty((h, w, nc), dtype=np.float32)
            t[:, :, :stored] = lut[b]
            if stored != nc:
                t[:, :, 3] = np.float32(alpha_v)
        else:
            raise ValueError(f"unknown texture encoding {kind}")
        textures.append(np.ascontiguousarray(t))
    n = u32()
    cubemap_path = take(n).decode()
    size = u32()
    has = u32()
    cube = None
    if has:
        cube = np.frombuffer(take(6 * size * size * 16), dtype=np.float32).reshape(6, size, size, 4).copy()
    return SceneBlob(camera, lights, mats, counts, faces, textures, cubemap_path, size, cube)


# ---------------------------------------------------------------------------------------------
# indexed meshes (include/ptcore.h ptc_indexed_desc): the OBJ loader's output BEFORE Face[]
# ---------------------------------------------------------------------------------------------
class PtcIndexedMesh(C.Structure):
    _fields_ = [("indices", C.c_void_p), ("n_indices", C.c_uint64), ("material_ids", C.c_void_p)]


class PtcIndexedDesc(C.Structure):
    _fields_ = [("vertices", C.c_void_p), ("n_vertex_floats", C.c_uint64),
                ("normals", C.c_void_p), ("n_normal_floats", C.c_uint64),
                ("texcoords", C.c_void_p), ("n_texcoord_floats", C.c_uint64),
                ("meshes", C.POINTER(PtcIndexedMesh)), ("n_meshes", C.c_uint32),
                ("index_stride_ints", C.c_uint32),
                ("materials", C.c_void_p), ("n_materials", C.c_uint32),
                ("lights", C.c_void_p), ("n_lights", C.c_uint32),
                ("textures", C.c_void_p), ("n_textures", C.c_uint32)]


class IndexedMesh:
    """Attribute arrays + per-shape index records and material ids (tinyobj's output)."""

    def __init__(self, vertices, normals, texcoords, shapes, stride=3):
        self.vertices = np.ascontiguousarray(vertices, np.float32).reshape(-1)
        self.normals = np.ascontiguousarray(normals, np.float32).reshape(-1)
        self.texcoords = np.ascontiguousarray(texcoords, np.float32).reshape(-1)
        self.stride = int(stride)
        # shapes: list of (indices int32 (n, stride), material_ids int32 (n // 3,))
        self.shapes = [(np.ascontiguousarray(i, np.int32).reshape(-1, self.stride),
                        np.ascontiguousarray(m, np.int32).reshape(-1)) for i, m in shapes]

    @property
    def n_faces(self):
        return sum(i.shape[0] // 3 for i, _ in self.shapes)

    def to_desc(self, scene=None) -> PtcIndexedDesc:
        d = PtcIndexedDesc()
        d.vertices, d.n_vertex_floats = self.vertices.ctypes.data, self.vertices.size
        d.normals, d.n_normal_floats = self.normals.ctypes.data, self.normals.size
        d.texcoords, d.n_texcoord_floats = self.texcoords.ctypes.data, self.texcoords.size
        arr = (PtcIndexedMesh * max(1, len(self.shapes)))()
        for k, (idx, mids) in enumerate(self.shapes):
            arr[k].indices, arr[k].n_indices, arr[k].material_ids = idx.ctypes.data, idx.shape[0], mids.ctypes.data
        d.meshes, d.n_meshes, d.index_stride_ints = arr, len(self.shapes), self.stride
        self._keep = [arr]
        if scene is not None:  # materials / lights / textures of a SceneBlob
            sd = scene.to_desc()
            self._keep.append((scene, sd))
            d.materials, d.n_materials = sd.materials, sd.n_materials
            d.lights, d.n_lights = sd.lights, sd.n_lights
            d.textures, d.n_textures = sd.textures, sd.n_textures
        return d


def load_ptix(path) -> IndexedMesh:
    """Reads a .ptix file: the reference loader's tinyobj output (written by the fixture tooling, see
    tests/golden/make_fixtures.py) -- attribute arrays, index records, per-face material ids."""
    import struct
    with open(path, "rb") as f:
        raw = f.read()
    assert raw[:4] == b"PTIX" and struct.unpack_from("<I", raw, 4)[0] == 1
    off = 8
    arrs = []
    for _ in range(3):
        n = struct.unpack_from("<Q", raw, off)[0]
        arrs.append(np.frombuffer(raw, np.float32, n, off + 8).copy())
        off += 8 + 4 * n
    n_shapes, stride = struct.unpack_from("<II", raw, off)
    off += 8
    shapes = []
    for _ in range(n_shapes):
        ni = struct.unpack_from("<Q", raw, off)[0]
        idx = np.frombuffer(raw, np.int32, ni * stride, off + 8).copy().reshape(-1, stride)
        off += 8 + 4 * ni * stride
        nm = struct.unpack_from("<Q", raw, off)[0]
        mids = np.frombuffer(raw, np.int32, nm, off + 8).copy()
        off += 8 + 4 * nm
        shapes.append((idx, mids))
    return IndexedMesh(arrs[0], arrs[1], arrs[2], shapes, stride)


# ---------------------------------------------------------------------------------------------
# material / texture table input (include/ptcore.h ptc_mtl_entry, ptc_image)
# ---------------------------------------------------------------------------------------------
class PtcImage(C.Structure):
    _fields_ = [("w", C.c_int32), ("h", C.c_int32), ("nb_chan", C.c_int32), ("_pad", C.c_int32),
                ("data", C.c_void_p), ("name", C.c_char_p)]


class PtcMtlEntry(C.Structure):
    _fields_ = [("diffuse", C.c_float * 3), ("specular", C.c_float * 3), ("ior", C.c_float), ("_pad", C.c_uint32),
                ("diffuse_map", C.POINTER(PtcImage)), ("specular_map", C.POINTER(PtcImage)),
                ("normal_map", C.POINTER(PtcImage))]


@dataclass
class MtlSet:
    """Materials of one OBJ as tinyobj parsed them + the decoded images they name."""
    materials: list   # dicts: diffuse (3,), specular (3,), ior, names = (diffuse, specular, normal)
    images: dict      # name -> float32 array (h, w, nb_chan), or None when the file failed to load


def load_ptmt(path) -> MtlSet:
    """Reads a .ptmt file (the reference loader's tinyobj materials + stb-decoded images, written
    by the fixture tooling, tests/golden/make_fixtures.py)."""
    raw = open(path, "rb").read()
    assert raw[:4] == b"PTMT" and struct.unpack_from("<I", raw, 4)[0] == 1
    off = 8

    def u32():
        nonlocal off
        v = struct.unpack_from("<I", raw, off)[0]
        off += 4
        return v

    def text():
        nonlocal off
        n = u32()
        s = raw[off:off + n].decode()
        off += n
        return s

    mats = []
    for _ in range(u32()):
        f = np.frombuffer(raw, np.float32, 7, off).copy()
        off += 28
        mats.append({"diffuse": f[0:3], "specular": f[3:6], "ior": float(f[6]), "names": (text(), text(), text())})
    images = {}
    for _ in range(u32()):
        name = text()
        w, h, nc = struct.unpack_from("<iii", raw, off)
        off += 12
        if nc:
            images[name] = np.frombuffer(raw, np.float32, w * h * nc, off).copy().reshape(h, w, nc)
            off += 4 * w * h * nc
        else:
            images[name] = None
    return MtlSet(mats, images)
