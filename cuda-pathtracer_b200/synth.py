"""Procedural scenes in the reference's device data contract (``Face[]``/``Material[]``/
``Texture[]``/``LightProp[]``/``Camera``, cuda_opengl/include/scene/scene_data.h:31-133).

BASELINE.json config 4 names a "synthetic procedurally tessellated 10M-triangle mesh with
textured materials, divergence-heavy glossy BRDFs"; the reference ships no such asset, so
it is generated here, deterministically (integer hash, seed 0x5EED), and fed through the
same ``ptc_scene_desc`` as a loaded OBJ.  The per-face tangent follows the reference
loader (cuda_opengl/src/scene/scene.cpp:251-262) so the normal-mapping path sees the same
kind of input.  Host-side numpy only -- plumbing, not the product.
"""
from __future__ import annotations

import numpy as np

from .ptscene import MATERIAL_DTYPE, SceneBlob

SEED = 0x5EED


def _hash_u32(a: np.ndarray) -> np.ndarray:
    """Wang hash (the reference seeds its RNG with the same mixer, raytrace.cu:275-285)."""
    a = a.astype(np.uint32, copy=True)
    a = (a ^ np.uint32(61)) ^ (a >> np.uint32(16))
    a = a + (a << np.uint32(3))
    a = a ^ (a >> np.uint32(4))
    a = a * np.uint32(0x27D4EB2D)
    a = a ^ (a >> np.uint32(15))
    return a


def _hash01(a: np.ndarray) -> np.ndarray:
    return (_hash_u32(a).astype(np.float64) / 4294967296.0).astype(np.float32)


def camera_array(position, direction, fov_x=np.pi / 2, aperture=0.01, focus_dist=3.0,
                 speed=1.4) -> np.ndarray:
    d = np.asarray(direction, np.float64)
    d = d / np.linalg.norm(d)
    cam = np.zeros(16, np.float32)
    cam[0:3] = position
    cam[3:6] = d
    cam[6:9] = (1, 0, 0)
    cam[9:12] = (0, -1, 0)
    cam[12], cam[13], cam[14], cam[15] = fov_x, speed, aperture, focus_dist
    return cam


def faces_from_arrays(v, n, uv, material_ids) -> np.ndarray:
    """Packs per-face vertex (F,3,3), normal (F,3,3), uv (F,3,2), material id (F,) into
    ``scene::Face`` records (112 B) and computes the loader's per-face tangent."""
    v = np.asarray(v, np.float32)
    n = np.asarray(n, np.float32)
    uv = np.asarray(uv, np.float32)
    F = v.shape[0]
    rec = np.zeros((F, 28), np.float32)
    rec[:, 0:9] = v.reshape(F, 9)
    rec[:, 9:18] = n.reshape(F, 9)
    rec[:, 18:24] = uv.reshape(F, 6)
    e1 = v[:, 1] - v[:, 0]
    e2 = v[:, 2] - v[:, 0]
    d1 = uv[:, 1] - uv[:, 0]
    d2 = uv[:, 2] - uv[:, 0]
    with np.errstate(divide="ignore", invalid="ignore"):
        f = np.float32(1.0) / (d1[:, 0] * d2[:, 1] - d2[:, 0] * d1[:, 1])
        tan = f[:, None] * (d2[:, 1:2] * e1 - d1[:, 1:2] * e2)
    rec[:, 24:27] = tan.astype(np.float32)
    rec.view(np.uint32)[:, 27] = np.asarray(material_ids, np.uint32)
    return rec.view(np.uint8).reshape(F, 112)


def make_blob(faces, mesh_face_counts, materials, textures, lights, camera, cubemap=None) -> SceneBlob:
    mats = np.zeros(len(materials), MATERIAL_DTYPE)
    for i, (d, nm, ior) in enumerate(materials):
        mats[i] = (d, nm, ior, 0)
    lights = np.asarray(lights, np.float32).reshape(-1, 8)
    if cubemap is None:
        cubemap = constant_cubemap((19 / 255.0, 27 / 255.0, 35 / 255.0))
    return SceneBlob(np.asarray(camera, np.float32), lights, mats,
                     np.asarray(mesh_face_counts, np.uint32), np.ascontiguousarray(faces),
                     [np.ascontiguousarray(t, np.float32) for t in textures], "",
                     int(cubemap.shape[1]), np.ascontiguousarray(cubemap, np.float32))


def constant_cubemap(rgb) -> np.ndarray:
    """The reference's fallback when the cubemap image is missing: 1x1x6 constant colour
    (cuda_opengl/src/gpu_processor.cpp:75,99-105,128-132)."""
    c = np.zeros((6, 1, 1, 4), np.float32)
    c[..., 0:3] = np.asarray(rgb, np.float32)
    return c


def cube_texel_dirs(size: int) -> np.ndarray:
    """Unit directions through the texel centres of a cubemap, faces +x,-x,+y,-y,+z,-z in
    CUDA's cubemap convention (the layout texCubemap addresses, raytrace.cu:60,197)."""
    c = (np.arange(size, dtype=np.float64) + 0.5) / size * 2.0 - 1.0
    s, t = np.meshgrid(c, c, indexing="xy")  # s along a row, t down the rows
    one = np.ones_like(s)
    faces = [(one, -t, -s), (-one, -t, s), (s, one, t), (s, -one, -t), (s, -t, one), (-s, -t, -one)]
    d = np.stack([np.stack(f, axis=-1) for f in faces])
    return d / np.linalg.norm(d, axis=-1, keepdims=True)


def gradient_cubemap(size=32) -> np.ndarray:
    """Smooth sky-like cubemap (a function of direction only, so it is continuous across
    face borders); exercises the linear-filtered lookup of raytrace.cu:60,197."""
    d = cube_texel_dirs(size)
    c = np.zeros((6, size, size, 4), np.float32)
    up = 0.5 + 0.5 * d[..., 1]
    c[..., 0] = 0.15 + 0.55 * up + 0.1 * d[..., 0]
    c[..., 1] = 0.25 + 0.5 * up + 0.1 * d[..., 2]
    c[..., 2] = 0.45 + 0.45 * up
    return c


def procedural_texture(size, nb_chan, tex_seed, spec=None) -> np.ndarray:
    """Integer-hash texture: value noise at 1/8 resolution, nearest upsampled (``nb_chan`` 4:
    rgb + specular-mix alpha; 3: a tangent-space normal encoded like a decoded normal map)."""
    cells = max(1, size // 8)
    idx = (np.arange(cells * cells * nb_chan, dtype=np.uint32) + np.uint32(tex_seed * 7919 + SEED))
    base = _hash01(idx).reshape(cells, cells, nb_chan)
    if nb_chan == 3:
        base = np.float32(0.5) + (base - np.float32(0.5)) * np.float32(0.3)
        base[..., 2] = 1.0
    else:
        base[..., 0:3] = np.float32(0.25) + base[..., 0:3] * np.float32(0.7)
        base[..., 3] = np.float32(spec if spec is not None else 0.0)
    rep = size // cells
    return np.ascontiguousarray(np.repeat(np.repeat(base, rep, 0), rep, 1)[:size, :size])


SPEC_CYCLE = (0.0, 0.25, 0.5, 0.75, 0.9, 1.0)


def terrain_room(grid: int = 64, tex_size: int = 64, n_materials: int = 16, room: float = 8.0,
                 seed: int = SEED, refractive: bool = True) -> SceneBlob:
    """Closed box room (12 triangles, inward facing) around a ``grid x grid`` height field
    (2*grid*grid triangles, smooth normals, UVs tiled), ``n_materials`` materials cycling the
    specular mix over SPEC_CYCLE (two of them refractive, ior 1.5), procedural RGBA + normal
    textures, 3 sphere lights, DOF camera inside.  ``grid=2236`` gives 10 000 000 triangles."""
    g = int(grid)
    half = room / 2.0
    # height field on [-half, half]^2, y = smooth hills + hashed roughness
    xs = np.linspace(-half * 0.98, half * 0.98, g + 1, dtype=np.float32)
    X, Z = np.meshgrid(xs, xs, indexing="xy")
    vid = (np.arange((g + 1) * (g + 1), dtype=np.uint32) + np.uint32(seed)).reshape(g + 1, g + 1)
    rough = (_hash01(vid) - np.float32(0.5)) * np.float32(min(0.02, 0.5 * room / g))
    Y = (np.float32(0.35) * np.sin(X * np.float32(1.7)) * np.cos(Z * np.float32(1.3))
         + np.float32(0.15) * np.sin(X * np.float32(5.1) + Z * np.float32(3.7)) + rough).astype(np.float32)
    P = np.stack([X, Y, Z], axis=-1)  # (g+1, g+1, 3)
    # smooth normals from central differences
    dx = np.zeros_like(P)
    dz = np.zeros_like(P)
    dx[:, 1:-1] = P[:, 2:] - P[:, :-2]
    dx[:, 0] = P[:, 1] - P[:, 0]
    dx[:, -1] = P[:, -1] - P[:, -2]
    dz[1:-1] = P[2:] - P[:-2]
    dz[0] = P[1] - P[0]
    dz[-1] = P[-1] - P[-2]
    N = np.cross(dz, dx)
    N /= np.linalg.norm(N, axis=-1, keepdims=True)
    tile = np.float32(max(1.0, g / 35.0))  # UVs tiled (x64 at 10 M triangles)
    U = ((X + half) / room * tile).astype(np.float32)
    V = ((Z + half) / room * tile).astype(np.float32)
    UVg = np.stack([U, V], axis=-1)

    def quad(a, i0, j0, i1, j1, i2, j2):
        return np.stack([a[i0:i0 + g, j0:j0 + g], a[i1:i1 + g, j1:j1 + g], a[i2:i2 + g, j2:j2 + g]], axis=2)

    # two CCW-from-above triangles per cell: (i,j),(i+1,j),(i,j+1) and (i+1,j),(i+1,j+1),(i,j+1)
    # rows = z index, cols = x index; facing +y requires cross(e1,e2).y > 0
    tv = np.concatenate([quad(P, 0, 0, 1, 0, 0, 1).reshape(-1, 3, 3),
                         quad(P, 1, 0, 1, 1, 0, 1).reshape(-1, 3, 3)])
    tn = np.concatenate([quad(N, 0, 0, 1, 0, 0, 1).reshape(-1, 3, 3),
                         quad(N, 1, 0, 1, 1, 0, 1).reshape(-1, 3, 3)])
    tuv = np.concatenate([quad(UVg, 0, 0, 1, 0, 0, 1).reshape(-1, 3, 2),
                          quad(UVg, 1, 0, 1, 1, 0, 1).reshape(-1, 3, 2)])
    cell = np.tile(np.arange(g * g, dtype=np.uint32), 2)
    patch = max(1, g // 16)
    ci, cj = (cell // g) // patch, (cell % g) // patch
    tm = (_hash_u32(ci * np.uint32(131) + cj + np.uint32(seed)) % np.uint32(n_materials - 1)) + np.uint32(1)

    # room: 6 inward-facing quads, material 0
    lo, hi = -half, half
    y0, y1 = np.float32(-1.0), np.float32(room * 0.6)
    c = np.array([[lo, y0, lo], [hi, y0, lo], [hi, y0, hi], [lo, y0, hi],
                  [lo, y1, lo], [hi, y1, lo], [hi, y1, hi], [lo, y1, hi]], np.float32)
    quads = [(0, 3, 2, 1, (0, 1, 0)), (4, 5, 6, 7, (0, -1, 0)), (0, 1, 5, 4, (0, 0, 1)),
             (3, 7, 6, 2, (0, 0, -1)), (0, 4, 7, 3, (1, 0, 0)), (1, 2, 6, 5, (-1, 0, 0))]
    rv, rn, ruv = [], [], []
    for a, b, cc, d, nrm in quads:
        for tri in ((a, b, cc), (a, cc, d)):
            p = c[list(tri)]
            # make the geometric normal agree with the inward normal (single-sided faces, Q1)
            if np.dot(np.cross(p[1] - p[0], p[2] - p[0]), nrm) < 0:
                p = p[[0, 2, 1]]
            rv.append(p)
            rn.append(np.tile(np.asarray(nrm, np.float32), (3, 1)))
            ruv.append(np.array([[0, 0], [3, 0], [0, 3]], np.float32) + np.float32(0.01 * len(rv)))
    rv, rn, ruv = np.array(rv), np.array(rn), np.array(ruv)

    v = np.concatenate([rv, tv])
    n = np.concatenate([rn, tn])
    uv = np.concatenate([ruv, tuv])
    mid = np.concatenate([np.zeros(len(rv), np.uint32), tm])
    faces = faces_from_arrays(v, n, uv, mid)

    textures, materials = [], []
    for m in range(n_materials):
        spec = SPEC_CYCLE[m % len(SPEC_CYCLE)]
        textures.append(procedural_texture(tex_size, 4, 2 * m, spec))
        d_id = len(textures) - 1
        n_id = -1
        if m % 2 == 1:
            textures.append(procedural_texture(tex_size, 3, 2 * m + 1))
            n_id = len(textures) - 1
        ior = 1.5 if (refractive and m in (5, 11) and m < n_materials) else 1.0
        materials.append((d_id, n_id, ior))
    lights = [[1.0, 1.0, 1.0, -half * 0.5, y1 * 0.7, -half * 0.4, 2.0, 0.3],
              [1.0, 0.9, 0.8, half * 0.5, y1 * 0.6, half * 0.3, 2.0, 0.3],
              [0.8, 0.9, 1.0, 0.0, y1 * 0.8, half * 0.6, 2.0, 0.3]]
    cam = camera_array((-half * 0.7, 2.0, half * 0.7), (0.62, -0.45, -0.7), aperture=0.01, focus_dist=3.5)
    return make_blob(faces, [len(rv), len(tv)], materials, textures, lights, cam)
