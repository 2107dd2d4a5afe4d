"""Pins the CPU oracle (oracle/pt_oracle.cpp) -- `-m "not gpu"`.

The reference has no tests of its own, so the pin is the reference ITSELF:
tests/golden/*.npz and rng_kat.json were produced by the reference's own raytrace.cu /
intersection.cuh / curand_kernel.h compiled for host cores (tests/golden/make_fixtures.py,
oracle/build_ref.py).  The oracle must reproduce them bit for bit; where oracle/_ref was
built (the container that has /root/reference) the two are additionally compared live on
scenes and settings the frozen fixtures do not cover.
"""
import json
import os

import numpy as np
import pytest

from pt_testlib import (GOLDEN, aimed_rays, load_scene, oracle_bvh_intersect, oracle_render_stats, pkg, random_rays,
                        soup_scene, triangle_soups)


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


@pytest.fixture(scope="module")
def kat():
    return json.load(open(os.path.join(GOLDEN, "rng_kat.json")))


def test_wang_hash_kat(oracle, kat):
    # raytrace.cu:275-285
    for k, v in kat["wang_hash"].items():
        assert oracle.wang_hash(int(k)) == v


def test_xorwow_uniform_kat(oracle, kat):
    # curand_kernel.h:772-799,863-874 + curand_uniform.h:69-72 as called at raytrace.cu:234-235
    for seed, vals in kat["uniforms"].items():
        want = np.array([float.fromhex(v) for v in vals], np.float32)
        got = oracle.uniforms(int(seed), len(want))
        assert (bits(got) == bits(want)).all(), seed
        assert (got > 0).all() and (got <= 1).all()


@pytest.mark.parametrize("tag", ["plain", "tex"])
def test_radiance_matches_reference_fixture(oracle, tag):
    sc = load_scene(f"indoor_{tag}")
    z = np.load(os.path.join(GOLDEN, f"ref_indoor_{tag}_160x90.npz"))
    w, h = 160, 90
    for key, frame, bounces, moved in [("radiance_b3_f1", 1, 3, False), ("radiance_b4_f1", 1, 4, False),
                                       ("radiance_b3_f7", 7, 3, False), ("radiance_preview", 1, 3, True)]:
        got = oracle.radiance(sc, w, h, frame, bounces, moved=moved)
        assert (bits(got) == bits(z[key])).all(), key


@pytest.mark.parametrize("tag", ["plain", "tex"])
def test_kernel_accumulation_matches_reference_fixture(oracle, tag):
    # raytrace.cu:248-270: clamp, accumulate bottom-up, tonemap, gamma, RGBA8 pack
    sc = load_scene(f"indoor_{tag}")
    z = np.load(os.path.join(GOLDEN, f"ref_indoor_{tag}_160x90.npz"))
    w, h = 160, 90
    accum = np.zeros((h, w, 3), np.float32)
    for k in range(1, 5):
        rgba = oracle.kernel(sc, w, h, k, 3, accum)
    assert (bits(accum) == bits(z["accum_b3_f1to4"])).all()
    assert (rgba == z["rgba_b3_f1to4"]).all()


@pytest.mark.parametrize("tag", ["plain", "tex"])
def test_primary_hits_match_reference_fixture(oracle, tag):
    sc = load_scene(f"indoor_{tag}")
    z = np.load(os.path.join(GOLDEN, f"ref_indoor_{tag}_160x90.npz"))
    o, d = oracle.primary_rays(sc, 160, 90, 1)
    ids, t = oracle.intersect(sc, o, d)
    assert (ids.reshape(90, 160) == z["primary_ids_f1"]).all()
    assert (bits(t.reshape(90, 160)) == bits(z["primary_t_f1"])).all()


def test_render_stats_counts(oracle):
    """segments = radiance-loop iterations; traced <= segments; first-hit ids = primary ids."""
    sc = load_scene("indoor_plain")
    z = np.load(os.path.join(GOLDEN, "ref_indoor_plain_160x90.npz"))
    rad, first, seg, tr = oracle_render_stats(oracle, sc, 160, 90, 1, 3)
    assert (bits(rad) == bits(z["radiance_b3_f1"])).all()
    assert (first == z["primary_ids_f1"]).all()
    assert 160 * 90 <= tr <= seg <= 160 * 90 * 3


# -- live against the reference's own code on host cores (only where oracle/_ref exists) ----
LIVE = [("color_sample_plain", 64, 36), ("island_plain", 64, 36), ("sss_crate_plain", 48, 27),
        ("sss_crate_plain+camera", 48, 27), ("indoor_tex", 80, 45)]


def _with_cubemap(sc):
    if sc.cubemap is None:
        sc.cubemap = pkg().synth.gradient_cubemap(16)
        sc.cubemap_size = 16
    return sc


@pytest.mark.parametrize("name,w,h", LIVE)
def test_oracle_equals_reference_live(oracle, ref_cpu, name, w, h):
    if name.endswith("+camera"):
        # sss_crate.scene's legacy camera line parses to fov 0 -> NaN rays (kept as a case of its
        # own above); this variant looks at the crate (translucent material, emission-40 light)
        import copy
        sc = _with_cubemap(copy.copy(load_scene(name[:-7])))
        sc.camera = pkg().synth.camera_array((-2.7, 2.06, 2.52), (0.62, -0.348, -0.7), aperture=0.02,
                                             focus_dist=3.555)
    else:
        sc = _with_cubemap(load_scene(name))
    for frame, bounces in [(1, 3), (3, 8), (2, 1)]:
        a = oracle.radiance(sc, w, h, frame, bounces)
        b = ref_cpu.radiance(sc, w, h, frame, bounces)
        assert (bits(a) == bits(b)).all(), (name, frame, bounces)
    acc_a = np.zeros((h, w, 3), np.float32)
    acc_b = np.zeros((h, w, 3), np.float32)
    for post in range(4):
        ra = oracle.kernel(sc, w, h, post + 1, 3, acc_a, moved=(post == 2), post_id=post)
        rb = ref_cpu.kernel(sc, w, h, post + 1, 3, acc_b, moved=(post == 2), post_id=post)
        assert (ra == rb).all(), (name, post)
    assert (bits(acc_a) == bits(acc_b)).all()


def test_oracle_equals_reference_live_synthetic(oracle, ref_cpu):
    """Textured, normal-mapped, glossy and refractive materials on a procedural mesh."""
    sc = pkg().synth.terrain_room(grid=12, tex_size=16)
    a = oracle.radiance(sc, 64, 36, 1, 6)
    b = ref_cpu.radiance(sc, 64, 36, 1, 6)
    assert (bits(a) == bits(b)).all()
    o, d = random_rays(sc, 5000, seed=3)
    ia, ta = oracle.intersect(sc, o, d)
    ib, tb = ref_cpu.intersect(sc, o, d)
    assert (ia == ib).all() and (bits(ta) == bits(tb)).all()


def test_rng_live(oracle, ref_cpu):
    for seed in [5, 12345, 0x80000000, 0xFFFFFFFE]:
        assert (bits(oracle.uniforms(seed, 64)) == bits(ref_cpu.uniforms(seed, 64))).all()
        assert oracle.wang_hash(seed) == ref_cpu.wang_hash(seed)


# -- O2: the oracle traversing the product's BVH must give O1's (brute force) answers ------
@pytest.mark.parametrize("name", ["indoor_plain", "color_sample_plain", "island_plain", "sss_crate_plain"])
def test_bvh_traversal_equals_brute_force(oracle, name):
    sc = load_scene(name)
    hb = pkg().host_bvh(sc)
    o, d = random_rays(sc, 40000, seed=7)
    i1, t1 = oracle.intersect(sc, o, d)
    i2, t2, nn, nt = oracle_bvh_intersect(oracle, sc, hb, o, d)
    assert (i1 == i2).all()
    assert (bits(t1) == bits(t2)).all()
    assert nt < sc.n_faces * len(o)  # the BVH actually culls


def test_bvh_traversal_synthetic_and_axis_aligned_rays(oracle):
    sc = pkg().synth.terrain_room(grid=40, tex_size=8)
    hb = pkg().host_bvh(sc)
    o, d = random_rays(sc, 20000, seed=11)
    # axis-aligned and zero-component directions stress the guarded reciprocal in the slab test
    d[:3000] = 0
    d[np.arange(3000), np.arange(3000) % 3] = np.where(np.arange(3000) % 2, 1.0, -1.0)
    d[3000:4000, 1] = 0
    d[3000:4000] /= np.linalg.norm(d[3000:4000], axis=1, keepdims=True)
    i1, t1 = oracle.intersect(sc, o, d)
    i2, t2, _, _ = oracle_bvh_intersect(oracle, sc, hb, o, d)
    assert (i1 == i2).all()
    assert (bits(t1) == bits(t2)).all()


def test_bvh_traversal_with_coincident_triangles(oracle):
    """Many copies of one triangle (one centroid, no separating plane): the positional split of the
    builder must lose none of them, and the FIRST copy must win the tie as in the reference's loop
    (intersection.cuh:183-184)."""
    base = pkg().synth.terrain_room(grid=12, tex_size=8)
    faces = np.concatenate([np.repeat(base.faces[5:7], 300, axis=0), base.faces])
    mats = [(int(m[0]), int(m[1]), float(m[2])) for m in base.materials]
    sc = pkg().synth.make_blob(faces, [faces.shape[0]], mats, base.textures, base.lights, base.camera)
    hb = pkg().host_bvh(sc)
    assert sorted(hb["face_ids"].tolist()) == list(range(sc.n_faces))
    o, d = random_rays(sc, 20000, seed=3)
    i1, t1 = oracle.intersect(sc, o, d)
    i2, t2, _, _ = oracle_bvh_intersect(oracle, sc, hb, o, d)
    assert (i1 == i2).all()
    assert (bits(t1) == bits(t2)).all()
    assert ((i1 >= 0) & (i1 < 600)).any()  # the copies are hit, and then only through their first copy
    assert set(np.unique(i1[(i1 >= 0) & (i1 < 600)]).tolist()) <= {0, 300}


@pytest.mark.parametrize("name,verts", list(triangle_soups()), ids=[n for n, _ in triangle_soups()])
def test_bvh_traversal_on_triangle_soups(oracle, name, verts):
    """O2 (the oracle walking the product's BVH) == O1 (the reference's loop over all faces,
    intersection.cuh:179-196) on geometry that stresses the builder and the quantised boxes:
    ids and t bit for bit, for random rays and for rays aimed at vertices / edges / centroids."""
    sc = soup_scene(verts)
    hb = pkg().host_bvh(sc)
    assert sorted(hb["face_ids"].tolist()) == list(range(sc.n_faces))
    rng = np.random.default_rng(len(name))
    o1, d1 = random_rays(sc, 3000, seed=len(name))
    o2, d2 = aimed_rays(sc, 5000, rng)
    o, d = np.concatenate([o1, o2]), np.concatenate([d1, d2])
    i1, t1 = oracle.intersect(sc, o, d)
    i2, t2, _, _ = oracle_bvh_intersect(oracle, sc, hb, o, d)
    assert (i1 == i2).all(), f"{name}: {int((i1 != i2).sum())} rays disagree"
    assert (bits(t1) == bits(t2)).all()
    assert (i1[3000:] >= 0).mean() > 0.1  # the aimed rays do hit (front faces only, intersection.cuh:110)
