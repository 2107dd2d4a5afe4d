// flatten.cpp -- indexed mesh -> scene::Face records, host side, parallel.
//
// Replaces the de-indexing loop of upload_meshes (cuda_opengl/src/scene/scene.cpp:202-283): for
// every index triple of every shape it gathers position / normal / texcoord into a 112-byte Face,
// copies the per-face material id and computes ONE tangent per face from the edges and the UV
// deltas (scene.cpp:250-261).  The reference does this serially, one std::vector<Face> and one
// cudaMalloc + cudaMemcpy per shape; at 10 M triangles that loop, not the BVH build, is the load
// time.  Here all faces of all shapes are one index space split over the worker threads, written
// straight into the caller's (or the upload's) single Face array.
//
// Bit parity with the reference's Face[]: every tangent component is evaluated with the
// reference's expression order, one rounding per operation (this file is compiled with
// -ffp-contract=off: no fused multiply-add may replace `a * b - c * d`).
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "bvh.h"
#include "ptcore.h"

namespace {

unsigned
flatten_threads(uint64_t n_faces)
{
  if (n_faces < 16384)
    return 1;
  if (const char* e = getenv("PTC_BUILD_THREADS"))
    return (unsigned)(atoi(e) > 0 ? atoi(e) : 1);
  const unsigned hw = std::thread::hardware_concurrency();
  return hw ? (hw < 32u ? hw : 32u) : 1u;
}

// One face (scene.cpp:236-261).  Returns false when an index points outside its attribute array
// (the reference reads out of bounds there: an OBJ without `vn` / `vt` lines has index -1).
inline bool
flatten_one(const ptc_indexed_desc* d, const ptc_indexed_mesh& m, uint64_t f, uint32_t stride, ptc_face& out)
{
  const uint64_t nv = d->n_vertex_floats / 3, nn = d->n_normal_floats / 3, nt = d->n_texcoord_floats / 2;
  for (int v = 0; v < 3; ++v) {
    const int32_t* idx = m.indices + (3 * f + v) * stride;
    const int64_t iv = idx[0], in = idx[1], it = idx[2];
    if (iv < 0 || (uint64_t)iv >= nv || in < 0 || (uint64_t)in >= nn || it < 0 || (uint64_t)it >= nt)
      return false;
    for (int a = 0; a < 3; ++a) {
      out.vertices[v][a] = d->vertices[3 * iv + a];
      out.normals[v][a] = d->normals[3 * in + a];
    }
    out.texcoords[v][0] = d->texcoords[2 * it];
    out.texcoords[v][1] = d->texcoords[2 * it + 1];
  }
  out.material_id = (uint32_t)m.material_ids[f];
  // one tangent for the whole face, scene.cpp:250-261
  float e1[3], e2[3];
  for (int a = 0; a < 3; ++a) {
    e1[a] = out.vertices[1][a] - out.vertices[0][a];
    e2[a] = out.vertices[2][a] - out.vertices[0][a];
  }
  const float du1 = out.texcoords[1][0] - out.texcoords[0][0], dv1 = out.texcoords[1][1] - out.texcoords[0][1];
  const float du2 = out.texcoords[2][0] - out.texcoords[0][0], dv2 = out.texcoords[2][1] - out.texcoords[0][1];
  const float fdet = 1.0f / (du1 * dv2 - du2 * dv1);
  for (int a = 0; a < 3; ++a)
    out.tangent[a] = fdet * (dv2 * e1[a] - dv1 * e2[a]);
  return true;
}

} // namespace

extern "C" {

uint64_t
ptc_indexed_face_count(const ptc_indexed_desc* d)
{
  if (!d)
    return 0;
  uint64_t n = 0;
  for (uint32_t m = 0; m < d->n_meshes; ++m)
    n += d->meshes[m].n_indices / 3; // scene.cpp:222: nb_faces = nb_indices / 3
  return n;
}

int
ptc_flatten_faces(const ptc_indexed_desc* d, ptc_face* faces_out, uint64_t capacity, uint32_t* mesh_face_counts)
{
  if (!d || (!d->meshes && d->n_meshes))
    return PTC_ERR_INVALID;
  const uint32_t stride = d->index_stride_ints ? d->index_stride_ints : 3u;
  if (stride < 3)
    return PTC_ERR_INVALID;
  // prefix of face counts: face g of the scene = face g - first[m] of mesh m
  std::vector<uint64_t> first(d->n_meshes + 1, 0);
  for (uint32_t m = 0; m < d->n_meshes; ++m) {
    const uint64_t nf = d->meshes[m].n_indices / 3;
    if (nf && (!d->meshes[m].indices || !d->meshes[m].material_ids))
      return PTC_ERR_INVALID;
    if (nf > 0xFFFFFFFFull)
      return PTC_ERR_INVALID;
    first[m + 1] = first[m] + nf;
    if (mesh_face_counts)
      mesh_face_counts[m] = (uint32_t)nf;
  }
  const uint64_t total = first[d->n_meshes];
  if (total > capacity || (total && !faces_out))
    return PTC_ERR_INVALID;
  if (total && (!d->vertices || !d->normals || !d->texcoords))
    return PTC_ERR_INVALID;
  const unsigned nt = flatten_threads(total);
  std::vector<int> bad(nt, 0);
  auto work = [&](unsigned t) {
    const uint64_t lo = total * t / nt, hi = total * (t + 1) / nt;
    uint32_t m = 0;
    while (m + 1 < d->n_meshes && first[m + 1] <= lo)
      ++m;
    for (uint64_t g = lo; g < hi; ++g) {
      while (g >= first[m + 1])
        ++m;
      if (!flatten_one(d, d->meshes[m], g - first[m], stride, faces_out[g]))
        bad[t] = 1;
    }
  };
  std::vector<std::thread> th;
  for (unsigned t = 1; t < nt; ++t)
    th.emplace_back(work, t);
  work(0);
  for (std::thread& x : th)
    x.join();
  for (int b : bad)
    if (b)
      return PTC_ERR_INVALID;
  return PTC_OK;
}

int
ptc_upload_scene_indexed(ptc_context* ctx, const ptc_indexed_desc* d, int32_t* scene_id)
{
  if (!ctx || !d || !scene_id)
    return PTC_ERR_INVALID;
  const uint64_t n = ptc_indexed_face_count(d);
  // (not zero-filled: 1.12 GB at 10 M triangles, written once by the flattening threads)
  ptc::RawVector<ptc_face> faces(n);
  std::vector<uint32_t> counts(d->n_meshes);
  const int rc = ptc_flatten_faces(d, faces.data(), n, counts.data());
  if (rc)
    return rc; // an index outside its attribute array (e.g. an OBJ without normals or texcoords)
  ptc_scene_desc s;
  memset(&s, 0, sizeof(s));
  s.faces = faces.data();
  s.mesh_face_counts = counts.data();
  s.n_meshes = d->n_meshes;
  s.materials = d->materials, s.n_materials = d->n_materials;
  s.lights = d->lights, s.n_lights = d->n_lights;
  s.textures = d->textures, s.n_textures = d->n_textures;
  return ptc_upload_scene(ctx, &s, scene_id);
}

} // extern "C"
