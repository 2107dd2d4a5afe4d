// device_types.cuh -- device-side views shared by the kernels and the C ABI.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "ptcore.h"

namespace ptc {

constexpr float kMaxDist = 100000.0f; // MAX_DIST, intersection.cuh:165
constexpr int kNodeVec = 8;           // uint4 per Node8 (128 B)
constexpr int kTriVec = 3;            // float4 per TriRecord
constexpr int kShadeVec = 5;          // float4 per ShadeRecord
// traversal stack per ray: node groups grow from the bottom (one per level of the wide tree:
// ptc_upload_scene rejects trees deeper than kStackSize - kTriBacklog), deferred triangle
// groups from the top (at most kTriBacklog; a lane with that many pending groups stops
// descending until it has tested some triangles)
constexpr int kStackSize = 40;
constexpr int kTriBacklog = 8;
// shared-memory copy of a small scene's BVH (SharedSceneStack, traverse.cuh): node / triangle record
// strides, and the deferred-triangle-group backlog of that form (8 KB less stack per CTA)
constexpr uint32_t kSmemNodeBytes = 144, kSmemTriBytes = 48;
constexpr int kSmemSceneBacklog = 4;
constexpr int kBlock = 256;           // threads per CTA of the path kernels (8 warps)
constexpr int kTileW = 8, kTileH = 4; // pixels one warp covers in the first wavefront stage

struct DTexture // arena-resident copy of scene::Texture (scene_data.h:31-37)
{
  const float* data;
  int32_t w, h, nb_chan, _pad;
};

// What shading needs to know about a material in ONE table entry (3 x LDG.128): the reference
// walks Material -> Texture[diffuse_spec_map] -> texels and Texture[normal_map] -> texels
// (intersection.cuh:216-242), i.e. two dependent loads before the texel address is known.
struct __align__(16) DMaterial
{
  const float* diffuse; // RGBA float texels (rgb = diffuse, a = specular mix)
  const float* normal;  // RGB float texels, or null
  int32_t dw, dh;       // diffuse/specular map size
  int32_t nw, nh;       // normal map size
  float ior;
  uint32_t _pad[3];
};
static_assert(sizeof(DMaterial) == 48, "DMaterial is read as three 16-byte loads");

struct DScene
{
  const uint4* nodes;
  const float4* tris;
  const float4* shade;
  const DMaterial* mats;
  const ptc_light* lights;
  uint32_t n_faces;
  uint32_t n_lights;
  uint32_t n_nodes;
  uint32_t n_tris;
};

// One compacted ray queue: structure of arrays over queue slots, 64 bytes per live path,
// written once by the stage that produced the ray and read once by the next stage.
//   o   : origin.xyz, w = carried specular_col (reference quirk Q3)
//   d   : dir.xyz (not renormalised after a specular mix, Q8), w = bits: RNG draws so far
//   thr : throughput.xyz, w = bits: band-local pixel index
//   acc : radiance of this sample so far .xyz
// The XORWOW state is NOT stored: it is a pure function of (pixel, frame, draws so far) and
// is replayed from the seed (24 B read + 24 B written per ray per bounce saved for ~8 integer
// instructions per replayed draw).
struct RayQueue
{
  float4* o;
  float4* d;
  float4* thr;
  float4* acc;
};

struct Counters // device mirror of the ray-count part of ptc_counters
{
  unsigned long long paths;
  unsigned long long segments;
  unsigned long long traced;
  unsigned long long rays_per_bounce[16];
  unsigned long long nodes_visited;
  unsigned long long tris_tested;
  // statistics build of the extend kernel only (ptc_profile.extend_lanes): warp iterations of the
  // traversal loop, and summed over them the lanes that held a ray, ran the node test, ran a
  // first / a second triangle test
  unsigned long long ext_iters, ext_have, ext_node, ext_tri1, ext_tri2;
};

struct FrameParams
{
  ptc_camera cam;
  uint32_t width, height;      // full image
  uint32_t row_begin;          // contiguous band: first global row (strip_stride == 1)
  uint32_t band_rows;          // rows rendered by this call
  uint32_t strip_rows;         // interleaved bands: rows per strip ...
  uint32_t strip_stride;       // ... every strip_stride-th strip ...
  uint32_t strip_phase;        // ... starting at strip strip_phase (1, x, 0 = contiguous)
  uint32_t part_row0;          // the launches of this pipeline generate band-local rows
  uint32_t part_rows;          // [part_row0, part_row0 + part_rows) only (a frame may run as two halves)
  uint32_t tiles_x;            // ceil(width / kTileW)
  uint32_t hash_seed;          // WangHash(frame_nb), raytrace.cu:321
  int32_t frame_nb;
  int32_t max_bounces;         // 1 + (static_samples + 1), raytrace.cu:66
  int32_t is_static;           // !moved
  uint32_t post_id;
};

// band-local row -> row of the full image
__host__ __device__ inline uint32_t
global_row(const FrameParams& fp, uint32_t lrow)
{
  if (fp.strip_stride <= 1)
    return fp.row_begin + lrow;
  const uint32_t k = lrow / fp.strip_rows;
  return (fp.strip_phase + k * fp.strip_stride) * fp.strip_rows + lrow % fp.strip_rows;
}

struct FrameOut
{
  float* accum;         // float3 per band pixel, rows bottom-up inside the band (raytrace.cu:252)
  uint32_t* out_rgba;   // RGBA8 per band pixel, row 0 = top; may be null
  float* radiance_out;  // optional un-clamped radiance of this sample, top-down; may be null
  uint32_t out_full;    // out_rgba is a FULL width x height image (PTC_OUT_FULL_IMAGE): a pixel goes to
                        // its GLOBAL row -- the image may live on another GPU (peer / IPC mapping), which
                        // is how image tiles are assembled without a collective
};

} // namespace ptc
