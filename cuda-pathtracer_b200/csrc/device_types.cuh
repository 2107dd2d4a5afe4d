// device_types.cuh -- device-side views shared by the kernels and the C ABI.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "ptcore.h"

namespace ptc {

constexpr float kMaxDist = 100000.0f; // MAX_DIST, intersection.cuh:165
constexpr int kNodeVec = 5;           // uint4 per Node8
constexpr int kTriVec = 3;            // float4 per TriRecord
constexpr int kShadeVec = 5;          // float4 per ShadeRecord

struct DTexture // arena-resident copy of scene::Texture (scene_data.h:31-37)
{
  const float* data;
  int32_t w, h, nb_chan, _pad;
};

struct DScene
{
  const uint4* nodes;
  const float4* tris;
  const float4* shade;
  const ptc_material* materials;
  const DTexture* textures;
  const ptc_light* lights;
  uint32_t n_faces;
  uint32_t n_lights;
  uint32_t n_nodes;
  uint32_t n_tris;
};

// Wavefront path state, structure of arrays over path slots (slot = pixel index
// inside the rendered band).  104 bytes per path.
struct PathState
{
  float4* ray_o;  // origin.xyz, w = carried specular_col (reference quirk Q3)
  float4* ray_d;  // dir.xyz (not normalised after a specular mix, Q8)
  uint4* hit;     // t bits, hit id (PTC_HIT_MISS = 0xffffffff), barycentric u, v bits
  float4* thr;    // throughput.xyz
  float4* acc;    // accumulated radiance.xyz of this sample
  uint4* rng_a;   // XORWOW v[0..3]
  uint2* rng_b;   // XORWOW v[4], d
};

struct Counters // device mirror of the ray-count part of ptc_counters
{
  unsigned long long paths;
  unsigned long long segments;
  unsigned long long traced;
  unsigned long long rays_per_bounce[16];
  unsigned long long nodes_visited;
  unsigned long long tris_tested;
};

struct FrameParams
{
  ptc_camera cam;
  uint32_t width, height;      // full image
  uint32_t row_begin, row_end; // band rendered by this context
  uint32_t hash_seed;          // WangHash(frame_nb), raytrace.cu:321
  int32_t frame_nb;
  int32_t max_bounces;         // 1 + (static_samples + 1), raytrace.cu:66
  int32_t is_static;           // !moved
  uint32_t post_id;
};

} // namespace ptc
