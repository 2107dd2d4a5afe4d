// bvh.h -- host/device-shared layouts of the acceleration structure and the
// per-triangle records, plus the host builder's interface.
//
// The reference has no acceleration structure: intersect() loops over every face
// (cuda_opengl/include/shaders/intersection.cuh:179-196).  What this structure must
// reproduce is the reference's ANSWER: the closest front-facing hit, first face
// index winning ties (intersection.cuh:110,183-184).
#pragma once

#include <stdint.h>

#include <memory>
#include <utility>
#include <vector>

#include "ptcore.h"

namespace ptc {

// 8-wide compressed node, 128 bytes, 16-byte aligned (8 x 128-bit loads).
// Child boxes are quantised to 8 bits per plane relative to the node origin,
//   plane_a = origin_a + q * 2^(e_a - 127)        (e_a << 23 are the float bits of the scale)
// rounded outwards, so decoded boxes contain the true ones.  The 8-bit q values are stored
// as IEEE half floats (exact for 0..255): the kernel turns them into fp32 with one
// HADD2.F32 on the FMA pipe, whereas unpacking bytes costs an ALU-pipe PRMT (or an XU-pipe
// I2F) each -- and the box test is ALU-pipe bound (48 planes per node).
//
// Slot s (0..7) is either an internal child (bit s of imask; children of a node are
// contiguous from child_base in slot order), a leaf of 1..3 triangles, or empty.  The
// triangles of leaf slot s own bits 3s..3s+2 of the node's 24-bit triangle space; valid24
// marks the ones that exist and the record of bit b is
//   tris[tri_base + popcount(valid24 & ((1 << b) - 1))].
// Slots are assigned by octant (bvh_builder.cpp) so that (s ^ (7 - ray_octant)) orders the
// children far-to-near.
struct alignas(16) Node8
{
  float origin[3];
  uint8_t e[3];
  uint8_t imask;       // bit s: slot s holds an internal child
  uint32_t child_base; // node index of the first internal child
  uint32_t tri_base;   // index of this node's first triangle record
  uint32_t valid24;    // bit 3s+j: leaf slot s has a j-th triangle
  uint32_t occupied;   // bit s: slot s is not empty (diagnostic; the kernel needs imask/valid24 only)
#if defined(PTC_NODE_LAYOUT) && PTC_NODE_LAYOUT == 1
  // [axis][0 = lower planes, 1 = upper planes][slot]: the 32 bytes of one axis are one load
  uint16_t planes[3][2][8];
  uint16_t& lo_q(int axis, int slot) { return planes[axis][0][slot]; }
  uint16_t& hi_q(int axis, int slot) { return planes[axis][1][slot]; }
#else
  uint16_t lo[3][8];   // half-float q of the lower planes, [axis][slot]
  uint16_t hi[3][8];   // half-float q of the upper planes
  uint16_t& lo_q(int axis, int slot) { return lo[axis][slot]; }
  uint16_t& hi_q(int axis, int slot) { return hi[axis][slot]; }
#endif
};
static_assert(sizeof(Node8) == 128, "Node8 must be 128 bytes");

// Intersection record, 48 bytes (3 x 128-bit loads).  Edges are precomputed with the
// same single fp32 subtraction the reference performs per ray
// (intersection.cuh:106-107), so Moeller-Trumbore sees identical operands.
struct alignas(16) TriRecord
{
  float v0[3];
  float e1[3]; // v1 - v0
  float e2[3]; // v2 - v0
  uint32_t face_id; // global face index (mesh order) -- the tie-break key
  uint32_t _pad[2];
};
static_assert(sizeof(TriRecord) == 48, "TriRecord must be 48 bytes");

// Shading record, 80 bytes, indexed by global face id (5 x 128-bit loads).
struct alignas(16) ShadeRecord
{
  float n[3][3];  // vertex normals
  float uv[3][2]; // vertex texcoords
  float tangent[3];
  uint32_t material_id;
};
static_assert(sizeof(ShadeRecord) == 80, "ShadeRecord must be 80 bytes");

struct BuildStats
{
  float sah_cost = 0.f;
  uint32_t max_depth = 0;  // binary tree
  uint32_t wide_depth = 0; // levels of 8-wide nodes (root = 1)
  uint32_t n_leaves = 0;
};

// std::vector whose resize(n) leaves new elements uninitialised (trivial types only): the builder
// sizes these arrays once and its worker threads fill them, so the pages are first touched by the
// thread that writes them instead of being zero-filled by one thread (480 MB of triangle records
// at 10 M triangles).
template <typename T>
struct default_init_alloc : std::allocator<T>
{
  template <typename U>
  struct rebind
  {
    using other = default_init_alloc<U>;
  };
  default_init_alloc() = default;
  template <typename U>
  default_init_alloc(const default_init_alloc<U>&) noexcept
  {
  }
  template <typename U>
  void construct(U* p)
  {
    ::new (static_cast<void*>(p)) U;
  }
  template <typename U, typename... A>
  void construct(U* p, A&&... a)
  {
    ::new (static_cast<void*>(p)) U(std::forward<A>(a)...);
  }
};
template <typename T>
using RawVector = std::vector<T, default_init_alloc<T>>;

struct HostBvh
{
  RawVector<Node8> nodes;       // nodes[0] is the root
  RawVector<TriRecord> tris;    // leaf order
  RawVector<uint32_t> face_ids; // leaf order -> global face id
  BuildStats stats;
};

// Builds the wide BVH for all faces of `desc` (host, single call, deterministic).
void build_bvh(const ptc_scene_desc* desc, HostBvh& out);

// Shading records in global face order.
// returns the largest material id any face names (0 for an empty scene), found in the same pass
uint32_t build_shade_records(const ptc_scene_desc* desc, RawVector<ShadeRecord>& out);

} // namespace ptc
