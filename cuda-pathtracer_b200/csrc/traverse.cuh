// traverse.cuh -- RNG, camera ray generation and closest hit (device functions).
//
//   rng_*        cuRAND XORWOW restated (curand_kernel.h:772-799,863-874)
//   primary_ray  RNG seed + camera ray + depth of field       (raytrace.cu:218-240)
//   closest_hit  what intersect() computes by brute force (intersection.cuh:161-212),
//                answered through the 8-wide compressed BVH
#pragma once

#include <cuda_fp16.h>

#include "device_types.cuh"
#include "vecmath.cuh"

namespace ptc {

// node layout: 0 = lo[3][8], hi[3][8] (six 16-byte plane loads addressed by ray octant),
// 1 = [axis][lo | hi][8] (three 32-byte loads + register selects); must match bvh_builder.cpp
// Measured on B200 (profiles/r2_experiment_log.md): layout 1 is 1-1.4 % SLOWER than layout 0 on the
// 28 KB and the 1.4 GB scene and equal once the triangle loads are batched -- the 24 register
// selects cost what the three fewer load instructions save -- so layout 0 stays the default.
#ifndef PTC_NODE_LAYOUT
#define PTC_NODE_LAYOUT 0
#endif

// ---------------------------------------------------------------------------
// cuRAND XORWOW, restated (curand_kernel.h:772-799 init with subsequence = offset = 0,
// :863-874 step, curand_uniform.h:69-72).  State is 6 words, carried per path.
// ---------------------------------------------------------------------------
struct Rng
{
  uint32_t v0, v1, v2, v3, v4, d;
};

__device__ __forceinline__ void
rng_init(Rng& s, uint32_t seed32)
{
  uint32_t s0 = seed32 ^ 0xaad26b49u;
  uint32_t s1 = 0xf7dcefddu; // high seed word is zero: hash_seed + tid wraps in 32 bits
  uint32_t t0 = 1099087573u * s0;
  uint32_t t1 = 2591861531u * s1;
  s.d = 6615241u + t1 + t0;
  s.v0 = 123456789u + t0;
  s.v1 = 362436069u ^ t0;
  s.v2 = 521288629u + t1;
  s.v3 = 88675123u ^ t1;
  s.v4 = 5783321u + t0;
}

__device__ __forceinline__ uint32_t
rng_next(Rng& s)
{
  uint32_t t = s.v0 ^ (s.v0 >> 2);
  s.v0 = s.v1;
  s.v1 = s.v2;
  s.v2 = s.v3;
  s.v3 = s.v4;
  s.v4 = (s.v4 ^ (s.v4 << 4)) ^ (t ^ (t << 1));
  s.d += 362437u;
  return s.v4 + s.d;
}

__device__ __forceinline__ float
rng_uniform(Rng& s)
{
  // x * 2^-32 is exact, so the fused and unfused forms agree bit for bit
  return __fmaf_rn((float)rng_next(s), 2.3283064e-10f, 2.3283064e-10f / 2.0f);
}

// Advances the generator by n draws (replay instead of storing 24 B of state per ray).
__device__ __forceinline__ void
rng_skip(Rng& s, uint32_t n)
{
  for (uint32_t k = 0; k < n; ++k)
    rng_next(s);
}

// raytrace.cu:227-229 under the reference's launch shape (16x16 blocks,
// gridDim.x = width/16 + 1, raytrace.cu:315-317) -- independent of OUR launch shape.
__device__ __forceinline__ uint32_t
reference_tid(uint32_t x, uint32_t y, uint32_t width)
{
  uint32_t gx = width / 16u + 1u;
  return ((x >> 4) + (y >> 4) * gx) * 256u + (y & 15u) * 16u + (x & 15u);
}

// The part of generateRay (intersection.cuh:75-90) that depends on the camera only: screen
// distance and the screen basis.  The stage kernel evaluates it once per CTA into shared memory
// (same instructions, so the same bits) instead of once per pixel: a tanf, an IEEE division and
// two normalisations, about 130 of the 2 100 instructions of a primary path.
struct CamBasis
{
  float screen_dist;
  float3 cu, cv;
};

__device__ __forceinline__ CamBasis
camera_basis(const FrameParams& fp)
{
  CamBasis B;
  const int half_w = (int)(fp.width / 2);
  const float3 cdir = mk3(fp.cam.dir[0], fp.cam.dir[1], fp.cam.dir[2]);
  B.screen_dist = half_w / tanf(fp.cam.fov_x * 0.5f);
  float3 cu = normalize(cross(cdir, mk3(0.0f, -1.0f, 0.0f)));
  B.cv = normalize(cross(cu, cdir));
  B.cu = cu * -1.0f;
  return B;
}

__device__ __forceinline__ void
primary_ray(const FrameParams& fp, const CamBasis& B, uint32_t x, uint32_t y, Rng& rng, float3& origin,
            float3& dir)
{
  rng_init(rng, fp.hash_seed + reference_tid(x, y, fp.width));

  // generateRay, intersection.cuh:75-97
  const int half_w = (int)(fp.width / 2), half_h = (int)(fp.height / 2);
  const float3 cpos = mk3(fp.cam.position[0], fp.cam.position[1], fp.cam.position[2]);
  const float3 cdir = mk3(fp.cam.dir[0], fp.cam.dir[1], fp.cam.dir[2]);
  const float screen_dist = B.screen_dist;
  const float3 cu = B.cu, cv = B.cv;
  float3 screen_pos =
    cpos + (cdir * screen_dist) + (cu * (float)((int)x - half_w)) + (cv * (float)((int)y - half_h));
  dir = normalize(screen_pos - cpos);
  origin = cpos;

  // camera_dof, post_process.cuh:49-67 (the * M_PI is a double multiply on Linux)
  float3 focal_point = fp.cam.focus_dist * dir;
  float random_angle = (float)((double)(rng_uniform(rng) * 2.0f) * 3.14159265358979323846);
  float random_radius = rng_uniform(rng) * fp.cam.aperture;
  float3 ap = (cosf(random_angle) * cu + sinf(random_angle) * cv) * random_radius;
  dir = normalize(focal_point - ap);
  origin = origin + ap;
}

__device__ __forceinline__ void
primary_ray(const FrameParams& fp, uint32_t x, uint32_t y, Rng& rng, float3& origin, float3& dir)
{
  primary_ray(fp, camera_basis(fp), x, y, rng, origin, dir);
}

// Moeller-Trumbore exactly as intersection.cuh:102-135 evaluates it on the device
// (fma placement per the reference build's SASS); e1/e2 are the reference's v0v1/v0v2.
// Returns t, or NaN-free "no hit" via the bool.
__device__ __forceinline__ bool
tri_test(const float4 t0, const float4 t1, const float4 t2, const float3 o, const float3 d, float& t,
         float& u_out, float& v_out)
{
  const float3 v0 = mk3(t0.x, t0.y, t0.z);
  const float3 e1 = mk3(t0.w, t1.x, t1.y);
  const float3 e2 = mk3(t1.z, t1.w, t2.x);
  const float3 p = cross_ref(d, e2);
  const float det = dot3_ref(e1, p);
  if (det < 1e-7f) // == ((double)det < 0.0000001): no float lies between the two constants
    return false;
  const float inv_det = __fdividef(1.f, det);
  const float3 tv = mk3(__fsub_rn(o.x, v0.x), __fsub_rn(o.y, v0.y), __fsub_rn(o.z, v0.z));
  const float u = __fmul_rn(dot3_ref(tv, p), inv_det);
  if (u < 0.f || u > 1.f)
    return false;
  const float3 q = cross_ref(tv, e1);
  const float v = __fmul_rn(dot3_ref(d, q), inv_det);
  if (v < 0.f || __fadd_rn(u, v) > 1.f)
    return false;
  t = __fmul_rn(dot3_ref(e2, q), inv_det);
  u_out = u;
  v_out = v;
  return true;
}

// intersectSphere, intersection.cuh:140-155, including the discarded-ternary
// behaviour at :152 (t keeps b + disc when both roots fail the epsilon test).
__device__ __forceinline__ bool
sphere_test(const ptc_light& L, const float3 o, const float3 d, float& t)
{
  const float3 op = mk3(__fsub_rn(L.vec[0], o.x), __fsub_rn(L.vec[1], o.y), __fsub_rn(L.vec[2], o.z));
  const float b = dot3_ref(op, d);
  float disc = __fmaf_rn(L.radius, L.radius, __fmaf_rn(b, b, -dot3_ref(op, op)));
  if (disc < 0.f)
    return false;
  disc = __fsqrt_rn(disc);
  t = __fsub_rn(b, disc);
  if (!(t > 0.01f))
    t = __fadd_rn(b, disc);
  return t != 0.f;
}

// Half-float pair -> fp32 (HADD2.F32 on the FMA pipe; the planes of a node are stored as
// halves for exactly this reason, see bvh.h).
__device__ __forceinline__ float
half_lo(uint32_t w)
{
  return __low2float(*reinterpret_cast<const __half2*>(&w));
}
__device__ __forceinline__ float
half_hi(uint32_t w)
{
  return __high2float(*reinterpret_cast<const __half2*>(&w));
}

// Packed fp32 pair helpers for fma.rn.f32x2 (sm_100 FFMA2).
__device__ __forceinline__ uint64_t
pack2(float lo, float hi)
{
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void
fma2(float& r0, float& r1, float q0, float q1, uint64_t a, uint64_t b)
{
  uint64_t r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(pack2(q0, q1)), "l"(a), "l"(b));
  asm("mov.b64 {%0, %1}, %2;" : "=f"(r0), "=f"(r1) : "l"(r));
}

// bit s of an 8-bit mask -> bit 3s (the triangle space of a node gives every slot 3 bits)
__device__ __forceinline__ uint32_t
spread3(uint32_t x)
{
  x = (x | (x << 8)) & 0x00F00Fu;
  x = (x | (x << 4)) & 0x0C30C3u;
  x = (x | (x << 2)) & 0x249249u;
  return x;
}

// bit s of an 8-bit mask -> bit s ^ c (children are visited in ray-octant order)
__device__ __forceinline__ uint32_t
xor_permute8(uint32_t m, uint32_t c)
{
  if (c & 1u)
    m = ((m & 0x55u) << 1) | ((m >> 1) & 0x55u);
  if (c & 2u)
    m = ((m & 0x33u) << 2) | ((m >> 2) & 0x33u);
  if (c & 4u)
    m = ((m & 0x0Fu) << 4) | (m >> 4);
  return m;
}

#ifndef PTC_NO_TABLES
#define PTC_TABLES 1
#endif
// Two small tables the stage kernels keep in shared memory (filled once per CTA, trav_tables_init):
// the octant permutation of a hit mask (8 x 256 bytes) and the slot mask -> triangle-bit mask
// spreading (256 words).  One LDS each instead of 12 and 8 ALU instructions per node test: the
// kernels are bound by instruction issue with the ALU pipe as the busiest one.
constexpr int kPermBytes = 8 * 256, kSpreadWords = 256;

// Per-ray traversal stack.  Two homes with one interface:
//  * SharedStack -- the stage kernels: entry e of thread t lives at smem[(e * kBlock + t) * 8],
//    so whatever depth each lane is at, a warp's push/pop is one conflict-free shared-memory
//    access (2 wavefronts for 64 bits).  In local memory the same access costs one L1 wavefront
//    per lane that sits at a different depth (measured on the 10 M-triangle scene: 37 % of all
//    L1 wavefronts of k_extend).  `cap` = levels of the scene's wide tree + 1 + kTriBacklog.
//  * LocalStack -- the debug / parity kernels (one thread per ray, no shared-memory budget).
// 32-byte read-only load (sm_100: LDG.E.256): one L1 tag request where two 16-byte loads make two.
__device__ __forceinline__ void
ldg256(const uint4* p, uint4& a, uint4& b)
{
  asm("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
      : "=r"(a.x), "=r"(a.y), "=r"(a.z), "=r"(a.w), "=r"(b.x), "=r"(b.y), "=r"(b.z), "=r"(b.w)
      : "l"(p));
}

// Where a stack type reads the BVH from (every stack type carries the same five loaders, so that
// trav_step is written once): global memory through the read-only path ...
struct GlobalBvhLoads
{
  static constexpr int kBacklog = kTriBacklog;
  __device__ __forceinline__ void ld_node_hdr(const DScene& sc, uint32_t node, uint4& a, uint4& b) const
  {
    ldg256(sc.nodes + node * kNodeVec, a, b);
  }
  __device__ __forceinline__ uint4 ld_node_vec(const DScene& sc, uint32_t node, uint32_t k) const
  {
    return __ldg(sc.nodes + node * kNodeVec + k);
  }
  __device__ __forceinline__ void ld_node_vec2(const DScene& sc, uint32_t node, uint32_t k, uint4& a, uint4& b) const
  {
    ldg256(sc.nodes + node * kNodeVec + k, a, b);
  }
  __device__ __forceinline__ float4 ld_tri(const DScene& sc, uint32_t rec, uint32_t k) const
  {
    return __ldg(sc.tris + (size_t)rec * kTriVec + k);
  }
  __device__ __forceinline__ void prefetch_node(const DScene& sc, uint32_t node) const
  {
    asm volatile("prefetch.global.L1 [%0];" ::"l"(sc.nodes + node * kNodeVec));
  }
};

struct LocalStack : GlobalBvhLoads
{
  uint2 a[kStackSize];
  __device__ __forceinline__ uint32_t permute(uint32_t m, uint32_t oct_inv) const { return xor_permute8(m, oct_inv); }
  __device__ __forceinline__ uint32_t spread7(uint32_t m) const { return spread3(m) * 7u; }
  __device__ __forceinline__ uint2 get(int i) const { return a[i]; }
  __device__ __forceinline__ void put(int i, uint2 v) { a[i] = v; }
  __device__ __forceinline__ int top() const { return kStackSize; }
};

struct SharedStack : GlobalBvhLoads
{
  uint2* base; // smem + threadIdx.x
  int cap;
#ifdef PTC_TABLES
  const uint8_t* perm;    // [oct_inv][mask]
  const uint32_t* spread; // [mask] -> spread3(mask) * 7
  __device__ __forceinline__ uint32_t permute(uint32_t m, uint32_t oct_inv) const { return perm[(oct_inv << 8) | m]; }
  __device__ __forceinline__ uint32_t spread7(uint32_t m) const { return spread[m]; }
#else
  __device__ __forceinline__ uint32_t permute(uint32_t m, uint32_t oct_inv) const { return xor_permute8(m, oct_inv); }
  __device__ __forceinline__ uint32_t spread7(uint32_t m) const { return spread3(m) * 7u; }
#endif
  __device__ __forceinline__ uint2 get(int i) const { return base[i * kBlock]; }
  __device__ __forceinline__ void put(int i, uint2 v) { base[i * kBlock] = v; }
  __device__ __forceinline__ int top() const { return cap; }
};

// ... or, for a scene whose nodes + triangle records fit beside the stacks of PTC_PATH_BLOCKS
// resident CTAs (every scene the reference ships: 10-30 KB), a per-CTA copy in SHARED memory.
// Lanes of a warp read different nodes, so a node load is one L1 wavefront per distinct line in
// global memory (l1tex__data_pipe_lsu_wavefronts was the busiest unit of the path kernels, 57-62 %
// of peak); here nodes are 144 bytes apart (128 + 16 of padding), so the same word of eight
// different nodes lies in eight different bank groups, and an LDS is a short-scoreboard wait.

struct SharedSceneStack
{
  static constexpr int kBacklog = kSmemSceneBacklog;
  uint2* base;
  int cap;
  const uint8_t* perm;
  const uint32_t* spread;
  uint32_t nodes_s, tris_s; // shared-window byte addresses of the copies
  __device__ __forceinline__ uint32_t permute(uint32_t m, uint32_t oct_inv) const { return perm[(oct_inv << 8) | m]; }
  __device__ __forceinline__ uint32_t spread7(uint32_t m) const { return spread[m]; }
  __device__ __forceinline__ uint2 get(int i) const { return base[i * kBlock]; }
  __device__ __forceinline__ void put(int i, uint2 v) { base[i * kBlock] = v; }
  __device__ __forceinline__ int top() const { return cap; }
  static __device__ __forceinline__ uint4 lds128(uint32_t a)
  {
    uint4 v;
    asm("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a));
    return v;
  }
  __device__ __forceinline__ void ld_node_hdr(const DScene&, uint32_t node, uint4& a, uint4& b) const
  {
    const uint32_t p = nodes_s + node * kSmemNodeBytes;
    a = lds128(p), b = lds128(p + 16u);
  }
  __device__ __forceinline__ uint4 ld_node_vec(const DScene&, uint32_t node, uint32_t k) const
  {
    return lds128(nodes_s + node * kSmemNodeBytes + k * 16u);
  }
  __device__ __forceinline__ void ld_node_vec2(const DScene&, uint32_t node, uint32_t k, uint4& a, uint4& b) const
  {
    const uint32_t p = nodes_s + node * kSmemNodeBytes + k * 16u;
    a = lds128(p), b = lds128(p + 16u);
  }
  __device__ __forceinline__ float4 ld_tri(const DScene&, uint32_t rec, uint32_t k) const
  {
    const uint4 v = lds128(tris_s + rec * kSmemTriBytes + k * 16u);
    return make_float4(__uint_as_float(v.x), __uint_as_float(v.y), __uint_as_float(v.z), __uint_as_float(v.w));
  }
  __device__ __forceinline__ void prefetch_node(const DScene&, uint32_t) const {}
};

// Copies the scene's nodes (padded to kSmemNodeBytes) and triangle records behind the stacks (all
// threads of the CTA; the caller synchronises afterwards).
__device__ __forceinline__ void
smem_scene_fill(const DScene& sc, uint4* nodes_dst, uint4* tris_dst)
{
  for (uint32_t i = threadIdx.x; i < sc.n_nodes * (uint32_t)kNodeVec; i += blockDim.x)
    nodes_dst[(i >> 3) * (kSmemNodeBytes / 16u) + (i & 7u)] = __ldg(sc.nodes + i);
  const uint4* tsrc = reinterpret_cast<const uint4*>(sc.tris);
  for (uint32_t i = threadIdx.x; i < sc.n_tris * (uint32_t)kTriVec; i += blockDim.x)
    tris_dst[i] = __ldg(tsrc + i);
}

// Fills the two tables (all threads of the CTA; the caller synchronises afterwards).
__device__ __forceinline__ void
trav_tables_init(uint8_t* perm, uint32_t* spread)
{
  for (uint32_t i = threadIdx.x; i < (uint32_t)kPermBytes; i += blockDim.x)
    perm[i] = (uint8_t)xor_permute8(i & 0xFFu, i >> 8);
  for (uint32_t i = threadIdx.x; i < (uint32_t)kSpreadWords; i += blockDim.x)
    spread[i] = spread3(i) * 7u;
}

// Traversal state of one ray (registers; the stack: see above).
struct Trav
{
  float3 o, d, idir;
  uint32_t oct, oct_inv;
#if PTC_NODE_LAYOUT == 0
  uint32_t vnx, vny, vnz; // uint4 index (inside a node) of the near plane set per axis
#endif
  uint2 ng; // node group: first child index, 8 hit bits (31..24) | imask (7..0)
  uint2 tg; // triangle group: tri_base, valid24 of its node
  uint32_t tmask; // triangles of tg still to test (bits of the node's 24-bit triangle space)
  int sp;  // node-group entries on stack[0 .. sp)
  int tsp; // deferred triangle groups on stack[top - tsp .. top)
  float best_t, best_u, best_v;
  uint32_t best_id;
  bool any;  // any-hit query: stop at the first accepted triangle (trav_init_any)
  bool incl; // any-hit: a triangle at exactly best_t also counts
};

__device__ __forceinline__ void
trav_init(Trav& T, const float3 o, const float3 d)
{
  T.o = o;
  T.d = d;
  T.best_t = kMaxDist;
  T.best_id = 0xFFFFFFFFu;
  T.best_u = T.best_v = 0.f;
  T.any = T.incl = false;
  // guarded reciprocal direction for the (conservative) slab tests only
  const float eps = 1e-20f;
  T.idir = mk3(1.0f / (fabsf(d.x) > eps ? d.x : copysignf(eps, d.x)),
               1.0f / (fabsf(d.y) > eps ? d.y : copysignf(eps, d.y)),
               1.0f / (fabsf(d.z) > eps ? d.z : copysignf(eps, d.z)));
  T.oct = (d.x < 0.f ? 1u : 0u) | (d.y < 0.f ? 2u : 0u) | (d.z < 0.f ? 4u : 0u);
  T.oct_inv = 7u - T.oct;
#if PTC_NODE_LAYOUT == 0
  T.vnx = d.x < 0.f ? 5u : 2u; // lo planes at 2,3,4 / hi planes at 5,6,7 (bvh.h)
  T.vny = d.y < 0.f ? 6u : 3u;
  T.vnz = d.z < 0.f ? 7u : 4u;
#endif
  T.sp = 0;
  T.tsp = 0;
  T.ng = make_uint2(0u, 0x80000000u); // root as a one-entry node group
  T.tg = make_uint2(0u, 0u);
  T.tmask = 0u;
}

// Last radiance-loop iteration of a path (raytrace.cu:67 with b == max_bounces - 1): whatever
// the ray hits, the loop ends afterwards, and only a LIGHT hit (emission, raytrace.cu:98-99)
// or a MISS (environment, :194-199) adds radiance -- a triangle hit adds nothing and its
// material is never looked at.  So the closest-hit query degenerates into: nearest sphere
// light first (intersection.cuh:199-212 accepts a light only if it is strictly nearer than
// the nearest triangle), then an ANY-hit query for a triangle with 0 < t <= t_light (or
// 0 < t < MAX_DIST when no light is on the ray), which stops at the first triangle found.
__device__ __forceinline__ void lights_hit(const DScene& sc, const float3 o, const float3 d, float& best_t,
                                           uint32_t& best_id);

__device__ __forceinline__ void
trav_init_any(const DScene& sc, Trav& T, const float3 o, const float3 d)
{
  trav_init(T, o, d);
  lights_hit(sc, o, d, T.best_t, T.best_id);
  T.any = true;
  T.incl = T.best_id != 0xFFFFFFFFu;
}

#ifndef PTC_TRI_PER_ITER
#define PTC_TRI_PER_ITER 2
#endif

// One traversal iteration: AT MOST PTC_TRI_PER_ITER triangle tests and one 8-wide node test.  Returns
// true when the ray has nothing left to visit.
//
// Traversal order never changes the answer (the winner is the lexicographic minimum of
// (t, face id) and boxes are only culled when strictly farther than the best hit), so the
// loop is shaped for SIMT convergence: a leaf's remaining triangles are deferred on the
// stack instead of being looped over while the other lanes of the warp idle.
//
// F2: the slab tests of two children share one packed fp32 FMA (fma.rn.f32x2 -> FFMA2, new on
// sm_100): each half is the same round-to-nearest fma as the scalar form, so the hit masks do not
// change, and the node test issues 24 FFMA2 instead of 48 FFMA.  Measured on B200: -2.5 % frame time
// on the 10 M-triangle scene (k_extend, 80 registers, no spills); +0.3 % in the fused path kernels,
// whose 64-register cap turns the register pairs into extra spills -- so only k_extend asks for it.
#ifndef PTC_PREFETCH_NODE
#define PTC_PREFETCH_NODE 2
#endif
// the same two switches for the traversal inside the fused path kernels (64-register cap)
#ifndef PTC_PATH_PF
#define PTC_PATH_PF 0
#endif
#ifndef PTC_PATH_TB
#define PTC_PATH_TB 0
#endif
template <bool STATS, typename Stack, bool F2 = false, int PF = 0, bool TB = false>
__device__ __forceinline__ bool
trav_step(const DScene& sc, Trav& T, Stack& stack, uint32_t& n_nodes, uint32_t& n_tris)
{
  // keep both tests fed: an exhausted group is replaced from its stack right away, so a lane
  // runs a triangle test AND a node test in the same iteration whenever it has either left
  if (!T.tmask && T.tsp) {
    const uint2 e = stack.get(stack.top() - T.tsp); // (tri_base, valid24 | leaf-slot hit mask << 24)
    --T.tsp;
    T.tg = make_uint2(e.x, e.y & 0x00FFFFFFu);
    T.tmask = stack.spread7(e.y >> 24) & T.tg.y;
  }
  if (!(T.ng.y & 0xFF000000u) && T.sp)
    T.ng = stack.get(--T.sp);

  // -- which node this iteration tests ------------------------------------------------------
  // A warp whose lanes hold both kinds of work waits for the triangle records, computes, and only
  // then asks for the nodes: dependent memory round trips per iteration although every address is
  // known here.  PF != 0 (the extend kernel) selects the node BEFORE the triangle tests and
  // requests its line from L1 (CCTL.E.PF1): 1 = the lanes that also hold triangles, 2 = every lane
  // with a node, 3 = the node header itself is loaded here.  PF == 0 selects it after them (the
  // early selection alone costs the fused path kernels 2 %).
  bool do_node = (T.ng.y & 0xFF000000u) && T.tsp < Stack::kBacklog;
  uint32_t node = 0;
  uint4 n0 = make_uint4(0, 0, 0, 0), n1 = n0;
  auto select_node = [&]() {
    const uint32_t bit = 31u - __clz(T.ng.y);
    T.ng.y &= ~(1u << bit);
    const uint32_t imask_parent = T.ng.y & 0xFFu;
    if (T.ng.y & 0xFF000000u)
      stack.put(T.sp++, T.ng);
    const uint32_t slot = (bit - 24u) ^ T.oct_inv;
    const uint32_t rel = __popc(imask_parent & ~(0xFFFFFFFFu << slot));
    node = T.ng.x + rel;
  };
  if (PF != 0 && do_node) {
    select_node();
    if (PF == 3)
      stack.ld_node_hdr(sc, node, n0, n1);
    else if (PF == 2 || (PF == 1 && T.tmask))
      stack.prefetch_node(sc, node);
  }

  // -- at most PTC_TRI_PER_ITER triangle tests --------------------------------------------------
  // (Holding triangle tests back until N lanes of the warp have one pending was measured for
  // N = 8..24: 4-7 % slower on both ends of the scene-size range -- the later a hit shrinks
  // best_t, the more boxes are entered.)
  // TB: the records of BOTH tests are requested before the first is evaluated.  The source-level
  // profile of the extend kernel on the 10 M-triangle scene showed three exposed memory waits per
  // iteration -- first use of triangle record 1 (17 % of all stall samples), of record 2 (13 %),
  // of the node header (15 %) -- because the second record's address was only formed after the
  // first test.  Same tests, same order, same bits.
  bool hit_any = false;
  auto accept = [&](const float4& t2, float t, float u, float v) {
    const uint32_t fid = __float_as_uint(t2.y);
    if (t > 0.f && (t < T.best_t || (t == T.best_t && (T.any ? T.incl : fid < T.best_id)))) {
      T.best_t = t;
      T.best_id = fid;
      T.best_u = u;
      T.best_v = v;
      if (T.any) { // occluded: nothing else about this ray matters
        T.ng.y = 0u, T.tmask = 0u;
        T.sp = T.tsp = 0;
        hit_any = true;
      }
    }
  };
  if (TB && PTC_TRI_PER_ITER == 2) {
    if (T.tmask) {
      const uint32_t bit0 = __ffs(T.tmask) - 1u;
      T.tmask &= T.tmask - 1u;
      const bool two = T.tmask != 0u;
      const uint32_t bit1 = two ? __ffs(T.tmask) - 1u : bit0;
      T.tmask &= T.tmask - 1u;
      const uint32_t r0 = T.tg.x + __popc(T.tg.y & ((1u << bit0) - 1u));
      const uint32_t r1 = T.tg.x + __popc(T.tg.y & ((1u << bit1) - 1u));
      const float4 a0 = stack.ld_tri(sc, r0, 0), a1 = stack.ld_tri(sc, r0, 1), a2 = stack.ld_tri(sc, r0, 2);
      float4 b0 = a0, b1 = a1, b2 = a2;
      if (two)
        b0 = stack.ld_tri(sc, r1, 0), b1 = stack.ld_tri(sc, r1, 1), b2 = stack.ld_tri(sc, r1, 2);
      if (STATS)
        n_tris += two ? 2u : 1u;
      float t, u, v;
      if (tri_test(a0, a1, a2, T.o, T.d, t, u, v))
        accept(a2, t, u, v);
      if (two && !hit_any && tri_test(b0, b1, b2, T.o, T.d, t, u, v))
        accept(b2, t, u, v);
    }
  } else {
#pragma unroll
    for (int rep = 0; rep < PTC_TRI_PER_ITER; ++rep)
      if (T.tmask) {
        const uint32_t bit = __ffs(T.tmask) - 1u;
        T.tmask &= T.tmask - 1u;
        const uint32_t rec = T.tg.x + __popc(T.tg.y & ((1u << bit) - 1u));
        const float4 t0 = stack.ld_tri(sc, rec, 0), t1 = stack.ld_tri(sc, rec, 1), t2 = stack.ld_tri(sc, rec, 2);
        if (STATS)
          n_tris++;
        float t, u, v;
        if (tri_test(t0, t1, t2, T.o, T.d, t, u, v))
          accept(t2, t, u, v);
      }
  }

  // -- at most one node test ------------------------------------------------------------
  // (A warp vote that holds the node test back while fewer than N = 4 / 8 / 12 lanes want one and
  // others still have triangles to test was measured: 3-5 % slower -- the held lanes sit on the
  // warp's critical path more often than not.)
  if (do_node && !hit_any) {
    if (PF == 0)
      select_node();
    if (PF != 3)
      stack.ld_node_hdr(sc, node, n0, n1);
#if PTC_NODE_LAYOUT == 0
    // near / far plane sets picked by ray octant through the load address, not by selects
    // (measured: deriving the far index as near ^ c instead of the subtraction is slower; so is
    // summing the indices in 32 bits first, +1.3 %, although it is two instructions shorter per set)
    const uint4 NX = stack.ld_node_vec(sc, node, T.vnx), FX = stack.ld_node_vec(sc, node, 7u - T.vnx);
    const uint4 NY = stack.ld_node_vec(sc, node, T.vny), FY = stack.ld_node_vec(sc, node, 9u - T.vny);
    const uint4 NZ = stack.ld_node_vec(sc, node, T.vnz), FZ = stack.ld_node_vec(sc, node, 11u - T.vnz);
#else
    // One 32-byte load per axis (lower and upper planes of the axis are adjacent, bvh.h): the node
    // is FOUR load instructions instead of seven.  Every lane of a warp reads a different node, so
    // each load instruction costs one L1 wavefront per lane, and the L1 data pipe -- not DRAM, not
    // occupancy (2, 3 and 4 CTAs per SM measured the same) -- is what the traversal kernels
    // saturate first (ncu: l1tex__data_pipe_lsu_wavefronts 62-73 % of peak on the L2-resident
    // 180 k-triangle scene).  The near / far set is then a register select on the ray's octant.
    uint4 X0, X1, Y0, Y1, Z0, Z1;
    stack.ld_node_vec2(sc, node, 2, X0, X1);
    stack.ld_node_vec2(sc, node, 4, Y0, Y1);
    stack.ld_node_vec2(sc, node, 6, Z0, Z1);
    const bool ngx = T.oct & 1u, ngy = T.oct & 2u, ngz = T.oct & 4u;
#define PTC_SEL4(c, A, B) make_uint4((c) ? (A).x : (B).x, (c) ? (A).y : (B).y, (c) ? (A).z : (B).z, (c) ? (A).w : (B).w)
    const uint4 NX = PTC_SEL4(ngx, X1, X0), FX = PTC_SEL4(ngx, X0, X1);
    const uint4 NY = PTC_SEL4(ngy, Y1, Y0), FY = PTC_SEL4(ngy, Y0, Y1);
    const uint4 NZ = PTC_SEL4(ngz, Z1, Z0), FZ = PTC_SEL4(ngz, Z0, Z1);
#undef PTC_SEL4
#endif
    if (STATS)
      n_nodes++;

    // plane = origin + q * 2^(e-127);  t = (plane - o) * idir = q * a + b
    const float sx = __uint_as_float((n0.w & 0xFFu) << 23);
    const float sy = __uint_as_float(((n0.w >> 8) & 0xFFu) << 23);
    const float sz = __uint_as_float(((n0.w >> 16) & 0xFFu) << 23);
    const float ax = sx * T.idir.x, ay = sy * T.idir.y, az = sz * T.idir.z;
    const float bx = (__uint_as_float(n0.x) - T.o.x) * T.idir.x;
    const float by = (__uint_as_float(n0.y) - T.o.y) * T.idir.y;
    const float bz = (__uint_as_float(n0.z) - T.o.z) * T.idir.z;

    uint32_t hit8 = 0;
    if (F2) {
    // two children per FMA (see above); the kernels are bound by instruction issue, not by the
    // FMA pipe (ncu: issue 65-73 %, FMA pipe 26-28 %)
    const uint64_t ax2 = pack2(ax, ax), ay2 = pack2(ay, ay), az2 = pack2(az, az);
    const uint64_t bx2 = pack2(bx, bx), by2 = pack2(by, by), bz2 = pack2(bz, bz);
#define PTC_PAIR(k, W)                                                                           \
  {                                                                                              \
    float tnx0, tnx1, tfx0, tfx1, tny0, tny1, tfy0, tfy1, tnz0, tnz1, tfz0, tfz1;                \
    fma2(tnx0, tnx1, half_lo(NX.W), half_hi(NX.W), ax2, bx2);                                    \
    fma2(tfx0, tfx1, half_lo(FX.W), half_hi(FX.W), ax2, bx2);                                    \
    fma2(tny0, tny1, half_lo(NY.W), half_hi(NY.W), ay2, by2);                                    \
    fma2(tfy0, tfy1, half_lo(FY.W), half_hi(FY.W), ay2, by2);                                    \
    fma2(tnz0, tnz1, half_lo(NZ.W), half_hi(NZ.W), az2, bz2);                                    \
    fma2(tfz0, tfz1, half_lo(FZ.W), half_hi(FZ.W), az2, bz2);                                    \
    const float tmin0 = fmaxf(fmaxf(tnx0, tny0), fmaxf(tnz0, 0.0f));                             \
    const float tmax0 = fminf(fminf(tfx0, tfy0), fminf(tfz0, T.best_t));                         \
    const float tmin1 = fmaxf(fmaxf(tnx1, tny1), fmaxf(tnz1, 0.0f));                             \
    const float tmax1 = fminf(fminf(tfx1, tfy1), fminf(tfz1, T.best_t));                         \
    if (tmin0 <= tmax0)                                                                          \
      hit8 |= 1u << (k);                                                                         \
    if (tmin1 <= tmax1)                                                                          \
      hit8 |= 2u << (k);                                                                         \
  }
    PTC_PAIR(0, x) PTC_PAIR(2, y) PTC_PAIR(4, z) PTC_PAIR(6, w)
#undef PTC_PAIR
    } else {
#define PTC_CHILD(k, W, H)                                                                     \
  {                                                                                              \
    const float tnx = fmaf(H(NX.W), ax, bx), tfx = fmaf(H(FX.W), ax, bx);                        \
    const float tny = fmaf(H(NY.W), ay, by), tfy = fmaf(H(FY.W), ay, by);                        \
    const float tnz = fmaf(H(NZ.W), az, bz), tfz = fmaf(H(FZ.W), az, bz);                        \
    const float tmin = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, 0.0f));                                 \
    const float tmax = fminf(fminf(tfx, tfy), fminf(tfz, T.best_t));                             \
    if (tmin <= tmax)                                                                            \
      hit8 |= 1u << k;                                                                           \
  }
    PTC_CHILD(0, x, half_lo) PTC_CHILD(1, x, half_hi) PTC_CHILD(2, y, half_lo) PTC_CHILD(3, y, half_hi)
    PTC_CHILD(4, z, half_lo) PTC_CHILD(5, z, half_hi) PTC_CHILD(6, w, half_lo) PTC_CHILD(7, w, half_hi)
#undef PTC_CHILD
    }

    // internal children: bit s -> 24 + (s ^ oct_inv), i.e. far-to-near from the top bit down
    const uint32_t imask = n0.w >> 24;
    const uint32_t ih = stack.permute(hit8 & imask, T.oct_inv);
    T.ng = make_uint2(n1.x, (ih << 24) | imask);
    // (Asking L2 for every hit internal child here -- prefetch.global.L2, DRAM runs at 10 % of its
    // bandwidth on the 10 M-triangle scene -- was measured: +2 % frame time; the extra issue slots
    // cost more than the shorter misses give back.)

    // leaf slots that were hit -> their triangles (3 bits per slot, masked by the existing ones)
    const uint32_t lh = hit8 & ~imask;
    const uint32_t tnew = stack.spread7(lh) & n1.z;
    if (tnew) {
      if (T.tmask) {
        stack.put(stack.top() - (++T.tsp), make_uint2(n1.y, n1.z | (lh << 24))); // defer, unstarted
      } else {
        T.tg = make_uint2(n1.y, n1.z);
        T.tmask = tnew;
      }
    }
  }

  return !((T.ng.y & 0xFF000000u) | T.tmask | (uint32_t)T.sp | (uint32_t)T.tsp);
}

// 32-byte light record in two 16-byte loads (the array comes from cudaMalloc; the C struct alone
// only promises 4-byte alignment, which makes the compiler emit eight scalar loads).
__device__ __forceinline__ ptc_light
load_light(const DScene& sc, uint32_t l)
{
  const float4* lp = reinterpret_cast<const float4*>(sc.lights + l);
  const float4 l0 = __ldg(lp), l1 = __ldg(lp + 1);
  ptc_light L;
  L.color[0] = l0.x, L.color[1] = l0.y, L.color[2] = l0.z;
  L.vec[0] = l0.w, L.vec[1] = l1.x, L.vec[2] = l1.y;
  L.emission = l1.z, L.radius = l1.w;
  return L;
}

// Sphere lights after the triangles, 0 <= t < best in light order (intersection.cuh:199-212).
__device__ __forceinline__ void
lights_hit(const DScene& sc, const float3 o, const float3 d, float& best_t, uint32_t& best_id)
{
  for (uint32_t l = 0; l < sc.n_lights; ++l) {
    float t;
    const ptc_light L = load_light(sc, l);
    if (sphere_test(L, o, d, t) && t < best_t && t >= 0.f) {
      best_t = t;
      best_id = sc.n_faces + l;
    }
  }
}

// One radiance-loop iteration's intersection query, ONE instance of the traversal loop for both
// forms (the stage kernels used to inline a closest-hit and an any-hit copy of it, and a third for
// the preview frame: ~450 instructions each in kernels whose first stage showed 19 % of its stall
// samples waiting for instructions):
//   last == false  closest hit with the reference's acceptance rules -- triangles front-facing
//                  only, 0 < t < best, lowest face index wins equal t (intersection.cuh:110,
//                  183-184) -- then the sphere lights (:199-212)
//   last == true   what the LAST iteration needs to know (trav_init_any): id = light id, miss, or
//                  SOME face id when a triangle is in front; u, v are not meaningful
template <bool STATS, typename Stack>
__device__ __forceinline__ void
trace_ray(const DScene& sc, Stack& stack, const float3 o, const float3 d, const bool last, float& best_t,
          uint32_t& best_id, float& best_u, float& best_v, uint32_t& n_nodes, uint32_t& n_tris)
{
  Trav T;
  trav_init(T, o, d);
  if (last) {
    lights_hit(sc, o, d, T.best_t, T.best_id);
    T.any = true;
    T.incl = T.best_id != 0xFFFFFFFFu;
  }
  while (!trav_step<STATS, Stack, false, PTC_PATH_PF, PTC_PATH_TB != 0>(sc, T, stack, n_nodes, n_tris)) {
  }
  best_t = T.best_t, best_id = T.best_id, best_u = T.best_u, best_v = T.best_v;
  if (!last)
    lights_hit(sc, o, d, best_t, best_id);
}

template <bool STATS, typename Stack>
__device__ __forceinline__ void
last_hit(const DScene& sc, Stack& stack, const float3 o, const float3 d, float& best_t, uint32_t& best_id,
         uint32_t& n_nodes, uint32_t& n_tris)
{
  float u, v;
  trace_ray<STATS>(sc, stack, o, d, true, best_t, best_id, u, v, n_nodes, n_tris);
}

template <bool STATS, typename Stack>
__device__ __forceinline__ void
closest_hit(const DScene& sc, Stack& stack, const float3 o, const float3 d, float& best_t, uint32_t& best_id,
            float& best_u, float& best_v, uint32_t& n_nodes, uint32_t& n_tris)
{
  trace_ray<STATS>(sc, stack, o, d, false, best_t, best_id, best_u, best_v, n_nodes, n_tris);
}

} // namespace ptc
