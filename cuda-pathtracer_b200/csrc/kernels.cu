// kernels.cu -- the wavefront path-tracing kernels for sm_100a.
//
//   generate  : RNG seed + camera ray + depth of field       (raytrace.cu:218-240)
//   extend    : closest hit through the 8-wide compressed BVH (answers what
//               intersect() computes by brute force, intersection.cuh:161-212)
//   shade     : material resolve, BRDF sampling, throughput, roulette, env on miss,
//               compaction of the surviving paths into the next queue
//                                                             (raytrace.cu:67-207,
//                                                              intersection.cuh:214-243)
//   accumulate: clamp, temporal accumulation, tonemap, gamma, post, RGBA8
//                                                             (raytrace.cu:248-270)
//
// No tensor cores (traversal is not a dense contraction), no RT cores (B200 has none).
#include "kernels.cuh"

#include "vecmath.cuh"

namespace ptc {

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

__device__ __forceinline__ void
rng_load(const PathState& st, uint32_t p, Rng& s)
{
  uint4 a = st.rng_a[p];
  uint2 b = st.rng_b[p];
  s.v0 = a.x, s.v1 = a.y, s.v2 = a.z, s.v3 = a.w, s.v4 = b.x, s.d = b.y;
}

__device__ __forceinline__ void
rng_store(const PathState& st, uint32_t p, const Rng& s)
{
  st.rng_a[p] = make_uint4(s.v0, s.v1, s.v2, s.v3);
  st.rng_b[p] = make_uint2(s.v4, s.d);
}

// raytrace.cu:227-229 under the reference's launch shape (16x16 blocks,
// gridDim.x = width/16 + 1, raytrace.cu:315-317) -- independent of OUR launch shape.
__device__ __forceinline__ uint32_t
reference_tid(uint32_t x, uint32_t y, uint32_t width)
{
  uint32_t gx = width / 16u + 1u;
  return ((x >> 4) + (y >> 4) * gx) * 256u + (y & 15u) * 16u + (x & 15u);
}

// ---------------------------------------------------------------------------
// generate: one thread per pixel of the band
// ---------------------------------------------------------------------------
__device__ __forceinline__ void
primary_ray(const FrameParams& fp, uint32_t x, uint32_t y, Rng& rng, float3& origin, float3& dir)
{
  rng_init(rng, fp.hash_seed + reference_tid(x, y, fp.width));

  // generateRay, intersection.cuh:75-97
  const int half_w = (int)(fp.width / 2), half_h = (int)(fp.height / 2);
  const float3 cpos = mk3(fp.cam.position[0], fp.cam.position[1], fp.cam.position[2]);
  const float3 cdir = mk3(fp.cam.dir[0], fp.cam.dir[1], fp.cam.dir[2]);
  float screen_dist = half_w / tanf(fp.cam.fov_x * 0.5f);
  float3 cu = normalize(cross(cdir, mk3(0.0f, -1.0f, 0.0f)));
  float3 cv = normalize(cross(cu, cdir));
  cu = cu * -1.0f;
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

__global__ void __launch_bounds__(256)
k_generate(FrameParams fp, PathState st)
{
  const uint32_t band_px = (fp.row_end - fp.row_begin) * fp.width;
  const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= band_px)
    return;
  const uint32_t x = p % fp.width, y = fp.row_begin + p / fp.width;
  Rng rng;
  float3 o, d;
  primary_ray(fp, x, y, rng, o, d);
  st.ray_o[p] = make_float4(o.x, o.y, o.z, 0.0f); // w: carried specular_col starts at 0 (Q3)
  st.ray_d[p] = make_float4(d.x, d.y, d.z, 0.0f);
  st.thr[p] = make_float4(1.0f, 1.0f, 1.0f, 0.0f);
  st.acc[p] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
  rng_store(st, p, rng);
}

// ---------------------------------------------------------------------------
// extend: closest hit
// ---------------------------------------------------------------------------

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

__device__ __forceinline__ float
byte_f(uint32_t w, int k)
{
  return (float)((w >> (8 * k)) & 0xFFu);
}

constexpr int kStackSize = 32;

// Closest hit with the reference's acceptance rules: triangles front-facing only,
// 0 < t < best, lowest face index wins equal t (intersection.cuh:110,183-184);
// then sphere lights with 0 <= t < best in light order (intersection.cuh:199-212).
template <bool STATS>
__device__ __forceinline__ void
closest_hit(const DScene& sc, const float3 o, const float3 d, float& best_t, uint32_t& best_id,
            float& best_u, float& best_v, uint32_t& n_nodes, uint32_t& n_tris)
{
  best_t = kMaxDist;
  best_id = 0xFFFFFFFFu;
  best_u = best_v = 0.f;

  // guarded reciprocal direction for the (conservative) slab tests only
  const float eps = 1e-20f;
  const float3 idir = mk3(1.0f / (fabsf(d.x) > eps ? d.x : copysignf(eps, d.x)),
                          1.0f / (fabsf(d.y) > eps ? d.y : copysignf(eps, d.y)),
                          1.0f / (fabsf(d.z) > eps ? d.z : copysignf(eps, d.z)));
  const uint32_t oct = (d.x < 0.f ? 1u : 0u) | (d.y < 0.f ? 2u : 0u) | (d.z < 0.f ? 4u : 0u);
  const uint32_t oct_inv = 7u - oct;

  uint2 stack[kStackSize];
  int sp = 0;
  uint2 ng = make_uint2(0u, 0x80000000u); // root as a one-entry node group
  uint2 tg = make_uint2(0u, 0u);

  for (;;) {
    if (ng.y & 0xFF000000u) {
      const uint32_t bit = 31u - __clz(ng.y);
      ng.y &= ~(1u << bit);
      const uint32_t imask = ng.y & 0xFFu;
      if (ng.y & 0xFF000000u)
        stack[sp++] = ng;
      const uint32_t slot = (bit - 24u) ^ oct_inv;
      const uint32_t rel = __popc(imask & ~(0xFFFFFFFFu << slot));
      const uint32_t idx = ng.x + rel;

      const uint4* np = sc.nodes + (size_t)idx * kNodeVec;
      const uint4 n0 = __ldg(np), n1 = __ldg(np + 1), n2 = __ldg(np + 2), n3 = __ldg(np + 3),
                  n4 = __ldg(np + 4);
      if (STATS)
        n_nodes++;

      const float sx = __uint_as_float((n0.w & 0xFFu) << 23);
      const float sy = __uint_as_float(((n0.w >> 8) & 0xFFu) << 23);
      const float sz = __uint_as_float(((n0.w >> 16) & 0xFFu) << 23);
      const float ax = sx * idir.x, ay = sy * idir.y, az = sz * idir.z;
      const float bx = (__uint_as_float(n0.x) - o.x) * idir.x;
      const float by = (__uint_as_float(n0.y) - o.y) * idir.y;
      const float bz = (__uint_as_float(n0.z) - o.z) * idir.z;

      uint32_t hitmask = 0;
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        const uint32_t meta4 = half ? n1.w : n1.z;
        const uint32_t lox = half ? n2.y : n2.x, loy = half ? n2.w : n2.z, loz = half ? n3.y : n3.x;
        const uint32_t hix = half ? n3.w : n3.z, hiy = half ? n4.y : n4.x, hiz = half ? n4.w : n4.z;
        const uint32_t nx = (oct & 1u) ? hix : lox, fx = (oct & 1u) ? lox : hix;
        const uint32_t ny = (oct & 2u) ? hiy : loy, fy = (oct & 2u) ? loy : hiy;
        const uint32_t nz = (oct & 4u) ? hiz : loz, fz = (oct & 4u) ? loz : hiz;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const float tnx = fmaf(byte_f(nx, k), ax, bx), tfx = fmaf(byte_f(fx, k), ax, bx);
          const float tny = fmaf(byte_f(ny, k), ay, by), tfy = fmaf(byte_f(fy, k), ay, by);
          const float tnz = fmaf(byte_f(nz, k), az, bz), tfz = fmaf(byte_f(fz, k), az, bz);
          const float tmin = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, 0.0f));
          const float tmax = fminf(fminf(tfx, tfy), fminf(tfz, best_t));
          if (tmin <= tmax) {
            const uint32_t m = (meta4 >> (8 * k)) & 0xFFu;
            const uint32_t inner = ((m & 0x18u) == 0x18u) ? oct_inv : 0u;
            hitmask |= (m >> 5) << ((m & 31u) ^ inner);
          }
        }
      }
      ng = make_uint2(n1.x, (hitmask & 0xFF000000u) | (n0.w >> 24));
      tg = make_uint2(n1.y, hitmask & 0x00FFFFFFu);
    } else {
      tg = ng;
      ng = make_uint2(0u, 0u);
    }

    while (tg.y) {
      const uint32_t bit = __ffs(tg.y) - 1u;
      tg.y &= tg.y - 1u;
      const float4* tp = sc.tris + (size_t)(tg.x + bit) * kTriVec;
      const float4 t0 = __ldg(tp), t1 = __ldg(tp + 1), t2 = __ldg(tp + 2);
      if (STATS)
        n_tris++;
      float t, u, v;
      if (tri_test(t0, t1, t2, o, d, t, u, v)) {
        const uint32_t fid = __float_as_uint(t2.y);
        if (t > 0.f && (t < best_t || (t == best_t && fid < best_id))) {
          best_t = t;
          best_id = fid;
          best_u = u;
          best_v = v;
        }
      }
    }

    if (!(ng.y & 0xFF000000u)) {
      if (sp == 0)
        break;
      ng = stack[--sp];
    }
  }

  for (uint32_t l = 0; l < sc.n_lights; ++l) {
    float t;
    if (sphere_test(sc.lights[l], o, d, t) && t < best_t && t >= 0.f) {
      best_t = t;
      best_id = sc.n_faces + l;
    }
  }
}

__global__ void __launch_bounds__(128)
k_extend(DScene sc, PathState st, const uint32_t* __restrict__ queue,
         const uint32_t* __restrict__ queue_count, uint32_t n_max, Counters* ctr, int bounce)
{
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t n = queue_count ? *queue_count : n_max;
  if (i >= n)
    return;
  const uint32_t p = queue ? queue[i] : i;
  const float4 o4 = st.ray_o[p], d4 = st.ray_d[p];
  float t, u, v;
  uint32_t id, nn = 0, nt = 0;
  closest_hit<false>(sc, mk3(o4.x, o4.y, o4.z), mk3(d4.x, d4.y, d4.z), t, id, u, v, nn, nt);
  st.hit[p] = make_uint4(__float_as_uint(t), id, __float_as_uint(u), __float_as_uint(v));
}

// Debug / parity entry: arbitrary rays in, hit ids out, with traversal statistics.
__global__ void __launch_bounds__(128)
k_intersect_rays(DScene sc, const float* __restrict__ origins, const float* __restrict__ dirs,
                 uint64_t n, int32_t* __restrict__ ids, float* __restrict__ ts, Counters* ctr)
{
  const uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
  if (i >= n)
    return;
  float t, u, v;
  uint32_t id, nn = 0, nt = 0;
  closest_hit<true>(sc, mk3(origins[3 * i], origins[3 * i + 1], origins[3 * i + 2]),
                    mk3(dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2]), t, id, u, v, nn, nt);
  if (ids)
    ids[i] = (int32_t)id;
  if (ts)
    ts[i] = t;
  if (ctr) {
    atomicAdd(&ctr->nodes_visited, (unsigned long long)nn);
    atomicAdd(&ctr->tris_tested, (unsigned long long)nt);
  }
}

// ---------------------------------------------------------------------------
// shade
// ---------------------------------------------------------------------------
struct Surface // what intersect() leaves in IntersectionData (intersection.cuh:7-18)
{
  float3 normal;
  float3 diffuse_col;
  float specular_col;
  float ior;
  bool is_light;
  float3 light_color;
  float light_emission;
};

// nearest texel, no filtering, no V flip: intersection.cuh:20-26
__device__ __forceinline__ int
texel_index(const DTexture& tex, float2 uv)
{
  int x = (int)(uv.x * (float)(tex.w - 1));
  int y = (int)(uv.y * (float)(tex.h - 1));
  return (y * tex.w + x) * tex.nb_chan;
}

// Material / normal resolve for the winning hit: intersection.cuh:183-243.
// `spec_carry` is IntersectionData.specular_col, which a light hit leaves untouched (Q3).
__device__ __forceinline__ void
resolve_hit(const DScene& sc, uint32_t id, float t, float u, float v, float3 dir, float spec_carry,
            Surface& s)
{
  if (id >= sc.n_faces) { // sphere light, intersection.cuh:199-212
    const ptc_light L = sc.lights[id - sc.n_faces];
    s.is_light = true;
    s.light_color = mk3(L.color[0], L.color[1], L.color[2]);
    s.light_emission = L.emission;
    s.diffuse_col = s.light_color;
    s.normal = normalize(mk3(L.vec[0], L.vec[1], L.vec[2]) - (t * dir)); // as written, :208
    s.specular_col = spec_carry;
    s.ior = 1.0f; // unused: the light test comes first (raytrace.cu:96)
    return;
  }
  s.is_light = false;
  const float4* sp = sc.shade + (size_t)id * kShadeVec;
  const float4 a = __ldg(sp), b = __ldg(sp + 1), c = __ldg(sp + 2), d = __ldg(sp + 3), e = __ldg(sp + 4);
  const float3 n0 = mk3(a.x, a.y, a.z), n1 = mk3(a.w, b.x, b.y), n2 = mk3(b.z, b.w, c.x);
  const float2 uv0 = make_float2(c.y, c.z), uv1 = make_float2(c.w, d.x), uv2 = make_float2(d.y, d.z);
  const float3 tangent = mk3(d.w, e.x, e.y);
  const uint32_t material_id = __float_as_uint(e.z);

  // intersection.cuh:124-131
  const float w0 = 1.0f - u - v;
  const float3 surface_normal = w0 * n0 + u * n1 + v * n2; // not normalised (Q8)
  float2 uv = make_float2(w0 * uv0.x + u * uv1.x + v * uv2.x, w0 * uv0.y + u * uv1.y + v * uv2.y);
  uv.x = uv.x - floorf(uv.x / 1.0f);
  uv.y = uv.y - floorf(uv.y / 1.0f);

  const ptc_material m = sc.materials[material_id];
  s.ior = m.ior;
  const DTexture dt = sc.textures[m.diffuse_spec_map];
  const int idx = texel_index(dt, uv);
  s.diffuse_col = mk3(dt.data[idx], dt.data[idx + 1], dt.data[idx + 2]);
  s.specular_col = dt.data[idx + 3];
  s.normal = surface_normal;
  if (m.normal_map >= 0) { // intersection.cuh:228-242
    const DTexture nt = sc.textures[m.normal_map];
    const int nidx = texel_index(nt, uv);
    float3 n = mk3(nt.data[nidx], nt.data[nidx + 1], nt.data[nidx + 2]);
    n = normalize((n * 2.0f) - 1.0f);
    const float3 binormal = normalize(cross(tangent, surface_normal));
    const float3 tx = tangent, ty = -binormal, tz = surface_normal;
    s.normal = mk3(tx.x * n.x + ty.x * n.y + tz.x * n.z, tx.y * n.x + ty.y * n.y + tz.y * n.z,
                   tx.z * n.x + ty.z * n.y + tz.z * n.z);
  }
}

__device__ __forceinline__ float3
env_lookup(cudaTextureObject_t cube, float3 dir)
{
  const float4 val = texCubemap<float4>(cube, dir.x, dir.y, -dir.z); // raytrace.cu:60,197
  return mk3(val.x, val.y, val.z);
}

__device__ __forceinline__ uint32_t
warp_sum(uint32_t v)
{
  return __reduce_add_sync(0xFFFFFFFFu, v);
}

// One radiance-loop iteration per live path (raytrace.cu:67-207), iteration `bounce`.
// A path that misses is finished here arithmetically: the reference keeps looping on
// the unchanged ray, re-adding the environment with renormalised throughput (Q5).
__global__ void __launch_bounds__(128)
k_shade(FrameParams fp, DScene sc, PathState st, cudaTextureObject_t cube,
        const uint32_t* __restrict__ queue, const uint32_t* __restrict__ queue_count, uint32_t n_max,
        uint32_t* __restrict__ queue_out, uint32_t* __restrict__ count_out, Counters* ctr, int bounce)
{
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t n = queue_count ? *queue_count : n_max;
  const bool valid = i < n;
  bool cont = false;
  uint32_t p = 0, segs = 0;

  if (valid) {
    p = queue ? queue[i] : i;
    const uint4 h = st.hit[p];
    const float4 o4 = st.ray_o[p], d4 = st.ray_d[p], t4 = st.thr[p], a4 = st.acc[p];
    Rng rng;
    rng_load(st, p, rng);
    float3 origin = mk3(o4.x, o4.y, o4.z), dir = mk3(d4.x, d4.y, d4.z);
    float3 throughput = mk3(t4.x, t4.y, t4.z), acc = mk3(a4.x, a4.y, a4.z);
    float spec_carry = o4.w;
    const float dist = __uint_as_float(h.x);
    const bool hit = h.y != 0xFFFFFFFFu;
    bool done = false;
    int b = bounce;

    float r1 = rng_uniform(rng); // raytrace.cu:70, drawn before the intersection
    segs = 1;
    if (hit) {
      Surface s;
      resolve_hit(sc, h.y, dist, __uint_as_float(h.z), __uint_as_float(h.w), dir, spec_carry, s);
      spec_carry = s.specular_col;
      const float cos_theta = dot(s.normal, dir);
      float3 oriented_normal = s.normal;
      const float3 spec = normalize(reflect(dir, s.normal));
      const float3 direct_light = s.diffuse_col / 0.5f; // brdf_lambert / pdf_lambert
      if (s.ior == 1.0f || s.is_light) {
        if (s.is_light)
          acc = acc + s.light_color * s.light_emission * throughput;
        const float phi = (float)((double)2.0f * 3.14159265358979323846 * (double)rng_uniform(rng));
        const float sin_t = __fsqrt_rn(r1);
        const float cos_t = __fsqrt_rn(1.f - r1);
        const float3 uu = normalize(cross((double)fabsf(oriented_normal.x) > .1 ? mk3(0.0f, 1.0f, 0.0f)
                                                                                : mk3(1.0f, 0.0f, 0.0f),
                                          oriented_normal));
        const float3 vv = cross(oriented_normal, uu);
        const float3 dd =
          normalize(vv * sin_t * __cosf(phi) + uu * __sinf(phi) * sin_t + oriented_normal * cos_t);
        origin = origin + dir * dist;
        dir = mix(dd, spec, s.specular_col); // not renormalised (Q8)
        origin = origin + dir * 0.03f;
        throughput = throughput * direct_light;
      } else { // refraction, raytrace.cu:135-193
        const float n1 = 1.0f, n2 = s.ior;
        oriented_normal = cos_theta < 0 ? s.normal : s.normal * -1.0f;
        const float c1 = dot(oriented_normal, dir);
        const bool entering = dot(s.normal, oriented_normal) > 0;
        const float eta = entering ? n1 / n2 : n2 / n1;
        const float eta_2 = eta * eta;
        const float c2_term = 1.0f - eta_2 * (1.0f - c1 * c1);
        if (c2_term < 0.0f) {
          origin = origin + oriented_normal * dist / 100.f; // origin not moved to the hit (Q10)
          dir = spec;
        } else {
          float R0 = (n2 - n1) / (n1 + n2);
          R0 *= R0;
          const float c2 = __fsqrt_rn(c2_term);
          const float3 T = normalize(eta * dir + (eta * c1 - c2) * oriented_normal);
          const float f_cos_theta = powf(cos_theta, 5.0f); // overwrites the Schlick term (Q9)
          const float f_r = R0 + (1.0f - R0) * f_cos_theta;
          if (rng_uniform(rng) < 0.25f) {
            throughput = throughput * (f_r * direct_light);
            origin = origin + oriented_normal * dist / 100.f;
            dir = spec;
          } else {
            const float f_t = 1.0f - f_r;
            throughput = throughput * (f_t * direct_light);
            origin = origin + oriented_normal * dist / 10000.f;
            dir = T;
          }
        }
      }
    } else {
      acc = acc + env_lookup(cube, dir) * throughput;
    }

    // roulette + renormalisation, raytrace.cu:201-206 (every iteration, Q6/Q7)
    float pmax = fmaxf(throughput.x, fmaxf(throughput.y, throughput.z));
    if (r1 > pmax && b > 1)
      done = true;
    else
      throughput = throughput * __fdividef(1.0f, pmax);

    if (!done && !hit) {
      // remaining iterations of a missed path: same ray, same miss
      const float3 env = env_lookup(cube, dir);
      for (b = bounce + 1; b < fp.max_bounces; ++b) {
        r1 = rng_uniform(rng);
        segs++;
        acc = acc + env * throughput;
        pmax = fmaxf(throughput.x, fmaxf(throughput.y, throughput.z));
        if (r1 > pmax && b > 1)
          break;
        throughput = throughput * __fdividef(1.0f, pmax);
      }
      done = true;
    }
    if (bounce + 1 >= fp.max_bounces)
      done = true;

    st.acc[p] = make_float4(acc.x, acc.y, acc.z, 0.f);
    if (!done) {
      st.ray_o[p] = make_float4(origin.x, origin.y, origin.z, spec_carry);
      st.ray_d[p] = make_float4(dir.x, dir.y, dir.z, 0.f);
      st.thr[p] = make_float4(throughput.x, throughput.y, throughput.z, 0.f);
      rng_store(st, p, rng);
      cont = true;
    }
  }

  // warp-aggregated append to the next queue
  const uint32_t mask = __ballot_sync(0xFFFFFFFFu, cont);
  const uint32_t lane = threadIdx.x & 31u;
  if (mask) {
    uint32_t base = 0;
    const uint32_t leader = __ffs(mask) - 1u;
    if (lane == leader)
      base = atomicAdd(count_out, (uint32_t)__popc(mask));
    base = __shfl_sync(0xFFFFFFFFu, base, leader);
    if (cont)
      queue_out[base + __popc(mask & ((1u << lane) - 1u))] = p;
  }
  const uint32_t wsegs = warp_sum(segs), wrays = warp_sum(valid ? 1u : 0u);
  if (lane == 0 && wrays) {
    atomicAdd(&ctr->segments, (unsigned long long)wsegs);
    atomicAdd(&ctr->traced, (unsigned long long)wrays);
    atomicAdd(&ctr->rays_per_bounce[bounce < 16 ? bounce : 15], (unsigned long long)wrays);
  }
}

// Preview frame (camera moved): one ray, flat colour, raytrace.cu:54-62.
__global__ void __launch_bounds__(128)
k_shade_preview(FrameParams fp, DScene sc, PathState st, cudaTextureObject_t cube, uint32_t n_max,
                Counters* ctr)
{
  const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  const bool valid = p < n_max;
  if (valid) {
    const uint4 h = st.hit[p];
    const float4 o4 = st.ray_o[p], d4 = st.ray_d[p];
    const float3 dir = mk3(d4.x, d4.y, d4.z);
    float3 rad;
    if (h.y != 0xFFFFFFFFu) {
      Surface s;
      resolve_hit(sc, h.y, __uint_as_float(h.x), __uint_as_float(h.z), __uint_as_float(h.w), dir,
                  o4.w, s);
      rad = s.diffuse_col;
    } else {
      rad = env_lookup(cube, dir);
    }
    st.acc[p] = make_float4(rad.x, rad.y, rad.z, 0.f);
  }
  const uint32_t w = warp_sum(valid ? 1u : 0u);
  if ((threadIdx.x & 31u) == 0 && w) {
    atomicAdd(&ctr->segments, (unsigned long long)w);
    atomicAdd(&ctr->traced, (unsigned long long)w);
    atomicAdd(&ctr->rays_per_bounce[0], (unsigned long long)w);
  }
}

// ---------------------------------------------------------------------------
// accumulate: raytrace.cu:248-270, post_process.cuh:14-41, raytrace.cu:327-357
// ---------------------------------------------------------------------------
__device__ __forceinline__ float3
uncharted_tonemap(float3 x)
{
  const float A = 0.15f, B = 0.50f, C = 0.10f, D = 0.20f, E = 0.02f, F = 0.30f;
  return ((x * (A * x + C * B) + D * E) / (x * (A * x + B) + D * F)) - E / F;
}

__device__ __forceinline__ float3
exposure(float3 color)
{
  const float exposure_bias = 2.0f;
  float3 curr = uncharted_tonemap(exposure_bias * color);
  float3 white_scale = 1.0f / uncharted_tonemap(mk3(11.2f, 11.2f, 11.2f));
  return curr * white_scale;
}

__device__ __forceinline__ float3
post_process(float3 c, uint32_t post_id)
{
  switch (post_id) {
    case 1: {
      const float gray = c.x * 0.3 + c.y * 0.59 + c.z * 0.11; // double literals, as the reference
      return mk3(gray, gray, gray);
    }
    case 2:
      return mk3(c.x * 0.393 + c.y * 0.769 + c.z * 0.189, c.x * 0.349 + c.y * 0.686 + c.z * 0.168,
                 c.x * 0.272 + c.y * 0.534 + c.z * 0.131);
    case 3:
      return mk3(1.0 - c.x, 1.0 - c.y, 1.0 - c.z);
    default:
      return c;
  }
}

__global__ void __launch_bounds__(256)
k_accumulate(FrameParams fp, PathState st, float* __restrict__ accum, uint32_t* __restrict__ out_rgba,
             float* __restrict__ radiance_out, Counters* ctr)
{
  const uint32_t band_h = fp.row_end - fp.row_begin;
  const uint32_t band_px = band_h * fp.width;
  const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= band_px)
    return;
  const uint32_t x = p % fp.width, ly = p / fp.width;
  const float4 a4 = st.acc[p];
  float3 rad = mk3(a4.x, a4.y, a4.z);
  if (radiance_out) {
    radiance_out[3 * (size_t)p] = rad.x;
    radiance_out[3 * (size_t)p + 1] = rad.y;
    radiance_out[3 * (size_t)p + 2] = rad.z;
  }
  rad = mk3(clamp_sel(rad.x, 0.0f, 1.0f), clamp_sel(rad.y, 0.0f, 1.0f), clamp_sel(rad.z, 0.0f, 1.0f));

  // bottom-up rows inside the band == (height - y - 1) * width + x of raytrace.cu:252
  float* a = accum + 3 * ((size_t)(band_h - ly - 1) * fp.width + x);
  const float fs = (float)fp.is_static;
  float3 t = mk3(a[0], a[1], a[2]);
  t = t * fs;
  t = t + rad;
  a[0] = t.x, a[1] = t.y, a[2] = t.z;
  if (!out_rgba)
    return;
  rad = t / (float)fp.frame_nb;
  rad = exposure(rad);
  rad = mk3(powf(rad.x, 1.0f / 2.2f), powf(rad.y, 1.0f / 2.2f), powf(rad.z, 1.0f / 2.2f));
  rad = post_process(rad, fp.post_id);
  const uint32_t R = (uint32_t)(rad.x * 255) & 0xFFu, G = (uint32_t)(rad.y * 255) & 0xFFu,
                 B = (uint32_t)(rad.z * 255) & 0xFFu;
  out_rgba[p] = R | (G << 8) | (B << 16); // a = 0, raytrace.cu:232
}

// ---------------------------------------------------------------------------
// parity / debug kernels
// ---------------------------------------------------------------------------
__global__ void
k_debug_rng(const uint32_t* __restrict__ seeds, uint32_t n_seeds, uint32_t n_draws,
            float* __restrict__ out)
{
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_seeds)
    return;
  Rng s;
  rng_init(s, seeds[i]);
  for (uint32_t k = 0; k < n_draws; ++k)
    out[(size_t)i * n_draws + k] = rng_uniform(s);
}

__global__ void
k_export_primary(PathState st, uint32_t n, int32_t* __restrict__ ids, float* __restrict__ ts,
                 float* __restrict__ origins, float* __restrict__ dirs)
{
  const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n)
    return;
  const uint4 h = st.hit[p];
  if (ids)
    ids[p] = (int32_t)h.y;
  if (ts)
    ts[p] = __uint_as_float(h.x);
  if (origins) {
    const float4 o = st.ray_o[p];
    origins[3 * (size_t)p] = o.x, origins[3 * (size_t)p + 1] = o.y, origins[3 * (size_t)p + 2] = o.z;
  }
  if (dirs) {
    const float4 d = st.ray_d[p];
    dirs[3 * (size_t)p] = d.x, dirs[3 * (size_t)p + 1] = d.y, dirs[3 * (size_t)p + 2] = d.z;
  }
}

} // namespace ptc

// ---------------------------------------------------------------------------
// launchers
// ---------------------------------------------------------------------------
#include <curand_kernel.h>

namespace ptc {

static __global__ void
k_debug_curand(const uint32_t* __restrict__ seeds, uint32_t n_seeds, uint32_t n_draws,
               float* __restrict__ out)
{
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_seeds)
    return;
  curandState st;
  curand_init(seeds[i], 0, 0, &st); // raytrace.cu:234-235
  for (uint32_t k = 0; k < n_draws; ++k)
    out[(size_t)i * n_draws + k] = curand_uniform(&st);
}

static inline uint32_t
blocks_for(uint64_t n, uint32_t block)
{
  return (uint32_t)((n + block - 1) / block);
}

void
launch_generate(const FrameParams& fp, const PathState& st, cudaStream_t s)
{
  const uint32_t n = (fp.row_end - fp.row_begin) * fp.width;
  if (n)
    k_generate<<<blocks_for(n, 256), 256, 0, s>>>(fp, st);
}

void
launch_extend(const DScene& sc, const PathState& st, const uint32_t* queue,
              const uint32_t* queue_count, uint32_t n_max, Counters* ctr, int bounce, cudaStream_t s)
{
  if (n_max)
    k_extend<<<blocks_for(n_max, 128), 128, 0, s>>>(sc, st, queue, queue_count, n_max, ctr, bounce);
}

void
launch_shade(const FrameParams& fp, const DScene& sc, const PathState& st, cudaTextureObject_t cube,
             const uint32_t* queue, const uint32_t* queue_count, uint32_t n_max, uint32_t* queue_out,
             uint32_t* count_out, Counters* ctr, int bounce, cudaStream_t s)
{
  if (n_max)
    k_shade<<<blocks_for(n_max, 128), 128, 0, s>>>(fp, sc, st, cube, queue, queue_count, n_max,
                                                   queue_out, count_out, ctr, bounce);
}

void
launch_shade_preview(const FrameParams& fp, const DScene& sc, const PathState& st,
                     cudaTextureObject_t cube, uint32_t n_max, Counters* ctr, cudaStream_t s)
{
  if (n_max)
    k_shade_preview<<<blocks_for(n_max, 128), 128, 0, s>>>(fp, sc, st, cube, n_max, ctr);
}

void
launch_accumulate(const FrameParams& fp, const PathState& st, float* accum, uint32_t* out_rgba,
                  float* radiance_out, Counters* ctr, cudaStream_t s)
{
  const uint32_t n = (fp.row_end - fp.row_begin) * fp.width;
  if (n)
    k_accumulate<<<blocks_for(n, 256), 256, 0, s>>>(fp, st, accum, out_rgba, radiance_out, ctr);
}

void
launch_intersect_rays(const DScene& sc, const float* origins, const float* dirs, uint64_t n,
                      int32_t* ids, float* ts, Counters* ctr, cudaStream_t s)
{
  if (n)
    k_intersect_rays<<<blocks_for(n, 128), 128, 0, s>>>(sc, origins, dirs, n, ids, ts, ctr);
}

void
launch_debug_rng(const uint32_t* seeds, uint32_t n_seeds, uint32_t n_draws, float* out, cudaStream_t s)
{
  if (n_seeds)
    k_debug_rng<<<blocks_for(n_seeds, 64), 64, 0, s>>>(seeds, n_seeds, n_draws, out);
}

void
launch_debug_curand(const uint32_t* seeds, uint32_t n_seeds, uint32_t n_draws, float* out,
                    cudaStream_t s)
{
  if (n_seeds)
    k_debug_curand<<<blocks_for(n_seeds, 64), 64, 0, s>>>(seeds, n_seeds, n_draws, out);
}

void
launch_export_primary(const PathState& st, uint32_t n, int32_t* ids, float* ts, float* origins,
                      float* dirs, cudaStream_t s)
{
  if (n)
    k_export_primary<<<blocks_for(n, 256), 256, 0, s>>>(st, n, ids, ts, origins, dirs);
}

} // namespace ptc
