// pt_oracle.cpp -- CPU ORACLE.  TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// A from-scratch, plain C++ restatement of the reference's per-pixel path loop
// (DavidPeicho/cuda-pathtracer).  Only tests/, __graft_entry__.smoke() and the
// cpu_baseline / --impl reference legs of bench.py may load this library, and
// only as the checker.  The product (cuda-pathtracer_b200/csrc) never links it.
//
// PINNING: every function below is checked bit-for-bit (tests/test_oracle.py)
// against oracle/_ref/libref_cpu.so -- the reference's OWN raytrace.cu and .cuh
// headers compiled for host cores (oracle/build_ref.py) -- and against the golden
// vectors that build froze under tests/golden/ (rng_kat.json, ref_indoor_*.npz).
// Both are compiled with -ffp-contract=off: each fp32 operation is rounded once,
// in the order the reference's C++ expressions evaluate.
//
// What cannot be bit-identical to a DEVICE run of the reference (stated once,
// here): libdevice vs libm for tanf/sinf/cosf/powf, the approximate intrinsics
// __fdividef/__sinf/__cosf (here: IEEE divide, libm), nvcc's fma contraction, and
// the texture unit's cubemap filtering.  Those are the tolerances the GPU parity
// tests state.
//
// Each function cites the reference lines it follows:
//   rt.cu  = cuda_opengl/src/shaders/raytrace.cu
//   ix.cuh = cuda_opengl/include/shaders/intersection.cuh
//   pp.cuh = cuda_opengl/include/shaders/post_process.cuh
//   brdf.cuh, cm.h = .../shaders/brdf.cuh, .../shaders/cutils_math.h
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

#include "ptcore.h"

namespace {

// ---------------------------------------------------------------------------
// fp32 vector helpers with the evaluation order of cm.h
// ---------------------------------------------------------------------------
struct V3
{
  float x, y, z;
};
inline V3 v3(float x, float y, float z) { return V3{ x, y, z }; }
inline V3 v3(const float* p) { return V3{ p[0], p[1], p[2] }; }
inline V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline V3 operator*(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline V3 operator*(V3 a, float b) { return v3(a.x * b, a.y * b, a.z * b); }
inline V3 operator*(float b, V3 a) { return v3(b * a.x, b * a.y, b * a.z); }
inline V3 operator/(V3 a, float b) { return v3(a.x / b, a.y / b, a.z / b); }
inline V3 operator/(V3 a, V3 b) { return v3(a.x / b.x, a.y / b.y, a.z / b.z); }
inline V3 operator/(float b, V3 a) { return v3(b / a.x, b / a.y, b / a.z); }
inline V3 operator+(V3 a, float b) { return v3(a.x + b, a.y + b, a.z + b); }
inline V3 operator-(V3 a, float b) { return v3(a.x - b, a.y - b, a.z - b); }
inline V3 neg(V3 a) { return v3(-a.x, -a.y, -a.z); }

// cm.h:1484
inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
// cm.h:1688-1692
inline V3 cross(V3 a, V3 b)
{
  return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
// cm.h:71-74, 1557-1561: v * (1.0f / sqrtf(dot(v,v)))
inline V3 normalize(V3 v)
{
  float inv_len = 1.0f / sqrtf(dot(v, v));
  return v * inv_len;
}
// cm.h:1678-1681
inline V3 reflect(V3 i, V3 n) { return i - 2.0f * n * dot(n, i); }
// cm.h:1722-1725
inline V3 mix(V3 x, V3 y, float a) { return (x * (1.0f - a)) + y * a; }
// cm.h:47-56, 1357-1360: compare-select clamp; NaN and +inf both become b (Q6)
inline float fminf_d(float a, float b) { return a < b ? a : b; }
inline float fmaxf_d(float a, float b) { return a > b ? a : b; }
inline float clampf(float f, float a, float b) { return fmaxf_d(a, fminf_d(f, b)); }

struct Ray
{
  V3 dir, origin; // sd.h:140-144
};

// ---------------------------------------------------------------------------
// RNG: cuRAND XORWOW as called by rt.cu:234-235 (subsequence = offset = 0)
// toolkit curand_kernel.h:772-799 (init), :863-874 (step), curand_uniform.h:69-72
// ---------------------------------------------------------------------------
struct Xorwow
{
  uint32_t v[5];
  uint32_t d;
};

inline void
xorwow_init(Xorwow& s, uint32_t seed32)
{
  // The reference passes an unsigned int (hash_seed + tid wraps in 32 bits), so
  // the high word of the 64-bit seed is zero.
  uint32_t s0 = seed32 ^ 0xaad26b49u;
  uint32_t s1 = 0u ^ 0xf7dcefddu;
  uint32_t t0 = 1099087573u * s0;
  uint32_t t1 = 2591861531u * s1;
  s.d = 6615241u + t1 + t0;
  s.v[0] = 123456789u + t0;
  s.v[1] = 362436069u ^ t0;
  s.v[2] = 521288629u + t1;
  s.v[3] = 88675123u ^ t1;
  s.v[4] = 5783321u + t0;
}

inline uint32_t
xorwow_next(Xorwow& s)
{
  uint32_t t = s.v[0] ^ (s.v[0] >> 2);
  s.v[0] = s.v[1];
  s.v[1] = s.v[2];
  s.v[2] = s.v[3];
  s.v[3] = s.v[4];
  s.v[4] = (s.v[4] ^ (s.v[4] << 4)) ^ (t ^ (t << 1));
  s.d += 362437u;
  return s.v[4] + s.d;
}

// (0,1]: x * 2^-32 + 2^-33
inline float
xorwow_uniform(Xorwow& s)
{
  return (float)xorwow_next(s) * 2.3283064e-10f + (2.3283064e-10f / 2.0f);
}

// rt.cu:275-285
inline uint32_t
wang_hash(uint32_t a)
{
  a = (a ^ 61u) ^ (a >> 16);
  a = a + (a << 3);
  a = a ^ (a >> 4);
  a = a * 0x27d4eb2du;
  a = a ^ (a >> 15);
  return a;
}

// rt.cu:227-229 with the reference's launch shape (16x16 blocks,
// gridDim.x = w/16 + 1, rt.cu:315-317).  Independent of how WE launch (F9, Q15).
inline uint32_t
reference_tid(uint32_t x, uint32_t y, uint32_t width)
{
  uint32_t gx = width / 16u + 1u;
  return ((x / 16u) + (y / 16u) * gx) * 256u + (y % 16u) * 16u + (x % 16u);
}

// ---------------------------------------------------------------------------
// Scene view over the flat C-ABI description (sd.h:31-133)
// ---------------------------------------------------------------------------
struct SceneView
{
  const ptc_scene_desc* d;
  uint32_t n_faces;
  const float* cube;  // 6 faces, RGBA float, size*size each (+x,-x,+y,-y,+z,-z)
  uint32_t cube_size;

  SceneView(const ptc_scene_desc* desc, const float* cube_faces, uint32_t cs)
    : d(desc), n_faces(0), cube(cube_faces), cube_size(cs)
  {
    for (uint32_t m = 0; m < d->n_meshes; ++m)
      n_faces += d->mesh_face_counts[m];
  }
};

// ix.cuh:7-18
struct Inter
{
  V3 normal, surface_normal, tangent, diffuse_col;
  float uv[2];
  int light; // index or -1 (reference: pointer / NULL)
  float dist, specular_col, ior;
};

// ix.cuh:102-135 -- Moeller-Trumbore, back-face culled (Q1); normal/uv are
// interpolated before the caller's distance test, as in the reference.
inline bool
intersect_triangle(const ptc_face& f, V3& out_normal, float out_uv[2], const Ray& ray, float& t)
{
  V3 v0 = v3(f.vertices[0]);
  V3 v0v1 = v3(f.vertices[1]) - v0;
  V3 v0v2 = v3(f.vertices[2]) - v0;
  V3 p_vec = cross(ray.dir, v0v2);
  float det = dot(v0v1, p_vec);
  if ((double)det < 0.0000001) // double literal in the reference
    return false;
  float inv_det = 1.f / det; // __fdividef(1.f, det) on the device
  V3 t_vec = ray.origin - v0;
  float u = dot(t_vec, p_vec) * inv_det;
  if (u < 0 || u > 1)
    return false;
  V3 qvec = cross(t_vec, v0v1);
  float v = dot(ray.dir, qvec) * inv_det;
  if (v < 0 || u + v > 1)
    return false;
  out_normal = (1.0f - u - v) * v3(f.normals[0]) + u * v3(f.normals[1]) + v * v3(f.normals[2]);
  float w0 = 1.0f - u - v;
  float ux = w0 * f.texcoords[0][0] + u * f.texcoords[1][0] + v * f.texcoords[2][0];
  float uy = w0 * f.texcoords[0][1] + u * f.texcoords[1][1] + v * f.texcoords[2][1];
  // mod(uv, 1.0): x - floorf(x / 1)  (cm.h:1728-1736)
  out_uv[0] = ux - floorf(ux / 1.0f);
  out_uv[1] = uy - floorf(uy / 1.0f);
  t = dot(v0v2, qvec) * inv_det;
  return true;
}

// ix.cuh:140-155
// NOTE on ix.cuh:152: `(t = b - disc) > eps ? t : ((t = b + disc) > eps ? t : 0);`
// discards the value of the conditional expression, so when both roots fail the
// epsilon test `t` keeps b + disc (not 0) and the function returns `t != 0`,
// i.e. TRUE for almost every ray whose line meets the sphere, with a possibly
// negative t.  The caller's `t >= 0` test (ix.cuh:203) then rejects negative t
// but accepts roots in [0, 0.01].  intersect_sphere_ref below states exactly that.
inline bool
intersect_sphere_ref(const Ray& r, const ptc_light& light, float& t)
{
  V3 op = v3(light.vec) - r.origin;
  float b = dot(op, r.dir);
  float disc = b * b - dot(op, op) + light.radius * light.radius;
  if (disc < 0.0)
    return false;
  disc = sqrtf(disc);
  t = b - disc;
  if (!(t > 0.01f))
    t = b + disc;
  return t != 0.0;
}

// ix.cuh:20-26
inline int
texture_idx(const ptc_texture& tex, const float uv[2])
{
  int x = (int)(uv[0] * (float)(tex.w - 1));
  int y = (int)(uv[1] * (float)(tex.h - 1));
  return (y * tex.w + x) * tex.nb_chan;
}

// Closest hit: ix.cuh:161-246.  `hit_id` (not in the reference) reports the
// winner: global face index, n_faces + light, or -1.
inline bool
intersect(const Ray& r, const SceneView& sc, Inter& in, int32_t* hit_id)
{
  const float MAX_DIST = 100000.0f;
  const ptc_scene_desc* d = sc.d;
  float inter_dist = MAX_DIST;
  in.dist = MAX_DIST;
  V3 normal = v3(0, 0, 0);
  float uv[2] = { 0, 0 };
  const ptc_material* inter_mat = nullptr;
  int32_t id = -1;

  const ptc_face* face = d->faces;
  int32_t g = 0;
  for (uint32_t m = 0; m < d->n_meshes; ++m)
    for (uint32_t i = 0; i < d->mesh_face_counts[m]; ++i, ++face, ++g)
      if (intersect_triangle(*face, normal, uv, r, inter_dist) && inter_dist < in.dist &&
          inter_dist > 0.0) { // strict '<': first index wins ties (Q2)
        inter_mat = &d->materials[face->material_id];
        in.ior = inter_mat->ior;
        in.normal = normal;
        in.surface_normal = normal;
        in.tangent = v3(face->tangent);
        in.uv[0] = uv[0];
        in.uv[1] = uv[1];
        in.dist = inter_dist;
        in.light = -1;
        id = g;
      }

  for (uint32_t l = 0; l < d->n_lights; ++l) {
    const ptc_light& light = d->lights[l];
    if (intersect_sphere_ref(r, light, inter_dist) && inter_dist < in.dist && inter_dist >= 0.0) {
      in.light = (int)l;
      in.dist = inter_dist;
      in.diffuse_col = v3(light.color);
      in.normal = normalize(v3(light.vec) - (inter_dist * r.dir)); // as written, ix.cuh:208
      inter_mat = nullptr;
      id = (int32_t)(sc.n_faces + l);
      // specular_col, ior, uv, tangent keep their previous values (Q3)
    }
  }

  if (inter_mat) {
    const ptc_texture& tex = d->textures[inter_mat->diffuse_spec_map];
    int idx = texture_idx(tex, in.uv);
    in.diffuse_col = v3(tex.data[idx], tex.data[idx + 1], tex.data[idx + 2]);
    in.specular_col = tex.data[idx + 3];
    if (inter_mat->normal_map >= 0) {
      const ptc_texture& nt = d->textures[inter_mat->normal_map];
      int nidx = texture_idx(nt, in.uv);
      V3 n = v3(nt.data[nidx], nt.data[nidx + 1], nt.data[nidx + 2]);
      in.normal = normalize((n * 2.0f) - 1.0f);
      V3 binormal = normalize(cross(in.tangent, in.surface_normal));
      // tbn columns: tangent, -binormal, surface_normal (cm.h:1134-1139)
      V3 tx = in.tangent, ty = neg(binormal), tz = in.surface_normal, a = in.normal;
      in.normal = v3(tx.x * a.x + ty.x * a.y + tz.x * a.z, tx.y * a.x + ty.y * a.y + tz.y * a.z,
                     tx.z * a.x + ty.z * a.y + tz.z * a.z);
    }
  }
  if (hit_id)
    *hit_id = id;
  return in.dist < MAX_DIST;
}

// Software cubemap fetch standing in for texCubemap (rt.cu:60,197,304-309):
// same face table / bilinear rule as oracle/ref_driver/ref_cpu_driver.cpp.
inline void
env_lookup(const SceneView& sc, float x, float y, float z, float out[4])
{
  const float ax = fabsf(x), ay = fabsf(y), az = fabsf(z);
  int face;
  float m, s, t;
  if (ax >= ay && ax >= az) {
    face = x >= 0 ? 0 : 1;
    m = ax;
    s = x >= 0 ? -z : z;
    t = -y;
  } else if (ay >= az) {
    face = y >= 0 ? 2 : 3;
    m = ay;
    s = x;
    t = y >= 0 ? z : -z;
  } else {
    face = z >= 0 ? 4 : 5;
    m = az;
    s = z >= 0 ? x : -x;
    t = -y;
  }
  const uint32_t N = sc.cube_size;
  const float* base = sc.cube + (size_t)face * N * N * 4;
  float u = (s / m + 1.0f) * 0.5f, v = (t / m + 1.0f) * 0.5f;
  float xb = u * N - 0.5f, yb = v * N - 0.5f;
  float fx = floorf(xb), fy = floorf(yb);
  float a = floorf((xb - fx) * 256.0f + 0.5f) / 256.0f;
  float b = floorf((yb - fy) * 256.0f + 0.5f) / 256.0f;
  auto clampi = [N](int i) { return i < 0 ? 0 : (i >= (int)N ? (int)N - 1 : i); };
  int x0 = clampi((int)fx), x1 = clampi((int)fx + 1);
  int y0 = clampi((int)fy), y1 = clampi((int)fy + 1);
  for (int c = 0; c < 4; ++c) {
    float t00 = base[((size_t)y0 * N + x0) * 4 + c], t10 = base[((size_t)y0 * N + x1) * 4 + c];
    float t01 = base[((size_t)y1 * N + x0) * 4 + c], t11 = base[((size_t)y1 * N + x1) * 4 + c];
    out[c] = (1 - a) * (1 - b) * t00 + a * (1 - b) * t10 + (1 - a) * b * t01 + a * b * t11;
  }
}

// ---------------------------------------------------------------------------
// Camera (ix.cuh:75-97) and thin-lens depth of field (pp.cuh:49-67)
// ---------------------------------------------------------------------------
struct Cam
{
  V3 position, dir, u, v;
  float fov_x, aperture, focus_dist;
};

inline Cam
to_cam(const ptc_camera* c)
{
  Cam cam;
  cam.position = v3(c->position);
  cam.dir = v3(c->dir);
  cam.u = v3(c->u);
  cam.v = v3(c->v);
  cam.fov_x = c->fov_x;
  cam.aperture = c->aperture;
  cam.focus_dist = c->focus_dist;
  return cam;
}

inline Ray
generate_ray(int x, int y, int half_w, int half_h, Cam& cam)
{
  float screen_dist = half_w / tanf(cam.fov_x * 0.5f);
  Ray ray;
  ray.origin = cam.position;
  cam.u = normalize(cross(cam.dir, v3(0.0f, -1.0f, 0.0f)));
  cam.v = normalize(cross(cam.u, cam.dir));
  cam.u = cam.u * -1.0f;
  V3 screen_pos = cam.position + (cam.dir * screen_dist) + (cam.u * (float)(x - half_w)) +
                  (cam.v * (float)(y - half_h));
  ray.dir = screen_pos - cam.position;
  ray.dir = normalize(ray.dir);
  return ray; // no +0.5, no jitter (Q14)
}

inline void
camera_dof(Ray& r, const Cam& cam, Xorwow& rng)
{
  V3 focal_point = cam.focus_dist * r.dir;
  // float * 2.0f in fp32, then * M_PI in fp64 (M_PI is the <math.h> double), rounded to float
  float random_angle = (float)((double)(xorwow_uniform(rng) * 2.0f) * M_PI);
  float random_radius = xorwow_uniform(rng) * cam.aperture;
  V3 random_aperture_pos = (cosf(random_angle) * cam.u + sinf(random_angle) * cam.v) * random_radius;
  V3 final_ray_dir = normalize(focal_point - random_aperture_pos);
  r.origin = r.origin + random_aperture_pos;
  r.dir = final_ray_dir;
}

// ---------------------------------------------------------------------------
// The path loop: rt.cu:41-210
// ---------------------------------------------------------------------------
struct PathStats
{
  uint64_t segments = 0; // loop iterations == reference intersect() calls
  uint64_t traced = 0;   // iterations that still needed a closest-hit query
};

inline V3
radiance(Ray& r, const SceneView& sc, Xorwow& rng, int is_static, int static_samples,
         PathStats* stats, int32_t* first_hit)
{
  V3 acc = v3(0, 0, 0);
  V3 throughput = v3(1, 1, 1);
  Inter inter;
  std::memset(&inter, 0, sizeof(inter)); // Q3: defined as zero-initialised
  inter.light = -1;

  if (!is_static) { // preview, rt.cu:54-62
    int32_t id;
    bool hit = intersect(r, sc, inter, &id);
    if (first_hit)
      *first_hit = id;
    if (stats)
      stats->segments++, stats->traced++;
    if (hit)
      return inter.diffuse_col;
    float val[4];
    env_lookup(sc, r.dir.x, r.dir.y, -r.dir.z, val);
    return v3(val[0], val[1], val[2]);
  }

  const int max_bounces = 1 + is_static * (static_samples + 1);
  bool missed = false;
  for (int b = 0; b < max_bounces; b++) {
    float r1 = xorwow_uniform(rng);
    int32_t id = -1;
    // After a miss the ray is unchanged, so the reference's next intersect() misses
    // again (Q5): no query is needed, which is what `traced` counts.
    bool hit = missed ? false : intersect(r, sc, inter, &id);
    if (stats) {
      stats->segments++;
      if (!missed)
        stats->traced++;
    }
    if (b == 0 && first_hit)
      *first_hit = id;
    if (hit) {
      float cos_theta = dot(inter.normal, r.dir);
      V3 oriented_normal = inter.normal;
      V3 spec = normalize(reflect(r.dir, inter.normal));
      V3 direct_light = inter.diffuse_col / 0.5f; // brdf_lambert / pdf_lambert, brdf.cuh:14-31
      if (inter.ior == 1.0f || inter.light >= 0) {
        if (inter.light >= 0) {
          const ptc_light& L = sc.d->lights[inter.light];
          V3 BRDF = v3(L.color);
          acc = acc + BRDF * L.emission * throughput;
        }
        float phi = (float)((double)2.0f * M_PI * (double)xorwow_uniform(rng));
        float sin_t = sqrtf(r1);
        float cos_t = sqrtf(1.f - r1);
        V3 u = normalize(cross((double)fabsf(oriented_normal.x) > .1 ? v3(0.0f, 1.0f, 0.0f)
                                                                     : v3(1.0f, 0.0f, 0.0f),
                               oriented_normal));
        V3 v = cross(oriented_normal, u);
        V3 d = normalize(v * sin_t * cosf(phi) + u * sinf(phi) * sin_t + oriented_normal * cos_t);
        r.origin = r.origin + r.dir * inter.dist;
        r.dir = mix(d, spec, inter.specular_col); // not renormalised (Q8)
        r.origin = r.origin + r.dir * 0.03f;
        throughput = throughput * direct_light;
      } else {
        float n1 = 1.0f;
        float n2 = inter.ior;
        oriented_normal = cos_theta < 0 ? inter.normal : inter.normal * -1.0f;
        float c1 = dot(oriented_normal, r.dir);
        bool entering = dot(inter.normal, oriented_normal) > 0;
        float eta = entering ? n1 / n2 : n2 / n1;
        float eta_2 = eta * eta;
        float c2_term = 1.0f - eta_2 * (1.0f - c1 * c1);
        if (c2_term < 0.0f) { // total internal reflection; origin not moved to the hit (Q10)
          r.origin = r.origin + oriented_normal * inter.dist / 100.f;
          r.dir = spec;
        } else {
          float R0 = (n2 - n1) / (n1 + n2);
          R0 *= R0;
          float c2 = sqrtf(c2_term);
          V3 T = normalize(eta * r.dir + (eta * c1 - c2) * oriented_normal);
          float f_cos_theta = powf(cos_theta, 5.0f); // overwrites the Schlick term (Q9)
          float f_r = R0 + (1.0f - R0) * f_cos_theta;
          if (xorwow_uniform(rng) < 0.25f) {
            throughput = throughput * (f_r * direct_light);
            r.origin = r.origin + oriented_normal * inter.dist / 100.f;
            r.dir = spec;
          } else {
            float f_t = 1.0f - f_r;
            throughput = throughput * (f_t * direct_light);
            r.origin = r.origin + oriented_normal * inter.dist / 10000.f;
            r.dir = T;
          }
        }
      }
    } else {
      missed = true;
      float val[4];
      env_lookup(sc, r.dir.x, r.dir.y, -r.dir.z, val);
      acc = acc + v3(val[0], val[1], val[2]) * throughput;
    }
    float p = fmaxf(throughput.x, fmaxf(throughput.y, throughput.z));
    if (r1 > p && b > 1) // Q7
      return acc;
    throughput = throughput * (1.0f / p); // __fdividef(1.0f, p), every iteration (Q6)
  }
  return acc;
}

// ---------------------------------------------------------------------------
// Tonemap / gamma / post-process: pp.cuh:14-41, rt.cu:260-268, 327-357
// ---------------------------------------------------------------------------
inline V3
uncharted_tonemap(V3 x)
{
  const float A = 0.15f, B = 0.50f, C = 0.10f, D = 0.20f, E = 0.02f, F = 0.30f;
  return ((x * (A * x + C * B) + D * E) / (x * (A * x + B) + D * F)) - E / F;
}

inline V3
exposure(V3 color)
{
  const float exposure_bias = 2.0f;
  V3 curr = uncharted_tonemap(exposure_bias * color);
  V3 white_scale = 1.0f / uncharted_tonemap(v3(11.2f, 11.2f, 11.2f));
  return curr * white_scale;
}

inline V3
post_process(V3 c, unsigned post_id)
{
  switch (post_id) {
    case 1: { // grayscale, double literals
      float gray = (float)(c.x * 0.3 + c.y * 0.59 + c.z * 0.11);
      return v3(gray, gray, gray);
    }
    case 2: // sepia
      return v3((float)(c.x * 0.393 + c.y * 0.769 + c.z * 0.189),
                (float)(c.x * 0.349 + c.y * 0.686 + c.z * 0.168),
                (float)(c.x * 0.272 + c.y * 0.534 + c.z * 0.131));
    case 3: // invert
      return v3((float)(1.0 - c.x), (float)(1.0 - c.y), (float)(1.0 - c.z));
    default:
      return c;
  }
}

// One pixel of rt.cu:212-271.  accum is the full-image float3 buffer.
inline void
pixel(const SceneView& sc, const ptc_camera* camera, uint32_t width, uint32_t height, uint32_t x,
      uint32_t y, uint32_t hash_seed, int frame_nb, int bounces, bool moved, unsigned post_id,
      float* accum, uint32_t* out_rgba, float* radiance_out, int32_t* first_hit, PathStats* stats)
{
  Xorwow rng;
  xorwow_init(rng, hash_seed + reference_tid(x, y, width));
  Cam cam = to_cam(camera);
  Ray r = generate_ray((int)x, (int)y, (int)(width / 2), (int)(height / 2), cam);
  camera_dof(r, cam, rng);
  int is_static = !moved;
  V3 rad = radiance(r, sc, rng, is_static, bounces - 2, stats, first_hit);
  if (radiance_out) {
    float* o = radiance_out + ((size_t)y * width + x) * 3;
    o[0] = rad.x, o[1] = rad.y, o[2] = rad.z;
  }
  if (!accum)
    return;
  rad = v3(clampf(rad.x, 0.0f, 1.0f), clampf(rad.y, 0.0f, 1.0f), clampf(rad.z, 0.0f, 1.0f)); // Q12
  size_t i = (size_t)(height - y - 1) * width + x; // bottom-up rows, rt.cu:252
  float* a = accum + i * 3;
  float fs = (float)is_static;
  a[0] = a[0] * fs, a[1] = a[1] * fs, a[2] = a[2] * fs;
  a[0] += rad.x, a[1] += rad.y, a[2] += rad.z;
  rad = v3(a[0], a[1], a[2]) / (float)frame_nb;
  rad = exposure(rad);
  rad = v3(powf(rad.x, 1.0f / 2.2f), powf(rad.y, 1.0f / 2.2f), powf(rad.z, 1.0f / 2.2f));
  rad = post_process(rad, post_id);
  if (out_rgba) {
    uint32_t R = (uint32_t)(rad.x * 255) & 0xFFu, G = (uint32_t)(rad.y * 255) & 0xFFu,
             B = (uint32_t)(rad.z * 255) & 0xFFu;
    out_rgba[(size_t)y * width + x] = R | (G << 8) | (B << 16); // a = 0
  }
}

template <typename F>
void
parallel_for(unsigned n, int n_threads, F&& body)
{
  int nt = n_threads > 0 ? n_threads : 1;
  if (nt == 1 || n < 2) {
    for (unsigned i = 0; i < n; ++i)
      body(i);
    return;
  }
  std::atomic<unsigned> next(0);
  std::vector<std::thread> pool;
  for (int t = 0; t < nt; ++t)
    pool.emplace_back([&] {
      for (;;) {
        unsigned i = next.fetch_add(1);
        if (i >= n)
          break;
        body(i);
      }
    });
  for (auto& th : pool)
    th.join();
}

} // namespace

// ---------------------------------------------------------------------------
// C entry points (same argument lists as oracle/ref_driver/ref_cpu_driver.cpp)
// ---------------------------------------------------------------------------
extern "C" {

// One accumulate-into-framebuffer launch for sample `frame_nb` (rt.cu:212-271,
// seeded as rt.cu:296-300,321 would for the frame_nb-th static frame).
int
oracle_kernel(const ptc_scene_desc* desc, const float* cube_faces, unsigned cube_size,
              const ptc_camera* camera, unsigned width, unsigned height, unsigned frame_nb,
              unsigned bounces, int moved, unsigned post_id, float* accum, uint32_t* out_rgba,
              int n_threads)
{
  SceneView sc(desc, cube_faces, cube_size);
  const uint32_t hash_seed = wang_hash(frame_nb);
  parallel_for(height, n_threads, [&](unsigned y) {
    for (unsigned x = 0; x < width; ++x)
      pixel(sc, camera, width, height, x, y, hash_seed, (int)frame_nb, (int)bounces, moved != 0,
            post_id, accum, out_rgba, nullptr, nullptr, nullptr);
  });
  return 0;
}

// Bounded SAMPLE of one launch, for timing on host cores (bench.py's cpu_baseline leg):
// only the reference's 16x16 blocks b = by * (width/16 + 1) + bx with
// b % block_stride == block_offset are rendered (rt.cu:315-317 defines the blocks).
// Reports the pixels rendered and the radiance-loop iterations (segments) they took.
int
oracle_kernel_sample(const ptc_scene_desc* desc, const float* cube_faces, unsigned cube_size,
                     const ptc_camera* camera, unsigned width, unsigned height, unsigned frame_nb,
                     unsigned bounces, unsigned block_stride, unsigned block_offset, float* accum,
                     uint32_t* out_rgba, uint64_t* pixels, uint64_t* segments, int n_threads)
{
  SceneView sc(desc, cube_faces, cube_size);
  const uint32_t hash_seed = wang_hash(frame_nb);
  const unsigned gx = width / 16 + 1, gy = height / 16 + 1;
  if (!block_stride)
    block_stride = 1;
  std::vector<unsigned> blocks;
  for (unsigned b = block_offset % block_stride; b < gx * gy; b += block_stride)
    blocks.push_back(b);
  std::vector<PathStats> st(blocks.size());
  std::vector<uint64_t> px(blocks.size(), 0);
  parallel_for((unsigned)blocks.size(), n_threads, [&](unsigned k) {
    const unsigned bx = blocks[k] % gx, by = blocks[k] / gx;
    for (unsigned ty = 0; ty < 16; ++ty)
      for (unsigned tx = 0; tx < 16; ++tx) {
        const unsigned x = bx * 16 + tx, y = by * 16 + ty;
        if (x >= width || y >= height)
          continue;
        pixel(sc, camera, width, height, x, y, hash_seed, (int)frame_nb, (int)bounces, false, 0,
              accum, out_rgba, nullptr, nullptr, &st[k]);
        px[k]++;
      }
  });
  uint64_t s = 0, n = 0;
  for (size_t k = 0; k < blocks.size(); ++k)
    s += st[k].segments, n += px[k];
  if (pixels)
    *pixels = n;
  if (segments)
    *segments = s;
  return 0;
}

// Per-pixel un-clamped radiance (row 0 = top).
int
oracle_radiance(const ptc_scene_desc* desc, const float* cube_faces, unsigned cube_size,
                const ptc_camera* camera, unsigned width, unsigned height, unsigned frame_nb,
                unsigned bounces, int moved, float* rad_out, int n_threads)
{
  SceneView sc(desc, cube_faces, cube_size);
  const uint32_t hash_seed = wang_hash(frame_nb);
  parallel_for(height, n_threads, [&](unsigned y) {
    for (unsigned x = 0; x < width; ++x)
      pixel(sc, camera, width, height, x, y, hash_seed, (int)frame_nb, (int)bounces, moved != 0, 0,
            nullptr, nullptr, rad_out, nullptr, nullptr);
  });
  return 0;
}

// Same as oracle_radiance, additionally reporting the primary hit id per pixel
// and the segment / traced ray counts (the units of BASELINE.json's Mrays/s).
int
oracle_render_stats(const ptc_scene_desc* desc, const float* cube_faces, unsigned cube_size,
                    const ptc_camera* camera, unsigned width, unsigned height, unsigned frame_nb,
                    unsigned bounces, float* rad_out, int32_t* first_hit, uint64_t* segments,
                    uint64_t* traced, int n_threads)
{
  SceneView sc(desc, cube_faces, cube_size);
  const uint32_t hash_seed = wang_hash(frame_nb);
  std::vector<PathStats> rows(height);
  parallel_for(height, n_threads, [&](unsigned y) {
    for (unsigned x = 0; x < width; ++x)
      pixel(sc, camera, width, height, x, y, hash_seed, (int)frame_nb, (int)bounces, false, 0,
            nullptr, nullptr, rad_out, first_hit ? first_hit + (size_t)y * width + x : nullptr,
            &rows[y]);
  });
  uint64_t s = 0, t = 0;
  for (const PathStats& p : rows)
    s += p.segments, t += p.traced;
  if (segments)
    *segments = s;
  if (traced)
    *traced = t;
  return 0;
}

// Brute-force closest hit for arbitrary rays: ix.cuh:161-212 (O1 -- defines hit ids).
int
oracle_intersect(const ptc_scene_desc* desc, const float* origins, const float* dirs, uint64_t n,
                 int32_t* hit_ids, float* t_out, int n_threads)
{
  SceneView sc(desc, nullptr, 0);
  const unsigned chunk = 4096;
  parallel_for((unsigned)((n + chunk - 1) / chunk), n_threads, [&](unsigned c) {
    uint64_t lo = (uint64_t)c * chunk, hi = lo + chunk < n ? lo + chunk : n;
    for (uint64_t k = lo; k < hi; ++k) {
      Ray r;
      r.origin = v3(origins + 3 * k);
      r.dir = v3(dirs + 3 * k);
      // geometry only: same acceptance rules as intersect(), no material resolve
      const float MAX_DIST = 100000.0f;
      float dist = MAX_DIST, inter_dist = MAX_DIST;
      int32_t id = -1, g = 0;
      V3 nrm;
      float uv[2];
      const ptc_face* face = desc->faces;
      for (uint32_t m = 0; m < desc->n_meshes; ++m)
        for (uint32_t i = 0; i < desc->mesh_face_counts[m]; ++i, ++face, ++g)
          if (intersect_triangle(*face, nrm, uv, r, inter_dist) && inter_dist < dist &&
              inter_dist > 0.0) {
            dist = inter_dist;
            id = g;
          }
      for (uint32_t l = 0; l < desc->n_lights; ++l)
        if (intersect_sphere_ref(r, desc->lights[l], inter_dist) && inter_dist < dist &&
            inter_dist >= 0.0) {
          dist = inter_dist;
          id = (int32_t)(sc.n_faces + l);
        }
      if (hit_ids)
        hit_ids[k] = id;
      if (t_out)
        t_out[k] = dist;
    }
  });
  return 0;
}

// Primary rays of sample `frame_nb` (rt.cu:227-240), row 0 = top.
int
oracle_primary_rays(const ptc_camera* camera, unsigned width, unsigned height, unsigned frame_nb,
                    float* origins, float* dirs)
{
  const uint32_t hash_seed = wang_hash(frame_nb);
  for (unsigned y = 0; y < height; ++y)
    for (unsigned x = 0; x < width; ++x) {
      Xorwow rng;
      xorwow_init(rng, hash_seed + reference_tid(x, y, width));
      Cam cam = to_cam(camera);
      Ray r = generate_ray((int)x, (int)y, (int)(width / 2), (int)(height / 2), cam);
      camera_dof(r, cam, rng);
      size_t k = (size_t)y * width + x;
      origins[3 * k] = r.origin.x, origins[3 * k + 1] = r.origin.y, origins[3 * k + 2] = r.origin.z;
      dirs[3 * k] = r.dir.x, dirs[3 * k + 1] = r.dir.y, dirs[3 * k + 2] = r.dir.z;
    }
  return 0;
}

// ---------------------------------------------------------------------------
// O2: closest hit through the PRODUCT's wide BVH (nodes/triangle records built by
// ptc_host_bvh_build and passed in by the test harness), traversed here by
// independent scalar code in the same near-to-far slot order as the device kernel,
// so the node / triangle visit counts are the algorithmic traffic the roofline uses.
// The triangle arithmetic is the oracle's own (unfused) Moeller-Trumbore.
// ---------------------------------------------------------------------------
namespace {

struct O2Node // mirrors the 128-byte layout documented in cuda-pathtracer_b200/csrc/bvh.h
{
  float origin[3];
  uint8_t e[3];
  uint8_t imask;
  uint32_t child_base;
  uint32_t tri_base;
  uint32_t valid24;
  uint32_t occupied;
  uint16_t lo[3][8]; // half floats holding the 8-bit quantised planes
  uint16_t hi[3][8];
};
static_assert(sizeof(O2Node) == 128, "node layout");

// IEEE half -> float (normal numbers and zero are all the node planes use)
inline float
o2_half(uint16_t h)
{
  const uint32_t e = (h >> 10) & 31u, m = h & 1023u;
  if (e == 0)
    return (float)m * (1.0f / 16777216.0f); // subnormal (only 0 occurs)
  return std::ldexp((float)(1024u + m), (int)e - 25);
}

struct O2Tri
{
  float v0[3], e1[3], e2[3];
  uint32_t face_id;
  uint32_t pad[2];
};
static_assert(sizeof(O2Tri) == 48, "tri layout");

inline bool
o2_tri(const O2Tri& T, const Ray& ray, float& t)
{
  V3 v0 = v3(T.v0), v0v1 = v3(T.e1), v0v2 = v3(T.e2);
  V3 p_vec = cross(ray.dir, v0v2);
  float det = dot(v0v1, p_vec);
  if ((double)det < 0.0000001)
    return false;
  float inv_det = 1.f / det;
  V3 t_vec = ray.origin - v0;
  float u = dot(t_vec, p_vec) * inv_det;
  if (u < 0 || u > 1)
    return false;
  V3 qvec = cross(t_vec, v0v1);
  float v = dot(ray.dir, qvec) * inv_det;
  if (v < 0 || u + v > 1)
    return false;
  t = dot(v0v2, qvec) * inv_det;
  return true;
}

inline void
o2_closest(const O2Node* nodes, const O2Tri* tris, const ptc_scene_desc* desc, uint32_t n_faces,
           const Ray& r, int32_t& best_id, float& best_t, uint64_t& n_nodes, uint64_t& n_tris)
{
  best_t = 100000.0f;
  best_id = -1;
  const float eps = 1e-20f;
  float idir[3], o[3] = { r.origin.x, r.origin.y, r.origin.z }, d[3] = { r.dir.x, r.dir.y, r.dir.z };
  uint32_t oct = 0;
  for (int a = 0; a < 3; ++a) {
    idir[a] = 1.0f / (fabsf(d[a]) > eps ? d[a] : copysignf(eps, d[a]));
    if (d[a] < 0.f)
      oct |= 1u << a;
  }
  const uint32_t oct_inv = 7u - oct;
  struct Entry
  {
    uint32_t base, mask;
  };
  Entry stack[64];
  int sp = 0;
  Entry ng = { 0u, 0x80000000u }, tg = { 0u, 0u };
  for (;;) {
    if (ng.mask & 0xFF000000u) {
      uint32_t bit = 31u - (uint32_t)__builtin_clz(ng.mask);
      ng.mask &= ~(1u << bit);
      uint32_t imask = ng.mask & 0xFFu;
      if (ng.mask & 0xFF000000u)
        stack[sp++] = ng;
      uint32_t slot = (bit - 24u) ^ oct_inv;
      uint32_t rel = (uint32_t)__builtin_popcount(imask & ~(0xFFFFFFFFu << slot));
      const O2Node& n = nodes[ng.base + rel];
      n_nodes++;
      uint32_t hit_inner = 0, hit_tris = 0;
      for (int i = 0; i < 8; ++i) {
        float tmin = 0.0f, tmax = best_t;
        for (int a = 0; a < 3; ++a) {
          uint32_t bits = (uint32_t)n.e[a] << 23;
          float sc;
          std::memcpy(&sc, &bits, 4);
          float lo = n.origin[a] + o2_half(n.lo[a][i]) * sc, hi = n.origin[a] + o2_half(n.hi[a][i]) * sc;
          if (lo > hi) { // empty slot (inverted box)
            tmin = 1.0f, tmax = 0.0f;
            break;
          }
          float t0 = (lo - o[a]) * idir[a], t1 = (hi - o[a]) * idir[a];
          if (t0 > t1)
            std::swap(t0, t1);
          // widen by a few ulp: this checker only has to be conservative
          t0 -= fabsf(t0) * 1e-6f;
          t1 += fabsf(t1) * 1e-6f;
          tmin = fmaxf(tmin, t0);
          tmax = fminf(tmax, t1);
        }
        if (tmin <= tmax) {
          if ((n.imask >> i) & 1u)
            hit_inner |= 1u << (24u + ((uint32_t)i ^ oct_inv));
          else
            hit_tris |= (7u << (3 * i)) & n.valid24;
        }
      }
      ng.base = n.child_base;
      ng.mask = hit_inner | n.imask;
      // dense record numbers of the hit triangles: popcount of the valid bits below
      tg.base = n.tri_base;
      tg.mask = 0;
      for (uint32_t b = 0; b < 24; ++b)
        if ((hit_tris >> b) & 1u)
          tg.mask |= 1u << (uint32_t)__builtin_popcount(n.valid24 & ((1u << b) - 1u));
    } else {
      tg = ng;
      ng = Entry{ 0u, 0u };
    }
    while (tg.mask) {
      uint32_t bit = (uint32_t)__builtin_ctz(tg.mask);
      tg.mask &= tg.mask - 1u;
      const O2Tri& T = tris[tg.base + bit];
      n_tris++;
      float t;
      if (o2_tri(T, r, t) && t > 0.0 &&
          (t < best_t || (t == best_t && (int32_t)T.face_id < best_id))) {
        best_t = t;
        best_id = (int32_t)T.face_id;
      }
    }
    if (!(ng.mask & 0xFF000000u)) {
      if (sp == 0)
        break;
      ng = stack[--sp];
    }
  }
  for (uint32_t l = 0; l < desc->n_lights; ++l) {
    float t;
    if (intersect_sphere_ref(r, desc->lights[l], t) && t < best_t && t >= 0.0) {
      best_t = t;
      best_id = (int32_t)(n_faces + l);
    }
  }
}

} // namespace

int
oracle_intersect_bvh(const ptc_scene_desc* desc, const void* nodes, const void* tris,
                     const float* origins, const float* dirs, uint64_t n, int32_t* hit_ids,
                     float* t_out, uint64_t* nodes_visited, uint64_t* tris_tested, int n_threads)
{
  SceneView sc(desc, nullptr, 0);
  const unsigned chunk = 4096;
  const unsigned n_chunks = (unsigned)((n + chunk - 1) / chunk);
  std::vector<uint64_t> nn(n_chunks, 0), nt(n_chunks, 0);
  parallel_for(n_chunks, n_threads, [&](unsigned c) {
    uint64_t lo = (uint64_t)c * chunk, hi = lo + chunk < n ? lo + chunk : n;
    for (uint64_t k = lo; k < hi; ++k) {
      Ray r;
      r.origin = v3(origins + 3 * k);
      r.dir = v3(dirs + 3 * k);
      int32_t id;
      float t;
      o2_closest((const O2Node*)nodes, (const O2Tri*)tris, desc, sc.n_faces, r, id, t, nn[c], nt[c]);
      if (hit_ids)
        hit_ids[k] = id;
      if (t_out)
        t_out[k] = t;
    }
  });
  uint64_t a = 0, b = 0;
  for (unsigned c = 0; c < n_chunks; ++c)
    a += nn[c], b += nt[c];
  if (nodes_visited)
    *nodes_visited = a;
  if (tris_tested)
    *tris_tested = b;
  return 0;
}

void
oracle_uniforms(uint32_t seed, unsigned n, float* out)
{
  Xorwow s;
  xorwow_init(s, seed);
  for (unsigned i = 0; i < n; ++i)
    out[i] = xorwow_uniform(s);
}

unsigned
oracle_wang_hash(unsigned a)
{
  return wang_hash(a);
}

} // extern "C"
