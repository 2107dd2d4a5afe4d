// shade.cuh -- material resolve, BRDF sampling and the per-pixel epilogue (device functions).
//
//   resolve_hit  material / normal-map resolve of the winning hit (intersection.cuh:183-243)
//   shade_step   ONE iteration of the radiance loop after its intersection
//                (raytrace.cu:67-207), including the arithmetic tail of a missed path (Q5)
//   epilogue     clamp, temporal accumulation, tonemap, gamma, post, RGBA8
//                (raytrace.cu:248-270, post_process.cuh:14-41, raytrace.cu:327-357)
//
// Expression forms are kept exactly as the verified first version had them: nvcc's fma
// contraction of these expressions is what makes the accumulator bit-identical to the
// reference kernel built for the same GPU.
#pragma once

#include "traverse.cuh"

namespace ptc {

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
texel_index(int w, int h, int nb_chan, float2 uv)
{
  int x = (int)(uv.x * (float)(w - 1));
  int y = (int)(uv.y * (float)(h - 1));
  return (y * w + x) * nb_chan;
}

// Material / normal resolve for the winning hit: intersection.cuh:183-243.
// `spec_carry` is IntersectionData.specular_col, which a light hit leaves untouched (Q3).
__device__ __forceinline__ void
resolve_hit(const DScene& sc, uint32_t id, float t, float u, float v, float3 dir, float spec_carry,
            Surface& s)
{
  if (id >= sc.n_faces) { // sphere light, intersection.cuh:199-212
    const ptc_light L = load_light(sc, id - sc.n_faces);
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

  const float4* mp = reinterpret_cast<const float4*>(sc.mats + material_id);
  const float4 m0 = __ldg(mp), m1 = __ldg(mp + 1), m2 = __ldg(mp + 2);
  const float* diffuse = reinterpret_cast<const float*>(
    (uintptr_t)__float_as_uint(m0.x) | ((uintptr_t)__float_as_uint(m0.y) << 32));
  const float* nmap = reinterpret_cast<const float*>(
    (uintptr_t)__float_as_uint(m0.z) | ((uintptr_t)__float_as_uint(m0.w) << 32));
  const int dw = __float_as_int(m1.x), dh = __float_as_int(m1.y);
  const int nw = __float_as_int(m1.z), nh = __float_as_int(m1.w);
  s.ior = m2.x;
  // RGBA texel in one 16-byte load (4 channels checked at upload, arena 256-byte aligned)
  const float4 tx4 = __ldg(reinterpret_cast<const float4*>(diffuse + texel_index(dw, dh, 4, uv)));
  s.diffuse_col = mk3(tx4.x, tx4.y, tx4.z);
  s.specular_col = tx4.w;
  s.normal = surface_normal;
  if (nmap) { // intersection.cuh:228-242
    const int nidx = texel_index(nw, nh, 3, uv);
    float3 n = mk3(__ldg(nmap + nidx), __ldg(nmap + nidx + 1), __ldg(nmap + nidx + 2));
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


// Registers of one path between two stages.
struct Path
{
  float3 origin, dir, throughput, acc;
  float spec_carry; // IntersectionData.specular_col carried across bounces (Q3)
  uint32_t draws;   // RNG draws consumed so far (the state is replayed from the seed)
};

// One radiance-loop iteration (raytrace.cu:67-207), iteration `bounce`, given the closest
// hit of the path's current ray.  Returns true when the path is finished.  A path that
// misses is finished here arithmetically: the reference keeps looping on the unchanged ray,
// re-adding the environment with renormalised throughput (Q5).  `segs` receives the number
// of radiance-loop iterations this call accounts for.
__device__ __forceinline__ bool
shade_step(const FrameParams& fp, const DScene& sc, cudaTextureObject_t cube, Path& P, Rng& rng,
           const float dist, const uint32_t hit_id, const float hu, const float hv, const int bounce,
           uint32_t& segs)
{
  float3 origin = P.origin, dir = P.dir;
  float3 throughput = P.throughput, acc = P.acc;
  float spec_carry = P.spec_carry;
  const bool hit = hit_id != 0xFFFFFFFFu;
  bool done = false;
  int b = bounce;

  float r1 = rng_uniform(rng); // raytrace.cu:70, drawn before the intersection
  P.draws++;
  segs = 1;
  if (hit) {
    Surface s;
    resolve_hit(sc, hit_id, dist, hu, hv, dir, spec_carry, s);
    spec_carry = s.specular_col;
    const float cos_theta = dot(s.normal, dir);
    float3 oriented_normal = s.normal;
    const float3 spec = normalize(reflect(dir, s.normal));
    const float3 direct_light = s.diffuse_col / 0.5f; // brdf_lambert / pdf_lambert
    if (s.ior == 1.0f || s.is_light) {
      if (s.is_light)
        acc = acc + s.light_color * s.light_emission * throughput;
      const float phi = (float)((double)2.0f * 3.14159265358979323846 * (double)rng_uniform(rng));
      P.draws++;
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
        P.draws++;
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

  P.origin = origin, P.dir = dir, P.throughput = throughput, P.acc = acc;
  P.spec_carry = spec_carry;
  return done;
}

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

// Per-pixel tail of the reference kernel (raytrace.cu:248-270) for band-local pixel
// (x, lrow): called exactly once per pixel per frame, by whichever stage finishes the path.
__device__ __forceinline__ void
epilogue(const FrameParams& fp, const FrameOut& fo, uint32_t x, uint32_t lrow, float3 rad)
{
  const size_t p = (size_t)lrow * fp.width + x;
  if (fo.radiance_out) {
    fo.radiance_out[3 * p] = rad.x;
    fo.radiance_out[3 * p + 1] = rad.y;
    fo.radiance_out[3 * p + 2] = rad.z;
  }
  rad = mk3(clamp_sel(rad.x, 0.0f, 1.0f), clamp_sel(rad.y, 0.0f, 1.0f), clamp_sel(rad.z, 0.0f, 1.0f));

  // bottom-up rows inside the band == (height - y - 1) * width + x of raytrace.cu:252
  float* a = fo.accum + 3 * ((size_t)(fp.band_rows - lrow - 1) * fp.width + x);
  const float fs = (float)fp.is_static;
  float3 t = mk3(a[0], a[1], a[2]);
  t = t * fs;
  t = t + rad;
  a[0] = t.x, a[1] = t.y, a[2] = t.z;
  if (!fo.out_rgba)
    return;
  rad = t / (float)fp.frame_nb;
  rad = exposure(rad);
  rad = mk3(powf(rad.x, 1.0f / 2.2f), powf(rad.y, 1.0f / 2.2f), powf(rad.z, 1.0f / 2.2f));
  rad = post_process(rad, fp.post_id);
  const uint32_t R = (uint32_t)(rad.x * 255) & 0xFFu, G = (uint32_t)(rad.y * 255) & 0xFFu,
                 B = (uint32_t)(rad.z * 255) & 0xFFu;
  const size_t po = fo.out_full ? (size_t)global_row(fp, lrow) * fp.width + x : p;
  __stcs(fo.out_rgba + po, R | (G << 8) | (B << 16)); // a = 0, raytrace.cu:232
}

} // namespace ptc
