// TEST INFRASTRUCTURE (oracle/_ref recipe) -- not product code.
//
// Executes the REFERENCE's own path-tracing code on host cores.  The file
// gen/raytrace_cpu.cpp included below is produced by oracle/build_ref.py from
// /root/reference/cuda_opengl/src/shaders/raytrace.cu with three mechanical
// substitutions (launch syntax -> ref_launch, static_samples -> runtime value,
// IntersectionData zero-initialised); everything else -- radiance(), intersect(),
// intersectTriangle(), camera_dof(), exposure(), the cuRAND XORWOW from the
// toolkit's curand_kernel.h -- is the reference's code, compiled by g++ with
// ref_shim/ref_cpu_shim.h force-included.
//
// The C entry points take the same flat scene description as the product's C
// ABI (include/ptcore.h) so the tests can feed all three implementations from
// one fixture.
#include "raytrace_cpu.cpp" // generated into oracle/_ref/gen/

#include <atomic>
#include <thread>
#include <vector>

#include "ptcore.h"

// ---- shim globals ---------------------------------------------------------
thread_local uint3 threadIdx, blockIdx;
dim3 blockDim, gridDim;
int g_ref_threads = 1;
RefCubemapHost g_ref_cubemap = { nullptr, 0 };
uint32_t* g_ref_out_rgba = nullptr;
unsigned g_ref_out_width = 0;
int g_ref_static_samples = 1;
// bounded sample of a launch (bench.py's reference arm): only blocks with
// b % stride == offset are executed
static unsigned g_ref_block_stride = 1, g_ref_block_offset = 0;

void
ref_parallel_blocks(unsigned n_blocks_all, void (*fn_all)(unsigned, void*), void* ctx_all)
{
  struct Strided
  {
    void (*fn)(unsigned, void*);
    void* ctx;
    unsigned stride, offset;
  } sd = { fn_all, ctx_all, g_ref_block_stride, g_ref_block_offset % g_ref_block_stride };
  const unsigned n_blocks =
    sd.stride > 1 ? (n_blocks_all > sd.offset ? (n_blocks_all - sd.offset + sd.stride - 1) / sd.stride : 0)
                  : n_blocks_all;
  void (*fn)(unsigned, void*) = [](unsigned b, void* c) {
    Strided* s = static_cast<Strided*>(c);
    s->fn(s->stride > 1 ? b * s->stride + s->offset : b, s->ctx);
  };
  void* ctx = &sd;
  int nt = g_ref_threads > 0 ? g_ref_threads : 1;
  if (nt == 1) {
    for (unsigned b = 0; b < n_blocks; ++b)
      fn(b, ctx);
    return;
  }
  std::atomic<unsigned> next(0);
  std::vector<std::thread> pool;
  for (int t = 0; t < nt; ++t)
    pool.emplace_back([&] {
      for (;;) {
        unsigned b = next.fetch_add(8);
        if (b >= n_blocks)
          break;
        for (unsigned i = b; i < b + 8 && i < n_blocks; ++i)
          fn(i, ctx);
      }
    });
  for (auto& th : pool)
    th.join();
}

// Software stand-in for the cubemap texture unit (raytrace.cu:60,197,304-309:
// normalised coordinates, linear filter).  Face selection follows the CUDA
// programming guide's cubemap table; bilinear weights are quantised to 8
// fractional bits like the hardware; texels clamp at face borders.
float4
ref_shim_texCubemap(float x, float y, float z)
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
  const unsigned N = g_ref_cubemap.size;
  const float* base = g_ref_cubemap.faces + (size_t)face * N * N * 4;
  float u = (s / m + 1.0f) * 0.5f, v = (t / m + 1.0f) * 0.5f;
  float xb = u * N - 0.5f, yb = v * N - 0.5f;
  float fx = floorf(xb), fy = floorf(yb);
  float a = floorf((xb - fx) * 256.0f + 0.5f) / 256.0f;
  float b = floorf((yb - fy) * 256.0f + 0.5f) / 256.0f;
  auto clampi = [N](int i) { return i < 0 ? 0 : (i >= (int)N ? (int)N - 1 : i); };
  int x0 = clampi((int)fx), x1 = clampi((int)fx + 1);
  int y0 = clampi((int)fy), y1 = clampi((int)fy + 1);
  float4 r;
  float* out = &r.x;
  for (int c = 0; c < 4; ++c) {
    float t00 = base[((size_t)y0 * N + x0) * 4 + c], t10 = base[((size_t)y0 * N + x1) * 4 + c];
    float t01 = base[((size_t)y1 * N + x0) * 4 + c], t11 = base[((size_t)y1 * N + x1) * 4 + c];
    out[c] = (1 - a) * (1 - b) * t00 + a * (1 - b) * t10 + (1 - a) * b * t01 + a * b * t11;
  }
  return r;
}

namespace {

// Rebuilds the reference's pointer tree Scenes -> SceneData*[] -> Mesh[] -> Face[]
// (scene_data.h:59-87) over the flat description; everything stays host memory.
struct HostScenes
{
  std::vector<scene::Mesh> meshes;
  std::vector<scene::Texture> textures;
  scene::SceneData sd;
  scene::SceneData* sd_ptr;
  scene::Scenes scenes;

  explicit HostScenes(const ptc_scene_desc* d)
  {
    static_assert(sizeof(scene::Face) == sizeof(ptc_face), "Face");
    static_assert(sizeof(scene::Material) == sizeof(ptc_material), "Material");
    static_assert(sizeof(scene::LightProp) == sizeof(ptc_light), "LightProp");
    static_assert(sizeof(scene::Texture) == sizeof(ptc_texture), "Texture");
    static_assert(sizeof(scene::Camera) == sizeof(ptc_camera), "Camera");
    meshes.resize(d->n_meshes);
    size_t off = 0;
    for (uint32_t m = 0; m < d->n_meshes; ++m) {
      meshes[m].faces.size = d->mesh_face_counts[m];
      meshes[m].faces.data = (scene::Face*)d->faces + off;
      off += d->mesh_face_counts[m];
    }
    textures.resize(d->n_textures);
    for (uint32_t i = 0; i < d->n_textures; ++i) {
      textures[i].w = d->textures[i].w;
      textures[i].h = d->textures[i].h;
      textures[i].nb_chan = d->textures[i].nb_chan;
      textures[i].data = (float*)d->textures[i].data;
    }
    sd.meshes.size = d->n_meshes;
    sd.meshes.data = meshes.data();
    sd.materials.size = d->n_materials;
    sd.materials.data = (scene::Material*)d->materials;
    sd.lights.size = d->n_lights;
    sd.lights.data = (scene::LightProp*)d->lights;
    sd_ptr = &sd;
    scenes.textures.size = d->n_textures;
    scenes.textures.data = textures.data();
    scenes.scenes = &sd_ptr;
  }
};

scene::Camera
to_cam(const ptc_camera* c)
{
  scene::Camera cam;
  std::memcpy(&cam, c, sizeof(cam));
  return cam;
}

} // namespace

extern "C" {

// One launch of the reference `kernel` (raytrace.cu:212-271) for sample
// `frame_nb`, exactly as raytrace() would issue it (raytrace.cu:315-322).
int
ref_cpu_kernel(const ptc_scene_desc* desc, const float* cube_faces, unsigned cube_size,
               const ptc_camera* camera, unsigned width, unsigned height,
               unsigned frame_nb, unsigned bounces, int moved, unsigned post_id,
               float* accum, uint32_t* out_rgba, int n_threads)
{
  HostScenes hs(desc);
  g_ref_cubemap.faces = cube_faces;
  g_ref_cubemap.size = cube_size;
  g_ref_out_rgba = out_rgba;
  g_ref_out_width = width;
  g_ref_static_samples = (int)bounces - 2; // max_bounces = 1 + (static_samples + 1)
  g_ref_threads = n_threads;
  static bool tables = false;
  if (!tables) {
    setupFunctionTables();
    tables = true;
  }
  dim3 threads_per_block(16, 16);
  dim3 nb_blocks(width / threads_per_block.x + 1, height / threads_per_block.y + 1);
  ref_launch(kernel, nb_blocks, threads_per_block, width, height, hs.scenes, 0u,
             to_cam(camera), WangHash(frame_nb), (int)frame_nb, (float3*)accum,
             moved != 0, h_post_process_table[post_id]);
  return 0;
}

// The same launch restricted to the 16x16 blocks b with b % block_stride == block_offset
// (a bounded sample for timing the reference on host cores).
int
ref_cpu_kernel_sample(const ptc_scene_desc* desc, const float* cube_faces, unsigned cube_size,
                      const ptc_camera* camera, unsigned width, unsigned height,
                      unsigned frame_nb, unsigned bounces, unsigned block_stride,
                      unsigned block_offset, float* accum, uint32_t* out_rgba, int n_threads)
{
  g_ref_block_stride = block_stride ? block_stride : 1;
  g_ref_block_offset = block_offset;
  int rc = ref_cpu_kernel(desc, cube_faces, cube_size, camera, width, height, frame_nb, bounces, 0,
                          0, accum, out_rgba, n_threads);
  g_ref_block_stride = 1;
  g_ref_block_offset = 0;
  return rc;
}

// Per-pixel un-clamped radiance of sample `frame_nb` (row 0 = top): the kernel's
// prelude (raytrace.cu:227-246) followed by the reference's radiance().
int
ref_cpu_radiance(const ptc_scene_desc* desc, const float* cube_faces, unsigned cube_size,
                 const ptc_camera* camera, unsigned width, unsigned height,
                 unsigned frame_nb, unsigned bounces, int moved, float* rad_out,
                 int n_threads)
{
  HostScenes hs(desc);
  g_ref_cubemap.faces = cube_faces;
  g_ref_cubemap.size = cube_size;
  g_ref_threads = n_threads;
  const unsigned gx = width / 16 + 1;
  const unsigned hash_seed = WangHash(frame_nb);
  const int static_samples = (int)bounces - 2;
  auto body = [&](unsigned y) {
    for (unsigned x = 0; x < width; ++x) {
      unsigned tid = ((x / 16) + (y / 16) * gx) * 256 + (y % 16) * 16 + (x % 16);
      curandState st;
      curand_init(hash_seed + tid, 0, 0, &st);
      scene::Camera cam = to_cam(camera);
      scene::Ray r = generateRay(x, y, width / 2, height / 2, cam);
      camera_dof(r, cam, &st);
      float3 rad = radiance(r, hs.scenes, 0, &cam, &st, !moved, static_samples);
      float* o = rad_out + ((size_t)y * width + x) * 3;
      o[0] = rad.x;
      o[1] = rad.y;
      o[2] = rad.z;
    }
  };
  using B = decltype(body);
  ref_parallel_blocks(height, [](unsigned b, void* c) { (*static_cast<B*>(c))(b); }, &body);
  return 0;
}

// Closest hit of arbitrary rays with the reference's intersectTriangle /
// intersectSphere and the acceptance rules of intersect()
// (intersection.cuh:179-212).  intersect() itself does not report which face
// won, so the 12-line loop is repeated here around the reference's own tests.
int
ref_cpu_intersect(const ptc_scene_desc* desc, const float* origins, const float* dirs,
                  uint64_t n, int32_t* hit_ids, float* t_out, int n_threads)
{
  HostScenes hs(desc);
  g_ref_threads = n_threads;
  uint32_t n_faces = 0;
  for (uint32_t m = 0; m < desc->n_meshes; ++m)
    n_faces += desc->mesh_face_counts[m];
  const unsigned chunk = 4096;
  auto body = [&](unsigned c) {
    uint64_t lo = (uint64_t)c * chunk, hi = lo + chunk < n ? lo + chunk : n;
    for (uint64_t k = lo; k < hi; ++k) {
      scene::Ray r;
      r.origin = make_float3(origins[3 * k], origins[3 * k + 1], origins[3 * k + 2]);
      r.dir = make_float3(dirs[3 * k], dirs[3 * k + 1], dirs[3 * k + 2]);
      float dist = 100000.0f, inter_dist = 100000.0f;
      int32_t id = -1, g = 0;
      float3 nrm;
      float2 uv;
      for (size_t m = 0; m < hs.sd.meshes.size; ++m) {
        const auto& mesh = hs.sd.meshes.data[m];
        for (size_t i = 0; i < mesh.faces.size; ++i, ++g)
          if (intersectTriangle(mesh.faces.data[i], nrm, uv, r, inter_dist) &&
              inter_dist < dist && inter_dist > 0.0) {
            dist = inter_dist;
            id = g;
          }
      }
      for (unsigned l = 0; l < hs.sd.lights.size; ++l)
        if (intersectSphere(r, hs.sd.lights.data[l], inter_dist) && inter_dist < dist &&
            inter_dist >= 0.0) {
          dist = inter_dist;
          id = (int32_t)(n_faces + l);
        }
      if (hit_ids)
        hit_ids[k] = id;
      if (t_out)
        t_out[k] = dist;
    }
  };
  using B = decltype(body);
  ref_parallel_blocks((unsigned)((n + chunk - 1) / chunk),
                      [](unsigned b, void* c) { (*static_cast<B*>(c))(b); }, &body);
  return 0;
}

// Primary rays of sample `frame_nb` as the kernel generates them
// (raytrace.cu:227-240): origins/dirs out (row 0 = top).
int
ref_cpu_primary_rays(const ptc_camera* camera, unsigned width, unsigned height,
                     unsigned frame_nb, float* origins, float* dirs)
{
  const unsigned gx = width / 16 + 1;
  const unsigned hash_seed = WangHash(frame_nb);
  for (unsigned y = 0; y < height; ++y)
    for (unsigned x = 0; x < width; ++x) {
      unsigned tid = ((x / 16) + (y / 16) * gx) * 256 + (y % 16) * 16 + (x % 16);
      curandState st;
      curand_init(hash_seed + tid, 0, 0, &st);
      scene::Camera cam = to_cam(camera);
      scene::Ray r = generateRay(x, y, width / 2, height / 2, cam);
      camera_dof(r, cam, &st);
      size_t k = (size_t)y * width + x;
      origins[3 * k] = r.origin.x, origins[3 * k + 1] = r.origin.y, origins[3 * k + 2] = r.origin.z;
      dirs[3 * k] = r.dir.x, dirs[3 * k + 1] = r.dir.y, dirs[3 * k + 2] = r.dir.z;
    }
  return 0;
}

// Known-answer generator: the toolkit's own XORWOW (curand_kernel.h:772-799,
// 863-874; curand_uniform.h:69-72) as the reference calls it (raytrace.cu:234-235).
void
ref_curand_uniforms(uint32_t seed, unsigned n, float* out)
{
  curandState st;
  curand_init(seed, 0, 0, &st);
  for (unsigned i = 0; i < n; ++i)
    out[i] = curand_uniform(&st);
}

unsigned
ref_wang_hash(unsigned a)
{
  return WangHash(a);
}

} // extern "C"
