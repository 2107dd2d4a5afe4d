// ptcore.cu -- implementation of the C ABI in include/ptcore.h.
//
// Replaces, behind the reference's render boundary (raytrace.h:9-15), the upload
// path (scene.cpp:370-391, gpu_processor.cpp:68-238) and the per-frame launch
// (raytrace.cu:287-325).  One context per GPU; every kernel of a frame is enqueued
// on the caller's stream with no hidden synchronisation.  There is no CPU fallback:
// without a usable CUDA device ptc_create fails with PTC_ERR_NO_DEVICE.
#include <cuda_runtime.h>

#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "bvh.h"
#include "kernels.cuh"
#include "ptcore.h"

using namespace ptc;

namespace {

struct SceneDev
{
  DScene d{};
  std::vector<void*> allocs;
  ptc_scene_info info{};
};

struct CubeDev
{
  cudaArray_t array = nullptr;
  cudaTextureObject_t tex = 0;
};

} // namespace

struct ptc_context
{
  int device = 0;
  std::vector<SceneDev> scenes;
  std::vector<CubeDev> cubemaps;
  PathState st{};
  uint32_t capacity = 0;       // path slots allocated
  uint32_t* queue[2] = { nullptr, nullptr };
  uint32_t* counts = nullptr;  // one counter per bounce
  Counters* ctr = nullptr;
  uint32_t seed = 0;           // the reference's function-static frame counter (raytrace.cu:296)
  // staging for ptc_render_host
  float* h_accum_dev = nullptr;
  uint32_t* h_out_dev = nullptr;
  float* h_rad_dev = nullptr;
  size_t h_px = 0;
  std::string err;
};

namespace {

constexpr int kMaxBounces = 64;

int
fail(ptc_context* ctx, int code, const std::string& msg)
{
  if (ctx)
    ctx->err = msg;
  return code;
}

#define CUDA_TRY(ctx, expr)                                                                    \
  do {                                                                                         \
    cudaError_t e_ = (expr);                                                                   \
    if (e_ != cudaSuccess)                                                                     \
      return fail(ctx, (int)e_, std::string(#expr) + ": " + cudaGetErrorString(e_));           \
  } while (0)

struct DeviceGuard
{
  int prev = 0;
  explicit DeviceGuard(int dev)
  {
    cudaGetDevice(&prev);
    if (prev != dev)
      cudaSetDevice(dev);
    else
      prev = -1;
  }
  ~DeviceGuard()
  {
    if (prev >= 0)
      cudaSetDevice(prev);
  }
};

template <typename T>
int
dev_upload(ptc_context* ctx, SceneDev& s, const T* host, size_t n, const T** out)
{
  *out = nullptr;
  void* p = nullptr;
  CUDA_TRY(ctx, cudaMalloc(&p, n ? n * sizeof(T) : 16));
  s.allocs.push_back(p);
  if (n)
    CUDA_TRY(ctx, cudaMemcpy(p, host, n * sizeof(T), cudaMemcpyHostToDevice));
  *out = (const T*)p;
  return 0;
}

void
free_state(ptc_context* ctx)
{
  cudaFree(ctx->st.ray_o);
  cudaFree(ctx->st.ray_d);
  cudaFree(ctx->st.hit);
  cudaFree(ctx->st.thr);
  cudaFree(ctx->st.acc);
  cudaFree(ctx->st.rng_a);
  cudaFree(ctx->st.rng_b);
  cudaFree(ctx->queue[0]);
  cudaFree(ctx->queue[1]);
  ctx->st = PathState{};
  ctx->queue[0] = ctx->queue[1] = nullptr;
  ctx->capacity = 0;
}

int
ensure_capacity(ptc_context* ctx, uint32_t n)
{
  if (n <= ctx->capacity)
    return 0;
  free_state(ctx);
  CUDA_TRY(ctx, cudaMalloc(&ctx->st.ray_o, (size_t)n * sizeof(float4)));
  CUDA_TRY(ctx, cudaMalloc(&ctx->st.ray_d, (size_t)n * sizeof(float4)));
  CUDA_TRY(ctx, cudaMalloc(&ctx->st.hit, (size_t)n * sizeof(uint4)));
  CUDA_TRY(ctx, cudaMalloc(&ctx->st.thr, (size_t)n * sizeof(float4)));
  CUDA_TRY(ctx, cudaMalloc(&ctx->st.acc, (size_t)n * sizeof(float4)));
  CUDA_TRY(ctx, cudaMalloc(&ctx->st.rng_a, (size_t)n * sizeof(uint4)));
  CUDA_TRY(ctx, cudaMalloc(&ctx->st.rng_b, (size_t)n * sizeof(uint2)));
  CUDA_TRY(ctx, cudaMalloc(&ctx->queue[0], (size_t)n * sizeof(uint32_t)));
  CUDA_TRY(ctx, cudaMalloc(&ctx->queue[1], (size_t)n * sizeof(uint32_t)));
  ctx->capacity = n;
  return 0;
}

// rt.cu:275-285 (host side of the seed, raytrace.cu:321)
uint32_t
wang_hash(uint32_t a)
{
  a = (a ^ 61u) ^ (a >> 16);
  a = a + (a << 3);
  a = a ^ (a >> 4);
  a = a * 0x27d4eb2du;
  a = a ^ (a >> 15);
  return a;
}

int
fill_frame(ptc_context* ctx, const ptc_render_args* a, FrameParams& fp, bool advance_counter)
{
  if (!a || !a->camera)
    return fail(ctx, PTC_ERR_INVALID, "null render args / camera");
  if (a->scene_id < 0 || (size_t)a->scene_id >= ctx->scenes.size())
    return fail(ctx, PTC_ERR_INVALID, "unknown scene_id");
  if (a->cubemap_id < 0 || (size_t)a->cubemap_id >= ctx->cubemaps.size())
    return fail(ctx, PTC_ERR_INVALID, "unknown cubemap_id (call ptc_set_cubemap first)");
  if (a->width == 0 || a->height == 0)
    return fail(ctx, PTC_ERR_INVALID, "empty image");
  fp.cam = *a->camera;
  fp.width = a->width;
  fp.height = a->height;
  fp.row_begin = a->row_begin;
  fp.row_end = a->row_end;
  if (fp.row_begin == 0 && fp.row_end == 0)
    fp.row_end = a->height;
  if (fp.row_begin >= fp.row_end || fp.row_end > a->height)
    return fail(ctx, PTC_ERR_INVALID, "bad row range");
  uint32_t bounces = a->bounces ? a->bounces : 3u; // 1 + (static_samples + 1), static_samples = 1
  if (bounces > (uint32_t)kMaxBounces)
    return fail(ctx, PTC_ERR_INVALID, "too many bounces");
  fp.max_bounces = (int)bounces;
  fp.is_static = a->moved ? 0 : 1;
  fp.post_id = a->post_id;
  if (fp.post_id > 3)
    return fail(ctx, PTC_ERR_INVALID, "post_id out of range");
  uint32_t frame = a->frame_nb;
  if (frame == 0) { // raytrace.cu:296-300
    uint32_t seed = ctx->seed;
    if (a->moved)
      seed = 0;
    seed++;
    if (advance_counter)
      ctx->seed = seed;
    frame = seed;
  }
  fp.frame_nb = (int)frame;
  fp.hash_seed = wang_hash(frame);
  return 0;
}

// Enqueues one frame; accum/out/rad are device pointers.
int
enqueue_frame(ptc_context* ctx, const ptc_render_args* a, const FrameParams& fp, float* accum,
              uint32_t* out, float* rad, cudaStream_t s)
{
  const SceneDev& sc = ctx->scenes[a->scene_id];
  const CubeDev& cm = ctx->cubemaps[a->cubemap_id];
  const uint32_t n = (fp.row_end - fp.row_begin) * fp.width;
  int rc = ensure_capacity(ctx, n);
  if (rc)
    return rc;
  CUDA_TRY(ctx, cudaMemsetAsync(ctx->counts, 0, sizeof(uint32_t) * (kMaxBounces + 1), s));
  launch_generate(fp, ctx->st, s);
  if (!fp.is_static) {
    launch_extend(sc.d, ctx->st, nullptr, nullptr, n, ctx->ctr, 0, s);
    launch_shade_preview(fp, sc.d, ctx->st, cm.tex, n, ctx->ctr, s);
  } else {
    for (int b = 0; b < fp.max_bounces; ++b) {
      const uint32_t* q = b ? ctx->queue[b & 1] : nullptr;
      const uint32_t* qc = b ? ctx->counts + b : nullptr;
      launch_extend(sc.d, ctx->st, q, qc, n, ctx->ctr, b, s);
      launch_shade(fp, sc.d, ctx->st, cm.tex, q, qc, n, ctx->queue[(b + 1) & 1], ctx->counts + b + 1,
                   ctx->ctr, b, s);
    }
  }
  launch_accumulate(fp, ctx->st, accum, out, rad, ctx->ctr, s);
  CUDA_TRY(ctx, cudaGetLastError());
  return 0;
}

} // namespace

extern "C" {

int
ptc_abi_version(void)
{
  return PTC_ABI_VERSION;
}

int
ptc_create(int device, ptc_context** out)
{
  if (!out)
    return PTC_ERR_INVALID;
  *out = nullptr;
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || count <= 0 || device < 0 || device >= count) {
    cudaGetLastError();
    return PTC_ERR_NO_DEVICE;
  }
  ptc_context* ctx = new ptc_context();
  ctx->device = device;
  DeviceGuard g(device);
  if (cudaMalloc(&ctx->ctr, sizeof(Counters)) != cudaSuccess ||
      cudaMalloc(&ctx->counts, sizeof(uint32_t) * (kMaxBounces + 1)) != cudaSuccess) {
    delete ctx;
    return PTC_ERR_ALLOC;
  }
  cudaMemset(ctx->ctr, 0, sizeof(Counters));
  cudaMemset(ctx->counts, 0, sizeof(uint32_t) * (kMaxBounces + 1));
  *out = ctx;
  return PTC_OK;
}

void
ptc_destroy(ptc_context* ctx)
{
  if (!ctx)
    return;
  DeviceGuard g(ctx->device);
  cudaDeviceSynchronize();
  free_state(ctx);
  for (SceneDev& s : ctx->scenes)
    for (void* p : s.allocs)
      cudaFree(p);
  for (CubeDev& c : ctx->cubemaps) {
    if (c.tex)
      cudaDestroyTextureObject(c.tex);
    if (c.array)
      cudaFreeArray(c.array);
  }
  cudaFree(ctx->ctr);
  cudaFree(ctx->counts);
  cudaFree(ctx->h_accum_dev);
  cudaFree(ctx->h_out_dev);
  cudaFree(ctx->h_rad_dev);
  delete ctx;
}

const char*
ptc_last_error(const ptc_context* ctx)
{
  return ctx ? ctx->err.c_str() : "null context";
}

int
ptc_upload_scene(ptc_context* ctx, const ptc_scene_desc* desc, int32_t* scene_id)
{
  if (!ctx || !desc || !scene_id)
    return PTC_ERR_INVALID;
  DeviceGuard g(ctx->device);
  uint32_t n_faces = 0;
  for (uint32_t m = 0; m < desc->n_meshes; ++m)
    n_faces += desc->mesh_face_counts[m];
  for (uint32_t f = 0; f < n_faces; ++f)
    if (desc->faces[f].material_id >= desc->n_materials)
      return fail(ctx, PTC_ERR_INVALID, "face references a material outside the table");
  for (uint32_t i = 0; i < desc->n_materials; ++i) {
    const ptc_material& m = desc->materials[i];
    if (m.diffuse_spec_map < 0 || (uint32_t)m.diffuse_spec_map >= desc->n_textures ||
        m.normal_map >= (int32_t)desc->n_textures)
      return fail(ctx, PTC_ERR_INVALID, "material references a texture outside the table");
    if (desc->textures[m.diffuse_spec_map].nb_chan != 4)
      return fail(ctx, PTC_ERR_INVALID, "diffuse/specular texture must have 4 channels");
    if (m.normal_map >= 0 && desc->textures[m.normal_map].nb_chan != 3)
      return fail(ctx, PTC_ERR_INVALID, "normal map must have 3 channels");
  }

  HostBvh bvh;
  build_bvh(desc, bvh);
  std::vector<ShadeRecord> shade;
  build_shade_records(desc, shade);

  ctx->scenes.emplace_back();
  SceneDev& s = ctx->scenes.back();
  int rc;
  const Node8* d_nodes;
  const TriRecord* d_tris;
  const ShadeRecord* d_shade;
  if ((rc = dev_upload(ctx, s, bvh.nodes.data(), bvh.nodes.size(), &d_nodes)) ||
      (rc = dev_upload(ctx, s, bvh.tris.data(), bvh.tris.size(), &d_tris)) ||
      (rc = dev_upload(ctx, s, shade.data(), shade.size(), &d_shade)) ||
      (rc = dev_upload(ctx, s, desc->materials, desc->n_materials, &s.d.materials)) ||
      (rc = dev_upload(ctx, s, desc->lights, desc->n_lights, &s.d.lights)))
    return rc;
  s.d.nodes = (const uint4*)d_nodes;
  s.d.tris = (const float4*)d_tris;
  s.d.shade = (const float4*)d_shade;
  s.d.n_faces = n_faces;
  s.d.n_lights = desc->n_lights;
  s.d.n_nodes = (uint32_t)bvh.nodes.size();
  s.d.n_tris = (uint32_t)bvh.tris.size();

  // one texel arena, 16-byte aligned sub-allocations (gpu_processor.cpp:205-238 made one
  // cudaMalloc per texture)
  std::vector<size_t> offs(desc->n_textures);
  size_t total = 0;
  for (uint32_t i = 0; i < desc->n_textures; ++i) {
    offs[i] = total;
    size_t bytes = (size_t)desc->textures[i].w * desc->textures[i].h * desc->textures[i].nb_chan * 4;
    total += (bytes + 255) & ~(size_t)255;
  }
  char* arena = nullptr;
  CUDA_TRY(ctx, cudaMalloc(&arena, total ? total : 256));
  s.allocs.push_back(arena);
  std::vector<DTexture> dt(desc->n_textures);
  for (uint32_t i = 0; i < desc->n_textures; ++i) {
    const ptc_texture& t = desc->textures[i];
    size_t bytes = (size_t)t.w * t.h * t.nb_chan * 4;
    if (bytes)
      CUDA_TRY(ctx, cudaMemcpy(arena + offs[i], t.data, bytes, cudaMemcpyHostToDevice));
    dt[i].data = (const float*)(arena + offs[i]);
    dt[i].w = t.w;
    dt[i].h = t.h;
    dt[i].nb_chan = t.nb_chan;
    dt[i]._pad = 0;
  }
  if ((rc = dev_upload(ctx, s, dt.data(), dt.size(), &s.d.textures)))
    return rc;

  s.info.n_faces = n_faces;
  s.info.n_lights = desc->n_lights;
  s.info.n_materials = desc->n_materials;
  s.info.n_textures = desc->n_textures;
  s.info.bvh_nodes = (uint32_t)bvh.nodes.size();
  s.info.bvh_node_bytes = sizeof(Node8);
  s.info.tri_bytes = sizeof(TriRecord);
  s.info.shade_bytes = sizeof(ShadeRecord);
  s.info.texture_bytes = total;
  s.info.smem_resident = 0;
  *scene_id = (int32_t)ctx->scenes.size() - 1;
  return PTC_OK;
}

int
ptc_scene_info_get(const ptc_context* ctx, int32_t scene_id, ptc_scene_info* out)
{
  if (!ctx || !out || scene_id < 0 || (size_t)scene_id >= ctx->scenes.size())
    return PTC_ERR_INVALID;
  *out = ctx->scenes[scene_id].info;
  return PTC_OK;
}

int
ptc_set_cubemap(ptc_context* ctx, const float* faces, uint32_t size, int32_t* cubemap_id)
{
  if (!ctx || !faces || !size || !cubemap_id)
    return PTC_ERR_INVALID;
  DeviceGuard g(ctx->device);
  CubeDev c;
  // same array type and sampler state as the reference (gpu_processor.cpp:134-152,
  // raytrace.cu:304-309): RGBA32F cubemap, wrap, linear, normalised coordinates
  cudaChannelFormatDesc desc = cudaCreateChannelDesc(32, 32, 32, 32, cudaChannelFormatKindFloat);
  CUDA_TRY(ctx, cudaMalloc3DArray(&c.array, &desc, make_cudaExtent(size, size, 6), cudaArrayCubemap));
  cudaMemcpy3DParms p;
  memset(&p, 0, sizeof(p));
  p.srcPtr = make_cudaPitchedPtr((void*)faces, size * sizeof(float) * 4, size, size);
  p.dstArray = c.array;
  p.extent = make_cudaExtent(size, size, 6);
  p.kind = cudaMemcpyHostToDevice;
  CUDA_TRY(ctx, cudaMemcpy3D(&p));
  cudaResourceDesc res;
  memset(&res, 0, sizeof(res));
  res.resType = cudaResourceTypeArray;
  res.res.array.array = c.array;
  cudaTextureDesc td;
  memset(&td, 0, sizeof(td));
  td.addressMode[0] = cudaAddressModeWrap;
  td.addressMode[1] = cudaAddressModeWrap;
  td.filterMode = cudaFilterModeLinear;
  td.normalizedCoords = 1;
  td.readMode = cudaReadModeElementType;
  CUDA_TRY(ctx, cudaCreateTextureObject(&c.tex, &res, &td, nullptr));
  ctx->cubemaps.push_back(c);
  *cubemap_id = (int32_t)ctx->cubemaps.size() - 1;
  return PTC_OK;
}

int
ptc_render(ptc_context* ctx, const ptc_render_args* args)
{
  if (!ctx)
    return PTC_ERR_INVALID;
  DeviceGuard g(ctx->device);
  FrameParams fp;
  int rc = fill_frame(ctx, args, fp, true);
  if (rc)
    return rc;
  if (!args->accum)
    return fail(ctx, PTC_ERR_INVALID, "accum is required");
  return enqueue_frame(ctx, args, fp, (float*)args->accum, (uint32_t*)args->out_rgba,
                       (float*)args->radiance_out, (cudaStream_t)args->stream);
}

int
ptc_render_host(ptc_context* ctx, const ptc_render_args* args, int upload_accum, int download_accum)
{
  if (!ctx)
    return PTC_ERR_INVALID;
  DeviceGuard g(ctx->device);
  FrameParams fp;
  int rc = fill_frame(ctx, args, fp, true);
  if (rc)
    return rc;
  if (!args->accum)
    return fail(ctx, PTC_ERR_INVALID, "accum is required");
  const size_t n = (size_t)(fp.row_end - fp.row_begin) * fp.width;
  cudaStream_t s = (cudaStream_t)args->stream;
  if (n != ctx->h_px) {
    cudaFree(ctx->h_accum_dev);
    cudaFree(ctx->h_out_dev);
    cudaFree(ctx->h_rad_dev);
    ctx->h_accum_dev = nullptr, ctx->h_out_dev = nullptr, ctx->h_rad_dev = nullptr;
    ctx->h_px = 0;
    CUDA_TRY(ctx, cudaMalloc(&ctx->h_accum_dev, n * 12));
    CUDA_TRY(ctx, cudaMalloc(&ctx->h_out_dev, n * 4));
    CUDA_TRY(ctx, cudaMalloc(&ctx->h_rad_dev, n * 12));
    ctx->h_px = n;
    upload_accum = 1; // a fresh device accumulator must be seeded from the caller's
  }
  if (upload_accum)
    CUDA_TRY(ctx, cudaMemcpyAsync(ctx->h_accum_dev, args->accum, n * 12, cudaMemcpyHostToDevice, s));
  rc = enqueue_frame(ctx, args, fp, ctx->h_accum_dev, args->out_rgba ? ctx->h_out_dev : nullptr,
                     args->radiance_out ? ctx->h_rad_dev : nullptr, s);
  if (rc)
    return rc;
  if (args->out_rgba)
    CUDA_TRY(ctx, cudaMemcpyAsync(args->out_rgba, ctx->h_out_dev, n * 4, cudaMemcpyDeviceToHost, s));
  if (args->radiance_out)
    CUDA_TRY(ctx, cudaMemcpyAsync(args->radiance_out, ctx->h_rad_dev, n * 12, cudaMemcpyDeviceToHost, s));
  if (download_accum)
    CUDA_TRY(ctx, cudaMemcpyAsync(args->accum, ctx->h_accum_dev, n * 12, cudaMemcpyDeviceToHost, s));
  CUDA_TRY(ctx, cudaStreamSynchronize(s));
  return PTC_OK;
}

int
ptc_reset_accumulation(ptc_context* ctx)
{
  if (!ctx)
    return PTC_ERR_INVALID;
  ctx->seed = 0;
  return PTC_OK;
}

int
ptc_counters_get(ptc_context* ctx, ptc_counters* out)
{
  if (!ctx || !out)
    return PTC_ERR_INVALID;
  DeviceGuard g(ctx->device);
  Counters c;
  CUDA_TRY(ctx, cudaDeviceSynchronize());
  CUDA_TRY(ctx, cudaMemcpy(&c, ctx->ctr, sizeof(c), cudaMemcpyDeviceToHost));
  memset(out, 0, sizeof(*out));
  out->paths = c.paths;
  out->segments = c.segments;
  out->traced = c.traced;
  for (int i = 0; i < 16; ++i)
    out->rays_per_bounce[i] = c.rays_per_bounce[i];
  out->nodes_visited = c.nodes_visited;
  out->tris_tested = c.tris_tested;
  out->paths = c.rays_per_bounce[0];
  return PTC_OK;
}

int
ptc_counters_reset(ptc_context* ctx)
{
  if (!ctx)
    return PTC_ERR_INVALID;
  DeviceGuard g(ctx->device);
  CUDA_TRY(ctx, cudaDeviceSynchronize());
  CUDA_TRY(ctx, cudaMemset(ctx->ctr, 0, sizeof(Counters)));
  return PTC_OK;
}

int
ptc_trace_primary(ptc_context* ctx, int32_t scene_id, const ptc_camera* camera, uint32_t width,
                  uint32_t height, uint32_t frame_nb, int32_t* hit_ids, float* t)
{
  if (!ctx || !camera || scene_id < 0 || (size_t)scene_id >= ctx->scenes.size() || !width || !height)
    return PTC_ERR_INVALID;
  DeviceGuard g(ctx->device);
  FrameParams fp;
  memset(&fp, 0, sizeof(fp));
  fp.cam = *camera;
  fp.width = width;
  fp.height = height;
  fp.row_begin = 0;
  fp.row_end = height;
  fp.frame_nb = (int)(frame_nb ? frame_nb : 1);
  fp.hash_seed = wang_hash((uint32_t)fp.frame_nb);
  fp.max_bounces = 1;
  fp.is_static = 1;
  const uint32_t n = width * height;
  int rc = ensure_capacity(ctx, n);
  if (rc)
    return rc;
  int32_t* d_ids = nullptr;
  float* d_t = nullptr;
  CUDA_TRY(ctx, cudaMalloc(&d_ids, (size_t)n * 4));
  CUDA_TRY(ctx, cudaMalloc(&d_t, (size_t)n * 4));
  launch_generate(fp, ctx->st, 0);
  launch_extend(ctx->scenes[scene_id].d, ctx->st, nullptr, nullptr, n, ctx->ctr, 0, 0);
  launch_export_primary(ctx->st, n, d_ids, d_t, nullptr, nullptr, 0);
  cudaError_t e = cudaDeviceSynchronize();
  if (e == cudaSuccess && hit_ids)
    e = cudaMemcpy(hit_ids, d_ids, (size_t)n * 4, cudaMemcpyDeviceToHost);
  if (e == cudaSuccess && t)
    e = cudaMemcpy(t, d_t, (size_t)n * 4, cudaMemcpyDeviceToHost);
  cudaFree(d_ids);
  cudaFree(d_t);
  if (e != cudaSuccess)
    return fail(ctx, (int)e, cudaGetErrorString(e));
  return PTC_OK;
}

int
ptc_intersect_rays(ptc_context* ctx, int32_t scene_id, const float* origins, const float* dirs,
                   uint64_t n, int32_t* hit_ids, float* t)
{
  if (!ctx || scene_id < 0 || (size_t)scene_id >= ctx->scenes.size() || (n && (!origins || !dirs)))
    return PTC_ERR_INVALID;
  if (!n)
    return PTC_OK;
  DeviceGuard g(ctx->device);
  float *d_o = nullptr, *d_d = nullptr, *d_t = nullptr;
  int32_t* d_i = nullptr;
  CUDA_TRY(ctx, cudaMalloc(&d_o, n * 12));
  CUDA_TRY(ctx, cudaMalloc(&d_d, n * 12));
  CUDA_TRY(ctx, cudaMalloc(&d_i, n * 4));
  CUDA_TRY(ctx, cudaMalloc(&d_t, n * 4));
  cudaError_t e = cudaMemcpy(d_o, origins, n * 12, cudaMemcpyHostToDevice);
  if (e == cudaSuccess)
    e = cudaMemcpy(d_d, dirs, n * 12, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) {
    launch_intersect_rays(ctx->scenes[scene_id].d, d_o, d_d, n, d_i, d_t, ctx->ctr, 0);
    e = cudaDeviceSynchronize();
  }
  if (e == cudaSuccess && hit_ids)
    e = cudaMemcpy(hit_ids, d_i, n * 4, cudaMemcpyDeviceToHost);
  if (e == cudaSuccess && t)
    e = cudaMemcpy(t, d_t, n * 4, cudaMemcpyDeviceToHost);
  cudaFree(d_o), cudaFree(d_d), cudaFree(d_i), cudaFree(d_t);
  if (e != cudaSuccess)
    return fail(ctx, (int)e, cudaGetErrorString(e));
  return PTC_OK;
}

int
ptc_debug_rng(ptc_context* ctx, const uint32_t* seeds, uint32_t n_seeds, uint32_t n_draws,
              int reference_impl, float* out)
{
  if (!ctx || !seeds || !out || !n_seeds || !n_draws)
    return PTC_ERR_INVALID;
  DeviceGuard g(ctx->device);
  uint32_t* d_s = nullptr;
  float* d_o = nullptr;
  CUDA_TRY(ctx, cudaMalloc(&d_s, (size_t)n_seeds * 4));
  CUDA_TRY(ctx, cudaMalloc(&d_o, (size_t)n_seeds * n_draws * 4));
  cudaError_t e = cudaMemcpy(d_s, seeds, (size_t)n_seeds * 4, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) {
    if (reference_impl)
      launch_debug_curand(d_s, n_seeds, n_draws, d_o, 0);
    else
      launch_debug_rng(d_s, n_seeds, n_draws, d_o, 0);
    e = cudaDeviceSynchronize();
  }
  if (e == cudaSuccess)
    e = cudaMemcpy(out, d_o, (size_t)n_seeds * n_draws * 4, cudaMemcpyDeviceToHost);
  cudaFree(d_s), cudaFree(d_o);
  if (e != cudaSuccess)
    return fail(ctx, (int)e, cudaGetErrorString(e));
  return PTC_OK;
}

} // extern "C"
