// TEST / BASELINE INFRASTRUCTURE (oracle/_ref recipe) -- not product code.
//
// GPU baseline "G0": the REFERENCE's own kernel (cuda_opengl/src/shaders/raytrace.cu
// and the .cuh headers it inlines) built for sm_100a.  gen/raytrace_gpu.cu is
// produced by oracle/build_ref.py with the mechanical CUDA-12 patch (texture
// object for the removed texture reference, linear RGBA8 store for the removed
// surface reference, runtime static_samples, zero-initialised IntersectionData);
// this file only adds the headless plumbing GPUProcessor used to provide
// (cuda_opengl/src/gpu_processor.cpp:68-238): the device pointer tree, the cubemap
// array, and entry points with the flat C-ABI scene description.
#include "raytrace_gpu.cu" // generated into oracle/_ref/gen/

#include <cstdio>
#include <vector>

#include "ptcore.h"
#define PTCORE_RAYTRACE_NO_SETUP // raytrace_gpu.cu above has the reference's own
#include "ptcore_raytrace.hpp" // the product's source-compatible raytrace(uchar4*, ...) overload

namespace {

struct RefGpu
{
  scene::Scenes scenes;
  std::vector<void*> allocs;
  cudaArray_t cube_array = nullptr;
  cudaTextureObject_t cube_tex = 0;
  uint32_t n_faces = 0, n_lights = 0;
  scene::SceneData h_sd;
  std::vector<scene::Mesh> h_meshes; // host copy of the device Mesh table (device Face pointers)
};

#define CK(x)                                                                                  \
  do {                                                                                         \
    cudaError_t e_ = (x);                                                                      \
    if (e_ != cudaSuccess) {                                                                   \
      fprintf(stderr, "ref_gpu: %s -> %s\n", #x, cudaGetErrorString(e_));                      \
      return (int)e_;                                                                          \
    }                                                                                          \
  } while (0)

template <typename T>
int
upload(RefGpu* g, const T* host, size_t n, T** dev)
{
  *dev = nullptr;
  if (!n)
    return 0;
  CK(cudaMalloc(dev, n * sizeof(T)));
  g->allocs.push_back(*dev);
  CK(cudaMemcpy(*dev, host, n * sizeof(T), cudaMemcpyHostToDevice));
  return 0;
}

// Brute-force closest hit around the reference's own device intersectTriangle /
// intersectSphere with the acceptance rules of intersect() (intersection.cuh:179-212);
// intersect() does not report which face won, so the loop is repeated to get ids.
__global__ void
ref_hits_kernel(const scene::SceneData* sd, uint32_t n_faces, const float3* origins,
                const float3* dirs, uint64_t n, int32_t* ids, float* ts)
{
  uint64_t k = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
  if (k >= n)
    return;
  scene::Ray r;
  r.origin = origins[k];
  r.dir = dirs[k];
  float dist = 100000.0f, inter_dist = 100000.0f;
  int32_t id = -1, gi = 0;
  float3 nrm;
  float2 uv;
  for (size_t m = 0; m < sd->meshes.size; ++m) {
    const auto& mesh = sd->meshes.data[m];
    for (size_t i = 0; i < mesh.faces.size; ++i, ++gi)
      if (intersectTriangle(mesh.faces.data[i], nrm, uv, r, inter_dist) && inter_dist < dist &&
          inter_dist > 0.0) {
        dist = inter_dist;
        id = gi;
      }
  }
  for (unsigned l = 0; l < sd->lights.size; ++l)
    if (intersectSphere(r, sd->lights.data[l], inter_dist) && inter_dist < dist &&
        inter_dist >= 0.0) {
      dist = inter_dist;
      id = (int32_t)(n_faces + l);
    }
  if (ids)
    ids[k] = id;
  if (ts)
    ts[k] = dist;
}

// The same answer for meshes too large for one thread per ray (10 M faces): a block of 256 rays
// walks ONE chunk of one mesh's faces, 64 at a time through shared memory, with the reference's
// own intersectTriangle; the winner of intersect()'s in-order scan (accept t < dist && t > 0,
// intersection.cuh:183-184) is the minimum t with the LOWEST face index among equal t, i.e. the
// minimum of the 64-bit key (bits(t) << 32 | global face index) -- positive floats order like
// their bit patterns -- so chunks combine with atomicMin in any order.
constexpr int kBfRays = 256, kBfTile = 64;

__global__ void __launch_bounds__(kBfRays)
ref_hits_tiled_kernel(const scene::Face* faces, uint32_t n_mesh_faces, uint32_t first_global_id,
                      uint32_t faces_per_chunk, const float3* origins, const float3* dirs, uint64_t n,
                      unsigned long long* keys)
{
  __shared__ uint4 s_faces[kBfTile * 7]; // 64 x scene::Face (112 B)
  static_assert(sizeof(scene::Face) == 112, "Face layout");
  const uint64_t k = blockIdx.x * (uint64_t)kBfRays + threadIdx.x;
  const bool live = k < n;
  scene::Ray r;
  if (live) {
    r.origin = origins[k];
    r.dir = dirs[k];
  }
  const uint32_t f0 = blockIdx.y * faces_per_chunk;
  const uint32_t f1 = f0 + faces_per_chunk < n_mesh_faces ? f0 + faces_per_chunk : n_mesh_faces;
  unsigned long long best = ~0ull;
  float3 nrm;
  float2 uv;
  for (uint32_t base = f0; base < f1; base += kBfTile) {
    const uint32_t cnt = f1 - base < (uint32_t)kBfTile ? f1 - base : (uint32_t)kBfTile;
    __syncthreads();
    const uint4* src = reinterpret_cast<const uint4*>(faces + base);
    for (uint32_t i = threadIdx.x; i < cnt * 7u; i += kBfRays)
      s_faces[i] = src[i];
    __syncthreads();
    if (!live)
      continue;
    const scene::Face* sf = reinterpret_cast<const scene::Face*>(s_faces);
    for (uint32_t i = 0; i < cnt; ++i) {
      float t;
      if (intersectTriangle(sf[i], nrm, uv, r, t) && t < 100000.0f && t > 0.0) {
        const unsigned long long key =
          ((unsigned long long)__float_as_uint(t) << 32) | (unsigned long long)(first_global_id + base + i);
        best = key < best ? key : best;
      }
    }
  }
  if (live && best != ~0ull)
    atomicMin(keys + k, best);
}

__global__ void
ref_hits_finish_kernel(const scene::SceneData* sd, uint32_t n_faces, const float3* origins, const float3* dirs,
                       uint64_t n, const unsigned long long* keys, int32_t* ids, float* ts)
{
  uint64_t k = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
  if (k >= n)
    return;
  scene::Ray r;
  r.origin = origins[k];
  r.dir = dirs[k];
  float dist = 100000.0f, inter_dist = 100000.0f;
  int32_t id = -1;
  if (keys[k] != ~0ull) {
    dist = __uint_as_float((unsigned)(keys[k] >> 32));
    id = (int32_t)(keys[k] & 0xFFFFFFFFull);
  }
  for (unsigned l = 0; l < sd->lights.size; ++l) // intersection.cuh:199-212
    if (intersectSphere(r, sd->lights.data[l], inter_dist) && inter_dist < dist && inter_dist >= 0.0) {
      dist = inter_dist;
      id = (int32_t)(n_faces + l);
    }
  if (ids)
    ids[k] = id;
  if (ts)
    ts[k] = dist;
}

// Primary rays exactly as `kernel` generates them (raytrace.cu:218-240).
__global__ void
ref_primary_kernel(unsigned width, unsigned height, scene::Camera cam, unsigned hash_seed,
                   float3* origins, float3* dirs)
{
  const int x = blockDim.x * blockIdx.x + threadIdx.x;
  const int y = blockDim.y * blockIdx.y + threadIdx.y;
  if (x >= width || y >= height)
    return;
  const unsigned int tid = (blockIdx.x + blockIdx.y * gridDim.x) * (blockDim.x * blockDim.y) +
                           (threadIdx.y * blockDim.x) + threadIdx.x;
  curandState rand_state;
  curand_init(hash_seed + tid, 0, 0, &rand_state);
  scene::Ray r = generateRay(x, y, width / 2, height / 2, cam);
  camera_dof(r, cam, &rand_state);
  origins[(size_t)y * width + x] = r.origin;
  dirs[(size_t)y * width + x] = r.dir;
}

__global__ void
ref_rng_kernel(const uint32_t* seeds, uint32_t n_seeds, uint32_t n_draws, float* out)
{
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_seeds)
    return;
  curandState st;
  curand_init(seeds[i], 0, 0, &st);
  for (uint32_t k = 0; k < n_draws; ++k)
    out[(size_t)i * n_draws + k] = curand_uniform(&st);
}

} // namespace

extern "C" {

int
ref_gpu_create(const ptc_scene_desc* d, const float* cube_faces, unsigned cube_size, void** out)
{
  RefGpu* g = new RefGpu();
  *out = g;
  // Scenes -> SceneData*[] -> Mesh[] -> Face[]   (scene.cpp:202-283, 370-391)
  std::vector<scene::Mesh> meshes(d->n_meshes);
  size_t off = 0;
  for (uint32_t m = 0; m < d->n_meshes; ++m) {
    meshes[m].faces.size = d->mesh_face_counts[m];
    int rc = upload(g, (const scene::Face*)d->faces + off, d->mesh_face_counts[m],
                    &meshes[m].faces.data);
    if (rc)
      return rc;
    off += d->mesh_face_counts[m];
  }
  g->n_faces = (uint32_t)off;
  g->n_lights = d->n_lights;
  g->h_meshes = meshes;
  scene::SceneData& sd = g->h_sd;
  sd.meshes.size = d->n_meshes;
  sd.materials.size = d->n_materials;
  sd.lights.size = d->n_lights;
  int rc;
  if ((rc = upload(g, meshes.data(), meshes.size(), &sd.meshes.data)))
    return rc;
  if ((rc = upload(g, (const scene::Material*)d->materials, d->n_materials, &sd.materials.data)))
    return rc;
  if ((rc = upload(g, (const scene::LightProp*)d->lights, d->n_lights, &sd.lights.data)))
    return rc;
  scene::SceneData* d_sd;
  if ((rc = upload(g, &sd, 1, &d_sd)))
    return rc;
  if ((rc = upload(g, &d_sd, 1, &g->scenes.scenes))) // gpu_processor.cpp:184-203
    return rc;
  // textures (gpu_processor.cpp:205-238)
  std::vector<scene::Texture> tex(d->n_textures);
  for (uint32_t i = 0; i < d->n_textures; ++i) {
    tex[i].w = d->textures[i].w;
    tex[i].h = d->textures[i].h;
    tex[i].nb_chan = d->textures[i].nb_chan;
    size_t n = (size_t)tex[i].w * tex[i].h * tex[i].nb_chan;
    if ((rc = upload(g, d->textures[i].data, n, &tex[i].data)))
      return rc;
  }
  g->scenes.textures.size = d->n_textures;
  if ((rc = upload(g, tex.data(), tex.size(), &g->scenes.textures.data)))
    return rc;
  // cubemap (gpu_processor.cpp:134-152) + the texture state raytrace() used to set
  // on the texture reference (raytrace.cu:304-309)
  cudaChannelFormatDesc desc = cudaCreateChannelDesc(32, 32, 32, 32, cudaChannelFormatKindFloat);
  CK(cudaMalloc3DArray(&g->cube_array, &desc, make_cudaExtent(cube_size, cube_size, 6),
                       cudaArrayCubemap));
  cudaMemcpy3DParms p;
  memset(&p, 0, sizeof(p));
  p.srcPtr = make_cudaPitchedPtr((void*)cube_faces, cube_size * sizeof(float) * 4, cube_size,
                                 cube_size);
  p.dstArray = g->cube_array;
  p.extent = make_cudaExtent(cube_size, cube_size, 6);
  p.kind = cudaMemcpyHostToDevice;
  CK(cudaMemcpy3D(&p));
  cudaResourceDesc res;
  memset(&res, 0, sizeof(res));
  res.resType = cudaResourceTypeArray;
  res.res.array.array = g->cube_array;
  cudaTextureDesc td;
  memset(&td, 0, sizeof(td));
  td.addressMode[0] = cudaAddressModeWrap;
  td.addressMode[1] = cudaAddressModeWrap;
  td.filterMode = cudaFilterModeLinear;
  td.normalizedCoords = 1;
  td.readMode = cudaReadModeElementType;
  CK(cudaCreateTextureObject(&g->cube_tex, &res, &td, nullptr));
  CK(cudaMemcpyToSymbol(cubemap_ref, &g->cube_tex, sizeof(g->cube_tex)));
  setupFunctionTables();
  return 0;
}

void
ref_gpu_destroy(void* h)
{
  RefGpu* g = (RefGpu*)h;
  if (!g)
    return;
  if (g->cube_tex)
    cudaDestroyTextureObject(g->cube_tex);
  if (g->cube_array)
    cudaFreeArray(g->cube_array);
  for (void* p : g->allocs)
    cudaFree(p);
  delete g;
}

static int
set_launch_state(RefGpu* g, unsigned* d_out, unsigned bounces)
{
  int static_samples = (int)bounces - 2; // max_bounces = 1 + (static_samples + 1)
  // the reference re-binds its cubemap on every raytrace() call (raytrace.cu:304-309);
  // several RefGpu handles may be alive, so bind this handle's texture per launch too
  CK(cudaMemcpyToSymbol(cubemap_ref, &g->cube_tex, sizeof(g->cube_tex)));
  CK(cudaMemcpyToSymbol(ref_out_rgba, &d_out, sizeof(d_out)));
  CK(cudaMemcpyToSymbol(g_ref_static_samples, &static_samples, sizeof(int)));
  return 0;
}

// The reference's public entry point raytrace() (raytrace.cu:287-325), including its
// function-static frame counter.  d_accum / d_out are device pointers.
int
ref_gpu_raytrace(void* h, const ptc_camera* camera, unsigned width, unsigned height,
                 unsigned bounces, int moved, unsigned post_id, float* d_accum, unsigned* d_out,
                 void* stream)
{
  RefGpu* g = (RefGpu*)h;
  int rc = set_launch_state(g, d_out, bounces);
  if (rc)
    return rc;
  std::vector<scene::Cubemap> cubemaps(1);
  cubemaps[0].cubemap = g->cube_array;
  scene::Camera cam;
  memcpy(&cam, camera, sizeof(cam));
  raytrace((cudaArray_const_t) nullptr, g->scenes, 0, cubemaps, 0, &cam, width, height, (cudaStream_t)stream,
           (float3*)d_accum, moved != 0, post_id);
  return (int)cudaGetLastError();
}

// The PRODUCT behind the reference's call: the same Scenes tree, cubemap vector and camera
// GPUProcessor::render passes (gpu_processor.cpp:375-377) go to the raytrace() overload of
// include/ptcore_raytrace.hpp, which derives its own state from this device tree.  Lets the
// parity tests check the drop-in boundary itself, not only the C ABI underneath.
int
ref_gpu_raytrace_compat(void* h, const ptc_camera* camera, unsigned width, unsigned height, int moved,
                        unsigned post_id, float* d_accum, unsigned* d_out, void* stream)
{
  RefGpu* g = (RefGpu*)h;
  std::vector<scene::Cubemap> cubemaps(1);
  cubemaps[0].cubemap = g->cube_array;
  scene::Camera cam;
  memcpy(&cam, camera, sizeof(cam));
  return (int)raytrace((uchar4*)d_out, g->scenes, 0, cubemaps, 0, &cam, width, height,
                       (cudaStream_t)stream, (float3*)d_accum, moved != 0, post_id);
}

void
ref_gpu_compat_reset()
{
  ptc_compat::invalidate();
}

// The overload's several-GPU opt-in (ptc_compat::use_devices) and the group's accumulator, so the
// parity test can compare the samples too.
int
ref_gpu_compat_use_devices(const int* devices, unsigned n)
{
  return (int)ptc_compat::use_devices(devices, n);
}

int
ref_gpu_compat_gather_accum(float* host_accum)
{
  ptc_group* g = ptc_compat::registry().group;
  if (!g)
    return -1;
  int rc = ptc_group_sync(g);
  return rc ? rc : ptc_group_gather_accum(g, host_accum);
}

// Same launch as raytrace.cu:315-322 but for an explicit sample number, so tests
// can compare frame k without replaying frames 1..k-1.
int
ref_gpu_kernel(void* h, const ptc_camera* camera, unsigned width, unsigned height,
               unsigned frame_nb, unsigned bounces, int moved, unsigned post_id, float* d_accum,
               unsigned* d_out, void* stream)
{
  RefGpu* g = (RefGpu*)h;
  int rc = set_launch_state(g, d_out, bounces);
  if (rc)
    return rc;
  scene::Camera cam;
  memcpy(&cam, camera, sizeof(cam));
  dim3 threads_per_block(16, 16);
  dim3 nb_blocks(width / threads_per_block.x + 1, height / threads_per_block.y + 1);
  kernel<<<nb_blocks, threads_per_block, 0, (cudaStream_t)stream>>>(
    width, height, g->scenes, 0, cam, WangHash(frame_nb), (int)frame_nb, (float3*)d_accum,
    moved != 0, h_post_process_table[post_id]);
  return (int)cudaGetLastError();
}

// Host-buffer conveniences for the parity tests.
int
ref_gpu_primary(void* h, const ptc_camera* camera, unsigned width, unsigned height,
                unsigned frame_nb, float* origins, float* dirs, int32_t* ids, float* ts)
{
  RefGpu* g = (RefGpu*)h;
  size_t n = (size_t)width * height;
  float3 *d_o, *d_d;
  int32_t* d_i;
  float* d_t;
  CK(cudaMalloc(&d_o, n * 12));
  CK(cudaMalloc(&d_d, n * 12));
  CK(cudaMalloc(&d_i, n * 4));
  CK(cudaMalloc(&d_t, n * 4));
  scene::Camera cam;
  memcpy(&cam, camera, sizeof(cam));
  dim3 tpb(16, 16), nb(width / 16 + 1, height / 16 + 1);
  ref_primary_kernel<<<nb, tpb>>>(width, height, cam, WangHash(frame_nb), d_o, d_d);
  scene::SceneData* d_sd;
  CK(cudaMemcpy(&d_sd, g->scenes.scenes, sizeof(d_sd), cudaMemcpyDeviceToHost));
  ref_hits_kernel<<<(unsigned)((n + 127) / 128), 128>>>(d_sd, g->n_faces, d_o, d_d, n, d_i, d_t);
  CK(cudaDeviceSynchronize());
  if (origins)
    CK(cudaMemcpy(origins, d_o, n * 12, cudaMemcpyDeviceToHost));
  if (dirs)
    CK(cudaMemcpy(dirs, d_d, n * 12, cudaMemcpyDeviceToHost));
  if (ids)
    CK(cudaMemcpy(ids, d_i, n * 4, cudaMemcpyDeviceToHost));
  if (ts)
    CK(cudaMemcpy(ts, d_t, n * 4, cudaMemcpyDeviceToHost));
  cudaFree(d_o), cudaFree(d_d), cudaFree(d_i), cudaFree(d_t);
  return 0;
}

int
ref_gpu_intersect(void* h, const float* origins, const float* dirs, uint64_t n, int32_t* ids,
                  float* ts)
{
  RefGpu* g = (RefGpu*)h;
  float3 *d_o, *d_d;
  int32_t* d_i;
  float* d_t;
  CK(cudaMalloc(&d_o, n * 12));
  CK(cudaMalloc(&d_d, n * 12));
  CK(cudaMalloc(&d_i, n * 4));
  CK(cudaMalloc(&d_t, n * 4));
  CK(cudaMemcpy(d_o, origins, n * 12, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(d_d, dirs, n * 12, cudaMemcpyHostToDevice));
  scene::SceneData* d_sd;
  CK(cudaMemcpy(&d_sd, g->scenes.scenes, sizeof(d_sd), cudaMemcpyDeviceToHost));
  ref_hits_kernel<<<(unsigned)((n + 127) / 128), 128>>>(d_sd, g->n_faces, d_o, d_d, n, d_i, d_t);
  CK(cudaDeviceSynchronize());
  if (ids)
    CK(cudaMemcpy(ids, d_i, n * 4, cudaMemcpyDeviceToHost));
  if (ts)
    CK(cudaMemcpy(ts, d_t, n * 4, cudaMemcpyDeviceToHost));
  cudaFree(d_o), cudaFree(d_d), cudaFree(d_i), cudaFree(d_t);
  return 0;
}

// Same contract as ref_gpu_intersect through the tiled brute-force kernels above (any mesh size).
int
ref_gpu_intersect_tiled(void* h, const float* origins, const float* dirs, uint64_t n, int32_t* ids,
                        float* ts)
{
  RefGpu* g = (RefGpu*)h;
  if (!n)
    return 0;
  float3 *d_o, *d_d;
  int32_t* d_i;
  float* d_t;
  unsigned long long* d_k;
  CK(cudaMalloc(&d_o, n * 12));
  CK(cudaMalloc(&d_d, n * 12));
  CK(cudaMalloc(&d_i, n * 4));
  CK(cudaMalloc(&d_t, n * 4));
  CK(cudaMalloc(&d_k, n * 8));
  CK(cudaMemcpy(d_o, origins, n * 12, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(d_d, dirs, n * 12, cudaMemcpyHostToDevice));
  CK(cudaMemset(d_k, 0xFF, n * 8));
  const unsigned ray_blocks = (unsigned)((n + kBfRays - 1) / kBfRays);
  uint32_t first = 0;
  for (const scene::Mesh& m : g->h_meshes) {
    const uint32_t nf = m.faces.size;
    if (nf) {
      // enough chunks to fill the machine whatever the ray count; chunk = whole tiles
      uint32_t chunks = (2048u + ray_blocks - 1) / ray_blocks;
      const uint32_t tiles = (nf + kBfTile - 1) / kBfTile;
      chunks = chunks > tiles ? tiles : chunks;
      const uint32_t per = ((tiles + chunks - 1) / chunks) * kBfTile;
      chunks = (nf + per - 1) / per;
      ref_hits_tiled_kernel<<<dim3(ray_blocks, chunks), kBfRays>>>(m.faces.data, nf, first, per, d_o, d_d, n, d_k);
    }
    first += nf;
  }
  scene::SceneData* d_sd;
  CK(cudaMemcpy(&d_sd, g->scenes.scenes, sizeof(d_sd), cudaMemcpyDeviceToHost));
  ref_hits_finish_kernel<<<(unsigned)((n + 127) / 128), 128>>>(d_sd, g->n_faces, d_o, d_d, n, d_k, d_i, d_t);
  CK(cudaDeviceSynchronize());
  if (ids)
    CK(cudaMemcpy(ids, d_i, n * 4, cudaMemcpyDeviceToHost));
  if (ts)
    CK(cudaMemcpy(ts, d_t, n * 4, cudaMemcpyDeviceToHost));
  cudaFree(d_o), cudaFree(d_d), cudaFree(d_i), cudaFree(d_t), cudaFree(d_k);
  return 0;
}

// cuRAND XORWOW known answers from the device (raytrace.cu:234-235 call pattern).
int
ref_gpu_rng(const uint32_t* seeds, uint32_t n_seeds, uint32_t n_draws, float* out)
{
  uint32_t* d_s;
  float* d_o;
  CK(cudaMalloc(&d_s, n_seeds * 4));
  CK(cudaMalloc(&d_o, (size_t)n_seeds * n_draws * 4));
  CK(cudaMemcpy(d_s, seeds, n_seeds * 4, cudaMemcpyHostToDevice));
  ref_rng_kernel<<<(n_seeds + 63) / 64, 64>>>(d_s, n_seeds, n_draws, d_o);
  CK(cudaDeviceSynchronize());
  CK(cudaMemcpy(out, d_o, (size_t)n_seeds * n_draws * 4, cudaMemcpyDeviceToHost));
  cudaFree(d_s), cudaFree(d_o);
  return 0;
}

} // extern "C"
