// TEST INFRASTRUCTURE (oracle/_ref recipe) -- not product code.
//
// Force-included (-include) ahead of the reference's raytrace.cu when that file
// is compiled by g++ as plain C++.  It supplies host stand-ins for the device
// built-ins the reference kernel touches so the reference's own radiance()/
// intersect()/kernel() code (cuda_opengl/src/shaders/raytrace.cu:41-271,
// cuda_opengl/include/shaders/intersection.cuh:20-246) can be executed on host
// cores as the ground truth the from-scratch oracle is pinned against.
//
// What is NOT the reference here (and therefore differs from a device run by
// a few ulp): __fdividef -> IEEE divide, __sinf/__cosf -> libm sinf/cosf,
// texCubemap -> nearest/bilinear software fetch.  The oracle states the same
// substitutions, so oracle == this build bit-for-bit.
#pragma once

#include <cuda_runtime.h>

#include <cmath>
#include <cstdint>
#include <cstring>

// ---- launch geometry built-ins (raytrace.cu:221-229) -----------------------
// curand_mtgp32_kernel.h declares `extern const` built-ins of the same names;
// include it first, then rename ours out of the way.
#include <curand_kernel.h>
#define threadIdx ref_threadIdx
#define blockIdx ref_blockIdx
#define blockDim ref_blockDim
#define gridDim ref_gridDim
extern thread_local uint3 threadIdx, blockIdx;
extern dim3 blockDim, gridDim;
extern int g_ref_threads; // host threads used by ref_launch

// Stand-in for `kernel<<<grid, block, 0, stream>>>(args...)` (raytrace.cu:320):
// runs every (block, thread) of the launch on host threads.
void ref_parallel_blocks(unsigned n_blocks, void (*fn)(unsigned, void*), void* ctx);

template <typename K, typename... A>
static inline void
ref_launch(K k, dim3 grid, dim3 block, A... args)
{
  gridDim = grid;
  blockDim = block;
  auto body = [&](unsigned b) {
    blockIdx.x = b % grid.x;
    blockIdx.y = b / grid.x;
    blockIdx.z = 0;
    for (unsigned ty = 0; ty < block.y; ++ty)
      for (unsigned tx = 0; tx < block.x; ++tx) {
        threadIdx.x = tx;
        threadIdx.y = ty;
        threadIdx.z = 0;
        k(args...);
      }
  };
  using B = decltype(body);
  ref_parallel_blocks(grid.x * grid.y,
                      [](unsigned b, void* c) { (*static_cast<B*>(c))(b); }, &body);
}

// ---- fast-math intrinsics (raytrace.cu:111-122,159,206; intersection.cuh:113,151)
static inline float __fdividef(float a, float b) { return a / b; }
static inline float __fsqrt_rn(float a) { return sqrtf(a); }
// glibc's <math.h> already declares extern __sinf/__cosf aliases: go through macros.
static inline float ref_fast_sinf(float a) { return sinf(a); }
static inline float ref_fast_cosf(float a) { return cosf(a); }
#define __sinf(x) ref_fast_sinf(x)
#define __cosf(x) ref_fast_cosf(x)

// ---- legacy texture / surface references (raytrace.cu:21-22,302-309) -------
template <typename T, int Dim = 1, int Mode = 0>
struct texture
{
  int addressMode[3];
  int filterMode;
  bool normalized;
};
template <typename T, int Dim = 1>
struct surface
{};

struct RefCubemapHost
{
  const float* faces; // +x,-x,+y,-y,+z,-z ; RGBA float ; size*size each
  unsigned size;
};
extern RefCubemapHost g_ref_cubemap;
extern uint32_t* g_ref_out_rgba; // row 0 = top, pitch = width
extern unsigned g_ref_out_width;
extern int g_ref_static_samples;

float4 ref_shim_texCubemap(float x, float y, float z);

template <typename Tex>
static inline float4
texCubemap(const Tex&, float x, float y, float z)
{
  return ref_shim_texCubemap(x, y, z);
}

template <typename Surf>
static inline void
surf2Dwrite(uint1 v, const Surf&, int x_bytes, int y, int)
{
  g_ref_out_rgba[(size_t)y * g_ref_out_width + (size_t)x_bytes / 4] = v.x;
}

template <typename A, typename B>
static inline cudaError_t
cudaBindSurfaceToArray(const A&, const B&)
{
  return cudaSuccess;
}
template <typename A, typename B, typename C>
static inline cudaError_t
cudaBindTextureToArray(const A&, const B&, const C&)
{
  return cudaSuccess;
}
template <typename T>
static inline cudaError_t
cudaMemcpyFromSymbol(void* dst, const T& sym, size_t n)
{
  std::memcpy(dst, &sym, n);
  return cudaSuccess;
}
