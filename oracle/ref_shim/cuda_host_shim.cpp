// TEST INFRASTRUCTURE (oracle/_ref recipe) -- not product code.
//
// Link-time stand-in for the handful of CUDA runtime entry points that the
// reference's *host* loader calls (cuda_opengl/src/scene/scene.cpp:165-168,
// 190-196, 269-281, 386-389, 402-419).  "Device" memory becomes plain host
// memory so the reference loader can run in a container without a GPU and its
// Face[]/Material[]/LightProp[] output can be dumped and compared.
#include <cuda_runtime.h>

#include <cstdlib>
#include <cstring>

extern "C" {

cudaError_t cudaMalloc(void** p, size_t n)
{
  *p = std::malloc(n ? n : 1);
  return *p ? cudaSuccess : cudaErrorMemoryAllocation;
}

cudaError_t cudaFree(void* p)
{
  std::free(p);
  return cudaSuccess;
}

cudaError_t cudaMemcpy(void* dst, const void* src, size_t n, cudaMemcpyKind)
{
  std::memcpy(dst, src, n);
  return cudaSuccess;
}

cudaError_t cudaMemset(void* dst, int v, size_t n)
{
  std::memset(dst, v, n);
  return cudaSuccess;
}

cudaError_t cudaGetLastError(void) { return cudaSuccess; }

const char* cudaGetErrorString(cudaError_t) { return "shim: no error"; }

} // extern "C"
