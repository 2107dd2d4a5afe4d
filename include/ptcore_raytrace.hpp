// ptcore_raytrace.hpp -- source-compatible C++ replacement for the reference's
//
//   cudaError_t raytrace(cudaArray_const_t array, const scene::Scenes&, unsigned scene_id,
//                        const std::vector<scene::Cubemap>&, int cubemap_id, const scene::Camera*,
//                        unsigned width, unsigned height, cudaStream_t, float3* temporal_framebuffer,
//                        bool moved, unsigned post_id);      (cuda_opengl/include/shaders/raytrace.h:9-15)
//   void setupFunctionTables();                              (raytrace.h:17)
//
// Include it from gpu_processor.cpp INSTEAD of "shaders/raytrace.h", after "scene/scene_data.h".
// The only signature change is the first argument: a linear device uchar4* (pitch = width * 4,
// row 0 = image top) instead of the GL-backed cudaArray -- there is no GL interop in a headless
// build.  Everything else, including the argument order and the reference's "always returns
// cudaSuccess, caller checks cudaGetLastError afterwards" convention (raytrace.cu:324,
// cuda_helper.h:9-17), is kept; a failure inside the core is reported on stderr and returned as
// cudaErrorUnknown so that a caller that DOES look at the result sees it.
//
// The core needs derived state (BVH, split records) the reference's `Scenes` tree cannot carry.
// It is built lazily, once per (device, scene_id) and per cubemap id, from the caller's own
// device tree (ptc_upload_scene_from_device / ptc_set_cubemap_from_array in ptcore.h), so no
// other line of the reference has to change.  Like the reference's function-static frame counter
// (raytrace.cu:296-300) the registry is process-wide and not thread-safe.
#pragma once

#include <cuda_runtime.h>

#include <cstdio>
#include <map>
#include <utility>
#include <vector>

#include "ptcore.h"

namespace ptc_compat {

struct Registry
{
  ptc_group* group = nullptr;                                // use_devices(): every call drives this group
  std::map<int, ptc_context*> ctx;                           // device -> context
  std::map<std::pair<int, unsigned>, int32_t> scene;         // (device, scene_id) -> core id
  std::map<std::pair<int, const void*>, int32_t> cubemap;    // (device, cudaArray) -> core id
  ~Registry()
  {
    for (auto& kv : ctx)
      ptc_destroy(kv.second);
    if (group)
      ptc_group_destroy(group);
  }
};

inline Registry&
registry()
{
  static Registry r;
  return r;
}

inline cudaError_t
report(ptc_context* c, const char* what, int rc)
{
  std::fprintf(stderr, "ptcore: %s failed (%d): %s\n", what, rc, c ? ptc_last_error(c) : "no context");
  return cudaErrorUnknown;
}

// Forget every derived scene / cubemap (call after the caller re-uploads its Scenes tree); also
// leaves the several-GPU mode.
inline void
invalidate()
{
  Registry& r = registry();
  for (auto& kv : r.ctx)
    ptc_destroy(kv.second);
  if (r.group)
    ptc_group_destroy(r.group);
  r.group = nullptr;
  r.ctx.clear();
  r.scene.clear();
  r.cubemap.clear();
}

// Opt-in: every later raytrace() call renders on ALL of `devices` (ptc_group_*, ptcore.h): the
// image is split into round-robin 16-row strips, devices[0] must be the device the caller's Scenes
// tree, cubemap arrays and output image live on (the current device of the reference's
// GPUProcessor), every device's kernels store their pixels straight into the caller's image there,
// and the result is bit-identical to one GPU.  The group owns one accumulator per device:
// `temporal_framebuffer` is then neither read nor written (read the samples back with
// ptc_group_gather_accum(ptc_compat::registry().group, ...)).  The frame counter follows
// raytrace.cu:296-300 as before.
inline cudaError_t
use_devices(const int* devices, unsigned n)
{
  invalidate();
  Registry& r = registry();
  const int rc = ptc_group_create(devices, n, &r.group);
  if (rc) {
    std::fprintf(stderr, "ptcore: ptc_group_create failed (%d)\n", rc);
    r.group = nullptr;
    return cudaErrorUnknown;
  }
  return cudaSuccess;
}

inline cudaError_t
report_group(ptc_group* g, const char* what, int rc)
{
  std::fprintf(stderr, "ptcore: %s failed (%d): %s\n", what, rc, ptc_group_last_error(g));
  return cudaErrorUnknown;
}

} // namespace ptc_compat

#ifndef PTCORE_RAYTRACE_NO_SETUP // define when the reference's raytrace.cu is in the same build
inline void
setupFunctionTables()
{
  // nothing to set up: post_id selects the post-process inside the kernels (raytrace.cu:354-375)
}
#endif

template <typename Scenes, typename Cubemap, typename Camera>
inline cudaError_t
raytrace(uchar4* out_rgba, const Scenes& scenes, unsigned int scene_id, const std::vector<Cubemap>& cubemaps,
         int cubemap_id, const Camera* const cam, const unsigned int width, const unsigned int height,
         cudaStream_t stream, float3* temporal_framebuffer, bool moved, unsigned int post_id)
{
  static_assert(sizeof(Scenes) == sizeof(ptc_ref_scenes), "scene::Scenes layout (scene_data.h:83-87)");
  static_assert(sizeof(Camera) == sizeof(ptc_camera), "scene::Camera layout (scene_data.h:123-133)");
  ptc_compat::Registry& r = ptc_compat::registry();
  if (r.group) { // several GPUs behind the same call (ptc_compat::use_devices)
    ptc_group* g = r.group;
    int rc;
    auto sk = std::make_pair(-1, scene_id);
    auto sit = r.scene.find(sk);
    if (sit == r.scene.end()) {
      int32_t id;
      if ((rc = ptc_group_upload_scene_from_device(g, reinterpret_cast<const ptc_ref_scenes*>(&scenes), scene_id, &id)))
        return ptc_compat::report_group(g, "ptc_group_upload_scene_from_device", rc);
      sit = r.scene.emplace(sk, id).first;
    }
    if (cubemap_id < 0 || (size_t)cubemap_id >= cubemaps.size())
      return ptc_compat::report_group(g, "cubemap_id out of range", PTC_ERR_INVALID);
    auto ck = std::make_pair(-1, (const void*)cubemaps[cubemap_id].cubemap);
    auto cit = r.cubemap.find(ck);
    if (cit == r.cubemap.end()) {
      int32_t id;
      if ((rc = ptc_group_set_cubemap_from_array(g, cubemaps[cubemap_id].cubemap, &id)))
        return ptc_compat::report_group(g, "ptc_group_set_cubemap_from_array", rc);
      cit = r.cubemap.emplace(ck, id).first;
    }
    ptc_render_args a{};
    a.scene_id = sit->second;
    a.cubemap_id = cit->second;
    a.camera = reinterpret_cast<const ptc_camera*>(cam);
    a.width = width;
    a.height = height;
    a.moved = moved ? 1u : 0u;
    a.post_id = post_id;
    a.out_rgba = out_rgba; // the caller's image on devices[0]; accumulators are the group's
    a.stream = stream;
    if ((rc = ptc_group_render(g, &a)))
      return ptc_compat::report_group(g, "ptc_group_render", rc);
    return cudaSuccess;
  }
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess)
    return cudaErrorUnknown;
  ptc_context*& ctx = r.ctx[dev];
  int rc;
  if (!ctx && (rc = ptc_create(dev, &ctx)))
    return ptc_compat::report(nullptr, "ptc_create", rc);
  auto sk = std::make_pair(dev, scene_id);
  auto sit = r.scene.find(sk);
  if (sit == r.scene.end()) {
    int32_t id;
    if ((rc = ptc_upload_scene_from_device(ctx, reinterpret_cast<const ptc_ref_scenes*>(&scenes), scene_id, &id)))
      return ptc_compat::report(ctx, "ptc_upload_scene_from_device", rc);
    sit = r.scene.emplace(sk, id).first;
  }
  if (cubemap_id < 0 || (size_t)cubemap_id >= cubemaps.size())
    return ptc_compat::report(ctx, "cubemap_id out of range", PTC_ERR_INVALID);
  auto ck = std::make_pair(dev, (const void*)cubemaps[cubemap_id].cubemap);
  auto cit = r.cubemap.find(ck);
  if (cit == r.cubemap.end()) {
    int32_t id;
    if ((rc = ptc_set_cubemap_from_array(ctx, cubemaps[cubemap_id].cubemap, &id)))
      return ptc_compat::report(ctx, "ptc_set_cubemap_from_array", rc);
    cit = r.cubemap.emplace(ck, id).first;
  }
  ptc_render_args a{};
  a.scene_id = sit->second;
  a.cubemap_id = cit->second;
  a.camera = reinterpret_cast<const ptc_camera*>(cam);
  a.width = width;
  a.height = height;
  a.bounces = 0; // the reference's 1 + (static_samples + 1) = 3 (raytrace.cu:66, 243)
  a.moved = moved ? 1u : 0u;
  a.post_id = post_id;
  a.frame_nb = 0; // the context's counter == the reference's function-static seed
  a.accum = temporal_framebuffer;
  a.out_rgba = out_rgba;
  a.stream = stream;
  if ((rc = ptc_render(ctx, &a)))
    return ptc_compat::report(ctx, "ptc_render", rc);
  return cudaSuccess;
}
