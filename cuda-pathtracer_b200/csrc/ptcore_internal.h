// ptcore_internal.h -- what group.cu needs from ptcore.cu beyond the public C ABI.
#pragma once

#include <functional>
#include <string>
#include <vector>

#include "bvh.h"
#include "ptcore.h"

namespace ptc {

// Host-side result of the scene preparation (validation, BVH build, shading records): built ONCE
// and uploaded to every GPU of a group (the reference uploads one scene to one GPU,
// scene.cpp:370-391; at 10 M triangles the build is 0.6 s of host time per copy otherwise).
struct PreparedScene
{
  HostBvh bvh;
  RawVector<ShadeRecord> shade;
  uint32_t n_faces = 0;
};

// returns PTC_OK or PTC_ERR_INVALID with a message
int prepare_scene(const ptc_scene_desc* desc, PreparedScene& out, std::string& err);
int upload_prepared(ptc_context* ctx, const ptc_scene_desc* desc, const PreparedScene& ps, int32_t* scene_id);
int context_device(const ptc_context* ctx);
// the reference's device tree / cubemap array read back through ctx's device and handed to `sink`
int scene_from_device(ptc_context* ctx, const ptc_ref_scenes* scenes, uint32_t scene_index,
                      const std::function<int(const ptc_scene_desc*)>& sink);
int cubemap_from_array(ptc_context* ctx, const void* cuda_array,
                       const std::function<int(const float*, uint32_t)>& sink);

} // namespace ptc
