/* ptcore.h -- C ABI of the B200-native path-tracing core.
 *
 * Drop-in boundary for the hot path of DavidPeicho/cuda-pathtracer: everything
 * that sits behind
 *
 *     cudaError_t raytrace(cudaArray_const_t, const scene::Scenes&, unsigned scene_id,
 *                          const std::vector<scene::Cubemap>&, int cubemap_id,
 *                          const scene::Camera*, unsigned w, unsigned h, cudaStream_t,
 *                          float3* temporal_framebuffer, bool moved, unsigned post_id)
 *     (reference: cuda_opengl/include/shaders/raytrace.h:9-15, sole caller
 *      cuda_opengl/src/gpu_processor.cpp:375-377)
 *
 * Plain C: pointers and sizes only, no C++/torch types.  All record types below
 * are layout-identical to the reference's device data contract
 * (cuda_opengl/include/scene/scene_data.h) so the reference's loader output can
 * be passed without repacking.  Every entry point returns 0 on success or a
 * negative ptc_status / positive cudaError_t; nothing exits or throws across the
 * boundary (the reference's cudaCheckError() exits, cuda_helper.h:9-17 -- the C
 * ABI reports instead and the C++ shim keeps the check-after convention).
 */
#ifndef PTCORE_H
#define PTCORE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* 3: device-tree upload, cube-cross cubemaps, PPM / accumulator files, read-bandwidth probe;
 * the structs of version 2 are unchanged.
 * 4: rgb_hex of the cube-cross calls: PTC_CUBEMAP_DEFAULT_COLOR selects the default colour, 0 is
 *    black; ptc_profile_frame no longer touches the caller's buffers and reports more;
 *    ptc_render_args._pad became `flags` (PTC_OUT_FULL_IMAGE); ptc_group_* (several GPUs behind
 *    one call), ptc_shared_* / ptc_flag_* (the same across processes).
 * 5: additions only -- ptc_render_static, ptc_flatten_faces / ptc_upload_scene_indexed,
 *    ptc_texpack_*, ptc_group_upload_scene_from_device / ptc_group_set_cubemap_from_array,
 *    ptc_host_image_* / ptc_memcpy_strips_to_host / ptc_memcpy_strips_to_device;
 *    ptc_scene_info.smem_resident now reports what it says.
 * 6: addition only -- ptc_write_png. */
#define PTC_ABI_VERSION 6

typedef enum ptc_status
{
  PTC_OK = 0,
  PTC_ERR_INVALID = -1,   /* bad argument / unknown id                      */
  PTC_ERR_NO_DEVICE = -2, /* no CUDA device: the core has NO CPU fallback    */
  PTC_ERR_ALLOC = -3,
  PTC_ERR_STATE = -4      /* e.g. render before upload                       */
} ptc_status;

/* == scene::Face, scene_data.h:46-53 (112 bytes, 16-aligned) */
typedef struct ptc_face
{
  float vertices[3][3];  /* @0   */
  float normals[3][3];   /* @36  */
  float texcoords[3][2]; /* @72  */
  float tangent[3];      /* @96  */
  uint32_t material_id;  /* @108 */
} ptc_face;

/* == scene::Material, scene_data.h:95-100 (12 bytes of payload, 8-aligned -> 16) */
typedef struct ptc_material
{
  int32_t diffuse_spec_map; /* RGBA float texture id: rgb = diffuse, a = specular mix */
  int32_t normal_map;       /* RGB float texture id or -1                            */
  float ior;                /* 1.0 = opaque                                          */
  uint32_t _pad;
} ptc_material;

/* == scene::LightProp, scene_data.h:109-115 (32 bytes) */
typedef struct ptc_light
{
  float color[3];
  float vec[3]; /* sphere centre */
  float emission;
  float radius;
} ptc_light;

/* == scene::Texture, scene_data.h:31-37 (32 bytes, 16-aligned).  `data` is a HOST
 * pointer here (linear floats as produced by stbi_loadf, nb_chan 3 or 4). */
typedef struct ptc_texture
{
  int32_t w, h, nb_chan;
  int32_t _pad0;
  const float* data;
  uint64_t _pad1;
} ptc_texture;

/* == scene::Camera, scene_data.h:123-133 (64 bytes).  u, v are outputs of ray
 * generation in the reference (intersection.cuh:84-87); ignored on input. */
typedef struct ptc_camera
{
  float position[3];
  float dir[3];
  float u[3];
  float v[3];
  float fov_x; /* radians */
  float speed;
  float aperture;
  float focus_dist;
} ptc_camera;

/* One scene = what scene::SceneData + the global texture table describe
 * (scene_data.h:71-87).  Faces are concatenated in mesh order; the global face
 * index g = sum(mesh_face_counts[0..m-1]) + i is the hit id reported by the
 * core, and the reference's first-index-wins tie rule
 * (intersection.cuh:179-196) is defined on this order. */
typedef struct ptc_scene_desc
{
  const ptc_face* faces;
  const uint32_t* mesh_face_counts;
  uint32_t n_meshes;
  const ptc_material* materials;
  uint32_t n_materials;
  const ptc_light* lights;
  uint32_t n_lights;
  const ptc_texture* textures; /* global texture table (Scenes.textures) */
  uint32_t n_textures;
} ptc_scene_desc;

/* Hit id encoding used by every entry point that reports hits. */
#define PTC_HIT_MISS (-1)
/* id in [0, n_faces)            : triangle, global face index
 * id in [n_faces, n_faces+n_l)  : sphere light (id - n_faces)               */

typedef struct ptc_render_args
{
  int32_t scene_id;
  int32_t cubemap_id;
  const ptc_camera* camera; /* host pointer, copied at call time             */
  uint32_t width, height;   /* full image size (defines RNG ids, see below)  */
  uint32_t row_begin, row_end; /* rows [row_begin,row_end) of the full image
                                  rendered by this call (image-tile sharding);
                                  0,0 means all rows                          */
  uint32_t bounces;   /* path segments per sample; 0 -> 3, the reference's
                         1 + (static_samples+1), raytrace.cu:66,243           */
  uint32_t moved;     /* !=0: preview frame + accumulator reset, raytrace.cu:54-62,255,298 */
  uint32_t post_id;   /* 0 none, 1 grayscale, 2 sepia, 3 invert, raytrace.cu:327-357 */
  uint32_t frame_nb;  /* 0: use the context's own counter (the reference's
                         function-static `seed`, raytrace.cu:296-300);
                         k>0: render sample k explicitly (hash_seed = WangHash(k)) */
  void* accum;        /* float3[(row_end-row_begin)*width] temporal framebuffer;
                         rows stored bottom-up exactly like raytrace.cu:252 within
                         the band.  DEVICE pointer for ptc_render, HOST (ideally
                         pinned) for ptc_render_host                           */
  void* out_rgba;     /* uchar4[(row_end-row_begin)*width], row 0 = top; may be NULL */
  void* radiance_out; /* optional float3 per pixel (top-down rows): this frame's
                         un-clamped radiance, for parity tests; may be NULL    */
  void* stream;       /* cudaStream_t; NULL = default stream                  */
  /* Interleaved image tiles (multi-GPU load balance).  strip_stride <= 1: the contiguous
   * band [row_begin,row_end) above.  strip_stride > 1: row_begin/row_end must be 0 and this
   * call renders the strips s = strip_phase, strip_phase + strip_stride, ... where strip s
   * is rows [s*strip_rows, min((s+1)*strip_rows, height)); the band seen through accum /
   * out_rgba / radiance_out is those strips stacked top-down (accum bottom-up inside it). */
  uint32_t strip_rows;   /* 0 -> 16, the reference's block height (raytrace.cu:315) */
  uint32_t strip_stride;
  uint32_t strip_phase;
  uint32_t flags;        /* PTC_OUT_FULL_IMAGE: out_rgba points at a FULL width x height image and every
                            pixel of the band is stored at its global row.  The image may be memory of
                            ANOTHER GPU (peer access or a CUDA IPC mapping, ptc_ipc_*): the tiles of
                            several GPUs then assemble one image with plain stores from the kernel's
                            epilogue, no collective (device-pointer calls only)                   */
} ptc_render_args;
#define PTC_OUT_FULL_IMAGE 1u

typedef struct ptc_counters
{
  uint64_t paths;          /* pixels rendered                                  */
  uint64_t segments;       /* radiance-loop iterations == reference intersect() calls */
  uint64_t traced;         /* closest-hit queries actually submitted to traversal */
  uint64_t rays_per_bounce[16];
  uint64_t nodes_visited;  /* only when built with PTC_STATS (debug entry)     */
  uint64_t tris_tested;
} ptc_counters;

typedef struct ptc_scene_info
{
  uint32_t n_faces, n_lights, n_materials, n_textures;
  uint32_t bvh_nodes;      /* wide nodes                                       */
  uint32_t bvh_node_bytes; /* bytes per node                                   */
  uint32_t tri_bytes;      /* bytes per intersection record                    */
  uint32_t shade_bytes;    /* bytes per shading record                         */
  uint64_t texture_bytes;  /* device bytes of texels                           */
  uint32_t smem_resident;  /* 1 if the fused path kernels read nodes + triangles from a shared-memory
                              copy (scene fits beside the stacks AND the context was created with
                              PTC_SMEM_SCENE=1; off by default: measured slower on B200) */
} ptc_scene_info;

typedef struct ptc_context ptc_context;

int ptc_abi_version(void);

/* One context per GPU.  Fails with PTC_ERR_NO_DEVICE when no CUDA device is
 * usable -- there is no CPU fallback. */
int ptc_create(int device, ptc_context** out);
void ptc_destroy(ptc_context* ctx);
const char* ptc_last_error(const ptc_context* ctx);

/* Replaces Scene::upload_gpu + uploadScenes/uploadTextures
 * (scene.cpp:370-391, gpu_processor.cpp:184-238): builds the wide BVH and the
 * split triangle / shading records on the host, uploads them and the texel
 * arena.  Returns the scene id to pass to ptc_render. */
int ptc_upload_scene(ptc_context* ctx, const ptc_scene_desc* desc, int32_t* scene_id);
int ptc_scene_info_get(const ptc_context* ctx, int32_t scene_id, ptc_scene_info* out);

/* -- the material / texture table (replaces MaterialLoader::load, material_loader.cpp:153-401) --
 * Input: per MTL material its constants and its DECODED maps (stbi_loadf output, linear floats;
 * `name` = the texture name in the MTL, which is the cache key; a NULL map pointer or an empty
 * name = the material names no such map; data == NULL or nb_chan == 0 = the file failed to load).
 * Output: scene::Material per material with the reference's texture ids, and the global texture
 * table: one RGBA texture per material (rgb = diffuse, a = specular mix; packed from a 3-channel
 * and a 1-channel map, from one map and the material's constant, or a 1x1 unit texture) plus the
 * normal maps.  Ids and caches live in the packer and continue across calls exactly like the
 * reference's singleton across scenes: both maps -> packed once per name pair; one map -> packed
 * again for every material naming it; failed diffuse / specular map -> unit texture; failed
 * normal map -> -1.  Diffuse and specular maps of different sizes are PTC_ERR_INVALID (the
 * reference rescales with stb_image_resize, third-party: resize first). */
typedef struct ptc_image
{
  int32_t w, h, nb_chan;
  int32_t _pad;
  const float* data;
  const char* name;
} ptc_image;

typedef struct ptc_mtl_entry
{
  float diffuse[3];  /* Kd */
  float specular[3]; /* Ks: default specular mix = (Ks.r + Ks.g + Ks.b) / 3 (:174-176) */
  float ior;         /* Ni */
  uint32_t _pad;
  const ptc_image* diffuse_map;  /* map_Kd */
  const ptc_image* specular_map; /* map_Ks */
  const ptc_image* normal_map;   /* map_bump if given, else norm (:184-187) */
} ptc_mtl_entry;

typedef struct ptc_texpack ptc_texpack;
int ptc_texpack_create(ptc_texpack** out);
void ptc_texpack_destroy(ptc_texpack* p);
const char* ptc_texpack_last_error(const ptc_texpack* p);
int ptc_texpack_add_materials(ptc_texpack* p, const ptc_mtl_entry* mats, uint32_t n, ptc_material* out);
uint32_t ptc_texpack_texture_count(const ptc_texpack* p);
int ptc_texpack_textures(const ptc_texpack* p, ptc_texture* out, uint32_t capacity);

/* -- the step BEFORE Face[]: the mesh as the OBJ loader leaves it ------------------------------
 * Replaces the de-indexing loop of upload_meshes (scene.cpp:202-283): per shape a list of index
 * records (tinyobj::index_t: vertex, normal, texcoord index first; `index_stride_ints` ints per
 * record, 0 -> 3) and a material id per face, over shared attribute arrays (tinyobj::attrib_t).
 * Every triple becomes one scene::Face with the reference's per-face tangent (scene.cpp:250-261,
 * same expression order, so the same bits); faces of all shapes are flattened in parallel into
 * ONE array.  An index outside its attribute array (an OBJ without `vn` / `vt`: the reference
 * reads out of bounds there) is PTC_ERR_INVALID. */
typedef struct ptc_indexed_mesh
{
  const int32_t* indices;      /* n_indices records of index_stride_ints ints                 */
  uint64_t n_indices;          /* faces = n_indices / 3 (scene.cpp:222)                       */
  const int32_t* material_ids; /* one per face (tinyobj mesh_t.material_ids)                  */
} ptc_indexed_mesh;

typedef struct ptc_indexed_desc
{
  const float* vertices;  uint64_t n_vertex_floats;   /* attrib_t.vertices,  xyz per vertex  */
  const float* normals;   uint64_t n_normal_floats;   /* attrib_t.normals,   xyz per normal  */
  const float* texcoords; uint64_t n_texcoord_floats; /* attrib_t.texcoords, uv per entry    */
  const ptc_indexed_mesh* meshes;
  uint32_t n_meshes;
  uint32_t index_stride_ints;
  const ptc_material* materials; uint32_t n_materials; /* as in ptc_scene_desc               */
  const ptc_light* lights;       uint32_t n_lights;
  const ptc_texture* textures;   uint32_t n_textures;
} ptc_indexed_desc;

/* host only (no GPU): faces_out[capacity] in mesh order, mesh_face_counts[n_meshes] (optional) */
uint64_t ptc_indexed_face_count(const ptc_indexed_desc* desc);
int ptc_flatten_faces(const ptc_indexed_desc* desc, ptc_face* faces_out, uint64_t capacity,
                      uint32_t* mesh_face_counts);
/* flatten + ptc_upload_scene */
int ptc_upload_scene_indexed(ptc_context* ctx, const ptc_indexed_desc* desc, int32_t* scene_id);

/* Replaces uploadCubemap (gpu_processor.cpp:68-161) after face extraction:
 * `faces` = 6 faces (+x,-x,+y,-y,+z,-z) of size*size RGBA floats, host memory. */
int ptc_set_cubemap(ptc_context* ctx, const float* faces, uint32_t size, int32_t* cubemap_id);

/* The other half of uploadCubemap (gpu_processor.cpp:84-132, texture_utils.cpp:5-52): a
 * 4:3 "cube cross" image of w x h pixels, nb_chan (>= 3) linear floats per pixel as
 * stbi_loadf returns it, is cut into the six faces (+x = cell (2,1), -x = (0,1), +y = (1,0),
 * -y = (1,2), +z = (1,1), -z = (3,1); cells of size = w/4; alpha 0) and uploaded.  Like the
 * reference, a cross with w/4 != h/3 or a non-power-of-two size is not an error: the constant
 * colour 0x131b23 (size 1) is used instead and *used_fallback (optional) is set.
 * `cross` == NULL selects that constant-colour path with `rgb_hex` = 0xRRGGBB (0 is black, as
 * strtol of a "0x000000" cubemap line gives, gpu_processor.cpp:88-93) or
 * PTC_CUBEMAP_DEFAULT_COLOR for the reference's default 0x131b23 (a scene without cubemap). */
#define PTC_CUBEMAP_DEFAULT_COLOR 0xFFFFFFFFu
int ptc_set_cubemap_from_cross(ptc_context* ctx, const float* cross, uint32_t w, uint32_t h,
                               uint32_t nb_chan, uint32_t rgb_hex, int32_t* cubemap_id,
                               int32_t* used_fallback);

/* Host-only half of the above (no context, no GPU): writes the faces to `faces_out`
 * (6 * size * size * 4 floats; capacity in floats) and the size to *size_out.  Returns
 * PTC_ERR_INVALID when the capacity is too small. */
int ptc_cubemap_faces_from_cross(const float* cross, uint32_t w, uint32_t h, uint32_t nb_chan,
                                 uint32_t rgb_hex, float* faces_out, uint64_t capacity,
                                 uint32_t* size_out, int32_t* used_fallback);

/* -- headless output and resume (the window is gone; raytrace.cu:260-270 wrote RGBA8 to a GL
 * surface that interop.cpp:67-72 blitted upside down) ------------------------------------
 * ptc_write_ppm : binary PPM (P6) of an RGBA8 HOST frame as ptc_render writes it (row 0 = top,
 *                 alpha dropped).
 * ptc_write_png : the same frame as a PNG (8-bit RGB, or RGBA when keep_alpha != 0; Sub filter,
 *                 one fixed-Huffman deflate block, written by the library itself: no libpng / zlib).
 * ptc_accum_save/load : the float3 temporal framebuffer of a band (HOST, rows bottom-up as the
 *                 core keeps them) with the sample count it holds, so that a long accumulation
 *                 can stop and resume: load, pass the buffer as `accum` with upload_accum = 1
 *                 and continue with frame_nb = samples + 1.
 *                 File: "PTAC" u32 version(1) u32 width u32 rows u32 samples, then w*rows*3 f32. */
int ptc_write_ppm(const char* path, const void* rgba, uint32_t width, uint32_t height);
int ptc_write_png(const char* path, const void* rgba, uint32_t width, uint32_t height, int keep_alpha);
int ptc_accum_save(const char* path, const float* accum, uint32_t width, uint32_t rows, uint32_t samples);
int ptc_accum_load(const char* path, float* accum, uint64_t capacity_floats, uint32_t* width,
                   uint32_t* rows, uint32_t* samples);

/* -- taking the reference's DEVICE data as GPUProcessor::init() leaves it --------------
 * Layout mirrors of scene::Buffer<T> (scene_data.h:10-15, 16 bytes) and scene::Scenes
 * (scene_data.h:83-87, 32 bytes): host structs holding device pointers.
 *   Scenes.scenes -> device SceneData*[] -> device SceneData{meshes, materials, lights}
 *   -> device Mesh[]{Buffer<Face>} -> device Face[]          (scene.cpp:202-283, 370-391)
 * A caller that only has this tree (the reference discards the host Face vector after
 * cudaMemcpy, scene.cpp:230-281) passes it here; the records are read back once, then the
 * call does what ptc_upload_scene does.  Synchronous; caller keeps owning the tree. */
typedef struct ptc_ref_buffer
{
  uint32_t size;
  uint32_t _pad;
  const void* data; /* DEVICE pointer */
} ptc_ref_buffer;

typedef struct ptc_ref_scenes
{
  ptc_ref_buffer textures;   /* device scene::Texture[] whose .data are device floats */
  const void* const* scenes; /* device array of device scene::SceneData*              */
  uint64_t _pad;
} ptc_ref_scenes;

int ptc_upload_scene_from_device(ptc_context* ctx, const ptc_ref_scenes* scenes, uint32_t scene_index,
                                 int32_t* scene_id);

/* Same for scene::Cubemap.cubemap (scene_data.h:21-25): `cuda_array` is the reference's
 * RGBA32F cubemap cudaArray (gpu_processor.cpp:134-152); its six faces are copied (once,
 * through the host) into an array owned by the context. */
int ptc_set_cubemap_from_array(ptc_context* ctx, const void* cuda_array, int32_t* cubemap_id);

/* The render call: one sample per pixel accumulated into `accum`, tonemapped
 * RGBA8 written to `out_rgba`.  Asynchronous on args->stream, no hidden syncs. */
int ptc_render(ptc_context* ctx, const ptc_render_args* args);

/* ptc_render for callers that cannot allocate (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * Same arguments, same frame counter and the same bits in accum / out_rgba /
 * radiance_out as ptc_render, but the frame's scratch (ray queues, hit records, parked pixels,
 * counters: 176 bytes per pixel) is a file-scope device array inside the library instead of
 * cudaMalloc'ed context state, so the call allocates nothing (no cudaMalloc / cudaFree / allocating
 * library routine) and runs the frame as one pipeline on args->stream.
 *   limit:  a band of at most PTC_STATIC_MAX_PIXELS pixels (rows rendered x width; 640x360 fits);
 *           larger calls fail with PTC_ERR_INVALID and a message in ptc_last_error, and leave the
 *           frame counter untouched.  The array is 44 MiB and every GPU that loads the library
 *           holds it until exit.
 *   shared: every call on a device -- from any context -- uses the same array.  It keeps the last
 *           call's contents; the counters that must start at zero are cleared on args->stream
 *           before the first launch.  Calls that overlap on the GPU (two streams, two contexts of
 *           one device) race: serialise them.  ptc_render is unaffected (its scratch is its own). */
#define PTC_STATIC_MAX_PIXELS 262144u
int ptc_render_static(ptc_context* ctx, const ptc_render_args* args);

/* Same, but accum / out_rgba / radiance_out are HOST buffers: accum is copied
 * host->device only when `upload_accum` != 0 (first frame of a resumed
 * accumulation), rendered, and out_rgba (and accum when `download_accum`)
 * copied back; synchronises before returning. */
int ptc_render_host(ptc_context* ctx, const ptc_render_args* args, int upload_accum,
                    int download_accum);

/* Pipelined form of ptc_render_host for a caller that consumes every frame on the host (what
 * the reference's blit-to-window does each frame, interop.cpp:67-72): enqueues the frame and
 * the device->host copy of its RGBA8 (on an internal copy stream, two device frames) and
 * returns without waiting, so the copy of frame k overlaps the rendering of frame k+1.
 * out_rgba must stay valid (and should be pinned) until ptc_wait or until the second next
 * call returns a frame into it; use two host buffers alternately.  The accumulator stays on
 * the device; args->accum is read only when upload_accum != 0 or the band size changed. */
int ptc_render_host_async(ptc_context* ctx, const ptc_render_args* args, int upload_accum);

/* Waits for every pending copy of ptc_render_host_async.  accum_out != NULL additionally
 * downloads the device-side accumulator (accum_bytes = band pixels * 12). */
int ptc_wait(ptc_context* ctx, void* accum_out, uint64_t accum_bytes);

/* Resets the context's frame counter (what `moved` does, without a preview frame). */
int ptc_reset_accumulation(ptc_context* ctx);

int ptc_counters_get(ptc_context* ctx, ptc_counters* out);
int ptc_counters_reset(ptc_context* ctx);

/* Number of rows a call with these args renders (band height), or 0 on bad args. */
uint32_t ptc_band_rows(const ptc_render_args* args);

/* Kernels this context has launched so far (host-side count). */
int ptc_launch_count(const ptc_context* ctx, uint64_t* out);

/* Wavefront schedule.  first_iters / next_iters: how many radiance-loop iterations the first
 * stage (generate + extend + shade) and each later stage run in registers before handing the
 * surviving rays to the next stage through the compacted queue (0 keeps the current value).
 * split_extend: 1 = every later stage is an extend kernel with per-lane refill followed by
 * a shade kernel (one iteration per stage), 0 = fused extend+shade, 2 = shifted stages (stage k
 * shades iteration k-1 and intersects iteration k, so only rays that hit are queued; one
 * iteration per stage, first_iters / next_iters ignored), < 0 keeps.
 * refill: idle lanes per warp that trigger a refill in the extend kernel (0 keeps; default 12).
 * Default: 1 iteration per stage; later stages fused when nodes + triangles fit 64 KB (L1),
 * split otherwise;
 * first_iters = 0xFFFFFFFF restores that default for all three.
 * Results do not depend on any of it
 * (tests/test_gpu_parity.py::test_wavefront_granularity_invariance). */
int ptc_set_wavefront(ptc_context* ctx, uint32_t first_iters, uint32_t next_iters,
                      int32_t split_extend, uint32_t refill);

/* Measurement: renders the frame described by `args` `repeats` times with CUDA events
 * around every kernel launch on args->stream (as ONE pipeline, so that every launch can be
 * timed alone), once more with traversal statistics (untimed), and `repeats` times as ptc_render
 * ships it (frame_ms_shipped).  All passes render into scratch buffers owned by the call:
 * args->accum / out_rgba / radiance_out are ignored and the caller's ray counters unchanged. */
typedef struct ptc_profile
{
  uint32_t n_launches;        /* kernel launches per frame                            */
  float launch_ms[130];       /* mean duration of each launch over the repeats        */
  uint32_t launch_kind[130];  /* 0 first stage (generate+extend+shade), 1 fused later
                                 stage, 2 extend (refilling traversal), 3 shade,
                                 4 / 5 shifted form: generate+intersect / shade+intersect */
  uint32_t launch_bounce0[130];/* first radiance-loop iteration each launch handles    */
  uint32_t launch_iters[130];
  float frame_ms;             /* mean duration of the whole frame (all launches)      */
  uint64_t traced, segments;  /* per frame                                            */
  uint64_t nodes_visited, tris_tested; /* per frame, from the statistics pass          */
  uint64_t rays_per_bounce[16];
  uint64_t queue_rays;        /* rays written to (and read back from) a queue          */
  uint32_t queue_bytes_per_ray; /* bytes per queue slot (written once + read once)     */
  uint32_t hit_bytes_per_ray; /* hit record between extend and shade (0 when fused)    */
  float frame_ms_shipped;     /* mean duration of the frame as ptc_render enqueues it  */
  uint32_t pipelines;         /* half-frame pipelines ptc_render uses for this frame   */
  uint64_t extend_lanes[8];   /* extend kernel, statistics pass: [0] warp iterations of the
                                 traversal loop, and summed over them the lanes that [1] held a
                                 ray, [2] ran a node test, [3] a triangle test, [4] a second one */
} ptc_profile;
int ptc_profile_frame(ptc_context* ctx, const ptc_render_args* args, uint32_t repeats,
                      ptc_profile* out);

/* -- several GPUs behind one call (single process) ------------------------------------------
 * The reference drives one GPU from one host thread (gpu_info.cpp:26-28,38-43;
 * GPUProcessor::render, gpu_processor.cpp:365-386).  A group keeps that call shape -- one upload,
 * one render call per frame, one image -- over N devices: the scene is prepared once and
 * replicated, the image is split into round-robin 16-row strips (strip s -> devices[s mod N]),
 * every device accumulates its strips in an accumulator the group owns, and the RGBA8 pixels are
 * stored by each device's kernels straight into the full image on devices[0] through the peer
 * mapping (PTC_OUT_FULL_IMAGE) -- no gather pass, no collective.  The N-device image and
 * accumulator are bit-identical to one device's.  `devices` may name the same device more than
 * once (several contexts on one GPU; how the path is tested on a one-GPU box).
 * ptc_group_render: args as for ptc_render except accum, radiance_out, rows and strips must be 0;
 *   out_rgba = device pointer ON devices[0] of the full width x height image, or NULL for an image
 *   the group owns (ptc_group_image / ptc_group_read_image); asynchronous: in the order of
 *   args->stream (a stream of devices[0]) the frame is complete when the call's work is.
 * ptc_group_gather_accum: the full float3 accumulator, rows bottom-up as one device keeps it
 *   (raytrace.cu:252), assembled in HOST memory; synchronises. */
typedef struct ptc_group ptc_group;
int ptc_group_create(const int* devices, uint32_t n, ptc_group** out);
void ptc_group_destroy(ptc_group* g);
const char* ptc_group_last_error(const ptc_group* g);
uint32_t ptc_group_size(const ptc_group* g);
ptc_context* ptc_group_context(ptc_group* g, uint32_t i);
int ptc_group_upload_scene(ptc_group* g, const ptc_scene_desc* desc, int32_t* scene_id);
int ptc_group_set_cubemap(ptc_group* g, const float* faces, uint32_t size, int32_t* cubemap_id);
/* ptc_upload_scene_from_device / ptc_set_cubemap_from_array for a group: the caller's device tree and
 * cubemap array live on devices[0] (where the reference's GPUProcessor put them,
 * gpu_processor.cpp:184-263); read back once, prepared once, replicated to every member. */
int ptc_group_upload_scene_from_device(ptc_group* g, const ptc_ref_scenes* scenes, uint32_t scene_index,
                                       int32_t* scene_id);
int ptc_group_set_cubemap_from_array(ptc_group* g, const void* cuda_array, int32_t* cubemap_id);
int ptc_group_render(ptc_group* g, const ptc_render_args* args);
int ptc_group_sync(ptc_group* g);
const void* ptc_group_image(const ptc_group* g);
int ptc_group_read_image(ptc_group* g, void* host_rgba, void* stream);
int ptc_group_gather_accum(ptc_group* g, float* host_accum);
int ptc_group_reset_accumulation(ptc_group* g);

/* -- the same direct-store assembly for one PROCESS per GPU (bench.py under torchrun) ----------
 * ptc_shared_alloc: device memory on `device` plus its CUDA IPC handle (64 bytes) to send to the
 *   other ranks; ptc_shared_open maps it there (pass the mapped pointer as out_rgba with
 *   PTC_OUT_FULL_IMAGE); ptc_shared_close / ptc_shared_free undo them.
 * ptc_flag_signal: after everything enqueued on `stream` so far, store `value` to the (possibly
 *   peer-mapped) 32-bit flag with system-scope release semantics ("my strips of frame k are in").
 * ptc_flag_wait: `stream` waits (a one-block kernel that spins) until flags[i * stride_words]
 *   has reached `value` for every i < n (n <= 256).  Enqueue it AFTER the producers that run in
 *   the same process on the same GPU (a spinning kernel enqueued before them can sit in front of
 *   them in a shared hardware queue) and never synchronise the device while it may be spinning;
 *   a wait of more than 20 s traps (the stream then reports a launch failure) instead of hanging
 *   the GPU.
 * ptc_memcpy_to_host: device -> host copy on `stream` (sync != 0: waits for it). */
int ptc_shared_alloc(int device, uint64_t bytes, void** ptr, uint8_t handle[64]);
int ptc_shared_open(int device, const uint8_t handle[64], void** ptr);
int ptc_shared_close(int device, void* ptr);
int ptc_shared_free(int device, void* ptr);
int ptc_flag_signal(int device, void* flag, uint32_t value, void* stream);
int ptc_flag_wait(int device, const void* flags, uint32_t n, uint32_t stride_words, uint32_t value, void* stream);
int ptc_memcpy_to_host(int device, void* host, const void* dev, uint64_t bytes, void* stream, int sync);

/* -- one process per GPU, image consumed on the HOST ---------------------------------------------
 * The reference's consumer of a frame is the host side of the GL interop (interop.cpp:67-72).  With
 * one process per GPU, the assembled image does not have to cross one GPU's PCIe link: every rank
 * copies ITS strips straight into one host image all processes map.
 * ptc_host_image_open: POSIX shared memory `name` ("/..." as for shm_open) of `bytes`, created
 *   (create != 0, one rank) or opened (the others, after the creator), mapped and page-locked for
 *   this process's CUDA context; ptc_host_image_close unmaps it (unlink_name != NULL also removes it).
 * ptc_memcpy_strips_to_host: device band of this rank (band-local rows = its strips one after the
 *   other, what ptc_render writes for strip_rows / strip_stride / strip_phase) -> the rows those
 *   strips have in the full width x height RGBA8 host image; asynchronous on `stream` (at most two
 *   DMA transfers: a pitched one for the whole strips, one for a short last strip). */
int ptc_host_image_open(const char* name, uint64_t bytes, int create, void** ptr);
int ptc_host_image_close(void* ptr, uint64_t bytes, const char* unlink_name);
int ptc_memcpy_strips_to_host(int device, void* host_image, const void* dev_band, uint32_t width, uint32_t height,
                              uint32_t strip_rows, uint32_t strip_stride, uint32_t strip_phase, void* stream);
/* Same scatter with a DEVICE image as the destination -- this device's or, mapped with
 * ptc_shared_open, another GPU's (DMA over NVLink): the copy-engine alternative to storing the pixels
 * from the kernels (PTC_OUT_FULL_IMAGE).  Measured on 8 B200: seven GPUs storing their 32-byte pixel
 * runs into one GPU during the short last stage of a frame cost 0.16 ms per 0.62 ms frame; seven
 * pitched DMA copies overlapped with the next frame cost nothing visible. */
int ptc_memcpy_strips_to_device(int device, void* image, const void* dev_band, uint32_t width, uint32_t height,
                                uint32_t strip_rows, uint32_t strip_stride, uint32_t strip_phase, void* stream);

/* -- parity / debugging entry points (used by tests, not by a renderer) -------- */

/* Primary rays of sample `frame_nb` (RNG + camera + DOF as raytrace.cu:227-240),
 * closest hit per pixel.  hit_ids[h*w] int32 (row 0 = top), t[h*w] float; HOST
 * buffers; either may be NULL. */
int ptc_trace_primary(ptc_context* ctx, int32_t scene_id, const ptc_camera* camera,
                      uint32_t width, uint32_t height, uint32_t frame_nb,
                      int32_t* hit_ids, float* t);

/* Closest hit for n arbitrary rays (HOST arrays, origins/dirs as xyz triples). */
int ptc_intersect_rays(ptc_context* ctx, int32_t scene_id, const float* origins,
                       const float* dirs, uint64_t n, int32_t* hit_ids, float* t);

/* First `n_draws` curand_uniform draws of curand_init(seed,0,0) for each of the
 * n_seeds seeds, computed by the core's own XORWOW and (reference_impl != 0) by
 * the toolkit's curand_kernel.h on the device. out[n_seeds*n_draws] HOST. */
int ptc_debug_rng(ptc_context* ctx, const uint32_t* seeds, uint32_t n_seeds,
                  uint32_t n_draws, int reference_impl, float* out);

/* Measured read bandwidth (GB/s) of the traversal's load path (16-byte read-only loads, all
 * SMs) over a zeroed working set of `working_set_bytes` (rounded down to a power of two),
 * reading `total_bytes` in all: a working set that fits L1 (<= 64 KB), L2 (e.g. 32 MB) or
 * neither gives the denominator of that tier for the roofline in bench.py. */
int ptc_debug_read_bandwidth(ptc_context* ctx, uint64_t working_set_bytes, uint64_t total_bytes,
                             float* gb_per_s);

/* -- host-only helpers (no GPU needed) ----------------------------------------- */

/* Builds the wide BVH for a scene on the host and returns its size; used by the
 * oracle (O2 traverses the same nodes) and by CPU tests.  Buffers are owned by
 * the library until ptc_host_bvh_free. */
typedef struct ptc_host_bvh
{
  const void* nodes;       /* bvh_nodes * node_bytes                              */
  uint32_t n_nodes;
  uint32_t node_bytes;
  const void* tris;        /* n_tris * tri_bytes, leaf order                      */
  uint32_t n_tris;
  uint32_t tri_bytes;
  const uint32_t* tri_face_ids; /* leaf order -> global face index              */
  float sah_cost;
  uint32_t max_depth;
  void* _owner;
} ptc_host_bvh;

int ptc_host_bvh_build(const ptc_scene_desc* desc, ptc_host_bvh* out);
void ptc_host_bvh_free(ptc_host_bvh* bvh);

#ifdef __cplusplus
}
#endif
#endif /* PTCORE_H */
