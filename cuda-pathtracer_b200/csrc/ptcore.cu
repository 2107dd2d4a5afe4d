// ptcore.cu -- implementation of the C ABI in include/ptcore.h.
//
// Replaces, behind the reference's render boundary (raytrace.h:9-15), the upload
// path (scene.cpp:370-391, gpu_processor.cpp:68-238) and the per-frame launch
// (raytrace.cu:287-325).  One context per GPU; every kernel of a frame is enqueued
// on the caller's stream with no hidden synchronisation.  There is no CPU fallback:
// without a usable CUDA device ptc_create fails with PTC_ERR_NO_DEVICE.
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <mutex>
#include <string>
#include <vector>

#include "bvh.h"
#include "kernels.cuh"
#include "ptcore.h"
#include "ptcore_internal.h"

using namespace ptc;


namespace {

struct SceneDev
{
  DScene d{};
  std::vector<void*> allocs;
  ptc_scene_info info{};
  // traversal stack entries per thread (levels of the wide tree + 1 + kTriBacklog) and the
  // persistent grids that follow from it: resident CTAs per SM x SMs, [0] plain / [1] statistics
  int stack_cap = 0;
  int grid_first[2] = { 0, 0 }, grid_next[2] = { 0, 0 }, grid_shade[2] = { 0, 0 }, grid_extend[2] = { 0, 0 };
  int grid_wave_first[2] = { 0, 0 }, grid_wave_next[2] = { 0, 0 }; // shifted schedule (k_wave)
  // fused path kernels reading the BVH from a per-CTA shared-memory copy (kernels.cuh): stack
  // entries and grids of that form; smem_scene = it keeps the path kernels' resident CTA count
  int stack_cap_s = 0;
  int grid_first_s[2] = { 0, 0 }, grid_next_s[2] = { 0, 0 };
  bool smem_scene = false;
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
  RayQueue queue[2]{};         // double-buffered compacted ray queues
  uint32_t capacity = 0;       // queue slots allocated
  uint32_t* counts = nullptr;  // one queue counter per wavefront step
  Counters* ctr = nullptr;
  int n_sms = 148;
  // persistent grids, [stats]: first stage, fused later stage, shade-after-extend, extend
  uint32_t first_iters = 0, next_iters = 1; // first_iters 0 = per scene (frame_schedule) // wavefront granularity (ptc_set_wavefront)
  int32_t split_extend = -1;   // form of the stages: 0 fused, 1 refilling extend kernel + shade kernel, 2 shifted (k_wave), -1 = per scene
  int wave_small = 0;          // automatic choice for L1-resident scenes: 1 shifted, 0 fused; PTC_WAVE overrides (experiments)
  uint32_t refill = 12;        // idle lanes of a warp that trigger a refill in k_extend (kDefaultRefill)
  uint4* hits = nullptr;       // closest triangle hit per queue slot (split stages; shifted stages: even queue)
  uint4* hits2 = nullptr;      // shifted stages: hit records of the odd queue
  float4* done = nullptr;      // (radiance, pixel) of paths that finished before the last stage
  int defer = 1;               // epilogue of those pixels runs in the last stage; PTC_DEFER=0: in place
  // scenes that fit: BVH read from a per-CTA shared-memory copy in the fused path kernels.  Measured on
  // B200 and NOT the default: +5 % frame time (indoor 1080p), +3 % (4K, 8 iterations) -- the LSU
  // wavefronts drop (56 -> 46 % of peak) but the L1 that is left (28 KB beside 4 x 50 KB of shared
  // memory) no longer holds shading records and texels (L1 hit rate 63 -> 25 %).  PTC_SMEM_SCENE=1.
  int smem_scene = 0;
  // a frame runs as TWO half-band pipelines on two streams, so that the tail of every stage of
  // one half (SMs draining while the last warps finish) is filled by the other half's stage
  int dual = 1;                // extra pipelines (0..3); PTC_DUAL overrides (experiments)
  bool dual_forced = false;    // PTC_DUAL was given: no per-frame choice (frame_pipelines)
  cudaStream_t side_stream[3] = { nullptr, nullptr, nullptr };
  cudaEvent_t ev_fork = nullptr, ev_join[3] = { nullptr, nullptr, nullptr };
  uint64_t launches = 0;
  uint32_t seed = 0;           // the reference's function-static frame counter (raytrace.cu:296)
  // staging for ptc_render_host
  float* h_accum_dev = nullptr;
  uint32_t* h_out_dev = nullptr;
  float* h_rad_dev = nullptr;
  size_t h_px = 0;
  // whose samples the staged device accumulator holds: host buffer, scene and band geometry of the
  // last host-buffer call.  A call for anything else re-seeds it from the caller's buffer.
  struct StagingOwner
  {
    const void* accum = nullptr;
    int32_t scene_id = -1;
    uint32_t geom[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };
    bool operator==(const StagingOwner& o) const
    {
      return accum == o.accum && scene_id == o.scene_id && !memcmp(geom, o.geom, sizeof(geom));
    }
  } h_owner;
  // pipelined host call: second device frame + copy stream so the device->host copy of
  // frame k overlaps the rendering of frame k+1
  uint32_t* h_out_dev2 = nullptr;
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t ev_render[2] = { nullptr, nullptr }, ev_copy[2] = { nullptr, nullptr };
  int async_slot = 0;
  std::string err;
};

namespace {

constexpr int kMaxBounces = 64;
constexpr uint32_t kDefaultRefill = 12;
// queue counters, then 2 work counters per step; the last two slots: parked-pixel count and the
// work counter of their drain (kernels.cuh, kDoneDefer / kDoneDrain)
constexpr int kCountSlots = 3 * (kMaxBounces + 2);

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

// Temporary device buffer of the parity / measurement entry points: freed on every return path.
struct DevBuf
{
  void* p = nullptr;
  ~DevBuf() { cudaFree(p); }
  cudaError_t alloc(size_t bytes) { return cudaMalloc(&p, bytes ? bytes : 16); }
  template <typename T>
  T* as() const
  {
    return (T*)p;
  }
};

struct EventList
{
  std::vector<cudaEvent_t> ev;
  ~EventList()
  {
    for (cudaEvent_t e : ev)
      if (e)
        cudaEventDestroy(e);
  }
  cudaError_t create(size_t n)
  {
    ev.assign(n, nullptr);
    for (cudaEvent_t& e : ev) {
      cudaError_t rc = cudaEventCreate(&e);
      if (rc != cudaSuccess)
        return rc;
    }
    return cudaSuccess;
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
  for (RayQueue& q : ctx->queue) {
    cudaFree(q.o);
    cudaFree(q.d);
    cudaFree(q.thr);
    cudaFree(q.acc);
    q = RayQueue{};
  }
  cudaFree(ctx->hits);
  cudaFree(ctx->hits2);
  cudaFree(ctx->done);
  ctx->hits = nullptr;
  ctx->hits2 = nullptr;
  ctx->done = nullptr;
  ctx->capacity = 0;
}

int
ensure_capacity(ptc_context* ctx, uint32_t n)
{
  if (n <= ctx->capacity)
    return 0;
  free_state(ctx);
  for (RayQueue& q : ctx->queue) {
    CUDA_TRY(ctx, cudaMalloc(&q.o, (size_t)n * sizeof(float4)));
    CUDA_TRY(ctx, cudaMalloc(&q.d, (size_t)n * sizeof(float4)));
    CUDA_TRY(ctx, cudaMalloc(&q.thr, (size_t)n * sizeof(float4)));
    CUDA_TRY(ctx, cudaMalloc(&q.acc, (size_t)n * sizeof(float4)));
  }
  CUDA_TRY(ctx, cudaMalloc(&ctx->hits, (size_t)n * sizeof(uint4)));
  CUDA_TRY(ctx, cudaMalloc(&ctx->hits2, (size_t)n * sizeof(uint4)));
  CUDA_TRY(ctx, cudaMalloc(&ctx->done, (size_t)n * sizeof(float4)));
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

// Rows this call renders: contiguous band or interleaved strips (ptcore.h).
bool
band_geometry(const ptc_render_args* a, FrameParams& fp)
{
  fp.width = a->width;
  fp.height = a->height;
  fp.tiles_x = (a->width + kTileW - 1) / kTileW;
  fp.strip_rows = a->strip_rows ? a->strip_rows : 16u;
  fp.strip_stride = a->strip_stride > 1 ? a->strip_stride : 1u;
  fp.strip_phase = a->strip_phase;
  if (fp.strip_stride == 1) {
    fp.strip_phase = 0;
    uint32_t r0 = a->row_begin, r1 = a->row_end;
    if (r0 == 0 && r1 == 0)
      r1 = a->height;
    if (r0 >= r1 || r1 > a->height)
      return false;
    fp.row_begin = r0;
    fp.band_rows = r1 - r0;
    return true;
  }
  if (a->row_begin || a->row_end || fp.strip_phase >= fp.strip_stride)
    return false;
  fp.row_begin = 0;
  const uint32_t n_strips = (a->height + fp.strip_rows - 1) / fp.strip_rows;
  uint32_t rows = 0;
  for (uint32_t s = fp.strip_phase; s < n_strips; s += fp.strip_stride) {
    const uint32_t b = s * fp.strip_rows, e = b + fp.strip_rows < a->height ? b + fp.strip_rows : a->height;
    rows += e - b;
  }
  fp.band_rows = rows;
  return rows > 0;
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
  if (!band_geometry(a, fp))
    return fail(ctx, PTC_ERR_INVALID, "bad row range / strip arguments");
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

size_t
scene_bvh_bytes(const ptc_context* ctx, int32_t scene_id)
{
  const ptc_scene_info& i = ctx->scenes[scene_id].info;
  return (size_t)i.bvh_nodes * i.bvh_node_bytes + (size_t)i.n_faces * i.tri_bytes;
}

// One launch of the frame's schedule.
struct Step
{
  bool first;
  int bounce0, n_iters;
};

// The wavefront schedule of a frame: the first stage generates the rays and runs
// `first_iters` radiance-loop iterations in registers; every later stage runs `next_iters`.
int
frame_schedule(const ptc_context* ctx, const FrameParams& fp, size_t bvh_bytes, Step* steps, bool* split_out,
               bool* wave_out = nullptr)
{
  int n = 0;
  // automatic (measured on B200 with the path kernels at 4 CTAs per SM and two half-frame
  // pipelines, tools/sweep2.py): ONE radiance-loop iteration per LATER stage everywhere (the first
  // stage: see below), and
  //  * nodes + triangles <= 64 KB (L1-resident; the reference's scenes are 10-30 KB): every
  //    later stage traverses AND shades in one launch -- 0.638 ms against 0.671 ms for the split
  //    form on the 1080p indoor frame;
  //  * larger: the refilling extend kernel + a shade kernel per iteration -- already 1.5 % ahead
  //    at 70 KB (1 164 triangles), 9 % at 1.2 MB, 2.59 ms against 3.09 ms fused on 10 M triangles
  //    (tools/sweep_split.py).
  const bool small = bvh_bytes <= (64u << 10);
  const bool split = ctx->split_extend >= 0 ? ctx->split_extend == 1 : !small;
  // shifted form (k_wave, kernels.cu): stage k intersects iteration k and shades iteration k-1, so
  // that only rays that hit are queued and shaded.  Selectable (split_extend = 2, or PTC_WAVE=1 for
  // the L1-resident scenes), NOT a default: it measured 0.2-4 % behind the fused form
  const bool wave = fp.is_static && (ctx->split_extend >= 0 ? ctx->split_extend == 2 : small && ctx->wave_small);
  if (split_out)
    *split_out = split;
  if (wave_out)
    *wave_out = wave;
  if (!fp.is_static) {
    steps[n++] = Step{ true, 0, 1 };
    return n;
  }
  if (wave) {
    for (int b = 0; b < fp.max_bounces; ++b)
      steps[n++] = Step{ b == 0, b, 1 };
    return n;
  }
  int first = (int)ctx->first_iters;
  if (first == 0) {
    // measured per frame size and iteration count on B200 (tools/sweep_first.py,
    // profiles/r2_sweep_first.jsonl; three reference scenes, 640x360 .. 3840x2160, 3 / 4 / 8 iterations):
    // in the fused form TWO iterations in the first stage beat one at every size (-6 % at 1080p, 3
    // iterations; -4 % at 4K, 8 iterations: a path that ends at its second segment -- 44 % of them on
    // the indoor view -- never enters a queue, and a frame is one launch shorter), and a frame small
    // enough to run as one pipeline (< 256 K pixels, 1.5 waves of warps) likes three (-25 % at 640x360);
    // in the split form (scenes beyond 64 KB) one stays ahead (+4 % / +8 % with two on 180 k / 10 M triangles)
    first = split ? 1 : ((uint64_t)fp.band_rows * fp.width < (1u << 18) ? 3 : 2);
  }
  int b = first < fp.max_bounces ? first : fp.max_bounces;
  steps[n++] = Step{ true, 0, b };
  while (b < fp.max_bounces) {
    int k = (int)ctx->next_iters < fp.max_bounces - b ? (int)ctx->next_iters : fp.max_bounces - b;
    if (split)
      k = 1;
    steps[n++] = Step{ false, b, k };
    b += k;
  }
  return n;
}

// Queues, hit records and counters one pipeline of launches works on.
struct Lane
{
  RayQueue q[2];
  uint4* hits;
  uint4* hits2;
  float4* done;
  uint32_t* counts;
};

// Enqueues the launches that render band-local rows [fp.part_row0, +fp.part_rows) of a frame.
// `ev` (optional) receives one event before the first launch and one after every launch.
int
enqueue_part(ptc_context* ctx, const ptc_render_args* a, const FrameParams& fp, const FrameOut& fo,
             cudaStream_t s, const Lane& L, bool stats, cudaEvent_t* ev)
{
  const SceneDev& sc = ctx->scenes[a->scene_id];
  const CubeDev& cm = ctx->cubemaps[a->cubemap_id];
  Step steps[kMaxBounces + 2];
  bool split = false, wave = false;
  const int n_steps = frame_schedule(ctx, fp, scene_bvh_bytes(ctx, a->scene_id), steps, &split, &wave);
  CUDA_TRY(ctx, cudaMemsetAsync(L.counts, 0, sizeof(uint32_t) * kCountSlots, s));
  int n_ev = 0;
  if (ev)
    CUDA_TRY(ctx, cudaEventRecord(ev[0], s));
  const uint32_t tiles = fp.tiles_x * ((fp.part_rows + kTileH - 1) / kTileH);
  for (int k = 0; k < n_steps; ++k) {
    const Step& st = steps[k];
    uint32_t grid;
    const uint4* hits = nullptr;
    uint32_t* work = L.counts + kMaxBounces + 2 + 2 * k;
    if (wave) {
      const uint32_t want = (tiles + kBlock / 32 - 1) / (kBlock / 32);
      const uint32_t cap = (uint32_t)sc.grid_wave_first[stats];
      grid = st.first ? (want < cap ? want : cap) : (uint32_t)sc.grid_wave_next[stats];
      const int done_mode = ctx->defer && n_steps > 1 ? (k + 1 < n_steps ? kDoneDefer : kDoneDrain) : 0;
      uint4* const hb[2] = { L.hits, L.hits2 };
      launch_wave(st.first, stats, grid, fp, sc.d, cm.tex, L.q[k & 1], hb[k & 1], L.counts + k, L.q[(k + 1) & 1],
                  hb[(k + 1) & 1], L.counts + k + 1, fo, ctx->ctr, work, st.bounce0, sc.stack_cap, L.done,
                  L.counts + kCountSlots - 2, done_mode, s);
      ctx->launches++;
      if (ev)
        CUDA_TRY(ctx, cudaEventRecord(ev[++n_ev], s));
      continue;
    }
    const bool sms = sc.smem_scene && ctx->smem_scene && !split;
    if (st.first) {
      const uint32_t want = (tiles + kBlock / 32 - 1) / (kBlock / 32);
      const uint32_t cap = (uint32_t)(sms ? sc.grid_first_s[stats] : sc.grid_first[stats]);
      grid = want < cap ? want : cap;
    } else if (split) {
      launch_extend(stats, (uint32_t)sc.grid_extend[stats], sc.d, L.q[k & 1], L.counts + k, L.hits, work + 1,
                    ctx->ctr, ctx->refill, st.bounce0 + 1 >= fp.max_bounces, sc.stack_cap, s);
      ctx->launches++;
      if (ev)
        CUDA_TRY(ctx, cudaEventRecord(ev[++n_ev], s));
      hits = L.hits;
      grid = (uint32_t)sc.grid_shade[stats];
    } else {
      grid = (uint32_t)(sms ? sc.grid_next_s[stats] : sc.grid_next[stats]);
    }
    // every pixel finishes exactly once per frame; in a stage that is not the last, finished pixels
    // are parked and the last stage runs their epilogue (kernels.cuh)
    int done_mode = 0;
    if (ctx->defer && n_steps > 1)
      done_mode = k + 1 < n_steps ? kDoneDefer : kDoneDrain;
    launch_path(st.first, stats, grid, fp, sc.d, cm.tex, L.q[k & 1], hits, L.counts + k, L.q[(k + 1) & 1],
                L.counts + k + 1, fo, ctx->ctr, work, st.bounce0, st.n_iters, sms ? sc.stack_cap_s : sc.stack_cap,
                L.done, L.counts + kCountSlots - 2, done_mode, s, sms);
    ctx->launches++;
    if (ev)
      CUDA_TRY(ctx, cudaEventRecord(ev[++n_ev], s));
  }
  CUDA_TRY(ctx, cudaGetLastError());
  return 0;
}

// Pipelines a frame of this size is rendered as by ptc_render (see enqueue_frame).
uint32_t
frame_pipelines(const ptc_context* ctx, const FrameParams& fp, size_t bvh_bytes)
{
  const uint32_t n = fp.band_rows * fp.width;
  uint32_t parts = ctx->dual && fp.is_static && n >= (1u << 18) ? (uint32_t)ctx->dual + 1u : 1u;
  // Overlapping the stage tails of two half frames pays when a frame has tails to overlap.  A frame
  // of one or two launches (the reference's three iterations in the fused form) is 0.5-2 % faster as
  // ONE pipeline at every size up to 2720x1530 and equal at 4K (profiles/r2_sweep_dual.jsonl); eight
  // iterations (-1.5 %) and the split form (10 M triangles: -6 %) keep two.  PTC_DUAL set: as asked.
  if (parts > 1 && !ctx->dual_forced) {
    Step steps[kMaxBounces + 2];
    bool split = false;
    if (frame_schedule(ctx, fp, bvh_bytes, steps, &split) <= 2 && !split)
      parts = 1;
  }
  const uint32_t strips = (fp.band_rows + 15u) / 16u;
  return parts > strips ? strips : parts;
}

// Enqueues one frame; accum/out/rad are device pointers.  Profiling calls (stats / ev) run
// the frame as one pipeline; otherwise a frame of at least 256 K pixels runs as two halves
// (whole 16-row strips) on the caller's stream and an internal one, joined before returning
// to the caller's stream order.
int
enqueue_frame(ptc_context* ctx, const ptc_render_args* a, const FrameParams& fp_in, float* accum,
              uint32_t* out, float* rad, cudaStream_t s, bool stats = false, cudaEvent_t* ev = nullptr)
{
  FrameParams fp = fp_in;
  const uint32_t n = fp.band_rows * fp.width;
  int rc = ensure_capacity(ctx, n);
  if (rc)
    return rc;
  const FrameOut fo{ accum, out, rad, (a->flags & PTC_OUT_FULL_IMAGE) ? 1u : 0u };
  Lane A{ { ctx->queue[0], ctx->queue[1] }, ctx->hits, ctx->hits2, ctx->done, ctx->counts };
  fp.part_row0 = 0;
  fp.part_rows = fp.band_rows;
  const uint32_t strips = (fp.band_rows + 15u) / 16u;
  const int parts = stats || ev ? 1 : (int)frame_pipelines(ctx, fp, scene_bvh_bytes(ctx, a->scene_id));
  if (parts <= 1)
    return enqueue_part(ctx, a, fp, fo, s, A, stats, ev);

  CUDA_TRY(ctx, cudaEventRecord(ctx->ev_fork, s));
  uint32_t row0 = 0;
  for (int i = 0; i < parts; ++i) {
    const uint32_t s0 = (uint32_t)((uint64_t)strips * i / parts), s1 = (uint32_t)((uint64_t)strips * (i + 1) / parts);
    const uint32_t rows = (i == parts - 1 ? fp.band_rows : s1 * 16u) - s0 * 16u;
    const size_t off = (size_t)row0 * fp.width; // queue slots of the parts before this one
    Lane L = A;
    for (RayQueue& q : L.q)
      q.o += off, q.d += off, q.thr += off, q.acc += off;
    L.hits += off;
    L.hits2 += off;
    L.done += off;
    L.counts += (size_t)kCountSlots * i;
    FrameParams f = fp;
    f.part_row0 = row0;
    f.part_rows = rows;
    cudaStream_t st = i == 0 ? s : ctx->side_stream[i - 1];
    if (i)
      CUDA_TRY(ctx, cudaStreamWaitEvent(st, ctx->ev_fork, 0));
    if ((rc = enqueue_part(ctx, a, f, fo, st, L, false, nullptr)))
      return rc;
    if (i) {
      CUDA_TRY(ctx, cudaEventRecord(ctx->ev_join[i - 1], st));
      CUDA_TRY(ctx, cudaStreamWaitEvent(s, ctx->ev_join[i - 1], 0));
    }
    row0 += rows;
  }
  return 0;
}

// Device-side staging of the host-buffer calls (accumulator, two RGBA8 frames, radiance).
// `fresh` is set when the staged accumulator does not hold the samples of THIS caller buffer /
// scene / band (first call, new size, or a different owner with the same pixel count -- e.g. rows
// 0-540 then 540-1080): the caller's accumulator must then be uploaded first.
int
ensure_staging(ptc_context* ctx, const ptc_render_args* a, const FrameParams& fp, size_t n, bool& fresh)
{
  ptc_context::StagingOwner who;
  who.accum = a->accum;
  who.scene_id = a->scene_id;
  const uint32_t g[8] = { fp.width, fp.height, fp.row_begin, fp.band_rows, fp.strip_rows, fp.strip_stride,
                          fp.strip_phase, 0u };
  memcpy(who.geom, g, sizeof(g));
  fresh = !(who == ctx->h_owner);
  ctx->h_owner = who;
  if (n == ctx->h_px)
    return 0;
  cudaFree(ctx->h_accum_dev);
  cudaFree(ctx->h_out_dev);
  cudaFree(ctx->h_out_dev2);
  cudaFree(ctx->h_rad_dev);
  ctx->h_accum_dev = nullptr, ctx->h_out_dev = nullptr, ctx->h_out_dev2 = nullptr, ctx->h_rad_dev = nullptr;
  ctx->h_px = 0;
  CUDA_TRY(ctx, cudaMalloc(&ctx->h_accum_dev, n * 12));
  CUDA_TRY(ctx, cudaMalloc(&ctx->h_out_dev, n * 4));
  CUDA_TRY(ctx, cudaMalloc(&ctx->h_out_dev2, n * 4));
  CUDA_TRY(ctx, cudaMalloc(&ctx->h_rad_dev, n * 12));
  ctx->h_px = n;
  fresh = true;
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
      cudaMalloc(&ctx->counts, sizeof(uint32_t) * kCountSlots * 4) != cudaSuccess) {
    delete ctx;
    return PTC_ERR_ALLOC;
  }
  cudaMemset(ctx->ctr, 0, sizeof(Counters));
  cudaMemset(ctx->counts, 0, sizeof(uint32_t) * kCountSlots * 4);
  cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming);
  for (int i = 0; i < 3; ++i) {
    cudaStreamCreateWithFlags(&ctx->side_stream[i], cudaStreamNonBlocking);
    cudaEventCreateWithFlags(&ctx->ev_join[i], cudaEventDisableTiming);
  }
  if (const char* e = getenv("PTC_WAVE"))
    ctx->wave_small = atoi(e) != 0;
  if (const char* e = getenv("PTC_DEFER"))
    ctx->defer = atoi(e) != 0;
  if (const char* e = getenv("PTC_SMEM_SCENE"))
    ctx->smem_scene = atoi(e) != 0;
  if (const char* e = getenv("PTC_DUAL")) {
    ctx->dual = atoi(e);
    ctx->dual = ctx->dual < 0 ? 0 : (ctx->dual > 3 ? 3 : ctx->dual);
    ctx->dual_forced = true;
  }
  cudaDeviceGetAttribute(&ctx->n_sms, cudaDevAttrMultiProcessorCount, device);
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
  for (int i = 0; i < 3; ++i) {
    if (ctx->side_stream[i])
      cudaStreamDestroy(ctx->side_stream[i]);
    if (ctx->ev_join[i])
      cudaEventDestroy(ctx->ev_join[i]);
  }
  if (ctx->ev_fork)
    cudaEventDestroy(ctx->ev_fork);
  cudaFree(ctx->h_accum_dev);
  cudaFree(ctx->h_out_dev);
  cudaFree(ctx->h_rad_dev);
  cudaFree(ctx->h_out_dev2);
  for (int i = 0; i < 2; ++i) {
    if (ctx->ev_render[i])
      cudaEventDestroy(ctx->ev_render[i]);
    if (ctx->ev_copy[i])
      cudaEventDestroy(ctx->ev_copy[i]);
  }
  if (ctx->copy_stream)
    cudaStreamDestroy(ctx->copy_stream);
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
  PreparedScene ps;
  std::string err;
  int rc = prepare_scene(desc, ps, err);
  if (rc)
    return fail(ctx, rc, err);
  return upload_prepared(ctx, desc, ps, scene_id);
}

} // extern "C"

namespace ptc {

int
context_device(const ptc_context* ctx)
{
  return ctx->device;
}

int
prepare_scene(const ptc_scene_desc* desc, PreparedScene& ps, std::string& err)
{
  uint32_t n_faces = 0;
  for (uint32_t m = 0; m < desc->n_meshes; ++m)
    n_faces += desc->mesh_face_counts[m];
  // shading records first: the (threaded) pass that splits them off also finds the largest
  // material id, which replaces a serial scan over 1.12 GB of Face records at 10 M triangles
  if (n_faces && build_shade_records(desc, ps.shade) >= desc->n_materials) {
    err = "face references a material outside the table";
    return PTC_ERR_INVALID;
  }
  if (!n_faces)
    ps.shade.clear();
  for (uint32_t i = 0; i < desc->n_materials; ++i) {
    const ptc_material& m = desc->materials[i];
    if (m.diffuse_spec_map < 0 || (uint32_t)m.diffuse_spec_map >= desc->n_textures ||
        m.normal_map >= (int32_t)desc->n_textures) {
      err = "material references a texture outside the table";
      return PTC_ERR_INVALID;
    }
    if (desc->textures[m.diffuse_spec_map].nb_chan != 4) {
      err = "diffuse/specular texture must have 4 channels";
      return PTC_ERR_INVALID;
    }
    if (m.normal_map >= 0 && desc->textures[m.normal_map].nb_chan != 3) {
      err = "normal map must have 3 channels";
      return PTC_ERR_INVALID;
    }
  }
  ps.n_faces = n_faces;
  build_bvh(desc, ps.bvh);
  if ((int)ps.bvh.stats.wide_depth + 1 > kStackSize - kTriBacklog) {
    err = "BVH too deep for the traversal stack";
    return PTC_ERR_INVALID;
  }
  return PTC_OK;
}

int
upload_prepared(ptc_context* ctx, const ptc_scene_desc* desc, const PreparedScene& ps, int32_t* scene_id)
{
  DeviceGuard g(ctx->device);
  const HostBvh& bvh = ps.bvh;
  const RawVector<ShadeRecord>& shade = ps.shade;
  const uint32_t n_faces = ps.n_faces;

  // built into a local record and published only when everything succeeded: a failed upload
  // leaves no half-initialised scene behind a valid scene id (and frees what it allocated)
  SceneDev s;
  struct Rollback
  {
    SceneDev& s;
    bool armed = true;
    ~Rollback()
    {
      if (armed)
        for (void* p : s.allocs)
          cudaFree(p);
    }
  } rollback{ s };
  int rc;
  const Node8* d_nodes;
  const TriRecord* d_tris;
  const ShadeRecord* d_shade;
  if ((rc = dev_upload(ctx, s, bvh.nodes.data(), bvh.nodes.size(), &d_nodes)) ||
      (rc = dev_upload(ctx, s, bvh.tris.data(), bvh.tris.size(), &d_tris)) ||
      (rc = dev_upload(ctx, s, shade.data(), shade.size(), &d_shade)) ||
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
  // one merged entry per material: texel pointers + sizes + ior (device_types.cuh DMaterial)
  std::vector<DMaterial> dm(desc->n_materials);
  for (uint32_t i = 0; i < desc->n_materials; ++i) {
    const ptc_material& m = desc->materials[i];
    memset(&dm[i], 0, sizeof(DMaterial));
    dm[i].diffuse = dt[m.diffuse_spec_map].data;
    dm[i].dw = dt[m.diffuse_spec_map].w;
    dm[i].dh = dt[m.diffuse_spec_map].h;
    if (m.normal_map >= 0) {
      dm[i].normal = dt[m.normal_map].data;
      dm[i].nw = dt[m.normal_map].w;
      dm[i].nh = dt[m.normal_map].h;
    }
    dm[i].ior = m.ior;
  }
  if ((rc = dev_upload(ctx, s, dm.data(), dm.size(), &s.d.mats)))
    return rc;

  // persistent grids: resident CTAs per SM x SMs (148 on B200), for this scene's stack size
  s.stack_cap = (int)bvh.stats.wide_depth + 1 + kTriBacklog;
  for (int st = 0; st < 2; ++st) {
    s.grid_first[st] = ctx->n_sms * path_kernel_blocks_per_sm(true, st != 0, false, s.stack_cap);
    s.grid_next[st] = ctx->n_sms * path_kernel_blocks_per_sm(false, st != 0, false, s.stack_cap);
    s.grid_shade[st] = ctx->n_sms * path_kernel_blocks_per_sm(false, st != 0, true, s.stack_cap);
    s.grid_extend[st] = ctx->n_sms * extend_kernel_blocks_per_sm(st != 0, s.stack_cap);
    s.grid_wave_first[st] = ctx->n_sms * wave_kernel_blocks_per_sm(true, st != 0, s.stack_cap);
    s.grid_wave_next[st] = ctx->n_sms * wave_kernel_blocks_per_sm(false, st != 0, s.stack_cap);
  }
  if (!s.grid_first[0] || !s.grid_next[0] || !s.grid_extend[0] || !s.grid_first[1] || !s.grid_extend[1])
    return fail(ctx, PTC_ERR_INVALID, "traversal stack does not fit in shared memory");
  // shared-memory copy of nodes + triangle records for the fused path kernels: only when as many
  // CTAs stay resident as without it
  s.stack_cap_s = (int)bvh.stats.wide_depth + 1 + kSmemSceneBacklog;
  for (int st = 0; st < 2; ++st) {
    s.grid_first_s[st] = ctx->n_sms * path_kernel_blocks_per_sm_scene(true, st != 0, s.stack_cap_s, s.d);
    s.grid_next_s[st] = ctx->n_sms * path_kernel_blocks_per_sm_scene(false, st != 0, s.stack_cap_s, s.d);
  }
  s.smem_scene = s.grid_first_s[0] && s.grid_first_s[0] >= s.grid_first[0] && s.grid_next_s[0] >= s.grid_next[0] &&
                 s.grid_first_s[1] && s.grid_next_s[1];

  s.info.n_faces = n_faces;
  s.info.n_lights = desc->n_lights;
  s.info.n_materials = desc->n_materials;
  s.info.n_textures = desc->n_textures;
  s.info.bvh_nodes = (uint32_t)bvh.nodes.size();
  s.info.bvh_node_bytes = sizeof(Node8);
  s.info.tri_bytes = sizeof(TriRecord);
  s.info.shade_bytes = sizeof(ShadeRecord);
  s.info.texture_bytes = total;
  s.info.smem_resident = s.smem_scene && ctx->smem_scene ? 1u : 0u;
  rollback.armed = false;
  ctx->scenes.push_back(std::move(s));
  *scene_id = (int32_t)ctx->scenes.size() - 1;
  return PTC_OK;
}

} // namespace ptc

extern "C" {

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
ptc_cubemap_faces_from_cross(const float* cross, uint32_t w, uint32_t h, uint32_t nb_chan, uint32_t rgb_hex,
                             float* faces_out, uint64_t capacity, uint32_t* size_out, int32_t* used_fallback)
{
  if (!faces_out || !size_out)
    return PTC_ERR_INVALID;
  uint32_t size = w / 4u; // gpu_processor.cpp:101
  const bool usable = cross && nb_chan >= 3 && size > 0 && size == h / 3u && !(size & (size - 1u));
  if (used_fallback)
    *used_fallback = (cross && !usable) ? 1 : 0;
  if (!usable) {
    // createUnitTexture (gpu_processor.cpp:38-58): six 1x1 faces of one colour, alpha 0
    // a cross that fails the checks, or no colour given: the reference's default 0x131b23
    // (gpu_processor.cpp:75,99-105,128-132); "0x000000" in a scene file is black (:88-93)
    const uint32_t color = (cross || rgb_hex == PTC_CUBEMAP_DEFAULT_COLOR) ? 0x131b23u : (rgb_hex & 0xFFFFFFu);
    if (capacity < 24)
      return PTC_ERR_INVALID;
    for (int f = 0; f < 6; ++f) {
      faces_out[4 * f + 0] = (float)((color >> 16) & 0xFFu) / 255.0f;
      faces_out[4 * f + 1] = (float)((color >> 8) & 0xFFu) / 255.0f;
      faces_out[4 * f + 2] = (float)(color & 0xFFu) / 255.0f;
      faces_out[4 * f + 3] = 0.0f;
    }
    *size_out = 1;
    return PTC_OK;
  }
  if (capacity < (uint64_t)size * size * 24u)
    return PTC_ERR_INVALID;
  // cell (column, row) of the cross for +x, -x, +y, -y, +z, -z
  static const uint32_t cell[6][2] = { { 2, 1 }, { 0, 1 }, { 1, 0 }, { 1, 2 }, { 1, 1 }, { 3, 1 } };
  for (int f = 0; f < 6; ++f) {
    float* dst = faces_out + (size_t)f * size * size * 4u;
    for (uint32_t y = 0; y < size; ++y) {
      const float* src = cross + ((size_t)(cell[f][1] * size + y) * w + cell[f][0] * size) * nb_chan;
      for (uint32_t x = 0; x < size; ++x, src += nb_chan, dst += 4) {
        dst[0] = src[0];
        dst[1] = src[1];
        dst[2] = src[2];
        dst[3] = 0.0f;
      }
    }
  }
  *size_out = size;
  return PTC_OK;
}

int
ptc_set_cubemap_from_cross(ptc_context* ctx, const float* cross, uint32_t w, uint32_t h, uint32_t nb_chan,
                           uint32_t rgb_hex, int32_t* cubemap_id, int32_t* used_fallback)
{
  if (!ctx || !cubemap_id)
    return PTC_ERR_INVALID;
  const uint32_t s = w / 4u;
  const bool usable = cross && nb_chan >= 3 && s > 0 && s == h / 3u && !(s & (s - 1u));
  std::vector<float> faces(usable ? (size_t)s * s * 24u : 24u);
  uint32_t size = 0;
  int rc = ptc_cubemap_faces_from_cross(cross, w, h, nb_chan, rgb_hex, faces.data(), faces.size(), &size,
                                        used_fallback);
  if (rc)
    return fail(ctx, rc, "cube cross extraction failed");
  return ptc_set_cubemap(ctx, faces.data(), size, cubemap_id);
}

int
ptc_write_ppm(const char* path, const void* rgba, uint32_t width, uint32_t height)
{
  if (!path || !rgba || !width || !height)
    return PTC_ERR_INVALID;
  FILE* f = fopen(path, "wb");
  if (!f)
    return PTC_ERR_INVALID;
  fprintf(f, "P6\n%u %u\n255\n", width, height);
  const unsigned char* px = (const unsigned char*)rgba;
  std::vector<unsigned char> row((size_t)width * 3);
  bool ok = true;
  for (uint32_t y = 0; y < height && ok; ++y) {
    for (uint32_t x = 0; x < width; ++x) {
      const unsigned char* p = px + ((size_t)y * width + x) * 4;
      row[3 * x] = p[0], row[3 * x + 1] = p[1], row[3 * x + 2] = p[2];
    }
    ok = fwrite(row.data(), 1, row.size(), f) == row.size();
  }
  ok = fclose(f) == 0 && ok;
  return ok ? PTC_OK : PTC_ERR_INVALID;
}

int
ptc_accum_save(const char* path, const float* accum, uint32_t width, uint32_t rows, uint32_t samples)
{
  if (!path || !accum || !width || !rows)
    return PTC_ERR_INVALID;
  FILE* f = fopen(path, "wb");
  if (!f)
    return PTC_ERR_INVALID;
  const uint32_t hdr[4] = { 1u, width, rows, samples };
  const size_t n = (size_t)width * rows * 3;
  bool ok = fwrite("PTAC", 1, 4, f) == 4 && fwrite(hdr, 4, 4, f) == 4 && fwrite(accum, 4, n, f) == n;
  ok = fclose(f) == 0 && ok;
  return ok ? PTC_OK : PTC_ERR_INVALID;
}

int
ptc_accum_load(const char* path, float* accum, uint64_t capacity_floats, uint32_t* width, uint32_t* rows,
               uint32_t* samples)
{
  if (!path || !width || !rows || !samples)
    return PTC_ERR_INVALID;
  FILE* f = fopen(path, "rb");
  if (!f)
    return PTC_ERR_INVALID;
  char magic[4];
  uint32_t hdr[4];
  bool ok = fread(magic, 1, 4, f) == 4 && memcmp(magic, "PTAC", 4) == 0 && fread(hdr, 4, 4, f) == 4 && hdr[0] == 1u;
  if (ok) {
    *width = hdr[1], *rows = hdr[2], *samples = hdr[3];
    const uint64_t n = (uint64_t)hdr[1] * hdr[2] * 3;
    if (accum) // accum == NULL: header query only
      ok = n <= capacity_floats && fread(accum, 4, (size_t)n, f) == (size_t)n;
  }
  fclose(f);
  return ok ? PTC_OK : PTC_ERR_INVALID;
}

// The reference's device tree, read back record by record (see ptcore.h).
int
ptc_upload_scene_from_device(ptc_context* ctx, const ptc_ref_scenes* scenes, uint32_t scene_index,
                             int32_t* scene_id)
{
  if (!ctx || !scenes || !scene_id)
    return PTC_ERR_INVALID;
  return ptc::scene_from_device(ctx, scenes, scene_index,
                                [&](const ptc_scene_desc* d) { return ptc_upload_scene(ctx, d, scene_id); });
}

int
ptc_set_cubemap_from_array(ptc_context* ctx, const void* cuda_array, int32_t* cubemap_id)
{
  if (!ctx || !cuda_array || !cubemap_id)
    return PTC_ERR_INVALID;
  return ptc::cubemap_from_array(ctx, cuda_array, [&](const float* faces, uint32_t size) {
    return ptc_set_cubemap(ctx, faces, size, cubemap_id);
  });
}

} // extern "C"

// The reference's device tree / cubemap array read back to the host (through `ctx`'s device) and
// handed to `sink` -- ptc_upload_scene for one context, ptc_group_upload_scene for a group.
int
ptc::scene_from_device(ptc_context* ctx, const ptc_ref_scenes* scenes, uint32_t scene_index,
                       const std::function<int(const ptc_scene_desc*)>& sink)
{
  if (!scenes->scenes)
    return fail(ctx, PTC_ERR_INVALID, "Scenes.scenes is null");
  DeviceGuard g(ctx->device);
  struct RefSceneData // scene::SceneData, scene_data.h:71-76
  {
    ptc_ref_buffer meshes, materials, lights;
  };
  static_assert(sizeof(ptc_ref_buffer) == 16 && sizeof(RefSceneData) == 48, "reference layout");
  const void* d_sd = nullptr;
  CUDA_TRY(ctx, cudaMemcpy(&d_sd, scenes->scenes + scene_index, sizeof(void*), cudaMemcpyDeviceToHost));
  if (!d_sd)
    return fail(ctx, PTC_ERR_INVALID, "scene pointer is null");
  RefSceneData sd;
  CUDA_TRY(ctx, cudaMemcpy(&sd, d_sd, sizeof(sd), cudaMemcpyDeviceToHost));
  std::vector<ptc_ref_buffer> meshes(sd.meshes.size); // scene::Mesh == Buffer<Face>
  if (sd.meshes.size)
    CUDA_TRY(ctx, cudaMemcpy(meshes.data(), sd.meshes.data, meshes.size() * sizeof(ptc_ref_buffer),
                             cudaMemcpyDeviceToHost));
  std::vector<uint32_t> counts(meshes.size());
  size_t n_faces = 0;
  for (size_t m = 0; m < meshes.size(); ++m)
    n_faces += (counts[m] = meshes[m].size);
  std::vector<ptc_face> faces(n_faces);
  size_t off = 0;
  for (size_t m = 0; m < meshes.size(); ++m) {
    if (counts[m])
      CUDA_TRY(ctx, cudaMemcpy(faces.data() + off, meshes[m].data, counts[m] * sizeof(ptc_face),
                               cudaMemcpyDeviceToHost));
    off += counts[m];
  }
  std::vector<ptc_material> mats(sd.materials.size);
  std::vector<ptc_light> lights(sd.lights.size);
  if (!mats.empty())
    CUDA_TRY(ctx, cudaMemcpy(mats.data(), sd.materials.data, mats.size() * sizeof(ptc_material),
                             cudaMemcpyDeviceToHost));
  if (!lights.empty())
    CUDA_TRY(ctx, cudaMemcpy(lights.data(), sd.lights.data, lights.size() * sizeof(ptc_light),
                             cudaMemcpyDeviceToHost));
  for (ptc_material& m : mats)
    m._pad = 0;
  // global texture table; texels come back to the host because ptc_upload_scene packs them
  // into the context's own arena (the reference keeps one cudaMalloc per texture)
  std::vector<ptc_texture> tex(scenes->textures.size);
  if (!tex.empty())
    CUDA_TRY(ctx, cudaMemcpy(tex.data(), scenes->textures.data, tex.size() * sizeof(ptc_texture),
                             cudaMemcpyDeviceToHost));
  std::vector<std::vector<float>> texels(tex.size());
  for (size_t i = 0; i < tex.size(); ++i) {
    if (tex[i].w < 0 || tex[i].h < 0 || tex[i].nb_chan < 0)
      return fail(ctx, PTC_ERR_INVALID, "negative texture dimensions in the device table");
    const size_t n = (size_t)tex[i].w * tex[i].h * tex[i].nb_chan;
    texels[i].resize(n);
    if (n)
      CUDA_TRY(ctx, cudaMemcpy(texels[i].data(), tex[i].data, n * sizeof(float), cudaMemcpyDeviceToHost));
    tex[i].data = texels[i].data();
  }
  ptc_scene_desc d;
  memset(&d, 0, sizeof(d));
  d.faces = faces.data();
  d.mesh_face_counts = counts.data();
  d.n_meshes = (uint32_t)counts.size();
  d.materials = mats.data();
  d.n_materials = (uint32_t)mats.size();
  d.lights = lights.data();
  d.n_lights = (uint32_t)lights.size();
  d.textures = tex.data();
  d.n_textures = (uint32_t)tex.size();
  return sink(&d);
}

int
ptc::cubemap_from_array(ptc_context* ctx, const void* cuda_array,
                        const std::function<int(const float*, uint32_t)>& sink)
{
  DeviceGuard g(ctx->device);
  cudaArray_t src = (cudaArray_t)const_cast<void*>(cuda_array);
  cudaChannelFormatDesc desc;
  cudaExtent ext;
  unsigned flags = 0;
  CUDA_TRY(ctx, cudaArrayGetInfo(&desc, &ext, &flags, src));
  if (!(flags & cudaArrayCubemap) || ext.depth != 6 || ext.width != ext.height ||
      desc.f != cudaChannelFormatKindFloat || desc.x != 32 || desc.y != 32 || desc.z != 32 || desc.w != 32)
    return fail(ctx, PTC_ERR_INVALID, "expected an RGBA32F cubemap array (gpu_processor.cpp:134-152)");
  // through the host: one-time, and keeps a single code path for array creation
  std::vector<float> faces((size_t)ext.width * ext.height * 6 * 4);
  cudaMemcpy3DParms p;
  memset(&p, 0, sizeof(p));
  p.srcArray = src;
  p.dstPtr = make_cudaPitchedPtr(faces.data(), ext.width * sizeof(float) * 4, ext.width, ext.height);
  p.extent = make_cudaExtent(ext.width, ext.height, 6);
  p.kind = cudaMemcpyDeviceToHost;
  CUDA_TRY(ctx, cudaMemcpy3D(&p));
  return sink(faces.data(), (uint32_t)ext.width);
}

extern "C" {

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

// ---- ptc_render_static: the frame with no allocation -------------------------------------------------
// Scratch of one pipeline of launches (two ray queues, hit records, parked pixels, queue / work
// counters) for PTC_STATIC_MAX_PIXELS pixels, in a file-scope device array: every device that loads
// the library holds it until exit.  Host code only ever names it in cudaGetSymbolAddress.
namespace {

constexpr size_t kStaticSlotBytes = (size_t)PTC_STATIC_MAX_PIXELS * 16; // one float4 / uint4 per queue slot
constexpr size_t kStaticCountBytes = (sizeof(uint32_t) * kCountSlots + 255) & ~(size_t)255;
constexpr size_t kStaticBytes = 11 * kStaticSlotBytes + kStaticCountBytes;

static __device__ __align__(256) unsigned char g_static_scratch[kStaticBytes];

constexpr int kStaticMaxDevices = 64;
std::mutex g_static_mu;
unsigned char* g_static_base[kStaticMaxDevices] = {};

// Device address of the array on `device` (which must be current); looked up once per device.
cudaError_t
static_scratch_base(int device, unsigned char** out)
{
  std::lock_guard<std::mutex> lock(g_static_mu);
  if (!g_static_base[device]) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
    if (e != cudaSuccess)
      return e;
    g_static_base[device] = (unsigned char*)p;
  }
  *out = g_static_base[device];
  return cudaSuccess;
}

} // namespace

int
ptc_render_static(ptc_context* ctx, const ptc_render_args* args)
{
  if (!ctx)
    return PTC_ERR_INVALID;
  if (ctx->device < 0 || ctx->device >= kStaticMaxDevices)
    return fail(ctx, PTC_ERR_INVALID, "ptc_render_static: device ordinal beyond the per-device address table");
  DeviceGuard g(ctx->device);
  // geometry first: an oversized frame is refused before the frame counter moves
  FrameParams fp;
  memset(&fp, 0, sizeof(fp));
  if (args && args->width && args->height && band_geometry(args, fp)) {
    const uint64_t px = (uint64_t)fp.band_rows * fp.width;
    if (px > PTC_STATIC_MAX_PIXELS) {
      char msg[160];
      snprintf(msg, sizeof(msg), "ptc_render_static: band of %llu pixels exceeds the static scratch limit of %u pixels",
               (unsigned long long)px, (unsigned)PTC_STATIC_MAX_PIXELS);
      return fail(ctx, PTC_ERR_INVALID, msg);
    }
  }
  int rc = fill_frame(ctx, args, fp, true);
  if (rc)
    return rc;
  if (!args->accum)
    return fail(ctx, PTC_ERR_INVALID, "accum is required");
  unsigned char* base = nullptr;
  CUDA_TRY(ctx, static_scratch_base(ctx->device, &base));
  Lane L;
  float4* slot = (float4*)base;
  const size_t P = PTC_STATIC_MAX_PIXELS;
  for (RayQueue& q : L.q) {
    q.o = slot, q.d = slot + P, q.thr = slot + 2 * P, q.acc = slot + 3 * P;
    slot += 4 * P;
  }
  L.hits = (uint4*)slot;
  L.hits2 = (uint4*)(slot + P);
  L.done = slot + 2 * P;
  L.counts = (uint32_t*)(base + 11 * kStaticSlotBytes); // cleared by enqueue_part on the caller's stream
  fp.part_row0 = 0;
  fp.part_rows = fp.band_rows;
  const FrameOut fo{ (float*)args->accum, (uint32_t*)args->out_rgba, (float*)args->radiance_out,
                     (args->flags & PTC_OUT_FULL_IMAGE) ? 1u : 0u };
  return enqueue_part(ctx, args, fp, fo, (cudaStream_t)args->stream, L, false, nullptr);
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
  const size_t n = (size_t)fp.band_rows * fp.width;
  cudaStream_t s = (cudaStream_t)args->stream;
  bool fresh = false;
  rc = ensure_staging(ctx, args, fp, n, fresh);
  if (rc)
    return rc;
  if (fresh)
    upload_accum = 1; // a fresh device accumulator must be seeded from the caller's
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
ptc_render_host_async(ptc_context* ctx, const ptc_render_args* args, int upload_accum)
{
  if (!ctx)
    return PTC_ERR_INVALID;
  DeviceGuard g(ctx->device);
  FrameParams fp;
  int rc = fill_frame(ctx, args, fp, true);
  if (rc)
    return rc;
  if (!args->accum || !args->out_rgba)
    return fail(ctx, PTC_ERR_INVALID, "accum and out_rgba are required");
  const size_t n = (size_t)fp.band_rows * fp.width;
  cudaStream_t s = (cudaStream_t)args->stream;
  if (!ctx->copy_stream) {
    CUDA_TRY(ctx, cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
    for (int i = 0; i < 2; ++i) {
      CUDA_TRY(ctx, cudaEventCreateWithFlags(&ctx->ev_render[i], cudaEventDisableTiming));
      CUDA_TRY(ctx, cudaEventCreateWithFlags(&ctx->ev_copy[i], cudaEventDisableTiming));
    }
  }
  bool fresh = false;
  rc = ensure_staging(ctx, args, fp, n, fresh);
  if (rc)
    return rc;
  if (fresh || upload_accum)
    CUDA_TRY(ctx, cudaMemcpyAsync(ctx->h_accum_dev, args->accum, n * 12, cudaMemcpyHostToDevice, s));
  const int slot = ctx->async_slot ^= 1;
  uint32_t* d_out = slot ? ctx->h_out_dev2 : ctx->h_out_dev;
  // the copy that last read this device frame must be done before it is overwritten
  CUDA_TRY(ctx, cudaStreamWaitEvent(s, ctx->ev_copy[slot], 0));
  rc = enqueue_frame(ctx, args, fp, ctx->h_accum_dev, d_out, nullptr, s);
  if (rc)
    return rc;
  CUDA_TRY(ctx, cudaEventRecord(ctx->ev_render[slot], s));
  CUDA_TRY(ctx, cudaStreamWaitEvent(ctx->copy_stream, ctx->ev_render[slot], 0));
  CUDA_TRY(ctx, cudaMemcpyAsync(args->out_rgba, d_out, n * 4, cudaMemcpyDeviceToHost, ctx->copy_stream));
  CUDA_TRY(ctx, cudaEventRecord(ctx->ev_copy[slot], ctx->copy_stream));
  return PTC_OK;
}

int
ptc_wait(ptc_context* ctx, void* accum_out, uint64_t accum_bytes)
{
  if (!ctx)
    return PTC_ERR_INVALID;
  DeviceGuard g(ctx->device);
  if (ctx->copy_stream)
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->copy_stream));
  if (accum_out) {
    if (accum_bytes != ctx->h_px * 12)
      return fail(ctx, PTC_ERR_INVALID, "accumulator size does not match the last host-buffer call");
    CUDA_TRY(ctx, cudaMemcpy(accum_out, ctx->h_accum_dev, accum_bytes, cudaMemcpyDeviceToHost));
  }
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

uint32_t
ptc_band_rows(const ptc_render_args* args)
{
  FrameParams fp;
  memset(&fp, 0, sizeof(fp));
  if (!args || !args->width || !args->height || !band_geometry(args, fp))
    return 0;
  return fp.band_rows;
}

int
ptc_launch_count(const ptc_context* ctx, uint64_t* out)
{
  if (!ctx || !out)
    return PTC_ERR_INVALID;
  *out = ctx->launches;
  return PTC_OK;
}

int
ptc_set_wavefront(ptc_context* ctx, uint32_t first_iters, uint32_t next_iters, int32_t split_extend,
                  uint32_t refill)
{
  if (!ctx)
    return PTC_ERR_INVALID;
  if (split_extend >= 0)
    ctx->split_extend = split_extend > 2 ? 2 : split_extend;
  if (refill)
    ctx->refill = refill > 32 ? 32 : refill;
  if (first_iters == 0xFFFFFFFFu) {
    ctx->first_iters = 0; // automatic (per scene size) ...
    ctx->split_extend = -1; // ... for the whole schedule
    ctx->next_iters = 1;
    ctx->refill = kDefaultRefill; // ... including the measured refill threshold
  } else if (first_iters)
    ctx->first_iters = first_iters;
  if (next_iters)
    ctx->next_iters = next_iters;
  return PTC_OK;
}

int
ptc_profile_frame(ptc_context* ctx, const ptc_render_args* args, uint32_t repeats, ptc_profile* out)
{
  if (!ctx || !out || !repeats)
    return PTC_ERR_INVALID;
  DeviceGuard g(ctx->device);
  memset(out, 0, sizeof(*out));
  FrameParams fp;
  int rc = fill_frame(ctx, args, fp, false);
  if (rc)
    return rc;
  cudaStream_t s = (cudaStream_t)args->stream;
  Step steps[kMaxBounces + 2];
  bool split = false, wave = false;
  const int n_steps = frame_schedule(ctx, fp, scene_bvh_bytes(ctx, args->scene_id), steps, &split, &wave);
  // the launches of one frame: kind 0 = first stage, 1 = fused later stage, 2 = extend, 3 = shade,
  // 4 / 5 = shifted form: generate + intersect / shade + intersect
  struct L { uint32_t kind, bounce0, iters; };
  std::vector<L> launches;
  for (int k = 0; k < n_steps; ++k) {
    if (wave)
      launches.push_back({ steps[k].first ? 4u : 5u, (uint32_t)steps[k].bounce0, 1u });
    else if (steps[k].first)
      launches.push_back({ 0u, 0u, (uint32_t)steps[k].n_iters });
    else if (split) {
      launches.push_back({ 2u, (uint32_t)steps[k].bounce0, 1u });
      launches.push_back({ 3u, (uint32_t)steps[k].bounce0, 1u });
    } else
      launches.push_back({ 1u, (uint32_t)steps[k].bounce0, (uint32_t)steps[k].n_iters });
  }
  const int n_l = (int)launches.size();
  EventList evl, evf;
  CUDA_TRY(ctx, evl.create((size_t)(n_l + 1) * repeats));
  CUDA_TRY(ctx, evf.create(2));
  std::vector<cudaEvent_t>& ev = evl.ev;
  // every pass renders into SCRATCH buffers: the caller's accumulator and image are not touched
  // (repeating one sample number into a live accumulator would bias it)
  const size_t n_px = (size_t)fp.band_rows * fp.width;
  DevBuf b_acc, b_out;
  CUDA_TRY(ctx, b_acc.alloc(n_px * 12));
  CUDA_TRY(ctx, b_out.alloc(n_px * 4));
  CUDA_TRY(ctx, cudaMemsetAsync(b_acc.p, 0, n_px * 12, s));
  float* accum = b_acc.as<float>();
  uint32_t* out_rgba = b_out.as<uint32_t>();
  // statistics pass (untimed): ray counts and BVH node / triangle visits of this frame
  Counters before, after;
  CUDA_TRY(ctx, cudaStreamSynchronize(s));
  CUDA_TRY(ctx, cudaMemcpy(&before, ctx->ctr, sizeof(before), cudaMemcpyDeviceToHost));
  rc = enqueue_frame(ctx, args, fp, accum, out_rgba, nullptr, s, true);
  if (rc)
    return rc;
  CUDA_TRY(ctx, cudaStreamSynchronize(s));
  CUDA_TRY(ctx, cudaMemcpy(&after, ctx->ctr, sizeof(after), cudaMemcpyDeviceToHost));
  std::vector<uint32_t> counts(kMaxBounces + 2);
  CUDA_TRY(ctx, cudaMemcpy(counts.data(), ctx->counts, counts.size() * 4, cudaMemcpyDeviceToHost));
  // the statistics pass must not show up in the caller's ray counters
  CUDA_TRY(ctx, cudaMemcpy(ctx->ctr, &before, sizeof(before), cudaMemcpyHostToDevice));
  out->traced = after.traced - before.traced;
  out->segments = after.segments - before.segments;
  out->nodes_visited = after.nodes_visited - before.nodes_visited;
  out->tris_tested = after.tris_tested - before.tris_tested;
  for (int i = 0; i < 16; ++i)
    out->rays_per_bounce[i] = after.rays_per_bounce[i] - before.rays_per_bounce[i];
  out->extend_lanes[0] = after.ext_iters - before.ext_iters;
  out->extend_lanes[1] = after.ext_have - before.ext_have;
  out->extend_lanes[2] = after.ext_node - before.ext_node;
  out->extend_lanes[3] = after.ext_tri1 - before.ext_tri1;
  out->extend_lanes[4] = after.ext_tri2 - before.ext_tri2;
  for (int k = 1; k < n_steps; ++k)
    out->queue_rays += counts[k];
  out->queue_bytes_per_ray = 4 * sizeof(float4);
  out->hit_bytes_per_ray = split || wave ? (uint32_t)sizeof(uint4) : 0u;
  // timed passes, one pipeline, an event after every launch
  for (uint32_t r = 0; r < repeats; ++r) {
    rc = enqueue_frame(ctx, args, fp, accum, out_rgba, nullptr, s, false, ev.data() + (size_t)r * (n_l + 1));
    if (rc)
      return rc;
  }
  CUDA_TRY(ctx, cudaStreamSynchronize(s));
  out->n_launches = (uint32_t)(n_l < 130 ? n_l : 130);
  for (int k = 0; k < n_l && k < 130; ++k) {
    float sum = 0.f;
    for (uint32_t r = 0; r < repeats; ++r) {
      float ms = 0.f;
      cudaEventElapsedTime(&ms, ev[(size_t)r * (n_l + 1) + k], ev[(size_t)r * (n_l + 1) + k + 1]);
      sum += ms;
    }
    out->launch_ms[k] = sum / repeats;
    out->launch_kind[k] = launches[k].kind;
    out->launch_bounce0[k] = launches[k].bounce0;
    out->launch_iters[k] = launches[k].iters;
  }
  float total = 0.f;
  for (uint32_t r = 0; r < repeats; ++r) {
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ev[(size_t)r * (n_l + 1)], ev[(size_t)r * (n_l + 1) + n_l]);
    total += ms;
  }
  out->frame_ms = total / repeats;
  // the frame as ptc_render ships it (half-frame pipelines on two streams), `repeats` frames back to back
  CUDA_TRY(ctx, cudaEventRecord(evf.ev[0], s));
  for (uint32_t r = 0; r < repeats; ++r) {
    rc = enqueue_frame(ctx, args, fp, accum, out_rgba, nullptr, s);
    if (rc)
      return rc;
  }
  CUDA_TRY(ctx, cudaEventRecord(evf.ev[1], s));
  CUDA_TRY(ctx, cudaStreamSynchronize(s));
  float shipped = 0.f;
  cudaEventElapsedTime(&shipped, evf.ev[0], evf.ev[1]);
  out->frame_ms_shipped = shipped / repeats;
  out->pipelines = frame_pipelines(ctx, fp, scene_bvh_bytes(ctx, args->scene_id));
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
  fp.band_rows = height;
  fp.strip_rows = 16;
  fp.strip_stride = 1;
  fp.tiles_x = (width + kTileW - 1) / kTileW;
  fp.frame_nb = (int)(frame_nb ? frame_nb : 1);
  fp.hash_seed = wang_hash((uint32_t)fp.frame_nb);
  fp.max_bounces = 1;
  fp.is_static = 1;
  const uint32_t n = width * height;
  DevBuf b_ids, b_t;
  CUDA_TRY(ctx, b_ids.alloc((size_t)n * 4));
  CUDA_TRY(ctx, b_t.alloc((size_t)n * 4));
  launch_primary_hits(fp, ctx->scenes[scene_id].d, b_ids.as<int32_t>(), b_t.as<float>(), 0);
  ctx->launches++;
  CUDA_TRY(ctx, cudaDeviceSynchronize());
  if (hit_ids)
    CUDA_TRY(ctx, cudaMemcpy(hit_ids, b_ids.p, (size_t)n * 4, cudaMemcpyDeviceToHost));
  if (t)
    CUDA_TRY(ctx, cudaMemcpy(t, b_t.p, (size_t)n * 4, cudaMemcpyDeviceToHost));
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
  DevBuf b_o, b_d, b_i, b_t;
  CUDA_TRY(ctx, b_o.alloc(n * 12));
  CUDA_TRY(ctx, b_d.alloc(n * 12));
  CUDA_TRY(ctx, b_i.alloc(n * 4));
  CUDA_TRY(ctx, b_t.alloc(n * 4));
  CUDA_TRY(ctx, cudaMemcpy(b_o.p, origins, n * 12, cudaMemcpyHostToDevice));
  CUDA_TRY(ctx, cudaMemcpy(b_d.p, dirs, n * 12, cudaMemcpyHostToDevice));
  launch_intersect_rays(ctx->scenes[scene_id].d, b_o.as<float>(), b_d.as<float>(), n, b_i.as<int32_t>(),
                        b_t.as<float>(), ctx->ctr, 0);
  ctx->launches++;
  CUDA_TRY(ctx, cudaDeviceSynchronize());
  if (hit_ids)
    CUDA_TRY(ctx, cudaMemcpy(hit_ids, b_i.p, n * 4, cudaMemcpyDeviceToHost));
  if (t)
    CUDA_TRY(ctx, cudaMemcpy(t, b_t.p, n * 4, cudaMemcpyDeviceToHost));
  return PTC_OK;
}

int
ptc_debug_read_bandwidth(ptc_context* ctx, uint64_t working_set_bytes, uint64_t total_bytes, float* gb_per_s)
{
  if (!ctx || !gb_per_s || working_set_bytes < 16 || !total_bytes)
    return PTC_ERR_INVALID;
  DeviceGuard g(ctx->device);
  uint64_t ws_vec = 1;
  while (ws_vec * 2 * 16 <= working_set_bytes && ws_vec < (1ull << 31))
    ws_vec *= 2; // power of two <= the request
  uint4* buf = nullptr;
  uint32_t* sink = nullptr;
  CUDA_TRY(ctx, cudaMalloc(&buf, ws_vec * 16));
  CUDA_TRY(ctx, cudaMalloc(&sink, 4));
  cudaMemset(buf, 0, ws_vec * 16);
  const uint32_t grid = (uint32_t)ctx->n_sms * 8u;
  const uint64_t threads = (uint64_t)grid * 256u;
  uint64_t per_thread = (total_bytes / 16 + threads - 1) / threads;
  per_thread = (per_thread + 3) & ~(uint64_t)3;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0), cudaEventCreate(&e1);
  float best = 0.f;
  for (int rep = 0; rep < 4; ++rep) { // rep 0 warms the cache tier under test
    cudaEventRecord(e0, 0);
    launch_debug_read_bw(buf, (uint32_t)ws_vec, per_thread, grid, sink, 0);
    cudaEventRecord(e1, 0);
    cudaEventSynchronize(e1);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    const float gbs = (float)((double)per_thread * threads * 16.0 / (ms * 1e-3) / 1e9);
    if (rep && gbs > best)
      best = gbs;
  }
  cudaEventDestroy(e0), cudaEventDestroy(e1);
  cudaError_t e = cudaDeviceSynchronize();
  cudaFree(buf), cudaFree(sink);
  if (e != cudaSuccess)
    return fail(ctx, (int)e, cudaGetErrorString(e));
  *gb_per_s = best;
  return PTC_OK;
}

int
ptc_debug_rng(ptc_context* ctx, const uint32_t* seeds, uint32_t n_seeds, uint32_t n_draws,
              int reference_impl, float* out)
{
  if (!ctx || !seeds || !out || !n_seeds || !n_draws)
    return PTC_ERR_INVALID;
  DeviceGuard g(ctx->device);
  DevBuf b_s, b_o;
  CUDA_TRY(ctx, b_s.alloc((size_t)n_seeds * 4));
  CUDA_TRY(ctx, b_o.alloc((size_t)n_seeds * n_draws * 4));
  CUDA_TRY(ctx, cudaMemcpy(b_s.p, seeds, (size_t)n_seeds * 4, cudaMemcpyHostToDevice));
  if (reference_impl)
    launch_debug_curand(b_s.as<uint32_t>(), n_seeds, n_draws, b_o.as<float>(), 0);
  else
    launch_debug_rng(b_s.as<uint32_t>(), n_seeds, n_draws, b_o.as<float>(), 0);
  CUDA_TRY(ctx, cudaDeviceSynchronize());
  CUDA_TRY(ctx, cudaMemcpy(out, b_o.p, (size_t)n_seeds * n_draws * 4, cudaMemcpyDeviceToHost));
  return PTC_OK;
}

} // extern "C"
