// group.cu -- several GPUs behind ONE call of the C ABI (include/ptcore.h, ptc_group_*).
//
// The reference drives one GPU (cuda_opengl/src/driver/gpu_info.cpp:26-28,38-43) from one host
// thread (GPUProcessor::render, cuda_opengl/src/gpu_processor.cpp:365-386).  A group keeps that
// call shape -- one upload, one render call per frame, one image -- and spreads the frame over
// N devices of one process:
//
//   * the scene is prepared once on the host (BVH build, shading records) and uploaded to every
//     device (replicated; pixels are the independent units);
//   * the image is split into round-robin 16-row strips (strip s -> device s mod N: the reference's
//     16x16 blocks stay whole, every device sees the same mix of cheap and expensive rows);
//   * every device accumulates its own strips in its own accumulator;
//   * the tonemapped RGBA8 pixel is stored by the epilogue of each device's last stage kernel
//     STRAIGHT INTO THE FULL IMAGE ON THE ROOT DEVICE through the peer mapping
//     (PTC_OUT_FULL_IMAGE): the frame is assembled by the stores that produce it, over NVLink,
//     with no gather pass, no staging copy and no collective;
//   * one worker thread per device enqueues that device's launches, so the host-side launch cost
//     of a frame does not grow with N; cross-device ordering is by CUDA events only (the root
//     stream waits for every device's frame), nothing synchronises with the host.
//
// Also here: the CUDA IPC helpers and the flag kernels that give the same direct-store assembly
// to a one-process-per-GPU job (bench.py under torchrun).
#include <cuda_runtime.h>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <atomic>
#include <condition_variable>
#include <cstring>
#include <functional>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "ptcore.h"
#include "ptcore_internal.h"

namespace {

struct Worker
{
  std::thread th;
  std::mutex m;
  std::condition_variable cv;
  std::function<int()> job;
  bool has_job = false, stop = false, done = true;
  int rc = 0;

  void run()
  {
    for (;;) {
      std::function<int()> j;
      {
        std::unique_lock<std::mutex> l(m);
        cv.wait(l, [&] { return has_job || stop; });
        if (stop)
          return;
        j = std::move(job);
        has_job = false;
      }
      const int r = j();
      {
        std::lock_guard<std::mutex> l(m);
        rc = r;
        done = true;
      }
      cv.notify_all();
    }
  }
  void submit(std::function<int()> j)
  {
    {
      std::lock_guard<std::mutex> l(m);
      job = std::move(j);
      has_job = true;
      done = false;
    }
    cv.notify_all();
  }
  int wait()
  {
    std::unique_lock<std::mutex> l(m);
    cv.wait(l, [&] { return done; });
    return rc;
  }
};

struct Member
{
  int device = 0;
  ptc_context* ctx = nullptr;
  cudaStream_t stream = nullptr;   // this member's launches (the root uses the caller's stream)
  cudaEvent_t ev_done = nullptr;   // recorded after the member's frame
  float* accum = nullptr;          // this member's strips, band layout of ptc_render
  size_t accum_px = 0;
  Worker* worker = nullptr;
};

} // namespace

struct ptc_group
{
  std::vector<Member> members;
  cudaEvent_t ev_start = nullptr;  // recorded on the root stream when a frame is issued
  uint32_t* image = nullptr;       // group-owned full image on the root device (out_rgba == NULL)
  size_t image_px = 0;
  uint32_t width = 0, height = 0;  // geometry of the accumulators
  uint32_t seed = 0;               // the reference's frame counter (raytrace.cu:296-300), one per group
  std::string err;
};

namespace {

int
gfail(ptc_group* g, int code, const std::string& msg)
{
  if (g)
    g->err = msg;
  return code;
}

struct DevGuard
{
  int prev = -1;
  explicit DevGuard(int d)
  {
    cudaGetDevice(&prev);
    cudaSetDevice(d);
  }
  ~DevGuard()
  {
    if (prev >= 0)
      cudaSetDevice(prev);
  }
};

// f(i) for every member on its worker thread; returns the first failure
int
for_all(ptc_group* g, const std::function<int(uint32_t)>& f)
{
  const uint32_t n = (uint32_t)g->members.size();
  for (uint32_t i = 1; i < n; ++i)
    g->members[i].worker->submit([&f, i] { return f(i); });
  int rc = f(0);
  for (uint32_t i = 1; i < n; ++i) {
    const int r = g->members[i].worker->wait();
    if (!rc && r)
      rc = r;
  }
  return rc;
}

int
member_fail(ptc_group* g, uint32_t i, int rc, const char* what)
{
  return gfail(g, rc, std::string(what) + " on group member " + std::to_string(i) + " (device " +
                        std::to_string(g->members[i].device) + "): " + ptc_last_error(g->members[i].ctx));
}

uint32_t
strips_rows(uint32_t height, uint32_t n, uint32_t phase)
{
  ptc_render_args a;
  memset(&a, 0, sizeof(a));
  a.width = 1, a.height = height;
  a.strip_rows = 16, a.strip_stride = n, a.strip_phase = phase;
  if (n <= 1)
    a.strip_stride = 0;
  return ptc_band_rows(&a);
}

int
ensure_buffers(ptc_group* g, uint32_t w, uint32_t h, bool need_image)
{
  const uint32_t n = (uint32_t)g->members.size();
  if (g->width != w || g->height != h) {
    for (uint32_t i = 0; i < n; ++i) {
      Member& m = g->members[i];
      DevGuard dg(m.device);
      cudaFree(m.accum);
      m.accum = nullptr;
      m.accum_px = (size_t)strips_rows(h, n, i) * w;
      if (m.accum_px) {
        if (cudaMalloc(&m.accum, m.accum_px * 12) != cudaSuccess)
          return gfail(g, PTC_ERR_ALLOC, "group accumulator allocation failed");
        cudaMemset(m.accum, 0, m.accum_px * 12); // cudaCalloc of gpu_processor.cpp:255
      }
    }
    g->width = w, g->height = h;
    g->seed = 0;
  }
  if (need_image && g->image_px != (size_t)w * h) {
    DevGuard dg(g->members[0].device);
    cudaFree(g->image);
    g->image = nullptr, g->image_px = 0;
    if (cudaMalloc(&g->image, (size_t)w * h * 4) != cudaSuccess)
      return gfail(g, PTC_ERR_ALLOC, "group image allocation failed");
    cudaMemset(g->image, 0, (size_t)w * h * 4);
    g->image_px = (size_t)w * h;
  }
  return PTC_OK;
}

} // namespace

extern "C" {

int
ptc_group_create(const int* devices, uint32_t n, ptc_group** out)
{
  if (!devices || !n || !out)
    return PTC_ERR_INVALID;
  *out = nullptr;
  ptc_group* g = new ptc_group();
  g->members.resize(n);
  for (uint32_t i = 0; i < n; ++i) {
    Member& m = g->members[i];
    m.device = devices[i];
    int rc = ptc_create(devices[i], &m.ctx);
    if (rc) {
      ptc_group_destroy(g);
      return rc;
    }
    DevGuard dg(m.device);
    cudaStreamCreateWithFlags(&m.stream, cudaStreamNonBlocking);
    cudaEventCreateWithFlags(&m.ev_done, cudaEventDisableTiming);
    // every member stores into the root's image: map the root's memory into its address space
    if (i && devices[i] != devices[0]) {
      int can = 0;
      cudaDeviceCanAccessPeer(&can, devices[i], devices[0]);
      if (!can) {
        ptc_group_destroy(g);
        return PTC_ERR_NO_DEVICE;
      }
      cudaError_t e = cudaDeviceEnablePeerAccess(devices[0], 0);
      if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
        ptc_group_destroy(g);
        return (int)e;
      }
      cudaGetLastError();
    }
    if (i) {
      m.worker = new Worker();
      m.worker->th = std::thread([w = m.worker] { w->run(); });
    }
  }
  {
    DevGuard dg(devices[0]);
    cudaEventCreateWithFlags(&g->ev_start, cudaEventDisableTiming);
  }
  *out = g;
  return PTC_OK;
}

void
ptc_group_destroy(ptc_group* g)
{
  if (!g)
    return;
  for (Member& m : g->members) {
    if (m.worker) {
      {
        std::lock_guard<std::mutex> l(m.worker->m);
        m.worker->stop = true;
      }
      m.worker->cv.notify_all();
      if (m.worker->th.joinable())
        m.worker->th.join();
      delete m.worker;
    }
    if (!m.ctx)
      continue;
    DevGuard dg(m.device);
    cudaDeviceSynchronize();
    cudaFree(m.accum);
    if (m.stream)
      cudaStreamDestroy(m.stream);
    if (m.ev_done)
      cudaEventDestroy(m.ev_done);
    ptc_destroy(m.ctx);
  }
  if (!g->members.empty() && g->members[0].ctx) {
    DevGuard dg(g->members[0].device);
    cudaFree(g->image);
    if (g->ev_start)
      cudaEventDestroy(g->ev_start);
  }
  delete g;
}

const char*
ptc_group_last_error(const ptc_group* g)
{
  return g ? g->err.c_str() : "null group";
}

uint32_t
ptc_group_size(const ptc_group* g)
{
  return g ? (uint32_t)g->members.size() : 0u;
}

ptc_context*
ptc_group_context(ptc_group* g, uint32_t i)
{
  return g && i < g->members.size() ? g->members[i].ctx : nullptr;
}

int
ptc_group_upload_scene(ptc_group* g, const ptc_scene_desc* desc, int32_t* scene_id)
{
  if (!g || !desc || !scene_id)
    return PTC_ERR_INVALID;
  ptc::PreparedScene ps;
  std::string err;
  int rc = ptc::prepare_scene(desc, ps, err); // once, whatever the number of devices
  if (rc)
    return gfail(g, rc, err);
  std::vector<int32_t> ids(g->members.size(), -1);
  rc = for_all(g, [&](uint32_t i) { return ptc::upload_prepared(g->members[i].ctx, desc, ps, &ids[i]); });
  if (rc)
    return gfail(g, rc, "scene upload failed on a group member");
  for (int32_t id : ids)
    if (id != ids[0])
      return gfail(g, PTC_ERR_STATE, "group members disagree on the scene id (contexts used outside the group?)");
  *scene_id = ids[0];
  return PTC_OK;
}

int
ptc_group_set_cubemap(ptc_group* g, const float* faces, uint32_t size, int32_t* cubemap_id)
{
  if (!g || !faces || !size || !cubemap_id)
    return PTC_ERR_INVALID;
  std::vector<int32_t> ids(g->members.size(), -1);
  int rc = for_all(g, [&](uint32_t i) { return ptc_set_cubemap(g->members[i].ctx, faces, size, &ids[i]); });
  if (rc)
    return gfail(g, rc, "cubemap upload failed on a group member");
  for (int32_t id : ids)
    if (id != ids[0])
      return gfail(g, PTC_ERR_STATE, "group members disagree on the cubemap id");
  *cubemap_id = ids[0];
  return PTC_OK;
}

int
ptc_group_upload_scene_from_device(ptc_group* g, const ptc_ref_scenes* scenes, uint32_t scene_index, int32_t* scene_id)
{
  if (!g || !scenes || !scene_id)
    return PTC_ERR_INVALID;
  ptc_context* root = g->members[0].ctx; // the caller's tree lives on devices[0]
  bool read_back = false; // a failure before the sink ran is the read-back's: its message is in the root context
  const int rc = ptc::scene_from_device(root, scenes, scene_index, [&](const ptc_scene_desc* d) {
    read_back = true;
    return ptc_group_upload_scene(g, d, scene_id);
  });
  if (rc && !read_back)
    g->err = ptc_last_error(root);
  return rc;
}

int
ptc_group_set_cubemap_from_array(ptc_group* g, const void* cuda_array, int32_t* cubemap_id)
{
  if (!g || !cuda_array || !cubemap_id)
    return PTC_ERR_INVALID;
  ptc_context* root = g->members[0].ctx;
  bool read_back = false;
  const int rc = ptc::cubemap_from_array(root, cuda_array, [&](const float* faces, uint32_t size) {
    read_back = true;
    return ptc_group_set_cubemap(g, faces, size, cubemap_id);
  });
  if (rc && !read_back)
    g->err = ptc_last_error(root);
  return rc;
}

int
ptc_group_render(ptc_group* g, const ptc_render_args* args)
{
  if (!g || !args)
    return PTC_ERR_INVALID;
  if (args->row_begin || args->row_end || args->strip_stride > 1 || args->radiance_out || args->accum)
    return gfail(g, PTC_ERR_INVALID,
                 "group render: the group owns the accumulators and the split (accum, radiance_out, rows and strips must be 0)");
  const uint32_t n = (uint32_t)g->members.size();
  int rc = ensure_buffers(g, args->width, args->height, args->out_rgba == nullptr);
  if (rc)
    return rc;
  // one frame counter for the whole group (raytrace.cu:296-300), so that every member renders
  // the same sample number
  uint32_t frame = args->frame_nb;
  if (frame == 0) {
    if (args->moved)
      g->seed = 0;
    frame = ++g->seed;
  }
  uint32_t* image = args->out_rgba ? (uint32_t*)args->out_rgba : g->image;
  cudaStream_t root = (cudaStream_t)args->stream;
  {
    DevGuard dg(g->members[0].device);
    if (cudaEventRecord(g->ev_start, root) != cudaSuccess)
      return gfail(g, PTC_ERR_STATE, "cudaEventRecord failed on the root stream");
  }
  rc = for_all(g, [&](uint32_t i) -> int {
    Member& m = g->members[i];
    if (!m.accum_px)
      return PTC_OK; // fewer strips than devices
    DevGuard dg(m.device);
    ptc_render_args a = *args;
    a.frame_nb = frame;
    a.accum = m.accum;
    a.out_rgba = image;
    a.flags |= PTC_OUT_FULL_IMAGE;
    a.strip_rows = 16;
    a.strip_stride = n > 1 ? n : 0;
    a.strip_phase = n > 1 ? i : 0;
    a.stream = i ? m.stream : root;
    if (i) {
      // whatever the caller ordered before this frame on the root stream (the consumer of the
      // previous image) comes first
      if (cudaStreamWaitEvent(m.stream, g->ev_start, 0) != cudaSuccess)
        return PTC_ERR_STATE;
    }
    const int r = ptc_render(m.ctx, &a);
    if (r)
      return r;
    if (i && cudaEventRecord(m.ev_done, m.stream) != cudaSuccess)
      return PTC_ERR_STATE;
    return PTC_OK;
  });
  if (rc) {
    for (uint32_t i = 0; i < n; ++i)
      if (*ptc_last_error(g->members[i].ctx))
        return member_fail(g, i, rc, "ptc_render");
    return gfail(g, rc, "group render failed");
  }
  // the frame is complete, in root-stream order, when every member's stores have landed
  DevGuard dg(g->members[0].device);
  for (uint32_t i = 1; i < n; ++i)
    if (g->members[i].accum_px && cudaStreamWaitEvent(root, g->members[i].ev_done, 0) != cudaSuccess)
      return gfail(g, PTC_ERR_STATE, "cudaStreamWaitEvent failed on the root stream");
  return PTC_OK;
}

int
ptc_group_sync(ptc_group* g)
{
  if (!g)
    return PTC_ERR_INVALID;
  for (Member& m : g->members) {
    DevGuard dg(m.device);
    if (cudaDeviceSynchronize() != cudaSuccess)
      return gfail(g, PTC_ERR_STATE, "device synchronisation failed");
  }
  return PTC_OK;
}

const void*
ptc_group_image(const ptc_group* g)
{
  return g ? g->image : nullptr;
}

int
ptc_group_read_image(ptc_group* g, void* host_rgba, void* stream)
{
  if (!g || !host_rgba || !g->image)
    return PTC_ERR_INVALID;
  DevGuard dg(g->members[0].device);
  cudaStream_t s = (cudaStream_t)stream;
  if (cudaMemcpyAsync(host_rgba, g->image, g->image_px * 4, cudaMemcpyDeviceToHost, s) != cudaSuccess ||
      cudaStreamSynchronize(s) != cudaSuccess)
    return gfail(g, PTC_ERR_STATE, "image download failed");
  return PTC_OK;
}

// Full accumulator in the single-GPU layout (rows bottom-up over the WHOLE image, raytrace.cu:252)
// assembled on the host from the members' strip accumulators; synchronises.
int
ptc_group_gather_accum(ptc_group* g, float* host_accum)
{
  if (!g || !host_accum || !g->width)
    return PTC_ERR_INVALID;
  const uint32_t n = (uint32_t)g->members.size(), w = g->width, h = g->height;
  int rc = ptc_group_sync(g);
  if (rc)
    return rc;
  std::vector<float> band;
  for (uint32_t i = 0; i < n; ++i) {
    Member& m = g->members[i];
    if (!m.accum_px)
      continue;
    DevGuard dg(m.device);
    band.resize(m.accum_px * 3);
    if (cudaMemcpy(band.data(), m.accum, m.accum_px * 12, cudaMemcpyDeviceToHost) != cudaSuccess)
      return gfail(g, PTC_ERR_STATE, "accumulator download failed");
    const uint32_t band_rows = (uint32_t)(m.accum_px / w);
    const uint32_t n_strips = (h + 15u) / 16u;
    uint32_t lrow = 0;
    for (uint32_t s = i; s < n_strips; s += n) {
      const uint32_t r0 = s * 16u, r1 = r0 + 16u < h ? r0 + 16u : h;
      for (uint32_t y = r0; y < r1; ++y, ++lrow) {
        // band rows are stored bottom-up inside the band, image rows bottom-up inside the image
        const float* src = band.data() + (size_t)(band_rows - lrow - 1) * w * 3;
        float* dst = host_accum + (size_t)(h - y - 1) * w * 3;
        memcpy(dst, src, (size_t)w * 12);
      }
    }
  }
  return PTC_OK;
}

int
ptc_group_reset_accumulation(ptc_group* g)
{
  if (!g)
    return PTC_ERR_INVALID;
  g->seed = 0;
  for (Member& m : g->members) {
    if (!m.accum)
      continue;
    DevGuard dg(m.device);
    cudaMemset(m.accum, 0, m.accum_px * 12);
  }
  return PTC_OK;
}

// -- one process per GPU: CUDA IPC mapping of the root's image + flags ----------------------

int
ptc_shared_alloc(int device, uint64_t bytes, void** ptr, uint8_t handle[64])
{
  if (!ptr || !handle || !bytes)
    return PTC_ERR_INVALID;
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  DevGuard dg(device);
  if (cudaMalloc(ptr, bytes) != cudaSuccess)
    return PTC_ERR_ALLOC;
  cudaMemset(*ptr, 0, bytes);
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, *ptr);
  if (e != cudaSuccess) {
    cudaFree(*ptr);
    *ptr = nullptr;
    return (int)e;
  }
  memcpy(handle, &h, 64);
  return PTC_OK;
}

int
ptc_shared_open(int device, const uint8_t handle[64], void** ptr)
{
  if (!ptr || !handle)
    return PTC_ERR_INVALID;
  DevGuard dg(device);
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, 64);
  cudaError_t e = cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess);
  return e == cudaSuccess ? PTC_OK : (int)e;
}

int
ptc_shared_close(int device, void* ptr)
{
  DevGuard dg(device);
  return cudaIpcCloseMemHandle(ptr) == cudaSuccess ? PTC_OK : PTC_ERR_INVALID;
}

int
ptc_shared_free(int device, void* ptr)
{
  DevGuard dg(device);
  return cudaFree(ptr) == cudaSuccess ? PTC_OK : PTC_ERR_INVALID;
}

int
ptc_memcpy_to_host(int device, void* host, const void* dev, uint64_t bytes, void* stream, int sync)
{
  DevGuard dg(device);
  cudaStream_t s = (cudaStream_t)stream;
  if (cudaMemcpyAsync(host, dev, bytes, cudaMemcpyDeviceToHost, s) != cudaSuccess)
    return PTC_ERR_STATE;
  if (sync && cudaStreamSynchronize(s) != cudaSuccess)
    return PTC_ERR_STATE;
  return PTC_OK;
}

// -- one process per GPU, image consumed on the HOST: a POSIX shared-memory image every process maps
// and page-locks, and every rank's own strips copied into it over that rank's own PCIe link --------

int
ptc_host_image_open(const char* name, uint64_t bytes, int create, void** ptr)
{
  if (!name || !bytes || !ptr)
    return PTC_ERR_INVALID;
  *ptr = nullptr;
  const int fd = shm_open(name, create ? (O_CREAT | O_RDWR) : O_RDWR, 0600);
  if (fd < 0)
    return PTC_ERR_STATE;
  if (create && ftruncate(fd, (off_t)bytes) != 0) {
    close(fd);
    shm_unlink(name);
    return PTC_ERR_ALLOC;
  }
  void* p = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_SHARED | MAP_POPULATE, fd, 0);
  close(fd);
  if (p == MAP_FAILED) {
    if (create)
      shm_unlink(name);
    return PTC_ERR_ALLOC;
  }
  const cudaError_t e = cudaHostRegister(p, bytes, cudaHostRegisterPortable);
  if (e != cudaSuccess) {
    cudaGetLastError();
    munmap(p, bytes);
    if (create)
      shm_unlink(name);
    return (int)e;
  }
  *ptr = p;
  return PTC_OK;
}

int
ptc_host_image_close(void* ptr, uint64_t bytes, const char* unlink_name)
{
  if (!ptr || !bytes)
    return PTC_ERR_INVALID;
  cudaHostUnregister(ptr);
  munmap(ptr, bytes);
  if (unlink_name)
    shm_unlink(unlink_name);
  return PTC_OK;
}

} // extern "C"

namespace {

// A rank's band (band-local rows = its strips one after the other) -> the rows those strips have
// in a full image, by DMA: a pitched copy for the whole strips, one more for a short last strip.
int
copy_strips(int device, void* host_image, const void* dev_band, uint32_t width, uint32_t height, uint32_t strip_rows,
            uint32_t strip_stride, uint32_t strip_phase, void* stream, cudaMemcpyKind kind)
{
  if (!host_image || !dev_band || !width || !height || !strip_rows || !strip_stride || strip_phase >= strip_stride)
    return PTC_ERR_INVALID;
  DevGuard dg(device);
  cudaStream_t s = (cudaStream_t)stream;
  const size_t strip_bytes = (size_t)strip_rows * width * 4;
  const uint32_t n_strips = (height + strip_rows - 1) / strip_rows;
  if (strip_phase >= n_strips)
    return PTC_OK; // this rank owns no strip of an image this small
  const uint32_t mine = (n_strips - strip_phase + strip_stride - 1) / strip_stride; // strips phase, phase + stride, ...
  const uint32_t last = strip_phase + (mine - 1) * strip_stride;
  const uint32_t last_rows = (last + 1) * strip_rows <= height ? strip_rows : height - last * strip_rows;
  const uint32_t full = last_rows == strip_rows ? mine : mine - 1;
  uint8_t* dst = (uint8_t*)host_image + (size_t)strip_phase * strip_bytes;
  if (full && cudaMemcpy2DAsync(dst, strip_bytes * strip_stride, dev_band, strip_bytes, strip_bytes, full, kind, s) !=
                cudaSuccess)
    return PTC_ERR_STATE;
  if (full != mine &&
      cudaMemcpyAsync((uint8_t*)host_image + (size_t)last * strip_bytes, (const uint8_t*)dev_band + (size_t)full * strip_bytes,
                      (size_t)last_rows * width * 4, kind, s) != cudaSuccess)
    return PTC_ERR_STATE;
  return PTC_OK;
}

} // namespace

extern "C" {

int
ptc_memcpy_strips_to_host(int device, void* host_image, const void* dev_band, uint32_t width, uint32_t height,
                          uint32_t strip_rows, uint32_t strip_stride, uint32_t strip_phase, void* stream)
{
  return copy_strips(device, host_image, dev_band, width, height, strip_rows, strip_stride, strip_phase, stream,
                     cudaMemcpyDeviceToHost);
}

int
ptc_memcpy_strips_to_device(int device, void* image, const void* dev_band, uint32_t width, uint32_t height,
                            uint32_t strip_rows, uint32_t strip_stride, uint32_t strip_phase, void* stream)
{
  return copy_strips(device, image, dev_band, width, height, strip_rows, strip_stride, strip_phase, stream,
                     cudaMemcpyDefault);
}

} // extern "C"

// Flags in (possibly peer-mapped) device memory: "my strips of frame k are stored" / "frame k has
// been consumed".  A release store at system scope after the frame's kernels (stream order makes
// their stores precede it; the fence makes them visible to the other GPU before the flag), an
// acquire spin on the consumer.
namespace {

constexpr unsigned long long kFlagWaitNs = 20ull * 1000000000ull; // 20 s

__global__ void
k_flag_signal(uint32_t* flag, uint32_t value)
{
  __threadfence_system();
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(flag), "r"(value) : "memory");
}

__global__ void
k_flag_wait(const uint32_t* flags, uint32_t n, uint32_t stride_words, uint32_t value)
{
  const uint32_t i = threadIdx.x;
  if (i < n) {
    uint32_t v;
    unsigned long long t0, t1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    do {
      asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(flags + (size_t)i * stride_words) : "memory");
      if ((int32_t)(v - value) >= 0)
        break;
      __nanosleep(200);
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
      // a producer that never signals (a crashed rank, a caller that synchronised the device before
      // enqueueing it) must not hang the GPU: give up loudly after kFlagWaitNs
      if (t1 - t0 > kFlagWaitNs)
        __trap();
    } while (true);
  }
}

} // namespace

extern "C" {

int
ptc_flag_signal(int device, void* flag, uint32_t value, void* stream)
{
  if (!flag)
    return PTC_ERR_INVALID;
  DevGuard dg(device);
  k_flag_signal<<<1, 1, 0, (cudaStream_t)stream>>>((uint32_t*)flag, value);
  return cudaGetLastError() == cudaSuccess ? PTC_OK : PTC_ERR_STATE;
}

// waits until flags[i * stride_words] >= value (wrap-safe) for i < n, n <= 256
int
ptc_flag_wait(int device, const void* flags, uint32_t n, uint32_t stride_words, uint32_t value, void* stream)
{
  if (!flags || !n || n > 256)
    return PTC_ERR_INVALID;
  DevGuard dg(device);
  k_flag_wait<<<1, 256, 0, (cudaStream_t)stream>>>((const uint32_t*)flags, n, stride_words ? stride_words : 1u, value);
  return cudaGetLastError() == cudaSuccess ? PTC_OK : PTC_ERR_STATE;
}

} // extern "C"
