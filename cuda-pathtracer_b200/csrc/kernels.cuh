// kernels.cuh -- host-callable launchers of the wavefront kernels (kernels.cu).
#pragma once

#include "device_types.cuh"

namespace ptc {

// The stage kernels keep the per-ray traversal stack in dynamic shared memory:
// stack_cap entries (levels of the scene's wide tree + 1 + kTriBacklog) x kBlock threads x 8 B.
size_t stack_smem_bytes(int stack_cap);

// resident CTAs per SM of the path / extend kernels for that stack size (size the persistent
// grids; 0 = the stack does not fit).  Also opts the kernel in to the shared memory it needs.
int path_kernel_blocks_per_sm(bool first, bool stats, bool pretraced, int stack_cap);
int extend_kernel_blocks_per_sm(bool stats, int stack_cap);
// the fused path kernels with the scene's nodes + triangle records copied into shared memory
// behind the stacks (SharedSceneStack, traverse.cuh): bytes of that copy, and resident CTAs per SM
// (0 = does not fit)
size_t scene_smem_bytes(const DScene& sc);
int path_kernel_blocks_per_sm_scene(bool first, bool stats, int stack_cap, const DScene& sc);

// One wavefront step: `n_iters` x [extend -> shade] starting at radiance-loop iteration
// `bounce0`.  first: rays are generated (8x4 pixel tile per warp) instead of read from
// `qin`; survivors are appended to `qout` / `qout_count`.  `work` is a zeroed counter the
// persistent warps pull their work items from.  hits != null: the triangle hits of `qin`
// were already found by launch_extend (n_iters must be 1).
// done_mode: kDoneDefer = paths that finish in this stage are parked in done_q (count in
// done_ctl[0]) instead of running the per-pixel epilogue; kDoneDrain = this is the frame's last
// stage, which also runs the epilogue of every parked pixel (done_ctl[1] = its work counter).
constexpr int kDoneDefer = 1, kDoneDrain = 2;
void launch_path(bool first, bool stats, uint32_t grid, const FrameParams& fp, const DScene& sc,
                 cudaTextureObject_t cube, const RayQueue& qin, const uint4* hits,
                 const uint32_t* qin_count, const RayQueue& qout, uint32_t* qout_count,
                 const FrameOut& fo, Counters* ctr, uint32_t* work, int bounce0, int n_iters,
                 int stack_cap, float4* done_q, uint32_t* done_ctl, int done_mode, cudaStream_t s,
                 bool scene_in_smem = false);

// One stage of the SHIFTED schedule (k_wave): shade iteration bounce-1 from the queued hit records
// `hin` (not for the first stage), intersect iteration `bounce`, finish the rays that miss, queue
// the rays that hit with their hit records in `qout` / `hout`.
int wave_kernel_blocks_per_sm(bool first, bool stats, int stack_cap);
void launch_wave(bool first, bool stats, uint32_t grid, const FrameParams& fp, const DScene& sc,
                 cudaTextureObject_t cube, const RayQueue& qin, const uint4* hin,
                 const uint32_t* qin_count, const RayQueue& qout, uint4* hout, uint32_t* qout_count,
                 const FrameOut& fo, Counters* ctr, uint32_t* work, int bounce, int stack_cap,
                 float4* done_q, uint32_t* done_ctl, int done_mode, cudaStream_t s);

// Closest triangle hit of every ray of `qin` -> hits[slot] = (t, face id | ~0, u, v), with
// per-lane refill once `refill` lanes of a warp are idle.  last: the rays belong to the last
// radiance-loop iteration -> lights + any-hit query (traverse.cuh, trav_init_any); the hit
// record then already includes the sphere lights.
void launch_extend(bool stats, uint32_t grid, const DScene& sc, const RayQueue& qin,
                   const uint32_t* qin_count, uint4* hits, uint32_t* work, Counters* ctr,
                   uint32_t refill, bool last, int stack_cap, cudaStream_t s);

void launch_primary_hits(const FrameParams& fp, const DScene& sc, int32_t* ids, float* ts,
                         cudaStream_t s);
void launch_intersect_rays(const DScene& sc, const float* origins, const float* dirs, uint64_t n,
                           int32_t* ids, float* ts, Counters* ctr, cudaStream_t s);
// read-bandwidth probe over a working set of ws_vec (power of two) uint4
void launch_debug_read_bw(const uint4* p, uint32_t ws_vec, uint64_t per_thread, uint32_t grid, uint32_t* sink,
                          cudaStream_t s);
void launch_debug_rng(const uint32_t* seeds, uint32_t n_seeds, uint32_t n_draws, float* out,
                      cudaStream_t s);
// toolkit cuRAND on the device, for the RNG known-answer test
void launch_debug_curand(const uint32_t* seeds, uint32_t n_seeds, uint32_t n_draws, float* out,
                         cudaStream_t s);

} // namespace ptc
