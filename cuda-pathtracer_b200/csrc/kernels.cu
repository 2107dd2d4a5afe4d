// kernels.cu -- the wavefront path-tracing kernels for sm_100a.
//
// The wavefront stages of the design (generate, extend, shade/sample, accumulate) are
// device functions (traverse.cuh, shade.cuh); what is LAUNCHED is one persistent kernel per
// wavefront step that runs extend + shade back to back on its rays and hands the survivors
// to the next step through a compacted structure-of-arrays ray queue:
//
//   k_path<FIRST>   generate (RNG seed, camera ray, DOF) for an 8x4 pixel tile per warp,
//                   then `n_iters` x [extend -> shade] in registers
//   k_path<!FIRST>  one warp per 32 queue slots: load ray, replay RNG, `n_iters` x
//                   [extend -> shade]
//   both            finished paths run the accumulate/tonemap/RGBA8 epilogue in place (each
//                   pixel finishes exactly once per frame); unfinished ones are appended to
//                   the output queue with one warp-aggregated atomic (ballot + popc + shfl)
//
// Persistent grid = resident CTAs per SM x 148 SMs, warps stride over the work items.
// No tensor cores (traversal is not a dense contraction), no RT cores (B200 has none).
#include "kernels.cuh"

#include <type_traits>

#include "shade.cuh"

namespace ptc {

__device__ __forceinline__ uint32_t
warp_sum(uint32_t v)
{
  return __reduce_add_sync(0xFFFFFFFFu, v);
}

// HITS: the closest triangle hit of every queue slot was already found by k_extend; only the
// sphere lights are tested here before shading (n_iters must be 1).
// Register budget of the path kernels = 65536 / (256 x PTC_PATH_BLOCKS).  The fused first stage
// wants 126 registers (2 CTAs per SM), but it is bound by latency (texel fetches, dependent
// traversal steps), not by registers: capped at 64 it spills a few words and runs 4 CTAs per SM --
// measured -3 % (indoor 1080p) to -10 % (color_sample 4K) frame time; 3 CTAs (80 registers) is
// within 0.5 % of that.
#ifndef PTC_PATH_BLOCKS
#define PTC_PATH_BLOCKS 4
#endif
// SMS: the scene's nodes + triangle records are copied into shared memory behind the stacks at
// kernel start and the traversal reads them with LDS (SharedSceneStack, traverse.cuh); chosen per
// scene by the host when PTC_PATH_BLOCKS CTAs still fit an SM that way.
template <bool FIRST, bool STATS, bool HITS, bool SMS = false>
__global__ void __launch_bounds__(kBlock, PTC_PATH_BLOCKS)
k_path(const FrameParams fp, const DScene sc, const cudaTextureObject_t cube, const RayQueue qin,
       const uint4* __restrict__ hits,
       const uint32_t* __restrict__ qin_count, const RayQueue qout, uint32_t* __restrict__ qout_count,
       const FrameOut fo, Counters* __restrict__ ctr, uint32_t* __restrict__ work, const int bounce0,
       const int n_iters, const int stack_cap, float4* __restrict__ done_q, uint32_t* __restrict__ done_ctl,
       const int done_mode)
{
#ifdef PTC_STAGE_LOCAL_STACK
  LocalStack stack;
#else
  extern __shared__ uint2 s_stack[]; // stack_cap entries x kBlock threads (none when HITS)
#ifdef PTC_TABLES
  __shared__ uint32_t s_spread[HITS ? 1 : kSpreadWords];
  __shared__ uint8_t s_perm[HITS ? 4 : kPermBytes];
  if (!HITS)
    trav_tables_init(s_perm, s_spread);
  typename std::conditional<SMS, SharedSceneStack, SharedStack>::type stack;
  stack.base = s_stack + threadIdx.x, stack.cap = stack_cap, stack.perm = s_perm, stack.spread = s_spread;
  if constexpr (SMS) {
    uint4* s_nodes = reinterpret_cast<uint4*>(s_stack + (size_t)stack_cap * kBlock);
    uint4* s_tris = s_nodes + sc.n_nodes * (kSmemNodeBytes / 16u);
    smem_scene_fill(sc, s_nodes, s_tris);
    stack.nodes_s = (uint32_t)__cvta_generic_to_shared(s_nodes);
    stack.tris_s = (uint32_t)__cvta_generic_to_shared(s_tris);
  }
#else
  SharedStack stack{ {}, s_stack + threadIdx.x, stack_cap };
#endif
#endif
  __shared__ uint32_t s_rays[16];
  __shared__ float s_cam[8]; // camera basis of the frame, evaluated once per CTA (FIRST)
  if (threadIdx.x < 16)
    s_rays[threadIdx.x] = 0;
#ifndef PTC_NO_CAM_HOIST
  if (FIRST && threadIdx.x == 32) {
    const CamBasis B = camera_basis(fp);
    s_cam[0] = B.screen_dist;
    s_cam[1] = B.cu.x, s_cam[2] = B.cu.y, s_cam[3] = B.cu.z;
    s_cam[4] = B.cv.x, s_cam[5] = B.cv.y, s_cam[6] = B.cv.z;
  }
#endif
  __syncthreads();

  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t tiles_y = (fp.part_rows + kTileH - 1) / kTileH;
  const uint32_t n_items = FIRST ? fp.tiles_x * tiles_y * 32u : *qin_count;

  uint32_t my_segs = 0, my_traced = 0, my_nodes = 0, my_tris = 0;

  // persistent warps pull 32 work items at a time from one counter (no tail imbalance)
  for (;;) {
    uint32_t base = 0;
    if (lane == 0)
      base = atomicAdd(work, 32u);
    base = __shfl_sync(0xFFFFFFFFu, base, 0);
    if (base >= n_items)
      break;
    const uint32_t i = base + lane;
    bool valid;
    uint32_t x = 0, lrow = 0;
    Path P;
    Rng rng;
    if (FIRST) {
      const uint32_t wi = base >> 5;
      x = (wi % fp.tiles_x) * kTileW + (lane & (kTileW - 1));
      lrow = fp.part_row0 + (wi / fp.tiles_x) * kTileH + (lane / kTileW);
      valid = x < fp.width && lrow < fp.part_row0 + fp.part_rows;
      if (valid) {
#ifndef PTC_NO_CAM_HOIST
        CamBasis B;
        B.screen_dist = s_cam[0];
        B.cu = mk3(s_cam[1], s_cam[2], s_cam[3]);
        B.cv = mk3(s_cam[4], s_cam[5], s_cam[6]);
        primary_ray(fp, B, x, global_row(fp, lrow), rng, P.origin, P.dir);
#else
        primary_ray(fp, x, global_row(fp, lrow), rng, P.origin, P.dir);
#endif
        P.throughput = mk3(1.0f, 1.0f, 1.0f);
        P.acc = mk3(0.0f, 0.0f, 0.0f);
        P.spec_carry = 0.0f; // IntersectionData zero-initialised at path start (Q3)
        P.draws = 2;
      }
    } else {
      valid = i < n_items;
      if (valid) {
        // queue traffic is streamed (evict-first): it must not push the texels out of L2
        const float4 o4 = __ldcs(qin.o + i), d4 = __ldcs(qin.d + i), t4 = __ldcs(qin.thr + i), a4 = __ldcs(qin.acc + i);
        P.origin = mk3(o4.x, o4.y, o4.z);
        P.dir = mk3(d4.x, d4.y, d4.z);
        P.throughput = mk3(t4.x, t4.y, t4.z);
        P.acc = mk3(a4.x, a4.y, a4.z);
        P.spec_carry = o4.w;
        P.draws = __float_as_uint(d4.w);
        const uint32_t p = __float_as_uint(t4.w);
        x = p % fp.width;
        lrow = p / fp.width;
        rng_init(rng, fp.hash_seed + reference_tid(x, global_row(fp, lrow), fp.width));
        rng_skip(rng, P.draws);
      }
    }

    bool done = !valid;
    // preview frame (camera moved, raytrace.cu:54-62): one ray, flat colour -- runs through the same
    // loop (frame_schedule gives it one iteration) so that the kernel holds ONE traversal instance
    const bool preview = FIRST && !fp.is_static;
    for (int it = 0; it < n_iters; ++it) {
      const uint32_t active = __ballot_sync(0xFFFFFFFFu, !done);
      if (!active)
        break;
      const int b = bounce0 + it;
      if (lane == 0)
        atomicAdd(&s_rays[b < 16 ? b : 15], (uint32_t)__popc(active));
      if (!done) {
        float t, u = 0.f, v = 0.f;
        uint32_t id, segs;
        const bool last = !preview && b + 1 >= fp.max_bounces; // nothing after this iteration: any-hit query
        if (HITS) {
          const uint4 h = __ldcs(hits + i);
          t = __uint_as_float(h.x), id = h.y, u = __uint_as_float(h.z), v = __uint_as_float(h.w);
          if (!last)
            lights_hit(sc, P.origin, P.dir, t, id);
        } else {
          trace_ray<STATS>(sc, stack, P.origin, P.dir, last, t, id, u, v, my_nodes, my_tris);
        }
        if (preview) {
          if (id != 0xFFFFFFFFu) {
            Surface s;
            resolve_hit(sc, id, t, u, v, P.dir, P.spec_carry, s);
            P.acc = s.diffuse_col;
          } else {
            P.acc = env_lookup(cube, P.dir);
          }
          done = true;
          segs = 1;
        } else if (last && id < sc.n_faces) {
          done = true; // a triangle in front on the last iteration adds no radiance
          segs = 1;
        } else {
          done = shade_step(fp, sc, cube, P, rng, t, id, u, v, b, segs);
        }
        my_segs += segs;
        my_traced++;
      }
    }

    // finished paths: accumulate + tonemap + RGBA8 -- in place in the frame's LAST stage, where every
    // path of the warp finishes together.  In an earlier stage only some lanes of a warp finish (on the
    // indoor view 44 % of the paths leave through a window at their second segment), and the ~400
    // instructions of the epilogue (three powf, six IEEE divisions) would run for them alone: those
    // pixels are parked as (radiance, pixel) and the last stage runs their epilogue 32 at a time.
    const bool fin = valid && done;
    if (done_mode & kDoneDefer) {
      const uint32_t fm = __ballot_sync(0xFFFFFFFFu, fin);
      if (fm) {
        uint32_t slot = 0;
        const uint32_t leader = __ffs(fm) - 1u;
        if (lane == leader)
          slot = atomicAdd(done_ctl, (uint32_t)__popc(fm));
        slot = __shfl_sync(0xFFFFFFFFu, slot, leader) + __popc(fm & ((1u << lane) - 1u));
        if (fin)
          __stcs(done_q + slot, make_float4(P.acc.x, P.acc.y, P.acc.z, __uint_as_float(lrow * fp.width + x)));
      }
    } else if (fin) {
      epilogue(fp, fo, x, lrow, P.acc);
    }

    // survivors: warp-aggregated append to the next queue
    const bool cont = valid && !done;
    const uint32_t mask = __ballot_sync(0xFFFFFFFFu, cont);
    if (mask) {
      uint32_t slot = 0;
      const uint32_t leader = __ffs(mask) - 1u;
      if (lane == leader)
        slot = atomicAdd(qout_count, (uint32_t)__popc(mask));
      slot = __shfl_sync(0xFFFFFFFFu, slot, leader) + __popc(mask & ((1u << lane) - 1u));
      if (cont) {
        __stcs(qout.o + slot, make_float4(P.origin.x, P.origin.y, P.origin.z, P.spec_carry));
        __stcs(qout.d + slot, make_float4(P.dir.x, P.dir.y, P.dir.z, __uint_as_float(P.draws)));
        __stcs(qout.thr + slot, make_float4(P.throughput.x, P.throughput.y, P.throughput.z,
                                            __uint_as_float(lrow * fp.width + x)));
        __stcs(qout.acc + slot, make_float4(P.acc.x, P.acc.y, P.acc.z, 0.0f));
      }
    }
  }

  // last stage: the pixels earlier stages parked (their count is final: stream order)
  if (done_mode & kDoneDrain) {
    const uint32_t n_done = done_ctl[0];
    for (;;) {
      uint32_t base = 0;
      if (lane == 0)
        base = atomicAdd(done_ctl + 1, 32u);
      base = __shfl_sync(0xFFFFFFFFu, base, 0);
      if (base >= n_done)
        break;
      if (base + lane < n_done) {
        const float4 e = __ldcs(done_q + base + lane);
        const uint32_t p = __float_as_uint(e.w);
        epilogue(fp, fo, p % fp.width, p / fp.width, mk3(e.x, e.y, e.z));
      }
    }
  }

  // ray counters: one set of atomics per warp / per CTA for the whole launch
  const uint32_t wsegs = warp_sum(my_segs), wtraced = warp_sum(my_traced);
  if (lane == 0 && wtraced) {
    atomicAdd(&ctr->segments, (unsigned long long)wsegs);
    atomicAdd(&ctr->traced, (unsigned long long)wtraced);
  }
  if (STATS) {
    const uint32_t wn = warp_sum(my_nodes), wt = warp_sum(my_tris);
    if (lane == 0 && (wn | wt)) {
      atomicAdd(&ctr->nodes_visited, (unsigned long long)wn);
      atomicAdd(&ctr->tris_tested, (unsigned long long)wt);
    }
  }
  __syncthreads();
  if (threadIdx.x < 16 && s_rays[threadIdx.x])
    atomicAdd(&ctr->rays_per_bounce[threadIdx.x], (unsigned long long)s_rays[threadIdx.x]);
}

// ---------------------------------------------------------------------------
// k_wave -- the stage of the SHIFTED schedule: the stage boundary sits between a path's
// intersection and its shading, not after the shading.
//
//   k_wave<FIRST>   generate -> intersect (iteration 0)
//   k_wave<!FIRST>  load ray + hit of iteration b-1 -> shade it -> intersect the new ray (iteration b)
//   both            a ray that MISSES is finished at once (environment + the arithmetic tail of the
//                   reference's loop, shade_step); only rays that HIT something are queued, with
//                   their hit record, for the next stage
//
// Why: with the boundary after the shading (k_path), every stage shades the rays it traced -- hits
// and misses side by side in a warp.  On the indoor view 44 % of the second segments leave through
// a window, so the ~600-instruction material/BRDF part ran with 56 % of a warp's lanes.  Here the
// queue holds hits only: the shading of a stage is fully populated, and the misses never occupy a
// queue slot.  The last iteration (any-hit query, lights and environment only) ends in the stage
// that traced it.  Same arithmetic per path, so the same bits as k_path.
// ---------------------------------------------------------------------------
template <bool FIRST, bool STATS>
__global__ void __launch_bounds__(kBlock, PTC_PATH_BLOCKS)
k_wave(const FrameParams fp, const DScene sc, const cudaTextureObject_t cube, const RayQueue qin,
       const uint4* __restrict__ hin, const uint32_t* __restrict__ qin_count, const RayQueue qout,
       uint4* __restrict__ hout, uint32_t* __restrict__ qout_count, const FrameOut fo,
       Counters* __restrict__ ctr, uint32_t* __restrict__ work, const int b, const int stack_cap,
       float4* __restrict__ done_q, uint32_t* __restrict__ done_ctl, const int done_mode)
{
  extern __shared__ uint2 s_stack[]; // stack_cap entries x kBlock threads
#ifdef PTC_TABLES
  __shared__ uint32_t s_spread[kSpreadWords];
  __shared__ uint8_t s_perm[kPermBytes];
  trav_tables_init(s_perm, s_spread);
  SharedStack stack{ {}, s_stack + threadIdx.x, stack_cap, s_perm, s_spread };
#else
  SharedStack stack{ {}, s_stack + threadIdx.x, stack_cap };
#endif
  __shared__ uint32_t s_nrays;
  __shared__ float s_cam[8];
  if (threadIdx.x == 0)
    s_nrays = 0;
  if (FIRST && threadIdx.x == 32) {
    const CamBasis B = camera_basis(fp);
    s_cam[0] = B.screen_dist;
    s_cam[1] = B.cu.x, s_cam[2] = B.cu.y, s_cam[3] = B.cu.z;
    s_cam[4] = B.cv.x, s_cam[5] = B.cv.y, s_cam[6] = B.cv.z;
  }
  __syncthreads();

  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t tiles_y = (fp.part_rows + kTileH - 1) / kTileH;
  const uint32_t n_items = FIRST ? fp.tiles_x * tiles_y * 32u : *qin_count;
  const bool last = b + 1 >= fp.max_bounces; // nothing after this iteration: any-hit query

  uint32_t my_segs = 0, my_traced = 0, my_nodes = 0, my_tris = 0;

  for (;;) {
    uint32_t base = 0;
    if (lane == 0)
      base = atomicAdd(work, 32u);
    base = __shfl_sync(0xFFFFFFFFu, base, 0);
    if (base >= n_items)
      break;
    const uint32_t i = base + lane;
    bool valid;
    uint32_t x = 0, lrow = 0;
    Path P;
    Rng rng;
    bool done;
    if (FIRST) {
      const uint32_t wi = base >> 5;
      x = (wi % fp.tiles_x) * kTileW + (lane & (kTileW - 1));
      lrow = fp.part_row0 + (wi / fp.tiles_x) * kTileH + (lane / kTileW);
      valid = x < fp.width && lrow < fp.part_row0 + fp.part_rows;
      if (valid) {
        CamBasis B;
        B.screen_dist = s_cam[0];
        B.cu = mk3(s_cam[1], s_cam[2], s_cam[3]);
        B.cv = mk3(s_cam[4], s_cam[5], s_cam[6]);
        primary_ray(fp, B, x, global_row(fp, lrow), rng, P.origin, P.dir);
        P.throughput = mk3(1.0f, 1.0f, 1.0f);
        P.acc = mk3(0.0f, 0.0f, 0.0f);
        P.spec_carry = 0.0f; // IntersectionData zero-initialised at path start (Q3)
        P.draws = 2;
      }
      done = !valid;
    } else {
      valid = i < n_items;
      done = true;
      if (valid) {
        const float4 o4 = __ldcs(qin.o + i), d4 = __ldcs(qin.d + i), t4 = __ldcs(qin.thr + i), a4 = __ldcs(qin.acc + i);
        const uint4 h = __ldcs(hin + i);
        P.origin = mk3(o4.x, o4.y, o4.z);
        P.dir = mk3(d4.x, d4.y, d4.z);
        P.throughput = mk3(t4.x, t4.y, t4.z);
        P.acc = mk3(a4.x, a4.y, a4.z);
        P.spec_carry = o4.w;
        P.draws = __float_as_uint(d4.w);
        const uint32_t p = __float_as_uint(t4.w);
        x = p % fp.width;
        lrow = p / fp.width;
        rng_init(rng, fp.hash_seed + reference_tid(x, global_row(fp, lrow), fp.width));
        rng_skip(rng, P.draws);
        // iteration b-1: every queued ray hit a face or a light
        uint32_t segs;
        done = shade_step(fp, sc, cube, P, rng, __uint_as_float(h.x), h.y, __uint_as_float(h.z),
                          __uint_as_float(h.w), b - 1, segs);
        my_segs += segs;
      }
    }

    // iteration b: intersect; a miss (and the whole last iteration) is finished here
    const uint32_t active = __ballot_sync(0xFFFFFFFFu, !done);
    if (lane == 0 && active)
      atomicAdd(&s_nrays, (uint32_t)__popc(active));
    float t = 0.f, u = 0.f, v = 0.f;
    uint32_t id = 0xFFFFFFFFu;
    if (!done) {
      my_traced++;
      trace_ray<STATS>(sc, stack, P.origin, P.dir, last, t, id, u, v, my_nodes, my_tris);
      if (last && id < sc.n_faces) {
        done = true; // a triangle in front on the last iteration adds no radiance
        my_segs++;
      } else if (last || id == 0xFFFFFFFFu) {
        uint32_t segs;
        done = shade_step(fp, sc, cube, P, rng, t, id, u, v, b, segs); // always finishes the path
        my_segs += segs;
      }
    }

    const bool fin = valid && done;
    if (done_mode & kDoneDefer) {
      const uint32_t fm = __ballot_sync(0xFFFFFFFFu, fin);
      if (fm) {
        uint32_t slot = 0;
        const uint32_t leader = __ffs(fm) - 1u;
        if (lane == leader)
          slot = atomicAdd(done_ctl, (uint32_t)__popc(fm));
        slot = __shfl_sync(0xFFFFFFFFu, slot, leader) + __popc(fm & ((1u << lane) - 1u));
        if (fin)
          __stcs(done_q + slot, make_float4(P.acc.x, P.acc.y, P.acc.z, __uint_as_float(lrow * fp.width + x)));
      }
    } else if (fin) {
      epilogue(fp, fo, x, lrow, P.acc);
    }

    // rays that hit: warp-aggregated append of ray + hit record to the next stage's queue
    const bool cont = valid && !done;
    const uint32_t mask = __ballot_sync(0xFFFFFFFFu, cont);
    if (mask) {
      uint32_t slot = 0;
      const uint32_t leader = __ffs(mask) - 1u;
      if (lane == leader)
        slot = atomicAdd(qout_count, (uint32_t)__popc(mask));
      slot = __shfl_sync(0xFFFFFFFFu, slot, leader) + __popc(mask & ((1u << lane) - 1u));
      if (cont) {
        __stcs(qout.o + slot, make_float4(P.origin.x, P.origin.y, P.origin.z, P.spec_carry));
        __stcs(qout.d + slot, make_float4(P.dir.x, P.dir.y, P.dir.z, __uint_as_float(P.draws)));
        __stcs(qout.thr + slot, make_float4(P.throughput.x, P.throughput.y, P.throughput.z,
                                            __uint_as_float(lrow * fp.width + x)));
        __stcs(qout.acc + slot, make_float4(P.acc.x, P.acc.y, P.acc.z, 0.0f));
        __stcs(hout + slot, make_uint4(__float_as_uint(t), id, __float_as_uint(u), __float_as_uint(v)));
      }
    }
  }

  if (done_mode & kDoneDrain) {
    const uint32_t n_done = done_ctl[0];
    for (;;) {
      uint32_t base = 0;
      if (lane == 0)
        base = atomicAdd(done_ctl + 1, 32u);
      base = __shfl_sync(0xFFFFFFFFu, base, 0);
      if (base >= n_done)
        break;
      if (base + lane < n_done) {
        const float4 e = __ldcs(done_q + base + lane);
        const uint32_t p = __float_as_uint(e.w);
        epilogue(fp, fo, p % fp.width, p / fp.width, mk3(e.x, e.y, e.z));
      }
    }
  }

  const uint32_t wsegs = warp_sum(my_segs), wtraced = warp_sum(my_traced);
  if (lane == 0 && (wsegs | wtraced)) {
    atomicAdd(&ctr->segments, (unsigned long long)wsegs);
    atomicAdd(&ctr->traced, (unsigned long long)wtraced);
  }
  if (STATS) {
    const uint32_t wn = warp_sum(my_nodes), wt = warp_sum(my_tris);
    if (lane == 0 && (wn | wt)) {
      atomicAdd(&ctr->nodes_visited, (unsigned long long)wn);
      atomicAdd(&ctr->tris_tested, (unsigned long long)wt);
    }
  }
  __syncthreads();
  if (threadIdx.x == 0 && s_nrays)
    atomicAdd(&ctr->rays_per_bounce[b < 16 ? b : 15], (unsigned long long)s_nrays);
}

// ---------------------------------------------------------------------------
// extend with per-lane refill: incoherent rays need very different numbers of node and
// triangle tests, so a warp that traverses 32 fixed rays runs at the pace of its slowest
// lane (measured: 13 of 32 lanes active).  Here a lane that finishes its ray stores the
// hit and, as soon as `refill` lanes of the warp are idle, the idle lanes pull new rays
// from the queue (one warp-aggregated atomic), so the traversal loop stays full.
// ---------------------------------------------------------------------------
// (A 4-CTAs-per-SM build at 64 registers was measured: no gain on either the 28 KB or the
// 1.4 GB scene -- the limit is issue slots / L1 wavefronts, not warps in flight.)
#ifndef PTC_EXTEND_BLOCKS
#define PTC_EXTEND_BLOCKS 3
#endif
// batched triangle loads + node-line prefetch by every lane (traverse.cuh TB / PF = 2): -5 % frame
// time on the 10 M-triangle scene, -2 % at 180 k triangles (profiles/r2_experiment_log.md)
#ifndef PTC_TRI_BATCH
#define PTC_TRI_BATCH 1
#endif
template <bool STATS>
__global__ void __launch_bounds__(kBlock, PTC_EXTEND_BLOCKS)
k_extend(const DScene sc, const RayQueue qin, const uint32_t* __restrict__ qin_count,
         uint4* __restrict__ hits, uint32_t* __restrict__ work, Counters* __restrict__ ctr,
         const uint32_t refill, const int last, const int stack_cap)
{
#ifdef PTC_STAGE_LOCAL_STACK
  LocalStack stack;
#else
  extern __shared__ uint2 s_stack[]; // stack_cap entries x kBlock threads
#ifdef PTC_TABLES
  __shared__ uint32_t s_spread[kSpreadWords];
  __shared__ uint8_t s_perm[kPermBytes];
  trav_tables_init(s_perm, s_spread);
  __syncthreads();
  SharedStack stack{ {}, s_stack + threadIdx.x, stack_cap, s_perm, s_spread };
#else
  SharedStack stack{ {}, s_stack + threadIdx.x, stack_cap };
#endif
#endif
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t n = *qin_count;
  Trav T;
  bool have = false, exhausted = false;
  uint32_t slot = 0, my_nodes = 0, my_tris = 0;
  unsigned long long st_iters = 0, st_have = 0, st_node = 0, st_tri1 = 0, st_tri2 = 0;

  for (;;) {
    const uint32_t idle = __ballot_sync(0xFFFFFFFFu, !have);
    if (idle == 0xFFFFFFFFu || (!exhausted && (uint32_t)__popc(idle) >= refill)) {
      if (!exhausted) {
        const uint32_t want = (uint32_t)__popc(idle);
        uint32_t base = 0;
        if (lane == 0)
          base = atomicAdd(work, want);
        base = __shfl_sync(0xFFFFFFFFu, base, 0);
        if (!have) {
          const uint32_t mine = base + __popc(idle & ((1u << lane) - 1u));
          if (mine < n) {
            const float4 o4 = __ldcs(qin.o + mine), d4 = __ldcs(qin.d + mine);
            if (last) // last radiance-loop iteration: lights first, then an any-hit query
              trav_init_any(sc, T, mk3(o4.x, o4.y, o4.z), mk3(d4.x, d4.y, d4.z));
            else
              trav_init(T, mk3(o4.x, o4.y, o4.z), mk3(d4.x, d4.y, d4.z));
            slot = mine;
            have = true;
          }
        }
        exhausted = base + want >= n;
      }
      if (!__any_sync(0xFFFFFFFFu, have))
        break;
    }
    const uint32_t n_before = my_nodes, t_before = my_tris;
    const bool had = have;
    if (have && trav_step<STATS, decltype(stack), true, PTC_PREFETCH_NODE, PTC_TRI_BATCH != 0>(sc, T, stack, my_nodes, my_tris)) {
      __stcs(hits + slot, make_uint4(__float_as_uint(T.best_t), T.best_id, __float_as_uint(T.best_u),
                                     __float_as_uint(T.best_v)));
      have = false;
    }
    if (STATS) { // where the lanes of a warp iteration go (lane 0 keeps the sums)
      const uint32_t mh = __ballot_sync(0xFFFFFFFFu, had), mn = __ballot_sync(0xFFFFFFFFu, my_nodes != n_before);
      const uint32_t m1 = __ballot_sync(0xFFFFFFFFu, my_tris != t_before);
      const uint32_t m2 = __ballot_sync(0xFFFFFFFFu, my_tris - t_before > 1u);
      st_iters++, st_have += __popc(mh), st_node += __popc(mn), st_tri1 += __popc(m1), st_tri2 += __popc(m2);
    }
  }
  if (STATS) {
    const uint32_t wn = warp_sum(my_nodes), wt = warp_sum(my_tris);
    if (lane == 0 && (wn | wt)) {
      atomicAdd(&ctr->nodes_visited, (unsigned long long)wn);
      atomicAdd(&ctr->tris_tested, (unsigned long long)wt);
    }
    if (lane == 0 && st_iters) {
      atomicAdd(&ctr->ext_iters, st_iters);
      atomicAdd(&ctr->ext_have, st_have);
      atomicAdd(&ctr->ext_node, st_node);
      atomicAdd(&ctr->ext_tri1, st_tri1);
      atomicAdd(&ctr->ext_tri2, st_tri2);
    }
  }
}

// ---------------------------------------------------------------------------
// parity / debug kernels
// ---------------------------------------------------------------------------

// Primary rays of a frame and their closest hits, one thread per pixel (row 0 = top).
__global__ void __launch_bounds__(128)
k_primary_hits(FrameParams fp, DScene sc, int32_t* __restrict__ ids, float* __restrict__ ts)
{
  const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= fp.band_rows * fp.width)
    return;
  const uint32_t x = p % fp.width, lrow = p / fp.width;
  Rng rng;
  float3 o, d;
  primary_ray(fp, x, global_row(fp, lrow), rng, o, d);
  float t, u, v;
  uint32_t id, nn = 0, nt = 0;
  LocalStack stack;
  closest_hit<false>(sc, stack, o, d, t, id, u, v, nn, nt);
  if (ids)
    ids[p] = (int32_t)id;
  if (ts)
    ts[p] = t;
}

// Arbitrary rays in, hit ids out, with traversal statistics.
__global__ void __launch_bounds__(128)
k_intersect_rays(DScene sc, const float* __restrict__ origins, const float* __restrict__ dirs,
                 uint64_t n, int32_t* __restrict__ ids, float* __restrict__ ts, Counters* ctr)
{
  const uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
  if (i >= n)
    return;
  float t, u, v;
  uint32_t id, nn = 0, nt = 0;
  LocalStack stack;
  closest_hit<true>(sc, stack, mk3(origins[3 * i], origins[3 * i + 1], origins[3 * i + 2]),
                    mk3(dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2]), t, id, u, v, nn, nt);
  if (ids)
    ids[i] = (int32_t)id;
  if (ts)
    ts[i] = t;
  if (ctr) {
    atomicAdd(&ctr->nodes_visited, (unsigned long long)nn);
    atomicAdd(&ctr->tris_tested, (unsigned long long)nt);
  }
}

__global__ void
k_debug_rng(const uint32_t* __restrict__ seeds, uint32_t n_seeds, uint32_t n_draws,
            float* __restrict__ out)
{
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_seeds)
    return;
  // second half of the draws goes through rng_skip, the replay path the queue stages use
  Rng s;
  rng_init(s, seeds[i]);
  const uint32_t half = n_draws / 2;
  for (uint32_t k = 0; k < half; ++k)
    out[(size_t)i * n_draws + k] = rng_uniform(s);
  Rng r;
  rng_init(r, seeds[i]);
  rng_skip(r, half);
  for (uint32_t k = half; k < n_draws; ++k)
    out[(size_t)i * n_draws + k] = rng_uniform(r);
}

} // namespace ptc

// ---------------------------------------------------------------------------
// launchers
// ---------------------------------------------------------------------------
#include <curand_kernel.h>

namespace ptc {

static __global__ void
k_debug_curand(const uint32_t* __restrict__ seeds, uint32_t n_seeds, uint32_t n_draws,
               float* __restrict__ out)
{
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_seeds)
    return;
  curandState st;
  curand_init(seeds[i], 0, 0, &st); // raytrace.cu:234-235
  for (uint32_t k = 0; k < n_draws; ++k)
    out[(size_t)i * n_draws + k] = curand_uniform(&st);
}

static inline uint32_t
blocks_for(uint64_t n, uint32_t block)
{
  return (uint32_t)((n + block - 1) / block);
}

size_t
stack_smem_bytes(int stack_cap)
{
#ifdef PTC_STAGE_LOCAL_STACK
  return 0;
#else
  return (size_t)stack_cap * kBlock * sizeof(uint2);
#endif
}

// every scene may need a different amount: opt in to the device maximum once (the attribute is
// per function, not per launch, and must not shrink when a smaller scene is uploaded later)
static int
max_optin_smem()
{
  int dev = 0, v = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&v, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
  return v - 128 - kPermBytes - 4 * kSpreadWords; // static shared memory of the stage kernels: counters, camera basis, tables
}

template <bool FIRST, bool STATS, bool HITS, bool SMS = false>
static int
occ(size_t smem)
{
  int n = 0;
  cudaFuncSetAttribute(k_path<FIRST, STATS, HITS, SMS>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_optin_smem());
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_path<FIRST, STATS, HITS, SMS>, kBlock, smem) != cudaSuccess)
    n = 0;
  return n;
}

size_t
scene_smem_bytes(const DScene& sc)
{
  return (size_t)sc.n_nodes * kSmemNodeBytes + (size_t)sc.n_tris * kSmemTriBytes;
}

int
path_kernel_blocks_per_sm_scene(bool first, bool stats, int stack_cap, const DScene& sc)
{
#if defined(PTC_TABLES) && !defined(PTC_STAGE_LOCAL_STACK) && !defined(PTC_NO_SMEM_SCENE)
  const size_t smem = stack_smem_bytes(stack_cap) + scene_smem_bytes(sc);
  if (smem > (size_t)max_optin_smem())
    return 0;
  if (first)
    return stats ? occ<true, true, false, true>(smem) : occ<true, false, false, true>(smem);
  return stats ? occ<false, true, false, true>(smem) : occ<false, false, false, true>(smem);
#else
  return 0;
#endif
}

// Resident CTAs per SM for a scene whose traversal stack needs `stack_cap` entries per thread
// (0 when the stack does not fit the SM's shared memory at all).  Also opts the kernel in to
// that much dynamic shared memory.
int
path_kernel_blocks_per_sm(bool first, bool stats, bool pretraced, int stack_cap)
{
  const size_t smem = stack_smem_bytes(stack_cap);
  if (first)
    return stats ? occ<true, true, false>(smem) : occ<true, false, false>(smem);
  if (pretraced) // shade-only stage: no traversal, no stack
    return stats ? occ<false, true, true>(0) : occ<false, false, true>(0);
  return stats ? occ<false, true, false>(smem) : occ<false, false, false>(smem);
}

template <bool FIRST, bool STATS>
static int
occ_wave(size_t smem)
{
  int n = 0;
  cudaFuncSetAttribute(k_wave<FIRST, STATS>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_optin_smem());
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_wave<FIRST, STATS>, kBlock, smem) != cudaSuccess)
    n = 0;
  return n;
}

int
wave_kernel_blocks_per_sm(bool first, bool stats, int stack_cap)
{
  const size_t smem = stack_smem_bytes(stack_cap);
  if (first)
    return stats ? occ_wave<true, true>(smem) : occ_wave<true, false>(smem);
  return stats ? occ_wave<false, true>(smem) : occ_wave<false, false>(smem);
}

int
extend_kernel_blocks_per_sm(bool stats, int stack_cap)
{
  const size_t smem = stack_smem_bytes(stack_cap);
  int n = 0;
  cudaError_t e;
  if (stats) {
    cudaFuncSetAttribute(k_extend<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_optin_smem());
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_extend<true>, kBlock, smem);
  } else {
    cudaFuncSetAttribute(k_extend<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_optin_smem());
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_extend<false>, kBlock, smem);
  }
  return e != cudaSuccess ? 0 : n;
}

#define PTC_LAUNCH_PATH(F, S, H, SMEM)                                                             \
  k_path<F, S, H><<<grid, kBlock, SMEM, s>>>(fp, sc, cube, qin, hits, qin_count, qout, qout_count, fo, ctr, \
                                             work, bounce0, n_iters, stack_cap, done_q, done_ctl, done_mode)
#define PTC_LAUNCH_PATH_SMS(F, S, SMEM)                                                            \
  k_path<F, S, false, true><<<grid, kBlock, SMEM, s>>>(fp, sc, cube, qin, hits, qin_count, qout, qout_count, fo, \
                                                       ctr, work, bounce0, n_iters, stack_cap, done_q, done_ctl,  \
                                                       done_mode)

void
launch_path(bool first, bool stats, uint32_t grid, const FrameParams& fp, const DScene& sc,
            cudaTextureObject_t cube, const RayQueue& qin, const uint4* hits, const uint32_t* qin_count,
            const RayQueue& qout, uint32_t* qout_count, const FrameOut& fo, Counters* ctr,
            uint32_t* work, int bounce0, int n_iters, int stack_cap, float4* done_q, uint32_t* done_ctl,
            int done_mode, cudaStream_t s, bool scene_in_smem)
{
  if (!grid)
    return;
  const size_t smem = stack_smem_bytes(stack_cap);
#if defined(PTC_TABLES) && !defined(PTC_STAGE_LOCAL_STACK) && !defined(PTC_NO_SMEM_SCENE)
  if (scene_in_smem && !hits) {
    const size_t both = smem + scene_smem_bytes(sc);
    if (first) {
      if (stats)
        PTC_LAUNCH_PATH_SMS(true, true, both);
      else
        PTC_LAUNCH_PATH_SMS(true, false, both);
    } else {
      if (stats)
        PTC_LAUNCH_PATH_SMS(false, true, both);
      else
        PTC_LAUNCH_PATH_SMS(false, false, both);
    }
    return;
  }
#endif
  if (first) {
    if (stats)
      PTC_LAUNCH_PATH(true, true, false, smem);
    else
      PTC_LAUNCH_PATH(true, false, false, smem);
  } else if (hits) {
    if (stats)
      PTC_LAUNCH_PATH(false, true, true, 0);
    else
      PTC_LAUNCH_PATH(false, false, true, 0);
  } else {
    if (stats)
      PTC_LAUNCH_PATH(false, true, false, smem);
    else
      PTC_LAUNCH_PATH(false, false, false, smem);
  }
}

#define PTC_LAUNCH_WAVE(F, S)                                                                      \
  k_wave<F, S><<<grid, kBlock, smem, s>>>(fp, sc, cube, qin, hin, qin_count, qout, hout, qout_count, fo, ctr, \
                                          work, bounce, stack_cap, done_q, done_ctl, done_mode)

void
launch_wave(bool first, bool stats, uint32_t grid, const FrameParams& fp, const DScene& sc,
            cudaTextureObject_t cube, const RayQueue& qin, const uint4* hin, const uint32_t* qin_count,
            const RayQueue& qout, uint4* hout, uint32_t* qout_count, const FrameOut& fo, Counters* ctr,
            uint32_t* work, int bounce, int stack_cap, float4* done_q, uint32_t* done_ctl, int done_mode,
            cudaStream_t s)
{
  if (!grid)
    return;
  const size_t smem = stack_smem_bytes(stack_cap);
  if (first) {
    if (stats)
      PTC_LAUNCH_WAVE(true, true);
    else
      PTC_LAUNCH_WAVE(true, false);
  } else {
    if (stats)
      PTC_LAUNCH_WAVE(false, true);
    else
      PTC_LAUNCH_WAVE(false, false);
  }
}

void
launch_extend(bool stats, uint32_t grid, const DScene& sc, const RayQueue& qin, const uint32_t* qin_count,
              uint4* hits, uint32_t* work, Counters* ctr, uint32_t refill, bool last, int stack_cap,
              cudaStream_t s)
{
  if (!grid)
    return;
  const size_t smem = stack_smem_bytes(stack_cap);
  const int l = last ? 1 : 0;
  if (stats)
    k_extend<true><<<grid, kBlock, smem, s>>>(sc, qin, qin_count, hits, work, ctr, refill, l, stack_cap);
  else
    k_extend<false><<<grid, kBlock, smem, s>>>(sc, qin, qin_count, hits, work, ctr, refill, l, stack_cap);
}

void
launch_primary_hits(const FrameParams& fp, const DScene& sc, int32_t* ids, float* ts, cudaStream_t s)
{
  const uint32_t n = fp.band_rows * fp.width;
  if (n)
    k_primary_hits<<<blocks_for(n, 128), 128, 0, s>>>(fp, sc, ids, ts);
}

void
launch_intersect_rays(const DScene& sc, const float* origins, const float* dirs, uint64_t n,
                      int32_t* ids, float* ts, Counters* ctr, cudaStream_t s)
{
  if (n)
    k_intersect_rays<<<blocks_for(n, 128), 128, 0, s>>>(sc, origins, dirs, n, ids, ts, ctr);
}

// Read-bandwidth probe: every thread streams 16-byte read-only loads (the traversal's load
// path) over a working set of ws_vec uint4 (power of two), wrapping around, four loads in
// flight.  A working set that fits L1 / L2 / neither measures that tier.
__global__ void __launch_bounds__(256)
k_debug_read_bw(const uint4* __restrict__ p, uint32_t ws_mask, uint64_t per_thread, uint32_t* sink)
{
  const uint32_t stride = gridDim.x * blockDim.x;
  uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  uint4 a = make_uint4(0, 0, 0, 0);
  for (uint64_t i = 0; i < per_thread; i += 4) {
    const uint4 v0 = __ldg(p + (k & ws_mask));
    const uint4 v1 = __ldg(p + ((k + stride) & ws_mask));
    const uint4 v2 = __ldg(p + ((k + 2u * stride) & ws_mask));
    const uint4 v3 = __ldg(p + ((k + 3u * stride) & ws_mask));
    k += 4u * stride;
    a.x ^= v0.x ^ v1.x ^ v2.x ^ v3.x, a.y ^= v0.y ^ v1.y ^ v2.y ^ v3.y;
    a.z ^= v0.z ^ v1.z ^ v2.z ^ v3.z, a.w ^= v0.w ^ v1.w ^ v2.w ^ v3.w;
  }
  if ((a.x ^ a.y ^ a.z ^ a.w) == 0x9E3779B9u) // never true for a zeroed buffer; keeps the loads
    *sink = a.x;
}

void
launch_debug_read_bw(const uint4* p, uint32_t ws_vec, uint64_t per_thread, uint32_t grid, uint32_t* sink,
                     cudaStream_t s)
{
  k_debug_read_bw<<<grid, 256, 0, s>>>(p, ws_vec - 1u, per_thread, sink);
}

void
launch_debug_rng(const uint32_t* seeds, uint32_t n_seeds, uint32_t n_draws, float* out, cudaStream_t s)
{
  if (n_seeds)
    k_debug_rng<<<blocks_for(n_seeds, 64), 64, 0, s>>>(seeds, n_seeds, n_draws, out);
}

void
launch_debug_curand(const uint32_t* seeds, uint32_t n_seeds, uint32_t n_draws, float* out,
                    cudaStream_t s)
{
  if (n_seeds)
    k_debug_curand<<<blocks_for(n_seeds, 64), 64, 0, s>>>(seeds, n_seeds, n_draws, out);
}

} // namespace ptc
