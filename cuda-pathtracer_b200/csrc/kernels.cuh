// kernels.cuh -- host-callable launchers of the wavefront kernels (kernels.cu).
#pragma once

#include "device_types.cuh"

namespace ptc {

void launch_generate(const FrameParams& fp, const PathState& st, cudaStream_t s);
void launch_extend(const DScene& sc, const PathState& st, const uint32_t* queue,
                   const uint32_t* queue_count, uint32_t n_max, Counters* ctr, int bounce,
                   cudaStream_t s);
void launch_shade(const FrameParams& fp, const DScene& sc, const PathState& st,
                  cudaTextureObject_t cube, const uint32_t* queue, const uint32_t* queue_count,
                  uint32_t n_max, uint32_t* queue_out, uint32_t* count_out, Counters* ctr, int bounce,
                  cudaStream_t s);
void launch_shade_preview(const FrameParams& fp, const DScene& sc, const PathState& st,
                          cudaTextureObject_t cube, uint32_t n_max, Counters* ctr, cudaStream_t s);
void launch_accumulate(const FrameParams& fp, const PathState& st, float* accum, uint32_t* out_rgba,
                       float* radiance_out, Counters* ctr, cudaStream_t s);
void launch_intersect_rays(const DScene& sc, const float* origins, const float* dirs, uint64_t n,
                           int32_t* ids, float* ts, Counters* ctr, cudaStream_t s);
void launch_debug_rng(const uint32_t* seeds, uint32_t n_seeds, uint32_t n_draws, float* out,
                      cudaStream_t s);
void launch_export_primary(const PathState& st, uint32_t n, int32_t* ids, float* ts, float* origins,
                           float* dirs, cudaStream_t s);
// toolkit cuRAND on the device, for the RNG known-answer test
void launch_debug_curand(const uint32_t* seeds, uint32_t n_seeds, uint32_t n_draws, float* out,
                         cudaStream_t s);

} // namespace ptc
