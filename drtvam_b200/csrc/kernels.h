// Host-side launch interface of the kernel translation units.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>
#include "device_scene.cuh"

namespace drtvam {

void count_launch();  // bumps the library-wide launch counter (abi.cu)

// general per-lane tracer (general_kernels.cu)
cudaError_t launch_forward_general(const SceneParams& P, const float* active_data, const uint32_t* active_pixels,
                                   uint64_t n_lanes, uint32_t spp, uint32_t seed, float ray_weight, float* film,
                                   cudaStream_t stream);
cudaError_t launch_adjoint_general(const SceneParams& P, const float* dL, const uint32_t* active_pixels,
                                   uint64_t n_lanes, uint32_t spp, uint32_t seed, float ray_weight, float* grad,
                                   cudaStream_t stream);
cudaError_t launch_raygen(const SceneParams& P, const uint32_t* active_pixels, uint64_t n_lanes, uint32_t spp,
                          uint32_t seed, float* out, cudaStream_t stream);
cudaError_t launch_count_general(const SceneParams& P, const uint32_t* active_pixels, uint64_t n_lanes, uint32_t spp,
                                 uint32_t seed, unsigned long long* total, cudaStream_t stream);

}  // namespace drtvam
