// Host-side launch interface of the kernel translation units.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>
#include "device_scene.cuh"
#include "planar.cuh"

namespace drtvam {

void count_launch();  // bumps the library-wide launch counter (abi.cu)

// general per-lane tracer (general_kernels.cu)
cudaError_t launch_forward_general(const SceneParams& P, const float* active_data, const uint32_t* active_pixels,
                                   uint64_t n_lanes, uint32_t spp, uint32_t seed, float ray_weight, float* film,
                                   cudaStream_t stream, const int* gate = nullptr);
cudaError_t launch_adjoint_general(const SceneParams& P, const float* dL, const uint32_t* active_pixels,
                                   uint64_t n_lanes, uint32_t spp, uint32_t seed, float ray_weight, float* grad,
                                   cudaStream_t stream, const int* gate = nullptr);
cudaError_t launch_raygen(const SceneParams& P, const uint32_t* active_pixels, uint64_t n_lanes, uint32_t spp,
                          uint32_t seed, float* out, cudaStream_t stream);
cudaError_t launch_count_general(const SceneParams& P, const uint32_t* active_pixels, uint64_t n_lanes, uint32_t spp,
                                 uint32_t seed, unsigned long long* total, cudaStream_t stream);


// planar z-separable fast path (planar_kernels.cu); `gate` (device int) is 0
// when the planar kernels own the call and non-zero when the general tracer does
struct PlanarHost {
    bool valid = false;
    int flags = 0;             // kPlanarFlag* reported by the table builder
    int n_rows_in_sensor = 0;  // DMD rows whose z lies inside the sensor box
    float4* seg_a = nullptr;
    float2* seg_b = nullptr;
    int* layer_row_begin = nullptr;
    int* layer_rows = nullptr;
    int* tile_off = nullptr;
    int* tile_clo = nullptr;
    int* gate = nullptr;  // [0] per-call pixel-layout gate, [1] build flags
    PlanarPlan plan{};
    size_t smem_bytes = 0;
};
cudaError_t planar_build(const SceneParams& P, PlanarHost& H);
void planar_release(PlanarHost& H);
cudaError_t launch_planar_check(const SceneParams& P, PlanarHost& H, const uint32_t* active_pixels, uint64_t n_active,
                                cudaStream_t stream);
cudaError_t launch_planar_forward(const SceneParams& P, PlanarHost& H, const float* active_data,
                                  const uint32_t* active_pixels, uint64_t n_active, float ray_weight, float* film,
                                  cudaStream_t stream);
cudaError_t launch_planar_adjoint(const SceneParams& P, PlanarHost& H, const float* dL, const uint32_t* active_pixels,
                                  uint64_t n_active, float ray_weight, float* grad, cudaStream_t stream);

}  // namespace drtvam
