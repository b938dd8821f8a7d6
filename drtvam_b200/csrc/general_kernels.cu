// General per-lane tracer kernels: one thread per (active pixel, sample).
// Handles every configuration of the path (jittered sampling, aperture
// projectors, scattering, reflections, any vial built from cylinders/boxes).
// Reference: VolumeIntegrator.render / render_backward
// (integrators/volume.py:18-56, 97-134).
#include "kernels.h"
#include "tracer.cuh"

namespace drtvam {

namespace {

constexpr int kThreads = 256;

template <bool kScatter>
__global__ void __launch_bounds__(kThreads)
forward_general_kernel(const __grid_constant__ SceneParams P, const float* __restrict__ active_data,
                       const uint32_t* __restrict__ active_pixels, uint64_t n_lanes, uint32_t spp, uint32_t seed,
                       float ray_weight, float* __restrict__ film, const int* __restrict__ gate) {
    const uint64_t lane = static_cast<uint64_t>(blockIdx.x) * kThreads + threadIdx.x;
    if (lane >= n_lanes) return;
    if (gate && *gate == 0) return;  // the planar fast path owns this call
    const uint64_t j = lane / spp;
    const float em = __ldg(active_data + j) * ray_weight;
    if (em == 0.f && !kScatter) return;  // a zero pixel deposits nothing
    Sampler rng;
    rng.seed(seed, static_cast<uint32_t>(lane));
    const Ray ray = generate_ray(P, __ldg(active_pixels + j), rng);
    ScatterSink sink{film, em};
    trace_lane<kScatter>(P, ray, rng, sink);
}

template <bool kScatter>
__global__ void __launch_bounds__(kThreads)
adjoint_general_kernel(const __grid_constant__ SceneParams P, const float* __restrict__ dL,
                       const uint32_t* __restrict__ active_pixels, uint64_t n_lanes, uint32_t spp, uint32_t seed,
                       float ray_weight, float* __restrict__ grad, const int* __restrict__ gate) {
    if (gate && *gate == 0) return;  // the planar fast path owns this call
    const uint64_t lane = static_cast<uint64_t>(blockIdx.x) * kThreads + threadIdx.x;
    const bool live = lane < n_lanes;
    const uint64_t j = live ? lane / spp : 0;
    float g = 0.f;
    if (live) {
        Sampler rng;
        rng.seed(seed, static_cast<uint32_t>(lane));
        const Ray ray = generate_ray(P, __ldg(active_pixels + j), rng);
        GatherSink sink{dL};
        trace_lane<kScatter>(P, ray, rng, sink);
        g = sink.acc * ray_weight;
    }
    // sum over the spp lanes of one pixel (the dr.repeat of common.py:82)
    if (spp == 1) {
        if (live) grad[j] = g;
    } else if ((spp & (spp - 1)) == 0 && spp <= 32) {
        // the spp lanes of a pixel are consecutive and aligned inside the warp
        for (uint32_t off = spp >> 1; off > 0; off >>= 1) g += __shfl_xor_sync(0xffffffffu, g, off);
        if (live && (lane % spp) == 0) grad[j] = g;
    } else {
        if (live) atomicAdd(grad + j, g);  // grad pre-zeroed by the host
    }
}

__global__ void __launch_bounds__(kThreads)
raygen_kernel(const __grid_constant__ SceneParams P, const uint32_t* __restrict__ active_pixels, uint64_t n_lanes,
              uint32_t spp, uint32_t seed, float* __restrict__ out) {
    const uint64_t lane = static_cast<uint64_t>(blockIdx.x) * kThreads + threadIdx.x;
    if (lane >= n_lanes) return;
    Sampler rng;
    rng.seed(seed, static_cast<uint32_t>(lane));
    const Ray r = generate_ray(P, __ldg(active_pixels + lane / spp), rng);
    float* o = out + 6 * lane;
    o[0] = r.ox; o[1] = r.oy; o[2] = r.oz; o[3] = r.dx; o[4] = r.dy; o[5] = r.dz;
}

template <bool kScatter>
__global__ void __launch_bounds__(kThreads)
count_general_kernel(const __grid_constant__ SceneParams P, const uint32_t* __restrict__ active_pixels,
                     uint64_t n_lanes, uint32_t spp, uint32_t seed, unsigned long long* __restrict__ total) {
    const uint64_t lane = static_cast<uint64_t>(blockIdx.x) * kThreads + threadIdx.x;
    unsigned long long n = 0;
    if (lane < n_lanes) {
        Sampler rng;
        rng.seed(seed, static_cast<uint32_t>(lane));
        const Ray ray = generate_ray(P, __ldg(active_pixels + lane / spp), rng);
        CountSink sink;
        trace_lane<kScatter>(P, ray, rng, sink);
        n = sink.n;
    }
    for (int off = 16; off > 0; off >>= 1) n += __shfl_xor_sync(0xffffffffu, n, off);
    if ((threadIdx.x & 31) == 0 && n) atomicAdd(total, n);
}

inline unsigned grid_for(uint64_t n_lanes) { return static_cast<unsigned>((n_lanes + kThreads - 1) / kThreads); }

}  // namespace

cudaError_t launch_forward_general(const SceneParams& P, const float* active_data, const uint32_t* active_pixels,
                                   uint64_t n_lanes, uint32_t spp, uint32_t seed, float ray_weight, float* film,
                                   cudaStream_t stream, const int* gate) {
    if (n_lanes == 0) return cudaSuccess;
    if (P.albedo != 0.f)
        forward_general_kernel<true><<<grid_for(n_lanes), kThreads, 0, stream>>>(P, active_data, active_pixels,
                                                                                  n_lanes, spp, seed, ray_weight, film, gate);
    else
        forward_general_kernel<false><<<grid_for(n_lanes), kThreads, 0, stream>>>(P, active_data, active_pixels,
                                                                                   n_lanes, spp, seed, ray_weight, film, gate);
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_adjoint_general(const SceneParams& P, const float* dL, const uint32_t* active_pixels,
                                   uint64_t n_lanes, uint32_t spp, uint32_t seed, float ray_weight, float* grad,
                                   cudaStream_t stream, const int* gate) {
    if (n_lanes == 0) return cudaSuccess;
    const bool pow2 = (spp & (spp - 1)) == 0 && spp <= 32;
    if (!pow2 && gate == nullptr) {  // a gated call hands in a zeroed gradient (abi.cu)
        cudaError_t e = cudaMemsetAsync(grad, 0, sizeof(float) * (n_lanes / spp), stream);
        if (e != cudaSuccess) return e;
    }
    if (P.albedo != 0.f)
        adjoint_general_kernel<true><<<grid_for(n_lanes), kThreads, 0, stream>>>(P, dL, active_pixels, n_lanes, spp,
                                                                                  seed, ray_weight, grad, gate);
    else
        adjoint_general_kernel<false><<<grid_for(n_lanes), kThreads, 0, stream>>>(P, dL, active_pixels, n_lanes, spp,
                                                                                   seed, ray_weight, grad, gate);
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_raygen(const SceneParams& P, const uint32_t* active_pixels, uint64_t n_lanes, uint32_t spp,
                          uint32_t seed, float* out, cudaStream_t stream) {
    if (n_lanes == 0) return cudaSuccess;
    raygen_kernel<<<grid_for(n_lanes), kThreads, 0, stream>>>(P, active_pixels, n_lanes, spp, seed, out);
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_count_general(const SceneParams& P, const uint32_t* active_pixels, uint64_t n_lanes, uint32_t spp,
                                 uint32_t seed, unsigned long long* total, cudaStream_t stream) {
    if (n_lanes == 0) return cudaSuccess;
    if (P.albedo != 0.f)
        count_general_kernel<true><<<grid_for(n_lanes), kThreads, 0, stream>>>(P, active_pixels, n_lanes, spp, seed,
                                                                                total);
    else
        count_general_kernel<false><<<grid_for(n_lanes), kThreads, 0, stream>>>(P, active_pixels, n_lanes, spp, seed,
                                                                                 total);
    count_launch();
    return cudaGetLastError();
}

}  // namespace drtvam
