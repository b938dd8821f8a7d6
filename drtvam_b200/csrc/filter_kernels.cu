// One-shot pre-filters that shrink the active pixel list before the optimisation
// (optimize.py:143-185): the Radon integrator (integrators/radon.py) and the corner
// integrator (integrators/filter_corner.py).  Both render one value per DMD pixel:
// image[pixel] += L * ray weight over the lanes of the pixel (radon.py:42-44).
#include "kernels.h"
#include "tracer.cuh"

namespace drtvam {

namespace {

constexpr int kThreads = 256;

// RadonIntegrator.sample (radon.py:48-106): the absorbed fraction along the parts of the path that
// lie inside the target mesh, under the purely-absorptive assumption of radon.py:76.  No Russian
// roulette and no free flight; `depth` of radon.py:93 is the argument that is never incremented,
// so every interface is sampled with the transmission lobe only.
__global__ void __launch_bounds__(kThreads)
radon_kernel(const __grid_constant__ SceneParams P, const uint32_t* __restrict__ active_pixels, uint64_t n_lanes,
             uint32_t spp, uint32_t seed, float ray_weight, float* __restrict__ image) {
    const uint64_t lane = static_cast<uint64_t>(blockIdx.x) * kThreads + threadIdx.x;
    if (lane >= n_lanes) return;
    Sampler rng;
    rng.seed(seed, static_cast<uint32_t>(lane));
    const uint32_t pixel = __ldg(active_pixels + lane / spp);
    Ray ray = generate_ray(P, pixel, rng);
    const float st = P.sigma_t;
    bool active_medium = false, inside_target = false;
    float throughput = 1.f, L = 0.f, t = 0.f;
    uint32_t d = 0;
    for (;;) {
        SurfaceHit H = nearest_surface(P, ray);
        bool hit_target = false;
        if (P.target.n_tri > 0) {
            float nx, ny, nz;
            const float tm = mesh_nearest(P.target, ray.ox, ray.oy, ray.oz, ray.dx, ray.dy, ray.dz, H.t, nx, ny, nz);
            if (tm < H.t) {
                hit_target = true;
                H.t = tm; H.shape = kMaxSurfaces;
                H.nx = nx; H.ny = ny; H.nz = nz;
                H.px = fmaf(ray.dx, tm, ray.ox); H.py = fmaf(ray.dy, tm, ray.oy); H.pz = fmaf(ray.dz, tm, ray.oz);
            }
        }
        const bool valid = H.shape >= 0;
        const float seg = valid ? H.t : INFINITY;
        const float contrib = throughput * expf(-st * t) * (1.f - expf(-st * seg));
        if (inside_target && active_medium) L += contrib;  // radon.py:79 is not masked by `active`
        if (!valid) break;
        t += H.t;
        inside_target = inside_target != hit_target;
        (void)rng.next(); (void)rng.next(); (void)rng.next();  // s1, s2 of radon.py:85-86
        if (hit_target) {
            spawn_through(H, ray);  // null BSDF, weight 1; no medium transition
        } else {
            throughput *= cross_interface(P, H, false, 0.f, ray);
            const Surface& S = P.surf[H.shape];
            if (S.interior_medium != S.exterior_medium) {  // is_medium_transition, radon.py:104
                const bool outward = (ray.dx * H.nx + ray.dy * H.ny + ray.dz * H.nz) > 0.f;
                active_medium = outward ? (S.exterior_medium != 0) : (S.interior_medium != 0);
            }
            d += 1;  // radon.py:101
        }
        if (!(d < static_cast<uint32_t>(P.max_depth))) break;
    }
    if (L != 0.f) atomicAdd(image + pixel, L * ray_weight);
}

// CornerIntegrator.sample (filter_corner.py:51-66): 1 unless the first hit lies within `radius` of
// the vertical edge |x| = |y| = dist of a square vial.
__global__ void __launch_bounds__(kThreads)
corner_kernel(const __grid_constant__ SceneParams P, const uint32_t* __restrict__ active_pixels, uint64_t n_lanes,
              uint32_t spp, uint32_t seed, float ray_weight, float dist, float radius, float* __restrict__ image) {
    const uint64_t lane = static_cast<uint64_t>(blockIdx.x) * kThreads + threadIdx.x;
    if (lane >= n_lanes) return;
    Sampler rng;
    rng.seed(seed, static_cast<uint32_t>(lane));
    const uint32_t pixel = __ldg(active_pixels + lane / spp);
    const Ray ray = generate_ray(P, pixel, rng);
    SurfaceHit H = nearest_surface(P, ray);
    if (P.target.n_tri > 0) {
        float nx, ny, nz;
        const float tm = mesh_nearest(P.target, ray.ox, ray.oy, ray.oz, ray.dx, ray.dy, ray.dz, H.t, nx, ny, nz);
        if (tm < H.t) {
            H.shape = kMaxSurfaces;
            H.px = fmaf(ray.dx, tm, ray.ox); H.py = fmaf(ray.dy, tm, ray.oy);
        }
    }
    if (H.shape < 0) return;
    const float ex = fabsf(H.px) - dist, ey = fabsf(H.py) - dist;
    if (sqrtf(ex * ex + ey * ey) < radius) return;
    atomicAdd(image + pixel, ray_weight);
}

}  // namespace

cudaError_t launch_radon(const SceneParams& P, const uint32_t* active_pixels, uint64_t n_lanes, uint32_t spp,
                         uint32_t seed, float ray_weight, float* image, cudaStream_t stream) {
    const size_t n_img = static_cast<size_t>(P.n_patterns) * P.res_x * P.res_y;
    cudaError_t e = cudaMemsetAsync(image, 0, sizeof(float) * n_img, stream);
    if (e != cudaSuccess || n_lanes == 0) return e;
    radon_kernel<<<static_cast<unsigned>((n_lanes + kThreads - 1) / kThreads), kThreads, 0, stream>>>(
        P, active_pixels, n_lanes, spp, seed, ray_weight, image);
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_corner(const SceneParams& P, const uint32_t* active_pixels, uint64_t n_lanes, uint32_t spp,
                          uint32_t seed, float ray_weight, float dist, float radius, float* image,
                          cudaStream_t stream) {
    const size_t n_img = static_cast<size_t>(P.n_patterns) * P.res_x * P.res_y;
    cudaError_t e = cudaMemsetAsync(image, 0, sizeof(float) * n_img, stream);
    if (e != cudaSuccess || n_lanes == 0) return e;
    corner_kernel<<<static_cast<unsigned>((n_lanes + kThreads - 1) / kThreads), kThreads, 0, stream>>>(
        P, active_pixels, n_lanes, spp, seed, ray_weight, dist, radius, image);
    count_launch();
    return cudaGetLastError();
}

}  // namespace drtvam
