// Surface-aware film on top of the planar kernels (C4 of BASELINE.json: square vial with
// refraction and surface-aware discretisation).
//
// The surface-aware film has two channels per voxel, selected by whether the ray is inside the
// target mesh while it crosses the voxel (sensor.py:405-409, volume.py:196,219).  Every crossing of
// the mesh ends a medium segment and re-spawns the ray a little past the surface along the normal
// (null BSDF, volume.py:228-247), so from its first crossing on a lane follows a slightly shifted
// ray: neither the flag nor the geometry fits the (angle, column) tables of the planar kernels.
// But for a deterministic configuration those sub-segments are the same in every pass, and only the
// lanes that meet the target have any.  So:
//
//   set-up   trace every lane once with the general tracer (tracer.cuh, BVH walk of mesh.cuh), with
//            and without the mesh.  A lane that crosses the mesh at t1 keeps records
//            {origin, direction, length, throughput at the origin, start}:
//              - the UNSPLIT segment with negative throughput, walked from t1 on only, and
//              - each sub-segment after the first crossing, tagged inside / outside,
//            stored lane after lane behind an offset table (count -> exclusive scan -> fill);
//   forward  plane A = planar kernels (unsplit rays)  +  DDA over the negative and the outside records
//            plane B = DDA over the inside records            (red.global.add into the planes)
//            film[v][0] += B[v],  film[v][1] += A[v]
//   adjoint  A = g[.][1], B = g[.][0];  grad = planar adjoint on A;
//            grad[lane] += DDA gather over the records, each from its plane
//
// The negative record keeps the origin of the unsplit segment so that its voxel boundaries are those
// of the walk it cancels.  An unsplit segment with a frozen DDA axis (dda.cuh, sensor.py:359) is
// cancelled in full and its head is added back as a record of its own.
// The BVH traversal leaves the per-pass work altogether, and lanes that miss the target cost nothing
// beyond the planar kernels.
#include <cub/device/device_scan.cuh>

#include "kernels.h"
#include "tracer.cuh"

namespace drtvam {

namespace {

constexpr int kThreads = 256;

#define IN_TRY(expr)                      \
    do {                                  \
        cudaError_t _e = (expr);          \
        if (_e != cudaSuccess) return _e; \
    } while (0)

// kFill = false: cnt[lane] = number of records of the lane.
// kFill = true : the records are written at off[lane].
// P is the full scene (mesh attached), P1 its one-channel view without the mesh.
// flag[0] |= 1 when a lane that meets the mesh has more than one unsplit medium segment (not planar).
template <bool kFill>
__global__ void __launch_bounds__(kThreads)
inside_trace_kernel(const __grid_constant__ SceneParams P, const __grid_constant__ SceneParams P1, uint64_t n_lanes,
                    uint32_t* __restrict__ cnt, const uint64_t* __restrict__ off, float4* __restrict__ seg,
                    float* __restrict__ t_lo, int* __restrict__ flag) {
    const uint64_t lane = static_cast<uint64_t>(blockIdx.x) * kThreads + threadIdx.x;
    if (lane >= n_lanes) return;
    Sampler rng;
    rng.seed(0u, 0u);  // never consumed: deterministic configuration
    const Ray ray0 = generate_ray(P, static_cast<uint32_t>(lane), rng);
    uint64_t at = kFill ? off[lane] : 0;
    auto emit = [&](const Ray& r, float maxt, float w0, bool inside, float start) {
        if (kFill) {
            seg[2 * at] = make_float4(r.ox, r.oy, r.oz, inside ? -maxt : maxt);  // sign of the length = plane
            seg[2 * at + 1] = make_float4(r.dx, r.dy, r.dz, w0);
            t_lo[at] = start;
            ++at;
        }
    };
    // with the mesh: sub-segment 0 is the head of the unsplit segment and stays with the planar kernels
    TraceState T;
    T.ray = ray0;
    uint32_t n_sub = 0;
    float t1 = 0.f;
    while (bounce_once<false>(P, T, rng, [&](const Ray& r, float maxt, float w0, bool inside, float) {
        if (n_sub == 0) t1 = maxt;
        else emit(r, maxt, w0, inside, 0.f);
        ++n_sub;
    })) {}
    if (n_sub < 2) {
        if (!kFill) cnt[lane] = 0;
        return;
    }
    // without the mesh: what the planar kernels deposit for this lane
    TraceState U;
    U.ray = ray0;
    uint32_t n_u = 0;
    Ray ur = ray0;
    float u_maxt = 0.f, u_w0 = 0.f;
    while (bounce_once<false>(P1, U, rng, [&](const Ray& r, float maxt, float w0, bool, float) {
        if (n_u == 0) { ur = r; u_maxt = maxt; u_w0 = w0; }
        ++n_u;
    })) {}
    if (n_u != 1) atomicOr(flag, 1);
    const bool whole = dda_has_frozen_axis(P1, ur, u_maxt) || !(t1 > 0.f);
    if (whole) emit(ur, t1, u_w0, false, 0.f);
    emit(ur, u_maxt, -u_w0, false, whole ? 0.f : t1);
    if (!kFill) cnt[lane] = n_sub + (whole ? 1u : 0u);
}

__global__ void widen_kernel(const uint32_t* __restrict__ in, uint64_t* __restrict__ out, uint64_t n) {
    const uint64_t i = static_cast<uint64_t>(blockIdx.x) * kThreads + threadIdx.x;
    if (i < n) out[i] = in[i];
}

__global__ void narrow_kernel(const uint64_t* __restrict__ in, uint32_t* __restrict__ out, uint64_t n) {
    const uint64_t i = static_cast<uint64_t>(blockIdx.x) * kThreads + threadIdx.x;
    if (i < n) out[i] = static_cast<uint32_t>(in[i]);
}

// P is the one-channel view of the scene (surface_aware = 0): the planes are indexed by voxel.
// One thread per listed lane walks the lane's records one after the other: neighbouring lanes
// (columns of one DMD row) have records of the same kind and nearly the same length, so a warp
// stays together, and the adjoint adds to its pattern pixel without an atomic.
template <int kMode>
__global__ void __launch_bounds__(kThreads)
inside_pass_kernel(const __grid_constant__ SceneParams P, const uint32_t* __restrict__ off,
                   const float4* __restrict__ seg, const float* __restrict__ t_lo, const float* __restrict__ in,
                   float* __restrict__ out, size_t n_vox, const uint32_t* __restrict__ active_pixels,
                   uint64_t n_active, float ray_weight, const int* __restrict__ gate) {
    if (*gate != 0) return;
    const uint64_t j = static_cast<uint64_t>(blockIdx.x) * kThreads + threadIdx.x;
    if (j >= n_active) return;
    const uint64_t lane = __ldg(active_pixels) + j;  // the pixel list is base + j (planar check)
    const uint32_t r0 = __ldg(off + lane), r1 = __ldg(off + lane + 1);
    if (r0 == r1) return;
    float em = 0.f;
    if (kMode == 0) {
        em = __ldg(in + j) * ray_weight;
        if (em == 0.f) return;
    }
    float acc = 0.f;
    for (uint32_t i = r0; i < r1; ++i) {
        const float4 a = __ldg(seg + 2 * static_cast<size_t>(i)), b = __ldg(seg + 2 * static_cast<size_t>(i) + 1);
        Ray r;
        r.ox = a.x; r.oy = a.y; r.oz = a.z;
        r.dx = b.x; r.dy = b.y; r.dz = b.z;
        const size_t plane = a.w < 0.f ? n_vox : 0;  // plane B (inside the target) after plane A
        const float start = __ldg(t_lo + i);
        DdaWalk W;
        const bool any = start > 0.f ? W.begin<true>(P, r, fabsf(a.w), b.w, false, start)
                                     : W.begin<false>(P, r, fabsf(a.w), b.w, false);
        if (!any) continue;
        if (kMode == 0) {
            // accurate deposits: the negative record must cancel the planar kernels' deposits, whose
            // coefficients are accurate too (planarz_kernels.cu::walk_entry), to well below 1e-4
            ScatterSink sink{out + plane, em};
            while (W.step<true>(sink)) {}
        } else {
            GatherSink sink{in + plane};
            while (W.step<true>(sink)) {}
            acc += sink.acc;
        }
    }
    if (kMode == 1) out[j] += acc * ray_weight;
}

__global__ void __launch_bounds__(kThreads)
inside_combine_kernel(const float* __restrict__ outside, const float* __restrict__ inside, float2* __restrict__ film,
                      size_t n_vox, const int* __restrict__ gate) {
    if (*gate != 0) return;
    const size_t v = static_cast<size_t>(blockIdx.x) * kThreads + threadIdx.x;
    if (v >= n_vox) return;
    float2 f = film[v];
    f.x += inside[v];
    f.y += outside[v];
    film[v] = f;
}

__global__ void __launch_bounds__(kThreads)
inside_split_kernel(const float2* __restrict__ g, float* __restrict__ g_out, float* __restrict__ g_in, size_t n_vox) {
    const size_t v = static_cast<size_t>(blockIdx.x) * kThreads + threadIdx.x;
    if (v >= n_vox) return;
    const float2 f = g[v];
    g_out[v] = f.y;
    g_in[v] = f.x;
}

unsigned blocks_for(uint64_t n) { return static_cast<unsigned>((n + kThreads - 1) / kThreads); }

}  // namespace

void inside_release(InsideHost& H) {
    if (H.off) cudaFree(H.off);
    if (H.seg) cudaFree(H.seg);
    if (H.t_lo) cudaFree(H.t_lo);
    if (H.planes) cudaFree(H.planes);
    H = InsideHost{};
}

// P: the full scene (target mesh attached, surface_aware = 1).  One-off set-up; synchronises.
cudaError_t inside_build(const SceneParams& P, const SceneParams& P1, InsideHost& H) {
    inside_release(H);
    const uint64_t n_lanes = static_cast<uint64_t>(P.n_patterns) * P.res_x * P.res_y;
    const size_t n_vox = static_cast<size_t>(P.res[0]) * P.res[1] * P.res[2];
    uint32_t* cnt = nullptr;
    int* flag = nullptr;
    uint64_t *wide = nullptr, *off = nullptr;
    void* tmp = nullptr;
    size_t tmp_bytes = 0;
    cudaError_t e = cudaSuccess;
    auto cleanup = [&]() {
        if (cnt) cudaFree(cnt);
        if (flag) cudaFree(flag);
        if (wide) cudaFree(wide);
        if (off) cudaFree(off);
        if (tmp) cudaFree(tmp);
    };
    auto run = [&]() -> cudaError_t {
        IN_TRY(cudaMalloc(&cnt, sizeof(uint32_t) * n_lanes));
        IN_TRY(cudaMalloc(&wide, sizeof(uint64_t) * (n_lanes + 1)));
        IN_TRY(cudaMalloc(&off, sizeof(uint64_t) * (n_lanes + 1)));
        IN_TRY(cudaMemset(wide + n_lanes, 0, sizeof(uint64_t)));
        IN_TRY(cudaMalloc(&flag, sizeof(int)));
        IN_TRY(cudaMemset(flag, 0, sizeof(int)));
        inside_trace_kernel<false><<<blocks_for(n_lanes), kThreads>>>(P, P1, n_lanes, cnt, nullptr, nullptr, nullptr,
                                                                     flag);
        count_launch();
        widen_kernel<<<blocks_for(n_lanes), kThreads>>>(cnt, wide, n_lanes);
        count_launch();
        IN_TRY(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, wide, off, n_lanes + 1));
        IN_TRY(cudaMalloc(&tmp, tmp_bytes));
        IN_TRY(cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, wide, off, n_lanes + 1));
        count_launch();
        uint64_t total = 0;
        IN_TRY(cudaMemcpy(&total, off + n_lanes, sizeof(total), cudaMemcpyDeviceToHost));
        int bad = 0;
        IN_TRY(cudaMemcpy(&bad, flag, sizeof(bad), cudaMemcpyDeviceToHost));
        if (bad) return cudaErrorNotSupported;
        H.n_rec = total;
        if (total) {
            size_t free_b = 0, all_b = 0;
            IN_TRY(cudaMemGetInfo(&free_b, &all_b));
            if (total * 40 > free_b / 2 || total >= (1ull << 32)) return cudaErrorMemoryAllocation;
            IN_TRY(cudaMalloc(&H.seg, sizeof(float4) * 2 * total));
            IN_TRY(cudaMalloc(&H.t_lo, sizeof(float) * total));
            inside_trace_kernel<true><<<blocks_for(n_lanes), kThreads>>>(P, P1, n_lanes, nullptr, off, H.seg, H.t_lo,
                                                                        flag);
            count_launch();
        }
        IN_TRY(cudaMalloc(&H.off, sizeof(uint32_t) * (n_lanes + 1)));
        narrow_kernel<<<blocks_for(n_lanes + 1), kThreads>>>(off, H.off, n_lanes + 1);
        count_launch();
        IN_TRY(cudaMalloc(&H.planes, sizeof(float) * 2 * n_vox));
        IN_TRY(cudaDeviceSynchronize());
        return cudaGetLastError();
    };
    e = run();
    cleanup();
    if (e != cudaSuccess) {
        inside_release(H);
        return e;
    }
    H.n_vox = n_vox;
    H.valid = true;
    return cudaSuccess;
}

// Forward, part 1: clears the two planes; the caller then runs the planar forward kernels into plane 0.
cudaError_t inside_forward_begin(InsideHost& H, cudaStream_t stream) {
    return cudaMemsetAsync(H.planes, 0, sizeof(float) * 2 * H.n_vox, stream);
}

// Forward, part 2: the records into the planes, then both planes into the two-channel film.
cudaError_t inside_forward_finish(const SceneParams& P1, InsideHost& H, const int* gate, const float* active_data,
                                  const uint32_t* active_pixels, uint64_t n_active, float ray_weight, float* film,
                                  cudaStream_t stream) {
    if (H.n_rec) {
        inside_pass_kernel<0><<<blocks_for(n_active), kThreads, 0, stream>>>(P1, H.off, H.seg, H.t_lo,
                                                                           active_data, H.planes, H.n_vox, active_pixels, n_active,
                                                                           ray_weight, gate);
        count_launch();
    }
    inside_combine_kernel<<<blocks_for(H.n_vox), kThreads, 0, stream>>>(H.planes, H.planes + H.n_vox,
                                                                       reinterpret_cast<float2*>(film), H.n_vox, gate);
    count_launch();
    return cudaGetLastError();
}

// Adjoint, part 1: plane A = outside-channel gradient (input of the planar adjoint), plane B = inside channel.
cudaError_t inside_adjoint_begin(InsideHost& H, const float* dL, cudaStream_t stream) {
    inside_split_kernel<<<blocks_for(H.n_vox), kThreads, 0, stream>>>(reinterpret_cast<const float2*>(dL), H.planes,
                                                                     H.planes + H.n_vox, H.n_vox);
    count_launch();
    return cudaGetLastError();
}

// Adjoint, part 2 (after the planar adjoint has written grad): adds the gather over the inside records.
cudaError_t inside_adjoint_finish(const SceneParams& P1, InsideHost& H, const int* gate, const uint32_t* active_pixels,
                                  uint64_t n_active, float ray_weight, float* grad, cudaStream_t stream) {
    if (!H.n_rec) return cudaSuccess;
    inside_pass_kernel<1><<<blocks_for(n_active), kThreads, 0, stream>>>(P1, H.off, H.seg, H.t_lo, H.planes,
                                                                       grad, H.n_vox, active_pixels, n_active, ray_weight,
                                                                       gate);
    count_launch();
    return cudaGetLastError();
}

}  // namespace drtvam
