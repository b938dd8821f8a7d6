// Voxel traversal of one medium segment with analytic per-voxel absorption.
// Reference: DDAVolumetricSensor.accumulate (sensor.py:306-440) and
// RatioVolumetricSensor.accumulate (sensor.py:205-295); film.write
// (film.py:40-41) is the Sink of the primal pass, the dr.gather of
// sensor.py:419 the Sink of the adjoint.
#pragma once
#include "device_scene.cuh"
#include "rng.cuh"

namespace drtvam {

// Primal sink: red.global.add.f32 into the film (no return value -> REDG).
struct ScatterSink {
    static constexpr bool kRecordOnly = false;
    float* film;
    float em;  // emitted radiance * ray weight
    float unused = 0.f;
    __device__ __forceinline__ void visit(size_t idx, float w) { atomicAdd(film + idx, w * em); }
};

// Adjoint sink: gather dL and accumulate d/d(em).
struct GatherSink {
    static constexpr bool kRecordOnly = false;
    const float* dL;
    float acc = 0.f;
    __device__ __forceinline__ void visit(size_t idx, float w) { acc = fmaf(w, __ldg(dL + idx), acc); }
};

// Counting sink (roofline units).
struct CountSink {
    static constexpr bool kRecordOnly = false;
    unsigned long long n = 0;
    __device__ __forceinline__ void visit(size_t, float) { ++n; }
};

// Amanatides-Woo walk over [0, maxt] of `ray`, clipped to the sensor box.
// `w0` is the throughput at the segment origin times sigma_a / sigma_t.
template <class Sink>
__device__ __forceinline__ void dda_segment(const SceneParams& P, const Ray& ray, float maxt, float w0,
                                            bool inside_target, Sink& sink) {
    const float o[3] = {ray.ox, ray.oy, ray.oz};
    const float d[3] = {ray.dx, ray.dy, ray.dz};

    // slab clip (sensor.py:328-337)
    float t_near = -INFINITY, t_far = INFINITY;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float t0 = (P.bmin[a] - o[a]) / d[a];
        const float t1 = (P.bmax[a] - o[a]) / d[a];
        t_near = fmaxf(t_near, fminf(t0, t1));
        t_far = fminf(t_far, fmaxf(t0, t1));
    }
    const float t_start = fmaxf(t_near, 0.f);
    const float t_end = fminf(t_far, maxt);
    if (!(isfinite(t_start) && isfinite(t_end) && t_start < t_end)) return;

    int cur[3], endv[3], stp[3];
    float dtmax[3], tstep[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float vs = (P.bmax[a] - P.bmin[a]) / static_cast<float>(P.res[a]);
        const float gs = fmaf(d[a], t_start, o[a]);
        const float ge = fmaf(d[a], t_end, o[a]);
        stp[a] = d[a] > 0.f ? 1 : -1;
        // truncation toward zero then clip (sensor.py:345-346)
        cur[a] = min(max(__float2int_rz((gs - P.bmin[a]) / vs), 0), P.res[a] - 1);
        endv[a] = min(max(__float2int_rz((ge - P.bmin[a]) / vs), 0), P.res[a] - 1);
        float boundary = P.bmin[a] + static_cast<float>(cur[a] + stp[a]) * vs;
        if (d[a] < 0.f) boundary += vs;
        const bool valid = fabsf(d[a]) > 1e-8f;
        float first = valid ? (boundary - gs) / d[a] : INFINITY;
        if (first < 0.f) first = INFINITY;  // sensor.py:359
        dtmax[a] = first;
        tstep[a] = valid ? vs / d[a] * static_cast<float>(stp[a]) : INFINITY;
    }

    const float st = P.sigma_t;
    float t = t_start;
    float remaining = t_end - t_start;
    const int rx = P.res[0], rxy = P.res[0] * P.res[1];
    bool active = true;
    while (active) {
        const float dt = fminf(fminf(fminf(dtmax[0], dtmax[1]), dtmax[2]), remaining);
        remaining -= dt;
        const float w = w0 * expf(-st * t) * (1.f - expf(-st * fmaxf(dt, 0.f)));
        const size_t flat = static_cast<size_t>(cur[0]) + static_cast<size_t>(cur[1]) * rx +
                            static_cast<size_t>(cur[2]) * rxy;
        const size_t idx = P.surface_aware ? (2 * flat + (inside_target ? 0 : 1)) : flat;
        sink.visit(idx, w);

        active = (endv[0] != cur[0] || endv[1] != cur[1] || endv[2] != cur[2]) && (remaining > 1e-6f);
        bool in_grid = true;
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const bool hit = dtmax[a] == dt;
            dtmax[a] = hit ? tstep[a] : dtmax[a] - dt;
            if (active && hit) cur[a] += stp[a];
            in_grid = in_grid && cur[a] >= 0 && cur[a] < P.res[a];
        }
        active = active && in_grid;
        t += dt;
    }
}

// The reference sets `dtmax[dtmax < 0] = inf` (sensor.py:359).  When fp32
// rounding puts the segment start a hair on the far side of the boundary of
// the voxel it was truncated into, that axis is frozen for the whole walk and
// the ray deposits along a straight line of voxels.  dda_segment reproduces
// this; the tiled planar kernels cannot, so they ask this predicate (same
// arithmetic as the head of dda_segment) and leave such rays to dda_segment.
__device__ __forceinline__ bool dda_has_frozen_axis(const SceneParams& P, const Ray& ray, float maxt) {
    const float o[3] = {ray.ox, ray.oy, ray.oz};
    const float d[3] = {ray.dx, ray.dy, ray.dz};
    float t_near = -INFINITY, t_far = INFINITY;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float t0 = (P.bmin[a] - o[a]) / d[a];
        const float t1 = (P.bmax[a] - o[a]) / d[a];
        t_near = fmaxf(t_near, fminf(t0, t1));
        t_far = fminf(t_far, fmaxf(t0, t1));
    }
    const float t_start = fmaxf(t_near, 0.f);
    const float t_end = fminf(t_far, maxt);
    if (!(isfinite(t_start) && isfinite(t_end) && t_start < t_end)) return false;
    bool frozen = false;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float vs = (P.bmax[a] - P.bmin[a]) / static_cast<float>(P.res[a]);
        const float gs = fmaf(d[a], t_start, o[a]);
        const int stp = d[a] > 0.f ? 1 : -1;
        const int cur = min(max(__float2int_rz((gs - P.bmin[a]) / vs), 0), P.res[a] - 1);
        float boundary = P.bmin[a] + static_cast<float>(cur + stp) * vs;
        if (d[a] < 0.f) boundary += vs;
        if (fabsf(d[a]) > 1e-8f) {
            // frozen axis, with a guard band: the decision is a knife edge of fp32 rounding that two
            // compilations of this arithmetic can take differently, so everything within 1e-4 voxel of
            // it goes to dda_segment as well (which then simply walks normally)
            if ((boundary - gs) / d[a] < 0.f || fabsf(boundary - gs) < 1e-4f * vs) frozen = true;
        }
        if (a < 2 && fabsf(d[a]) > 1e-8f && fabsf(d[a]) < 1e-5f * fabsf(d[1 - a])) {
            const float ge = fmaf(d[a], t_end, o[a]);
            const float fs = (gs - P.bmin[a]) / vs, fe = (ge - P.bmin[a]) / vs;
            const int endv = min(max(__float2int_rz(fe), 0), P.res[a] - 1);
            const bool near_edge = fabsf(fs - rintf(fs)) < 1e-3f || fabsf(fe - rintf(fe)) < 1e-3f;
            if (endv != cur || near_edge) frozen = true;
        }
    }
    return frozen;
}

// Ratio-tracking estimator (sensor.py:243-293): exponential steps with the
// majorant, one deposit per tentative collision.
template <class Sink>
__device__ __forceinline__ void ratio_segment(const SceneParams& P, const Ray& ray, float maxt, float w0,
                                              bool inside_target, Sampler& rng, Sink& sink) {
    const float st = P.sigma_t, mu = P.majorant;
    const float base = w0 * st / mu;
    const float ratio = 1.f - st / mu;
    float t = 0.f, k = 0.f;
    for (;;) {
        t += -logf(1.f - rng.next()) / mu;
        if (!(t < maxt)) break;
        const float p[3] = {fmaf(ray.dx, t, ray.ox), fmaf(ray.dy, t, ray.oy), fmaf(ray.dz, t, ray.oz)};
        int v[3];
        bool in = true;
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const float vs = (P.bmax[a] - P.bmin[a]) / static_cast<float>(P.res[a]);
            v[a] = __float2int_rd((p[a] - P.bmin[a]) / vs);
            in = in && v[a] >= 0 && v[a] < P.res[a];
        }
        if (in) {
            const size_t flat = static_cast<size_t>(v[0]) + static_cast<size_t>(v[1]) * P.res[0] +
                                static_cast<size_t>(v[2]) * P.res[0] * P.res[1];
            const size_t idx = P.surface_aware ? (2 * flat + (inside_target ? 0 : 1)) : flat;
            sink.visit(idx, base * powf(ratio, k));
        }
        k += 1.f;
    }
}

}  // namespace drtvam
