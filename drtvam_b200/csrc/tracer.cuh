// The bounce loop of one lane: interfaces, free flight, deposit, phase sampling.
// Reference: VolumeIntegrator.sample (integrators/volume.py:136-282).
#pragma once
#include "dda.cuh"
#include "device_scene.cuh"
#include "raygen.cuh"
#include "rng.cuh"
#include "surfaces.cuh"

namespace drtvam {

// Branchless orthonormal basis around n (Duff et al.), as Mitsuba's Frame3f.
__device__ __forceinline__ void frame_from_normal(float nx, float ny, float nz, float s[3], float t[3]) {
    const float sign = nz >= 0.f ? 1.f : -1.f;
    const float a = -1.f / (sign + nz);
    const float b = nx * ny * a;
    s[0] = nx * nx * a * sign + 1.f;  // mulsign(x, nz) = x * sign(nz)
    s[1] = b * sign;
    s[2] = -nx * sign;
    t[0] = b;
    t[1] = ny * ny * a + sign;
    t[2] = -ny;
}

// Samples the phase function around the propagation direction (dx,dy,dz)
// (volume.py:252-260).  HG follows Mitsuba's hg.cpp, Rayleigh rayleigh.cpp.
__device__ __forceinline__ void sample_phase(const SceneParams& P, float u1, float u2, Ray& ray) {
    float cos_theta;
    if (P.phase_kind == 2) {
        const float g = P.hg_g;
        const float sq = (1.f - g * g) / (1.f - g + 2.f * g * u1);
        cos_theta = (1.f + g * g - sq * sq) / (2.f * g);
        if (fabsf(g) < 1e-7f) cos_theta = 1.f - 2.f * u1;
        cos_theta = -cos_theta;  // expressed in the frame of wi = -d
    } else if (P.phase_kind == 1) {
        const float z = 2.f * (2.f * u1 - 1.f), tmp = sqrtf(z * z + 1.f);
        cos_theta = cbrtf(z + tmp) + cbrtf(z - tmp);
    } else {
        cos_theta = 1.f - 2.f * u1;
    }
    const float sin_theta = sqrtf(fmaxf(1.f - cos_theta * cos_theta, 0.f));
    float sp, cp;
    sincosf(6.28318530717958647692f * u2, &sp, &cp);
    const float wi[3] = {-ray.dx, -ray.dy, -ray.dz};
    float s[3], t[3];
    frame_from_normal(wi[0], wi[1], wi[2], s, t);
    const float a = sin_theta * cp, b = sin_theta * sp;
    ray.dx = s[0] * a + t[0] * b + wi[0] * cos_theta;
    ray.dy = s[1] * a + t[1] * b + wi[1] * cos_theta;
    ray.dz = s[2] * a + t[2] * b + wi[2] * cos_theta;
}

// State of one lane between two trips of the bounce loop (volume.py:165-179).
struct TraceState {
    Ray ray;
    float attenuation = 1.f;
    uint32_t depth = 0;
    bool in_medium = false;
    bool inside_target = false;  // every ray starts outside the target (volume.py:173-174)
};

// ONE trip of the bounce loop of volume.py:179-272: Russian roulette, nearest interface (vial
// surfaces and, in surface-aware runs, the target mesh), free flight, the medium segment of this
// trip handed to `on_segment(ray, maxt, w0, inside_target, t_coll)` with the throughput at its origin
// (t_coll: free-flight distance of a lane that collides inside the segment, else +inf), then
// the interface / phase event and the bookkeeping.  Nothing after the hand-over depends on how the
// segment is deposited, so the caller may walk it later.  Returns false when the lane is finished.
// kScatter is the scalar `has_scattering` branch of volume.py:154-155,199.
template <bool kScatter, class OnSegment>
__device__ __forceinline__ bool bounce_once(const SceneParams& P, TraceState& T, Sampler& rng,
                                            OnSegment&& on_segment) {
    const float st = P.sigma_t;
    const float ss = st * P.albedo;
    const float absorb = (st - ss) / st;  // sigma_a / sigma_t
    Ray& ray = T.ray;

    // Russian roulette (volume.py:181-186); the sample is always consumed
    const float q = fminf(0.99f, T.attenuation);
    const bool rr = T.depth > static_cast<uint32_t>(P.rr_depth);
    const float u_rr = rng.next();
    if (!((u_rr < q) || !rr)) return false;
    if (rr) T.attenuation *= 1.f / q;
    if (T.attenuation == 0.f) return false;

    SurfaceHit H = nearest_surface(P, ray);
    bool hit_target = false;  // volume.py:196
    if (P.target.n_tri > 0) {
        float nx, ny, nz;
        const float t = mesh_nearest(P.target, ray.ox, ray.oy, ray.oz, ray.dx, ray.dy, ray.dz, H.t, nx, ny, nz);
        if (t < H.t) {
            hit_target = true;
            H.t = t; H.shape = kMaxSurfaces;
            H.nx = nx; H.ny = ny; H.nz = nz;
            H.px = fmaf(ray.dx, t, ray.ox); H.py = fmaf(ray.dy, t, ray.oy); H.pz = fmaf(ray.dz, t, ray.oz);
        }
    }
    if (H.shape < 0) return false;  // volume.py:193

    float weight = 1.f;
    float t_coll = INFINITY;
    bool reached_surface = T.in_medium;
    if (kScatter) {
        const float xi = rng.next();  // unmasked draw, volume.py:200
        if (T.in_medium) {
            t_coll = -logf(1.f - xi) / st;
            reached_surface = H.t < t_coll;
            if (reached_surface) t_coll = INFINITY;
            const float tr = expf(-fminf(t_coll, H.t) * st);
            const float pdf = reached_surface ? tr : tr * st;
            weight *= tr * (pdf > 0.f ? 1.f / pdf : 0.f);
        }
    }
    const bool collide = T.in_medium && !reached_surface;

    // both collision and surface lanes deposit along the whole [0, si.t] (volume.py:216)
    if (T.in_medium) on_segment(ray, H.t, T.attenuation * absorb, T.inside_target, t_coll);

    T.inside_target = T.inside_target != hit_target;  // volume.py:219

    bool enters_medium = false;
    if (!collide) {
        // interface: BSDF sample + spawn (volume.py:228-247)
        const float s1 = rng.next();
        (void)rng.next();
        (void)rng.next();
        if (hit_target) {
            // null BSDF (optimize.py:70-72): straight through, weight 1; the ray stays in the medium
            spawn_through(H, ray);
            enters_medium = true;  // volume.py:268
        } else {
            const bool allow_reflection = !P.transmission_only && T.depth != 0;
            weight *= cross_interface(P, H, allow_reflection, s1, ray);
            const Surface& S = P.surf[H.shape];
            const bool outward = (ray.dx * H.nx + ray.dy * H.ny + ray.dz * H.nz) > 0.f;
            enters_medium = outward ? (S.exterior_medium != 0) : (S.interior_medium != 0);  // volume.py:268
        }
        if (!kScatter && reached_surface) weight *= expf(-st * H.t);  // volume.py:263
    } else {
        // medium collision: albedo weight + phase sampling (volume.py:249-261)
        weight *= ss;
        (void)rng.next();
        const float u1 = rng.next(), u2 = rng.next();
        ray.ox = fmaf(ray.dx, t_coll, ray.ox);
        ray.oy = fmaf(ray.dy, t_coll, ray.oy);
        ray.oz = fmaf(ray.dz, t_coll, ray.oz);
        sample_phase(P, u1, u2, ray);
    }
    T.attenuation *= weight;
    T.in_medium = collide || enters_medium;
    if (!hit_target) T.depth += 1;  // volume.py:271
    return T.depth < static_cast<uint32_t>(P.max_depth);
}

// Walks one lane through the scene from start to end, depositing every medium segment as it is met.
template <bool kScatter, class Sink>
__device__ __forceinline__ void trace_lane(const SceneParams& P, Ray ray, Sampler& rng, Sink& sink) {
    TraceState T;
    T.ray = ray;
    while (bounce_once<kScatter>(P, T, rng, [&](const Ray& r, float maxt, float w0, bool inside, float t_coll) {
        if constexpr (Sink::kRecordOnly) sink.record(r, maxt, w0);
        else if (P.sensor_kind == 2) delta_deposit(P, r, t_coll, w0, inside, sink);
        else if (P.sensor_kind == 1) ratio_segment(P, r, maxt, w0, inside, rng, sink);
        else dda_segment(P, r, maxt, w0, inside, sink);
    })) {}
}

// Deposits / gathers only the FIRST medium segment of a lane: in a scattering scene the segments after the first
// collision belong to the general tracer, the unscattered chords to the planar per-lane kernels
// (planarj_kernels.cu), whose odd lanes come here.  The geometry up to the first medium segment does not depend
// on the random numbers of the bounce loop (transmission only), so the non-scattering trip serves.
template <class Sink>
__device__ __forceinline__ void trace_first_segment(const SceneParams& P, Ray ray, Sampler& rng, Sink& sink) {
    TraceState T;
    T.ray = ray;
    int n = 0;
    while (bounce_once<false>(P, T, rng, [&](const Ray& r, float maxt, float w0, bool inside, float) {
        if (n++ == 0) dda_segment(P, r, maxt, w0, inside, sink);
    }) && n == 0) {}
}

}  // namespace drtvam
