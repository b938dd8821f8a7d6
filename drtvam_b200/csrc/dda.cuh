// Voxel traversal of one medium segment with analytic per-voxel absorption.
// Reference: DDAVolumetricSensor.accumulate (sensor.py:306-440) and
// RatioVolumetricSensor.accumulate (sensor.py:205-295); film.write
// (film.py:40-41) is the Sink of the primal pass, the dr.gather of
// sensor.py:419 the Sink of the adjoint.
#pragma once
#include "device_scene.cuh"
#include "expm.cuh"
#include "rng.cuh"

namespace drtvam {

__device__ __forceinline__ float ex2_approx_dda(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

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

// Adjoint sink of the persistent tracer: the product of a visit is formed one visit later, so the
// thread does not wait for its own gather before it computes the next voxel step.
struct PipelinedGatherSink {
    static constexpr bool kRecordOnly = false;
    const float* dL;
    float acc = 0.f, w_prev = 0.f, v_prev = 0.f;
    __device__ __forceinline__ void visit(size_t idx, float w) {
        acc = fmaf(w_prev, v_prev, acc);
        v_prev = __ldg(dL + idx);
        w_prev = w;
    }
    __device__ __forceinline__ float finish() {
        acc = fmaf(w_prev, v_prev, acc);
        const float r = acc;
        acc = w_prev = v_prev = 0.f;
        return r;
    }
};

// Counting sink (roofline units).
struct CountSink {
    static constexpr bool kRecordOnly = false;
    unsigned long long n = 0;
    __device__ __forceinline__ void visit(size_t, float) { ++n; }
};

// Amanatides-Woo walk over [0, maxt] of `ray`, clipped to the sensor box, as a
// resumable state: begin() clips and sets up, step() deposits one voxel.  The
// general tracer interleaves the steps of the 32 lanes of a warp whatever
// segment each of them is on (general_kernels.cu).
//
// Same walk as sensor.py:328-438, arranged for the issue slots of one thread:
//  - optical depth u = sigma_t log2(e) t, so the deposit of a voxel is
//    w0 (2^-u_in - 2^-u_out): one ex2.approx per visit, the previous one is carried;
//  - boundary times are absolute, u_b = b0 + count * du (no accumulated rounding);
//  - the flat film index is stepped by +-stride of the axis that was crossed;
//  - an axis takes as many steps as separate the start voxel from the end voxel (sensor.py:427)
//    and the walk ends at the segment end u_end.
// An axis whose first boundary time is negative never steps, as sensor.py:359 has it.
struct DdaWalk {
    int idx, sx, sy, sz;               // film index (32 bit) and the signed strides of the three axes
    float ux, uy, uz;                  // next boundary of every axis (inf = none left)
    float fx, fy, fz, lx, ly, lz;      // steps taken / allowed
    float bx, by, bz, dx, dy, dz;      // first boundary and boundary spacing
    float u_end, e_prev, w0;
    float u_prev;                      // kAccurate steps: optical depth at the entry of the current voxel

    // `w0_` is the throughput at the segment origin times sigma_a / sigma_t.  False: no overlap.
    // kSub: the walk covers [t_lo, maxt] of the segment only - the tail of a walk that something else
    // has made in full (inside_kernels.cu) - so a start that rounding puts a hair past a voxel
    // boundary must not freeze the axis (sensor.py:359 applies to the start of a segment).
    template <bool kSub = false>
    __device__ __forceinline__ bool begin(const SceneParams& P, const Ray& ray, float maxt, float w0_,
                                          bool inside_target, float t_lo = 0.f) {
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
        const float t_start = fmaxf(t_near, kSub ? t_lo : 0.f);
        const float t_end = fminf(t_far, maxt);
        if (!(isfinite(t_start) && isfinite(t_end) && t_start < t_end)) return false;

        const float sig = P.sigma_t * 1.4426950408889634f;
        const float u_start = sig * t_start;
        u_end = sig * t_end;
        int cur[3], stp[3];
        float b0[3], du[3], lim[3];
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const float vs = (P.bmax[a] - P.bmin[a]) / static_cast<float>(P.res[a]);
            const float gs = fmaf(d[a], t_start, o[a]);
            const float ge = fmaf(d[a], t_end, o[a]);
            stp[a] = d[a] > 0.f ? 1 : -1;
            // truncation toward zero then clip (sensor.py:345-346)
            cur[a] = min(max(__float2int_rz((gs - P.bmin[a]) / vs), 0), P.res[a] - 1);
            const int endv = min(max(__float2int_rz((ge - P.bmin[a]) / vs), 0), P.res[a] - 1);
            float boundary = P.bmin[a] + static_cast<float>(cur[a] + stp[a]) * vs;
            if (d[a] < 0.f) boundary += vs;
            const bool valid = fabsf(d[a]) > 1e-8f;
            float first = valid ? (boundary - gs) / d[a] : INFINITY;
            if (first < 0.f) first = kSub ? 0.f : INFINITY;  // sensor.py:359: this axis never steps
            const int n = (first < INFINITY) ? max(stp[a] > 0 ? endv - cur[a] : cur[a] - endv, 0) : 0;
            lim[a] = static_cast<float>(n);
            b0[a] = n > 0 ? fmaf(sig, first, u_start) : INFINITY;
            du[a] = valid ? fabsf(sig * vs / d[a]) : 0.f;
        }
        // 32-bit film index: drtvam_scene_create rejects films of 2^31 elements or more
        const int C = P.surface_aware ? 2 : 1;
        sx = C * stp[0]; sy = C * stp[1] * P.res[0]; sz = C * stp[2] * P.res[0] * P.res[1];
        idx = C * (cur[0] + cur[1] * P.res[0] + cur[2] * P.res[0] * P.res[1]) +
              ((P.surface_aware && !inside_target) ? 1 : 0);
        bx = b0[0]; by = b0[1]; bz = b0[2];
        ux = bx; uy = by; uz = bz;
        dx = du[0]; dy = du[1]; dz = du[2];
        lx = lim[0]; ly = lim[1]; lz = lim[2];
        fx = fy = fz = 0.f;
        e_prev = ex2_approx_dda(-u_start);
        u_prev = u_start;
        w0 = w0_;
        return true;
    }

    // Deposits the current voxel and moves to the next one.  False: that was the last voxel.
    // kAccurate: the deposit is 2^-u_in (1 - 2^-(u_out - u_in)) with the small factor evaluated directly
    // (expm.cuh) instead of the difference of two ex2.approx of absolute depths - 5 more instructions per
    // visit, for walks whose results are SUBTRACTED from others (the negative records of
    // inside_kernels.cu), where the 1e-4 noise of the difference form does not average out.
    template <bool kAccurate = false, class Sink>
    __device__ __forceinline__ bool step(Sink& sink) {
        const float un = fminf(fminf(ux, uy), uz);
        const bool more = un < u_end;  // false also when every axis has used up its steps (all inf)
        const float u_out = more ? un : u_end;
        const float e = ex2_approx_dda(-u_out);
        if (kAccurate) {
            sink.visit(static_cast<size_t>(static_cast<unsigned>(idx)), w0 * (e_prev * one_minus_ex2_neg(fmaxf(u_out - u_prev, 0.f))));
            u_prev = u_out;
        } else {
            sink.visit(static_cast<size_t>(static_cast<unsigned>(idx)), w0 * (e_prev - e));
        }
        e_prev = e;
        // branch-free choice of the axis that was crossed (lanes of a warp cross different axes)
        const bool px = ux == un;
        const bool py = !px && (uy == un);
        const bool pz = !px && !py;
        // (written as two selects: nvcc 12.9 drops the y stride from `px ? sx : (py ? sy : sz)`)
        int stride = sz;
        stride = py ? sy : stride;
        stride = px ? sx : stride;
        idx += more ? stride : 0;
        const float fxn = fx + 1.f, fyn = fy + 1.f, fzn = fz + 1.f;
        const float uxn = fxn < lx ? fmaf(fxn, dx, bx) : INFINITY;
        const float uyn = fyn < ly ? fmaf(fyn, dy, by) : INFINITY;
        const float uzn = fzn < lz ? fmaf(fzn, dz, bz) : INFINITY;
        fx = px ? fxn : fx; ux = px ? uxn : ux;
        fy = py ? fyn : fy; uy = py ? uyn : uy;
        fz = pz ? fzn : fz; uz = pz ? uzn : uz;
        return more;
    }
};

template <class Sink>
__device__ __forceinline__ void dda_segment(const SceneParams& P, const Ray& ray, float maxt, float w0,
                                            bool inside_target, Sink& sink) {
    DdaWalk W;
    if (!W.begin(P, ray, maxt, w0, inside_target)) return;
    while (W.step(sink)) {}
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

// Delta-tracking sensor (sensor.py:141-178): one deposit at the collision point of a lane whose free
// flight ended inside the medium.  contrib = throughput * sigma_a * tr / (tr * sigma_t), zero when
// the pdf tr * sigma_t underflows; voxel by floor, points outside the grid dropped (sensor.py:144-146).
template <class Sink>
__device__ __forceinline__ void delta_deposit(const SceneParams& P, const Ray& ray, float t_coll, float w0,
                                              bool inside_target, Sink& sink) {
    if (!(t_coll < INFINITY)) return;  // mei.is_valid()
    if (expf(-P.sigma_t * t_coll) * P.sigma_t == 0.f) return;
    const float p[3] = {fmaf(ray.dx, t_coll, ray.ox), fmaf(ray.dy, t_coll, ray.oy), fmaf(ray.dz, t_coll, ray.oz)};
    int v[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float vs = (P.bmax[a] - P.bmin[a]) / static_cast<float>(P.res[a]);
        const float f = floorf((p[a] - P.bmin[a]) / vs);
        if (!(f >= 0.f && f < static_cast<float>(P.res[a]))) return;
        v[a] = static_cast<int>(f);
    }
    const size_t flat = static_cast<size_t>(v[0]) + static_cast<size_t>(v[1]) * P.res[0] +
                        static_cast<size_t>(v[2]) * P.res[0] * P.res[1];
    sink.visit(P.surface_aware ? (2 * flat + (inside_target ? 0 : 1)) : flat, w0);
}

// Ratio-tracking estimator (sensor.py:243-293): exponential steps with the
// majorant, one deposit per tentative collision.
template <class Sink>
__device__ __forceinline__ void ratio_segment(const SceneParams& P, const Ray& ray, float maxt, float w0,
                                              bool inside_target, Sampler& rng, Sink& sink) {
    const float st = P.sigma_t, mu = P.majorant;
    const float base = w0 * st / mu;
    const float ratio = 1.f - st / mu;
    float t = 0.f, wk = base;  // base * ratio^k as a running product
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
            sink.visit(idx, wk);
        }
        wk *= ratio;
    }
}

}  // namespace drtvam
