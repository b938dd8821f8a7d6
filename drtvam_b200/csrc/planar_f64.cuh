// Double-precision set-up of the planar tables.
//
// The planar kernels run on tables made once per scene: one medium segment per (angle, DMD column) and,
// from it, the clipped voxel walks / coefficient blocks of every tile.  Made in fp32 they carry the
// rounding of positions of ~10 mm (an ulp is 4e-5 of a 25 um voxel) and of refracted directions into
// EVERY deposit of every pass - 5e-5 relative L2 on the dose and 1e-4 on a white-noise gradient at the
// full 400^3 size (profiles/r02a_*), the whole parity budget.  Here the same arithmetic runs in double:
// pixel centre -> world ray (projector.py:184-188, motion.py:26-36), the closed-form vial interfaces with
// Snell / Fresnel (integrators/volume.py:191,228-247 for the null and dielectric BSDFs of
// geometry.py:82-96,152-183,195-219), the spawn offset, up to the first medium segment.  It only serves
// deterministic planar scenes (collimated projector, motion table, no mesh, transmission only): the fp32
// tracer (tracer.cuh) stays the one that decides WHETHER a scene is planar and which rays have a frozen
// DDA axis; this one refines the numbers.  The per-pass kernels keep computing in fp32, on optical depths
// relative to the tile entry, where fp32 is exact enough.
#pragma once
#include "device_scene.cuh"

namespace drtvam {

struct SegmentD {
    double ox, oy, dx, dy, tmax, w0;
};

namespace f64 {

struct RayD {
    double ox, oy, oz, dx, dy, dz;
};

struct HitD {
    double t;
    int shape;
    double px, py, pz, nx, ny, nz;
};

__device__ __forceinline__ bool quadratic_roots(double a, double b, double c, double& lo, double& hi) {
    if (a == 0.0) {
        if (b == 0.0) return false;
        lo = hi = -c / b;
        return true;
    }
    const double disc = b * b - 4.0 * a * c;
    if (disc < 0.0) return false;
    const double q = -0.5 * (b + copysign(sqrt(disc), b));
    double r0 = q / a, r1 = c / q;
    if (q == 0.0) r0 = r1 = 0.0;
    lo = fmin(r0, r1);
    hi = fmax(r0, r1);
    return true;
}

__device__ __forceinline__ double hit_cylinder(const Surface& S, const RayD& r) {
    const double R = static_cast<double>(S.radius);
    const double a = r.dx * r.dx + r.dy * r.dy;
    const double b = 2.0 * (r.dx * r.ox + r.dy * r.oy);
    const double c = r.ox * r.ox + r.oy * r.oy - R * R;
    double tn, tf;
    if (!quadratic_roots(a, b, c, tn, tf)) return INFINITY;
    if (!(tf >= 0.0)) return INFINITY;
    const double zn = r.oz + r.dz * tn, zf = r.oz + r.dz * tf;
    if (tn >= 0.0 && zn >= S.z0 && zn <= S.z1) return tn;
    if (zf >= S.z0 && zf <= S.z1) return tf;
    return INFINITY;
}

__device__ __forceinline__ double hit_box(const Surface& S, const RayD& r, double& nx, double& ny, double& nz) {
    const double h[3] = {static_cast<double>(S.hx), static_cast<double>(S.hy), static_cast<double>(S.hz)};
    const double o[3] = {r.ox, r.oy, r.oz};
    const double d[3] = {r.dx, r.dy, r.dz};
    double t_in = -INFINITY, t_out = INFINITY;
    int a_in = -1, a_out = -1;
    double s_in = 0.0, s_out = 0.0;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        if (d[a] == 0.0) {
            if (o[a] < -h[a] || o[a] > h[a]) return INFINITY;
            continue;
        }
        const double inv = 1.0 / d[a];
        double t0 = (-h[a] - o[a]) * inv, t1 = (h[a] - o[a]) * inv;
        double s0 = -1.0, s1 = 1.0;
        if (t0 > t1) {
            const double tt = t0; t0 = t1; t1 = tt;
            s0 = 1.0; s1 = -1.0;
        }
        if (t0 > t_in) { t_in = t0; a_in = a; s_in = s0; }
        if (t1 < t_out) { t_out = t1; a_out = a; s_out = s1; }
    }
    if (t_in > t_out) return INFINITY;
    nx = ny = nz = 0.0;
    if (t_in >= 0.0 && a_in >= 0) {
        if (a_in == 0) nx = s_in; else if (a_in == 1) ny = s_in; else nz = s_in;
        return t_in;
    }
    if (t_out >= 0.0 && a_out >= 0) {
        if (a_out == 0) nx = s_out; else if (a_out == 1) ny = s_out; else nz = s_out;
        return t_out;
    }
    return INFINITY;
}

__device__ __forceinline__ HitD nearest_surface(const SceneParams& P, const RayD& r) {
    HitD H;
    H.t = INFINITY; H.shape = -1;
    H.px = H.py = H.pz = 0.0;
    H.nx = 0.0; H.ny = 0.0; H.nz = 1.0;
    for (int i = 0; i < P.n_surf; ++i) {
        const Surface& S = P.surf[i];
        double nx = 0.0, ny = 0.0, nz = 1.0, t = INFINITY;
        if (S.kind == 0) t = hit_cylinder(S, r);
        else if (S.kind == 1) t = hit_box(S, r, nx, ny, nz);
        if (t < H.t) {
            H.t = t; H.shape = i;
            H.nx = nx; H.ny = ny; H.nz = nz;
        }
    }
    if (H.shape >= 0) {
        H.px = fma(r.dx, H.t, r.ox);
        H.py = fma(r.dy, H.t, r.oy);
        H.pz = fma(r.dz, H.t, r.oz);
        const Surface& S = P.surf[H.shape];
        if (S.kind == 0) {
            const double rr = sqrt(H.px * H.px + H.py * H.py);
            H.nx = H.px / rr; H.ny = H.py / rr; H.nz = 0.0;
            const double corr = static_cast<double>(S.radius) - rr;
            H.px = fma(H.nx, corr, H.px);
            H.py = fma(H.ny, corr, H.py);
        }
    }
    return H;
}

// transmission through one interface (transmission_only: weight (1 - F) eta_ti^2 for a dielectric); rewrites
// the ray to the spawned one
__device__ __forceinline__ double cross_interface(const SceneParams& P, const HitD& H, RayD& ray) {
    const Surface& S = P.surf[H.shape];
    double wx = ray.dx, wy = ray.dy, wz = ray.dz, weight = 1.0;
    if (S.bsdf == 1) {
        const double eta = static_cast<double>(S.eta);
        const double cos_i = -(H.nx * ray.dx + H.ny * ray.dy + H.nz * ray.dz);
        const bool outside = cos_i >= 0.0;
        const double rcp_eta = 1.0 / eta;
        const double eta_it = outside ? eta : rcp_eta;
        const double eta_ti = outside ? rcp_eta : eta;
        const double cos_t_sq = 1.0 - (1.0 - cos_i * cos_i) * (eta_ti * eta_ti);
        const double ci = fabs(cos_i);
        const double ct = sqrt(fmax(cos_t_sq, 0.0));
        const double a_s = (ci - eta_it * ct) / (ci + eta_it * ct);
        const double a_p = (ct - eta_it * ci) / (ct + eta_it * ci);
        double R = 0.5 * (a_s * a_s + a_p * a_p);
        if (S.eta == 1.f) R = 0.0;
        else if (ci == 0.0) R = 1.0;
        const double cos_t = outside ? -ct : ct;
        const double kk = eta_ti * cos_i + cos_t;
        wx = ray.dx * eta_ti + H.nx * kk;
        wy = ray.dy * eta_ti + H.ny * kk;
        wz = ray.dz * eta_ti + H.nz * kk;
        weight = (1.0 - R) * (eta_ti * eta_ti);
    } else if (S.bsdf == 2) {
        weight = 0.0;
    }
    const double m = fmax(fmax(fabs(H.px), fabs(H.py)), fabs(H.pz));
    double mag = (1.0 + m) * (1500.0 * 5.9604644775390625e-8);
    if (H.nx * wx + H.ny * wy + H.nz * wz < 0.0) mag = -mag;
    ray.ox = fma(H.nx, mag, H.px);
    ray.oy = fma(H.ny, mag, H.py);
    ray.oz = fma(H.nz, mag, H.pz);
    ray.dx = wx; ray.dy = wy; ray.dz = wz;
    return weight;
}

// Centre ray of pixel (k, row, c) of a collimated projector with a motion table, traced to its first
// medium segment.  False when the double-precision walk does not arrive at exactly one planar segment
// (the caller then keeps the promoted fp32 numbers).
__device__ __forceinline__ bool trace_segment(const SceneParams& P, const double* __restrict__ motion, int k, int row,
                                              int c, SegmentD& out) {
    const double w = P.res_x, h = P.res_y;
    const double sx = (static_cast<double>(c) + 0.5) / w, sy = (static_cast<double>(row) + 0.5) / h;
    const double W = w * static_cast<double>(P.pixel_x), H = h * static_cast<double>(P.pixel_y);
    const double lox = (0.5 - sx) * W, loy = (0.5 - sy) * H, loz = 1e-2;
    const double* M = motion + 12 * static_cast<size_t>(k);
    RayD ray;
    ray.ox = M[0] * lox + M[1] * loy + M[2] * loz + M[3];
    ray.oy = M[4] * lox + M[5] * loy + M[6] * loz + M[7];
    ray.oz = M[8] * lox + M[9] * loy + M[10] * loz + M[11];
    ray.dx = M[2]; ray.dy = M[6]; ray.dz = M[10];
    if (ray.dz != 0.0) return false;
    const double absorb = 1.0 - static_cast<double>(P.albedo);   // sigma_a / sigma_t
    double attenuation = 1.0;
    bool in_medium = false;
    for (int depth = 0; depth < P.max_depth; ++depth) {
        if (attenuation == 0.0) return false;
        const HitD Hh = nearest_surface(P, ray);
        if (Hh.shape < 0) return false;
        if (in_medium) {
            if (ray.dz != 0.0 || !(Hh.t > 0.0) || !isfinite(Hh.t)) return false;
            out.ox = ray.ox; out.oy = ray.oy; out.dx = ray.dx; out.dy = ray.dy;
            out.tmax = Hh.t;
            out.w0 = attenuation * absorb;
            return out.w0 != 0.0;
        }
        attenuation *= cross_interface(P, Hh, ray);
        const Surface& S = P.surf[Hh.shape];
        const bool outward = (ray.dx * Hh.nx + ray.dy * Hh.ny + ray.dz * Hh.nz) > 0.0;
        in_medium = outward ? (S.exterior_medium != 0) : (S.interior_medium != 0);
    }
    return false;
}

}  // namespace f64
}  // namespace drtvam
