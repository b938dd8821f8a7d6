// Closed-form vial interfaces: ray/cylinder and ray/box intersection,
// Snell refraction + Fresnel weight, ray re-spawn.
// Replaces scene.ray_intersect / si.bsdf().sample / si.spawn_ray of the
// reference loop (integrators/volume.py:191,230-247) for the vials of
// geometry.py:82-96 (index matched), :152-183 (cylindrical), :195-219 (square),
// :257-308 (double cylindrical); triangle-mesh interfaces (custom vials :98-138,
// occluders :55-72) go through the flat BVH of mesh.cuh.
#pragma once
#include "device_scene.cuh"

namespace drtvam {

struct SurfaceHit {
    float t;           // distance along the ray, +inf when nothing is hit
    int shape;         // index into SceneParams::surf, -1 when nothing is hit
    float px, py, pz;  // hit point
    float nx, ny, nz;  // outward geometric normal
};

// Stable quadratic roots (the larger-magnitude root first, the other by Vieta).
__device__ __forceinline__ bool quadratic_roots(float a, float b, float c, float& lo, float& hi) {
    if (a == 0.f) {
        if (b == 0.f) return false;
        lo = hi = -c / b;
        return true;
    }
    const float disc = b * b - 4.f * a * c;
    if (disc < 0.f) return false;
    const float q = -0.5f * (b + copysignf(sqrtf(disc), b));
    float r0 = q / a, r1 = c / q;
    if (q == 0.f) r0 = r1 = 0.f;
    lo = fminf(r0, r1);
    hi = fmaxf(r0, r1);
    return true;
}

// Open tube about z: nearest root >= 0 whose axial coordinate is in [z0, z1].
__device__ __forceinline__ float hit_cylinder(const Surface& S, const Ray& r) {
    const float a = r.dx * r.dx + r.dy * r.dy;
    const float b = 2.f * (r.dx * r.ox + r.dy * r.oy);
    const float c = r.ox * r.ox + r.oy * r.oy - S.radius * S.radius;
    float tn, tf;
    if (!quadratic_roots(a, b, c, tn, tf)) return INFINITY;
    if (!(tf >= 0.f)) return INFINITY;
    const float zn = r.oz + r.dz * tn, zf = r.oz + r.dz * tf;
    if (tn >= 0.f && zn >= S.z0 && zn <= S.z1) return tn;
    if (zf >= S.z0 && zf <= S.z1) return tf;
    return INFINITY;
}

// Slab test against [-h, h]^3; returns the nearest face crossing with t >= 0
// and its outward normal.
__device__ __forceinline__ float hit_box(const Surface& S, const Ray& r, float& nx, float& ny, float& nz) {
    const float h[3] = {S.hx, S.hy, S.hz};
    const float o[3] = {r.ox, r.oy, r.oz};
    const float d[3] = {r.dx, r.dy, r.dz};
    float t_in = -INFINITY, t_out = INFINITY;
    int a_in = -1, a_out = -1;
    float s_in = 0.f, s_out = 0.f;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        if (d[a] == 0.f) {
            if (o[a] < -h[a] || o[a] > h[a]) return INFINITY;
            continue;
        }
        const float inv = 1.f / d[a];
        float t0 = (-h[a] - o[a]) * inv, t1 = (h[a] - o[a]) * inv;
        float s0 = -1.f, s1 = 1.f;
        if (t0 > t1) {
            const float tt = t0; t0 = t1; t1 = tt;
            s0 = 1.f; s1 = -1.f;
        }
        if (t0 > t_in) { t_in = t0; a_in = a; s_in = s0; }
        if (t1 < t_out) { t_out = t1; a_out = a; s_out = s1; }
    }
    if (t_in > t_out) return INFINITY;
    nx = ny = nz = 0.f;
    if (t_in >= 0.f && a_in >= 0) {
        if (a_in == 0) nx = s_in; else if (a_in == 1) ny = s_in; else nz = s_in;
        return t_in;
    }
    if (t_out >= 0.f && a_out >= 0) {
        if (a_out == 0) nx = s_out; else if (a_out == 1) ny = s_out; else nz = s_out;
        return t_out;
    }
    return INFINITY;
}

__device__ __forceinline__ SurfaceHit nearest_surface(const SceneParams& P, const Ray& r) {
    SurfaceHit H;
    H.t = INFINITY; H.shape = -1;
    H.px = H.py = H.pz = 0.f;
    H.nx = 0.f; H.ny = 0.f; H.nz = 1.f;
    for (int i = 0; i < P.n_surf; ++i) {
        const Surface& S = P.surf[i];
        float nx = 0.f, ny = 0.f, nz = 1.f;
        float t;
        if (S.kind == 0) t = hit_cylinder(S, r);
        else if (S.kind == 1) t = hit_box(S, r, nx, ny, nz);
        else t = mesh_nearest(P.mesh_surf[S.mesh], r.ox, r.oy, r.oz, r.dx, r.dy, r.dz, H.t, nx, ny, nz);  // face normals
        if (t < H.t) {
            H.t = t; H.shape = i;
            H.nx = nx; H.ny = ny; H.nz = nz;
        }
    }
    if (H.shape >= 0) {
        H.px = fmaf(r.dx, H.t, r.ox);
        H.py = fmaf(r.dy, H.t, r.oy);
        H.pz = fmaf(r.dz, H.t, r.oz);
        const Surface& S = P.surf[H.shape];
        if (S.kind == 0) {
            const float rr = sqrtf(H.px * H.px + H.py * H.py);
            H.nx = H.px / rr; H.ny = H.py / rr; H.nz = 0.f;
            const float corr = S.radius - rr;  // pull the point back onto the tube
            H.px = fmaf(H.nx, corr, H.px);
            H.py = fmaf(H.ny, corr, H.py);
        }
    }
    return H;
}

// Unpolarised Fresnel reflectance for a dielectric interface with relative
// index eta = int_ior / ext_ior; cos_i measured against the outward normal.
__device__ __forceinline__ float fresnel_dielectric(float cos_i, float eta, float& cos_t, float& eta_ti) {
    const bool outside = cos_i >= 0.f;
    const float rcp_eta = 1.f / eta;
    const float eta_it = outside ? eta : rcp_eta;
    eta_ti = outside ? rcp_eta : eta;
    const float cos_t_sq = 1.f - (1.f - cos_i * cos_i) * (eta_ti * eta_ti);
    const float ci = fabsf(cos_i);
    const float ct = sqrtf(fmaxf(cos_t_sq, 0.f));
    const float a_s = (ci - eta_it * ct) / (ci + eta_it * ct);
    const float a_p = (ct - eta_it * ci) / (ct + eta_it * ci);
    float R = 0.5f * (a_s * a_s + a_p * a_p);
    if (eta == 1.f) R = 0.f;
    else if (ci == 0.f) R = 1.f;
    cos_t = outside ? -ct : ct;
    return R;
}

// Interface event: returns the BSDF weight and rewrites `ray` to the spawned
// ray.  `u` is the lobe-selection sample (only used when reflections are on).
__device__ __forceinline__ float cross_interface(const SceneParams& P, const SurfaceHit& H, bool allow_reflection,
                                                  float u, Ray& ray) {
    const Surface& S = P.surf[H.shape];
    float wx = ray.dx, wy = ray.dy, wz = ray.dz, weight = 1.f;
    if (S.bsdf == 1) {
        const float cos_i = -(H.nx * ray.dx + H.ny * ray.dy + H.nz * ray.dz);
        float cos_t, eta_ti;
        const float R = fresnel_dielectric(cos_i, S.eta, cos_t, eta_ti);
        const bool reflect = allow_reflection && (u <= R);
        if (reflect) {
            const float k = 2.f * cos_i;
            wx = ray.dx + H.nx * k; wy = ray.dy + H.ny * k; wz = ray.dz + H.nz * k;
        } else {
            const float k = eta_ti * cos_i + cos_t;
            wx = ray.dx * eta_ti + H.nx * k;
            wy = ray.dy * eta_ti + H.ny * k;
            wz = ray.dz * eta_ti + H.nz * k;
            // radiance transport: solid-angle compression eta_ti^2 (Mitsuba dielectric)
            weight = (allow_reflection ? 1.f : (1.f - R)) * (eta_ti * eta_ti);
        }
    } else if (S.bsdf == 2) {
        weight = 0.f;
    }
    // spawn_ray: origin offset (1 + max|p|) * 1500 * 2^-24 along the normal, on the side of w
    const float m = fmaxf(fmaxf(fabsf(H.px), fabsf(H.py)), fabsf(H.pz));
    float mag = (1.f + m) * (1500.0f * 5.9604644775390625e-8f);
    if (H.nx * wx + H.ny * wy + H.nz * wz < 0.f) mag = -mag;
    ray.ox = fmaf(H.nx, mag, H.px);
    ray.oy = fmaf(H.ny, mag, H.py);
    ray.oz = fmaf(H.nz, mag, H.pz);
    ray.dx = wx; ray.dy = wy; ray.dz = wz;
    return weight;
}

// Null-BSDF crossing (the target mesh): direction unchanged, origin re-spawned past the surface.
__device__ __forceinline__ void spawn_through(const SurfaceHit& H, Ray& ray) {
    const float m = fmaxf(fmaxf(fabsf(H.px), fabsf(H.py)), fabsf(H.pz));
    float mag = (1.f + m) * (1500.0f * 5.9604644775390625e-8f);
    if (H.nx * ray.dx + H.ny * ray.dy + H.nz * ray.dz < 0.f) mag = -mag;
    ray.ox = fmaf(H.nx, mag, H.px);
    ray.oy = fmaf(H.ny, mag, H.py);
    ray.oz = fmaf(H.nz, mag, H.pz);
}

}  // namespace drtvam
