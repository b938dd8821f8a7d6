// Projector ray generation: active-pixel index -> world-space ray.
// Reference: TVAMIntegrator.sample_rays (integrators/common.py:70-116),
// TVAMProjector.sample_ray (projector.py:148-165), get_ray of the collimated /
// telecentric / lens projectors (projector.py:184-188, 220-234, 273-286) and
// CircularMotion.eval (motion.py:26-36).
#pragma once
#include "device_scene.cuh"
#include "rng.cuh"

namespace drtvam {

// Shirley-Chiu concentric square -> disk map (projector.py:224,278).
__device__ __forceinline__ void concentric_disk(float u1, float u2, float& x, float& y) {
    const float a = 2.f * u1 - 1.f, b = 2.f * u2 - 1.f;
    const bool quadrant = fabsf(a) < fabsf(b);
    const float r = quadrant ? b : a;
    const float rp = quadrant ? a : b;
    float phi = 0.78539816339744830962f * rp / r;
    if (quadrant) phi = 1.57079632679489661923f - phi;
    if (a == 0.f && b == 0.f) phi = 0.f;
    float s, c;
    sincosf(phi, &s, &c);
    x = r * c;
    y = r * s;
}

// look_at(origin = dist (cos a, sin a, 0), target = 0, up = z): columns
// [left, up, dir, origin] (motion.py:33-36).
__device__ __forceinline__ void circular_to_world(const SceneParams& P, float time, float M[12]) {
    float alpha = 6.28318530717958647692f * time;
    if (P.clockwise) alpha = -alpha;
    float sa, ca;
    sincosf(alpha, &sa, &ca);
    const float ox = P.distance * ca, oy = P.distance * sa;
    const float inv = 1.f / sqrtf(ox * ox + oy * oy);
    const float dx = -ox * inv, dy = -oy * inv;  // dir = normalize(target - origin)
    // left = normalize(cross(up = z, dir)) = normalize(-dy, dx, 0)
    const float ln = 1.f / sqrtf(dy * dy + dx * dx);
    const float lx = -dy * ln, ly = dx * ln;
    // new_up = cross(dir, left) = (0, 0, dx*ly - dy*lx)
    const float uz = dx * ly - dy * lx;
    M[0] = lx; M[1] = 0.f; M[2] = dx; M[3] = ox;
    M[4] = ly; M[5] = 0.f; M[6] = dy; M[7] = oy;
    M[8] = 0.f; M[9] = uz; M[10] = 0.f; M[11] = 0.f;
}

// Generates the ray of one lane and consumes the sampler exactly like
// common.py:92-108: [2 floats pixel jitter] [1 float time] 2 floats aperture.
__device__ __forceinline__ Ray generate_ray(const SceneParams& P, uint32_t pixel, Sampler& rng) {
    const uint32_t w = P.res_x, h = P.res_y;
    const uint32_t hw = h * w;
    const uint32_t k = pixel / hw;
    const uint32_t rem = pixel - k * hw;
    const uint32_t row = rem / w;
    const uint32_t col = rem - row * w;

    float px, py;
    if (P.regular_sampling) {
        px = static_cast<float>(col) + 0.5f;
        py = static_cast<float>(row) + 0.5f;
    } else {
        px = static_cast<float>(col) + rng.next();
        py = static_cast<float>(row) + rng.next();
    }
    const float sx = px * (1.f / static_cast<float>(w));
    const float sy = py * (1.f / static_cast<float>(h));
    float time = static_cast<float>(k);
    if (P.sample_time) time += rng.next();
    time /= static_cast<float>(P.n_patterns);
    const float a1 = rng.next(), a2 = rng.next();  // always drawn (common.py:108)

    const float W = static_cast<float>(P.res_x) * P.pixel_x;
    const float H = static_cast<float>(P.res_y) * P.pixel_y;
    float lox, loy, loz, ldx, ldy, ldz;
    if (P.proj_kind == 0) {
        lox = (0.5f - sx) * W; loy = (0.5f - sy) * H; loz = 1e-2f;
        ldx = 0.f; ldy = 0.f; ldz = 1.f;
    } else if (P.proj_kind == 1) {
        float ax, ay;
        concentric_disk(a1, a2, ax, ay);
        ax *= P.aperture_radius; ay *= P.aperture_radius;
        const float cx = (0.5f - sx) * W, cy = (0.5f - sy) * H;
        lox = cx + ax; loy = cy + ay; loz = 0.f;
        // normalize(origin - (origin.xy + a, -focus)), origin.z = near plane 0.01
        const float vx = cx - lox, vy = cy - loy, vz = 1e-2f + P.focus_distance;
        const float inv = 1.f / sqrtf(vx * vx + vy * vy + vz * vz);
        ldx = vx * inv; ldy = vy * inv; ldz = vz * inv;
    } else {
        const float th = P.pixel_x * static_cast<float>(P.res_x) / (2.f * P.focus_distance);
        const float aspect = static_cast<float>(P.res_x) / static_cast<float>(P.res_y);
        const float nx = (1.f - 2.f * sx) * th * 1e-2f;
        const float ny = (1.f - 2.f * sy) * th * 1e-2f / aspect;
        float ax, ay;
        concentric_disk(a1, a2, ax, ay);
        lox = ax * P.aperture_radius; loy = ay * P.aperture_radius; loz = 0.f;
        const float sc = P.focus_distance / 1e-2f;
        const float vx = nx * sc - lox, vy = ny * sc - loy, vz = 1e-2f * sc;
        const float inv = 1.f / sqrtf(vx * vx + vy * vy + vz * vz);
        ldx = vx * inv; ldy = vy * inv; ldz = vz * inv;
    }

    float M[12];
    if (P.motion_analytic) {
        circular_to_world(P, time, M);
    } else {
        const float4* T = reinterpret_cast<const float4*>(P.motion_table) + 3 * static_cast<size_t>(k);
        const float4 r0 = __ldg(T), r1 = __ldg(T + 1), r2 = __ldg(T + 2);
        M[0] = r0.x; M[1] = r0.y; M[2] = r0.z; M[3] = r0.w;
        M[4] = r1.x; M[5] = r1.y; M[6] = r1.z; M[7] = r1.w;
        M[8] = r2.x; M[9] = r2.y; M[10] = r2.z; M[11] = r2.w;
    }
    Ray r;
    r.ox = M[0] * lox + M[1] * loy + M[2] * loz + M[3];
    r.oy = M[4] * lox + M[5] * loy + M[6] * loz + M[7];
    r.oz = M[8] * lox + M[9] * loy + M[10] * loz + M[11];
    r.dx = M[0] * ldx + M[1] * ldy + M[2] * ldz;
    r.dy = M[4] * ldx + M[5] * ldy + M[6] * ldz;
    r.dz = M[8] * ldx + M[9] * ldy + M[10] * ldz;
    return r;
}

}  // namespace drtvam
