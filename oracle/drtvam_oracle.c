/*
 * drtvam_oracle.c -- CPU restatement of Dr.TVAM's forward projection and its
 * adjoint.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * `--impl reference` legs may load this library.  The product path
 * (drtvam_b200/) never links, imports or calls anything in oracle/.
 *
 * PARITY UNPINNED: the arithmetic of this path lives in Mitsuba 3 (>= 3.6.4)
 * and Dr.Jit, which are not vendored in /root/reference and cannot be
 * installed in this image (no network).  The reference's own tests hold no
 * golden dose volumes or gradients.  What this file is
 * pinned against: the ray-origin bound of tests/test_projector.py:7-38, the
 * adjoint/FD consistency test of tests/test_integrators.py:69-110 (restated
 * in tests/), analytic single-ray and energy identities, the f32 build is
 * compared with the f64 build, orc_discretize reproduces the reference's own
 * tests/files/target_hollow_gear.npy exactly, and the CUDA path held to this
 * oracle passes the nine end-to-end known answers of tests/test_optimization.py
 * at the reference's thresholds.  Absolute dose values stay unpinned.  The
 * Mitsuba pieces are restated from the
 * published Mitsuba 3.6 algorithms (look_at, orthographic/perspective
 * projection, cylinder/cube intersection, dielectric Fresnel + refraction
 * with radiance scaling, spawn_ray offset, homogeneous medium free flight,
 * Henyey-Greenstein / Rayleigh sampling, PCG32 + TEA independent sampler).
 *
 * Each function cites the reference file:line (relative to
 * /root/reference/src/drtvam/) it follows.
 *
 * Build: see oracle/Makefile.  -DORC_REAL=float gives the voxel-exact fp32
 * build (the reference computes in fp32), -DORC_REAL=double the "truth".
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#ifndef ORC_REAL
#define ORC_REAL double
#endif
typedef ORC_REAL real;

#define ORC_MAX_SURF 8

/* ------------------------------------------------------------------ scene */

typedef struct {
    int32_t kind;            /* 0 cylinder (axis = z through origin), 1 axis-aligned box centred at 0,
                              * 2 triangle mesh with face normals (custom vial walls geometry.py:113-134,
                              * occluders geometry.py:55-72) */
    int32_t bsdf;            /* 0 null, 1 dielectric, 2 absorber (diffuse occluder: no transmission lobe) */
    int32_t interior_medium; /* shape.interior_medium() is the printing medium */
    int32_t exterior_medium; /* shape.exterior_medium() is the printing medium */
    double radius, z0, z1;   /* cylinder: geometry.py:85-89,157-160,168-171 */
    double hx, hy, hz;       /* box half extents: geometry.py:200,208 */
    double eta;              /* int_ior / ext_ior (Mitsuba dielectric m_eta) */
    int32_t n_tri;           /* kind 2: world-space triangles, 9 doubles each */
    int32_t _pad_tri;
    const double *tri;
} OrcSurface;

typedef struct {
    /* projector: projector.py:41-296 */
    int32_t proj_kind;       /* 0 collimated, 1 telecentric, 2 lens */
    int32_t n_patterns, res_x, res_y;
    double pixel_x, pixel_y; /* collimated / telecentric pixel size; lens: pixel_x = pixel_y = pixel_size */
    double aperture_radius, focus_distance;
    /* motion: motion.py:19-36; table of n_patterns row-major 4x4 used when !sample_time */
    int32_t motion_analytic; /* 1: circular evaluated per lane from `time` */
    int32_t clockwise;
    double distance;
    const double *motion_table; /* [n_patterns][16] or NULL */
    /* integrator: common.py:10-22, optimize.py:121-128 */
    int32_t max_depth, rr_depth, transmission_only, regular_sampling, sample_time;
    double print_time;
    /* medium: geometry.py:37-45 */
    double sigma_t, albedo;
    int32_t phase_kind;      /* 0 isotropic, 1 rayleigh, 2 hg */
    double hg_g;
    /* surfaces */
    int32_t n_surf;
    OrcSurface surf[ORC_MAX_SURF];
    /* sensor + film: sensor.py:11-19, film.py:9-21 */
    double bbox_min[3], bbox_max[3];
    int32_t res[3];          /* x, y, z as seen by the sensor (film.res) */
    int32_t surface_aware;   /* 2-channel film */
    int32_t sensor_kind;     /* 0 dda, 1 ratio, 2 delta */
    double majorant;
    /* target mesh (optimize.py:66-73: shape 'target', null BSDF), world space, 9 doubles per
     * triangle.  Only present while the run is surface-aware: otherwise optimize.py:188-191 moves it
     * 1e5 away and no ray meets it. */
    int32_t n_tri;
    int32_t _pad_tri;
    const double *tri;
} OrcScene;

#define ORC_SHAPE_TARGET 1000

/* ------------------------------------------------------------------ RNG */
/* Mitsuba independent sampler: one PCG32 per lane seeded through TEA
 * (call sites common.py:46-67,95,104,108). */

typedef struct { uint64_t state, inc; } Pcg32;

static void tea32(uint32_t v0, uint32_t v1, uint32_t *o0, uint32_t *o1) {
    uint32_t sum = 0;
    for (int i = 0; i < 4; ++i) {
        sum += 0x9e3779b9u;
        v0 += ((v1 << 4) + 0xa341316cu) ^ (v1 + sum) ^ ((v1 >> 5) + 0xc8013ea4u);
        v1 += ((v0 << 4) + 0xad90777du) ^ (v0 + sum) ^ ((v0 >> 5) + 0x7e95761eu);
    }
    *o0 = v0; *o1 = v1;
}

static uint32_t pcg_next(Pcg32 *r) {
    uint64_t old = r->state;
    r->state = old * 0x5851f42d4c957f2dull + r->inc;
    uint32_t xs = (uint32_t)(((old >> 18u) ^ old) >> 27u);
    uint32_t rot = (uint32_t)(old >> 59u);
    return (xs >> rot) | (xs << ((0u - rot) & 31u));
}

static void pcg_seed(Pcg32 *r, uint64_t initstate, uint64_t initseq) {
    r->state = 0; r->inc = (initseq << 1u) | 1u;
    pcg_next(r); r->state += initstate; pcg_next(r);
}

static void sampler_seed(Pcg32 *r, uint32_t seed, uint32_t lane) {
    uint32_t v0, v1;
    tea32(seed, lane, &v0, &v1);
    pcg_seed(r, v0, v1);
}

static real next_1d(Pcg32 *r) {
    union { uint32_t u; float f; } x;
    x.u = (pcg_next(r) >> 9) | 0x3f800000u;
    return (real)(x.f - 1.0f);
}

/* ------------------------------------------------------------------ math */

#define RINF ((real)INFINITY)
static real rmin(real a, real b) { return a < b ? a : b; }
static real rmax(real a, real b) { return a > b ? a : b; }
static real rexp(real x) { return sizeof(real) == 4 ? (real)expf((float)x) : (real)exp((double)x); }
static real rsqrt_(real x) { return sizeof(real) == 4 ? (real)sqrtf((float)x) : (real)sqrt((double)x); }
static real rlog(real x) { return sizeof(real) == 4 ? (real)logf((float)x) : (real)log((double)x); }
static real rcbrt(real x) { return sizeof(real) == 4 ? (real)cbrtf((float)x) : (real)cbrt((double)x); }
static real rabs(real x) { return x < 0 ? -x : x; }
static real safe_sqrt(real x) { return rsqrt_(rmax(x, (real)0)); }
static real mulsign(real a, real s) { return (s < 0 || (s == 0 && signbit((double)s))) ? -a : a; }

typedef struct { real x, y, z; } V3;
static V3 v3(real x, real y, real z) { V3 r = {x, y, z}; return r; }
static V3 vadd(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
static V3 vsub(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
static V3 vscale(V3 a, real s) { return v3(a.x * s, a.y * s, a.z * s); }
static real vdot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static V3 vcross(V3 a, V3 b) { return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
static V3 vnorm(V3 a) { real n = rsqrt_(vdot(a, a)); return v3(a.x / n, a.y / n, a.z / n); }

/* ------------------------------------------------------------------ projector + motion */

/* Shirley-Chiu concentric map (Mitsuba warp::square_to_uniform_disk_concentric;
 * call sites projector.py:224,278). */
static void disk_concentric(real u1, real u2, real *dx, real *dy) {
    real x = 2 * u1 - 1, y = 2 * u2 - 1;
    int is_zero = (x == 0 && y == 0);
    int quad = rabs(x) < rabs(y);
    real r = quad ? y : x;
    real rp = quad ? x : y;
    const real pi = (real)3.14159265358979323846;
    real phi = (real)0.25 * pi * rp / r;
    if (quad) phi = (real)0.5 * pi - phi;
    if (is_zero) phi = 0;
    real s = sizeof(real) == 4 ? (real)sinf((float)phi) : (real)sin((double)phi);
    real c = sizeof(real) == 4 ? (real)cosf((float)phi) : (real)cos((double)phi);
    *dx = r * c; *dy = r * s;
}

/* motion.py:26-36 CircularMotion.eval + Mitsuba Transform4f::look_at:
 * columns [left, new_up, dir, origin]. Stored row-major 3x4. */
static void circular_motion(const OrcScene *s, real time, real M[12]) {
    const real pi = (real)3.14159265358979323846;
    real alpha = 2 * pi * time;
    if (s->clockwise) alpha = -alpha;
    real ca = sizeof(real) == 4 ? (real)cosf((float)alpha) : (real)cos((double)alpha);
    real sa = sizeof(real) == 4 ? (real)sinf((float)alpha) : (real)sin((double)alpha);
    V3 origin = v3((real)s->distance * ca, (real)s->distance * sa, 0);
    V3 dir = vnorm(vsub(v3(0, 0, 0), origin));
    V3 left = vnorm(vcross(v3(0, 0, 1), dir));
    V3 up = vcross(dir, left);
    M[0] = left.x; M[1] = up.x; M[2] = dir.x; M[3] = origin.x;
    M[4] = left.y; M[5] = up.y; M[6] = dir.y; M[7] = origin.y;
    M[8] = left.z; M[9] = up.z; M[10] = dir.z; M[11] = origin.z;
}

typedef struct { V3 o, d; } Ray;

/* common.py:70-116 sample_rays + projector.py:148-165 sample_ray +
 * get_ray of each projector (projector.py:184-188, 220-234, 273-286). */
static Ray gen_ray(const OrcScene *s, uint32_t pix, Pcg32 *rng) {
    uint32_t w = (uint32_t)s->res_x, h = (uint32_t)s->res_y;
    uint32_t k = pix / (h * w);
    uint32_t pixel_idx = pix % (h * w);
    uint32_t row = pixel_idx / w;
    uint32_t col = pixel_idx - row * w;
    real px, py;
    if (s->regular_sampling) { px = (real)col + (real)0.5; py = (real)row + (real)0.5; }
    else { px = (real)col + next_1d(rng); py = (real)row + next_1d(rng); }
    real sx = px * ((real)1 / (real)w), sy = py * ((real)1 / (real)h);
    real time = (real)k;
    if (s->sample_time) time += next_1d(rng);
    time /= (real)s->n_patterns;
    real a1 = next_1d(rng), a2 = next_1d(rng); /* aperture sample is always drawn: common.py:108 */

    real W = (real)s->res_x * (real)s->pixel_x, H = (real)s->res_y * (real)s->pixel_y;
    V3 o, d;
    if (s->proj_kind == 0) {
        /* sample_to_camera @ (sx, sy, 0) = ((.5 - sx) W, (.5 - sy) H, near = 1e-2): SURVEY App. A.4 */
        o = v3(((real)0.5 - sx) * W, ((real)0.5 - sy) * H, (real)1e-2);
        d = v3(0, 0, 1);
    } else if (s->proj_kind == 1) {
        V3 org = v3(((real)0.5 - sx) * W, ((real)0.5 - sy) * H, (real)1e-2);
        real ax, ay; disk_concentric(a1, a2, &ax, &ay);
        ax *= (real)s->aperture_radius; ay *= (real)s->aperture_radius;
        o = v3(org.x + ax, org.y + ay, 0);
        V3 moved = v3(org.x + ax, org.y + ay, -(real)s->focus_distance);
        d = vnorm(vsub(org, moved));
    } else {
        /* perspective: near_p = ((1-2sx) tan(fov/2) near, (1-2sy) tan(fov/2) near / aspect, near),
         * tan(fov/2) = pixel_size res_x / (2 focus): projector.py:252-265 */
        real th = (real)s->pixel_x * (real)s->res_x / (2 * (real)s->focus_distance);
        real aspect = (real)s->res_x / (real)s->res_y;
        real near = (real)1e-2;
        V3 near_p = v3((1 - 2 * sx) * th * near, (1 - 2 * sy) * th * near / aspect, near);
        real ax, ay; disk_concentric(a1, a2, &ax, &ay);
        o = v3(ax * (real)s->aperture_radius, ay * (real)s->aperture_radius, 0);
        V3 focus_p = vscale(near_p, (real)s->focus_distance / near_p.z);
        d = vnorm(vsub(focus_p, o));
    }
    real M[12];
    if (s->motion_analytic) circular_motion(s, time, M);
    else { const double *T = s->motion_table + 16 * (size_t)k; for (int i = 0; i < 12; ++i) M[i] = (real)T[i]; }
    Ray r;
    r.o = v3(M[0] * o.x + M[1] * o.y + M[2] * o.z + M[3],
             M[4] * o.x + M[5] * o.y + M[6] * o.z + M[7],
             M[8] * o.x + M[9] * o.y + M[10] * o.z + M[11]);
    r.d = v3(M[0] * d.x + M[1] * d.y + M[2] * d.z,
             M[4] * d.x + M[5] * d.y + M[6] * d.z,
             M[8] * d.x + M[9] * d.y + M[10] * d.z);
    return r;
}

/* ------------------------------------------------------------------ intersection */

typedef struct { int valid; int shape; real t; V3 p, n; } Hit;

/* Mitsuba math::solve_quadratic (numerically stable form). */
static int solve_quadratic(real a, real b, real c, real *x0, real *x1) {
    if (a == 0) { if (b == 0) return 0; *x0 = *x1 = -c / b; return 1; }
    real disc = b * b - 4 * a * c;
    if (disc < 0) return 0;
    real temp = (real)-0.5 * (b + mulsign(rsqrt_(disc), b));
    real r0 = temp / a, r1 = c / temp;
    if (temp == 0) { r0 = r1 = 0; }
    *x0 = rmin(r0, r1); *x1 = rmax(r0, r1);
    return 1;
}

/* Mitsuba cylinder.cpp ray_intersect_preliminary (open tube, hit from either
 * side, nearest non-negative root inside the axial range): SURVEY App. A.7. */
static real isect_cylinder(const OrcSurface *c, Ray r) {
    real A = r.d.x * r.d.x + r.d.y * r.d.y;
    real B = 2 * (r.d.x * r.o.x + r.d.y * r.o.y);
    real C = r.o.x * r.o.x + r.o.y * r.o.y - (real)c->radius * (real)c->radius;
    real tn, tf;
    if (!solve_quadratic(A, B, C, &tn, &tf)) return RINF;
    if (!(tf >= 0)) return RINF;
    real zn = r.o.z + r.d.z * tn, zf = r.o.z + r.d.z * tf;
    int near_ok = zn >= (real)c->z0 && zn <= (real)c->z1 && tn >= 0;
    int far_ok = zf >= (real)c->z0 && zf <= (real)c->z1;
    if (near_ok) return tn;
    if (far_ok) return tf;
    return RINF;
}

/* Mitsuba 'cube' = 12 triangles with outward face normals; closed-form slab
 * equivalent.  Returns nearest t >= 0 and the face normal. */
static real isect_box(const OrcSurface *b, Ray r, V3 *n) {
    real h[3] = {(real)b->hx, (real)b->hy, (real)b->hz};
    real o[3] = {r.o.x, r.o.y, r.o.z}, d[3] = {r.d.x, r.d.y, r.d.z};
    real tmin = -RINF, tmax = RINF; int amin = -1, amax = -1; real smin = 0, smax = 0;
    for (int a = 0; a < 3; ++a) {
        if (d[a] == 0) { if (o[a] < -h[a] || o[a] > h[a]) return RINF; continue; }
        real t0 = (-h[a] - o[a]) / d[a], t1 = (h[a] - o[a]) / d[a];
        real s0 = -1, s1 = 1;
        if (t0 > t1) { real tt = t0; t0 = t1; t1 = tt; s0 = 1; s1 = -1; }
        if (t0 > tmin) { tmin = t0; amin = a; smin = s0; }
        if (t1 < tmax) { tmax = t1; amax = a; smax = s1; }
    }
    if (tmin > tmax) return RINF;
    real nn[3] = {0, 0, 0};
    if (tmin >= 0 && amin >= 0) { nn[amin] = smin; *n = v3(nn[0], nn[1], nn[2]); return tmin; }
    if (tmax >= 0 && amax >= 0) { nn[amax] = smax; *n = v3(nn[0], nn[1], nn[2]); return tmax; }
    return RINF;
}

/* Mitsuba Mesh::ray_intersect_triangle (Moeller-Trumbore): returns t or inf, and the unnormalised
 * geometric normal e1 x e2. */
static real isect_triangle(const double *T, Ray r, V3 *ng) {
    V3 p0 = v3((real)T[0], (real)T[1], (real)T[2]);
    V3 e1 = vsub(v3((real)T[3], (real)T[4], (real)T[5]), p0);
    V3 e2 = vsub(v3((real)T[6], (real)T[7], (real)T[8]), p0);
    V3 pvec = vcross(r.d, e2);
    real det = vdot(e1, pvec);
    if (det == 0) return RINF;
    real inv_det = 1 / det;
    V3 tvec = vsub(r.o, p0);
    real u = vdot(tvec, pvec) * inv_det;
    if (!(u >= 0 && u <= 1)) return RINF;
    V3 qvec = vcross(tvec, e1);
    real v = vdot(r.d, qvec) * inv_det;
    if (!(v >= 0 && u + v <= 1)) return RINF;
    real t = vdot(e2, qvec) * inv_det;
    if (!(t >= 0)) return RINF;
    *ng = vcross(e1, e2);
    return t;
}

/* nearest hit on the target mesh alone (sensor.py:66-69 target_scene, and part of the full scene) */
static real mesh_intersect(const OrcScene *s, Ray r, V3 *n) {
    real best = RINF;
    for (int i = 0; i < s->n_tri; ++i) {
        V3 ng; real t = isect_triangle(s->tri + 9 * (size_t)i, r, &ng);
        if (t < best) { best = t; *n = ng; }
    }
    if (best < RINF) *n = vnorm(*n);
    return best;
}

/* nearest hit on a mesh interface (brute force over its triangles), geometric face normal */
static real surface_mesh_intersect(const OrcSurface *sf, Ray r, V3 *n) {
    real best = RINF;
    for (int i = 0; i < sf->n_tri; ++i) {
        V3 ng; real t = isect_triangle(sf->tri + 9 * (size_t)i, r, &ng);
        if (t < best) { best = t; *n = ng; }
    }
    if (best < RINF) *n = vnorm(*n);
    return best;
}

/* scene.ray_intersect (volume.py:191): nearest hit over the vial surfaces and the target mesh. */
static Hit scene_intersect(const OrcScene *s, Ray r) {
    Hit h; h.valid = 0; h.shape = -1; h.t = RINF; h.p = v3(0, 0, 0); h.n = v3(0, 0, 1);
    for (int i = 0; i < s->n_surf; ++i) {
        const OrcSurface *sf = &s->surf[i];
        V3 n = v3(0, 0, 1); real t;
        if (sf->kind == 0) t = isect_cylinder(sf, r);
        else if (sf->kind == 1) t = isect_box(sf, r, &n);
        else t = surface_mesh_intersect(sf, r, &n);
        if (t < h.t) { h.t = t; h.shape = i; h.valid = 1; if (sf->kind != 0) h.n = n; }
    }
    if (s->n_tri > 0) {
        V3 n = v3(0, 0, 1); real t = mesh_intersect(s, r, &n);
        if (t < h.t) { h.t = t; h.shape = ORC_SHAPE_TARGET; h.valid = 1; h.n = n; h.p = vadd(r.o, vscale(r.d, t)); return h; }
    }
    if (h.valid) {
        h.p = vadd(r.o, vscale(r.d, h.t));
        const OrcSurface *sf = &s->surf[h.shape];
        if (sf->kind == 0) {
            real rr = rsqrt_(h.p.x * h.p.x + h.p.y * h.p.y);
            h.n = v3(h.p.x / rr, h.p.y / rr, 0);
            /* cylinder.cpp: normal shift of the hit point onto the surface */
            real corr = (real)sf->radius - rr;
            h.p = vadd(h.p, vscale(h.n, corr));
        }
    }
    return h;
}

/* Mitsuba fresnel() + dielectric.cpp sample() + refract()/reflect():
 * SURVEY App. A.9.  Returns world-space wo and the BSDF weight. */
static real fresnel(real cos_i, real eta, real *cos_t, real *eta_it, real *eta_ti) {
    int outside = cos_i >= 0;
    real rcp_eta = 1 / eta;
    *eta_it = outside ? eta : rcp_eta;
    *eta_ti = outside ? rcp_eta : eta;
    real cos_t_sqr = 1 - (1 - cos_i * cos_i) * (*eta_ti * *eta_ti);
    real ci = rabs(cos_i), ct = safe_sqrt(cos_t_sqr);
    real a_s = (ci - *eta_it * ct) / (ci + *eta_it * ct);
    real a_p = (ct - *eta_it * ci) / (ct + *eta_it * ci);
    real r = (real)0.5 * (a_s * a_s + a_p * a_p);
    if (eta == 1) r = 0; else if (ci == 0) r = 1;
    *cos_t = cos_i >= 0 ? -ct : ct; /* mulsign_neg */
    return r;
}

/* spawn_ray: SurfaceInteraction::offset_p, RayEpsilon = 1500 * 2^-24 (App. A.8). */
static V3 offset_p(V3 p, V3 n, V3 d) {
    real m = rmax(rmax(rabs(p.x), rabs(p.y)), rabs(p.z));
    real eps = sizeof(real) == 4 ? (real)(1500.0f * 5.9604644775390625e-8f) : (real)(1500.0 * 5.9604644775390625e-8);
    real mag = (1 + m) * eps;
    if (vdot(n, d) < 0) mag = -mag;
    return vadd(p, vscale(n, mag));
}

/* Branchless ONB (Mitsuba coordinate_system, Duff et al.). */
static void onb(V3 n, V3 *s, V3 *t) {
    real sign = n.z >= 0 ? (real)1 : (real)-1;
    real a = -1 / (sign + n.z);
    real b = n.x * n.y * a;
    *s = v3(n.x * n.x * a * sign + 1, b * sign, -n.x * sign); /* mulsign */
    *t = v3(b, n.y * n.y * a + sign, -n.y);
}

/* ------------------------------------------------------------------ sensor (DDA) */

typedef struct {
    const OrcScene *s;
    real *film;        /* primal: scatter target */
    const real *dL;    /* backward: gathered */
    int mode;          /* 0 primal, 1 backward */
    uint64_t visits;
} Acc;

static void film_add(real *film, size_t idx, real v) {
#ifdef _OPENMP
#pragma omp atomic
#endif
    film[idx] += v;
}

/* sensor.py:306-440 DDAVolumetricSensor.accumulate, op for op. Returns g_em. */
static real dda_accumulate(Acc *A, Ray ray, real em, int inside_target, real attenuation,
                           real t_prev, real n_scat, real maxt) {
    const OrcScene *s = A->s;
    real bmin[3] = {(real)s->bbox_min[0], (real)s->bbox_min[1], (real)s->bbox_min[2]};
    real bmax[3] = {(real)s->bbox_max[0], (real)s->bbox_max[1], (real)s->bbox_max[2]};
    real o[3] = {ray.o.x, ray.o.y, ray.o.z}, d[3] = {ray.d.x, ray.d.y, ray.d.z};
    int res[3] = {s->res[0], s->res[1], s->res[2]};
    real vs[3];
    for (int a = 0; a < 3; ++a) vs[a] = (bmax[a] - bmin[a]) / (real)res[a];

    real mint_box = -RINF, maxt_box = RINF;
    for (int a = 0; a < 3; ++a) {
        real t0 = (bmin[a] - o[a]) / d[a], t1 = (bmax[a] - o[a]) / d[a];
        real lo = t0 < t1 ? t0 : t1, hi = t0 > t1 ? t0 : t1; /* NaN falls through like dr.minimum */
        if (t0 != t0 || t1 != t1) { lo = t0 != t0 ? t1 : t0; hi = lo; }
        if (lo > mint_box) mint_box = lo;
        if (hi < maxt_box) maxt_box = hi;
    }
    mint_box = rmax(mint_box, 0);
    real t_start = rmax(mint_box, 0), t_end = rmin(maxt_box, maxt);
    if (!(isfinite((double)t_start) && isfinite((double)t_end) && t_start < t_end)) return 0;

    real gs[3], ge[3]; int step[3], cur[3], endv[3]; real dtmax[3], tstep[3];
    for (int a = 0; a < 3; ++a) {
        gs[a] = o[a] + d[a] * t_start; /* ray(t) = fma(d, t, o) */
        ge[a] = o[a] + d[a] * t_end;
        step[a] = d[a] > 0 ? 1 : -1;
        int sv = (int)((gs[a] - bmin[a]) / vs[a]); /* truncation toward zero: sensor.py:345 */
        int ev = (int)((ge[a] - bmin[a]) / vs[a]);
        cur[a] = sv < 0 ? 0 : (sv > res[a] - 1 ? res[a] - 1 : sv);
        endv[a] = ev < 0 ? 0 : (ev > res[a] - 1 ? res[a] - 1 : ev);
        real next = bmin[a] + (real)(cur[a] + step[a]) * vs[a];
        if (d[a] < 0) next += vs[a];
        int valid = rabs(d[a]) > (real)1e-8;
        dtmax[a] = valid ? (next - gs[a]) / d[a] : RINF;
        if (dtmax[a] < 0) dtmax[a] = RINF;
        tstep[a] = valid ? vs[a] / d[a] * (real)step[a] : RINF;
    }
    real t = t_start, remaining = t_end - t_start;
    real st = (real)s->sigma_t, ss = (real)s->sigma_t * (real)s->albedo;
    real sa = st - ss;
    /* sensor.py:368-369, 401-402: throughput*exp(-st t_prev) == attenuation up to rounding */
    real throughput = attenuation * rexp(st * t_prev);
    if (ss != 0) throughput *= 1 / (sizeof(real) == 4 ? (real)powf((float)ss, (float)n_scat) : (real)pow((double)ss, (double)n_scat));
    real weight = throughput * rexp(-st * t_prev);
    if (ss != 0) weight *= (sizeof(real) == 4 ? (real)powf((float)ss, (float)n_scat) : (real)pow((double)ss, (double)n_scat));

    real g_em = 0;
    int active = 1;
    while (active) {
        real dt = rmin(rmin(rmin(dtmax[0], dtmax[1]), dtmax[2]), remaining);
        remaining -= dt;
        real cw = weight * sa / st * rexp(-st * t) * (1 - rexp(-st * rmax(dt, 0)));
        size_t flat = (size_t)cur[0] + (size_t)cur[1] * (size_t)res[0] + (size_t)cur[2] * (size_t)res[0] * (size_t)res[1];
        size_t idx = s->surface_aware ? (inside_target ? 2 * flat : 2 * flat + 1) : flat;
        A->visits++;
        if (A->mode == 0) film_add(A->film, idx, cw * em);
        else g_em += cw * A->dL[idx];
        active = (endv[0] != cur[0] || endv[1] != cur[1] || endv[2] != cur[2]) && (remaining > (real)1e-6);
        int upd[3];
        for (int a = 0; a < 3; ++a) {
            int m = dtmax[a] == dt;
            dtmax[a] = m ? tstep[a] : dtmax[a] - dt;
            upd[a] = m ? step[a] : 0;
        }
        if (active) for (int a = 0; a < 3; ++a) cur[a] += upd[a];
        for (int a = 0; a < 3; ++a) if (cur[a] < 0 || cur[a] >= res[a]) active = 0;
        if (active) t += dt;
    }
    return g_em;
}

/* sensor.py:205-295 RatioVolumetricSensor.accumulate. */
static real ratio_accumulate(Acc *A, Ray ray, real em, int inside_target, real attenuation,
                             real t_prev, real n_scat, real maxt, Pcg32 *rng) {
    const OrcScene *s = A->s;
    real st = (real)s->sigma_t, ss = st * (real)s->albedo, sa = st - ss, mu = (real)s->majorant;
    real throughput = attenuation * rexp(st * t_prev);
    real weight = throughput * rexp(-st * t_prev);
    (void)n_scat;
    real t = 0, g_em = 0, n_int = 0;
    real bmin[3] = {(real)s->bbox_min[0], (real)s->bbox_min[1], (real)s->bbox_min[2]};
    real bmax[3] = {(real)s->bbox_max[0], (real)s->bbox_max[1], (real)s->bbox_max[2]};
    for (;;) {
        real st_ = -rlog(1 - next_1d(rng)) / mu;
        t += st_;
        if (!(t < maxt)) break;
        real p[3] = {ray.o.x + ray.d.x * t, ray.o.y + ray.d.y * t, ray.o.z + ray.d.z * t};
        int v[3], in = 1;
        for (int a = 0; a < 3; ++a) {
            real vs = (bmax[a] - bmin[a]) / (real)s->res[a];
            v[a] = (int)floor((double)((p[a] - bmin[a]) / vs));
            if (v[a] < 0 || v[a] >= s->res[a]) in = 0;
        }
        if (in) {
            size_t flat = (size_t)v[0] + (size_t)v[1] * s->res[0] + (size_t)v[2] * s->res[0] * s->res[1];
            size_t idx = s->surface_aware ? (inside_target ? 2 * flat : 2 * flat + 1) : flat;
            real cw = weight * sa / st * (sizeof(real) == 4 ? (real)powf((float)(1 - st / mu), (float)n_int) : (real)pow((double)(1 - st / mu), (double)n_int)) * st / mu;
            A->visits++;
            if (A->mode == 0) film_add(A->film, idx, cw * em); else g_em += cw * A->dL[idx];
        }
        n_int += 1;
    }
    return g_em;
}

/* sensor.py:112-191 DeltaVolumetricSensor.accumulate: one deposit at the collision point ray(mei.t)
 * of the lanes whose free flight ended inside the medium (mei.is_valid(), sensor.py:141):
 * contrib = throughput * sigma_a * tr / (tr * sigma_t) * em (sensor.py:169-178), zero when the pdf
 * tr * sigma_t underflows.  Voxel by floor, lanes outside the grid dropped (sensor.py:144-146). */
static real delta_accumulate(Acc *A, Ray ray, real em, int inside_target, real attenuation,
                             real t_prev, real mei_t) {
    const OrcScene *s = A->s;
    if (!(mei_t < RINF)) return 0;
    real st = (real)s->sigma_t, ss = st * (real)s->albedo, sa = st - ss;
    real tr = rexp(-st * mei_t), tr_pdf = tr * st;
    real inv_pdf = tr_pdf != 0 ? 1 / tr_pdf : 0;
    real throughput = attenuation * rexp(st * t_prev) * rexp(-st * t_prev);
    real p[3] = {ray.o.x + ray.d.x * mei_t, ray.o.y + ray.d.y * mei_t, ray.o.z + ray.d.z * mei_t};
    size_t v[3];
    for (int a = 0; a < 3; ++a) {
        real lo = (real)s->bbox_min[a], hi = (real)s->bbox_max[a];
        real vs = (hi - lo) / (real)s->res[a];
        double f = floor((double)((p[a] - lo) / vs));
        if (!(f >= 0 && f < (double)s->res[a])) return 0;
        v[a] = (size_t)f;
    }
    size_t flat = v[0] + v[1] * s->res[0] + v[2] * (size_t)s->res[0] * s->res[1];
    size_t idx = s->surface_aware ? (inside_target ? 2 * flat : 2 * flat + 1) : flat;
    real cw = throughput * sa * tr * inv_pdf;
    A->visits++;
    if (A->mode == 0) { film_add(A->film, idx, cw * em); return 0; }
    return cw * A->dL[idx];
}

/* ------------------------------------------------------------------ integrator */

/* volume.py:136-282 VolumeIntegrator.sample for one lane.  Returns em_grad
 * (backward) and deposits into the film (primal). */
static real trace_lane(Acc *A, Ray ray, real Le, Pcg32 *rng) {
    const OrcScene *s = A->s;
    real st = (real)s->sigma_t, ss = st * (real)s->albedo;
    int has_scattering = ss != 0;
    real attenuation = 1, total_t = 0, n_scat = 0, em_grad = 0;
    int inside_target = 0, active = 1, active_medium = 0;
    uint32_t depth = 0;
    while (active) {
        real q = rmin((real)0.99, attenuation);
        int perform_rr = depth > (uint32_t)s->rr_depth;
        real u = next_1d(rng);
        active = active && ((u < q) || !perform_rr);
        if (perform_rr) attenuation *= 1 / q;
        active = active && (attenuation != 0);
        active_medium = active_medium && active;

        Hit si; si.valid = 0; si.t = RINF; si.shape = -1; si.p = v3(0, 0, 0); si.n = v3(0, 0, 1);
        if (active) si = scene_intersect(s, ray);
        active = active && si.valid;
        active_medium = active_medium && active;
        int hit_target = active && si.shape == ORC_SHAPE_TARGET; /* volume.py:196 */

        real weight = 1, mei_t = RINF;
        int reached_surface;
        if (has_scattering) {
            real xi = next_1d(rng);
            if (active_medium) mei_t = -rlog(1 - xi) / st;
            reached_surface = active_medium && si.valid && (si.t < mei_t);
            if (reached_surface) mei_t = RINF;
            real tmin_ = rmin(mei_t, si.t);
            real tr = rexp(-tmin_ * st);
            real pdf = si.t < mei_t ? tr : tr * st;
            real inv_pdf = pdf > 0 ? 1 / pdf : 0;
            if (active_medium) weight *= tr * inv_pdf;
        } else {
            reached_surface = active_medium;
        }
        active_medium = active_medium && !reached_surface;

        if (active_medium || reached_surface) {
            real g;
            if (s->sensor_kind == 2) g = delta_accumulate(A, ray, Le, inside_target, attenuation, total_t, active_medium ? mei_t : RINF);
            else if (s->sensor_kind == 1) g = ratio_accumulate(A, ray, Le, inside_target, attenuation, total_t, n_scat, si.t, rng);
            else g = dda_accumulate(A, ray, Le, inside_target, attenuation, total_t, n_scat, si.t);
            em_grad += g;
        }
        inside_target = inside_target != hit_target;

        int active_surface = active && si.valid && !active_medium;
        Ray new_ray = ray;
        if (active_surface) {
            real s1 = next_1d(rng); (void)next_1d(rng); (void)next_1d(rng);
            const OrcSurface *sf = hit_target ? NULL : &s->surf[si.shape];
            V3 wo; real bw;
            if (hit_target || sf->bsdf == 0) { wo = ray.d; bw = 1; } /* null BSDF: optimize.py:70-72 */
            else if (sf->bsdf == 2) { wo = ray.d; bw = 0; }
            else {
                real cos_i = -vdot(si.n, ray.d), cos_t, eta_it, eta_ti;
                real F = fresnel(cos_i, (real)sf->eta, &cos_t, &eta_it, &eta_ti);
                int reflect = 0;
                if (!s->transmission_only && depth != 0) reflect = s1 <= F;
                if (reflect) {
                    wo = vadd(ray.d, vscale(si.n, 2 * cos_i));
                    bw = 1;
                } else {
                    /* to_world(refract(wi, cos_t, eta_ti)) = eta_ti d + n (eta_ti cos_i + cos_t) */
                    wo = vadd(vscale(ray.d, eta_ti), vscale(si.n, eta_ti * cos_i + cos_t));
                    bw = (s->transmission_only || depth == 0) ? (1 - F) : (real)1;
                    bw *= eta_ti * eta_ti;
                }
            }
            weight *= bw;
            new_ray.o = offset_p(si.p, si.n, wo);
            new_ray.d = wo;
        }
        if (has_scattering) {
            if (active_medium) {
                weight *= ss;
                (void)next_1d(rng); /* phase.sample sample1 */
                real u1 = next_1d(rng), u2 = next_1d(rng);
                real cos_theta;
                if (s->phase_kind == 2) {
                    real g = (real)s->hg_g;
                    real sq = (1 - g * g) / (1 - g + 2 * g * u1);
                    cos_theta = (1 + g * g - sq * sq) / (2 * g);
                    if (rabs(g) < (real)1e-7) cos_theta = 1 - 2 * u1;
                    cos_theta = -cos_theta; /* hg.cpp: wo = to_world(.., .., -cos_theta) in the frame of wi = -d */
                } else if (s->phase_kind == 1) {
                    real z = 2 * (2 * u1 - 1), tmp = rsqrt_(z * z + 1);
                    cos_theta = rcbrt(z + tmp) + rcbrt(z - tmp);
                } else cos_theta = 1 - 2 * u1;
                real sin_theta = safe_sqrt(1 - cos_theta * cos_theta);
                const real pi = (real)3.14159265358979323846;
                real phi = 2 * pi * u2;
                real sp = sizeof(real) == 4 ? (real)sinf((float)phi) : (real)sin((double)phi);
                real cp = sizeof(real) == 4 ? (real)cosf((float)phi) : (real)cos((double)phi);
                V3 wi = vscale(ray.d, -1), fs, ft; onb(wi, &fs, &ft);
                V3 wo = vadd(vadd(vscale(fs, sin_theta * cp), vscale(ft, sin_theta * sp)), vscale(wi, cos_theta));
                new_ray.o = vadd(ray.o, vscale(ray.d, mei_t));
                new_ray.d = wo;
                n_scat += 1;
            }
        } else if (active && reached_surface) {
            weight *= rexp(-st * si.t);
        }
        if (active_medium || reached_surface) total_t += reached_surface ? si.t : mei_t;
        if (active) attenuation *= weight;
        ray = new_ray;
        if (active_surface) {
            if (hit_target) active_medium = 1; /* volume.py:268 */
            else {
                const OrcSurface *sf = &s->surf[si.shape];
                int transition = sf->interior_medium || sf->exterior_medium;
                int target_is_medium = vdot(ray.d, si.n) > 0 ? sf->exterior_medium : sf->interior_medium;
                if (transition && target_is_medium) active_medium = 1;
            }
        }
        active = active && (active_surface || active_medium);
        if (active && !hit_target) depth += 1;
        active = active && (depth < (uint32_t)s->max_depth);
    }
    return em_grad;
}

/* ------------------------------------------------------------------ entry points */

static real ray_weight(const OrcScene *s, uint64_t n_active, uint32_t spp) {
    /* projector.py:187,164-165 and common.py:111: pixel_area * n_active / (n_active spp) * print_time */
    real area = (real)s->pixel_x * (real)s->pixel_y;
    real inv_pdf = area * (real)n_active;
    return inv_pdf / (real)(n_active * (uint64_t)spp) * (real)s->print_time;
}

int orc_sizeof_real(void) { return (int)sizeof(real); }

/* volume.py:18-56 render (without the inv_vol scale, which the caller applies). */
int orc_forward(const OrcScene *s, const float *active_data, const uint32_t *active_pixels,
                uint64_t n_active, uint32_t spp, uint32_t seed, real *film, uint64_t *n_visits, int n_threads) {
    if (s->regular_sampling) spp = 1; /* common.py:49-51 */
    uint64_t lanes = n_active * spp, visits = 0;
    real w = ray_weight(s, n_active, spp);
#ifdef _OPENMP
    if (n_threads > 0) omp_set_num_threads(n_threads);
#pragma omp parallel for schedule(dynamic, 4096) reduction(+ : visits)
#endif
    for (int64_t lane = 0; lane < (int64_t)lanes; ++lane) {
        uint64_t j = (uint64_t)lane / spp;
        Pcg32 rng; sampler_seed(&rng, seed, (uint32_t)lane);
        Ray ray = gen_ray(s, active_pixels[j], &rng);
        Acc A; A.s = s; A.film = film; A.dL = NULL; A.mode = 0; A.visits = 0;
        trace_lane(&A, ray, (real)active_data[j] * w, &rng);
        visits += A.visits;
    }
    if (n_visits) *n_visits = visits;
    return 0;
}

/* volume.py:97-134 render_backward; dL is grad_in * inv_vol. */
int orc_adjoint(const OrcScene *s, const real *dL, const uint32_t *active_pixels,
                uint64_t n_active, uint32_t spp, uint32_t seed, real *grad, int n_threads) {
    if (s->regular_sampling) spp = 1;
    real w = ray_weight(s, n_active, spp);
#ifdef _OPENMP
    if (n_threads > 0) omp_set_num_threads(n_threads);
#pragma omp parallel for schedule(dynamic, 1024)
#endif
    for (int64_t j = 0; j < (int64_t)n_active; ++j) {
        real g = 0;
        for (uint32_t sidx = 0; sidx < spp; ++sidx) {
            uint64_t lane = (uint64_t)j * spp + sidx;
            Pcg32 rng; sampler_seed(&rng, seed, (uint32_t)lane);
            Ray ray = gen_ray(s, active_pixels[j], &rng);
            Acc A; A.s = s; A.film = NULL; A.dL = dL; A.mode = 1; A.visits = 0;
            /* em_grad is d/d(Le); Le = L * w, and contrib is linear in em: evaluate with em = 1 */
            g += trace_lane(&A, ray, 1, &rng) * w;
        }
        grad[j] = g;
    }
    return 0;
}

/* common.py:70-116: the rays alone (tests/test_projector.py). out = [n_active*spp][6]. */
/* integrators/radon.py:48-106 RadonIntegrator.sample for one lane: the absorbed fraction along the
 * parts of the path that lie inside the target mesh (purely absorptive assumption, radon.py:76).
 * No Russian roulette, no free flight; `depth` of radon.py:93 is the never-incremented argument,
 * so every interface is sampled with the transmission lobe only. */
static real radon_lane(const OrcScene *s, Ray ray, Pcg32 *rng) {
    real st = (real)s->sigma_t;
    int active = 1, active_medium = 0, inside_target = 0;
    real throughput = 1, L = 0, t = 0;
    uint32_t d = 0;
    while (active) {
        Hit si = scene_intersect(s, ray);
        active = active && si.valid;
        real seg = si.valid ? si.t : RINF;
        real contrib = throughput * rexp(-st * t) * (1 - rexp(-st * seg));
        int hit_target = active && si.shape == ORC_SHAPE_TARGET;
        if (inside_target && active_medium) L += contrib; /* radon.py:79: not masked by `active` */
        if (!active) break;
        t += si.t;
        inside_target = inside_target != hit_target;
        (void)next_1d(rng); (void)next_1d(rng); (void)next_1d(rng);
        const OrcSurface *sf = hit_target ? NULL : &s->surf[si.shape];
        V3 wo = ray.d; real bw = 1;
        if (!hit_target && sf->bsdf == 2) bw = 0;
        else if (!hit_target && sf->bsdf == 1) {
            real cos_i = -vdot(si.n, ray.d), cos_t, eta_it, eta_ti;
            real F = fresnel(cos_i, (real)sf->eta, &cos_t, &eta_it, &eta_ti);
            wo = vadd(vscale(ray.d, eta_ti), vscale(si.n, eta_ti * cos_i + cos_t));
            bw = (1 - F) * eta_ti * eta_ti;
        }
        throughput *= bw;
        ray.o = offset_p(si.p, si.n, wo);
        ray.d = wo;
        if (!hit_target) d += 1;
        active = d < (uint32_t)s->max_depth;
        if (!hit_target) { /* radon.py:104: only shapes that carry a medium are transitions */
            int outward = vdot(wo, si.n) > 0;
            if (sf->interior_medium != sf->exterior_medium)
                active_medium = outward ? sf->exterior_medium != 0 : sf->interior_medium != 0;
        }
    }
    return L;
}

/* RadonIntegrator.render (radon.py:13-46): image[pixel] += L * weight over the spp lanes of a pixel.
 * `image` has n_patterns * res_y * res_x entries and is cleared first. */
int orc_radon(const OrcScene *s, const uint32_t *active_pixels, uint64_t n_active, uint32_t spp, uint32_t seed,
              real *image) {
    if (s->regular_sampling) spp = 1;
    real w = ray_weight(s, n_active, spp);
    size_t n_img = (size_t)s->n_patterns * s->res_x * s->res_y;
    for (size_t i = 0; i < n_img; ++i) image[i] = 0;
    for (uint64_t lane = 0; lane < n_active * spp; ++lane) {
        Pcg32 rng; sampler_seed(&rng, seed, (uint32_t)lane);
        uint32_t px = active_pixels[lane / spp];
        Ray ray = gen_ray(s, px, &rng);
        image[px] += radon_lane(s, ray, &rng) * w;
    }
    return 0;
}

/* CornerIntegrator (filter_corner.py:17-66): 1 * weight unless the first hit lies within `radius` of
 * the vertical edge |x| = |y| = dist of a square vial. */
int orc_corner(const OrcScene *s, const uint32_t *active_pixels, uint64_t n_active, uint32_t spp, uint32_t seed,
               double dist, double radius, real *image) {
    if (s->regular_sampling) spp = 1;
    real w = ray_weight(s, n_active, spp);
    size_t n_img = (size_t)s->n_patterns * s->res_x * s->res_y;
    for (size_t i = 0; i < n_img; ++i) image[i] = 0;
    for (uint64_t lane = 0; lane < n_active * spp; ++lane) {
        Pcg32 rng; sampler_seed(&rng, seed, (uint32_t)lane);
        uint32_t px = active_pixels[lane / spp];
        Ray ray = gen_ray(s, px, &rng);
        Hit si = scene_intersect(s, ray);
        if (!si.valid) continue;
        real ex = rabs(si.p.x) - (real)dist, ey = rabs(si.p.y) - (real)dist;
        real nrm = sizeof(real) == 4 ? (real)sqrtf((float)(ex * ex + ey * ey)) : (real)sqrt((double)(ex * ex + ey * ey));
        if (!(nrm < (real)radius)) image[px] += w;
    }
    return 0;
}

int orc_raygen(const OrcScene *s, const uint32_t *active_pixels, uint64_t n_active, uint32_t spp,
               uint32_t seed, real *out) {
    if (s->regular_sampling) spp = 1;
    for (uint64_t lane = 0; lane < n_active * spp; ++lane) {
        Pcg32 rng; sampler_seed(&rng, seed, (uint32_t)lane);
        Ray r = gen_ray(s, active_pixels[lane / spp], &rng);
        real *o = out + 6 * lane;
        o[0] = r.o.x; o[1] = r.o.y; o[2] = r.o.z; o[3] = r.d.x; o[4] = r.d.y; o[5] = r.d.z;
    }
    return 0;
}

/* sensor.py:47-110 VolumetricSensor.compute_volume, surface-aware branch: per voxel, `sample_count`
 * random rays from random points of the voxel; a point is inside when the first hit on the target
 * is a back face.  volumes = [rz][ry][rx][2] (in, out) in units of volume. */
int orc_compute_volume(const OrcScene *s, uint32_t sample_count, real *volumes, int n_threads) {
    if (s->n_tri <= 0) return -1;
    real mbmin[3] = {RINF, RINF, RINF}, mbmax[3] = {-RINF, -RINF, -RINF};
    for (int i = 0; i < s->n_tri * 3; ++i)
        for (int a = 0; a < 3; ++a) {
            real v = (real)s->tri[3 * (size_t)i + a];
            if (v < mbmin[a]) mbmin[a] = v;
            if (v > mbmax[a]) mbmax[a] = v;
        }
    real bmin[3], vs[3];
    for (int a = 0; a < 3; ++a) {
        bmin[a] = (real)s->bbox_min[a];
        vs[a] = ((real)s->bbox_max[a] - (real)s->bbox_min[a]) / (real)s->res[a];
    }
    real voxel_vol = vs[0] * vs[1] * vs[2];
    int64_t n_vox = (int64_t)s->res[0] * s->res[1] * s->res[2];
#ifdef _OPENMP
    if (n_threads > 0) omp_set_num_threads(n_threads);
#pragma omp parallel for schedule(dynamic, 256)
#endif
    for (int64_t idx = 0; idx < n_vox; ++idx) {
        int x = (int)(idx % s->res[0]), y = (int)((idx / s->res[0]) % s->res[1]), z = (int)(idx / ((int64_t)s->res[0] * s->res[1]));
        Pcg32 rng; sampler_seed(&rng, 0u, (uint32_t)idx); /* sensor.py:83-84 */
        uint32_t cin = 0, cout = 0;
        for (uint32_t d = 0; d < sample_count; ++d) {
            real o0 = next_1d(&rng), o1 = next_1d(&rng), o2 = next_1d(&rng);
            real u0 = next_1d(&rng), u1 = next_1d(&rng);
            Ray r;
            r.o = v3(bmin[0] + vs[0] * ((real)x + o0), bmin[1] + vs[1] * ((real)y + o1), bmin[2] + vs[2] * ((real)z + o2));
            /* warp::square_to_uniform_sphere */
            real cz = 1 - 2 * u1, rr = safe_sqrt(1 - cz * cz);
            const real pi = (real)3.14159265358979323846;
            real phi = 2 * pi * u0;
            real sp = sizeof(real) == 4 ? (real)sinf((float)phi) : (real)sin((double)phi);
            real cp = sizeof(real) == 4 ? (real)cosf((float)phi) : (real)cos((double)phi);
            r.d = v3(rr * cp, rr * sp, cz);
            int in_bbox = r.o.x > mbmin[0] && r.o.x < mbmax[0] && r.o.y > mbmin[1] && r.o.y < mbmax[1] &&
                          r.o.z > mbmin[2] && r.o.z < mbmax[2];
            int inside = 0;
            if (in_bbox) {
                V3 n; real t = mesh_intersect(s, r, &n);
                inside = t < RINF && vdot(r.d, n) > 0;
            }
            if (inside) ++cin; else ++cout;
        }
        volumes[2 * idx] = (real)cin * voxel_vol / (real)sample_count;
        volumes[2 * idx + 1] = (real)cout * voxel_vol / (real)sample_count;
    }
    return 0;
}

/* utils.py:83-128 discretize: voxels[rz][ry][rx] = 1 where the voxel centre lies in the target's
 * bounding box and a ray from it, in the direction lane idx of sampler.seed(0, n_voxels) draws
 * (square_to_uniform_sphere(next_2d), utils.py:114-117), first hits a back face. */
int orc_discretize(const OrcScene *s, real *voxels, int n_threads) {
    if (s->n_tri <= 0) return -1;
    real mbmin[3] = {RINF, RINF, RINF}, mbmax[3] = {-RINF, -RINF, -RINF};
    for (int i = 0; i < s->n_tri * 3; ++i)
        for (int a = 0; a < 3; ++a) {
            real v = (real)s->tri[3 * (size_t)i + a];
            if (v < mbmin[a]) mbmin[a] = v;
            if (v > mbmax[a]) mbmax[a] = v;
        }
    real bmin[3], vs[3];
    for (int a = 0; a < 3; ++a) {
        bmin[a] = (real)s->bbox_min[a];
        vs[a] = ((real)s->bbox_max[a] - (real)s->bbox_min[a]) / (real)s->res[a];
    }
    int64_t n_vox = (int64_t)s->res[0] * s->res[1] * s->res[2];
#ifdef _OPENMP
    if (n_threads > 0) omp_set_num_threads(n_threads);
#pragma omp parallel for schedule(dynamic, 1024)
#endif
    for (int64_t idx = 0; idx < n_vox; ++idx) {
        int x = (int)(idx % s->res[0]), y = (int)((idx / s->res[0]) % s->res[1]), z = (int)(idx / ((int64_t)s->res[0] * s->res[1]));
        Pcg32 rng; sampler_seed(&rng, 0u, (uint32_t)idx);
        real u0 = next_1d(&rng), u1 = next_1d(&rng);
        Ray r;
        r.o = v3(bmin[0] + vs[0] * ((real)x + (real)0.5), bmin[1] + vs[1] * ((real)y + (real)0.5),
                 bmin[2] + vs[2] * ((real)z + (real)0.5));
        real cz = 1 - 2 * u1, rr = safe_sqrt(1 - cz * cz);
        const real pi = (real)3.14159265358979323846;
        real phi = 2 * pi * u0;
        real sp = sizeof(real) == 4 ? (real)sinf((float)phi) : (real)sin((double)phi);
        real cp = sizeof(real) == 4 ? (real)cosf((float)phi) : (real)cos((double)phi);
        r.d = v3(rr * cp, rr * sp, cz);
        int in_bbox = r.o.x > mbmin[0] && r.o.x < mbmax[0] && r.o.y > mbmin[1] && r.o.y < mbmax[1] &&
                      r.o.z > mbmin[2] && r.o.z < mbmax[2];
        int inside = 0;
        if (in_bbox) {
            V3 n; real t = mesh_intersect(s, r, &n);
            inside = t < RINF && vdot(r.d, n) > 0;
        }
        voxels[idx] = inside ? (real)1 : (real)0;
    }
    return 0;
}

/* first `n` floats of lane `lane`'s stream (RNG unit test). */
int orc_rng_stream(uint32_t seed, uint32_t lane, int n, float *out) {
    Pcg32 rng; sampler_seed(&rng, seed, lane);
    for (int i = 0; i < n; ++i) out[i] = (float)next_1d(&rng);
    return 0;
}
