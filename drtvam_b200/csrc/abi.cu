// C-ABI of the hot path (include/drtvam_b200.h): handle management, argument
// validation, kernel-family selection, host-buffer convenience calls.
#include <dlfcn.h>

#include <atomic>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>

#include "../../include/drtvam_b200.h"
#include "kernels.h"

namespace drtvam {

static std::atomic<uint64_t> g_launches{0};
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

static thread_local char g_error[512] = "";

static int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_error, sizeof(g_error), fmt, ap);
    va_end(ap);
    return code;
}

#define CUDA_TRY(expr)                                                                       \
    do {                                                                                     \
        cudaError_t _e = (expr);                                                             \
        if (_e != cudaSuccess)                                                               \
            return fail(DRTVAM_E_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), \
                        __FILE__, __LINE__);                                                 \
    } while (0)

struct Scratch {
    void* ptr = nullptr;
    size_t bytes = 0;
    cudaError_t reserve(size_t n) {
        if (n <= bytes) return cudaSuccess;
        if (ptr) cudaFree(ptr);
        ptr = nullptr; bytes = 0;
        cudaError_t e = cudaMalloc(&ptr, n);
        if (e == cudaSuccess) bytes = n;
        return e;
    }
    void release() { if (ptr) cudaFree(ptr); ptr = nullptr; bytes = 0; }
};

}  // namespace drtvam

using namespace drtvam;

struct NcclId { char internal[128]; };   // ncclUniqueId (nccl.h: NCCL_UNIQUE_ID_BYTES = 128), passed by value

struct DrtvamScene {
    int device = 0;
    SceneParams P{};
    SceneParams P1{};                // one-channel view of P (no target, plain film) for the planar kernels
    InsideHost inside;               // inside-target segment records of a surface-aware planar scene
    float print_time = 1.f;
    float* motion_dev = nullptr;
    double* motion_f64_dev = nullptr;
    int force_mode = 0;              // 0 auto, 1 general tracer, 2 planar v1 (per-ray tile DDA)
    std::vector<float> motion_host;  // copy of the motion table for the plan check
    PlanarHost planar;
    PlanarZHost zplan;               // layer-parallel kernels on top of the planar tables
    PlanarGHost gplan;               // patch-gather forward on top of the layer-parallel plan
    PlanarJHost jplan;               // jittered planar passes (regular_sampling = false)
    // pinned pixel list (drtvam_scene_pin_pixels): verified once, then the per-call layout check and the
    // stand-by launch of the other kernel family are skipped
    const uint32_t* pinned_pixels = nullptr;
    uint64_t pinned_n = 0;
    int pinned_gate = -1;            // -1 = to be verified at the next call, 0 = planar layout, 1 = general
    // collectives for bindings without torch.distributed (drtvam_comm_init)
    void* nccl_comm = nullptr;
    int comm_rank = 0, comm_world = 1;
    bool planar_tried = false;       // planar_build ran for the current depth setting
    bool planar_ok = false;
    bool hybrid = false;             // scattering scene: planar per-lane kernels for the first medium segments only
    MeshHost target;                 // target mesh (surface-aware runs only)
    MeshHost mesh_surf[kMaxMeshSurfaces];  // triangle-mesh interfaces (custom vial walls, occluders)
    int n_mesh_surf = 0;
    const float* inv_volume = nullptr;  // caller-owned [film_elems] 1/sub-voxel volume (volume.py:41-42), or NULL
    Scratch dl_scaled;               // grad_in * inv_vol of the adjoint
    Scratch host_io[3];      // device staging for the *_host calls
    unsigned long long* counter_dev = nullptr;
};

namespace {

struct DeviceGuard {
    int prev = 0;
    bool ok;
    explicit DeviceGuard(int dev) { ok = cudaGetDevice(&prev) == cudaSuccess && cudaSetDevice(dev) == cudaSuccess; }
    ~DeviceGuard() { cudaSetDevice(prev); }
};

float ray_weight(const DrtvamScene* s, uint64_t n_active, uint32_t spp) {
    // pixel_area * n_active / (n_active * spp) * print_time
    // (projector.py:187,164-165; integrators/common.py:111)
    const float area = s->P.pixel_x * s->P.pixel_y;
    const float inv_pdf = area * static_cast<float>(n_active);
    return inv_pdf / static_cast<float>(n_active * static_cast<uint64_t>(spp)) * s->print_time;
}

int check_call(const DrtvamScene* s, uint64_t n_active, uint32_t& spp) {
    if (!s) return fail(DRTVAM_E_INVALID, "scene handle is NULL");
    if (spp == 0) return fail(DRTVAM_E_INVALID, "spp must be >= 1");
    if (s->P.regular_sampling) spp = 1;  // integrators/common.py:49-51
    int medium_surfaces = 0;
    for (int i = 0; i < s->P.n_surf; ++i)
        medium_surfaces += (s->P.surf[i].interior_medium != 0) + (s->P.surf[i].exterior_medium != 0);
    if (medium_surfaces == 0)
        return fail(DRTVAM_E_INVALID, "No printing medium found in the scene");  // common.py:36-37
    const unsigned __int128 lanes = static_cast<unsigned __int128>(n_active) * spp;
    if (lanes > (static_cast<unsigned __int128>(1) << 32))
        return fail(DRTVAM_E_TOO_MANY_RAYS,
                    "The total number of Monte Carlo samples required by this rendering task (%llu) exceeds "
                    "2^32 = 4294967296. Please use fewer samples per pixel or render using multiple passes.",
                    static_cast<unsigned long long>(lanes));
    return DRTVAM_OK;
}

// Host-side eligibility of the planar z-separable plan (see planar.cuh): every
// ray must have d_z = 0, one deterministic medium segment that depends only on
// (angle, column), and the film must be the plain single-channel DDA film.
// A SCATTERING medium (albedo != 0) is eligible as a hybrid: only the first medium segment of a lane - the
// unscattered chord, which every lane deposits in full (volume.py:216) - has the planar structure; the planar
// per-lane kernels take it and the general tracer everything after the first collision.
bool planar_static_eligible(const DrtvamScene* s) {
    const SceneParams& P = s->P;
    if (P.proj_kind != DRTVAM_PROJ_COLLIMATED || P.sample_time) return false;  // jitter keeps d_z = 0
    if (P.sensor_kind != DRTVAM_SENSOR_DDA) return false;
    // surface-aware film: the planar kernels make the channel sum, the inside records split it (inside_kernels.cu)
    if (P.surface_aware && !P.regular_sampling) return false;
    if (!P.transmission_only) return false;
    if (P.albedo != 0.f) {
        if (P.surface_aware || !general_supports_hybrid(P) || getenv("DRTVAM_SCATTER_GENERAL")) return false;
        // Russian roulette must not fire before the first medium segment (its draws differ from the table kernel's)
        if (P.max_depth < 1 || P.rr_depth < P.n_surf) return false;
    } else if (P.max_depth < 1 || P.max_depth - 1 > P.rr_depth) {
        return false;  // Russian roulette never fires
    }
    if (!P.motion_analytic) {
        // rotation about z: world z = const * local y, no z component in the beam direction
        const size_t n = static_cast<size_t>(P.n_patterns);
        if (s->motion_host.size() < 12 * n) return false;
        const float* M = s->motion_host.data();
        for (size_t k = 0; k < n; ++k) {
            const float* R = M + 12 * k + 8;
            if (R[0] != 0.f || R[2] != 0.f) return false;
            if (fabsf(R[1] - M[9]) > 1e-6f || fabsf(R[3] - M[11]) > 1e-6f) return false;
        }
    }
    float zmax = fmaxf(fabsf(P.bmin[2]), fabsf(P.bmax[2]));
    for (int i = 0; i < P.n_surf; ++i) {
        const Surface& S = P.surf[i];
        if (S.kind == DRTVAM_SURF_MESH) return false;  // arbitrary interface: general tracer
        if (S.kind == DRTVAM_SURF_CYLINDER) {
            if (!(S.z0 < P.bmin[2] && S.z1 > P.bmax[2])) return false;
            // spawn offset (1 + max|p|) eps must not depend on z
            if (!(S.radius * 0.70710678f >= zmax)) return false;
        } else {
            if (!(S.hz > zmax && S.hx >= zmax && S.hy >= zmax)) return false;
        }
    }
    return true;
}

// The sensor as the film buffers of a call see it: the z window [z_begin, z_end) of a z-slab shard as a
// box of its own.  Used by everything that walks with the general 3-D DDA into / out of those buffers.
SceneParams windowed(const SceneParams& P) {
    SceneParams W = P;
    if (P.z_begin == 0 && P.z_end == P.res[2]) return W;
    const float vs = (P.bmax[2] - P.bmin[2]) / static_cast<float>(P.res[2]);
    W.bmin[2] = P.bmin[2] + static_cast<float>(P.z_begin) * vs;
    W.bmax[2] = P.bmin[2] + static_cast<float>(P.z_end) * vs;
    W.res[2] = P.z_end - P.z_begin;
    W.z_begin = 0;
    W.z_end = W.res[2];
    return W;
}

bool has_window(const SceneParams& P) { return P.z_begin != 0 || P.z_end != P.res[2] || P.n_rows != P.res_y; }

int ensure_planar(DrtvamScene* s) {
    if (s->planar_tried) return DRTVAM_OK;
    s->planar_tried = true;
    s->planar_ok = false;
    if (s->pinned_pixels) s->pinned_gate = -1;   // the plan changed: verify the pinned list again
    inside_release(s->inside);
    s->P1 = s->P;
    s->P1.surface_aware = 0;
    s->P1.target = MeshView{};
    if (!planar_static_eligible(s)) return DRTVAM_OK;
    if (s->P.surface_aware && getenv("DRTVAM_SURFACE_AWARE_GENERAL")) return DRTVAM_OK;  // A/B tests
    cudaError_t e = planar_build(s->P1, s->planar);
    if (e != cudaSuccess) {
        planar_release(s->planar);
        return fail(DRTVAM_E_CUDA, "planar plan build failed: %s", cudaGetErrorString(e));
    }
    s->planar_ok = s->planar.valid && s->planar.flags == 0 && s->planar.n_rows_in_sensor > 0;
    planarz_release(s->zplan);
    planarg_release(s->gplan);
    planarj_release(s->jplan);
    s->hybrid = s->planar_ok && s->P.albedo != 0.f;
    if (s->planar_ok && (!s->P.regular_sampling || s->hybrid)) {
        // jittered passes: the layer tables of the layer-parallel plan (no walks) + the per-lane kernels
        e = planarz_build(s->P1, s->planar, s->planar.row_layer, s->zplan, false);
        if (e == cudaSuccess) e = planarj_build(s->P1, s->planar, s->zplan, s->jplan);
        if (e != cudaSuccess) {
            planarz_release(s->zplan);
            planarj_release(s->jplan);
            return fail(DRTVAM_E_CUDA, "jittered planar plan build failed: %s", cudaGetErrorString(e));
        }
        if (!s->jplan.valid) s->planar_ok = s->hybrid = false;
    }
    if (s->planar_ok && s->P.regular_sampling && !s->hybrid) {
        e = planarz_build(s->P1, s->planar, s->planar.row_layer, s->zplan);
        if (e != cudaSuccess) {
            planarz_release(s->zplan);
            return fail(DRTVAM_E_CUDA, "layer-parallel plan build failed: %s", cudaGetErrorString(e));
        }
        e = planarg_build(s->P1, s->planar, s->zplan, s->gplan);
        if (e != cudaSuccess) {  // e.g. out of memory: the scatter kernel serves the forward pass
            cudaGetLastError();
            planarg_release(s->gplan);
        }
    }
    if (s->planar_ok && s->P.surface_aware) {
        // tile lanes need kernels that can mask them: the patch-gather forward and the layer-parallel adjoint
        const bool allow_tiles = s->zplan.valid && s->gplan.valid && !has_window(s->P);
        e = inside_build(s->P, s->P1, s->planar, s->zplan, allow_tiles, s->inside);
        if (e != cudaSuccess) {  // the general tracer serves the scene
            cudaGetLastError();
            s->planar_ok = false;
        }
    }
    if (getenv("DRTVAM_DEBUG"))
        fprintf(stderr, "[drtvam] planar_ok %d zplan %d gplan %d (pairs %lld of %lld) frozen %d inside: %llu records of %u odd lanes, "
                        "tiles %d with %llu table rows\n",
                int(s->planar_ok), int(s->zplan.valid), int(s->gplan.valid), s->gplan.pairs, s->gplan.candidates, s->planar.n_frozen,
                static_cast<unsigned long long>(s->inside.n_rec), s->inside.n_odd, int(s->inside.tiles),
                static_cast<unsigned long long>(s->inside.n_rows));
    return DRTVAM_OK;
}

bool use_planar(DrtvamScene* s, uint64_t n_active) {
    if (s->force_mode == 1 || !s->planar_ok || n_active == 0) return false;
    // the per-ray tile DDA (A/B tests) knows no window, no hybrid and no masked lanes
    if (s->force_mode == 2 && (has_window(s->P) || s->hybrid || s->inside.tiles)) return false;
    const uint64_t rw = static_cast<uint64_t>(s->P.res_x) * s->P.n_rows;
    return n_active % rw == 0;
}

// elements of the film buffers of a call: the layers of the z window (all of them by default)
size_t film_elems(const DrtvamScene* s) {
    return static_cast<size_t>(s->P.res[0]) * s->P.res[1] * (s->P.z_end - s->P.z_begin) * (s->P.surface_aware ? 2 : 1);
}

// Layout gate of a planar call.  Returns the device gate word the kernels of the call read and sets
// `host_gate`: 0 = the list is known to have the planar layout (only planar kernels are launched),
// 1 = known not to (only the general tracer), -1 = decided on the device (both families are launched,
// the one that does not own the call exits in its first instruction).
cudaError_t planar_gate(DrtvamScene* s, const uint32_t* active_pixels, uint64_t n_active, cudaStream_t st,
                        const int** gate, int* host_gate) {
    *gate = s->planar.gate;
    *host_gate = -1;
    // the pinned list itself, or whole patterns out of it (blocks of angles of a streamed call): a
    // sub-range of a dense layout is dense
    const uint64_t rw = static_cast<uint64_t>(s->P.res_x) * s->P.n_rows;
    const bool inside = s->pinned_pixels && active_pixels >= s->pinned_pixels &&
                        active_pixels + n_active <= s->pinned_pixels + s->pinned_n &&
                        static_cast<uint64_t>(active_pixels - s->pinned_pixels) % rw == 0 && n_active % rw == 0;
    if (inside) {
        if (s->pinned_gate < 0) {   // first call after the pin or after a plan rebuild: verify once, synchronously
            cudaError_t e = launch_planar_check(s->P1, s->planar, s->pinned_pixels, s->pinned_n, st);
            if (e != cudaSuccess) return e;
            int g = 0;
            e = cudaMemcpyAsync(&g, s->planar.gate, sizeof(int), cudaMemcpyDeviceToHost, st);
            if (e == cudaSuccess) e = cudaStreamSynchronize(st);
            if (e != cudaSuccess) return e;
            s->pinned_gate = g != 0 ? 1 : 0;
        }
        *host_gate = s->pinned_gate;
        *gate = s->planar.gate + 3;   // a word that is always 0: the planar kernels own the call
        return cudaSuccess;
    }
    return launch_planar_check(s->P1, s->planar, active_pixels, n_active, st);
}

}  // namespace

extern "C" {

int drtvam_abi_version(void) { return DRTVAM_ABI_VERSION; }
const char* drtvam_last_error(void) { return g_error; }
uint64_t drtvam_launch_count(void) { return g_launches.load(); }

int drtvam_scene_create(const DrtvamSceneDesc* d, int device, DrtvamScene** out) {
    if (!d || !out) return fail(DRTVAM_E_INVALID, "NULL argument");
    *out = nullptr;
    if (d->struct_size != sizeof(DrtvamSceneDesc) || d->abi_version != DRTVAM_ABI_VERSION)
        return fail(DRTVAM_E_INVALID, "ABI mismatch: desc size %u (want %zu), version %u (want %d)", d->struct_size,
                    sizeof(DrtvamSceneDesc), d->abi_version, DRTVAM_ABI_VERSION);
    if (d->proj_kind < 0 || d->proj_kind > 2) return fail(DRTVAM_E_INVALID, "unknown projector kind %d", d->proj_kind);
    if (d->n_patterns <= 0 || d->res_x <= 0 || d->res_y <= 0)
        return fail(DRTVAM_E_INVALID, "projector size must be positive");
    // n_surf = 0: every interface is a mesh added afterwards (custom vial)
    if (d->n_surf < 0 || d->n_surf > DRTVAM_MAX_SURFACES)
        return fail(DRTVAM_E_INVALID, "surface count %d outside [0, %d]", d->n_surf, DRTVAM_MAX_SURFACES);
    if (!(d->sigma_t > 0.f)) return fail(DRTVAM_E_INVALID, "medium extinction must be positive");
    if (d->albedo < 0.f || d->albedo > 1.f) return fail(DRTVAM_E_INVALID, "albedo outside [0, 1]");
    if (d->sensor_kind != DRTVAM_SENSOR_DDA && d->sensor_kind != DRTVAM_SENSOR_RATIO &&
        d->sensor_kind != DRTVAM_SENSOR_DELTA)
        return fail(DRTVAM_E_INVALID, "unsupported sensor kind %d", d->sensor_kind);
    if (d->sensor_kind == DRTVAM_SENSOR_DELTA && !(d->albedo > 0.f))  // volume.py:156-157
        return fail(DRTVAM_E_INVALID, "Tried to render a purely absorptive volume with a delta tracking sensor. "
                                      "This is not supported.");
    if (d->sensor_kind == DRTVAM_SENSOR_RATIO && !(d->majorant > 0.f))
        return fail(DRTVAM_E_INVALID, "ratio sensor needs a positive majorant");
    if (!d->motion_analytic && !d->motion_table_host)
        return fail(DRTVAM_E_INVALID, "motion table missing");
    if (d->sample_time && !d->motion_analytic)
        return fail(DRTVAM_E_INVALID, "sample_time needs an analytic motion");
    int medium_surfaces = 0;
    for (int i = 0; i < d->n_surf; ++i) {
        const DrtvamSurface& s = d->surf[i];
        if (s.kind != DRTVAM_SURF_CYLINDER && s.kind != DRTVAM_SURF_BOX)
            return fail(DRTVAM_E_INVALID, "surface %d: unknown kind %d", i, s.kind);
        if (s.bsdf < 0 || s.bsdf > 2) return fail(DRTVAM_E_INVALID, "surface %d: unknown bsdf %d", i, s.bsdf);
        if (s.bsdf == DRTVAM_BSDF_DIELECTRIC && !(s.eta > 0.f))
            return fail(DRTVAM_E_INVALID, "surface %d: dielectric needs eta > 0", i);
        medium_surfaces += (s.interior_medium != 0) + (s.exterior_medium != 0);
    }
    if (medium_surfaces == 0 && d->n_surf > 0)
        return fail(DRTVAM_E_INVALID, "No printing medium found in the scene");  // common.py:36-37
    for (int a = 0; a < 3; ++a) {
        if (d->res[a] <= 0) return fail(DRTVAM_E_INVALID, "film resolution must be positive");
        if (!(d->bbox_max[a] > d->bbox_min[a])) return fail(DRTVAM_E_INVALID, "empty sensor bounding box");
    }

    if (static_cast<unsigned long long>(d->res[0]) * d->res[1] * d->res[2] * (d->surface_aware ? 2 : 1) >= (1ull << 31))
        return fail(DRTVAM_E_INVALID, "film of 2^31 elements or more is not supported (32-bit voxel index)");

    int n_dev = 0;
    CUDA_TRY(cudaGetDeviceCount(&n_dev));
    if (device < 0 || device >= n_dev) return fail(DRTVAM_E_INVALID, "device %d not available (%d devices)", device, n_dev);
    DeviceGuard guard(device);
    if (!guard.ok) return fail(DRTVAM_E_CUDA, "cudaSetDevice(%d) failed", device);

    DrtvamScene* s = new (std::nothrow) DrtvamScene();
    if (!s) return fail(DRTVAM_E_INVALID, "out of host memory");
    s->device = device;
    SceneParams& P = s->P;
    P.proj_kind = d->proj_kind; P.n_patterns = d->n_patterns; P.res_x = d->res_x; P.res_y = d->res_y;
    P.pixel_x = d->pixel_x; P.pixel_y = d->pixel_y;
    P.aperture_radius = d->aperture_radius; P.focus_distance = d->focus_distance;
    P.motion_analytic = d->motion_analytic; P.clockwise = d->clockwise; P.distance = d->distance;
    P.max_depth = d->max_depth; P.rr_depth = d->rr_depth; P.transmission_only = d->transmission_only;
    P.regular_sampling = d->regular_sampling; P.sample_time = d->sample_time;
    s->print_time = d->print_time;
    P.sigma_t = d->sigma_t; P.albedo = d->albedo; P.phase_kind = d->phase_kind; P.hg_g = d->hg_g;
    P.n_surf = d->n_surf;
    for (int i = 0; i < d->n_surf; ++i) {
        const DrtvamSurface& a = d->surf[i];
        Surface& b = P.surf[i];
        b.kind = a.kind; b.bsdf = a.bsdf; b.interior_medium = a.interior_medium; b.exterior_medium = a.exterior_medium;
        b.radius = a.radius; b.z0 = a.z0; b.z1 = a.z1; b.hx = a.hx; b.hy = a.hy; b.hz = a.hz; b.eta = a.eta;
        b.mesh = -1;
    }
    for (int a = 0; a < 3; ++a) { P.bmin[a] = d->bbox_min[a]; P.bmax[a] = d->bbox_max[a]; P.res[a] = d->res[a]; }
    P.surface_aware = d->surface_aware; P.sensor_kind = d->sensor_kind; P.majorant = d->majorant;
    P.z_begin = 0; P.z_end = d->res[2]; P.row_begin = 0; P.n_rows = d->res_y;
    P.sampler_index = nullptr;
    P.motion_table = nullptr;
    P.motion_table_f64 = nullptr;
    if (d->motion_table_host) {
        const size_t bytes = sizeof(float) * 12 * static_cast<size_t>(d->n_patterns);
        cudaError_t e = cudaMalloc(&s->motion_dev, bytes);
        if (e == cudaSuccess) e = cudaMemcpy(s->motion_dev, d->motion_table_host, bytes, cudaMemcpyHostToDevice);
        if (e != cudaSuccess) {
            drtvam_scene_destroy(s);
            return fail(DRTVAM_E_CUDA, "motion table upload failed: %s", cudaGetErrorString(e));
        }
        P.motion_table = s->motion_dev;
        s->motion_host.assign(d->motion_table_host, d->motion_table_host + 12 * static_cast<size_t>(d->n_patterns));
        // double-precision copy for the set-up of the planar tables (given, or the fp32 one promoted)
        std::vector<double> m64(12 * static_cast<size_t>(d->n_patterns));
        for (size_t i = 0; i < m64.size(); ++i)
            m64[i] = d->motion_table_host_f64 ? d->motion_table_host_f64[i] : static_cast<double>(d->motion_table_host[i]);
        e = cudaMalloc(&s->motion_f64_dev, sizeof(double) * m64.size());
        if (e == cudaSuccess) e = cudaMemcpy(s->motion_f64_dev, m64.data(), sizeof(double) * m64.size(), cudaMemcpyHostToDevice);
        if (e != cudaSuccess) {
            drtvam_scene_destroy(s);
            return fail(DRTVAM_E_CUDA, "motion table upload failed: %s", cudaGetErrorString(e));
        }
        P.motion_table_f64 = s->motion_f64_dev;
    }
    cudaError_t e = cudaMalloc(&s->counter_dev, sizeof(unsigned long long));
    if (e != cudaSuccess) {
        drtvam_scene_destroy(s);
        return fail(DRTVAM_E_CUDA, "counter allocation failed: %s", cudaGetErrorString(e));
    }
    *out = s;
    return DRTVAM_OK;
}

void drtvam_scene_destroy(DrtvamScene* s) {
    if (!s) return;
    drtvam_comm_destroy(s);
    DeviceGuard guard(s->device);
    if (s->motion_dev) cudaFree(s->motion_dev);
    if (s->motion_f64_dev) cudaFree(s->motion_f64_dev);
    if (s->counter_dev) cudaFree(s->counter_dev);
    planar_release(s->planar);
    planarz_release(s->zplan);
    planarg_release(s->gplan);
    planarj_release(s->jplan);
    inside_release(s->inside);
    mesh_release(s->target);
    for (auto& m : s->mesh_surf) mesh_release(m);
    s->dl_scaled.release();
    for (auto& b : s->host_io) b.release();
    delete s;
}

int drtvam_scene_set_depth(DrtvamScene* s, int32_t max_depth, int32_t rr_depth) {
    if (!s) return fail(DRTVAM_E_INVALID, "scene handle is NULL");
    if (s->P.max_depth != max_depth || s->P.rr_depth != rr_depth) s->planar_tried = false;
    s->P.max_depth = max_depth;
    s->P.rr_depth = rr_depth;
    return DRTVAM_OK;
}

int drtvam_scene_force_general(DrtvamScene* s, int32_t on) {
    if (!s) return fail(DRTVAM_E_INVALID, "scene handle is NULL");
    if (on < 0 || on > 2) return fail(DRTVAM_E_INVALID, "force mode must be 0 (auto), 1 (general) or 2 (planar v1)");
    s->force_mode = on;
    return DRTVAM_OK;
}

int drtvam_scene_stat(DrtvamScene* s, int32_t which, uint64_t* value) {
    if (!s) return fail(DRTVAM_E_INVALID, "scene handle is NULL");
    if (!value) return fail(DRTVAM_E_INVALID, "NULL buffer");
    DeviceGuard guard(s->device);
    if (int rc = ensure_planar(s)) return rc;
    const bool tables = s->zplan.valid && s->gplan.valid && s->P.regular_sampling;
    switch (which) {
        case 0: *value = tables ? static_cast<uint64_t>(s->gplan.pairs) : 0; break;
        case 1: *value = s->zplan.valid ? static_cast<uint64_t>(s->zplan.nz) : 0; break;
        case 2: *value = s->inside.valid ? s->inside.n_rec + s->inside.n_rows * 32 : 0; break;
        case 3: *value = tables ? static_cast<uint64_t>(s->gplan.pairs) * 68u : 0; break;
        case 4: {
            unsigned n = 0;
            if (s->jplan.valid && s->jplan.wmax)
                CUDA_TRY(cudaMemcpy(&n, s->jplan.wmax + 1, sizeof(n), cudaMemcpyDeviceToHost));
            unsigned m = 0;
            if (s->inside.valid && s->inside.stat)
                CUDA_TRY(cudaMemcpy(&m, s->inside.stat + 2, sizeof(m), cudaMemcpyDeviceToHost));
            *value = static_cast<uint64_t>(n) + m;
            break;
        }
        default: return fail(DRTVAM_E_INVALID, "unknown statistic %d", which);
    }
    return DRTVAM_OK;
}

int drtvam_scene_set_target(DrtvamScene* s, const float* triangles_host, uint32_t n_triangles) {
    if (!s) return fail(DRTVAM_E_INVALID, "scene handle is NULL");
    if (n_triangles && !triangles_host) return fail(DRTVAM_E_INVALID, "NULL triangle buffer");
    if (n_triangles > (1u << 28)) return fail(DRTVAM_E_INVALID, "too many triangles");
    DeviceGuard guard(s->device);
    CUDA_TRY(cudaDeviceSynchronize());
    CUDA_TRY(mesh_build(triangles_host, n_triangles, s->target));
    s->P.target = s->target.view;
    s->planar_tried = false;
    return DRTVAM_OK;
}

int drtvam_scene_add_mesh_surface(DrtvamScene* s, const float* triangles_host, uint32_t n_triangles, int32_t bsdf,
                                  float eta, int32_t interior_medium, int32_t exterior_medium) {
    if (!s) return fail(DRTVAM_E_INVALID, "scene handle is NULL");
    if (!triangles_host || n_triangles == 0) return fail(DRTVAM_E_INVALID, "empty triangle buffer");
    if (n_triangles > (1u << 28)) return fail(DRTVAM_E_INVALID, "too many triangles");
    if (bsdf < 0 || bsdf > 2) return fail(DRTVAM_E_INVALID, "unknown bsdf %d", bsdf);
    if (bsdf == DRTVAM_BSDF_DIELECTRIC && !(eta > 0.f)) return fail(DRTVAM_E_INVALID, "dielectric needs eta > 0");
    if (s->n_mesh_surf >= kMaxMeshSurfaces || s->P.n_surf >= kMaxSurfaces)
        return fail(DRTVAM_E_INVALID, "too many surfaces (%d of %d, %d of %d meshes)", s->P.n_surf, kMaxSurfaces,
                    s->n_mesh_surf, kMaxMeshSurfaces);
    DeviceGuard guard(s->device);
    CUDA_TRY(cudaDeviceSynchronize());
    const int m = s->n_mesh_surf;
    CUDA_TRY(mesh_build(triangles_host, n_triangles, s->mesh_surf[m]));
    s->P.mesh_surf[m] = s->mesh_surf[m].view;
    Surface& b = s->P.surf[s->P.n_surf];
    b = Surface{};
    b.kind = DRTVAM_SURF_MESH; b.bsdf = bsdf; b.interior_medium = interior_medium; b.exterior_medium = exterior_medium;
    b.eta = bsdf == DRTVAM_BSDF_DIELECTRIC ? eta : 1.f;
    b.mesh = m;
    s->n_mesh_surf = m + 1;
    s->P.n_surf += 1;
    s->planar_tried = false;
    return DRTVAM_OK;
}

int drtvam_scene_set_inv_volume(DrtvamScene* s, const float* inv_volume) {
    if (!s) return fail(DRTVAM_E_INVALID, "scene handle is NULL");
    s->inv_volume = inv_volume;
    return DRTVAM_OK;
}

int drtvam_scene_set_z_window(DrtvamScene* s, int32_t z_begin, int32_t z_end, int32_t* row_begin, int32_t* row_end) {
    if (!s) return fail(DRTVAM_E_INVALID, "scene handle is NULL");
    const int rz = s->P.res[2];
    if (z_begin < 0 || z_end > rz || z_begin >= z_end)
        return fail(DRTVAM_E_INVALID, "z window [%d, %d) outside [0, %d)", z_begin, z_end, rz);
    DeviceGuard guard(s->device);
    // the row -> layer table of the whole film decides which rows the window owns
    s->P.z_begin = 0; s->P.z_end = rz; s->P.row_begin = 0; s->P.n_rows = s->P.res_y;
    s->planar_tried = false;
    if (z_begin == 0 && z_end == rz) {
        if (row_begin) *row_begin = 0;
        if (row_end) *row_end = s->P.res_y;
        return DRTVAM_OK;
    }
    if (s->P.surface_aware) return fail(DRTVAM_E_INVALID, "a z window is not supported with a surface-aware film");
    if (int rc = ensure_planar(s)) return rc;
    // (a scattering medium is planar only in its first segments: the scattered ones leave their layers)
    if (!s->planar_ok || s->hybrid)
        return fail(DRTVAM_E_INVALID, "a z window needs a planar scene (collimated projector, rotation about z, "
                                      "vertical vial walls, no scattering): shard by projection angle instead");
    const std::vector<int>& rl = s->planar.row_layer;
    const std::vector<float>& rzv = s->planar.row_z;
    int r0 = -1, r1 = -1;
    for (int r = 0; r < s->P.res_y; ++r) {
        if (rl[r] < z_begin || rl[r] >= z_end) continue;
        if (r0 < 0) r0 = r;
        else if (r != r1) return fail(DRTVAM_E_INVALID, "the DMD rows of z window [%d, %d) are not contiguous", z_begin, z_end);
        r1 = r + 1;
    }
    if (r0 < 0) return fail(DRTVAM_E_INVALID, "no DMD row deposits in z window [%d, %d)", z_begin, z_end);
    if (!s->P.regular_sampling) {
        // jittered rays of a row spread over the row's height: every row must lie inside ONE layer, otherwise
        // its pattern gradient would need contributions from two windows
        const float vs = (s->P.bmax[2] - s->P.bmin[2]) / static_cast<float>(rz);
        const float dz = s->P.res_y > 1 ? fabsf(rzv[1] - rzv[0]) : s->P.pixel_y;
        for (int r = r0; r < r1; ++r) {
            const float lo = (rzv[r] - 0.5f * dz - s->P.bmin[2]) / vs, hi = (rzv[r] + 0.5f * dz - s->P.bmin[2]) / vs;
            if (static_cast<int>(floorf(lo + 1e-3f)) != rl[r] || static_cast<int>(floorf(hi - 1e-3f)) != rl[r])
                return fail(DRTVAM_E_INVALID, "DMD row %d straddles two voxel layers: a z window cannot own it with "
                                              "jittered sampling; shard by projection angle instead", r);
        }
    }
    s->P.z_begin = z_begin; s->P.z_end = z_end; s->P.row_begin = r0; s->P.n_rows = r1 - r0;
    s->planar_tried = false;   // tables of the window
    if (row_begin) *row_begin = r0;
    if (row_end) *row_end = r1;
    return DRTVAM_OK;
}

int drtvam_scene_pin_pixels(DrtvamScene* s, const uint32_t* active_pixels, uint64_t n_active) {
    if (!s) return fail(DRTVAM_E_INVALID, "scene handle is NULL");
    s->pinned_pixels = active_pixels;
    s->pinned_n = active_pixels ? n_active : 0;
    s->pinned_gate = -1;
    s->jplan.cache_valid = false;   // a segment table kept for the previous list is void
    return DRTVAM_OK;
}

int drtvam_scene_set_sampler_index(DrtvamScene* s, const uint32_t* global_index) {
    if (!s) return fail(DRTVAM_E_INVALID, "scene handle is NULL");
    s->P.sampler_index = global_index;
    s->P1.sampler_index = global_index;
    return DRTVAM_OK;
}

// ---- collectives for bindings that have no torch.distributed: NCCL, resolved at run time ------------
namespace {
struct NcclApi {
    void* lib = nullptr;
    int (*GetUniqueId)(void*) = nullptr;
    int (*CommInitRank)(void**, int, NcclId, int) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
    int (*CommGetAsyncError)(void*, int*) = nullptr;
    int (*CommDestroy)(void*) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
};
NcclApi g_nccl;

int load_nccl() {
    if (g_nccl.lib) return DRTVAM_OK;
    // the copy the process already uses (torch's bundled NCCL, if any) wins over the system one
    void* lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
    if (!lib) lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) return fail(DRTVAM_E_INVALID, "NCCL not found: %s", dlerror());
    NcclApi a;
    a.lib = lib;
    a.GetUniqueId = reinterpret_cast<decltype(a.GetUniqueId)>(dlsym(lib, "ncclGetUniqueId"));
    a.CommInitRank = reinterpret_cast<decltype(a.CommInitRank)>(dlsym(lib, "ncclCommInitRank"));
    a.AllReduce = reinterpret_cast<decltype(a.AllReduce)>(dlsym(lib, "ncclAllReduce"));
    a.CommGetAsyncError = reinterpret_cast<decltype(a.CommGetAsyncError)>(dlsym(lib, "ncclCommGetAsyncError"));
    a.CommDestroy = reinterpret_cast<decltype(a.CommDestroy)>(dlsym(lib, "ncclCommDestroy"));
    a.GetErrorString = reinterpret_cast<decltype(a.GetErrorString)>(dlsym(lib, "ncclGetErrorString"));
    if (!a.GetUniqueId || !a.CommInitRank || !a.AllReduce || !a.CommGetAsyncError || !a.CommDestroy || !a.GetErrorString)
        return fail(DRTVAM_E_INVALID, "NCCL library lacks a required symbol");
    g_nccl = a;
    return DRTVAM_OK;
}

int nccl_fail(const char* what, int rc) {
    return fail(DRTVAM_E_NCCL, "%s failed: %s", what, g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "?");
}

// ncclCommGetAsyncError: a peer that died or a network error surfaces here, not in the enqueue call
int comm_check(DrtvamScene* s) {
    int async = 0;
    const int rc = g_nccl.CommGetAsyncError(s->nccl_comm, &async);
    if (rc != 0) return nccl_fail("ncclCommGetAsyncError", rc);
    if (async != 0 && async != 7 /* ncclInProgress */) return nccl_fail("NCCL communicator (asynchronous error)", async);
    return DRTVAM_OK;
}

int comm_allreduce(DrtvamScene* s, void* buf, uint64_t n, int dtype, void* stream) {
    if (!s) return fail(DRTVAM_E_INVALID, "scene handle is NULL");
    if (!s->nccl_comm) return fail(DRTVAM_E_INVALID, "drtvam_comm_init has not been called on this handle");
    if (!buf && n) return fail(DRTVAM_E_INVALID, "NULL buffer");
    DeviceGuard guard(s->device);
    if (int rc = comm_check(s)) return rc;
    const int rc = g_nccl.AllReduce(buf, buf, n, dtype, 0 /* ncclSum */, s->nccl_comm, static_cast<cudaStream_t>(stream));
    if (rc != 0) return nccl_fail("ncclAllReduce", rc);
    return comm_check(s);
}
}  // namespace

int drtvam_comm_unique_id(void* id_out) {
    if (!id_out) return fail(DRTVAM_E_INVALID, "NULL buffer");
    if (int rc = load_nccl()) return rc;
    const int rc = g_nccl.GetUniqueId(id_out);
    if (rc != 0) return nccl_fail("ncclGetUniqueId", rc);
    return DRTVAM_OK;
}

int drtvam_comm_init(DrtvamScene* s, const void* unique_id, int32_t rank, int32_t world) {
    if (!s) return fail(DRTVAM_E_INVALID, "scene handle is NULL");
    if (!unique_id || world < 1 || rank < 0 || rank >= world) return fail(DRTVAM_E_INVALID, "bad rank / world / id");
    if (s->nccl_comm) return fail(DRTVAM_E_INVALID, "the handle already has a communicator");
    if (int rc = load_nccl()) return rc;
    DeviceGuard guard(s->device);
    NcclId id;
    memcpy(&id, unique_id, sizeof(id));
    void* comm = nullptr;
    const int rc = g_nccl.CommInitRank(&comm, world, id, rank);
    if (rc != 0) return nccl_fail("ncclCommInitRank", rc);
    s->nccl_comm = comm;
    s->comm_rank = rank;
    s->comm_world = world;
    return comm_check(s);
}

int drtvam_comm_destroy(DrtvamScene* s) {
    if (!s) return fail(DRTVAM_E_INVALID, "scene handle is NULL");
    if (!s->nccl_comm) return DRTVAM_OK;
    DeviceGuard guard(s->device);
    const int rc = g_nccl.CommDestroy(s->nccl_comm);
    s->nccl_comm = nullptr;
    if (rc != 0) return nccl_fail("ncclCommDestroy", rc);
    return DRTVAM_OK;
}

int drtvam_allreduce_dose(DrtvamScene* s, float* dose, uint64_t n, void* stream) {
    return comm_allreduce(s, dose, n, 7 /* ncclFloat32 */, stream);
}

int drtvam_allreduce_scalars(DrtvamScene* s, double* values, uint64_t n, void* stream) {
    return comm_allreduce(s, values, n, 8 /* ncclFloat64 */, stream);
}

int drtvam_compute_volume(DrtvamScene* s, uint32_t sample_count, float* volumes, void* stream) {
    if (!s) return fail(DRTVAM_E_INVALID, "scene handle is NULL");
    if (!volumes || sample_count == 0) return fail(DRTVAM_E_INVALID, "NULL buffer or zero sample count");
    if (s->P.target.n_tri <= 0) return fail(DRTVAM_E_INVALID, "No target shape found in the scene");  // sensor.py:62-63
    DeviceGuard guard(s->device);
    CUDA_TRY(launch_compute_volume(s->P, sample_count, volumes, static_cast<cudaStream_t>(stream)));
    return DRTVAM_OK;
}

int drtvam_discretize(DrtvamScene* s, float* voxels, void* stream) {
    if (!s) return fail(DRTVAM_E_INVALID, "scene handle is NULL");
    if (!voxels) return fail(DRTVAM_E_INVALID, "NULL buffer");
    if (s->P.target.n_tri <= 0) return fail(DRTVAM_E_INVALID, "No target shape found in the scene");  // utils.py:95-96
    DeviceGuard guard(s->device);
    CUDA_TRY(launch_discretize(s->P, voxels, static_cast<cudaStream_t>(stream)));
    return DRTVAM_OK;
}

int drtvam_scene_plan(DrtvamScene* s, uint64_t n_active, uint32_t) {
    if (!s) return fail(DRTVAM_E_INVALID, "scene handle is NULL");
    DeviceGuard guard(s->device);
    if (int rc = ensure_planar(s)) return rc;
    if (!use_planar(s, n_active)) return 0;
    if (s->hybrid) return 4;
    if (!s->P.regular_sampling) return 3;
    return (s->zplan.valid && s->force_mode != 2) ? 2 : 1;
}

int drtvam_forward(DrtvamScene* s, const float* active_data, const uint32_t* active_pixels, uint64_t n_active,
                   uint32_t spp, uint32_t seed, float out_scale, float* dose, int32_t zero_first, void* stream) {
    if (int rc = check_call(s, n_active, spp)) return rc;
    if (!dose || (n_active && (!active_data || !active_pixels))) return fail(DRTVAM_E_INVALID, "NULL buffer");
    DeviceGuard guard(s->device);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (zero_first) CUDA_TRY(cudaMemsetAsync(dose, 0, sizeof(float) * film_elems(s), st));
    if (n_active == 0) return DRTVAM_OK;
    const float w = ray_weight(s, n_active, spp) * out_scale;
    if (int rc = ensure_planar(s)) return rc;
    if (s->inv_volume && !zero_first)
        return fail(DRTVAM_E_INVALID, "an accumulating forward call cannot apply the per-voxel inverse volume");
    const SceneParams Pw = windowed(s->P);
    if (s->hybrid && use_planar(s, n_active)) {
        // scattering: the unscattered chords through the planar per-lane kernels, the rest through the tracer
        const int* gate = nullptr;
        int host_gate = -1;
        CUDA_TRY(planar_gate(s, active_pixels, n_active, st, &gate, &host_gate));
        if (host_gate != 1)
            CUDA_TRY(launch_planarj_forward(s->P1, windowed(s->P1), s->zplan, s->jplan, gate, active_data, active_pixels,
                                            n_active, spp, seed, w, dose, host_gate == 0, st));
        CUDA_TRY(launch_forward_general(Pw, active_data, active_pixels, n_active * spp, spp, seed, w, dose, st,
                                        host_gate != 1 ? gate : nullptr, host_gate != 1));
        if (s->inv_volume) CUDA_TRY(launch_mul_inplace(dose, s->inv_volume, film_elems(s), st));  // volume.py:54
        return DRTVAM_OK;
    }
    if (use_planar(s, n_active)) {
        const int* gate = nullptr;
        int host_gate = -1;
        CUDA_TRY(planar_gate(s, active_pixels, n_active, st, &gate, &host_gate));
        if (host_gate != 1) {
            // gate word 0 -> the planar kernels own the call (decided on the device unless the list is pinned)
            const bool sa = s->inside.valid;  // surface-aware: the planar kernels fill a one-channel plane
            const SceneParams& P1 = s->P1;
            float* plane = sa ? s->inside.planes : dose;
            if (sa) CUDA_TRY(inside_forward_begin(s->inside, st));
            if (!P1.regular_sampling)
                // a pinned list stays as it is, so the segment table of this pass can serve the next forward
                // pass with the same seed (A d after A p, lbfgs.py:248)
                CUDA_TRY(launch_planarj_forward(P1, windowed(P1), s->zplan, s->jplan, gate, active_data, active_pixels,
                                                n_active, spp, seed, w, plane, host_gate == 0, st));
            else if (s->zplan.valid && s->gplan.valid && s->force_mode != 2)
                // fresh: the plane was cleared above (zero_first, or the planes of a surface-aware scene) and
                // the patch gather is its first writer, so it stores instead of read-add-write
                CUDA_TRY(launch_planarg_forward(P1, s->zplan, s->gplan, gate, active_data, active_pixels, n_active, w,
                                                plane, sa || zero_first != 0, st,
                                                sa && s->inside.tiles ? s->inside.mask : nullptr));
            else if (s->zplan.valid && s->force_mode != 2)
                CUDA_TRY(launch_planarz_forward(P1, s->zplan, gate, active_data, active_pixels, n_active, w, plane, st));
            else
                CUDA_TRY(launch_planar_forward(P1, s->planar, active_data, active_pixels, n_active, w, plane, st, gate));
            if (P1.regular_sampling)
                CUDA_TRY(launch_planar_frozen(windowed(P1), s->planar, 0, active_data, plane, active_pixels, n_active, w,
                                              st, gate));
            if (sa)
                CUDA_TRY(inside_forward_finish(P1, s->inside, s->zplan, gate, active_data, active_pixels, n_active, w, dose,
                                               st));
        }
        if (host_gate != 0)
            CUDA_TRY(launch_forward_general(Pw, active_data, active_pixels, n_active * spp, spp, seed, w, dose, st,
                                            host_gate < 0 ? s->planar.gate : nullptr));
        if (s->inv_volume) CUDA_TRY(launch_mul_inplace(dose, s->inv_volume, film_elems(s), st));  // volume.py:54
        return DRTVAM_OK;
    }
    CUDA_TRY(launch_forward_general(Pw, active_data, active_pixels, n_active * spp, spp, seed, w, dose, st));
    if (s->inv_volume) CUDA_TRY(launch_mul_inplace(dose, s->inv_volume, film_elems(s), st));  // volume.py:54
    return DRTVAM_OK;
}

// The deterministic planar forward with no allocation anywhere on the call path: the repacked patterns go to the
// static scratch of the library image (planarz_kernels.cu) instead of the handle's cudaMalloc'ed buffer.
int drtvam_forward_noalloc(DrtvamScene* s, const float* active_data, const uint32_t* active_pixels, uint64_t n_active,
                           uint32_t spp, uint32_t seed, float out_scale, float* dose, int32_t zero_first, void* stream) {
    if (int rc = check_call(s, n_active, spp)) return rc;
    if (!dose || (n_active && (!active_data || !active_pixels))) return fail(DRTVAM_E_INVALID, "NULL buffer");
    if (n_active == 0) {   // nothing to project: only the clearing of the film is left
        DeviceGuard guard0(s->device);
        if (zero_first) CUDA_TRY(cudaMemsetAsync(dose, 0, sizeof(float) * film_elems(s), static_cast<cudaStream_t>(stream)));
        return DRTVAM_OK;
    }
    if (!s->planar_tried)
        return fail(DRTVAM_E_INVALID, "drtvam_forward_noalloc: the tables of the scene are not built yet (building them "
                                      "allocates): call drtvam_scene_plan or drtvam_forward once first");
    if (!s->P.regular_sampling || !use_planar(s, n_active) || !s->zplan.valid || !s->gplan.valid || s->force_mode != 0 ||
        s->inside.valid || s->inv_volume)
        return fail(DRTVAM_E_INVALID, "drtvam_forward_noalloc serves the deterministic planar plan with its patch-gather "
                                      "tables only (drtvam_scene_plan == 2, one film channel, whole angles)");
    const uint64_t rw = static_cast<uint64_t>(s->P.res_x) * s->P.n_rows;
    const uint64_t need = (n_active / rw) * s->P.res_x * static_cast<uint64_t>(s->zplan.nz_pad);
    if (need > planarg_noalloc_capacity())
        return fail(DRTVAM_E_INVALID, "drtvam_forward_noalloc: %llu angles x %d columns x %d layers = %llu floats of scratch "
                                      "exceed the static maximum of %llu (16 MiB): split the call by angles",
                    static_cast<unsigned long long>(n_active / rw), s->P.res_x, s->zplan.nz_pad,
                    static_cast<unsigned long long>(need), static_cast<unsigned long long>(planarg_noalloc_capacity()));
    DeviceGuard guard(s->device);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (zero_first) CUDA_TRY(cudaMemsetAsync(dose, 0, sizeof(float) * film_elems(s), st));
    const float w = ray_weight(s, n_active, spp) * out_scale;
    const int* gate = nullptr;
    int host_gate = -1;
    CUDA_TRY(planar_gate(s, active_pixels, n_active, st, &gate, &host_gate));
    if (host_gate != 1) {
        CUDA_TRY(launch_planarg_forward_noalloc(s->P1, s->zplan, s->gplan, s->device, gate, active_data, active_pixels,
                                                n_active, w, dose, zero_first != 0, st));
        CUDA_TRY(launch_planar_frozen(windowed(s->P1), s->planar, 0, active_data, dose, active_pixels, n_active, w, st, gate));
    }
    if (host_gate != 0)   // a list without the planar layout: the general tracer owns the call (it allocates nothing)
        CUDA_TRY(launch_forward_general(windowed(s->P), active_data, active_pixels, n_active * spp, spp, seed, w, dose, st,
                                        host_gate < 0 ? s->planar.gate : nullptr));
    return DRTVAM_OK;
}

int drtvam_adjoint(DrtvamScene* s, const float* dose_grad, const uint32_t* active_pixels, uint64_t n_active,
                   uint32_t spp, uint32_t seed, float in_scale, float* pattern_grad, void* stream) {
    if (int rc = check_call(s, n_active, spp)) return rc;
    if (n_active == 0) return DRTVAM_OK;
    if (!dose_grad || !active_pixels || !pattern_grad) return fail(DRTVAM_E_INVALID, "NULL buffer");
    DeviceGuard guard(s->device);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const float w = ray_weight(s, n_active, spp) * in_scale;
    if (int rc = ensure_planar(s)) return rc;
    if (s->inv_volume) {  // volume.py:130: dL = grad_in * inv_vol
        CUDA_TRY(s->dl_scaled.reserve(sizeof(float) * film_elems(s)));
        CUDA_TRY(launch_mul(dose_grad, s->inv_volume, static_cast<float*>(s->dl_scaled.ptr), film_elems(s), st));
        dose_grad = static_cast<const float*>(s->dl_scaled.ptr);
    }
    const SceneParams Pw = windowed(s->P);
    if (s->hybrid && use_planar(s, n_active)) {
        const int* gate = nullptr;
        int host_gate = -1;
        CUDA_TRY(planar_gate(s, active_pixels, n_active, st, &gate, &host_gate));
        if (host_gate != 1) {
            // the planar kernels write the gradient of the first segments (or nothing when the gate says the list
            // is not planar), the tracer adds the rest
            CUDA_TRY(cudaMemsetAsync(pattern_grad, 0, sizeof(float) * n_active, st));
            CUDA_TRY(launch_planarj_adjoint(s->P1, windowed(s->P1), s->zplan, s->jplan, gate, dose_grad, active_pixels,
                                            n_active, spp, seed, w, pattern_grad, st));
        }
        CUDA_TRY(launch_adjoint_general(Pw, dose_grad, active_pixels, n_active * spp, spp, seed, w, pattern_grad, st,
                                        host_gate != 1 ? gate : nullptr, host_gate != 1));
        return DRTVAM_OK;
    }
    if (use_planar(s, n_active)) {
        const int* gate = nullptr;
        int host_gate = -1;
        CUDA_TRY(planar_gate(s, active_pixels, n_active, st, &gate, &host_gate));
        if (host_gate != 1) {
            const bool sa = s->inside.valid;
            const SceneParams& P1 = s->P1;
            const float* dl1 = dose_grad;  // one-channel gradient plane the planar kernels gather from
            if (sa) {
                CUDA_TRY(inside_adjoint_begin(s->inside, dose_grad, st));
                dl1 = s->inside.planes;
            }
            if (!P1.regular_sampling) {
                CUDA_TRY(launch_planarj_adjoint(P1, windowed(P1), s->zplan, s->jplan, gate, dl1, active_pixels, n_active,
                                                spp, seed, w, pattern_grad, st));
            } else if (s->zplan.valid && s->force_mode != 2) {
                CUDA_TRY(launch_planarz_adjoint(P1, s->zplan, gate, dl1, active_pixels, n_active, w, pattern_grad, st,
                                                sa ? &s->inside : nullptr));
            } else {
                CUDA_TRY(cudaMemsetAsync(pattern_grad, 0, sizeof(float) * n_active, st));
                CUDA_TRY(launch_planar_adjoint(P1, s->planar, dl1, active_pixels, n_active, w, pattern_grad, st, gate));
            }
            if (P1.regular_sampling)
                CUDA_TRY(launch_planar_frozen(windowed(P1), s->planar, 1, dl1, pattern_grad, active_pixels, n_active, w,
                                              st, gate));
            if (sa)
                CUDA_TRY(inside_adjoint_finish(P1, s->inside, gate, active_pixels, n_active, w, pattern_grad, st));
        }
        if (host_gate != 0)
            CUDA_TRY(launch_adjoint_general(Pw, dose_grad, active_pixels, n_active * spp, spp, seed, w, pattern_grad, st,
                                            host_gate < 0 ? s->planar.gate : nullptr));
        return DRTVAM_OK;
    }
    CUDA_TRY(launch_adjoint_general(Pw, dose_grad, active_pixels, n_active * spp, spp, seed, w, pattern_grad, st));
    return DRTVAM_OK;
}

int drtvam_raygen(DrtvamScene* s, const uint32_t* active_pixels, uint64_t n_active, uint32_t spp, uint32_t seed,
                  float* rays_out, void* stream) {
    if (int rc = check_call(s, n_active, spp)) return rc;
    if (n_active == 0) return DRTVAM_OK;
    if (!active_pixels || !rays_out) return fail(DRTVAM_E_INVALID, "NULL buffer");
    DeviceGuard guard(s->device);
    CUDA_TRY(launch_raygen(s->P, active_pixels, n_active * spp, spp, seed, rays_out,
                           static_cast<cudaStream_t>(stream)));
    return DRTVAM_OK;
}

int drtvam_radon(DrtvamScene* s, const uint32_t* active_pixels, uint64_t n_active, uint32_t spp, uint32_t seed,
                 float* image, void* stream) {
    if (int rc = check_call(s, n_active, spp)) return rc;
    if (!image || (n_active && !active_pixels)) return fail(DRTVAM_E_INVALID, "NULL buffer");
    if (s->P.target.n_tri <= 0) return fail(DRTVAM_E_INVALID, "No target shape found in the scene");  // radon.py:57-58
    DeviceGuard guard(s->device);
    CUDA_TRY(launch_radon(s->P, active_pixels, n_active * spp, spp, seed, ray_weight(s, n_active ? n_active : 1, spp),
                          image, static_cast<cudaStream_t>(stream)));
    return DRTVAM_OK;
}

int drtvam_corner_filter(DrtvamScene* s, const uint32_t* active_pixels, uint64_t n_active, uint32_t spp, uint32_t seed,
                         float dist, float radius, float* image, void* stream) {
    if (int rc = check_call(s, n_active, spp)) return rc;
    if (!image || (n_active && !active_pixels)) return fail(DRTVAM_E_INVALID, "NULL buffer");
    DeviceGuard guard(s->device);
    CUDA_TRY(launch_corner(s->P, active_pixels, n_active * spp, spp, seed, ray_weight(s, n_active ? n_active : 1, spp),
                           dist, radius, image, static_cast<cudaStream_t>(stream)));
    return DRTVAM_OK;
}

int drtvam_count_visits(DrtvamScene* s, const uint32_t* active_pixels, uint64_t n_active, uint32_t spp,
                        uint32_t seed, uint64_t* out_host) {
    if (int rc = check_call(s, n_active, spp)) return rc;
    if (!out_host) return fail(DRTVAM_E_INVALID, "NULL buffer");
    *out_host = 0;
    if (n_active == 0) return DRTVAM_OK;
    DeviceGuard guard(s->device);
    CUDA_TRY(cudaMemset(s->counter_dev, 0, sizeof(unsigned long long)));
    CUDA_TRY(launch_count_general(s->P, active_pixels, n_active * spp, spp, seed, s->counter_dev, nullptr));
    unsigned long long n = 0;
    CUDA_TRY(cudaMemcpy(&n, s->counter_dev, sizeof(n), cudaMemcpyDeviceToHost));
    *out_host = n;
    return DRTVAM_OK;
}

int drtvam_thresholded_loss(int device, const float* dose, const float* dose_b, float alpha, const float* target,
                            uint64_t n, float tl, float tu, float weight_object, float weight_void,
                            float weight_limit, int32_t K, float grad_scale, float* grad, double* loss_out,
                            void* stream) {
    if (!dose || !target || !loss_out) return fail(DRTVAM_E_INVALID, "NULL buffer");
    if (K < 1 || K > 8) return fail(DRTVAM_E_INVALID, "exponent K = %d outside [1, 8]", K);
    if (!(tl < tu))  // loss.py:114-115
        return fail(DRTVAM_E_INVALID, "[ThresholdedLoss] Lower threshold (%g) must be smaller than upper threshold (%g)",
                    tl, tu);
    DeviceGuard guard(device);
    if (!guard.ok) return fail(DRTVAM_E_CUDA, "cudaSetDevice(%d) failed", device);
    CUDA_TRY(launch_thresholded_loss(dose, dose_b, alpha, target, n, tl, tu, weight_object, weight_void, weight_limit,
                                     K, grad_scale, grad, loss_out, static_cast<cudaStream_t>(stream)));
    return DRTVAM_OK;
}

int drtvam_thresholded_loss_sa(int device, const float* dose, const float* dose_b, float alpha, const float* target,
                               uint64_t n_voxels, float tl, float tu, float weight_object, float weight_void,
                               float weight_limit, int32_t K, float grad_scale, float* grad, double* loss_out,
                               void* stream) {
    if (!dose || !target || !loss_out) return fail(DRTVAM_E_INVALID, "NULL buffer");
    if (K < 1 || K > 8) return fail(DRTVAM_E_INVALID, "exponent K = %d outside [1, 8]", K);
    if (!(tl < tu))  // loss.py:114-115
        return fail(DRTVAM_E_INVALID, "[ThresholdedLoss] Lower threshold (%g) must be smaller than upper threshold (%g)",
                    tl, tu);
    DeviceGuard guard(device);
    if (!guard.ok) return fail(DRTVAM_E_CUDA, "cudaSetDevice(%d) failed", device);
    CUDA_TRY(launch_thresholded_loss_sa(dose, dose_b, alpha, target, n_voxels, tl, tu, weight_object, weight_void,
                                        weight_limit, K, grad_scale, grad, loss_out, static_cast<cudaStream_t>(stream)));
    return DRTVAM_OK;
}

// ---- LinearLBFGS on the device (lbfgs.py:146-275) ---------------------------------------------------
#define LB_GUARD(dev)                                                                 \
    DeviceGuard guard(dev);                                                           \
    if (!guard.ok) return fail(DRTVAM_E_CUDA, "cudaSetDevice(%d) failed", dev)

int drtvam_lbfgs_history(int device, const float* p, const float* g, float* p_old, float* g_old, float* s_new,
                         float* y_new, uint64_t n, int32_t first, void* stream) {
    if (!p || !g || !p_old || !g_old || (!first && (!s_new || !y_new))) return fail(DRTVAM_E_INVALID, "NULL buffer");
    LB_GUARD(device);
    CUDA_TRY(launch_lbfgs_history(p, g, p_old, g_old, s_new, y_new, n, first, static_cast<cudaStream_t>(stream)));
    return DRTVAM_OK;
}

int drtvam_lbfgs_direction(int device, const float* const* vectors, int32_t h, uint64_t n, double* gram, int32_t solve_only,
                           double* coef, void* stream) {
    if (!vectors || !gram || !coef) return fail(DRTVAM_E_INVALID, "NULL buffer");
    if (h < 0 || h > 6) return fail(DRTVAM_E_INVALID, "history length %d outside [0, 6]", h);
    LB_GUARD(device);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (solve_only != 1) CUDA_TRY(launch_lbfgs_gram(vectors, 2 * h + 1, n, gram, st));
    if (solve_only != 2) CUDA_TRY(launch_lbfgs_solve(gram, h, h == 0, coef, st));
    return DRTVAM_OK;
}

int drtvam_lbfgs_combine(int device, const float* const* vectors, int32_t h, const double* coef, uint64_t n,
                         float* direction, void* stream) {
    if (!vectors || !coef || !direction) return fail(DRTVAM_E_INVALID, "NULL buffer");
    if (h < 0 || h > 6) return fail(DRTVAM_E_INVALID, "history length %d outside [0, 6]", h);
    LB_GUARD(device);
    CUDA_TRY(launch_lbfgs_combine(vectors, 2 * h + 1, coef, n, direction, static_cast<cudaStream_t>(stream)));
    return DRTVAM_OK;
}

int drtvam_armijo_losses(int device, const float* dose, const float* ddose, const float* target, uint64_t n, float tl,
                         float tu, float weight_object, float weight_void, float weight_limit, int32_t K, int32_t first,
                         int32_t count, double* losses, const float* skip, void* stream) {
    if (!dose || !ddose || !target || !losses) return fail(DRTVAM_E_INVALID, "NULL buffer");
    if (K < 1 || K > 4) return fail(DRTVAM_E_INVALID, "exponent K = %d outside [1, 4]", K);
    if (first < 0 || count < 1 || count > 16 || first + count > 32)
        return fail(DRTVAM_E_INVALID, "trials [%d, %d) outside [0, 32) or more than 16 at once", first, first + count);
    LB_GUARD(device);
    CUDA_TRY(launch_armijo_losses(dose, ddose, target, n, tl, tu, weight_object, weight_void, weight_limit, K, first, count,
                                  losses, skip, static_cast<cudaStream_t>(stream)));
    return DRTVAM_OK;
}

int drtvam_armijo_select(int device, const double* losses, int32_t first, int32_t last, double scale, const double* f0,
                         const double* pattern_term, const double* g_dot_d, double c1, float* out, void* stream) {
    if (!losses || !f0 || !g_dot_d || !out) return fail(DRTVAM_E_INVALID, "NULL buffer");
    if (first < 0 || last <= first || last > 32) return fail(DRTVAM_E_INVALID, "trials [%d, %d) outside [0, 32)", first, last);
    LB_GUARD(device);
    CUDA_TRY(launch_armijo_select(losses, first, last, last, scale, f0, pattern_term, g_dot_d, c1, out,
                                  static_cast<cudaStream_t>(stream)));
    return DRTVAM_OK;
}

int drtvam_lbfgs_update(int device, const float* p, const float* direction, const float* alpha, uint64_t n, float* out,
                        void* stream) {
    if (!p || !direction || !alpha || !out) return fail(DRTVAM_E_INVALID, "NULL buffer");
    LB_GUARD(device);
    CUDA_TRY(launch_lbfgs_update(p, direction, alpha, n, out, static_cast<cudaStream_t>(stream)));
    return DRTVAM_OK;
}

int drtvam_forward_host(DrtvamScene* s, const float* active_data_host, const uint32_t* active_pixels_host,
                        uint64_t n_active, uint32_t spp, uint32_t seed, float out_scale, float* dose_host) {
    if (!s) return fail(DRTVAM_E_INVALID, "scene handle is NULL");
    if (!dose_host || (n_active && (!active_data_host || !active_pixels_host)))
        return fail(DRTVAM_E_INVALID, "NULL buffer");
    DeviceGuard guard(s->device);
    const size_t nb = sizeof(float) * n_active, fb = sizeof(float) * film_elems(s);
    CUDA_TRY(s->host_io[0].reserve(nb ? nb : 4));
    CUDA_TRY(s->host_io[1].reserve(nb ? nb : 4));
    CUDA_TRY(s->host_io[2].reserve(fb));
    CUDA_TRY(cudaMemcpyAsync(s->host_io[0].ptr, active_data_host, nb, cudaMemcpyHostToDevice, nullptr));
    CUDA_TRY(cudaMemcpyAsync(s->host_io[1].ptr, active_pixels_host, nb, cudaMemcpyHostToDevice, nullptr));
    if (int rc = drtvam_forward(s, static_cast<float*>(s->host_io[0].ptr), static_cast<uint32_t*>(s->host_io[1].ptr),
                                n_active, spp, seed, out_scale, static_cast<float*>(s->host_io[2].ptr), 1, nullptr))
        return rc;
    CUDA_TRY(cudaMemcpyAsync(dose_host, s->host_io[2].ptr, fb, cudaMemcpyDeviceToHost, nullptr));
    CUDA_TRY(cudaStreamSynchronize(nullptr));
    return DRTVAM_OK;
}

int drtvam_adjoint_host(DrtvamScene* s, const float* dose_grad_host, const uint32_t* active_pixels_host,
                        uint64_t n_active, uint32_t spp, uint32_t seed, float in_scale, float* pattern_grad_host) {
    if (!s) return fail(DRTVAM_E_INVALID, "scene handle is NULL");
    if (n_active == 0) return DRTVAM_OK;
    if (!dose_grad_host || !active_pixels_host || !pattern_grad_host) return fail(DRTVAM_E_INVALID, "NULL buffer");
    DeviceGuard guard(s->device);
    const size_t nb = sizeof(float) * n_active, fb = sizeof(float) * film_elems(s);
    CUDA_TRY(s->host_io[0].reserve(nb));
    CUDA_TRY(s->host_io[1].reserve(nb));
    CUDA_TRY(s->host_io[2].reserve(fb));
    CUDA_TRY(cudaMemcpyAsync(s->host_io[2].ptr, dose_grad_host, fb, cudaMemcpyHostToDevice, nullptr));
    CUDA_TRY(cudaMemcpyAsync(s->host_io[1].ptr, active_pixels_host, nb, cudaMemcpyHostToDevice, nullptr));
    if (int rc = drtvam_adjoint(s, static_cast<float*>(s->host_io[2].ptr), static_cast<uint32_t*>(s->host_io[1].ptr),
                                n_active, spp, seed, in_scale, static_cast<float*>(s->host_io[0].ptr), nullptr))
        return rc;
    CUDA_TRY(cudaMemcpyAsync(pattern_grad_host, s->host_io[0].ptr, nb, cudaMemcpyDeviceToHost, nullptr));
    CUDA_TRY(cudaStreamSynchronize(nullptr));
    return DRTVAM_OK;
}

}  // extern "C"
