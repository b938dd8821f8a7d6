/*
 * drtvam_b200.h -- C-ABI of the B200-native Dr.TVAM hot path
 * (forward projection A and its adjoint A^T).
 *
 * This is the drop-in boundary.  In the reference the path sits behind
 * Mitsuba's Python plugin ABI and is reached through
 *     vol = mi.render(scene, params, integrator, sensor, spp, spp_grad, seed)
 *                                         (src/drtvam/optimize.py:216,294,328)
 *     integrator.render_backward(scene, params, grad_in, sensor, seed, spp)
 *                                         (src/drtvam/integrators/volume.py:97-134)
 * which Dr.Jit turns into JIT megakernels.  There is no FFI in the reference;
 * the entry points below are what a ctypes binding of that path binds
 * (INTEGRATION.md shows the stub).  Each one cites the reference interface it
 * replaces.
 *
 * Conventions
 *  - return 0 on success, a negative DRTVAM_E_* otherwise; the message is
 *    available through drtvam_last_error() (thread local).
 *  - all data pointers are DEVICE pointers on the scene's device unless the
 *    name says host; the library never allocates or frees caller-visible
 *    buffers and never synchronises the stream it is given (the *_host
 *    convenience calls are the exception: they copy and synchronise).
 *  - `stream` is a cudaStream_t passed as void* (NULL = default stream).
 *  - one handle per device; a handle is not thread safe.
 *  - fp32 arithmetic throughout, like the reference's *_ad_mono variants.
 */
#ifndef DRTVAM_B200_H
#define DRTVAM_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DRTVAM_ABI_VERSION 13
#define DRTVAM_MAX_SURFACES 8
#define DRTVAM_MAX_MESH_SURFACES 4

#define DRTVAM_OK 0
#define DRTVAM_E_INVALID -1      /* bad argument / unsupported configuration */
#define DRTVAM_E_CUDA -2         /* CUDA runtime error */
#define DRTVAM_E_TOO_MANY_RAYS -3 /* wavefront > 2^32, common.py:60-65 */
#define DRTVAM_E_NCCL -4         /* NCCL error, incl. asynchronous ones (ncclCommGetAsyncError) */

/* surface kinds / bsdf kinds */
#define DRTVAM_SURF_CYLINDER 0   /* open tube about the z axis: geometry.py:85-89,157-160 */
#define DRTVAM_SURF_BOX 1        /* origin-centred axis-aligned box: geometry.py:198-208 */
#define DRTVAM_SURF_MESH 2       /* triangle mesh added with drtvam_scene_add_mesh_surface: geometry.py:55-72,113-134 */
#define DRTVAM_BSDF_NULL 0       /* geometry.py:90 */
#define DRTVAM_BSDF_DIELECTRIC 1 /* geometry.py:161-165,172-176,201-204,209-213 */
#define DRTVAM_BSDF_ABSORBER 2   /* diffuse occluder, no transmission lobe: geometry.py:57-63 */

#define DRTVAM_PROJ_COLLIMATED 0  /* projector.py:167-196 */
#define DRTVAM_PROJ_TELECENTRIC 1 /* projector.py:199-237 */
#define DRTVAM_PROJ_LENS 2        /* projector.py:241-296 */

#define DRTVAM_PHASE_ISOTROPIC 0
#define DRTVAM_PHASE_RAYLEIGH 1
#define DRTVAM_PHASE_HG 2

#define DRTVAM_SENSOR_DDA 0   /* sensor.py:297-440 */
#define DRTVAM_SENSOR_RATIO 1 /* sensor.py:193-295 */
#define DRTVAM_SENSOR_DELTA 2 /* sensor.py:112-191; scattering media only (volume.py:156-157) */

typedef struct DrtvamSurface {
    int32_t kind, bsdf;
    int32_t interior_medium; /* shape 'interior' is the printing medium */
    int32_t exterior_medium; /* shape 'exterior' is the printing medium */
    float radius, z0, z1;    /* cylinder */
    float hx, hy, hz;        /* box half extents */
    float eta;               /* int_ior / ext_ior */
    float _pad;
} DrtvamSurface;

/* Flat description of what mi.load_dict(scene_dict) builds for this path
 * (optimize.py:61-73, geometry.py to_dict(), projector/sensor/film props). */
typedef struct DrtvamSceneDesc {
    uint32_t struct_size;    /* = sizeof(DrtvamSceneDesc), ABI check */
    uint32_t abi_version;    /* = DRTVAM_ABI_VERSION */
    /* projector (projector.py:41-296) */
    int32_t proj_kind, n_patterns, res_x, res_y;
    float pixel_x, pixel_y, aperture_radius, focus_distance;
    /* motion (motion.py:19-36): analytic circular, or a HOST table of
     * n_patterns row-major 3x4 to_world matrices ("arbitrary motion") */
    int32_t motion_analytic, clockwise;
    float distance;
    float _pad0;
    const float *motion_table_host;
    /* integrator (integrators/common.py:10-22, optimize.py:121-128) */
    int32_t max_depth, rr_depth, transmission_only, regular_sampling, sample_time;
    float print_time;
    /* medium (geometry.py:37-45) */
    float sigma_t, albedo;
    int32_t phase_kind;
    float hg_g;
    /* vial surfaces */
    int32_t n_surf;
    int32_t _pad1;
    DrtvamSurface surf[DRTVAM_MAX_SURFACES];
    /* sensor + film (sensor.py:11-19, film.py:9-21) */
    float bbox_min[3], bbox_max[3];
    int32_t res[3];          /* x, y, z as film.resolution() returns them */
    int32_t surface_aware;
    int32_t sensor_kind;
    float majorant;
    /* optional: the motion table in double precision (HOST, n_patterns row-major 3x4), or NULL.  The
     * per-pass kernels compute in fp32 like the reference; the ONE-OFF set-up of the planar walk tables
     * runs in double, so that the rounding of ~10 mm positions (4e-5 of a 25 um voxel per fp32 ulp) is
     * not baked into every deposit of every pass.  NULL: the fp32 table is promoted. */
    const double *motion_table_host_f64;
} DrtvamSceneDesc;

typedef struct DrtvamScene DrtvamScene;

/* replaces mi.load_dict(scene_dict) for this path (optimize.py:90). */
int drtvam_scene_create(const DrtvamSceneDesc *desc, int device, DrtvamScene **out);
void drtvam_scene_destroy(DrtvamScene *scene);

/* integrator.max_depth = ... (optimize.py:288-289) and the *_ref integrator
 * (optimize.py:231-238) without rebuilding the scene. */
int drtvam_scene_set_depth(DrtvamScene *scene, int32_t max_depth, int32_t rr_depth);

/* A triangle-mesh interface of the traced scene: the 'ply' shapes of a custom
 * vial (CustomVial.to_dict, geometry.py:113-134: dielectric outer / inner wall,
 * face normals) and the occluders of Container.add_occlusions
 * (geometry.py:55-72: by default a black diffuse BSDF, i.e. an absorber, whose
 * exterior is the printing medium).  n_triangles x 9 floats (three world-space
 * vertices each, counter-clockwise seen from outside) in HOST memory.  The
 * surface is appended after the closed-form ones of the scene description (at
 * most DRTVAM_MAX_SURFACES in total, DRTVAM_MAX_MESH_SURFACES meshes); scenes
 * with a mesh interface run on the general tracer.  Builds a BVH; synchronises
 * the device. */
int drtvam_scene_add_mesh_surface(DrtvamScene *scene, const float *triangles_host, uint32_t n_triangles,
                                  int32_t bsdf, float eta, int32_t interior_medium, int32_t exterior_medium);

/* Target mesh of a surface-aware run: the shape 'target' of optimize.py:66-73
 * (null BSDF), n_triangles x 9 floats (three world-space vertices each) in HOST
 * memory.  Its crossings flip the inside/outside flag that selects the film
 * channel (volume.py:196,219; sensor.py:407-408) and do not count as bounces
 * (volume.py:271).  Only pass it while film.surface_aware is on: otherwise the
 * reference moves the mesh 1e5 away (optimize.py:188-191).  n_triangles = 0
 * removes it.  Builds the BVH; synchronises the device. */
int drtvam_scene_set_target(DrtvamScene *scene, const float *triangles_host, uint32_t n_triangles);

/* VolumetricSensor.compute_volume, surface-aware branch (sensor.py:53-110):
 * volumes[res_z][res_y][res_x][2] = (inside, outside) sub-voxel volumes by
 * Monte Carlo with sample_count rays per voxel (reference default 2^14),
 * per-voxel PCG32 streams seeded as sensor.py:83-84.  Synchronises. */
int drtvam_compute_volume(DrtvamScene *scene, uint32_t sample_count, float *volumes, void *stream);

/* utils.discretize (utils.py:83-128): binary occupancy grid of the target,
 * voxels[res_z][res_y][res_x] = 1 where a ray from the voxel centre in the
 * random direction of lane idx of sampler.seed(0, n_voxels) first hits a back
 * face of the target (and the centre lies in the target's bounding box), else
 * 0.  The reference for the loss (optimize.py:131-139).  Asynchronous on
 * `stream`. */
int drtvam_discretize(DrtvamScene *scene, float *voxels, void *stream);

/* inv_vol of volume.py:41-42,54,130 when it is a tensor (surface-aware): a
 * caller-owned DEVICE array of film size that drtvam_forward multiplies the
 * film by and drtvam_adjoint multiplies dose_grad by.  NULL (default) = the
 * scalar out_scale / in_scale only. */
int drtvam_scene_set_inv_volume(DrtvamScene *scene, const float *inv_volume);

/* ---- multi-GPU (SURVEY.md section 8(b),(e)) -------------------------------------------------------
 * The path shards in two ways.  By projection angle (A = sum_k A_k; each ray belongs to one pattern,
 * integrators/common.py:83): every rank passes the slice of active_pixels / active_data of its angles
 * and the full-size partial doses are summed with drtvam_allreduce_dose.  Or, for planar scenes
 * (every ray keeps d_z = 0 and never leaves its voxel layer, sensor.py:345), by z-slab: */

/* z-slab shard: this handle projects the film layers [z_begin, z_end) only.  From now on the dose /
 * dose_grad buffers of drtvam_forward / drtvam_adjoint hold those layers only, i.e.
 * (z_end - z_begin) * res_y * res_x floats, and the caller lists in active_pixels, for every pattern,
 * every column of the DMD rows [*row_begin, *row_end) - the rows whose rays deposit in the window.  No
 * dose collective is needed: every rank owns its layers outright; the loss is a sum over ranks
 * (drtvam_allreduce_scalars) and pattern vectors are sharded by row.  DRTVAM_E_INVALID when the scene is
 * not planar, the rows of the window are not contiguous, or (jittered sampling) a row straddles two
 * layers: shard by angle then.  z_begin = 0, z_end = res_z restores the whole film. */
int drtvam_scene_set_z_window(DrtvamScene *scene, int32_t z_begin, int32_t z_end, int32_t *row_begin,
                              int32_t *row_end);

/* Position of every entry of a shard's active list in the unsharded list (DEVICE array, caller-owned,
 * n_active entries; NULL = identity).  sampler.seed(seed, wavefront_size) gives lane i of the wavefront
 * the i-th PCG32 stream (integrators/common.py:67); with the index a shard draws exactly the samples the
 * unsharded run draws for the same pixels, so N-rank results equal 1-rank results for jittered and
 * scattering runs too. */
int drtvam_scene_set_sampler_index(DrtvamScene *scene, const uint32_t *global_index);

/* A promise that the n_active entries at active_pixels do not change until the next pin (NULL unpins):
 * the planar kernels index the list implicitly and check its layout on the device in every call
 * (0.3 ms at 6.4e7 pixels); for a pinned list the check runs once (synchronising the stream of the
 * first call) and later calls skip it and the stand-by launch of the other kernel family.  The host
 * mirror pins projector.active_pixels, which only params.update() replaces (optimize.py:157-159). */
int drtvam_scene_pin_pixels(DrtvamScene *scene, const uint32_t *active_pixels, uint64_t n_active);

/* NCCL plumbing for bindings without torch.distributed (the Python mirror uses torch.distributed and
 * never calls these).  libnccl is resolved at run time (the copy already loaded in the process wins), so
 * the library has no link-time dependency on it.  drtvam_comm_unique_id fills 128 bytes (ncclUniqueId) on
 * rank 0; the caller hands them to every rank by its own means; drtvam_comm_init joins the communicator
 * (one per handle, on the handle's device).  Both collectives reduce in place with ncclSum on `stream`
 * and query ncclCommGetAsyncError before and after enqueueing, so a dead peer surfaces as
 * DRTVAM_E_NCCL instead of a hang at the next synchronisation. */
int drtvam_comm_unique_id(void *id_out_128_bytes);
int drtvam_comm_init(DrtvamScene *scene, const void *unique_id_128_bytes, int32_t rank, int32_t world);
int drtvam_comm_destroy(DrtvamScene *scene);
/* sum of the per-rank partial doses of an angle-sharded forward (n floats, DEVICE) */
int drtvam_allreduce_dose(DrtvamScene *scene, float *dose, uint64_t n, void *stream);
/* sum of n doubles (DEVICE): the loss of a z-slab shard, batched L-BFGS dot products (lbfgs.py:191-231) */
int drtvam_allreduce_scalars(DrtvamScene *scene, double *values, uint64_t n, void *stream);

/* Which kernel family a call will use for this scene and spp:
 * 0 = general per-lane tracer, 1 = planar per-ray tile DDA,
 * 2 = planar layer-parallel kernels (deterministic fast path),
 * 3 = planar layer-lane kernels with a per-lane segment table (jittered sampling),
 * 4 = scattering medium: the kernels of 3 for the first medium segment of every lane (the unscattered chord,
 *     deposited in full whether or not the lane collides on it, volume.py:216) and the general tracer for the
 *     segments after the first collision. */
int drtvam_scene_plan(DrtvamScene *scene, uint64_t n_active, uint32_t spp);
/* 0 = let the library choose, 1 = force the general tracer, 2 = force the
 * planar per-ray tile DDA where the scene is planar (A/B tests). */
int drtvam_scene_force_general(DrtvamScene *scene, int32_t on);
/* Sizes of the per-scene tables behind the planar kernels (bench.py's roofline is counted from them; the
 * reference has no counterpart, its kernels are JIT-generated per call).  which:
 *   0 = (4x4 patch, segment) coefficient pairs of the patch-gather forward (0 = tables not in use),
 *   1 = voxel layers that receive dose from the DMD rows of this handle (inside the z window),
 *   2 = target-crossing records of a surface-aware deterministic scene,
 *   3 = bytes of the coefficient table of the patch-gather forward (64 B of coefficients + 4 B ray index per pair),
 *   4 = fixed-point tile words of the jittered forward kernels that came within a factor 2 of overflow since the
 *       scene was created (synchronises; 0 unless rays are focused 16 times denser than the scale assumes). */
int drtvam_scene_stat(DrtvamScene *scene, int32_t which, uint64_t *value);

/* A p : VolumeIntegrator.render (volume.py:18-56).  dose has
 * res_z*res_y*res_x*C floats (C = 2 when surface_aware; film.py:17-21,
 * flat index sensor.py:405-409) and receives film * out_scale, where the
 * caller passes out_scale = 1/voxel_volume (volume.py:41-42,54).  active_data
 * may be signed (lbfgs.py:240-248).  zero_first != 0 clears the film first
 * (film.clear(), volume.py:33); otherwise the call accumulates. */
int drtvam_forward(DrtvamScene *scene, const float *active_data, const uint32_t *active_pixels,
                   uint64_t n_active, uint32_t spp, uint32_t seed, float out_scale,
                   float *dose, int32_t zero_first, void *stream);

/* drtvam_forward for callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid
 * cudaMalloc).  No counterpart in the reference.  Same
 * arguments and result as drtvam_forward, for the deterministic planar plan only (drtvam_scene_plan() == 2
 * with its patch-gather tables, one film channel, whole angles; anything else is DRTVAM_E_INVALID).  The
 * scratch drtvam_forward keeps in a cudaMalloc'ed buffer - the [angle][column][layer] repack of the
 * patterns - lives in a static __device__ array of the library image instead: 4 Mi floats = 16 MiB, held by
 * every GPU that loads the library until the process exits.  A call needs (angles of the call) x res_x x
 * (active layers padded to 32) floats; larger calls are rejected with DRTVAM_E_INVALID: split them by
 * angles (zero_first = 0 accumulates).  The scene's tables must exist already (call drtvam_scene_plan or
 * drtvam_forward once: building them allocates).  The call itself makes no cudaMalloc / cudaFree; its
 * first use on a device resolves the array's address with cudaGetSymbolAddress.  All calls on a device
 * share the array: each call overwrites the part it uses before reading it (nothing needs clearing), and
 * calls that overlap on the GPU (different streams, concurrent graphs) race on it. */
int drtvam_forward_noalloc(DrtvamScene *scene, const float *active_data, const uint32_t *active_pixels,
                           uint64_t n_active, uint32_t spp, uint32_t seed, float out_scale, float *dose,
                           int32_t zero_first, void *stream);

/* A^T g : VolumeIntegrator.render_backward (volume.py:97-134).  dose_grad is
 * grad_in; in_scale = 1/voxel_volume (volume.py:130).  Overwrites
 * pattern_grad[n_active] with d loss / d active_data. */
int drtvam_adjoint(DrtvamScene *scene, const float *dose_grad, const uint32_t *active_pixels,
                   uint64_t n_active, uint32_t spp, uint32_t seed, float in_scale,
                   float *pattern_grad, void *stream);

/* TVAMIntegrator.sample_rays (common.py:70-116): rays_out[lanes][6] = o,d. */
int drtvam_raygen(DrtvamScene *scene, const uint32_t *active_pixels, uint64_t n_active,
                  uint32_t spp, uint32_t seed, float *rays_out, void *stream);

/* Pixel pre-filters of optimize.py:143-185.  Both write
 * image[n_patterns * res_y * res_x] (cleared first): image[pixel] += L * ray weight
 * over the spp lanes of each listed pixel (radon.py:42-44); the caller keeps the
 * pixels with image > 0 (optimize.py:151,173).
 * drtvam_radon: RadonIntegrator.render / sample (integrators/radon.py:13-106), the
 * absorbed fraction along the parts of the path inside the target mesh; needs the
 * target (drtvam_scene_set_target) and uses the scene's max_depth
 * (optimize.py:145-148 sets 5).
 * drtvam_corner_filter: CornerIntegrator.render / sample
 * (integrators/filter_corner.py:17-66), 1 unless the first hit lies within `radius`
 * of the vial edge |x| = |y| = dist. */
int drtvam_radon(DrtvamScene *scene, const uint32_t *active_pixels, uint64_t n_active, uint32_t spp,
                 uint32_t seed, float *image, void *stream);
int drtvam_corner_filter(DrtvamScene *scene, const uint32_t *active_pixels, uint64_t n_active, uint32_t spp,
                         uint32_t seed, float dist, float radius, float *image, void *stream);

/* Number of DDA trips (sensor.py:383) one forward pass makes; the unit of
 * work of the roofline.  Synchronises and returns the count in *out_host. */
int drtvam_count_visits(DrtvamScene *scene, const uint32_t *active_pixels, uint64_t n_active,
                        uint32_t spp, uint32_t seed, uint64_t *out_host);

/* HOST-buffer convenience calls: copy in, run, copy out, synchronise. */
int drtvam_forward_host(DrtvamScene *scene, const float *active_data_host,
                        const uint32_t *active_pixels_host, uint64_t n_active, uint32_t spp,
                        uint32_t seed, float out_scale, float *dose_host);
int drtvam_adjoint_host(DrtvamScene *scene, const float *dose_grad_host,
                        const uint32_t *active_pixels_host, uint64_t n_active, uint32_t spp,
                        uint32_t seed, float in_scale, float *pattern_grad_host);

/* Caller of the path (next row): ThresholdedLoss on the
 * dose volume, loss.py:102-132 with Loss.__call__'s reduction (loss.py:58-59),
 * binary target (target > 0 = inside).  Evaluates at dose + alpha * dose_b
 * when dose_b != NULL (the Armijo trial of lbfgs.py:258).  *loss_out (device
 * double) receives the SUM over voxels; grad (may be NULL) receives
 * grad_scale * d sum / d dose (grad_scale = 1/N for reduction 'mean'). */
int drtvam_thresholded_loss(int device, const float *dose, const float *dose_b, float alpha, const float *target,
                            uint64_t n, float tl, float tu, float weight_object, float weight_void,
                            float weight_limit, int32_t K, float grad_scale, float *grad, double *loss_out,
                            void *stream);

/* The same loss on a SURFACE-AWARE film (two channels per voxel: dose deposited inside / outside the target,
 * film.py:17-21) against the two-channel target of sub-voxel volumes (sensor.py:53-110): per voxel
 * w_in eval_in(dose[0]) + w_out eval_out(dose[1]) with w = target / (target[0] + target[1]), the weighting of
 * Loss.__call__ (loss.py:44-53).  n_voxels voxels; dose, dose_b, target and grad hold 2 n_voxels floats. */
int drtvam_thresholded_loss_sa(int device, const float *dose, const float *dose_b, float alpha, const float *target,
                               uint64_t n_voxels, float tl, float tu, float weight_object, float weight_void,
                               float weight_limit, int32_t K, float grad_scale, float *grad, double *loss_out,
                               void *stream);

/* ---- LinearLBFGS on the device (lbfgs.py:146-275) --------------------------
 * All buffers are DEVICE buffers of fp32 pattern vectors of n elements unless stated; nothing below
 * synchronises or reads a result back, so a whole iteration is queued without a host round trip.
 *
 * history (lbfgs.py:176-196): s_new = p - p_old, y_new = g - g_old (skipped when first != 0), then
 * p_old = p, g_old = g. */
int drtvam_lbfgs_history(int device, const float *p, const float *g, float *p_old, float *g_old, float *s_new,
                         float *y_new, uint64_t n, int32_t first, void *stream);
/* two-loop recursion (lbfgs.py:213-234).  vectors: HOST array of 2h + 1 device pointers
 * [s_0 .. s_{h-1}, y_0 .. y_{h-1}, g], oldest pair first.  gram: DEVICE double[(2h+1)^2], the Gram matrix of
 * the vectors - every dot product of the recursion, each vector read once (13 separate passes in the
 * reference).  coef: DEVICE double[16]: coef[j] with d = - sum_j coef[j] vectors[j], coef[13] = g . d
 * (lbfgs.py:251-254), coef[14] = 1 when a curvature y_i . s_i was zero or NaN.  solve_only: 0 = both
 * steps; 2 = only form the Gram matrix (the caller sums it over ranks: pattern vectors are sharded, this is
 * the ONE collective of the iteration); 1 = only run the recursion on `gram`. */
int drtvam_lbfgs_direction(int device, const float *const *vectors, int32_t h, uint64_t n, double *gram,
                           int32_t solve_only, double *coef, void *stream);
int drtvam_lbfgs_combine(int device, const float *const *vectors, int32_t h, const double *coef, uint64_t n,
                         float *direction, void *stream);
/* Armijo back-tracking (lbfgs.py:252-266) without a host decision per trial, in stages of trials
 * [first, first + count), count <= 16: losses[k] (DEVICE double[32], cleared by the caller before the first
 * stage) += sum over voxels of the thresholded loss (loss.py:102-132) at dose + 2^-k ddose, every trial of
 * the stage in ONE sweep of the two volumes.  select looks at trials [first, last): the first k with
 * scale * losses[k] + pattern_term <= f0 + c1 * 2^-k * g_dot_d (all DEVICE doubles; pattern_term may be
 * NULL) ends the search: out[0] = alpha, out[1] = trials used, out[2] = the loss at alpha, out[3] = 1
 * (DEVICE floats).  Passing out + 3 as `skip` makes the next stage return in its first instruction once
 * the search is over - alpha = 1 is accepted in almost every iteration, so the usual cost is the first
 * stage of 4 trials.  With the dose sharded over ranks (z-slab) the caller sums `losses` between the two. */
int drtvam_armijo_losses(int device, const float *dose, const float *ddose, const float *target, uint64_t n, float tl,
                         float tu, float weight_object, float weight_void, float weight_limit, int32_t K,
                         int32_t first, int32_t count, double *losses, const float *skip, void *stream);
int drtvam_armijo_select(int device, const double *losses, int32_t first, int32_t last, double scale, const double *f0,
                         const double *pattern_term, const double *g_dot_d, double c1, float *out, void *stream);
/* out = p + alpha * direction with alpha read from DEVICE memory (lbfgs.py:271-274) */
int drtvam_lbfgs_update(int device, const float *p, const float *direction, const float *alpha, uint64_t n, float *out,
                        void *stream);

/* number of kernels this library has launched since load (bench gpu_launches) */
uint64_t drtvam_launch_count(void);

const char *drtvam_last_error(void);
int drtvam_abi_version(void);

#ifdef __cplusplus
}
#endif
#endif /* DRTVAM_B200_H */
