// Host-side launch interface of the kernel translation units.
#pragma once
#include <cstdint>
#include <vector>
#include <cuda_runtime.h>
#include "device_scene.cuh"
#include "planar.cuh"

namespace drtvam {

void count_launch();  // bumps the library-wide launch counter (abi.cu)

// general per-lane tracer (general_kernels.cu)
cudaError_t launch_forward_general(const SceneParams& P, const float* active_data, const uint32_t* active_pixels,
                                   uint64_t n_lanes, uint32_t spp, uint32_t seed, float ray_weight, float* film,
                                   cudaStream_t stream, const int* gate = nullptr, bool hybrid = false);
cudaError_t launch_adjoint_general(const SceneParams& P, const float* dL, const uint32_t* active_pixels,
                                   uint64_t n_lanes, uint32_t spp, uint32_t seed, float ray_weight, float* grad,
                                   cudaStream_t stream, const int* gate = nullptr, bool hybrid = false);
// hybrid = true (needs a gate word): where the gate reads 0 the planar per-lane kernels have served the first
// medium segment of every lane, so the tracer skips it and ADDS to the gradient; elsewhere it does the whole lane
bool general_supports_hybrid(const SceneParams& P);
cudaError_t launch_raygen(const SceneParams& P, const uint32_t* active_pixels, uint64_t n_lanes, uint32_t spp,
                          uint32_t seed, float* out, cudaStream_t stream);
cudaError_t launch_count_general(const SceneParams& P, const uint32_t* active_pixels, uint64_t n_lanes, uint32_t spp,
                                 uint32_t seed, unsigned long long* total, cudaStream_t stream);


// planar z-separable fast path (planar_kernels.cu); `gate` (device int) is 0
// when the planar kernels own the call and non-zero when the general tracer does
struct PlanarHost {
    bool valid = false;
    int flags = 0;             // kPlanarFlag* reported by the table builder
    int n_rows_in_sensor = 0;  // DMD rows whose z lies inside the sensor box
    float4* seg_a = nullptr;
    float2* seg_b = nullptr;
    SegmentD* seg_d = nullptr;   // double-precision copy for the set-up kernels (planar_f64.cuh)
    int* layer_row_begin = nullptr;
    int* layer_rows = nullptr;
    int* tile_off = nullptr;
    int* tile_clo = nullptr;
    int* gate = nullptr;  // [0] per-call pixel-layout gate, [1] build flags
    int* frozen_list = nullptr;    // (angle, column) rays with a frozen DDA axis (sensor.py:359), walked generally
    int n_frozen = 0;
    PlanarPlan plan{};
    size_t smem_bytes = 0;
    std::vector<int> row_layer;  // host copy: film layer of every DMD row, -1 = outside the sensor
    std::vector<float> row_z;    // host copy: world z of every DMD row's pixel centres
};
cudaError_t planar_build(const SceneParams& P, PlanarHost& H);
void planar_release(PlanarHost& H);
cudaError_t launch_planar_check(const SceneParams& P, PlanarHost& H, const uint32_t* active_pixels, uint64_t n_active,
                                cudaStream_t stream);
cudaError_t launch_planar_forward(const SceneParams& P, PlanarHost& H, const float* active_data,
                                  const uint32_t* active_pixels, uint64_t n_active, float ray_weight, float* film,
                                  cudaStream_t stream, const int* gate = nullptr);
cudaError_t launch_planar_adjoint(const SceneParams& P, PlanarHost& H, const float* dL, const uint32_t* active_pixels,
                                  uint64_t n_active, float ray_weight, float* grad, cudaStream_t stream,
                                  const int* gate = nullptr);

// per-(tile, angle) first crossing column and prefix sums of the crossing-column counts
cudaError_t planar_tile_ranges(const SceneParams& P, const float4* seg_a, const float2* seg_b, int tiles_x, int tiles_y,
                               int tile_w, int tile_h, int* tile_off_dev, int* tile_clo_dev, float margin = 0.f);
// per-tile column lists instead of ranges (exact where a range would span the gap between two ray families)
cudaError_t planar_tile_lists(const SceneParams& P, const float4* seg_a, const float2* seg_b, int tiles_x, int tiles_y,
                              int tile_w, int tile_h, int* tile_off_dev, int* tile_ebase_dev, int** list_dev,
                              long long* n_entries, float margin = 0.f);
// mode 0: deposit the frozen-axis rays into the film; mode 1: write their pattern gradients
cudaError_t launch_planar_frozen(const SceneParams& P, PlanarHost& H, int mode, const float* in, float* out,
                                 const uint32_t* active_pixels, uint64_t n_active, float ray_weight,
                                 cudaStream_t stream, const int* gate = nullptr);

// layer-parallel planar kernels (planarz_kernels.cu), the deterministic fast path
struct PlanarZSub {      // a run of z-blocks that all hold 32 * lanes_l active layers
    PlanarZPlan plan{};
    int* tile_off = nullptr;
    int* tile_clo = nullptr;
    int* tile_ebase = nullptr;
    float4* entries = nullptr;
    // visit lists of the adjoint (planarv_kernel): per entry a head {weight, start | n, offset} and, from
    // `offset` on, the n deposit factors of its voxels with the step direction in the lowest mantissa bit
    float4* vhead = nullptr;
    float* vde = nullptr;
    int* v_off = nullptr;     // prefix tables of the kept (non-empty) entries
    int* v_ebase = nullptr;
    size_t smem_bytes = 0;
    int n_blocks = 0;
};
struct PlanarZHost {
    bool valid = false;
    int n_sub = 0;
    PlanarZSub sub[3];       // e.g. 400 layers = 3 blocks of 128 (4 per lane) + 1 block of 32 (1 per lane)
    int nz = 0, nz_pad = 0;
    int rows_max = 0;        // most DMD rows that deposit in one layer
    int* zl = nullptr;
    int* a_row_begin = nullptr;
    int* a_rows = nullptr;
    int* row_a = nullptr;
    int* layer_a = nullptr;    // [res_z] active layer of film layer z, -1 = no DMD row deposits there (or outside the z window)
    int* row_ri = nullptr;     // [res_y] position of DMD row r among the rows of its layer
    unsigned* pmax = nullptr;  // largest |pattern value| of the current forward call (bits of a float): fp16 scale
    float* scratch = nullptr;  // [angles][columns][active layers] repack of patterns / gradients
    size_t scratch_bytes = 0;
};
// patch-gather forward (planarz_kernels.cu): per (4x4 voxel patch, angle, column) a dense block of
// 16 deposit coefficients, so that the forward pass keeps its accumulators in registers
struct PlanarGHost {
    bool valid = false;
    PlanarZPlan plan{};        // 4x4 tiles; entries unused
    int* tile_off = nullptr;
    int* tile_clo = nullptr;
    int* tile_ebase = nullptr;
    float4* coef = nullptr;    // [pairs][4] float4 = 16 coefficients, voxel (yy, xx) at 4 * yy + xx
    uint32_t* coef_h = nullptr;   // the same as {hi, lo} fp16 pairs of coefficient / coef_inv_scale (planarg_h16_kernel)
    float coef_inv_scale = 1.f;
    int* ray = nullptr;        // [pairs] k * res_x + c
    long long pairs = 0;       // after the pairs that touch nothing are dropped
    long long candidates = 0;  // before
};
cudaError_t planarz_build(const SceneParams& P, const PlanarHost& base, const std::vector<int>& row_layer,
                          PlanarZHost& H, bool walks = true);

// jittered planar passes (planarj_kernels.cu): regular_sampling = false, every lane its own segment
struct PlanarJHost {
    bool valid = false;
    int tiles_x = 0, tiles_y = 0, tile_w = 0, tile_h = 0;
    int* tile_off = nullptr;       // [n_tiles][n_patterns + 1] prefix of the (conservative) column counts
    int* tile_clo = nullptr;       // the column list: angle * res_x + column of every (tile, angle, column) entry
    int* tile_ebase = nullptr;     // [n_tiles + 1] first entry of every tile in the list
    void* table = nullptr;         // per-lane segment table of the current angle chunk (forward passes)
    size_t table_bytes = 0;
    void* table_adj = nullptr;     // the same for adjoint passes
    size_t table_adj_bytes = 0;
    float* gT = nullptr;           // adjoint: per-lane gradients in table order
    size_t gT_bytes = 0;
    size_t smem_bytes = 0;
    unsigned* wmax = nullptr;      // [0] largest |weight x pattern value| of the current forward table (float bits),
                                   // [1] fixed-point tile words that came close to overflow (never seen)
    // the table still holds the geometry of the last forward pass (same pixels, seed, spp, weight)
    bool cache_valid = false;
    uint32_t cache_seed = 0, cache_spp = 0;
    const uint32_t* cache_pixels = nullptr;
    uint64_t cache_n = 0;
    float cache_weight = 0.f;
    const uint32_t* cache_index = nullptr;
};
cudaError_t planarj_build(const SceneParams& P, const PlanarHost& base, const PlanarZHost& Zh, PlanarJHost& J);
void planarj_release(PlanarJHost& J);
cudaError_t launch_planarj_forward(const SceneParams& P, const SceneParams& Pw, PlanarZHost& Zh, PlanarJHost& J,
                                   const int* gate, const float* active_data, const uint32_t* active_pixels,
                                   uint64_t n_active, uint32_t spp, uint32_t seed, float ray_weight, float* film,
                                   bool cacheable, cudaStream_t stream);
cudaError_t launch_planarj_adjoint(const SceneParams& P, const SceneParams& Pw, PlanarZHost& Zh, PlanarJHost& J,
                                   const int* gate, const float* dL, const uint32_t* active_pixels, uint64_t n_active,
                                   uint32_t spp, uint32_t seed, float ray_weight, float* grad, cudaStream_t stream);
cudaError_t planarg_build(const SceneParams& P, const PlanarHost& base, const PlanarZHost& Zh, PlanarGHost& G);
void planarg_release(PlanarGHost& G);
// `lane_mask` (may be null): bit ((angle * res_x + column) * nz_pad + active layer) set = the lane is not this
// kernel's (the tile lanes of a surface-aware scene, inside_kernels.cu); its pattern value is cleared after the repack
cudaError_t launch_planarg_forward(const SceneParams& P, PlanarZHost& H, PlanarGHost& G, const int* gate,
                                   const float* active_data, const uint32_t* active_pixels, uint64_t n_active,
                                   float ray_weight, float* film, bool fresh, cudaStream_t stream,
                                   const uint32_t* lane_mask = nullptr);
// the same forward on the static scratch of the library image: no allocation (planarz_kernels.cu)
size_t planarg_noalloc_capacity();   // floats: angles of the call x res_x x padded layers must fit
cudaError_t launch_planarg_forward_noalloc(const SceneParams& P, const PlanarZHost& H, const PlanarGHost& G, int device,
                                           const int* gate, const float* active_data, const uint32_t* active_pixels,
                                           uint64_t n_active, float ray_weight, float* film, bool fresh,
                                           cudaStream_t stream);
void planarz_release(PlanarZHost& H);
cudaError_t launch_planarz_forward(const SceneParams& P, PlanarZHost& H, const int* gate, const float* active_data,
                                   const uint32_t* active_pixels, uint64_t n_active, float ray_weight, float* film,
                                   cudaStream_t stream);
struct InsideHost;
// `inside` (may be null): a surface-aware scene whose tile lanes are cleared in the repacked gradient and then
// receive the gather over their own records (inside_adjoint_tiles) before the unpack
cudaError_t launch_planarz_adjoint(const SceneParams& P, PlanarZHost& H, const int* gate, const float* dL,
                                   const uint32_t* active_pixels, uint64_t n_active, float ray_weight, float* grad,
                                   cudaStream_t stream, InsideHost* inside = nullptr);

// target mesh + surface-aware sub-voxel volumes (mesh_kernels.cu)
struct MeshHost {
    BvhNode* nodes = nullptr;
    float4* tri = nullptr;
    MeshView view{};
};
cudaError_t mesh_build(const float* tri_host, uint32_t n_tri, MeshHost& H);
void mesh_release(MeshHost& H);
// sensor.py:47-110; synchronises the stream (one-off set-up work)
cudaError_t launch_compute_volume(const SceneParams& P, uint32_t sample_count, float* volumes, cudaStream_t stream);
// utils.py:83-128: binary occupancy of the target at the voxel centres, voxels[res_z][res_y][res_x].
cudaError_t launch_discretize(const SceneParams& P, float* voxels, cudaStream_t stream);
cudaError_t launch_mul_inplace(float* a, const float* b, size_t n, cudaStream_t stream);
cudaError_t launch_mul(const float* a, const float* b, float* out, size_t n, cudaStream_t stream);

// inside-target segment records: the surface-aware film on top of the planar kernels (inside_kernels.cu)
struct InsideHost {
    bool valid = false;
    // ODD lanes (a frozen DDA axis, a crossing that leaves the lane's layer ...): per-lane record lists, walked by
    // one thread per lane with global atomics; the planar kernels deposit the lane's unsplit segment and a
    // NEGATIVE record cancels its tail
    uint32_t* off = nullptr;    // [lanes + 1] first record of every lane of the full pixel grid
    float4* seg = nullptr;      // [n_rec][2] {origin, length}, {direction, throughput at the origin}
    float* t_lo = nullptr;      // [n_rec] > 0: walk the record from here on only (the negative record)
    uint64_t n_rec = 0;
    uint32_t* odd_lanes = nullptr;  // the lanes that have such records
    uint32_t n_odd = 0;
    float* planes = nullptr;    // [2][n_vox] plane A (outside channel), plane B (inside channel)
    size_t n_vox = 0;
    // TILE lanes (all other lanes that meet the target): masked out of the planar kernels altogether; all their
    // sub-segments, head included, sit in a table ordered [angle][column][32-layer block][record][layer lane]
    // and are marched by a shared-memory tile kernel
    bool tiles = false;
    uint2* grp = nullptr;       // [angle][column][block] {first record row of the group, record rows}
    float4* table = nullptr;    // [rows][32 lanes][2] {ox, oy, dx, dy}, {+-length (sign: plane B), throughput, -, -}
    float2* rbox = nullptr;     // [rows] stretch of the row along the centre ray of its (angle, column)
    const float4* seg_a = nullptr;  // the centre rays (owned by the planar plan)
    unsigned* stat = nullptr;   // fixed-point scale inputs and the near-overflow counter of the forward tile kernel
    uint64_t n_rows = 0;
    uint64_t n_tile_rec = 0;    // records in the table (statistics)
    uint32_t* mask = nullptr;   // bit ((angle * res_x + column) * nz_pad + active layer): a tile lane
    int* tile_off = nullptr;    // per-tile column lists (planar_tile_lists)
    int* tile_clo = nullptr;
    int* tile_ebase = nullptr;
    int tiles_x = 0, tiles_y = 0, tile_w = 0, tile_h = 0;
    size_t smem_bytes = 0;
};
// `allow_tiles`: the deterministic forward / adjoint of this scene run on the patch-gather and layer-parallel
// kernels, which can mask lanes (no z window, one DMD row per layer)
cudaError_t inside_build(const SceneParams& P, const SceneParams& P1, const PlanarHost& base, const PlanarZHost& Zh,
                         bool allow_tiles, InsideHost& H);
void inside_release(InsideHost& H);
cudaError_t inside_forward_begin(InsideHost& H, cudaStream_t stream);
cudaError_t inside_forward_finish(const SceneParams& P1, InsideHost& H, const PlanarZHost& Zh, const int* gate,
                                  const float* active_data, const uint32_t* active_pixels, uint64_t n_active,
                                  float ray_weight, float* film, cudaStream_t stream);
cudaError_t inside_adjoint_begin(InsideHost& H, const float* dL, cudaStream_t stream);
// the tile lanes' part of the adjoint: adds to gT ([angle of the call][column][active layer], the repacked gradient
// of the layer-parallel adjoint) between its gather and its unpack
cudaError_t inside_adjoint_tiles(const SceneParams& P1, InsideHost& H, const PlanarZHost& Zh, const int* gate,
                                 const uint32_t* active_pixels, uint64_t n_active, float ray_weight, float* gT,
                                 cudaStream_t stream);
cudaError_t inside_adjoint_finish(const SceneParams& P1, InsideHost& H, const int* gate, const uint32_t* active_pixels,
                                  uint64_t n_active, float ray_weight, float* grad, cudaStream_t stream);

// one-shot pixel pre-filters (filter_kernels.cu); `image` has n_patterns * res_y * res_x floats and is cleared
cudaError_t launch_radon(const SceneParams& P, const uint32_t* active_pixels, uint64_t n_lanes, uint32_t spp,
                         uint32_t seed, float ray_weight, float* image, cudaStream_t stream);
cudaError_t launch_corner(const SceneParams& P, const uint32_t* active_pixels, uint64_t n_lanes, uint32_t spp,
                          uint32_t seed, float ray_weight, float dist, float radius, float* image,
                          cudaStream_t stream);

// fused thresholded loss + gradient on the dose volume (loss_kernels.cu)
cudaError_t launch_thresholded_loss(const float* x, const float* x_b, float alpha, const float* target, size_t n,
                                    float tl, float tu, float w_obj, float w_void, float w_lim, int K, float grad_scale,
                                    float* grad, double* loss_out, cudaStream_t stream);
// the same for a surface-aware (two-channel) film and target; n_vox voxels, 2 n_vox floats per array
cudaError_t launch_thresholded_loss_sa(const float* x, const float* x_b, float alpha, const float* target, size_t n_vox,
                                    float tl, float tu, float w_obj, float w_void, float w_lim, int K, float grad_scale,
                                    float* grad, double* loss_out, cudaStream_t stream);


// LinearLBFGS on the device (lbfgs_kernels.cu)
cudaError_t launch_lbfgs_history(const float* p, const float* g, float* p_old, float* g_old, float* s, float* y, size_t n,
                                 int first, cudaStream_t st);
cudaError_t launch_lbfgs_gram(const float* const* vectors, int n_vec, size_t n, double* gram, cudaStream_t st);
cudaError_t launch_lbfgs_solve(const double* gram, int h, int first_step, double* coef, cudaStream_t st);
cudaError_t launch_lbfgs_combine(const float* const* vectors, int n_vec, const double* coef, size_t n, float* out,
                                 cudaStream_t st);
cudaError_t launch_armijo_losses(const float* dose, const float* ddose, const float* target, size_t n, float tl, float tu,
                                 float w_obj, float w_void, float w_lim, int K, int k0, int count, double* losses,
                                 const float* skip, cudaStream_t st);
cudaError_t launch_armijo_select(const double* losses, int k0, int k1, int trials, double scale, const double* f0,
                                 const double* pattern_term, const double* gd, double c1, float* out, cudaStream_t st);
cudaError_t launch_lbfgs_update(const float* p, const float* d, const float* alpha, size_t n, float* out, cudaStream_t st);

}  // namespace drtvam
