// Device-side scene parameters for the Dr.TVAM hot path (sm_100a).
// Filled by abi.cu from DrtvamSceneDesc (include/drtvam_b200.h).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "mesh.cuh"

namespace drtvam {

constexpr int kMaxSurfaces = 8;
constexpr int kMaxMeshSurfaces = 4;

struct Surface {
    int kind;             // 0 cylinder about z, 1 origin-centred box, 2 triangle mesh (mesh_surf[mesh])
    int bsdf;             // 0 null, 1 dielectric, 2 absorber
    int interior_medium;  // printing medium inside
    int exterior_medium;  // printing medium outside
    float radius, z0, z1;
    float hx, hy, hz;
    float eta;            // int_ior / ext_ior
    int mesh;             // kind 2: index into SceneParams::mesh_surf
};

struct SceneParams {
    // projector
    int proj_kind, n_patterns, res_x, res_y;
    float pixel_x, pixel_y, aperture_radius, focus_distance;
    // motion
    int motion_analytic, clockwise;
    float distance;
    const float* motion_table;  // device, [n_patterns][12] row-major 3x4
    const double* motion_table_f64;  // the same in double precision (set-up of the planar tables), or NULL
    // integrator
    int max_depth, rr_depth, transmission_only, regular_sampling, sample_time;
    // medium
    float sigma_t, albedo;
    int phase_kind;
    float hg_g;
    // surfaces
    int n_surf;
    Surface surf[kMaxSurfaces];
    // triangle-mesh interfaces: custom vials (geometry.py:98-138) and occluders (geometry.py:55-72)
    MeshView mesh_surf[kMaxMeshSurfaces];
    // sensor / film
    float bmin[3], bmax[3];
    int res[3];
    int surface_aware, sensor_kind;
    float majorant;
    // target mesh; n_tri = 0 unless the run is surface-aware (otherwise the reference moves the mesh
    // 1e5 away, optimize.py:188-191, and no ray meets it)
    MeshView target;
    // z-slab shard of a planar scene (drtvam_scene_set_z_window): the film buffers of a call hold the
    // layers [z_begin, z_end) only and the active list holds the DMD rows [row_begin, row_begin + n_rows)
    // of every pattern.  Defaults: the whole film / every row.
    int z_begin, z_end, row_begin, n_rows;
    // position of every entry of the call's active list in the UNSHARDED list (device, n_active entries),
    // or NULL = the list is the whole one.  A lane's PCG32 stream is seeded with its position in the
    // wavefront (integrators/common.py:67); with this a shard draws exactly the samples the unsharded run
    // draws for the same pixels.
    const uint32_t* sampler_index;
};

// PCG32 stream of lane `lane` (= position in the active list x spp + sample) of a call
__device__ __forceinline__ uint32_t sampler_lane(const SceneParams& P, uint64_t lane, uint32_t spp) {
    if (!P.sampler_index) return static_cast<uint32_t>(lane);
    const uint64_t j = lane / spp;
    return static_cast<uint32_t>(static_cast<uint64_t>(__ldg(P.sampler_index + j)) * spp + (lane - j * spp));
}

// index of (pattern kk of the call, DMD row, column) in the active list of a planar call
__host__ __device__ __forceinline__ size_t active_index(const SceneParams& P, int kk, int row, int col) {
    return (static_cast<size_t>(kk) * P.n_rows + (row - P.row_begin)) * P.res_x + col;
}

struct Ray {
    float ox, oy, oz, dx, dy, dz;
};

}  // namespace drtvam
