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
};

struct Ray {
    float ox, oy, oz, dx, dy, dz;
};

}  // namespace drtvam
