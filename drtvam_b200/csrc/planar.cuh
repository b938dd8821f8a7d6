// Planar (z-separable) fast path: shared declarations.
//
// When every ray of the projector has d_z = 0 (collimated beam, rotation about
// z, vertical-walled vial, no scattering) a ray never leaves the z-layer of
// voxels it starts in, and its single medium segment depends only on
// (angle, DMD column), not on the DMD row.  The 3-D scatter of
// DDAVolumetricSensor.accumulate (sensor.py:306-440) then decomposes into
// res_z independent 2-D scatters, which lets a CTA own a 2-D tile of one layer
// in shared memory for all angles.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace drtvam {

// One medium segment per (angle, column): origin, direction, length and the
// throughput at its origin times sigma_a / sigma_t (0 = the ray never enters
// the medium).
struct PlanarPlan {
    const float4* seg_a;         // [n_patterns * res_x]  ox, oy, dx, dy
    const float2* seg_b;         // [n_patterns * res_x]  tmax, w0
    const int* layer_row_begin;  // [res_z + 1] CSR over layer_rows
    const int* layer_rows;       // DMD rows sorted by the layer they deposit in
    const int* tile_off;         // [n_tiles][n_patterns + 1] prefix of per-angle column counts
    const int* tile_clo;         // [n_tiles][n_patterns] first column crossing the tile
    int tiles_x, tiles_y, tile_w, tile_h, pitch, groups;
};

// flags written by the table builder
constexpr int kPlanarFlagMultiSegment = 1;  // a ray had more than one medium segment
constexpr int kPlanarFlagNotPlanar = 2;     // a segment had d_z != 0

}  // namespace drtvam
