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

#include "device_scene.cuh"
#include "planar_f64.cuh"

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

// Layer-parallel ("z-lane") plan: a warp marches ONE (angle, column) ray and
// its 32 lanes deposit into 32 different z-layers at the same (x, y), so the
// traversal is warp-uniform, shared-memory banks never conflict and the
// pattern / gradient values are read and written as 128-byte lines from a
// [angle][column][active layer] repack of active_data.
struct PlanarZPlan {
    const float4* seg_a;     // shared with PlanarPlan
    const float2* seg_b;
    const SegmentD* seg_d;   // the same segments in double precision (set-up kernels only)
    const int* tile_off;     // [n_tiles][n_patterns + 1] prefix of per-angle column counts
    const int* tile_clo;     // [n_tiles][n_patterns]; the jittered kernels: list of angle * res_x + column per entry
    const int* tile_ebase;   // [n_tiles + 1] first walk entry of every tile
    const float4* entries;   // 2 x float4 per (tile, angle, column): the clipped walk, ready to march
    const float4* vhead;     // visit lists of the adjoint: head per KEPT entry (.w = angle * res_x + column) ...
    const float* vde;        // ... and the deposit factors of its voxels
    const int* v_off;        // [n_tiles][n_patterns + 1] prefix of the kept entries of a tile, per angle
    const int* v_ebase;      // [n_tiles] kept entries before the tile
    const int* zl;           // [nz_pad] film layer of active layer a (-1 = padding)
    const int* a_row_begin;  // [nz + 1] CSR over a_rows: DMD rows that deposit in active layer a
    const int* a_rows;
    const int* row_a;        // [res_y] active layer of DMD row r, -1 = outside the sensor
    int tiles_x, tiles_y, tile_w, tile_h;
    int nz, nz_pad;          // active layers, padded to a multiple of 32
    int lanes_l;             // layers per lane of this sub-plan (1, 2 or 4)
    int a_begin;             // first active layer of this sub-plan
    int groups;
};

#ifdef __CUDACC__
// Records the medium segments of one lane instead of depositing them.
struct RecordSink {
    static constexpr bool kRecordOnly = true;
    Ray seg;
    float tmax = 0.f, w0 = 0.f;
    int count = 0;
    __device__ __forceinline__ void record(const Ray& r, float maxt, float w) {
        if (count == 0) { seg = r; tmax = maxt; w0 = w; }
        ++count;
    }
    __device__ __forceinline__ void visit(size_t, float) {}
};

// first pattern and pattern count of a planar call: the active list holds rows
// [row_begin, row_begin + n_rows) of patterns [k0, k0 + nk), every column
__device__ __forceinline__ void call_angles(const SceneParams& P, const uint32_t* __restrict__ active_pixels,
                                            uint64_t n_active, int& k0, int& nk) {
    k0 = static_cast<int>(__ldg(active_pixels) / (static_cast<uint32_t>(P.res_x) * P.res_y));
    nk = static_cast<int>(n_active / (static_cast<uint64_t>(P.n_rows) * P.res_x));
}

__device__ __forceinline__ float ex2_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

__device__ __forceinline__ float rcp_approx(float x) {
    float y;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d)
                 : "memory");
}

// Clip of a 2-D segment against the voxel rectangle [x0, x0+w) x [y0, y0+h).
template <class Params>
__device__ __forceinline__ bool clip_to_tile(const Params& P, const float4 a, float tmax, int x0, int y0, int w,
                                             int h, float& t0, float& t1, float margin = 0.f) {
    const float vsx = (P.bmax[0] - P.bmin[0]) / static_cast<float>(P.res[0]);
    const float vsy = (P.bmax[1] - P.bmin[1]) / static_cast<float>(P.res[1]);
    const float xl = P.bmin[0] + static_cast<float>(x0) * vsx - margin;
    const float xh = P.bmin[0] + static_cast<float>(x0 + w) * vsx + margin;
    const float yl = P.bmin[1] + static_cast<float>(y0) * vsy - margin;
    const float yh = P.bmin[1] + static_cast<float>(y0 + h) * vsy + margin;
    float tn = 0.f, tf = tmax;
    if (fabsf(a.z) > 1e-8f) {
        const float ta = (xl - a.x) / a.z, tb = (xh - a.x) / a.z;
        tn = fmaxf(tn, fminf(ta, tb));
        tf = fminf(tf, fmaxf(ta, tb));
    } else if (a.x < xl || a.x >= xh) {
        return false;
    }
    if (fabsf(a.w) > 1e-8f) {
        const float ta = (yl - a.y) / a.w, tb = (yh - a.y) / a.w;
        tn = fmaxf(tn, fminf(ta, tb));
        tf = fminf(tf, fmaxf(ta, tb));
    } else if (a.y < yl || a.y >= yh) {
        return false;
    }
    t0 = tn;
    t1 = tf;
    return tn < tf;
}
#endif

}  // namespace drtvam
