// Surface-aware film on top of the planar kernels (C4 of BASELINE.json: square vial with
// refraction and surface-aware discretisation).
//
// The surface-aware film has two channels per voxel, selected by whether the ray is inside the
// target mesh while it crosses the voxel (sensor.py:405-409, volume.py:196,219).  Every crossing of
// the mesh ends a medium segment and re-spawns the ray a little past the surface along the normal
// (null BSDF, volume.py:228-247), so from its first crossing on a lane follows a slightly shifted
// ray: neither the flag nor the geometry fits the (angle, column) tables of the planar kernels.
// But for a deterministic configuration those sub-segments are the same in every pass, and only the
// lanes that meet the target have any.  So:
//
//   set-up   trace every lane once with the general tracer (tracer.cuh, BVH walk of mesh.cuh), with
//            and without the mesh.  A lane that crosses the mesh at t1 keeps records
//            {origin, direction, length, throughput at the origin, start}:
//              - the UNSPLIT segment with negative throughput, walked from t1 on only, and
//              - each sub-segment after the first crossing, tagged inside / outside,
//            stored lane after lane behind an offset table (count -> exclusive scan -> fill);
//   forward  plane A = planar kernels (unsplit rays)  +  DDA over the negative and the outside records
//            plane B = DDA over the inside records            (red.global.add into the planes)
//            film[v][0] += B[v],  film[v][1] += A[v]
//   adjoint  A = g[.][1], B = g[.][0];  grad = planar adjoint on A;
//            grad[lane] += DDA gather over the records, each from its plane
//
// The negative record keeps the origin of the unsplit segment so that its voxel boundaries are those
// of the walk it cancels.  An unsplit segment with a frozen DDA axis (dda.cuh, sensor.py:359) is
// cancelled in full and its head is added back as a record of its own.
// The BVH traversal leaves the per-pass work altogether, and lanes that miss the target cost nothing
// beyond the planar kernels.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>

#include <cub/device/device_scan.cuh>

#include "kernels.h"
#include "planar.cuh"
#include "tracer.cuh"

namespace drtvam {

namespace {

constexpr int kThreads = 256;

#define IN_TRY(expr)                      \
    do {                                  \
        cudaError_t _e = (expr);          \
        if (_e != cudaSuccess) return _e; \
    } while (0)

// Layer tables of the layer-parallel plan, for the classification of lanes (enabled = 0: every lane is odd).
struct TileInfo {
    const int* row_a;   // [res_y] active layer of DMD row r, -1 = none
    const int* zl;      // film layer of active layer a
    int nz_pad;         // active layers padded to 32
    int enabled;
};

constexpr int kMaxTileRecords = 255;

// One thread per lane of the full pixel grid, run twice.
//   kFill = false: cnt_old[lane] / cnt_tile[lane] = number of records of the lane as an odd lane / as a tile lane
//                  (one of the two is 0; both 0: the lane misses the target).
//   kFill = true : the records are written - odd lanes at off[lane] into seg / t_lo, tile lanes into the table
//                  rows of their (angle, column, block) group, and the lane's bit is set in `mask`.
// A lane is a TILE lane when every one of its sub-segments, the head included, has d_z = 0, lies in the voxel layer
// of the lane's DMD row and starts without a frozen DDA axis (sensor.py:359), and its unsplit segment is the plain
// planar one.  P is the full scene (mesh attached), P1 its one-channel view without the mesh.
// flag[0] |= 1 when a lane that meets the mesh has more than one unsplit medium segment (not planar).
template <bool kFill>
__global__ void __launch_bounds__(kThreads)
inside_trace_kernel(const __grid_constant__ SceneParams P, const __grid_constant__ SceneParams P1, uint64_t n_lanes,
                    const TileInfo TI, uint32_t* __restrict__ cnt_old, uint8_t* __restrict__ cnt_tile,
                    const uint64_t* __restrict__ off, float4* __restrict__ seg, float* __restrict__ t_lo,
                    const uint2* __restrict__ grp, float4* __restrict__ table, uint32_t* __restrict__ mask,
                    int* __restrict__ flag) {
    const uint64_t lane = static_cast<uint64_t>(blockIdx.x) * kThreads + threadIdx.x;
    if (lane >= n_lanes) return;
    const uint32_t hw = static_cast<uint32_t>(P.res_x) * P.res_y;
    const uint32_t k = static_cast<uint32_t>(lane / hw), rem = static_cast<uint32_t>(lane - static_cast<uint64_t>(k) * hw);
    const uint32_t row = rem / P.res_x, c = rem - row * P.res_x;
    const int a = TI.enabled ? __ldg(TI.row_a + row) : -1;
    const int z_row = a >= 0 ? __ldg(TI.zl + a) : -1;
    const float vsz = (P.bmax[2] - P.bmin[2]) / static_cast<float>(P.res[2]);

    bool as_tile = false, as_odd = false;
    size_t trow = 0;   // first table row of the lane's group
    uint32_t n_mine = 0, n_rows_g = 0;
    if (kFill) {
        as_tile = cnt_tile[lane] != 0;
        as_odd = cnt_old[lane] != 0;
        if (!as_tile && !as_odd) return;
        if (as_tile) {
            const size_t g = (static_cast<size_t>(k) * P.res_x + c) * (TI.nz_pad >> 5) + (a >> 5);
            const uint2 hd = __ldg(grp + g);
            trow = hd.x;
            n_rows_g = hd.y;
            n_mine = cnt_tile[lane];
            const size_t m = (static_cast<size_t>(k) * P.res_x + c) * TI.nz_pad + a;
            atomicOr(mask + (m >> 5), 1u << (m & 31));
        }
    }
    Sampler rng;
    rng.seed(0u, 0u);  // never consumed: deterministic configuration
    const Ray ray0 = generate_ray(P, static_cast<uint32_t>(lane), rng);
    uint64_t at = kFill && as_odd ? off[lane] : 0;
    uint32_t n_tile = 0;
    auto emit_old = [&](const Ray& r, float maxt, float w0, bool inside, float start) {
        if (kFill && as_odd) {
            seg[2 * at] = make_float4(r.ox, r.oy, r.oz, inside ? -maxt : maxt);  // sign of the length = plane
            seg[2 * at + 1] = make_float4(r.dx, r.dy, r.dz, w0);
            t_lo[at] = start;
            ++at;
        }
    };
    auto emit_tile = [&](const Ray& r, float maxt, float w0, bool inside) {
        if (kFill && as_tile) {
            // the LAST record of a lane (its tail, the longest) goes to the last row of the group whatever the
            // number of crossings before it, so that the tails of the 32 lanes are marched together
            const uint32_t trow_i = n_tile + 1 == n_mine ? n_rows_g - 1 : n_tile;
            const size_t slot = ((trow + trow_i) << 5) + (a & 31);
            table[2 * slot] = make_float4(r.ox, r.oy, r.dx, r.dy);
            table[2 * slot + 1] = make_float4(inside ? -maxt : maxt, w0, 0.f, 0.f);
        }
        ++n_tile;
    };
    // is this sub-segment one the tile kernel can march?
    bool tile_ok = z_row >= 0;
    auto check = [&](const Ray& r, float maxt) {
        const int z = min(max(__float2int_rz((r.oz - P.bmin[2]) / vsz), 0), P.res[2] - 1);
        tile_ok = tile_ok && r.dz == 0.f && z == z_row && r.oz > P.bmin[2] && r.oz < P.bmax[2] && maxt > 0.f &&
                  isfinite(maxt) && !dda_has_frozen_axis(P1, r, maxt);
    };
    // with the mesh: sub-segment 0 is the head of the unsplit segment
    TraceState T;
    T.ray = ray0;
    uint32_t n_sub = 0;
    float t1 = 0.f;
    while (bounce_once<false>(P, T, rng, [&](const Ray& r, float maxt, float w0, bool inside, float) {
        if (n_sub == 0) t1 = maxt;
        else emit_old(r, maxt, w0, inside, 0.f);
        if (!kFill) check(r, maxt);
        emit_tile(r, maxt, w0, inside);
        ++n_sub;
    })) {}
    if (n_sub < 2) {
        if (!kFill) { cnt_old[lane] = 0; cnt_tile[lane] = 0; }
        return;
    }
    // without the mesh: what the planar kernels deposit for this lane
    TraceState U;
    U.ray = ray0;
    uint32_t n_u = 0;
    Ray ur = ray0;
    float u_maxt = 0.f, u_w0 = 0.f;
    while (bounce_once<false>(P1, U, rng, [&](const Ray& r, float maxt, float w0, bool, float) {
        if (n_u == 0) { ur = r; u_maxt = maxt; u_w0 = w0; }
        ++n_u;
    })) {}
    if (n_u != 1) atomicOr(flag, 1);
    const bool whole = dda_has_frozen_axis(P1, ur, u_maxt) || !(t1 > 0.f);
    if (!kFill) {
        const bool tile = tile_ok && !whole && n_u == 1 && n_tile <= kMaxTileRecords;
        cnt_tile[lane] = tile ? static_cast<uint8_t>(n_tile) : 0;
        cnt_old[lane] = tile ? 0u : n_sub + (whole ? 1u : 0u);
        return;
    }
    // odd lane: the planar kernels keep the unsplit segment, its tail is cancelled by a negative record
    if (whole) emit_old(ur, t1, u_w0, false, 0.f);
    emit_old(ur, u_maxt, -u_w0, false, whole ? 0.f : t1);
}

// rows[g] = most records any of the 32 lanes of group g = (angle, column, 32-layer block) holds
__global__ void __launch_bounds__(kThreads)
inside_group_rows_kernel(const __grid_constant__ SceneParams P, const int* __restrict__ a_row_begin,
                         const int* __restrict__ a_rows, int nz, int n_blk, const uint8_t* __restrict__ cnt_tile,
                         uint32_t* __restrict__ rows, size_t n_groups) {
    const size_t g = static_cast<size_t>(blockIdx.x) * kThreads + threadIdx.x;
    if (g >= n_groups) return;
    const int blk = static_cast<int>(g % n_blk);
    const size_t kc = g / n_blk;
    const uint32_t c = static_cast<uint32_t>(kc % P.res_x), k = static_cast<uint32_t>(kc / P.res_x);
    uint32_t m = 0;
    for (int l = 0; l < 32; ++l) {
        const int a = blk * 32 + l;
        if (a >= nz) break;
        const int row = a_rows[a_row_begin[a]];   // one DMD row per layer (checked by the host)
        m = max(m, static_cast<uint32_t>(cnt_tile[(static_cast<size_t>(k) * P.res_y + row) * P.res_x + c]));
    }
    rows[g] = m;
}

__global__ void __launch_bounds__(kThreads)
inside_group_pack_kernel(const uint32_t* __restrict__ rows, const uint32_t* __restrict__ first, uint2* __restrict__ grp,
                         size_t n_groups) {
    const size_t g = static_cast<size_t>(blockIdx.x) * kThreads + threadIdx.x;
    if (g < n_groups) grp[g] = make_uint2(first[g], rows[g]);
}

// rbox[r] = (smin, smax): the stretch of table row r along the centre ray of its (angle, column) - all records of
// a group lie on that line up to the sub-micrometre re-spawn offsets - as s = (p - o_c) . d_c over the end points of
// its 32 records.  One warp per group.  The rows of a group are the consecutive sub-segments of its lanes - head,
// first stretch inside the target, ... - so a row is short, and a tile reads only the rows whose stretch meets
// the stretch of the centre ray inside the tile.
__global__ void __launch_bounds__(kThreads)
inside_row_range_kernel(const float4* __restrict__ seg_a, const uint2* __restrict__ grp, int n_blk,
                        const float4* __restrict__ table, size_t n_groups, float2* __restrict__ rbox,
                        unsigned* __restrict__ stat) {
    const size_t g = (static_cast<size_t>(blockIdx.x) * kThreads + threadIdx.x) >> 5;
    if (g >= n_groups) return;
    const uint2 hd = grp[g];
    if (hd.y == 0) return;
    const int lane = threadIdx.x & 31;
    const float4 c = seg_a[g / n_blk];   // (ox, oy, dx, dy) of the centre ray, unit direction
    float wm = 0.f;
    for (uint32_t i = 0; i < hd.y; ++i) {
        const size_t slot = ((static_cast<size_t>(hd.x) + i) << 5) + lane;
        const float4 ea = table[2 * slot], eb = table[2 * slot + 1];
        if (eb.x != 0.f && isfinite(eb.y)) wm = fmaxf(wm, fabsf(eb.y));
        float s0 = INFINITY, s1 = -INFINITY;
        if (eb.x != 0.f) {
            const float t = fabsf(eb.x);
            const float sa = (ea.x - c.x) * c.z + (ea.y - c.y) * c.w;
            const float sb = (fmaf(ea.z, t, ea.x) - c.x) * c.z + (fmaf(ea.w, t, ea.y) - c.y) * c.w;
            s0 = fminf(sa, sb);
            s1 = fmaxf(sa, sb);
        }
        for (int o = 16; o > 0; o >>= 1) {
            s0 = fminf(s0, __shfl_xor_sync(0xffffffffu, s0, o));
            s1 = fmaxf(s1, __shfl_xor_sync(0xffffffffu, s1, o));
        }
        if (lane == 0) rbox[hd.x + i] = make_float2(s0, s1);
    }
    // stat[0] = largest throughput of the table (bits of a non-negative float order like unsigned integers)
    for (int o = 16; o > 0; o >>= 1) wm = fmaxf(wm, __shfl_xor_sync(0xffffffffu, wm, o));
    if (lane == 0 && __float_as_uint(wm) > *reinterpret_cast<volatile unsigned*>(stat)) atomicMax(stat, __float_as_uint(wm));
}

// stat[1] = largest |pattern value| of a forward call: with stat[0] the scale of the fixed-point tiles
__global__ void __launch_bounds__(kThreads)
inside_absmax_kernel(const float* __restrict__ v, uint64_t n, unsigned* __restrict__ stat) {
    float m = 0.f;
    for (uint64_t i = static_cast<uint64_t>(blockIdx.x) * kThreads + threadIdx.x; i < n; i += static_cast<uint64_t>(gridDim.x) * kThreads) {
        const float a = fabsf(__ldg(v + i));
        if (a < INFINITY) m = fmaxf(m, a);
    }
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0 && __float_as_uint(m) > *reinterpret_cast<volatile unsigned*>(stat + 1))
        atomicMax(stat + 1, __float_as_uint(m));
}

// lanes with cnt[lane] != 0 -> list (positions from an exclusive scan of the flags)
__global__ void __launch_bounds__(kThreads)
inside_flag_kernel(const uint32_t* __restrict__ cnt, uint32_t* __restrict__ flagged, uint64_t n) {
    const uint64_t i = static_cast<uint64_t>(blockIdx.x) * kThreads + threadIdx.x;
    if (i < n) flagged[i] = cnt[i] != 0 ? 1u : 0u;
}

__global__ void __launch_bounds__(kThreads)
inside_list_kernel(const uint32_t* __restrict__ flagged, const uint32_t* __restrict__ pos, uint64_t n,
                   uint32_t* __restrict__ list) {
    const uint64_t i = static_cast<uint64_t>(blockIdx.x) * kThreads + threadIdx.x;
    if (i < n && flagged[i]) list[pos[i]] = static_cast<uint32_t>(i);
}

__global__ void widen_kernel(const uint32_t* __restrict__ in, uint64_t* __restrict__ out, uint64_t n) {
    const uint64_t i = static_cast<uint64_t>(blockIdx.x) * kThreads + threadIdx.x;
    if (i < n) out[i] = in[i];
}

__global__ void narrow_kernel(const uint64_t* __restrict__ in, uint32_t* __restrict__ out, uint64_t n) {
    const uint64_t i = static_cast<uint64_t>(blockIdx.x) * kThreads + threadIdx.x;
    if (i < n) out[i] = static_cast<uint32_t>(in[i]);
}

// P is the one-channel view of the scene (surface_aware = 0): the planes are indexed by voxel.
// One thread per listed lane walks the lane's records one after the other: neighbouring lanes
// (columns of one DMD row) have records of the same kind and nearly the same length, so a warp
// stays together, and the adjoint adds to its pattern pixel without an atomic.
template <int kMode>
__global__ void __launch_bounds__(kThreads)
inside_pass_kernel(const __grid_constant__ SceneParams P, const uint32_t* __restrict__ off,
                   const float4* __restrict__ seg, const float* __restrict__ t_lo, const float* __restrict__ in,
                   float* __restrict__ out, size_t n_vox, const uint32_t* __restrict__ active_pixels,
                   uint64_t n_active, float ray_weight, const int* __restrict__ gate,
                   const uint32_t* __restrict__ odd_lanes, uint32_t n_odd) {
    if (*gate != 0) return;
    // one thread per odd lane of the scene; the ones that are part of this call (the pixel list is base + j,
    // verified by the planar check) do the work
    const uint64_t t = static_cast<uint64_t>(blockIdx.x) * kThreads + threadIdx.x;
    if (t >= n_odd) return;
    const uint64_t lane = __ldg(odd_lanes + t);
    const uint64_t base = __ldg(active_pixels);
    if (lane < base || lane - base >= n_active) return;
    const uint64_t j = lane - base;
    const uint32_t r0 = __ldg(off + lane), r1 = __ldg(off + lane + 1);
    if (r0 == r1) return;
    float em = 0.f;
    if (kMode == 0) {
        em = __ldg(in + j) * ray_weight;
        if (em == 0.f) return;
    }
    float acc = 0.f;
    for (uint32_t i = r0; i < r1; ++i) {
        const float4 a = __ldg(seg + 2 * static_cast<size_t>(i)), b = __ldg(seg + 2 * static_cast<size_t>(i) + 1);
        Ray r;
        r.ox = a.x; r.oy = a.y; r.oz = a.z;
        r.dx = b.x; r.dy = b.y; r.dz = b.z;
        const size_t plane = a.w < 0.f ? n_vox : 0;  // plane B (inside the target) after plane A
        const float start = __ldg(t_lo + i);
        DdaWalk W;
        const bool any = start > 0.f ? W.begin<true>(P, r, fabsf(a.w), b.w, false, start)
                                     : W.begin<false>(P, r, fabsf(a.w), b.w, false);
        if (!any) continue;
        if (kMode == 0) {
            // accurate deposits: the negative record must cancel the planar kernels' deposits, whose
            // coefficients are accurate too (planarz_kernels.cu::walk_entry), to well below 1e-4
            ScatterSink sink{out + plane, em};
            while (W.step<true>(sink)) {}
        } else {
            GatherSink sink{in + plane};
            while (W.step<true>(sink)) {}
            acc += sink.acc;
        }
    }
    if (kMode == 1) out[j] += acc * ray_weight;
}

// ---- tile lanes -------------------------------------------------------------------------------------------
// Built like the jittered planar kernels (planarj_kernels.cu): a CTA owns [tile_h + 2][tile_w + 2][2 planes][32 layers]
// floats of shared memory for a range of angles, a warp takes one (angle, column) group and its 32 lanes are the
// 32 voxel layers of the block, every lane marching ITS OWN records (neighbouring layers meet the target in
// nearly the same places, so the warp stays together).  lane = layer = bank: the red.shared.add of a deposit never
// meets a bank conflict; a tile is flushed once with red.global.add.v4.f32.  The table rows of a group are
// [record][lane], so a warp reads one coalesced kilobyte per record row; groups without records cost one 8-byte
// header.  (A first version kept the records in per-lane lists: 32 scattered 32-byte loads per row, 9 long-
// scoreboard stalls per issue, slower than the per-lane kernel - session r02w.)
constexpr int kTThreads = 1024;
constexpr int kTWarps = kTThreads / 32;
constexpr float kLog2e = 1.4426950408889634f;

struct InsideTilePlan {
    const float4* seg_a;     // centre rays of the (angle, column) pairs (planar.cuh)
    const int* tile_off;     // [n_tiles][n_patterns + 1] prefix of the entries of a tile, per angle
    const int* tile_clo;     // the column list: angle * res_x + column of every entry (planar_tile_lists)
    const int* tile_ebase;   // [n_tiles + 1] first entry of every tile
    const int* zl;           // film layer of active layer a
    const int* a_row_begin;  // CSR of the DMD rows of active layer a (exactly one row per layer here)
    const int* a_rows;
    int tiles_x, tiles_y, tile_w, tile_h;
    int nz, nz_pad, groups;
};

// kMode 0 (forward): in = active_data; planes = plane A, plane B at + n_vox.
// kMode 1 (adjoint): planes = the two gradient planes; out = gT [angle of the call][column][active layer] +=.
// kInt (forward): the tile holds fixed-point sums and a deposit is one native red.shared.add.s32, as in the jittered
// kernels (planarj_kernels.cu); scale from stat[0] x stat[1] (largest throughput of the table x largest |pattern
// value| of the call), stat[2] counts words that came within a factor 2 of overflow.
template <int kMode, bool kInt>
__global__ void __launch_bounds__(kTThreads, 1)
inside_tile_kernel(const __grid_constant__ SceneParams P, const __grid_constant__ InsideTilePlan Z,
                   const uint2* __restrict__ grp, const float4* __restrict__ table, const float2* __restrict__ rbox,
                   const float* __restrict__ in, float* __restrict__ out, float* __restrict__ planes, size_t n_vox,
                   const uint32_t* __restrict__ active_pixels, uint64_t n_active, float ray_weight,
                   const int* __restrict__ gate, unsigned* __restrict__ stat) {
    extern __shared__ __align__(16) float tile[];
    __shared__ int touched;
    if (*gate != 0) return;
    const int tile_id = blockIdx.x, blk = blockIdx.y, group = blockIdx.z;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int x0 = (tile_id % Z.tiles_x) * Z.tile_w, y0 = (tile_id / Z.tiles_x) * Z.tile_h;
    const int w = min(Z.tile_w, P.res[0] - x0), h = min(Z.tile_h, P.res[1] - y0);
    const int pitch = Z.tile_w + 2;                        // voxels per tile row, rim included
    const int tile_floats = (Z.tile_h + 2) * pitch * 64;   // [y][x][plane][layer lane]
    const int n_blk = Z.nz_pad >> 5;
    const int a = blk * 32 + lane;
    const size_t layer_elems = static_cast<size_t>(P.res[0]) * P.res[1];
    const int z_mine = a < Z.nz ? __ldg(Z.zl + a) : -1;
    const int row = a < Z.nz ? __ldg(Z.a_rows + __ldg(Z.a_row_begin + a)) : -1;
    const bool vec = ((x0 | w | P.res[0]) & 3) == 0 && (reinterpret_cast<uintptr_t>(planes) & 15) == 0 && (n_vox & 3) == 0;

    int k0, nk;
    call_angles(P, active_pixels, n_active, k0, nk);
    const int ka = k0 + static_cast<int>(static_cast<long long>(nk) * group / Z.groups);
    const int kb = k0 + static_cast<int>(static_cast<long long>(nk) * (group + 1) / Z.groups);
    const int* __restrict__ off = Z.tile_off + static_cast<size_t>(tile_id) * (P.n_patterns + 1);
    const int* __restrict__ lst = Z.tile_clo + __ldg(Z.tile_ebase + tile_id);
    const int g_begin = __ldg(off + ka), g_end = __ldg(off + kb);

    // ---- does any group of this CTA hold records?  (most tiles lie outside the target or above / below it)
    if (threadIdx.x == 0) touched = 0;
    __syncthreads();
    {
        int k = ka;
        bool any = false;
        for (int g = g_begin + static_cast<int>(threadIdx.x); g < g_end && !any; g += kTThreads) {
            any = __ldg(grp + static_cast<size_t>(__ldg(lst + g)) * n_blk + blk).y != 0;
        }
        if (any) touched = 1;
    }
    __syncthreads();
    if (!touched) return;

    // (the adjoint too: the last step of a walk may land in the rim or in a padding layer, which the staging
    // below does not write - a stale NaN there would survive the multiplication by a zero factor)
    for (int i = threadIdx.x; i < tile_floats; i += kTThreads) tile[i] = 0.f;
    __syncthreads();
    if (kMode == 1 && z_mine >= 0) {
        // stage both gradient planes: lane = layer, a warp takes 4 voxels of a row at a time
        for (int pl = 0; pl < 2; ++pl) {
            const float* src = planes + pl * n_vox + z_mine * layer_elems;
            if (vec) {
                const int w4 = w >> 2;
                for (int i = warp; i < h * w4; i += kTWarps) {
                    const int yy = i / w4, xx = (i - yy * w4) << 2;
                    const float4 v = __ldg(reinterpret_cast<const float4*>(src + static_cast<size_t>(y0 + yy) * P.res[0] + x0 + xx));
                    float* d = tile + (((yy + 1) * pitch + xx + 1) * 2 + pl) * 32 + lane;
                    d[0] = v.x; d[64] = v.y; d[128] = v.z; d[192] = v.w;
                }
            } else {
                for (int i = warp; i < h * w; i += kTWarps) {
                    const int yy = i / w, xx = i - yy * w;
                    tile[(((yy + 1) * pitch + xx + 1) * 2 + pl) * 32 + lane] =
                        __ldg(src + static_cast<size_t>(y0 + yy) * P.res[0] + x0 + xx);
                }
            }
        }
    }
    __syncthreads();

    const float sig = P.sigma_t * kLog2e;
    const float vsx = (P.bmax[0] - P.bmin[0]) / static_cast<float>(P.res[0]);
    const float vsy = (P.bmax[1] - P.bmin[1]) / static_cast<float>(P.res[1]);
    const float rvx = 1.f / vsx, rvy = 1.f / vsy;
    const float xl = P.bmin[0] + static_cast<float>(x0) * vsx, xh = P.bmin[0] + static_cast<float>(x0 + w) * vsx;
    const float yl = P.bmin[1] + static_cast<float>(y0) * vsy, yh = P.bmin[1] + static_cast<float>(y0 + h) * vsy;
    const uint32_t tile_s = static_cast<uint32_t>(__cvta_generic_to_shared(tile));
    const int sxb = 4 * 64, syb = 4 * 64 * pitch;
    // fixed-point scale: 2^25 / (angles of this CTA x 2 records that can share a voxel x largest deposit)
    float fx_scale = 1.f, fx_inv = 1.f;
    if (kInt) {
        const float de_max = 1.f - ex2_approx(-sig * (vsx + vsy));
        const float d_max = __uint_as_float(*reinterpret_cast<volatile unsigned*>(stat)) *
                            __uint_as_float(*reinterpret_cast<volatile unsigned*>(stat + 1)) * fabsf(ray_weight) * de_max;
        fx_scale = d_max > 0.f ? 33554432.f / (2.f * static_cast<float>(max(kb - ka, 1)) * d_max) : 0.f;
        fx_inv = fx_scale > 0.f ? 1.f / fx_scale : 0.f;
    }

    int k = ka;
    // (issuing the header chain of the NEXT group before the current one is marched was measured: 98.5 instead
    // of 96.8 ms per C4 step, session r02zj - the registers it holds cost more than the latency it hides)
    for (int g = g_begin + warp; g < g_end; g += kTWarps) {
        const int kc = __ldg(lst + g);
        k = kc / P.res_x;
        const int c = kc - k * P.res_x;
        const uint2 hd = __ldg(grp + static_cast<size_t>(kc) * n_blk + blk);
        if (hd.y == 0) continue;
        float em = 1.f;
        if (kMode == 0) em = row >= 0 ? __ldg(in + active_index(P, k - k0, row, c)) * ray_weight : 0.f;
        float acc = 0.f;
        // the stretch of the centre ray of (k, c) inside the tile, one voxel of margin
        float T0 = 0.f, T1 = 0.f;
        {
            const float4 cs = __ldg(Z.seg_a + static_cast<size_t>(k) * P.res_x + c);
            if (!clip_to_tile(P, cs, INFINITY, x0, y0, w, h, T0, T1, fmaxf(vsx, vsy))) continue;
        }
        // which rows reach this tile?  One coalesced load of (up to 32) row stretches, one ballot.
        unsigned reach = 0xffffffffu;
        if (hd.y <= 32) {
            bool r = false;
            if (static_cast<unsigned>(lane) < hd.y) {
                const float2 rs = __ldg(rbox + hd.x + lane);
                r = !(rs.x > T1 || rs.y < T0);
            }
            reach = __ballot_sync(0xffffffffu, r);
            if (reach == 0) continue;
        }
        for (unsigned i = 0; i < hd.y; ++i) {
            if (hd.y <= 32) {
                if (!((reach >> i) & 1u)) continue;
            } else {
                const float2 rs = __ldg(rbox + hd.x + i);
                if (rs.x > T1 || rs.y < T0) continue;
            }
            const size_t slot = ((static_cast<size_t>(hd.x) + i) << 5) + lane;
            const float4 eb = __ldg(table + 2 * slot + 1);
            bool live = eb.x != 0.f && em != 0.f;
            if (!__any_sync(0xffffffffu, live)) continue;
            float4 ea = make_float4(0.f, 0.f, 1.f, 0.f);
            if (live) ea = __ldg(table + 2 * slot);
            // ---- clip the record to the tile
            float t0 = 0.f, t1 = fabsf(eb.x);
            const bool mx = fabsf(ea.z) > 1e-8f, my = fabsf(ea.w) > 1e-8f;
            const float rdx = rcp_approx(ea.z), rdy = rcp_approx(ea.w);
            if (mx) {
                const float ta = (xl - ea.x) * rdx, tb = (xh - ea.x) * rdx;
                t0 = fmaxf(t0, fminf(ta, tb));
                t1 = fminf(t1, fmaxf(ta, tb));
            } else if (ea.x < xl || ea.x >= xh) {
                live = false;
            }
            if (my) {
                const float ta = (yl - ea.y) * rdy, tb = (yh - ea.y) * rdy;
                t0 = fmaxf(t0, fminf(ta, tb));
                t1 = fminf(t1, fmaxf(ta, tb));
            } else if (ea.y < yl || ea.y >= yh) {
                live = false;
            }
            live = live && t0 < t1;
            if (!live) continue;
            // entry voxel by truncation then clip (sensor.py:345-346), tile-local
            const float gx = fmaf(ea.z, t0, ea.x), gy = fmaf(ea.w, t0, ea.y);
            const int ix = min(max(__float2int_rz((gx - P.bmin[0]) * rvx) - x0, 0), w - 1);
            const int iy = min(max(__float2int_rz((gy - P.bmin[1]) * rvy) - y0, 0), h - 1);
            // optical depths relative to the entry into the tile (see planarz_kernels.cu)
            const float u1 = sig * (t1 - t0);
            float ux = INFINITY, uy = INFINITY, ax = 0.f, ay = 0.f;
            int sx = 0, sy = 0;
            if (mx) {
                const float inv = sig * rdx;
                const bool pos = ea.z > 0.f;
                sx = pos ? sxb : -sxb;
                ax = fabsf(vsx * inv);
                ux = (fmaf(static_cast<float>(x0 + ix + (pos ? 1 : 0)), vsx, P.bmin[0]) - gx) * inv;
            }
            if (my) {
                const float inv = sig * rdy;
                const bool pos = ea.w > 0.f;
                sy = pos ? syb : -syb;
                ay = fabsf(vsy * inv);
                uy = (fmaf(static_cast<float>(y0 + iy + (pos ? 1 : 0)), vsy, P.bmin[1]) - gy) * inv;
            }
            uint32_t addr = tile_s + 4u * static_cast<uint32_t>((((iy + 1) * pitch + ix + 1) * 2 + (eb.x < 0.f ? 1 : 0)) * 32 + lane);
            const float wgt = eb.y * ex2_approx(-sig * t0) * em * (kInt ? fx_scale : 1.f);
            float e_prev = 1.f, part = 0.f;
            for (;;) {
                const float un = fminf(ux, uy);
                const bool px = ux <= uy;
                const float e_out = ex2_approx(-fminf(un, u1));
                const float de = e_prev - e_out;
                if (kMode == 1) {
                    float v;
                    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
                    part = fmaf(v, de, part);
                } else if (kInt) {
                    asm volatile("red.shared.add.s32 [%0], %1;" ::"r"(addr), "r"(__float2int_rn(wgt * de)) : "memory");
                } else {
                    asm volatile("red.shared.add.f32 [%0], %1;" ::"r"(addr), "f"(wgt * de) : "memory");
                }
                if (!(un < u1)) break;
                e_prev = e_out;
                addr += px ? sx : sy;
                if (px) ux += ax; else uy += ay;
            }
            acc = fmaf(part, wgt, acc);
        }
        if (kMode == 1 && acc != 0.f)
            atomicAdd(out + (static_cast<size_t>(k - k0) * P.res_x + c) * Z.nz_pad + a, acc * ray_weight);
    }
    if (kMode == 1) return;
    __syncthreads();
    if (z_mine < 0) return;

    // ---- flush: lane = layer; every lane adds 4 voxels of a row of its own layer and plane with one red.global.add.v4
    for (int pl = 0; pl < 2; ++pl) {
        float* dst = planes + pl * n_vox + z_mine * layer_elems;
        if (vec) {
            const int w4 = w >> 2;
            for (int i = warp; i < h * w4; i += kTWarps) {
                const int yy = i / w4, xx = (i - yy * w4) << 2;
                const float* src = tile + (((yy + 1) * pitch + xx + 1) * 2 + pl) * 32 + lane;
                float v0 = src[0], v1 = src[64], v2 = src[128], v3 = src[192];
                if (kInt) {
                    const int i0 = __float_as_int(v0), i1 = __float_as_int(v1), i2 = __float_as_int(v2), i3 = __float_as_int(v3);
                    if (max(max(abs(i0), abs(i1)), max(abs(i2), abs(i3))) > (1 << 30)) atomicAdd(stat + 2, 1u);
                    v0 = static_cast<float>(i0) * fx_inv; v1 = static_cast<float>(i1) * fx_inv;
                    v2 = static_cast<float>(i2) * fx_inv; v3 = static_cast<float>(i3) * fx_inv;
                }
                if (v0 != 0.f || v1 != 0.f || v2 != 0.f || v3 != 0.f)
                    red_add_v4(dst + static_cast<size_t>(y0 + yy) * P.res[0] + (x0 + xx), v0, v1, v2, v3);
            }
        } else {
            for (int i = warp; i < h * w; i += kTWarps) {
                const int yy = i / w, xx = i - yy * w;
                float v = tile[(((yy + 1) * pitch + xx + 1) * 2 + pl) * 32 + lane];
                if (kInt) {
                    const int iv = __float_as_int(v);
                    if (abs(iv) > (1 << 30)) atomicAdd(stat + 2, 1u);
                    v = static_cast<float>(iv) * fx_inv;
                }
                if (v != 0.f) atomicAdd(dst + static_cast<size_t>(y0 + yy) * P.res[0] + (x0 + xx), v);
            }
        }
    }
}

__global__ void __launch_bounds__(kThreads)
inside_combine_kernel(const float* __restrict__ outside, const float* __restrict__ inside, float2* __restrict__ film,
                      size_t n_vox, const int* __restrict__ gate) {
    if (*gate != 0) return;
    const size_t v = static_cast<size_t>(blockIdx.x) * kThreads + threadIdx.x;
    if (v >= n_vox) return;
    float2 f = film[v];
    f.x += inside[v];
    f.y += outside[v];
    film[v] = f;
}

__global__ void __launch_bounds__(kThreads)
inside_split_kernel(const float2* __restrict__ g, float* __restrict__ g_out, float* __restrict__ g_in, size_t n_vox) {
    const size_t v = static_cast<size_t>(blockIdx.x) * kThreads + threadIdx.x;
    if (v >= n_vox) return;
    const float2 f = g[v];
    g_out[v] = f.y;
    g_in[v] = f.x;
}

unsigned blocks_for(uint64_t n) { return static_cast<unsigned>((n + kThreads - 1) / kThreads); }

}  // namespace

void inside_release(InsideHost& H) {
    void* ptrs[] = {H.off, H.seg, H.t_lo, H.odd_lanes, H.planes, H.grp, H.table, H.rbox, H.stat, H.mask, H.tile_off, H.tile_clo, H.tile_ebase};
    for (void* q : ptrs)
        if (q) cudaFree(q);
    H = InsideHost{};
}

// P: the full scene (target mesh attached, surface_aware = 1).  One-off set-up; synchronises.
cudaError_t inside_build(const SceneParams& P, const SceneParams& P1, const PlanarHost& base, const PlanarZHost& Zh,
                         bool allow_tiles, InsideHost& H) {
    inside_release(H);
    const uint64_t n_lanes = static_cast<uint64_t>(P.n_patterns) * P.res_x * P.res_y;
    const size_t n_vox = static_cast<size_t>(P.res[0]) * P.res[1] * P.res[2];
    if (n_lanes >= (1ull << 32)) return cudaErrorNotSupported;
    const char* env = std::getenv("DRTVAM_INSIDE_TILES");
    const bool tiles = allow_tiles && Zh.valid && Zh.rows_max == 1 && Zh.nz > 0 && !(env && std::atoi(env) == 0);
    TileInfo TI{Zh.row_a, Zh.zl, Zh.nz_pad, tiles ? 1 : 0};
    const int n_blk = tiles ? Zh.nz_pad / 32 : 0;
    const size_t n_groups = static_cast<size_t>(P.n_patterns) * P.res_x * n_blk;
    const size_t mask_words = tiles ? (static_cast<size_t>(P.n_patterns) * P.res_x * Zh.nz_pad) >> 5 : 0;

    uint32_t *cnt = nullptr, *flagged = nullptr, *pos = nullptr, *rows = nullptr, *first = nullptr;
    uint8_t* cnt_tile = nullptr;
    int* flag = nullptr;
    uint64_t *wide = nullptr, *off = nullptr;
    void* tmp = nullptr;
    size_t tmp_bytes = 0;
    auto cleanup = [&]() {
        void* ptrs[] = {cnt, flagged, pos, rows, first, cnt_tile, flag, wide, off, tmp};
        for (void* q : ptrs)
            if (q) cudaFree(q);
    };
    auto scan32 = [&](const uint32_t* in, uint32_t* out, size_t n) -> cudaError_t {
        size_t need = 0;
        IN_TRY(cub::DeviceScan::ExclusiveSum(nullptr, need, in, out, n));
        if (need > tmp_bytes) {
            if (tmp) cudaFree(tmp);
            tmp = nullptr;
            IN_TRY(cudaMalloc(&tmp, need));
            tmp_bytes = need;
        }
        IN_TRY(cub::DeviceScan::ExclusiveSum(tmp, need, in, out, n));
        count_launch();
        return cudaSuccess;
    };
    auto run = [&]() -> cudaError_t {
        IN_TRY(cudaMalloc(&cnt, sizeof(uint32_t) * (n_lanes + 1)));
        IN_TRY(cudaMalloc(&cnt_tile, n_lanes));
        IN_TRY(cudaMalloc(&wide, sizeof(uint64_t) * (n_lanes + 1)));
        IN_TRY(cudaMalloc(&off, sizeof(uint64_t) * (n_lanes + 1)));
        IN_TRY(cudaMemset(cnt + n_lanes, 0, sizeof(uint32_t)));
        IN_TRY(cudaMemset(wide + n_lanes, 0, sizeof(uint64_t)));
        IN_TRY(cudaMalloc(&flag, sizeof(int)));
        IN_TRY(cudaMemset(flag, 0, sizeof(int)));
        inside_trace_kernel<false><<<blocks_for(n_lanes), kThreads>>>(P, P1, n_lanes, TI, cnt, cnt_tile, nullptr, nullptr,
                                                                     nullptr, nullptr, nullptr, nullptr, flag);
        count_launch();
        int bad = 0;
        IN_TRY(cudaMemcpy(&bad, flag, sizeof(bad), cudaMemcpyDeviceToHost));
        if (bad) return cudaErrorNotSupported;
        // ---- odd lanes: offsets of their record lists, and the list of the lanes themselves
        widen_kernel<<<blocks_for(n_lanes), kThreads>>>(cnt, wide, n_lanes);
        count_launch();
        size_t need = 0;
        IN_TRY(cub::DeviceScan::ExclusiveSum(nullptr, need, wide, off, n_lanes + 1));
        IN_TRY(cudaMalloc(&tmp, need));
        tmp_bytes = need;
        IN_TRY(cub::DeviceScan::ExclusiveSum(tmp, need, wide, off, n_lanes + 1));
        count_launch();
        uint64_t total = 0;
        IN_TRY(cudaMemcpy(&total, off + n_lanes, sizeof(total), cudaMemcpyDeviceToHost));
        if (total >= (1ull << 32)) return cudaErrorMemoryAllocation;
        H.n_rec = total;
        IN_TRY(cudaMalloc(&flagged, sizeof(uint32_t) * (n_lanes + 1)));
        IN_TRY(cudaMalloc(&pos, sizeof(uint32_t) * (n_lanes + 1)));
        IN_TRY(cudaMemset(flagged + n_lanes, 0, sizeof(uint32_t)));
        inside_flag_kernel<<<blocks_for(n_lanes), kThreads>>>(cnt, flagged, n_lanes);
        count_launch();
        IN_TRY(scan32(flagged, pos, n_lanes + 1));
        IN_TRY(cudaMemcpy(&H.n_odd, pos + n_lanes, sizeof(uint32_t), cudaMemcpyDeviceToHost));
        IN_TRY(cudaMalloc(&H.odd_lanes, sizeof(uint32_t) * std::max<uint32_t>(H.n_odd, 1u)));
        if (H.n_odd) {
            inside_list_kernel<<<blocks_for(n_lanes), kThreads>>>(flagged, pos, n_lanes, H.odd_lanes);
            count_launch();
        }
        // ---- tile lanes: rows per group, first row of every group
        if (tiles) {
            IN_TRY(cudaMalloc(&rows, sizeof(uint32_t) * (n_groups + 1)));
            IN_TRY(cudaMalloc(&first, sizeof(uint32_t) * (n_groups + 1)));
            IN_TRY(cudaMemset(rows + n_groups, 0, sizeof(uint32_t)));
            inside_group_rows_kernel<<<blocks_for(n_groups), kThreads>>>(P, Zh.a_row_begin, Zh.a_rows, Zh.nz, n_blk, cnt_tile,
                                                                       rows, n_groups);
            count_launch();
            IN_TRY(scan32(rows, first, n_groups + 1));
            uint32_t n_rows = 0;
            IN_TRY(cudaMemcpy(&n_rows, first + n_groups, sizeof(uint32_t), cudaMemcpyDeviceToHost));
            H.n_rows = n_rows;
        }
        size_t free_b = 0, all_b = 0;
        IN_TRY(cudaMemGetInfo(&free_b, &all_b));
        if (total * 40 + H.n_rows * 32 * 32 > free_b / 2) return cudaErrorMemoryAllocation;
        if (total) {
            IN_TRY(cudaMalloc(&H.seg, sizeof(float4) * 2 * total));
            IN_TRY(cudaMalloc(&H.t_lo, sizeof(float) * total));
        }
        if (tiles) {
            IN_TRY(cudaMalloc(&H.grp, sizeof(uint2) * n_groups));
            inside_group_pack_kernel<<<blocks_for(n_groups), kThreads>>>(rows, first, H.grp, n_groups);
            count_launch();
            IN_TRY(cudaMalloc(&H.table, sizeof(float4) * 2 * 32 * std::max<uint64_t>(H.n_rows, 1)));
            IN_TRY(cudaMemset(H.table, 0, sizeof(float4) * 2 * 32 * std::max<uint64_t>(H.n_rows, 1)));
            IN_TRY(cudaMalloc(&H.mask, sizeof(uint32_t) * std::max<size_t>(mask_words, 1)));
            IN_TRY(cudaMemset(H.mask, 0, sizeof(uint32_t) * std::max<size_t>(mask_words, 1)));
        }
        if (total || H.n_rows) {
            inside_trace_kernel<true><<<blocks_for(n_lanes), kThreads>>>(P, P1, n_lanes, TI, cnt, cnt_tile, off, H.seg, H.t_lo,
                                                                        H.grp, H.table, H.mask, flag);
            count_launch();
        }
        if (tiles) {
            IN_TRY(cudaMalloc(&H.rbox, sizeof(float2) * std::max<uint64_t>(H.n_rows, 1)));
            IN_TRY(cudaMalloc(&H.stat, 4 * sizeof(unsigned)));
            IN_TRY(cudaMemset(H.stat, 0, 4 * sizeof(unsigned)));
            if (H.n_rows) {
                inside_row_range_kernel<<<blocks_for(n_groups * 32), kThreads>>>(base.seg_a, H.grp, n_blk, H.table, n_groups,
                                                                               H.rbox, H.stat);
                count_launch();
            }
        }
        IN_TRY(cudaMalloc(&H.off, sizeof(uint32_t) * (n_lanes + 1)));
        narrow_kernel<<<blocks_for(n_lanes + 1), kThreads>>>(off, H.off, n_lanes + 1);
        count_launch();
        IN_TRY(cudaMalloc(&H.planes, sizeof(float) * 2 * n_vox));
        if (tiles && H.n_rows) {
            // the largest xy tile whose [h + 2][w + 2][2 planes][32 layers] float slab fits 227 KB
            const int np = P1.n_patterns, rx = P1.res[0], ry = P1.res[1];
            const int area = (227 * 1024 / 4) / 64;
            const int side = static_cast<int>(std::floor(std::sqrt(static_cast<double>(area)))) - 2;
            const int tw_max = std::max(4, side & ~3);
            H.tiles_x = (rx + tw_max - 1) / tw_max;
            H.tile_w = (((rx + H.tiles_x - 1) / H.tiles_x) + 3) & ~3;
            H.tiles_x = (rx + H.tile_w - 1) / H.tile_w;
            const int th_max = std::max(1, area / (H.tile_w + 2) - 2);
            H.tiles_y = (ry + th_max - 1) / th_max;
            H.tile_h = (ry + H.tiles_y - 1) / H.tiles_y;
            H.smem_bytes = sizeof(float) * static_cast<size_t>(H.tile_h + 2) * (H.tile_w + 2) * 64;
            const int n_tiles = H.tiles_x * H.tiles_y;
            IN_TRY((cudaFuncSetAttribute(inside_tile_kernel<0, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(H.smem_bytes))));
            IN_TRY((cudaFuncSetAttribute(inside_tile_kernel<0, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(H.smem_bytes))));
            IN_TRY((cudaFuncSetAttribute(inside_tile_kernel<1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(H.smem_bytes))));
            IN_TRY(cudaMalloc(&H.tile_off, sizeof(int) * static_cast<size_t>(n_tiles) * (np + 1)));
            IN_TRY(cudaMalloc(&H.tile_ebase, sizeof(int) * (n_tiles + 1)));
            // a record follows the centre ray of its (angle, column) up to the re-spawn offsets of its crossings
            // (under a micrometre each): one pixel of margin
            const float margin = std::fmax(P1.pixel_x, P1.pixel_y);
            // column lists, not ranges: behind the refracting box the range of a (tile, angle) spans the gap between
            // the two ray families that cross the tile (60 % of the groups a CTA looked at missed its tile, r03m)
            long long n_entries = 0;
            IN_TRY(planar_tile_lists(P1, base.seg_a, base.seg_b, H.tiles_x, H.tiles_y, H.tile_w, H.tile_h, H.tile_off,
                                     H.tile_ebase, &H.tile_clo, &n_entries, margin));
            H.seg_a = base.seg_a;
            H.tiles = true;
        }
        IN_TRY(cudaDeviceSynchronize());
        return cudaGetLastError();
    };
    const cudaError_t e = run();
    cleanup();
    if (e != cudaSuccess) {
        inside_release(H);
        return e;
    }
    H.n_vox = n_vox;
    H.valid = true;
    return cudaSuccess;
}

// Forward, part 1: clears the two planes; the caller then runs the planar forward kernels into plane 0.
cudaError_t inside_forward_begin(InsideHost& H, cudaStream_t stream) {
    return cudaMemsetAsync(H.planes, 0, sizeof(float) * 2 * H.n_vox, stream);
}

static InsideTilePlan tile_plan(const InsideHost& H, const PlanarZHost& Zh) {
    InsideTilePlan Z{};
    Z.seg_a = H.seg_a;
    Z.tile_off = H.tile_off; Z.tile_clo = H.tile_clo; Z.tile_ebase = H.tile_ebase;
    Z.zl = Zh.zl; Z.a_row_begin = Zh.a_row_begin; Z.a_rows = Zh.a_rows;
    Z.tiles_x = H.tiles_x; Z.tiles_y = H.tiles_y; Z.tile_w = H.tile_w; Z.tile_h = H.tile_h;
    Z.nz = Zh.nz; Z.nz_pad = Zh.nz_pad;
    // the tiles over the target hold all the work: several CTAs (angle groups) each
    static const int groups = std::getenv("DRTVAM_INSIDE_GROUPS") ? std::max(1, std::atoi(std::getenv("DRTVAM_INSIDE_GROUPS"))) : 4;
    Z.groups = groups;
    return Z;
}

// Forward, part 2: the records into the planes, then both planes into the two-channel film.
cudaError_t inside_forward_finish(const SceneParams& P1, InsideHost& H, const PlanarZHost& Zh, const int* gate,
                                  const float* active_data, const uint32_t* active_pixels, uint64_t n_active,
                                  float ray_weight, float* film, cudaStream_t stream) {
    if (H.tiles) {
        const InsideTilePlan Z = tile_plan(H, Zh);
        const dim3 grid(H.tiles_x * H.tiles_y, Zh.nz_pad / 32, Z.groups);
        static const bool fixed_point = !(std::getenv("DRTVAM_JITTER_ATOMICS") &&
                                          std::strcmp(std::getenv("DRTVAM_JITTER_ATOMICS"), "float") == 0);
        if (fixed_point) {
            IN_TRY(cudaMemsetAsync(H.stat + 1, 0, sizeof(unsigned), stream));
            inside_absmax_kernel<<<148 * 8, kThreads, 0, stream>>>(active_data, n_active, H.stat);
            count_launch();
            inside_tile_kernel<0, true><<<grid, kTThreads, H.smem_bytes, stream>>>(P1, Z, H.grp, H.table, H.rbox, active_data,
                                                                                  nullptr, H.planes, H.n_vox, active_pixels,
                                                                                  n_active, ray_weight, gate, H.stat);
        } else {
            inside_tile_kernel<0, false><<<grid, kTThreads, H.smem_bytes, stream>>>(P1, Z, H.grp, H.table, H.rbox, active_data,
                                                                                   nullptr, H.planes, H.n_vox, active_pixels,
                                                                                   n_active, ray_weight, gate, H.stat);
        }
        count_launch();
    }
    if (H.n_odd) {
        inside_pass_kernel<0><<<blocks_for(H.n_odd), kThreads, 0, stream>>>(P1, H.off, H.seg, H.t_lo, active_data, H.planes,
                                                                          H.n_vox, active_pixels, n_active, ray_weight, gate,
                                                                          H.odd_lanes, H.n_odd);
        count_launch();
    }
    inside_combine_kernel<<<blocks_for(H.n_vox), kThreads, 0, stream>>>(H.planes, H.planes + H.n_vox,
                                                                       reinterpret_cast<float2*>(film), H.n_vox, gate);
    count_launch();
    return cudaGetLastError();
}

// Adjoint, part 1: plane A = outside-channel gradient (input of the planar adjoint), plane B = inside channel.
cudaError_t inside_adjoint_begin(InsideHost& H, const float* dL, cudaStream_t stream) {
    inside_split_kernel<<<blocks_for(H.n_vox), kThreads, 0, stream>>>(reinterpret_cast<const float2*>(dL), H.planes,
                                                                     H.planes + H.n_vox, H.n_vox);
    count_launch();
    return cudaGetLastError();
}

// Adjoint of the tile lanes: between the gather of the layer-parallel adjoint (whose values for these lanes have
// just been cleared) and its unpack.
cudaError_t inside_adjoint_tiles(const SceneParams& P1, InsideHost& H, const PlanarZHost& Zh, const int* gate,
                                 const uint32_t* active_pixels, uint64_t n_active, float ray_weight, float* gT,
                                 cudaStream_t stream) {
    if (!H.tiles) return cudaSuccess;
    const InsideTilePlan Z = tile_plan(H, Zh);
    const dim3 grid(H.tiles_x * H.tiles_y, Zh.nz_pad / 32, Z.groups);
    inside_tile_kernel<1, false><<<grid, kTThreads, H.smem_bytes, stream>>>(P1, Z, H.grp, H.table, H.rbox, nullptr, gT, H.planes,
                                                                           H.n_vox, active_pixels, n_active, ray_weight, gate,
                                                                           H.stat);
    count_launch();
    return cudaGetLastError();
}

// Adjoint, part 2 (after the planar adjoint has written grad): adds the gather over the records of the odd lanes.
cudaError_t inside_adjoint_finish(const SceneParams& P1, InsideHost& H, const int* gate, const uint32_t* active_pixels,
                                  uint64_t n_active, float ray_weight, float* grad, cudaStream_t stream) {
    if (!H.n_odd) return cudaSuccess;
    inside_pass_kernel<1><<<blocks_for(H.n_odd), kThreads, 0, stream>>>(P1, H.off, H.seg, H.t_lo, H.planes, grad, H.n_vox,
                                                                      active_pixels, n_active, ray_weight, gate,
                                                                      H.odd_lanes, H.n_odd);
    count_launch();
    return cudaGetLastError();
}

}  // namespace drtvam
