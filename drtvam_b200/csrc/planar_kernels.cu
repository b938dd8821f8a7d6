// Planar (z-separable) fast path of the forward projection and its adjoint.
//
// Forward: CTA = (tile of one z-layer, angle group).  The tile lives in shared
// memory; every thread takes one (angle, column) ray whose medium segment
// crosses the tile, clips it to the tile and marches a 2-D voxel DDA with
// shared-memory atomics (tile privatisation); the tile is flushed once with
// vectorised red.global.add.v4.f32.  Adjoint: same walk in gather form over a
// staged tile of the dose gradient, one red.global per ray and tile.
// Reference: DDAVolumetricSensor.accumulate (sensor.py:306-440), film.write
// (film.py:40-41), VolumeIntegrator.render / render_backward
// (integrators/volume.py:18-56, 97-134).
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "kernels.h"
#include "planar.cuh"
#include "tracer.cuh"

namespace drtvam {

namespace {

constexpr int kPlanarThreads = 256;
constexpr float kLog2e = 1.4426950408889634f;

// Segment table: one thread per (angle, column), traced at row `row_ref`
// through the general interface loop so both plans share the same arithmetic.
__global__ void __launch_bounds__(kPlanarThreads)
planar_segments_kernel(const __grid_constant__ SceneParams P, int row_ref, float4* __restrict__ seg_a,
                       float2* __restrict__ seg_b, SegmentD* __restrict__ seg_d, int* __restrict__ flags,
                       int* __restrict__ frozen_list, int* __restrict__ frozen_count) {
    const int i = blockIdx.x * kPlanarThreads + threadIdx.x;
    const int n = P.n_patterns * P.res_x;
    if (i >= n) return;
    const int k = i / P.res_x, c = i - k * P.res_x;
    const uint32_t pixel = (static_cast<uint32_t>(k) * P.res_y + row_ref) * P.res_x + c;
    Sampler rng;
    rng.seed(0u, static_cast<uint32_t>(i));  // never consumed for a deterministic configuration
    const Ray ray = generate_ray(P, pixel, rng);
    RecordSink sink;
    sink.seg = ray;
    trace_lane<false>(P, ray, rng, sink);
    int f = 0;
    if (sink.count > 1) f |= kPlanarFlagMultiSegment;
    if (sink.count >= 1 && sink.seg.dz != 0.f) f |= kPlanarFlagNotPlanar;
    if (ray.dz != 0.f) f |= kPlanarFlagNotPlanar;
    if (f) atomicOr(flags, f);
    bool ok = sink.count >= 1 && sink.w0 != 0.f && sink.tmax > 0.f && isfinite(sink.tmax);
    if (ok && dda_has_frozen_axis(P, sink.seg, sink.tmax)) {
        // sensor.py:359 quirk: this (angle, column) is walked by dda_segment in planar_frozen_kernel
        frozen_list[atomicAdd(frozen_count, 1)] = i;
        ok = false;
    }
    seg_a[i] = make_float4(sink.seg.ox, sink.seg.oy, sink.seg.dx, sink.seg.dy);
    seg_b[i] = ok ? make_float2(sink.tmax, sink.w0) : make_float2(0.f, 0.f);
    // the same segment in double precision for the set-up of the walk tables (planar_f64.cuh); the fp32
    // tracer above decides, this one refines, and a walk that disagrees is not trusted
    SegmentD D{sink.seg.ox, sink.seg.oy, sink.seg.dx, sink.seg.dy, ok ? sink.tmax : 0.f, ok ? sink.w0 : 0.f};
    if (ok && P.motion_table_f64 && P.proj_kind == 0 && !P.sample_time) {
        SegmentD R;
        if (f64::trace_segment(P, P.motion_table_f64, k, row_ref, c, R) &&
            fabs(R.tmax - D.tmax) < 1e-3 * (1.0 + D.tmax) && fabs(R.ox - D.ox) < 1e-3 && fabs(R.oy - D.oy) < 1e-3 &&
            fabs(R.w0 - D.w0) < 1e-3 * fabs(D.w0))
            D = R;
    }
    seg_d[i] = D;
}

// Layer each DMD row deposits in (the z part of sensor.py:328-346), -1 = none.
__global__ void planar_row_layer_kernel(const __grid_constant__ SceneParams P, int* __restrict__ row_layer,
                                        float* __restrict__ row_z) {
    const int row = blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= P.res_y) return;
    Sampler rng;
    rng.seed(0u, 0u);
    const Ray ray = generate_ray(P, static_cast<uint32_t>(row) * P.res_x, rng);
    int layer = -1;
    if (ray.oz > P.bmin[2] && ray.oz < P.bmax[2]) {
        const float vs = (P.bmax[2] - P.bmin[2]) / static_cast<float>(P.res[2]);
        layer = min(max(__float2int_rz((ray.oz - P.bmin[2]) / vs), 0), P.res[2] - 1);
    }
    row_layer[row] = layer;
    row_z[row] = ray.oz;
}

// Column range of every (tile, angle): one thread each, scanning the columns.
__global__ void planar_ranges_kernel(const __grid_constant__ SceneParams P, const PlanarPlan L, int* __restrict__ clo,
                                     int* __restrict__ cnt, float margin) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int n_tiles = L.tiles_x * L.tiles_y;
    if (i >= n_tiles * P.n_patterns) return;
    const int tile = i / P.n_patterns, k = i - tile * P.n_patterns;
    const int x0 = (tile % L.tiles_x) * L.tile_w, y0 = (tile / L.tiles_x) * L.tile_h;
    const int w = min(L.tile_w, P.res[0] - x0), h = min(L.tile_h, P.res[1] - y0);
    int lo = P.res_x, hi = -1;
    for (int c = 0; c < P.res_x; ++c) {
        const float2 b = L.seg_b[k * P.res_x + c];
        if (b.y == 0.f) continue;
        float t0, t1;
        if (clip_to_tile(P, L.seg_a[k * P.res_x + c], b.x, x0, y0, w, h, t0, t1, margin)) {
            lo = min(lo, c);
            hi = max(hi, c);
        }
    }
    if (margin > 0.f && hi >= lo) {  // jittered rays of the neighbouring columns may cross too
        lo = max(lo - 1, 0);
        hi = min(hi + 1, P.res_x - 1);
    }
    clo[i] = hi >= lo ? lo : 0;
    cnt[i] = hi >= lo ? hi - lo + 1 : 0;
}

// The same question per column instead of per range: kFill = false counts, kFill = true writes angle * res_x + column
// of every column of angle k whose centre ray (or, with a margin, one of its two neighbours') crosses the tile.
// One column range per (tile, angle) is enough for a single family of rays; behind a refracting box two families
// cross a tile and the range spans everything between them (most of it empty).
template <bool kFill>
__global__ void planar_lists_kernel(const __grid_constant__ SceneParams P, const PlanarPlan L, const int* __restrict__ off,
                                    const int* __restrict__ ebase, int* __restrict__ cnt, int* __restrict__ list,
                                    float margin) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int n_tiles = L.tiles_x * L.tiles_y;
    if (i >= n_tiles * P.n_patterns) return;
    const int tile = i / P.n_patterns, k = i - tile * P.n_patterns;
    const int x0 = (tile % L.tiles_x) * L.tile_w, y0 = (tile / L.tiles_x) * L.tile_h;
    const int w = min(L.tile_w, P.res[0] - x0), h = min(L.tile_h, P.res[1] - y0);
    int n = 0;
    int* dst = kFill ? list + ebase[tile] + off[tile * (P.n_patterns + 1) + k] : nullptr;
    bool prev = false, cur = false;   // column c - 1, column c
    auto crosses = [&](int c) {
        if (c >= P.res_x) return false;
        const float2 b = L.seg_b[k * P.res_x + c];
        float t0, t1;
        return b.y != 0.f && clip_to_tile(P, L.seg_a[k * P.res_x + c], b.x, x0, y0, w, h, t0, t1, margin);
    };
    cur = crosses(0);
    for (int c = 0; c < P.res_x; ++c) {
        const bool next = crosses(c + 1);
        if (cur || (margin > 0.f && (prev || next))) {
            if (kFill) dst[n] = k * P.res_x + c;
            ++n;
        }
        prev = cur;
        cur = next;
    }
    if (!kFill) cnt[i] = n;
}

// gate[0] != 0  <=>  active_pixels is not the dense layout the planar kernels index implicitly - DMD rows
// [row_begin, row_begin + n_rows) of consecutive whole patterns, every column, in order - i.e. the planar
// kernels must stand down and the general tracer runs instead.
__global__ void planar_check_pixels_kernel(const uint32_t* __restrict__ active_pixels, uint64_t n_active, uint32_t hw,
                                           uint32_t rw, uint32_t first, int* __restrict__ gate) {
    const uint64_t j = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    const uint32_t base = __ldg(active_pixels);
    bool bad = false;
    if (j == 0) bad = (base % hw) != first;
    if (j < n_active) {
        const uint64_t kk = j / rw;
        bad = bad || (__ldg(active_pixels + j) != base + static_cast<uint32_t>(kk * hw + (j - kk * rw)));
    }
    if (__any_sync(0xffffffffu, bad) && (threadIdx.x & 31) == 0) atomicOr(gate, 1);
}


// Rays of the deterministic table whose walk has a frozen axis (sensor.py:359):
// one visit of a gather walk: the weight and the (not yet consumed) gathered value
struct FrozenVisit {
    static constexpr bool kRecordOnly = false;
    const float* dL;
    float w = 0.f, v = 0.f;
    __device__ __forceinline__ void visit(size_t idx, float wt) {
        w = wt;
        v = __ldg(dL + idx);
    }
};

// one thread per (listed (angle, column), DMD row), general walk.
// kMode 0: in = active_data, out = film.  kMode 1: in = dL, out = pattern_grad (written, not added).
template <int kMode>
__global__ void __launch_bounds__(kPlanarThreads)
planar_frozen_kernel(const __grid_constant__ SceneParams P, const int* __restrict__ frozen_list, int n_frozen,
                     const float* __restrict__ in, float* __restrict__ out, const uint32_t* __restrict__ active_pixels,
                     uint64_t n_active, float ray_weight, const int* __restrict__ gate) {
    if (*gate != 0) return;
    const int i = blockIdx.x * kPlanarThreads + threadIdx.x;
    if (i >= n_frozen * P.n_rows) return;
    const int e = __ldg(frozen_list + i / P.n_rows), row = P.row_begin + i % P.n_rows;
    const int k = e / P.res_x, c = e - k * P.res_x;
    int k0, nk;
    call_angles(P, active_pixels, n_active, k0, nk);
    if (k < k0 || k >= k0 + nk) return;
    const size_t j = active_index(P, k - k0, row, c);
    Sampler rng;
    rng.seed(0u, 0u);  // never consumed: deterministic configuration
    const Ray ray = generate_ray(P, __ldg(active_pixels + j), rng);
    if (kMode == 1) {
        // one medium segment (the scene is planar): record it, then walk it with 8 gathers in flight - walked with
        // one dependent load per voxel this kernel took 0.31 ms for the 2 400 lanes of C2 (profiles/r04c_*)
        RecordSink rec;
        rec.seg = ray;
        trace_lane<false>(P, ray, rng, rec);
        float acc = 0.f;
        DdaWalk W;
        if (rec.count > 1) {   // not a planar lane after all: the plain walk
            Sampler rng2;
            rng2.seed(0u, 0u);
            GatherSink g{in};
            trace_lane<false>(P, ray, rng2, g);
            acc = g.acc;
        } else if (rec.count == 1 && W.begin(P, rec.seg, rec.tmax, rec.w0, false)) {
            bool walking = true;
            while (walking) {
                float w[8], v[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    FrozenVisit one{in};
                    if (walking) walking = W.step(one);
                    w[i] = one.w;
                    v[i] = one.v;
                }
#pragma unroll
                for (int i = 0; i < 8; ++i) acc = fmaf(w[i], v[i], acc);
            }
        }
        out[j] = acc * ray_weight;
    } else {
        const float em = __ldg(in + j) * ray_weight;
        if (em == 0.f) return;
        ScatterSink sc{out, em};
        trace_lane<false>(P, ray, rng, sc);
    }
}

}  // namespace

namespace {

// Kernel modes.  The forward tile accumulates with fp32 atomicAdd, which is an
// LDS + ATOMS.CAST.SPIN loop on shared memory.  An int32 fixed-point tile with
// native ATOMS.ADD.S32 was measured 2.4x slower on B200 (76.9 ms against
// 31.5 ms per forward pass at C2, profiles/r01_notes.md) and was dropped.
constexpr int kScatter = 0, kGather = 1;

// 2-D DDA of one clipped segment inside a tile.  Times are carried as
// u = sigma_t * log2(e) * t so that the transmittance is ex2(-u); the deposit
// of a voxel is the telescoped difference  E(u_in) - E(u_out)  which equals
// e^{-sigma_t t} (1 - e^{-sigma_t dt})  of sensor.py:400-404.
// Scatter modes add em * dE to tile[v]; the gather mode returns sum tile[v] * dE.
template <int kMode>
__device__ __forceinline__ float march_tile(const SceneParams& P, float* __restrict__ tile, int pitch, const float4 a,
                                            float t0, float t1, int x0, int y0, int w, int h, float em) {
    const float sig = P.sigma_t * kLog2e;
    const float vsx = (P.bmax[0] - P.bmin[0]) / static_cast<float>(P.res[0]);
    const float vsy = (P.bmax[1] - P.bmin[1]) / static_cast<float>(P.res[1]);
    // entry / exit voxel: truncation then clip (sensor.py:345-346), tile-local
    const float gx = fmaf(a.z, t0, a.x), gy = fmaf(a.w, t0, a.y);
    const float hx = fmaf(a.z, t1, a.x), hy = fmaf(a.w, t1, a.y);
    const int ix = min(max(__float2int_rz((gx - P.bmin[0]) / vsx) - x0, 0), w - 1);
    const int iy = min(max(__float2int_rz((gy - P.bmin[1]) / vsy) - y0, 0), h - 1);
    const int jx = min(max(__float2int_rz((hx - P.bmin[0]) / vsx) - x0, 0), w - 1);
    const int jy = min(max(__float2int_rz((hy - P.bmin[1]) / vsy) - y0, 0), h - 1);

    float ux = INFINITY, uy = INFINITY, ax = 0.f, bx = 0.f, ay = 0.f, by = 0.f, fx = 0.f, fy = 0.f;
    int sx = 0, sy = 0, nx = 0, ny = 0;
    if (fabsf(a.z) > 1e-8f) {
        const float inv = sig / a.z;
        const bool pos = a.z > 0.f;
        sx = pos ? 4 : -4;  // byte strides
        nx = max(pos ? jx - ix : ix - jx, 0);
        // boundary times relative to the tile entry, u = (boundary - g) * sig / d; later boundaries are
        // fx whole voxels further (fx counts the steps taken)
        ax = fabsf(vsx * inv);
        bx = (fmaf(static_cast<float>(x0 + ix + (pos ? 1 : 0)), vsx, P.bmin[0]) - gx) * inv;
        fx = 0.f;
        if (nx > 0) ux = bx;
    }
    if (fabsf(a.w) > 1e-8f) {
        const float inv = sig / a.w;
        const bool pos = a.w > 0.f;
        sy = pos ? 4 * pitch : -4 * pitch;
        ny = max(pos ? jy - iy : iy - jy, 0);
        ay = fabsf(vsy * inv);
        by = (fmaf(static_cast<float>(y0 + iy + (pos ? 1 : 0)), vsy, P.bmin[1]) - gy) * inv;
        fy = 0.f;
        if (ny > 0) uy = by;
    }
    // one-voxel halo around the tile; shared-window byte address
    uint32_t addr = static_cast<uint32_t>(__cvta_generic_to_shared(tile)) + 4u * ((iy + 1) * pitch + ix + 1);
    // optical depths relative to the tile entry; the throughput at the entry scales the result
    const float e_in = ex2_approx(-sig * t0);
    em *= e_in;
    float e_prev = 1.f;
    float acc = 0.f;
    auto deposit = [&](float de) {
        if (kMode == kGather) {
            float v;
            asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
            acc = fmaf(v, de, acc);
        } else {
            // explicit state space: LDS + FADD + ATOMS.CAST.SPIN, no generic-address detour
            asm volatile("red.shared.add.f32 [%0], %1;" ::"r"(addr), "f"(em * de) : "memory");
        }
    };
    for (int n = nx + ny; n > 0; --n) {
        const float e = ex2_approx(-fminf(ux, uy));
        deposit(e_prev - e);
        e_prev = e;
        if (ux <= uy) {
            fx += 1.f;
            ux = fmaf(fx, ax, bx);
            addr += sx;
        } else {
            fy += 1.f;
            uy = fmaf(fy, ay, by);
            addr += sy;
        }
    }
    deposit(e_prev - ex2_approx(-sig * (t1 - t0)));
    return acc * e_in;
}

// kScatter: in = active_data [n_active], out = film (accumulated with red.global)
// kGather  : in = dL (film layout), out = pattern_grad [n_active] (pre-zeroed, red.global)
template <int kMode>
__global__ void __launch_bounds__(kPlanarThreads)
planar_kernel(const __grid_constant__ SceneParams P, const __grid_constant__ PlanarPlan L,
              const float* __restrict__ in, float* __restrict__ out, const uint32_t* __restrict__ active_pixels,
              uint64_t n_active, float ray_weight, const int* __restrict__ gate) {
    extern __shared__ __align__(16) float smem[];
    if (*gate != 0) return;
    const int tile_id = blockIdx.x, layer = blockIdx.y, group = blockIdx.z;
    const int rb = L.layer_row_begin[layer], re = L.layer_row_begin[layer + 1];
    if (rb == re) return;

    const int x0 = (tile_id % L.tiles_x) * L.tile_w, y0 = (tile_id / L.tiles_x) * L.tile_h;
    const int w = min(L.tile_w, P.res[0] - x0), h = min(L.tile_h, P.res[1] - y0);
    const int pitch = L.pitch;
    float* tile = smem;
    int* off = reinterpret_cast<int*>(smem + (L.tile_h + 2) * pitch);
    int* clo = off + P.n_patterns + 1;

    int k0, nk;
    call_angles(P, active_pixels, n_active, k0, nk);
    const int ka = k0 + static_cast<int>(static_cast<long long>(nk) * group / L.groups);
    const int kb = k0 + static_cast<int>(static_cast<long long>(nk) * (group + 1) / L.groups);

    for (int i = threadIdx.x; i <= P.n_patterns; i += kPlanarThreads)
        off[i] = L.tile_off[tile_id * (P.n_patterns + 1) + i];
    for (int i = threadIdx.x; i < P.n_patterns; i += kPlanarThreads) clo[i] = L.tile_clo[tile_id * P.n_patterns + i];
    const size_t layer_base = static_cast<size_t>(layer) * P.res[0] * P.res[1];
    for (int i = threadIdx.x; i < (h + 2) * pitch; i += kPlanarThreads) {
        float v = 0.f;
        if (kMode == kGather) {
            const int yy = i / pitch - 1, xx = i - (yy + 1) * pitch - 1;
            if (yy >= 0 && yy < h && xx >= 0 && xx < w)
                v = __ldg(in + layer_base + static_cast<size_t>(y0 + yy) * P.res[0] + (x0 + xx));
        }
        tile[i] = v;
    }
    __syncthreads();

    const int f_base = off[ka], total = off[kb] - f_base;
    for (int ri = rb; ri < re; ++ri) {
        const int row = L.layer_rows[ri];
        int k = ka;
        for (int f = threadIdx.x; f < total; f += kPlanarThreads) {
            const int g = f + f_base;
            while (off[k + 1] <= g) ++k;
            const int c = clo[k] + (g - off[k]);
            const int s = k * P.res_x + c;
            const float2 b = __ldg(L.seg_b + s);
            if (b.y == 0.f) continue;
            const size_t j = active_index(P, k - k0, row, c);
            float em = 0.f;
            if (kMode != kGather) {
                em = __ldg(in + j) * (ray_weight * b.y);
                if (em == 0.f) continue;
            }
            const float4 a = __ldg(L.seg_a + s);
            float t0, t1;
            if (!clip_to_tile(P, a, b.x, x0, y0, w, h, t0, t1)) continue;
            const float acc = march_tile<kMode>(P, tile, pitch, a, t0, t1, x0, y0, w, h, em);
            if (kMode == kGather && acc != 0.f) atomicAdd(out + j, acc * (ray_weight * b.y));
        }
    }
    if (kMode == kGather) return;
    __syncthreads();

    // flush: vectorised red.global.add when the tile rows are 16-byte aligned in the film
    auto value = [&](int i) { return tile[i]; };
    float* dst = out + layer_base;
    if (((x0 | w | P.res[0]) & 3) == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0) {
        const int w4 = w >> 2;
        for (int i = threadIdx.x; i < h * w4; i += kPlanarThreads) {
            const int yy = i / w4, xx = (i - yy * w4) << 2;
            const int t = (yy + 1) * pitch + xx + 1;
            const float v0 = value(t), v1 = value(t + 1), v2 = value(t + 2), v3 = value(t + 3);
            if (v0 != 0.f || v1 != 0.f || v2 != 0.f || v3 != 0.f)
                red_add_v4(dst + static_cast<size_t>(y0 + yy) * P.res[0] + (x0 + xx), v0, v1, v2, v3);
        }
    } else {
        for (int i = threadIdx.x; i < h * w; i += kPlanarThreads) {
            const int yy = i / w, xx = i - yy * w;
            const float v = value((yy + 1) * pitch + xx + 1);
            if (v != 0.f) atomicAdd(dst + static_cast<size_t>(y0 + yy) * P.res[0] + (x0 + xx), v);
        }
    }
}


}  // namespace

// ---------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------
#define PL_TRY(expr)                        \
    do {                                    \
        cudaError_t _e = (expr);            \
        if (_e != cudaSuccess) return _e;   \
    } while (0)

void planar_release(PlanarHost& H) {
    void* ptrs[] = {H.seg_a, H.seg_b, H.seg_d, H.layer_row_begin, H.layer_rows, H.tile_off, H.tile_clo, H.gate, H.frozen_list};
    for (void* p : ptrs)
        if (p) cudaFree(p);
    H = PlanarHost();
}

cudaError_t planar_tile_ranges(const SceneParams& P, const float4* seg_a, const float2* seg_b, int tiles_x, int tiles_y,
                               int tile_w, int tile_h, int* tile_off_dev, int* tile_clo_dev, float margin) {
    const int np = P.n_patterns, n_tiles = tiles_x * tiles_y;
    PlanarPlan L{};
    L.seg_a = seg_a; L.seg_b = seg_b;
    L.tiles_x = tiles_x; L.tiles_y = tiles_y; L.tile_w = tile_w; L.tile_h = tile_h;
    int* cnt_dev = nullptr;
    PL_TRY(cudaMalloc(&cnt_dev, sizeof(int) * n_tiles * np));
    planar_ranges_kernel<<<(n_tiles * np + 127) / 128, 128>>>(P, L, tile_clo_dev, cnt_dev, margin);
    count_launch();
    std::vector<int> cnt(static_cast<size_t>(n_tiles) * np), off(static_cast<size_t>(n_tiles) * (np + 1));
    cudaError_t e = cudaMemcpy(cnt.data(), cnt_dev, sizeof(int) * cnt.size(), cudaMemcpyDeviceToHost);
    cudaFree(cnt_dev);
    PL_TRY(e);
    for (int t = 0; t < n_tiles; ++t) {
        int acc = 0;
        for (int k = 0; k < np; ++k) { off[t * (np + 1) + k] = acc; acc += cnt[t * np + k]; }
        off[t * (np + 1) + np] = acc;
    }
    return cudaMemcpy(tile_off_dev, off.data(), sizeof(int) * off.size(), cudaMemcpyHostToDevice);
}

// Per-tile column LISTS: off[t][k] = entries of tile t before angle k, ebase[t] = entries before tile t,
// *list_dev (allocated here) = angle * res_x + column of every entry.
cudaError_t planar_tile_lists(const SceneParams& P, const float4* seg_a, const float2* seg_b, int tiles_x, int tiles_y,
                              int tile_w, int tile_h, int* tile_off_dev, int* tile_ebase_dev, int** list_dev,
                              long long* n_entries, float margin) {
    const int n_tiles = tiles_x * tiles_y, np = P.n_patterns;
    PlanarPlan L{};
    L.seg_a = seg_a; L.seg_b = seg_b;
    L.tiles_x = tiles_x; L.tiles_y = tiles_y; L.tile_w = tile_w; L.tile_h = tile_h;
    int* cnt_dev = nullptr;
    PL_TRY(cudaMalloc(&cnt_dev, sizeof(int) * n_tiles * np));
    planar_lists_kernel<false><<<(n_tiles * np + 127) / 128, 128>>>(P, L, nullptr, nullptr, cnt_dev, nullptr, margin);
    count_launch();
    std::vector<int> cnt(static_cast<size_t>(n_tiles) * np), off(static_cast<size_t>(n_tiles) * (np + 1)), ebase(n_tiles + 1);
    cudaError_t e = cudaMemcpy(cnt.data(), cnt_dev, sizeof(int) * cnt.size(), cudaMemcpyDeviceToHost);
    cudaFree(cnt_dev);
    PL_TRY(e);
    long long total = 0;
    for (int t = 0; t < n_tiles; ++t) {
        int acc = 0;
        for (int k = 0; k < np; ++k) { off[t * (np + 1) + k] = acc; acc += cnt[t * np + k]; }
        off[t * (np + 1) + np] = acc;
        ebase[t] = static_cast<int>(total);
        total += acc;
    }
    if (total >= (1ll << 31)) return cudaErrorInvalidValue;
    ebase[n_tiles] = static_cast<int>(total);
    PL_TRY(cudaMemcpy(tile_off_dev, off.data(), sizeof(int) * off.size(), cudaMemcpyHostToDevice));
    PL_TRY(cudaMemcpy(tile_ebase_dev, ebase.data(), sizeof(int) * ebase.size(), cudaMemcpyHostToDevice));
    PL_TRY(cudaMalloc(list_dev, sizeof(int) * static_cast<size_t>(std::max<long long>(total, 1))));
    planar_lists_kernel<true><<<(n_tiles * np + 127) / 128, 128>>>(P, L, tile_off_dev, tile_ebase_dev, nullptr, *list_dev,
                                                                 margin);
    count_launch();
    *n_entries = total;
    return cudaDeviceSynchronize();
}

// Builds the segment table, the layer -> rows CSR and the per-tile column
// ranges.  Synchronous (runs once per scene and depth setting).
cudaError_t planar_build(const SceneParams& P, PlanarHost& H) {
    planar_release(H);
    const int np = P.n_patterns, W = P.res_x, Hh = P.res_y;
    const int rx = P.res[0], ry = P.res[1], rz = P.res[2];
    constexpr int kMaxTile = 128;
    PlanarPlan& L = H.plan;
    L.tiles_x = (rx + kMaxTile - 1) / kMaxTile;
    L.tiles_y = (ry + kMaxTile - 1) / kMaxTile;
    L.tile_w = (((rx + L.tiles_x - 1) / L.tiles_x) + 3) & ~3;
    L.tile_h = (ry + L.tiles_y - 1) / L.tiles_y;
    L.tiles_x = (rx + L.tile_w - 1) / L.tile_w;
    L.pitch = (L.tile_w + 2) | 1;
    L.groups = 1;
    const int n_tiles = L.tiles_x * L.tiles_y;
    H.smem_bytes = sizeof(float) * static_cast<size_t>(L.tile_h + 2) * L.pitch + sizeof(int) * (2 * np + 1);

    PL_TRY(cudaMalloc(&H.seg_a, sizeof(float4) * np * W));
    PL_TRY(cudaMalloc(&H.seg_b, sizeof(float2) * np * W));
    PL_TRY(cudaMalloc(&H.seg_d, sizeof(SegmentD) * np * W));
    PL_TRY(cudaMalloc(&H.layer_row_begin, sizeof(int) * (rz + 1)));
    PL_TRY(cudaMalloc(&H.layer_rows, sizeof(int) * Hh));
    PL_TRY(cudaMalloc(&H.tile_off, sizeof(int) * n_tiles * (np + 1)));
    PL_TRY(cudaMalloc(&H.tile_clo, sizeof(int) * n_tiles * np));
    PL_TRY(cudaMalloc(&H.gate, sizeof(int) * 4));
    PL_TRY(cudaMemset(H.gate, 0, sizeof(int) * 4));
    PL_TRY(cudaMalloc(&H.frozen_list, sizeof(int) * np * W));
    L.seg_a = H.seg_a; L.seg_b = H.seg_b;
    L.layer_row_begin = H.layer_row_begin; L.layer_rows = H.layer_rows;
    L.tile_off = H.tile_off; L.tile_clo = H.tile_clo;

    // rows -> layers
    int* row_layer_dev = nullptr;
    float* row_z_dev = nullptr;
    PL_TRY(cudaMalloc(&row_layer_dev, sizeof(int) * Hh));
    PL_TRY(cudaMalloc(&row_z_dev, sizeof(float) * Hh));
    // the tables describe the pixel-centre rays, also when the passes themselves are jittered
    SceneParams R = P;
    R.regular_sampling = 1;
    planar_row_layer_kernel<<<(Hh + 127) / 128, 128>>>(R, row_layer_dev, row_z_dev);
    count_launch();
    std::vector<int> row_layer(Hh);
    std::vector<float> row_z(Hh);
    cudaError_t e = cudaMemcpy(row_layer.data(), row_layer_dev, sizeof(int) * Hh, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess) e = cudaMemcpy(row_z.data(), row_z_dev, sizeof(float) * Hh, cudaMemcpyDeviceToHost);
    cudaFree(row_layer_dev);
    cudaFree(row_z_dev);
    PL_TRY(e);
    std::vector<int> begin(rz + 1, 0), rows;
    int row_ref = -1;
    for (int r = 0; r < Hh; ++r)
        if (row_layer[r] >= 0) { ++begin[row_layer[r] + 1]; if (row_ref < 0) row_ref = r; }
    for (int z = 0; z < rz; ++z) begin[z + 1] += begin[z];
    rows.assign(Hh, 0);
    {
        std::vector<int> cur(begin.begin(), begin.end() - 1);
        for (int r = 0; r < Hh; ++r)
            if (row_layer[r] >= 0) rows[cur[row_layer[r]]++] = r;
    }
    H.n_rows_in_sensor = begin[rz];
    PL_TRY(cudaMemcpy(H.layer_row_begin, begin.data(), sizeof(int) * (rz + 1), cudaMemcpyHostToDevice));
    PL_TRY(cudaMemcpy(H.layer_rows, rows.data(), sizeof(int) * Hh, cudaMemcpyHostToDevice));
    if (row_ref < 0) row_ref = Hh / 2;

    // segment table (flags live in gate[1])
    planar_segments_kernel<<<(np * W + kPlanarThreads - 1) / kPlanarThreads, kPlanarThreads>>>(
        R, row_ref, H.seg_a, H.seg_b, H.seg_d, H.gate + 1, H.frozen_list, H.gate + 2);
    count_launch();
    PL_TRY(cudaGetLastError());
    int flags = 0;
    PL_TRY(cudaMemcpy(&flags, H.gate + 1, sizeof(int), cudaMemcpyDeviceToHost));
    H.flags = flags;
    PL_TRY(cudaMemcpy(&H.n_frozen, H.gate + 2, sizeof(int), cudaMemcpyDeviceToHost));

    PL_TRY(planar_tile_ranges(P, H.seg_a, H.seg_b, L.tiles_x, L.tiles_y, L.tile_w, L.tile_h, H.tile_off, H.tile_clo));
    H.row_layer = row_layer;
    H.row_z = row_z;

    PL_TRY(cudaFuncSetAttribute(planar_kernel<kScatter>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                static_cast<int>(H.smem_bytes)));
    PL_TRY(cudaFuncSetAttribute(planar_kernel<kGather>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                static_cast<int>(H.smem_bytes)));
    H.valid = true;
    return cudaSuccess;
}

static int planar_groups(const SceneParams& P, const PlanarPlan& L, uint64_t n_active) {
    const long long ctas = static_cast<long long>(L.tiles_x) * L.tiles_y * P.res[2];
    const long long nk = static_cast<long long>(n_active / (static_cast<uint64_t>(P.res_x) * P.n_rows));
    long long g = (148 * 8 + ctas - 1) / ctas;
    if (g > nk) g = nk;
    if (g > 32) g = 32;
    return g < 1 ? 1 : static_cast<int>(g);
}

cudaError_t launch_planar_check(const SceneParams& P, PlanarHost& H, const uint32_t* active_pixels, uint64_t n_active,
                                cudaStream_t stream) {
    PL_TRY(cudaMemsetAsync(H.gate, 0, sizeof(int), stream));
    const uint32_t hw = static_cast<uint32_t>(P.res_x) * P.res_y, rw = static_cast<uint32_t>(P.res_x) * P.n_rows;
    planar_check_pixels_kernel<<<static_cast<unsigned>((n_active + 255) / 256), 256, 0, stream>>>(
        active_pixels, n_active, hw, rw, static_cast<uint32_t>(P.row_begin) * P.res_x, H.gate);
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_planar_forward(const SceneParams& P, PlanarHost& H, const float* active_data,
                                  const uint32_t* active_pixels, uint64_t n_active, float ray_weight, float* film,
                                  cudaStream_t stream, const int* gate) {
    if (!gate) gate = H.gate;
    PlanarPlan L = H.plan;
    L.groups = planar_groups(P, L, n_active);
    const dim3 grid(L.tiles_x * L.tiles_y, P.res[2], L.groups);
    planar_kernel<kScatter><<<grid, kPlanarThreads, H.smem_bytes, stream>>>(P, L, active_data, film, active_pixels,
                                                                           n_active, ray_weight, gate);
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_planar_adjoint(const SceneParams& P, PlanarHost& H, const float* dL, const uint32_t* active_pixels,
                                  uint64_t n_active, float ray_weight, float* grad, cudaStream_t stream,
                                  const int* gate) {
    if (!gate) gate = H.gate;
    PlanarPlan L = H.plan;
    L.groups = planar_groups(P, L, n_active);
    const dim3 grid(L.tiles_x * L.tiles_y, P.res[2], L.groups);
    planar_kernel<kGather><<<grid, kPlanarThreads, H.smem_bytes, stream>>>(P, L, dL, grad, active_pixels, n_active,
                                                                          ray_weight, gate);
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_planar_frozen(const SceneParams& P, PlanarHost& H, int mode, const float* in, float* out,
                                 const uint32_t* active_pixels, uint64_t n_active, float ray_weight,
                                 cudaStream_t stream, const int* gate) {
    if (H.n_frozen <= 0) return cudaSuccess;
    if (!gate) gate = H.gate;
    const unsigned blocks = static_cast<unsigned>((static_cast<size_t>(H.n_frozen) * P.n_rows + kPlanarThreads - 1) /
                                                  kPlanarThreads);
    if (mode == 1)
        planar_frozen_kernel<1><<<blocks, kPlanarThreads, 0, stream>>>(P, H.frozen_list, H.n_frozen, in, out,
                                                                       active_pixels, n_active, ray_weight, gate);
    else
        planar_frozen_kernel<0><<<blocks, kPlanarThreads, 0, stream>>>(P, H.frozen_list, H.n_frozen, in, out,
                                                                       active_pixels, n_active, ray_weight, gate);
    count_launch();
    return cudaGetLastError();
}

}  // namespace drtvam
