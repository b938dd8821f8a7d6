// Layer-parallel ("z-lane") planar kernels: the fast path of the deterministic
// (regular_sampling) forward projection and adjoint.
//
// In a planar scene (planar.cuh) the medium segment of a ray depends only on
// (angle, DMD column); all DMD rows share it and differ only in the z-layer
// they deposit in.  So a WARP marches ONE (angle, column) segment through a
// small xy-tile and its 32 lanes are 32 different z-layers (times kL layers per
// lane in registers):
//   - the 2-D voxel DDA is warp-uniform: no divergence, one traversal is
//     amortised over 32 * kL rays;
//   - the shared tile is laid out [y][x][33 lanes][kL] floats, so a lane's kL
//     layers are one 4 / 8 / 16-byte word and every shared access of a warp is
//     a run of consecutive words (no bank conflicts); the forward accumulates
//     with ONE compare-and-swap per lane and voxel (red.shared.add.f32,
//     atom.shared.cas.b64 or atom.shared.cas.b128 for kL = 1, 2, 4);
//   - pattern values / pattern gradients are read / written as 128-byte lines
//     from a [angle][column][active layer] repack of active_data.
// Forward: red.shared.add.f32 into the tile, one red.global.add.v4.f32 flush.
// Adjoint: the tile holds the dose gradient, plain ld.shared gathers, one
// coalesced red.global per (ray, tile) into the repacked gradient.
// Reference: DDAVolumetricSensor.accumulate (sensor.py:306-440), film.write
// (film.py:40-41), VolumeIntegrator.render / render_backward
// (integrators/volume.py:18-56, 97-134).
#include <type_traits>
#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

#include <cuda_fp16.h>
#include <cub/device/device_scan.cuh>

#include "expm.cuh"
#include "kernels.h"
#include "planar.cuh"

namespace drtvam {

namespace {

constexpr int kZThreads = 1024;
constexpr int kZWarps = kZThreads / 32;
constexpr int kZLaneStride = 33;  // floats per voxel: 32 lanes + 1 pad (conflict-free flush)
constexpr float kLog2e = 1.4426950408889634f;
constexpr int kScatter = 0, kGather = 1;

// active_data [kk][row][col]  ->  pT [kk][col][a]  (sum over the rows of layer a)
__global__ void __launch_bounds__(256)
zpack_kernel(const __grid_constant__ SceneParams P, const __grid_constant__ PlanarZPlan Z,
             const float* __restrict__ active_data, float* __restrict__ pT, int nk, const int* __restrict__ gate) {
    __shared__ float tr[32][33];
    if (*gate != 0) return;
    const int kk = blockIdx.z;
    const int c0 = blockIdx.x * 32, a0 = blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
    for (int j = ty; j < 32; j += 8) {
        const int a = a0 + j, c = c0 + tx;
        float v = 0.f;
        if (a < Z.nz && c < P.res_x) {
            const int rb = __ldg(Z.a_row_begin + a), re = __ldg(Z.a_row_begin + a + 1);
            for (int ri = rb; ri < re; ++ri) {
                const int r = __ldg(Z.a_rows + ri);
                v += __ldg(active_data + active_index(P, kk, r, c));
            }
        }
        tr[j][tx] = v;
    }
    __syncthreads();
    for (int j = ty; j < 32; j += 8) {
        const int c = c0 + j, a = a0 + tx;
        if (c < P.res_x && a < Z.nz_pad)
            pT[(static_cast<size_t>(kk) * P.res_x + c) * Z.nz_pad + a] = tr[tx][j];
    }
    (void)nk;
}

// gT [kk][col][a]  ->  pattern_grad [kk][row][col]  (every row of a layer gets the layer's value)
__global__ void __launch_bounds__(256)
zunpack_kernel(const __grid_constant__ SceneParams P, const __grid_constant__ PlanarZPlan Z,
               const float* __restrict__ gT, float* __restrict__ grad, const int* __restrict__ gate) {
    __shared__ float tr[32][33];
    if (*gate != 0) return;
    const int kk = blockIdx.z;
    const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int j = ty; j < 32; j += 8) {
        const int c = c0 + j, r = P.row_begin + r0 + tx;
        float v = 0.f;
        if (c < P.res_x && r < P.row_begin + P.n_rows) {
            const int a = __ldg(Z.row_a + r);
            if (a >= 0) v = __ldg(gT + (static_cast<size_t>(kk) * P.res_x + c) * Z.nz_pad + a);
        }
        tr[j][tx] = v;
    }
    __syncthreads();
    for (int j = ty; j < 32; j += 8) {
        const int r = P.row_begin + r0 + j, c = c0 + tx;
        if (r < P.row_begin + P.n_rows && c < P.res_x) grad[active_index(P, kk, r, c)] = tr[tx][j];
    }
}

// ---- split fp16 operands of the patch gather (planarg_h16_kernel) -----------------------------------------
// x ~ hi + lo with hi = fp16(x), lo = fp16(x - hi): 22 bits of x when x is scaled into fp16's normal range.
__device__ __forceinline__ uint32_t pack_split_h16(float x) {
    const __half hi = __float2half_rn(x);
    const __half lo = __float2half_rn(x - __half2float(hi));
    return static_cast<uint32_t>(__half_as_ushort(hi)) | (static_cast<uint32_t>(__half_as_ushort(lo)) << 16);
}

// power of two that brings values of magnitude <= m into [1024, 2048) (m = 0: 1)
__host__ __device__ __forceinline__ float h16_scale(float m) {
    if (!(m > 0.f) || !isfinite(m)) return 1.f;
    int e;
    frexpf(m, &e);   // m = f 2^e, f in [0.5, 1)
    return ldexpf(1.f, 11 - e);
}

// pmax[0] = largest |v[i]| as the bits of a non-negative float
__global__ void __launch_bounds__(256)
absmax_kernel(const float* __restrict__ v, size_t n, unsigned* __restrict__ pmax, const int* __restrict__ gate) {
    if (gate && *gate != 0) return;
    float m = 0.f;
    for (size_t i = static_cast<size_t>(blockIdx.x) * 256 + threadIdx.x; i < n; i += static_cast<size_t>(gridDim.x) * 256) {
        const float a = fabsf(__ldg(v + i));
        if (a < INFINITY) m = fmaxf(m, a);
    }
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0 && __float_as_uint(m) > *reinterpret_cast<volatile unsigned*>(pmax)) atomicMax(pmax, __float_as_uint(m));
}

// active_data [kk][row][col]  ->  pT [kk][col][a] as {hi, lo} fp16 pairs of (sum over the rows of layer a) x scale,
// scale = h16_scale(rows_max x largest |pattern value| of the call)
__global__ void __launch_bounds__(256)
zpack_h16_kernel(const __grid_constant__ SceneParams P, const __grid_constant__ PlanarZPlan Z,
                 const float* __restrict__ active_data, uint32_t* __restrict__ pT, const unsigned* __restrict__ pmax,
                 float rows_max, const int* __restrict__ gate) {
    __shared__ float tr[32][33];
    if (*gate != 0) return;
    const float scale = h16_scale(__uint_as_float(__ldg(pmax)) * rows_max);
    const int kk = blockIdx.z;
    const int c0 = blockIdx.x * 32, a0 = blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
    for (int j = ty; j < 32; j += 8) {
        const int a = a0 + j, c = c0 + tx;
        float v = 0.f;
        if (a < Z.nz && c < P.res_x) {
            const int rb = __ldg(Z.a_row_begin + a), re = __ldg(Z.a_row_begin + a + 1);
            for (int ri = rb; ri < re; ++ri) {
                const int r = __ldg(Z.a_rows + ri);
                v += __ldg(active_data + active_index(P, kk, r, c));
            }
        }
        tr[j][tx] = v;
    }
    __syncthreads();
    for (int j = ty; j < 32; j += 8) {
        const int c = c0 + j, a = a0 + tx;
        if (c < P.res_x && a < Z.nz_pad)
            pT[(static_cast<size_t>(kk) * P.res_x + c) * Z.nz_pad + a] = pack_split_h16(tr[tx][j] * scale);
    }
}

// coef_h[i] = {hi, lo} fp16 pair of coef[i] x scale
__global__ void __launch_bounds__(256)
gcoef_h16_kernel(const float* __restrict__ coef, size_t n, float scale, uint32_t* __restrict__ coef_h) {
    const size_t i = static_cast<size_t>(blockIdx.x) * 256 + threadIdx.x;
    if (i < n) coef_h[i] = pack_split_h16(coef[i] * scale);
}

// t [kk][col][a] = 0 where the lane (k0 + kk, col, a) is marked: lanes that other kernels serve (the tile lanes
// of a surface-aware scene, inside_kernels.cu), in the repacked patterns before the forward pass and in the
// repacked gradient after the gather
__global__ void __launch_bounds__(256)
zmask_kernel(const __grid_constant__ SceneParams P, const __grid_constant__ PlanarZPlan Z, float* __restrict__ t,
             const uint32_t* __restrict__ mask, const uint32_t* __restrict__ active_pixels, uint64_t n_active,
             const int* __restrict__ gate) {
    if (*gate != 0) return;
    int k0, nk;
    call_angles(P, active_pixels, n_active, k0, nk);
    const size_t words_per_angle = (static_cast<size_t>(P.res_x) * Z.nz_pad) >> 5;   // nz_pad is a multiple of 32
    const size_t wi = static_cast<size_t>(blockIdx.x) * 256 + threadIdx.x;
    if (wi >= words_per_angle * nk) return;
    uint32_t m = __ldg(mask + static_cast<size_t>(k0) * words_per_angle + wi);
    while (m) {
        const int b = __ffs(m) - 1;
        m &= m - 1;
        t[(wi << 5) + b] = 0.f;
    }
}

// tile[addr .. addr + kL) += em[0..kL) * de, atomically with respect to the other warps of the CTA.
template <int kL>
__device__ __forceinline__ void z_scatter(uint32_t addr, const float (&em)[kL], float de) {
    if constexpr (kL == 1) {
        asm volatile("red.shared.add.f32 [%0], %1;" ::"r"(addr), "f"(em[0] * de) : "memory");
    } else if constexpr (kL == 2) {
        float o0, o1;
        asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(o0), "=f"(o1) : "r"(addr));
        bool again;
        do {
            const float n0 = fmaf(em[0], de, o0), n1 = fmaf(em[1], de, o1);
            float r0, r1;
            asm volatile(
                "{\n\t.reg .b64 c, u, r;\n\t"
                "mov.b64 c, {%2, %3};\n\t"
                "mov.b64 u, {%4, %5};\n\t"
                "atom.shared.cas.b64 r, [%6], c, u;\n\t"
                "mov.b64 {%0, %1}, r;\n\t}"
                : "=f"(r0), "=f"(r1)
                : "f"(o0), "f"(o1), "f"(n0), "f"(n1), "r"(addr)
                : "memory");
            again = (__float_as_uint(r0) != __float_as_uint(o0)) | (__float_as_uint(r1) != __float_as_uint(o1));
            o0 = r0; o1 = r1;
        } while (again);
    } else {
        float o0, o1, o2, o3;
        asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(o0), "=f"(o1), "=f"(o2), "=f"(o3) : "r"(addr));
        bool again;
        do {
            const float n0 = fmaf(em[0], de, o0), n1 = fmaf(em[1], de, o1);
            const float n2 = fmaf(em[2], de, o2), n3 = fmaf(em[3], de, o3);
            float r0, r1, r2, r3;
            asm volatile(
                "{\n\t.reg .b128 c, u, r;\n\t"
                "mov.b128 c, {%4, %5, %6, %7};\n\t"
                "mov.b128 u, {%8, %9, %10, %11};\n\t"
                "atom.shared.cas.b128 r, [%12], c, u;\n\t"
                "mov.b128 {%0, %1, %2, %3}, r;\n\t}"
                : "=f"(r0), "=f"(r1), "=f"(r2), "=f"(r3)
                : "f"(o0), "f"(o1), "f"(o2), "f"(o3), "f"(n0), "f"(n1), "f"(n2), "f"(n3), "r"(addr)
                : "memory");
            again = (__float_as_uint(r0) != __float_as_uint(o0)) | (__float_as_uint(r1) != __float_as_uint(o1)) |
                    (__float_as_uint(r2) != __float_as_uint(o2)) | (__float_as_uint(r3) != __float_as_uint(o3));
            o0 = r0; o1 = r1; o2 = r2; o3 = r3;
        } while (again);
    }
}

// packed fp32 pairs (Blackwell FFMA2)
__device__ __forceinline__ uint64_t pack_f32x2(float lo, float hi) {
    uint64_t r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ float lo_f32x2(uint64_t v) {
    float lo, hi;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
    return lo;
}
__device__ __forceinline__ float hi_f32x2(uint64_t v) {
    float lo, hi;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
    return hi;
}
__device__ __forceinline__ void fma_f32x2(uint64_t& acc, uint64_t a, uint64_t b) {
    asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(acc) : "l"(a), "l"(b));
}

__device__ __forceinline__ void unpack_f32x2(uint64_t v, float& lo, float& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}

// acc[0..kL) += tile[addr .. addr + kL) * de
template <int kL>
__device__ __forceinline__ void z_gather(uint32_t addr, float (&acc)[kL], float de) {
    if constexpr (kL == 1) {
        float v;
        asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
        acc[0] = fmaf(v, de, acc[0]);
    } else if constexpr (kL == 2) {
        // packed pair: one FFMA2 (two IEEE fp32 fmas, same bits as two fmaf) per two layers
        uint64_t v01, a01 = pack_f32x2(acc[0], acc[1]);
        asm volatile("ld.shared.b64 %0, [%1];" : "=l"(v01) : "r"(addr));
        fma_f32x2(a01, v01, pack_f32x2(de, de));
        unpack_f32x2(a01, acc[0], acc[1]);
    } else {
        uint64_t v01, v23, a01 = pack_f32x2(acc[0], acc[1]), a23 = pack_f32x2(acc[2], acc[3]);
        asm volatile("ld.shared.v2.b64 {%0, %1}, [%2];" : "=l"(v01), "=l"(v23) : "r"(addr));
        const uint64_t dd = pack_f32x2(de, de);
        fma_f32x2(a01, v01, dd);
        fma_f32x2(a23, v23, dd);
        unpack_f32x2(a01, acc[0], acc[1]);
        unpack_f32x2(a23, acc[2], acc[3]);
    }
}

// Two deposits in flight: the loads and the compare-and-swaps of two consecutive voxels of the
// walk are issued back to back, so the shared-memory latency of one hides behind the other (a warp
// never visits the same voxel twice in a row, so the two never alias within the warp).
template <int kL>
__device__ __forceinline__ void z_scatter2(uint32_t addr_a, float de_a, uint32_t addr_b, float de_b,
                                           const float (&em)[kL]) {
    if constexpr (kL == 4) {
        float a0, a1, a2, a3, b0, b1, b2, b3;
        asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(a0), "=f"(a1), "=f"(a2), "=f"(a3) : "r"(addr_a));
        asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(b0), "=f"(b1), "=f"(b2), "=f"(b3) : "r"(addr_b));
        float ra0, ra1, ra2, ra3, rb0, rb1, rb2, rb3;
        asm volatile(
            "{\n\t.reg .b128 c, u, r;\n\t"
            "mov.b128 c, {%4, %5, %6, %7};\n\t"
            "mov.b128 u, {%8, %9, %10, %11};\n\t"
            "atom.shared.cas.b128 r, [%12], c, u;\n\t"
            "mov.b128 {%0, %1, %2, %3}, r;\n\t}"
            : "=f"(ra0), "=f"(ra1), "=f"(ra2), "=f"(ra3)
            : "f"(a0), "f"(a1), "f"(a2), "f"(a3), "f"(fmaf(em[0], de_a, a0)), "f"(fmaf(em[1], de_a, a1)),
              "f"(fmaf(em[2], de_a, a2)), "f"(fmaf(em[3], de_a, a3)), "r"(addr_a)
            : "memory");
        asm volatile(
            "{\n\t.reg .b128 c, u, r;\n\t"
            "mov.b128 c, {%4, %5, %6, %7};\n\t"
            "mov.b128 u, {%8, %9, %10, %11};\n\t"
            "atom.shared.cas.b128 r, [%12], c, u;\n\t"
            "mov.b128 {%0, %1, %2, %3}, r;\n\t}"
            : "=f"(rb0), "=f"(rb1), "=f"(rb2), "=f"(rb3)
            : "f"(b0), "f"(b1), "f"(b2), "f"(b3), "f"(fmaf(em[0], de_b, b0)), "f"(fmaf(em[1], de_b, b1)),
              "f"(fmaf(em[2], de_b, b2)), "f"(fmaf(em[3], de_b, b3)), "r"(addr_b)
            : "memory");
        const bool fail_a = (__float_as_uint(ra0) != __float_as_uint(a0)) | (__float_as_uint(ra1) != __float_as_uint(a1)) |
                            (__float_as_uint(ra2) != __float_as_uint(a2)) | (__float_as_uint(ra3) != __float_as_uint(a3));
        const bool fail_b = (__float_as_uint(rb0) != __float_as_uint(b0)) | (__float_as_uint(rb1) != __float_as_uint(b1)) |
                            (__float_as_uint(rb2) != __float_as_uint(b2)) | (__float_as_uint(rb3) != __float_as_uint(b3));
        if (fail_a) z_scatter<kL>(addr_a, em, de_a);  // another warp got in between: the plain retry loop
        if (fail_b) z_scatter<kL>(addr_b, em, de_b);
    } else {
        z_scatter<kL>(addr_a, em, de_a);
        z_scatter<kL>(addr_b, em, de_b);
    }
}

// ---------------------------------------------------------------------------
// Walk entries.  The clipped 2-D walk of a (tile, angle, column) triple depends
// only on the scene, not on the patterns, so it is computed once per scene
// (and depth setting) and every pass of every z-block just reads it:
//   e[0] = { w, u1, ux, uy }    w = throughput at the TILE ENTRY (w0 2^-u(t0)); u = sigma_t log2(e) (t - t0)
//                               at the tile exit and at the first x / y voxel boundary (inf = none)
//   e[1] = { ax, ay, -, bits }  boundary-to-boundary increments of u and
//                               ix | iy<<7 | nx<<14 | ny<<21 | xneg<<28 | yneg<<29 | valid<<30
// Optical depths are RELATIVE to the tile entry, u = (boundary - g) sig / d: the deposit of a voxel is
// w (2^-u_in - 2^-u_out), and the rounding of an absolute u ~ 2 (2.4e-7) would be divided by the
// per-voxel extinction ~2.5e-3 - 1e-4 of noise per deposit; relative to the tile it is 16 x smaller.
// ---------------------------------------------------------------------------
constexpr unsigned kEntryValid = 1u << 30;

// The walk entry of one (angle k, column c) segment clipped to the tile [x0, x0+w) x [y0, y0+h).  Runs once
// per scene, in double precision on the double-precision segment (planar_f64.cuh): positions of ~10 mm have
// an fp32 ulp of 4e-5 voxels of 25 um, which would sit in the first and last deposit of EVERY tile crossing.
__device__ __forceinline__ void make_entry(const SceneParams& P, const SegmentD* __restrict__ seg_d, int k, int c, int x0,
                                           int y0, int w, int h, float4& e0, float4& e1) {
    e0 = make_float4(0.f, 0.f, INFINITY, INFINITY);
    e1 = make_float4(0.f, 0.f, 0.f, __uint_as_float(0u));
    const SegmentD S = seg_d[k * P.res_x + c];
    if (S.w0 == 0.0) return;
    const double bx0 = P.bmin[0], by0 = P.bmin[1];
    const double vsx = (static_cast<double>(P.bmax[0]) - bx0) / P.res[0];
    const double vsy = (static_cast<double>(P.bmax[1]) - by0) / P.res[1];
    // clip to the tile rectangle
    const double xl = bx0 + x0 * vsx, xh = bx0 + (x0 + w) * vsx, yl = by0 + y0 * vsy, yh = by0 + (y0 + h) * vsy;
    double t0 = 0.0, t1 = S.tmax;
    if (fabs(S.dx) > 1e-8) {
        const double ta = (xl - S.ox) / S.dx, tb = (xh - S.ox) / S.dx;
        t0 = fmax(t0, fmin(ta, tb));
        t1 = fmin(t1, fmax(ta, tb));
    } else if (S.ox < xl || S.ox >= xh) {
        return;
    }
    if (fabs(S.dy) > 1e-8) {
        const double ta = (yl - S.oy) / S.dy, tb = (yh - S.oy) / S.dy;
        t0 = fmax(t0, fmin(ta, tb));
        t1 = fmin(t1, fmax(ta, tb));
    } else if (S.oy < yl || S.oy >= yh) {
        return;
    }
    if (!(t0 < t1)) return;
    const double sig = static_cast<double>(P.sigma_t) * 1.4426950408889634;
    // entry / exit voxel: truncation then clip (sensor.py:345-346), tile-local
    const double gx = fma(S.dx, t0, S.ox), gy = fma(S.dy, t0, S.oy);
    const double hx = fma(S.dx, t1, S.ox), hy = fma(S.dy, t1, S.oy);
    const int ix = min(max(static_cast<int>((gx - bx0) / vsx) - x0, 0), w - 1);
    const int iy = min(max(static_cast<int>((gy - by0) / vsy) - y0, 0), h - 1);
    const int jx = min(max(static_cast<int>((hx - bx0) / vsx) - x0, 0), w - 1);
    const int jy = min(max(static_cast<int>((hy - by0) / vsy) - y0, 0), h - 1);
    unsigned bits = kEntryValid | static_cast<unsigned>(ix) | (static_cast<unsigned>(iy) << 7);
    double ux = INFINITY, uy = INFINITY, ax = 0.0, ay = 0.0;
    if (fabs(S.dx) > 1e-8) {
        const double inv = sig / S.dx;
        const bool pos = S.dx > 0.0;
        const int nx = max(pos ? jx - ix : ix - jx, 0);
        ax = fabs(vsx * inv);
        if (nx > 0) ux = (bx0 + (x0 + ix + (pos ? 1 : 0)) * vsx - gx) * inv;
        bits |= (static_cast<unsigned>(nx) << 14) | (pos ? 0u : (1u << 28));
    }
    if (fabs(S.dy) > 1e-8) {
        const double inv = sig / S.dy;
        const bool pos = S.dy > 0.0;
        const int ny = max(pos ? jy - iy : iy - jy, 0);
        ay = fabs(vsy * inv);
        if (ny > 0) uy = (by0 + (y0 + iy + (pos ? 1 : 0)) * vsy - gy) * inv;
        bits |= (static_cast<unsigned>(ny) << 21) | (pos ? 0u : (1u << 29));
    }
    e0 = make_float4(static_cast<float>(S.w0 * exp2(-sig * t0)), static_cast<float>(sig * (t1 - t0)),
                     static_cast<float>(ux), static_cast<float>(uy));
    e1 = make_float4(static_cast<float>(ax), static_cast<float>(ay), 0.f, __uint_as_float(bits));
}

// (tile, angle, column) of walk entry e of a plan (binary searches in the prefix tables)
__device__ __forceinline__ void locate_entry(const SceneParams& P, const PlanarZPlan& Z, int e, int& tile_id, int& k,
                                             int& c) {
    const int n_tiles = Z.tiles_x * Z.tiles_y;
    int lo = 0, hi = n_tiles;  // last tile with ebase <= e
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (__ldg(Z.tile_ebase + mid) <= e) lo = mid; else hi = mid;
    }
    tile_id = lo;
    const int g = e - __ldg(Z.tile_ebase + tile_id);
    const int* off = Z.tile_off + static_cast<size_t>(tile_id) * (P.n_patterns + 1);
    lo = 0; hi = P.n_patterns;  // last angle with off <= g
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (__ldg(off + mid) <= g) lo = mid; else hi = mid;
    }
    k = lo;
    c = __ldg(Z.tile_clo + static_cast<size_t>(tile_id) * P.n_patterns + k) + (g - __ldg(off + k));
}

__global__ void __launch_bounds__(256)
zsetup_kernel(const __grid_constant__ SceneParams P, const __grid_constant__ PlanarZPlan Z, float4* __restrict__ entries,
              int total) {
    const int e = blockIdx.x * 256 + threadIdx.x;
    if (e >= total) return;
    int tile_id, k, c;
    locate_entry(P, Z, e, tile_id, k, c);
    const int x0 = (tile_id % Z.tiles_x) * Z.tile_w, y0 = (tile_id / Z.tiles_x) * Z.tile_h;
    const int w = min(Z.tile_w, P.res[0] - x0), h = min(Z.tile_h, P.res[1] - y0);
    float4 e0, e1;
    make_entry(P, Z.seg_d, k, c, x0, y0, w, h, e0, e1);
    entries[2 * static_cast<size_t>(e)] = e0;
    entries[2 * static_cast<size_t>(e) + 1] = e1;
}

// kScatter: in = pT, out = film (red.global flush).  kGather: in = dL (film layout), out = gT.
template <int kMode, int kL>
__global__ void __launch_bounds__(kZThreads, 1)
planarz_kernel(const __grid_constant__ SceneParams P, const __grid_constant__ PlanarZPlan Z,
               const float* __restrict__ in, float* __restrict__ out, const uint32_t* __restrict__ active_pixels,
               uint64_t n_active, float ray_weight, const int* __restrict__ gate) {
    extern __shared__ __align__(16) float tile[];
    if (*gate != 0) return;
    const int tile_id = blockIdx.x, zb = blockIdx.y, group = blockIdx.z;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int x0 = (tile_id % Z.tiles_x) * Z.tile_w, y0 = (tile_id / Z.tiles_x) * Z.tile_h;
    const int w = min(Z.tile_w, P.res[0] - x0), h = min(Z.tile_h, P.res[1] - y0);
    const int tw = Z.tile_w;
    constexpr int kVox = kZLaneStride * kL;  // floats per voxel: [33 lanes][kL layers]
    const int tile_floats = Z.tile_h * tw * kVox;
    const int a_base = Z.a_begin + zb * (32 * kL);
    const size_t layer_elems = static_cast<size_t>(P.res[0]) * P.res[1];

    // ---- stage the tile
    if (kMode == kScatter) {
        for (int i = threadIdx.x; i < tile_floats; i += kZThreads) tile[i] = 0.f;
    } else {
        const int per_layer = h * w;
        for (int i = threadIdx.x; i < kL * 32 * per_layer; i += kZThreads) {
            const int al = i / per_layer, rem = i - al * per_layer;
            const int yy = rem / w, xx = rem - yy * w;
            const int z = __ldg(Z.zl + a_base + al);
            float v = 0.f;
            if (z >= 0) v = __ldg(in + z * layer_elems + static_cast<size_t>(y0 + yy) * P.res[0] + (x0 + xx));
            tile[(yy * tw + xx) * kVox + (al & 31) * kL + (al >> 5)] = v;
        }
    }
    __syncthreads();

    int k0, nk;
    call_angles(P, active_pixels, n_active, k0, nk);
    const int ka = k0 + static_cast<int>(static_cast<long long>(nk) * group / Z.groups);
    const int kb = k0 + static_cast<int>(static_cast<long long>(nk) * (group + 1) / Z.groups);
    const int* __restrict__ off = Z.tile_off + static_cast<size_t>(tile_id) * (P.n_patterns + 1);
    const int* __restrict__ clo = Z.tile_clo + static_cast<size_t>(tile_id) * P.n_patterns;
    const float4* __restrict__ entries = Z.entries + 2 * static_cast<size_t>(__ldg(Z.tile_ebase + tile_id));

    const uint32_t tile_s = static_cast<uint32_t>(__cvta_generic_to_shared(tile)) + 4u * kL * lane;
    const int sxb = 4 * kVox, syb = 4 * kVox * tw;

    const int f_base = __ldg(off + ka), total = __ldg(off + kb) - f_base;
    int k = ka;
    for (int f = warp; f < total; f += kZWarps) {
        const int g = f + f_base;
        const float4 e1 = __ldg(entries + 2 * static_cast<size_t>(g) + 1);
        const unsigned bits = __float_as_uint(e1.w);
        if (!(bits & kEntryValid)) continue;
        const float4 e0 = __ldg(entries + 2 * static_cast<size_t>(g));
        while (__ldg(off + k + 1) <= g) ++k;
        const int c = __ldg(clo + k) + (g - __ldg(off + k));

        const float wray = ray_weight * e0.x;
        const size_t tbase = (static_cast<size_t>(k - k0) * P.res_x + c) * Z.nz_pad + a_base + lane;
        float em[kL], acc[kL];
#pragma unroll
        for (int m = 0; m < kL; ++m) { em[m] = 0.f; acc[m] = 0.f; }
        if (kMode == kScatter) {
            bool any = false;
#pragma unroll
            for (int m = 0; m < kL; ++m) {
                em[m] = __ldg(in + tbase + 32 * m) * wray;
                any = any || (em[m] != 0.f);
            }
            if (!__any_sync(0xffffffffu, any)) continue;  // dark ray in all of this block's layers
        }

        // ---- warp-uniform 2-D DDA inside the tile (sensor.py:345-438 restricted to z = const)
        const int ix = bits & 127, iy = (bits >> 7) & 127;
        int nx = (bits >> 14) & 127, ny = (bits >> 21) & 127;
        const int sx = (bits & (1u << 28)) ? -sxb : sxb, sy = (bits & (1u << 29)) ? -syb : syb;
        float ux = e0.z, uy = e0.w;
        const float ax = e1.x, ay = e1.y, bx = e0.z, by = e0.w;
        float fx = 0.f, fy = 0.f;
        uint32_t addr = tile_s + static_cast<uint32_t>((iy * tw + ix) * (4 * kVox));
        float e_prev = 1.f;   // 2^-0: optical depths are relative to the tile entry
        // every branch below is warp-uniform: all lanes walk the same (angle, column) segment.
        // `next` yields the voxel address and the deposit E(u_in) - E(u_out) of the current voxel
        // and steps to the neighbour across the nearer boundary (the last voxel ends at u1).
        int n = nx + ny;
        auto next = [&](uint32_t& a_out, float& de_out) {
            a_out = addr;
            float u_out = e0.y;
            if (n > 0) {
                if (ux <= uy) {
                    u_out = ux;
                    addr += sx;
                    fx += 1.f;
                    ux = fmaf(fx, ax, bx);
                    if (--nx == 0) ux = INFINITY;
                } else {
                    u_out = uy;
                    addr += sy;
                    fy += 1.f;
                    uy = fmaf(fy, ay, by);
                    if (--ny == 0) uy = INFINITY;
                }
            }
            --n;
            const float e = ex2_approx(-u_out);
            de_out = e_prev - e;
            e_prev = e;
        };
        if (kMode == kGather) {
            // plain loads: the compiler already overlaps them, keep the loop minimal
            for (; n > 0; --n) {
                if (ux <= uy) {
                    const float e = ex2_approx(-ux);
                    z_gather<kL>(addr, acc, e_prev - e);
                    e_prev = e;
                    addr += sx;
                    fx += 1.f;
                    ux = fmaf(fx, ax, bx);
                    if (--nx == 0) ux = INFINITY;
                } else {
                    const float e = ex2_approx(-uy);
                    z_gather<kL>(addr, acc, e_prev - e);
                    e_prev = e;
                    addr += sy;
                    fy += 1.f;
                    uy = fmaf(fy, ay, by);
                    if (--ny == 0) uy = INFINITY;
                }
            }
            z_gather<kL>(addr, acc, e_prev - ex2_approx(-e0.y));
        } else {
            int deposits = n + 1;
            for (; deposits >= 2; deposits -= 2) {
                uint32_t a_a, a_b;
                float de_a, de_b;
                next(a_a, de_a);
                next(a_b, de_b);
                z_scatter2<kL>(a_a, de_a, a_b, de_b, em);
            }
            if (deposits) {
                uint32_t a_a;
                float de_a;
                next(a_a, de_a);
                z_scatter<kL>(a_a, em, de_a);
            }
        }

        if (kMode == kGather) {
#pragma unroll
            for (int m = 0; m < kL; ++m)
                if (acc[m] != 0.f) atomicAdd(out + tbase + 32 * m, acc[m] * wray);
        }
    }
    if (kMode == kGather) return;
    __syncthreads();

    // ---- flush: one red.global per voxel (v4 when the rows are 16-byte aligned in the film)
    const bool vec = ((x0 | w | P.res[0]) & 3) == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0;
    if (vec) {
        const int w4 = w >> 2, per_layer = h * w4;
        for (int i = threadIdx.x; i < kL * 32 * per_layer; i += kZThreads) {
            const int al = i / per_layer, rem = i - al * per_layer;
            const int yy = rem / w4, xx = (rem - yy * w4) << 2;
            const int z = __ldg(Z.zl + a_base + al);
            if (z < 0) continue;
            const float* src = tile + (yy * tw + xx) * kVox + (al & 31) * kL + (al >> 5);
            const float v0 = src[0], v1 = src[kVox], v2 = src[2 * kVox], v3 = src[3 * kVox];
            if (v0 != 0.f || v1 != 0.f || v2 != 0.f || v3 != 0.f)
                red_add_v4(out + z * layer_elems + static_cast<size_t>(y0 + yy) * P.res[0] + (x0 + xx), v0, v1, v2, v3);
        }
    } else {
        const int per_layer = h * w;
        for (int i = threadIdx.x; i < kL * 32 * per_layer; i += kZThreads) {
            const int al = i / per_layer, rem = i - al * per_layer;
            const int yy = rem / w, xx = rem - yy * w;
            const int z = __ldg(Z.zl + a_base + al);
            if (z < 0) continue;
            const float v = tile[(yy * tw + xx) * kVox + (al & 31) * kL + (al >> 5)];
            if (v != 0.f) atomicAdd(out + z * layer_elems + static_cast<size_t>(y0 + yy) * P.res[0] + (x0 + xx), v);
        }
    }
}

// ---------------------------------------------------------------------------
// Visit lists: the adjoint without a DDA.  The walk of a (tile, angle, column) entry is the same in every
// pass, so the set-up also writes it out: the deposit factor 2^-u_in (1 - 2^-(u_out - u_in)) of every voxel
// the segment crosses in the tile - evaluated without cancellation (expm.cuh) - with the direction of the
// step that follows (0 = across an x boundary, 1 = across a y boundary) in the lowest mantissa bit.  The
// gather kernel then spends, per voxel and 32 * kL layers, one shuffle, one predicate and one address
// add next to the shared-memory load and the FMAs, instead of a compare, an ex2.approx, two subtractions,
// a counter and a select: 8 instructions instead of 25 (profiles/r02i_*: the DDA form was issue bound at
// 85 %).  3.6e8 bytes at C2, streamed once per z-block.
// ---------------------------------------------------------------------------
constexpr int kVisitMax = 64;   // a head carries no more; longer walks keep the DDA kernel

__global__ void __launch_bounds__(256)
vcount_kernel(const float4* __restrict__ entries, int total, int* __restrict__ count) {
    const int e = blockIdx.x * 256 + threadIdx.x;
    if (e > total) return;
    int n = 0;
    if (e < total) {
        const unsigned bits = __float_as_uint(__ldg(entries + 2 * static_cast<size_t>(e) + 1).w);
        if (bits & kEntryValid) n = static_cast<int>(((bits >> 14) & 127) + ((bits >> 21) & 127) + 1);
    }
    count[e] = (n + 3) & ~3;   // 16-byte aligned lists
}

__global__ void __launch_bounds__(256)
vfill_kernel(const float4* __restrict__ entries, const int* __restrict__ voff, int total, int tw, float4* __restrict__ vhead,
             float* __restrict__ vde, int* __restrict__ too_long) {
    const int e = blockIdx.x * 256 + threadIdx.x;
    if (e >= total) return;
    const float4 e0 = __ldg(entries + 2 * static_cast<size_t>(e)), e1 = __ldg(entries + 2 * static_cast<size_t>(e) + 1);
    const unsigned bits = __float_as_uint(e1.w);
    const int off = __ldg(voff + e);
    if (!(bits & kEntryValid)) {
        vhead[e] = make_float4(0.f, __int_as_float(0), __int_as_float(off), 0.f);
        return;
    }
    const int ix = bits & 127, iy = (bits >> 7) & 127;
    int nx = (bits >> 14) & 127, ny = (bits >> 21) & 127;
    const int n = nx + ny + 1;
    if (n > kVisitMax) atomicOr(too_long, 1);
    float ux = e0.z, uy = e0.w;
    const float ax = e1.x, ay = e1.y, bx = e0.z, by = e0.w;
    float fx = 0.f, fy = 0.f, u_prev = 0.f;
    for (int i = 0; i < n; ++i) {
        float u_out = e0.y;
        unsigned step_y = 0;
        if (i < n - 1) {
            if (ux <= uy) {
                u_out = ux; fx += 1.f;
                ux = fmaf(fx, ax, bx);
                if (--nx == 0) ux = INFINITY;
            } else {
                u_out = uy; fy += 1.f; step_y = 1;
                uy = fmaf(fy, ay, by);
                if (--ny == 0) uy = INFINITY;
            }
        }
        const float de = exp2f(-u_prev) * one_minus_ex2_neg(fmaxf(u_out - u_prev, 0.f));
        vde[off + i] = __uint_as_float((__float_as_uint(de) & ~1u) | step_y);
        u_prev = u_out;
    }
    // start voxel | n << 16 | x step negative << 30 | y step negative << 31
    const unsigned packed = static_cast<unsigned>(iy * tw + ix) | (static_cast<unsigned>(n) << 16) | (bits & (3u << 28)) << 2;
    vhead[e] = make_float4(e0.x, __uint_as_float(packed), __int_as_float(off), 0.f);
}

// ---- dropping the entries that visit nothing ---------------------------------------------------------------
// planar_tile_ranges gives ONE column range per (tile, angle); behind a refracting box two ray families cross a
// tile and the range spans everything between them, so at C4 most entries of a tile are empty.  The heads are
// compacted to the entries with visits (keep -> exclusive scan -> copy); a kept head carries its
// angle * res_x + column in .w, and the per-(tile, angle) prefix table is rewritten for the kept entries, so the
// gather kernel walks a dense list and needs no range tables to find (angle, column).
__global__ void __launch_bounds__(256)
vkeep_kernel(const int* __restrict__ count, int total, int* __restrict__ keep) {
    const int e = blockIdx.x * 256 + threadIdx.x;
    if (e <= total) keep[e] = (e < total && count[e] > 0) ? 1 : 0;
}

__global__ void __launch_bounds__(256)
vcompact_kernel(const __grid_constant__ SceneParams P, const __grid_constant__ PlanarZPlan Z, const int* __restrict__ keep,
                const int* __restrict__ kpos, int total, const float4* __restrict__ vhead, float4* __restrict__ vhead_c) {
    const int e = blockIdx.x * 256 + threadIdx.x;
    if (e >= total || !keep[e]) return;
    int tile_id, k, c;
    locate_entry(P, Z, e, tile_id, k, c);
    float4 h = vhead[e];
    h.w = __int_as_float(k * P.res_x + c);
    vhead_c[kpos[e]] = h;
}

// v_off[t][k] = kept entries of tile t before angle k; v_ebase[t] = kept entries before tile t
__global__ void __launch_bounds__(256)
vremap_kernel(const __grid_constant__ SceneParams P, const __grid_constant__ PlanarZPlan Z, const int* __restrict__ kpos,
              int* __restrict__ v_off, int* __restrict__ v_ebase) {
    const int n_tiles = Z.tiles_x * Z.tiles_y, np1 = P.n_patterns + 1;
    const long long i = static_cast<long long>(blockIdx.x) * 256 + threadIdx.x;
    if (i >= static_cast<long long>(n_tiles) * np1) return;
    const int t = static_cast<int>(i / np1), k = static_cast<int>(i - static_cast<long long>(t) * np1);
    const int eb = __ldg(Z.tile_ebase + t);
    const int first = kpos[eb];
    v_off[i] = kpos[eb + __ldg(Z.tile_off + i)] - first;
    if (k == 0) v_ebase[t] = first;
}

// in = dL (film layout), out = gT.  Same CTA / tile / z-block structure as planarz_kernel<kGather>.
template <int kL>
__global__ void __launch_bounds__(kZThreads, 1)
planarv_kernel(const __grid_constant__ SceneParams P, const __grid_constant__ PlanarZPlan Z,
               const float* __restrict__ in, float* __restrict__ out, const uint32_t* __restrict__ active_pixels,
               uint64_t n_active, float ray_weight, const int* __restrict__ gate) {
    extern __shared__ __align__(16) float tile[];
    if (*gate != 0) return;
    const int tile_id = blockIdx.x, zb = blockIdx.y, group = blockIdx.z;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int x0 = (tile_id % Z.tiles_x) * Z.tile_w, y0 = (tile_id / Z.tiles_x) * Z.tile_h;
    const int w = min(Z.tile_w, P.res[0] - x0), h = min(Z.tile_h, P.res[1] - y0);
    const int tw = Z.tile_w;
    constexpr int kVox = kZLaneStride * kL;
    const int a_base = Z.a_begin + zb * (32 * kL);
    const size_t layer_elems = static_cast<size_t>(P.res[0]) * P.res[1];

    // ---- stage the dose-gradient tile (as planarz_kernel)
    {
        const int per_layer = h * w;
        for (int i = threadIdx.x; i < kL * 32 * per_layer; i += kZThreads) {
            const int al = i / per_layer, rem = i - al * per_layer;
            const int yy = rem / w, xx = rem - yy * w;
            const int z = __ldg(Z.zl + a_base + al);
            float v = 0.f;
            if (z >= 0) v = __ldg(in + z * layer_elems + static_cast<size_t>(y0 + yy) * P.res[0] + (x0 + xx));
            tile[(yy * tw + xx) * kVox + (al & 31) * kL + (al >> 5)] = v;
        }
    }
    __syncthreads();

    int k0, nk;
    call_angles(P, active_pixels, n_active, k0, nk);
    const int ka = k0 + static_cast<int>(static_cast<long long>(nk) * group / Z.groups);
    const int kb = k0 + static_cast<int>(static_cast<long long>(nk) * (group + 1) / Z.groups);
    const int* __restrict__ off = Z.v_off + static_cast<size_t>(tile_id) * (P.n_patterns + 1);
    const size_t ebase = static_cast<size_t>(__ldg(Z.v_ebase + tile_id));
    const float4* __restrict__ vhead = Z.vhead + ebase;
    const uint32_t tile_s = static_cast<uint32_t>(__cvta_generic_to_shared(tile)) + 4u * kL * lane;
    const int sxb = 4 * kVox, syb = 4 * kVox * tw;

    const int f_base = __ldg(off + ka), total = __ldg(off + kb) - f_base;
    // the head and the list of the NEXT entry are fetched while the current one is walked
    float4 head = make_float4(0.f, 0.f, 0.f, 0.f);
    float d_lo = 0.f, d_hi = 0.f;
    auto fetch = [&](int f, float4& hd, float& lo, float& hi) {
        hd = make_float4(0.f, 0.f, 0.f, 0.f);
        lo = hi = 0.f;
        if (f < total) {
            hd = __ldg(vhead + f + f_base);
            const int n = (__float_as_uint(hd.y) >> 16) & 0x7f;
            const float* lst = Z.vde + __float_as_int(hd.z);
            if (lane < n) lo = __ldg(lst + lane);
            if (lane + 32 < n) hi = __ldg(lst + lane + 32);
        }
    };
    fetch(warp, head, d_lo, d_hi);
    for (int f = warp; f < total; f += kZWarps) {
        const float4 hd = head;
        const float my_lo = d_lo, my_hi = d_hi;
        fetch(f + kZWarps, head, d_lo, d_hi);
        const unsigned packed = __float_as_uint(hd.y);
        const int n = (packed >> 16) & 0x7f;
        if (n == 0) continue;
        // angle * res_x + column of the entry, relative to the first angle of the call
        const size_t tbase = (static_cast<size_t>(__float_as_int(hd.w)) - static_cast<size_t>(k0) * P.res_x) * Z.nz_pad + a_base + lane;
        const int sx = (packed & (1u << 30)) ? -sxb : sxb, sy = (packed & (1u << 31)) ? -syb : syb;
        uint32_t addr = tile_s + static_cast<uint32_t>((packed & 0xffffu) * (4 * kVox));
        float acc[kL];
#pragma unroll
        for (int m = 0; m < kL; ++m) acc[m] = 0.f;
        const int n_lo = min(n, 32);
#pragma unroll 4
        for (int i = 0; i < n_lo; ++i) {
            const float de = __shfl_sync(0xffffffffu, my_lo, i);
            z_gather<kL>(addr, acc, de);
            addr += (__float_as_uint(de) & 1u) ? sy : sx;
        }
        for (int i = 32; i < n; ++i) {
            const float de = __shfl_sync(0xffffffffu, my_hi, i - 32);
            z_gather<kL>(addr, acc, de);
            addr += (__float_as_uint(de) & 1u) ? sy : sx;
        }
        const float wray = ray_weight * hd.x;
#pragma unroll
        for (int m = 0; m < kL; ++m)
            if (acc[m] != 0.f) atomicAdd(out + tbase + 32 * m, acc[m] * wray);
    }
}

// ---------------------------------------------------------------------------
// Patch-gather forward.  The scatter kernel above spends its time on the
// read-modify-write of the shared tile (12 shared wavefronts and ~50 issue slots
// per 128 visits, profiles/r01a_planarz_ncu_details.txt).  Here a warp OWNS a
// 4x4 patch of voxels (lanes = z-layers, kL layers per lane) and keeps its
// 16 * kL accumulators in registers for the whole pass; for every (angle,
// column) segment that crosses the patch it reads a dense block of 16
// precomputed deposit coefficients (warp-uniform 128-bit loads; a segment
// visits 3 of the 16 voxels on average, the others are zero) and the kL pattern
// values of its layers, and issues 16 * kL FMAs.  No atomics, no shared memory,
// and the result is deterministic.  The coefficients come from the same clipped
// walk as the scatter kernel (make_entry + the walk below).
// ---------------------------------------------------------------------------
constexpr int kPatch = 4;

// Marches the clipped walk of entry (e0, e1) and calls visit(yy * tw + xx, de) for every voxel, with
// de = 2^-u_in (1 - 2^-(u_out - u_in)) evaluated without cancellation (expm.cuh): this runs once per
// scene, to make the coefficient blocks of the patch gather.
template <class Visit>
__device__ __forceinline__ void walk_entry(const float4 e0, const float4 e1, int tw, Visit&& visit) {
    const unsigned bits = __float_as_uint(e1.w);
    if (!(bits & kEntryValid)) return;
    const int ix = bits & 127, iy = (bits >> 7) & 127;
    int nx = (bits >> 14) & 127, ny = (bits >> 21) & 127;
    const int sx = (bits & (1u << 28)) ? -1 : 1, sy = (bits & (1u << 29)) ? -tw : tw;
    float ux = e0.z, uy = e0.w;
    const float ax = e1.x, ay = e1.y, bx = e0.z, by = e0.w;
    float fx = 0.f, fy = 0.f;
    int v = iy * tw + ix;
    float u_prev = 0.f;
    for (int n = nx + ny; n >= 0; --n) {
        float u_out = e0.y;
        const int v_here = v;
        if (n > 0) {
            if (ux <= uy) {
                u_out = ux; v += sx; fx += 1.f;
                ux = fmaf(fx, ax, bx);
                if (--nx == 0) ux = INFINITY;
            } else {
                u_out = uy; v += sy; fy += 1.f;
                uy = fmaf(fy, ay, by);
                if (--ny == 0) uy = INFINITY;
            }
        }
        visit(v_here, exp2f(-u_prev) * one_minus_ex2_neg(fmaxf(u_out - u_prev, 0.f)));
        u_prev = u_out;
    }
}

// coef[g][16] and ray[g] of every (patch, angle, column) pair g
__global__ void __launch_bounds__(256)
gsetup_kernel(const __grid_constant__ SceneParams P, const __grid_constant__ PlanarZPlan Z, float4* __restrict__ coef,
              int* __restrict__ ray, int total) {
    const int g = blockIdx.x * 256 + threadIdx.x;
    if (g >= total) return;
    int tile_id, k, c;
    locate_entry(P, Z, g, tile_id, k, c);
    const int x0 = (tile_id % Z.tiles_x) * kPatch, y0 = (tile_id / Z.tiles_x) * kPatch;
    const int w = min(kPatch, P.res[0] - x0), h = min(kPatch, P.res[1] - y0);
    float4 e0, e1;
    make_entry(P, Z.seg_d, k, c, x0, y0, w, h, e0, e1);
    float cf[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) cf[i] = 0.f;
    const float w0 = e0.x;  // throughput at the patch entry times sigma_a / sigma_t
    walk_entry(e0, e1, kPatch, [&](int v, float de) {
#pragma unroll
        for (int i = 0; i < 16; ++i)
            if (i == v) cf[i] += w0 * de;
    });
    float4* o = coef + 4 * static_cast<size_t>(g);
    o[0] = make_float4(cf[0], cf[1], cf[2], cf[3]);
    o[1] = make_float4(cf[4], cf[5], cf[6], cf[7]);
    o[2] = make_float4(cf[8], cf[9], cf[10], cf[11]);
    o[3] = make_float4(cf[12], cf[13], cf[14], cf[15]);
    int mask = 0;
#pragma unroll
    for (int q = 0; q < 4; ++q)
        if (cf[4 * q] != 0.f || cf[4 * q + 1] != 0.f || cf[4 * q + 2] != 0.f || cf[4 * q + 3] != 0.f) mask |= 1 << q;
    ray[g] = (k * P.res_x + c) | (mask << 24);  // planarg_build guarantees n_patterns * res_x < 2^24
}

// ---- dropping the pairs that touch nothing -------------------------------------------------------
// planar_tile_ranges gives ONE column range per (patch, angle).  Behind a refracting box (C4) two
// faces are lit at most angles and the two ray families cross inside the resin, so the range spans
// everything between them: 7 of 8 candidate pairs miss the patch.  The tables are compacted to the
// pairs with a non-empty row mask; the kernel only ever sees a contiguous range of pairs per
// (patch, angle range), so nothing else changes.
__global__ void __launch_bounds__(256) gkeep_kernel(const int* __restrict__ ray, int total, int* __restrict__ keep) {
    const int g = blockIdx.x * 256 + threadIdx.x;
    if (g <= total) keep[g] = (g < total && (__ldg(ray + g) >> 24) != 0) ? 1 : 0;
}

__global__ void __launch_bounds__(256)
gcompact_kernel(const float4* __restrict__ coef_in, const int* __restrict__ ray_in, const int* __restrict__ pos,
                int total, float4* __restrict__ coef_out, int* __restrict__ ray_out) {
    const int g = blockIdx.x * 256 + threadIdx.x;
    if (g >= total) return;
    const int r = __ldg(ray_in + g);
    if ((r >> 24) == 0) return;
    const int q = __ldg(pos + g);
    ray_out[q] = r;
#pragma unroll
    for (int i = 0; i < 4; ++i) coef_out[4 * static_cast<size_t>(q) + i] = __ldg(coef_in + 4 * static_cast<size_t>(g) + i);
}

// off[t][a] (prefix inside the tile) and ebase[t] follow the kept pairs; one thread per (tile, angle slot)
__global__ void __launch_bounds__(256)
gremap_kernel(int n_tiles, int np, const int* __restrict__ ebase_old, const int* __restrict__ off_old,
              const int* __restrict__ pos, int* __restrict__ ebase_new, int* __restrict__ off_new) {
    const long long i = static_cast<long long>(blockIdx.x) * 256 + threadIdx.x;
    if (i >= static_cast<long long>(n_tiles) * (np + 1)) return;
    const int t = static_cast<int>(i / (np + 1));
    const int e = __ldg(ebase_old + t);
    const int p0 = __ldg(pos + e);
    off_new[i] = __ldg(pos + e + __ldg(off_old + i)) - p0;
    if (i % (np + 1) == 0) {
        ebase_new[t] = p0;
        if (t == n_tiles - 1) ebase_new[n_tiles] = __ldg(pos + __ldg(ebase_old + n_tiles));
    }
}

constexpr int kGWarpsX = 4, kGWarpsY = 2;  // patches per CTA: a 16 x 8 voxel region, one warp per patch

// ---- the forward pass proper --------------------------------------------------------------------
// The two streams a warp consumes - 64 B of coefficients per pair (contiguous in g) and the
// 128 * kL byte pattern line of the pair's ray (lane l holds layers kL*l .. kL*l + kL-1) - are copied
// to shared memory one chunk of kGChunk pairs ahead with cp.async, so the FMA loop only reads shared
// memory (a first version that prefetched into registers sat at 42 % issue activity on
// long-scoreboard stalls, profiles/r01l_*).  ray[g] carries a 4-bit mask of the patch rows the
// segment touches (2.2 of 4 on average), so the FMAs and coefficient loads of untouched rows are
// skipped with warp-uniform branches.
constexpr int kGChunk = 8;
constexpr int g_buf_floats(int kL) { return kGChunk * 32 * kL + kGChunk * 16 + kGChunk; }
constexpr size_t g_smem_bytes(int kL) {
    return (sizeof(float) * 2 * g_buf_floats(kL) + 2 * sizeof(uint64_t)) * kGWarpsX * kGWarpsY;
}

// 16-byte asynchronous copies (LDGSTS): one warp instruction moves a 512-byte line, lane by lane
__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(static_cast<uint32_t>(__cvta_generic_to_shared(smem))),
                 "l"(gmem)
                 : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// TMA bulk copy (one instruction of one thread moves a whole line) completing on an mbarrier
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* smem, const void* gmem, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(smem)),
                 "l"(gmem), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

template <int kL>
__global__ void __launch_bounds__(32 * kGWarpsX * kGWarpsY, 3)
planarg_kernel(const __grid_constant__ SceneParams P, const __grid_constant__ PlanarZPlan Z,
               const float4* __restrict__ coef, const int* __restrict__ ray, const float* __restrict__ in,
               float* __restrict__ out, const uint32_t* __restrict__ active_pixels, uint64_t n_active, float ray_weight,
               const int* __restrict__ gate, int a_begin, int fresh) {
    extern __shared__ __align__(16) float g_smem[];
    if (*gate != 0) return;
    constexpr int kBuf = g_buf_floats(kL), kLine = 32 * kL;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int regions_x = (Z.tiles_x + kGWarpsX - 1) / kGWarpsX;
    const int px = (blockIdx.x % regions_x) * kGWarpsX + (warp % kGWarpsX);
    const int py = (blockIdx.x / regions_x) * kGWarpsY + (warp / kGWarpsX);
    if (px >= Z.tiles_x || py >= Z.tiles_y) return;
    const int tile_id = py * Z.tiles_x + px;
    const int a_base = a_begin + blockIdx.y * kLine;
    float* buf = g_smem + warp * (2 * kBuf);
    uint64_t* bars = reinterpret_cast<uint64_t*>(g_smem + kGWarpsX * kGWarpsY * 2 * kBuf) + 2 * warp;
    if (lane == 0) {
        mbar_init(bars, 1);
        mbar_init(bars + 1, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();

    int k0, nk;
    call_angles(P, active_pixels, n_active, k0, nk);
    const int* __restrict__ off = Z.tile_off + static_cast<size_t>(tile_id) * (P.n_patterns + 1);
    const int ebase = __ldg(Z.tile_ebase + tile_id);
    const int g0 = ebase + __ldg(off + k0), g1 = ebase + __ldg(off + k0 + nk);
    // pattern line of ray r for this block's layers: xb + r * nz_pad, 128 * kL bytes
    const float* __restrict__ xb = in + a_base - static_cast<long long>(k0) * P.res_x * Z.nz_pad;
    const int n_chunks = (g1 - g0 + kGChunk - 1) / kGChunk;

    float acc[16][kL];
#pragma unroll
    for (int i = 0; i < 16; ++i)
#pragma unroll
        for (int m = 0; m < kL; ++m) acc[i][m] = 0.f;

    // ray word of pair (lane % kGChunk) of a chunk, loaded one chunk before it is needed
    auto load_rays = [&](int c) {
        const int g = g0 + c * kGChunk + (lane & (kGChunk - 1));
        return (c < n_chunks && g < g1) ? __ldg(ray + g) : 0;  // mask 0: nothing to do
    };
    // lane j < kGChunk copies the pattern line of pair j, lane kGChunk the coefficient blocks of the chunk
    auto issue = [&](int c, int rl) {
        float* b = buf + (c & 1) * kBuf;
        uint64_t* bar = bars + (c & 1);
        const int g = g0 + c * kGChunk;
        const int n_here = min(kGChunk, g1 - g);
        const bool wants = lane < kGChunk && (rl >> 24) != 0;
        const unsigned lines = __popc(__ballot_sync(0xffffffffu, wants));
        if (lane == 0) mbar_expect_tx(bar, lines * (4u * kLine) + 64u * n_here);
        __syncwarp();
        if (wants) bulk_g2s(b + lane * kLine, xb + static_cast<size_t>(rl & 0xffffff) * Z.nz_pad, 4u * kLine, bar);
        if (lane == kGChunk) bulk_g2s(b + kGChunk * kLine, coef + 4 * static_cast<size_t>(g), 64u * n_here, bar);
        if (lane < kGChunk) reinterpret_cast<int*>(b)[kGChunk * kLine + kGChunk * 16 + lane] = rl;
    };

    int rl_next = load_rays(0);
    if (n_chunks > 0) issue(0, rl_next);
    rl_next = load_rays(1);
    for (int c = 0; c < n_chunks; ++c) {
        if (c + 1 < n_chunks) {
            issue(c + 1, rl_next);
            rl_next = load_rays(c + 2);
        }
        mbar_wait(bars + (c & 1), (c >> 1) & 1);
        __syncwarp();
        const float* b = buf + (c & 1) * kBuf;
        const int n_here = min(kGChunk, g1 - (g0 + c * kGChunk));
        for (int j = 0; j < n_here; ++j) {
            const int mask = reinterpret_cast<const int*>(b)[kGChunk * kLine + kGChunk * 16 + j] >> 24;
            if (mask == 0) continue;  // the segment does not deposit in this patch
            float x[kL];
            bool any = false;
#pragma unroll
            for (int m = 0; m < kL; ++m) { x[m] = b[j * kLine + kL * lane + m]; any = any || (x[m] != 0.f); }
            if (!__any_sync(0xffffffffu, any)) continue;  // dark in all of this block's layers
            const float4* cj = reinterpret_cast<const float4*>(b + kGChunk * kLine + 16 * j);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                if (!(mask & (1 << q))) continue;  // warp-uniform
                const float4 cq = cj[q];           // warp-uniform address: one shared wavefront
                const float cs[4] = {cq.x, cq.y, cq.z, cq.w};
                // (packed FFMA2 accumulators were measured here too: half the FMA issue slots, same
                // 11.2 ms - the loop is bound by dependent shared-memory latencies at 6 warps per
                // scheduler, not by issue slots or the FMA pipe (28 % busy); fetching pair j + 1's mask
                // and pattern values ahead of pair j's FMAs spills at 80 registers: 16.1 ms.
                // profiles/r01o_*, r01s_*)
#pragma unroll
                for (int t = 0; t < 4; ++t)
#pragma unroll
                    for (int m = 0; m < kL; ++m) acc[4 * q + t][m] = fmaf(cs[t], x[m], acc[4 * q + t][m]);
            }
        }
        __syncwarp();  // the buffer is refilled by the issue of the next iteration
    }

    // the warp is the only writer of its (patch, layers) in this launch.  `fresh`: the film was
    // cleared in this call and nothing has been added yet, so the rows are stored without reading
    // them back (no load latency at the end of the warp - with one eighth of the angles per rank
    // that wait was a third of the kernel, profiles/r01x2_*); otherwise read-add-write.  A patch row
    // is 16 aligned bytes when res_x is a multiple of 4.
    const int x0 = px * kPatch, y0 = py * kPatch;
    const size_t layer_elems = static_cast<size_t>(P.res[0]) * P.res[1];
    const bool vec = ((P.res[0] & 3) == 0) && ((reinterpret_cast<uintptr_t>(out) & 15) == 0) && (x0 + 3 < P.res[0]);
#pragma unroll
    for (int m = 0; m < kL; ++m) {
        const int z = __ldg(Z.zl + a_base + kL * lane + m);
        if (z < 0) continue;
        float* o = out + z * layer_elems;
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const int yy = y0 + r;
            if (yy >= P.res[1]) continue;
            float* row = o + static_cast<size_t>(yy) * P.res[0] + x0;
            if (vec) {
                float4 v = make_float4(acc[4 * r][m] * ray_weight, acc[4 * r + 1][m] * ray_weight,
                                       acc[4 * r + 2][m] * ray_weight, acc[4 * r + 3][m] * ray_weight);
                if (!fresh) {
                    if (v.x == 0.f && v.y == 0.f && v.z == 0.f && v.w == 0.f) continue;
                    const float4 old = *reinterpret_cast<const float4*>(row);
                    v.x += old.x; v.y += old.y; v.z += old.z; v.w += old.w;
                }
                *reinterpret_cast<float4*>(row) = v;
            } else {
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                    if (x0 + t >= P.res[0]) continue;
                    const float a = acc[4 * r + t][m] * ray_weight;
                    if (fresh) row[t] = a;
                    else if (a != 0.f) row[t] += a;
                }
            }
        }
    }
}

// ---- the patch gather on the tensor cores -------------------------------------------------------
// acc[16 voxels][layers] += coef[16 voxels][8 segments] . x[8 segments][layers], summed over the ~2 440
// segments that cross a patch, IS a matrix product (M = 16 voxels of the patch, K = the segments of a
// chunk, N = the layers, the rows of B gathered by ray index).  Here it runs as
// mma.sync.aligned.m16n8k8 tf32 (SASS HMMA: 276 dense TFLOP/s measured on this B200,
// profiles/r02g_microbench.json), with both operands split into a tf32 head and tail and three products
// per tile (hi.hi + lo.hi + hi.lo, error ~2^-20 per product) so that the 1e-4 bound holds for signed
// pattern vectors too.  Same staging as the FFMA kernel: TMA bulk copies on mbarriers, one chunk ahead.
//
// Fragment layout (PTX ISA, m16n8k8 .tf32): with g = lane / 4, q = lane % 4
//   A (16 x 8):  a0 = A[g][q]  a1 = A[g + 8][q]  a2 = A[g][q + 4]  a3 = A[g + 8][q + 4]
//   B ( 8 x 8):  b0 = B[q][g]  b1 = B[q + 4][g]
//   C (16 x 8):  c0 = C[g][2q] c1 = C[g][2q + 1] c2 = C[g + 8][2q] c3 = C[g + 8][2q + 1]
// Column n of n-tile t stands for layer NT * n + t of the block, so the NT values a lane needs of one
// pattern line are contiguous (LDS.128) and the lines sit 8 * NT + 4 floats apart (no bank conflicts).
__device__ __forceinline__ void split_tf32(float x, uint32_t& hi, uint32_t& lo) {
    hi = __float_as_uint(x) & 0xffffe000u;
    lo = __float_as_uint(x - __uint_as_float(hi));
}

__device__ __forceinline__ void mma_tf32(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

constexpr int gm_line(int NT) { return 8 * NT + 4; }   // floats between two staged pattern lines
constexpr int gm_buf_floats(int NT) { return kGChunk * gm_line(NT) + kGChunk * 16 + kGChunk; }
constexpr size_t gm_smem_bytes(int NT) {
    return (sizeof(float) * 2 * gm_buf_floats(NT) + 2 * sizeof(uint64_t)) * kGWarpsX * kGWarpsY;
}

// kTma = false (default since r02zt): the chunk is staged with 16-byte cp.async instead of bulk copies.  A bulk
// copy is a uniform-datapath instruction: the eight lanes that each fetch one gathered line are serialised by the
// compiler into an ELECT loop of 9 instructions per line, 26 % of the kernel's instructions and 30 % of its stall
// samples (profiles/r02zs_*); with LDGSTS one warp instruction moves a whole 512-byte line.
template <int NT, bool kTma>
__global__ void __launch_bounds__(32 * kGWarpsX * kGWarpsY, 2)
planarg_mma_kernel(const __grid_constant__ SceneParams P, const __grid_constant__ PlanarZPlan Z,
                   const float4* __restrict__ coef, const int* __restrict__ ray, const float* __restrict__ in,
                   float* __restrict__ out, const uint32_t* __restrict__ active_pixels, uint64_t n_active,
                   float ray_weight, const int* __restrict__ gate, int a_begin, int fresh) {
    extern __shared__ __align__(16) float g_smem[];
    if (*gate != 0) return;
    constexpr int kLayers = 8 * NT, kLine = gm_line(NT), kBuf = gm_buf_floats(NT);
    static_assert(kGChunk == 8, "one chunk is one k-step of the m16n8k8 product");
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int regions_x = (Z.tiles_x + kGWarpsX - 1) / kGWarpsX;
    const int px = (blockIdx.x % regions_x) * kGWarpsX + (warp % kGWarpsX);
    const int py = (blockIdx.x / regions_x) * kGWarpsY + (warp / kGWarpsX);
    if (px >= Z.tiles_x || py >= Z.tiles_y) return;
    const int tile_id = py * Z.tiles_x + px;
    // region fastest, z-block slowest: the CTAs in flight then share their pattern lines in L2.  The other
    // order (a region's z-blocks together, so that its coefficient stream is read from DRAM once) was
    // measured: 8.2 instead of 6.1 GB of DRAM reads and 6.55 instead of 6.09 ms (profiles/r02t_*)
    const int a_base = a_begin + blockIdx.y * kLayers;
    float* buf = g_smem + warp * (2 * kBuf);
    uint64_t* bars = reinterpret_cast<uint64_t*>(g_smem + kGWarpsX * kGWarpsY * 2 * kBuf) + 2 * warp;
    if (kTma && lane == 0) {
        mbar_init(bars, 1);
        mbar_init(bars + 1, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();

    int k0, nk;
    call_angles(P, active_pixels, n_active, k0, nk);
    const int* __restrict__ off = Z.tile_off + static_cast<size_t>(tile_id) * (P.n_patterns + 1);
    const int ebase = __ldg(Z.tile_ebase + tile_id);
    const int g0 = ebase + __ldg(off + k0), g1 = ebase + __ldg(off + k0 + nk);
    const float* __restrict__ xb = in + a_base - static_cast<long long>(k0) * P.res_x * Z.nz_pad;
    const int n_chunks = (g1 - g0 + kGChunk - 1) / kGChunk;

    float acc[NT][4];
#pragma unroll
    for (int t = 0; t < NT; ++t)
#pragma unroll
        for (int i = 0; i < 4; ++i) acc[t][i] = 0.f;

    auto load_rays = [&](int c) {
        const int g = g0 + c * kGChunk + (lane & (kGChunk - 1));
        return (c < n_chunks && g < g1) ? __ldg(ray + g) : 0;
    };
    // lane j < 8 copies the pattern line of pair j (8 * NT layers), lane 8 the coefficient blocks of the chunk
    auto issue = [&](int c, int rl) {
        float* b = buf + (c & 1) * kBuf;
        uint64_t* bar = bars + (c & 1);
        const int g = g0 + c * kGChunk;
        const int n_here = min(kGChunk, g1 - g);
        if constexpr (kTma) {
            const bool wants = lane < kGChunk && (rl >> 24) != 0;
            const unsigned lines = __popc(__ballot_sync(0xffffffffu, wants));
            if (lane == 0) mbar_expect_tx(bar, lines * (4u * kLayers) + 64u * n_here);
            __syncwarp();
            if (wants) bulk_g2s(b + lane * kLine, xb + static_cast<size_t>(rl & 0xffffff) * Z.nz_pad, 4u * kLayers, bar);
            if (lane == kGChunk) bulk_g2s(b + kGChunk * kLine, coef + 4 * static_cast<size_t>(g), 64u * n_here, bar);
        } else {
            // every lane has finished reading this buffer (chunk c - 2) before it is overwritten
            __syncwarp();
            constexpr int kPieces = kLayers / 4;          // 16-byte pieces of a line: 32, 16 or 8
            constexpr int kLinesPerPass = 32 / kPieces;   // lines one warp instruction covers
            const int piece = lane % kPieces;
#pragma unroll
            for (int p = 0; p < kGChunk / kLinesPerPass; ++p) {
                const int j = p * kLinesPerPass + lane / kPieces;
                const int rj = __shfl_sync(0xffffffffu, rl, j);   // every lane holds the ray word of pair lane & 7
                if ((rj >> 24) != 0)
                    cp_async16(b + j * kLine + 4 * piece, xb + static_cast<size_t>(rj & 0xffffff) * Z.nz_pad + 4 * piece);
            }
            if (lane < 4 * n_here)
                cp_async16(b + kGChunk * kLine + 4 * lane, reinterpret_cast<const float*>(coef + 4 * static_cast<size_t>(g)) + 4 * lane);
            cp_async_commit();
        }
        if (lane < kGChunk) reinterpret_cast<int*>(b)[kGChunk * kLine + kGChunk * 16 + lane] = rl;
    };

    const int gq = lane >> 2, q = lane & 3;
    int rl_next = load_rays(0);
    if (n_chunks > 0) issue(0, rl_next);
    rl_next = load_rays(1);
    for (int c = 0; c < n_chunks; ++c) {
        if (c + 1 < n_chunks) {
            issue(c + 1, rl_next);
            rl_next = load_rays(c + 2);
            if constexpr (!kTma) cp_async_wait<1>();
        } else if constexpr (!kTma) {
            cp_async_wait<0>();
        }
        if constexpr (kTma) mbar_wait(bars + (c & 1), (c >> 1) & 1);
        __syncwarp();
        const float* b = buf + (c & 1) * kBuf;
        const int n_here = min(kGChunk, g1 - (g0 + c * kGChunk));
        const int* rw = reinterpret_cast<const int*>(b) + kGChunk * kLine + kGChunk * 16;
        // pairs beyond the end of the list, and pairs that touch nothing (their line is not fetched),
        // enter with zero coefficients: whatever their stale line holds is multiplied by 0
        const bool v_lo = q < n_here && (rw[q] >> 24) != 0, v_hi = q + 4 < n_here && (rw[q + 4] >> 24) != 0;
        const float* cf = b + kGChunk * kLine;
        const float a_f[4] = {v_lo ? cf[16 * q + gq] : 0.f, v_lo ? cf[16 * q + gq + 8] : 0.f,
                              v_hi ? cf[16 * (q + 4) + gq] : 0.f, v_hi ? cf[16 * (q + 4) + gq + 8] : 0.f};
        uint32_t a_hi[4], a_lo[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) split_tf32(a_f[i], a_hi[i], a_lo[i]);
        const float* x_lo = b + q * kLine + gq * NT;          // layers NT * gq + t of pair q
        const float* x_hi = b + (q + 4) * kLine + gq * NT;    // ... of pair q + 4
        if constexpr (NT >= 4) {
#pragma unroll
            for (int t4 = 0; t4 < NT / 4; ++t4) {
                float4 xl = *reinterpret_cast<const float4*>(x_lo + 4 * t4);
                float4 xh = *reinterpret_cast<const float4*>(x_hi + 4 * t4);
                if (!v_lo) xl = make_float4(0.f, 0.f, 0.f, 0.f);   // a stale line may hold inf / nan
                if (!v_hi) xh = make_float4(0.f, 0.f, 0.f, 0.f);
                const float bl[4] = {xl.x, xl.y, xl.z, xl.w}, bh[4] = {xh.x, xh.y, xh.z, xh.w};
                // (issuing the three products of a tile four tiles apart, and a non-volatile asm that lets
                // ptxas interleave them, were measured: 6.50 ms against 6.08 ms for this order, r02j)
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    uint32_t b0h, b0l, b1h, b1l;
                    split_tf32(bl[i], b0h, b0l);
                    split_tf32(bh[i], b1h, b1l);
                    mma_tf32(acc[4 * t4 + i], a_hi, b0h, b1h);
                    mma_tf32(acc[4 * t4 + i], a_lo, b0h, b1h);
                    mma_tf32(acc[4 * t4 + i], a_hi, b0l, b1l);
                }
            }
        } else {
#pragma unroll
            for (int t = 0; t < NT; ++t) {
                const float xl = v_lo ? x_lo[t] : 0.f, xh = v_hi ? x_hi[t] : 0.f;
                uint32_t b0h, b0l, b1h, b1l;
                split_tf32(xl, b0h, b0l);
                split_tf32(xh, b1h, b1l);
                mma_tf32(acc[t], a_hi, b0h, b1h);
                mma_tf32(acc[t], a_lo, b0h, b1h);
                mma_tf32(acc[t], a_hi, b0l, b1l);
            }
        }
        __syncwarp();  // the buffer is refilled by the issue of the next iteration
    }

    // lane (g, q) holds voxels g and g + 8 of the patch (row g / 4 resp. g / 4 + 2, column g % 4) for the
    // layers NT * 2q + t and NT * (2q + 1) + t; the four lanes of consecutive g write 16 contiguous bytes
    const int x0 = px * kPatch, y0 = py * kPatch;
    const size_t layer_elems = static_cast<size_t>(P.res[0]) * P.res[1];
    const int xx = x0 + (gq & 3), yy0 = y0 + (gq >> 2), yy1 = yy0 + 2;
    if (xx < P.res[0]) {
#pragma unroll
        for (int t = 0; t < NT; ++t) {
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int z = __ldg(Z.zl + a_base + NT * (2 * q + h) + t);
                if (z < 0) continue;
                float* o = out + z * layer_elems + xx;
                if (yy0 < P.res[1]) {
                    const float v = acc[t][h] * ray_weight;
                    float* d = o + static_cast<size_t>(yy0) * P.res[0];
                    if (fresh) *d = v; else if (v != 0.f) *d += v;
                }
                if (yy1 < P.res[1]) {
                    const float v = acc[t][2 + h] * ray_weight;
                    float* d = o + static_cast<size_t>(yy1) * P.res[0];
                    if (fresh) *d = v; else if (v != 0.f) *d += v;
                }
            }
        }
    }
}

// ---- the patch gather in split fp16 --------------------------------------------------------------------------
// The 3xTF32 kernel above is bound by the tensor pipe (54 % busy at a third of its instructions; an m16n8k8 tf32
// HMMA occupies the pipe for ~8 cycles per SM sub-partition, profiles/r02zs_*).  fp16 runs at twice the rate, and a
// pair of fp16 numbers holds x to 22 bits once x is scaled into fp16's normal range - which both operands allow:
// the coefficient table is static (scaled once, planarg_build) and the patterns of a call are scaled by a power of
// two from their largest magnitude (absmax_kernel -> zpack_h16_kernel).  So the same three products
//     hi.hi + lo.hi + hi.lo     (the dropped lo.lo term is 2^-22 relative)
// run as mma.sync.m16n8k8 .f16 with fp32 accumulators, and both operands arrive ALREADY split ({hi, lo} packed in
// the 32 bits an fp32 value took, so tables and staging keep their size): a fragment register is one PRMT of two
// staged words instead of a mask, a subtraction and a select per value.
// Fragment layout (PTX ISA, m16n8k8 .f16): with g = lane / 4, q = lane % 4
//   A (16 x 8):  a0 = {A[g][2q], A[g][2q+1]}   a1 = {A[g+8][2q], A[g+8][2q+1]}
//   B ( 8 x 8):  b0 = {B[2q][g], B[2q+1][g]}
//   C (16 x 8):  as for tf32
__device__ __forceinline__ void mma_h16(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t b0) {
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a0), "r"(a1), "r"(b0));
}

__device__ __forceinline__ void mma_h16k16(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                           uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

constexpr bool kK16 = true;   // false: three m16n8k8 products per tile (measured slower)

template <int NT>
__global__ void __launch_bounds__(32 * kGWarpsX * kGWarpsY, 2)
planarg_h16_kernel(const __grid_constant__ SceneParams P, const __grid_constant__ PlanarZPlan Z,
                   const uint32_t* __restrict__ coef, const int* __restrict__ ray, const uint32_t* __restrict__ in,
                   float* __restrict__ out, const uint32_t* __restrict__ active_pixels, uint64_t n_active,
                   float ray_weight, const int* __restrict__ gate, int a_begin, int fresh,
                   const unsigned* __restrict__ pmax, float rows_max, float coef_inv_scale) {
    extern __shared__ __align__(16) float g_smem[];
    if (*gate != 0) return;
    constexpr int kLayers = 8 * NT, kLine = gm_line(NT), kBuf = gm_buf_floats(NT);
    static_assert(kGChunk == 8, "one chunk is one k-step of the m16n8k8 product");
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int regions_x = (Z.tiles_x + kGWarpsX - 1) / kGWarpsX;
    const int px = (blockIdx.x % regions_x) * kGWarpsX + (warp % kGWarpsX);
    const int py = (blockIdx.x / regions_x) * kGWarpsY + (warp / kGWarpsX);
    if (px >= Z.tiles_x || py >= Z.tiles_y) return;
    const int tile_id = py * Z.tiles_x + px;
    const int a_base = a_begin + blockIdx.y * kLayers;
    uint32_t* buf = reinterpret_cast<uint32_t*>(g_smem) + warp * (2 * kBuf);
    // a line that is never fetched (a pair that touches nothing) enters with zero coefficients: what it holds must
    // be finite, so the buffers start as zeros (afterwards they only ever hold staged fp16 pairs)
    for (int i = lane; i < 2 * kBuf; i += 32) buf[i] = 0u;
    __syncwarp();

    int k0, nk;
    call_angles(P, active_pixels, n_active, k0, nk);
    const int* __restrict__ off = Z.tile_off + static_cast<size_t>(tile_id) * (P.n_patterns + 1);
    const int ebase = __ldg(Z.tile_ebase + tile_id);
    const int g0 = ebase + __ldg(off + k0), g1 = ebase + __ldg(off + k0 + nk);
    const uint32_t* __restrict__ xb = in + a_base - static_cast<long long>(k0) * P.res_x * Z.nz_pad;
    const int n_chunks = (g1 - g0 + kGChunk - 1) / kGChunk;

    float acc[NT][4];
#pragma unroll
    for (int t = 0; t < NT; ++t)
#pragma unroll
        for (int i = 0; i < 4; ++i) acc[t][i] = 0.f;

    auto load_rays = [&](int c) {
        const int g = g0 + c * kGChunk + (lane & (kGChunk - 1));
        return (c < n_chunks && g < g1) ? __ldg(ray + g) : 0;
    };
    // Address arithmetic of the staging, hoisted: per line the loop pays a shuffle, a mask, one 32-bit multiply, a
    // compare, one wide add and the copy (the first version re-derived 64-bit addresses and re-read kernel
    // parameters for every line: 124 of the 269 instructions of a chunk were integer work, profiles/r02zw_*).
    constexpr int kPieces = kLayers / 4;          // 16-byte pieces of a line: 32, 16 or 8
    constexpr int kLinesPerPass = 32 / kPieces;   // lines one warp instruction covers
    const uint32_t nzp4 = static_cast<uint32_t>(Z.nz_pad) * 4u;                    // bytes between two pattern lines
    const char* const xsrc = reinterpret_cast<const char*>(xb) + 16 * (lane % kPieces);
    const uint32_t sdst0 = static_cast<uint32_t>(__cvta_generic_to_shared(buf)) + 4u * ((lane / kPieces) * kLine) + 16u * (lane % kPieces);
    const char* const csrc = reinterpret_cast<const char*>(coef + 16 * static_cast<size_t>(g0)) + 16 * lane;
    const uint32_t cdst0 = static_cast<uint32_t>(__cvta_generic_to_shared(buf)) + 4u * (kGChunk * kLine) + 16u * lane;
    const uint32_t rdst0 = static_cast<uint32_t>(__cvta_generic_to_shared(buf)) + 4u * (kGChunk * kLine + kGChunk * 16 + lane);
    auto issue = [&](int c, int rl) {
        const uint32_t bsel = (c & 1) ? 4u * kBuf : 0u;
        const int n_here = min(kGChunk, g1 - (g0 + c * kGChunk));
        __syncwarp();   // every lane has finished reading this buffer (chunk c - 2)
#pragma unroll
        for (int p = 0; p < kGChunk / kLinesPerPass; ++p) {
            const unsigned rj = static_cast<unsigned>(__shfl_sync(0xffffffffu, rl, p * kLinesPerPass + lane / kPieces));
            if (rj >= 0x1000000u)   // a non-empty row mask: the line is needed
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sdst0 + bsel + 4u * (p * kLinesPerPass * kLine)),
                             "l"(xsrc + static_cast<size_t>((rj & 0xffffffu) * nzp4))
                             : "memory");
        }
        if (lane < 4 * n_here)
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(cdst0 + bsel), "l"(csrc + static_cast<size_t>(c) * (64 * kGChunk))
                         : "memory");
        cp_async_commit();
        if (lane < kGChunk) asm volatile("st.shared.b32 [%0], %1;" ::"r"(rdst0 + bsel), "r"(rl) : "memory");
    };

    const int gq = lane >> 2, q = lane & 3;
    int rl_next = load_rays(0);
    if (n_chunks > 0) issue(0, rl_next);
    rl_next = load_rays(1);
    for (int c = 0; c < n_chunks; ++c) {
        if (c + 1 < n_chunks) {
            issue(c + 1, rl_next);
            rl_next = load_rays(c + 2);
            cp_async_wait<1>();
        } else {
            cp_async_wait<0>();
        }
        __syncwarp();
        const uint32_t* b = buf + (c & 1) * kBuf;
        const int n_here = min(kGChunk, g1 - (g0 + c * kGChunk));
        const int* rw = reinterpret_cast<const int*>(b) + kGChunk * kLine + kGChunk * 16;
        // this lane's two k indices are the pairs 2q and 2q + 1 of the chunk; pairs beyond the end of the list and
        // pairs that touch nothing (their line is not fetched) enter with zero coefficients
        const bool v0 = 2 * q < n_here && (rw[2 * q] >> 24) != 0, v1 = 2 * q + 1 < n_here && (rw[2 * q + 1] >> 24) != 0;
        const uint32_t* cf = b + kGChunk * kLine;
        const uint32_t u00 = v0 ? cf[16 * (2 * q) + gq] : 0u, u01 = v1 ? cf[16 * (2 * q + 1) + gq] : 0u;
        const uint32_t u10 = v0 ? cf[16 * (2 * q) + gq + 8] : 0u, u11 = v1 ? cf[16 * (2 * q + 1) + gq + 8] : 0u;
        const uint32_t a0h = __byte_perm(u00, u01, 0x5410), a0l = __byte_perm(u00, u01, 0x7632);
        const uint32_t a1h = __byte_perm(u10, u11, 0x5410), a1l = __byte_perm(u10, u11, 0x7632);
        const uint32_t* x0 = b + (2 * q) * kLine + gq * NT;       // layers NT * gq + t of pair 2q
        const uint32_t* x1 = b + (2 * q + 1) * kLine + gq * NT;   // ... of pair 2q + 1
#pragma unroll
        for (int t4 = 0; t4 < NT / 4; ++t4) {
            const uint4 xa = *reinterpret_cast<const uint4*>(x0 + 4 * t4);
            const uint4 xc = *reinterpret_cast<const uint4*>(x1 + 4 * t4);
            const uint32_t ba[4] = {xa.x, xa.y, xa.z, xa.w}, bc[4] = {xc.x, xc.y, xc.z, xc.w};
            // (product-major order over the four tiles - the three products of a tile four HMMAs apart - was
            // measured: no difference, r02zx)
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const uint32_t bh = __byte_perm(ba[i], bc[i], 0x5410), bl = __byte_perm(ba[i], bc[i], 0x7632);
                if (kK16) {
                    // [a_hi | a_lo] (16 x 16) times [b_hi ; b_hi], then times [b_lo ; b_lo]: all four products of
                    // the split operands in two m16n8k16 instructions (k = 0..7: the hi parts of the eight pairs,
                    // k = 8..15: their lo parts)
                    mma_h16k16(acc[4 * t4 + i], a0h, a1h, a0l, a1l, bh, bh);
                    mma_h16k16(acc[4 * t4 + i], a0h, a1h, a0l, a1l, bl, bl);
                } else {
                    mma_h16(acc[4 * t4 + i], a0h, a1h, bh);
                    mma_h16(acc[4 * t4 + i], a0l, a1l, bh);
                    mma_h16(acc[4 * t4 + i], a0h, a1h, bl);
                }
            }
        }
        __syncwarp();  // the buffer is refilled by the issue of the next iteration
    }

    // undo the two scalings (powers of two: exact)
    const float wscale = ray_weight * coef_inv_scale / h16_scale(__uint_as_float(__ldg(pmax)) * rows_max);
    // lane (g, q) holds voxels g and g + 8 of the patch (row g / 4 resp. g / 4 + 2, column g % 4) for the
    // layers NT * 2q + t and NT * (2q + 1) + t; the four lanes of consecutive g write 16 contiguous bytes
    const int x0v = px * kPatch, y0v = py * kPatch;
    const size_t layer_elems = static_cast<size_t>(P.res[0]) * P.res[1];
    const int xx = x0v + (gq & 3), yy0 = y0v + (gq >> 2), yy1 = yy0 + 2;
    if (xx < P.res[0]) {
#pragma unroll
        for (int t = 0; t < NT; ++t) {
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int z = __ldg(Z.zl + a_base + NT * (2 * q + h) + t);
                if (z < 0) continue;
                float* o = out + z * layer_elems + xx;
                if (yy0 < P.res[1]) {
                    const float v = acc[t][h] * wscale;
                    float* d = o + static_cast<size_t>(yy0) * P.res[0];
                    if (fresh) *d = v; else if (v != 0.f) *d += v;
                }
                if (yy1 < P.res[1]) {
                    const float v = acc[t][2 + h] * wscale;
                    float* d = o + static_cast<size_t>(yy1) * P.res[0];
                    if (fresh) *d = v; else if (v != 0.f) *d += v;
                }
            }
        }
    }
}

}  // namespace

#define PZ_TRY(expr)                      \
    do {                                  \
        cudaError_t _e = (expr);          \
        if (_e != cudaSuccess) return _e; \
    } while (0)

void planarz_release(PlanarZHost& H) {
    for (PlanarZSub& S : H.sub) {
        void* ptrs[] = {S.tile_off, S.tile_clo, S.tile_ebase, S.entries, S.vhead, S.vde, S.v_off, S.v_ebase};
        for (void* p : ptrs)
            if (p) cudaFree(p);
    }
    void* ptrs[] = {H.zl, H.a_row_begin, H.a_rows, H.row_a, H.layer_a, H.row_ri, H.scratch, H.pmax};
    for (void* p : ptrs)
        if (p) cudaFree(p);
    H = PlanarZHost();
}

// Tile geometry for kL layers per lane: the largest xy tile whose
// [h][w][33][kL] float slab fits the 227 KB of opt-in shared memory.
static void z_tile_dims(int rx, int ry, int kL, int& tw, int& th) {
    const int max_floats = (227 * 1024) / 4;
    const int area = max_floats / (kZLaneStride * kL);
    int side = static_cast<int>(std::floor(std::sqrt(static_cast<double>(area))));
    side = std::min(side, 32);   // a walk crosses at most w + h - 1 voxels: the visit lists of the adjoint hold 64
    int tw_max = std::max(4, side & ~3);
    int tiles_x = (rx + tw_max - 1) / tw_max;
    tw = (((rx + tiles_x - 1) / tiles_x) + 3) & ~3;
    int th_max = std::max(1, std::min(area / tw, 64 - tw));
    int tiles_y = (ry + th_max - 1) / th_max;
    th = (ry + tiles_y - 1) / tiles_y;
}

template <int kL>
static cudaError_t z_set_smem(size_t bytes) {
    PZ_TRY(cudaFuncSetAttribute(planarz_kernel<kScatter, kL>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                static_cast<int>(bytes)));
    PZ_TRY(cudaFuncSetAttribute(planarv_kernel<kL>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(bytes)));
    return cudaFuncSetAttribute(planarz_kernel<kGather, kL>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                static_cast<int>(bytes));
}

static cudaError_t z_build_sub(const SceneParams& P, const PlanarHost& base, const PlanarZHost& H, int kL, int a_begin,
                               int n_blocks, PlanarZSub& S) {
    const int np = P.n_patterns;
    PlanarZPlan& Z = S.plan;
    Z.seg_a = base.seg_a;
    Z.seg_b = base.seg_b;
    Z.seg_d = base.seg_d;
    z_tile_dims(P.res[0], P.res[1], kL, Z.tile_w, Z.tile_h);
    Z.tiles_x = (P.res[0] + Z.tile_w - 1) / Z.tile_w;
    Z.tiles_y = (P.res[1] + Z.tile_h - 1) / Z.tile_h;
    Z.nz = H.nz;
    Z.nz_pad = H.nz_pad;
    Z.lanes_l = kL;
    Z.a_begin = a_begin;
    Z.groups = 1;
    Z.zl = H.zl; Z.a_row_begin = H.a_row_begin; Z.a_rows = H.a_rows; Z.row_a = H.row_a;
    S.n_blocks = n_blocks;
    const int n_tiles = Z.tiles_x * Z.tiles_y;
    S.smem_bytes = sizeof(float) * static_cast<size_t>(kL) * Z.tile_h * Z.tile_w * kZLaneStride;
    PZ_TRY(cudaMalloc(&S.tile_off, sizeof(int) * n_tiles * (np + 1)));
    PZ_TRY(cudaMalloc(&S.tile_clo, sizeof(int) * n_tiles * np));
    PZ_TRY(cudaMalloc(&S.tile_ebase, sizeof(int) * (n_tiles + 1)));
    PZ_TRY(planar_tile_ranges(P, base.seg_a, base.seg_b, Z.tiles_x, Z.tiles_y, Z.tile_w, Z.tile_h, S.tile_off,
                              S.tile_clo));
    // entry base of every tile = prefix of the per-tile totals (last element of each tile's prefix row)
    std::vector<int> last(n_tiles), ebase(n_tiles + 1, 0);
    PZ_TRY(cudaMemcpy2D(last.data(), sizeof(int), S.tile_off + np, sizeof(int) * (np + 1), sizeof(int), n_tiles,
                        cudaMemcpyDeviceToHost));
    long long total = 0;
    for (int t = 0; t < n_tiles; ++t) { ebase[t] = static_cast<int>(total); total += last[t]; }
    if (total >= (1ll << 30)) return cudaErrorInvalidValue;
    ebase[n_tiles] = static_cast<int>(total);
    PZ_TRY(cudaMemcpy(S.tile_ebase, ebase.data(), sizeof(int) * (n_tiles + 1), cudaMemcpyHostToDevice));
    PZ_TRY(cudaMalloc(&S.entries, sizeof(float4) * 2 * static_cast<size_t>(std::max<long long>(total, 1))));
    Z.tile_off = S.tile_off; Z.tile_clo = S.tile_clo; Z.tile_ebase = S.tile_ebase; Z.entries = S.entries;
    if (total > 0) {
        zsetup_kernel<<<static_cast<unsigned>((total + 255) / 256), 256>>>(P, Z, S.entries, static_cast<int>(total));
        count_launch();
        PZ_TRY(cudaGetLastError());
    }
    // visit lists of the adjoint (left out when they would not fit comfortably, or on request)
    const char* env = std::getenv("DRTVAM_PLANAR_ADJOINT");
    if (total > 0 && !(env && std::strcmp(env, "dda") == 0)) {
        int *count = nullptr, *voff = nullptr, *flag = nullptr;
        void* tmp = nullptr;
        auto run = [&]() -> cudaError_t {
            const int n1 = static_cast<int>(total) + 1;
            PZ_TRY(cudaMalloc(&count, sizeof(int) * n1));
            PZ_TRY(cudaMalloc(&voff, sizeof(int) * n1));
            PZ_TRY(cudaMalloc(&flag, sizeof(int)));
            PZ_TRY(cudaMemset(flag, 0, sizeof(int)));
            vcount_kernel<<<(n1 + 255) / 256, 256>>>(S.entries, static_cast<int>(total), count);
            count_launch();
            size_t tmp_bytes = 0;
            PZ_TRY(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, count, voff, n1));
            PZ_TRY(cudaMalloc(&tmp, tmp_bytes));
            PZ_TRY(cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, count, voff, n1));
            count_launch();
            int visits = 0;
            PZ_TRY(cudaMemcpy(&visits, voff + total, sizeof(int), cudaMemcpyDeviceToHost));
            size_t free_b = 0, total_b = 0;
            PZ_TRY(cudaMemGetInfo(&free_b, &total_b));
            if (visits <= 0 || (sizeof(float) * static_cast<size_t>(visits) + sizeof(float4) * total) > free_b / 4) return cudaSuccess;
            PZ_TRY(cudaMalloc(&S.vhead, sizeof(float4) * static_cast<size_t>(total)));
            PZ_TRY(cudaMalloc(&S.vde, sizeof(float) * static_cast<size_t>(visits)));
            PZ_TRY(cudaMemset(S.vde, 0, sizeof(float) * static_cast<size_t>(visits)));
            vfill_kernel<<<static_cast<unsigned>((total + 255) / 256), 256>>>(S.entries, voff, static_cast<int>(total), Z.tile_w,
                                                                              S.vhead, S.vde, flag);
            count_launch();
            PZ_TRY(cudaGetLastError());
            int too_long = 0;
            PZ_TRY(cudaMemcpy(&too_long, flag, sizeof(int), cudaMemcpyDeviceToHost));
            if (too_long) {   // a walk of more than kVisitMax voxels: this sub-plan keeps the DDA kernel
                cudaFree(S.vhead); cudaFree(S.vde);
                S.vhead = nullptr; S.vde = nullptr;
                return cudaSuccess;
            }
            // compact the heads to the entries with visits (`count` becomes the keep flags, `voff` their positions)
            vkeep_kernel<<<(n1 + 255) / 256, 256>>>(count, static_cast<int>(total), count);
            count_launch();
            PZ_TRY(cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, count, voff, n1));
            count_launch();
            int kept = 0;
            PZ_TRY(cudaMemcpy(&kept, voff + total, sizeof(int), cudaMemcpyDeviceToHost));
            float4* vhead_c = nullptr;
            PZ_TRY(cudaMalloc(&vhead_c, sizeof(float4) * static_cast<size_t>(std::max(kept, 1))));
            vcompact_kernel<<<static_cast<unsigned>((total + 255) / 256), 256>>>(P, Z, count, voff, static_cast<int>(total),
                                                                                 S.vhead, vhead_c);
            count_launch();
            PZ_TRY(cudaMalloc(&S.v_off, sizeof(int) * static_cast<size_t>(n_tiles) * (np + 1)));
            PZ_TRY(cudaMalloc(&S.v_ebase, sizeof(int) * n_tiles));
            const long long cells = static_cast<long long>(n_tiles) * (np + 1);
            vremap_kernel<<<static_cast<unsigned>((cells + 255) / 256), 256>>>(P, Z, voff, S.v_off, S.v_ebase);
            count_launch();
            PZ_TRY(cudaDeviceSynchronize());
            cudaFree(S.vhead);
            S.vhead = vhead_c;
            if (std::getenv("DRTVAM_DEBUG"))
                fprintf(stderr, "[drtvam] visit lists (kL %d): %d of %lld entries kept, %d visits\n", kL, kept, total, visits);
            return cudaSuccess;
        };
        const cudaError_t ve = run();
        void* ptrs[] = {count, voff, flag, tmp};
        for (void* q : ptrs)
            if (q) cudaFree(q);
        if (ve != cudaSuccess) {
            cudaGetLastError();
            if (S.vhead) cudaFree(S.vhead);
            if (S.vde) cudaFree(S.vde);
            S.vhead = nullptr; S.vde = nullptr;
        }
        Z.vhead = S.vhead; Z.vde = S.vde; Z.v_off = S.v_off; Z.v_ebase = S.v_ebase;
    }
    if (kL == 1) return z_set_smem<1>(S.smem_bytes);
    if (kL == 2) return z_set_smem<2>(S.smem_bytes);
    return z_set_smem<4>(S.smem_bytes);
}

cudaError_t planarz_build(const SceneParams& P, const PlanarHost& base, const std::vector<int>& row_layer,
                          PlanarZHost& H, bool walks) {
    planarz_release(H);
    int max_l = 4;
    if (const char* env = std::getenv("DRTVAM_Z_LAYERS")) {
        const int v = std::atoi(env);
        if (v == 1 || v == 2 || v == 4) max_l = v;
    }
    const int rz = P.res[2], Hh = P.res_y;
    // active layers and their rows
    std::vector<int> zl, a_row_begin(1, 0), a_rows, row_a(Hh, -1), layer_a(rz, -1), row_ri(Hh, 0);
    std::vector<std::vector<int>> rows_of(rz);
    // only rows of the call's row range and layers of its z window (drtvam_scene_set_z_window) take part;
    // zl holds the layer index inside the window, i.e. inside the film buffers of a call
    for (int r = P.row_begin; r < P.row_begin + P.n_rows && r < Hh; ++r)
        if (row_layer[r] >= P.z_begin && row_layer[r] < P.z_end) rows_of[row_layer[r]].push_back(r);
    // jittered passes (walks == false): a lane of row r lands anywhere in the row's height, which can
    // reach layers that hold no row CENTRE (e.g. 0.05 mm pixels over 0.025 mm layers); those layers are
    // active too, with no rows of their own
    std::vector<char> reach(rz, 0);
    if (!walks && base.row_z.size() == static_cast<size_t>(Hh)) {
        const float vs = (P.bmax[2] - P.bmin[2]) / static_cast<float>(rz);
        const float dz = Hh > 1 ? std::fabs(base.row_z[1] - base.row_z[0]) : P.pixel_y;
        for (int r = P.row_begin; r < P.row_begin + P.n_rows && r < Hh; ++r) {
            if (row_layer[r] < P.z_begin || row_layer[r] >= P.z_end) continue;
            const float lo = (base.row_z[r] - 0.5f * dz - P.bmin[2]) / vs, hi = (base.row_z[r] + 0.5f * dz - P.bmin[2]) / vs;
            const int la = std::max(static_cast<int>(std::floor(lo - 1e-3f)), P.z_begin);
            const int lb = std::min(static_cast<int>(std::floor(hi + 1e-3f)), P.z_end - 1);
            for (int z = la; z <= lb; ++z) reach[z] = 1;
        }
    }
    for (int z = 0; z < rz; ++z) {
        if (rows_of[z].empty() && !reach[z]) continue;
        int ri = 0;
        for (int r : rows_of[z]) { a_rows.push_back(r); row_a[r] = static_cast<int>(zl.size()); row_ri[r] = ri++; }
        layer_a[z] = static_cast<int>(zl.size());
        zl.push_back(z - P.z_begin);
        a_row_begin.push_back(static_cast<int>(a_rows.size()));
    }
    const int nz = static_cast<int>(zl.size());
    if (nz == 0) return cudaSuccess;  // nothing deposits: leave the plan invalid
    const int units = (nz + 31) / 32;  // warps' worth of layers
    H.nz = nz;
    H.nz_pad = units * 32;
    zl.resize(H.nz_pad, -1);
    if (a_rows.empty()) a_rows.push_back(0);
    PZ_TRY(cudaMalloc(&H.zl, sizeof(int) * H.nz_pad));
    PZ_TRY(cudaMalloc(&H.a_row_begin, sizeof(int) * (nz + 1)));
    PZ_TRY(cudaMalloc(&H.a_rows, sizeof(int) * a_rows.size()));
    PZ_TRY(cudaMalloc(&H.row_a, sizeof(int) * Hh));
    PZ_TRY(cudaMalloc(&H.layer_a, sizeof(int) * rz));
    PZ_TRY(cudaMalloc(&H.row_ri, sizeof(int) * Hh));
    PZ_TRY(cudaMemcpy(H.layer_a, layer_a.data(), sizeof(int) * rz, cudaMemcpyHostToDevice));
    PZ_TRY(cudaMemcpy(H.row_ri, row_ri.data(), sizeof(int) * Hh, cudaMemcpyHostToDevice));
    PZ_TRY(cudaMemcpy(H.zl, zl.data(), sizeof(int) * H.nz_pad, cudaMemcpyHostToDevice));
    PZ_TRY(cudaMemcpy(H.a_row_begin, a_row_begin.data(), sizeof(int) * (nz + 1), cudaMemcpyHostToDevice));
    PZ_TRY(cudaMemcpy(H.a_rows, a_rows.data(), sizeof(int) * a_rows.size(), cudaMemcpyHostToDevice));
    PZ_TRY(cudaMemcpy(H.row_a, row_a.data(), sizeof(int) * Hh, cudaMemcpyHostToDevice));

    H.rows_max = 0;
    for (int a = 0; a < nz; ++a) H.rows_max = std::max(H.rows_max, a_row_begin[a + 1] - a_row_begin[a]);
    // split the layer units into runs of blocks with 4, 2 and 1 layers per lane (no padded blocks);
    // the jittered kernels (planarj_kernels.cu) only need the layer tables above
    int done = walks ? 0 : units;
    for (int kL = max_l; kL >= 1 && done < units; kL >>= 1) {
        const int n_blocks = (units - done) / kL;
        if (n_blocks == 0) continue;
        PZ_TRY(z_build_sub(P, base, H, kL, done * 32, n_blocks, H.sub[H.n_sub]));
        ++H.n_sub;
        done += n_blocks * kL;
    }
    PZ_TRY(cudaDeviceSynchronize());
    H.valid = true;
    return cudaSuccess;
}

static cudaError_t z_scratch(PlanarZHost& H, size_t bytes) {
    if (bytes <= H.scratch_bytes) return cudaSuccess;
    if (H.scratch) cudaFree(H.scratch);
    H.scratch = nullptr;
    H.scratch_bytes = 0;
    PZ_TRY(cudaMalloc(&H.scratch, bytes));
    H.scratch_bytes = bytes;
    return cudaSuccess;
}

static cudaError_t launch_zmask(const SceneParams& P, PlanarZHost& H, const uint32_t* mask, const uint32_t* active_pixels,
                                uint64_t n_active, const int* gate, cudaStream_t stream) {
    const size_t nk = n_active / (static_cast<uint64_t>(P.res_x) * P.n_rows);
    const size_t words = ((static_cast<size_t>(P.res_x) * H.nz_pad) >> 5) * nk;
    zmask_kernel<<<static_cast<unsigned>((words + 255) / 256), 256, 0, stream>>>(P, H.sub[0].plan, H.scratch, mask,
                                                                              active_pixels, n_active, gate);
    count_launch();
    return cudaGetLastError();
}

// Angle groups of one sub-plan: enough CTAs for two waves of the 148 SMs.  The gather (adjoint) kernel runs one
// 1024-thread CTA per SM, so a grid just past a multiple of 148 pays a whole extra wave (the 1-layer tail of C2:
// 100 tiles x 3 groups = 300 CTAs = 3 waves for 2.03 waves of work); there the group count is raised to the one
// with the fewest waves per group.  The scatter kernel keeps the plain rule (its group count fixes the order of
// its atomic adds).
static int z_groups(const PlanarZSub& S, int nk, bool gather) {
    const long long ctas = static_cast<long long>(S.plan.tiles_x) * S.plan.tiles_y * S.n_blocks;
    const long long g_max = std::min<long long>(64, std::max(1, nk / 8));
    long long g = (148 * 2 + ctas - 1) / std::max<long long>(ctas, 1);
    g = std::max<long long>(1, std::min(g, g_max));
    if (!gather) return static_cast<int>(g);
    auto waves_per_group = [&](long long n) { return static_cast<double>((ctas * n + 147) / 148) / n; };
    const double ideal = static_cast<double>(ctas) / 148.0;
    if (waves_per_group(g) <= 1.15 * ideal) return static_cast<int>(g);
    long long best = g;
    for (long long n = g + 1; n <= std::min<long long>(g_max, 16); ++n)
        if (waves_per_group(n) < 0.96 * waves_per_group(best)) best = n;
    return static_cast<int>(best);
}

template <int kMode>
static cudaError_t z_launch(const SceneParams& P, PlanarZHost& H, const float* in, float* out,
                            const uint32_t* active_pixels, uint64_t n_active, float w, const int* gate,
                            cudaStream_t stream) {
    const int nk = static_cast<int>(n_active / (static_cast<uint64_t>(P.res_x) * P.n_rows));
    for (int i = 0; i < H.n_sub; ++i) {
        const PlanarZSub& S = H.sub[i];
        PlanarZPlan Z = S.plan;
        const int groups = z_groups(S, nk, kMode == kGather);
        Z.groups = groups;
        const dim3 grid(Z.tiles_x * Z.tiles_y, S.n_blocks, groups);
        if (kMode == kGather && Z.vhead) {   // the visit-list adjoint
            if (Z.lanes_l == 1)
                planarv_kernel<1><<<grid, kZThreads, S.smem_bytes, stream>>>(P, Z, in, out, active_pixels, n_active, w, gate);
            else if (Z.lanes_l == 2)
                planarv_kernel<2><<<grid, kZThreads, S.smem_bytes, stream>>>(P, Z, in, out, active_pixels, n_active, w, gate);
            else
                planarv_kernel<4><<<grid, kZThreads, S.smem_bytes, stream>>>(P, Z, in, out, active_pixels, n_active, w, gate);
        } else if (Z.lanes_l == 1)
            planarz_kernel<kMode, 1><<<grid, kZThreads, S.smem_bytes, stream>>>(P, Z, in, out, active_pixels, n_active,
                                                                               w, gate);
        else if (Z.lanes_l == 2)
            planarz_kernel<kMode, 2><<<grid, kZThreads, S.smem_bytes, stream>>>(P, Z, in, out, active_pixels, n_active,
                                                                               w, gate);
        else
            planarz_kernel<kMode, 4><<<grid, kZThreads, S.smem_bytes, stream>>>(P, Z, in, out, active_pixels, n_active,
                                                                               w, gate);
        count_launch();
        PZ_TRY(cudaGetLastError());
    }
    return cudaSuccess;
}

cudaError_t launch_planarz_forward(const SceneParams& P, PlanarZHost& H, const int* gate, const float* active_data,
                                   const uint32_t* active_pixels, uint64_t n_active, float ray_weight, float* film,
                                   cudaStream_t stream) {
    const int nk = static_cast<int>(n_active / (static_cast<uint64_t>(P.res_x) * P.n_rows));
    const PlanarZPlan& Z0 = H.sub[0].plan;
    PZ_TRY(z_scratch(H, sizeof(float) * static_cast<size_t>(nk) * P.res_x * H.nz_pad));
    const dim3 pgrid((P.res_x + 31) / 32, H.nz_pad / 32, nk);
    zpack_kernel<<<pgrid, 256, 0, stream>>>(P, Z0, active_data, H.scratch, nk, gate);
    count_launch();
    PZ_TRY(cudaGetLastError());
    return z_launch<kScatter>(P, H, H.scratch, film, active_pixels, n_active, ray_weight, gate, stream);
}

cudaError_t launch_planarz_adjoint(const SceneParams& P, PlanarZHost& H, const int* gate, const float* dL,
                                   const uint32_t* active_pixels, uint64_t n_active, float ray_weight, float* grad,
                                   cudaStream_t stream, InsideHost* inside) {
    const int nk = static_cast<int>(n_active / (static_cast<uint64_t>(P.res_x) * P.n_rows));
    const size_t bytes = sizeof(float) * static_cast<size_t>(nk) * P.res_x * H.nz_pad;
    PZ_TRY(z_scratch(H, bytes));
    PZ_TRY(cudaMemsetAsync(H.scratch, 0, bytes, stream));
    PZ_TRY(z_launch<kGather>(P, H, dL, H.scratch, active_pixels, n_active, ray_weight, gate, stream));
    if (inside && inside->tiles) {
        PZ_TRY(launch_zmask(P, H, inside->mask, active_pixels, n_active, gate, stream));
        PZ_TRY(inside_adjoint_tiles(P, *inside, H, gate, active_pixels, n_active, ray_weight, H.scratch, stream));
    }
    const dim3 ugrid((P.res_x + 31) / 32, (P.n_rows + 31) / 32, nk);
    zunpack_kernel<<<ugrid, 256, 0, stream>>>(P, H.sub[0].plan, H.scratch, grad, gate);
    count_launch();
    return cudaGetLastError();
}

template <int kL>
static cudaError_t g_set_smem() {
    PZ_TRY((cudaFuncSetAttribute(planarg_mma_kernel<4 * kL, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 static_cast<int>(gm_smem_bytes(4 * kL)))));
    PZ_TRY((cudaFuncSetAttribute(planarg_mma_kernel<4 * kL, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 static_cast<int>(gm_smem_bytes(4 * kL)))));
    PZ_TRY(cudaFuncSetAttribute(planarg_h16_kernel<4 * kL>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                static_cast<int>(gm_smem_bytes(4 * kL))));
    return cudaFuncSetAttribute(planarg_kernel<kL>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                static_cast<int>(g_smem_bytes(kL)));
}

void planarg_release(PlanarGHost& G) {
    void* ptrs[] = {G.tile_off, G.tile_clo, G.tile_ebase, G.coef, G.ray, G.coef_h};
    for (void* p : ptrs)
        if (p) cudaFree(p);
    G = PlanarGHost();
}

// Builds the coefficient blocks of the patch-gather forward.  Leaves G.valid = false (no error) when
// the tables would not fit comfortably: the scatter kernel then serves the forward pass.
cudaError_t planarg_build(const SceneParams& P, const PlanarHost& base, const PlanarZHost& Zh, PlanarGHost& G) {
    planarg_release(G);
    if (const char* env = std::getenv("DRTVAM_PLANAR_FORWARD"))
        if (std::strcmp(env, "scatter") == 0) return cudaSuccess;
    if (!Zh.valid) return cudaSuccess;
    const int np = P.n_patterns;
    PlanarZPlan& Z = G.plan;
    Z = Zh.sub[0].plan;
    Z.seg_a = base.seg_a;
    Z.seg_b = base.seg_b;
    Z.seg_d = base.seg_d;
    Z.tile_w = Z.tile_h = kPatch;
    Z.tiles_x = (P.res[0] + kPatch - 1) / kPatch;
    Z.tiles_y = (P.res[1] + kPatch - 1) / kPatch;
    Z.entries = nullptr;
    const int n_tiles = Z.tiles_x * Z.tiles_y;
    if (static_cast<long long>(n_tiles) * (np + 1) >= (1ll << 29)) return cudaSuccess;
    if (static_cast<long long>(np) * P.res_x >= (1ll << 24)) return cudaSuccess;  // ray index + row mask share a word
    if (g_set_smem<1>() != cudaSuccess || g_set_smem<2>() != cudaSuccess || g_set_smem<4>() != cudaSuccess) {
        cudaGetLastError();
        return cudaSuccess;
    }
    PZ_TRY(cudaMalloc(&G.tile_off, sizeof(int) * static_cast<size_t>(n_tiles) * (np + 1)));
    PZ_TRY(cudaMalloc(&G.tile_clo, sizeof(int) * static_cast<size_t>(n_tiles) * np));
    PZ_TRY(cudaMalloc(&G.tile_ebase, sizeof(int) * (n_tiles + 1)));
    PZ_TRY(planar_tile_ranges(P, base.seg_a, base.seg_b, Z.tiles_x, Z.tiles_y, kPatch, kPatch, G.tile_off, G.tile_clo));
    std::vector<int> last(n_tiles), ebase(n_tiles + 1, 0);
    PZ_TRY(cudaMemcpy2D(last.data(), sizeof(int), G.tile_off + np, sizeof(int) * (np + 1), sizeof(int), n_tiles,
                        cudaMemcpyDeviceToHost));
    long long total = 0;
    for (int t = 0; t < n_tiles; ++t) { ebase[t] = static_cast<int>(total); total += last[t]; }
    ebase[n_tiles] = static_cast<int>(std::min<long long>(total, INT32_MAX));
    size_t free_b = 0, total_b = 0;
    PZ_TRY(cudaMemGetInfo(&free_b, &total_b));
    const unsigned long long need = static_cast<unsigned long long>(total) * (sizeof(float4) * 4 + sizeof(int));
    if (total <= 0 || total >= (1ll << 31) - 1 || need > free_b / 4) {  // too large: keep the scatter kernel
        planarg_release(G);
        return cudaSuccess;
    }
    PZ_TRY(cudaMemcpy(G.tile_ebase, ebase.data(), sizeof(int) * (n_tiles + 1), cudaMemcpyHostToDevice));
    PZ_TRY(cudaMalloc(&G.coef, sizeof(float4) * 4 * static_cast<size_t>(total)));
    PZ_TRY(cudaMalloc(&G.ray, sizeof(int) * static_cast<size_t>(total)));
    Z.tile_off = G.tile_off; Z.tile_clo = G.tile_clo; Z.tile_ebase = G.tile_ebase;
    gsetup_kernel<<<static_cast<unsigned>((total + 255) / 256), 256>>>(P, Z, G.coef, G.ray, static_cast<int>(total));
    count_launch();
    PZ_TRY(cudaGetLastError());
    // compaction (see gkeep_kernel)
    {
        int *keep = nullptr, *pos = nullptr, *ebase_new = nullptr, *off_new = nullptr, *ray_new = nullptr;
        float4* coef_new = nullptr;
        void* tmp = nullptr;
        auto run = [&]() -> cudaError_t {
            const int n1 = static_cast<int>(total) + 1;
            PZ_TRY(cudaMalloc(&keep, sizeof(int) * n1));
            PZ_TRY(cudaMalloc(&pos, sizeof(int) * n1));
            gkeep_kernel<<<(n1 + 255) / 256, 256>>>(G.ray, static_cast<int>(total), keep);
            count_launch();
            size_t tmp_bytes = 0;
            PZ_TRY(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, keep, pos, n1));
            PZ_TRY(cudaMalloc(&tmp, tmp_bytes));
            PZ_TRY(cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, keep, pos, n1));
            count_launch();
            int kept = 0;
            PZ_TRY(cudaMemcpy(&kept, pos + total, sizeof(int), cudaMemcpyDeviceToHost));
            if (kept <= 0 || kept > total - total / 16) return cudaSuccess;  // nothing worth dropping
            PZ_TRY(cudaMalloc(&coef_new, sizeof(float4) * 4 * static_cast<size_t>(kept)));
            PZ_TRY(cudaMalloc(&ray_new, sizeof(int) * static_cast<size_t>(kept)));
            PZ_TRY(cudaMalloc(&ebase_new, sizeof(int) * (n_tiles + 1)));
            PZ_TRY(cudaMalloc(&off_new, sizeof(int) * static_cast<size_t>(n_tiles) * (np + 1)));
            gcompact_kernel<<<static_cast<unsigned>((total + 255) / 256), 256>>>(G.coef, G.ray, pos,
                                                                                 static_cast<int>(total), coef_new, ray_new);
            count_launch();
            const long long slots = static_cast<long long>(n_tiles) * (np + 1);
            gremap_kernel<<<static_cast<unsigned>((slots + 255) / 256), 256>>>(n_tiles, np, G.tile_ebase, G.tile_off, pos,
                                                                               ebase_new, off_new);
            count_launch();
            PZ_TRY(cudaGetLastError());
            PZ_TRY(cudaDeviceSynchronize());
            std::swap(G.coef, coef_new);
            std::swap(G.ray, ray_new);
            std::swap(G.tile_ebase, ebase_new);
            std::swap(G.tile_off, off_new);
            Z.tile_off = G.tile_off; Z.tile_ebase = G.tile_ebase;
            G.candidates = total;
            total = kept;
            return cudaSuccess;
        };
        const cudaError_t ce = run();
        void* ptrs[] = {keep, pos, ebase_new, off_new, ray_new, coef_new, tmp};
        for (void* q : ptrs)
            if (q) cudaFree(q);
        if (ce != cudaSuccess) return ce;
    }
    PZ_TRY(cudaDeviceSynchronize());
    G.pairs = total;
    G.valid = true;
    // the same table as {hi, lo} fp16 pairs of the coefficients scaled into fp16's range (planarg_h16_kernel);
    // left out (the tf32 kernel then serves) when it does not fit comfortably
    if (total > 0) {
        size_t free_b = 0, total_b = 0;
        PZ_TRY(cudaMemGetInfo(&free_b, &total_b));
        const size_t n = 16 * static_cast<size_t>(total);
        unsigned* cmax = nullptr;
        if (sizeof(uint32_t) * n < free_b / 4 && cudaMalloc(&cmax, sizeof(unsigned)) == cudaSuccess) {
            unsigned bits = 0;
            cudaError_t e = cudaMemset(cmax, 0, sizeof(unsigned));
            if (e == cudaSuccess) {
                absmax_kernel<<<148 * 8, 256>>>(reinterpret_cast<const float*>(G.coef), n, cmax, nullptr);
                count_launch();
                e = cudaMemcpy(&bits, cmax, sizeof(bits), cudaMemcpyDeviceToHost);
            }
            cudaFree(cmax);
            float cm;
            std::memcpy(&cm, &bits, sizeof(cm));
            if (e == cudaSuccess && cm > 0.f && cudaMalloc(&G.coef_h, sizeof(uint32_t) * n) == cudaSuccess) {
                const float sc = h16_scale(cm);
                gcoef_h16_kernel<<<static_cast<unsigned>((n + 255) / 256), 256>>>(reinterpret_cast<const float*>(G.coef), n, sc,
                                                                                  G.coef_h);
                count_launch();
                G.coef_inv_scale = 1.f / sc;
                PZ_TRY(cudaDeviceSynchronize());
            } else {
                cudaGetLastError();
            }
        }
    }
    return cudaSuccess;
}

// DRTVAM_PLANAR_STAGING=tma stages the chunks of the tensor-core patch gather with bulk copies (A/B measurements)
static const bool mma_tma = std::getenv("DRTVAM_PLANAR_STAGING") && std::strcmp(std::getenv("DRTVAM_PLANAR_STAGING"), "tma") == 0;

template <int NT, class... Args>
static void launch_mma(bool tma, dim3 grid, int threads, cudaStream_t stream, Args... args) {
    if (tma) planarg_mma_kernel<NT, true><<<grid, threads, gm_smem_bytes(NT), stream>>>(args...);
    else planarg_mma_kernel<NT, false><<<grid, threads, gm_smem_bytes(NT), stream>>>(args...);
}

cudaError_t launch_planarg_forward(const SceneParams& P, PlanarZHost& H, PlanarGHost& G, const int* gate,
                                   const float* active_data, const uint32_t* active_pixels, uint64_t n_active,
                                   float ray_weight, float* film, bool fresh, cudaStream_t stream,
                                   const uint32_t* lane_mask) {
    const int nk = static_cast<int>(n_active / (static_cast<uint64_t>(P.res_x) * P.n_rows));
    PZ_TRY(z_scratch(H, sizeof(float) * static_cast<size_t>(nk) * P.res_x * H.nz_pad));
    const dim3 pgrid((P.res_x + 31) / 32, H.nz_pad / 32, nk);
    const PlanarZPlan& Z = G.plan;
    const int regions = ((Z.tiles_x + kGWarpsX - 1) / kGWarpsX) * ((Z.tiles_y + kGWarpsY - 1) / kGWarpsY);
    constexpr int threads = 32 * kGWarpsX * kGWarpsY;
    // DRTVAM_PLANAR_FORWARD = tf32 | ffma keep the 3xTF32 tensor-core kernel / the FMA-pipe kernel (A/B measurements)
    static const bool use_h16 = std::getenv("DRTVAM_PLANAR_FORWARD") == nullptr;
    if (use_h16 && G.coef_h != nullptr) {
        if (H.pmax == nullptr) PZ_TRY(cudaMalloc(&H.pmax, sizeof(unsigned)));
        const float rows_max = static_cast<float>(std::max(1, H.rows_max));
        PZ_TRY(cudaMemsetAsync(H.pmax, 0, sizeof(unsigned), stream));
        absmax_kernel<<<148 * 8, 256, 0, stream>>>(active_data, n_active, H.pmax, gate);
        count_launch();
        zpack_h16_kernel<<<pgrid, 256, 0, stream>>>(P, H.sub[0].plan, active_data, reinterpret_cast<uint32_t*>(H.scratch), H.pmax,
                                                    rows_max, gate);
        count_launch();
        PZ_TRY(cudaGetLastError());
        if (lane_mask) PZ_TRY(launch_zmask(P, H, lane_mask, active_pixels, n_active, gate, stream));
        const uint32_t* xin = reinterpret_cast<const uint32_t*>(H.scratch);
        for (int i = 0; i < H.n_sub; ++i) {
            const PlanarZSub& S = H.sub[i];
            const dim3 grid(regions, S.n_blocks);
            const int kL = S.plan.lanes_l, a_begin = S.plan.a_begin;
            if (kL == 1)
                planarg_h16_kernel<4><<<grid, threads, gm_smem_bytes(4), stream>>>(P, Z, G.coef_h, G.ray, xin, film, active_pixels,
                                                                                  n_active, ray_weight, gate, a_begin, fresh ? 1 : 0,
                                                                                  H.pmax, rows_max, G.coef_inv_scale);
            else if (kL == 2)
                planarg_h16_kernel<8><<<grid, threads, gm_smem_bytes(8), stream>>>(P, Z, G.coef_h, G.ray, xin, film, active_pixels,
                                                                                  n_active, ray_weight, gate, a_begin, fresh ? 1 : 0,
                                                                                  H.pmax, rows_max, G.coef_inv_scale);
            else
                planarg_h16_kernel<16><<<grid, threads, gm_smem_bytes(16), stream>>>(P, Z, G.coef_h, G.ray, xin, film, active_pixels,
                                                                                    n_active, ray_weight, gate, a_begin, fresh ? 1 : 0,
                                                                                    H.pmax, rows_max, G.coef_inv_scale);
            count_launch();
            PZ_TRY(cudaGetLastError());
        }
        return cudaSuccess;
    }
    zpack_kernel<<<pgrid, 256, 0, stream>>>(P, H.sub[0].plan, active_data, H.scratch, nk, gate);
    count_launch();
    PZ_TRY(cudaGetLastError());
    if (lane_mask) PZ_TRY(launch_zmask(P, H, lane_mask, active_pixels, n_active, gate, stream));
    // DRTVAM_PLANAR_FORWARD=ffma keeps the products on the FMA pipe (A/B measurements)
    static const bool use_mma = !(std::getenv("DRTVAM_PLANAR_FORWARD") && std::strcmp(std::getenv("DRTVAM_PLANAR_FORWARD"), "ffma") == 0);
    for (int i = 0; i < H.n_sub; ++i) {
        const PlanarZSub& S = H.sub[i];
        const dim3 grid(regions, S.n_blocks);
        const int kL = S.plan.lanes_l, a_begin = S.plan.a_begin;
        if (use_mma) {
            if (kL == 1)
                launch_mma<4>(mma_tma, grid, threads, stream, P, Z, G.coef, G.ray, H.scratch, film,
                                                                                  active_pixels, n_active, ray_weight, gate, a_begin, fresh ? 1 : 0);
            else if (kL == 2)
                launch_mma<8>(mma_tma, grid, threads, stream, P, Z, G.coef, G.ray, H.scratch, film,
                                                                                  active_pixels, n_active, ray_weight, gate, a_begin, fresh ? 1 : 0);
            else
                launch_mma<16>(mma_tma, grid, threads, stream, P, Z, G.coef, G.ray, H.scratch, film,
                                                                                    active_pixels, n_active, ray_weight, gate, a_begin, fresh ? 1 : 0);
        } else if (kL == 1)
            planarg_kernel<1><<<grid, threads, g_smem_bytes(1), stream>>>(P, Z, G.coef, G.ray, H.scratch, film,
                                                                         active_pixels, n_active, ray_weight, gate, a_begin, fresh ? 1 : 0);
        else if (kL == 2)
            planarg_kernel<2><<<grid, threads, g_smem_bytes(2), stream>>>(P, Z, G.coef, G.ray, H.scratch, film,
                                                                         active_pixels, n_active, ray_weight, gate, a_begin, fresh ? 1 : 0);
        else
            planarg_kernel<4><<<grid, threads, g_smem_bytes(4), stream>>>(P, Z, G.coef, G.ray, H.scratch, film,
                                                                         active_pixels, n_active, ray_weight, gate, a_begin, fresh ? 1 : 0);
        count_launch();
        PZ_TRY(cudaGetLastError());
    }
    return cudaSuccess;
}

// ---- the patch-gather forward without any allocation ------------------------------------------------------
// launch_planarg_forward keeps the [angle][column][layer] repack of the patterns in a scratch buffer that it
// (re)allocates with cudaMalloc when a call needs more than the last one did.  Callers that may not allocate
// (CUDA graph capture, serving loops, harnesses that forbid it) get the same two kernels - zpack_kernel, then
// planarg_mma_kernel per z-block - on a scratch that lives in the library image: kNoAllocFloats floats
// (16 MiB) of device memory on every GPU that loads the library, until the process exits.  Every call on a
// device shares it: zpack_kernel overwrites all of the part a call uses before the gather reads it, so nothing
// has to start at zero, but two calls that overlap on the GPU (different streams) race on it.
constexpr size_t kNoAllocFloats = size_t(4) << 20;
static __device__ __align__(256) float g_noalloc_scratch[kNoAllocFloats];

size_t planarg_noalloc_capacity() { return kNoAllocFloats; }

// Device address of the scratch on `device`, which must be the current device: in host code the array's name
// is only a handle for cudaGetSymbolAddress, not an address a kernel can use.  Resolved once per device.
static cudaError_t noalloc_scratch(int device, float** out) {
    constexpr int kMaxDevices = 64;
    static float* ptrs[kMaxDevices] = {};
    static std::mutex lock;
    if (device < 0 || device >= kMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> hold(lock);
    if (!ptrs[device]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess) return e;
        ptrs[device] = static_cast<float*>(p);
    }
    *out = ptrs[device];
    return cudaSuccess;
}

cudaError_t launch_planarg_forward_noalloc(const SceneParams& P, const PlanarZHost& H, const PlanarGHost& G, int device,
                                           const int* gate, const float* active_data, const uint32_t* active_pixels,
                                           uint64_t n_active, float ray_weight, float* film, bool fresh,
                                           cudaStream_t stream) {
    const int nk = static_cast<int>(n_active / (static_cast<uint64_t>(P.res_x) * P.n_rows));
    if (static_cast<size_t>(nk) * P.res_x * H.nz_pad > kNoAllocFloats) return cudaErrorInvalidValue;
    float* scratch = nullptr;
    PZ_TRY(noalloc_scratch(device, &scratch));
    const dim3 pgrid((P.res_x + 31) / 32, H.nz_pad / 32, nk);
    zpack_kernel<<<pgrid, 256, 0, stream>>>(P, H.sub[0].plan, active_data, scratch, nk, gate);
    count_launch();
    PZ_TRY(cudaGetLastError());
    const PlanarZPlan& Z = G.plan;
    const int regions = ((Z.tiles_x + kGWarpsX - 1) / kGWarpsX) * ((Z.tiles_y + kGWarpsY - 1) / kGWarpsY);
    constexpr int threads = 32 * kGWarpsX * kGWarpsY;
    for (int i = 0; i < H.n_sub; ++i) {
        const PlanarZSub& S = H.sub[i];
        const dim3 grid(regions, S.n_blocks);
        const int kL = S.plan.lanes_l, a_begin = S.plan.a_begin;
        if (kL == 1)
            launch_mma<4>(mma_tma, grid, threads, stream, P, Z, G.coef, G.ray, scratch, film, active_pixels,
                                                                              n_active, ray_weight, gate, a_begin, fresh ? 1 : 0);
        else if (kL == 2)
            launch_mma<8>(mma_tma, grid, threads, stream, P, Z, G.coef, G.ray, scratch, film, active_pixels,
                                                                              n_active, ray_weight, gate, a_begin, fresh ? 1 : 0);
        else
            launch_mma<16>(mma_tma, grid, threads, stream, P, Z, G.coef, G.ray, scratch, film, active_pixels,
                                                                                n_active, ray_weight, gate, a_begin, fresh ? 1 : 0);
        count_launch();
        PZ_TRY(cudaGetLastError());
    }
    return cudaSuccess;
}

}  // namespace drtvam
