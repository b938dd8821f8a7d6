// Jittered planar passes: regular_sampling = false, the reference's default
// (integrators/common.py:22,92-95).
//
// Every lane (pixel, sample) has its own sub-pixel jitter, so no two rays share a segment and the
// coefficient tables of the deterministic kernels (planarz_kernels.cu) do not apply.  What survives is
// the planar structure: a ray never leaves its voxel layer (d_z = 0, sensor.py:345), the 32 DMD rows
// above one another at one (angle, column) walk almost the same 2-D path, and the dose of a layer tile
// can live in shared memory for all angles.  So, as in the layer-parallel kernels, a WARP takes one
// (angle, column) and its 32 lanes are 32 voxel layers - but every lane marches its OWN jittered segment:
//   - a pre-pass (jtable_kernel) draws each lane's samples from its PCG32 stream, traces the vial
//     interfaces once per pass and writes the medium segment (origin, direction, length, weight) into a
//     table ordered [angle][column][32-layer block][item][layer lane], so a warp's 32 entries are one
//     coalesced kilobyte.  Lanes that are not one planar segment inside their block (a frozen DDA axis,
//     sensor.py:359; a corner of a square vial; a jitter sample that lands in the next block of layers)
//     are walked right there by the general voxel walk;
//   - the tile kernel (planarj_kernel) owns [tile_h][tile_w][32 layers] floats of shared memory for all
//     angles; lane = layer puts the 32 lanes of a warp on 32 different banks, so the red.shared.add of
//     a deposit never meets a bank conflict (the per-ray tile kernel of round 1 serialised ~3.5 lanes per
//     bank on average); the tile is flushed once with red.global.add.v4.f32.  The adjoint stages the dose
//     gradient in the same tile and gathers with plain ld.shared.
// Reference: TVAMIntegrator.sample_rays (integrators/common.py:70-116), DDAVolumetricSensor.accumulate
// (sensor.py:306-440), VolumeIntegrator.render / render_backward (integrators/volume.py:18-56, 97-134).
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "kernels.h"
#include "planar.cuh"
#include "tracer.cuh"

namespace drtvam {

namespace {

constexpr int kJThreads = 1024;
constexpr int kJWarps = kJThreads / 32;
constexpr float kLog2e = 1.4426950408889634f;
constexpr int kScatter = 0, kGather = 1;

struct JTable {
    float4* a;  // ox, oy, dx, dy
    float4* b;  // tmax, ray weight x throughput at the origin, layer lane inside the block (int bits; -1 =
                // nothing to march, -2 = walked by the table kernel), pattern value (forward passes)
};

// layout of the table: slot = ((((kk * res_x + c) * n_blk + blk) * items + q) * 32 + lane
struct JLayout {
    int n_blk;   // 32-layer blocks
    int items;   // rows of the fullest layer x samples per pixel
    int kc;      // first angle of the chunk, relative to the first angle of the call
    int nkc;     // angles in the chunk
};

__device__ __forceinline__ uint32_t sampler_stream(const SceneParams& P, size_t j, uint32_t smp, uint32_t spp) {
    const size_t g = P.sampler_index ? static_cast<size_t>(__ldg(P.sampler_index + j)) : j;
    return static_cast<uint32_t>(g * spp + smp);
}

// wmax[0] = largest |weight x pattern value| of the table of a forward pass, as the bits of a non-negative float
// (ordered like unsigned integers): the scale of the fixed-point tile of planarj_kernel.  The plain read first
// keeps all but the first few thousand lanes off the atomic.
__device__ __forceinline__ void note_weight(unsigned* wmax, float v) {
    const unsigned b = __float_as_uint(fabsf(v));
    if (b < 0x7f800000u && b > *reinterpret_cast<volatile unsigned*>(wmax)) atomicMax(wmax, b);
}

// One thread per table slot.  kScatter: in = active_data, out = film (only lanes walked here touch it).
// kGather: in = dL, gT = per-slot gradient (lanes walked here write theirs, every other slot is zeroed).
// kRefresh (forward only): the geometry in the table is that of this very call - same pixels, seed and
// spp, as for A d after A p of one L-BFGS iteration (lbfgs.py:248 renders the direction with the seed of
// the primal pass) - so only the pattern values are written again and the odd lanes walked again.
template <int kMode, bool kRefresh>
__global__ void __launch_bounds__(256)
jtable_kernel(const __grid_constant__ SceneParams P, const __grid_constant__ SceneParams Pw,
              const __grid_constant__ PlanarZPlan Z, const int* __restrict__ layer_a, const JLayout Ly,
              const float* __restrict__ in, float* __restrict__ out, float* __restrict__ gT,
              const uint32_t* __restrict__ active_pixels, uint64_t n_active, uint32_t spp, uint32_t seed,
              float ray_weight, JTable T, const int* __restrict__ gate, unsigned* __restrict__ wmax) {
    if (*gate != 0) return;
    const size_t t = static_cast<size_t>(blockIdx.x) * 256 + threadIdx.x;
    const size_t total = static_cast<size_t>(Ly.nkc) * P.res_x * Ly.n_blk * Ly.items * 32;
    if (t >= total) return;
    const int lane = static_cast<int>(t & 31);
    size_t r = t >> 5;
    const int q = static_cast<int>(r % Ly.items); r /= Ly.items;
    const int blk = static_cast<int>(r % Ly.n_blk); r /= Ly.n_blk;
    const int c = static_cast<int>(r % P.res_x);
    const int kk = Ly.kc + static_cast<int>(r / P.res_x);

    int state = -1;   // layer lane inside the block, -1 = nothing to march, -2 = walked here (odd lane)
    if (kRefresh) {
        state = __float_as_int(__ldg(&T.b[t].z));
        if (state == -1) return;
    }
    float4 ea = make_float4(0.f, 0.f, 1.f, 0.f);
    float4 eb = make_float4(0.f, 0.f, __int_as_float(-1), 0.f);
    float g_here = 0.f;
    const int a = blk * 32 + lane;
    const int ri = q / static_cast<int>(spp), smp = q - ri * static_cast<int>(spp);
    bool live = a < Z.nz;
    int row = 0;
    if (live) {
        const int rb = __ldg(Z.a_row_begin + a), re = __ldg(Z.a_row_begin + a + 1);
        live = ri < re - rb;
        if (live) row = __ldg(Z.a_rows + rb + ri);
    }
    if (live) {
        int k0, nk;
        call_angles(P, active_pixels, n_active, k0, nk);
        const size_t j = active_index(P, kk, row, c);
        float em = 1.f;
        if (kMode == kScatter) em = __ldg(in + j);
        if (kRefresh && state >= 0) {
            T.b[t].w = em;
            note_weight(wmax, T.b[t].y * em);
            return;
        }
        // the gate has verified the layout, so the pixel index follows from (angle, row, column)
        const uint32_t pixel = (static_cast<uint32_t>(k0 + kk) * P.res_y + row) * P.res_x + c;
        const uint32_t stream = sampler_stream(P, j, smp, spp);
        Sampler rng;
        rng.seed(seed, stream);
        const Ray ray = generate_ray(P, pixel, rng);
        bool odd = kRefresh;
        int lyr = -1;
        RecordSink sink;
        sink.seg = ray;
        if (!kRefresh) {
            trace_lane<false>(P, ray, rng, sink);
            const bool one = sink.count == 1 && sink.w0 != 0.f && sink.tmax > 0.f && isfinite(sink.tmax);
            odd = sink.count > 1 || ray.dz != 0.f || (sink.count == 1 && sink.seg.dz != 0.f) ||
                  (sink.count == 1 && dda_has_frozen_axis(P, sink.seg, sink.tmax));
            if (!odd && one) {
                const float z = sink.seg.oz;
                if (z > P.bmin[2] && z < P.bmax[2]) {
                    const float vs = (P.bmax[2] - P.bmin[2]) / static_cast<float>(P.res[2]);
                    const int layer = min(max(__float2int_rz((z - P.bmin[2]) / vs), 0), P.res[2] - 1);
                    const int aa = __ldg(layer_a + layer);   // -1: outside the z window of this handle
                    if (aa >= 0) {
                        if ((aa >> 5) == blk) lyr = aa & 31;
                        else odd = true;   // the jitter moved the lane into another block of layers
                    }
                }
            }
        }
        if (odd) {
            // not a single planar segment of this block: general walk, deposited / gathered here
            Sampler rng2;
            rng2.seed(seed, stream);
            const Ray ray2 = generate_ray(P, pixel, rng2);
            // (a scattering scene: only the first medium segment is this kernel's, the tracer does the rest)
            const bool first_only = P.albedo != 0.f;
            if (kMode == kGather) {
                GatherSink gs{in};
                if (first_only) trace_first_segment(Pw, ray2, rng2, gs);
                else trace_lane<false>(Pw, ray2, rng2, gs);
                g_here = gs.acc * ray_weight;
            } else if (em != 0.f) {
                ScatterSink sc{out, em * ray_weight};
                if (first_only) trace_first_segment(Pw, ray2, rng2, sc);
                else trace_lane<false>(Pw, ray2, rng2, sc);
            }
            eb.z = __int_as_float(-2);
        } else if (lyr >= 0) {
            ea = make_float4(sink.seg.ox, sink.seg.oy, sink.seg.dx, sink.seg.dy);
            eb = make_float4(sink.tmax, ray_weight * sink.w0, __int_as_float(lyr), em);
        }
    }
    if (kRefresh) return;
    T.a[t] = ea;
    T.b[t] = eb;
    if (kMode == kScatter && __float_as_int(eb.z) >= 0) note_weight(wmax, eb.y * eb.w);
    if (kMode == kGather) gT[t] = g_here;
}

// kScatter: out = film (red.global flush).  kGather: in = dL (film layout), out = gT (red.global per lane).
//
// Shared tile: [(tile_h + 2)][(tile_w + 2)][32] floats - lane = layer, so the bank of an access is the
// lane's layer whatever voxel the lane stands in (lanes of a warp stand in neighbouring voxels: with the
// 33-float voxel pitch of the deterministic kernels they met 2.7 bank conflicts per deposit,
// profiles/r02b_*); a one-voxel rim catches the step a walk takes when rounding puts its last boundary a
// hair before the tile exit.
//
// Walk of one lane: optical depths relative to the tile entry; each trip deposits the current voxel up
// to min(next boundary, tile exit) and steps across the nearer boundary, so the last voxel needs no
// code of its own and the loop carries no step counters: 17 instructions per visit instead of 27.
//
// kInt (forward only): the tile holds 32-bit FIXED-POINT sums and a deposit is one red.shared.add.s32 - a native
// shared-memory atomic, fire and forget - instead of the LDS + FADD + ATOMS.CAST.SPIN loop that red.shared.add.f32
// compiles to, which a warp has to wait out (37 % of the stall samples of the march, profiles/r02zf_*).  The scale
// is chosen per CTA from the largest |weight x pattern value| of the table (wmax, written by jtable_kernel) so that
// 64 times the deposits a voxel can receive from 4 columns per angle and item still fit 31 bits; integer addition is
// exact, and a deposit is rounded to nearest at 2^-25 / (angles x items) of the largest possible one - far below
// fp32's own rounding of a sum of that many terms.  A word whose magnitude passes 2^30 at the flush is counted
// in wmax[1] (drtvam_scene_stat): never seen, a tripwire for rays focused 16 times denser than assumed.
template <int kMode, bool kInt>
__global__ void __launch_bounds__(kJThreads, 1)
planarj_kernel(const __grid_constant__ SceneParams P, const __grid_constant__ PlanarZPlan Z, const JLayout Ly,
               const JTable T, const float* __restrict__ in, float* __restrict__ out,
               const uint32_t* __restrict__ active_pixels, uint64_t n_active, const int* __restrict__ gate,
               unsigned* __restrict__ wmax) {
    extern __shared__ __align__(16) float tile[];
    if (*gate != 0) return;
    const int tile_id = blockIdx.x, blk = blockIdx.y, group = blockIdx.z;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int x0 = (tile_id % Z.tiles_x) * Z.tile_w, y0 = (tile_id / Z.tiles_x) * Z.tile_h;
    const int w = min(Z.tile_w, P.res[0] - x0), h = min(Z.tile_h, P.res[1] - y0);
    const int pitch = Z.tile_w + 2;                       // voxels per tile row, rim included
    const int tile_floats = (Z.tile_h + 2) * pitch * 32;
    const int a_base = blk * 32;
    const size_t layer_elems = static_cast<size_t>(P.res[0]) * P.res[1];
    const int z_mine = (a_base + lane < Z.nz) ? __ldg(Z.zl + a_base + lane) : -1;   // film layer of this lane
    const bool vec = ((x0 | w | P.res[0]) & 3) == 0 && (reinterpret_cast<uintptr_t>(kMode == kScatter ? out : in) & 15) == 0;

    // ---- stage the tile: lane = layer, a warp takes 4 voxels of a row at a time (conflict-free in shared
    // memory; in global memory every lane reads 16 contiguous bytes of its own layer)
    for (int i = threadIdx.x; i < tile_floats; i += kJThreads) tile[i] = 0.f;
    __syncthreads();
    if (kMode == kGather && z_mine >= 0) {
        const float* src = in + z_mine * layer_elems;
        if (vec) {
            const int w4 = w >> 2;
            for (int i = warp; i < h * w4; i += kJWarps) {
                const int yy = i / w4, xx = (i - yy * w4) << 2;
                const float4 v = __ldg(reinterpret_cast<const float4*>(src + static_cast<size_t>(y0 + yy) * P.res[0] + x0 + xx));
                float* d = tile + ((yy + 1) * pitch + xx + 1) * 32 + lane;
                d[0] = v.x; d[32] = v.y; d[64] = v.z; d[96] = v.w;
            }
        } else {
            for (int i = warp; i < h * w; i += kJWarps) {
                const int yy = i / w, xx = i - yy * w;
                tile[((yy + 1) * pitch + xx + 1) * 32 + lane] = __ldg(src + static_cast<size_t>(y0 + yy) * P.res[0] + x0 + xx);
            }
        }
    }
    if (kMode == kGather) __syncthreads();

    int k0, nk;
    call_angles(P, active_pixels, n_active, k0, nk);
    const int kfirst = k0 + Ly.kc;   // first angle held by the table
    const int ka = kfirst + static_cast<int>(static_cast<long long>(Ly.nkc) * group / Z.groups);
    const int kb = kfirst + static_cast<int>(static_cast<long long>(Ly.nkc) * (group + 1) / Z.groups);
    const int* __restrict__ off = Z.tile_off + static_cast<size_t>(tile_id) * (P.n_patterns + 1);
    const int* __restrict__ lst = Z.tile_clo + __ldg(Z.tile_ebase + tile_id);   // angle * res_x + column per entry

    const float sig = P.sigma_t * kLog2e;
    const float vsx = (P.bmax[0] - P.bmin[0]) / static_cast<float>(P.res[0]);
    const float vsy = (P.bmax[1] - P.bmin[1]) / static_cast<float>(P.res[1]);
    const float rvx = 1.f / vsx, rvy = 1.f / vsy;
    // tile rectangle in world units
    const float xl = P.bmin[0] + static_cast<float>(x0) * vsx, xh = P.bmin[0] + static_cast<float>(x0 + w) * vsx;
    const float yl = P.bmin[1] + static_cast<float>(y0) * vsy, yh = P.bmin[1] + static_cast<float>(y0 + h) * vsy;
    const uint32_t tile_s = static_cast<uint32_t>(__cvta_generic_to_shared(tile));
    const int sxb = 4 * 32, syb = 4 * 32 * pitch;
    // fixed-point scale of this CTA's tile: 2^25 / (angles x items x largest deposit)
    float fx_scale = 0.f, fx_inv = 0.f;
    if (kInt) {
        const float de_max = 1.f - ex2_approx(-sig * (vsx + vsy));     // no chord through a voxel is longer
        const float d_max = __uint_as_float(*reinterpret_cast<volatile unsigned*>(wmax)) * de_max;
        if (d_max > 0.f) {
            fx_scale = 33554432.f / (static_cast<float>(max(kb - ka, 1)) * static_cast<float>(Ly.items) * d_max);
            fx_inv = 1.f / fx_scale;
        }
    }

    const int f_base = __ldg(off + ka);
    const int tasks = (__ldg(off + kb) - f_base) * Ly.items;   // < 2^31: checked by planarj_pass
    const unsigned slots_kc = static_cast<unsigned>(P.res_x) * Ly.n_blk * Ly.items;   // slots / 32 per angle
    int k = ka;
    for (int f = warp; f < tasks; f += kJWarps) {
        const int g = f / Ly.items + f_base;
        const int q = f - (g - f_base) * Ly.items;
        const int kc = __ldg(lst + g);
        k = kc / P.res_x;
        const int c = kc - k * P.res_x;
        // < 2^31 slots per chunk (planarj_pass): 32-bit index arithmetic
        const unsigned slot = ((static_cast<unsigned>(k - kfirst) * slots_kc + (static_cast<unsigned>(c) * Ly.n_blk + blk) * Ly.items + q) << 5) + lane;
        const float4 eb = __ldg(T.b + slot);
        const int lyr = __float_as_int(eb.z);
        float wgt = kMode == kScatter ? eb.y * eb.w : eb.y;
        bool live = lyr >= 0 && wgt != 0.f;
        if (!__any_sync(0xffffffffu, live)) continue;
        float4 ea = make_float4(0.f, 0.f, 1.f, 0.f);
        if (live) ea = __ldg(T.a + slot);

        // ---- clip to the tile (slab test with reciprocals: the clip only places the entry and the exit)
        float t0 = 0.f, t1 = eb.x;
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
        float acc = 0.f;
        if (live) {
            // entry voxel by truncation then clip (sensor.py:345-346), tile-local
            const float gx = fmaf(ea.z, t0, ea.x), gy = fmaf(ea.w, t0, ea.y);
            const int ix = min(max(__float2int_rz((gx - P.bmin[0]) * rvx) - x0, 0), w - 1);
            const int iy = min(max(__float2int_rz((gy - P.bmin[1]) * rvy) - y0, 0), h - 1);
            // optical depths relative to the tile entry (see planarz_kernels.cu): the throughput at the
            // entry goes into the weight, the differences below are then 16 x less noisy
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
            uint32_t addr = tile_s + 4u * static_cast<uint32_t>(((iy + 1) * pitch + ix + 1) * 32 + lyr);
            wgt *= ex2_approx(-sig * t0);
            if (kInt) wgt *= fx_scale;
            float e_prev = 1.f;
            // one trip: deposit the current voxel up to min(next boundary, tile exit), step across the
            // nearer boundary.  Unrolled by two so that the carried 2^-u needs no register move.
            auto trip = [&](float e_in, float& e_out) -> bool {
                const float un = fminf(ux, uy);
                const bool px = ux <= uy;
                e_out = ex2_approx(-fminf(un, u1));
                const float de = e_in - e_out;
                if (kMode == kGather) {
                    float v;
                    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
                    acc = fmaf(v, de, acc);
                } else if (kInt) {
                    asm volatile("red.shared.add.s32 [%0], %1;" ::"r"(addr), "r"(__float2int_rn(wgt * de)) : "memory");
                } else {
                    asm volatile("red.shared.add.f32 [%0], %1;" ::"r"(addr), "f"(wgt * de) : "memory");
                }
                addr += px ? sx : sy;
                if (px) ux += ax; else uy += ay;
                return un < u1;
            };
            for (;;) {
                float e_mid;
                if (!trip(e_prev, e_mid)) break;
                if (!trip(e_mid, e_prev)) break;
            }
        }
        if (kMode == kGather && live && acc != 0.f) atomicAdd(out + slot, acc * wgt);
    }
    if (kMode == kGather) return;
    __syncthreads();

    // ---- flush: lane = layer; every lane adds 4 voxels of a row of its own layer with one red.global.add.v4
    if (z_mine < 0) return;
    float* dst = out + z_mine * layer_elems;
    if (vec) {
        const int w4 = w >> 2;
        for (int i = warp; i < h * w4; i += kJWarps) {
            const int yy = i / w4, xx = (i - yy * w4) << 2;
            const float* src = tile + ((yy + 1) * pitch + xx + 1) * 32 + lane;
            float v0 = src[0], v1 = src[32], v2 = src[64], v3 = src[96];
            if (kInt) {
                const int i0 = __float_as_int(v0), i1 = __float_as_int(v1), i2 = __float_as_int(v2), i3 = __float_as_int(v3);
                if (max(max(abs(i0), abs(i1)), max(abs(i2), abs(i3))) > (1 << 30)) atomicAdd(wmax + 1, 1u);
                v0 = static_cast<float>(i0) * fx_inv; v1 = static_cast<float>(i1) * fx_inv;
                v2 = static_cast<float>(i2) * fx_inv; v3 = static_cast<float>(i3) * fx_inv;
            }
            if (v0 != 0.f || v1 != 0.f || v2 != 0.f || v3 != 0.f)
                red_add_v4(dst + static_cast<size_t>(y0 + yy) * P.res[0] + (x0 + xx), v0, v1, v2, v3);
        }
    } else {
        for (int i = warp; i < h * w; i += kJWarps) {
            const int yy = i / w, xx = i - yy * w;
            float v = tile[((yy + 1) * pitch + xx + 1) * 32 + lane];
            if (kInt) {
                const int iv = __float_as_int(v);
                if (abs(iv) > (1 << 30)) atomicAdd(wmax + 1, 1u);
                v = static_cast<float>(iv) * fx_inv;
            }
            if (v != 0.f) atomicAdd(dst + static_cast<size_t>(y0 + yy) * P.res[0] + (x0 + xx), v);
        }
    }
}

// gT [slot]  ->  pattern_grad [kk][row][col] = sum over the samples of the pixel (the dr.repeat of common.py:82)
__global__ void __launch_bounds__(256)
junpack_kernel(const __grid_constant__ SceneParams P, const __grid_constant__ PlanarZPlan Z, const int* __restrict__ row_ri,
               const JLayout Ly, const float* __restrict__ gT, float* __restrict__ grad, uint32_t spp,
               const int* __restrict__ gate) {
    __shared__ float tr[32][33];
    if (*gate != 0) return;
    const int kk = blockIdx.z;   // relative to the chunk
    const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int j = ty; j < 32; j += 8) {
        const int c = c0 + j, r = P.row_begin + r0 + tx;
        float v = 0.f;
        if (c < P.res_x && r < P.row_begin + P.n_rows) {
            const int a = __ldg(Z.row_a + r);
            if (a >= 0) {
                const int ri = __ldg(row_ri + r);
                const size_t base = ((((static_cast<size_t>(kk) * P.res_x + c) * Ly.n_blk + (a >> 5)) * Ly.items +
                                      static_cast<size_t>(ri) * spp) << 5) + (a & 31);
                for (uint32_t s = 0; s < spp; ++s) v += __ldg(gT + base + (static_cast<size_t>(s) << 5));
            }
        }
        tr[j][tx] = v;
    }
    __syncthreads();
    for (int j = ty; j < 32; j += 8) {
        const int r = P.row_begin + r0 + j, c = c0 + tx;
        if (r < P.row_begin + P.n_rows && c < P.res_x) grad[active_index(P, Ly.kc + kk, r, c)] = tr[tx][j];
    }
}

}  // namespace

#define PJ_TRY(expr)                      \
    do {                                  \
        cudaError_t _e = (expr);          \
        if (_e != cudaSuccess) return _e; \
    } while (0)

void planarj_release(PlanarJHost& J) {
    void* ptrs[] = {J.tile_off, J.tile_clo, J.tile_ebase, J.table, J.table_adj, J.gT, J.wmax};
    for (void* p : ptrs)
        if (p) cudaFree(p);
    J = PlanarJHost();
}

// Tile geometry (the largest xy tile whose [h][w][33] float slab fits the 227 KB of opt-in shared
// memory) and the conservative per-(tile, angle) column ranges: a jittered ray of column c lies within
// half a pixel of the centre ray of c, refraction can magnify that, hence the margin of 1.5 pixels.
cudaError_t planarj_build(const SceneParams& P, const PlanarHost& base, const PlanarZHost& Zh, PlanarJHost& J) {
    planarj_release(J);
    if (!Zh.valid || Zh.nz == 0) return cudaSuccess;
    const int np = P.n_patterns, rx = P.res[0], ry = P.res[1];
    const int area = (227 * 1024 / 4) / 32;   // voxels of [32 layers], the one-voxel rim included
    const int side = static_cast<int>(std::floor(std::sqrt(static_cast<double>(area)))) - 2;
    const int tw_max = std::max(4, side & ~3);
    J.tiles_x = (rx + tw_max - 1) / tw_max;
    J.tile_w = (((rx + J.tiles_x - 1) / J.tiles_x) + 3) & ~3;
    J.tiles_x = (rx + J.tile_w - 1) / J.tile_w;
    const int th_max = std::max(1, area / (J.tile_w + 2) - 2);
    J.tiles_y = (ry + th_max - 1) / th_max;
    J.tile_h = (ry + J.tiles_y - 1) / J.tiles_y;
    J.smem_bytes = sizeof(float) * static_cast<size_t>(J.tile_h + 2) * (J.tile_w + 2) * 32;
    const int n_tiles = J.tiles_x * J.tiles_y;
    PJ_TRY(cudaMalloc(&J.tile_off, sizeof(int) * static_cast<size_t>(n_tiles) * (np + 1)));
    PJ_TRY(cudaMalloc(&J.tile_ebase, sizeof(int) * (n_tiles + 1)));
    const float margin = 1.5f * std::fmax(P.pixel_x, P.pixel_y);
    // column LISTS, not ranges: behind a refracting box the range of a (tile, angle) spans the gap between the two
    // ray families that cross the tile, and every column in it costs a table read and a clip per item
    long long n_entries = 0;
    PJ_TRY(planar_tile_lists(P, base.seg_a, base.seg_b, J.tiles_x, J.tiles_y, J.tile_w, J.tile_h, J.tile_off, J.tile_ebase,
                             &J.tile_clo, &n_entries, margin));
    PJ_TRY(cudaFuncSetAttribute(planarj_kernel<kScatter, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                static_cast<int>(J.smem_bytes)));
    PJ_TRY(cudaFuncSetAttribute(planarj_kernel<kScatter, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                static_cast<int>(J.smem_bytes)));
    PJ_TRY(cudaFuncSetAttribute(planarj_kernel<kGather, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                static_cast<int>(J.smem_bytes)));
    PJ_TRY(cudaMalloc(&J.wmax, 2 * sizeof(unsigned)));
    PJ_TRY(cudaMemset(J.wmax, 0, 2 * sizeof(unsigned)));
    J.valid = true;
    return cudaSuccess;
}

static cudaError_t j_reserve(void*& ptr, size_t& have, size_t need) {
    if (need <= have) return cudaSuccess;
    if (ptr) cudaFree(ptr);
    ptr = nullptr;
    have = 0;
    PJ_TRY(cudaMalloc(&ptr, need));
    have = need;
    return cudaSuccess;
}

template <int kMode>
static cudaError_t planarj_pass(const SceneParams& P, const SceneParams& Pw, PlanarZHost& Zh, PlanarJHost& J,
                                const int* gate, const float* in, float* out, const uint32_t* active_pixels,
                                uint64_t n_active, uint32_t spp, uint32_t seed, float ray_weight, bool cacheable,
                                cudaStream_t stream) {
    const int nk = static_cast<int>(n_active / (static_cast<uint64_t>(P.res_x) * P.n_rows));
    JLayout Ly{};
    Ly.n_blk = Zh.nz_pad / 32;
    Ly.items = std::max(1, Zh.rows_max) * static_cast<int>(spp);
    if (static_cast<long long>(P.n_patterns) * P.res_x * Ly.items >= (1ll << 31)) return cudaErrorInvalidValue;
    const size_t slots_per_angle = static_cast<size_t>(P.res_x) * Ly.n_blk * Ly.items * 32;
    // at most ~4 GB of table (32 bytes per slot, + 4 for the adjoint's per-slot gradients)
    size_t cap = 4ull << 30;
    if (const char* env = std::getenv("DRTVAM_JTABLE_MB")) cap = static_cast<size_t>(std::atoll(env)) << 20;
    const size_t per_slot = 32 + (kMode == kGather ? 4 : 0);
    int chunk = static_cast<int>(std::max<size_t>(1, cap / (slots_per_angle * per_slot)));
    chunk = std::min<long long>(chunk, ((1ll << 31) - 64) / static_cast<long long>(slots_per_angle));   // 32-bit slot indices
    if (chunk < 1) return cudaErrorInvalidValue;
    chunk = std::min(chunk, nk);
    const size_t slots_max = slots_per_angle * chunk;
    // forward and adjoint passes keep tables of their own: the forward one is reused by the next forward
    // pass of the same iteration (A d after A p), which the adjoint pass in between must not overwrite
    JTable T;
    if (kMode == kGather) {
        PJ_TRY(j_reserve(J.table_adj, J.table_adj_bytes, slots_max * 32));
        void* g = J.gT;
        PJ_TRY(j_reserve(g, J.gT_bytes, slots_max * 4));
        J.gT = static_cast<float*>(g);
        T.a = static_cast<float4*>(J.table_adj);
    } else {
        if (slots_max * 32 > J.table_bytes) J.cache_valid = false;
        PJ_TRY(j_reserve(J.table, J.table_bytes, slots_max * 32));
        T.a = static_cast<float4*>(J.table);
    }
    T.b = T.a + slots_max;

    PlanarZPlan Z{};
    Z.zl = Zh.zl; Z.a_row_begin = Zh.a_row_begin; Z.a_rows = Zh.a_rows; Z.row_a = Zh.row_a;
    Z.nz = Zh.nz; Z.nz_pad = Zh.nz_pad;
    Z.tiles_x = J.tiles_x; Z.tiles_y = J.tiles_y; Z.tile_w = J.tile_w; Z.tile_h = J.tile_h;
    Z.tile_off = J.tile_off; Z.tile_clo = J.tile_clo; Z.tile_ebase = J.tile_ebase;
    Z.lanes_l = 1;
    const long long ctas = static_cast<long long>(J.tiles_x) * J.tiles_y * Ly.n_blk;

    // the geometry of a forward pass survives in the table when the whole call fits one chunk
    const bool reuse = kMode == kScatter && cacheable && chunk == nk && J.cache_valid && J.cache_seed == seed && J.cache_spp == spp &&
                       J.cache_pixels == active_pixels && J.cache_n == n_active && J.cache_weight == ray_weight &&
                       J.cache_index == P.sampler_index && !std::getenv("DRTVAM_JTABLE_NOCACHE");
    if (kMode == kScatter) J.cache_valid = false;
    for (int kc = 0; kc < nk; kc += chunk) {
        Ly.kc = kc;
        Ly.nkc = std::min(chunk, nk - kc);
        const size_t slots = slots_per_angle * Ly.nkc;
        const unsigned tblocks = static_cast<unsigned>((slots + 255) / 256);
        if (kMode == kScatter) PJ_TRY(cudaMemsetAsync(J.wmax, 0, sizeof(unsigned), stream));   // the weight maximum of this table
        if (reuse)
            jtable_kernel<kScatter, true><<<tblocks, 256, 0, stream>>>(P, Pw, Z, Zh.layer_a, Ly, in, out, J.gT, active_pixels,
                                                                      n_active, spp, seed, ray_weight, T, gate, J.wmax);
        else
            jtable_kernel<kMode, false><<<tblocks, 256, 0, stream>>>(P, Pw, Z, Zh.layer_a, Ly, in, out, J.gT, active_pixels,
                                                                    n_active, spp, seed, ray_weight, T, gate, J.wmax);
        count_launch();
        PJ_TRY(cudaGetLastError());
        // enough CTAs for two waves of the 148 SMs: the tile flush is a red.global, so angle groups are free
        long long groups = (148 * 2 + ctas - 1) / std::max<long long>(ctas, 1);
        groups = std::max<long long>(1, std::min<long long>(groups, std::max(1, Ly.nkc / 4)));
        Z.groups = static_cast<int>(groups);
        const dim3 grid(J.tiles_x * J.tiles_y, Ly.n_blk, Z.groups);
        // DRTVAM_JITTER_ATOMICS=float keeps the fp32 compare-and-swap tile (A/B measurements)
        static const bool fixed_point = !(std::getenv("DRTVAM_JITTER_ATOMICS") &&
                                          std::strcmp(std::getenv("DRTVAM_JITTER_ATOMICS"), "float") == 0);
        if (kMode == kGather)
            planarj_kernel<kGather, false><<<grid, kJThreads, J.smem_bytes, stream>>>(P, Z, Ly, T, in, J.gT, active_pixels,
                                                                                     n_active, gate, J.wmax);
        else if (fixed_point)
            planarj_kernel<kScatter, true><<<grid, kJThreads, J.smem_bytes, stream>>>(P, Z, Ly, T, in, out, active_pixels,
                                                                                     n_active, gate, J.wmax);
        else
            planarj_kernel<kScatter, false><<<grid, kJThreads, J.smem_bytes, stream>>>(P, Z, Ly, T, in, out, active_pixels,
                                                                                      n_active, gate, J.wmax);
        count_launch();
        PJ_TRY(cudaGetLastError());
        if (kMode == kGather) {
            const dim3 ugrid((P.res_x + 31) / 32, (P.n_rows + 31) / 32, Ly.nkc);
            junpack_kernel<<<ugrid, 256, 0, stream>>>(P, Z, Zh.row_ri, Ly, J.gT, out, spp, gate);
            count_launch();
            PJ_TRY(cudaGetLastError());
        }
    }
    if (kMode == kScatter && cacheable && chunk == nk) {
        J.cache_valid = true;
        J.cache_seed = seed; J.cache_spp = spp; J.cache_pixels = active_pixels; J.cache_n = n_active;
        J.cache_weight = ray_weight;
        J.cache_index = P.sampler_index;
    }
    return cudaSuccess;
}

cudaError_t launch_planarj_forward(const SceneParams& P, const SceneParams& Pw, PlanarZHost& Zh, PlanarJHost& J,
                                   const int* gate, const float* active_data, const uint32_t* active_pixels,
                                   uint64_t n_active, uint32_t spp, uint32_t seed, float ray_weight, float* film,
                                   bool cacheable, cudaStream_t stream) {
    return planarj_pass<kScatter>(P, Pw, Zh, J, gate, active_data, film, active_pixels, n_active, spp, seed, ray_weight,
                                  cacheable, stream);
}

cudaError_t launch_planarj_adjoint(const SceneParams& P, const SceneParams& Pw, PlanarZHost& Zh, PlanarJHost& J,
                                   const int* gate, const float* dL, const uint32_t* active_pixels, uint64_t n_active,
                                   uint32_t spp, uint32_t seed, float ray_weight, float* grad, cudaStream_t stream) {
    return planarj_pass<kGather>(P, Pw, Zh, J, gate, dL, grad, active_pixels, n_active, spp, seed, ray_weight, false,
                                 stream);
}

}  // namespace drtvam
