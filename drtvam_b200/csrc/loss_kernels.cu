// Fused thresholded loss (Wechsler et al. 2024) on the dose volume: one pass
// computes the reduced loss and, optionally, d loss / d dose.
// Reference: ThresholdedLoss.eval_in / eval_out / eval (loss.py:115-132) and
// Loss.__call__ reduction (loss.py:58-59); the reference evaluates it with
// ~20 elementwise Dr.Jit ops + AD, here it is one read of (dose, target) and
// one write of the gradient.  HBM bound: 12 B per voxel with gradient.
#include <algorithm>
#include <cstdint>
#include <cuda_runtime.h>

#include "kernels.h"

namespace drtvam {

namespace {

constexpr int kLossThreads = 256;

// d/db relu(b)^(k+1) / (k+1) shape: relu(b)^k with relu(b)^0 = [b > 0], so that K = 1 has the step derivative
__device__ __forceinline__ float powi(float b, int k) {  // b >= 0, small integer exponent
    float r = b > 0.f ? 1.f : 0.f;
    for (int i = 0; i < k; ++i) r *= b;
    return r;
}

struct LossParams {
    float tl, tu, w_obj, w_void, w_lim;
    int K;
    float grad_scale;  // 1 (sum) or 1 / N (mean): applied to the gradient
};

__device__ __forceinline__ float loss_term(const LossParams& L, float x, float t, float& g) {
    if (x != x) {  // relu(NaN) is NaN in the reference; fmaxf would return the other operand and report loss 0
        g = x;
        return x;
    }
    if (t > 0.f) {  // inside the object: under-exposure below tu, over-exposure above 1
        const float a = fmaxf(L.tu - x, 0.f), b = fmaxf(x - 1.f, 0.f);
        const float a1 = powi(a, L.K - 1), b1 = powi(b, L.K - 1);
        g = static_cast<float>(L.K) * (L.w_lim * b1 - L.w_obj * a1);
        return L.w_obj * a1 * a + L.w_lim * b1 * b;
    }
    const float c = fmaxf(x - L.tl, 0.f);
    const float c1 = powi(c, L.K - 1);
    g = static_cast<float>(L.K) * L.w_void * c1;
    return L.w_void * c1 * c;
}

// x_b may be null; otherwise the loss is evaluated at x + alpha * x_b (Armijo trial, lbfgs.py:258).
template <bool kGrad>
__global__ void __launch_bounds__(kLossThreads)
thresholded_loss_kernel(const float* __restrict__ x, const float* __restrict__ x_b, float alpha,
                        const float* __restrict__ target, size_t n, LossParams L, float* __restrict__ grad,
                        double* __restrict__ loss_out) {
    const size_t stride = static_cast<size_t>(gridDim.x) * kLossThreads;
    size_t i = static_cast<size_t>(blockIdx.x) * kLossThreads + threadIdx.x;
    float acc = 0.f;
    const size_t n4 = n >> 2;
    const bool aligned = ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(target) |
                           reinterpret_cast<uintptr_t>(grad) | reinterpret_cast<uintptr_t>(x_b)) & 15) == 0;
    if (aligned) {
        for (; i < n4; i += stride) {
            float4 v = __ldcs(reinterpret_cast<const float4*>(x) + i);
            if (x_b) {
                const float4 d = __ldcs(reinterpret_cast<const float4*>(x_b) + i);
                v.x = fmaf(alpha, d.x, v.x); v.y = fmaf(alpha, d.y, v.y);
                v.z = fmaf(alpha, d.z, v.z); v.w = fmaf(alpha, d.w, v.w);
            }
            const float4 t = __ldcs(reinterpret_cast<const float4*>(target) + i);
            float4 g;
            acc += loss_term(L, v.x, t.x, g.x) + loss_term(L, v.y, t.y, g.y);
            acc += loss_term(L, v.z, t.z, g.z) + loss_term(L, v.w, t.w, g.w);
            if (kGrad) {
                g.x *= L.grad_scale; g.y *= L.grad_scale; g.z *= L.grad_scale; g.w *= L.grad_scale;
                __stcs(reinterpret_cast<float4*>(grad) + i, g);
            }
        }
        i = n4 * 4 + static_cast<size_t>(blockIdx.x) * kLossThreads + threadIdx.x;  // tail
    }
    for (; i < n; i += stride) {
        float v = x[i];
        if (x_b) v = fmaf(alpha, x_b[i], v);
        float g;
        acc += loss_term(L, v, target[i], g);
        if (kGrad) grad[i] = g * L.grad_scale;
    }
    // block reduction in double, one atomic per block
    double d = static_cast<double>(acc);
    for (int o = 16; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
    __shared__ double part[kLossThreads / 32];
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = d;
    __syncthreads();
    if (threadIdx.x < 32) {
        d = threadIdx.x < kLossThreads / 32 ? part[threadIdx.x] : 0.0;
        for (int o = 4; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
        if (threadIdx.x == 0) atomicAdd(loss_out, d);
    }
}

// Surface-aware film (two channels per voxel: dose received inside / outside the target, film.py:17-21) against the
// two-channel target of sub-voxel volumes (sensor.py:53-110): per voxel
//     w_in eval_in(x[0]) + w_out eval_out(x[1]),   w_in = t[0] / (t[0] + t[1]),  w_out = t[1] / (t[0] + t[1])
// (Loss.__call__, loss.py:44-53).  n = number of voxels; x, x_b, target and grad hold 2 n floats.
template <bool kGrad>
__global__ void __launch_bounds__(kLossThreads)
thresholded_loss_sa_kernel(const float2* __restrict__ x, const float2* __restrict__ x_b, float alpha,
                           const float2* __restrict__ target, size_t n, LossParams L, float2* __restrict__ grad,
                           double* __restrict__ loss_out) {
    const size_t stride = static_cast<size_t>(gridDim.x) * kLossThreads;
    float acc = 0.f;
    for (size_t i = static_cast<size_t>(blockIdx.x) * kLossThreads + threadIdx.x; i < n; i += stride) {
        float2 v = __ldcs(x + i);
        if (x_b) {
            const float2 d = __ldcs(x_b + i);
            v.x = fmaf(alpha, d.x, v.x);
            v.y = fmaf(alpha, d.y, v.y);
        }
        const float2 t = __ldcs(target + i);
        const float tot = t.x + t.y;
        const float w_in = t.x / tot, w_out = t.y / tot;
        float g_in, g_out;
        acc += w_in * loss_term(L, v.x, 1.f, g_in) + w_out * loss_term(L, v.y, 0.f, g_out);
        if (kGrad) __stcs(grad + i, make_float2(w_in * g_in * L.grad_scale, w_out * g_out * L.grad_scale));
    }
    double d = static_cast<double>(acc);
    for (int o = 16; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
    __shared__ double part[kLossThreads / 32];
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = d;
    __syncthreads();
    if (threadIdx.x < 32) {
        d = threadIdx.x < kLossThreads / 32 ? part[threadIdx.x] : 0.0;
        for (int o = 4; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
        if (threadIdx.x == 0) atomicAdd(loss_out, d);
    }
}

}  // namespace

cudaError_t launch_thresholded_loss_sa(const float* x, const float* x_b, float alpha, const float* target, size_t n_vox,
                                       float tl, float tu, float w_obj, float w_void, float w_lim, int K, float grad_scale,
                                       float* grad, double* loss_out, cudaStream_t stream) {
    cudaError_t e = cudaMemsetAsync(loss_out, 0, sizeof(double), stream);
    if (e != cudaSuccess) return e;
    if (n_vox == 0) return cudaSuccess;
    LossParams L{tl, tu, w_obj, w_void, w_lim, K, grad_scale};
    size_t blocks = (n_vox / 2 + kLossThreads - 1) / kLossThreads;
    blocks = std::min<size_t>(std::max<size_t>(blocks, 1), 148 * 8);
    const float2 *x2 = reinterpret_cast<const float2*>(x), *b2 = reinterpret_cast<const float2*>(x_b),
                 *t2 = reinterpret_cast<const float2*>(target);
    if (grad)
        thresholded_loss_sa_kernel<true><<<static_cast<unsigned>(blocks), kLossThreads, 0, stream>>>(
            x2, b2, alpha, t2, n_vox, L, reinterpret_cast<float2*>(grad), loss_out);
    else
        thresholded_loss_sa_kernel<false><<<static_cast<unsigned>(blocks), kLossThreads, 0, stream>>>(x2, b2, alpha, t2, n_vox, L,
                                                                                                    nullptr, loss_out);
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_thresholded_loss(const float* x, const float* x_b, float alpha, const float* target, size_t n,
                                    float tl, float tu, float w_obj, float w_void, float w_lim, int K, float grad_scale,
                                    float* grad, double* loss_out, cudaStream_t stream) {
    cudaError_t e = cudaMemsetAsync(loss_out, 0, sizeof(double), stream);
    if (e != cudaSuccess) return e;
    if (n == 0) return cudaSuccess;
    LossParams L{tl, tu, w_obj, w_void, w_lim, K, grad_scale};
    size_t blocks = (n / 4 + kLossThreads - 1) / kLossThreads;
    if (blocks > 148 * 8) blocks = 148 * 8;
    if (blocks < 1) blocks = 1;
    if (grad)
        thresholded_loss_kernel<true><<<static_cast<unsigned>(blocks), kLossThreads, 0, stream>>>(x, x_b, alpha, target,
                                                                                                   n, L, grad, loss_out);
    else
        thresholded_loss_kernel<false><<<static_cast<unsigned>(blocks), kLossThreads, 0, stream>>>(
            x, x_b, alpha, target, n, L, nullptr, loss_out);
    count_launch();
    return cudaGetLastError();
}

}  // namespace drtvam
