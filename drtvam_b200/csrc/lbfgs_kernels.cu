// LinearLBFGS on the device (reference: src/drtvam/lbfgs.py:146-275).
//
// The reference's step does ~13 dot products and ~10 axpys over the pattern vector one after the other
// (two-loop recursion, lbfgs.py:213-234) and up to 20 loss evaluations with a host decision after each
// (Armijo search, lbfgs.py:252-266).  Here, per iteration:
//   lbfgs_history   s = p - p_old, y = g - g_old, p_old = p, g_old = g        one pass over four vectors
//   lbfgs_gram      G = B^T B for B = [s_0 .. s_{h-1}, y_0 .. y_{h-1}, g]     every vector read ONCE; with
//                   several GPUs the (2h+1)^2 doubles are the single all-reduce of the iteration
//   lbfgs_solve     the two-loop recursion on the coefficients of the search direction in the basis B
//                   (one thread; everything it needs is in G), g.d as a by-product
//   lbfgs_combine   d = - sum_j coef_j b_j                                    every vector read once
//   armijo_losses   the thresholded loss at dose + 2^-k ddose for a whole stage of trials in one sweep (k < 4,
//                   then k < 20; a stage that finds the search already over returns in its first instruction)
//   armijo_select   the first k that meets the Armijo condition, on the device (no host round trip)
//   lbfgs_update    p <- p + alpha d with alpha read from device memory
#include <cstdint>
#include <cuda_runtime.h>

#include "kernels.h"

namespace drtvam {

namespace {

constexpr int kLbThreads = 256;
constexpr int kMaxVec = 13;      // 2 * 6 + 1
constexpr int kMaxTrials = 32;

struct VecTable {
    const float* v[kMaxVec];
};

__device__ __forceinline__ double block_sum(double d, double* part) {
    for (int o = 16; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = d;
    __syncthreads();
    d = threadIdx.x < kLbThreads / 32 ? part[threadIdx.x] : 0.0;
    if (threadIdx.x < 32)
        for (int o = 4; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
    return d;   // valid in thread 0
}

__global__ void __launch_bounds__(kLbThreads)
history_kernel(const float* __restrict__ p, const float* __restrict__ g, float* __restrict__ p_old,
               float* __restrict__ g_old, float* __restrict__ s, float* __restrict__ y, size_t n, int first) {
    const size_t stride = static_cast<size_t>(gridDim.x) * kLbThreads;
    for (size_t i = static_cast<size_t>(blockIdx.x) * kLbThreads + threadIdx.x; i < n; i += stride) {
        const float pi = p[i], gi = g[i];
        if (!first) {
            s[i] = pi - p_old[i];
            y[i] = gi - g_old[i];
        }
        p_old[i] = pi;
        g_old[i] = gi;
    }
}

// upper triangle of the Gram matrix: fp32 products accumulated per thread in fp32 (a thread sees ~200
// elements of each vector), across threads and blocks in double
template <int kN>
__global__ void __launch_bounds__(kLbThreads)
gram_kernel(const VecTable T, size_t n, double* __restrict__ gram) {
    __shared__ double part[kLbThreads / 32];
    float acc[kN * (kN + 1) / 2];
#pragma unroll
    for (int i = 0; i < kN * (kN + 1) / 2; ++i) acc[i] = 0.f;
    const size_t stride = static_cast<size_t>(gridDim.x) * kLbThreads;
    const size_t tid = static_cast<size_t>(blockIdx.x) * kLbThreads + threadIdx.x;
    bool aligned = true;
#pragma unroll
    for (int a = 0; a < kN; ++a) aligned = aligned && (reinterpret_cast<uintptr_t>(T.v[a]) & 15) == 0;
    size_t done = 0;
    if (aligned) {   // 16 bytes per load and vector: kN loads in flight per thread
        const size_t n4 = n >> 2;
        for (size_t i = tid; i < n4; i += stride) {
            float4 x[kN];
#pragma unroll
            for (int a = 0; a < kN; ++a) x[a] = __ldcs(reinterpret_cast<const float4*>(T.v[a]) + i);
            int q = 0;
#pragma unroll
            for (int a = 0; a < kN; ++a)
#pragma unroll
                for (int b = a; b < kN; ++b) {
                    acc[q] = fmaf(x[a].x, x[b].x, acc[q]);
                    acc[q] = fmaf(x[a].y, x[b].y, acc[q]);
                    acc[q] = fmaf(x[a].z, x[b].z, acc[q]);
                    acc[q] = fmaf(x[a].w, x[b].w, acc[q]);
                    ++q;
                }
        }
        done = n4 << 2;
    }
    for (size_t i = done + tid; i < n; i += stride) {
        float x[kN];
#pragma unroll
        for (int a = 0; a < kN; ++a) x[a] = T.v[a][i];
        int q = 0;
#pragma unroll
        for (int a = 0; a < kN; ++a)
#pragma unroll
            for (int b = a; b < kN; ++b) acc[q] = fmaf(x[a], x[b], acc[q]), ++q;
    }
    int q = 0;
#pragma unroll
    for (int a = 0; a < kN; ++a)
#pragma unroll
        for (int b = a; b < kN; ++b) {
            const double d = block_sum(static_cast<double>(acc[q]), part);
            if (threadIdx.x == 0 && d != 0.0) {
                atomicAdd(gram + a * kN + b, d);
                if (a != b) atomicAdd(gram + b * kN + a, d);
            }
            ++q;
        }
}

// Two-loop recursion (lbfgs.py:213-234) on coefficients.  Basis: b_j = s_j (j < h), y_{j-h} (h <= j < 2h),
// g (j = 2h), oldest pair first.  q = sum_j delta_j b_j starts as g; s_i . q = sum_j delta_j G[i][j].
// out: coef[0 .. 2h] with d = - sum_j coef_j b_j, coef[kMaxVec] = g . d, coef[kMaxVec + 1] = 1 when a
// curvature y_i . s_i is not positive and finite (the caller resets the history, the step is -g).
__global__ void solve_kernel(const double* __restrict__ G, int h, int first_step, double* __restrict__ coef) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    const int N = 2 * h + 1;
    double delta[kMaxVec], alpha[kMaxVec];
    for (int j = 0; j < N; ++j) delta[j] = 0.0;
    delta[2 * h] = 1.0;
    bool bad = false;
    // A pair with y.s = 0 (both vectors vanish when the loss has all but reached 0) would make 1 / (y.s) infinite
    // and the direction NaN - in the reference too (lbfgs.py:221-222 takes dr.rcp(ys) unguarded).  Such a pair
    // carries no curvature information and is left out; the reference's arithmetic is unchanged wherever it is finite.
    auto usable = [](double ys) { return ys == ys && ys != 0.0 && fabs(ys) < 1.0e300; };
    for (int i = h - 1; i >= 0; --i) {
        const double ys = G[(h + i) * N + i];
        alpha[i] = 0.0;
        if (!usable(ys)) { bad = true; continue; }
        double sq = 0.0;
        for (int j = 0; j < N; ++j) sq += delta[j] * G[i * N + j];
        const double a = sq / ys;
        alpha[i] = a;
        delta[h + i] -= a;
    }
    double gamma = 1.0;
    if (!first_step && h > 0) {
        const double yy = G[(2 * h - 1) * N + (2 * h - 1)], ys = G[(2 * h - 1) * N + (h - 1)];
        if (usable(ys) && usable(yy)) gamma = ys / yy;
    }
    for (int j = 0; j < N; ++j) delta[j] *= gamma;
    for (int i = 0; i < h; ++i) {
        const double ys = G[(h + i) * N + i];
        if (!usable(ys)) continue;
        double yz = 0.0;
        for (int j = 0; j < N; ++j) yz += delta[j] * G[(h + i) * N + j];
        delta[i] += alpha[i] - yz / ys;
    }
    double gd = 0.0;
    for (int j = 0; j < N; ++j) {
        coef[j] = delta[j];
        gd -= delta[j] * G[2 * h * N + j];
    }
    coef[kMaxVec] = gd;
    coef[kMaxVec + 1] = bad ? 1.0 : 0.0;
}

template <int kN>
__global__ void __launch_bounds__(kLbThreads)
combine_kernel(const VecTable T, const double* __restrict__ coef, size_t n, float* __restrict__ out) {
    float c[kN];
#pragma unroll
    for (int a = 0; a < kN; ++a) c[a] = -static_cast<float>(coef[a]);
    const size_t stride = static_cast<size_t>(gridDim.x) * kLbThreads;
    for (size_t i = static_cast<size_t>(blockIdx.x) * kLbThreads + threadIdx.x; i < n; i += stride) {
        float d = 0.f;
#pragma unroll
        for (int a = 0; a < kN; ++a) d = fmaf(c[a], __ldcs(T.v[a] + i), d);
        out[i] = d;
    }
}

struct ArmijoParams {
    float tl, tu, w_obj, w_void, w_lim;
    int k0, trials;   // trials [k0, k0 + trials)
};

template <int K>
__device__ __forceinline__ float powk(float b) {   // relu(b)^K for b >= 0, relu(b)^1 = b
    float r = b;
#pragma unroll
    for (int i = 1; i < K; ++i) r *= b;
    return r;
}

template <int K>
__device__ __forceinline__ float armijo_term(const ArmijoParams& L, float x, float t) {
    if (x != x) return x;   // relu(NaN) is NaN in the reference (fmaxf would return the other operand)
    if (t > 0.f) return L.w_obj * powk<K>(fmaxf(L.tu - x, 0.f)) + L.w_lim * powk<K>(fmaxf(x - 1.f, 0.f));
    return L.w_void * powk<K>(fmaxf(x - L.tl, 0.f));
}

// losses[k0 + j] += sum_voxels loss(dose + 2^-(k0 + j) ddose), j < kT: several trials of the back-tracking
// search in one sweep of the two volumes.  `skip` (device float, may be null): non-zero = the search has
// ended in an earlier stage, nothing to do.
template <int K, int kT>
__global__ void __launch_bounds__(kLbThreads)
armijo_losses_kernel(const float* __restrict__ x, const float* __restrict__ dx, const float* __restrict__ target, size_t n,
                     ArmijoParams L, double* __restrict__ losses, const float* __restrict__ skip) {
    __shared__ double part[kLbThreads / 32];
    if (skip && *skip != 0.f) return;
    float acc[kT];
#pragma unroll
    for (int k = 0; k < kT; ++k) acc[k] = 0.f;
    const float a0 = exp2f(-static_cast<float>(L.k0));
    const size_t stride = static_cast<size_t>(gridDim.x) * kLbThreads;
    const size_t tid = static_cast<size_t>(blockIdx.x) * kLbThreads + threadIdx.x;
    size_t done = 0;
    if (((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(dx) | reinterpret_cast<uintptr_t>(target)) & 15) == 0) {
        const size_t n4 = n >> 2;
        for (size_t i = tid; i < n4; i += stride) {
            const float4 v = __ldcs(reinterpret_cast<const float4*>(x) + i);
            const float4 t = __ldcs(reinterpret_cast<const float4*>(target) + i);
            float4 d = __ldcs(reinterpret_cast<const float4*>(dx) + i);
            d.x *= a0; d.y *= a0; d.z *= a0; d.w *= a0;
#pragma unroll
            for (int k = 0; k < kT; ++k) {
                acc[k] += armijo_term<K>(L, v.x + d.x, t.x) + armijo_term<K>(L, v.y + d.y, t.y) +
                          armijo_term<K>(L, v.z + d.z, t.z) + armijo_term<K>(L, v.w + d.w, t.w);
                d.x *= 0.5f; d.y *= 0.5f; d.z *= 0.5f; d.w *= 0.5f;
            }
        }
        done = n4 << 2;
    }
    for (size_t i = done + tid; i < n; i += stride) {
        const float v = x[i], t = target[i];
        float d = dx[i] * a0;
#pragma unroll
        for (int k = 0; k < kT; ++k) {
            acc[k] += armijo_term<K>(L, v + d, t);
            d *= 0.5f;
        }
    }
#pragma unroll
    for (int k = 0; k < kT; ++k) {
        const double d = block_sum(static_cast<double>(acc[k]), part);
        if (threadIdx.x == 0 && k < L.trials) atomicAdd(losses + L.k0 + k, d);
    }
}

// lbfgs.py:256-266: alpha = 1, halved until f(alpha) <= f(0) + c1 alpha g.d, at most `trials` evaluations;
// this stage looks at trials [k0, k1).  in: losses[k] (dose term, summed over ranks already), scale (1 or
// 1 / N), f0, pattern term of the trial (the reference evaluates it on the DIRECTION, lbfgs.py:260: a
// constant), g.d.  out[0] = alpha, out[1] = trials used, out[2] = the loss at alpha, out[3] = 1 once a
// trial has been accepted (later stages then return at once).
__global__ void armijo_select_kernel(const double* __restrict__ losses, int k0, int k1, int trials, double scale,
                                     const double* __restrict__ f0, const double* __restrict__ pattern_term,
                                     const double* __restrict__ gd, double c1, float* __restrict__ out) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    if (k0 == 0) out[3] = 0.f;
    else if (out[3] != 0.f) return;
    double alpha = exp2(-static_cast<double>(k0));
    for (int k = k0; k < k1; ++k) {
        const double f_new = static_cast<double>(static_cast<float>(losses[k] * scale)) + (pattern_term ? *pattern_term : 0.0);
        out[1] = static_cast<float>(k + 1);
        out[2] = static_cast<float>(f_new);
        if (f_new <= *f0 + c1 * alpha * (*gd)) {
            out[0] = static_cast<float>(alpha);
            out[3] = 1.f;
            return;
        }
        alpha *= 0.5;
    }
    out[0] = static_cast<float>(alpha);   // after the last trial the reference has halved once more
    (void)trials;
}

__global__ void __launch_bounds__(kLbThreads)
update_kernel(const float* __restrict__ p, const float* __restrict__ d, const float* __restrict__ alpha, size_t n,
              float* __restrict__ out) {
    const float a = *alpha;
    const size_t stride = static_cast<size_t>(gridDim.x) * kLbThreads;
    for (size_t i = static_cast<size_t>(blockIdx.x) * kLbThreads + threadIdx.x; i < n; i += stride)
        out[i] = fmaf(a, d[i], p[i]);
}

unsigned blocks_for(size_t n) {
    size_t b = (n + kLbThreads - 1) / kLbThreads;
    if (b > 148 * 8) b = 148 * 8;
    return static_cast<unsigned>(b < 1 ? 1 : b);
}

template <int kN>
void gram_launch(const VecTable& T, size_t n, double* gram, cudaStream_t st) {
    gram_kernel<kN><<<blocks_for(n), kLbThreads, 0, st>>>(T, n, gram);
}
template <int kN>
void combine_launch(const VecTable& T, const double* coef, size_t n, float* out, cudaStream_t st) {
    combine_kernel<kN><<<blocks_for(n), kLbThreads, 0, st>>>(T, coef, n, out);
}

}  // namespace

cudaError_t launch_lbfgs_history(const float* p, const float* g, float* p_old, float* g_old, float* s, float* y, size_t n,
                                 int first, cudaStream_t st) {
    history_kernel<<<blocks_for(n), kLbThreads, 0, st>>>(p, g, p_old, g_old, s, y, n, first);
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_lbfgs_gram(const float* const* vectors, int n_vec, size_t n, double* gram, cudaStream_t st) {
    if (n_vec < 1 || n_vec > kMaxVec || (n_vec & 1) == 0) return cudaErrorInvalidValue;
    VecTable T{};
    for (int i = 0; i < n_vec; ++i) T.v[i] = vectors[i];
    cudaError_t e = cudaMemsetAsync(gram, 0, sizeof(double) * n_vec * n_vec, st);
    if (e != cudaSuccess) return e;
    switch (n_vec) {
        case 1: gram_launch<1>(T, n, gram, st); break;
        case 3: gram_launch<3>(T, n, gram, st); break;
        case 5: gram_launch<5>(T, n, gram, st); break;
        case 7: gram_launch<7>(T, n, gram, st); break;
        case 9: gram_launch<9>(T, n, gram, st); break;
        case 11: gram_launch<11>(T, n, gram, st); break;
        default: gram_launch<13>(T, n, gram, st); break;
    }
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_lbfgs_solve(const double* gram, int h, int first_step, double* coef, cudaStream_t st) {
    if (h < 0 || 2 * h + 1 > kMaxVec) return cudaErrorInvalidValue;
    solve_kernel<<<1, 32, 0, st>>>(gram, h, first_step, coef);
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_lbfgs_combine(const float* const* vectors, int n_vec, const double* coef, size_t n, float* out,
                                 cudaStream_t st) {
    if (n_vec < 1 || n_vec > kMaxVec || (n_vec & 1) == 0) return cudaErrorInvalidValue;
    VecTable T{};
    for (int i = 0; i < n_vec; ++i) T.v[i] = vectors[i];
    switch (n_vec) {
        case 1: combine_launch<1>(T, coef, n, out, st); break;
        case 3: combine_launch<3>(T, coef, n, out, st); break;
        case 5: combine_launch<5>(T, coef, n, out, st); break;
        case 7: combine_launch<7>(T, coef, n, out, st); break;
        case 9: combine_launch<9>(T, coef, n, out, st); break;
        case 11: combine_launch<11>(T, coef, n, out, st); break;
        default: combine_launch<13>(T, coef, n, out, st); break;
    }
    count_launch();
    return cudaGetLastError();
}

template <int K>
static void armijo_dispatch(const float* dose, const float* ddose, const float* target, size_t n, const ArmijoParams& L,
                            double* losses, const float* skip, cudaStream_t st) {
    if (L.trials <= 4)
        armijo_losses_kernel<K, 4><<<blocks_for(n / 4), kLbThreads, 0, st>>>(dose, ddose, target, n, L, losses, skip);
    else
        armijo_losses_kernel<K, 16><<<blocks_for(n / 4), kLbThreads, 0, st>>>(dose, ddose, target, n, L, losses, skip);
}

// trials [k0, k0 + count), count <= 16; the caller has cleared `losses` before the first stage
cudaError_t launch_armijo_losses(const float* dose, const float* ddose, const float* target, size_t n, float tl, float tu,
                                 float w_obj, float w_void, float w_lim, int K, int k0, int count, double* losses,
                                 const float* skip, cudaStream_t st) {
    if (count < 1 || count > 16 || k0 < 0 || k0 + count > kMaxTrials) return cudaErrorInvalidValue;
    ArmijoParams L{tl, tu, w_obj, w_void, w_lim, k0, count};
    switch (K) {
        case 1: armijo_dispatch<1>(dose, ddose, target, n, L, losses, skip, st); break;
        case 2: armijo_dispatch<2>(dose, ddose, target, n, L, losses, skip, st); break;
        case 3: armijo_dispatch<3>(dose, ddose, target, n, L, losses, skip, st); break;
        case 4: armijo_dispatch<4>(dose, ddose, target, n, L, losses, skip, st); break;
        default: return cudaErrorInvalidValue;
    }
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_armijo_select(const double* losses, int k0, int k1, int trials, double scale, const double* f0,
                                 const double* pattern_term, const double* gd, double c1, float* out, cudaStream_t st) {
    armijo_select_kernel<<<1, 32, 0, st>>>(losses, k0, k1, trials, scale, f0, pattern_term, gd, c1, out);
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_lbfgs_update(const float* p, const float* d, const float* alpha, size_t n, float* out, cudaStream_t st) {
    update_kernel<<<blocks_for(n), kLbThreads, 0, st>>>(p, d, alpha, n, out);
    count_launch();
    return cudaGetLastError();
}

}  // namespace drtvam
