// Micro-benchmarks that measure the denominators MEASURED_PEAKS.json does not
// have: fp32 reduction-atomic throughput into a > L2 array (REDG, scalar and
// v4), and shared-memory accumulate throughput (native integer ATOMS, the
// CAS-loop fp32 atomicAdd, and plain LDS+FADD+STS).  Bench-only library
// (libdrtvam_microbench.so); not part of the product path.
#include <cstdint>
#include <cuda_runtime.h>

namespace {

__device__ __forceinline__ uint32_t hash32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16;
    return x;
}

// pattern 0: a warp adds to 32 consecutive floats (one 128 B line)
// pattern 1: a warp adds to 32 floats `stride` elements apart (one sector each)
// pattern 2: every lane adds to an independent pseudo-random float
__global__ void red_scalar_kernel(float* buf, uint32_t n_mask, int pattern, uint32_t stride, int iters) {
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t warp = tid >> 5, lane = tid & 31;
    uint32_t base = hash32(warp);
    for (int i = 0; i < iters; ++i) {
        base = hash32(base + i);
        uint32_t idx;
        if (pattern == 0) idx = (base & ~31u) + lane;
        else if (pattern == 1) idx = base + lane * stride;
        else idx = hash32(base ^ (lane * 0x9e3779b9u));
        atomicAdd(buf + (idx & n_mask), 1.0f);
    }
}

__global__ void red_v4_kernel(float* buf, uint32_t n_mask, int iters) {
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t warp = tid >> 5, lane = tid & 31;
    uint32_t base = hash32(warp);
    for (int i = 0; i < iters; ++i) {
        base = hash32(base + i);
        const uint32_t idx = ((base & ~127u) + lane * 4) & n_mask;
        asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(buf + idx), "f"(1.f), "f"(1.f), "f"(1.f),
                     "f"(1.f)
                     : "memory");
    }
}

// mode 0: atomicAdd(int) ; 1: atomicAdd(float) (CAS loop) ; 2: LDS + FADD + STS (no atomicity)
// addresses: each lane hits its own bank, a different row per iteration
__global__ void smem_accumulate_kernel(float* out, int mode, int iters) {
    extern __shared__ float tile[];
    const int n = 12288;  // 48 KB
    for (int i = threadIdx.x; i < n; i += blockDim.x) tile[i] = 0.f;
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31;
    uint32_t row = hash32(blockIdx.x * blockDim.x + threadIdx.x) >> 8;
    int* itile = reinterpret_cast<int*>(tile);
    for (int i = 0; i < iters; ++i) {
        row = row * 1664525u + 1013904223u;
        const uint32_t idx = (((row >> 10) % (n / 32)) * 32 + lane);
        if (mode == 0) atomicAdd(itile + idx, 1);
        else if (mode == 1) atomicAdd(tile + idx, 1.0f);
        else tile[idx] += 1.0f;
    }
    __syncthreads();
    if (threadIdx.x == 0) out[blockIdx.x] = tile[lane];
}


// Refined shared-memory accumulate modes with cheap address generation (two
// integer instructions per access), so that the shared-memory pipe and not
// the issue slots bound the rate.  `share` lanes use the same address
// (1 = conflict free, 2 = pairs of lanes collide, ...).
// mode 0: red.shared.add.f32 (LDS + FADD + ATOMS.CAST.SPIN)
// mode 1: red.shared.add.u32 (native ATOMS.ADD)
// mode 2: 64-bit CAS loop adding a float2
// mode 3: plain LDS + FADD + STS
// mode 4: ld.shared only (gather)
__global__ void smem_accumulate2_kernel(float* out, int mode, int iters, int share) {
    extern __shared__ float tile[];
    const int n = 12288;  // 48 KB
    for (int i = threadIdx.x; i < n; i += blockDim.x) tile[i] = 0.f;
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t base = static_cast<uint32_t>(__cvta_generic_to_shared(tile));
    uint32_t row = hash32(blockIdx.x * blockDim.x + (threadIdx.x >> 5)) & 127u;
    const uint32_t col = lane / share;
    float acc = 0.f;
    if (mode == 2) {
        for (int i = 0; i < iters; ++i) {
            row = (row + 37u) & 127u;
            const uint32_t a = base + 8u * (row * 32u + col);
            unsigned long long old, assumed, upd;
            asm volatile("ld.shared.b64 %0, [%1];" : "=l"(old) : "r"(a));
            do {
                assumed = old;
                const float lo = __uint_as_float(static_cast<uint32_t>(assumed)) + 1.0f;
                const float hi = __uint_as_float(static_cast<uint32_t>(assumed >> 32)) + 1.0f;
                upd = (static_cast<unsigned long long>(__float_as_uint(hi)) << 32) | __float_as_uint(lo);
                asm volatile("atom.shared.cas.b64 %0, [%1], %2, %3;" : "=l"(old) : "r"(a), "l"(assumed), "l"(upd) : "memory");
            } while (old != assumed);
        }
    } else {
        for (int i = 0; i < iters; ++i) {
            row = (row + 37u) & 255u;
            const uint32_t a = base + 4u * (row * 32u + col);
            if (mode == 0) asm volatile("red.shared.add.f32 [%0], %1;" ::"r"(a), "f"(1.0f) : "memory");
            else if (mode == 1) asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(a), "r"(1u) : "memory");
            else if (mode == 3) {
                float v;
                asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a));
                v += 1.0f;
                asm volatile("st.shared.f32 [%0], %1;" ::"r"(a), "f"(v) : "memory");
            } else {
                float v;
                asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a));
                acc += v;
            }
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) out[blockIdx.x] = tile[lane] + acc;
}

float time_launches(cudaEvent_t a, cudaEvent_t b) {
    float ms = 0.f;
    cudaEventSynchronize(b);
    cudaEventElapsedTime(&ms, a, b);
    return ms;
}

// Legacy tensor path: mma.sync.aligned.m16n8k8 tf32 (SASS HMMA), kIlp independent accumulator tiles per warp.
// The question it answers (patch gather as a GEMM): how many of these can a B200 SM retire?
template <int kIlp>
__global__ void mma_tf32_kernel(float* out, int iters) {
    float c[kIlp][4];
#pragma unroll
    for (int i = 0; i < kIlp; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) c[i][j] = 0.f;
    uint32_t a[4], b[2];
    const uint32_t lane = threadIdx.x & 31;
#pragma unroll
    for (int j = 0; j < 4; ++j) a[j] = __float_as_uint(1.0f + 0.125f * ((lane + j) & 7));
    b[0] = __float_as_uint(0.5f + 0.0625f * (lane & 3));
    b[1] = __float_as_uint(0.25f);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < kIlp; ++i)
            asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+f"(c[i][0]), "+f"(c[i][1]), "+f"(c[i][2]), "+f"(c[i][3])
                         : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
    }
    float acc = 0.f;
#pragma unroll
    for (int i = 0; i < kIlp; ++i) acc += c[i][0] + c[i][1] + c[i][2] + c[i][3];
    if (acc == 123.456f) out[0] = acc;
}

// The same for mma.sync.aligned.m16n8k8 f16 with fp32 accumulators (SASS HMMA.1688.F32), the instruction of the
// split-fp16 patch gather (planarg_h16_kernel).
template <int kIlp>
__global__ void mma_f16_kernel(float* out, int iters) {
    float c[kIlp][4];
#pragma unroll
    for (int i = 0; i < kIlp; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) c[i][j] = 0.f;
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t a0 = 0x3c003800u + (lane & 7), a1 = 0x38003c00u + (lane & 3), b0 = 0x34003400u + (lane & 1);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < kIlp; ++i)
            asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
                         : "+f"(c[i][0]), "+f"(c[i][1]), "+f"(c[i][2]), "+f"(c[i][3])
                         : "r"(a0), "r"(a1), "r"(b0));
    }
    float acc = 0.f;
#pragma unroll
    for (int i = 0; i < kIlp; ++i) acc += c[i][0] + c[i][1] + c[i][2] + c[i][3];
    if (acc == 123.456f) out[0] = acc;
}

// mma.sync.aligned.m16n8k16 f16 with fp32 accumulators (SASS HMMA.16816.F32): twice the flops per instruction
template <int kIlp>
__global__ void mma_f16k16_kernel(float* out, int iters) {
    float c[kIlp][4];
#pragma unroll
    for (int i = 0; i < kIlp; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) c[i][j] = 0.f;
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t a0 = 0x3c003800u + (lane & 7), a1 = 0x38003c00u + (lane & 3), b0 = 0x34003400u + (lane & 1);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < kIlp; ++i)
            asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+f"(c[i][0]), "+f"(c[i][1]), "+f"(c[i][2]), "+f"(c[i][3])
                         : "r"(a0), "r"(a1), "r"(a1), "r"(a0), "r"(b0), "r"(b0));
    }
    float acc = 0.f;
#pragma unroll
    for (int i = 0; i < kIlp; ++i) acc += c[i][0] + c[i][1] + c[i][2] + c[i][3];
    if (acc == 123.456f) out[0] = acc;
}

}  // namespace

extern "C" {

// mma.sync m16n8k16 f16 -> f32: dense TFLOP/s (2 * 16 * 8 * 16 flop per instruction and warp)
double drtvam_mb_mma_f16k16(float* out, int ilp, int iters, int blocks, int threads) {
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    for (int rep = 0; rep < 2; ++rep) {
        cudaEventRecord(a);
        if (ilp == 1) mma_f16k16_kernel<1><<<blocks, threads>>>(out, iters);
        else if (ilp == 4) mma_f16k16_kernel<4><<<blocks, threads>>>(out, iters);
        else mma_f16k16_kernel<16><<<blocks, threads>>>(out, iters);
        cudaEventRecord(b);
    }
    if (cudaGetLastError() != cudaSuccess) return -1.0;
    const float ms = time_launches(a, b);
    cudaEventDestroy(a); cudaEventDestroy(b);
    const double mmas = static_cast<double>(blocks) * (threads / 32) * iters * (ilp > 4 ? 16 : ilp);
    return mmas * 4096.0 / (ms * 1e-3) * 1e-12;
}

// mma.sync m16n8k8 f16 -> f32: dense TFLOP/s (2 * 16 * 8 * 8 flop per instruction and warp)
double drtvam_mb_mma_f16(float* out, int ilp, int iters, int blocks, int threads) {
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    for (int rep = 0; rep < 2; ++rep) {
        cudaEventRecord(a);
        if (ilp == 1) mma_f16_kernel<1><<<blocks, threads>>>(out, iters);
        else if (ilp == 4) mma_f16_kernel<4><<<blocks, threads>>>(out, iters);
        else mma_f16_kernel<16><<<blocks, threads>>>(out, iters);
        cudaEventRecord(b);
    }
    if (cudaGetLastError() != cudaSuccess) return -1.0;
    const float ms = time_launches(a, b);
    cudaEventDestroy(a); cudaEventDestroy(b);
    const double mmas = static_cast<double>(blocks) * (threads / 32) * iters * (ilp > 4 ? 16 : ilp);
    return mmas * 2048.0 / (ms * 1e-3) * 1e-12;
}

// mma.sync m16n8k8 tf32: returns dense TFLOP/s (2 * 16 * 8 * 8 flop per instruction and warp)
double drtvam_mb_mma_tf32(float* out, int ilp, int iters, int blocks, int threads) {
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    for (int rep = 0; rep < 2; ++rep) {
        cudaEventRecord(a);
        if (ilp == 1) mma_tf32_kernel<1><<<blocks, threads>>>(out, iters);
        else if (ilp == 2) mma_tf32_kernel<2><<<blocks, threads>>>(out, iters);
        else if (ilp == 4) mma_tf32_kernel<4><<<blocks, threads>>>(out, iters);
        else if (ilp == 8) mma_tf32_kernel<8><<<blocks, threads>>>(out, iters);
        else mma_tf32_kernel<16><<<blocks, threads>>>(out, iters);
        cudaEventRecord(b);
    }
    if (cudaGetLastError() != cudaSuccess) return -1.0;
    const float ms = time_launches(a, b);
    cudaEventDestroy(a); cudaEventDestroy(b);
    const double mmas = static_cast<double>(blocks) * (threads / 32) * iters * (ilp > 8 ? 16 : ilp);
    return mmas * 2048.0 / (ms * 1e-3) * 1e-12;
}

// Returns giga-adds per second (fp32 adds landed in memory), or < 0 on error.
// buf must hold (n_mask + 1 + 32*stride) floats.
double drtvam_mb_red(float* buf, uint32_t n_mask, int pattern, uint32_t stride, int iters, int blocks, int v4) {
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    const int threads = 256;
    for (int rep = 0; rep < 2; ++rep) {
        cudaEventRecord(a);
        if (v4) red_v4_kernel<<<blocks, threads>>>(buf, n_mask, iters);
        else red_scalar_kernel<<<blocks, threads>>>(buf, n_mask, pattern, stride, iters);
        cudaEventRecord(b);
    }
    if (cudaGetLastError() != cudaSuccess) return -1.0;
    const float ms = time_launches(a, b);
    cudaEventDestroy(a); cudaEventDestroy(b);
    const double adds = static_cast<double>(blocks) * threads * iters * (v4 ? 4 : 1);
    return adds / (ms * 1e-3) * 1e-9;
}

// Returns giga-accumulates per second over the whole chip.
double drtvam_mb_smem(float* out, int mode, int iters, int blocks) {
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    const int threads = 256;
    for (int rep = 0; rep < 2; ++rep) {
        cudaEventRecord(a);
        smem_accumulate_kernel<<<blocks, threads, 49152>>>(out, mode, iters);
        cudaEventRecord(b);
    }
    if (cudaGetLastError() != cudaSuccess) return -1.0;
    const float ms = time_launches(a, b);
    cudaEventDestroy(a); cudaEventDestroy(b);
    return static_cast<double>(blocks) * threads * iters / (ms * 1e-3) * 1e-9;
}

// Refined variant (cheap addressing); returns giga-accumulates per second
// (mode 2 counts two fp32 adds per access).
double drtvam_mb_smem2(float* out, int mode, int iters, int blocks, int threads, int share) {
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    for (int rep = 0; rep < 2; ++rep) {
        cudaEventRecord(a);
        smem_accumulate2_kernel<<<blocks, threads, 49152>>>(out, mode, iters, share);
        cudaEventRecord(b);
    }
    if (cudaGetLastError() != cudaSuccess) return -1.0;
    const float ms = time_launches(a, b);
    cudaEventDestroy(a); cudaEventDestroy(b);
    return static_cast<double>(blocks) * threads * iters * (mode == 2 ? 2 : 1) / (ms * 1e-3) * 1e-9;
}

}  // extern "C"
