// Per-lane sampler: PCG32 seeded through 4-round TEA, the stream layout of
// Mitsuba's independent sampler (reference call sites:
// integrators/common.py:46-67,95,104,108; volume.py:183,200,232-233,256-257).
#pragma once
#include <cstdint>

namespace drtvam {

struct Sampler {
    uint64_t state, inc;

    __device__ __forceinline__ uint32_t next_u32() {
        const uint64_t old = state;
        state = old * 0x5851f42d4c957f2dull + inc;
        const uint32_t xorshifted = static_cast<uint32_t>(((old >> 18u) ^ old) >> 27u);
        const uint32_t rot = static_cast<uint32_t>(old >> 59u);
        return __funnelshift_r(xorshifted, xorshifted, rot);
    }
    // uniform in [0,1): mantissa trick, 23 random bits
    __device__ __forceinline__ float next() {
        return __uint_as_float((next_u32() >> 9) | 0x3f800000u) - 1.0f;
    }
    __device__ __forceinline__ void seed(uint32_t seed_value, uint32_t lane) {
        uint32_t v0 = seed_value, v1 = lane, sum = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            sum += 0x9e3779b9u;
            v0 += ((v1 << 4) + 0xa341316cu) ^ (v1 + sum) ^ ((v1 >> 5) + 0xc8013ea4u);
            v1 += ((v0 << 4) + 0xad90777du) ^ (v0 + sum) ^ ((v0 >> 5) + 0x7e95761eu);
        }
        state = 0;
        inc = (static_cast<uint64_t>(v1) << 1u) | 1u;
        next_u32();
        state += static_cast<uint64_t>(v0);
        next_u32();
    }
};

}  // namespace drtvam
