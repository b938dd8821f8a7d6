// General per-lane tracer kernels: one lane per (active pixel, sample).
// Handles every configuration of the path (jittered sampling, aperture
// projectors, scattering, reflections, any vial built from cylinders/boxes,
// the target mesh of surface-aware runs).
//
// The production kernels are the nested ones (forward_general_kernel, adjoint_general_kernel).
// general_persistent_kernel is the experimental alternative: a warp owns a chunk of
// consecutive lanes and runs a small state machine per thread
//     {no ray} -> ray generation -> {needs a bounce} -> bounce_once -> {walking} -> ... -> retire
// so that the 32 threads spend their issue slots on voxel steps whatever
// segment each of them is on.  With the naive nesting (a DDA loop inside the
// bounce loop, forward_general_kernel below) ncu shows 7.2 of 32 threads active
// per instruction on the scattering configuration (profiles/r01d_*): every lane
// waits for the longest segment of every bounce and for the deepest path of
// the warp.  Bounces are expensive and rare (one per ~450 voxel steps), so they
// are batched: the warp only runs the bounce code when at least kSetupMin
// lanes wait for it, nobody is walking, or the waiting lanes have sat out
// kMaxWait bursts.
// Reference: VolumeIntegrator.render / render_backward
// (integrators/volume.py:18-56, 97-134).
#include <cstdlib>
#include <cstring>

#include "kernels.h"
#include "tracer.cuh"

namespace drtvam {

namespace {

constexpr int kThreads = 256;

template <bool kScatter>
__global__ void __launch_bounds__(kThreads)
forward_general_kernel(const __grid_constant__ SceneParams P, const float* __restrict__ active_data,
                       const uint32_t* __restrict__ active_pixels, uint64_t n_lanes, uint32_t spp, uint32_t seed,
                       float ray_weight, float* __restrict__ film, const int* __restrict__ gate) {
    const uint64_t lane = static_cast<uint64_t>(blockIdx.x) * kThreads + threadIdx.x;
    if (lane >= n_lanes) return;
    if (gate && *gate == 0) return;  // the planar fast path owns this call
    const uint64_t j = lane / spp;
    const float em = __ldg(active_data + j) * ray_weight;
    if (em == 0.f && !kScatter) return;  // a zero pixel deposits nothing
    Sampler rng;
    rng.seed(seed, sampler_lane(P, lane, spp));
    const Ray ray = generate_ray(P, __ldg(active_pixels + j), rng);
    ScatterSink sink{film, em};
    trace_lane<kScatter>(P, ray, rng, sink);
}

template <bool kScatter>
__global__ void __launch_bounds__(kThreads)
adjoint_general_kernel(const __grid_constant__ SceneParams P, const float* __restrict__ dL,
                       const uint32_t* __restrict__ active_pixels, uint64_t n_lanes, uint32_t spp, uint32_t seed,
                       float ray_weight, float* __restrict__ grad, const int* __restrict__ gate) {
    if (gate && *gate == 0) return;  // the planar fast path owns this call
    const uint64_t lane = static_cast<uint64_t>(blockIdx.x) * kThreads + threadIdx.x;
    const bool live = lane < n_lanes;
    const uint64_t j = live ? lane / spp : 0;
    float g = 0.f;
    if (live) {
        Sampler rng;
        rng.seed(seed, sampler_lane(P, lane, spp));
        const Ray ray = generate_ray(P, __ldg(active_pixels + j), rng);
        PipelinedGatherSink sink{dL};
        trace_lane<kScatter>(P, ray, rng, sink);
        g = sink.finish() * ray_weight;
    }
    // sum over the spp lanes of one pixel (the dr.repeat of common.py:82)
    if (spp == 1) {
        if (live) grad[j] = g;
    } else if ((spp & (spp - 1)) == 0 && spp <= 32) {
        // the spp lanes of a pixel are consecutive and aligned inside the warp
        for (uint32_t off = spp >> 1; off > 0; off >>= 1) g += __shfl_xor_sync(0xffffffffu, g, off);
        if (live && (lane % spp) == 0) grad[j] = g;
    } else {
        if (live) atomicAdd(grad + j, g);  // grad pre-zeroed by the host
    }
}

__global__ void __launch_bounds__(kThreads)
raygen_kernel(const __grid_constant__ SceneParams P, const uint32_t* __restrict__ active_pixels, uint64_t n_lanes,
              uint32_t spp, uint32_t seed, float* __restrict__ out) {
    const uint64_t lane = static_cast<uint64_t>(blockIdx.x) * kThreads + threadIdx.x;
    if (lane >= n_lanes) return;
    Sampler rng;
    rng.seed(seed, sampler_lane(P, lane, spp));
    const Ray r = generate_ray(P, __ldg(active_pixels + lane / spp), rng);
    float* o = out + 6 * lane;
    o[0] = r.ox; o[1] = r.oy; o[2] = r.oz; o[3] = r.dx; o[4] = r.dy; o[5] = r.dz;
}

template <bool kScatter>
__global__ void __launch_bounds__(kThreads)
count_general_kernel(const __grid_constant__ SceneParams P, const uint32_t* __restrict__ active_pixels,
                     uint64_t n_lanes, uint32_t spp, uint32_t seed, unsigned long long* __restrict__ total) {
    const uint64_t lane = static_cast<uint64_t>(blockIdx.x) * kThreads + threadIdx.x;
    unsigned long long n = 0;
    if (lane < n_lanes) {
        Sampler rng;
        rng.seed(seed, sampler_lane(P, lane, spp));
        const Ray ray = generate_ray(P, __ldg(active_pixels + lane / spp), rng);
        CountSink sink;
        trace_lane<kScatter>(P, ray, rng, sink);
        n = sink.n;
    }
    for (int off = 16; off > 0; off >>= 1) n += __shfl_xor_sync(0xffffffffu, n, off);
    if ((threadIdx.x & 31) == 0 && n) atomicAdd(total, n);
}

constexpr int kChunk = 512;     // consecutive lanes (rays) owned by one warp
struct PersistTuning {
    int burst;       // voxel steps between two looks at the lane states
    int setup_min;   // lanes that must wait for a bounce before the warp runs one ...
    int max_wait;    // ... unless they have waited this many bursts
};

// kMode 0: in = active_data, out = film (red.global).  kMode 1: in = dL, out = pattern gradient.
template <bool kScatter, int kMode, int kMinBlocks>
__global__ void __launch_bounds__(kThreads, kMinBlocks)
general_persistent_kernel(const __grid_constant__ SceneParams P, const float* __restrict__ in,
                          const uint32_t* __restrict__ active_pixels, uint64_t n_lanes, uint32_t spp, uint32_t seed,
                          float ray_weight, float* __restrict__ out, const int* __restrict__ gate,
                          const PersistTuning tune) {
    const int kBurst = tune.burst, kSetupMin = tune.setup_min, kMaxWait = tune.max_wait;
    if (gate && *gate == 0) return;  // the planar fast path owns this call
    constexpr unsigned kFull = 0xffffffffu;
    const unsigned lane_id = threadIdx.x & 31;
    const uint64_t warp_global = (static_cast<uint64_t>(blockIdx.x) * kThreads + threadIdx.x) >> 5;
    uint64_t next = warp_global * kChunk;
    if (next >= n_lanes) return;
    const uint64_t chunk_end = min(next + static_cast<uint64_t>(kChunk), n_lanes);

    TraceState T;
    Sampler rng;
    DdaWalk W;
    bool has_ray = false, walking = false, ray_done = false;
    uint64_t item = 0;
    float em = 0.f;
    PipelinedGatherSink gsink{in};
    int waited = 0;

    for (;;) {
        const bool need = !walking && (has_ray || next < chunk_end);
        const unsigned nm = __ballot_sync(kFull, need), km = __ballot_sync(kFull, walking);
        if (nm == 0 && km == 0) break;
        if (nm != 0 && (km == 0 || __popc(nm) >= kSetupMin || waited >= kMaxWait)) {
            waited = 0;
            // ---- hand out the next lanes of the chunk to the threads without a ray
            for (int attempt = 0; attempt < 4; ++attempt) {
                const bool want = need && !has_ray;
                const unsigned wm = __ballot_sync(kFull, want);
                if (wm == 0 || next >= chunk_end) break;
                const uint64_t mine = next + __popc(wm & ((1u << lane_id) - 1u));
                next += __popc(wm);
                if (want && mine < chunk_end) {
                    const uint64_t j = mine / spp;
                    bool take = true;
                    if (kMode == 0) {
                        em = __ldg(in + j) * ray_weight;
                        take = em != 0.f;  // a zero pixel deposits nothing
                    }
                    if (take) {
                        item = mine;
                        rng.seed(seed, sampler_lane(P, mine, spp));
                        T = TraceState();
                        T.ray = generate_ray(P, __ldg(active_pixels + j), rng);
                        has_ray = true;
                        ray_done = false;
                    }
                }
            }
            // ---- one or more bounces, until a segment has been set up or the ray is over
            if (need && has_ray) {
                bool alive = true, begun = false;
                do {
                    alive = bounce_once<kScatter>(P, T, rng, [&](const Ray& r, float maxt, float w0, bool inside,
                                                                float t_coll) {
                        if (P.sensor_kind == 2) {  // delta sensor: a single deposit, nothing to walk
                            if (kMode == 0) {
                                ScatterSink sink{out, em};
                                delta_deposit(P, r, t_coll, w0, inside, sink);
                            } else {
                                delta_deposit(P, r, t_coll, w0, inside, gsink);
                            }
                        } else if (P.sensor_kind == 1) {  // ratio tracking draws from the lane's stream: not deferred
                            if (kMode == 0) {
                                ScatterSink sink{out, em};
                                ratio_segment(P, r, maxt, w0, inside, rng, sink);
                            } else {
                                ratio_segment(P, r, maxt, w0, inside, rng, gsink);
                            }
                        } else {
                            begun = W.begin(P, r, maxt, w0, inside);
                        }
                    });
                } while (alive && !begun);
                walking = begun;
                ray_done = !alive;
            }
        } else if (nm != 0) {
            ++waited;
        }
        // ---- a burst of voxel steps
        if (kMode == 0) {
            ScatterSink sink{out, em};
#pragma unroll 1
            for (int i = 0; i < kBurst; ++i) {
                if (walking) walking = W.step(sink);
            }
        } else {
#pragma unroll 1
            for (int i = 0; i < kBurst; ++i) {
                if (walking) walking = W.step(gsink);
            }
        }
        // ---- retire the lanes whose ray is over and whose last segment has been walked
        if (has_ray && !walking && ray_done) {
            if (kMode == 1) {
                const float g = gsink.finish() * ray_weight;
                if (spp == 1) out[item] = g;
                else atomicAdd(out + item / spp, g);  // the dr.repeat of common.py:82; zeroed by the host
            }
            has_ray = false;
            ray_done = false;
        }
    }
}

// ---- two-phase tracer for scattering media (DDA sensor) ------------------------------------------
// Phase 1: every thread runs the bounce loops of kQLanes lanes WITHOUT walking anything; each medium
// segment (origin, direction, length, weight) goes to a queue in shared memory (bounce_once hands
// the segment over before anything depends on how it is deposited).  Phase 2: the threads of the
// block drain the queue, one segment per thread at a time, in bursts of voxel steps; a thread whose
// segment ends takes the next one, so the 32 threads of a warp keep stepping whatever the depth of
// the path or the length of the segment they came from.  The nested kernels above sit at 7.2 of 32
// threads active per instruction on C5 because every lane waits for the longest segment of each
// bounce and the deepest path of its warp (profiles/r01d_*).  Same lanes, same PCG32 streams, same
// deposits as the nested kernels; only the order of the (unordered) atomics differs.
constexpr int kQLanes = 2;   // lanes traced per thread
constexpr int kQCap = 3;     // queue slots per lane; a segment that finds the queue full is walked in place
constexpr int kQBurst = 16;  // voxel steps between two looks at the queue
constexpr int kQDepth = 8;   // adjoint: gathers in flight per thread

// one visit of the adjoint walk: the weight and the (not yet consumed) gathered value
struct OneVisitSink {
    static constexpr bool kRecordOnly = false;
    const float* dL;
    float w = 0.f, v = 0.f;
    __device__ __forceinline__ void visit(size_t idx, float wt) {
        w = wt;
        v = __ldg(dL + idx);
    }
};
constexpr int kQSlots = kThreads * kQLanes * kQCap;
constexpr size_t kQSmemBytes = static_cast<size_t>(kQSlots) * (2 * sizeof(float4) + sizeof(int)) +
                               sizeof(float) * kThreads * kQLanes;

// kMode 0: in = active_data, out = film (red.global).  kMode 1: in = dL, out = pattern gradient.
template <bool kScatter, int kMode>
__global__ void __launch_bounds__(kThreads)
general_queue_kernel(const __grid_constant__ SceneParams P, const float* __restrict__ in,
                     const uint32_t* __restrict__ active_pixels, uint64_t n_lanes, uint32_t spp, uint32_t seed,
                     float ray_weight, float* __restrict__ out, const int* __restrict__ gate, int hybrid) {
    extern __shared__ __align__(16) float4 q_rec[];  // two float4 per slot
    int* q_flag = reinterpret_cast<int*>(q_rec + 2 * kQSlots);
    float* grad_s = reinterpret_cast<float*>(q_flag + kQSlots);
    __shared__ int q_count, q_head;
    // hybrid (scattering scenes whose FIRST medium segments - the unscattered chords - the planar per-lane
    // kernels have already deposited / gathered, planarj_kernels.cu): gate word 0 = skip the first medium
    // segment of every lane and ADD to the gradient the planar kernels wrote; anything else = the whole lane
    const bool skip_first = hybrid && gate && *gate == 0;
    if (!hybrid && gate && *gate == 0) return;  // the planar fast path owns this call
    const int tid = threadIdx.x;
    if (tid == 0) { q_count = 0; q_head = 0; }
    if (kMode == 1)
        for (int l = 0; l < kQLanes; ++l) grad_s[l * kThreads + tid] = 0.f;
    __syncthreads();
    const uint64_t block_base = static_cast<uint64_t>(blockIdx.x) * (kThreads * kQLanes);

    // ---- phase 1: bounce loops, segments to the queue
    for (int l = 0; l < kQLanes; ++l) {
        const int local = l * kThreads + tid;
        const uint64_t lane = block_base + local;
        if (lane >= n_lanes) continue;
        const uint64_t j = lane / spp;
        float em = 1.f;
        if (kMode == 0) {
            em = __ldg(in + j) * ray_weight;
            if (em == 0.f) continue;  // every deposit of the lane is proportional to em
        }
        Sampler rng;
        rng.seed(seed, sampler_lane(P, lane, spp));
        TraceState T;
        T.ray = generate_ray(P, __ldg(active_pixels + j), rng);
        bool first = skip_first;
        while (bounce_once<kScatter>(P, T, rng, [&](const Ray& r, float maxt, float w0, bool inside, float) {
            if (first) { first = false; return; }
            const int slot = atomicAdd(&q_count, 1);
            if (slot < kQSlots) {
                q_rec[2 * slot] = make_float4(r.ox, r.oy, r.oz, maxt);
                q_rec[2 * slot + 1] = make_float4(r.dx, r.dy, r.dz, w0 * em);
                q_flag[slot] = local | (inside ? (1 << 24) : 0);
            } else if (kMode == 0) {
                ScatterSink sink{out, em};
                dda_segment(P, r, maxt, w0, inside, sink);
            } else {
                GatherSink sink{in};
                dda_segment(P, r, maxt, w0, inside, sink);
                atomicAdd(grad_s + local, sink.acc);
            }
        })) {}
    }
    __syncthreads();
    const int n_seg = min(q_count, kQSlots);

    // ---- phase 2: drain the queue
    {
        DdaWalk W;
        bool walking = false, exhausted = false, open = false;
        int owner = 0;
        ScatterSink ssink{out, 1.f};  // the record's weight carries em
        float g_acc = 0.f;
        for (;;) {
            while (!walking && !exhausted) {
                const int slot = atomicAdd(&q_head, 1);
                if (slot >= n_seg) { exhausted = true; break; }
                const float4 a = q_rec[2 * slot], b = q_rec[2 * slot + 1];
                const int flag = q_flag[slot];
                owner = flag & 0xffffff;
                Ray r;
                r.ox = a.x; r.oy = a.y; r.oz = a.z; r.dx = b.x; r.dy = b.y; r.dz = b.z;
                walking = W.begin(P, r, a.w, b.w, (flag >> 24) != 0);
                open = walking;
            }
            if (__all_sync(0xffffffffu, !walking)) break;
            if (kMode == 0) {
#pragma unroll 1
                for (int i = 0; i < kQBurst; ++i)
                    if (walking) walking = W.step(ssink);
            } else {
                // the voxel walk does not depend on what it gathers: the loads of kQDepth steps are issued before
                // the first of them is consumed, so a thread keeps kQDepth sectors in flight instead of one (the
                // gathers of scattered segments miss L1 and often L2: 1.4e11 visits/s with one load in flight per
                // thread, profiles/r02za_*)
#pragma unroll 1
                for (int r = 0; r < kQBurst / kQDepth; ++r) {
                    float w[kQDepth], v[kQDepth];
#pragma unroll
                    for (int i = 0; i < kQDepth; ++i) {
                        OneVisitSink one{in};
                        if (walking) walking = W.step(one);
                        w[i] = one.w;
                        v[i] = one.v;
                    }
#pragma unroll
                    for (int i = 0; i < kQDepth; ++i) g_acc = fmaf(w[i], v[i], g_acc);
                }
                if (open && !walking) {
                    atomicAdd(grad_s + owner, g_acc);
                    g_acc = 0.f;
                    open = false;
                }
            }
        }
    }
    if (kMode == 0) return;
    __syncthreads();

    // ---- sum over the spp lanes of one pixel (the dr.repeat of common.py:82)
    for (int l = 0; l < kQLanes; ++l) {
        const int local = l * kThreads + tid;
        const uint64_t lane = block_base + local;
        const bool live = lane < n_lanes;
        const uint64_t j = live ? lane / spp : 0;
        float g = live ? grad_s[local] * ray_weight : 0.f;
        if (spp == 1) {
            if (live) out[j] = hybrid ? out[j] + g : g;
        } else if ((spp & (spp - 1)) == 0 && spp <= 32) {
            for (uint32_t off = spp >> 1; off > 0; off >>= 1) g += __shfl_xor_sync(0xffffffffu, g, off);
            if (live && (lane % spp) == 0) out[j] = hybrid ? out[j] + g : g;
        } else {
            if (live) atomicAdd(out + j, g);  // grad pre-zeroed by the host
        }
    }
}

// 0 nested (default for absorptive media), 1 persistent (opt-in), 2 two-phase queue (default for
// scattering media with the DDA sensor).  DRTVAM_GENERAL_KERNELS = nested | persistent | queue.
inline int general_kernel_family(const SceneParams& P) {
    static const int forced = [] {
        const char* e = std::getenv("DRTVAM_GENERAL_KERNELS");
        if (e == nullptr) return -1;
        if (std::strcmp(e, "persistent") == 0) return 1;
        if (std::strcmp(e, "queue") == 0) return 2;
        if (std::strcmp(e, "nested") == 0) return 0;
        return -1;
    }();
    if (forced == 2 || forced == -1) {
        const bool fits = P.sensor_kind == 0 && P.albedo != 0.f;
        if (fits) return 2;
        return 0;
    }
    return forced;
}

template <int kMode>
cudaError_t launch_queue(const SceneParams& P, const float* in, const uint32_t* active_pixels, uint64_t n_lanes,
                         uint32_t spp, uint32_t seed, float ray_weight, float* out, cudaStream_t stream,
                         const int* gate, bool hybrid) {
    static const cudaError_t attr = cudaFuncSetAttribute(general_queue_kernel<true, kMode>,
                                                         cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                         static_cast<int>(kQSmemBytes));
    if (attr != cudaSuccess) return attr;
    const uint64_t per_block = static_cast<uint64_t>(kThreads) * kQLanes;
    const unsigned grid = static_cast<unsigned>((n_lanes + per_block - 1) / per_block);
    general_queue_kernel<true, kMode><<<grid, kThreads, kQSmemBytes, stream>>>(P, in, active_pixels, n_lanes, spp, seed,
                                                                             ray_weight, out, gate, hybrid ? 1 : 0);
    count_launch();
    return cudaGetLastError();
}

// The persistent kernel is opt-in (DRTVAM_GENERAL_KERNELS=persistent): on B200 it is not yet faster
// than the nested kernels (profiles/r01_general_tracer_sweep.txt), see DESIGN.md section 5.
inline bool use_simple_kernels() {
    static const bool simple = [] {
        const char* e = std::getenv("DRTVAM_GENERAL_KERNELS");
        return !(e != nullptr && std::strcmp(e, "persistent") == 0);
    }();
    return simple;
}

inline int env_int(const char* name, int fallback) {
    const char* e = std::getenv(name);
    return e ? std::atoi(e) : fallback;
}

template <bool kScatter, int kMode>
cudaError_t launch_persistent(const SceneParams& P, const float* in, const uint32_t* active_pixels, uint64_t n_lanes,
                              uint32_t spp, uint32_t seed, float ray_weight, float* out, cudaStream_t stream,
                              const int* gate) {
    static const int blocks = env_int("DRTVAM_PERSIST_BLOCKS", 2);
    static const PersistTuning tune{env_int("DRTVAM_PERSIST_BURST", 16), env_int("DRTVAM_PERSIST_SETUP_MIN", 8),
                                    env_int("DRTVAM_PERSIST_MAX_WAIT", 3)};
    const uint64_t warps = (n_lanes + kChunk - 1) / kChunk;
    const unsigned grid = static_cast<unsigned>((warps * 32 + kThreads - 1) / kThreads);
    if (blocks >= 4)
        general_persistent_kernel<kScatter, kMode, 4><<<grid, kThreads, 0, stream>>>(P, in, active_pixels, n_lanes, spp,
                                                                                    seed, ray_weight, out, gate, tune);
    else if (blocks == 3)
        general_persistent_kernel<kScatter, kMode, 3><<<grid, kThreads, 0, stream>>>(P, in, active_pixels, n_lanes, spp,
                                                                                    seed, ray_weight, out, gate, tune);
    else
        general_persistent_kernel<kScatter, kMode, 2><<<grid, kThreads, 0, stream>>>(P, in, active_pixels, n_lanes, spp,
                                                                                    seed, ray_weight, out, gate, tune);
    count_launch();
    return cudaGetLastError();
}

inline unsigned grid_persistent(uint64_t n_lanes) {
    const uint64_t warps = (n_lanes + kChunk - 1) / kChunk;
    return static_cast<unsigned>((warps * 32 + kThreads - 1) / kThreads);
}

inline unsigned grid_for(uint64_t n_lanes) { return static_cast<unsigned>((n_lanes + kThreads - 1) / kThreads); }

}  // namespace

cudaError_t launch_forward_general(const SceneParams& P, const float* active_data, const uint32_t* active_pixels,
                                   uint64_t n_lanes, uint32_t spp, uint32_t seed, float ray_weight, float* film,
                                   cudaStream_t stream, const int* gate, bool hybrid) {
    if (n_lanes == 0) return cudaSuccess;
    if (hybrid && !(general_kernel_family(P) == 2 && gate)) return cudaErrorInvalidValue;
    if (general_kernel_family(P) == 2)
        return launch_queue<0>(P, active_data, active_pixels, n_lanes, spp, seed, ray_weight, film, stream, gate, hybrid);
    if (!use_simple_kernels()) {
        if (P.albedo != 0.f)
            return launch_persistent<true, 0>(P, active_data, active_pixels, n_lanes, spp, seed, ray_weight, film, stream,
                                              gate);
        return launch_persistent<false, 0>(P, active_data, active_pixels, n_lanes, spp, seed, ray_weight, film, stream,
                                           gate);
    }
    if (P.albedo != 0.f)
        forward_general_kernel<true><<<grid_for(n_lanes), kThreads, 0, stream>>>(P, active_data, active_pixels,
                                                                                  n_lanes, spp, seed, ray_weight, film, gate);
    else
        forward_general_kernel<false><<<grid_for(n_lanes), kThreads, 0, stream>>>(P, active_data, active_pixels,
                                                                                   n_lanes, spp, seed, ray_weight, film, gate);
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_adjoint_general(const SceneParams& P, const float* dL, const uint32_t* active_pixels,
                                   uint64_t n_lanes, uint32_t spp, uint32_t seed, float ray_weight, float* grad,
                                   cudaStream_t stream, const int* gate, bool hybrid) {
    if (n_lanes == 0) return cudaSuccess;
    const bool queue = general_kernel_family(P) == 2;
    if (hybrid && !(queue && gate)) return cudaErrorInvalidValue;
    const bool simple = use_simple_kernels() || queue;
    const bool pow2 = (spp & (spp - 1)) == 0 && spp <= 32;
    // lanes of one pixel are summed with atomics (always in the persistent kernel, for odd spp in the
    // simple one); a gated call hands in a zeroed gradient (abi.cu)
    if (((simple && !pow2) || (!simple && spp > 1)) && gate == nullptr) {
        cudaError_t e = cudaMemsetAsync(grad, 0, sizeof(float) * (n_lanes / spp), stream);
        if (e != cudaSuccess) return e;
    }
    if (queue)
        return launch_queue<1>(P, dL, active_pixels, n_lanes, spp, seed, ray_weight, grad, stream, gate, hybrid);
    if (!simple) {
        if (P.albedo != 0.f)
            return launch_persistent<true, 1>(P, dL, active_pixels, n_lanes, spp, seed, ray_weight, grad, stream, gate);
        return launch_persistent<false, 1>(P, dL, active_pixels, n_lanes, spp, seed, ray_weight, grad, stream, gate);
    }
    if (P.albedo != 0.f)
        adjoint_general_kernel<true><<<grid_for(n_lanes), kThreads, 0, stream>>>(P, dL, active_pixels, n_lanes, spp,
                                                                                  seed, ray_weight, grad, gate);
    else
        adjoint_general_kernel<false><<<grid_for(n_lanes), kThreads, 0, stream>>>(P, dL, active_pixels, n_lanes, spp,
                                                                                   seed, ray_weight, grad, gate);
    count_launch();
    return cudaGetLastError();
}

// scattering scenes with the DDA sensor run on the two-phase queue kernels, which can leave the first medium
// segment of every lane to the planar kernels
bool general_supports_hybrid(const SceneParams& P) { return general_kernel_family(P) == 2; }

cudaError_t launch_raygen(const SceneParams& P, const uint32_t* active_pixels, uint64_t n_lanes, uint32_t spp,
                          uint32_t seed, float* out, cudaStream_t stream) {
    if (n_lanes == 0) return cudaSuccess;
    raygen_kernel<<<grid_for(n_lanes), kThreads, 0, stream>>>(P, active_pixels, n_lanes, spp, seed, out);
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_count_general(const SceneParams& P, const uint32_t* active_pixels, uint64_t n_lanes, uint32_t spp,
                                 uint32_t seed, unsigned long long* total, cudaStream_t stream) {
    if (n_lanes == 0) return cudaSuccess;
    if (P.albedo != 0.f)
        count_general_kernel<true><<<grid_for(n_lanes), kThreads, 0, stream>>>(P, active_pixels, n_lanes, spp, seed,
                                                                                total);
    else
        count_general_kernel<false><<<grid_for(n_lanes), kThreads, 0, stream>>>(P, active_pixels, n_lanes, spp, seed,
                                                                                 total);
    count_launch();
    return cudaGetLastError();
}

}  // namespace drtvam
