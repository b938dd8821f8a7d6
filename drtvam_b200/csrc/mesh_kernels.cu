// Target mesh: host-side BVH build + upload, and the surface-aware sub-voxel
// volumes of VolumetricSensor.compute_volume (sensor.py:47-110).
//
// compute_volume in the reference shoots `sample_count` (2^14) random rays from
// random points of EVERY voxel and counts the points whose first hit on the
// target is a back face.  Here:
//   1. mesh_mark_kernel flags the voxels whose box overlaps a triangle's
//      bounding box (the only voxels that can hold both inside and outside
//      points);
//   2. volume_uniform_kernel classifies every unflagged voxel with ONE inside
//      test at its centre - all of its samples would agree - and appends the
//      flagged ones to a list;
//   3. volume_mc_kernel gives each listed voxel a warp; the lanes take samples
//      lane, lane + 32, ... of the voxel's own PCG32 stream (seed 0, stream =
//      flat voxel index, sensor.py:83-84) by jumping ahead 5 draws per sample,
//      so the counts are those of the reference's sequential loop.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <numeric>
#include <vector>

#include "kernels.h"
#include "rng.cuh"

namespace drtvam {

namespace {

#define MK_TRY(expr)                      \
    do {                                  \
        cudaError_t _e = (expr);          \
        if (_e != cudaSuccess) return _e; \
    } while (0)

struct BuildTri {
    float c[3];
    float lo[3], hi[3];
};

// Median split on the widest centroid axis; leaves hold up to 4 triangles.
void build_node(std::vector<BvhNode>& nodes, int node, std::vector<int>& order, const std::vector<BuildTri>& T, int lo,
                int hi) {
    float bmin[3] = {INFINITY, INFINITY, INFINITY}, bmax[3] = {-INFINITY, -INFINITY, -INFINITY};
    float cmin[3] = {INFINITY, INFINITY, INFINITY}, cmax[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int i = lo; i < hi; ++i) {
        const BuildTri& t = T[order[i]];
        for (int a = 0; a < 3; ++a) {
            bmin[a] = std::min(bmin[a], t.lo[a]); bmax[a] = std::max(bmax[a], t.hi[a]);
            cmin[a] = std::min(cmin[a], t.c[a]); cmax[a] = std::max(cmax[a], t.c[a]);
        }
    }
    for (int a = 0; a < 3; ++a) {
        // pad: a ray lying exactly in an axis-aligned face must not be culled by rounding
        const float pad = 1e-5f * (1.f + std::max(std::fabs(bmin[a]), std::fabs(bmax[a])));
        nodes[node].bmin[a] = bmin[a] - pad;
        nodes[node].bmax[a] = bmax[a] + pad;
    }
    const int n = hi - lo;
    int axis = 0;
    for (int a = 1; a < 3; ++a)
        if (cmax[a] - cmin[a] > cmax[axis] - cmin[axis]) axis = a;
    if (n <= 4 || !(cmax[axis] > cmin[axis])) {
        nodes[node].first = lo;
        nodes[node].count = n;
        return;
    }
    const int mid = lo + n / 2;
    std::nth_element(order.begin() + lo, order.begin() + mid, order.begin() + hi,
                     [&](int a, int b) { return T[a].c[axis] < T[b].c[axis]; });
    const int left = static_cast<int>(nodes.size());
    nodes.push_back(BvhNode{});
    nodes.push_back(BvhNode{});
    nodes[node].first = left;
    nodes[node].count = 0;
    build_node(nodes, left, order, T, lo, mid);
    build_node(nodes, left + 1, order, T, mid, hi);
}

// One warp per triangle: flag the voxels its bounding box touches.
__global__ void mesh_mark_kernel(const __grid_constant__ SceneParams P, unsigned char* __restrict__ flags) {
    const int tri = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (tri >= P.target.n_tri) return;
    const float4 p0 = P.target.tri[3 * tri], e1 = P.target.tri[3 * tri + 1], e2 = P.target.tri[3 * tri + 2];
    const float v[3][3] = {{p0.x, p0.y, p0.z}, {p0.x + e1.x, p0.y + e1.y, p0.z + e1.z},
                           {p0.x + e2.x, p0.y + e2.y, p0.z + e2.z}};
    int lo[3], hi[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float vs = (P.bmax[a] - P.bmin[a]) / static_cast<float>(P.res[a]);
        const float mn = fminf(fminf(v[0][a], v[1][a]), v[2][a]), mx = fmaxf(fmaxf(v[0][a], v[1][a]), v[2][a]);
        // one voxel of slack on either side absorbs every rounding of the voxel <-> position maps
        lo[a] = max(static_cast<int>(floorf((mn - P.bmin[a]) / vs)) - 1, 0);
        hi[a] = min(static_cast<int>(floorf((mx - P.bmin[a]) / vs)) + 1, P.res[a] - 1);
        if (hi[a] < lo[a]) return;
    }
    const int nx = hi[0] - lo[0] + 1, ny = hi[1] - lo[1] + 1, nz = hi[2] - lo[2] + 1;
    // plane of the triangle: a voxel whose centre is further from it than half a (padded) diagonal is clear
    float n[3] = {e1.y * e2.z - e1.z * e2.y, e1.z * e2.x - e1.x * e2.z, e1.x * e2.y - e1.y * e2.x};
    const float nl = sqrtf(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
    float half_diag = 0.f, vs3[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        vs3[a] = (P.bmax[a] - P.bmin[a]) / static_cast<float>(P.res[a]);
        half_diag += vs3[a] * vs3[a];
    }
    half_diag = 1.01f * sqrtf(half_diag);  // a full diagonal: generous
    const long long total = static_cast<long long>(nx) * ny * nz;
    for (long long i = lane; i < total; i += 32) {
        const int x = lo[0] + static_cast<int>(i % nx), y = lo[1] + static_cast<int>((i / nx) % ny);
        const int z = lo[2] + static_cast<int>(i / (static_cast<long long>(nx) * ny));
        if (nl > 0.f) {
            const float cx = P.bmin[0] + (x + 0.5f) * vs3[0] - p0.x, cy = P.bmin[1] + (y + 0.5f) * vs3[1] - p0.y;
            const float cz = P.bmin[2] + (z + 0.5f) * vs3[2] - p0.z;
            if (fabsf(cx * n[0] + cy * n[1] + cz * n[2]) > half_diag * nl) continue;
        }
        flags[static_cast<size_t>(x) + static_cast<size_t>(y) * P.res[0] +
              static_cast<size_t>(z) * P.res[0] * P.res[1]] = 1;
    }
}

// sensor.py:92-97: a point is inside when it lies in the mesh bounding box and the first hit of a
// ray from it is a back face.
__device__ __forceinline__ bool point_inside(const MeshView& M, float ox, float oy, float oz, float dx, float dy,
                                             float dz) {
    const bool in_bbox = ox > M.bmin[0] && ox < M.bmax[0] && oy > M.bmin[1] && oy < M.bmax[1] && oz > M.bmin[2] &&
                         oz < M.bmax[2];
    if (!in_bbox) return false;
    float nx, ny, nz;
    const float t = mesh_nearest(M, ox, oy, oz, dx, dy, dz, INFINITY, nx, ny, nz);
    return t < INFINITY && (dx * nx + dy * ny + dz * nz) > 0.f;
}

__global__ void __launch_bounds__(256)
volume_uniform_kernel(const __grid_constant__ SceneParams P, const unsigned char* __restrict__ flags,
                      float* __restrict__ volumes, unsigned* __restrict__ list, unsigned* __restrict__ list_count) {
    const size_t n_vox = static_cast<size_t>(P.res[0]) * P.res[1] * P.res[2];
    const size_t idx = static_cast<size_t>(blockIdx.x) * 256 + threadIdx.x;
    if (idx >= n_vox) return;
    if (flags[idx]) {
        list[atomicAdd(list_count, 1u)] = static_cast<unsigned>(idx);
        return;
    }
    const int x = static_cast<int>(idx % P.res[0]), y = static_cast<int>((idx / P.res[0]) % P.res[1]);
    const int z = static_cast<int>(idx / (static_cast<size_t>(P.res[0]) * P.res[1]));
    float vs[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) vs[a] = (P.bmax[a] - P.bmin[a]) / static_cast<float>(P.res[a]);
    const float vol = vs[0] * vs[1] * vs[2];
    // no triangle near this voxel: every point of it is on the same side as its centre
    const bool in = point_inside(P.target, P.bmin[0] + (x + 0.5f) * vs[0], P.bmin[1] + (y + 0.5f) * vs[1],
                                 P.bmin[2] + (z + 0.5f) * vs[2], 0.5377f, 0.3012f, 0.7875f);
    volumes[2 * idx] = in ? vol : 0.f;
    volumes[2 * idx + 1] = in ? 0.f : vol;
}

// PCG32 jump-ahead by `delta` draws (O(log delta)); used once per lane.
__device__ __forceinline__ void pcg_advance(Sampler& s, uint64_t delta) {
    uint64_t cur_mult = 0x5851f42d4c957f2dull, cur_plus = s.inc, acc_mult = 1u, acc_plus = 0u;
    while (delta > 0) {
        if (delta & 1) {
            acc_mult *= cur_mult;
            acc_plus = acc_plus * cur_mult + cur_plus;
        }
        cur_plus = (cur_mult + 1) * cur_plus;
        cur_mult *= cur_mult;
        delta >>= 1;
    }
    s.state = acc_mult * s.state + acc_plus;
}

__global__ void __launch_bounds__(256)
volume_mc_kernel(const __grid_constant__ SceneParams P, const unsigned* __restrict__ list, unsigned n_list,
                 uint32_t sample_count, float* __restrict__ volumes) {
    const unsigned w = (blockIdx.x * 256 + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (w >= n_list) return;
    const unsigned idx = list[w];
    const int x = idx % P.res[0], y = (idx / P.res[0]) % P.res[1], z = idx / (P.res[0] * P.res[1]);
    float vs[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) vs[a] = (P.bmax[a] - P.bmin[a]) / static_cast<float>(P.res[a]);
    Sampler rng;
    rng.seed(0u, idx);  // sensor.py:83-84
    pcg_advance(rng, 5ull * lane);
    // constants of a jump over the 5 * 31 draws the other lanes consume between two of this lane's samples
    uint64_t jm = 1u, jp = 0u;
    {
        uint64_t cm = 0x5851f42d4c957f2dull, cp = rng.inc, d = 5ull * 31;
        while (d > 0) {
            if (d & 1) { jm *= cm; jp = jp * cm + cp; }
            cp = (cm + 1) * cp;
            cm *= cm;
            d >>= 1;
        }
    }
    unsigned cin = 0;
    for (uint32_t s = lane; s < sample_count; s += 32) {
        const float o0 = rng.next(), o1 = rng.next(), o2 = rng.next();
        const float u0 = rng.next(), u1 = rng.next();
        rng.state = jm * rng.state + jp;
        // warp::square_to_uniform_sphere
        const float cz = 1.f - 2.f * u1, rr = sqrtf(fmaxf(1.f - cz * cz, 0.f));
        float sp, cp;
        sincosf(6.28318530717958647692f * u0, &sp, &cp);
        if (point_inside(P.target, P.bmin[0] + vs[0] * (x + o0), P.bmin[1] + vs[1] * (y + o1),
                         P.bmin[2] + vs[2] * (z + o2), rr * cp, rr * sp, cz))
            ++cin;
    }
    for (int off = 16; off > 0; off >>= 1) cin += __shfl_xor_sync(0xffffffffu, cin, off);
    if (lane == 0) {
        const float vol = vs[0] * vs[1] * vs[2];
        volumes[2 * static_cast<size_t>(idx)] = static_cast<float>(cin) * vol / static_cast<float>(sample_count);
        volumes[2 * static_cast<size_t>(idx) + 1] =
            static_cast<float>(sample_count - cin) * vol / static_cast<float>(sample_count);
    }
}

__global__ void mul_inplace_kernel(float* __restrict__ a, const float* __restrict__ b, size_t n) {
    const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < n) a[i] *= b[i];
}

__global__ void mul_kernel(const float* __restrict__ a, const float* __restrict__ b, float* __restrict__ out, size_t n) {
    const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < n) out[i] = a[i] * b[i];
}


// utils.py:83-128 discretize: one ray in a random direction from every voxel centre; the voxel is
// occupied when the centre lies in the mesh bounding box and the first hit is a back face.  Lane
// idx of sampler.seed(0, n_voxels) (utils.py:114-115) draws the direction (next_2d).
__global__ void __launch_bounds__(256)
discretize_kernel(const __grid_constant__ SceneParams P, float* __restrict__ voxels) {
    const size_t n_vox = static_cast<size_t>(P.res[0]) * P.res[1] * P.res[2];
    const size_t idx = static_cast<size_t>(blockIdx.x) * 256 + threadIdx.x;
    if (idx >= n_vox) return;
    const int x = static_cast<int>(idx % P.res[0]), y = static_cast<int>((idx / P.res[0]) % P.res[1]);
    const int z = static_cast<int>(idx / (static_cast<size_t>(P.res[0]) * P.res[1]));
    float vs[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) vs[a] = (P.bmax[a] - P.bmin[a]) / static_cast<float>(P.res[a]);
    Sampler rng;
    rng.seed(0u, static_cast<uint32_t>(idx));
    const float u0 = rng.next(), u1 = rng.next();
    const float cz = 1.f - 2.f * u1, rr = sqrtf(fmaxf(1.f - cz * cz, 0.f));
    float sp, cp;
    sincosf(6.28318530717958647692f * u0, &sp, &cp);
    const bool in = point_inside(P.target, P.bmin[0] + (x + 0.5f) * vs[0], P.bmin[1] + (y + 0.5f) * vs[1],
                                 P.bmin[2] + (z + 0.5f) * vs[2], rr * cp, rr * sp, cz);
    voxels[idx] = in ? 1.f : 0.f;
}
}  // namespace

void mesh_release(MeshHost& H) {
    if (H.nodes) cudaFree(H.nodes);
    if (H.tri) cudaFree(H.tri);
    H = MeshHost();
}

cudaError_t mesh_build(const float* tri_host, uint32_t n_tri, MeshHost& H) {
    mesh_release(H);
    if (n_tri == 0) return cudaSuccess;
    std::vector<BuildTri> T(n_tri);
    for (uint32_t i = 0; i < n_tri; ++i) {
        const float* v = tri_host + 9 * static_cast<size_t>(i);
        for (int a = 0; a < 3; ++a) {
            T[i].lo[a] = std::min(std::min(v[a], v[3 + a]), v[6 + a]);
            T[i].hi[a] = std::max(std::max(v[a], v[3 + a]), v[6 + a]);
            T[i].c[a] = (v[a] + v[3 + a] + v[6 + a]) / 3.f;
        }
    }
    std::vector<int> order(n_tri);
    std::iota(order.begin(), order.end(), 0);
    std::vector<BvhNode> nodes;
    nodes.reserve(2 * static_cast<size_t>(n_tri));
    nodes.push_back(BvhNode{});
    build_node(nodes, 0, order, T, 0, static_cast<int>(n_tri));
    std::vector<float4> tri(3 * static_cast<size_t>(n_tri));
    for (int a = 0; a < 3; ++a) { H.view.bmin[a] = INFINITY; H.view.bmax[a] = -INFINITY; }
    for (uint32_t i = 0; i < n_tri; ++i) {
        const float* v = tri_host + 9 * static_cast<size_t>(order[i]);
        tri[3 * i] = make_float4(v[0], v[1], v[2], 0.f);
        tri[3 * i + 1] = make_float4(v[3] - v[0], v[4] - v[1], v[5] - v[2], 0.f);
        tri[3 * i + 2] = make_float4(v[6] - v[0], v[7] - v[1], v[8] - v[2], 0.f);
        for (int a = 0; a < 3; ++a) {
            H.view.bmin[a] = std::min(H.view.bmin[a], T[order[i]].lo[a]);
            H.view.bmax[a] = std::max(H.view.bmax[a], T[order[i]].hi[a]);
        }
    }
    MK_TRY(cudaMalloc(&H.nodes, sizeof(BvhNode) * nodes.size()));
    MK_TRY(cudaMalloc(&H.tri, sizeof(float4) * tri.size()));
    MK_TRY(cudaMemcpy(H.nodes, nodes.data(), sizeof(BvhNode) * nodes.size(), cudaMemcpyHostToDevice));
    MK_TRY(cudaMemcpy(H.tri, tri.data(), sizeof(float4) * tri.size(), cudaMemcpyHostToDevice));
    H.view.nodes = H.nodes;
    H.view.tri = H.tri;
    H.view.n_tri = static_cast<int>(n_tri);
    return cudaSuccess;
}

cudaError_t launch_compute_volume(const SceneParams& P, uint32_t sample_count, float* volumes, cudaStream_t stream) {
    const size_t n_vox = static_cast<size_t>(P.res[0]) * P.res[1] * P.res[2];
    unsigned char* flags = nullptr;
    unsigned *list = nullptr, *count = nullptr;
    cudaError_t e = cudaMalloc(&flags, n_vox);
    if (e == cudaSuccess) e = cudaMalloc(&list, sizeof(unsigned) * n_vox);
    if (e == cudaSuccess) e = cudaMalloc(&count, sizeof(unsigned));
    unsigned n_list = 0;
    if (e == cudaSuccess) e = cudaMemsetAsync(flags, 0, n_vox, stream);
    if (e == cudaSuccess) e = cudaMemsetAsync(count, 0, sizeof(unsigned), stream);
    if (e == cudaSuccess) {
        const unsigned blocks = (static_cast<unsigned>(P.target.n_tri) * 32 + 255) / 256;
        mesh_mark_kernel<<<blocks, 256, 0, stream>>>(P, flags);
        count_launch();
        volume_uniform_kernel<<<static_cast<unsigned>((n_vox + 255) / 256), 256, 0, stream>>>(P, flags, volumes, list,
                                                                                                count);
        count_launch();
        e = cudaMemcpyAsync(&n_list, count, sizeof(unsigned), cudaMemcpyDeviceToHost, stream);
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
    if (e == cudaSuccess && n_list > 0) {
        const unsigned blocks = static_cast<unsigned>((static_cast<size_t>(n_list) * 32 + 255) / 256);
        volume_mc_kernel<<<blocks, 256, 0, stream>>>(P, list, n_list, sample_count, volumes);
        count_launch();
        e = cudaGetLastError();
        if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
    }
    if (flags) cudaFree(flags);
    if (list) cudaFree(list);
    if (count) cudaFree(count);
    return e;
}

cudaError_t launch_discretize(const SceneParams& P, float* voxels, cudaStream_t stream) {
    const size_t n_vox = static_cast<size_t>(P.res[0]) * P.res[1] * P.res[2];
    if (n_vox == 0) return cudaSuccess;
    discretize_kernel<<<static_cast<unsigned>((n_vox + 255) / 256), 256, 0, stream>>>(P, voxels);
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_mul_inplace(float* a, const float* b, size_t n, cudaStream_t stream) {
    if (n == 0) return cudaSuccess;
    mul_inplace_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, stream>>>(a, b, n);
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_mul(const float* a, const float* b, float* out, size_t n, cudaStream_t stream) {
    if (n == 0) return cudaSuccess;
    mul_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, stream>>>(a, b, out, n);
    count_launch();
    return cudaGetLastError();
}

}  // namespace drtvam
