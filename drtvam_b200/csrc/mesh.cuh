// Target mesh as seen by the path: a triangle soup with a null BSDF whose
// crossings flip the inside/outside flag of the surface-aware film
// (optimize.py:66-73 'target'; integrators/volume.py:196,219,268,271) and whose
// inside test defines the sub-voxel volumes (sensor.py:47-110).
// Replaces the OptiX / Embree acceleration structure behind scene.ray_intersect
// with a flat binary BVH (built on the host in mesh_kernels.cu) walked with a
// per-thread stack.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace drtvam {

// 32 bytes.  count > 0: leaf over triangles [first, first + count).
// count == 0: inner node, children at `first` and `first + 1`.
struct BvhNode {
    float bmin[3];
    int first;
    float bmax[3];
    int count;
};

struct MeshView {
    const BvhNode* nodes;  // node 0 is the root
    const float4* tri;     // 3 per triangle: p0, e1 = p1 - p0, e2 = p2 - p0
    int n_tri;
    float bmin[3], bmax[3];
};

#ifdef __CUDACC__
// Mitsuba's Mesh::ray_intersect_triangle (Moeller-Trumbore): u, v in [0,1], u + v <= 1, t >= 0.
__device__ __forceinline__ float hit_triangle(const float4 p0, const float4 e1, const float4 e2, float ox, float oy,
                                              float oz, float dx, float dy, float dz) {
    const float px = dy * e2.z - dz * e2.y, py = dz * e2.x - dx * e2.z, pz = dx * e2.y - dy * e2.x;
    const float det = e1.x * px + e1.y * py + e1.z * pz;
    if (det == 0.f) return INFINITY;
    const float inv = 1.f / det;
    const float tx = ox - p0.x, ty = oy - p0.y, tz = oz - p0.z;
    const float u = (tx * px + ty * py + tz * pz) * inv;
    if (!(u >= 0.f && u <= 1.f)) return INFINITY;
    const float qx = ty * e1.z - tz * e1.y, qy = tz * e1.x - tx * e1.z, qz = tx * e1.y - ty * e1.x;
    const float v = (dx * qx + dy * qy + dz * qz) * inv;
    if (!(v >= 0.f && u + v <= 1.f)) return INFINITY;
    const float t = (e2.x * qx + e2.y * qy + e2.z * qz) * inv;
    return t >= 0.f ? t : INFINITY;
}

// Nearest hit with t < tmax; returns t (or +inf) and the unit geometric normal e1 x e2.
__device__ __forceinline__ float mesh_nearest(const MeshView& M, float ox, float oy, float oz, float dx, float dy,
                                              float dz, float tmax, float& nx, float& ny, float& nz) {
    float best = tmax;
    int best_tri = -1;
    const float ix = 1.f / dx, iy = 1.f / dy, iz = 1.f / dz;
    int stack[32];
    int sp = 0;
    stack[sp++] = 0;
    while (sp > 0) {
        const BvhNode nd = M.nodes[stack[--sp]];
        // slab test; the box is padded a little so that a ray lying in a face plane is not lost to rounding
        float t0 = (nd.bmin[0] - ox) * ix, t1 = (nd.bmax[0] - ox) * ix;
        float tn = fminf(t0, t1), tf = fmaxf(t0, t1);
        t0 = (nd.bmin[1] - oy) * iy; t1 = (nd.bmax[1] - oy) * iy;
        tn = fmaxf(tn, fminf(t0, t1)); tf = fminf(tf, fmaxf(t0, t1));
        t0 = (nd.bmin[2] - oz) * iz; t1 = (nd.bmax[2] - oz) * iz;
        tn = fmaxf(tn, fminf(t0, t1)); tf = fminf(tf, fmaxf(t0, t1));
        if (!(tn <= tf && tf >= 0.f && tn < best)) continue;
        if (nd.count > 0) {
            for (int i = nd.first; i < nd.first + nd.count; ++i) {
                const float t = hit_triangle(__ldg(M.tri + 3 * i), __ldg(M.tri + 3 * i + 1), __ldg(M.tri + 3 * i + 2),
                                             ox, oy, oz, dx, dy, dz);
                if (t < best) { best = t; best_tri = i; }
            }
        } else if (sp <= 30) {
            stack[sp++] = nd.first;
            stack[sp++] = nd.first + 1;
        }
    }
    if (best_tri < 0) return INFINITY;
    const float4 e1 = __ldg(M.tri + 3 * best_tri + 1), e2 = __ldg(M.tri + 3 * best_tri + 2);
    const float cx = e1.y * e2.z - e1.z * e2.y, cy = e1.z * e2.x - e1.x * e2.z, cz = e1.x * e2.y - e1.y * e2.x;
    const float inv = rsqrtf(cx * cx + cy * cy + cz * cz);
    nx = cx * inv; ny = cy * inv; nz = cz * inv;
    return best;
}
#endif

}  // namespace drtvam
