// geometry.cu -- the step before the rasteriser: vertex normals of a batch of
// meshes with common topology, forward and backward, replacing dirt/lighting.py:34-93.  The
// reference builds a sparse tensor with 9 * F * N entries and reduces it (dirt/lighting.py:74-88);
// here a face adds its unit normal to its three corners with vector reductions.
//
//   forward   k_face_normals    thread = (element, face): n = (v1 - v0) x (v2 - v0); n /= |n| + 1e-12
//                               (_get_face_normals, dirt/lighting.py:24-31); sums[v_k] += n
//             k_unit_normals    thread = (element, vertex): normal = s / (|s| + 1e-12)  (:89)
//   backward  k_unit_normals_grad   d/ds of s / (|s| + eps): g / (r + eps) - s (s.g) / (r (r + eps)^2)
//             k_face_normals_grad   gathers the three corner gradients, goes back through the
//                                   normalisation and the cross product, scatters to the corners
// HBM-bound: forward moves 12 B/face of indices + 3 gathers + 9 reductions per face, all L2-resident
// for meshes up to a few million vertices.
#include <cstdint>
#include <cuda_runtime.h>
#include "../../include/dirt_b200.h"

namespace dirt {

namespace {

constexpr float kEps = 1.e-12f;

struct V3 { float x, y, z; };
__device__ __forceinline__ V3 sub(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ V3 add(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ V3 scale(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
__device__ __forceinline__ float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ V3 cross(V3 a, V3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
__device__ __forceinline__ V3 load3(const float *p) { return {__ldg(p), __ldg(p + 1), __ldg(p + 2)}; }
__device__ __forceinline__ void red3(float *p, V3 v) { atomicAdd(p, v.x); atomicAdd(p + 1, v.y); atomicAdd(p + 2, v.z); }

// gradient of u / (|u| + eps) with respect to u, applied to g
__device__ __forceinline__ V3 unit_grad(V3 u, V3 g)
{
    const float r = sqrtf(dot(u, u));
    const float inv = 1.0f / (r + kEps);
    if (r == 0.0f) return scale(g, inv);
    return sub(scale(g, inv), scale(u, dot(u, g) * inv * inv / r));
}

__device__ __forceinline__ bool face_corners(const int32_t *faces, long long f, int V, int &i0, int &i1, int &i2)
{
    i0 = __ldg(faces + 3 * f); i1 = __ldg(faces + 3 * f + 1); i2 = __ldg(faces + 3 * f + 2);
    return (unsigned)i0 < (unsigned)V && (unsigned)i1 < (unsigned)V && (unsigned)i2 < (unsigned)V;
}

__global__ void __launch_bounds__(256) k_face_normals(const float *vertices, int B, int V, int D, const int32_t *faces, int F, float *sums)
{
    const long long total = (long long)B * F;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int b = (int)(i / F);
        const long long f = i - (long long)b * F;
        int i0, i1, i2;
        if (!face_corners(faces, f, V, i0, i1, i2)) continue;
        const float *vb = vertices + (size_t)b * V * D;
        const V3 p0 = load3(vb + (size_t)i0 * D), p1 = load3(vb + (size_t)i1 * D), p2 = load3(vb + (size_t)i2 * D);
        const V3 m = cross(sub(p1, p0), sub(p2, p0));
        const V3 n = scale(m, 1.0f / (sqrtf(dot(m, m)) + kEps));
        float *sb = sums + (size_t)b * V * 3;
        red3(sb + (size_t)i0 * 3, n); red3(sb + (size_t)i1 * 3, n); red3(sb + (size_t)i2 * 3, n);
    }
}

__global__ void __launch_bounds__(256) k_unit_normals(const float *sums, long long n, float *normals)
{
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const V3 s = load3(sums + 3 * i);
        const V3 u = scale(s, 1.0f / (sqrtf(dot(s, s)) + kEps));
        normals[3 * i] = u.x; normals[3 * i + 1] = u.y; normals[3 * i + 2] = u.z;
    }
}

__global__ void __launch_bounds__(256) k_unit_normals_grad(const float *sums, const float *grad_normals, long long n, float *grad_sums)
{
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const V3 g = unit_grad(load3(sums + 3 * i), load3(grad_normals + 3 * i));
        grad_sums[3 * i] = g.x; grad_sums[3 * i + 1] = g.y; grad_sums[3 * i + 2] = g.z;
    }
}

__global__ void __launch_bounds__(256) k_face_normals_grad(const float *vertices, int B, int V, int D, const int32_t *faces, int F,
                                                           const float *grad_sums, float *grad_vertices)
{
    const long long total = (long long)B * F;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int b = (int)(i / F);
        const long long f = i - (long long)b * F;
        int i0, i1, i2;
        if (!face_corners(faces, f, V, i0, i1, i2)) continue;
        const float *vb = vertices + (size_t)b * V * D;
        const float *gs = grad_sums + (size_t)b * V * 3;
        const V3 p0 = load3(vb + (size_t)i0 * D), p1 = load3(vb + (size_t)i1 * D), p2 = load3(vb + (size_t)i2 * D);
        const V3 e1 = sub(p1, p0), e2 = sub(p2, p0);
        // the face normal went to all three corners
        const V3 gn = add(add(load3(gs + (size_t)i0 * 3), load3(gs + (size_t)i1 * 3)), load3(gs + (size_t)i2 * 3));
        const V3 gm = unit_grad(cross(e1, e2), gn);
        // m = e1 x e2:  dL/de1 = e2 x gm,  dL/de2 = gm x e1
        const V3 g1 = cross(e2, gm), g2 = cross(gm, e1);
        float *gv = grad_vertices + (size_t)b * V * D;
        red3(gv + (size_t)i1 * D, g1);
        red3(gv + (size_t)i2 * D, g2);
        red3(gv + (size_t)i0 * D, scale(add(g1, g2), -1.0f));
    }
}

inline int grid_for(long long work, int sm_count)
{
    long long g = (work + 255) / 256;
    const long long cap = (long long)sm_count * 16;
    if (g > cap) g = cap;
    return g < 1 ? 1 : (int)g;
}

int sm_count_or(int fallback)
{
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return fallback;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) return fallback;
    return n;
}

}  // namespace

cudaError_t launch_vertex_normals(const float *vertices, int B, int V, int D, const int32_t *faces, int F,
                                  float *sums, float *normals, cudaStream_t stream)
{
    const int sms = sm_count_or(148);
    cudaError_t e = cudaMemsetAsync(sums, 0, (size_t)B * V * 3 * sizeof(float), stream);
    if (e != cudaSuccess) return e;
    if ((long long)B * F > 0) k_face_normals<<<grid_for((long long)B * F, sms), 256, 0, stream>>>(vertices, B, V, D, faces, F, sums);
    k_unit_normals<<<grid_for((long long)B * V, sms), 256, 0, stream>>>(sums, (long long)B * V, normals);
    return cudaGetLastError();
}

cudaError_t launch_vertex_normals_grad(const float *vertices, int B, int V, int D, const int32_t *faces, int F, const float *sums,
                                       const float *grad_normals, float *grad_sums, float *grad_vertices, cudaStream_t stream)
{
    const int sms = sm_count_or(148);
    cudaError_t e = cudaMemsetAsync(grad_vertices, 0, (size_t)B * V * D * sizeof(float), stream);
    if (e != cudaSuccess) return e;
    // sums == nullptr: the caller differentiates the un-normalised sums themselves
    // (vertex_normals_pre_split), so grad_normals already is the gradient w.r.t. the sums
    const float *gs = grad_normals;
    if (sums) {
        k_unit_normals_grad<<<grid_for((long long)B * V, sms), 256, 0, stream>>>(sums, grad_normals, (long long)B * V, grad_sums);
        gs = grad_sums;
    }
    if ((long long)B * F > 0)
        k_face_normals_grad<<<grid_for((long long)B * F, sms), 256, 0, stream>>>(vertices, B, V, D, faces, F, gs, grad_vertices);
    return cudaGetLastError();
}

}  // namespace dirt
