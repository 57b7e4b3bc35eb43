// backward.cu -- the `RasteriseGrad` backward pass as an sm_100a CUDA kernel.
//
// Takes the role of launch_grad_assembly, which the reference fork declares but never defines
// (csrc/rasterise_grad_common.h:19-24): per pixel, from the G-buffer the backward shaders
// would render (barycentrics + clip-space w + vertex indices, csrc/shaders.cpp:2187-2221) and
// from pixels / grad_pixels, scatter-add into grad_vertices [B,V,4] and grad_vertex_colors
// [B,V,C] and write grad_background [B,H,W,C].  The gradient definition is
// docs/GRADIENT_SPEC.md ("filter-based derivatives", README.md:146-147).
//
// One thread per pixel.  Contributions to the same vertex are first summed across the warp
// (__match_any_sync on the vertex index + shuffle tree), so a warp issues one atomic per
// distinct vertex; the 4-wide position gradient goes out as a single 16-byte vector atomic.
#include <cfloat>
#include "kernels.h"

namespace dirt {

constexpr unsigned kFull = 0xFFFFFFFFu;

// Sums v[] over the lanes that hold the same key.  Returns true on the lane that ends up
// holding the group's total (the lowest lane of the group).
template <int NV>
__device__ __forceinline__ bool peer_reduce(int key, unsigned same_elem, float (&v)[NV])
{
    const int lane = threadIdx.x & 31;
    const unsigned peers = __match_any_sync(kFull, key) & same_elem;
    int rank = __popc(peers & ((1u << lane) - 1u));
    const bool leader = rank == 0;
    unsigned above = peers & ~((2u << lane) - 1u);
    while (__any_sync(kFull, above != 0u)) {
        const int next = __ffs(above) - 1;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            float t = __shfl_sync(kFull, v[i], next & 31);
            if (next >= 0) v[i] += t;
        }
        above &= __ballot_sync(kFull, (rank & 1) == 0);
        rank >>= 1;
    }
    return leader;
}

__device__ __forceinline__ float finite_or_zero(float v) { return fabsf(v) <= FLT_MAX ? v : 0.0f; }

// Which pixel's surface owns the image change seen at p along one axis (GRADIENT_SPEC 3):
// a covered neighbour on a different face with strictly smaller window depth is an occluder;
// the nearer neighbour wins, the lower-coordinate one on a tie.  Returns -1 if nobody does.
__device__ __forceinline__ long long pick_owner(const int32_t *face_id, const float4 *gbuf,
                                                long long p, int fp, float zp_self, long long m, long long q)
{
    const float zp = fp >= 0 ? zp_self : 2.0f;
    long long best = fp >= 0 ? p : -1;
    float zbest = zp;
    if (m >= 0) {
        int fc = __ldg(face_id + m);
        if (fc >= 0 && fc != fp) {
            float zc = __ldg(&gbuf[m].w);
            if (zc < zp && zc < zbest) { best = m; zbest = zc; }
        }
    }
    if (q >= 0) {
        int fc = __ldg(face_id + q);
        if (fc >= 0 && fc != fp) {
            float zc = __ldg(&gbuf[q].w);
            if (zc < zp && zc < zbest) { best = q; zbest = zc; }
        }
    }
    return best;
}

// Adds (gx, gy, 0, gw) * B_k (+ colour gradients when NCOL > 0) to the three vertices of
// `face`, warp-aggregated.  Every lane of the warp must call this; lanes with face < 0 only
// take part in the votes.
template <int NCOL>
__device__ __forceinline__ void scatter_face(const BwdParams &p, int b, unsigned same_elem, int face, float B0, float B1,
                                             float gx, float gy, float gw, const float *gcol, int C)
{
    const int32_t *fp = p.faces + (int64_t)b * p.faces_bs + 3 * (int64_t)(face < 0 ? 0 : face);
    const float Bk[3] = {B0, B1, (1.0f - B0) - B1};
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        // key: vertex index within this batch element; distinct negative keys for idle lanes.
        // same_elem keeps lanes of different batch elements apart (a warp straddles two
        // elements only when H*W is not a multiple of 32).
        int vi = face >= 0 ? __ldg(fp + k) : -1 - (int)(threadIdx.x & 31);
        float v[3 + (NCOL > 0 ? NCOL : 1)];
        v[0] = gx * Bk[k]; v[1] = gy * Bk[k]; v[2] = gw * Bk[k];
#pragma unroll
        for (int c = 0; c < NCOL; ++c) v[3 + c] = gcol[c] * Bk[k];
        bool leader = peer_reduce(vi, same_elem, v);
        if (leader && face >= 0) {
            if (v[0] != 0.0f || v[1] != 0.0f || v[2] != 0.0f) {
                float4 *dst = reinterpret_cast<float4 *>(p.grad_vertices + (int64_t)b * p.grad_vertices_bs) + vi;
                atomicAdd(dst, make_float4(v[0], v[1], 0.0f, v[2]));
            }
            if (NCOL > 0) {
                float *dc = p.grad_vertex_colors + (int64_t)b * p.grad_vertex_colors_bs + (int64_t)vi * C;
#pragma unroll
                for (int c = 0; c < NCOL; ++c) atomicAdd(dc + c, v[3 + c]);
            }
        }
    }
    (void)C;
}

// Generic-channel colour scatter (C not 1 or 3): plain atomics, no aggregation.
__device__ __forceinline__ void scatter_colors_generic(const BwdParams &p, int b, int face, float B0, float B1,
                                                       const float *g, int C)
{
    const int32_t *fp = p.faces + (int64_t)b * p.faces_bs + 3 * (int64_t)face;
    const float Bk[3] = {B0, B1, (1.0f - B0) - B1};
    for (int k = 0; k < 3; ++k) {
        float *dc = p.grad_vertex_colors + (int64_t)b * p.grad_vertex_colors_bs + (int64_t)__ldg(fp + k) * C;
        for (int c = 0; c < C; ++c) atomicAdd(dc + c, g[c] * Bk[k]);
    }
}

template <int C>
__global__ void __launch_bounds__(256) k_grad_assemble(BwdParams p)
{
    const int CC = C > 0 ? C : p.C;
    const long long npix = (long long)p.H * p.W, total = npix * p.B;
    const long long total_padded = (total + 31) & ~31ll;     // whole warps stay in the loop together
    const float4 *gbuf = reinterpret_cast<const float4 *>(p.gbuffer);
    const float halfW = 0.5f * (float)p.W, halfH = 0.5f * (float)p.H;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total_padded;
         i += (long long)gridDim.x * blockDim.x) {
        const bool live = i < total;
        const int b = live ? (int)(i / npix) : 0;
        const long long o = live ? i - (long long)b * npix : 0;
        const int r = (int)(o / p.W), x = (int)(o - (long long)r * p.W);
        const long long base = (long long)b * npix;

        int fself = -1;
        float4 gb = make_float4(0.f, 0.f, 1.f, 1.f);
        float g[C > 0 ? C : 16];
        float sx = 0.f, sr = 0.f;
        if (live) {
            fself = __ldg(p.face_id + i);
            if (fself >= 0) gb = __ldg(gbuf + i);
            const float *gp = p.grad_pixels + i * CC;
            for (int c = 0; c < CC; ++c) g[c] = __ldg(gp + c);
            // (1) background: exact
            if (p.grad_background) {
                float *gbg = p.grad_background + i * CC;
                for (int c = 0; c < CC; ++c) gbg[c] = fself < 0 ? g[c] : 0.0f;
            }
            // (2) Sobel response of `pixels`, replicate border
            const int xm = max(x - 1, 0), xp = min(x + 1, p.W - 1), rm = max(r - 1, 0), rp = min(r + 1, p.H - 1);
            const float *img = p.pixels + base * CC;
            for (int c = 0; c < CC; ++c) {
#define PX(rr, xx) __ldg(img + ((long long)(rr) * p.W + (xx)) * CC + c)
                float dx = ((PX(rm, xp) - PX(rm, xm)) + 2.0f * (PX(r, xp) - PX(r, xm)) + (PX(rp, xp) - PX(rp, xm))) * 0.125f;
                float dr = ((PX(rp, xm) - PX(rm, xm)) + 2.0f * (PX(rp, x) - PX(rm, x)) + (PX(rp, xp) - PX(rm, xp))) * 0.125f;
#undef PX
                sx += g[c] * finite_or_zero(dx);
                sr += g[c] * finite_or_zero(dr);
            }
        } else {
            for (int c = 0; c < CC; ++c) g[c] = 0.f;
        }

        // (3) owners of the x- and row-direction changes
        long long ox = -1, orow = -1;
        if (live && sx != 0.0f)
            ox = pick_owner(p.face_id + base, gbuf + base, o, fself, gb.w, x > 0 ? o - 1 : -1, x < p.W - 1 ? o + 1 : -1);
        if (live && sr != 0.0f)
            orow = pick_owner(p.face_id + base, gbuf + base, o, fself, gb.w, r > 0 ? o - p.W : -1, r < p.H - 1 ? o + p.W : -1);

        // position-gradient coefficients (GRADIENT_SPEC 4): d/dx_clip, d/dy_clip, d/dw_clip per unit B_k
        float self_gx = 0.f, self_gy = 0.f, self_gw = 0.f;
        int xface = -1, yface = -1;
        float4 xg = gb, yg = gb;
        float x_gx = 0.f, x_gw = 0.f, y_gy = 0.f, y_gw = 0.f;
        if (ox >= 0) {
            const float4 og = ox == o ? gb : __ldg(gbuf + base + ox);
            const int oxx = (int)(ox % p.W);
            const float xn = ((float)oxx + 0.5f) / halfW - 1.0f;
            const float a = sx * halfW / og.z;
            if (ox == o) { self_gx = -a; self_gw += a * xn; }
            else { xface = __ldg(p.face_id + base + ox); xg = og; x_gx = -a; x_gw = a * xn; }
        }
        if (orow >= 0) {
            const float4 og = orow == o ? gb : __ldg(gbuf + base + orow);
            const int orr = (int)(orow / p.W);
            const float yn = ((float)(p.H - 1 - orr) + 0.5f) / halfH - 1.0f;
            const float a = sr * halfH / og.z;
            if (orow == o) { self_gy = a; self_gw += -a * yn; }
            else { yface = __ldg(p.face_id + base + orow); yg = og; y_gy = a; y_gw = -a * yn; }
        }

        // (4) scatter: own face (colours + own position part), then the occluders' faces
        const unsigned same_elem = __match_any_sync(kFull, b);
        if (C == 1 || C == 3) {
            scatter_face<(C > 0 ? C : 1)>(p, b, same_elem, fself, gb.x, gb.y, self_gx, self_gy, self_gw, g, CC);
        } else {
            if (fself >= 0) scatter_colors_generic(p, b, fself, gb.x, gb.y, g, CC);
            scatter_face<0>(p, b, same_elem, fself, gb.x, gb.y, self_gx, self_gy, self_gw, nullptr, CC);
        }
        if (__any_sync(kFull, xface >= 0))
            scatter_face<0>(p, b, same_elem, xface, xg.x, xg.y, x_gx, 0.f, x_gw, nullptr, CC);
        if (__any_sync(kFull, yface >= 0))
            scatter_face<0>(p, b, same_elem, yface, yg.x, yg.y, 0.f, y_gy, y_gw, nullptr, CC);
    }
}

cudaError_t launch_backward(const BwdParams &p, int sm_count, cudaStream_t stream, uint64_t *launches, StageTimer &tm)
{
    cudaError_t e;
    const size_t nbv = p.grad_vertices_bs == 0 ? 1 : (size_t)p.B, nbc = p.grad_vertex_colors_bs == 0 ? 1 : (size_t)p.B;
    tm.begin(kStageGradClear, stream);
    if ((e = cudaMemsetAsync(p.grad_vertices, 0, nbv * p.V * 4 * sizeof(float), stream)) != cudaSuccess) return e;
    if ((e = cudaMemsetAsync(p.grad_vertex_colors, 0, nbc * p.V * p.C * sizeof(float), stream)) != cudaSuccess) return e;
    tm.end(kStageGradClear, stream);
    const long long total = (long long)p.B * p.H * p.W;
    if (total == 0) return cudaSuccess;
    long long g = (total + 255) / 256;
    tm.begin(kStageGradAssemble, stream);
    const long long cap = (long long)sm_count * 64;
    if (g > cap) g = cap;
    switch (p.C) {
        case 1: k_grad_assemble<1><<<(int)g, 256, 0, stream>>>(p); break;
        case 3: k_grad_assemble<3><<<(int)g, 256, 0, stream>>>(p); break;
        default: k_grad_assemble<0><<<(int)g, 256, 0, stream>>>(p); break;
    }
    tm.end(kStageGradAssemble, stream);
    *launches += 1;
    return cudaGetLastError();
}

}  // namespace dirt
