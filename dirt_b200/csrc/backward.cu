// backward.cu -- the `RasteriseGrad` backward pass as an sm_100a CUDA kernel.
//
// Takes the role of launch_grad_assembly, which the reference fork declares but never defines
// (csrc/rasterise_grad_common.h:19-24): per pixel, from the G-buffer the backward shaders
// would render (barycentrics + clip-space w + vertex indices, csrc/shaders.cpp:2187-2221) and
// from pixels / grad_pixels, scatter-add into grad_vertices [B,V,4] and grad_vertex_colors
// [B,V,C] and write grad_background [B,H,W,C].  The gradient definition is
// docs/GRADIENT_SPEC.md ("filter-based derivatives", README.md:146-147).
//
// One CTA of 256 threads per 16x12-pixel output tile.  Warps 0..5 own the interior (six 8x4 pixel
// blocks, every lane an output pixel); warps 6..7 own the 60 pixels of the one-pixel ring around it,
// which only take part in phase 1 -- so the expensive phase 2 (shuffles, atomics) runs on six full
// warps instead of eight partly idle ones.
//   phase 0  the 20x16 neighbourhood (pixels, face id, window depth) is staged in shared memory
//   phase 1  every thread: Sobel response of `pixels` weighted by grad_pixels, then the OWNER of
//            the x- and of the row-direction change (itself or the occluding neighbour) and a
//            plain store of the response into the owner's inbox slot -- a pixel has exactly
//            three possible senders per axis, so no atomics are needed and the sum is
//            deterministic
//   phase 2  every covered interior pixel folds its inbox into (d/dx, d/dy, d/dw) of its own
//            surface point and scatters ONE contribution to its own face's three vertices.
//            Lanes of a warp that hit the same face are first summed with __match_any_sync +
//            a shuffle tree (visible faces span many pixels, and a warp is an 8x4 pixel block,
//            so this removes most atomics); the position gradient goes out as one 16-byte
//            vector reduction, colours as 8-byte + 4-byte reductions when C == 3.
#include <cfloat>
#include "kernels.h"

namespace dirt {

constexpr unsigned kFull = 0xFFFFFFFFu;
#ifndef KBWD_TY
#define KBWD_TY 12
#endif
#ifndef KBWD_THREADS
#define KBWD_THREADS 256
#endif
constexpr int kThreads = KBWD_THREADS;
constexpr int kTX = 16, kTY = KBWD_TY;       // interior (output) pixels per CTA
constexpr int kRX = kTX + 2, kRY = kTY + 2;   // pixels that compute a response (interior + ring)
constexpr int kSX = kTX + 4, kSY = kTY + 4;   // staged region
constexpr int kInteriorWarps = (kTX / 8) * (kTY / 4);
constexpr int kBP = 24;                 // row pitch of the shared arrays: a warp is an 8x4 pixel block, and
                                        // 4 rows at a pitch of 24 words land in 4 disjoint groups of 8 banks
static_assert(kInteriorWarps * 32 + 2 * kRX + 2 * kTY <= kThreads && kSX <= kBP, "tile does not fit the CTA");
constexpr int kMaxC = 16;

__device__ __forceinline__ float finite_or_zero(float v) { return fabsf(v) <= FLT_MAX ? v : 0.0f; }

// Sums v[] over the lanes of `peers` (which contains this lane).  Returns true on the lowest
// lane of the group, which ends up holding the total.
template <int NV>
__device__ __forceinline__ bool peer_reduce(unsigned peers, float (&v)[NV])
{
    const int lane = threadIdx.x & 31;
    int rank = __popc(peers & ((1u << lane) - 1u));
    const bool leader = rank == 0;
    unsigned above = peers & ~((2u << lane) - 1u);      // group members on higher lanes
    while (__any_sync(kFull, above != 0u)) {
        const int next = __ffs(above) - 1;              // -1: nothing to fetch
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const float t = __shfl_sync(kFull, v[i], next & 31);
            if (next >= 0) v[i] += t;
        }
        // survivors of this round are the even-ranked members; drop the odd-ranked ones
        above &= __ballot_sync(kFull, (rank & 1) == 0);
        if (rank & 1) above = 0u;
        rank >>= 1;
    }
    return leader;
}

// Which pixel's surface owns the image change seen at a pixel along one axis (GRADIENT_SPEC 3):
// a covered neighbour on a different face with strictly smaller window depth is an occluder;
// the nearer neighbour wins, the lower-coordinate one (m) on a tie.
// Returns 0 = nobody, 1 = the lower neighbour m, 2 = the pixel itself, 3 = the upper neighbour q.
__device__ __forceinline__ int pick_owner(int fp, float zself, int fm, float zm, int fq, float zq)
{
    const float zp = fp >= 0 ? zself : 2.0f;
    int best = fp >= 0 ? 2 : 0;
    float zbest = zp;
    if (fm >= 0 && fm != fp && zm < zp && zm < zbest) { best = 1; zbest = zm; }
    if (fq >= 0 && fq != fp && zq < zp && zq < zbest) { best = 3; zbest = zq; }
    return best;
}

// 8-byte / 4-byte reductions for a 3-float colour gradient at a 4-byte aligned address.
__device__ __forceinline__ void red_add3(float *dst, float a, float b, float c)
{
    if ((reinterpret_cast<uintptr_t>(dst) & 7u) == 0) {
        atomicAdd(reinterpret_cast<float2 *>(dst), make_float2(a, b));
        atomicAdd(dst + 2, c);
    } else {
        atomicAdd(dst, a);
        atomicAdd(reinterpret_cast<float2 *>(dst + 1), make_float2(b, c));
    }
}

// Deterministic mode (DirtBackwardArgs.flags & DIRT_BACKWARD_DETERMINISTIC; GRADIENT_SPEC 5): fp32 atomics make the
// sums depend on the order in which warps arrive.  Integer addition does not: pass DET == 1 finds the largest
// |contribution| of the call (a maximum is order-free too), pass DET == 2 converts every contribution to fixed point
// with the power-of-two scale that maximum allows and accumulates in 64-bit integers, k_det_finish converts back.
// Within a warp the per-face sums are a fixed shuffle tree in every mode.
__device__ __forceinline__ float det_scale(unsigned max_bits)
{
    // |value| * scale < 2^35 and at most 2^27 contributions per sum: no overflow of the 63-bit accumulator
    int e = (int)(max_bits >> 23) - 127;
    e = e < -90 ? -90 : (e > 127 ? 127 : e);
    return ldexpf(1.0f, 33 - e);
}
__device__ __forceinline__ void det_add(long long *dst, float v, float scale)
{
    atomicAdd(reinterpret_cast<unsigned long long *>(dst), (unsigned long long)__float2ll_rn(v * scale));
}

template <int C, int DET>
#ifndef KBWD_CTAS
#define KBWD_CTAS 5
#endif
__global__ void __launch_bounds__(kThreads, KBWD_CTAS) k_grad_assemble(BwdParams p)
{
    constexpr int CS = C > 0 ? C : kMaxC;
    const int CC = C > 0 ? C : p.C;
    __shared__ float s_px[CS][kSY * kBP];
    __shared__ int s_fid[kSY * kBP];
    __shared__ float s_zw[kSY * kBP];
    __shared__ float s_in[2][3][kRY * kBP];

    const int tid = threadIdx.x;
    // a warp is an 8x4 pixel block: better odds that its lanes share faces than a 32x1 strip
    const int warp = tid >> 5, lane = tid & 31;
    // (lx, ly) in the kRX x kRY region of responding pixels; interior = [1, kTX] x [1, kTY]
    int lx, ly;
    bool has = true;
    if (warp < kInteriorWarps) {
        lx = 1 + (warp % (kTX / 8)) * 8 + (lane & 7);
        ly = 1 + (warp / (kTX / 8)) * 4 + (lane >> 3);
    } else {
        const int q = tid - kInteriorWarps * 32;      // ring: top row, bottom row, left column, right column
        if (q < kRX) { lx = q; ly = 0; }
        else if (q < 2 * kRX) { lx = q - kRX; ly = kRY - 1; }
        else if (q < 2 * kRX + kTY) { lx = 0; ly = 1 + (q - 2 * kRX); }
        else { lx = kRX - 1; ly = 1 + (q - 2 * kRX - kTY); }
        has = q < 2 * kRX + 2 * kTY;
        if (!has) { lx = 0; ly = 0; }
    }
    const int W = p.W, H = p.H;
    const int x = (int)blockIdx.x * kTX - 1 + lx, r = (int)blockIdx.y * kTY - 1 + ly;
    const float halfW = 0.5f * (float)W, halfH = 0.5f * (float)H;
    const float4 *gbuf_all = reinterpret_cast<const float4 *>(p.gbuffer);
    float det_mpos = 0.0f, det_mcol = 0.0f;                       // DET == 1: largest |contribution| seen by this thread
    const float det_spos = DET == 2 ? det_scale(p.det_max[0]) : 1.0f, det_scol = DET == 2 ? det_scale(p.det_max[1]) : 1.0f;

    for (int b = blockIdx.z; b < p.B; b += gridDim.z) {
        const size_t base = (size_t)b * H * W;
        const float *img = p.pixels + base * CC;
        // ---- phase 0: stage the 20x16 neighbourhood (replicated at the image border) --------
        for (int i = tid; i < kSX * kSY; i += kThreads) {
            const int ry = i / kSX, rx = i - ry * kSX;
            const int gx = (int)blockIdx.x * kTX - 2 + rx, gy = (int)blockIdx.y * kTY - 2 + ry;
            const bool in = gx >= 0 && gx < W && gy >= 0 && gy < H;
            const int cx = min(max(gx, 0), W - 1), cy = min(max(gy, 0), H - 1);
            const size_t o = (size_t)cy * W + cx;
            const int si = ry * kBP + rx;
            for (int c = 0; c < CC; ++c) s_px[c][si] = __ldg(img + o * CC + c);
            // both loads are issued before either is used (the depth of an uncovered pixel is ignored)
            const int f = in ? __ldg(p.face_id + base + o) : -1;
            const float zw = __ldg(&gbuf_all[base + o].w);
            s_fid[si] = f;
            s_zw[si] = f >= 0 ? zw : 2.0f;
        }
        __syncthreads();

        // ---- phase 1: Sobel response, owners, inbox -----------------------------------------
        const bool valid = has && x >= 0 && x < W && r >= 0 && r < H;
        const bool interior = valid && warp < kInteriorWarps;
        const int ri = (ly + 1) * kBP + (lx + 1);               // own position in the staged region
        const size_t o = valid ? (size_t)r * W + x : 0;
        const int fself = s_fid[ri];
        float g[CS];
        float sx = 0.f, sr = 0.f;
        if (valid) {
            const float *gp = p.grad_pixels + (base + o) * CC;
            for (int c = 0; c < CC; ++c) {
                g[c] = __ldg(gp + c);
                const float *q = &s_px[c][ri];
                // rows go down the image: q[-kBP] is the row above
                const float dx = ((q[-kBP + 1] - q[-kBP - 1]) + 2.0f * (q[1] - q[-1]) + (q[kBP + 1] - q[kBP - 1])) * 0.125f;
                const float dr = ((q[kBP - 1] - q[-kBP - 1]) + 2.0f * (q[kBP] - q[-kBP]) + (q[kBP + 1] - q[-kBP + 1])) * 0.125f;
                sx += g[c] * finite_or_zero(dx);
                sr += g[c] * finite_or_zero(dr);
            }
            if (interior && p.grad_background) {
                float *gbg = p.grad_background + (base + o) * CC;
                for (int c = 0; c < CC; ++c) gbg[c] = fself < 0 ? g[c] : 0.0f;       // exact
            }
        }
        if (has) {
            const float zs = s_zw[ri];
            const int ox = sx != 0.0f ? pick_owner(fself, zs, s_fid[ri - 1], s_zw[ri - 1], s_fid[ri + 1], s_zw[ri + 1]) : 0;
            const int oy = sr != 0.0f ? pick_owner(fself, zs, s_fid[ri - kBP], s_zw[ri - kBP], s_fid[ri + kBP], s_zw[ri + kBP]) : 0;
            const int li = ly * kBP + lx;
            // candidate c = this pixel + d, d in {-1, 0, +1}; this pixel is sender slot 1 - d of c
            if (lx > 0) s_in[0][2][li - 1] = ox == 1 ? sx : 0.0f;
            s_in[0][1][li] = ox == 2 ? sx : 0.0f;
            if (lx < kRX - 1) s_in[0][0][li + 1] = ox == 3 ? sx : 0.0f;
            if (ly > 0) s_in[1][2][li - kBP] = oy == 1 ? sr : 0.0f;
            s_in[1][1][li] = oy == 2 ? sr : 0.0f;
            if (ly < kRY - 1) s_in[1][0][li + kBP] = oy == 3 ? sr : 0.0f;
        }
        __syncthreads();

        // ---- phase 2: one scatter per covered interior pixel, summed per face across the warp --
        if (warp < kInteriorWarps) {      // warp-uniform: the ring warps have nothing to scatter
            const bool act = interior && fself >= 0;
            float B0 = 0.f, B1 = 0.f, gxv = 0.f, gyv = 0.f, gwv = 0.f;
            if (act) {
                const int li = ly * kBP + lx;
                const float4 gb = __ldg(gbuf_all + base + o);
                B0 = gb.x; B1 = gb.y;
                const float totx = (s_in[0][0][li] + s_in[0][1][li]) + s_in[0][2][li];
                const float toty = (s_in[1][0][li] + s_in[1][1][li]) + s_in[1][2][li];
                // GRADIENT_SPEC 4: a clip-space shift of the surface point by dx moves it by
                // dx * W / (2 w) pixels; +y_clip moves it UP, towards smaller rows
                const float xn = ((float)x + 0.5f) / halfW - 1.0f;
                const float yn = ((float)(H - 1 - r) + 0.5f) / halfH - 1.0f;
                const float ax = totx * halfW / gb.z, ay = toty * halfH / gb.z;
                gxv = -ax; gyv = ay; gwv = ax * xn - ay * yn;
            }
            const unsigned peers = __match_any_sync(kFull, act ? fself : -1 - lane);
            const float B2 = (1.0f - B0) - B1;
            const int32_t *fp = p.faces + (int64_t)b * p.faces_bs + 3 * (int64_t)(act ? fself : 0);
            float4 *gv = reinterpret_cast<float4 *>(p.grad_vertices + (int64_t)b * p.grad_vertices_bs);
            float *gc = p.grad_vertex_colors + (int64_t)b * p.grad_vertex_colors_bs;
            const bool packed = p.grad_packed != nullptr;
            // shared mesh: clip = world @ M_b, so grad_world[j] = sum_i grad_clip[i] * M_b[j][i]; the
            // z column never receives a gradient (GRADIENT_SPEC 4)
            const float *M = p.matrices + 16 * (int64_t)b;
            auto to_world = [&](float gx, float gy, float gw) {
                return make_float4(gx * __ldg(M + 0) + gy * __ldg(M + 1) + gw * __ldg(M + 3),
                                   gx * __ldg(M + 4) + gy * __ldg(M + 5) + gw * __ldg(M + 7),
                                   gx * __ldg(M + 8) + gy * __ldg(M + 9) + gw * __ldg(M + 11),
                                   gx * __ldg(M + 12) + gy * __ldg(M + 13) + gw * __ldg(M + 15));
            };
            if (DET == 0 && (C == 1 || C == 3)) {
                constexpr int NQ = 3 + (C > 0 ? C : 1);
                float v[3 * NQ];
    #pragma unroll
                for (int k = 0; k < 3; ++k) {
                    const float bk = k == 0 ? B0 : (k == 1 ? B1 : B2);
                    v[k * NQ] = gxv * bk; v[k * NQ + 1] = gyv * bk; v[k * NQ + 2] = gwv * bk;
    #pragma unroll
                    for (int c = 0; c < NQ - 3; ++c) v[k * NQ + 3 + c] = act ? g[c] * bk : 0.0f;
                }
                const bool leader = peer_reduce(peers, v);
                if (leader && act) {
    #pragma unroll
                    for (int k = 0; k < 3; ++k) {
                        const int vi = __ldg(fp + k);
                        const bool moved = v[k * NQ] != 0.0f || v[k * NQ + 1] != 0.0f || v[k * NQ + 2] != 0.0f;
                        if (packed) {
                            float *row = p.grad_packed + (int64_t)vi * p.packed_stride;
                            if (moved) atomicAdd(reinterpret_cast<float4 *>(row), to_world(v[k * NQ], v[k * NQ + 1], v[k * NQ + 2]));
                            if (C == 3) atomicAdd(reinterpret_cast<float4 *>(row + 4), make_float4(v[k * NQ + 3], v[k * NQ + 4], v[k * NQ + 5], 0.0f));
                            else atomicAdd(row + 4, v[k * NQ + 3]);
                        } else {
                            if (moved) atomicAdd(gv + vi, make_float4(v[k * NQ], v[k * NQ + 1], 0.0f, v[k * NQ + 2]));
                            if (C == 3) red_add3(gc + (int64_t)vi * 3, v[k * NQ + 3], v[k * NQ + 4], v[k * NQ + 5]);
                            else atomicAdd(gc + vi, v[k * NQ + 3]);
                        }
                    }
                }
            } else {
                float v[9];
    #pragma unroll
                for (int k = 0; k < 3; ++k) {
                    const float bk = k == 0 ? B0 : (k == 1 ? B1 : B2);
                    v[k * 3] = gxv * bk; v[k * 3 + 1] = gyv * bk; v[k * 3 + 2] = gwv * bk;
                }
                const bool leader = peer_reduce(peers, v) && act;
                int vi[3] = {0, 0, 0};
                if (leader) {
    #pragma unroll
                    for (int k = 0; k < 3; ++k) {
                        vi[k] = __ldg(fp + k);
                        if (v[k * 3] != 0.0f || v[k * 3 + 1] != 0.0f || v[k * 3 + 2] != 0.0f) {
                            const float4 q = packed ? to_world(v[k * 3], v[k * 3 + 1], v[k * 3 + 2]) : make_float4(v[k * 3], v[k * 3 + 1], 0.0f, v[k * 3 + 2]);
                            if (DET == 0) {
                                if (packed) atomicAdd(reinterpret_cast<float4 *>(p.grad_packed + (int64_t)vi[k] * p.packed_stride), q);
                                else atomicAdd(gv + vi[k], q);
                            } else if (DET == 1) {
                                det_mpos = fmaxf(det_mpos, fmaxf(fmaxf(fabsf(q.x), fabsf(q.y)), fmaxf(fabsf(q.z), fabsf(q.w))));
                            } else {
                                long long *a = p.det_pos + 4 * ((packed || p.grad_vertices_bs == 0 ? 0 : (int64_t)b * p.V) + vi[k]);
                                det_add(a, q.x, det_spos); det_add(a + 1, q.y, det_spos); det_add(a + 2, q.z, det_spos); det_add(a + 3, q.w, det_spos);
                            }
                        }
                    }
                }
                // colours, three channels at a time: summed per face across the warp like the position part (a deferred
                // G-buffer call has 9+ channels and few, large faces: one scalar atomic per pixel, vertex and channel
                // would all land on the same few addresses)
                for (int c0 = 0; c0 < CC; c0 += 3) {
    #pragma unroll
                    for (int k = 0; k < 3; ++k) {
                        const float bk = k == 0 ? B0 : (k == 1 ? B1 : B2);
    #pragma unroll
                        for (int j = 0; j < 3; ++j) v[k * 3 + j] = (act && c0 + j < CC) ? g[c0 + j] * bk : 0.0f;
                    }
                    peer_reduce(peers, v);
                    if (leader) {
    #pragma unroll
                        for (int k = 0; k < 3; ++k) {
                            float *dst = packed ? p.grad_packed + (int64_t)vi[k] * p.packed_stride + 4 + c0 : gc + (int64_t)vi[k] * CC + c0;
                            long long *a = DET == 2 ? p.det_col + ((packed || p.grad_vertex_colors_bs == 0 ? 0 : (int64_t)b * p.V) + vi[k]) * CC + c0 : nullptr;
    #pragma unroll
                            for (int j = 0; j < 3; ++j)
                                if (c0 + j < CC) {
                                    if (DET == 0) atomicAdd(dst + j, v[k * 3 + j]);
                                    else if (DET == 1) det_mcol = fmaxf(det_mcol, fabsf(v[k * 3 + j]));
                                    else det_add(a + j, v[k * 3 + j], det_scol);
                                }
                        }
                    }
                }
            }
        }
        __syncthreads();        // the staged region is about to be overwritten
    }
    if (DET == 1) {
        // non-negative floats order like their bit patterns
        const unsigned mp = __reduce_max_sync(kFull, __float_as_uint(det_mpos)), mc = __reduce_max_sync(kFull, __float_as_uint(det_mcol));
        if (lane == 0) {
            if (mp) atomicMax(p.det_max, mp);
            if (mc) atomicMax(p.det_max + 1, mc);
        }
    }
}

// Deterministic mode, last pass: fixed point -> float.  One thread per output value.
__global__ void __launch_bounds__(256) k_det_finish(BwdParams p)
{
    const float rpos = 1.0f / det_scale(p.det_max[0]), rcol = 1.0f / det_scale(p.det_max[1]);
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (p.grad_packed) {
        const long long n = (long long)p.V * p.packed_stride;
        if (i >= n) return;
        const int v = (int)(i / p.packed_stride), j = (int)(i - (long long)v * p.packed_stride);
        float out = 0.0f;
        if (j < 4) out = __ll2float_rn(p.det_pos[4 * (long long)v + j]) * rpos;
        else if (j < 4 + p.C) out = __ll2float_rn(p.det_col[(long long)v * p.C + (j - 4)]) * rcol;
        p.grad_packed[i] = out;
        return;
    }
    const long long nbv = p.grad_vertices_bs == 0 ? 1 : p.B, nbc = p.grad_vertex_colors_bs == 0 ? 1 : p.B;
    const long long npos = nbv * p.V * 4, ncol = nbc * p.V * p.C;
    if (i < npos) {
        const long long b = i / ((long long)p.V * 4), r = i - b * (long long)p.V * 4;
        p.grad_vertices[b * p.grad_vertices_bs + r] = __ll2float_rn(p.det_pos[i]) * rpos;
    } else if (i < npos + ncol) {
        const long long k = i - npos, b = k / ((long long)p.V * p.C), r = k - b * (long long)p.V * p.C;
        p.grad_vertex_colors[b * p.grad_vertex_colors_bs + r] = __ll2float_rn(p.det_col[k]) * rcol;
    }
}

cudaError_t launch_backward(const BwdParams &p, int sm_count, cudaStream_t stream, uint64_t *launches, StageTimer &tm)
{
    cudaError_t e;
    const size_t nbv = p.grad_vertices_bs == 0 ? 1 : (size_t)p.B, nbc = p.grad_vertex_colors_bs == 0 ? 1 : (size_t)p.B;
    tm.begin(kStageGradClear, stream);
    if (p.grad_packed) {
        if ((e = cudaMemsetAsync(p.grad_packed, 0, (size_t)p.V * p.packed_stride * sizeof(float), stream)) != cudaSuccess) return e;
    } else {
        // one [V,*] block per batch element, `stride` elements apart (the blocks may be padded: clear each one)
        const size_t wv = (size_t)p.V * 4 * sizeof(float), wc = (size_t)p.V * p.C * sizeof(float);
        if (wv && (e = cudaMemset2DAsync(p.grad_vertices, nbv > 1 ? (size_t)p.grad_vertices_bs * sizeof(float) : wv, 0, wv, nbv, stream)) != cudaSuccess) return e;
        if (wc && (e = cudaMemset2DAsync(p.grad_vertex_colors, nbc > 1 ? (size_t)p.grad_vertex_colors_bs * sizeof(float) : wc, 0, wc, nbc, stream)) != cudaSuccess) return e;
    }
    tm.end(kStageGradClear, stream);
    if ((long long)p.B * p.H * p.W == 0) return cudaSuccess;
    // each CTA walks several batch elements: tiles are small, and a grid of ~96 CTAs per SM balances
    // without paying a block launch per 192 pixels (32 per SM: 2.7 % slower, one CTA per tile and element: 4 % slower;
    // profiles/r2_setup_notes.md)
    const long long tiles = (long long)((p.W + kTX - 1) / kTX) * ((p.H + kTY - 1) / kTY);
#ifndef KBWD_WAVES
#define KBWD_WAVES 96
#endif
    long long gz = ((long long)sm_count * KBWD_WAVES + tiles - 1) / tiles;
    if (gz > p.B) gz = p.B;
    if (gz > 65535) gz = 65535;
    if (gz < 1) gz = 1;
    gz = (p.B + (p.B + gz - 1) / gz - 1) / ((p.B + gz - 1) / gz);       // same number of loop trips, spread evenly
    const dim3 grid((p.W + kTX - 1) / kTX, (p.H + kTY - 1) / kTY, (unsigned)gz);
    tm.begin(kStageGradAssemble, stream);
    if (p.det_pos) {
        // deterministic mode: measure, accumulate in fixed point, convert (the clears above are redundant but harmless)
        const size_t nacc = (p.grad_packed ? (size_t)p.V * (4 + p.C) : nbv * p.V * 4 + nbc * p.V * p.C);
        if ((e = cudaMemsetAsync(p.det_max, 0, 2 * sizeof(unsigned), stream)) != cudaSuccess) return e;
        if ((e = cudaMemsetAsync(p.det_pos, 0, nacc * sizeof(long long), stream)) != cudaSuccess) return e;
        k_grad_assemble<0, 1><<<grid, kThreads, 0, stream>>>(p);
        k_grad_assemble<0, 2><<<grid, kThreads, 0, stream>>>(p);
        const size_t nout = p.grad_packed ? (size_t)p.V * p.packed_stride : nacc;
        if (nout) k_det_finish<<<(unsigned)((nout + 255) / 256), 256, 0, stream>>>(p);
        tm.end(kStageGradAssemble, stream);
        *launches += 3;
        return cudaGetLastError();
    }
    switch (p.C) {
        case 1: k_grad_assemble<1, 0><<<grid, kThreads, 0, stream>>>(p); break;
        case 3: k_grad_assemble<3, 0><<<grid, kThreads, 0, stream>>>(p); break;
        case 9: k_grad_assemble<9, 0><<<grid, kThreads, 0, stream>>>(p); break;      // the deferred G-buffer of samples/deferred.py in one call
        default: k_grad_assemble<0, 0><<<grid, kThreads, 0, stream>>>(p); break;
    }
    tm.end(kStageGradAssemble, stream);
    *launches += 1;
    return cudaGetLastError();
}

}  // namespace dirt
