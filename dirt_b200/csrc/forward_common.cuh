// forward_common.cuh -- pieces shared by the forward kernels (setup_raster.cu, forward.cu).
#pragma once
#include "raster_core.cuh"
#include "kernels.h"

namespace dirt {

constexpr int kSmallExtent = (1 << 15) - 512;   // sub-pixel extent (126 px) below which a triangle is line-sampled.  Edge
                                                // functions fit in int32 (two products below 2^30 each), and at a point
                                                // of the bounding box |e| <= extent_x * extent_y < 2^30 (twice the area of
                                                // a triangle inside a rectangle), which the exact float crossing of
                                                // setup_raster.cu relies on: |e| / 256 < 2^22
constexpr int kProjSlow = (int)0x80000000;   // ProjVert.X of a vertex that needs the general (clipping) path
constexpr int kTile = 16;             // tile edge in pixels
constexpr int kTileThreads = kTile * kTile;
constexpr int kChunk = 128;           // queue entries staged per round in tile_raster
constexpr uint32_t kSubShift = 30;    // queue entry = face | fan-index << 30

struct BigTri {            // 16 bytes
    uint32_t elem;         // batch element
    uint32_t ref;          // face | fan index << 30
    uint16_t tx0, ty0, tx1, ty1;   // inclusive tile bbox (y up)
};

__device__ __forceinline__ const float4 *vertex_ptr(const float *vertices, int64_t bstride, int b)
{
    return reinterpret_cast<const float4 *>(vertices + (int64_t)b * bstride);
}

// Loads the three corners of a face; returns false if an index is outside [0, V).
__device__ __forceinline__ bool load_face(const int32_t *faces, int64_t f_bs, const float *vertices, int64_t v_bs,
                                          int b, int f, int V, float4 &a, float4 &bb, float4 &c)
{
    const int32_t *fp = faces + (int64_t)b * f_bs + 3 * (int64_t)f;
    int i0 = __ldg(fp), i1 = __ldg(fp + 1), i2 = __ldg(fp + 2);
    if ((unsigned)i0 >= (unsigned)V || (unsigned)i1 >= (unsigned)V || (unsigned)i2 >= (unsigned)V) return false;
    const float4 *vp = vertex_ptr(vertices, v_bs, b);
    a = __ldg(vp + i0); bb = __ldg(vp + i1); c = __ldg(vp + i2);
    return true;
}

// Projected vertex record written by k_project: snapped window position, window depth, 1/w.
// X == kProjSlow marks a vertex that is not plainly inside the depth range with w > 0 (or is
// not finite); faces touching one go through classify_face / clip_face on the clip-space data.
__device__ __forceinline__ const int4 *proj_ptr(const FwdParams &p, int b)
{
    return p.proj + (int64_t)b * p.proj_bs;
}

// The depth-ordered passes (depth_order.cu) pay off only where faces hide one another.
// k_depth_pivots leaves the batch sum of the per-element estimates of fragments per pixel in
// counters[kCtrOverdraw]; every later kernel of the forward pass reads the same verdict.
__device__ __forceinline__ bool plan_is_ordered(const FwdParams &p, float min_overdraw)
{
    return __uint_as_float(p.counters[kCtrOverdraw]) >= min_overdraw * (float)p.B;
}

// Generation of a usable visibility hint (1..255), or 0: none given, or its header is not that of
// this shape (first use, another mesh) -- then the faces are ordered by depth as without a hint.
__device__ __forceinline__ uint32_t hint_generation(const FwdParams &p)
{
    if (!p.hint_hdr) return 0u;
    const uint4 h = *reinterpret_cast<const uint4 *>(p.hint_hdr);
    return (h.x == kHintMagic && h.y == (uint32_t)p.B && h.z == (uint32_t)p.F && h.w >= 1u && h.w <= 255u) ? h.w : 0u;
}
__device__ __forceinline__ uint32_t hint_next_generation(uint32_t gen) { return gen % 255u + 1u; }

// TriSetup from three projected records.  Returns false if culled; `slow` is set when a corner
// needs the general path (then t is untouched).
__device__ __forceinline__ bool setup_from_proj(const int4 &a, const int4 &b, const int4 &c, int W, int H,
                                                TriSetup &t, bool &slow)
{
    slow = a.x == kProjSlow || b.x == kProjSlow || c.x == kProjSlow;
    if (slow) return false;
    t.X0 = a.x; t.Y0 = a.y; t.zw0 = __int_as_float(a.z); t.rw0 = __int_as_float(a.w);
    t.X1 = b.x; t.Y1 = b.y; t.zw1 = __int_as_float(b.z); t.rw1 = __int_as_float(b.w);
    t.X2 = c.x; t.Y2 = c.y; t.zw2 = __int_as_float(c.z); t.rw2 = __int_as_float(c.w);
    return finish_setup(t, W, H);
}

// Samples every pixel of t's bbox from one thread and merges into the key buffer.
__device__ inline void raster_serial(const TriSetup &t, uint32_t face, unsigned long long *keys_elem, int W, int H,
                              int x0, int x1, int y0, int y1)
{
    long long sx[3], sy[3], r0, r1, r2;
    edge_steps(t, sx, sy);
    edge_values(t, x0, y0, r0, r1, r2);
    for (int py = y0; py <= y1; ++py) {
        long long e0 = r0, e1 = r1, e2 = r2;
        for (int px = x0; px <= x1; ++px) {
            if (covered(t, e0, e1, e2)) {
                unsigned long long key;
                if (depth_key(sample_depth(t, px, py), face, key))
                    atomicMin(keys_elem + (size_t)(H - 1 - py) * W + px, key);
            }
            e0 += sx[0]; e1 += sx[1]; e2 += sx[2];
        }
        r0 += sy[0]; r1 += sy[1]; r2 += sy[2];
    }
}

// Appends a triangle to the large-triangle list (sampled in-thread if the list is full).
__device__ inline void push_big(const TriSetup &t, int b, uint32_t face, uint32_t fan, const FwdParams &p)
{
    uint32_t slot = atomicAdd(&p.counters[kCtrBig], 1u);
    if (slot < p.cap_big) {
        BigTri bt;
        bt.elem = (uint32_t)b;
        bt.ref = face | (fan << kSubShift);
        bt.tx0 = (uint16_t)(t.xmin / kTile); bt.tx1 = (uint16_t)(t.xmax / kTile);
        bt.ty0 = (uint16_t)(t.ymin / kTile); bt.ty1 = (uint16_t)(t.ymax / kTile);
        reinterpret_cast<uint4 *>(p.big)[slot] = *reinterpret_cast<uint4 *>(&bt);
    } else {
        // list full: still correct, just slow
        raster_serial(t, face, p.keys + (size_t)b * p.H * p.W, p.W, p.H, t.xmin, t.xmax, t.ymin, t.ymax);
    }
}


// Re-derives the TriSetup a queue / list reference stands for.
__device__ inline bool setup_from_ref(const FwdParams &p, int b, uint32_t ref, TriSetup &t)
{
    uint32_t face = ref & ((1u << kSubShift) - 1u), fan = ref >> kSubShift;
    if (fan == 0) {
        const int32_t *fp = p.faces + (int64_t)b * p.faces_bs + 3 * (int64_t)face;
        const int i0 = __ldg(fp), i1 = __ldg(fp + 1), i2 = __ldg(fp + 2);
        if ((unsigned)i0 >= (unsigned)p.V || (unsigned)i1 >= (unsigned)p.V || (unsigned)i2 >= (unsigned)p.V) return false;
        const int4 *pp = proj_ptr(p, b);
        bool slow;
        const bool ok = setup_from_proj(__ldg(pp + i0), __ldg(pp + i1), __ldg(pp + i2), p.W, p.H, t, slow);
        if (!slow) return ok;
    }
    float4 va, vb, vc;
    if (!load_face(p.faces, p.faces_bs, p.vertices, p.vertices_bs, b, (int)face, p.V, va, vb, vc)) return false;
    if (fan == 0) return setup_face(va, vb, vc, p.W, p.H, t);
    ClipVert poly[6];
    int n = clip_face(va, vb, vc, poly);
    if ((int)fan + 1 >= n) return false;
    return setup_fan(poly, (int)fan, p.W, p.H, t);
}

void launch_project(const FwdParams &p, int sm_count, cudaStream_t stream, StageTimer &tm);
void launch_setup(const FwdParams &p, int sm_count, cudaStream_t stream, uint64_t *launches, StageTimer &tm);

}  // namespace dirt
