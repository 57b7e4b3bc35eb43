// raster_core.cuh -- device-side arithmetic shared by every forward kernel: clip, viewport
// snap, integer edge functions, tie rule, depth key, perspective-correct weights.
//
// This is the CUDA statement of docs/RASTER_SPEC.md.  It restates, as explicit arithmetic,
// what the reference delegates to the OpenGL driver through glDrawElementsBaseVertex
// (csrc/rasterise_egl.cpp:451-457) under the state set at :174-195, :248, :446-449 and the
// pass-through `smooth` shader csrc/shaders.cpp:16-34.
//
// Every floating-point operation goes through a round-to-nearest intrinsic (__fadd_rn,
// __fmul_rn, __fdiv_rn) so that nvcc never contracts a*b+c into an FMA: coverage and depth
// keys must be bit-identical to a plain IEEE evaluation on any other machine.
#pragma once
#include <cstdint>
#include <cfloat>
#include <cuda_runtime.h>

namespace dirt {

constexpr int kSubpixBits = 8;
constexpr int kSubpix = 1 << kSubpixBits;     // 256 steps per pixel
constexpr int kSubpixHalf = kSubpix / 2;
constexpr float kGuardPx = 2097152.0f;        // 2^21 px clamp on window coordinates
constexpr uint32_t kDepthMax = 16777215u;     // 2^24 - 1 (GL_DEPTH24_STENCIL8, rasterise_egl.cpp:248)
constexpr unsigned long long kKeyEmpty = 0xFFFFFFFFFFFFFFFFull;
constexpr float kDepthScaleInv = 1.0f / 16777215.0f;   // depth24 -> [0, 1]; the product is individually rounded

__device__ __forceinline__ float fadd(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float fsub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float fmul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float fdiv(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ bool is_finite(float v) { return fabsf(v) <= FLT_MAX; }

// A triangle ready to be sampled.  Corner k carries snapped window position, window depth
// and 1/w; edge k is the one opposite corner k.
struct TriSetup {
    int X0, Y0, X1, Y1, X2, Y2;   // 1/256 px, y up
    float zw0, zw1, zw2;
    float rw0, rw1, rw2;
    float zref, dzdx, dzdy;       // depth plane: zw at the centre of pixel (xmin, ymin); step per pixel in x, y
    float rcp_area;               // 1 / |2*area|
    int sign;                     // +1 / -1: sign * edge >= 0 inside
    int bias0, bias1, bias2;      // 0 = edge owns its samples, -1 = it does not
    int xmin, xmax, ymin, ymax;   // inclusive pixel bbox inside the viewport (y up)
};

__device__ __forceinline__ int snap_coord(float ndc, float half_extent)
{
    // glViewport (rasterise_egl.cpp:446): win = (ndc + 1) * extent / 2, then snap to 1/256 px
    float f = fmul(fadd(ndc, 1.0f), half_extent);
    f = f > kGuardPx ? kGuardPx : f;
    f = f < -kGuardPx ? -kGuardPx : f;
    return (int)floorf(fadd(fmul(f, (float)kSubpix), 0.5f));
}

__device__ __forceinline__ long long orient2d(int ax, int ay, int bx, int by, int px, int py)
{
    return (long long)(bx - ax) * (long long)(py - ay) - (long long)(by - ay) * (long long)(px - ax);
}

// Edge-ownership rule for samples exactly on an edge (docs/RASTER_SPEC.md section 4).
__device__ __forceinline__ int edge_bias(int dx, int dy)
{
    return ((dy < 0) || (dy == 0 && dx < 0)) ? 0 : -1;
}

// Perspective divide + viewport + snap of one corner.  Returns false when the corner cannot
// be projected (w <= 0 after clipping, or overflow).
__device__ __forceinline__ bool project_corner(float x, float y, float z, float w, float halfW, float halfH,
                                               int &X, int &Y, float &zw, float &rw)
{
    if (!(w > 0.0f)) return false;
    rw = fdiv(1.0f, w);
    if (!(rw <= FLT_MAX)) return false;
    float xn = fmul(x, rw), yn = fmul(y, rw), zn = fmul(z, rw);
    if (!is_finite(xn) || !is_finite(yn) || !is_finite(zn)) return false;
    X = snap_coord(xn, halfW);
    Y = snap_coord(yn, halfH);
    zw = fadd(fmul(zn, 0.5f), 0.5f);
    return true;
}

// Window depth is affine in window x, y (OpenGL 3.3 core eq. 3.10).  It is evaluated as a plane
// through corner 0 (docs/RASTER_SPEC.md section 6): gradients per sub-pixel unit from the two
// edge vectors, then re-based to the centre of the bbox's first pixel so that a sample costs
// two multiplies and two adds.  Every operation is individually rounded.
__device__ __forceinline__ void depth_plane(TriSetup &t, long long area2)
{
    const float fA = __ll2float_rn(area2);
    const float dz1 = fsub(t.zw1, t.zw0), dz2 = fsub(t.zw2, t.zw0);
    const float dX1 = __int2float_rn(t.X1 - t.X0), dY1 = __int2float_rn(t.Y1 - t.Y0);
    const float dX2 = __int2float_rn(t.X2 - t.X0), dY2 = __int2float_rn(t.Y2 - t.Y0);
    const float gx = fdiv(fsub(fmul(dz1, dY2), fmul(dz2, dY1)), fA);
    const float gy = fdiv(fsub(fmul(dz2, dX1), fmul(dz1, dX2)), fA);
    const float rx = __int2float_rn(t.xmin * kSubpix + kSubpixHalf - t.X0);
    const float ry = __int2float_rn(t.ymin * kSubpix + kSubpixHalf - t.Y0);
    t.zref = fadd(t.zw0, fadd(fmul(gx, rx), fmul(gy, ry)));
    t.dzdx = fmul(gx, (float)kSubpix);
    t.dzdy = fmul(gy, (float)kSubpix);
}

// Window depth at the centre of pixel (px, py), clamped to the depth range.
__device__ __forceinline__ float plane_depth(float zref, float dzdx, float dzdy, int dx, int dy)
{
    float zw = fadd(fadd(zref, fmul(dzdx, __int2float_rn(dx))), fmul(dzdy, __int2float_rn(dy)));
    return fminf(fmaxf(zw, 0.0f), 1.0f);
}

// Finish a TriSetup whose corners are already projected.  Returns false if culled.
__device__ __forceinline__ bool finish_setup(TriSetup &t, int W, int H)
{
    long long area2 = orient2d(t.X0, t.Y0, t.X1, t.Y1, t.X2, t.Y2);
    if (area2 == 0) return false;
    t.sign = area2 > 0 ? 1 : -1;
    t.rcp_area = fdiv(1.0f, __ll2float_rn(area2 > 0 ? area2 : -area2));
    t.bias0 = edge_bias((t.X2 - t.X1) * t.sign, (t.Y2 - t.Y1) * t.sign);
    t.bias1 = edge_bias((t.X0 - t.X2) * t.sign, (t.Y0 - t.Y2) * t.sign);
    t.bias2 = edge_bias((t.X1 - t.X0) * t.sign, (t.Y1 - t.Y0) * t.sign);
    int minX = min(t.X0, min(t.X1, t.X2)), maxX = max(t.X0, max(t.X1, t.X2));
    int minY = min(t.Y0, min(t.Y1, t.Y2)), maxY = max(t.Y0, max(t.Y1, t.Y2));
    // pixels whose centre px*256+128 lies inside [min, max]; scissor = viewport (:447)
    t.xmin = max((minX + kSubpixHalf - 1) >> kSubpixBits, 0);
    t.xmax = min((maxX - kSubpixHalf) >> kSubpixBits, W - 1);
    t.ymin = max((minY + kSubpixHalf - 1) >> kSubpixBits, 0);
    t.ymax = min((maxY - kSubpixHalf) >> kSubpixBits, H - 1);
    if (t.xmin > t.xmax || t.ymin > t.ymax) return false;
    depth_plane(t, area2);
    return true;
}

// Frustum classification of a face from its three clip-space corners.
//   returns 0: invisible (non-finite input or entirely outside one clip plane)
//           1: needs no depth clipping (all corners inside -w <= z <= w)
//           2: crosses the near and/or far plane -> polygon clipping path
__device__ __forceinline__ int classify_face(const float4 &a, const float4 &b, const float4 &c)
{
    bool fin = is_finite(a.x) && is_finite(a.y) && is_finite(a.z) && is_finite(a.w) &&
               is_finite(b.x) && is_finite(b.y) && is_finite(b.z) && is_finite(b.w) &&
               is_finite(c.x) && is_finite(c.y) && is_finite(c.z) && is_finite(c.w);
    if (!fin) return 0;
    int nearo = (fadd(a.z, a.w) < 0.0f) + (fadd(b.z, b.w) < 0.0f) + (fadd(c.z, c.w) < 0.0f);
    int faro = (fsub(a.w, a.z) < 0.0f) + (fsub(b.w, b.z) < 0.0f) + (fsub(c.w, c.z) < 0.0f);
    int l = (a.x < -a.w) + (b.x < -b.w) + (c.x < -c.w);
    int r = (a.x > a.w) + (b.x > b.w) + (c.x > c.w);
    int d = (a.y < -a.w) + (b.y < -b.w) + (c.y < -c.w);
    int u = (a.y > a.w) + (b.y > b.w) + (c.y > c.w);
    if (nearo == 3 || faro == 3 || l == 3 || r == 3 || d == 3 || u == 3) return 0;
    return (nearo | faro) ? 2 : 1;
}

// Setup of an unclipped face.
__device__ __forceinline__ bool setup_face(const float4 &a, const float4 &b, const float4 &c,
                                           int W, int H, TriSetup &t)
{
    float halfW = fmul(0.5f, (float)W), halfH = fmul(0.5f, (float)H);
    if (!project_corner(a.x, a.y, a.z, a.w, halfW, halfH, t.X0, t.Y0, t.zw0, t.rw0)) return false;
    if (!project_corner(b.x, b.y, b.z, b.w, halfW, halfH, t.X1, t.Y1, t.zw1, t.rw1)) return false;
    if (!project_corner(c.x, c.y, c.z, c.w, halfW, halfH, t.X2, t.Y2, t.zw2, t.rw2)) return false;
    return finish_setup(t, W, H);
}

// ---------------------------------------------------------------------------------------
// Depth-plane clipping (rare path: only faces that cross z = -w or z = w).
// ---------------------------------------------------------------------------------------
struct ClipVert {
    float x, y, z, w;
    float b0, b1, b2;       // weights w.r.t. the face's original corners
};

static __device__ __noinline__ int clip_depth_plane(const ClipVert *in, int n, ClipVert *out, int plane)
{
    float d[6];
    int m = 0;
    for (int i = 0; i < n; ++i) d[i] = plane == 0 ? fadd(in[i].z, in[i].w) : fsub(in[i].w, in[i].z);
    for (int i = 0; i < n; ++i) {
        int j = (i + 1 == n) ? 0 : i + 1;
        bool in_i = d[i] >= 0.0f, in_j = d[j] >= 0.0f;
        if (in_i) out[m++] = in[i];
        if (in_i != in_j) {
            // always walk from the inside corner to the outside one: neighbouring faces that
            // share the crossing edge then compute the identical intersection (no cracks)
            const ClipVert &s = in_i ? in[i] : in[j];
            const ClipVert &o = in_i ? in[j] : in[i];
            float ds = in_i ? d[i] : d[j], dO = in_i ? d[j] : d[i];
            float t = fdiv(ds, fsub(ds, dO));
            ClipVert v;
            v.x = fadd(s.x, fmul(t, fsub(o.x, s.x)));
            v.y = fadd(s.y, fmul(t, fsub(o.y, s.y)));
            v.z = fadd(s.z, fmul(t, fsub(o.z, s.z)));
            v.w = fadd(s.w, fmul(t, fsub(o.w, s.w)));
            v.b0 = fadd(s.b0, fmul(t, fsub(o.b0, s.b0)));
            v.b1 = fadd(s.b1, fmul(t, fsub(o.b1, s.b1)));
            v.b2 = fadd(s.b2, fmul(t, fsub(o.b2, s.b2)));
            out[m++] = v;
        }
    }
    return m;
}

// Clips a face against near then far; poly must hold 6 entries.  Returns the corner count
// (0 or 3..5).  The clipped polygon is sampled as the fan (0,s,s+1), s = 1..n-2.
static __device__ __noinline__ int clip_face(const float4 &a, const float4 &b, const float4 &c, ClipVert *poly)
{
    ClipVert tmp[6];
    poly[0] = {a.x, a.y, a.z, a.w, 1.f, 0.f, 0.f};
    poly[1] = {b.x, b.y, b.z, b.w, 0.f, 1.f, 0.f};
    poly[2] = {c.x, c.y, c.z, c.w, 0.f, 0.f, 1.f};
    int n = 3;
    bool nearo = false, faro = false;
    for (int k = 0; k < 3; ++k) {
        nearo |= fadd(poly[k].z, poly[k].w) < 0.0f;
        faro |= fsub(poly[k].w, poly[k].z) < 0.0f;
    }
    if (nearo) {
        n = clip_depth_plane(poly, n, tmp, 0);
        for (int i = 0; i < n; ++i) poly[i] = tmp[i];
        if (n < 3) return 0;
    }
    if (faro) {
        n = clip_depth_plane(poly, n, tmp, 1);
        for (int i = 0; i < n; ++i) poly[i] = tmp[i];
        if (n < 3) return 0;
    }
    return n;
}

__device__ __forceinline__ bool setup_fan(const ClipVert *poly, int s, int W, int H, TriSetup &t)
{
    float halfW = fmul(0.5f, (float)W), halfH = fmul(0.5f, (float)H);
    const ClipVert &a = poly[0], &b = poly[s], &c = poly[s + 1];
    if (!project_corner(a.x, a.y, a.z, a.w, halfW, halfH, t.X0, t.Y0, t.zw0, t.rw0)) return false;
    if (!project_corner(b.x, b.y, b.z, b.w, halfW, halfH, t.X1, t.Y1, t.zw1, t.rw1)) return false;
    if (!project_corner(c.x, c.y, c.z, c.w, halfW, halfH, t.X2, t.Y2, t.zw2, t.rw2)) return false;
    return finish_setup(t, W, H);
}

// ---------------------------------------------------------------------------------------
// Sampling
// ---------------------------------------------------------------------------------------
// The three (sign-normalised) edge values at the centre of pixel (px, py).
__device__ __forceinline__ void edge_values(const TriSetup &t, int px, int py,
                                            long long &e0, long long &e1, long long &e2)
{
    int cx = px * kSubpix + kSubpixHalf, cy = py * kSubpix + kSubpixHalf;
    e0 = orient2d(t.X1, t.Y1, t.X2, t.Y2, cx, cy) * t.sign;
    e1 = orient2d(t.X2, t.Y2, t.X0, t.Y0, cx, cy) * t.sign;
    e2 = orient2d(t.X0, t.Y0, t.X1, t.Y1, cx, cy) * t.sign;
}

// Per-pixel steps of the edge values: +1 px in x adds (sx0,sx1,sx2), +1 px in y adds (sy*).
__device__ __forceinline__ void edge_steps(const TriSetup &t, long long *sx, long long *sy)
{
    sx[0] = -(long long)(t.Y2 - t.Y1) * t.sign * kSubpix;  sy[0] = (long long)(t.X2 - t.X1) * t.sign * kSubpix;
    sx[1] = -(long long)(t.Y0 - t.Y2) * t.sign * kSubpix;  sy[1] = (long long)(t.X0 - t.X2) * t.sign * kSubpix;
    sx[2] = -(long long)(t.Y1 - t.Y0) * t.sign * kSubpix;  sy[2] = (long long)(t.X1 - t.X0) * t.sign * kSubpix;
}

__device__ __forceinline__ bool covered(const TriSetup &t, long long e0, long long e1, long long e2)
{
    return (e0 + t.bias0 >= 0) && (e1 + t.bias1 >= 0) && (e2 + t.bias2 >= 0);
}

// Window depth of the sample at pixel (px, py).
__device__ __forceinline__ float sample_depth(const TriSetup &t, int px, int py)
{
    return plane_depth(t.zref, t.dzdx, t.dzdy, px - t.xmin, py - t.ymin);
}

// Screen-space barycentrics of a covered sample from its edge values.
__device__ __forceinline__ void screen_bary(const TriSetup &t, long long e0, long long e1, long long e2,
                                            float &l0, float &l1, float &l2)
{
    l0 = fmul(__ll2float_rn(e0), t.rcp_area);
    l1 = fmul(__ll2float_rn(e1), t.rcp_area);
    l2 = fmul(__ll2float_rn(e2), t.rcp_area);
}

// Depth test key: depth24 << 32 | face (zw already clamped to [0, 1]).  Returns false when the
// fragment cannot pass GL_LESS against the cleared value 1.0 (rasterise_egl.cpp:449).
__device__ __forceinline__ bool depth_key(float zw, uint32_t face, unsigned long long &key)
{
    uint32_t d = __float2uint_rz(fadd(fmul(zw, 16777215.0f), 0.5f));
    if (d >= kDepthMax) return false;
    key = ((unsigned long long)d << 32) | face;
    return true;
}

// Perspective-correct weights of the triangle's own corners (eq. 3.9) and clip-space w.
__device__ __forceinline__ void perspective_weights(const TriSetup &t, float l0, float l1, float l2,
                                                    float &b0, float &b1, float &b2, float &w_clip)
{
    float q0 = fmul(l0, t.rw0), q1 = fmul(l1, t.rw1), q2 = fmul(l2, t.rw2);
    float qs = fadd(fadd(q0, q1), q2);
    float rq = fdiv(1.0f, qs);
    b0 = fmul(q0, rq); b1 = fmul(q1, rq); b2 = fmul(q2, rq);
    w_clip = rq;
}

}  // namespace dirt
