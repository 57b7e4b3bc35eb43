// setup_raster.cu -- stages S0b/S1 of the forward pass: vertex projection, triangle setup and
// sampling of every triangle that is not large enough to be worth binning.
//
// k_project  one thread per (element, vertex): perspective divide, viewport transform, snap to
//            1/256 px, window depth, 1/w -> a 16-byte record.  A vertex is projected once
//            instead of once per incident face (~6x fewer divides on a closed mesh).
// k_setup    lane t of a warp sets up face t of a group of 32 faces from three gathered
//            records (int32 edge functions: every triangle on this path fits in 126 px).
//            Sampling is then re-distributed inside the warp, because "one lane walks one
//            triangle" leaves most lanes idle on long thin triangles
//            (profiles/r1_setup_notes.md): the unit of work is a scan LINE of a triangle --
//            a row or a column, whichever makes the lines' spans short (slivers are scanned
//            across their narrow side).  A warp prefix sum over the lines per triangle, one
//            OR-reduction of start bits and two population counts hand 32 lines at a time to the 32 lanes.  Each lane
//            computes its line's EXACT covered span (each edge crossing is a ceiling division
//            evaluated in float in a form that is provably exact for these operand ranges, so
//            coverage stays an integer decision) and walks it: plane depth, depth|face key, a plain load
//            of the current key, and a 64-bit atomicMin only when the fragment would win.
//            With depth complexity d only ~H(d) fragments per pixel reach the atomic unit.
// Triangles wider than 126 px go to the large-triangle list (S2/S3, forward.cu); faces that
// cross the near / far plane take the general clipping path.
#include <cstdio>
#include "forward_common.cuh"

namespace dirt {

constexpr unsigned kFullMask = 0xFFFFFFFFu;
constexpr int kWarpsPerCta = 8;
// Resident CTAs per SM and fragments per round, per launch kind (A/B in profiles/r2_setup_notes.md):
// the blind launch is bound by the reductions it issues, not by latency, and runs best with 64
// registers (4 CTAs); the pyramid-tested launches wait on depth loads and want occupancy (6 CTAs).
#ifndef KCTAS
#define KCTAS 4
#endif
#ifndef KCTAS_HIZ
#define KCTAS_HIZ 6
#endif
#ifndef KFRAG
#define KFRAG 4
#endif
#ifndef KFRAG_HIZ
#define KFRAG_HIZ 4
#endif
#ifndef KPRE0
#define KPRE0 0
#endif
#ifndef KCHUNK_TWO
#define KCHUNK_TWO 1     // chunk test: the pyramid words of the first two tiles are fetched together, before either is compared
#endif
#ifndef KCHUNK_MIN
#define KCHUNK_MIN 0     // faces of at most this many lines skip the chunk test (their lines are sampled directly)
#endif
#ifndef KLDCG
#define KLDCG 0      // 1: the depth pre-load bypasses L1 (sees the reductions of this launch)
#endif
constexpr int kCtasPerSm = KCTAS;                     // launches without the pyramid test (the blind first launch, unordered meshes)
constexpr int kCtasPerSmHiz = KCTAS_HIZ;              // the launches that test against the pyramid
constexpr int ctas_per_sm(bool hiz) { return hiz ? kCtasPerSmHiz : kCtasPerSm; }
constexpr int kFragRoundBlind = KFRAG;                // launch 0: no pre-load, the round only unrolls the loop
constexpr int kFragRoundHiz = KFRAG_HIZ;              // depths fetched together per round; spans are 1-2 pixels on 3 lines out of 4 (profiles/r1_setup_notes.md)
constexpr long long kFewFaces = 4096;
constexpr float kLineRcpAreaMin = 1.0f / (2.0f * 65536.0f * 1024.0f);   // 1 / |2 area| of a 1024 px triangle: fatter ones go to the tile path

// ---------------------------------------------------------------------------------------
// S0b: vertex projection
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_project(FwdParams p, int nbatch)
{
    const float halfW = fmul(0.5f, (float)p.W), halfH = fmul(0.5f, (float)p.H);
    const long long total = (long long)nbatch * p.V;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int b = (int)(i / p.V), v = (int)(i - (long long)b * p.V);
        const float4 c = __ldg(vertex_ptr(p.vertices, p.vertices_bs, b) + v);
        int4 r = make_int4(kProjSlow, 0, 0, 0);
        const bool fin = is_finite(c.x) && is_finite(c.y) && is_finite(c.z) && is_finite(c.w);
        // inside the depth range (same tests as classify_face) and in front of the eye
        if (fin && !(fadd(c.z, c.w) < 0.0f) && !(fsub(c.w, c.z) < 0.0f)) {
            int X, Y;
            float zw, rw;
            if (project_corner(c.x, c.y, c.z, c.w, halfW, halfH, X, Y, zw, rw))
                r = make_int4(X, Y, __float_as_int(zw), __float_as_int(rw));
        }
        if (r.x == kProjSlow) p.counters[kCtrSlow] = 1u;      // k_resolve_clipped has work (benign race: everyone writes 1)
        const_cast<int4 *>(p.proj)[(int64_t)b * p.V + v] = r;
        p.zvert[(int64_t)b * p.V + v] = r.x == kProjSlow ? 0.0f : __int_as_float(r.z);
    }
}

// ---------------------------------------------------------------------------------------
// S1: setup + line-parallel sampling
// ---------------------------------------------------------------------------------------
// Shared-memory record of a triangle being sampled by the warp: five 16-byte rows, read with
// LDS.128 (20-word stride: the 8 lanes of a quarter-warp phase hit disjoint bank groups).
// The three edges are stored in SLOT order: slot 0 is an edge whose value grows along a line
// (it bounds the span from below), slot 1 one that falls (bounds it from above), slot 2 the
// remaining edge of either kind -- the steps of a triangle's edges along any direction sum to
// zero, so a non-degenerate triangle always has one of each.
//   row 0: A0 A1 A2 base_lo  A_k = sign-normalised edge value at line 0, position 0, plus the tie
//                            bias (inside: >= 0); base = key index of pixel (xmin, ymin)
//   row 1: sa0 sa1 sa2 dims  sa_k = edge step per +1 position ALONG a line; dims = m1 | axis << 8 |
//                            first_line << 9 (m1 = last position of a line; axis 1: lines are columns)
//   row 2: sc0 sc1 sc2 face  sc_k = edge step per +1 line
//   row 3: inv0 inv1 inv2 base_hi   inv_k = 256 / |sa_k| as float (0 when sa_k == 0)
//   row 4: zref dzdx dzdy xy  xy = xmin | ymin << 16 (both below 2^15)
constexpr int kRecWords = 20;

__device__ __forceinline__ uint32_t pick3(const uint32_t (&v)[3], int i) { return i == 0 ? v[0] : (i == 1 ? v[1] : v[2]); }

// Fills the record of a triangle whose extents are below kSmallExtent (int32 edge math) and
// returns its dims word; the caller adds first_line once the warp's line counts are summed.
__device__ __forceinline__ uint32_t write_record(uint32_t *rec, const TriSetup &t, uint32_t face, uint32_t elem, int H, int W)
{
    const int cx = t.xmin * kSubpix + kSubpixHalf, cy = t.ymin * kSubpix + kSubpixHalf;
    const int sg = t.sign;
    const int ax[3] = {t.X1, t.X2, t.X0}, ay[3] = {t.Y1, t.Y2, t.Y0};
    const int bx[3] = {t.X2, t.X0, t.X1}, by[3] = {t.Y2, t.Y0, t.Y1};
    const int bias[3] = {t.bias0, t.bias1, t.bias2};
    const int bw = t.xmax - t.xmin + 1, bh = t.ymax - t.ymin + 1;
    const int axis = bw > bh ? 1 : 0;
    uint32_t A[3], sa[3], sc[3], inv[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const int ex = bx[k] - ax[k], ey = by[k] - ay[k];
        const int e = (ex * (cy - ay[k]) - ey * (cx - ax[k])) * sg;      // orient2d in int32
        const int sx = -ey * sg * kSubpix, sy = ex * sg * kSubpix;
        const int along = axis ? sy : sx;
        A[k] = (uint32_t)(e + bias[k]);
        sa[k] = (uint32_t)along;
        sc[k] = (uint32_t)(axis ? sx : sy);
        const int mag = (along < 0 ? -along : along) >> 8;           // |ey| or |ex|
        inv[k] = __float_as_uint(mag != 0 ? __frcp_rn((float)mag) : 0.0f);
    }
    const int up = (int)sa[0] > 0 ? 0 : ((int)sa[1] > 0 ? 1 : 2);        // slot 0: a rising edge
    const int dn = (int)sa[0] < 0 ? 0 : ((int)sa[1] < 0 ? 1 : 2);        // slot 1: a falling edge
    const int re = 3 - up - dn;
    const unsigned long long base = ((unsigned long long)elem * H + (unsigned)(H - 1 - t.ymin)) * W + (unsigned)t.xmin;
    uint4 *row = reinterpret_cast<uint4 *>(rec);
    row[0] = make_uint4(pick3(A, up), pick3(A, dn), pick3(A, re), (uint32_t)base);
    const uint32_t dims = (uint32_t)((axis ? bh : bw) - 1) | ((uint32_t)axis << 8);
    row[1] = make_uint4(pick3(sa, up), pick3(sa, dn), pick3(sa, re), dims);
    row[2] = make_uint4(pick3(sc, up), pick3(sc, dn), pick3(sc, re), face);
    row[3] = make_uint4(pick3(inv, up), pick3(inv, dn), pick3(inv, re), (uint32_t)(base >> 32));
    row[4] = make_uint4(__float_as_uint(t.zref), __float_as_uint(t.dzdx), __float_as_uint(t.dzdy), (uint32_t)t.xmin | (uint32_t)t.ymin << 16);
    return dims;
}

// 64-bit minimum without a return value, explicitly in the global space.
__device__ __forceinline__ void red_min_key(unsigned long long *kp, unsigned long long key)
{
    asm volatile("red.global.min.u64 [%0], %1;" :: "l"(kp), "l"(key) : "memory");
}

// The depth half (high word) of a key.  Cached in L1: staleness is harmless (see sample_line).
__device__ __forceinline__ uint32_t load_depth_word(const unsigned long long *kp)
{
    uint32_t v;
#if KLDCG
    asm volatile("ld.global.cg.u32 %0, [%1];" : "=r"(v) : "l"(reinterpret_cast<const uint32_t *>(kp) + 1));
#else
    asm volatile("ld.global.ca.u32 %0, [%1];" : "=r"(v) : "l"(reinterpret_cast<const uint32_t *>(kp) + 1));
#endif
    return v;
}

// Samples line `j - (first >> 9)` of the triangle in `rec` (a row when axis == 0, a column when
// axis == 1); FIRST: j counts the lines of the warp's batch and the record's dims word carries the
// number of the triangle's first line, otherwise j is the line of the triangle itself.
template <bool PRELOAD, bool FIRST>
__device__ __forceinline__ void sample_line(const uint32_t *rec, int j, const FwdParams &p)
{
    const uint4 *row = reinterpret_cast<const uint4 *>(rec);
    const uint4 rA = row[0], rSA = row[1], rSC = row[2], rINV = row[3];
    const uint32_t face = rSC.w;
    const int axis = (int)(rSA.w >> 8) & 1, m1 = (int)(rSA.w & 0xFFu);      // m1: last position along the line
    if (FIRST) j -= (int)(rSA.w >> 9);
    const float fj = __int2float_rn(j);
    // Edge along the line: a + sa * c >= 0 with sa = +-256 * s.  The bound on c is
    // ceil(na / s) with na = ceil(-a / 256) (nested ceilings), and
    // ceil((float(na) - 0.5) * (1 / s)) evaluates it EXACTLY in floating point: |na| < 2^22
    // (kSmallExtent) keeps the rounding error below the 0.5 / s that separates the shifted
    // quotient from the nearest integers (checked exhaustively-at-random in tests/test_span_math.py),
    // so coverage is still decided by integer arithmetic alone.
    auto bound = [&](uint32_t A, uint32_t SC, uint32_t INV, int &a) {
        a = (int)A + (int)SC * j;                    // edge value (+ bias) at position 0 of this line
        const int na = (255 - a) >> 8;
        return __float2int_ru(fmul(fadd(__int2float_rn(na), -0.5f), __uint_as_float(INV)));
    };
    int a0, a1, a2;
    int lo = max(0, bound(rA.x, rSC.x, rINV.x, a0));         // slot 0 rises: smallest c inside
    int hi = min(m1, -bound(rA.y, rSC.y, rINV.y, a1));       // slot 1 falls: largest c inside
    {
        const int sa = (int)rSA.z, m = bound(rA.z, rSC.z, rINV.z, a2);
        if (sa > 0) lo = max(lo, m);
        else if (sa < 0) hi = min(hi, -m);
        else if (a2 < 0) hi = -1;
    }
    if (lo > hi) return;
    const uint4 rZ = row[4];
    const float zref = __uint_as_float(rZ.x), dzdx = __uint_as_float(rZ.y), dzdy = __uint_as_float(rZ.z);
    // zw = (zref + dzdx * dx) + dzdy * dy, the operation order of plane_depth(), written for both
    // axes as (zp + zq * fi) + zr with fi the position along the line: rows have dy = j fixed,
    // columns dx = j (adding zr = 0 is exact).
    const float zp = axis ? fadd(zref, fmul(dzdx, fj)) : zref;
    const float zq = axis ? dzdy : dzdx;
    const float zr = axis ? 0.0f : fmul(dzdy, fj);
    // pixel (xmin + dx, ymin + dy) has key index base + dx - dy * W (tensor rows run downwards)
    const long long base = (long long)(((unsigned long long)rINV.w << 32) | rA.w);
    const int off = axis ? j - lo * p.W : lo - j * p.W;
    unsigned long long *kp = p.keys + (base + off);
    const long long stride = axis ? -(long long)p.W : 1ll;
    float fi = __int2float_rn(lo);
    // kFragRound fragments per round: their current depths are fetched together (independent
    // loads), so a line pays the memory latency once per round instead of once per fragment.
    constexpr int kFragRound = PRELOAD ? kFragRoundHiz : kFragRoundBlind;
    for (int i = lo; i <= hi; i += kFragRound) {
        const int n = hi - i + 1;
        uint32_t cur[kFragRound];
#pragma unroll
        for (int u = 0; u < kFragRound; ++u)
            cur[u] = !PRELOAD ? 0xFFFFFFFFu : (u < n ? load_depth_word(kp + u * stride) : 0u);
#pragma unroll
        for (int u = 0; u < kFragRound; ++u) {
            if (u < n) {
                float zw = fadd(fadd(zp, fmul(zq, fi)), zr);
                zw = fminf(fmaxf(zw, 0.0f), 1.0f);
                const uint32_t d = __float2uint_rz(fadd(fmul(zw, 16777215.0f), 0.5f));
                // keys only ever decrease, so a stale depth can only cost a redundant atomic;
                // on equal depth the atomic decides (lowest face index wins)
                if (d < kDepthMax && d <= cur[u])
                    red_min_key(kp + u * stride, ((unsigned long long)d << 32) | face);
                fi += 1.0f;
            }
        }
        kp += kFragRound * stride;
    }
}

// Rare path: a corner of the face is outside the depth range or behind the eye.  Clips the
// face and samples the fan triangles serially / via the large-triangle list.
static __device__ __noinline__ void setup_clipped(const FwdParams &p, int b, int f, int i0, int i1, int i2)
{
    const float4 *vp = vertex_ptr(p.vertices, p.vertices_bs, b);
    const float4 va = __ldg(vp + i0), vb = __ldg(vp + i1), vc = __ldg(vp + i2);
    if (classify_face(va, vb, vc) != 2) return;
    ClipVert poly[6];
    const int n = clip_face(va, vb, vc, poly);
    TriSetup t;
    for (int s = 1; s + 1 < n; ++s) {
        if (!setup_fan(poly, s, p.W, p.H, t)) continue;
        if ((t.xmax - t.xmin + 1) * (t.ymax - t.ymin + 1) <= 64)
            raster_serial(t, (uint32_t)f, p.keys + (size_t)b * p.H * p.W, p.W, p.H, t.xmin, t.xmax, t.ymin, t.ymax);
        else
            push_big(t, b, (uint32_t)f, (uint32_t)s, p);
    }
}

// True when no fragment of t can win a pixel: its smallest possible depth is beyond the
// hierarchical depth buffer in every tile the bounding box touches.  The plane depth is monotone
// in dx and in dy under round-to-nearest, so its minimum over the box is taken at a corner.
__device__ __forceinline__ uint32_t hiz_dmin(const TriSetup &t)
{
    const float dx = t.dzdx >= 0.0f ? 0.0f : __int2float_rn(t.xmax - t.xmin);
    const float dy = t.dzdy >= 0.0f ? 0.0f : __int2float_rn(t.ymax - t.ymin);
    float zw = fadd(fadd(t.zref, fmul(t.dzdx, dx)), fmul(t.dzdy, dy));
    zw = fminf(fmaxf(zw, 0.0f), 1.0f);
    return __float2uint_rz(fadd(fmul(zw, 16777215.0f), 0.5f));
}

__device__ __forceinline__ bool hidden_by_hiz(const FwdParams &p, int b, const TriSetup &t)
{
    const uint32_t dmin = hiz_dmin(t);
    // the level at which the box spans at most 3 x 3 tiles: nine independent loads
    const int ext = max(t.xmax - t.xmin, t.ymax - t.ymin);
    int l = 0;
    while (l < kHizLevels && (ext >> (kHizLog2 + l)) > 1) ++l;
    if (l == kHizLevels) return false;
    const int sh = kHizLog2 + l;
    const int ntx = (p.W + (1 << sh) - 1) >> sh, nty = (p.H + (1 << sh) - 1) >> sh;
    const int tx0 = t.xmin >> sh, ty0 = t.ymin >> sh, nx = (t.xmax >> sh) - tx0, ny = (t.ymax >> sh) - ty0;
    const uint32_t *hz = p.hiz + p.hiz_off[l] + ((size_t)b * nty + ty0) * ntx + tx0;
    uint32_t m = 0;
#pragma unroll
    for (int dy = 0; dy < 3; ++dy)
#pragma unroll
        for (int dx = 0; dx < 3; ++dx)
            if (dx <= nx && dy <= ny) m = max(m, __ldg(hz + dy * ntx + dx));
    return m < dmin;          // equal depth: the face index decides
}

// Second-level occlusion test (launches that have a depth pyramid).  Can any sample of lines ja .. jb
// of the triangle in `rec` still win a pixel?  The lines are eight (or fewer) neighbouring rows or
// columns that share one row / column of 8-px pyramid tiles.  Their samples lie inside the rectangle
// [lo, hi] x [ja, jb]: lo / hi are bounds on the edge crossings over the chunk -- an edge's crossing is
// linear in the line number, so it is bounded by its values at the two end lines -- evaluated in
// float with a margin (conservative: never narrower than the exact spans of sample_line).  Per tile
// the rectangle touches, the smallest depth the plane takes on the rectangle's part inside the tile
// (at a corner: the plane evaluation is monotone in dx and in dy under round-to-nearest, and it is
// the evaluation sample_line uses) is compared with the tile's maximum: strictly greater everywhere
// means no fragment of these lines can pass the depth test (equal depth: the face index decides).
__device__ __forceinline__ bool chunk_visible(const uint32_t *rec, int ja, int jb, const uint32_t *hz0, int ntx)
{
    const uint4 *row = reinterpret_cast<const uint4 *>(rec);
    const uint4 rA = row[0], rSA = row[1], rSC = row[2], rINV = row[3], rZ = row[4];
    const int axis = (int)(rSA.w >> 8) & 1, m1 = (int)(rSA.w & 0xFFu);
    constexpr float k256th = 1.0f / 256.0f;
    // slot 0 rises along the line (inside: c >= -a / sa), slot 1 falls (inside: c <= a / |sa|)
    float flo, fhi;
    {
        const int a0 = (int)rA.x + (int)rSC.x * ja, a1 = (int)rA.x + (int)rSC.x * jb;
        flo = -__int2float_rn(max(a0, a1)) * (__uint_as_float(rINV.x) * k256th);
    }
    {
        const int a0 = (int)rA.y + (int)rSC.y * ja, a1 = (int)rA.y + (int)rSC.y * jb;
        fhi = __int2float_rn(max(a0, a1)) * (__uint_as_float(rINV.y) * k256th);
    }
    {
        const int a0 = (int)rA.z + (int)rSC.z * ja, a1 = (int)rA.z + (int)rSC.z * jb;
        const float q = __int2float_rn(max(a0, a1)) * (__uint_as_float(rINV.z) * k256th);
        const int sa = (int)rSA.z;
        if (sa > 0) flo = fmaxf(flo, -q);
        else if (sa < 0) fhi = fminf(fhi, q);
        else if (max(a0, a1) < 0) return false;          // outside this edge on every line of the chunk
    }
    // |error| of the float crossings is below 2^-12 for |c| < 2^10; beyond that the clamp decides
    const int lo = max(0, __float2int_rd(flo - 0.01f)), hi = min(m1, __float2int_ru(fhi + 0.01f));
    if (lo > hi) return false;
    const float zref = __uint_as_float(rZ.x), dzdx = __uint_as_float(rZ.y), dzdy = __uint_as_float(rZ.z);
    const int xmin = (int)(rZ.w & 0xFFFFu), ymin = (int)(rZ.w >> 16);
    const int o = axis ? ymin : xmin;                                  // absolute coordinate of position 0
    const int tl = ((axis ? xmin : ymin) + ja) >> kHizLog2;            // the chunk's row / column of tiles
    const float zl = axis ? dzdx : dzdy, zq = axis ? dzdy : dzdx;
    const float fl = __int2float_rn(zl >= 0.0f ? ja : jb);
    const uint32_t *hz = hz0 + (axis ? tl : tl * ntx);
    const int step = axis ? ntx : 1;
    // smallest depth key of the rectangle's part inside tile t
    auto tile_dmin = [&](int t) {
        const int q0 = max(lo, (t << kHizLog2) - o), q1 = min(hi, (t << kHizLog2) + (1 << kHizLog2) - 1 - o);
        const float fc = __int2float_rn(zq >= 0.0f ? q0 : q1);
        // (zref + dzdx * dx) + dzdy * dy, the operation order of sample_line / plane_depth()
        float zw = axis ? fadd(fadd(zref, fmul(dzdx, fl)), fmul(dzdy, fc)) : fadd(fadd(zref, fmul(dzdx, fc)), fmul(dzdy, fl));
        zw = fminf(fmaxf(zw, 0.0f), 1.0f);
        return __float2uint_rz(fadd(fmul(zw, 16777215.0f), 0.5f));
    };
    int t = (o + lo) >> kHizLog2;
    const int t1 = (o + hi) >> kHizLog2;
#if KCHUNK_TWO
    {
        // most chunks touch one or two tiles: both words are in flight before either is looked at
        const int tb = min(t + 1, t1);
        const uint32_t ha = __ldg(hz + t * step), hb = __ldg(hz + tb * step);
        if (ha >= tile_dmin(t) || hb >= tile_dmin(tb)) return true;
        t += 2;
    }
#endif
    for (; t <= t1; ++t)
        if (__ldg(hz + t * step) >= tile_dmin(t)) return true;
    return false;
}

// ORDERED: the faces of depth buckets k0 .. k1 (depth_order.cu), nearest bucket first because work
// is claimed in the order of `order`.  HIZ (every ordered launch but the first): faces are tested
// against the hierarchical depth buffer, and those that survive are sampled in CHUNKS of eight lines
// that are tested again, each against the 8-px tiles it touches (chunk_visible).  Otherwise all
// faces in index order.
template <bool ORDERED, bool HIZ>
__global__ void __launch_bounds__(kWarpsPerCta * 32, ctas_per_sm(HIZ)) k_setup(FwdParams p, int k0, int k1, int launch_index, float min_overdraw)
{
    __shared__ __align__(16) uint32_t s_recs[kWarpsPerCta][32 * kRecWords];
    __shared__ uint8_t s_owner[kWarpsPerCta][32];
    __shared__ uint32_t s_aux[HIZ ? kWarpsPerCta : 1][32];     // HIZ: element | (first line & 7) << 16 | (lines - 1) << 19
    __shared__ uint16_t s_list[HIZ ? kWarpsPerCta : 1][32];    // HIZ: surviving chunks: owner lane | first line << 5 | (lines - 1) << 12
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t *recs = s_recs[warp];
    uint8_t *owner_lane = s_owner[warp];
    // ORDERED launches check the verdict of k_depth_pivots: without enough overdraw the launch of
    // bucket 0 samples every face in index order and the other buckets' launches do nothing.
    const bool ordered = ORDERED && plan_is_ordered(p, min_overdraw);
    if (HIZ && !ordered) return;
    long long first_entry = 0, total = (long long)p.B * p.F;
    // entries of the very nearest bucket: nearly every fragment wins there, reading the depth
    // first only adds latency; later buckets (of this or a later launch) look before they reduce
    long long no_preload_below = 0;
    if (ordered) {
        first_entry = p.bucket_base[k0];
        total = (long long)p.bucket_base[k1 + 1] - first_entry;
        if (k0 == 0 && KPRE0 == 0) no_preload_below = (long long)p.bucket_base[1];
    }
    // Batches of 32 faces are claimed dynamically (their line counts differ by orders of
    // magnitude); the claim for the next batch is issued before the current one is sampled.
    // Smaller batches shorten the tail of a launch but repeat the latency-bound setup phase more
    // often: 16 and 8 faces per claim measured slower (profiles/r1_setup_notes.md); several batches
    // per claim (fewer atomics on the one counter) lengthen the tail: slower too (profiles/r2_setup_notes.md).
    unsigned long long *work = p.work_next + launch_index;
    // a handful of faces cannot fill the machine line by line (one warp would walk every span in
    // turn): then anything but tiny triangles goes to the pixel-parallel tile path
    const bool many_faces = (long long)p.B * p.F >= kFewFaces;
    const int hz_ntx = (p.W + (1 << kHizLog2) - 1) >> kHizLog2, hz_nty = (p.H + (1 << kHizLog2) - 1) >> kHizLog2;
    unsigned long long claim = 0;
    if (lane == 0) claim = atomicAdd(work, 32ull);
    for (;;) {
        const long long i0 = (long long)__shfl_sync(kFullMask, claim, 0);
        if (i0 >= total) break;
        const long long i = i0 + lane;
        const long long i_end = min(total, i0 + 32ll);
        const bool preload = HIZ || i0 >= no_preload_below;
        if (lane == 0) claim = atomicAdd(work, 32ull);
        // ---- setup: one face per lane -------------------------------------------------------
        int nlines = 0;                  // > 0: this lane's face is line-sampled by the warp
        int nunits = 0;                  // what the warp distributes: lines, or (HIZ) chunks of eight lines
        uint32_t dims = 0;
        if (i < i_end) {
            // flat face numbers of an ordered plan are below 2^32 (capi.cu): one 32-bit division
            int b, f;
            if (ordered) {
                const uint32_t flat = p.order[first_entry + i];
                b = (int)(flat / (uint32_t)p.F); f = (int)(flat - (uint32_t)b * (uint32_t)p.F);
            } else {
                b = (int)(i / p.F); f = (int)(i - (long long)b * p.F);
            }
            const int32_t *fp = p.faces + (int64_t)b * p.faces_bs + 3 * (int64_t)f;
            const int i0 = __ldg(fp), i1 = __ldg(fp + 1), i2 = __ldg(fp + 2);
            if ((unsigned)i0 < (unsigned)p.V && (unsigned)i1 < (unsigned)p.V && (unsigned)i2 < (unsigned)p.V) {
                const int4 *pp = proj_ptr(p, b);
                TriSetup t;
                bool slow;
                if (setup_from_proj(__ldg(pp + i0), __ldg(pp + i1), __ldg(pp + i2), p.W, p.H, t, slow)) {
                    const int ex = max(t.X0, max(t.X1, t.X2)) - min(t.X0, min(t.X1, t.X2));
                    const int ey = max(t.Y0, max(t.Y1, t.Y2)) - min(t.Y0, min(t.Y1, t.Y2));
                    if (HIZ && hidden_by_hiz(p, b, t)) {
                        // dropped: cannot win any pixel
#ifdef DIRT_STATS
                        atomicAdd(p.counters + 4, 1u);
                        atomicAdd(p.counters + 5, (uint32_t)max(t.xmax - t.xmin + 1, t.ymax - t.ymin + 1));
#endif
                    } else if (ex < kSmallExtent && ey < kSmallExtent && ((many_faces && t.rcp_area > kLineRcpAreaMin) || (t.xmax - t.xmin + 1) * (t.ymax - t.ymin + 1) <= 256))
                    {   // few-but-fat triangles are better off pixel-parallel in the tile path
                        nlines = max(t.xmax - t.xmin, t.ymax - t.ymin) + 1;     // the longer side of the bounding box
                        dims = write_record(recs + lane * kRecWords, t, (uint32_t)f, (uint32_t)b, p.H, p.W);
                        nunits = nlines;
                        if (HIZ) {
                            // chunks are aligned to the pyramid's tiles along the axis the lines are counted on
                            const int l0 = ((dims >> 8) & 1u) ? t.xmin : t.ymin;
                            nunits = ((l0 + nlines - 1) >> kHizLog2) - (l0 >> kHizLog2) + 1;
                            s_aux[warp][lane] = (uint32_t)b | (uint32_t)(l0 & 7) << 16 | (uint32_t)(nlines - 1) << 19;
                        }
                    } else
                        push_big(t, b, (uint32_t)f, 0u, p);
                } else if (slow) {
                    // a private copy: taking the address of the kernel parameter block itself would move
                    // it to local memory for the whole kernel (and make every key atomic a generic one)
                    const FwdParams pc = p;
                    setup_clipped(pc, b, f, i0, i1, i2);
                }
            }
        }
        // the r-th line-sampled face of the batch, in lane order, is the face of lane owner_lane[r]
        const bool sampled = nunits > 0;
        const unsigned owners = __ballot_sync(kFullMask, sampled);
        if (sampled) owner_lane[__popc(owners & ((1u << lane) - 1u))] = (uint8_t)lane;
#ifdef DIRT_STATS
        if (nlines) { atomicAdd(p.counters + 6, 1u); atomicAdd(p.counters + 7, (uint32_t)nlines); }
#endif
        // running sum of the unit counts over the warp: record r owns units excl .. excl + nunits - 1
        int incl = nunits;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int n = __shfl_up_sync(kFullMask, incl, o);
            if (lane >= o) incl += n;
        }
        const int nall = __shfl_sync(kFullMask, incl, 31);
        const int excl = incl - nunits;
        if (sampled) recs[lane * kRecWords + 7] = dims | ((uint32_t)excl << 9);
        __syncwarp();
        // ---- sampling: 32 units at a time ---------------------------------------------------
        // Unit g belongs to the last record that starts at or before g.  Records start at
        // distinct units (every record has at least one), so one OR-reduction of "my record
        // starts at unit base + d" bits and two population counts find the owner of every lane.
        const unsigned upto_lane = (2u << lane) - 1u;
        for (int base = 0; base < nall; base += 32) {
            const int d = excl - base;
            const unsigned starts = __reduce_or_sync(kFullMask, (sampled && (unsigned)d < 32u) ? 1u << d : 0u);
            const int before = __popc(__ballot_sync(kFullMask, sampled && d < 0));
            const int g = base + lane;
            if (!HIZ) {
                if (g < nall) {
                    const uint32_t *rec = recs + owner_lane[before + __popc(starts & upto_lane) - 1] * kRecWords;
                    if (preload) sample_line<true, true>(rec, g, p);
                    else sample_line<false, true>(rec, g, p);
                }
            } else {
                // the lane's chunk: lines ja .. jb of its triangle; tested, and the survivors listed
                bool visible = false;
                uint32_t entry = 0;
                if (g < nall) {
                    const int ol = owner_lane[before + __popc(starts & upto_lane) - 1];
                    const uint32_t *rec = recs + ol * kRecWords;
                    const uint32_t aux = s_aux[warp][ol];
                    const int ci = g - (int)(rec[7] >> 9);
                    const int first = (ci << kHizLog2) - (int)((aux >> 16) & 7u);
                    const int ja = max(first, 0), jb = min(first + 7, (int)(aux >> 19));
                    const uint32_t *hz0 = p.hiz + p.hiz_off[0] + (size_t)(aux & 0xFFFFu) * hz_nty * hz_ntx;
                    visible = (KCHUNK_MIN > 0 && (int)(aux >> 19) < KCHUNK_MIN) || chunk_visible(rec, ja, jb, hz0, hz_ntx);
                    entry = (uint32_t)ol | (uint32_t)ja << 5 | (uint32_t)(jb - ja) << 12;
#ifdef DIRT_STATS
                    atomicAdd(p.counters + 9, 1u);
                    if (visible) atomicAdd(p.counters + 10, (uint32_t)(jb - ja + 1));
#endif
                }
                const unsigned alive = __ballot_sync(kFullMask, visible);
                if (visible) s_list[warp][__popc(alive & ((1u << lane) - 1u))] = (uint16_t)entry;
                __syncwarp();
                const int nl = __popc(alive) << 3;
                for (int e = lane; e < nl; e += 32) {
                    const uint32_t en = s_list[warp][e >> 3];
                    const int dj = e & 7;
                    if (dj <= (int)(en >> 12)) sample_line<true, false>(recs + (en & 31u) * kRecWords, (int)((en >> 5) & 127u) + dj, p);
                }
                __syncwarp();     // the list is about to be overwritten
            }
        }
        __syncwarp();     // the records are about to be overwritten
    }
}

void launch_project(const FwdParams &p, int sm_count, cudaStream_t stream, StageTimer &tm)
{
    const int nbatch = p.vertices_bs == 0 ? 1 : p.B;
    const long long nverts = (long long)nbatch * p.V;
    tm.begin(kStageProject, stream);
    {
        long long gv = (nverts + 255) / 256;
        if (gv > (long long)sm_count * 16) gv = (long long)sm_count * 16;
        if (gv < 1) gv = 1;
        k_project<<<(int)gv, 256, 0, stream>>>(p, nbatch);
    }
    tm.end(kStageProject, stream);
}

void launch_setup(const FwdParams &p, int sm_count, cudaStream_t stream, uint64_t *launches, StageTimer &tm)
{
    const long long nfaces = (long long)p.B * p.F;
    launch_project(p, sm_count, stream, tm);
    long long g = (nfaces + kWarpsPerCta * 32 - 1) / (kWarpsPerCta * 32);
    const long long gall = g < 1 ? 1 : g;
    auto grid_for = [&](bool hiz) {                                 // resident CTAs: work is claimed dynamically
        const long long cap = (long long)sm_count * ctas_per_sm(hiz);
        return (int)(gall > cap ? cap : gall);
    };
    const BucketPlan plan = bucket_plan();
    if (p.order && plan.n > 0) {
        // front-to-back buckets; the hierarchical depth buffer is refreshed between them
        tm.begin(kStageDepthOrder, stream);
        launch_depth_order(p, plan, sm_count, stream);
        tm.end(kStageDepthOrder, stream);
        for (int j = 0; j < plan.nlaunch; ++j) {
            if (j > 0) {
                tm.begin(kStageHizBuild, stream);
                launch_hiz_build(p, plan, sm_count, stream);
                tm.end(kStageHizBuild, stream);
            }
            tm.begin(kStageSetup, stream);
            if (j == 0) k_setup<true, false><<<grid_for(false), kWarpsPerCta * 32, 0, stream>>>(p, plan.launch_first[j], plan.launch_first[j + 1] - 1, j, plan.min_overdraw);
            else k_setup<true, true><<<grid_for(true), kWarpsPerCta * 32, 0, stream>>>(p, plan.launch_first[j], plan.launch_first[j + 1] - 1, j, plan.min_overdraw);
            tm.end(kStageSetup, stream);
#ifdef DIRT_STATS
            uint32_t c[12], bb[kMaxBuckets + 1];
            cudaStreamSynchronize(stream);
            cudaMemcpy(c, p.counters, sizeof(c), cudaMemcpyDeviceToHost);
            cudaMemcpy(bb, p.bucket_base, sizeof(bb), cudaMemcpyDeviceToHost);
            printf("launch %d: faces %u | cumulative: culled tris %u (lines %u), sampled tris %u (lines %u), chunks tested %u, lines in surviving chunks %u\n", j,
                   bb[plan.launch_first[j + 1]] - bb[plan.launch_first[j]], c[4], c[5], c[6], c[7], c[9], c[10]);
#endif
        }
        *launches += (plan.n == 2 ? 2 : 4) + 2 * plan.nlaunch - 1;     // ordering kernels + setup launches + pyramid builds
    } else {
        tm.begin(kStageSetup, stream);
        k_setup<false, false><<<grid_for(false), kWarpsPerCta * 32, 0, stream>>>(p, 0, 0, 0, 0.0f);
        tm.end(kStageSetup, stream);
        *launches += 1;
    }
}

}  // namespace dirt
