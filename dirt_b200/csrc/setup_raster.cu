// setup_raster.cu -- stage S1 of the forward pass: triangle setup plus sampling of every
// triangle that is not large enough to be worth binning.
//
// Thread t of a warp sets up face t of a batch of 32 faces.  Sampling is then organised as a
// producer / consumer pair inside the warp, because "one lane walks one triangle" leaves most
// lanes idle on long thin triangles (profiles/r1_setup_lane_per_triangle.md):
//   producers  each lane walks rows of a triangle and computes the row's EXACT covered span:
//              a float estimate of where each edge crosses the row, then an integer fix-up
//              against the exact edge values, so float arithmetic never decides coverage;
//              the span's pixels are pushed as 32-bit fragment references into a per-warp
//              shared-memory queue (warp prefix sum for the slots).
//   consumers  whenever 32 fragments are queued, every lane takes one: edge values from the
//              triangle's shared-memory record, window depth, depth|face key, 64-bit atomicMin
//              into the key buffer.  All 32 lanes are busy regardless of triangle shape.
// Triangles of at most kInlineRows rows are walked by their own lane (32 triangles in
// flight); taller ones, up to kSpanDim rows, by the whole warp (lane = row).
#include "forward_common.cuh"

namespace dirt {

constexpr unsigned kFullMask = 0xFFFFFFFFu;
constexpr int kWarpsPerCta = 8;
constexpr int kPushMax = 8;                       // fragments one lane may push per round
constexpr int kQueueCap = 32 + 32 * kPushMax;     // < 32 left over + one round of pushes
constexpr int kMaxExtent = 1 << 22;               // span route needs edge steps that fit in int32

// Shared-memory record of a triangle being sampled by the warp (23-word stride: odd, so lanes
// reading the same field of different records hit different banks).
enum {
    rE0 = 0,        // 3 x int64 (lo, hi words): sign-normalised edge values at pixel (xmin, ymin)
    rSX = 6,        // 3 x int32: edge step per +1 pixel in x
    rSY = 9,        // 3 x int32: edge step per +1 pixel in y
    rZW = 12,       // 3 x float: window depth of the corners
    rRCP = 15,      // float: 1 / |2 area|
    rXMIN = 16, rYMIN = 17,
    rFACE = 18, rELEM = 19,
    rBIAS = 20,     // bit k set: edge k does NOT own its samples (bias -1)
    rDIMS = 21,     // (xmax - xmin) | (ymax - ymin) << 16
    kRecWords = 23
};

struct Producer {
    long long A[3];       // edge value at (xmin, ymin) plus tie bias: row constraint is A + sy*dy + sx*d >= 0
    int sx[3], sy[3];
    float inv[3];         // 1 / sx (0 when sx == 0)
    int bw;               // xmax - xmin
    int dy, dy_end, dy_step;
    int slot;
    int span_dy, d_cur, d_hi;   // pending span (empty when d_cur > d_hi)
};

__device__ __forceinline__ bool span_route_ok(const TriSetup &t)
{
    int ex = max(t.X0, max(t.X1, t.X2)) - min(t.X0, min(t.X1, t.X2));
    int ey = max(t.Y0, max(t.Y1, t.Y2)) - min(t.Y0, min(t.Y1, t.Y2));
    return ex < kMaxExtent && ey < kMaxExtent && (t.xmax - t.xmin) < kSpanDim && (t.ymax - t.ymin) < kSpanDim;
}

__device__ __forceinline__ void write_record(uint32_t *rec, const TriSetup &t, uint32_t face, uint32_t elem)
{
    long long e[3], sx[3], sy[3];
    edge_values(t, t.xmin, t.ymin, e[0], e[1], e[2]);
    edge_steps(t, sx, sy);
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        rec[rE0 + 2 * k] = (uint32_t)(unsigned long long)e[k];
        rec[rE0 + 2 * k + 1] = (uint32_t)((unsigned long long)e[k] >> 32);
        rec[rSX + k] = (uint32_t)(int)sx[k];
        rec[rSY + k] = (uint32_t)(int)sy[k];
    }
    rec[rZW] = __float_as_uint(t.zw0); rec[rZW + 1] = __float_as_uint(t.zw1); rec[rZW + 2] = __float_as_uint(t.zw2);
    rec[rRCP] = __float_as_uint(t.rcp_area);
    rec[rXMIN] = (uint32_t)t.xmin; rec[rYMIN] = (uint32_t)t.ymin;
    rec[rFACE] = face; rec[rELEM] = elem;
    rec[rBIAS] = (t.bias0 ? 1u : 0u) | (t.bias1 ? 2u : 0u) | (t.bias2 ? 4u : 0u);
    rec[rDIMS] = (uint32_t)(t.xmax - t.xmin) | ((uint32_t)(t.ymax - t.ymin) << 16);
}

__device__ __forceinline__ long long rec_e0(const uint32_t *rec, int k)
{
    return (long long)((unsigned long long)rec[rE0 + 2 * k] | ((unsigned long long)rec[rE0 + 2 * k + 1] << 32));
}

// Makes this lane a producer for rows dy0, dy0 + step, ... of the triangle in `rec`.
__device__ __forceinline__ void start_producer(Producer &pr, const uint32_t *rec, int slot, int dy0, int step)
{
    const uint32_t bias = rec[rBIAS], dims = rec[rDIMS];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        pr.A[k] = rec_e0(rec, k) - (long long)((bias >> k) & 1u);
        pr.sx[k] = (int)rec[rSX + k];
        pr.sy[k] = (int)rec[rSY + k];
        pr.inv[k] = pr.sx[k] != 0 ? __frcp_rn((float)pr.sx[k]) : 0.0f;
    }
    pr.bw = (int)(dims & 0xFFFFu);
    pr.dy = dy0; pr.dy_end = (int)(dims >> 16); pr.dy_step = step;
    pr.slot = slot;
    pr.d_cur = 0; pr.d_hi = -1; pr.span_dy = 0;
}

// Exact covered span [lo, hi] (offsets from xmin) of row dy.  The float quotient only seeds
// the search; the while loops settle on the exact integers (they iterate 0 or 1 times when
// the estimate is within one pixel, which DESIGN.md "span estimate" shows it is).
__device__ __forceinline__ void row_span(const Producer &pr, int dy, int &lo, int &hi)
{
    lo = 0; hi = pr.bw;
    bool empty = false;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const long long a = pr.A[k] + (long long)pr.sy[k] * dy;
        const int sx = pr.sx[k];
        if (sx == 0) { empty |= a < 0; continue; }
        const float d = -__ll2float_rn(a) * pr.inv[k];
        if (sx > 0) {           // smallest c >= 0 with a + sx*c >= 0
            int c = (int)fminf(fmaxf(ceilf(d), 0.0f), (float)(pr.bw + 1));
            while (c <= pr.bw && a + (long long)sx * c < 0) ++c;
            while (c > 0 && a + (long long)sx * (c - 1) >= 0) --c;
            lo = max(lo, c);
        } else {                // largest c <= bw with a + sx*c >= 0
            int c = (int)fminf(fmaxf(floorf(d), -1.0f), (float)pr.bw);
            while (c >= 0 && a + (long long)sx * c < 0) --c;
            while (c < pr.bw && a + (long long)sx * (c + 1) >= 0) ++c;
            hi = min(hi, c);
        }
    }
    if (empty) hi = lo - 1;
}

// One queued fragment: slot | dx << 5 | dy << 13.
__device__ __forceinline__ void consume(uint32_t frag, const uint32_t *recs, const FwdParams &p)
{
    const uint32_t *rec = recs + (frag & 31u) * kRecWords;
    const int dx = (int)((frag >> 5) & 0xFFu), dy = (int)((frag >> 13) & 0xFFu);
    const long long e0 = rec_e0(rec, 0) + (long long)(int)rec[rSX + 0] * dx + (long long)(int)rec[rSY + 0] * dy;
    const long long e1 = rec_e0(rec, 1) + (long long)(int)rec[rSX + 1] * dx + (long long)(int)rec[rSY + 1] * dy;
    const long long e2 = rec_e0(rec, 2) + (long long)(int)rec[rSX + 2] * dx + (long long)(int)rec[rSY + 2] * dy;
    // same arithmetic as sample_depth() in raster_core.cuh
    const float rcp = __uint_as_float(rec[rRCP]);
    const float l0 = fmul(__ll2float_rn(e0), rcp), l1 = fmul(__ll2float_rn(e1), rcp), l2 = fmul(__ll2float_rn(e2), rcp);
    const float zw = fadd(fadd(fmul(l0, __uint_as_float(rec[rZW])), fmul(l1, __uint_as_float(rec[rZW + 1]))),
                          fmul(l2, __uint_as_float(rec[rZW + 2])));
    unsigned long long key;
    if (!depth_key(zw, rec[rFACE], key)) return;
    const int px = (int)rec[rXMIN] + dx, py = (int)rec[rYMIN] + dy;
    unsigned long long *dst = p.keys + ((size_t)rec[rELEM] * p.H + (size_t)(p.H - 1 - py)) * p.W + px;
    atomicMin(dst, key);
}

// Runs the producers of this warp until they are out of rows, consuming full batches of 32
// fragments as they become available.  qcount (< 32 on entry and exit) is warp-uniform.
__device__ __forceinline__ void pump(Producer &pr, uint32_t *queue, const uint32_t *recs, int &qcount,
                                     const FwdParams &p, int lane)
{
    for (;;) {
        if (pr.d_cur > pr.d_hi && pr.dy <= pr.dy_end) {
            row_span(pr, pr.dy, pr.d_cur, pr.d_hi);
            pr.span_dy = pr.dy;
            pr.dy += pr.dy_step;
        }
        const int L = pr.d_cur <= pr.d_hi ? min(pr.d_hi - pr.d_cur + 1, kPushMax) : 0;
        int incl = L;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int n = __shfl_up_sync(kFullMask, incl, o);
            if (lane >= o) incl += n;
        }
        const int total = __shfl_sync(kFullMask, incl, 31);
        const int base = qcount + incl - L;
        const uint32_t tag = (uint32_t)pr.slot | ((uint32_t)pr.span_dy << 13);
        for (int j = 0; j < L; ++j) queue[base + j] = tag | ((uint32_t)(pr.d_cur + j) << 5);
        pr.d_cur += L;
        qcount += total;
        const bool more = __any_sync(kFullMask, pr.d_cur <= pr.d_hi || pr.dy <= pr.dy_end);
        __syncwarp();
        while (qcount >= 32) {
            qcount -= 32;
            consume(queue[qcount + lane], recs, p);
        }
        __syncwarp();
        if (!more) break;
    }
}

__global__ void __launch_bounds__(kWarpsPerCta * 32) k_setup(FwdParams p)
{
    __shared__ uint32_t s_recs[kWarpsPerCta][32 * kRecWords];
    __shared__ uint32_t s_queue[kWarpsPerCta][kQueueCap];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t *recs = s_recs[warp], *queue = s_queue[warp];
    const long long total = (long long)p.B * p.F;
    const long long total_padded = (total + 31) & ~31ll;       // warps stay together
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total_padded;
         i += (long long)gridDim.x * blockDim.x) {
        // ---- setup: one face per lane -------------------------------------------------------
        bool spans = false;
        int rows = 0;
        if (i < total) {
            const int b = (int)(i / p.F), f = (int)(i - (long long)b * p.F);
            float4 va, vb, vc;
            if (load_face(p.faces, p.faces_bs, p.vertices, p.vertices_bs, b, f, p.V, va, vb, vc)) {
                TriSetup t;
                const int cls = classify_face(va, vb, vc);
                if (cls == 1) {
                    if (setup_face(va, vb, vc, p.W, p.H, t)) {
                        if (span_route_ok(t)) {
                            write_record(recs + lane * kRecWords, t, (uint32_t)f, (uint32_t)b);
                            spans = true;
                            rows = t.ymax - t.ymin + 1;
                        } else {
                            push_big(t, b, (uint32_t)f, 0u, p);
                        }
                    }
                } else if (cls == 2) {
                    // rare: the face crosses the near / far plane; fan triangles go the slow ways
                    ClipVert poly[6];
                    const int n = clip_face(va, vb, vc, poly);
                    for (int s = 1; s + 1 < n; ++s) {
                        if (!setup_fan(poly, s, p.W, p.H, t)) continue;
                        if ((t.xmax - t.xmin + 1) * (t.ymax - t.ymin + 1) <= 64)
                            raster_serial(t, (uint32_t)f, p.keys + (size_t)b * p.H * p.W, p.W, p.H, t.xmin, t.xmax, t.ymin, t.ymax);
                        else
                            push_big(t, b, (uint32_t)f, (uint32_t)s, p);
                    }
                }
            }
        }
        __syncwarp();
        // ---- sampling ------------------------------------------------------------------------
        int qcount = 0;
        Producer pr;
        const bool own = spans && rows <= kInlineRows;
        if (own) {
            start_producer(pr, recs + lane * kRecWords, lane, 0, 1);
        } else {
            pr.dy = 1; pr.dy_end = 0; pr.dy_step = 1; pr.d_cur = 0; pr.d_hi = -1; pr.slot = 0; pr.span_dy = 0; pr.bw = 0;
        }
        if (__any_sync(kFullMask, own)) pump(pr, queue, recs, qcount, p, lane);
        unsigned tall = __ballot_sync(kFullMask, spans && !own);
        while (tall) {
            const int src = __ffs(tall) - 1;
            tall &= tall - 1;
            start_producer(pr, recs + src * kRecWords, src, lane, 32);
            pump(pr, queue, recs, qcount, p, lane);
        }
        // the records are about to be overwritten: drain what is left (< 32 fragments)
        if (lane < qcount) consume(queue[lane], recs, p);
        __syncwarp();
    }
}

void launch_setup(const FwdParams &p, int sm_count, cudaStream_t stream)
{
    const long long nfaces = (long long)p.B * p.F;
    long long g = (nfaces + kWarpsPerCta * 32 - 1) / (kWarpsPerCta * 32);
    const long long cap = (long long)sm_count * 32;
    if (g > cap) g = cap;
    if (g < 1) g = 1;
    k_setup<<<(int)g, kWarpsPerCta * 32, 0, stream>>>(p);
}

}  // namespace dirt
