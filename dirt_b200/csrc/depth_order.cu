// depth_order.cu -- front-to-back passes with a hierarchical depth buffer, for meshes with depth
// complexity (the headline mesh draws ~13 fragments per pixel of which one survives).
//
// The result of the forward pass is min over fragments of (depth24 << 32 | face), so the ORDER in
// which faces are sampled cannot change it.  That freedom is used here: the faces of every
// element are split into a few depth buckets (nearest first), the buckets are sampled one launch
// after the other, and between launches the key buffer is reduced to the maximum depth of each
// 8x8 pixel tile.  A triangle of a later bucket whose smallest possible fragment depth is still
// greater than that maximum in every tile its bounding box touches cannot win a pixel and is
// dropped before any of its lines is generated (setup_raster.cu).  The bound is exact for the
// arithmetic actually used: the plane depth of RASTER_SPEC section 6 is monotone in dx and in dy
// under round-to-nearest, so its minimum over the bounding box is its value at one corner.
//
//   k_depth_pivots    256 evenly spaced sample faces per element, sorted by nearest corner depth
//                     -> the bucket boundaries of the element (sample quantiles)
//   k_depth_split2    two buckets (the default): classify + count + scatter in one pass, bucket 0 filled
//                     from the front of `order`, bucket 1 from its back
//   k_depth_classify  (three buckets and more) bucket of every face; faces per (bucket, element)
//   k_depth_offsets   exclusive scan, bucket-major: bucket k of all elements is one contiguous
//                     range of `order`, so one launch samples bucket k of the whole batch
//   k_depth_scatter   flat face numbers into their ranges (order inside a range is arbitrary)
//   k_hiz_build       maxima of the depth words over 8, 16, 32 and 64 px tiles
#include <cstdlib>
#include "forward_common.cuh"

namespace dirt {

namespace {

constexpr unsigned kFull = 0xFFFFFFFFu;
constexpr int kSamples = 256;

// Nearest corner depth of a face; 0 (first bucket, never culled) for faces on the general path.
// `frags` (optional) receives an estimate of the fragments the face generates.
__device__ __forceinline__ float face_near_depth(const FwdParams &p, int b, int f, float *frags = nullptr, int *facing = nullptr)
{
    if (frags) *frags = 0.0f;
    if (facing) *facing = 0;
    const int32_t *fp = p.faces + (int64_t)b * p.faces_bs + 3 * (int64_t)f;
    const int i0 = __ldg(fp), i1 = __ldg(fp + 1), i2 = __ldg(fp + 2);
    if ((unsigned)i0 >= (unsigned)p.V || (unsigned)i1 >= (unsigned)p.V || (unsigned)i2 >= (unsigned)p.V) return 0.0f;
    if (!frags && !facing) {
        // only the depths: three 4-byte gathers from the packed copy (neighbouring faces share vertices, and 4-byte
        // values of neighbouring vertices share cache lines four times as often as the 16-byte records do)
        const float *zv = p.zvert + (int64_t)b * p.proj_bs;
        return fminf(__ldg(zv + i0), fminf(__ldg(zv + i1), __ldg(zv + i2)));
    }
    const int4 *pp = proj_ptr(p, b);
    const int4 a = __ldg(pp + i0), c = __ldg(pp + i1), d = __ldg(pp + i2);
    if (a.x == kProjSlow || c.x == kProjSlow || d.x == kProjSlow) return 0.0f;
    if (facing) {
        // winding in the window: sign of twice the area of the snapped corners
        const long long a2 = (long long)(c.x - a.x) * (d.y - a.y) - (long long)(c.y - a.y) * (d.x - a.x);
        *facing = a2 > 0 ? 1 : (a2 < 0 ? -1 : 0);
    }
    if (frags) {
        TriSetup t;
        bool slow;
        if (setup_from_proj(a, c, d, p.W, p.H, t, slow)) {
            // area in pixels, but no more than the part of the bounding box inside the image
            const float area = 0.5f / (t.rcp_area * (float)(kSubpix * kSubpix));
            *frags = fminf(area, (float)((t.xmax - t.xmin + 1) * (t.ymax - t.ymin + 1)));
        }
    }
    return fminf(__int_as_float(a.z), fminf(__int_as_float(c.z), __int_as_float(d.z)));
}

// ascending bitonic sort of s_z[kSamples] by the kSamples threads of the CTA
__device__ __forceinline__ void sort_samples(float *s_z, int t)
{
    for (int k = 2; k <= kSamples; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1) {
            const int q = t ^ j;
            if (q > t) {
                const float x = s_z[t], y = s_z[q];
                const bool up = (t & k) == 0;
                if ((x > y) == up) { s_z[t] = y; s_z[q] = x; }
            }
            __syncthreads();
        }
}

__global__ void __launch_bounds__(kSamples) k_depth_pivots(FwdParams p, BucketPlan plan)
{
    __shared__ float s_z[kSamples];
    __shared__ float s_frags;
    __shared__ int s_vote, s_front;
    const int b = blockIdx.x, t = threadIdx.x;
    const int f = (int)(((long long)t * p.F) / kSamples);
    if (t == 0) { s_frags = 0.0f; s_vote = 0; s_front = 0; }
    __syncthreads();
    float frags;
    int facing;
    const float z = face_near_depth(p, b, f, &frags, &facing);
    s_z[t] = z;
    for (int o = 16; o >= 1; o >>= 1) frags += __shfl_xor_sync(kFull, frags, o);
    if ((t & 31) == 0) atomicAdd(&s_frags, frags);
    __syncthreads();
    // fragments per pixel of this element, extrapolated from the sample; summed over the batch
    if (t == 0)
        atomicAdd(reinterpret_cast<float *>(p.counters + kCtrOverdraw), s_frags * ((float)p.F / kSamples) / ((float)p.H * (float)p.W));
    sort_samples(s_z, t);
    int nfront = kSamples;
    if (plan.facing) {
        // Which winding faces the viewer?  The one most of the NEARER half of the sample has: on a
        // closed surface the far half hides behind the near half, and the near half's other winding
        // is the inside of bumps.  Faces of the other winding go to the last bucket, behind the
        // pyramid built from everything before it (a guess only orders the work; results are the same).
        const float median = s_z[kSamples / 2];
        __syncthreads();
        if (z < median && facing) atomicAdd(&s_vote, facing);
        __syncthreads();
        const int front = s_vote >= 0 ? 1 : -1;
        const bool is_front = facing == front || z == 0.0f;      // clipped / invalid faces stay in the first bucket
        s_z[t] = is_front ? z : 3.0f;                            // window depths are <= 1
        if (is_front) atomicAdd(&s_front, 1);
        __syncthreads();
        nfront = s_front;
        sort_samples(s_z, t);
        if (t == 0) p.pivots[b * kMaxBuckets + kMaxBuckets - 1] = (float)front;
    }
    // the union mode of the visibility hint sends the nearest tenth of the faces along with the hinted ones
    if (t == 0 && plan.n == 2) p.pivots[b * kMaxBuckets + kMaxBuckets - 2] = s_z[min(kSamples - 1, (int)(0.10f * nfront))];
    if (t < plan.n - 1) {
        int at = (int)(plan.upto[t] * nfront);
        at = at < 0 ? 0 : (at > kSamples - 1 ? kSamples - 1 : at);
        p.pivots[b * kMaxBuckets + t] = s_z[at];
    }
}

__device__ __forceinline__ int bucket_for(const float *pv, int nb, float z)
{
    int k = 0;
    for (int j = 0; j < nb - 1; ++j) k += z >= pv[j] ? 1 : 0;
    return k;
}

constexpr int kChunkFaces = 2048;          // faces of ONE element per CTA in classify / scatter

// blockIdx.y = element, blockIdx.x = chunk of kChunkFaces faces.  Counts go through shared
// memory: kMaxBuckets global reductions per CTA instead of one per warp and bucket.
__global__ void __launch_bounds__(256) k_depth_classify(FwdParams p, int nb, float min_overdraw, int use_facing)
{
    if (!plan_is_ordered(p, min_overdraw)) return;
    __shared__ uint32_t s_cnt[kMaxBuckets];
    __shared__ float s_pv[kMaxBuckets];
    const int b = blockIdx.y, t = threadIdx.x;
    if (t < kMaxBuckets) { s_cnt[t] = 0; s_pv[t] = t < nb - 1 ? p.pivots[b * kMaxBuckets + t] : 0.0f; }
    __syncthreads();
    const int f0 = blockIdx.x * kChunkFaces, f1 = min(f0 + kChunkFaces, p.F);
    const int front = use_facing ? (int)p.pivots[b * kMaxBuckets + kMaxBuckets - 1] : 0;
    for (int f = f0 + t; f < f1; f += 256) {
        int facing;
        const float z = face_near_depth(p, b, f, nullptr, &facing);
        const int k = (front == 0 || facing == front || z == 0.0f) ? bucket_for(s_pv, nb, z) : nb - 1;
        p.bucket_of[(size_t)b * p.F + f] = (uint8_t)k;
        atomicAdd(&s_cnt[k], 1u);
    }
    __syncthreads();
    if (t < nb && s_cnt[t]) atomicAdd(p.bucket_count + (size_t)t * p.B + b, s_cnt[t]);
}

// Exclusive scan of bucket_count (bucket-major) -> bucket_cursor; bucket_base[k] = start of bucket k.
__global__ void __launch_bounds__(1024) k_depth_offsets(FwdParams p, int nb, float min_overdraw)
{
    if (!plan_is_ordered(p, min_overdraw)) return;
    __shared__ uint32_t s_part[1024];
    const int n = nb * p.B, t = threadIdx.x;
    const int per = (n + 1023) / 1024;
    const int lo = min(t * per, n), hi = min(lo + per, n);
    uint32_t sum = 0;
    for (int i = lo; i < hi; ++i) sum += p.bucket_count[i];
    s_part[t] = sum;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {
        const uint32_t v = t >= o ? s_part[t - o] : 0u;
        __syncthreads();
        s_part[t] += v;
        __syncthreads();
    }
    uint32_t run = s_part[t] - sum;
    for (int i = lo; i < hi; ++i) {
        p.bucket_cursor[i] = run;
        if (i % p.B == 0) p.bucket_base[i / p.B] = run;
        run += p.bucket_count[i];
    }
    if (t == 1023) p.bucket_base[nb] = s_part[1023];
}

// Same chunks as k_depth_classify: a CTA reserves one range per bucket and fills it with its
// flat face numbers (ranks inside the CTA come from shared-memory counters).
__global__ void __launch_bounds__(256) k_depth_scatter(FwdParams p, int nb, float min_overdraw)
{
    if (!plan_is_ordered(p, min_overdraw)) return;
    __shared__ uint32_t s_cnt[kMaxBuckets], s_base[kMaxBuckets];
    const int b = blockIdx.y, t = threadIdx.x;
    if (t < kMaxBuckets) s_cnt[t] = 0;
    __syncthreads();
    const int f0 = blockIdx.x * kChunkFaces, f1 = min(f0 + kChunkFaces, p.F);
    uint32_t rank[kChunkFaces / 256];
    uint8_t kk[kChunkFaces / 256];
#pragma unroll
    for (int j = 0; j < kChunkFaces / 256; ++j) {
        const int f = f0 + t + j * 256;
        if (f < f1) {
            kk[j] = p.bucket_of[(size_t)b * p.F + f];
            rank[j] = atomicAdd(&s_cnt[kk[j]], 1u);
        }
    }
    __syncthreads();
    if (t < nb && s_cnt[t]) s_base[t] = atomicAdd(p.bucket_cursor + (size_t)t * p.B + b, s_cnt[t]);
    __syncthreads();
#pragma unroll
    for (int j = 0; j < kChunkFaces / 256; ++j) {
        const int f = f0 + t + j * 256;
        if (f < f1) p.order[s_base[kk[j]] + rank[j]] = (uint32_t)((size_t)b * p.F + f);
    }
}

// Two buckets (the default plan): classify, count and scatter in ONE pass.  Bucket 0 grows up from
// the start of `order`, bucket 1 down from its end -- every face lands in exactly one of them, so
// the two ranges meet and [0, n0) / [n0, B*F) are the bucket-major ranges the setup launches expect.
// A CTA reserves its two ranges with one atomic each (ranks inside the CTA from shared counters);
// the last CTA to finish publishes the boundaries.  bucket_count[0..2] serve as the two cursors and
// the ticket (zeroed with the other counters before every forward pass).
__global__ void __launch_bounds__(256) k_depth_split2(FwdParams p, float min_overdraw, int use_facing, int hint_union)
{
    if (!plan_is_ordered(p, min_overdraw)) return;
    __shared__ uint32_t s_cnt[2], s_base[2];
    const int b = blockIdx.y, t = threadIdx.x;
    if (t < 2) s_cnt[t] = 0;
    __syncthreads();
    const float pivot = p.pivots[b * kMaxBuckets];
    const uint32_t gen = hint_generation(p);
    const int front = use_facing ? (int)p.pivots[b * kMaxBuckets + kMaxBuckets - 1] : 0;
    const int f0 = blockIdx.x * kChunkFaces, f1 = min(f0 + kChunkFaces, p.F);
    uint32_t rank[kChunkFaces / 256];
    uint32_t kbits = 0;
#pragma unroll
    for (int j = 0; j < kChunkFaces / 256; ++j) {
        const int f = f0 + t + j * 256;
        if (f < f1) {
            int k;
            if (gen && !hint_union) {
                // hinted: the faces that won a pixel in the previous call go first, everything else behind the
                // pyramid they leave (no depth is looked at: one byte per face)
                k = p.vis_flags[(size_t)b * p.F + f] == (uint8_t)gen ? 0 : 1;
            } else if (gen) {
                // hinted faces and the nearest tenth together go first (measured better than the hinted faces alone)
                const float z = face_near_depth(p, b, f, nullptr, nullptr);
                k = (p.vis_flags[(size_t)b * p.F + f] == (uint8_t)gen || z < p.pivots[b * kMaxBuckets + kMaxBuckets - 2]) ? 0 : 1;
            } else {
                int facing = 0;
                const float z = face_near_depth(p, b, f, nullptr, use_facing ? &facing : nullptr);
                k = (front == 0 || facing == front || z == 0.0f) ? (z >= pivot ? 1 : 0) : 1;
            }
            kbits |= (uint32_t)k << j;
            rank[j] = atomicAdd(&s_cnt[k], 1u);
        }
    }
    __syncthreads();
    if (t < 2 && s_cnt[t]) s_base[t] = atomicAdd(p.bucket_count + t, s_cnt[t]);
    __syncthreads();
    const uint32_t total = (uint32_t)((size_t)p.B * p.F);
#pragma unroll
    for (int j = 0; j < kChunkFaces / 256; ++j) {
        const int f = f0 + t + j * 256;
        if (f < f1) {
            const int k = (kbits >> j) & 1;
            const uint32_t at = s_base[k] + rank[j];
            p.order[k == 0 ? at : total - 1u - at] = (uint32_t)((size_t)b * p.F + f);
        }
    }
    // publish the boundaries once every CTA has reserved its ranges
    __syncthreads();
    if (t == 0) {
        __threadfence();
        const uint32_t done = atomicAdd(p.bucket_count + 2, 1u);
        if (done == gridDim.x * gridDim.y - 1) {
            __threadfence();
            p.bucket_base[0] = 0u;
            p.bucket_base[1] = atomicAdd(p.bucket_count + 0, 0u);
            p.bucket_base[2] = total;
        }
    }
}

// Depth pyramid of one 64 x 64 pixel block per CTA: warp w reduces pixel rows 8w .. 8w+7 to
// eight 8x8 tiles, then levels 1..3 are reduced from shared memory.
__global__ void __launch_bounds__(256) k_hiz_build(FwdParams p, float min_overdraw)
{
    if (!plan_is_ordered(p, min_overdraw)) return;
    __shared__ uint32_t s_t[8][8], s_u[4][4], s_v[2][2];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int bx = blockIdx.x, by = blockIdx.y, b = blockIdx.z;
    uint32_t m0 = 0, m1 = 0;
#pragma unroll
    for (int dy = 0; dy < 8; ++dy) {
        const int py = by * 64 + warp * 8 + dy;                      // window row, y up
        if (py < p.H) {
            const unsigned long long *row = p.keys + ((size_t)b * p.H + (size_t)(p.H - 1 - py)) * p.W;
            const int x0 = bx * 64 + lane, x1 = x0 + 32;
            if (x0 < p.W) m0 = max(m0, reinterpret_cast<const uint32_t *>(row + x0)[1]);
            if (x1 < p.W) m1 = max(m1, reinterpret_cast<const uint32_t *>(row + x1)[1]);
        }
    }
#pragma unroll
    for (int o = 1; o < 8; o <<= 1) {
        m0 = max(m0, __shfl_xor_sync(kFull, m0, o));
        m1 = max(m1, __shfl_xor_sync(kFull, m1, o));
    }
    if ((lane & 7) == 0) { s_t[warp][lane >> 3] = m0; s_t[warp][4 + (lane >> 3)] = m1; }
    __syncthreads();
    const int t = threadIdx.x;
    auto store = [&](int level, int ty, int tx, uint32_t v) {
        const int sh = kHizLog2 + level;
        const int ntx = (p.W + (1 << sh) - 1) >> sh, nty = (p.H + (1 << sh) - 1) >> sh;
        const int gx = (bx << (3 - level)) + tx, gy = (by << (3 - level)) + ty;
        if (gx < ntx && gy < nty) p.hiz[p.hiz_off[level] + ((size_t)b * nty + gy) * ntx + gx] = v;
    };
    if (t < 64) store(0, t >> 3, t & 7, s_t[t >> 3][t & 7]);
    if (t < 16) {
        const int y = t >> 2, x = t & 3;
        const uint32_t v = max(max(s_t[2 * y][2 * x], s_t[2 * y][2 * x + 1]), max(s_t[2 * y + 1][2 * x], s_t[2 * y + 1][2 * x + 1]));
        s_u[y][x] = v;
        store(1, y, x, v);
    }
    __syncthreads();
    if (t < 4) {
        const int y = t >> 1, x = t & 1;
        const uint32_t v = max(max(s_u[2 * y][2 * x], s_u[2 * y][2 * x + 1]), max(s_u[2 * y + 1][2 * x], s_u[2 * y + 1][2 * x + 1]));
        s_v[y][x] = v;
        store(2, y, x, v);
    }
    __syncthreads();
    if (t == 0) store(3, 0, 0, max(max(s_v[0][0], s_v[0][1]), max(s_v[1][0], s_v[1][1])));
}

}  // namespace

// Cumulative face fractions of the buckets.  DIRT_B200_BUCKETS="0.1,0.3,1" overrides the
// default (tuning aid); "1" selects a single ordered pass, "0" the unordered single pass.
// DIRT_B200_MIN_OVERDRAW overrides the estimated fragments-per-pixel threshold below which the
// kernels of an ordered plan fall back to one unordered pass (0 = always ordered).
BucketPlan bucket_plan()
{
    static const BucketPlan plan = [] {
        BucketPlan q;
        q.min_overdraw = 4.0f;
        if (const char *e = std::getenv("DIRT_B200_MIN_OVERDRAW")) q.min_overdraw = std::strtof(e, nullptr);
        // measured best on the 100k-face workloads (profiles/r1_setup_notes.md)
        q.n = 2; q.upto[0] = 0.35f; q.upto[1] = 1.0f;
        q.nlaunch = 2; q.launch_first[0] = 0; q.launch_first[1] = 1; q.launch_first[2] = 2;
        q.facing = false;
        if (const char *e = std::getenv("DIRT_B200_FACING")) q.facing = e[0] == '1';
        if (const char *e = std::getenv("DIRT_B200_BUCKETS")) {
            if (e[0] == '0' && e[1] == 0) { q.n = 0; q.nlaunch = 0; q.facing = false; return q; }
            // "a,b;c,1": ';' ends a launch, ',' only a bucket inside it; without any ';' every bucket is a launch
            bool grouped = false;
            for (const char *c = e; *c; ++c) grouped = grouped || *c == ';';
            BucketPlan r; r.n = 0; r.min_overdraw = q.min_overdraw; r.facing = q.facing;
            r.nlaunch = 0; r.launch_first[0] = 0;
            float last = 0.0f;
            while (*e && r.n < kMaxBuckets) {
                char *end = nullptr;
                const float v = std::strtof(e, &end);
                if (end == e) break;
                if (v > last && v <= 1.0f) { r.upto[r.n++] = v; last = v; }
                if ((*end == ';' || !grouped) && r.n > r.launch_first[r.nlaunch]) r.launch_first[++r.nlaunch] = r.n;
                e = (*end == ',' || *end == ';') ? end + 1 : end;
                if (*end != ',' && *end != ';') break;
            }
            if (r.n > 0) {
                r.upto[r.n - 1] = 1.0f;
                if (r.launch_first[r.nlaunch] != r.n) r.launch_first[++r.nlaunch] = r.n;
                q = r;
            }
        }
        return q;
    }();
    return plan;
}

void launch_depth_order(const FwdParams &p, const BucketPlan &plan, int sm_count, cudaStream_t stream)
{
    (void)sm_count;
    const dim3 chunks((unsigned)((p.F + kChunkFaces - 1) / kChunkFaces), (unsigned)p.B);
    k_depth_pivots<<<p.B, kSamples, 0, stream>>>(p, plan);
    if (plan.n == 2) {
        // with a hint: the hinted faces AND the nearest tenth go first (DIRT_B200_HINT_UNION=0: the hinted faces alone)
        static const int hint_union = [] { const char *e = std::getenv("DIRT_B200_HINT_UNION"); return (e && e[0] == '0') ? 0 : 1; }();
        k_depth_split2<<<chunks, 256, 0, stream>>>(p, plan.min_overdraw, plan.facing ? 1 : 0, hint_union);
        return;
    }
    k_depth_classify<<<chunks, 256, 0, stream>>>(p, plan.n, plan.min_overdraw, plan.facing ? 1 : 0);
    k_depth_offsets<<<1, 1024, 0, stream>>>(p, plan.n, plan.min_overdraw);
    k_depth_scatter<<<chunks, 256, 0, stream>>>(p, plan.n, plan.min_overdraw);
}

void launch_hiz_build(const FwdParams &p, const BucketPlan &plan, int sm_count, cudaStream_t stream)
{
    (void)sm_count;
    const dim3 g((unsigned)((p.W + 63) / 64), (unsigned)((p.H + 63) / 64), (unsigned)p.B);
    k_hiz_build<<<g, 256, 0, stream>>>(p, plan.min_overdraw);
}

}  // namespace dirt
