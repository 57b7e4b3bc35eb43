// kernels.h -- parameter blocks passed by value to the kernels, and the launch entry points
// capi.cu calls.  Internal to libdirt_b200.so; the public surface is include/dirt_b200.h.
#pragma once
#include <cstddef>
#include <cstdint>
#include <cuda_runtime.h>

namespace dirt {

enum { kCtrBig = 0, kCtrQueue = 1, kCtrOverdraw = 2, kCtrActiveTiles = 3, kCtrSlow = 8, kCtrCount = 12 };   // 4..7, 9, 10: -DDIRT_STATS; kCtrSlow: some vertex needs the clip path; kCtrOverdraw: float bits, see depth_order.cu

struct FwdParams {
    int B, H, W, C, V, F;
    int tiles_x, tiles_y;
    const float *background;     int64_t background_bs;
    const float *vertices;       int64_t vertices_bs;
    const float *vertex_colors;  int64_t vertex_colors_bs;
    const int32_t *faces;        int64_t faces_bs;
    float *pixels;
    int32_t *face_id;            // optional (with gbuffer)
    float *gbuffer;
    // workspace carve-up (see workspace_layout in capi.cu)
    const int4 *proj;            // [mB*V] projected vertices (k_project); mB = 1 when vertices are broadcast
    int64_t proj_bs;             // V, or 0 when vertices are broadcast
    float *zvert;                // [mB*V] the window depth of every projected vertex again, packed (0 for a vertex that needs the
                                 // clipper): the ordering pass gathers three of these per face instead of three 16-byte records
    unsigned long long *keys;    // [B*H*W] depth24 << 32 | face
    uint32_t *counters;          // [kCtrCount]; counters, tile_count, tile_cursor are contiguous
    uint32_t *tile_count;        // [B*tiles]
    uint32_t *tile_cursor;       // [B*tiles]
    uint32_t *tile_offset;       // [B*tiles]
    uint32_t *active_tiles;      // [B*tiles] tiles with a non-empty queue, counters[kCtrActiveTiles] of them
    void *big;                   // [cap_big] BigTri (16 B each)
    uint32_t *queue;             // [cap_queue + 4]
    uint32_t cap_big;
    unsigned long long cap_queue;
    size_t zero_bytes;           // bytes to clear starting at `counters`
    // depth-ordered passes with a hierarchical depth buffer (depth_order.cu); order == nullptr: off
    uint32_t *order;             // [B*F] flat face numbers (b * F + f), grouped by depth bucket
    uint8_t *bucket_of;          // [B*F]
    float *pivots;               // [B * kMaxBuckets] bucket boundaries per element
    uint32_t *bucket_count;      // [kMaxBuckets * B], bucket-major; inside the zeroed region
    unsigned long long *work_next; // [kMaxBuckets] next unclaimed entry of each setup launch; inside the zeroed region
    uint32_t *bucket_cursor;     // [kMaxBuckets * B]
    uint32_t *bucket_base;       // [kMaxBuckets + 1] first entry of each bucket in `order`
    // visibility hint (optional, caller-owned, persistent between calls; see include/dirt_b200.h): which faces won a
    // pixel in the previous call on this mesh.  Only ORDERS the work (those faces are sampled first), never decides it.
    uint32_t *hint_hdr;          // [kHintHeaderWords] magic, B, F, generation; nullptr: no hint
    uint8_t *vis_flags;          // [B*F] generation in which the face last won a pixel
    uint32_t *hiz;               // max depth24 per tile (window space, y up); level l has (8 << l)-px tiles and
    uint32_t hiz_off[4];         // starts at hiz + hiz_off[l], laid out [B][tiles_y(l)][tiles_x(l)]
};

constexpr int kMaxBuckets = 8;
constexpr int kHintHeaderWords = 64;             // 256 bytes: the flags behind it stay 256-byte aligned
constexpr uint32_t kHintMagic = 0x54524944u;     // "DIRT"
constexpr int kHizLog2 = 3;      // hierarchical-depth tile edge at level 0: 8 px
constexpr int kHizLevels = 4;    // 8, 16, 32, 64 px
constexpr int kOrderMinFaces = 4096;   // per element; below this one unordered pass is cheaper

struct BwdParams {
    int B, H, W, C, V, F;
    const int32_t *faces;        int64_t faces_bs;
    const float *pixels;
    const float *grad_pixels;
    const int32_t *face_id;
    const float *gbuffer;
    float *grad_vertices;        int64_t grad_vertices_bs;
    float *grad_vertex_colors;   int64_t grad_vertex_colors_bs;
    float *grad_background;      // optional
    // shared-mesh mode (grad_packed != nullptr): world-space position + colour gradients of all
    // batch elements accumulate into grad_packed [V, packed_stride]; grad_vertices / colours unused
    const float *matrices;       // [B,4,4] clip = world @ M_b
    float *grad_packed;
    int packed_stride;
    // deterministic mode (det_pos != nullptr): fixed-point accumulators and the call's largest |contribution|
    long long *det_pos;          // [nbv*V*4] (packed: [V*4]); det_col follows it in the same allocation
    long long *det_col;          // [nbc*V*C] (packed: [V*C])
    unsigned *det_max;           // [2] float bits: position, colour
};

enum Stage {
    kStageClear = 0, kStageProject, kStageDepthOrder, kStageSetup, kStageHizBuild, kStageBinCount, kStageBinAlloc, kStageBinFill,
    kStageTileRaster, kStageResolve, kStageGradClear, kStageGradAssemble, kStageCount
};

// Records CUDA event pairs around each stage when enabled (capi.cu owns the storage).  A stage
// may run several times per call (the depth-bucket launches): each run takes the next slot.
struct StageTimer {
    static constexpr int kSlots = 8;
    bool enabled = false;
    cudaEvent_t ev[kStageCount][kSlots][2] = {};
    int used[kStageCount] = {};
    void begin(int s, cudaStream_t st) { if (enabled && used[s] < kSlots) cudaEventRecord(ev[s][used[s]][0], st); }
    void end(int s, cudaStream_t st) { if (enabled && used[s] < kSlots) { cudaEventRecord(ev[s][used[s]][1], st); ++used[s]; } }
};

cudaError_t launch_forward(const FwdParams &p, int sm_count, cudaStream_t stream, uint64_t *launches, StageTimer &tm);
// depth_order.cu
struct BucketPlan {
    int n; float upto[kMaxBuckets];      // cumulative face fractions, upto[n-1] = 1
    int nlaunch;                         // buckets are sampled in `nlaunch` launches (pyramid rebuilt between them);
    int launch_first[kMaxBuckets + 1];   // launch j covers buckets launch_first[j] .. launch_first[j+1]-1, nearest first
    float min_overdraw;                  // estimated fragments per pixel below which ordering does not pay
    bool facing;                         // faces of the winding that looks away from the viewer go to the last bucket
};
BucketPlan bucket_plan();
void launch_depth_order(const FwdParams &p, const BucketPlan &plan, int sm_count, cudaStream_t stream);
void launch_hiz_build(const FwdParams &p, const BucketPlan &plan, int sm_count, cudaStream_t stream);
cudaError_t launch_dilate(const float *x, float *out, int B, int H, int W, int C, int sm_count, cudaStream_t stream);
cudaError_t launch_dilate_grad(const float *x, const float *grad_out, float *grad_x, int B, int H, int W, int C,
                               int sm_count, cudaStream_t stream);
cudaError_t launch_vertex_normals(const float *vertices, int B, int V, int D, const int32_t *faces, int F,
                                  float *sums, float *normals, cudaStream_t stream);
cudaError_t launch_vertex_normals_grad(const float *vertices, int B, int V, int D, const int32_t *faces, int F, const float *sums,
                                       const float *grad_normals, float *grad_sums, float *grad_vertices, cudaStream_t stream);
cudaError_t launch_backward(const BwdParams &p, int sm_count, cudaStream_t stream, uint64_t *launches, StageTimer &tm);
// capi.cu: adds to the counter behind dirt_kernel_launch_count (scene stages launch outside capi.cu)
void note_launches(uint64_t n);

}  // namespace dirt
