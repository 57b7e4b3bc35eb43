// capi.cu -- the extern "C" surface declared in include/dirt_b200.h: argument validation
// (the OP_REQUIRES block of csrc/rasterise_egl.cpp:309-336), workspace carve-up, and the
// host-buffer convenience calls.
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>
#include "../../include/dirt_b200.h"
#include "kernels.h"

namespace {

thread_local char g_error[512] = "";
std::atomic<uint64_t> g_launches{0};

// stage timing (dirt_profile_*): one global timer, guarded by a mutex when (re)configured
std::mutex g_prof_mu;
dirt::StageTimer g_timer;
double g_stage_ms[dirt::kStageCount] = {};
uint64_t g_stage_calls[dirt::kStageCount] = {};
const char *const kStageNames[dirt::kStageCount] = {
    "clear", "project", "depth_order", "setup", "hiz_build", "bin_count", "bin_alloc", "bin_fill", "tile_raster", "resolve",
    "grad_clear", "grad_assemble"};

int fail(int status, const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_error, sizeof(g_error), fmt, ap);
    va_end(ap);
    return status;
}

int cuda_fail(cudaError_t e, const char *what)
{
    return fail(DIRT_STATUS_CUDA_ERROR, "%s: %s", what, cudaGetErrorString(e));
}

constexpr int kTilePx = 16;
constexpr size_t kAlign = 256;
inline size_t align_up(size_t v) { return (v + kAlign - 1) / kAlign * kAlign; }

// Workspace carve-up.  Offsets are a pure function of the shape.
struct Layout {
    size_t keys, proj, counters, tile_count, tile_cursor, bucket_count, tile_offset, big, queue, total;
    size_t order, bucket_of, pivots, bucket_cursor, bucket_base, hiz, work_next, active_tiles, zvert;
    uint32_t hiz_off[dirt::kHizLevels];
    size_t zero_bytes;
    uint32_t cap_big;
    unsigned long long cap_queue;
    int tiles_x, tiles_y;
};

Layout workspace_layout(const DirtShape &s)
{
    Layout L;
    L.tiles_x = (s.width + kTilePx - 1) / kTilePx;
    L.tiles_y = (s.height + kTilePx - 1) / kTilePx;
    const size_t npix = (size_t)s.batch * s.height * s.width;
    const size_t ntiles = (size_t)s.batch * L.tiles_x * L.tiles_y;
    const size_t nfaces = (size_t)s.batch * s.face_count;
    // Large-triangle list: one slot per face is enough unless many faces are split by the
    // near/far planes; the kernels fall back to in-thread sampling when a list is full, so
    // these capacities bound memory, not correctness.
    size_t cap_big = nfaces + 1024;
    if (cap_big > 0xFFFFFFF0ull) cap_big = 0xFFFFFFF0ull;
    L.cap_big = (uint32_t)cap_big;
    L.cap_queue = 8ull * ntiles + 4ull * nfaces + 4096ull;
    size_t off = 0;
    L.keys = off;        off = align_up(off + npix * sizeof(unsigned long long));
    L.proj = off;        off = align_up(off + (size_t)s.batch * s.vertex_count * 16);
    L.zvert = off;       off = align_up(off + (size_t)s.batch * s.vertex_count * sizeof(float));
    L.counters = off;    off += dirt::kCtrCount * sizeof(uint32_t);
    L.tile_count = off;  off += ntiles * sizeof(uint32_t);
    L.tile_cursor = off; off += ntiles * sizeof(uint32_t);
    L.bucket_count = off; off += (size_t)dirt::kMaxBuckets * s.batch * sizeof(uint32_t);
    off = (off + 7) & ~(size_t)7;
    L.work_next = off;   off += dirt::kMaxBuckets * sizeof(unsigned long long);
    L.zero_bytes = off - L.counters;
    off = align_up(off);
    L.tile_offset = off; off = align_up(off + ntiles * sizeof(uint32_t));
    L.active_tiles = off; off = align_up(off + ntiles * sizeof(uint32_t));
    L.big = off;         off = align_up(off + (size_t)L.cap_big * 16);
    L.queue = off;       off = align_up(off + (size_t)(L.cap_queue + 4) * sizeof(uint32_t));
    // depth-ordered passes (depth_order.cu)
    size_t hiz_words = 0;
    for (int l = 0; l < dirt::kHizLevels; ++l) {
        const int sh = dirt::kHizLog2 + l;
        L.hiz_off[l] = (uint32_t)hiz_words;
        hiz_words += (size_t)s.batch * ((s.width + (1 << sh) - 1) >> sh) * ((s.height + (1 << sh) - 1) >> sh);
    }
    L.order = off;         off = align_up(off + nfaces * sizeof(uint32_t));
    L.bucket_of = off;     off = align_up(off + nfaces);
    L.pivots = off;        off = align_up(off + (size_t)s.batch * dirt::kMaxBuckets * sizeof(float));
    L.bucket_cursor = off; off = align_up(off + (size_t)dirt::kMaxBuckets * s.batch * sizeof(uint32_t));
    L.bucket_base = off;   off = align_up(off + (dirt::kMaxBuckets + 1) * sizeof(uint32_t));
    L.hiz = off;           off = align_up(off + hiz_words * sizeof(uint32_t));
    L.total = off;
    return L;
}

int check_shape(const DirtShape &s)
{
    // csrc/hwc.h:27-28: CHECK(channels == 1 || channels == 3), CHECK(width > 0 && height > 0);
    // channels is widened to [1, DIRT_MAX_CHANNELS] here (deferred G-buffers in one call)
    if (s.height <= 0 || s.width <= 0) return fail(DIRT_STATUS_INVALID_ARGUMENT, "Rasterise expects height > 0 and width > 0");
    if (s.height > 32768 || s.width > 32768) return fail(DIRT_STATUS_UNSUPPORTED, "height/width above 32768 are not supported");
    if (s.channels < 1 || s.channels > DIRT_MAX_CHANNELS) return fail(DIRT_STATUS_INVALID_ARGUMENT, "Rasterise expects 1 <= channels <= %d", DIRT_MAX_CHANNELS);
    if (s.batch < 0 || s.vertex_count < 0 || s.face_count < 0) return fail(DIRT_STATUS_INVALID_ARGUMENT, "negative batch / vertex / face count");
    if ((unsigned long long)s.face_count >= (1ull << 30)) return fail(DIRT_STATUS_UNSUPPORTED, "face_count must be below 2^30");
    if ((unsigned long long)s.batch * s.face_count >= (1ull << 40)) return fail(DIRT_STATUS_UNSUPPORTED, "batch * face_count too large");
    return DIRT_STATUS_OK;
}

int check_stride(int64_t stride, int64_t block, const char *name)
{
    if (stride != 0 && stride < block)
        return fail(DIRT_STATUS_INVALID_ARGUMENT, "%s batch stride %lld is neither 0 (broadcast) nor >= %lld", name, (long long)stride, (long long)block);
    return DIRT_STATUS_OK;
}

int current_sm_count(int *out)
{
    static std::mutex m;
    static int cached[64];
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetDevice (no CUDA device: libdirt_b200 has no CPU fallback)");
    std::lock_guard<std::mutex> lock(m);
    if (dev >= 0 && dev < 64 && cached[dev] > 0) { *out = cached[dev]; return DIRT_STATUS_OK; }
    int n = 0;
    e = cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) return cuda_fail(e, "cudaDeviceGetAttribute");
    if (dev >= 0 && dev < 64) cached[dev] = n;
    *out = n;
    return DIRT_STATUS_OK;
}

}  // namespace

namespace dirt { void note_launches(uint64_t n) { g_launches += n; } }
namespace dirt {
int set_error(int status, const char *what, cudaError_t e)
{
    return e == cudaSuccess ? fail(status, "%s", what) : fail(status, "%s: %s", what, cudaGetErrorString(e));
}
}

extern "C" {

int dirt_abi_version(void) { return DIRT_B200_ABI_VERSION; }

const char *dirt_last_error(void) { return g_error; }

uint64_t dirt_kernel_launch_count(void) { return g_launches.load(); }

size_t dirt_workspace_bytes(const DirtShape *shape)
{
    if (!shape || check_shape(*shape) != DIRT_STATUS_OK) return 0;
    return workspace_layout(*shape).total;
}

size_t dirt_visibility_hint_bytes(const DirtShape *shape)
{
    if (!shape || check_shape(*shape) != DIRT_STATUS_OK) return 0;
    return align_up(dirt::kHintHeaderWords * sizeof(uint32_t) + (size_t)shape->batch * shape->face_count);
}

}  // extern "C"

namespace {

// aux_only: the stateless backward re-derives face_id / gbuffer; no colour is read or written
int forward_impl(const DirtForwardArgs *a, void *stream, bool aux_only)
{
    if (!a) return fail(DIRT_STATUS_INVALID_ARGUMENT, "null arguments");
    const DirtShape &s = a->shape;
    int rc = check_shape(s);
    if (rc) return rc;
    const int64_t npix = (int64_t)s.height * s.width;
    // the four OP_REQUIRES shape checks of rasterise_egl.cpp:310-321 become stride checks here
    if (!aux_only && (rc = check_stride(a->background_bstride, npix * s.channels, "background"))) return rc;
    if ((rc = check_stride(a->vertices_bstride, (int64_t)s.vertex_count * 4, "vertices"))) return rc;
    if (!aux_only && (rc = check_stride(a->vertex_colors_bstride, (int64_t)s.vertex_count * s.channels, "vertex_colors"))) return rc;
    if ((rc = check_stride(a->faces_bstride, (int64_t)s.face_count * 3, "faces"))) return rc;
    if (s.batch == 0) return DIRT_STATUS_OK;
    if (!aux_only && (!a->background || !a->pixels)) return fail(DIRT_STATUS_INVALID_ARGUMENT, "background and pixels must not be null");
    if (s.face_count > 0 && (!a->vertices || (!aux_only && !a->vertex_colors) || !a->faces))
        return fail(DIRT_STATUS_INVALID_ARGUMENT, "vertices, vertex_colors and faces must not be null when face_count > 0");
    if ((a->face_id == nullptr) != (a->gbuffer == nullptr))
        return fail(DIRT_STATUS_INVALID_ARGUMENT, "face_id and gbuffer must be both null or both set");
    if (((uintptr_t)a->vertices & 15u) || ((uintptr_t)a->gbuffer & 15u) || ((a->vertices_bstride * 4) & 15))
        return fail(DIRT_STATUS_INVALID_ARGUMENT, "vertices and gbuffer must be 16-byte aligned");
    const Layout L = workspace_layout(s);
    if (!a->workspace || a->workspace_bytes < L.total)
        return fail(DIRT_STATUS_WORKSPACE_TOO_SMALL, "workspace of %zu bytes given, %zu needed", a->workspace_bytes, L.total);
    if ((uintptr_t)a->workspace & (kAlign - 1))
        return fail(DIRT_STATUS_INVALID_ARGUMENT, "workspace must be %zu-byte aligned", kAlign);
    if (a->visibility_hint) {
        const size_t need = dirt_visibility_hint_bytes(&s);
        if (a->visibility_hint_bytes < need)
            return fail(DIRT_STATUS_INVALID_ARGUMENT, "visibility_hint of %zu bytes given, %zu needed", a->visibility_hint_bytes, need);
        if ((uintptr_t)a->visibility_hint & 15u) return fail(DIRT_STATUS_INVALID_ARGUMENT, "visibility_hint must be 16-byte aligned");
    }
    int sms = 0;
    if ((rc = current_sm_count(&sms))) return rc;

    dirt::FwdParams p;
    p.B = s.batch; p.H = s.height; p.W = s.width; p.C = s.channels; p.V = s.vertex_count; p.F = s.face_count;
    p.tiles_x = L.tiles_x; p.tiles_y = L.tiles_y;
    p.background = a->background;       p.background_bs = a->background_bstride;
    p.vertices = a->vertices;           p.vertices_bs = a->vertices_bstride;
    p.vertex_colors = a->vertex_colors; p.vertex_colors_bs = a->vertex_colors_bstride;
    p.faces = a->faces;                 p.faces_bs = a->faces_bstride;
    p.pixels = aux_only ? nullptr : a->pixels; p.face_id = a->face_id; p.gbuffer = a->gbuffer;
    char *ws = static_cast<char *>(a->workspace);
    p.keys = reinterpret_cast<unsigned long long *>(ws + L.keys);
    p.proj = reinterpret_cast<const int4 *>(ws + L.proj);
    p.proj_bs = a->vertices_bstride == 0 ? 0 : s.vertex_count;
    p.zvert = reinterpret_cast<float *>(ws + L.zvert);
    p.counters = reinterpret_cast<uint32_t *>(ws + L.counters);
    p.tile_count = reinterpret_cast<uint32_t *>(ws + L.tile_count);
    p.tile_cursor = reinterpret_cast<uint32_t *>(ws + L.tile_cursor);
    p.tile_offset = reinterpret_cast<uint32_t *>(ws + L.tile_offset);
    p.active_tiles = reinterpret_cast<uint32_t *>(ws + L.active_tiles);
    p.big = ws + L.big;
    p.queue = reinterpret_cast<uint32_t *>(ws + L.queue);
    p.cap_big = L.cap_big; p.cap_queue = L.cap_queue; p.zero_bytes = L.zero_bytes;
    // depth-ordered passes pay off on meshes with real depth complexity; flat face numbers are 32-bit
    const bool ordered = s.face_count >= dirt::kOrderMinFaces && s.batch <= 65535 && (unsigned long long)s.batch * s.face_count < (1ull << 32);
    p.order = ordered ? reinterpret_cast<uint32_t *>(ws + L.order) : nullptr;
    // the hint orders the depth-ordered passes; without them it is neither read nor written
    const bool hinted = ordered && a->visibility_hint != nullptr;
    p.hint_hdr = hinted ? static_cast<uint32_t *>(a->visibility_hint) : nullptr;
    p.vis_flags = hinted ? static_cast<uint8_t *>(a->visibility_hint) + dirt::kHintHeaderWords * sizeof(uint32_t) : nullptr;
    p.bucket_of = reinterpret_cast<uint8_t *>(ws + L.bucket_of);
    p.pivots = reinterpret_cast<float *>(ws + L.pivots);
    p.bucket_count = reinterpret_cast<uint32_t *>(ws + L.bucket_count);
    p.work_next = reinterpret_cast<unsigned long long *>(ws + L.work_next);
    p.bucket_cursor = reinterpret_cast<uint32_t *>(ws + L.bucket_cursor);
    p.bucket_base = reinterpret_cast<uint32_t *>(ws + L.bucket_base);
    p.hiz = reinterpret_cast<uint32_t *>(ws + L.hiz);
    for (int l = 0; l < dirt::kHizLevels; ++l) p.hiz_off[l] = L.hiz_off[l];

    uint64_t n = 0;
    cudaError_t e = dirt::launch_forward(p, sms, static_cast<cudaStream_t>(stream), &n, g_timer);
    g_launches += n;
    if (e != cudaSuccess) return cuda_fail(e, "rasterise forward launch");
    return DIRT_STATUS_OK;
}

// Deterministic mode: carves the fixed-point accumulators out of `ws` (dirt_deterministic_workspace_bytes).
int place_deterministic(dirt::BwdParams &p, const DirtShape &s, void *ws, size_t ws_bytes, size_t offset, bool packed)
{
    const size_t need = offset + align_up(kAlign + (size_t)s.batch * s.vertex_count * (4 + s.channels) * sizeof(long long));
    if (!ws || ws_bytes < need)
        return fail(DIRT_STATUS_WORKSPACE_TOO_SMALL, "deterministic backward: workspace of %zu bytes given, %zu needed", ws ? ws_bytes : (size_t)0, need);
    if ((uintptr_t)ws & (kAlign - 1)) return fail(DIRT_STATUS_INVALID_ARGUMENT, "workspace must be %zu-byte aligned", kAlign);
    char *base = static_cast<char *>(ws) + offset;
    p.det_max = reinterpret_cast<unsigned *>(base);
    p.det_pos = reinterpret_cast<long long *>(base + kAlign);
    const size_t nbv = (packed || p.grad_vertices_bs == 0) ? 1 : (size_t)s.batch;
    p.det_col = p.det_pos + nbv * s.vertex_count * 4;
    return DIRT_STATUS_OK;
}

}  // namespace

extern "C" {

int dirt_rasterise_forward(const DirtForwardArgs *a, void *stream) { return forward_impl(a, stream, false); }

size_t dirt_backward_workspace_bytes(const DirtShape *shape)
{
    if (!shape || check_shape(*shape) != DIRT_STATUS_OK) return 0;
    const size_t npix = (size_t)shape->batch * shape->height * shape->width;
    return align_up(workspace_layout(*shape).total) + align_up(npix * sizeof(int32_t)) + align_up(npix * 4 * sizeof(float));
}

size_t dirt_deterministic_workspace_bytes(const DirtShape *shape)
{
    if (!shape || check_shape(*shape) != DIRT_STATUS_OK) return 0;
    return align_up(kAlign + (size_t)shape->batch * shape->vertex_count * (4 + shape->channels) * sizeof(long long));
}

int dirt_workspace_create(const DirtShape *shape, void **workspace, size_t *workspace_bytes)
{
    if (!workspace) return fail(DIRT_STATUS_INVALID_ARGUMENT, "null arguments");
    *workspace = nullptr;
    // covers dirt_workspace_bytes(shape), the stateless and the deterministic backward
    const size_t n = dirt_backward_workspace_bytes(shape) + dirt_deterministic_workspace_bytes(shape);
    if (n == 0) return fail(DIRT_STATUS_INVALID_ARGUMENT, "invalid shape");
    const cudaError_t e = cudaMalloc(workspace, n);
    if (e != cudaSuccess) return cuda_fail(e, "cudaMalloc (workspace)");
    if (workspace_bytes) *workspace_bytes = n;
    return DIRT_STATUS_OK;
}

int dirt_workspace_destroy(void *workspace)
{
    if (!workspace) return DIRT_STATUS_OK;
    const cudaError_t e = cudaFree(workspace);
    return e == cudaSuccess ? DIRT_STATUS_OK : cuda_fail(e, "cudaFree (workspace)");
}

int dirt_rasterise_backward(const DirtBackwardArgs *a, void *stream)
{
    if (!a) return fail(DIRT_STATUS_INVALID_ARGUMENT, "null arguments");
    const DirtShape &s = a->shape;
    int rc = check_shape(s);
    if (rc) return rc;
    if ((rc = check_stride(a->faces_bstride, (int64_t)s.face_count * 3, "faces"))) return rc;
    if ((rc = check_stride(a->grad_vertices_bstride, (int64_t)s.vertex_count * 4, "grad_vertices"))) return rc;
    if ((rc = check_stride(a->grad_vertex_colors_bstride, (int64_t)s.vertex_count * s.channels, "grad_vertex_colors"))) return rc;
    if (s.batch == 0) return DIRT_STATUS_OK;
    if (!a->pixels || !a->grad_pixels) return fail(DIRT_STATUS_INVALID_ARGUMENT, "pixels and grad_pixels must not be null");
    if ((a->face_id == nullptr) != (a->gbuffer == nullptr))
        return fail(DIRT_STATUS_INVALID_ARGUMENT, "face_id and gbuffer must be both given (from the forward) or both null (re-derived from vertices)");
    if (s.vertex_count > 0 && (!a->grad_vertices || !a->grad_vertex_colors))
        return fail(DIRT_STATUS_INVALID_ARGUMENT, "grad_vertices and grad_vertex_colors must not be null");
    if (s.face_count > 0 && !a->faces) return fail(DIRT_STATUS_INVALID_ARGUMENT, "faces must not be null");
    const int32_t *face_id = a->face_id;
    const float *gbuffer = a->gbuffer;
    if (!face_id) {
        // stateless mode: visibility is re-derived with the forward's kernels (same keys, same G-buffer)
        const size_t need = dirt_backward_workspace_bytes(&s);
        if (!a->workspace || a->workspace_bytes < need)
            return fail(DIRT_STATUS_WORKSPACE_TOO_SMALL, "backward without face_id / gbuffer: workspace of %zu bytes given, %zu needed",
                        a->workspace ? a->workspace_bytes : (size_t)0, need);
        if ((uintptr_t)a->workspace & (kAlign - 1)) return fail(DIRT_STATUS_INVALID_ARGUMENT, "workspace must be %zu-byte aligned", kAlign);
        if (s.face_count > 0 && !a->vertices)
            return fail(DIRT_STATUS_INVALID_ARGUMENT, "backward without face_id / gbuffer needs vertices");
        const size_t npix = (size_t)s.batch * s.height * s.width;
        char *ws = static_cast<char *>(a->workspace);
        const size_t fwd_bytes = align_up(workspace_layout(s).total);
        DirtForwardArgs f = {};
        f.shape = s;
        f.vertices = a->vertices; f.vertices_bstride = a->vertices_bstride;
        f.faces = a->faces;       f.faces_bstride = a->faces_bstride;
        f.face_id = reinterpret_cast<int32_t *>(ws + fwd_bytes);
        f.gbuffer = reinterpret_cast<float *>(ws + fwd_bytes + align_up(npix * sizeof(int32_t)));
        f.workspace = ws; f.workspace_bytes = fwd_bytes;
        if ((rc = forward_impl(&f, stream, true))) return rc;
        face_id = f.face_id; gbuffer = f.gbuffer;
    }
    if (((uintptr_t)gbuffer & 15u) || ((uintptr_t)a->grad_vertices & 15u) || ((a->grad_vertices_bstride * 4) & 15))
        return fail(DIRT_STATUS_INVALID_ARGUMENT, "gbuffer and grad_vertices must be 16-byte aligned");
    int sms = 0;
    if ((rc = current_sm_count(&sms))) return rc;

    dirt::BwdParams p;
    p.B = s.batch; p.H = s.height; p.W = s.width; p.C = s.channels; p.V = s.vertex_count; p.F = s.face_count;
    p.faces = a->faces; p.faces_bs = a->faces_bstride;
    p.pixels = a->pixels; p.grad_pixels = a->grad_pixels; p.face_id = face_id; p.gbuffer = gbuffer;
    p.grad_vertices = a->grad_vertices;           p.grad_vertices_bs = a->grad_vertices_bstride;
    p.grad_vertex_colors = a->grad_vertex_colors; p.grad_vertex_colors_bs = a->grad_vertex_colors_bstride;
    p.grad_background = a->grad_background;
    p.matrices = nullptr; p.grad_packed = nullptr; p.packed_stride = 0;
    p.det_pos = nullptr; p.det_col = nullptr; p.det_max = nullptr;
    if (a->flags & DIRT_BACKWARD_DETERMINISTIC)
        if ((rc = place_deterministic(p, s, a->workspace, a->workspace_bytes, a->face_id ? 0 : align_up(dirt_backward_workspace_bytes(&s)), false))) return rc;
    uint64_t n = 0;
    cudaError_t e = dirt::launch_backward(p, sms, static_cast<cudaStream_t>(stream), &n, g_timer);
    g_launches += n;
    if (e != cudaSuccess) return cuda_fail(e, "rasterise backward launch");
    return DIRT_STATUS_OK;
}

int dirt_shared_grad_stride(int channels) { return channels < 1 ? 0 : (4 + channels + 3) / 4 * 4; }

int dirt_rasterise_backward_shared(const DirtSharedBackwardArgs *a, void *stream)
{
    if (!a) return fail(DIRT_STATUS_INVALID_ARGUMENT, "null arguments");
    const DirtShape &s = a->shape;
    int rc = check_shape(s);
    if (rc) return rc;
    if ((rc = check_stride(a->faces_bstride, (int64_t)s.face_count * 3, "faces"))) return rc;
    if (s.vertex_count > 0 && !a->grad_shared) return fail(DIRT_STATUS_INVALID_ARGUMENT, "grad_shared must not be null");
    if ((uintptr_t)a->grad_shared & 15u) return fail(DIRT_STATUS_INVALID_ARGUMENT, "grad_shared must be 16-byte aligned");
    const int stride = dirt_shared_grad_stride(s.channels);
    if (s.batch == 0) {
        // nothing renders the mesh: the sum over an empty batch is zero
        if (s.vertex_count > 0) {
            cudaError_t e = cudaMemsetAsync(a->grad_shared, 0, (size_t)s.vertex_count * stride * sizeof(float), static_cast<cudaStream_t>(stream));
            if (e != cudaSuccess) return cuda_fail(e, "cudaMemsetAsync");
        }
        return DIRT_STATUS_OK;
    }
    if (!a->pixels || !a->grad_pixels || !a->face_id || !a->gbuffer || !a->clip_from_world)
        return fail(DIRT_STATUS_INVALID_ARGUMENT, "pixels, grad_pixels, face_id, gbuffer and clip_from_world must not be null");
    if (s.face_count > 0 && !a->faces) return fail(DIRT_STATUS_INVALID_ARGUMENT, "faces must not be null");
    if ((uintptr_t)a->gbuffer & 15u) return fail(DIRT_STATUS_INVALID_ARGUMENT, "gbuffer must be 16-byte aligned");
    int sms = 0;
    if ((rc = current_sm_count(&sms))) return rc;

    dirt::BwdParams p;
    p.B = s.batch; p.H = s.height; p.W = s.width; p.C = s.channels; p.V = s.vertex_count; p.F = s.face_count;
    p.faces = a->faces; p.faces_bs = a->faces_bstride;
    p.pixels = a->pixels; p.grad_pixels = a->grad_pixels; p.face_id = a->face_id; p.gbuffer = a->gbuffer;
    p.grad_vertices = nullptr;      p.grad_vertices_bs = 0;
    p.grad_vertex_colors = nullptr; p.grad_vertex_colors_bs = 0;
    p.grad_background = a->grad_background;
    p.matrices = a->clip_from_world; p.grad_packed = a->grad_shared; p.packed_stride = stride;
    p.det_pos = nullptr; p.det_col = nullptr; p.det_max = nullptr;
    if (a->flags & DIRT_BACKWARD_DETERMINISTIC)
        if ((rc = place_deterministic(p, s, a->workspace, a->workspace_bytes, 0, true))) return rc;
    uint64_t n = 0;
    cudaError_t e = dirt::launch_backward(p, sms, static_cast<cudaStream_t>(stream), &n, g_timer);
    g_launches += n;
    if (e != cudaSuccess) return cuda_fail(e, "rasterise shared backward launch");
    return DIRT_STATUS_OK;
}

int dirt_vertex_normals(const float *vertices, int batch, int vertex_count, int vertex_dim, const int32_t *faces, int face_count,
                        float *sums, float *normals, void *stream)
{
    if (batch < 0 || vertex_count < 0 || face_count < 0 || (vertex_dim != 3 && vertex_dim != 4))
        return fail(DIRT_STATUS_INVALID_ARGUMENT, "vertex_normals: vertices must be [batch, V, 3|4], counts non-negative");
    if ((long long)batch * vertex_count == 0) return DIRT_STATUS_OK;
    if (!vertices || !sums || !normals || (face_count > 0 && !faces)) return fail(DIRT_STATUS_INVALID_ARGUMENT, "vertex_normals: null pointer");
    int sms = 0, rc;
    if ((rc = current_sm_count(&sms))) return rc;
    cudaError_t e = dirt::launch_vertex_normals(vertices, batch, vertex_count, vertex_dim, faces, face_count, sums, normals,
                                                static_cast<cudaStream_t>(stream));
    g_launches += face_count > 0 ? 2 : 1;
    if (e != cudaSuccess) return cuda_fail(e, "vertex_normals launch");
    return DIRT_STATUS_OK;
}

int dirt_vertex_normals_grad(const float *vertices, int batch, int vertex_count, int vertex_dim, const int32_t *faces, int face_count,
                             const float *sums, const float *grad_normals, float *grad_sums, float *grad_vertices, void *stream)
{
    if (batch < 0 || vertex_count < 0 || face_count < 0 || (vertex_dim != 3 && vertex_dim != 4))
        return fail(DIRT_STATUS_INVALID_ARGUMENT, "vertex_normals_grad: vertices must be [batch, V, 3|4], counts non-negative");
    if ((long long)batch * vertex_count == 0) return DIRT_STATUS_OK;
    if (!vertices || !grad_normals || !grad_vertices || (sums && !grad_sums) || (face_count > 0 && !faces))
        return fail(DIRT_STATUS_INVALID_ARGUMENT, "vertex_normals_grad: null pointer");
    int sms = 0, rc;
    if ((rc = current_sm_count(&sms))) return rc;
    cudaError_t e = dirt::launch_vertex_normals_grad(vertices, batch, vertex_count, vertex_dim, faces, face_count, sums, grad_normals,
                                                     grad_sums, grad_vertices, static_cast<cudaStream_t>(stream));
    g_launches += face_count > 0 ? 2 : 1;
    if (e != cudaSuccess) return cuda_fail(e, "vertex_normals_grad launch");
    return DIRT_STATUS_OK;
}

int dirt_dilate_silhouette(const float *x, float *out, int batch, int height, int width, int channels, void *stream)
{
    if (batch < 0 || height <= 0 || width <= 0 || channels <= 0) return fail(DIRT_STATUS_INVALID_ARGUMENT, "dilate: bad shape");
    if (batch == 0) return DIRT_STATUS_OK;
    if (!x || !out) return fail(DIRT_STATUS_INVALID_ARGUMENT, "dilate: null pointer");
    int sms = 0, rc;
    if ((rc = current_sm_count(&sms))) return rc;
    cudaError_t e = dirt::launch_dilate(x, out, batch, height, width, channels, sms, static_cast<cudaStream_t>(stream));
    g_launches += 1;
    return e == cudaSuccess ? DIRT_STATUS_OK : cuda_fail(e, "dilate launch");
}

int dirt_dilate_silhouette_grad(const float *x, const float *grad_out, float *grad_x, int batch, int height, int width,
                                int channels, void *stream)
{
    if (batch < 0 || height <= 0 || width <= 0 || channels <= 0) return fail(DIRT_STATUS_INVALID_ARGUMENT, "dilate: bad shape");
    if (batch == 0) return DIRT_STATUS_OK;
    if (!x || !grad_out || !grad_x) return fail(DIRT_STATUS_INVALID_ARGUMENT, "dilate: null pointer");
    int sms = 0, rc;
    if ((rc = current_sm_count(&sms))) return rc;
    cudaError_t e = dirt::launch_dilate_grad(x, grad_out, grad_x, batch, height, width, channels, sms, static_cast<cudaStream_t>(stream));
    g_launches += 1;
    return e == cudaSuccess ? DIRT_STATUS_OK : cuda_fail(e, "dilate grad launch");
}

int g_prof_device = -1;      // the events live on the device that was current when profiling was first enabled

int dirt_profile_enable(int enable)
{
    std::lock_guard<std::mutex> lock(g_prof_mu);
    if (enable) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess) return fail(DIRT_STATUS_CUDA_ERROR, "dirt_profile_enable: no CUDA device");
        if (g_prof_device >= 0 && dev != g_prof_device)
            return fail(DIRT_STATUS_UNSUPPORTED, "stage profiling is per process and bound to device %d; current device is %d", g_prof_device, dev);
        g_prof_device = dev;
    }
    if (enable && !g_timer.ev[0][0][0]) {
        for (int s = 0; s < dirt::kStageCount; ++s)
            for (int r = 0; r < dirt::StageTimer::kSlots; ++r)
                for (int k = 0; k < 2; ++k) {
                    cudaError_t e = cudaEventCreate(&g_timer.ev[s][r][k]);
                    if (e != cudaSuccess) return cuda_fail(e, "cudaEventCreate");
                }
    }
    g_timer.enabled = enable != 0;
    return DIRT_STATUS_OK;
}

int dirt_profile_collect(void)
{
    std::lock_guard<std::mutex> lock(g_prof_mu);
    for (int s = 0; s < dirt::kStageCount; ++s) {
        if (!g_timer.used[s]) continue;
        for (int r = 0; r < g_timer.used[s]; ++r) {
            float ms = 0.f;
            cudaError_t e = cudaEventElapsedTime(&ms, g_timer.ev[s][r][0], g_timer.ev[s][r][1]);
            if (e != cudaSuccess) return cuda_fail(e, "cudaEventElapsedTime (synchronise the stream before dirt_profile_collect)");
            g_stage_ms[s] += ms;
        }
        g_stage_calls[s] += (uint64_t)g_timer.used[s];   // timed runs of the stage (the depth-bucket launches count separately)
        g_timer.used[s] = 0;
    }
    return DIRT_STATUS_OK;
}

int dirt_profile_read(double *total_ms, uint64_t *calls, int n)
{
    std::lock_guard<std::mutex> lock(g_prof_mu);
    for (int s = 0; s < n && s < dirt::kStageCount; ++s) {
        if (total_ms) total_ms[s] = g_stage_ms[s];
        if (calls) calls[s] = g_stage_calls[s];
    }
    return DIRT_STATUS_OK;
}

int dirt_profile_reset(void)
{
    std::lock_guard<std::mutex> lock(g_prof_mu);
    for (int s = 0; s < dirt::kStageCount; ++s) { g_stage_ms[s] = 0; g_stage_calls[s] = 0; g_timer.used[s] = 0; }
    return DIRT_STATUS_OK;
}

const char *dirt_profile_stage_name(int stage)
{
    return (stage >= 0 && stage < dirt::kStageCount) ? kStageNames[stage] : "";
}

}  // extern "C"
