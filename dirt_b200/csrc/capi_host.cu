// capi_host.cu -- host-buffer variants of the two entry points (include/dirt_b200.h):
// stage inputs to the current device, run the device entry points, copy the outputs back.
// The staging buffers are library-owned and grow-only per device, the role RegisteredBuffer
// plays in the reference (csrc/rasterise_egl.cpp:55-111).
#include <cstring>
#include <mutex>
#include "../../include/dirt_b200.h"
#include <cuda_runtime.h>

namespace dirt { int set_error(int status, const char *what, cudaError_t e); }    // capi.cu

namespace {

struct Arena {
    char *base = nullptr;
    size_t cap = 0, used = 0;
};
std::mutex g_mu;
Arena g_arena[64];

// All staging for one call is carved from one grow-only device allocation.
cudaError_t arena_reserve(Arena &a, size_t bytes)
{
    a.used = 0;
    if (bytes <= a.cap) return cudaSuccess;
    if (a.base) { cudaError_t e = cudaFree(a.base); a.base = nullptr; a.cap = 0; if (e != cudaSuccess) return e; }
    cudaError_t e = cudaMalloc(&a.base, bytes);
    if (e == cudaSuccess) a.cap = bytes;
    return e;
}
inline size_t pad(size_t b) { return (b + 255) / 256 * 256; }
template <typename T> T *arena_take(Arena &a, size_t count)
{
    T *p = reinterpret_cast<T *>(a.base + a.used);
    a.used += pad(count * sizeof(T));
    return p;
}
inline size_t extent(int64_t stride, int64_t block, int batch) { return (size_t)(stride == 0 ? block : stride * (int64_t)(batch - 1) + block); }

#define HOST_TRY(expr) do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) { cudaStreamDestroy(st); return dirt::set_error(DIRT_STATUS_CUDA_ERROR, "host staging copy", e_); } } while (0)

// the current device, or a recorded error: a process without a usable GPU is told so in words
int current_device(int *dev)
{
    const cudaError_t e = cudaGetDevice(dev);
    if (e != cudaSuccess) return dirt::set_error(DIRT_STATUS_CUDA_ERROR, "no usable CUDA device (there is no CPU path)", e);
    if (*dev < 0 || *dev >= 64) return dirt::set_error(DIRT_STATUS_CUDA_ERROR, "device ordinal out of range", cudaSuccess);
    return DIRT_STATUS_OK;
}

}  // namespace

extern "C" {

int dirt_rasterise_forward_host(const DirtForwardArgs *h)
{
    if (!h) return dirt::set_error(DIRT_STATUS_INVALID_ARGUMENT, "null arguments", cudaSuccess);
    const DirtShape &s = h->shape;
    const size_t ws_bytes = dirt_workspace_bytes(&s);
    if (ws_bytes == 0) return dirt::set_error(DIRT_STATUS_INVALID_ARGUMENT, "invalid shape", cudaSuccess);
    if (s.batch == 0) return DIRT_STATUS_OK;
    int dev = 0;
    if (int rc_ = current_device(&dev)) return rc_;
    std::lock_guard<std::mutex> lock(g_mu);
    Arena &A = g_arena[dev];
    const size_t npix = (size_t)s.batch * s.height * s.width;
    const size_t n_bg = extent(h->background_bstride, (int64_t)s.height * s.width * s.channels, s.batch);
    const size_t n_v = extent(h->vertices_bstride, (int64_t)s.vertex_count * 4, s.batch);
    const size_t n_c = extent(h->vertex_colors_bstride, (int64_t)s.vertex_count * s.channels, s.batch);
    const size_t n_f = extent(h->faces_bstride, (int64_t)s.face_count * 3, s.batch);
    const bool aux = h->face_id != nullptr;
    size_t need = pad(ws_bytes) + pad(n_bg * 4) + pad(n_v * 4) + pad(n_c * 4) + pad(n_f * 4) + pad(npix * s.channels * 4) +
                  (aux ? pad(npix * 4) + pad(npix * 16) : 0);
    if (cudaError_t e_ = arena_reserve(A, need)) return dirt::set_error(DIRT_STATUS_CUDA_ERROR, "device staging allocation", e_);
    cudaStream_t st;
    if (cudaError_t e_ = cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking)) return dirt::set_error(DIRT_STATUS_CUDA_ERROR, "stream creation", e_);
    DirtForwardArgs d = *h;
    d.workspace = arena_take<char>(A, ws_bytes); d.workspace_bytes = ws_bytes;
    d.visibility_hint = nullptr; d.visibility_hint_bytes = 0;      // a device-side, persistent buffer: not part of the host-buffer calls
    float *bg = arena_take<float>(A, n_bg), *v = arena_take<float>(A, n_v), *c = arena_take<float>(A, n_c);
    int32_t *f = arena_take<int32_t>(A, n_f);
    d.pixels = arena_take<float>(A, npix * s.channels);
    if (aux) { d.face_id = arena_take<int32_t>(A, npix); d.gbuffer = arena_take<float>(A, npix * 4); }
    HOST_TRY(cudaMemcpyAsync(bg, h->background, n_bg * 4, cudaMemcpyHostToDevice, st));
    if (n_v) HOST_TRY(cudaMemcpyAsync(v, h->vertices, n_v * 4, cudaMemcpyHostToDevice, st));
    if (n_c) HOST_TRY(cudaMemcpyAsync(c, h->vertex_colors, n_c * 4, cudaMemcpyHostToDevice, st));
    if (n_f) HOST_TRY(cudaMemcpyAsync(f, h->faces, n_f * 4, cudaMemcpyHostToDevice, st));
    d.background = bg; d.vertices = v; d.vertex_colors = c; d.faces = f;
    int rc = dirt_rasterise_forward(&d, st);
    if (rc != DIRT_STATUS_OK) { cudaStreamDestroy(st); return rc; }
    HOST_TRY(cudaMemcpyAsync(h->pixels, d.pixels, npix * s.channels * 4, cudaMemcpyDeviceToHost, st));
    if (aux) {
        HOST_TRY(cudaMemcpyAsync(h->face_id, d.face_id, npix * 4, cudaMemcpyDeviceToHost, st));
        HOST_TRY(cudaMemcpyAsync(h->gbuffer, d.gbuffer, npix * 16, cudaMemcpyDeviceToHost, st));
    }
    HOST_TRY(cudaStreamSynchronize(st));
    cudaStreamDestroy(st);
    return DIRT_STATUS_OK;
}

int dirt_rasterise_backward_host(const DirtBackwardArgs *h)
{
    if (!h) return dirt::set_error(DIRT_STATUS_INVALID_ARGUMENT, "null arguments", cudaSuccess);
    const DirtShape &s = h->shape;
    if (s.batch == 0) return DIRT_STATUS_OK;
    if (s.batch < 0 || s.height <= 0 || s.width <= 0 || s.channels <= 0) return dirt::set_error(DIRT_STATUS_INVALID_ARGUMENT, "invalid shape", cudaSuccess);
    int dev = 0;
    if (int rc_ = current_device(&dev)) return rc_;
    std::lock_guard<std::mutex> lock(g_mu);
    Arena &A = g_arena[dev];
    const size_t npix = (size_t)s.batch * s.height * s.width, nimg = npix * s.channels;
    const size_t n_f = extent(h->faces_bstride, (int64_t)s.face_count * 3, s.batch);
    const size_t n_gv = extent(h->grad_vertices_bstride, (int64_t)s.vertex_count * 4, s.batch);
    const size_t n_gc = extent(h->grad_vertex_colors_bstride, (int64_t)s.vertex_count * s.channels, s.batch);
    const bool stateless = h->face_id == nullptr && h->gbuffer == nullptr;      // visibility re-derived from vertices
    const size_t n_v = stateless ? extent(h->vertices_bstride, (int64_t)s.vertex_count * 4, s.batch) : 0;
    const size_t ws_bytes = stateless ? dirt_backward_workspace_bytes(&s) : 0;
    if (stateless && (!h->vertices || ws_bytes == 0)) return dirt::set_error(DIRT_STATUS_INVALID_ARGUMENT, "backward without face_id / gbuffer needs vertices", cudaSuccess);
    size_t need = pad(n_f * 4) + 3 * pad(nimg * 4) + pad(npix * 4) + pad(npix * 16) + pad(n_gv * 4) + pad(n_gc * 4) + pad(n_v * 4) + pad(ws_bytes);
    if (cudaError_t e_ = arena_reserve(A, need)) return dirt::set_error(DIRT_STATUS_CUDA_ERROR, "device staging allocation", e_);
    cudaStream_t st;
    if (cudaError_t e_ = cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking)) return dirt::set_error(DIRT_STATUS_CUDA_ERROR, "stream creation", e_);
    DirtBackwardArgs d = *h;
    int32_t *f = arena_take<int32_t>(A, n_f);
    float *px = arena_take<float>(A, nimg), *g = arena_take<float>(A, nimg), *gbg = arena_take<float>(A, nimg);
    int32_t *fid = arena_take<int32_t>(A, npix);
    float *gb = arena_take<float>(A, npix * 4);
    float *gv = arena_take<float>(A, n_gv), *gc = arena_take<float>(A, n_gc);
    if (n_f) HOST_TRY(cudaMemcpyAsync(f, h->faces, n_f * 4, cudaMemcpyHostToDevice, st));
    HOST_TRY(cudaMemcpyAsync(px, h->pixels, nimg * 4, cudaMemcpyHostToDevice, st));
    HOST_TRY(cudaMemcpyAsync(g, h->grad_pixels, nimg * 4, cudaMemcpyHostToDevice, st));
    d.faces = f; d.pixels = px; d.grad_pixels = g; d.face_id = fid; d.gbuffer = gb;
    d.vertices = nullptr; d.workspace = nullptr; d.workspace_bytes = 0;
    d.flags = h->flags & ~DIRT_BACKWARD_DETERMINISTIC;      // the host-buffer calls offer the default summation only
    if (stateless) {
        float *v = arena_take<float>(A, n_v);
        d.workspace = arena_take<char>(A, ws_bytes); d.workspace_bytes = ws_bytes;
        if (n_v) HOST_TRY(cudaMemcpyAsync(v, h->vertices, n_v * 4, cudaMemcpyHostToDevice, st));
        d.vertices = v; d.face_id = nullptr; d.gbuffer = nullptr;
    } else {
        if (!h->face_id || !h->gbuffer) { cudaStreamDestroy(st); return dirt::set_error(DIRT_STATUS_INVALID_ARGUMENT, "face_id and gbuffer must be both given or both null", cudaSuccess); }
        HOST_TRY(cudaMemcpyAsync(fid, h->face_id, npix * 4, cudaMemcpyHostToDevice, st));
        HOST_TRY(cudaMemcpyAsync(gb, h->gbuffer, npix * 16, cudaMemcpyHostToDevice, st));
    }
    d.grad_vertices = gv; d.grad_vertex_colors = gc; d.grad_background = h->grad_background ? gbg : nullptr;
    int rc = dirt_rasterise_backward(&d, st);
    if (rc != DIRT_STATUS_OK) { cudaStreamDestroy(st); return rc; }
    if (n_gv) HOST_TRY(cudaMemcpyAsync(h->grad_vertices, gv, n_gv * 4, cudaMemcpyDeviceToHost, st));
    if (n_gc) HOST_TRY(cudaMemcpyAsync(h->grad_vertex_colors, gc, n_gc * 4, cudaMemcpyDeviceToHost, st));
    if (h->grad_background) HOST_TRY(cudaMemcpyAsync(h->grad_background, gbg, nimg * 4, cudaMemcpyDeviceToHost, st));
    HOST_TRY(cudaStreamSynchronize(st));
    cudaStreamDestroy(st);
    return DIRT_STATUS_OK;
}

int dirt_release_host_staging(void)
{
    int dev = 0;
    if (int rc_ = current_device(&dev)) return rc_;
    std::lock_guard<std::mutex> lock(g_mu);
    Arena &A = g_arena[dev];
    if (A.base) cudaFree(A.base);
    A = Arena();
    return DIRT_STATUS_OK;
}

}  // extern "C"
