// dilate.cu -- silhouette dilation of a deferred G-buffer channel set, forward and backward.
//
// samples/deferred.py:87-91 (reference) dilates the world-position and normal G-buffers at the
// silhouette so that pixels just outside it see finite values:
//     mask = (x == -inf);  out = x * (1 - mask) + max_pool3x3_SAME(x) * mask
// There it costs a framework max-pool, two multiplies and an add per channel set on top of three
// rasterise calls; here it is one pass per direction, per channel, fused with the masking.
// Backward: a foreground pixel passes its gradient through; a background pixel sends it to the
// FIRST maximal element (row-major) of its 3x3 window, nothing if the whole window is -inf.
// Written as a gather (each pixel collects from the <= 9 background pixels whose window it wins),
// so there are no atomics and the result is deterministic.
#include <cmath>
#include <cstdint>
#include <cuda_runtime.h>
#include "kernels.h"

namespace dirt {

__device__ __forceinline__ bool is_neg_inf(float v) { return v == -INFINITY; }

// Offset (dy * 3 + dx, 0..8) of the first maximal in-image element of the 3x3 window centred
// on (r, x); -1 if every element is -inf (or NaN).
__device__ __forceinline__ int window_argmax(const float *img, int H, int W, int C, int c, int r, int x, float &best)
{
    best = -INFINITY;
    int arg = -1;
#pragma unroll
    for (int k = 0; k < 9; ++k) {
        const int rr = r + k / 3 - 1, xx = x + k % 3 - 1;
        if (rr < 0 || rr >= H || xx < 0 || xx >= W) continue;      // SAME padding = -inf
        const float v = __ldg(img + ((size_t)rr * W + xx) * C + c);
        if (v > best) { best = v; arg = k; }
    }
    return arg;
}

__global__ void __launch_bounds__(256) k_dilate(const float *x, float *out, int B, int H, int W, int C)
{
    const size_t total = (size_t)B * H * W * C;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const float v = __ldg(x + i);
        float o = v;
        if (is_neg_inf(v)) {
            const int c = (int)(i % C);
            const size_t pix = i / C;
            const int xx = (int)(pix % W), r = (int)((pix / W) % H);
            const size_t b = pix / ((size_t)W * H);
            window_argmax(x + b * H * W * C, H, W, C, c, r, xx, o);
        }
        out[i] = o;
    }
}

__global__ void __launch_bounds__(256) k_dilate_grad(const float *x, const float *grad_out, float *grad_x, int B, int H, int W, int C)
{
    const size_t total = (size_t)B * H * W * C;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const int c = (int)(i % C);
        const size_t pix = i / C;
        const int xx = (int)(pix % W), r = (int)((pix / W) % H);
        const size_t b = pix / ((size_t)W * H);
        const float *img = x + b * H * W * C;
        const float *go = grad_out + b * H * W * C;
        const float v = __ldg(x + i);
        float g = 0.0f;
        if (!is_neg_inf(v)) {
            g = __ldg(grad_out + i);
            // background neighbours q whose window's first maximum is this pixel
            for (int k = 0; k < 9; ++k) {
                const int qr = r + k / 3 - 1, qx = xx + k % 3 - 1;
                if (k == 4 || qr < 0 || qr >= H || qx < 0 || qx >= W) continue;
                if (!is_neg_inf(__ldg(img + ((size_t)qr * W + qx) * C + c))) continue;
                float best;
                const int arg = window_argmax(img, H, W, C, c, qr, qx, best);
                // this pixel sits at offset (r - qr + 1) * 3 + (xx - qx + 1) of q's window
                if (arg == (r - qr + 1) * 3 + (xx - qx + 1)) g += __ldg(go + ((size_t)qr * W + qx) * C + c);
            }
        }
        grad_x[i] = g;
    }
}

static inline int dilate_grid(size_t total, int sm_count)
{
    size_t g = (total + 255) / 256;
    const size_t cap = (size_t)sm_count * 32;
    if (g > cap) g = cap;
    return g < 1 ? 1 : (int)g;
}

cudaError_t launch_dilate(const float *x, float *out, int B, int H, int W, int C, int sm_count, cudaStream_t stream)
{
    const size_t total = (size_t)B * H * W * C;
    if (total == 0) return cudaSuccess;
    k_dilate<<<dilate_grid(total, sm_count), 256, 0, stream>>>(x, out, B, H, W, C);
    return cudaGetLastError();
}

cudaError_t launch_dilate_grad(const float *x, const float *grad_out, float *grad_x, int B, int H, int W, int C,
                               int sm_count, cudaStream_t stream)
{
    const size_t total = (size_t)B * H * W * C;
    if (total == 0) return cudaSuccess;
    k_dilate_grad<<<dilate_grid(total, sm_count), 256, 0, stream>>>(x, grad_out, grad_x, B, H, W, C);
    return cudaGetLastError();
}

}  // namespace dirt
