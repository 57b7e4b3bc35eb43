// lighting.cu -- Lambertian shading under one directional light as ONE pass forward and ONE backward,
// replacing the framework ops of dirt/lighting.py:182-225 (matmul with the light direction, abs / relu,
// two broadcast multiplies, and their autodiff).  Measured on the c5 optimisation step the torch
// composites of this formula took a third of the step (profiles/r2_c5_step_breakdown.json); per
// vertex the work is a dot product and C multiplies, so both kernels are HBM-bound element-wise passes:
//   forward   out[v, c] = light_color[c] * colours[v, c] * f(n_v . (-d)),  f = |.| (double sided) or max(., 0)
//   backward  grad_colours[v, c] = g[v, c] * light_color[c] * f(cos)
//             grad_normals[v]   = (sum_c g[v, c] * light_color[c] * colours[v, c]) * f'(cos) * (-d)
// `rows` = batch * vertices; light_direction / light_color are DEVICE pointers with a row stride of
// `*_bstride` elements per batch element (0: one light for the whole batch).
#include <cstdint>
#include <cuda_runtime.h>
#include "../../include/dirt_b200.h"

namespace dirt {

namespace {

constexpr int kMaxShadeC = 16;

template <bool BACKWARD>
__global__ void __launch_bounds__(256) k_diffuse_directional(const float *normals, const float *colours, const float *direction,
                                                             int64_t direction_bs, const float *light_color, int64_t light_color_bs,
                                                             int double_sided, long long rows, int V, int C, float *out,
                                                             const float *grad_out, float *grad_normals, float *grad_colours)
{
    for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < rows; r += (long long)gridDim.x * blockDim.x) {
        const long long b = r / V;
        const float *d = direction + b * direction_bs, *lc = light_color + b * light_color_bs;
        const float dx = -__ldg(d), dy = -__ldg(d + 1), dz = -__ldg(d + 2);
        const float *n = normals + 3 * r;
        const float cosine = __ldg(n) * dx + __ldg(n + 1) * dy + __ldg(n + 2) * dz;
        const float f = double_sided ? fabsf(cosine) : fmaxf(cosine, 0.0f);
        const float *c = colours + (size_t)r * C;
        if (!BACKWARD) {
            float *o = out + (size_t)r * C;
            for (int k = 0; k < C; ++k) o[k] = __ldg(lc + k) * __ldg(c + k) * f;
        } else {
            const float *g = grad_out + (size_t)r * C;
            float s = 0.0f;
            for (int k = 0; k < C; ++k) {
                const float gl = __ldg(g + k) * __ldg(lc + k);
                if (grad_colours) grad_colours[(size_t)r * C + k] = gl * f;
                s += gl * __ldg(c + k);
            }
            if (grad_normals) {
                // d|x|/dx = sign(x) (0 at 0, as torch.abs), d max(x, 0)/dx = [x > 0]
                const float df = double_sided ? (cosine > 0.0f ? 1.0f : (cosine < 0.0f ? -1.0f : 0.0f)) : (cosine > 0.0f ? 1.0f : 0.0f);
                const float t = s * df;
                grad_normals[3 * r] = t * dx; grad_normals[3 * r + 1] = t * dy; grad_normals[3 * r + 2] = t * dz;
            }
        }
    }
}

int grid_for_rows(long long rows)
{
    long long g = (rows + 255) / 256;
    if (g > 148 * 16) g = 148 * 16;
    return (int)(g < 1 ? 1 : g);
}

}  // namespace

int set_error(int status, const char *what, cudaError_t e);      // capi.cu
void note_launches(uint64_t n);

}  // namespace dirt

extern "C" {

int dirt_diffuse_directional(const float *normals, const float *colours, const float *light_direction, int64_t light_direction_bstride,
                             const float *light_color, int64_t light_color_bstride, int double_sided, int batch, int vertex_count,
                             int channels, float *out, void *stream)
{
    if (batch < 0 || vertex_count < 0 || channels < 1 || channels > dirt::kMaxShadeC)
        return dirt::set_error(DIRT_STATUS_INVALID_ARGUMENT, "diffuse_directional: bad shape", cudaSuccess);
    const long long rows = (long long)batch * vertex_count;
    if (rows == 0) return DIRT_STATUS_OK;
    if (!normals || !colours || !light_direction || !light_color || !out)
        return dirt::set_error(DIRT_STATUS_INVALID_ARGUMENT, "diffuse_directional: null pointer", cudaSuccess);
    dirt::k_diffuse_directional<false><<<dirt::grid_for_rows(rows), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        normals, colours, light_direction, light_direction_bstride, light_color, light_color_bstride, double_sided, rows, vertex_count,
        channels, out, nullptr, nullptr, nullptr);
    dirt::note_launches(1);
    const cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? DIRT_STATUS_OK : dirt::set_error(DIRT_STATUS_CUDA_ERROR, "diffuse_directional launch", e);
}

int dirt_diffuse_directional_grad(const float *normals, const float *colours, const float *light_direction, int64_t light_direction_bstride,
                                  const float *light_color, int64_t light_color_bstride, int double_sided, int batch, int vertex_count,
                                  int channels, const float *grad_out, float *grad_normals, float *grad_colours, void *stream)
{
    if (batch < 0 || vertex_count < 0 || channels < 1 || channels > dirt::kMaxShadeC)
        return dirt::set_error(DIRT_STATUS_INVALID_ARGUMENT, "diffuse_directional_grad: bad shape", cudaSuccess);
    const long long rows = (long long)batch * vertex_count;
    if (rows == 0 || (!grad_normals && !grad_colours)) return DIRT_STATUS_OK;
    if (!normals || !colours || !light_direction || !light_color || !grad_out)
        return dirt::set_error(DIRT_STATUS_INVALID_ARGUMENT, "diffuse_directional_grad: null pointer", cudaSuccess);
    dirt::k_diffuse_directional<true><<<dirt::grid_for_rows(rows), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        normals, colours, light_direction, light_direction_bstride, light_color, light_color_bstride, double_sided, rows, vertex_count,
        channels, nullptr, grad_out, grad_normals, grad_colours);
    dirt::note_launches(1);
    const cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? DIRT_STATUS_OK : dirt::set_error(DIRT_STATUS_CUDA_ERROR, "diffuse_directional_grad launch", e);
}

}  // extern "C"
