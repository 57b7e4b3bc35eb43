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
#include <cstdio>
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

// One point light (dirt/lighting.py:291-343) and the Phong term of one directional light (:228-288), forward
// (BACKWARD = false) or backward in one pass per vertex.  The formulas, including where the reference adds its 1e-12:
//   diffuse_point          inc = (p - light) / (|p - light| + 1e-12);  out = lc * colour * f(n . inc)
//   specular_directional   L = -d;  R = -L + 2 (n . L) n;  T = (cam - p) / |cam - p| + 1e-12 (per component);
//                          out = lc * reflectivity * f(T . R) ^ shininess
template <bool SPECULAR, bool BACKWARD>
__global__ void __launch_bounds__(256) k_vertex_shade(DirtVertexShadeArgs a, long long rows)
{
    const int V = a.vertex_count, C = a.channels;
    for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < rows; r += (long long)gridDim.x * blockDim.x) {
        const long long b = r / V;
        const float *lv = a.light_vector + b * a.light_vector_bstride, *lc = a.light_color + b * a.light_color_bstride;
        const float px = __ldg(a.positions + 3 * r), py = __ldg(a.positions + 3 * r + 1), pz = __ldg(a.positions + 3 * r + 2);
        const float nx = __ldg(a.normals + 3 * r), ny = __ldg(a.normals + 3 * r + 1), nz = __ldg(a.normals + 3 * r + 2);
        const float *c = a.colours + (size_t)r * C;
        float cosine, weight, dweight;            // weight = f(cos) [^ shininess]; dweight = d weight / d cos
        // quantities the backward pass needs again
        float ix = 0.f, iy = 0.f, iz = 0.f, len = 0.f;          // diffuse_point: rel / (len + eps); specular: T / |T|
        float Rx = 0.f, Ry = 0.f, Rz = 0.f, Lx = 0.f, Ly = 0.f, Lz = 0.f, ndl = 0.f;
        if (!SPECULAR) {
            const float rx = px - __ldg(lv), ry = py - __ldg(lv + 1), rz = pz - __ldg(lv + 2);
            len = sqrtf(rx * rx + ry * ry + rz * rz);
            const float inv = 1.0f / (len + 1.e-12f);
            ix = rx * inv; iy = ry * inv; iz = rz * inv;
            cosine = nx * ix + ny * iy + nz * iz;
            weight = a.double_sided ? fabsf(cosine) : fmaxf(cosine, 0.0f);
            dweight = a.double_sided ? (cosine > 0.0f ? 1.0f : (cosine < 0.0f ? -1.0f : 0.0f)) : (cosine > 0.0f ? 1.0f : 0.0f);
        } else {
            Lx = -__ldg(lv); Ly = -__ldg(lv + 1); Lz = -__ldg(lv + 2);
            ndl = nx * Lx + ny * Ly + nz * Lz;
            Rx = -Lx + 2.0f * ndl * nx; Ry = -Ly + 2.0f * ndl * ny; Rz = -Lz + 2.0f * ndl * nz;
            const float *cam = a.camera_position + b * a.camera_position_bstride;
            const float tx = __ldg(cam) - px, ty = __ldg(cam + 1) - py, tz = __ldg(cam + 2) - pz;
            len = sqrtf(tx * tx + ty * ty + tz * tz);
            ix = tx / len; iy = ty / len; iz = tz / len;
            cosine = (ix + 1.e-12f) * Rx + (iy + 1.e-12f) * Ry + (iz + 1.e-12f) * Rz;
            const float base = a.double_sided ? fabsf(cosine) : fmaxf(cosine, 0.0f);
            const float dbase = a.double_sided ? (cosine > 0.0f ? 1.0f : (cosine < 0.0f ? -1.0f : 0.0f)) : (cosine > 0.0f ? 1.0f : 0.0f);
            const float sh = __ldg(a.shininess + b * a.shininess_bstride);
            weight = powf(base, sh);
            // d base^sh / d base as torch.pow differentiates it: sh * base^(sh - 1), and 0 for sh == 0
            dweight = sh == 0.0f ? 0.0f : sh * powf(base, sh - 1.0f) * dbase;
        }
        if (!BACKWARD) {
            float *o = a.out + (size_t)r * C;
            for (int k = 0; k < C; ++k) o[k] = __ldg(lc + k) * __ldg(c + k) * weight;
            continue;
        }
        const float *g = a.grad_out + (size_t)r * C;
        float s = 0.0f;
        for (int k = 0; k < C; ++k) {
            const float gl = __ldg(g + k) * __ldg(lc + k);
            if (a.grad_colours) a.grad_colours[(size_t)r * C + k] = gl * weight;
            s += gl * __ldg(c + k);
        }
        const float dcos = s * dweight;
        if (!SPECULAR) {
            if (a.grad_normals) { a.grad_normals[3 * r] = dcos * ix; a.grad_normals[3 * r + 1] = dcos * iy; a.grad_normals[3 * r + 2] = dcos * iz; }
            if (a.grad_positions) {
                // inc = rel / (len + eps):  d rel = d inc / (len + eps) - inc (d inc . inc) / len   (second term 0 at len = 0)
                const float inv = 1.0f / (len + 1.e-12f);
                const float dix = dcos * nx, diy = dcos * ny, diz = dcos * nz;
                const float along = len > 0.0f ? (dix * ix + diy * iy + diz * iz) / len : 0.0f;
                a.grad_positions[3 * r] = dix * inv - along * ix;
                a.grad_positions[3 * r + 1] = diy * inv - along * iy;
                a.grad_positions[3 * r + 2] = diz * inv - along * iz;
            }
        } else {
            if (a.grad_normals) {
                // R = -L + 2 (n . L) n with dR = dcos * (T + eps)
                const float dRx = dcos * (ix + 1.e-12f), dRy = dcos * (iy + 1.e-12f), dRz = dcos * (iz + 1.e-12f);
                const float dRn = dRx * nx + dRy * ny + dRz * nz;
                a.grad_normals[3 * r] = 2.0f * (Lx * dRn + ndl * dRx);
                a.grad_normals[3 * r + 1] = 2.0f * (Ly * dRn + ndl * dRy);
                a.grad_normals[3 * r + 2] = 2.0f * (Lz * dRn + ndl * dRz);
            }
            if (a.grad_positions) {
                // T = t / |t| (+ eps), t = cam - p: dt = (dT - That (dT . That)) / |t|, dp = -dt
                const float dTx = dcos * Rx, dTy = dcos * Ry, dTz = dcos * Rz;
                const float along = dTx * ix + dTy * iy + dTz * iz;
                a.grad_positions[3 * r] = -(dTx - ix * along) / len;
                a.grad_positions[3 * r + 1] = -(dTy - iy * along) / len;
                a.grad_positions[3 * r + 2] = -(dTz - iz * along) / len;
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

namespace dirt {
namespace {
template <bool SPECULAR, bool BACKWARD>
int vertex_shade(const DirtVertexShadeArgs *a, void *stream, const char *name)
{
    char msg[96];
    if (!a || a->batch < 0 || a->vertex_count < 0 || a->channels < 1 || a->channels > kMaxShadeC) {
        snprintf(msg, sizeof(msg), "%s: bad shape", name);
        return set_error(DIRT_STATUS_INVALID_ARGUMENT, msg, cudaSuccess);
    }
    const long long rows = (long long)a->batch * a->vertex_count;
    if (rows == 0) return DIRT_STATUS_OK;
    if (BACKWARD && !a->grad_positions && !a->grad_normals && !a->grad_colours) return DIRT_STATUS_OK;
    const bool ok = a->positions && a->normals && a->colours && a->light_vector && a->light_color &&
                    (!SPECULAR || (a->camera_position && a->shininess)) && (BACKWARD ? a->grad_out != nullptr : a->out != nullptr);
    if (!ok) {
        snprintf(msg, sizeof(msg), "%s: null pointer", name);
        return set_error(DIRT_STATUS_INVALID_ARGUMENT, msg, cudaSuccess);
    }
    k_vertex_shade<SPECULAR, BACKWARD><<<grid_for_rows(rows), 256, 0, static_cast<cudaStream_t>(stream)>>>(*a, rows);
    note_launches(1);
    const cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) return DIRT_STATUS_OK;
    snprintf(msg, sizeof(msg), "%s launch", name);
    return set_error(DIRT_STATUS_CUDA_ERROR, msg, e);
}
}  // namespace
}  // namespace dirt

extern "C" {

int dirt_diffuse_point(const DirtVertexShadeArgs *a, void *stream) { return dirt::vertex_shade<false, false>(a, stream, "diffuse_point"); }
int dirt_diffuse_point_grad(const DirtVertexShadeArgs *a, void *stream) { return dirt::vertex_shade<false, true>(a, stream, "diffuse_point_grad"); }
int dirt_specular_directional(const DirtVertexShadeArgs *a, void *stream) { return dirt::vertex_shade<true, false>(a, stream, "specular_directional"); }
int dirt_specular_directional_grad(const DirtVertexShadeArgs *a, void *stream) { return dirt::vertex_shade<true, true>(a, stream, "specular_directional_grad"); }


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
