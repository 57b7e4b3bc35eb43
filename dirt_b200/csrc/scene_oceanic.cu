// scene_oceanic.cu -- the procedural fragment stage the fork's `Rasterise` op runs on every
// covered fragment (shaders::oceanic_horizon, csrc/shaders.cpp:1668-1919; uniforms
// csrc/rasterise_egl.cpp:399-433), as a CUDA per-pixel kernel behind the new rasteriser.
//
// The rasteriser (forward.cu) decides coverage and delivers the perspective-correct weights; this
// kernel rebuilds the `smooth` varying texCoordV (= clip-space xy of the face's corners,
// csrc/shaders.cpp:16-31) from the G-buffer and evaluates the shader: a 20-step march of the view
// ray against a sum-of-sines ocean, output .x = sky/sea mask, .y = sun term (sky) or mirrored
// sun glint (sea), .z = 0.  camera_pos stays on the device: the reference's blocking D2H copy
// (csrc/rasterise_egl.cpp:323) has no counterpart.  Dead shader code (godrays, the zero-weighted
// small-wave octaves) is listed in DESIGN.md and not evaluated.
#include "forward_common.cuh"
#include "kernels.h"

namespace dirt {

struct SceneParams {
    int B, H, W, C, V, F;
    const float *background;  int64_t background_bs;
    const float *vertices;    int64_t vertices_bs;
    const int32_t *faces;     int64_t faces_bs;
    const int32_t *face_id;
    const float *gbuffer;
    const float *camera_pos;  // device, >= 8 floats
    float *pixels;
};

__device__ __forceinline__ float ocean_height(float px, float pz, float s2x)
{
    // water(), csrc/shaders.cpp:1762-1793, with small_waveheight = 0 (:1692)
    float wave = sinf(px * 0.021f + s2x) * 4.5f;
    wave += sinf(px * 0.0172f + pz * 0.010f + s2x * 1.121f) * 4.0f;
    wave -= sinf(px * 0.00104f + pz * 0.005f + s2x * 0.121f) * 4.0f;
    wave += sinf(px * 0.02221f + pz * 0.01233f + s2x * 3.437f) * 5.0f;
    wave += sinf(px * 0.03112f + pz * 0.01122f + s2x * 4.269f) * 2.5f;
    return 70.0f + wave * 1.0f;
}

__global__ void __launch_bounds__(256) k_oceanic_horizon(SceneParams p)
{
    const int npix = p.H * p.W;
    const int o = (int)(blockIdx.x * blockDim.x + threadIdx.x);
    if (o >= npix) return;
    // per-launch constants (uniforms)
    const float *cam = p.camera_pos;
    const float cx = __ldg(cam), cy = __ldg(cam + 1), cz = __ldg(cam + 2);
    float s1, c1, s2, c2, s3, c3;
    s1 = sinf(__ldg(cam + 3)); c1 = cosf(__ldg(cam + 3));
    s2 = sinf(__ldg(cam + 4)); c2 = cosf(__ldg(cam + 4));
    s3 = sinf(__ldg(cam + 5)); c3 = cosf(__ldg(cam + 5));
    const float time = __ldg(cam + 6), lz = __ldg(cam + 7);
    const float ll = sqrtf(0.1f * 0.1f + 0.25f * 0.25f + lz * lz);
    const float lx = 0.1f / ll, ly = 0.25f / ll, lzz = lz / ll;
    const float s2x = 0.001f * (time * 190.0f * 2.0f);
    const float width = (float)p.W, height = (float)p.H;
    for (int b = blockIdx.y; b < p.B; b += gridDim.y) {
        const size_t i = (size_t)b * npix + o;
        const int f = __ldg(p.face_id + i);
        if (f < 0) continue;                         // background shows through (resolve wrote it)
        // texCoordV: the smooth varying, same arithmetic as the colour interpolation in k_resolve
        const float4 gb = __ldg(reinterpret_cast<const float4 *>(p.gbuffer) + i);
        const int32_t *fp = p.faces + (int64_t)b * p.faces_bs + 3 * (int64_t)f;
        const float4 *vp = vertex_ptr(p.vertices, p.vertices_bs, b);
        const float4 v0 = __ldg(vp + __ldg(fp)), v1 = __ldg(vp + __ldg(fp + 1)), v2 = __ldg(vp + __ldg(fp + 2));
        const float tx = fadd(fadd(v2.x, fmul(gb.x, fsub(v0.x, v2.x))), fmul(gb.y, fsub(v1.x, v2.x)));
        const float ty = fadd(fadd(v2.y, fmul(gb.x, fsub(v0.y, v2.y))), fmul(gb.y, fsub(v1.y, v2.y)));
        const float *bg = p.background + (int64_t)b * p.background_bs + (size_t)o * p.C;
        const float nx = __ldg(bg), ny = p.C > 1 ? __ldg(bg + 1) : nx;
        const float x = tx + nx / width, y = ty + ny / height;
        // view ray (:1875-1894)
        float rx = (x + 1.0f) * width / 2.0f - width / 2.0f, ry = (y + 1.0f) * height / 2.0f - height / 2.0f, rz = 1.73f * width / 2.0f;
        const float rl = sqrtf(rx * rx + ry * ry + rz * rz);
        rx /= rl; ry /= rl; rz /= rl;
        const float dx = c2 * c3 * rx + (-c1 * s3 + s1 * s2 * c3) * ry + (s1 * s3 + c1 * s2 * c3) * rz;
        const float dy = c2 * s3 * rx + (c1 * c3 + s1 * s2 * s3) * ry + (-s1 * c3 + c1 * s2 * s3) * rz;
        const float dz = -s2 * rx + s1 * c2 * ry + c1 * c2 * rz;
        float sundot = fminf(fmaxf(dx * lx + dy * ly + dz * lzz, 0.0f), 1.0f);
        // trace (:1816-1856)
        float t = -cy / dy, st = 0.5f, old_h = 0.0f;
        for (int j = 0; j < 20; ++j) {
            if (t > 500.0f) st = 1.0f;
            if (t > 800.0f) st = 2.0f;
            if (t > 1500.0f) st = 3.0f;
            const float h = (cy + t * dy) - ocean_height(cx + t * dx, cz + t * dz, s2x);
            const float sg = h > 0.0f ? 1.0f : (h < 0.0f ? -1.0f : 0.0f);
            t += fmaxf(1.0f, fabsf(h)) * sg * st;
            if (old_h * h < 0.0f) st /= 2.0f;
            old_h = h;
        }
        float o0, o1;
        if (dy > 0.0f) {                             // sky
            o0 = 1.0f;
            o1 = powf(sundot, 350.0f);
        } else {                                     // sea: mirror the ray at the wave normal
            o0 = 0.0f;
            const float wx = cx + t * dx, wz = cz + t * dz, e = 0.1f * 1.0f * 4.0f;
            float nxv = ocean_height(wx - e, wz, s2x) - ocean_height(wx + e, wz, s2x), nyv = 1.0f;
            float nzv = ocean_height(wx, wz - e, s2x) - ocean_height(wx, wz + e, s2x);
            const float nl = sqrtf(nxv * nxv + nyv * nyv + nzv * nzv);
            nxv /= nl; nyv /= nl; nzv /= nl;
            const float k = 2.0f * (nxv * dx + nyv * dy + nzv * dz);
            sundot = fminf(fmaxf((dx - k * nxv) * lx + (dy - k * nyv) * ly + (dz - k * nzv) * lzz, 0.0f), 1.0f);
            o1 = 0.5f * powf(sundot, 10.0f) + 0.25f * powf(sundot, 3.5f) + 0.75f * powf(sundot, 300.0f);
        }
        float *out = p.pixels + i * p.C;
        out[0] = o0;
        if (p.C > 1) out[1] = o1;
        if (p.C > 2) out[2] = 0.0f;
        for (int c = 3; c < p.C; ++c) out[c] = 1.0f;
    }
}

cudaError_t launch_oceanic_horizon(const SceneParams &p, cudaStream_t stream)
{
    if ((long long)p.B * p.H * p.W == 0) return cudaSuccess;
    const dim3 g((unsigned)((p.H * p.W + 255) / 256), (unsigned)(p.B < 65535 ? p.B : 65535));
    k_oceanic_horizon<<<g, 256, 0, stream>>>(p);
    note_launches(1);
    return cudaGetLastError();
}

}  // namespace dirt

// ---- C ABI ---------------------------------------------------------------------------------
#include "../../include/dirt_b200.h"

extern "C" int dirt_scene_oceanic_horizon(const DirtSceneArgs *a, void *stream)
{
    if (!a) return DIRT_STATUS_INVALID_ARGUMENT;
    const DirtShape &s = a->shape;
    if (s.height <= 0 || s.width <= 0 || s.channels < 1 || s.channels > DIRT_MAX_CHANNELS || s.batch < 0) return DIRT_STATUS_INVALID_ARGUMENT;
    if (s.batch == 0) return DIRT_STATUS_OK;
    if (!a->background || !a->vertices || !a->faces || !a->face_id || !a->gbuffer || !a->camera_pos || !a->pixels)
        return DIRT_STATUS_INVALID_ARGUMENT;
    dirt::SceneParams p;
    p.B = s.batch; p.H = s.height; p.W = s.width; p.C = s.channels; p.V = s.vertex_count; p.F = s.face_count;
    p.background = a->background; p.background_bs = a->background_bstride;
    p.vertices = a->vertices; p.vertices_bs = a->vertices_bstride;
    p.faces = a->faces; p.faces_bs = a->faces_bstride;
    p.face_id = a->face_id; p.gbuffer = a->gbuffer; p.camera_pos = a->camera_pos; p.pixels = a->pixels;
    return dirt::launch_oceanic_horizon(p, static_cast<cudaStream_t>(stream)) == cudaSuccess ? DIRT_STATUS_OK : DIRT_STATUS_CUDA_ERROR;
}
