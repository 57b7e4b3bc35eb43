// scene_shaders.cu -- the fork's other procedural fragment stages as CUDA per-pixel kernels behind
// the new rasteriser:
//   DIRT_SCENE_OCEANIC              shaders::oceanic             csrc/shaders.cpp:556-865   (op RasteriseGrad, csrc/rasterise_grad_egl.cpp:399)
//   DIRT_SCENE_OCEANIC_STILL_CLOUD  shaders::oceanic_still_cloud csrc/shaders.cpp:866-1175  (csrc/oceanic_still_cloud.cpp)
//   DIRT_SCENE_OCEANIC_NO_CLOUD     shaders::oceanic_no_cloud    csrc/shaders.cpp:1402-1665 (csrc/oceanic_no_cloud.cpp)
//   DIRT_SCENE_OCEANIC_SIMPLE_PROXY shaders::oceanic_simple_proxy csrc/shaders.cpp:1921-2185 (csrc/oceanic_simple_proxy.cpp)
//   DIRT_SCENE_OCEANIC_OPT_FLOW     shaders::oceanic_opt_flow    csrc/shaders.cpp:1178-1398 (csrc/oceanic_opt_flow.cpp)
//   DIRT_SCENE_HILL                 shaders::hill                csrc/shaders.cpp:123-552   (csrc/hill.cpp)
// (DIRT_SCENE_OCEANIC_HORIZON lives in scene_oceanic.cu.)
//
// As in scene_oceanic.cu the rasteriser decides coverage and hands over the perspective-correct
// weights; each thread rebuilds the `smooth` varying texCoordV for its pixel and evaluates the
// shader.  The five ocean programs differ only in constants and in which blocks exist, so they are
// one template whose traits are compile-time: a variant without clouds carries no cloud code.
// The uniforms stay on the device (no counterpart of the blocking D2H copy, csrc/hill.cpp:323).
// How under-specified GLSL is read (dead god-rays, self-sampled render target, hill's texture
// orientation and filtering) is listed in DESIGN.md.
#include "forward_common.cuh"
#include "kernels.h"
#include "../../include/dirt_b200.h"

namespace dirt {

struct ShadeParams {
    int B, H, W, C, Cb;
    const float *background;  int64_t background_bs;
    const float *vertices;    int64_t vertices_bs;
    const int32_t *faces;     int64_t faces_bs;
    const int32_t *face_id;
    const float *gbuffer;
    const float *camera_pos;
    float *pixels;
};

struct f2 { float x, y; };
struct f3 { float x, y, z; };
struct f4 { float x, y, z, w; };
#define DI __device__ __forceinline__
DI f2 mk2(float x, float y) { f2 r; r.x = x; r.y = y; return r; }
DI f3 mk3(float x, float y, float z) { f3 r; r.x = x; r.y = y; r.z = z; return r; }
DI f2 operator+(f2 a, f2 b) { return mk2(a.x + b.x, a.y + b.y); }
DI f2 operator-(f2 a, f2 b) { return mk2(a.x - b.x, a.y - b.y); }
DI f2 operator*(f2 a, float s) { return mk2(a.x * s, a.y * s); }
DI f3 operator+(f3 a, f3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
DI f3 operator-(f3 a, f3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
DI f3 operator*(f3 a, float s) { return mk3(a.x * s, a.y * s, a.z * s); }
DI f3 operator*(f3 a, f3 b) { return mk3(a.x * b.x, a.y * b.y, a.z * b.z); }
DI float dot2(f2 a, f2 b) { return a.x * b.x + a.y * b.y; }
DI float dot3(f3 a, f3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
// sin / cos correctly rounded to binary32: evaluated in double, rounded once.  The noise hashes scale
// them by ~4e4, which turns a one-ulp disagreement between maths libraries into a visibly different
// wave field; this way the kernels reproduce the CPU restatement bit for bit (tests/test_scene_shaders.py).
DI float rsin(float v) { return (float)sin((double)v); }
DI float rcos(float v) { return (float)cos((double)v); }
DI float fractf(float v) { return v - floorf(v); }
DI float clamp01(float v) { return fminf(fmaxf(v, 0.0f), 1.0f); }
DI float sgn(float v) { return v > 0.0f ? 1.0f : (v < 0.0f ? -1.0f : 0.0f); }
DI float lerp(float a, float b, float t) { return a * (1.0f - t) + b * t; }
DI f3 lerp3(f3 a, f3 b, float t) { return a * (1.0f - t) + b * t; }
DI float hermite(float e0, float e1, float v)
{
    const float t = clamp01((v - e0) / (e1 - e0));
    return t * t * (3.0f - 2.0f * t);
}
DI f3 unit(f3 v)
{
    const float l = sqrtf(dot3(v, v));
    return mk3(v.x / l, v.y / l, v.z / l);
}

// texCoordV of pixel i: the same arithmetic as the colour interpolation of k_resolve
DI f2 texcoord_at(const ShadeParams &p, int b, size_t i, int f)
{
    const float4 gb = __ldg(reinterpret_cast<const float4 *>(p.gbuffer) + i);
    const int32_t *fp = p.faces + (int64_t)b * p.faces_bs + 3 * (int64_t)f;
    const float4 *vp = vertex_ptr(p.vertices, p.vertices_bs, b);
    const float4 v0 = __ldg(vp + __ldg(fp)), v1 = __ldg(vp + __ldg(fp + 1)), v2 = __ldg(vp + __ldg(fp + 2));
    return mk2(fadd(fadd(v2.x, fmul(gb.x, fsub(v0.x, v2.x))), fmul(gb.y, fsub(v1.x, v2.x))),
               fadd(fadd(v2.y, fmul(gb.x, fsub(v0.y, v2.y))), fmul(gb.y, fsub(v1.y, v2.y))));
}

// ---- ocean family ---------------------------------------------------------------------------

enum { kCloudNone = 0, kCloudDrift = 1, kCloudStill = 2 };

template <int S> struct OceanTraits;
struct OceanDefaults {                                    // shaders::oceanic, csrc/shaders.cpp:575-584
    static constexpr float wavegain = 1.0f, large_wave = 1.0f, small_wave = 1.0f;
    static constexpr float fog0 = 0.5f, fog1 = 0.7f, fog2 = 1.1f,
                           skyb0 = 0.6f, skyb1 = 0.8f, skyb2 = 1.2f,
                           skyt0 = 0.05f, skyt1 = 0.2f, skyt2 = 0.5f,
                           refl0 = 0.025f, refl1 = 0.10f, refl2 = 0.20f,
                           wcol0 = 0.2f, wcol1 = 0.25f, wcol2 = 0.3f;
    static constexpr float s1x = 160.0f, s1y = 120.0f, s2x = 190.0f, s2y = 130.0f;
    static constexpr bool cosine = false, flat_sea = false, jitter = true;
    static constexpr int octaves = 7, steps = 20, clouds = kCloudDrift;
};
template <> struct OceanTraits<DIRT_SCENE_OCEANIC> : OceanDefaults {};
template <> struct OceanTraits<DIRT_SCENE_OCEANIC_STILL_CLOUD> : OceanDefaults { static constexpr int clouds = kCloudStill; };
template <> struct OceanTraits<DIRT_SCENE_OCEANIC_NO_CLOUD> : OceanDefaults { static constexpr int clouds = kCloudNone; };
template <> struct OceanTraits<DIRT_SCENE_OCEANIC_SIMPLE_PROXY> : OceanDefaults {   // csrc/shaders.cpp:1940-1949, :2017-2034, :2060
    static constexpr float wavegain = 0.75f, large_wave = 0.75f, small_wave = 1.5f;
    static constexpr float fog0 = 0.4f, fog1 = 0.4f, fog2 = 1.2f,
                           skyb0 = 0.5f, skyb1 = 0.5f, skyb2 = 1.3f,
                           skyt0 = 0.15f, skyt1 = 0.1f, skyt2 = 0.7f,
                           refl0 = 0.1f, refl1 = 0.1f, refl2 = 0.15f,
                           wcol0 = 0.1f, wcol1 = 0.2f, wcol2 = 0.5f;
    static constexpr float s1x = 260.0f, s1y = 100.0f, s2x = 150.0f, s2y = 230.0f;
    static constexpr bool cosine = true;
    static constexpr int octaves = 3, steps = 10, clouds = kCloudNone;
};
template <> struct OceanTraits<DIRT_SCENE_OCEANIC_OPT_FLOW> : OceanDefaults {       // csrc/shaders.cpp:1207-1208, :1280, :1325-1327
    static constexpr float large_wave = 0.0f, small_wave = 0.0f;
    static constexpr bool flat_sea = true, jitter = false;
    static constexpr int clouds = kCloudNone;
};

DI float cos_hash(float n) { return fractf(rcos(n) * 41415.92653f); }
DI float sin_hash(f2 n) { return fractf(rsin(dot2(n, mk2(12.9898f, 4.1414f))) * 43758.5453f); }

DI float value_noise2(f2 p)
{
    const f2 ip = mk2(floorf(p.x), floorf(p.y));
    f2 u = mk2(fractf(p.x), fractf(p.y));
    u = mk2(u.x * u.x * (3.0f - 2.0f * u.x), u.y * u.y * (3.0f - 2.0f * u.y));
    return lerp(lerp(sin_hash(ip), sin_hash(ip + mk2(1.0f, 0.0f)), u.x),
                lerp(sin_hash(ip + mk2(0.0f, 1.0f)), sin_hash(ip + mk2(1.0f, 1.0f)), u.x), u.y);
}

DI float value_noise3(f3 x)
{
    const f3 p = mk3(floorf(x.x), floorf(x.y), floorf(x.z));
    const f3 f = mk3(hermite(0.0f, 1.0f, fractf(x.x)), hermite(0.0f, 1.0f, fractf(x.y)), hermite(0.0f, 1.0f, fractf(x.z)));
    const float n = p.x + p.y * 57.0f + 113.0f * p.z;
    return lerp(lerp(lerp(cos_hash(n + 0.0f), cos_hash(n + 1.0f), f.x), lerp(cos_hash(n + 57.0f), cos_hash(n + 58.0f), f.x), f.y),
                lerp(lerp(cos_hash(n + 113.0f), cos_hash(n + 114.0f), f.x), lerp(cos_hash(n + 170.0f), cos_hash(n + 171.0f), f.x), f.y), f.z);
}

DI f3 cloud_rot(f3 p)      // mat3 m, csrc/shaders.cpp:624 (column major)
{
    return mk3(0.00f * p.x + -1.60f * p.y + -1.20f * p.z, 1.60f * p.x + 0.72f * p.y + -0.96f * p.z,
               1.20f * p.x + -0.96f * p.y + 1.28f * p.z);
}
DI f2 chop_rot(f2 p) { return mk2(1.6f * p.x + 1.2f * p.y, -1.2f * p.x + 1.6f * p.y); }   // mat2 m2, :636

__device__ __noinline__ float cloud_fbm(f3 p)
{
    float f = 0.5000f * value_noise3(p); p = cloud_rot(p) * 1.1f;
    f += 0.2500f * value_noise3(p); p = cloud_rot(p) * 1.2f;
    f += 0.1666f * value_noise3(p); p = cloud_rot(p);
    f += 0.0834f * value_noise3(p);
    return f;
}

DI float chop_fbm(f2 p)
{
    float f = 0.5000f * value_noise2(p); p = chop_rot(p);
    f += 0.2500f * value_noise2(p); p = chop_rot(p);
    f += 0.1666f * value_noise2(p); p = chop_rot(p);
    f += 0.0834f * value_noise2(p);
    return f;
}

struct OceanFrame {           // per-launch uniforms
    float cam[16];
    float width, height;
};

template <int S> __device__ __noinline__ float sea_height(const OceanFrame &u, f2 p)
{
    typedef OceanTraits<S> T;
    if (T::flat_sea) return 70.0f - 12.0f;
    const float time = u.cam[6];
    f2 shift1 = mk2(0.001f * (time * T::s1x * 2.0f), 0.001f * (time * T::s1y * 2.0f));
    const f2 shift2 = mk2(0.001f * (time * T::s2x * 2.0f), 0.001f * (-time * T::s2y * 2.0f));
    const float a0 = p.x * 0.021f + shift2.x, a1 = p.x * 0.0172f + p.y * 0.010f + shift2.x * 1.121f,
                a2 = p.x * 0.00104f + p.y * 0.005f + shift2.x * 0.121f, a3 = p.x * 0.02221f + p.y * 0.01233f + shift2.x * 3.437f,
                a4 = p.x * 0.03112f + p.y * 0.01122f + shift2.x * 4.269f;
    float wave = 0.0f;
    if (T::cosine) {
        wave += rcos(a0) * 4.5f; wave += rcos(a1) * 4.0f; wave -= rcos(a2) * 4.0f; wave += rcos(a3) * 5.0f; wave += rcos(a4) * 2.5f;
    } else {
        wave += rsin(a0) * 4.5f; wave += rsin(a1) * 4.0f; wave -= rsin(a2) * 4.0f; wave += rsin(a3) * 5.0f; wave += rsin(a4) * 2.5f;
    }
    wave *= T::large_wave;
    wave -= chop_fbm(p * 0.004f - shift2 * 0.5f) * T::small_wave * 24.0f;
    float amp = 6.0f * T::small_wave;
    shift1 = shift1 * 0.3f;
    const float m00 = 1.6f * 0.9331f, m01 = -1.2f * 0.9331f, m10 = 1.2f * 0.9331f, m11 = 1.6f * 0.9331f;
#pragma unroll 1
    for (int i = 0; i < T::octaves; ++i) {
        wave -= fabsf(rsin((value_noise2(p * 0.01f + shift1) - 0.5f) * 3.14f)) * amp;
        amp *= 0.51f;
        shift1 = shift1 * 1.841f;
        p = mk2(p.x * m00 + p.y * m01, p.x * m10 + p.y * m11);
    }
    return 70.0f + wave;
}

template <int S> DI float cloud_shadow(const OceanFrame &u, f3 start, f3 dir)   // trace_fog
{
    typedef OceanTraits<S> T;
    if (T::clouds == kCloudNone) return 1.0f;
    const float ct = T::clouds == kCloudStill ? u.cam[8] : u.cam[6];
    const float shx = ct * 80.0f, shy = ct * 60.0f;
    float sum = 0.0f, q2 = 0.0f, q3 = 0.0f;
#pragma unroll 1
    for (int q = 0; q < 10; ++q) {
        const f3 off = mk3(831.0f, 321.0f + q3 - shx * 0.2f, 1330.0f + shy * 3.0f);
        f3 cpos;
        if (T::clouds == kCloudDrift) cpos = start + dir * ((q2 + 350.0f - start.y) / dir.y) + off;
        else cpos = dir * ((q2 + 350.0f) / dir.y) + off;
        const float alpha = hermite(0.5f, 1.0f, cloud_fbm(cpos * 0.0015f));
        sum += (1.0f - sum) * alpha;
        if (sum > 0.98f) break;
        q2 += 120.0f;
        q3 += 0.15f;
    }
    return clamp01(1.0f - sum);
}

DI f3 view_rotate(float s1, float c1, float s2, float c2, float s3, float c3, f3 r)
{
    return mk3(c2 * c3 * r.x + (-c1 * s3 + s1 * s2 * c3) * r.y + (s1 * s3 + c1 * s2 * c3) * r.z,
               c2 * s3 * r.x + (c1 * c3 + s1 * s2 * s3) * r.y + (-s1 * c3 + c1 * s2 * s3) * r.z,
               -s2 * r.x + s1 * c2 * r.y + c1 * c2 * r.z);
}

template <int S> __device__ f3 shade_ocean(const OceanFrame &u, f2 tc, float nx, float ny)
{
    typedef OceanTraits<S> T;
    const float width = u.width, height = u.height;
    f2 xy = tc;
    if (T::jitter) { xy.x += nx / width; xy.y += ny / height; }
    const f3 campos = mk3(u.cam[0], u.cam[1], u.cam[2]);
    const f3 ray = unit(mk3((xy.x + 1.0f) * width / 2.0f - width / 2.0f, (xy.y + 1.0f) * height / 2.0f - height / 2.0f,
                            1.73f * width / 2.0f));
    float s1 = rsin(u.cam[3]), c1 = rcos(u.cam[3]), s2 = rsin(u.cam[4]), c2 = rcos(u.cam[4]), s3 = rsin(u.cam[5]), c3 = rcos(u.cam[5]);
    f3 rd = view_rotate(s1, c1, s2, c2, s3, c3, ray);
    const f3 light = unit(mk3(0.1f, 0.25f, u.cam[7]));
    float sundot = clamp01(dot3(rd, light));

    // march (trace; the god-ray block is compiled out in the reference, so fog == 0)
    float t = -campos.y / rd.y, st = 0.5f, old_h = 0.0f;
#pragma unroll 1
    for (int j = 0; j < T::steps; ++j) {
        if (t > 500.0f) st = 1.0f;
        if (t > 800.0f) st = 2.0f;
        if (t > 1500.0f) st = 3.0f;
        const f3 p = campos + rd * t;
        const float h = p.y - sea_height<S>(u, mk2(p.x, p.z));
        t += fmaxf(1.0f, fabsf(h)) * sgn(h) * st;
        if (old_h * h < 0.0f) st /= 2.0f;
        old_h = h;
    }
    const float dist = t, fog = 0.0f;
    const bool hit = !(rd.y > 0.0f);

    if (S == DIRT_SCENE_OCEANIC_OPT_FLOW) {
        const float dt = u.cam[9];
        const float b1 = u.cam[3] - u.cam[13] * dt, b2 = u.cam[4] - u.cam[14] * dt, b3 = u.cam[5] - u.cam[15] * dt;
        s1 = rsin(b1); c1 = rcos(b1); s2 = rsin(b2); c2 = rcos(b2); s3 = rsin(b3); c3 = rcos(b3);
        f3 old_rd = rd;
        if (hit) old_rd = (campos + rd * dist) - (campos - mk3(u.cam[10], u.cam[11], u.cam[12]) * dt);
        f3 o = mk3(c2 * c3 * old_rd.x + c2 * s3 * old_rd.y - s2 * old_rd.z,
                   (-c1 * s3 + s1 * s2 * c3) * old_rd.x + (c1 * c3 + s1 * s2 * s3) * old_rd.y + s1 * c2 * old_rd.z,
                   (s1 * s3 + c1 * s2 * c3) * old_rd.x + (-s1 * c3 + c1 * s2 * s3) * old_rd.y + c1 * c2 * old_rd.z);
        const float oz = o.z;
        o = mk3(o.x / oz, o.y / oz, o.z / oz) * (1.73f * width / 2.0f);
        return mk3(o.x + width / 2.0f, o.y + height / 2.0f, 0.0f);
    }

    const f3 skyb = mk3(T::skyb0, T::skyb1, T::skyb2), skyt = mk3(T::skyt0, T::skyt1, T::skyt2);
    f3 col;
    if (!hit) {
        const float tt = powf(1.0f - 0.7f * rd.y, 15.0f);
        col = (skyb * tt + skyt * (1.0f - tt)) * 0.8f;
        col = col + mk3(1.6f, 1.4f, 1.0f) * 0.47f * powf(sundot, 350.0f);
        col = col + mk3(0.8f, 0.9f, 1.0f) * 0.4f * powf(sundot, 2.0f);
        if (T::clouds != kCloudNone) {
            const float ct = T::clouds == kCloudStill ? u.cam[8] : u.cam[6];
            const float shx = ct * 80.0f, shy = ct * 60.0f;
            f3 acc = mk3(0.0f, 0.0f, 0.0f);
            float cover = 0.0f;
#pragma unroll 1
            for (int q = 0; q < 100; ++q) {
                const float fq = (float)q;
                const f3 off = mk3(831.0f, 321.0f + fq * 0.15f - shx * 0.2f, 1330.0f + shy * 3.0f);
                f3 cpos;
                if (T::clouds == kCloudDrift) cpos = campos + rd * ((fq * 12.0f + 350.0f - campos.y) / rd.y) + off;
                else cpos = rd * ((fq * 12.0f + 350.0f) / rd.y) + off;
                float alpha = hermite(0.5f, 1.0f, cloud_fbm(cpos * 0.0015f)) * 0.9f;
                const f3 local = lerp3(mk3(1.1f, 1.05f, 1.0f), mk3(0.4f, 0.4f, 0.3f) * 0.7f, alpha);
                alpha = (1.0f - cover) * alpha;
                acc = acc + local * alpha;
                cover += alpha;
                if (cover > 0.98f) break;
            }
            const float alpha = hermite(0.7f, 1.0f, cover);
            const float den = cover + 0.0001f;
            f3 rgb = mk3(acc.x / den, acc.y / den, acc.z / den);
            rgb = rgb - mk3(0.8f, 0.75f, 0.7f) * 0.6f * powf(sundot, 13.0f) * alpha;
            rgb = rgb + mk3(1.3f, 1.2f, 1.0f) * 0.2f * powf(sundot, 5.0f) * (1.0f - alpha);
            col = lerp3(col, rgb, cover * (1.0f - tt));
        }
        col = col + mk3(0.5f, 0.4f, 0.3f) * fog;
    } else {
        const f3 wpos = campos + rd * dist;
        const float e = 0.1f * T::wavegain * 4.0f;
        const f2 w = mk2(wpos.x, wpos.z);
        const f3 n = unit(mk3(sea_height<S>(u, w - mk2(e, 0.0f)) - sea_height<S>(u, w + mk2(e, 0.0f)), 1.0f,
                              sea_height<S>(u, w - mk2(0.0f, e)) - sea_height<S>(u, w + mk2(0.0f, e))));
        rd = rd - n * (2.0f * dot3(n, rd));
        const float refl = 1.0f - clamp01(dot3(rd, mk3(0.0f, 1.0f, 0.0f)));
        const float sh = hermite(0.2f, 1.0f, cloud_shadow<S>(u, wpos + rd * 20.0f, rd)) * 0.7f + 0.3f;
        const float wsky = refl * sh, wwater = (1.0f - refl) * sh;
        sundot = clamp01(dot3(rd, light));
        col = mk3(T::refl0, T::refl1, T::refl2) * wsky;
        col = col + mk3(T::wcol0, T::wcol1, T::wcol2) * wwater;
        col = col + mk3(0.003f, 0.005f, 0.005f) * (wpos.y - 70.0f + 30.0f);
        const float wsunrefl = wsky * (0.5f * powf(sundot, 10.0f) + 0.25f * powf(sundot, 3.5f) + 0.75f * powf(sundot, 300.0f));
        col = col + mk3(1.5f, 1.3f, 1.0f) * wsunrefl;
        const float fo = 1.0f - expf(-powf(0.0003f * dist, 1.5f));
        const f3 fco = mk3(T::fog0, T::fog1, T::fog2) + mk3(0.6f, 0.5f, 0.4f) * 0.6f * powf(sundot, 4.0f);
        col = lerp3(col, fco, fo);
        col = col + mk3(0.5f, 0.4f, 0.3f) * fog;
    }
    return col;
}

template <int S> __global__ void __launch_bounds__(128) k_scene_ocean(ShadeParams p)
{
    const int npix = p.H * p.W;
    const int o = (int)(blockIdx.x * blockDim.x + threadIdx.x);
    if (o >= npix) return;
    OceanFrame u;
#pragma unroll
    for (int k = 0; k < 16; ++k) u.cam[k] = 0.0f;
    const int ncam = S == DIRT_SCENE_OCEANIC_OPT_FLOW ? 16 : (S == DIRT_SCENE_OCEANIC_STILL_CLOUD ? 9 : 8);
#pragma unroll
    for (int k = 0; k < ncam; ++k) u.cam[k] = __ldg(p.camera_pos + k);
    u.width = (float)p.W; u.height = (float)p.H;
    for (int b = blockIdx.y; b < p.B; b += gridDim.y) {
        const size_t i = (size_t)b * npix + o;
        const int f = __ldg(p.face_id + i);
        if (f < 0) continue;
        const f2 tc = texcoord_at(p, b, i, f);
        float nx = 0.0f, ny = 0.0f;
        if (OceanTraits<S>::jitter) {
            const float *bg = p.background + (int64_t)b * p.background_bs + (size_t)o * p.Cb;
            nx = __ldg(bg); ny = p.Cb > 1 ? __ldg(bg + 1) : nx;
        }
        const f3 col = shade_ocean<S>(u, tc, nx, ny);
        float *out = p.pixels + i * p.C;
        out[0] = col.x;
        if (p.C > 1) out[1] = col.y;
        if (p.C > 2) out[2] = col.z;
        for (int c = 3; c < p.C; ++c) out[c] = 1.0f;
    }
}

// ---- hill -----------------------------------------------------------------------------------

struct HillTex { const float *tex; int H, W, Cb; };

// GL_LINEAR + CLAMP_TO_EDGE; texture row j is tensor row H-1-j (upload_background's flip).
// 1 channel uploads as (v,v,v,1), 3 as (r,g,b,1), 4 as is (csrc/rasterise_egl.cu:29-48).
template <int K0, int K1> DI void bilinear(const HillTex &c, float s, float t, float *r)
{
    const float u = s * (float)c.W - 0.5f, v = t * (float)c.H - 0.5f;
    const float fu = floorf(u), fv = floorf(v), a = u - fu, b = v - fv;
    int i0 = (int)fu, j0 = (int)fv, i1 = i0 + 1, j1 = j0 + 1;
    i0 = min(max(i0, 0), c.W - 1); i1 = min(max(i1, 0), c.W - 1);
    j0 = min(max(j0, 0), c.H - 1); j1 = min(max(j1, 0), c.H - 1);
    const float *r0 = c.tex + (size_t)(c.H - 1 - j0) * c.W * c.Cb, *r1 = c.tex + (size_t)(c.H - 1 - j1) * c.W * c.Cb;
#pragma unroll
    for (int k = K0; k <= K1; ++k) {
        if (k == 3 && c.Cb < 4) { r[k - K0] = 1.0f; continue; }
        const int ch = c.Cb == 1 ? 0 : k;
        const float t00 = __ldg(r0 + i0 * c.Cb + ch), t10 = __ldg(r0 + i1 * c.Cb + ch);
        const float t01 = __ldg(r1 + i0 * c.Cb + ch), t11 = __ldg(r1 + i1 * c.Cb + ch);
        r[k - K0] = lerp(lerp(t00, t10, a), lerp(t01, t11, a), b);
    }
}

DI f2 terrain_coord(f2 p)     // Terrain(): x in [-14,14] -> t, z in [5,25] -> s (csrc/shaders.cpp:222-236)
{
    return mk2(clamp01((p.y - 5.0f) / (-5.0f + 25.0f)), clamp01((p.x - -14.0f) / (14.0f + 14.0f)));
}
DI float terrain_height(const HillTex &c, f2 p)
{
    const f2 st = terrain_coord(p);
    float h;
    bilinear<0, 0>(c, st.x, st.y, &h);
    return h * 10.3f - 6.1f;
}

DI float hill_hash(float p)
{
    f2 q = mk2(fractf(p / 3.07965f), fractf(p / 7.4235f));
    const float d = q.y * (q.x + 19.19f) + q.x * (q.y + 19.19f);
    q = mk2(q.x + d, q.y + d);
    return fractf(q.x * q.y);
}
DI float hill_hash(f2 p)
{
    f2 q = mk2(fractf(p.x / 3.07965f), fractf(p.y / 7.4235f));
    const float d = q.x * (q.y + 19.19f) + q.y * (q.x + 19.19f);
    q = mk2(q.x + d, q.y + d);
    return fractf(q.x * q.y);
}
DI float hill_noise(f2 x)
{
    const f2 p = mk2(floorf(x.x), floorf(x.y));
    f2 f = mk2(fractf(x.x), fractf(x.y));
    f = mk2(f.x * f.x * (3.0f - 2.0f * f.x), f.y * f.y * (3.0f - 2.0f * f.y));
    const float n = p.x + p.y * 57.0f;
    return lerp(lerp(hill_hash(n + 0.0f), hill_hash(n + 1.0f), f.x), lerp(hill_hash(n + 57.0f), hill_hash(n + 58.0f), f.x), f.y);
}
DI f2 hill_voronoi(f2 x)
{
    const f2 p = mk2(floorf(x.x), floorf(x.y)), f = mk2(fractf(x.x), fractf(x.y));
    float res = 100.0f, id = 0.0f;
#pragma unroll 1
    for (int j = -1; j <= 1; ++j)
#pragma unroll
        for (int i = -1; i <= 1; ++i) {
            const f2 b = mk2((float)i, (float)j);
            const float hh = hill_hash(p + b);
            const f2 r = mk2(b.x - f.x + hh, b.y - f.y + hh);
            const float d = dot2(r, r);
            if (d < res) { res = d; id = hh; }
        }
    return mk2(fmaxf(0.4f - sqrtf(res), 0.0f), id);
}

DI f3 hill_sun() { return unit(mk3(0.35f, 0.2f, 0.3f)); }
DI f3 hill_sky(f3 rd)
{
    const f3 sun_col = mk3(1.0f, 0.75f, 0.6f);
    const float sun = fmaxf(dot3(rd, hill_sun()), 0.0f);
    const float v = powf(1.0f - fmaxf(rd.y, 0.0f), 6.0f);
    f3 sky = lerp3(mk3(0.1f, 0.2f, 0.3f), mk3(0.32f, 0.32f, 0.32f), v);
    sky = sky + sun_col * sun * sun * 0.25f;
    sky = sky + sun_col * fminf(powf(sun, 800.0f) * 1.5f, 0.3f);
    return mk3(clamp01(sky.x), clamp01(sky.y), clamp01(sky.z));
}

__device__ __noinline__ f3 hill_grass_field(const HillTex &c, f3 p)    // DE(), iTime = 0
{
    const float base = terrain_height(c, mk2(p.x, p.z)) - 1.3f;
    p.x *= 4.0f; p.z *= 4.0f;
    const f2 q = mk2(p.x, p.z);
    const float bump = hill_noise(q * 2.0f) * 0.75f + hill_noise(q) * 0.35f + hill_noise(q * 0.5f) * 0.2f;
    float y = p.y - base - bump;
    y = y * y;
    const f2 arg = mk2(p.x * 2.5f + rsin(y * 4.0f + p.z * 12.3f) * 0.12f + rsin(0.0f * 2.3f + 1.5f * p.z) * y * 0.5f,
                       p.z * 2.5f + rsin(y * 4.0f + p.x * 12.3f) * 0.12f + rsin(0.0f * 3.6f + 1.5f * p.x) * y * 0.5f);
    const f2 ret = hill_voronoi(arg);
    const float f = ret.x * 0.6f + y * 0.58f;
    return mk3(y - f * 1.4f, clamp01(f * 1.5f), ret.y);
}

__device__ f3 shade_hill(const HillTex &c, const float *cam, f2 tc, float width, float height)
{
    const f2 xy = mk2((tc.x + 1.0f) / 2.0f, (tc.y * -1.0f + 1.0f) / 2.0f);
    if (fabsf(xy.y * height - height / 2.0f) / (width / 2.0f) >= 0.5625f) return mk3(0.0f, 0.0f, 0.0f);
    const f3 ray = unit(mk3(xy.x * width - width / 2.0f, xy.y * height - height / 2.0f, 0.85f * width));
    const f3 campos = mk3(-__ldg(cam + 9), __ldg(cam + 11), __ldg(cam + 10));
    // the shader's local constants shadow the r00..r22 uniforms (csrc/shaders.cpp:437-445)
    const float r00 = 0.999999573f, r01 = -0.0000933038802f, r02 = 0.000919791287f, r10 = 0.000918443273f, r11 = -0.0135434586f,
                r12 = -0.999907861f, r20 = 0.000105752439f, r21 = 0.999908279f, r22 = -0.0135433672f;
    f3 dir;
    dir.x = -(r00 * ray.x + r10 * ray.y + r20 * ray.z);
    dir.z = (r01 * ray.x + r11 * ray.y + r21 * ray.z);
    dir.y = (r02 * ray.x + r12 * ray.y + r22 * ray.z);

    // Scene(): 100 relaxed steps against the height texture
    float t = -(campos.y + 1.0f) / dir.y, t_inc = 0.0f, h = 0.0f, st = 1.0f, old_h = 0.0f;
    if (dir.y > -0.015f) t = 80.0f;
#pragma unroll 1
    for (int j = 0; j < 100; ++j) {
        t += t_inc;
        const f3 p = campos + dir * t;
        h = p.y - terrain_height(c, mk2(p.x, p.z));
        t_inc = fmaxf(1.0f, fabsf(h)) * sgn(h) * st;
        if (h * old_h < 0.0f) st /= 2.0f;
        old_h = h;
    }
    f3 col;
    if (!(fabsf(h) < 0.05f)) {
        col = hill_sky(dir);
    } else {
        const float dist = t;
        const f3 pos = campos + dir * dist;
        const f2 st2 = terrain_coord(mk2(pos.x, pos.z));
        float nrm[3];
        bilinear<1, 3>(c, st2.x, st2.y, nrm);
        const f3 nor = mk3(nrm[0] * 2.0f - 1.0f, nrm[1] * 2.0f - 1.0f, nrm[2] * 2.0f - 1.0f);
        f3 mat = lerp3(mk3(0.0f, 0.3f, 0.0f), mk3(0.2f, 0.3f, 0.0f), hill_noise(mk2(pos.x, pos.z) * 0.025f));
        f2 fx = mk2(pos.x, pos.z) * 0.1f;
        float w = 0.7f, shade = 0.0f;
#pragma unroll
        for (int k = 0; k < 3; ++k) { shade += hill_noise(fx) * w; w = w * 0.6f; fx = fx * 2.0f; }
        shade = shade + 0.5f;
        // GrassBlades()
        const float tt = dist * 0.3f;
        const float rcoc = fmaxf(tt * 0.04f, (2.0f / height) * (1.0f + tt));
        f3 acc = mat * 0.15f;
        float cover = 0.0f, d = 0.0f;
#pragma unroll 1
        for (int k = 0; k < 15; ++k) {
            if (cover > 0.99f) break;
            f3 ret = hill_grass_field(c, pos + dir * d);
            ret.x += 0.5f * rcoc;
            if (ret.x < rcoc) {
                const float alpha = (1.0f - acc.y) * clamp01((-ret.x - -rcoc) / (rcoc - -rcoc));
                const f3 gra = lerp3(mat, mk3(0.35f, 0.35f, fminf(powf(ret.z, 4.0f) * 35.0f, 0.35f)), powf(ret.y, 9.0f) * 0.7f) * ret.y;
                acc = acc + gra * alpha;
                cover += alpha;
            }
            d += fmaxf(ret.x * 0.7f, 0.1f);
        }
        if (cover < 0.2f) acc = mk3(0.1f, 0.15f, 0.05f);
        mat = acc * shade;
        const float dl = dot3(hill_sun(), nor);
        mat = mat * mk3(1.0f, 0.75f, 0.6f) * (dl * dl * 4.0f);
        col = lerp3(mat, hill_sky(dir), clamp01(dist * dist * 0.0000012f));
    }
    // PostEffects()
    col = mk3(powf(col.x, 0.45f), powf(col.y, 0.45f), powf(col.z, 0.45f));
    const f3 bright = col * 1.3f;
    const float lum = dot3(mk3(0.2125f, 0.7154f, 0.0721f), bright);
    col = lerp3(mk3(0.5f, 0.5f, 0.5f), lerp3(mk3(lum, lum, lum), bright, 1.3f), 1.1f);
    return col * (0.4f + 0.5f * powf(40.0f * xy.x * xy.y * (1.0f - xy.x) * (1.0f - xy.y), 0.2f));
}

__global__ void __launch_bounds__(128) k_scene_hill(ShadeParams p)
{
    const int npix = p.H * p.W;
    const int o = (int)(blockIdx.x * blockDim.x + threadIdx.x);
    if (o >= npix) return;
    for (int b = blockIdx.y; b < p.B; b += gridDim.y) {
        const size_t i = (size_t)b * npix + o;
        const int f = __ldg(p.face_id + i);
        if (f < 0) continue;
        HillTex tex;
        tex.tex = p.background + (int64_t)b * p.background_bs; tex.H = p.H; tex.W = p.W; tex.Cb = p.Cb;
        const f3 col = shade_hill(tex, p.camera_pos, texcoord_at(p, b, i, f), (float)p.W, (float)p.H);
        float *out = p.pixels + i * p.C;
        out[0] = col.x;
        if (p.C > 1) out[1] = col.y;
        if (p.C > 2) out[2] = col.z;
        for (int c = 3; c < p.C; ++c) out[c] = 1.0f;
    }
}

cudaError_t launch_scene(int scene, const ShadeParams &p, cudaStream_t stream)
{
    if ((long long)p.B * p.H * p.W == 0) return cudaSuccess;
    const dim3 g((unsigned)((p.H * p.W + 127) / 128), (unsigned)(p.B < 65535 ? p.B : 65535));
    switch (scene) {
    case DIRT_SCENE_OCEANIC: k_scene_ocean<DIRT_SCENE_OCEANIC><<<g, 128, 0, stream>>>(p); break;
    case DIRT_SCENE_OCEANIC_STILL_CLOUD: k_scene_ocean<DIRT_SCENE_OCEANIC_STILL_CLOUD><<<g, 128, 0, stream>>>(p); break;
    case DIRT_SCENE_OCEANIC_NO_CLOUD: k_scene_ocean<DIRT_SCENE_OCEANIC_NO_CLOUD><<<g, 128, 0, stream>>>(p); break;
    case DIRT_SCENE_OCEANIC_SIMPLE_PROXY: k_scene_ocean<DIRT_SCENE_OCEANIC_SIMPLE_PROXY><<<g, 128, 0, stream>>>(p); break;
    case DIRT_SCENE_OCEANIC_OPT_FLOW: k_scene_ocean<DIRT_SCENE_OCEANIC_OPT_FLOW><<<g, 128, 0, stream>>>(p); break;
    case DIRT_SCENE_HILL: k_scene_hill<<<g, 128, 0, stream>>>(p); break;
    default: return cudaErrorInvalidValue;
    }
    note_launches(1);
    return cudaGetLastError();
}

}  // namespace dirt

// ---- C ABI ---------------------------------------------------------------------------------

extern "C" int dirt_scene_camera_floats(int scene)
{
    switch (scene) {
    case DIRT_SCENE_OCEANIC_HORIZON: case DIRT_SCENE_OCEANIC: case DIRT_SCENE_OCEANIC_NO_CLOUD: case DIRT_SCENE_OCEANIC_SIMPLE_PROXY: return 8;
    case DIRT_SCENE_OCEANIC_STILL_CLOUD: return 9;
    case DIRT_SCENE_OCEANIC_OPT_FLOW: return 16;
    case DIRT_SCENE_HILL: return 12;
    default: return -1;
    }
}

extern "C" int dirt_scene_shade(int scene, const DirtSceneArgs *a, int background_channels, int camera_pos_count, void *stream)
{
    if (!a) return DIRT_STATUS_INVALID_ARGUMENT;
    const int need = dirt_scene_camera_floats(scene);
    if (need < 0 || camera_pos_count < need) return DIRT_STATUS_INVALID_ARGUMENT;
    if (scene == DIRT_SCENE_OCEANIC_HORIZON) {
        if (background_channels != a->shape.channels) return DIRT_STATUS_INVALID_ARGUMENT;
        return dirt_scene_oceanic_horizon(a, stream);
    }
    const DirtShape &s = a->shape;
    if (s.height <= 0 || s.width <= 0 || s.channels < 1 || s.channels > DIRT_MAX_CHANNELS || s.batch < 0) return DIRT_STATUS_INVALID_ARGUMENT;
    // upload_background knows 1, 3 and 4 channels (csrc/rasterise_egl.cu:29-48)
    if (background_channels != 1 && background_channels != 3 && background_channels != 4) return DIRT_STATUS_INVALID_ARGUMENT;
    if (s.batch == 0) return DIRT_STATUS_OK;
    if (!a->background || !a->vertices || !a->faces || !a->face_id || !a->gbuffer || !a->camera_pos || !a->pixels)
        return DIRT_STATUS_INVALID_ARGUMENT;
    dirt::ShadeParams p;
    p.B = s.batch; p.H = s.height; p.W = s.width; p.C = s.channels; p.Cb = background_channels;
    p.background = a->background; p.background_bs = a->background_bstride;
    p.vertices = a->vertices; p.vertices_bs = a->vertices_bstride;
    p.faces = a->faces; p.faces_bs = a->faces_bstride;
    p.face_id = a->face_id; p.gbuffer = a->gbuffer; p.camera_pos = a->camera_pos; p.pixels = a->pixels;
    return dirt::launch_scene(scene, p, static_cast<cudaStream_t>(stream)) == cudaSuccess ? DIRT_STATUS_OK : DIRT_STATUS_CUDA_ERROR;
}
