/*
 * scene_shaders.cpp -- CPU restatement of the fork's other procedural fragment stages
 * (SURVEY.md 8(f) row 2).  TEST INFRASTRUCTURE ONLY (see golden_raster.c): nothing under
 * dirt_b200/ includes, links or calls this file.
 *
 *   scene 1  oceanic            /root/reference csrc/shaders.cpp:556-865   (op RasteriseGrad, csrc/rasterise_grad_egl.cpp:399)
 *   scene 2  oceanic_still_cloud   csrc/shaders.cpp:866-1175   (csrc/oceanic_still_cloud.cpp:385, cloud_t = camera_pos[8], :407)
 *   scene 3  oceanic_no_cloud      csrc/shaders.cpp:1402-1665  (csrc/oceanic_no_cloud.cpp:385)
 *   scene 4  oceanic_simple_proxy  csrc/shaders.cpp:1921-2185  (csrc/oceanic_simple_proxy.cpp:385)
 *   scene 5  oceanic_opt_flow      csrc/shaders.cpp:1178-1398  (csrc/oceanic_opt_flow.cpp:385, uniforms :408-414)
 *   scene 6  hill                  csrc/shaders.cpp:123-552    (csrc/hill.cpp:378, uniforms :395-406)
 * (scene 0, oceanic_horizon, is oracle/scene_oceanic.c.)
 *
 * The five ocean shaders are one program with different constants, so they are restated once,
 * driven by a table (OceanSpec) that lists exactly what differs between the GLSL sources.
 * Arithmetic is binary32, every operation individually rounded (-ffp-contract=off), with sin / cos
 * correctly rounded (see rsin) and libm's powf / expf.  GLSL built-ins are written out by their specification formulas
 * (mix = x(1-a)+ya, smoothstep = Hermite of the clamped ratio, reflect = I - 2 dot(N,I) N,
 * normalize = v / length).  Matrix products are summed left to right.
 *
 * Parity status: UNPINNED -- the reference stores no rendered image for any of these ops and a
 * GL driver's sin / cos / pow are approximations no CPU code reproduces bit for bit.
 *
 * Reading of under-specified GLSL:
 *   - RENDER_GODRAYS is never defined (shaders.cpp:10-14): the god-ray block is compiled out, so
 *     `fog` is always 0 and trace() never leaves its loop early.
 *   - `texture(backgroundTexture, (texCoordV+1)/2)` samples the attachment being rendered with no
 *     filter set; read as "the background at this pixel" (same reading as scene_oceanic.c).
 *   - uninitialised locals (`old_h` in hill's Scene()) start at 0.
 *   - hill: pow(x, 2.0) of a possibly negative x (DoLighting) is taken as x*x, its evident meaning
 *     (GLSL leaves pow of a negative base undefined).  The r00..r22 uniforms are shadowed by local
 *     constants in main() (shaders.cpp:437-445) and therefore dead; o0..o2 are live.
 *   - hill: TerrainLookup is the background tensor uploaded with a vertical flip
 *     (csrc/rasterise_egl.cu:27) and sampled GL_LINEAR / CLAMP_TO_EDGE (csrc/hill.cpp:233-237) with
 *     full float weights; each batch element samples its own frame (the reference samples the
 *     whole atlas, which only coincides for a batch of one -- the only way the fork calls it).
 *   - hill renders without a depth test into a target that is never cleared: the value depends on
 *     the fragment position only, so for the full-screen quads the fork draws the depth-tested
 *     rasteriser gives the same picture; uncovered pixels are defined as 0.
 */
#include <cmath>
#include <cstdint>

namespace {

struct vec2 { float x, y; };
struct vec3 { float x, y, z; };
struct vec4 { float x, y, z, w; };

inline vec2 V2(float x, float y) { return vec2{x, y}; }
inline vec3 V3(float x, float y, float z) { return vec3{x, y, z}; }
inline vec2 operator+(vec2 a, vec2 b) { return V2(a.x + b.x, a.y + b.y); }
inline vec2 operator-(vec2 a, vec2 b) { return V2(a.x - b.x, a.y - b.y); }
inline vec2 operator*(vec2 a, float s) { return V2(a.x * s, a.y * s); }
inline vec3 operator+(vec3 a, vec3 b) { return V3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline vec3 operator-(vec3 a, vec3 b) { return V3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline vec3 operator*(vec3 a, float s) { return V3(a.x * s, a.y * s, a.z * s); }
inline vec3 operator*(vec3 a, vec3 b) { return V3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline float dot(vec2 a, vec2 b) { return a.x * b.x + a.y * b.y; }
inline float dot(vec3 a, vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
/* sin / cos correctly rounded to binary32 (double evaluation, one rounding): the noise hashes multiply
 * them by ~4e4, so a one-ulp difference between two maths libraries would move a hash by 3e-3 and the
 * chaotic march with it; computed this way the CUDA kernels and this file agree to the last bit in all
 * but ~1e-9 of the calls.  A GLSL implementation may return any value this close to sin(x). */
inline float rsin(float v) { return (float)sin((double)v); }
inline float rcos(float v) { return (float)cos((double)v); }
inline float fract(float v) { return v - floorf(v); }
inline float clampf(float v, float lo, float hi) { return fminf(fmaxf(v, lo), hi); }
inline float signf(float v) { return v > 0.0f ? 1.0f : (v < 0.0f ? -1.0f : 0.0f); }
inline float mixf(float a, float b, float t) { return a * (1.0f - t) + b * t; }
inline vec3 mix3(vec3 a, vec3 b, float t) { return a * (1.0f - t) + b * t; }
inline float smoothstepf(float e0, float e1, float v)
{
    float t = clampf((v - e0) / (e1 - e0), 0.0f, 1.0f);
    return t * t * (3.0f - 2.0f * t);
}
inline vec3 normalize(vec3 v)
{
    float l = sqrtf(dot(v, v));
    return V3(v.x / l, v.y / l, v.z / l);
}
inline vec3 reflect(vec3 i, vec3 n) { return i - n * (2.0f * dot(n, i)); }

/* ------------------------------------------------------------------------------------------ */
/* ocean family                                                                               */
/* ------------------------------------------------------------------------------------------ */

enum CloudMode { CLOUD_NONE, CLOUD_DRIFT, CLOUD_STILL };

struct OceanSpec {
    float wavegain, large_wave, small_wave;            /* shaders.cpp:576-578 / :1941-1943 / :1207-1208 */
    vec3 fogcolor, skybottom, skytop, reflsky, watercolor;
    float s1x, s1y, s2x, s2y;                          /* |multipliers| of time in shift1 / shift2 (:652-653, :2017-2018) */
    bool cosine;                                       /* simple_proxy swaps sin for cos in the swell (:2022-2027) */
    int octaves, steps;                                /* chop octaves (:669 / :2034), march steps (:712 / :2060) */
    CloudMode clouds;
    bool flat_sea;                                     /* opt_flow: water() = waterlevel - 12 (:1280) */
};

const vec3 FOG0 = {0.5f, 0.7f, 1.1f}, SKYB0 = {0.6f, 0.8f, 1.2f}, SKYT0 = {0.05f, 0.2f, 0.5f}, REFL0 = {0.025f, 0.10f, 0.20f},
           WATER0 = {0.2f, 0.25f, 0.3f};
const OceanSpec SPECS[6] = {
    {},
    {1.0f, 1.0f, 1.0f, FOG0, SKYB0, SKYT0, REFL0, WATER0, 160.0f, 120.0f, 190.0f, 130.0f, false, 7, 20, CLOUD_DRIFT, false},
    {1.0f, 1.0f, 1.0f, FOG0, SKYB0, SKYT0, REFL0, WATER0, 160.0f, 120.0f, 190.0f, 130.0f, false, 7, 20, CLOUD_STILL, false},
    {1.0f, 1.0f, 1.0f, FOG0, SKYB0, SKYT0, REFL0, WATER0, 160.0f, 120.0f, 190.0f, 130.0f, false, 7, 20, CLOUD_NONE, false},
    {0.75f, 0.75f, 1.5f, {0.4f, 0.4f, 1.2f}, {0.5f, 0.5f, 1.3f}, {0.15f, 0.1f, 0.7f}, {0.1f, 0.1f, 0.15f}, {0.1f, 0.2f, 0.5f},
     260.0f, 100.0f, 150.0f, 230.0f, true, 3, 10, CLOUD_NONE, false},
    {1.0f, 0.0f, 0.0f, FOG0, SKYB0, SKYT0, REFL0, WATER0, 160.0f, 120.0f, 190.0f, 130.0f, false, 7, 20, CLOUD_NONE, true},
};

const float WATERLEVEL = 70.0f;

inline float hash1(float n) { return fract(rcos(n) * 41415.92653f); }                         /* :590-593 */
inline float rand2(vec2 n) { return fract(rsin(dot(n, V2(12.9898f, 4.1414f))) * 43758.5453f); }  /* :595-597 */

float noise2(vec2 p)                                                                          /* :599-608 */
{
    vec2 ip = V2(floorf(p.x), floorf(p.y));
    vec2 u = V2(fract(p.x), fract(p.y));
    u = V2(u.x * u.x * (3.0f - 2.0f * u.x), u.y * u.y * (3.0f - 2.0f * u.y));
    return mixf(mixf(rand2(ip), rand2(ip + V2(1.0f, 0.0f)), u.x),
                mixf(rand2(ip + V2(0.0f, 1.0f)), rand2(ip + V2(1.0f, 1.0f)), u.x), u.y);
}

float noise3(vec3 x)                                                                          /* :611-621 */
{
    vec3 p = V3(floorf(x.x), floorf(x.y), floorf(x.z));
    vec3 f = V3(smoothstepf(0.0f, 1.0f, fract(x.x)), smoothstepf(0.0f, 1.0f, fract(x.y)), smoothstepf(0.0f, 1.0f, fract(x.z)));
    float n = p.x + p.y * 57.0f + 113.0f * p.z;
    return mixf(mixf(mixf(hash1(n + 0.0f), hash1(n + 1.0f), f.x), mixf(hash1(n + 57.0f), hash1(n + 58.0f), f.x), f.y),
                mixf(mixf(hash1(n + 113.0f), hash1(n + 114.0f), f.x), mixf(hash1(n + 170.0f), hash1(n + 171.0f), f.x), f.y), f.z);
}

/* mat3 m, column major (:624): columns (0,1.6,1.2), (-1.6,.72,-.96), (-1.2,-.96,1.28) */
inline vec3 mul_m(vec3 p)
{
    return V3(0.00f * p.x + -1.60f * p.y + -1.20f * p.z, 1.60f * p.x + 0.72f * p.y + -0.96f * p.z,
              1.20f * p.x + -0.96f * p.y + 1.28f * p.z);
}
/* mat2 m2 = (1.6,-1.2 | 1.2,1.6), column major (:636) */
inline vec2 mul_m2(vec2 p) { return V2(1.6f * p.x + 1.2f * p.y, -1.2f * p.x + 1.6f * p.y); }

float fbm3(vec3 p)                                                                            /* :627-634 */
{
    float f = 0.5000f * noise3(p); p = mul_m(p) * 1.1f;
    f += 0.2500f * noise3(p); p = mul_m(p) * 1.2f;
    f += 0.1666f * noise3(p); p = mul_m(p);
    f += 0.0834f * noise3(p);
    return f;
}

float fbm2(vec2 p)                                                                            /* :639-646 */
{
    float f = 0.5000f * noise2(p); p = mul_m2(p);
    f += 0.2500f * noise2(p); p = mul_m2(p);
    f += 0.1666f * noise2(p); p = mul_m2(p);
    f += 0.0834f * noise2(p);
    return f;
}

struct OceanUniforms { float cam[16]; float width, height; };

float water(const OceanSpec &s, const OceanUniforms &u, vec2 p)                               /* :648-679 */
{
    if (s.flat_sea) return WATERLEVEL - 12.0f;
    const float time = u.cam[6];
    float height = WATERLEVEL;
    vec2 shift1 = V2(0.001f * (time * s.s1x * 2.0f), 0.001f * (time * s.s1y * 2.0f));
    vec2 shift2 = V2(0.001f * (time * s.s2x * 2.0f), 0.001f * (-time * s.s2y * 2.0f));
    float wave = 0.0f;
    float a0 = p.x * 0.021f + shift2.x, a1 = p.x * 0.0172f + p.y * 0.010f + shift2.x * 1.121f,
          a2 = p.x * 0.00104f + p.y * 0.005f + shift2.x * 0.121f, a3 = p.x * 0.02221f + p.y * 0.01233f + shift2.x * 3.437f,
          a4 = p.x * 0.03112f + p.y * 0.01122f + shift2.x * 4.269f;
    if (s.cosine) {
        wave += rcos(a0) * 4.5f; wave += rcos(a1) * 4.0f; wave -= rcos(a2) * 4.0f; wave += rcos(a3) * 5.0f; wave += rcos(a4) * 2.5f;
    } else {
        wave += rsin(a0) * 4.5f; wave += rsin(a1) * 4.0f; wave -= rsin(a2) * 4.0f; wave += rsin(a3) * 5.0f; wave += rsin(a4) * 2.5f;
    }
    wave *= s.large_wave;
    wave -= fbm2(p * 0.004f - shift2 * 0.5f) * s.small_wave * 24.0f;
    float amp = 6.0f * s.small_wave;
    shift1 = shift1 * 0.3f;
    const float m00 = 1.6f * 0.9331f, m01 = -1.2f * 0.9331f, m10 = 1.2f * 0.9331f, m11 = 1.6f * 0.9331f;   /* m2*0.9331, columns */
    for (int i = 0; i < s.octaves; ++i) {
        wave -= fabsf(rsin((noise2(p * 0.01f + shift1) - 0.5f) * 3.14f)) * amp;
        amp *= 0.51f;
        shift1 = shift1 * 1.841f;
        p = V2(p.x * m00 + p.y * m01, p.x * m10 + p.y * m11);      /* p *= M : row vector times matrix */
    }
    height += wave;
    return height;
}

float trace_fog(const OceanSpec &s, const OceanUniforms &u, vec3 start, vec3 dir)            /* :681-700 / :990-1009 / :1527-1529 */
{
    if (s.clouds == CLOUD_NONE) return 1.0f;
    const float ct = s.clouds == CLOUD_STILL ? u.cam[8] : u.cam[6];
    const vec2 shift = V2(ct * 80.0f, ct * 60.0f);
    float sum = 0.0f, q2 = 0.0f, q3 = 0.0f;
    for (int q = 0; q < 10; ++q) {
        float c;
        vec3 cpos;
        const vec3 off = V3(831.0f, 321.0f + q3 - shift.x * 0.2f, 1330.0f + shift.y * 3.0f);
        if (s.clouds == CLOUD_DRIFT) {
            c = (q2 + 350.0f - start.y) / dir.y;
            cpos = start + dir * c + off;
        } else {
            c = (q2 + 350.0f) / dir.y;
            cpos = dir * c + off;
        }
        float alpha = smoothstepf(0.5f, 1.0f, fbm3(cpos * 0.0015f));
        sum += (1.0f - sum) * alpha;
        if (sum > 0.98f) break;
        q2 += 120.0f;
        q3 += 0.15f;
    }
    return clampf(1.0f - sum, 0.0f, 1.0f);
}

bool trace(const OceanSpec &s, const OceanUniforms &u, vec3 start, vec3 dir, float *dist)     /* :702-746 */
{
    float t = -start.y / dir.y, st = 0.5f, old_h = 0.0f;
    for (int j = 0; j < s.steps; ++j) {
        if (t > 500.0f) st = 1.0f;
        if (t > 800.0f) st = 2.0f;
        if (t > 1500.0f) st = 3.0f;
        vec3 p = start + dir * t;
        float h = p.y - water(s, u, V2(p.x, p.z));
        t += fmaxf(1.0f, fabsf(h)) * signf(h) * st;
        if (old_h * h < 0.0f) st /= 2.0f;
        old_h = h;
    }
    *dist = t;
    return !(dir.y > 0.0f);
}

inline vec3 rotate_view(float s1, float c1, float s2, float c2, float s3, float c3, vec3 r)   /* :776-779 */
{
    return V3(c2 * c3 * r.x + (-c1 * s3 + s1 * s2 * c3) * r.y + (s1 * s3 + c1 * s2 * c3) * r.z,
              c2 * s3 * r.x + (c1 * c3 + s1 * s2 * s3) * r.y + (-s1 * c3 + c1 * s2 * s3) * r.z,
              -s2 * r.x + s1 * c2 * r.y + c1 * c2 * r.z);
}

void shade_ocean(int scene, const OceanUniforms &u, float tx, float ty, float nx, float ny, float out[3])
{
    const OceanSpec &s = SPECS[scene];
    const float width = u.width, height = u.height, time = u.cam[6];
    vec2 xy = V2(tx, ty);
    if (scene != 5) { xy.x += nx / width; xy.y += ny / height; }                               /* :751-754; commented out at :1325-1327 */
    const vec3 campos = V3(u.cam[0], u.cam[1], u.cam[2]);
    vec3 ray = V3((xy.x + 1.0f) * width / 2.0f - width / 2.0f, (xy.y + 1.0f) * height / 2.0f - height / 2.0f, 1.73f * width / 2.0f);
    ray = normalize(ray);
    float s1 = rsin(u.cam[3]), c1 = rcos(u.cam[3]), s2 = rsin(u.cam[4]), c2 = rcos(u.cam[4]), s3 = rsin(u.cam[5]), c3 = rcos(u.cam[5]);
    vec3 rd = rotate_view(s1, c1, s2, c2, s3, c3, ray);
    const vec3 light = normalize(V3(0.1f, 0.25f, u.cam[7]));                                   /* :586 */
    float sundot = clampf(dot(rd, light), 0.0f, 1.0f);
    float dist = 0.0f;
    const float fog = 0.0f;
    const bool hit = trace(s, u, campos, rd, &dist);

    if (scene == 5) {                                                                          /* :1359-1398 */
        const float dt = u.cam[9];
        const vec3 vel = V3(u.cam[10], u.cam[11], u.cam[12]);
        const float b1 = u.cam[3] - u.cam[13] * dt, b2 = u.cam[4] - u.cam[14] * dt, b3 = u.cam[5] - u.cam[15] * dt;
        s1 = rsin(b1); c1 = rcos(b1); s2 = rsin(b2); c2 = rcos(b2); s3 = rsin(b3); c3 = rcos(b3);
        vec3 old_rd;
        if (!hit) {
            old_rd = rd;
        } else {
            vec3 wpos = campos + rd * dist;
            vec3 old_campos = campos - vel * dt;
            old_rd = V3(wpos.x, wpos.y, wpos.z) - old_campos;
        }
        /* transpose of rotate_view */
        vec3 o = V3(c2 * c3 * old_rd.x + c2 * s3 * old_rd.y - s2 * old_rd.z,
                    (-c1 * s3 + s1 * s2 * c3) * old_rd.x + (c1 * c3 + s1 * s2 * s3) * old_rd.y + s1 * c2 * old_rd.z,
                    (s1 * s3 + c1 * s2 * c3) * old_rd.x + (-s1 * c3 + c1 * s2 * s3) * old_rd.y + c1 * c2 * old_rd.z);
        const float oz = o.z;
        o = V3(o.x / oz, o.y / oz, o.z / oz);
        o = o * (1.73f * width / 2.0f);
        out[0] = o.x + width / 2.0f;
        out[1] = o.y + height / 2.0f;
        out[2] = 0.0f;
        return;
    }

    vec3 col;
    if (!hit) {                                                                                /* sky: :788-829 */
        float t = powf(1.0f - 0.7f * rd.y, 15.0f);
        col = (s.skybottom * t + s.skytop * (1.0f - t)) * 0.8f;
        col = col + V3(1.6f, 1.4f, 1.0f) * 0.47f * powf(sundot, 350.0f);
        col = col + V3(0.8f, 0.9f, 1.0f) * 0.4f * powf(sundot, 2.0f);
        if (s.clouds != CLOUD_NONE) {
            const float ct = s.clouds == CLOUD_STILL ? u.cam[8] : time;
            const vec2 shift = V2(ct * 80.0f, ct * 60.0f);
            vec4 sum = {0.0f, 0.0f, 0.0f, 0.0f};
            for (int q = 0; q < 100; ++q) {
                const float fq = (float)q;
                const vec3 off = V3(831.0f, 321.0f + fq * 0.15f - shift.x * 0.2f, 1330.0f + shift.y * 3.0f);
                float c;
                vec3 cpos;
                if (s.clouds == CLOUD_DRIFT) {
                    c = (fq * 12.0f + 350.0f - campos.y) / rd.y;
                    cpos = campos + rd * c + off;
                } else {
                    c = (fq * 12.0f + 350.0f) / rd.y;
                    cpos = rd * c + off;
                }
                float alpha = smoothstepf(0.5f, 1.0f, fbm3(cpos * 0.0015f)) * 0.9f;
                vec3 local = mix3(V3(1.1f, 1.05f, 1.0f), V3(0.4f, 0.4f, 0.3f) * 0.7f, alpha);
                alpha = (1.0f - sum.w) * alpha;
                sum.x += local.x * alpha; sum.y += local.y * alpha; sum.z += local.z * alpha; sum.w += alpha;
                if (sum.w > 0.98f) break;
            }
            float alpha = smoothstepf(0.7f, 1.0f, sum.w);
            const float den = sum.w + 0.0001f;
            vec3 rgb = V3(sum.x / den, sum.y / den, sum.z / den);
            rgb = rgb - V3(0.8f, 0.75f, 0.7f) * 0.6f * powf(sundot, 13.0f) * alpha;
            rgb = rgb + V3(1.3f, 1.2f, 1.0f) * 0.2f * powf(sundot, 5.0f) * (1.0f - alpha);
            col = mix3(col, rgb, sum.w * (1.0f - t));
        }
        col = col + V3(0.5f, 0.4f, 0.3f) * fog;
    } else {                                                                                   /* sea: :830-861 */
        vec3 wpos = campos + rd * dist;
        const float e = 0.1f * s.wavegain * 4.0f;
        vec2 w = V2(wpos.x, wpos.z);
        vec3 n = normalize(V3(water(s, u, w - V2(e, 0.0f)) - water(s, u, w + V2(e, 0.0f)), 1.0f,
                              water(s, u, w - V2(0.0f, e)) - water(s, u, w + V2(0.0f, e))));
        rd = reflect(rd, n);
        float refl = 1.0f - clampf(dot(rd, V3(0.0f, 1.0f, 0.0f)), 0.0f, 1.0f);
        float sh = smoothstepf(0.2f, 1.0f, trace_fog(s, u, wpos + rd * 20.0f, rd)) * 0.7f + 0.3f;
        float wsky = refl * sh, wwater = (1.0f - refl) * sh;
        sundot = clampf(dot(rd, light), 0.0f, 1.0f);
        col = s.reflsky * wsky;
        col = col + s.watercolor * wwater;
        col = col + V3(0.003f, 0.005f, 0.005f) * (wpos.y - WATERLEVEL + 30.0f);
        float wsunrefl = wsky * (0.5f * powf(sundot, 10.0f) + 0.25f * powf(sundot, 3.5f) + 0.75f * powf(sundot, 300.0f));
        col = col + V3(1.5f, 1.3f, 1.0f) * wsunrefl;
        float fo = 1.0f - expf(-powf(0.0003f * dist, 1.5f));
        vec3 fco = s.fogcolor + V3(0.6f, 0.5f, 0.4f) * 0.6f * powf(sundot, 4.0f);
        col = mix3(col, fco, fo);
        col = col + V3(0.5f, 0.4f, 0.3f) * fog;
    }
    out[0] = col.x; out[1] = col.y; out[2] = col.z;
}

/* ------------------------------------------------------------------------------------------ */
/* hill                                                                                       */
/* ------------------------------------------------------------------------------------------ */

struct HillCtx {
    const float *tex;   /* this frame's background [H,W,Cb], top row first */
    int H, W, Cb;
    float width, height;
};

/* GL_LINEAR, CLAMP_TO_EDGE lookup of channel set (r,g,b,a); texture row j = tensor row H-1-j. */
vec4 texture_lookup(const HillCtx &c, float s, float t)
{
    float u = s * (float)c.W - 0.5f, v = t * (float)c.H - 0.5f;
    float fu = floorf(u), fv = floorf(v);
    float a = u - fu, b = v - fv;
    int i0 = (int)fu, j0 = (int)fv, i1 = i0 + 1, j1 = j0 + 1;
    i0 = i0 < 0 ? 0 : (i0 > c.W - 1 ? c.W - 1 : i0); i1 = i1 < 0 ? 0 : (i1 > c.W - 1 ? c.W - 1 : i1);
    j0 = j0 < 0 ? 0 : (j0 > c.H - 1 ? c.H - 1 : j0); j1 = j1 < 0 ? 0 : (j1 > c.H - 1 ? c.H - 1 : j1);
    float r[4];
    for (int k = 0; k < 4; ++k) {
        /* upload_background: 1 channel -> (v,v,v,1); 3 -> (r,g,b,1); 4 -> as is (csrc/rasterise_egl.cu:29-48) */
        float t00, t10, t01, t11;
        if (k == 3 && c.Cb < 4) { r[k] = 1.0f; continue; }
        const int ch = c.Cb == 1 ? 0 : k;
        t00 = c.tex[((long)(c.H - 1 - j0) * c.W + i0) * c.Cb + ch];
        t10 = c.tex[((long)(c.H - 1 - j0) * c.W + i1) * c.Cb + ch];
        t01 = c.tex[((long)(c.H - 1 - j1) * c.W + i0) * c.Cb + ch];
        t11 = c.tex[((long)(c.H - 1 - j1) * c.W + i1) * c.Cb + ch];
        r[k] = mixf(mixf(t00, t10, a), mixf(t01, t11, a), b);
    }
    return vec4{r[0], r[1], r[2], r[3]};
}

inline vec2 fract_div_mod2(vec2 p) { return V2(fract(p.x / 3.07965f), fract(p.y / 7.4235f)); }   /* MOD2, :160 */

float hill_hash1(float p)                                                                     /* :171-176 */
{
    vec2 p2 = fract_div_mod2(V2(p, p));
    float d = dot(V2(p2.y, p2.x), V2(p2.x + 19.19f, p2.y + 19.19f));
    p2 = V2(p2.x + d, p2.y + d);
    return fract(p2.x * p2.y);
}

float hill_hash2(vec2 p)                                                                      /* :179-184 */
{
    p = fract_div_mod2(p);
    float d = dot(V2(p.x, p.y), V2(p.y + 19.19f, p.x + 19.19f));
    p = V2(p.x + d, p.y + d);
    return fract(p.x * p.y);
}

float hill_noise(vec2 x)                                                                      /* :188-197 */
{
    vec2 p = V2(floorf(x.x), floorf(x.y));
    vec2 f = V2(fract(x.x), fract(x.y));
    f = V2(f.x * f.x * (3.0f - 2.0f * f.x), f.y * f.y * (3.0f - 2.0f * f.y));
    float n = p.x + p.y * 57.0f;
    return mixf(mixf(hill_hash1(n + 0.0f), hill_hash1(n + 1.0f), f.x), mixf(hill_hash1(n + 57.0f), hill_hash1(n + 58.0f), f.x), f.y);
}

vec2 voronoi(vec2 x)                                                                          /* :199-217 */
{
    vec2 p = V2(floorf(x.x), floorf(x.y));
    vec2 f = V2(fract(x.x), fract(x.y));
    float res = 100.0f, id = 0.0f;
    for (int j = -1; j <= 1; ++j)
        for (int i = -1; i <= 1; ++i) {
            vec2 b = V2((float)i, (float)j);
            float hh = hill_hash2(p + b);
            vec2 r = V2(b.x - f.x + hh, b.y - f.y + hh);
            float d = dot(r, r);
            if (d < res) { res = d; id = hh; }
        }
    return V2(fmaxf(0.4f - sqrtf(res), 0.0f), id);
}

inline vec2 terrain_uv(vec2 p)                                                                /* :229-236: range x [-14,14], z [5,25] */
{
    const vec2 tr = V2(-14.0f, 5.0f), sc = V2(14.0f + 14.0f, -5.0f + 25.0f);
    return V2((p.x - tr.x) / sc.x, (p.y - tr.y) / sc.y);
}
float terrain(const HillCtx &c, vec2 p)
{
    vec2 q = terrain_uv(p);
    return texture_lookup(c, clampf(q.y, 0.0f, 1.0f), clampf(q.x, 0.0f, 1.0f)).x * 10.3f - 6.1f;
}
vec3 terrain_normal(const HillCtx &c, vec2 p)                                                 /* :239-247 */
{
    vec2 q = terrain_uv(p);
    vec4 t = texture_lookup(c, clampf(q.y, 0.0f, 1.0f), clampf(q.x, 0.0f, 1.0f));
    return V3(t.y * 2.0f - 1.0f, t.z * 2.0f - 1.0f, t.w * 2.0f - 1.0f);
}

float fractal_noise(vec2 xy)                                                                  /* :257-269 */
{
    float w = 0.7f, f = 0.0f;
    for (int i = 0; i < 3; ++i) {
        f += hill_noise(xy) * w;
        w = w * 0.6f;
        xy = xy * 2.0f;
    }
    return f;
}

const vec3 SUN_COLOUR = {1.0f, 0.75f, 0.6f};
inline vec3 sun_light() { return normalize(V3(0.35f, 0.2f, 0.3f)); }

vec3 get_sky(vec3 rd)                                                                         /* :273-281 */
{
    float sun = fmaxf(dot(rd, sun_light()), 0.0f);
    float v = powf(1.0f - fmaxf(rd.y, 0.0f), 6.0f);
    vec3 sky = mix3(V3(0.1f, 0.2f, 0.3f), V3(0.32f, 0.32f, 0.32f), v);
    sky = sky + SUN_COLOUR * sun * sun * 0.25f;
    sky = sky + SUN_COLOUR * fminf(powf(sun, 800.0f) * 1.5f, 0.3f);
    return V3(clampf(sky.x, 0.0f, 1.0f), clampf(sky.y, 0.0f, 1.0f), clampf(sky.z, 0.0f, 1.0f));
}

vec3 grass_de(const HillCtx &c, vec3 p)                                                       /* DE, :293-305; iTime = 0 */
{
    float base = terrain(c, V2(p.x, p.z)) - 1.3f;
    p.x *= 4.0f; p.z *= 4.0f;
    vec2 pxz = V2(p.x, p.z);
    float height = hill_noise(pxz * 2.0f) * 0.75f + hill_noise(pxz) * 0.35f + hill_noise(pxz * 0.5f) * 0.2f;
    float y = p.y - base - height;
    y = y * y;
    const float iTime = 0.0f;
    vec2 arg = V2(p.x * 2.5f + rsin(y * 4.0f + p.z * 12.3f) * 0.12f + rsin(iTime * 2.3f + 1.5f * p.z) * y * 0.5f,
                  p.z * 2.5f + rsin(y * 4.0f + p.x * 12.3f) * 0.12f + rsin(iTime * 3.6f + 1.5f * p.x) * y * 0.5f);
    vec2 ret = voronoi(arg);
    float f = ret.x * 0.6f + y * 0.58f;
    return V3(y - f * 1.4f, clampf(f * 1.5f, 0.0f, 1.0f), ret.y);
}

inline float linstep(float a, float b, float t) { return clampf((t - a) / (b - a), 0.0f, 1.0f); }

vec3 grass_blades(const HillCtx &c, vec3 ro, vec3 rd, vec3 mat, float dist)                   /* :323-352 */
{
    float d = 0.0f;
    const float tt = dist * 0.3f;
    float rcoc = fmaxf(tt * 0.04f, (2.0f / c.height) * (1.0f + tt));                          /* CircleOfConfusion, :310-313 */
    vec4 col = {mat.x * 0.15f, mat.y * 0.15f, mat.z * 0.15f, 0.0f};
    for (int i = 0; i < 15; ++i) {
        if (col.w > 0.99f) break;
        vec3 p = ro + rd * d;
        vec3 ret = grass_de(c, p);
        ret.x += 0.5f * rcoc;
        if (ret.x < rcoc) {
            float alpha = (1.0f - col.y) * linstep(-rcoc, rcoc, -ret.x);
            vec3 gra = mix3(mat, V3(0.35f, 0.35f, fminf(powf(ret.z, 4.0f) * 35.0f, 0.35f)), powf(ret.y, 9.0f) * 0.7f) * ret.y;
            col.x += gra.x * alpha; col.y += gra.y * alpha; col.z += gra.z * alpha; col.w += alpha;
        }
        d += fmaxf(ret.x * 0.7f, 0.1f);
    }
    if (col.w < 0.2f) return V3(0.1f, 0.15f, 0.05f);
    return V3(col.x, col.y, col.z);
}

bool hill_scene(const HillCtx &c, vec3 ro, vec3 rd, float *res_t)                             /* Scene, :392-421 */
{
    float t = -(ro.y + 1.0f) / rd.y, t_inc = 0.0f, h = 0.0f, st = 1.0f, old_h = 0.0f;
    if (rd.y > -0.015f) t = 80.0f;
    for (int j = 0; j < 100; ++j) {
        t += t_inc;
        vec3 p = ro + rd * t;
        h = p.y - terrain(c, V2(p.x, p.z));
        t_inc = fmaxf(1.0f, fabsf(h)) * signf(h) * st;
        if (h * old_h < 0.0f) st /= 2.0f;
        old_h = h;
    }
    *res_t = t;
    return fabsf(h) < 0.05f;
}

vec3 post_effects(vec3 rgb, vec2 xy)                                                          /* :433-447 */
{
    rgb = V3(powf(rgb.x, 0.45f), powf(rgb.y, 0.45f), powf(rgb.z, 0.45f));
    vec3 bright = rgb * 1.3f;
    float lum = dot(V3(0.2125f, 0.7154f, 0.0721f), bright);
    vec3 sat = mix3(V3(lum, lum, lum), bright, 1.3f);
    rgb = mix3(V3(0.5f, 0.5f, 0.5f), sat, 1.1f);
    float vig = 0.4f + 0.5f * powf(40.0f * xy.x * xy.y * (1.0f - xy.x) * (1.0f - xy.y), 0.2f);
    return rgb * vig;
}

void shade_hill(const HillCtx &c, const float *cam, float tx, float ty, float out[3])        /* main, :450-548 */
{
    const float width = c.width, height = c.height;
    vec2 xy = V2((tx + 1.0f) / 2.0f, (ty * -1.0f + 1.0f) / 2.0f);
    out[0] = out[1] = out[2] = 0.0f;
    if (fabsf(xy.y * height - height / 2.0f) / (width / 2.0f) >= 0.5625f) return;
    vec3 ray = normalize(V3(xy.x * width - width / 2.0f, xy.y * height - height / 2.0f, 0.85f * width));
    const vec3 campos = V3(-cam[9], cam[11], cam[10]);
    const float r00 = 0.999999573f, r01 = -0.0000933038802f, r02 = 0.000919791287f, r10 = 0.000918443273f, r11 = -0.0135434586f,
                r12 = -0.999907861f, r20 = 0.000105752439f, r21 = 0.999908279f, r22 = -0.0135433672f;
    vec3 dir;
    dir.x = -(r00 * ray.x + r10 * ray.y + r20 * ray.z);
    dir.z = (r01 * ray.x + r11 * ray.y + r21 * ray.z);
    dir.y = (r02 * ray.x + r12 * ray.y + r22 * ray.z);
    vec3 col;
    float dist = 0.0f;
    if (!hill_scene(c, campos, dir, &dist)) {
        col = get_sky(dir);
    } else {
        vec3 pos = campos + dir * dist;
        vec3 nor = terrain_normal(c, V2(pos.x, pos.z));
        /* TerrainColour, :365-380 (type == 0) */
        vec3 mat = mix3(V3(0.0f, 0.3f, 0.0f), V3(0.2f, 0.3f, 0.0f), hill_noise(V2(pos.x, pos.z) * 0.025f));
        float t = fractal_noise(V2(pos.x, pos.z) * 0.1f) + 0.5f;
        mat = grass_blades(c, pos, dir, mat, dist) * t;
        float dl = dot(sun_light(), nor);
        float h = dl * dl * 4.0f;                                                             /* DoLighting, :356-361 */
        mat = mat * SUN_COLOUR * h;
        float fogamount = clampf(dist * dist * 0.0000012f, 0.0f, 1.0f);                        /* ApplyFog, :285-289 */
        col = mix3(mat, get_sky(dir), fogamount);
    }
    col = post_effects(col, xy);
    out[0] = col.x; out[1] = col.y; out[2] = col.z;
}

}  // namespace

/* texcoord [B,H,W,2] (interpolated clip xy), covered [B,H,W] (face id, < 0 = uncovered),
 * background [B,H,W,Cb], pixels [B,H,W,C] (already holding what uncovered pixels should show),
 * camera_pos [n_cam].  scene 1..6 as listed in the header. */
extern "C" int golden_scene_shade(int scene, int B, int H, int W, int C, int Cb, const float *texcoord, const int32_t *covered,
                                  const float *background, const float *camera_pos, int n_cam, float *pixels)
{
    static const int need[7] = {8, 8, 9, 8, 8, 16, 12};
    if (scene < 1 || scene > 6 || B < 0 || H <= 0 || W <= 0 || C <= 0 || Cb <= 0 || n_cam < need[scene]) return -1;
    OceanUniforms u;
    for (int i = 0; i < 16; ++i) u.cam[i] = i < n_cam ? camera_pos[i] : 0.0f;
    u.width = (float)W; u.height = (float)H;
    for (int b = 0; b < B; ++b) {
        HillCtx hc = {background + (long)b * H * W * Cb, H, W, Cb, (float)W, (float)H};
        for (long o = 0; o < (long)H * W; ++o) {
            const long i = (long)b * H * W + o;
            if (covered[i] < 0) continue;
            float out[3];
            const float *bg = background + i * Cb;
            if (scene == 6) shade_hill(hc, camera_pos, texcoord[2 * i], texcoord[2 * i + 1], out);
            else shade_ocean(scene, u, texcoord[2 * i], texcoord[2 * i + 1], bg[0], Cb > 1 ? bg[1] : bg[0], out);
            for (int c = 0; c < C && c < 3; ++c) pixels[i * C + c] = out[c];
            for (int c = 3; c < C; ++c) pixels[i * C + c] = 1.0f;
        }
    }
    return 0;
}
