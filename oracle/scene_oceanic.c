/*
 * scene_oceanic.c -- CPU restatement of the fragment shader the fork's `Rasterise` op actually
 * runs: shaders::oceanic_horizon, /root/reference csrc/shaders.cpp:1668-1919, with the uniforms
 * of csrc/rasterise_egl.cpp:399-433 (camera_pos[0..7] = cam_x, cam_y, cam_z, ang1, ang2, ang3,
 * time, light_z; width, height = frame size).  TEST INFRASTRUCTURE ONLY (see golden_raster.c).
 *
 * It is the GLSL transcribed to C, statement by statement, in binary32 with libm's sinf / cosf /
 * powf.  Parity status: unpinned -- the reference stores no image, and a GL driver's sin / pow are
 * approximations this file cannot reproduce bit for bit.
 *   - RENDER_GODRAYS is not defined in the shader header (shaders.cpp:10-14), so `#if
 *     RENDER_GODRAYS` (:1833) is false and trace_fog / fbm(vec3) are dead code.
 *   - small_waveheight = 0.0 (:1692): the fbm term (:1777) and the 7-octave loop (:1782-1788) are
 *     multiplied by an exact 0 and contribute exactly 0 (noise() is finite); they are skipped.
 *   - `texture(backgroundTexture, (texCoordV + 1) / 2)` (:1864): the fork binds the colour
 *     attachment it is rendering to, without setting a filter (csrc/rasterise_egl.cpp has no
 *     glTextureParameteri), so what GL returns there is undefined.  The evident intent
 *     (tests/render_random_video_seq.py:16, :170-174 feed a random-normal background to jitter
 *     the samples) is "the background value at this pixel"; that is what is implemented.
 * The per-pixel inputs are the interpolated texCoordV (= clip-space xy of the quad, `smooth`,
 * csrc/shaders.cpp:16-31) and the background at the pixel; uncovered pixels are left alone.
 */
#include <math.h>
#include <stdint.h>

typedef struct { float x, y; } vec2;
typedef struct { float x, y, z; } vec3;

static const float waterlevel = 70.0f, wavegain = 1.0f, large_waveheight = 1.0f;   /* :1689-1691 */

static float clampf(float v, float lo, float hi) { return fminf(fmaxf(v, lo), hi); }
static float signf(float v) { return v > 0.0f ? 1.0f : (v < 0.0f ? -1.0f : 0.0f); }
static vec3 normalize3(vec3 v)
{
    float l = sqrtf(v.x * v.x + v.y * v.y + v.z * v.z);
    vec3 r = {v.x / l, v.y / l, v.z / l};
    return r;
}
static float dot3(vec3 a, vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }

/* water(): csrc/shaders.cpp:1762-1793 with small_waveheight = 0 */
static float water(vec2 p, float time)
{
    float height = waterlevel;
    vec2 shift2 = {0.001f * (time * 190.0f * 2.0f), 0.001f * (-time * 130.0f * 2.0f)};
    float wave = 0.0f;
    wave += sinf(p.x * 0.021f + shift2.x) * 4.5f;
    wave += sinf(p.x * 0.0172f + p.y * 0.010f + shift2.x * 1.121f) * 4.0f;
    wave -= sinf(p.x * 0.00104f + p.y * 0.005f + shift2.x * 0.121f) * 4.0f;
    wave += sinf(p.x * 0.02221f + p.y * 0.01233f + shift2.x * 3.437f) * 5.0f;
    wave += sinf(p.x * 0.03112f + p.y * 0.01122f + shift2.x * 4.269f) * 2.5f;
    wave *= large_waveheight;
    height += wave;
    return height;
}

/* trace(): csrc/shaders.cpp:1816-1856 (godrays compiled out) */
static int trace(vec3 rStart, vec3 rd, float time, float *dist)
{
    float h, t = -rStart.y / rd.y, st = 0.5f, old_h = 0.0f;
    for (int j = 1000; j < 1020; ++j) {
        if (t > 500.0f) st = 1.0f;
        if (t > 800.0f) st = 2.0f;
        if (t > 1500.0f) st = 3.0f;
        vec3 p = {rStart.x + t * rd.x, rStart.y + t * rd.y, rStart.z + t * rd.z};
        vec2 pxz = {p.x, p.z};
        h = p.y - water(pxz, time);
        t += fmaxf(1.0f, fabsf(h)) * signf(h) * st;
        if (old_h * h < 0.0f) st /= 2.0f;
        old_h = h;
    }
    *dist = t;
    return rd.y > 0.0f ? 0 : 1;
}

/* main(): csrc/shaders.cpp:1858-1917.  out = fragColor.xyz */
static void shade(float tx, float ty, float nx, float ny, const float *cam, float width, float height, float *out)
{
    vec2 xy = {tx + nx / width, ty + ny / height};
    vec3 campos = {cam[0], cam[1], cam[2]};
    vec3 ray_dir = {(xy.x + 1.0f) * width / 2.0f - width / 2.0f, (xy.y + 1.0f) * height / 2.0f - height / 2.0f,
                    1.73f * width / 2.0f};
    ray_dir = normalize3(ray_dir);
    float sin1 = sinf(cam[3]), cos1 = cosf(cam[3]), sin2 = sinf(cam[4]), cos2 = cosf(cam[4]), sin3 = sinf(cam[5]), cos3 = cosf(cam[5]);
    vec3 rd;
    rd.x = cos2 * cos3 * ray_dir.x + (-cos1 * sin3 + sin1 * sin2 * cos3) * ray_dir.y + (sin1 * sin3 + cos1 * sin2 * cos3) * ray_dir.z;
    rd.y = cos2 * sin3 * ray_dir.x + (cos1 * cos3 + sin1 * sin2 * sin3) * ray_dir.y + (-sin1 * cos3 + cos1 * sin2 * sin3) * ray_dir.z;
    rd.z = -sin2 * ray_dir.x + sin1 * cos2 * ray_dir.y + cos1 * cos2 * ray_dir.z;
    vec3 l0 = {0.1f, 0.25f, cam[7]};
    vec3 light = normalize3(l0);                                               /* :1699 */
    float time = cam[6];
    float sundot = clampf(dot3(rd, light), 0.0f, 1.0f);
    float dist = 0.0f;
    out[0] = out[1] = out[2] = 0.0f;
    if (!trace(campos, rd, time, &dist)) {
        out[0] = 1.0f;
        out[1] = powf(sundot, 350.0f);
    } else {
        vec3 wpos = {campos.x + dist * rd.x, campos.y + dist * rd.y, campos.z + dist * rd.z};
        float dx = 0.1f * wavegain * 4.0f;
        vec2 a0 = {wpos.x - dx, wpos.z}, a1 = {wpos.x + dx, wpos.z}, b0 = {wpos.x, wpos.z - dx}, b1 = {wpos.x, wpos.z + dx};
        vec3 n0 = {water(a0, time) - water(a1, time), 1.0f, water(b0, time) - water(b1, time)};
        vec3 n = normalize3(n0);
        float k = 2.0f * dot3(n, rd);                                          /* reflect(I, N) = I - 2 dot(N, I) N */
        vec3 r = {rd.x - k * n.x, rd.y - k * n.y, rd.z - k * n.z};
        sundot = clampf(dot3(r, light), 0.0f, 1.0f);
        out[1] = 0.5f * powf(sundot, 10.0f) + 0.25f * powf(sundot, 3.5f) + 0.75f * powf(sundot, 300.0f);
    }
}

/* texcoord [B,H,W,2] (interpolated clip xy), covered [B,H,W] (face id, < 0 = uncovered),
 * background / pixels [B,H,W,C] (pixels must already hold the background), camera_pos [>=8]. */
int golden_oceanic_horizon(int B, int H, int W, int C, const float *texcoord, const int32_t *covered,
                           const float *background, const float *camera_pos, float *pixels)
{
    if (B < 0 || H <= 0 || W <= 0 || C <= 0) return -1;
    for (long i = 0; i < (long)B * H * W; ++i) {
        if (covered[i] < 0) continue;
        float out[3];
        const float *bg = background + i * C;
        shade(texcoord[2 * i], texcoord[2 * i + 1], bg[0], C > 1 ? bg[1] : bg[0], camera_pos, (float)W, (float)H, out);
        for (int c = 0; c < C && c < 3; ++c) pixels[i * C + c] = out[c];
        for (int c = 3; c < C; ++c) pixels[i * C + c] = 1.0f;                   /* fragColor.w */
    }
    return 0;
}
