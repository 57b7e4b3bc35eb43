/*
 * golden_raster.c -- scalar CPU restatement of the `rasterise` forward op and its
 * `rasterise_grad` backward.  TEST INFRASTRUCTURE ONLY: nothing under dirt_b200/ may
 * import, link or execute this file; only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs use it, and only as the checker / CPU baseline.
 *
 * Parity status
 *   - coverage / face-id: PINNED by the reference's two analytic known-answer tests
 *     (README.md:46-70 "KAT-1", tests/multi_gpu_test.py:8-17 "KAT-2"); see tests/test_oracle_kat.py.
 *   - interpolated attribute values: parity unpinned (the reference stores no images).
 *   - gradients: parity unpinned (the reference fork has no gradient code at all:
 *     csrc/rasterise_grad_common.h:13-24 are declarations only).  The gradient implemented
 *     here is this project's written spec, docs/GRADIENT_SPEC.md.
 *
 * Why a restatement: the reference performs rasterisation inside the OpenGL driver
 * (glDrawElementsBaseVertex, csrc/rasterise_egl.cpp:451-457), which has no source in the
 * reference tree and cannot run on this box.  What this file restates is the GL state the
 * reference configures plus the OpenGL 3.3 core rules that state selects:
 *   - vertex format: 4 x f32 clip-space position, C x f32 colour     csrc/rasterise_egl.cpp:177-189
 *   - pass-through vertex shader, `smooth` (perspective-correct) colour csrc/shaders.cpp:16-34
 *   - viewport = the element's own HxW cell, scissor = same           csrc/rasterise_egl.cpp:446-447
 *   - depth test on, default func GL_LESS, depth cleared to 1.0 per element  :194, :449
 *   - 24-bit unorm depth buffer (GL_DEPTH24_STENCIL8)                csrc/rasterise_egl.cpp:248
 *   - no face culling, no blending; colour buffer is NOT cleared, so the background shows
 *     through at uncovered pixels                                     :194-195, :370-377
 *   - triangles drawn in index order, faces index element-local vertices  :451-457
 *   - output rows flipped: tensor row 0 is the TOP of the GL image    csrc/rasterise_egl.cu:29, :93
 *   - float RGBA32F colour storage => no clamping of values           csrc/rasterise_egl.cpp:219
 * Fixed-point choices that GL leaves to the implementation (sub-pixel snapping to 1/256 px,
 * tie rule for samples exactly on an edge) are written down in docs/RASTER_SPEC.md.
 *
 * Build: gcc -O2 -ffp-contract=off -fno-fast-math -fopenmp -shared -fPIC  (oracle/Makefile).
 * -ffp-contract=off matters: every float operation below is individually rounded, which is
 * what makes face-id maps comparable bit-for-bit with the CUDA path.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <float.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define SUBPIX_BITS 8
#define SUBPIX (1 << SUBPIX_BITS)          /* 256 sub-pixel steps per pixel          */
#define SUBPIX_HALF (SUBPIX / 2)           /* pixel centre offset                    */
#define GUARD_PX 2097152.0f                /* 2^21 px: window coords are clamped here */
#define DEPTH_MAX 16777215u                /* 2^24 - 1                               */
#define KEY_EMPTY 0xFFFFFFFFFFFFFFFFull

typedef struct {
    float x, y, z, w;      /* clip-space position                                 */
    float beta[3];         /* barycentric weights w.r.t. the ORIGINAL 3 vertices  */
} ClipVert;

/* A triangle ready for sampling: snapped window positions + per-vertex depth and 1/w. */
typedef struct {
    int32_t X[3], Y[3];    /* window position in 1/256 px, y up                   */
    float zw[3];           /* window depth in [0,1]                               */
    float rw[3];           /* 1 / clip w                                          */
    float beta[3][3];      /* beta[k][j]: weight of original vertex j at corner k */
    int64_t area2;         /* twice the signed area (before sign normalisation)   */
    int sign;              /* +1 / -1 so that sign * edge >= 0 inside             */
    float rcp_area;        /* 1 / |area2| as float                                */
    float zref, dzdx, dzdy; /* depth plane re-based to the centre of pixel (xmin, ymin);
                              steps per pixel (docs/RASTER_SPEC.md section 6)       */
    int bias[3];           /* 0 if edge k is inclusive (top-left), else -1        */
    int xmin, xmax, ymin, ymax;   /* inclusive pixel bbox clipped to viewport, y up */
} SubTri;

/* What a sample of a SubTri yields at a covered pixel. */
typedef struct {
    float zw;              /* interpolated window depth                           */
    float B0, B1;          /* perspective-correct barycentrics w.r.t. original v0, v1 */
    float w_clip;          /* clip-space w of the surface point (= 1/gl_FragCoord.w,
                              csrc/shaders.cpp:2216)                              */
} Sample;

static int finite4(const float *v)
{
    return fabsf(v[0]) <= FLT_MAX && fabsf(v[1]) <= FLT_MAX &&
           fabsf(v[2]) <= FLT_MAX && fabsf(v[3]) <= FLT_MAX;
}

/* Sutherland-Hodgman against one depth plane.  plane 0: z >= -w (near), plane 1: z <= w (far).
 * GL clips primitives to -w <= z <= w before the perspective divide (OpenGL 3.3 core 2.20);
 * the reference never enables depth clamp (csrc/rasterise_egl.cpp:194-195).
 * The intersection point is always computed from the inside vertex towards the outside one,
 * so two faces sharing the crossing edge get the identical new vertex (no cracks). */
static int clip_depth_plane(const ClipVert *in, int n, ClipVert *out, int plane)
{
    float d[6];
    int m = 0;
    for (int i = 0; i < n; ++i)
        d[i] = plane == 0 ? in[i].z + in[i].w : in[i].w - in[i].z;
    for (int i = 0; i < n; ++i) {
        int j = (i + 1 == n) ? 0 : i + 1;
        int in_i = d[i] >= 0.0f, in_j = d[j] >= 0.0f;
        if (in_i) out[m++] = in[i];
        if (in_i != in_j) {
            int a = in_i ? i : j, o = in_i ? j : i;
            float t = d[a] / (d[a] - d[o]);
            ClipVert v;
            v.x = in[a].x + t * (in[o].x - in[a].x);
            v.y = in[a].y + t * (in[o].y - in[a].y);
            v.z = in[a].z + t * (in[o].z - in[a].z);
            v.w = in[a].w + t * (in[o].w - in[a].w);
            for (int k = 0; k < 3; ++k)
                v.beta[k] = in[a].beta[k] + t * (in[o].beta[k] - in[a].beta[k]);
            out[m++] = v;
        }
    }
    return m;
}

/* Clip one face; returns the polygon (0, or 3..5 vertices). */
static int clip_face(const float *v0, const float *v1, const float *v2, ClipVert *poly)
{
    const float *v[3] = {v0, v1, v2};
    if (!finite4(v0) || !finite4(v1) || !finite4(v2)) return 0;
    int near_out = 0, far_out = 0, l = 0, r = 0, b = 0, t = 0;
    for (int k = 0; k < 3; ++k) {
        float x = v[k][0], y = v[k][1], z = v[k][2], w = v[k][3];
        near_out += (z + w < 0.0f);
        far_out += (w - z < 0.0f);
        l += (x < -w); r += (x > w); b += (y < -w); t += (y > w);
    }
    /* trivially outside one frustum plane: nothing can be visible */
    if (near_out == 3 || far_out == 3 || l == 3 || r == 3 || b == 3 || t == 3) return 0;
    ClipVert a[6], c[6];
    for (int k = 0; k < 3; ++k) {
        a[k].x = v[k][0]; a[k].y = v[k][1]; a[k].z = v[k][2]; a[k].w = v[k][3];
        a[k].beta[0] = k == 0; a[k].beta[1] = k == 1; a[k].beta[2] = k == 2;
    }
    int n = 3;
    if (near_out) { n = clip_depth_plane(a, n, c, 0); memcpy(a, c, sizeof(ClipVert) * n); }
    if (n < 3) return 0;
    if (far_out) { n = clip_depth_plane(a, n, c, 1); memcpy(a, c, sizeof(ClipVert) * n); }
    if (n < 3) return 0;
    memcpy(poly, a, sizeof(ClipVert) * n);
    return n;
}

static int32_t snap(float ndc, float half_extent)
{
    /* viewport transform (glViewport, csrc/rasterise_egl.cpp:446): win = (ndc + 1) * extent / 2 */
    float f = (ndc + 1.0f) * half_extent;
    if (f > GUARD_PX) f = GUARD_PX;
    if (f < -GUARD_PX) f = -GUARD_PX;
    return (int32_t)floorf(f * (float)SUBPIX + 0.5f);
}

static int64_t orient(int32_t ax, int32_t ay, int32_t bx, int32_t by, int32_t px, int32_t py)
{
    return (int64_t)(bx - ax) * (int64_t)(py - ay) - (int64_t)(by - ay) * (int64_t)(px - ax);
}

static int floor_div_subpix(int32_t v) { return v >> SUBPIX_BITS; }   /* arithmetic shift = floor */

/* Build the sampling record for corners (p0,p1,p2) of the clipped polygon.  Returns 0 if culled. */
static int setup_subtri(const ClipVert *p0, const ClipVert *p1, const ClipVert *p2,
                        int W, int H, SubTri *t)
{
    const ClipVert *p[3] = {p0, p1, p2};
    float halfW = 0.5f * (float)W, halfH = 0.5f * (float)H;
    for (int k = 0; k < 3; ++k) {
        float w = p[k]->w;
        if (!(w > 0.0f)) return 0;
        float rw = 1.0f / w;
        if (!(rw <= FLT_MAX)) return 0;
        float xn = p[k]->x * rw, yn = p[k]->y * rw, zn = p[k]->z * rw;
        if (!(fabsf(xn) <= FLT_MAX) || !(fabsf(yn) <= FLT_MAX) || !(fabsf(zn) <= FLT_MAX)) return 0;
        t->X[k] = snap(xn, halfW);
        t->Y[k] = snap(yn, halfH);
        t->zw[k] = zn * 0.5f + 0.5f;
        t->rw[k] = rw;
        for (int j = 0; j < 3; ++j) t->beta[k][j] = p[k]->beta[j];
    }
    t->area2 = orient(t->X[0], t->Y[0], t->X[1], t->Y[1], t->X[2], t->Y[2]);
    if (t->area2 == 0) return 0;
    t->sign = t->area2 > 0 ? 1 : -1;
    t->rcp_area = 1.0f / (float)(t->area2 > 0 ? t->area2 : -t->area2);
    /* edge k is opposite corner k: from corner (k+1)%3 to corner (k+2)%3 */
    for (int k = 0; k < 3; ++k) {
        int a = (k + 1) % 3, b = (k + 2) % 3;
        int32_t dx = (t->X[b] - t->X[a]) * t->sign, dy = (t->Y[b] - t->Y[a]) * t->sign;
        /* tie rule for a sample exactly on an edge (docs/RASTER_SPEC.md section 4): with the
         * interior on the left of the directed edge (y up), the edge owns its samples iff it
         * runs downwards, or is horizontal and runs towards -x. */
        int inclusive = (dy < 0) || (dy == 0 && dx < 0);
        t->bias[k] = inclusive ? 0 : -1;
    }
    int32_t minX = t->X[0], maxX = t->X[0], minY = t->Y[0], maxY = t->Y[0];
    for (int k = 1; k < 3; ++k) {
        if (t->X[k] < minX) minX = t->X[k];
        if (t->X[k] > maxX) maxX = t->X[k];
        if (t->Y[k] < minY) minY = t->Y[k];
        if (t->Y[k] > maxY) maxY = t->Y[k];
    }
    /* pixels whose centre (px*256+128) lies in [min,max]; scissor = viewport (:447) */
    t->xmin = floor_div_subpix(minX + SUBPIX_HALF - 1);
    t->xmax = floor_div_subpix(maxX - SUBPIX_HALF);
    t->ymin = floor_div_subpix(minY + SUBPIX_HALF - 1);
    t->ymax = floor_div_subpix(maxY - SUBPIX_HALF);
    if (t->xmin < 0) t->xmin = 0;
    if (t->ymin < 0) t->ymin = 0;
    if (t->xmax > W - 1) t->xmax = W - 1;
    if (t->ymax > H - 1) t->ymax = H - 1;
    if (t->xmin > t->xmax || t->ymin > t->ymax) return 0;
    /* window depth is affine in window x,y (OpenGL 3.3 core 3.6.1, eq. 3.10): plane through
     * corner 0, gradients per sub-pixel unit, every operation individually rounded */
    {
        float fA = (float)t->area2;
        float dz1 = t->zw[1] - t->zw[0], dz2 = t->zw[2] - t->zw[0];
        float dX1 = (float)(t->X[1] - t->X[0]), dY1 = (float)(t->Y[1] - t->Y[0]);
        float dX2 = (float)(t->X[2] - t->X[0]), dY2 = (float)(t->Y[2] - t->Y[0]);
        float a = dz1 * dY2, b = dz2 * dY1, c = dz2 * dX1, d = dz1 * dX2;
        float gx = (a - b) / fA, gy = (c - d) / fA;
        float rx = (float)(t->xmin * SUBPIX + SUBPIX_HALF - t->X[0]);
        float ry = (float)(t->ymin * SUBPIX + SUBPIX_HALF - t->Y[0]);
        float u = gx * rx, v = gy * ry;
        float uv = u + v;
        t->zref = t->zw[0] + uv;
        t->dzdx = gx * (float)SUBPIX;
        t->dzdy = gy * (float)SUBPIX;
    }
    return 1;
}

/* Sample SubTri t at the centre of pixel (px, py) (y up).  Returns 1 when the sample is
 * covered AND passes the depth-range clip; fills s. */
static int sample_subtri(const SubTri *t, int px, int py, Sample *s)
{
    int32_t cx = px * SUBPIX + SUBPIX_HALF, cy = py * SUBPIX + SUBPIX_HALF;
    int64_t e[3];
    for (int k = 0; k < 3; ++k) {
        int a = (k + 1) % 3, b = (k + 2) % 3;
        e[k] = orient(t->X[a], t->Y[a], t->X[b], t->Y[b], cx, cy) * t->sign;
        if (e[k] + t->bias[k] < 0) return 0;
    }
    float l0 = (float)e[0] * t->rcp_area;
    float l1 = (float)e[1] * t->rcp_area;
    float l2 = (float)e[2] * t->rcp_area;
    /* depth from the plane; clamped to the depth range (fragments of a clipped polygon lie
     * inside it up to rounding) */
    float zx = t->dzdx * (float)(px - t->xmin), zy = t->dzdy * (float)(py - t->ymin);
    float zw = (t->zref + zx) + zy;
    zw = fminf(fmaxf(zw, 0.0f), 1.0f);
    /* perspective-correct weights (eq. 3.9): b_k = (l_k / w_k) / sum_m (l_m / w_m) */
    float q0 = l0 * t->rw[0], q1 = l1 * t->rw[1], q2 = l2 * t->rw[2];
    float qs = (q0 + q1) + q2;
    float rq = 1.0f / qs;
    float b0 = q0 * rq, b1 = q1 * rq, b2 = q2 * rq;
    s->zw = zw;
    s->w_clip = rq;
    s->B0 = (b0 * t->beta[0][0] + b1 * t->beta[1][0]) + b2 * t->beta[2][0];
    s->B1 = (b0 * t->beta[0][1] + b1 * t->beta[1][1]) + b2 * t->beta[2][1];
    return 1;
}

static uint32_t depth24(float zw)
{
    /* GL_DEPTH24_STENCIL8 (csrc/rasterise_egl.cpp:248): unorm24, round to nearest */
    return (uint32_t)(zw * 16777215.0f + 0.5f);
}

/* ------------------------------------------------------------------------------------ *
 * Forward: one batch element.  Mirrors the per-element draw of csrc/rasterise_egl.cpp:440-458
 * (viewport = own cell, depth cleared, all F faces drawn in index order, GL_LESS).
 * key = depth24 << 32 | face index, minimum wins  =>  on equal quantised depth the lowest
 * face index wins, which is what in-order drawing with GL_LESS produces.
 * ------------------------------------------------------------------------------------ */
static void forward_element(int H, int W, int C, int V, int F,
                            const float *bg, const float *verts, const float *cols,
                            const int32_t *faces,
                            float *pixels, int32_t *face_id, float *gbuf,
                            uint64_t *key, float *samp /* [H*W*4] scratch */)
{
    const size_t npix = (size_t)H * W;
    for (size_t i = 0; i < npix; ++i) key[i] = KEY_EMPTY;
    for (int f = 0; f < F; ++f) {
        int32_t i0 = faces[3 * f], i1 = faces[3 * f + 1], i2 = faces[3 * f + 2];
        if (i0 < 0 || i1 < 0 || i2 < 0 || i0 >= V || i1 >= V || i2 >= V) continue;
        ClipVert poly[6];
        int n = clip_face(verts + 4 * (size_t)i0, verts + 4 * (size_t)i1, verts + 4 * (size_t)i2, poly);
        for (int s = 1; s + 1 < n; ++s) {           /* triangle fan of the clipped polygon */
            SubTri t;
            if (!setup_subtri(&poly[0], &poly[s], &poly[s + 1], W, H, &t)) continue;
            for (int py = t.ymin; py <= t.ymax; ++py)
                for (int px = t.xmin; px <= t.xmax; ++px) {
                    Sample sm;
                    if (!sample_subtri(&t, px, py, &sm)) continue;
                    uint32_t d = depth24(sm.zw);
                    if (d >= DEPTH_MAX) continue;                 /* GL_LESS against cleared 1.0 */
                    uint64_t k = ((uint64_t)d << 32) | (uint32_t)f;
                    size_t row = (size_t)(H - 1 - py);            /* y flip, rasterise_egl.cu:93 */
                    size_t o = row * W + px;
                    if (k < key[o]) {
                        key[o] = k;
                        samp[4 * o + 0] = sm.B0; samp[4 * o + 1] = sm.B1;
                        samp[4 * o + 2] = sm.w_clip;
                        /* the window depth as the 24-bit depth buffer holds it */
                        samp[4 * o + 3] = (float)d * (1.0f / 16777215.0f);
                    }
                }
        }
    }
    for (size_t o = 0; o < npix; ++o) {
        if (key[o] == KEY_EMPTY) {
            /* colour buffer is never cleared: background shows through (:370-377) */
            for (int c = 0; c < C; ++c) pixels[o * C + c] = bg[o * C + c];
            if (face_id) face_id[o] = -1;
            if (gbuf) { gbuf[4 * o] = 0.f; gbuf[4 * o + 1] = 0.f; gbuf[4 * o + 2] = 0.f; gbuf[4 * o + 3] = 1.0f; }
            continue;
        }
        int32_t f = (int32_t)(key[o] & 0xFFFFFFFFu);
        const int32_t *fi = faces + 3 * (size_t)f;
        float B0 = samp[4 * o], B1 = samp[4 * o + 1];
        for (int c = 0; c < C; ++c) {
            float c0 = cols[(size_t)fi[0] * C + c], c1 = cols[(size_t)fi[1] * C + c], c2 = cols[(size_t)fi[2] * C + c];
            /* smooth varying (csrc/shaders.cpp:23,29); written so that constant colours are exact */
            float t0 = c0 - c2, t1 = c1 - c2;
            pixels[o * C + c] = (c2 + B0 * t0) + B1 * t1;
        }
        if (face_id) face_id[o] = f;
        if (gbuf) memcpy(gbuf + 4 * o, samp + 4 * o, 4 * sizeof(float));
    }
}

int golden_rasterise_forward(int B, int H, int W, int C, int V, int F,
                             const float *background, int64_t bg_bs,
                             const float *vertices, int64_t v_bs,
                             const float *colors, int64_t c_bs,
                             const int32_t *faces, int64_t f_bs,
                             float *pixels, int32_t *face_id, float *gbuf, int threads)
{
    if (B < 0 || H <= 0 || W <= 0 || C <= 0 || V < 0 || F < 0) return -1;
    const size_t npix = (size_t)H * W;
    int err = 0;
    (void)threads;
#ifdef _OPENMP
#pragma omp parallel num_threads(threads > 0 ? threads : 1)
#endif
    {
        uint64_t *key = (uint64_t *)malloc(npix * sizeof(uint64_t));
        float *samp = (float *)malloc(npix * 4 * sizeof(float));
        if (!key || !samp) {
            err = 1;
        } else {
#ifdef _OPENMP
#pragma omp for schedule(dynamic, 1)
#endif
            for (int b = 0; b < B; ++b)
                forward_element(H, W, C, V, F,
                                background + (size_t)b * bg_bs, vertices + (size_t)b * v_bs,
                                colors + (size_t)b * c_bs, faces + (size_t)b * f_bs,
                                pixels + (size_t)b * npix * C,
                                face_id ? face_id + (size_t)b * npix : NULL,
                                gbuf ? gbuf + (size_t)b * npix * 4 : NULL, key, samp);
        }
        free(key); free(samp);
    }
    return err ? -2 : 0;
}

/* ------------------------------------------------------------------------------------ *
 * Backward (docs/GRADIENT_SPEC.md).  Inputs are exactly the per-pixel state the reference's
 * vestigial backward design makes available (csrc/rasterise_grad_common.h:19-24,
 * csrc/shaders.cpp:2187-2221): barycentrics, clip-space w and the face's vertex indices per
 * pixel, plus `pixels` and `grad_pixels`.  Accumulation is in double; results are cast to
 * float at the end.
 * ------------------------------------------------------------------------------------ */
static float finite_or_zero(float v) { return fabsf(v) <= FLT_MAX ? v : 0.0f; }

static int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

/* Pick the pixel whose surface "owns" the image change seen at pixel p along one axis:
 * a covered neighbour (m or q) on a different face with strictly smaller window depth is an
 * occluder and takes ownership; the nearer one wins, the lower-coordinate one on a tie. */
static long pick_owner(const int32_t *face_id, const float *gbuf, long p, long m, long q)
{
    int32_t fp = face_id[p];
    float zp = fp >= 0 ? gbuf[4 * p + 3] : 2.0f;       /* uncovered = farther than anything */
    long best = fp >= 0 ? p : -1;
    float zbest = zp;
    long cand[2] = {m, q};
    for (int i = 0; i < 2; ++i) {
        long c = cand[i];
        if (c < 0 || c == p) continue;
        int32_t fc = face_id[c];
        if (fc < 0 || fc == fp) continue;
        float zc = gbuf[4 * c + 3];
        if (zc < zp && zc < zbest) { best = c; zbest = zc; }
    }
    return best;
}

static void backward_element(int H, int W, int C, int V,
                             const int32_t *faces, const float *pixels, const float *grad_pixels,
                             const int32_t *face_id, const float *gbuf,
                             double *gv /* [V,4] */, double *gc /* [V,C] */, float *grad_bg)
{
    (void)V;
    const float halfW = 0.5f * (float)W, halfH = 0.5f * (float)H;
    for (int r = 0; r < H; ++r)
        for (int x = 0; x < W; ++x) {
            long p = (long)r * W + x;
            const float *g = grad_pixels + p * C;
            int32_t fp = face_id[p];
            /* (1) exact parts: colours and background */
            if (fp < 0) {
                for (int c = 0; c < C; ++c) grad_bg[p * C + c] = g[c];
            } else {
                const int32_t *fi = faces + 3 * (size_t)fp;
                float B0 = gbuf[4 * p], B1 = gbuf[4 * p + 1], B2 = (1.0f - B0) - B1;
                for (int c = 0; c < C; ++c) {
                    grad_bg[p * C + c] = 0.0f;
                    gc[(size_t)fi[0] * C + c] += (double)(B0 * g[c]);
                    gc[(size_t)fi[1] * C + c] += (double)(B1 * g[c]);
                    gc[(size_t)fi[2] * C + c] += (double)(B2 * g[c]);
                }
            }
            /* (2) filter-based position part: Sobel response of `pixels`, replicate border */
            int xm = clampi(x - 1, 0, W - 1), xp = clampi(x + 1, 0, W - 1);
            int rm = clampi(r - 1, 0, H - 1), rp = clampi(r + 1, 0, H - 1);
            float sx = 0.0f, sr = 0.0f;
            for (int c = 0; c < C; ++c) {
#define PX(rr, xx) pixels[((long)(rr) * W + (xx)) * C + c]
                float dx = ((PX(rm, xp) - PX(rm, xm)) + 2.0f * (PX(r, xp) - PX(r, xm)) + (PX(rp, xp) - PX(rp, xm))) * 0.125f;
                float dr = ((PX(rp, xm) - PX(rm, xm)) + 2.0f * (PX(rp, x) - PX(rm, x)) + (PX(rp, xp) - PX(rm, xp))) * 0.125f;
#undef PX
                sx += g[c] * finite_or_zero(dx);
                sr += g[c] * finite_or_zero(dr);
            }
            for (int axis = 0; axis < 2; ++axis) {
                float s = axis == 0 ? sx : sr;
                if (s == 0.0f) continue;
                long m = axis == 0 ? (x > 0 ? p - 1 : -1) : (r > 0 ? p - W : -1);
                long q = axis == 0 ? (x < W - 1 ? p + 1 : -1) : (r < H - 1 ? p + W : -1);
                long o = pick_owner(face_id, gbuf, p, m, q);
                if (o < 0) continue;
                const int32_t *fi = faces + 3 * (size_t)face_id[o];
                float B[3];
                B[0] = gbuf[4 * o]; B[1] = gbuf[4 * o + 1]; B[2] = (1.0f - B[0]) - B[1];
                float wc = gbuf[4 * o + 2];
                int ox = (int)(o % W), orow = (int)(o / W);
                /* NDC of the owner's pixel centre; window y is up, rows go down */
                float xn = ((float)ox + 0.5f) / halfW - 1.0f;
                float yn = ((float)(H - 1 - orow) + 0.5f) / halfH - 1.0f;
                for (int k = 0; k < 3; ++k) {
                    double *d = gv + 4 * (size_t)fi[k];
                    if (axis == 0) {
                        /* moving vertex k by dx_clip shifts the surface by B_k*dx*W/(2 w) px in +x;
                         * a shift of +1 px changes the pixel by -dI/dx */
                        float a = s * B[k] * halfW / wc;
                        d[0] += (double)(-a);
                        d[3] += (double)(a * xn);
                    } else {
                        /* +y_clip moves the surface UP = towards smaller rows */
                        float a = s * B[k] * halfH / wc;
                        d[1] += (double)a;
                        d[3] += (double)(-a * yn);
                    }
                }
            }
        }
}

int golden_rasterise_backward(int B, int H, int W, int C, int V, int F,
                              const int32_t *faces, int64_t f_bs,
                              const float *pixels, const float *grad_pixels,
                              const int32_t *face_id, const float *gbuf,
                              float *grad_vertices, int64_t gv_bs,
                              float *grad_colors, int64_t gc_bs,
                              float *grad_background, int threads)
{
    (void)F;
    if (B < 0 || H <= 0 || W <= 0 || C <= 0 || V < 0) return -1;
    const size_t npix = (size_t)H * W;
    /* one double accumulator per batch element, folded afterwards according to the strides
     * (stride 0 = shared mesh: gradients of all elements sum into one [V,*] block) */
    double *gv = (double *)calloc((size_t)B * V * 4 + 1, sizeof(double));
    double *gc = (double *)calloc((size_t)B * V * C + 1, sizeof(double));
    if (!gv || !gc) { free(gv); free(gc); return -2; }
    (void)threads;
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic, 1) num_threads(threads > 0 ? threads : 1)
#endif
    for (int b = 0; b < B; ++b)
        backward_element(H, W, C, V, faces + (size_t)b * f_bs,
                         pixels + (size_t)b * npix * C, grad_pixels + (size_t)b * npix * C,
                         face_id + (size_t)b * npix, gbuf + (size_t)b * npix * 4,
                         gv + (size_t)b * V * 4, gc + (size_t)b * V * C,
                         grad_background + (size_t)b * npix * C);
    int nbv = gv_bs == 0 ? 1 : B, nbc = gc_bs == 0 ? 1 : B;
    for (int b = 0; b < nbv; ++b)
        for (size_t i = 0; i < (size_t)V * 4; ++i) {
            double s = 0.0;
            if (gv_bs == 0) for (int bb = 0; bb < B; ++bb) s += gv[(size_t)bb * V * 4 + i];
            else s = gv[(size_t)b * V * 4 + i];
            grad_vertices[(size_t)b * gv_bs + i] = (float)s;
        }
    for (int b = 0; b < nbc; ++b)
        for (size_t i = 0; i < (size_t)V * C; ++i) {
            double s = 0.0;
            if (gc_bs == 0) for (int bb = 0; bb < B; ++bb) s += gc[(size_t)bb * V * C + i];
            else s = gc[(size_t)b * V * C + i];
            grad_colors[(size_t)b * gc_bs + i] = (float)s;
        }
    free(gv); free(gc);
    return 0;
}

int golden_num_threads(void)
{
#ifdef _OPENMP
    return omp_get_num_procs();      /* not omp_get_max_threads(): torchrun exports OMP_NUM_THREADS=1 */
#else
    return 1;
#endif
}
