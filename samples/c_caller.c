/* A plain C99 caller of libdirt_b200.so through the host-buffer entry points: renders the reference
 * README's square (README.md:46-70: 128 x 128, one channel, a 16-pixel square centred on (32, 64)),
 * checks the 256 ones, then back-propagates an all-ones image gradient and prints the gradient of
 * the four corner colours (they sum to the 256 covered pixels).  No CUDA headers are needed: the
 * library stages host buffers itself.
 *
 *   gcc -std=c99 -Iinclude samples/c_caller.c -Ldirt_b200 -ldirt_b200 -Wl,-rpath,$PWD/dirt_b200 -o c_caller
 */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "dirt_b200.h"

enum { W = 128, H = 128, V = 4, F = 2 };

int main(void)
{
    static float background[H * W], pixels[H * W], gbuffer[H * W * 4], grad_pixels[H * W], grad_background[H * W];
    static int32_t face_id[H * W];
    float vertices[V * 4], colours[V] = {1.f, 1.f, 1.f, 1.f}, grad_vertices[V * 4], grad_colours[V];
    const int32_t faces[F * 3] = {0, 1, 2, 0, 2, 3};
    const float cx = 32.f, cy = 64.f, side = 16.f;
    const float corner[V][2] = {{0.f, 0.f}, {0.f, 1.f}, {1.f, 1.f}, {1.f, 0.f}};
    for (int k = 0; k < V; ++k) {
        vertices[4 * k + 0] = (corner[k][0] * side - side / 2.f + cx) * 2.f / W - 1.f;
        vertices[4 * k + 1] = (corner[k][1] * side - side / 2.f + cy) * 2.f / H - 1.f;
        vertices[4 * k + 2] = 0.f;
        vertices[4 * k + 3] = 1.f;
    }
    if (dirt_abi_version() != DIRT_B200_ABI_VERSION) {
        fprintf(stderr, "header / library mismatch\n");
        return 2;
    }

    DirtForwardArgs fwd;
    memset(&fwd, 0, sizeof fwd);
    fwd.shape.batch = 1; fwd.shape.height = H; fwd.shape.width = W; fwd.shape.channels = 1;
    fwd.shape.vertex_count = V; fwd.shape.face_count = F;
    fwd.background = background;  fwd.background_bstride = (int64_t)H * W;
    fwd.vertices = vertices;      fwd.vertices_bstride = V * 4;
    fwd.vertex_colors = colours;  fwd.vertex_colors_bstride = V;
    fwd.faces = faces;            fwd.faces_bstride = F * 3;
    fwd.pixels = pixels; fwd.face_id = face_id; fwd.gbuffer = gbuffer;
    if (dirt_rasterise_forward_host(&fwd) != DIRT_STATUS_OK) {
        fprintf(stderr, "forward: %s\n", dirt_last_error());
        return 1;
    }
    int ones = 0, wrong = 0;
    for (int r = 0; r < H; ++r)
        for (int x = 0; x < W; ++x) {
            const int inside = r >= 56 && r < 72 && x >= 24 && x < 40;
            ones += pixels[r * W + x] == 1.f;
            wrong += pixels[r * W + x] != (inside ? 1.f : 0.f);
        }
    printf("covered pixels: %d (expected 256), mismatches: %d\n", ones, wrong);

    DirtBackwardArgs bwd;
    memset(&bwd, 0, sizeof bwd);
    bwd.shape = fwd.shape;
    bwd.faces = faces; bwd.faces_bstride = F * 3;
    bwd.pixels = pixels; bwd.grad_pixels = grad_pixels; bwd.face_id = face_id; bwd.gbuffer = gbuffer;
    bwd.grad_vertices = grad_vertices;        bwd.grad_vertices_bstride = V * 4;
    bwd.grad_vertex_colors = grad_colours;    bwd.grad_vertex_colors_bstride = V;
    bwd.grad_background = grad_background;
    for (int i = 0; i < H * W; ++i) grad_pixels[i] = 1.f;
    if (dirt_rasterise_backward_host(&bwd) != DIRT_STATUS_OK) {
        fprintf(stderr, "backward: %s\n", dirt_last_error());
        return 1;
    }
    /* the same backward in its stateless form: no face_id / gbuffer from the forward, visibility is re-derived
     * from the vertices (what the TF op pair does: RasteriseGrad receives the forward's inputs) */
    static float grad_colours2[4], grad_vertices2[16];
    DirtBackwardArgs bwd2 = bwd;
    bwd2.face_id = NULL; bwd2.gbuffer = NULL;
    bwd2.vertices = vertices; bwd2.vertices_bstride = V * 4;
    bwd2.grad_vertices = grad_vertices2; bwd2.grad_vertex_colors = grad_colours2;
    if (dirt_rasterise_backward_host(&bwd2) != DIRT_STATUS_OK) {
        fprintf(stderr, "stateless backward: %s\n", dirt_last_error());
        return 1;
    }
    int same = 1;       /* equal up to the summation order of the fp32 atomics */
    for (int k = 0; k < V; ++k) same = same && fabsf(grad_colours2[k] - grad_colours[k]) <= 1e-3f;
    for (int k = 0; k < 4 * V; ++k) same = same && fabsf(grad_vertices2[k] - grad_vertices[k]) <= 1e-3f * (1.f + fabsf(grad_vertices[k]));
    printf("stateless backward %s the one fed with the forward's buffers\n", same ? "matches" : "DIFFERS FROM");
    if (!same) return 1;
    float sum = 0.f;
    for (int k = 0; k < V; ++k) sum += grad_colours[k];
    printf("grad_vertex_colors: %.3f %.3f %.3f %.3f (sum %.3f, expected 256)\n",
           grad_colours[0], grad_colours[1], grad_colours[2], grad_colours[3], sum);
    dirt_release_host_staging();
    return (ones == 256 && wrong == 0 && sum > 255.9f && sum < 256.1f) ? 0 : 1;
}
