/*
 * dirt_b200.h -- C ABI of libdirt_b200.so: the `Rasterise` forward op and the `RasteriseGrad`
 * backward of yyuting/dirt as hand-written sm_100a CUDA kernels.
 *
 * This is the drop-in boundary: plain pointers and sizes, no TensorFlow / torch types.  Each
 * entry point names the reference interface it replaces (paths relative to the reference
 * tree).  The reference-side binding a maintainer would add is shown in INTEGRATION.md.
 *
 * Conventions (identical to the reference op, csrc/rasterise_egl.cpp:33-53, :309-339):
 *   - all tensors dense, row-major, fp32 / int32;
 *   - background, pixels: [B, H, W, C], row 0 = TOP of the image (csrc/rasterise_egl.cu:29, :93);
 *   - vertices: [B, V, 4] OpenGL clip space; vertex_colors: [B, V, C]; faces: [B, F, 3] int32,
 *     indices local to the batch element (csrc/rasterise_egl.cpp:451-457);
 *   - per-input batch strides are given in ELEMENTS; 0 broadcasts one [*, ...] block over the
 *     batch (the relaxation the reference wishes for at csrc/rasterise_egl.cpp:334).
 *
 * Device-pointer entry points are asynchronous on the caller's stream: no host
 * synchronisation, no device->host copy, no threads (the reference blocks on a GL thread and a
 * synchronous cudaMemcpy, csrc/gl_dispatcher.h:58-67, csrc/rasterise_egl.cpp:323).
 * There is no CPU fallback: without a CUDA device every compute entry point returns
 * DIRT_STATUS_CUDA_ERROR (the reference registers a DEVICE_GPU kernel only, :520).
 */
#ifndef DIRT_B200_H
#define DIRT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define DIRT_API __attribute__((visibility("default")))
#else
#define DIRT_API
#endif

#define DIRT_B200_ABI_VERSION 2          /* 2: DirtForwardArgs.visibility_hint, DirtBackwardArgs recompute inputs */
#define DIRT_MAX_CHANNELS 16          /* reference: channels in {1,3} (csrc/hwc.h:27); wider
                                         attribute sets serve deferred G-buffers in one call */

typedef enum DirtStatus {
    DIRT_STATUS_OK = 0,
    DIRT_STATUS_INVALID_ARGUMENT = -1,   /* replaces OP_REQUIRES -> InvalidArgument, rasterise_egl.cpp:310-336 */
    DIRT_STATUS_CUDA_ERROR = -2,         /* replaces LOG(FATAL) aborts, rasterise_egl.cpp:85-109 */
    DIRT_STATUS_WORKSPACE_TOO_SMALL = -3,
    DIRT_STATUS_UNSUPPORTED = -4
} DirtStatus;

/* Problem shape: the op attrs height/width/channels (csrc/rasterise_egl.cpp:34-36, csrc/hwc.h:21-30)
 * plus the dynamic dims the reference reads from its input tensors (:313-336). */
typedef struct DirtShape {
    int32_t batch;        /* B  = vertices.dim0                          */
    int32_t height;       /* H  > 0                                       */
    int32_t width;        /* W  > 0                                       */
    int32_t channels;     /* C  in [1, DIRT_MAX_CHANNELS]                 */
    int32_t vertex_count; /* V                                            */
    int32_t face_count;   /* F                                            */
} DirtShape;

/* Forward arguments.  Replaces the inputs/outputs of RasteriseOpGpu::Compute
 * (csrc/rasterise_egl.cpp:284-514).  `camera_pos` (input 4 of the fork's op, :41) only
 * parameterises the fork's procedural fragment shaders (:399-433) and is not an input of the
 * mesh path; adapters accept and drop it.
 * `background` is read only at pixels no face covers, so it may also point to PINNED HOST memory
 * mapped into the device address space (cudaHostAlloc / cudaHostRegister under unified
 * addressing): the kernel then fetches just those pixels over PCIe, where the reference uploads
 * the whole image first (upload_background, csrc/rasterise_egl.cu:16-50). */
typedef struct DirtForwardArgs {
    DirtShape shape;
    const float *background;      int64_t background_bstride;    /* [B,H,W,C]              */
    const float *vertices;        int64_t vertices_bstride;      /* [B,V,4]                */
    const float *vertex_colors;   int64_t vertex_colors_bstride; /* [B,V,C]                */
    const int32_t *faces;         int64_t faces_bstride;         /* [B,F,3]                */
    float *pixels;                /* out [B,H,W,C], required                               */
    /* Optional per-pixel state for the backward pass (both NULL, or both set).  This is the
     * G-buffer the reference's backward shaders render (csrc/shaders.cpp:2187-2221):
     * face_id -> (i0,i1,i2) through `faces`; gbuffer = (b0, b1, clip-space w, window depth). */
    int32_t *face_id;             /* out [B,H,W], -1 = background                          */
    float *gbuffer;               /* out [B,H,W,4]                                         */
    void *workspace;              /* device scratch, >= dirt_workspace_bytes(&shape)       */
    size_t workspace_bytes;
    /* Optional visibility hint: caller-owned DEVICE memory of >= dirt_visibility_hint_bytes(&shape)
     * bytes that PERSISTS between calls which render the same mesh (same B and F) with slowly
     * changing vertices -- the inverse-rendering loop of tests/optimize_sequence.py:293-352.  Every
     * call records which faces won a pixel; the next call samples those faces first and tests all
     * others against the nearly final depth buffer they leave.  The hint only ORDERS work: the
     * output is bit-identical with a fresh, a stale or a wrong hint, or none (NULL).  Zero it once
     * after allocating (any content is safe, zero merely avoids a slow first call); one buffer per
     * mesh and per stream, not shared between calls in flight.  Used by meshes of >= 4096 faces.
     * Measured on the 100k-face sphere workloads: 2 % faster at 512 x 512, 5 % slower at 256 x 256 than
     * the depth-ordered plan the library uses without a hint: an option, not a
     * recommendation. */
    void *visibility_hint;
    size_t visibility_hint_bytes;
} DirtForwardArgs;

/* Backward arguments.  Replaces launch_grad_assembly (declared, never defined, in the fork:
 * csrc/rasterise_grad_common.h:19-24): same outputs (grad_vertices, grad_vertex_colors,
 * grad_background), same inputs (G-buffer, pixels, grad_pixels) minus the debug tensor.
 * Gradient outputs are zero-initialised by the callee.  A 0 batch stride on an output sums the
 * gradients of all batch elements into one [V,*] block (shared mesh). */
typedef struct DirtBackwardArgs {
    DirtShape shape;
    const int32_t *faces;         int64_t faces_bstride;         /* [B,F,3]                */
    const float *pixels;          /* [B,H,W,C] forward output                              */
    const float *grad_pixels;     /* [B,H,W,C]                                             */
    const int32_t *face_id;       /* [B,H,W]   from forward                                */
    const float *gbuffer;         /* [B,H,W,4] from forward                                */
    float *grad_vertices;         int64_t grad_vertices_bstride;      /* out [B,V,4]       */
    float *grad_vertex_colors;    int64_t grad_vertex_colors_bstride; /* out [B,V,C]       */
    float *grad_background;       /* out [B,H,W,C]; may be NULL                            */
    /* Stateless mode, as the reference's graph sees the op pair (RasteriseGrad receives the forward's
     * INPUTS, csrc/rasterise_grad_egl.cpp:38-42, and re-renders what it needs): with face_id == NULL
     * and gbuffer == NULL the backward re-derives both from `vertices` and `faces` with the forward's
     * own kernels into `workspace` (>= dirt_backward_workspace_bytes(&shape)) -- identical values, no
     * buffer carried from the forward.  Ignored when face_id / gbuffer are given. */
    const float *vertices;        int64_t vertices_bstride;      /* [B,V,4] clip space     */
    void *workspace;              size_t workspace_bytes;
    /* DIRT_BACKWARD_DETERMINISTIC: bit-identical gradients from run to run (the default sums with fp32
     * atomics, whose order varies: reproducible to ~1e-6 relative only).  A debugging mode, about three
     * times slower; needs `workspace` of dirt_deterministic_workspace_bytes(&shape) more bytes (after the
     * dirt_backward_workspace_bytes of the stateless mode, when both are used). */
    uint32_t flags;
} DirtBackwardArgs;
#define DIRT_BACKWARD_DETERMINISTIC 1u

/* Shared-mesh backward: every batch element renders the SAME mesh, element b
 * through its own matrix (clip_b = world @ M_b, row vectors: the convention of dirt/matrices.py and
 * the inverse-rendering loop of tests/optimize_sequence.py:293-352).  The clip-space vertex
 * gradient of each element is folded to world space (grad_world += grad_clip_b @ M_b^T) and
 * accumulated, together with the colour gradient, straight into ONE packed buffer
 *     grad_shared [V, stride],  stride = dirt_shared_grad_stride(C) = 4 + C rounded up to 4:
 *     columns 0..3 world-space position gradient, 4..4+C-1 colour gradient, rest zero,
 * which is also the message of the cross-GPU all-reduce.  Neither [B,V,4] nor [B,V,C] exists. */
typedef struct DirtSharedBackwardArgs {
    DirtShape shape;
    const int32_t *faces;         int64_t faces_bstride;         /* [B,F,3] (0 = shared)   */
    const float *pixels;          /* [B,H,W,C] forward output                              */
    const float *grad_pixels;     /* [B,H,W,C]                                             */
    const int32_t *face_id;       /* [B,H,W]   from forward                                */
    const float *gbuffer;         /* [B,H,W,4] from forward                                */
    const float *clip_from_world; /* [B,4,4] row-major M_b                                 */
    float *grad_shared;           /* out [V, stride], 16-byte aligned, zeroed by the callee */
    float *grad_background;       /* out [B,H,W,C]; may be NULL                            */
    void *workspace;              size_t workspace_bytes;        /* only for DIRT_BACKWARD_DETERMINISTIC */
    uint32_t flags;
} DirtSharedBackwardArgs;
DIRT_API int dirt_shared_grad_stride(int channels);
DIRT_API int dirt_rasterise_backward_shared(const DirtSharedBackwardArgs *args, void *stream);

/* The sum of `grad_shared` over the GPUs of one box, taken over peer memory in one kernel (no counterpart in the reference,
 * whose loop runs on one device): every rank's buffer must be mapped into every process -- symmetric memory,
 * cudaIpc or the VMM API over NVLink.  Rank r sums slice r of all buffers in rank order (all ranks end up with identical
 * bits) and stores it into slice r of every buffer; with `multicast` (an NVSwitch multicast mapping of the same buffers) the
 * switch does the additions (multimem.ld_reduce / multimem.st).  Two handshakes on flag words bracket the sums.  In place,
 * asynchronous on `stream`.  Every rank of the group must make the call, once per reduction. */
typedef struct DirtPeerAllreduceArgs {
    float *const *buffers;        /* DEVICE array [world]: rank p's buffer as this process addresses it; 16-byte aligned   */
    uint32_t *const *signals;     /* DEVICE array [world]: rank p's flag pad, >= 2 * world words, zeroed once before use  */
    int32_t rank, world;          /* 1 <= world <= 32                                                                      */
    int64_t count;                /* floats per buffer, a multiple of 4                                                    */
    uint32_t *epoch;              /* DEVICE, FOUR words of this rank, zeroed once before use (together with the pads):
                                     [0] calls completed; [1] stays 0, or 1 + the rank a handshake gave up waiting for
                                     (~30 s: that peer never made its call; the buffer's contents are then undefined);
                                     [2], [3] internal                                                                     */
    float *multicast;             /* optional: multicast address of the buffers (all ranks' copies), or NULL               */
} DirtPeerAllreduceArgs;
DIRT_API int dirt_allreduce_peer(const DirtPeerAllreduceArgs *args, void *stream);

/* Vertex normals of `batch` meshes with common topology: replaces
 * vertex_normals / vertex_normals_pre_split, dirt/lighting.py:34-93, :101-133 -- the normalised
 * sum, per vertex, of the unit normals (v1 - v0) x (v2 - v0) of the faces that use it, with the
 * reference's + 1e-12 in both normalisations.  vertices [batch, V, D], D = 3 or 4 (w ignored);
 * faces [F,3] shared by the batch; sums [batch,V,3] (un-normalised, kept for the backward) and
 * normals [batch,V,3] are outputs.  Faces with an index outside [0,V) are skipped. */
DIRT_API int dirt_vertex_normals(const float *vertices, int batch, int vertex_count, int vertex_dim, const int32_t *faces,
                                 int face_count, float *sums, float *normals, void *stream);
/* Its backward: grad_normals [batch,V,3] -> grad_vertices [batch,V,D] (zero-initialised by the
 * callee; the w column stays zero).  grad_sums [batch,V,3] is scratch.  With sums == NULL the
 * final normalisation is left out: grad_normals is taken as the gradient w.r.t. the `sums`
 * output (the un-renormalised normals of vertex_normals_pre_split). */
DIRT_API int dirt_vertex_normals_grad(const float *vertices, int batch, int vertex_count, int vertex_dim, const int32_t *faces,
                                      int face_count, const float *sums, const float *grad_normals, float *grad_sums,
                                      float *grad_vertices, void *stream);

/* Lambertian shading under one directional light, forward and backward, each ONE element-wise pass: replaces the
 * framework ops (and their autodiff) of diffuse_directional, dirt/lighting.py:182-225:
 *     out[b,v,c] = light_color[b,c] * colours[b,v,c] * f(normals[b,v] . (-light_direction[b])),
 *     f = |.| when double_sided, else max(., 0).
 * normals [batch,V,3], colours / out / grad_out / grad_colours [batch,V,C], grad_normals [batch,V,3]; light_direction
 * and light_color are DEVICE pointers, `*_bstride` elements apart per batch element (0: one light for all).
 * grad_normals / grad_colours may be NULL (not wanted).  Gradients w.r.t. the light itself are left to the caller. */
DIRT_API int dirt_diffuse_directional(const float *normals, const float *colours, const float *light_direction,
                                      int64_t light_direction_bstride, const float *light_color, int64_t light_color_bstride,
                                      int double_sided, int batch, int vertex_count, int channels, float *out, void *stream);
DIRT_API int dirt_diffuse_directional_grad(const float *normals, const float *colours, const float *light_direction,
                                           int64_t light_direction_bstride, const float *light_color, int64_t light_color_bstride,
                                           int double_sided, int batch, int vertex_count, int channels, const float *grad_out,
                                           float *grad_normals, float *grad_colours, void *stream);

/* The other two per-vertex shading helpers of dirt/lighting.py, each ONE element-wise pass forward and ONE backward:
 *   dirt_diffuse_point         Lambertian reflectance under one point light (dirt/lighting.py:291-343):
 *                              inc = (p - light_position) / (|p - light_position| + 1e-12),
 *                              out = light_color * colours * f(normals . inc)
 *   dirt_specular_directional  Phong reflectance under one directional light (dirt/lighting.py:228-288), with the
 *                              reference's operator precedence (1e-12 is added AFTER the division by the norm):
 *                              L = -light_direction, R = -L + 2 (n . L) n, T = (camera - p) / |camera - p| + 1e-12,
 *                              out = light_color * reflectivities * f(T . R) ^ shininess
 * f = |.| when double_sided, else max(., 0).  positions / normals [batch,V,3]; colours (vertex_colors resp.
 * vertex_reflectivities), out, grad_out, grad_colours [batch,V,C]; light_vector = light_position resp. light_direction
 * [*,3]; light_color [*,C]; camera_position [*,3] and shininess [*] (specular only): DEVICE pointers, `*_bstride`
 * elements apart per batch element (0: one for all).  The *_grad calls read grad_out and write the gradients that are
 * not NULL; gradients w.r.t. lights, camera and shininess are left to the caller. */
typedef struct DirtVertexShadeArgs {
    int batch, vertex_count, channels, double_sided;
    const float *positions, *normals, *colours;
    const float *light_vector;      int64_t light_vector_bstride;
    const float *light_color;       int64_t light_color_bstride;
    const float *camera_position;   int64_t camera_position_bstride;
    const float *shininess;         int64_t shininess_bstride;
    float *out;
    const float *grad_out;
    float *grad_positions, *grad_normals, *grad_colours;
} DirtVertexShadeArgs;
DIRT_API int dirt_diffuse_point(const DirtVertexShadeArgs *args, void *stream);
DIRT_API int dirt_diffuse_point_grad(const DirtVertexShadeArgs *args, void *stream);
DIRT_API int dirt_specular_directional(const DirtVertexShadeArgs *args, void *stream);
DIRT_API int dirt_specular_directional_grad(const DirtVertexShadeArgs *args, void *stream);

/* Library identity. */
DIRT_API int dirt_abi_version(void);
/* Thread-local message of the last failing call on this thread ("" if none). */
DIRT_API const char *dirt_last_error(void);

/* Scratch the forward pass needs for `shape` (depth|face-id keys, projected vertices, tile counters
 * and queues, large-triangle records).  Replaces the grow-only GL framebuffer / buffers the reference
 * keeps per GL thread (csrc/rasterise_egl.cpp:73-89, :199-274, :347-354); here the caller
 * owns it, so torch's / TF's caching allocator can. */
DIRT_API size_t dirt_workspace_bytes(const DirtShape *shape);

/* Scratch of the backward's stateless mode (face_id == gbuffer == NULL): the forward's workspace plus
 * the two per-pixel buffers it re-derives. */
DIRT_API size_t dirt_backward_workspace_bytes(const DirtShape *shape);
/* Extra scratch of DIRT_BACKWARD_DETERMINISTIC (64-bit fixed-point accumulators, one per gradient value). */
DIRT_API size_t dirt_deterministic_workspace_bytes(const DirtShape *shape);
/* Convenience for callers without an allocator of their own (C programs): a workspace of
 * max(dirt_workspace_bytes, dirt_backward_workspace_bytes)(shape) bytes on the current device, and its
 * release.  Counterpart of the reference's per-thread grow-only buffers (csrc/rasterise_egl.cpp:73-89). */
DIRT_API int dirt_workspace_create(const DirtShape *shape, void **workspace, size_t *workspace_bytes);
DIRT_API int dirt_workspace_destroy(void *workspace);

/* Size of DirtForwardArgs.visibility_hint for `shape` (one byte per face and batch element plus a
 * 256-byte header); 0 for an invalid shape.  No reference counterpart: the reference's GL driver
 * keeps no state between draws either (csrc/rasterise_egl.cpp:440-458 redraws from scratch). */
DIRT_API size_t dirt_visibility_hint_bytes(const DirtShape *shape);

/* Forward: replaces op "Rasterise" (REGISTER_OP csrc/rasterise_egl.cpp:33-53; kernel :284-514;
 * upload_background / download_pixels csrc/rasterise_egl.cu:16-129 are folded in).
 * `stream` is a cudaStream_t passed as void* (NULL = legacy default stream). */
DIRT_API int dirt_rasterise_forward(const DirtForwardArgs *args, void *stream);

/* Forward without any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  Same arguments and results as dirt_rasterise_forward, except that
 * args->workspace / workspace_bytes are IGNORED: the scratch is a file-scope device array inside the
 * library (DIRT_NOALLOC_WORKSPACE_BYTES, held by every GPU that loads the library until exit).  Shapes
 * whose dirt_workspace_bytes() exceeds it are rejected with DIRT_STATUS_UNSUPPORTED and a message; the
 * call itself issues no cudaMalloc / cudaFree.  All calls on a device share the array and it keeps the
 * last call's contents (what must start empty is cleared by every call): calls that overlap on the GPU
 * race -- serialise them, or use dirt_rasterise_forward with a workspace per stream. */
#define DIRT_NOALLOC_WORKSPACE_BYTES (8u << 20)   /* 8 MiB: e.g. one 512x512 frame of a 100k-face mesh */
DIRT_API size_t dirt_noalloc_workspace_limit(void);
DIRT_API int dirt_rasterise_forward_noalloc(const DirtForwardArgs *args, void *stream);

/* Backward: replaces op "RasteriseGrad" in its intended role (csrc/rasterise_grad_egl.cpp:33-53
 * carries the name; csrc/rasterise_grad_common.h:13-24 the intended kernels). */
DIRT_API int dirt_rasterise_backward(const DirtBackwardArgs *args, void *stream);

/* Silhouette dilation of deferred G-buffer channels, fused with its masking: replaces the
 * framework ops of samples/deferred.py:87-91 (mask = x == -inf; 3x3 SAME max-pool; blend).
 * x, out, grad_*: [batch, height, width, channels] device fp32.  The backward sends a background
 * pixel's gradient to the first maximal element (row-major) of its 3x3 window. */
DIRT_API int dirt_dilate_silhouette(const float *x, float *out, int batch, int height, int width, int channels, void *stream);
DIRT_API int dirt_dilate_silhouette_grad(const float *x, const float *grad_out, float *grad_x, int batch, int height,
                                         int width, int channels, void *stream);

/* Procedural scene stage of the fork's own `Rasterise` op: shaders::oceanic_horizon
 * (csrc/shaders.cpp:1668-1919) evaluated on every pixel the rasteriser covered.  Call after
 * dirt_rasterise_forward with face_id / gbuffer requested, on the same stream: covered pixels of
 * `pixels` are overwritten with (sky mask, sun term, 0); uncovered ones keep the background.
 * camera_pos: DEVICE pointer to >= 8 floats (cam_x, cam_y, cam_z, ang1, ang2, ang3, time,
 * light_z -- input 4 of the op, csrc/rasterise_egl.cpp:41, :399-433); no host copy is made. */
typedef struct DirtSceneArgs {
    DirtShape shape;
    const float *background;      int64_t background_bstride;    /* [B,H,W,C]: sample jitter      */
    const float *vertices;        int64_t vertices_bstride;      /* [B,V,4]                       */
    const int32_t *faces;         int64_t faces_bstride;         /* [B,F,3]                       */
    const int32_t *face_id;       /* [B,H,W]   from forward                                       */
    const float *gbuffer;         /* [B,H,W,4] from forward                                       */
    const float *camera_pos;      /* device, >= 8 floats                                          */
    float *pixels;                /* in/out [B,H,W,C]                                             */
} DirtSceneArgs;
DIRT_API int dirt_scene_oceanic_horizon(const DirtSceneArgs *args, void *stream);

/* The fork's other procedural scene ops (dirt/rasterise_ops.py:91-205): same inputs and outputs as
 * `Rasterise`, a different fragment program per op.  One entry point, selected by `scene`:
 *   OCEANIC              op "RasteriseGrad"      csrc/rasterise_grad_egl.cpp:33, :399  shaders.cpp:556-865
 *   OCEANIC_STILL_CLOUD  op "OceanicStillCloud"  csrc/oceanic_still_cloud.cpp:33, :385 shaders.cpp:866-1175
 *   OCEANIC_NO_CLOUD     op "OceanicNoCloud"     csrc/oceanic_no_cloud.cpp:33, :385    shaders.cpp:1402-1665
 *   OCEANIC_SIMPLE_PROXY op "OceanicSimpleProxy" csrc/oceanic_simple_proxy.cpp:33, :385 shaders.cpp:1921-2185
 *   OCEANIC_OPT_FLOW     op "OceanicOptFlow"     csrc/oceanic_opt_flow.cpp:33, :385    shaders.cpp:1178-1398
 *   HILL                 op "Hill"               csrc/hill.cpp:33, :378                shaders.cpp:123-552
 * Call after dirt_rasterise_forward (face_id / gbuffer requested) on the same stream; covered
 * pixels of `pixels` [B,H,W,shape.channels] receive fragColor.xyz, the rest keep what they hold.
 * `background` is [B,H,W,background_channels] with background_channels in {1,3,4} (the upload
 * kernel's cases, csrc/rasterise_egl.cu:29-48): the sample jitter of the ocean ops, the terrain
 * height + normal texture of HILL (bilinear, clamp to edge, each frame its own texture).
 * camera_pos is a DEVICE pointer to camera_pos_count floats; the count each scene reads is
 * dirt_scene_camera_floats(scene) (8, 8, 9, 8, 8, 16, 12 -- the cudaMemcpy sizes at :323 of each
 * op file); fewer is DIRT_STATUS_INVALID_ARGUMENT.  OPT_FLOW writes the previous-frame pixel
 * coordinates into channels 0 and 1. */
enum {
    DIRT_SCENE_OCEANIC_HORIZON = 0,
    DIRT_SCENE_OCEANIC = 1,
    DIRT_SCENE_OCEANIC_STILL_CLOUD = 2,
    DIRT_SCENE_OCEANIC_NO_CLOUD = 3,
    DIRT_SCENE_OCEANIC_SIMPLE_PROXY = 4,
    DIRT_SCENE_OCEANIC_OPT_FLOW = 5,
    DIRT_SCENE_HILL = 6
};
DIRT_API int dirt_scene_camera_floats(int scene);
DIRT_API int dirt_scene_shade(int scene, const DirtSceneArgs *args, int background_channels, int camera_pos_count, void *stream);

/* Host-buffer convenience calls: every pointer in `args` is HOST memory (pinned or pageable);
 * the library stages inputs to the current device, runs the kernels and copies the outputs
 * back, then synchronises.  workspace / workspace_bytes are ignored (library-owned, grow-only,
 * like the reference's RegisteredBuffer, csrc/rasterise_egl.cpp:55-111).  This is what a
 * host-side caller of the reference effectively pays through session.run feeds/fetches. */
DIRT_API int dirt_rasterise_forward_host(const DirtForwardArgs *args);
DIRT_API int dirt_rasterise_backward_host(const DirtBackwardArgs *args);
/* Frees the library-owned staging buffers of the *_host calls on the current device. */
DIRT_API int dirt_release_host_staging(void);

/* Number of kernel launches issued by this library since load (all threads); lets a
 * benchmark report `gpu_launches` from a count, not a guess. */
DIRT_API uint64_t dirt_kernel_launch_count(void);

/* Per-stage device timing, the counterpart of the reference's compile-time TIME_SECTIONS
 * stamps (csrc/rasterise_egl.cpp:292-294, :489-512), but measured on the device with CUDA
 * events recorded on the caller's stream around each kernel.  Off by default.
 *   dirt_profile_enable(1)  -> subsequent forward/backward calls record events per stage
 *   dirt_profile_collect()  -> call after synchronising the stream: folds the elapsed time of
 *                              every stage recorded since the last collect into the totals
 *   dirt_profile_read()     -> cumulative milliseconds and timed runs per stage (a stage that
 *                              runs several times per call, e.g. one setup launch per depth
 *                              bucket, counts each run)
 * Stage order: see dirt_profile_stage_name(i), i in [0, DIRT_STAGE_COUNT).
 * The timer is ONE per process: profile from a single thread on a single device (the one current at
 * dirt_profile_enable(1)); dirt_profile_enable refuses other devices while it is on. */
#define DIRT_STAGE_COUNT 12
DIRT_API int dirt_profile_enable(int enable);
DIRT_API int dirt_profile_collect(void);
DIRT_API int dirt_profile_read(double *total_ms, uint64_t *calls, int n);
DIRT_API int dirt_profile_reset(void);
DIRT_API const char *dirt_profile_stage_name(int stage);

#ifdef __cplusplus
}
#endif
#endif /* DIRT_B200_H */
