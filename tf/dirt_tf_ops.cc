// dirt_tf_ops.cc -- TensorFlow shim (UNCOMPILED here: no TensorFlow in this image): re-registers ops "Rasterise" and "RasteriseGrad" on top of
// libdirt_b200.so's C ABI (include/dirt_b200.h), with the attrs / input order / output of the
// reference registration (csrc/rasterise_egl.cpp:33-53, :520; csrc/rasterise_grad_egl.cpp:33-53),
// so that `tf.load_op_library` in dirt/rasterise_ops.py:6-7 keeps working unchanged.
//
// NOT BUILT OR TESTED IN THIS REPOSITORY: the build image has no TensorFlow headers.  It is kept
// paper-thin (shape checks are done by the C ABI; this file only unpacks tensors) so that the
// torch-driven tests cover everything but this glue.  Build (on a machine with TF):
//   g++ -std=c++17 -shared -fPIC tf/dirt_tf_ops.cc -Iinclude $(python -c "import tensorflow as tf; \
//       print(' '.join(tf.sysconfig.get_compile_flags() + tf.sysconfig.get_link_flags()))") \
//       -Ldirt_b200 -ldirt_b200 -o librasterise.so
#include "tensorflow/core/framework/op.h"
#include "tensorflow/core/framework/op_kernel.h"
#include "tensorflow/core/framework/shape_inference.h"
#include "tensorflow/core/common_runtime/gpu/gpu_event_mgr.h"
#include "tensorflow/core/platform/stream_executor.h"

#include "dirt_b200.h"

using namespace tensorflow;

namespace {

Status status_from(int rc)
{
    if (rc == DIRT_STATUS_OK) return Status::OK();
    if (rc == DIRT_STATUS_INVALID_ARGUMENT) return errors::InvalidArgument(dirt_last_error());
    return errors::Internal(dirt_last_error());
}

void *cuda_stream_of(OpKernelContext *ctx)
{
    // the reference takes TF's compute stream the same way (csrc/rasterise_egl.cpp:359-361)
    return *reinterpret_cast<void **>(ctx->op_device_context()->stream()->implementation()->GpuStreamMemberHack());
}

DirtShape shape_of(const Tensor &background, const Tensor &vertices, const Tensor &faces, int channels, int64 batch)
{
    DirtShape s;
    s.batch = static_cast<int32_t>(batch);
    s.height = static_cast<int32_t>(background.dim_size(1));
    s.width = static_cast<int32_t>(background.dim_size(2));
    s.channels = channels;
    s.vertex_count = static_cast<int32_t>(vertices.dim_size(1));
    s.face_count = static_cast<int32_t>(faces.dim_size(1));
    return s;
}

// The batch of a call is the largest leading dim of its inputs; every input must have that leading dim or 1
// (1 broadcasts: the relaxation wished for at csrc/rasterise_egl.cpp:334, whose check this generalises).
bool common_batch(std::initializer_list<const Tensor *> ts, int64 *batch)
{
    int64 B = 0;
    for (const Tensor *t : ts) B = std::max<int64>(B, t->dim_size(0));
    for (const Tensor *t : ts)
        if (t->dim_size(0) != B && t->dim_size(0) != 1) return false;
    *batch = B;
    return true;
}

int64_t batch_stride(const Tensor &t, int64 B)
{
    return t.dim_size(0) == 1 && B > 1 ? 0 : t.NumElements() / std::max<int64>(t.dim_size(0), 1);
}

}  // namespace

// The registration of csrc/rasterise_egl.cpp:33-53, unchanged: five inputs, ONE output.  The per-pixel state the
// backward needs is not an output: RasteriseGrad re-derives it from the forward's inputs (stateless mode of
// dirt_rasterise_backward), the way the reference's graph sees the pair (csrc/rasterise_grad_egl.cpp:38-42).
REGISTER_OP("Rasterise")
    .Attr("height: int").Attr("width: int").Attr("channels: int = 3")
    .Input("background: float32").Input("vertices: float32").Input("vertex_colors: float32")
    .Input("faces: int32").Input("camera_pos: float32")
    .Output("pixels: float32")
    .SetShapeFn([](shape_inference::InferenceContext *c) {
        int h, w, ch;
        TF_RETURN_IF_ERROR(c->GetAttr("height", &h));
        TF_RETURN_IF_ERROR(c->GetAttr("width", &w));
        TF_RETURN_IF_ERROR(c->GetAttr("channels", &ch));
        c->set_output(0, c->MakeShape({c->Dim(c->input(1), 0), h, w, ch}));      // as csrc/rasterise_egl.cpp:44-52
        return Status::OK();
    });

class RasteriseOpB200 : public OpKernel {
public:
    explicit RasteriseOpB200(OpKernelConstruction *c) : OpKernel(c)
    {
        OP_REQUIRES_OK(c, c->GetAttr("height", &height_));
        OP_REQUIRES_OK(c, c->GetAttr("width", &width_));
        OP_REQUIRES_OK(c, c->GetAttr("channels", &channels_));
    }
    void Compute(OpKernelContext *ctx) override
    {
        const Tensor &bg = ctx->input(0), &v = ctx->input(1), &col = ctx->input(2), &f = ctx->input(3);
        // the checks of csrc/rasterise_egl.cpp:310-336, message for message (the C ABI re-checks extents and strides)
        OP_REQUIRES(ctx, bg.dims() == 4 && bg.dim_size(1) == height_ && bg.dim_size(2) == width_ && bg.dim_size(3) == channels_,
                    errors::InvalidArgument("Rasterise expects background_tensor to be 4D, and bgcolor.shape == [None, height, width, channels]"));
        OP_REQUIRES(ctx, v.dims() == 3 && v.dim_size(2) == 4, errors::InvalidArgument("Rasterise expects vertices to be 3D, and vertices.shape[2] == 4"));
        OP_REQUIRES(ctx, col.dims() == 3 && col.dim_size(1) == v.dim_size(1) && col.dim_size(2) == channels_,
                    errors::InvalidArgument("Rasterise expects vertex_colors to be 3D, and vertex_colors.shape == [None, vertices.shape[1], channels]"));
        OP_REQUIRES(ctx, f.dims() == 3 && f.dim_size(2) == 3, errors::InvalidArgument("Rasterise expects faces to be 3D, and faces.shape[2] == 3"));
        int64 B = 0;
        OP_REQUIRES(ctx, common_batch({&bg, &v, &col, &f}, &B),
                    errors::InvalidArgument("Rasterise expects all arguments to have same leading (batch) dimension"));
        DirtForwardArgs a = {};
        a.shape = shape_of(bg, v, f, channels_, B);
        a.background = bg.flat<float>().data();        a.background_bstride = batch_stride(bg, B);
        a.vertices = v.flat<float>().data();           a.vertices_bstride = batch_stride(v, B);
        a.vertex_colors = col.flat<float>().data();    a.vertex_colors_bstride = batch_stride(col, B);
        a.faces = f.flat<int32>().data();              a.faces_bstride = batch_stride(f, B);
        Tensor *pixels, ws;
        OP_REQUIRES_OK(ctx, ctx->allocate_output(0, {B, height_, width_, channels_}, &pixels));
        a.workspace_bytes = dirt_workspace_bytes(&a.shape);
        OP_REQUIRES_OK(ctx, ctx->allocate_temp(DT_UINT8, {static_cast<int64>(a.workspace_bytes)}, &ws));
        a.pixels = pixels->flat<float>().data();
        a.workspace = ws.flat<uint8>().data();
        OP_REQUIRES_OK(ctx, status_from(dirt_rasterise_forward(&a, cuda_stream_of(ctx))));
    }
private:
    int height_, width_, channels_;
};
REGISTER_KERNEL_BUILDER(Name("Rasterise").Device(DEVICE_GPU), RasteriseOpB200);   // GPU only, as csrc/rasterise_egl.cpp:520

// The intended backward (csrc/rasterise_grad_common.h:19-24): the forward's vertices and faces, its output and the
// incoming gradient in; grad_background / grad_vertices / grad_vertex_colors out.  Stateless: nothing but graph
// tensors connects it to the forward op.
REGISTER_OP("RasteriseGrad")
    .Input("vertices: float32").Input("faces: int32").Input("pixels: float32").Input("grad_pixels: float32")
    .Output("grad_background: float32").Output("grad_vertices: float32").Output("grad_vertex_colors: float32");

class RasteriseGradOpB200 : public OpKernel {
public:
    using OpKernel::OpKernel;
    void Compute(OpKernelContext *ctx) override
    {
        const Tensor &v = ctx->input(0), &f = ctx->input(1), &px = ctx->input(2), &g = ctx->input(3);
        OP_REQUIRES(ctx, px.dims() == 4 && g.shape() == px.shape(), errors::InvalidArgument("RasteriseGrad expects pixels and grad_pixels of one 4D shape"));
        OP_REQUIRES(ctx, v.dims() == 3 && v.dim_size(2) == 4, errors::InvalidArgument("RasteriseGrad expects vertices to be 3D, and vertices.shape[2] == 4"));
        OP_REQUIRES(ctx, f.dims() == 3 && f.dim_size(2) == 3, errors::InvalidArgument("RasteriseGrad expects faces to be 3D, and faces.shape[2] == 3"));
        int64 B = 0;
        OP_REQUIRES(ctx, common_batch({&px, &v, &f}, &B) && px.dim_size(0) == B,
                    errors::InvalidArgument("RasteriseGrad expects vertices and faces to have the leading (batch) dimension of pixels, or 1"));
        DirtBackwardArgs a = {};
        a.shape = shape_of(px, v, f, static_cast<int>(px.dim_size(3)), B);
        const int64 V = a.shape.vertex_count, C = a.shape.channels;
        a.faces = f.flat<int32>().data();              a.faces_bstride = batch_stride(f, B);
        a.vertices = v.flat<float>().data();           a.vertices_bstride = batch_stride(v, B);
        a.pixels = px.flat<float>().data(); a.grad_pixels = g.flat<float>().data();
        const int64 Bv = v.dim_size(0);            // 1 = shared mesh: gradients summed over the batch
        Tensor *gbg, *gv, *gc, ws;
        OP_REQUIRES_OK(ctx, ctx->allocate_output(0, px.shape(), &gbg));
        OP_REQUIRES_OK(ctx, ctx->allocate_output(1, {Bv, V, 4}, &gv));
        OP_REQUIRES_OK(ctx, ctx->allocate_output(2, {Bv, V, C}, &gc));
        a.workspace_bytes = dirt_backward_workspace_bytes(&a.shape);
        OP_REQUIRES_OK(ctx, ctx->allocate_temp(DT_UINT8, {static_cast<int64>(a.workspace_bytes)}, &ws));
        a.workspace = ws.flat<uint8>().data();
        a.grad_background = gbg->flat<float>().data();
        a.grad_vertices = gv->flat<float>().data();            a.grad_vertices_bstride = Bv == 1 && B > 1 ? 0 : V * 4;
        a.grad_vertex_colors = gc->flat<float>().data();       a.grad_vertex_colors_bstride = Bv == 1 && B > 1 ? 0 : V * C;
        OP_REQUIRES_OK(ctx, status_from(dirt_rasterise_backward(&a, cuda_stream_of(ctx))));
    }
};
REGISTER_KERNEL_BUILDER(Name("RasteriseGrad").Device(DEVICE_GPU), RasteriseGradOpB200);

// ---- the fork's procedural scene ops --------------------------------------------------------
// Same inputs / attrs / output as "Rasterise" (csrc/hill.cpp:33-53, csrc/oceanic_*.cpp:33-53); the
// fragment program is selected by the op name.  Each runs the rasteriser for coverage, then
// dirt_scene_shade on the same stream.  camera_pos stays on the device (no cudaMemcpy as at :323).
template <int SCENE> class SceneOpB200 : public OpKernel {
public:
    explicit SceneOpB200(OpKernelConstruction *c) : OpKernel(c)
    {
        OP_REQUIRES_OK(c, c->GetAttr("height", &height_));
        OP_REQUIRES_OK(c, c->GetAttr("width", &width_));
        OP_REQUIRES_OK(c, c->GetAttr("channels", &channels_));
    }
    void Compute(OpKernelContext *ctx) override
    {
        const Tensor &bg = ctx->input(0), &v = ctx->input(1), &col = ctx->input(2), &f = ctx->input(3), &cam = ctx->input(4);
        const bool hill = SCENE == DIRT_SCENE_HILL;   // csrc/hill.cpp:310 drops the channel check: the background is the terrain texture
        OP_REQUIRES(ctx, bg.dims() == 4 && bg.dim_size(1) == height_ && bg.dim_size(2) == width_ && (hill || bg.dim_size(3) == channels_),
                    errors::InvalidArgument("Rasterise expects background_tensor to be 4D, and bgcolor.shape == [None, height, width, channels]"));
        OP_REQUIRES(ctx, v.dims() == 3 && v.dim_size(2) == 4, errors::InvalidArgument("Rasterise expects vertices to be 3D, and vertices.shape[2] == 4"));
        OP_REQUIRES(ctx, f.dims() == 3 && f.dim_size(2) == 3, errors::InvalidArgument("Rasterise expects faces to be 3D, and faces.shape[2] == 3"));
        int64 B = 0;
        OP_REQUIRES(ctx, common_batch({&bg, &v, &col, &f}, &B),
                    errors::InvalidArgument("Rasterise expects all arguments to have same leading (batch) dimension"));
        DirtForwardArgs a = {};
        a.shape = shape_of(bg, v, f, channels_, B);
        auto stride = [B](const Tensor &t) -> int64_t { return batch_stride(t, B); };
        Tensor *pixels, face_id, gbuffer, ws, blank;
        OP_REQUIRES_OK(ctx, ctx->allocate_output(0, {B, height_, width_, channels_}, &pixels));
        OP_REQUIRES_OK(ctx, ctx->allocate_temp(DT_INT32, {B, height_, width_}, &face_id));
        OP_REQUIRES_OK(ctx, ctx->allocate_temp(DT_FLOAT, {B, height_, width_, 4}, &gbuffer));
        if (hill) {   // uncovered pixels of the hill target are 0: rasterise over a blank background
            OP_REQUIRES_OK(ctx, ctx->allocate_temp(DT_FLOAT, {1, height_, width_, channels_}, &blank));
            cudaMemsetAsync(blank.flat<float>().data(), 0, blank.NumElements() * sizeof(float), static_cast<cudaStream_t>(cuda_stream_of(ctx)));
            a.background = blank.flat<float>().data();  a.background_bstride = B > 1 ? 0 : blank.NumElements();
        } else {
            a.background = bg.flat<float>().data();     a.background_bstride = stride(bg);
        }
        a.vertices = v.flat<float>().data();           a.vertices_bstride = stride(v);
        a.vertex_colors = col.flat<float>().data();    a.vertex_colors_bstride = stride(col);
        a.faces = f.flat<int32>().data();              a.faces_bstride = stride(f);
        a.workspace_bytes = dirt_workspace_bytes(&a.shape);
        OP_REQUIRES_OK(ctx, ctx->allocate_temp(DT_UINT8, {static_cast<int64>(a.workspace_bytes)}, &ws));
        a.pixels = pixels->flat<float>().data();
        a.face_id = face_id.flat<int32>().data();
        a.gbuffer = gbuffer.flat<float>().data();
        a.workspace = ws.flat<uint8>().data();
        OP_REQUIRES_OK(ctx, status_from(dirt_rasterise_forward(&a, cuda_stream_of(ctx))));
        DirtSceneArgs s = {};
        s.shape = a.shape;
        s.background = bg.flat<float>().data();        s.background_bstride = stride(bg);
        s.vertices = a.vertices;                       s.vertices_bstride = a.vertices_bstride;
        s.faces = a.faces;                             s.faces_bstride = a.faces_bstride;
        s.face_id = a.face_id; s.gbuffer = a.gbuffer; s.camera_pos = cam.flat<float>().data(); s.pixels = a.pixels;
        OP_REQUIRES_OK(ctx, status_from(dirt_scene_shade(SCENE, &s, static_cast<int>(bg.dim_size(3)),
                                                         static_cast<int>(cam.NumElements()), cuda_stream_of(ctx))));
    }
private:
    int height_, width_, channels_;
};

#define DIRT_REGISTER_SCENE_OP(NAME, SCENE)                                                              \
    REGISTER_OP(NAME)                                                                                    \
        .Attr("height: int").Attr("width: int").Attr("channels: int = 3")                                \
        .Input("background: float32").Input("vertices: float32").Input("vertex_colors: float32")         \
        .Input("faces: int32").Input("camera_pos: float32")                                              \
        .Output("pixels: float32")                                                                       \
        .SetShapeFn([](shape_inference::InferenceContext *c) {                                           \
            int h, w, ch;                                                                                \
            TF_RETURN_IF_ERROR(c->GetAttr("height", &h));                                                \
            TF_RETURN_IF_ERROR(c->GetAttr("width", &w));                                                 \
            TF_RETURN_IF_ERROR(c->GetAttr("channels", &ch));                                             \
            c->set_output(0, c->MakeShape({c->Dim(c->input(1), 0), h, w, ch}));                          \
            return Status::OK();                                                                         \
        });                                                                                              \
    REGISTER_KERNEL_BUILDER(Name(NAME).Device(DEVICE_GPU), SceneOpB200<SCENE>)

DIRT_REGISTER_SCENE_OP("Hill", DIRT_SCENE_HILL);                                   // csrc/hill.cpp:33, :503
DIRT_REGISTER_SCENE_OP("OceanicNoCloud", DIRT_SCENE_OCEANIC_NO_CLOUD);             // csrc/oceanic_no_cloud.cpp:33, :519
DIRT_REGISTER_SCENE_OP("OceanicSimpleProxy", DIRT_SCENE_OCEANIC_SIMPLE_PROXY);     // csrc/oceanic_simple_proxy.cpp:33, :519
DIRT_REGISTER_SCENE_OP("OceanicStillCloud", DIRT_SCENE_OCEANIC_STILL_CLOUD);       // csrc/oceanic_still_cloud.cpp:33, :523
DIRT_REGISTER_SCENE_OP("OceanicOptFlow", DIRT_SCENE_OCEANIC_OPT_FLOW);             // csrc/oceanic_opt_flow.cpp:33, :544
// The fork's own "Rasterise" / "RasteriseGrad" pictures, for callers that want them back under new names:
DIRT_REGISTER_SCENE_OP("OceanicHorizon", DIRT_SCENE_OCEANIC_HORIZON);              // csrc/rasterise_egl.cpp:385
DIRT_REGISTER_SCENE_OP("Oceanic", DIRT_SCENE_OCEANIC);                             // csrc/rasterise_grad_egl.cpp:399
