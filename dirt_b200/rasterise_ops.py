"""Host-side mirror of the reference's Python API for the rasterise path.

Same names, positional order and argument meaning as dirt/rasterise_ops.py in the reference
(`rasterise` :10-54, `rasterise_batch` :57-88, `rasterise_grad` :91-108), on torch CUDA tensors
instead of TF graph tensors.  All compute goes through the C ABI of libdirt_b200.so
(include/dirt_b200.h); torch only owns device memory, streams and autograd bookkeeping.

Differences from the fork, all documented in DESIGN.md:
  * `camera_pos` is optional: it only parameterises the fork's procedural fragment shaders
    (csrc/rasterise_egl.cpp:399-433) and is not an input of the mesh path.
  * gradients exist: `rasterise` / `rasterise_batch` are differentiable w.r.t. background,
    vertices and vertex_colors (the fork registers no gradient at all).
  * a leading batch dim of 1 on any input broadcasts over the batch without copying (the
    relaxation wished for at csrc/rasterise_egl.cpp:334).
"""
import ctypes

import numpy as np
import torch

from . import _lib
from ._lib import DirtBackwardArgs, DirtForwardArgs, DirtShape, InvalidArgumentError

_workspaces = {}


def _is_host_mapped(t):
    """A pinned host tensor: under unified addressing the kernels can read it in place over PCIe."""
    return isinstance(t, torch.Tensor) and not t.is_cuda and t.dtype == torch.float32 and t.is_pinned()


def _as_tensor(x, dtype, device, name, keep_pinned=False):
    # tf.convert_to_tensor(x, dtype=...) of dirt/rasterise_ops.py:40-43
    if keep_pinned and _is_host_mapped(x) and dtype == torch.float32:
        return x
    if isinstance(x, torch.Tensor):
        t = x
        if t.device != device:
            t = t.to(device, non_blocking=True)
        if t.dtype != dtype:
            if dtype == torch.int32 and t.dtype.is_floating_point:
                raise InvalidArgumentError(-1, "%s must be an integer tensor" % name)
            t = t.to(dtype)
    else:
        t = torch.as_tensor(np.asarray(x), dtype=dtype).to(device, non_blocking=True)
    return t


def _device_of(*xs):
    for x in xs:
        if isinstance(x, torch.Tensor) and x.is_cuda:
            return x.device
    if not torch.cuda.is_available():
        # the reference registers a DEVICE_GPU kernel only (csrc/rasterise_egl.cpp:520)
        raise RuntimeError("dirt_b200.rasterise needs a CUDA device; there is no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device())


def _workspace(device, nbytes):
    stream = torch.cuda.current_stream(device)
    key = (device.index, stream.cuda_stream)
    ws = _workspaces.get(key)
    if ws is None or ws.numel() < nbytes:
        _workspaces.pop(key, None)
        ws = torch.empty(int(nbytes), dtype=torch.uint8, device=device)
        _workspaces[key] = ws
    return ws


def release_workspaces():
    _workspaces.clear()
    _hints.clear()


# Visibility hints (DirtForwardArgs.visibility_hint): one persistent device buffer per mesh that is
# rendered repeatedly -- keyed by device, stream, batch, face count and frame size (not by tensor
# identity: a double-buffered training loop alternates between input buffers), the way the workspace
# is keyed by device and stream.  A hint only orders the kernels' work (results are bit-identical
# with any hint), so a key that stands for another mesh costs time, not correctness.
HINT_MIN_FACES = 4096          # kOrderMinFaces of the library: smaller meshes take one unordered pass
_HINT_SLOTS = 8
_hints = {}


def hint_default():
    """Whether rasterise / rasterise_batch carry a visibility hint from call to call on their own.  Off unless
    DIRT_B200_HINT=1: on the measured workloads the hinted order is no faster than the depth-ordered one
    (DESIGN.md section 6); `forward_raw(..., hint=buffer)` uses it explicitly."""
    import os
    return os.environ.get("DIRT_B200_HINT", "0") not in ("", "0")


def visibility_hint(device, B, F, key=None):
    """The persistent hint buffer for (device, current stream, B, F, key); zeroed when first created."""
    lib = _lib.load()
    shape = DirtShape(B, 1, 1, 1, 0, F)
    nbytes = int(lib.dirt_visibility_hint_bytes(ctypes.byref(shape)))
    k = (device.index, torch.cuda.current_stream(device).cuda_stream, B, F, key)
    h = _hints.pop(k, None)
    if h is None or h.numel() < nbytes:
        h = torch.zeros(nbytes, dtype=torch.uint8, device=device)
    _hints[k] = h                       # most recently used last
    while len(_hints) > _HINT_SLOTS:
        _hints.pop(next(iter(_hints)))
    return h


def _bstride(t, B):
    return 0 if (t.shape[0] == 1 and B > 1) else int(t[0].numel())


def _validate(background, vertices, vertex_colors, faces, height, width, channels):
    # the OP_REQUIRES block of csrc/rasterise_egl.cpp:309-336, message for message
    if channels < 1 or channels > _lib.MAX_CHANNELS:
        raise InvalidArgumentError(-1, "Rasterise expects 1 <= channels <= %d" % _lib.MAX_CHANNELS)
    if height <= 0 or width <= 0:
        raise InvalidArgumentError(-1, "Rasterise expects height > 0 and width > 0")
    if background.dim() != 4 or tuple(background.shape[1:]) != (height, width, channels):
        raise InvalidArgumentError(-1, "Rasterise expects background_tensor to be 4D, and bgcolor.shape == [None, height, width, channels]")
    if vertices.dim() != 3 or vertices.shape[2] != 4:
        raise InvalidArgumentError(-1, "Rasterise expects vertices to be 3D, and vertices.shape[2] == 4")
    if vertex_colors.dim() != 3 or vertex_colors.shape[1] != vertices.shape[1] or vertex_colors.shape[2] != channels:
        raise InvalidArgumentError(-1, "Rasterise expects vertex_colors to be 3D, and vertex_colors.shape == [None, vertices.shape[1], channels]")
    if faces.dim() != 3 or faces.shape[2] != 3:
        raise InvalidArgumentError(-1, "Rasterise expects faces to be 3D, and faces.shape[2] == 3")
    B = max(background.shape[0], vertices.shape[0], vertex_colors.shape[0], faces.shape[0])
    for t in (background, vertices, vertex_colors, faces):
        if t.shape[0] != B and t.shape[0] != 1:
            raise InvalidArgumentError(-1, "Rasterise expects all arguments to have same leading (batch) dimension")
    return B


def forward_raw(background, vertices, vertex_colors, faces, want_aux=True, hint=None):
    """One C-ABI forward call on batched, contiguous CUDA tensors.
    Returns (pixels [B,H,W,C], face_id [B,H,W] | None, gbuffer [B,H,W,4] | None).
    `hint`: None (no visibility hint), a uint8 CUDA tensor of dirt_visibility_hint_bytes() bytes owned by
    the caller, or "auto" (the module keeps one per mesh, see visibility_hint()).
    `background` may also be a PINNED HOST tensor: the resolve kernel reads the background only
    where no face covers the pixel, so reading it in place over PCIe moves just those bytes
    instead of the whole image (the reference uploads all of it, csrc/rasterise_egl.cu:16-50)."""
    lib = _lib.load()
    device = vertices.device              # `background` may be pinned host memory, read in place
    B = max(background.shape[0], vertices.shape[0], vertex_colors.shape[0], faces.shape[0])
    _, H, W, C = background.shape
    V, F = vertices.shape[1], faces.shape[1]
    shape = DirtShape(B, H, W, C, V, F)
    pixels = torch.empty((B, H, W, C), dtype=torch.float32, device=device)
    face_id = torch.empty((B, H, W), dtype=torch.int32, device=device) if want_aux else None
    gbuffer = torch.empty((B, H, W, 4), dtype=torch.float32, device=device) if want_aux else None
    if B == 0:
        return pixels, face_id, gbuffer
    nbytes = lib.dirt_workspace_bytes(ctypes.byref(shape))
    if nbytes == 0:
        _lib.check(-1)
    ws = _workspace(device, nbytes)
    if isinstance(hint, str):
        hint = visibility_hint(device, B, F, key=(H, W)) if (hint == "auto" and F >= HINT_MIN_FACES and hint_default()) else None
    args = DirtForwardArgs(
        shape,
        background.data_ptr(), _bstride(background, B),
        vertices.data_ptr(), _bstride(vertices, B),
        vertex_colors.data_ptr(), _bstride(vertex_colors, B),
        faces.data_ptr(), _bstride(faces, B),
        pixels.data_ptr(),
        face_id.data_ptr() if want_aux else None, gbuffer.data_ptr() if want_aux else None,
        ws.data_ptr(), ws.numel(),
        hint.data_ptr() if hint is not None else None, hint.numel() if hint is not None else 0)
    with torch.cuda.device(device):
        _lib.check(lib.dirt_rasterise_forward(ctypes.byref(args), torch.cuda.current_stream(device).cuda_stream))
    return pixels, face_id, gbuffer


def deterministic_default():
    """DIRT_B200_DETERMINISTIC=1 switches the autograd path to the deterministic backward (debugging aid)."""
    import os
    return os.environ.get("DIRT_B200_DETERMINISTIC", "0") not in ("", "0")


def backward_raw(faces, pixels, grad_pixels, face_id, gbuffer, V, shared_vertices=False, shared_colors=False,
                 want_grad_background=True, vertices=None, deterministic=False):
    """One C-ABI backward call.  Returns (grad_vertices [B|1,V,4], grad_vertex_colors [B|1,V,C],
    grad_background [B,H,W,C] | None).  With face_id = gbuffer = None the call is stateless: it re-derives
    both from `vertices` [B|1,V,4] with the forward's kernels (the mode the TF op pair uses)."""
    lib = _lib.load()
    device = pixels.device
    B, H, W, C = pixels.shape
    F = faces.shape[1]
    shape = DirtShape(B, H, W, C, V, F)
    gv = torch.empty((1 if shared_vertices else B, V, 4), dtype=torch.float32, device=device)
    gc = torch.empty((1 if shared_colors else B, V, C), dtype=torch.float32, device=device)
    gbg = torch.empty((B, H, W, C), dtype=torch.float32, device=device) if want_grad_background else None
    if B == 0:
        return gv.zero_(), gc.zero_(), gbg
    stateless = face_id is None and gbuffer is None
    ws = None
    if stateless or deterministic:
        ws = _workspace(device, (lib.dirt_backward_workspace_bytes(ctypes.byref(shape)) if stateless else 0) +
                        (lib.dirt_deterministic_workspace_bytes(ctypes.byref(shape)) if deterministic else 0))
    args = DirtBackwardArgs(
        shape, faces.data_ptr(), _bstride(faces, B),
        pixels.data_ptr(), grad_pixels.data_ptr(),
        None if face_id is None else face_id.data_ptr(), None if gbuffer is None else gbuffer.data_ptr(),
        gv.data_ptr(), 0 if (shared_vertices and B > 1) else V * 4,
        gc.data_ptr(), 0 if (shared_colors and B > 1) else V * C,
        gbg.data_ptr() if want_grad_background else None,
        vertices.data_ptr() if stateless else None, _bstride(vertices, B) if stateless else 0,
        ws.data_ptr() if ws is not None else None, ws.numel() if ws is not None else 0,
        _lib.BACKWARD_DETERMINISTIC if deterministic else 0)
    with torch.cuda.device(device):
        _lib.check(lib.dirt_rasterise_backward(ctypes.byref(args), torch.cuda.current_stream(device).cuda_stream))
    return gv, gc, gbg


def backward_shared_raw(faces, pixels, grad_pixels, face_id, gbuffer, V, clip_from_world, want_grad_background=True,
                        deterministic=False, out=None):
    """Shared-mesh backward: one C-ABI call that folds every element's
    clip-space vertex gradient to world space with its matrix (clip_b = world @ M_b) and sums
    position and colour gradients over the batch into ONE packed buffer [V, stride]: columns 0..3
    world-space position gradient, 4..4+C-1 colour gradient.  The buffer is the all-reduce message
    of `sharding.allreduce_packed`.  Returns (grad_shared, grad_background | None)."""
    lib = _lib.load()
    device = pixels.device
    B, H, W, C = pixels.shape
    F = faces.shape[1]
    stride = lib.dirt_shared_grad_stride(C)
    # `out`: a caller-owned [V, stride] buffer, e.g. one in symmetric memory (sharding.SymmetricPacked) that the
    # collective reduces in place without a copy
    if out is not None and (tuple(out.shape) != (V, stride) or out.dtype != torch.float32 or not out.is_contiguous()):
        raise _lib.InvalidArgumentError(-1, "out must be a contiguous float32 [V, %d] tensor" % stride)
    packed = out if out is not None else torch.empty((V, stride), dtype=torch.float32, device=device)
    gbg = torch.empty((B, H, W, C), dtype=torch.float32, device=device) if want_grad_background else None
    m = clip_from_world.to(torch.float32).contiguous()
    if tuple(m.shape) != (B, 4, 4):
        raise _lib.InvalidArgumentError(-1, "clip_from_world must be [batch, 4, 4]")
    shape = DirtShape(B, H, W, C, V, F)
    ws = _workspace(device, lib.dirt_deterministic_workspace_bytes(ctypes.byref(shape))) if deterministic else None
    args = _lib.DirtSharedBackwardArgs(
        shape, faces.data_ptr(), _bstride(faces, B),
        pixels.data_ptr(), grad_pixels.data_ptr(), face_id.data_ptr(), gbuffer.data_ptr(),
        m.data_ptr(), packed.data_ptr(), gbg.data_ptr() if want_grad_background else None,
        ws.data_ptr() if deterministic else None, ws.numel() if deterministic else 0,
        _lib.BACKWARD_DETERMINISTIC if deterministic else 0)
    with torch.cuda.device(device):
        _lib.check(lib.dirt_rasterise_backward_shared(ctypes.byref(args), torch.cuda.current_stream(device).cuda_stream))
    return packed, gbg


class _Rasterise(torch.autograd.Function):
    """Forward = op "Rasterise"; backward = op "RasteriseGrad".  In the TF shim the same pairing
    is `@ops.RegisterGradient("Rasterise")`."""

    @staticmethod
    def forward(ctx, background, vertices, vertex_colors, faces):
        need_grad = any(ctx.needs_input_grad[:3])
        ctx.bg_on_host = not background.is_cuda
        pixels, face_id, gbuffer = forward_raw(background, vertices, vertex_colors, faces, want_aux=need_grad, hint="auto")
        if need_grad:
            ctx.save_for_backward(faces, pixels, face_id, gbuffer)
            ctx.meta = (background.shape[0], vertices.shape[0], vertex_colors.shape[0], vertices.shape[1])
        return pixels

    @staticmethod
    def backward(ctx, grad_pixels):
        faces, pixels, face_id, gbuffer = ctx.saved_tensors
        bg_b, v_b, c_b, V = ctx.meta
        B = pixels.shape[0]
        grad_pixels = grad_pixels.contiguous().to(torch.float32)
        gv, gc, gbg = backward_raw(faces, pixels, grad_pixels, face_id, gbuffer, V,
                                   shared_vertices=(v_b == 1 and B > 1), shared_colors=(c_b == 1 and B > 1),
                                   want_grad_background=ctx.needs_input_grad[0], deterministic=deterministic_default())
        if gbg is not None and bg_b == 1 and B > 1:
            gbg = gbg.sum(dim=0, keepdim=True)
        if gbg is not None and ctx.bg_on_host:
            gbg = gbg.cpu()                   # autograd wants the gradient where the leaf lives
        return (gbg if ctx.needs_input_grad[0] else None,
                gv if ctx.needs_input_grad[1] else None,
                gc if ctx.needs_input_grad[2] else None,
                None)


def _prepare(background, vertices, vertex_colors, faces, batched, height, width, channels):
    device = _device_of(background, vertices, vertex_colors, faces)
    background = _as_tensor(background, torch.float32, device, "background", keep_pinned=True)
    vertices = _as_tensor(vertices, torch.float32, device, "vertices")
    vertex_colors = _as_tensor(vertex_colors, torch.float32, device, "vertex_colors")
    faces = _as_tensor(faces, torch.int32, device, "faces")
    if not batched:
        # dirt/rasterise_ops.py:51: every input gets a leading [np.newaxis]
        background, vertices, vertex_colors, faces = (t.unsqueeze(0) for t in (background, vertices, vertex_colors, faces))
    o = 1
    if height is None:
        height = int(background.shape[o]) if background.dim() > o else -1
    if width is None:
        width = int(background.shape[o + 1]) if background.dim() > o + 1 else -1
    if channels is None:
        channels = int(background.shape[o + 2]) if background.dim() > o + 2 else -1
    _validate(background, vertices, vertex_colors, faces, int(height), int(width), int(channels))
    return background.contiguous(), vertices.contiguous(), vertex_colors.contiguous(), faces.contiguous()


def rasterise(background, vertices, vertex_colors, faces, camera_pos=None, height=None, width=None, channels=None,
              name=None):
    """Rasterises `vertices` / `faces` over `background` (reference: dirt/rasterise_ops.py:10-54).

    background [H,W,C] f32; vertices [V,4] f32 clip space; vertex_colors [V,C] f32, interpolated
    perspective-correctly; faces [F,3] int32.  Returns pixels [H,W,C], top row first.
    `camera_pos` and `name` are accepted for call compatibility and unused by the mesh path.
    """
    del camera_pos, name
    bg, v, c, f = _prepare(background, vertices, vertex_colors, faces, False, height, width, channels)
    return _Rasterise.apply(bg, v, c, f)[0]


def rasterise_batch(background, vertices, vertex_colors, faces, height=None, width=None, channels=None, name=None):
    """Batched `rasterise` (reference: dirt/rasterise_ops.py:57-88); every input has a leading
    batch dim (1 broadcasts).  Returns pixels [B,H,W,C]."""
    del name
    bg, v, c, f = _prepare(background, vertices, vertex_colors, faces, True, height, width, channels)
    return _Rasterise.apply(bg, v, c, f)


class _RasteriseShared(torch.autograd.Function):
    """One mesh seen by B cameras: forward projects and rasterises, backward is the fused
    shared-mesh backward (gradients w.r.t. the world-space vertices and the colours, already
    summed over the frames)."""

    @staticmethod
    def forward(ctx, background, world_vertices, clip_from_world, vertex_colors, faces):
        if ctx.needs_input_grad[2]:
            raise NotImplementedError("rasterise_batch_shared has no gradient w.r.t. clip_from_world: project the "
                                      "vertices yourself and call rasterise_batch for that")
        vertices = torch.matmul(world_vertices.unsqueeze(0), clip_from_world).contiguous()      # [B,V,4]
        need_grad = any(ctx.needs_input_grad)
        ctx.bg_on_host = not background.is_cuda
        pixels, face_id, gbuffer = forward_raw(background, vertices, vertex_colors.unsqueeze(0), faces.unsqueeze(0),
                                               want_aux=need_grad, hint="auto")
        if need_grad:
            ctx.save_for_backward(faces, pixels, face_id, gbuffer, clip_from_world)
            ctx.meta = (world_vertices.shape[0], background.shape[0])
        return pixels

    @staticmethod
    def backward(ctx, grad_pixels):
        faces, pixels, face_id, gbuffer, clip_from_world = ctx.saved_tensors
        V, bg_b = ctx.meta
        C = pixels.shape[-1]
        packed, gbg = backward_shared_raw(faces.unsqueeze(0), pixels, grad_pixels.contiguous().to(torch.float32), face_id,
                                          gbuffer, V, clip_from_world, want_grad_background=ctx.needs_input_grad[0],
                                          deterministic=deterministic_default())
        if gbg is not None and bg_b == 1 and pixels.shape[0] > 1:
            gbg = gbg.sum(dim=0, keepdim=True)
        if gbg is not None and ctx.bg_on_host:
            gbg = gbg.cpu()
        return gbg, packed[:, :4], None, packed[:, 4:4 + C], None


def rasterise_batch_shared(background, world_vertices, clip_from_world, vertex_colors, faces):
    """The shared-mesh form of `rasterise_batch` (the inverse-rendering loop of
    tests/optimize_sequence.py:293-352): ONE mesh `world_vertices` [V,4], `vertex_colors` [V,C],
    `faces` [F,3], rendered by B cameras `clip_from_world` [B,4,4] (row vectors: clip = world @ M_b)
    over `background` [B|1,H,W,C].  Returns pixels [B,H,W,C]; the backward delivers the gradients
    w.r.t. world_vertices and vertex_colors summed over the frames without materialising the
    per-frame [B,V,4] / [B,V,C] gradients.  Across GPUs, all-reduce `.grad` of the two leaves (or use
    `rasterise_ops.backward_shared_raw` + `sharding.allreduce_packed` for a single collective)."""
    dev = _device_of(background, world_vertices, clip_from_world, vertex_colors, faces)
    bg = _as_tensor(background, torch.float32, dev, "background", keep_pinned=True).contiguous()
    wv = _as_tensor(world_vertices, torch.float32, dev, "world_vertices").contiguous()
    m = _as_tensor(clip_from_world, torch.float32, dev, "clip_from_world").contiguous()
    col = _as_tensor(vertex_colors, torch.float32, dev, "vertex_colors").contiguous()
    f = _as_tensor(faces, torch.int32, dev, "faces").contiguous()
    if bg.dim() != 4 or wv.dim() != 2 or wv.shape[1] != 4 or m.dim() != 3 or tuple(m.shape[1:]) != (4, 4) \
            or col.dim() != 2 or col.shape[0] != wv.shape[0] or f.dim() != 2 or f.shape[1] != 3:
        raise InvalidArgumentError(-1, "rasterise_batch_shared expects background [B|1,H,W,C], world_vertices [V,4], "
                                       "clip_from_world [B,4,4], vertex_colors [V,C], faces [F,3]")
    if col.shape[1] != bg.shape[3]:
        raise InvalidArgumentError(-1, "vertex_colors and background must have the same number of channels")
    B = m.shape[0]
    if bg.shape[0] not in (1, B):
        raise InvalidArgumentError(-1, "background batch must be 1 or the number of cameras")
    return _RasteriseShared.apply(bg, wv, m, col, f)


def rasterise_grad(background, vertices, vertex_colors, faces, camera_pos=None, height=None, width=None,
                   channels=None, name=None, grad_pixels=None):
    """Explicit backward (op name "RasteriseGrad", csrc/rasterise_grad_egl.cpp:33-53) with the
    reference wrapper's leading arguments (dirt/rasterise_ops.py:91-108): runs the forward for
    the given inputs and returns the gradients of sum(pixels * grad_pixels) as
    (grad_background, grad_vertices, grad_vertex_colors).  Inputs may be unbatched or batched;
    `grad_pixels` defaults to ones."""
    del camera_pos, name
    batched = (torch.as_tensor(vertices).dim() if not isinstance(vertices, torch.Tensor) else vertices.dim()) == 3
    bg, v, c, f = _prepare(background, vertices, vertex_colors, faces, batched, height, width, channels)
    pixels, face_id, gbuffer = forward_raw(bg, v, c, f, want_aux=True)
    if grad_pixels is None:
        g = torch.ones_like(pixels)
    else:
        g = _as_tensor(grad_pixels, torch.float32, pixels.device, "grad_pixels")
        g = (g if batched else g.unsqueeze(0)).expand_as(pixels).contiguous()
    B = pixels.shape[0]
    gv, gc, gbg = backward_raw(f, pixels, g, face_id, gbuffer, v.shape[1],
                               shared_vertices=(v.shape[0] == 1 and B > 1), shared_colors=(c.shape[0] == 1 and B > 1))
    if bg.shape[0] == 1 and B > 1:
        gbg = gbg.sum(dim=0, keepdim=True)
    if not batched:
        gbg, gv, gc = gbg[0], gv[0], gc[0]
    return gbg, gv, gc
