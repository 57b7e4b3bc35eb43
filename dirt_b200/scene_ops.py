"""Procedural scene ops of the fork that sit on top of the rasteriser.

In yyuting/dirt the op behind `dirt.rasterise` does not output interpolated vertex colours: it
runs the `oceanic_horizon` fragment shader on every covered fragment (csrc/rasterise_egl.cpp:
384-386; csrc/shaders.cpp:1668-1919), driven by the extra `camera_pos` input
(dirt/rasterise_ops.py:10, tests/optimize_horizon.py:251-268).  `oceanic_horizon` below is that
behaviour: the new rasteriser decides coverage, then a CUDA per-pixel kernel evaluates the shader.
The other scene ops of the fork (hill, oceanic_no_cloud / simple_proxy / still_cloud / opt_flow)
are not ported yet.  Like the fork's op, this one has no gradient (the fork fits cameras with a
derivative-free Powell search, tests/optimize_sequence.py:159-189).
"""
import ctypes

import torch

from . import _lib, rasterise_ops


def oceanic_horizon(background, vertices, vertex_colors, faces, camera_pos, height=None, width=None, channels=None,
                    name=None):
    """Same signature as the fork's `rasterise` (dirt/rasterise_ops.py:10): background [H,W,C],
    vertices [V,4] clip space, vertex_colors [V,C] (ignored by the shader, as in the fork), faces
    [F,3], camera_pos [>=8] = (cam_x, cam_y, cam_z, ang1, ang2, ang3, time, light_z).
    Returns pixels [H,W,C]: (1, sun term, 0) on sky, (0, sun glint, 0) on sea, background elsewhere."""
    del name
    bg, v, c, f = rasterise_ops._prepare(background, vertices, vertex_colors, faces, False, height, width, channels)
    return oceanic_horizon_batch(bg, v, c, f, camera_pos)[0]


def oceanic_horizon_batch(background, vertices, vertex_colors, faces, camera_pos):
    """Batched form; camera_pos [>=8] is shared by the batch (one set of uniforms per op call,
    csrc/rasterise_egl.cpp:399-433)."""
    lib = _lib.load()
    bg, v, c, f = rasterise_ops._prepare(background, vertices, vertex_colors, faces, True, None, None, None)
    cam = rasterise_ops._as_tensor(camera_pos, torch.float32, bg.device, "camera_pos").reshape(-1).contiguous()
    if cam.numel() < 8:
        raise _lib.InvalidArgumentError(-1, "Rasterise expects camera_pos to hold at least 8 values")
    with torch.no_grad():
        pixels, face_id, gbuffer = rasterise_ops.forward_raw(bg, v, c, f, want_aux=True)
        B, H, W, C = pixels.shape
        if B:
            bs = rasterise_ops._bstride
            args = _lib.DirtSceneArgs(
                _lib.DirtShape(B, H, W, C, v.shape[1], f.shape[1]),
                bg.data_ptr(), bs(bg, B), v.data_ptr(), bs(v, B), f.data_ptr(), bs(f, B),
                face_id.data_ptr(), gbuffer.data_ptr(), cam.data_ptr(), pixels.data_ptr())
            with torch.cuda.device(bg.device):
                _lib.check(lib.dirt_scene_oceanic_horizon(ctypes.byref(args), torch.cuda.current_stream(bg.device).cuda_stream))
    return pixels
