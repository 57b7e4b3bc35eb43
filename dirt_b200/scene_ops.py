"""Procedural scene ops of the fork that sit on top of the rasteriser.

In yyuting/dirt none of the compiled ops outputs interpolated vertex colours: each one runs a
procedural fragment shader on every covered fragment, driven by the extra `camera_pos` input.
Here the new rasteriser decides coverage, then one CUDA per-pixel kernel evaluates the shader:

    function                reference wrapper (dirt/rasterise_ops.py)   op file / shader (csrc/)
    oceanic_horizon         rasterise, :10                              rasterise_egl.cpp:385, shaders.cpp:1668-1919
    oceanic                 rasterise_grad, :91                         rasterise_grad_egl.cpp:399, shaders.cpp:556-865
    oceanic_no_cloud        :110                                        oceanic_no_cloud.cpp, shaders.cpp:1402-1665
    oceanic_simple_proxy    :129                                        oceanic_simple_proxy.cpp, shaders.cpp:1921-2185
    oceanic_still_cloud     :148                                        oceanic_still_cloud.cpp, shaders.cpp:866-1175
    oceanic_opt_flow        :167                                        oceanic_opt_flow.cpp, shaders.cpp:1178-1398
    hill                    :186                                        hill.cpp, shaders.cpp:123-552

(`dirt_b200.rasterise_grad` is the backward the op name promises; what the fork's op of that name
really renders is `oceanic` here.)  Like the fork's ops these have no gradient: the fork fits
cameras with a derivative-free Powell search (tests/optimize_sequence.py:159-189).
"""
import ctypes

import torch

from . import _lib, rasterise_ops

SCENE_OCEANIC_HORIZON, SCENE_OCEANIC, SCENE_OCEANIC_STILL_CLOUD, SCENE_OCEANIC_NO_CLOUD = 0, 1, 2, 3
SCENE_OCEANIC_SIMPLE_PROXY, SCENE_OCEANIC_OPT_FLOW, SCENE_HILL = 4, 5, 6


def _scene_batch(scene, background, vertices, vertex_colors, faces, camera_pos, height=None, width=None, channels=None,
                 batched=True):
    lib = _lib.load()
    if scene == SCENE_HILL:
        # hill.cpp:310 drops the channel check on the background: it is the terrain texture
        # ([H,W,1|3|4]: height, normal xyz), while `channels` sizes the output.
        device = rasterise_ops._device_of(background, vertices, vertex_colors, faces)
        tex = rasterise_ops._as_tensor(background, torch.float32, device, "background")
        if not batched:
            tex = tex.unsqueeze(0)
        if tex.dim() != 4 or tex.shape[3] not in (1, 3, 4):
            raise _lib.InvalidArgumentError(-1, "Rasterise expects background_tensor to be 4D, and bgcolor.shape == [None, height, width, channels]")
        tex = tex.contiguous()
        if channels is None:
            channels = 3 if tex.shape[3] == 4 else int(tex.shape[3])
        H = int(tex.shape[1]) if height is None else int(height)
        W = int(tex.shape[2]) if width is None else int(width)
        blank = torch.zeros((tex.shape[0], H, W, int(channels)), dtype=torch.float32, device=tex.device)
        bg, v, c, f = rasterise_ops._prepare(blank if batched else blank[0], vertices, vertex_colors, faces, batched, H, W, channels)
        if tuple(tex.shape[1:3]) != (H, W):
            raise _lib.InvalidArgumentError(-1, "Rasterise expects background_tensor to be 4D, and bgcolor.shape == [None, height, width, channels]")
    else:
        bg, v, c, f = rasterise_ops._prepare(background, vertices, vertex_colors, faces, batched, height, width, channels)
        tex = bg
    cam = rasterise_ops._as_tensor(camera_pos, torch.float32, v.device, "camera_pos").reshape(-1).contiguous()      # bg may be pinned host memory
    need = lib.dirt_scene_camera_floats(scene)
    if cam.numel() < need:
        raise _lib.InvalidArgumentError(-1, "this scene op expects camera_pos to hold at least %d values" % need)
    with torch.no_grad():
        pixels, face_id, gbuffer = rasterise_ops.forward_raw(bg, v, c, f, want_aux=True)
        B, H, W, C = pixels.shape
        if B:
            bs = rasterise_ops._bstride
            args = _lib.DirtSceneArgs(
                _lib.DirtShape(B, H, W, C, v.shape[1], f.shape[1]),
                tex.data_ptr(), bs(tex, B), v.data_ptr(), bs(v, B), f.data_ptr(), bs(f, B),
                face_id.data_ptr(), gbuffer.data_ptr(), cam.data_ptr(), pixels.data_ptr())
            with torch.cuda.device(v.device):
                _lib.check(lib.dirt_scene_shade(scene, ctypes.byref(args), int(tex.shape[3]), int(cam.numel()),
                                                torch.cuda.current_stream(v.device).cuda_stream))
    return pixels


def _make(scene, name, doc):
    def single(background, vertices, vertex_colors, faces, camera_pos, height=None, width=None, channels=None, name=None):
        del name
        return _scene_batch(scene, background, vertices, vertex_colors, faces, camera_pos, height, width, channels, batched=False)[0]

    def batch(background, vertices, vertex_colors, faces, camera_pos, height=None, width=None, channels=None):
        return _scene_batch(scene, background, vertices, vertex_colors, faces, camera_pos, height, width, channels, batched=True)

    single.__name__, batch.__name__ = name, name + "_batch"
    single.__doc__ = doc + ("\n\nSame signature as the fork's wrapper: background [H,W,C], vertices [V,4] clip space, vertex_colors "
                            "[V,C] (ignored by the shader, as in the fork), faces [F,3], camera_pos on host or device.  Returns pixels [H,W,C].")
    batch.__doc__ = "Batched form of `%s` (leading B on the four tensors); camera_pos is shared by the batch, one set of uniforms per op call." % name
    return single, batch


oceanic_horizon, oceanic_horizon_batch = _make(
    SCENE_OCEANIC_HORIZON, "oceanic_horizon",
    "The fork's `rasterise` (dirt/rasterise_ops.py:10): camera_pos [>=8] = (cam_x, cam_y, cam_z, ang1, ang2, ang3, time, light_z); "
    "(1, sun term, 0) on sky, (0, sun glint, 0) on sea, background where nothing is drawn.")
oceanic, oceanic_batch = _make(
    SCENE_OCEANIC, "oceanic",
    "What the fork's `rasterise_grad` really renders (dirt/rasterise_ops.py:91, csrc/rasterise_grad_egl.cpp:399): the full ocean "
    "shader with drifting clouds; camera_pos [>=8] as for oceanic_horizon.")
oceanic_no_cloud, oceanic_no_cloud_batch = _make(
    SCENE_OCEANIC_NO_CLOUD, "oceanic_no_cloud", "dirt/rasterise_ops.py:110: the ocean shader without the cloud layers; camera_pos [>=8].")
oceanic_simple_proxy, oceanic_simple_proxy_batch = _make(
    SCENE_OCEANIC_SIMPLE_PROXY, "oceanic_simple_proxy",
    "dirt/rasterise_ops.py:129: the cheap look-alike (cosine swell, 3 chop octaves, 10 march steps, other palette); camera_pos [>=8].")
oceanic_still_cloud, oceanic_still_cloud_batch = _make(
    SCENE_OCEANIC_STILL_CLOUD, "oceanic_still_cloud",
    "dirt/rasterise_ops.py:148: clouds placed by camera_pos[8] (cloud_t) instead of time and not tied to the camera position; camera_pos [>=9].")
oceanic_opt_flow, oceanic_opt_flow_batch = _make(
    SCENE_OCEANIC_OPT_FLOW, "oceanic_opt_flow",
    "dirt/rasterise_ops.py:167: where each pixel's flat-sea (or sky) point was one step earlier: channels 0, 1 = previous pixel "
    "coordinates; camera_pos [>=16] = (..., unused, dt, dx, dy, dz, dang1, dang2, dang3) (csrc/oceanic_opt_flow.cpp:408-414).")
hill, hill_batch = _make(
    SCENE_HILL, "hill",
    "dirt/rasterise_ops.py:186: rolling grass hills ray-marched against a height texture.  `background` [H,W,1|3|4] is the terrain "
    "lookup (height, normal xyz; bilinear, clamp to edge), `channels` (default 3) sizes the output; camera_pos [>=12]: slots 9..11 = "
    "camera origin (o0, o1, o2), slots 0..8 are shadowed by constants in the shader.  Pixels outside the drawn faces are 0.")
