"""ctypes front-end of oracle/golden_raster.c (numpy in, numpy out).

TEST INFRASTRUCTURE ONLY.  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this module; the product package
dirt_b200/ never does.  Parity status is stated in golden_raster.c's header:
coverage pinned by the reference's KAT-1 / KAT-2, attribute values and gradients unpinned.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libgolden_raster.so")
_lib = None


_SO_SCENE = os.path.join(_HERE, "_build", "libscene_oceanic.so")
_scene = None
_SO_SHADERS = os.path.join(_HERE, "_build", "libscene_shaders.so")
_shaders = None
SCENES = {"oceanic_horizon": 0, "oceanic": 1, "oceanic_still_cloud": 2, "oceanic_no_cloud": 3, "oceanic_simple_proxy": 4,
          "oceanic_opt_flow": 5, "hill": 6}


def build(force=False):
    stale = force
    for so, src in ((_SO, "golden_raster.c"), (_SO_SCENE, "scene_oceanic.c"), (_SO_SHADERS, "scene_shaders.cpp")):
        src = os.path.join(_HERE, src)
        stale = stale or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src)
    if stale:
        subprocess.check_call(["make", "-s", "-C", _HERE, "all"] + (["-B"] if force else []))
    return _SO


def _load():
    global _lib
    if _lib is None:
        build()          # no-op when the library is newer than golden_raster.c
        lib = ctypes.CDLL(_SO)
        i32, i64, vp = ctypes.c_int, ctypes.c_int64, ctypes.c_void_p
        lib.golden_rasterise_forward.argtypes = [i32] * 6 + [vp, i64, vp, i64, vp, i64, vp, i64, vp, vp, vp, i32]
        lib.golden_rasterise_forward.restype = i32
        lib.golden_rasterise_backward.argtypes = [i32] * 6 + [vp, i64, vp, vp, vp, vp, vp, i64, vp, i64, vp, i32]
        lib.golden_rasterise_backward.restype = i32
        lib.golden_num_threads.restype = i32
        _lib = lib
    return _lib


def _prep(a, dtype, ndim, name):
    a = np.ascontiguousarray(a, dtype=dtype)
    if a.ndim == ndim - 1:
        a = a[None]
    if a.ndim != ndim:
        raise ValueError("%s must have %d dims (or %d, unbatched)" % (name, ndim, ndim - 1))
    return a


def _bstride(a, B):
    if a.shape[0] == B:
        return int(np.prod(a.shape[1:]))
    if a.shape[0] == 1:
        return 0
    raise ValueError("batch dim must be %d or 1" % B)


def max_threads():
    return int(_load().golden_num_threads())


def forward(background, vertices, vertex_colors, faces, threads=1):
    """Batched forward.  Inputs [B|1, ...]; returns pixels [B,H,W,C], face_id [B,H,W] (-1 =
    background), gbuf [B,H,W,4] = (B0, B1, w_clip, zw)."""
    lib = _load()
    bg = _prep(background, np.float32, 4, "background")
    v = _prep(vertices, np.float32, 3, "vertices")
    c = _prep(vertex_colors, np.float32, 3, "vertex_colors")
    f = _prep(faces, np.int32, 3, "faces")
    B = max(bg.shape[0], v.shape[0], c.shape[0], f.shape[0])
    _, H, W, C = bg.shape
    V, F = v.shape[1], f.shape[1]
    assert v.shape[2] == 4 and f.shape[2] == 3 and c.shape[1] == V and c.shape[2] == C
    pixels = np.empty((B, H, W, C), np.float32)
    face_id = np.empty((B, H, W), np.int32)
    gbuf = np.empty((B, H, W, 4), np.float32)
    rc = lib.golden_rasterise_forward(
        B, H, W, C, V, F, bg.ctypes.data, _bstride(bg, B), v.ctypes.data, _bstride(v, B),
        c.ctypes.data, _bstride(c, B), f.ctypes.data, _bstride(f, B),
        pixels.ctypes.data, face_id.ctypes.data, gbuf.ctypes.data, int(threads))
    if rc != 0:
        raise RuntimeError("golden_rasterise_forward failed: %d" % rc)
    return pixels, face_id, gbuf


def backward(faces, pixels, grad_pixels, face_id, gbuf, V, shared_vertices=False, shared_colors=False,
             threads=1):
    """Batched backward.  Returns grad_vertices [B|1,V,4], grad_vertex_colors [B|1,V,C],
    grad_background [B,H,W,C]."""
    lib = _load()
    px = _prep(pixels, np.float32, 4, "pixels")
    g = _prep(grad_pixels, np.float32, 4, "grad_pixels")
    f = _prep(faces, np.int32, 3, "faces")
    fid = _prep(face_id, np.int32, 3, "face_id")
    gb = _prep(gbuf, np.float32, 4, "gbuf")
    B, H, W, C = px.shape
    F = f.shape[1]
    gv = np.zeros((1 if shared_vertices else B, V, 4), np.float32)
    gc = np.zeros((1 if shared_colors else B, V, C), np.float32)
    gbg = np.empty((B, H, W, C), np.float32)
    rc = lib.golden_rasterise_backward(
        B, H, W, C, V, F, f.ctypes.data, _bstride(f, B), px.ctypes.data, g.ctypes.data,
        fid.ctypes.data, gb.ctypes.data,
        gv.ctypes.data, 0 if shared_vertices else V * 4, gc.ctypes.data, 0 if shared_colors else V * C,
        gbg.ctypes.data, int(threads))
    if rc != 0:
        raise RuntimeError("golden_rasterise_backward failed: %d" % rc)
    return gv, gc, gbg


def oceanic_horizon(background, vertices, faces, camera_pos, threads=1):
    """The fork's `Rasterise` output (oracle/scene_oceanic.c): rasterises `vertices` / `faces` over
    `background` [B,H,W,C] for coverage and the interpolated texCoordV (= clip xy), then evaluates
    shaders::oceanic_horizon on the covered pixels.  Returns pixels [B,H,W,C]."""
    global _scene
    build()
    if _scene is None:
        _scene = ctypes.CDLL(_SO_SCENE)
        vp = ctypes.c_void_p
        _scene.golden_oceanic_horizon.argtypes = [ctypes.c_int] * 4 + [vp] * 5
        _scene.golden_oceanic_horizon.restype = ctypes.c_int
    bg = _prep(background, np.float32, 4, "background")
    v = _prep(vertices, np.float32, 3, "vertices")
    B = max(bg.shape[0], v.shape[0])
    _, H, W, C = bg.shape
    bgB = np.ascontiguousarray(np.broadcast_to(bg, (B, H, W, C)))
    tc, fid, _ = forward(np.zeros((B, H, W, 2), np.float32), v, v[..., :2], faces, threads=threads)
    cam = np.ascontiguousarray(camera_pos, np.float32).reshape(-1)
    assert cam.size >= 8
    pixels = bgB.copy()
    rc = _scene.golden_oceanic_horizon(B, H, W, C, tc.ctypes.data, fid.ctypes.data, bgB.ctypes.data, cam.ctypes.data,
                                       pixels.ctypes.data)
    if rc != 0:
        raise RuntimeError("golden_oceanic_horizon failed: %d" % rc)
    return pixels


def scene(name, background, vertices, faces, camera_pos, channels=None, threads=1):
    """Any of the fork's scene ops (SCENES) on the CPU: coverage and the interpolated texCoordV from
    the golden rasteriser, then the shader restated in oracle/scene_shaders.cpp (scene_oceanic.c
    for oceanic_horizon).  background [B,H,W,Cb]; for hill it is the terrain texture and
    `channels` (default 3) sizes the output, whose uncovered pixels are 0."""
    global _shaders
    sid = SCENES[name]
    if sid == 0:
        return oceanic_horizon(background, vertices, faces, camera_pos, threads=threads)
    build()
    if _shaders is None:
        _shaders = ctypes.CDLL(_SO_SHADERS)
        vp = ctypes.c_void_p
        _shaders.golden_scene_shade.argtypes = [ctypes.c_int] * 6 + [vp] * 4 + [ctypes.c_int, vp]
        _shaders.golden_scene_shade.restype = ctypes.c_int
    bg = _prep(background, np.float32, 4, "background")
    v = _prep(vertices, np.float32, 3, "vertices")
    B = max(bg.shape[0], v.shape[0])
    _, H, W, Cb = bg.shape
    bgB = np.ascontiguousarray(np.broadcast_to(bg, (B, H, W, Cb)))
    tc, fid, _ = forward(np.zeros((B, H, W, 2), np.float32), v, v[..., :2], faces, threads=threads)
    cam = np.ascontiguousarray(camera_pos, np.float32).reshape(-1)
    if sid == 6:
        C = 3 if channels is None else int(channels)
        pixels = np.zeros((B, H, W, C), np.float32)
    else:
        C = Cb
        pixels = bgB.copy()
    rc = _shaders.golden_scene_shade(sid, B, H, W, C, Cb, tc.ctypes.data, fid.ctypes.data, bgB.ctypes.data, cam.ctypes.data,
                                     int(cam.size), pixels.ctypes.data)
    if rc != 0:
        raise ValueError("golden_scene_shade rejected its arguments (scene %s, %d camera values)" % (name, cam.size))
    return pixels
