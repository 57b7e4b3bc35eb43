"""ctypes binding of libdirt_b200.so (include/dirt_b200.h).

There is no CPU fallback and no alternative implementation: if the library is missing this
module raises on first use, exactly as the reference fails to import without
librasterise.so (dirt/rasterise_ops.py:6-7).
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# DIRT_B200_LIB selects an alternative build of the same library (kernel tuning experiments)
LIB_PATH = os.environ.get("DIRT_B200_LIB") or os.path.join(_HERE, "libdirt_b200.so")

STATUS_NAMES = {0: "OK", -1: "INVALID_ARGUMENT", -2: "CUDA_ERROR", -3: "WORKSPACE_TOO_SMALL", -4: "UNSUPPORTED"}
MAX_CHANNELS = 16
ABI_VERSION = 2
BACKWARD_DETERMINISTIC = 1

_f32p = ctypes.c_void_p
_i64 = ctypes.c_int64


class DirtShape(ctypes.Structure):
    _fields_ = [("batch", ctypes.c_int32), ("height", ctypes.c_int32), ("width", ctypes.c_int32),
                ("channels", ctypes.c_int32), ("vertex_count", ctypes.c_int32), ("face_count", ctypes.c_int32)]


class DirtForwardArgs(ctypes.Structure):
    _fields_ = [("shape", DirtShape),
                ("background", _f32p), ("background_bstride", _i64),
                ("vertices", _f32p), ("vertices_bstride", _i64),
                ("vertex_colors", _f32p), ("vertex_colors_bstride", _i64),
                ("faces", _f32p), ("faces_bstride", _i64),
                ("pixels", _f32p), ("face_id", _f32p), ("gbuffer", _f32p),
                ("workspace", ctypes.c_void_p), ("workspace_bytes", ctypes.c_size_t),
                ("visibility_hint", ctypes.c_void_p), ("visibility_hint_bytes", ctypes.c_size_t)]


class DirtBackwardArgs(ctypes.Structure):
    _fields_ = [("shape", DirtShape),
                ("faces", _f32p), ("faces_bstride", _i64),
                ("pixels", _f32p), ("grad_pixels", _f32p), ("face_id", _f32p), ("gbuffer", _f32p),
                ("grad_vertices", _f32p), ("grad_vertices_bstride", _i64),
                ("grad_vertex_colors", _f32p), ("grad_vertex_colors_bstride", _i64),
                ("grad_background", _f32p),
                ("vertices", _f32p), ("vertices_bstride", _i64),
                ("workspace", ctypes.c_void_p), ("workspace_bytes", ctypes.c_size_t),
                ("flags", ctypes.c_uint32)]


class DirtSharedBackwardArgs(ctypes.Structure):
    _fields_ = [("shape", DirtShape),
                ("faces", _f32p), ("faces_bstride", _i64),
                ("pixels", _f32p), ("grad_pixels", _f32p), ("face_id", _f32p), ("gbuffer", _f32p),
                ("clip_from_world", _f32p), ("grad_shared", _f32p), ("grad_background", _f32p),
                ("workspace", ctypes.c_void_p), ("workspace_bytes", ctypes.c_size_t), ("flags", ctypes.c_uint32)]


class DirtPeerAllreduceArgs(ctypes.Structure):
    _fields_ = [("buffers", ctypes.c_void_p), ("signals", ctypes.c_void_p), ("rank", ctypes.c_int32), ("world", ctypes.c_int32),
                ("count", _i64), ("epoch", ctypes.c_void_p), ("multicast", ctypes.c_void_p)]


class DirtVertexShadeArgs(ctypes.Structure):
    _fields_ = [("batch", ctypes.c_int), ("vertex_count", ctypes.c_int), ("channels", ctypes.c_int), ("double_sided", ctypes.c_int),
                ("positions", ctypes.c_void_p), ("normals", ctypes.c_void_p), ("colours", ctypes.c_void_p),
                ("light_vector", ctypes.c_void_p), ("light_vector_bstride", _i64),
                ("light_color", ctypes.c_void_p), ("light_color_bstride", _i64),
                ("camera_position", ctypes.c_void_p), ("camera_position_bstride", _i64),
                ("shininess", ctypes.c_void_p), ("shininess_bstride", _i64),
                ("out", ctypes.c_void_p), ("grad_out", ctypes.c_void_p),
                ("grad_positions", ctypes.c_void_p), ("grad_normals", ctypes.c_void_p), ("grad_colours", ctypes.c_void_p)]


class DirtSceneArgs(ctypes.Structure):
    _fields_ = [("shape", DirtShape),
                ("background", _f32p), ("background_bstride", _i64),
                ("vertices", _f32p), ("vertices_bstride", _i64),
                ("faces", _f32p), ("faces_bstride", _i64),
                ("face_id", _f32p), ("gbuffer", _f32p), ("camera_pos", _f32p), ("pixels", _f32p)]


# every symbol include/dirt_b200.h declares: name -> (restype, argtypes)
EXPORTS = {
    "dirt_abi_version": (ctypes.c_int, []),
    "dirt_last_error": (ctypes.c_char_p, []),
    "dirt_workspace_bytes": (ctypes.c_size_t, [ctypes.POINTER(DirtShape)]),
    "dirt_visibility_hint_bytes": (ctypes.c_size_t, [ctypes.POINTER(DirtShape)]),
    "dirt_backward_workspace_bytes": (ctypes.c_size_t, [ctypes.POINTER(DirtShape)]),
    "dirt_deterministic_workspace_bytes": (ctypes.c_size_t, [ctypes.POINTER(DirtShape)]),
    "dirt_workspace_create": (ctypes.c_int, [ctypes.POINTER(DirtShape), ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_size_t)]),
    "dirt_workspace_destroy": (ctypes.c_int, [ctypes.c_void_p]),
    "dirt_rasterise_forward": (ctypes.c_int, [ctypes.POINTER(DirtForwardArgs), ctypes.c_void_p]),
    "dirt_noalloc_workspace_limit": (ctypes.c_size_t, []),
    "dirt_rasterise_forward_noalloc": (ctypes.c_int, [ctypes.POINTER(DirtForwardArgs), ctypes.c_void_p]),
    "dirt_rasterise_backward": (ctypes.c_int, [ctypes.POINTER(DirtBackwardArgs), ctypes.c_void_p]),
    "dirt_shared_grad_stride": (ctypes.c_int, [ctypes.c_int]),
    "dirt_rasterise_backward_shared": (ctypes.c_int, [ctypes.POINTER(DirtSharedBackwardArgs), ctypes.c_void_p]),
    "dirt_allreduce_peer": (ctypes.c_int, [ctypes.POINTER(DirtPeerAllreduceArgs), ctypes.c_void_p]),
    "dirt_rasterise_forward_host": (ctypes.c_int, [ctypes.POINTER(DirtForwardArgs)]),
    "dirt_rasterise_backward_host": (ctypes.c_int, [ctypes.POINTER(DirtBackwardArgs)]),
    "dirt_release_host_staging": (ctypes.c_int, []),
    "dirt_scene_oceanic_horizon": (ctypes.c_int, [ctypes.POINTER(DirtSceneArgs), ctypes.c_void_p]),
    "dirt_scene_camera_floats": (ctypes.c_int, [ctypes.c_int]),
    "dirt_scene_shade": (ctypes.c_int, [ctypes.c_int, ctypes.POINTER(DirtSceneArgs), ctypes.c_int, ctypes.c_int, ctypes.c_void_p]),
    "dirt_vertex_normals": (ctypes.c_int, [ctypes.c_void_p] + [ctypes.c_int] * 3 + [ctypes.c_void_p, ctypes.c_int] + [ctypes.c_void_p] * 3),
    "dirt_vertex_normals_grad": (ctypes.c_int, [ctypes.c_void_p] + [ctypes.c_int] * 3 + [ctypes.c_void_p, ctypes.c_int] + [ctypes.c_void_p] * 5),
    "dirt_diffuse_directional": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, _i64, ctypes.c_void_p, _i64] +
                                 [ctypes.c_int] * 4 + [ctypes.c_void_p, ctypes.c_void_p]),
    "dirt_diffuse_directional_grad": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, _i64, ctypes.c_void_p, _i64] +
                                      [ctypes.c_int] * 4 + [ctypes.c_void_p] * 4),
    "dirt_diffuse_point": (ctypes.c_int, [ctypes.POINTER(DirtVertexShadeArgs), ctypes.c_void_p]),
    "dirt_diffuse_point_grad": (ctypes.c_int, [ctypes.POINTER(DirtVertexShadeArgs), ctypes.c_void_p]),
    "dirt_specular_directional": (ctypes.c_int, [ctypes.POINTER(DirtVertexShadeArgs), ctypes.c_void_p]),
    "dirt_specular_directional_grad": (ctypes.c_int, [ctypes.POINTER(DirtVertexShadeArgs), ctypes.c_void_p]),
    "dirt_dilate_silhouette": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p] + [ctypes.c_int] * 4 + [ctypes.c_void_p]),
    "dirt_dilate_silhouette_grad": (ctypes.c_int, [ctypes.c_void_p] * 3 + [ctypes.c_int] * 4 + [ctypes.c_void_p]),
    "dirt_kernel_launch_count": (ctypes.c_uint64, []),
    "dirt_profile_enable": (ctypes.c_int, [ctypes.c_int]),
    "dirt_profile_collect": (ctypes.c_int, []),
    "dirt_profile_read": (ctypes.c_int, [ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_uint64), ctypes.c_int]),
    "dirt_profile_reset": (ctypes.c_int, []),
    "dirt_profile_stage_name": (ctypes.c_char_p, [ctypes.c_int]),
}
STAGE_COUNT = 12

_lib = None


class DirtError(RuntimeError):
    """A failing C-ABI call; `.status` carries the DirtStatus code."""

    def __init__(self, status, message):
        super().__init__("%s: %s" % (STATUS_NAMES.get(status, status), message))
        self.status = status


class InvalidArgumentError(DirtError, ValueError):
    """Mirrors tf.errors.InvalidArgumentError raised by the reference's OP_REQUIRES checks
    (csrc/rasterise_egl.cpp:310-336)."""


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            "libdirt_b200.so not found at %s: build it with `python -m dirt_b200.build`. "
            "dirt_b200 has no CPU or pure-Python fallback." % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    for name, (restype, argtypes) in EXPORTS.items():
        fn = getattr(lib, name)
        fn.restype = restype
        fn.argtypes = argtypes
    if lib.dirt_abi_version() != ABI_VERSION:
        raise ImportError("libdirt_b200.so ABI %d, expected %d" % (lib.dirt_abi_version(), ABI_VERSION))
    _lib = lib
    return lib


def check(status):
    if status == 0:
        return
    msg = (load().dirt_last_error() or b"").decode("utf-8", "replace")
    if status == -1:
        raise InvalidArgumentError(status, msg)
    raise DirtError(status, msg)


def launch_count():
    return int(load().dirt_kernel_launch_count())


def stage_times():
    """{stage name: (total ms, launches)} accumulated by dirt_profile_collect()."""
    lib = load()
    ms = (ctypes.c_double * STAGE_COUNT)()
    calls = (ctypes.c_uint64 * STAGE_COUNT)()
    check(lib.dirt_profile_read(ms, calls, STAGE_COUNT))
    return {lib.dirt_profile_stage_name(i).decode(): (ms[i], int(calls[i])) for i in range(STAGE_COUNT)}
