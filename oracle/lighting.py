"""TEST INFRASTRUCTURE ONLY (see oracle/README or DESIGN.md 4): numpy float64 restatement of the
reference's geometry / lighting helpers, dirt/lighting.py and dirt/matrices.py, function by function.
Used by tests/ to check dirt_b200/lighting.py, dirt_b200/matrices.py and the vertex-normal kernels.
Parity status: pinned only by hand-derived cases (a unit cube, axis rotations, a known projection)
in tests/test_lighting.py -- the reference module needs TensorFlow 1.x and cannot be imported here.
"""
import numpy as np

EPS = 1.e-12


def face_normals(vertices, faces):
    """_get_face_normals, dirt/lighting.py:24-31.  vertices [*, V, 3|4], faces [F, 3] -> [*, F, 3]."""
    p = np.asarray(vertices, np.float64)[..., :3]
    f = np.asarray(faces)
    m = np.cross(p[..., f[:, 1], :] - p[..., f[:, 0], :], p[..., f[:, 2], :] - p[..., f[:, 0], :])
    return m / (np.linalg.norm(m, axis=-1, keepdims=True) + EPS)


def vertex_normal_sums(vertices, faces):
    """The per-vertex sum of unit face normals (dirt/lighting.py:66-88; also the whole of
    vertex_normals_pre_split, :101-133, for a pre-split mesh)."""
    n = face_normals(vertices, faces)
    f = np.asarray(faces)
    out = np.zeros(np.asarray(vertices).shape[:-1] + (3,), np.float64)
    flat_out = out.reshape(-1, out.shape[-2], 3)
    flat_n = n.reshape(-1, n.shape[-2], 3)
    for b in range(flat_out.shape[0]):
        for k in range(3):
            np.add.at(flat_out[b], f[:, k], flat_n[b])
    return out


def vertex_normals(vertices, faces):
    """dirt/lighting.py:34-93."""
    s = vertex_normal_sums(vertices, faces)
    return s / (np.linalg.norm(s, axis=-1, keepdims=True) + EPS)


def split_vertices_by_face(vertices, faces):
    """dirt/lighting.py:136-179."""
    v = np.asarray(vertices)
    f = np.asarray(faces)
    return v[..., f.reshape(-1), :], np.arange(f.shape[0] * 3, dtype=np.int32).reshape(-1, 3)


def _cos(c, double_sided):
    return np.abs(c) if double_sided else np.maximum(c, 0.)


def diffuse_directional(normals, colors, light_direction, light_color, double_sided=True):
    """dirt/lighting.py:182-225."""
    n, c, d, lc = (np.asarray(x, np.float64) for x in (normals, colors, light_direction, light_color))
    cosines = n @ (-d[..., None])
    return lc[..., None, :] * c * _cos(cosines, double_sided)


def specular_directional(positions, normals, reflectivities, light_direction, light_color, camera_position, shininess,
                         double_sided=True):
    """dirt/lighting.py:228-288 (the 1e-12 lands after the division, as written there)."""
    p, n, r, d, lc, cam, sh = (np.asarray(x, np.float64) for x in (positions, normals, reflectivities, light_direction,
                                                                   light_color, camera_position, shininess))
    to_light = -d
    reflected = -to_light[..., None, :] + 2. * (n @ to_light[..., None]) * n
    to_camera = cam[..., None, :] - p
    cosines = ((to_camera / np.linalg.norm(to_camera, axis=-1, keepdims=True) + EPS) * reflected).sum(-1, keepdims=True)
    return lc[..., None, :] * r * np.power(_cos(cosines, double_sided), sh[..., None, None])


def diffuse_point(positions, normals, colors, light_position, light_color, double_sided=True):
    """dirt/lighting.py:291-343."""
    p, n, c, lp, lc = (np.asarray(x, np.float64) for x in (positions, normals, colors, light_position, light_color))
    rel = p - lp[..., None, :]
    inc = rel / (np.linalg.norm(rel, axis=-1, keepdims=True) + EPS)
    cosines = (n * inc).sum(-1)
    return lc[..., None, :] * c * _cos(cosines, double_sided)[..., None]


# ---- dirt/matrices.py -----------------------------------------------------------------------

def pad_3x3_to_4x4(m):
    """dirt/matrices.py:119-143."""
    m = np.asarray(m, np.float64)
    out = np.zeros(m.shape[:-2] + (4, 4))
    out[..., :3, :3] = m
    out[..., 3, 3] = 1.
    return out


def rodrigues(vectors):
    """dirt/matrices.py:15-56."""
    v = np.asarray(vectors, np.float64) + EPS
    norms = np.linalg.norm(v, axis=-1, keepdims=True)
    v = v / norms
    norms = norms[..., 0]
    z = np.zeros_like(v[..., 0])
    K = np.stack([np.stack([z, -v[..., 2], v[..., 1]], -1),
                  np.stack([v[..., 2], z, -v[..., 0]], -1),
                  np.stack([-v[..., 1], v[..., 0], z], -1)], -2)
    c, s = np.cos(norms)[..., None, None], np.sin(norms)[..., None, None]
    return pad_3x3_to_4x4(c * np.eye(3) + (1 - c) * v[..., :, None] * v[..., None, :] + s * K)


def translation(x):
    """dirt/matrices.py:59-83."""
    x = np.asarray(x, np.float64)
    out = np.zeros(x.shape[:-1] + (4, 4))
    out[..., np.arange(4), np.arange(4)] = 1.
    out[..., 3, :3] = x
    return out


def perspective_projection(near, far, right, aspect):
    """dirt/matrices.py:86-116."""
    top = right * aspect
    e = np.array([[near / right, 0., 0., 0.],
                  [0., near / top, 0., 0.],
                  [0., 0., -(far + near) / (far - near), -2. * far * near / (far - near)],
                  [0., 0., -1., 0.]])
    return e.T


def compose(*matrices):
    """dirt/matrices.py:146-170."""
    out = np.eye(4)
    for m in matrices:
        out = out @ np.asarray(m, np.float64)
    return out
