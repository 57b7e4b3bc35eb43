"""Deterministic scene builders (numpy) for tests, smoke() and bench.py.

Each builder returns a dict with batched arrays: background [B,H,W,C] f32, vertices [B,V,4] f32
(clip space), vertex_colors [B,V,C] f32, faces [B,F,3] i32.  Sources in the reference:
  kat1_square      README.md:46-70
  kat2_square      tests/multi_gpu_test.py:8-17
  cylinder         tests/rasterise_tests.py:11-89
  cube             samples/simple.py:15-74 (matrices: dirt/matrices.py:15-170)
  sphere           SURVEY.md section 8(d), c4 / headline recipe
"""
import numpy as np


# ---- row-vector matrix helpers, restating dirt/matrices.py for numpy fixtures --------------
def rodrigues(axis_angle):
    """dirt/matrices.py:15-56 (3x3, row-vector convention: x' = x @ R)."""
    a = np.asarray(axis_angle, np.float64)
    theta = np.linalg.norm(a)
    if theta < 1e-12:
        return np.eye(3)
    k = a / theta
    K = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
    R = np.eye(3) + np.sin(theta) * K + (1 - np.cos(theta)) * (K @ K)
    return R.T


def translation(t):
    """dirt/matrices.py:59-83."""
    M = np.eye(4)
    M[3, :3] = t
    return M


def pad_3x3_to_4x4(R):
    M = np.eye(4)
    M[:3, :3] = R
    return M


def perspective_projection(near, far, right, aspect):
    """dirt/matrices.py:86-116 (camera looks along -z; row-vector convention)."""
    top = right * aspect
    return np.array([
        [near / right, 0, 0, 0],
        [0, near / top, 0, 0],
        [0, 0, -(far + near) / (far - near), -1],
        [0, 0, -2 * far * near / (far - near), 0]], np.float64)


def _pack(background, vertices, colors, faces):
    f32, i32 = np.float32, np.int32
    return dict(background=np.ascontiguousarray(background, f32), vertices=np.ascontiguousarray(vertices, f32),
                vertex_colors=np.ascontiguousarray(colors, f32), faces=np.ascontiguousarray(faces, i32))


def kat1_square():
    W = H = 128
    cx, cy, s = 32, 64, 16
    sq = np.array([[0, 0], [0, 1], [1, 1], [1, 0]], np.float32) * s - s / 2. + [cx, cy]
    sq = sq * 2. / np.array([W, H], np.float32) - 1.
    v = np.concatenate([sq, np.zeros((4, 1)), np.ones((4, 1))], 1)
    expected = np.zeros((H, W), np.float32)
    expected[56:72, 24:40] = 1
    sc = _pack(np.zeros((1, H, W, 1)), v[None], np.ones((1, 4, 1)), np.array([[[0, 1, 2], [0, 2, 3]]]))
    sc["expected"] = expected[None, :, :, None]
    return sc


def kat2_square():
    sq = np.array([[0, 0], [0, 1], [1, 1], [1, 0]], np.float32)
    v = np.concatenate([sq, np.zeros((4, 1)), np.ones((4, 1))], 1)
    expected = np.zeros((1, 256, 256, 3), np.float32)
    expected[0, 0:128, 128:256] = 1
    sc = _pack(np.zeros((1, 256, 256, 3)), v[None], np.ones((1, 4, 3)), np.array([[[0, 1, 2], [0, 2, 3]]]))
    sc["expected"] = expected
    return sc


def cube_mesh():
    """samples/simple.py:15-23, split by face -> V=36? no: 24 corners would do, the sample splits
    per triangle (split_vertices_by_face, dirt/lighting.py:136-179) -> V = 36, F = 12."""
    corners = np.array([[x, y, z] for z in (-1, 1) for y in (-1, 1) for x in (-1, 1)], np.float64)
    quads = [(0, 1, 3, 2), (4, 6, 7, 5), (0, 4, 5, 1), (2, 3, 7, 6), (1, 5, 7, 3), (0, 2, 6, 4)]
    faces = []
    for a, b, c, d in quads:
        faces += [(a, b, c), (a, c, d)]
    faces = np.array(faces)
    verts = corners[faces.reshape(-1)]
    return verts, np.arange(36).reshape(12, 3)


def cube(B=1, H=256, W=256, C=3, seed=0, rot0=0.5, drot=0.1):
    """Gouraud-lit cube (config c2 / c3 geometry): per-element y rotation rot0 + drot*b."""
    verts, faces = cube_mesh()
    rng = np.random.RandomState(seed)
    # Lambert shading per vertex with face normals (samples/simple.py:63-66)
    tri = verts.reshape(12, 3, 3)
    n = np.cross(tri[:, 1] - tri[:, 0], tri[:, 2] - tri[:, 0])
    n /= np.linalg.norm(n, axis=1, keepdims=True)
    normals = np.repeat(n, 3, axis=0)
    base = rng.uniform(0.3, 1.0, size=(36, C))
    P = perspective_projection(0.1, 20., 0.1, H / W)
    vs, cs = [], []
    for b in range(B):
        R = pad_3x3_to_4x4(rodrigues([0., rot0 + drot * b, 0.])) @ pad_3x3_to_4x4(rodrigues([0.6, 0., 0.]))
        M = R @ translation([0., 0., -4.5]) @ P
        vh = np.concatenate([verts, np.ones((36, 1))], 1) @ M
        light = np.array([1., 1., 1.]) / np.sqrt(3.)
        nn = normals @ R[:3, :3]
        lam = np.clip(nn @ light, 0, None)[:, None] * 0.8 + 0.2
        vs.append(vh)
        cs.append(base * lam)
    bg = rng.uniform(0, 1, size=(B, H, W, C))
    return _pack(bg, np.stack(vs), np.stack(cs), np.broadcast_to(faces, (B, 12, 3)))


def cylinder(H=36, W=48, segments=16, seed=0):
    """Open cylinder seen slightly from above (tests/rasterise_tests.py:11-89 geometry)."""
    ang = np.linspace(0, 2 * np.pi, segments, endpoint=False)
    ring = np.stack([np.cos(ang), np.zeros_like(ang), np.sin(ang)], 1)
    verts = np.concatenate([ring + [0, -1, 0], ring + [0, 1, 0]], 0)
    faces = []
    for i in range(segments):
        j = (i + 1) % segments
        faces += [(i, j, segments + i), (j, segments + j, segments + i)]
    faces = np.array(faces)
    R = pad_3x3_to_4x4(rodrigues([0.5, 0., 0.]))
    M = R @ translation([0., 0., -5.]) @ perspective_projection(0.1, 20., 0.05, H / W)
    vh = np.concatenate([verts, np.ones((len(verts), 1))], 1) @ M
    rng = np.random.RandomState(seed)
    cols = rng.uniform(0, 1, size=(len(verts), 3))
    bg = rng.uniform(0, 1, size=(1, H, W, 3))
    return _pack(bg, vh[None], cols[None], faces[None])


def uv_sphere_mesh(n_lat, n_lon):
    """Closed UV sphere: V = (n_lat-1)*n_lon + 2, F = 2*n_lon*(n_lat-1)."""
    lat = np.linspace(0, np.pi, n_lat + 1)[1:-1]
    lon = np.linspace(0, 2 * np.pi, n_lon, endpoint=False)
    la, lo = np.meshgrid(lat, lon, indexing="ij")
    ring = np.stack([np.sin(la) * np.cos(lo), np.cos(la), np.sin(la) * np.sin(lo)], -1).reshape(-1, 3)
    verts = np.concatenate([[[0, 1, 0]], ring, [[0, -1, 0]]], 0)
    idx = lambda i, j: 1 + i * n_lon + (j % n_lon)
    faces = []
    for j in range(n_lon):
        faces.append((0, idx(0, j + 1), idx(0, j)))
    i, j = np.meshgrid(np.arange(n_lat - 2), np.arange(n_lon), indexing="ij")
    a = 1 + i * n_lon + j
    b = 1 + i * n_lon + (j + 1) % n_lon
    c = 1 + (i + 1) * n_lon + j
    d = 1 + (i + 1) * n_lon + (j + 1) % n_lon
    body = np.concatenate([np.stack([a, b, c], -1).reshape(-1, 3), np.stack([b, d, c], -1).reshape(-1, 3)], 0)
    south = len(verts) - 1
    cap = [(south, idx(n_lat - 2, j), idx(n_lat - 2, j + 1)) for j in range(n_lon)]
    faces = np.concatenate([np.array(faces), body, np.array(cap)], 0)
    return verts, faces.astype(np.int64)


def sphere(B, H, W, C=3, n_lat=224, n_lon=224, fill=0.6, seeds=(1, 2, 3, 4)):
    """Bumpy sphere, random per-element rotation (SURVEY 8(d) c4/headline recipe):
    n_lat = n_lon = 224 gives F = 99,904 ~ 100k faces, V = 49,954."""
    return sphere_frames(B, H, W, C, n_lat, n_lon, fill, seeds, frame=0)


def sphere_frames(B, H, W, C=3, n_lat=224, n_lon=224, fill=0.6, seeds=(1, 2, 3, 4), frame=0, step_deg=0.25, noise_px=0.0,
                  vertices_only=False):
    """Frame `frame` of a sequence that starts at sphere(): every element's object turns by a further
    frame * step_deg degrees about the camera's y axis, and (noise_px > 0) every vertex is displaced by
    Gaussian noise worth ~noise_px pixels on screen, fresh per frame -- the way an inverse-rendering loop
    (tests/optimize_sequence.py:293-352) re-renders a scene whose parameters moved a little.  Colours,
    faces and background are those of frame 0."""
    verts, faces = uv_sphere_mesh(n_lat, n_lon)
    V = len(verts)
    verts = verts * (1 + 0.1 * np.random.RandomState(seeds[0]).randn(V, 1).clip(-2.5, 2.5))
    rr = np.random.RandomState(seeds[1])
    z = 3.0
    # choose `right` so a unit sphere's silhouette covers ~fill of the frame area
    r_ndc = np.sqrt(fill * 4 / np.pi) if fill * 4 / np.pi < 1 else 0.98
    near = 0.1
    right = near * (1.0 / np.sqrt(z * z - 1)) / r_ndc
    P = perspective_projection(near, 20., right, H / W)
    if frame and noise_px > 0:
        verts = verts + np.random.RandomState(1000 + frame).randn(V, 3) * (noise_px * 2.0 / (W * r_ndc))
    vh = np.concatenate([verts, np.ones((V, 1))], 1)
    vs = np.empty((B, V, 4), np.float32)
    turn = pad_3x3_to_4x4(rodrigues([0., np.deg2rad(step_deg * frame), 0.]))
    for b in range(B):
        M = pad_3x3_to_4x4(rodrigues(rr.randn(3))) @ turn @ translation([0., 0., -z]) @ P
        vs[b] = vh @ M
    if vertices_only:
        return vs
    cols = np.random.RandomState(seeds[2]).uniform(0, 1, size=(B, V, C)).astype(np.float32)
    bg = np.random.RandomState(seeds[3]).uniform(0, 1, size=(B, H, W, C)).astype(np.float32)
    return _pack(bg, vs, cols, np.broadcast_to(faces, (B,) + faces.shape))


def random_triangles(B, H, W, C, F, seed, size=0.5, depth_ties=False, w_range=(0.5, 2.0)):
    """Triangle soup: random centres, sizes up to `size` NDC, random w (perspective)."""
    rng = np.random.RandomState(seed)
    V = 3 * F
    ctr = rng.uniform(-1.1, 1.1, size=(B, F, 1, 2))
    off = rng.uniform(-size, size, size=(B, F, 3, 2)) * rng.uniform(0.02, 1.0, size=(B, F, 1, 1))
    xy = (ctr + off).reshape(B, V, 2)
    if depth_ties:
        zf = np.round(rng.uniform(-0.9, 0.9, size=(B, F, 1)) * 4) / 4
        z = np.repeat(zf, 3, axis=2).reshape(B, V, 1)
    else:
        z = rng.uniform(-0.95, 0.95, size=(B, V, 1))
    w = rng.uniform(w_range[0], w_range[1], size=(B, V, 1))
    v = np.concatenate([xy * w, z * w, w], 2)
    cols = rng.uniform(-1, 2, size=(B, V, C))
    bg = rng.uniform(0, 1, size=(B, H, W, C))
    faces = np.broadcast_to(np.arange(V).reshape(F, 3), (B, F, 3))
    return _pack(bg, v, cols, faces)


def near_crossing(B=2, H=96, W=128, C=3, seed=5, n=24):
    """A ground-plane grid the camera stands on: many faces cross the near plane or lie behind
    the camera (w <= 0), the config-c5 situation."""
    xs, zs = np.meshgrid(np.linspace(-6, 6, n), np.linspace(-12, 3, n), indexing="ij")
    rng = np.random.RandomState(seed)
    verts = np.stack([xs, -0.7 + 0.15 * rng.randn(n, n), zs], -1).reshape(-1, 3)
    i, j = np.meshgrid(np.arange(n - 1), np.arange(n - 1), indexing="ij")
    a, b_, c, d = i * n + j, i * n + j + 1, (i + 1) * n + j, (i + 1) * n + j + 1
    faces = np.concatenate([np.stack([a, b_, c], -1).reshape(-1, 3), np.stack([b_, d, c], -1).reshape(-1, 3)], 0)
    P = perspective_projection(0.3, 10., 0.25, H / W)   # far plane at 10 also clips the far rows
    vh = np.concatenate([verts, np.ones((len(verts), 1))], 1)
    vs = np.stack([vh @ (pad_3x3_to_4x4(rodrigues([0., 0.3 * b, 0.])) @ P) for b in range(B)])
    cols = rng.uniform(0, 1, size=(B, len(verts), C))
    bg = rng.uniform(0, 1, size=(B, H, W, C))
    return _pack(bg, vs, cols, np.broadcast_to(faces, (B,) + faces.shape))


def _value_noise(x, z, seed=0, octaves=5):
    """Fractal value noise in [0, 1] (the kind of terrain function the fork's hill shader uses,
    csrc/shaders.cpp:188-269, restated with numpy's generator -- not the shader's hash)."""
    rng = np.random.RandomState(seed)
    table = rng.uniform(0, 1, size=(256, 256))
    out = np.zeros_like(x, dtype=np.float64)
    amp, freq, norm = 0.5, 0.25, 0.0
    for _ in range(octaves):
        xs, zs = x * freq, z * freq
        x0, z0 = np.floor(xs).astype(np.int64), np.floor(zs).astype(np.int64)
        fx, fz = xs - x0, zs - z0
        fx, fz = fx * fx * (3 - 2 * fx), fz * fz * (3 - 2 * fz)
        t = lambda i, j: table[i & 255, j & 255]
        v = (t(x0, z0) * (1 - fx) + t(x0 + 1, z0) * fx) * (1 - fz) + (t(x0, z0 + 1) * (1 - fx) + t(x0 + 1, z0 + 1) * fx) * fz
        out += amp * v
        norm += amp
        amp *= 0.5
        freq *= 2.0
    return out / norm


def hill(B, H, W, C=3, n=708, seed=0):
    """Config c5: one heightfield mesh shared by B frames on a smooth camera path
    (SURVEY.md 8(d)): n x n vertices over x in [-14, 14], z in [-25, -5] (the camera looks along
    -z, dirt/matrices.py:86-116), height = noise * 10.3 - 6.1 (csrc/shaders.cpp:222-236); pinhole
    focal 0.85 W (csrc/shaders.cpp:469-471).  n = 708 gives V = 501,264, F = 999,698.
    faces and vertex_colors have batch dim 1 (broadcast); returns also `world` [V,4] and
    `matrices` [B,4,4] so that clip-space vertex gradients can be folded back to world space."""
    xs, zs = np.meshgrid(np.linspace(-14, 14, n), np.linspace(-25, -5, n), indexing="ij")
    h = _value_noise(xs + 40.0, zs + 40.0, seed) * 10.3 - 6.1
    world = np.stack([xs, h * 0.35 - 1.0, zs, np.ones_like(xs)], -1).reshape(-1, 4)
    i, j = np.meshgrid(np.arange(n - 1), np.arange(n - 1), indexing="ij")
    a, b_, c, d = i * n + j, i * n + j + 1, (i + 1) * n + j, (i + 1) * n + j + 1
    faces = np.concatenate([np.stack([a, b_, c], -1).reshape(-1, 3), np.stack([b_, d, c], -1).reshape(-1, 3)], 0)
    near = 0.1
    P = perspective_projection(near, 60., near * 0.5 / 0.85, H / W)
    Ms = np.stack([pad_3x3_to_4x4(rodrigues([0.0, 0.02 * (b - B / 2.0), 0.0])) @ translation([0.05 * b, -1.5, 0.0]) @ P
                   for b in range(B)])
    vs = np.einsum("vi,bij->bvj", world, Ms)
    rng = np.random.RandomState(seed + 1)
    cols = rng.uniform(0, 1, size=(1, len(world), C))
    bg = rng.uniform(0, 1, size=(B, H, W, C))
    sc = _pack(bg, vs, cols, faces[None])
    sc["world"] = world.astype(np.float32)
    sc["matrices"] = Ms.astype(np.float32)
    return sc


# Three faces with a corner behind the eye (w < 0: outside the near AND the far plane).  Clipping leaves
# polygon corners with w -> 0 whose window positions hit the guard-band clamp, so the fan folds over
# itself and two of its triangles cover the same pixels (found by tests/test_fuzz_gpu.py, seeds 18 and 23).
FOLDED_FAN_FACES = [
    [[-4.520687580108643, -6.046390533447266, -1.3254375457763672, 1.7913875579833984],
     [0.820976972579956, 0.5507926344871521, 0.3445732891559601, 0.39526867866516113],
     [0.5961543321609497, 0.04113024100661278, -0.1284199357032776, -0.191458597779274]],
    [[0.04086809232831001, 0.1765381395816803, -0.10965538769960403, -0.1991819143295288],
     [0.2568352222442627, 0.22824278473854065, 0.24783971905708313, -0.2613210082054138],
     [-0.15913863480091095, 0.3746587634086609, -0.5689247250556946, 0.8287256956100464]],
    [[0.7668004035949707, -0.1253456324338913, -1.4441827535629272, 1.92557692527771],
     [0.86167973279953, 0.03097062185406685, -1.4628005027770996, 1.9504005908966064],
     [-0.02990591526031494, 0.011887836270034313, 0.031696148216724396, -0.04226153343915939]],
]


def folded_fans(H=74, W=76, C=3, seed=0):
    rng = np.random.RandomState(seed)
    v = np.asarray(FOLDED_FAN_FACES, np.float32).reshape(1, 9, 4)
    return _pack(rng.uniform(0, 1, (1, H, W, C)), v, rng.uniform(0, 1, (1, 9, C)), np.arange(9).reshape(1, 3, 3))
