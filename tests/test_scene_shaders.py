"""The fork's other scene ops (oceanic, oceanic_still_cloud, oceanic_no_cloud,
oceanic_simple_proxy, oceanic_opt_flow, hill): CUDA per-pixel stages behind the new rasteriser
(dirt_b200/csrc/scene_shaders.cu) against the C++ restatement of the GLSL (oracle/scene_shaders.cpp).

What pins the restatement (the reference stores no image of any of these ops):
  - oceanic_opt_flow with dt = 0 must return every pixel's own window coordinates -- an analytic
    known answer that exercises the ray set-up, the rotation and its transpose;
  - hill over a flat terrain: sky pixels must equal GetSky + PostEffects written independently in
    numpy float64, and the ground/sky boundary must sit on the analytic horizon row.

Tolerance of the CUDA comparison: the shaders hash with fract(cos(n) * 41415.9) and
fract(sin(n) * 43758.5), so one ulp of difference between two sin / cos implementations moves a
hash by ~3e-3 and the chaotic march with it (measured with libm sinf vs CUDA sinf: 30 % of the
pixels off by more than 2e-2).  Both sides therefore take sin / cos correctly rounded (double
evaluation, one rounding); what is left is powf / expf and is measured at <= 4e-7.  The bar: mean
absolute difference <= 1e-6 and at most 0.2 % of the pixels beyond 1e-4 (room for the ~1e-9 of
calls where the two double sines straddle a rounding boundary and a march takes another branch).
"""
import numpy as np
import pytest

from oracle import golden

QUAD_V = np.array([[-1, -1, 0, 1], [-1, 1, 0, 1], [1, 1, 0, 1], [1, -1, 0, 1]], np.float32)
QUAD_F = np.array([[0, 1, 2], [0, 2, 3]], np.int32)


def quad(B, H, W, C, seed=0, noise=0.0):
    rng = np.random.RandomState(seed)
    bg = (rng.randn(B, H, W, C) * noise).astype(np.float32)
    return bg, np.broadcast_to(QUAD_V, (B, 4, 4)).copy(), np.ones((B, 4, C), np.float32), np.broadcast_to(QUAD_F, (B, 2, 3)).copy()


def terrain_texture(H, W, flat=None):
    """[H,W,4] = (height in [0,1], normal * 0.5 + 0.5).  Texture s (columns) spans world z in
    [5,25]; t (rows, bottom row = t 0 because the upload flips) spans world x in [-14,14]."""
    t = np.linspace(1, 0, H)[:, None]
    s = np.linspace(0, 1, W)[None, :]
    if flat is not None:
        h = np.full((H, W), flat)
    else:
        h = 0.55 + 0.08 * np.sin(6 * s + 1) * np.cos(5 * t) + 0.05 * np.sin(13 * s * t)
    hw = h * 10.3 - 6.1
    dz = np.gradient(hw, axis=1) / (20.0 / (W - 1))
    dx = np.gradient(hw, axis=0) / (-28.0 / (H - 1))
    n = np.stack([-dx, np.ones_like(hw), -dz], -1)
    n /= np.linalg.norm(n, axis=-1, keepdims=True)
    return np.concatenate([h[..., None], (n + 1) / 2], -1).astype(np.float32)


# cam_x, cam_y, cam_z, ang1, ang2, ang3, time, light_z, cloud_t, dt, dx, dy, dz, dang1, dang2, dang3
CAM_A = [0.0, 100.0, 0.0, 0.15, 0.3, 0.0, 1.0, 0.9, 0.5, 1 / 30.0, 30.0, 14.0, 15.0, 0.1, 0.1, 0.1]
CAM_B = [30.0, 140.0, -20.0, 0.12, 0.05, -0.08, 3.5, 0.4, 2.0, 0.1, -10.0, 5.0, 40.0, -0.2, 0.05, 0.3]
OCEAN_SCENES = ["oceanic", "oceanic_still_cloud", "oceanic_no_cloud", "oceanic_simple_proxy", "oceanic_opt_flow"]
HILL_CAM = np.array([0] * 9 + [0.0, 0.0, 3.0], np.float32)            # o0, o1, o2 -> camera at (0, 3, 0)


# ---- CPU: the oracle itself ---------------------------------------------------------------------

def test_opt_flow_without_motion_returns_each_pixels_own_coordinates():
    H, W = 40, 64
    bg, v, c, f = quad(1, H, W, 3)
    for cam in (CAM_A, CAM_B):
        cam = list(cam)
        cam[9] = 0.0                                               # dt = 0: previous frame == this frame
        px = golden.scene("oceanic_opt_flow", bg, v, f, cam)[0]
        xs = np.arange(W) + 0.5
        ys = H - (np.arange(H) + 0.5)                              # tensors are top row first, GL window y is bottom up
        assert np.abs(px[..., 0] - xs[None, :]).max() < 2e-3
        assert np.abs(px[..., 1] - ys[:, None]).max() < 2e-3
        assert (px[..., 2] == 0).all()


def test_opt_flow_pure_translation_moves_sea_pixels_only():
    H, W = 40, 64
    bg, v, c, f = quad(1, H, W, 3)
    cam = [0.0, 100.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.9, 0.0, 0.5, 0.0, 0.0, 20.0, 0.0, 0.0, 0.0]   # forward motion, level camera
    px = golden.scene("oceanic_opt_flow", bg, v, f, cam)[0]
    xs = np.arange(W) + 0.5
    ys = H - (np.arange(H) + 0.5)
    sky, sea = slice(0, H // 2), slice(H // 2, H)
    assert np.abs(px[sky, :, 0] - xs).max() < 2e-3 and np.abs(px[sky, :, 1] - ys[sky, None]).max() < 2e-3   # sky is at infinity
    # the camera was 10 units further back one step ago: sea points were closer to the horizon (image centre)
    assert (np.abs(px[sea, :, 1] - H / 2) < np.abs(ys[sea, None] - H / 2)).all()
    assert (np.abs(px[sea, :, 0] - W / 2) <= np.abs(xs - W / 2) + 1e-3).all()


def test_ocean_scenes_basic_properties():
    H, W = 36, 64
    bg, v, c, f = quad(1, H, W, 3)
    cam = [0.0, 100.0, 0.0, 0.0, 0.0, 0.0, 1.0, 0.9, 0.5]
    out = {n: golden.scene(n, bg, v, f, cam)[0] for n in OCEAN_SCENES[:4]}
    for n, px in out.items():
        assert np.isfinite(px).all(), n
        assert px[: H // 2, :, 2].mean() > px[H // 2:, :, 2].mean(), n      # blue sky over a darker sea
    # without cloud layers the sky is a closed form of the ray (csrc/shaders.cpp:1617-1624): check it in float64
    ys, xs = np.meshgrid(np.arange(H) + 0.5, np.arange(W) + 0.5, indexing="ij")
    ray = np.stack([xs - W / 2, (H - ys) - H / 2, np.full_like(xs, 1.73 * W / 2)], -1)
    ray /= np.linalg.norm(ray, axis=-1, keepdims=True)              # level camera: rd == ray
    light = np.array([0.1, 0.25, 0.9]); light /= np.linalg.norm(light)
    sd = np.clip(ray @ light, 0, 1)[..., None]
    t = ((1 - 0.7 * ray[..., 1]) ** 15)[..., None]
    sky = 0.8 * (np.array([0.6, 0.8, 1.2]) * t + np.array([0.05, 0.2, 0.5]) * (1 - t))
    sky = sky + 0.47 * np.array([1.6, 1.4, 1.0]) * sd ** 350 + 0.4 * np.array([0.8, 0.9, 1.0]) * sd ** 2
    up = ray[..., 1] > 0
    assert np.abs(out["oceanic_no_cloud"][up] - sky[up]).max() < 1e-4
    # the two cloud placements differ in the sky but share the water surface geometry
    assert not np.array_equal(out["oceanic"][: H // 2], out["oceanic_still_cloud"][: H // 2])
    # half-covered frame: background survives where nothing is drawn
    v2 = v.copy(); v2[..., 0] = v2[..., 0] * 0.5 - 0.5
    px2 = golden.scene("oceanic_no_cloud", np.full_like(bg, 7.0), v2, f, cam)[0]
    assert (px2[:, 40:, :] == 7.0).all() and (px2[:, :30, :] != 7.0).all()
    with pytest.raises(ValueError):
        golden.scene("oceanic_still_cloud", bg, v, f, cam[:8])                  # needs cloud_t


def _hill_sky_numpy(H, W):
    """GetSky + PostEffects (csrc/shaders.cpp:273-281, :433-447) in float64 for every pixel."""
    ys, xs = np.meshgrid(np.arange(H) + 0.5, np.arange(W) + 0.5, indexing="ij")
    tx = xs / W * 2 - 1
    ty = (H - ys) / H * 2 - 1                                       # clip y of the pixel centre (row 0 is the top)
    x01, y01 = (tx + 1) / 2, (-ty + 1) / 2
    ray = np.stack([x01 * W - W / 2, y01 * H - H / 2, np.full_like(x01, 0.85 * W)], -1)
    ray /= np.linalg.norm(ray, axis=-1, keepdims=True)
    R = np.array([[0.999999573, -0.0000933038802, 0.000919791287], [0.000918443273, -0.0135434586, -0.999907861],
                  [0.000105752439, 0.999908279, -0.0135433672]])
    d = ray @ R                                                     # (r0k x + r1k y + r2k z) for k = 0, 1, 2
    dirv = np.stack([-d[..., 0], d[..., 2], d[..., 1]], -1)
    sun = np.array([0.35, 0.2, 0.3]); sun /= np.linalg.norm(sun)
    sa = np.maximum(dirv @ sun, 0)
    vv = (1 - np.maximum(dirv[..., 1], 0)) ** 6
    sky = np.array([.1, .2, .3]) * (1 - vv[..., None]) + np.array([.32, .32, .32]) * vv[..., None]
    sc = np.array([1.0, .75, .6])
    sky = sky + sc * (sa * sa * .25)[..., None] + sc * np.minimum(sa ** 800 * 1.5, .3)[..., None]
    sky = np.clip(sky, 0, 1) ** 0.45
    b = sky * 1.3
    lum = b @ np.array([.2125, .7154, .0721])
    sat = lum[..., None] * (1 - 1.3) + b * 1.3
    rgb = 0.5 * (1 - 1.1) + sat * 1.1
    vig = .4 + .5 * (40 * x01 * y01 * (1 - x01) * (1 - y01)) ** 0.2
    return rgb * vig[..., None], dirv


def test_hill_flat_terrain_sky_and_horizon():
    H, W = 54, 96
    _, v, c, f = quad(1, H, W, 3)
    tex = terrain_texture(H, W, flat=0.5)[None]                     # ground plane at y = 0.5 * 10.3 - 6.1 = -0.95
    px = golden.scene("hill", tex, v, f, HILL_CAM)[0]
    assert px.shape == (H, W, 3) and np.isfinite(px).all()
    sky, dirv = _hill_sky_numpy(H, W)
    up = dirv[..., 1] > 0.02                                        # rays that cannot reach the ground
    assert up.sum() > H * W // 3
    assert np.abs(px[up] - sky[up]).max() < 2e-5
    # below the horizon, inside the texture's extent, the march converges on the plane: grass (green dominant)
    down = (dirv[..., 1] < -0.12) & (np.abs(dirv[..., 0]) < 0.2)
    assert down.sum() > 50 and (px[down][:, 1] > px[down][:, 2]).mean() > 0.95
    # a quad over the left half only: the rest of the frame is 0
    v2 = v.copy(); v2[..., 0] = v2[..., 0] * 0.5 - 0.5
    px2 = golden.scene("hill", tex, v2, f, HILL_CAM)[0]
    left = up.copy(); left[:, W // 2 - 2:] = False
    assert (px2[:, W // 2 + 2:, :] == 0).all() and np.abs(px2[left] - px[left]).max() < 1e-5


def test_hill_texture_orientation_and_channel_expansion():
    """The terrain is sampled at (s, t) = (z, x) scaled, rows flipped: raising the terrain on the
    camera's right (world x > 0 is screen left, dir.x = -ray.x) must raise the left side of the picture."""
    H, W = 54, 96
    _, v, c, f = quad(1, H, W, 3)
    tex = terrain_texture(H, W, flat=0.5)
    bump = tex.copy()
    bump[: H // 3, :, 0] = 0.75                                     # tensor rows near the top = texture t near 1 = world x near +14
    a = golden.scene("hill", tex[None], v, f, HILL_CAM)[0]
    b = golden.scene("hill", bump[None], v, f, HILL_CAM)[0]
    changed = np.abs(a - b).max(axis=-1) > 1e-3
    assert changed[:, : W // 2].sum() > 20 * max(changed[:, W // 2:].sum(), 1)
    # 1-channel texture = (h, h, h, 1): same heights, other normal -> same ground/sky split
    g1 = golden.scene("hill", tex[None, ..., :1], v, f, HILL_CAM)[0]
    sky, dirv = _hill_sky_numpy(H, W)
    up = dirv[..., 1] > 0.02
    assert np.abs(g1[up] - a[up]).max() == 0


# ---- GPU: CUDA stage vs the oracle ----------------------------------------------------------------

def _stats(out, ref, bound):
    """Mean absolute difference and the share of pixels beyond `bound`.  The shaders can produce NaN
    (pow of a negative base in hill's PostEffects, csrc/shaders.cpp:436): NaN in both counts as equal,
    NaN in one of them as a pixel beyond the bound."""
    both = np.isnan(out) & np.isnan(ref)
    diff = np.where(both, 0.0, np.abs(out - ref))
    diff = np.where(np.isnan(diff), np.inf, diff)
    d = diff.max(axis=-1)
    finite = np.isfinite(diff)
    return float(diff[finite].mean()), float((d > bound).mean())


BARS = {n: (1e-6, 1e-4, 0.002) for n in OCEAN_SCENES}


@pytest.mark.gpu
@pytest.mark.parametrize("cam", [CAM_A, CAM_B])
@pytest.mark.parametrize("name", OCEAN_SCENES)
def test_cuda_ocean_scenes_match_oracle(name, cam):
    import torch
    import dirt_b200
    B, H, W, C = 2, 72, 128, 3
    bg, v, c, f = quad(B, H, W, C, seed=1, noise=0.5)
    ref = golden.scene(name, bg, v, f, cam)
    fn = getattr(dirt_b200, name + "_batch")
    out = fn(torch.from_numpy(bg).cuda(), torch.from_numpy(v).cuda(), torch.from_numpy(c).cuda(), torch.from_numpy(f).cuda(), cam).cpu().numpy()
    assert out.shape == ref.shape and np.isfinite(out).all()
    mean_bar, px_bar, share = BARS[name]
    scale = max(1.0, float(np.abs(ref).max())) if name == "oceanic_opt_flow" else 1.0
    mean, beyond = _stats(out / scale, ref / scale, px_bar)
    assert mean <= mean_bar and beyond <= share, (name, mean, beyond)
    one = getattr(dirt_b200, name)(torch.from_numpy(bg[0]).cuda(), v[0], c[0], f[0], np.array(cam, np.float32))
    assert np.array_equal(one.cpu().numpy(), out[0])


@pytest.mark.gpu
def test_cuda_ocean_scene_single_channel_and_partial_cover():
    import torch
    import dirt_b200
    B, H, W = 1, 48, 80
    bg, v, c, f = quad(B, H, W, 1, seed=2, noise=0.3)
    v[..., 0] = v[..., 0] * 0.5 - 0.5
    bg += 5.0
    ref = golden.scene("oceanic_no_cloud", bg, v, f, CAM_A)
    out = dirt_b200.oceanic_no_cloud_batch(torch.from_numpy(bg).cuda(), v, c, f, CAM_A).cpu().numpy()
    assert np.array_equal(out[:, :, W // 2 + 1:], bg[:, :, W // 2 + 1:])
    mean, beyond = _stats(out, ref, 1e-4)
    assert mean <= 1e-6 and beyond <= 0.002, (mean, beyond)


@pytest.mark.gpu
@pytest.mark.parametrize("Cb,C", [(4, 3), (1, 3), (3, 1)])
def test_cuda_hill_matches_oracle(Cb, C):
    import torch
    import dirt_b200
    B, H, W = 2, 72, 128
    _, v, c, f = quad(B, H, W, C)
    tex = np.stack([terrain_texture(H, W), terrain_texture(H, W)[::-1].copy()])[..., :Cb]
    if Cb == 3:
        tex = np.ascontiguousarray(tex)
    ref = golden.scene("hill", tex, v, f, HILL_CAM, channels=C)
    out = dirt_b200.hill_batch(torch.from_numpy(np.ascontiguousarray(tex)).cuda(), torch.from_numpy(v).cuda(), torch.from_numpy(c).cuda(),
                               torch.from_numpy(f).cuda(), HILL_CAM, channels=C).cpu().numpy()
    assert out.shape == ref.shape == (B, H, W, C) and np.isfinite(out).all()
    mean, beyond = _stats(out, ref, 1e-4)
    assert mean <= 1e-6 and beyond <= 0.002, (mean, beyond)
    one = dirt_b200.hill(torch.from_numpy(np.ascontiguousarray(tex[0])).cuda(), v[0], c[0], f[0], HILL_CAM, channels=C)
    assert np.array_equal(one.cpu().numpy(), out[0])


@pytest.mark.gpu
def test_cuda_scenes_on_perspective_geometry_and_broadcast_inputs():
    """texCoordV is a `smooth` (perspective-correct) varying: a quad with different w per corner, drawn
    twice at different depths, exercises the interpolation and the depth test under the shader; the
    terrain texture and the faces are given once and broadcast over a batch of two."""
    import torch
    import dirt_b200
    H, W = 60, 96
    v = np.array([[[-1.6, -0.9, 0.3, 2.0], [-0.7, 0.9, 0.1, 1.0], [0.9, 0.8, 0.2, 1.0], [2.4, -2.1, 0.9, 3.0],
                   [-0.5, -0.5, -0.4, 1.0], [-0.5, 0.5, -0.4, 1.0], [0.5, 0.5, -0.4, 1.0], [0.5, -0.5, -0.4, 1.0]]], np.float32)
    v = np.concatenate([v, v * np.array([0.8, 0.8, 1.0, 1.0], np.float32)], 0)                  # two elements
    f = np.array([[[0, 1, 2], [0, 2, 3], [4, 5, 6], [4, 6, 7]]], np.int32)                      # shared by both
    c = np.ones((2, 8, 3), np.float32)
    bg = (np.random.RandomState(3).randn(2, H, W, 3) * 0.2).astype(np.float32)
    for name, tex, cam in (("oceanic_simple_proxy", bg, CAM_B), ("oceanic_opt_flow", bg, CAM_A), ("hill", terrain_texture(H, W)[None], HILL_CAM)):
        ref = golden.scene(name, tex, v, f, cam)
        out = getattr(dirt_b200, name + "_batch")(torch.from_numpy(tex).cuda(), torch.from_numpy(v).cuda(), torch.from_numpy(c).cuda(),
                                                  torch.from_numpy(f).cuda(), cam).cpu().numpy()
        assert out.shape == ref.shape == (2, H, W, 3)
        scale = max(1.0, float(np.abs(ref).max())) if name == "oceanic_opt_flow" else 1.0
        mean, beyond = _stats(out / scale, ref / scale, 1e-4)
        assert mean <= 1e-6 and beyond <= 0.004, (name, mean, beyond)
        assert not np.array_equal(out[0], out[1])


@pytest.mark.gpu
def test_cuda_scene_argument_errors():
    import torch
    import dirt_b200
    bg, v, c, f = quad(1, 16, 16, 3)
    t = [torch.from_numpy(a).cuda() for a in (bg, v, c, f)]
    with pytest.raises(dirt_b200.InvalidArgumentError):
        dirt_b200.oceanic_still_cloud_batch(*t, CAM_A[:8])
    with pytest.raises(dirt_b200.InvalidArgumentError):
        dirt_b200.oceanic_opt_flow_batch(*t, CAM_A[:15])
    with pytest.raises(dirt_b200.InvalidArgumentError):
        dirt_b200.hill_batch(*t, CAM_A[:11])
    with pytest.raises(dirt_b200.InvalidArgumentError):
        dirt_b200.hill_batch(torch.zeros(1, 16, 16, 2).cuda(), *t[1:], HILL_CAM)
