"""The fork's own `Rasterise` behaviour (full-screen quad + oceanic_horizon shader driven by
camera_pos: tests/optimize_horizon.py:251-268, tests/square_test.py:39-56) -- CUDA scene stage
behind the new rasteriser vs the C transcription of the GLSL (oracle/scene_oceanic.c).
Tolerance: the shader is binary32 with sin / pow, and its ray march branches on signs, so a
1-ulp difference between libm and CUDA's sinf can move a rare pixel a lot: the bar is 1e-3
absolute on >= 99.5 % of the pixels and an identical sky / sea mask on >= 99.9 %."""
import numpy as np
import pytest

from oracle import golden


def quad_scene(B, H, W, C, seed=0, noise=0.0):
    # full-screen quad, clockwise faces, as every caller in the fork draws it
    v = np.array([[-1, -1, 0, 1], [-1, 1, 0, 1], [1, 1, 0, 1], [1, -1, 0, 1]], np.float32)
    faces = np.array([[0, 1, 2], [0, 2, 3]], np.int32)
    rng = np.random.RandomState(seed)
    bg = (rng.randn(B, H, W, C) * noise).astype(np.float32)
    return bg, np.broadcast_to(v, (B, 4, 4)).copy(), np.ones((B, 4, C), np.float32), np.broadcast_to(faces, (B, 2, 3)).copy()


CAMS = [
    # cam_x, cam_y, cam_z, ang1, ang2, ang3, time, light_z   (tests/optimize_horizon.py:251-268 style)
    [0.0, 100.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.9],
    [30.0, 140.0, -20.0, 0.12, 0.05, -0.08, 3.5, 0.4],
    [-5.0, 85.0, 10.0, -0.25, 0.3, 0.1, 10.0, -0.3],
]


def test_oracle_scene_basic_properties():
    bg, v, c, f = quad_scene(1, 48, 64, 3)
    px = golden.oceanic_horizon(bg, v, f, CAMS[0])
    assert px.shape == (1, 48, 64, 3) and np.isfinite(px).all()
    mask = px[0, :, :, 0]
    assert set(np.unique(mask)) == {0.0, 1.0}                 # sky / sea
    assert mask[0].all() and not mask[-1].any()                # level camera: sky on top, sea below
    assert (px[..., 2] == 0).all() and (px[..., 1] >= 0).all()
    # a quad covering only half the frame leaves the background elsewhere
    v2 = v.copy(); v2[..., 0] = v2[..., 0] * 0.5 - 0.5
    bg2 = np.full_like(bg, 7.0)
    px2 = golden.oceanic_horizon(bg2, v2, f, CAMS[0])
    assert (px2[0, :, 40:, :] == 7.0).all() and (px2[0, :, :30, 2] == 0).all()


@pytest.mark.gpu
@pytest.mark.parametrize("cam", CAMS)
@pytest.mark.parametrize("C,noise", [(3, 0.0), (3, 0.5), (1, 0.0)])
def test_cuda_scene_matches_oracle(cam, C, noise):
    import torch
    import dirt_b200
    B, H, W = 2, 90, 160
    bg, v, c, f = quad_scene(B, H, W, C, seed=1, noise=noise)
    ref = golden.oceanic_horizon(bg, v, f, cam)
    out = dirt_b200.oceanic_horizon_batch(torch.from_numpy(bg).cuda(), torch.from_numpy(v).cuda(), torch.from_numpy(c).cuda(),
                                          torch.from_numpy(f).cuda(), cam).cpu().numpy()
    assert out.shape == ref.shape
    same_mask = (out[..., 0] == ref[..., 0]).mean()
    close = (np.abs(out - ref) <= 1e-3).all(axis=-1).mean()
    assert same_mask >= 0.999 and close >= 0.995, (same_mask, close)
    # unbatched form, the fork's call shape
    one = dirt_b200.oceanic_horizon(torch.from_numpy(bg[0]).cuda(), v[0], c[0], f[0], np.array(cam, np.float32))
    assert np.array_equal(one.cpu().numpy(), out[0])


@pytest.mark.gpu
def test_scene_op_with_a_pinned_host_background():
    """the rasterise path reads a pinned host background in place; the scene ops must take their
    device from the vertices, not from that background."""
    import torch
    import dirt_b200
    B, H, W, C = 2, 60, 96, 3
    bg, v, c, f = quad_scene(B, H, W, C, seed=2, noise=0.3)
    dev = [torch.from_numpy(a).cuda() for a in (bg, v, c, f)]
    ref = dirt_b200.oceanic_horizon_batch(dev[0], dev[1], dev[2], dev[3], CAMS[1])
    out = dirt_b200.oceanic_horizon_batch(torch.from_numpy(bg).pin_memory(), dev[1], dev[2], dev[3], CAMS[1])
    assert out.is_cuda and torch.equal(out, ref)
