"""Ground truth for the position gradient that does not come from this project's own code: cases whose
answer is an area integral known in closed form (docs/GRADIENT_SPEC.md sections 2-4).

A constant-colour axis-aligned square (colour cs) over a constant surround (colour cb), loss = sum of the
pixels of a window whose boundary crosses the square: translating the square by t in clip x moves its
vertical edges by t * W / (2 w) pixels, and only the edge inside the window changes the loss, so
    d loss / d t = (cs - cb) * edge_height_px * W / (2 w)        exactly.
The filter-based gradient reproduces it up to the two corner rows (a relative error of ~0.5 / height);
the bar is 2 %.  The same under perspective (w != 1) pins the W / (2 w) factor, a uniform change of w pins
the w column, and an occluder in front of a second surface pins the ownership rule.  Each case runs against
the oracle on the CPU and against the CUDA path through the C ABI on the GPU."""
import numpy as np
import pytest

from oracle import golden

H = W = 256
TOL = 0.02


def _square(x0, x1, y0, y1, z, w, first=0):
    """Two triangles covering NDC [x0,x1] x [y0,y1] at clip depth z*w, clip w."""
    ndc = np.array([[x0, y0], [x0, y1], [x1, y1], [x1, y0]], np.float64)
    v = np.concatenate([ndc * w, np.full((4, 1), z * w), np.full((4, 1), w)], 1)
    f = np.array([[0, 1, 2], [0, 2, 3]], np.int32) + first
    return v, f


def _oracle(bg, v, col, f, g):
    px, fid, gb = golden.forward(bg, v, col, f)
    gv, gc, gbg = golden.backward(f, px, g, fid, gb, v.shape[1])
    return px, gv


def _gpu(bg, v, col, f, g):
    import torch
    from dirt_b200 import rasterise_ops as ops
    t = [torch.from_numpy(np.ascontiguousarray(a)).cuda() for a in (bg, v, col, f, g)]
    px, fid, gb = ops.forward_raw(t[0], t[1], t[2], t[3])
    gv, gc, gbg = ops.backward_raw(t[3], px, t[4], fid, gb, v.shape[1])
    return px.cpu().numpy(), gv.cpu().numpy()


IMPLS = [pytest.param(_oracle, id="oracle"), pytest.param(_gpu, id="cuda", marks=pytest.mark.gpu)]


def _scene(squares, cb=0.25, C=3):
    """squares: list of (x0, x1, y0, y1, z, w, colour)."""
    vs, fs, cs = [], [], []
    for (x0, x1, y0, y1, z, w, colour) in squares:
        v, f = _square(x0, x1, y0, y1, z, w, first=4 * len(vs))
        vs.append(v); fs.append(f); cs.append(np.full((4, C), colour))
    bg = np.full((1, H, W, C), cb, np.float32)
    return (bg, np.concatenate(vs)[None].astype(np.float32), np.concatenate(cs)[None].astype(np.float32),
            np.concatenate(fs)[None].astype(np.int32))


def _window_right_half(C=3):
    g = np.zeros((1, H, W, C), np.float32)
    g[:, :, W // 2:, :] = 1.0
    return g


@pytest.mark.parametrize("impl", IMPLS)
@pytest.mark.parametrize("w", [1.0, 2.5])
def test_translation_of_a_square_over_the_background(impl, w):
    cs, cb = 0.9, 0.25
    # edges off the pixel grid; the right edge (x = 0.43) lies inside the window, the left one outside
    bg, v, col, f = _scene([(-0.52, 0.43, -0.37, 0.41, 0.2, w, cs)], cb)
    g = _window_right_half()
    px, gv = impl(bg, v, col, f, g)
    rows = (px[0, :, W - 40, 0] == np.float32(cb)).sum(), (px[0, :, W // 2 + 5, 0] != np.float32(cb)).sum()
    height_px = rows[1]
    assert abs(height_px - (0.41 + 0.37) * H / 2) <= 1
    expect = 3 * (cs - cb) * height_px * W / (2 * w)          # three channels
    got = gv[0, :, 0].sum()
    assert abs(got - expect) <= TOL * abs(expect), (got, expect)
    # nothing moves the loss in y (both horizontal edges cross the window symmetrically: up minus down)
    # and z never receives a gradient
    assert np.all(gv[0, :, 2] == 0)


@pytest.mark.parametrize("impl", IMPLS)
def test_uniform_change_of_w_pins_the_w_column(impl):
    cs, cb, w = 0.9, 0.25, 2.0
    x0, x1, y0, y1 = -0.52, 0.43, -0.37, 0.41
    bg, v, col, f = _scene([(x0, x1, y0, y1, 0.2, w, cs)], cb)
    g = _window_right_half()
    px, gv = impl(bg, v, col, f, g)
    # raising every vertex's w by dw at fixed clip x, y scales the NDC square by (1 - dw / w): inside the window
    # the right edge moves left by x1 * W / (2 w) px per unit w, the top edge down by y1 * H / (2 w), the bottom
    # edge up by |y0| * H / (2 w), along the part of each edge that lies inside the window
    edge_h = (y1 - y0) * H / 2
    width_in_window = x1 * W / 2
    expect = -3 * (cs - cb) * (edge_h * x1 * W / (2 * w) + width_in_window * (y1 - y0) * H / (2 * w))
    got = gv[0, :, 3].sum()
    # the spec evaluates x_ndc, y_ndc at the centre of the owning pixel, half a pixel inside the edge (~1 % of the
    # ~50 px lever arms here), on top of the corner rows: the bar for this column is 4 %
    assert abs(got - expect) <= 2 * TOL * abs(expect), (got, expect)


@pytest.mark.parametrize("impl", IMPLS)
def test_occluder_edge_over_a_second_surface_goes_to_the_occluder(impl):
    c_near, c_far, w_near, w_far = 0.8, 0.3, 1.5, 3.0
    # the far square fills the frame and beyond; the near one ends inside the window
    bg, v, col, f = _scene([(-0.52, 0.43, -0.37, 0.41, -0.1, w_near, c_near), (-1.5, 1.5, -1.5, 1.5, 0.6, w_far, c_far)], 0.0)
    g = _window_right_half()
    px, gv = impl(bg, v, col, f, g)
    height_px = (px[0, :, W // 2 + 5, 0] == np.float32(c_near)).sum()
    expect = 3 * (c_near - c_far) * height_px * W / (2 * w_near)
    near, far = gv[0, :4, 0].sum(), gv[0, 4:, 0].sum()
    assert abs(near - expect) <= TOL * abs(expect), (near, expect)
    # the surface behind does not move when the edge in front of it does
    assert abs(far) <= TOL * abs(expect), (far, expect)
