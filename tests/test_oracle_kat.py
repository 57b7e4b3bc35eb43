"""CPU: pins the oracle against the reference's own known-answer tests and checks the raster
contract's micro-cases."""
import numpy as np
import pytest

import scenes
from oracle import golden


def _fwd(sc, **kw):
    return golden.forward(sc["background"], sc["vertices"], sc["vertex_colors"], sc["faces"], **kw)


def test_kat1_readme_square():
    # README.md:46-70; upstream asserted "successful: all pixels agree" (README.md:104)
    sc = scenes.kat1_square()
    px, fid, _ = _fwd(sc)
    assert np.array_equal(px, sc["expected"])
    assert px.sum() == 256.0
    assert set(np.unique(fid)) == {-1, 0, 1}


def test_kat2_multi_gpu_square():
    # tests/multi_gpu_test.py:8-17: asymmetric in y => pins the row flip; clockwise faces => no culling
    sc = scenes.kat2_square()
    px, _, _ = _fwd(sc)
    assert np.array_equal(px, sc["expected"])


def test_both_windings_render():
    sc = scenes.kat1_square()
    a, fa, _ = _fwd(sc)
    sc["faces"] = sc["faces"][:, :, ::-1].copy()
    b, fb, _ = _fwd(sc)
    assert np.array_equal(a, b) and np.array_equal(fa, fb)


def test_shared_edge_covered_exactly_once():
    # two triangles sharing the diagonal of a pixel-aligned square whose diagonal passes through
    # pixel centres: every centre on the diagonal must belong to exactly one face
    W = H = 16
    v = np.array([[-1, -1, 0, 1], [1, -1, 0, 1], [1, 1, 0, 1], [-1, 1, 0, 1]], np.float32)
    faces = np.array([[0, 1, 2], [0, 2, 3]], np.int32)
    px, fid, _ = golden.forward(np.zeros((H, W, 1)), v, np.ones((4, 1)), faces)
    assert (fid >= 0).all()
    cnt = np.zeros((H, W), int)
    for f in range(2):
        _, fid1, _ = golden.forward(np.zeros((H, W, 1)), v, np.ones((4, 1)), faces[f:f + 1])
        cnt += (fid1[0] >= 0)
    assert (cnt == 1).all()


def test_depth_tie_lowest_face_wins_and_nearer_wins():
    W = H = 8
    quad = lambda z: [[-1, -1, z, 1], [1, -1, z, 1], [1, 1, z, 1], [-1, 1, z, 1]]
    v = np.array(quad(0.0) + quad(0.0) + quad(-0.5), np.float32)
    faces = np.array([[4, 5, 6], [0, 1, 2], [8, 9, 10]], np.int32)
    cols = np.repeat(np.array([[1.], [2.], [3.]], np.float32), 4, axis=0)
    px, fid, _ = golden.forward(np.zeros((H, W, 1)), v, cols, faces[:2])
    assert set(np.unique(fid)) <= {-1, 0}            # equal depth: face 0 (drawn first) wins
    px, fid, _ = golden.forward(np.zeros((H, W, 1)), v, cols, faces)
    assert set(np.unique(fid)) <= {-1, 2}            # z = -0.5 is nearer


def test_depth_range_clip_and_background_passthrough():
    W = H = 8
    tri = lambda z: [[-1, -1, z, 1], [3, -1, z, 1], [-1, 3, z, 1]]
    bg = np.full((H, W, 3), -np.inf, np.float32)     # samples/deferred.py:67 uses -inf backgrounds
    for z, visible in ((-1.5, False), (1.5, False), (1.0, False), (0.999, True), (-1.0, True)):
        px, fid, _ = golden.forward(bg, np.array(tri(z), np.float32), np.ones((3, 3)), [[0, 1, 2]])
        assert (fid >= 0).all() == visible, z
        if not visible:
            assert np.isneginf(px).all()


def test_near_plane_crossing_keeps_barycentrics_of_original_vertices():
    # a triangle with one vertex behind the camera; colour = clip-space w, which is linear over
    # the triangle, so every covered pixel must read back its own interpolated w (gbuf[...,2])
    W = H = 64
    v = np.array([[-0.5, -0.5, -0.8, 1.0], [0.5, -0.5, -0.8, 1.0], [0.0, 3.0, -2.0, -0.5]], np.float32)
    cols = v[:, 3:4].copy()
    px, fid, gb = golden.forward(np.zeros((H, W, 1)), v, cols, [[0, 1, 2]])
    m = fid[0] >= 0
    assert m.sum() > 100
    np.testing.assert_allclose(px[0, :, :, 0][m], gb[0, :, :, 2][m], rtol=2e-4, atol=1e-5)
    assert (gb[0, :, :, 2][m] > 0).all()


@pytest.mark.parametrize("winding", [(0, 1, 2), (0, 2, 1)])
@pytest.mark.parametrize("plane", ["near", "far"])
def test_depth_plane_clip_of_a_known_triangle_has_the_coverage_derived_by_hand(winding, plane):
    """A known answer that owes nothing to this project's code: a triangle with w = 1 whose third corner lies
    beyond the near (far) plane.  Clipping at z = -w (z = w) cuts the two edges that reach it at t = 3/4, i.e. at
    NDC (-1/8, -1/4) and (1/8, -1/4); what is left is the trapezoid between y = -1/4 and y = 1/2 whose sides run from
    (-1/2, 1/2) to (-1/8, -1/4) and from (1/2, 1/2) to (1/8, -1/4).  At 64 x 64 no pixel centre lies on any of its edges,
    so the covered set follows from the OpenGL rule "centre inside the polygon" alone (OpenGL 3.3 core, 3.6.1 / 2.20)."""
    W = H = 64
    z_out = -1.5 if plane == "near" else 1.5
    z_in = 0.5 if plane == "near" else -0.5
    v = np.array([[-0.5, 0.5, z_in, 1.0], [0.5, 0.5, z_in, 1.0], [0.0, -0.5, z_out, 1.0]], np.float32)
    px, fid, gb = golden.forward(np.zeros((H, W, 1)), v, np.ones((3, 1)), [list(winding)])
    expect = np.zeros((H, W), bool)
    for row in range(H):
        y = (H - 1 - row) + 0.5                       # window y of the row's pixel centres, y up
        if not (24.0 < y < 48.0):                     # -1/4 and 1/2 in NDC are 24 and 48 in window coordinates
            continue
        half = 16.0 - (48.0 - y) * 0.5                # half width: 16 px at the top edge, 4 px at the cut
        xs = np.arange(W) + 0.5
        expect[row] = np.abs(xs - 32.0) < half
    assert expect.sum() > 400
    assert np.array_equal(fid[0] >= 0, expect)
    # window depth along the cut is the clip plane's: 0 (near) or 1 (far); the row of centres just above the cut is
    # 1/48 of the way up the trapezoid's 24-px height in z, i.e. within 0.05 of it
    cut_row = H - 1 - 24
    d = gb[0, cut_row][expect[cut_row], 3]
    assert np.all(np.abs(d - (0.0 if plane == "near" else 1.0)) < 0.05)


def test_batch_local_indices_and_broadcast():
    sc = scenes.cube(B=3, H=48, W=48)
    px, fid, _ = _fwd(sc)
    for b in range(3):
        p1, f1, _ = golden.forward(sc["background"][b], sc["vertices"][b], sc["vertex_colors"][b], sc["faces"][b])
        assert np.array_equal(p1[0], px[b]) and np.array_equal(f1[0], fid[b])
    # faces given once, broadcast over the batch
    p2, f2, _ = golden.forward(sc["background"], sc["vertices"], sc["vertex_colors"], sc["faces"][:1])
    assert np.array_equal(p2, px) and np.array_equal(f2, fid)


def test_constant_colour_is_exact_and_vertex_rotation_invariant_coverage():
    sc = scenes.random_triangles(1, 64, 64, 3, 200, seed=3)
    sc["vertex_colors"][:] = 0.7
    px, fid, _ = _fwd(sc)
    assert np.all(px[fid >= 0] == np.float32(0.7))
    sc2 = dict(sc)
    sc2["faces"] = np.roll(sc["faces"], 1, axis=2).copy()
    _, fid2, _ = _fwd(sc2)
    assert np.array_equal(fid, fid2)


def test_threads_do_not_change_results():
    sc = scenes.random_triangles(4, 40, 56, 3, 150, seed=9)
    a = _fwd(sc, threads=1)
    b = _fwd(sc, threads=4)
    for x, y in zip(a, b):
        assert np.array_equal(x, y)


def test_backward_colour_and_background_parts_are_exact():
    sc = scenes.cube(B=2, H=40, W=40)
    px, fid, gb = _fwd(sc)
    rng = np.random.RandomState(0)
    g = rng.randn(*px.shape).astype(np.float32)
    V = sc["vertices"].shape[1]
    gv, gc, gbg = golden.backward(sc["faces"], px, g, fid, gb, V)
    assert np.array_equal(gbg, np.where((fid < 0)[..., None], g, 0))
    # colours: pixels are linear in vertex_colors => sum(g * pixels(colours=e)) reproduces gc
    for b in range(2):
        for vi in (0, 7, 20):
            for c in range(3):
                e = np.zeros_like(sc["vertex_colors"])
                e[b, vi, c] = 1
                p1, _, _ = golden.forward(np.zeros_like(sc["background"]), sc["vertices"], e, sc["faces"])
                np.testing.assert_allclose((p1 * g).sum(), gc[b, vi, c], rtol=1e-4, atol=1e-5)


def test_backward_position_gradient_is_boundary_only_for_flat_square():
    # README.md:41: gradients w.r.t. the square's centre "should be non-zero at the boundary"
    sc = scenes.kat1_square()
    px, fid, gb = _fwd(sc)
    H, W = px.shape[1:3]
    for (r, x, expect_nonzero) in ((64, 32, False), (64, 24, True), (64, 23, True), (10, 100, False)):
        g = np.zeros_like(px)
        g[0, r, x, 0] = 1
        gv, _, _ = golden.backward(sc["faces"], px, g, fid, gb, 4)
        assert (np.abs(gv).sum() > 0) == expect_nonzero, (r, x)
    # moving the square right (+x) brightens pixels just right of its right edge
    g = np.zeros_like(px)
    g[0, 64, 40, 0] = 1           # first background pixel right of the square
    gv, _, _ = golden.backward(sc["faces"], px, g, fid, gb, 4)
    assert gv[0, :, 0].sum() > 0


def test_backward_finite_difference_translation():
    # d/dt of a smooth loss under a rigid x-translation of a shaded cube, against central
    # differences of 4x supersampled renders (band: same sign, within 35 %)
    H = W = 48
    ss = 4
    sc = scenes.cube(B=1, H=H, W=W)
    sc["background"][:] = 0.1
    big = scenes.cube(B=1, H=H * ss, W=W * ss)

    def render_ss(dx):
        v = big["vertices"].copy()
        v[..., 0] += dx * v[..., 3]
        p, _, _ = golden.forward(np.full((1, H * ss, W * ss, 3), 0.1, np.float32), v, sc["vertex_colors"], sc["faces"])
        return p.reshape(1, H, ss, W, ss, 3).mean(axis=(2, 4))

    yy, xx = np.mgrid[0:H, 0:W]
    wgt = (np.sin(xx / 7.0) + np.cos(yy / 5.0))[None, :, :, None].astype(np.float32) * np.ones((1, 1, 1, 3), np.float32)
    eps = 0.5 * 2.0 / W           # half a pixel in NDC
    fd = ((render_ss(eps) - render_ss(-eps)) * wgt).sum() / (2 * eps)
    px, fid, gb = _fwd(sc)
    gv, _, _ = golden.backward(sc["faces"], px, wgt, fid, gb, 36)
    # x_clip = x_ndc * w  =>  d/dt sum over vertices of gx * w
    an = (gv[0, :, 0] * sc["vertices"][0, :, 3]).sum()
    assert np.sign(fd) == np.sign(an)
    assert abs(an - fd) <= 0.35 * abs(fd), (an, fd)


def test_attributes_match_the_closed_form_of_perspective_correct_interpolation():
    """Pins the VALUES (not just coverage) to the OpenGL definition the reference's `smooth`
    varyings follow (csrc/shaders.cpp:16-31): with screen-space barycentrics l_i of the pixel
    centre, attribute = sum(l_i a_i / w_i) / sum(l_i / w_i), clip w = 1 / sum(l_i / w_i), window
    depth = (sum(l_i z_i / w_i) + 1) / 2 -- evaluated here in float64 from the raw clip coordinates,
    independently of the oracle's arithmetic."""
    H, W, C = 40, 56, 3
    v = np.array([[[-0.9, -0.8, 0.2, 1.0], [0.7, -0.5, 0.9, 2.5], [0.1, 2.4, -0.6, 3.0]]], np.float32)   # different w per corner
    col = np.array([[[1.0, 0.0, 0.25], [0.0, 1.0, 0.5], [0.3, 0.2, 1.0]]], np.float32)
    f = np.array([[[0, 1, 2]]], np.int32)
    bg = np.full((1, H, W, C), -7.0, np.float32)
    px, fid, gb = golden.forward(bg, v, col, f)
    cov = fid[0] >= 0
    assert 200 < cov.sum() < H * W - 200
    vd = v[0].astype(np.float64)
    ndc = vd[:, :3] / vd[:, 3:4]
    sx = (ndc[:, 0] + 1) * W / 2
    sy = (ndc[:, 1] + 1) * H / 2
    ys, xs = np.nonzero(cov)
    cx, cy = xs + 0.5, (H - 1 - ys) + 0.5                               # window coordinates, y up
    def edge(ax, ay, bx, by, qx, qy):
        return (bx - ax) * (qy - ay) - (by - ay) * (qx - ax)
    area = edge(sx[0], sy[0], sx[1], sy[1], sx[2], sy[2])
    l0 = edge(sx[1], sy[1], sx[2], sy[2], cx, cy) / area
    l1 = edge(sx[2], sy[2], sx[0], sy[0], cx, cy) / area
    l2 = 1 - l0 - l1
    # points within a hair of an edge may fall on the other side after snapping to 1/256 px
    inside = (np.minimum(np.minimum(l0, l1), l2) > 2e-3)
    assert inside.mean() > 0.9
    iw = l0 / vd[0, 3] + l1 / vd[1, 3] + l2 / vd[2, 3]
    b0, b1, b2 = l0 / vd[0, 3] / iw, l1 / vd[1, 3] / iw, l2 / vd[2, 3] / iw
    want = b0[:, None] * col[0, 0] + b1[:, None] * col[0, 1] + b2[:, None] * col[0, 2]
    got = px[0][cov]
    # vertices are snapped to 1/256 px before interpolation (docs/RASTER_SPEC.md), which moves a weight by ~1e-5 here
    assert np.abs(got - want)[inside].max() < 5e-5
    g = gb[0][cov]
    assert np.abs(g[:, 0] - b0)[inside].max() < 5e-5 and np.abs(g[:, 1] - b1)[inside].max() < 5e-5
    assert np.abs(g[:, 2] * iw - 1)[inside].max() < 5e-5                 # clip-space w
    zw = (l0 * ndc[0, 2] + l1 * ndc[1, 2] + l2 * ndc[2, 2] + 1) / 2
    assert np.abs(g[:, 3] - zw)[inside].max() < 5e-5                     # window depth is linear on the screen
    assert (px[0][~cov] == -7.0).all()
