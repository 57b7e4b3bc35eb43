"""Fill-rule properties (SURVEY.md 8(c) "property tests"): vertices on a 1/16-pixel lattice, so that
edges pass EXACTLY through pixel centres and shared vertices sit on them -- the cases a random
continuous soup never produces.
  CPU (hypothesis, oracle): two triangles sharing an edge tile their union -- no pixel twice, none
  lost -- whichever way either is wound; a fan around a pixel centre covers that pixel exactly once.
  GPU: lattice soups (depth ties included) match the oracle bit for bit, gradients to the usual bar."""
import numpy as np
import pytest
from hypothesis import given, settings, assume, strategies as st

from oracle import golden

N = 24              # frame is N x N pixels
SUB = 16            # lattice: 1/16 pixel


def _clip(points):
    """Lattice points (units of 1/SUB pixel, y up) -> clip-space vertices [n,4]; exact in binary32."""
    p = np.asarray(points, np.float64) / SUB
    v = np.zeros((len(points), 4), np.float32)
    v[:, 0] = p[:, 0] * 2.0 / N - 1.0
    v[:, 1] = p[:, 1] * 2.0 / N - 1.0
    v[:, 3] = 1.0
    return v


def _coverage(v, faces):
    _, fid, _ = golden.forward(np.zeros((N, N, 1), np.float32), v, np.ones((len(v), 1), np.float32), np.asarray(faces, np.int32))
    return fid[0]


def _cross(o, a, b):
    return (a[0] - o[0]) * (b[1] - o[1]) - (a[1] - o[1]) * (b[0] - o[0])


coord = st.integers(0, N * SUB)
# half of the draws land on pixel centres exactly (odd multiples of SUB / 2)
centre = st.integers(0, N - 1).map(lambda i: i * SUB + SUB // 2)
point = st.tuples(st.one_of(coord, centre), st.one_of(coord, centre))


@settings(max_examples=150, deadline=None)
@given(point, point, point, point, st.booleans(), st.booleans())
def test_two_triangles_sharing_an_edge_tile_their_union(p0, p1, p2, p3, flip_a, flip_b):
    c1, c3 = _cross(p0, p2, p1), _cross(p0, p2, p3)
    assume(c1 != 0 and c3 != 0 and (c1 > 0) != (c3 > 0))          # p1 and p3 strictly on opposite sides of p0-p2
    v = _clip([p0, p1, p2, p3])
    ta = [0, 2, 1] if flip_a else [0, 1, 2]
    tb = [0, 3, 2] if flip_b else [0, 2, 3]
    a, b = _coverage(v, [ta]) >= 0, _coverage(v, [tb]) >= 0
    both = _coverage(v, [ta, tb])
    assert not (a & b).any(), "a pixel on the shared edge belongs to both triangles"
    assert np.array_equal(both >= 0, a | b)
    assert np.array_equal(both[a], np.zeros(a.sum(), np.int32)) and np.array_equal(both[b], np.ones(b.sum(), np.int32))
    # winding does not change coverage
    assert np.array_equal(a, _coverage(v, [ta[::-1]]) >= 0)


@settings(max_examples=60, deadline=None)
@given(st.integers(2, N - 3), st.integers(2, N - 3), st.integers(3, 9), st.integers(0, 1000))
def test_fan_around_a_pixel_centre_covers_it_exactly_once(ix, iy, n, seed):
    rng = np.random.RandomState(seed)
    hub = (ix * SUB + SUB // 2, iy * SUB + SUB // 2)                # exactly a pixel centre
    ang = np.sort(rng.uniform(0, 2 * np.pi, n))
    assume(np.all(np.diff(np.concatenate([ang, [ang[0] + 2 * np.pi]])) < np.pi - 0.2))    # star-shaped around the hub
    rim = [(int(round(hub[0] + 2.2 * SUB * np.cos(a))), int(round(hub[1] + 2.2 * SUB * np.sin(a)))) for a in ang]
    assume(all(_cross(hub, rim[k], rim[(k + 1) % n]) > 0 for k in range(n)))
    v = _clip([hub] + rim)
    faces = [[0, 1 + k, 1 + (k + 1) % n] for k in range(n)]
    count = np.zeros((N, N), int)
    for f in faces:
        count += _coverage(v, [f]) >= 0
    assert count.max() == 1, "a pixel is covered by two triangles of the fan"
    assert count[N - 1 - iy, ix] == 1, "the hub's pixel is lost"
    assert np.array_equal(count == 1, _coverage(v, faces) >= 0)


def lattice_soup(B, H, W, F, seed, sub=2, depth_levels=3):
    """Triangles whose vertices sit on a 1/sub-pixel lattice (sub = 2: pixel corners and centres),
    depths from a few discrete levels: edges through centres, coincident vertices, depth ties."""
    rng = np.random.RandomState(seed)
    V = 3 * F
    ctr = rng.randint(0, [W * sub, H * sub], size=(B, F, 1, 2))
    off = rng.randint(-6 * sub, 6 * sub + 1, size=(B, F, 3, 2))
    p = (ctr + off).reshape(B, V, 2).astype(np.float64) / sub
    xy = p * 2.0 / np.array([W, H], np.float64) - 1.0
    z = np.repeat(rng.randint(0, depth_levels, size=(B, F, 1)), 3, axis=2).reshape(B, V, 1) / float(depth_levels) - 0.5
    v = np.concatenate([xy, z, np.ones((B, V, 1))], 2).astype(np.float32)
    faces = np.broadcast_to(np.arange(V, dtype=np.int32).reshape(F, 3), (B, F, 3)).copy()
    return {"background": rng.uniform(0, 1, size=(B, H, W, 3)).astype(np.float32), "vertices": v,
            "vertex_colors": rng.uniform(0, 1, size=(B, V, 3)).astype(np.float32), "faces": faces}


@pytest.mark.gpu
@pytest.mark.parametrize("H,W,F,sub", [(64, 64, 400, 2), (48, 80, 1500, 1), (128, 128, 3000, 16), (32, 32, 40000, 2)])
def test_lattice_soup_matches_the_oracle_bit_for_bit(H, W, F, sub):
    import torch
    from dirt_b200 import rasterise_ops as ops
    sc = lattice_soup(3, H, W, F, seed=F + sub, sub=sub)
    t = {k: torch.from_numpy(v).cuda() for k, v in sc.items()}
    px, fid, gb = ops.forward_raw(t["background"], t["vertices"], t["vertex_colors"], t["faces"])
    opx, ofid, ogb = golden.forward(sc["background"], sc["vertices"], sc["vertex_colors"], sc["faces"], threads=8)
    assert (ofid >= 0).mean() > 0.3
    assert np.array_equal(fid.cpu().numpy(), ofid), "face ids differ at %d pixels" % (fid.cpu().numpy() != ofid).sum()
    np.testing.assert_allclose(px.cpu().numpy(), opx, rtol=1e-5, atol=1e-6)
    assert np.array_equal(gb.cpu().numpy()[..., 3].view(np.int32), ogb[..., 3].view(np.int32))
    g = np.random.RandomState(0).randn(*opx.shape).astype(np.float32)
    V = sc["vertices"].shape[1]
    gv, gc, gbg = ops.backward_raw(t["faces"], px, torch.from_numpy(g).cuda(), fid, gb, V)
    ogv, ogc, ogbg = golden.backward(sc["faces"], opx, g, ofid, ogb, V, threads=8)
    assert np.array_equal(gbg.cpu().numpy(), ogbg)
    for a, b in ((gv.cpu().numpy(), ogv), (gc.cpu().numpy(), ogc)):
        assert np.abs(a - b).max() <= 1e-4 * max(np.abs(b).max(), 1e-12)


@settings(max_examples=80, deadline=None)
@given(point, point, point, st.integers(-6, 6), st.integers(-6, 6))
def test_coverage_moves_with_the_triangle_by_whole_pixels(p0, p1, p2, kx, ky):
    """Shifting every corner by whole pixels shifts the covered set by the same pixels (where both
    stay inside the frame): the fill rule depends on sub-pixel position only, never on where in the
    frame a sample lies."""
    assume(_cross(p0, p1, p2) != 0)
    pts = np.array([p0, p1, p2])
    a = _coverage(_clip(pts), [[0, 1, 2]]) >= 0
    b = _coverage(_clip(pts + np.array([kx, ky]) * SUB), [[0, 1, 2]]) >= 0
    # rows run downwards: +ky pixels up is -ky rows
    ys, xs = np.nonzero(a)
    ys2, xs2 = ys - ky, xs + kx
    keep = (ys2 >= 0) & (ys2 < N) & (xs2 >= 0) & (xs2 < N)
    assert b[ys2[keep], xs2[keep]].all(), "a covered pixel was lost by the shift"
    ys, xs = np.nonzero(b)
    ys0, xs0 = ys + ky, xs - kx
    keep = (ys0 >= 0) & (ys0 < N) & (xs0 >= 0) & (xs0 < N)
    assert a[ys0[keep], xs0[keep]].all(), "the shift covered a pixel it should not"
