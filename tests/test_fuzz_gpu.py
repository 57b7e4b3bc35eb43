"""Seeded fuzz of the C-ABI forward + backward against the oracle over the argument space the
reference op accepts: ragged frames, any channel count up to 16, zero to a few thousand faces of any
size, depth ties, broadcast (batch-1) background / mesh / colours in every combination, faces that
repeat vertices or point outside the vertex array."""
import numpy as np
import pytest

import scenes
from oracle import golden

pytestmark = pytest.mark.gpu


def _case(seed):
    rng = np.random.RandomState(1000 + seed)
    B = int(rng.randint(1, 5))
    H, W = int(rng.randint(1, 97)), int(rng.randint(1, 97))
    C = int(rng.choice([1, 1, 2, 3, 3, 3, 4, 5, 9, 16]))
    F = int(rng.choice([0, 1, 2, 7, 50, 300, 2500]))
    size = float(rng.choice([0.02, 0.1, 0.5, 1.5, 4.0]))
    if F == 0:
        sc = scenes.random_triangles(B, H, W, C, 1, seed=seed, size=size)
        sc["faces"] = sc["faces"][:, :0]
    else:
        sc = scenes.random_triangles(B, H, W, C, F, seed=seed, size=size, depth_ties=bool(rng.randint(2)),
                                     w_range=(0.5, 2.0) if rng.randint(2) else (-0.3, 2.0))
    sc = {k: np.ascontiguousarray(v).copy() for k, v in sc.items()}
    if F > 2:
        f = sc["faces"]
        V = sc["vertices"].shape[1]
        k = max(1, F // 10)
        f[:, rng.randint(0, F, k), rng.randint(0, 3, k)] = rng.randint(0, V, k)          # shared / repeated corners
        f[:, rng.randint(0, F), 0] = V + 3                                              # out of range: skipped
        f[:, rng.randint(0, F), 2] = -1
    bcast = {k: bool(rng.randint(3) == 0) for k in ("background", "vertices", "vertex_colors", "faces")}
    if B > 1:
        for k, on in bcast.items():
            if on:
                sc[k] = sc[k][:1].copy()
    return sc, max(v.shape[0] for v in sc.values())      # every input broadcast: the batch really is 1


@pytest.mark.parametrize("seed", range(160))
def test_fuzz_forward_backward(seed):
    import torch
    from dirt_b200 import rasterise_ops as ops
    sc, B = _case(seed)
    t = {k: torch.from_numpy(v).cuda() for k, v in sc.items()}
    px, fid, gb = ops.forward_raw(t["background"], t["vertices"], t["vertex_colors"], t["faces"])
    opx, ofid, ogb = golden.forward(sc["background"], sc["vertices"], sc["vertex_colors"], sc["faces"], threads=4)
    assert px.shape[0] == B
    assert np.array_equal(fid.cpu().numpy(), ofid), "face ids differ at %d pixels" % (fid.cpu().numpy() != ofid).sum()
    np.testing.assert_allclose(px.cpu().numpy(), opx, rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(gb.cpu().numpy(), ogb, rtol=1e-5, atol=1e-6)
    g = np.random.RandomState(seed).randn(*opx.shape).astype(np.float32)
    V = sc["vertices"].shape[1]
    sv, scol = sc["vertices"].shape[0] == 1 and B > 1, sc["vertex_colors"].shape[0] == 1 and B > 1
    gv, gc, gbg = ops.backward_raw(t["faces"], px, torch.from_numpy(g).cuda(), fid, gb, V, shared_vertices=sv, shared_colors=scol)
    ogv, ogc, ogbg = golden.backward(sc["faces"], opx, g, ofid, ogb, V, shared_vertices=sv, shared_colors=scol, threads=4)
    assert np.array_equal(gbg.cpu().numpy(), ogbg)
    for name, a, b in (("grad_vertices", gv.cpu().numpy(), ogv), ("grad_vertex_colors", gc.cpu().numpy(), ogc)):
        assert a.shape == b.shape, name
        assert np.abs(a - b).max() <= 1e-4 * max(np.abs(b).max(), 1e-12), name


def _ordered_case(seed):
    """Meshes past the threshold of the depth-ordered passes (>= 4096 faces per element) with real
    overdraw, so that the nearest-first buckets, the depth pyramid and its culling all run."""
    rng = np.random.RandomState(5000 + seed)
    B = int(rng.randint(1, 4))
    H, W = int(rng.randint(40, 200)), int(rng.randint(40, 200))
    C = int(rng.choice([1, 3, 3, 4]))
    F = int(rng.choice([4096, 5000, 9000, 16384]))
    size = float(rng.choice([0.15, 0.3, 0.6, 1.2]))
    sc = scenes.random_triangles(B, H, W, C, F, seed=seed, size=size, depth_ties=bool(rng.randint(2)),
                                 w_range=(0.5, 2.0) if rng.randint(3) else (-0.2, 2.0))
    sc = {k: np.ascontiguousarray(v).copy() for k, v in sc.items()}
    if B > 1:
        for k in ("background", "vertices", "vertex_colors", "faces"):
            if rng.randint(4) == 0:
                sc[k] = sc[k][:1].copy()
    return sc, max(v.shape[0] for v in sc.values())


@pytest.mark.parametrize("seed", range(32))
def test_fuzz_depth_ordered_path(seed):
    import torch
    from dirt_b200 import rasterise_ops as ops
    sc, B = _ordered_case(seed)
    t = {k: torch.from_numpy(v).cuda() for k, v in sc.items()}
    px, fid, gb = ops.forward_raw(t["background"], t["vertices"], t["vertex_colors"], t["faces"])
    opx, ofid, ogb = golden.forward(sc["background"], sc["vertices"], sc["vertex_colors"], sc["faces"], threads=8)
    assert np.array_equal(fid.cpu().numpy(), ofid), "face ids differ at %d pixels" % (fid.cpu().numpy() != ofid).sum()
    np.testing.assert_allclose(px.cpu().numpy(), opx, rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(gb.cpu().numpy(), ogb, rtol=1e-5, atol=1e-6)
    g = np.random.RandomState(seed).randn(*opx.shape).astype(np.float32)
    V = sc["vertices"].shape[1]
    sv, scol = sc["vertices"].shape[0] == 1 and B > 1, sc["vertex_colors"].shape[0] == 1 and B > 1
    gv, gc, gbg = ops.backward_raw(t["faces"], px, torch.from_numpy(g).cuda(), fid, gb, V, shared_vertices=sv, shared_colors=scol)
    ogv, ogc, ogbg = golden.backward(sc["faces"], opx, g, ofid, ogb, V, shared_vertices=sv, shared_colors=scol, threads=8)
    assert np.array_equal(gbg.cpu().numpy(), ogbg)
    for name, a, b in (("grad_vertices", gv.cpu().numpy(), ogv), ("grad_vertex_colors", gc.cpu().numpy(), ogc)):
        assert np.abs(a - b).max() <= 1e-4 * max(np.abs(b).max(), 1e-12), name


@pytest.mark.parametrize("seed", range(24))
def test_fuzz_dilate(seed):
    """dirt_dilate_silhouette (+grad) against oracle/dilate.py over ragged shapes and hole densities."""
    import torch
    import dirt_b200
    from oracle import dilate as odil
    rng = np.random.RandomState(7000 + seed)
    shape = (int(rng.randint(1, 4)), int(rng.randint(1, 70)), int(rng.randint(1, 70)), int(rng.choice([1, 2, 3, 4, 6, 9, 16])))
    x = rng.randn(*shape).astype(np.float32)
    x[rng.uniform(size=shape[:3]) < rng.choice([0.0, 0.1, 0.5, 0.9, 1.0])] = -np.inf
    xt = torch.from_numpy(x).cuda().requires_grad_(True)
    out = dirt_b200.dilate_silhouette(xt)
    assert np.array_equal(out.detach().cpu().numpy(), odil.dilate(x))
    g = rng.randn(*shape).astype(np.float32)
    out.backward(torch.from_numpy(g).cuda())
    np.testing.assert_allclose(xt.grad.cpu().numpy(), odil.dilate_grad(x, g), rtol=1e-6, atol=1e-6)


@pytest.mark.parametrize("seed", range(16))
def test_fuzz_shared_mesh_backward(seed):
    """dirt_rasterise_backward_shared against the oracle's per-frame gradients folded on the host."""
    import torch
    from dirt_b200 import rasterise_ops as ops
    rng = np.random.RandomState(9000 + seed)
    B, H, W = int(rng.randint(1, 5)), int(rng.randint(8, 90)), int(rng.randint(8, 90))
    C = int(rng.choice([1, 3, 3, 4, 7]))
    F = int(rng.choice([1, 12, 200, 5000]))
    one = scenes.random_triangles(1, H, W, C, F, seed=seed, size=float(rng.choice([0.05, 0.4, 1.5])))
    world = one["vertices"][0].astype(np.float64)
    V = world.shape[0]
    M = np.stack([np.eye(4) + 0.05 * rng.randn(4, 4) for _ in range(B)])
    sc = {"background": rng.uniform(0, 1, (B, H, W, C)).astype(np.float32),
          "vertices": np.einsum("vi,bij->bvj", world, M).astype(np.float32),
          "vertex_colors": one["vertex_colors"].copy(), "faces": np.ascontiguousarray(one["faces"]).copy()}
    t = {k: torch.from_numpy(v).cuda() for k, v in sc.items()}
    px, fid, gb = ops.forward_raw(t["background"], t["vertices"], t["vertex_colors"], t["faces"])
    opx, ofid, ogb = golden.forward(sc["background"], sc["vertices"], sc["vertex_colors"], sc["faces"], threads=4)
    assert np.array_equal(fid.cpu().numpy(), ofid)
    g = rng.randn(B, H, W, C).astype(np.float32)
    packed, gbg = ops.backward_shared_raw(t["faces"], px, torch.from_numpy(g).cuda(), fid, gb, V,
                                          torch.from_numpy(M.astype(np.float32)).cuda())
    ogv, ogc, ogbg = golden.backward(sc["faces"], opx, g, ofid, ogb, V, shared_colors=B > 1, threads=4)
    want_v = np.einsum("bvj,bij->vi", ogv.astype(np.float64), M.astype(np.float32).astype(np.float64))
    want_c = ogc.astype(np.float64).sum(0)
    got = packed.cpu().numpy()
    assert np.array_equal(gbg.cpu().numpy(), ogbg)
    assert np.abs(got[:, :4] - want_v).max() <= 1e-4 * max(np.abs(want_v).max(), 1e-12)
    assert np.abs(got[:, 4:4 + C] - want_c.reshape(V, C)).max() <= 1e-4 * max(np.abs(want_c).max(), 1e-12)
    assert not got[:, 4 + C:].any()


def _extreme_case(seed):
    """Geometry at the edges of the clip volume and of the number range: corners exactly on the near
    / far plane, w a hair above zero (window positions hit the guard-band clamp), huge coordinates,
    degenerate faces (repeated corner, zero area), frames that are one pixel wide or very wide."""
    rng = np.random.RandomState(11000 + seed)
    B = int(rng.randint(1, 3))
    H, W = [(1, 300), (300, 1), (5, 1031), (64, 64), (97, 33), (17, 200)][seed % 6]
    C = int(rng.choice([1, 3]))
    F = int(rng.choice([6, 60, 600]))
    sc = scenes.random_triangles(B, H, W, C, F, seed=seed, size=float(rng.choice([0.3, 1.0, 3.0])), w_range=(-0.1, 2.0))
    sc = {k: np.ascontiguousarray(v).copy() for k, v in sc.items()}
    v = sc["vertices"]
    V = v.shape[1]
    pick = lambda frac: rng.uniform(size=(B, V)) < frac
    m = pick(0.08); v[m, 2] = v[m, 3]                                   # exactly on the far plane
    m = pick(0.08); v[m, 2] = -v[m, 3]                                  # exactly on the near plane
    m = pick(0.05); v[m, 3] = rng.choice([1e-7, 1e-5, 1e-3], size=m.sum()).astype(np.float32)   # w -> 0+
    m = pick(0.03); v[m, :2] *= 1e6                                     # far outside the guard band
    m = pick(0.02); v[m, 3] = 0.0
    f = sc["faces"]
    k = max(1, F // 12)
    idx = rng.randint(0, F, k)
    f[:, idx, 1] = f[:, idx, 0]                                         # repeated corner
    idx = rng.randint(0, F, k)
    v[:, f[0, idx, 2]] = v[:, f[0, idx, 1]]                             # coincident corners: zero area
    return sc


@pytest.mark.parametrize("seed", range(48))
def test_fuzz_extreme_geometry(seed):
    import torch
    from dirt_b200 import rasterise_ops as ops
    sc = _extreme_case(seed)
    t = {k: torch.from_numpy(v).cuda() for k, v in sc.items()}
    px, fid, gb = ops.forward_raw(t["background"], t["vertices"], t["vertex_colors"], t["faces"])
    opx, ofid, ogb = golden.forward(sc["background"], sc["vertices"], sc["vertex_colors"], sc["faces"], threads=4)
    assert np.array_equal(fid.cpu().numpy(), ofid), "face ids differ at %d pixels" % (fid.cpu().numpy() != ofid).sum()
    fin = np.isfinite(opx)
    assert np.array_equal(np.isfinite(px.cpu().numpy()), fin)
    np.testing.assert_allclose(px.cpu().numpy()[fin], opx[fin], rtol=1e-5, atol=1e-6)
    fin = np.isfinite(ogb)
    assert np.array_equal(np.isfinite(gb.cpu().numpy()), fin)
    np.testing.assert_allclose(gb.cpu().numpy()[fin], ogb[fin], rtol=1e-5, atol=1e-6)
    g = np.random.RandomState(seed).randn(*opx.shape).astype(np.float32)
    V = sc["vertices"].shape[1]
    gv, gc, gbg = ops.backward_raw(t["faces"], px, torch.from_numpy(g).cuda(), fid, gb, V)
    ogv, ogc, ogbg = golden.backward(sc["faces"], opx, g, ofid, ogb, V, threads=4)
    assert np.array_equal(gbg.cpu().numpy(), ogbg)
    for name, a, b in (("grad_vertices", gv.cpu().numpy(), ogv), ("grad_vertex_colors", gc.cpu().numpy(), ogc)):
        fin = np.isfinite(b)
        assert np.array_equal(np.isfinite(a), fin), name
        assert np.abs(a[fin] - b[fin]).max() <= 1e-4 * max(np.abs(b[fin]).max(), 1e-12), name


@pytest.mark.parametrize("seed", range(4))
@pytest.mark.parametrize("name", ["oceanic", "oceanic_still_cloud", "oceanic_no_cloud", "oceanic_simple_proxy", "oceanic_opt_flow", "hill"])
def test_fuzz_scene_cameras(name, seed):
    """The scene stages under random cameras, frame sizes and jitter, against the C/C++ oracle with the
    bars of tests/test_scene_shaders.py (mean <= 1e-6, <= 0.2 % of the pixels beyond 1e-4)."""
    import torch
    import dirt_b200
    import test_scene_shaders as ts
    rng = np.random.RandomState(13000 + 17 * seed + len(name))
    B, H, W = int(rng.randint(1, 3)), int(rng.randint(8, 60)), int(rng.randint(8, 90))
    if name == "hill":
        C, Cb = int(rng.choice([1, 3])), int(rng.choice([1, 3, 4]))
        _, v, c, f = ts.quad(B, H, W, C)
        tex = np.stack([ts.terrain_texture(H, W) for _ in range(B)])[..., :Cb]
        tex = np.ascontiguousarray(tex + 0.02 * rng.randn(*tex.shape).astype(np.float32))
        cam = ts.HILL_CAM.copy()
        cam[9:12] = [rng.uniform(-3, 3), rng.uniform(-1, 1), rng.uniform(1.5, 6.0)]
        ref = golden.scene("hill", tex, v, f, cam, channels=C)
        out = dirt_b200.hill_batch(torch.from_numpy(tex).cuda(), torch.from_numpy(v).cuda(), torch.from_numpy(c).cuda(),
                                   torch.from_numpy(f).cuda(), cam, channels=C).cpu().numpy()
        scale = 1.0
    else:
        C = 3
        bg, v, c, f = ts.quad(B, H, W, C, seed=seed, noise=float(rng.choice([0.0, 0.3, 1.0])))
        base = np.array(ts.CAM_A if rng.randint(2) else ts.CAM_B, np.float64)
        cam = base + rng.randn(16) * np.array([40, 30, 40, 0.1, 0.2, 0.1, 2, 0.3, 1, 0.05, 20, 5, 10, 0.1, 0.1, 0.1])
        cam[1] = max(cam[1], 85.0)                       # stay above the swell
        cam = cam.astype(np.float32)
        ref = golden.scene(name, bg, v, f, cam)
        out = getattr(dirt_b200, name + "_batch")(torch.from_numpy(bg).cuda(), torch.from_numpy(v).cuda(), torch.from_numpy(c).cuda(),
                                                  torch.from_numpy(f).cuda(), cam).cpu().numpy()
        scale = max(1.0, float(np.nanmax(np.abs(ref)))) if name == "oceanic_opt_flow" else 1.0
    assert out.shape == ref.shape
    mean, beyond = ts._stats(out / scale, ref / scale, 1e-4)
    assert mean <= 1e-6 and beyond <= 0.002, (name, seed, mean, beyond)


@pytest.mark.parametrize("F,size", [(4000, 3.0), (3000, 30.0), (9000, 2.0)])
def test_work_list_overflow_paths(F, size):
    """More large / clipped triangles than the workspace's lists hold (their capacities bound memory,
    not correctness): the in-thread fallbacks of the large-triangle list and of the tile queues
    must give the same image."""
    import torch
    from dirt_b200 import rasterise_ops as ops
    sc = scenes.random_triangles(1, 64, 80, 3, F, seed=F, size=size, w_range=(-0.6, 1.0))
    sc = {k: np.ascontiguousarray(v).copy() for k, v in sc.items()}
    t = {k: torch.from_numpy(v).cuda() for k, v in sc.items()}
    px, fid, gb = ops.forward_raw(t["background"], t["vertices"], t["vertex_colors"], t["faces"])
    opx, ofid, ogb = golden.forward(sc["background"], sc["vertices"], sc["vertex_colors"], sc["faces"], threads=8)
    assert (ofid >= 0).mean() > 0.9
    assert np.array_equal(fid.cpu().numpy(), ofid), "face ids differ at %d pixels" % (fid.cpu().numpy() != ofid).sum()
    np.testing.assert_allclose(px.cpu().numpy(), opx, rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(gb.cpu().numpy(), ogb, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("seed", range(16))
def test_fuzz_vertex_normals(seed):
    """dirt_vertex_normals (+grad) on random topology against oracle/lighting.py (float64) and, for
    the gradient, against float64 autograd of the reference's formula."""
    import torch
    from dirt_b200 import lighting
    from oracle import lighting as ol
    from test_lighting import _torch_reference
    rng = np.random.RandomState(15000 + seed)
    Bn, V, D = int(rng.randint(1, 4)), int(rng.randint(3, 400)), int(rng.choice([3, 4]))
    F = int(rng.randint(1, 4 * V))
    v = rng.randn(Bn, V, D).astype(np.float32)
    f = np.stack([rng.permutation(V)[:3] for _ in range(F)]).astype(np.int32)            # three distinct corners
    vt = torch.tensor(v, device="cuda", requires_grad=True)
    ft = torch.tensor(f, device="cuda")
    out = lighting.vertex_normals(vt, ft)
    np.testing.assert_allclose(out.detach().cpu().numpy(), ol.vertex_normals(v[..., :3], f), atol=5e-6)
    g = torch.tensor(rng.randn(Bn, V, 3).astype(np.float32), device="cuda")
    out.backward(g)
    v64 = torch.tensor(v, dtype=torch.float64, device="cuda", requires_grad=True)
    _torch_reference(v64, ft, True).backward(g.double())
    exp, got = v64.grad.cpu().numpy(), vt.grad.cpu().numpy()
    assert np.abs(got[..., :3] - exp[..., :3]).max() <= 2e-4 * np.abs(exp).max()
    if D == 4:
        assert not got[..., 3].any()


def _spiky_case(seed):
    """The headline's kind of geometry -- a UV sphere with per-vertex radial noise: long slivers seen edge-on, depth
    complexity ~10 -- at random sizes, zooms and offsets: partly outside the frame, so bounding boxes are cut by the
    viewport and the 8-line chunks of the second occlusion test start and end inside pyramid tiles; sometimes so close
    that faces cross the near plane."""
    rng = np.random.RandomState(9000 + seed)
    B = int(rng.randint(1, 3))
    H, W = int(rng.randint(48, 260)), int(rng.randint(48, 260))
    n_lat, n_lon = int(rng.randint(46, 90)), int(rng.randint(46, 90))
    sc = scenes.sphere_frames(B, H, W, 3, n_lat, n_lon, fill=float(rng.uniform(0.3, 0.75)),
                              seeds=(seed + 1, seed + 2, seed + 3, seed + 4))
    sc = {k: np.ascontiguousarray(v).copy() for k, v in sc.items()}
    zoom = float(rng.choice([0.6, 1.0, 1.7, 2.6]))
    off = rng.uniform(-0.7, 0.7, size=2) if zoom > 1.0 else rng.uniform(-0.3, 0.3, size=2)
    v = sc["vertices"]
    v[..., 0] = v[..., 0] * zoom + off[0] * v[..., 3]
    v[..., 1] = v[..., 1] * zoom + off[1] * v[..., 3]
    if rng.randint(4) == 0:
        v[..., 3] -= 2.2          # the camera moves into the object: w <= 0 for part of it, faces cross the near plane
        v[..., 2] -= 2.2
    return sc


@pytest.mark.parametrize("seed", range(24))
def test_fuzz_spiky_spheres_with_and_without_a_hint(seed):
    import torch
    from dirt_b200 import rasterise_ops as ops
    sc = _spiky_case(seed)
    t = {k: torch.from_numpy(v).cuda() for k, v in sc.items()}
    B, V, F = sc["vertices"].shape[0], sc["vertices"].shape[1], sc["faces"].shape[1]
    px, fid, gb = ops.forward_raw(t["background"], t["vertices"], t["vertex_colors"], t["faces"])
    opx, ofid, ogb = golden.forward(sc["background"], sc["vertices"], sc["vertex_colors"], sc["faces"], threads=8)
    assert np.array_equal(fid.cpu().numpy(), ofid), "face ids differ at %d pixels" % (fid.cpu().numpy() != ofid).sum()
    np.testing.assert_allclose(px.cpu().numpy(), opx, rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(gb.cpu().numpy(), ogb, rtol=1e-5, atol=1e-6)
    # a hint of random bytes, then the hint that call left behind: same bits both times
    hint = ops.visibility_hint(t["vertices"].device, B, F, key="fuzz-spiky")
    hint.copy_(torch.randint(0, 256, (hint.numel(),), dtype=torch.uint8, device=hint.device,
                             generator=torch.Generator(device=hint.device).manual_seed(seed)))
    for _ in range(2):
        hp, hf, hg = ops.forward_raw(t["background"], t["vertices"], t["vertex_colors"], t["faces"], hint=hint)
        assert torch.equal(hf, fid) and torch.equal(hp, px) and torch.equal(hg, gb)
    if seed % 4 == 0:
        g = np.random.RandomState(seed).randn(*opx.shape).astype(np.float32)
        gv, gc, gbg = ops.backward_raw(t["faces"], px, torch.from_numpy(g).cuda(), fid, gb, V)
        ogv, ogc, ogbg = golden.backward(sc["faces"], opx, g, ofid, ogb, V, threads=8)
        assert np.array_equal(gbg.cpu().numpy(), ogbg)
        for a, b in ((gv.cpu().numpy(), ogv), (gc.cpu().numpy(), ogc)):
            assert np.abs(a - b).max() <= 1e-4 * max(np.abs(b).max(), 1e-12)
