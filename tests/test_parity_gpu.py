"""GPU: the CUDA path (through the C ABI) against the CPU oracle on identical seeded inputs.

Bars (BASELINE.json north_star): face-id / coverage maps bit-exact; interpolated attributes
<= 1e-5 relative (abs floor 1e-6); gradients <= 1e-4 relative to max|grad| per tensor.
"""
import numpy as np
import pytest
import torch

import scenes
from oracle import golden

pytestmark = pytest.mark.gpu

ATTR_RTOL, ATTR_ATOL = 1e-5, 1e-6
GRAD_RTOL = 1e-4


def _cuda(sc):
    return {k: torch.from_numpy(np.ascontiguousarray(v)).cuda() for k, v in sc.items() if k != "expected"}


def _run_forward(sc, aux=True):
    from dirt_b200 import rasterise_ops as ops
    t = _cuda(sc)
    px, fid, gb = ops.forward_raw(t["background"], t["vertices"], t["vertex_colors"], t["faces"], want_aux=aux)
    torch.cuda.synchronize()
    return px, fid, gb, t


def _check_forward(sc, threads=8):
    px, fid, gb, t = _run_forward(sc)
    opx, ofid, ogb = golden.forward(sc["background"], sc["vertices"], sc["vertex_colors"], sc["faces"], threads=threads)
    fid_n = fid.cpu().numpy()
    assert np.array_equal(fid_n, ofid), "face-id maps differ at %d pixels" % (fid_n != ofid).sum()
    px_n, gb_n = px.cpu().numpy(), gb.cpu().numpy()
    fin = np.isfinite(opx)
    assert np.array_equal(np.isfinite(px_n), fin)
    assert np.array_equal(px_n[~fin], opx[~fin])
    np.testing.assert_allclose(px_n[fin], opx[fin], rtol=ATTR_RTOL, atol=ATTR_ATOL)
    np.testing.assert_allclose(gb_n, ogb, rtol=ATTR_RTOL, atol=ATTR_ATOL)
    # window depth feeds the depth|face keys: it is specified operation by operation, so bit-exact
    assert np.array_equal(gb_n[..., 3].view(np.int32), ogb[..., 3].view(np.int32)), "window depth differs in bits"
    return (px, fid, gb, t), (opx, ofid, ogb)


def _check_backward(sc, dev, ora, seed=0, shared=False):
    from dirt_b200 import rasterise_ops as ops
    px, fid, gb, t = dev
    opx, ofid, ogb = ora
    g = np.random.RandomState(seed).randn(*opx.shape).astype(np.float32)
    g[~np.isfinite(opx)] = 0
    V = sc["vertices"].shape[1]
    gv, gc, gbg = ops.backward_raw(t["faces"], px, torch.from_numpy(g).cuda(), fid, gb, V,
                                   shared_vertices=shared, shared_colors=shared)
    torch.cuda.synchronize()
    ogv, ogc, ogbg = golden.backward(sc["faces"], opx, g, ofid, ogb, V, shared_vertices=shared, shared_colors=shared, threads=8)
    assert np.array_equal(gbg.cpu().numpy(), ogbg)
    for name, a, b in (("grad_vertices", gv.cpu().numpy(), ogv), ("grad_vertex_colors", gc.cpu().numpy(), ogc)):
        scale = max(np.abs(b).max(), 1e-12)
        err = np.abs(a - b).max() / scale
        assert err <= GRAD_RTOL, "%s: max err %.3g of max|grad|" % (name, err)


def test_kat1_and_kat2_bit_exact():
    for sc in (scenes.kat1_square(), scenes.kat2_square()):
        px, _, _, _ = _run_forward(sc)
        assert np.array_equal(px.cpu().numpy(), sc["expected"])


def test_c1_square_forward_only_without_aux():
    sc = scenes.kat1_square()
    px, fid, gb, _ = _run_forward(sc, aux=False)
    assert fid is None and gb is None
    assert np.array_equal(px.cpu().numpy(), sc["expected"])


@pytest.mark.parametrize("C", [1, 3])
def test_cube_c2_forward_backward(C):
    sc = scenes.cube(B=1, H=256, W=256, C=C)
    dev, ora = _check_forward(sc)
    _check_backward(sc, dev, ora)


def test_cube_c3_deferred_shapes():
    # config c3 geometry: B=16, 512x512, three attribute sets incl. a -inf background
    sc = scenes.cube(B=16, H=512, W=512, C=3)
    world = np.random.RandomState(1).randn(*sc["vertex_colors"].shape).astype(np.float32) * 3
    for cols, bgval in ((world, -np.inf), (sc["vertex_colors"], 0.0)):
        s2 = dict(sc)
        s2["vertex_colors"] = cols
        s2["background"] = np.full_like(sc["background"], bgval)
        dev, ora = _check_forward(s2)
        _check_backward(s2, dev, ora)


def test_cylinder_scene():
    sc = scenes.cylinder()
    dev, ora = _check_forward(sc)
    _check_backward(sc, dev, ora)


@pytest.mark.parametrize("F,size,ties", [(300, 0.6, False), (3000, 0.08, False), (400, 0.5, True), (20000, 0.02, False)])
def test_triangle_soup(F, size, ties):
    sc = scenes.random_triangles(3, 120, 200, 3, F, seed=F, size=size, depth_ties=ties)
    dev, ora = _check_forward(sc)
    _check_backward(sc, dev, ora)


def test_ragged_frame_sizes_and_single_channel():
    for (H, W) in ((1, 1), (7, 33), (17, 16), (31, 250)):
        sc = scenes.random_triangles(2, H, W, 1, 64, seed=H * W, size=0.9)
        dev, ora = _check_forward(sc)
        _check_backward(sc, dev, ora)


def test_wide_channels_generic_path():
    sc = scenes.random_triangles(2, 64, 64, 9, 100, seed=4)
    dev, ora = _check_forward(sc)
    _check_backward(sc, dev, ora)


def test_near_and_far_plane_clipping():
    sc = scenes.near_crossing()
    dev, ora = _check_forward(sc)
    assert (ora[1] >= 0).mean() > 0.2
    _check_backward(sc, dev, ora)


def test_clipped_fans_that_fold_over_themselves_at_the_guard_band():
    """Found by tests/test_fuzz_gpu.py: a corner behind the eye (w < 0, outside near AND far) leaves
    polygon corners with w -> 0 whose window positions are clamped; the fan then overlaps itself
    and the attributes must come from the fan triangle that won the depth test (RASTER_SPEC 2.3)."""
    faces3 = scenes.FOLDED_FAN_FACES
    rng = np.random.RandomState(0)
    for (H, W) in ((74, 76), (36, 92)):
        v = np.asarray(faces3, np.float32).reshape(1, 9, 4)
        sc = {"background": rng.uniform(0, 1, (1, H, W, 3)).astype(np.float32), "vertices": v,
              "vertex_colors": rng.uniform(0, 1, (1, 9, 3)).astype(np.float32),
              "faces": np.arange(9, dtype=np.int32).reshape(1, 3, 3)}
        dev, ora = _check_forward(sc)
        assert (ora[1] >= 0).sum() > 100
        _check_backward(sc, dev, ora)


def test_sphere_small_triangles():
    sc = scenes.sphere(2, 256, 256, n_lat=96, n_lon=96)
    dev, ora = _check_forward(sc)
    _check_backward(sc, dev, ora)


def test_empty_inputs():
    from dirt_b200 import rasterise_ops as ops
    bg = torch.rand(2, 16, 16, 3, device="cuda")
    px, fid, gb = ops.forward_raw(bg, torch.zeros(2, 0, 4, device="cuda"), torch.zeros(2, 0, 3, device="cuda"),
                                  torch.zeros(2, 0, 3, dtype=torch.int32, device="cuda"))
    assert torch.equal(px, bg) and (fid == -1).all()
    px, fid, gb = ops.forward_raw(bg[:0], torch.zeros(0, 3, 4, device="cuda"), torch.zeros(0, 3, 3, device="cuda"),
                                  torch.zeros(0, 1, 3, dtype=torch.int32, device="cuda"))
    assert px.shape == (0, 16, 16, 3)


def test_out_of_range_and_nonfinite_faces_are_skipped():
    sc = scenes.random_triangles(1, 64, 64, 3, 50, seed=11)
    sc["faces"][0, 3] = [0, 1, 10 ** 6]
    sc["faces"][0, 4] = [-1, 1, 2]
    sc["vertices"][0, 30] = np.nan
    sc["vertices"][0, 33, 0] = np.inf
    _check_forward(sc)


def test_broadcast_shared_mesh_and_summed_gradients():
    sc = scenes.cube(B=4, H=64, W=64)
    sc["faces"] = sc["faces"][:1].copy()
    sc["vertices"] = np.repeat(sc["vertices"][:1], 1, 0)
    sc["vertex_colors"] = sc["vertex_colors"][:1].copy()
    dev, ora = _check_forward(sc)
    _check_backward(sc, dev, ora, shared=True)


@pytest.mark.parametrize("C", [1, 3, 5])
def test_shared_mesh_backward_folds_to_world_space(C):
    """one call sums grad_clip[b] @ M_b^T and the colour gradients of all frames
    into the packed [V, stride] buffer.  Checked against the oracle's per-frame gradients folded
    with numpy; bar as for gradients (1e-4 of max|grad|)."""
    from dirt_b200 import rasterise_ops as ops
    sc = scenes.hill(3, 96, 128, C=C, n=40, seed=2)
    M = sc.pop("matrices")
    sc.pop("world")
    dev, ora = _check_forward(sc)
    px, fid, gb, t = dev
    opx, ofid, ogb = ora
    g = np.random.RandomState(3).randn(*opx.shape).astype(np.float32)
    V = sc["vertices"].shape[1]
    packed, gbg = ops.backward_shared_raw(t["faces"], px, torch.from_numpy(g).cuda(), fid, gb, V, torch.from_numpy(M).cuda())
    torch.cuda.synchronize()
    faces_b = np.broadcast_to(sc["faces"], (3,) + sc["faces"].shape[1:])
    ogv, ogc, ogbg = golden.backward(np.ascontiguousarray(faces_b), opx, g, ofid, ogb, V, threads=8)
    exp_w = np.einsum("bvj,bij->vi", ogv.astype(np.float64), M.astype(np.float64))
    exp_c = ogc.astype(np.float64).sum(0)
    got = packed.cpu().numpy()
    assert got.shape == (V, (4 + C + 3) // 4 * 4)
    assert np.array_equal(gbg.cpu().numpy(), ogbg)
    for name, a, b in (("grad_world", got[:, :4], exp_w), ("grad_vertex_colors", got[:, 4:4 + C], exp_c)):
        err = np.abs(a - b).max() / max(np.abs(b).max(), 1e-12)
        assert err <= GRAD_RTOL, "%s: max err %.3g of max|grad|" % (name, err)
    assert not got[:, 4 + C:].any()


def test_public_shared_mesh_api_matches_the_per_frame_autograd_path():
    """dirt_b200.rasterise_batch_shared (fused backward) against rasterise_batch fed with
    world @ M_b through torch autograd (per-frame gradients, then matmul backward)."""
    import dirt_b200
    sc = scenes.hill(3, 80, 96, C=3, n=36, seed=4)
    dev = {k: torch.from_numpy(np.ascontiguousarray(v)).cuda() for k, v in sc.items()}
    target = torch.rand(3, 80, 96, 3, device="cuda")

    def run(fused):
        w = dev["world"].clone().requires_grad_(True)
        c = dev["vertex_colors"][0].clone().requires_grad_(True)
        bg = dev["background"].clone().requires_grad_(True)
        if fused:
            px = dirt_b200.rasterise_batch_shared(bg, w, dev["matrices"], c, dev["faces"][0])
        else:
            px = dirt_b200.rasterise_batch(bg, torch.matmul(w.unsqueeze(0), dev["matrices"]), c.unsqueeze(0), dev["faces"])
        (0.5 * ((px - target) ** 2).sum()).backward()
        return px.detach(), w.grad, c.grad, bg.grad

    pa, wa, ca, ba = run(True)
    pb, wb, cb, bb = run(False)
    assert torch.equal(pa, pb) and torch.equal(ba, bb)
    for a, b in ((wa, wb), (ca, cb)):
        assert float((a - b).abs().max()) <= GRAD_RTOL * float(b.abs().max())


def test_determinism_run_to_run():
    sc = scenes.sphere(2, 128, 128, n_lat=64, n_lon=64)
    a = _run_forward(sc)
    b = _run_forward(sc)
    for x, y in zip(a[:3], b[:3]):
        assert torch.equal(x, y)


def test_forward_and_backward_replay_from_a_cuda_graph():
    """Every C-ABI call is asynchronous on the caller's stream and never synchronises with the
    host (the reference blocks on a GL thread and a cudaMemcpy, csrc/gl_dispatcher.h:58-67,
    csrc/rasterise_egl.cpp:323), so a whole fwd+bwd step can be captured in a CUDA graph and
    replayed on new inputs -- the way to run launch-bound small frames (configs c1, c2)."""
    from dirt_b200 import rasterise_ops as ops
    for sc in (scenes.cube(B=1, H=96, W=96), scenes.sphere(2, 128, 128, n_lat=64, n_lon=64)):
        t = _cuda(sc)
        V = sc["vertices"].shape[1]
        g = torch.randn(sc["background"].shape, device="cuda")
        static_v = t["vertices"].clone()
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):                         # warm-up grows the workspace outside the capture
            px, fid, gb = ops.forward_raw(t["background"], static_v, t["vertex_colors"], t["faces"])
            ops.backward_raw(t["faces"], px, g, fid, gb, V)
        torch.cuda.current_stream().wait_stream(side)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            px, fid, gb = ops.forward_raw(t["background"], static_v, t["vertex_colors"], t["faces"])
            gv, gc, gbg = ops.backward_raw(t["faces"], px, g, fid, gb, V)
        for k in range(3):
            moved = t["vertices"].clone()
            moved[..., 0] += 0.05 * k * moved[..., 3]         # shift the mesh in clip space
            static_v.copy_(moved)
            graph.replay()
            torch.cuda.synchronize()
            epx, efid, egb = ops.forward_raw(t["background"], moved, t["vertex_colors"], t["faces"])
            egv, egc, egbg = ops.backward_raw(t["faces"], epx, g, efid, egb, V)
            assert torch.equal(fid, efid) and torch.equal(px, epx) and torch.equal(gbg, egbg)
            for a, b in ((gv, egv), (gc, egc)):
                assert float((a - b).abs().max()) <= GRAD_RTOL * max(float(b.abs().max()), 1e-12)


def test_background_read_in_place_from_pinned_host_memory():
    """The resolve kernel touches the background only where no face covers the pixel, so the
    caller may leave it in pinned host memory: same pixels bit for bit, the gradient w.r.t. the
    background comes back on the host."""
    import dirt_b200
    from dirt_b200 import rasterise_ops as ops
    sc = scenes.sphere(3, 96, 128, n_lat=32, n_lon=32)
    t = _cuda(sc)
    bg_host = torch.from_numpy(sc["background"]).pin_memory()
    ref, fid, _ = ops.forward_raw(t["background"], t["vertices"], t["vertex_colors"], t["faces"])
    px, fid2, _ = ops.forward_raw(bg_host, t["vertices"], t["vertex_colors"], t["faces"])
    assert torch.equal(fid, fid2) and torch.equal(px, ref)
    assert 0.05 < float((fid < 0).float().mean()) < 0.95          # both kinds of pixel are exercised
    bg_leaf = bg_host.clone().pin_memory().requires_grad_(True)
    v = t["vertices"].clone().requires_grad_(True)
    out = dirt_b200.rasterise_batch(bg_leaf, v, t["vertex_colors"], t["faces"])
    assert out.is_cuda and torch.equal(out.detach(), ref)
    (out * 2.0).sum().backward()
    assert not bg_leaf.grad.is_cuda and v.grad.is_cuda
    assert torch.equal(bg_leaf.grad, (2.0 * (fid < 0)[..., None].float().expand_as(ref)).cpu())
    one = dirt_b200.rasterise(bg_host[0], sc["vertices"][0], sc["vertex_colors"][0], sc["faces"][0])
    assert torch.equal(one, ref[0])


def test_public_api_autograd_matches_raw_calls():
    import dirt_b200
    sc = scenes.cube(B=2, H=64, W=64)
    t = _cuda(sc)
    v = t["vertices"].clone().requires_grad_(True)
    c = t["vertex_colors"].clone().requires_grad_(True)
    bg = t["background"].clone().requires_grad_(True)
    px = dirt_b200.rasterise_batch(bg, v, c, t["faces"])
    g = torch.randn_like(px)
    (px * g).sum().backward()
    gbg, gv, gc = dirt_b200.rasterise_grad(t["background"], t["vertices"], t["vertex_colors"], t["faces"], grad_pixels=g)
    assert torch.equal(bg.grad, gbg)
    torch.testing.assert_close(v.grad, gv, rtol=1e-4, atol=1e-5 * float(gv.abs().max()))
    torch.testing.assert_close(c.grad, gc, rtol=1e-4, atol=1e-5 * float(gc.abs().max()))
    # unbatched call = batch of one (dirt/rasterise_ops.py:50-54)
    p1 = dirt_b200.rasterise(t["background"][0], t["vertices"][0], t["vertex_colors"][0], t["faces"][0], None)
    assert torch.equal(p1, px[0].detach())


def test_reference_validation_errors():
    import dirt_b200
    bg = torch.zeros(8, 8, 3, device="cuda")
    v = torch.zeros(3, 4, device="cuda")
    with pytest.raises(dirt_b200.InvalidArgumentError):
        dirt_b200.rasterise(bg, v[:, :3], torch.zeros(3, 3, device="cuda"), [[0, 1, 2]])
    with pytest.raises(dirt_b200.InvalidArgumentError):
        dirt_b200.rasterise(bg, v, torch.zeros(3, 1, device="cuda"), [[0, 1, 2]])
    with pytest.raises(dirt_b200.InvalidArgumentError):
        dirt_b200.rasterise(bg, v, torch.zeros(3, 3, device="cuda"), [[0, 1]])
    with pytest.raises(dirt_b200.InvalidArgumentError):
        dirt_b200.rasterise_batch(bg[None].repeat(2, 1, 1, 1), v[None].repeat(3, 1, 1), torch.zeros(3, 3, 3, device="cuda"),
                                  torch.zeros(3, 1, 3, dtype=torch.int32, device="cuda"))


def test_host_buffer_entry_points():
    import ctypes
    from dirt_b200 import _lib
    lib = _lib.load()
    sc = scenes.cube(B=2, H=48, W=48)
    B, H, W, C = sc["background"].shape
    V, F = sc["vertices"].shape[1], sc["faces"].shape[1]
    px = np.empty((B, H, W, C), np.float32); fid = np.empty((B, H, W), np.int32); gb = np.empty((B, H, W, 4), np.float32)
    p = lambda a: a.ctypes.data
    fa = _lib.DirtForwardArgs(_lib.DirtShape(B, H, W, C, V, F), p(sc["background"]), H * W * C, p(sc["vertices"]), V * 4,
                              p(sc["vertex_colors"]), V * C, p(sc["faces"]), F * 3, p(px), p(fid), p(gb), None, 0)
    _lib.check(lib.dirt_rasterise_forward_host(ctypes.byref(fa)))
    opx, ofid, ogb = golden.forward(sc["background"], sc["vertices"], sc["vertex_colors"], sc["faces"])
    assert np.array_equal(fid, ofid)
    np.testing.assert_allclose(px, opx, rtol=ATTR_RTOL, atol=ATTR_ATOL)
    g = np.random.RandomState(0).randn(B, H, W, C).astype(np.float32)
    gv = np.empty((B, V, 4), np.float32); gc = np.empty((B, V, C), np.float32); gbg = np.empty((B, H, W, C), np.float32)
    ba = _lib.DirtBackwardArgs(_lib.DirtShape(B, H, W, C, V, F), p(sc["faces"]), F * 3, p(px), p(g), p(fid), p(gb),
                               p(gv), V * 4, p(gc), V * C, p(gbg))
    _lib.check(lib.dirt_rasterise_backward_host(ctypes.byref(ba)))
    ogv, ogc, ogbg = golden.backward(sc["faces"], opx, g, ofid, ogb, V)
    assert np.array_equal(gbg, ogbg)
    assert np.abs(gv - ogv).max() <= GRAD_RTOL * np.abs(ogv).max()
    assert np.abs(gc - ogc).max() <= GRAD_RTOL * np.abs(ogc).max()
    _lib.check(lib.dirt_release_host_staging())


_PLAN_CHECK = """
import sys
sys.path[:0] = [{root!r}, {tests!r}]
import scenes, test_parity_gpu as t
t._check_forward(scenes.sphere(2, 192, 160, n_lat=80, n_lon=80))
t._check_forward(scenes.random_triangles(2, 200, 136, 3, 6000, seed=11, size=0.12, depth_ties=True))
t._check_forward(scenes.near_crossing(n=56))      # 6050 faces, many clipped by the near / far plane
print("plan ok")
"""


@pytest.mark.parametrize("plan,min_overdraw,facing", [("0", "0", "0"), ("1", "0", "0"), ("0.5,1", "0", "0"),
                                                      ("0.05,0.1,0.2,0.3,0.4,0.6,0.8,1", "0", "0"),
                                                      ("0.35,1", "1.5", "0"), ("0.35,1", "1e9", "0"),
                                                      ("0.1,0.2,0.35;0.5,0.7,1", "0", "0"), ("0.2,0.4;1", "0", "1"), ("0.5,1", "0", "1")])
def test_depth_bucket_plans_do_not_change_the_result(plan, min_overdraw, facing):
    """Meshes of >= 4096 faces are sampled front to back in depth buckets, with triangles hidden
    behind the depth pyramid dropped (dirt_b200/csrc/depth_order.cu).  The cull is conservative by
    construction, so every plan -- unordered ("0"), one ordered pass, two and eight buckets --
    must reproduce the oracle bit for bit (face ids, window depth), depth ties included.
    DIRT_B200_MIN_OVERDRAW = 0 forces the ordered kernels whatever overdraw the scene is
    estimated to have; 1.5 lets the estimate choose per scene, 1e9 always falls back to the
    in-kernel unordered pass.  "a,b;c,1" puts several buckets into one launch (';' ends a launch);
    DIRT_B200_FACING=1 sends the faces of the winding that looks away to the last bucket.
    The plan is read once per process, hence the subprocess."""
    import os
    import subprocess
    import sys
    tests = os.path.dirname(os.path.abspath(__file__))
    env = dict(os.environ, DIRT_B200_BUCKETS=plan, DIRT_B200_MIN_OVERDRAW=min_overdraw, DIRT_B200_FACING=facing)
    r = subprocess.run([sys.executable, "-c", _PLAN_CHECK.format(root=os.path.dirname(tests), tests=tests)],
                       env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "plan ok" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]


def test_concurrent_streams_do_not_share_state():
    """Two forward + backward pairs in flight on two streams (each stream has its own workspace; the
    library keeps no per-call state of its own): results equal those of the same calls run alone."""
    from dirt_b200 import rasterise_ops as ops
    a = _cuda(scenes.sphere(2, 200, 216, n_lat=80, n_lon=80))
    b = _cuda(scenes.random_triangles(3, 160, 120, 3, 5000, seed=3, size=0.3))
    ga = torch.randn(2, 200, 216, 3, device="cuda")
    gb_ = torch.randn(3, 160, 120, 3, device="cuda")

    def run(t, g):
        px, fid, gbuf = ops.forward_raw(t["background"], t["vertices"], t["vertex_colors"], t["faces"])
        gv, gc, gbg = ops.backward_raw(t["faces"], px, g, fid, gbuf, t["vertices"].shape[1])
        return px, fid, gbuf, gv, gc, gbg

    ref_a, ref_b = run(a, ga), run(b, gb_)
    torch.cuda.synchronize()
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    for _ in range(6):
        with torch.cuda.stream(s1):
            out_a = run(a, ga)
        with torch.cuda.stream(s2):
            out_b = run(b, gb_)
        torch.cuda.synchronize()
        for ref, out in ((ref_a, out_a), (ref_b, out_b)):
            for k in (0, 1, 2, 5):                       # pixels, face ids, G-buffer, grad_background: bit for bit
                assert torch.equal(ref[k], out[k])
            for k in (3, 4):                             # scattered float sums
                assert (ref[k] - out[k]).abs().max() <= 1e-4 * ref[k].abs().max()
    ops.release_workspaces()
