"""Memory safety of the C-ABI calls without a sanitizer (compute-sanitizer is closed on this pool):
every buffer of a forward + backward pair -- inputs, outputs and a workspace of EXACTLY
dirt_workspace_bytes() -- is carved out of one arena with sentinel-filled guard zones on both
sides.  After the calls every guard must be untouched, the inputs unchanged, and the results must
not depend on what the workspace held before (all-ones bits vs zeros): nothing is read before the
library initialised it.  Frame, mesh and batch sizes are odd on purpose."""
import ctypes

import numpy as np
import pytest

import scenes

pytestmark = pytest.mark.gpu

GUARD = 4096            # bytes, multiple of 256 so that the carved buffers stay 256-byte aligned
SENTINEL = 0xA5


class Arena:
    def __init__(self, nbytes):
        import torch
        self.buf = torch.full((nbytes,), SENTINEL, dtype=torch.uint8, device="cuda")
        self.off = GUARD
        self.spans = []                     # (offset, nbytes) of carved buffers

    def carve(self, nbytes):
        off = self.off
        self.spans.append((off, nbytes))
        self.off = (off + nbytes + GUARD + 255) // 256 * 256
        assert self.off <= self.buf.numel()
        return self.buf[off:off + nbytes]

    def put(self, array):
        import torch
        t = self.carve(array.nbytes)
        t.copy_(torch.from_numpy(np.array(array, copy=True, order="C")).view(-1).view(torch.uint8))
        return t

    def guards_intact(self):
        import torch
        mask = torch.ones(self.buf.numel(), dtype=torch.bool, device="cuda")
        for off, n in self.spans:
            mask[off:off + n] = False
        return bool((self.buf[mask] == SENTINEL).all())


def _run(sc, poison):
    import torch
    from dirt_b200 import _lib
    from dirt_b200._lib import DirtShape, DirtForwardArgs, DirtBackwardArgs
    lib = _lib.load()
    B, H, W, C = sc["background"].shape
    V, F = sc["vertices"].shape[1], sc["faces"].shape[1]
    shape = DirtShape(B, H, W, C, V, F)
    ws_bytes = lib.dirt_workspace_bytes(ctypes.byref(shape))
    assert ws_bytes > 0
    g = np.random.RandomState(0).randn(B, H, W, C).astype(np.float32)
    out_bytes = {"pixels": B * H * W * C * 4, "face_id": B * H * W * 4, "gbuffer": B * H * W * 16,
                 "gv": B * V * 16, "gc": B * V * C * 4, "gbg": B * H * W * C * 4}
    total = (sum(v.nbytes for k, v in sc.items() if k != "expected") + g.nbytes + sum(out_bytes.values()) + ws_bytes
             + 16 * (GUARD + 256))
    arena = Arena(total)
    ins = {k: arena.put(sc[k]) for k in ("background", "vertices", "vertex_colors", "faces")}
    ws = arena.carve(ws_bytes)
    ws.fill_(poison)
    outs = {k: arena.carve(n) for k, n in out_bytes.items()}
    gp = arena.put(g)
    for t in outs.values():
        t.fill_(0x5A)                       # outputs start as junk too
    stream = torch.cuda.current_stream().cuda_stream
    fwd = DirtForwardArgs(shape,
                          ins["background"].data_ptr(), H * W * C, ins["vertices"].data_ptr(), V * 4,
                          ins["vertex_colors"].data_ptr(), V * C, ins["faces"].data_ptr(), F * 3,
                          outs["pixels"].data_ptr(), outs["face_id"].data_ptr(), outs["gbuffer"].data_ptr(),
                          ws.data_ptr(), ws_bytes)
    _lib.check(lib.dirt_rasterise_forward(ctypes.byref(fwd), stream))
    bwd = DirtBackwardArgs(shape, ins["faces"].data_ptr(), F * 3,
                           outs["pixels"].data_ptr(), gp.data_ptr(), outs["face_id"].data_ptr(), outs["gbuffer"].data_ptr(),
                           outs["gv"].data_ptr(), V * 4, outs["gc"].data_ptr(), V * C, outs["gbg"].data_ptr())
    _lib.check(lib.dirt_rasterise_backward(ctypes.byref(bwd), stream))
    torch.cuda.synchronize()
    assert arena.guards_intact(), "a kernel wrote outside its buffers"
    for k, t in ins.items():
        assert np.array_equal(t.cpu().numpy(), np.ascontiguousarray(sc[k]).view(np.uint8).reshape(-1)), k + " was modified"
    assert np.array_equal(gp.cpu().numpy(), g.view(np.uint8).reshape(-1))
    return {k: t.cpu().numpy().copy() for k, t in outs.items()}


CASES = {
    "soup_odd": lambda: scenes.random_triangles(3, 37, 53, 3, 199, seed=7, size=0.7),
    "soup_big_triangles": lambda: scenes.random_triangles(2, 129, 257, 3, 61, seed=8, size=2.5),
    "soup_many_small": lambda: scenes.random_triangles(3, 120, 200, 3, 20001, seed=9, size=0.02),
    "single_channel_1x1": lambda: scenes.random_triangles(2, 1, 1, 1, 5, seed=10, size=0.9),
    "seven_channels": lambda: scenes.random_triangles(2, 33, 65, 7, 77, seed=11),
    "near_crossing": lambda: scenes.near_crossing(),
    "sphere_ordered": lambda: scenes.sphere(3, 250, 230, n_lat=90, n_lon=70),
}


@pytest.mark.parametrize("case", sorted(CASES))
def test_guard_zones_inputs_and_workspace_independence(case):
    from oracle import golden
    sc = CASES[case]()
    a = _run(sc, 0xFF)
    b = _run(sc, 0x00)
    for k in ("pixels", "face_id", "gbuffer", "gbg"):
        assert np.array_equal(a[k], b[k]), "%s depends on what the workspace held before the call" % k
    for k in ("gv", "gc"):                  # scattered float reductions: the order of the sums is not fixed
        x, y = a[k].view(np.float32), b[k].view(np.float32)
        assert np.isfinite(x).all() and np.abs(x - y).max() <= 1e-4 * max(np.abs(x).max(), 1e-12), k
    B, H, W, C = sc["background"].shape
    _, ofid, _ = golden.forward(sc["background"], sc["vertices"], sc["vertex_colors"], sc["faces"], threads=8)
    assert np.array_equal(a["face_id"].view(np.int32).reshape(B, H, W), ofid)


def _f32(t):
    import torch
    return t.view(torch.float32)


def test_guard_zones_dilate_and_vertex_normals():
    import torch
    from dirt_b200 import _lib
    lib = _lib.load()
    stream = torch.cuda.current_stream().cuda_stream
    rs = np.random.RandomState(3)
    # silhouette dilation, odd frame
    B, H, W, C = 2, 19, 45, 3
    x = rs.randn(B, H, W, C).astype(np.float32)
    x[rs.rand(B, H, W) < 0.5] = -np.inf
    arena = Arena(6 * x.nbytes + 16 * (GUARD + 256))
    dx, go = arena.put(x), arena.put(rs.randn(B, H, W, C).astype(np.float32))
    out, gx = arena.carve(x.nbytes), arena.carve(x.nbytes)
    _lib.check(lib.dirt_dilate_silhouette(dx.data_ptr(), out.data_ptr(), B, H, W, C, stream))
    _lib.check(lib.dirt_dilate_silhouette_grad(dx.data_ptr(), go.data_ptr(), gx.data_ptr(), B, H, W, C, stream))
    torch.cuda.synchronize()
    assert arena.guards_intact(), "dilate wrote outside its buffers"
    assert np.array_equal(dx.cpu().numpy(), x.view(np.uint8).reshape(-1))
    assert torch.isfinite(_f32(gx)).all()

    # vertex normals, D = 3 and 4, with out-of-range faces
    for D in (3, 4):
        Bn, V, F = 3, 101, 333
        v = rs.randn(Bn, V, D).astype(np.float32)
        f = rs.randint(0, V, size=(F, 3)).astype(np.int32)
        f[5] = (-1, 2, 3)
        f[9] = (1, V, 3)
        arena = Arena(v.nbytes * 2 + f.nbytes + 5 * Bn * V * 12 + 16 * (GUARD + 256))
        dv, df = arena.put(v), arena.put(f)
        sums, normals, gs = arena.carve(Bn * V * 12), arena.carve(Bn * V * 12), arena.carve(Bn * V * 12)
        gn = arena.put(rs.randn(Bn, V, 3).astype(np.float32))
        gvt = arena.carve(v.nbytes)
        _lib.check(lib.dirt_vertex_normals(dv.data_ptr(), Bn, V, D, df.data_ptr(), F, sums.data_ptr(), normals.data_ptr(), stream))
        _lib.check(lib.dirt_vertex_normals_grad(dv.data_ptr(), Bn, V, D, df.data_ptr(), F, sums.data_ptr(), gn.data_ptr(),
                                                gs.data_ptr(), gvt.data_ptr(), stream))
        torch.cuda.synchronize()
        assert arena.guards_intact(), "vertex normals wrote outside their buffers (D=%d)" % D
        assert np.array_equal(dv.cpu().numpy(), v.view(np.uint8).reshape(-1)) and np.array_equal(df.cpu().numpy(), f.view(np.uint8).reshape(-1))
        assert torch.isfinite(_f32(normals)).all() and torch.isfinite(_f32(gvt)).all()


@pytest.mark.parametrize("scene,Cb,C", [(0, 3, 3), (3, 3, 3), (5, 1, 3), (6, 4, 3), (6, 1, 1), (4, 3, 4)])
def test_guard_zones_scene_shade(scene, Cb, C):
    """The scene stage after a forward over a full-screen quad: camera_pos holds exactly the floats
    the scene reads, the texture / jitter has its own channel count."""
    import torch
    from dirt_b200 import _lib
    from dirt_b200._lib import DirtShape, DirtForwardArgs, DirtSceneArgs
    lib = _lib.load()
    stream = torch.cuda.current_stream().cuda_stream
    rs = np.random.RandomState(scene)
    B, H, W, V, F = 2, 27, 41, 4, 2
    n_cam = lib.dirt_scene_camera_floats(scene)
    quad = np.tile(np.array([[-1, -1, 0, 1], [-1, 1, 0, 1], [1, 1, 0, 1], [1, -1, 0, 1]], np.float32), (B, 1, 1))
    faces = np.tile(np.array([[0, 1, 2], [0, 2, 3]], np.int32), (B, 1, 1))
    tex = rs.rand(B, H, W, Cb).astype(np.float32)
    cam = np.zeros(n_cam, np.float32)
    cam[1] = 150.0 if scene != 6 else 0.0
    if scene == 6:
        cam[11] = 3.0
    shape = DirtShape(B, H, W, C, V, F)
    ws_bytes = lib.dirt_workspace_bytes(ctypes.byref(shape))
    npx = B * H * W
    arena = Arena(tex.nbytes + npx * (C * 8 + 4 + 16 + 4 * C) + quad.nbytes + faces.nbytes + ws_bytes + B * V * C * 4 + 4096 + 20 * (GUARD + 256))
    d_tex, d_quad, d_faces, d_cam = arena.put(tex), arena.put(quad), arena.put(faces), arena.put(cam)
    d_col = arena.put(np.ones((B, V, C), np.float32))
    d_blank = arena.put(np.zeros((B, H, W, C), np.float32))
    px, fid, gb, ws = arena.carve(npx * C * 4), arena.carve(npx * 4), arena.carve(npx * 16), arena.carve(ws_bytes)
    fwd = DirtForwardArgs(shape, d_blank.data_ptr(), H * W * C, d_quad.data_ptr(), V * 4, d_col.data_ptr(), V * C,
                          d_faces.data_ptr(), F * 3, px.data_ptr(), fid.data_ptr(), gb.data_ptr(), ws.data_ptr(), ws_bytes)
    _lib.check(lib.dirt_rasterise_forward(ctypes.byref(fwd), stream))
    sa = DirtSceneArgs(shape, d_tex.data_ptr(), H * W * Cb, d_quad.data_ptr(), V * 4, d_faces.data_ptr(), F * 3,
                       fid.data_ptr(), gb.data_ptr(), d_cam.data_ptr(), px.data_ptr())
    _lib.check(lib.dirt_scene_shade(scene, ctypes.byref(sa), Cb, n_cam, stream))
    torch.cuda.synchronize()
    assert arena.guards_intact(), "scene %d wrote outside its buffers" % scene
    assert np.array_equal(d_tex.cpu().numpy(), tex.view(np.uint8).reshape(-1))
    assert np.array_equal(d_cam.cpu().numpy(), cam.view(np.uint8).reshape(-1))
    assert (_f32(fid).view(torch.int32) >= 0).all()
    # one float fewer than the scene reads is refused
    assert lib.dirt_scene_shade(scene, ctypes.byref(sa), Cb, n_cam - 1, stream) != 0


def test_guard_zones_shared_mesh_backward():
    import torch
    from dirt_b200 import _lib
    from dirt_b200._lib import DirtShape, DirtForwardArgs, DirtSharedBackwardArgs
    lib = _lib.load()
    stream = torch.cuda.current_stream().cuda_stream
    sc = scenes.hill(3, 61, 83, 3, n=23, seed=1)
    B, H, W, C = 3, 61, 83, 3
    V, F = sc["vertices"].shape[1], sc["faces"].shape[1]
    shape = DirtShape(B, H, W, C, V, F)
    ws_bytes = lib.dirt_workspace_bytes(ctypes.byref(shape))
    stride = lib.dirt_shared_grad_stride(C)
    npx = B * H * W
    g = np.random.RandomState(0).randn(B, H, W, C).astype(np.float32)
    M = np.random.RandomState(1).randn(B, 4, 4).astype(np.float32)
    bg = np.ascontiguousarray(np.broadcast_to(sc["background"], (B, H, W, C)))
    total = (bg.nbytes + sc["vertices"].nbytes + sc["vertex_colors"].nbytes + sc["faces"].nbytes + g.nbytes + M.nbytes
             + npx * (C * 8 + 20) + V * stride * 4 + ws_bytes + 20 * (GUARD + 256))
    arena = Arena(total)
    d_bg, d_v, d_c, d_f = arena.put(bg), arena.put(sc["vertices"]), arena.put(sc["vertex_colors"]), arena.put(sc["faces"])
    d_g, d_M = arena.put(g), arena.put(M)
    px, fid, gb, ws = arena.carve(npx * C * 4), arena.carve(npx * 4), arena.carve(npx * 16), arena.carve(ws_bytes)
    packed, gbg = arena.carve(V * stride * 4), arena.carve(npx * C * 4)
    packed.fill_(0x7F)                      # "zeroed by the callee"
    cb = 0 if sc["vertex_colors"].shape[0] == 1 else V * C
    fb = 0 if sc["faces"].shape[0] == 1 else F * 3
    fwd = DirtForwardArgs(shape, d_bg.data_ptr(), H * W * C, d_v.data_ptr(), V * 4, d_c.data_ptr(), cb, d_f.data_ptr(), fb,
                          px.data_ptr(), fid.data_ptr(), gb.data_ptr(), ws.data_ptr(), ws_bytes)
    _lib.check(lib.dirt_rasterise_forward(ctypes.byref(fwd), stream))
    bwd = DirtSharedBackwardArgs(shape, d_f.data_ptr(), fb, px.data_ptr(), d_g.data_ptr(), fid.data_ptr(), gb.data_ptr(),
                                 d_M.data_ptr(), packed.data_ptr(), gbg.data_ptr())
    _lib.check(lib.dirt_rasterise_backward_shared(ctypes.byref(bwd), stream))
    torch.cuda.synchronize()
    assert arena.guards_intact(), "the shared-mesh backward wrote outside its buffers"
    p = _f32(packed).view(V, stride)
    assert torch.isfinite(p).all() and p.abs().max() > 0 and (p[:, 4 + C:] == 0).all()
    assert (_f32(fid).view(torch.int32) >= 0).float().mean() > 0.2


def test_guard_zones_host_entry_points():
    """dirt_rasterise_forward_host / _backward_host: HOST buffers carved from one numpy arena."""
    from dirt_b200 import _lib
    from dirt_b200._lib import DirtShape, DirtForwardArgs, DirtBackwardArgs
    lib = _lib.load()
    sc = scenes.random_triangles(3, 35, 51, 3, 157, seed=21, size=0.6)
    B, H, W, C = sc["background"].shape
    V, F = sc["vertices"].shape[1], sc["faces"].shape[1]
    g = np.random.RandomState(0).randn(B, H, W, C).astype(np.float32)
    npx = B * H * W
    sizes = {"pixels": npx * C * 4, "face_id": npx * 4, "gbuffer": npx * 16, "gv": B * V * 16, "gc": B * V * C * 4, "gbg": npx * C * 4}
    ins = {k: np.array(sc[k], copy=True, order="C") for k in ("background", "vertices", "vertex_colors", "faces")}
    ins["g"] = g
    arena = np.full(sum(sizes.values()) + sum(a.nbytes for a in ins.values()) + 16 * (GUARD + 256), SENTINEL, np.uint8)
    base = arena.ctypes.data
    off = (-base) % 256 + GUARD
    spans, ptr = [], {}
    for k, n in list((k, a.nbytes) for k, a in ins.items()) + list(sizes.items()):
        spans.append((off, n))
        ptr[k] = base + off
        if k in ins:
            arena[off:off + n] = ins[k].view(np.uint8).reshape(-1)
        off = (off + n + GUARD + 255) // 256 * 256
    assert off <= arena.size
    shape = DirtShape(B, H, W, C, V, F)
    fa = DirtForwardArgs(shape, ptr["background"], H * W * C, ptr["vertices"], V * 4, ptr["vertex_colors"], V * C, ptr["faces"], F * 3,
                         ptr["pixels"], ptr["face_id"], ptr["gbuffer"], None, 0)
    _lib.check(lib.dirt_rasterise_forward_host(ctypes.byref(fa)))
    ba = DirtBackwardArgs(shape, ptr["faces"], F * 3, ptr["pixels"], ptr["g"], ptr["face_id"], ptr["gbuffer"],
                          ptr["gv"], V * 4, ptr["gc"], V * C, ptr["gbg"])
    _lib.check(lib.dirt_rasterise_backward_host(ctypes.byref(ba)))
    _lib.check(lib.dirt_release_host_staging())
    mask = np.ones(arena.size, bool)
    for o, n in spans:
        mask[o:o + n] = False
    assert (arena[mask] == SENTINEL).all(), "a host entry point wrote outside its buffers"
    for k, a in ins.items():
        o = ptr[k] - base
        assert np.array_equal(arena[o:o + a.nbytes], a.view(np.uint8).reshape(-1)), k + " was modified"
    o = ptr["face_id"] - base
    from oracle import golden
    _, ofid, _ = golden.forward(sc["background"], sc["vertices"], sc["vertex_colors"], sc["faces"])
    assert np.array_equal(arena[o:o + npx * 4].view(np.int32).reshape(B, H, W), ofid)


@pytest.mark.parametrize("poison", [0x00, 0xFF])
def test_guard_zones_of_the_hint_and_of_the_stateless_deterministic_backward(poison):
    """Round-2 buffers: a visibility hint of EXACTLY dirt_visibility_hint_bytes() (whatever it held before: the output
    must not change), and the backward in its stateless + deterministic form on a workspace of exactly
    dirt_backward_workspace_bytes() + dirt_deterministic_workspace_bytes()."""
    import torch
    from dirt_b200 import _lib, rasterise_ops as ops
    from dirt_b200._lib import DirtShape, DirtForwardArgs, DirtBackwardArgs
    lib = _lib.load()
    sc = scenes.sphere(3, 170, 210, n_lat=80, n_lon=60)                 # 9,440 faces per element: the ordered path
    B, H, W, C = sc["background"].shape
    V, F = sc["vertices"].shape[1], sc["faces"].shape[1]
    shape = DirtShape(B, H, W, C, V, F)
    ws_bytes = lib.dirt_workspace_bytes(ctypes.byref(shape))
    hint_bytes = lib.dirt_visibility_hint_bytes(ctypes.byref(shape))
    bws_bytes = lib.dirt_backward_workspace_bytes(ctypes.byref(shape)) + lib.dirt_deterministic_workspace_bytes(ctypes.byref(shape))
    g = np.random.RandomState(0).randn(B, H, W, C).astype(np.float32)
    out_bytes = {"pixels": B * H * W * C * 4, "face_id": B * H * W * 4, "gbuffer": B * H * W * 16,
                 "gv": B * V * 16, "gc": B * V * C * 4, "gbg": B * H * W * C * 4}
    total = (sum(v.nbytes for k, v in sc.items() if k != "expected") + g.nbytes + sum(out_bytes.values()) + ws_bytes + hint_bytes
             + bws_bytes + 20 * (GUARD + 256))
    arena = Arena(total)
    ins = {k: arena.put(sc[k]) for k in ("background", "vertices", "vertex_colors", "faces")}
    ws, hint, bws = arena.carve(ws_bytes), arena.carve(hint_bytes), arena.carve(bws_bytes)
    ws.fill_(poison); hint.fill_(poison); bws.fill_(poison)
    outs = {k: arena.carve(n) for k, n in out_bytes.items()}
    gp = arena.put(g)
    stream = torch.cuda.current_stream().cuda_stream
    t = {k: torch.from_numpy(np.ascontiguousarray(v)).cuda() for k, v in sc.items()}
    ref = ops.forward_raw(t["background"], t["vertices"], t["vertex_colors"], t["faces"])
    for call in range(3):                   # the hint is written by call 1 and read by calls 2 and 3
        fwd = DirtForwardArgs(shape,
                              ins["background"].data_ptr(), H * W * C, ins["vertices"].data_ptr(), V * 4,
                              ins["vertex_colors"].data_ptr(), V * C, ins["faces"].data_ptr(), F * 3,
                              outs["pixels"].data_ptr(), outs["face_id"].data_ptr(), outs["gbuffer"].data_ptr(),
                              ws.data_ptr(), ws_bytes, hint.data_ptr(), hint_bytes)
        _lib.check(lib.dirt_rasterise_forward(ctypes.byref(fwd), stream))
        torch.cuda.synchronize()
        assert arena.guards_intact(), "the forward wrote outside its buffers (call %d)" % call
        assert torch.equal(outs["face_id"].view(torch.int32).view(B, H, W), ref[1])
        assert torch.equal(outs["pixels"].view(torch.float32).view(B, H, W, C), ref[0])
    hdr = hint[:16].view(torch.int32).tolist()
    assert hdr[1:3] == [B, F] and hdr[3] in (2, 3)          # poison 0xFF is no valid header either: first call starts over
    results = []
    for _ in range(2):
        for k in ("gv", "gc", "gbg"):
            outs[k].fill_(0x5A)
        bwd = DirtBackwardArgs(shape, ins["faces"].data_ptr(), F * 3, outs["pixels"].data_ptr(), gp.data_ptr(), None, None,
                               outs["gv"].data_ptr(), V * 4, outs["gc"].data_ptr(), V * C, outs["gbg"].data_ptr(),
                               ins["vertices"].data_ptr(), V * 4, bws.data_ptr(), bws_bytes, _lib.BACKWARD_DETERMINISTIC)
        _lib.check(lib.dirt_rasterise_backward(ctypes.byref(bwd), stream))
        torch.cuda.synchronize()
        assert arena.guards_intact(), "the backward wrote outside its buffers"
        results.append({k: outs[k].clone() for k in ("gv", "gc", "gbg")})
    for k in ("gv", "gc", "gbg"):
        assert torch.equal(results[0][k], results[1][k]), k + ": deterministic mode is not reproducible"
    want = ops.backward_raw(t["faces"], ref[0], torch.from_numpy(g).cuda(), ref[1], ref[2], V)
    for k, wt in (("gv", want[0]), ("gc", want[1])):
        got = results[0][k].view(torch.float32).view(wt.shape)
        assert float((got - wt).abs().max()) <= 1e-4 * float(wt.abs().max())
    for k, tns in ins.items():
        assert np.array_equal(tns.cpu().numpy(), np.ascontiguousarray(sc[k]).view(np.uint8).reshape(-1)), k + " was modified"
