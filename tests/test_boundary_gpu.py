"""The drop-in boundary: the backward in its stateless form -- no face_id /
gbuffer carried from the forward, visibility re-derived from the vertices with the forward's kernels, as the
reference's op pair would see it (RasteriseGrad receives the forward's inputs, csrc/rasterise_grad_egl.cpp:38-42) --
gradient outputs with padded batch strides, and dirt_workspace_create / destroy."""
import ctypes

import numpy as np
import pytest

import scenes
from dirt_b200 import _lib
from oracle import golden

pytestmark = pytest.mark.gpu
GRAD_RTOL = 1e-4


def _cuda(sc):
    import torch
    return {k: torch.from_numpy(np.ascontiguousarray(v)).cuda() for k, v in sc.items() if k != "expected"}


@pytest.mark.parametrize("scene", ["cube", "sphere", "near_crossing", "shared"])
def test_stateless_backward_matches_the_oracle_and_the_stateful_call(scene):
    import torch
    from dirt_b200 import rasterise_ops as ops
    sc = {"cube": lambda: scenes.cube(B=2, H=72, W=56),
          "sphere": lambda: scenes.sphere(2, 160, 128, n_lat=64, n_lon=64),          # depth-ordered path
          "near_crossing": lambda: scenes.near_crossing(n=24),                          # faces through the clipper
          "shared": lambda: scenes.cube(B=3, H=48, W=64)}[scene]()
    t = _cuda(sc)
    if scene == "shared":                      # one mesh, per-element cameras would be per-element vertices; here: all shared
        t = {k: (v[:1] if k in ("vertices", "faces", "vertex_colors") else v) for k, v in t.items()}
        sc = {k: (v[:1] if k in ("vertices", "faces", "vertex_colors") else v) for k, v in sc.items()}
    px, fid, gb = ops.forward_raw(t["background"], t["vertices"], t["vertex_colors"], t["faces"])
    g = torch.from_numpy(np.random.RandomState(1).randn(*px.shape).astype(np.float32)).cuda()
    V = sc["vertices"].shape[1]
    sv = scene == "shared"
    a = ops.backward_raw(t["faces"], px, g, fid, gb, V, shared_vertices=sv, shared_colors=sv)
    b = ops.backward_raw(t["faces"], px, g, None, None, V, shared_vertices=sv, shared_colors=sv, vertices=t["vertices"])
    assert torch.equal(a[2], b[2])
    for x, y in zip(a[:2], b[:2]):
        assert float((x - y).abs().max()) <= GRAD_RTOL * max(float(x.abs().max()), 1e-12)
    opx, ofid, ogb = golden.forward(sc["background"], sc["vertices"], sc["vertex_colors"], sc["faces"])
    ogv, ogc, ogbg = golden.backward(sc["faces"], opx, g.cpu().numpy(), ofid, ogb, V, shared_vertices=sv, shared_colors=sv)
    assert np.array_equal(b[2].cpu().numpy(), ogbg)
    for x, y in ((b[0].cpu().numpy(), ogv), (b[1].cpu().numpy(), ogc)):
        assert np.abs(x.reshape(y.shape) - y).max() <= GRAD_RTOL * max(np.abs(y).max(), 1e-12)


def test_stateless_backward_needs_vertices_and_a_workspace():
    import torch
    lib = _lib.load()
    sc = scenes.cube(B=1, H=32, W=32)
    t = _cuda(sc)
    B, H, W, C = sc["background"].shape
    V, F = sc["vertices"].shape[1], sc["faces"].shape[1]
    shape = _lib.DirtShape(B, H, W, C, V, F)
    px = torch.zeros(B, H, W, C, device="cuda")
    gv = torch.zeros(B, V, 4, device="cuda"); gc = torch.zeros(B, V, C, device="cuda")
    need = lib.dirt_backward_workspace_bytes(ctypes.byref(shape))
    assert need > lib.dirt_workspace_bytes(ctypes.byref(shape)) + B * H * W * 20 - 1
    ws = torch.empty(need, dtype=torch.uint8, device="cuda")

    def call(fid, gb, vertices, wsp, wsn):
        a = _lib.DirtBackwardArgs(shape, t["faces"].data_ptr(), F * 3, px.data_ptr(), px.data_ptr(), fid, gb,
                                  gv.data_ptr(), V * 4, gc.data_ptr(), V * C, None, vertices, V * 4, wsp, wsn)
        return lib.dirt_rasterise_backward(ctypes.byref(a), None)

    assert call(None, None, t["vertices"].data_ptr(), ws.data_ptr(), need - 1) == -3           # workspace too small
    assert call(None, None, None, ws.data_ptr(), need) == -1 and b"vertices" in lib.dirt_last_error()
    assert call(px.data_ptr(), None, t["vertices"].data_ptr(), ws.data_ptr(), need) == -1      # one of the pair only
    assert call(None, None, t["vertices"].data_ptr(), ws.data_ptr(), need) == 0
    torch.cuda.synchronize()


def test_gradient_outputs_with_padded_batch_strides_are_cleared_per_element():
    """a stride larger than the dense block left the blocks of elements b >= 1 partly uncleared."""
    import torch
    from dirt_b200 import rasterise_ops as ops
    lib = _lib.load()
    sc = scenes.cube(B=3, H=40, W=48)
    t = _cuda(sc)
    B, H, W, C = sc["background"].shape
    V, F = sc["vertices"].shape[1], sc["faces"].shape[1]
    px, fid, gb = ops.forward_raw(t["background"], t["vertices"], t["vertex_colors"], t["faces"])
    g = torch.randn_like(px)
    ref = ops.backward_raw(t["faces"], px, g, fid, gb, V)
    pad_v, pad_c = V * 4 + 20, V * C + 7                       # strides in elements; 16-byte alignment kept for grad_vertices
    gv = torch.full((B, pad_v), 123.0, device="cuda"); gc = torch.full((B, pad_c), 123.0, device="cuda")
    a = _lib.DirtBackwardArgs(_lib.DirtShape(B, H, W, C, V, F), t["faces"].data_ptr(), F * 3, px.data_ptr(), g.data_ptr(),
                              fid.data_ptr(), gb.data_ptr(), gv.data_ptr(), pad_v, gc.data_ptr(), pad_c, None, None, 0, None, 0)
    _lib.check(lib.dirt_rasterise_backward(ctypes.byref(a), torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    assert bool((gv[:, V * 4:] == 123.0).all()) and bool((gc[:, V * C:] == 123.0).all())          # padding untouched
    for got, want in ((gv[:, :V * 4].reshape(B, V, 4), ref[0]), (gc[:, :V * C].reshape(B, V, C), ref[1])):
        assert float((got - want).abs().max()) <= GRAD_RTOL * float(want.abs().max())


def test_workspace_create_and_destroy():
    import torch
    lib = _lib.load()
    sc = scenes.cube(B=1, H=32, W=32)
    t = _cuda(sc)
    B, H, W, C = sc["background"].shape
    V, F = sc["vertices"].shape[1], sc["faces"].shape[1]
    shape = _lib.DirtShape(B, H, W, C, V, F)
    ws, n = ctypes.c_void_p(), ctypes.c_size_t()
    _lib.check(lib.dirt_workspace_create(ctypes.byref(shape), ctypes.byref(ws), ctypes.byref(n)))
    assert ws.value and n.value >= lib.dirt_backward_workspace_bytes(ctypes.byref(shape))
    px = torch.empty(B, H, W, C, device="cuda")
    a = _lib.DirtForwardArgs(shape, t["background"].data_ptr(), H * W * C, t["vertices"].data_ptr(), V * 4,
                             t["vertex_colors"].data_ptr(), V * C, t["faces"].data_ptr(), F * 3, px.data_ptr(), None, None,
                             ws, n.value, None, 0)
    _lib.check(lib.dirt_rasterise_forward(ctypes.byref(a), torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    opx, _, _ = golden.forward(sc["background"], sc["vertices"], sc["vertex_colors"], sc["faces"])
    np.testing.assert_allclose(px.cpu().numpy(), opx, rtol=1e-5, atol=1e-6)
    _lib.check(lib.dirt_workspace_destroy(ws))
    assert lib.dirt_workspace_create(ctypes.byref(_lib.DirtShape(1, 0, 4, 3, 1, 1)), ctypes.byref(ws), None) == -1


@pytest.mark.parametrize("scene", ["sphere", "cube9", "shared_strides"])
def test_deterministic_backward_is_bit_identical_run_to_run_and_matches_the_oracle(scene):
    """the default backward sums with fp32 atomics (order varies run to run); with
    DIRT_BACKWARD_DETERMINISTIC every sum is an integer sum, so repeated calls agree bit for bit."""
    import torch
    from dirt_b200 import rasterise_ops as ops
    if scene == "sphere":
        sc = scenes.sphere(3, 192, 160, n_lat=72, n_lon=72)
    elif scene == "cube9":
        sc = scenes.cube(B=2, H=96, W=96)
        rng = np.random.RandomState(4)
        sc["vertex_colors"] = rng.uniform(0, 1, sc["vertex_colors"].shape[:2] + (9,)).astype(np.float32)
        sc["background"] = rng.uniform(0, 1, sc["background"].shape[:3] + (9,)).astype(np.float32)
    else:
        sc = scenes.cube(B=3, H=64, W=80)
        sc = {k: (v[:1] if k in ("vertices", "faces", "vertex_colors") else v) for k, v in sc.items()}
    shared = scene == "shared_strides"
    t = _cuda(sc)
    px, fid, gb = ops.forward_raw(t["background"], t["vertices"], t["vertex_colors"], t["faces"])
    g = torch.from_numpy(np.random.RandomState(2).randn(*px.shape).astype(np.float32)).cuda()
    V = sc["vertices"].shape[1]
    runs = [ops.backward_raw(t["faces"], px, g, fid, gb, V, shared_vertices=shared, shared_colors=shared, deterministic=True) for _ in range(4)]
    # other work in between changes the arrival order of the warps, not the result
    for r in runs[1:]:
        assert torch.equal(r[0], runs[0][0]) and torch.equal(r[1], runs[0][1]) and torch.equal(r[2], runs[0][2])
    stateless = ops.backward_raw(t["faces"], px, g, None, None, V, shared_vertices=shared, shared_colors=shared,
                                 vertices=t["vertices"], deterministic=True)
    assert torch.equal(stateless[0], runs[0][0]) and torch.equal(stateless[1], runs[0][1])
    opx, ofid, ogb = golden.forward(sc["background"], sc["vertices"], sc["vertex_colors"], sc["faces"])
    ogv, ogc, ogbg = golden.backward(sc["faces"], opx, g.cpu().numpy(), ofid, ogb, V, shared_vertices=shared, shared_colors=shared)
    for x, y in ((runs[0][0].cpu().numpy(), ogv), (runs[0][1].cpu().numpy(), ogc)):
        assert np.abs(x.reshape(y.shape) - y).max() <= GRAD_RTOL * max(np.abs(y).max(), 1e-12)
    assert np.array_equal(runs[0][2].cpu().numpy(), ogbg)


def test_deterministic_shared_mesh_backward():
    import torch
    from dirt_b200 import rasterise_ops as ops
    sc = scenes.hill(3, 96, 128, 3, n=40, seed=1)
    t = _cuda(sc)
    px, fid, gb = ops.forward_raw(t["background"], t["vertices"], t["vertex_colors"], t["faces"])
    g = torch.randn_like(px)
    V = sc["vertices"].shape[1]
    ref, _ = ops.backward_shared_raw(t["faces"], px, g, fid, gb, V, t["matrices"])
    runs = [ops.backward_shared_raw(t["faces"], px, g, fid, gb, V, t["matrices"], deterministic=True)[0] for _ in range(3)]
    assert torch.equal(runs[0], runs[1]) and torch.equal(runs[0], runs[2])
    assert float((runs[0] - ref).abs().max()) <= GRAD_RTOL * float(ref.abs().max())


def test_peer_allreduce_protocol_between_two_streams_of_one_gpu(monkeypatch):
    """dirt_allreduce_peer (csrc/peer_allreduce.cu) needs every rank's buffer and flag pad mapped into every process; on one
    GPU two "ranks" are two streams of this process that address each other's buffers directly -- the same kernels, the
    same handshakes.  (Between real GPUs: tools/peer_allreduce_check.py under torchrun, profiles/r2_peer_allreduce_n*.json.)"""
    import torch
    lib = _lib.load()
    monkeypatch.setenv("DIRT_B200_PEER_TIMEOUT_MS", "3000")     # should the two streams be run one after the other: give up, skip
    n = 4 * 50021                                     # not a multiple of the slice size
    world = 2
    streams = [torch.cuda.Stream() for _ in range(world)]
    for it in range(3):
        src = [torch.randn(n, device="cuda") for _ in range(world)]
        bufs = [s.clone() for s in src]
        if it == 0:
            pads = [torch.zeros(2 * world, dtype=torch.int32, device="cuda") for _ in range(world)]
            epochs = [torch.zeros(4, dtype=torch.int32, device="cuda") for _ in range(world)]
        bt = torch.tensor([b.data_ptr() for b in bufs], dtype=torch.int64, device="cuda")
        pt = torch.tensor([p.data_ptr() for p in pads], dtype=torch.int64, device="cuda")
        torch.cuda.synchronize()
        for r in range(world):
            a = _lib.DirtPeerAllreduceArgs(bt.data_ptr(), pt.data_ptr(), r, world, n, epochs[r].data_ptr(), None)
            _lib.check(lib.dirt_allreduce_peer(ctypes.byref(a), streams[r].cuda_stream))
        torch.cuda.synchronize()
        if any(e.tolist()[1] != 0 for e in epochs):
            pytest.skip("the two streams' kernels did not run concurrently here (a handshake timed out)")
        want = src[0] + src[1]
        for r in range(world):
            assert torch.equal(bufs[r], want)
            assert epochs[r].tolist()[:2] == [it + 1, 0]
    # argument checks need no second rank
    a = _lib.DirtPeerAllreduceArgs(bt.data_ptr(), pt.data_ptr(), 2, 2, n, epochs[0].data_ptr(), None)
    assert lib.dirt_allreduce_peer(ctypes.byref(a), None) == -1
    a = _lib.DirtPeerAllreduceArgs(bt.data_ptr(), pt.data_ptr(), 0, 2, n + 1, epochs[0].data_ptr(), None)
    assert lib.dirt_allreduce_peer(ctypes.byref(a), None) == -1
    a = _lib.DirtPeerAllreduceArgs(None, None, 0, 1, n, None, None)           # a world of one: nothing to do
    assert lib.dirt_allreduce_peer(ctypes.byref(a), None) == 0
