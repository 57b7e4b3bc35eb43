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
        t.copy_(torch.from_numpy(np.ascontiguousarray(array)).view(-1).view(torch.uint8))
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
