"""dirt_rasterise_forward_noalloc (include/dirt_b200.h): the forward pass with its scratch in a static
device array inside the library -- no cudaMalloc on the call.  Collected last on purpose (the file name
sorts after every other test file): the entry point shares one array per device between all calls.

The GPU test calls the new entry point through the C ABI twice in a row (two different scenes, so the
second call finds the first one's leftovers in the scratch) and compares every output with the existing
entry point's on the same inputs: same kernels, so bit-identical."""
import ctypes

import numpy as np
import pytest

import scenes
from dirt_b200 import _lib


def _shape_args(B, H, W, C, V, F):
    a = _lib.DirtForwardArgs()
    a.shape = _lib.DirtShape(B, H, W, C, V, F)
    return a


def test_limit_is_exported_and_oversized_shapes_are_rejected_with_a_message():
    lib = _lib.load()
    limit = lib.dirt_noalloc_workspace_limit()
    assert limit == 8 << 20
    # the headline batch needs ~400 MB of scratch: refused before anything touches a GPU
    a = _shape_args(64, 512, 512, 3, 49954, 99904)
    assert lib.dirt_workspace_bytes(ctypes.byref(a.shape)) > limit
    assert lib.dirt_rasterise_forward_noalloc(ctypes.byref(a), None) == -4
    msg = lib.dirt_last_error().decode()
    assert "static scratch" in msg and str(limit) in msg
    assert lib.dirt_rasterise_forward_noalloc(None, None) == -1
    bad = _shape_args(1, 0, 8, 3, 3, 1)
    assert lib.dirt_rasterise_forward_noalloc(ctypes.byref(bad), None) == -1


@pytest.mark.gpu
def test_noalloc_forward_matches_the_workspace_entry_point_on_two_calls_in_a_row():
    import torch
    from dirt_b200 import rasterise_ops as ops

    lib = _lib.load()
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    stream = torch.cuda.current_stream(dev).cuda_stream
    # a mesh on the depth-ordered path (>= 4096 faces), then a small-mesh scene with another frame size
    for sc in (scenes.sphere(1, 160, 192, n_lat=56, n_lon=56), scenes.cube(B=2, H=96, W=80), scenes.sphere(2, 128, 128, n_lat=48, n_lon=48)):
        t = {k: torch.from_numpy(np.ascontiguousarray(v)).to(dev) for k, v in sc.items() if k != "expected"}
        bg, v, c, f = t["background"], t["vertices"], t["vertex_colors"], t["faces"]
        B, H, W, C = bg.shape
        V, F = v.shape[1], f.shape[1]
        ref_px, ref_fid, ref_gb = ops.forward_raw(bg, v, c, f)
        px = torch.full((B, H, W, C), float("nan"), device=dev)
        fid = torch.full((B, H, W), -7, dtype=torch.int32, device=dev)
        gb = torch.full((B, H, W, 4), float("nan"), device=dev)
        a = _lib.DirtForwardArgs(
            _lib.DirtShape(B, H, W, C, V, F),
            bg.data_ptr(), H * W * C, v.data_ptr(), V * 4, c.data_ptr(), V * C, f.data_ptr(), F * 3,
            px.data_ptr(), fid.data_ptr(), gb.data_ptr(), None, 0)      # no workspace passed
        assert lib.dirt_workspace_bytes(ctypes.byref(a.shape)) <= lib.dirt_noalloc_workspace_limit()
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated(dev)
        for _ in range(2):
            _lib.check(lib.dirt_rasterise_forward_noalloc(ctypes.byref(a), stream))
            torch.cuda.synchronize()
            assert torch.cuda.memory_allocated(dev) == before
            assert torch.equal(fid, ref_fid)
            assert torch.equal(px, ref_px)
            assert torch.equal(gb, ref_gb)
            px.fill_(float("nan")); fid.fill_(-7); gb.fill_(float("nan"))
