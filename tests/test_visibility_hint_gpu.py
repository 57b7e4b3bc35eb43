"""DirtForwardArgs.visibility_hint: the persistent record of which faces won a pixel in the previous call
only ORDERS the next call's work (hinted faces first, everything else behind the depth pyramid they leave,
face by face and then in chunks of eight scan lines).  Whatever the hint holds, the output is bit-identical."""
import ctypes

import numpy as np
import pytest

import scenes
from dirt_b200 import _lib

pytestmark = pytest.mark.gpu

HDR = 256      # bytes of header in front of the flags


def _frames(B, H, W, n, frames, deg=0.5, noise_px=0.0):
    import torch
    sc = scenes.sphere_frames(B, H, W, 3, n, n, frame=0)
    t = {k: torch.from_numpy(np.ascontiguousarray(v)).cuda() for k, v in sc.items()}
    vs = [torch.from_numpy(scenes.sphere_frames(B, H, W, 3, n, n, frame=k, step_deg=deg, noise_px=noise_px, vertices_only=True)).cuda()
          for k in range(frames)]
    return sc, t, vs


def _render(t, v, hint):
    from dirt_b200 import rasterise_ops as ops
    return ops.forward_raw(t["background"], v, t["vertex_colors"], t["faces"], hint=hint)


def _same(a, b):
    import torch
    return all(torch.equal(x, y) for x, y in zip(a, b))


def test_hint_never_changes_the_output_and_tracks_the_visible_set():
    import torch
    from dirt_b200 import rasterise_ops as ops
    from oracle import golden
    B, H, W, n = 3, 192, 160, 72                     # F = 10,224 per element: the depth-ordered path
    sc, t, vs = _frames(B, H, W, n, 5, noise_px=0.3)
    F = sc["faces"].shape[1]
    assert F >= ops.HINT_MIN_FACES
    lib = _lib.load()
    nbytes = lib.dirt_visibility_hint_bytes(ctypes.byref(_lib.DirtShape(B, H, W, 3, sc["vertices"].shape[1], F)))
    assert nbytes >= HDR + B * F
    hint = torch.zeros(nbytes, dtype=torch.uint8, device="cuda")
    for k, v in enumerate(vs):
        ref = _render(t, v, None)
        got = _render(t, v, hint)
        assert _same(ref, got), "frame %d" % k
        hdr = hint[:16].view(torch.int32).cpu().numpy()
        assert hdr[1] == B and hdr[2] == F and hdr[3] == k + 1          # generation advances once per call
        flagged = (hint[HDR:HDR + B * F] == k + 1).view(B, F)
        fid = ref[1]
        for b in range(B):
            vis = torch.unique(fid[b][fid[b] >= 0])
            assert torch.equal(torch.nonzero(flagged[b]).flatten(), vis.to(torch.int64))
    # and the reference of it all is the oracle
    opx, ofid, ogb = golden.forward(sc["background"], vs[-1].cpu().numpy(), sc["vertex_colors"], sc["faces"], threads=4)
    assert np.array_equal(ref[1].cpu().numpy(), ofid)
    np.testing.assert_allclose(ref[0].cpu().numpy(), opx, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("kind", ["empty", "garbage", "all_faces", "no_faces", "stale", "other_shape"])
def test_wrong_hints_are_harmless(kind):
    import torch
    B, H, W, n = 2, 128, 224, 64                     # F = 8,064
    sc, t, vs = _frames(B, H, W, n, 2, deg=40.0)     # frame 1 is a very different view
    F = sc["faces"].shape[1]
    lib = _lib.load()
    nbytes = lib.dirt_visibility_hint_bytes(ctypes.byref(_lib.DirtShape(B, H, W, 3, sc["vertices"].shape[1], F)))
    hint = torch.zeros(nbytes, dtype=torch.uint8, device="cuda")
    header = torch.tensor([0x54524944, B, F, 7], dtype=torch.int32, device="cuda").view(torch.uint8)
    g = torch.Generator(device="cuda").manual_seed(3)
    if kind == "garbage":
        hint.copy_(torch.randint(0, 256, (nbytes,), dtype=torch.uint8, device="cuda", generator=g))
    elif kind == "all_faces":
        hint[:16] = header
        hint[HDR:] = 7
    elif kind == "no_faces":
        hint[:16] = header
    elif kind == "stale":
        _render(t, vs[1], hint)                      # the visible set of another view
    elif kind == "other_shape":
        hint[:16] = torch.tensor([0x54524944, B + 1, F - 3, 9], dtype=torch.int32, device="cuda").view(torch.uint8)
        hint[HDR:] = 9
    ref = _render(t, vs[0], None)
    for _ in range(2):
        assert _same(ref, _render(t, vs[0], hint))


def test_hint_buffer_is_validated():
    import torch
    from dirt_b200 import rasterise_ops as ops
    sc, t, vs = _frames(1, 64, 64, 48, 1)
    small = torch.zeros(300, dtype=torch.uint8, device="cuda")
    with pytest.raises(_lib.InvalidArgumentError):
        _render(t, vs[0], small)
    # meshes below the ordering threshold neither read nor write the hint
    cube = {k: torch.from_numpy(np.ascontiguousarray(v)).cuda() for k, v in scenes.cube(B=1, H=64, W=64).items()}
    h = torch.full((4096,), 5, dtype=torch.uint8, device="cuda")
    ops.forward_raw(cube["background"], cube["vertices"], cube["vertex_colors"], cube["faces"], hint=h)
    assert bool((h == 5).all())


def test_autograd_path_keeps_a_hint_per_mesh_and_matches_the_unhinted_result(monkeypatch):
    import torch
    import dirt_b200
    from dirt_b200 import rasterise_ops as ops
    B, H, W, n = 2, 128, 128, 64
    sc, t, vs = _frames(B, H, W, n, 3)
    ops.release_workspaces()
    out = dirt_b200.rasterise_batch(t["background"], vs[0], t["vertex_colors"], t["faces"])
    assert len(ops._hints) == 0                      # opt-in: DIRT_B200_HINT=1
    monkeypatch.setenv("DIRT_B200_HINT", "1")
    for v in vs:
        ref = _render(t, v, None)[0]
        out = dirt_b200.rasterise_batch(t["background"], v, t["vertex_colors"], t["faces"])
        assert torch.equal(out, ref)
    assert len(ops._hints) == 1
