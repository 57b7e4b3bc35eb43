"""The committed fixtures of tests/golden/ (written by tests/golden/make_golden.py).
CPU: the oracle reproduces them bit for bit (kat1/kat2 are the reference's analytic known-answer
cases; the others are outputs of this project's own oracle and pin the written spec against drift -- they are
regression guards, NOT evidence of parity with the reference: attribute values and gradients stay "parity unpinned").  GPU: the CUDA path, through the C ABI,
matches them at the parity bars."""
import os

import numpy as np
import pytest

from oracle import golden

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
ORACLE_CASES = ["cylinder", "cube", "clip", "folded_fan"]


def _load(name):
    with np.load(os.path.join(GOLDEN, name + ".npz")) as z:
        return {k: z[k] for k in z.files}


@pytest.mark.parametrize("name", ["kat1", "kat2"])
def test_oracle_reproduces_reference_known_answers(name):
    d = _load(name)
    px, fid, _ = golden.forward(d["background"], d["vertices"], d["vertex_colors"], d["faces"])
    assert np.array_equal(px, d["expected"])


@pytest.mark.parametrize("name", ORACLE_CASES)
def test_oracle_reproduces_spec_fixtures(name):
    d = _load(name)
    px, fid, gb = golden.forward(d["background"], d["vertices"], d["vertex_colors"], d["faces"])
    assert np.array_equal(fid, d["face_id"])
    assert np.array_equal(px.view(np.int32), d["pixels"].view(np.int32))
    assert np.array_equal(gb.view(np.int32), d["gbuffer"].view(np.int32))
    gv, gc, gbg = golden.backward(d["faces"], px, d["grad_pixels"], fid, gb, d["vertices"].shape[1])
    assert np.array_equal(gbg, d["grad_background"])
    np.testing.assert_allclose(gv, d["grad_vertices"], rtol=0, atol=1e-6 * np.abs(d["grad_vertices"]).max())
    np.testing.assert_allclose(gc, d["grad_vertex_colors"], rtol=0, atol=1e-6 * np.abs(d["grad_vertex_colors"]).max())


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["kat1", "kat2"] + ORACLE_CASES)
def test_cuda_matches_fixtures(name):
    import torch
    from dirt_b200 import rasterise_ops as ops
    d = _load(name)
    t = {k: torch.from_numpy(np.ascontiguousarray(d[k])).cuda() for k in ("background", "vertices", "vertex_colors", "faces")}
    px, fid, gb = ops.forward_raw(t["background"], t["vertices"], t["vertex_colors"], t["faces"])
    if "expected" in d:
        assert np.array_equal(px.cpu().numpy(), d["expected"])        # bit-exact coverage vs the reference's KAT
        return
    assert np.array_equal(fid.cpu().numpy(), d["face_id"])
    np.testing.assert_allclose(px.cpu().numpy(), d["pixels"], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(gb.cpu().numpy(), d["gbuffer"], rtol=1e-5, atol=1e-6)
    gv, gc, gbg = ops.backward_raw(t["faces"], px, torch.from_numpy(d["grad_pixels"]).cuda(), fid, gb, d["vertices"].shape[1])
    assert np.array_equal(gbg.cpu().numpy(), d["grad_background"])
    for a, b in ((gv.cpu().numpy(), d["grad_vertices"]), (gc.cpu().numpy(), d["grad_vertex_colors"])):
        assert np.abs(a - b).max() <= 1e-4 * max(np.abs(b).max(), 1e-12)


SCENE_NAMES = ["oceanic_horizon", "oceanic", "oceanic_still_cloud", "oceanic_no_cloud", "oceanic_simple_proxy", "oceanic_opt_flow", "hill"]


def _scene_inputs(name):
    import test_scene_shaders as T
    H, W = 36, 64
    bg, v, c, f = T.quad(1, H, W, 3, seed=5, noise=0.4)
    if name == "hill":
        return T.terrain_texture(H, W)[None], v, c, f, T.HILL_CAM
    return bg, v, c, f, T.CAM_A


@pytest.mark.parametrize("name", SCENE_NAMES)
def test_oracle_reproduces_scene_fixtures(name):
    """Drift pin of the scene restatements (tests/golden/scenes.npz, written by make_golden.py on
    this image's libm; a libm whose powf / expf round differently may move the last bits)."""
    bg, v, c, f, cam = _scene_inputs(name)
    px = golden.scene(name, bg, v, f, cam)
    np.testing.assert_allclose(px, _load("scenes")[name], rtol=0, atol=2e-6)


@pytest.mark.gpu
@pytest.mark.parametrize("name", SCENE_NAMES)
def test_cuda_matches_scene_fixtures(name):
    import torch
    import dirt_b200
    bg, v, c, f, cam = _scene_inputs(name)
    out = getattr(dirt_b200, name + "_batch")(torch.from_numpy(bg).cuda(), v, c, f, cam).cpu().numpy()
    ref = _load("scenes")[name]
    d = np.abs(out - ref).max(axis=-1)
    bound = 1e-3 if name == "oceanic_horizon" else 1e-4            # horizon keeps float sinf (tests/test_scene_oceanic.py)
    assert (d > bound).mean() <= 0.005 and np.abs(out - ref).mean() <= 1e-5, (name, d.max())
