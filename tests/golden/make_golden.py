"""Writes the committed golden fixtures tests/golden/*.npz.

    python tests/golden/make_golden.py

Two kinds of fixture:
  kat1.npz, kat2.npz    ANALYTIC expectations of the reference's own known-answer cases
                        (README.md:46-70 / README.md:104, tests/multi_gpu_test.py:8-17): inputs
                        and the expected image, written down from the reference text -- these
                        pin the oracle AND the CUDA path to the reference.
  cylinder.npz, cube.npz, clip.npz, folded_fan.npz
                        outputs of the oracle (oracle/golden_raster.c) on the reference's
                        deterministic scenes that store no expectation (tests/rasterise_tests.py:
                        11-89, samples/simple.py:15-74), on a near-plane-crossing scene and on three faces whose
                        clipped fans fold over themselves at the guard band (tests/scenes.py).
                        The reference cannot run here (no TensorFlow, no GL), so these are
                        regression pins of the written spec (docs/RASTER_SPEC.md,
                        docs/GRADIENT_SPEC.md), not reference outputs; parity of attribute values
                        and gradients with the reference stays "unpinned".
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import scenes  # noqa: E402
from oracle import golden  # noqa: E402


def oracle_case(sc, seed):
    px, fid, gb = golden.forward(sc["background"], sc["vertices"], sc["vertex_colors"], sc["faces"])
    g = np.random.RandomState(seed).randn(*px.shape).astype(np.float32)
    gv, gc, gbg = golden.backward(sc["faces"], px, g, fid, gb, sc["vertices"].shape[1])
    out = {k: v for k, v in sc.items() if k != "expected"}
    out.update(pixels=px, face_id=fid, gbuffer=gb, grad_pixels=g, grad_vertices=gv, grad_vertex_colors=gc,
               grad_background=gbg)
    return out


def scene_cases():
    """One small render of every scene op by the oracle (oracle/scene_oceanic.c, scene_shaders.cpp):
    inputs are rebuilt from tests/test_scene_shaders.py, only the images are stored."""
    import test_scene_shaders as T
    H, W = 36, 64
    out = {}
    bg, v, c, f = T.quad(1, H, W, 3, seed=5, noise=0.4)
    for name in ["oceanic_horizon"] + T.OCEAN_SCENES:
        out[name] = golden.scene(name, bg, v, f, T.CAM_A)
    out["hill"] = golden.scene("hill", T.terrain_texture(H, W)[None], v, f, T.HILL_CAM)
    return out


def main():
    for name, sc in (("kat1", scenes.kat1_square()), ("kat2", scenes.kat2_square())):
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **sc)
    np.savez_compressed(os.path.join(HERE, "cylinder.npz"), **oracle_case(scenes.cylinder(), 0))
    np.savez_compressed(os.path.join(HERE, "cube.npz"), **oracle_case(scenes.cube(B=2, H=64, W=64), 1))
    np.savez_compressed(os.path.join(HERE, "clip.npz"), **oracle_case(scenes.near_crossing(B=1, H=48, W=64), 2))
    np.savez_compressed(os.path.join(HERE, "folded_fan.npz"), **oracle_case(scenes.folded_fans(), 3))
    np.savez_compressed(os.path.join(HERE, "scenes.npz"), **scene_cases())
    for f in sorted(os.listdir(HERE)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(HERE, f)), "bytes")


if __name__ == "__main__":
    main()
