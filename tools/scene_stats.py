"""Prints CUDA-vs-oracle difference statistics for every scene op (run on a GPU box)."""
import os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import dirt_b200
from oracle import golden
import test_scene_shaders as T

B, H, W = 1, 72, 128
for name in T.OCEAN_SCENES + ["oceanic_horizon", "hill"]:
    for cam in (T.CAM_A, T.CAM_B):
        if name == "hill":
            _, v, c, f = T.quad(B, H, W, 3)
            bg = T.terrain_texture(H, W)[None]
            cam = T.HILL_CAM
        else:
            bg, v, c, f = T.quad(B, H, W, 3, seed=1, noise=0.5)
        ref = golden.scene(name, bg, v, f, cam)
        out = getattr(dirt_b200, name + "_batch")(torch.from_numpy(bg).cuda(), v, c, f, cam).cpu().numpy()
        d = np.abs(out - ref).max(axis=-1)
        print("%-22s mean %.2e  p99 %.2e  max %.2e  >1e-3 %.4f  >2e-2 %.4f  exact %.3f" % (
            name, np.abs(out - ref).mean(), np.quantile(d, 0.99), d.max(), (d > 1e-3).mean(), (d > 2e-2).mean(), (d == 0).mean()))
