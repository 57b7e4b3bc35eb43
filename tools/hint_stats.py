"""Cull statistics of the setup launches (build with -DDIRT_STATS and select it with DIRT_B200_LIB): frame 0
without a usable hint (depth buckets), frame 1 with the hint frame 0 left.  Development aid.
    DIRT_B200_LIB=$PWD/dirt_b200/libvar_stats.so python tools/hint_stats.py [B=8] [H=512] [deg=0.25]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import scenes  # noqa: E402
from dirt_b200 import rasterise_ops as ops  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
H = W = int(sys.argv[2]) if len(sys.argv) > 2 else 512
deg = float(sys.argv[3]) if len(sys.argv) > 3 else 0.25
sc = scenes.sphere(B, H, W)
d = {k: torch.from_numpy(np.ascontiguousarray(v)).cuda() for k, v in sc.items()}
v1 = torch.from_numpy(scenes.sphere_frames(B, H, W, frame=1, step_deg=deg, vertices_only=True)).cuda()
hint = ops.visibility_hint(torch.device("cuda", 0), B, sc["faces"].shape[1], key="stats")
print("== frame 0, B=%d, no usable hint (depth buckets)" % B, flush=True)
ops.forward_raw(d["background"], d["vertices"], d["vertex_colors"], d["faces"], hint=hint)
torch.cuda.synchronize()
print("== frame 1, hinted", flush=True)
ops.forward_raw(d["background"], v1, d["vertex_colors"], d["faces"], hint=hint)
torch.cuda.synchronize()
