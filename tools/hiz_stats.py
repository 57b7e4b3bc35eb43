"""Development aid: one forward of a workload with the DIRT_STATS build (prints cull counts per bucket)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tools")]
import torch
import stages
from dirt_b200 import rasterise_ops as ops
sc = stages.scene(sys.argv[1] if len(sys.argv) > 1 else "headline")
import numpy as np
d = {k: torch.from_numpy(np.ascontiguousarray(v)).cuda() for k, v in sc.items()}
ops.forward_raw(d["background"], d["vertices"], d["vertex_colors"], d["faces"])
torch.cuda.synchronize()
