import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import scenes
from oracle import golden
from dirt_b200 import rasterise_ops as ops
for name, sc in (("cube", scenes.cube(B=1, H=256, W=256)), ("sphere", scenes.sphere(1, 128, 128, n_lat=48, n_lon=48))):
    t = {k: torch.from_numpy(np.ascontiguousarray(v)).cuda() for k, v in sc.items()}
    px, fid, gb = ops.forward_raw(t["background"], t["vertices"], t["vertex_colors"], t["faces"])
    opx, ofid, ogb = golden.forward(sc["background"], sc["vertices"], sc["vertex_colors"], sc["faces"])
    gbn = gb.cpu().numpy()
    for ch, nm in enumerate(("B0", "B1", "wc", "zw")):
        d = gbn[..., ch].view(np.int32) != ogb[..., ch].view(np.int32)
        print(name, nm, "bit mismatches", d.sum(), "of", d.size)
        if d.sum():
            idx = np.argwhere(d)[:5]
            for i in idx:
                print("  ", tuple(i), repr(gbn[tuple(i)][ch]), repr(ogb[tuple(i)][ch]), "face", ofid[tuple(i)])
