"""development aid: minimal repro search for a fuzz mismatch"""
import sys, os
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import test_fuzz_gpu as t
from oracle import golden
from dirt_b200 import rasterise_ops as ops

def run(sc):
    tt = {k: torch.from_numpy(np.ascontiguousarray(v)).cuda() for k, v in sc.items()}
    px, fid, gb = ops.forward_raw(tt["background"], tt["vertices"], tt["vertex_colors"], tt["faces"])
    opx, ofid, ogb = golden.forward(sc["background"], sc["vertices"], sc["vertex_colors"], sc["faces"], threads=4)
    return px.cpu().numpy(), fid.cpu().numpy(), gb.cpu().numpy(), opx, ofid, ogb

for seed in [int(a) for a in sys.argv[1:]] or [18]:
    sc, B = t._case(seed)
    px, fid, gb, opx, ofid, ogb = run(sc)
    bad = np.abs(px - opx).max(-1) > 1e-5 * np.abs(opx).max(-1) + 1e-6
    print("seed", seed, "bad px", bad.sum(), "fid equal", np.array_equal(fid, ofid), "gb maxdiff", np.abs(gb - ogb).max(0).max(0).max(0))
    faces_bad = np.unique(fid[bad])
    print("faces involved", faces_bad[:20], "count", len(faces_bad))
    shown = 0
    for b in range(B):
        for f in np.unique(fid[b][bad[b]]):
            if shown >= 4: break
            F = sc["faces"][b if sc["faces"].shape[0] > 1 else 0][f]
            V = sc["vertices"][b if sc["vertices"].shape[0] > 1 else 0][F]
            n_bad = (bad[b] & (fid[b] == f)).sum(); n_all = (fid[b] == f).sum()
            print(" b", b, "face", f, "idx", F, "bad", n_bad, "of", n_all)
            print("  verts", V.tolist())
            # single-face repro
            one = {"background": np.zeros((1,) + sc["background"].shape[1:], np.float32), "vertices": V[None].copy(),
                   "vertex_colors": np.eye(3, sc["background"].shape[-1], dtype=np.float32)[None].copy(), "faces": np.array([[[0, 1, 2]]], np.int32)}
            p1, f1, g1, o1, of1, og1 = run(one)
            bad1 = np.abs(g1 - og1).max(-1) > 1e-5
            print("  alone: fid equal", np.array_equal(f1, of1), "covered", (of1 >= 0).sum(), "gb bad", bad1.sum(), "maxdiff", np.abs(g1 - og1).max(0).max(0).max(0))
            if bad1.any():
                yx = np.argwhere(bad1[0])[0]
                print("  first bad px", yx, "gpu", g1[0, yx[0], yx[1]], "oracle", og1[0, yx[0], yx[1]])
            shown += 1
