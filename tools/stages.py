"""Per-stage device times (ms) of fwd+bwd for several workloads (c1: forward only); development aid.
    python tools/stages.py [headline c4 c3 c2 c1 ...]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import scenes  # noqa: E402
from dirt_b200 import rasterise_ops as ops, _lib  # noqa: E402


def scene(name):
    if name == "headline":
        return scenes.sphere(64, 512, 512)
    if name == "h8":
        return scenes.sphere(8, 512, 512)
    if name == "c4":
        return scenes.sphere(256, 256, 256)
    if name == "c3":
        return scenes.cube(B=16, H=512, W=512)
    if name == "c3g":                     # the deferred G-buffer call of samples/deferred.py as ONE 9-channel call
        sc = scenes.cube(B=16, H=512, W=512)
        rng = np.random.RandomState(3)
        sc["vertex_colors"] = rng.uniform(0, 1, sc["vertex_colors"].shape[:2] + (9,)).astype(np.float32)
        sc["background"] = np.zeros(sc["background"].shape[:3] + (9,), np.float32)
        return sc
    if name == "c2":
        return scenes.cube(B=1, H=256, W=256)
    if name == "c1":                      # README.md:46-70, forward only
        sc = scenes.kat1_square()
        sc.pop("expected")
        return sc
    if name == "soup":
        return scenes.random_triangles(16, 512, 512, 3, 2000, seed=1, size=0.3)
    if name == "quad":
        sc = scenes.kat2_square()
        sc.pop("expected")
        sc["vertices"][..., :2] = sc["vertices"][..., :2] * 2 - 1
        return {k: np.repeat(v, 16, 0) for k, v in sc.items()}
    raise SystemExit("unknown workload " + name)


def main():
    lib = _lib.load()
    for name in (sys.argv[1:] or ["headline", "c4", "c3", "c2", "c1"]):
        sc = scene(name)
        d = {k: torch.from_numpy(np.ascontiguousarray(v)).cuda() for k, v in sc.items()}
        B, H, W, C = sc["background"].shape
        V = sc["vertices"].shape[1]
        g = torch.randn(B, H, W, C, device="cuda")

        def step():
            if name == "c1":
                return ops.forward_raw(d["background"], d["vertices"], d["vertex_colors"], d["faces"], want_aux=False)
            px, fid, gb = ops.forward_raw(d["background"], d["vertices"], d["vertex_colors"], d["faces"])
            return ops.backward_raw(d["faces"], px, g, fid, gb, V)

        for _ in range(3):
            step()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        n = 10
        for _ in range(n):
            step()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / n
        lib.dirt_profile_reset(); lib.dirt_profile_enable(1)
        for _ in range(5):
            step(); torch.cuda.synchronize(); lib.dirt_profile_collect()
        lib.dirt_profile_enable(0)
        st = {k: m / 5 for k, (m, c) in _lib.stage_times().items()}      # per step
        graph_ms = None
        if os.environ.get("STAGES_GRAPH", "1") != "0" and ms < 0.5:
            # launch-bound sizes: the same step captured once in a CUDA graph and replayed
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                step()
            for _ in range(3):
                graph.replay()
            torch.cuda.synchronize()
            e0.record()
            for _ in range(50):
                graph.replay()
            e1.record()
            torch.cuda.synchronize()
            graph_ms = e0.elapsed_time(e1) / 50
        print("%-9s B=%d %dx%d F=%d  step %.3f ms  %.0f Mpix/s%s | %s" % (
            name, B, H, W, sc["faces"].shape[1], ms, B * H * W / ms / 1e3,
            "  (CUDA graph replay %.3f ms)" % graph_ms if graph_ms else "",
            " ".join("%s=%.3f" % (k, v) for k, v in st.items())), flush=True)


if __name__ == "__main__":
    main()
