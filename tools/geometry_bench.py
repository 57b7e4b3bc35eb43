"""Times dirt_b200.lighting.vertex_normals (forward + backward) on the c5 mesh (1M-face height
field, B frames) against a plain torch index_add formulation on the same GPU and the numpy
oracle on the host.   python tools/geometry_bench.py [B]"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import scenes  # noqa: E402
from dirt_b200 import lighting  # noqa: E402
from oracle import lighting as ol  # noqa: E402


def torch_formulation(v, f):
    p = v[..., :3]
    idx = f.long()
    m = torch.cross(p[:, idx[:, 1]] - p[:, idx[:, 0]], p[:, idx[:, 2]] - p[:, idx[:, 0]], dim=-1)
    n = m / (m.norm(dim=-1, keepdim=True) + 1.e-12)
    s = torch.zeros_like(p)
    for k in range(3):
        s = s.index_add(1, idx[:, k], n)
    return s / (s.norm(dim=-1, keepdim=True) + 1.e-12)


def timed(fn, reps=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 4
    sc = scenes.hill(B, 64, 64, 3)
    v = torch.tensor(sc["vertices"], device="cuda")
    f = torch.tensor(sc["faces"][0], device="cuda")
    g = torch.randn(B, v.shape[1], 3, device="cuda")

    def ours():
        x = v.clone().requires_grad_(True)
        lighting.vertex_normals(x, f).backward(g)

    def plain():
        x = v.clone().requires_grad_(True)
        torch_formulation(x, f).backward(g)

    t0 = time.perf_counter()
    ol.vertex_normals(sc["vertices"][:1], sc["faces"][0])
    cpu_s = time.perf_counter() - t0
    out = {"workload": "vertex_normals fwd+bwd, B=%d, V=%d, F=%d" % (B, v.shape[1], f.shape[0]),
           "dirt_b200_ms": timed(ours), "torch_index_add_ms": timed(plain),
           "numpy_oracle_forward_only_one_frame_s": cpu_s}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
