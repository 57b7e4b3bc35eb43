"""Where a c5 optimisation step spends its time when run as the reference's caller runs it
(tests/optimize_sequence.py:293-352: normals -> diffuse_directional -> project -> rasterise -> loss -> backward):
decides whether the shading helpers (torch composites in dirt_b200/lighting.py) deserve a fused kernel
(only if they exceed 15 % of the step).  One JSON line.
    python tools/c5_step_breakdown.py [B=4] [H=1024] [n=708]"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import scenes  # noqa: E402
import dirt_b200  # noqa: E402
from dirt_b200 import lighting  # noqa: E402


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 4
    H = W = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
    n = int(sys.argv[3]) if len(sys.argv) > 3 else 708
    sc = scenes.hill(B, H, W, 3, n=n, seed=0)
    dev = torch.device("cuda", 0)
    world = torch.from_numpy(sc["world"]).to(dev).requires_grad_(True)
    albedo = torch.from_numpy(sc["vertex_colors"][0]).to(dev).requires_grad_(True)
    faces = torch.from_numpy(sc["faces"][0]).to(dev)
    M = torch.from_numpy(sc["matrices"]).to(dev)
    bg = torch.from_numpy(sc["background"]).to(dev)
    target = torch.rand(B, H, W, 3, device=dev)
    light_dir = torch.tensor([0.3, -0.8, -0.5], device=dev)
    light_dir = light_dir / light_dir.norm()
    light_col = torch.tensor([1.0, 0.95, 0.9], device=dev)

    def shade():
        normals = lighting.vertex_normals(world[:, :3], faces)
        return lighting.diffuse_directional(normals, albedo, light_dir, light_col) + 0.2 * albedo

    def full_step(with_shading=True):
        world.grad = None; albedo.grad = None
        colours = shade() if with_shading else albedo
        px = dirt_b200.rasterise_batch_shared(bg, world, M, colours, faces)
        loss = 0.5 * ((px - target) ** 2).mean()
        loss.backward()
        return loss

    def timed(fn, reps=10):
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

    def shade_fwd_bwd():
        world.grad = None; albedo.grad = None
        c = shade()
        c.backward(torch.ones_like(c))

    def loss_only():
        px = target.detach().clone().requires_grad_(True)
        (0.5 * ((px - target) ** 2).mean()).backward()

    out = {"workload": "c5: B=%d frames %dx%d, hill n=%d (V=%d, F=%d)" % (B, H, W, n, world.shape[0], faces.shape[0]),
           "step_with_shading_ms": timed(lambda: full_step(True)),
           "step_without_shading_ms": timed(lambda: full_step(False)),
           "shading_fwd_bwd_alone_ms": timed(shade_fwd_bwd),
           "loss_fwd_bwd_alone_ms": timed(loss_only)}
    out["shading_share"] = 1.0 - out["step_without_shading_ms"] / out["step_with_shading_ms"]
    print(json.dumps(out))


if __name__ == "__main__":
    main()
