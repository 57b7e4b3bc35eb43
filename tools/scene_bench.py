"""Times the fork-style workload: full-screen quad + oceanic_horizon scene stage (the fork renders
1024x768 and 960x540 frames this way, tests/optimize_sequence.py:24, tests/square_test.py:10).
    python tools/scene_bench.py  -> one JSON line (Mpix/s, forward only: the fork's op has no gradient)
The CPU figure is the C transcription of the GLSL (oracle/scene_oceanic.c), 1 thread, one frame."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import dirt_b200  # noqa: E402
from oracle import golden  # noqa: E402
from test_scene_oceanic import quad_scene, CAMS  # noqa: E402


def main():
    B, H, W, C = 16, 768, 1024, 3
    bg, v, c, f = quad_scene(B, H, W, C, seed=0, noise=0.3)
    t = [torch.from_numpy(x).cuda() for x in (bg, v, c, f)]
    cam = torch.tensor(CAMS[1], dtype=torch.float32, device="cuda")
    for _ in range(3):
        dirt_b200.oceanic_horizon_batch(*t, cam)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 20
    e0.record()
    for _ in range(n):
        dirt_b200.oceanic_horizon_batch(*t, cam)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    t0 = time.perf_counter()
    golden.oceanic_horizon(bg[:1], v[:1], f[:1], CAMS[1])
    cpu_s = time.perf_counter() - t0
    print(json.dumps({"workload": "oceanic_horizon B=%d %dx%dx%d full-screen quad" % (B, W, H, C), "ms_per_call": ms,
                      "Mpix_per_s": B * H * W / ms / 1e3, "cpu_port_Mpix_per_s_1thread": H * W / cpu_s / 1e6}))
    other_scenes()


def other_scenes():
    """The other scene ops at the fork's own canvas (960x540, tests/square_test.py:10), B=4; the CPU
    figure is oracle/scene_shaders.cpp on one thread over a 240x135 frame."""
    import test_scene_shaders as T
    B, H, W = 4, 540, 960
    for name in T.OCEAN_SCENES + ["hill"]:
        if name == "hill":
            _, v, c, f = T.quad(B, H, W, 3)
            bg = np.broadcast_to(T.terrain_texture(H, W), (B, H, W, 4)).copy()
            cam = T.HILL_CAM
        else:
            bg, v, c, f = T.quad(B, H, W, 3, seed=1, noise=0.3)
            cam = T.CAM_A
        t = [torch.from_numpy(x).cuda() for x in (bg, v, c, f)]
        camd = torch.tensor(np.asarray(cam, np.float32), device="cuda")
        fn = getattr(dirt_b200, name + "_batch")
        for _ in range(2):
            fn(*t, camd)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 5
        e0.record()
        for _ in range(n):
            fn(*t, camd)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / n
        h, w = H // 4, W // 4
        if name == "hill":
            sb = T.terrain_texture(h, w)[None]
        else:
            sb = bg[:1, :h, :w]
        t0 = time.perf_counter()
        golden.scene(name, sb, v[:1], f[:1], cam)
        cpu_s = time.perf_counter() - t0
        print(json.dumps({"workload": "%s B=%d %dx%dx3 full-screen quad" % (name, B, W, H), "ms_per_call": ms,
                          "Mpix_per_s": B * H * W / ms / 1e3, "cpu_port_Mpix_per_s_1thread": h * w / cpu_s / 1e6}))


if __name__ == "__main__":
    main()
