"""Per-stage device times (ms) of fwd+bwd on a frame SEQUENCE (the geometry moves between steps), with and
without the visibility hint; checks on the way that the hinted output is bit-identical.  Development aid.
    python tools/stages_seq.py [headline|c4|h8] [deg=0.25] [noise_px=0] [frames=4]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import scenes  # noqa: E402
from dirt_b200 import rasterise_ops as ops, _lib  # noqa: E402

SHAPES = {"headline": (64, 512, 512), "c4": (256, 256, 256), "h8": (8, 512, 512)}


def main():
    argv = [a for a in sys.argv if not a.startswith("--")]
    name = argv[1] if len(argv) > 1 else "headline"
    deg = float(argv[2]) if len(argv) > 2 else 0.25
    noise = float(argv[3]) if len(argv) > 3 else 0.0
    nfr = int(argv[4]) if len(argv) > 4 else 4
    B, H, W = SHAPES[name]
    lib = _lib.load()
    sc = scenes.sphere(B, H, W)
    d = {k: torch.from_numpy(np.ascontiguousarray(v)).cuda() for k, v in sc.items()}
    V = sc["vertices"].shape[1]
    frames = [d["vertices"]] + [torch.from_numpy(scenes.sphere_frames(B, H, W, frame=k, step_deg=deg, noise_px=noise, vertices_only=True)).cuda()
                                for k in range(1, nfr)]
    order = list(range(nfr)) + list(range(nfr - 2, 0, -1))        # 0 1 2 3 2 1 | 0 1 ...: every step moves by one frame
    g = torch.randn(B, H, W, 3, device="cuda")
    hint = ops.visibility_hint(torch.device("cuda", 0), B, sc["faces"].shape[1], key="seq")

    def step(i, h):
        v = frames[order[i % len(order)]]
        px, fid, gb = ops.forward_raw(d["background"], v, d["vertex_colors"], d["faces"], hint=h)
        return (px, fid, gb), ops.backward_raw(d["faces"], px, g, fid, gb, V)

    if "--ncu" in sys.argv:            # a few hinted steps for a profiler: no checks, no timing
        for i in range(6):
            step(i, hint)
        torch.cuda.synchronize()
        return
    # parity of the hinted path, frame by frame
    for i in range(len(order) + 2):
        a, _ = step(i, None)
        b, _ = step(i, hint)
        assert all(torch.equal(x, y) for x, y in zip(a, b)), "hinted output differs at step %d" % i
    print("%s: hinted output bit-identical over %d steps" % (name, len(order) + 2))
    only = os.environ.get("SEQ_ONLY")
    for label, h, moving in (("static, no hint", None, False), ("moving, no hint", None, True), ("moving, hint", hint, True), ("static, hint", hint, False)):
        if only and label not in only.split(";"):
            continue
        idx = (lambda i: i) if moving else (lambda i: 0)
        for i in range(4):
            step(idx(i), h)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 12
        e0.record()
        for i in range(n):
            step(idx(i + 4), h)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / n
        lib.dirt_profile_reset(); lib.dirt_profile_enable(1)
        for i in range(6):
            step(idx(i + 4), h); torch.cuda.synchronize(); lib.dirt_profile_collect()
        lib.dirt_profile_enable(0)
        st = {k: m / 6 for k, (m, c) in _lib.stage_times().items()}
        print("%-8s %-16s step %.3f ms = %.2f Gpix/s | %s" % (name, label, ms, B * H * W / ms / 1e6,
              " ".join("%s=%.3f" % (k, v) for k, v in st.items() if v > 0.0005)))


if __name__ == "__main__":
    main()
