"""Triangle-size statistics of the sphere workloads (one element): how many faces a "bbox <= 2x2 pixels" fast path
 would serve.  CPU only.
    python tools/tri_stats.py            # c4 (256^2) and headline (512^2)"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import scenes  # noqa: E402

for H in (256, 512):
    W = H
    sc = scenes.sphere(1, H, W)
    v = sc["vertices"][0].astype(np.float64)
    f = sc["faces"][0]
    win = np.stack([(v[:, 0] / v[:, 3] + 1) * W / 2, (v[:, 1] / v[:, 3] + 1) * H / 2], 1)
    tri = win[f]
    xmin = np.clip(np.ceil(tri[:, :, 0].min(1) - 0.5), 0, W - 1).astype(int)
    xmax = np.clip(np.floor(tri[:, :, 0].max(1) - 0.5), 0, W - 1).astype(int)
    ymin = np.clip(np.ceil(tri[:, :, 1].min(1) - 0.5), 0, H - 1).astype(int)
    ymax = np.clip(np.floor(tri[:, :, 1].max(1) - 0.5), 0, H - 1).astype(int)
    bw, bh = xmax - xmin + 1, ymax - ymin + 1
    valid = (bw > 0) & (bh > 0)
    area = np.abs((tri[:, 1, 0] - tri[:, 0, 0]) * (tri[:, 2, 1] - tri[:, 0, 1]) - (tri[:, 1, 1] - tri[:, 0, 1]) * (tri[:, 2, 0] - tri[:, 0, 0])) / 2
    print("%dx%d: %d faces, %.1f %% with no pixel centre in their box" % (H, W, len(f), 100 * (~valid).mean()))
    bw, bh = bw[valid], bh[valid]
    for k in (1, 2, 3, 4):
        print("  box <= %dx%d px: %5.1f %% of the faces   (shorter side <= %d: %5.1f %%)"
              % (k, k, 100 * ((bw <= k) & (bh <= k)).mean(), k, 100 * (np.minimum(bw, bh) <= k).mean()))
    print("  mean: %.1f lines per face, shorter side %.1f px, box %.0f px, area %.1f px, %.2f fragments per line"
          % (np.maximum(bw, bh).mean(), np.minimum(bw, bh).mean(), (bw * bh).mean(), area.mean(), area[valid].sum() / np.maximum(bw, bh).sum()))
