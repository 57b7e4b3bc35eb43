"""Development aid (CPU only): how many scan lines of the SECOND depth bucket survive three culling
tests on one element of the headline scene -- (a) today's: the pyramid level where the face's box
spans <= 3x3 tiles; (b) exact 8-px tiles over the whole box; (c) 8-px tiles that the TRIANGLE touches
(a walk along the sliver instead of its box).  Depths: vertex minimum (the kernel's plane bound is a
little more conservative), pyramid from an oracle render of the nearest 35 % of the faces.
    python tools/sim_cull.py [first_fraction]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import scenes                      # noqa: E402
from oracle import golden          # noqa: E402

frac = float(sys.argv[1]) if len(sys.argv) > 1 else 0.35
H = W = 512
sc = scenes.sphere(1, H, W, 3, 224, 224, seeds=(1, 2, 3, 4))
v, f = sc["vertices"][0].astype(np.float64), sc["faces"][0]
win = np.stack([(v[:, 0] / v[:, 3] + 1) * W / 2, (v[:, 1] / v[:, 3] + 1) * H / 2, v[:, 2] / v[:, 3] * 0.5 + 0.5], 1)   # x, y (up), depth
tri = win[f]                                                    # [F,3,3]
znear = tri[:, :, 2].min(1)
order = np.argsort(znear, kind="stable")
n0 = int(frac * len(f))
first, rest = order[:n0], order[n0:]
_, fid, gb = golden.forward(sc["background"], sc["vertices"], sc["vertex_colors"], sc["faces"][:, first], threads=8)
depth = np.where(fid[0] >= 0, gb[0, :, :, 3], 1.0)[::-1]        # row 0 = bottom, like window y


def pyramid(d, t):
    hh, ww = (H + t - 1) // t, (W + t - 1) // t
    p = np.full((hh * t, ww * t), 0.0)
    p[:H, :W] = d
    return p.reshape(hh, t, ww, t).max(axis=(1, 3))


pyr = {t: pyramid(depth, t) for t in (8, 16, 32, 64)}
xmin = np.clip(np.ceil(tri[:, :, 0].min(1) - 0.5), 0, W - 1).astype(int)
xmax = np.clip(np.floor(tri[:, :, 0].max(1) - 0.5), 0, W - 1).astype(int)
ymin = np.clip(np.ceil(tri[:, :, 1].min(1) - 0.5), 0, H - 1).astype(int)
ymax = np.clip(np.floor(tri[:, :, 1].max(1) - 0.5), 0, H - 1).astype(int)
lines = np.maximum(xmax - xmin, ymax - ymin) + 1
valid = (xmax >= xmin) & (ymax >= ymin)

surv = {"a_box_3x3_level": 0, "b_box_8px": 0, "c_triangle_8px": 0}
faces_surv = dict.fromkeys(surv, 0)
total_lines = 0
for i in rest:
    if not valid[i]:
        continue
    total_lines += lines[i]
    z = znear[i]
    ext = max(xmax[i] - xmin[i], ymax[i] - ymin[i])
    # (a) the level where the box spans at most 3x3 tiles
    t = 8
    while t <= 64 and (ext // t) > 1:
        t *= 2
    hid_a = t <= 64 and z > pyr[t][ymin[i] // t:ymax[i] // t + 1, xmin[i] // t:xmax[i] // t + 1].max()
    # (b) every 8-px tile of the box
    blk = pyr[8][ymin[i] // 8:ymax[i] // 8 + 1, xmin[i] // 8:xmax[i] // 8 + 1]
    hid_b = z > blk.max()
    # (c) only the 8-px tiles the triangle itself can touch (edge test against the tile's best corner)
    p = tri[i, :, :2]
    area = (p[1, 0] - p[0, 0]) * (p[2, 1] - p[0, 1]) - (p[1, 1] - p[0, 1]) * (p[2, 0] - p[0, 0])
    sg = 1.0 if area > 0 else -1.0
    ty, tx = np.mgrid[ymin[i] // 8:ymax[i] // 8 + 1, xmin[i] // 8:xmax[i] // 8 + 1]
    touch = np.ones(tx.shape, bool)
    for k in range(3):
        a, b = p[(k + 1) % 3], p[(k + 2) % 3]
        ex, ey = (b[0] - a[0]) * sg, (b[1] - a[1]) * sg
        cx = np.where(-ey > 0, tx * 8 + 8, tx * 8)            # the corner that maximises the edge function
        cy = np.where(ex > 0, ty * 8 + 8, ty * 8)
        touch &= ex * (cy - a[1]) - ey * (cx - a[0]) >= 0
    hid_c = (not touch.any()) or z > blk[touch].max()
    for name, hid in (("a_box_3x3_level", hid_a), ("b_box_8px", hid_b), ("c_triangle_8px", hid_c)):
        if not hid:
            surv[name] += lines[i]
            faces_surv[name] += 1
print("first bucket %.2f: second bucket has %d faces, %d lines" % (frac, len(rest), total_lines))
for name in surv:
    print("  %-18s surviving faces %6d  lines %8d (%.1f %%)" % (name, faces_surv[name], surv[name], 100.0 * surv[name] / total_lines))
