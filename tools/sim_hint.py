"""Development aid (CPU only): what ordering the faces buys on one element of the headline scene when the
geometry moves between frames, and what a two-level occlusion test leaves to sample.

Frame 1 = frame 0 turned by `deg` degrees.  Bucket 0 is either the faces visible in frame 0 (the
visibility hint) or the nearest `frac` of the faces; the pyramid comes from an oracle render of bucket 0
alone.  Level 1 tests each remaining face (vertex-minimum depth) at the pyramid level where its box spans
<= 3 tiles per axis; level 2 tests every group of 8 scan lines of a survivor against the 8-px tiles its
covered samples touch.  Prints faces / scan lines left after each level.
    python tools/sim_hint.py [hint|frac=0.35] [deg=0.25] [H=512]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import scenes                      # noqa: E402
from oracle import golden          # noqa: E402

mode = sys.argv[1] if len(sys.argv) > 1 else "hint"
deg = float(sys.argv[2]) if len(sys.argv) > 2 else 0.25
H = W = int(sys.argv[3]) if len(sys.argv) > 3 else 512
sc0, sc1 = (scenes.sphere_frames(1, H, W, 3, 224, 224, frame=k, step_deg=deg) for k in (0, 1))
f = sc1["faces"][0]
v = sc1["vertices"][0].astype(np.float64)
win = np.stack([(v[:, 0] / v[:, 3] + 1) * W / 2, (v[:, 1] / v[:, 3] + 1) * H / 2, v[:, 2] / v[:, 3] * 0.5 + 0.5], 1)
tri = win[f]
znear = tri[:, :, 2].min(1)
xmin = np.clip(np.ceil(tri[:, :, 0].min(1) - 0.5), 0, W - 1).astype(int)
xmax = np.clip(np.floor(tri[:, :, 0].max(1) - 0.5), 0, W - 1).astype(int)
ymin = np.clip(np.ceil(tri[:, :, 1].min(1) - 0.5), 0, H - 1).astype(int)
ymax = np.clip(np.floor(tri[:, :, 1].max(1) - 0.5), 0, H - 1).astype(int)
valid = (xmax >= xmin) & (ymax >= ymin)
lines = np.where(valid, np.maximum(xmax - xmin, ymax - ymin) + 1, 0)
first = np.zeros(len(f), bool)
if mode == "hint":
    _, fid0, _ = golden.forward(sc0["background"], sc0["vertices"], sc0["vertex_colors"], sc0["faces"], threads=8)
    first[np.unique(fid0[fid0 >= 0])] = True
else:
    first[np.argsort(np.where(valid, znear, 9.0), kind="stable")[:int(float(mode) * len(f))]] = True
first &= valid
_, fid, gb = golden.forward(sc1["background"], sc1["vertices"], sc1["vertex_colors"], np.ascontiguousarray(sc1["faces"][:, first]), threads=8)
depth = np.where(fid[0] >= 0, gb[0, :, :, 3], 1.0)[::-1]        # row 0 = bottom, like window y


def pyramid(t):
    hh, ww = (H + t - 1) // t, (W + t - 1) // t
    p = np.zeros((hh * t, ww * t))
    p[:H, :W] = depth
    return p.reshape(hh, t, ww, t).max(axis=(1, 3))


pyr = {t: pyramid(t) for t in (8, 16, 32, 64)}
rest = np.nonzero(~first & valid)[0]
print("bucket 0: %d faces, %d lines; rest: %d faces, %d lines" % (first.sum(), lines[first].sum(), len(rest), lines[rest].sum()))
s1f = s1l = s2l = chunks = 0
for i in rest:
    t = 8
    while t <= 64 and max(xmax[i] // t - xmin[i] // t, ymax[i] // t - ymin[i] // t) + 1 > 3:
        t *= 2
    if t <= 64 and znear[i] > pyr[t][ymin[i] // t:ymax[i] // t + 1, xmin[i] // t:xmax[i] // t + 1].max():
        continue
    s1f += 1
    s1l += lines[i]
    p = tri[i]
    x0, x1, y0, y1 = xmin[i], xmax[i], ymin[i], ymax[i]
    X, Y = np.meshgrid(np.arange(x0, x1 + 1) + 0.5, np.arange(y0, y1 + 1) + 0.5)
    a2 = (p[1, 0] - p[0, 0]) * (p[2, 1] - p[0, 1]) - (p[1, 1] - p[0, 1]) * (p[2, 0] - p[0, 0])
    if a2 == 0:
        continue
    l1 = ((X - p[0, 0]) * (p[2, 1] - p[0, 1]) - (Y - p[0, 1]) * (p[2, 0] - p[0, 0])) / a2
    l2 = ((p[1, 0] - p[0, 0]) * (Y - p[0, 1]) - (p[1, 1] - p[0, 1]) * (X - p[0, 0])) / a2
    cov = (l1 >= 0) & (l2 >= 0) & (1 - l1 - l2 >= 0)
    z = p[0, 2] * (1 - l1 - l2) + p[1, 2] * l1 + p[2, 2] * l2
    cols = (x1 - x0) > (y1 - y0)                    # lines are columns
    L0 = x0 if cols else y0
    if cols:
        cov, z = cov.T, z.T
    for c0 in range(-(L0 % 8), cov.shape[0], 8):    # chunks aligned to the 8-px tile rows
        a, b = max(c0, 0), min(c0 + 8, cov.shape[0])
        cc = cov[a:b]
        if not cc.any():
            continue
        chunks += 1
        rr, pp = np.nonzero(cc)
        la, lb, pa, pb = a + rr.min(), a + rr.max(), pp.min(), pp.max()
        bx0, bx1, by0, by1 = (x0 + la, x0 + lb, y0 + pa, y0 + pb) if cols else (x0 + pa, x0 + pb, y0 + la, y0 + lb)
        if not z[a:b][cc].min() > pyr[8][by0 // 8:by1 // 8 + 1, bx0 // 8:bx1 // 8 + 1].max():
            s2l += b - a
print("level 1 (face vs pyramid): %d faces, %d lines survive" % (s1f, s1l))
print("level 2 (8-line chunks):   %d lines survive (of %d non-empty chunks tested)" % (s2l, chunks))
