"""Development aid (CPU only): what the chunk-level occlusion test of k_setup leaves to sample on one element
of the headline scene, modelled the way the kernel does it (conservative rectangle of an 8-line chunk from the
edge crossings at its two end lines, per 8-px tile the plane's minimum over the rectangle's part in that tile).
    python tools/sim_chunk.py [hint|frac=0.35] [deg=0.25] [H=512] [chunk=8]"""
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
CH = int(sys.argv[4]) if len(sys.argv) > 4 else 8
nlat = int(os.environ.get("NLAT", 224))
sc0, sc1 = (scenes.sphere_frames(1, H, W, 3, nlat, nlat, frame=k, step_deg=deg) for k in (0, 1))
f = sc1["faces"][0]
v = sc1["vertices"][0].astype(np.float64)
win = np.stack([(v[:, 0] / v[:, 3] + 1) * W / 2, (v[:, 1] / v[:, 3] + 1) * H / 2, v[:, 2] / v[:, 3] * 0.5 + 0.5], 1)
win[:, :2] = np.floor(win[:, :2] * 256 + 0.5) / 256          # the 1/256 px snap
tri = win[f]
znear = tri[:, :, 2].min(1)
xmin = np.clip(np.ceil(tri[:, :, 0].min(1) - 0.5), 0, W - 1).astype(int)
xmax = np.clip(np.floor(tri[:, :, 0].max(1) - 0.5), 0, W - 1).astype(int)
ymin = np.clip(np.ceil(tri[:, :, 1].min(1) - 0.5), 0, H - 1).astype(int)
ymax = np.clip(np.floor(tri[:, :, 1].max(1) - 0.5), 0, H - 1).astype(int)
a2 = (tri[:, 1, 0] - tri[:, 0, 0]) * (tri[:, 2, 1] - tri[:, 0, 1]) - (tri[:, 1, 1] - tri[:, 0, 1]) * (tri[:, 2, 0] - tri[:, 0, 0])
valid = (xmax >= xmin) & (ymax >= ymin) & (a2 != 0)
lines = np.where(valid, np.maximum(xmax - xmin, ymax - ymin) + 1, 0)
first = np.zeros(len(f), bool)
if mode == "hint":
    _, fid0, _ = golden.forward(sc0["background"], sc0["vertices"], sc0["vertex_colors"], sc0["faces"], threads=8)
    first[np.unique(fid0[fid0 >= 0])] = True
elif mode.startswith("tile="):
    # per screen tile (centroid of the face's box) its own depth quantile: "tile=<fraction>,<tile px>"
    q, tpx = mode[5:].split(",")
    q, tpx = float(q), int(tpx)
    cxs = np.clip((xmin + xmax) // 2 // tpx, 0, (W - 1) // tpx)
    cys = np.clip((ymin + ymax) // 2 // tpx, 0, (H - 1) // tpx)
    tid = cys * ((W - 1) // tpx + 1) + cxs
    for t_ in np.unique(tid[valid]):
        idx = np.nonzero(valid & (tid == t_))[0]
        k = max(1, int(np.ceil(q * len(idx))))
        first[idx[np.argsort(znear[idx], kind="stable")[:k]]] = True
else:
    first[np.argsort(np.where(valid, znear, 9.0), kind="stable")[:int(float(mode) * len(f))]] = True
first &= valid
_, fid, gb = golden.forward(sc1["background"], sc1["vertices"], sc1["vertex_colors"], np.ascontiguousarray(sc1["faces"][:, first]), threads=8)
depth = np.where(fid[0] >= 0, gb[0, :, :, 3], 1.0)[::-1]        # row 0 = bottom, like window y
_, fidall, _ = golden.forward(sc1["background"], sc1["vertices"], sc1["vertex_colors"], sc1["faces"], threads=8)
vis1 = np.zeros(len(f), bool); vis1[np.unique(fidall[fidall >= 0])] = True


def pyramid(t):
    hh, ww = (H + t - 1) // t, (W + t - 1) // t
    p = np.zeros((hh * t, ww * t))
    p[:H, :W] = depth
    return p.reshape(hh, t, ww, t).max(axis=(1, 3))


pyr = {t: pyramid(t) for t in (2, 4, 8, 16, 32, 64)}
TT = int(os.environ.get("TILE", 8))
rest = np.nonzero(~first & valid)[0]
print("bucket 0: %d faces, %d lines (visible in this frame: %d faces, of which %d in bucket 0); rest: %d faces, %d lines"
      % (first.sum(), lines[first].sum(), vis1.sum(), (vis1 & first).sum(), len(rest), lines[rest].sum()))
s1f = s1l = 0
chunks = ch_empty = ch_vis = ch_lines = tiles_tested = 0
vis_lines_rest = 0
for i in rest:
    t = 8
    while t <= 64 and max(xmax[i] // t - xmin[i] // t, ymax[i] // t - ymin[i] // t) + 1 > 3:
        t *= 2
    p = tri[i]
    # plane z = z0 + gx (x - x0) + gy (y - y0)
    d1, d2 = p[1] - p[0], p[2] - p[0]
    gx = (d1[2] * d2[1] - d2[2] * d1[1]) / a2[i]
    gy = (d2[2] * d1[0] - d1[2] * d2[0]) / a2[i]
    zat = lambda x, y: p[0, 2] + gx * (x + 0.5 - p[0, 0]) + gy * (y + 0.5 - p[0, 1])
    cx = xmin[i] if gx >= 0 else xmax[i]
    cy = ymin[i] if gy >= 0 else ymax[i]
    zface = max(zat(cx, cy), znear[i]) if os.environ.get("VMIN", "1") == "1" else zat(cx, cy)
    if t <= 64 and zface > pyr[t][ymin[i] // t:ymax[i] // t + 1, xmin[i] // t:xmax[i] // t + 1].max():
        continue
    s1f += 1
    s1l += lines[i]
    if vis1[i]:
        vis_lines_rest += lines[i]
    cols = (xmax[i] - xmin[i]) > (ymax[i] - ymin[i])                    # lines are columns
    L0, L1 = (xmin[i], xmax[i]) if cols else (ymin[i], ymax[i])         # line range (absolute)
    P0, P1 = (ymin[i], ymax[i]) if cols else (xmin[i], xmax[i])         # position range (absolute)
    sg = 1.0 if a2[i] > 0 else -1.0
    edges = []
    for a, b in ((1, 2), (2, 0), (0, 1)):
        ex, ey = p[b, 0] - p[a, 0], p[b, 1] - p[a, 1]
        # e(x, y) = ex (y - ay) - ey (x - ax), inside: sg * e >= 0
        if cols:      # line = x, position = y
            edges.append((sg * ex, -sg * ey, sg * (-ex * p[a, 1] + ey * p[a, 0])))       # coefficient of position, of line, constant
        else:
            edges.append((-sg * ey, sg * ex, sg * (-ex * p[a, 1] + ey * p[a, 0])))
    for c0 in range(L0 - L0 % CH, L1 + 1, CH):
        ja, jb = max(c0, L0), min(c0 + CH - 1, L1)
        chunks += 1
        lo, hi = float(P0), float(P1)
        empty = False
        for (kp, kl, kc) in edges:
            va, vb = kl * (ja + 0.5) + kc, kl * (jb + 0.5) + kc        # edge value at position coordinate 0 (continuous), lines ja / jb
            if kp > 0:
                lo = max(lo, min(-va / kp, -vb / kp) - 0.5)           # pixel centre c + 0.5 >= crossing
            elif kp < 0:
                hi = min(hi, max(-va / kp, -vb / kp) - 0.5)
            elif va < 0 and vb < 0:
                empty = True
        ilo, ihi = int(np.floor(lo - 0.01)), int(np.ceil(hi + 0.01))
        ilo, ihi = max(ilo, P0), min(ihi, P1)
        if empty or ilo > ihi:
            ch_empty += 1
            continue
        seen = False
        for tt in range(ilo // TT, ihi // TT + 1):
            q0, q1 = max(ilo, tt * TT), min(ihi, tt * TT + TT - 1)
            tiles_tested += 1
            if cols:
                zx = ja if gx >= 0 else jb
                zy = q0 if gy >= 0 else q1
                hz = pyr[TT][tt, ja // TT]
            else:
                zx = q0 if gx >= 0 else q1
                zy = ja if gy >= 0 else jb
                hz = pyr[TT][ja // TT, tt]
            if not max(zat(zx, zy), znear[i]) > hz:
                seen = True
                break
        if seen:
            ch_vis += 1
            ch_lines += jb - ja + 1
print("level 1 (face vs pyramid): %d faces, %d lines survive (lines of truly visible faces among them: %d)" % (s1f, s1l, vis_lines_rest))
print("level 2 (%d-line chunks, conservative rectangles): %d chunks, %d empty, %d visible -> %d lines; %d tile tests"
      % (CH, chunks, ch_empty, ch_vis, ch_lines, tiles_tested))
