"""numpy restatement of the silhouette dilation of samples/deferred.py:87-91 (reference):
    mask = (x == -inf);  out = x * (1 - mask) + max_pool3x3_SAME(x) * mask
and of its gradient (background pixel -> first maximal element, row-major, of its 3x3 window).
TEST INFRASTRUCTURE ONLY, like the rest of oracle/.  Pure loops: small cases only."""
import numpy as np


def _window(x, r, c, H, W):
    for k in range(9):
        rr, cc = r + k // 3 - 1, c + k % 3 - 1
        if 0 <= rr < H and 0 <= cc < W:
            yield k, rr, cc


def dilate(x):
    x = np.asarray(x, np.float32)
    B, H, W, C = x.shape
    out = x.copy()
    for b, r, c, ch in zip(*np.nonzero(x == -np.inf)):
        out[b, r, c, ch] = max(x[b, rr, cc, ch] for _, rr, cc in _window(x, r, c, H, W))
    return out


def dilate_grad(x, grad_out):
    x = np.asarray(x, np.float32)
    g = np.asarray(grad_out, np.float32)
    B, H, W, C = x.shape
    fg = x != -np.inf
    gx = np.where(fg, g, 0).astype(np.float64)
    for b, r, c, ch in zip(*np.nonzero(~fg)):
        best, arg = -np.inf, None
        for _, rr, cc in _window(x, r, c, H, W):
            if x[b, rr, cc, ch] > best:
                best, arg = x[b, rr, cc, ch], (rr, cc)
        if arg is not None:
            gx[b, arg[0], arg[1], ch] += g[b, r, c, ch]
    return gx.astype(np.float32)
