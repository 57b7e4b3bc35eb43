"""The line-span arithmetic of dirt_b200/csrc/setup_raster.cu (sample_line): the covered span of a
scan line is bounded per edge by ceil(na / s), na = ceil(-a / 256), which the kernel evaluates as
ceil((float(na) - 0.5) * (1 / s)) in binary32 with every operation rounded to nearest.  That form
must be EXACT for |na| <= 32256^2 / 256 + 1 (< 2^22) and 1 <= s <= 32256 (kSmallExtent: an edge value at a
point of the bounding box is at most extent_x * extent_y), otherwise float rounding
would decide coverage.  numpy float32 arithmetic rounds each operation the same way."""
import numpy as np


def _kernel_form(na, s):
    inv = (np.float32(1) / s.astype(np.float32)).astype(np.float32)
    x = (na.astype(np.float32) - np.float32(0.5)).astype(np.float32)
    return np.ceil((x * inv).astype(np.float32)).astype(np.int64)


def _exact(na, s):
    return -((-na) // s)


def test_ceiling_division_is_exact_on_random_and_adversarial_operands():
    rng = np.random.default_rng(0)
    lim = 32256 * 32256 // 256 + 2
    assert lim < 2 ** 22
    n = 2_000_000
    cases = []
    s = rng.integers(1, 32256 + 1, n).astype(np.int64)
    cases.append((rng.integers(-lim, lim + 1, n), s))
    q = rng.integers(-(lim // s), lim // s + 1)                       # exact multiples and their neighbours
    cases.append((np.clip(q * s + rng.integers(-1, 2, n), -lim, lim), s))
    s_small = rng.integers(1, 64, n).astype(np.int64)                  # largest quotients
    cases.append((rng.integers(-lim, lim + 1, n), s_small))
    cases.append(((rng.integers(0, 2, n) * 2 - 1) * (lim - rng.integers(0, 4096, n)), s))   # largest operands
    for na, s in cases:
        assert np.array_equal(_kernel_form(na, s), _exact(na, s))


def test_all_small_operands_exhaustively():
    na = np.arange(-4096, 4097, dtype=np.int64)[:, None]
    s = np.arange(1, 1025, dtype=np.int64)[None, :]
    na, s = np.broadcast_arrays(na, s)
    assert np.array_equal(_kernel_form(na.ravel(), s.ravel()), _exact(na.ravel(), s.ravel()))
