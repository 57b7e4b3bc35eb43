"""Minimal PNG writer (the reference samples display through OpenCV, which this image does not have)."""
import struct
import zlib

import numpy as np


def save_png(path, image):
    """image: float [H,W,3] in [0,1] (clipped) or uint8."""
    a = np.asarray(image)
    if a.dtype != np.uint8:
        a = (np.clip(np.nan_to_num(a), 0, 1) * 255 + 0.5).astype(np.uint8)
    h, w, _ = a.shape
    raw = b"".join(b"\x00" + a[y].tobytes() for y in range(h))

    def chunk(tag, data):
        return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)

    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 2, 0, 0, 0)) +
                chunk(b"IDAT", zlib.compress(raw, 6)) + chunk(b"IEND", b""))
