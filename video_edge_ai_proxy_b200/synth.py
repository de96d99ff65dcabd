"""Synthetic decoded frames (the stand-in for `p.decode()` at reference `python/read_image.py:87`).

Generator fixed by SURVEY.md §8(c): `rng = np.random.default_rng(seed)`, then Y (H,W), U (H/2,W/2),
V (H/2,W/2) drawn in that order as uniform uint8; NV12 interleaves U,V into one (H/2,W) plane.
The golden checksums under `tests/golden/` are tied to this exact draw order.
"""
from __future__ import annotations

import numpy as np

FORMATS = ("yuv420p", "yuvj420p", "nv12", "nv21", "yuv422p", "yuvj422p", "yuv420p10le", "yuv444p", "yuvj444p", "gray")


def make_planes(fmt: str, w: int, h: int, seed: int, dist: str = "uniform"):
    """Return the plane tuple for one frame: (Y,U,V) for planar formats, (Y,UV) for nv12 / (Y,VU) for nv21."""
    if fmt not in FORMATS:
        raise ValueError(fmt)
    if w % 2 or h % 2:
        raise ValueError("4:2:0 frames need even width and height")
    rng = np.random.default_rng(seed)
    if dist == "uniform":
        lo_y, hi_y, lo_c, hi_c = 0, 256, 0, 256
    elif dist == "video":  # legal limited-range levels
        lo_y, hi_y, lo_c, hi_c = 16, 236, 16, 241
    else:
        raise ValueError(dist)
    if fmt == "gray":             # monochrome decode: one full-range luma plane
        return (rng.integers(lo_y, hi_y, (h, w), dtype=np.uint8),)
    if fmt == "yuv420p10le":      # 10-bit samples in little-endian 16-bit containers (HEVC Main10 decodes to this)
        s4 = lambda v: min(v * 4, 1024)
        Y = rng.integers(s4(lo_y), s4(hi_y), (h, w), dtype=np.uint16)
        U = rng.integers(s4(lo_c), s4(hi_c), (h // 2, w // 2), dtype=np.uint16)
        V = rng.integers(s4(lo_c), s4(hi_c), (h // 2, w // 2), dtype=np.uint16)
        return (Y, U, V)
    ch = h if fmt in ("yuv422p", "yuvj422p", "yuv444p", "yuvj444p") else h // 2      # 4:2:2 / 4:4:4 keep every chroma row
    cw = w if fmt in ("yuv444p", "yuvj444p") else w // 2                            # 4:4:4 keeps every chroma column
    Y = rng.integers(lo_y, hi_y, (h, w), dtype=np.uint8)
    U = rng.integers(lo_c, hi_c, (ch, cw), dtype=np.uint8)
    V = rng.integers(lo_c, hi_c, (ch, cw), dtype=np.uint8)
    if fmt in ("nv12", "nv21"):
        UV = np.empty((h // 2, w), np.uint8)
        UV[:, 0::2] = U if fmt == "nv12" else V
        UV[:, 1::2] = V if fmt == "nv12" else U
        return (Y, UV)
    return (Y, U, V)


def exhaustive_yuv_frame():
    """One 4096x4096 yuv420p frame whose 2x2 blocks enumerate all 2^16 (U,V) pairs 256 times and
    whose luma walks 0..255 against every pair: every (Y,U,V) triple occurs at least once."""
    h = w = 4096
    cy, cx = np.meshgrid(np.arange(h // 2), np.arange(w // 2), indexing="ij")
    # chroma block (cy,cx): u = cx & 255, v = cy & 255; 8x8 tiling index selects a Y quad
    U = (cx & 255).astype(np.uint8)
    V = (cy & 255).astype(np.uint8)
    quad = ((cy >> 8) * 8 + (cx >> 8)).astype(np.int32)  # 0..63 -> four luma values each
    Y = np.empty((h, w), np.uint8)
    Y[0::2, 0::2] = quad * 4 + 0
    Y[0::2, 1::2] = quad * 4 + 1
    Y[1::2, 0::2] = quad * 4 + 2
    Y[1::2, 1::2] = quad * 4 + 3
    return (Y, U, V)
