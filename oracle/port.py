"""TEST INFRASTRUCTURE ONLY — ctypes loader for the plain-C oracle (oracle/yuv2bgr_port.c).

Restates reference `python/read_image.py:94`'s conversion on one core with no SIMD.  Built by
`oracle/Makefile` (also from `__graft_entry__.build()`).  Only tests/, smoke() and bench.py's
cpu_baseline leg may import this.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libyuv2bgr_port.so")
_lib = None


def build() -> str:
    subprocess.check_call(["make", "-s", "-C", _HERE])
    return _SO


def _load():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        _lib = ctypes.CDLL(_SO)
        p, i = ctypes.c_void_p, ctypes.c_int
        _lib.port_yuv420p_to_bgr24.argtypes = [p, i, p, i, p, i, i, i, i, p, i]
        _lib.port_yuv422p_to_bgr24.argtypes = [p, i, p, i, p, i, i, i, i, p, i]
        _lib.port_nv12_to_bgr24.argtypes = [p, i, p, i, i, i, p, i, i]
        _lib.port_yuv420p10_to_bgr24.argtypes = [p, i, p, i, p, i, i, i, p, i]
        _lib.port_yuv444p_to_bgr24.argtypes = [p, i, p, i, p, i, i, i, i, p, i]
    return _lib


def to_bgr24(fmt: str, planes) -> np.ndarray:
    lib = _load()
    planes = [np.ascontiguousarray(p) for p in planes]
    h, w = planes[0].shape
    out = np.empty((h, w, 3), np.uint8)
    if fmt == "gray":                       # B = G = R = Y; nothing to restate in C
        if w <= 0 or h <= 0 or w % 2 or h % 2:
            raise ValueError(f"port rejected geometry {w}x{h} ({fmt})")
        out[...] = planes[0][..., None]
        return out
    if fmt in ("yuv420p", "yuvj420p"):
        Y, U, V = planes
        r = lib.port_yuv420p_to_bgr24(Y.ctypes.data, Y.strides[0], U.ctypes.data, U.strides[0],
                                      V.ctypes.data, V.strides[0], w, h,
                                      1 if fmt == "yuvj420p" else 0, out.ctypes.data, w * 3)
    elif fmt in ("yuv444p", "yuvj444p"):
        Y, U, V = planes
        r = lib.port_yuv444p_to_bgr24(Y.ctypes.data, Y.strides[0], U.ctypes.data, U.strides[0],
                                      V.ctypes.data, V.strides[0], w, h,
                                      1 if fmt == "yuvj444p" else 0, out.ctypes.data, w * 3)
    elif fmt == "yuv420p10le":
        Y, U, V = planes
        r = lib.port_yuv420p10_to_bgr24(Y.ctypes.data, Y.strides[0], U.ctypes.data, U.strides[0],
                                        V.ctypes.data, V.strides[0], w, h, out.ctypes.data, w * 3)
    elif fmt in ("yuv422p", "yuvj422p"):
        Y, U, V = planes
        r = lib.port_yuv422p_to_bgr24(Y.ctypes.data, Y.strides[0], U.ctypes.data, U.strides[0],
                                      V.ctypes.data, V.strides[0], w, h,
                                      1 if fmt == "yuvj422p" else 0, out.ctypes.data, w * 3)
    elif fmt in ("nv12", "nv21"):
        Y, UV = planes
        r = lib.port_nv12_to_bgr24(Y.ctypes.data, Y.strides[0], UV.ctypes.data, UV.strides[0],
                                   w, h, out.ctypes.data, w * 3, 1 if fmt == "nv21" else 0)
    else:
        raise ValueError(fmt)
    if r != 0:
        raise ValueError(f"port rejected geometry {w}x{h} ({fmt})")
    return out
