"""TEST INFRASTRUCTURE ONLY — live libswscale binding used as the parity authority.

The reference obtains its BGR24 pixels from PyAV (`av=8.0.2`, pinned at
reference `python/environment.yml:10`) at `python/read_image.py:94`
(`frame.to_ndarray(format='bgr24')`).  PyAV is not in the reference tree and is not
installed here; what it does for that call is
`sws_getCachedContext(w, h, src_fmt, w, h, AV_PIX_FMT_BGR24, SWS_BILINEAR)` followed by
`sws_scale`.  This module makes exactly those two libswscale calls through ctypes on the
libswscale that ships inside the `opencv-python-headless` wheel of this image
(libswscale 9.1.100 / FFmpeg 8.0.1), so the numbers it returns are the real library's, not a
restatement.  The same image runs on the GPU box, so the library is present there too.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s CPU-baseline / `--impl reference`
legs may import this file.  Nothing under `video_edge_ai_proxy_b200/` does.
"""
from __future__ import annotations

import ctypes
import glob
import os
import threading

import numpy as np

# AVPixelFormat ids of the bundled FFmpeg (checked in `_load()` via av_get_pix_fmt).
PIX_FMT = {"gray": 8, "yuv420p": 0, "bgr24": 3, "yuv422p": 4, "yuv444p": 5, "yuvj420p": 12, "yuvj422p": 13, "yuvj444p": 14, "nv12": 23, "nv21": 24, "yuv420p10le": 62}
SWS_BILINEAR = 2  # PyAV's default `interpolation` for VideoFrame.reformat

# swscale's SIMD row loops run in 16/32-pixel steps and touch bytes past the visible row end
# (SURVEY §8c), so every oracle buffer row is padded by this much and has this alignment.
_ROW_PAD = 128

_lock = threading.Lock()
_libs = None


class SwscaleUnavailable(RuntimeError):
    pass


def _load():
    global _libs
    with _lock:
        if _libs is not None:
            return _libs
        try:
            import cv2  # noqa: F401  (resolves the wheel's bundled sibling .so files)
        except Exception as e:  # pragma: no cover
            raise SwscaleUnavailable(f"cv2 import failed: {e}")
        root = os.path.join(os.path.dirname(os.path.dirname(cv2.__file__)),
                            "opencv_python_headless.libs")
        sw = sorted(glob.glob(os.path.join(root, "libswscale-*.so*")))
        au = sorted(glob.glob(os.path.join(root, "libavutil-*.so*")))
        if not sw or not au:
            raise SwscaleUnavailable(f"no bundled libswscale under {root}")
        avutil = ctypes.CDLL(au[0])
        sws = ctypes.CDLL(sw[0])
        sws.swscale_version.restype = ctypes.c_uint
        sws.sws_getContext.restype = ctypes.c_void_p
        sws.sws_getContext.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                       ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                       ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p,
                                       ctypes.c_void_p]
        sws.sws_freeContext.argtypes = [ctypes.c_void_p]
        sws.sws_freeContext.restype = None
        sws.sws_scale.restype = ctypes.c_int
        sws.sws_scale.argtypes = [ctypes.c_void_p,
                                  ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_int),
                                  ctypes.c_int, ctypes.c_int,
                                  ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_int)]
        avutil.av_get_cpu_flags.restype = ctypes.c_int
        avutil.av_force_cpu_flags.argtypes = [ctypes.c_int]
        avutil.av_force_cpu_flags.restype = None
        avutil.av_get_pix_fmt.argtypes = [ctypes.c_char_p]
        avutil.av_get_pix_fmt.restype = ctypes.c_int
        for name, want in PIX_FMT.items():
            got = avutil.av_get_pix_fmt(name.encode())
            if got != want:
                raise SwscaleUnavailable(f"pix fmt id of {name}: {got} != {want}")
        _libs = (sws, avutil)
        return _libs


def available() -> bool:
    try:
        _load()
        return True
    except SwscaleUnavailable:
        return False


def version() -> str:
    sws, _ = _load()
    v = sws.swscale_version()
    return f"{v >> 16}.{(v >> 8) & 255}.{v & 255}"


def cpu_flags() -> int:
    return _load()[1].av_get_cpu_flags()


AV_CPU_FLAG_SSSE3 = 0x0080


def on_simd_path() -> bool:
    """True when libswscale will take its x86 SIMD yuv2rgb path (the parity target)."""
    return bool(cpu_flags() & AV_CPU_FLAG_SSSE3)


def force_cpu_flags(flags: int) -> None:
    """`av_force_cpu_flags`; 0 selects the plain-C table path, -1 restores auto-detect.
    Process-global: only the oracle self-tests use it, and they restore it."""
    _load()[1].av_force_cpu_flags(flags)


def _padded(h: int, row_bytes: int):
    """(array, stride) with 64-byte aligned rows and slack after each row and at the end."""
    stride = (row_bytes + _ROW_PAD + 63) // 64 * 64
    raw = np.zeros(h * stride + _ROW_PAD + 64, dtype=np.uint8)
    off = (-raw.ctypes.data) % 64
    arr = raw[off:off + h * stride].reshape(h, stride)
    return arr, stride, raw


class Converter:
    """One SwsContext + reusable padded buffers for a fixed (fmt, w, h).  Not thread-safe; make
    one per thread (this is what the CPU baseline does)."""

    def __init__(self, fmt: str, w: int, h: int):
        sws, _ = _load()
        self._sws = sws
        self.fmt, self.w, self.h = fmt, w, h
        self.ctx = sws.sws_getContext(w, h, PIX_FMT[fmt], w, h, PIX_FMT["bgr24"],
                                      SWS_BILINEAR, None, None, None)
        if not self.ctx:
            raise RuntimeError("sws_getContext failed")
        self.bps = 2 if fmt == "yuv420p10le" else 1          # bytes per sample
        self.y, self.sy, self._ky = _padded(h, w * self.bps)
        full_c = fmt in ("yuv444p", "yuvj444p")
        ch = h if (full_c or fmt in ("yuv422p", "yuvj422p")) else h // 2
        self.cw = w if full_c else w // 2
        if fmt == "gray":
            self.c0 = self.c1 = None
            self.s0 = 0
        elif fmt in ("nv12", "nv21"):
            self.c0, self.s0, self._k0 = _padded(h // 2, w)
            self.c1 = None
        else:
            self.c0, self.s0, self._k0 = _padded(ch, self.cw * self.bps)
            self.c1, self.s1, self._k1 = _padded(ch, self.cw * self.bps)
        self.dst, self.sd, self._kd = _padded(h, w * 3)
        self._src = (ctypes.c_void_p * 4)(self.y.ctypes.data, self.c0.ctypes.data if self.c0 is not None else None,
                                          self.c1.ctypes.data if self.c1 is not None else None,
                                          None)
        self._sstr = (ctypes.c_int * 4)(self.sy, self.s0,
                                        self.s1 if self.c1 is not None else 0, 0)
        self._dst = (ctypes.c_void_p * 4)(self.dst.ctypes.data, None, None, None)
        self._dstr = (ctypes.c_int * 4)(self.sd, 0, 0, 0)

    def renew_context(self):
        """Free and re-create the SwsContext (PyAV makes a new reformatter for every decoded frame)."""
        self._sws.sws_freeContext(self.ctx)
        self.ctx = self._sws.sws_getContext(self.w, self.h, PIX_FMT[self.fmt], self.w, self.h, PIX_FMT["bgr24"],
                                            SWS_BILINEAR, None, None, None)
        if not self.ctx:
            raise RuntimeError("sws_getContext failed")

    def load(self, planes):
        w, h = self.w, self.h
        as_bytes = lambda p: np.ascontiguousarray(p).view(np.uint8).reshape(p.shape[0], -1)   # u16 samples -> LE bytes
        self.y[:, :w * self.bps] = as_bytes(planes[0])
        if self.fmt == "gray":
            return
        if self.fmt in ("nv12", "nv21"):
            self.c0[:, :w] = planes[1]
        else:
            self.c0[:, :self.cw * self.bps] = as_bytes(planes[1])
            self.c1[:, :self.cw * self.bps] = as_bytes(planes[2])

    def scale(self):
        """sws_scale on the loaded planes; result stays in the padded dst buffer."""
        r = self._sws.sws_scale(self.ctx, self._src, self._sstr, 0, self.h,
                                self._dst, self._dstr)
        if r != self.h:
            raise RuntimeError(f"sws_scale returned {r}")

    def result(self) -> np.ndarray:
        """Contiguous (H, W, 3) copy — PyAV's `useful_array` padding strip."""
        return np.ascontiguousarray(self.dst[:, :self.w * 3]).reshape(self.h, self.w, 3)

    def convert(self, planes) -> np.ndarray:
        self.load(planes)
        self.scale()
        return self.result()

    def close(self):
        if self.ctx:
            self._sws.sws_freeContext(self.ctx)
            self.ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def to_bgr24(fmt: str, planes, w: int, h: int) -> np.ndarray:
    """What `frame.to_ndarray(format='bgr24')` returns for a decoded frame with these planes:
    a fresh context per call (PyAV builds a reformatter per frame), one sws_scale, padding
    stripped.  `planes` are 2-D uint8 arrays: (Y, U, V) for yuv420p/yuvj420p, (Y, UV) for nv12."""
    c = Converter(fmt, w, h)
    try:
        return c.convert(planes)
    finally:
        c.close()
