"""TEST INFRASTRUCTURE ONLY — numpy restatement of the arithmetic behind reference
`python/read_image.py:94` (`frame.to_ndarray(format='bgr24')`).

That call lands in libswscale (via PyAV `av=8.0.2`, `python/environment.yml:10`), which is not in
the reference tree, so the algorithm is restated here from its published behaviour and pinned
against the live library (`oracle/swscale_live.py`) and the golden table in SURVEY.md §8(c) by
`tests/test_oracle.py`.  Parity status: the reference itself holds no golden vector or test for
this path ("parity unpinned" by the reference's own tests); the pin is the live libswscale
9.1.100 run in this image plus the fixtures under `tests/golden/` generated from it.

Three closed forms (all integer, `>>` is an arithmetic/floor shift):

  A  planar yuv420p / yuvj420p -> BGR24, x86 SIMD path (what PyAV produces on x86-64): 13-bit
     coefficients, int16 `pmulhw` products floored separately, nearest chroma, saturating pack.
  B  same conversion on swscale's plain-C table path (`av_force_cpu_flags(0)`); differs from A
     by up to 3 LSB.  Needed because C below falls back to it for two rows.
  C  nv12 -> BGR24: generic vertical scaler in front of the same SIMD output stage; even rows
     average the chroma row above with the current one, the last two rows use B's arithmetic on
     a 12-bit blended chroma.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s CPU-baseline leg may import this.
"""
from __future__ import annotations

import numpy as np

# ---------------------------------------------------------------------------------------------
# Coefficients.  swscale starts from the ITU-601 inverse table {crv, cbu, cgu, cgv} =
# {104597, 132201, 25675, 53279} (16.16 fixed point) and, for limited-range sources, the luma
# gain cy = 65536*255/219 with offset oy = 16<<16; for full-range (yuvj) sources cy = 65536,
# oy = 0 and the chroma gains are scaled by 224/255.  The x86 path then keeps 13 fractional
# bits: coeff = clip_int16((c << 13 + 32768) >> 16) and y offset = (oy << 3 + 32768) >> 16.
# ---------------------------------------------------------------------------------------------
_INV_601 = (104597, 132201, 25675, 53279)


def _round16(v: int) -> int:
    return (v + 32768) >> 16


def _tdiv(a: int, b: int) -> int:
    """C integer division (truncates toward zero)."""
    q = abs(a) // abs(b)
    return q if (a >= 0) == (b >= 0) else -q


def _coeffs_raw(full_range: bool):
    """16.16 gains after swscale's range handling (contrast = saturation = 1.0, brightness 0)."""
    crv, cbu, cgu, cgv = _INV_601
    cgu, cgv = -cgu, -cgv
    if full_range:
        cy, oy = 1 << 16, 0
        crv = _tdiv(crv * 224, 255)
        cbu = _tdiv(cbu * 224, 255)
        cgu = _tdiv(cgu * 224, 255)
        cgv = _tdiv(cgv * 224, 255)
    else:
        cy, oy = (65536 * 255) // 219, 16 << 16
    return dict(cy=cy, oy=oy, crv=crv, cbu=cbu, cgu=cgu, cgv=cgv)


def _coeffs_c_path(full_range: bool):
    """Coefficients of the plain-C table path (formula B): the chroma gains are divided by the
    luma gain because the table lookup applies cy to (Y + chroma offset) as a whole."""
    c = dict(_coeffs_raw(full_range))
    cy = max(c["cy"], 1)
    for n in ("crv", "cbu", "cgu", "cgv"):
        c[n] = _tdiv(c[n] * 65536 + 0x8000, cy)
    return c


def coeffs_simd(full_range: bool = False):
    """13-bit coefficient set of the x86 SIMD path (formula A)."""
    c = _coeffs_raw(full_range)
    clip16 = lambda v: max(-32768, min(32767, v))
    return dict(
        yC=clip16(_round16(c["cy"] << 13)),
        vr=clip16(_round16(c["crv"] << 13)),
        ub=clip16(_round16(c["cbu"] << 13)),
        ug=clip16(_round16(c["cgu"] << 13)),
        vg=clip16(_round16(c["cgv"] << 13)),
        yOff=_round16(c["oy"] << 3),
    )


# SURVEY.md §8(c) states these values; test_oracle pins coeffs_simd() to them.
SIMD_LIMITED = dict(yC=9539, vr=13075, ub=16525, ug=-3209, vg=-6660, yOff=128)
SIMD_FULL = dict(yC=8192, vr=11485, ub=14516, ug=-2819, vg=-5850, yOff=0)
C_LIMITED = dict(cy=76309, oy=1048576, crv=89830, cbu=113537, cgu=-22049, cgv=-45756)


def _sat_u8(a):
    return np.clip(a, 0, 255).astype(np.uint8)


def _up2(c, h, w):
    """Nearest-neighbour 2x2 replication of a chroma plane."""
    return np.repeat(np.repeat(c, 2, axis=0), 2, axis=1)[:h, :w]


def _simd_output_stage(Y, Uw, Vw, k):
    """Shared tail of A and C: Y (H,W) u8; Uw,Vw (H,W) int32 already `(u<<3)-1024` style."""
    Y = Y.astype(np.int32)
    y = (((Y << 3) - k["yOff"]) * k["yC"]) >> 16
    cb = (Uw * k["ub"]) >> 16
    cr = (Vw * k["vr"]) >> 16
    cg = ((Uw * k["ug"]) >> 16) + ((Vw * k["vg"]) >> 16)
    out = np.empty(Y.shape + (3,), np.uint8)
    out[..., 0] = _sat_u8(y + cb)
    out[..., 1] = _sat_u8(y + cg)
    out[..., 2] = _sat_u8(y + cr)
    return out


def yuv420p_to_bgr24(Y, U, V, full_range: bool = False) -> np.ndarray:
    """Formula A.  Y (H,W), U/V (H/2,W/2) uint8 -> (H,W,3) uint8 in B,G,R order."""
    h, w = Y.shape
    k = coeffs_simd(full_range)
    Uw = (_up2(U, h, w).astype(np.int32) << 3) - 1024
    Vw = (_up2(V, h, w).astype(np.int32) << 3) - 1024
    return _simd_output_stage(Y, Uw, Vw, k)


def _c_path_pixels(Y, u, v, c):
    """Formula B on per-pixel (already upsampled) chroma u, v (int32 arrays shaped like Y)."""
    Y = Y.astype(np.int32)
    cy = c["cy"]
    base = -(384 << 16) - c["oy"] + 326 * cy + 0x8000

    def lut(i):
        return _sat_u8((base + i * cy) >> 16)

    dr = ((v * c["crv"]) >> 16) - (c["crv"] >> 9)
    db = ((u * c["cbu"]) >> 16) - (c["cbu"] >> 9)
    dg = ((u * c["cgu"]) >> 16) - (c["cgu"] >> 9) + ((v * c["cgv"]) >> 16) - (c["cgv"] >> 9)
    out = np.empty(Y.shape + (3,), np.uint8)
    out[..., 0] = lut(Y + db)
    out[..., 1] = lut(Y + dg)
    out[..., 2] = lut(Y + dr)
    return out


def yuv420p_to_bgr24_cpath(Y, U, V, full_range: bool = False) -> np.ndarray:
    """Formula B (secondary: libswscale with SIMD disabled)."""
    h, w = Y.shape
    c = _coeffs_c_path(full_range)
    return _c_path_pixels(Y, _up2(U, h, w).astype(np.int32), _up2(V, h, w).astype(np.int32), c)


def nv12_to_bgr24(Y, UV) -> np.ndarray:
    """Formula C.  Y (H,W), UV (H/2,W) interleaved U,V uint8 -> (H,W,3) uint8 B,G,R."""
    h, w = Y.shape
    if h < 8:
        # libswscale shrinks the vertical chroma filter to nearest for H = 4, 6 (probed against
        # the live library); the reference never meets such frames and the C-ABI rejects them.
        raise ValueError("nv12 restatement covers H >= 8 only")
    hc = h // 2
    U = UV[:, 0::2].astype(np.int32)
    V = UV[:, 1::2].astype(np.int32)
    k = coeffs_simd(False)
    out = np.empty((h, w, 3), np.uint8)

    # rows 0 .. H-3 : SIMD output stage on vertically filtered chroma
    kk = np.arange(hc)
    up = np.maximum(kk - 1, 0)
    # odd rows: chroma row k as is
    Uw_odd = (U << 3) - 1024
    Vw_odd = (V << 3) - 1024
    # even rows: sum of row k-1 (clamped at the top) and row k, two fractional bits instead of three
    Uw_even = ((U[up] + U) << 2) - 1024
    Vw_even = ((V[up] + V) << 2) - 1024
    rep = lambda a: np.repeat(a, 2, axis=1)[:, :w]
    out[0::2] = _simd_output_stage(Y[0::2], rep(Uw_even), rep(Vw_even), k)
    out[1::2] = _simd_output_stage(Y[1::2], rep(Uw_odd), rep(Vw_odd), k)

    # rows H-2, H-1 : plain-C output stage on 12-bit blended chroma
    c = _coeffs_c_path(False)
    for j in (h - 2, h - 1):
        kr = j >> 1
        if j & 1:
            k0, k1, a = kr, min(kr + 1, hc - 1), 1024
        else:
            k0, k1, a = kr - 1, kr, 3072
        k0 = max(k0, 0)
        if k0 == k1:
            a = 0
        u = (U[k0] * (4096 - a) + U[k1] * a + 2048) >> 12
        v = (V[k0] * (4096 - a) + V[k1] * a + 2048) >> 12
        out[j] = _c_path_pixels(Y[j], np.repeat(u, 2)[:w], np.repeat(v, 2)[:w], c)
    return out


def yuv422p_to_bgr24(Y, U, V, full_range: bool = False) -> np.ndarray:
    """4:2:2 planar (MJPEG cameras decode to yuvj422p).  Probed against the live library: libswscale keeps
    every chroma row — luma row j uses chroma row j, horizontally nearest — and feeds the same SIMD
    output stage as formula A (all rows, including the last two).  U/V are (H, W/2)."""
    h, w = Y.shape
    k = coeffs_simd(full_range)
    rep = lambda a: np.repeat(a, 2, axis=1)[:, :w]
    return _simd_output_stage(Y, (rep(U).astype(np.int32) << 3) - 1024, (rep(V).astype(np.int32) << 3) - 1024, k)


def nv21_to_bgr24(Y, VU) -> np.ndarray:
    """NV21 = NV12 with V first in each chroma pair; identical arithmetic (probed)."""
    UV = np.empty_like(VU)
    UV[:, 0::2] = VU[:, 1::2]
    UV[:, 1::2] = VU[:, 0::2]
    return nv12_to_bgr24(Y, UV)


def yuv420p10_to_bgr24(Y, U, V) -> np.ndarray:
    """yuv420p10le (10-bit samples, limited range) -> BGR24.  Probed against the live library, exact for H >= 8
    on uniform and extreme sample values: libswscale sends 10-bit input through its generic vertical scaler (15-bit
    intermediates = sample << 5), so the rule is formula C's with 11-bit terms: odd rows use chroma row k
    (`(u << 1) - 1024`), even rows the sum with the row above (`(u_up + u) - 1024`), luma `(Y << 1)`; rows H-2 and
    H-1 drop to the plain-C stage on 8-bit values y = (Y + 2) >> 2, u = min((blend + 8192) >> 14, 255)."""
    h, w = Y.shape
    if h < 8:
        raise ValueError("yuv420p10le restatement covers H >= 8 only")
    hc = h // 2
    Yi, Ui, Vi = Y.astype(np.int32), U.astype(np.int32), V.astype(np.int32)
    k = coeffs_simd(False)
    rep = lambda a: np.repeat(a, 2, axis=1)[:, :w]
    out = np.empty((h, w, 3), np.uint8)

    def stage(Yr, Uw, Vw):
        y = (((Yr << 1) - k["yOff"]) * k["yC"]) >> 16
        cb = (Uw * k["ub"]) >> 16
        cr = (Vw * k["vr"]) >> 16
        cg = ((Uw * k["ug"]) >> 16) + ((Vw * k["vg"]) >> 16)
        return np.stack([_sat_u8(y + cb), _sat_u8(y + cg), _sat_u8(y + cr)], -1)

    up = np.maximum(np.arange(hc) - 1, 0)
    out[1::2] = stage(Yi[1::2], rep((Ui << 1) - 1024), rep((Vi << 1) - 1024))
    out[0::2] = stage(Yi[0::2], rep((Ui[up] + Ui) - 1024), rep((Vi[up] + Vi) - 1024))
    c = _coeffs_c_path(False)
    for j in (h - 2, h - 1):
        kr = j >> 1
        if j & 1:
            k0, k1, a = kr, min(kr + 1, hc - 1), 1024
        else:
            k0, k1, a = kr - 1, kr, 3072
        if k0 == k1:
            a = 0
        u = np.minimum((Ui[k0] * (4096 - a) + Ui[k1] * a + 8192) >> 14, 255)   # the C tables clip the chroma index to 8 bits
        v = np.minimum((Vi[k0] * (4096 - a) + Vi[k1] * a + 8192) >> 14, 255)
        out[j] = _c_path_pixels((Yi[j] + 2) >> 2, np.repeat(u, 2)[:w], np.repeat(v, 2)[:w], c)
    return out


def yuv444p_to_bgr24(Y, U, V, full_range: bool = False) -> np.ndarray:
    """4:4:4 planar.  libswscale turns on full chroma interpolation for it and runs its plain-C `yuv2rgb_full`
    stage: per pixel, in wrapping 32-bit arithmetic with the SAME 13-bit coefficients as formula A,
        Yt = ((Y << 9) - (yOff << 6)) * yC + (1 << 21)
        B = Yt + ((U - 128) << 9) * ub,  G = Yt + ((U-128) << 9) * ug + ((V-128) << 9) * vg,  R = Yt + ((V-128) << 9) * vr
    each clipped to [0, 2^30) (negative after wrap -> 0) and shifted right by 22.  One rounding per channel instead
    of formula A's three floors; the limited-range blue term can wrap past 2^31 for bright saturated pixels and then
    comes out 0 — reproduced, since the bar is the library's bytes.  Probed exact on uniform and extreme inputs."""
    k = coeffs_simd(full_range)
    Yi, Ui, Vi = Y.astype(np.int64), U.astype(np.int64), V.astype(np.int64)
    Yt = ((Yi << 9) - (k["yOff"] << 6)) * k["yC"] + (1 << 21)
    Uc, Vc = (Ui - 128) << 9, (Vi - 128) << 9
    wrap = lambda a: ((a + (1 << 31)) % (1 << 32)) - (1 << 31)
    fin = lambda a: np.clip(wrap(a) >> 22, 0, 255).astype(np.uint8)
    return np.stack([fin(Yt + Uc * k["ub"]), fin(Yt + Uc * k["ug"] + Vc * k["vg"]), fin(Yt + Vc * k["vr"])], -1)


def gray_to_bgr24(Y) -> np.ndarray:
    """gray (monochrome decode): libswscale treats the plane as full-range luma and replicates it, B = G = R = Y (probed)."""
    return np.repeat(Y[..., None], 3, axis=2)


def to_bgr24(fmt: str, planes) -> np.ndarray:
    if fmt == "gray":
        return gray_to_bgr24(planes[0])
    if fmt == "yuv444p":
        return yuv444p_to_bgr24(planes[0], planes[1], planes[2], False)
    if fmt == "yuvj444p":
        return yuv444p_to_bgr24(planes[0], planes[1], planes[2], True)
    if fmt == "yuv420p10le":
        return yuv420p10_to_bgr24(planes[0], planes[1], planes[2])
    if fmt == "yuv422p":
        return yuv422p_to_bgr24(planes[0], planes[1], planes[2], False)
    if fmt == "yuvj422p":
        return yuv422p_to_bgr24(planes[0], planes[1], planes[2], True)
    if fmt == "nv21":
        return nv21_to_bgr24(planes[0], planes[1])
    if fmt == "yuv420p":
        return yuv420p_to_bgr24(planes[0], planes[1], planes[2], False)
    if fmt == "yuvj420p":
        return yuv420p_to_bgr24(planes[0], planes[1], planes[2], True)
    if fmt == "nv12":
        return nv12_to_bgr24(planes[0], planes[1])
    raise ValueError(fmt)
