"""CPU: pins the oracle (oracle/formulas.py, oracle/yuv2bgr_port.c) to the golden vectors that came
from the live libswscale, and the live libswscale itself to SURVEY §8(c)."""
import hashlib

import numpy as np
import pytest

from oracle import formulas as F
from oracle import port as P
from oracle import swscale_live as L
from video_edge_ai_proxy_b200.synth import exhaustive_yuv_frame, make_planes


def _digest(a):
    return int(a.sum(dtype=np.int64)), hashlib.sha256(a.tobytes()).hexdigest()


def test_coefficients_derive_from_itu601_table():
    assert F.coeffs_simd(False) == F.SIMD_LIMITED
    assert F.coeffs_simd(True) == F.SIMD_FULL
    assert F._coeffs_c_path(False) == F.C_LIMITED


def test_gray_known_answers(golden_pixels):
    for yv, want in golden_pixels["gray_kat"].items():
        Y = np.full((2, 16), int(yv), np.uint8)
        C = np.full((1, 8), 128, np.uint8)
        out = F.yuv420p_to_bgr24(Y, C, C)
        assert (out == want).all(), (yv, want)
    # SURVEY §8(c) known answers, stated independently of the fixture
    assert {k: golden_pixels["gray_kat"][k] for k in ("16", "17", "18", "128", "200", "234", "235", "255")} == \
        {"16": 0, "17": 1, "18": 2, "128": 130, "200": 214, "234": 253, "235": 255, "255": 255}


def test_small_fixtures_numpy_and_c_port(small_frames):
    for fmt, w, h, seed, planes, out in small_frames:
        ref = [np.ascontiguousarray(p) for p in make_planes(fmt, w, h, seed)]
        for a, b in zip(ref, planes):
            assert np.array_equal(a, b), "synthetic generator drifted from the fixture"
        assert np.array_equal(F.to_bgr24(fmt, planes), out), (fmt, w, h)
        assert np.array_equal(P.to_bgr24(fmt, planes), out), (fmt, w, h)


@pytest.mark.parametrize("row", range(42))
def test_numpy_restatement_matches_golden_table(golden_pixels, row):
    rows = golden_pixels["rows"]
    if row >= len(rows):
        pytest.skip("no such row")
    r = rows[row]
    planes = make_planes(r["fmt"], r["w"], r["h"], r["seed"])
    if r["path"] == "c":
        if r["fmt"] == "nv12":
            pytest.skip("formula B is restated for planar input only (nv12 uses it for two rows)")
        out = F.yuv420p_to_bgr24_cpath(*planes)
    else:
        out = F.to_bgr24(r["fmt"], planes)
    assert _digest(out) == (r["sum"], r["sha256"]), r


def test_ten_bit_extreme_samples_against_live_library():
    if not L.available():
        pytest.skip("bundled libswscale not found")
    rng = np.random.default_rng(11)
    for w, h in ((64, 8), (130, 18), (640, 480)):
        Y = rng.choice(np.array([0, 1, 2, 3, 64, 940, 1020, 1021, 1022, 1023], np.uint16), (h, w))
        U = rng.choice(np.array([0, 1, 512, 1022, 1023], np.uint16), (h // 2, w // 2))
        V = rng.choice(np.array([0, 3, 512, 1020, 1023], np.uint16), (h // 2, w // 2))
        assert np.array_equal(L.to_bgr24("yuv420p10le", (Y, U, V), w, h), F.to_bgr24("yuv420p10le", (Y, U, V))), (w, h)
    with pytest.raises(ValueError):
        F.to_bgr24("yuv420p10le", make_planes("yuv420p10le", 16, 6, 0))


def test_444_blue_wraparound_is_reproduced():
    """Bright + saturated limited-range 4:4:4 pixels overflow swscale's 32-bit blue sum and come out 0, not 255."""
    Y = np.full((2, 2), 255, np.uint8)
    U = np.full((2, 2), 255, np.uint8)
    V = np.full((2, 2), 128, np.uint8)
    out = F.yuv444p_to_bgr24(Y, U, V, False)
    assert out[0, 0, 0] == 0 and out[0, 0, 2] == 255
    if L.available():
        assert np.array_equal(L.to_bgr24("yuv444p", (Y, U, V), 2, 2), out)
    assert np.array_equal(P.to_bgr24("yuv444p", (Y, U, V)), out)


def test_c_port_matches_numpy_at_config_sizes():
    for fmt, w, h in (("yuv420p", 1920, 1080), ("yuvj420p", 1280, 720), ("nv12", 1920, 1080), ("nv12", 642, 10),
                      ("nv21", 1280, 720), ("yuv422p", 642, 10), ("yuvj422p", 1920, 1080), ("yuv420p10le", 1280, 720),
                      ("yuv444p", 1280, 720), ("yuvj444p", 642, 10)):
        planes = make_planes(fmt, w, h, 3)
        assert np.array_equal(P.to_bgr24(fmt, planes), F.to_bgr24(fmt, planes)), (fmt, w, h)


def test_exhaustive_yuv_triples_cover_and_range():
    Y, U, V = exhaustive_yuv_frame()
    # every (Y,U,V) triple occurs: 256 u x 256 v per tile, 64 tiles x 4 luma values
    assert len(np.unique(Y)) == 256 and len(np.unique(U)) == 256 and len(np.unique(V)) == 256
    out = F.yuv420p_to_bgr24(Y[:512], U[:256], V[:256])
    assert out.shape == (512, 4096, 3)


def test_nv12_small_heights_rejected():
    Y, UV = make_planes("nv12", 16, 6, 0)
    with pytest.raises(ValueError):
        F.nv12_to_bgr24(Y, UV)
    with pytest.raises(ValueError):
        P.to_bgr24("nv12", (Y, UV))


def test_odd_geometry_rejected():
    with pytest.raises(ValueError):
        make_planes("yuv420p", 15, 8, 0)


# ---- the live library (same image on the GPU box) -------------------------------------------------
needs_live = pytest.mark.skipif(not L.available(), reason="bundled libswscale not found")


@needs_live
def test_live_library_is_on_simd_path_and_version(golden_pixels):
    assert L.on_simd_path(), "libswscale would not take its x86 SIMD path on this host"
    assert L.version() == golden_pixels["libswscale"]


@needs_live
@pytest.mark.parametrize("fmt,w,h,seed", [("yuv420p", 640, 480, 0), ("yuvj420p", 640, 480, 0),
                                          ("nv12", 640, 480, 0), ("yuv420p", 1920, 1080, 0)])
def test_live_library_reproduces_survey_checksums(golden_pixels, fmt, w, h, seed):
    row = next(r for r in golden_pixels["rows"] if (r["fmt"], r["w"], r["h"], r["seed"], r["path"]) == (fmt, w, h, seed, "simd"))
    assert row["survey_8c"]
    out = L.to_bgr24(fmt, make_planes(fmt, w, h, seed), w, h)
    assert _digest(out) == (row["sum"], row["sha256"])


@needs_live
def test_live_vs_restatement_ragged_widths_and_heights():
    for fmt in ("yuv420p", "yuvj420p", "nv12", "nv21", "yuv422p", "yuvj422p", "yuv420p10le", "yuv444p", "yuvj444p", "gray"):
        for w in (2, 6, 14, 16, 18, 30, 34, 62, 130, 642):
            for h in (2, 4, 8, 10, 18):
                if fmt in ("nv12", "nv21", "yuv420p10le") and h < 8:
                    continue
                planes = make_planes(fmt, w, h, 1000 * w + h)
                assert np.array_equal(L.to_bgr24(fmt, planes, w, h), F.to_bgr24(fmt, planes)), (fmt, w, h)


@needs_live
def test_live_video_legal_levels():
    for fmt in ("yuv420p", "nv12"):
        planes = make_planes(fmt, 320, 64, 5, dist="video")
        assert np.array_equal(L.to_bgr24(fmt, planes, 320, 64), F.to_bgr24(fmt, planes))


@needs_live
def test_simd_and_c_paths_differ_which_is_why_the_target_is_named():
    planes = make_planes("yuv420p", 64, 16, 9)
    a = F.yuv420p_to_bgr24(*planes)
    b = F.yuv420p_to_bgr24_cpath(*planes)
    d = np.abs(a.astype(int) - b.astype(int))
    assert 0 < d.max() <= 3
