"""GPU: the BASELINE.json configurations as parity cases, at their full sizes, through the C ABI.

config[1] 16 x 1080p one launch per tick        -> test_gpu_parity.py::test_linearity_free_properties_at_full_size
config[2] 64 x 1080p sharded 8 per GPU           -> here (one rank's 8 cameras; the partition itself is tests/test_shard.py)
config[3] 8 x 3840x2160 NV12, key-frame-only     -> here, through FrameProducer + the decode gate
config[4] 128-camera mixed fleet, interval-skip  -> here, heterogeneous descriptor table in one launch per tick
"""
from fractions import Fraction

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

from oracle import port as P  # noqa: E402
from oracle import swscale_live as L  # noqa: E402
from video_edge_ai_proxy_b200 import CameraControl, FramePath, FrameProducer, assign_cameras, make_meta  # noqa: E402
from video_edge_ai_proxy_b200.synth import make_planes  # noqa: E402


def oracle(fmt, planes):
    """Live libswscale when the image has it (it does on the box), else the plain-C restatement."""
    h, w = planes[0].shape
    return L.to_bgr24(fmt, planes, w, h) if L.available() else P.to_bgr24(fmt, planes)


@pytest.fixture()
def path():
    if not torch.cuda.is_available():
        pytest.fail("marked gpu but no CUDA device")
    p = FramePath(0)
    yield p
    p.close()


def test_config_64_cameras_sharded_8_per_gpu(path):
    shard = assign_cameras([(1920, 1080)] * 64, 8)[5]          # the cameras of GPU 5
    assert shard == list(range(40, 48))
    for c in shard:
        path.open_stream(f"cam{c}", "yuv420p", 1920, 1080)
    for tick in range(2):
        for c in shard:
            path.submit(f"cam{c}", make_planes("yuv420p", 1920, 1080, 100 * tick + c), make_meta(packet=tick))
        before = path.stats()["kernel_launches"]
        assert path.tick() == 8
        assert path.stats()["kernel_launches"] == before + 1
        for c in shard:
            assert np.array_equal(path.result(f"cam{c}"), oracle("yuv420p", make_planes("yuv420p", 1920, 1080, 100 * tick + c))), (tick, c)


class _Frame:
    def __init__(self, planes, pts):
        self.planes, self.pts, self.dts = planes, pts, pts
        self.time_base, self.time, self.pict_type = Fraction(1, 90000), pts / 90000, "I" if pts % 30 == 0 else "P"


class _Packet:
    def __init__(self, cam, idx, fmt, w, h):
        self.cam, self.idx, self.fmt, self.w, self.h = cam, idx, fmt, w, h
        self.is_keyframe, self.is_corrupt, self.decoded = idx % 30 == 0, False, 0

    def decode(self):       # synthetic "libavcodec": planes are produced only when the gate asks for them
        self.decoded += 1
        return [_Frame(make_planes(self.fmt, self.w, self.h, 1000 * self.cam + self.idx), self.idx)]


class _Redis:
    def __init__(self):
        self.streams = {}

    def xadd(self, name, fields, maxlen=None):
        s = self.streams.setdefault(name, [])
        s.append(bytes(fields["data"]))
        if maxlen is not None:
            del s[:-maxlen]


def test_config_8_cameras_4k_nv12_keyframe_only(path):
    w, h, ncam, npk = 3840, 2160, 8, 61
    r, now = _Redis(), [5_000_000]
    prod = FrameProducer(path, r, now_ms=lambda: now[0])
    for c in range(ncam):
        prod.add_camera(f"k{c}", "nv12", w, h, memory_buffer=1)
    packets = {c: [_Packet(c, i, "nv12", w, h) for i in range(npk)] for c in range(ncam)}
    published_per_tick = []
    for i in range(npk):
        now[0] += 33
        ctl = CameraControl(last_query=now[0] - 1, key_frame_only=True)    # clients asked for key frames only
        for c in range(ncam):
            prod.on_packet(f"k{c}", packets[c][i], ctl)
        published_per_tick.append(prod.flush())
    st = path.stats()
    # offered: 8 cameras x 61 packets; converted: GOP length 30 -> frames 0, 30, 60 of each camera
    assert sum(published_per_tick) == 3 * ncam and [n for n in published_per_tick if n] == [ncam] * 3
    assert st["frames_converted"] == 3 * ncam and st["kernel_launches"] == 3
    for c in range(ncam):
        assert [p.idx for p in packets[c] if p.decoded] == [0, 30, 60]
        assert len(r.streams[f"k{c}"]) == 1
    for c in (0, 3, 7):      # pixel parity of the newest key frame (a 4K NV12 oracle frame costs ~60 ms)
        want = oracle("nv12", make_planes("nv12", w, h, 1000 * c + 60))
        assert np.array_equal(path.result(f"k{c}"), want), c
        msg = r.streams[f"k{c}"][0]
        assert msg[:11].hex() == "08801e10f0101a80e0ee0b" and len(msg) > w * h * 3     # width 3840, height 2160, data tag + length
        assert msg[11:11 + 64] == want.tobytes()[:64]


def test_config_128_camera_mixed_fleet_interval_skip(path):
    sizes = [(640, 480)] * 43 + [(1280, 720)] * 43 + [(1920, 1080)] * 42
    # this rank's share when the fleet is spread over 8 GPUs, pixel-balanced ...
    shard = assign_cameras(sizes, 8)[2]
    assert 14 <= len(shard) <= 18
    # ... and the whole fleet on ONE GPU for the heterogeneous-table test proper
    for c, (w, h) in enumerate(sizes):
        path.open_stream(f"m{c}", "yuv420p", w, h)
    total = 0
    for tick in range(3):
        sel = np.random.default_rng(tick).random(len(sizes)) < 0.25       # seeded Bernoulli(0.25), seed = tick index
        for c, (w, h) in enumerate(sizes):
            if sel[c]:
                path.submit(f"m{c}", make_planes("yuv420p", w, h, 7 * tick + c), make_meta(packet=tick))
            else:
                path.submit(f"m{c}", None, None, convert=False)               # gate-skipped: counted, never copied
        before = path.stats()
        n = path.tick()
        after = path.stats()
        assert n == int(sel.sum()) and after["kernel_launches"] == before["kernel_launches"] + 1
        assert len({sizes[c] for c in np.nonzero(sel)[0]}) == 3              # all three resolutions in the one launch
        for c in np.nonzero(sel)[0]:
            w, h = sizes[c]
            assert np.array_equal(path.result(f"m{c}"), P.to_bgr24("yuv420p", make_planes("yuv420p", w, h, 7 * tick + c))), (tick, c)
        total += n
    st = path.stats()
    assert st["frames_offered"] == 3 * len(sizes) and st["frames_converted"] == total
