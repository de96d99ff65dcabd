"""GPU: host pipeline through the C ABI — pinned staging, tick, result ring, packed VideoFrame,
producer loop with the decode gate and a fake Redis."""
import threading
from fractions import Fraction

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

from oracle import formulas as F  # noqa: E402
from oracle import videoframe_wire as W  # noqa: E402
from video_edge_ai_proxy_b200 import (CameraControl, FramePath, FrameProducer, VepError, make_meta,  # noqa: E402
                                      to_ndarray_bgr24)
from video_edge_ai_proxy_b200 import _abi  # noqa: E402
from video_edge_ai_proxy_b200.synth import make_planes  # noqa: E402


@pytest.fixture()
def path():
    if not torch.cuda.is_available():
        pytest.fail("marked gpu but no CUDA device")
    p = FramePath(0)
    yield p
    p.close()


def test_to_ndarray_drop_in(path):
    for fmt, w, h in (("yuv420p", 640, 480), ("yuvj420p", 1280, 720), ("nv12", 1920, 1080), ("yuv420p", 50, 6),
                      ("yuvj422p", 1280, 720), ("yuv422p", 50, 6), ("nv21", 640, 480)):
        planes = make_planes(fmt, w, h, 3)
        img = to_ndarray_bgr24(path, fmt, planes)
        assert img.dtype == np.uint8 and img.shape == (h, w, 3) and img.flags["C_CONTIGUOUS"]
        assert np.array_equal(img, F.to_bgr24(fmt, planes))


def test_submit_with_decoder_strides(path):
    """AVFrame planes arrive with linesize > width; the staging copy strips it."""
    w, h = 1920, 1080
    Y, U, V = make_planes("yuv420p", w, h, 8)
    big = [np.zeros((p.shape[0], p.shape[1] + 64), np.uint8) for p in (Y, U, V)]
    for b, p in zip(big, (Y, U, V)):
        b[:, :p.shape[1]] = p
        b[:, p.shape[1]:] = 0xEE
    path.open_stream("cam", "yuv420p", w, h)
    path.submit("cam", [b[:, :p.shape[1]] for b, p in zip(big, (Y, U, V))])
    assert path.tick() == 1
    assert np.array_equal(path.result("cam"), F.to_bgr24("yuv420p", (Y, U, V)))


def test_tick_batches_many_cameras_in_one_launch(path):
    cams = [(f"c{i}", fmt, w, h, i) for i, (fmt, w, h) in enumerate(
        [("yuv420p", 1920, 1080)] * 4 + [("nv12", 1280, 720)] * 3 + [("yuvj420p", 640, 480)] * 3 + [("yuv420p", 642, 482)])]
    for name, fmt, w, h, seed in cams:
        path.open_stream(name, fmt, w, h)
        path.submit(name, make_planes(fmt, w, h, seed))
    s0 = path.stats()
    assert path.tick() == len(cams)
    s1 = path.stats()
    assert s1["kernel_launches"] - s0["kernel_launches"] == 1
    assert s1["frames_converted"] - s0["frames_converted"] == len(cams)
    assert s1["h2d_bytes"] > 0 and s1["d2h_bytes"] == sum(w * h * 3 for _, _, w, h, _ in cams)
    for name, fmt, w, h, seed in cams:
        assert np.array_equal(path.result(name), F.to_bgr24(fmt, make_planes(fmt, w, h, seed))), name
    assert path.tick() == 0      # nothing staged -> no launch
    assert path.stats()["kernel_launches"] == s1["kernel_launches"]


def test_packed_videoframe_is_the_reference_message(path):
    w, h = 640, 480
    planes = make_planes("yuv420p", w, h, 21)
    fields = dict(timestamp=1604660000123, pts=-3003, dts=897000, packet=4, keyframe=17, time_base=1 / 90000,
                  is_keyframe=True, is_corrupt=False, frame_type="I")
    path.open_stream("cam", "yuv420p", w, h)
    path.submit("cam", planes, make_meta(**fields))
    path.tick()
    msg = bytes(path.result_packed("cam"))
    bgr = F.to_bgr24("yuv420p", planes)
    want = W.serialize_videoframe(width=w, height=h, data=bgr.tobytes(), shape=(h, w, 3), **fields)
    assert msg == want
    assert msg == W.serialize_with_protobuf_runtime(width=w, height=h, data=bgr.tobytes(), shape=(h, w, 3), **fields)
    # consumer side, examples/opencv_display.py:48-53
    VideoFrame, _ = W.dynamic_message_class()
    vf = VideoFrame()
    vf.ParseFromString(msg)
    img = np.frombuffer(vf.data, dtype=np.uint8).reshape([d.size for d in vf.shape.dim])
    assert np.array_equal(img, bgr) and [d.name for d in vf.shape.dim] == ["0", "1", "2"] and vf.device_id == ""


def test_result_ring_semantics(path):
    w, h = 64, 16
    path.open_stream("cam", "yuv420p", w, h, memory_buffer=3)
    with pytest.raises(VepError) as e:
        path.result("cam")
    assert e.value.status == _abi.VEP_ERR_NO_FRAME
    views = []
    for s in range(3):
        path.submit("cam", make_planes("yuv420p", w, h, s), make_meta(packet=s))
        path.tick()
        views.append(path.result("cam"))
    # the three newest results are all still valid and distinct buffers
    for s in range(3):
        assert np.array_equal(views[s], F.to_bgr24("yuv420p", make_planes("yuv420p", w, h, s)))
        assert np.array_equal(path.result("cam", age=2 - s), views[s])
    # a burst of three frames in one tick (catch-up, read_image.py:81-85): all retrievable, oldest first
    for s in (10, 11, 12):
        path.submit("cam", make_planes("yuv420p", w, h, s), make_meta(packet=s))
    assert path.tick() == 3
    for age, s in ((2, 10), (1, 11), (0, 12)):
        assert np.array_equal(path.result("cam", age=age), F.to_bgr24("yuv420p", make_planes("yuv420p", w, h, s)))
    # producer outrunning the ticks is refused, not overwritten
    with pytest.raises(VepError) as e:
        for s in range(20):
            path.submit("cam", make_planes("yuv420p", w, h, s))
    assert e.value.status == _abi.VEP_ERR_BUSY
    path.tick()


def test_two_ticks_in_flight_and_zero_copy_staging(path):
    w, h = 1280, 720
    path.open_stream("cam", "nv12", w, h)
    frames = [make_planes("nv12", w, h, s) for s in range(4)]
    got = []
    for i, pl in enumerate(frames):
        dst = path.acquire_input("cam")
        for d, p in zip(dst, pl):
            d[...] = p
        path.commit_input("cam", make_meta(packet=i))
        assert path.tick_begin() == 1
        if i >= 1:
            path.tick_wait()
            got.append(path.result("cam").copy())
    path.tick_wait()
    got.append(path.result("cam").copy())
    for g, pl in zip(got, frames):
        assert np.array_equal(g, F.to_bgr24("nv12", pl))


def test_bounded_ticks_in_flight(path):
    w, h = 64, 16
    path.open_stream("cam", "yuv420p", w, h)
    for s in range(4):          # VEP_MAX_TICKS_IN_FLIGHT
        path.submit("cam", make_planes("yuv420p", w, h, s))
        assert path.tick_begin() == 1
    path.submit("cam", make_planes("yuv420p", w, h, 4))
    with pytest.raises(VepError) as e:
        path.tick_begin()
    assert e.value.status == _abi.VEP_ERR_BUSY
    with pytest.raises(VepError) as e:
        path.close_stream("cam")          # refused while ticks are in flight
    assert e.value.status == _abi.VEP_ERR_BUSY
    for _ in range(4):
        path.tick_wait()
    assert path.tick() == 1
    assert np.array_equal(path.result("cam"), F.to_bgr24("yuv420p", make_planes("yuv420p", w, h, 4)))
    assert path.tick_begin() == 0         # an empty tick is not counted as in flight
    path.tick_wait()


def test_concurrent_producers_one_thread_per_camera(path):
    n, w, h = 8, 1280, 720
    for i in range(n):
        path.open_stream(f"c{i}", "yuv420p", w, h)
    planes = [make_planes("yuv420p", w, h, 100 + i) for i in range(n)]
    errs = []

    def prod(i):
        try:
            path.submit(f"c{i}", planes[i], make_meta(packet=i))
        except Exception as e:  # pragma: no cover
            errs.append(e)

    th = [threading.Thread(target=prod, args=(i,)) for i in range(n)]
    [t.start() for t in th]
    [t.join() for t in th]
    assert not errs
    assert path.tick() == n
    for i in range(n):
        assert np.array_equal(path.result(f"c{i}"), F.to_bgr24("yuv420p", planes[i]))


def test_device_result_is_zero_copy_torch_tensor(path):
    w, h = 640, 480
    planes = make_planes("yuv420p", w, h, 5)
    path.open_stream("cam", "yuv420p", w, h)
    path.submit("cam", planes)
    path.tick()
    t = path.result_tensor("cam")
    assert t.is_cuda and t.dtype == torch.uint8 and tuple(t.shape) == (h, w, 3)
    assert np.array_equal(t.cpu().numpy(), F.to_bgr24("yuv420p", planes))


def test_stream_errors(path):
    with pytest.raises(VepError):
        path.open_stream("x", "yuv420p", 641, 480)
    with pytest.raises(VepError):
        path.open_stream("x", "nv12", 640, 6)
    with pytest.raises(VepError):
        path.open_stream("x", "rgb24", 640, 480)
    path.open_stream("x", "yuv420p", 640, 480)
    with pytest.raises(VepError):
        path.open_stream("x", "yuv420p", 640, 480)
    with pytest.raises(VepError):
        path.submit("nope", make_planes("yuv420p", 640, 480, 0))
    path.close_stream("x")


# ---- producer loop: gate + conversion + XADD against a fake Redis -----------------------------------
class FakeRedis:
    def __init__(self):
        self.streams = {}

    def xadd(self, name, fields, maxlen=None):
        s = self.streams.setdefault(name, [])
        s.append(bytes(fields["data"]))       # Redis copies the payload off the socket
        if maxlen is not None:
            del s[:-maxlen]


class FakeFrame:
    def __init__(self, planes, pts):
        self.planes, self.pts, self.dts = planes, pts, pts - 1
        self.width, self.height = planes[0].shape[1], planes[0].shape[0]
        self.time_base = Fraction(1, 90000)
        self.time = pts / 90000
        self.pict_type = "I" if pts % 30 == 0 else "P"


class FakePacket:
    def __init__(self, idx, planes):
        self.is_keyframe, self.is_corrupt, self.idx, self._planes = idx % 30 == 0, False, idx, planes
        self.decoded = 0

    def decode(self):
        self.decoded += 1
        return [FakeFrame(self._planes, self.idx)]


def test_producer_keyframe_only_and_interval_skip(path):
    r = FakeRedis()
    now = [1_000_000]
    prod = FrameProducer(path, r, now_ms=lambda: now[0])
    w, h = 640, 480
    prod.add_camera("kf", "yuv420p", w, h, memory_buffer=1)     # key-frame-only client
    prod.add_camera("live", "yuv420p", w, h, memory_buffer=4)   # client polling every 10 packets
    prod.add_camera("idle", "yuv420p", w, h)                    # nobody asked for >10 s
    planes = {i: make_planes("yuv420p", w, h, i % 5) for i in range(60)}
    pk = {c: [FakePacket(i, planes[i]) for i in range(60)] for c in ("kf", "live", "idle")}
    for i in range(60):
        now[0] += 40
        ctl = {"kf": CameraControl(last_query=now[0] - 5, key_frame_only=True),
               "live": CameraControl(last_query=now[0] - ((i + 1) % 10) * 40 - 1),   # fresh query at i = 9, 19, ...
               "idle": CameraControl(last_query=now[0] - 60_000)}
        for c in ("kf", "live", "idle"):
            prod.on_packet(c, pk[c][i], ctl[c])
        prod.flush()
    VideoFrame, _ = W.dynamic_message_class()

    def parse(b):
        vf = VideoFrame()
        vf.ParseFromString(b)
        return vf

    # key-frame-only: exactly the two keyframes were decoded; MAXLEN 1 keeps the newest
    assert [p.idx for p in pk["kf"] if p.decoded] == [0, 30]
    assert len(r.streams["kf"]) == 1
    vf = parse(r.streams["kf"][0])
    assert (vf.pts, vf.packet, vf.keyframe, vf.is_keyframe, vf.frame_type) == (30, 0, 2, True, "I")
    assert np.array_equal(np.frombuffer(vf.data, np.uint8).reshape(h, w, 3), F.to_bgr24("yuv420p", planes[30]))
    # live camera: every packet decoded exactly once (bursts catch up), ring keeps the newest 4
    assert all(p.decoded == 1 for p in pk["live"])
    assert [parse(b).pts for b in r.streams["live"]] == [56, 57, 58, 59]
    last = parse(r.streams["live"][-1])
    assert last.packet == 29 and last.keyframe == 2 and not last.is_keyframe
    assert last.timestamp == int(59 / 90000 * 90000)
    # idle camera: nothing decoded, nothing published
    assert not any(p.decoded for p in pk["idle"]) and "idle" not in r.streams


def test_producers_and_ticker_run_concurrently(path):
    """Eight producer threads keep submitting while the main thread ticks: every published frame is a
    correct conversion of some frame its camera submitted, in submission order (no tearing, no mix-ups)."""
    import threading
    n, w, h, per = 8, 640, 480, 12
    want = {}
    for c in range(n):
        path.open_stream(f"t{c}", "yuv420p", w, h, memory_buffer=2)
        for k in range(per):
            want[(c, k)] = F.to_bgr24("yuv420p", make_planes("yuv420p", w, h, 100 * c + k))
    stop, errs = threading.Event(), []

    def producer(c):
        try:
            k = 0
            while k < per:
                try:
                    path.submit(f"t{c}", make_planes("yuv420p", w, h, 100 * c + k), make_meta(packet=k))
                    k += 1
                except VepError as e:
                    if e.status != _abi.VEP_ERR_BUSY:
                        raise
        except Exception as e:  # pragma: no cover
            errs.append(e)

    th = [threading.Thread(target=producer, args=(c,)) for c in range(n)]
    [t.start() for t in th]
    seen = {c: -1 for c in range(n)}
    VideoFrame, _ = W.dynamic_message_class()
    total = 0
    while any(t.is_alive() for t in th) or total < n * per:
        got = path.tick()
        total += got
        for c in range(n):
            try:
                msg = bytes(path.result_packed(f"t{c}"))
            except VepError:
                continue
            vf = VideoFrame(); vf.ParseFromString(msg)
            k = int(vf.packet)
            assert k >= seen[c]
            seen[c] = k
            assert np.array_equal(np.frombuffer(vf.data, np.uint8).reshape(h, w, 3), want[(c, k)]), (c, k)
        if not got and not any(t.is_alive() for t in th):
            break
    [t.join() for t in th]
    assert not errs and total == n * per and all(v == per - 1 for v in seen.values())


# ---- drop-in ReadImage: driven exactly the way reference rtsp_to_rtmp.py:191-251 drives read_image.ReadImage ----
class _AvPlane:
    """Stand-in for av.VideoPlane: buffer protocol + line_size/width/height, rows padded like libavcodec's."""
    def __init__(self, arr, pad=32):
        self.height, self.width = arr.shape
        self.line_size = self.width + pad
        buf = np.full((self.height, self.line_size), 0xEE, np.uint8)
        buf[:, :self.width] = arr
        self._b = buf.tobytes()

    def __buffer__(self, flags):
        return memoryview(self._b)


class _AvFormat:
    def __init__(self, name):
        self.name = name


class _AvPict:
    def __init__(self, name):
        self.name = name


class _AvFrame:
    def __init__(self, fmt, planes, pts):
        self.format, self.planes = _AvFormat(fmt), [_AvPlane(p) for p in planes]
        self.width, self.height = planes[0].shape[1], planes[0].shape[0]
        self.pts, self.dts, self.time, self.time_base = pts, pts, pts / 90000, Fraction(1, 90000)
        self.pict_type = _AvPict("I" if pts % 5 == 0 else "P")


class _AvPacket:
    def __init__(self, idx, fmt, w, h):
        self.idx, self.is_keyframe, self.is_corrupt = idx, idx % 5 == 0, False
        self._args = (fmt, w, h)

    def decode(self):
        fmt, w, h = self._args
        return [_AvFrame(fmt, make_planes(fmt, w, h, 500 + self.idx), self.idx)]


class _RedisKV(FakeRedis):
    def __init__(self, kv):
        super().__init__()
        self.kv = kv

    def get(self, k):
        return self.kv.get(k)


def test_read_image_drop_in_thread():
    import queue
    import threading
    import time as _time
    from video_edge_ai_proxy_b200 import global_vars
    from video_edge_ai_proxy_b200.read_image import ReadImage
    if not torch.cuda.is_available():
        pytest.fail("marked gpu but no CUDA device")
    fmt, w, h = "yuvj420p", 640, 480
    r = _RedisKV({"is_key_frame_only_camA": b"false"})
    packet_queue, ev, cond = queue.Queue(), threading.Event(), threading.Condition()
    th = ReadImage(packet_queue=packet_queue, device_id="camA", memory_buffer=3, redis_conn=r,
                   is_decode_packets_event=ev, lock_condition=cond)          # reference argument names (rtsp_to_rtmp.py:239-244)
    th.daemon = True
    th.start()
    global_vars.query_timestamp = None
    try:
        for i in range(10):                                                   # producer side, rtsp_to_rtmp.py:141-160
            with cond:
                global_vars.query_timestamp = 1000 + i
                cond.notify_all()
            ev.set()
            packet_queue.put(_AvPacket(i, fmt, w, h))
            with cond:
                cond.notify_all()
            packet_queue.join()
        deadline = _time.time() + 10
        while len(r.streams.get("camA", [])) < 3 and _time.time() < deadline:
            _time.sleep(0.01)
    finally:
        th.stop()
        th.join(timeout=10)
    msgs = r.streams["camA"]
    assert len(msgs) == 3                                                     # MAXLEN = memory_buffer
    VideoFrame, _ = W.dynamic_message_class()
    for msg, idx in zip(msgs, (7, 8, 9)):
        vf = VideoFrame(); vf.ParseFromString(msg)
        assert (vf.width, vf.height, vf.pts, vf.packet, vf.keyframe) == (w, h, idx, idx % 5, idx // 5 + 1)
        assert vf.frame_type == ("I" if idx % 5 == 0 else "P") and vf.is_keyframe == (idx % 5 == 0)
        img = np.frombuffer(vf.data, np.uint8).reshape([d.size for d in vf.shape.dim])
        assert np.array_equal(img, F.to_bgr24(fmt, make_planes(fmt, w, h, 500 + idx)))
    # a bad frame is dropped with the reference's message and the thread survives (read_image.py:130-131)


def test_ragged_widths_through_the_pipeline(path):
    """854 x 480, 1366 x 768 and friends: the staged planes and the device output get padded rows (vector path),
    the copy back to the host removes the padding — results are the tight (H, W, 3) arrays and messages."""
    cams = [("a", "yuv420p", 854, 480), ("b", "nv12", 1366, 768), ("c", "yuvj422p", 854, 480), ("d", "gray", 1366, 10),
            ("e", "yuv444p", 130, 8), ("f", "yuv420p10le", 854, 480), ("g", "yuv420p", 1928, 1080), ("h", "nv21", 50, 10)]
    for i, (name, fmt, w, h) in enumerate(cams):
        path.open_stream(name, fmt, w, h)
        path.submit(name, make_planes(fmt, w, h, 40 + i), make_meta(timestamp=i + 1, frame_type="I"))
    assert path.tick() == len(cams)
    for i, (name, fmt, w, h) in enumerate(cams):
        want = F.to_bgr24(fmt, make_planes(fmt, w, h, 40 + i))
        got = path.result(name)
        assert got.shape == (h, w, 3) and got.flags["C_CONTIGUOUS"] and np.array_equal(got, want), name
        msg = bytes(path.result_packed(name))
        assert want.tobytes() in msg and len(msg) < h * w * 3 + 256
        t = path.result_tensor(name)                      # strided view when the device rows are padded
        assert tuple(t.shape) == (h, w, 3) and np.array_equal(t.cpu().numpy(), want), name
    # zero-copy staging hands out the padded pitch
    planes = make_planes("yuv420p", 854, 480, 99)
    views = path.acquire_input("a")
    for v, p in zip(views, planes):
        assert v.shape == p.shape
        v[...] = p
    path.commit_input("a")
    path.tick()
    assert np.array_equal(path.result("a"), F.to_bgr24("yuv420p", planes))
    # the unpitched device accessor refuses padded rows instead of handing out a wrong shape
    import ctypes as C
    p, shp = C.c_void_p(), (C.c_int32 * 3)()
    assert _abi.lib().vep_result_device(path._ctx, path._sid("a"), C.byref(p), shp) == _abi.VEP_ERR_GEOMETRY
    assert _abi.lib().vep_result_device(path._ctx, path._sid("g"), C.byref(p), shp) == _abi.VEP_ERR_GEOMETRY   # 1928*3 % 16 != 0


def test_tick_waits_for_the_tick_it_began_with_two_tickers(path):
    """Two threads share one FramePath and each runs submit -> tick -> result for its own camera (what two
    ReadImage threads of one process do).  Every result must be the frame that very thread just submitted."""
    w, h, rounds = 640, 480, 40
    names = ("left", "right")
    frames = {n: [make_planes("yuv420p", w, h, 1000 * k + i) for i in range(4)] for k, n in enumerate(names)}
    want = {n: [F.to_bgr24("yuv420p", p) for p in frames[n]] for n in names}
    for n in names:
        path.open_stream(n, "yuv420p", w, h)
    errors = []
    start = threading.Barrier(2)

    def worker(name):
        try:
            start.wait()
            for r in range(rounds):
                path.submit(name, frames[name][r % 4], make_meta(timestamp=r + 1))
                path.tick()
                if not np.array_equal(path.result(name), want[name][r % 4]):
                    errors.append((name, r))
        except Exception as e:      # pragma: no cover
            errors.append((name, repr(e)))

    ts = [threading.Thread(target=worker, args=(n,)) for n in names]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    assert not errors, errors[:5]


def test_tickets_and_empty_ticks(path):
    w, h = 320, 240
    path.open_stream("cam", "yuv420p", w, h, memory_buffer=2)
    n, t0 = path.tick_begin_ticket()
    assert (n, t0) == (0, 0)                              # nothing staged, no tick ever: ticket 0, waiting is a no-op
    path.tick_wait_ticket(t0)
    planes = [make_planes("yuv420p", w, h, s) for s in range(3)]
    path.submit("cam", planes[0])
    n1, t1 = path.tick_begin_ticket()
    path.submit("cam", planes[1])
    n2, t2 = path.tick_begin_ticket()
    n3, t3 = path.tick_begin_ticket()                     # empty: carries the newest real ticket
    assert (n1, n2, n3) == (1, 1, 0) and t2 == t1 + 1 and t3 == t2
    path.tick_wait_ticket(t3)                             # both real ticks have landed
    assert np.array_equal(path.result("cam"), F.to_bgr24("yuv420p", planes[1]))
    assert np.array_equal(path.result("cam", age=1), F.to_bgr24("yuv420p", planes[0]))
    path.tick_wait_ticket(t1)                             # already retired: returns at once
    path.tick_wait()                                      # nothing in flight: fine


def test_staging_helper_threads_split_a_4k_frame(path):
    w, h = 3840, 2160
    path.set_staging_threads(3)
    for fmt in ("nv12", "yuv420p"):
        planes = make_planes(fmt, w, h, 21)
        # decoder-style padded linesize on the source side as well
        padded = [np.full((p.shape[0], p.shape[1] + 64), 0xEE, np.uint8) for p in planes]
        for b, p in zip(padded, planes):
            b[:, :p.shape[1]] = p
        path.open_stream(fmt, fmt, w, h)
        path.submit(fmt, [b[:, :p.shape[1]] for b, p in zip(padded, planes)])
        assert path.tick() == 1
        assert np.array_equal(path.result(fmt), F.to_bgr24(fmt, planes)), fmt
    path.set_staging_threads(0)
    planes = make_planes("nv12", w, h, 22)
    path.submit("nv12", planes)
    path.tick()
    assert np.array_equal(path.result("nv12"), F.to_bgr24("nv12", planes))
    with pytest.raises(VepError):
        path.set_staging_threads(1000)


def test_submit_many_stages_a_whole_tick_with_helper_threads(path):
    """vep_submit_batch: every camera's frame of a tick in one call, copies shared with the helper threads; also
    gate-skipped items, padded decoder strides, ragged widths, and the error position on a full ring."""
    cams = [("s%d" % i, fmt, w, h) for i, (fmt, w, h) in enumerate(
        [("yuv420p", 1920, 1080)] * 3 + [("nv12", 3840, 2160), ("yuv420p", 854, 480), ("yuvj422p", 640, 480), ("gray", 50, 6), ("yuv420p10le", 640, 480)])]
    for name, fmt, w, h in cams:
        path.open_stream(name, fmt, w, h)
    for helpers in (0, 3):
        path.set_staging_threads(helpers)
        frames, want = [], {}
        for i, (name, fmt, w, h) in enumerate(cams):
            if i == 1:
                frames.append((name, None, None))          # the gate skipped this camera's frame
                continue
            planes = make_planes(fmt, w, h, 10 * helpers + i)
            padded = [np.full((p.shape[0], p.shape[1] + 24), 0xEE, p.dtype) for p in planes]      # decoder-style linesize
            for b, p in zip(padded, planes):
                b[:, :p.shape[1]] = p
            frames.append((name, [b[:, :p.shape[1]] for b, p in zip(padded, planes)], make_meta(timestamp=i + 1, packet=i)))
            want[name] = F.to_bgr24(fmt, planes)
        off0 = path.stats()["frames_offered"]
        assert path.submit_many(frames) == len(cams)
        assert path.stats()["frames_offered"] - off0 == len(cams)
        assert path.tick() == len(cams) - 1
        for name, w_ in want.items():
            assert np.array_equal(path.result(name), w_), (helpers, name)
    # a stream whose staging ring is full stops the batch there: items before it are staged, the status says why
    planes = make_planes("yuv420p", 1920, 1080, 1)
    for _ in range(6):                                       # 7 slots, one holds the newest result (ring_depth 1)
        path.submit("s0", planes)
    with pytest.raises(VepError) as e:
        path.submit_many([("s2", planes, None), ("s0", planes, None), ("s6", make_planes("gray", 50, 6, 0), None)])
    assert e.value.status == _abi.VEP_ERR_BUSY
    assert path.tick() == 7                                  # 6 frames of s0 + the one of s2 staged before the failure
    path.set_staging_threads(0)


def test_pinned_submit_reads_the_callers_planes_in_place(path):
    """VEP_SUBMIT_PINNED: frames that live in a registered (page-locked) decoder arena are converted without a CPU
    copy, whatever their linesize; pageable planes are refused in that mode; the staging slot's own memory stays
    untouched (sentinel) — proof that the H2D copy came from the caller's memory."""
    cams = [("p%d" % i, fmt, w, h) for i, (fmt, w, h) in enumerate(
        [("yuv420p", 1920, 1080), ("nv12", 1280, 720), ("yuv420p", 854, 480), ("yuvj422p", 640, 480), ("gray", 50, 6),
         ("yuv420p10le", 640, 480), ("yuv444p", 130, 8)])]
    arena = np.zeros(64 << 20, np.uint8)                     # the decoder's frame pool: one allocation, registered once
    path.register_host(arena)
    try:
        off, frames, want = 0, [], {}
        for i, (name, fmt, w, h) in enumerate(cams):
            path.open_stream(name, fmt, w, h)
            planes = make_planes(fmt, w, h, 60 + i)
            views = []
            for p in planes:                                 # decoder-style padded linesize (+32 bytes), 64-byte aligned planes
                pb = p.view(np.uint8).reshape(p.shape[0], -1)
                ls = pb.shape[1] + 32
                a = arena[off:off + pb.shape[0] * ls].reshape(pb.shape[0], ls)
                a[:, :pb.shape[1]] = pb
                a[:, pb.shape[1]:] = 0xEE
                views.append(a[:, :pb.shape[1]].view(p.dtype))
                off = (off + a.size + 63) // 64 * 64
            frames.append((name, views, make_meta(timestamp=i + 1)))
            want[name] = F.to_bgr24(fmt, planes)
        # poison the staging slots first: a pinned submit must not read them
        for name, fmt, w, h in cams:
            for v in path.acquire_input(name):
                v[...] = 0x77 if v.dtype == np.uint8 else 0x0177
            path.commit_input(name)
        path.tick()
        h2d0 = path.stats()["h2d_bytes"]
        assert path.submit_many(frames, pinned=True) == len(cams)
        assert path.tick() == len(cams)
        moved = path.stats()["h2d_bytes"] - h2d0
        visible = sum(sum(v.shape[0] * v.shape[1] * v.itemsize for v in f[1]) for f in frames)
        assert visible <= moved <= visible * 1.6              # rows travel with the decoder's padding at most; never the staging slots
        for name, w_ in want.items():
            assert np.array_equal(path.result(name), w_), name
        # single-frame call, and the frame may change once its tick has been waited for
        name, views, meta = frames[0]
        views[0][...] = 255 - views[0]
        path.submit(name, views, meta, pinned=True)
        path.tick()
        assert np.array_equal(path.result(name), F.to_bgr24("yuv420p", views))
        # pageable memory is refused in this mode, with the reason
        with pytest.raises(VepError) as e:
            path.submit("p0", make_planes("yuv420p", 1920, 1080, 1), None, pinned=True)
        assert e.value.status == _abi.VEP_ERR_INVALID_ARG and "page-locked" in str(e.value)
        assert path.tick() == 0
    finally:
        path.unregister_host(arena)


def test_staged_bytes_is_what_a_tick_copies_in(path):
    for i, (fmt, w, h) in enumerate([("yuv420p", 1920, 1080), ("nv12", 1366, 768), ("yuv420p", 854, 480), ("gray", 50, 6),
                                     ("yuvj422p", 640, 480), ("yuv444p", 130, 8), ("yuv420p10le", 640, 480)]):
        name = f"b{i}"
        path.open_stream(name, fmt, w, h)
        before = path.stats()["h2d_bytes"]
        path.submit(name, make_planes(fmt, w, h, i))
        path.tick()
        assert path.stats()["h2d_bytes"] - before == path.staged_bytes(name), (fmt, w, h)


def test_stream_groups_move_a_lock_step_fleet_with_one_copy_per_direction(path):
    """vep_stream_open_many: slots laid out [slot][camera]; neighbours that tick in lock step ride in one H2D and one D2H
    copy.  Results, packed messages (the framing is written after the merged copy swept its room) and copy counts."""
    cams = [("g%d" % i, fmt, w, h) for i, (fmt, w, h) in enumerate(
        [("yuv420p", 1920, 1080)] * 3 + [("nv12", 1280, 720)] * 2 + [("yuv420p", 854, 480)] + [("yuvj420p", 640, 480)] * 2)]
    path.open_streams(cams, memory_buffer=2)
    VideoFrame, _ = W.dynamic_message_class()

    def tick(selected, seed):
        frames, want = [], {}
        for name, fmt, w, h in cams:
            if name not in path.streams():
                continue
            if name in selected:
                planes = make_planes(fmt, w, h, seed + len(frames))
                frames.append((name, planes, make_meta(timestamp=seed, pts=seed + 1, packet=len(frames), frame_type="P")))
                want[name] = (F.to_bgr24(fmt, planes), seed)
            else:
                frames.append((name, None, None))
        s0 = path.stats()
        path.submit_many(frames)
        assert path.tick() == len(selected)
        s1 = path.stats()
        for name, (img, sd) in want.items():
            assert np.array_equal(path.result(name), img), name
            vf = VideoFrame(); vf.ParseFromString(bytes(path.result_packed(name)))
            assert vf.data == img.tobytes() and vf.timestamp == sd and vf.pts == sd + 1 and vf.frame_type == "P"
            assert [(d.size, d.name) for d in vf.shape.dim] == [(img.shape[0], "0"), (img.shape[1], "1"), (3, "2")]
        return s1["copies_in"] - s0["copies_in"], s1["copies_out"] - s0["copies_out"]

    names = [c[0] for c in cams]
    # all eight in lock step: one copy in; out: the 854-wide camera (padded device rows, 2-D copy) splits the run in three
    assert tick(set(names), 100) == (1, 3)
    assert tick(set(names), 200) == (1, 3)
    path.set_copy_run_limit(2)                              # at most two cameras per copy: {0,1} {2,3} {4,5} {6,7} in; {4} and {5} split out
    assert tick(set(names), 250) == (4, 5)
    path.set_copy_run_limit(0)
    # cameras 0-4 only: one run each way
    assert tick(set(names[:5]), 300) == (1, 1)
    # a gap splits the run: {0,1} and {3,4}; then they are out of step with 2 by one slot, and stay split
    assert tick({names[0], names[1], names[3], names[4]}, 400) == (2, 2)
    ci, co = tick(set(names[:5]), 500)
    assert (ci, co) == (3, 3)                               # {0,1}, {2}, {3,4}: camera 2 sits one slot index behind
    # a burst: two frames of one camera in one tick (different slot indices), neighbours one each
    planes_a, planes_b = make_planes("yuv420p", 1920, 1080, 601), make_planes("yuv420p", 1920, 1080, 602)
    path.submit(names[0], planes_a, make_meta(timestamp=1))
    path.submit(names[0], planes_b, make_meta(timestamp=2))
    path.submit(names[1], planes_a, make_meta(timestamp=3))
    assert path.tick() == 3
    assert np.array_equal(path.result(names[0]), F.to_bgr24("yuv420p", planes_b))
    assert np.array_equal(path.result(names[0], age=1), F.to_bgr24("yuv420p", planes_a))
    assert np.array_equal(path.result(names[1]), F.to_bgr24("yuv420p", planes_a))
    # members close one by one; the shared memory goes with the last; the names can be opened again
    for n in names:
        path.close_stream(n)
    path.open_streams(cams[:2])
    assert tick({names[0], names[1]}, 700) == (1, 1)
    with pytest.raises(VepError):
        path.open_streams([("dup", "yuv420p", 64, 48), ("dup", "yuv420p", 64, 48)])
    with pytest.raises(VepError):
        path.open_streams([("ok", "yuv420p", 64, 48), ("bad", "nv12", 64, 6)])
    assert "ok" not in path.streams()
