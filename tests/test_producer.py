"""CPU: FrameProducer's publishing side (read_image.py:121) with a FramePath stand-in — one pipelined
round trip per tick, reference XADD arguments per message, oldest first."""
from video_edge_ai_proxy_b200 import CameraControl, FrameProducer

import readimage_script as RS


class _Path:
    """Stand-in for FramePath: remembers staged frames per camera, `tick` makes them readable by age."""

    def __init__(self):
        self.staged, self.ready, self.opened = {}, {}, {}

    def open_stream(self, device_id, fmt, w, h, ring):
        self.opened[device_id] = ring

    def submit(self, device_id, planes, meta, pinned=False):
        self.staged.setdefault(device_id, []).append(b"msg-%d" % meta.pts)

    def tick(self):
        n = sum(len(v) for v in self.staged.values())
        for k, v in self.staged.items():
            self.ready[k] = (self.ready.get(k, []) + v)[-8:]
        self.staged = {}
        return n

    def result_packed(self, device_id, age=0):
        return memoryview(self.ready[device_id][-1 - age])


class _Pipe:
    def __init__(self, owner):
        self.owner, self.cmds = owner, []

    def xadd(self, name, fields, maxlen=None):
        self.cmds.append((name, bytes(fields["data"]), maxlen))

    def execute(self):
        self.owner.round_trips += 1
        self.owner.log += self.cmds
        self.cmds = []


class _Redis:
    def __init__(self):
        self.round_trips, self.log, self.pipe_args = 0, [], []

    def pipeline(self, transaction=True):
        self.pipe_args.append(transaction)
        return _Pipe(self)


class _PlainRedis:
    def __init__(self):
        self.log = []

    def xadd(self, name, fields, maxlen=None):
        self.log.append((name, bytes(fields["data"]), maxlen))


def _feed(prod, now):
    # camera a: a query mid-GOP produces a burst of three frames in one tick; camera b: one frame
    pk = {c: [RS.FakePacket("yuv420p", 64, 48, i, i == 0, False, 1) for i in range(3)] for c in "ab"}
    for i in range(3):
        now[0] += 40
        prod.on_packet("a", pk["a"][i], CameraControl(last_query=1 if i < 2 else now[0]))
        if i == 2:
            prod.on_packet("b", pk["b"][0], CameraControl(last_query=now[0]))
        if i == 0:
            prod.flush()


def test_flush_is_one_pipelined_round_trip_per_tick_oldest_first():
    r, now = _Redis(), [5000]
    prod = FrameProducer(_Path(), r, now_ms=lambda: now[0])
    prod.add_camera("a", "yuv420p", 64, 48, memory_buffer=3)
    prod.add_camera("b", "yuv420p", 64, 48, memory_buffer=100)       # ring depth is clamped, MAXLEN is not
    assert prod.path.opened == {"a": 3, "b": 64}
    _feed(prod, now)
    assert r.round_trips == 1 and r.log == [("a", b"msg-0", 3)]     # the key frame of the first tick
    prod.flush()
    assert r.round_trips == 2 and r.pipe_args == [False, False]      # non-transactional, as separate XADDs would be
    assert r.log[1:] == [("a", b"msg-1000", 3), ("a", b"msg-2000", 3), ("b", b"msg-0", 100)]
    prod.flush()                                                     # nothing staged: no round trip at all
    assert r.round_trips == 2


def test_flush_without_pipeline_support_issues_the_reference_calls():
    r, now = _PlainRedis(), [5000]
    prod = FrameProducer(_Path(), r, now_ms=lambda: now[0])
    prod.add_camera("a", "yuv420p", 64, 48, memory_buffer=3)
    prod.add_camera("b", "yuv420p", 64, 48, memory_buffer=1)
    _feed(prod, now)
    prod.flush()
    assert r.log == [("a", b"msg-0", 3), ("a", b"msg-1000", 3), ("a", b"msg-2000", 3), ("b", b"msg-0", 1)]
