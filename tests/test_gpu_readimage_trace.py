"""GPU: this repo's `ReadImage` (drop-in thread class) and `FrameProducer` publish, for the scripts of
tests/readimage_script.py, byte-identical messages to those the REFERENCE's own `ReadImage.run` published
(tests/golden/readimage_trace.json, generator tests/golden/make_readimage_trace.py): same XADD key, MAXLEN,
order, and sha256 of every serialized VideoFrame — pixels (libswscale there, sm_100a kernels here) included."""
import hashlib
import json
import os
import queue
import threading

import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

import readimage_script as RS  # noqa: E402
from video_edge_ai_proxy_b200 import CameraControl, FramePath, FrameProducer, global_vars  # noqa: E402
from video_edge_ai_proxy_b200.read_image import ReadImage  # noqa: E402

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "readimage_trace.json")


class _Redis:
    def __init__(self, pipelined):
        self.kv, self.calls, self.streams, self.pipelined = {}, [], {}, pipelined

    def get(self, key):
        return self.kv.get(key)

    def xadd(self, name, fields, maxlen=None):
        data = bytes(fields["data"])
        self.calls.append((name, maxlen, hashlib.sha256(data).hexdigest(), len(data)))
        s = self.streams.setdefault(name, [])
        s.append(self.calls[-1][2])
        if maxlen is not None:
            del s[:-maxlen]

    def __getattr__(self, name):
        if name == "pipeline" and self.pipelined:
            outer = self

            class _Pipe:
                def __init__(self):
                    self.cmds = []

                def xadd(self, name, fields, maxlen=None):
                    self.cmds.append((name, {"data": bytes(fields["data"])}, maxlen))

                def execute(self):
                    for n, f, m in self.cmds:
                        outer.xadd(n, f, maxlen=m)
                    self.cmds = []
            return lambda transaction=True: _Pipe()
        raise AttributeError(name)


@pytest.fixture(scope="module")
def trace():
    with open(GOLDEN) as f:
        return json.load(f)["scenarios"]


@pytest.fixture()
def path():
    if not torch.cuda.is_available():
        pytest.fail("marked gpu but no CUDA device")
    p = FramePath(0)
    yield p
    p.close()


def _want(t):
    return [(c["key"], c["maxlen"], c["sha256"], c["len"]) for c in t["xadd"]]


def test_read_image_class_publishes_the_reference_messages(trace, path):
    """process_packet() is the body of the reference loop for one popped packet; played like the generator plays the reference."""
    for name, t in trace.items():
        sc = RS.scenarios()[name]
        r = _Redis(pipelined=False)
        ri = ReadImage(queue.Queue(), "cam-" + name, sc["memory_buffer"], r, threading.Event(), threading.Condition(), frame_path=path)
        per_packet = []
        for pkt, (_kf, _c, kfo, query, _n) in zip(RS.packets_of(sc), sc["events"]):
            r.kv[global_vars.RedisIsKeyFrameOnlyPrefix + "cam-" + name] = b"true" if kfo else b"false"
            global_vars.query_timestamp = query
            before = len(r.calls)
            ri.process_packet(pkt)
            per_packet.append(len(r.calls) - before)
        assert r.calls == _want(t), name
        assert per_packet == t["published_per_packet"], name
        assert r.streams.get("cam-" + name, []) == t["stream_after"], name


@pytest.mark.parametrize("pipelined", [False, True])
def test_frame_producer_publishes_the_reference_messages(trace, path, pipelined):
    """All scenarios as cameras of ONE producer, advanced in lock step: one flush (one launch, one pipelined
    round trip) per step serves every camera; per camera the messages are the reference's, in order."""
    r = _Redis(pipelined)
    prod = FrameProducer(path, r, now_ms=lambda: 9000)          # every script's query stamps are < 10 s old at t = 9000
    scs = RS.scenarios()
    packets = {}
    for name, sc in scs.items():
        prod.add_camera("cam-" + name, sc["fmt"], sc["w"], sc["h"], sc["memory_buffer"])
        packets[name] = RS.packets_of(sc)
    for step in range(max(len(sc["events"]) for sc in scs.values())):
        for name, sc in scs.items():
            if step < len(sc["events"]):
                _kf, _c, kfo, query, _n = sc["events"][step]
                prod.on_packet("cam-" + name, packets[name][step], CameraControl(last_query=query, key_frame_only=kfo))
        prod.flush()
    for name, t in trace.items():
        assert [c for c in r.calls if c[0] == "cam-" + name] == _want(t), name
        assert r.streams.get("cam-" + name, []) == t["stream_after"], name
        assert [p.decode_calls for p in packets[name]] == t["decode_calls"], name


def test_two_read_image_threads_share_one_frame_path(trace, path):
    """Two `ReadImage` objects of one process on ONE FramePath, each fed from its own thread (the deployment
    INTEGRATION.md describes for batching cameras in one process): every camera still publishes exactly the reference's
    messages — a tick waits for the tick it began, whoever else ticks."""
    picks = ["query_mid_gop_burst", "multi_frame_packets"]
    results, errors = {}, []

    def camera(name):
        try:
            sc = RS.scenarios()[name]
            r = _Redis(pipelined=False)
            ri = ReadImage(queue.Queue(), "cam2-" + name, sc["memory_buffer"], r, threading.Event(), threading.Condition(), frame_path=path)
            for _ in range(3):            # three passes over the script: more interleavings
                ri._gate.__init__()
                r.calls.clear(); r.streams.clear()
                for pkt, (_kf, _c, kfo, query, _n) in zip(RS.packets_of(sc), sc["events"]):
                    r.kv[global_vars.RedisIsKeyFrameOnlyPrefix + "cam2-" + name] = b"true" if kfo else b"false"
                    ri.process_packet(pkt, query_timestamp=query)        # per-camera query stamp: the threads run freely
                results[name] = [(k.replace("cam2-", "cam-"), m, h, n) for k, m, h, n in r.calls]
        except Exception as e:      # pragma: no cover
            errors.append((name, repr(e)))

    th = [threading.Thread(target=camera, args=(n,)) for n in picks]
    [t.start() for t in th]
    [t.join() for t in th]
    assert not errors, errors
    for name in picks:
        assert results[name] == _want(trace[name]), name
