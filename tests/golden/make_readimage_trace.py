"""Generates tests/golden/readimage_trace.json by running the REFERENCE's own consumer thread,
/root/reference/python/read_image.py `ReadImage.run` (:47-133), unmodified, on the scripts of
tests/readimage_script.py.  Build container only (/root/reference does not exist on the GPU box); the
tests consume the committed JSON.

How the reference class is made to run here:
  * it imports `redis` only for the type of its connection -> a stub module is put into sys.modules and
    a fake connection (`get`, `xadd`) is handed to the constructor;
  * `from proto import video_streaming_pb2` is the reference's generated protobuf-3 module -> needs
    PROTOCOL_BUFFERS_PYTHON_IMPLEMENTATION=python under protobuf 6;
  * `frame.to_ndarray(format='bgr24')` (:94) is PyAV -> the fake frame makes the two libswscale calls PyAV
    makes (oracle/swscale_live.py), on the real library;
  * `query_timestamp` is set on the reference MODULE before each packet: that is the value the producer
    thread would have published (rtsp_to_rtmp.py:148) had the two modules shared the variable;
  * this script plays the producer thread deterministically: one packet into the queue, event set, notify
    until the consumer has called task_done() — one pop per wake-up, exactly the reference's rhythm.

Recorded per XADD: key, maxlen, sha256 and length of the serialized VideoFrame, and its scalar fields.
"""
import hashlib
import json
import os
import queue
import sys
import threading
import time
import types

os.environ["PROTOCOL_BUFFERS_PYTHON_IMPLEMENTATION"] = "python"
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
REF = "/root/reference/python"
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), REF]
sys.modules.setdefault("redis", types.ModuleType("redis"))     # read_image.py:19 only imports it

import read_image as ref_read_image                             # noqa: E402  the reference module
from proto import video_streaming_pb2                           # noqa: E402  the reference's generated code

import readimage_script as RS                                   # noqa: E402

assert ref_read_image.__file__.startswith(REF), ref_read_image.__file__


class FakeRedis:
    def __init__(self):
        self.kv, self.calls, self.streams = {}, [], {}

    def get(self, key):
        return self.kv.get(key)

    def xadd(self, name, fields, maxlen=None):
        data = bytes(fields["data"])
        vf = video_streaming_pb2.VideoFrame()
        vf.ParseFromString(data)
        self.calls.append(dict(
            key=name, maxlen=maxlen, sha256=hashlib.sha256(data).hexdigest(), len=len(data),
            fields=dict(width=vf.width, height=vf.height, timestamp=vf.timestamp, frame_type=vf.frame_type, pts=vf.pts, dts=vf.dts,
                        packet=vf.packet, keyframe=vf.keyframe, time_base=vf.time_base, is_keyframe=vf.is_keyframe,
                        is_corrupt=vf.is_corrupt, device_id=vf.device_id, data_sha256=hashlib.sha256(vf.data).hexdigest(),
                        shape=[[d.size, d.name] for d in vf.shape.dim])))
        s = self.streams.setdefault(name, [])
        s.append(hashlib.sha256(data).hexdigest())
        if maxlen is not None:
            del s[:-maxlen]


def play(name, sc):
    r = FakeRedis()
    q, ev, cond = queue.Queue(), threading.Event(), threading.Condition()
    th = ref_read_image.ReadImage(q, "cam-" + name, sc["memory_buffer"], r, ev, cond)
    th.daemon = True                                            # run() never returns (while True)
    th.start()
    packets = RS.packets_of(sc)
    per_packet = []
    for pkt, (_kf, _c, kfo, query, _n) in zip(packets, sc["events"]):
        r.kv[ref_read_image.RedisIsKeyFrameOnlyPrefix + "cam-" + name] = b"true" if kfo else b"false"
        ref_read_image.query_timestamp = query                  # the repair of SURVEY §3.2
        before = len(r.calls)
        q.put(pkt)
        ev.set()
        t0 = time.time()
        while q.unfinished_tasks:                               # notify until the consumer popped and finished it
            with cond:
                cond.notify_all()
            time.sleep(0.0005)
            assert time.time() - t0 < 30, "reference consumer did not finish the packet"
        per_packet.append(len(r.calls) - before)
    return dict(fmt=sc["fmt"], w=sc["w"], h=sc["h"], memory_buffer=sc["memory_buffer"],
                events=[list(e) for e in sc["events"]], published_per_packet=per_packet,
                decode_calls=[p.decode_calls for p in packets], xadd=r.calls, stream_after=r.streams.get("cam-" + name, []))


def main():
    out = {"generator": "tests/golden/make_readimage_trace.py", "reference": "python/read_image.py ReadImage.run, unmodified",
           "libswscale": "9.1.100 (bundled with opencv-python-headless)", "scenarios": {}}
    for name, sc in RS.scenarios().items():
        out["scenarios"][name] = play(name, sc)
        print(name, "packets", len(sc["events"]), "published", len(out["scenarios"][name]["xadd"]))
    with open(os.path.join(ROOT, "tests", "golden", "readimage_trace.json"), "w") as f:
        json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
