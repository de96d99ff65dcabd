"""CPU: `Image.VideoLatestImage` pass-through (video_edge_ai_proxy_b200/image_service.py) against the behaviour of
reference server/grpcapi/grpc_api.go:133-235, in-process over a real grpc channel and a fake Redis; and, where the
reference tree is mounted, through the reference's own generated client stub (examples/basic_usage.py:50-62)."""
import os
import subprocess
import sys

import numpy as np
import pytest

grpc = pytest.importorskip("grpc")

from conftest import wire_payload  # noqa: E402
from video_edge_ai_proxy_b200 import make_meta, videoframe_framing  # noqa: E402
from video_edge_ai_proxy_b200.image_service import (METHOD, SERVICE, ImageService, parse_video_frame_request,  # noqa: E402
                                                    serve)

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class FakeRedis:
    """set / hset / xread with redis-py's shapes; streams are lists of (id, {b'data': bytes})."""

    def __init__(self):
        self.kv, self.hashes, self.streams, self.xreads = {}, {}, {}, []

    def set(self, k, v):
        self.kv[k] = v

    def hset(self, k, mapping=None):
        self.hashes.setdefault(k, {}).update(mapping)

    def xadd(self, name, data):
        s = self.streams.setdefault(name, [])
        s.append((b"%d-0" % (len(s) + 1), {b"data": data}))

    def xread(self, streams, count=None, block=None):
        self.xreads.append((dict(streams), count, block))
        out = []
        for name, last in streams.items():
            last_n = int(str(last).split("-")[0])
            new = [(i, f) for i, f in self.streams.get(name, []) if int(i.split(b"-")[0]) > last_n]
            if new:
                out.append((name.encode(), new[:count]))
        return out


def packed(w, h, pts, seed=0):
    meta = make_meta(timestamp=1000 + pts, pts=pts, dts=pts - 1, packet=3, keyframe=7, time_base=1 / 90000, is_keyframe=True, frame_type="I")
    pre, suf = videoframe_framing(meta, w, h)
    return pre + wire_payload(w * h * 3)[seed:] + wire_payload(w * h * 3)[:seed] + suf


def request_bytes(device_id, key_frame_only):
    b = b"\x08\x01" if key_frame_only else b""
    d = device_id.encode()
    return b + b"\x12" + bytes([len(d)]) + d


def test_request_parser():
    assert parse_video_frame_request(request_bytes("cam-1", True)) == (True, "cam-1")
    assert parse_video_frame_request(request_bytes("x", False)) == (False, "x")
    assert parse_video_frame_request(b"") == (False, "")
    assert parse_video_frame_request(b"\x12\x03abc\x18\x05\x08\x01") == (True, "abc")       # unknown field 3 skipped, any order
    with pytest.raises(ValueError):
        parse_video_frame_request(b"\x12\x09abc")


def test_latest_follows_the_reference_request_handling():
    r = FakeRedis()
    svc = ImageService(r, now_ms=lambda: 1604660000123)
    for pts in (1, 2, 3):
        r.xadd("cam", packed(64, 48, pts))
    got = svc.latest("cam", True)
    assert got == packed(64, 48, 3)                                              # the LAST message of the read (grpc_api.go:207)
    assert r.kv["is_key_frame_only_cam"] == "true"                               # :159-160, strconv.FormatBool
    assert r.hashes["last_access_time_cam"] == {"last_query": 1604660000123}     # :166-171
    assert r.xreads[0] == ({"cam": "0"}, 60, 1000)                               # :189-193
    assert svc.latest("cam", False) == b"" and r.kv["is_key_frame_only_cam"] == "false"    # nothing new: empty VideoFrame after 3 tries
    assert [x[0] for x in r.xreads[1:]] == [{"cam": "3-0"}] * 3                  # resumes after the id it forwarded (:181-184, :208)
    r.xadd("cam", packed(64, 48, 4))
    assert svc.latest("cam", False) == packed(64, 48, 4)


@pytest.fixture()
def server():
    r = FakeRedis()
    srv, port, svc = serve(r, "127.0.0.1:0")
    yield r, port, svc
    srv.stop(0)


def test_pass_through_over_grpc_forwards_the_packed_bytes_untouched(server):
    r, port, svc = server
    big = packed(1920, 1080, 5)                                                  # 6.2 MB: beyond gRPC's 4 MiB default limits
    r.xadd("cam-hd", big)
    r.xadd("cam-sd", packed(640, 480, 9, seed=11))
    ch = grpc.insecure_channel(f"127.0.0.1:{port}", options=[("grpc.max_receive_message_length", 64 << 20)])
    call = ch.stream_stream(f"/{SERVICE}/{METHOD}", request_serializer=lambda t: request_bytes(*t), response_deserializer=lambda b: b)
    replies = list(call(iter([("cam-hd", False), ("cam-sd", True), ("nobody", False)])))
    assert replies[0] == big and replies[1] == packed(640, 480, 9, seed=11) and replies[2] == b""
    assert svc.forwarded_bytes == len(big) + len(replies[1])
    assert r.kv == {"is_key_frame_only_cam-hd": "false", "is_key_frame_only_cam-sd": "true", "is_key_frame_only_nobody": "false"}
    ch.close()


_REF_CLIENT = r'''
import sys
sys.path[:0] = ["/root/reference/examples"]
import grpc, numpy as np
import video_streaming_pb2, video_streaming_pb2_grpc           # the reference's generated client code
port, want_pts = int(sys.argv[1]), int(sys.argv[2])
channel = grpc.insecure_channel("127.0.0.1:%d" % port, options=[("grpc.max_receive_message_length", 64 << 20)])
stub = video_streaming_pb2_grpc.ImageStub(channel)
def gen():                                                       # examples/basic_usage.py:30-38
    req = video_streaming_pb2.VideoFrameRequest()
    req.device_id = "cam-hd"
    req.key_frame_only = True
    yield req
n = 0
for frame in stub.VideoLatestImage(gen()):
    img = np.frombuffer(frame.data, dtype=np.uint8).reshape([d.size for d in frame.shape.dim])    # examples/opencv_display.py:48-53
    assert img.shape == (1080, 1920, 3) and frame.pts == want_pts and frame.is_keyframe and frame.frame_type == "I"
    assert (frame.width, frame.height, frame.packet, frame.keyframe) == (1920, 1080, 3, 7)
    n += 1
assert n == 1
print("REF_CLIENT_OK")
'''


def test_reference_client_stub_receives_frames_from_the_pass_through(server, tmp_path):
    if not os.path.exists("/root/reference/examples/video_streaming_pb2_grpc.py"):
        pytest.skip("reference tree not mounted on this machine")
    r, port, svc = server
    r.xadd("cam-hd", packed(1920, 1080, 77))
    script = tmp_path / "client.py"
    script.write_text(_REF_CLIENT)
    env = dict(os.environ, PROTOCOL_BUFFERS_PYTHON_IMPLEMENTATION="python")
    out = subprocess.run([sys.executable, str(script), str(port), "77"], capture_output=True, text=True, env=env, timeout=120)
    assert out.returncode == 0 and "REF_CLIENT_OK" in out.stdout, out.stdout + out.stderr
    assert r.kv["is_key_frame_only_cam-hd"] == "true"
