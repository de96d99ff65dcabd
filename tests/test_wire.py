"""CPU: VideoFrame wire framing — the C packer (libvepb200.so, host code, no GPU needed) against
(i) goldens produced by the reference's own generated protobuf module, (ii) the python restatement,
(iii) the installed protobuf runtime through a dynamically built descriptor."""
import hashlib
import random

import pytest

from conftest import wire_payload
from oracle import videoframe_wire as W
from video_edge_ai_proxy_b200 import make_meta, videoframe_framing


def _kw(c, data=b""):
    return dict(width=c["width"], height=c["height"], data=data, timestamp=c["timestamp"],
                is_keyframe=c["is_keyframe"], pts=c["pts"], dts=c["dts"], frame_type=c["frame_type"],
                is_corrupt=c["is_corrupt"], time_base=c["time_base"], shape=c["shape"], packet=c["packet"],
                keyframe=c["keyframe"])


def _meta(c):
    return make_meta(timestamp=c["timestamp"], pts=c["pts"], dts=c["dts"], packet=c["packet"], keyframe=c["keyframe"],
                     time_base=c["time_base"], is_keyframe=c["is_keyframe"], is_corrupt=c["is_corrupt"],
                     frame_type=c["frame_type"])


def test_c_packer_matches_reference_protobuf_goldens(golden_wire):
    for case in golden_wire["cases"]:
        c = case["fields"]
        pre, suf = videoframe_framing(_meta(c), c["width"], c["height"])
        assert pre.hex() == case["prefix_hex"]
        assert suf.hex() == case["suffix_hex"]
        msg = pre + wire_payload(c["payload_len"]) + suf
        assert len(msg) == case["total_len"]
        assert hashlib.sha256(msg).hexdigest() == case["sha256"]


def test_python_restatement_matches_goldens(golden_wire):
    for case in golden_wire["cases"]:
        c = case["fields"]
        msg = W.serialize_videoframe(**_kw(c, wire_payload(c["payload_len"])))
        assert hashlib.sha256(msg).hexdigest() == case["sha256"]


def test_survey_prefix_and_shape_bytes():
    m = make_meta(timestamp=1, frame_type="I")
    pre, suf = videoframe_framing(m, 1920, 1080)
    assert pre.hex() == "08800f10b8081a80d8fb02"
    assert "5a17120608b808120130120608800f12013112050803120132" in suf.hex()


def test_random_fields_against_protobuf_runtime():
    rnd = random.Random(7)
    for _ in range(200):
        w, h = rnd.choice([(2, 2), (640, 480), (1280, 720), (1920, 1080), (3840, 2160), (16384, 2)])
        c = dict(width=w, height=h, timestamp=rnd.choice([0, 1, rnd.getrandbits(40), rnd.getrandbits(62)]),
                 is_keyframe=rnd.random() < .5, pts=rnd.choice([0, rnd.getrandbits(33), -rnd.getrandbits(33)]),
                 dts=rnd.choice([0, -1, rnd.getrandbits(50)]), frame_type=rnd.choice(["", "I", "P", "B", "NONE", "SI"]),
                 is_corrupt=rnd.random() < .3, time_base=rnd.choice([0.0, 1 / 90000, 0.04, 1001 / 30000, -0.0]),
                 shape=[h, w, 3], packet=rnd.choice([0, 1, 255, 70000]), keyframe=rnd.choice([0, 1, 999999]))
        pre, suf = videoframe_framing(_meta(c), w, h)
        # payload replaced by a short stand-in: splice the length varint to compare the rest
        data = bytes(rnd.getrandbits(8) for _ in range(12)) if (w, h) == (2, 2) else None
        if data is not None:
            assert pre + data + suf == W.serialize_with_protobuf_runtime(**_kw(c, data))
            assert pre + data + suf == W.serialize_videoframe(**_kw(c, data))
        else:
            ref = W.serialize_videoframe(**_kw(c, b"\x00"))   # 1-byte payload: tag+len = 2 bytes
            k = ref.index(b"\x1a\x01\x00")
            assert pre[:-len(W.varint(w * h * 3)) - 1] == ref[:k]
            assert pre[-len(W.varint(w * h * 3)):] == W.varint(w * h * 3)
            assert suf == ref[k + 3:]


def test_bad_arguments():
    from video_edge_ai_proxy_b200 import VepError
    with pytest.raises(VepError):
        videoframe_framing(make_meta(), 0, 10)
    with pytest.raises(ValueError):
        make_meta(frame_type="TOOLONGNAME")


_REF_SCRIPT = r'''
import os, sys
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, "/root/reference/python")
import numpy as np
from proto import video_streaming_pb2 as pb            # the reference's own generated module
from video_edge_ai_proxy_b200 import make_meta, videoframe_framing
rng = np.random.default_rng(5)
for w, h in ((640, 480), (2, 2), (1920, 1080)):
    f = dict(timestamp=int(rng.integers(0, 1 << 50)), pts=int(rng.integers(-1 << 40, 1 << 40)), dts=int(rng.integers(-5, 5)),
             packet=int(rng.integers(0, 300)), keyframe=int(rng.integers(0, 1 << 20)), time_base=1 / 90000,
             is_keyframe=bool(rng.integers(2)), is_corrupt=bool(rng.integers(2)), frame_type=str(rng.choice(["I", "P", "B", ""])))
    pre, suf = videoframe_framing(make_meta(**f), w, h)
    data = rng.integers(0, 256, w * h * 3, dtype=np.uint8).tobytes()
    msg = pre + data + suf
    vf = pb.VideoFrame(); vf.ParseFromString(msg)                       # what grpc_api.go:212-216 does with proto.Unmarshal
    assert (vf.width, vf.height, vf.data) == (w, h, data)
    for k, v in f.items():
        assert getattr(vf, k) == v, (k, getattr(vf, k), v)
    assert [(d.size, d.name) for d in vf.shape.dim] == [(h, "0"), (w, "1"), (3, "2")] and vf.device_id == ""
    assert vf.SerializeToString() == msg                                # and what it re-sends is the same bytes
    img = np.frombuffer(vf.data, dtype=np.uint8).reshape([d.size for d in vf.shape.dim])   # examples/opencv_display.py:48-53
    assert img.shape == (h, w, 3)
print("REF_PB2_OK")
'''


def test_reference_generated_module_parses_our_messages(tmp_path):
    """Only where the reference tree is mounted (the build container): its own video_streaming_pb2 must parse and
    re-serialize the C packer's output unchanged."""
    import os
    import subprocess
    import sys
    if not os.path.exists("/root/reference/python/proto/video_streaming_pb2.py"):
        pytest.skip("reference tree not mounted on this machine")
    script = tmp_path / "ref_parse.py"
    script.write_text(_REF_SCRIPT)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, PROTOCOL_BUFFERS_PYTHON_IMPLEMENTATION="python")
    out = subprocess.run([sys.executable, str(script), root], capture_output=True, text=True, env=env, timeout=120)
    assert out.returncode == 0 and "REF_PB2_OK" in out.stdout, out.stdout + out.stderr
