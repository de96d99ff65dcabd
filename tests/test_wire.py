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
