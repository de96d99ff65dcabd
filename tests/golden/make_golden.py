"""Regenerates the fixtures in this directory.  Run in the BUILD container only:

    PROTOCOL_BUFFERS_PYTHON_IMPLEMENTATION=python python tests/golden/make_golden.py

* pixels.json / small_frames.npz — outputs of the live libswscale (oracle/swscale_live.py: the two
  libswscale calls PyAV makes for reference python/read_image.py:94) on the seeded synthetic frames
  of video_edge_ai_proxy_b200/synth.py.  The rows that SURVEY.md §8(c) lists must reproduce its
  checksums; the script asserts that before writing.
* videoframe_wire.json — `SerializeToString()` of the reference's own generated module
  /root/reference/python/proto/video_streaming_pb2.py filled the way read_image.py:99-117 fills it
  (needs the pure-python protobuf backend because that generated code predates protobuf 4).
"""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import swscale_live as live  # noqa: E402
from video_edge_ai_proxy_b200.synth import make_planes  # noqa: E402

SURVEY_8C = {  # (fmt, w, h, seed, path) -> (sum, sha256[:16])
    ("yuv420p", 640, 480, 0, "simd"): (117728602, "e5778b72f362c875"),
    ("yuv420p", 640, 480, 0, "c"): (117582077, "8fc1135e10205fa9"),
    ("yuv420p", 1280, 720, 0, "simd"): (353742863, "22c3eed6b7702bb4"),
    ("yuv420p", 1920, 1080, 0, "simd"): (795964703, "66cdc051bce00a10"),
    ("yuv420p", 1920, 1080, 0, "c"): (794987645, "4fc1b33c7542662a"),
    ("yuv420p", 3840, 2160, 0, "simd"): (3184974613, "79c85329d1a09a84"),
    ("yuvj420p", 640, 480, 0, "simd"): (116690493, "c4b28365a378166a"),
    ("nv12", 640, 480, 0, "simd"): (117735317, "31c35f5256f5bc6a"),
    ("nv12", 3840, 2160, 0, "simd"): (3185534842, "bbaeece1f04701dc"),
    ("nv12", 3840, 2160, 0, "c"): (3182628851, "5ae47f8e96f3ae9c"),
}

EXTRA = [  # more rows (simd path), sizes/seeds the parity tests use
    ("yuv420p", 1920, 1080, 1), ("yuv420p", 1920, 1080, 15), ("yuvj420p", 1920, 1080, 0),
    ("nv12", 1920, 1080, 0), ("nv12", 1280, 720, 2), ("yuv420p", 1922, 1080, 0),
    ("yuv420p", 642, 482, 0), ("yuvj420p", 1280, 720, 5), ("nv12", 642, 482, 0),
    ("yuv420p", 2, 2, 0), ("yuv420p", 4096, 2, 0),
    # formats beyond SURVEY's three (probed: 4:2:2 keeps every chroma row; nv21 = nv12 with V first)
    ("yuvj422p", 1920, 1080, 0), ("yuv422p", 640, 480, 1), ("nv21", 1280, 720, 0), ("yuvj422p", 50, 6, 2), ("nv21", 642, 10, 3),
    ("yuv420p10le", 1920, 1080, 0), ("yuv420p10le", 642, 10, 1), ("yuv420p10le", 3840, 2160, 2),
    ("yuv444p", 1920, 1080, 0), ("yuvj444p", 1280, 720, 1), ("yuv444p", 50, 6, 2), ("gray", 1280, 720, 0), ("gray", 50, 6, 1),
]

SMALL = [  # stored raw (inputs + outputs) so the oracle can be checked without libswscale
    ("yuv420p", 32, 8, 11), ("yuv420p", 50, 6, 12), ("yuvj420p", 34, 4, 13),
    ("nv12", 32, 8, 14), ("nv12", 38, 12, 15), ("yuv420p", 2, 2, 16), ("yuvj422p", 34, 6, 17), ("nv21", 32, 8, 18), ("yuv420p10le", 34, 10, 19), ("yuv444p", 18, 4, 20),
]


def digest(a):
    return int(a.sum(dtype=np.int64)), hashlib.sha256(a.tobytes()).hexdigest()


def main():
    assert live.on_simd_path(), "libswscale is not on its x86 SIMD path on this host"
    rows = []
    for (fmt, w, h, seed, path), (s, sha16) in SURVEY_8C.items():
        live.force_cpu_flags(0 if path == "c" else -1)
        out = live.to_bgr24(fmt, make_planes(fmt, w, h, seed), w, h)
        live.force_cpu_flags(-1)
        gs, gsha = digest(out)
        assert (gs, gsha[:16]) == (s, sha16), (fmt, w, h, path, gs, gsha)
        rows.append(dict(fmt=fmt, w=w, h=h, seed=seed, path=path, sum=gs, sha256=gsha,
                         survey_8c=True))
    for fmt, w, h, seed in EXTRA:
        out = live.to_bgr24(fmt, make_planes(fmt, w, h, seed), w, h)
        gs, gsha = digest(out)
        rows.append(dict(fmt=fmt, w=w, h=h, seed=seed, path="simd", sum=gs, sha256=gsha,
                         survey_8c=False))
    # known answers: flat frames, U = V = 128, limited range
    kat = {}
    for yv in (0, 16, 17, 18, 128, 200, 234, 235, 255):
        Y = np.full((2, 16), yv, np.uint8)
        C = np.full((1, 8), 128, np.uint8)
        kat[str(yv)] = int(live.to_bgr24("yuv420p", (Y, C, C), 16, 2)[0, 0, 0])
    meta = dict(libswscale=live.version(), cpu_flags=hex(live.cpu_flags()), rows=rows,
                gray_kat=kat)
    with open(os.path.join(HERE, "pixels.json"), "w") as f:
        json.dump(meta, f, indent=1)

    small = {}
    for n, (fmt, w, h, seed) in enumerate(SMALL):
        planes = make_planes(fmt, w, h, seed)
        small[f"{n}_{fmt}_{w}x{h}_s{seed}_out"] = live.to_bgr24(fmt, planes, w, h)
        for i, p in enumerate(planes):
            small[f"{n}_{fmt}_{w}x{h}_s{seed}_p{i}"] = p
    np.savez_compressed(os.path.join(HERE, "small_frames.npz"), **small)

    # ---- wire goldens from the reference's generated protobuf module ----
    os.environ.setdefault("PROTOCOL_BUFFERS_PYTHON_IMPLEMENTATION", "python")
    sys.path.insert(0, "/root/reference/python")
    from proto import video_streaming_pb2 as pb  # noqa: E402

    cases = [
        dict(width=1920, height=1080, payload_len=1920 * 1080 * 3, timestamp=1604660000123,
             is_keyframe=True, pts=900000, dts=897000, frame_type="I", is_corrupt=False,
             time_base=1 / 90000, shape=[1080, 1920, 3], packet=0, keyframe=1),
        dict(width=640, height=480, payload_len=640 * 480 * 3, timestamp=77, is_keyframe=False,
             pts=-3003, dts=-6006, frame_type="P", is_corrupt=True, time_base=0.04,
             shape=[480, 640, 3], packet=17, keyframe=300),
        dict(width=2, height=2, payload_len=12, timestamp=0, is_keyframe=False, pts=0, dts=0,
             frame_type="", is_corrupt=False, time_base=0.0, shape=[2, 2, 3], packet=0,
             keyframe=0),
        dict(width=3840, height=2160, payload_len=3840 * 2160 * 3, timestamp=(1 << 62) + 5,
             is_keyframe=True, pts=(1 << 40), dts=-(1 << 40), frame_type="B", is_corrupt=False,
             time_base=1001 / 30000, shape=[2160, 3840, 3], packet=127, keyframe=128),
    ]
    out_cases = []
    for c in cases:
        n = c["payload_len"]
        data = (np.arange(n, dtype=np.uint32) * 2654435761 >> 24).astype(np.uint8).tobytes()
        vf = pb.VideoFrame()
        vf.data = data
        vf.width, vf.height, vf.timestamp = c["width"], c["height"], c["timestamp"]
        vf.frame_type = c["frame_type"]
        vf.pts, vf.dts, vf.packet, vf.keyframe = c["pts"], c["dts"], c["packet"], c["keyframe"]
        vf.time_base = c["time_base"]
        vf.is_keyframe, vf.is_corrupt = c["is_keyframe"], c["is_corrupt"]
        for i, d in enumerate(c["shape"]):
            nd = pb.ShapeProto.Dim()
            nd.size = d
            nd.name = str(i)
            vf.shape.dim.append(nd)
        msg = vf.SerializeToString()
        at = msg.index(data[:12]) if n >= 12 else 0
        assert msg[at:at + n] == data
        out_cases.append(dict(fields=c, prefix_hex=msg[:at].hex(), suffix_hex=msg[at + n:].hex(),
                              total_len=len(msg), sha256=hashlib.sha256(msg).hexdigest()))
    with open(os.path.join(HERE, "videoframe_wire.json"), "w") as f:
        json.dump(dict(payload_rule="data[i] = ((i * 2654435761) mod 2^32) >> 24",
                       cases=out_cases), f, indent=1)
    print("wrote pixels.json, small_frames.npz, videoframe_wire.json")


if __name__ == "__main__":
    main()
