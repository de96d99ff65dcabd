"""CPU: the decode gate and the per-frame VideoFrame fields against tests/golden/readimage_trace.json —
the XADD payloads the REFERENCE's own `ReadImage.run` (python/read_image.py:47-133) produced for the
scripts of tests/readimage_script.py (generator: tests/golden/make_readimage_trace.py)."""
import hashlib
import json
import os

import numpy as np
import pytest

import readimage_script as RS
from oracle import gate_ref
from oracle import swscale_live as L
from video_edge_ai_proxy_b200 import DecodeGate, make_meta, videoframe_framing

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "readimage_trace.json")


@pytest.fixture(scope="module")
def trace():
    with open(GOLDEN) as f:
        return json.load(f)["scenarios"]


def test_trace_was_generated_from_the_committed_scripts(trace):
    sc = RS.scenarios()
    assert set(sc) == set(trace)
    for name, s in sc.items():
        t = trace[name]
        assert [list(e) for e in s["events"]] == t["events"] and (s["fmt"], s["w"], s["h"], s["memory_buffer"]) == (t["fmt"], t["w"], t["h"], t["memory_buffer"])
        assert all(c["key"] == "cam-" + name and c["maxlen"] == s["memory_buffer"] for c in t["xadd"])
        assert all(c["fields"]["device_id"] == "" for c in t["xadd"])          # field 12 is never set (read_image.py:99-111)


def _gate_walk(events):
    """(source packet, frame, packet, keyframe, is_keyframe, is_corrupt) per published frame, and decode() calls per packet."""
    g = DecodeGate()
    group, out, calls, per_packet = [], [], [0] * len(events), []
    for n, (kf, corrupt, kfo, q, nframes) in enumerate(events):
        if kf:
            group = []
        group.append((n, nframes))
        before = len(out)
        for o in g.on_packet(kf, corrupt, kfo, q):
            src, nf = group[o.group_index]
            calls[src] += 1
            for f in range(nf):
                out.append((src, f, o.packet, o.keyframe, o.is_keyframe, o.is_corrupt))
                if o.first_frame_only:
                    break
                g.frame_published(kfo, q)
            g.packet_done()
        per_packet.append(len(out) - before)
    return out, calls, per_packet


def _trace_rows(t):
    return [(c["fields"]["pts"] // 1000, c["fields"]["pts"] % 1000, c["fields"]["packet"], c["fields"]["keyframe"],
             c["fields"]["is_keyframe"], c["fields"]["is_corrupt"]) for c in t["xadd"]]


def test_decode_gate_reproduces_the_reference_run(trace):
    for name, t in trace.items():
        events = [tuple(e) for e in t["events"]]
        out, calls, per_packet = _gate_walk(events)
        assert out == _trace_rows(t), name
        assert calls == t["decode_calls"], name                  # which packets were decoded, and how often
        assert per_packet == t["published_per_packet"], name     # and at which wake-up each message appeared


def test_gate_ref_restatement_is_pinned_by_the_same_trace(trace):
    for name, t in trace.items():
        got = [(p["source_packet"], p["frame"], p["packet"], p["keyframe"], p["is_keyframe"], p["is_corrupt"])
               for p in gate_ref.run([tuple(e) for e in t["events"]])]
        assert got == _trace_rows(t), name


def test_idle_gops_are_skipped_and_activity_replays_the_gop_from_its_key_frame():
    """The producer side of the gate (rtsp_to_rtmp.py:141-160) as FrameProducer models it, with a FramePath stand-in."""
    from video_edge_ai_proxy_b200 import CameraControl, FrameProducer

    class _Path:
        def __init__(self):
            self.staged = []

        def open_stream(self, *a):
            pass

        def submit(self, device_id, planes, meta, pinned=False):
            self.staged.append((meta.pts, meta.packet, meta.keyframe, bool(meta.is_keyframe)))

        def tick(self):
            return 0

    def run(active_at, n=24, gop=6):
        path, now = _Path(), [100000]
        prod = FrameProducer(path, None, now_ms=lambda: now[0])
        prod.add_camera("c", "yuv420p", 64, 48)
        pk = [RS.FakePacket("yuv420p", 64, 48, i, i % gop == 0 and i >= 2, False, 1) for i in range(n)]   # starts mid-GOP
        for i, p in enumerate(pk):
            now[0] += 40
            ctl = CameraControl(last_query=now[0] - 1 if active_at(i) else now[0] - 60000)
            prod.on_packet("c", p, ctl)
        return path.staged, [p.decode_calls for p in pk]

    # cold start while idle, a client shows up at packet 9 (mid-GOP 6..11): the GOP is replayed from its key frame 6
    staged, calls = run(lambda i: i >= 9)
    assert staged[0] == (6000, 0, 1, True)               # first GOP the gate ever saw: keyframe counter 1, is_keyframe of packet 6
    assert [s[0] // 1000 for s in staged] == list(range(6, 24))    # 7, 8, 9 follow with the next query (packet 10), then one per packet
    assert staged[1] == (7000, 1, 1, False)
    assert calls[:6] == [0] * 6                          # packets before the first key frame / idle GOP: never decoded
    assert all(c == 1 for c in calls[6:])                # every later packet exactly once
    # GOP 6..11 active, idle through the key frame 12, active again at 15: no IndexError, replay from 12
    staged, calls = run(lambda i: 6 <= i < 10 or i >= 15)
    pts = [s[0] // 1000 for s in staged]
    assert pts == list(range(6, 10)) + list(range(12, 24)), pts
    assert [s for s in staged if s[0] == 12000][0] == (12000, 0, 2, True)
    # a GOP that passes entirely while idle is never seen: its key frame does not count
    staged, calls = run(lambda i: i < 12 or i >= 18)
    assert [s for s in staged if s[0] == 18000][0] == (18000, 0, 2, True) and calls[12:18] == [0] * 6


@pytest.mark.skipif(not L.available(), reason="bundled libswscale not found")
def test_c_packer_and_field_fill_reproduce_the_reference_messages(trace):
    """The whole message, byte for byte: prefix + BGR payload (live libswscale) + suffix from vep_videoframe_frame, with
    the fields filled the way producer.py / read_image.py fill them, against the sha256 the reference published."""
    for name, t in trace.items():
        sc = RS.scenarios()[name]
        packets = RS.packets_of(sc)
        for c in t["xadd"]:
            f = c["fields"]
            frame = packets[f["pts"] // 1000].decode()[f["pts"] % 1000]
            ts = int(frame.time * frame.time_base.denominator)
            meta = make_meta(timestamp=ts, pts=frame.pts, dts=frame.dts, packet=f["packet"], keyframe=f["keyframe"],
                             time_base=float(frame.time_base), is_keyframe=f["is_keyframe"], is_corrupt=f["is_corrupt"],
                             frame_type=frame.pict_type.name)
            pre, suf = videoframe_framing(meta, frame.width, frame.height)
            data = np.ascontiguousarray(frame.to_ndarray(format="bgr24")).tobytes()
            assert hashlib.sha256(data).hexdigest() == f["data_sha256"], (name, f["pts"])
            msg = pre + data + suf
            assert len(msg) == c["len"] and hashlib.sha256(msg).hexdigest() == c["sha256"], (name, f["pts"])
            assert (f["timestamp"], f["frame_type"], f["shape"]) == (ts, frame.pict_type.name, [[frame.height, "0"], [frame.width, "1"], [3, "2"]])
