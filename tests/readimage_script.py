"""Scripted packet sequences for pinning the decode gate and the per-frame VideoFrame fields to the
reference's own `ReadImage.run` (python/read_image.py:47-133).

The same scripts are played (a) through the reference class itself by tests/golden/make_readimage_trace.py,
whose recorded XADD payloads are committed as tests/golden/readimage_trace.json, and (b) through this
repo's DecodeGate / ReadImage / FrameProducer by the tests.  A script is a list of events, one per packet
the consumer pops: (is_keyframe, is_corrupt, key_frame_only, query_timestamp, frames_in_packet).
Packet n decodes to frames whose planes are `make_planes(fmt, w, h, seed=100*n + f)` and whose
pts is 1000*n + f, so every published message names the packet and frame it came from.
"""
from fractions import Fraction

from video_edge_ai_proxy_b200.synth import make_planes

TIME_BASE = Fraction(1, 90000)


class _Named:
    def __init__(self, name):
        self.name = name


class FakeFrame:
    """The subset of av.VideoFrame the reference touches (read_image.py:89-109), plus `.planes` as arrays."""

    def __init__(self, fmt, w, h, packet_index, frame_index, key):
        self.format = _Named(fmt)
        self.width, self.height = w, h
        self.planes = list(make_planes(fmt, w, h, 100 * packet_index + frame_index))
        self.pts = 1000 * packet_index + frame_index
        self.dts = self.pts - 1
        self.time_base = TIME_BASE
        self.time = self.pts / 90000
        self.pict_type = _Named("I" if key and frame_index == 0 else ("B" if frame_index % 2 else "P"))

    def to_ndarray(self, format="bgr24"):          # what PyAV does here: sws_getCachedContext + sws_scale
        assert format == "bgr24"
        from oracle import swscale_live            # only the golden generator (and tests) come here
        return swscale_live.to_bgr24(self.format.name, self.planes, self.width, self.height)


class FakePacket:
    """The subset of av.Packet the reference touches (:63, :87, :110-111)."""

    def __init__(self, fmt, w, h, index, is_keyframe, is_corrupt, nframes):
        self.index, self.is_keyframe, self.is_corrupt = index, bool(is_keyframe), bool(is_corrupt)
        self._args, self._nframes = (fmt, w, h), nframes
        self.decode_calls = 0

    def decode(self):
        self.decode_calls += 1
        return [FakeFrame(*self._args, self.index, f, self.is_keyframe) for f in range(self._nframes)]


def _gop(n, gop, query=lambda i: 1000, kfo=lambda i: False, corrupt=lambda i: False, frames=lambda i: 1):
    return [(i % gop == 0, corrupt(i), kfo(i), query(i), frames(i)) for i in range(n)]


def scenarios():
    """name -> dict(fmt, w, h, memory_buffer, events)."""
    S = {}
    # one query before the stream starts, never renewed: only the key frames are converted (read_image.py:80)
    S["no_new_query"] = dict(fmt="yuv420p", w=64, h=48, memory_buffer=1, events=_gop(14, 6))
    # a client query arrives mid-GOP: every not yet decoded packet of the GOP is decoded in order (:81-85),
    # the query is served once (:126), the next one triggers the next burst
    S["query_mid_gop_burst"] = dict(fmt="yuv420p", w=64, h=48, memory_buffer=3,
                                    events=_gop(16, 8, query=lambda i: 1000 if i < 5 else (2000 if i < 11 else 3000 + i)))
    # key-frame-only on, then off mid-GOP with a fresh query, then on again (:77-78, :123-124)
    S["keyframe_only_toggle"] = dict(fmt="yuv420p", w=64, h=48, memory_buffer=1,
                                     events=_gop(20, 5, query=lambda i: 1000 + 10 * i, kfo=lambda i: i < 7 or i >= 14))
    # packets that decode to 0, 1, 2 or 3 frames; key-frame-only takes the first frame of a packet only (:123-124)
    S["multi_frame_packets"] = dict(fmt="yuv420p", w=64, h=48, memory_buffer=3,
                                    events=_gop(12, 4, query=lambda i: 1000 + i, frames=lambda i: (2, 1, 0, 3)[i % 4],
                                                kfo=lambda i: i >= 8))
    # corrupt flags are those of the packet just popped, also for the older packets of a burst (:110-111)
    S["corrupt_flags"] = dict(fmt="yuv420p", w=64, h=48, memory_buffer=3,
                              events=_gop(12, 6, query=lambda i: 1000 if i < 4 else 5000 + (i >= 9), corrupt=lambda i: i in (2, 4, 9)))
    # every packet queried (a client polling faster than the frame rate), NV12 and full-range planar input
    S["every_packet_nv12"] = dict(fmt="nv12", w=64, h=16, memory_buffer=3, events=_gop(7, 3, query=lambda i: 1000 + i))
    S["every_packet_yuvj420p"] = dict(fmt="yuvj420p", w=50, h=6, memory_buffer=1, events=_gop(5, 5, query=lambda i: 1000 + i))
    return S


def packets_of(sc):
    return [FakePacket(sc["fmt"], sc["w"], sc["h"], n, kf, corrupt, nframes)
            for n, (kf, corrupt, _kfo, _q, nframes) in enumerate(sc["events"])]
