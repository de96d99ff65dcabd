"""Drop-in for reference python/read_image.py: class `ReadImage` with the reference's constructor,
thread protocol and Redis side effects, whose per-frame work (reference lines 94-119: convert to
BGR24, tobytes, fill and serialize VideoFrame) runs on the B200 through libvepb200.so.

    from video_edge_ai_proxy_b200.read_image import ReadImage      # instead of `from read_image import ReadImage`
    th = ReadImage(packet_queue, device_id, memory_buffer, redis_conn, is_decode_packets_event, lock_condition)

Everything the reference's `rtsp_to_rtmp.py:191-251` main does with the object keeps working: it is a
`threading.Thread`, it waits on `lock_condition`, pops one packet per wake-up while the event is set,
reads `is_key_frame_only_<device_id>` from Redis, decodes with `packet.decode()`, and publishes with
`redis_conn.xadd(device_id, {'data': vfData}, maxlen=memory_buffer)`.  Differences, all deliberate:
`query_timestamp` is read from the `global_vars` MODULE (the reference reads a stale by-value copy,
SURVEY §3.2); several ReadImage threads of one process may share one `FramePath` (pass `frame_path=`)
so that their frames can be converted in one launch by whoever ticks; a conversion error drops the
frame with the reference's message and the loop continues (read_image.py:130-131).
"""
from __future__ import annotations

import threading
import time

import numpy as np

from . import global_vars
from .frame_path import FramePath, make_meta
from .gate import DecodeGate

_SUPPORTED = ("yuv420p", "yuvj420p", "nv12", "nv21", "yuv422p", "yuvj422p", "yuv420p10le", "yuv444p", "yuvj444p", "gray")


def planes_of(frame):
    """2-D uint8 views of a decoded frame's planes.  Accepts PyAV `av.VideoFrame` (planes expose the
    buffer protocol, `.line_size`, `.height`, `.width`) or anything whose `.planes` already are 2-D arrays."""
    out = []
    for p in frame.planes:
        if isinstance(p, np.ndarray):
            out.append(p)
            continue
        a = np.frombuffer(p, np.uint8)
        ls = int(p.line_size)
        rows = int(getattr(p, "height", a.size // ls))
        a = a[:rows * ls].reshape(rows, ls)
        if _name(getattr(frame.format, "name", frame.format)) == "yuv420p10le":
            a = a.view(np.uint16)                       # 16-bit little-endian samples
        out.append(a[:, :int(p.width)])
    return out


def _name(x):
    return x if isinstance(x, str) else getattr(x, "name", str(x))


class ReadImage(threading.Thread):

    def __init__(self, packet_queue, device_id, memory_buffer, redis_conn, is_decode_packets_event, lock_condition,
                 frame_path: FramePath = None, device: int = 0):
        threading.Thread.__init__(self)
        self._packet_queue = packet_queue
        self.device_id = device_id
        self._memory_buffer = memory_buffer
        self.redis_conn = redis_conn
        self.is_decode_packets_event = is_decode_packets_event
        self.lock_condition = lock_condition
        self.last_query_timestamp = 0
        self.packet_group = []
        self._path = frame_path
        self._own_path = frame_path is None
        self._device = device
        self._opened = None          # (fmt, w, h) the stream was opened with
        self._gate = DecodeGate()
        self._stop_flag = threading.Event()

    # checks if only keyframes requested (reference read_image.py:36-45)
    def check_decode_only_keyframes(self):
        decode_only_keyframes = False
        only_keyframes = self.redis_conn.get(global_vars.RedisIsKeyFrameOnlyPrefix + self.device_id)
        if only_keyframes is not None:
            okeys = only_keyframes.decode("utf-8") if isinstance(only_keyframes, (bytes, bytearray)) else str(only_keyframes)
            if okeys.lower() == "true":
                decode_only_keyframes = True
        return decode_only_keyframes

    def stop(self):
        """Not in the reference (its thread runs until the container dies); lets tests and hosts end the loop."""
        self._stop_flag.set()
        with self.lock_condition:
            self.lock_condition.notify_all()

    def _ensure_stream(self, fmt, w, h):
        if self._path is None:
            self._path = FramePath(self._device)
        if self._opened != (fmt, w, h):
            if fmt not in _SUPPORTED:
                raise ValueError(f"pixel format {fmt!r} is not supported by the B200 frame path")
            if self._opened is not None:
                self._path.close_stream(self.device_id)     # the camera changed resolution/format mid-stream
            # the library keeps at most 64 frames readable per camera; Redis MAXLEN holds the history beyond that
            self._path.open_stream(self.device_id, fmt, w, h, min(max(int(self._memory_buffer), 1), 64))
            self._opened = (fmt, w, h)

    def process_packet(self, packet, query_timestamp=None):
        """Body of the reference loop for one popped packet (read_image.py:61-128).  `query_timestamp`: this camera's
        last client query; by default the module global the reference's producer thread sets (one camera per process
        there) — a process that serves several cameras passes each camera's own value."""
        decode_only_keyframes = self.check_decode_only_keyframes()
        if packet.is_keyframe:
            self.packet_group = []
        self.packet_group.append(packet)
        if query_timestamp is None:
            query_timestamp = global_vars.query_timestamp
        orders = self._gate.on_packet(packet.is_keyframe, packet.is_corrupt, decode_only_keyframes, query_timestamp)
        for o in orders:
            for frame in self.packet_group[o.group_index].decode() or ():
                timestamp = int(round(time.time() * 1000))
                if frame.time is not None:
                    timestamp = int(frame.time * frame.time_base.denominator)
                fmt = _name(getattr(frame.format, "name", frame.format))
                self._ensure_stream(fmt, frame.width, frame.height)
                meta = make_meta(timestamp=timestamp, pts=frame.pts, dts=frame.dts, packet=o.packet, keyframe=o.keyframe,
                                 time_base=float(frame.time_base), is_keyframe=o.is_keyframe, is_corrupt=o.is_corrupt,
                                 frame_type=_name(frame.pict_type))
                self._path.submit(self.device_id, planes_of(frame), meta)
                self._path.tick()
                vfData = self._path.result_packed(self.device_id)      # == vf.SerializeToString() (:119), zero-copy
                self.redis_conn.xadd(self.device_id, {'data': vfData}, maxlen=self._memory_buffer)
                if decode_only_keyframes:
                    break
                self._gate.frame_published(decode_only_keyframes, query_timestamp)
            self._gate.packet_done()          # counted only once it went through (read_image.py:128)

    def run(self):
        while not self._stop_flag.is_set():
            with self.lock_condition:
                self.lock_condition.wait(timeout=0.5)
                if self._stop_flag.is_set():
                    break
                if not self._packet_queue.empty() and self.is_decode_packets_event.is_set():
                    try:
                        packet = self._packet_queue.get()
                        self.process_packet(packet)
                    except Exception as e:
                        print("failed to deode packet", e)
                    finally:
                        self._packet_queue.task_done()
        if self._own_path and self._path is not None:
            self._path.close()
