"""Frame producer: the replacement for the body of `ReadImage.run` (reference
python/read_image.py:57-128) serving MANY cameras from one process.

Per packet the reference pops it, consults the gate, decodes, converts, packs and XADDs one frame
at a time.  Here packets of all cameras are fed in, the same gate decides what is decoded, decoded
frames are staged through `FramePath.submit`, and `flush()` converts everything staged with one
kernel launch and publishes each message with the reference's Redis call
`xadd(device_id, {'data': <VideoFrame bytes>}, maxlen=memory_buffer)` (:121).

Decoding itself (libavcodec, :87) stays outside: packets are duck-typed — `.is_keyframe`,
`.is_corrupt`, `.decode()` returning frame objects with `.planes` (list of 2-D uint8 arrays),
`.width`, `.height`, `.pts`, `.dts`, `.time`, `.time_base`, `.pict_type` (str) — which is the subset
of av.Packet / av.VideoFrame the reference touches.
"""
from __future__ import annotations

import time
from typing import Dict, List, Optional

from ._abi import VEP_ERR_BUSY, VepError
from .frame_path import FramePath, make_meta
from .gate import CameraControl, DecodeGate, camera_active


class _Camera:
    def __init__(self, memory_buffer: int):
        self.gate = DecodeGate()
        self.group: List[object] = []      # packets of the current GOP (needed for catch-up bursts)
        self.memory_buffer = memory_buffer
        self.staged = 0                    # frames staged since the last flush
        self.offered = 0
        self.published = 0


class FrameProducer:
    def __init__(self, path: FramePath, redis_conn, now_ms=None):
        self.path = path
        self.redis = redis_conn
        self._cams: Dict[str, _Camera] = {}
        self._now_ms = now_ms or (lambda: int(round(time.time() * 1000)))

    def add_camera(self, device_id: str, fmt: str, width: int, height: int, memory_buffer: int = 1):
        self.path.open_stream(device_id, fmt, width, height, memory_buffer)
        self._cams[device_id] = _Camera(memory_buffer)

    def on_packet(self, device_id: str, packet, control: Optional[CameraControl] = None) -> int:
        """Feed one demuxed packet of a camera.  Returns how many frames were staged for the next
        flush().  `control` carries the camera's Redis control state (gate.poll_control)."""
        cam = self._cams[device_id]
        control = control or CameraControl()
        cam.offered += 1
        if packet.is_keyframe:
            cam.group = []
        cam.group.append(packet)
        # rtsp_to_rtmp.py:141-153: packets reach the decoder only while a client is listening
        if not camera_active(control.last_query, self._now_ms()):
            # the gate still has to see the packet so GOP bookkeeping stays aligned, but with no
            # live query nothing beyond the bookkeeping happens (the reference never pops it)
            return 0
        orders = cam.gate.on_packet(packet.is_keyframe, packet.is_corrupt, control.key_frame_only,
                                    control.last_query)
        staged = 0
        for o in orders:
            for frame in cam.group[o.group_index].decode() or ():
                ts = self._now_ms()
                if getattr(frame, "time", None) is not None:
                    ts = int(frame.time * frame.time_base.denominator)      # read_image.py:89-91
                meta = make_meta(timestamp=ts, pts=frame.pts, dts=frame.dts, packet=o.packet,
                                 keyframe=o.keyframe, time_base=float(frame.time_base),
                                 is_keyframe=o.is_keyframe, is_corrupt=o.is_corrupt,
                                 frame_type=frame.pict_type)
                try:
                    self.path.submit(device_id, frame.planes, meta)
                except VepError as e:
                    if e.status != VEP_ERR_BUSY:
                        raise
                    self.flush()      # a long catch-up burst filled this camera's staging ring:
                    self.path.submit(device_id, frame.planes, meta)   # convert + publish, then go on
                cam.staged += 1
                staged += 1
                if o.first_frame_only:
                    break
                cam.gate.frame_published(control.key_frame_only, control.last_query)
        return staged

    def flush(self) -> int:
        """One tick: convert all staged frames (one launch) and publish them oldest-first."""
        n = self.path.tick()
        for device_id, cam in self._cams.items():
            for age in range(cam.staged - 1, -1, -1):
                msg = self.path.result_packed(device_id, age)
                self.redis.xadd(device_id, {"data": msg}, maxlen=cam.memory_buffer)
                cam.published += 1
            cam.staged = 0
        return n
