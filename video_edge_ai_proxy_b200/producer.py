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


MAX_RING_DEPTH = 64     # vep_stream_open's limit; Redis MAXLEN keeps the history beyond it


class _Camera:
    def __init__(self, memory_buffer: int):
        self.gate = DecodeGate()
        self.group: List[object] = []      # packets since the last key frame: the reference's packet queue
        self.fed = 0                       # how many of them the gate has seen (the consumer has popped)
        self.seen_keyframe = False
        self.memory_buffer = memory_buffer
        self.staged = 0                    # frames staged since the last flush
        self.offered = 0
        self.published = 0


class FrameProducer:
    def __init__(self, path: FramePath, redis_conn, now_ms=None, pinned_frames: bool = False):
        """pinned_frames=True: the decoder hands out frames whose planes lie in page-locked memory (its buffer pool
        was registered with FramePath.register_host); they are submitted without a CPU copy (VEP_SUBMIT_PINNED) and
        this object keeps a reference to every such frame until the tick that converts it has finished."""
        self.path = path
        self.redis = redis_conn
        self._cams: Dict[str, _Camera] = {}
        self._now_ms = now_ms or (lambda: int(round(time.time() * 1000)))
        self._pinned = bool(pinned_frames)
        self._held: List[object] = []      # frames the GPU may still read (pinned_frames only)

    def add_camera(self, device_id: str, fmt: str, width: int, height: int, memory_buffer: int = 1):
        # the library keeps min(memory_buffer, 64) frames readable; XADD MAXLEN carries the full value
        self.path.open_stream(device_id, fmt, width, height, max(1, min(int(memory_buffer), MAX_RING_DEPTH)))
        self._cams[device_id] = _Camera(memory_buffer)

    def add_cameras(self, cameras, memory_buffer: int = 1):
        """Several cameras at once — [(device_id, fmt, width, height), ...] — with shared slot memory, so that a fleet
        whose cameras convert in lock step is copied to and from the GPU with one copy per direction per tick
        (FramePath.open_streams)."""
        self.path.open_streams(list(cameras), max(1, min(int(memory_buffer), MAX_RING_DEPTH)))
        for device_id, _fmt, _w, _h in cameras:
            self._cams[device_id] = _Camera(memory_buffer)

    def on_packet(self, device_id: str, packet, control: Optional[CameraControl] = None) -> int:
        """Feed one demuxed packet of a camera.  Returns how many frames were staged for the next
        flush().  `control` carries the camera's Redis control state (gate.poll_control).

        Queue semantics of the reference's producer thread: packets before the first key frame are dropped
        (rtsp_to_rtmp.py:111-113), a key frame empties the queue (:155-157), every packet is queued (:160) but
        the consumer is woken only while a client is listening (:141-153).  So while a camera is idle its GOP
        piles up un-popped, and when a client comes back the consumer works through it FROM THE KEY FRAME —
        which is what happens here on the first active packet: every packet of the current GOP the gate has
        not seen yet is fed to it, oldest first.  GOPs that pass entirely while idle are never seen by the
        gate (their key frames do not count in `VideoFrame.keyframe`), as in the reference."""
        cam = self._cams[device_id]
        control = control or CameraControl()
        cam.offered += 1
        if packet.is_keyframe:
            cam.group, cam.fed, cam.seen_keyframe = [], 0, True
        if not cam.seen_keyframe:
            return 0
        cam.group.append(packet)
        if not camera_active(control.last_query, self._now_ms()):
            return 0
        staged = 0
        while cam.fed < len(cam.group):
            popped = cam.group[cam.fed]
            cam.fed += 1
            orders = cam.gate.on_packet(popped.is_keyframe, popped.is_corrupt, control.key_frame_only, control.last_query)
            for o in orders:
                staged += self._decode_and_stage(device_id, cam, o, control)
                cam.gate.packet_done()
        return staged

    def _decode_and_stage(self, device_id: str, cam: _Camera, o, control: CameraControl) -> int:
        staged = 0
        for frame in cam.group[o.group_index].decode() or ():
            ts = self._now_ms()
            if getattr(frame, "time", None) is not None:
                ts = int(frame.time * frame.time_base.denominator)      # read_image.py:89-91
            meta = make_meta(timestamp=ts, pts=frame.pts, dts=frame.dts, packet=o.packet,
                             keyframe=o.keyframe, time_base=float(frame.time_base),
                             is_keyframe=o.is_keyframe, is_corrupt=o.is_corrupt,
                             frame_type=getattr(frame.pict_type, "name", frame.pict_type))
            try:
                self.path.submit(device_id, frame.planes, meta, pinned=self._pinned)
            except VepError as e:
                if e.status != VEP_ERR_BUSY:
                    raise
                self.flush()      # a long catch-up burst filled this camera's staging ring:
                self.path.submit(device_id, frame.planes, meta, pinned=self._pinned)   # convert + publish, then go on
            if self._pinned:
                self._held.append(frame)
            cam.staged += 1
            staged += 1
            if o.first_frame_only:
                break
            cam.gate.frame_published(control.key_frame_only, control.last_query)
        return staged

    def flush(self) -> int:
        """One tick: convert all staged frames (one launch) and publish them, per camera oldest-first, each with
        the reference's call `xadd(device_id, {'data': vfData}, maxlen=memory_buffer)` (read_image.py:121) —
        queued on ONE non-transactional pipeline and sent with one `execute()`, i.e. one round trip per tick
        instead of one per message.  Connections without `pipeline()` get the direct calls."""
        n = self.path.tick()
        self._held.clear()       # the tick has read every submitted plane
        make_pipe = getattr(self.redis, "pipeline", None)
        sink = make_pipe(transaction=False) if make_pipe is not None else self.redis
        queued = 0
        for device_id, cam in self._cams.items():
            for age in range(cam.staged - 1, -1, -1):
                msg = self.path.result_packed(device_id, age)      # zero-copy view of the pinned ring
                sink.xadd(device_id, {"data": msg}, maxlen=cam.memory_buffer)
                cam.published += 1
                queued += 1
            cam.staged = 0
        if make_pipe is not None and queued:
            sink.execute()       # the views stay valid until the next tick recycles their slots
        return n
