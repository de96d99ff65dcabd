"""Decode gate: which packets of a camera get decoded and converted (host logic, stays in Python).

Mirrors the selection logic of reference python/read_image.py:57-128 and its producer side
python/rtsp_to_rtmp.py:117-157, as INTENDED: at the surveyed snapshot the two modules each import
their own copy of `query_timestamp` (global_vars.py:15), so the consumer compares `None > 0` and
drops every frame.  The rule implemented here is the one the code spells out:

  * the first packet of every GOP is always decoded            (read_image.py:80, len(group) == 1)
  * later packets only if a client asked since the last conversion (:73-74) and key-frame-only
    mode is off (:77-78); then every not-yet-decoded packet of the GOP is decoded in order (:81-85)
  * in key-frame-only mode only the first frame of the packet is taken (:123-124)
  * a camera is active while its last client query is younger than 10 s (rtsp_to_rtmp.py:141-145)

Redis key names come from python/global_vars.py:16-17 / server/models/RedisConstants.go:18-28.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, Iterable, List, Optional

REDIS_LAST_ACCESS_PREFIX = "last_access_time_"       # hash: last_query, proxy_rtmp, store
REDIS_KEY_FRAME_ONLY_PREFIX = "is_key_frame_only_"   # string "true"/"false"
ACTIVITY_WINDOW_MS = 10000                           # rtsp_to_rtmp.py:145


@dataclass
class DecodeOrder:
    """One packet the caller must decode and convert now."""
    group_index: int        # position of the packet inside its GOP (what gets decoded)
    packet: int             # VideoFrame.packet   (read_image.py:107)
    keyframe: int           # VideoFrame.keyframe (:108)
    is_keyframe: bool       # flags of the packet that TRIGGERED the burst (:110-111)
    is_corrupt: bool
    first_frame_only: bool  # key-frame-only mode: take one frame of this packet (:123-124)


class DecodeGate:
    """Per-camera state machine equivalent to the bookkeeping in ReadImage.run (read_image.py:50-128).

    Protocol per popped packet: `orders = on_packet(...)`; for each order decode that packet of the GOP,
    call `frame_published()` after every published frame and `packet_done()` once the packet's frames are
    through.  `packet_done()` is what advances the count of decoded packets — as in the reference, where
    `packet_count` grows only after a packet went through without an exception (:128), so a packet whose
    decode or conversion failed is tried again at the next burst instead of being lost."""

    def __init__(self):
        self.group_len = 0            # len(packet_group)
        self.decoded = 0              # packet_count: packets of this GOP already decoded
        self.keyframes = 0            # keyframes_count
        self.last_query_timestamp = 0

    def on_packet(self, is_keyframe: bool, is_corrupt: bool = False, key_frame_only: bool = False,
                  query_timestamp: Optional[int] = None) -> List[DecodeOrder]:
        if is_keyframe:
            self.group_len = 0
            self.decoded = 0
            self.keyframes += 1
        self.group_len += 1
        wanted = (query_timestamp is not None and query_timestamp > self.last_query_timestamp
                  and not key_frame_only)
        if self.group_len != 1 and not wanted:
            return []
        return [DecodeOrder(gi, gi, self.keyframes, bool(is_keyframe), bool(is_corrupt), bool(key_frame_only))
                for gi in range(self.decoded, self.group_len)]

    def packet_done(self):
        """One ordered packet went through (read_image.py:128)."""
        self.decoded += 1

    def frame_published(self, key_frame_only: bool, query_timestamp: Optional[int]):
        """Call once per published frame (read_image.py:126): remembers the query that was served.
        In key-frame-only mode the reference breaks out before this line (:123-124)."""
        if not key_frame_only and query_timestamp is not None:
            self.last_query_timestamp = query_timestamp


def camera_active(last_query_ms: Optional[int], now_ms: int) -> bool:
    """rtsp_to_rtmp.py:141-145: decode only while a client queried within the last 10 s."""
    return last_query_ms is not None and (now_ms - last_query_ms) < ACTIVITY_WINDOW_MS


@dataclass
class CameraControl:
    last_query: Optional[int] = None
    key_frame_only: bool = False
    proxy_rtmp: bool = False


def poll_control(redis_conn, device_ids: Iterable[str]) -> Dict[str, CameraControl]:
    """One pipelined read of the control keys of ALL cameras (the reference does two round trips
    per packet per camera: rtsp_to_rtmp.py:117 HGETALL, read_image.py:40 GET)."""
    ids = list(device_ids)
    pipe = redis_conn.pipeline()
    for d in ids:
        pipe.hgetall(REDIS_LAST_ACCESS_PREFIX + d)
        pipe.get(REDIS_KEY_FRAME_ONLY_PREFIX + d)
    res = pipe.execute()
    out = {}
    dec = lambda v: v.decode("utf-8") if isinstance(v, (bytes, bytearray)) else v
    for i, d in enumerate(ids):
        h, k = res[2 * i] or {}, res[2 * i + 1]
        h = {dec(a): dec(b) for a, b in h.items()}
        ctl = CameraControl()
        if "last_query" in h:
            ctl.last_query = int(h["last_query"])
        ctl.proxy_rtmp = h.get("proxy_rtmp") == "1"
        ctl.key_frame_only = k is not None and dec(k).lower() == "true"
        out[d] = ctl
    return out
