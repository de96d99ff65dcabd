"""Consumer side of the frame path: `Image.VideoLatestImage` as a pass-through of the already packed bytes.

Reference: server/grpcapi/grpc_api.go:133-235 (Go; no Go toolchain in this image, so the host side is Python
over grpcio, which the image has).  Per request the reference
  * SETs  is_key_frame_only_<device_id>            = "true" / "false"            (:159-160)
  * HSETs last_access_time_<device_id>.last_query  = now in ms                   (:166-171)
  * XREADs the camera's stream from the last id it saw (Block 1 s, Count 60), up to 3 tries 16 ms apart, and keeps
    the LAST message (:187-229)
  * proto.Unmarshal()s its `data` field into a VideoFrame (:212-216) and stream.Send()s it (:231) — two more copies
    of a 6 MB payload, through a server whose gRPC message limits are the 4 MiB defaults, which a 1080p frame
    exceeds.
Here the bytes the producer wrote (FramePath.result_packed -> XADD) are forwarded UNTOUCHED: the response
serializer of the handler is the identity, so nothing is parsed or re-encoded, and the server's send limit is
raised so that 1080p and 4K frames fit.  When no frame turns up, the reference sends an empty VideoFrame (:179,
:231) — an empty message on the wire, which is what is sent here as well.

Registered under the reference's service name (proto/video_streaming.proto:140-146), so the reference's own
generated client stub (examples/basic_usage.py:50-62) talks to it unchanged.  Only VideoLatestImage is served; the
other four RPCs are control plane (out of scope).
"""
from __future__ import annotations

import threading
import time
from concurrent import futures
from typing import Iterable, Optional

from .gate import REDIS_KEY_FRAME_ONLY_PREFIX, REDIS_LAST_ACCESS_PREFIX

SERVICE = "chrys.cloud.videostreaming.v1beta1.Image"
METHOD = "VideoLatestImage"
MAX_MESSAGE_BYTES = 64 << 20          # 4K BGR24 is 24.9 MB; the reference leaves gRPC's 4 MiB default in place
XREAD_BLOCK_MS = 1000                 # grpc_api.go:192
XREAD_COUNT = 60                      # :193
XREAD_TRIES = 3                       # :186
XREAD_RETRY_S = 0.016                 # :227
STREAM_DEADLINE_S = 15.0              # :135


def parse_video_frame_request(data: bytes):
    """(key_frame_only, device_id) from a serialized VideoFrameRequest (proto/video_streaming.proto:95-98):
    field 1 = bool key_frame_only (varint), field 2 = string device_id.  Unknown fields are skipped."""
    key_frame_only, device_id = False, ""
    i, n = 0, len(data)

    def varint(i):
        v, shift = 0, 0
        while True:
            if i >= n:
                raise ValueError("truncated varint")
            b = data[i]
            i += 1
            v |= (b & 0x7F) << shift
            if not b & 0x80:
                return v, i
            shift += 7
            if shift > 63:
                raise ValueError("varint too long")

    while i < n:
        tag, i = varint(i)
        field, wt = tag >> 3, tag & 7
        if wt == 0:
            v, i = varint(i)
            if field == 1:
                key_frame_only = bool(v)
        elif wt == 2:
            ln, i = varint(i)
            if i + ln > n:
                raise ValueError("truncated field")
            if field == 2:
                device_id = bytes(data[i:i + ln]).decode("utf-8")
            i += ln
        elif wt == 1:
            i += 8
        elif wt == 5:
            i += 4
        else:
            raise ValueError(f"unsupported wire type {wt}")
    return key_frame_only, device_id


class ImageService:
    """The handler object; `redis_conn` needs `set`, `hset` and `xread` (redis-py's signatures)."""

    def __init__(self, redis_conn, now_ms=None):
        self.redis = redis_conn
        self._now_ms = now_ms or (lambda: int(time.time() * 1000))
        self._last_id = {}                       # device_id -> id of the last message forwarded (grpc_api.go:40, :181-184)
        self._lock = threading.Lock()
        self.forwarded_bytes = 0

    # ---- one request (grpc_api.go:155-231) -----------------------------------------------------------------
    def latest(self, device_id: str, key_frame_only: bool) -> bytes:
        """The newest packed VideoFrame of `device_id` not forwarded yet, as it lies in Redis; b"" if none came."""
        try:
            self.redis.set(REDIS_KEY_FRAME_ONLY_PREFIX + device_id, "true" if key_frame_only else "false")     # :159-160
        except Exception:
            pass                                                                                                # logged and ignored there
        try:
            self.redis.hset(REDIS_LAST_ACCESS_PREFIX + device_id, mapping={"last_query": self._now_ms()})      # :166-171
        except Exception:
            return b""                                                                                          # `continue` there: nothing is sent
        with self._lock:
            last = self._last_id.get(device_id, "0")
        for attempt in range(XREAD_TRIES):
            try:
                res = self.redis.xread({device_id: last}, count=XREAD_COUNT, block=XREAD_BLOCK_MS)
            except Exception:
                res = None
            for _stream, messages in res or ():
                if not messages:
                    continue
                msg_id, fields = messages[-1]                                                                   # only the latest image (:207)
                with self._lock:
                    self._last_id[device_id] = msg_id.decode() if isinstance(msg_id, (bytes, bytearray)) else msg_id
                data = fields.get("data", fields.get(b"data"))
                if data is not None:
                    return bytes(data) if not isinstance(data, bytes) else data                                 # forwarded as is: no Unmarshal
            if attempt + 1 < XREAD_TRIES:
                time.sleep(XREAD_RETRY_S)
        return b""

    # ---- the bidirectional stream (:133-153) ---------------------------------------------------------------
    def VideoLatestImage(self, request_iterator: Iterable, context) -> Iterable[bytes]:
        deadline = time.time() + STREAM_DEADLINE_S
        for key_frame_only, device_id in request_iterator:
            if time.time() > deadline:
                return
            out = self.latest(device_id, key_frame_only)
            self.forwarded_bytes += len(out)
            yield out

    def handler(self):
        import grpc
        return grpc.method_handlers_generic_handler(SERVICE, {
            METHOD: grpc.stream_stream_rpc_method_handler(
                self.VideoLatestImage,
                request_deserializer=parse_video_frame_request,
                response_serializer=lambda packed: packed)})         # identity: the packed message IS the wire format


def serve(redis_conn, address: str = "0.0.0.0:50001", max_workers: int = 8, service: Optional[ImageService] = None):
    """Start a grpc server with the pass-through handler (the reference listens on 0.0.0.0:50001, server/main.go:143).
    Returns (server, bound port, service)."""
    import grpc
    svc = service or ImageService(redis_conn)
    server = grpc.server(futures.ThreadPoolExecutor(max_workers=max_workers),
                         options=[("grpc.max_send_message_length", MAX_MESSAGE_BYTES),
                                  ("grpc.max_receive_message_length", 4 << 20)])
    server.add_generic_rpc_handlers((svc.handler(),))
    port = server.add_insecure_port(address)
    server.start()
    return server, port, svc
