"""TEST INFRASTRUCTURE ONLY — proto3 wire encoding of the reference's `VideoFrame`.

Restates what `vf.SerializeToString()` at reference `python/read_image.py:119` emits for the
message filled at `:99-117`, whose schema is reference `proto/video_streaming.proto:67-93`:
fields in field-number order, scalar fields equal to their proto3 default omitted, int64 as
varint (negative values sign-extended to 10 bytes), `time_base` as little-endian fixed64 double,
`shape` as a length-delimited ShapeProto holding `repeated Dim dim = 2`, `device_id` (12) never
set.  Pinned by tests/golden/videoframe_wire.json, which was produced in this container by the
reference's own generated module `python/proto/video_streaming_pb2.py`
(tests/golden/make_golden.py), and cross-checked against the installed protobuf runtime through a
dynamically built descriptor (`dynamic_message_class`).
"""
from __future__ import annotations

import struct


def varint(v: int) -> bytes:
    if v < 0:
        v += 1 << 64
    out = bytearray()
    while True:
        b = v & 0x7F
        v >>= 7
        if v:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _tag(field: int, wt: int) -> bytes:
    return varint((field << 3) | wt)


def _int(field: int, v: int) -> bytes:
    return b"" if v == 0 else _tag(field, 0) + varint(v)


def _bool(field: int, v: bool) -> bytes:
    return _tag(field, 0) + b"\x01" if v else b""


def _bytes(field: int, v: bytes) -> bytes:
    return b"" if len(v) == 0 else _tag(field, 2) + varint(len(v)) + v


def _double(field: int, v: float) -> bytes:
    raw = struct.pack("<d", v)
    return b"" if raw == b"\x00" * 8 else _tag(field, 1) + raw


def shape_proto(dims) -> bytes:
    """ShapeProto with one Dim{size=dim, name=str(i)} per entry (read_image.py:113-117)."""
    body = b""
    for i, d in enumerate(dims):
        dim = _int(1, int(d)) + _bytes(2, str(i).encode())
        body += _tag(2, 2) + varint(len(dim)) + dim
    return body


def serialize_videoframe(*, width, height, data, timestamp, is_keyframe, pts, dts, frame_type,
                         is_corrupt, time_base, shape, packet, keyframe) -> bytes:
    sp = shape_proto(shape)
    return b"".join([
        _int(1, width), _int(2, height), _bytes(3, data), _int(4, timestamp),
        _bool(5, is_keyframe), _int(6, pts), _int(7, dts), _bytes(8, frame_type.encode()),
        _bool(9, is_corrupt), _double(10, time_base),
        (_tag(11, 2) + varint(len(sp)) + sp) if len(shape) else b"",
        _int(13, packet), _int(14, keyframe),
    ])


def dynamic_message_class():
    """Build VideoFrame/ShapeProto message classes in the installed protobuf runtime from a
    descriptor restating proto/video_streaming.proto:67-93 (no generated code, no protoc)."""
    from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

    F = descriptor_pb2.FieldDescriptorProto
    fd = descriptor_pb2.FileDescriptorProto()
    fd.name = "vep_oracle_video_streaming.proto"
    fd.package = "vep_oracle"
    fd.syntax = "proto3"
    sp = fd.message_type.add()
    sp.name = "ShapeProto"
    dim = sp.nested_type.add()
    dim.name = "Dim"
    for n, num, t in (("size", 1, F.TYPE_INT64), ("name", 2, F.TYPE_STRING)):
        f = dim.field.add()
        f.name, f.number, f.type, f.label = n, num, t, F.LABEL_OPTIONAL
    f = sp.field.add()
    f.name, f.number, f.type, f.label = "dim", 2, F.TYPE_MESSAGE, F.LABEL_REPEATED
    f.type_name = ".vep_oracle.ShapeProto.Dim"
    vf = fd.message_type.add()
    vf.name = "VideoFrame"
    for n, num, t in (("width", 1, F.TYPE_INT64), ("height", 2, F.TYPE_INT64),
                      ("data", 3, F.TYPE_BYTES), ("timestamp", 4, F.TYPE_INT64),
                      ("is_keyframe", 5, F.TYPE_BOOL), ("pts", 6, F.TYPE_INT64),
                      ("dts", 7, F.TYPE_INT64), ("frame_type", 8, F.TYPE_STRING),
                      ("is_corrupt", 9, F.TYPE_BOOL), ("time_base", 10, F.TYPE_DOUBLE),
                      ("shape", 11, F.TYPE_MESSAGE), ("device_id", 12, F.TYPE_STRING),
                      ("packet", 13, F.TYPE_INT64), ("keyframe", 14, F.TYPE_INT64)):
        f = vf.field.add()
        f.name, f.number, f.type, f.label = n, num, t, F.LABEL_OPTIONAL
        if n == "shape":
            f.type_name = ".vep_oracle.ShapeProto"
    pool = descriptor_pool.DescriptorPool()
    pool.Add(fd)
    get = message_factory.GetMessageClass
    return (get(pool.FindMessageTypeByName("vep_oracle.VideoFrame")),
            get(pool.FindMessageTypeByName("vep_oracle.ShapeProto.Dim")))


def serialize_with_protobuf_runtime(**kw) -> bytes:
    """Fill the dynamic message exactly as read_image.py:99-117 does and serialise it."""
    VideoFrame, Dim = dynamic_message_class()
    vf = VideoFrame()
    vf.data = kw["data"]
    vf.width = kw["width"]
    vf.height = kw["height"]
    vf.timestamp = kw["timestamp"]
    vf.frame_type = kw["frame_type"]
    vf.pts = kw["pts"]
    vf.dts = kw["dts"]
    vf.packet = kw["packet"]
    vf.keyframe = kw["keyframe"]
    vf.time_base = kw["time_base"]
    vf.is_keyframe = kw["is_keyframe"]
    vf.is_corrupt = kw["is_corrupt"]
    for i, d in enumerate(kw["shape"]):
        nd = Dim()
        nd.size = d
        nd.name = str(i)
        vf.shape.dim.append(nd)
    return vf.SerializeToString()
