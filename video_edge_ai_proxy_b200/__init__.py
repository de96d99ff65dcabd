"""b200 frame path: batched YUV 4:2:0 -> BGR24 + VideoFrame packing for video-edge-ai-proxy.

Public surface (host side mirrors reference python/read_image.py; the arithmetic is in
libvepb200.so, hand-written sm_100a CUDA behind the C ABI of include/vepb200.h):

    FramePath, Plan, make_meta, videoframe_framing, to_ndarray_bgr24   (frame_path.py)
    DecodeGate, poll_control, camera_active                             (gate.py)
    FrameProducer                                                       (producer.py)
    ReadImage (drop-in for the reference's class of that name)          (read_image.py)
    assign_cameras                                                      (shard.py)
"""
from ._abi import VepError, build_library  # noqa: F401
from .frame_path import (FramePath, Plan, convert_batch_noalloc, desc_from_device_ptrs, desc_from_tensors,  # noqa: F401
                         launch_sequence, make_meta, to_ndarray_bgr24, videoframe_framing)
from .gate import CameraControl, DecodeGate, camera_active, poll_control  # noqa: F401
from .producer import FrameProducer  # noqa: F401
from .shard import assign_cameras  # noqa: F401

__all__ = ["FramePath", "Plan", "make_meta", "videoframe_framing", "to_ndarray_bgr24", "DecodeGate",
           "CameraControl", "camera_active", "poll_control", "FrameProducer", "assign_cameras",
           "VepError", "build_library", "desc_from_device_ptrs", "desc_from_tensors", "convert_batch_noalloc",
           "launch_sequence"]
