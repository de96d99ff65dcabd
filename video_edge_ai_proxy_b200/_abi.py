"""ctypes binding of libvepb200.so — one declaration per entry point of include/vepb200.h.

This is the stub INTEGRATION.md shows for the reference side.  The library is built in-tree by
`video_edge_ai_proxy_b200/csrc/Makefile`; if it is missing the import fails loudly (there is no
CPU or PyTorch fallback for the conversion).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libvepb200.so")

VEP_FMT = {"yuv420p": 0, "yuvj420p": 1, "nv12": 2, "nv21": 3, "yuv422p": 4, "yuvj422p": 5, "yuv420p10le": 6, "yuv444p": 7, "yuvj444p": 8, "gray": 9}
VEP_OK, VEP_ERR_INVALID_ARG, VEP_ERR_GEOMETRY, VEP_ERR_CUDA = 0, -1, -2, -3
VEP_ORDER_STRICT, VEP_ORDER_WRITES, VEP_ORDER_NONE = 0, 1, 2
VEP_SUBMIT_SKIP, VEP_SUBMIT_COPY, VEP_SUBMIT_PINNED = 0, 1, 2
VEP_ERR_NOMEM, VEP_ERR_STATE, VEP_ERR_NO_FRAME, VEP_ERR_BUSY = -4, -5, -6, -7


class FrameDesc(C.Structure):
    _fields_ = [("plane", C.c_void_p * 3), ("stride", C.c_int32 * 3), ("dst", C.c_void_p),
                ("dst_stride", C.c_int32), ("width", C.c_int32), ("height", C.c_int32),
                ("fmt", C.c_int32), ("convert", C.c_int32)]


class FrameMeta(C.Structure):
    _fields_ = [("timestamp", C.c_int64), ("pts", C.c_int64), ("dts", C.c_int64),
                ("packet", C.c_int64), ("keyframe", C.c_int64), ("time_base", C.c_double),
                ("is_keyframe", C.c_int32), ("is_corrupt", C.c_int32),
                ("frame_type", C.c_char * 8)]


class SubmitItem(C.Structure):
    _fields_ = [("stream_id", C.c_int32), ("convert", C.c_int32), ("plane", C.c_void_p * 3), ("stride", C.c_int32 * 3),
                ("reserved", C.c_int32), ("meta", FrameMeta)]


class Stats(C.Structure):
    _fields_ = [("frames_offered", C.c_int64), ("frames_converted", C.c_int64),
                ("ticks", C.c_int64), ("kernel_launches", C.c_int64),
                ("h2d_bytes", C.c_int64), ("d2h_bytes", C.c_int64),
                ("last_kernel_ms", C.c_float), ("reserved", C.c_int32),
                ("copies_in", C.c_int64), ("copies_out", C.c_int64)]


class StreamSpec(C.Structure):
    _fields_ = [("stream_id", C.c_int32), ("fmt", C.c_int32), ("width", C.c_int32), ("height", C.c_int32)]


_p = C.c_void_p
_pp = C.POINTER(C.c_void_p)
_i = C.c_int
_ip = C.POINTER(C.c_int)
_i32p = C.POINTER(C.c_int32)

# name -> (restype, argtypes); must list every symbol include/vepb200.h declares
SIGNATURES = {
    "vep_abi_version": (_i, []),
    "vep_last_error": (C.c_char_p, []),
    "vep_status_name": (C.c_char_p, [_i]),
    "vep_create": (_i, [_i, _pp]),
    "vep_destroy": (None, [_p]),
    "vep_convert_batch": (_i, [_p, C.POINTER(FrameDesc), _i, _p]),
    "vep_convert_batch_ordered": (_i, [_p, C.POINTER(FrameDesc), _i, _p, _i]),
    "vep_convert_batch_noalloc": (_i, [C.POINTER(FrameDesc), _i, _p]),
    "vep_plan_create": (_i, [_p, C.POINTER(FrameDesc), _i, _pp]),
    "vep_plan_launch": (_i, [_p, _p]),
    "vep_plan_launch_overlapped": (_i, [_p, _p]),
    "vep_plan_launch_ordered": (_i, [_p, _p, _i]),
    "vep_plan_launch_sequence": (_i, [C.POINTER(C.c_void_p), _i, _p, _i, _p, _p]),
    "vep_plan_info": (_i, [_p, _ip, _ip, C.POINTER(C.c_int64)]),
    "vep_plan_destroy": (None, [_p]),
    "vep_stream_open": (_i, [_p, _i, _i, _i, _i, _i]),
    "vep_stream_open_many": (_i, [_p, C.POINTER(StreamSpec), _i, _i]),
    "vep_stream_close": (_i, [_p, _i]),
    "vep_set_staging_threads": (_i, [_p, _i]),
    "vep_set_copy_run_limit": (_i, [_p, _i]),
    "vep_host_register": (_i, [_p, C.c_size_t]),
    "vep_host_unregister": (_i, [_p]),
    "vep_submit": (_i, [_p, _i, C.POINTER(C.c_void_p * 3), C.POINTER(C.c_int32 * 3),
                        C.POINTER(FrameMeta), _i]),
    "vep_submit_batch": (_i, [_p, C.POINTER(SubmitItem), _i, _ip]),
    "vep_acquire_input": (_i, [_p, _i, C.POINTER(C.c_void_p * 3), C.POINTER(C.c_int32 * 3)]),
    "vep_commit_input": (_i, [_p, _i, C.POINTER(FrameMeta)]),
    "vep_tick_begin": (_i, [_p, _ip]),
    "vep_tick_wait": (_i, [_p]),
    "vep_tick": (_i, [_p, _ip]),
    "vep_tick_begin_ticket": (_i, [_p, _ip, C.POINTER(C.c_uint64)]),
    "vep_tick_wait_ticket": (_i, [_p, C.c_uint64]),
    "vep_result": (_i, [_p, _i, _pp, C.POINTER(C.c_size_t), _i32p]),
    "vep_result_packed": (_i, [_p, _i, _pp, C.POINTER(C.c_size_t)]),
    "vep_result_packed_at": (_i, [_p, _i, _i, _pp, C.POINTER(C.c_size_t)]),
    "vep_result_at": (_i, [_p, _i, _i, _pp, C.POINTER(C.c_size_t), _i32p]),
    "vep_result_device": (_i, [_p, _i, _pp, _i32p]),
    "vep_result_device_pitched": (_i, [_p, _i, _pp, _i32p, _i32p]),
    "vep_videoframe_frame": (_i, [C.POINTER(FrameMeta), C.c_int32, C.c_int32, _p, _i32p, _p, _i32p]),
    "vep_get_stats": (_i, [_p, C.POINTER(Stats)]),
}

_lib = None


class VepError(RuntimeError):
    def __init__(self, status: int, where: str, detail: str):
        self.status = status
        super().__init__(f"{where}: {detail} (status {status})")


def build_library(verbose: bool = False) -> str:
    """Compile libvepb200.so in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    subprocess.check_call(["make", "-C", os.path.join(_HERE, "csrc")] + ([] if verbose else ["-s"]))
    return LIB_PATH


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                "(the conversion has no CPU or PyTorch fallback)")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(status: int, where: str) -> int:
    if status < 0:
        L = lib()
        raise VepError(status, where, f"{L.vep_status_name(status).decode()}: {L.vep_last_error().decode()}")
    return status
