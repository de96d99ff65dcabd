"""Host-side mirror of the reference's per-frame path, on top of the C ABI (include/vepb200.h).

What the reference does per frame, per camera process (python/read_image.py):
    img = frame.to_ndarray(format='bgr24')      # :94   PyAV -> libswscale
    shape = img.shape                            # :95
    img_bytes = np.ndarray.tobytes(img)          # :97
    vf = VideoFrame(...); vf.SerializeToString() # :99-119
    redis_conn.xadd(device_id, {'data': ...})    # :121
`FramePath` does the first four for MANY cameras per tick on one B200: submit() stages decoded
planes in pinned memory, tick() runs H2D -> one batched kernel -> D2H, result()/result_packed()
hand back views of the pinned output ring (no copies), result_tensor() a zero-copy CUDA tensor.
All arithmetic runs in libvepb200.so's sm_100a kernels; nothing here converts pixels on the CPU.
"""
from __future__ import annotations

import ctypes as C
import threading
from typing import Dict, Optional, Sequence

import numpy as np

from . import _abi
from ._abi import FrameDesc, FrameMeta, VEP_FMT, VepError, check


def make_meta(timestamp=0, pts=0, dts=0, packet=0, keyframe=0, time_base=0.0, is_keyframe=False,
              is_corrupt=False, frame_type="") -> FrameMeta:
    """The scalar VideoFrame fields the reference fills at read_image.py:101-111."""
    m = FrameMeta()
    m.timestamp, m.pts, m.dts, m.packet, m.keyframe = int(timestamp), int(pts), int(dts), int(packet), int(keyframe)
    m.time_base = float(time_base)
    m.is_keyframe, m.is_corrupt = int(bool(is_keyframe)), int(bool(is_corrupt))
    ft = frame_type.encode() if isinstance(frame_type, str) else bytes(frame_type)
    if len(ft) > 7:
        raise ValueError("frame_type longer than 7 bytes")
    m.frame_type = ft
    return m


def videoframe_framing(meta: FrameMeta, width: int, height: int):
    """(prefix, suffix) bytes that surround `data` in the serialized VideoFrame
    (read_image.py:99-119 / proto/video_streaming.proto:78-93)."""
    L = _abi.lib()
    pre = (C.c_uint8 * 32)()
    suf = (C.c_uint8 * 192)()
    pl, sl = C.c_int32(0), C.c_int32(0)
    check(L.vep_videoframe_frame(C.byref(meta), width, height, pre, C.byref(pl), suf, C.byref(sl)),
          "vep_videoframe_frame")
    return bytes(pre[:pl.value]), bytes(suf[:sl.value])


def _plane_ptrs(planes: Sequence[np.ndarray]):
    ptrs = (C.c_void_p * 3)()
    strides = (C.c_int32 * 3)()
    for i, p in enumerate(planes):
        if p.dtype not in (np.uint8, np.uint16) or p.ndim != 2 or p.strides[1] != p.itemsize:
            raise ValueError("planes must be 2-D uint8 (uint16 for 10-bit formats) arrays with contiguous rows")
        ptrs[i] = p.ctypes.data
        strides[i] = p.strides[0]          # bytes, as in AVFrame.linesize
    return ptrs, strides


def desc_from_device_ptrs(fmt: str, width: int, height: int, planes, strides, dst, dst_stride=None,
                          convert=True) -> FrameDesc:
    d = FrameDesc()
    for i, (p, s) in enumerate(zip(planes, strides)):
        d.plane[i] = int(p)
        d.stride[i] = int(s)
    d.dst = int(dst)
    d.dst_stride = int(dst_stride if dst_stride is not None else 3 * width)
    d.width, d.height, d.fmt, d.convert = width, height, VEP_FMT[fmt], int(bool(convert))
    return d


def desc_from_tensors(fmt: str, planes, out, convert=True) -> FrameDesc:
    """Descriptor over torch CUDA uint8 tensors (zero-copy: only data_ptr()/stride() are read).
    planes: (Y,U,V) or (Y,UV) 2-D tensors; out: (H,W,3) tensor."""
    h, w = int(planes[0].shape[0]), int(planes[0].shape[1])
    for t in list(planes) + [out]:
        if not t.is_cuda or t.stride(-1) != 1:
            raise ValueError("tensors must be CUDA tensors with unit innermost stride")
    if str(out.dtype) != "torch.uint8" or any(t.element_size() != (2 if fmt == "yuv420p10le" else 1) for t in planes):
        raise ValueError("planes must be uint8 (16-bit elements for yuv420p10le), out uint8")
    if tuple(out.shape) != (h, w, 3) or out.stride(1) != 3:
        raise ValueError("out must be (H, W, 3) with packed pixels")
    return desc_from_device_ptrs(fmt, w, h, [p.data_ptr() for p in planes], [p.stride(0) * p.element_size() for p in planes],
                                 out.data_ptr(), out.stride(0), convert)


class Plan:
    """A validated descriptor table resident on the GPU: one kernel launch per `launch()`."""

    def __init__(self, path: "FramePath", descs: Sequence[FrameDesc]):
        self._L = _abi.lib()
        self._path = path   # keeps the context alive
        arr = (FrameDesc * len(descs))(*descs)
        self._h = C.c_void_p()
        check(self._L.vep_plan_create(path._ctx, arr, len(descs), C.byref(self._h)), "vep_plan_create")
        n, ctas, ab = C.c_int(), C.c_int(), C.c_int64()
        check(self._L.vep_plan_info(self._h, C.byref(n), C.byref(ctas), C.byref(ab)), "vep_plan_info")
        self.n_converted, self.n_ctas, self.alg_bytes = n.value, ctas.value, ab.value

    def launch(self, cuda_stream: int = 0, overlapped: bool = False, order: Optional[int] = None):
        """overlapped=True: the caller vouches that this launch is independent of the previous kernel
        launch on the stream (disjoint buffers), so the GPU may overlap the two.  `order` selects any of
        _abi.VEP_ORDER_STRICT / VEP_ORDER_WRITES / VEP_ORDER_NONE explicitly (see vepb200.h)."""
        if order is None:
            order = _abi.VEP_ORDER_NONE if overlapped else _abi.VEP_ORDER_STRICT
        check(self._L.vep_plan_launch_ordered(self._h, C.c_void_p(cuda_stream), int(order)), "vep_plan_launch_ordered")

    def close(self):
        if self._h:
            self._L.vep_plan_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def launch_sequence(plans: Sequence["Plan"], cuda_stream: int = 0, order: int = _abi.VEP_ORDER_STRICT,
                    event_before: int = 0, event_after: int = 0):
    """Launch `plans` back to back on one stream from native code (vep_plan_launch_sequence): one ctypes
    call instead of len(plans), for producers that convert tick after tick.  event_before / event_after are raw
    cudaEvent_t handles (e.g. torch.cuda.Event(...).cuda_event of an event that has been recorded once) recorded
    directly around the launches."""
    arr = (C.c_void_p * len(plans))(*[p._h for p in plans])
    check(_abi.lib().vep_plan_launch_sequence(arr, len(plans), C.c_void_p(cuda_stream), int(order),
                                              C.c_void_p(event_before or None), C.c_void_p(event_after or None)), "vep_plan_launch_sequence")


class _CudaView:
    """Minimal __cuda_array_interface__ carrier so torch can wrap a device pointer without a copy."""

    def __init__(self, ptr: int, shape, owner, strides=None):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": "|u1", "data": (ptr, False),
                                         "version": 2, "strides": strides}
        self._owner = owner


class FramePath:
    """One per GPU.  Camera streams are named by the reference's `device_id` strings."""

    def __init__(self, device: int = 0):
        self._L = _abi.lib()
        self._ctx = C.c_void_p()
        check(self._L.vep_create(int(device), C.byref(self._ctx)), "vep_create")
        self.device = int(device)
        self._ids: Dict[str, int] = {}
        self._geom: Dict[str, tuple] = {}
        self._next_id = 0
        self._open_lock = threading.Lock()      # stream ids are handed out here: cameras may be opened from several threads

    # ---- lifecycle -----------------------------------------------------------------------------
    def close(self):
        if self._ctx:
            self._L.vep_destroy(self._ctx)
            self._ctx = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- camera streams ------------------------------------------------------------------------
    def open_stream(self, device_id: str, fmt: str, width: int, height: int, memory_buffer: int = 1):
        """Declare a camera.  memory_buffer is the reference's --memory_buffer (XADD MAXLEN)."""
        if device_id in self._ids:
            raise VepError(_abi.VEP_ERR_STATE, "open_stream", f"{device_id!r} already open")
        if fmt not in VEP_FMT:
            raise VepError(_abi.VEP_ERR_INVALID_ARG, "open_stream", f"unknown format {fmt!r}")
        with self._open_lock:
            if device_id in self._ids:
                raise VepError(_abi.VEP_ERR_STATE, "open_stream", f"{device_id!r} already open")
            sid = self._next_id
            check(self._L.vep_stream_open(self._ctx, sid, VEP_FMT[fmt], width, height, memory_buffer), "vep_stream_open")
            self._next_id += 1
            self._ids[device_id] = sid
            self._geom[device_id] = (fmt, width, height)

    def open_streams(self, cameras, memory_buffer: int = 1):
        """Declare several cameras at once: `cameras` = [(device_id, fmt, width, height), ...].  Their slots share one
        allocation laid out [slot][camera], so cameras that are neighbours in the list and tick in lock step are copied to
        and from the GPU with one copy per direction for the whole run (vep_stream_open_many)."""
        with self._open_lock:
            self._open_streams_locked(cameras, memory_buffer)

    def _open_streams_locked(self, cameras, memory_buffer):
        specs = (_abi.StreamSpec * len(cameras))()
        for i, (device_id, fmt, w, h) in enumerate(cameras):
            if device_id in self._ids or device_id in [c[0] for c in cameras[:i]]:
                raise VepError(_abi.VEP_ERR_STATE, "open_streams", f"{device_id!r} already open")
            if fmt not in VEP_FMT:
                raise VepError(_abi.VEP_ERR_INVALID_ARG, "open_streams", f"unknown format {fmt!r}")
            specs[i].stream_id, specs[i].fmt, specs[i].width, specs[i].height = self._next_id + i, VEP_FMT[fmt], w, h
        check(self._L.vep_stream_open_many(self._ctx, specs, len(cameras), memory_buffer), "vep_stream_open_many")
        for i, (device_id, fmt, w, h) in enumerate(cameras):
            self._ids[device_id] = self._next_id + i
            self._geom[device_id] = (fmt, w, h)
        self._next_id += len(cameras)

    def close_stream(self, device_id: str):
        with self._open_lock:
            check(self._L.vep_stream_close(self._ctx, self._sid(device_id)), "vep_stream_close")
            del self._ids[device_id], self._geom[device_id]

    def _sid(self, device_id: str) -> int:
        try:
            return self._ids[device_id]
        except KeyError:
            raise VepError(_abi.VEP_ERR_INVALID_ARG, "stream", f"{device_id!r} not open") from None

    def streams(self):
        return list(self._ids)

    # ---- per-frame -----------------------------------------------------------------------------
    def submit(self, device_id: str, planes: Sequence[np.ndarray], meta: Optional[FrameMeta] = None,
               convert: bool = True, pinned: bool = False):
        """Stage one decoded frame (host planes, any row stride).  ctypes drops the GIL for the
        copy, so one producer thread per camera scales like the reference's one process per camera.
        pinned=True (VEP_SUBMIT_PINNED): the planes lie in page-locked memory (see register_host) and are NOT
        copied — keep them alive and unchanged until the tick that converts them has been waited for."""
        sid = self._sid(device_id)
        if not convert:
            check(self._L.vep_submit(self._ctx, sid, None, None, None, 0), "vep_submit")
            return
        ptrs, strides = _plane_ptrs(planes)
        check(self._L.vep_submit(self._ctx, sid, C.byref(ptrs), C.byref(strides),
                                 C.byref(meta) if meta is not None else None,
                                 _abi.VEP_SUBMIT_PINNED if pinned else _abi.VEP_SUBMIT_COPY), "vep_submit")

    @staticmethod
    def register_host(buf: np.ndarray):
        """Page-lock the memory of `buf` (a contiguous array a decoder's frames live in) so that frames inside it can
        be submitted with pinned=True.  Once per buffer, not per frame; unregister_host before the array is freed."""
        check(_abi.lib().vep_host_register(C.c_void_p(buf.ctypes.data), C.c_size_t(buf.nbytes)), "vep_host_register")

    @staticmethod
    def unregister_host(buf: np.ndarray):
        check(_abi.lib().vep_host_unregister(C.c_void_p(buf.ctypes.data)), "vep_host_unregister")

    def submit_many(self, frames, pinned: bool = False) -> int:
        """Stage the frames of one tick with ONE call: `frames` is a sequence of (device_id, planes, meta) — at most
        one per camera — or (device_id, None, None) for a gate-skipped frame.  The copies are shared between this
        thread and the helper threads of set_staging_threads(); with pinned=True nothing is copied (see submit).
        Returns the number of items staged; raises on the first one that could not be."""
        arr = (_abi.SubmitItem * len(frames))()
        for it, (device_id, planes, meta) in zip(arr, frames):
            it.stream_id = self._sid(device_id)
            if planes is None:
                it.convert = 0
                continue
            it.convert = _abi.VEP_SUBMIT_PINNED if pinned else _abi.VEP_SUBMIT_COPY
            for i, p in enumerate(planes):
                if p.dtype not in (np.uint8, np.uint16) or p.ndim != 2 or p.strides[1] != p.itemsize:
                    raise ValueError("planes must be 2-D uint8 (uint16 for 10-bit formats) arrays with contiguous rows")
                it.plane[i] = p.ctypes.data
                it.stride[i] = p.strides[0]
            if meta is not None:
                it.meta = meta
        done = C.c_int()
        check(self._L.vep_submit_batch(self._ctx, arr, len(frames), C.byref(done)), "vep_submit_batch")
        return done.value

    def acquire_input(self, device_id: str):
        """Writable numpy views of the next pinned staging slot (zero-copy submit); follow with
        commit_input()."""
        fmt, w, h = self._geom[device_id]
        ptrs, strides = (C.c_void_p * 3)(), (C.c_int32 * 3)()
        check(self._L.vep_acquire_input(self._ctx, self._sid(device_id), C.byref(ptrs), C.byref(strides)), "vep_acquire_input")
        ch = h if fmt in ("yuv422p", "yuvj422p", "yuv444p", "yuvj444p") else h // 2
        cw = w if fmt in ("yuv444p", "yuvj444p") else w // 2
        shapes = [(h, w)] if fmt == "gray" else [(h, w), (h // 2, w)] if fmt in ("nv12", "nv21") else [(h, w), (ch, cw), (ch, cw)]
        out = []
        for i, shp in enumerate(shapes):
            buf = (C.c_uint8 * (shp[0] * strides[i])).from_address(ptrs[i])
            rows = np.frombuffer(buf, np.uint8).reshape(shp[0], strides[i])
            out.append(rows.view(np.uint16)[:, :shp[1]] if fmt == "yuv420p10le" else rows[:, :shp[1]])
        return out

    def staged_bytes(self, device_id: str) -> int:
        """Bytes one staged frame of this camera occupies in its pinned slot = what the tick's H2D copy moves
        (planes with rows padded to 16-byte pitches where the width needs it, planes 256-byte aligned)."""
        fmt, w, h = self._geom[device_id]
        bps = 2 if fmt == "yuv420p10le" else 1
        pitch = lambda rb: rb if rb % 16 == 0 else (rb + 63) // 64 * 64
        a256 = lambda v: (v + 255) // 256 * 256
        ysz = pitch(w * bps) * h
        if fmt == "gray":
            return ysz
        semi, c444, c422 = fmt in ("nv12", "nv21"), fmt in ("yuv444p", "yuvj444p"), fmt in ("yuv422p", "yuvj422p")
        csz = pitch((w if (semi or c444) else w // 2) * bps) * (h if (c422 or c444) else h // 2)
        return a256(ysz) + csz if semi else a256(ysz) + a256(csz) + csz

    def acquire_raw(self, device_id: str):
        """(pointers, strides) of the next pinned staging slot as ctypes arrays — for producers that hand the
        addresses to native code (a decoder's get_buffer callback) instead of numpy."""
        ptrs, strides = (C.c_void_p * 3)(), (C.c_int32 * 3)()
        check(self._L.vep_acquire_input(self._ctx, self._sid(device_id), C.byref(ptrs), C.byref(strides)), "vep_acquire_input")
        return ptrs, strides

    def commit_input(self, device_id: str, meta: Optional[FrameMeta] = None):
        check(self._L.vep_commit_input(self._ctx, self._sid(device_id),
                                       C.byref(meta) if meta is not None else None), "vep_commit_input")

    def set_staging_threads(self, n: int):
        """Helper threads that share the staging copy of large (4K) frames with the submitting thread."""
        check(self._L.vep_set_staging_threads(self._ctx, int(n)), "vep_set_staging_threads")

    def set_copy_run_limit(self, max_frames_per_copy: int):
        """Most frames a merged copy of neighbouring cameras may carry (0: no limit, 1: one copy per frame)."""
        check(self._L.vep_set_copy_run_limit(self._ctx, int(max_frames_per_copy)), "vep_set_copy_run_limit")

    def tick_begin(self) -> int:
        n = C.c_int()
        check(self._L.vep_tick_begin(self._ctx, C.byref(n)), "vep_tick_begin")
        return n.value

    def tick_wait(self):
        """Wait for the oldest tick in flight."""
        check(self._L.vep_tick_wait(self._ctx), "vep_tick_wait")

    def tick_begin_ticket(self):
        """(frames in this tick, ticket): the ticket names this tick for tick_wait_ticket (an empty tick
        returns the ticket of the newest real tick)."""
        n, t = C.c_int(), C.c_uint64()
        check(self._L.vep_tick_begin_ticket(self._ctx, C.byref(n), C.byref(t)), "vep_tick_begin_ticket")
        return n.value, t.value

    def tick_wait_ticket(self, ticket: int):
        """Return once the tick `ticket` and every older one have landed, whoever else ticks meanwhile."""
        check(self._L.vep_tick_wait_ticket(self._ctx, C.c_uint64(ticket)), "vep_tick_wait_ticket")

    def tick(self) -> int:
        """Convert every staged frame of every camera with one kernel launch; returns the count."""
        n = C.c_int()
        check(self._L.vep_tick(self._ctx, C.byref(n)), "vep_tick")
        return n.value

    def result(self, device_id: str, age: int = 0) -> np.ndarray:
        """(H, W, 3) uint8 BGR view of the newest converted frame — the `img` of read_image.py:94.
        Zero-copy view of pinned memory: valid until memory_buffer newer frames of this camera
        have been converted; copy it (`.copy()`/`.tobytes()`) to keep it longer."""
        p, n, shp = C.c_void_p(), C.c_size_t(), (C.c_int32 * 3)()
        check(self._L.vep_result_at(self._ctx, self._sid(device_id), age, C.byref(p), C.byref(n), shp), "vep_result_at")
        buf = (C.c_uint8 * n.value).from_address(p.value)
        return np.frombuffer(buf, np.uint8).reshape(shp[0], shp[1], shp[2])

    def result_packed(self, device_id: str, age: int = 0) -> memoryview:
        """The serialized VideoFrame (`vfData`, read_image.py:119) as a zero-copy memoryview,
        ready for `redis_conn.xadd(device_id, {'data': ...})` (:121)."""
        p, n = C.c_void_p(), C.c_size_t()
        check(self._L.vep_result_packed_at(self._ctx, self._sid(device_id), age, C.byref(p), C.byref(n)), "vep_result_packed_at")
        return memoryview((C.c_uint8 * n.value).from_address(p.value)).cast("B")

    def result_tensor(self, device_id: str):
        """Zero-copy torch CUDA uint8 tensor (H, W, 3) over the device-side result (rows are padded on the
        device when width*3 is not a multiple of 16: the tensor is then a strided view)."""
        import torch
        p, shp, pitch = C.c_void_p(), (C.c_int32 * 3)(), C.c_int32()
        check(self._L.vep_result_device_pitched(self._ctx, self._sid(device_id), C.byref(p), shp, C.byref(pitch)), "vep_result_device_pitched")
        strides = None if pitch.value == 3 * shp[1] else (pitch.value, 3, 1)
        return torch.as_tensor(_CudaView(p.value, (shp[0], shp[1], shp[2]), self, strides), device=f"cuda:{self.device}")

    # ---- device-resident batches (torch tensors in, torch tensors out) --------------------------
    def convert_batch(self, descs: Sequence[FrameDesc], cuda_stream: int = 0, order: int = _abi.VEP_ORDER_STRICT):
        arr = (FrameDesc * len(descs))(*descs)
        check(self._L.vep_convert_batch_ordered(self._ctx, arr, len(descs), C.c_void_p(cuda_stream), int(order)), "vep_convert_batch_ordered")

    def plan(self, descs: Sequence[FrameDesc]) -> Plan:
        return Plan(self, descs)

    def stats(self) -> dict:
        s = _abi.Stats()
        check(self._L.vep_get_stats(self._ctx, C.byref(s)), "vep_get_stats")
        return {k: getattr(s, k) for k, _ in s._fields_}


def to_ndarray_bgr24(path: FramePath, fmt: str, planes: Sequence[np.ndarray]) -> np.ndarray:
    """Single-frame convenience with the reference's call shape: what
    `frame.to_ndarray(format='bgr24')` (read_image.py:94) returns, computed on the GPU.
    Opens a temporary stream; for throughput use open_stream/submit/tick."""
    h, w = planes[0].shape
    name = f"__oneshot_{id(planes)}"
    path.open_stream(name, fmt, w, h, 1)
    try:
        path.submit(name, planes)
        path.tick()
        return path.result(name).copy()
    finally:
        path.close_stream(name)


def convert_batch_noalloc(descs: Sequence[FrameDesc], cuda_stream: int = 0):
    """Batched conversion without any device allocation and without a FramePath (vep_convert_batch_noalloc).

    For producers that may not allocate on the GPU while running: a fixed memory slice, a host
    application that owns the allocator, a real-time loop.  The descriptor tables go into a fixed
    array inside the library (limits: 256 frames / 65536 thread blocks per call, larger batches
    raise VepError).  Runs on the current CUDA device and the given stream; calls that could overlap
    on the GPU share that array and must be serialised by the caller.  The first call on a device
    also does a one-time address lookup, so make it before starting a stream capture."""
    arr = (FrameDesc * len(descs))(*descs)
    check(_abi.lib().vep_convert_batch_noalloc(arr, len(descs), C.c_void_p(cuda_stream)), "vep_convert_batch_noalloc")
