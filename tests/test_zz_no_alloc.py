"""GPU: the allocation-free entry point (vep_convert_batch_noalloc), called through the C ABI.

It exists for callers that cannot allocate device memory while they run — a fixed GPU memory slice,
a host application that owns the allocator, a real-time loop — so its descriptor tables sit in a
fixed array inside the library instead of cudaMalloc'ed buffers.  The check: on the same inputs its
output must equal, byte for byte (the bar every test here uses), what the ordinary entry point
vep_convert_batch writes — on two calls in a row, because the second call re-uses the array that
still holds the first call's tables (a larger batch first, then a smaller, different one) — and, so
that the comparison is not only kernel against kernel, what the CPU oracle computes for those planes.

Collected last on purpose (file name), so a run that stops at the first failure has already run
everything else.  Needs no state from other tests and leaves none behind.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

from oracle import formulas  # noqa: E402
from video_edge_ai_proxy_b200 import FramePath, VepError, _abi, desc_from_tensors  # noqa: E402
from video_edge_ai_proxy_b200.synth import make_planes  # noqa: E402

FIRST = [("yuv420p", 1920, 1080, 1), ("nv12", 1280, 720, 2), ("yuvj420p", 640, 480, 3), ("yuv420p", 50, 6, 4),
         ("nv12", 642, 10, 5), ("yuv420p", 1920, 1080, 6), ("yuv420p", 528, 8, 7)]
SECOND = [("nv12", 1920, 1080, 11), ("yuv420p", 640, 480, 12), ("yuvj420p", 34, 4, 13), ("yuvj422p", 640, 480, 14), ("nv21", 320, 240, 15)]


def _batch(specs, fill):
    descs, keep = [], []
    for fmt, w, h, seed in specs:
        planes = [torch.from_numpy(p).cuda() for p in make_planes(fmt, w, h, seed)]
        out = torch.full((h, w, 3), fill, dtype=torch.uint8, device="cuda")
        descs.append(desc_from_tensors(fmt, planes, out))
        keep.append((planes, out))
    return descs, keep


def _noalloc(descs, stream):
    arr = (_abi.FrameDesc * len(descs))(*descs)
    _abi.check(_abi.lib().vep_convert_batch_noalloc(arr, len(descs), C.c_void_p(stream)), "vep_convert_batch_noalloc")


def test_noalloc_entry_point_matches_convert_batch_twice_in_a_row():
    if not torch.cuda.is_available():
        pytest.fail("marked gpu but no CUDA device")
    torch.cuda.set_device(0)
    stream = torch.cuda.current_stream().cuda_stream
    with FramePath(0) as path:
        for specs in (FIRST, SECOND):                      # two calls in a row share the one scratch array
            d_ref, k_ref = _batch(specs, 0x00)
            d_new, k_new = _batch(specs, 0xFF)
            path.convert_batch(d_ref, stream)              # the existing entry point
            _noalloc(d_new, stream)                        # the new one, same inputs, other outputs
            torch.cuda.synchronize()
            for (fmt, w, h, _), (_, a), (_, b) in zip(specs, k_ref, k_new):
                ref, new = a.cpu().numpy(), b.cpu().numpy()
                assert ref.any(), "reference output was not written"
                assert np.array_equal(new, ref), (fmt, w, h)   # bit-exact, as everywhere in this suite
            for (fmt, w, h, seed), (_, b) in zip(specs, k_new):    # and against the oracle itself
                assert np.array_equal(b.cpu().numpy(), formulas.to_bgr24(fmt, make_planes(fmt, w, h, seed))), (fmt, w, h, "oracle")


def test_noalloc_limits_and_skips():
    if not torch.cuda.is_available():
        pytest.fail("marked gpu but no CUDA device")
    torch.cuda.set_device(0)
    stream = torch.cuda.current_stream().cuda_stream
    planes = [torch.from_numpy(p).cuda() for p in make_planes("yuv420p", 16, 2, 0)]
    out = torch.full((2, 16, 3), 7, dtype=torch.uint8, device="cuda")
    # one frame too many -> refused with a clear message, nothing launched
    descs = [desc_from_tensors("yuv420p", planes, out) for _ in range(257)]
    with pytest.raises(VepError) as e:
        _noalloc(descs, stream)
    assert e.value.status == _abi.VEP_ERR_INVALID_ARG and "VEP_NOALLOC_MAX_FRAMES" in str(e.value)
    # too many thread blocks: 61 frames of 4096x4096 would need > 65536
    big = [torch.empty((4096, 4096), dtype=torch.uint8, device="cuda"), torch.empty((2048, 2048), dtype=torch.uint8, device="cuda"),
           torch.empty((2048, 2048), dtype=torch.uint8, device="cuda")]
    bout = torch.empty((4096, 4096, 3), dtype=torch.uint8, device="cuda")
    with pytest.raises(VepError) as e:
        _noalloc([desc_from_tensors("yuv420p", big, bout) for _ in range(61)], stream)
    assert e.value.status == _abi.VEP_ERR_INVALID_ARG and "VEP_NOALLOC_MAX_CTAS" in str(e.value)
    torch.cuda.synchronize()
    assert (out == 7).all()
    # gate-skipped descriptors do not count and are not touched; an empty call is fine
    d = desc_from_tensors("yuv420p", planes, out, convert=False)
    _noalloc([d] * 300, stream)
    _noalloc([], stream)
    torch.cuda.synchronize()
    assert (out == 7).all()
