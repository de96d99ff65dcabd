"""GPU: parity of the sm_100a kernels, called through the C ABI, against the oracle and the goldens."""
import hashlib
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

from oracle import formulas as F  # noqa: E402
from oracle import swscale_live as L  # noqa: E402
from video_edge_ai_proxy_b200 import FramePath, VepError, desc_from_device_ptrs, desc_from_tensors, launch_sequence  # noqa: E402
from video_edge_ai_proxy_b200 import _abi  # noqa: E402
from video_edge_ai_proxy_b200.synth import exhaustive_yuv_frame, make_planes  # noqa: E402


@pytest.fixture(scope="module")
def path():
    if not torch.cuda.is_available():
        pytest.fail("marked gpu but no CUDA device")
    p = FramePath(0)
    yield p
    p.close()


def gpu_convert(path, fmt, planes, out=None, stream=None):
    t = [torch.from_numpy(np.ascontiguousarray(p)).cuda() for p in planes]
    h, w = planes[0].shape
    if out is None:
        out = torch.full((h, w, 3), 0xA5, dtype=torch.uint8, device="cuda")
    path.convert_batch([desc_from_tensors(fmt, t, out)], torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    return out.cpu().numpy()


def _digest(a):
    return int(a.sum(dtype=np.int64)), hashlib.sha256(a.tobytes()).hexdigest()


def test_golden_table_every_simd_row(path, golden_pixels):
    for r in golden_pixels["rows"]:
        if r["path"] != "simd":
            continue
        out = gpu_convert(path, r["fmt"], make_planes(r["fmt"], r["w"], r["h"], r["seed"]))
        assert _digest(out) == (r["sum"], r["sha256"]), r


def test_small_fixtures(path, small_frames):
    for fmt, w, h, seed, planes, want in small_frames:
        assert np.array_equal(gpu_convert(path, fmt, planes), want), (fmt, w, h)


def test_exhaustive_yuv_triples_planar_both_ranges(path):
    planes = exhaustive_yuv_frame()
    for fmt in ("yuv420p", "yuvj420p"):
        got = gpu_convert(path, fmt, planes)
        # oracle in 512-row bands to bound memory
        for r0 in range(0, 4096, 512):
            want = F.to_bgr24(fmt, (planes[0][r0:r0 + 512], planes[1][r0 // 2:r0 // 2 + 256], planes[2][r0 // 2:r0 // 2 + 256]))
            assert np.array_equal(got[r0:r0 + 512], want), (fmt, r0)


ALL_FORMATS = ["yuv420p", "yuvj420p", "nv12", "nv21", "yuv422p", "yuvj422p", "yuv420p10le", "yuv444p", "yuvj444p", "gray"]
TALL_ONLY = ("nv12", "nv21", "yuv420p10le")     # formats whose rule needs H >= 8


@pytest.mark.parametrize("fmt", ALL_FORMATS)
def test_seeded_random_config_sizes(path, fmt):
    for (w, h) in ((640, 480), (1280, 720), (1920, 1080), (3840, 2160)):
        for seed in (1, 2):
            planes = make_planes(fmt, w, h, seed)
            assert np.array_equal(gpu_convert(path, fmt, planes), F.to_bgr24(fmt, planes)), (fmt, w, h, seed)


@pytest.mark.parametrize("fmt", ALL_FORMATS)
def test_ragged_geometries_in_tight_buffers(path, fmt):
    for w in (2, 6, 14, 18, 30, 34, 62, 130, 510, 514, 642, 1922):
        for h in (2, 4, 8, 10, 18, 50):
            if fmt in TALL_ONLY and h < 8:
                continue
            planes = make_planes(fmt, w, h, 31 * w + h)
            assert np.array_equal(gpu_convert(path, fmt, planes), F.to_bgr24(fmt, planes)), (fmt, w, h)


@pytest.mark.parametrize("fmt", ["yuv420p", "nv12", "nv21", "yuvj422p", "yuv420p10le"])
def test_fast_path_widths_and_partial_segments(path, fmt):
    # multiples of 16 around the 512-px segment boundary: full, partial and single-lane segments
    for w in (16, 32, 496, 512, 528, 1008, 1024, 1040, 2048, 4096):
        planes = make_planes(fmt, w, 16, w)
        assert np.array_equal(gpu_convert(path, fmt, planes), F.to_bgr24(fmt, planes)), (fmt, w)


def test_live_libswscale_agrees_on_the_box(path):
    if not L.available():
        pytest.skip("bundled libswscale not found")
    assert L.on_simd_path()
    for fmt, w, h in (("yuv420p", 1920, 1080), ("yuvj420p", 640, 480), ("nv12", 1280, 720), ("yuv420p", 50, 6),
                      ("nv21", 1920, 1080), ("yuv422p", 1280, 720), ("yuvj422p", 1920, 1080), ("yuvj422p", 50, 6),
                      ("yuv420p10le", 1920, 1080), ("yuv420p10le", 50, 10), ("yuv444p", 1920, 1080), ("yuvj444p", 50, 6), ("gray", 1920, 1080), ("gray", 50, 6)):
        planes = make_planes(fmt, w, h, 77)
        assert np.array_equal(gpu_convert(path, fmt, planes), L.to_bgr24(fmt, planes, w, h)), (fmt, w, h)


def test_padded_strides_and_unaligned_pointers(path):
    rng = np.random.default_rng(5)
    for fmt, w, h, pad, off in (("yuv420p", 1920, 64, 64, 0), ("yuv420p", 640, 32, 48, 16), ("yuv420p", 640, 32, 7, 3),
                                ("nv12", 640, 32, 32, 0), ("nv12", 640, 32, 10, 2), ("yuvj420p", 320, 16, 0, 1)):
        planes = make_planes(fmt, w, h, 9)
        dev, ptrs, strides = [], [], []
        for p in planes:
            ph, pw = p.shape
            buf = torch.from_numpy(rng.integers(0, 256, (ph * (pw + pad) + off + 64,), dtype=np.uint8)).cuda()
            view = buf[off:off + ph * (pw + pad)].view(ph, pw + pad)
            view[:, :pw] = torch.from_numpy(p).cuda()
            dev.append(buf)
            ptrs.append(view.data_ptr())
            strides.append(pw + pad)
        dstride = 3 * w + (pad // 16) * 16 if off == 0 else 3 * w + pad
        obuf = torch.full((h * dstride + off + 64,), 0x5A, dtype=torch.uint8, device="cuda")
        d = desc_from_device_ptrs(fmt, w, h, ptrs, strides, obuf.data_ptr() + off, dstride)
        path.convert_batch([d], torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        o = obuf.cpu().numpy()
        rows = o[off:off + h * dstride].reshape(h, dstride)
        assert np.array_equal(rows[:, :3 * w].reshape(h, w, 3), F.to_bgr24(fmt, planes)), (fmt, w, h, pad, off)
        # nothing outside the visible W*3 bytes of each row is written (swscale overwrites padding; we must not)
        assert (rows[:, 3 * w:] == 0x5A).all() and (o[:off] == 0x5A).all() and (o[off + h * dstride:] == 0x5A).all()


def test_heterogeneous_batch_one_launch_with_skips(path):
    specs = [("yuv420p", 640, 480, 1, True), ("nv12", 1280, 720, 2, True), ("yuv420p", 1920, 1080, 3, False),
             ("yuvj420p", 642, 482, 4, True), ("yuv420p", 1920, 1080, 5, True), ("nv12", 50, 10, 6, True)]
    descs, keep = [], []
    for fmt, w, h, seed, conv in specs:
        planes = make_planes(fmt, w, h, seed)
        t = [torch.from_numpy(p).cuda() for p in planes]
        out = torch.full((h, w, 3), 0x11, dtype=torch.uint8, device="cuda")
        descs.append(desc_from_tensors(fmt, t, out, convert=conv))
        keep.append((fmt, planes, t, out, conv))
    before = path.stats()["kernel_launches"]
    plan = path.plan(descs)
    assert plan.n_converted == 5
    assert plan.alg_bytes == sum(int(4.5 * w * h) for _, w, h, _, c in specs if c)
    plan.launch(torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert path.stats()["kernel_launches"] == before + 1
    for fmt, planes, t, out, conv in keep:
        if conv:
            assert np.array_equal(out.cpu().numpy(), F.to_bgr24(fmt, planes)), fmt
        else:
            assert (out == 0x11).all()      # gate-skipped frame: untouched
    plan.close()


def test_linearity_free_properties_at_full_size(path):
    """Size-independent checks at 16 x 1080p: per-frame independence and determinism."""
    frames = [make_planes("yuv420p", 1920, 1080, s) for s in range(16)]
    descs, outs = [], []
    for pl in frames:
        t = [torch.from_numpy(p).cuda() for p in pl]
        out = torch.empty((1080, 1920, 3), dtype=torch.uint8, device="cuda")
        descs.append(desc_from_tensors("yuv420p", t, out))
        outs.append((t, out))
    plan = path.plan(descs)
    plan.launch(torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    batch = [o.cpu().numpy().copy() for _, o in outs]
    for i in (0, 7, 15):
        assert np.array_equal(batch[i], F.to_bgr24("yuv420p", frames[i]))
    plan.launch(torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    for i, (_, o) in enumerate(outs):
        assert np.array_equal(o.cpu().numpy(), batch[i])
    # gray input (U = V = 128) gives B = G = R everywhere
    Y = frames[0][0]
    C = np.full((540, 960), 128, np.uint8)
    g = gpu_convert(path, "yuv420p", (Y, C, C))
    assert (g[..., 0] == g[..., 1]).all() and (g[..., 1] == g[..., 2]).all()
    plan.close()


def _padded_case(fmt, w, h, seed, extra=0, off=0, fill_in=0x33, fill_out=0x5A):
    """Planes and output in buffers whose rows are padded to multiples of 16 bytes (+ extra) and start 16-byte
    aligned (+ off): the layout of an AVFrame with padded linesize, and what the host pipeline stages."""
    planes = make_planes(fmt, w, h, seed)
    bufs, ptrs, strides = [], [], []
    for p in planes:
        pb = np.ascontiguousarray(p).view(np.uint8).reshape(p.shape[0], -1)     # rows as bytes
        ph, pw = pb.shape
        st = (pw + 15) // 16 * 16 + extra
        buf = torch.full((ph * st + off + 64,), fill_in, dtype=torch.uint8, device="cuda")
        view = buf[off:off + ph * st].view(ph, st)
        view[:, :pw] = torch.from_numpy(pb).cuda()
        bufs.append(buf); ptrs.append(view.data_ptr()); strides.append(st)
    dstride = (3 * w + 15) // 16 * 16 + extra
    obuf = torch.full((h * dstride + off + 64,), fill_out, dtype=torch.uint8, device="cuda")
    desc = desc_from_device_ptrs(fmt, w, h, ptrs, strides, obuf.data_ptr() + off, dstride)
    return desc, (fmt, w, h, off, dstride, planes, bufs, obuf)


def _check_padded(case, fill_out=0x5A):
    fmt, w, h, off, dstride, planes, bufs, obuf = case
    o = obuf.cpu().numpy()
    rows = o[off:off + h * dstride].reshape(h, dstride)
    assert np.array_equal(rows[:, :3 * w].reshape(h, w, 3), F.to_bgr24(fmt, planes)), (fmt, w, h, off, dstride)
    # nothing outside the visible W*3 bytes of each row is written (swscale overwrites padding; we must not)
    assert (rows[:, 3 * w:] == fill_out).all() and (o[:off] == fill_out).all() and (o[off + h * dstride:] == fill_out).all(), (fmt, w, h, off, dstride)


@pytest.mark.parametrize("fmt", ALL_FORMATS)
def test_ragged_widths_on_the_vector_path(path, fmt):
    """Widths that are not a multiple of 16 with 16-byte padded rows (what the host pipeline stages and what a
    decoder's AVFrame looks like): the vector path with a partial last lane.  Visible bytes equal the oracle, the
    row padding of the output keeps its sentinel, whatever garbage the input padding holds."""
    for w in (2, 6, 18, 34, 50, 130, 510, 514, 854, 1366, 1922):
        for h, extra, off in ((8, 0, 0), (10, 16, 0), (18, 32, 16)):
            for fill_in in (0x00, 0xFF):                 # the input padding is read: it must not matter what it holds
                d, case = _padded_case(fmt, w, h, 7 * w + h, extra, off, fill_in=fill_in)
                path.convert_batch([d], torch.cuda.current_stream().cuda_stream)
                torch.cuda.synchronize()
                _check_padded(case)


@pytest.mark.parametrize("fmt", ["yuv420p", "yuvj420p", "nv12", "nv21", "yuv422p", "yuvj422p"])
def test_tight_rows_of_any_alignment_keep_the_vector_tiling(path, fmt):
    """Contiguous tensors whose width is not a multiple of 16 (every row starts at another alignment), and planes / outputs
    at odd byte offsets: loads and stores are assembled from the aligned chunks around them.  Oracle-exact, and nothing
    outside the visible bytes of the output is touched."""
    for w, h in ((854, 480), (1366, 768), (1922, 1080), (2, 8), (34, 10), (530, 8)):
        planes = make_planes(fmt, w, h, w + h)
        assert np.array_equal(gpu_convert(path, fmt, planes), F.to_bgr24(fmt, planes)), (fmt, w, h)
    rng = np.random.default_rng(11)
    for w, h, off, pad in ((640, 16, 1, 0), (640, 16, 3, 5), (854, 24, 7, 1), (1920, 8, 13, 3), (50, 10, 15, 9)):
        planes = make_planes(fmt, w, h, off)
        ptrs, strides, keep = [], [], []
        for p in planes:
            ph, pw = p.shape
            buf = torch.from_numpy(rng.integers(0, 256, (ph * (pw + pad) + off + 64,), dtype=np.uint8)).cuda()
            view = buf[off:off + ph * (pw + pad)].view(ph, pw + pad)
            view[:, :pw] = torch.from_numpy(p).cuda()
            keep.append(buf); ptrs.append(view.data_ptr()); strides.append(pw + pad)
        dstride = 3 * w + pad
        obuf = torch.full((h * dstride + off + 64,), 0x5A, dtype=torch.uint8, device="cuda")
        path.convert_batch([desc_from_device_ptrs(fmt, w, h, ptrs, strides, obuf.data_ptr() + off, dstride)], torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        o = obuf.cpu().numpy()
        rows = o[off:off + h * dstride].reshape(h, dstride)
        assert np.array_equal(rows[:, :3 * w].reshape(h, w, 3), F.to_bgr24(fmt, planes)), (fmt, w, h, off, pad)
        assert (rows[:, 3 * w:] == 0x5A).all() and (o[:off] == 0x5A).all() and (o[off + h * dstride:] == 0x5A).all(), (fmt, w, h, off, pad)


def test_ragged_widths_mixed_launch_and_device_table(path):
    """The same layouts in one mixed launch (parameter table) and in a 140-frame launch (device table)."""
    fmts_w = [("yuv420p", 854, 480), ("nv12", 1366, 768), ("yuvj420p", 50, 6), ("yuv422p", 854, 16), ("gray", 1366, 10),
              ("yuv444p", 130, 8), ("nv21", 514, 12), ("yuv420p", 1920, 1080), ("yuv420p10le", 50, 10), ("yuvj422p", 18, 4)]
    for n in (len(fmts_w), 140):
        descs, cases = [], []
        for i in range(n):
            fmt, w, h = fmts_w[i % len(fmts_w)]
            if n > 100:
                h = min(h, 24)
            d, case = _padded_case(fmt, w, h, 100 + i, extra=16 * (i % 2))
            descs.append(d); cases.append(case)
        path.convert_batch(descs, torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        for case in cases:
            _check_padded(case)


def test_large_batches_parameter_table_and_device_table(path, small_frames):
    for fmt, w, h, seed, planes, want in small_frames:
        assert np.array_equal(gpu_convert(path, fmt, planes), want), (fmt, w, h)
    # 40 frames in one launch (128-entry parameter table) and 150 frames (falls back to the device table)
    for n in (40, 150):
        descs, keep = [], []
        for i in range(n):
            fmt, w, h = (("yuv420p", 320, 240), ("nv12", 352, 288), ("yuvj420p", 160, 120), ("yuv420p", 50, 6),
                         ("yuvj422p", 320, 240), ("nv21", 64, 48), ("yuv420p10le", 160, 120), ("yuvj444p", 96, 64), ("gray", 64, 32))[i % 9]
            planes = make_planes(fmt, w, h, i)
            t = [torch.from_numpy(q).cuda() for q in planes]
            out = torch.empty((h, w, 3), dtype=torch.uint8, device="cuda")
            descs.append(desc_from_tensors(fmt, t, out))
            keep.append((fmt, planes, t, out))
        path.convert_batch(descs, torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        for fmt, planes, t, out in keep:
            assert np.array_equal(out.cpu().numpy(), F.to_bgr24(fmt, planes)), (n, fmt)
        plan = path.plan(descs)                          # the same batch as a plan, launched twice in one native call
        for _, _, _, out in keep:
            out.zero_()
        launch_sequence([plan, plan], torch.cuda.current_stream().cuda_stream, _abi.VEP_ORDER_STRICT)
        torch.cuda.synchronize()
        for fmt, planes, t, out in keep:
            assert np.array_equal(out.cpu().numpy(), F.to_bgr24(fmt, planes)), (n, fmt, "plan")
        plan.close()


def test_overlapped_and_ordered_launches(path):
    """Back-to-back launches: overlapped ones over disjoint buffers stay exact; ordered ones that
    re-use one output buffer leave the LAST launch's result (stream order is kept)."""
    st = torch.cuda.current_stream().cuda_stream
    sets = []
    for s in range(6):
        planes = [make_planes("yuv420p", 1920, 1080, 10 * s + i) for i in range(4)]
        descs, outs = [], []
        for pl in planes:
            t = [torch.from_numpy(q).cuda() for q in pl]
            out = torch.zeros((1080, 1920, 3), dtype=torch.uint8, device="cuda")
            descs.append(desc_from_tensors("yuv420p", t, out))
            outs.append((t, out))
        sets.append((planes, outs, path.plan(descs)))
    for _ in range(3):
        for planes, outs, plan in sets:
            plan.launch(st, overlapped=True)
    torch.cuda.synchronize()
    for planes, outs, plan in sets:
        for pl, (_, out) in zip(planes, outs):
            assert np.array_equal(out.cpu().numpy(), F.to_bgr24("yuv420p", pl))
    # ordered: 6 different inputs into ONE output buffer, default launch
    shared = torch.zeros((1080, 1920, 3), dtype=torch.uint8, device="cuda")
    plans, keep = [], []
    for s in range(6):
        pl = make_planes("yuv420p", 1920, 1080, 100 + s)
        t = [torch.from_numpy(q).cuda() for q in pl]
        keep.append((pl, t))
        plans.append(path.plan([desc_from_tensors("yuv420p", t, shared)]))
    for order in (_abi.VEP_ORDER_STRICT, _abi.VEP_ORDER_WRITES):     # both keep write-after-write order
        shared.zero_()
        for _ in range(3):
            for pln in plans:
                pln.launch(st, order=order)
        torch.cuda.synchronize()
        assert np.array_equal(shared.cpu().numpy(), F.to_bgr24("yuv420p", keep[-1][0])), order
    # the same through one native call per round (vep_plan_launch_sequence) and through vep_convert_batch_ordered
    shared.zero_()
    for _ in range(3):
        launch_sequence(plans, st, _abi.VEP_ORDER_WRITES)
    torch.cuda.synchronize()
    assert np.array_equal(shared.cpu().numpy(), F.to_bgr24("yuv420p", keep[-1][0]))
    shared.zero_()
    for pl, t in keep:
        path.convert_batch([desc_from_tensors("yuv420p", t, shared)], st, order=_abi.VEP_ORDER_WRITES)
    torch.cuda.synchronize()
    assert np.array_equal(shared.cpu().numpy(), F.to_bgr24("yuv420p", keep[-1][0]))
    launch_sequence([], st)
    with pytest.raises(VepError):
        plans[0].launch(st, order=7)
    with pytest.raises(VepError):
        launch_sequence(plans, st, 7)
    for _, _, pln in sets:
        pln.close()
    for pln in plans:
        pln.close()


def test_rejections(path):
    planes = make_planes("yuv420p", 64, 16, 0)
    t = [torch.from_numpy(p).cuda() for p in planes]
    out = torch.empty((16, 64, 3), dtype=torch.uint8, device="cuda")
    good = desc_from_tensors("yuv420p", t, out)
    for mutate, status in ((lambda d: setattr(d, "width", 63), _abi.VEP_ERR_GEOMETRY),
                           (lambda d: setattr(d, "height", 15), _abi.VEP_ERR_GEOMETRY),
                           (lambda d: setattr(d, "fmt", 99), _abi.VEP_ERR_INVALID_ARG),
                           (lambda d: setattr(d, "dst_stride", 100), _abi.VEP_ERR_GEOMETRY),
                           (lambda d: setattr(d, "dst", 0), _abi.VEP_ERR_INVALID_ARG)):
        d = desc_from_tensors("yuv420p", t, out)
        mutate(d)
        with pytest.raises(VepError) as e:
            path.convert_batch([good, d])
        assert e.value.status == status
    nv = make_planes("nv12", 16, 8, 0)
    tn = [torch.from_numpy(p).cuda() for p in nv]
    d = desc_from_tensors("nv12", tn, torch.empty((8, 16, 3), dtype=torch.uint8, device="cuda"))
    d.height = 6
    with pytest.raises(VepError) as e:
        path.convert_batch([d])
    assert e.value.status == _abi.VEP_ERR_GEOMETRY
    path.convert_batch([])          # empty batch is fine


def test_random_geometry_fuzz_one_launch_per_round(path):
    """Seeded fuzz: random formats, even sizes (biased to segment/vector boundaries), random row padding and
    pointer offsets, 24 frames per launch — every visible byte equals the oracle, no byte outside is touched."""
    rng = np.random.default_rng(2026)
    widths = [2, 4, 14, 16, 18, 30, 32, 34, 48, 62, 64, 96, 130, 496, 510, 512, 514, 528, 1008, 1024, 1040, 1056, 1922]
    for rnd in range(6):
        descs, keep = [], []
        for i in range(24):
            fmt = ALL_FORMATS[int(rng.integers(len(ALL_FORMATS)))]
            w = int(widths[int(rng.integers(len(widths)))])
            h = int(rng.choice([2, 4, 6, 8, 10, 12, 18, 34, 50]))
            if fmt in TALL_ONLY:
                h = max(h, 8)
            pad = int(rng.choice([0, 0, 16, 32, 7, 48]))
            off = int(rng.choice([0, 0, 16, 8, 3]))
            align_rows = bool(rng.integers(2))            # half the frames: rows padded up to 16-byte multiples (vector path for any width)
            if align_rows:
                pad, off = pad & ~15, off & ~15
            if fmt == "yuv420p10le":                      # 16-bit samples: planes and strides stay 2-byte aligned
                pad, off = pad & ~1, off & ~1
            planes = make_planes(fmt, w, h, 1000 * rnd + i)
            bufs, ptrs, strides = [], [], []
            for p in planes:
                p = np.ascontiguousarray(p).view(np.uint8).reshape(p.shape[0], -1)     # rows as bytes
                ph, pw = p.shape
                st_ = ((pw + 15) // 16 * 16 if align_rows else pw) + pad
                buf = torch.full((ph * st_ + off + 64,), 0x33, dtype=torch.uint8, device="cuda")
                view = buf[off:off + ph * st_].view(ph, st_)
                view[:, :pw] = torch.from_numpy(p).cuda()
                bufs.append(buf); ptrs.append(view.data_ptr()); strides.append(st_)
            dstride = ((3 * w + 15) // 16 * 16 if align_rows else 3 * w) + pad
            obuf = torch.full((h * dstride + off + 64,), 0x5A, dtype=torch.uint8, device="cuda")
            descs.append(desc_from_device_ptrs(fmt, w, h, ptrs, strides, obuf.data_ptr() + off, dstride))
            keep.append((fmt, w, h, pad, off, dstride, planes, bufs, obuf))
        path.convert_batch(descs, torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        for fmt, w, h, pad, off, dstride, planes, bufs, obuf in keep:
            o = obuf.cpu().numpy()
            rows = o[off:off + h * dstride].reshape(h, dstride)
            assert np.array_equal(rows[:, :3 * w].reshape(h, w, 3), F.to_bgr24(fmt, planes)), (rnd, fmt, w, h, pad, off)
            assert (rows[:, 3 * w:] == 0x5A).all() and (o[:off] == 0x5A).all() and (o[off + h * dstride:] == 0x5A).all(), (rnd, fmt, w, h, pad, off)


@pytest.mark.parametrize("fmt", ALL_FORMATS)
def test_known_answer_frames_and_video_legal_levels(path, fmt):
    """SURVEY §8(d) value distributions beyond uniform noise: all-0, all-max, a gray ramp with neutral chroma,
    and video-legal levels — each against the oracle, on a vector-path size and a ragged one."""
    ten = fmt == "yuv420p10le"
    top, mid = (1023, 512) if ten else (255, 128)
    for w, h in ((1040, 16), (50, 10)):
        ref = make_planes(fmt, w, h, 0)
        for kind in ("zero", "max", "ramp"):
            planes = []
            for i, p in enumerate(ref):
                if kind == "zero":
                    a = np.zeros_like(p)
                elif kind == "max":
                    a = np.full_like(p, top)
                elif i == 0:
                    a = (np.arange(p.shape[1], dtype=np.int64)[None, :] * (top + 1) // p.shape[1] + np.zeros((p.shape[0], 1), np.int64)).astype(p.dtype)
                else:
                    a = np.full_like(p, mid)
                planes.append(a)
            got = gpu_convert(path, fmt, planes)
            assert np.array_equal(got, F.to_bgr24(fmt, planes)), (fmt, w, h, kind)
            if kind == "ramp" and fmt not in ("yuv444p", "yuvj444p", "yuv420p10le"):   # (those round per channel / differ in the tail rows)
                assert (got[..., 0] == got[..., 1]).all() and (got[..., 1] == got[..., 2]).all()    # neutral chroma -> gray
        legal = make_planes(fmt, w, h, 3, dist="video")
        assert np.array_equal(gpu_convert(path, fmt, legal), F.to_bgr24(fmt, legal)), (fmt, w, h, "video")


def test_8k_frame_and_maximum_dimension(path):
    planes = make_planes("yuv420p", 7680, 4320, 8)
    got = gpu_convert(path, "yuv420p", planes)
    for r0 in range(0, 4320, 1080):      # oracle in bands to bound memory
        want = F.to_bgr24("yuv420p", (planes[0][r0:r0 + 1080], planes[1][r0 // 2:r0 // 2 + 540], planes[2][r0 // 2:r0 // 2 + 540]))
        assert np.array_equal(got[r0:r0 + 1080], want), r0
    # the widest and the tallest frame the ABI admits (16384): thin slices, both paths
    for fmt, w, h in (("yuv420p", 16384, 2), ("nv12", 16384, 8), ("yuv420p", 2, 16384), ("yuv420p", 16, 16384)):
        pl = make_planes(fmt, w, h, 9)
        assert np.array_equal(gpu_convert(path, fmt, pl), F.to_bgr24(fmt, pl)), (fmt, w, h)
    t = [torch.zeros((2, 16400), dtype=torch.uint8, device="cuda"), torch.zeros((1, 8200), dtype=torch.uint8, device="cuda"),
         torch.zeros((1, 8200), dtype=torch.uint8, device="cuda")]
    out = torch.zeros((2, 16400, 3), dtype=torch.uint8, device="cuda")
    with pytest.raises(VepError) as e:
        path.convert_batch([desc_from_tensors("yuv420p", t, out)])
    assert e.value.status == _abi.VEP_ERR_GEOMETRY
