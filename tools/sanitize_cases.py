"""Small multi-geometry run for compute-sanitizer (memcheck / racecheck): every path of the selected kernel
(VEP_KERNEL / VEP_RPT env), device batch + host pipeline, compared with the oracle."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import formulas as F
from video_edge_ai_proxy_b200 import FramePath, desc_from_tensors, make_meta, convert_batch_noalloc
from video_edge_ai_proxy_b200.synth import make_planes

cases = [("yuv420p", 1040, 6), ("yuv420p", 528, 10), ("yuvj420p", 64, 4), ("nv12", 1040, 10), ("nv12", 64, 8), ("yuv420p", 50, 6),
         ("nv12", 38, 12), ("yuv420p", 2, 2), ("yuv420p", 48, 8), ("nv12", 528, 18)]
with FramePath(0) as path:
    st = torch.cuda.current_stream().cuda_stream
    descs, keep = [], []
    for i, (fmt, w, h) in enumerate(cases):
        pl = make_planes(fmt, w, h, i)
        t = [torch.from_numpy(p).cuda() for p in pl]
        out = torch.zeros((h, w, 3), dtype=torch.uint8, device="cuda")
        descs.append(desc_from_tensors(fmt, t, out)); keep.append((fmt, pl, t, out))
    path.convert_batch(descs, st); torch.cuda.synchronize()
    ok = all(np.array_equal(o.cpu().numpy(), F.to_bgr24(f, pl)) for f, pl, t, o in keep)
    plan = path.plan(descs)
    plan.launch(st, True); plan.launch(st, False); torch.cuda.synchronize()
    ok &= all(np.array_equal(o.cpu().numpy(), F.to_bgr24(f, pl)) for f, pl, t, o in keep)
    for f, pl, t, o in keep: o.zero_()
    convert_batch_noalloc(descs, st); torch.cuda.synchronize()
    ok &= all(np.array_equal(o.cpu().numpy(), F.to_bgr24(f, pl)) for f, pl, t, o in keep)
    for i, (fmt, w, h) in enumerate(cases[:5]):
        path.open_stream(f"c{i}", fmt, w, h)
        path.submit(f"c{i}", make_planes(fmt, w, h, i), make_meta(packet=i))
    path.tick()
    ok &= all(np.array_equal(path.result(f"c{i}"), F.to_bgr24(fmt, make_planes(fmt, w, h, i))) for i, (fmt, w, h) in enumerate(cases[:5]))
print("SANITIZE_CASES", "OK" if ok else "MISMATCH")
