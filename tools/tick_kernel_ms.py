"""Device time of the conversion kernel INSIDE the library's own tick (vep_stats.last_kernel_ms: CUDA events around
the launch on the context stream), for a few tick shapes, with ticks issued back to back (4 in flight) so that the
launch order the tick uses (VEP_ORDER_NONE: programmatic dependent launch, no wait on the previous tick's kernel) can
show.  Beside it the same frames as a device-resident plan launched in strict order and with VEP_ORDER_NONE.
usage: python tools/tick_kernel_ms.py"""
import os
import statistics
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from video_edge_ai_proxy_b200 import FramePath, _abi, desc_from_tensors, launch_sequence, make_meta
from video_edge_ai_proxy_b200.synth import make_planes

SHAPES = [("one 4K NV12 frame per tick (config 4)", [("nv12", 3840, 2160)]),
          ("16 x 1080p per tick (config 2)", [("yuv420p", 1920, 1080)] * 16),
          ("32 mixed frames per tick (config 5's average tick)", [("yuv420p", 640, 480)] * 11 + [("yuv420p", 1280, 720)] * 11 + [("yuv420p", 1920, 1080)] * 10)]
meta = make_meta(frame_type="I")
for label, cams in SHAPES:
    with FramePath(0) as path:
        names = [f"c{i}" for i in range(len(cams))]
        path.open_streams([(n, f, w, h) for n, (f, w, h) in zip(names, cams)])
        planes = [make_planes(f, w, h, i) for i, (f, w, h) in enumerate(cams)]
        alg = sum(4.5 * w * h for _, w, h in cams)
        for _ in range(8):                                   # every slot gets its frame once
            for n, p in zip(names, planes):
                for v, q in zip(path.acquire_input(n), p):
                    v[...] = q
                path.commit_input(n, meta)
            path.tick()
        ms = []
        pending = 0
        for i in range(200):
            for n in names:
                path.acquire_raw(n); path.commit_input(n, meta)
            path.tick_begin(); pending += 1
            if pending == 4:
                path.tick_wait(); pending -= 1
                ms.append(path.stats()["last_kernel_ms"])
        while pending:
            path.tick_wait(); pending -= 1
        tick_ms = statistics.median(ms[20:])
        # the same frames, device resident, as a plan: strict order and no order, 200 launches each
        t = [[torch.from_numpy(q).cuda() for q in p] for p in planes]
        outs = [torch.empty((h, w, 3), dtype=torch.uint8, device="cuda") for _, w, h in cams]
        plan = path.plan([desc_from_tensors(f, tt, o) for (f, _, _), tt, o in zip(cams, t, outs)])
        st = torch.cuda.current_stream().cuda_stream
        res = {}
        for order, nm in ((_abi.VEP_ORDER_STRICT, "strict"), (_abi.VEP_ORDER_NONE, "none")):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); e1.record()
            launch_sequence([plan] * 20, st, order)
            torch.cuda.synchronize()
            launch_sequence([plan] * 200, st, order, e0.cuda_event, e1.cuda_event)
            torch.cuda.synchronize()
            res[nm] = e0.elapsed_time(e1) / 200
        # one launch at a time after the GPU sat idle for 2 ms (what the PCIe-bound tick pipeline looks like to the SMs)
        import time
        idle = []
        for _ in range(30):
            torch.cuda.synchronize(); time.sleep(0.002)
            launch_sequence([plan], st, _abi.VEP_ORDER_STRICT, e0.cuda_event, e1.cuda_event)
            torch.cuda.synchronize()
            idle.append(e0.elapsed_time(e1))
        res["idle"] = statistics.median(idle)
        plan.close()
        print(f"  the same plan, one launch after 2 ms of idle GPU: {res['idle'] * 1e3:.1f} us")
        print(f"{label}: kernel inside the tick {tick_ms * 1e3:.1f} us = {alg / tick_ms / 1e6:.0f} GB/s algorithmic; "
              f"plan back to back: strict {res['strict'] * 1e3:.1f} us ({alg / res['strict'] / 1e6:.0f} GB/s), no order {res['none'] * 1e3:.1f} us ({alg / res['none'] / 1e6:.0f} GB/s)")
