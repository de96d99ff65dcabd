"""Soak: many ticks, streams opened and closed, plans created and destroyed — device/host memory must not grow and
results must stay exact.  usage: soak.py [seconds]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import formulas as F
from video_edge_ai_proxy_b200 import FramePath, desc_from_tensors, make_meta
from video_edge_ai_proxy_b200.synth import make_planes

secs = float(sys.argv[1]) if len(sys.argv) > 1 else 30
fmts = ["yuv420p", "nv12", "yuvj422p", "yuv420p10le", "yuv444p", "gray", "nv21", "yuvj420p"]
rng = np.random.default_rng(0)
path = FramePath(0)
want = {}
def free_mb(): return torch.cuda.mem_get_info()[0] / 1e6
def rss_mb():      # CURRENT resident set (not the peak): sampled after every stream of the iteration has been closed
    with open("/proc/self/statm") as f:
        return int(f.read().split()[1]) * os.sysconf("SC_PAGE_SIZE") / 1e6
t_end = time.time() + secs
it = ticks = 0
base_free = base_rss = None
while time.time() < t_end:
    names = []
    for c in range(6):
        fmt = fmts[(it + c) % len(fmts)]
        w, h = [(640, 480), (1280, 720), (50, 10), (1920, 1080), (854, 480)][(it + c) % 5]
        names.append((f"s{it}_{c}", fmt, w, h))
    if it % 2:                                     # alternately one by one and as a group with shared slot memory
        for nm, fmt, w, h in names:
            path.open_stream(nm, fmt, w, h, memory_buffer=2)
    else:
        path.open_streams(names, memory_buffer=2)
    path.set_staging_threads(it % 3)
    for k in range(20):
        for nm, fmt, w, h in names:
            key = (fmt, w, h, k % 3)
            if key not in want:
                want[key] = (make_planes(fmt, w, h, k % 3), None)
                want[key] = (want[key][0], F.to_bgr24(fmt, want[key][0]))
            if k % 4 != 3:
                path.submit(nm, want[key][0], make_meta(packet=k))
        if k % 4 == 3:                             # every fourth tick through the batch call
            path.submit_many([(nm, want[(fmt, w, h, k % 3)][0], make_meta(packet=k)) for nm, fmt, w, h in names])
        path.tick_begin()
        if k % 2:
            path.tick_wait(); path.tick_wait()
        ticks += 1
    path.tick_wait(); path.tick_wait(); path.tick_wait()
    for nm, fmt, w, h in names:
        assert np.array_equal(path.result(nm), want[(fmt, w, h, 19 % 3)][1]), (nm, fmt, w, h)
    # a plan per iteration, created and destroyed
    pl = want[("yuv420p", 640, 480, 0)][0] if ("yuv420p", 640, 480, 0) in want else make_planes("yuv420p", 640, 480, 0)
    t = [torch.from_numpy(p).cuda() for p in pl]
    out = torch.empty((480, 640, 3), dtype=torch.uint8, device="cuda")
    plan = path.plan([desc_from_tensors("yuv420p", t, out)])
    plan.launch(torch.cuda.current_stream().cuda_stream); torch.cuda.synchronize(); plan.close()
    for nm, *_ in names:
        path.close_stream(nm)
    it += 1
    if it == 40:       # past every (format, size, group/single) combination's first use: allocator pools have their final size
        torch.cuda.empty_cache(); base_free, base_rss = free_mb(), rss_mb()
torch.cuda.empty_cache()
print(f"SOAK iterations={it} ticks={ticks} device_free_delta_mb={free_mb() - base_free:.1f} host_rss_delta_mb={rss_mb() - base_rss:.1f} (current RSS after close, iteration 40 -> {it})")
path.close()
