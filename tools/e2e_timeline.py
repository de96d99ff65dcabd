"""Diagnostic: where does a tick's wall time go in the host pipeline (submit / begin / wait)?"""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from concurrent.futures import ThreadPoolExecutor
from video_edge_ai_proxy_b200 import FramePath, make_meta
from video_edge_ai_proxy_b200.synth import make_planes

n = 16
path = FramePath(0)
planes = [make_planes("yuv420p", 1920, 1080, i) for i in range(n)]
names = [f"c{i}" for i in range(n)]
for nm in names:
    path.open_stream(nm, "yuv420p", 1920, 1080)
meta = make_meta(frame_type="I")
nthreads = int(sys.argv[1]) if len(sys.argv) > 1 else 16
pool = ThreadPoolExecutor(nthreads)
def submit_all():
    list(pool.map(lambda a: path.submit(a[0], a[1], meta), zip(names, planes)))
for _ in range(3):
    submit_all(); path.tick()
# serial phases
for label, fn in (("submit_all", submit_all), ("tick (sync)", path.tick)):
    ts = []
    for _ in range(10):
        if label != "submit_all":
            submit_all()
        t = time.perf_counter(); fn(); ts.append(time.perf_counter() - t)
        if label == "submit_all":
            path.tick()
    print(label, "ms median", sorted(ts)[5] * 1e3)
print("kernel ms", path.stats()["last_kernel_ms"])
# pipelined
T = {"submit": [], "begin": [], "wait": []}
t0 = time.perf_counter()
K = 40
DEPTH = int(sys.argv[2]) if len(sys.argv) > 2 else 3
waiting = 0
for i in range(K):
    a = time.perf_counter(); submit_all(); b = time.perf_counter(); path.tick_begin(); c = time.perf_counter(); waiting += 1
    if waiting == DEPTH:
        path.tick_wait(); waiting -= 1
    d = time.perf_counter()
    T["submit"].append(b - a); T["begin"].append(c - b); T["wait"].append(d - c)
while waiting:
    path.tick_wait(); waiting -= 1
tot = time.perf_counter() - t0
print("pipelined: fps", n * K / tot, "ms/tick", tot / K * 1e3, {k: round(sorted(v)[len(v) // 2] * 1e3, 3) for k, v in T.items()})
# zero-copy staging (decoder writes straight into pinned slots): acquire/commit only
t0 = time.perf_counter()
for nm in names: path.acquire_input(nm); path.commit_input(nm, meta)
path.tick_begin()
for i in range(1, K):
    for nm in names: path.acquire_input(nm); path.commit_input(nm, meta)
    path.tick_begin(); path.tick_wait()
path.tick_wait()
tot = time.perf_counter() - t0
print("no staging copy: fps", n * K / tot, "ms/tick", tot / K * 1e3)
