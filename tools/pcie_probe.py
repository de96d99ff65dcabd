"""Measures what bounds the end-to-end number on a box: pinned H2D / D2H bandwidth (alone and
concurrently) and the host memcpy rate of the staging copy.  Diagnostic only (not part of bench.py)."""
import time
import numpy as np
import torch

dev = torch.device("cuda:0")
N = 100 * 1024 * 1024
hp = torch.empty(N, dtype=torch.uint8).pin_memory()
hp2 = torch.empty(N, dtype=torch.uint8).pin_memory()
d1 = torch.empty(N, dtype=torch.uint8, device=dev)
d2 = torch.empty(N, dtype=torch.uint8, device=dev)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def timeit(fn, n=10):
    fn(); torch.cuda.synchronize()
    t = time.perf_counter()
    for _ in range(n):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t) / n


def h2d():
    with torch.cuda.stream(s1):
        d1.copy_(hp, non_blocking=True)


def d2h():
    with torch.cuda.stream(s2):
        hp2.copy_(d2, non_blocking=True)


def both():
    h2d(); d2h()


print("H2D  GB/s", N / timeit(h2d) / 1e9)
print("D2H  GB/s", N / timeit(d2h) / 1e9)
t = timeit(both)
print("both GB/s per dir", N / t / 1e9, "sum", 2 * N / t / 1e9)
# many small-ish copies (3.1 MB / 6.2 MB like one 1080p frame)
chunks_in = [(hp[i * 3110400:(i + 1) * 3110400], d1[i * 3110400:(i + 1) * 3110400]) for i in range(16)]
chunks_out = [(hp2[i * 6220800:(i + 1) * 6220800], d2[i * 6220800:(i + 1) * 6220800]) for i in range(16)]
streams = [torch.cuda.Stream() for _ in range(16)]


def frames():
    for i in range(16):
        with torch.cuda.stream(streams[i]):
            chunks_in[i][1].copy_(chunks_in[i][0], non_blocking=True)
            chunks_out[i][0].copy_(chunks_out[i][1], non_blocking=True)


t = timeit(frames)
print("16 x (3.1MB H2D + 6.2MB D2H) on 16 streams: ms", t * 1e3, "fps", 16 / t, "GB/s", 16 * 9331200 / t / 1e9)
src = np.random.randint(0, 255, N, dtype=np.uint8)
hn = hp.numpy()
t = time.perf_counter()
for _ in range(5):
    hn[:] = src
print("host memcpy pageable->pinned GB/s (1 thread)", 5 * N / (time.perf_counter() - t) / 1e9)
from concurrent.futures import ThreadPoolExecutor
pool = ThreadPoolExecutor(16)
parts = [(hn[i * (N // 16):(i + 1) * (N // 16)], src[i * (N // 16):(i + 1) * (N // 16)]) for i in range(16)]


def cp(a):
    a[0][:] = a[1]


t = time.perf_counter()
for _ in range(5):
    list(pool.map(cp, parts))
print("host memcpy pageable->pinned GB/s (16 threads)", 5 * N / (time.perf_counter() - t) / 1e9)
