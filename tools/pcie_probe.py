"""What bounds the end-to-end number on a box: the DMA path between pinned host memory and the GPUs, measured with
all ranks copying at once.  Diagnostic, not part of bench.py (which carries a short version of phase `frames_both`).

    python tools/pcie_probe.py                                            # one GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29533 tools/pcie_probe.py

Phases (every rank runs the same one at the same time; GB/s are whole-box sums, time = max over ranks):
  h2d_only / d2h_only   100 MB chunks, one direction
  both_large            100 MB chunks, both directions at once
  frames_both           the 16 x 1080p tick: 16 x 3,110,400 B in + 16 x 6,220,800 B out, two streams   <- the e2e ceiling
  frames_both_pinned    the same after pinning each rank to its own share of the host cores
  host_copy             no DMA: T threads copy pageable -> pinned (the staging copy's memory traffic alone)
  frames_both+copy      frames_both while T threads keep copying (what the staging copy costs the DMA path)
One JSON line per phase on rank 0.
"""
import json
import os
import sys
import threading
import time

import numpy as np
import torch
import torch.distributed as dist

rank, local, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)


def red(v, op):
    if world == 1:
        return float(v)
    t = torch.tensor([float(v)], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=op)
    return float(t.item())


def bar():
    if world > 1:
        dist.barrier()


s_in, s_out = torch.cuda.Stream(dev), torch.cuda.Stream(dev)


def phase(name, in_sizes, out_sizes, target_s=0.5, note=""):
    h_in = [torch.empty(s, dtype=torch.uint8).pin_memory() for s in in_sizes]
    h_out = [torch.empty(s, dtype=torch.uint8).pin_memory() for s in out_sizes]
    d_in = [torch.empty(s, dtype=torch.uint8, device=dev) for s in in_sizes]
    d_out = [torch.zeros(s, dtype=torch.uint8, device=dev) for s in out_sizes]
    for t in h_in:
        t.fill_(3)

    def run(n):
        for _ in range(n):
            with torch.cuda.stream(s_in):
                for a, b in zip(d_in, h_in):
                    a.copy_(b, non_blocking=True)
            with torch.cuda.stream(s_out):
                for a, b in zip(h_out, d_out):
                    a.copy_(b, non_blocking=True)

    run(2)
    torch.cuda.synchronize(dev)
    t0 = time.perf_counter(); run(2); torch.cuda.synchronize(dev)
    one = red((time.perf_counter() - t0) / 2, dist.ReduceOp.MAX if world > 1 else None)
    reps = max(3, int(target_s / max(one, 1e-5)))
    bar()
    t0 = time.perf_counter()
    run(reps)
    torch.cuda.synchronize(dev)
    dt = red(time.perf_counter() - t0, dist.ReduceOp.MAX if world > 1 else None)
    bar()
    nb_in, nb_out = sum(in_sizes) * reps * world, sum(out_sizes) * reps * world
    if rank == 0:
        print(json.dumps({"phase": name, "n_gpus": world, "h2d_gbs": round(nb_in / dt / 1e9, 2), "d2h_gbs": round(nb_out / dt / 1e9, 2),
                          "sum_gbs": round((nb_in + nb_out) / dt / 1e9, 2), "per_gpu_sum_gbs": round((nb_in + nb_out) / dt / 1e9 / world, 2),
                          "frames_per_s_1080p": round(len(out_sizes) * reps * world / dt, 1) if out_sizes and out_sizes[0] == 6220800 else None,
                          "reps": reps, "seconds": round(dt, 3), "note": note}), flush=True)
    return dt


BIG = 100 * 1024 * 1024
F_IN, F_OUT = [3110400] * 16, [6220800] * 16
cores = sorted(os.sched_getaffinity(0))
if rank == 0:
    topo = os.popen("nvidia-smi topo -m 2>/dev/null | head -14").read()
    numa = os.popen("lscpu 2>/dev/null | grep -i -E 'numa|model name|^cpu\\(s\\)|socket'").read()
    print(json.dumps({"phase": "host", "n_gpus": world, "cores_visible": len(cores), "lscpu": numa.strip().splitlines(), "topo": topo.strip().splitlines()}), flush=True)

if os.environ.get("PROBE_ONLY_MERGE"):      # short form: does the copy granularity move the box ceiling?
    phase("frames_both", F_IN, F_OUT, note="16 + 16 copies per tick")
    phase("tick_merged", [sum(F_IN)], [sum(F_OUT)], note="one 49.8 MB copy in + one 99.5 MB copy out per tick")
    phase("tick_merged_4", [sum(F_IN) // 4] * 4, [sum(F_OUT) // 4] * 4, note="four copies per direction and tick")
    if world > 1:
        dist.destroy_process_group()
    sys.exit(0)
phase("h2d_only", [BIG], [])
phase("d2h_only", [], [BIG])
phase("both_large", [BIG], [BIG])
phase("frames_both", F_IN, F_OUT, note="the 16 x 1080p tick; this is the e2e ceiling")
if os.environ.get("PROBE_EXTRA"):
    phase("frames_h2d_only", F_IN, [], note="16 x 3.1 MB per tick, H2D alone")
    phase("frames_d2h_only", [], F_OUT, note="16 x 6.2 MB per tick, D2H alone")
    phase("tick_merged", [sum(F_IN)], [sum(F_OUT)], note="one 49.8 MB copy in + one 99.5 MB copy out per tick")
    phase("frames_3_copies_in", [2073600, 518400, 518400] * 16, F_OUT, note="H2D as three copies per frame (Y, U, V separately)")

# ---- the same with each rank on its own cores ----
per = max(1, len(cores) // world)
mine = cores[local * per:(local + 1) * per]
os.sched_setaffinity(0, mine)
phase("frames_both_pinned", F_IN, F_OUT, note=f"each rank pinned to {per} own cores")

# ---- host copies: pageable -> pinned with this rank's cores, alone and under DMA ----
T = max(1, per - 1)
src = [np.random.randint(0, 255, 3110400, dtype=np.uint8) for _ in range(16)]
dst_t = [torch.empty(3110400, dtype=torch.uint8).pin_memory() for _ in range(16)]
dst = [t.numpy() for t in dst_t]
stop = threading.Event()
copied = [0] * T


def copier(k):
    i = k
    while not stop.is_set():
        np.copyto(dst[i % 16], src[i % 16])      # memcpy, GIL released
        copied[k] += 3110400
        i += T


def with_copiers(fn):
    stop.clear()
    for k in range(T):
        copied[k] = 0
    th = [threading.Thread(target=copier, args=(k,)) for k in range(T)]
    [t.start() for t in th]
    time.sleep(0.05)
    c0, t0 = sum(copied), time.perf_counter()
    r = fn()
    c1, t1 = sum(copied), time.perf_counter()
    stop.set()
    [t.join() for t in th]
    return r, (c1 - c0) / (t1 - t0) / 1e9


bar()
_, gbs = with_copiers(lambda: time.sleep(0.5))
tot = red(gbs, dist.ReduceOp.SUM if world > 1 else None)
if rank == 0:
    print(json.dumps({"phase": "host_copy", "n_gpus": world, "threads_per_rank": T, "copy_gbs_whole_box": round(tot, 2),
                      "note": "numpy memcpy pageable -> pinned, 3.1 MB pieces; memory traffic is about twice this (read + write)"}), flush=True)
bar()
_, gbs = with_copiers(lambda: phase("frames_both+copy", F_IN, F_OUT, note=f"while {T} threads per rank copy pageable -> pinned"))
tot = red(gbs, dist.ReduceOp.SUM if world > 1 else None)
if rank == 0:
    print(json.dumps({"phase": "copy_during_dma", "n_gpus": world, "copy_gbs_whole_box": round(tot, 2)}), flush=True)
if world > 1:
    dist.destroy_process_group()
