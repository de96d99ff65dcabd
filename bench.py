#!/usr/bin/env python
"""bench.py — headline benchmark of the B200 frame path.

Metric (BASELINE.json): 1080p BGR24 frames/s, whole job, and achieved HBM GB/s vs the measured peak.
Default workload = BASELINE.json configs[1]: 16 synthetic 1080p yuv420p cameras per GPU, one step =
one tick = ONE batched kernel launch converting all 16 frames.  Under torchrun every rank owns its
own 16 cameras on its own GPU (weak scaling, no data-path collective); times are max over ranks.

One JSON line on rank 0:
  value        frames/s with the decoded planes already resident in HBM (kernel-only, CUDA events on
               the launching stream, K steps issued back to back by ONE native call, inputs rotate over
               buffer sets > L2)
  e2e          frames/s through the public API with HOST buffers: FramePath.submit (pinned staging) ->
               tick (H2D, kernel, D2H) -> result_packed, copies inside the timed region; beside it the
               box's DMA ceiling for the same byte mix, probed in the same run on all ranks at once
               (pinned H2D + D2H only, no kernel, no staging copy), and e2e as a fraction of it
  parity       before anything is timed: every camera's output of buffer set 0 and every message of the
               first end-to-end tick are compared byte for byte with the live libswscale; all other
               buffer sets must equal set 0 exactly.  A mismatch aborts the run.
  roofline     algorithmic bytes (4.5 B/px) per launch / measured launch duration vs MEASURED_PEAKS.json; the headline
               launches consecutive ticks with VEP_ORDER_NONE (disjoint buffer sets); the same K ticks with only the
               stores ordered (`ordered_writes`) and in full stream order (`strict_stream_order`) are reported beside it
  cpu_baseline the reference's CPU path (libswscale sws_scale + tobytes, read_image.py:94-97) timed on
               this box's host cores on a bounded sample of the same frames
`--impl reference` prints the CPU arm alone in the same shape (rank 0 only under torchrun).
"""
from __future__ import annotations

import argparse
import json
import math
import os
import statistics
import sys
import threading
import time
from collections import deque
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "1080p_bgr24_frames_per_sec"
UNIT = "frames/s"
ALG_BYTES_PER_PX = 4.5          # 1.5 B/px read (Y + U/4 + V/4) + 3 B/px written
ALG_BYTES_BY_FMT = {"yuv422p": 5.0, "yuvj422p": 5.0, "yuv420p10le": 6.0, "yuv444p": 6.0, "yuvj444p": 6.0}   # formats that read more
FALLBACK_HBM_GBS = 6650.0       # B200_PROFILING.md fallback when MEASURED_PEAKS.json is absent

WORKLOADS = {
    # name: (cameras per GPU, fmt, w, h, description)
    "16x1080p": (16, "yuv420p", 1920, 1080, "16 synthetic 1080p yuv420p cameras per GPU, one batched launch per tick (BASELINE configs[1])"),
    "8x1080p": (8, "yuv420p", 1920, 1080, "8 synthetic 1080p yuv420p cameras per GPU (BASELINE configs[2] layout: 64 cameras over 8 GPUs)"),
    "8x4k-nv12": (8, "nv12", 3840, 2160, "8 synthetic 3840x2160 NV12 cameras per GPU, every frame converted"),
    "8x4k-nv12-keyframe": (8, "nv12", 3840, 2160, "8 synthetic 3840x2160 NV12 cameras per GPU, key-frame-only: GOP 30, staggered phases, only key frames are converted (BASELINE configs[3])"),
    "mixed128-skip": (128, "yuv420p", 0, 0, "128-camera fleet per GPU: 43 x 640x480, 43 x 1280x720, 42 x 1920x1080 yuv420p, interval-skip: each camera converts a frame with p = 0.25 per tick (seeded Bernoulli, seed = tick) (BASELINE configs[4])"),
    "16x1080p-yuvj422p": (16, "yuvj422p", 1920, 1080, "16 synthetic 1080p yuvj422p (MJPEG) cameras per GPU; 5 B/px"),
    "16x1080p-yuv420p10le": (16, "yuv420p10le", 1920, 1080, "16 synthetic 1080p 10-bit 4:2:0 cameras per GPU; 6 B/px"),
    "16x1080p-yuv444p": (16, "yuv444p", 1920, 1080, "16 synthetic 1080p 4:4:4 cameras per GPU; 6 B/px"),
    "64x1080p": (64, "yuv420p", 1920, 1080, "64 synthetic 1080p yuv420p cameras on one GPU (launch-overhead study, not a BASELINE config)"),
    "1x480p": (1, "yuv420p", 640, 480, "one 640x480 yuv420p stream (BASELINE configs[0])"),
    "16x854x480": (16, "yuv420p", 854, 480, "16 synthetic 854x480 yuv420p cameras per GPU: width not a multiple of 16 (rows padded to 16-byte strides, partial last lane)"),
    "16x1366x768": (16, "yuv420p", 1366, 768, "16 synthetic 1366x768 yuv420p cameras per GPU: width not a multiple of 16 (rows padded to 16-byte strides, partial last lane)"),
    "16x854x480-tight": (16, "yuv420p", 854, 480, "16 synthetic 854x480 yuv420p cameras per GPU in TIGHT device buffers (stride = width, rows at every alignment): loads and stores assembled from aligned chunks"),
}
MIXED_SIZES = [(640, 480)] * 43 + [(1280, 720)] * 43 + [(1920, 1080)] * 42


# ------------------------------------------------------------------------------------------------
# rank plumbing (gloo on CPU in tests, nccl on the box)
# ------------------------------------------------------------------------------------------------
def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


def cameras_for_rank(workload: str, rank: int, world: int):
    """Static partition: rank r owns cameras [r*n, (r+1)*n) — disjoint sets, no exchange step."""
    n, fmt, w, h, _ = WORKLOADS[workload]
    if workload == "mixed128-skip":
        return [dict(camera=rank * n + i, fmt=fmt, w=MIXED_SIZES[i][0], h=MIXED_SIZES[i][1], seed=rank * n + i) for i in range(n)]
    return [dict(camera=rank * n + i, fmt=fmt, w=w, h=h, seed=rank * n + i) for i in range(n)]


def schedule_for(workload: str, n_cams: int):
    """The repeating tick pattern: for each tick of the cycle, the local camera indices whose frame is
    converted (the decode gate's decision; every camera is OFFERED one frame per tick)."""
    if workload == "8x4k-nv12-keyframe":       # GOP 30, camera c's key frames at ticks = phase_c (mod 30)
        phase = [(c * 30) // n_cams for c in range(n_cams)]
        return [[c for c in range(n_cams) if phase[c] == t] for t in range(30)]
    if workload == "mixed128-skip":            # 16 distinct ticks, selection seeded by the tick index
        return [[int(c) for c in np.nonzero(np.random.default_rng(t).random(n_cams) < 0.25)[0]] for t in range(16)]
    return [list(range(n_cams))]


def config_for(workload: str, steps: int):
    """The `config` object of the JSON line — the same for both arms (it names the workload, not the arm)."""
    n, fmt, w, h, desc = WORKLOADS[workload]
    sched = schedule_for(workload, n)
    conv = sum(len(sched[i % len(sched)]) for i in range(steps))
    return {"workload": f"{workload}: {desc}", "cameras_per_gpu": n, "format": fmt,
            "converted_frames_per_step_per_gpu": round(conv / max(steps, 1), 3),
            "parallelism": f"cameras sharded {n}/GPU, no collective",
            "l2": "b200 arm: inputs and outputs rotate over buffer sets whose footprint is >= 3x the GPU's L2 (at least 9 sets; "
                  "count in roofline.buffer_sets); reference arm: one buffer set per camera in host memory"}


def _reduce(value: float, op, device: str):
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=op)
    return float(t.item())


def max_over_ranks(value: float, device: str = "cuda") -> float:
    import torch.distributed as dist
    return _reduce(value, dist.ReduceOp.MAX, device)


def sum_over_ranks(value: float, device: str = "cuda") -> float:
    import torch.distributed as dist
    return _reduce(value, dist.ReduceOp.SUM, device)


def barrier():
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        dist.barrier()


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def pin_rank_cores(local_rank: int, world: int):
    """Give this rank its own, disjoint share of the host cores (one host process per GPU with its
    own cores): producer threads, the ticker, the CUDA and NCCL helper threads of a rank then never migrate onto
    another rank's cores.  Threads started later inherit the mask.  Returns the core list (None: left alone)."""
    if world <= 1 or os.environ.get("VEP_BENCH_NO_AFFINITY"):
        return None
    try:
        cores = sorted(os.sched_getaffinity(0))
        per = len(cores) // world
        if per < 1:
            return None
        mine = cores[local_rank * per:(local_rank + 1) * per]
        os.sched_setaffinity(0, mine)
        return mine
    except Exception:
        return None


# ------------------------------------------------------------------------------------------------
# clocks (NVML in-process; nvidia-smi fallback)
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    REASONS = {0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
               0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
               0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting"}

    def __init__(self, index: int, period_s: float = 0.004):
        self.samples = []   # (t, sm_mhz, reasons_mask, power_w)
        self.period = period_s
        self._stop = threading.Event()
        self._thr = None
        self.sm_max = None
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            # honour CUDA_VISIBLE_DEVICES ordering when it is a plain index list
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = index
            if vis:
                try:
                    phys = int(vis.split(",")[index])
                except Exception:
                    phys = index
            self._nv = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.sm_max = pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception as e:  # pragma: no cover
            self.err = str(e)

    def _run(self):
        nv, h = self._nv, self._h
        while not self._stop.is_set():
            try:
                sm = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
                try:
                    rs = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                except Exception:
                    rs = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                try:
                    pw = nv.nvmlDeviceGetPowerUsage(h) / 1000.0
                except Exception:
                    pw = 0.0
                self.samples.append((time.perf_counter(), sm, rs, pw))
            except Exception:
                pass
            time.sleep(self.period)

    def start(self):
        if self.ok:
            self._thr = threading.Thread(target=self._run, daemon=True)
            self._thr.start()

    def stop(self):
        self._stop.set()
        if self._thr:
            self._thr.join(timeout=1)

    def summary(self, windows):
        """windows: list of (t0, t1) perf_counter intervals that were under load."""
        if not self.ok:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "note": "NVML unavailable"}
        sel = [s for s in self.samples if any(a <= s[0] <= b for a, b in windows)]
        note = "sampled inside the timed regions"
        if not sel:
            sel, note = self.samples, "timed regions shorter than the sampling period; all samples of the run"
        mask = 0
        for s in sel:
            mask |= s[2]
        reasons = [n for b, n in self.REASONS.items() if mask & b and n != "gpu_idle"]
        return {"sm_mhz": statistics.median(s[1] for s in sel) if sel else None, "sm_max_mhz": self.sm_max,
                "reasons": reasons, "samples": len(sel), "power_w_max": round(max((s[3] for s in sel), default=0.0), 1),
                "note": note}


# ------------------------------------------------------------------------------------------------
# CPU arm: the reference's own path on host cores (libswscale via the oracle binding)
# ------------------------------------------------------------------------------------------------
class CpuArm:
    """One SwsContext + buffers per camera, a thread pool, and `run(steps)`: each step converts the frames the tick's
    schedule selects (default: one frame of every camera — the same tick the GPU arm converts), spread over the
    threads; ctypes releases the GIL inside sws_scale."""

    def __init__(self, cams, threads: int, sched=None):
        from oracle import swscale_live as live   # bench.py's CPU legs are allowed to execute oracle/
        from video_edge_ai_proxy_b200.synth import make_planes
        if not live.available():
            raise RuntimeError("bundled libswscale not found")
        self.version = live.version()
        self.threads = max(1, min(threads, len(cams)))
        self.convs = []
        for c in cams:
            cv = live.Converter(c["fmt"], c["w"], c["h"])
            cv.load(make_planes(c["fmt"], c["w"], c["h"], c["seed"]))
            self.convs.append(cv)
        self.sched = sched or [list(range(len(cams)))]
        self.pool = ThreadPoolExecutor(self.threads)

    def run(self, steps: int, warmup: int = 0, with_tobytes: bool = True, fresh_ctx: bool = False, threads: int = 0):
        def one(cv):
            if fresh_ctx:                                # PyAV builds a reformatter (SwsContext) per decoded frame
                cv.renew_context()
            cv.scale()                                   # frame.to_ndarray(format='bgr24')  read_image.py:94
            if with_tobytes:
                return cv.result().nbytes                # one H*W*3 copy = np.ndarray.tobytes(img)  :97
            return 0

        pool = self.pool if not threads else ThreadPoolExecutor(threads)
        done = 0
        for i in range(warmup):
            list(pool.map(one, [self.convs[c] for c in self.sched[i % len(self.sched)]]))
        t0 = time.perf_counter()
        for i in range(steps):
            sel = self.sched[i % len(self.sched)]
            list(pool.map(one, [self.convs[c] for c in sel]))
            done += len(sel)
        dt = time.perf_counter() - t0
        if threads:
            pool.shutdown()
        return done / dt, dt

    def close(self):
        self.pool.shutdown()
        for cv in self.convs:
            cv.close()


def cpu_model():
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except Exception:
        pass
    return "unknown"


def run_reference(args):
    rank, _, world = dist_env()
    if rank != 0:
        return 0
    n, fmt, w, h, desc = WORKLOADS[args.workload]
    # the whole job's cameras (weak scaling: n per GPU); the reference runs one worker per camera
    # (rtsp_process_manager.go:106-111), so it can use min(host threads, cameras) cores
    G = max(args.gpus, 1)
    cams = [c for r in range(G) for c in cameras_for_rank(args.workload, r, G)]
    local = schedule_for(args.workload, n)                 # every rank runs the same pattern on its own cameras
    sched = [[r * n + c for r in range(G) for c in pat] for pat in local]
    T = host_threads()
    try:
        arm = CpuArm(cams, T, sched)
        fps, dt = arm.run(args.steps, args.warmup)
    except Exception as e:
        print(json.dumps({"impl": "reference", "unavailable": str(e)[:200]}))
        return 0
    used = arm.threads
    extra = ""
    try:
        few = max(3, min(args.steps, 40))
        f1, _ = arm.run(few, 1, threads=1)
        ff, _ = arm.run(few, 1, fresh_ctx=True)
        extra = f"; 1 thread: {f1:.0f} frames/s; new SwsContext per frame (as PyAV does): {ff:.0f} frames/s"
    except Exception:
        pass
    line = {
        "impl": "reference", "metric": METRIC if (w, h) == (1920, 1080) else "bgr24_frames_per_sec", "value": round(fps, 2), "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(dt / args.steps * 1e3, 4),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": config_for(args.workload, args.steps),
        "cpu_baseline": {"value": round(fps, 2), "unit": UNIT, "cores": used, "kind": "reference",
                         "sample": f"{args.steps} ticks of the workload ({len(cams)} cameras over {G} GPU share(s), {fmt}, frames selected per tick as in the b200 arm): "
                                   f"libswscale {arm.version} sws_scale (SWS_BILINEAR, the call PyAV makes for read_image.py:94) + tobytes (:97), "
                                   f"{used} threads of {T} available{extra}; cpu: {cpu_model()}; host cores only, the reference has no GPU path"},
        "e2e": {"value": round(fps, 2), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    arm.close()
    print(json.dumps(line))
    return 0


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md)"


def ncu_traffic(kernel: str, workload: str):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel, from the committed
    ncu --set full capture of this same command (profiles/ncu_summary.json), or None."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_summary.json")) as f:
            e = json.load(f)[f"{kernel}:{workload}"]
        return e["dram_bytes_read"] + e["dram_bytes_write"]
    except Exception:
        return None


def _pitch(row_bytes: int) -> int:
    """Row pitch of device-resident planes/outputs: tight when that is a multiple of 16 bytes, else padded to 64
    (the layout the library's own staging uses, csrc/vep_abi.cu pitch_of)."""
    return row_bytes if row_bytes % 16 == 0 else (row_bytes + 63) // 64 * 64


def _resident_frame(torch, dev, cam, host_planes, tight: bool):
    """Device copies of one camera's planes + its output buffer; returns (descriptor, keep-alive, output view (H, W, 3))."""
    from video_edge_ai_proxy_b200 import desc_from_device_ptrs
    w, h, fmt = cam["w"], cam["h"], cam["fmt"]
    planes, ptrs, strides = [], [], []
    for p in host_planes:
        pb = np.ascontiguousarray(p).view(np.uint8).reshape(p.shape[0], -1)
        pitch = pb.shape[1] if tight else _pitch(pb.shape[1])
        t = torch.zeros((pb.shape[0], pitch), dtype=torch.uint8, device=dev)
        t[:, :pb.shape[1]] = torch.from_numpy(pb).to(dev)
        planes.append(t); ptrs.append(t.data_ptr()); strides.append(pitch)
    opitch = 3 * w if tight else _pitch(3 * w)
    out = torch.zeros((h, opitch), dtype=torch.uint8, device=dev)
    d = desc_from_device_ptrs(fmt, w, h, ptrs, strides, out.data_ptr(), opitch)
    return d, (planes, out), out[:, :3 * w].unflatten(1, (w, 3)) if opitch != 3 * w else out.view(h, w, 3)


def dma_ceiling(torch, dev, in_sizes, out_sizes, world, target_bytes=12e9):
    """What the box's DMA path can carry for this tick's byte mix: every rank copies, from and to pinned host memory,
    the same H2D and D2H transfers one end-to-end tick makes (one per frame and direction, both directions at once on
    two streams), back to back, with no kernel and no CPU copy in between.  All ranks run it together, so at N GPUs it
    is the ceiling of the whole box (PCIe switches, root complexes, host memory), not N times one link."""
    if not in_sizes:
        return None
    reps = max(4, int(math.ceil(target_bytes / float(sum(in_sizes) + sum(out_sizes)))))
    h_in = [torch.empty(s, dtype=torch.uint8).pin_memory() for s in in_sizes]
    h_out = [torch.empty(s, dtype=torch.uint8).pin_memory() for s in out_sizes]
    d_in = [torch.empty(s, dtype=torch.uint8, device=dev) for s in in_sizes]
    d_out = [torch.zeros(s, dtype=torch.uint8, device=dev) for s in out_sizes]
    for t in h_in:
        t.fill_(7)
    s_in, s_out = torch.cuda.Stream(dev), torch.cuda.Stream(dev)

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
    best = None
    for _ in range(3):                                       # a ceiling is a maximum: best of three runs
        barrier()
        t0 = time.perf_counter()
        run(reps)
        torch.cuda.synchronize(dev)
        dt = time.perf_counter() - t0
        barrier()
        dt = max_over_ranks(dt) if world > 1 else dt
        best = dt if best is None else min(best, dt)
    dt = best
    frames = (sum_over_ranks(len(out_sizes)) if world > 1 else len(out_sizes)) * reps
    nbytes = (sum_over_ranks(sum(in_sizes) + sum(out_sizes)) if world > 1 else sum(in_sizes) + sum(out_sizes)) * reps
    return {"frames_per_s": frames / dt, "gbs": nbytes / dt / 1e9, "reps": reps, "seconds": dt}


def run_gpu(args):
    rank, local_rank, world = dist_env()
    my_cores = pin_rank_cores(local_rank, world)     # before torch starts any thread
    import torch
    import torch.distributed as dist
    from video_edge_ai_proxy_b200 import FramePath, launch_sequence, make_meta, videoframe_framing
    from video_edge_ai_proxy_b200 import _abi
    from video_edge_ai_proxy_b200.synth import make_planes

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the GPU arm has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)

    n, fmt, w, h, desc = WORKLOADS[args.workload]
    tight = args.workload.endswith("-tight")
    cams = cameras_for_rank(args.workload, rank, world)
    sched = schedule_for(args.workload, len(cams))
    n_pat = len(sched)
    alg = [int(ALG_BYTES_BY_FMT.get(c["fmt"], ALG_BYTES_PER_PX) * c["w"] * c["h"]) for c in cams]
    per_set_bytes = sum(alg)
    l2_bytes = torch.cuda.get_device_properties(dev).L2_cache_size
    n_sets = max(args.sets, 2, -(-3 * l2_bytes // per_set_bytes))   # footprint of the rotation >= 3x L2
    while math.gcd(n_sets, n_pat) != 1:
        n_sets += 1                                                   # coprime: every pattern meets every buffer set
    conv_in = lambda first, count: sum(len(sched[i % n_pat]) for i in range(first, first + count))
    bytes_in = lambda first, count: sum(alg[c] for i in range(first, first + count) for c in sched[i % n_pat])

    sampler = ClockSampler(local_rank)
    sampler.start()
    path = FramePath(local_rank)
    host_planes = [make_planes(c["fmt"], c["w"], c["h"], c["seed"]) for c in cams]
    touched = sorted({c for pat in sched for c in pat})

    # ---- parity authority for this run: the live libswscale on this box (outside every timed region) ----
    parity = {"checked_frames": 0, "checked_messages": 0, "against": None}
    want = {}
    if not args.no_parity:
        from oracle import swscale_live as live        # the checker, never the thing measured
        if not live.available():
            raise SystemExit("bench.py: bundled libswscale not found — cannot run the parity gate (use --no-parity to time anyway)")
        parity["against"] = f"libswscale {live.version()} (sws_scale, SWS_BILINEAR: the call PyAV makes at read_image.py:94), x86 SIMD path"
        for c in touched:
            want[c] = live.to_bgr24(cams[c]["fmt"], host_planes[c], cams[c]["w"], cams[c]["h"])

    line = {"metric": METRIC if (w, h) == (1920, 1080) else "bgr24_frames_per_sec", "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": config_for(args.workload, args.steps)}
    windows = []
    peak, peak_src = measured_peak()

    if not args.e2e_only:
        # ---- resident inputs: n_sets copies of every camera's planes, separate outputs per set ----------
        plans, outs = [], []          # plans[s][p]: buffer set s, tick pattern p (None when the pattern converts nothing)
        for s in range(n_sets):
            descs, set_outs = [], []
            for c, hp in zip(cams, host_planes):
                d, keep, view = _resident_frame(torch, dev, c, hp, tight)
                descs.append(d)
                set_outs.append((keep, view))
            plans.append([path.plan([descs[c] for c in pat]) if pat else None for pat in sched])
            outs.append(set_outs)
        stream = torch.cuda.current_stream(dev)
        sptr = stream.cuda_stream

        def tick_plans(first, count):
            return [pl for pl in (plans[i % n_sets][i % n_pat] for i in range(first, first + count)) if pl is not None]

        warm = max(args.warmup, 3, n_sets * n_pat)        # every (set, pattern) pair is converted at least once
        launch_sequence(tick_plans(0, warm), sptr, _abi.VEP_ORDER_STRICT)
        torch.cuda.synchronize(dev)
        # parity gate: set 0 against the live library, every other set against set 0 (same inputs), byte for byte
        for c in touched:
            got0 = outs[0][c][1]
            if want:
                if not np.array_equal(got0.cpu().numpy(), want[c]):
                    raise SystemExit(f"bench.py: camera {c} differs from the live libswscale — refusing to time a wrong kernel")
                parity["checked_frames"] += 1
            for s in range(1, n_sets):
                if not torch.equal(outs[s][c][1], got0) or not bool(got0.any()):
                    raise SystemExit(f"bench.py: buffer set {s} camera {c} disagrees with set 0 — refusing to time a broken kernel")
        parity["buffer_sets_equal_to_checked_set"] = n_sets - 1

        def timed(order):
            seq = tick_plans(warm, args.steps)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            # W untimed warm-up steps directly in front of the timed region (the parity checks above left the GPU idle)
            launch_sequence(tick_plans(warm - max(args.warmup, 3), max(args.warmup, 3)), sptr, order)
            e0.record(stream); e1.record(stream)        # creates the handles; both are re-recorded below
            barrier()
            torch.cuda.synchronize(dev)
            t0 = time.perf_counter()
            # one native call: record e0, K launches, record e1 — on the launching stream, nothing else in between
            launch_sequence(seq, sptr, order, e0.cuda_event, e1.cuda_event)
            torch.cuda.synchronize(dev)
            t1 = time.perf_counter()
            barrier()
            ms = e0.elapsed_time(e1)
            return (max_over_ranks(ms) if world > 1 else ms), len(seq), (t0, t1)

        # ---- value: K steps, device-timed on the launching stream; consecutive ticks use disjoint buffer sets ----
        head_order = _abi.VEP_ORDER_STRICT if args.strict_order else _abi.VEP_ORDER_NONE
        dev_ms, gpu_launches, win = timed(head_order)
        windows.append(win)
        converted = conv_in(warm, args.steps)
        total_frames = sum_over_ranks(converted) if world > 1 else converted          # whole job, K steps
        offered_frames = (sum_over_ranks(len(cams)) if world > 1 else len(cams)) * args.steps
        value = total_frames / (dev_ms / 1e3)
        timed_bytes = bytes_in(warm, args.steps)
        achieved = timed_bytes / (dev_ms / 1e3) / 1e9      # this rank's GPU (all ranks run the same schedule)
        # the same K ticks in full stream order (what vep_plan_launch / vep_convert_batch do by default) and with only the
        # stores ordered after the previous tick (VEP_ORDER_WRITES); reported beside the headline, not instead of it
        strict_ms, _, _ = timed(_abi.VEP_ORDER_STRICT)
        ow_ms, _, _ = timed(_abi.VEP_ORDER_WRITES)
        strict_gbs, ow_gbs = timed_bytes / (strict_ms / 1e3) / 1e9, timed_bytes / (ow_ms / 1e3) / 1e9
        written = sum(3 * cams[c]["w"] * cams[c]["h"] for i in range(warm, warm + args.steps) for c in sched[i % n_pat])

        for row in plans:
            for p in row:
                if p is not None:
                    p.close()
        del outs, plans
        torch.cuda.empty_cache()
        kname = "vep::yuv_to_bgr24_tick"
        line.update({
            "value": round(value, 1), "ms_per_step": round(dev_ms / args.steps, 5), "gpu_launches": gpu_launches,
            "roofline": {"bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s", "frac": round(achieved / peak, 4),
                         "traffic": ncu_traffic("tick", args.workload),
                         "traffic_note": "ncu, isolated launch: reads = algorithmic; part of the written bytes is still dirty in the L2 at kernel end and is written back during the next launch",
                         "peak_source": peak_src, "frac_of_nominal_8TBs": round(achieved / 8000.0, 4),
                         "alg_bytes_per_launch": int(timed_bytes / max(gpu_launches, 1)), "buffer_sets": n_sets,
                         "launch": ("strict stream order" if args.strict_order else "VEP_ORDER_NONE: consecutive ticks touch disjoint buffer sets") + "; K launches issued by one vep_plan_launch_sequence call",
                         "unflushed_write_fraction_upper_bound": round(min(1.0, l2_bytes / max(written, 1)), 4),
                         "offered_frames_per_s": round(offered_frames / (dev_ms / 1e3), 1),
                         "strict_stream_order": {"achieved": round(strict_gbs, 1), "frac": round(strict_gbs / peak, 4),
                                                 "frames_per_s": round(total_frames / (strict_ms / 1e3), 1)},
                         "ordered_writes": {"achieved": round(ow_gbs, 1), "frac": round(ow_gbs / peak, 4),
                                            "frames_per_s": round(total_frames / (ow_ms / 1e3), 1)},
                         "kernel": kname}})
    else:
        converted = conv_in(0, args.steps)

    # ---- e2e: host planes -> submit -> tick (H2D, kernel, D2H) -> packed VideoFrame in pinned memory ------------
    if not args.skip_e2e:
        e2e_steps = max(args.steps if args.steps <= args.e2e_steps else args.e2e_steps, min(300, args.e2e_steps))   # at least 300 ticks per leg: shorter runs scatter by +-10 %
        names = [f"cam{c['camera']}" for c in cams]
        # the rank's cameras are opened together: shared slot memory, one copy per direction for cameras in lock step
        path.open_streams([(name, c["fmt"], c["w"], c["h"]) for name, c in zip(names, cams)], memory_buffer=1)
        meta = make_meta(timestamp=1604660000123, pts=900000, dts=897000, time_base=1 / 90000, is_keyframe=True, frame_type="I", keyframe=1)
        share = len(my_cores) if my_cores else host_threads() // max(world, 1)
        # The "decoder's buffer pool": one arena per rank that holds every camera's decoded planes, page-locked ONCE at
        # set-up (vep_host_register, what a get_buffer callback over a registered arena gives a real decoder).  Frames
        # in it are submitted without a CPU copy (VEP_SUBMIT_PINNED); the pageable originals serve the secondary leg.
        # Rows have the linesize a decoder gives them (libavcodec pads linesizes to 64 bytes).
        ls_of = lambda p: (p.shape[1] * p.itemsize + 63) // 64 * 64
        need = sum(((ls_of(p) * p.shape[0] + 255) // 256) * 256 for hp in host_planes for p in hp)
        if args.pool == "register":
            arena = np.empty(need, np.uint8)
            path.register_host(arena)                     # cudaHostRegister over memory the producer allocated itself
        else:
            arena_t = torch.empty(need, dtype=torch.uint8).pin_memory()       # cudaHostAlloc'ed by the producer
            arena = arena_t.numpy()
        pinned_planes, off = [], 0
        for hp in host_planes:
            views = []
            for p in hp:
                ls, rb = ls_of(p), p.shape[1] * p.itemsize
                rows = arena[off:off + ls * p.shape[0]].reshape(p.shape[0], ls)
                rows[:, :rb] = p.view(np.uint8).reshape(p.shape[0], rb)
                views.append(rows[:, :rb].view(p.dtype))
                off += ((ls * p.shape[0] + 255) // 256) * 256
            pinned_planes.append(views)
        per_camera_threads = args.e2e_submit == "threads"
        if per_camera_threads:
            # pageable leg, variant: one producer thread per camera like the reference (each calls vep_submit), never more
            # than this rank's share of the host minus the ticker; 4K frames: ONE frame's copy is split over helper threads
            n_prod = max(1, min(len(cams), share - 1 if share > 1 else 1))
            big = max(int(1.5 * c["w"] * c["h"]) for c in cams) >= (4 << 20)
            helpers = max(0, min(3, share - 2)) if big else 0
        else:
            # pageable leg: one producer thread hands the whole tick to the library (vep_submit_batch); the library's
            # helper threads share the copies.  Measured on the 16-core box at N = 1: 1-3 helpers 6.87 k frames/s,
            # 5-7 helpers 6.7 k, none 5.8 k
            n_prod = 1
            helpers = max(0, min(3, share - 2)) if args.staging_threads < 0 else args.staging_threads
        pool = ThreadPoolExecutor(n_prod)
        path.set_staging_threads(helpers)
        if args.run_limit >= 0:
            path.set_copy_run_limit(args.run_limit)

        class _Done:                      # a finished "future" for work done inline
            def result(self):
                return None

        def stage_async(tick, pinned):
            """Stage tick `tick`'s frames; returns (selected cameras, futures to join before tick_begin).
            pinned: "slot" = the decoder wrote the frame into the library's pinned slot (vep_acquire_input/vep_commit_input),
            True = the frame lies in the producer's own page-locked pool (VEP_SUBMIT_PINNED), False = pageable (VEP_SUBMIT_COPY)."""
            sel = sorted(sched[tick % n_pat])
            chosen = set(sel)
            if pinned == "slot":
                for c in sel:
                    path.acquire_raw(names[c]); path.commit_input(names[c], meta)
                if len(sel) < len(names):         # gate-skipped frames of the other cameras: counted, one call
                    path.submit_many([(names[c], None, None) for c in range(len(names)) if c not in chosen])
                return sel, [_Done()]
            if pinned:                    # no copy: one cheap native call, made inline by the ticker thread
                path.submit_many([(names[c], pinned_planes[c], meta) if c in chosen else (names[c], None, None) for c in range(len(names))], pinned=True)
                return sel, [_Done()]
            if not per_camera_threads:
                items = [(names[c], host_planes[c], meta) if c in chosen else (names[c], None, None) for c in range(len(names))]
                return sel, [pool.submit(path.submit_many, items)]
            for c in range(len(names)):
                if c not in chosen:
                    path.submit(names[c], None, None, convert=False)          # gate-skipped: counted, nothing copied
            return sel, [pool.submit(path.submit, names[c], host_planes[c], meta) for c in sel]

        def join(futs):
            for f in futs:
                f.result()

        def consume(sel):
            tot = 0
            for c in sel:
                tot += len(path.result_packed(names[c]))      # the message a producer hands to XADD
            return tot

        def warm_and_check(pinned, ticks, lat=None):
            """Synchronous ticks: warm-up, latency samples, and the end-to-end parity gate (first occurrence of every pattern)."""
            checked = set()
            for t in range(ticks):
                sel, futs = stage_async(t, pinned); join(futs)
                l0 = time.perf_counter(); path.tick(); consume(sel)
                if lat is not None and t >= 3 and sel:
                    lat.append((time.perf_counter() - l0) * 1e3)
                if want and t < n_pat:
                    for c in sel:
                        if c in checked:
                            continue
                        pre, suf = videoframe_framing(meta, cams[c]["w"], cams[c]["h"])
                        if bytes(path.result_packed(names[c])) != pre + want[c].tobytes() + suf:
                            raise SystemExit(f"bench.py: end-to-end message of camera {c} (source mode {pinned!r}) differs from reference framing + live libswscale pixels")
                        checked.add(c)
                        parity["checked_messages"] += 1

        def timed_e2e(pinned):
            barrier()
            torch.cuda.synchronize(dev)
            s0 = path.stats()
            x0 = time.perf_counter()
            depth, pending, msg_bytes = 3, deque(), 0         # ticks in flight: H2D of later ticks overlaps D2H of earlier ones
            sel, futs = stage_async(0, pinned)
            for i in range(e2e_steps):
                join(futs)                                    # tick i fully staged
                nconv, ticket = path.tick_begin_ticket()
                if nconv:
                    pending.append((ticket, sel))
                if i + 1 < e2e_steps:
                    sel, futs = stage_async(i + 1, pinned)    # pageable source: the copy runs while this thread waits below
                while len(pending) >= depth:
                    tk, s_ = pending.popleft()
                    path.tick_wait_ticket(tk); msg_bytes = consume(s_) or msg_bytes
            while pending:
                tk, s_ = pending.popleft()
                path.tick_wait_ticket(tk); msg_bytes = consume(s_) or msg_bytes
            torch.cuda.synchronize(dev)
            x1 = time.perf_counter()
            barrier()
            s1 = path.stats()
            secs = max_over_ranks(x1 - x0) if world > 1 else (x1 - x0)
            return secs, (x0, x1), msg_bytes, (s1["h2d_bytes"] - s0["h2d_bytes"]) // e2e_steps, (s1["d2h_bytes"] - s0["d2h_bytes"]) // e2e_steps

        # the decoder's writes into the library's pinned slots: every slot of every camera receives its camera's frame once
        # (the timed region then re-submits them; decoding itself is outside this path in both arms)
        for _ in range(8 + 2):                              # ring 1 + 4 in flight + 2 spare slots, and then some
            for c in range(len(names)):
                for v, p in zip(path.acquire_input(names[c]), host_planes[c]):
                    v[...] = p
                path.commit_input(names[c], meta)
            path.tick()
        lat = []
        warm_and_check("slot", 3 + 30, lat)                # warm-up, then 30 synchronous ticks for the latency figure
        lat.sort()
        # byte mix of the fullest tick for the ceiling probe: per frame what the library copies in and out
        fullest = max(sched, key=len)
        in_sizes = [path.staged_bytes(names[c]) for c in fullest]
        out_sizes = [3 * cams[c]["w"] * cams[c]["h"] for c in fullest]
        ceiling = None
        if not args.no_ceiling:
            # both copy granularities the pipeline can end up with; the ceiling is the better one (on one GPU fewer, larger
            # copies carry ~12 % more; with several GPUs the box's limit is elsewhere and they are equal)
            per_frame = dma_ceiling(torch, dev, in_sizes, out_sizes, world)
            merged = dma_ceiling(torch, dev, [sum(in_sizes)], [sum(out_sizes)], world)
            merged["frames_per_s"] *= len(out_sizes)          # it counted one "frame" per tick
            ceiling = max(per_frame, merged, key=lambda r: r["frames_per_s"])
            ceiling["per_frame_copies_frames_per_s"] = per_frame["frames_per_s"]
            ceiling["one_copy_per_direction_frames_per_s"] = merged["frames_per_s"]

        e2e_s, win, msg_bytes, h2d_per_step, d2h_per_step = timed_e2e("slot")
        windows.append(win)
        warm_and_check(False, max(3, min(n_pat, 30)))
        pg_s, _, _, pg_h2d, _ = timed_e2e(False)
        warm_and_check(True, max(3, min(n_pat, 30)))
        pp_s, _, _, pp_h2d, _ = timed_e2e(True)
        e2e_conv = conv_in(0, e2e_steps)
        e2e_total = sum_over_ranks(e2e_conv) if world > 1 else e2e_conv
        e2e_value, pg_value, pp_value = e2e_total / e2e_s, e2e_total / pg_s, e2e_total / pp_s
        pcie_job = (sum_over_ranks(h2d_per_step + d2h_per_step) if world > 1 else (h2d_per_step + d2h_per_step)) * e2e_steps / e2e_s / 1e9
        pool.shutdown()
        path.set_staging_threads(0)
        line["e2e"] = {"value": round(e2e_value, 1), "unit": UNIT, "h2d_bytes_per_step": int(h2d_per_step), "d2h_bytes_per_step": int(d2h_per_step),
                       "steps": e2e_steps,
                       "api": "decoded planes written by the decoder into the library's pinned staging slots (vep_acquire_input hands the plane pointers to its get_buffer callback, "
                              "vep_commit_input submits the frame: no CPU copy) -> tick_begin_ticket/tick_wait_ticket (H2D, one kernel launch, D2H; 3 ticks in flight) -> result_packed",
                       "pcie_gbs": round((h2d_per_step + d2h_per_step) * e2e_steps / e2e_s / 1e9, 2), "pcie_gbs_whole_job": round(pcie_job, 2),
                       "message_bytes": msg_bytes,
                       "host_cores_of_this_rank": len(my_cores) if my_cores else host_threads(),
                       "core_pinning": "each rank pinned to its own disjoint share of the host cores" if my_cores else "none (single rank)",
                       "sync_tick_latency_ms": {"p50": round(lat[len(lat) // 2], 3), "max": round(lat[-1], 3),
                                                "what": "one tick alone, planes in pinned memory -> packed messages readable (H2D + kernel + D2H, no overlap)"} if lat else None,
                       "pageable_source": {"value": round(pg_value, 1), "unit": UNIT, "h2d_bytes_per_step": int(pg_h2d),
                                           "api": ("FramePath.submit on one producer thread per camera" if per_camera_threads else "FramePath.submit_many: one vep_submit_batch call per tick on a producer thread, copies shared with the library's helper threads")
                                                  + " [VEP_SUBMIT_COPY: pageable planes (what PyAV hands out) -> staging copy into pinned slots] -> same tick pipeline",
                                           "producer_threads": n_prod, "staging_helper_threads": helpers},
                       "caller_pinned_pool": {"value": round(pp_value, 1), "unit": UNIT, "h2d_bytes_per_step": int(pp_h2d),
                                              "api": "planes in the producer's OWN page-locked frame pool (decoder linesizes, 64-byte aligned) -> FramePath.submit_many(pinned=True) "
                                                     "[vep_submit_batch, VEP_SUBMIT_PINNED: no CPU copy, one H2D copy per plane] -> same tick pipeline"}}
        if ceiling:
            line["e2e"].update({"box_dma_ceiling_gbs": round(ceiling["gbs"], 2), "box_dma_ceiling_frames_per_s": round(ceiling["frames_per_s"], 1),
                                "frac_of_ceiling": round(e2e_value / ceiling["frames_per_s"], 4),
                                "ceiling_probe": f"all {world} rank(s) at once: the tick's H2D and D2H bytes from/to pinned memory on two streams, ticks back to back (best of 3 runs of {ceiling['seconds']:.2f} s), no kernel, no CPU copy; "
                                                 f"measured with one copy per frame and direction ({ceiling['per_frame_copies_frames_per_s']:.0f} frames/s) and with one copy per direction and tick ({ceiling['one_copy_per_direction_frames_per_s']:.0f} frames/s); the ceiling is the larger"})
            line["e2e"]["pageable_source"]["frac_of_ceiling"] = round(pg_value / ceiling["frames_per_s"], 4)
            line["e2e"]["caller_pinned_pool"]["frac_of_ceiling"] = round(pp_value / ceiling["frames_per_s"], 4)
        if args.pool == "register":
            path.unregister_host(arena)
        line["e2e"]["caller_pinned_pool"]["frame_pool"] = "cudaHostRegister (vep_host_register) over the producer's own allocation" if args.pool == "register" else "cudaHostAlloc'ed by the producer"
    sampler.stop()
    line["clocks"] = sampler.summary(windows)
    line["parity"] = parity
    path.close()

    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        T = host_threads()
        try:
            # bounded sample: time three ticks, then size the sample to ~15 s of wall time on all host threads
            arm = CpuArm(cams, T, sched if n_pat > 1 else None)
            _, t3 = arm.run(max(3, n_pat if n_pat <= 30 else 3), 1)
            per_tick = t3 / max(3, n_pat if n_pat <= 30 else 3)
            cpu_steps = max(3, min(int(15.0 / max(per_tick, 1e-4)), 20000))
            fps, dt = arm.run(cpu_steps, 2)
            fps_scale, _ = arm.run(max(3, cpu_steps // 4), 1, with_tobytes=False)
            fps_one, _ = arm.run(max(3, cpu_steps // 16), 1, threads=1)
            line["cpu_baseline"] = {"value": round(fps, 1), "unit": UNIT, "cores": arm.threads, "kind": "reference",
                                    "sample": f"{cpu_steps} ticks of the workload ({n} cameras, {fmt}, frames selected per tick as in the b200 arm) ({dt:.1f} s wall; sized from three timed ticks): "
                                              f"libswscale {arm.version} sws_scale (what PyAV calls for read_image.py:94) + tobytes (:97) on {arm.threads} threads of {T}; "
                                              f"sws_scale alone: {fps_scale:.0f} frames/s; 1 thread: {fps_one:.0f} frames/s; cpu: {cpu_model()}"}
            arm.close()
        except Exception as e:
            line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": 0, "kind": "reference", "sample": f"unavailable: {e}"}
    if rank == 0:
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5000)
    ap.add_argument("--warmup", type=int, default=50)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="16x1080p", choices=sorted(WORKLOADS))
    ap.add_argument("--sets", type=int, default=9, help="minimum number of rotating buffer sets")
    ap.add_argument("--e2e-steps", type=int, default=600, help="upper bound of end-to-end ticks per leg (at least 300 are run)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="skip the libswscale parity gate (diagnostic runs only)")
    ap.add_argument("--no-ceiling", action="store_true", help="skip the DMA ceiling probe")
    ap.add_argument("--e2e-only", action="store_true", help="diagnostic: skip the device-resident legs")
    ap.add_argument("--skip-e2e", action="store_true", help="diagnostic: skip the end-to-end legs")
    ap.add_argument("--run-limit", type=int, default=-1, help="diagnostic: most frames per merged H2D/D2H copy (library default: no limit)")
    ap.add_argument("--staging-threads", type=int, default=-1, help="library helper threads for the staging copy (default: by this rank's share of the cores)")
    ap.add_argument("--pool", default="register", choices=["register", "alloc"], help="how the producer's frame pool of the e2e leg is page-locked")
    ap.add_argument("--e2e-submit", default="batch", choices=["batch", "threads"],
                    help="how host frames reach the library: one vep_submit_batch call per tick (default) or one Python producer thread per camera calling vep_submit")
    ap.add_argument("--strict-order", action="store_true", help="launch every tick fully ordered after the previous one (no overlap)")
    args = ap.parse_args()
    if args.impl == "reference":
        # the CPU arm times one tick per step as well; cap so the run ends within minutes
        return run_reference(args)
    return run_gpu(args)


if __name__ == "__main__":
    sys.exit(main())
