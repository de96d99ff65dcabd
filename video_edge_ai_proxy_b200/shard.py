"""Camera -> GPU assignment.  Cameras are independent (the reference runs one container per camera,
server/services/rtsp_process_manager.go:106-111), so the multi-GPU layout is a static partition
with no collective on the data path: every rank converts its own cameras on its own GPU.
"""
from __future__ import annotations

from typing import List, Sequence, Tuple


def assign_cameras(sizes: Sequence[Tuple[int, int]], world_size: int, mode: str = "auto") -> List[List[int]]:
    """Return, per rank, the list of camera indices it owns.

    sizes: (width, height) per camera.  mode "block": contiguous blocks (camera c -> rank
    c // ceil(n/world)), the layout of the 64-camera config; "balanced": greedy longest-processing-
    time on W*H so a mixed-resolution fleet gives every GPU about the same number of pixels;
    "auto": block when all cameras have one size, else balanced.  Deterministic."""
    n = len(sizes)
    if world_size < 1:
        raise ValueError("world_size must be >= 1")
    if mode == "auto":
        mode = "block" if len(set(sizes)) <= 1 else "balanced"
    out: List[List[int]] = [[] for _ in range(world_size)]
    if mode == "block":
        per = -(-n // world_size) if n else 0
        for c in range(n):
            out[c // per].append(c)
        return out
    if mode != "balanced":
        raise ValueError(mode)
    load = [0] * world_size
    order = sorted(range(n), key=lambda c: (-sizes[c][0] * sizes[c][1], c))
    for c in order:
        r = min(range(world_size), key=lambda k: (load[k], k))
        out[r].append(c)
        load[r] += sizes[c][0] * sizes[c][1]
    for r in out:
        r.sort()
    return out
