"""CPU: camera->GPU partition and the N>1 bench plumbing over gloo (world_size 2)."""
import os
import subprocess
import sys

from video_edge_ai_proxy_b200 import assign_cameras

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_block_assignment_64_cameras_8_gpus():
    a = assign_cameras([(1920, 1080)] * 64, 8)
    assert [len(r) for r in a] == [8] * 8
    assert a[3] == list(range(24, 32))          # camera c -> GPU c // 8


def test_balanced_assignment_mixed_fleet():
    sizes = [(640, 480)] * 43 + [(1280, 720)] * 43 + [(1920, 1080)] * 42
    a = assign_cameras(sizes, 8)
    assert sorted(c for r in a for c in r) == list(range(128))
    load = [sum(sizes[c][0] * sizes[c][1] for c in r) for r in a]
    assert max(load) - min(load) <= 1920 * 1080
    assert assign_cameras(sizes, 8) == a        # deterministic


def test_uneven_and_degenerate():
    assert assign_cameras([(2, 2)] * 3, 2) == [[0, 1], [2]]
    assert assign_cameras([], 4) == [[], [], [], []]
    assert assign_cameras([(2, 2)], 1) == [[0]]


_WORKER = r'''
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
import bench
dist.init_process_group("gloo", rank=int(os.environ["RANK"]), world_size=int(os.environ["WORLD_SIZE"]))
rank, world = dist.get_rank(), dist.get_world_size()
cams = bench.cameras_for_rank("16x1080p", rank, world)
# every rank owns a disjoint, equal-sized camera set; no data-path collective
gathered = [None] * world
dist.all_gather_object(gathered, cams)
ids = [c["camera"] for g in gathered for c in g]
assert sorted(ids) == list(range(16 * world)), ids
# timing reduction used by bench.py: max over ranks
t = bench.max_over_ranks(1.0 + rank, device="cpu")
assert t == float(world), t
tot = bench.sum_over_ranks(len(cams), device="cpu")
assert tot == 16 * world
if rank == 0:
    print("GLOO_OK")
dist.destroy_process_group()
'''


def test_world_size_2_gloo_sharding(tmp_path):
    script = tmp_path / "w.py"
    script.write_text(_WORKER)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                          "--master-addr", "127.0.0.1", "--master-port", "29611", str(script), ROOT],
                         capture_output=True, text=True, env=env, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "GLOO_OK" in out.stdout


def test_bench_schedules_match_the_baseline_configs():
    import numpy as np
    sys.path.insert(0, ROOT)
    import bench
    # key-frame-only: GOP 30, every camera converts exactly once per 30 ticks, phases staggered
    s = bench.schedule_for("8x4k-nv12-keyframe", 8)
    assert len(s) == 30 and sorted(c for t in s for c in t) == list(range(8)) and max(len(t) for t in s) == 1
    # interval skip: seeded Bernoulli(0.25) per camera per tick, seed = tick index
    s = bench.schedule_for("mixed128-skip", 128)
    assert len(s) == 16
    for t, sel in enumerate(s):
        assert sel == [int(c) for c in np.nonzero(np.random.default_rng(t).random(128) < 0.25)[0]]
    assert 0.2 < sum(len(t) for t in s) / (16 * 128) < 0.3
    cams = bench.cameras_for_rank("mixed128-skip", 0, 1)
    assert [(c["w"], c["h"]) for c in cams].count((1920, 1080)) == 42 and len(cams) == 128
    # default: every camera every tick; ranks own disjoint cameras
    assert bench.schedule_for("16x1080p", 16) == [list(range(16))]
    a, b = bench.cameras_for_rank("16x1080p", 0, 2), bench.cameras_for_rank("16x1080p", 1, 2)
    assert {c["camera"] for c in a}.isdisjoint({c["camera"] for c in b}) and len(a) == len(b) == 16


def test_both_bench_arms_print_the_same_config_object_and_ranks_get_disjoint_cores():
    import bench
    for wl in ("16x1080p", "8x4k-nv12-keyframe", "mixed128-skip"):
        a, b = bench.config_for(wl, 20), bench.config_for(wl, 20)
        assert a == b and set(a) == {"workload", "cameras_per_gpu", "format", "converted_frames_per_step_per_gpu", "parallelism", "l2"}
    assert bench.config_for("8x4k-nv12-keyframe", 30)["converted_frames_per_step_per_gpu"] == round(8 / 30, 3)
    # core pinning: a child process per "rank" so this process keeps its own mask
    code = ("import os,sys,json; sys.path.insert(0,%r); import bench; "
            "print(json.dumps(bench.pin_rank_cores(int(sys.argv[1]), int(sys.argv[2]))))" % ROOT)
    import json
    n = len(os.sched_getaffinity(0))
    world = 2 if n >= 2 else 1
    sets = [json.loads(subprocess.run([sys.executable, "-c", code, str(r), str(world)], capture_output=True, text=True).stdout) for r in range(world)]
    if world == 2:
        assert sets[0] and sets[1] and set(sets[0]).isdisjoint(sets[1]) and len(sets[0]) == len(sets[1]) == n // 2
    assert bench.pin_rank_cores(0, 1) is None          # a single rank is left alone
