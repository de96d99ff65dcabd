#!/bin/bash
# round 2, one GPU: tests, smoke, the two arms as the driver runs them, every workload, launch list, soak, tick kernel time
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2f1_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2f1_pytest.log
tail -3 gpurun_out/r2f1_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2f1_smoke.log 2>&1; tail -1 gpurun_out/r2f1_smoke.log
python bench.py --gpus 1 --steps 20 --warmup 5 --impl reference > gpurun_out/r2f1_ref_16x1080p.json 2>/dev/null; echo "ref rc=$?"
python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2f1_bench_16x1080p.json 2> gpurun_out/r2f1_bench_16x1080p.err; echo "bench20 rc=$?"
python bench.py --gpus 1 --steps 5000 --warmup 50 --no-cpu-baseline > gpurun_out/r2f1_bench_16x1080p_5000steps.json 2> gpurun_out/r2f1_bench_16x1080p_5000steps.err; echo "bench5000 rc=$?"
for wl in 1x480p 8x1080p 8x4k-nv12 8x4k-nv12-keyframe mixed128-skip 16x854x480 16x1366x768 16x854x480-tight 16x1080p-yuvj422p 16x1080p-yuv420p10le 16x1080p-yuv444p; do
python bench.py --gpus 1 --steps 300 --warmup 5 --workload $wl > gpurun_out/r2f1_bench_$wl.json 2> gpurun_out/r2f1_bench_$wl.err; echo "$wl rc=$?"
done
python bench.py --gpus 1 --steps 300 --warmup 5 --workload 1x480p --impl reference > gpurun_out/r2f1_ref_1x480p.json 2>/dev/null
python tools/tick_kernel_ms.py > gpurun_out/r2f1_tick_kernel_ms.log 2>&1; cat gpurun_out/r2f1_tick_kernel_ms.log | tail -4
python tools/soak.py 40 > gpurun_out/r2f1_soak.log 2>&1; tail -1 gpurun_out/r2f1_soak.log
