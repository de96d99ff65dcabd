#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2k_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2k_pytest.log
tail -3 gpurun_out/r2k_pytest.log
python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2k_bench_20.json 2> gpurun_out/r2k_bench_20.err; echo "bench20 rc=$?"
python bench.py --gpus 1 --steps 20 --warmup 5 --impl reference > gpurun_out/r2k_ref_20.json 2> /dev/null; echo "ref20 rc=$?"
for wl in 8x4k-nv12-keyframe mixed128-skip 16x854x480; do
python bench.py --gpus 1 --steps 300 --warmup 5 --workload $wl --no-cpu-baseline > gpurun_out/r2k_bench_$wl.json 2> gpurun_out/r2k_bench_$wl.err; echo "$wl rc=$?"
done
