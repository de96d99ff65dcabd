#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2b_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2b_pytest.log
tail -3 gpurun_out/r2b_pytest.log
for i in 1 2 3; do
python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2b_bench_20_$i.json 2> gpurun_out/r2b_bench_20_$i.err; echo "bench20 rc=$?"
done
python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu-baseline --e2e-submit threads > gpurun_out/r2b_bench_20_threads.json 2> gpurun_out/r2b_bench_20_threads.err; echo "bench20t rc=$?"
python bench.py --gpus 1 --steps 300 --warmup 5 --no-cpu-baseline --workload 16x854x480 > gpurun_out/r2b_bench_854.json 2> gpurun_out/r2b_bench_854.err; echo "854 rc=$?"
python bench.py --gpus 1 --steps 300 --warmup 5 --no-cpu-baseline --workload mixed128-skip > gpurun_out/r2b_bench_mixed.json 2> gpurun_out/r2b_bench_mixed.err; echo "mixed rc=$?"
python bench.py --gpus 1 --steps 300 --warmup 5 --no-cpu-baseline --workload 8x4k-nv12-keyframe > gpurun_out/r2b_bench_4kkf.json 2> gpurun_out/r2b_bench_4kkf.err; echo "4kkf rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/r2b_launches.csv python bench.py --steps 20 --warmup 5 --no-cpu-baseline --skip-e2e --no-parity > gpurun_out/r2b_ncu.log 2>&1; echo "ncu rc=$?"
