#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2c_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2c_pytest.log
tail -3 gpurun_out/r2c_pytest.log
for i in 1 2; do
python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2c_bench_20_$i.json 2> gpurun_out/r2c_bench_20_$i.err; echo "bench20 rc=$?"
done
for k in 0 1 3 5; do
python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu-baseline --e2e-only --staging-threads $k > gpurun_out/r2c_e2e_st$k.json 2> gpurun_out/r2c_e2e_st$k.err; echo "st$k rc=$?"
done
