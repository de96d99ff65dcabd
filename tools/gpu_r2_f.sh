#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_pipeline.py -m gpu -x -q > gpurun_out/r2f_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2f_pytest.log
tail -3 gpurun_out/r2f_pytest.log
for pool in register alloc; do
python bench.py --gpus 1 --steps 20 --warmup 5 --e2e-only --no-cpu-baseline --pool $pool > gpurun_out/r2f_e2e_$pool.json 2> gpurun_out/r2f_e2e_$pool.err; echo "$pool rc=$?"
python bench.py --gpus 1 --steps 20 --warmup 5 --e2e-only --no-cpu-baseline --pool $pool --workload 16x854x480 > gpurun_out/r2f_e2e_854_$pool.json 2> gpurun_out/r2f_e2e_854_$pool.err; echo "$pool 854 rc=$?"
done
