#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2l_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2l_pytest.log
tail -3 gpurun_out/r2l_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
python bench.py --gpus 1 --steps 300 --warmup 5 --workload 16x854x480-tight > gpurun_out/r2l_bench_tight.json 2> gpurun_out/r2l_bench_tight.err; echo "tight rc=$?"
python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2l_bench_20.json 2> gpurun_out/r2l_bench_20.err; echo "bench20 rc=$?"
python bench.py --gpus 1 --steps 20 --warmup 5 --impl reference > gpurun_out/r2l_ref_20.json 2>/dev/null; echo "ref rc=$?"
