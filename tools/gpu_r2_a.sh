#!/bin/bash
# round 2, first GPU call: tests, driver-shaped bench, new workloads, launch list
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2a_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2a_pytest.log
tail -5 gpurun_out/r2a_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2a_smoke.log 2>&1; tail -2 gpurun_out/r2a_smoke.log
python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2a_bench_20.json 2> gpurun_out/r2a_bench_20.err; echo "bench20 rc=$?"
python bench.py --gpus 1 --steps 20 --warmup 5 --impl reference > gpurun_out/r2a_ref_20.json 2>/dev/null
python bench.py --gpus 1 --steps 2000 --warmup 20 --no-cpu-baseline > gpurun_out/r2a_bench_2000.json 2> gpurun_out/r2a_bench_2000.err; echo "bench2000 rc=$?"
for wl in 16x854x480 16x1366x768 16x854x480-tight 8x4k-nv12-keyframe; do
  python bench.py --gpus 1 --steps 600 --warmup 10 --workload $wl > gpurun_out/r2a_bench_$wl.json 2> gpurun_out/r2a_bench_$wl.err; echo "$wl rc=$?"
done
python tools/pcie_probe.py > gpurun_out/r2a_pcie_probe_n1.log 2>&1; echo "probe rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r2a_launches.csv python bench.py --steps 20 --warmup 5 --no-cpu-baseline --skip-e2e --no-parity > gpurun_out/r2a_ncu.log 2>&1; echo "ncu rc=$?"
