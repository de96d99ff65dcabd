#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r2un_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2un_pytest.log
tail -4 gpurun_out/r2un_pytest.log
python bench.py --gpus 1 --steps 300 --warmup 5 --no-cpu-baseline --skip-e2e --workload 16x854x480-tight > gpurun_out/r2un_bench_tight.json 2> gpurun_out/r2un_bench_tight.err; echo "tight rc=$?"
python bench.py --gpus 1 --steps 300 --warmup 5 --no-cpu-baseline --skip-e2e --workload mixed128-skip > gpurun_out/r2un_bench_mixed.json 2> gpurun_out/r2un_bench_mixed.err; echo "mixed rc=$?"
python - <<'PY'
import json
for f in ("r2un_bench_tight","r2un_bench_mixed"):
    try:
        d=json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1]); r=d["roofline"]
        print(f, d["value"], r["frac"], r["ordered_writes"]["frac"], r["strict_stream_order"]["frac"], d["parity"])
    except Exception as e: print(f, "ERR", open(f"gpurun_out/{f}.err").read()[-600:])
PY
