#!/bin/bash
mkdir -p gpurun_out
for k in 0 8 4 2 1; do
python bench.py --gpus 1 --steps 20 --warmup 5 --e2e-only --no-cpu-baseline --no-parity --run-limit $k > gpurun_out/r2rl_n1_k$k.json 2> gpurun_out/r2rl_n1_k$k.err; echo "k=$k rc=$?"
done
