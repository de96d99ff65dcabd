#!/bin/bash
mkdir -p gpurun_out
for i in 1 2; do
for pool in register alloc; do
python bench.py --gpus 1 --steps 20 --warmup 5 --e2e-only --no-cpu-baseline --no-parity --pool $pool > gpurun_out/r2g_e2e_${pool}_$i.json 2> gpurun_out/r2g_e2e_${pool}_$i.err; echo "$pool rc=$?"
done
done
python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu-baseline --pool alloc > gpurun_out/r2g_full_alloc.json 2> gpurun_out/r2g_full_alloc.err; echo "full rc=$?"
