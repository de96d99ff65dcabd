#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for k in 0 4 1; do
$TR --nproc-per-node 8 --master-port $((29900+k)) bench.py --gpus 8 --steps 20 --warmup 5 --e2e-only --no-parity --run-limit $k > gpurun_out/r2rl_n8_k$k.json 2> gpurun_out/r2rl_n8_k$k.err; echo "k=$k rc=$?"
done
