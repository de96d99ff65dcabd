#!/bin/bash
# round 2, 8-GPU box: both arms as the driver runs them at N = 2, 4, 8, and BASELINE configs[2] (8 cameras per GPU) at N = 8
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for n in 8 4 2; do
$TR --nproc-per-node $n --master-port $((29800+n)) bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/r2f8_bench_16x1080p_n$n.json 2> gpurun_out/r2f8_bench_16x1080p_n$n.err; echo "bench n$n rc=$?"
python bench.py --impl reference --gpus $n --steps 20 --warmup 5 > gpurun_out/r2f8_ref_16x1080p_n$n.json 2>/dev/null; echo "ref n$n rc=$?"
done
$TR --nproc-per-node 8 --master-port 29811 bench.py --gpus 8 --steps 20 --warmup 5 --workload 8x1080p > gpurun_out/r2f8_bench_8x1080p_n8.json 2> gpurun_out/r2f8_bench_8x1080p_n8.err; echo "8x1080p rc=$?"
python bench.py --impl reference --gpus 8 --steps 20 --warmup 5 --workload 8x1080p > gpurun_out/r2f8_ref_8x1080p_n8.json 2>/dev/null
