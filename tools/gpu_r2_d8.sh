#!/bin/bash
# 8-GPU box: DMA probe at N = 1, 2, 4, 8 and the end-to-end variants at N = 8 (and the full line at N = 2, 4, 8)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
nvidia-smi topo -m > gpurun_out/r2d_topo.txt 2>&1
lscpu > gpurun_out/r2d_lscpu.txt 2>&1
numactl -H >> gpurun_out/r2d_lscpu.txt 2>&1
for n in 1 2 4 8; do
  $TR --nproc-per-node $n --master-port $((29600+n)) tools/pcie_probe.py > gpurun_out/r2d_pcie_probe_n$n.log 2> gpurun_out/r2d_pcie_probe_n$n.err; echo "probe n=$n rc=$?"
done
$TR --nproc-per-node 8 --master-port 29611 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/r2d_bench_n8.json 2> gpurun_out/r2d_bench_n8.err; echo "bench n8 rc=$?"
python bench.py --impl reference --gpus 8 --steps 20 --warmup 5 > gpurun_out/r2d_ref_n8.json 2>/dev/null; echo "ref n8 rc=$?"
VEP_BENCH_NO_AFFINITY=1 $TR --nproc-per-node 8 --master-port 29612 bench.py --gpus 8 --steps 20 --warmup 5 --e2e-only --no-parity > gpurun_out/r2d_e2e_n8_noaff.json 2> gpurun_out/r2d_e2e_n8_noaff.err; echo "noaff rc=$?"
$TR --nproc-per-node 8 --master-port 29613 bench.py --gpus 8 --steps 20 --warmup 5 --e2e-only --no-parity --e2e-submit threads > gpurun_out/r2d_e2e_n8_threads.json 2> gpurun_out/r2d_e2e_n8_threads.err; echo "threads rc=$?"
for k in 0 1; do
$TR --nproc-per-node 8 --master-port $((29620+k)) bench.py --gpus 8 --steps 20 --warmup 5 --e2e-only --no-parity --staging-threads $k > gpurun_out/r2d_e2e_n8_st$k.json 2> gpurun_out/r2d_e2e_n8_st$k.err; echo "st$k rc=$?"
done
$TR --nproc-per-node 8 --master-port 29614 bench.py --gpus 8 --steps 20 --warmup 5 --workload 8x1080p > gpurun_out/r2d_bench_n8_8x1080p.json 2> gpurun_out/r2d_bench_n8_8x1080p.err; echo "8x1080p rc=$?"
for n in 2 4; do
$TR --nproc-per-node $n --master-port $((29630+n)) bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/r2d_bench_n$n.json 2> gpurun_out/r2d_bench_n$n.err; echo "bench n$n rc=$?"
done
