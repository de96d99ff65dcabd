#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for n in 2 4 8; do
  PROBE_ONLY_MERGE=1 $TR --nproc-per-node $n --master-port $((29700+n)) tools/pcie_probe.py > gpurun_out/r2i_merge_probe_n$n.log 2> gpurun_out/r2i_merge_probe_n$n.err; echo "probe n=$n rc=$?"
done
cat gpurun_out/r2i_merge_probe_n*.log | cut -c1-230
