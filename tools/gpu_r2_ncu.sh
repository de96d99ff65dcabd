#!/bin/bash
mkdir -p gpurun_out
B="python bench.py --steps 20 --warmup 5 --no-cpu-baseline --skip-e2e --no-parity"
$B > gpurun_out/r2_ncu_plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/r2_launches.csv $B > gpurun_out/r2_ncu_list.log 2>&1; echo "list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:yuv_to_bgr24_tick -s 12 -c 3 -o gpurun_out/r2_prof_tick_1080p $B > gpurun_out/r2_ncu_full_1080p.log 2>&1; echo "full 1080p rc=$?"
ncu --set full --clock-control none --import-source on -k regex:yuv_to_bgr24_tick -s 16 -c 3 -o gpurun_out/r2_prof_tick_854 $B --workload 16x854x480 > gpurun_out/r2_ncu_full_854.log 2>&1; echo "full 854 rc=$?"
ncu --set full --clock-control none --import-source on -k regex:yuv_to_bgr24_tick -s 16 -c 2 -o gpurun_out/r2_prof_tick_854_tight $B --workload 16x854x480-tight > gpurun_out/r2_ncu_full_854t.log 2>&1; echo "full 854 tight rc=$?"
ls -la gpurun_out/*.ncu-rep
