#!/bin/bash
# kernel-variant sweep on the box: prints frames/s and roofline fraction per configuration
for cfg in "VEP_KERNEL=oneshot" "VEP_RPT=1 VEP_MINB=5" "VEP_RPT=1 VEP_MINB=6" "VEP_RPT=1 VEP_MINB=8" "VEP_RPT=2 VEP_MINB=3" "VEP_RPT=2 VEP_MINB=4" $EXTRA_CFGS; do
  out=$(env $cfg timeout 200 python bench.py --no-cpu-baseline --e2e-steps 3 --steps ${STEPS:-3000} ${BENCH_ARGS} 2>&1 | grep '^{' | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['roofline']['achieved'], d['roofline']['frac'], d['ms_per_step'])")
  echo "$cfg -> $out"
done
