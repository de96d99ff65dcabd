#!/bin/bash
# variant sweep: overlapped + strict-order GB/s per workload
for wl in ${WLS:-16x1080p 8x4k-nv12}; do
for cfg in "VEP_RPT=2 VEP_MINB=4" "VEP_RPT=2 VEP_MINB=3" "VEP_RPT=1 VEP_MINB=6" "VEP_RPT=1 VEP_MINB=5" $EXTRA_CFGS; do
  out=$(env $cfg timeout 200 python bench.py --workload $wl --no-cpu-baseline --e2e-steps 3 --steps ${STEPS:-3000} 2>&1 | grep '^{' | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['roofline']['achieved'], d['roofline']['frac'], 'strict', d['roofline']['strict_stream_order']['achieved'], d['roofline']['strict_stream_order']['frac'])")
  echo "$wl $cfg -> $out"
done; done
