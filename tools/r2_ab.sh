#!/bin/bash
# A/B of the implementations behind the ABI on one GPU: device-resident timing only (no CPU leg, no e2e).
# usage: tools/r2_ab.sh "cfg2 cfg3" "token general"  [extra env assignments]
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for wl in $1; do
  for path in $2; do
    echo "== $wl $path $3"
    env QB200_PATH=$path $3 python bench.py --workload $wl --steps 20 --warmup 3 --no-cpu --no-e2e --only-headline --no-sustained 2>&1 | tail -1 | python -c "
import json,sys
l=json.loads(sys.stdin.readline())
print('ms', round(l['ms_per_step'],4), 'GB/s', round(l['value'],1), 'frac', round(l['roofline']['frac'],3))
for k in l['roofline'].get('kernels',[]): print('   ', k['name'], round(k['ms_per_step'],4))
"
  done
done
