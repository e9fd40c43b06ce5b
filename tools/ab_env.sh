#!/bin/bash
# A/B helper: tools/ab_env.sh VAR v1 v2 ... -> cfg2 ms_per_step for each value of the environment variable
VAR=$1; shift
for v in "$@"; do
  echo -n "$VAR=$v: "
  env $VAR=$v python bench.py --workload ${WORKLOAD:-cfg2} --no-cpu --no-e2e --steps 30 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['ms_per_step'], d['value'])"
done
