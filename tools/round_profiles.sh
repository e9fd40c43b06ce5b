#!/bin/bash
# Regenerates the measured artefacts of a round on a GPU box: the bench line (all workloads, both arms) and the ncu launch
# lists (durations + DRAM bytes per launch) the roofline traffic figures come from.  Usage: tools/round_profiles.sh r2
R=${1:-r2}
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python bench.py --steps 20 --warmup 5 > gpurun_out/${R}_bench.json 2> gpurun_out/${R}_bench.err || echo "bench failed"
python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/${R}_bench_reference.json 2> gpurun_out/${R}_bench_reference.err
for w in cfg2 cfg3 cfg4 cfg5; do
  ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv \
      --log-file gpurun_out/${R}_ncu_launches_$w.csv python bench.py --workload $w --steps 2 --warmup 3 --no-cpu --no-e2e \
      --only-headline --no-sustained > gpurun_out/${R}_ncu_$w.log 2>&1
done
