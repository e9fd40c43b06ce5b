#!/bin/bash
# Regenerates the measured artefacts of a round on a GPU box: bench lines for every workload, the reference arm, and the
# ncu launch lists (durations + DRAM bytes) the roofline traffic figures come from.  Usage: tools/round_profiles.sh r1
R=${1:-r1}
mkdir -p gpurun_out
for w in cfg2 cfg3 cfg4 cfg5; do
  python bench.py --workload $w > gpurun_out/${R}_bench_$w.json 2> gpurun_out/${R}_bench_$w.err || echo "bench $w failed"
done
python bench.py --impl reference > gpurun_out/${R}_bench_reference_cfg2.json 2> gpurun_out/${R}_bench_reference_cfg2.err
for w in cfg2 cfg3 cfg4; do
  ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv \
      --log-file gpurun_out/${R}_ncu_launches_$w.csv python bench.py --workload $w --steps 2 --warmup 3 --no-cpu --no-e2e \
      > gpurun_out/${R}_ncu_$w.log 2>&1
done
