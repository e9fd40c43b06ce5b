cd /root/repo
python -m pytest tests -x -q -m gpu > gpurun_out/r2f_gputests.log 2>&1; tail -2 gpurun_out/r2f_gputests.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
bash tools/round_profiles.sh r2
python tools/bench_widened.py > gpurun_out/r2_bench_widened.json 2> gpurun_out/r2_bench_widened.err
for w in cfg2 cfg3 cfg5; do
  ncu --set full --clock-control none --import-source on -k regex:"tokenize|strip_|stream_|request_" -c 6 -f -o gpurun_out/r2_full_$w python bench.py --workload $w --steps 1 --warmup 0 --no-cpu --no-e2e --only-headline --no-sustained > gpurun_out/ncu_full_$w.log 2>&1
done
ls -la gpurun_out/*.ncu-rep
