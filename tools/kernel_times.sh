#!/bin/bash
# per-kernel duration + instruction counts of one cfg2 call (scratch helper, 16k streams)
U=${1:-16384}
python bench.py --workload ${2:-cfg2} --units $U --steps 2 --warmup 1 --no-cpu --no-e2e > /dev/null 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none -k regex:"tokenize|stream_|scan|strip_|request|sse" --csv --log-file gpurun_out/kt.csv python bench.py --workload ${2:-cfg2} --units $U --steps 2 --warmup 1 --no-cpu --no-e2e > /dev/null 2>&1
python - <<'PY'
import csv,collections
rows=[r for r in csv.reader(open('gpurun_out/kt.csv')) if len(r)>8]
h=rows[0]; ki=h.index('Kernel Name'); mi=h.index('Metric Name'); vi=h.index('Metric Value')
d=collections.OrderedDict()
for r in rows[1:]:
    try: v=float(r[vi].replace(',',''))
    except: continue
    d.setdefault((r[ki][:30],r[mi]),[]).append(v)
for (k,m),v in d.items():
    v=v[len(v)//2:]  # later launches (warm)
    print(f"{k:32s} {m:28s} n={len(v)} avg={sum(v)/len(v):.1f}")
PY
