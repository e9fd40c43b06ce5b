#!/bin/bash
# rows f1/f2 of tools/bench_widened.py in short form
cd "$(dirname "$0")/.."
python tools/bench_widened.py 2>/dev/null | head -2 | python -c "
import json,sys
for ln in sys.stdin:
    l=json.loads(ln); print(l['row'], round(l['ms_per_step'],4), 'frac', round(l['roofline']['frac'],3), l['roofline']['kernels_ms'])"
