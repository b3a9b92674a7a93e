#!/bin/bash
run() { python bench.py --steps 50 --warmup 10 --no-cpu --no-also 2>/dev/null | python -c "
import sys,json
try:
    d=json.loads(sys.stdin.read()); print('$1', round(d['ms_per_step'],4), 'e2e', round(d['e2e']['ms_per_step'],4))
except Exception as e: print('$1 failed')"; }
run base
B2NN_DYNAMIC_TILES=1 run dyn_base
for c in 8 16 32 64; do
B2NN_PROLOGUE_CTAS=$c B2NN_PROLOGUE_THREADS=256 run static_c${c}
B2NN_DYNAMIC_TILES=1 B2NN_PROLOGUE_CTAS=$c B2NN_PROLOGUE_THREADS=256 run dyn_c${c}
done
