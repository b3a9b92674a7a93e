#!/bin/bash
run() { name=$1; shift; python bench.py --steps 50 --warmup 10 --no-cpu --no-also "$@" 2>/dev/null | python -c "
import sys,json
try:
    d=json.loads(sys.stdin.read()); print('$name', round(d['ms_per_step'],4), 'e2e', round(d['e2e']['ms_per_step'],4))
except Exception as e: print('$name failed')"; }
for i in 1 2 3; do
run vec
B2NN_PROLOGUE_SCALAR=1 run scalar
done
