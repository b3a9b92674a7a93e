#!/bin/bash
run() { python bench.py --steps 50 --warmup 10 --no-cpu --no-also 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$1', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], d['e2e']['last_loss'])"; }
B2NN_PROLOGUE_BULK=1 timeout 200 python -m pytest tests/test_gpu_wide.py -m gpu -x -q 2>&1 | tail -3
run base
B2NN_PROLOGUE_BULK=1 B2NN_PROLOGUE_CTAS=16 run bulk16
B2NN_PROLOGUE_BULK=1 B2NN_PROLOGUE_CTAS=32 run bulk32
B2NN_PROLOGUE_BULK=1 B2NN_PROLOGUE_CTAS=74 run bulk74
B2NN_PROLOGUE_BULK=1 B2NN_PROLOGUE_CTAS=148 run bulk148
run base
