#!/bin/bash
run() { name=$1; shift; timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 30 --warmup 5 "$@" 2>gpurun_out/f8_$name.err | tee gpurun_out/f8_$name.json | python -c "
import sys,json
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: continue
    print('$name', d['config'].get('exchange'), round(d['ms_per_step'],4), round(d['value']), 'e2e', round(d['e2e']['ms_per_step'],4), round(d['e2e']['value']), d['e2e'].get('rows'))
"; grep -iE "error|Traceback" gpurun_out/f8_$name.err | head -3; }
run default
run u8 --e2e-rows u8
