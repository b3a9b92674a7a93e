#!/bin/bash
# 8-GPU comparison: peer-memory exchange vs in-switch reduction
mkdir -p gpurun_out
run() { # name env... -- workload
  name=$1; wl=$2; shift 2
  env "$@" timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --workload $wl --steps 30 --warmup 5 --no-cpu 2>gpurun_out/n8_${name}.err | tee gpurun_out/n8_${name}.json | python -c "
import sys,json
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: continue
    print('$name', d['config'].get('exchange'), round(d['ms_per_step'],4), round(d['value']), round(d['e2e']['value']) if d.get('e2e') else None)
"
  grep -iE "error|Traceback" gpurun_out/n8_${name}.err | head -3
}
run p2p_c4 c4 B2NN_DP=p2p
run nvls_c4 c4 B2NN_DP=nvls
run nvls_c4_74 c4 B2NN_DP=nvls B2NN_DP_CTAS=74
run p2p_c5 c5 B2NN_DP=p2p
run nvls_c5 c5 B2NN_DP=nvls
