#!/bin/bash
# 2-GPU check of the NVSwitch multicast exchange
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_dp.py -m gpu -x -q -s 2>&1 | tail -15 > gpurun_out/nvls_test.log
cat gpurun_out/nvls_test.log
for mode in p2p nvls; do
 for wl in c4 c5; do
  B2NN_DP=$mode timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --workload $wl --steps 30 --warmup 5 --no-cpu 2>gpurun_out/nvls_${mode}_${wl}.err | tee gpurun_out/nvls_${mode}_${wl}.json | python -c "
import sys,json
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: continue
    print('$mode $wl', d['config'].get('exchange'), d['ms_per_step'], d['value'], d['e2e']['value'] if d.get('e2e') else None)
"
  tail -3 gpurun_out/nvls_${mode}_${wl}.err
 done
done
