#!/bin/bash
mkdir -p gpurun_out/r02n4b
O=gpurun_out/r02n4b
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
i=0
for cfg in "B2NN_DP_CTAS=74" "B2NN_DP_CTAS=296" "B2NN_DP_CTAS=148" "B2NN_BENCH_WIRE=bf16" "B2NN_DP=p2p" "B2NN_DP=nccl"; do
  i=$((i+1))
  env $cfg $TR --master-port $((29560+i)) bench.py --gpus 4 --no-cpu --no-also --no-sustained > $O/bench_$i.json 2> $O/bench_$i.err
  python -c "
import json; d=json.load(open('$O/bench_$i.json')); print('$cfg', 'ms/step %.4f'%d['ms_per_step'], d['config']['exchange'], d['config'].get('gradient_wire'), 'dp_check %.2e'%d['dp_check']['rel'], 'e2e %.3f'%d['e2e']['ms_per_step'])"
done
