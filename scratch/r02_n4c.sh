#!/bin/bash
# 4 GPUs, final check of the default bench line with the in-switch exchange beside programmatic dependent launches and the
# banded tile order (dp_check against the single-GPU run; C5 strong scaling in `also`)
#   gpurun --gpus 4 --timeout 600 -- 'bash scratch/r02_n4c.sh'
mkdir -p gpurun_out/r02n4c
O=gpurun_out/r02n4c
( time timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29543 bench.py --gpus 4 --no-cpu > $O/bench_n4.json 2> $O/bench_n4.err ) 2>&1 | grep real; echo bench rc=$?
python -c "
import json; d=json.load(open('$O/bench_n4.json')); print('N=4: ms/step %.4f'%d['ms_per_step'], 'value %.4g'%d['value'], d.get('dp_check'), 'e2e %.4f'%d['e2e']['ms_per_step'], d['e2e'].get('rows'), 'exchange', d['config'].get('exchange'));
c=d.get('also',{}).get('c5',{}); print('c5:', c.get('ms_per_step'), c.get('value'), c.get('error'))"
