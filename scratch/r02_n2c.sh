#!/bin/bash
# 2 GPUs, final check of the data-parallel path on the current default build (programmatic dependent launch on, banded
# tile order): the DP parity file, then the default 2-GPU bench line (dp_check, C5 strong scaling in `also`)
#   gpurun --gpus 2 --timeout 900 -- 'bash scratch/r02_n2c.sh'
mkdir -p gpurun_out/r02n2c
O=gpurun_out/r02n2c
timeout 600 python -m pytest tests/test_dp.py tests/test_gpu_zz_dp_tf32.py -x -q -m gpu > $O/pytest_dp.log 2>&1; echo pytest rc=$?; tail -3 $O/pytest_dp.log
( time python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --no-cpu > $O/bench_n2.json 2> $O/bench_n2.err ) 2>&1 | grep real; echo bench rc=$?
python -c "
import json; d=json.load(open('$O/bench_n2.json')); print('N=2: ms/step %.4f'%d['ms_per_step'], 'value %.4g'%d['value'], d.get('dp_check'), 'e2e %.4f'%d['e2e']['ms_per_step'], d['e2e'].get('rows'), 'exchange', d['config'].get('exchange'));
c=d.get('also',{}).get('c5',{}); print('c5:', c.get('ms_per_step'), c.get('value'), c.get('dp_check'), c.get('error'))"
