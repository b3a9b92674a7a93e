#!/bin/bash
# 2 GPUs: the tests added since the first 2-GPU call (BF16 wire, bounded wait, two devices in one process), the whole DP file
# again on the current kernels, and the TF32 comparator at the batch size that failed at world 8 (single GPU batch 2048)
mkdir -p gpurun_out/r02n2b
O=gpurun_out/r02n2b
timeout 900 python -m pytest tests/test_dp.py -q -m gpu -s > $O/pytest_dp.log 2>&1; echo pytest rc=$?; grep -v "^$" $O/pytest_dp.log | grep "passed\|failed\|dp worker\|bf16 wire\|exchange:\|FAILED" | cut -c1-250
B2NN_TEST_ROWS_PER_RANK=1024 timeout 300 python -m pytest tests/test_dp.py -q -m gpu -s -k "world_matches and tf32 and 2" > $O/pytest_dp_1024.log 2>&1; echo pytest1024 rc=$?; grep "passed\|failed\|dp worker\|FAILED" $O/pytest_dp_1024.log | cut -c1-250
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 2 --no-cpu --no-also --no-sustained --wire bf16 > $O/bench_n2_wire.json 2> $O/bench_n2_wire.err; echo rc=$?
python -c "
import json; d=json.load(open('$O/bench_n2_wire.json')); print('wire bf16 N=2: ms/step', d['ms_per_step'], d['config'].get('gradient_wire'), d.get('dp_check'))"
