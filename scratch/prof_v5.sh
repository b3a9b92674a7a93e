#!/bin/bash
CMD="python bench.py --steps 3 --warmup 3 --no-cpu --no-also"
$CMD > gpurun_out/plain_v5.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_v5.csv $CMD > gpurun_out/ncu1.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemm_bf16 -s 24 -c 8 -o gpurun_out/c4_gemm_v5 $CMD > gpurun_out/ncu2.log 2>&1 &&
ncu --set full --clock-control none -k regex:"output_delta|sgd_update|input_prologue" -s 12 -c 5 -o gpurun_out/c4_hbm_v5 $CMD > gpurun_out/ncu3.log 2>&1
echo rc=$?
tail -2 gpurun_out/ncu1.log gpurun_out/ncu2.log gpurun_out/ncu3.log
