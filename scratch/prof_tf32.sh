#!/bin/bash
# round 2, first GPU call: TF32 throughput of C4, the TF32 cuBLAS rate to put it against, launch list and one full
# capture of the 4096^3 TF32 contraction.  Outputs under gpurun_out/ (copy the summaries to profiles/).
mkdir -p gpurun_out
python scratch/tf32_peak.py > gpurun_out/tf32_peak.log 2>&1
CMD="python bench.py --precision tf32 --steps 3 --warmup 3 --no-cpu --no-also"
python bench.py --precision tf32 --no-cpu --no-also > gpurun_out/bench_c4_tf32.json 2> gpurun_out/bench_c4_tf32.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_tf32.csv $CMD > gpurun_out/ncu_tf32_1.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemm_tf32 -s 24 -c 8 -o gpurun_out/c4_gemm_tf32 $CMD > gpurun_out/ncu_tf32_2.log 2>&1
echo rc=$?
tail -3 gpurun_out/tf32_peak.log gpurun_out/bench_c4_tf32.json gpurun_out/ncu_tf32_1.log gpurun_out/ncu_tf32_2.log
