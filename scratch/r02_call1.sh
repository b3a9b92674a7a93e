#!/bin/bash
# round 2, GPU call 1 (1 GPU): numbers that round 1 never took -- the TF32 cuBLAS rate, C4 in TF32 (bench line, launch
# list, one full capture), the C1 / C2 / C3 bench lines with the K1 phase counters, and a B2NN_TRACE timeline of C4.
#   gpurun --timeout 840 -- 'bash scratch/r02_call1.sh'
mkdir -p gpurun_out/r02a
O=gpurun_out/r02a
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > $O/gpu.txt 2>&1
python scratch/tf32_peak.py > $O/tf32_peak.log 2>&1; cp gpurun_out/tf32_peaks.json $O/ 2>/dev/null
mkdir -p profiles; cp gpurun_out/tf32_peaks.json profiles/tf32_peaks.json 2>/dev/null
for w in c1 c3 c2; do
  python bench.py --workload $w --steps 20000 --warmup 200 > $O/bench_$w.json 2> $O/bench_$w.err; echo "bench $w rc=$?"
  B2NN_K1_PROF=1 python bench.py --workload $w --steps 2000 --warmup 200 --no-cpu > $O/bench_${w}_prof.json 2> $O/bench_${w}_prof.err
  grep "K1 prof" $O/bench_${w}_prof.err | tail -2
done
python bench.py --precision tf32 --no-cpu --no-also > $O/bench_c4_tf32.json 2> $O/bench_c4_tf32.err; echo "bench tf32 rc=$?"
B2NN_TRACE=1 python bench.py --steps 8 --warmup 3 --no-cpu --no-also > $O/bench_c4_trace.json 2> $O/bench_c4_trace.err
B2NN_GEMM_STALLS=1 python scratch/stalls.py > $O/stalls_c4.log 2>&1
CMD="python bench.py --precision tf32 --steps 3 --warmup 3 --no-cpu --no-also"
$CMD > $O/plain_tf32.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_tf32.csv $CMD > $O/ncu_tf32_1.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemm_tf32 -s 24 -c 8 -o $O/c4_gemm_tf32 $CMD > $O/ncu_tf32_2.log 2>&1
echo ncu rc=$?
tail -3 $O/tf32_peak.log
for f in $O/bench_c1.json $O/bench_c3.json $O/bench_c2.json $O/bench_c4_tf32.json; do echo $f; head -c 1500 $f; echo; done
