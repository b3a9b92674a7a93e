#!/bin/bash
# First GPU call of round 2: everything that was written after round 1's GPU budget ran out, then the TF32 numbers.
#   gpurun --timeout 900 -- 'bash scratch/round2_first_call.sh'
# Not yet run on a B200 at the end of round 1: smoke()'s TF32 step, tests/test_gpu_zz_fullsize.py,
# bench.py --precision tf32, scratch/tf32_peak.py (2 GPUs: tests/test_gpu_zz_dp_tf32.py).
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
B2NN_TEST_REPORT=gpurun_out/tf32_report.txt timeout 600 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1
echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"; tail -c 600 gpurun_out/bench_default.json
bash scratch/prof_tf32.sh
