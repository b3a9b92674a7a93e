#!/bin/bash
# round 2, 1 GPU, validation of HEAD after the in-kernel timer / PDL-default commits: smoke, the whole GPU suite, the default bench line
#   gpurun --timeout 900 -- 'bash scratch/r02_final3.sh'
mkdir -p gpurun_out/r02y
O=gpurun_out/r02y
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo smoke rc=$?; tail -1 $O/smoke.log
B2NN_TEST_REPORT=$O/parity_report.txt timeout 600 python -m pytest tests -x -q -m gpu > $O/pytest_gpu.log 2>&1; echo pytest rc=$?; tail -4 $O/pytest_gpu.log
( time python bench.py > $O/bench_default.json 2> $O/bench_default.err ) 2>&1 | grep real; echo bench rc=$?
python - <<'PY'
import json
d=json.load(open('gpurun_out/r02y/bench_default.json'))
r=d["roofline"]
print("main ms/step %.4f value %.4g frac %.3f (burst %.3f sust %.3f) step_frac_burst %.3f e2e %.4f"%(d["ms_per_step"], d["value"], r["frac"], r["frac_burst"], r["frac_sustained"], r["step_frac_burst"], d["e2e"]["ms_per_step"]), r.get("in_kernel_timer"), d.get("sustained"), d["clocks"])
PY
( time python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_reference.json 2> $O/bench_reference.err ) 2>&1 | grep real; tail -c 600 $O/bench_reference.json
