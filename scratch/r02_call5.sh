#!/bin/bash
# GPU suite, the reference's experiment on the surrogate set, the default bench line
mkdir -p gpurun_out/r02e
O=gpurun_out/r02e
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo smoke rc=$?; tail -1 $O/smoke.log
timeout 600 python -m pytest tests -q -m gpu --durations=5 > $O/pytest_gpu.log 2>&1; echo pytest rc=$?; tail -12 $O/pytest_gpu.log
timeout 600 python scratch/mnist_experiment.py --out $O/mnist > $O/mnist.log 2>&1; echo mnist rc=$?; cat $O/mnist.log | cut -c1-600
rm -rf $O/mnist/data
( time timeout 900 python bench.py > $O/bench_default.json 2> $O/bench_default.err ) 2>&1 | tail -4; echo bench rc=$?
tail -3 $O/bench_default.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r02e/bench_default.json'))
print("main", d["ms_per_step"], d["roofline"]["frac"], d["e2e"]["ms_per_step"], d.get("sustained"))
for k,v in d.get('also',{}).items():
    print(k, {q: v.get(q) for q in ("ms_per_step","value","error")}, (v.get("e2e") or {}).get("gradient_step_call",{}).get("us_per_call"))
PY
