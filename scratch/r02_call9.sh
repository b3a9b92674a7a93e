#!/bin/bash
# K1 with the deferred W1/W2 updates only: parity, then C2 / C1 / C3
mkdir -p gpurun_out/r02i
O=gpurun_out/r02i
timeout 600 python -m pytest tests/test_gpu_f64.py tests/test_gpu_mnist.py -q -m gpu > $O/pytest_f64.log 2>&1; echo pytest rc=$?; tail -3 $O/pytest_f64.log
for w in c2 c1 c3; do
  python bench.py --workload $w --steps 20000 --warmup 200 --no-cpu > $O/bench_$w.json 2> $O/bench_$w.err
  python - <<PY
import json
d=json.load(open("$O/bench_$w.json"))
r=d["roofline"]; p=r["phase_cycles_per_step"]
print("$w: %.2f us/step, cycles %.0f, %s; grad_step call %.1f us; k2 %s" % (d["ms_per_step"]*1e3, r["cycles_per_step"], {k:int(v) for k,v in p.items() if v>50}, d["e2e"]["gradient_step_call"]["us_per_call"], d.get("k2_evaluate")))
PY
done
python bench.py --no-cpu --no-also --no-sustained > $O/bench_c4.json 2> $O/bench_c4.err; python -c "
import json; d=json.load(open('$O/bench_c4.json')); print('c4 ms/step', d['ms_per_step'], 'gemm avg', d['roofline']['avg_launch_ms'], 'e2e', d['e2e']['ms_per_step'])"
