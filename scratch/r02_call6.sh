#!/bin/bash
# K1 after the sharded barrier + early staging: parity tests, then C1/C2/C3 with 1 / 4 / 16 / 32 arrival counters
mkdir -p gpurun_out/r02f
O=gpurun_out/r02f
timeout 600 python -m pytest tests/test_gpu_f64.py tests/test_gpu_mnist.py -q -m gpu > $O/pytest_f64.log 2>&1; echo pytest rc=$?; tail -4 $O/pytest_f64.log
for s in 1 4 16 32; do
  for w in c2 c1 c3; do
    B2NN_K1_SHARDS=$s python bench.py --workload $w --steps 20000 --warmup 200 --no-cpu > $O/bench_${w}_s$s.json 2> $O/bench_${w}_s$s.err
    python - <<PY
import json
d=json.load(open("$O/bench_${w}_s$s.json"))
r=d["roofline"]; p=r["phase_cycles_per_step"]
print("$w shards $s: %.2f us/step, cycles %.0f, barrier1 %.0f barrier2 %.0f stage %.0f wait %.0f phase1 %.0f p5 %.0f; grad_step call %.1f us" % (d["ms_per_step"]*1e3, r["cycles_per_step"], p["barrier1"], p["barrier2"], p["stage"], p["stage:wait"], p["phase1"], p["p5:W0"]+p["p5:W1c"]+p["p5:W1r"]+p["p5:W2"], d["e2e"]["gradient_step_call"]["us_per_call"]))
PY
  done
done
