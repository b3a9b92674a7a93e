#!/bin/bash
# final 1-GPU pass of the round: smoke, the whole GPU suite, the default bench line, then the ncu evidence of the same command
mkdir -p gpurun_out/r02z
O=gpurun_out/r02z
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo smoke rc=$?; tail -1 $O/smoke.log
B2NN_TEST_REPORT=$O/parity_report.txt timeout 900 python -m pytest tests -q -m gpu > $O/pytest_gpu.log 2>&1; echo pytest rc=$?; tail -4 $O/pytest_gpu.log
( time python bench.py > $O/bench_default.json 2> $O/bench_default.err ) 2>&1 | grep real; echo bench rc=$?
python - <<'PY'
import json
d=json.load(open('gpurun_out/r02z/bench_default.json'))
r=d["roofline"]
print("main ms/step %.4f value %.4g frac %.3f (burst %.3f sust %.3f) step_frac_burst %.3f e2e %.4f"%(d["ms_per_step"], d["value"], r["frac"], r["frac_burst"], r["frac_sustained"], r["step_frac_burst"], d["e2e"]["ms_per_step"]), r["non_gemm_us"], d.get("sustained"), d["cpu_baseline"], d["clocks"])
for k,v in d.get('also',{}).items():
    rr=v.get("roofline") or {}
    print(k, {q: v.get(q) for q in ("ms_per_step","value","error")}, rr.get("frac"), rr.get("cycles_per_step"), (v.get("e2e") or {}).get("gradient_step_call",{}).get("us_per_call"), v.get("k2_evaluate",{}).get("rows_per_s"), v.get("cpu_baseline",{}))
PY
B2NN_GEMM_STALLS=1 python scratch/stalls.py 2>&1 | tail -8 | cut -c1-330 > $O/stalls.log; cat $O/stalls.log
CMD="python bench.py --steps 3 --warmup 3 --no-cpu --no-also --no-sustained"
$CMD > $O/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_c4.csv $CMD > $O/ncu_1.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemm_bf16 -s 21 -c 8 -o $O/c4_gemm $CMD > $O/ncu_2.log 2>&1
echo ncu rc=$?
