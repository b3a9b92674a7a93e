#!/bin/bash
# final 1-GPU pass of the round on the banded build: smoke, the whole GPU suite, the default bench line, the reference arm,
# then the ncu evidence of the same command (launch list, one full capture of the contraction kernels)
#   gpurun --timeout 1200 -- 'bash scratch/r02_final4.sh'
mkdir -p gpurun_out/r02z2
O=gpurun_out/r02z2
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo smoke rc=$?; tail -1 $O/smoke.log
B2NN_TEST_REPORT=$O/parity_report.txt timeout 600 python -m pytest tests -x -q -m gpu > $O/pytest_gpu.log 2>&1; echo pytest rc=$?; tail -4 $O/pytest_gpu.log
( time python bench.py > $O/bench_default.json 2> $O/bench_default.err ) 2>&1 | grep real; echo bench rc=$?
python - <<'PY'
import json
d=json.load(open('gpurun_out/r02z2/bench_default.json'))
r=d["roofline"]; k=r["in_kernel_timer"]
print("main ms/step %.4f value %.4g frac %.3f (burst %.3f sust %.3f) step_frac_burst %.3f"%(d["ms_per_step"], d["value"], r["frac"], r["frac_burst"], r["frac_sustained"], r["step_frac_burst"]), "eff MHz %.0f"%k["sm_mhz_effective"], "in-kernel frac %.3f"%k["frac_burst"])
print("e2e", {q: d["e2e"].get(q) for q in ("rows","ms_per_step","value","h2d_bytes_per_step")}, "bf16 rows", d["e2e"].get("bf16_rows",{}).get("ms_per_step"))
print("sustained", d.get("sustained")); print("clocks", d["clocks"]); print("cpu", d["cpu_baseline"])
for k,v in d.get('also',{}).items():
    rr=v.get("roofline") or {}
    print(k, {q: v.get(q) for q in ("ms_per_step","value","error")}, rr.get("frac"), (v.get("e2e") or {}).get("gradient_step_call",{}).get("us_per_call"), v.get("k2_evaluate",{}).get("rows_per_s"))
PY
python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_reference.json 2> $O/bench_reference.err; echo ref rc=$?
CMD="python bench.py --steps 3 --warmup 3 --no-cpu --no-also --no-sustained"
$CMD > $O/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_c4.csv $CMD > $O/ncu_1.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemm_bf16 -s 21 -c 8 -o $O/c4_gemm $CMD > $O/ncu_2.log 2>&1
echo ncu rc=$?; ls -la $O | head -20
