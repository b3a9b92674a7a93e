#!/bin/bash
# coalesced GEMM epilogue + K1 (deferred W1/W2 updates, warp-per-row phase 3/4): parity first, then numbers
mkdir -p gpurun_out/r02h
O=gpurun_out/r02h
timeout 900 python -m pytest tests -q -m gpu -x > $O/pytest_gpu.log 2>&1; echo pytest rc=$?; tail -8 $O/pytest_gpu.log
python bench.py --no-cpu --no-sustained > $O/bench.json 2> $O/bench.err; echo bench rc=$?
B2NN_LIB=$PWD/graph-neural-net_b200/lib/libb2nn_s6.so python bench.py --no-cpu --no-also --no-sustained > $O/bench_s6.json 2> $O/bench_s6.err; echo bench s6 rc=$?
B2NN_GEMM_STALLS=1 python scratch/stalls.py 2>&1 | tail -8 > $O/stalls.log
python - <<'PY'
import json
for v in ("","_s6"):
    try:
        d=json.load(open("gpurun_out/r02h/bench%s.json"%v)); r=d["roofline"]
        print("variant %s: ms/step %.4f gemm avg %.4f serialised %.4f e2e %.3f frac %.3f"%(v or "product", d["ms_per_step"], r["avg_launch_ms"], r["serialised_step_ms"], d["e2e"]["ms_per_step"], r["frac"]), r["non_gemm_us"])
        for k,a in d.get("also",{}).items():
            rr=a.get("roofline",{})
            print("  also", k, {q: a.get(q) for q in ("ms_per_step","value","error")}, rr.get("phase_cycles_per_step"), (a.get("e2e") or {}).get("gradient_step_call",{}).get("us_per_call"), a.get("k2_evaluate"))
    except Exception as e: print(v, "failed", e)
PY
cat $O/stalls.log | cut -c1-330
