#!/bin/bash
# A/B on one box: round-1 bench.py against the new one (serialised-pass GEMM times, e2e), the general FP64 path per call,
# then the GPU suite with the one-epoch test
mkdir -p gpurun_out/r02d
O=gpurun_out/r02d
python scratch/bench_r01.py --no-cpu --no-also > $O/old.json 2> $O/old.err; echo old rc=$?
python bench.py --no-cpu --no-also --no-sustained > $O/new_nosus.json 2> $O/new_nosus.err; echo new rc=$?
python bench.py --no-cpu --no-also > $O/new_sus.json 2> $O/new_sus.err; echo new rc=$?
B2NN_F64_FUSED=0 python bench.py --workload c2 --steps 2000 --no-cpu > $O/c2_general.json 2> $O/c2_general.err; echo c2 general rc=$?
python - <<'PY'
import json
for f in ("old","new_nosus","new_sus"):
    try:
        d=json.load(open("gpurun_out/r02d/%s.json"%f))
        r=d["roofline"]
        print(f, "ms/step %.4f"%d["ms_per_step"], "gemm avg %.4f ms"%r["avg_launch_ms"], "serialised %.4f"%r["serialised_step_ms"], "e2e ms %.3f"%d["e2e"]["ms_per_step"], "loss", d["e2e"]["last_loss"], r.get("non_gemm_us"), d.get("sustained"))
    except Exception as e:
        print(f, "failed", e)
try:
    d=json.load(open("gpurun_out/r02d/c2_general.json"))
    print("c2 general path: us/step", d["ms_per_step"]*1e3, d["e2e"].get("gradient_step_call"))
except Exception as e:
    print("c2 general failed", e)
PY
tail -3 $O/*.err
timeout 600 python -m pytest tests -q -m gpu --durations=8 > $O/pytest_gpu.log 2>&1; echo pytest rc=$?; tail -15 $O/pytest_gpu.log
