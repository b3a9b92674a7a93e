#!/bin/bash
mkdir -p gpurun_out/r02y
O=gpurun_out/r02y
timeout 900 python -m pytest tests -q -m gpu > $O/pytest_gpu.log 2>&1; echo pytest rc=$?; tail -4 $O/pytest_gpu.log
python bench.py > $O/bench_default.json 2> $O/bench_default.err; echo bench rc=$?
python - <<'PY'
import json
d=json.load(open('gpurun_out/r02y/bench_default.json'))
r=d["roofline"]
print("main ms/step %.4f value %.4g frac %.3f e2e %.4f"%(d["ms_per_step"], d["value"], r["frac"], d["e2e"]["ms_per_step"]), d["e2e"].get("pixel_byte_rows"), d.get("sustained"))
PY
python bench.py --impl reference --steps 5 --warmup 3 | head -c 400; echo
