#!/bin/bash
mkdir -p gpurun_out/r02n4
O=gpurun_out/r02n4
python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29551 bench.py --gpus 4 --no-cpu > $O/bench_n4.json 2> $O/bench_n4.err; echo bench rc=$?
python - <<'PY'
import json
d=json.load(open("gpurun_out/r02n4/bench_n4.json"))
print("ms/step %.4f value %.4g"%(d["ms_per_step"], d["value"]), d["config"]["exchange"], d.get("dp_check"), "e2e %.3f"%d["e2e"]["ms_per_step"], d.get("sustained"))
for k,v in d.get("also",{}).items():
    print("  also", k, {q: v.get(q) for q in ("ms_per_step","value","batch_per_gpu","exchange","step_tflops_per_gpu","error")})
PY
