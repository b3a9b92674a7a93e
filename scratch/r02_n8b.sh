#!/bin/bash
# 8 GPUs, final code: the driver's SCALE command, then the TF32 in-switch parity test with diagnostics
mkdir -p gpurun_out/r02n8b
O=gpurun_out/r02n8b
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
$TR --master-port 29541 bench.py --gpus 8 --no-cpu > $O/bench_n8.json 2> $O/bench_n8.err; echo bench rc=$?
python - <<'PY'
import json
d=json.load(open("gpurun_out/r02n8b/bench_n8.json"))
print("ms/step %.4f value %.4g"%(d["ms_per_step"], d["value"]), d["config"]["exchange"], d.get("dp_check"), "e2e %.3f"%d["e2e"]["ms_per_step"], d.get("sustained"))
for k,v in d.get("also",{}).items():
    print("  also", k, {q: v.get(q) for q in ("ms_per_step","value","batch_per_gpu","exchange","step_tflops_per_gpu","error")})
PY
B2NN_TEST_DIAG=1 timeout 300 python -m pytest tests/test_dp.py -q -m gpu -s -k "(tf32-nvls-8 or tf32-p2p-8 or bf16_wire and nvls-8)" > $O/pytest_dp.log 2>&1; echo pytest rc=$?
grep "passed\|failed\|dp worker\|bf16 wire\|FAILED\|^E " $O/pytest_dp.log | cut -c1-300
